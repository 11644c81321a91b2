// Stand-alone probe of the TMA wrappers used by functors/volumetric_packed.cuh.  usage: tma_probe <variant>
//   0: tensor map nested in a __grid_constant__ struct, 4-D box     1: tensor map as its own __grid_constant__ parameter
//   2: like 1 but a plain 2-D map / 2-D box
#include <cstdio>
#include <vector>
#include <map>
#include "../../opt_b200/csrc/engine.cuh"
#include "../../opt_b200/csrc/functors/volumetric_packed.cuh"
using namespace optk;

struct Holder { int a, b; alignas(64) CUtensorMap m0, m1; float w; };
constexpr int BOXF = 32 * 10 * 6;

__device__ void body(const CUtensorMap* m, float* out, int c0, int c1, int c2, int rank, int nf) {
    extern __shared__ unsigned char raw[];
    float* ring = reinterpret_cast<float*>((reinterpret_cast<unsigned long long>(raw) + 127) & ~127ull);
    unsigned long long* bar = reinterpret_cast<unsigned long long*>(ring + BOXF);
    if (threadIdx.x == 0) {
        tma::mbar_init(bar, 1);
        asm volatile("fence.mbarrier_init.release.cluster;" ::: "memory");
        asm volatile("fence.proxy.async.shared::cta;" ::: "memory");
    }
    __syncthreads();
    if (threadIdx.x == 0) {
        tma::mbar_expect_tx(bar, nf * 4);
        if (rank == 4) tma::load4d(ring, m, bar, c0, c1, c2, 0);
        else asm volatile("cp.async.bulk.tensor.2d.shared::cluster.global.mbarrier::complete_tx::bytes [%0], [%1, {%2, %3}], [%4];"
                          ::"r"(tma::smem_u32(ring)), "l"((unsigned long long)m), "r"(c0), "r"(c1), "r"(tma::smem_u32(bar)) : "memory");
    }
    tma::mbar_wait(bar, 0);
    for (int i = threadIdx.x; i < nf; i += blockDim.x) out[i] = ring[i];
}
__global__ void k0(const __grid_constant__ Holder h, float* out, int c0, int c1, int c2) { body(&h.m1, out, c0, c1, c2, 4, BOXF); }
__global__ void k1(const __grid_constant__ CUtensorMap m, float* out, int c0, int c1, int c2) { body(&m, out, c0, c1, c2, 4, BOXF); }
__global__ void k2(const __grid_constant__ CUtensorMap m, float* out, int c0, int c1, int c2) { body(&m, out, c0, c1, c2, 2, 32 * 10); }

int main(int argc, char** argv) {
    const int variant = argc > 1 ? atoi(argv[1]) : 0;
    const int W = 64, H = 20, D = 5, P = 6; const long plane = 64 * ((long)W * H * D / 64 + 1);
    std::vector<float> h(P * plane);
    for (size_t i = 0; i < h.size(); ++i) h[i] = (float)i;
    float *d, *o; cudaMalloc(&d, h.size() * 4); cudaMalloc(&o, BOXF * 4);
    cudaMemset(o, 0, BOXF * 4);
    cudaMemcpy(d, h.data(), h.size() * 4, cudaMemcpyHostToDevice);
    Holder hd; hd.a = 1; hd.b = 2; hd.w = 3;
    hd.m1 = optk_grid_map(d, W, H, D, P, plane);
    hd.m0 = hd.m1;
    const int smem = BOXF * 4 + 256;
    int planes = P;
    if (variant == 0) { cudaFuncSetAttribute(k0, cudaFuncAttributeMaxDynamicSharedMemorySize, smem); k0<<<1, 128, smem>>>(hd, o, 29, -1, 2); }
    if (variant == 1) { cudaFuncSetAttribute(k1, cudaFuncAttributeMaxDynamicSharedMemorySize, smem); k1<<<1, 128, smem>>>(hd.m1, o, 29, -1, 2); }
    if (variant == 2) {
        CUtensorMap m;
        const cuuint64_t dims[2] = {(cuuint64_t)W, (cuuint64_t)H * D};
        const cuuint64_t strides[1] = {(cuuint64_t)W * 4};
        const cuuint32_t box[2] = {32, 10}, es[2] = {1, 1};
        CUresult r = optk_encode_tiled()(&m, CU_TENSOR_MAP_DATA_TYPE_FLOAT32, 2, d, dims, strides, box, es, CU_TENSOR_MAP_INTERLEAVE_NONE,
                                         CU_TENSOR_MAP_SWIZZLE_NONE, CU_TENSOR_MAP_L2_PROMOTION_NONE, CU_TENSOR_MAP_FLOAT_OOB_FILL_NONE);
        printf("encode 2d: %d\n", (int)r);
        cudaFuncSetAttribute(k2, cudaFuncAttributeMaxDynamicSharedMemorySize, smem); k2<<<1, 128, smem>>>(m, o, 29, 2 * H - 1, 0);
        planes = 1;
    }
    cudaError_t e = cudaDeviceSynchronize();
    printf("variant %d sync: %s\n", variant, cudaGetErrorString(e));
    std::vector<float> r(BOXF);
    cudaMemcpy(r.data(), o, r.size() * 4, cudaMemcpyDeviceToHost);
    int bad = 0;
    for (int pl = 0; pl < planes; ++pl) for (int row = 0; row < 10; ++row) for (int x = 0; x < 32; ++x) {
        int gx = 29 + x, gy = -1 + row, gz = 2;
        float want = (gx < 0 || gx >= W || gy < 0 || gy >= H) ? 0.f : (float)(pl * plane + ((long)gz * H + gy) * W + gx);
        if (variant == 2) want = gx >= W ? 0.f : (float)(((long)gz * H + gy) * W + gx);
        if (r[(pl * 10 + row) * 32 + x] != want) { if (bad < 3) printf("mismatch pl %d row %d x %d: %f vs %f\n", pl, row, x, r[(pl * 10 + row) * 32 + x], want); ++bad; }
    }
    printf("bad = %d\n", bad);
    return 0;
}
