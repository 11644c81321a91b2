// the CUDA programming guide's TMA sample (libcu++ API), to compare against the hand-written PTX wrappers
#include <cstdio>
#include <vector>
#include <cuda.h>
#include <cuda_runtime.h>
#include <cuda/barrier>
using barrier = cuda::barrier<cuda::thread_scope_block>;
namespace cde = cuda::device::experimental;

__global__ void kref(const __grid_constant__ CUtensorMap tensor_map, float* out, int x, int y) {
    __shared__ alignas(128) float smem_buffer[10][32];
#pragma nv_diag_suppress static_var_with_dynamic_init
    __shared__ barrier bar;
    if (threadIdx.x == 0) { init(&bar, blockDim.x); cde::fence_proxy_async_shared_cta(); }
    __syncthreads();
    barrier::arrival_token token;
    if (threadIdx.x == 0) {
        cde::cp_async_bulk_tensor_2d_global_to_shared(&smem_buffer, &tensor_map, x, y, bar);
        token = cuda::device::barrier_arrive_tx(bar, 1, sizeof(smem_buffer));
    } else token = bar.arrive();
    bar.wait(std::move(token));
    for (int i = threadIdx.x; i < 320; i += blockDim.x) out[i] = (&smem_buffer[0][0])[i];
}
typedef CUresult (*Enc)(CUtensorMap*, CUtensorMapDataType, cuuint32_t, void*, const cuuint64_t*, const cuuint64_t*, const cuuint32_t*, const cuuint32_t*,
                        CUtensorMapInterleave, CUtensorMapSwizzle, CUtensorMapL2promotion, CUtensorMapFloatOOBfill);
int main() {
    const int W = 64, H = 100;
    std::vector<float> h(W * H);
    for (size_t i = 0; i < h.size(); ++i) h[i] = (float)i;
    float *d, *o; cudaMalloc(&d, h.size() * 4); cudaMalloc(&o, 320 * 4);
    cudaMemcpy(d, h.data(), h.size() * 4, cudaMemcpyHostToDevice);
    void* p = nullptr; cudaDriverEntryPointQueryResult q;
    cudaGetDriverEntryPoint("cuTensorMapEncodeTiled", &p, cudaEnableDefault, &q);
    CUtensorMap m;
    const cuuint64_t dims[2] = {W, H}; const cuuint64_t strides[1] = {W * 4};
    const cuuint32_t box[2] = {32, 10}, es[2] = {1, 1};
    CUresult r = ((Enc)p)(&m, CU_TENSOR_MAP_DATA_TYPE_FLOAT32, 2, d, dims, strides, box, es, CU_TENSOR_MAP_INTERLEAVE_NONE,
                          CU_TENSOR_MAP_SWIZZLE_NONE, CU_TENSOR_MAP_L2_PROMOTION_NONE, CU_TENSOR_MAP_FLOAT_OOB_FILL_NONE);
    printf("encode %d  q=%d d=%p\n", (int)r, (int)q, (void*)d);
    for (int i = 0; i < 16; ++i) printf("%016llx%c", ((unsigned long long*)&m)[i], i % 4 == 3 ? '\n' : ' ');
    int drv = 0, rt = 0; cudaDriverGetVersion(&drv); cudaRuntimeGetVersion(&rt); printf("driver %d runtime %d\n", drv, rt);
    cudaDeviceProp pr; cudaGetDeviceProperties(&pr, 0); printf("%s cc %d.%d\n", pr.name, pr.major, pr.minor);
    kref<<<1, 128>>>(m, o, 29, 39);
    printf("sync: %s\n", cudaGetErrorString(cudaDeviceSynchronize()));
    std::vector<float> res(320);
    cudaMemcpy(res.data(), o, 320 * 4, cudaMemcpyDeviceToHost);
    int bad = 0;
    for (int row = 0; row < 10; ++row) for (int x = 0; x < 32; ++x) {
        int gx = 29 + x, gy = 39 + row;
        float want = gx >= W ? 0.f : (float)(gy * W + gx);
        if (res[row * 32 + x] != want) ++bad;
    }
    printf("bad = %d (res[0]=%f res[1]=%f)\n", bad, res[0], res[1]);
}
