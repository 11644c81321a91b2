// probe: mbarrier alone (variant 0) and the descriptor-less 1-D bulk copy (variant 1)
#include <cstdio>
#include <vector>
#include <cuda_runtime.h>
__device__ __forceinline__ unsigned smem_u32(const void* p) { return (unsigned)__cvta_generic_to_shared(p); }
__global__ void k(const float* src, float* out, int variant) {
    __shared__ alignas(128) float buf[320];
    __shared__ alignas(8) unsigned long long bar;
    if (threadIdx.x == 0) {
        asm volatile("mbarrier.init.shared::cta.b64 [%0], %1;" ::"r"(smem_u32(&bar)), "r"(1) : "memory");
        asm volatile("fence.mbarrier_init.release.cluster;" ::: "memory");
        asm volatile("fence.proxy.async.shared::cta;" ::: "memory");
    }
    for (int i = threadIdx.x; i < 320; i += blockDim.x) buf[i] = -1.f;
    __syncthreads();
    if (threadIdx.x == 0) {
        if (variant == 0) {
            asm volatile("mbarrier.arrive.shared::cta.b64 _, [%0];" ::"r"(smem_u32(&bar)) : "memory");
        } else {
            asm volatile("mbarrier.arrive.expect_tx.shared::cta.b64 _, [%0], %1;" ::"r"(smem_u32(&bar)), "r"(10 * 128) : "memory");
            for (int r = 0; r < 10; ++r)
                asm volatile("cp.async.bulk.shared::cluster.global.mbarrier::complete_tx::bytes [%0], [%1], %2, [%3];"
                             ::"r"(smem_u32(buf + r * 32)), "l"(src + r * 64 + 4), "r"(128), "r"(smem_u32(&bar)) : "memory");
        }
    }
    unsigned ok;
    do {
        asm volatile("{\n .reg .pred p;\n mbarrier.try_wait.parity.shared::cta.b64 p, [%1], %2;\n selp.b32 %0, 1, 0, p;\n}"
                     : "=r"(ok) : "r"(smem_u32(&bar)), "r"(0) : "memory");
    } while (!ok);
    for (int i = threadIdx.x; i < 320; i += blockDim.x) out[i] = buf[i];
}
int main(int argc, char** argv) {
    const int variant = argc > 1 ? atoi(argv[1]) : 0;
    std::vector<float> h(6400);
    for (size_t i = 0; i < h.size(); ++i) h[i] = (float)i;
    float *d, *o; cudaMalloc(&d, h.size() * 4); cudaMalloc(&o, 320 * 4);
    cudaMemcpy(d, h.data(), h.size() * 4, cudaMemcpyHostToDevice);
    k<<<1, 128>>>(d, o, variant);
    printf("variant %d sync: %s\n", variant, cudaGetErrorString(cudaDeviceSynchronize()));
    std::vector<float> res(320);
    cudaMemcpy(res.data(), o, 320 * 4, cudaMemcpyDeviceToHost);
    int bad = 0;
    for (int r = 0; r < 10; ++r) for (int x = 0; x < 32; ++x) if (res[r * 32 + x] != (variant ? (float)(r * 64 + 4 + x) : -1.f)) ++bad;
    printf("bad = %d (res[0]=%f res[33]=%f)\n", bad, res[0], res[33]);
}
