// common.cuh -- device/host helpers shared by every kernel of the engine.
#pragma once
#include <cuda_runtime.h>
#include <cstdio>
#include <cstdlib>
#include <cstring>
#include <cmath>

// Fatal-error convention of the reference: print and exit (backend_cuda.t:26-40, 767-779).
#define OPTK_CUDA(call)                                                                        \
    do {                                                                                       \
        cudaError_t e_ = (call);                                                               \
        if (e_ != cudaSuccess) {                                                               \
            printf("Cuda reported error %d: %s\nIn call: %s\nIn file: %s:%d\n", (int)e_,       \
                   cudaGetErrorString(e_), #call, __FILE__, __LINE__);                         \
            fflush(stdout);                                                                    \
            exit((int)e_);                                                                     \
        }                                                                                      \
    } while (0)

namespace optk {

// Device-resident reduction scalars (reference: scanAlphaNumerator/Denominator, scanBetaNumerator,
// modelCost, scratch, q; solverGPUGaussNewton.t:213-235).  They are kept in double and are
// rotated through small rings so the PCG loop needs no memset between kernels:
//   iteration l:  alphaNum = RZ[l%3], betaNum accumulates into RZ[(l+1)%3], alphaDen = ADEN[l&1]
enum Scal { S_COST = 0, S_MODEL = 1, S_Q = 2, S_ADEN0 = 3, S_ADEN1 = 4, S_RZ0 = 5, S_RZ1 = 6, S_RZ2 = 7, S_DUMMY = 8, S_QPREV = 9, S_STOP = 10, S_COUNT = 16 };
// S_QPREV / S_STOP: Levenberg-Marquardt inner loop run from the device side (Engine::pcg_run_lm): the previous
// iteration's q and the "q_tolerance reached" flag that turns the step of a speculatively queued iteration into a no-op.

// View of one solver vector.  UB == 0: SoA, plane u holds channel u of every site (x fastest).
// UB == U > 0: blocked layout for functors whose J^T J kernel stages tiles with the bulk-copy engine
// (VEC_UB in the functor): sites are grouped in blocks of 32 consecutive sites and a block stores its U channels
// back to back, [site / 32][u][site % 32], so the U channels of a 32-site row segment are ONE contiguous 128 U-byte
// chunk.  The flat streaming kernels do not care (all vectors of a plan share the layout); site kernels go
// through operator().
template <class T, int UB = 0>
struct Vec {
    T* d;
    long plane;
    __device__ __forceinline__ static long index(int u, long s, long plane) {
        if constexpr (UB == 0) return (long)u * plane + s;
        else return (s >> 5) * (32 * UB) + u * 32 + (s & 31);
    }
    __device__ __forceinline__ T& operator()(int u, long s) const { return d[index(u, s, plane)]; }
    static long host_index(int u, long s, long plane) {
        if (UB == 0) return (long)u * plane + s;
        return (s >> 5) * (32 * UB) + u * 32 + (s & 31);
    }
};

__device__ __forceinline__ double warp_sum(double v) {
#pragma unroll
    for (int o = 16; o > 0; o >>= 1) v += __shfl_down_sync(0xffffffffu, v, o);
    return v;
}

__device__ __forceinline__ int linear_tid() {
    return threadIdx.x + blockDim.x * (threadIdx.y + blockDim.y * threadIdx.z);
}

// Block-wide sum of one per-thread partial, then ONE atomicAdd(double) per block.
// (reference: warpReduce + lane-0 red.global.add per warp, solver:466-472.)  Must be reached by
// every thread of the block.
__device__ __forceinline__ void block_accumulate(double v, double* target) {
    __shared__ double part[32];
    const int tid = linear_tid(), lane = tid & 31, warp = tid >> 5;
    const int nwarps = (blockDim.x * blockDim.y * blockDim.z + 31) >> 5;
    v = warp_sum(v);
    if (lane == 0) part[warp] = v;
    __syncthreads();
    if (warp == 0) {
        double t = lane < nwarps ? part[lane] : 0.0;
        t = warp_sum(t);
        if (lane == 0) atomicAdd(target, t);
    }
    __syncthreads();
}

// CERES-style guarded inverse of the Jacobi preconditioner (solver:478-487)
template <class T> __device__ __forceinline__ T guarded_invert(T d) {
    T t = T(1) + sqrt(d);
    return T(1) / (t * t);
}

template <class T> struct VecIO;
template <> struct VecIO<float> {
    static constexpr int W = 4;
    using V = float4;
};
template <> struct VecIO<double> {
    static constexpr int W = 2;
    using V = double2;
};
__device__ __forceinline__ float& comp(float4& v, int i) { return (&v.x)[i]; }
__device__ __forceinline__ double& comp(double2& v, int i) { return (&v.x)[i]; }

inline long round_up(long n, long m) { return (n + m - 1) / m * m; }

}  // namespace optk
