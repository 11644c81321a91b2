// common.cuh -- device/host helpers shared by every kernel of the engine.
#pragma once
#include <cuda_runtime.h>
#include <cstdio>
#include <cstdlib>
#include <cstring>
#include <cmath>

// Fatal-error convention.  The reference prints and exits (backend_cuda.t:26-40, 767-779); a shared library that does so kills its
// host application, so every fatal path goes through optk_fatal() (registry.cu).  With no handler installed
// (OptB200_SetErrorHandler, include/OptB200.h) that is still the reference's behaviour: the message on stdout, then exit(code).
// With a handler it throws optk::Fatal instead, which the extern "C" layer catches (optk_guard): the handler is called once with
// the code and the message, the plan is marked failed, and that call and every later call on the plan return at once.
namespace optk { struct Fatal { int code; }; }
extern "C" void optk_fatal(int code, const char* fmt, ...) __attribute__((format(printf, 2, 3)));
#define OPTK_CUDA(call)                                                                        \
    do {                                                                                       \
        cudaError_t e_ = (call);                                                               \
        if (e_ != cudaSuccess)                                                                 \
            optk_fatal((int)e_, "Cuda reported error %d: %s\nIn call: %s\nIn file: %s:%d\n", (int)e_, \
                       cudaGetErrorString(e_), #call, __FILE__, __LINE__);                     \
    } while (0)

namespace optk {

// Device-resident reduction scalars (reference: scanAlphaNumerator/Denominator, scanBetaNumerator,
// modelCost, scratch, q; solverGPUGaussNewton.t:213-235).  They are kept in double and are
// rotated through small rings so the PCG loop needs no memset between kernels:
//   iteration l:  alphaNum = RZ[l%3], betaNum accumulates into RZ[(l+1)%3], alphaDen = ADEN[l&1]
enum Scal { S_COST = 0, S_MODEL = 1, S_Q = 2, S_ADEN0 = 3, S_ADEN1 = 4, S_RZ0 = 5, S_RZ1 = 6, S_RZ2 = 7, S_DUMMY = 8, S_QPREV = 9, S_STOP = 10, S_APREV = 11, S_COUNT = 16 };
// S_APREV: alpha of the previous PCG iteration, kept while its delta update is deferred (k_pcg_step2, DM)
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
    // UB == -U < 0: AoS, the U channels of a site are adjacent ([site][u]).  For graph functors whose J^T J kernel GATHERS the
    // direction of a vertex' neighbours (arap): a neighbour is one 24-byte record (1-2 sectors) instead of U sectors in U planes.
    __device__ __forceinline__ static long index(int u, long s, long plane) {
        if constexpr (UB == 0) return (long)u * plane + s;
        else if constexpr (UB > 0) return (s >> 5) * (32 * UB) + u * 32 + (s & 31);
        else return s * (-UB) + u;
    }
    __device__ __forceinline__ T& operator()(int u, long s) const { return d[index(u, s, plane)]; }
    static long host_index(int u, long s, long plane) {
        if (UB == 0) return (long)u * plane + s;
        if (UB > 0) return (s >> 5) * (32 * UB) + u * 32 + (s & 31);
        return s * (-UB) + u;
    }
};

// all U channels of a site: generic accessors (SoA / blocked / lazily fused directions) read them one by one, the AoS
// layout with 64-bit (float) / 128-bit (double) accesses
template <int U, class P, class T> __device__ __forceinline__ void load_channels(const P& p, long s, T* o) {
#pragma unroll
    for (int u = 0; u < U; ++u) o[u] = p(u, s);
}
template <int U> __device__ __forceinline__ void load_channels(const Vec<float, -U>& p, long s, float* o) {
    static_assert(U % 2 == 0, "AoS rows of an even number of floats");
    const float2* q = reinterpret_cast<const float2*>(p.d + s * U);
#pragma unroll
    for (int k = 0; k < U / 2; ++k) { const float2 v = q[k]; o[2 * k] = v.x; o[2 * k + 1] = v.y; }
}
template <int U> __device__ __forceinline__ void load_channels(const Vec<double, -U>& p, long s, double* o) {
    static_assert(U % 2 == 0, "AoS rows of an even number of doubles");
    const double2* q = reinterpret_cast<const double2*>(p.d + s * U);
#pragma unroll
    for (int k = 0; k < U / 2; ++k) { const double2 v = q[k]; o[2 * k] = v.x; o[2 * k + 1] = v.y; }
}
template <int U, class V, class T> __device__ __forceinline__ void store_channels(const V& p, long s, const T* o) {
#pragma unroll
    for (int u = 0; u < U; ++u) p(u, s) = o[u];
}
template <int U> __device__ __forceinline__ void store_channels(const Vec<float, -U>& p, long s, const float* o) {
    float2* q = reinterpret_cast<float2*>(p.d + s * U);
#pragma unroll
    for (int k = 0; k < U / 2; ++k) q[k] = make_float2(o[2 * k], o[2 * k + 1]);
}
template <int U> __device__ __forceinline__ void store_channels(const Vec<double, -U>& p, long s, const double* o) {
    double2* q = reinterpret_cast<double2*>(p.d + s * U);
#pragma unroll
    for (int k = 0; k < U / 2; ++k) q[k] = make_double2(o[2 * k], o[2 * k + 1]);
}

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
