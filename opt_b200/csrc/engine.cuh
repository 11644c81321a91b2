// engine.cuh -- the generic GN/LM + PCG engine, templated over a per-example functor.
//
// Replaces, for one example, the kernel set the reference JIT-compiles from
// solverGPUGaussNewton.t:521-1345 plus the generated fmap.* functions.  Layout:
//   * every solver vector (delta r b Adelta z p Ap CtC pre SSq prevX) is SoA: U planes of Npad
//     elements, x fastest inside a plane (the reference is AoS per image, o.t:1056-1100);
//   * the streaming kernels (PCGStep2/3 and friends) treat a vector as one flat array of
//     U*Npad elements and move it with 128-bit loads/stores.  They need no exclude test:
//     excluded sites hold 0 in delta/r/p/z/Ap from allocation on and every update maps 0 to 0;
//   * reductions are block-level (warp shuffle -> smem -> one double atomicAdd per block) into
//     device-resident double scalars that the next kernel reads (no host sync in the GN loop).
#pragma once
#include <algorithm>
#include <map>
#include <string>
#include <vector>
#include <nccl.h>
#include <nvtx3/nvToolsExt.h>      // header-only; ranges are no-ops unless a profiler injects itself
#include <cub/device/device_radix_sort.cuh>
#include <cub/device/device_select.cuh>
#include "common.cuh"
#include "csr.cuh"
#include "functors/math.cuh"
#include "optk.h"

#define OPTK_NCCL(call)                                                                        \
    do {                                                                                       \
        ncclResult_t r_ = (call);                                                              \
        if (r_ != ncclSuccess)                                                                 \
            optk_fatal(200 + (int)r_, "NCCL reported error %d: %s\nIn call: %s\nIn file: %s:%d\n", (int)r_, \
                       ncclGetErrorString(r_), #call, __FILE__, __LINE__);                     \
    } while (0)

// The C handle is the abstract base class of every Engine<F>.
struct OptK_Plan {
    virtual ~OptK_Plan() {}
    virtual void bind(void** pp) = 0;
    virtual void set_lm(double radius, double min_diag, double max_diag) = 0;
    virtual void precompute() = 0;
    virtual double cost() = 0;
    virtual void pcg_begin(int nIter) = 0;
    virtual void pcg_iter(int lIter, int reset) = 0;
    virtual void pcg_run(int lIterations) = 0;
    virtual int pcg_run_lm(int lIterations, int reset_period, double q_tolerance, double* zeta, int* fired) = 0;
    virtual double fetch_q() = 0;
    virtual double model_cost() = 0;
    virtual void save_unknowns() = 0;
    virtual void linear_update() = 0;
    virtual void revert_update() = 0;
    virtual int unknown_channels() const = 0;
    virtual long site_count() const = 0;
    virtual int debug_read(const char* name, void* host_out) = 0;
    virtual int debug_read_range(const char* name, long first_site, long nsites, void* host_out) = 0;
    virtual void debug_apply_jtj(const void* host_p, void* host_out) = 0;
    virtual int set_slab(int rank, int nranks, int halo, const void* nccl_unique_id) = 0;
    virtual int set_materialized(int mode) = 0;          // 0 matrix-free, 1 J and J^T, 2 fused J^T J; returns the mode in effect
    virtual long materialized_nnz(int which) const = 0;   // 0 J, 1 J^T J

    bool failed = false;          // a fatal error was reported to the installed handler: every later call returns at once

    // ---- launch bookkeeping shared by all engines -------------------------------------------
    cudaStream_t stream = nullptr;
    cudaEvent_t ev_in = nullptr, ev_out = nullptr;
    long launches = 0;
    bool timing = false;
    struct Timed { const char* name; cudaEvent_t a, b; };
    std::vector<Timed> timed;

    void base_init(bool collect_timing) {
        timing = collect_timing;
        OPTK_CUDA(cudaStreamCreateWithFlags(&stream, cudaStreamNonBlocking));
        OPTK_CUDA(cudaEventCreateWithFlags(&ev_in, cudaEventDisableTiming));
        OPTK_CUDA(cudaEventCreateWithFlags(&ev_out, cudaEventDisableTiming));
    }
    void base_destroy() {
        for (auto& t : timed) { cudaEventDestroy(t.a); cudaEventDestroy(t.b); }
        if (ev_in) cudaEventDestroy(ev_in);
        if (ev_out) cudaEventDestroy(ev_out);
        if (stream) cudaStreamDestroy(stream);
    }
    // The reference runs on the default stream (SURVEY 8b "Threading").  The engine runs on its
    // own stream (so the PCG loop can be captured in a CUDA graph) and fences it against the
    // legacy default stream at both ends of every API call, which preserves those semantics.
    void begin_call() {
        OPTK_CUDA(cudaEventRecord(ev_in, cudaStreamLegacy));
        OPTK_CUDA(cudaStreamWaitEvent(stream, ev_in, 0));
    }
    void end_call() {
        OPTK_CUDA(cudaEventRecord(ev_out, stream));
        OPTK_CUDA(cudaStreamWaitEvent(cudaStreamLegacy, ev_out, 0));
    }
    void pre_launch(const char* name) {
        ++launches;
        if (timing) {
            Timed t; t.name = name;
            OPTK_CUDA(cudaEventCreate(&t.a)); OPTK_CUDA(cudaEventCreate(&t.b));
            OPTK_CUDA(cudaEventRecord(t.a, stream));
            timed.push_back(t);
        }
    }
    void post_launch() {
        OPTK_CUDA(cudaGetLastError());                      // backend_cuda.t:827
        if (timing) OPTK_CUDA(cudaEventRecord(timed.back().b, stream));
    }
    // every kernel of the engine goes through here: launch counter, optional event pair, error check
    template <class... KA, class... A>
    void launch(const char* name, dim3 grid, dim3 block, void (*kernel)(KA...), A... args) {
        pre_launch(name);
        kernel<<<grid, block, 0, stream>>>(args...);
        post_launch();
    }
    // total milliseconds and launch count of all timed launches called `name` (timing must be on)
    int kernel_time(const char* name, double* total_ms) {
        OPTK_CUDA(cudaStreamSynchronize(stream));
        int n = 0; double tot = 0;
        for (auto& t : timed) if (!strcmp(t.name, name)) { float ms = 0; cudaEventElapsedTime(&ms, t.a, t.b); tot += ms; ++n; }
        *total_ms = tot;
        return n;
    }
    void reset_timing() {
        OPTK_CUDA(cudaStreamSynchronize(stream));
        for (auto& t : timed) { cudaEventDestroy(t.a); cudaEventDestroy(t.b); }
        timed.clear();
    }
    // persistent kernels: as many CTAs as can be co-resident (occupancy x SM count), never more than
    // there are tiles -- a larger grid would run in waves and leave a tail
    int n_sm = 0;
    std::map<const void*, int> occ_cache;
    template <class... KA>
    dim3 persistent_grid(void (*kernel)(KA...), int threads, long ntiles, size_t smem = 0) {
        if (!n_sm) { int dev = 0; cudaGetDevice(&dev); cudaDeviceGetAttribute(&n_sm, cudaDevAttrMultiProcessorCount, dev); }
        auto it = occ_cache.find((const void*)kernel);
        int nb;
        if (it == occ_cache.end()) {
            nb = 1;
            OPTK_CUDA(cudaOccupancyMaxActiveBlocksPerMultiprocessor(&nb, kernel, threads, smem));
            if (nb < 1) nb = 1;
            occ_cache[(const void*)kernel] = nb;
        } else nb = it->second;
        long g = (long)n_sm * nb;
        if (g > ntiles) g = ntiles;
        if (g < 1) g = 1;
        return dim3((unsigned)g);
    }
    template <class... KA, class... A>
    void launch_tiles(const char* name, long ntiles, dim3 block, void (*kernel)(KA...), A... args) {
        launch(name, persistent_grid(kernel, (int)(block.x * block.y * block.z), ntiles), block, kernel, args...);
    }
    // same with dynamic shared memory (TMA-staged kernels): opt in to the large carve-out once per kernel
    template <class... KA, class... A>
    void launch_tiles_smem(const char* name, long ntiles, dim3 block, size_t smem, void (*kernel)(KA...), A... args) {
        if (smem == 0) { launch_tiles(name, ntiles, block, kernel, args...); return; }
        if (!occ_cache.count((const void*)kernel))
            OPTK_CUDA(cudaFuncSetAttribute(kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem));
        const dim3 grid = persistent_grid(kernel, (int)(block.x * block.y * block.z), ntiles, smem);
        pre_launch(name);
        kernel<<<grid, block, smem, stream>>>(args...);
        post_launch();
    }
    // The reference brackets a whole solve with an event pair called "overall" (solver:1813, 1906): started at Init,
    // appended to the event list when the solve is cleaned up, so it is the last row of the table.
    Timed overall_{nullptr, nullptr, nullptr};
    void overall_begin() {
        if (!timing) return;
        if (overall_.a) { cudaEventDestroy(overall_.a); cudaEventDestroy(overall_.b); }
        overall_.name = "overall";
        OPTK_CUDA(cudaEventCreate(&overall_.a)); OPTK_CUDA(cudaEventCreate(&overall_.b));
        OPTK_CUDA(cudaEventRecord(overall_.a, stream));
    }
    void overall_end() {
        if (!timing || !overall_.a) return;
        OPTK_CUDA(cudaEventRecord(overall_.b, stream));
        timed.push_back(overall_);
        overall_ = Timed{nullptr, nullptr, nullptr};
    }
    // Timer:evaluate (backend_cuda.t:231-297): the kernel table, the `TIMING` line (totals of the kernels whose name starts
    // with PCGInit1 / PCGStep1 / overall, in table order) and `Per-iter times ms (nonlinear,linear)` with the reference's own
    // aggregation rule: the summed totals of all kernels launched as often as PCGInit1, and as often as PCGStep1
    // (examples/outParse.py reads these two lines).
    void print_timing() {
        if (!timing || timed.empty()) return;
        OPTK_CUDA(cudaStreamSynchronize(stream));
        std::map<std::string, std::pair<int, double>> agg;
        std::vector<std::string> order;
        for (auto& t : timed) {
            float ms = 0; cudaEventElapsedTime(&ms, t.a, t.b);
            if (!agg.count(t.name)) order.push_back(t.name);
            agg[t.name].first++; agg[t.name].second += ms;
        }
        auto prefix = [](const char* pre, const std::string& s) { return s.compare(0, strlen(pre), pre) == 0; };
        printf("--------------------------------------------------------\n");
        printf("        Kernel        |   Count  |   Total   | Average \n");
        printf("----------------------+----------+-----------+----------\n");
        for (auto& n : order) {
            printf("----------------------+----------+-----------+----------\n");
            printf(" %-20s |   %4d   | %8.3fms| %7.4fms\n", n.c_str(), agg[n].first, agg[n].second, agg[n].second / agg[n].first);
        }
        printf("--------------------------------------------------------\n");
        printf("TIMING ");
        for (auto& n : order)
            if (prefix("PCGInit1", n) || prefix("PCGStep1", n) || prefix("overall", n)) printf("%f ", agg[n].second);
        printf("\n");
        int linIters = 0, nonLinIters = 0;
        for (auto& n : order) {
            if (prefix("PCGInit1", n)) linIters = agg[n].first;
            if (prefix("PCGStep1", n)) nonLinIters = agg[n].first;
        }
        float linAggregate = 0.f, nonLinAggregate = 0.f;
        for (auto& n : order) {
            if (agg[n].first == linIters) linAggregate += (float)agg[n].second;
            if (agg[n].first == nonLinIters) nonLinAggregate += (float)agg[n].second;
        }
        printf("Per-iter times ms (nonlinear,linear): %7.4f\t%7.4f\n", linAggregate, nonLinAggregate);
    }
};

namespace optk {

template <class F> using VecOf = Vec<typename F::Scalar, F::VEC_UB>;     // solver-vector view in the functor's layout

// ------------------------------------------------------------------------------ site mapping
// Site kernels are persistent: a fixed grid (a few CTAs per SM) loops over the virtual tiles of the
// index space, so a kernel does one reduction atomic per resident CTA instead of one per tile (65 536
// same-address atomics per launch on a 4096^2 image otherwise) and pays no CTA-dispatch cost for
// tiles that are entirely masked out.
struct TileGrid {
    int gx, gy, gz;     // tiles per dimension
    long n;             // gx*gy*gz
};
#define OPTK_TILE_LOOP(tg) for (long tile_ = blockIdx.x; tile_ < (tg).n; tile_ += gridDim.x)

template <class F>
__device__ __forceinline__ bool thread_site(const F& f, const TileGrid& tg, long tile, int& x, int& y, int& z, long& s) {
    const int bx = (int)(tile % tg.gx), by = (int)((tile / tg.gx) % tg.gy), bz = (int)(tile / ((long)tg.gx * tg.gy));
    x = bx * blockDim.x + threadIdx.x;
    y = by * blockDim.y + threadIdx.y;
    z = bz * blockDim.z + threadIdx.z;
    s = x + (long)f.W * (y + (long)f.H * z);
    // slab mode: only rows (2-D) / planes (3-D) [o_lo, o_hi) are owned by this rank; the rest of the
    // local array are ghost copies of the neighbours' boundary and are only ever read
    const int slow = F::ND == 3 ? z : y;
    return x < f.W && y < f.H && z < f.D && slow >= f.o_lo && slow < f.o_hi;
}

// ------------------------------------------------------------------------------ site kernels
// PCGInit1 (solver:521-563).  For graph problems the raw diagonal is stored and inverted later.
template <class F>
__global__ void k_pcg_init1(F f, TileGrid tg, VecOf<F> delta, VecOf<F> r,
                            VecOf<F> p, VecOf<F> pre, double* aNum) {
    using T = typename F::Scalar;
    constexpr int U = F::U;
    int x, y, z; long s;
    double acc = 0.0;
    OPTK_TILE_LOOP(tg)
    if (thread_site(f, tg, tile_, x, y, z, s)) {
        T g[U], d[U];
#pragma unroll
        for (int u = 0; u < U; ++u) g[u] = d[u] = T(0);
        const bool ex = f.exclude(x, y, z, s);
        if (!ex) {
            f.evalJTF(x, y, z, s, g, d);
#pragma unroll
            for (int u = 0; u < U; ++u) {
                delta(u, s) = T(0);
                g[u] = -g[u];
                r(u, s) = g[u];
                if (!F::USE_PRE) d[u] = T(1);
            }
            if (F::GV == 0) {
#pragma unroll
                for (int u = 0; u < U; ++u) {
                    d[u] = guarded_invert(d[u]);
                    T pp = d[u] * g[u];
                    p(u, s) = pp;
                    acc += (double)(g[u] * pp);
                }
            }
        }
#pragma unroll
        for (int u = 0; u < U; ++u) pre(u, s) = d[u];
    }
    if (F::GV == 0) block_accumulate(acc, aNum);
}

// Fused PCGStep3 + PCGStep1 (solver:814-845 + 607-632): the new search direction p' = z + beta p is
// formed on the fly wherever the stencil reads it, written once at the thread's own site into the
// OTHER p buffer (ping-pong: neighbours still read the old p), and Ap' = J^T J p' is produced in the
// same pass.  Saves one read of p and one kernel per PCG iteration.
template <class V, class T>
struct FusedDir {
    V z, p;
    T beta;
    __device__ __forceinline__ T operator()(int u, long s) const { return z(u, s) + beta * p(u, s); }
};
// all channels of the fused direction at a site of AoS vectors: two vectorised record reads instead of 2 U scalar ones
template <int U, class T> __device__ __forceinline__ void load_channels(const FusedDir<Vec<T, -U>, T>& d, long s, T* o) {
    T a[U], b[U];
    load_channels<U>(d.z, s, a);
    load_channels<U>(d.p, s, b);
#pragma unroll
    for (int u = 0; u < U; ++u) o[u] = a[u] + d.beta * b[u];
}
template <class F, bool LM>
__global__ void k_apply_jtj_fused(F f, TileGrid tg, VecOf<F> z, VecOf<F> p_old,
                                  VecOf<F> p_new, VecOf<F> out,
                                  VecOf<F> CtC, const double* bNum, const double* aNum,
                                  double* den, double* zero_a, double* zero_b) {
    using T = typename F::Scalar;
    constexpr int U = F::U;
    int x, y, zc; long s;
    double acc = 0.0;
    FusedDir<VecOf<F>, T> dir{z, p_old, (T)(*bNum) / (T)(*aNum)};
    if (blockIdx.x == 0 && blockIdx.y == 0 && blockIdx.z == 0 && linear_tid() == 0) {
        if (zero_a) *zero_a = 0.0;
        if (zero_b) *zero_b = 0.0;
    }
    OPTK_TILE_LOOP(tg)
    if (thread_site(f, tg, tile_, x, y, zc, s) && !f.exclude(x, y, zc, s)) {
        T o[U];
        f.applyJTJ(x, y, zc, s, dir, o);
#pragma unroll
        for (int u = 0; u < U; ++u) {
            T pv = dir(u, s);
            p_new(u, s) = pv;
            if (LM) o[u] += CtC(u, s) * pv;
            out(u, s) = o[u];
            acc += (double)(pv * o[u]);
        }
    }
    block_accumulate(acc, den);
}

// PCGStep1 / computeAdelta (solver:607-632, 873-878): out = (J^T J [+ CtC]) dir over non-excluded sites
template <class F, bool LM>
__global__ void k_apply_jtj(F f, TileGrid tg, VecOf<F> dir, VecOf<F> out,
                            VecOf<F> CtC, double* den, double* zero_slot) {
    using T = typename F::Scalar;
    constexpr int U = F::U;
    int x, y, z; long s;
    double acc = 0.0;
    if (zero_slot && blockIdx.x == 0 && blockIdx.y == 0 && blockIdx.z == 0 && linear_tid() == 0) *zero_slot = 0.0;
    OPTK_TILE_LOOP(tg)
    if (thread_site(f, tg, tile_, x, y, z, s) && !f.exclude(x, y, z, s)) {
        T o[U];
        f.applyJTJ(x, y, z, s, dir, o);
#pragma unroll
        for (int u = 0; u < U; ++u) {
            T pv = dir(u, s);
            if (LM) o[u] += CtC(u, s) * pv;
            out(u, s) = o[u];
            acc += (double)(pv * o[u]);
        }
    }
    block_accumulate(acc, den);
}

// computeCost (solver:971-997)
template <class F>
__global__ void k_cost(F f, TileGrid tg, double* target) {
    using T = typename F::Scalar;
    int x, y, z; long s;
    double acc = 0.0;
    OPTK_TILE_LOOP(tg)
    if (thread_site(f, tg, tile_, x, y, z, s) && !f.exclude(x, y, z, s)) acc += (double)(T)(0.5 * (double)f.cost(x, y, z, s));
    block_accumulate(acc, target);
}

// computeModelCost (solver:1105-1118)
template <class F>
__global__ void k_model_cost(F f, TileGrid tg, VecOf<F> delta, double* target) {
    using T = typename F::Scalar;
    int x, y, z; long s;
    double acc = 0.0;
    OPTK_TILE_LOOP(tg)
    if (thread_site(f, tg, tile_, x, y, z, s) && !f.exclude(x, y, z, s)) acc += (double)(T(0.5) * f.modelCost(x, y, z, s, delta));
    block_accumulate(acc, target);
}

// PCGComputeCtC (solver:1042-1048): CtC = diag(J^T J) / radius
template <class F, bool GENERIC = false>
__global__ void k_compute_ctc(F f, TileGrid tg, VecOf<F> CtC, typename F::Scalar inv_radius) {   // GENERIC: always through evalJTF
    using T = typename F::Scalar;
    constexpr int U = F::U;
    int x, y, z; long s;
    OPTK_TILE_LOOP(tg)
    if (thread_site(f, tg, tile_, x, y, z, s) && !f.exclude(x, y, z, s)) {
        T g[U], d[U];
#pragma unroll
        for (int u = 0; u < U; ++u) g[u] = d[u] = T(0);
        if constexpr (F::HAS_DIAG_JTJ && !GENERIC) f.diagJTJ(x, y, z, s, d);
        else f.evalJTF(x, y, z, s, g, d);
#pragma unroll
        for (int u = 0; u < U; ++u) CtC(u, s) = d[u] * inv_radius;
    }
}

// PCGFinalizeDiagonal (solver:1061-1099)
template <class F>
__global__ void k_finalize_diagonal(F f, TileGrid tg, VecOf<F> CtC, VecOf<F> SSq,
                                    VecOf<F> pre, VecOf<F> r,
                                    VecOf<F> b, VecOf<F> p,
                                    VecOf<F> delta, typename F::Scalar radius,
                                    typename F::Scalar min_diag, typename F::Scalar max_diag,
                                    double* aNum, double* q) {
    using T = typename F::Scalar;
    constexpr int U = F::U;
    int x, y, z; long s;
    double acc = 0.0, qacc = 0.0;
    OPTK_TILE_LOOP(tg)
    if (thread_site(f, tg, tile_, x, y, z, s) && !f.exclude(x, y, z, s)) {
        T qq = T(0);
#pragma unroll
        for (int u = 0; u < U; ++u) {
            T unclamped = CtC(u, s);
            T invS = T(1) / SSq(u, s);
            T mult = invS / radius;
            T c = fmin(fmax(unclamped, min_diag * mult), max_diag * mult);
            CtC(u, s) = c;
            T m = T(1) / (c + radius * unclamped);
            pre(u, s) = m;
            T res = r(u, s);
            b(u, s) = res;
            T pp = m * res;
            p(u, s) = pp;
            acc += (double)(res * pp);
            qq += delta(u, s) * (res + res);
        }
        qacc += (double)(T(0.5) * qq);
    }
    block_accumulate(acc, aNum);
    block_accumulate(qacc, q);
}

// PCGLinearUpdate / savePreviousUnknowns / revertUpdate (solver:854-887); MODE 0: X += delta,
// 1: prevX = X, 2: X = prevX
template <class F, int MODE>
__global__ void k_unknown_update(F f, TileGrid tg, VecOf<F> v) {
    constexpr int U = F::U;
    int x, y, z; long s;
    OPTK_TILE_LOOP(tg)
    if (thread_site(f, tg, tile_, x, y, z, s) && !f.exclude(x, y, z, s)) {
#pragma unroll
        for (int u = 0; u < U; ++u) {
            typename F::Scalar* xp = f.xptr(s, u);
            if (MODE == 0) *xp = *xp + v(u, s);
            else if (MODE == 1) v(u, s) = *xp;
            else *xp = v(u, s);
        }
    }
}

// Single-channel unknowns (U = 1): the unknown image and the solver vectors share one flat layout, and excluded sites
// carry delta = 0 / prevX = X, so the three updates are plain 128-bit streaming passes over all sites.
template <class T, int MODE>
__global__ void __launch_bounds__(256) k_flat_update(long n4, T* __restrict__ X, T* __restrict__ v) {
    for (long i = blockIdx.x * (long)blockDim.x + threadIdx.x; i < n4; i += (long)gridDim.x * blockDim.x) {
        T a[4], b[4];
        if (MODE != 2) load4(X + 4 * i, a);
        if (MODE != 1) load4(v + 4 * i, b);
        if (MODE == 0) {
#pragma unroll
            for (int j = 0; j < 4; ++j) a[j] += b[j];
            st4v(X + 4 * i, a);
        } else if (MODE == 1) st4v(v + 4 * i, a);
        else st4v(X + 4 * i, b);
    }
}

// precompute (solver:1028-1039) runs over EVERY local site, ghost rows included: the per-site tables are
// read over the stencil, and the ghost unknowns were refreshed by the halo exchange just before
template <class F>
__global__ void k_precompute(F f, TileGrid tg) {
    int x, y, z; long s;
    f.o_lo = 0; f.o_hi = 0x7fffffff;
    OPTK_TILE_LOOP(tg)
    if (thread_site(f, tg, tile_, x, y, z, s)) f.precompute(x, y, z, s);
}

// ------------------------------------------------------------------------------ graph kernels
// PCGInit1_Graph (solver:1188-1197): scatter -dr/du * r into r, (dr/du)^2 into pre
template <class F>
__global__ void k_graph_jtf(F f, long E, VecOf<F> r, VecOf<F> pre) {
    long e = blockIdx.x * (long)blockDim.x + threadIdx.x;
    if (e < E) f.graphJTF(e, r, pre);
}
// PCGInit1_Finish (solver:570-598): no exclude test
template <class F>
__global__ void k_pcg_init1_finish(F f, TileGrid tg, VecOf<F> r, VecOf<F> pre,
                                   VecOf<F> p, double* aNum) {
    using T = typename F::Scalar;
    constexpr int U = F::U;
    int x, y, z; long s;
    double acc = 0.0;
    OPTK_TILE_LOOP(tg)
    if (thread_site(f, tg, tile_, x, y, z, s)) {
#pragma unroll
        for (int u = 0; u < U; ++u) {
            T res = r(u, s);
            T m = guarded_invert(pre(u, s));
            if (!F::USE_PRE) m = T(1);
            T pp = m * res;
            pre(u, s) = m;
            p(u, s) = pp;
            acc += (double)(res * pp);
        }
    }
    block_accumulate(acc, aNum);
}
// PCGStep1_Graph / computeAdelta_Graph (solver:1216-1244)
template <class F>
__global__ void k_graph_jtj(F f, long E, VecOf<F> dir, VecOf<F> out, double* den) {
    long e = blockIdx.x * (long)blockDim.x + threadIdx.x;
    double acc = 0.0;
    if (e < E) acc = (double)f.graphJTJ(e, dir, out);
    block_accumulate(acc, den);
}
template <class F>
__global__ void k_graph_cost(F f, long E, double* target) {
    using T = typename F::Scalar;
    long e = blockIdx.x * (long)blockDim.x + threadIdx.x;
    double acc = 0.0;
    if (e < E) acc = (double)(T)(0.5 * (double)f.graphCost(e));
    block_accumulate(acc, target);
}
template <class F>
__global__ void k_graph_model_cost(F f, long E, VecOf<F> delta, double* target) {
    using T = typename F::Scalar;
    long e = blockIdx.x * (long)blockDim.x + threadIdx.x;
    double acc = 0.0;
    if (e < E) acc = (double)(T(0.5) * f.graphModelCost(e, delta));
    block_accumulate(acc, target);
}
template <class F>
__global__ void k_graph_ctc(F f, long E, VecOf<F> CtC, typename F::Scalar inv_radius) {
    long e = blockIdx.x * (long)blockDim.x + threadIdx.x;
    if (e < E) f.graphCtC(e, CtC, inv_radius);
}

// ------------------------------------------------------------------------------ graph terms as a segmented scatter
// (functors with HAS_SEGSCATTER: the closed-form hyper-edge energies of functors/graph_generic.cuh)
template <class F>
__global__ void k_segment_keys(F f, long n, unsigned* vertices, unsigned* ids) {
    const long i = blockIdx.x * (long)blockDim.x + threadIdx.x;
    if (i >= n) return;
    vertices[i] = (unsigned)f.edgeVertex(i / F::GV, (int)(i % F::GV));
    ids[i] = (unsigned)i;
}
// PCGStep1_Graph, part 1: one thread per hyper-edge, partials into the edge's own GV * U slots (coalesced rows)
// rank[id] = position of incidence id = edge * GV + slot in the vertex-sorted list (inverse of the sorted id array)
static __global__ void k_segment_rank(long n, const unsigned* __restrict__ sorted_ids, unsigned* __restrict__ rank) {
    const long i = blockIdx.x * (long)blockDim.x + threadIdx.x;
    if (i < n) rank[sorted_ids[i]] = (unsigned)i;
}
// PCGStep1_Graph, part 1: one thread per hyper-edge evaluates the GV * U partials of its edge; the row of incidence (edge, slot)
// is stored AT ITS SORTED POSITION, so that the rows of a vertex are one contiguous range for part 2.  A thread storing its own
// rows would make every store instruction of the warp touch 32 sectors (r2n: embedded, 1 M vertices, 0.69 ms), so the 32 GV
// rows of a warp are staged in shared memory (row pitch GV U + 1: conflict-free) and written out with consecutive lanes on
// consecutive channels of a row (U x 4 contiguous bytes per row).
template <class F>
__global__ void __launch_bounds__(128) k_graph_contrib(F f, long E, VecOf<F> dir, const unsigned* __restrict__ rank,
                                                       typename F::Scalar* __restrict__ contrib) {
    using T = typename F::Scalar;
    constexpr int U = F::U, GV = F::GV, C = GV * U, PITCH = C + 1;
    __shared__ T stage[4][32 * PITCH];                                           // 128 threads: <= 25 KB also for U = 12 in fp64
    __shared__ unsigned pos[4][32 * GV];
    const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
    const long e0 = (blockIdx.x * (long)blockDim.x + threadIdx.x) - lane;       // first edge of this warp
    const long e = e0 + lane;
    if (e0 >= E) return;                                                         // warp-uniform
    if (e < E) {
        T c[C];
        f.graphContrib(e, dir, c);
#pragma unroll
        for (int k = 0; k < C; ++k) stage[warp][lane * PITCH + k] = c[k];
#pragma unroll
        for (int s = 0; s < GV; ++s) pos[warp][lane * GV + s] = rank[e * GV + s];
    }
    __syncwarp();
    const int rows = (int)(E - e0 < 32 ? E - e0 : 32) * GV;                      // (edge, slot) rows of this warp
    for (int i = lane; i < rows * U; i += 32) {
        const int row = i / U, u = i - row * U;
        contrib[(long)pos[warp][row] * U + u] = stage[warp][(row / GV) * PITCH + (row % GV) * U + u];
    }
}
// PCGStep1 + PCGStep1_Graph, part 2: the vertex terms plus the sum of the vertex' partial rows, which lie contiguously at
// [off[v], off[v + 1]) in sorted (= fixed) order; Ap written once, alphaDenominator = <p, Ap> from the same pass
template <class F, bool LM>
__global__ void k_apply_jtj_seg(F f, TileGrid tg, VecOf<F> dir, VecOf<F> out, VecOf<F> CtC, const int* __restrict__ off,
                                const typename F::Scalar* __restrict__ contrib, double* den, double* zero_slot) {
    using T = typename F::Scalar;
    constexpr int U = F::U;
    int x, y, z; long s;
    double acc = 0.0;
    if (zero_slot && blockIdx.x == 0 && linear_tid() == 0) *zero_slot = 0.0;
    OPTK_TILE_LOOP(tg)
    if (thread_site(f, tg, tile_, x, y, z, s) && !f.exclude(x, y, z, s)) {
        T o[U];
        f.applyJTJ(x, y, z, s, dir, o);
        const int b = off[s], e = off[s + 1];
        if constexpr (U % 4 == 0 && sizeof(T) == 4) {                            // rows of whole 128-bit words
            const float4* row = reinterpret_cast<const float4*>(contrib + (long)b * U);
            for (int i = 0; i < (e - b); ++i)
#pragma unroll
                for (int q = 0; q < U / 4; ++q) {
                    const float4 v = row[i * (U / 4) + q];
                    o[4 * q] += v.x; o[4 * q + 1] += v.y; o[4 * q + 2] += v.z; o[4 * q + 3] += v.w;
                }
        } else {
            const T* row = contrib + (long)b * U;
            for (int i = 0; i < (e - b); ++i)
#pragma unroll
                for (int u = 0; u < U; ++u) o[u] += row[i * U + u];
        }
#pragma unroll
        for (int u = 0; u < U; ++u) {
            const T pv = dir(u, s);
            if (LM) o[u] += CtC(u, s) * pv;
            out(u, s) = o[u];
            acc += (double)(pv * o[u]);
        }
    }
    block_accumulate(acc, den);
}

// ------------------------------------------------------------------------------ graph terms as a gather
// Instead of one global atomicAdd per edge and channel (the reference, o.t:2586-2600), J^T J p over the
// graph domain is gathered per vertex: at Init the engine sorts every (edge, slot) incidence by
// (vertex, other endpoint, slot) -- a segmented, deterministic adjacency in which the two directions of a mesh edge
// are neighbours -- and a functor with HAS_GATHER then evaluates, for its own vertex, the contribution of every
// incident edge.  A sorted entry is 2 * payload + slot; k_incidence_merge turns it into 4 * payload + direction
// bits.  No atomics, no memset of Ap, Ap is written once, and
// alphaDen = <p, Ap> comes out of the same kernel.
template <class F>
__global__ void k_incidence_keys(F f, long n, unsigned* entries, unsigned* vertices) {
    static_assert(F::GV <= 2, "entry = 2 * payload + slot");
    const long i = blockIdx.x * (long)blockDim.x + threadIdx.x;
    if (i >= n) return;
    const long e = i / F::GV; const int slot = (int)(i % F::GV);
    entries[i] = f.incidence_payload(e, slot) * 2u + (unsigned)slot;
    vertices[i] = (unsigned)f.graph_vertex(e, slot);
}
// Merge the two directions of a mesh edge: sorted entries 2n, 2n+1 of the same vertex become one entry 4n + 3 (the
// second is marked SKIP, in both arrays, and compacted away); a lone entry 2n + slot becomes 4n + 1 (edge v -> n) or
// 4n + 2 (edge n -> v).
static constexpr unsigned OPTK_INC_SKIP = 0xffffffffu;
struct IncidenceKept { __host__ __device__ bool operator()(unsigned e) const { return e != OPTK_INC_SKIP; } };
static __global__ void k_incidence_merge(const unsigned* __restrict__ entries, const unsigned* __restrict__ vertices, long n,
                                         unsigned* out, unsigned* vout) {
    const long i = blockIdx.x * (long)blockDim.x + threadIdx.x;
    if (i >= n) return;
    const unsigned e = entries[i], v = vertices[i];
    const unsigned nb = e >> 1;
    unsigned o;
    if (e & 1u) o = (i > 0 && vertices[i - 1] == v && entries[i - 1] == e - 1u) ? OPTK_INC_SKIP : nb * 4u + 2u;
    else        o = (i + 1 < n && vertices[i + 1] == v && entries[i + 1] == e + 1u) ? nb * 4u + 3u : nb * 4u + 1u;
    out[i] = o;
    vout[i] = o == OPTK_INC_SKIP ? OPTK_INC_SKIP : v;
}
// off[k] = first position whose key is >= k, for k in [0, nkeys]
static __global__ void k_incidence_offsets(const unsigned* __restrict__ keys, long n, long nkeys, int* off) {
    const long i = blockIdx.x * (long)blockDim.x + threadIdx.x;
    if (i > n) return;
    const long lo = i == 0 ? 0 : (long)keys[i - 1] + 1;
    const long hi = i == n ? nkeys : (long)keys[i];
    for (long k = lo; k <= hi; ++k) off[k] = (int)i;
}
template <class F, bool LM, bool SIMPLE = false>
__global__ void __launch_bounds__(128, sizeof(typename F::Scalar) == 4 ? (SIMPLE ? 8 : 5) : 2) k_gather_jtj(F f, long N, const int* __restrict__ off, const unsigned* __restrict__ inc,
                                                    VecOf<F> dir, VecOf<F> out,
                                                    VecOf<F> CtC, double* den, double* zero_slot) {
    using T = typename F::Scalar;
    constexpr int U = F::U;
    double acc = 0.0;
    if (zero_slot && blockIdx.x == 0 && threadIdx.x == 0) *zero_slot = 0.0;
    for (long v = blockIdx.x * (long)blockDim.x + threadIdx.x; v < N; v += (long)gridDim.x * blockDim.x) {
        T o[U], pv[U];
        f.template gatherJTJ<SIMPLE>(v, off + v, inc, dir, o);
        load_channels<U>(dir, v, pv);
        T dsum = T(0);
#pragma unroll
        for (int u = 0; u < U; ++u) {
            if (LM) o[u] += CtC(u, v) * pv[u];
            dsum += pv[u] * o[u];
        }
        store_channels<U>(out, v, o);
        acc += (double)dsum;
    }
    block_accumulate(acc, den);
}

// The same gather fused with PCGStep3: the direction p' = z + beta p is formed on the fly for the vertex and for every
// neighbour it reads (FusedDir), written once at the thread's own vertex into the other p buffer, and Ap' = J^T J p' is
// produced in the same pass (one kernel and one pass over z, p less per PCG iteration).
template <class F, bool LM>
__global__ void __launch_bounds__(128, sizeof(typename F::Scalar) == 4 ? 5 : 2) k_gather_jtj_fused(F f, long N, const int* __restrict__ off, const unsigned* __restrict__ inc,
                                                    VecOf<F> z, VecOf<F> p_old, VecOf<F> p_new, VecOf<F> out, VecOf<F> CtC,
                                                    const double* bNum, const double* aNum, double* den, double* zero_a, double* zero_b) {
    using T = typename F::Scalar;
    constexpr int U = F::U;
    double acc = 0.0;
    FusedDir<VecOf<F>, T> dir{z, p_old, (T)(*bNum) / (T)(*aNum)};
    if (blockIdx.x == 0 && threadIdx.x == 0) {
        if (zero_a) *zero_a = 0.0;
        if (zero_b) *zero_b = 0.0;
    }
    for (long v = blockIdx.x * (long)blockDim.x + threadIdx.x; v < N; v += (long)gridDim.x * blockDim.x) {
        T o[U], pv[U];
        f.gatherJTJ(v, off + v, inc, dir, o);
        load_channels<U>(dir, v, pv);
        store_channels<U>(p_new, v, pv);
        T dsum = T(0);
#pragma unroll
        for (int u = 0; u < U; ++u) {
            if (LM) o[u] += CtC(u, v) * pv[u];
            dsum += pv[u] * o[u];
        }
        store_channels<U>(out, v, o);
        acc += (double)dsum;
    }
    block_accumulate(acc, den);
}

// ------------------------------------------------------------------------------ flat streaming kernels
// PCGStep2 (solver:665-731): alpha = aNum/aDen; delta += alpha p; r -= alpha Ap; z = M^-1 r;
// betaNum += z.r ; LM: q += 0.5 delta.(r+b).   One 128-bit vector per thread and array.
// Exclusion-aware streaming: `active4[g]` is nonzero iff any of the sites 4g..4g+3 is an unknown
// (not excluded).  Groups without unknowns hold zeros in every solver vector and are skipped, so a
// masked-out half of an image costs no HBM traffic (N/4 bytes of flags per plane pass instead).
template <class T>
__device__ __forceinline__ bool group_active(const unsigned char* active4, long i, long plane) {
    if (!active4) return true;
    const long e = (i * VecIO<T>::W) % plane;
    return active4[e >> 2] != 0;
}
// `changed` is raised when the Exclude mask differs from the one the previous PCGInit1 saw (a caller that re-uses a plan with
// another mask image): sites that were unknowns and are excluded now still hold their old delta / r / p / z / Ap, which the
// streaming kernels, the stencils and the U = 1 flat updates assume to be 0 -- k_zero_if then clears the solver vectors.
template <class F>
__global__ void k_active4(F f, long ngroups, unsigned char* active4, int* changed) {
    const long g = blockIdx.x * (long)blockDim.x + threadIdx.x;
    if (g >= ngroups) return;
    const long N = (long)f.W * f.H * f.D;
    unsigned char a = 0;
    for (int j = 0; j < 4; ++j) {
        const long s = g * 4 + j;
        if (s < N) {
            const int x = (int)(s % f.W), y = (int)((s / f.W) % f.H), z = (int)(s / ((long)f.W * f.H));
            if (!f.exclude(x, y, z, s)) a |= (unsigned char)(1u << j);     // one bit per site of the group
        }
    }
    if (active4[g] != a) { active4[g] = a; *changed = 1; }
}
template <class V4>
__global__ void __launch_bounds__(256) k_zero_if(const int* __restrict__ flag, V4* __restrict__ v, long n) {
    if (*flag == 0) return;
    V4 zero{};
    for (long i = blockIdx.x * (long)blockDim.x + threadIdx.x; i < n; i += (long)gridDim.x * blockDim.x) v[i] = zero;
}

// LM inner-loop exit test (solver:2180-2200) evaluated on the device.  One thread: zeta = (l+1)(Q1-Q0)/Q1 in
// opt_float (the same roundings as the reference's host code), sets the stop flag the step kernels of a
// speculatively queued iteration look at, rolls Q0 forward and posts the result into a pinned host mailbox
// (payload, system-scope fence, sequence number) that the host polls without any stream synchronisation.
struct LmMailbox { double zeta, q; long long stop; long long seq; };
template <class T>
__global__ void k_lm_begin(const double* q, double* qprev) { *qprev = (double)(T)(*q); }
template <class T>
__device__ __forceinline__ void lm_zeta_body(int l, double q_tol, double q_now, double* qprev, double* stop, LmMailbox* box, long long seq) {
    const double Q1 = (double)(T)q_now, Q0 = *qprev;
    const double zeta = (double)(T)((double)(T)((double)(T)(l + 1) * (double)(T)(Q1 - Q0)) / Q1);
    const bool was = *stop != 0.0;
    const bool now = was || zeta < q_tol;
    if (!was) { *qprev = Q1; if (now) *stop = 1.0; }
    volatile LmMailbox* b = box;
    b->zeta = zeta; b->q = Q1; b->stop = now ? 1 : 0;
    __threadfence_system();
    b->seq = seq;
}
template <class T>
__global__ void k_lm_zeta(int l, double q_tol, const double* q, double* qprev, double* stop, LmMailbox* box, long long seq) {
    lm_zeta_body<T>(l, q_tol, *q, qprev, stop, box, seq);
}
// The exit test folded into the tail of PCGStep2 (single GPU, no residual reset): the last block to retire -- arrival counter
// behind a device-scope fence, so every block's contribution to q is visible to it -- evaluates the test.  One launch less per
// LM iteration (the separate one-thread kernel cost 8.5 us of a 80 us shape_from_shading iteration, profiles/r1k_timing_sfs_lm.txt).
struct LmFold { int on; int l; double q_tol; double* qprev; double* stop_w; LmMailbox* box; long long seq; unsigned* counter; };


// DM (Gauss-Newton with the fused PCGStep3+1 kernel only): delta is not read inside the inner loop, and the fused kernel's
// ping-pong keeps p_l and p_{l-1} both alive, so the update delta += alpha p is applied every SECOND iteration for two
// directions at once.  DM = 0: delta += alpha p (every iteration);  DM = 1: delta untouched, alpha parked in *aprev;
// DM = 2: delta = fma(alpha, p, fma(*aprev, p_prev, delta)) -- the same two roundings per element as two DM = 0 passes, so the
// result is bit-identical -- and 14 instead of 16 vector passes over two iterations (10 instead of 12 without preconditioner).
// This is the dominant kernel of the big grids by time (512^3: 51.6 % of a solve, profiles/r2m_launches_volumetric512.csv)
// and runs at the copy peak, so fewer bytes is the only way to make it faster.
template <class T, bool USE_PRE, bool LM, int DM = 0>
__global__ void __launch_bounds__(256) k_pcg_step2(long i_begin, long nvec, const unsigned char* __restrict__ active4, long plane,
                                                   T* __restrict__ delta, const T* __restrict__ p, const T* __restrict__ p_prev, double* aprev,
                                                   T* __restrict__ r, const T* __restrict__ Ap,
                                                   const T* __restrict__ pre, const T* __restrict__ b,
                                                   T* __restrict__ z, const double* aNum, const double* aDen,
                                                   double* bNum, double* q, double* zero_slot, const double* stop, LmFold fold) {
    using V = typename VecIO<T>::V;
    constexpr int VW = VecIO<T>::W;
    const long i0 = blockIdx.x * (long)blockDim.x + threadIdx.x, istride = (long)gridDim.x * blockDim.x;
    // LM: once the q_tolerance test has fired, an iteration that was queued ahead of the host's
    // decision must leave delta untouched (pcg_run_lm); nothing else of it is read afterwards
    const T alpha = (LM && stop && *stop != 0.0) ? T(0) : (T)(*aNum) / (T)(*aDen);
    const T alpha_prev = DM == 2 ? (T)(*aprev) : T(0);           // written by the previous launch of this kernel (DM = 1)
    if (i0 == 0 && zero_slot) *zero_slot = 0.0;
    if (DM == 1 && i0 == 0) *aprev = (double)alpha;
    double bacc = 0.0, qacc = 0.0;
    for (long i = i_begin + i0; i < nvec; i += istride)
    if (group_active<T>(active4, i, plane)) {
        V d, pv, pq;
        if (DM != 1) { d = reinterpret_cast<const V*>(delta)[i]; pv = reinterpret_cast<const V*>(p)[i]; }
        if (DM == 2) pq = reinterpret_cast<const V*>(p_prev)[i];
        V rv = reinterpret_cast<const V*>(r)[i];
        V av = reinterpret_cast<const V*>(Ap)[i];
        V mv, bv, zv;
        if (USE_PRE) mv = reinterpret_cast<const V*>(pre)[i];
        if (LM) bv = reinterpret_cast<const V*>(b)[i];
        T bsum = T(0), qsum = T(0);
#pragma unroll
        for (int k = 0; k < VW; ++k) {
            T dl = T(0);
            if (DM == 2) dl = comp(d, k) + alpha_prev * comp(pq, k);
            if (DM == 0) dl = comp(d, k);
            if (DM != 1) { dl = dl + alpha * comp(pv, k); comp(d, k) = dl; }
            T rr = comp(rv, k) - alpha * comp(av, k);
            comp(rv, k) = rr;
            T zz = USE_PRE ? comp(mv, k) * rr : rr;
            comp(zv, k) = zz;
            bsum += zz * rr;
            if (LM) qsum += dl * (rr + comp(bv, k));
        }
        if (DM != 1) reinterpret_cast<V*>(delta)[i] = d;
        reinterpret_cast<V*>(r)[i] = rv;
        if (USE_PRE) reinterpret_cast<V*>(z)[i] = zv;
        bacc += (double)bsum;
        qacc += (double)(T(0.5) * qsum);
    }
    block_accumulate(bacc, bNum);
    if (LM) block_accumulate(qacc, q);
    if (LM && fold.on && threadIdx.x == 0) {                  // thread 0 issued this block's atomics just above
        __threadfence();
        if (atomicAdd(fold.counter, 1u) == gridDim.x - 1) {
            *fold.counter = 0;                                // next launch
            __threadfence();
            lm_zeta_body<T>(fold.l, fold.q_tol, *(volatile double*)q, fold.qprev, fold.stop_w, fold.box, fold.seq);
        }
    }
}

// PCGStep3 (solver:814-845): beta = bNum/aNum; p = z + beta p
template <class T>
__global__ void __launch_bounds__(256) k_pcg_step3(long i_begin, long nvec, const unsigned char* __restrict__ active4, long plane,
                                                   const T* __restrict__ z, T* __restrict__ p,
                                                   const double* bNum, const double* aNum, double* zero_slot) {
    using V = typename VecIO<T>::V;
    constexpr int VW = VecIO<T>::W;
    const long i0 = blockIdx.x * (long)blockDim.x + threadIdx.x, istride = (long)gridDim.x * blockDim.x;
    const T beta = (T)(*bNum) / (T)(*aNum);
    if (i0 == 0 && zero_slot) *zero_slot = 0.0;
    for (long i = i_begin + i0; i < nvec; i += istride)
    if (group_active<T>(active4, i, plane)) {
        V zv = reinterpret_cast<const V*>(z)[i];
        V pv = reinterpret_cast<const V*>(p)[i];
#pragma unroll
        for (int k = 0; k < VW; ++k) comp(pv, k) = comp(zv, k) + beta * comp(pv, k);
        reinterpret_cast<V*>(p)[i] = pv;
    }
}

// Slab mode with the fused PCGStep3+PCGStep1 kernel: the new search direction p' = z + beta p of the owned rows next to a
// neighbouring slab, computed ahead of the fused kernel so that it can be sent into the neighbour's ghost rows of p'
// (the fused kernel then takes its ghost rows from p' instead of forming z + beta p there).  SoA vectors only.
template <class T>
__global__ void __launch_bounds__(256) k_boundary_dir(const T* __restrict__ z, const T* __restrict__ p, T* __restrict__ p_new, long plane, int U,
                                                      long rowsz, int lo, int hi, int halo, int has_lo, int has_hi,
                                                      const double* bNum, const double* aNum) {
    const T beta = (T)(*bNum) / (T)(*aNum);
    const long per_side = rowsz * halo, total = per_side * 2 * U;
    for (long i = blockIdx.x * (long)blockDim.x + threadIdx.x; i < total; i += (long)gridDim.x * blockDim.x) {
        const int u = (int)(i / (2 * per_side));
        const long r = i % (2 * per_side);
        const bool upper = r >= per_side;
        if (upper ? !has_hi : !has_lo) continue;
        const long e = (long)u * plane + (upper ? (long)(hi - halo) * rowsz + (r - per_side) : (long)lo * rowsz + r);
        p_new[e] = z[e] + beta * p[e];
    }
}

// PCGStep2_1stHalf (solver:738-762): delta += alpha p
template <class T>
__global__ void __launch_bounds__(256) k_axpy_alpha(long i_begin, long nvec, const unsigned char* __restrict__ active4, long plane,
                                                    T* __restrict__ delta, const T* __restrict__ p,
                                                    const double* aNum, const double* aDen, const double* stop) {
    using V = typename VecIO<T>::V;
    constexpr int VW = VecIO<T>::W;
    const long i0 = blockIdx.x * (long)blockDim.x + threadIdx.x, istride = (long)gridDim.x * blockDim.x;
    const T alpha = (stop && *stop != 0.0) ? T(0) : (T)(*aNum) / (T)(*aDen);
    for (long i = i_begin + i0; i < nvec; i += istride)
    if (group_active<T>(active4, i, plane)) {
        V d = reinterpret_cast<const V*>(delta)[i];
        V pv = reinterpret_cast<const V*>(p)[i];
#pragma unroll
        for (int k = 0; k < VW; ++k) comp(d, k) = comp(d, k) + alpha * comp(pv, k);
        reinterpret_cast<V*>(delta)[i] = d;
    }
}

// PCGStep2_2ndHalf (solver:766-799): r = b - A delta; z = M^-1 r; betaNum += z.r; q += 0.5 delta.(r+b)
template <class T, bool USE_PRE>
__global__ void __launch_bounds__(256) k_residual_reset(long i_begin, long nvec, const unsigned char* __restrict__ active4, long plane,
                                                        const T* __restrict__ delta, T* __restrict__ r,
                                                        const T* __restrict__ Adelta, const T* __restrict__ pre,
                                                        const T* __restrict__ b, T* __restrict__ z,
                                                        double* bNum, double* q, double* zero_slot) {
    using V = typename VecIO<T>::V;
    constexpr int VW = VecIO<T>::W;
    const long i0 = blockIdx.x * (long)blockDim.x + threadIdx.x, istride = (long)gridDim.x * blockDim.x;
    if (i0 == 0 && zero_slot) *zero_slot = 0.0;
    double bacc = 0.0, qacc = 0.0;
    for (long i = i_begin + i0; i < nvec; i += istride)
    if (group_active<T>(active4, i, plane)) {
        V d = reinterpret_cast<const V*>(delta)[i];
        V av = reinterpret_cast<const V*>(Adelta)[i];
        V bv = reinterpret_cast<const V*>(b)[i];
        V mv, rv, zv;
        if (USE_PRE) mv = reinterpret_cast<const V*>(pre)[i];
        T bsum = T(0), qsum = T(0);
#pragma unroll
        for (int k = 0; k < VW; ++k) {
            T rr = comp(bv, k) - comp(av, k);
            comp(rv, k) = rr;
            T zz = USE_PRE ? comp(mv, k) * rr : rr;
            comp(zv, k) = zz;
            bsum += zz * rr;
            qsum += comp(d, k) * (rr + comp(bv, k));
        }
        reinterpret_cast<V*>(r)[i] = rv;
        if (USE_PRE) reinterpret_cast<V*>(z)[i] = zv;
        bacc += (double)bsum;
        qacc += (double)(T(0.5) * qsum);
    }
    block_accumulate(bacc, bNum);
    block_accumulate(qacc, q);
}

// Scalar read-back without the copy engine: one thread stores the device scalar (and the peer-spin error flag) straight into
// pinned, device-mapped host memory.  A cudaMemcpyAsync D2H of 8 bytes queues behind whatever else the D2H copy engine is doing:
// a caller that streams problems through the API has the previous result's download in flight (bench.py's e2e leg: 268 MB / 3.2 GB),
// and every cost read-back of the solve stalled for one such download (r2s: +6.9 ms on a 24.9 ms poisson solve, +68 ms on a 951 ms
// volumetric one).
static __global__ void k_publish_scalar(const double* __restrict__ src, const int* __restrict__ err, volatile double* host_val, volatile int* host_err) {
    *host_val = *src;
    *host_err = err ? *err : 0;
    __threadfence_system();
}

// ---- scalar all-reduce over NVLink peer memory ------------------------------------------------------------
// The PCG loop needs two dependent scalar all-reduces per iteration (alphaDen after PCGStep1, betaNum after PCGStep2).
// ncclAllReduce of 8 bytes costs a full collective launch each (~10-15 us); here every rank stores its partial sums
// straight into a mailbox in EVERY peer's memory (cudaIpc-mapped at OptB200_PlanSetSlab; NVSwitch gives every pair a
// direct path), fences, publishes an epoch number, and then waits for the N epochs in its own mailbox: one tiny
// kernel, one NVLink store latency.  Sums are formed in rank order, so every rank gets the bit-identical result.
// Two mailbox sets alternate by epoch parity: a rank cannot run two epochs ahead of a peer that has not finished
// reading, because finishing epoch k+1 needs that peer's epoch-k+1 contribution.
constexpr int XR_MAX_RANKS = 16, XR_MAX_VALS = 4;
struct XSlot { double v[XR_MAX_VALS]; long long epoch; long long pad[3]; };
struct XPeers { XSlot* p[XR_MAX_RANKS]; };
static __global__ void k_xreduce(double* vals, int count, XPeers peers, int rank, int nranks, long long epoch, int* error_flag) {
    // one warp, lane t talks to rank t: stores and polls of the N peers proceed in parallel
    const int t = threadIdx.x;
    const int par = (int)(epoch & 1);
    __shared__ double got[XR_MAX_RANKS][XR_MAX_VALS];
    if (t < nranks) {
        volatile XSlot* dst = peers.p[t] + par * XR_MAX_RANKS + rank;
        for (int c = 0; c < count; ++c) dst->v[c] = vals[c];
        __threadfence_system();
        dst->epoch = epoch;
        volatile XSlot* src = peers.p[rank] + par * XR_MAX_RANKS + t;
        long long spins = 0;
        bool dead = false;
        while (src->epoch != epoch) {
            if (++spins > (1LL << 28)) { dead = true; break; }             // minutes: a peer died; give up instead of hanging the GPU
        }
        __threadfence_system();
        if (dead) *error_flag = 1;
        for (int c = 0; c < count; ++c) got[t][c] = dead ? 0.0 : src->v[c];
    }
    __syncwarp();
    if (t < count) {                                                        // rank order: identical bits on every rank
        double s = 0.0;
        for (int r2 = 0; r2 < nranks; ++r2) s += got[r2][t];
        vals[t] = s;
    }
}

// Halo exchange of the fused slab path through peer memory: the new search direction of the owned boundary rows is
// computed and stored STRAIGHT into the neighbours' ghost rows (their solver-vector pools are cudaIpc-mapped); the
// last block to finish publishes an epoch in both neighbours' mailboxes and waits for theirs (system-scope fences order
// every block's data stores before the flag store); the fused kernel is queued behind this one.
template <class T>
struct HaloPeer { T* base; long plane; long row; };          // neighbour's p' vector, its plane stride, first ghost row to fill
__device__ __forceinline__ void halo_signal_and_wait(const XPeers& peers, int rank, int nranks, long long epoch, int* error_flag) {
    // mailbox words: slot [2*XR_MAX_RANKS + 1].v[0..1] reinterpreted as two epochs: [0] written by rank-1, [1] by rank+1
    __threadfence_system();
    if (rank > 0) ((volatile long long*)peers.p[rank - 1][2 * XR_MAX_RANKS + 1].v)[1] = epoch;
    if (rank < nranks - 1) ((volatile long long*)peers.p[rank + 1][2 * XR_MAX_RANKS + 1].v)[0] = epoch;
    volatile long long* mine = (volatile long long*)peers.p[rank][2 * XR_MAX_RANKS + 1].v;
    for (int side = 0; side < 2; ++side) {
        if (side == 0 ? rank == 0 : rank == nranks - 1) continue;
        long long spins = 0;
        while (mine[side] < epoch) { if (++spins > (1LL << 28)) { *error_flag = 1; return; } }
    }
    __threadfence_system();
}
// One launch: every block stores its share of the boundary rows into the neighbours' ghost rows; the block that
// finishes last (device-wide counter) publishes the epoch and waits for the neighbours' epochs.
template <class T>
__global__ void __launch_bounds__(256) k_boundary_push(const T* __restrict__ z, const T* __restrict__ p, long plane, int U, long rowsz,
                                                       int lo, int hi, int halo, HaloPeer<T> below, HaloPeer<T> above,
                                                       const double* bNum, const double* aNum,
                                                       XPeers peers, int rank, int nranks, long long epoch, int* error_flag, unsigned* done_counter) {
    const T beta = (T)(*bNum) / (T)(*aNum);
    const long per_side = rowsz * halo, total = per_side * 2 * U;
    for (long i = blockIdx.x * (long)blockDim.x + threadIdx.x; i < total; i += (long)gridDim.x * blockDim.x) {
        const int u = (int)(i / (2 * per_side));
        const long r = i % (2 * per_side);
        const bool upper = r >= per_side;
        const HaloPeer<T>& nb = upper ? above : below;
        if (!nb.base) continue;
        const long off = upper ? r - per_side : r;
        const long e = (long)u * plane + (upper ? (long)(hi - halo) * rowsz : (long)lo * rowsz) + off;
        nb.base[(long)u * nb.plane + nb.row * rowsz + off] = z[e] + beta * p[e];
    }
    __threadfence_system();                                  // this thread's remote stores before the block's arrival
    __syncthreads();
    if (threadIdx.x == 0) {
        const unsigned arrived = atomicAdd(done_counter, 1u);
        if (arrived == gridDim.x - 1) {
            *done_counter = 0;                               // next launch
            halo_signal_and_wait(peers, rank, nranks, epoch, error_flag);
        }
    }
}

// Halo exchange of a blocked solver vector (one contiguous slice per face) through peer memory: the owned boundary
// slices are copied straight into the neighbours' ghost slices, then the same last-block handshake as k_boundary_push.
template <class V4>
__global__ void __launch_bounds__(256) k_halo_push(const V4* __restrict__ src, long cnt, long src_lo, long src_hi, V4* dst_below, V4* dst_above,
                                                   XPeers peers, int rank, int nranks, long long epoch, int* error_flag, unsigned* done_counter) {
    for (long i = blockIdx.x * (long)blockDim.x + threadIdx.x; i < 2 * cnt; i += (long)gridDim.x * blockDim.x) {
        const bool upper = i >= cnt;
        V4* dst = upper ? dst_above : dst_below;
        if (!dst) continue;
        const long k = upper ? i - cnt : i;
        dst[k] = src[(upper ? src_hi : src_lo) + k];
    }
    __threadfence_system();
    __syncthreads();
    if (threadIdx.x == 0) {
        const unsigned arrived = atomicAdd(done_counter, 1u);
        if (arrived == gridDim.x - 1) {
            *done_counter = 0;
            halo_signal_and_wait(peers, rank, nranks, epoch, error_flag);
        }
    }
}

// ------------------------------------------------------------------------------ materialized Jacobian
// saveJToCRS / saveJToCRS_Graph (solver:1004-1022, 1287-1307, generateDumpJ :385-442): every site / edge writes its residual
// rows through a RowSink (csr.cuh).  COUNT pass: row lengths; fill pass: columns and values at the scanned row offsets.
template <class F, bool COUNT>
__global__ void k_save_j(F f, TileGrid tg, int* cnt_or_ptr, int* colInd, typename F::Scalar* val, long plane) {
    int x, y, z; long s;
    f.o_lo = 0; f.o_hi = 0x7fffffff;
    OPTK_TILE_LOOP(tg)
    if (thread_site(f, tg, tile_, x, y, z, s)) {
        RowSink<typename F::Scalar, VecOf<F>, COUNT> sink{cnt_or_ptr, colInd, val, s * F::MAT_ROWS, plane};
        f.rowsJ(x, y, z, s, sink);
        sink.close();
    }
}
template <class F, bool COUNT>
__global__ void k_save_j_graph(F f, long E, long row_base, int* cnt_or_ptr, int* colInd, typename F::Scalar* val, long plane) {
    const long e = blockIdx.x * (long)blockDim.x + threadIdx.x;
    if (e >= E) return;
    RowSink<typename F::Scalar, VecOf<F>, COUNT> sink{cnt_or_ptr, colInd, val, row_base + e * F::MAT_GROWS, plane};
    f.graphRowsJ(e, sink);
    sink.close();
}
// PCGStep1_Finish (solver:646-660): alphaDenominator = <p, Ap> over the non-excluded unknowns after the sparse products.  The
// products also fill the entries of Ap that belong to excluded unknowns (their columns exist in J); those are cleared here,
// because the streaming kernels rely on excluded sites holding 0.  LM: + CtC p, which the matrix-free PCGStep1 adds
// (o.t:2817-2823) -- the reference's materialized path leaves it out; here the three modes give the same iterates.
template <class F, bool LM>
__global__ void k_pcg_step1_finish(F f, TileGrid tg, VecOf<F> dir, VecOf<F> out, VecOf<F> CtC, double* den, double* zero_slot) {
    using T = typename F::Scalar;
    constexpr int U = F::U;
    int x, y, z; long s;
    double acc = 0.0;
    if (zero_slot && blockIdx.x == 0 && linear_tid() == 0) *zero_slot = 0.0;
    OPTK_TILE_LOOP(tg)
    if (thread_site(f, tg, tile_, x, y, z, s)) {
        const bool ex = f.exclude(x, y, z, s);
#pragma unroll
        for (int u = 0; u < U; ++u) {
            if (ex) { out(u, s) = T(0); continue; }
            const T pv = dir(u, s);
            T o = out(u, s);
            if (LM) { o += CtC(u, s) * pv; out(u, s) = o; }
            acc += (double)(pv * o);
        }
    }
    block_accumulate(acc, den);
}

// Fused slab path of a blocked solver vector: the new search direction p' = z + beta p of the two owned boundary slices
// (the PCGStep3 of those planes, ahead of the fused PCGStep3+1 kernel) is stored straight into the neighbours' ghost
// slices of THEIR p' buffer; the fused kernel then stages ghost planes from p' (volumetric_packed.cuh).  `local_out`
// (NCCL fallback): p' goes into this rank's own p' buffer and is exchanged with send/recv afterwards.
template <class T>
__global__ void __launch_bounds__(256) k_halo_push_dir(const typename VecIO<T>::V* __restrict__ z, const typename VecIO<T>::V* __restrict__ p,
                                                       typename VecIO<T>::V* local_out, long cnt, long src_lo, long src_hi, int has_lo, int has_hi,
                                                       typename VecIO<T>::V* dst_below, typename VecIO<T>::V* dst_above,
                                                       const double* bNum, const double* aNum,
                                                       XPeers peers, int rank, int nranks, long long epoch, int* error_flag, unsigned* done_counter, int handshake) {
    using V4 = typename VecIO<T>::V;
    const T beta = (T)(*bNum) / (T)(*aNum);
    for (long i = blockIdx.x * (long)blockDim.x + threadIdx.x; i < 2 * cnt; i += (long)gridDim.x * blockDim.x) {
        const bool upper = i >= cnt;
        if (upper ? !has_hi : !has_lo) continue;
        const long k = upper ? i - cnt : i;
        const long e = (upper ? src_hi : src_lo) + k;
        V4 zv = z[e], pv = p[e];
#pragma unroll
        for (int c = 0; c < VecIO<T>::W; ++c) comp(zv, c) = fma(beta, comp(pv, c), comp(zv, c));   // the fused kernel's rounding
        if (local_out) local_out[e] = zv;
        V4* dst = upper ? dst_above : dst_below;
        if (dst) dst[k] = zv;
    }
    if (!handshake) return;
    __threadfence_system();
    __syncthreads();
    if (threadIdx.x == 0) {
        const unsigned arrived = atomicAdd(done_counter, 1u);
        if (arrived == gridDim.x - 1) {
            *done_counter = 0;
            halo_signal_and_wait(peers, rank, nranks, epoch, error_flag);
        }
    }
}

// ------------------------------------------------------------------------------ the engine
template <class F>
struct Engine : OptK_Plan {
    using T = typename F::Scalar;
    using V = VecOf<F>;
    static constexpr int U = F::U;
    static constexpr bool GRAPH = F::GV > 0;

    F f;
    bool lm;
    long N, Npad, E;
    T* pool = nullptr;
    V delta, r, b, Adelta, z, p, Ap, CtC, pre, SSq, prevX;
    V p_alt;                                           // second p buffer of the fused Step3+Step1 kernel
    unsigned char* active4 = nullptr;                       // 1 flag per 4 sites (HAS_EXCLUDE functors only)
    int* mask_changed = nullptr;                            // raised by k_active4 when the Exclude mask is not the previous one
    unsigned* inc = nullptr;                                // HAS_GATHER: incidence payloads sorted by (vertex, slot)
    int* inc_off = nullptr;                                 //             N*GV + 1 segment offsets
    void* sort_tmp = nullptr; size_t sort_tmp_bytes = 0;
    int* inc_count = nullptr;                               //             entries kept after merging edge directions
    unsigned *key_a = nullptr, *key_b = nullptr, *val_a = nullptr;
    bool fuse = true;                                       // fused PCGStep3+PCGStep1 (single rank)
    double* scal = nullptr;
    double* h_scal = nullptr;    // pinned + device-mapped
    double* h_scal_dev = nullptr; // its device address (read_scalar publishes through it); nullptr: cudaMemcpyAsync instead
    LmMailbox* lm_box = nullptr; // pinned + device-mapped, ring of 2 (pcg_run_lm)
    long long lm_seq = 0;
    T radius = 0, min_diag = 0, max_diag = 0;
    dim3 sgrid, sblock;
    TileGrid tg;
    int nsm = 148;
    // slab decomposition along the slowest dimension; nranks == 1: single GPU
    int rank = 0, nranks = 1, halo = 0;
    ncclComm_t comm = nullptr;
    // scalar all-reduce through peer memory (k_xreduce); xr_on == false: ncclAllReduce
    bool xr_on = false;
    XSlot* xr_buf = nullptr; XPeers xr_peers{}; long long xr_epoch = 0; int* xr_error = nullptr;
    std::vector<void*> xr_opened;
    // halo exchange through peer memory (fused slab path): neighbours' pools and geometry
    bool xh_on = false; long long xh_epoch = 0;
    T* nb_pool[2] = {nullptr, nullptr}; long nb_npad[2] = {0, 0}; long nb_row[2] = {0, 0};

    explicit Engine(bool is_lm) : lm(is_lm) {}
    // All allocation happens here, not in the constructor: if it fails part-way with an error handler installed (optk_fatal
    // throws), make_engine deletes the object and the destructor releases what had been allocated.
    void init(const unsigned* dims, bool collect_timing) {
        base_init(collect_timing);
        f.plan(dims);
        N = (long)f.W * f.H * f.D;
        E = GRAPH ? (long)f.E : 0;
        Npad = round_up(N, 64);
        const int nvec = lm ? 12 : 8;
        size_t bytes = (size_t)nvec * U * Npad * sizeof(T);
        OPTK_CUDA(cudaMalloc(&pool, bytes));
        OPTK_CUDA(cudaMemsetAsync(pool, 0, bytes, stream));
        V* vs[12] = {&delta, &r, &z, &p, &Ap, &pre, &prevX, &p_alt, &b, &Adelta, &CtC, &SSq};
        for (int i = 0; i < 12; ++i) {
            vs[i]->plane = Npad;
            vs[i]->d = i < nvec ? pool + (size_t)i * U * Npad : nullptr;
        }
        if (!F::USE_PRE) z = r;                             // M^-1 = 1: z aliases r (solver:705-711)
        if (F::HAS_EXCLUDE) {
            OPTK_CUDA(cudaMalloc(&active4, Npad / 4));
            OPTK_CUDA(cudaMemsetAsync(active4, 0, Npad / 4, stream));
            OPTK_CUDA(cudaMalloc(&mask_changed, sizeof(int)));
        }
        // the gather kernel is issue-bound: forming z + beta p for every neighbour costs what the separate PCGStep3 pass costs
        // (arap 1 M vertices: 0.0683 ms fused vs 0.0535 + 0.0148 ms; r1s), so it stays unfused unless OPTB200_GATHER_FUSED is set
        if (getenv("OPTB200_NO_FUSE") || !F::FUSE_STEP3 || (use_gather() && !getenv("OPTB200_GATHER_FUSED")) || use_seg()) fuse = false;
        OPTK_CUDA(cudaMalloc(&scal, S_COUNT * sizeof(double)));
        OPTK_CUDA(cudaMemsetAsync(scal, 0, S_COUNT * sizeof(double), stream));
        OPTK_CUDA(cudaHostAlloc(&h_scal, S_COUNT * sizeof(double), cudaHostAllocMapped | cudaHostAllocPortable));
        memset(h_scal, 0, S_COUNT * sizeof(double));
        if (getenv("OPTB200_SCALAR_MEMCPY") == nullptr) {
            void* dp = nullptr;
            if (cudaHostGetDevicePointer(&dp, h_scal, 0) == cudaSuccess) h_scal_dev = (double*)dp; else cudaGetLastError();
        }
        if (lm) {
            OPTK_CUDA(cudaHostAlloc(&lm_box, 2 * sizeof(LmMailbox), cudaHostAllocMapped | cudaHostAllocPortable));
            memset(lm_box, 0, 2 * sizeof(LmMailbox));
            OPTK_CUDA(cudaMalloc(&lm_counter, sizeof(unsigned)));
            OPTK_CUDA(cudaMemsetAsync(lm_counter, 0, sizeof(unsigned), stream));
        }
        if (F::ND == 1) { sblock = dim3(256, 1, 1); }
        else if (F::ND == 2) { sblock = dim3(32, 8, 1); }
        else { sblock = dim3(32, 4, 2); }
        tg.gx = (f.W + sblock.x - 1) / sblock.x; tg.gy = (f.H + sblock.y - 1) / sblock.y; tg.gz = (f.D + sblock.z - 1) / sblock.z;
        tg.n = (long)tg.gx * tg.gy * tg.gz;
        { int dev = 0; cudaGetDevice(&dev); cudaDeviceGetAttribute(&nsm, cudaDevAttrMultiProcessorCount, dev); }
        sgrid = dim3((unsigned)std::min<long>(tg.n, (long)nsm * 8));        // persistent: <= 8 CTAs of 256 threads per SM
        f.alloc_temporaries(stream);
        OPTK_CUDA(cudaStreamSynchronize(stream));
    }
    ~Engine() override {
        cudaStreamSynchronize(stream);
        f.free_temporaries();
        for (void* q : xr_opened) cudaIpcCloseMemHandle(q);
        cudaFree(xr_buf); cudaFree(xr_error);
        if (comm) ncclCommDestroy(comm);
        release_materialized(); m_tmp.release();
        cudaFree(pool); cudaFree(scal); cudaFreeHost(h_scal); cudaFreeHost(lm_box); cudaFree(lm_counter); cudaFree(active4); cudaFree(mask_changed);
        cudaFree(seg_ids); cudaFree(seg_rank); cudaFree(seg_off); cudaFree(seg_contrib);
        cudaFree(inc); cudaFree(inc_off); cudaFree(inc_count); cudaFree(sort_tmp); cudaFree(key_a); cudaFree(key_b); cudaFree(val_a);
        base_destroy();
    }
    // ---- materialized Jacobian: J (and J^T, J^T J) as CSR over the solver-vector index space -----------------------
    int mat_mode = 0;
    CsrMat<T> mJ, mJT, mJTJ;
    int *m_perm = nullptr, *m_rowOf = nullptr;
    T* m_Jp = nullptr;
    Scratch m_tmp;
    void release_materialized() {
        mJ.release(); mJT.release(); mJTJ.release();
        cudaFree(m_perm); cudaFree(m_rowOf); cudaFree(m_Jp); m_perm = m_rowOf = nullptr; m_Jp = nullptr;
    }
    int set_materialized(int mode) override {
        if constexpr (!F::HAS_ROWS_J) {
            if (mode) fprintf(stderr, "Opt: this energy has no Jacobian-row emitter; useMaterializedJTJ is served by the matrix-free kernels (same iterates)\n");
            return 0;
        } else {
            if (mode && (nranks > 1 || F::VEC_UB > 0)) { fprintf(stderr, "Opt: materialized Jacobian is single-GPU / SoA only; using the matrix-free kernels\n"); return 0; }
            mat_mode = mode;
            if (mode) fuse = false;                          // PCGStep3 stays its own kernel, as in the reference's materialized path
            return mode;
        }
    }
    long materialized_nnz(int which) const override { return which == 0 ? mJ.nnz : mJTJ.nnz; }
    // cusparseOuter (solver:1532-1657).  Pattern (row offsets, columns, transpose permutation, J^T J pattern) once per solve:
    // the graph index arrays belong to the caller; values at every nonlinear iteration.
    void materialize(int nIter) {
        if constexpr (F::HAS_ROWS_J) {
            const long rows = N * F::MAT_ROWS + E * F::MAT_GROWS;
            if (nIter == 0 || !mJ.rowPtr) {
                release_materialized();
                mJ.nrows = (int)rows; mJ.ncols = (int)((long)U * Npad);
                int* cnt = nullptr;
                OPTK_CUDA(cudaMalloc(&cnt, (size_t)(rows + 1) * sizeof(int)));
                OPTK_CUDA(cudaMalloc(&mJ.rowPtr, (size_t)(rows + 1) * sizeof(int)));
                OPTK_CUDA(cudaMemsetAsync(cnt, 0, (size_t)(rows + 1) * sizeof(int), stream));
                if constexpr (F::MAT_ROWS > 0) this->launch_tiles("saveJToCRS_count", tg.n, sblock, k_save_j<F, true>, f, tg, cnt, (int*)nullptr, (T*)nullptr, Npad);
                if constexpr (F::MAT_GROWS > 0) if (E > 0) this->launch("saveJToCRS_Graph_count", edge_grid(), dim3(256), k_save_j_graph<F, true>, f, E, N * F::MAT_ROWS, cnt, (int*)nullptr, (T*)nullptr, Npad);
                size_t bytes = 0;
                cub::DeviceScan::ExclusiveSum(nullptr, bytes, cnt, mJ.rowPtr, (int)(rows + 1), stream);
                cub::DeviceScan::ExclusiveSum(m_tmp.get(bytes), bytes, cnt, mJ.rowPtr, (int)(rows + 1), stream);
                int nnz = 0;
                OPTK_CUDA(cudaMemcpyAsync(&nnz, mJ.rowPtr + rows, sizeof(int), cudaMemcpyDeviceToHost, stream));
                OPTK_CUDA(cudaStreamSynchronize(stream));
                cudaFree(cnt);
                mJ.nnz = nnz;
                OPTK_CUDA(cudaMalloc(&mJ.colInd, (size_t)std::max(nnz, 1) * sizeof(int)));
                OPTK_CUDA(cudaMalloc(&mJ.val, (size_t)std::max(nnz, 1) * sizeof(T)));
                OPTK_CUDA(cudaMalloc(&m_Jp, (size_t)std::max<long>(rows, 1) * sizeof(T)));
                fill_j();
                this->pre_launch("JT alloc");
                csr_transpose_pattern(mJ, mJT, &m_perm, m_tmp, stream);
                this->post_launch();
                if (mat_mode == 2) {
                    this->pre_launch("J^TJ alloc");
                    csr_ata_pattern(mJ, mJTJ, &m_rowOf, m_tmp, stream);
                    this->post_launch();
                }
            } else fill_j();
            this->pre_launch("J_transpose");
            csr_transpose_values(mJ, mJT, m_perm, stream);
            this->post_launch();
            if (mat_mode == 2) {
                this->pre_launch("JTJ multiply");
                csr_ata_values(mJ, mJT, mJTJ, m_rowOf, stream);
                this->post_launch();
            }
        }
    }
    void fill_j() {
        if constexpr (F::HAS_ROWS_J) {
            if constexpr (F::MAT_ROWS > 0) this->launch_tiles("saveJToCRS", tg.n, sblock, k_save_j<F, false>, f, tg, mJ.rowPtr, mJ.colInd, mJ.val, Npad);
            if constexpr (F::MAT_GROWS > 0) if (E > 0) this->launch("saveJToCRS_Graph", edge_grid(), dim3(256), k_save_j_graph<F, false>, f, E, N * F::MAT_ROWS, mJ.rowPtr, mJ.colInd, mJ.val, Npad);
        }
    }
    // cusparseInner + PCGStep1_Finish (solver:1660-1756, 646-660)
    void apply_materialized(V dir, V out, double* den, double* zero_slot) {
        if (mat_mode == 2) {
            this->pre_launch("J^TJp");
            csr_spmv(mJTJ, (const T*)dir.d, out.d, nsm, stream);
            this->post_launch();
        } else {
            this->pre_launch("Jp");
            csr_spmv(mJ, (const T*)dir.d, m_Jp, nsm, stream);
            this->post_launch();
            this->pre_launch("J^T");
            csr_spmv(mJT, (const T*)m_Jp, out.d, nsm, stream);
            this->post_launch();
        }
        if (lm) this->launch_tiles("PCGStep1_Finish", tg.n, sblock, k_pcg_step1_finish<F, true>, f, tg, dir, out, CtC, den, zero_slot);
        else    this->launch_tiles("PCGStep1_Finish", tg.n, sblock, k_pcg_step1_finish<F, false>, f, tg, dir, out, CtC, den, zero_slot);
    }

    int unknown_channels() const override { return U; }
    long site_count() const override { return N; }
    void bind(void** pp) override { f.bind(pp); }
    void set_lm(double rad, double mn, double mx) override { radius = (T)rad; min_diag = (T)mn; max_diag = (T)mx; }

    double* S(int slot) const { return scal + slot; }
    long flat_vecs() const { return (long)U * Npad / VecIO<T>::W; }
    dim3 flat_grid() const { return dim3((unsigned)std::min<long>((flat_vecs() + 255) / 256, (long)nsm * 8)); }
    dim3 edge_grid() const { return dim3((unsigned)((E + 255) / 256)); }

    // ---- multi-GPU: ownership, scalar all-reduce, halo exchange ---------------------------------
    int set_slab(int rk, int nr, int h, const void* uid) override {
        if (F::ND < 2 || GRAPH || !F::SLAB_OK) { fprintf(stderr, "Opt: slab decomposition is not available for this energy\n"); return -1; }
        const int L = F::ND == 3 ? f.D : f.H;
        rank = rk; nranks = nr; halo = h;
        // ghost rows of p are refreshed after PCGStep3; the fused PCGStep3+1 kernel stays on only for functors whose
        // kernel takes its ghost rows from the exchanged p' (FUSE_SLAB, see k_boundary_dir)
        if (nranks > 1 && !(F::FUSE_SLAB && f.fuse_slab_ok())) fuse = false;
        f.o_lo = rank > 0 ? halo : 0;
        f.o_hi = rank < nranks - 1 ? L - halo : L;
        if (f.o_hi - f.o_lo < halo) { fprintf(stderr, "Opt: slab thinner than its halo\n"); return -1; }
        long slow_off = 0;
        if (nranks > 1) {
            ncclUniqueId id; memcpy(&id, uid, sizeof(id));
            OPTK_NCCL(ncclCommInitRank(&comm, nranks, id, rank));
            // global index of local row 0 = owned rows of the ranks before this one - own ghost rows
            int* cnt = nullptr;
            OPTK_CUDA(cudaMalloc(&cnt, (size_t)(nranks + 1) * sizeof(int)));
            const int own = f.o_hi - f.o_lo;
            OPTK_CUDA(cudaMemcpyAsync(cnt + nranks, &own, sizeof(int), cudaMemcpyHostToDevice, stream));
            OPTK_NCCL(ncclAllGather(cnt + nranks, cnt, 1, ncclInt, comm, stream));
            std::vector<int> h(nranks);
            OPTK_CUDA(cudaMemcpyAsync(h.data(), cnt, (size_t)nranks * sizeof(int), cudaMemcpyDeviceToHost, stream));
            OPTK_CUDA(cudaStreamSynchronize(stream));
            cudaFree(cnt);
            setup_peer_reduce();
            for (int r2 = 0; r2 < rank; ++r2) slow_off += h[r2];
            slow_off -= f.o_lo;
        }
        f.set_slab_geometry(slow_off, rank > 0, rank < nranks - 1);
        return 0;
    }
    void reduce(int slot, int count = 1) {                  // sum over ranks, in place, device-resident
        if (nranks <= 1) return;
        if (xr_on && count <= XR_MAX_VALS) {
            this->launch("AllReduceP2P", dim3(1), dim3(32), k_xreduce, scal + slot, count, xr_peers, rank, nranks, ++xr_epoch, xr_error);
            return;
        }
        OPTK_NCCL(ncclAllReduce(scal + slot, scal + slot, count, ncclDouble, ncclSum, comm, stream));
    }
    // Map a mailbox of every peer (one process per GPU on one node): the cudaIpc handles travel through an NCCL
    // all-gather; a magic number written by the owner is read back through the mapping before the path is enabled.
    void setup_peer_reduce() {
        if (getenv("OPTB200_NO_P2P") || nranks > XR_MAX_RANKS) return;
        const size_t bytes = 4u << 20;                      // its own allocation block: the IPC handle maps the block base
        OPTK_CUDA(cudaMalloc(&xr_buf, bytes));
        OPTK_CUDA(cudaMemsetAsync(xr_buf, 0, bytes, stream));
        OPTK_CUDA(cudaMalloc(&xr_error, 2 * sizeof(int)));   // [0] error flag, [1] arrival counter of k_boundary_push
        OPTK_CUDA(cudaMemsetAsync(xr_error, 0, 2 * sizeof(int), stream));
        const long long magic = 0x4f70744232303000LL + rank;
        OPTK_CUDA(cudaMemcpyAsync(&xr_buf[2 * XR_MAX_RANKS].epoch, &magic, sizeof(magic), cudaMemcpyHostToDevice, stream));
        cudaIpcMemHandle_t mine;
        OPTK_CUDA(cudaIpcGetMemHandle(&mine, xr_buf));
        char* dh = nullptr;
        OPTK_CUDA(cudaMalloc(&dh, sizeof(mine) * (size_t)(nranks + 1)));
        OPTK_CUDA(cudaMemcpyAsync(dh + sizeof(mine) * (size_t)nranks, &mine, sizeof(mine), cudaMemcpyHostToDevice, stream));
        OPTK_NCCL(ncclAllGather(dh + sizeof(mine) * (size_t)nranks, dh, sizeof(mine), ncclChar, comm, stream));
        std::vector<cudaIpcMemHandle_t> all(nranks);
        OPTK_CUDA(cudaMemcpyAsync(all.data(), dh, sizeof(mine) * (size_t)nranks, cudaMemcpyDeviceToHost, stream));
        OPTK_CUDA(cudaStreamSynchronize(stream));
        cudaFree(dh);
        bool ok = true;
        for (int r2 = 0; r2 < nranks && ok; ++r2) {
            if (r2 == rank) { xr_peers.p[r2] = xr_buf; continue; }
            void* ptr = nullptr;
            if (cudaIpcOpenMemHandle(&ptr, all[r2], cudaIpcMemLazyEnablePeerAccess) != cudaSuccess) { cudaGetLastError(); ok = false; break; }
            xr_opened.push_back(ptr);
            xr_peers.p[r2] = (XSlot*)ptr;
            long long seen = 0;
            if (cudaMemcpy(&seen, &xr_peers.p[r2][2 * XR_MAX_RANKS].epoch, sizeof(seen), cudaMemcpyDeviceToHost) != cudaSuccess ||
                seen != 0x4f70744232303000LL + r2) { cudaGetLastError(); ok = false; }
        }
        // every rank must take the same path
        int* flag = nullptr;
        OPTK_CUDA(cudaMalloc(&flag, sizeof(int)));
        const int mine_ok = ok ? 1 : 0;
        OPTK_CUDA(cudaMemcpyAsync(flag, &mine_ok, sizeof(int), cudaMemcpyHostToDevice, stream));
        OPTK_NCCL(ncclAllReduce(flag, flag, 1, ncclInt, ncclMin, comm, stream));
        int all_ok = 0;
        OPTK_CUDA(cudaMemcpyAsync(&all_ok, flag, sizeof(int), cudaMemcpyDeviceToHost, stream));
        OPTK_CUDA(cudaStreamSynchronize(stream));
        cudaFree(flag);
        xr_on = all_ok == 1;
        if (!xr_on && rank == 0) fprintf(stderr, "Opt: peer-memory all-reduce unavailable, using ncclAllReduce\n");
        if (xr_on && ((fuse && F::VEC_UB == 0) || F::VEC_UB > 0)) setup_peer_halo();
    }
    // neighbours' solver-vector pools (for k_boundary_push) and where their ghost rows are
    void setup_peer_halo() {
        struct Geo { cudaIpcMemHandle_t h; long npad; long ghost_from_below; long ghost_from_above; };
        Geo mine;
        OPTK_CUDA(cudaIpcGetMemHandle(&mine.h, pool));
        mine.npad = Npad;
        mine.ghost_from_below = f.o_lo - halo;              // rows [o_lo - halo, o_lo) are filled by rank - 1
        mine.ghost_from_above = f.o_hi;                     // rows [o_hi, o_hi + halo) by rank + 1
        char* dg = nullptr;
        OPTK_CUDA(cudaMalloc(&dg, sizeof(Geo) * (size_t)(nranks + 1)));
        OPTK_CUDA(cudaMemcpyAsync(dg + sizeof(Geo) * (size_t)nranks, &mine, sizeof(Geo), cudaMemcpyHostToDevice, stream));
        OPTK_NCCL(ncclAllGather(dg + sizeof(Geo) * (size_t)nranks, dg, sizeof(Geo), ncclChar, comm, stream));
        std::vector<Geo> all(nranks);
        OPTK_CUDA(cudaMemcpyAsync(all.data(), dg, sizeof(Geo) * (size_t)nranks, cudaMemcpyDeviceToHost, stream));
        OPTK_CUDA(cudaStreamSynchronize(stream));
        cudaFree(dg);
        bool ok = true;
        for (int side = 0; side < 2 && ok; ++side) {
            const int nb = side == 0 ? rank - 1 : rank + 1;
            if (nb < 0 || nb >= nranks) continue;
            void* ptr = nullptr;
            if (cudaIpcOpenMemHandle(&ptr, all[nb].h, cudaIpcMemLazyEnablePeerAccess) != cudaSuccess) { cudaGetLastError(); ok = false; break; }
            xr_opened.push_back(ptr);
            nb_pool[side] = (T*)ptr;
            nb_npad[side] = all[nb].npad;
            nb_row[side] = side == 0 ? all[nb].ghost_from_above : all[nb].ghost_from_below;   // I am above rank-1, below rank+1
        }
        int* flag = nullptr;
        OPTK_CUDA(cudaMalloc(&flag, sizeof(int)));
        const int mine_ok = ok ? 1 : 0;
        OPTK_CUDA(cudaMemcpyAsync(flag, &mine_ok, sizeof(int), cudaMemcpyHostToDevice, stream));
        OPTK_NCCL(ncclAllReduce(flag, flag, 1, ncclInt, ncclMin, comm, stream));
        int all_ok = 0;
        OPTK_CUDA(cudaMemcpyAsync(&all_ok, flag, sizeof(int), cudaMemcpyDeviceToHost, stream));
        OPTK_CUDA(cudaStreamSynchronize(stream));
        cudaFree(flag);
        xh_on = all_ok == 1;
    }
    // ghost rows/planes of a contiguous array with `nplanes` planes of stride `pstride` elements,
    // `rowsz` elements per row (2-D) or per plane (3-D) of the slowest dimension
    template <class E_>
    void exchange_raw(E_* base, int nplanes, size_t pstride, size_t rowsz) {
        if (nranks <= 1) return;
        const size_t cnt = rowsz * halo * sizeof(E_);
        const int lo = f.o_lo, hi = f.o_hi;
        OPTK_NCCL(ncclGroupStart());
        for (int u = 0; u < nplanes; ++u) {
            char* pl = (char*)(base + (size_t)u * pstride);
            if (rank > 0) {
                OPTK_NCCL(ncclSend(pl + (size_t)lo * rowsz * sizeof(E_), cnt, ncclChar, rank - 1, comm, stream));
                OPTK_NCCL(ncclRecv(pl + (size_t)(lo - halo) * rowsz * sizeof(E_), cnt, ncclChar, rank - 1, comm, stream));
            }
            if (rank < nranks - 1) {
                OPTK_NCCL(ncclSend(pl + (size_t)(hi - halo) * rowsz * sizeof(E_), cnt, ncclChar, rank + 1, comm, stream));
                OPTK_NCCL(ncclRecv(pl + (size_t)hi * rowsz * sizeof(E_), cnt, ncclChar, rank + 1, comm, stream));
            }
        }
        OPTK_NCCL(ncclGroupEnd());
    }
    size_t slow_rowsz() const { return F::ND == 3 ? (size_t)f.W * f.H : (size_t)f.W; }
    void exchange(V v) {
        if constexpr (F::VEC_UB > 0) {                      // blocked layout: a slow-dimension slice is contiguous
            const long slice = (long)slow_rowsz() * U;
            if (xh_on && nranks > 1 && v.d >= pool && (slice * halo) % VecIO<T>::W == 0) {
                using V4 = typename VecIO<T>::V;
                const long per = slice / VecIO<T>::W * 1;   // vectors per slow index (slice is a multiple of the vector width: rows of 32 sites)
                const long slot = (long)((v.d - pool) / ((long)U * Npad));
                V4* below = rank > 0 ? (V4*)(nb_pool[0] + slot * U * nb_npad[0]) + nb_row[0] * per : nullptr;
                V4* above = rank < nranks - 1 ? (V4*)(nb_pool[1] + slot * U * nb_npad[1]) + nb_row[1] * per : nullptr;
                const long cnt = per * halo;
                this->launch("HaloPushP2P", dim3((unsigned)std::min<long>((2 * cnt + 255) / 256, (long)nsm * 8)), dim3(256), k_halo_push<V4>,
                             (const V4*)v.d, cnt, (long)f.o_lo * per, (long)(f.o_hi - halo) * per, below, above,
                             xr_peers, rank, nranks, ++xh_epoch, xr_error, (unsigned*)(xr_error + 1));
                return;
            }
            exchange_raw(v.d, 1, 0, slow_rowsz() * U);
        }
        else exchange_raw(v.d, U, (size_t)v.plane, slow_rowsz());
    }
    void exchange_unknowns() {
        if (nranks <= 1) return;
        void* ptrs[8]; int ch[8];
        const int n = f.unknown_images(ptrs, ch);
        for (int i = 0; i < n; ++i) exchange_raw((T*)ptrs[i], 1, 0, slow_rowsz() * ch[i]);
    }

    double read_scalar(int slot) {                          // blocking D2H (backend_cuda.t:353-355)
        if (h_scal_dev) {                                    // zero-copy publish (no copy engine), then wait for the stream
            k_publish_scalar<<<1, 1, 0, stream>>>(scal + slot, xr_on ? xr_error : nullptr, h_scal_dev, (volatile int*)(h_scal_dev + 1));
            OPTK_CUDA(cudaGetLastError());
        } else {
            OPTK_CUDA(cudaMemcpyAsync(h_scal, scal + slot, sizeof(double), cudaMemcpyDeviceToHost, stream));
            if (xr_on) OPTK_CUDA(cudaMemcpyAsync(h_scal + 1, xr_error, sizeof(int), cudaMemcpyDeviceToHost, stream));
        }
        OPTK_CUDA(cudaStreamSynchronize(stream));
        if (xr_on && *(const int*)(h_scal + 1) != 0) {
            optk_fatal(3, "Opt: a peer rank never arrived at a scalar all-reduce (rank %d of %d)\n", rank, nranks);
        }
        return h_scal[0];
    }

    // (vertex, slot)-sorted incidence lists of the graph (HAS_GATHER functors); rebuilt at the first
    // nonlinear iteration of every solve because the index arrays belong to the caller
    void build_incidence() {
        if constexpr (F::HAS_GATHER) {
            const long n = E * F::GV, nkeys = N;
            if (!inc) {
                OPTK_CUDA(cudaMalloc(&inc, (size_t)std::max<long>(n, 1) * sizeof(unsigned)));
                OPTK_CUDA(cudaMalloc(&inc_off, (size_t)(nkeys + 1) * sizeof(int)));
                OPTK_CUDA(cudaMalloc(&key_a, (size_t)std::max<long>(n, 1) * sizeof(unsigned)));
                OPTK_CUDA(cudaMalloc(&key_b, (size_t)std::max<long>(n, 1) * sizeof(unsigned)));
                OPTK_CUDA(cudaMalloc(&val_a, (size_t)std::max<long>(n, 1) * sizeof(unsigned)));
                size_t sel_bytes = 0;
                OPTK_CUDA(cudaMalloc(&inc_count, sizeof(int)));
                cub::DeviceRadixSort::SortPairs(nullptr, sort_tmp_bytes, key_a, key_b, val_a, inc, (int)n);
                cub::DeviceSelect::If(nullptr, sel_bytes, key_a, key_b, inc_count, (int)n, IncidenceKept());
                sort_tmp_bytes = std::max(sort_tmp_bytes, sel_bytes);
                OPTK_CUDA(cudaMalloc(&sort_tmp, sort_tmp_bytes));
            }
            int vbits = 1;
            while ((1L << vbits) < nkeys) ++vbits;
            if (n > 0) {
                // two stable passes: by entry (other endpoint, slot), then by vertex
                this->launch("graphIncidenceKeys", dim3((unsigned)((n + 255) / 256)), dim3(256), k_incidence_keys<F>, f, n, key_a, val_a);
                this->pre_launch("graphIncidenceSort");
                cub::DeviceRadixSort::SortPairs(sort_tmp, sort_tmp_bytes, key_a, key_b, val_a, inc, (int)n, 0, vbits + 1, stream);   // entries -> key_b, vertices -> inc
                cub::DeviceRadixSort::SortPairs(sort_tmp, sort_tmp_bytes, inc, key_a, key_b, val_a, (int)n, 0, vbits, stream);      // vertices -> key_a, entries -> val_a
                this->post_launch();
                this->launch("graphIncidenceMerge", dim3((unsigned)((n + 255) / 256)), dim3(256), k_incidence_merge, (const unsigned*)val_a, (const unsigned*)key_a, n, inc, key_b);
                this->pre_launch("graphIncidenceCompact");
                cub::DeviceSelect::If(sort_tmp, sort_tmp_bytes, inc, val_a, inc_count, (int)n, IncidenceKept(), stream);      // entries
                cub::DeviceSelect::If(sort_tmp, sort_tmp_bytes, key_b, key_a, inc_count, (int)n, IncidenceKept(), stream);    // their vertices
                this->post_launch();
                std::swap(inc, val_a);
            }
            int kept = 0;
            if (n > 0) {
                OPTK_CUDA(cudaMemcpyAsync(&kept, inc_count, sizeof(int), cudaMemcpyDeviceToHost, stream));
                OPTK_CUDA(cudaStreamSynchronize(stream));
            }
            this->launch("graphIncidenceOffsets", dim3((unsigned)((kept + 1 + 255) / 256)), dim3(256), k_incidence_offsets, (const unsigned*)key_a, (long)kept, nkeys, inc_off);
        }
    }
    bool use_gather() const { return F::HAS_GATHER && getenv("OPTB200_GRAPH_ATOMICS") == nullptr; }
    // segmented scatter of the hyper-edge energies: (edge, slot) ids sorted by vertex, offsets per vertex, partial rows
    bool use_seg() const { return F::HAS_SEGSCATTER && GRAPH && getenv("OPTB200_GRAPH_ATOMICS") == nullptr; }
    unsigned* seg_ids = nullptr; unsigned* seg_rank = nullptr; int* seg_off = nullptr; T* seg_contrib = nullptr;
    void build_segments() {
        if constexpr (F::HAS_SEGSCATTER) {
            const long n = E * F::GV;
            if (!seg_off) {
                OPTK_CUDA(cudaMalloc(&seg_ids, (size_t)std::max<long>(n, 1) * sizeof(unsigned)));
                OPTK_CUDA(cudaMalloc(&seg_rank, (size_t)std::max<long>(n, 1) * sizeof(unsigned)));
                OPTK_CUDA(cudaMalloc(&seg_off, (size_t)(N + 1) * sizeof(int)));
                OPTK_CUDA(cudaMalloc(&seg_contrib, (size_t)std::max<long>(n, 1) * U * sizeof(T)));
                OPTK_CUDA(cudaMalloc(&key_a, (size_t)std::max<long>(n, 1) * sizeof(unsigned)));
                OPTK_CUDA(cudaMalloc(&key_b, (size_t)std::max<long>(n, 1) * sizeof(unsigned)));
                OPTK_CUDA(cudaMalloc(&val_a, (size_t)std::max<long>(n, 1) * sizeof(unsigned)));
                cub::DeviceRadixSort::SortPairs(nullptr, sort_tmp_bytes, key_a, key_b, val_a, seg_ids, (int)n);
                OPTK_CUDA(cudaMalloc(&sort_tmp, sort_tmp_bytes));
            }
            int vbits = 1;
            while ((1L << vbits) < N) ++vbits;
            if (n > 0) {
                this->launch("graphSegmentKeys", dim3((unsigned)((n + 255) / 256)), dim3(256), k_segment_keys<F>, f, n, key_a, val_a);
                this->pre_launch("graphSegmentSort");
                cub::DeviceRadixSort::SortPairs(sort_tmp, sort_tmp_bytes, key_a, key_b, val_a, seg_ids, (int)n, 0, vbits, stream);   // stable
                this->post_launch();
                this->launch("graphSegmentRank", dim3((unsigned)((n + 255) / 256)), dim3(256), k_segment_rank, n, (const unsigned*)seg_ids, seg_rank);
            }
            this->launch("graphSegmentOffsets", dim3((unsigned)((n + 1 + 255) / 256)), dim3(256), k_incidence_offsets, (const unsigned*)key_b, n, N, seg_off);
        }
    }

    void precompute() override {
        if constexpr (F::HAS_CUSTOM_PRECOMPUTE) {           // functor owns a table in its own layout
            this->pre_launch("precompute");
            f.custom_precompute(stream, nsm);
            this->post_launch();
        } else if (F::HAS_PRECOMPUTE) {
            if constexpr (F::HAS_FAST_PRECOMPUTE) {
                if (f.fast_ok()) {
                    this->pre_launch("precompute");
                    f.fast_precompute(stream, nsm);
                    this->post_launch();
                    return;
                }
            }
            this->launch_tiles("precompute", tg.n, sblock, k_precompute<F>, f, tg);
        }
    }
    double cost() override {
        OPTK_CUDA(cudaMemsetAsync(S(S_COST), 0, sizeof(double), stream));
        bool fast_cost = false;
        if constexpr (F::HAS_FAST_COST) {
            if (f.fast_ok()) {
                this->launch_tiles_smem("computeCost", f.fast_tiles(), f.fast_block(), f.fast_smem(), f.fast_cost_kernel(), f, S(S_COST));
                fast_cost = true;
            }
        }
        if (!fast_cost) this->launch_tiles("computeCost", tg.n, sblock, k_cost<F>, f, tg, S(S_COST));
        if (GRAPH) this->launch("computeCost_Graph", edge_grid(), dim3(256), k_graph_cost<F>, f, E, S(S_COST));
        reduce(S_COST);
        return (double)(T)read_scalar(S_COST);
    }
    double model_cost() override {
        exchange(delta);
        OPTK_CUDA(cudaMemsetAsync(S(S_MODEL), 0, sizeof(double), stream));
        bool fast_mc = false;
        if constexpr (F::HAS_FAST_MODEL_COST) {
            if (f.fast_ok() && getenv("OPTB200_GENERIC_MODEL_COST") == nullptr) {
                this->launch_tiles_smem("computeModelCost", f.fast_tiles(), f.fast_block(), f.fast_smem(), f.fast_model_cost_kernel(), f, delta, S(S_MODEL));
                fast_mc = true;
            }
        }
        if (!fast_mc) this->launch_tiles("computeModelCost", tg.n, sblock, k_model_cost<F>, f, tg, delta, S(S_MODEL));
        if (GRAPH) this->launch("computeModelCost_Graph", edge_grid(), dim3(256), k_graph_model_cost<F>, f, E, delta, S(S_MODEL));
        reduce(S_MODEL);
        return (double)(T)read_scalar(S_MODEL);
    }
    double fetch_q() override { return (double)(T)read_scalar(S_Q); }

    // active4 (one bit per site of a group of 4) and, for functors whose J^T J kernel walks a work list, that list
    void active_groups() {
        if constexpr (F::HAS_EXCLUDE) {
        OPTK_CUDA(cudaMemsetAsync(mask_changed, 0, sizeof(int), stream));
        this->launch("activeGroups", dim3((unsigned)((Npad / 4 + 255) / 256)), dim3(256), k_active4<F>, f, Npad / 4, active4, mask_changed);
        {   // delta r z p Ap pre prevX p_alt are the first eight vectors of the pool; all but pre / prevX must be 0 at excluded sites
            using V4 = typename VecIO<T>::V;
            const long per = flat_vecs();
            this->launch("clearAfterMaskChange", flat_grid(), dim3(256), k_zero_if<V4>, (const int*)mask_changed, (V4*)pool, 5 * per);
            this->launch("clearAfterMaskChange", flat_grid(), dim3(256), k_zero_if<V4>, (const int*)mask_changed, (V4*)(pool + (size_t)7 * U * Npad), per);   // slot 7: whichever of p / p_alt lives there
        }
        if constexpr (F::HAS_ACTIVE_ITEMS) {
            if (f.fast_ok()) {
                this->pre_launch("activeItems");
                f.build_active_items(stream, active4);
                this->post_launch();
            }
        }
    }
    }
    void pcg_begin(int nIter) override {
        OPTK_CUDA(cudaMemsetAsync(scal, 0, S_COUNT * sizeof(double), stream));
        if (nIter == 0 && use_gather()) build_incidence();
        if (nIter == 0 && use_seg()) build_segments();
        active_groups();                                    // Exclude() may read inputs the caller changed
        bool fast_init = false;
        if constexpr (F::HAS_FAST_INIT) {
            if (f.fast_ok()) {
                this->launch_tiles_smem("PCGInit1", f.fast_tiles(), f.fast_block(), f.fast_smem(), f.fast_init_kernel(), f, delta, r, p, pre, S(S_RZ0));
                fast_init = true;
            }
        }
        if (!fast_init) this->launch_tiles("PCGInit1", tg.n, sblock, k_pcg_init1<F>, f, tg, delta, r, p, pre, S(S_RZ0));
        if (GRAPH) {
            this->launch("PCGInit1_Graph", edge_grid(), dim3(256), k_graph_jtf<F>, f, E, r, pre);
            this->launch_tiles("PCGInit1_Finish", tg.n, sblock, k_pcg_init1_finish<F>, f, tg, r, pre, p, S(S_RZ0));
        }
        if (lm) {
            OPTK_CUDA(cudaMemsetAsync(S(S_RZ0), 0, sizeof(double), stream));
            if (nIter == 0)                                  // PCGSaveSSq (solver:1052-1057), flat copy
                OPTK_CUDA(cudaMemcpyAsync(SSq.d, pre.d, (size_t)U * Npad * sizeof(T), cudaMemcpyDeviceToDevice, stream));
            const T inv_radius = T(1) / radius;
            if (getenv("OPTB200_GENERIC_CTC")) this->launch_tiles("PCGComputeCtC", tg.n, sblock, k_compute_ctc<F, true>, f, tg, CtC, inv_radius);
            else this->launch_tiles("PCGComputeCtC", tg.n, sblock, k_compute_ctc<F, false>, f, tg, CtC, inv_radius);
            if (GRAPH) this->launch("PCGComputeCtC_Graph", edge_grid(), dim3(256), k_graph_ctc<F>, f, E, CtC, inv_radius);
            this->launch_tiles("PCGFinalizeDiagonal", tg.n, sblock, k_finalize_diagonal<F>, f, tg, CtC, SSq, pre, r, b, p, delta, radius, min_diag, max_diag, S(S_RZ0), S(S_Q));
            reduce(S_Q);
            this->launch("LMBegin", dim3(1), dim3(1), k_lm_begin<T>, (const double*)S(S_Q), S(S_QPREV));   // Q0 (solver:2060)
        }
        reduce(S_RZ0);
        exchange(p);                                        // Step1 reads p over the stencil
        if (mat_mode) materialize(nIter);                   // cusparseOuter (solver:2068)
    }

    void apply_A(const char* name, const char* gname, V dir, V out, double* den, double* zero_slot) {
        if (mat_mode && mJ.rowPtr) { apply_materialized(dir, out, den, zero_slot); return; }
        if constexpr (F::HAS_GATHER) {
            if (use_gather()) {
                const long nb = (N + 127) / 128;
                if (lm) this->launch_tiles(name, nb, dim3(128), k_gather_jtj<F, true>, f, N, (const int*)inc_off, (const unsigned*)inc, dir, out, CtC, den, zero_slot);
                else if (getenv("OPTB200_GATHER_SIMPLE")) this->launch_tiles(name, nb, dim3(128), k_gather_jtj<F, false, true>, f, N, (const int*)inc_off, (const unsigned*)inc, dir, out, CtC, den, zero_slot);
                else    this->launch_tiles(name, nb, dim3(128), k_gather_jtj<F, false>, f, N, (const int*)inc_off, (const unsigned*)inc, dir, out, CtC, den, zero_slot);
                return;
            }
        }
        if constexpr (F::HAS_SEGSCATTER) {
            if (use_seg()) {
                if (E > 0) this->launch(gname, dim3((unsigned)((E + 127) / 128)), dim3(128), k_graph_contrib<F>, f, E, dir, (const unsigned*)seg_rank, seg_contrib);
                if (lm) this->launch_tiles(name, tg.n, sblock, k_apply_jtj_seg<F, true>, f, tg, dir, out, CtC, (const int*)seg_off, (const T*)seg_contrib, den, zero_slot);
                else    this->launch_tiles(name, tg.n, sblock, k_apply_jtj_seg<F, false>, f, tg, dir, out, CtC, (const int*)seg_off, (const T*)seg_contrib, den, zero_slot);
                return;
            }
        }
        if constexpr (F::HAS_FAST_APPLY) {
            if (f.fast_ok()) {
                if (lm) this->launch_tiles_smem(name, f.fast_tiles(), f.fast_block(), f.fast_smem(), f.template fast_kernel<false, true>(), f, dir, dir, dir, out, CtC, (const unsigned char*)active4, (const double*)nullptr, (const double*)nullptr, den, zero_slot, (double*)nullptr);
                else    this->launch_tiles_smem(name, f.fast_tiles(), f.fast_block(), f.fast_smem(), f.template fast_kernel<false, false>(), f, dir, dir, dir, out, CtC, (const unsigned char*)active4, (const double*)nullptr, (const double*)nullptr, den, zero_slot, (double*)nullptr);
                return;
            }
        }
        if (lm) this->launch_tiles(name, tg.n, sblock, k_apply_jtj<F, true>, f, tg, dir, out, CtC, den, zero_slot);
        else    this->launch_tiles(name, tg.n, sblock, k_apply_jtj<F, false>, f, tg, dir, out, CtC, den, zero_slot);
        if (GRAPH) this->launch(gname, edge_grid(), dim3(256), k_graph_jtj<F>, f, E, dir, out, den);
    }

    void pcg_iter(int l, int reset) override {
        struct Range { Range() { nvtxRangePushA("PCG iteration"); } ~Range() { nvtxRangePop(); } } nvtx_range;
        double* aNum = S(S_RZ0 + l % 3);
        double* bNum = S(S_RZ0 + (l + 1) % 3);
        double* nextB = S(S_RZ0 + (l + 2) % 3);
        double* aDen = S(S_ADEN0 + (l & 1));
        double* nextDen = S(S_ADEN0 + ((l + 1) & 1));
        // flat streaming range: everything, except for blocked vectors in slab mode, where a slab's owned planes are one
        // contiguous element range and the ghost planes are left to the halo exchange alone (nothing else may write
        // them once neighbours push into them directly)
        long fb = 0, nv = flat_vecs();
        if (F::VEC_UB > 0 && nranks > 1) {
            const long per = (long)slow_rowsz() * U / VecIO<T>::W;
            fb = f.o_lo * per; nv = f.o_hi * per;
        }
        const unsigned char* act = active4;
        // PCGStep1 also clears q for this iteration (solver:2090)
        if (fuse && l > 0) {
            // p' = z + beta p with beta = RZ[l%3] / RZ[(l-1)%3] (the PCGStep3 of iteration l-1), then Ap'
            double* prevA = S(S_RZ0 + (l + 2) % 3);
            if constexpr (F::VEC_UB > 0) {
                if (nranks > 1) {                           // blocked vectors: a boundary plane is one contiguous slice
                    using V4 = typename VecIO<T>::V;
                    const long per = (long)slow_rowsz() * U / VecIO<T>::W, cnt = per * halo;
                    const long slot = (long)((p_alt.d - pool) / ((long)U * Npad));
                    V4 *below = nullptr, *above = nullptr;
                    if (xh_on && rank > 0) below = (V4*)(nb_pool[0] + slot * U * nb_npad[0]) + nb_row[0] * per;
                    if (xh_on && rank < nranks - 1) above = (V4*)(nb_pool[1] + slot * U * nb_npad[1]) + nb_row[1] * per;
                    this->launch("PCGStep3_BoundaryPush", dim3((unsigned)std::min<long>((2 * cnt + 255) / 256, (long)nsm * 4)), dim3(256), k_halo_push_dir<T>,
                                 (const V4*)z.d, (const V4*)p.d, xh_on ? (V4*)nullptr : (V4*)p_alt.d, cnt, (long)f.o_lo * per, (long)(f.o_hi - halo) * per,
                                 (int)(rank > 0), (int)(rank < nranks - 1), below, above, (const double*)aNum, (const double*)prevA,
                                 xr_peers, rank, nranks, xh_on ? ++xh_epoch : 0LL, xr_error, (unsigned*)(xr_error + 1), (int)xh_on);
                    if (!xh_on) exchange(p_alt);
                }
            } else
            if (nranks > 1) {                               // only reached with F::FUSE_SLAB
                const long rowsz = (long)slow_rowsz();
                const long total = rowsz * halo * 2 * U;
                if (xh_on) {
                    // which of the two p buffers is p' this iteration: the same slot on every rank (they swap in lockstep)
                    const long slot = (long)((p_alt.d - pool) / ((long)U * Npad));
                    HaloPeer<T> below{nullptr, 0, 0}, above{nullptr, 0, 0};
                    if (rank > 0) below = HaloPeer<T>{nb_pool[0] + slot * U * nb_npad[0], nb_npad[0], nb_row[0]};
                    if (rank < nranks - 1) above = HaloPeer<T>{nb_pool[1] + slot * U * nb_npad[1], nb_npad[1], nb_row[1]};
                    this->launch("PCGStep3_BoundaryPush", dim3((unsigned)std::min<long>((total + 255) / 256, (long)nsm * 4)), dim3(256), k_boundary_push<T>,
                                 (const T*)z.d, (const T*)p.d, (long)p.plane, (int)U, rowsz, f.o_lo, f.o_hi, halo, below, above,
                                 (const double*)aNum, (const double*)prevA, xr_peers, rank, nranks, ++xh_epoch, xr_error, (unsigned*)(xr_error + 1));
                } else {
                this->launch("PCGStep3_Boundary", dim3((unsigned)std::min<long>((total + 255) / 256, (long)nsm * 4)), dim3(256), k_boundary_dir<T>,
                             (const T*)z.d, (const T*)p.d, p_alt.d, (long)p.plane, (int)U, rowsz, f.o_lo, f.o_hi, halo, (int)(rank > 0), (int)(rank < nranks - 1),
                             (const double*)aNum, (const double*)prevA);
                exchange(p_alt);
                }
            }
            if constexpr (F::HAS_GATHER) {
                if (use_gather()) {
                    const long nb = (N + 127) / 128;
                    if (lm) this->launch_tiles("PCGStep3+1", nb, dim3(128), k_gather_jtj_fused<F, true>, f, N, (const int*)inc_off, (const unsigned*)inc, z, p, p_alt, Ap, CtC, (const double*)aNum, (const double*)prevA, aDen, bNum, S(S_Q));
                    else    this->launch_tiles("PCGStep3+1", nb, dim3(128), k_gather_jtj_fused<F, false>, f, N, (const int*)inc_off, (const unsigned*)inc, z, p, p_alt, Ap, CtC, (const double*)aNum, (const double*)prevA, aDen, bNum, (double*)nullptr);
                    std::swap(p.d, p_alt.d);
                    goto applied;
                }
            }
            if constexpr (F::HAS_FAST_APPLY) {
                if (f.fast_ok()) {
                    if (lm) this->launch_tiles_smem("PCGStep3+1", f.fast_tiles(), f.fast_block(), f.fast_smem(), f.template fast_kernel<true, true>(), f, z, p, p_alt, Ap, CtC, (const unsigned char*)active4, (const double*)aNum, (const double*)prevA, aDen, bNum, S(S_Q));
                    else    this->launch_tiles_smem("PCGStep3+1", f.fast_tiles(), f.fast_block(), f.fast_smem(), f.template fast_kernel<true, false>(), f, z, p, p_alt, Ap, CtC, (const unsigned char*)active4, (const double*)aNum, (const double*)prevA, aDen, bNum, (double*)nullptr);
                    std::swap(p.d, p_alt.d);
                    goto applied;
                }
            }
            if (lm) this->launch_tiles("PCGStep3+1", tg.n, sblock, k_apply_jtj_fused<F, true>, f, tg, z, p, p_alt, Ap, CtC, aNum, prevA, aDen, bNum, S(S_Q));
            else    this->launch_tiles("PCGStep3+1", tg.n, sblock, k_apply_jtj_fused<F, false>, f, tg, z, p, p_alt, Ap, CtC, aNum, prevA, aDen, bNum, (double*)nullptr);
            std::swap(p.d, p_alt.d);
            if (GRAPH) this->launch("PCGStep1_Graph", edge_grid(), dim3(256), k_graph_jtj<F>, f, E, p, Ap, aDen);
        } else {
            apply_A("PCGStep1", "PCGStep1_Graph", p, Ap, aDen, lm ? S(S_Q) : nullptr);
        }
    applied:
        reduce(int(aDen - scal));
        if (lm && reset) {
            this->launch_tiles("PCGStep2_1stHalf", (flat_vecs() + 255) / 256, dim3(256), k_axpy_alpha<T>, fb, nv, act, Npad, delta.d, p.d, aNum, aDen, (const double*)S(S_STOP));
            exchange(delta);
            apply_A("computeAdelta", "computeAdelta_Graph", delta, Adelta, S(S_DUMMY), nullptr);
            this->launch_tiles("PCGStep2_2ndHalf", (flat_vecs() + 255) / 256, dim3(256), k_residual_reset<T, F::USE_PRE>, fb, nv, act, Npad, delta.d, r.d, Adelta.d, pre.d, b.d, z.d, bNum, S(S_Q), nextDen);
        } else if (lm) {
            this->launch_tiles("PCGStep2", (flat_vecs() + 255) / 256, dim3(256), k_pcg_step2<T, F::USE_PRE, true>, fb, nv, act, Npad, delta.d, p.d, (const T*)nullptr, (double*)nullptr, r.d, Ap.d, pre.d, b.d, z.d, aNum, aDen, bNum, S(S_Q), nextDen, (const double*)S(S_STOP), lm_fold);
            lm_fold_used = lm_fold.on != 0;
        } else {
            // deferred delta update (see k_pcg_step2): only inside pcg_run's Gauss-Newton loop with the fused kernel, where the
            // other p buffer still holds the previous direction; `lazy_last` = this is the loop's final iteration
            const int dm = !lazy_delta ? 0 : (l & 1) ? 2 : (lazy_last ? 0 : 1);
            auto go = [&](auto kernel) {
                this->launch_tiles("PCGStep2", (flat_vecs() + 255) / 256, dim3(256), kernel, fb, nv, act, Npad, delta.d, p.d, (const T*)p_alt.d, S(S_APREV), r.d, Ap.d, pre.d, b.d, z.d, aNum, aDen, bNum, S(S_Q), nextDen, (const double*)nullptr, LmFold{});
            };
            if (dm == 0) go(k_pcg_step2<T, F::USE_PRE, false, 0>);
            else if (dm == 1) go(k_pcg_step2<T, F::USE_PRE, false, 1>);
            else go(k_pcg_step2<T, F::USE_PRE, false, 2>);
        }
        reduce(int(bNum - scal));
        if (lm) reduce(S_Q);
        if (!fuse) {
            this->launch_tiles("PCGStep3", (flat_vecs() + 255) / 256, dim3(256), k_pcg_step3<T>, fb, nv, act, Npad, z.d, p.d, bNum, aNum, nextB);
            exchange(p);
        }
        // fused mode: PCGStep3 of this iteration happens inside the next iteration's first kernel; after
        // the last iteration its result (the next search direction) is never used (solver:2203-2227)
    }
    // (A cooperative single-kernel form of this loop -- grid-wide barriers instead of two launches per iteration -- was built
    // and measured in r2o on image_warping 512^2: 18.64 us per PCG iteration against 18.59 us for the launches below.  That
    // problem is bound by the sincos arithmetic of its generic J^T J functor, not by launch latency, so the kernel was removed.)
    LmFold lm_fold{}; bool lm_fold_used = false; unsigned* lm_counter = nullptr;
    bool lm_fold_ok = getenv("OPTB200_LM_SEPARATE_TEST") == nullptr;
    bool lazy_delta = false, lazy_last = false;
    void pcg_run(int lIterations) override {
        lazy_delta = fuse && !lm && !mat_mode && getenv("OPTB200_NO_LAZY_DELTA") == nullptr;
        for (int l = 0; l < lIterations; ++l) { lazy_last = l + 1 == lIterations; pcg_iter(l, 0); }
        lazy_delta = false;
    }
    // LM inner loop (solver:2072-2201) without a blocking fetchQ per iteration.  The exit test runs on the
    // device (k_lm_zeta); the host queues iteration l+1 BEFORE it looks at the verdict of iteration l, so the
    // GPU never waits for the host.  If iteration l fired the test, iteration l+1 is already in the stream: its
    // step kernels see the stop flag and use alpha = 0, so delta -- the only result of the loop that is read
    // afterwards (model cost, linear update; the next PCGInit1 rebuilds everything else) -- is exactly what the
    // reference has after its `break`.  Returns the number of iterations the reference would have executed.
    bool wait_box(int slot, long long expect) {
        for (long spins = 1;; ++spins) {
            if (__atomic_load_n(&lm_box[slot].seq, __ATOMIC_ACQUIRE) == expect) return true;
            if ((spins & 0x3fff) == 0) {
                cudaError_t e = cudaStreamQuery(stream);
                if (e == cudaSuccess) {                      // stream drained: the mailbox must be visible now
                    if (__atomic_load_n(&lm_box[slot].seq, __ATOMIC_ACQUIRE) == expect) return true;
                    optk_fatal(1, "Opt: LM mailbox was not written (sequence %lld)\n", expect);
                }
                if (e != cudaErrorNotReady) OPTK_CUDA(e);
            }
        }
    }
    int pcg_run_lm(int lIterations, int reset_period, double q_tol, double* zeta_out, int* fired) override {
        long long seq_of[2] = {0, 0};
        int executed = lIterations;
        *zeta_out = 0.0; *fired = 0;
        for (int l = 0; l <= lIterations; ++l) {
            if (l < lIterations) {
                const bool reset = ((l + 1) % reset_period) == 0;
                seq_of[l & 1] = ++lm_seq;
                // single GPU, plain PCGStep2: the test rides in that kernel's tail (LmFold); otherwise its own launch
                lm_fold = LmFold{};
                if (lm_fold_ok && nranks == 1 && !reset)
                    lm_fold = LmFold{1, l, q_tol, S(S_QPREV), S(S_STOP), lm_box + (l & 1), seq_of[l & 1], lm_counter};
                lm_fold_used = false;
                pcg_iter(l, reset);
                lm_fold = LmFold{};
                if (!lm_fold_used)
                    this->launch("LMExitTest", dim3(1), dim3(1), k_lm_zeta<T>, l, q_tol, (const double*)S(S_Q), S(S_QPREV), S(S_STOP), lm_box + (l & 1), seq_of[l & 1]);
            }
            if (l > 0) {                                    // verdict of iteration l-1, while iteration l runs
                const int k = (l - 1) & 1;
                wait_box(k, seq_of[k]);
                if (lm_box[k].stop) { executed = l; *zeta_out = lm_box[k].zeta; *fired = 1; break; }
            }
        }
        return executed;
    }

    static constexpr bool FLAT_X = F::U == 1 && F::VEC_UB == 0 && !GRAPH;
    template <int MODE> bool flat_update(const char* name, V v) {
        if constexpr (FLAT_X) {
            if (N % 4 == 0 && comm == nullptr) {
                const long n4 = N / 4;
                void* img[8]; int ch[8];
                f.unknown_images(img, ch);
                this->launch(name, dim3((unsigned)std::min<long>((n4 + 255) / 256, (long)nsm * 16)), dim3(256), k_flat_update<T, MODE>, n4, (T*)img[0], v.d);
                return true;
            }
        }
        return false;
    }
    void save_unknowns() override {
        if (flat_update<1>("savePreviousUnknowns", prevX)) return;
        this->launch_tiles("savePreviousUnknowns", tg.n, sblock, k_unknown_update<F, 1>, f, tg, prevX);
    }
    void linear_update() override {
        if (!flat_update<0>("PCGLinearUpdate", delta))
            this->launch_tiles("PCGLinearUpdate", tg.n, sblock, k_unknown_update<F, 0>, f, tg, delta);
        exchange_unknowns();                                // neighbours read X over the stencil
    }
    void revert_update() override {
        if (!flat_update<2>("revertUpdate", prevX))
            this->launch_tiles("revertUpdate", tg.n, sblock, k_unknown_update<F, 2>, f, tg, prevX);
        exchange_unknowns();
    }

    V* vec_by_name(const char* n) {
        struct { const char* n; V* v; } tab[] = {{"delta", &delta}, {"r", &r}, {"p", &p}, {"Ap", &Ap}, {"z", &z},
                                                      {"pre", &pre}, {"CtC", &CtC}, {"b", &b}, {"prevX", &prevX}};
        for (auto& t : tab) if (!strcmp(t.n, n)) return t.v;
        return nullptr;
    }
    int debug_read(const char* name, void* host_out) override {
        V* v = vec_by_name(name);
        if (!v || !v->d) return -1;
        std::vector<T> tmp((size_t)U * Npad);
        OPTK_CUDA(cudaStreamSynchronize(stream));
        OPTK_CUDA(cudaMemcpy(tmp.data(), v->d, tmp.size() * sizeof(T), cudaMemcpyDeviceToHost));
        T* o = (T*)host_out;
        for (long s = 0; s < N; ++s) for (int u = 0; u < U; ++u) o[s * U + u] = tmp[(size_t)V::host_index(u, s, Npad)];
        return 0;
    }
    int debug_read_range(const char* name, long s0, long n, void* host_out) override {
        V* v = vec_by_name(name);
        if (!v || !v->d || s0 < 0 || n < 0 || s0 + n > N) return -1;
        OPTK_CUDA(cudaStreamSynchronize(stream));
        T* o = (T*)host_out;
        if (F::VEC_UB < 0) {
            OPTK_CUDA(cudaMemcpy(o, v->d + (size_t)s0 * U, (size_t)n * U * sizeof(T), cudaMemcpyDeviceToHost));
        } else if (F::VEC_UB == 0) {
            std::vector<T> tmp((size_t)n);
            for (int u = 0; u < U; ++u) {
                OPTK_CUDA(cudaMemcpy(tmp.data(), v->d + (size_t)u * Npad + s0, (size_t)n * sizeof(T), cudaMemcpyDeviceToHost));
                for (long s = 0; s < n; ++s) o[s * U + u] = tmp[(size_t)s];
            }
        } else {
            const long b0 = s0 >> 5, b1 = (s0 + n + 31) >> 5;
            std::vector<T> tmp((size_t)(b1 - b0) * 32 * U);
            OPTK_CUDA(cudaMemcpy(tmp.data(), v->d + (size_t)b0 * 32 * U, tmp.size() * sizeof(T), cudaMemcpyDeviceToHost));
            for (long s = s0; s < s0 + n; ++s) for (int u = 0; u < U; ++u) o[(s - s0) * U + u] = tmp[(size_t)(((s >> 5) - b0) * 32 * U + u * 32 + (s & 31))];
        }
        return 0;
    }
    void debug_apply_jtj(const void* host_p, void* host_out) override {
        std::vector<T> tmp((size_t)U * Npad, T(0));
        const T* hp = (const T*)host_p;
        for (long s = 0; s < N; ++s) for (int u = 0; u < U; ++u) tmp[(size_t)V::host_index(u, s, Npad)] = hp[s * U + u];
        // on the plan's stream: a pageable cudaMemcpy on the legacy stream may still be in flight when the kernels
        // below start on the (non-blocking) plan stream
        active_groups();                                    // first: a changed Exclude mask clears the solver vectors
        OPTK_CUDA(cudaMemcpyAsync(p.d, tmp.data(), tmp.size() * sizeof(T), cudaMemcpyHostToDevice, stream));
        OPTK_CUDA(cudaMemsetAsync(Ap.d, 0, tmp.size() * sizeof(T), stream));
        bool saved = lm; lm = false;
        if (use_gather()) build_incidence();
        if (use_seg()) build_segments();
        if (mat_mode) materialize(0);
        apply_A("PCGStep1", "PCGStep1_Graph", p, Ap, S(S_DUMMY), nullptr);
        lm = saved;
        debug_read("Ap", host_out);
    }
};

// functors without ComputedArrays / graphs inherit these no-ops
template <class T>
struct FunctorDefaults {
    static constexpr bool HAS_PRECOMPUTE = false;
    static constexpr bool HAS_CUSTOM_PRECOMPUTE = false; // precompute is a functor-launched kernel (own table layout)
    static constexpr bool HAS_EXCLUDE = false;           // functors with an Exclude() override this
    static constexpr bool HAS_FAST_APPLY = false;        // functors with a specialised J^T J kernel override this
    static constexpr bool FUSE_SLAB = false;             // ... whose fused kernel reads ghost rows from the exchanged p' (slab mode)
    bool fuse_slab_ok() const { return false; }
    static constexpr bool HAS_ACTIVE_ITEMS = false;      // ... whose kernel walks a compacted list of work items (poisson)
    static constexpr bool HAS_FAST_INIT = false;         // ... and with a specialised PCGInit1 (fast_init_kernel())
    static constexpr bool HAS_FAST_COST = false;         // ... and computeCost (fast_cost_kernel())
    static constexpr bool HAS_FAST_MODEL_COST = false;   // ... and computeModelCost (fast_model_cost_kernel())
    static constexpr bool HAS_FAST_PRECOMPUTE = false;   // ... and precompute (fast_precompute(stream, nsm))
    static constexpr bool HAS_DIAG_JTJ = false;          // functors with diagJTJ(): diag(J^T J) alone, without the gradient (PCGComputeCtC)
    static constexpr bool SLAB_OK = true;                // grid functors whose bounds predicates survive a slab cut
    static constexpr bool HAS_GATHER = false;            // graph functors with a per-vertex gather form of J^T J p
    static constexpr bool HAS_SEGSCATTER = false;        // graph functors whose J^T J scatter runs as a segmented sum (graph_generic.cuh)
    static constexpr bool HAS_ROWS_J = false;            // functors that emit their Jacobian rows (materialized-J mode, csr.cuh RowSink)
    static constexpr int MAT_ROWS = 0, MAT_GROWS = 0;    // residual rows per site / per edge
    static constexpr int VEC_UB = 0;                     // solver-vector layout: 0 = SoA, U = blocked, -U = AoS (common.cuh, Vec)
    static constexpr bool FUSE_STEP3 = true;             // form p' = z + beta p inside the J^T J kernel (engine.cuh, FusedDir)
    bool fast_ok() const { return false; }
    size_t fast_smem() const { return 0; }                 // dynamic shared memory of the fast J^T J kernel
    static constexpr int GV = 0;
    int E = 0;
    int o_lo = 0, o_hi = 0x7fffffff;                       // owned range along the slowest dimension (slab mode)
    void set_slab_geometry(long, bool, bool) {}            // global index of local row 0, slab continues below / above
    int unknown_images(void**, int*) const { return 0; }   // grid functors list their AoS unknown images (halo exchange)
    void alloc_temporaries(cudaStream_t) {}
    void free_temporaries() {}
    __device__ void precompute(int, int, int, long) const {}
    template <class V> __device__ void graphJTF(long, V, V) const {}
    template <class V> __device__ T graphJTJ(long, V, V) const { return T(0); }
    __device__ T graphCost(long) const { return T(0); }
    template <class V> __device__ T graphModelCost(long, V) const { return T(0); }
    template <class V> __device__ void graphCtC(long, V, T) const {}
};

// new + init; a throw out of init (optk_fatal with an error handler installed) releases what init had allocated
template <class Fn>
OptK_Plan* build_engine(int is_lm, const unsigned* dims, int timing) {
    auto* e = new Engine<Fn>(is_lm != 0);
    try { e->init(dims, timing != 0); } catch (...) { delete e; throw; }
    return e;
}
template <template <class> class F>
OptK_Plan* make_engine(int is_double, int is_lm, const unsigned* dims, int timing) {
    if (is_double) return build_engine<F<double>>(is_lm, dims, timing);
    return build_engine<F<float>>(is_lm, dims, timing);
}

}  // namespace optk
