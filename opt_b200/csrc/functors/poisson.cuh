// functors/poisson.cuh -- poisson_image_editing (examples/poisson_image_editing/poisson_image_editing.t:1-13)
//   r_o(i) = inb(i+o) * [ (X_i - X_{i+o}) - (T_i - T_{i+o}) ],  o in {(1,0),(-1,0),(0,1),(0,-1)}, 4 channels
//   Exclude(M != 0), UsePreconditioner(false).
// Closed forms (both the instance centred at i and the one centred at i+o touch X_i):
//   J^T F(i) = 2 sum_o inb(n) e_o(i)      diag = 2 #inb      J^T J p (i) = 2 sum_o inb(n) (p_i - p_n)
#pragma once
#include "math.cuh"

namespace optk {

template <class T>
struct PoissonFunctor : FunctorDefaults<T> {
    using Scalar = T;
    static constexpr int U = 4, ND = 2;
    static constexpr bool HAS_EXCLUDE = true;
    static constexpr bool USE_PRE = false;
    int W, H, D;
    T* X; const T* Tm; const T* M;
    void plan(const unsigned* d) { W = d[0]; H = d[1]; D = 1; }
    void bind(void** pp) { X = (T*)pp[0]; Tm = (const T*)pp[1]; M = (const T*)pp[2]; }
    __device__ bool exclude(int, int, int, long s) const { return M[s] != T(0); }
    __device__ T* xptr(long s, int u) const { return X + s * 4 + u; }
    int unknown_images(void** ptrs, int* ch) const { ptrs[0] = X; ch[0] = 4; return 1; }
    // vectorised / fused J^T J apply (k_poisson_apply below); needs rows that are a multiple of 4 pixels
    static constexpr bool HAS_FAST_APPLY = true;
    bool fast_ok() const { return W % 4 == 0; }
    long fast_tiles() const { return (long)((W / 4 + 31) / 32) * ((H + 7) / 8); }
    dim3 fast_block() const { return dim3(32, 8, 1); }
    template <bool FUSED, bool LM> static auto fast_kernel();

    template <class Fn> __device__ __forceinline__ void neighbours(int x, int y, long s, Fn&& fn) const {
        if (x + 1 < W) fn(s + 1);
        if (x > 0) fn(s - 1);
        if (y + 1 < H) fn(s + W);
        if (y > 0) fn(s - W);
    }
    __device__ T cost(int x, int y, int, long s) const {
        T xi[4], ti[4]; load4(X + s * 4, xi); load4(Tm + s * 4, ti);
        T sum = T(0);
        neighbours(x, y, s, [&](long n) {
            T xn[4], tn[4]; load4(X + n * 4, xn); load4(Tm + n * 4, tn);
#pragma unroll
            for (int c = 0; c < 4; ++c) { T e = (xi[c] - xn[c]) - (ti[c] - tn[c]); sum += e * e; }
        });
        return sum;
    }
    __device__ void evalJTF(int x, int y, int, long s, T* g, T* d) const {
        T xi[4], ti[4]; load4(X + s * 4, xi); load4(Tm + s * 4, ti);
        T cnt = T(0);
        neighbours(x, y, s, [&](long n) {
            T xn[4], tn[4]; load4(X + n * 4, xn); load4(Tm + n * 4, tn);
#pragma unroll
            for (int c = 0; c < 4; ++c) g[c] += (xi[c] - xn[c]) - (ti[c] - tn[c]);
            cnt += T(1);
        });
#pragma unroll
        for (int c = 0; c < 4; ++c) { g[c] = T(2) * g[c]; d[c] = T(2) * cnt; }
    }
    template <class P> __device__ void applyJTJ(int x, int y, int, long s, const P& p, T* out) const {
        T pi[4];
#pragma unroll
        for (int c = 0; c < 4; ++c) { pi[c] = p(c, s); out[c] = T(0); }
        neighbours(x, y, s, [&](long n) {
#pragma unroll
            for (int c = 0; c < 4; ++c) out[c] += pi[c] - p(c, n);
        });
#pragma unroll
        for (int c = 0; c < 4; ++c) out[c] = T(2) * out[c];
    }
    __device__ T modelCost(int x, int y, int, long s, const Vec<T>& dl) const {
        T xi[4], ti[4], di[4]; load4(X + s * 4, xi); load4(Tm + s * 4, ti);
#pragma unroll
        for (int c = 0; c < 4; ++c) di[c] = dl(c, s);
        T sum = T(0);
        neighbours(x, y, s, [&](long n) {
            T xn[4], tn[4]; load4(X + n * 4, xn); load4(Tm + n * 4, tn);
#pragma unroll
            for (int c = 0; c < 4; ++c) { T m = ((xi[c] - xn[c]) - (ti[c] - tn[c])) + (di[c] - dl(c, n)); sum += m * m; }
        });
        return sum;
    }
};


// ---------------------------------------------------------------------------------------------
// Vectorised J^T J apply for poisson (optionally fused with PCGStep3, see engine.cuh).
// J^T J is four independent 5-point Laplacians (one per channel plane), so each thread owns FOUR
// consecutive pixels of a row: per plane it issues 128-bit loads for the centre / upper / lower
// quads and gets the left/right neighbours from the adjacent lanes by warp shuffle (only lanes 0 and
// 31 touch memory for them).  ~2 LDG.128 per pixel instead of 25 scalar loads: the generic kernel
// was LSU-issue bound at 2.3 TB/s (profiles/r1_a).  Quads without unknowns are skipped via active4.
template <class T> __device__ __forceinline__ void ld4v(const T* p, T* o) { load4(p, o); }

template <class T, bool FUSED, bool LM>
__global__ void __launch_bounds__(256) k_poisson_apply(PoissonFunctor<T> f, Vec<T> z, Vec<T> p, Vec<T> p_new, Vec<T> out,
                                                       Vec<T> CtC, const unsigned char* __restrict__ active4,
                                                       const double* bNum, const double* aNum, double* den,
                                                       double* zero_a, double* zero_b) {
    const int W = f.W, H = f.H;
    const int tiles_x = (W / 4 + 31) / 32;
    const long ntiles = (long)tiles_x * ((H + 7) / 8);
    const int lane = threadIdx.x;
    T beta = T(0);
    if (FUSED) beta = (T)(*bNum) / (T)(*aNum);
    if (blockIdx.x == 0 && linear_tid() == 0) {
        if (zero_a) *zero_a = 0.0;
        if (zero_b) *zero_b = 0.0;
    }
    double acc = 0.0;
    for (long tile = blockIdx.x; tile < ntiles; tile += gridDim.x) {      // persistent CTAs over 128x8-pixel tiles
    const int gx = ((int)(tile % tiles_x) * 32 + threadIdx.x) * 4;
    const int y = (int)(tile / tiles_x) * 8 + threadIdx.y;
    const bool in_img = gx < W && y < H;
    const bool own = in_img && y >= f.o_lo && y < f.o_hi;
    const long s = (long)y * W + gx;
    // a quad takes part in loads if it is inside the image and holds at least one unknown
    const bool act = in_img && active4[s >> 2] != 0;
    bool ex[4] = {true, true, true, true};
    if (act && own) {
        T m[4]; ld4v(f.M + s, m);
#pragma unroll
        for (int j = 0; j < 4; ++j) ex[j] = m[j] != T(0);
    }
    const bool up = y > 0, dn = y + 1 < H;
#pragma unroll
    for (int c = 0; c < 4; ++c) {
        const T* pc = p.d + (long)c * p.plane;
        const T* zc = z.d + (long)c * z.plane;
        auto dirq = [&](long at, T* o) {                 // direction at a quad
            T a[4]; ld4v(pc + at, a);
            if (FUSED) { T b[4]; ld4v(zc + at, b);
#pragma unroll
                for (int j = 0; j < 4; ++j) o[j] = b[j] + beta * a[j];
            } else {
#pragma unroll
                for (int j = 0; j < 4; ++j) o[j] = a[j];
            }
        };
        auto dir1 = [&](long at) -> T { return FUSED ? zc[at] + beta * pc[at] : pc[at]; };
        // all loads of the channel are issued before the shuffles consume the centre quad: one exposed memory
        // latency per channel instead of two (the kernel is latency-bound, ncu r1i)
        T ctr[4] = {T(0), T(0), T(0), T(0)}, u4[4] = {T(0), T(0), T(0), T(0)}, d4[4] = {T(0), T(0), T(0), T(0)};
        T lmem = T(0), rmem = T(0);
        if (act) dirq(s, ctr);
        if (act && own) {
            if (up) dirq(s - W, u4);
            if (dn) dirq(s + W, d4);
            if (lane == 0 && gx > 0) lmem = dir1(s - 1);
            if (lane == 31 && gx + 4 < W) rmem = dir1(s + 4);
        }
        // lateral neighbours: from the adjacent lanes, from memory at the warp's ends
        T lft = __shfl_up_sync(0xffffffffu, ctr[3], 1);
        T rgt = __shfl_down_sync(0xffffffffu, ctr[0], 1);
        if (act && own) {
            if (lane == 0) lft = lmem;
            if (lane == 31) rgt = rmem;
            T o[4], pn[4];
#pragma unroll
            for (int j = 0; j < 4; ++j) {
                const T v = ctr[j];
                T a = T(0);
                if (gx + j + 1 < W) a += v - (j < 3 ? ctr[j + 1] : rgt);      // {1,0}
                if (gx + j > 0) a += v - (j > 0 ? ctr[j - 1] : lft);          // {-1,0}
                if (dn) a += v - d4[j];                                       // {0,1}
                if (up) a += v - u4[j];                                       // {0,-1}
                a = T(2) * a;
                pn[j] = ex[j] ? T(0) : v;
                o[j] = ex[j] ? T(0) : a;
            }
            if (LM) {
                T cc[4]; ld4v(CtC.d + (long)c * CtC.plane + s, cc);
#pragma unroll
                for (int j = 0; j < 4; ++j) if (!ex[j]) o[j] += cc[j] * pn[j];
            }
#pragma unroll
            for (int j = 0; j < 4; ++j) acc += (double)(pn[j] * o[j]);
            if (FUSED) st4v(p_new.d + (long)c * p_new.plane + s, pn);
            st4v(out.d + (long)c * out.plane + s, o);
        }
    }
    }   // tile loop
    block_accumulate(acc, den);
}

template <class T> template <bool FUSED, bool LM> auto PoissonFunctor<T>::fast_kernel() { return &k_poisson_apply<T, FUSED, LM>; }

}  // namespace optk
