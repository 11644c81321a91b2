// functors/optical_flow.cuh -- optical_flow (examples/optical_flow/optical_flow.t:1-20)
//   fit(i)   = w_f ( I_i - Ihat(x + X_i.x, y + X_i.y) )        bilinear sample, OOB taps = 0 (o.t:856-876)
//   reg_o(i) = inb(n) w_r (X_i - X_n)                           (2 channels, 4 offsets)
//   d fit / d X_i = -w_f (Ihat_dx, Ihat_dy) sampled at the same position (o.t:3267-3282)
//   no Exclude, UsePreconditioner(false)
#pragma once
#include "math.cuh"

namespace optk {

template <class T>
struct OpticalFlowFunctor : FunctorDefaults<T> {
    using Scalar = T;
    static constexpr int U = 2, ND = 2;
    static constexpr bool USE_PRE = false;
    int W, H, D;
    T* X; const T* I; const T* Ih; const T* Idx; const T* Idy;
    T w_fit, w_reg;
    // The Jacobian of the fit term, -w_f (Ihat_dx, Ihat_dy) sampled at x + X(x), only changes with X: it is sampled
    // once per nonlinear iteration into two planes (precompute, like the reference's ComputedArrays) and the
    // vectorised J^T J kernel below reads those instead of gathering 8 taps per site and PCG iteration.
    T* GX = nullptr; T* GY = nullptr;
    static constexpr bool HAS_PRECOMPUTE = true, HAS_FAST_APPLY = true;
    void alloc_temporaries(cudaStream_t) {
        const size_t n = (size_t)round_up((long)W * H, 64);
        OPTK_CUDA(cudaMalloc(&GX, 2 * n * sizeof(T)));
        GY = GX + n;
    }
    void free_temporaries() { cudaFree(GX); GX = GY = nullptr; }
    __device__ void precompute(int x, int y, int, long s) const {
        T xi[2]; load2(X + s * 2, xi);
        const T px = (T)x + xi[0], py = (T)y + xi[1];
        GX[s] = sample(Idx, px, py); GY[s] = sample(Idy, px, py);
    }
    bool fast_ok() const { return W % 4 == 0 && getenv("OPTB200_GENERIC_JTJ") == nullptr; }
    long fast_tiles() const { return (long)((W / 4 + 31) / 32) * ((H + 7) / 8); }
    dim3 fast_block() const { return dim3(32, 8, 1); }
    template <bool FUSED, bool LM> static auto fast_kernel();
    void plan(const unsigned* d) { W = d[0]; H = d[1]; D = 1; }
    void bind(void** pp) {
        w_fit = (T) * (const float*)pp[0]; w_reg = (T) * (const float*)pp[1];
        X = (T*)pp[2]; I = (const T*)pp[3]; Ih = (const T*)pp[4]; Idx = (const T*)pp[5]; Idy = (const T*)pp[6];
    }
    __device__ bool exclude(int, int, int, long) const { return false; }
    __device__ T* xptr(long s, int u) const { return X + s * 2 + u; }
    int unknown_images(void** ptrs, int* ch) const { ptrs[0] = X; ch[0] = 2; return 1; }

    template <class Fn> __device__ __forceinline__ void neighbours(int x, int y, long s, Fn&& fn) const {
        if (x + 1 < W) fn(s + 1);
        if (x > 0) fn(s - 1);
        if (y + 1 < H) fn(s + W);
        if (y > 0) fn(s - W);
    }
    __device__ __forceinline__ T get(const T* im, int x, int y) const {
        return (x >= 0 && y >= 0 && x < W && y < H) ? im[x + (long)W * y] : T(0);
    }
    // Image:sample (o.t:868-875): x0=floor, x1=ceil, lerp (1-t) a + t b
    __device__ __forceinline__ T sample(const T* im, T px, T py) const {
        const int x0 = (int)floor(px), x1 = (int)ceil(px), y0 = (int)floor(py), y1 = (int)ceil(py);
        const T xn = px - (T)x0, yn = py - (T)y0;
        const T u = (T(1) - xn) * get(im, x0, y0) + xn * get(im, x1, y0);
        const T b = (T(1) - xn) * get(im, x0, y1) + xn * get(im, x1, y1);
        return (T(1) - yn) * u + yn * b;
    }
    __device__ T cost(int x, int y, int, long s) const {
        T xi[2]; load2(X + s * 2, xi);
        const T px = (T)x + xi[0], py = (T)y + xi[1];
        T r = w_fit * (I[s] - sample(Ih, px, py));
        T sum = r * r;
        neighbours(x, y, s, [&](long n) {
            T xn[2]; load2(X + n * 2, xn);
            T r0 = w_reg * (xi[0] - xn[0]), r1 = w_reg * (xi[1] - xn[1]);
            sum += r0 * r0 + r1 * r1;
        });
        return sum;
    }
    __device__ void evalJTF(int x, int y, int, long s, T* g, T* d) const {
        T xi[2]; load2(X + s * 2, xi);
        const T px = (T)x + xi[0], py = (T)y + xi[1];
        const T wf2 = w_fit * w_fit, wr2 = w_reg * w_reg;
        const T gx = sample(Idx, px, py), gy = sample(Idy, px, py);
        const T e = I[s] - sample(Ih, px, py);
        g[0] += -wf2 * gx * e; g[1] += -wf2 * gy * e;
        d[0] += wf2 * gx * gx; d[1] += wf2 * gy * gy;
        neighbours(x, y, s, [&](long n) {
            T xn[2]; load2(X + n * 2, xn);
            g[0] += T(2) * wr2 * (xi[0] - xn[0]); g[1] += T(2) * wr2 * (xi[1] - xn[1]);
            d[0] += T(2) * wr2; d[1] += T(2) * wr2;
        });
    }
    template <class P> __device__ void applyJTJ(int x, int y, int, long s, const P& p, T* out) const {
        T xi[2]; load2(X + s * 2, xi);
        const T px = (T)x + xi[0], py = (T)y + xi[1];
        const T wf2 = w_fit * w_fit, wr2 = w_reg * w_reg;
        const T gx = sample(Idx, px, py), gy = sample(Idy, px, py);
        const T p0 = p(0, s), p1 = p(1, s);
        const T jp = gx * p0 + gy * p1;
        T a0 = T(0), a1 = T(0);
        neighbours(x, y, s, [&](long n) { a0 += p0 - p(0, n); a1 += p1 - p(1, n); });
        out[0] = wf2 * gx * jp + T(2) * wr2 * a0;
        out[1] = wf2 * gy * jp + T(2) * wr2 * a1;
    }
    __device__ T modelCost(int x, int y, int, long s, const Vec<T>& dl) const {
        T xi[2]; load2(X + s * 2, xi);
        const T px = (T)x + xi[0], py = (T)y + xi[1];
        const T gx = sample(Idx, px, py), gy = sample(Idy, px, py);
        const T d0 = dl(0, s), d1 = dl(1, s);
        T m = w_fit * (I[s] - sample(Ih, px, py)) - w_fit * (gx * d0 + gy * d1);
        T sum = m * m;
        neighbours(x, y, s, [&](long n) {
            T xn[2]; load2(X + n * 2, xn);
            T m0 = w_reg * (xi[0] - xn[0]) + w_reg * (d0 - dl(0, n));
            T m1 = w_reg * (xi[1] - xn[1]) + w_reg * (d1 - dl(1, n));
            sum += m0 * m0 + m1 * m1;
        });
        return sum;
    }
};

// ---------------------------------------------------------------------------------------------
// Vectorised J^T J apply for optical_flow, optionally fused with PCGStep3 (p' = z + beta p, engine.cuh):
//   (J^T J p)(i) = w_f^2 g_i (g_i . p_i) + 2 w_r^2 sum_{n in image} (p_i - p_n),   g_i = (GX, GY)(i) from precompute
// Same structure as k_poisson_apply: a thread owns four consecutive pixels of a row, 128-bit loads for the centre /
// upper / lower quads of both channel planes, lateral neighbours from the adjacent lanes by shuffle.
template <class T, bool FUSED, bool LM>
__global__ void __launch_bounds__(256) k_flow_apply(OpticalFlowFunctor<T> f, Vec<T> z, Vec<T> p, Vec<T> p_new, Vec<T> out, Vec<T> CtC,
                                                    const unsigned char* __restrict__, const double* bNum, const double* aNum,
                                                    double* den, double* zero_a, double* zero_b) {
    const int W = f.W, H = f.H;
    const int tiles_x = (W / 4 + 31) / 32;
    const long ntiles = (long)tiles_x * ((H + 7) / 8);
    const int lane = threadIdx.x;
    const T wf2 = f.w_fit * f.w_fit, wr2 = T(2) * f.w_reg * f.w_reg;
    T beta = T(0);
    if (FUSED) beta = (T)(*bNum) / (T)(*aNum);
    if (blockIdx.x == 0 && linear_tid() == 0) {
        if (zero_a) *zero_a = 0.0;
        if (zero_b) *zero_b = 0.0;
    }
    double acc = 0.0;
    for (long tile = blockIdx.x; tile < ntiles; tile += gridDim.x) {
        const int gx = ((int)(tile % tiles_x) * 32 + lane) * 4;
        const int y = (int)(tile / tiles_x) * 8 + threadIdx.y;
        const bool in_img = gx < W && y < H;
        const bool own = in_img && y >= f.o_lo && y < f.o_hi;
        const long s = (long)y * W + gx;
        const bool up = y > 0, dn = y + 1 < H;
        // Phase 1: issue every load of the tile before anything is consumed (latency-bound kernel, see k_poisson_apply)
        T pv[2][4], u4[2][4], d4[2][4], lmem[2], rmem[2], g0[4] = {T(0), T(0), T(0), T(0)}, g1[4] = {T(0), T(0), T(0), T(0)};
        const bool ledge = own && lane == 0 && gx > 0, redge = own && lane == 31 && gx + 4 < W;
#pragma unroll
        for (int c = 0; c < 2; ++c) {
            const T* pc = p.d + (long)c * p.plane;
            const T* zc = z.d + (long)c * z.plane;
            auto dirq = [&](bool on, long at, T* o) {          // direction at a quad (0 when not loaded)
                T a[4] = {T(0), T(0), T(0), T(0)}, b[4] = {T(0), T(0), T(0), T(0)};
                if (on) { load4(pc + at, a); if (FUSED) load4(zc + at, b); }
#pragma unroll
                for (int j = 0; j < 4; ++j) o[j] = FUSED ? b[j] + beta * a[j] : a[j];
            };
            auto dir1 = [&](bool on, long at) -> T { return on ? (FUSED ? zc[at] + beta * pc[at] : pc[at]) : T(0); };
            dirq(in_img, s, pv[c]);
            dirq(own && up, s - W, u4[c]);
            dirq(own && dn, s + W, d4[c]);
            lmem[c] = dir1(ledge, s - 1);
            rmem[c] = dir1(redge, s + 4);
        }
        if (own) { load4(f.GX + s, g0); load4(f.GY + s, g1); }
        // Phase 2
        T lap[2][4];
#pragma unroll
        for (int c = 0; c < 2; ++c) {
            T lft = __shfl_up_sync(0xffffffffu, pv[c][3], 1);
            T rgt = __shfl_down_sync(0xffffffffu, pv[c][0], 1);
            if (lane == 0) lft = lmem[c];
            if (lane == 31) rgt = rmem[c];
#pragma unroll
            for (int j = 0; j < 4; ++j) {
                const T v = pv[c][j];
                T a = T(0);
                if (gx + j + 1 < W) a += v - (j < 3 ? pv[c][j + 1] : rgt);
                if (gx + j > 0) a += v - (j > 0 ? pv[c][j - 1] : lft);
                if (dn) a += v - d4[c][j];
                if (up) a += v - u4[c][j];
                lap[c][j] = a;
            }
        }
        if (own) {
            T o0[4], o1[4];
#pragma unroll
            for (int j = 0; j < 4; ++j) {
                const T jp = g0[j] * pv[0][j] + g1[j] * pv[1][j];
                o0[j] = wf2 * g0[j] * jp + wr2 * lap[0][j];
                o1[j] = wf2 * g1[j] * jp + wr2 * lap[1][j];
            }
            if (LM) {
                T c0[4], c1[4]; load4(CtC.d + s, c0); load4(CtC.d + CtC.plane + s, c1);
#pragma unroll
                for (int j = 0; j < 4; ++j) { o0[j] += c0[j] * pv[0][j]; o1[j] += c1[j] * pv[1][j]; }
            }
#pragma unroll
            for (int j = 0; j < 4; ++j) acc += (double)(pv[0][j] * o0[j] + pv[1][j] * o1[j]);
            if (FUSED) { st4v(p_new.d + s, pv[0]); st4v(p_new.d + p_new.plane + s, pv[1]); }
            st4v(out.d + s, o0); st4v(out.d + out.plane + s, o1);
        }
    }
    block_accumulate(acc, den);
}

template <class T> template <bool FUSED, bool LM> auto OpticalFlowFunctor<T>::fast_kernel() { return &k_flow_apply<T, FUSED, LM>; }

}  // namespace optk
