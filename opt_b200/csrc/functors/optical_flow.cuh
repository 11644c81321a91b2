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

}  // namespace optk
