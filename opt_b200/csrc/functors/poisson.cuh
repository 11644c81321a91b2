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
    static constexpr bool USE_PRE = false;
    int W, H, D;
    T* X; const T* Tm; const T* M;
    void plan(const unsigned* d) { W = d[0]; H = d[1]; D = 1; }
    void bind(void** pp) { X = (T*)pp[0]; Tm = (const T*)pp[1]; M = (const T*)pp[2]; }
    __device__ bool exclude(int, int, int, long s) const { return M[s] != T(0); }
    __device__ T* xptr(long s, int u) const { return X + s * 4 + u; }
    int unknown_images(void** ptrs, int* ch) const { ptrs[0] = X; ch[0] = 4; return 1; }

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
    __device__ void applyJTJ(int x, int y, int, long s, const Vec<T>& p, T* out) const {
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

}  // namespace optk
