// functors/intrinsic.cuh -- intrinsic_image_decomposition
// (examples/intrinsic_image_decomposition/intrinsic_image_decomposition.t:1-32, L_p: API/src/lib.t:112-121)
//   unknown element = (r.xyz, s), U = 4 (log-albedo and log-shading); no Exclude; preconditioner off (default)
//   albedo_o(c)  = [inb(c+o)] w_a q_o(c) (r(c) - r(c+o))          3 channels, o in {(1,0),(-1,0),(0,1),(0,-1)}
//   shading_o(c) = [inb(c+o)] w_s (s(c) - s(c+o))
//   fit(c)       = w_f (r(c) + s(c) - i(c))                       3 channels
//   q_o(c) = sqrt(pow(|r_const(c) - r_const(c+o)| + 1e-7, pNorm - 2)) is a ComputedArray of the constant view
//   r_const of r (the same buffer): refreshed by precompute once per nonlinear iteration, no derivative.
// |a - b| = |b - a| bit for bit, so q_{-o}(c+o) = q_o(c): the reference's four L_p arrays hold two distinct
// images, QX(c) for the pair (c, c+ex) and QY(c) for (c, c+ey); only those two are stored.
// Closed forms (both the residual centred at c and the one centred at c+o touch the unknowns of c):
//   J^T J p (c)[k] = 2 w_a^2 sum_o q_o^2 (p_k(c) - p_k(c+o)) + w_f^2 (p_k(c) + p_s(c))
//   J^T J p (c)[s] = 2 w_s^2 sum_o (p_s(c) - p_s(c+o)) + w_f^2 sum_k (p_k(c) + p_s(c))
#pragma once
#include "math.cuh"

namespace optk {

template <class T>
struct IntrinsicFunctor : FunctorDefaults<T> {
    using Scalar = T;
    static constexpr int U = 4, ND = 2;
    static constexpr bool USE_PRE = false, HAS_PRECOMPUTE = true;
    // slab mode with a halo of one row: every predicate is an in-image test, and the weight images are recomputed
    // locally on the ghost rows from the exchanged r (engine.cuh: precompute runs over every local site)
    int W, H, D;
    T* R; T* S; const T* Im;
    T w_fit, w_alb, w_shd, pnorm;
    T* QX = nullptr; T* QY = nullptr;
    void plan(const unsigned* d) { W = d[0]; H = d[1]; D = 1; }
    void bind(void** pp) {
        w_fit = (T) * (const float*)pp[0]; w_alb = (T) * (const float*)pp[1]; w_shd = (T) * (const float*)pp[2];
        pnorm = *(const T*)pp[3];                                   // Param("pNorm", opt_float, 3)
        R = (T*)pp[4]; Im = (const T*)pp[5]; S = (T*)pp[6];
    }
    void alloc_temporaries(cudaStream_t) {
        const size_t n = (size_t)W * H;
        OPTK_CUDA(cudaMalloc(&QX, 2 * n * sizeof(T)));
        QY = QX + n;
    }
    void free_temporaries() { cudaFree(QX); QX = QY = nullptr; }
    __device__ bool exclude(int, int, int, long) const { return false; }
    __device__ T* xptr(long s, int u) const { return u < 3 ? R + s * 3 + u : S + s; }
    int unknown_images(void** ptrs, int* ch) const { ptrs[0] = R; ch[0] = 3; ptrs[1] = S; ch[1] = 1; return 2; }

    __device__ __forceinline__ T lp_weight(const T* a, const T* b) const {
        const T d0 = a[0] - b[0], d1 = a[1] - b[1], d2 = a[2] - b[2];
        const T dist = sqrt(d0 * d0 + d1 * d1 + d2 * d2);
        return sqrt(pow(dist + T(0.0000001), pnorm - T(2)));
    }
    __device__ void precompute(int x, int y, int, long s) const {
        T a[3]; load3(R + s * 3, a);
        T qx = T(0), qy = T(0);
        if (x + 1 < W) { T b[3]; load3(R + (s + 1) * 3, b); qx = lp_weight(a, b); }
        if (y + 1 < H) { T b[3]; load3(R + (s + W) * 3, b); qy = lp_weight(a, b); }
        QX[s] = qx; QY[s] = qy;
    }
    // fn(neighbour site, q of the pair)
    template <class Fn> __device__ __forceinline__ void neighbours(int x, int y, long s, Fn&& fn) const {
        if (x + 1 < W) fn(s + 1, QX[s]);
        if (x > 0) fn(s - 1, QX[s - 1]);
        if (y + 1 < H) fn(s + W, QY[s]);
        if (y > 0) fn(s - W, QY[s - W]);
    }
    __device__ T cost(int x, int y, int, long s) const {
        T rc[3]; load3(R + s * 3, rc);
        const T sc = S[s];
        T sum = T(0);
        neighbours(x, y, s, [&](long n, T q) {
            T rn[3]; load3(R + n * 3, rn);
#pragma unroll
            for (int k = 0; k < 3; ++k) { const T e = w_alb * (q * (rc[k] - rn[k])); sum += e * e; }
            const T e = w_shd * (sc - S[n]); sum += e * e;
        });
#pragma unroll
        for (int k = 0; k < 3; ++k) { const T e = w_fit * (rc[k] + sc - Im[s * 3 + k]); sum += e * e; }
        return sum;
    }
    __device__ void evalJTF(int x, int y, int, long s, T* g, T* d) const {
        T rc[3]; load3(R + s * 3, rc);
        const T sc = S[s];
        const T wa2 = w_alb * w_alb, ws2 = w_shd * w_shd, wf2 = w_fit * w_fit;
        neighbours(x, y, s, [&](long n, T q) {
            T rn[3]; load3(R + n * 3, rn);
            const T c = T(2) * wa2 * (q * q);
#pragma unroll
            for (int k = 0; k < 3; ++k) { g[k] += c * (rc[k] - rn[k]); d[k] += c; }
            g[3] += T(2) * ws2 * (sc - S[n]); d[3] += T(2) * ws2;
        });
#pragma unroll
        for (int k = 0; k < 3; ++k) {
            const T e = rc[k] + sc - Im[s * 3 + k];
            g[k] += wf2 * e; d[k] += wf2; g[3] += wf2 * e; d[3] += wf2;
        }
    }
    template <class P> __device__ void applyJTJ(int x, int y, int, long s, const P& p, T* out) const {
        const T wa2 = w_alb * w_alb, ws2 = w_shd * w_shd, wf2 = w_fit * w_fit;
        T pc[4], a[4] = {T(0), T(0), T(0), T(0)};
#pragma unroll
        for (int k = 0; k < 4; ++k) pc[k] = p(k, s);
        neighbours(x, y, s, [&](long n, T q) {
            const T c = q * q;
#pragma unroll
            for (int k = 0; k < 3; ++k) a[k] += c * (pc[k] - p(k, n));
            a[3] += pc[3] - p(3, n);
        });
        T fs = T(0);
#pragma unroll
        for (int k = 0; k < 3; ++k) { const T f = wf2 * (pc[k] + pc[3]); out[k] = T(2) * wa2 * a[k] + f; fs += f; }
        out[3] = T(2) * ws2 * a[3] + fs;
    }
    __device__ T modelCost(int x, int y, int, long s, const Vec<T>& dl) const {
        T rc[3], dc[4]; load3(R + s * 3, rc);
        const T sc = S[s];
#pragma unroll
        for (int k = 0; k < 4; ++k) dc[k] = dl(k, s);
        T sum = T(0);
        neighbours(x, y, s, [&](long n, T q) {
            T rn[3]; load3(R + n * 3, rn);
#pragma unroll
            for (int k = 0; k < 3; ++k) {
                const T m = w_alb * (q * (rc[k] - rn[k])) + w_alb * q * (dc[k] - dl(k, n));
                sum += m * m;
            }
            const T m = w_shd * (sc - S[n]) + w_shd * (dc[3] - dl(3, n));
            sum += m * m;
        });
#pragma unroll
        for (int k = 0; k < 3; ++k) {
            const T m = w_fit * (rc[k] + sc - Im[s * 3 + k]) + w_fit * (dc[k] + dc[3]);
            sum += m * m;
        }
        return sum;
    }
};

}  // namespace optk
