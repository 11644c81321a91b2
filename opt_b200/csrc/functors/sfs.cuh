// functors/sfs.cuh -- shape_from_shading (examples/shape_from_shading/shape_from_shading.t:1-90)
//   unknown X = refined depth (1 channel); Exclude(D_i <= 0); preconditioner off (no UsePreconditioner()).
//   ComputedArrays (refreshed by precompute, stale inside a PCG solve by design, o.t:1551-1563, 3131-3153):
//     BI(c)   = [inbexp1(c) dv(c-ex) dv(c) dv(c-ey)] ( B(normal(X(c), X(c-ex), X(c-ey))) - I(c) )
//     g0,g1,g2(c) = dBI(c)/dX(c), dBI(c)/dX(c-ex), dBI(c)/dX(c-ey)           (gradient images)
//     valid(c) = depth valid at c and its 4 neighbours, |X(c)-X(n)| < 0.01, inbexp1(c)
//   residuals at centre c (a(c) = ((i-u_x)/f_x, (j-u_y)/f_y, 1) is the back-projection ray):
//     r_p = [D_i(c) > 0] w_p (X(c) - D_i(c))
//     r_h = [inbexp1(c)] w_g maskR(c) (BI(c) - BI(c+ex)),   r_v likewise with maskC, c+ey
//     r_s[ch] = [valid(c) == 1, inbexp1(c)] w_s ( 4 a_ch(c) X(c) - sum_o a_ch(c+o) X(c+o) )
// Closed forms.  With dB(q) = g0(q) p(q) + g1(q) p(q-ex) + g2(q) p(q-ey):
//     J_h(c) p = k_h(c) (dB(c) - dB(c+ex)),   J_v(c) p = k_v(c) (dB(c) - dB(c+ey)),   k = [inbexp1] w_g mask
//     T(q) = k_h(q) J_h(q)p - k_h(q-ex) J_h(q-ex)p + k_v(q) J_v(q)p - k_v(q-ey) J_v(q-ey)p      (= sum_r dr/dBI(q) J_r p)
//     (J^T J p)(s) = g0(s) T(s) + g1(s+ex) T(s+ex) + g2(s+ey) T(s+ey)
//                  + sum_ch a_ch(s) ( 4 k_s(s) J_s(s)p[ch] - sum_o k_s(s-o) J_s(s-o)p[ch] ) + [D_i(s)>0] w_p^2 p(s)
// J^T F has the same shape with J_r p replaced by r.
#pragma once
#include "math.cuh"

namespace optk {

template <class T>
struct SfsFunctor : FunctorDefaults<T> {
    using Scalar = T;
    static constexpr int U = 1, ND = 2;
    static constexpr bool USE_PRE = false, HAS_EXCLUDE = true, HAS_PRECOMPUTE = true, FUSE_STEP3 = false;
    static constexpr bool SLAB_OK = false;     // the expanded-bounds predicates would need global row numbers
    int W, H, D;
    T* X; const T* Di; const T* Im; const unsigned char* mR; const unsigned char* mC;
    T w_p, w_s, w_g, f_x, f_y, u_x, u_y, L[9];
    T* BI = nullptr; T* G0 = nullptr; T* G1 = nullptr; T* G2 = nullptr; unsigned char* valid = nullptr;
    void plan(const unsigned* d) { W = d[0]; H = d[1]; D = 1; }
    void bind(void** pp) {
        // sqrt(Param(...)): the float parameter is cast to opt_float, then the square root is taken (o.t:2430)
        w_p = std::sqrt((T) * (const float*)pp[0]); w_s = std::sqrt((T) * (const float*)pp[1]); w_g = std::sqrt((T) * (const float*)pp[2]);
        f_x = (T) * (const float*)pp[3]; f_y = (T) * (const float*)pp[4]; u_x = (T) * (const float*)pp[5]; u_y = (T) * (const float*)pp[6];
        for (int i = 0; i < 9; ++i) L[i] = (T) * (const float*)pp[7 + i];
        X = (T*)pp[16]; Di = (const T*)pp[17]; Im = (const T*)pp[18];
        mR = (const unsigned char*)pp[19]; mC = (const unsigned char*)pp[20];
    }
    void alloc_temporaries(cudaStream_t) {
        const size_t n = (size_t)W * H;
        OPTK_CUDA(cudaMalloc(&BI, 4 * n * sizeof(T)));
        G0 = BI + n; G1 = G0 + n; G2 = G1 + n;
        OPTK_CUDA(cudaMalloc(&valid, n));
    }
    void free_temporaries() { cudaFree(BI); cudaFree(valid); BI = nullptr; valid = nullptr; }
    int unknown_images(void** ptrs, int* ch) const { ptrs[0] = X; ch[0] = 1; return 1; }
    __device__ T* xptr(long s, int) const { return X + s; }

    __device__ __forceinline__ bool inimg(int x, int y) const { return x >= 0 && y >= 0 && x < W && y < H; }
    __device__ __forceinline__ bool inbexp1(int x, int y) const { return x >= 1 && y >= 1 && x + 1 < W && y + 1 < H; }
    template <class A> __device__ __forceinline__ A get(const A* im, int x, int y) const { return inimg(x, y) ? im[x + (long)W * y] : A(0); }
    __device__ __forceinline__ bool dv(int x, int y) const { return get(Di, x, y) > T(0); }
    __device__ bool exclude(int x, int y, int, long) const { return !dv(x, y); }
    __device__ __forceinline__ T ax(int x) const { return ((T)x - u_x) / f_x; }
    __device__ __forceinline__ T ay(int y) const { return ((T)y - u_y) / f_y; }
    // residual weights including their bounds predicates (0 when the instance does not exist)
    __device__ __forceinline__ T kh(int x, int y) const { return inbexp1(x, y) ? w_g * (T)mR[x + (long)W * y] : T(0); }
    __device__ __forceinline__ T kv(int x, int y) const { return inbexp1(x, y) ? w_g * (T)mC[x + (long)W * y] : T(0); }
    __device__ __forceinline__ T ks(int x, int y) const { return (inbexp1(x, y) && valid[x + (long)W * y] == 1) ? w_s : T(0); }

    // ---- precompute: B_I, its three gradient images, valid -------------------------------------------
    __device__ void precompute(int x, int y, int, long s) const {
        const T d00 = X[s], dm0 = get(X, x - 1, y), d0m = get(X, x, y - 1);
        const bool ok = inbexp1(x, y) && dv(x - 1, y) && dv(x, y) && dv(x, y - 1);
        T b = T(0), g[3] = {T(0), T(0), T(0)};
        if (ok) {
            const T ify = T(1) / f_y, ifx = T(1) / f_x, cx = (u_x - (T)x) / f_x, cy = (u_y - (T)y) / f_y, ifxy = T(1) / (f_x * f_y);
            // n and its derivatives wrt (d00, dm0, d0m)
            T n[3], dn[3][3];
            n[0] = d0m * (d00 - dm0) * ify;
            dn[0][0] = d0m * ify; dn[0][1] = -d0m * ify; dn[0][2] = (d00 - dm0) * ify;
            n[1] = dm0 * (d00 - d0m) * ifx;
            dn[1][0] = dm0 * ifx; dn[1][1] = (d00 - d0m) * ifx; dn[1][2] = -dm0 * ifx;
            n[2] = n[0] * cx + n[1] * cy - dm0 * d0m * ifxy;
            dn[2][0] = dn[0][0] * cx + dn[1][0] * cy;
            dn[2][1] = dn[0][1] * cx + dn[1][1] * cy - d0m * ifxy;
            dn[2][2] = dn[0][2] * cx + dn[1][2] * cy - dm0 * ifxy;
            const T sq = n[0] * n[0] + n[1] * n[1] + n[2] * n[2];
            const bool pos = sq > T(0);
            const T inv = pos ? T(1) / sqrt(sq) : T(1);
            const T N0 = inv * n[0], N1 = inv * n[1], N2 = inv * n[2];
            b = L[0] + L[1] * N1 + L[2] * N2 + L[3] * N0 + L[4] * N0 * N1 + L[5] * N1 * N2 +
                L[6] * (-(N0 * N0) - N1 * N1 + T(2) * N2 * N2) + L[7] * N2 * N0 + L[8] * (N0 * N0 - N1 * N1);
            const T bN0 = L[3] + L[4] * N1 - T(2) * L[6] * N0 + L[7] * N2 + T(2) * L[8] * N0;
            const T bN1 = L[1] + L[4] * N0 + L[5] * N2 - T(2) * L[6] * N1 - T(2) * L[8] * N1;
            const T bN2 = L[2] + L[5] * N1 + T(4) * L[6] * N2 + L[7] * N0;
#pragma unroll
            for (int v = 0; v < 3; ++v) {
                // dN/dv = inv (dn/dv - N (N . dn/dv)) when the normal was normalised, dn/dv otherwise
                const T nd = pos ? N0 * dn[0][v] + N1 * dn[1][v] + N2 * dn[2][v] : T(0);
                g[v] = inv * (bN0 * (dn[0][v] - N0 * nd) + bN1 * (dn[1][v] - N1 * nd) + bN2 * (dn[2][v] - N2 * nd));
            }
            b -= Im[s] * T(0.5) + T(0.25) * (get(Im, x - 1, y) + get(Im, x, y - 1));
        }
        BI[s] = b; G0[s] = g[0]; G1[s] = g[1]; G2[s] = g[2];
        auto cont = [&](int ox, int oy) { return fabs(d00 - get(X, x + ox, y + oy)) < T(0.01); };
        const bool v = dv(x, y) && dv(x, y - 1) && dv(x, y + 1) && dv(x - 1, y) && dv(x + 1, y) &&
                       cont(0, -1) && cont(0, 1) && cont(-1, 0) && cont(1, 0) && inbexp1(x, y);
        valid[s] = v ? 1 : 0;
    }

    // ---- building blocks -------------------------------------------------------------------------
    // chain rule through the gradient images at pixel q for a field f (p or delta), out-of-image = 0
    template <class Fld> __device__ __forceinline__ T dB(const Fld& f, int x, int y) const {
        if (!inimg(x, y)) return T(0);
        const long q = x + (long)W * y;
        return G0[q] * f(x, y) + G1[q] * f(x - 1, y) + G2[q] * f(x, y - 1);
    }
    template <class Fld> __device__ __forceinline__ T Jh(const Fld& f, int x, int y) const {   // J of r_h centred at (x,y) times f
        const T k = kh(x, y);
        return k != T(0) ? k * (dB(f, x, y) - dB(f, x + 1, y)) : T(0);
    }
    template <class Fld> __device__ __forceinline__ T Jv(const Fld& f, int x, int y) const {
        const T k = kv(x, y);
        return k != T(0) ? k * (dB(f, x, y) - dB(f, x, y + 1)) : T(0);
    }
    __device__ __forceinline__ T rh(int x, int y) const {
        const T k = kh(x, y);
        return k != T(0) ? k * (BI[x + (long)W * y] - BI[x + 1 + (long)W * y]) : T(0);
    }
    __device__ __forceinline__ T rv(int x, int y) const {
        const T k = kv(x, y);
        return k != T(0) ? k * (BI[x + (long)W * y] - BI[x + (long)W * (y + 1)]) : T(0);
    }
    // the linear regularisation operator of centre (x,y) applied to a field: k_s (4 a f - sum_o a_o f_o), 3 channels
    template <class Fld> __device__ __forceinline__ void Js(const Fld& f, int x, int y, T* o) const {
        const T k = inimg(x, y) ? ks(x, y) : T(0);
        o[0] = o[1] = o[2] = T(0);
        if (k == T(0)) return;
        const T c = f(x, y), l = f(x - 1, y), r = f(x + 1, y), u = f(x, y - 1), d = f(x, y + 1);
        const T a0 = ax(x), a1 = ay(y);
        o[0] = k * (T(4) * a0 * c - (ax(x - 1) * l + a0 * u + ax(x + 1) * r + a0 * d));
        o[1] = k * (T(4) * a1 * c - (a1 * l + ay(y - 1) * u + a1 * r + ay(y + 1) * d));
        o[2] = k * (T(4) * c - (l + u + r + d));
    }
    // sum over the shading residuals of d r / d BI(q) times a per-residual quantity (J_r p or r itself)
    template <class FH, class FV> __device__ __forceinline__ T gatherT(FH&& h, FV&& v, int x, int y) const {
        if (!inimg(x, y)) return T(0);
        return kh(x, y) * h(x, y) - kh(x - 1, y) * h(x - 1, y) + kv(x, y) * v(x, y) - kv(x, y - 1) * v(x, y - 1);
    }
    struct XField {                       // the unknown image as a field (out-of-image loads are 0, o.t:856-862)
        const SfsFunctor* f;
        __device__ __forceinline__ T operator()(int x, int y) const { return f->get(f->X, x, y); }
    };
    template <class P> struct PField {    // a solver vector as a field
        const SfsFunctor* f; const P* p;
        __device__ __forceinline__ T operator()(int x, int y) const { return f->inimg(x, y) ? (*p)(0, x + (long)f->W * y) : T(0); }
    };

    // ---- per-site functions ----------------------------------------------------------------------
    __device__ T cost(int x, int y, int, long s) const {
        T sum = T(0);
        if (Di[s] > T(0)) { const T r = w_p * (X[s] - Di[s]); sum += r * r; }
        const T a = rh(x, y), b = rv(x, y);
        sum += a * a + b * b;
        T e[3]; Js(XField{this}, x, y, e);
        return sum + e[0] * e[0] + e[1] * e[1] + e[2] * e[2];
    }
    __device__ void evalJTF(int x, int y, int, long s, T* g, T* d) const {
        T gg = T(0), dd = T(0);
        if (Di[s] > T(0)) { gg += w_p * w_p * (X[s] - Di[s]); dd += w_p * w_p; }
        // d BI(q) / d X(s): nonzero for q = s, s+ex, s+ey
        const T b0 = G0[s], b1 = get(G1, x + 1, y), b2 = get(G2, x, y + 1);
        auto dBI = [&](int qx, int qy) -> T {
            return (qx == x && qy == y) ? b0 : (qx == x + 1 && qy == y) ? b1 : (qx == x && qy == y + 1) ? b2 : T(0);
        };
        const int hx[5] = {x - 1, x, x + 1, x - 1, x}, hy[5] = {y, y, y, y + 1, y + 1};       // centres of r_h touching X(s)
#pragma unroll
        for (int i = 0; i < 5; ++i) {
            const T k = inimg(hx[i], hy[i]) ? kh(hx[i], hy[i]) : T(0);
            if (k != T(0)) { const T dr = k * (dBI(hx[i], hy[i]) - dBI(hx[i] + 1, hy[i])); gg += dr * rh(hx[i], hy[i]); dd += dr * dr; }
        }
        const int vx[5] = {x, x, x + 1, x + 1, x}, vy[5] = {y - 1, y, y - 1, y, y + 1};       // centres of r_v touching X(s)
#pragma unroll
        for (int i = 0; i < 5; ++i) {
            const T k = inimg(vx[i], vy[i]) ? kv(vx[i], vy[i]) : T(0);
            if (k != T(0)) { const T dr = k * (dBI(vx[i], vy[i]) - dBI(vx[i], vy[i] + 1)); gg += dr * rv(vx[i], vy[i]); dd += dr * dr; }
        }
        // regularisation: d r_s(c)[ch] / d X(s) = k_s(c) a_ch(s) (4 [c == s] - [c is a 4-neighbour of s])
        const T a0 = ax(x), a1 = ay(y), an = a0 * a0 + a1 * a1 + T(1);
        XField xf{this};
        T e[3]; Js(xf, x, y, e);
        const T k0 = ks(x, y);
        T acc[3] = {T(4) * k0 * e[0], T(4) * k0 * e[1], T(4) * k0 * e[2]};
        T k2 = T(16) * k0 * k0;
        const int ox[4] = {1, -1, 0, 0}, oy[4] = {0, 0, 1, -1};
#pragma unroll
        for (int i = 0; i < 4; ++i) {
            const int cx = x - ox[i], cy = y - oy[i];
            if (!inimg(cx, cy)) continue;
            const T k = ks(cx, cy);
            Js(xf, cx, cy, e);
            acc[0] -= k * e[0]; acc[1] -= k * e[1]; acc[2] -= k * e[2];
            k2 += k * k;
        }
        gg += a0 * acc[0] + a1 * acc[1] + acc[2];
        dd += an * k2;
        g[0] += gg; d[0] += dd;
    }
    template <class P> __device__ void applyJTJ(int x, int y, int, long s, const P& p, T* out) const {
        PField<P> pf{this, &p};
        T o = Di[s] > T(0) ? w_p * w_p * p(0, s) : T(0);
        auto h = [&](int qx, int qy) { return Jh(pf, qx, qy); };
        auto v = [&](int qx, int qy) { return Jv(pf, qx, qy); };
        o += G0[s] * gatherT(h, v, x, y) + get(G1, x + 1, y) * gatherT(h, v, x + 1, y) + get(G2, x, y + 1) * gatherT(h, v, x, y + 1);
        T e[3]; Js(pf, x, y, e);
        const T k0 = ks(x, y);
        T acc[3] = {T(4) * k0 * e[0], T(4) * k0 * e[1], T(4) * k0 * e[2]};
        const int ox[4] = {1, -1, 0, 0}, oy[4] = {0, 0, 1, -1};
#pragma unroll
        for (int i = 0; i < 4; ++i) {
            const int cx = x - ox[i], cy = y - oy[i];
            if (!inimg(cx, cy)) continue;
            const T k = ks(cx, cy);
            Js(pf, cx, cy, e);
            acc[0] -= k * e[0]; acc[1] -= k * e[1]; acc[2] -= k * e[2];
        }
        out[0] = o + ax(x) * acc[0] + ay(y) * acc[1] + acc[2];
    }
    __device__ T modelCost(int x, int y, int, long s, const Vec<T>& dl) const {
        PField<Vec<T>> df{this, &dl};
        T sum = T(0);
        if (Di[s] > T(0)) { const T m = w_p * (X[s] - Di[s]) + w_p * dl(0, s); sum += m * m; }
        const T a = rh(x, y) + Jh(df, x, y), b = rv(x, y) + Jv(df, x, y);
        sum += a * a + b * b;
        T e[3], je[3]; Js(XField{this}, x, y, e); Js(df, x, y, je);
#pragma unroll
        for (int k = 0; k < 3; ++k) { const T m = e[k] + je[k]; sum += m * m; }
        return sum;
    }
};

}  // namespace optk
