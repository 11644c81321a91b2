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
    // tiled J^T J apply (k_sfs_apply below): J p per residual once per pixel in shared memory, then J^T
    static constexpr bool HAS_FAST_APPLY = true;
    static constexpr int FTW = 32, FTH = 16;
    bool fast_ok() const { return getenv("OPTB200_GENERIC_JTJ") == nullptr; }
    long fast_tiles() const { return (long)((W + FTW - 1) / FTW) * ((H + FTH - 1) / FTH); }
    dim3 fast_block() const { return dim3(32, 8, 1); }
    template <bool FUSED, bool LM> static auto fast_kernel();
    __device__ long fast_tiles_dev() const { return (long)((W + FTW - 1) / FTW) * ((H + FTH - 1) / FTH); }
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

// ---------------------------------------------------------------------------------------------
// Tiled J^T J apply for shape_from_shading.  The per-site functor form above is a gather of gathers: every site
// re-evaluates J p of the 10 shading residuals and 5 regularisation residuals that touch it (~200 loads and a dozen
// divisions per site; 0.47 ms for 2048^2, 5 % of the HBM roofline).  Here a 32 x 16 tile does it in three
// shared-memory stages, each quantity once per pixel:
//   1. p on the tile + 2 halo pixels;  dB(q) = g0(q) p(q) + g1(q) p(q-ex) + g2(q) p(q-ey)
//   2. per residual centre q (tile + 1 halo):  th(q) = k_h(q)^2 (dB(q) - dB(q+ex)),  tv(q) = k_v(q)^2 (dB(q) - dB(q+ey)),
//      ts_ch(q) = k_s(q)^2 (4 a_ch(q) p(q) - sum_o a_ch(q+o) p(q+o))                     (k_r J_r p of every residual)
//   3. J^T:  T(q) = th(q) - th(q-ex) + tv(q) - tv(q-ey);
//      out(s) = g0(s) T(s) + g1(s+ex) T(s+ex) + g2(s+ey) T(s+ey) + sum_ch a_ch(s) (4 ts_ch(s) - sum_o ts_ch(s-o)) + [D_i(s)>0] w_p^2 p(s)
// Global traffic per site: p, three gradient images, two uint8 masks, valid, D_i -- the algorithmic minimum.
template <class T, bool FUSED, bool LM>
__global__ void __launch_bounds__(256) k_sfs_apply(SfsFunctor<T> f, Vec<T> zv, Vec<T> p, Vec<T> p_new, Vec<T> out, Vec<T> CtC,
                                                   const unsigned char* __restrict__, const double* bNum, const double* aNum,
                                                   double* den, double* zero_a, double* zero_b) {
    using F = SfsFunctor<T>;
    constexpr int TW = F::FTW, TH = F::FTH;
    constexpr int PW = TW + 4, PH = TH + 4;            // p / dB region: origin (X0-2, Y0-2)
    constexpr int RW = TW + 2, RH = TH + 2;            // residual-centre region: origin (X0-1, Y0-1)
    __shared__ T sp[PH][PW], sdb[PH][PW], sth[RH][RW], stv[RH][RW], sts[3][RH][RW];
    const int W = f.W, H = f.H;
    const int tiles_x = (W + TW - 1) / TW;
    const long ntiles = f.fast_tiles_dev();
    const int tid = linear_tid();
    const T ifx = T(1) / f.f_x, ify = T(1) / f.f_y;
    if (blockIdx.x == 0 && tid == 0) {
        if (zero_a) *zero_a = 0.0;
        if (zero_b) *zero_b = 0.0;
    }
    double acc = 0.0;
    for (long tile = blockIdx.x; tile < ntiles; tile += gridDim.x) {
        const int X0 = (int)(tile % tiles_x) * TW, Y0 = (int)(tile / tiles_x) * TH;
        // ---- stage 1a: p with a halo of 2 (0 outside the image) ----
        for (int i = tid; i < PW * PH; i += 256) {
            const int lx = i % PW, ly = i / PW, x = X0 - 2 + lx, y = Y0 - 2 + ly;
            sp[ly][lx] = f.inimg(x, y) ? p.d[x + (long)W * y] : T(0);
        }
        __syncthreads();
        // ---- stage 1b: dB on the same region except its first row / column ----
        for (int i = tid; i < PW * PH; i += 256) {
            const int lx = i % PW, ly = i / PW, x = X0 - 2 + lx, y = Y0 - 2 + ly;
            T v = T(0);
            if (lx > 0 && ly > 0 && f.inimg(x, y)) {
                const long q = x + (long)W * y;
                v = f.G0[q] * sp[ly][lx] + f.G1[q] * sp[ly][lx - 1] + f.G2[q] * sp[ly - 1][lx];
            }
            sdb[ly][lx] = v;
        }
        __syncthreads();
        // ---- stage 2: k_r J_r p of the residuals centred on the tile + 1 halo pixel ----
        for (int i = tid; i < RW * RH; i += 256) {
            const int lx = i % RW, ly = i / RW, x = X0 - 1 + lx, y = Y0 - 1 + ly;
            const int px = lx + 1, py = ly + 1;                          // same pixel in the p / dB region
            T th = T(0), tv = T(0), t0 = T(0), t1 = T(0), t2 = T(0);
            if (f.inbexp1(x, y)) {
                const long q = x + (long)W * y;
                const T kh = f.w_g * (T)f.mR[q], kv = f.w_g * (T)f.mC[q];
                const T db = sdb[py][px];
                th = kh * (kh * (db - sdb[py][px + 1]));
                tv = kv * (kv * (db - sdb[py + 1][px]));
                if (f.valid[q] == 1) {
                    const T c = sp[py][px], l = sp[py][px - 1], r = sp[py][px + 1], u = sp[py - 1][px], d = sp[py + 1][px];
                    const T a0 = ((T)x - f.u_x) * ifx, a1 = ((T)y - f.u_y) * ify;
                    const T ws2 = f.w_s * f.w_s;
                    t0 = ws2 * (T(4) * a0 * c - ((a0 - ifx) * l + a0 * u + (a0 + ifx) * r + a0 * d));
                    t1 = ws2 * (T(4) * a1 * c - (a1 * l + (a1 - ify) * u + a1 * r + (a1 + ify) * d));
                    t2 = ws2 * (T(4) * c - (l + u + r + d));
                }
            }
            sth[ly][lx] = th; stv[ly][lx] = tv; sts[0][ly][lx] = t0; sts[1][ly][lx] = t1; sts[2][ly][lx] = t2;
        }
        __syncthreads();
        // ---- stage 3: J^T at the sites of the tile (two rows per thread) ----
#pragma unroll
        for (int rr = 0; rr < 2; ++rr) {
            const int lx = threadIdx.x, ly = threadIdx.y + rr * 8;
            const int x = X0 + lx, y = Y0 + ly;
            if (x < W && y < H) {
                const long s = x + (long)W * y;
                const T di = f.Di[s];
                if (di > T(0)) {                                           // Exclude(D_i <= 0)
                    const int rx = lx + 1, ry = ly + 1;                    // this site in the residual-centre region
                    auto Tq = [&](int qx, int qy) { return sth[qy][qx] - sth[qy][qx - 1] + stv[qy][qx] - stv[qy - 1][qx]; };
                    const T pv = sp[ly + 2][lx + 2];
                    T o = f.w_p * f.w_p * pv + f.G0[s] * Tq(rx, ry);
                    if (x + 1 < W) o += f.G1[s + 1] * Tq(rx + 1, ry);
                    if (y + 1 < H) o += f.G2[s + W] * Tq(rx, ry + 1);
                    T a3[3];
#pragma unroll
                    for (int ch = 0; ch < 3; ++ch)
                        a3[ch] = T(4) * sts[ch][ry][rx] - (sts[ch][ry][rx - 1] + sts[ch][ry][rx + 1] + sts[ch][ry - 1][rx] + sts[ch][ry + 1][rx]);
                    o += ((T)x - f.u_x) * ifx * a3[0] + ((T)y - f.u_y) * ify * a3[1] + a3[2];
                    if (LM) o += CtC.d[s] * pv;
                    out.d[s] = o;
                    acc += (double)(pv * o);
                }
            }
        }
        __syncthreads();
    }
    block_accumulate(acc, den);
}

template <class T> template <bool FUSED, bool LM> auto SfsFunctor<T>::fast_kernel() { return &k_sfs_apply<T, FUSED, LM>; }

}  // namespace optk
