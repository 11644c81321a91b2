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
    static constexpr bool USE_PRE = false, HAS_EXCLUDE = true, HAS_PRECOMPUTE = true;
    // slab mode: the expanded-bounds predicate and the back-projection ray need GLOBAL row numbers; the engine passes the
    // global index of local row 0 and whether the slab continues above / below (set_slab_geometry), halo = 2 rows
    long y_off = 0;
    bool lo_open = false, hi_open = false;
    void set_slab_geometry(long off, bool lo, bool hi) { y_off = off; lo_open = lo; hi_open = hi; }
    __host__ __device__ int own_lo() const { return this->o_lo; }
    __host__ __device__ int own_hi() const { return this->o_hi < H ? this->o_hi : H; }
    // marching J^T J apply (k_sfs_march below): a warp owns a strip of STRIP written pixels and a band of bandR rows
    static constexpr bool HAS_FAST_APPLY = true;
    static constexpr int STRIP = 120;
    int bandR = 16;
    bool fast_ok() const { return W % 4 == 0 && getenv("OPTB200_GENERIC_JTJ") == nullptr; }
    __host__ __device__ int strips() const { return (W + STRIP - 1) / STRIP; }
    long fast_tiles() const { return ((long)strips() * ((own_hi() - own_lo() + bandR - 1) / bandR) + 7) / 8; }
    dim3 fast_block() const { return dim3(32, 8, 1); }
    template <bool FUSED, bool LM> static auto fast_kernel();
    static constexpr bool HAS_FAST_INIT = true;        // PCGInit1 by the same pipeline (k_sfs_init)
    static auto fast_init_kernel();
    static constexpr bool HAS_FAST_COST = true;        // computeCost likewise (k_sfs_cost)
    static auto fast_cost_kernel();
    static constexpr bool HAS_FAST_MODEL_COST = true;  // computeModelCost likewise (k_sfs_model_cost, MODE 4 of the march)
    static constexpr bool HAS_DIAG_JTJ = true;         // PCGComputeCtC without the gradient half of evalJTF (diagJTJ)
    static auto fast_model_cost_kernel();
    static constexpr bool HAS_FAST_PRECOMPUTE = true;  // one thread per quad of a row (k_sfs_precompute4)
    void fast_precompute(cudaStream_t stream, int nsm) const;
    size_t fast_smem() const;
    int W, H, D;
    T* X; const T* Di; const T* Im; const unsigned char* mR; const unsigned char* mC;
    T w_p, w_s, w_g, f_x, f_y, u_x, u_y, L[9];
    T* BI = nullptr; T* G0 = nullptr; T* G1 = nullptr; T* G2 = nullptr; unsigned char* valid = nullptr;
    void plan(const unsigned* d) {
        W = d[0]; H = d[1]; D = 1;
        // all (strip, band) items resident at once: 8 warps per CTA, 2 CTAs per SM in fp32, 1 in fp64
        int dev = 0, nsm = 148;
        cudaGetDevice(&dev); cudaDeviceGetAttribute(&nsm, cudaDevAttrMultiProcessorCount, dev);
        const long slots = (long)nsm * 8 * (sizeof(T) == 4 ? 2 : 1);
        bandR = (int)std::min<long>(256, std::max<long>(8, ((long)H * strips() + slots - 1) / slots));
        if (const char* e = getenv("OPTB200_SFS_BAND")) bandR = std::max(4, atoi(e));      // measurement hook: band height sweep
    }
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
    __device__ __forceinline__ bool row_inb(int y) const { return (y >= 1 || lo_open) && (y + 1 < H || hi_open); }
    __device__ __forceinline__ bool inbexp1(int x, int y) const { return x >= 1 && x + 1 < W && row_inb(y); }
    template <class A> __device__ __forceinline__ A get(const A* im, int x, int y) const { return inimg(x, y) ? im[x + (long)W * y] : A(0); }
    __device__ __forceinline__ bool dv(int x, int y) const { return get(Di, x, y) > T(0); }
    __device__ bool exclude(int x, int y, int, long) const { return !dv(x, y); }
    __device__ __forceinline__ T ax(int x) const { return ((T)x - u_x) / f_x; }
    __device__ __forceinline__ T ay(int y) const { return ((T)(y + y_off) - u_y) / f_y; }
    // residual weights including their bounds predicates (0 when the instance does not exist)
    __device__ __forceinline__ T kh(int x, int y) const { return inbexp1(x, y) ? w_g * (T)mR[x + (long)W * y] : T(0); }
    __device__ __forceinline__ T kv(int x, int y) const { return inbexp1(x, y) ? w_g * (T)mC[x + (long)W * y] : T(0); }
    __device__ __forceinline__ T ks(int x, int y) const { return (inbexp1(x, y) && valid[x + (long)W * y] == 1) ? w_s : T(0); }

    // ---- precompute: B_I, its three gradient images, valid -------------------------------------------
    // B(normal(d00, dm0, d0m)) - shading target and its derivatives wrt (d00, dm0, d0m) at pixel (x, y)
    __device__ __forceinline__ void shade(int x, int y, T d00, T dm0, T d0m, T target, T& b, T* g) const {
        const T ify = T(1) / f_y, ifx = T(1) / f_x, cx = (u_x - (T)x) / f_x, cy = (u_y - (T)(y + y_off)) / f_y, ifxy = T(1) / (f_x * f_y);
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
        b -= target;
    }
    __device__ void precompute(int x, int y, int, long s) const {
        const T d00 = X[s], dm0 = get(X, x - 1, y), d0m = get(X, x, y - 1);
        const bool ok = inbexp1(x, y) && dv(x - 1, y) && dv(x, y) && dv(x, y - 1);
        T b = T(0), g[3] = {T(0), T(0), T(0)};
        if (ok) shade(x, y, d00, dm0, d0m, Im[s] * T(0.5) + T(0.25) * (get(Im, x - 1, y) + get(Im, x, y - 1)), b, g);
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
    // diag(J^T J) alone, in evalJTF's summation order: needs neither X nor B_I, only the gradient images and the weights
    __device__ void diagJTJ(int x, int y, int, long s, T* d) const {
        T dd = Di[s] > T(0) ? w_p * w_p : T(0);
        const T b0 = G0[s], b1 = get(G1, x + 1, y), b2 = get(G2, x, y + 1);
        // d r_h(c) / d X(s) = k_h(c) (dBI(c) - dBI(c + ex)) over the five centres touching X(s); likewise r_v
        const T th[5] = {-b0, b0 - b1, b1, -b2, b2};
        const int hx[5] = {x - 1, x, x + 1, x - 1, x}, hy[5] = {y, y, y, y + 1, y + 1};
#pragma unroll
        for (int i = 0; i < 5; ++i) {
            const T k = inimg(hx[i], hy[i]) ? kh(hx[i], hy[i]) : T(0);
            const T dr = k * th[i];
            dd += dr * dr;
        }
        const T tv[5] = {-b0, b0 - b2, -b1, b1, b2};
        const int vx[5] = {x, x, x + 1, x + 1, x}, vy[5] = {y - 1, y, y - 1, y, y + 1};
#pragma unroll
        for (int i = 0; i < 5; ++i) {
            const T k = inimg(vx[i], vy[i]) ? kv(vx[i], vy[i]) : T(0);
            const T dr = k * tv[i];
            dd += dr * dr;
        }
        const T a0 = ax(x), a1 = ay(y), an = a0 * a0 + a1 * a1 + T(1);
        const T k0 = ks(x, y);
        T k2 = T(16) * k0 * k0;
        const int ox[4] = {1, -1, 0, 0}, oy[4] = {0, 0, 1, -1};
#pragma unroll
        for (int i = 0; i < 4; ++i) {
            const int cx = x - ox[i], cy = y - oy[i];
            if (!inimg(cx, cy)) continue;
            const T k = ks(cx, cy);
            k2 += k * k;
        }
        d[0] += dd + an * k2;
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
// Marching J^T J apply for shape_from_shading, optionally fused with PCGStep3 (p' = z + beta p, engine.cuh).
// The per-site functor form above is a gather of gathers (every site re-evaluates J p of the 10 shading and 5
// regularisation residuals that touch it: ~200 loads per site, 0.47 ms for 2048^2); a shared-memory tile version
// with three block barriers per tile ran at 82 us, bound by instruction issue.  Here a warp owns a strip of
// 128 pixels (lane = 4 consecutive pixels, 128-bit loads; lanes 0 and 31 are halo lanes, 120 pixels are written)
// and marches down a band of rows; every intermediate lives in registers for the one or two rows it is needed
// and crosses lanes by shuffle.  Row n is loaded once (p, g0, g1, g2, D_i, masks, valid [+ z, CtC]) and then
//   dB[n]   = g0 p + g1 p(x-1) + g2 p[n-1]                       th[n] = k_h^2 (dB - dB(x+1))
//   tv[n-1] = k_v^2[n-1] (dB[n-1] - dB[n])                       T[n-1] = th - th(x-1) + tv[n-1] - tv[n-2]
//   out[n-1] += g0[n-1] T[n-1] + (g1 T)[n-1](x+1),  out[n-2] += g2[n-1] T[n-1]                (J^T of the shading terms)
//   regularisation, row q = n-1:  lap = 4 p - p(x-1) - p(x+1) - p[q-1] - p[q+1],
//       t0 = k_s^2 (a0 lap - (p(x+1) - p(x-1)) / f_x),  t1 = k_s^2 (a1 lap - (p[q+1] - p[q-1]) / f_y),  t2 = k_s^2 lap
//       (= k_s^2 J_s p per channel, using a0(x +- 1) = a0 +- 1/f_x and a1(y +- 1) = a1 +- 1/f_y), and with
//       w = a0 t0 + a1 t1 + t2:   out[q] += 4 w - (w + t0/f_x)(x-1) - (w - t0/f_x)(x+1),
//       out[q-1] -= w - t1/f_y,   out[q+1] = [D_i > 0] w_p^2 p (+ CtC p) - (w + t1/f_y)
// so row n-2 is complete at the end of step n.  A band of R rows costs R + 4 steps; R is chosen at plan time so that
// all bands are resident at once (2 CTAs of 8 warps per SM in fp32).  Global traffic per site: the algorithmic minimum
// (p, three gradient images, D_i, two masks, valid) plus (4 / R) of it for the warm-up rows, which hit in L2.
// Rows are prefetched three or four steps ahead with per-thread cp.async (16 bytes per lane and array, zero-filled
// outside the image) into a shared-memory ring that only the issuing thread reads back: no barrier, no
// registers held by data in flight.  MODE 0: out = J^T J p;  1: p' = z + beta p, out = J^T J p';  2: PCGInit1
// (r = -J^T F with dB replaced by B_I and p by X in the pipeline above, p = pre r, delta = 0).
namespace sfsm {
__device__ __forceinline__ void cp16(void* dst, const void* src, bool on) {
    const unsigned d = (unsigned)__cvta_generic_to_shared(dst);
    const int n = on ? 16 : 0;
    asm volatile("cp.async.cg.shared.global [%0], [%1], 16, %2;" ::"r"(d), "l"(src), "r"(n) : "memory");
}
__device__ __forceinline__ void cp4(void* dst, const void* src, bool on) {
    const unsigned d = (unsigned)__cvta_generic_to_shared(dst);
    const int n = on ? 4 : 0;
    asm volatile("cp.async.ca.shared.global [%0], [%1], 4, %2;" ::"r"(d), "l"(src), "r"(n) : "memory");
}
__device__ __forceinline__ void commit() { asm volatile("cp.async.commit_group;" ::: "memory"); }
template <int N> __device__ __forceinline__ void wait() { asm volatile("cp.async.wait_group %0;" ::"n"(N) : "memory"); }
__device__ __forceinline__ float u8f(unsigned char v, float) { return (float)v; }
__device__ __forceinline__ double u8f(unsigned char v, double) { return (double)v; }
}  // namespace sfsm

template <class T>
struct SfsMarchArgs {
    const T* z; const T* p; T* p_new; T* out; const T* ctc;     // MODE 0 / 1
    T* delta; T* r; T* pre; T* p0;                              // MODE 2
    T beta;
};

template <class T, int MODE, bool LM>
struct SfsMarch {
    using F = SfsFunctor<T>;
    static constexpr bool HAS_G = MODE != 3;                                        // the cost pass needs no gradient images
    static constexpr bool USE_C = LM && MODE < 2;
    static constexpr bool COSTLIKE = MODE >= 3;                                     // 3: computeCost, 4: computeModelCost
    static constexpr int NF = (HAS_G ? 5 : 2) + (MODE == 1) + USE_C + (MODE >= 2) + (MODE == 4);  // float arrays per row
    static constexpr int QB = 4 * (int)sizeof(T);                                   // bytes of a lane's quad
    static constexpr int STAGE = NF * 256 * QB + 3 * 256 * 4;
    static constexpr int D = NF <= 5 ? 4 : 3;                                       // ring depth
    static constexpr int A_P = 0, A_G0 = 1, A_G1 = 2, A_G2 = 3, A_DI = HAS_G ? 4 : 1, A_Z = A_DI + 1, A_C = A_DI + 1 + (MODE == 1), A_BI = A_DI + 1, A_DL = A_DI + 2;

    static __device__ __forceinline__ unsigned char* slot(unsigned char* ring, int stage, int arr, int tid) {
        return ring + (size_t)stage * STAGE + (size_t)arr * 256 * QB + (size_t)tid * QB;
    }
    static __device__ __forceinline__ unsigned char* bslot(unsigned char* ring, int stage, int arr, int tid) {
        return ring + (size_t)stage * STAGE + (size_t)NF * 256 * QB + (size_t)arr * 1024 + (size_t)tid * 4;
    }
    static __device__ __forceinline__ void cpq(unsigned char* dst, const T* src, bool on) {
#pragma unroll
        for (int k = 0; k < QB / 16; ++k) sfsm::cp16(dst + 16 * k, (const unsigned char*)src + 16 * k, on);
    }
    // one row of this lane's quad into ring stage `st` (always commits a group, possibly empty)
    static __device__ __forceinline__ void issue(const F& f, const SfsMarchArgs<T>& a, unsigned char* ring, int st, int tid,
                                                 bool on, long s) {
        if (!on) s = 0;
        cpq(slot(ring, st, A_P, tid), (MODE >= 2 ? f.X : a.p) + s, on);
        if (HAS_G) {
            cpq(slot(ring, st, A_G0, tid), f.G0 + s, on);
            cpq(slot(ring, st, A_G1, tid), f.G1 + s, on);
            cpq(slot(ring, st, A_G2, tid), f.G2 + s, on);
        }
        cpq(slot(ring, st, A_DI, tid), f.Di + s, on);
        if (MODE == 1) cpq(slot(ring, st, A_Z, tid), a.z + s, on);
        if (USE_C) cpq(slot(ring, st, A_C, tid), a.ctc + s, on);
        if (MODE >= 2) cpq(slot(ring, st, A_BI, tid), f.BI + s, on);
        if (MODE == 4) cpq(slot(ring, st, A_DL, tid), a.delta + s, on);
        sfsm::cp4(bslot(ring, st, 0, tid), f.mR + s, on);
        sfsm::cp4(bslot(ring, st, 1, tid), f.mC + s, on);
        sfsm::cp4(bslot(ring, st, 2, tid), f.valid + s, on);
        sfsm::commit();
    }
    static __device__ __forceinline__ void rd(unsigned char* ring, int st, int arr, int tid, T* o) {
        load4(reinterpret_cast<const T*>(slot(ring, st, arr, tid)), o);
    }
    static __device__ __forceinline__ uchar4 rdb(unsigned char* ring, int st, int arr, int tid) {
        return *reinterpret_cast<const uchar4*>(bslot(ring, st, arr, tid));
    }
    static __device__ void run(const F& f, const SfsMarchArgs<T>& a, unsigned char* ring, double& acc);
};

template <class T, int MODE, bool LM>
__device__ void SfsMarch<T, MODE, LM>::run(const F& f, const SfsMarchArgs<T>& a, unsigned char* ring, double& acc) {
    constexpr unsigned FULL = 0xffffffffu;
    const int W = f.W, H = f.H, R = f.bandR;
    const int strips = f.strips();
    const int row_lo = f.own_lo(), row_hi = f.own_hi();             // slab mode: only the owned rows are produced
    const long nitems = (long)strips * ((row_hi - row_lo + R - 1) / R);
    const long ntiles = (nitems + 7) / 8;
    const int lane = threadIdx.x, warp = threadIdx.y, tid = warp * 32 + lane;
    const T ifx = T(1) / f.f_x, ify = T(1) / f.f_y;
    const T wp2 = f.w_p * f.w_p, ws2 = f.w_s * f.w_s, wg2 = f.w_g * f.w_g;
    const T beta = a.beta;
    for (long tile = blockIdx.x; tile < ntiles; tile += gridDim.x) {
        const long item = tile * 8 + warp;
        if (item >= nitems) continue;                                  // warp-uniform
        const int strip = (int)(item % strips), band = (int)(item / strips);
        const int x0 = strip * F::STRIP - 4 + lane * 4;               // first of this lane's four pixels
        const int y0 = row_lo + band * R, y1 = min(y0 + R, row_hi);
        const bool col_in = x0 >= 0 && x0 < W;                        // W % 4 == 0: a quad is inside or outside as a whole
        const bool owner = col_in && lane >= 1 && lane <= 30;
        T a0[4], xin[4];
#pragma unroll
        for (int j = 0; j < 4; ++j) {
            a0[j] = ((T)(x0 + j) - f.u_x) * ifx;
            xin[j] = (x0 + j >= 1 && x0 + j + 1 < W) ? T(1) : T(0);   // x part of the expanded-bounds predicate
        }
        // rows n-1 ("1") and n-2 ("2") carried between steps
        T p1[4], p2[4], db1[4], th1[4], tvp[4], g0p[4], g1p[4], g2p[4], kv1[4], ks1[4], oA[4], oB[4], p1l = T(0);
        T dl1[4] = {T(0), T(0), T(0), T(0)};                           // MODE 4: delta of row n - 1
        bool e1[4], e2[4];
#pragma unroll
        for (int j = 0; j < 4; ++j) {
            p1[j] = p2[j] = db1[j] = th1[j] = tvp[j] = g0p[j] = g1p[j] = g2p[j] = kv1[j] = ks1[j] = oA[j] = oB[j] = T(0);
            e1[j] = e2[j] = false;
        }
        const int nbeg = y0 - (COSTLIKE ? 1 : 2), nend = y1 + (COSTLIKE ? 1 : 2);
        auto row_on = [&](int n) { return col_in && n >= 0 && n < H && n < nend; };
#pragma unroll
        for (int k = 0; k < D - 1; ++k) issue(f, a, ring, k, tid, row_on(nbeg + k), x0 + (long)W * (nbeg + k));
        int st = 0;
        for (int n = nbeg; n < nend; ++n) {
            const long s = x0 + (long)W * n;
            {
                const int sf = st == 0 ? D - 1 : st - 1;              // the stage consumed in the previous step
                issue(f, a, ring, sf, tid, row_on(n + D - 1), s + (long)W * (D - 1));
            }
            sfsm::wait<D - 1>();
            T pn[4], g0[4] = {}, g1[4] = {}, g2[4] = {}, di[4];
            rd(ring, st, A_P, tid, pn);
            if (HAS_G) { rd(ring, st, A_G0, tid, g0); rd(ring, st, A_G1, tid, g1); rd(ring, st, A_G2, tid, g2); }
            rd(ring, st, A_DI, tid, di);
            const uchar4 mr = rdb(ring, st, 0, tid), mc = rdb(ring, st, 1, tid), va = rdb(ring, st, 2, tid);
            if (MODE == 1) {
                T z[4]; rd(ring, st, A_Z, tid, z);
#pragma unroll
                for (int j = 0; j < 4; ++j) pn[j] = z[j] + beta * pn[j];
                if (owner && n >= y0 && n < y1) st4v(a.p_new + s, pn);
            }
            const T yin = f.row_inb(n) ? T(1) : T(0);
            // ---- shading terms: dB[n], th[n], tv[n-1], T[n-1] ----
            T pl = __shfl_up_sync(FULL, pn[3], 1);
            T db[4], thn[4], kvn[4], ksn[4], oC[4];
            bool en[4];
            const unsigned char mrb[4] = {mr.x, mr.y, mr.z, mr.w}, mcb[4] = {mc.x, mc.y, mc.z, mc.w}, vab[4] = {va.x, va.y, va.z, va.w};
            if (MODE >= 2) rd(ring, st, A_BI, tid, db);
            if (MODE == 4) {
                // computeModelCost (solver:1105-1118): 0.5 sum (r + J delta)^2 is the cost of the linearised residuals, i.e. the
                // cost pass with B_I -> B_I + g0 d + g1 d(x-1) + g2 d[n-1] in the shading terms and X -> X + delta in the terms
                // that are linear in X (r_p and the regulariser); delta is 0 at excluded pixels
                T dl[4]; rd(ring, st, A_DL, tid, dl);
                const T dll = __shfl_up_sync(FULL, dl[3], 1);
#pragma unroll
                for (int j = 0; j < 4; ++j) {
                    db[j] += g0[j] * dl[j] + g1[j] * (j > 0 ? dl[j > 0 ? j - 1 : 0] : dll) + g2[j] * dl1[j];
                    pn[j] += dl[j];
                    dl1[j] = dl[j];
                }
                pl = __shfl_up_sync(FULL, pn[3], 1);                   // left neighbour of the linearised row
            }
            if (MODE < 2) {
#pragma unroll
                for (int j = 0; j < 4; ++j) db[j] = g0[j] * pn[j] + g1[j] * (j > 0 ? pn[j > 0 ? j - 1 : 0] : pl) + g2[j] * p1[j];
            }
            const T dbr = __shfl_down_sync(FULL, db[0], 1);
#pragma unroll
            for (int j = 0; j < 4; ++j) {
                const T in = xin[j] * yin;
                const T mh = sfsm::u8f(mrb[j], T(0)), mv = sfsm::u8f(mcb[j], T(0));
                const T dh = db[j] - (j < 3 ? db[j < 3 ? j + 1 : 3] : dbr);
                thn[j] = (wg2 * mh * mh * in) * dh;
                kvn[j] = wg2 * mv * mv * in;
                ksn[j] = vab[j] == 1 ? ws2 * in : T(0);
                en[j] = di[j] > T(0);
                oC[j] = en[j] ? wp2 * (MODE >= 2 ? pn[j] - di[j] : pn[j]) : T(0);
                if (COSTLIKE) oC[j] = oC[j] * (pn[j] - di[j]) + (en[j] ? thn[j] * dh : T(0));      // r_p^2 + r_h^2 of row n
            }
            if (COSTLIKE) {
                // cost pass: sum of squared residuals centred on non-excluded pixels (k_cost in engine.cuh); rows n and n-1
                T cs = T(0);
                const T a1c = ((T)(n - 1 + f.y_off) - f.u_y) * ify;
                const T p1rc = __shfl_down_sync(FULL, p1[0], 1);
#pragma unroll
                for (int j = 0; j < 4; ++j) {
                    if (n >= y0 && n < y1) cs += oC[j];
                    if (n - 1 >= y0 && n - 1 < y1 && e1[j]) {
                        const T dv = db1[j] - db[j];
                        const T l = j > 0 ? p1[j > 0 ? j - 1 : 0] : p1l, r = j < 3 ? p1[j < 3 ? j + 1 : 3] : p1rc;
                        const T lap = T(4) * p1[j] - (l + p2[j] + r + pn[j]);
                        const T u0 = a0[j] * lap - ifx * (r - l), u1 = a1c * lap - ify * (pn[j] - p2[j]);
                        cs += kv1[j] * dv * dv + ks1[j] * (u0 * u0 + u1 * u1 + lap * lap);
                    }
                }
                if (owner) acc += (double)(T(0.5) * cs);
#pragma unroll
                for (int j = 0; j < 4; ++j) { p2[j] = p1[j]; p1[j] = pn[j]; db1[j] = db[j]; kv1[j] = kvn[j]; ks1[j] = ksn[j]; e1[j] = en[j]; }
                p1l = pl;
                st = st + 1 == D ? 0 : st + 1;
                continue;
            }
            if (USE_C) {
                T c[4]; rd(ring, st, A_C, tid, c);
#pragma unroll
                for (int j = 0; j < 4; ++j) oC[j] += c[j] * pn[j];
            }
            const T th1l = __shfl_up_sync(FULL, th1[3], 1);
            T tv[4], gT[4];
#pragma unroll
            for (int j = 0; j < 4; ++j) {
                tv[j] = kv1[j] * (db1[j] - db[j]);
                const T Tq = th1[j] - (j > 0 ? th1[j > 0 ? j - 1 : 0] : th1l) + tv[j] - tvp[j];
                gT[j] = g1p[j] * Tq;
                oB[j] += g0p[j] * Tq;
                oA[j] += g2p[j] * Tq;
            }
            const T gTr = __shfl_down_sync(FULL, gT[0], 1);
#pragma unroll
            for (int j = 0; j < 4; ++j) oB[j] += j < 3 ? gT[j < 3 ? j + 1 : 3] : gTr;
            // ---- regularisation terms of row q = n-1 ----
            const T p1r = __shfl_down_sync(FULL, p1[0], 1);
            const T a1q = ((T)(n - 1 + f.y_off) - f.u_y) * ify;
            T w[4], t0x[4], t1y[4];
#pragma unroll
            for (int j = 0; j < 4; ++j) {
                const T l = j > 0 ? p1[j > 0 ? j - 1 : 0] : p1l, r = j < 3 ? p1[j < 3 ? j + 1 : 3] : p1r;
                const T lap = T(4) * p1[j] - (l + p2[j] + r + pn[j]);
                const T t2 = ks1[j] * lap;
                const T t0 = a0[j] * t2 - ks1[j] * (ifx * (r - l));
                const T t1 = a1q * t2 - ks1[j] * (ify * (pn[j] - p2[j]));
                w[j] = a0[j] * t0 + a1q * t1 + t2;
                t0x[j] = ifx * t0;
                t1y[j] = ify * t1;
            }
            const T fromL = __shfl_up_sync(FULL, w[3] + t0x[3], 1);
            const T fromR = __shfl_down_sync(FULL, w[0] - t0x[0], 1);
#pragma unroll
            for (int j = 0; j < 4; ++j) {
                const T lft = j > 0 ? w[j > 0 ? j - 1 : 0] + t0x[j > 0 ? j - 1 : 0] : fromL;
                const T rgt = j < 3 ? w[j < 3 ? j + 1 : 3] - t0x[j < 3 ? j + 1 : 3] : fromR;
                oB[j] += T(4) * w[j] - lft - rgt;
                oA[j] -= w[j] - t1y[j];
                oC[j] -= w[j] + t1y[j];
            }
            // ---- row n-2 is complete ----
            if (owner && n - 2 >= y0 && n - 2 < y1) {
                const long so = s - 2 * (long)W;
                T rs = T(0);
                if (MODE == 2) {
                    // PCGInit1 (engine.cuh k_pcg_init1): preconditioner off -> diagonal 1, guarded inverse 1/4
                    const T m = guarded_invert(T(1));
                    T rr[4], pp[4], dd[4], zz[4];
#pragma unroll
                    for (int j = 0; j < 4; ++j) {
                        rr[j] = e2[j] ? -oA[j] : T(0); pp[j] = m * rr[j]; dd[j] = e2[j] ? m : T(0); zz[j] = T(0);
                        rs += rr[j] * pp[j];
                    }
                    st4v(a.r + so, rr); st4v(a.p0 + so, pp); st4v(a.pre + so, dd); st4v(a.delta + so, zz);
                } else {
                    T o[4];
#pragma unroll
                    for (int j = 0; j < 4; ++j) { o[j] = e2[j] ? oA[j] : T(0); rs += p2[j] * o[j]; }
                    st4v(a.out + so, o);
                }
                acc += (double)rs;
            }
#pragma unroll
            for (int j = 0; j < 4; ++j) {
                p2[j] = p1[j]; p1[j] = pn[j]; db1[j] = db[j]; th1[j] = thn[j]; tvp[j] = tv[j];
                g0p[j] = g0[j]; g1p[j] = g1[j]; g2p[j] = g2[j]; kv1[j] = kvn[j]; ks1[j] = ksn[j];
                oA[j] = oB[j]; oB[j] = oC[j]; e2[j] = e1[j]; e1[j] = en[j];
            }
            p1l = pl;
            st = st + 1 == D ? 0 : st + 1;
        }
        sfsm::wait<0>();
    }
}

template <class T, bool FUSED, bool LM>
__global__ void __launch_bounds__(256, sizeof(T) == 4 ? 2 : 1)
k_sfs_march(SfsFunctor<T> f, Vec<T> zv, Vec<T> p, Vec<T> p_new, Vec<T> out, Vec<T> CtC,
            const unsigned char* __restrict__, const double* bNum, const double* aNum,
            double* den, double* zero_a, double* zero_b) {
    extern __shared__ __align__(16) unsigned char sfs_ring[];
    SfsMarchArgs<T> a{};
    a.z = zv.d; a.p = p.d; a.p_new = p_new.d; a.out = out.d; a.ctc = CtC.d;
    a.beta = FUSED ? (T)(*bNum) / (T)(*aNum) : T(0);
    if (blockIdx.x == 0 && linear_tid() == 0) {
        if (zero_a) *zero_a = 0.0;
        if (zero_b) *zero_b = 0.0;
    }
    double acc = 0.0;
    SfsMarch<T, FUSED ? 1 : 0, LM>::run(f, a, sfs_ring, acc);
    block_accumulate(acc, den);
}

template <class T>
__global__ void __launch_bounds__(256, sizeof(T) == 4 ? 2 : 1)
k_sfs_init(SfsFunctor<T> f, Vec<T> delta, Vec<T> r, Vec<T> p, Vec<T> pre, double* aNum) {
    extern __shared__ __align__(16) unsigned char sfs_ring[];
    SfsMarchArgs<T> a{};
    a.delta = delta.d; a.r = r.d; a.p0 = p.d; a.pre = pre.d; a.beta = T(0);
    double acc = 0.0;
    SfsMarch<T, 2, false>::run(f, a, sfs_ring, acc);
    block_accumulate(acc, aNum);
}

template <class T>
__global__ void __launch_bounds__(256, sizeof(T) == 4 ? 2 : 1) k_sfs_cost(SfsFunctor<T> f, double* target) {
    extern __shared__ __align__(16) unsigned char sfs_ring[];
    SfsMarchArgs<T> a{};
    double acc = 0.0;
    SfsMarch<T, 3, false>::run(f, a, sfs_ring, acc);
    block_accumulate(acc, target);
}

// computeModelCost (LM, once per nonlinear iteration) on the same pipeline: MODE 4
template <class T>
__global__ void __launch_bounds__(256, sizeof(T) == 4 ? 2 : 1) k_sfs_model_cost(SfsFunctor<T> f, Vec<T> delta, double* target) {
    extern __shared__ __align__(16) unsigned char sfs_ring[];
    SfsMarchArgs<T> a{};
    a.delta = delta.d;
    double acc = 0.0;
    SfsMarch<T, 4, false>::run(f, a, sfs_ring, acc);
    block_accumulate(acc, target);
}

// precompute with one thread per quad of a row: 128-bit loads of the three X / D_i rows and two image rows, lateral
// neighbours from the adjacent lanes (the first and last lane of a warp fetch theirs), 128-bit stores of B_I, g0..g2.
template <class T>
__global__ void __launch_bounds__(256) k_sfs_precompute4(SfsFunctor<T> f) {
    constexpr unsigned FULL = 0xffffffffu;
    const int W = f.W, H = f.H;
    const int tiles_x = (W / 4 + 31) / 32;
    const long ntiles = (long)tiles_x * ((H + 7) / 8);
    const int lane = threadIdx.x;
    for (long tile = blockIdx.x; tile < ntiles; tile += gridDim.x) {
        const int x0 = ((int)(tile % tiles_x) * 32 + lane) * 4;
        const int y = (int)(tile / tiles_x) * 8 + threadIdx.y;
        const bool on = x0 < W && y < H;
        const long s = (long)y * W + x0;
        const bool up = on && y > 0, dn = on && y + 1 < H;
        T xc[4] = {}, xu[4] = {}, xd[4] = {}, dc[4] = {}, du[4] = {}, dd[4] = {}, ic[4] = {}, iu[4] = {};
        if (on) { load4(f.X + s, xc); load4(f.Di + s, dc); load4(f.Im + s, ic); }
        if (up) { load4(f.X + s - W, xu); load4(f.Di + s - W, du); load4(f.Im + s - W, iu); }
        if (dn) { load4(f.X + s + W, xd); load4(f.Di + s + W, dd); }
        const bool ledge = on && lane == 0 && x0 > 0, redge = on && (lane == 31 || x0 + 4 >= W) && x0 + 4 < W;
        T xl = __shfl_up_sync(FULL, xc[3], 1), dl = __shfl_up_sync(FULL, dc[3], 1), il = __shfl_up_sync(FULL, ic[3], 1);
        T xr = __shfl_down_sync(FULL, xc[0], 1), dr = __shfl_down_sync(FULL, dc[0], 1);
        if (lane == 0) { xl = ledge ? f.X[s - 1] : T(0); dl = ledge ? f.Di[s - 1] : T(0); il = ledge ? f.Im[s - 1] : T(0); }
        if (lane == 31) { xr = redge ? f.X[s + 4] : T(0); dr = redge ? f.Di[s + 4] : T(0); }
        if (!on) continue;
        T b[4], g0[4], g1[4], g2[4];
        unsigned char vb[4];
#pragma unroll
        for (int j = 0; j < 4; ++j) {
            const int x = x0 + j;
            const T d00 = xc[j], dm0 = j > 0 ? xc[j > 0 ? j - 1 : 0] : xl, dp0 = j < 3 ? xc[j < 3 ? j + 1 : 3] : xr;
            const T d0m = xu[j], d0p = xd[j];
            const T vl = j > 0 ? dc[j > 0 ? j - 1 : 0] : dl, vr = j < 3 ? dc[j < 3 ? j + 1 : 3] : dr;
            const T iml = j > 0 ? ic[j > 0 ? j - 1 : 0] : il;
            const bool inb = f.inbexp1(x, y);
            const bool ok = inb && vl > T(0) && dc[j] > T(0) && du[j] > T(0);
            T bb = T(0), g[3] = {T(0), T(0), T(0)};
            if (ok) f.shade(x, y, d00, dm0, d0m, ic[j] * T(0.5) + T(0.25) * (iml + iu[j]), bb, g);
            b[j] = bb; g0[j] = g[0]; g1[j] = g[1]; g2[j] = g[2];
            const bool v = inb && dc[j] > T(0) && du[j] > T(0) && dd[j] > T(0) && vl > T(0) && vr > T(0) &&
                           fabs(d00 - d0m) < T(0.01) && fabs(d00 - d0p) < T(0.01) && fabs(d00 - dm0) < T(0.01) && fabs(d00 - dp0) < T(0.01);
            vb[j] = v ? 1 : 0;
        }
        st4v(f.BI + s, b); st4v(f.G0 + s, g0); st4v(f.G1 + s, g1); st4v(f.G2 + s, g2);
        *reinterpret_cast<uchar4*>(f.valid + s) = make_uchar4(vb[0], vb[1], vb[2], vb[3]);
    }
}

template <class T> void SfsFunctor<T>::fast_precompute(cudaStream_t stream, int nsm) const {
    const long ntiles = (long)((W / 4 + 31) / 32) * ((H + 7) / 8);
    k_sfs_precompute4<T><<<(unsigned)std::min<long>(ntiles, (long)nsm * 8), dim3(32, 8, 1), 0, stream>>>(*this);
}

template <class T> auto SfsFunctor<T>::fast_cost_kernel() { return &k_sfs_cost<T>; }
template <class T> auto SfsFunctor<T>::fast_model_cost_kernel() { return &k_sfs_model_cost<T>; }
template <class T> template <bool FUSED, bool LM> auto SfsFunctor<T>::fast_kernel() { return &k_sfs_march<T, FUSED, LM>; }
template <class T> auto SfsFunctor<T>::fast_init_kernel() { return &k_sfs_init<T>; }
template <class T> size_t SfsFunctor<T>::fast_smem() const {
    size_t m = 0;
    m = std::max<size_t>(m, (size_t)SfsMarch<T, 0, false>::STAGE * SfsMarch<T, 0, false>::D);
    m = std::max<size_t>(m, (size_t)SfsMarch<T, 0, true>::STAGE * SfsMarch<T, 0, true>::D);
    m = std::max<size_t>(m, (size_t)SfsMarch<T, 1, false>::STAGE * SfsMarch<T, 1, false>::D);
    m = std::max<size_t>(m, (size_t)SfsMarch<T, 1, true>::STAGE * SfsMarch<T, 1, true>::D);
    m = std::max<size_t>(m, (size_t)SfsMarch<T, 2, false>::STAGE * SfsMarch<T, 2, false>::D);
    m = std::max<size_t>(m, (size_t)SfsMarch<T, 3, false>::STAGE * SfsMarch<T, 3, false>::D);
    m = std::max<size_t>(m, (size_t)SfsMarch<T, 4, false>::STAGE * SfsMarch<T, 4, false>::D);
    return m;
}

}  // namespace optk
