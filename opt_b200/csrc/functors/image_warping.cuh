// functors/image_warping.cuh -- image_warping (examples/image_warping/image_warping.t:12-43,105-107)
//   reg_o(i) = v(i,n) w_r [ (O_i - O_n) - R(A_i)(U_i - U_n) ],  v = inb(n) [M_n==0] [M_i==0]   (2 channels, 4 offsets)
//   fit(i)   = w_f [C_i.x>=0][C_i.y>=0] (O_i - C_i)
//   Exclude(M != 0), UsePreconditioner(true); unknown element = (Offset.xy, Angle)
// Closed forms: SURVEY.md App. D, with D_ab = -R'(A_a)(U_a - U_b).
#pragma once
#include "math.cuh"

namespace optk {

template <class T>
struct ImageWarpingFunctor : FunctorDefaults<T> {
    using Scalar = T;
    static constexpr int U = 3, ND = 2;
    static constexpr bool HAS_EXCLUDE = true;
    static constexpr bool USE_PRE = true;
    int W, H, D;
    T* O; T* A; const T* Ur; const T* Cn; const T* M;
    T w_fit, w_reg;
    // (sin A, cos A) per pixel, refreshed by precompute (once per nonlinear iteration, ghost rows included).  J^T J p reads the
    // rotation of the site and of its four neighbours: five sincos per site and PCG iteration made the generic kernel
    // arithmetic-bound on the 512^2 configuration (14.8 us per launch on an L2-resident problem); with the table it is five
    // 8-byte loads.  Same sincos_t, same operands, so the values are the ones the inline form computes (OPTB200_IW_SINCOS=1 at plan
    // time restores it; tests compare both).
    T* SC = nullptr;
    bool use_table = true;
    static constexpr bool HAS_PRECOMPUTE = true;
    void plan(const unsigned* d) { W = d[0]; H = d[1]; D = 1; use_table = getenv("OPTB200_IW_SINCOS") == nullptr; }
    void alloc_temporaries(cudaStream_t) { OPTK_CUDA(cudaMalloc(&SC, (size_t)2 * round_up((long)W * H, 64) * sizeof(T))); }
    void free_temporaries() { cudaFree(SC); SC = nullptr; }
    __device__ void precompute(int, int, int, long s) const {
        T sc[2]; sincos_t(A[s], sc[0], sc[1]);
        store2(SC + s * 2, sc);
    }
    __device__ __forceinline__ void rotation(long s, T& sa, T& ca) const {
        if (use_table) { T sc[2]; load2(SC + s * 2, sc); sa = sc[0]; ca = sc[1]; }
        else sincos_t(A[s], sa, ca);
    }
    void bind(void** pp) {
        O = (T*)pp[0]; A = (T*)pp[1]; Ur = (const T*)pp[2]; Cn = (const T*)pp[3]; M = (const T*)pp[4];
        w_fit = (T) * (const float*)pp[5]; w_reg = (T) * (const float*)pp[6];
    }
    __device__ bool exclude(int, int, int, long s) const { return M[s] != T(0); }
    __device__ T* xptr(long s, int u) const { return u < 2 ? O + s * 2 + u : A + s; }
    int unknown_images(void** ptrs, int* ch) const { ptrs[0] = O; ch[0] = 2; ptrs[1] = A; ch[1] = 1; return 2; }

    template <class Fn> __device__ __forceinline__ void neighbours(int x, int y, long s, Fn&& fn) const {
        const bool mi = M[s] == T(0);
        if (x + 1 < W && mi && M[s + 1] == T(0)) fn(s + 1);
        if (x > 0 && mi && M[s - 1] == T(0)) fn(s - 1);
        if (y + 1 < H && mi && M[s + W] == T(0)) fn(s + W);
        if (y > 0 && mi && M[s - W] == T(0)) fn(s - W);
    }
    // d = (O_a - O_b) - R(A_a) u ; Dv = -R'(A_a) u, u = U_a - U_b
    __device__ __forceinline__ static void rot(T sa, T ca, const T* u, T* Ru, T* Dv) {
        Ru[0] = ca * u[0] - sa * u[1]; Ru[1] = sa * u[0] + ca * u[1];
        Dv[0] = sa * u[0] + ca * u[1]; Dv[1] = -ca * u[0] + sa * u[1];
    }
    __device__ T cost(int x, int y, int, long s) const {
        T oi[2], ui[2], ci[2]; load2(O + s * 2, oi); load2(Ur + s * 2, ui); load2(Cn + s * 2, ci);
        T sa, ca; sincos_t(A[s], sa, ca);
        T sum = T(0);
        if (ci[0] >= T(0) && ci[1] >= T(0)) { T r0 = w_fit * (oi[0] - ci[0]), r1 = w_fit * (oi[1] - ci[1]); sum += r0 * r0 + r1 * r1; }
        neighbours(x, y, s, [&](long n) {
            T on[2], un[2]; load2(O + n * 2, on); load2(Ur + n * 2, un);
            T u[2] = {ui[0] - un[0], ui[1] - un[1]}, Ru[2], Dv[2]; rot(sa, ca, u, Ru, Dv);
            T r0 = w_reg * ((oi[0] - on[0]) - Ru[0]), r1 = w_reg * ((oi[1] - on[1]) - Ru[1]);
            sum += r0 * r0 + r1 * r1;
        });
        return sum;
    }
    __device__ void evalJTF(int x, int y, int, long s, T* g, T* d) const {
        T oi[2], ui[2], ci[2]; load2(O + s * 2, oi); load2(Ur + s * 2, ui); load2(Cn + s * 2, ci);
        T sa, ca; rotation(s, sa, ca);
        const T wf2 = w_fit * w_fit, wr2 = w_reg * w_reg;
        if (ci[0] >= T(0) && ci[1] >= T(0)) {
            g[0] += wf2 * (oi[0] - ci[0]); g[1] += wf2 * (oi[1] - ci[1]);
            d[0] += wf2; d[1] += wf2;
        }
        neighbours(x, y, s, [&](long n) {
            T on[2], un[2]; load2(O + n * 2, on); load2(Ur + n * 2, un);
            T sn, cn; rotation(n, sn, cn);
            T u[2] = {ui[0] - un[0], ui[1] - un[1]}, Ru[2], Dv[2], Rnu[2], Dn[2];
            rot(sa, ca, u, Ru, Dv); rot(sn, cn, u, Rnu, Dn);
            T din[2] = {(oi[0] - on[0]) - Ru[0], (oi[1] - on[1]) - Ru[1]};          // d(i,n)
            T dni[2] = {(on[0] - oi[0]) + Rnu[0], (on[1] - oi[1]) + Rnu[1]};        // d(n,i) = (O_n-O_i) - R(A_n)(U_n-U_i)
            g[0] += wr2 * (din[0] - dni[0]); g[1] += wr2 * (din[1] - dni[1]);
            g[2] += wr2 * (Dv[0] * din[0] + Dv[1] * din[1]);
            d[0] += T(2) * wr2; d[1] += T(2) * wr2;
            d[2] += wr2 * (Dv[0] * Dv[0] + Dv[1] * Dv[1]);
        });
    }
    // Jacobian rows (materialized-J mode): 4 offsets x 2 channels of the regulariser, then the 2 fit rows.  The structure
    // depends on the bounds only; a residual that the masks or the constraint test switch off keeps its entries with value 0.
    static constexpr bool HAS_ROWS_J = true;
    static constexpr int MAT_ROWS = 10;
    template <class Sink> __device__ void rowsJ(int x, int y, int, long s, Sink& sink) const {
        T ui[2], ci[2]; load2(Ur + s * 2, ui); load2(Cn + s * 2, ci);
        T sa, ca; sincos_t(A[s], sa, ca);
        const bool inb[4] = {x + 1 < W, x > 0, y + 1 < H, y > 0};
        const long nb[4] = {s + 1, s - 1, s + W, s - W};
        const bool mi = M[s] == T(0);
#pragma unroll
        for (int o = 0; o < 4; ++o) {
            T w = T(0), Dv[2] = {T(0), T(0)};
            if (inb[o]) {
                T un[2]; load2(Ur + nb[o] * 2, un);
                T u[2] = {ui[0] - un[0], ui[1] - un[1]}, Ru[2];
                rot(sa, ca, u, Ru, Dv);
                w = (mi && M[nb[o]] == T(0)) ? w_reg : T(0);
            }
#pragma unroll
            for (int k = 0; k < 2; ++k) {
                sink.begin(o * 2 + k);
                if (inb[o]) { sink.add(s, k, w); sink.add(nb[o], k, -w); sink.add(s, 2, w * Dv[k]); }
            }
        }
        const T wf = (ci[0] >= T(0) && ci[1] >= T(0)) ? w_fit : T(0);
        sink.begin(8); sink.add(s, 0, wf);
        sink.begin(9); sink.add(s, 1, wf);
    }
    template <class P> __device__ void applyJTJ(int x, int y, int, long s, const P& p, T* out) const {
        T ui[2], ci[2]; load2(Ur + s * 2, ui); load2(Cn + s * 2, ci);
        T sa, ca; rotation(s, sa, ca);
        const T wf2 = w_fit * w_fit, wr2 = w_reg * w_reg;
        const T p0 = p(0, s), p1 = p(1, s), pa = p(2, s);
        T a0 = T(0), a1 = T(0), a2 = T(0);
        neighbours(x, y, s, [&](long n) {
            T un[2]; load2(Ur + n * 2, un);
            T sn, cn; rotation(n, sn, cn);
            T u[2] = {ui[0] - un[0], ui[1] - un[1]}, Ru[2], Dv[2], Rnu[2], Dn[2];
            rot(sa, ca, u, Ru, Dv);
            T um[2] = {-u[0], -u[1]}; rot(sn, cn, um, Rnu, Dn);                     // D_ni = -R'(A_n)(U_n - U_i)
            const T q0 = p0 - p(0, n), q1 = p1 - p(1, n), pan = p(2, n);
            a0 += T(2) * q0 + Dv[0] * pa - Dn[0] * pan;
            a1 += T(2) * q1 + Dv[1] * pa - Dn[1] * pan;
            a2 += Dv[0] * (q0 + Dv[0] * pa) + Dv[1] * (q1 + Dv[1] * pa);
        });
        const T vf = (ci[0] >= T(0) && ci[1] >= T(0)) ? wf2 : T(0);
        out[0] = vf * p0 + wr2 * a0; out[1] = vf * p1 + wr2 * a1; out[2] = wr2 * a2;
    }
    __device__ T modelCost(int x, int y, int, long s, const Vec<T>& dl) const {
        T oi[2], ui[2], ci[2]; load2(O + s * 2, oi); load2(Ur + s * 2, ui); load2(Cn + s * 2, ci);
        T sa, ca; sincos_t(A[s], sa, ca);
        const T d0 = dl(0, s), d1 = dl(1, s), da = dl(2, s);
        T sum = T(0);
        if (ci[0] >= T(0) && ci[1] >= T(0)) {
            T m0 = w_fit * (oi[0] - ci[0]) + w_fit * d0, m1 = w_fit * (oi[1] - ci[1]) + w_fit * d1;
            sum += m0 * m0 + m1 * m1;
        }
        neighbours(x, y, s, [&](long n) {
            T on[2], un[2]; load2(O + n * 2, on); load2(Ur + n * 2, un);
            T u[2] = {ui[0] - un[0], ui[1] - un[1]}, Ru[2], Dv[2]; rot(sa, ca, u, Ru, Dv);
            T m0 = w_reg * ((oi[0] - on[0]) - Ru[0]) + w_reg * ((d0 - dl(0, n)) + Dv[0] * da);
            T m1 = w_reg * ((oi[1] - on[1]) - Ru[1]) + w_reg * ((d1 - dl(1, n)) + Dv[1] * da);
            sum += m0 * m0 + m1 * m1;
        });
        return sum;
    }
};

}  // namespace optk
