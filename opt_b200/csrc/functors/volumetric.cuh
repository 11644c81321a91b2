// functors/volumetric.cuh -- volumetric_mesh_deformation
// (examples/volumetric_mesh_deformation/volumetric_mesh_deformation.t:1-21)
//   fit(i)   = [C_i.x >= -999999.9] w_f (O_i - C_i)                                   (3 channels)
//   reg_o(i) = inb(i) inb(n) w_r [ (O_i - O_n) - R3D(A_i)(U_i - U_n) ]                (3 channels, 6 offsets)
//   no Exclude, UsePreconditioner(true); unknown element = (Offset.xyz, Angle.xyz)
// Closed forms: SURVEY.md App. D with D_ab = -[dR/da u | dR/db u | dR/dg u], u = U_a - U_b.
#pragma once
#include "math.cuh"

namespace optk {

template <class T>
struct VolumetricFunctor : FunctorDefaults<T> {
    using Scalar = T;
    static constexpr int U = 6, ND = 3;
    static constexpr bool USE_PRE = true;
    // the generic lazy p' = z + beta p accessor doubles the 42 stencil loads of this functor and the kernel
    // is issue-bound, not DRAM-bound (measured: fused 2.28 ms vs 1.44 + 0.17 ms at 256^3)
    static constexpr bool FUSE_STEP3 = false;
    int W, H, D;
    T* O; T* A; const T* Ur; const T* Cn;
    T w_fit, w_reg;
    void plan(const unsigned* d) { W = d[0]; H = d[1]; D = d[2]; }
    void bind(void** pp) {
        O = (T*)pp[0]; A = (T*)pp[1]; Ur = (const T*)pp[2]; Cn = (const T*)pp[3];
        w_fit = (T) * (const float*)pp[4]; w_reg = (T) * (const float*)pp[5];
    }
    __device__ bool exclude(int, int, int, long) const { return false; }
    __device__ T* xptr(long s, int u) const { return u < 3 ? O + s * 3 + u : A + s * 3 + (u - 3); }
    int unknown_images(void** ptrs, int* ch) const { ptrs[0] = O; ch[0] = 3; ptrs[1] = A; ch[1] = 3; return 2; }

    template <class Fn> __device__ __forceinline__ void neighbours(int x, int y, int z, long s, Fn&& fn) const {
        const long sy = W, sz = (long)W * H;
        if (x + 1 < W) fn(s + 1);
        if (x > 0) fn(s - 1);
        if (y + 1 < H) fn(s + sy);
        if (y > 0) fn(s - sy);
        if (z + 1 < D) fn(s + sz);
        if (z > 0) fn(s - sz);
    }
    __device__ bool fit_valid(long s) const { return Cn[s * 3] >= T(-999999.9); }

    __device__ T cost(int x, int y, int z, long s) const {
        T oi[3], ui[3], ai[3]; load3(O + s * 3, oi); load3(Ur + s * 3, ui); load3(A + s * 3, ai);
        Rot3<T> R; R.set(ai);
        T sum = T(0);
        if (fit_valid(s)) {
#pragma unroll
            for (int k = 0; k < 3; ++k) { T r = w_fit * (oi[k] - Cn[s * 3 + k]); sum += r * r; }
        }
        neighbours(x, y, z, s, [&](long n) {
            T on[3], un[3]; load3(O + n * 3, on); load3(Ur + n * 3, un);
            T u[3] = {ui[0] - un[0], ui[1] - un[1], ui[2] - un[2]}, Ru[3]; R.apply(u, Ru);
#pragma unroll
            for (int k = 0; k < 3; ++k) { T r = w_reg * ((oi[k] - on[k]) - Ru[k]); sum += r * r; }
        });
        return sum;
    }
    __device__ void evalJTF(int x, int y, int z, long s, T* g, T* d) const {
        T oi[3], ui[3], ai[3]; load3(O + s * 3, oi); load3(Ur + s * 3, ui); load3(A + s * 3, ai);
        Rot3<T> R; R.set(ai);
        const T wf2 = w_fit * w_fit, wr2 = w_reg * w_reg;
        if (fit_valid(s)) {
#pragma unroll
            for (int k = 0; k < 3; ++k) { g[k] += wf2 * (oi[k] - Cn[s * 3 + k]); d[k] += wf2; }
        }
        neighbours(x, y, z, s, [&](long n) {
            T on[3], un[3], an[3]; load3(O + n * 3, on); load3(Ur + n * 3, un); load3(A + n * 3, an);
            Rot3<T> Rn; Rn.set(an);
            T u[3] = {ui[0] - un[0], ui[1] - un[1], ui[2] - un[2]}, Ru[3], Rnu[3], Dm[3][3];
            R.apply(u, Ru); Rn.apply(u, Rnu); R.jac(u, Dm);
            T din[3], dni[3];
#pragma unroll
            for (int k = 0; k < 3; ++k) { din[k] = (oi[k] - on[k]) - Ru[k]; dni[k] = (on[k] - oi[k]) + Rnu[k]; }
#pragma unroll
            for (int k = 0; k < 3; ++k) { g[k] += wr2 * (din[k] - dni[k]); d[k] += T(2) * wr2; }
#pragma unroll
            for (int j = 0; j < 3; ++j) {
                g[3 + j] += wr2 * (Dm[0][j] * din[0] + Dm[1][j] * din[1] + Dm[2][j] * din[2]);
                d[3 + j] += wr2 * (Dm[0][j] * Dm[0][j] + Dm[1][j] * Dm[1][j] + Dm[2][j] * Dm[2][j]);
            }
        });
    }
    template <class P> __device__ void applyJTJ(int x, int y, int z, long s, const P& p, T* out) const {
        T ui[3], ai[3]; load3(Ur + s * 3, ui); load3(A + s * 3, ai);
        Rot3<T> R; R.set(ai);
        const T wf2 = w_fit * w_fit, wr2 = w_reg * w_reg;
        T pi[6], acc[6];
#pragma unroll
        for (int k = 0; k < 6; ++k) { pi[k] = p(k, s); acc[k] = T(0); }
        neighbours(x, y, z, s, [&](long n) {
            T un[3], an[3]; load3(Ur + n * 3, un); load3(A + n * 3, an);
            Rot3<T> Rn; Rn.set(an);
            T u[3] = {ui[0] - un[0], ui[1] - un[1], ui[2] - un[2]}, um[3] = {-u[0], -u[1], -u[2]}, Dm[3][3], Dn[3][3];
            R.jac(u, Dm); Rn.jac(um, Dn);                                           // D_in and D_ni
            T pn[6];
#pragma unroll
            for (int k = 0; k < 6; ++k) pn[k] = p(k, n);
            T jp[3];                                                                // (pO_i - pO_n) + D_in pA_i
#pragma unroll
            for (int k = 0; k < 3; ++k) {
                const T q = pi[k] - pn[k];
                const T da = Dm[k][0] * pi[3] + Dm[k][1] * pi[4] + Dm[k][2] * pi[5];
                const T dn = Dn[k][0] * pn[3] + Dn[k][1] * pn[4] + Dn[k][2] * pn[5];
                jp[k] = q + da;
                acc[k] += T(2) * q + da - dn;
            }
#pragma unroll
            for (int j = 0; j < 3; ++j) acc[3 + j] += Dm[0][j] * jp[0] + Dm[1][j] * jp[1] + Dm[2][j] * jp[2];
        });
        const T vf = fit_valid(s) ? wf2 : T(0);
#pragma unroll
        for (int k = 0; k < 3; ++k) { out[k] = vf * pi[k] + wr2 * acc[k]; out[3 + k] = wr2 * acc[3 + k]; }
    }
    __device__ T modelCost(int x, int y, int z, long s, const Vec<T>& dl) const {
        T oi[3], ui[3], ai[3]; load3(O + s * 3, oi); load3(Ur + s * 3, ui); load3(A + s * 3, ai);
        Rot3<T> R; R.set(ai);
        T di[6];
#pragma unroll
        for (int k = 0; k < 6; ++k) di[k] = dl(k, s);
        T sum = T(0);
        if (fit_valid(s)) {
#pragma unroll
            for (int k = 0; k < 3; ++k) { T m = w_fit * (oi[k] - Cn[s * 3 + k]) + w_fit * di[k]; sum += m * m; }
        }
        neighbours(x, y, z, s, [&](long n) {
            T on[3], un[3]; load3(O + n * 3, on); load3(Ur + n * 3, un);
            T u[3] = {ui[0] - un[0], ui[1] - un[1], ui[2] - un[2]}, Ru[3], Dm[3][3]; R.apply(u, Ru); R.jac(u, Dm);
#pragma unroll
            for (int k = 0; k < 3; ++k) {
                T jd = (di[k] - dl(k, n)) + Dm[k][0] * di[3] + Dm[k][1] * di[4] + Dm[k][2] * di[5];
                T m = w_reg * ((oi[k] - on[k]) - Ru[k]) + w_reg * jd;
                sum += m * m;
            }
        });
        return sum;
    }
};

}  // namespace optk
