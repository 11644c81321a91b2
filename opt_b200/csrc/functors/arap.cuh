// functors/arap.cuh -- arap_mesh_deformation (examples/arap_mesh_deformation/arap_mesh_deformation.t:11-48)
//   vertex domain: fit(i) = [C_i.x >= -999999.9] w_f (O_i - C_i)
//   edge domain:   r_e = w_r [ (O_a - O_b) - R3D(A_a)(U_a - U_b) ],  a = v0[e], b = v1[e]
//   unknown support of r_e: O_a (+w_r I), O_b (-w_r I), A_a (w_r D_ab)
//   no Exclude, UsePreconditioner(true); unknown element = (Offset.xyz, Angle.xyz)
// Graph scatter: the reference issues one global atomic per edge and channel (o.t:2586-2600).
#pragma once
#include "math.cuh"

namespace optk {

template <class T>
struct ArapFunctor : FunctorDefaults<T> {
    using Scalar = T;
    static constexpr int U = 6, ND = 1, GV = 2;
    static constexpr bool USE_PRE = true;
    int W, H, D, E;
    T* O; T* A; const T* Ur; const T* Cn; const int* v0; const int* v1;
    T w_fit, w_reg;
    void plan(const unsigned* d) { W = d[0]; H = 1; D = 1; E = d[1]; }
    void bind(void** pp) {
        w_fit = (T) * (const float*)pp[0]; w_reg = (T) * (const float*)pp[1];
        O = (T*)pp[2]; A = (T*)pp[3]; Ur = (const T*)pp[4]; Cn = (const T*)pp[5];
        v0 = (const int*)pp[7]; v1 = (const int*)pp[8];                       // pp[6]: edge count, ignored (util.t:696)
    }
    __device__ bool exclude(int, int, int, long) const { return false; }
    __device__ T* xptr(long s, int u) const { return u < 3 ? O + s * 3 + u : A + s * 3 + (u - 3); }
    __device__ bool fit_valid(long s) const { return Cn[s * 3] >= T(-999999.9); }

    // ---- vertex domain
    __device__ T cost(int, int, int, long s) const {
        T sum = T(0);
        if (fit_valid(s)) {
#pragma unroll
            for (int k = 0; k < 3; ++k) { T r = w_fit * (O[s * 3 + k] - Cn[s * 3 + k]); sum += r * r; }
        }
        return sum;
    }
    __device__ void evalJTF(int, int, int, long s, T* g, T* d) const {
        const T wf2 = w_fit * w_fit;
        if (fit_valid(s)) {
#pragma unroll
            for (int k = 0; k < 3; ++k) { g[k] += wf2 * (O[s * 3 + k] - Cn[s * 3 + k]); d[k] += wf2; }
        }
    }
    template <class P> __device__ void applyJTJ(int, int, int, long s, const P& p, T* out) const {
        const T vf = fit_valid(s) ? w_fit * w_fit : T(0);
#pragma unroll
        for (int k = 0; k < 3; ++k) { out[k] = vf * p(k, s); out[3 + k] = T(0); }
    }
    __device__ T modelCost(int, int, int, long s, const Vec<T>& dl) const {
        T sum = T(0);
        if (fit_valid(s)) {
#pragma unroll
            for (int k = 0; k < 3; ++k) { T m = w_fit * (O[s * 3 + k] - Cn[s * 3 + k]) + w_fit * dl(k, s); sum += m * m; }
        }
        return sum;
    }
    // ---- edge domain
    struct Edge { long a, b; T e[3]; T Dm[3][3]; };
    __device__ __forceinline__ void edge(long ei, Edge& ed, bool need_e) const {
        ed.a = v0[ei]; ed.b = v1[ei];
        T ua[3], ub[3], aa[3]; load3(Ur + ed.a * 3, ua); load3(Ur + ed.b * 3, ub); load3(A + ed.a * 3, aa);
        Rot3<T> R; R.set(aa);
        T u[3] = {ua[0] - ub[0], ua[1] - ub[1], ua[2] - ub[2]};
        R.jac(u, ed.Dm);
        if (need_e) {
            T Ru[3]; R.apply(u, Ru);
#pragma unroll
            for (int k = 0; k < 3; ++k) ed.e[k] = (O[ed.a * 3 + k] - O[ed.b * 3 + k]) - Ru[k];
        }
    }
    __device__ T graphCost(long ei) const {
        Edge ed; edge(ei, ed, true);
        T sum = T(0);
#pragma unroll
        for (int k = 0; k < 3; ++k) { T r = w_reg * ed.e[k]; sum += r * r; }
        return sum;
    }
    __device__ void graphJTF(long ei, Vec<T> r, Vec<T> pre) const {
        Edge ed; edge(ei, ed, true);
        const T wr2 = w_reg * w_reg;
#pragma unroll
        for (int k = 0; k < 3; ++k) {
            atomicAdd(&r(k, ed.a), -wr2 * ed.e[k]);
            atomicAdd(&r(k, ed.b), wr2 * ed.e[k]);
            atomicAdd(&pre(k, ed.a), wr2);
            atomicAdd(&pre(k, ed.b), wr2);
        }
#pragma unroll
        for (int j = 0; j < 3; ++j) {
            atomicAdd(&r(3 + j, ed.a), -wr2 * (ed.Dm[0][j] * ed.e[0] + ed.Dm[1][j] * ed.e[1] + ed.Dm[2][j] * ed.e[2]));
            atomicAdd(&pre(3 + j, ed.a), wr2 * (ed.Dm[0][j] * ed.Dm[0][j] + ed.Dm[1][j] * ed.Dm[1][j] + ed.Dm[2][j] * ed.Dm[2][j]));
        }
    }
    __device__ T graphJTJ(long ei, Vec<T> p, Vec<T> out) const {
        Edge ed; edge(ei, ed, false);
        const T wr2 = w_reg * w_reg;
        T pa[3] = {p(3, ed.a), p(4, ed.a), p(5, ed.a)};
        T jp[3], den = T(0);
#pragma unroll
        for (int k = 0; k < 3; ++k) {
            jp[k] = (p(k, ed.a) - p(k, ed.b)) + ed.Dm[k][0] * pa[0] + ed.Dm[k][1] * pa[1] + ed.Dm[k][2] * pa[2];
            den += jp[k] * jp[k];
            atomicAdd(&out(k, ed.a), wr2 * jp[k]);
            atomicAdd(&out(k, ed.b), -wr2 * jp[k]);
        }
#pragma unroll
        for (int j = 0; j < 3; ++j)
            atomicAdd(&out(3 + j, ed.a), wr2 * (ed.Dm[0][j] * jp[0] + ed.Dm[1][j] * jp[1] + ed.Dm[2][j] * jp[2]));
        return wr2 * den;
    }
    __device__ T graphModelCost(long ei, Vec<T> dl) const {
        Edge ed; edge(ei, ed, true);
        T da[3] = {dl(3, ed.a), dl(4, ed.a), dl(5, ed.a)};
        T sum = T(0);
#pragma unroll
        for (int k = 0; k < 3; ++k) {
            T jd = (dl(k, ed.a) - dl(k, ed.b)) + ed.Dm[k][0] * da[0] + ed.Dm[k][1] * da[1] + ed.Dm[k][2] * da[2];
            T m = w_reg * ed.e[k] + w_reg * jd;
            sum += m * m;
        }
        return sum;
    }
    __device__ void graphCtC(long ei, Vec<T> CtC, T inv_radius) const {
        Edge ed; edge(ei, ed, false);
        const T wr2 = w_reg * w_reg;
#pragma unroll
        for (int k = 0; k < 3; ++k) { atomicAdd(&CtC(k, ed.a), wr2 * inv_radius); atomicAdd(&CtC(k, ed.b), wr2 * inv_radius); }
#pragma unroll
        for (int j = 0; j < 3; ++j)
            atomicAdd(&CtC(3 + j, ed.a), wr2 * (ed.Dm[0][j] * ed.Dm[0][j] + ed.Dm[1][j] * ed.Dm[1][j] + ed.Dm[2][j] * ed.Dm[2][j]) * inv_radius);
    }
};

}  // namespace optk
