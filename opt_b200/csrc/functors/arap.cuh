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
    // AoS solver vectors: the gather reads the direction of ~6 neighbours per vertex; [vertex][6] makes that one 24-byte record
    // per neighbour instead of six 4-byte reads in six planes (r1: L1 hit 68 %, 8 sectors per neighbour)
    static constexpr int VEC_UB = -6;
    int W, H, D, E;
    T* O; T* A; const T* Ur; const T* Cn; const int* v0; const int* v1;
    T w_fit, w_reg;
    // per-vertex table refreshed by precompute (once per nonlinear iteration), 12 scalars per vertex so that a
    // neighbour costs three 128-bit loads: [sa ca sb cb | sg cg ux uy | uz fit 0 0]
    T* tab = nullptr;
    static constexpr int TS = 12;
    static constexpr bool HAS_PRECOMPUTE = true, HAS_GATHER = true;
    void plan(const unsigned* d) { W = d[0]; H = 1; D = 1; E = d[1]; }
    void alloc_temporaries(cudaStream_t) { OPTK_CUDA(cudaMalloc(&tab, (size_t)TS * round_up((long)W, 64) * sizeof(T))); }
    void free_temporaries() { cudaFree(tab); tab = nullptr; }
    __device__ void precompute(int, int, int, long s) const {
        T a[3], u[3], t[TS]; load3(A + s * 3, a); load3(Ur + s * 3, u);
#pragma unroll
        for (int k = 0; k < 3; ++k) { sincos_t(a[k], t[2 * k], t[2 * k + 1]); t[6 + k] = u[k]; }
        t[9] = fit_valid(s) ? T(1) : T(0); t[10] = t[11] = T(0);
#pragma unroll
        for (int k = 0; k < TS; k += 4) st4v(tab + s * TS + k, t + k);
    }
    // ---- gather form of J^T J p (engine.cuh, k_gather_jtj).  The incidences of a vertex are sorted by the other
    // endpoint n of the edge and the two directions of a mesh edge merged: entry = 4 n + [edge v -> n] + 2 [edge n -> v]
    // (or SKIP), so both share one fetch of the neighbour's table row and direction; the index arrays are not touched
    // inside the PCG loop.
    __device__ long graph_vertex(long e, int slot) const { return slot == 0 ? v0[e] : v1[e]; }
    __device__ unsigned incidence_payload(long e, int slot) const { return (unsigned)(slot == 0 ? v1[e] : v0[e]); }
    // With M_a = sum_j pA_a[j] dR/da_j(A_a): J p of edge a->b is jp_ab = (pO_a - pO_b) - M_a (U_a - U_b);
    //   Ap_O(v) = vf pO_v + w_r^2 ( sum_{v->n} jp_vn - sum_{n->v} jp_nv ),  Ap_A(v)[j] = -w_r^2 <dR/da_j(A_v), sum_{v->n} jp_vn u^T>
    template <bool SIMPLE = false, class P> __device__ void gatherJTJ(long v, const int* off, const unsigned* inc, const P& p, T* o) const {
        T tr[TS], pc[6];
#pragma unroll
        for (int k = 0; k < TS; k += 4) load4(tab + v * TS + k, tr + k);
        load_channels<6>(p, v, pc);
        const T* uc = tr + 6;
        T M[3][3];
        { RotM<T> rm; rm.set(tr); rm.combine(pc[3], pc[4], pc[5], M); }
        T S[3][3] = {{T(0), T(0), T(0)}, {T(0), T(0), T(0)}, {T(0), T(0), T(0)}}, acc[3] = {T(0), T(0), T(0)};
        const int b0 = off[0], b1 = off[1];
        // software pipeline: entry i + 2 and the neighbour data of entry i + 1 are in flight while entry i is evaluated;
        // two buffers alternate (loop unrolled by two) so that nothing is copied between iterations
        T tA[TS], qA[6], tB[TS], qB[6];
        auto fetch = [&](unsigned ent, T* t, T* q6) {
            const long n = ent >> 2;
#pragma unroll
            for (int k = 0; k < TS; k += 4) load4(tab + n * TS + k, t + k);
            load_channels<6>(p, n, q6);
        };
        auto eval = [&](unsigned e, const T* tn, const T* pn) {
            T u[3], q[3];
#pragma unroll
            for (int k = 0; k < 3; ++k) { u[k] = uc[k] - tn[6 + k]; q[k] = pc[k] - pn[k]; }
            if (e & 1u) {                                     // edge v -> n
#pragma unroll
                for (int k = 0; k < 3; ++k) {
                    const T jp = nfma(M[k][0], u[0], nfma(M[k][1], u[1], nfma(M[k][2], u[2], q[k])));
                    acc[k] += jp;
                    S[k][0] = fma(jp, u[0], S[k][0]); S[k][1] = fma(jp, u[1], S[k][1]); S[k][2] = fma(jp, u[2], S[k][2]);
                }
            }
            if (e & 2u) {                                     // edge n -> v: acc -= jp_nv = -(q - M_n u)
                T mu[3]; rot_deriv_apply(tn, pn[3], pn[4], pn[5], u, mu);
#pragma unroll
                for (int k = 0; k < 3; ++k) acc[k] += q[k] - mu[k];
            }
        };
        if (SIMPLE) {                                         // no software pipeline: fewer registers, more resident warps
            for (int i = b0; i < b1; ++i) {
                const unsigned e = inc[i];
                fetch(e, tA, qA);
                eval(e, tA, qA);
            }
        } else {
        unsigned e0 = 0, e1 = 0;
        if (b0 < b1) { e0 = inc[b0]; fetch(e0, tA, qA); }
        if (b0 + 1 < b1) e1 = inc[b0 + 1];
        for (int i = b0; i < b1; i += 2) {
            unsigned e2 = 0, e3 = 0;
            if (i + 1 < b1) fetch(e1, tB, qB);
            if (i + 2 < b1) e2 = inc[i + 2];
            eval(e0, tA, qA);
            if (i + 1 >= b1) break;
            if (i + 2 < b1) fetch(e2, tA, qA);
            if (i + 3 < b1) e3 = inc[i + 3];
            eval(e1, tB, qB);
            e0 = e2; e1 = e3;
        }
        }
        const T wr2 = w_reg * w_reg, vf = tr[9] * w_fit * w_fit;
        RotM<T> rm; rm.set(tr);
        T oA[3]; rm.project(S, oA);
#pragma unroll
        for (int k = 0; k < 3; ++k) { o[k] = vf * pc[k] + wr2 * acc[k]; o[3 + k] = wr2 * oA[k]; }
    }
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
    __device__ T modelCost(int, int, int, long s, const Vec<T, VEC_UB>& dl) const {
        T sum = T(0);
        if (fit_valid(s)) {
#pragma unroll
            for (int k = 0; k < 3; ++k) { T m = w_fit * (O[s * 3 + k] - Cn[s * 3 + k]) + w_fit * dl(k, s); sum += m * m; }
        }
        return sum;
    }
    // ---- Jacobian rows (materialized-J mode): 3 fit rows per vertex; 3 rows per edge a -> b with +w_r at O_a, -w_r at O_b and
    // w_r D_ab[k][j] at A_a[j] (saveJToCRS_Graph, solver:1287-1307)
    static constexpr bool HAS_ROWS_J = true;
    static constexpr int MAT_ROWS = 3, MAT_GROWS = 3;
    template <class Sink> __device__ void rowsJ(int, int, int, long s, Sink& sink) const {
        const T wf = fit_valid(s) ? w_fit : T(0);
#pragma unroll
        for (int k = 0; k < 3; ++k) { sink.begin(k); sink.add(s, k, wf); }
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
    template <class Sink> __device__ void graphRowsJ(long ei, Sink& sink) const {
        Edge ed; edge(ei, ed, false);
#pragma unroll
        for (int k = 0; k < 3; ++k) {
            sink.begin(k);
            sink.add(ed.a, k, w_reg); sink.add(ed.b, k, -w_reg);
#pragma unroll
            for (int j = 0; j < 3; ++j) sink.add(ed.a, 3 + j, w_reg * ed.Dm[k][j]);
        }
    }
    __device__ T graphCost(long ei) const {
        Edge ed; edge(ei, ed, true);
        T sum = T(0);
#pragma unroll
        for (int k = 0; k < 3; ++k) { T r = w_reg * ed.e[k]; sum += r * r; }
        return sum;
    }
    __device__ void graphJTF(long ei, Vec<T, VEC_UB> r, Vec<T, VEC_UB> pre) const {
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
    __device__ T graphJTJ(long ei, Vec<T, VEC_UB> p, Vec<T, VEC_UB> out) const {
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
    __device__ T graphModelCost(long ei, Vec<T, VEC_UB> dl) const {
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
    __device__ void graphCtC(long ei, Vec<T, VEC_UB> CtC, T inv_radius) const {
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
