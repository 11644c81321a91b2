// functors/graph_generic.cuh -- scaffolding shared by the mesh energies whose unknown-domain terms only read the
// vertex itself and whose graph terms are small hyper-edges (cotangent_mesh_smoothing, embedded_mesh_deformation,
// robust_nonrigid_alignment).  A functor derived from GraphEnergyBase supplies two closed forms:
//     site_eval(s, r, J):  the NRS vertex residuals at vertex s and their Jacobian wrt the vertex' own U unknowns
//     edge_eval(e, v, r, J): the NRE residuals of hyper-edge e, its GV vertices and the Jacobian wrt the GV*U
//                            unknowns of those vertices (slot-major)
// and inherits every function the engine needs (the reference generates them from the energy: cost o.t:3119-3129,
// evalJTF o.t:2870-2913 / 2969-2994, applyJTJ o.t:2770-2867, modelcost o.t:2915-2966, computeCtC o.t:2996-3057).
// J^T J p over the graph domain is a SEGMENTED scatter (engine.cuh, build_segments / k_graph_contrib / k_apply_jtj_seg): the
// edge kernel stores the partial of every (edge, slot) into its own row of a buffer and the vertex kernel adds up the rows
// of its incidences, which were sorted by vertex once per solve -- no atomics, deterministic; the reference issues one
// global atomic per edge, slot and channel (o.t:2586-2600, 2833-2867).  OPTB200_GRAPH_ATOMICS=1 restores that form
// (graphJTJ below; tests compare the two).  The once-per-nonlinear-iteration scatters (J^T F, CtC) still use atomics.
// Entries of J that are structurally zero are compile-time zeros after inlining.
#pragma once
#include "math.cuh"

namespace optk {

template <class Derived, class T, int U_, int GV_, int NRS, int NRE>
struct GraphEnergyBase : FunctorDefaults<T> {
    using Scalar = T;
    static constexpr int U = U_, ND = 1, GV = GV_;
    int W = 0, H = 1, D = 1, E = 0;
    __device__ __forceinline__ const Derived& self() const { return *static_cast<const Derived*>(this); }
    __device__ bool exclude(int, int, int, long) const { return false; }

    // ---- vertex domain ---------------------------------------------------------------------------
    __device__ __forceinline__ void site(long s, T* r, T (*J)[U]) const {
#pragma unroll
        for (int k = 0; k < NRS; ++k) {
            r[k] = T(0);
#pragma unroll
            for (int u = 0; u < U; ++u) J[k][u] = T(0);
        }
        self().site_eval(s, r, J);
    }
    __device__ T cost(int, int, int, long s) const {
        T r[NRS], J[NRS][U]; site(s, r, J);
        T sum = T(0);
#pragma unroll
        for (int k = 0; k < NRS; ++k) sum += r[k] * r[k];
        return sum;
    }
    __device__ void evalJTF(int, int, int, long s, T* g, T* d) const {
        T r[NRS], J[NRS][U]; site(s, r, J);
#pragma unroll
        for (int k = 0; k < NRS; ++k)
#pragma unroll
            for (int u = 0; u < U; ++u) { g[u] += J[k][u] * r[k]; d[u] += J[k][u] * J[k][u]; }
    }
    template <class P> __device__ void applyJTJ(int, int, int, long s, const P& p, T* out) const {
        T r[NRS], J[NRS][U]; site(s, r, J);
        T pv[U];
#pragma unroll
        for (int u = 0; u < U; ++u) { pv[u] = p(u, s); out[u] = T(0); }
#pragma unroll
        for (int k = 0; k < NRS; ++k) {
            T jp = T(0);
#pragma unroll
            for (int u = 0; u < U; ++u) jp += J[k][u] * pv[u];
#pragma unroll
            for (int u = 0; u < U; ++u) out[u] += J[k][u] * jp;
        }
    }
    __device__ T modelCost(int, int, int, long s, const Vec<T>& dl) const {
        T r[NRS], J[NRS][U]; site(s, r, J);
        T sum = T(0);
#pragma unroll
        for (int k = 0; k < NRS; ++k) {
            T jd = T(0);
#pragma unroll
            for (int u = 0; u < U; ++u) jd += J[k][u] * dl(u, s);
            const T m = r[k] + jd;
            sum += m * m;
        }
        return sum;
    }

    // ---- edge domain -----------------------------------------------------------------------------
    __device__ __forceinline__ void edge(long e, long* v, T* r, T (*J)[GV * U]) const {
#pragma unroll
        for (int k = 0; k < NRE; ++k) {
            r[k] = T(0);
#pragma unroll
            for (int c = 0; c < GV * U; ++c) J[k][c] = T(0);
        }
        self().edge_eval(e, v, r, J);
    }
    __device__ T graphCost(long e) const {
        long v[GV]; T r[NRE], J[NRE][GV * U]; edge(e, v, r, J);
        T sum = T(0);
#pragma unroll
        for (int k = 0; k < NRE; ++k) sum += r[k] * r[k];
        return sum;
    }
    __device__ void graphJTF(long e, Vec<T> R, Vec<T> pre) const {
        long v[GV]; T r[NRE], J[NRE][GV * U]; edge(e, v, r, J);
#pragma unroll
        for (int s = 0; s < GV; ++s)
#pragma unroll
            for (int u = 0; u < U; ++u) {
                T g = T(0), d = T(0);
                bool any = false;
#pragma unroll
                for (int k = 0; k < NRE; ++k) {
                    const T j = J[k][s * U + u];
                    if (j != T(0)) { any = true; g += j * r[k]; d += j * j; }
                }
                if (any) { atomicAdd(&R(u, v[s]), -g); atomicAdd(&pre(u, v[s]), d); }
            }
    }
    __device__ T graphJTJ(long e, Vec<T> p, Vec<T> out) const {
        long v[GV]; T r[NRE], J[NRE][GV * U]; edge(e, v, r, J);
        T jp[NRE], den = T(0);
#pragma unroll
        for (int k = 0; k < NRE; ++k) jp[k] = T(0);
#pragma unroll
        for (int s = 0; s < GV; ++s)
#pragma unroll
            for (int u = 0; u < U; ++u) {
                const T pv = p(u, v[s]);
#pragma unroll
                for (int k = 0; k < NRE; ++k) jp[k] += J[k][s * U + u] * pv;
            }
#pragma unroll
        for (int k = 0; k < NRE; ++k) den += jp[k] * jp[k];
#pragma unroll
        for (int s = 0; s < GV; ++s)
#pragma unroll
            for (int u = 0; u < U; ++u) {
                T a = T(0);
                bool any = false;
#pragma unroll
                for (int k = 0; k < NRE; ++k) {
                    const T j = J[k][s * U + u];
                    if (j != T(0)) { any = true; a += j * jp[k]; }
                }
                if (any) atomicAdd(&out(u, v[s]), a);
            }
        return den;
    }
    static constexpr bool HAS_SEGSCATTER = true;
    // slot-major partials of edge e: contrib[s * U + u] = sum_k J[k][s U + u] (J p)[k]; every entry is written
    template <class P> __device__ void graphContrib(long e, const P& p, T* contrib) const {
        long v[GV]; T r[NRE], J[NRE][GV * U]; edge(e, v, r, J);
        T jp[NRE];
#pragma unroll
        for (int k = 0; k < NRE; ++k) jp[k] = T(0);
#pragma unroll
        for (int s = 0; s < GV; ++s)
#pragma unroll
            for (int u = 0; u < U; ++u) {
                const T pv = p(u, v[s]);
#pragma unroll
                for (int k = 0; k < NRE; ++k) jp[k] += J[k][s * U + u] * pv;
            }
#pragma unroll
        for (int c = 0; c < GV * U; ++c) {
            T a = T(0);
#pragma unroll
            for (int k = 0; k < NRE; ++k) a += J[k][c] * jp[k];
            contrib[c] = a;
        }
    }
    __device__ long edgeVertex(long e, int slot) const {
        long v[GV]; T r[NRE], J[NRE][GV * U]; edge(e, v, r, J);
        return v[slot];
    }
    __device__ T graphModelCost(long e, Vec<T> dl) const {
        long v[GV]; T r[NRE], J[NRE][GV * U]; edge(e, v, r, J);
        T sum = T(0);
#pragma unroll
        for (int k = 0; k < NRE; ++k) {
            T jd = T(0);
#pragma unroll
            for (int s = 0; s < GV; ++s)
#pragma unroll
                for (int u = 0; u < U; ++u) jd += J[k][s * U + u] * dl(u, v[s]);
            const T m = r[k] + jd;
            sum += m * m;
        }
        return sum;
    }
    __device__ void graphCtC(long e, Vec<T> CtC, T inv_radius) const {
        long v[GV]; T r[NRE], J[NRE][GV * U]; edge(e, v, r, J);
#pragma unroll
        for (int s = 0; s < GV; ++s)
#pragma unroll
            for (int u = 0; u < U; ++u) {
                T d = T(0);
#pragma unroll
                for (int k = 0; k < NRE; ++k) d += J[k][s * U + u] * J[k][s * U + u] * inv_radius;
                if (d != T(0)) atomicAdd(&CtC(u, v[s]), d);
            }
    }
};

}  // namespace optk
