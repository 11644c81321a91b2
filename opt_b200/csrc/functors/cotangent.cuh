// functors/cotangent.cuh -- cotangent_mesh_smoothing (examples/cotangent_mesh_smoothing/cotangent_mesh_smoothing.t:1-37)
//   unknown X (float3), UsePreconditioner(true), no Exclude
//   vertex: w_f (X - A)
//   hyper-edge (v0 current, v1 neighbour, v2 previous, v3 next neighbour):
//       a = nrm(X0-X2), b = nrm(X1-X2), c = nrm(X0-X3), d = nrm(X1-X3)
//       cot(p,q) = p.q / sqrt(sel(disc > 0, disc, 1e-4)),  disc = (p.p)(q.q) - (p.q)^2      (the literal expression:
//                  p.p and q.q are differentiated too, as the reference's AD does)
//       w = sqrt(sel(om > 0, om, 1e-4)),  om = (cot(a,b) + cot(c,d)) / 2
//       r = w_r w (X1 - X0)
//   d r / d X_s = w_r ( (X1-X0) (grad_{X_s} w)^T + w d(X1-X0)/dX_s ),  grad w = [om > 0] grad om / (2 w)
#pragma once
#include "graph_generic.cuh"

namespace optk {

template <class T>
struct CotangentFunctor : GraphEnergyBase<CotangentFunctor<T>, T, 3, 4, 3, 3> {
    static constexpr bool USE_PRE = true;
    T* X; const T* A; const int* vi[4];
    T w_fit, w_reg;
    void plan(const unsigned* d) { this->W = d[0]; this->E = d[1]; }
    void bind(void** pp) {
        w_fit = (T) * (const float*)pp[0]; w_reg = (T) * (const float*)pp[1];
        X = (T*)pp[2]; A = (const T*)pp[3];
        for (int k = 0; k < 4; ++k) vi[k] = (const int*)pp[5 + k];                 // pp[4]: edge count, ignored (util.t:696)
    }
    __device__ T* xptr(long s, int u) const { return X + s * 3 + u; }

    __device__ __forceinline__ void site_eval(long s, T* r, T (*J)[3]) const {
#pragma unroll
        for (int k = 0; k < 3; ++k) { r[k] = w_fit * (X[s * 3 + k] - A[s * 3 + k]); J[k][k] = w_fit; }
    }
    // n = p / |p| and the pull-back of a gradient wrt n to a gradient wrt p: (I - n n^T) g / |p|
    struct Unit { T n[3], inv; };
    __device__ __forceinline__ static Unit unit(const T* p0, const T* p1) {
        Unit u; T p[3] = {p0[0] - p1[0], p0[1] - p1[1], p0[2] - p1[2]};
        u.inv = T(1) / sqrt(p[0] * p[0] + p[1] * p[1] + p[2] * p[2]);
#pragma unroll
        for (int k = 0; k < 3; ++k) u.n[k] = p[k] * u.inv;
        return u;
    }
    __device__ __forceinline__ static void pull(const Unit& u, const T* g, T* o) {
        const T ng = u.n[0] * g[0] + u.n[1] * g[1] + u.n[2] * g[2];
#pragma unroll
        for (int k = 0; k < 3; ++k) o[k] = (g[k] - u.n[k] * ng) * u.inv;
    }
    // value of cot(p,q) and its gradients wrt p and q
    __device__ __forceinline__ static T cot(const T* p, const T* q, T* gp, T* gq) {
        const T s = p[0] * q[0] + p[1] * q[1] + p[2] * q[2];
        const T Ap = p[0] * p[0] + p[1] * p[1] + p[2] * p[2], Bq = q[0] * q[0] + q[1] * q[1] + q[2] * q[2];
        const T disc = Ap * Bq - s * s;
        T fs, fA, fB, val;
        if (disc > T(0)) {
            const T i1 = T(1) / sqrt(disc), i3 = i1 / disc;
            val = s * i1;
            fs = i1 + s * s * i3; fA = T(-0.5) * s * Bq * i3; fB = T(-0.5) * s * Ap * i3;
        } else {
            const T i1 = T(1) / sqrt(T(0.0001));
            val = s * i1; fs = i1; fA = T(0); fB = T(0);
        }
#pragma unroll
        for (int k = 0; k < 3; ++k) { gp[k] = fs * q[k] + T(2) * fA * p[k]; gq[k] = fs * p[k] + T(2) * fB * q[k]; }
        return val;
    }
    __device__ __forceinline__ void edge_eval(long e, long* v, T* r, T (*J)[12]) const {
        T x[4][3];
#pragma unroll
        for (int s = 0; s < 4; ++s) {
            v[s] = vi[s][e];
#pragma unroll
            for (int k = 0; k < 3; ++k) x[s][k] = X[v[s] * 3 + k];
        }
        const Unit a = unit(x[0], x[2]), b = unit(x[1], x[2]), c = unit(x[0], x[3]), d = unit(x[1], x[3]);
        T ga[3], gb[3], gc[3], gd[3];
        const T om = T(0.5) * (cot(a.n, b.n, ga, gb) + cot(c.n, d.n, gc, gd));
        const bool pos = om > T(0);
        const T w = sqrt(pos ? om : T(0.0001));
        T Ga[3], Gb[3], Gc[3], Gd[3];
        pull(a, ga, Ga); pull(b, gb, Gb); pull(c, gc, Gc); pull(d, gd, Gd);
        const T hw = pos ? T(0.25) / w : T(0);                 // (1/2) from om, 1/(2w) from the square root
        T gw[4][3];
#pragma unroll
        for (int k = 0; k < 3; ++k) {
            gw[0][k] = hw * (Ga[k] + Gc[k]);
            gw[1][k] = hw * (Gb[k] + Gd[k]);
            gw[2][k] = -hw * (Ga[k] + Gb[k]);
            gw[3][k] = -hw * (Gc[k] + Gd[k]);
        }
#pragma unroll
        for (int k = 0; k < 3; ++k) {
            const T ek = x[1][k] - x[0][k];
            r[k] = w_reg * w * ek;
#pragma unroll
            for (int s = 0; s < 4; ++s)
#pragma unroll
                for (int j = 0; j < 3; ++j) J[k][s * 3 + j] = w_reg * ek * gw[s][j];
            J[k][k] -= w_reg * w;
            J[k][3 + k] += w_reg * w;
        }
    }
};

}  // namespace optk
