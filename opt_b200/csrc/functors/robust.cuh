// functors/robust.cuh -- robust_nonrigid_alignment (examples/robust_nonrigid_alignment/robust_nonrigid_alignment.t:1-30)
//   unknown element = (Offset.xyz, Angle.xyz, RobustWeight), U = 7; UsePreconditioner(true); no Exclude
//   vertex (the comparison greatereq(Constraints(0), -999999.9) is a 3-vector, so ad broadcasts both terms to three
//   residuals, one per constraint component k; ad.t:373-395):
//       fit_k  = [C_k >= -999999.9] w_f  w (N . (O - C))
//       conf_k = [C_k >= -999999.9] 0.1 (1 - w^2)
//   edge a=v0 -> b=v1:  w_r [ (O_a - O_b) - R3D(A_a)(U_a - U_b) ]                  (ARAP, lib.t:84-98)
#pragma once
#include "graph_generic.cuh"

namespace optk {

template <class T>
struct RobustFunctor : GraphEnergyBase<RobustFunctor<T>, T, 7, 2, 6, 3> {
    static constexpr bool USE_PRE = true;
    T* O; T* A; T* Rw; const T* Ur; const T* Cn; const T* Nn; const int* v0; const int* v1;
    T w_fit, w_reg;
    void plan(const unsigned* d) { this->W = d[0]; this->E = d[1]; }
    void bind(void** pp) {
        w_fit = (T) * (const float*)pp[0]; w_reg = (T) * (const float*)pp[1];
        O = (T*)pp[2]; A = (T*)pp[3]; Rw = (T*)pp[4]; Ur = (const T*)pp[5]; Cn = (const T*)pp[6]; Nn = (const T*)pp[7];
        v0 = (const int*)pp[9]; v1 = (const int*)pp[10];                         // pp[8]: edge count, ignored (util.t:696)
    }
    __device__ T* xptr(long s, int u) const { return u < 3 ? O + s * 3 + u : u < 6 ? A + s * 3 + (u - 3) : Rw + s; }

    __device__ __forceinline__ void site_eval(long s, T* r, T (*J)[7]) const {
        const T w = Rw[s];
        T nd = T(0), n[3];
        bool any = false, valid[3];
#pragma unroll
        for (int k = 0; k < 3; ++k) { valid[k] = Cn[s * 3 + k] >= T(-999999.9); any = any || valid[k]; n[k] = Nn[s * 3 + k]; }
        if (!any) return;
#pragma unroll
        for (int k = 0; k < 3; ++k) nd += n[k] * (O[s * 3 + k] - Cn[s * 3 + k]);
#pragma unroll
        for (int k = 0; k < 3; ++k) {
            if (!valid[k]) continue;
            r[k] = w_fit * (w * nd);
#pragma unroll
            for (int j = 0; j < 3; ++j) J[k][j] = w_fit * w * n[j];
            J[k][6] = w_fit * nd;
            r[3 + k] = T(0.1) * (T(1) - w * w);
            J[3 + k][6] = T(-0.2) * w;
        }
    }
    __device__ __forceinline__ void edge_eval(long e, long* v, T* r, T (*J)[14]) const {
        const long a = v0[e], b = v1[e];
        v[0] = a; v[1] = b;
        T aa[3], u[3], Ru[3], Dm[3][3];
        load3(A + a * 3, aa);
#pragma unroll
        for (int k = 0; k < 3; ++k) u[k] = Ur[a * 3 + k] - Ur[b * 3 + k];
        Rot3<T> R; R.set(aa); R.apply(u, Ru); R.jac(u, Dm);
#pragma unroll
        for (int k = 0; k < 3; ++k) {
            r[k] = w_reg * ((O[a * 3 + k] - O[b * 3 + k]) - Ru[k]);
            J[k][k] = w_reg;
            J[k][7 + k] = -w_reg;
#pragma unroll
            for (int j = 0; j < 3; ++j) J[k][3 + j] = w_reg * Dm[k][j];
        }
    }
};

}  // namespace optk
