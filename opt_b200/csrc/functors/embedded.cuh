// functors/embedded.cuh -- embedded_mesh_deformation (examples/embedded_mesh_deformation/embedded_mesh_deformation.t:1-34)
//   unknown element = (Offset.xyz, RotMatrix[9] row-major), U = 12; UsePreconditioner(true); no Exclude
//   vertex:  fit  [C.x >= -999999.9] w_f (O - C)                                  (3)
//            rot  w_t c0.c1, w_t c0.c2, w_t c1.c2, w_t (c_j.c_j - 1), c_j = column j of R     (6)
//   edge a=v0 -> b=v1:  w_r [ (O_b - O_a) - R_a (U_b - U_a) ]                     (3)
#pragma once
#include "graph_generic.cuh"

namespace optk {

template <class T>
struct EmbeddedFunctor : GraphEnergyBase<EmbeddedFunctor<T>, T, 12, 2, 9, 3> {
    static constexpr bool USE_PRE = true;
    T* O; T* R; const T* Ur; const T* Cn; const int* v0; const int* v1;
    T w_fit, w_reg, w_rot;
    void plan(const unsigned* d) { this->W = d[0]; this->E = d[1]; }
    void bind(void** pp) {
        w_fit = (T) * (const float*)pp[0]; w_reg = (T) * (const float*)pp[1]; w_rot = (T) * (const float*)pp[2];
        O = (T*)pp[3]; R = (T*)pp[4]; Ur = (const T*)pp[5]; Cn = (const T*)pp[6];
        v0 = (const int*)pp[8]; v1 = (const int*)pp[9];                         // pp[7]: edge count, ignored (util.t:696)
    }
    __device__ T* xptr(long s, int u) const { return u < 3 ? O + s * 3 + u : R + s * 9 + (u - 3); }

    __device__ __forceinline__ void site_eval(long s, T* r, T (*J)[12]) const {
        if (Cn[s * 3] >= T(-999999.9)) {
#pragma unroll
            for (int k = 0; k < 3; ++k) { r[k] = w_fit * (O[s * 3 + k] - Cn[s * 3 + k]); J[k][k] = w_fit; }
        }
        T m[9];
#pragma unroll
        for (int k = 0; k < 9; ++k) m[k] = R[s * 9 + k];
        const int pa[3] = {0, 0, 1}, pb[3] = {1, 2, 2};                           // column pairs (c0,c1), (c0,c2), (c1,c2)
#pragma unroll
        for (int q = 0; q < 3; ++q) {
            const int a = pa[q], b = pb[q];
            r[3 + q] = w_rot * (m[a] * m[b] + m[3 + a] * m[3 + b] + m[6 + a] * m[6 + b]);
#pragma unroll
            for (int i = 0; i < 3; ++i) { J[3 + q][3 + 3 * i + a] = w_rot * m[3 * i + b]; J[3 + q][3 + 3 * i + b] = w_rot * m[3 * i + a]; }
        }
#pragma unroll
        for (int j = 0; j < 3; ++j) {
            r[6 + j] = w_rot * (m[j] * m[j] + m[3 + j] * m[3 + j] + m[6 + j] * m[6 + j] - T(1));
#pragma unroll
            for (int i = 0; i < 3; ++i) J[6 + j][3 + 3 * i + j] = T(2) * w_rot * m[3 * i + j];
        }
    }
    __device__ __forceinline__ void edge_eval(long e, long* v, T* r, T (*J)[24]) const {
        const long a = v0[e], b = v1[e];
        v[0] = a; v[1] = b;
        T du[3];
#pragma unroll
        for (int k = 0; k < 3; ++k) du[k] = Ur[b * 3 + k] - Ur[a * 3 + k];
#pragma unroll
        for (int k = 0; k < 3; ++k) {
            const T* row = R + a * 9 + 3 * k;
            r[k] = w_reg * ((O[b * 3 + k] - O[a * 3 + k]) - (row[0] * du[0] + row[1] * du[1] + row[2] * du[2]));
            J[k][k] = -w_reg;                                                   // O_a
            J[k][12 + k] = w_reg;                                               // O_b
#pragma unroll
            for (int j = 0; j < 3; ++j) J[k][3 + 3 * k + j] = -w_reg * du[j];   // R_a row k
        }
    }
};

}  // namespace optk
