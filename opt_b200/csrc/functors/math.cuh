// functors/math.cuh -- small device math shared by the per-example functors.
#pragma once
#include "../common.cuh"

namespace optk {

__device__ __forceinline__ void sincos_t(float a, float& s, float& c) { sincosf(a, &s, &c); }
__device__ __forceinline__ void sincos_t(double a, double& s, double& c) { sincos(a, &s, &c); }

// AoS element loads (caller arrays are AoS per element, x fastest; o.t:564-579)
__device__ __forceinline__ void load2(const float* p, float* o) { float2 v = *reinterpret_cast<const float2*>(p); o[0] = v.x; o[1] = v.y; }
__device__ __forceinline__ void load2(const double* p, double* o) { double2 v = *reinterpret_cast<const double2*>(p); o[0] = v.x; o[1] = v.y; }
__device__ __forceinline__ void load4(const float* p, float* o) { float4 v = *reinterpret_cast<const float4*>(p); o[0] = v.x; o[1] = v.y; o[2] = v.z; o[3] = v.w; }
__device__ __forceinline__ void load4(const double* p, double* o) {
    double2 a = reinterpret_cast<const double2*>(p)[0], b = reinterpret_cast<const double2*>(p)[1];
    o[0] = a.x; o[1] = a.y; o[2] = b.x; o[3] = b.y;
}
template <class T> __device__ __forceinline__ void load3(const T* p, T* o) { o[0] = p[0]; o[1] = p[1]; o[2] = p[2]; }

// Rotate3D of lib.t:84-98 (Z-Y-X Euler angles a = (alpha,beta,gamma), row-major R) applied to u,
// and D = -[dR/dalpha u | dR/dbeta u | dR/dgamma u]  (D[k][j] = d(residual_k)/d(angle_j) for the
// ARAP residual (O_a - O_b) - R(A_a) u).
template <class T>
struct Rot3 {
    T sa, ca, sb, cb, sg, cg;
    __device__ __forceinline__ void set(const T* a) { sincos_t(a[0], sa, ca); sincos_t(a[1], sb, cb); sincos_t(a[2], sg, cg); }
    __device__ __forceinline__ void apply(const T* u, T* o) const {
        o[0] = (cg * cb) * u[0] + (-sg * ca + cg * sb * sa) * u[1] + (sg * sa + cg * sb * ca) * u[2];
        o[1] = (sg * cb) * u[0] + (cg * ca + sg * sb * sa) * u[1] + (-cg * sa + sg * sb * ca) * u[2];
        o[2] = (-sb) * u[0] + (cb * sa) * u[1] + (cb * ca) * u[2];
    }
    __device__ __forceinline__ void jac(const T* u, T D[3][3]) const {
        // d/dalpha
        D[0][0] = -((sg * sa + cg * sb * ca) * u[1] + (sg * ca - cg * sb * sa) * u[2]);
        D[1][0] = -((-cg * sa + sg * sb * ca) * u[1] + (-cg * ca - sg * sb * sa) * u[2]);
        D[2][0] = -((cb * ca) * u[1] + (-cb * sa) * u[2]);
        // d/dbeta
        D[0][1] = -((-cg * sb) * u[0] + (cg * cb * sa) * u[1] + (cg * cb * ca) * u[2]);
        D[1][1] = -((-sg * sb) * u[0] + (sg * cb * sa) * u[1] + (sg * cb * ca) * u[2]);
        D[2][1] = -((-cb) * u[0] + (-sb * sa) * u[1] + (-sb * ca) * u[2]);
        // d/dgamma
        D[0][2] = -((-sg * cb) * u[0] + (-cg * ca - sg * sb * sa) * u[1] + (cg * sa - sg * sb * ca) * u[2]);
        D[1][2] = -((cg * cb) * u[0] + (-sg * ca + cg * sb * sa) * u[1] + (sg * sa + cg * sb * ca) * u[2]);
        D[2][2] = T(0);
    }
};

}  // namespace optk
