// functors/math.cuh -- small device math shared by the per-example functors.
#pragma once
#include "../common.cuh"

namespace optk {

__device__ __forceinline__ void sincos_t(float a, float& s, float& c) { sincosf(a, &s, &c); }
__device__ __forceinline__ void sincos_t(double a, double& s, double& c) { sincos(a, &s, &c); }

// AoS element loads (caller arrays are AoS per element, x fastest; o.t:564-579)
__device__ __forceinline__ void load2(const float* p, float* o) { float2 v = *reinterpret_cast<const float2*>(p); o[0] = v.x; o[1] = v.y; }
__device__ __forceinline__ void load2(const double* p, double* o) { double2 v = *reinterpret_cast<const double2*>(p); o[0] = v.x; o[1] = v.y; }
__device__ __forceinline__ void store2(float* p, const float* v) { *reinterpret_cast<float2*>(p) = make_float2(v[0], v[1]); }
__device__ __forceinline__ void store2(double* p, const double* v) { *reinterpret_cast<double2*>(p) = make_double2(v[0], v[1]); }
__device__ __forceinline__ void load4(const float* p, float* o) { float4 v = *reinterpret_cast<const float4*>(p); o[0] = v.x; o[1] = v.y; o[2] = v.z; o[3] = v.w; }
__device__ __forceinline__ void load4(const double* p, double* o) {
    double2 a = reinterpret_cast<const double2*>(p)[0], b = reinterpret_cast<const double2*>(p)[1];
    o[0] = a.x; o[1] = a.y; o[2] = b.x; o[3] = b.y;
}
// 4 consecutive elements, 16-byte aligned
__device__ __forceinline__ void st4v(float* p, const float* v) { *reinterpret_cast<float4*>(p) = make_float4(v[0], v[1], v[2], v[3]); }
__device__ __forceinline__ void st4v(double* p, const double* v) {
    reinterpret_cast<double2*>(p)[0] = make_double2(v[0], v[1]);
    reinterpret_cast<double2*>(p)[1] = make_double2(v[2], v[3]);
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

// Pieces of the rearranged ARAP J^T J (functors/volumetric.cuh, functors/arap.cuh): with the sin/cos of the
// Euler angles of a site, R = Rotate3D, B = dR/dbeta, and M = sum_j pA[j] dR/da_j.
__device__ __forceinline__ float nfma(float a, float b, float c) { return fmaf(-a, b, c); }     // c - a b
__device__ __forceinline__ double nfma(double a, double b, double c) { return fma(-a, b, c); }
template <class T>
struct RotM {
    T R[3][3], B[3][3];         // R = Rotate3D(a) (lib.t:84-98), B = dR/dbeta
    __device__ __forceinline__ void set(const T* t) {   // t = sa ca sb cb sg cg
        const T sa = t[0], ca = t[1], sb = t[2], cb = t[3], sg = t[4], cg = t[5];
        const T cgsb = cg * sb, sgsb = sg * sb;
        R[0][0] = cg * cb; R[0][1] = cgsb * sa - sg * ca; R[0][2] = cgsb * ca + sg * sa;
        R[1][0] = sg * cb; R[1][1] = sgsb * sa + cg * ca; R[1][2] = sgsb * ca - cg * sa;
        R[2][0] = -sb;     R[2][1] = cb * sa;             R[2][2] = cb * ca;
        B[0][0] = -cgsb;   B[0][1] = R[0][0] * sa;        B[0][2] = R[0][0] * ca;
        B[1][0] = -sgsb;   B[1][1] = R[1][0] * sa;        B[1][2] = R[1][0] * ca;
        B[2][0] = -cb;     B[2][1] = -sb * sa;            B[2][2] = -sb * ca;
    }
    // M = pa dR/dalpha + pb dR/dbeta + pg dR/dgamma, with dR/dalpha = [0 | R[:,2] | -R[:,1]] and
    // dR/dgamma = [-R[1,:]; R[0,:]; 0]
    __device__ __forceinline__ void combine(T pa, T pb, T pg, T M[3][3]) const {
        M[0][0] = pb * B[0][0] - pg * R[1][0];
        M[0][1] = pa * R[0][2] + pb * B[0][1] - pg * R[1][1];
        M[0][2] = pb * B[0][2] - pa * R[0][1] - pg * R[1][2];
        M[1][0] = pb * B[1][0] + pg * R[0][0];
        M[1][1] = pa * R[1][2] + pb * B[1][1] + pg * R[0][1];
        M[1][2] = pb * B[1][2] - pa * R[1][1] + pg * R[0][2];
        M[2][0] = pb * B[2][0];
        M[2][1] = pa * R[2][2] + pb * B[2][1];
        M[2][2] = pb * B[2][2] - pa * R[2][1];
    }
    // -< dR/da_j, S >_F for j = alpha, beta, gamma
    __device__ __forceinline__ void project(const T S[3][3], T* o) const {
        T a = T(0), b = T(0), g = T(0);
#pragma unroll
        for (int k = 0; k < 3; ++k) {
            a += R[k][2] * S[k][1] - R[k][1] * S[k][2];
            b += B[k][0] * S[k][0] + B[k][1] * S[k][1] + B[k][2] * S[k][2];
            g += R[0][k] * S[1][k] - R[1][k] * S[0][k];
        }
        o[0] = -a; o[1] = -b; o[2] = -g;
    }
};

// M u for M = pa dR/dalpha + pb dR/dbeta + pg dR/dgamma of R = Rz(gamma) Ry(beta) Rx(alpha) (lib.t:84-98), from the
// factors instead of the assembled matrix: with a = Rx u, b = Ry a,  dRx/dalpha u = (0, -a2, a1),
// dRy/dbeta a = (b2, 0, -b0),  dRz/dgamma b = (-(Rz b)_1, (Rz b)_0, 0).   t = sa ca sb cb sg cg.
template <class T>
__device__ __forceinline__ void rot_deriv_apply(const T* t, T pa, T pb, T pg, const T* u, T* o) {
    const T sa = t[0], ca = t[1], sb = t[2], cb = t[3], sg = t[4], cg = t[5];
    const T a1 = ca * u[1] - sa * u[2], a2 = sa * u[1] + ca * u[2];
    const T b0 = cb * u[0] + sb * a2, b2 = cb * a2 - sb * u[0];
    const T d0 = pa * (sb * a1) + pb * b2, d1 = -pa * a2, d2 = pa * (cb * a1) - pb * b0;
    const T z0 = cg * b0 - sg * a1, z1 = sg * b0 + cg * a1;               // (Rz b)_0, (Rz b)_1 with b1 = a1
    o[0] = cg * d0 - sg * d1 - pg * z1;
    o[1] = sg * d0 + cg * d1 + pg * z0;
    o[2] = d2;
}

}  // namespace optk
