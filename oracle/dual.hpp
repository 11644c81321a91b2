// TEST INFRASTRUCTURE ONLY -- part of the CPU oracle (see oracle/README.md).
// Forward-mode dual numbers used to obtain the residual Jacobians that the reference
// obtains symbolically (reference: API/src/ad.t:665-758 `d()/gradient()`, op partials
// ad.t:760-889).  Semantics that matter for parity:
//   * select(c,a,b)' = c ? a' : b'            (ad.t:776-791)
//   * comparisons / eq / not have zero derivative (they only look at the value)
//   * abs' = select(x>=0, 1, -1)              (ad.t:858-889)
#pragma once
#include <cmath>

namespace oracle {

template <class T, int N>
struct Dual {
    T v;
    T d[N];
    Dual() : v(0) { for (int i = 0; i < N; ++i) d[i] = 0; }
    Dual(T x) : v(x) { for (int i = 0; i < N; ++i) d[i] = 0; }
};

// value extraction works for plain scalars and duals
inline float  val(float x) { return x; }
inline double val(double x) { return x; }
template <class T, int N> inline T val(const Dual<T, N>& x) { return x.v; }

#define ORACLE_DUAL_BIN(OP, VEXPR, DEXPR)                                              \
    template <class T, int N> inline Dual<T, N> operator OP(const Dual<T, N>& a, const Dual<T, N>& b) { \
        Dual<T, N> r; r.v = VEXPR; for (int i = 0; i < N; ++i) r.d[i] = DEXPR; return r; }

ORACLE_DUAL_BIN(+, a.v + b.v, a.d[i] + b.d[i])
ORACLE_DUAL_BIN(-, a.v - b.v, a.d[i] - b.d[i])
ORACLE_DUAL_BIN(*, a.v * b.v, a.d[i] * b.v + a.v * b.d[i])
#undef ORACLE_DUAL_BIN

template <class T, int N> inline Dual<T, N> operator/(const Dual<T, N>& a, const Dual<T, N>& b) {
    Dual<T, N> r; T inv = T(1) / b.v; r.v = a.v * inv;
    for (int i = 0; i < N; ++i) r.d[i] = (a.d[i] - r.v * b.d[i]) * inv;
    return r;
}
template <class T, int N> inline Dual<T, N> operator-(const Dual<T, N>& a) {
    Dual<T, N> r; r.v = -a.v; for (int i = 0; i < N; ++i) r.d[i] = -a.d[i]; return r;
}
// mixed scalar forms
template <class T, int N> inline Dual<T, N> operator+(const Dual<T, N>& a, T b) { Dual<T, N> r = a; r.v += b; return r; }
template <class T, int N> inline Dual<T, N> operator+(T b, const Dual<T, N>& a) { Dual<T, N> r = a; r.v += b; return r; }
template <class T, int N> inline Dual<T, N> operator-(const Dual<T, N>& a, T b) { Dual<T, N> r = a; r.v -= b; return r; }
template <class T, int N> inline Dual<T, N> operator-(T b, const Dual<T, N>& a) { Dual<T, N> r = -a; r.v += b; return r; }
template <class T, int N> inline Dual<T, N> operator*(const Dual<T, N>& a, T b) {
    Dual<T, N> r; r.v = a.v * b; for (int i = 0; i < N; ++i) r.d[i] = a.d[i] * b; return r; }
template <class T, int N> inline Dual<T, N> operator*(T b, const Dual<T, N>& a) { return a * b; }
template <class T, int N> inline Dual<T, N> operator/(const Dual<T, N>& a, T b) { return a * (T(1) / b); }
template <class T, int N> inline Dual<T, N> operator/(T a, const Dual<T, N>& b) { return Dual<T, N>(a) / b; }

template <class T, int N> inline Dual<T, N> sqrt(const Dual<T, N>& a) {
    Dual<T, N> r; r.v = std::sqrt(a.v); T k = T(0.5) / r.v;     // ad.t: d sqrt(x) = 1/(2 sqrt(x))
    for (int i = 0; i < N; ++i) r.d[i] = a.d[i] * k; return r;
}
template <class T, int N> inline Dual<T, N> sin(const Dual<T, N>& a) {
    Dual<T, N> r; r.v = std::sin(a.v); T c = std::cos(a.v);
    for (int i = 0; i < N; ++i) r.d[i] = a.d[i] * c; return r;
}
template <class T, int N> inline Dual<T, N> cos(const Dual<T, N>& a) {
    Dual<T, N> r; r.v = std::cos(a.v); T s = -std::sin(a.v);
    for (int i = 0; i < N; ++i) r.d[i] = a.d[i] * s; return r;
}
template <class T, int N> inline Dual<T, N> abs(const Dual<T, N>& a) {
    return a.v >= T(0) ? a : -a;
}
// pow with a constant (non-differentiated) exponent, as used by lib.t:113-122
template <class T, int N> inline Dual<T, N> pow(const Dual<T, N>& a, T e) {
    Dual<T, N> r; r.v = std::pow(a.v, e); T k = e * std::pow(a.v, e - T(1));
    for (int i = 0; i < N; ++i) r.d[i] = a.d[i] * k; return r;
}
inline float  sqrt(float x) { return std::sqrt(x); }
inline double sqrt(double x) { return std::sqrt(x); }
inline float  sin(float x) { return std::sin(x); }
inline double sin(double x) { return std::sin(x); }
inline float  cos(float x) { return std::cos(x); }
inline double cos(double x) { return std::cos(x); }
inline float  abs(float x) { return std::fabs(x); }
inline double abs(double x) { return std::fabs(x); }
inline float  pow(float x, float e) { return std::pow(x, e); }
inline double pow(double x, double e) { return std::pow(x, e); }

template <class S> inline S select(bool c, const S& a, const S& b) { return c ? a : b; }

// small fixed vector of scalars S
template <class S, int K>
struct Vec {
    S e[K];
    S& operator[](int i) { return e[i]; }
    const S& operator[](int i) const { return e[i]; }
};
template <class S, int K> inline Vec<S, K> operator+(const Vec<S, K>& a, const Vec<S, K>& b) { Vec<S, K> r; for (int i = 0; i < K; ++i) r[i] = a[i] + b[i]; return r; }
template <class S, int K> inline Vec<S, K> operator-(const Vec<S, K>& a, const Vec<S, K>& b) { Vec<S, K> r; for (int i = 0; i < K; ++i) r[i] = a[i] - b[i]; return r; }
template <class S, int K, class W> inline Vec<S, K> operator*(const W& w, const Vec<S, K>& a) { Vec<S, K> r; for (int i = 0; i < K; ++i) r[i] = w * a[i]; return r; }
template <class S> inline S dot3(const Vec<S, 3>& a, const Vec<S, 3>& b) { return a[0] * b[0] + a[1] * b[1] + a[2] * b[2]; }

}  // namespace oracle
