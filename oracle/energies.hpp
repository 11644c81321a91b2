// TEST INFRASTRUCTURE ONLY -- CPU oracle (see oracle/README.md).
// The shipped example energies, restated one-to-one from the reference's energy files
// (examples/<ex>/<ex>.t).  Each `residual`/`gresidual` body is the expression handed to
// `Energy(...)` in the file, written over a scalar type S that is either opt_float or a dual
// number; derivatives therefore come from AD exactly as in the reference, and the CUDA
// functors (hand-derived closed forms) are checked against them.
#pragma once
#include "core.hpp"

namespace oracle {

#define ORACLE_OFFS(...)                                   \
    static const int tab[][3] = {__VA_ARGS__};             \
    *o = tab;                                              \
    return (int)(sizeof(tab) / sizeof(tab[0]));

// 4-neighbourhood in the order every 2-D example lists it: Stencil{{1,0},{-1,0},{0,1},{0,-1}}
static const int kSt2[4][2] = {{1, 0}, {-1, 0}, {0, 1}, {0, -1}};
// volumetric_mesh_deformation.t:16
static const int kSt3[6][3] = {{1, 0, 0}, {-1, 0, 0}, {0, 1, 0}, {0, -1, 0}, {0, 0, 1}, {0, 0, -1}};

// lib.t:99-103
template <class S> inline void rotate2d(const S& a, const S& v0, const S& v1, S& o0, S& o1) {
    S c = cos(a), s = sin(a);
    o0 = c * v0 + (-s) * v1;
    o1 = s * v0 + c * v1;
}
// lib.t:84-98 (Z-Y-X Euler, row-major) followed by Matrix3x3Mul (lib.t:40-45)
template <class S> inline Vec<S, 3> rotate3d(const Vec<S, 3>& a, const Vec<S, 3>& v) {
    S ca = cos(a[0]), cb = cos(a[1]), cg = cos(a[2]), sa = sin(a[0]), sb = sin(a[1]), sg = sin(a[2]);
    S m[9] = {cg * cb, -sg * ca + cg * sb * sa, sg * sa + cg * sb * ca,
              sg * cb, cg * ca + sg * sb * sa,  -cg * sa + sg * sb * ca,
              -sb,     cb * sa,                 cb * ca};
    Vec<S, 3> r;
    r[0] = m[0] * v[0] + m[1] * v[1] + m[2] * v[2];
    r[1] = m[3] * v[0] + m[4] * v[1] + m[5] * v[2];
    r[2] = m[6] * v[0] + m[7] * v[1] + m[8] * v[2];
    return r;
}

template <class T> struct CenteredOnly {
    static constexpr int GV = 0, NGT = 0;
    template <class S, class G> int gresidual(int, G&, S*) const { return 0; }
};

// ------------------------------------------------------------------------------------------
// examples/poisson_image_editing/poisson_image_editing.t:1-13
template <class T>
struct PoissonImageEditing : SpecBase<T>, CenteredOnly<T> {
    using Scalar = T;
    static constexpr int U = 4, NT = 4, RMAX = 4;
    static constexpr bool USE_PRE = false;
    static const char* name() { return "poisson_image_editing"; }
    const T* Tm = nullptr; const T* M = nullptr;
    void plan(const unsigned* d) { this->dim[0] = d[0]; this->dim[1] = d[1]; this->N = (long)d[0] * d[1]; }
    void bind(void** pp) { this->nimg = 0; this->add_unknown(pp[0], 4); Tm = (const T*)pp[1]; M = (const T*)pp[2]; }
    bool exclude(int x, int y, int z) const { return !(M[this->lin(x, y, z)] == T(0)); }
    static int offsets(int k, const int (**o)[3]) {
        static const int tab[4][2][3] = {{{0, 0, 0}, {1, 0, 0}}, {{0, 0, 0}, {-1, 0, 0}}, {{0, 0, 0}, {0, 1, 0}}, {{0, 0, 0}, {0, -1, 0}}};
        *o = tab[k]; return 2;
    }
    template <class S, class C> int residual(int k, C& c, S* r) const {
        const int ox = kSt2[k][0], oy = kSt2[k][1];
        for (int ch = 0; ch < 4; ++ch) {
            S e = (c.unk(0, ch) - c.unk(0, ch, ox, oy)) - S(c.arr(Tm, 4, ch) - c.arr(Tm, 4, ch, ox, oy));
            r[ch] = select(c.inb(ox, oy), e, S(T(0)));
        }
        return 4;
    }
    void precompute_all(int) {}
};

// ------------------------------------------------------------------------------------------
// examples/image_warping/image_warping.t:12-43,105-107
template <class T>
struct ImageWarping : SpecBase<T>, CenteredOnly<T> {
    using Scalar = T;
    static constexpr int U = 3, NT = 5, RMAX = 2;
    static constexpr bool USE_PRE = true;
    static const char* name() { return "image_warping"; }
    const T *Ur = nullptr, *Cn = nullptr, *Mask = nullptr;
    T w_fit = 0, w_reg = 0;
    void plan(const unsigned* d) { this->dim[0] = d[0]; this->dim[1] = d[1]; this->N = (long)d[0] * d[1]; }
    void bind(void** pp) {
        this->nimg = 0; this->add_unknown(pp[0], 2); this->add_unknown(pp[1], 1);
        Ur = (const T*)pp[2]; Cn = (const T*)pp[3]; Mask = (const T*)pp[4];
        w_fit = (T) * (const float*)pp[5]; w_reg = (T) * (const float*)pp[6];
    }
    bool exclude(int x, int y, int z) const { return !(Mask[this->lin(x, y, z)] == T(0)); }
    static int offsets(int k, const int (**o)[3]) {
        static const int tab[5][2][3] = {{{0, 0, 0}, {1, 0, 0}}, {{0, 0, 0}, {-1, 0, 0}}, {{0, 0, 0}, {0, 1, 0}}, {{0, 0, 0}, {0, -1, 0}}, {{0, 0, 0}, {0, 0, 0}}};
        *o = tab[k]; return k < 4 ? 2 : 1;
    }
    template <class S, class C> int residual(int k, C& c, S* r) const {
        if (k < 4) {
            const int ox = kSt2[k][0], oy = kSt2[k][1];
            S d0 = c.unk(0, 0) - c.unk(0, 0, ox, oy), d1 = c.unk(0, 1) - c.unk(0, 1, ox, oy);
            S u0 = S(c.arr(Ur, 2, 0) - c.arr(Ur, 2, 0, ox, oy)), u1 = S(c.arr(Ur, 2, 1) - c.arr(Ur, 2, 1, ox, oy));
            S q0, q1; rotate2d(c.unk(1, 0), u0, u1, q0, q1);
            bool valid = c.inb(ox, oy) && (c.arr(Mask, 1, 0, ox, oy) == T(0)) && (c.arr(Mask, 1, 0) == T(0));
            r[0] = select(valid, w_reg * (d0 - q0), S(T(0)));
            r[1] = select(valid, w_reg * (d1 - q1), S(T(0)));
        } else {
            T c0 = c.arr(Cn, 2, 0), c1 = c.arr(Cn, 2, 1);
            bool valid = (c0 >= T(0)) && (c1 >= T(0));
            r[0] = w_fit * select(valid, c.unk(0, 0) - c0, S(T(0)));
            r[1] = w_fit * select(valid, c.unk(0, 1) - c1, S(T(0)));
        }
        return 2;
    }
    void precompute_all(int) {}
};

// ------------------------------------------------------------------------------------------
// examples/volumetric_mesh_deformation/volumetric_mesh_deformation.t:1-21
template <class T>
struct VolumetricMeshDeformation : SpecBase<T>, CenteredOnly<T> {
    using Scalar = T;
    static constexpr int U = 6, NT = 7, RMAX = 3;
    static constexpr bool USE_PRE = true;
    static const char* name() { return "volumetric_mesh_deformation"; }
    const T *Ur = nullptr, *Cn = nullptr;
    T w_fit = 0, w_reg = 0;
    void plan(const unsigned* d) { for (int i = 0; i < 3; ++i) this->dim[i] = d[i]; this->N = (long)d[0] * d[1] * d[2]; }
    void bind(void** pp) {
        this->nimg = 0; this->add_unknown(pp[0], 3); this->add_unknown(pp[1], 3);
        Ur = (const T*)pp[2]; Cn = (const T*)pp[3];
        w_fit = (T) * (const float*)pp[4]; w_reg = (T) * (const float*)pp[5];
    }
    bool exclude(int, int, int) const { return false; }
    static int offsets(int k, const int (**o)[3]) {
        static const int tab[7][2][3] = {{{0, 0, 0}, {0, 0, 0}},
                                         {{0, 0, 0}, {1, 0, 0}}, {{0, 0, 0}, {-1, 0, 0}}, {{0, 0, 0}, {0, 1, 0}},
                                         {{0, 0, 0}, {0, -1, 0}}, {{0, 0, 0}, {0, 0, 1}}, {{0, 0, 0}, {0, 0, -1}}};
        *o = tab[k]; return k == 0 ? 1 : 2;
    }
    template <class S, class C> int residual(int k, C& c, S* r) const {
        if (k == 0) {
            bool valid = c.arr(Cn, 3, 0) >= T(-999999.9);
            for (int ch = 0; ch < 3; ++ch) r[ch] = select(valid, w_fit * (c.unk(0, ch) - c.arr(Cn, 3, ch)), S(T(0)));
        } else {
            const int* o = kSt3[k - 1];
            Vec<S, 3> dO, dU, A;
            for (int ch = 0; ch < 3; ++ch) {
                dO[ch] = c.unk(0, ch) - c.unk(0, ch, o[0], o[1], o[2]);
                dU[ch] = S(c.arr(Ur, 3, ch) - c.arr(Ur, 3, ch, o[0], o[1], o[2]));
                A[ch] = c.unk(1, ch);
            }
            Vec<S, 3> e = dO - rotate3d(A, dU);
            bool valid = c.inb(0, 0, 0) && c.inb(o[0], o[1], o[2]);
            for (int ch = 0; ch < 3; ++ch) r[ch] = w_reg * select(valid, e[ch], S(T(0)));
        }
        return 3;
    }
    void precompute_all(int) {}
};

// ------------------------------------------------------------------------------------------
// examples/optical_flow/optical_flow.t:1-20 ; sampling o.t:856-876, derivative o.t:3267-3282
template <class T>
struct OpticalFlow : SpecBase<T>, CenteredOnly<T> {
    using Scalar = T;
    static constexpr int U = 2, NT = 5, RMAX = 2;
    static constexpr bool USE_PRE = false;
    static const char* name() { return "optical_flow"; }
    const T *I = nullptr, *Ih = nullptr, *Idx = nullptr, *Idy = nullptr;
    T w_fit = 0, w_reg = 0;
    void plan(const unsigned* d) { this->dim[0] = d[0]; this->dim[1] = d[1]; this->N = (long)d[0] * d[1]; }
    void bind(void** pp) {
        w_fit = (T) * (const float*)pp[0]; w_reg = (T) * (const float*)pp[1];
        this->nimg = 0; this->add_unknown(pp[2], 2);
        I = (const T*)pp[3]; Ih = (const T*)pp[4]; Idx = (const T*)pp[5]; Idy = (const T*)pp[6];
    }
    bool exclude(int, int, int) const { return false; }
    static int offsets(int k, const int (**o)[3]) {
        static const int tab[5][2][3] = {{{0, 0, 0}, {0, 0, 0}}, {{0, 0, 0}, {1, 0, 0}}, {{0, 0, 0}, {-1, 0, 0}}, {{0, 0, 0}, {0, 1, 0}}, {{0, 0, 0}, {0, -1, 0}}};
        *o = tab[k]; return k == 0 ? 1 : 2;
    }
    T get(const T* im, int x, int y) const { return this->inside(x, y, 0) ? im[this->lin(x, y, 0)] : T(0); }
    T sample(const T* im, T x, T y) const {
        int x0 = (int)std::floor(x), x1 = (int)std::ceil(x), y0 = (int)std::floor(y), y1 = (int)std::ceil(y);
        T xn = x - x0, yn = y - y0;
        T u = (T(1) - xn) * get(im, x0, y0) + xn * get(im, x1, y0);
        T b = (T(1) - xn) * get(im, x0, y1) + xn * get(im, x1, y1);
        return (T(1) - yn) * u + yn * b;
    }
    // sampled image applied to scalar or dual coordinates: value from `im`, partials from dx/dy images
    T sampled(const T& x, const T& y) const { return sample(Ih, x, y); }
    template <int N> Dual<T, N> sampled(const Dual<T, N>& x, const Dual<T, N>& y) const {
        Dual<T, N> r(sample(Ih, x.v, y.v));
        T gx = sample(Idx, x.v, y.v), gy = sample(Idy, x.v, y.v);
        for (int i = 0; i < N; ++i) r.d[i] = gx * x.d[i] + gy * y.d[i];
        return r;
    }
    template <class S, class C> int residual(int k, C& c, S* r) const {
        if (k == 0) {
            S px = T(c.cx) + c.unk(0, 0), py = T(c.cy) + c.unk(0, 1);
            r[0] = w_fit * (c.arr(I, 1, 0) - sampled(px, py));
            return 1;
        }
        const int ox = kSt2[k - 1][0], oy = kSt2[k - 1][1];
        for (int ch = 0; ch < 2; ++ch)
            r[ch] = select(c.inb(ox, oy), w_reg * (c.unk(0, ch) - c.unk(0, ch, ox, oy)), S(T(0)));
        return 2;
    }
    void precompute_all(int) {}
};

// ------------------------------------------------------------------------------------------
// examples/arap_mesh_deformation/arap_mesh_deformation.t:11-48
template <class T>
struct ArapMeshDeformation : SpecBase<T> {
    using Scalar = T;
    static constexpr int U = 6, NT = 1, RMAX = 3, GV = 2, NGT = 1;
    static constexpr bool USE_PRE = true;
    static const char* name() { return "arap_mesh_deformation"; }
    const T *Ur = nullptr, *Cn = nullptr;
    T w_fit = 0, w_reg = 0;
    void plan(const unsigned* d) { this->dim[0] = d[0]; this->N = d[0]; this->E = d[1]; }
    void bind(void** pp) {
        w_fit = (T) * (const float*)pp[0]; w_reg = (T) * (const float*)pp[1];
        this->nimg = 0; this->add_unknown(pp[2], 3); this->add_unknown(pp[3], 3);
        Ur = (const T*)pp[4]; Cn = (const T*)pp[5];
        this->gidx[0] = (const int*)pp[7]; this->gidx[1] = (const int*)pp[8];   // pp[6] = edge count, ignored (util.t:696)
    }
    bool exclude(int, int, int) const { return false; }
    static int offsets(int, const int (**o)[3]) { ORACLE_OFFS({0, 0, 0}) }
    template <class S, class C> int residual(int, C& c, S* r) const {
        bool valid = c.arr(Cn, 3, 0) >= T(-999999.9);
        for (int ch = 0; ch < 3; ++ch) r[ch] = select(valid, w_fit * (c.unk(0, ch) - c.arr(Cn, 3, ch)), S(T(0)));
        return 3;
    }
    template <class S, class G> int gresidual(int, G& g, S* r) const {
        Vec<S, 3> dO, dU, A;
        for (int ch = 0; ch < 3; ++ch) {
            dO[ch] = g.unk(0, ch, 0) - g.unk(0, ch, 1);
            dU[ch] = S(g.arr(Ur, 3, ch, 0) - g.arr(Ur, 3, ch, 1));
            A[ch] = g.unk(1, ch, 0);
        }
        Vec<S, 3> e = dO - rotate3d(A, dU);
        for (int ch = 0; ch < 3; ++ch) r[ch] = w_reg * e[ch];
        return 3;
    }
    void precompute_all(int) {}
};

}  // namespace oracle

#include "energies_sfs.hpp"
#include "energies_mesh.hpp"

#define ORACLE_ENERGY_LIST(X)      \
    X(oracle::ShapeFromShading)    \
    X(oracle::EmbeddedMeshDeformation) \
    X(oracle::CotangentMeshSmoothing) \
    X(oracle::RobustNonrigidAlignment) \
    X(oracle::IntrinsicImageDecomposition) \
    X(oracle::PoissonImageEditing) \
    X(oracle::ImageWarping)        \
    X(oracle::VolumetricMeshDeformation) \
    X(oracle::OpticalFlow)         \
    X(oracle::ArapMeshDeformation)
