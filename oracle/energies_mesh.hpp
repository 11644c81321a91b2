// TEST INFRASTRUCTURE ONLY -- CPU oracle (see oracle/README.md).
// The mesh (graph) example energies beyond ARAP, restated one-to-one from the reference's energy files.
#pragma once
#include "core.hpp"

namespace oracle {

// ------------------------------------------------------------------------------------------
// examples/embedded_mesh_deformation/embedded_mesh_deformation.t:1-34 ; Matrix3x3Mul lib.t:45-50, Dot3 lib.t:52-54
template <class T>
struct EmbeddedMeshDeformation : SpecBase<T> {
    using Scalar = T;
    static constexpr int U = 12, NT = 2, RMAX = 6, GV = 2, NGT = 1;
    static constexpr bool USE_PRE = true;
    static const char* name() { return "embedded_mesh_deformation"; }
    const T *Ur = nullptr, *Cn = nullptr;
    T w_fit = 0, w_reg = 0, w_rot = 0;
    void plan(const unsigned* d) { this->dim[0] = d[0]; this->N = d[0]; this->E = d[1]; }
    void bind(void** pp) {
        w_fit = (T) * (const float*)pp[0]; w_reg = (T) * (const float*)pp[1]; w_rot = (T) * (const float*)pp[2];
        this->nimg = 0; this->add_unknown(pp[3], 3); this->add_unknown(pp[4], 9);
        Ur = (const T*)pp[5]; Cn = (const T*)pp[6];
        this->gidx[0] = (const int*)pp[8]; this->gidx[1] = (const int*)pp[9];   // pp[7] = edge count, ignored (util.t:696)
    }
    bool exclude(int, int, int) const { return false; }
    static int offsets(int, const int (**o)[3]) { ORACLE_OFFS({0, 0, 0}) }
    template <class S, class C> int residual(int k, C& c, S* r) const {
        if (k == 0) {
            bool valid = c.arr(Cn, 3, 0) >= T(-999999.9);
            for (int ch = 0; ch < 3; ++ch) r[ch] = select(valid, w_fit * (c.unk(0, ch) - c.arr(Cn, 3, ch)), S(T(0)));
            return 3;
        }
        Vec<S, 3> c0, c1, c2;
        for (int i = 0; i < 3; ++i) { c0[i] = c.unk(1, 3 * i); c1[i] = c.unk(1, 3 * i + 1); c2[i] = c.unk(1, 3 * i + 2); }
        r[0] = w_rot * dot3(c0, c1);
        r[1] = w_rot * dot3(c0, c2);
        r[2] = w_rot * dot3(c1, c2);
        r[3] = w_rot * (dot3(c0, c0) - T(1));
        r[4] = w_rot * (dot3(c1, c1) - T(1));
        r[5] = w_rot * (dot3(c2, c2) - T(1));
        return 6;
    }
    template <class S, class G> int gresidual(int, G& g, S* r) const {
        S m[9]; Vec<S, 3> dO; T dU[3];
        for (int i = 0; i < 9; ++i) m[i] = g.unk(1, i, 0);
        for (int ch = 0; ch < 3; ++ch) {
            dO[ch] = g.unk(0, ch, 1) - g.unk(0, ch, 0);
            dU[ch] = g.arr(Ur, 3, ch, 1) - g.arr(Ur, 3, ch, 0);
        }
        for (int k = 0; k < 3; ++k)
            r[k] = w_reg * (dO[k] - (m[3 * k] * dU[0] + m[3 * k + 1] * dU[1] + m[3 * k + 2] * dU[2]));
        return 3;
    }
    void precompute_all(int) {}
};

// ------------------------------------------------------------------------------------------
// examples/cotangent_mesh_smoothing/cotangent_mesh_smoothing.t:1-37 ; normalize lib.t:61-63
template <class T>
struct CotangentMeshSmoothing : SpecBase<T> {
    using Scalar = T;
    static constexpr int U = 3, NT = 1, RMAX = 3, GV = 4, NGT = 1;
    static constexpr bool USE_PRE = true;
    static const char* name() { return "cotangent_mesh_smoothing"; }
    const T* A = nullptr;
    T w_fit = 0, w_reg = 0;
    void plan(const unsigned* d) { this->dim[0] = d[0]; this->N = d[0]; this->E = d[1]; }
    void bind(void** pp) {
        w_fit = (T) * (const float*)pp[0]; w_reg = (T) * (const float*)pp[1];
        this->nimg = 0; this->add_unknown(pp[2], 3);
        A = (const T*)pp[3];
        for (int k = 0; k < 4; ++k) this->gidx[k] = (const int*)pp[5 + k];      // pp[4] = edge count, ignored (util.t:696)
    }
    bool exclude(int, int, int) const { return false; }
    static int offsets(int, const int (**o)[3]) { ORACLE_OFFS({0, 0, 0}) }
    template <class S, class C> int residual(int, C& c, S* r) const {
        for (int ch = 0; ch < 3; ++ch) r[ch] = w_fit * (c.unk(0, ch) - c.arr(A, 3, ch));
        return 3;
    }
    template <class S> static Vec<S, 3> normalize(const Vec<S, 3>& v) {
        S l = sqrt(dot3(v, v));
        Vec<S, 3> r; for (int i = 0; i < 3; ++i) r[i] = v[i] / l; return r;
    }
    template <class S> static S cot(const Vec<S, 3>& v0, const Vec<S, 3>& v1) {
        S adotb = dot3(v0, v1);
        S disc = dot3(v0, v0) * dot3(v1, v1) - adotb * adotb;
        disc = select(val(disc) > T(0), disc, S(T(0.0001)));
        return dot3(v0, v1) / sqrt(disc);
    }
    template <class S, class G> int gresidual(int, G& g, S* r) const {
        Vec<S, 3> x[4];
        for (int s = 0; s < 4; ++s) for (int ch = 0; ch < 3; ++ch) x[s][ch] = g.unk(0, ch, s);
        Vec<S, 3> a = normalize(x[0] - x[2]), b = normalize(x[1] - x[2]), c = normalize(x[0] - x[3]), d = normalize(x[1] - x[3]);
        S w = T(0.5) * (cot(a, b) + cot(c, d));
        w = sqrt(select(val(w) > T(0), w, S(T(0.0001))));
        for (int ch = 0; ch < 3; ++ch) r[ch] = w_reg * w * (x[1][ch] - x[0][ch]);
        return 3;
    }
    void precompute_all(int) {}
};

// ------------------------------------------------------------------------------------------
// examples/robust_nonrigid_alignment/robust_nonrigid_alignment.t:1-30.  greatereq(Constraints(0), -999999.9) compares a
// 3-vector: ad broadcasts (ad.t:373-395), so the fit and the confidence term are THREE residuals each.
template <class T>
struct RobustNonrigidAlignment : SpecBase<T> {
    using Scalar = T;
    static constexpr int U = 7, NT = 2, RMAX = 3, GV = 2, NGT = 1;
    static constexpr bool USE_PRE = true;
    static const char* name() { return "robust_nonrigid_alignment"; }
    const T *Ur = nullptr, *Cn = nullptr, *Nn = nullptr;
    T w_fit = 0, w_reg = 0;
    void plan(const unsigned* d) { this->dim[0] = d[0]; this->N = d[0]; this->E = d[1]; }
    void bind(void** pp) {
        w_fit = (T) * (const float*)pp[0]; w_reg = (T) * (const float*)pp[1];
        this->nimg = 0; this->add_unknown(pp[2], 3); this->add_unknown(pp[3], 3); this->add_unknown(pp[4], 1);
        Ur = (const T*)pp[5]; Cn = (const T*)pp[6]; Nn = (const T*)pp[7];
        this->gidx[0] = (const int*)pp[9]; this->gidx[1] = (const int*)pp[10];  // pp[8] = edge count, ignored (util.t:696)
    }
    bool exclude(int, int, int) const { return false; }
    static int offsets(int, const int (**o)[3]) { ORACLE_OFFS({0, 0, 0}) }
    template <class S, class C> int residual(int k, C& c, S* r) const {
        S w = c.unk(2, 0);
        if (k == 0) {
            S e_fit = w * (c.arr(Nn, 3, 0) * (c.unk(0, 0) - c.arr(Cn, 3, 0)) + c.arr(Nn, 3, 1) * (c.unk(0, 1) - c.arr(Cn, 3, 1)) +
                           c.arr(Nn, 3, 2) * (c.unk(0, 2) - c.arr(Cn, 3, 2)));
            for (int ch = 0; ch < 3; ++ch) r[ch] = w_fit * select(c.arr(Cn, 3, ch) >= T(-999999.9), e_fit, S(T(0)));
            return 3;
        }
        S e_conf = T(1) - (w * w);
        for (int ch = 0; ch < 3; ++ch) r[ch] = T(0.1) * select(c.arr(Cn, 3, ch) >= T(-999999.9), e_conf, S(T(0)));
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
