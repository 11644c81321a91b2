// TEST INFRASTRUCTURE ONLY -- CPU oracle (see oracle/README.md).
// examples/shape_from_shading/shape_from_shading.t:1-90, restated over a scalar type S (opt_float or dual).
// ComputedArrays (`B_I`, `valid`) are materialised by precompute_all exactly like fmap.precompute
// (o.t:3131-3153): value image + one gradient image per distinct unknown access of the expression
// (o.t:1686-1719); residuals read them back through CenterCtx::computed, so derivatives are STALE between
// precomputes by construction, as in the reference.
#pragma once
#include "core.hpp"

namespace oracle {

template <class T>
struct ShapeFromShading : SpecBase<T>, CenteredOnly<T> {
    using Scalar = T;
    static constexpr int U = 1, NT = 4, RMAX = 3;
    static constexpr bool USE_PRE = false;                  // no UsePreconditioner() in the file: default off
    static const char* name() { return "shape_from_shading"; }
    T w_p = 0, w_s = 0, w_g = 0, f_x = 0, f_y = 0, u_x = 0, u_y = 0, L[9];
    const T *D_i = nullptr, *Im = nullptr;
    const unsigned char *maskR = nullptr, *maskC = nullptr;
    std::vector<T> BI, G[3], validA;                        // ComputedArray B_I (+ gradients wrt X(0,0), X(-1,0), X(0,-1)), valid
    void plan(const unsigned* d) { this->dim[0] = d[0]; this->dim[1] = d[1]; this->N = (long)d[0] * d[1]; }
    void bind(void** pp) {
        // sqrt(Param(...)) is evaluated in opt_float on the float parameter (o.t:2430)
        w_p = std::sqrt((T) * (const float*)pp[0]); w_s = std::sqrt((T) * (const float*)pp[1]); w_g = std::sqrt((T) * (const float*)pp[2]);
        f_x = (T) * (const float*)pp[3]; f_y = (T) * (const float*)pp[4]; u_x = (T) * (const float*)pp[5]; u_y = (T) * (const float*)pp[6];
        for (int i = 0; i < 9; ++i) L[i] = (T) * (const float*)pp[7 + i];
        this->nimg = 0; this->add_unknown(pp[16], 1);
        D_i = (const T*)pp[17]; Im = (const T*)pp[18];
        maskR = (const unsigned char*)pp[19]; maskC = (const unsigned char*)pp[20];
    }
    T get(const T* im, int x, int y) const { return this->inside(x, y, 0) ? im[this->lin(x, y, 0)] : T(0); }
    bool depth_valid(int x, int y) const { return get(D_i, x, y) > T(0); }
    bool exclude(int x, int y, int) const { return !depth_valid(x, y); }     // Exclude(Not(DepthValid(0,0)))
    bool inb_exp1(int x, int y) const { return this->inside(x - 1, y - 1, 0) && this->inside(x + 1, y + 1, 0); }

    static int offsets(int k, const int (**o)[3]) {
        static const int tab[4][5][3] = {
            {{0, 0, 0}},
            {{0, 0, 0}, {-1, 0, 0}, {0, -1, 0}, {1, 0, 0}, {1, -1, 0}},      // B_I(0,0), B_I(1,0)
            {{0, 0, 0}, {-1, 0, 0}, {0, -1, 0}, {0, 1, 0}, {-1, 1, 0}},      // B_I(0,0), B_I(0,1)
            {{0, 0, 0}, {-1, 0, 0}, {0, -1, 0}, {1, 0, 0}, {0, 1, 0}}};
        *o = tab[k]; return k == 0 ? 1 : 5;
    }

    // shape_from_shading.t:33-66 at pixel (i,j) with d00 = X(0,0), dm0 = X(-1,0), d0m = X(0,-1)
    template <class S> S shading_minus_intensity(int i, int j, const S& d00, const S& dm0, const S& d0m) const {
        S n_x = d0m * (d00 - dm0) / f_y;
        S n_y = dm0 * (d00 - d0m) / f_x;
        S n_z = (n_x * (u_x - T(i)) / f_x) + (n_y * (u_y - T(j)) / f_y) - (dm0 * d0m / (f_x * f_y));
        S sq = n_x * n_x + n_y * n_y + n_z * n_z;
        S inv = select(val(sq) > T(0), T(1) / sqrt(sq), S(T(1)));
        n_x = inv * n_x; n_y = inv * n_y; n_z = inv * n_z;
        S B = L[0] + L[1] * n_y + L[2] * n_z + L[3] * n_x + L[4] * n_x * n_y + L[5] * n_y * n_z +
              L[6] * (-(n_x * n_x) - n_y * n_y + T(2) * n_z * n_z) + L[7] * n_z * n_x + L[8] * (n_x * n_x - n_y * n_y);
        T I = get(Im, i, j) * T(0.5) + T(0.25) * (get(Im, i - 1, j) + get(Im, i, j - 1));
        return B - I;
    }
    void precompute_all(int threads) {
        const int W = this->dim[0], H = this->dim[1];
        BI.assign(this->N, T(0)); validA.assign(this->N, T(0));
        for (auto& g : G) g.assign(this->N, T(0));
        const T* X = (const T*)this->img[0].ptr;
#pragma omp parallel for num_threads(threads) schedule(static)
        for (int j = 0; j < H; ++j)
            for (int i = 0; i < W; ++i) {
                using D3 = Dual<T, 3>;
                D3 d00(get(X, i, j)), dm0(get(X, i - 1, j)), d0m(get(X, i, j - 1));
                d00.d[0] = T(1); dm0.d[1] = T(1); d0m.d[2] = T(1);
                D3 bi = shading_minus_intensity(i, j, d00, dm0, d0m);
                bool ok = inb_exp1(i, j) && depth_valid(i - 1, j) && depth_valid(i, j) && depth_valid(i, j - 1);
                long q = this->lin(i, j, 0);
                BI[q] = ok ? bi.v : T(0);
                for (int k = 0; k < 3; ++k) G[k][q] = ok ? bi.d[k] : T(0);
                // shape_from_shading.t:80-84
                auto cont = [&](int x, int y) { return std::fabs(get(X, i, j) - get(X, i + x, j + y)) < T(0.01); };
                bool v = depth_valid(i, j) && depth_valid(i, j - 1) && depth_valid(i, j + 1) && depth_valid(i - 1, j) && depth_valid(i + 1, j) &&
                         cont(0, -1) && cont(0, 1) && cont(-1, 0) && cont(1, 0) && inb_exp1(i, j);
                validA[q] = v ? T(1) : T(0);
            }
    }
    template <class S, class C> S B_I(C& c, int ox, int oy) const {
        static const int goff[3][2] = {{0, 0}, {-1, 0}, {0, -1}};
        const T* g[3] = {G[0].data(), G[1].data(), G[2].data()};
        return c.computed(BI.data(), 3, g, goff, 0, 0, ox, oy);
    }
    template <class S, class C> int residual(int k, C& c, S* r) const {
        const S zero(T(0));
        if (k == 0) {                                           // fitting term
            bool v = c.arr(D_i, 1, 0) > T(0);
            r[0] = select(v, w_p * (c.unk(0, 0) - c.arr(D_i, 1, 0)), zero);
            return 1;
        }
        if (k == 1 || k == 2) {                                 // shading terms
            const int ox = k == 1 ? 1 : 0, oy = k == 1 ? 0 : 1;
            const unsigned char m = c.arr(k == 1 ? maskR : maskC, 1, 0);
            S e = (B_I<S>(c, 0, 0) - B_I<S>(c, ox, oy)) * T(m);
            r[0] = select(c.inb_expanded(0, 0, 1), w_g * e, zero);
            return 1;
        }
        // regularisation term: 4 p(0,0) - (p(-1,0) + p(0,-1) + p(1,0) + p(0,1)), p = back-projection (eq. 8)
        const bool v = c.arr(validA.data(), 1, 0) == T(1) && c.inb_expanded(0, 0, 1);   // + default bbox wrap (o.t:2704-2708)
        auto P = [&](int ox, int oy) {
            S d = c.unk(0, 0, ox, oy);
            Vec<S, 3> q;
            q[0] = ((T(c.cx + ox) - u_x) / f_x) * d; q[1] = ((T(c.cy + oy) - u_y) / f_y) * d; q[2] = d;
            return q;
        };
        Vec<S, 3> E = T(4) * P(0, 0) - (P(-1, 0) + P(0, -1) + P(1, 0) + P(0, 1));
        for (int ch = 0; ch < 3; ++ch) r[ch] = select(v, w_s * E[ch], zero);
        return 3;
    }
};

// ------------------------------------------------------------------------------------------
// examples/intrinsic_image_decomposition/intrinsic_image_decomposition.t:1-32.  L_p (lib.t:112-121) stores
// sqrt(pow(|r_const(0,0) - r_const(x,y)| + 1e-7, pNorm - 2)) in one ComputedArray per stencil offset; r_const is the
// SAME buffer as the unknown r (both parameter 4) read as a plain Array, so the weights carry no derivative and
// are refreshed by every precompute (o.t:3131-3153).  No UsePreconditioner() call: off.  pNorm is an opt_float Param.
template <class T>
struct IntrinsicImageDecomposition : SpecBase<T>, CenteredOnly<T> {
    using Scalar = T;
    static constexpr int U = 4, NT = 9, RMAX = 3;
    static constexpr bool USE_PRE = false;
    static const char* name() { return "intrinsic_image_decomposition"; }
    T w_fit = 0, w_alb = 0, w_shd = 0, pnorm = 0;
    const T *rc = nullptr, *im = nullptr;
    std::vector<T> Lp[4];
    void plan(const unsigned* d) { this->dim[0] = d[0]; this->dim[1] = d[1]; this->N = (long)d[0] * d[1]; }
    void bind(void** pp) {
        w_fit = (T) * (const float*)pp[0]; w_alb = (T) * (const float*)pp[1]; w_shd = (T) * (const float*)pp[2];
        pnorm = *(const T*)pp[3];
        this->nimg = 0; this->add_unknown(pp[4], 3); this->add_unknown(pp[6], 1);
        rc = (const T*)pp[4]; im = (const T*)pp[5];
    }
    bool exclude(int, int, int) const { return false; }
    static int offsets(int k, const int (**o)[3]) {
        static const int tab[5][2][3] = {{{0, 0, 0}, {1, 0, 0}}, {{0, 0, 0}, {-1, 0, 0}}, {{0, 0, 0}, {0, 1, 0}}, {{0, 0, 0}, {0, -1, 0}}, {{0, 0, 0}, {0, 0, 0}}};
        *o = tab[k < 8 ? k % 4 : 4]; return k < 8 ? 2 : 1;
    }
    T rget(int x, int y, int ch) const { return this->inside(x, y, 0) ? rc[this->lin(x, y, 0) * 3 + ch] : T(0); }
    void precompute_all(int threads) {
        const int W = this->dim[0], H = this->dim[1];
        for (auto& v : Lp) v.assign(this->N, T(0));
#pragma omp parallel for num_threads(threads) schedule(static)
        for (int j = 0; j < H; ++j)
            for (int i = 0; i < W; ++i)
                for (int k = 0; k < 4; ++k) {
                    T d[3];
                    for (int ch = 0; ch < 3; ++ch) d[ch] = rget(i, j, ch) - rget(i + kSt2[k][0], j + kSt2[k][1], ch);
                    T dist = std::sqrt(d[0] * d[0] + d[1] * d[1] + d[2] * d[2]);
                    T C = std::pow(dist + T(0.0000001), pnorm - T(2));
                    Lp[k][this->lin(i, j, 0)] = std::sqrt(C);
                }
    }
    template <class S, class C> int residual(int k, C& c, S* r) const {
        const S zero(T(0));
        if (k < 4) {                                            // reg Albedo
            const int ox = kSt2[k][0], oy = kSt2[k][1];
            const T sq = c.arr(Lp[k].data(), 1, 0);
            const bool v = c.inb(0, 0) && c.inb(ox, oy);
            for (int ch = 0; ch < 3; ++ch) r[ch] = w_alb * select(v, sq * (c.unk(0, ch) - c.unk(0, ch, ox, oy)), zero);
            return 3;
        }
        if (k < 8) {                                            // reg Shading
            const int ox = kSt2[k - 4][0], oy = kSt2[k - 4][1];
            const bool v = c.inb(0, 0) && c.inb(ox, oy);
            r[0] = w_shd * select(v, c.unk(1, 0) - c.unk(1, 0, ox, oy), zero);
            return 1;
        }
        for (int ch = 0; ch < 3; ++ch) r[ch] = w_fit * (c.unk(0, ch) + c.unk(1, 0) - c.arr(im, 3, ch));   // fit
        return 3;
    }
};

}  // namespace oracle
