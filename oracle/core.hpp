// TEST INFRASTRUCTURE ONLY -- CPU oracle for the GN/LM + PCG hot path (see oracle/README.md).
// Nothing in the product (opt_b200/, include/) includes or links this file.
//
// Generic restatement of the reference solver skeleton and of the way the reference derives
// the per-problem functions from an energy:
//   cost        o.t:3119-3129            evalJTF   o.t:2870-2913 (centred) / 2969-2994 (graph)
//   applyJTJ    o.t:2770-2830 / 2833-2867 modelcost o.t:2915-2966  computeCtC o.t:2996-3057
//   kernels     solverGPUGaussNewton.t:521-1345 ; init/step :1766-1897 / :1913-2349
//   loop nest   backend_cpu.t:718-777 (x innermost), scatter = plain += in edge order (:428-436)
// Derivatives come from forward-mode duals (dual.hpp) instead of symbolic AD; everything else
// (which residual instances touch X(i), bounds, exclude, OOB loads = 0) follows SURVEY.md App. B.
//
// Deliberate deviation (documented in DESIGN.md): the six reduction scalars are accumulated in
// `double` regardless of opt_float, so the oracle is summation-order independent; per-element
// arithmetic is done in opt_float exactly like the reference.
#pragma once
#include <cstdio>
#include <cstdlib>
#include <cstring>
#include <cmath>
#include <vector>
#include <string>
#include "dual.hpp"

namespace oracle {

struct UnknownImage { void* ptr; int ch; int uoff; };

template <class T>
struct SpecBase {
    int dim[3] = {1, 1, 1};     // unknown index space (x fastest; o.t:564-579)
    long N = 0;                 // sites
    long E = 0;                 // hyper-edges (graph problems)
    int nimg = 0;
    UnknownImage img[4];
    const int* gidx[4] = {nullptr, nullptr, nullptr, nullptr};
    // LM parameters (opt_float Params injected by o.t:1575-1580)
    T trust_region_radius = 0, radius_decrease_factor = 0, min_lm_diagonal = 0, max_lm_diagonal = 0;

    bool inside(int x, int y, int z) const {
        return x >= 0 && y >= 0 && z >= 0 && x < dim[0] && y < dim[1] && z < dim[2];
    }
    long lin(int x, int y, int z) const { return x + (long)dim[0] * (y + (long)dim[1] * z); }
    T& X(long site, int u) const {
        for (int i = 0; i < nimg; ++i)
            if (u < img[i].uoff + img[i].ch) return ((T*)img[i].ptr)[site * img[i].ch + (u - img[i].uoff)];
        abort();
    }
    void add_unknown(void* p, int ch) {
        int off = nimg ? img[nimg - 1].uoff + img[nimg - 1].ch : 0;
        img[nimg++] = UnknownImage{p, ch, off};
    }
};

// ---- seeding of unknown accesses ------------------------------------------------------------
template <class T, class S> struct Seeder;
template <class T> struct Seeder<T, T> {
    static T make(T v, int, T) { return v; }
    static void add(T&, int, T, T) {}
};
template <class T, int N> struct Seeder<T, Dual<T, N>> {
    // tangent slot `slot` (or -1) gets 1; the last tangent carries the direction (J*dir)
    static Dual<T, N> make(T v, int slot, T dirv) {
        Dual<T, N> r(v);
        if (slot >= 0) r.d[slot] = T(1);
        r.d[N - 1] = dirv;
        return r;
    }
    // chain rule through a stored gradient image: d/d(unknown in `slot`) += g, J*dir += g * dir(unknown)
    static void add(Dual<T, N>& r, int slot, T g, T dirv) {
        if (slot >= 0) r.d[slot] += g;
        r.d[N - 1] += g * dirv;
    }
};

// Context for a centred residual template evaluated at centre (cx,cy,cz).
template <class T, class S, class Spec>
struct CenterCtx {
    const Spec& sp;
    int cx, cy, cz;
    long seed;        // site whose channels get unit tangents (-1: none)
    const T* dir;     // direction field, AoS [site][U] (nullptr: none)
    CenterCtx(const Spec& s, int x, int y, int z, long sd, const T* d) : sp(s), cx(x), cy(y), cz(z), seed(sd), dir(d) {}

    bool inb(int ox, int oy = 0, int oz = 0) const { return sp.inside(cx + ox, cy + oy, cz + oz); }
    bool inb_expanded(int ox, int oy, int e) const {       // o.t:1794-1803 (2-D use only)
        return sp.inside(cx + ox - e, cy + oy - e, cz) && sp.inside(cx + ox + e, cy + oy + e, cz);
    }
    // unknown image `im`, channel ch at offset; out-of-bounds loads give 0 (o.t:856-862)
    S unk(int im, int ch, int ox = 0, int oy = 0, int oz = 0) const {
        int x = cx + ox, y = cy + oy, z = cz + oz;
        if (!sp.inside(x, y, z)) return S(T(0));
        long q = sp.lin(x, y, z);
        const UnknownImage& I = sp.img[im];
        T v = ((const T*)I.ptr)[q * I.ch + ch];
        int u = I.uoff + ch;
        return Seeder<T, S>::make(v, q == seed ? u : -1, dir ? dir[q * Spec::U + u] : T(0));
    }
    // seed bookkeeping for values rebuilt from gradient images (ComputedArrays)
    int  slot_of(int im, int ch, int ox, int oy) const {
        int x = cx + ox, y = cy + oy;
        if (!sp.inside(x, y, cz)) return -1;
        return sp.lin(x, y, cz) == seed ? sp.img[im].uoff + ch : -1;
    }
    T dir_of(int im, int ch, int ox, int oy) const {
        int x = cx + ox, y = cy + oy;
        if (!dir || !sp.inside(x, y, cz)) return T(0);
        return dir[sp.lin(x, y, cz) * Spec::U + sp.img[im].uoff + ch];
    }
    // ComputedArray access (o.t:1551-1563): the value image at (ox,oy) [0 out of bounds], differentiated through
    // its stored gradient images: image j holds d value / d X_{im,ch}(. + goff[j]) (ImageAccess:gradient)
    S computed(const T* val, int ng, const T* const* grad, const int (*goff)[2], int im, int ch, int ox, int oy) const {
        int x = cx + ox, y = cy + oy;
        if (!sp.inside(x, y, cz)) return S(T(0));
        long q = sp.lin(x, y, cz);
        S r(val[q]);
        for (int j = 0; j < ng; ++j) {
            int ux = ox + goff[j][0], uy = oy + goff[j][1];
            Seeder<T, S>::add(r, slot_of(im, ch, ux, uy), grad[j][q], dir_of(im, ch, ux, uy));
        }
        return r;
    }
    template <class A>
    A arr(const A* p, int nch, int ch, int ox = 0, int oy = 0, int oz = 0) const {
        int x = cx + ox, y = cy + oy, z = cz + oz;
        if (!sp.inside(x, y, z)) return A(0);
        return p[sp.lin(x, y, z) * nch + ch];
    }
};

// Context for a graph residual template evaluated on hyper-edge e.
template <class T, class S, class Spec>
struct GraphCtx {
    const Spec& sp;
    long v[4];
    bool seeded;
    const T* dir;
    GraphCtx(const Spec& s, long e, bool sd, const T* d) : sp(s), seeded(sd), dir(d) {
        for (int k = 0; k < Spec::GV; ++k) v[k] = s.gidx[k][e];
    }
    S unk(int im, int ch, int slot) const {
        const UnknownImage& I = sp.img[im];
        long q = v[slot];
        T val = ((const T*)I.ptr)[q * I.ch + ch];
        int u = I.uoff + ch;
        return Seeder<T, S>::make(val, seeded ? slot * Spec::U + u : -1, dir ? dir[q * Spec::U + u] : T(0));
    }
    template <class A> A arr(const A* p, int nch, int ch, int slot) const { return p[v[slot] * nch + ch]; }
};

// ---- solver parameters (solverGPUGaussNewton.t:41-55 defaults, :186-201 types) ----------------
struct SolverParameters {
    float min_relative_decrease = 1e-3f;
    float min_trust_region_radius = 1e-32f;
    float max_trust_region_radius = 1e16f;
    float q_tolerance = 0.0001f;
    float function_tolerance = 0.000001f;
    float trust_region_radius = 1e4f;
    float radius_decrease_factor = 2.0f;
    float min_lm_diagonal = 1e-6f;
    float max_lm_diagonal = 1e32f;
    int residual_reset_period = 10;
    int nIter = 0;
    int nIterations = 10;
    int lIterations = 10;

    bool set(const char* name, const void* value) {      // solverGPUGaussNewton.t:2382-2398
#define ORACLE_SP(f, ty) if (!strcmp(name, #f)) { f = *(const ty*)value; return true; }
        ORACLE_SP(min_relative_decrease, float) ORACLE_SP(min_trust_region_radius, float)
        ORACLE_SP(max_trust_region_radius, float) ORACLE_SP(q_tolerance, float)
        ORACLE_SP(function_tolerance, float) ORACLE_SP(trust_region_radius, float)
        ORACLE_SP(radius_decrease_factor, float) ORACLE_SP(min_lm_diagonal, float)
        ORACLE_SP(max_lm_diagonal, float) ORACLE_SP(residual_reset_period, int)
        ORACLE_SP(nIter, int) ORACLE_SP(nIterations, int) ORACLE_SP(lIterations, int)
#undef ORACLE_SP
        return false;
    }
};

struct PlanBase {
    SolverParameters sp;
    int verbosity = 0;
    int threads = 1;
    std::vector<double> cost_history;   // cost after init and after every step (test hook)
    long pcg_iterations = 0;            // PCG iterations executed since the plan was made (test hook)
    virtual ~PlanBase() {}
    virtual void init(void** params) = 0;
    virtual int step(void** params) = 0;
    virtual double cost() const = 0;
    // test hooks: apply J^T J to an arbitrary direction / fetch J^T F and the preconditioner at the current X
    virtual void debug_apply_jtj(void** params, const void* p_in, void* ap_out) = 0;
    virtual void debug_eval_jtf(void** params, void* g_out, void* pre_out) = 0;
    virtual long unknown_count() const = 0;
};

template <class Spec>
struct Solver : PlanBase {
    using T = typename Spec::Scalar;
    static constexpr int U = Spec::U;
    static constexpr int GV = Spec::GV;
    using DC = Dual<T, U + 1>;
    using DG = Dual<T, (GV > 0 ? GV : 1) * U + 1>;
    using Acc = double;

    Spec spec;
    bool lm;
    std::vector<T> delta, r, b, Adelta, z, p, Ap, CtC, pre, SSq, prevX;
    T prevCost = 0;

    Solver(const unsigned* dims, bool use_lm) : lm(use_lm) {
        spec.plan(dims);
        size_t n = (size_t)spec.N * U;
        for (auto* v : {&delta, &r, &b, &Adelta, &z, &p, &Ap, &CtC, &pre, &SSq, &prevX}) v->assign(n, T(0));
    }
    long unknown_count() const override { return spec.N * U; }
    double cost() const override { return (double)prevCost; }

    // ---------------------------------------------------------------- per-site functions
    // loop over every residual instance that reads X(site): template k centred at site - o
    template <class F>
    void for_instances(int x, int y, int zc, const T* dir, F&& f) const {
        long site = spec.lin(x, y, zc);
        for (int k = 0; k < Spec::NT; ++k) {
            const int(*offs)[3]; int no = Spec::offsets(k, &offs);
            for (int j = 0; j < no; ++j) {
                int cx = x - offs[j][0], cy = y - offs[j][1], cz = zc - offs[j][2];
                if (!spec.inside(cx, cy, cz)) continue;     // instance exists iff its centre is in the image
                CenterCtx<T, DC, Spec> c(spec, cx, cy, cz, site, dir);
                DC res[Spec::RMAX];
                int n = spec.template residual<DC>(k, c, res);
                for (int q = 0; q < n; ++q) f(res[q]);
            }
        }
    }
    T site_cost(int x, int y, int zc) const {                 // o.t:3119-3129
        T s = 0;
        for (int k = 0; k < Spec::NT; ++k) {
            CenterCtx<T, T, Spec> c(spec, x, y, zc, -1, nullptr);
            T res[Spec::RMAX];
            int n = spec.template residual<T>(k, c, res);
            for (int q = 0; q < n; ++q) s += res[q] * res[q];
        }
        return (T)(0.5 * (double)s);                           // test_fmapcost.t:32
    }
    void site_jtf(int x, int y, int zc, T* g, T* d) const {    // o.t:2870-2913
        for (int u = 0; u < U; ++u) g[u] = d[u] = 0;
        for_instances(x, y, zc, nullptr, [&](const DC& rr) {
            for (int u = 0; u < U; ++u) { g[u] += rr.d[u] * rr.v; d[u] += rr.d[u] * rr.d[u]; }
        });
        if (!Spec::USE_PRE) for (int u = 0; u < U; ++u) d[u] = T(1);
    }
    void site_jtj(int x, int y, int zc, const T* dir, T* out) const {   // o.t:2770-2830
        for (int u = 0; u < U; ++u) out[u] = 0;
        for_instances(x, y, zc, dir, [&](const DC& rr) {
            for (int u = 0; u < U; ++u) out[u] += rr.d[u] * rr.d[U];
        });
        if (lm) {
            long site = spec.lin(x, y, zc);
            for (int u = 0; u < U; ++u) out[u] += CtC[site * U + u] * dir[site * U + u];
        }
    }
    void site_ctc(int x, int y, int zc, T* out) const {       // o.t:2996-3029
        T inv_radius = T(1) / spec.trust_region_radius;
        for (int u = 0; u < U; ++u) out[u] = 0;
        for_instances(x, y, zc, nullptr, [&](const DC& rr) {
            for (int u = 0; u < U; ++u) out[u] += rr.d[u] * rr.d[u] * inv_radius;
        });
    }
    T site_modelcost(int x, int y, int zc) const {            // o.t:2915-2943
        T s = 0;
        for (int k = 0; k < Spec::NT; ++k) {
            CenterCtx<T, DC, Spec> c(spec, x, y, zc, -1, delta.data());
            DC res[Spec::RMAX];
            int n = spec.template residual<DC>(k, c, res);
            for (int q = 0; q < n; ++q) { T m = res[q].v + res[q].d[U]; s += m * m; }
        }
        return T(0.5) * s;
    }
    static T ginv(T d) { T t = T(1) + std::sqrt(d); return T(1) / (t * t); }   // solver:478-487 (CERES)

    // ---------------------------------------------------------------- site loops
    template <class F> Acc sites(F&& f) const {                // backend_cpu.t:718-777
        Acc total = 0;
        const int W = spec.dim[0], H = spec.dim[1], D = spec.dim[2];
        const long rows = (long)H * D;
#pragma omp parallel for reduction(+ : total) num_threads(threads) schedule(static)
        for (long row = 0; row < rows; ++row) {
            int y = (int)(row % H), zc = (int)(row / H);
            for (int x = 0; x < W; ++x) total += (Acc)f(x, y, zc, spec.lin(x, y, zc));
        }
        return total;
    }
    template <class F> Acc edges(F&& f) const {                 // serial: plain += in edge order
        Acc total = 0;
        for (long e = 0; e < spec.E; ++e) total += (Acc)f(e);
        return total;
    }

    void precompute() { spec.precompute_all(threads); }

    T compute_cost() {                                          // solver:1389-1441
        Acc c = sites([&](int x, int y, int zc, long) -> T {
            if (spec.exclude(x, y, zc)) return T(0);
            return site_cost(x, y, zc);
        });
        if (GV > 0) c += edges([&](long e) -> T {
            T s = 0;
            for (int k = 0; k < Spec::NGT; ++k) {
                GraphCtx<T, T, Spec> g(spec, e, false, nullptr);
                T res[Spec::RMAX];
                int n = spec.template gresidual<T>(k, g, res);
                for (int q = 0; q < n; ++q) s += res[q] * res[q];
            }
            return (T)(0.5 * (double)s);
        });
        return (T)c;
    }
    T compute_model_cost() {                                    // solver:1445-1476
        Acc c = sites([&](int x, int y, int zc, long) -> T {
            if (spec.exclude(x, y, zc)) return T(0);
            return site_modelcost(x, y, zc);
        });
        if (GV > 0) c += edges([&](long e) -> T {
            T s = 0;
            for (int k = 0; k < Spec::NGT; ++k) {
                GraphCtx<T, DG, Spec> g(spec, e, false, delta.data());
                DG res[Spec::RMAX];
                int n = spec.template gresidual<DG>(k, g, res);
                for (int q = 0; q < n; ++q) { T m = res[q].v + res[q].d[GV * U]; s += m * m; }
            }
            return T(0.5) * s;
        });
        return (T)c;
    }
    // graph scatter of J^T J dir into out; returns sum of dir . (J^T J dir) contributions (o.t:2833-2867)
    Acc graph_jtj(const T* dir, T* out) {
        return edges([&](long e) -> T {
            T d = 0;
            for (int k = 0; k < Spec::NGT; ++k) {
                GraphCtx<T, DG, Spec> g(spec, e, true, dir);
                DG res[Spec::RMAX];
                int n = spec.template gresidual<DG>(k, g, res);
                for (int q = 0; q < n; ++q) {
                    T Jp = res[q].d[GV * U];
                    for (int s = 0; s < GV; ++s)
                        for (int u = 0; u < U; ++u) {
                            T partial = res[q].d[s * U + u];
                            if (partial == T(0)) continue;
                            T jtjp = Jp * partial;
                            out[g.v[s] * U + u] += jtjp;
                            d += dir[g.v[s] * U + u] * jtjp;
                        }
                }
            }
            return d;
        });
    }
    void graph_jtf(T* R, T* Pre) {                               // o.t:2969-2994
        edges([&](long e) -> T {
            for (int k = 0; k < Spec::NGT; ++k) {
                GraphCtx<T, DG, Spec> g(spec, e, true, nullptr);
                DG res[Spec::RMAX];
                int n = spec.template gresidual<DG>(k, g, res);
                for (int q = 0; q < n; ++q)
                    for (int s = 0; s < GV; ++s)
                        for (int u = 0; u < U; ++u) {
                            T partial = res[q].d[s * U + u];
                            R[g.v[s] * U + u] += T(-1) * partial * res[q].v;
                            Pre[g.v[s] * U + u] += partial * partial;
                        }
            }
            return T(0);
        });
    }
    void graph_ctc(T* out) {                                     // o.t:3031-3057
        T inv_radius = T(1) / spec.trust_region_radius;
        edges([&](long e) -> T {
            for (int k = 0; k < Spec::NGT; ++k) {
                GraphCtx<T, DG, Spec> g(spec, e, true, nullptr);
                DG res[Spec::RMAX];
                int n = spec.template gresidual<DG>(k, g, res);
                for (int q = 0; q < n; ++q)
                    for (int s = 0; s < GV; ++s)
                        for (int u = 0; u < U; ++u) {
                            T partial = res[q].d[s * U + u];
                            out[g.v[s] * U + u] += partial * partial * inv_radius;
                        }
            }
            return T(0);
        });
    }

    // PCGInit1 (+_Graph, _Finish): returns alphaNumerator        solver:521-598, 1188-1197
    Acc pcg_init1() {
        constexpr bool isGraph = GV > 0;
        Acc aNum = sites([&](int x, int y, int zc, long s) -> T {
            T g[U], d[U]; T acc = 0;
            for (int u = 0; u < U; ++u) g[u] = d[u] = 0;
            bool ex = spec.exclude(x, y, zc);
            if (!ex) {
                site_jtf(x, y, zc, g, d);
                for (int u = 0; u < U; ++u) { delta[s * U + u] = 0; g[u] = -g[u]; r[s * U + u] = g[u]; }
            }
            if (!ex && !isGraph)
                for (int u = 0; u < U; ++u) {
                    d[u] = ginv(d[u]);
                    T pp = d[u] * g[u];
                    p[s * U + u] = pp;
                    acc += g[u] * pp;
                }
            for (int u = 0; u < U; ++u) pre[s * U + u] = d[u];
            return acc;
        });
        if (isGraph) {
            graph_jtf(r.data(), pre.data());
            aNum = sites([&](int, int, int, long s) -> T {       // PCGInit1_Finish: no exclude test
                T acc = 0;
                for (int u = 0; u < U; ++u) {
                    T m = ginv(pre[s * U + u]);
                    if (!Spec::USE_PRE) m = T(1);
                    T pp = m * r[s * U + u];
                    pre[s * U + u] = m; p[s * U + u] = pp;
                    acc += r[s * U + u] * pp;
                }
                return acc;
            });
        }
        return aNum;
    }
    // applies (J^T J [+ CtC]) to `dir` over non-excluded sites; returns dir . out
    Acc apply_A(const T* dir, T* out) {
        Acc den = sites([&](int x, int y, int zc, long s) -> T {
            if (spec.exclude(x, y, zc)) return T(0);
            T o[U]; site_jtj(x, y, zc, dir, o);
            T d = 0;
            for (int u = 0; u < U; ++u) { out[s * U + u] = o[u]; d += dir[s * U + u] * o[u]; }
            return d;
        });
        if (GV > 0) den += graph_jtj(dir, out);
        return den;
    }

    void bind(void** params, bool isInit) {
        spec.bind(params);
        (void)isInit;
    }

    void init(void** params) override {                          // solver:1766-1897
        bind(params, true);
        sp.nIter = 0;
        if (lm) {
            spec.trust_region_radius = (T)sp.trust_region_radius;
            spec.radius_decrease_factor = (T)sp.radius_decrease_factor;
            spec.min_lm_diagonal = (T)sp.min_lm_diagonal;
            spec.max_lm_diagonal = (T)sp.max_lm_diagonal;
        }
        precompute();
        prevCost = compute_cost();
        cost_history.clear();
        cost_history.push_back((double)prevCost);
    }
    void cleanup() {
        if (verbosity > 0) printf("final cost=%.16f\n", (double)prevCost);
    }

    int step(void** params) override {                           // solver:1913-2349
        bind(params, false);
        if (sp.nIter >= sp.nIterations) { cleanup(); return 0; }
        const T q_tolerance = (T)sp.q_tolerance;
        Acc aNum = pcg_init1(), aDen = 0, bNum = 0, qv = 0;
        T Q0 = 0, Q1 = 0;
        if (lm) {
            aNum = 0; qv = 0;
            if (sp.nIter == 0)                                   // PCGSaveSSq
                sites([&](int x, int y, int zc, long s) -> T {
                    if (!spec.exclude(x, y, zc)) for (int u = 0; u < U; ++u) SSq[s * U + u] = pre[s * U + u];
                    return T(0);
                });
            sites([&](int x, int y, int zc, long s) -> T {       // PCGComputeCtC
                if (spec.exclude(x, y, zc)) return T(0);
                T o[U]; site_ctc(x, y, zc, o);
                for (int u = 0; u < U; ++u) CtC[s * U + u] = o[u];
                return T(0);
            });
            if (GV > 0) graph_ctc(CtC.data());
            Acc qacc = 0;
            aNum = 0;
            // PCGFinalizeDiagonal                                 solver:1061-1099
            const long n = spec.N;
            for (long s = 0; s < n; ++s) {
                int x = (int)(s % spec.dim[0]), y = (int)((s / spec.dim[0]) % spec.dim[1]), zc = (int)(s / ((long)spec.dim[0] * spec.dim[1]));
                if (spec.exclude(x, y, zc)) continue;
                for (int u = 0; u < U; ++u) {
                    long i = s * U + u;
                    T unclamped = CtC[i];
                    T invS = T(1) / SSq[i];
                    T mult = invS / spec.trust_region_radius;
                    T lo = spec.min_lm_diagonal * mult, hi = spec.max_lm_diagonal * mult;
                    T c = std::fmin(std::fmax(unclamped, lo), hi);
                    CtC[i] = c;
                    T m = T(1) / (c + spec.trust_region_radius * unclamped);
                    pre[i] = m;
                    T res = r[i];
                    b[i] = res;
                    T pp = m * res;
                    p[i] = pp;
                    aNum += (Acc)(res * pp);
                    qacc += (Acc)(T(0.5) * (delta[i] * (res + res)));
                }
            }
            Q0 = (T)qacc;
        }
        for (int lIter = 0; lIter < sp.lIterations; ++lIter) {
            ++pcg_iterations;
            aDen = apply_A(p.data(), Ap.data());                 // PCGStep1 (+_Graph)
            // graph scatter accumulates into Ap; the centred pass *overwrites* its own entries first,
            // and the scatter target must start from that (solver:1216-1231) -- see apply_A ordering.
            bNum = 0; qv = 0;
            const T alpha = (T)aNum / (T)aDen;
            if (lm && ((lIter + 1) % sp.residual_reset_period) == 0) {
                // PCGStep2_1stHalf / computeAdelta / PCGStep2_2ndHalf     solver:738-799
                sites([&](int x, int y, int zc, long s) -> T {
                    if (spec.exclude(x, y, zc)) return T(0);
                    for (int u = 0; u < U; ++u) delta[s * U + u] = delta[s * U + u] + alpha * p[s * U + u];
                    return T(0);
                });
                if (GV > 0) {
                    // Adelta is scattered into by the graph pass: the reference never clears it
                    // between resets (solver:1237-1244 adds into pd.Adelta after the centred pass
                    // overwrote non-excluded entries), which apply_A reproduces.
                }
                apply_A(delta.data(), Adelta.data());
                Acc bq[2] = {0, 0};
                bNum = sites([&](int x, int y, int zc, long s) -> T {
                    if (spec.exclude(x, y, zc)) return T(0);
                    T acc = 0;
                    for (int u = 0; u < U; ++u) {
                        long i = s * U + u;
                        T rr = b[i] - Adelta[i];
                        r[i] = rr;
                        T m = Spec::USE_PRE ? pre[i] : T(1);
                        T zz = m * rr; z[i] = zz;
                        acc += zz * rr;
                    }
                    return acc;
                });
                qv = sites([&](int x, int y, int zc, long s) -> T {
                    if (spec.exclude(x, y, zc)) return T(0);
                    T d = 0;
                    for (int u = 0; u < U; ++u) { long i = s * U + u; d += delta[i] * (r[i] + b[i]); }
                    return T(0.5) * d;
                });
                (void)bq;
            } else {
                // PCGStep2                                        solver:665-731
                Acc qq = 0;
                bNum = 0;
                const int W = spec.dim[0], H = spec.dim[1], D = spec.dim[2];
                const long rows = (long)H * D;
#pragma omp parallel for reduction(+ : bNum, qq) num_threads(threads) schedule(static)
                for (long row = 0; row < rows; ++row) {
                    int y = (int)(row % H), zc = (int)(row / H);
                    for (int x = 0; x < W; ++x) {
                        if (spec.exclude(x, y, zc)) continue;
                        long s = spec.lin(x, y, zc);
                        T bacc = 0, qacc = 0;
                        for (int u = 0; u < U; ++u) {
                            long i = s * U + u;
                            T dl = delta[i] + alpha * p[i];
                            delta[i] = dl;
                            T rr = r[i] - alpha * Ap[i];
                            r[i] = rr;
                            T m = Spec::USE_PRE ? pre[i] : T(1);
                            T zz = m * rr; z[i] = zz;
                            bacc += zz * rr;
                            if (lm) qacc += dl * (rr + b[i]);
                        }
                        bNum += (Acc)bacc;
                        qq += (Acc)(T(0.5) * qacc);
                    }
                }
                qv = qq;
            }
            // PCGStep3                                            solver:814-845
            const T beta = (T)bNum / (T)aNum;
            sites([&](int x, int y, int zc, long s) -> T {
                if (spec.exclude(x, y, zc)) return T(0);
                for (int u = 0; u < U; ++u) { long i = s * U + u; p[i] = z[i] + beta * p[i]; }
                return T(0);
            });
            aNum = bNum;
            if (lm) {
                Q1 = (T)qv;
                T zeta = (T)(lIter + 1) * (Q1 - Q0) / Q1;
                if (zeta < q_tolerance) {
                    if (verbosity > 0) printf("zeta=%.18g < q_tol(%f), breaking at iteration: %d\n", (double)zeta, (double)q_tolerance, lIter + 1);
                    break;
                }
                Q0 = Q1;
            }
        }
        T model_cost_change = 0;
        if (lm) {
            T model_cost = compute_model_cost();
            model_cost_change = prevCost - model_cost;
            sites([&](int x, int y, int zc, long s) -> T {       // savePreviousUnknowns
                if (!spec.exclude(x, y, zc)) for (int u = 0; u < U; ++u) prevX[s * U + u] = spec.X(s, u);
                return T(0);
            });
        }
        sites([&](int x, int y, int zc, long s) -> T {           // PCGLinearUpdate
            if (!spec.exclude(x, y, zc)) for (int u = 0; u < U; ++u) spec.X(s, u) = spec.X(s, u) + delta[s * U + u];
            return T(0);
        });
        precompute();
        T newCost = compute_cost();
        if (lm) {
            T cost_change = prevCost - newCost;
            T relative_decrease = cost_change / model_cost_change;
            if (cost_change >= 0 && relative_decrease > (T)sp.min_relative_decrease) {
                T absolute_function_tolerance = prevCost * (T)sp.function_tolerance;
                if (cost_change <= absolute_function_tolerance) {
                    if (verbosity > 0) printf("\nFunction tolerance reached, exiting\n");
                    // the reference leaves prevCost untouched here (solver:2257-2262)
                    cost_history.push_back((double)prevCost);
                    cleanup();
                    return 0;
                }
                T step_quality = relative_decrease;
                double min_factor = 1.0 / 3.0;
                double tmp_factor = 1.0 - std::pow(2.0 * (double)step_quality - 1.0, 3.0);
                spec.trust_region_radius = (T)((double)spec.trust_region_radius / std::fmax(min_factor, tmp_factor));
                spec.trust_region_radius = (T)std::fmin((double)spec.trust_region_radius, (double)(T)sp.max_trust_region_radius);
                spec.radius_decrease_factor = T(2.0);
                prevCost = newCost;
            } else {
                sites([&](int x, int y, int zc, long s) -> T {   // revertUpdate
                    if (!spec.exclude(x, y, zc)) for (int u = 0; u < U; ++u) spec.X(s, u) = prevX[s * U + u];
                    return T(0);
                });
                spec.trust_region_radius = spec.trust_region_radius / spec.radius_decrease_factor;
                if (verbosity > 0) printf(" trust_region_radius=%f \n", (double)spec.trust_region_radius);
                spec.radius_decrease_factor = T(2.0) * spec.radius_decrease_factor;
                if (spec.trust_region_radius <= (T)sp.min_trust_region_radius) {
                    if (verbosity > 0) printf("\nTrust_region_radius is less than the min, exiting\n");
                    cost_history.push_back((double)prevCost);
                    cleanup();
                    return 0;
                }
                if (verbosity > 0) printf("REVERT\n");
                precompute();
            }
        } else {
            prevCost = newCost;
        }
        cost_history.push_back((double)prevCost);
        sp.nIter++;
        return 1;
    }

    void debug_apply_jtj(void** params, const void* p_in, void* ap_out) override {
        bind(params, false);
        precompute();
        std::vector<T> out((size_t)spec.N * U, T(0));
        bool saved = lm; lm = false;
        apply_A((const T*)p_in, out.data());
        lm = saved;
        memcpy(ap_out, out.data(), out.size() * sizeof(T));
    }
    void debug_eval_jtf(void** params, void* g_out, void* pre_out) override {
        bind(params, false);
        precompute();
        pcg_init1();
        memcpy(g_out, r.data(), r.size() * sizeof(T));        // r = -J^T F
        memcpy(pre_out, pre.data(), pre.size() * sizeof(T));
    }
};

}  // namespace oracle
