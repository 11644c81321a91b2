// registry.cu -- energy-name -> engine factory table and the extern "C" forwarding layer of optk.h.
// Opt_ProblemDefine keys on the basename of the energy file (SURVEY.md 3.1 "New design"): each of
// the shipped example energies is bound to a hand-written functor set; anything else is rejected.
#include "engine.cuh"

extern "C" {
#define OPTK_DECL(name) OptK_Plan* optk_make_##name(int, int, const unsigned*, int);
OPTK_DECL(poisson_image_editing)
OPTK_DECL(image_warping)
OPTK_DECL(volumetric_mesh_deformation)
OPTK_DECL(optical_flow)
OPTK_DECL(arap_mesh_deformation)
OPTK_DECL(shape_from_shading)
OPTK_DECL(embedded_mesh_deformation)
OPTK_DECL(cotangent_mesh_smoothing)
OPTK_DECL(robust_nonrigid_alignment)
OPTK_DECL(intrinsic_image_decomposition)
#undef OPTK_DECL
}

namespace {
struct Entry { const char* name; OptK_Plan* (*make)(int, int, const unsigned*, int); };
const Entry kTable[] = {
    {"poisson_image_editing", optk_make_poisson_image_editing},
    {"image_warping", optk_make_image_warping},
    {"volumetric_mesh_deformation", optk_make_volumetric_mesh_deformation},
    {"optical_flow", optk_make_optical_flow},
    {"arap_mesh_deformation", optk_make_arap_mesh_deformation},
    {"shape_from_shading", optk_make_shape_from_shading},
    {"embedded_mesh_deformation", optk_make_embedded_mesh_deformation},
    {"cotangent_mesh_smoothing", optk_make_cotangent_mesh_smoothing},
    {"robust_nonrigid_alignment", optk_make_robust_nonrigid_alignment},
    {"intrinsic_image_decomposition", optk_make_intrinsic_image_decomposition},
};
const Entry* find(const char* energy) {
    for (const Entry& e : kTable) if (!strcmp(e.name, energy)) return &e;
    return nullptr;
}
}  // namespace

// ---- fatal errors (common.cuh: optk_fatal) -------------------------------------------------------------------------
#include <cstdarg>
#include <mutex>
namespace {
std::mutex g_err_mutex;
void (*g_handler)(int, const char*, void*) = nullptr;
void* g_handler_user = nullptr;
thread_local char g_message[1024] = "";
thread_local int g_code = 0;

// The catch side of optk_fatal: hand the error to the handler (once per throw) and mark the plan failed.
void report(OptK_Plan* p, const optk::Fatal& f) {
    if (p) p->failed = true;
    cudaGetLastError();                           // a non-sticky error (a refused cudaMalloc) must not surface again in the next launch check
    void (*h)(int, const char*, void*); void* u;
    { std::lock_guard<std::mutex> lock(g_err_mutex); h = g_handler; u = g_handler_user; }
    if (h) h(f.code, g_message, u);
}
// Every entry that can reach a CUDA / NCCL call goes through this: a failed plan does nothing, a throw ends the call.
template <class R, class Fn> R guard(OptK_Plan* p, R fail, Fn&& fn) {
    if (p && p->failed) return fail;
    try { return fn(); } catch (const optk::Fatal& f) { report(p, f); return fail; }
}
template <class Fn> void guard(OptK_Plan* p, Fn&& fn) {
    if (p && p->failed) return;
    try { fn(); } catch (const optk::Fatal& f) { report(p, f); }
}
}  // namespace

extern "C" {

void optk_fatal(int code, const char* fmt, ...) {
    va_list ap;
    va_start(ap, fmt);
    vsnprintf(g_message, sizeof g_message, fmt, ap);
    va_end(ap);
    g_code = code;
    bool have;
    { std::lock_guard<std::mutex> lock(g_err_mutex); have = g_handler != nullptr; }
    if (!have) {                                  // the reference's convention (backend_cuda.t:26-40): print and exit
        fputs(g_message, stdout);
        fflush(stdout);
        exit(code);
    }
    throw optk::Fatal{code};
}
void optk_set_error_handler(void (*handler)(int, const char*, void*), void* user) {
    std::lock_guard<std::mutex> lock(g_err_mutex);
    g_handler = handler; g_handler_user = user;
}
const char* optk_last_error(int* code) { if (code) *code = g_code; return g_message; }
int optk_failed(OptK_Plan* p) { return p->failed ? 1 : 0; }
// for the entry points outside this file (csr_api.cu): run the catch side without a plan
void optk_report_fatal(int code) { report(nullptr, optk::Fatal{code}); }

int optk_has_energy(const char* energy) { return find(energy) != nullptr; }

OptK_Plan* optk_create(const char* energy, int is_double, int is_lm, const unsigned* dims, int collect_timing) {
    const Entry* e = find(energy);
    if (!e) { fprintf(stderr, "Opt: energy '%s' has no built-in functor set\n", energy); return nullptr; }
    int ndev = 0;
    if (cudaGetDeviceCount(&ndev) != cudaSuccess || ndev == 0) {
        fprintf(stderr, "Opt: no CUDA device available; this library has no CPU fallback\n");
        return nullptr;
    }
    return guard<OptK_Plan*>(nullptr, nullptr, [&] { return e->make(is_double, is_lm, dims, collect_timing); });
}
void optk_destroy(OptK_Plan* p) { delete p; }
void optk_bind(OptK_Plan* p, void** pp) { guard(p, [&] { p->bind(pp); }); }
void optk_set_lm(OptK_Plan* p, double r, double mn, double mx) { guard(p, [&] { p->set_lm(r, mn, mx); }); }
void optk_range_push(const char* name) { nvtxRangePushA(name); }
void optk_range_pop(void) { nvtxRangePop(); }
void optk_overall_begin(OptK_Plan* p) { guard(p, [&] { p->overall_begin(); }); }
void optk_overall_end(OptK_Plan* p) { guard(p, [&] { p->overall_end(); }); }
void optk_begin_call(OptK_Plan* p) { guard(p, [&] { p->begin_call(); }); }
void optk_end_call(OptK_Plan* p) { guard(p, [&] { p->end_call(); }); }
void optk_precompute(OptK_Plan* p) { guard(p, [&] { p->precompute(); }); }
double optk_cost(OptK_Plan* p) { return guard<double>(p, NAN, [&] { return p->cost(); }); }
void optk_pcg_begin(OptK_Plan* p, int nIter) { guard(p, [&] { p->pcg_begin(nIter); }); }
void optk_pcg_iter(OptK_Plan* p, int l, int reset) { guard(p, [&] { p->pcg_iter(l, reset); }); }
void optk_pcg_run(OptK_Plan* p, int n) { guard(p, [&] { p->pcg_run(n); }); }
int optk_pcg_run_lm(OptK_Plan* p, int n, int reset, double qtol, double* zeta, int* fired) {
    *zeta = 0.0; *fired = 0;
    return guard<int>(p, 0, [&] { return p->pcg_run_lm(n, reset, qtol, zeta, fired); });
}
double optk_fetch_q(OptK_Plan* p) { return guard<double>(p, NAN, [&] { return p->fetch_q(); }); }
double optk_model_cost(OptK_Plan* p) { return guard<double>(p, NAN, [&] { return p->model_cost(); }); }
void optk_save_unknowns(OptK_Plan* p) { guard(p, [&] { p->save_unknowns(); }); }
void optk_linear_update(OptK_Plan* p) { guard(p, [&] { p->linear_update(); }); }
void optk_revert_update(OptK_Plan* p) { guard(p, [&] { p->revert_update(); }); }
int optk_set_slab(OptK_Plan* p, int rank, int nranks, int halo, const void* uid) {
    return guard<int>(p, -1, [&] { return p->set_slab(rank, nranks, halo, uid); });
}
void optk_nccl_unique_id(void* out128) {
    memset(out128, 0, 128);
    guard(nullptr, [&] {
        ncclUniqueId id;
        OPTK_NCCL(ncclGetUniqueId(&id));
        static_assert(sizeof(id) == 128, "ncclUniqueId is 128 bytes");
        memcpy(out128, &id, sizeof(id));
    });
}
int optk_set_materialized(OptK_Plan* p, int mode) { return guard<int>(p, 0, [&] { return p->set_materialized(mode); }); }
long optk_materialized_nnz(OptK_Plan* p, int which) { return p->materialized_nnz(which); }
long optk_launch_count(OptK_Plan* p) { return p->launches; }
void optk_print_timing(OptK_Plan* p) { guard(p, [&] { p->print_timing(); }); }
int optk_kernel_time(OptK_Plan* p, const char* k, double* ms) { *ms = 0.0; return guard<int>(p, 0, [&] { return p->kernel_time(k, ms); }); }
void optk_reset_timing(OptK_Plan* p) { guard(p, [&] { p->reset_timing(); }); }
int optk_unknown_channels(OptK_Plan* p) { return p->unknown_channels(); }
long optk_site_count(OptK_Plan* p) { return p->site_count(); }
int optk_debug_read(OptK_Plan* p, const char* name, void* out) { return guard<int>(p, -1, [&] { return p->debug_read(name, out); }); }
int optk_debug_read_range(OptK_Plan* p, const char* name, long s0, long n, void* out) {
    return guard<int>(p, -1, [&] { return p->debug_read_range(name, s0, n, out); });
}
void optk_debug_apply_jtj(OptK_Plan* p, const void* hp, void* out) { guard(p, [&] { p->debug_apply_jtj(hp, out); }); }

}  // extern "C"
