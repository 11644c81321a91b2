/* opt_api.c -- the Opt.h entry points (plain C host) over the thin CUDA C-ABI of optk.h.
 *
 * Replaces the reference's C wrapper and Lua front end for the solver path:
 *   Opt_NewState          API/src/createwrapper.t:150-287
 *   Opt_ProblemDefine     API/src/o.t:3305-3309, 169-180   (energy file -> built-in functor set)
 *   Opt_ProblemPlan       API/src/o.t:1351-1528, solverGPUGaussNewton.t:2409-2502 (makePlan)
 *   Opt_ProblemInit/Step  solverGPUGaussNewton.t:1766-1897 / 1913-2349
 *   Opt_SetSolverParameter solverGPUGaussNewton.t:2382-2398
 * The nonlinear-iteration control flow (Gauss-Newton, Levenberg-Marquardt trust region) lives
 * here; every per-site / per-edge computation is a CUDA kernel behind optk_*.
 */
#include <math.h>
#include <stdio.h>
#include <stdlib.h>
#include <string.h>

#include "../../include/Opt.h"
#include "../../include/OptB200.h"
#include "optk.h"

int optb200_energy_verify(const char* path, const char* energy, char* msg, size_t cap);   /* energy_check.c */

struct Opt_State {
    Opt_InitializationParameters ip;
};

struct Opt_Problem {
    char energy[128]; /* basename of the energy file without ".t" */
    int is_lm;
};

/* solverGPUGaussNewton.t:186-201 (types), :41-55 (defaults) */
typedef struct {
    float min_relative_decrease, min_trust_region_radius, max_trust_region_radius, q_tolerance,
        function_tolerance, trust_region_radius, radius_decrease_factor, min_lm_diagonal, max_lm_diagonal;
    int residual_reset_period, nIter, nIterations, lIterations;
} SolverParameters;

struct Opt_Plan {
    OptK_Plan* k;
    SolverParameters sp;
    int is_lm, is_double, verbosity, timing;
    double prevCost;
    /* LM state the reference keeps as injected problem parameters (o.t:1575-1580) */
    double trust_region_radius, radius_decrease_factor, min_lm_diagonal, max_lm_diagonal;
    int finished_reported;
    long pcg_iterations;   /* PCG iterations executed since creation (LM leaves the inner loop early) */
    int lm_host_loop;      /* OPTB200_LM_HOSTLOOP=1: blocking fetchQ per iteration, as the reference does */
    int materialized;      /* 0 matrix-free, 1 J and J^T as CSR, 2 fused J^T J (Opt_InitializationParameters) */
};

/* opt_float rounding: the reference keeps these host scalars in opt_float */
static double of(const Opt_Plan* p, double v) { return p->is_double ? v : (double)(float)v; }

#define logSolver(plan, ...) do { if ((plan)->verbosity > 0) { printf(__VA_ARGS__); } } while (0)   /* o.t:95-104 */

Opt_State* Opt_NewState(Opt_InitializationParameters params) {
    char backend[21];
    memcpy(backend, params.backend, 20);
    backend[20] = 0;
    if (backend[0] != 0 && strcmp(backend, "backend_cuda") != 0) {
        fprintf(stderr, "Opt: backend '%s' is not available in this library: it is CUDA-only and has no CPU "
                        "fallback (use \"backend_cuda\")\n", backend);
        return NULL;
    }
    Opt_State* s = (Opt_State*)calloc(1, sizeof(Opt_State));
    s->ip = params;
    return s;
}

void OptB200_FreeState(Opt_State* state) { free(state); }

static void energy_basename(const char* filename, char* out, size_t cap) {
    const char* b = filename;
    for (const char* c = filename; *c; ++c)
        if (*c == '/' || *c == '\\') b = c + 1;
    strncpy(out, b, cap - 1);
    out[cap - 1] = 0;
    size_t n = strlen(out);
    if (n >= 2 && strcmp(out + n - 2, ".t") == 0) out[n - 2] = 0;
}

Opt_Problem* Opt_ProblemDefine(Opt_State* state, const char* filename, const char* solverkind) {
    if (!state || !filename || !solverkind) return NULL;
    Opt_Problem* p = (Opt_Problem*)calloc(1, sizeof(Opt_Problem));
    energy_basename(filename, p->energy, sizeof(p->energy));
    if (!optk_has_energy(p->energy)) {
        fprintf(stderr, "Opt: energy '%s' has no built-in functor set; only the shipped example energies are "
                        "supported (there is no energy compiler in this library)\n", filename);
        free(p);
        return NULL;
    }
    /* The reference compiles the file; here it is checked against the description the functor was written for
       (energy_check.c).  A path that cannot be read (drivers run from other directories, tests pass bare names) is
       bound by its name alone. */
    {
        char why[512];
        const int v = optb200_energy_verify(filename, p->energy, why, sizeof(why));
        if (v == -2 && getenv("OPTB200_ALLOW_MODIFIED_ENERGY")) {
            fprintf(stderr, "Opt: warning: energy file '%s': %s (accepted: OPTB200_ALLOW_MODIFIED_ENERGY is set)\n", filename, why);
        } else if (v == 1) {
            /* not readable: nothing could be checked.  Say so instead of binding silently (OPTB200_REQUIRE_ENERGY_FILE=1 refuses) */
            if (getenv("OPTB200_REQUIRE_ENERGY_FILE")) {
                fprintf(stderr, "Opt: energy file '%s' cannot be read and OPTB200_REQUIRE_ENERGY_FILE is set\n", filename);
                free(p);
                return NULL;
            }
            if (state->ip.verbosityLevel > 0)
                fprintf(stderr, "Opt: note: energy file '%s' cannot be read; bound to the built-in functor set '%s' by name, contents unverified\n",
                        filename, p->energy);
        } else if (v < 0) {
            fprintf(stderr, "Opt: energy file '%s' is not the one the built-in functor set for '%s' implements: %s\n",
                    filename, p->energy, why);
            free(p);
            return NULL;
        }
    }
    p->is_lm = strstr(solverkind, "LM") != NULL; /* o.t:474 */
    return p;
}

void Opt_ProblemDelete(Opt_State* state, Opt_Problem* problem) {
    (void)state;
    free(problem);
}

static void default_parameters(SolverParameters* sp) {
    sp->residual_reset_period = 10;
    sp->min_relative_decrease = 1e-3f;
    sp->min_trust_region_radius = 1e-32f;
    sp->max_trust_region_radius = 1e16f;
    sp->q_tolerance = 0.0001f;
    sp->function_tolerance = 0.000001f;
    sp->trust_region_radius = 1e4f;
    sp->radius_decrease_factor = 2.0f;
    sp->min_lm_diagonal = 1e-6f;
    sp->max_lm_diagonal = 1e32f;
    sp->nIter = 0;
    sp->nIterations = 10;
    sp->lIterations = 10;
}

Opt_Plan* Opt_ProblemPlan(Opt_State* state, Opt_Problem* problem, unsigned int* dimensions) {
    if (!state || !problem || !dimensions) return NULL;
    OptK_Plan* k = optk_create(problem->energy, state->ip.doublePrecision != 0, problem->is_lm, dimensions,
                               state->ip.collectPerKernelTimingInfo != 0);
    if (!k) return NULL;
    Opt_Plan* plan = (Opt_Plan*)calloc(1, sizeof(Opt_Plan));
    plan->k = k;
    plan->is_lm = problem->is_lm;
    plan->is_double = state->ip.doublePrecision != 0;
    plan->verbosity = state->ip.verbosityLevel;
    plan->timing = state->ip.collectPerKernelTimingInfo != 0;
    plan->lm_host_loop = getenv("OPTB200_LM_HOSTLOOP") != NULL;
    /* solver:1523-1761: useMaterializedJTJ stores J (and J^T); with useFusedJTJ the product J^T J as well */
    if (state->ip.useMaterializedJTJ) plan->materialized = optk_set_materialized(k, state->ip.useFusedJTJ ? 2 : 1);
    default_parameters(&plan->sp);
    return plan;
}

void Opt_PlanFree(Opt_State* state, Opt_Plan* plan) {
    (void)state;
    if (!plan) return;
    optk_destroy(plan->k);
    free(plan);
}

void Opt_SetSolverParameter(Opt_State* state, Opt_Plan* plan, const char* name, void* value) {
    (void)state;
    SolverParameters* sp = &plan->sp;
#define SP_FLOAT(f) if (strcmp(name, #f) == 0) { sp->f = *(float*)value; return; }
#define SP_INT(f) if (strcmp(name, #f) == 0) { sp->f = *(int*)value; return; }
    SP_FLOAT(min_relative_decrease) SP_FLOAT(min_trust_region_radius) SP_FLOAT(max_trust_region_radius)
    SP_FLOAT(q_tolerance) SP_FLOAT(function_tolerance) SP_FLOAT(trust_region_radius)
    SP_FLOAT(radius_decrease_factor) SP_FLOAT(min_lm_diagonal) SP_FLOAT(max_lm_diagonal)
    SP_INT(residual_reset_period) SP_INT(nIter) SP_INT(nIterations) SP_INT(lIterations)
#undef SP_FLOAT
#undef SP_INT
    logSolver(plan, "Warning: tried to set nonexistent solver parameter %s\n", name);
}

/* After a fatal error that was handed to the application's handler (OptB200_SetErrorHandler) the plan is dead: the engine
 * calls below do nothing, so end the solve here - no "final cost" line, the current cost reads as NaN. */
static int plan_failed(Opt_Plan* plan) {
    if (!optk_failed(plan->k)) return 0;
    plan->prevCost = NAN;
    return 1;
}
void Opt_ProblemInit(Opt_State* state, Opt_Plan* plan, void** problemparams) {
    (void)state;
    OptK_Plan* k = plan->k;
    optk_range_push("Opt_ProblemInit");
    optk_begin_call(k);
    optk_overall_begin(k);
    optk_bind(k, problemparams);
    plan->sp.nIter = 0;
    plan->finished_reported = 0;
    if (plan->is_lm) { /* solver:1866-1869 */
        plan->trust_region_radius = of(plan, plan->sp.trust_region_radius);
        plan->radius_decrease_factor = of(plan, plan->sp.radius_decrease_factor);
        plan->min_lm_diagonal = of(plan, plan->sp.min_lm_diagonal);
        plan->max_lm_diagonal = of(plan, plan->sp.max_lm_diagonal);
        optk_set_lm(k, plan->trust_region_radius, plan->min_lm_diagonal, plan->max_lm_diagonal);
    }
    optk_precompute(k);
    plan->prevCost = optk_cost(k);
    (void)plan_failed(plan);
    /* the reference launches one more PCGInit1 here whose results the first Step overwrites
       (solver:1886); it has no observable effect and is not replayed */
    optk_end_call(k);
    optk_range_pop();
}

static void cleanup(Opt_Plan* plan) { /* solver:1902-1910 */
    logSolver(plan, "final cost=%.16f\n", plan->prevCost);
    optk_overall_end(plan->k);
    if (plan->timing && plan->verbosity > 0) optk_print_timing(plan->k);
}

static int problem_step(Opt_Plan* plan, void** problemparams);
int Opt_ProblemStep(Opt_State* state, Opt_Plan* plan, void** problemparams) {
    (void)state;
    optk_range_push("Opt_ProblemStep");
    const int more = problem_step(plan, problemparams);
    optk_range_pop();
    return more;
}
static int problem_step(Opt_Plan* plan, void** problemparams) {
    OptK_Plan* k = plan->k;
    SolverParameters* sp = &plan->sp;
    if (plan_failed(plan)) return 0;
    optk_begin_call(k);
    optk_bind(k, problemparams); /* solver:2001: the caller may change pointers/scalars between steps */
    if (sp->nIter >= sp->nIterations) {
        cleanup(plan);
        optk_end_call(k);
        return 0;
    }
    const double q_tolerance = of(plan, sp->q_tolerance);
    if (plan->is_lm) optk_set_lm(k, plan->trust_region_radius, plan->min_lm_diagonal, plan->max_lm_diagonal);
    optk_pcg_begin(k, sp->nIter);
    if (plan->is_lm && plan->lm_host_loop) {
        double Q0 = optk_fetch_q(k), Q1 = 0.0;
        int lIter = 0;
        for (; lIter < sp->lIterations; ++lIter) {
            int reset = ((lIter + 1) % sp->residual_reset_period) == 0;
            optk_pcg_iter(k, lIter, reset);
            Q1 = optk_fetch_q(k);
            double zeta = of(plan, of(plan, of(plan, (double)(lIter + 1)) * of(plan, Q1 - Q0)) / Q1);
            if (zeta < q_tolerance) {
                logSolver(plan, "zeta=%.18g < q_tol(%f), breaking at iteration: %d\n", zeta, q_tolerance, lIter + 1);
                ++lIter;
                break;
            }
            Q0 = Q1;
        }
        plan->pcg_iterations += lIter;
    } else if (plan->is_lm) {
        /* same loop, exit test on the device, no blocking fetchQ (engine.cuh: pcg_run_lm) */
        double zeta = 0.0;
        int fired = 0;
        int executed = optk_pcg_run_lm(k, sp->lIterations, sp->residual_reset_period, q_tolerance, &zeta, &fired);
        if (fired)
            logSolver(plan, "zeta=%.18g < q_tol(%f), breaking at iteration: %d\n", zeta, q_tolerance, executed);
        plan->pcg_iterations += executed;
    } else {
        optk_pcg_run(k, sp->lIterations); /* no host round trip inside the GN inner loop */
        plan->pcg_iterations += sp->lIterations;
    }

    double model_cost_change = 0.0;
    if (plan->is_lm) {
        double model_cost = optk_model_cost(k);
        logSolver(plan, " cost=%f \n", plan->prevCost);
        logSolver(plan, " model_cost=%f \n", model_cost);
        model_cost_change = of(plan, plan->prevCost - model_cost);
        logSolver(plan, " model_cost_change=%f \n", model_cost_change);
        optk_save_unknowns(k);
    }
    optk_linear_update(k);
    optk_precompute(k);
    double newCost = optk_cost(k);
    if (plan_failed(plan)) return 0;

    if (plan->is_lm) { /* solver:2240-2300 */
        double cost_change = of(plan, plan->prevCost - newCost);
        double relative_decrease = of(plan, cost_change / model_cost_change);
        if (cost_change >= 0 && relative_decrease > of(plan, sp->min_relative_decrease)) {
            double absolute_function_tolerance = of(plan, plan->prevCost * of(plan, sp->function_tolerance));
            if (cost_change <= absolute_function_tolerance) {
                logSolver(plan, "\nFunction tolerance reached, exiting\n");
                cleanup(plan);
                optk_end_call(k);
                return 0;
            }
            double step_quality = relative_decrease;
            double min_factor = 1.0 / 3.0;
            double tmp_factor = 1.0 - pow(2.0 * step_quality - 1.0, 3.0);
            plan->trust_region_radius = of(plan, plan->trust_region_radius / fmax(min_factor, tmp_factor));
            plan->trust_region_radius = of(plan, fmin(plan->trust_region_radius, of(plan, sp->max_trust_region_radius)));
            plan->radius_decrease_factor = 2.0;
            plan->prevCost = newCost;
        } else {
            optk_revert_update(k);
            plan->trust_region_radius = of(plan, plan->trust_region_radius / plan->radius_decrease_factor);
            logSolver(plan, " trust_region_radius=%f \n", plan->trust_region_radius);
            plan->radius_decrease_factor = of(plan, 2.0 * plan->radius_decrease_factor);
            if (plan->trust_region_radius <= of(plan, sp->min_trust_region_radius)) {
                logSolver(plan, "\nTrust_region_radius is less than the min, exiting\n");
                cleanup(plan);
                optk_end_call(k);
                return 0;
            }
            logSolver(plan, "REVERT\n");
            optk_precompute(k);
        }
    } else {
        plan->prevCost = newCost;
    }
    sp->nIter++;
    optk_end_call(k);
    return 1;
}

void Opt_ProblemSolve(Opt_State* state, Opt_Plan* plan, void** problemparams) { /* o.t:3328-3333 */
    Opt_ProblemInit(state, plan, problemparams);
    while (Opt_ProblemStep(state, plan, problemparams)) {
    }
}

double Opt_ProblemCurrentCost(Opt_State* state, Opt_Plan* plan) { /* solver:2354-2357 */
    (void)state;
    return plan->prevCost;
}

/* ---- OptB200.h extensions (no counterpart in the reference) -------------------------------- */
void OptB200_SetErrorHandler(OptB200_ErrorHandler handler, void* user) { optk_set_error_handler(handler, user); }
const char* OptB200_LastError(int* code) { return optk_last_error(code); }
int OptB200_PlanFailed(Opt_Plan* plan) { return plan ? optk_failed(plan->k) : 1; }
void OptB200_GetNcclUniqueId(void* out128) { optk_nccl_unique_id(out128); }
int OptB200_PlanSetSlab(Opt_Plan* plan, int rank, int nranks, int halo, const void* nccl_unique_id) {
    return optk_set_slab(plan->k, rank, nranks, halo, nccl_unique_id);
}
long OptB200_PlanLaunchCount(Opt_Plan* plan) { return optk_launch_count(plan->k); }
int OptB200_PlanMaterializedMode(Opt_Plan* plan) { return plan->materialized; }
long OptB200_PlanMaterializedNnz(Opt_Plan* plan, int which) { return optk_materialized_nnz(plan->k, which); }
long OptB200_PlanPcgIterations(Opt_Plan* plan) { return plan->pcg_iterations; }
int OptB200_PlanKernelTime(Opt_Plan* plan, const char* kernel, double* total_ms) {
    return optk_kernel_time(plan->k, kernel, total_ms);
}
void OptB200_PlanResetTiming(Opt_Plan* plan) { optk_reset_timing(plan->k); }
int OptB200_PlanUnknownChannels(Opt_Plan* plan) { return optk_unknown_channels(plan->k); }
long OptB200_PlanSiteCount(Opt_Plan* plan) { return optk_site_count(plan->k); }
int OptB200_DebugReadVector(Opt_Plan* plan, const char* name, void* host_out) {
    return optk_debug_read(plan->k, name, host_out);
}
int OptB200_DebugReadVectorRange(Opt_Plan* plan, const char* name, long first_site, long nsites, void* host_out) {
    return optk_debug_read_range(plan->k, name, first_site, nsites, host_out);
}
void OptB200_DebugApplyJTJ(Opt_Plan* plan, void** problemparams, const void* host_p, void* host_out) {
    optk_begin_call(plan->k);
    optk_bind(plan->k, problemparams);
    optk_precompute(plan->k);
    optk_debug_apply_jtj(plan->k, host_p, host_out);
    optk_end_call(plan->k);
}
void OptB200_DebugPcgBegin(Opt_Plan* plan, void** problemparams) {
    optk_begin_call(plan->k);
    optk_bind(plan->k, problemparams);
    optk_precompute(plan->k);
    optk_pcg_begin(plan->k, 0);
    optk_end_call(plan->k);
}
