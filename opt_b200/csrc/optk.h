/* optk.h -- the thin C ABI between the C host (opt_api.c) and the CUDA engine (engine.cuh and the
 * per-example translation units).  Plain pointers and scalars only.
 *
 * One OptK_Plan corresponds to the reference's PlanData (solverGPUGaussNewton.t:208-272): the
 * solver vectors, the device-resident reduction scalars and the bound problem parameters.
 * Each function launches the kernel(s) that replace one reference kernel or kernel group; the
 * citation next to each gives the reference code it stands in for. */
#ifndef OPTK_H
#define OPTK_H
#ifdef __cplusplus
extern "C" {
#endif

typedef struct OptK_Plan OptK_Plan;

/* Returns NULL (after printing why) if `energy` has no built-in functor set. */
OptK_Plan* optk_create(const char* energy, int is_double, int is_lm, const unsigned* dims,
                       int collect_timing);
void  optk_destroy(OptK_Plan*);
int   optk_has_energy(const char* energy);

/* util.initParameters (util.t:677-721): image pointers, host scalars, graph index arrays */
void  optk_bind(OptK_Plan*, void** problemparams);
/* LM parameters that the reference injects as problem parameters (o.t:1575-1580) */
void  optk_set_lm(OptK_Plan*, double trust_region_radius, double min_lm_diagonal, double max_lm_diagonal);

void   optk_range_push(const char* name);     /* NVTX range (no-op without a profiler) */
void   optk_range_pop(void);
void   optk_overall_begin(OptK_Plan*);        /* the reference's "overall" timer event (solver:1813): started at Init ... */
void   optk_overall_end(OptK_Plan*);          /* ... and closed when the solve is cleaned up (solver:1906) */
void   optk_begin_call(OptK_Plan*);           /* order the plan's stream after the caller's default stream */
void   optk_end_call(OptK_Plan*);             /* ... and the default stream after the plan's stream */

void   optk_precompute(OptK_Plan*);           /* solver:1028-1039 */
double optk_cost(OptK_Plan*);                 /* computeCost + computeCost_Graph + D2H (solver:1389-1441) */
void   optk_pcg_begin(OptK_Plan*, int nIter); /* PCGInit1[_Graph,_Finish]; LM: SaveSSq, ComputeCtC[_Graph], FinalizeDiagonal */
void   optk_pcg_iter(OptK_Plan*, int lIter, int residual_reset);  /* PCGStep1[_Graph], Step2 (or 1stHalf/Adelta/2ndHalf), Step3 */
void   optk_pcg_run(OptK_Plan*, int lIterations);                  /* GN: the whole inner loop, no host sync */
/* LM: the whole inner loop with the q_tolerance exit test (solver:2180-2200) evaluated on the device; the host
 * polls a pinned mailbox one iteration behind the GPU.  Returns the iterations executed (as the reference counts
 * them); *fired = 1 and *zeta = the value below q_tolerance when the test ended the loop. */
int    optk_pcg_run_lm(OptK_Plan*, int lIterations, int residual_reset_period, double q_tolerance, double* zeta, int* fired);
double optk_fetch_q(OptK_Plan*);              /* fetchQ (solver:1480-1501) */
double optk_model_cost(OptK_Plan*);           /* computeModelCost[_Graph] + D2H (solver:1445-1476) */
void   optk_save_unknowns(OptK_Plan*);        /* savePreviousUnknowns (solver:882-887) */
void   optk_linear_update(OptK_Plan*);        /* PCGLinearUpdate (solver:854-859) */
void   optk_revert_update(OptK_Plan*);        /* revertUpdate (solver:864-869) */

/* Multi-GPU slab mode (one process per GPU).  The plan's dimensions are the LOCAL slab including
 * `halo` ghost rows/planes on every interior side of the slowest dimension; `nccl_unique_id` is the
 * 128-byte ncclUniqueId every rank received from rank 0.  Collective: every rank must call it. */
int    optk_set_slab(OptK_Plan*, int rank, int nranks, int halo, const void* nccl_unique_id);
void   optk_nccl_unique_id(void* out128);

/* Materialized Jacobian (Opt_InitializationParameters.useMaterializedJTJ / useFusedJTJ; solver:1523-1761): mode 1 =
 * PCGStep1 as J^T (J p), mode 2 = (J^T J) p, both through the CSR kernels of csr.cuh.  Returns the mode in effect (0 if the
 * energy has no Jacobian-row emitter: matrix-free, same iterates). */
int    optk_set_materialized(OptK_Plan*, int mode);
long   optk_materialized_nnz(OptK_Plan*, int which);   /* 0: nnz(J), 1: nnz(J^T J) */

/* Fatal errors (common.cuh): with no handler a CUDA / NCCL failure prints and exits as in the reference (backend_cuda.t:26-40);
 * with one, the failing call unwinds to this layer, the handler is called once with the error code and message, the plan is
 * marked failed (optk_failed) and every later call on it returns at once (costs read as NaN, counts as 0, status as -1). */
void   optk_set_error_handler(void (*handler)(int code, const char* message, void* user), void* user);
const char* optk_last_error(int* code);       /* message (and code) of the calling thread's most recent fatal error; "" if none */
int    optk_failed(OptK_Plan*);

/* statistics and debugging */
long   optk_launch_count(OptK_Plan*);         /* kernels launched since creation */
void   optk_print_timing(OptK_Plan*);         /* Kernel | Count | Total | Average table (backend_cuda.t:152-297) */
int    optk_kernel_time(OptK_Plan*, const char* kernel, double* total_ms);   /* returns launch count */
void   optk_reset_timing(OptK_Plan*);
int    optk_unknown_channels(OptK_Plan*);
long   optk_site_count(OptK_Plan*);
/* copies an internal solver vector ("delta","r","p","Ap","z","pre","CtC") to the host as AoS [site][U] of opt_float */
int    optk_debug_read(OptK_Plan*, const char* name, void* host_out);
int    optk_debug_read_range(OptK_Plan*, const char* name, long first_site, long nsites, void* host_out);
/* out = J^T J p for a host AoS [site][U] direction (test hook; runs PCGStep1 on it) */
void   optk_debug_apply_jtj(OptK_Plan*, const void* host_p, void* host_out);

#ifdef __cplusplus
}
#endif
#endif
