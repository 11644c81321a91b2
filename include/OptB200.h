/* OptB200.h -- extensions of the B200 engine that have no counterpart in the reference's Opt.h.
 * Nothing here is needed by a drop-in user of Opt.h; these entry points exist for resource
 * clean-up, for measurement (bench.py) and for the parity tests (tests/), which compare
 * individual solver vectors with the CPU oracle. */
#ifndef OPT_B200_H
#define OPT_B200_H
#include "Opt.h"
#ifdef __cplusplus
extern "C" {
#endif

/* The reference never frees an Opt_State (createwrapper.t has no destructor); this one can be. */
void OptB200_FreeState(Opt_State* state);

/* ---- fatal errors -----------------------------------------------------------------------------
 * The reference prints "Cuda reported error ..." and calls exit() on any CUDA failure (backend_cuda.t:26-40, 767-779), and
 * so does this library as long as no handler is installed - drivers written for the reference see what they always saw.
 * An application that must survive (a server, a host with its own clean-up) installs a handler: the failing call then unwinds
 * to the C-ABI boundary, the handler is called once - (code, message, user): the cudaError_t value, 200 + ncclResult_t, or a
 * small library code, with the text the reference would have printed - and the call returns: Opt_ProblemPlan returns NULL
 * (everything allocated so far is released), Opt_ProblemInit / Step / Solve return at once (Step with 0),
 * Opt_ProblemCurrentCost reads NaN and OptB200_PlanFailed 1.  A failed plan can only be freed (Opt_PlanFree); other plans are
 * not affected unless the CUDA context itself is gone.  The handler may be called from any thread that calls into the library
 * and must not call back into it, except OptB200_LastError. */
typedef void (*OptB200_ErrorHandler)(int code, const char* message, void* user);
void OptB200_SetErrorHandler(OptB200_ErrorHandler handler, void* user);   /* NULL restores print-and-exit */
const char* OptB200_LastError(int* code);     /* the calling thread's most recent fatal error ("" if none); code may be NULL */
int  OptB200_PlanFailed(Opt_Plan* plan);

/* ---- multi-GPU domain decomposition (new; the reference is single-device) ---------------------
 * One process per GPU.  A 2-D / 3-D grid problem is cut into slabs along its slowest dimension
 * (d1 for 2-D, d2 for 3-D).  Each rank plans its LOCAL slab: its owned rows/planes plus `halo`
 * ghost rows/planes on every side that faces another rank, i.e. `dimensions` passed to
 * Opt_ProblemPlan and every array in `problemparams` cover owned + ghost rows.  The caller fills
 * all arrays (ghost rows included) before Opt_ProblemInit; afterwards the engine keeps the ghost
 * rows of the unknown images current (NCCL send/recv after every update), exchanges the PCG
 * search direction once per PCG iteration and all-reduces the PCG scalars and the costs.
 * halo = 1 for every shipped grid example except shape_from_shading (2).
 * Usage: rank 0 calls OptB200_GetNcclUniqueId and sends the 128 bytes to every rank (MPI,
 * torch.distributed, a file ...); then EVERY rank calls OptB200_PlanSetSlab (collective) once,
 * after Opt_ProblemPlan and before Opt_ProblemInit.  Returns 0 on success.
 * The ranks are expected on one node: PlanSetSlab maps a mailbox (and, for energies whose fused kernel supports it, the
 * solver-vector pool) of every peer with cudaIpc and then all-reduces the PCG scalars and pushes boundary rows by direct
 * stores over NVLink; if the mapping fails on any rank, all ranks fall back to NCCL (OPTB200_NO_P2P=1 forces that). */
void OptB200_GetNcclUniqueId(void* out128);
int  OptB200_PlanSetSlab(Opt_Plan* plan, int rank, int nranks, int halo, const void* nccl_unique_id);

/* ---- energy-file verification (a restricted front end; the reference compiles the file, o.t:3305-3309) ----------
 * OptB200_EnergyDescribe parses the Lua/Terra energy file at `path` far enough to list what fixes the C-ABI binding
 * -- Dim / Unknown / Array / Param / Graph declarations with types and problemparams indices, ComputedArrays,
 * UsePreconditioner, Exclude -- followed by a fingerprint of the comment- and space-free source, one item per line
 * into `out`.  Returns 0, 1 if the file cannot be read, -1 on overflow / parse failure.  OptB200_EnergyExpected
 * returns the description the built-in functor of `energy` (basename without ".t") was written for, or NULL.
 * Opt_ProblemDefine rejects a readable file whose description differs (OPTB200_ALLOW_MODIFIED_ENERGY=1 accepts a file
 * whose declarations match but whose expressions were edited); a path that cannot be read is bound by its name. */
int OptB200_EnergyDescribe(const char* path, char* out, unsigned long cap);
const char* OptB200_EnergyExpected(const char* energy);

/* Materialized-Jacobian mode in effect for this plan (Opt_InitializationParameters.useMaterializedJTJ / useFusedJTJ,
 * solverGPUGaussNewton.t:1523-1761): 0 = matrix-free kernels, 1 = PCGStep1 as J^T (J p) over stored CSR matrices, 2 = as
 * (J^T J) p.  Energies without a Jacobian-row emitter stay at 0 (the iterates are the same in all three modes, as
 * examples/test_final_cost.py:91-93 requires).  Nnz: 0 = nnz(J), 1 = nnz(J^T J), valid after the first Opt_ProblemStep. */
int  OptB200_PlanMaterializedMode(Opt_Plan* plan);
long OptB200_PlanMaterializedNnz(Opt_Plan* plan, int which);

/* Kernels launched by this plan since it was created (bench.py's "gpu_launches"). */
long OptB200_PlanLaunchCount(Opt_Plan* plan);
/* PCG iterations executed by this plan since it was created.  Gauss-Newton always runs nIterations x lIterations;
 * Levenberg-Marquardt leaves the inner loop when its q_tolerance test fires (solverGPUGaussNewton.t:2180-2200). */
long OptB200_PlanPcgIterations(Opt_Plan* plan);
/* With collectPerKernelTimingInfo set: number of launches of `kernel` (the reference's kernel
 * names, e.g. "PCGStep1") since the last reset and their summed cudaEvent duration in ms. */
int  OptB200_PlanKernelTime(Opt_Plan* plan, const char* kernel, double* total_ms);
void OptB200_PlanResetTiming(Opt_Plan* plan);
int  OptB200_PlanUnknownChannels(Opt_Plan* plan);   /* U */
long OptB200_PlanSiteCount(Opt_Plan* plan);         /* N */

/* Test hooks.  `name` is one of "delta","r","p","Ap","z","pre","CtC","b","prevX"; the vector is
 * written to host memory as AoS [site][U] of opt_float.  Returns 0 on success. */
int  OptB200_DebugReadVector(Opt_Plan* plan, const char* name, void* host_out);
/* The same for the sites [first_site, first_site + nsites) only: full-size checks (512^3) read a few planes. */
int  OptB200_DebugReadVectorRange(Opt_Plan* plan, const char* name, long first_site, long nsites, void* host_out);
/* host_out = J^T J host_p at the current unknowns (host AoS [site][U]); runs the PCGStep1 kernel(s). */
void OptB200_DebugApplyJTJ(Opt_Plan* plan, void** problemparams, const void* host_p, void* host_out);
/* Runs PCGInit1 (+graph/finish, +LM diagonal) at the current unknowns so that r = -J^T F, pre, p can be read back. */
void OptB200_DebugPcgBegin(Opt_Plan* plan, void** problemparams);

#ifdef __cplusplus
}
#endif
#endif
