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

/* Kernels launched by this plan since it was created (bench.py's "gpu_launches"). */
long OptB200_PlanLaunchCount(Opt_Plan* plan);
/* With collectPerKernelTimingInfo set: number of launches of `kernel` (the reference's kernel
 * names, e.g. "PCGStep1") since the last reset and their summed cudaEvent duration in ms. */
int  OptB200_PlanKernelTime(Opt_Plan* plan, const char* kernel, double* total_ms);
void OptB200_PlanResetTiming(Opt_Plan* plan);
int  OptB200_PlanUnknownChannels(Opt_Plan* plan);   /* U */
long OptB200_PlanSiteCount(Opt_Plan* plan);         /* N */

/* Test hooks.  `name` is one of "delta","r","p","Ap","z","pre","CtC","b","prevX"; the vector is
 * written to host memory as AoS [site][U] of opt_float.  Returns 0 on success. */
int  OptB200_DebugReadVector(Opt_Plan* plan, const char* name, void* host_out);
/* host_out = J^T J host_p at the current unknowns (host AoS [site][U]); runs the PCGStep1 kernel(s). */
void OptB200_DebugApplyJTJ(Opt_Plan* plan, void** problemparams, const void* host_p, void* host_out);
/* Runs PCGInit1 (+graph/finish, +LM diagonal) at the current unknowns so that r = -J^T F, pre, p can be read back. */
void OptB200_DebugPcgBegin(Opt_Plan* plan, void** problemparams);

#ifdef __cplusplus
}
#endif
#endif
