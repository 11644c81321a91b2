/* TEST INFRASTRUCTURE ONLY -- C ABI of the CPU oracle (see oracle/README.md).
 * Same call sequence and argument meaning as include/Opt.h (reference:
 * API/release/include/Opt.h:40-76), but every pointer in `problemparams` is a HOST pointer
 * and the functions carry the OptOracle_ prefix so the oracle can never be mistaken for, or
 * linked in place of, the product library.  Only tests/, __graft_entry__.smoke() and
 * bench.py's cpu_baseline / --impl reference legs may load this library. */
#ifndef OPT_ORACLE_H
#define OPT_ORACLE_H
#include "../include/Opt.h"
#ifdef __cplusplus
extern "C" {
#endif
typedef struct OptOracle_State OptOracle_State;
typedef struct OptOracle_Problem OptOracle_Problem;
typedef struct OptOracle_Plan OptOracle_Plan;

OptOracle_State*   OptOracle_NewState(Opt_InitializationParameters params);   /* numthreads = OpenMP threads */
void               OptOracle_FreeState(OptOracle_State*);
OptOracle_Problem* OptOracle_ProblemDefine(OptOracle_State*, const char* filename, const char* solverkind);
void               OptOracle_ProblemDelete(OptOracle_State*, OptOracle_Problem*);
OptOracle_Plan*    OptOracle_ProblemPlan(OptOracle_State*, OptOracle_Problem*, unsigned int* dimensions);
void               OptOracle_PlanFree(OptOracle_State*, OptOracle_Plan*);
void               OptOracle_SetSolverParameter(OptOracle_State*, OptOracle_Plan*, const char* name, void* value);
void               OptOracle_ProblemSolve(OptOracle_State*, OptOracle_Plan*, void** problemparams);
void               OptOracle_ProblemInit(OptOracle_State*, OptOracle_Plan*, void** problemparams);
int                OptOracle_ProblemStep(OptOracle_State*, OptOracle_Plan*, void** problemparams);
double             OptOracle_ProblemCurrentCost(OptOracle_State*, OptOracle_Plan*);

/* test hooks (no counterpart in Opt.h) */
long   OptOracle_UnknownCount(OptOracle_Plan*);                       /* N * U */
long   OptOracle_PcgIterations(OptOracle_Plan*);                     /* PCG iterations executed so far (LM exits early) */
int    OptOracle_CostHistory(OptOracle_Plan*, double* out, int cap);  /* cost after Init and after each Step */
void   OptOracle_ApplyJTJ(OptOracle_Plan*, void** problemparams, const void* p, void* Ap);   /* AoS [site][U] */
void   OptOracle_EvalJTF(OptOracle_Plan*, void** problemparams, void* minusJTF, void* preconditioner);
#ifdef __cplusplus
}
#endif
#endif
