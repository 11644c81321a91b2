/* Opt.h -- C API of the B200-native Gauss-Newton / Levenberg-Marquardt + PCG engine.
 *
 * This is the drop-in boundary: every declaration below has the same name, argument
 * order, argument meaning and struct layout as the reference's public header
 * (reference: API/release/include/Opt.h:3-76; struct mirrored by
 * API/src/createwrapper.t:98-134), so that the reference's example drivers
 * (examples/shared/OptSolver.h:46-106, examples/shared/OptUtils.h:47-64,108-112)
 * link against libOptB200.so without source changes.
 *
 * Behavioural contract (see DESIGN.md and SURVEY.md section 8b):
 *  - images in `problemparams` are raw device arrays, AoS per element, x fastest;
 *    unknown images are updated in place;
 *  - scalar parameters are host `float*` and are re-read at Init and at every Step;
 *  - a graph contributes `int* edgeCount` followed by one `int*` index array per
 *    hyper-edge vertex; the edge count actually used is dimensions[k] of the graph;
 *  - only `backend_cuda` is accepted (empty string means backend_cuda); there is no
 *    CPU fallback in this library;
 *  - an energy file that is not one of the ten shipped examples is rejected:
 *    Opt_ProblemDefine prints a message on stderr and returns NULL.
 */
#ifndef OPT_H_B200
#define OPT_H_B200

#ifdef __cplusplus
extern "C" {
#endif

typedef struct Opt_State   Opt_State;
typedef struct Opt_Plan    Opt_Plan;
typedef struct Opt_Problem Opt_Problem;

/* Set once per Opt_NewState. A zeroed struct is a valid default.
 * (reference: API/release/include/Opt.h:10-35) */
struct Opt_InitializationParameters {
    int  doublePrecision;            /* nonzero: unknowns, opt_float images and all solver vectors are double */
    int  verbosityLevel;             /* 0 silent, 1 prints `final cost=` and LM diagnostics */
    int  collectPerKernelTimingInfo; /* nonzero: cudaEvent pair per kernel launch, table printed at the end */
    char backend[20];                /* "backend_cuda" (or empty). "backend_cpu", "backend_cpu_mt" are rejected */
    int  numthreads;                 /* ignored (CPU-MT backend only in the reference) */
    int  useMaterializedJTJ;         /* PCGStep1 over a stored Jacobian: J^T (J p) through CSR kernels (solver:1523-1761) */
    int  useFusedJTJ;                /* with useMaterializedJTJ: the product J^T J is formed and PCGStep1 = (J^T J) p */
};
typedef struct Opt_InitializationParameters Opt_InitializationParameters;

/* reference: Opt.h:40 / createwrapper.t:150-287 */
Opt_State* Opt_NewState(Opt_InitializationParameters params);

/* `filename` is an energy file name; its basename selects a built-in functor set.
 * `solverkind`: "gaussNewtonGPU" or any string containing "LM" for Levenberg-Marquardt
 * (reference: Opt.h:45, o.t:169-180, o.t:474). */
Opt_Problem* Opt_ProblemDefine(Opt_State* state, const char* filename, const char* solverkind);
void         Opt_ProblemDelete(Opt_State* state, Opt_Problem* problem);

/* Allocates all solver vectors for `dimensions` (indexed like Dim(name,i) in the energy
 * file; graph examples pass {N, numEdges}). (reference: Opt.h:51, o.t:1351-1528) */
Opt_Plan* Opt_ProblemPlan(Opt_State* state, Opt_Problem* problem, unsigned int* dimensions);
void      Opt_PlanFree(Opt_State* state, Opt_Plan* plan);

/* Name-keyed solver parameter; `value` points to a float or an int depending on the
 * parameter (reference: Opt.h:56, solverGPUGaussNewton.t:186-201, 2382-2398). */
void Opt_SetSolverParameter(Opt_State* state, Opt_Plan* plan, const char* name, void* value);

/* Init followed by Step until it returns 0 (reference: Opt.h:61, o.t:3328-3333). */
void Opt_ProblemSolve(Opt_State* state, Opt_Plan* plan, void** problemparams);

/* reference: Opt.h:68-71, solverGPUGaussNewton.t:1766-1897 (init), 1913-2349 (step) */
void Opt_ProblemInit(Opt_State* state, Opt_Plan* plan, void** problemparams);
int  Opt_ProblemStep(Opt_State* state, Opt_Plan* plan, void** problemparams);

/* Cost at the current unknowns as of the last Init/Step (reference: Opt.h:76,
 * solverGPUGaussNewton.t:2354-2357). */
double Opt_ProblemCurrentCost(Opt_State* state, Opt_Plan* plan);

#ifdef __cplusplus
}
#endif
#endif /* OPT_H_B200 */
