/* sfs_driver.c -- the reference's shape_from_shading example driver, restated in plain C against the UNCHANGED Opt.h
 * API, to show the library dropping into a C/C++ driver: it is linked against libOptB200.so exactly as the reference's
 * driver is linked against libOpt.a.
 *
 * Follows examples/shape_from_shading/src/main.cpp:6-86 (load the data set, solve, save the refined depth),
 * SFSSolverInput.h:27-90 (which file feeds which problem parameter, in which order), CombinedSolver.h:26-75 (the
 * unknown starts from the initialUnknown image) and examples/shared/OptSolver.h:43-106 (NewState / ProblemDefine /
 * ProblemPlan / SetSolverParameter / ProblemSolve).  mLib / OpenMesh are not needed: the .imagedump and
 * .SFSSolverParameters files are read with include/OptB200Formats.h.
 *
 *   sfs_driver <prefix> [--nIterations N] [--lIterations L] [--solver gaussNewtonGPU|LMGPU] [--out file.imagedump]
 * reads <prefix>_{targetIntensity,targetDepth,initialUnknown,maskEdgeMap}.imagedump and <prefix>.SFSSolverParameters.
 * The library prints "final cost=..." (verbosityLevel 1), the line the reference's tests parse. */
#include <cuda_runtime.h>
#include <stdio.h>
#include <stdlib.h>
#include <string.h>

#include "../include/Opt.h"
#include "../include/OptB200Formats.h"

#define CHECK_CUDA(call) do { cudaError_t e_ = (call); if (e_ != cudaSuccess) { fprintf(stderr, "%s: %s\n", #call, cudaGetErrorString(e_)); return 2; } } while (0)

static int load(const char* prefix, const char* suffix, OptB200_ImageDump* im) {
    char path[1024];
    snprintf(path, sizeof(path), "%s_%s.imagedump", prefix, suffix);
    const int rc = OptB200_ImageDumpRead(path, 1, im);          /* SimpleBuffer's default: clampInfinity = true */
    if (rc) fprintf(stderr, "cannot read %s (%d)\n", path, rc);
    return rc;
}

int main(int argc, char** argv) {
    if (argc < 2 || !strcmp(argv[1], "--help")) {
        printf("usage: %s <prefix> [--nIterations N] [--lIterations L] [--solver gaussNewtonGPU|LMGPU] [--out file.imagedump]\n", argv[0]);
        return argc < 2 ? 1 : 0;
    }
    const char* prefix = argv[1];
    const char* solver = "gaussNewtonGPU";
    const char* out = "sfsOutput.imagedump";
    unsigned int nIterations = 60, lIterations = 10;             /* the performance-run setting, main.cpp:60-61 */
    for (int i = 2; i + 1 < argc; i += 2) {
        if (!strcmp(argv[i], "--nIterations")) nIterations = (unsigned)atoi(argv[i + 1]);
        else if (!strcmp(argv[i], "--lIterations")) lIterations = (unsigned)atoi(argv[i + 1]);
        else if (!strcmp(argv[i], "--solver")) solver = argv[i + 1];
        else if (!strcmp(argv[i], "--out")) out = argv[i + 1];
        else { fprintf(stderr, "unknown option %s\n", argv[i]); return 1; }
    }

    OptB200_ImageDump intensity, depth, initial, masks;
    if (load(prefix, "targetIntensity", &intensity) || load(prefix, "targetDepth", &depth) ||
        load(prefix, "initialUnknown", &initial) || load(prefix, "maskEdgeMap", &masks)) return 1;
    OptB200_SFSParameters P;
    char ppath[1024];
    snprintf(ppath, sizeof(ppath), "%s.SFSSolverParameters", prefix);
    if (OptB200_SFSParametersRead(ppath, &P)) { fprintf(stderr, "cannot read %s\n", ppath); return 1; }
    const int W = initial.width, H = initial.height;
    const size_t n = (size_t)W * H;
    if (masks.datatype != 1 || masks.width != W || masks.height != 2 * H || depth.width != W || intensity.width != W) {
        fprintf(stderr, "inconsistent data set\n");
        return 1;
    }
    {   /* SFSSolverInput.h:76-84 */
        size_t active = 0;
        for (size_t i = 0; i < n; ++i) active += ((const float*)depth.data)[i] > 0.0f;
        printf("Num Active Unknowns: %zu\n", active);
    }

    float *dX, *dDepth, *dIm;
    unsigned char *dMaskR, *dMaskC;
    CHECK_CUDA(cudaMalloc((void**)&dX, n * sizeof(float)));
    CHECK_CUDA(cudaMalloc((void**)&dDepth, n * sizeof(float)));
    CHECK_CUDA(cudaMalloc((void**)&dIm, n * sizeof(float)));
    CHECK_CUDA(cudaMalloc((void**)&dMaskR, n));
    CHECK_CUDA(cudaMalloc((void**)&dMaskC, n));
    CHECK_CUDA(cudaMemcpy(dX, initial.data, n * sizeof(float), cudaMemcpyHostToDevice));         /* resetGPUMemory */
    CHECK_CUDA(cudaMemcpy(dDepth, depth.data, n * sizeof(float), cudaMemcpyHostToDevice));
    CHECK_CUDA(cudaMemcpy(dIm, intensity.data, n * sizeof(float), cudaMemcpyHostToDevice));
    CHECK_CUDA(cudaMemcpy(dMaskR, masks.data, n, cudaMemcpyHostToDevice));                         /* rows 0..H-1 */
    CHECK_CUDA(cudaMemcpy(dMaskC, (const unsigned char*)masks.data + n, n, cudaMemcpyHostToDevice));

    /* NamedParameters::data() order = first-set order of SFSSolverInput::setParameters (shape_from_shading.t:4-19) */
    void* problemparams[21];
    problemparams[0] = &P.weightFitting;      /* w_p */
    problemparams[1] = &P.weightRegularizer;  /* w_s */
    problemparams[2] = &P.weightShading;      /* w_g */
    problemparams[3] = &P.fx; problemparams[4] = &P.fy; problemparams[5] = &P.ux; problemparams[6] = &P.uy;
    for (int i = 0; i < 9; ++i) problemparams[7 + i] = &P.lightingCoefficients[i];
    problemparams[16] = dX; problemparams[17] = dDepth; problemparams[18] = dIm;
    problemparams[19] = dMaskR; problemparams[20] = dMaskC;

    Opt_InitializationParameters ip;
    memset(&ip, 0, sizeof(ip));
    ip.verbosityLevel = 1;
    strcpy(ip.backend, "backend_cuda");
    Opt_State* state = Opt_NewState(ip);
    Opt_Problem* problem = state ? Opt_ProblemDefine(state, "shape_from_shading.t", solver) : NULL;
    unsigned int dims[2] = {(unsigned)W, (unsigned)H};
    Opt_Plan* plan = problem ? Opt_ProblemPlan(state, problem, dims) : NULL;
    if (!plan) { fprintf(stderr, "Opt_ProblemPlan failed\n"); return 3; }                         /* OptSolver.h:67-69 asserts */
    Opt_SetSolverParameter(state, plan, "nIterations", &nIterations);
    Opt_SetSolverParameter(state, plan, "lIterations", &lIterations);
    printf("Solving\n");
    Opt_ProblemSolve(state, plan, problemparams);
    CHECK_CUDA(cudaDeviceSynchronize());
    printf("Solved\n");

    OptB200_ImageDump result = {W, H, 1, 0, malloc(n * sizeof(float))};
    CHECK_CUDA(cudaMemcpy(result.data, dX, n * sizeof(float), cudaMemcpyDeviceToHost));
    if (OptB200_ImageDumpWrite(out, &result)) { fprintf(stderr, "cannot write %s\n", out); return 1; }   /* result->save("sfsOutput.imagedump") */
    printf("Opt_ProblemCurrentCost=%.9g\n", Opt_ProblemCurrentCost(state, plan));

    Opt_PlanFree(state, plan);
    Opt_ProblemDelete(state, problem);
    free(result.data);
    OptB200_FormatFree(intensity.data); OptB200_FormatFree(depth.data); OptB200_FormatFree(initial.data); OptB200_FormatFree(masks.data);
    cudaFree(dX); cudaFree(dDepth); cudaFree(dIm); cudaFree(dMaskR); cudaFree(dMaskC);
    return 0;
}
