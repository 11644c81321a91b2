// TEST INFRASTRUCTURE: as pre.h, for the reference's hand-written .cu solvers (linked because CombinedSolver.h
// constructs them; never run by the tests).  They use the pre-CUDA-9 warp shuffles, which no longer exist.
#include "pre.h"
#define __shfl_down(...) __shfl_down_sync(0xffffffffu, __VA_ARGS__)
#define __shfl_xor(...) __shfl_xor_sync(0xffffffffu, __VA_ARGS__)
#define __shfl(...) __shfl_sync(0xffffffffu, __VA_ARGS__)
