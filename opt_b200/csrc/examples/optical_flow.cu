// Instantiates the generic engine for optical_flow (fp32 and fp64).
#include "../engine.cuh"
#include "../functors/optical_flow.cuh"

extern "C" OptK_Plan* optk_make_optical_flow(int is_double, int is_lm, const unsigned* dims, int timing) {
    return optk::make_engine<optk::OpticalFlowFunctor>(is_double, is_lm, dims, timing);
}
