// Instantiates the generic engine for cotangent_mesh_smoothing (fp32 and fp64).
#include "../engine.cuh"
#include "../functors/cotangent.cuh"

extern "C" OptK_Plan* optk_make_cotangent_mesh_smoothing(int is_double, int is_lm, const unsigned* dims, int timing) {
    return optk::make_engine<optk::CotangentFunctor>(is_double, is_lm, dims, timing);
}
