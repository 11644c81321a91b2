// Instantiates the generic engine for intrinsic_image_decomposition (fp32 and fp64).
#include "../engine.cuh"
#include "../functors/intrinsic.cuh"

extern "C" OptK_Plan* optk_make_intrinsic_image_decomposition(int is_double, int is_lm, const unsigned* dims, int timing) {
    return optk::make_engine<optk::IntrinsicFunctor>(is_double, is_lm, dims, timing);
}
