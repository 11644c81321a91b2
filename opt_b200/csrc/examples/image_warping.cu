// Instantiates the generic engine for image_warping (fp32 and fp64).
#include "../engine.cuh"
#include "../functors/image_warping.cuh"

extern "C" OptK_Plan* optk_make_image_warping(int is_double, int is_lm, const unsigned* dims, int timing) {
    return optk::make_engine<optk::ImageWarpingFunctor>(is_double, is_lm, dims, timing);
}
