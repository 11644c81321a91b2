// Instantiates the generic engine for shape_from_shading (fp32 and fp64).
#include "../engine.cuh"
#include "../functors/sfs.cuh"

extern "C" OptK_Plan* optk_make_shape_from_shading(int is_double, int is_lm, const unsigned* dims, int timing) {
    return optk::make_engine<optk::SfsFunctor>(is_double, is_lm, dims, timing);
}
