// Instantiates the generic engine for poisson_image_editing (fp32 and fp64).
#include "../engine.cuh"
#include "../functors/poisson.cuh"

extern "C" OptK_Plan* optk_make_poisson_image_editing(int is_double, int is_lm, const unsigned* dims, int timing) {
    return optk::make_engine<optk::PoissonFunctor>(is_double, is_lm, dims, timing);
}
