// Instantiates the generic engine for robust_nonrigid_alignment (fp32 and fp64).
#include "../engine.cuh"
#include "../functors/robust.cuh"

extern "C" OptK_Plan* optk_make_robust_nonrigid_alignment(int is_double, int is_lm, const unsigned* dims, int timing) {
    return optk::make_engine<optk::RobustFunctor>(is_double, is_lm, dims, timing);
}
