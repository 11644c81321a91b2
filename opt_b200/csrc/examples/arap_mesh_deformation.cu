// Instantiates the generic engine for arap_mesh_deformation (fp32 and fp64).
#include "../engine.cuh"
#include "../functors/arap.cuh"

extern "C" OptK_Plan* optk_make_arap_mesh_deformation(int is_double, int is_lm, const unsigned* dims, int timing) {
    return optk::make_engine<optk::ArapFunctor>(is_double, is_lm, dims, timing);
}
