// Instantiates the generic engine for volumetric_mesh_deformation (fp32 and fp64).
#include "../engine.cuh"
#include "../functors/volumetric_packed.cuh"

extern "C" OptK_Plan* optk_make_volumetric_mesh_deformation(int is_double, int is_lm, const unsigned* dims, int timing) {
    return optk::make_engine<optk::VolumetricFunctor>(is_double, is_lm, dims, timing);
}
