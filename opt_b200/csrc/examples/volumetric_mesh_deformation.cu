// Instantiates the generic engine for volumetric_mesh_deformation: fp64 and fp32 with SoA solver vectors and the
// scalar marching kernel, and fp32 with blocked vectors and the packed kernel fed by bulk copies for grids whose rows
// are a multiple of 32 sites (OPTB200_VOL_SCALAR=1 at plan time forces the former).
#include "../engine.cuh"
#include "../functors/volumetric_packed.cuh"

extern "C" OptK_Plan* optk_make_volumetric_mesh_deformation(int is_double, int is_lm, const unsigned* dims, int timing) {
    if (!is_double && dims[0] % 32 == 0 && getenv("OPTB200_VOL_SCALAR") == nullptr)
        return optk::build_engine<optk::VolumetricPackedFunctor<float>>(is_lm, dims, timing);
    return optk::make_engine<optk::VolumetricScalarFunctor>(is_double, is_lm, dims, timing);
}
