// Instantiates the generic engine for embedded_mesh_deformation (fp32 and fp64).
#include "../engine.cuh"
#include "../functors/embedded.cuh"

extern "C" OptK_Plan* optk_make_embedded_mesh_deformation(int is_double, int is_lm, const unsigned* dims, int timing) {
    return optk::make_engine<optk::EmbeddedFunctor>(is_double, is_lm, dims, timing);
}
