// TEST INFRASTRUCTURE (oracle/_ref/iw_raster): the reference's OWN rasteriser of the warped image_warping mesh --
// CombinedSolver::copyResultToCPU / rasterizeTriangle / PointInTriangleLK, examples/image_warping/src/CombinedSolver.h:63-99,
// 231-308 -- compiled from the reference's header where it lies (nothing copied) and driven with arrays from a file.  The
// solver object is built on the CPU location with every solver switched off; only its image members are used.
//   iw_raster in.bin out.bin      in: warp [H,W,2] f32, color [H,W,3] f32, mask [H,W] f32     out: result [H,W,3] f32
#include <algorithm>
#include <fstream>
#include <iostream>
#include <memory>
#include <sstream>
#include <string>
#include <vector>
#include "io.h"
#define private public
#define protected public
#include "CombinedSolver.h"          // -I <reference>/examples/image_warping/src
#undef private
#undef protected

int main(int argc, char** argv) {
    if (argc < 3) return 2;
    std::vector<Arr> in = read_arrays(argv[1]);
    const int H = in[0].dims[0], W = in[0].dims[1];
    const float* warp = in[0].as<float>(); const float* color = in[1].as<float>(); const float* mask = in[2].as<float>();
    ColorImageR32 image(W, H), imageMask(W, H);
    ColorImageR32G32B32 imageColor(W, H);
    for (int y = 0; y < H; ++y)
        for (int x = 0; x < W; ++x) {
            image(x, y) = 0.0f;
            imageMask(x, y) = mask[y * W + x];
            imageColor(x, y) = vec3f(color[(y * W + x) * 3], color[(y * W + x) * 3 + 1], color[(y * W + x) * 3 + 2]);
        }
    std::vector<std::vector<int>> constraints;
    CombinedSolverParameters params;
    params.useOpt = false; params.useOptLM = false; params.useCeres = false; params.useCUDA = false;
    CombinedSolver solver(image, imageColor, imageMask, constraints, params, OptImage::Location::CPU, "backend_cpu", 1);
    std::vector<float2> h_warp((size_t)W * H);
    for (size_t i = 0; i < h_warp.size(); ++i) h_warp[i] = make_float2(warp[2 * i], warp[2 * i + 1]);
    solver.m_warpField->update(h_warp);
    solver.copyResultToCPU();
    Arr out; out.dims = {H, W, 3}; out.elem = 4; out.data.resize((size_t)H * W * 3 * 4);
    float* o = out.as<float>();
    for (int y = 0; y < H; ++y)
        for (int x = 0; x < W; ++x) {
            const vec3f c = solver.m_resultColor(x, y);
            o[(y * W + x) * 3] = c.x; o[(y * W + x) * 3 + 1] = c.y; o[(y * W + x) * 3 + 2] = c.z;
        }
    write_arrays(argv[2], {out});
    return 0;
}
