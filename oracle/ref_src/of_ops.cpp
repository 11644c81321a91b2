// TEST INFRASTRUCTURE (oracle/_ref/of_ops): the reference's OWN optical_flow driver code -- ImageHelper::filterGaussian
// (examples/optical_flow/src/ImageHelper.h:63-114), HierarchyLevel::computeDU / computeDV (CombinedSolver.h:152-178) and
// renderFlowVecotors + ImageHelper::drawLine (main.cpp:7-24, ImageHelper.h:52-66, 118-141) -- compiled from the reference's
// main.cpp where it lies (its main() renamed by a macro; nothing copied) and driven with arrays from a file.
//   of_ops gauss   in.bin out.bin     in: image [H,W] f32, sigma [1] f32                 out: filtered [H,W] f32
//   of_ops prewitt in.bin out.bin     in: image [H,W] f32                                out: du, dv [H,W] f32
//   of_ops lines   in.bin out.bin     in: flow [H,W,2] f32, image [H,W,4] u8             out: image [H,W,4] u8, colours [n,4] u8
#include <algorithm>
#include <fstream>
#include <iostream>
#include <memory>
#include <sstream>
#include <string>
#include <vector>
#include <cstring>
#include "io.h"
#define private public
#define protected public
#define main reference_main
#include "main.cpp"                  // -I <reference>/examples/optical_flow/src
#undef main
#undef private
#undef protected

static Arr f32_image(const ColorImageR32& im) {
    Arr a; a.dims = {(int)im.getHeight(), (int)im.getWidth()}; a.elem = 4; a.data.resize((size_t)im.getWidth() * im.getHeight() * 4);
    float* o = a.as<float>();
    for (unsigned y = 0; y < im.getHeight(); ++y) for (unsigned x = 0; x < im.getWidth(); ++x) o[y * im.getWidth() + x] = im(x, y);
    return a;
}
static ColorImageR32 to_image(Arr& a) {
    const int H = a.dims[0], W = a.dims[1];
    ColorImageR32 im(W, H);
    for (int y = 0; y < H; ++y) for (int x = 0; x < W; ++x) im(x, y) = a.as<float>()[y * W + x];
    return im;
}

int main(int argc, char** argv) {
    if (argc < 4) return 2;
    std::vector<Arr> in = read_arrays(argv[2]);
    if (!strcmp(argv[1], "gauss")) {
        ColorImageR32 im = to_image(in[0]);
        ImageHelper::filterGaussian(im, in[1].as<float>()[0]);
        write_arrays(argv[3], {f32_image(im)});
    } else if (!strcmp(argv[1], "prewitt")) {
        ColorImageR32 im = to_image(in[0]);
        CombinedSolver::HierarchyLevel level;
        write_arrays(argv[3], {f32_image(level.computeDU(im)), f32_image(level.computeDV(im))});
    } else if (!strcmp(argv[1], "lines")) {
        const int H = in[0].dims[0], W = in[0].dims[1];
        BaseImage<float2> flow(W, H);
        ColorImageR8G8B8A8 image(W, H);
        for (int y = 0; y < H; ++y) for (int x = 0; x < W; ++x) {
            flow(x, y) = make_float2(in[0].as<float>()[(y * W + x) * 2], in[0].as<float>()[(y * W + x) * 2 + 1]);
            const unsigned char* c = in[1].as<unsigned char>() + (y * W + x) * 4;
            image(x, y) = vec4uc(c[0], c[1], c[2], c[3]);
        }
        renderFlowVecotors(image, flow);
        Arr out; out.dims = {H, W, 4}; out.elem = 1; out.data.resize((size_t)H * W * 4);
        for (int y = 0; y < H; ++y) for (int x = 0; x < W; ++x) {
            const vec4uc c = image(x, y);
            unsigned char* o = out.as<unsigned char>() + (y * W + x) * 4;
            o[0] = c.x; o[1] = c.y; o[2] = c.z; o[3] = c.w;
        }
        // the colour renderFlowVecotors gives every line (the CUDA entry point takes it as an input): the driver's own
        // expression, evaluated with mLib for the same samples in the same order
        std::vector<unsigned char> cols;
        for (unsigned int j = 1; j < (unsigned)H - 1; j += 5)
            for (unsigned int i = 1; i < (unsigned)W - 1; i += 5) {
                const float2& fv = flow(i, j);
                const float len = vec2f(fv.x, fv.y).length();
                vec4uc color = math::round(255.0f * BaseImageHelper::convertDepthToRGBA(len, 0.0f, 5.0f) * 2.0f); color.w = 255;
                cols.push_back(color.x); cols.push_back(color.y); cols.push_back(color.z); cols.push_back(color.w);
            }
        Arr ca; ca.dims = {(int)cols.size() / 4, 4}; ca.elem = 1; ca.data.assign((char*)cols.data(), (char*)cols.data() + cols.size());
        write_arrays(argv[3], {out, ca});
    } else return 2;
    return 0;
}
