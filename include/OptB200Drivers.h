/* OptB200Drivers.h -- the per-pixel / per-vertex steps the reference's example DRIVERS run on the CPU either side
 * of the solver path, as CUDA kernels behind a C ABI so that a driver can keep its
 * images on the device between pre-processing, Opt_ProblemSolve and post-processing.
 *
 * Every pointer is a DEVICE pointer unless stated otherwise; `stream` is a cudaStream_t passed as void* (NULL =
 * the legacy default stream, which is where Opt_Problem* calls are ordered, Opt.h).  All functions return 0 on
 * success and a cudaError_t value otherwise.  Arithmetic is IEEE single precision with separate multiply and add
 * (no contraction), in the reference's order of operations, so results match a plain C++ build of the reference
 * code bit for bit (the Gaussian up to the last-ulp difference between expf implementations).
 *
 *   entry point                 replaces (reference file:line)
 *   OptB200_FilterGaussian      examples/optical_flow/src/ImageHelper.h:68-114  (ImageHelper::filterGaussian)
 *   OptB200_PrewittDU / DV      examples/optical_flow/src/CombinedSolver.h:152-178 (computeDU / computeDV)
 *   OptB200_RasterizeWarp       examples/image_warping/src/CombinedSolver.h:63-99, 231-308 (PointInTriangleLK,
 *                               rasterizeTriangle, copyResultToCPU)
 *   OptB200_SkinTrilinear       examples/volumetric_mesh_deformation/src/CombinedSolver.h:302-335
 *                               (copyResultToCPUFromFloat3)
 *   OptB200_RenderFlowLines     examples/optical_flow/src/main.cpp:7-24 + ImageHelper.h:52-66, 118-141
 *                               (renderFlowVecotors / drawLine), line colour supplied by the caller
 */
#ifndef OPT_B200_DRIVERS_H
#define OPT_B200_DRIVERS_H
#ifdef __cplusplus
extern "C" {
#endif

/* In-place separable Gaussian of a W x H float image (x fastest): kernel radius ceil(2 sigma), weights
 * exp(-x^2 / (2 sigma^2)), rows first then columns, each output divided by the sum of the weights that fell inside
 * the image.  `scratch` holds W*H floats. */
int OptB200_FilterGaussian(float* image, unsigned W, unsigned H, float sigma, float* scratch, void* stream);

/* The same filter without any allocation and without a caller-owned scratch image, for callers that cannot call cudaMalloc
 * (code under CUDA graph capture, tight serving loops, harnesses that forbid it).  No counterpart in the reference.
 * The intermediate image of the row pass lives in a static device array inside the
 * library, OPTB200_GAUSS_NOALLOC_MAX_PIXELS floats (16 MiB per GPU that loads the library, held until the process exits);
 * W * H beyond that is rejected with cudaErrorInvalidValue and a message on stderr.  Calls no cudaMalloc / cudaFree and no
 * allocating library routine; same kernels and bit-identical output as OptB200_FilterGaussian.  All calls on a device share
 * the array (nothing in it needs clearing between calls: the row pass rewrites all that the column pass reads), so calls
 * that overlap on the GPU -- different streams, concurrently replayed graphs -- race: serialise them. */
#define OPTB200_GAUSS_NOALLOC_MAX_PIXELS (2048L * 2048L)
int OptB200_FilterGaussianNoAlloc(float* image, unsigned W, unsigned H, float sigma, void* stream);

/* 3x3 Prewitt derivative / 8 along x (DU) or y (DV); the one-pixel border of the output is 0. */
int OptB200_PrewittDU(const float* image, float* out, unsigned W, unsigned H, void* stream);
int OptB200_PrewittDV(const float* image, float* out, unsigned W, unsigned H, void* stream);

/* Rasterises the warped quad mesh of image_warping: pixel (x, y) of the W x H source grid moves to
 * warp[y*W + x] (float2); every grid cell whose four corners are unmasked (mask == 0) is drawn as the two
 * triangles (00,01,10) and (10,01,11) with barycentrically interpolated float3 vertex colours into the W x H
 * float3 image `result`, which is first cleared to (255,255,255).  The reference draws the triangles one after
 * another and the last writer of a pixel wins; here every triangle marks the pixels it covers with an atomicMax
 * of its sequence number and a second pass shades each pixel from its winner: the same image, deterministically.
 * `winner` is scratch for W*H ints. */
int OptB200_RasterizeWarp(const float* warp, const float* color, const float* mask, unsigned W, unsigned H,
                          float* result, int* winner, void* stream);

/* Trilinear skinning of n mesh vertices by the deformed control lattice: grid holds (dims[0]+1) x (dims[1]+1) x
 * (dims[2]+1) float3 node positions with index x*(dims[1]+1)*(dims[2]+1) + y*(dims[2]+1) + z (the driver's
 * getIndex1D, CombinedSolver.h:273-276); voxel[i] (int3) is the cell of vertex i and rel[i] (float3) its
 * coordinates inside the cell; out[i] (float3) is the skinned position. */
int OptB200_SkinTrilinear(const float* grid, const int* dims, const int* voxel, const float* rel, long n,
                          float* out, void* stream);

/* Flow-vector overlay: for every `skip`-th interior pixel (start at 1, as the driver does) a line from the pixel
 * to pixel + round(flow) is drawn into the W x H RGBA8 image with the colour colors[(j/skip... )] given per
 * line in `colors` (RGBA8, one per sampled pixel in row-major order of the samples).  Lines are drawn in the
 * driver's order (later lines overwrite earlier ones); implemented with the same two-pass winner scheme as the
 * rasteriser.  `winner` is scratch for W*H ints. */
int OptB200_RenderFlowLines(unsigned char* image, const float* flow, const unsigned char* colors, unsigned W,
                            unsigned H, unsigned skip, int* winner, void* stream);

#ifdef __cplusplus
}
#endif
#endif
