// driver_ops.cu -- CUDA versions of the per-pixel / per-vertex steps that the reference's example drivers run on
// the CPU around the solver path (include/OptB200Drivers.h lists the reference code each one replaces).
//
// All of them are memory-bound image passes.  Arithmetic uses the explicit round-to-nearest intrinsics
// (__fmul_rn / __fadd_rn / __fsub_rn / __fdiv_rn), which the compiler never contracts into FMAs, in the
// reference's order of operations: the outputs are the ones a plain C++ build of the driver code produces.
#include <cstdio>
#include <mutex>
#include <cuda_runtime.h>
#include <cmath>
#include <vector>
#include "../../include/OptB200Drivers.h"

namespace {

__device__ __forceinline__ float mul(float a, float b) { return __fmul_rn(a, b); }
__device__ __forceinline__ float add(float a, float b) { return __fadd_rn(a, b); }
__device__ __forceinline__ float sub(float a, float b) { return __fsub_rn(a, b); }

constexpr int GAUSS_MAX_R = 64;
struct GaussKernel { float w[GAUSS_MAX_R + 1]; };            // passed by value: no constant-memory upload, no host sync

// ---- separable Gaussian (ImageHelper.h:68-114) ---------------------------------------------------------------
// Row pass: a block owns 1024 consecutive pixels of one row (4 per thread, 128-bit stores) and stages them plus R
// pixels either side in shared memory; every pixel of the image is read from DRAM once.
constexpr int ROW_PX = 1024;
__global__ void __launch_bounds__(256) k_gauss_rows(const float* __restrict__ in, float* __restrict__ out, int W, int H, int R, const GaussKernel gk) {
    extern __shared__ float tile[];
    const int y = blockIdx.y, x0 = blockIdx.x * ROW_PX;
    const float* row = in + (size_t)y * W;
    for (int t = threadIdx.x; t < ROW_PX + 2 * R; t += 256) {
        const int x = x0 - R + t;
        tile[t] = (x >= 0 && x < W) ? row[x] : 0.f;
    }
    __syncthreads();
    float res[4];
#pragma unroll
    for (int q = 0; q < 4; ++q) {
        const int lx = threadIdx.x * 4 + q, x = x0 + lx;
        float v = 0.f, w = 0.f;
        for (int k = -R; k <= R; ++k) {
            const int xk = x + k;
            if (xk >= 0 && xk < W) {
                const float g = gk.w[k < 0 ? -k : k];
                v = add(v, mul(g, tile[lx + R + k]));
                w = add(w, g);
            }
        }
        if (w > 0.f) v = __fdiv_rn(v, w);
        res[q] = v;
    }
    const int x = x0 + threadIdx.x * 4;
    float* o = out + (size_t)y * W + x;
    if (x + 3 < W && (((size_t)y * W + x) & 3) == 0) *reinterpret_cast<float4*>(o) = make_float4(res[0], res[1], res[2], res[3]);
    else for (int q = 0; q < 4; ++q) if (x + q < W) o[q] = res[q];
}
// Column pass: a block of 32 x 8 threads owns a 32 x 64 tile (8 output rows per thread) and stages 64 + 2R rows.
constexpr int COL_ROWS = 64;
__global__ void __launch_bounds__(256) k_gauss_cols(const float* __restrict__ in, float* __restrict__ out, int W, int H, int R, const GaussKernel gk) {
    extern __shared__ float tile[];                          // [COL_ROWS + 2R][32]
    const int x = blockIdx.x * 32 + threadIdx.x, y0 = blockIdx.y * COL_ROWS;
    for (int t = threadIdx.y; t < COL_ROWS + 2 * R; t += 8) {
        const int y = y0 - R + t;
        tile[t * 32 + threadIdx.x] = (x < W && y >= 0 && y < H) ? in[(size_t)y * W + x] : 0.f;
    }
    __syncthreads();
    if (x >= W) return;
    for (int j = threadIdx.y; j < COL_ROWS; j += 8) {
        const int y = y0 + j;
        if (y >= H) break;
        float v = 0.f, w = 0.f;
        for (int k = -R; k <= R; ++k) {
            const int yk = y + k;
            if (yk >= 0 && yk < H) {
                const float g = gk.w[k < 0 ? -k : k];
                v = add(v, mul(g, tile[(j + R + k) * 32 + threadIdx.x]));
                w = add(w, g);
            }
        }
        if (w > 0.f) v = __fdiv_rn(v, w);
        out[(size_t)y * W + x] = v;
    }
}

// ---- Prewitt / 8 (optical_flow CombinedSolver.h:152-178) ----------------------------------------------------
template <bool DU>
__global__ void __launch_bounds__(256) k_prewitt(const float* __restrict__ im, float* __restrict__ out, int W, int H) {
    const int x = blockIdx.x * 32 + threadIdx.x, y = blockIdx.y * 8 + threadIdx.y;
    if (x >= W || y >= H) return;
    float d = 0.f;
    if (x >= 1 && y >= 1 && x < W - 1 && y < H - 1) {
        auto I = [&](int i, int j) { return im[(size_t)j * W + i]; };
        if (DU) d = add(add(add(sub(sub(-I(x - 1, y - 1), I(x - 1, y)), I(x - 1, y + 1)), I(x + 1, y - 1)), I(x + 1, y)), I(x + 1, y + 1));
        else    d = add(add(add(sub(sub(-I(x - 1, y - 1), I(x, y - 1)), I(x + 1, y - 1)), I(x - 1, y + 1)), I(x, y + 1)), I(x + 1, y + 1));
        d = __fdiv_rn(d, 8.0f);
    }
    out[(size_t)y * W + x] = d;
}

// ---- image_warping rasteriser --------------------------------------------------------------------------------
// PointInTriangleLK with w = 1 (image_warping CombinedSolver.h:63-99)
__device__ __forceinline__ bool point_in_triangle_lk(float x0, float y0, float x1, float y1, float x2, float y2, float sx, float sy,
                                                     float* wt0, float* wt1, float* wt2) {
    const float X0 = sub(x0, mul(sx, 1.0f)), X1 = sub(x1, mul(sx, 1.0f)), X2 = sub(x2, mul(sx, 1.0f));
    const float Y0 = sub(y0, mul(sy, 1.0f)), Y1 = sub(y1, mul(sy, 1.0f)), Y2 = sub(y2, mul(sy, 1.0f));
    float d01 = sub(mul(X0, Y1), mul(Y0, X1));
    float d12 = sub(mul(X1, Y2), mul(Y1, X2));
    float d20 = sub(mul(X2, Y0), mul(Y2, X0));
    if ((d01 < 0) & (d12 < 0) & (d20 < 0)) return false;   // back-facing
    const float inv = __fdiv_rn(1.f, add(add(d01, d12), d20));
    d01 = mul(d01, inv); d12 = mul(d12, inv); d20 = mul(d20, inv);
    *wt0 = d12; *wt1 = d20; *wt2 = d01;
    return d01 >= 0 && d12 >= 0 && d20 >= 0;
}
// triangle t = 2 * (y*W + x) + k of cell (x, y); false if the cell is not drawn (copyResultToCPU :262-300)
__device__ __forceinline__ bool warp_triangle(const float2* __restrict__ warp, const float* __restrict__ mask, int W, int H, int t,
                                              int* i0, int* i1, int* i2) {
    const int cell = t >> 1, x = cell % W, y = cell / W;
    if (x + 1 >= W || y + 1 >= H) return false;
    const int a = y * W + x, b = a + 1, c = a + W, d = c + 1;   // 00, 01 (x+1), 10 (y+1), 11
    if (!(mask[a] == 0 && mask[b] == 0 && mask[c] == 0 && mask[d] == 0)) return false;
    if (t & 1) { *i0 = c; *i1 = b; *i2 = d; } else { *i0 = a; *i1 = b; *i2 = c; }
    return true;
}
__global__ void k_raster_clear(float* __restrict__ result, int* __restrict__ winner, long n) {
    const long i = blockIdx.x * (long)blockDim.x + threadIdx.x;
    if (i >= n) return;
    result[3 * i] = 255.f; result[3 * i + 1] = 255.f; result[3 * i + 2] = 255.f;
    winner[i] = -1;
}
__global__ void __launch_bounds__(128) k_raster_mark(const float2* __restrict__ warp, const float* __restrict__ mask, int W, int H, int* __restrict__ winner) {
    const int t = blockIdx.x * blockDim.x + threadIdx.x;
    if (t >= 2 * W * H) return;
    int i0, i1, i2;
    if (!warp_triangle(warp, mask, W, H, t, &i0, &i1, &i2)) return;
    const float2 p0 = warp[i0], p1 = warp[i1], p2 = warp[i2];                  // m_scale = 1
    const float lox = floorf(fminf(p0.x, fminf(p1.x, p2.x))), loy = floorf(fminf(p0.y, fminf(p1.y, p2.y)));
    const float hix = ceilf(fmaxf(p0.x, fmaxf(p1.x, p2.x))), hiy = ceilf(fmaxf(p0.y, fmaxf(p1.y, p2.y)));
    if (!(lox <= hix && loy <= hiy)) return;                                   // NaN positions draw nothing
    const int xa = (int)fmaxf(lox, 0.f), ya = (int)fmaxf(loy, 0.f);
    const int xb = (int)fminf(hix, (float)(W - 1)), yb = (int)fminf(hiy, (float)(H - 1));
    for (int y = ya; y <= yb; ++y)
        for (int x = xa; x <= xb; ++x) {
            float b0, b1, b2;
            if (point_in_triangle_lk(p0.x, p0.y, p1.x, p1.y, p2.x, p2.y, (float)x, (float)y, &b0, &b1, &b2))
                atomicMax(&winner[y * W + x], t);
        }
}
__global__ void __launch_bounds__(256) k_raster_shade(const float2* __restrict__ warp, const float* __restrict__ color, const float* __restrict__ mask,
                                                      int W, int H, const int* __restrict__ winner, float* __restrict__ result) {
    const int x = blockIdx.x * 32 + threadIdx.x, y = blockIdx.y * 8 + threadIdx.y;
    if (x >= W || y >= H) return;
    const int t = winner[y * W + x];
    if (t < 0) return;
    int i0, i1, i2;
    warp_triangle(warp, mask, W, H, t, &i0, &i1, &i2);
    const float2 p0 = warp[i0], p1 = warp[i1], p2 = warp[i2];
    float b0, b1, b2;
    point_in_triangle_lk(p0.x, p0.y, p1.x, p1.y, p2.x, p2.y, (float)x, (float)y, &b0, &b1, &b2);
#pragma unroll
    for (int c = 0; c < 3; ++c)     // vec3f color = c0*b0 + c1*b1 + c2*b2
        result[3 * (y * W + x) + c] = add(add(mul(color[3 * i0 + c], b0), mul(color[3 * i1 + c], b1)), mul(color[3 * i2 + c], b2));
}

// ---- volumetric trilinear skinning (volumetric CombinedSolver.h:302-335) ------------------------------------
__global__ void __launch_bounds__(256) k_skin(const float* __restrict__ grid, int ny1, int nz1, const int* __restrict__ voxel,
                                              const float* __restrict__ rel, long n, float* __restrict__ out) {
    const long i = blockIdx.x * (long)blockDim.x + threadIdx.x;
    if (i >= n) return;
    const int vx = voxel[3 * i], vy = voxel[3 * i + 1], vz = voxel[3 * i + 2];
    const float rx = rel[3 * i], ry = rel[3 * i + 1], rz = rel[3 * i + 2];
    const float ox = sub(1.0f, rx), oy = sub(1.0f, ry), oz = sub(1.0f, rz);
    auto node = [&](int dx, int dy, int dz, int c) {
        return grid[3 * ((size_t)(vx + dx) * ny1 * nz1 + (size_t)(vy + dy) * nz1 + (vz + dz)) + c];
    };
#pragma unroll
    for (int c = 0; c < 3; ++c) {
        const float px00 = add(mul(ox, node(0, 0, 0, c)), mul(rx, node(1, 0, 0, c)));
        const float px01 = add(mul(ox, node(0, 0, 1, c)), mul(rx, node(1, 0, 1, c)));
        const float px10 = add(mul(ox, node(0, 1, 0, c)), mul(rx, node(1, 1, 0, c)));
        const float px11 = add(mul(ox, node(0, 1, 1, c)), mul(rx, node(1, 1, 1, c)));
        const float pxx0 = add(mul(oy, px00), mul(ry, px10));
        const float pxx1 = add(mul(oy, px01), mul(ry, px11));
        out[3 * i + c] = add(mul(oz, pxx0), mul(rz, pxx1));
    }
}

// ---- flow-vector overlay (optical_flow main.cpp:7-24, ImageHelper.h:52-66, 118-141) -------------------------
// Walks the driver's drawLine / drawLineInternal for sample `s` and calls put(x, y) for every pixel it writes.
template <class Put>
__device__ __forceinline__ void walk_flow_line(const float2* __restrict__ flow, int W, int H, int skip, int s, int nsx, Put put) {
    const int i = 1 + (s % nsx) * skip, j = 1 + (s / nsx) * skip;
    const float2 f = flow[(size_t)j * W + i];
    // math::round: half away from zero (mLib core-util/utility.h:130-133)
    auto rnd = [](float v) { return v > 0.0f ? (int)floorf(add(v, 0.5f)) : (int)ceilf(sub(v, 0.5f)); };
    const int ex = i + rnd(f.x), ey = j + rnd(f.y);
    auto clampi = [](int v, int lo, int hi) { return v < lo ? lo : (v > hi ? hi : v); };
    int sx = clampi(i, 0, W - 1), sy = clampi(j, 0, H - 1), tx = clampi(ex, 0, W - 1), ty = clampi(ey, 0, H - 1);
    if (!(ex >= i)) { int a = sx; sx = tx; tx = a; a = sy; sy = ty; ty = a; }
    const float dx = (float)(tx - sx);
    if (dx == 0.0f) { for (int y = sy; y < ty; ++y) put(sx, y); return; }
    const float dy = (float)(ty - sy);
    float error = 0.0f;
    const float dErr = fabsf(__fdiv_rn(dy, dx));
    const int dir = (ty - sy) > 0 ? 1 : ((ty - sy) < 0 ? -1 : 0);
    int y = sy;
    for (int x = sx; x < tx; ++x) {
        put(x, y);
        error = add(error, dErr);
        while (error >= 0.5f) { put(x, y); y += dir; error = sub(error, 1.0f); }
    }
}
__global__ void k_flow_mark(const float2* __restrict__ flow, int W, int H, int skip, int nsx, int ns, int* __restrict__ winner) {
    const int s = blockIdx.x * blockDim.x + threadIdx.x;
    if (s >= ns) return;
    walk_flow_line(flow, W, H, skip, s, nsx, [&](int x, int y) { if (x >= 0 && x < W && y >= 0 && y < H) atomicMax(&winner[y * W + x], s); });
}
__global__ void k_fill_int(int* p, long n, int v) {
    const long i = blockIdx.x * (long)blockDim.x + threadIdx.x;
    if (i < n) p[i] = v;
}
__global__ void k_flow_paint(const int* __restrict__ winner, const uchar4* __restrict__ colors, long n, uchar4* __restrict__ image) {
    const long i = blockIdx.x * (long)blockDim.x + threadIdx.x;
    if (i >= n) return;
    const int s = winner[i];
    if (s >= 0) image[i] = colors[s];
}

inline int done() { return (int)cudaGetLastError(); }

}  // namespace

extern "C" {

int OptB200_FilterGaussian(float* image, unsigned W, unsigned H, float sigma, float* scratch, void* stream_) {
    cudaStream_t st = (cudaStream_t)stream_;
    const int R = (int)std::ceil(2.0f * sigma);                // math::ceil(2.0f*sigma)
    if (R < 0 || R > GAUSS_MAX_R || W == 0 || H == 0) return (int)cudaErrorInvalidValue;
    GaussKernel gk;
    for (int i = 0; i <= GAUSS_MAX_R; ++i) { const float x = (float)i; gk.w[i] = i <= R ? expf(-((x * x) / (2.0f * sigma * sigma))) : 0.f; }
    k_gauss_rows<<<dim3((W + ROW_PX - 1) / ROW_PX, H), 256, (ROW_PX + 2 * R) * sizeof(float), st>>>(image, scratch, (int)W, (int)H, R, gk);
    k_gauss_cols<<<dim3((W + 31) / 32, (H + COL_ROWS - 1) / COL_ROWS), dim3(32, 8), (COL_ROWS + 2 * R) * 32 * sizeof(float), st>>>(scratch, image, (int)W, (int)H, R, gk);
    return done();
}

// ---- OptB200_FilterGaussianNoAlloc: the same filter for callers that can neither allocate nor hand over a scratch image
// (CUDA graph capture, serving loops, harnesses that forbid cudaMalloc).  No counterpart in the reference.
// The row pass's intermediate image lives in a file-scope device array of a fixed maximum
// size (OPTB200_GAUSS_NOALLOC_MAX_PIXELS floats = 16 MiB, held by every GPU that loads the library until the process exits).
// In HOST code the array's name is only a stand-in for the symbol, not a device address: its address is taken once per
// device, with that device current, by cudaGetSymbolAddress, and that pointer is what the kernels get.  The row pass writes
// every element of the scratch image that the column pass reads, so nothing has to be cleared between calls; all calls share
// the array, so two calls that overlap on the GPU (different streams, or graphs replayed concurrently) race.
static __device__ __align__(256) float g_gauss_scratch[OPTB200_GAUSS_NOALLOC_MAX_PIXELS];
static float* g_gauss_scratch_ptr[64];                         // per device ordinal; nullptr = address not taken yet
static std::mutex g_gauss_scratch_mutex;

int OptB200_FilterGaussianNoAlloc(float* image, unsigned W, unsigned H, float sigma, void* stream) {
    if (W == 0 || H == 0) return (int)cudaErrorInvalidValue;
    if ((unsigned long long)W * H > (unsigned long long)OPTB200_GAUSS_NOALLOC_MAX_PIXELS) {
        fprintf(stderr, "OptB200_FilterGaussianNoAlloc: image of %u x %u pixels exceeds the static scratch of %ld pixels "
                        "(use OptB200_FilterGaussian with a caller-owned scratch image)\n", W, H, (long)OPTB200_GAUSS_NOALLOC_MAX_PIXELS);
        return (int)cudaErrorInvalidValue;
    }
    int dev = 0;
    cudaError_t e = cudaGetDevice(&dev);
    if (e != cudaSuccess) return (int)e;
    if (dev < 0 || dev >= 64) return (int)cudaErrorInvalidDevice;
    float* scratch = nullptr;
    {
        std::lock_guard<std::mutex> lock(g_gauss_scratch_mutex);
        if (!g_gauss_scratch_ptr[dev]) {                       // set-up, once per device, with that device current
            void* p = nullptr;
            e = cudaGetSymbolAddress(&p, g_gauss_scratch);
            if (e != cudaSuccess) {
                fprintf(stderr, "OptB200_FilterGaussianNoAlloc: cudaGetSymbolAddress failed on device %d: %s\n", dev, cudaGetErrorString(e));
                return (int)e;
            }
            g_gauss_scratch_ptr[dev] = (float*)p;
        }
        scratch = g_gauss_scratch_ptr[dev];
    }
    return OptB200_FilterGaussian(image, W, H, sigma, scratch, stream);     // two kernel launches, no allocation
}

int OptB200_PrewittDU(const float* image, float* out, unsigned W, unsigned H, void* stream) {
    k_prewitt<true><<<dim3((W + 31) / 32, (H + 7) / 8), dim3(32, 8), 0, (cudaStream_t)stream>>>(image, out, (int)W, (int)H);
    return done();
}
int OptB200_PrewittDV(const float* image, float* out, unsigned W, unsigned H, void* stream) {
    k_prewitt<false><<<dim3((W + 31) / 32, (H + 7) / 8), dim3(32, 8), 0, (cudaStream_t)stream>>>(image, out, (int)W, (int)H);
    return done();
}

int OptB200_RasterizeWarp(const float* warp, const float* color, const float* mask, unsigned W, unsigned H,
                          float* result, int* winner, void* stream) {
    cudaStream_t st = (cudaStream_t)stream;
    const long n = (long)W * H;
    if (n == 0 || 2 * n > 0x7fffffffL) return (int)cudaErrorInvalidValue;
    k_raster_clear<<<(unsigned)((n + 255) / 256), 256, 0, st>>>(result, winner, n);
    k_raster_mark<<<(unsigned)((2 * n + 127) / 128), 128, 0, st>>>((const float2*)warp, mask, (int)W, (int)H, winner);
    k_raster_shade<<<dim3((W + 31) / 32, (H + 7) / 8), dim3(32, 8), 0, st>>>((const float2*)warp, color, mask, (int)W, (int)H, winner, result);
    return done();
}

int OptB200_SkinTrilinear(const float* grid, const int* dims, const int* voxel, const float* rel, long n, float* out, void* stream) {
    if (n <= 0) return 0;
    k_skin<<<(unsigned)((n + 255) / 256), 256, 0, (cudaStream_t)stream>>>(grid, dims[1] + 1, dims[2] + 1, voxel, rel, n, out);
    return done();
}

int OptB200_RenderFlowLines(unsigned char* image, const float* flow, const unsigned char* colors, unsigned W, unsigned H,
                            unsigned skip, int* winner, void* stream) {
    cudaStream_t st = (cudaStream_t)stream;
    if (W < 3 || H < 3 || skip == 0) return (int)cudaErrorInvalidValue;
    const int nsx = (int)((W - 2 + skip - 1) / skip), nsy = (int)((H - 2 + skip - 1) / skip), ns = nsx * nsy;   // i = 1, 1+skip, ... < W-1
    const long n = (long)W * H;
    k_fill_int<<<(unsigned)((n + 255) / 256), 256, 0, st>>>(winner, n, -1);
    k_flow_mark<<<(unsigned)((ns + 127) / 128), 128, 0, st>>>((const float2*)flow, (int)W, (int)H, (int)skip, nsx, ns, winner);
    k_flow_paint<<<(unsigned)((n + 255) / 256), 256, 0, st>>>(winner, (const uchar4*)colors, n, (uchar4*)image);
    return done();
}

}  // extern "C"
