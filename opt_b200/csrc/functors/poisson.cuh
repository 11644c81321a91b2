// functors/poisson.cuh -- poisson_image_editing (examples/poisson_image_editing/poisson_image_editing.t:1-13)
//   r_o(i) = inb(i+o) * [ (X_i - X_{i+o}) - (T_i - T_{i+o}) ],  o in {(1,0),(-1,0),(0,1),(0,-1)}, 4 channels
//   Exclude(M != 0), UsePreconditioner(false).
// Closed forms (both the instance centred at i and the one centred at i+o touch X_i):
//   J^T F(i) = 2 sum_o inb(n) e_o(i)      diag = 2 #inb      J^T J p (i) = 2 sum_o inb(n) (p_i - p_n)
#pragma once
#include "math.cuh"

namespace optk {

template <class T>
struct PoissonFunctor : FunctorDefaults<T> {
    using Scalar = T;
    static constexpr int U = 4, ND = 2;
    static constexpr bool HAS_EXCLUDE = true;
    static constexpr bool USE_PRE = false;
    int W, H, D;
    T* X; const T* Tm; const T* M;
    void plan(const unsigned* d) {
        W = d[0]; H = d[1]; D = 1;
        tiled = getenv("OPTB200_POISSON_TILED") != nullptr;
        // (channel, strip, band) items are dealt round-robin to the resident warps (8 per CTA, 3 CTAs per SM in fp32,
        // 2 in fp64); items without unknowns are dropped from the list (build_active_items), so the others cost about
        // the same and ~8 of them per warp keep the quantisation loss small.  A band re-reads 2 neighbour rows.
        int dev = 0, nsm = 148;
        cudaGetDevice(&dev); cudaDeviceGetAttribute(&nsm, cudaDevAttrMultiProcessorCount, dev);
        const long slots = (long)nsm * 8 * (sizeof(T) == 4 ? 3 : 2) * 8;
        bandR = (int)std::min<long>(256, std::max<long>(16, ((long)H * 4 * strips() + slots - 1) / slots));
    }
    void bind(void** pp) { X = (T*)pp[0]; Tm = (const T*)pp[1]; M = (const T*)pp[2]; }
    __device__ bool exclude(int, int, int, long s) const { return M[s] != T(0); }
    __device__ T* xptr(long s, int u) const { return X + s * 4 + u; }
    int unknown_images(void** ptrs, int* ch) const { ptrs[0] = X; ch[0] = 4; return 1; }
    // vectorised / fused J^T J apply; needs rows that are a multiple of 4 pixels.  Two kernels: k_poisson_march (a
    // warp marches one channel plane down a band of rows, rows prefetched by cp.async; the default) and the older
    // tile kernel k_poisson_apply (OPTB200_POISSON_TILED=1 when the plan is made; tests compare the two).
    static constexpr bool HAS_FAST_APPLY = true;
    static constexpr bool HAS_ACTIVE_ITEMS = true;
    static constexpr bool FUSE_SLAB = true;          // k_poisson_march takes ghost rows from the exchanged p' (engine.cuh k_boundary_dir)
    bool fuse_slab_ok() const { return fast_ok() && !tiled; }
    bool tiled = false;
    int bandR = 16;
    unsigned* items = nullptr;       // (band * strips + strip) of every item that holds an unknown, compacted
    unsigned* nactive = nullptr;     // their number
    void alloc_temporaries(cudaStream_t) {
        OPTK_CUDA(cudaMalloc(&items, sizeof(unsigned) * ((size_t)strips() * ((H + bandR - 1) / bandR) + 1)));
        OPTK_CUDA(cudaMalloc(&nactive, sizeof(unsigned)));
    }
    void free_temporaries() { cudaFree(items); cudaFree(nactive); items = nullptr; nactive = nullptr; }
    void build_active_items(cudaStream_t stream, const unsigned char* active4) const;
    bool fast_ok() const { return W % 4 == 0; }
    __host__ __device__ int own_lo() const { return this->o_lo > 0 ? this->o_lo : 0; }
    __host__ __device__ int own_hi() const { return this->o_hi < H ? this->o_hi : H; }
    __host__ __device__ int strips() const { return (W / 4 + 31) / 32; }
    __host__ __device__ int bands() const { return (own_hi() - own_lo() + bandR - 1) / bandR; }
    long fast_tiles() const {
        if (tiled) return (long)((W / 4 + 31) / 32) * ((H + 7) / 8);
        return ((long)4 * strips() * bands() + 7) / 8;
    }
    dim3 fast_block() const { return dim3(32, 8, 1); }
    size_t fast_smem() const;
    template <bool FUSED, bool LM> auto fast_kernel() const;

    template <class Fn> __device__ __forceinline__ void neighbours(int x, int y, long s, Fn&& fn) const {
        if (x + 1 < W) fn(s + 1);
        if (x > 0) fn(s - 1);
        if (y + 1 < H) fn(s + W);
        if (y > 0) fn(s - W);
    }
    __device__ T cost(int x, int y, int, long s) const {
        T xi[4], ti[4]; load4(X + s * 4, xi); load4(Tm + s * 4, ti);
        T sum = T(0);
        neighbours(x, y, s, [&](long n) {
            T xn[4], tn[4]; load4(X + n * 4, xn); load4(Tm + n * 4, tn);
#pragma unroll
            for (int c = 0; c < 4; ++c) { T e = (xi[c] - xn[c]) - (ti[c] - tn[c]); sum += e * e; }
        });
        return sum;
    }
    __device__ void evalJTF(int x, int y, int, long s, T* g, T* d) const {
        T xi[4], ti[4]; load4(X + s * 4, xi); load4(Tm + s * 4, ti);
        T cnt = T(0);
        neighbours(x, y, s, [&](long n) {
            T xn[4], tn[4]; load4(X + n * 4, xn); load4(Tm + n * 4, tn);
#pragma unroll
            for (int c = 0; c < 4; ++c) g[c] += (xi[c] - xn[c]) - (ti[c] - tn[c]);
            cnt += T(1);
        });
#pragma unroll
        for (int c = 0; c < 4; ++c) { g[c] = T(2) * g[c]; d[c] = T(2) * cnt; }
    }
    // Jacobian rows (materialized-J mode, engine.cuh k_save_j): 4 offsets x 4 channels; residual (o, c) at site s has
    // +1 at X(s)[c] and -1 at X(n)[c] when the neighbour is in bounds, and is identically 0 otherwise (Select(InBounds, e, 0)).
    static constexpr bool HAS_ROWS_J = true;
    static constexpr int MAT_ROWS = 16;
    template <class Sink> __device__ void rowsJ(int x, int y, int, long s, Sink& sink) const {
        const bool inb[4] = {x + 1 < W, x > 0, y + 1 < H, y > 0};
        const long nb[4] = {s + 1, s - 1, s + W, s - W};
#pragma unroll
        for (int o = 0; o < 4; ++o)
#pragma unroll
            for (int c = 0; c < 4; ++c) {
                sink.begin(o * 4 + c);
                if (inb[o]) { sink.add(s, c, T(1)); sink.add(nb[o], c, T(-1)); }
            }
    }
    template <class P> __device__ void applyJTJ(int x, int y, int, long s, const P& p, T* out) const {
        T pi[4];
#pragma unroll
        for (int c = 0; c < 4; ++c) { pi[c] = p(c, s); out[c] = T(0); }
        neighbours(x, y, s, [&](long n) {
#pragma unroll
            for (int c = 0; c < 4; ++c) out[c] += pi[c] - p(c, n);
        });
#pragma unroll
        for (int c = 0; c < 4; ++c) out[c] = T(2) * out[c];
    }
    __device__ T modelCost(int x, int y, int, long s, const Vec<T>& dl) const {
        T xi[4], ti[4], di[4]; load4(X + s * 4, xi); load4(Tm + s * 4, ti);
#pragma unroll
        for (int c = 0; c < 4; ++c) di[c] = dl(c, s);
        T sum = T(0);
        neighbours(x, y, s, [&](long n) {
            T xn[4], tn[4]; load4(X + n * 4, xn); load4(Tm + n * 4, tn);
#pragma unroll
            for (int c = 0; c < 4; ++c) { T m = ((xi[c] - xn[c]) - (ti[c] - tn[c])) + (di[c] - dl(c, n)); sum += m * m; }
        });
        return sum;
    }
};


// ---------------------------------------------------------------------------------------------
// Vectorised J^T J apply for poisson (optionally fused with PCGStep3, see engine.cuh).
// J^T J is four independent 5-point Laplacians (one per channel plane), so each thread owns FOUR
// consecutive pixels of a row: per plane it issues 128-bit loads for the centre / upper / lower
// quads and gets the left/right neighbours from the adjacent lanes by warp shuffle (only lanes 0 and
// 31 touch memory for them).  ~2 LDG.128 per pixel instead of 25 scalar loads: the generic kernel
// was LSU-issue bound at 2.3 TB/s (profiles/r1_a).  Quads without unknowns are skipped via active4.
template <class T> __device__ __forceinline__ void ld4v(const T* p, T* o) { load4(p, o); }

template <class T, bool FUSED, bool LM>
__global__ void __launch_bounds__(256) k_poisson_apply(PoissonFunctor<T> f, Vec<T> z, Vec<T> p, Vec<T> p_new, Vec<T> out,
                                                       Vec<T> CtC, const unsigned char* __restrict__ active4,
                                                       const double* bNum, const double* aNum, double* den,
                                                       double* zero_a, double* zero_b) {
    const int W = f.W, H = f.H;
    const int tiles_x = (W / 4 + 31) / 32;
    const long ntiles = (long)tiles_x * ((H + 7) / 8);
    const int lane = threadIdx.x;
    T beta = T(0);
    if (FUSED) beta = (T)(*bNum) / (T)(*aNum);
    if (blockIdx.x == 0 && linear_tid() == 0) {
        if (zero_a) *zero_a = 0.0;
        if (zero_b) *zero_b = 0.0;
    }
    double acc = 0.0;
    for (long tile = blockIdx.x; tile < ntiles; tile += gridDim.x) {      // persistent CTAs over 128x8-pixel tiles
    const int gx = ((int)(tile % tiles_x) * 32 + threadIdx.x) * 4;
    const int y = (int)(tile / tiles_x) * 8 + threadIdx.y;
    const bool in_img = gx < W && y < H;
    const bool own = in_img && y >= f.o_lo && y < f.o_hi;
    const long s = (long)y * W + gx;
    // a quad takes part in loads if it is inside the image and holds at least one unknown
    const bool act = in_img && active4[s >> 2] != 0;
    bool ex[4] = {true, true, true, true};
    if (act && own) {
        T m[4]; ld4v(f.M + s, m);
#pragma unroll
        for (int j = 0; j < 4; ++j) ex[j] = m[j] != T(0);
    }
    const bool up = y > 0, dn = y + 1 < H;
#pragma unroll
    for (int c = 0; c < 4; ++c) {
        const T* pc = p.d + (long)c * p.plane;
        const T* zc = z.d + (long)c * z.plane;
        auto dirq = [&](long at, T* o) {                 // direction at a quad
            T a[4]; ld4v(pc + at, a);
            if (FUSED) { T b[4]; ld4v(zc + at, b);
#pragma unroll
                for (int j = 0; j < 4; ++j) o[j] = b[j] + beta * a[j];
            } else {
#pragma unroll
                for (int j = 0; j < 4; ++j) o[j] = a[j];
            }
        };
        auto dir1 = [&](long at) -> T { return FUSED ? zc[at] + beta * pc[at] : pc[at]; };
        // all loads of the channel are issued before the shuffles consume the centre quad: one exposed memory
        // latency per channel instead of two (the kernel is latency-bound, ncu r1i)
        T ctr[4] = {T(0), T(0), T(0), T(0)}, u4[4] = {T(0), T(0), T(0), T(0)}, d4[4] = {T(0), T(0), T(0), T(0)};
        T lmem = T(0), rmem = T(0);
        if (act) dirq(s, ctr);
        if (act && own) {
            if (up) dirq(s - W, u4);
            if (dn) dirq(s + W, d4);
            if (lane == 0 && gx > 0) lmem = dir1(s - 1);
            if (lane == 31 && gx + 4 < W) rmem = dir1(s + 4);
        }
        // lateral neighbours: from the adjacent lanes, from memory at the warp's ends
        T lft = __shfl_up_sync(0xffffffffu, ctr[3], 1);
        T rgt = __shfl_down_sync(0xffffffffu, ctr[0], 1);
        if (act && own) {
            if (lane == 0) lft = lmem;
            if (lane == 31) rgt = rmem;
            T o[4], pn[4];
#pragma unroll
            for (int j = 0; j < 4; ++j) {
                const T v = ctr[j];
                T a = T(0);
                if (gx + j + 1 < W) a += v - (j < 3 ? ctr[j + 1] : rgt);      // {1,0}
                if (gx + j > 0) a += v - (j > 0 ? ctr[j - 1] : lft);          // {-1,0}
                if (dn) a += v - d4[j];                                       // {0,1}
                if (up) a += v - u4[j];                                       // {0,-1}
                a = T(2) * a;
                pn[j] = ex[j] ? T(0) : v;
                o[j] = ex[j] ? T(0) : a;
            }
            if (LM) {
                T cc[4]; ld4v(CtC.d + (long)c * CtC.plane + s, cc);
#pragma unroll
                for (int j = 0; j < 4; ++j) if (!ex[j]) o[j] += cc[j] * pn[j];
            }
#pragma unroll
            for (int j = 0; j < 4; ++j) acc += (double)(pn[j] * o[j]);
            if (FUSED) st4v(p_new.d + (long)c * p_new.plane + s, pn);
            st4v(out.d + (long)c * out.plane + s, o);
        }
    }
    }   // tile loop
    block_accumulate(acc, den);
}

// ---------------------------------------------------------------------------------------------
// Marching J^T J apply (optionally fused with PCGStep3).  The tile kernel above reads every row of p (and z) three
// times (as centre, upper and lower neighbour: 6 LDG.128 per quad and channel, 4 of them L1/L2 hits) and is bound by
// memory latency at 4.0 TB/s (profiles/r1i).  Here the four channel planes are separate work items -- they are
// independent 5-point Laplacians -- and a warp owns (plane, strip of 128 pixels, band of bandR rows): lane = 4
// consecutive pixels, the warp marches down the band holding rows n-2, n-1, n of the search direction in registers,
// so every quad is fetched once (plus 2 / bandR for the band's first and last neighbour rows).  Rows are prefetched
// D-1 steps ahead by per-thread cp.async (zero-filled for quads without unknowns and for rows outside the image)
// into a shared-memory ring that only the issuing thread reads back: no barriers, nothing in flight holds registers.
// Lateral neighbours come from the adjacent lanes by shuffle; lanes 0 and 31 fetch the pixel beyond the strip.
// The per-pixel mask M is not read at all: active4 holds one bit per pixel of a quad (k_active4, engine.cuh).
template <class T, bool FUSED, bool LM>
struct PoissonMarch {
    static constexpr int QB = 4 * (int)sizeof(T);
    static constexpr int NA = 1 + (FUSED ? 1 : 0) + (LM ? 1 : 0);
    static constexpr int STAGE = NA * 256 * QB + 2 * 256 * (int)sizeof(T);
    static constexpr int D = sizeof(T) == 4 ? 4 : 3;
    static constexpr int A_P = 0, A_Z = 1, A_C = 1 + (FUSED ? 1 : 0);
    static __device__ __forceinline__ unsigned char* slot(unsigned char* ring, int st, int arr, int tid) {
        return ring + (size_t)st * STAGE + (size_t)arr * 256 * QB + (size_t)tid * QB;
    }
    static __device__ __forceinline__ unsigned char* eslot(unsigned char* ring, int st, int arr, int tid) {
        return ring + (size_t)st * STAGE + (size_t)NA * 256 * QB + (size_t)arr * 256 * sizeof(T) + (size_t)tid * sizeof(T);
    }
    static __device__ __forceinline__ void cp16(void* dst, const void* src, bool on) {
        const unsigned d = (unsigned)__cvta_generic_to_shared(dst);
        const int n = on ? 16 : 0;
        asm volatile("cp.async.cg.shared.global [%0], [%1], 16, %2;" ::"r"(d), "l"(src), "r"(n) : "memory");
    }
    static __device__ __forceinline__ void cp1(void* dst, const T* src, bool on) {          // one scalar
        const unsigned d = (unsigned)__cvta_generic_to_shared(dst);
        if constexpr (sizeof(T) == 4) { const int n = on ? 4 : 0; asm volatile("cp.async.ca.shared.global [%0], [%1], 4, %2;" ::"r"(d), "l"(src), "r"(n) : "memory"); }
        else                          { const int n = on ? 8 : 0; asm volatile("cp.async.ca.shared.global [%0], [%1], 8, %2;" ::"r"(d), "l"(src), "r"(n) : "memory"); }
    }
    static __device__ __forceinline__ void cpq(unsigned char* dst, const T* src, bool on) {
#pragma unroll
        for (int k = 0; k < QB / 16; ++k) cp16(dst + 16 * k, (const unsigned char*)src + 16 * k, on);
    }
};

template <class T, bool FUSED, bool LM>
__global__ void __launch_bounds__(256, sizeof(T) == 4 ? 3 : 2)
k_poisson_march(PoissonFunctor<T> f, Vec<T> z, Vec<T> p, Vec<T> p_new, Vec<T> out, Vec<T> CtC,
                const unsigned char* __restrict__ active4, const double* bNum, const double* aNum, double* den,
                double* zero_a, double* zero_b) {
    using PM = PoissonMarch<T, FUSED, LM>;
    constexpr int D = PM::D;
    extern __shared__ __align__(16) unsigned char pm_ring[];
    const int W = f.W, H = f.H, qpr = W / 4;
    const int tid = linear_tid(), lane = tid & 31;
    const T beta = FUSED ? (T)(*bNum) / (T)(*aNum) : T(0);
    if (blockIdx.x == 0 && tid == 0) {
        if (zero_a) *zero_a = 0.0;
        if (zero_b) *zero_b = 0.0;
    }
    const int strips = f.strips(), bandR = f.bandR, row_lo = f.own_lo(), row_hi = f.own_hi();
    const long nitems = (long)4 * (*f.nactive);
    double acc = 0.0;
    for (long item = blockIdx.x * 8L + (tid >> 5); item < nitems; item += (long)gridDim.x * 8) {     // warp-uniform
        const int c = (int)(item & 3);
        const unsigned sb = f.items[item >> 2];
        const int strip = (int)(sb % strips), band = (int)(sb / strips);
        const int q = strip * 32 + lane, gx = q * 4;
        const bool col_in = q < qpr;
        const int y0 = row_lo + band * bandR, y1 = (y0 + bandR < row_hi) ? y0 + bandR : row_hi;
        const T* pc = p.d + (long)c * p.plane;
        const T* zc = z.d + (long)c * z.plane;
        const T* nc = p_new.d + (long)c * p_new.plane;                       // FUSED, slab mode: ghost rows of p' (exchanged)
        const T* cc = CtC.d + (long)c * CtC.plane;
        T* pnew_c = p_new.d + (long)c * p_new.plane;
        T* out_c = out.d + (long)c * out.plane;
        const bool has_left = gx > 0, has_right = gx + 4 < W;
        const bool has_l = lane == 0 && has_left, has_r = lane == 31 && has_right;
        const long eoff = lane == 0 ? -1 : 4;                               // pixel beyond the strip's end
        // Everything that moves with the row is a running offset / pointer (no per-step 64-bit multiplies, no modulo):
        //   issue side: row ri, element offset oi = ri*W + gx, ring stage si at shared address shi
        //   consume side: stage sc at byte offset shc of the ring; output row y = n-1 at element offset s_out
        const int nbeg = y0 - 1, nend = y1;                                  // rows y0-1 .. y1 pass through the pipeline
        const int flo = nbeg < 0 ? 0 : nbeg, fhi = nend < H - 1 ? nend : H - 1;   // rows whose flags exist
        const unsigned char* fbase = active4 + (gx >> 2);
        const long Wq = W >> 2;
        auto flag_in = [&](int row) -> unsigned { return (col_in && row >= flo && row <= fhi) ? fbase[(long)row * Wq] : 0u; };
        const unsigned sh_base = (unsigned)__cvta_generic_to_shared(pm_ring) + (unsigned)tid * PM::QB;
        const unsigned sh_ebase = (unsigned)__cvta_generic_to_shared(pm_ring) + (unsigned)(PM::NA * 256 * PM::QB) + (unsigned)tid * (unsigned)sizeof(T);
        int ri = nbeg, si = 0;
        long oi = (long)nbeg * W + gx;
        auto cp16a = [](unsigned dst, const void* src, bool on) {
            const int nb = on ? 16 : 0;
            asm volatile("cp.async.cg.shared.global [%0], [%1], 16, %2;" ::"r"(dst), "l"(src), "r"(nb) : "memory");
        };
        auto cpqa = [&](unsigned dst, const T* src, bool on) {
#pragma unroll
            for (int k = 0; k < PM::QB / 16; ++k) cp16a(dst + 16 * k, (const unsigned char*)src + 16 * k, on);
        };
        auto cp1a = [](unsigned dst, const T* src, bool on) {
            if constexpr (sizeof(T) == 4) { const int nb = on ? 4 : 0; asm volatile("cp.async.ca.shared.global [%0], [%1], 4, %2;" ::"r"(dst), "l"(src), "r"(nb) : "memory"); }
            else                          { const int nb = on ? 8 : 0; asm volatile("cp.async.ca.shared.global [%0], [%1], 8, %2;" ::"r"(dst), "l"(src), "r"(nb) : "memory"); }
        };
        // issue the next row (ri) into ring stage si; fl = active bits of this lane's quad in that row
        auto issue = [&](unsigned fl) {
            const bool on = fl != 0;
            // rows of a neighbouring slab (slab mode, fused): the direction is the received p' itself; it goes through
            // the z slot with a zero-filled p slot, so the consumer's z + beta p yields it unchanged
            const bool ghost = FUSED && (ri < row_lo || ri >= row_hi);
            const long o = on ? oi : 0;                                      // a zero-size copy still gets a valid address
            const T* zsrc = ghost ? nc : zc;
            const unsigned sh = sh_base + (unsigned)si * PM::STAGE;
            cpqa(sh + PM::A_P * 256 * PM::QB, pc + o, on && !ghost);
            if (FUSED) cpqa(sh + PM::A_Z * 256 * PM::QB, zsrc + o, on);
            if (LM) cpqa(sh + PM::A_C * 256 * PM::QB, cc + o, on);
            if (lane == 0 || lane == 31) {
                // the neighbour quad may be inactive (all zeros) -- loading it is still correct
                const bool eon = on && (lane == 0 ? has_l : has_r);
                const long eo = eon ? o + eoff : 0;
                const unsigned she = sh_ebase + (unsigned)si * PM::STAGE;
                cp1a(she, pc + eo, eon && !ghost);
                if (FUSED) cp1a(she + 256 * (unsigned)sizeof(T), zsrc + eo, eon);
            }
            asm volatile("cp.async.commit_group;" ::: "memory");
            ++ri; oi += W; si = si + 1 == D ? 0 : si + 1;
        };
        // fq[k] = active bits of row n+k at the top of step n: fq[D-1] gates the issue of row n+D-1, fq[0] is consumed
        // with row n; the flag of row n+D is loaded at the top of the step and shifted in at its end
        unsigned fq[D];
#pragma unroll
        for (int k = 0; k < D; ++k) fq[k] = flag_in(nbeg + k);
#pragma unroll
        for (int k = 0; k < D - 1; ++k) issue(fq[k]);
        T prev[4] = {T(0), T(0), T(0), T(0)}, cur[4] = {T(0), T(0), T(0), T(0)}, ccur[4] = {T(0), T(0), T(0), T(0)};
        T ecur = T(0);
        unsigned fcur = 0;
        long s_out = (long)(nbeg - 1) * W + gx;                              // offset of output row n-1
        int sc = 0;
        const unsigned char* fnext_p = fbase + (long)(nbeg + D) * Wq;        // flag of row n+D
        for (int n = nbeg; n <= nend; ++n) {
            const int rn = n + D;
            const unsigned f_next = (col_in && rn >= flo && rn <= fhi) ? *fnext_p : 0u;
            fnext_p += Wq;
            issue(fq[D - 1]);
            if constexpr (D == 4) asm volatile("cp.async.wait_group 3;" ::: "memory");
            else                  asm volatile("cp.async.wait_group 2;" ::: "memory");
            // row n has landed in stage sc
            const unsigned char* stage = pm_ring + (size_t)sc * PM::STAGE + (size_t)tid * PM::QB;
            const unsigned char* estage = pm_ring + (size_t)sc * PM::STAGE + (size_t)(PM::NA * 256 * PM::QB) + (size_t)tid * sizeof(T);
            sc = sc + 1 == D ? 0 : sc + 1;
            const unsigned fnew = fq[0];
            T nw[4] = {T(0), T(0), T(0), T(0)}, cnew[4] = {T(0), T(0), T(0), T(0)};
            T enew = T(0);
            {
                T a[4]; load4(reinterpret_cast<const T*>(stage + PM::A_P * 256 * PM::QB), a);
                if (FUSED) {
                    T b[4]; load4(reinterpret_cast<const T*>(stage + PM::A_Z * 256 * PM::QB), b);
#pragma unroll
                    for (int j = 0; j < 4; ++j) nw[j] = b[j] + beta * a[j];
                } else {
#pragma unroll
                    for (int j = 0; j < 4; ++j) nw[j] = a[j];
                }
                if (LM) load4(reinterpret_cast<const T*>(stage + PM::A_C * 256 * PM::QB), cnew);
                if (lane == 0 || lane == 31) {
                    const T ea = *reinterpret_cast<const T*>(estage);
                    enew = FUSED ? *reinterpret_cast<const T*>(estage + 256 * sizeof(T)) + beta * ea : ea;
                }
            }
            // output row y = n-1 (cur) with prev above and nw below
            const int y = n - 1;
            T lft = __shfl_up_sync(0xffffffffu, cur[3], 1);
            T rgt = __shfl_down_sync(0xffffffffu, cur[0], 1);
            if (lane == 0) lft = ecur;
            if (lane == 31) rgt = ecur;
            if (y >= y0 && y < y1 && fcur != 0) {
                const bool up = y > 0, dn = y + 1 < H;
                // gx is a multiple of 4: only the first pixel of a row lacks its left neighbour and only the last one its
                // right neighbour (two item-invariant predicates instead of eight per step)
                const T nbr[4] = {cur[1], cur[2], cur[3], rgt}, nbl[4] = {lft, cur[0], cur[1], cur[2]};
                T o[4], pn[4];
                T rowdot = T(0);
#pragma unroll
                for (int j = 0; j < 4; ++j) {
                    const T v = cur[j];
                    T a = T(0);
                    if (j < 3 || has_right) a += v - nbr[j];                      // {1,0}
                    if (j > 0 || has_left) a += v - nbl[j];                       // {-1,0}
                    if (dn) a += v - nw[j];                                       // {0,1}
                    if (up) a += v - prev[j];                                     // {0,-1}
                    a = T(2) * a;
                    const bool exj = ((fcur >> j) & 1u) == 0;
                    pn[j] = exj ? T(0) : v;
                    o[j] = exj ? T(0) : a;
                    if (LM) { if (!exj) o[j] += ccur[j] * pn[j]; }
                    rowdot += pn[j] * o[j];
                }
                acc += (double)rowdot;
                if (FUSED) st4v(pnew_c + s_out, pn);
                st4v(out_c + s_out, o);
            }
            s_out += W;
#pragma unroll
            for (int j = 0; j < 4; ++j) { prev[j] = cur[j]; cur[j] = nw[j]; ccur[j] = cnew[j]; }
            ecur = enew; fcur = fnew;
#pragma unroll
            for (int k = 0; k + 1 < D; ++k) fq[k] = fq[k + 1];
            fq[D - 1] = f_next;
        }
        asm volatile("cp.async.wait_group 0;" ::: "memory");
    }
    block_accumulate(acc, den);
}

// One warp per (strip, band): does any quad of its owned rows hold an unknown?  Kept items are appended to the list
// (the four channel planes of an item stay adjacent in the march kernel's item numbering).
template <class T>
__global__ void __launch_bounds__(256) k_poisson_items(PoissonFunctor<T> f, const unsigned char* __restrict__ active4) {
    const int lane = threadIdx.x & 31;
    const long w = (blockIdx.x * (long)blockDim.x + threadIdx.x) >> 5;
    const int strips = f.strips(), bands = f.bands();
    if (w >= (long)strips * bands) return;
    const int strip = (int)(w % strips), band = (int)(w / strips);
    const int q = strip * 32 + lane, y0 = f.own_lo() + band * f.bandR, y1 = min(y0 + f.bandR, f.own_hi());
    unsigned any = 0;
    if (q < f.W / 4)
        for (int y = y0; y < y1; ++y) any |= active4[((long)y * f.W >> 2) + q];
    any = __any_sync(0xffffffffu, any != 0);
    if (any && lane == 0) f.items[atomicAdd(f.nactive, 1u)] = (unsigned)w;
}
template <class T> void PoissonFunctor<T>::build_active_items(cudaStream_t stream, const unsigned char* active4) const {
    OPTK_CUDA(cudaMemsetAsync(nactive, 0, sizeof(unsigned), stream));
    const long warps = (long)strips() * bands();
    k_poisson_items<T><<<(unsigned)((warps * 32 + 255) / 256), 256, 0, stream>>>(*this, active4);
}

template <class T> size_t PoissonFunctor<T>::fast_smem() const {
    if (tiled) return 0;
    return (size_t)PoissonMarch<T, true, true>::STAGE * PoissonMarch<T, true, true>::D;
}
template <class T> template <bool FUSED, bool LM> auto PoissonFunctor<T>::fast_kernel() const {
    return tiled ? &k_poisson_apply<T, FUSED, LM> : &k_poisson_march<T, FUSED, LM>;
}

}  // namespace optk
