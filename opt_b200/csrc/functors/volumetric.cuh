// functors/volumetric.cuh -- volumetric_mesh_deformation
// (examples/volumetric_mesh_deformation/volumetric_mesh_deformation.t:1-21)
//   fit(i)   = [C_i.x >= -999999.9] w_f (O_i - C_i)                                   (3 channels)
//   reg_o(i) = inb(i) inb(n) w_r [ (O_i - O_n) - R3D(A_i)(U_i - U_n) ]                (3 channels, 6 offsets)
//   no Exclude, UsePreconditioner(true); unknown element = (Offset.xyz, Angle.xyz)
// Closed forms: SURVEY.md App. D with D_ab = -[dR/da u | dR/db u | dR/dg u], u = U_a - U_b.
#pragma once
#include "math.cuh"

namespace optk {

// UB = 0: SoA solver vectors, scalar marching kernel (fp64, and fp32 grids whose rows are not a multiple of 32 sites);
// UB = 6: blocked solver vectors + the packed fp32 kernel of functors/volumetric_packed.cuh.  The plan chooses
// (examples/volumetric_mesh_deformation.cu).
template <class T, int UB = 0>
struct VolumetricFunctor : FunctorDefaults<T> {
    using Scalar = T;
    static constexpr int U = 6, ND = 3;
    static constexpr bool USE_PRE = true;
    // SoA / scalar engine: the generic lazy p' = z + beta p accessor doubles the 42 stencil loads of this functor and the
    // kernel is issue-bound, not DRAM-bound (measured: fused 2.28 ms vs 1.44 + 0.17 ms at 256^3), so PCGStep3 stays a
    // separate pass there.  Blocked / packed engine: z is staged next to p and p' is formed in registers
    // (volumetric_packed.cuh), in slab mode with the ghost planes taken from the pushed p' (engine.cuh k_halo_push_dir).
    static constexpr bool FUSE_STEP3 = UB > 0;
    static constexpr bool FUSE_SLAB = UB > 0;
    bool fuse_slab_ok() const { return UB > 0 && fast_ok(); }
    static constexpr int VEC_UB = UB;                          // blocked layout (common.cuh, Vec) for the packed kernel's bulk copies
    int W, H, D;
    T* O; T* A; const T* Ur; const T* Cn;
    T w_fit, w_reg;
    // Table refreshed by precompute (once per nonlinear iteration, like the reference's ComputedArrays,
    // o.t:3131-3153): sin/cos of the three Euler angles, UrShape and the fit-term predicate of every site,
    // stored in the tile order of the J^T J kernels: one record per 30 x FTY x 1 tile, laid out
    // [ty][plane 10][32 lanes], where lanes 0 and 31 DUPLICATE the x-neighbours of the tile.  A warp row
    // is therefore one aligned 128-byte line per plane (and all planes of a row, or of the whole record, are one
    // contiguous chunk for the bulk-copy engine), and every table access of a thread is
    // (one pointer) + (compile-time offset).  The kernel reads ONLY this table and p inside the PCG loop:
    // no sincos, no 12-byte-strided AoS loads.  Planes: sa ca sb cb sg cg ux uy uz fit.
    T* tab = nullptr;
    static constexpr int NP = 10;
    static constexpr bool HAS_PRECOMPUTE = true, HAS_CUSTOM_PRECOMPUTE = true;
    static constexpr bool HAS_FAST_APPLY = true;
    static constexpr int FTY = 8, FTZ = 1, FTX = 30, FNT = 32 * FTY * FTZ;
    static constexpr int ZC = 32;                               // planes one warp marches through per work unit
    static constexpr long REC = (long)NP * FNT;                 // elements per tile record
    static constexpr int PLS = 32, RWS = NP * 32;               // plane / row stride inside a record: [ty][plane][32 lanes]
    unsigned mbar_hint_ns = 0;   // suspend-time hint (ns) of the packed kernel's mbarrier waits; measured r2w: 0 / 2000 / 20000 ns all 1007-1013 it/s at 256^3
    void plan(const unsigned* d) {
        W = d[0]; H = d[1]; D = d[2];
        if (const char* e = getenv("OPTB200_MBAR_HINT_NS")) mbar_hint_ns = (unsigned)atoi(e);
    }
    __host__ __device__ int ntx() const { return (W + FTX - 1) / FTX; }
    __host__ __device__ int nty() const { return (H + FTY - 1) / FTY; }
    __host__ __device__ int ntz() const { return (D + FTZ - 1) / FTZ; }
    void alloc_temporaries(cudaStream_t) {
        OPTK_CUDA(cudaMalloc(&tab, (size_t)ntx() * nty() * ntz() * REC * sizeof(T)));
    }
    void free_temporaries() { cudaFree(tab); tab = nullptr; }
    size_t fast_smem() const;
    __device__ void precompute(int, int, int, long) const {}   // table is written by custom_precompute
    void custom_precompute(cudaStream_t stream, int n_sm) const;
    static constexpr bool use_packed() { return UB > 0; }
    bool fast_ok() const { return getenv("OPTB200_GENERIC_JTJ") == nullptr; }
    // a slab rank marches through its owned planes only
    __host__ __device__ int z_lo() const { return this->o_lo; }
    __host__ __device__ int z_hi() const { return this->o_hi < D ? this->o_hi : D; }
    __host__ __device__ int nseg() const { return (z_hi() - z_lo() + ZC - 1) / ZC; }
    // the packed kernel works on 30 x 16 tiles (two table records stacked in y)
    long fast_tiles() const { return (long)ntx() * (use_packed() ? (nty() + 1) / 2 : nty()) * nseg(); }
    dim3 fast_block() const { return dim3(32, FTY, FTZ); }
    template <bool FUSED, bool LM> auto fast_kernel() const;
    void bind(void** pp) {
        O = (T*)pp[0]; A = (T*)pp[1]; Ur = (const T*)pp[2]; Cn = (const T*)pp[3];
        w_fit = (T) * (const float*)pp[4]; w_reg = (T) * (const float*)pp[5];
    }
    __device__ bool exclude(int, int, int, long) const { return false; }
    __device__ T* xptr(long s, int u) const { return u < 3 ? O + s * 3 + u : A + s * 3 + (u - 3); }
    int unknown_images(void** ptrs, int* ch) const { ptrs[0] = O; ch[0] = 3; ptrs[1] = A; ch[1] = 3; return 2; }

    template <class Fn> __device__ __forceinline__ void neighbours(int x, int y, int z, long s, Fn&& fn) const {
        const long sy = W, sz = (long)W * H;
        if (x + 1 < W) fn(s + 1);
        if (x > 0) fn(s - 1);
        if (y + 1 < H) fn(s + sy);
        if (y > 0) fn(s - sy);
        if (z + 1 < D) fn(s + sz);
        if (z > 0) fn(s - sz);
    }
    __device__ bool fit_valid(long s) const { return Cn[s * 3] >= T(-999999.9); }

    __device__ T cost(int x, int y, int z, long s) const {
        T oi[3], ui[3], ai[3]; load3(O + s * 3, oi); load3(Ur + s * 3, ui); load3(A + s * 3, ai);
        Rot3<T> R; R.set(ai);
        T sum = T(0);
        if (fit_valid(s)) {
#pragma unroll
            for (int k = 0; k < 3; ++k) { T r = w_fit * (oi[k] - Cn[s * 3 + k]); sum += r * r; }
        }
        neighbours(x, y, z, s, [&](long n) {
            T on[3], un[3]; load3(O + n * 3, on); load3(Ur + n * 3, un);
            T u[3] = {ui[0] - un[0], ui[1] - un[1], ui[2] - un[2]}, Ru[3]; R.apply(u, Ru);
#pragma unroll
            for (int k = 0; k < 3; ++k) { T r = w_reg * ((oi[k] - on[k]) - Ru[k]); sum += r * r; }
        });
        return sum;
    }
    __device__ void evalJTF(int x, int y, int z, long s, T* g, T* d) const {
        T oi[3], ui[3], ai[3]; load3(O + s * 3, oi); load3(Ur + s * 3, ui); load3(A + s * 3, ai);
        Rot3<T> R; R.set(ai);
        const T wf2 = w_fit * w_fit, wr2 = w_reg * w_reg;
        if (fit_valid(s)) {
#pragma unroll
            for (int k = 0; k < 3; ++k) { g[k] += wf2 * (oi[k] - Cn[s * 3 + k]); d[k] += wf2; }
        }
        neighbours(x, y, z, s, [&](long n) {
            T on[3], un[3], an[3]; load3(O + n * 3, on); load3(Ur + n * 3, un); load3(A + n * 3, an);
            Rot3<T> Rn; Rn.set(an);
            T u[3] = {ui[0] - un[0], ui[1] - un[1], ui[2] - un[2]}, Ru[3], Rnu[3], Dm[3][3];
            R.apply(u, Ru); Rn.apply(u, Rnu); R.jac(u, Dm);
            T din[3], dni[3];
#pragma unroll
            for (int k = 0; k < 3; ++k) { din[k] = (oi[k] - on[k]) - Ru[k]; dni[k] = (on[k] - oi[k]) + Rnu[k]; }
#pragma unroll
            for (int k = 0; k < 3; ++k) { g[k] += wr2 * (din[k] - dni[k]); d[k] += T(2) * wr2; }
#pragma unroll
            for (int j = 0; j < 3; ++j) {
                g[3 + j] += wr2 * (Dm[0][j] * din[0] + Dm[1][j] * din[1] + Dm[2][j] * din[2]);
                d[3 + j] += wr2 * (Dm[0][j] * Dm[0][j] + Dm[1][j] * Dm[1][j] + Dm[2][j] * Dm[2][j]);
            }
        });
    }
    template <class P> __device__ void applyJTJ(int x, int y, int z, long s, const P& p, T* out) const {
        T ui[3], ai[3]; load3(Ur + s * 3, ui); load3(A + s * 3, ai);
        Rot3<T> R; R.set(ai);
        const T wf2 = w_fit * w_fit, wr2 = w_reg * w_reg;
        T pi[6], acc[6];
#pragma unroll
        for (int k = 0; k < 6; ++k) { pi[k] = p(k, s); acc[k] = T(0); }
        neighbours(x, y, z, s, [&](long n) {
            T un[3], an[3]; load3(Ur + n * 3, un); load3(A + n * 3, an);
            Rot3<T> Rn; Rn.set(an);
            T u[3] = {ui[0] - un[0], ui[1] - un[1], ui[2] - un[2]}, um[3] = {-u[0], -u[1], -u[2]}, Dm[3][3], Dn[3][3];
            R.jac(u, Dm); Rn.jac(um, Dn);                                           // D_in and D_ni
            T pn[6];
#pragma unroll
            for (int k = 0; k < 6; ++k) pn[k] = p(k, n);
            T jp[3];                                                                // (pO_i - pO_n) + D_in pA_i
#pragma unroll
            for (int k = 0; k < 3; ++k) {
                const T q = pi[k] - pn[k];
                const T da = Dm[k][0] * pi[3] + Dm[k][1] * pi[4] + Dm[k][2] * pi[5];
                const T dn = Dn[k][0] * pn[3] + Dn[k][1] * pn[4] + Dn[k][2] * pn[5];
                jp[k] = q + da;
                acc[k] += T(2) * q + da - dn;
            }
#pragma unroll
            for (int j = 0; j < 3; ++j) acc[3 + j] += Dm[0][j] * jp[0] + Dm[1][j] * jp[1] + Dm[2][j] * jp[2];
        });
        const T vf = fit_valid(s) ? wf2 : T(0);
#pragma unroll
        for (int k = 0; k < 3; ++k) { out[k] = vf * pi[k] + wr2 * acc[k]; out[3 + k] = wr2 * acc[3 + k]; }
    }
    template <class V> __device__ T modelCost(int x, int y, int z, long s, const V& dl) const {
        T oi[3], ui[3], ai[3]; load3(O + s * 3, oi); load3(Ur + s * 3, ui); load3(A + s * 3, ai);
        Rot3<T> R; R.set(ai);
        T di[6];
#pragma unroll
        for (int k = 0; k < 6; ++k) di[k] = dl(k, s);
        T sum = T(0);
        if (fit_valid(s)) {
#pragma unroll
            for (int k = 0; k < 3; ++k) { T m = w_fit * (oi[k] - Cn[s * 3 + k]) + w_fit * di[k]; sum += m * m; }
        }
        neighbours(x, y, z, s, [&](long n) {
            T on[3], un[3]; load3(O + n * 3, on); load3(Ur + n * 3, un);
            T u[3] = {ui[0] - un[0], ui[1] - un[1], ui[2] - un[2]}, Ru[3], Dm[3][3]; R.apply(u, Ru); R.jac(u, Dm);
#pragma unroll
            for (int k = 0; k < 3; ++k) {
                T jd = (di[k] - dl(k, n)) + Dm[k][0] * di[3] + Dm[k][1] * di[4] + Dm[k][2] * di[5];
                T m = w_reg * ((oi[k] - on[k]) - Ru[k]) + w_reg * jd;
                sum += m * m;
            }
        });
        return sum;
    }
};


// ---------------------------------------------------------------------------------------------
// Marching J^T J apply for the 3-D ARAP energy.
//
// With M_a = sum_j pA_a[j] dR/da_j(A_a) (3x3, one per site) and u = U_i - U_n, the J p of the regularisation
// residual centred at i toward n is  jp_in = (pO_i - pO_n) - M_i u, and
//     Ap_O(i) = vf pO_i + w_r^2 sum_n (jp_in - jp_ni)
//     Ap_A(i)[j] = -w_r^2 < dR/da_j(A_i), S_i >_F ,   S_i = sum_n jp_in u^T
// (SURVEY.md App. D rearranged: D_in pA_i = -M_i u, D_ni pA_n = +M_n u).  A warp is one x-row of 32 sites (the
// outer two lanes are halo lanes that only produce the jp toward the interior, so no lane diverges) and
// marches through ZC consecutive z-planes.  jp_ni comes from the neighbour lane by shuffle along x, from the
// thread's own previous / next step along z (registers), and is recomputed from the neighbour's table entry
// along y.  No shared memory and no barrier: warps are independent and prefetch their next plane (a variant
// that exchanged the y direction through shared memory executed 23 % fewer instructions but was slower: the
// per-plane barrier put every warp's loads in lockstep, profiles/r1_e).  The generic functor kernel needed
// 21 sincos and ~1850 instructions per site; this needs no sincos and ~40 % of the instructions.
// Writes the tile-ordered table (same thread <-> site mapping as k_volumetric_apply; slots outside the
// grid are zero-filled so that the apply kernel can read them unconditionally).
template <class F>
__global__ void __launch_bounds__(F::FNT) k_volumetric_table(F f) {
    using T = typename F::Scalar;
    constexpr int TX = F::FTX, TY = F::FTY, TZ = F::FTZ, NT = F::FNT;
    const int W = f.W, H = f.H, D = f.D;
    const int ntx = f.ntx(), nty = f.nty();
    const int ntiles = ntx * nty * f.ntz();
    const int lane = threadIdx.x, tid = linear_tid();
    for (int tile = blockIdx.x; tile < ntiles; tile += gridDim.x) {
        const int bx = tile % ntx, by = (tile / ntx) % nty, bz = tile / (ntx * nty);
        const int x = bx * TX + lane - 1, y = by * TY + (int)threadIdx.y, z = bz * TZ + (int)threadIdx.z;
        T v[F::NP];
#pragma unroll
        for (int k = 0; k < F::NP; ++k) v[k] = T(0);
        if (x >= 0 && x < W && y < H && z < D) {
            const long s = x + (long)W * (y + (long)H * z);
            T a[3]; load3(f.A + s * 3, a); load3(f.Ur + s * 3, v + 6);
#pragma unroll
            for (int k = 0; k < 3; ++k) sincos_t(a[k], v[2 * k], v[2 * k + 1]);
            v[9] = f.Cn[s * 3] >= T(-999999.9) ? T(1) : T(0);
        }
        T* rec = f.tab + (long)tile * F::REC + (int)threadIdx.y * F::RWS + lane;
#pragma unroll
        for (int k = 0; k < F::NP; ++k) rec[k * F::PLS] = v[k];
    }
}
template <class T, int UB> void VolumetricFunctor<T, UB>::custom_precompute(cudaStream_t stream, int n_sm) const {
    const long ntiles = (long)ntx() * nty() * ntz();
    const unsigned grid = (unsigned)(ntiles < (long)n_sm * 4 ? ntiles : (long)n_sm * 4);
    k_volumetric_table<VolumetricFunctor<T, UB>><<<grid, dim3(32, FTY, FTZ), 0, stream>>>(*this);
}




template <class T>
struct VolSite {                 // what a thread keeps of one plane while it is the next / current / pending one
    T tr[6], pc[6], uc[3], fit;  // table entry and direction p
    T acc[3];                    // sum_n (jp_in - jp_ni) collected so far
    T jn[3];                     // jp toward the plane above
};

template <class T, bool FUSED, bool LM>
__global__ void __launch_bounds__(VolumetricFunctor<T>::FNT, sizeof(T) == 4 ? 2 : 1)
k_volumetric_apply(VolumetricFunctor<T> f, Vec<T> zv, Vec<T> p, Vec<T> p_new, Vec<T> out, Vec<T> CtC,
                   const unsigned char* __restrict__, const double* bNum, const double* aNum, double* den,
                   double* zero_a, double* zero_b) {
    using F = VolumetricFunctor<T>;
    constexpr int TX = F::FTX, TY = F::FTY, NT = F::FNT, ZC = F::ZC;
    const int W = f.W, H = f.H, D = f.D;
    const int sy = W;
    const long sz = (long)W * H, pp = p.plane, op = out.plane;
    const int zlo = f.z_lo(), zhi = f.z_hi();
    const int ntx = f.ntx(), nty = f.nty();
    const int nunits = ntx * nty * f.nseg();
    const int lane = threadIdx.x, ty = threadIdx.y, tid = linear_tid();
    const T wr2 = f.w_reg * f.w_reg, wf2 = f.w_fit * f.w_fit;
    // table: record of (bx, by, z) = [plane][ty][lane]; the y neighbours are in this record or in the facing
    // row of the adjacent one (warp-uniform choice), the z neighbours one z-stride away
    const long rec_y = (long)ntx * F::REC, rec_z = (long)ntx * nty * F::REC;
    const bool typ = ty + 1 < TY, tym = ty > 0;
    const long oyp = typ ? F::RWS : rec_y - (TY - 1) * F::RWS, oym = tym ? -F::RWS : -rec_y + (TY - 1) * F::RWS;
    if (blockIdx.x == 0 && tid == 0) {
        if (zero_a) *zero_a = 0.0;
        if (zero_b) *zero_b = 0.0;
    }
    double dot = 0.0;
    for (int unit = blockIdx.x; unit < nunits; unit += gridDim.x) {
        const int bx = unit % ntx, by = (unit / ntx) % nty, seg = unit / (ntx * nty);
        const int x = bx * TX + lane - 1, y = by * TY + ty;
        const int z0 = zlo + seg * ZC, z1 = z0 + ZC < zhi ? z0 + ZC : zhi;
        const bool inside = x >= 0 && x < W && y < H;
        const bool inner = inside && lane >= 1 && lane <= TX;          // not a halo lane
        const bool hasR = inside && lane < 31 && x + 1 < W, hasL = inside && lane > 0 && x > 0;
        const bool eyp = inner && y + 1 < H, eym = inner && y > 0;
        long s = x + (long)sy * y + sz * z0;
        const T* __restrict__ ts = f.tab + ((long)(z0 * nty + by) * ntx + bx) * F::REC + ty * F::RWS + lane;
        const T* __restrict__ ps = p.d + s;
        bool havePrev = z0 > 0, writePrev = false;

        auto load_site = [&](VolSite<T>& c, const T* __restrict__ tn, const T* __restrict__ pn) {
#pragma unroll
            for (int k = 0; k < 6; ++k) { c.tr[k] = tn[k * F::PLS]; c.pc[k] = inside ? pn[k * pp] : T(0); }
#pragma unroll
            for (int k = 0; k < 3; ++k) c.uc[k] = tn[(6 + k) * F::PLS];
            c.fit = tn[9 * F::PLS];
        };
        // Ap_O of a completed plane (all six jp_in and jp_ni collected in acc)
        auto write_O = [&](const VolSite<T>& c, const T* acc, long sp) {
            T dsum = T(0);
#pragma unroll
            for (int k = 0; k < 3; ++k) {
                T o = (c.fit * wf2) * c.pc[k] + wr2 * acc[k];
                if (LM) o += CtC.d[k * CtC.plane + sp] * c.pc[k];
                out.d[k * op + sp] = o;
                dsum += c.pc[k] * o;
            }
            dot += (double)dsum;
        };
        // one plane: P = pending plane below, C = this plane, N = plane above (filled here).  Returns false
        // after the step that only completes the pending plane.
        auto step = [&](VolSite<T>& P, VolSite<T>& C, VolSite<T>& N, int z) -> bool {
            const bool last = z == z1;
            if (last && z1 >= D) {                                    // top plane of the grid: nothing above it
                if (writePrev && inner) write_O(P, P.acc, s - sz);
                return false;
            }
            const bool hasNext = !last && z + 1 < D;
            if (hasNext) load_site(N, ts + rec_z, ps + sz);           // prefetch the plane above
            RotM<T> rm; rm.set(C.tr);
            T M[3][3]; rm.combine(C.pc[3], C.pc[4], C.pc[5], M);
            T S[3][3] = {{T(0), T(0), T(0)}, {T(0), T(0), T(0)}, {T(0), T(0), T(0)}};
#pragma unroll
            for (int k = 0; k < 3; ++k) C.acc[k] = T(0);
            // jp of the residual centred here toward a neighbour with UrShape un and direction pn
            auto edge = [&](const T* un, const T* pn, T* u, T* q, T* jp) {
#pragma unroll
                for (int k = 0; k < 3; ++k) { u[k] = C.uc[k] - un[k]; q[k] = C.pc[k] - pn[k]; }
#pragma unroll
                for (int k = 0; k < 3; ++k) {
                    jp[k] = nfma(M[k][0], u[0], nfma(M[k][1], u[1], nfma(M[k][2], u[2], q[k])));
                    C.acc[k] += jp[k];
                    S[k][0] = fma(jp[k], u[0], S[k][0]); S[k][1] = fma(jp[k], u[1], S[k][1]); S[k][2] = fma(jp[k], u[2], S[k][2]);
                }
            };
            // ---- z-: toward the pending plane below; completes it --------------------------------------
            if (havePrev) {
                T u[3], q[3], jd[3];
                edge(P.uc, P.pc, u, q, jd);
#pragma unroll
                for (int k = 0; k < 3; ++k) C.acc[k] -= P.jn[k];
                if (writePrev && inner) {
                    T a[3] = {P.acc[0] - jd[0], P.acc[1] - jd[1], P.acc[2] - jd[2]};
                    write_O(P, a, s - sz);
                }
            }
            if (last) return false;
            // ---- x: both directions through shuffles -------------------------------------------------
            {
                T unr[3], unl[3], pnr[3], pnl[3];
#pragma unroll
                for (int k = 0; k < 3; ++k) {
                    unr[k] = __shfl_down_sync(0xffffffffu, C.uc[k], 1); unl[k] = __shfl_up_sync(0xffffffffu, C.uc[k], 1);
                    pnr[k] = __shfl_down_sync(0xffffffffu, C.pc[k], 1); pnl[k] = __shfl_up_sync(0xffffffffu, C.pc[k], 1);
                }
                T u[3], q[3], jr[3] = {T(0), T(0), T(0)}, jl[3] = {T(0), T(0), T(0)};
                if (hasR) edge(unr, pnr, u, q, jr);
                if (hasL) edge(unl, pnl, u, q, jl);
#pragma unroll
                for (int k = 0; k < 3; ++k) {
                    const T fromR = __shfl_down_sync(0xffffffffu, jl[k], 1);   // the right neighbour's jp toward me
                    const T fromL = __shfl_up_sync(0xffffffffu, jr[k], 1);
                    if (hasR) C.acc[k] -= fromR;
                    if (hasL) C.acc[k] -= fromL;
                }
            }
            // ---- y: own jp into the exchange slot; a neighbour row of another CTA is recomputed ----------
            auto yside = [&](bool exists, long toff, long poff) {
                if (exists) {
                    T jp[3];
                    const T* __restrict__ tn = ts + toff;
                    const T* __restrict__ pn_ = ps + poff;
                    T un[3], pn[3], u[3], q[3];
#pragma unroll
                    for (int k = 0; k < 3; ++k) { un[k] = tn[(6 + k) * F::PLS]; pn[k] = pn_[k * pp]; }
                    edge(un, pn, u, q, jp);
                    {
                        T t6[6], Mn[3][3];
#pragma unroll
                        for (int k = 0; k < 6; ++k) t6[k] = tn[k * F::PLS];
                        RotM<T> rn; rn.set(t6);
                        rn.combine(pn_[3 * pp], pn_[4 * pp], pn_[5 * pp], Mn);
#pragma unroll
                        for (int k = 0; k < 3; ++k)                   // acc -= jp_ni, jp_ni = M_n u - q
                            C.acc[k] = nfma(Mn[k][0], u[0], nfma(Mn[k][1], u[1], nfma(Mn[k][2], u[2], C.acc[k] + q[k])));
                    }
                }
            };
            yside(eyp, oyp, sy);
            yside(eym, oym, -sy);
            // ---- z+: toward the plane above (its jp toward me arrives in the next step) -----------------
#pragma unroll
            for (int k = 0; k < 3; ++k) C.jn[k] = T(0);
            if (hasNext) { T u[3], q[3]; edge(N.uc, N.pc, u, q, C.jn); }
            // ---- the angle block of this plane is complete ------------------------------------------
            if (inner) {
                T oA[3];
                rm.project(S, oA);
                T dsum = T(0);
#pragma unroll
                for (int k = 0; k < 3; ++k) {
                    T o = wr2 * oA[k];
                    if (LM) o += CtC.d[(3 + k) * CtC.plane + s] * C.pc[3 + k];
                    out.d[(3 + k) * op + s] = o;
                    dsum += C.pc[3 + k] * o;
                }
                dot += (double)dsum;
            }
            havePrev = true; writePrev = true;
            s += sz; ts += rec_z; ps += sz;
            return true;
        };

        VolSite<T> A, B, C;
        load_site(A, ts, ps);
        if (havePrev) {                                               // jp of the plane below toward plane z0
            load_site(C, ts - rec_z, ps - sz);
            RotM<T> rn; rn.set(C.tr);
            T Mn[3][3]; rn.combine(C.pc[3], C.pc[4], C.pc[5], Mn);
            const T u[3] = {C.uc[0] - A.uc[0], C.uc[1] - A.uc[1], C.uc[2] - A.uc[2]};
#pragma unroll
            for (int k = 0; k < 3; ++k)
                C.jn[k] = nfma(Mn[k][0], u[0], nfma(Mn[k][1], u[1], nfma(Mn[k][2], u[2], C.pc[k] - A.pc[k])));
        }
        for (int z = z0;; z += 3) {                                   // roles rotate: no register shuffling
            if (!step(C, A, B, z)) break;
            if (!step(A, B, C, z + 1)) break;
            if (!step(B, C, A, z + 2)) break;
        }
    }
    block_accumulate(dot, den);
}

}  // namespace optk
