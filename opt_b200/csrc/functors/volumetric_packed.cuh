// functors/volumetric_packed.cuh -- fp32 J^T J apply of volumetric_mesh_deformation: packed fp32 math fed by TMA.
//
// Same algorithm as k_volumetric_apply (functors/volumetric.cuh): a warp is one x-row of 32 sites (lanes 0 and 31
// are halo lanes) that marches through ZC z-planes, x neighbours by shuffle, z neighbours carried in registers.
// That kernel is issue-bound (profiles/r1_f: 737 instructions per site and plane at 65 % issue utilisation, DRAM
// at 40 %), and a third of its instructions are 64-bit address arithmetic for ~50 scalar loads.  Here
//   * every thread owns TWO y-adjacent rows and all arithmetic is done on float2 pairs (lo = row y0, hi = row
//     y0+1) with FFMA2 / FADD2 / FMUL2 (sm_100 `fma.rn.f32x2`): one instruction does the work of two.  ptxas folds
//     negation, half swaps and scalar broadcasts into the operand modifiers, so building "the other row"
//     operands costs nothing;
//   * the inputs of a plane (table: 10 planes; direction: 6 channels of p and, in the fused kernel, of z; 18 rows
//     each, i.e. a 30 x 16 tile plus one halo row above and below) are brought into shared memory asynchronously
//     into a ring of NS stages guarded by full/empty mbarriers.  The table is tile-ordered [ty][plane][32 lanes]
//     (a whole 8-row record is one 10 KB chunk) and travels by the bulk-copy engine (`cp.async.bulk`, <= 4 copies
//     per stage).  The solver vectors use the blocked layout [site/32][channel][site%32] (common.cuh Vec, VEC_UB):
//     the 32-site window of a tile row starts at x = 30 bx - 1 and straddles two blocks, so it is fetched by
//     per-lane 4-byte `cp.async` (one coalesced 128-byte request per channel and row, zero-filled outside the grid)
//     whose completion is tied to the stage's `full` barrier (`cp.async.mbarrier.arrive.noinc`).  (r1: the window was
//     over-fetched as two whole blocks per row by the bulk engine, 1536 B per row and array; staging z next to p for
//     the fused kernel would not have fitted into 227 KB that way.)  The warps take turns issuing a stage.  Every
//     read of the march is then an LDS at (one base register) + (compile-time offset) and every store an STG at
//     pointer + immediate: no address arithmetic, no predicates, and the loads run three planes ahead of the math
//     without holding registers.  Rows outside the grid are simply not copied: the ring is zero-filled once and
//     only ever holds finite data, and every edge that leaves the grid is masked;
//   * FUSED: PCGStep3 of the previous iteration is folded in (solver:814-845 + 607-632): z is staged next to p, the
//     new search direction p' = z + beta p is formed in registers wherever the stencil reads it (own rows, the two
//     halo rows, the plane above) and written once at the thread's own sites into the other p buffer.  In slab mode
//     the ghost planes hold the neighbours' p' already (pushed before this kernel, engine.cuh k_halo_push_dir):
//     they are staged from p_new into the z slot and combined with beta = 0;
//   * the y edge between the two rows of a thread is internal (its J p in both directions comes out of the two
//     packed y edges); only the rows above and below the pair are recomputed from the table: ONE packed
//     recompute per two sites instead of two scalar ones per site;
//   * edges that do not exist (grid boundary, ragged tiles) are removed by a 0/1 mask pair on J p -- the mask of
//     an edge is the same on both of its end points, so incoming J p need no mask.
#pragma once
#include "volumetric.cuh"

namespace optk {

struct f2 { float2 v; };
__device__ __forceinline__ f2 mk2(float a, float b) { f2 r; r.v = make_float2(a, b); return r; }
__device__ __forceinline__ f2 bc2(float a) { return mk2(a, a); }
__device__ __forceinline__ f2 operator-(f2 a) { return mk2(-a.v.x, -a.v.y); }
__device__ __forceinline__ f2 operator+(f2 a, f2 b) { f2 r; r.v = __fadd2_rn(a.v, b.v); return r; }
__device__ __forceinline__ f2 operator-(f2 a, f2 b) { f2 r; r.v = __fadd2_rn(a.v, (-b).v); return r; }
__device__ __forceinline__ f2 operator*(f2 a, f2 b) { f2 r; r.v = __fmul2_rn(a.v, b.v); return r; }
__device__ __forceinline__ f2 fma2(f2 a, f2 b, f2 c) { f2 r; r.v = __ffma2_rn(a.v, b.v, c.v); return r; }      // a b + c
__device__ __forceinline__ f2 nfma2(f2 a, f2 b, f2 c) { f2 r; r.v = __ffma2_rn((-a).v, b.v, c.v); return r; }  // c - a b
__device__ __forceinline__ f2 shfl_down2(f2 a) { return mk2(__shfl_down_sync(0xffffffffu, a.v.x, 1), __shfl_down_sync(0xffffffffu, a.v.y, 1)); }
__device__ __forceinline__ f2 shfl_up2(f2 a) { return mk2(__shfl_up_sync(0xffffffffu, a.v.x, 1), __shfl_up_sync(0xffffffffu, a.v.y, 1)); }

// RotM (functors/math.cuh) on pairs, with explicit packed FMAs
struct RotM2 {
    f2 R[3][3], B[3][3];
    __device__ __forceinline__ void set(const f2* t) {   // t = sa ca sb cb sg cg
        const f2 sa = t[0], ca = t[1], sb = t[2], cb = t[3], sg = t[4], cg = t[5];
        const f2 cgsb = cg * sb, sgsb = sg * sb;
        R[0][0] = cg * cb; R[0][1] = nfma2(sg, ca, cgsb * sa); R[0][2] = fma2(cgsb, ca, sg * sa);
        R[1][0] = sg * cb; R[1][1] = fma2(sgsb, sa, cg * ca);  R[1][2] = nfma2(cg, sa, sgsb * ca);
        R[2][0] = -sb;     R[2][1] = cb * sa;                  R[2][2] = cb * ca;
        B[0][0] = -cgsb;   B[0][1] = R[0][0] * sa;             B[0][2] = R[0][0] * ca;
        B[1][0] = -sgsb;   B[1][1] = R[1][0] * sa;             B[1][2] = R[1][0] * ca;
        B[2][0] = -cb;     B[2][1] = -(sb * sa);               B[2][2] = -(sb * ca);
    }
    __device__ __forceinline__ void combine(f2 pa, f2 pb, f2 pg, f2 M[3][3]) const {
        M[0][0] = nfma2(pg, R[1][0], pb * B[0][0]);
        M[0][1] = nfma2(pg, R[1][1], fma2(pb, B[0][1], pa * R[0][2]));
        M[0][2] = nfma2(pg, R[1][2], nfma2(pa, R[0][1], pb * B[0][2]));
        M[1][0] = fma2(pg, R[0][0], pb * B[1][0]);
        M[1][1] = fma2(pg, R[0][1], fma2(pb, B[1][1], pa * R[1][2]));
        M[1][2] = fma2(pg, R[0][2], nfma2(pa, R[1][1], pb * B[1][2]));
        M[2][0] = pb * B[2][0];
        M[2][1] = fma2(pb, B[2][1], pa * R[2][2]);
        M[2][2] = nfma2(pa, R[2][1], pb * B[2][2]);
    }
    // < dR/da_j, S >_F for j = alpha, beta, gamma (the caller negates)
    __device__ __forceinline__ void project(const f2 S[3][3], f2* o) const {
        f2 a = nfma2(R[0][1], S[0][2], R[0][2] * S[0][1]);
        f2 b = fma2(B[0][2], S[0][2], fma2(B[0][1], S[0][1], B[0][0] * S[0][0]));
        f2 g = nfma2(R[1][0], S[0][0], R[0][0] * S[1][0]);
#pragma unroll
        for (int k = 1; k < 3; ++k) {
            a = nfma2(R[k][1], S[k][2], fma2(R[k][2], S[k][1], a));
            b = fma2(B[k][2], S[k][2], fma2(B[k][1], S[k][1], fma2(B[k][0], S[k][0], b)));
            g = nfma2(R[1][k], S[0][k], fma2(R[0][k], S[1][k], g));
        }
        o[0] = a; o[1] = b; o[2] = g;
    }
};


struct VolCarry {                // what a thread keeps of a plane after its own step (lo = row y0, hi = row y0 + 1)
    f2 uc[3], pO[3], fit;
    f2 acc[3];                   // sum_n (jp_in - jp_ni) collected so far
    f2 jn[3];                    // jp toward the plane above
};

namespace bulk {
__device__ __forceinline__ unsigned smem_u32(const void* p) { return (unsigned)__cvta_generic_to_shared(p); }
__device__ __forceinline__ void mbar_init(unsigned long long* b, unsigned count) {
    asm volatile("mbarrier.init.shared::cta.b64 [%0], %1;" ::"r"(smem_u32(b)), "r"(count) : "memory");
}
__device__ __forceinline__ void mbar_expect_tx(unsigned long long* b, unsigned bytes) {
    asm volatile("mbarrier.arrive.expect_tx.shared::cta.b64 _, [%0], %1;" ::"r"(smem_u32(b)), "r"(bytes) : "memory");
}
__device__ __forceinline__ void mbar_arrive(unsigned long long* b) {
    asm volatile("mbarrier.arrive.shared::cta.b64 _, [%0];" ::"r"(smem_u32(b)) : "memory");
}
// try_wait may suspend the thread in hardware up to `hint_ns` before it reports "not yet" (0: return at once, i.e. spin): with
// the short default limit the full-barrier waits of the march spun ~35 times per step (profiles/README.md r2d opcode mix:
// SYNCS 55 + BRA + YIELD), issue slots and power that a power-capped part pays for in clock.
__device__ __forceinline__ void mbar_wait(unsigned long long* b, unsigned parity, unsigned hint_ns) {
    unsigned ok;
    do {
        asm volatile("{\n .reg .pred p;\n mbarrier.try_wait.parity.shared::cta.b64 p, [%1], %2, %3;\n selp.b32 %0, 1, 0, p;\n}"
                     : "=r"(ok) : "r"(smem_u32(b)), "r"(parity), "r"(hint_ns) : "memory");
    } while (!ok);
}
// contiguous global -> shared copy by the bulk-copy engine; 16-byte aligned addresses and size
__device__ __forceinline__ void copy(void* dst, const void* src, unsigned bytes, unsigned long long* bar) {
    asm volatile("cp.async.bulk.shared::cluster.global.mbarrier::complete_tx::bytes [%0], [%1], %2, [%3];"
                 ::"r"(smem_u32(dst)), "l"(src), "r"(bytes), "r"(smem_u32(bar)) : "memory");
}
// per-thread 16-byte asynchronous copy (LDGSTS.128, L2 only); src_bytes = 0 zero-fills the destination instead
__device__ __forceinline__ void copy16(void* dst, const void* src, unsigned src_bytes) {
    asm volatile("cp.async.cg.shared.global [%0], [%1], 16, %2;" ::"r"(smem_u32(dst)), "l"(src), "r"(src_bytes) : "memory");
}
// the barrier receives one arrival from this thread once all its earlier cp.async have landed (no pending-count
// increment: the barrier's arrival count must include it)
__device__ __forceinline__ void arrive_on_copies(unsigned long long* bar) {
    asm volatile("cp.async.mbarrier.arrive.noinc.shared::cta.b64 [%0];" ::"r"(smem_u32(bar)) : "memory");
}
}  // namespace bulk

struct VolRing {
    static constexpr int NS = 4;                         // ring stages (planes in flight or in use)
    static constexpr int AHEAD = NS - 1;                 // elements issued ahead of the one being released
    static constexpr int NWARPS = 8, TILEY = 2 * NWARPS; // a warp owns two rows: 30 x 16 tiles
    static constexpr int BOXR = TILEY + 2;               // rows per stage: the tile + one halo row on each side
    static constexpr int TROW = 10 * 32;                 // table: floats per stage row, [plane][32 lanes]
    // direction (p and z): [row][channel][DW floats]: the 32-site window of a tile row starts at x = 30 bx - 1, which is never
    // 16-byte aligned; the aligned superset of 36 floats (9 chunks of 16 bytes, none of which straddles a 32-site block)
    // is staged instead and the march reads at (30 bx - 1) % 4 + lane
    static constexpr int DW = 36, DROW = 6 * DW;
    static constexpr int TABF = BOXR * TROW, DIRF = BOXR * DROW, STAGEF = TABF + 2 * DIRF;   // table, p, z
    static constexpr size_t SMEM = (size_t)NS * STAGEF * 4 + 2 * NS * 8 + 128;
};

template <bool FUSED, bool LM>
__global__ void __launch_bounds__(256, 1)
k_volumetric_apply_packed(VolumetricFunctor<float, 6> f, Vec<float, 6> zv, Vec<float, 6> p, Vec<float, 6> p_new, Vec<float, 6> out,
                          Vec<float, 6> CtC, const unsigned char* __restrict__, const double* bNum, const double* aNum, double* den,
                          double* zero_a, double* zero_b) {
    using F = VolumetricFunctor<float, 6>;
    using G = VolRing;
    constexpr int TX = F::FTX, TY = G::TILEY, ZC = F::ZC, NS = G::NS, PL = 32, DL = G::DW;     // PL / DL: plane / channel stride inside a table row / a direction row
    const int W = f.W, H = f.H, D = f.D;
    const long sz = (long)W * H;
    const int zlo = f.z_lo(), zhi = f.z_hi();
    const unsigned hint = f.mbar_hint_ns;
    const int ntx = f.ntx(), nty = (f.nty() + 1) / 2, nrec_y = f.nty();
    const int nunits = ntx * nty * f.nseg();
    const int lane = threadIdx.x, wy = threadIdx.y, ty0 = 2 * wy, tid = linear_tid();
    const f2 wr2 = bc2(f.w_reg * f.w_reg), wf2 = bc2(f.w_fit * f.w_fit);
    const long orow = 6L * W, oplane = 6L * sz;           // blocked layout: floats from a site to the one a row / a plane further
    // FUSED: beta of the PCGStep3 folded in here (engine.cuh FusedDir: the same float division as k_pcg_step3)
    const float beta = FUSED ? (float)(*bNum) / (float)(*aNum) : 0.f;

    extern __shared__ __align__(1024) float ring[];
    unsigned long long* full = reinterpret_cast<unsigned long long*>(ring + NS * G::STAGEF);
    unsigned long long* empty = full + NS;
    for (int i = tid; i < NS * G::STAGEF; i += 256) ring[i] = 0.f;
    if (tid == 0) {
#pragma unroll
        for (int i = 0; i < NS; ++i) { bulk::mbar_init(full + i, 1 + 32); bulk::mbar_init(empty + i, G::NWARPS); }   // full: expect_tx + the 32 lanes' copies
        asm volatile("fence.mbarrier_init.release.cluster;" ::: "memory");
        if (blockIdx.x == 0) {
            if (zero_a) *zero_a = 0.0;
            if (zero_b) *zero_b = 0.0;
        }
    }
    asm volatile("fence.proxy.async.shared::cta;" ::: "memory");     // the zero fill above vs the bulk copies below
    __syncthreads();

    // ---- producer: every warp walks the same (unit, plane) element sequence as the march, AHEAD elements in front
    // of its own releases; element e is issued by warp e % NWARPS, one copy per lane ----------------------------------
    int pr_unit = blockIdx.x, pr_z = 0, pr_ze = -1, pr_e = 0;
    auto pr_open = [&]() {                                // first / last plane of the current producer unit
        if (pr_unit >= nunits) return;
        const int seg = pr_unit / (ntx * nty);
        const int z0 = zlo + seg * ZC, z1 = z0 + ZC < zhi ? z0 + ZC : zhi;
        pr_z = z0 > 0 ? z0 - 1 : z0; pr_ze = z1 < D ? z1 : D - 1;
    };
    auto issue = [&]() {                                  // all lanes of the issuing warp
        const int st = pr_e % NS;
        if (pr_e >= NS) bulk::mbar_wait(empty + st, ((pr_e / NS) - 1) & 1, hint);
        const int bx = pr_unit % ntx, by = (pr_unit / ntx) % nty;
        float* tabs = ring + st * G::STAGEF;
        float* dirs = tabs + G::TABF;
        // table: records 2 by - 1 (its last row), 2 by, 2 by + 1, 2 by + 2 (its first row)
        const long rec_y = (long)ntx * F::REC;
        const float* rec = f.tab + ((long)(pr_z * nrec_y + 2 * by) * ntx + bx) * F::REC;
        const bool t0 = by > 0, t2 = 2 * by + 1 < nrec_y, t3 = 2 * by + 2 < nrec_y;
        if (lane == 0) {
            bulk::mbar_expect_tx(full + st, 10240u + (t0 ? 1280u : 0u) + (t2 ? 10240u : 0u) + (t3 ? 1280u : 0u));
            bulk::copy(tabs + G::TROW, rec, 10240, full + st);
        } else if (lane == 1) { if (t2) bulk::copy(tabs + 9 * G::TROW, rec + rec_y, 10240, full + st); }
        else if (lane == 2) { if (t0) bulk::copy(tabs, rec - rec_y + 7 * F::RWS, 1280, full + st); }
        else if (lane == 3) { if (t3) bulk::copy(tabs + 17 * G::TROW, rec + 2 * rec_y, 1280, full + st); }
        // direction: 16-byte chunks.  Lanes 0..26 = 3 channels x 9 chunks; two copies per row and array cover the six
        // channels of the 36-float superset [xa, xa + 36), xa = (30 bx - 1) & ~3, zero-filled outside the grid in x.
        // Fused: p into the p slot and z into the z slot; a ghost plane of a slab (its p' was pushed by the neighbour
        // before this kernel) goes from p_new into the z slot and is combined with beta = 0.
        const int g = lane / 9, ck = lane - 9 * g;                          // channel group member, chunk
        const int xc = ((bx * TX - 1) & ~3) + 4 * ck;                       // first site of this lane's chunk
        const unsigned okx = lane < 27 && xc >= 0 && xc + 3 < W ? 16u : 0u;
        const long lane_off = okx ? (long)(xc >> 5) * 192 + (xc & 31) + g * 32 : 0;
        const int ya = by * TY - 1;                                         // grid row of stage row 0
        const int r_lo = ya < 0 ? 1 : 0, r_hi = ya + G::BOXR <= H ? G::BOXR : H - ya;   // stage rows [r_lo, r_hi) exist
        const bool ghost = FUSED && (pr_z < zlo || pr_z >= zhi);
        const long row0 = ((long)pr_z * H + (ya + r_lo)) * orow + lane_off;
        const float* sa = (ghost ? p_new.d : p.d) + row0;                   // -> p slot (z slot for a ghost plane)
        const float* sb = zv.d + row0;                                      // -> z slot
        float* da = dirs + (ghost ? G::DIRF : 0) + r_lo * G::DROW + g * G::DW + 4 * ck;
        if (lane < 27) {
            for (int r = r_lo; r < r_hi; ++r) {
                bulk::copy16(da, sa, okx);
                bulk::copy16(da + 3 * G::DW, sa + 96, okx);
                if (FUSED && !ghost) {
                    bulk::copy16(da + G::DIRF, sb, okx);
                    bulk::copy16(da + G::DIRF + 3 * G::DW, sb + 96, okx);
                }
                sa += orow; sb += orow; da += G::DROW;
            }
        }
        bulk::arrive_on_copies(full + st);
    };
    auto produce_one = [&]() {
        if (pr_unit >= nunits) return;
        if (pr_e % G::NWARPS == wy) issue();
        ++pr_e;
        if (++pr_z > pr_ze) { pr_unit += gridDim.x; pr_open(); }
    };
    pr_open();
    for (int i = 0; i < G::AHEAD; ++i) produce_one();

    int e0 = 0;                                           // element index of the first plane of the current unit
    double dot = 0.0;
    for (int unit = blockIdx.x; unit < nunits; unit += gridDim.x) {
        const int bx = unit % ntx, by = (unit / ntx) % nty, seg = unit / (ntx * nty);
        const int x = bx * TX + lane - 1, y0 = by * TY + ty0, y1 = y0 + 1;
        const int z0 = zlo + seg * ZC, z1 = z0 + ZC < zhi ? z0 + ZC : zhi;
        const int zs = z0 > 0 ? z0 - 1 : z0, ze = z1 < D ? z1 : D - 1;
        const bool inx = x >= 0 && x < W, in0 = inx && y0 < H, in1 = inx && y1 < H;
        const bool innerl = lane >= 1 && lane <= TX;
        const bool w0 = in0 && innerl, w1 = in1 && innerl;                 // sites this thread writes
        const bool xr = lane < 31 && x + 1 < W, xl = lane > 0 && x > 0;
        const f2 m_xr = mk2(in0 && xr ? 1.f : 0.f, in1 && xr ? 1.f : 0.f), m_xl = mk2(in0 && xl ? 1.f : 0.f, in1 && xl ? 1.f : 0.f);
        // packed y edges: "plus" = (row0 -> row1, row1 -> row above), "minus" = (row0 -> row below, row1 -> row0)
        const bool eup = in1 && y1 + 1 < H, edn = in0 && y0 > 0;
        const f2 m_yp = mk2(in1 ? 1.f : 0.f, eup ? 1.f : 0.f), m_ym = mk2(edn ? 1.f : 0.f, in1 ? 1.f : 0.f);
        const f2 m_ext = mk2(edn ? 1.f : 0.f, eup ? 1.f : 0.f);
        // a thread's rows of a stage: table at tb[k PL], tb[k PL + TROW]; direction at db[k DL], db[k DL + DROW], z slot DIRF further
        const int toff = (1 + ty0) * G::TROW + lane;
        const int doff = G::TABF + (1 + ty0) * G::DROW + ((bx * TX - 1) & 3) + lane;   // the thread's first row; channel k at + DW k, second row DROW further
        // out (and CtC) in the blocked layout: channel k of this thread's row-0 site at os[k * 32], row 1 at os[k * 32 + orow]
        const long s0 = x + (long)W * y0 + sz * z0;
        float* os = out.d + ((s0 >> 5) * 192 + (s0 & 31));
        const float* cs = CtC.d + ((s0 >> 5) * 192 + (s0 & 31));
        float* ns = p_new.d + ((s0 >> 5) * 192 + (s0 & 31));            // FUSED: where this thread's p' goes
        bool havePrev = z0 > 0, writePrev = false;

        auto acquire = [&](int z) -> const float* {                       // stage base of plane z, once its copies have landed
            const int e = e0 + (z - zs);
            bulk::mbar_wait(full + (e % NS), (e / NS) & 1, hint);
            return ring + (e % NS) * G::STAGEF;
        };
        auto release = [&](int z) {
            const int e = e0 + (z - zs);
            __syncwarp();
            if (lane == 0) bulk::mbar_arrive(empty + (e % NS));
            produce_one();
        };
        auto tpair = [&](const float* q) { return mk2(q[0], q[G::TROW]); };
        // direction of the thread's two rows: p, or p' = z + beta p (b2 = 0 for a ghost plane, whose z slot holds p')
        auto dpair = [&](const float* q, f2 b2) {
            if (FUSED) return fma2(b2, mk2(q[0], q[G::DROW]), mk2(q[G::DIRF], q[G::DIRF + G::DROW]));
            return mk2(q[0], q[G::DROW]);
        };
        // ... of the stage rows below / above the pair (lo = below, hi = above)
        auto dext = [&](const float* q, f2 b2) {
            if (FUSED) return fma2(b2, mk2(q[-G::DROW], q[2 * G::DROW]), mk2(q[G::DIRF - G::DROW], q[G::DIRF + 2 * G::DROW]));
            return mk2(q[-G::DROW], q[2 * G::DROW]);
        };
        auto plane_beta = [&](int z) { return bc2(z >= zlo && z < zhi ? beta : 0.f); };

        // Ap_O of a completed plane; `back` = 1 for the plane below the current one
        auto write_O = [&](const VolCarry& c, const f2* acc, int back) {
            float* o_ = os - back * oplane;
            const float* c_ = cs - back * oplane;
            f2 dsum = bc2(0.f);
            const f2 fw = c.fit * wf2;
#pragma unroll
            for (int k = 0; k < 3; ++k) {
                f2 o = fma2(wr2, acc[k], fw * c.pO[k]);
                if (LM) o = fma2(mk2(w0 ? c_[k * 32] : 0.f, w1 ? c_[k * 32 + orow] : 0.f), c.pO[k], o);
                if (w0) o_[k * 32] = o.v.x;
                if (w1) o_[k * 32 + orow] = o.v.y;
                dsum = fma2(c.pO[k], o, dsum);
            }
            dot += (double)((w0 ? dsum.v.x : 0.f) + (w1 ? dsum.v.y : 0.f));
        };
        // one plane: P = carry of the plane below (pending), C = carry of this plane (filled here)
        auto step = [&](VolCarry& P, VolCarry& C, int z) -> bool {
            const bool last = z == z1;
            if (last && z1 >= D) {                                    // top plane of the grid: nothing above it
                if (writePrev) write_O(P, P.acc, 1);
                return false;
            }
            const float* __restrict__ stg = acquire(z);
            const float* __restrict__ tb = stg + toff;
            const float* __restrict__ db = stg + doff;
            f2 tr[6], pA[3];
            const f2 bz = plane_beta(z);
#pragma unroll
            for (int k = 0; k < 6; ++k) tr[k] = tpair(tb + k * PL);
#pragma unroll
            for (int k = 0; k < 3; ++k) { C.uc[k] = tpair(tb + (6 + k) * PL); C.pO[k] = dpair(db + k * DL, bz); pA[k] = dpair(db + (3 + k) * DL, bz); }
            C.fit = tpair(tb + 9 * PL);
            if (FUSED && !last) {                                      // PCGStep3's store: p' once, at the thread's own sites
#pragma unroll
                for (int k = 0; k < 3; ++k) {
                    if (w0) { ns[k * 32] = C.pO[k].v.x; ns[(3 + k) * 32] = pA[k].v.x; }
                    if (w1) { ns[k * 32 + orow] = C.pO[k].v.y; ns[(3 + k) * 32 + orow] = pA[k].v.y; }
                }
            }
            RotM2 rm; rm.set(tr);
            f2 M[3][3]; rm.combine(pA[0], pA[1], pA[2], M);
            f2 S[3][3];
#pragma unroll
            for (int k = 0; k < 3; ++k) { C.acc[k] = bc2(0.f); S[k][0] = S[k][1] = S[k][2] = bc2(0.f); }
            // jp of the residuals centred here toward neighbours with UrShape un and direction pn
            auto edge = [&](const f2* un, const f2* pn, f2* u, f2* q, f2* jp, bool masked, f2 m) {
#pragma unroll
                for (int k = 0; k < 3; ++k) { u[k] = C.uc[k] - un[k]; q[k] = C.pO[k] - pn[k]; }
#pragma unroll
                for (int k = 0; k < 3; ++k) {
                    jp[k] = nfma2(M[k][0], u[0], nfma2(M[k][1], u[1], nfma2(M[k][2], u[2], q[k])));
                    if (masked) jp[k] = jp[k] * m;
                    C.acc[k] = C.acc[k] + jp[k];
                    S[k][0] = fma2(jp[k], u[0], S[k][0]); S[k][1] = fma2(jp[k], u[1], S[k][1]); S[k][2] = fma2(jp[k], u[2], S[k][2]);
                }
            };
            // ---- z-: toward the pending plane below; completes it --------------------------------------
            if (havePrev) {
                f2 u[3], q[3], jd[3];
                edge(P.uc, P.pO, u, q, jd, false, bc2(1.f));
#pragma unroll
                for (int k = 0; k < 3; ++k) C.acc[k] = C.acc[k] - P.jn[k];
                if (writePrev) {
                    f2 a[3] = {P.acc[0] - jd[0], P.acc[1] - jd[1], P.acc[2] - jd[2]};
                    write_O(P, a, 1);
                }
            }
            if (!last) {              // (one release site per step instantiation: the inlined stage-issue code is large)
            // ---- x: both directions through shuffles -------------------------------------------------
            {
                f2 unr[3], unl[3], pnr[3], pnl[3];
#pragma unroll
                for (int k = 0; k < 3; ++k) {
                    unr[k] = shfl_down2(C.uc[k]); unl[k] = shfl_up2(C.uc[k]);
                    pnr[k] = shfl_down2(C.pO[k]); pnl[k] = shfl_up2(C.pO[k]);
                }
                f2 u[3], q[3], jr[3], jl[3];
                edge(unr, pnr, u, q, jr, true, m_xr);
                edge(unl, pnl, u, q, jl, true, m_xl);
#pragma unroll
                for (int k = 0; k < 3; ++k) C.acc[k] = (C.acc[k] - shfl_down2(jl[k])) - shfl_up2(jr[k]);
            }
            // ---- y: the edge between the two rows is internal; the rows above / below are recomputed ------
            {
                const float* __restrict__ tu = tb + 2 * G::TROW; const float* __restrict__ td = tb - G::TROW;   // stage rows above / below the pair
                f2 unp[3], pnp[3], unm[3], pnm[3], tn[6], pan[3];
#pragma unroll
                for (int k = 0; k < 3; ++k) {
                    const f2 pe = dext(db + k * DL, bz);                         // (row below, row above)
                    unp[k] = mk2(C.uc[k].v.y, tu[(6 + k) * PL]); pnp[k] = mk2(C.pO[k].v.y, pe.v.y);
                    unm[k] = mk2(td[(6 + k) * PL], C.uc[k].v.x); pnm[k] = mk2(pe.v.x, C.pO[k].v.x);
                    pan[k] = dext(db + (3 + k) * DL, bz);
                }
#pragma unroll
                for (int k = 0; k < 6; ++k) tn[k] = mk2(td[k * PL], tu[k * PL]);
                f2 up[3], qp[3], jpp[3], um[3], qm[3], jpm[3];
                edge(unp, pnp, up, qp, jpp, true, m_yp);
                edge(unm, pnm, um, qm, jpm, true, m_ym);
                RotM2 rn; rn.set(tn);
                f2 Mn[3][3]; rn.combine(pan[0], pan[1], pan[2], Mn);
                const f2 ue[3] = {mk2(um[0].v.x, up[0].v.y), mk2(um[1].v.x, up[1].v.y), mk2(um[2].v.x, up[2].v.y)};
#pragma unroll
                for (int k = 0; k < 3; ++k) {
                    C.acc[k] = C.acc[k] - mk2(jpm[k].v.y, jpp[k].v.x);           // the partner row's jp toward me
                    const f2 qe = mk2(qm[k].v.x, qp[k].v.y);
                    // acc -= jp_ni, jp_ni = M_n u - q
                    const f2 t = nfma2(Mn[k][0], ue[0], nfma2(Mn[k][1], ue[1], nfma2(Mn[k][2], ue[2], qe)));
                    C.acc[k] = fma2(m_ext, t, C.acc[k]);
                }
            }
            // ---- z+: toward the plane above (its jp toward me arrives in the next step) -----------------
#pragma unroll
            for (int k = 0; k < 3; ++k) C.jn[k] = bc2(0.f);
            if (z + 1 < D) {
                const float* __restrict__ nx = acquire(z + 1);
                const f2 bn = plane_beta(z + 1);
                f2 un[3], pn[3], u[3], q[3];
#pragma unroll
                for (int k = 0; k < 3; ++k) { un[k] = tpair(nx + toff + (6 + k) * PL); pn[k] = dpair(nx + doff + k * DL, bn); }
                edge(un, pn, u, q, C.jn, false, bc2(1.f));
            }
            // ---- the angle block of this plane is complete ------------------------------------------
            {
                f2 oA[3];
                rm.project(S, oA);
                f2 dsum = bc2(0.f);
#pragma unroll
                for (int k = 0; k < 3; ++k) {
                    f2 o = -(wr2 * oA[k]);
                    if (LM) o = fma2(mk2(w0 ? cs[(3 + k) * 32] : 0.f, w1 ? cs[(3 + k) * 32 + orow] : 0.f), pA[k], o);
                    if (w0) os[(3 + k) * 32] = o.v.x;
                    if (w1) os[(3 + k) * 32 + orow] = o.v.y;
                    dsum = fma2(pA[k], o, dsum);
                }
                dot += (double)((w0 ? dsum.v.x : 0.f) + (w1 ? dsum.v.y : 0.f));
            }
            }
            release(z);
            if (last) return false;
            havePrev = true; writePrev = true;
            os += oplane; cs += oplane; ns += oplane;
            return true;
        };

        VolCarry A, B;
        if (havePrev) {                                               // jp of the plane below toward plane z0
            const float* __restrict__ sb = acquire(z0 - 1);
            const float* __restrict__ sa = acquire(z0);
            const float* __restrict__ tb = sb + toff; const float* __restrict__ db = sb + doff;
            f2 tr[6];
#pragma unroll
            for (int k = 0; k < 6; ++k) tr[k] = tpair(tb + k * PL);
            RotM2 rn; rn.set(tr);
            const f2 bb = plane_beta(z0 - 1), ba = plane_beta(z0);
            f2 Mn[3][3]; rn.combine(dpair(db + 3 * DL, bb), dpair(db + 4 * DL, bb), dpair(db + 5 * DL, bb), Mn);
            f2 u[3];
#pragma unroll
            for (int k = 0; k < 3; ++k) {
                B.uc[k] = tpair(tb + (6 + k) * PL); B.pO[k] = dpair(db + k * DL, bb);
                u[k] = B.uc[k] - tpair(sa + toff + (6 + k) * PL);
            }
#pragma unroll
            for (int k = 0; k < 3; ++k)
                B.jn[k] = nfma2(Mn[k][0], u[0], nfma2(Mn[k][1], u[1], nfma2(Mn[k][2], u[2], B.pO[k] - dpair(sa + doff + k * DL, ba))));
            release(z0 - 1);
        }
        for (int z = z0;; z += 2) {                                   // roles alternate: no register shuffling
            if (!step(B, A, z)) break;
            if (!step(A, B, z + 1)) break;
        }
        e0 += ze - zs + 1;
    }
    block_accumulate(dot, den);
}

template <class T, int UB> size_t VolumetricFunctor<T, UB>::fast_smem() const { return UB > 0 ? VolRing::SMEM : 0; }

template <class T, int UB> template <bool FUSED, bool LM> auto VolumetricFunctor<T, UB>::fast_kernel() const {
    if constexpr (UB > 0) return &k_volumetric_apply_packed<FUSED, LM>;
    else return &k_volumetric_apply<T, FUSED, LM>;
}

template <class T> using VolumetricPackedFunctor = VolumetricFunctor<T, 6>;
template <class T> using VolumetricScalarFunctor = VolumetricFunctor<T, 0>;

}  // namespace optk
