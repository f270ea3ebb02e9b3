// Filter gradients of con_/cont_ (cpu_dlib.cpp:2327-2355, layers.h:222,547) on tcgen05:
//
//     G[tap][y][x] = sum_pos DY[pos][y] * X[pos + dy[tap]*P + dx[tap]][x]            (fp32 accumulate in TMEM)
//
// The contraction runs over POSITIONS, so both operands are [position][channel] tiles: MN-major operands.  TMA stages
// them with the 128-byte swizzle (box = 64 channels x rows) and the MMAs read them through MN-major shared-memory
// descriptors (SBO = 1024: the next 8 positions; LBO = distance to the next 64 rows of M) -- layout and the legality of
// arbitrary row offsets measured with experiments/umma_mn_probe.cu.
//   * A = X tile (M = channels of X), B = DY tile (N = channels of DY, padded to 16).
//   * A tap shift is a row offset into the staged halo segment (as in the forward kernel); TWO taps share one M = 128
//     MMA: rows 0-63 are the 64 X channels at tap t1, rows 64-127 the same channels at tap t2, reached through
//     LBO = (offset(t2) - offset(t1)) * 128 bytes.  Nine taps = five MMAs per 16 positions.
//   * Shift mode (3x3 stencils, <= 32 DY channels): the three dx taps ride on B instead.  DY is staged once per run as a
//     [kw + 8][16 | 32] tile (32- or 64-byte rows, matching swizzle); its N = 48 | 96 operand is three atoms ONE ROW apart
//     (LBO = one row: shift s reads DY[p + s]), against X at (dy, dx_max): G[dy][dx_max - s] for s = 0, 1, 2 from one MMA,
//     two MMAs (dy pair + single dy) per 16 positions instead of five.
//   * A CTA owns one 64-channel tile of X, all taps, and a contiguous range of position chunks (split-K); partial sums
//     leave TMEM once, at the end, to scratch[split][tile][tap][y][x]; wgrad_tc_reduce adds the splits and scatters
//     into the gradient arena in dlib's filter layout.
//   warp 0: TMA producer, warp 1: TMEM allocation + MMA issue, warps 2-5: epilogue.
#include "kcommon.cuh"
#include "tc.h"

#include <algorithm>
#include <stdexcept>

namespace ab {
namespace tc {
namespace {

constexpr int kWThreads = 192;
constexpr int kWMaxStages = 4;

__device__ __forceinline__ uint32_t smem_u32(const void* p) { return (uint32_t)__cvta_generic_to_shared(p); }
__device__ __forceinline__ void mbar_init(uint64_t* b, int n) { asm volatile("mbarrier.init.shared::cta.b64 [%0], %1;" ::"r"(smem_u32(b)), "r"(n)); }
__device__ __forceinline__ void mbar_expect(uint64_t* b, uint32_t tx)
{
    asm volatile("mbarrier.arrive.expect_tx.shared::cta.b64 _, [%0], %1;" ::"r"(smem_u32(b)), "r"(tx) : "memory");
}
__device__ __forceinline__ void mbar_wait(uint64_t* b, uint32_t parity)
{
    asm volatile(
        "{\n"
        ".reg .pred p;\n"
        "WG_WAIT:\n"
        "mbarrier.try_wait.parity.shared::cta.b64 p, [%0], %1;\n"
        "@p bra WG_DONE;\n"
        "bra WG_WAIT;\n"
        "WG_DONE:\n"
        "}" ::"r"(smem_u32(b)), "r"(parity) : "memory");
}
__device__ __forceinline__ void tma_2d(uint32_t dst, const CUtensorMap* m, int c0, int c1, uint64_t* bar)
{
    asm volatile("cp.async.bulk.tensor.2d.shared::cluster.global.mbarrier::complete_tx::bytes [%0], [%1, {%2, %3}], [%4];"
                 ::"r"(dst), "l"(m), "r"(c0), "r"(c1), "r"(smem_u32(bar)) : "memory");
}
__device__ __forceinline__ bool elect_one()
{
    uint32_t pred;
    asm volatile(
        "{\n"
        ".reg .pred p;\n"
        "elect.sync _|p, 0xffffffff;\n"
        "selp.u32 %0, 1, 0, p;\n"
        "}" : "=r"(pred));
    return pred != 0;
}
__device__ __forceinline__ void umma_commit(uint64_t* bar)
{
    asm volatile("tcgen05.commit.cta_group::1.mbarrier::arrive::one.shared::cluster.b64 [%0];" ::"r"(smem_u32(bar)) : "memory");
}
__device__ __forceinline__ void umma_bf16(uint32_t tmem_d, uint32_t a_lo, uint32_t b_lo, uint32_t desc_hi, uint32_t b_hi, uint32_t idesc, uint32_t accumulate)
{
    asm volatile(
        "{\n"
        ".reg .pred p;\n"
        ".reg .b64 da, db;\n"
        "setp.ne.b32 p, %6, 0;\n"
        "mov.b64 da, {%1, %3};\n"
        "mov.b64 db, {%2, %4};\n"
        "tcgen05.mma.cta_group::1.kind::f16 [%0], da, db, %5, p;\n"
        "}" ::"r"(tmem_d), "r"(a_lo), "r"(b_lo), "r"(desc_hi), "r"(b_hi), "r"(idesc), "r"(accumulate) : "memory");
}
__device__ __forceinline__ void tmem_ld16(uint32_t taddr, float v[16])
{
    asm volatile("tcgen05.ld.sync.aligned.32x32b.x16.b32 {%0,%1,%2,%3,%4,%5,%6,%7,%8,%9,%10,%11,%12,%13,%14,%15}, [%16];"
                 : "=f"(v[0]), "=f"(v[1]), "=f"(v[2]), "=f"(v[3]), "=f"(v[4]), "=f"(v[5]), "=f"(v[6]), "=f"(v[7]), "=f"(v[8]),
                   "=f"(v[9]), "=f"(v[10]), "=f"(v[11]), "=f"(v[12]), "=f"(v[13]), "=f"(v[14]), "=f"(v[15])
                 : "r"(taddr));
    asm volatile("tcgen05.wait::ld.sync.aligned;" ::: "memory");
    asm volatile("" : "+f"(v[0]), "+f"(v[1]), "+f"(v[2]), "+f"(v[3]), "+f"(v[4]), "+f"(v[5]), "+f"(v[6]), "+f"(v[7]), "+f"(v[8]),
                      "+f"(v[9]), "+f"(v[10]), "+f"(v[11]), "+f"(v[12]), "+f"(v[13]), "+f"(v[14]), "+f"(v[15]) :: "memory");
}

__global__ void __launch_bounds__(kWThreads, 1) wgrad_tc_kernel(const __grid_constant__ WgradTcLaunch L)
{
    extern __shared__ __align__(1024) uint8_t smem_raw[];
    __shared__ uint64_t full_bar[kWMaxStages], empty_bar[kWMaxStages], acc_full;
    __shared__ uint32_t tmem_base_slot;

    const uint32_t smem0 = (smem_u32(smem_raw) + 1023u) & ~1023u;
    const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
    const int split = blockIdx.x, cxt = blockIdx.y;
    const long long c_begin = L.chunks * split / L.nsplit, c_end = L.chunks * (split + 1) / L.nsplit;

    if (threadIdx.x == 0) {
        for (int s = 0; s < L.stages; ++s) { mbar_init(&full_bar[s], 1); mbar_init(&empty_bar[s], 1); }
        mbar_init(&acc_full, 1);
        asm volatile("fence.mbarrier_init.release.cluster;" ::: "memory");
    }
    if (warp == 1) {
        asm volatile("tcgen05.alloc.cta_group::1.sync.aligned.shared::cta.b32 [%0], %1;" ::"r"(smem_u32(&tmem_base_slot)), "r"(L.tmem_cols));
        asm volatile("tcgen05.relinquish_alloc_permit.cta_group::1.sync.aligned;");
    }
    asm volatile("tcgen05.fence::before_thread_sync;" ::: "memory");
    __syncthreads();
    asm volatile("tcgen05.fence::after_thread_sync;" ::: "memory");
    k::pdl_enter();   // shared memory and TMEM only above; the operands (and the scratch the previous reduce still reads) below
    const uint32_t tmem = tmem_base_slot;

    if (warp == 0) {
        if (elect_one()) {
            const uint32_t tx = L.x_stage_bytes + (uint32_t)L.th * (L.shift ? (uint32_t)(L.kw + 8) * (uint32_t)(2 * L.N) : (uint32_t)L.kw * 128u);
            uint32_t it = 0;
            for (long long chunk = c_begin; chunk < c_end; ++chunk, ++it) {
                const int s = it % L.stages;
                mbar_wait(&empty_bar[s], ((it / L.stages) & 1) ^ 1);
                mbar_expect(&full_bar[s], tx);
                int o0;
                if (L.rows_mode) {
                    const int rg = (int)(chunk / L.ncx);
                    o0 = rg * L.th * L.P + (int)(chunk - (long long)rg * L.ncx) * L.kw;
                } else o0 = (int)chunk * L.kw;
                const uint32_t xs = smem0 + (uint32_t)s * L.stage_bytes, ys = xs + L.x_stage_bytes;
                for (int sg = 0; sg < L.nseg; ++sg)
                    tma_2d(xs + (uint32_t)(sg * L.seg_alloc) * 128u, &L.tmX, L.x_c0 + cxt * 64, o0 + L.seg_pos[sg], &full_bar[s]);
                for (int m = 0; m < L.th; ++m) {
                    if (L.shift) tma_2d(ys + (uint32_t)(m * (L.kw + 8)) * (uint32_t)(2 * L.N), &L.tmY, L.dy_c0, o0 + m * L.run_stride, &full_bar[s]);
                    else tma_2d(ys + (uint32_t)(m * L.kw) * 128u, &L.tmY, L.dy_c0, o0 + m * L.run_stride, &full_bar[s]);
                }
            }
        }
    } else if (warp == 1) {
        if (elect_one()) {
            // instruction descriptor: D f32, A/B fp16 (format 0), both MN-major (bits 15, 16), N, M = 128
            const uint32_t idesc = (1u << 4) | (1u << 15) | (1u << 16) | ((uint32_t)(L.Nmma >> 3) << 17) | ((128u >> 4) << 24);
            // descriptor high word: SBO = 1024 bytes (8 positions), version 1, 128-byte swizzle
            const uint32_t desc_hi = (1024u >> 4) | (1u << 14) | (2u << 29);
            // B in shift mode: SBO = 256 bytes (8 positions of 32 bytes), 32-byte swizzle
            // (N = 16: 32-byte rows and swizzle, layout type 6;  N = 32: 64-byte rows and swizzle, layout type 4;  SBO = 8 rows)
            const uint32_t b_hi = L.shift ? ((uint32_t)(16 * L.N) >> 4) | (1u << 14) | ((L.N == 16 ? 6u : 4u) << 29) : desc_hi;
            const uint32_t row16 = (uint32_t)(2 * L.N) >> 4;   // shift mode: 16-byte units per staged DY row
            const int ksteps = L.kw / 16, npairs = L.npairs;
            uint32_t pa_off[5], pa_lbo[5];
            for (int a = 0; a < 5; ++a) {
                pa_off[a] = (uint32_t)(L.pair_seg[a] * L.seg_alloc + L.pair_off[a]) * 8u;   // in 16-byte units (one position = 128 bytes)
                pa_lbo[a] = ((uint32_t)L.pair_lbo[a] >> 4) << 16;
            }
            uint32_t it = 0;
            for (long long chunk = c_begin; chunk < c_end; ++chunk, ++it) {
                const int s = it % L.stages;
                mbar_wait(&full_bar[s], (it / L.stages) & 1);
                asm volatile("tcgen05.fence::after_thread_sync;" ::: "memory");
                const uint32_t xs16 = (smem0 + (uint32_t)s * L.stage_bytes) >> 4, ys16 = xs16 + (L.x_stage_bytes >> 4);
                for (int m = 0; m < L.th; ++m)
                    for (int kk = 0; kk < ksteps; ++kk) {
                        // shift mode: three N-channel atoms (one per shift) ONE ROW apart (LBO = one row) in the same staged tile;
                        // the swizzle works on absolute shared-memory addresses, so a row offset keeps matching what TMA wrote
                        const uint32_t b_lo = L.shift ? (ys16 + (uint32_t)(m * (L.kw + 8) + kk * 16) * row16) | (row16 << 16)
                                                      : (ys16 + (uint32_t)(m * L.kw + kk * 16) * 8u) | (1u << 16);
                        const uint32_t a_base = xs16 + (uint32_t)(m * L.seg_alloc + kk * 16) * 8u;
                        const uint32_t acc = (uint32_t)(it | m | kk);
#pragma unroll
                        for (int a = 0; a < 5; ++a)
                            if (a < npairs) umma_bf16(tmem + (uint32_t)(a * L.Nmma), (a_base + pa_off[a]) | pa_lbo[a], b_lo, desc_hi, b_hi, idesc, acc);
                    }
                umma_commit(&empty_bar[s]);
            }
            umma_commit(&acc_full);
        }
    } else {
        // ---- epilogue: thread <-> TMEM lane = X channel (rows 0-63: tap t1 of the pair, rows 64-127: tap t2)
        const int q = warp & 3;
        mbar_wait(&acc_full, 0);
        asm volatile("tcgen05.fence::after_thread_sync;" ::: "memory");
        const int x = (q & 1) * 32 + lane;
        float* sc = L.scratch + ((size_t)split * L.cxtiles + cxt) * L.ntaps * 4096;
        for (int a = 0; a < L.npairs; ++a) {
            const int tap = q < 2 ? L.pair_t1[a] : L.pair_t2[a];
            for (int j = 0; j < L.Nmma; j += 16) {
                float v[16];
                tmem_ld16(tmem + ((uint32_t)(q * 32) << 16) + (uint32_t)(a * L.Nmma + j), v);
                // shift mode: column block j / 16 is the dx tap of this row half's dy; else 16 more DY channels of `tap`
                const int tp = L.shift ? L.sh_tap[a][q >> 1][j / L.N] : tap;
                if (tp >= 0) {
                    float* o = sc + (size_t)tp * 4096 + (size_t)(L.shift ? j % L.N : j) * 64 + x;
#pragma unroll
                    for (int i = 0; i < 16; ++i) o[i * 64] = v[i];
                }
            }
        }
    }
    asm volatile("tcgen05.fence::before_thread_sync;" ::: "memory");
    __syncthreads();
    if (warp == 1) asm volatile("tcgen05.dealloc.cta_group::1.sync.aligned.b32 %0, %1;" ::"r"(tmem), "r"(L.tmem_cols));
}

// out[y*sy + x*sx + tapmap[t]] = sum over splits of scratch[split][tile][t][y][x];  block = 32 outputs x kRL split lanes
template <int kRL>
__global__ void __launch_bounds__(32 * kRL) wgrad_tc_reduce_kernel(const float* __restrict__ scratch, int nsplit, int cxtiles, int ntaps, int cy, int cx,
                                                                   float* __restrict__ out, long long sy, long long sx, float unscale, int t0,
                                                                   int t1, int t2, int t3, int t4, int t5, int t6, int t7, int t8)
{
    k::pdl_enter();
    const int tapmap[9] = {t0, t1, t2, t3, t4, t5, t6, t7, t8};
    __shared__ float sh[kRL][33];
    const int ox = threadIdx.x & 31, sl = threadIdx.x >> 5;
    const long long total = (long long)ntaps * cy * cx;
    const long long i = (long long)blockIdx.x * 32 + ox;
    float s = 0.f;
    int xg = 0, y = 0, t = 0;
    if (i < total) {
        xg = (int)(i % cx);
        y = (int)((i / cx) % cy);
        t = (int)(i / ((long long)cx * cy));
        const int cxt = xg >> 6, x = xg & 63;
        const float* p = scratch + ((size_t)cxt * ntaps + t) * 4096 + y * 64 + x;
        const size_t stride = (size_t)cxtiles * ntaps * 4096;
#pragma unroll 5
        for (int sp = sl; sp < nsplit; sp += kRL) s += p[sp * stride];
    }
    sh[sl][ox] = s;
    __syncthreads();
    if (sl == 0 && i < total) {
        float r = 0.f;
#pragma unroll
        for (int k = 0; k < kRL; ++k) r += sh[k][ox];
        out[y * sy + xg * sx + tapmap[t]] = r * unscale;
    }
}

}  // namespace

// scratch[split][tile][tap][64][64]: at most num_sms() (split, tile) pairs, 9 taps
size_t wgrad_tc_scratch_floats() { return (size_t)k::num_sms() * 9 * 4096; }

void setup_wgrad_tc(WgradTcLaunch& L, const WgradJob& job, float* scratch)
{
    if (job.cy % 8 || job.cx % 8 || job.dy_ctot % 8 || job.x_ctot % 8 || job.dy_c0 % 8 || job.x_c0 % 8)
        throw std::invalid_argument("wgrad_tc needs channel counts that are multiples of 8");
    if (job.cy > 64 || job.cy < 8) throw std::invalid_argument("wgrad_tc: 8 <= cy <= 64");
    if (job.ntaps < 1 || job.ntaps > 9) throw std::invalid_argument("wgrad_tc: 1..9 taps");
    if (job.g.positions() + 4096 >= (1ll << 31)) throw std::invalid_argument("tensor too large for the wgrad kernel (positions must fit 31 bits)");
    L = WgradTcLaunch();
    const int P = job.g.P();
    const long long total = (long long)job.g.N * job.g.S();
    L.P = P;
    L.x_c0 = job.x_c0; L.dy_c0 = job.dy_c0;
    L.ntaps = job.ntaps;
    L.N = padded_n(job.cy);
    for (int t = 0; t < job.ntaps; ++t) L.tapmap[t] = job.tapmap[t];
    // shift mode: a full 3x3 stencil with at most 16 output channels; the A side then sees three taps (dy, +1) only
    static const bool no_shift = getenv("AB_WGRAD_NOSHIFT") != nullptr;
    int grid3[3][3];
    for (auto& r : grid3) for (int& v : r) v = -1;
    bool full3 = job.ntaps == 9 && (L.N == 16 || L.N == 32) && !no_shift;
    int dy0 = 1 << 30, dx0 = 1 << 30;   // corner of the 3 x 3 stencil (-1, -1 for con_, -2, -2 for the zero-stuffed cont_)
    for (int t = 0; t < job.ntaps; ++t) { dy0 = std::min(dy0, job.tap_dy[t]); dx0 = std::min(dx0, job.tap_dx[t]); }
    for (int t = 0; t < job.ntaps && full3; ++t) {
        const int a = job.tap_dy[t] - dy0, b = job.tap_dx[t] - dx0;
        if (a > 2 || b > 2 || grid3[a][b] >= 0) full3 = false;
        else grid3[a][b] = t;
    }
    int ntaps = job.ntaps, tdy[9], tdx[9];
    for (int t = 0; t < job.ntaps; ++t) { tdy[t] = job.tap_dy[t]; tdx[t] = job.tap_dx[t]; }
    if (full3) {   // A sees (dy, dx_max) only; shift s of DY turns it into dx_max - s
        ntaps = 3;
        for (int t = 0; t < 3; ++t) { tdy[t] = dy0 + t; tdx[t] = dx0 + 2; }
    }
    int dylo = 0, dyhi = 0, dxlo = 0, dxhi = 0;
    for (int t = 0; t < ntaps; ++t) {
        dylo = std::min(dylo, tdy[t]); dyhi = std::max(dyhi, tdy[t]);
        dxlo = std::min(dxlo, tdx[t]); dxhi = std::max(dxhi, tdx[t]);
    }
    const int dyr = dyhi - dylo, dxr = dxhi - dxlo;
    // chunk geometry: runs of kw positions.  When kw divides the row pitch, `th` consecutive image rows share their halo
    // segments (ROWS); otherwise a run is kw consecutive flat positions with one segment per dy (th = 1).
    // Candidates are scored by the rows TMA has to stage per position (halo amplification; ~2.4 cycles per 128-byte row with
    // every SM loading, experiments/tma_bench.cu) with a preference for a ring of at least three stages.
    const size_t budget = 200u * 1024u;
    double best = 1e300;
    auto consider = [&](int rows_mode, int kw, int th) {
        const int nseg = th + dyr, seg_alloc = (kw + dxr + 7) / 8 * 8;
        const size_t xb = (size_t)nseg * seg_alloc * 128, sb = (xb + (size_t)th * (full3 ? (kw + 8) * 2 * L.N : kw * 128) + 1023) / 1024 * 1024;
        if (nseg > 12 || seg_alloc > 256 || 2 * sb + 1024 > budget) return;
        const int stages = (int)std::min<size_t>(kWMaxStages, (budget - 1024) / sb);
        const double score = (double)(nseg * seg_alloc + th * (full3 ? kw + 8 : kw)) / (th * kw) * (stages >= 3 ? 1.0 : 1.15);
        if (score < best) {
            best = score;
            L.rows_mode = rows_mode; L.kw = kw; L.th = th; L.nseg = nseg; L.seg_alloc = seg_alloc;
            L.x_stage_bytes = (unsigned)xb; L.stage_bytes = (unsigned)sb;
        }
    };
    for (int kw : {128, 96, 64, 48, 32})
        if (P % kw == 0)
            for (int th : {8, 4, 2, 1}) consider(1, kw, th);
    if (best == 1e300) consider(0, (int)std::min<long long>(128, (total + 15) / 16 * 16), 1);
    if (best == 1e300) throw std::logic_error("wgrad_tc: tile does not fit shared memory");
    L.stages = (int)std::min<size_t>(kWMaxStages, (200u * 1024u - 1024) / L.stage_bytes);
    L.smem_bytes = L.stages * L.stage_bytes + 1024;
    L.run_stride = L.rows_mode ? P : 0;
    L.ncx = L.rows_mode ? P / L.kw : 1;
    for (int s = 0; s < L.nseg; ++s) L.seg_pos[s] = (s + dylo) * P + dxlo;
    if (L.rows_mode) {
        const long long rows_total = (long long)job.g.N * (job.g.H + 1);
        L.chunks = (rows_total + L.th - 1) / L.th * L.ncx;
    } else L.chunks = (total + L.kw - 1) / L.kw;
    // taps sorted by their address inside the staged segments, then paired
    int order[9], key[9];
    for (int t = 0; t < ntaps; ++t) { order[t] = t; key[t] = (tdy[t] - dylo) * L.seg_alloc + (tdx[t] - dxlo); }
    std::sort(order, order + ntaps, [&](int a, int b) { return key[a] < key[b]; });
    L.npairs = (ntaps + 1) / 2;
    for (int a = 0; a < L.npairs; ++a) {
        const int t1 = order[2 * a], t2 = 2 * a + 1 < ntaps ? order[2 * a + 1] : -1;
        L.pair_t1[a] = t1; L.pair_t2[a] = t2;
        L.pair_seg[a] = tdy[t1] - dylo; L.pair_off[a] = tdx[t1] - dxlo;
        L.pair_lbo[a] = t2 >= 0 ? (key[t2] - key[t1]) * 128 : 1024;
        if (L.pair_lbo[a] <= 0 || L.pair_lbo[a] >= (1 << 18)) throw std::logic_error("wgrad_tc: tap pair distance out of range");
        if (full3)   // position p of DY at shift s pairs with X[p - s + dy*P + dx_max]: dx = dx_max - s
            for (int h = 0; h < 2; ++h)
                for (int sft = 0; sft < 3; ++sft) {
                    const int ts = h ? t2 : t1;
                    L.sh_tap[a][h][sft] = ts >= 0 ? grid3[tdy[ts] - dy0][2 - sft] : -1;
                }
    }
    L.shift = full3 ? 1 : 0;
    L.Nmma = full3 ? 3 * L.N : L.N;
    int cols = 32;
    while (cols < L.npairs * L.Nmma) cols *= 2;
    L.tmem_cols = cols;
    L.cxtiles = (job.cx + 63) / 64;
    L.nsplit = (int)std::max<long long>(1, std::min<long long>(L.chunks, k::num_sms() / L.cxtiles));
    L.scratch = scratch;
    L.out = job.out; L.sy = job.sy; L.sx = job.sx; L.unscale = job.unscale;
    L.cy_real = job.cy_real ? job.cy_real : job.cy; L.cx_real = job.cx_real ? job.cx_real : job.cx;
    make_tensor_map_2d(&L.tmX, job.x, job.g.positions(), job.x_ctot, 64, L.seg_alloc);
    // shift mode: one [kw + 8][16] tile per run (32-byte rows, 32-byte swizzle); the three shifts are row offsets into it
    make_tensor_map_2d(&L.tmY, job.dy, job.g.positions(), job.dy_ctot, L.shift ? L.N : 64, L.shift ? L.kw + 8 : L.kw);
    const double px = (double)job.g.N * job.g.H * job.g.W;
    L.flops = 2.0 * px * L.cx_real * L.cy_real * job.ntaps;
    L.bytes = 2.0 * px * (job.cx + job.cy);
}

void launch_wgrad_tc(const WgradTcLaunch& L, cudaStream_t st)
{
    static bool attr_set = false;
    if (!attr_set) {
        cudaFuncSetAttribute(wgrad_tc_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, 220 * 1024);
        attr_set = true;
    }
    k::ProfScope ps(k::F_WGRAD_TC, L.flops, L.bytes, st);
    k::launch_pdl(wgrad_tc_kernel, dim3(L.nsplit, L.cxtiles), dim3(kWThreads), (size_t)L.smem_bytes, st, L);
    const long long total = (long long)L.ntaps * L.cy_real * L.cx_real;
    static const int rl = getenv("AB_RL") ? atoi(getenv("AB_RL")) : 4;   // split lanes per 32 outputs: 4 measured best (2: 7.88, 4: 7.85, 8: 7.88, 16: 7.98, 32: 8.47 ms per step)
#define AB_REDUCE(R) k::launch_pdl(wgrad_tc_reduce_kernel<R>, dim3((unsigned)((total + 31) / 32)), dim3(32 * R), 0, st, L.scratch, L.nsplit, L.cxtiles, L.ntaps, L.cy_real, L.cx_real, L.out, \
        L.sy, L.sx, L.unscale, L.tapmap[0], L.tapmap[1], L.tapmap[2], L.tapmap[3], L.tapmap[4], L.tapmap[5], L.tapmap[6], L.tapmap[7], L.tapmap[8])
    if (rl == 2) AB_REDUCE(2); else if (rl == 16) AB_REDUCE(16); else if (rl == 8) AB_REDUCE(8); else AB_REDUCE(4);
#undef AB_REDUCE
    k::count_launch(2);
}

}  // namespace tc
}  // namespace ab
