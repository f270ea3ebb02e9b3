// fp16 tcgen05 / TMEM / TMA implicit-GEMM convolution (AB_PRECISION_F16): con_ / cont_ forward and
// backward-data (dnn/layers.h:196-229,521-554; cpu_dlib.cpp:2250-2325) as ONE shifted-window GEMM kernel
// over the padded flat NHWC layout described in tc.h:
//
//     D[pos][n] = sum_{tap, k} A[pos + dy[tap]*P + dx[tap]][a_c0 + k] * B[tap][n][k]
//
// A CTA owns `th` consecutive 128-position M tiles and all (or a <=256 wide block of) the N columns.
// Per K chunk of kc channels one halo segment of 128*th + (hi - lo) positions is TMA-loaded once and
// serves every tap of every M tile through row-offset shared-memory descriptors.
//   warp 0  : TMA producer (one elected lane)          full/empty mbarrier ring
//   warp 1  : TMEM allocation + tcgen05.mma issue (one elected lane), tcgen05.commit releases stages
//   warps 2-5: epilogue, tcgen05.ld of the fp32 accumulators -> bias / mish / affine -> global
#include "h16.cuh"
#include "kcommon.cuh"
#include "tc.h"

#include <cstdio>
#include <cstdlib>
#include <stdexcept>
#include <string>

namespace ab {
namespace tc {
namespace {

__device__ __forceinline__ uint32_t smem_u32(const void* p) { return (uint32_t)__cvta_generic_to_shared(p); }
__device__ __forceinline__ void mbar_init(uint64_t* b, int n)
{
    asm volatile("mbarrier.init.shared::cta.b64 [%0], %1;" ::"r"(smem_u32(b)), "r"(n));
}
__device__ __forceinline__ void mbar_expect(uint64_t* b, uint32_t tx)
{
    asm volatile("mbarrier.arrive.expect_tx.shared::cta.b64 _, [%0], %1;" ::"r"(smem_u32(b)), "r"(tx) : "memory");
}
__device__ __forceinline__ void mbar_wait(uint64_t* b, uint32_t parity)
{
    asm volatile(
        "{\n"
        ".reg .pred p;\n"
        "WAIT_LOOP:\n"
        "mbarrier.try_wait.parity.shared::cta.b64 p, [%0], %1;\n"
        "@p bra WAIT_DONE;\n"
        "bra WAIT_LOOP;\n"
        "WAIT_DONE:\n"
        "}" ::"r"(smem_u32(b)), "r"(parity) : "memory");
}
__device__ __forceinline__ void tma_2d(uint32_t dst, const CUtensorMap* m, int c0, int c1, uint64_t* bar)
{
    asm volatile("cp.async.bulk.tensor.2d.shared::cluster.global.mbarrier::complete_tx::bytes [%0], [%1, {%2, %3}], [%4];"
                 ::"r"(dst), "l"(m), "r"(c0), "r"(c1), "r"(smem_u32(bar)) : "memory");
}
// true in exactly one lane of the (converged) warp; ptxas treats the guarded region as single-threaded, so uniform-datapath
// operands (tcgen05 / TMA descriptors) need no per-lane "waterfall" loop
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
// 16 consecutive fp32 accumulator columns of this thread's TMEM lane; results are valid after tmem_ld_wait()
__device__ __forceinline__ void tmem_ld16_issue(uint32_t taddr, float v[16])
{
    asm volatile("tcgen05.ld.sync.aligned.32x32b.x16.b32 {%0,%1,%2,%3,%4,%5,%6,%7,%8,%9,%10,%11,%12,%13,%14,%15}, [%16];"
                 : "=f"(v[0]), "=f"(v[1]), "=f"(v[2]), "=f"(v[3]), "=f"(v[4]), "=f"(v[5]), "=f"(v[6]), "=f"(v[7]), "=f"(v[8]),
                   "=f"(v[9]), "=f"(v[10]), "=f"(v[11]), "=f"(v[12]), "=f"(v[13]), "=f"(v[14]), "=f"(v[15])
                 : "r"(taddr));
}
__device__ __forceinline__ void tmem_ld_wait() { asm volatile("tcgen05.wait::ld.sync.aligned;" ::: "memory"); }
// ties the loaded registers to the wait so that no use of them can be scheduled ahead of it
__device__ __forceinline__ void tmem_ld_fence(float v[16])
{
    asm volatile("" : "+f"(v[0]), "+f"(v[1]), "+f"(v[2]), "+f"(v[3]), "+f"(v[4]), "+f"(v[5]), "+f"(v[6]), "+f"(v[7]), "+f"(v[8]),
                      "+f"(v[9]), "+f"(v[10]), "+f"(v[11]), "+f"(v[12]), "+f"(v[13]), "+f"(v[14]), "+f"(v[15]) :: "memory");
}

// K-major shared-memory matrix descriptor (cute::UMMA::SmemDescriptor): start address >> 4, LBO (unused for
// swizzled K-major) = 1, SBO = 8 rows * pitch, version 1, layout type 6/4/2 for 32/64/128-byte swizzle.
__device__ __forceinline__ uint64_t smem_desc_hi(int pitch)
{
    const uint64_t sbo = (uint64_t)(8 * pitch) >> 4;
    const uint64_t layout = pitch == 128 ? 2 : pitch == 64 ? 4 : 6;
    return (sbo << 32) | (1ull << 46) | (layout << 61) | (1ull << 16);
}

constexpr int kEpiWarps = 8;
constexpr int kEpiThreads = 32 * kEpiWarps;
constexpr int kThreads = 64 + kEpiThreads;   // + TMA producer warp + MMA issuer warp
constexpr int kMaxStages = 8;

// D[tmem] (+)= A[smem desc] * B[smem desc]; descriptors assembled from 32-bit halves (hi is loop-invariant)
__device__ __forceinline__ void umma_bf16(uint32_t tmem_d, uint32_t a_lo, uint32_t b_lo, uint32_t desc_hi, uint32_t idesc, uint32_t accumulate)
{
    asm volatile(
        "{\n"
        ".reg .pred p;\n"
        ".reg .b64 da, db;\n"
        "setp.ne.b32 p, %5, 0;\n"
        "mov.b64 da, {%1, %3};\n"
        "mov.b64 db, {%2, %3};\n"
        "tcgen05.mma.cta_group::1.kind::f16 [%0], da, db, %4, p;\n"
        "}" ::"r"(tmem_d), "r"(a_lo), "r"(b_lo), "r"(desc_hi), "r"(idesc), "r"(accumulate) : "memory");
}

// mish for the h16 path: x - 2x / (e^2 + 2e + 2), e = exp(x), on the raw MUFU approximations (results are rounded
// to h16 anyway).  e = inf -> rcp = 0 -> x;  e = 0 -> x - x = 0.
__device__ __forceinline__ float mish_fast(float x)
{
    float e, rc;
    asm("ex2.approx.ftz.f32 %0, %1;" : "=f"(e) : "f"(x * 1.4426950408889634f));
    asm("rcp.approx.ftz.f32 %0, %1;" : "=f"(rc) : "f"(fmaf(e, e + 2.f, 2.f)));
    return fmaf(-2.f * x, rc, x);
}

// Persistent, warp-specialised: one CTA per SM walks work items (a group of `th` M tiles x one N block) with a
// shared-memory stage ring (TMA -> MMA) and two TMEM accumulator buffers (MMA -> epilogue), so the TMA loads, the
// tensor-core work and the epilogue of consecutive items overlap.
// KS: K steps (of 16 channels) per chunk, compile-time (kc / 16) so that the MMA issue loop of a tile is straight-line code --
// one thread issues every tcgen05.mma of the CTA, and with run-time trip counts it spent ~140 cycles per MMA on branches and
// vector-to-uniform register moves against the ~40 the tensor pipe needs (ncu source view: 80 % of the issuer warp's samples
// inside the issue loop).  KS = 0: run-time K steps (the rare tap counts).
template <int NTAPS, int KS>
__global__ void __launch_bounds__(kThreads, 1) conv_tc_kernel(const __grid_constant__ ConvLaunch L, const int rev)
{
    extern __shared__ __align__(1024) uint8_t smem_raw[];
    __shared__ uint64_t full_bar[kMaxStages], empty_bar[kMaxStages], acc_full[2], acc_empty[2];
    __shared__ uint32_t tmem_base_slot;
    __shared__ __align__(16) float s_bias[512], s_scale[512], s_shift[512];
    __shared__ double s_stat[32];

    const uint32_t smem0 = (smem_u32(smem_raw) + 1023u) & ~1023u;
    const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
    const int pitch = L.kc * 2;
    const int grid_x = (int)L.grid_x, items = grid_x * (L.npad_total / L.nblk);   // (host checks that positions fit 31 bits)
    const int nbuf = L.nbuf;

    if (threadIdx.x < 32) s_stat[threadIdx.x] = 0.0;
    if (threadIdx.x == 0) {
        for (int s = 0; s < L.stages; ++s) { mbar_init(&full_bar[s], 1); mbar_init(&empty_bar[s], 1); }
        for (int b = 0; b < 2; ++b) { mbar_init(&acc_full[b], 1); mbar_init(&acc_empty[b], kEpiThreads); }
        asm volatile("fence.mbarrier_init.release.cluster;" ::: "memory");
    }
    if (warp == 1) {
        asm volatile("tcgen05.alloc.cta_group::1.sync.aligned.shared::cta.b32 [%0], %1;" ::"r"(smem_u32(&tmem_base_slot)), "r"(L.tmem_cols));
        asm volatile("tcgen05.relinquish_alloc_permit.cta_group::1.sync.aligned;");
    }
    if (warp == 0 && lane == 0) {   // the TMA descriptors live in the kernel parameters: fetch them while the predecessor drains
        asm volatile("prefetch.tensormap [%0];" ::"l"(reinterpret_cast<uint64_t>(&L.tmA)) : "memory");
        asm volatile("prefetch.tensormap [%0];" ::"l"(reinterpret_cast<uint64_t>(&L.tmB)) : "memory");
    }
    asm volatile("tcgen05.fence::before_thread_sync;" ::: "memory");
    __syncthreads();
    asm volatile("tcgen05.fence::after_thread_sync;" ::: "memory");
    // everything above touched only shared memory, TMEM and the kernel parameters: from here on the predecessor's results are
    // needed (kcommon.cuh).  The TMA and MMA warps start at once; the epilogue warps stage their per-channel parameters behind
    // a barrier of their own.
    k::pdl_enter();
    if (warp >= 2) {  // per-channel epilogue parameters
        for (int co = threadIdx.x - 64; co < L.npad_total; co += kEpiThreads) {
            const bool in = co < L.cout;
            s_bias[co] = (in && L.bias) ? __ldg(L.bias + co) : 0.f;
            s_scale[co] = (in && L.act != ACT_NONE) ? __ldg(L.scale + co) : 1.f;
            s_shift[co] = (in && L.act != ACT_NONE) ? __ldg(L.shift + co) : 0.f;
        }
        asm volatile("bar.sync 1, %0;" ::"n"(kEpiThreads) : "memory");
    }
    const uint32_t tmem = tmem_base_slot;

    if (warp == 0) {
        if (elect_one()) {
            const uint32_t tx = (uint32_t)(L.nseg * L.seg_alloc + NTAPS * L.nblk) * pitch;
            uint32_t it = 0;
            for (int it0 = blockIdx.x; it0 < items; it0 += gridDim.x) {
                const int item = rev ? items - 1 - it0 : it0;
                const int by = item / grid_x, bx = item - by * grid_x;
                const int n0 = by * L.nblk;
                const int o0 = L.rows_mode ? (bx / L.ncx) * L.th * L.P + (bx % L.ncx) * 128 : bx * (128 * L.th);
                for (int ch = 0; ch < L.nchunks; ++ch, ++it) {
                    const int s = it % L.stages;
                    mbar_wait(&empty_bar[s], ((it / L.stages) & 1) ^ 1);
                    if (L.dbg & 2) {   // timing experiment (AB_TC_DBG): no loads, the stage is handed over empty
                        asm volatile("mbarrier.arrive.shared::cta.b64 _, [%0];" ::"r"(smem_u32(&full_bar[s])) : "memory");
                        continue;
                    }
                    mbar_expect(&full_bar[s], tx);
                    const uint32_t a_s = smem0 + (uint32_t)s * L.stage_bytes, b_s = a_s + L.a_stage_bytes;
                    // split operands: part 0 = the hi halves of A, part 1 = the lo halves, both against the same [W_hi | W_lo] (tc.h)
                    const int part = ch / L.nch1, c1 = ch - part * L.nch1;
                    const int c = L.a_c0 + c1 * L.kc;
                    const int arow = o0 + (part == 1 ? (int)L.a_lo_rows : 0);
                    for (int sg = 0; sg < L.nseg; ++sg)
                        for (int l = 0; l < L.nloads; ++l)
                            tma_2d(a_s + (uint32_t)(sg * L.seg_alloc + l * L.box_rows) * pitch, &L.tmA, c,
                                   arow + L.seg_pos[sg] + l * L.box_rows, &full_bar[s]);
                    const int bk = c1 * L.kc;
#pragma unroll
                    for (int t = 0; t < NTAPS; ++t)
                        tma_2d(b_s + (uint32_t)(t * L.nblk) * pitch, &L.tmB, bk, t * L.npad_total + n0, &full_bar[s]);
                }
            }
        }
    } else if (warp == 1) {
        if (elect_one()) {
            // instruction descriptor: D f32 (bit 4), A/B fp16 (format 0 in bits 7-9 / 10-12), both K-major, N = nblk, M = 128
            const uint32_t idesc = (1u << 4) | ((uint32_t)(L.nblk >> 3) << 17) | ((128u >> 4) << 24);
            const uint32_t desc_hi = (uint32_t)(smem_desc_hi(pitch) >> 32);
            const uint32_t lbo = 1u << 16;
            const int ksteps = KS ? KS : L.kc / 16, th = L.th, nblk = L.nblk, nchunks = L.nchunks, stages = L.stages;
            uint32_t tap_a[NTAPS], tap_b[NTAPS];
#pragma unroll
            for (int t = 0; t < NTAPS; ++t) {
                tap_a[t] = (uint32_t)(L.tap_row[t] * pitch) >> 4;
                tap_b[t] = (uint32_t)(t * nblk * pitch) >> 4;
            }
            const uint32_t m_step = (uint32_t)(L.m_row_stride * pitch) >> 4;
            const uint32_t stage16 = L.stage_bytes >> 4, b_off16 = L.a_stage_bytes >> 4;
            const int kym = L.kym;
            const uint32_t row16 = (uint32_t)pitch >> 4, nb16 = (uint32_t)(nblk * pitch) >> 4;
            uint32_t it = 0, ai = 0;
            for (int item = blockIdx.x; item < items; item += gridDim.x, ++ai) {
                const uint32_t buf = nbuf == 2 ? (ai & 1) : 0, use = ai / nbuf;
                mbar_wait(&acc_empty[buf], (use & 1) ^ 1);   // the epilogue has drained this accumulator buffer
                asm volatile("tcgen05.fence::after_thread_sync;" ::: "memory");
                const uint32_t tm = tmem + buf * (uint32_t)L.cols_buf;
                for (int ch = 0; ch < nchunks; ++ch, ++it) {
                    const int s = it % stages;
                    mbar_wait(&full_bar[s], (it / stages) & 1);
                    asm volatile("tcgen05.fence::after_thread_sync;" ::: "memory");
                    const uint32_t a0 = ((smem0 >> 4) + (uint32_t)s * stage16) | lbo, b0 = a0 + b_off16;
                    if (L.dbg & 1) {   // timing experiment (AB_TC_DBG): no MMAs
                    } else if (NTAPS == 9 && kym) {
                        // input row segment sg feeds output rows sg - u (u = dy + 1) of this item: u in [ulo, uhi]
                        for (int sg = 0; sg < th + 2; ++sg) {
                            const int ulo = max(0, sg - (th - 1)), uhi = min(2, sg);
                            const uint32_t nn = (uint32_t)((uhi - ulo + 1) * nblk);
                            const uint32_t idesc_n = (idesc & ~(0x3fu << 17)) | ((nn >> 3) << 17);
                            const uint32_t d = tm + (uint32_t)((sg * 3 + ulo) * nblk);
                            const uint32_t as = a0 + (uint32_t)sg * m_step;
                            const uint32_t bs = b0 + (uint32_t)ulo * nb16;
#pragma unroll
                            for (int dxi = 0; dxi < 3; ++dxi)
#pragma unroll
                                for (int k = 0; k < (KS ? KS : 4); ++k)
                                    if (KS || k < ksteps)
                                        umma_bf16(d, as + dxi * row16 + 2 * k, bs + (uint32_t)(dxi * 3) * nb16 + 2 * k, desc_hi, idesc_n,
                                                  (dxi | k) ? 1u : (uint32_t)ch);
                        }
                    } else {
                        for (int m = 0; m < th; ++m) {
                            const uint32_t am = a0 + (uint32_t)m * m_step, d = tm + (uint32_t)(m * nblk);
#pragma unroll
                            for (int t = 0; t < NTAPS; ++t)
#pragma unroll
                                for (int k = 0; k < (KS ? KS : 4); ++k)
                                    if (KS || k < ksteps)
                                        umma_bf16(d, am + tap_a[t] + 2 * k, b0 + tap_b[t] + 2 * k, desc_hi, idesc, (t | k) ? 1u : (uint32_t)ch);
                        }
                    }
                    umma_commit(&empty_bar[s]);
                }
                umma_commit(&acc_full[buf]);
            }
        }
    } else {
        // ---- epilogue (8 warps): thread <-> TMEM lane (position); warps 2-5 take the even M tiles, 6-9 the odd ones
        const int q = warp & 3;              // TMEM lane quarter this warp may access
        const int half = (warp - 2) >> 2;    // which of every kEpiWarps/4 consecutive M tiles this warp takes
        constexpr int kTileStep = kEpiWarps / 4;
        const bool up = L.mode == OUT_BF16_UP || L.mode == OUT_BF16_UP_ACCUM;
        const bool down = L.mode == OUT_BF16_DOWN_ACCUM || L.mode == OUT_BF16_DOWN;
        const bool accum = L.mode == OUT_BF16_ACCUM || L.mode == OUT_BF16_UP_ACCUM || L.mode == OUT_BF16_DOWN_ACCUM;
        const int act = L.act, nblk = L.nblk, P = L.P, rows_per_sample = L.H + 1;
        // position step between this thread's consecutive tiles (kTileStep M tiles), as rows + columns of the [rows][P] array
        const int step_r = (kTileStep * L.m_pos_stride) / P, step_c = (kTileStep * L.m_pos_stride) % P;
        const bool stats = L.stat_sums != nullptr;   // (host: only with OUT_F32_NHWC and nblk == npad_total == 16)
        float ssum[16] = {}, ssq[16] = {};
        uint32_t ai = 0;
        for (int it0 = blockIdx.x; it0 < items; it0 += gridDim.x, ++ai) {
            const int item = rev ? items - 1 - it0 : it0;
            const int by = item / grid_x, bx = item - by * grid_x;
            const int n0 = by * nblk;
            // (sample, row, column) of this thread's first position; later tiles advance incrementally
            int n, r, c;
            auto locate = [&](int m) {
                if (L.rows_mode) {
                    const int rg = bx / L.ncx, cx = bx - rg * L.ncx;
                    const int R = rg * L.th + m;               // global row of the [N*(H+1)][P] array
                    n = R / rows_per_sample; r = R - n * rows_per_sample;
                    c = cx * 128 + q * 32 + lane;
                } else {
                    const int pos0 = bx * (128 * L.th) + m * 128 + q * 32 + lane;
                    n = pos0 / (int)L.S;
                    const int rem = pos0 - n * (int)L.S;
                    r = rem / P; c = rem - r * P;
                }
            };
            locate(half);
            const uint32_t buf = nbuf == 2 ? (ai & 1) : 0, use = ai / nbuf;
            mbar_wait(&acc_full[buf], use & 1);
            asm volatile("tcgen05.fence::after_thread_sync;" ::: "memory");
            const uint32_t tm = tmem + buf * (uint32_t)L.cols_buf + ((uint32_t)(q * 32) << 16);
            // an odd tile count leaves one tile over: the two warp sets split its 16-channel blocks (even / odd)
            // split operands: the tile's columns are nsum groups of nout (filters hi | lo) that the epilogue adds
            const int nsum = L.nsum, nout = nblk / nsum, sstride = L.nsum_stride;
            const bool share = (L.th & 1) && nout >= 32;
            for (int m = half;; m += kTileStep) {
                int jpar = -1;
                if (m >= L.th) {
                    if (!share || half == 0) break;
                    m = L.th - 1; jpar = 1;
                    locate(m);
                } else if (share && m == L.th - 1) jpar = 0;
                bool valid = n < L.Nsamples && c < P;
                long long opos;
                if (up) {
                    // M grid = coarse tensor; row r <-> output rows 2r + a (r in [0, H] for a = 0, [0, H) for a = 1)
                    valid = valid && r <= L.H - L.par_a && c <= L.W - L.par_b;
                    opos = (long long)n * L.out_S + (long long)(2 * r + L.par_a + 1) * L.out_P + (2 * c + L.par_b + 1);
                } else if (down) {
                    // M grid = fine tensor; pixel (y, x) = (r - 1, c - 1) with y, x odd <-> coarse pixel ((y - 1)/2, (x - 1)/2)
                    valid = valid && r >= 2 && c >= 2 && !((r | c) & 1) && r <= L.H - 1 && c <= L.W - 1;
                    opos = (long long)n * L.out_S + (long long)(r >> 1) * L.out_P + (c >> 1);
                } else {
                    valid = valid && r >= 1 && c >= 1;
                    opos = (long long)n * L.S + (long long)r * P + c;
                }
                // Stores go out in lane PAIRS: lanes 2k and 2k+1 together write a whole 32-byte sector of row A (lane 2k's
                // position), then of row B, instead of a 16-byte half sector of their own row each -- half the L1 -> crossbar
                // sector requests (ncu: l1tex2xbar request cycles 71 % on the level-0 forward convolutions).
                const bool odd = lane & 1;
                const long long opos_p = __shfl_xor_sync(0xffffffffu, opos, 1);
                const bool valid_p = __shfl_xor_sync(0xffffffffu, (int)valid, 1) != 0;
                const long long posA = odd ? opos_p : opos, posB = odd ? opos : opos_p;
                const bool okA = odd ? valid_p : valid, okB = odd ? valid : valid_p;
                for (int j = jpar > 0 ? 16 : 0; j < nout; j += jpar >= 0 ? 32 : 16) {
                    float v[16];
                    if (NTAPS == 9 && L.kym) {
                        float w0[16], w1[16];
                        tmem_ld16_issue(tm + (uint32_t)((m * 3 + 0) * nblk + j), v);
                        tmem_ld16_issue(tm + (uint32_t)(((m + 1) * 3 + 1) * nblk + j), w0);
                        tmem_ld16_issue(tm + (uint32_t)(((m + 2) * 3 + 2) * nblk + j), w1);
                        tmem_ld_wait();
                        tmem_ld_fence(v); tmem_ld_fence(w0); tmem_ld_fence(w1);
#pragma unroll
                        for (int i = 0; i < 16; ++i) v[i] += w0[i] + w1[i];
                        if (nsum == 2) {
                            float x0[16];
                            tmem_ld16_issue(tm + (uint32_t)((m * 3 + 0) * nblk + sstride + j), x0);
                            tmem_ld16_issue(tm + (uint32_t)(((m + 1) * 3 + 1) * nblk + sstride + j), w0);
                            tmem_ld16_issue(tm + (uint32_t)(((m + 2) * 3 + 2) * nblk + sstride + j), w1);
                            tmem_ld_wait();
                            tmem_ld_fence(x0); tmem_ld_fence(w0); tmem_ld_fence(w1);
#pragma unroll
                            for (int i = 0; i < 16; ++i) v[i] += (x0[i] + w0[i]) + w1[i];
                        }
                    } else {
                        tmem_ld16_issue(tm + (uint32_t)(m * nblk + j), v);
                        if (nsum == 2) {
                            float x0[16];
                            tmem_ld16_issue(tm + (uint32_t)(m * nblk + sstride + j), x0);
                            tmem_ld_wait();
                            tmem_ld_fence(v); tmem_ld_fence(x0);
#pragma unroll
                            for (int i = 0; i < 16; ++i) v[i] += x0[i];
                        } else {
                            tmem_ld_wait();
                            tmem_ld_fence(v);
                        }
                    }
                    const int co0 = n0 + j;
                    if (co0 >= L.cout || (L.dbg & 4) || ((L.dbg & 8) && L.mode == OUT_F32_NHWC)) continue;   // (uniform) padding columns only; AB_TC_DBG & 4: no stores, & 8: no fp32 z stores
#pragma unroll
                    for (int i4 = 0; i4 < 4; ++i4) {
                        const float4 b = *reinterpret_cast<const float4*>(&s_bias[co0 + 4 * i4]);
                        v[4 * i4 + 0] += b.x; v[4 * i4 + 1] += b.y; v[4 * i4 + 2] += b.z; v[4 * i4 + 3] += b.w;
                    }
                    if (act != ACT_NONE) {
#pragma unroll
                        for (int i4 = 0; i4 < 4; ++i4) {
                            const float4 sc = *reinterpret_cast<const float4*>(&s_scale[co0 + 4 * i4]);
                            const float4 sh = *reinterpret_cast<const float4*>(&s_shift[co0 + 4 * i4]);
                            const float scv[4] = {sc.x, sc.y, sc.z, sc.w}, shv[4] = {sh.x, sh.y, sh.z, sh.w};
#pragma unroll
                            for (int e = 0; e < 4; ++e) {
                                float& x = v[4 * i4 + e];
                                x = act == ACT_MISH_AFFINE ? fmaf(mish_fast(x), scv[e], shv[e]) : mish_fast(fmaf(x, scv[e], shv[e]));
                            }
                        }
                    }
                    if (L.mode == OUT_F32_NCHW) {
                        // fp32 planes (the logits): consecutive lanes are consecutive pixels of a plane already
                        if (valid) {
                            const long long hw = (long long)L.H * L.W;
                            float* o = (float*)L.out + ((long long)n * L.cout) * hw + (long long)(r - 1) * L.W + (c - 1);
#pragma unroll
                            for (int i = 0; i < 16; ++i)
                                if (co0 + i < L.cout) o[(long long)(co0 + i) * hw] = v[i];
                        }
                    } else if (L.mode == OUT_F32_NHWC) {
                        if (stats && valid) {
#pragma unroll
                            for (int i = 0; i < 16; ++i) {
                                const float x = L.stat_mish ? mish_fast(v[i]) : v[i];
                                ssum[i] += x;
                                ssq[i] = fmaf(x, x, ssq[i]);
                            }
                        }
                        // row = four 16-byte quarters Q0..Q3; even lane writes Q0, Q2 of both rows, odd lane Q1, Q3
                        float rcv[8];
#pragma unroll
                        for (int i = 0; i < 4; ++i) {
                            rcv[i] = __shfl_xor_sync(0xffffffffu, odd ? v[i] : v[4 + i], 1);
                            rcv[4 + i] = __shfl_xor_sync(0xffffffffu, odd ? v[8 + i] : v[12 + i], 1);
                        }
                        const int qo = odd ? 4 : 0;
                        float* oa = (float*)L.out + posA * L.out_ctot + L.out_c0 + co0 + qo;
                        float* ob = (float*)L.out + posB * L.out_ctot + L.out_c0 + co0 + qo;
                        const bool c0ok = co0 + qo < L.cout, c1ok = co0 + 8 + qo < L.cout;
                        if (okA && c0ok) *reinterpret_cast<float4*>(oa) = odd ? make_float4(rcv[0], rcv[1], rcv[2], rcv[3]) : make_float4(v[0], v[1], v[2], v[3]);
                        if (okA && c1ok) *reinterpret_cast<float4*>(oa + 8) = odd ? make_float4(rcv[4], rcv[5], rcv[6], rcv[7]) : make_float4(v[8], v[9], v[10], v[11]);
                        if (okB && c0ok) *reinterpret_cast<float4*>(ob) = odd ? make_float4(v[4], v[5], v[6], v[7]) : make_float4(rcv[0], rcv[1], rcv[2], rcv[3]);
                        if (okB && c1ok) *reinterpret_cast<float4*>(ob + 8) = odd ? make_float4(v[12], v[13], v[14], v[15]) : make_float4(rcv[4], rcv[5], rcv[6], rcv[7]);
                    } else {
                        // h16 row block = two 16-byte halves (8 channels each); even lane writes half 0 of both rows, odd lane half 1
                        const int ho = odd ? 8 : 0;
                        const bool cok = co0 + ho < L.cout;
                        h16* oa = (h16*)L.out + posA * L.out_ctot + L.out_c0 + co0 + ho;
                        h16* ob = (h16*)L.out + posB * L.out_ctot + L.out_c0 + co0 + ho;
                        if (accum && !(L.dbg & 16)) {
                            // accumulate with ONE 16-byte reduction per half row (REDG.E.ADD.F16x8): the add happens in L2, nothing comes
                            // back to the SM, so the epilogue does not wait for the old value (the read-modify-write below cost ~1.1 ms of
                            // the step's conv time; AB_TC_DBG & 16 selects it again).  The addend is saturated, the sum is rounded once more
                            // in fp16 -- gradients only, which the loss scale keeps far from the fp16 range limits.
                            float rcv[8];
#pragma unroll
                            for (int i = 0; i < 8; ++i) rcv[i] = __shfl_xor_sync(0xffffffffu, odd ? v[i] : v[8 + i], 1);
                            __half2 ra[4], rb[4];
#pragma unroll
                            for (int i = 0; i < 4; ++i) {
                                const float a0 = odd ? rcv[2 * i] : v[2 * i], a1 = odd ? rcv[2 * i + 1] : v[2 * i + 1];
                                const float b0 = odd ? v[8 + 2 * i] : rcv[2 * i], b1 = odd ? v[8 + 2 * i + 1] : rcv[2 * i + 1];
                                ra[i] = sat_half2(a0, a1);
                                rb[i] = sat_half2(b0, b1);
                            }
                            if (okA && cok) red_add_h8(oa, reinterpret_cast<const uint32_t*>(ra));
                            if (okB && cok) red_add_h8(ob, reinterpret_cast<const uint32_t*>(rb));
                        } else if (accum) {
                            uint4 oldA = make_uint4(0, 0, 0, 0), oldB = make_uint4(0, 0, 0, 0);
                            if (okA && cok) oldA = *reinterpret_cast<const uint4*>(oa);
                            if (okB && cok) oldB = *reinterpret_cast<const uint4*>(ob);
                            float rcv[8];
#pragma unroll
                            for (int i = 0; i < 8; ++i) rcv[i] = __shfl_xor_sync(0xffffffffu, odd ? v[i] : v[8 + i], 1);
                            const __half2* pa = reinterpret_cast<const __half2*>(&oldA);
                            const __half2* pb = reinterpret_cast<const __half2*>(&oldB);
                            __half2 ra[4], rb[4];
#pragma unroll
                            for (int i = 0; i < 4; ++i) {
                                const float2 fa = __half22float2(pa[i]), fb = __half22float2(pb[i]);
                                // row A's values: odd lane got them from its partner; row B's: even lane did
                                const float a0 = odd ? rcv[2 * i] : v[2 * i], a1 = odd ? rcv[2 * i + 1] : v[2 * i + 1];
                                const float b0 = odd ? v[8 + 2 * i] : rcv[2 * i], b1 = odd ? v[8 + 2 * i + 1] : rcv[2 * i + 1];
                                ra[i] = sat_half2(fa.x + a0, fa.y + a1);
                                rb[i] = sat_half2(fb.x + b0, fb.y + b1);
                            }
                            if (okA && cok) *reinterpret_cast<uint4*>(oa) = *reinterpret_cast<const uint4*>(ra);
                            if (okB && cok) *reinterpret_cast<uint4*>(ob) = *reinterpret_cast<const uint4*>(rb);
                        } else {
                            __half2 pk[8], pl[8];
#pragma unroll
                            for (int i = 0; i < 8; ++i) split_half2(v[2 * i], v[2 * i + 1], pk[i], pl[i]);
                            // hi halves, then (split tensors) the rounding residuals lo_off elements behind
                            for (int half_i = 0; half_i < (L.out_lo_off ? 2 : 1); ++half_i) {
                                const uint32_t* pw = reinterpret_cast<const uint32_t*>(half_i ? pl : pk);
                                uint32_t rc[4];
#pragma unroll
                                for (int i = 0; i < 4; ++i) rc[i] = __shfl_xor_sync(0xffffffffu, odd ? pw[i] : pw[4 + i], 1);
                                const uint4 mine0 = make_uint4(pw[0], pw[1], pw[2], pw[3]), mine1 = make_uint4(pw[4], pw[5], pw[6], pw[7]);
                                const uint4 got = make_uint4(rc[0], rc[1], rc[2], rc[3]);
                                const long long lo = half_i ? L.out_lo_off : 0;
                                if (okA && cok) *reinterpret_cast<uint4*>(oa + lo) = odd ? got : mine0;
                                if (okB && cok) *reinterpret_cast<uint4*>(ob + lo) = odd ? mine1 : got;
                            }
                        }
                    }
                }
                if (jpar > 0) break;
                // next tile of this thread
                c += step_c; r += step_r;
                // FLAT tiles walk on through the rows.  ROWS tiles keep their columns: a column at or beyond P (a partial last
                // column tile) must stay invalid for every row -- wrapped into the next row it would be written a second time
                // (harmless: the flat layout makes it the same value) and COUNTED a second time by the fused statistics.
                if (!L.rows_mode && c >= P) { c -= P; ++r; }
                while (r >= rows_per_sample) { r -= rows_per_sample; ++n; }
            }
            // this thread's TMEM reads of the buffer are complete (tcgen05.wait::ld in tmem_ld16): hand it back to the MMA warp
            asm volatile("tcgen05.fence::before_thread_sync;" ::: "memory");
            asm volatile("mbarrier.arrive.shared::cta.b64 _, [%0];" ::"r"(smem_u32(&acc_empty[buf])) : "memory");
        }
        if (stats) {   // warp totals -> shared doubles (8 warps) -> one global atomic per channel and CTA below
#pragma unroll
            for (int i = 0; i < 16; ++i) {
                float a = ssum[i], b = ssq[i];
#pragma unroll
                for (int o = 16; o; o >>= 1) { a += __shfl_xor_sync(0xffffffffu, a, o); b += __shfl_xor_sync(0xffffffffu, b, o); }
                if (lane == 0) { atomicAdd(&s_stat[i], (double)a); atomicAdd(&s_stat[16 + i], (double)b); }
            }
        }
    }
    asm volatile("tcgen05.fence::before_thread_sync;" ::: "memory");
    __syncthreads();
    if (L.stat_sums && threadIdx.x < 32) {
        const int c = threadIdx.x & 15;
        if (c < L.cout) atomicAdd(L.stat_sums + (threadIdx.x >> 4) * L.cout + c, s_stat[threadIdx.x]);
    }
    if (warp == 1) asm volatile("tcgen05.dealloc.cta_group::1.sync.aligned.b32 %0, %1;" ::"r"(tmem), "r"(L.tmem_cols));
}

__global__ void pack_filters_kernel(const float* __restrict__ w, h16* __restrict__ packed, int nreal, int npad, int kreal, int kp,
                                    int ntaps, long long sn, long long sk, int t0, int t1, int t2, int t3, int t4, int t5, int t6,
                                    int t7, int t8)
{
    const int tapmap[9] = {t0, t1, t2, t3, t4, t5, t6, t7, t8};
    const long long total = (long long)ntaps * npad * kp;
    for (long long i = (long long)blockIdx.x * blockDim.x + threadIdx.x; i < total; i += (long long)gridDim.x * blockDim.x) {
        const int kidx = (int)(i % kp);
        const int n = (int)((i / kp) % npad);
        const int t = (int)(i / ((long long)kp * npad));
        float v = 0.f;
        if (n < nreal && kidx < kreal) v = w[n * sn + kidx * sk + tapmap[t]];
        packed[i] = __float2half_rn(fminf(fmaxf(v, -65504.f), 65504.f));
    }
}

typedef CUresult (*EncodeTiledFn)(CUtensorMap*, CUtensorMapDataType, cuuint32_t, void*, const cuuint64_t*, const cuuint64_t*,
                                  const cuuint32_t*, const cuuint32_t*, CUtensorMapInterleave, CUtensorMapSwizzle,
                                  CUtensorMapL2promotion, CUtensorMapFloatOOBfill);

EncodeTiledFn encode_fn()
{
    static EncodeTiledFn fn = [] {
        void* p = nullptr;
        cudaDriverEntryPointQueryResult q;
        if (cudaGetDriverEntryPoint("cuTensorMapEncodeTiled", &p, cudaEnableDefault, &q) != cudaSuccess || !p)
            throw std::runtime_error("cuTensorMapEncodeTiled is not available from this driver");
        return (EncodeTiledFn)p;
    }();
    return fn;
}

void make_map_2d(CUtensorMap* m, const void* base, long long rows, int ctot, int kc, int box_rows, CUtensorMapL2promotion promo)
{
    cuuint64_t dims[2] = {(cuuint64_t)ctot, (cuuint64_t)rows};
    cuuint64_t strides[1] = {(cuuint64_t)ctot * 2};
    cuuint32_t box[2] = {(cuuint32_t)kc, (cuuint32_t)box_rows}, es[2] = {1, 1};
    const CUtensorMapSwizzle sw = kc == 64 ? CU_TENSOR_MAP_SWIZZLE_128B : kc == 32 ? CU_TENSOR_MAP_SWIZZLE_64B : CU_TENSOR_MAP_SWIZZLE_32B;
    const CUresult r = encode_fn()(m, CU_TENSOR_MAP_DATA_TYPE_FLOAT16, 2, const_cast<void*>(base), dims, strides, box, es,
                                   CU_TENSOR_MAP_INTERLEAVE_NONE, sw, promo, CU_TENSOR_MAP_FLOAT_OOB_FILL_NONE);
    if (r != CUDA_SUCCESS)
        throw std::runtime_error("cuTensorMapEncodeTiled failed (" + std::to_string((int)r) + "): rows " + std::to_string(rows) + " ctot " +
                                 std::to_string(ctot) + " kc " + std::to_string(kc) + " box_rows " + std::to_string(box_rows));
}

}  // namespace

void make_tensor_map_2d(CUtensorMap* m, const void* base, long long rows, int ctot, int box_ch, int box_rows)
{
    // L2 promotion of the TMA fetches.  A box row is box_ch * 2 bytes of a ctot * 2-byte row: for the narrow channel slices of the
    // full-resolution levels (32 of 160 bytes) a 128-byte promotion fetches neighbours nobody asked for.  AB_TMA_PROMO: 0 none,
    // 1 64 B, 2 128 B, 3 256 B, 4 (default) = the smallest promotion covering the box row.
    static const int promo_env = getenv("AB_TMA_PROMO") ? atoi(getenv("AB_TMA_PROMO")) : 2;
    CUtensorMapL2promotion promo = CU_TENSOR_MAP_L2_PROMOTION_L2_128B;
    const int row_bytes = box_ch * 2;
    switch (promo_env) {
        case 0: promo = CU_TENSOR_MAP_L2_PROMOTION_NONE; break;
        case 1: promo = CU_TENSOR_MAP_L2_PROMOTION_L2_64B; break;
        case 3: promo = CU_TENSOR_MAP_L2_PROMOTION_L2_256B; break;
        case 4: promo = row_bytes <= 32 ? CU_TENSOR_MAP_L2_PROMOTION_NONE : row_bytes <= 64 ? CU_TENSOR_MAP_L2_PROMOTION_L2_64B : CU_TENSOR_MAP_L2_PROMOTION_L2_128B; break;
        default: break;
    }
    make_map_2d(m, base, rows, ctot, box_ch, box_rows, promo);
}

int padded_n(int cout) { return (cout + 15) / 16 * 16; }
int padded_k(int cin) { return (cin + 63) / 64 * 64; }

size_t packed_filter_elems(int cin, int npad_total, int ntaps) { return (size_t)ntaps * npad_total * padded_k(cin); }

void pack_filters_bf16(const float* w, h16* packed, int nreal, int npad_total, int kreal, int ntaps, const int* tapmap, long long sn,
                       long long sk, cudaStream_t st)
{
    const int kp = padded_k(kreal);
    const long long total = (long long)ntaps * npad_total * kp;
    int tm[9] = {0};
    for (int i = 0; i < ntaps; ++i) tm[i] = tapmap[i];
    k::ProfScope ps(k::F_PACK, 0, 6.0 * total, st);
    pack_filters_kernel<<<(unsigned)std::min<long long>((total + 255) / 256, 1024), 256, 0, st>>>(
        w, packed, nreal, npad_total, kreal, kp, ntaps, sn, sk, tm[0], tm[1], tm[2], tm[3], tm[4], tm[5], tm[6], tm[7], tm[8]);
    k::count_launch();
}

void setup_conv(ConvLaunch& L, const ConvSetup& s)
{
    if (s.cin % 8 || s.src_ctot % 8 || s.a_c0 % 8)
        throw std::invalid_argument("the tensor-core path needs channel counts that are multiples of 8");
    const int parts = s.split ? 2 : 1;
    L.a_c0 = s.a_c0;
    L.ntaps = s.ntaps;
    // split operands: the hi and the lo filters sit side by side along N (tc.h): the GEMM's N doubles, the epilogue adds the halves
    const int npad_total = s.npad_total * parts;
    L.npad_total = npad_total;
    L.nsum = parts; L.nsum_stride = s.npad_total;
    // N block: whole N when it fits one MMA (<= 256), else the smallest even split
    int nblocks = (npad_total + 255) / 256;
    L.nblk = ((npad_total + nblocks - 1) / nblocks + 15) / 16 * 16;
    nblocks = (npad_total + L.nblk - 1) / L.nblk;
    if (nblocks * L.nblk != npad_total) throw std::logic_error("N blocks must tile the padded N");
    if (s.split && nblocks != 1) throw std::invalid_argument("split operands need the doubled N to fit one MMA (at most 128 output channels)");
    if (s.grid.positions() * (s.split ? 2 : 1) + 4096 >= (1ll << 31))
        throw std::invalid_argument("tensor too large for the tcgen05 conv kernel (positions must fit 31 bits)");
    const int P = s.grid.P();
    L.total = (long long)s.grid.N * s.grid.S();
    L.S = s.grid.S(); L.P = P; L.H = s.grid.H; L.W = s.grid.W; L.Nsamples = s.grid.N;
    int dylo = 0, dyhi = 0, dxlo = 0, dxhi = 0, lo = 0, hi = 0;
    for (int t = 0; t < s.ntaps; ++t) {
        dylo = std::min(dylo, s.tap_dy[t]); dyhi = std::max(dyhi, s.tap_dy[t]);
        dxlo = std::min(dxlo, s.tap_dx[t]); dxhi = std::max(dxhi, s.tap_dx[t]);
        const int off = s.tap_dy[t] * P + s.tap_dx[t];
        lo = std::min(lo, off); hi = std::max(hi, off);
    }
    bool kym_ok = s.ntaps == 9 && nblocks == 1;
    for (int t = 0; t < s.ntaps && kym_ok; ++t) kym_ok = s.tap_dx[t] == t / 3 - 1 && s.tap_dy[t] == t % 3 - 1;
    static const bool no_kym = getenv("AB_TC_NO_KYM") != nullptr;
    const long long rows_total = (long long)s.grid.N * (s.grid.H + 1);
    const long long tiles_flat = (L.total + 127) / 128;
    const int ncx = (P + 127) / 128;

    // candidate search: K chunk, tiling mode, M tiles per CTA (see the cost model in DESIGN.md)
    struct Cand { double cost; int kc, rows_mode, th, nseg, seg_rows, nloads, box_rows, stages; unsigned a_bytes, stage_bytes; long long ctas; };
    Cand best{1e300, 0, 0, 0, 0, 0, 0, 0, 0, 0, 0, 0};
    static const int force_kc = getenv("AB_TC_KC") ? atoi(getenv("AB_TC_KC")) : 0, force_th = getenv("AB_TC_TH") ? atoi(getenv("AB_TC_TH")) : 0;
    static const int force_mode = getenv("AB_TC_MODE") ? atoi(getenv("AB_TC_MODE")) : -1;
    for (int kc : {16, 32, 64}) {
        if (force_kc && kc != force_kc && s.cin >= force_kc) continue;
        const int kpad = (s.cin + kc - 1) / kc * kc, nchunks = kpad / kc * parts, pitch = kc * 2;
        for (int rows_mode = 0; rows_mode < 3; ++rows_mode) {   // 0 FLAT, 1 ROWS, 2 ROWS + merged dy taps
            if (rows_mode && P < 128) continue;
            if (rows_mode == 2 && (!kym_ok || no_kym)) continue;
            if (force_mode >= 0 && rows_mode != force_mode && P >= 128) continue;
            for (int th : {8, 4, 2, 1}) {
                if (rows_mode == 2 && th < 2) continue;
                if (force_th && th > force_th) continue;
                const int nseg = rows_mode ? th + dyhi - dylo : 1;
                if (nseg > 10) continue;
                const int seg_rows = rows_mode ? 128 + dxhi - dxlo : 128 * th + hi - lo;
                const int nloads = (seg_rows + 255) / 256;
                const int box_rows = ((seg_rows + nloads - 1) / nloads + 7) / 8 * 8;
                const unsigned a_bytes = (unsigned)(((size_t)nseg * nloads * box_rows * pitch + 1023) / 1024 * 1024);
                const unsigned stage = a_bytes + (unsigned)(((size_t)s.ntaps * L.nblk * pitch + 1023) / 1024 * 1024);
                const int cols_needed = (rows_mode == 2 ? (th + 2) * 3 : th) * L.nblk;
                if (cols_needed > 512) continue;
                const bool two = cols_needed <= 256;   // double-buffered accumulators: MMAs of item i+1 overlap the epilogue of item i
                const size_t budget = 200u * 1024u;
                const int want = std::min(nchunks, 2);
                if ((size_t)stage * want + 1024 > budget) continue;
                const int stages = (int)std::min<size_t>(kMaxStages, (budget - 1024) / stage);
                const long long ctas = rows_mode ? ((rows_total + th - 1) / th) * ncx : (tiles_flat + th - 1) / th;
                const double waste = rows_mode ? (double)ncx * 128 / P * ((double)((rows_total + th - 1) / th) * th / rows_total)
                                               : (double)((tiles_flat + th - 1) / th) * th / tiles_flat;
                const double loadf = (double)nseg * nloads * box_rows / (128.0 * th);
                // cycle model of one work item on one SM (measured constants: experiments/umma_bench.cu, tma_bench.cu):
                //   tcgen05.mma M128 K16: 32 + N/4 cycles (shared-memory operand reads) or N/2 (tensor pipe);  TMA: ~2.4 cycles
                //   per box row with every SM loading;  epilogue: ~600 cycles per 128 x 16 accumulator block.
                const int ksteps = kc / 16;
                auto mma_cyc = [](int n) { return std::max(32.0 + n / 4.0, n / 2.0); };
                double t_mma;
                if (rows_mode == 2) {
                    t_mma = 0;
                    for (int sg = 0; sg < th + 2; ++sg)
                        t_mma += 3.0 * ksteps * mma_cyc((std::min(2, sg) - std::max(0, sg - (th - 1)) + 1) * L.nblk);
                } else t_mma = (double)th * s.ntaps * ksteps * mma_cyc(L.nblk);
                t_mma *= nchunks;
                const double t_tma = 2.4 * nchunks * ((double)nseg * nloads * box_rows + (double)s.ntaps * L.nblk);
                const double t_epi = 600.0 * th * (L.nblk / 16) * (rows_mode == 2 ? 1.15 : 1.0);
                // front end (TMA -> MMA) through `stages` shared-memory buffers: one buffer's cycle is fill + ~1 us of TMA / commit
                // latency + consume, and `stages` of them overlap.  A single stage serialises all three (measured: the level-0
                // data-gradient conv at kc 64 / 1 stage took 5.7k cycles per item against 2.4k modelled without this term).
                const double t_front = std::max(std::max(t_tma, t_mma), (t_tma + t_mma + 1800.0 * nchunks) / stages);
                const double t_item = two ? std::max(t_front, t_epi) : std::max(t_front, t_mma + t_epi);
                const long long rounds = (ctas * nblocks + k::num_sms() - 1) / k::num_sms();
                (void)waste; (void)loadf;
                double cost = (double)rounds * t_item + 3000.0;   // + pipeline fill / drain
                cost *= 1.0 + 0.01 * (kc == 16 ? 2 : kc == 32 ? 1 : 0);  // ties: fewer, larger TMA boxes
                if (cost < best.cost) best = Cand{cost, kc, rows_mode, th, nseg, seg_rows, nloads, box_rows, stages, a_bytes, stage, ctas};
            }
        }
    }
    if (best.kc == 0) throw std::logic_error("no tcgen05 conv tiling fits shared memory");
    L.kc = best.kc;
    L.nch1 = (s.cin + L.kc - 1) / L.kc;
    L.nchunks = L.nch1 * parts;
    L.a_lo_rows = s.split ? s.src_positions : 0;
    L.rows_mode = best.rows_mode != 0; L.kym = best.rows_mode == 2; L.th = best.th; L.nseg = best.nseg; L.nloads = best.nloads; L.box_rows = best.box_rows;
    L.seg_alloc = best.nloads * best.box_rows;
    L.a_stage_bytes = best.a_bytes; L.stage_bytes = best.stage_bytes; L.stages = std::max(best.stages, 1);
    L.smem_bytes = L.stages * L.stage_bytes + 1024;
    L.ncx = ncx;
    if (L.rows_mode) {
        L.m_pos_stride = P; L.m_row_stride = L.seg_alloc;
        for (int sg = 0; sg < L.nseg; ++sg) L.seg_pos[sg] = (sg + dylo) * P + dxlo;
        for (int t = 0; t < s.ntaps; ++t) L.tap_row[t] = (s.tap_dy[t] - dylo) * L.seg_alloc + (s.tap_dx[t] - dxlo);
    } else {
        L.m_pos_stride = 128; L.m_row_stride = 128;
        L.seg_pos[0] = lo;
        for (int t = 0; t < s.ntaps; ++t) L.tap_row[t] = s.tap_dy[t] * P + s.tap_dx[t] - lo;
    }
    L.cols_buf = (L.kym ? (L.th + 2) * 3 : L.th) * L.nblk;
    L.nbuf = L.cols_buf <= 256 ? 2 : 1;
    int cols = 32;
    while (cols < L.nbuf * L.cols_buf) cols *= 2;
    L.tmem_cols = cols;
    L.grid_x = best.ctas;
    static const bool debug = getenv("AB_TC_DEBUG") != nullptr;
    if (debug)
        fprintf(stderr, "tc conv: P %d cin %d npad %d taps %d -> kc %d %s th %d nseg %d seg %d stages %d smem %u ctas %lld cost %.1f\n", P, s.cin,
                s.npad_total, s.ntaps, L.kc, L.kym ? "KYM" : L.rows_mode ? "ROWS" : "FLAT", L.th, L.nseg, L.seg_alloc, L.stages, L.smem_bytes, best.ctas, best.cost);
    // activations: 64-byte L2 promotion -- a narrow channel slice costs 64 B of HBM traffic per position instead of 128 (experiments/tma_bench.cu)
    static const int promo_env = getenv("AB_TMA_PROMO") ? atoi(getenv("AB_TMA_PROMO")) : 2;   // (see make_tensor_map_2d)
    CUtensorMapL2promotion promoA = L.kc * 2 < 128 ? CU_TENSOR_MAP_L2_PROMOTION_L2_64B : CU_TENSOR_MAP_L2_PROMOTION_L2_128B;
    if (promo_env == 0) promoA = CU_TENSOR_MAP_L2_PROMOTION_NONE;
    if (promo_env == 4 && L.kc * 2 <= 32) promoA = CU_TENSOR_MAP_L2_PROMOTION_NONE;
    make_map_2d(&L.tmA, s.src, s.src_positions * (s.split ? 2 : 1), s.src_ctot, L.kc, L.box_rows, promoA);
    make_map_2d(&L.tmB, s.packed, (long long)s.ntaps * npad_total, padded_k(s.cin), L.kc, L.nblk, CU_TENSOR_MAP_L2_PROMOTION_L2_128B);
}

template <int NTAPS, int KS>
static void launch_conv_t(const ConvLaunch& L, cudaStream_t st, int rev)
{
    static bool attr_set = false;
    if (!attr_set) {
        cudaFuncSetAttribute(conv_tc_kernel<NTAPS, KS>, cudaFuncAttributeMaxDynamicSharedMemorySize, 220 * 1024);
        attr_set = true;
    }
    const long long items = L.grid_x * (L.npad_total / L.nblk);
    k::launch_pdl(conv_tc_kernel<NTAPS, KS>, dim3((unsigned)std::min<long long>(items, k::num_sms())), dim3(kThreads), (size_t)L.smem_bytes, st, L, rev);
}

template <int NTAPS>
static void launch_conv_k(const ConvLaunch& L, cudaStream_t st, int rev)
{
    switch (L.kc) {
    case 16: launch_conv_t<NTAPS, 1>(L, st, rev); break;
    case 32: launch_conv_t<NTAPS, 2>(L, st, rev); break;
    default: launch_conv_t<NTAPS, 4>(L, st, rev); break;
    }
}

void launch_conv(const ConvLaunch& L0, cudaStream_t st, int rev)
{
    static const int dbg = getenv("AB_TC_DBG") ? atoi(getenv("AB_TC_DBG")) : 0;   // timing experiments: 1 no MMAs, 2 no TMA loads, 4 no stores
    ConvLaunch Lc;
    if (dbg) { Lc = L0; Lc.dbg = dbg; }
    const ConvLaunch& L = dbg ? Lc : L0;
    k::ProfScope ps(k::F_CONV_TC, L.flops, L.bytes, st);
    switch (L.ntaps) {
    case 1: launch_conv_k<1>(L, st, rev); break;
    case 2: launch_conv_t<2, 0>(L, st, rev); break;
    case 4: launch_conv_t<4, 0>(L, st, rev); break;
    case 9: launch_conv_k<9>(L, st, rev); break;
    default: throw std::logic_error("unsupported tap count");
    }
    k::count_launch();
}

}  // namespace tc
}  // namespace ab
