// bf16 tcgen05 / TMEM / TMA implicit-GEMM convolution (AB_PRECISION_BF16): con_ / cont_ forward and
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
__device__ __forceinline__ void umma_commit(uint64_t* bar)
{
    asm volatile("tcgen05.commit.cta_group::1.mbarrier::arrive::one.shared::cluster.b64 [%0];" ::"r"(smem_u32(bar)) : "memory");
}
__device__ __forceinline__ void tmem_ld16(uint32_t taddr, float v[16])
{
    uint32_t r[16];
    asm volatile("tcgen05.ld.sync.aligned.32x32b.x16.b32 {%0,%1,%2,%3,%4,%5,%6,%7,%8,%9,%10,%11,%12,%13,%14,%15}, [%16];"
                 : "=r"(r[0]), "=r"(r[1]), "=r"(r[2]), "=r"(r[3]), "=r"(r[4]), "=r"(r[5]), "=r"(r[6]), "=r"(r[7]), "=r"(r[8]),
                   "=r"(r[9]), "=r"(r[10]), "=r"(r[11]), "=r"(r[12]), "=r"(r[13]), "=r"(r[14]), "=r"(r[15])
                 : "r"(taddr));
    asm volatile("tcgen05.wait::ld.sync.aligned;" ::: "memory");
#pragma unroll
    for (int i = 0; i < 16; ++i) v[i] = __uint_as_float(r[i]);
}

// K-major shared-memory matrix descriptor (cute::UMMA::SmemDescriptor): start address >> 4, LBO (unused for
// swizzled K-major) = 1, SBO = 8 rows * pitch, version 1, layout type 6/4/2 for 32/64/128-byte swizzle.
__device__ __forceinline__ uint64_t smem_desc_hi(int pitch)
{
    const uint64_t sbo = (uint64_t)(8 * pitch) >> 4;
    const uint64_t layout = pitch == 128 ? 2 : pitch == 64 ? 4 : 6;
    return (sbo << 32) | (1ull << 46) | (layout << 61) | (1ull << 16);
}

constexpr int kThreads = 192;
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

// mish for the bf16 path: x - 2x / (e^2 + 2e + 2) with the fast exp / divide (results are rounded to bf16 anyway)
__device__ __forceinline__ float mish_fast(float x)
{
    const float e = __expf(x);
    return x - __fdividef(2.f * x, fmaf(e, e + 2.f, 2.f));
}

template <int NTAPS>
__global__ void __launch_bounds__(kThreads) conv_tc_kernel(const __grid_constant__ ConvLaunch L)
{
    extern __shared__ __align__(1024) uint8_t smem_raw[];
    __shared__ uint64_t full_bar[kMaxStages], empty_bar[kMaxStages], acc_bar;
    __shared__ uint32_t tmem_base_slot;
    __shared__ __align__(16) float s_bias[256], s_scale[256], s_shift[256];

    const uint32_t smem0 = (smem_u32(smem_raw) + 1023u) & ~1023u;
    const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
    const int pitch = L.kc * 2;
    long long o0;
    int cx = 0;
    if (L.rows_mode) {
        cx = (int)(blockIdx.x % L.ncx);
        o0 = (long long)(blockIdx.x / L.ncx) * L.th * L.P + cx * 128;
    } else {
        o0 = (long long)blockIdx.x * (128 * L.th);
    }
    const int n0 = blockIdx.y * L.nblk;

    if (threadIdx.x == 0) {
        for (int s = 0; s < L.stages; ++s) { mbar_init(&full_bar[s], 1); mbar_init(&empty_bar[s], 1); }
        mbar_init(&acc_bar, 1);
        asm volatile("fence.mbarrier_init.release.cluster;" ::: "memory");
    }
    if (warp == 1) {
        asm volatile("tcgen05.alloc.cta_group::1.sync.aligned.shared::cta.b32 [%0], %1;" ::"r"(smem_u32(&tmem_base_slot)), "r"(L.tmem_cols));
        asm volatile("tcgen05.relinquish_alloc_permit.cta_group::1.sync.aligned;");
    }
    if (warp >= 2) {  // per-channel epilogue parameters of this CTA's N block
        for (int i = threadIdx.x - 64; i < L.nblk; i += 128) {
            const int co = n0 + i;
            const bool in = co < L.cout;
            s_bias[i] = (in && L.bias) ? __ldg(L.bias + co) : 0.f;
            s_scale[i] = (in && L.act != ACT_NONE) ? __ldg(L.scale + co) : 1.f;
            s_shift[i] = (in && L.act != ACT_NONE) ? __ldg(L.shift + co) : 0.f;
        }
    }
    asm volatile("tcgen05.fence::before_thread_sync;" ::: "memory");
    __syncthreads();
    asm volatile("tcgen05.fence::after_thread_sync;" ::: "memory");
    const uint32_t tmem = tmem_base_slot;

    if (warp == 0) {
        if (lane == 0) {
            const uint32_t tx = (uint32_t)(L.nseg * L.seg_alloc + NTAPS * L.nblk) * pitch;
            for (int ch = 0; ch < L.nchunks; ++ch) {
                const int s = ch % L.stages;
                mbar_wait(&empty_bar[s], ((ch / L.stages) & 1) ^ 1);
                mbar_expect(&full_bar[s], tx);
                const uint32_t a_s = smem0 + (uint32_t)s * L.stage_bytes, b_s = a_s + L.a_stage_bytes;
                const int c = L.a_c0 + ch * L.kc;
                for (int sg = 0; sg < L.nseg; ++sg)
                    for (int l = 0; l < L.nloads; ++l)
                        tma_2d(a_s + (uint32_t)(sg * L.seg_alloc + l * L.box_rows) * pitch, &L.tmA, c,
                               (int)(o0 + L.seg_pos[sg]) + l * L.box_rows, &full_bar[s]);
#pragma unroll
                for (int t = 0; t < NTAPS; ++t)
                    tma_2d(b_s + (uint32_t)(t * L.nblk) * pitch, &L.tmB, ch * L.kc, t * L.npad_total + n0, &full_bar[s]);
            }
        }
    } else if (warp == 1) {
        if (lane == 0) {
            // instruction descriptor: D f32, A/B bf16, both K-major, N = nblk, M = 128
            const uint32_t idesc = (1u << 4) | (1u << 7) | (1u << 10) | ((uint32_t)(L.nblk >> 3) << 17) | ((128u >> 4) << 24);
            const uint32_t desc_hi = (uint32_t)(smem_desc_hi(pitch) >> 32);
            const uint32_t lbo = 1u << 16;
            const int ksteps = L.kc / 16, th = L.th, nblk = L.nblk, nchunks = L.nchunks, stages = L.stages;
            uint32_t tap_a[NTAPS], tap_b[NTAPS];
#pragma unroll
            for (int t = 0; t < NTAPS; ++t) {
                tap_a[t] = (uint32_t)(L.tap_row[t] * pitch) >> 4;
                tap_b[t] = (uint32_t)(t * nblk * pitch) >> 4;
            }
            const uint32_t m_step = (uint32_t)(L.m_row_stride * pitch) >> 4;
            const uint32_t stage16 = L.stage_bytes >> 4, b_off16 = L.a_stage_bytes >> 4;
            for (int ch = 0; ch < nchunks; ++ch) {
                const int s = ch % stages;
                mbar_wait(&full_bar[s], (ch / stages) & 1);
                asm volatile("tcgen05.fence::after_thread_sync;" ::: "memory");
                const uint32_t a0 = ((smem0 >> 4) + (uint32_t)s * stage16) | lbo, b0 = a0 + b_off16;
                for (int m = 0; m < th; ++m) {
                    const uint32_t am = a0 + (uint32_t)m * m_step, d = tmem + (uint32_t)(m * nblk);
#pragma unroll
                    for (int t = 0; t < NTAPS; ++t)
                        for (int k = 0; k < ksteps; ++k)
                            umma_bf16(d, am + tap_a[t] + 2 * k, b0 + tap_b[t] + 2 * k, desc_hi, idesc, (uint32_t)(ch | t | k));
                }
                umma_commit(&empty_bar[s]);
            }
            umma_commit(&acc_bar);
        }
    } else {
        // ---- epilogue: thread <-> TMEM lane (position), 16 columns per tcgen05.ld
        mbar_wait(&acc_bar, 0);
        asm volatile("tcgen05.fence::after_thread_sync;" ::: "memory");
        const int q = warp & 3;  // TMEM lane quarter this warp may access
        const bool up = L.mode == OUT_BF16_UP || L.mode == OUT_BF16_UP_ACCUM;
        const bool accum = L.mode == OUT_BF16_ACCUM || L.mode == OUT_BF16_UP_ACCUM;
        const int act = L.act;
        const bool col_ok = !L.rows_mode || cx * 128 + q * 32 + lane < L.P;
        for (int m = 0; m < L.th; ++m) {
            const long long pos = o0 + (long long)m * L.m_pos_stride + q * 32 + lane;
            const int n = (int)(pos / L.S);
            const int rem = (int)(pos - (long long)n * L.S);
            const int r = rem / L.P, c = rem - r * L.P;
            bool valid = pos < L.total && col_ok;
            long long opos = pos;
            if (up) {
                // M grid = coarse tensor; row r <-> output rows 2r + a (r in [0, H] for a = 0, [0, H) for a = 1)
                valid = valid && r <= L.H - L.par_a && c <= L.W - L.par_b;
                opos = (long long)n * L.out_S + (long long)(2 * r + L.par_a + 1) * L.out_P + (2 * c + L.par_b + 1);
            } else {
                valid = valid && r >= 1 && c >= 1;
            }
            for (int j = 0; j < L.nblk; j += 16) {
                float v[16];
                tmem_ld16(tmem + ((uint32_t)(q * 32) << 16) + (uint32_t)(m * L.nblk + j), v);
                const int co0 = n0 + j;
                if (!valid || co0 >= L.cout) continue;
#pragma unroll
                for (int i4 = 0; i4 < 4; ++i4) {
                    const float4 b = *reinterpret_cast<const float4*>(&s_bias[j + 4 * i4]);
                    v[4 * i4 + 0] += b.x; v[4 * i4 + 1] += b.y; v[4 * i4 + 2] += b.z; v[4 * i4 + 3] += b.w;
                }
                if (act != ACT_NONE) {
#pragma unroll
                    for (int i4 = 0; i4 < 4; ++i4) {
                        const float4 sc = *reinterpret_cast<const float4*>(&s_scale[j + 4 * i4]);
                        const float4 sh = *reinterpret_cast<const float4*>(&s_shift[j + 4 * i4]);
                        const float scv[4] = {sc.x, sc.y, sc.z, sc.w}, shv[4] = {sh.x, sh.y, sh.z, sh.w};
#pragma unroll
                        for (int e = 0; e < 4; ++e) {
                            float& x = v[4 * i4 + e];
                            x = act == ACT_MISH_AFFINE ? fmaf(mish_fast(x), scv[e], shv[e]) : mish_fast(fmaf(x, scv[e], shv[e]));
                        }
                    }
                }
                if (L.mode == OUT_F32_NCHW) {
                    const long long hw = (long long)L.H * L.W;
                    float* o = (float*)L.out + ((long long)n * L.cout) * hw + (long long)(r - 1) * L.W + (c - 1);
#pragma unroll
                    for (int i = 0; i < 16; ++i)
                        if (co0 + i < L.cout) o[(long long)(co0 + i) * hw] = v[i];
                } else {
                    // channel counts are multiples of 8: 16-byte vector stores of 8 channels
                    bf16* o = (bf16*)L.out + opos * L.out_ctot + L.out_c0 + co0;
#pragma unroll
                    for (int h8 = 0; h8 < 2; ++h8) {
                        if (co0 + h8 * 8 >= L.cout) break;
                        uint4* dst = reinterpret_cast<uint4*>(o + h8 * 8);
                        __nv_bfloat162 pk[4];
                        if (accum) {
                            const uint4 old = *dst;
                            const __nv_bfloat162* ob = reinterpret_cast<const __nv_bfloat162*>(&old);
#pragma unroll
                            for (int i = 0; i < 4; ++i) {
                                const float2 f = __bfloat1622float2(ob[i]);
                                pk[i] = __floats2bfloat162_rn(f.x + v[h8 * 8 + 2 * i], f.y + v[h8 * 8 + 2 * i + 1]);
                            }
                        } else {
#pragma unroll
                            for (int i = 0; i < 4; ++i) pk[i] = __floats2bfloat162_rn(v[h8 * 8 + 2 * i], v[h8 * 8 + 2 * i + 1]);
                        }
                        *dst = *reinterpret_cast<const uint4*>(pk);
                    }
                }
            }
        }
    }
    asm volatile("tcgen05.fence::before_thread_sync;" ::: "memory");
    __syncthreads();
    if (warp == 1) asm volatile("tcgen05.dealloc.cta_group::1.sync.aligned.b32 %0, %1;" ::"r"(tmem), "r"(L.tmem_cols));
}

__global__ void pack_filters_kernel(const float* __restrict__ w, bf16* __restrict__ packed, int nreal, int npad, int kreal, int kp,
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
        packed[i] = __float2bfloat16(v);
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

void make_map_2d(CUtensorMap* m, const void* base, long long rows, int ctot, int kc, int box_rows)
{
    cuuint64_t dims[2] = {(cuuint64_t)ctot, (cuuint64_t)rows};
    cuuint64_t strides[1] = {(cuuint64_t)ctot * 2};
    cuuint32_t box[2] = {(cuuint32_t)kc, (cuuint32_t)box_rows}, es[2] = {1, 1};
    const CUtensorMapSwizzle sw = kc == 64 ? CU_TENSOR_MAP_SWIZZLE_128B : kc == 32 ? CU_TENSOR_MAP_SWIZZLE_64B : CU_TENSOR_MAP_SWIZZLE_32B;
    const CUresult r = encode_fn()(m, CU_TENSOR_MAP_DATA_TYPE_BFLOAT16, 2, const_cast<void*>(base), dims, strides, box, es,
                                   CU_TENSOR_MAP_INTERLEAVE_NONE, sw, CU_TENSOR_MAP_L2_PROMOTION_L2_128B, CU_TENSOR_MAP_FLOAT_OOB_FILL_NONE);
    if (r != CUDA_SUCCESS)
        throw std::runtime_error("cuTensorMapEncodeTiled failed (" + std::to_string((int)r) + "): rows " + std::to_string(rows) + " ctot " +
                                 std::to_string(ctot) + " kc " + std::to_string(kc) + " box_rows " + std::to_string(box_rows));
}

}  // namespace

int padded_n(int cout) { return (cout + 15) / 16 * 16; }
int padded_k(int cin) { return (cin + 63) / 64 * 64; }

size_t packed_filter_elems(int cin, int npad_total, int ntaps) { return (size_t)ntaps * npad_total * padded_k(cin); }

void pack_filters_bf16(const float* w, bf16* packed, int nreal, int npad_total, int kreal, int ntaps, const int* tapmap, long long sn,
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
        throw std::invalid_argument("the bf16 tensor-core path needs channel counts that are multiples of 8");
    L.a_c0 = s.a_c0;
    L.ntaps = s.ntaps;
    L.npad_total = s.npad_total;
    // N block: whole N when it fits one MMA (<= 256), else the smallest even split
    int nblocks = (s.npad_total + 255) / 256;
    L.nblk = ((s.npad_total + nblocks - 1) / nblocks + 15) / 16 * 16;
    nblocks = (s.npad_total + L.nblk - 1) / L.nblk;
    if (nblocks * L.nblk != s.npad_total) throw std::logic_error("N blocks must tile the padded N");
    const int P = s.grid.P();
    L.total = (long long)s.grid.N * s.grid.S();
    L.S = s.grid.S(); L.P = P; L.H = s.grid.H; L.W = s.grid.W;
    int dylo = 0, dyhi = 0, dxlo = 0, dxhi = 0, lo = 0, hi = 0;
    for (int t = 0; t < s.ntaps; ++t) {
        dylo = std::min(dylo, s.tap_dy[t]); dyhi = std::max(dyhi, s.tap_dy[t]);
        dxlo = std::min(dxlo, s.tap_dx[t]); dxhi = std::max(dxhi, s.tap_dx[t]);
        const int off = s.tap_dy[t] * P + s.tap_dx[t];
        lo = std::min(lo, off); hi = std::max(hi, off);
    }
    const long long rows_total = (long long)s.grid.N * (s.grid.H + 1);
    const long long tiles_flat = (L.total + 127) / 128;
    const int ncx = (P + 127) / 128;

    // candidate search: K chunk, tiling mode, M tiles per CTA (see the cost model in DESIGN.md)
    struct Cand { double cost; int kc, rows_mode, th, nseg, seg_rows, nloads, box_rows, stages; unsigned a_bytes, stage_bytes; long long ctas; };
    Cand best{1e300, 0, 0, 0, 0, 0, 0, 0, 0, 0, 0, 0};
    for (int kc : {16, 32, 64}) {
        const int kpad = (s.cin + kc - 1) / kc * kc, nchunks = kpad / kc, pitch = kc * 2;
        for (int rows_mode = 0; rows_mode < 2; ++rows_mode) {
            if (rows_mode && P < 128) continue;
            for (int th : {8, 4, 2, 1}) {
                if (th * L.nblk > 512) continue;
                const int nseg = rows_mode ? th + dyhi - dylo : 1;
                if (nseg > 10) continue;
                const int seg_rows = rows_mode ? 128 + dxhi - dxlo : 128 * th + hi - lo;
                const int nloads = (seg_rows + 255) / 256;
                const int box_rows = ((seg_rows + nloads - 1) / nloads + 7) / 8 * 8;
                const unsigned a_bytes = (unsigned)(((size_t)nseg * nloads * box_rows * pitch + 1023) / 1024 * 1024);
                const unsigned stage = a_bytes + (unsigned)(((size_t)s.ntaps * L.nblk * pitch + 1023) / 1024 * 1024);
                const size_t budget = (th * L.nblk <= 256 ? 100u : 200u) * 1024u;
                const int want = std::min(nchunks, 2);
                if ((size_t)stage * want + 1024 > budget) continue;
                const int stages = (int)std::min<size_t>(std::min(nchunks, kMaxStages), (budget - 1024) / stage);
                const long long ctas = rows_mode ? ((rows_total + th - 1) / th) * ncx : (tiles_flat + th - 1) / th;
                const double waste = rows_mode ? (double)ncx * 128 / P * ((double)((rows_total + th - 1) / th) * th / rows_total)
                                               : (double)((tiles_flat + th - 1) / th) * th / tiles_flat;
                const double loadf = (double)nseg * nloads * box_rows / (128.0 * th);
                double cost = waste * kpad * (s.ntaps + 1.0 * loadf);
                if (ctas < k::kNumSMs) cost *= (double)k::kNumSMs / ctas;
                cost *= 1.0 + 0.02 * (kc == 16 ? 2 : kc == 32 ? 1 : 0);  // ties: fewer, larger TMA boxes
                if (cost < best.cost) best = Cand{cost, kc, rows_mode, th, nseg, seg_rows, nloads, box_rows, stages, a_bytes, stage, ctas};
            }
        }
    }
    if (best.kc == 0) throw std::logic_error("no tcgen05 conv tiling fits shared memory");
    L.kc = best.kc;
    L.nchunks = (s.cin + L.kc - 1) / L.kc;
    L.rows_mode = best.rows_mode; L.th = best.th; L.nseg = best.nseg; L.nloads = best.nloads; L.box_rows = best.box_rows;
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
    int cols = 32;
    while (cols < L.th * L.nblk) cols *= 2;
    L.tmem_cols = cols;
    L.grid_x = best.ctas;
    static const bool debug = getenv("AB_TC_DEBUG") != nullptr;
    if (debug)
        fprintf(stderr, "tc conv: P %d cin %d npad %d taps %d -> kc %d %s th %d nseg %d seg %d stages %d smem %u ctas %lld cost %.1f\n", P, s.cin,
                s.npad_total, s.ntaps, L.kc, L.rows_mode ? "ROWS" : "FLAT", L.th, L.nseg, L.seg_alloc, L.stages, L.smem_bytes, best.ctas, best.cost);
    make_map_2d(&L.tmA, s.src, s.src_positions, s.src_ctot, L.kc, L.box_rows);
    make_map_2d(&L.tmB, s.packed, (long long)s.ntaps * s.npad_total, padded_k(s.cin), L.kc, L.nblk);
}

template <int NTAPS>
static void launch_conv_t(const ConvLaunch& L, cudaStream_t st)
{
    static bool attr_set = false;
    if (!attr_set) {
        cudaFuncSetAttribute(conv_tc_kernel<NTAPS>, cudaFuncAttributeMaxDynamicSharedMemorySize, 220 * 1024);
        attr_set = true;
    }
    dim3 grid((unsigned)L.grid_x, (unsigned)(L.npad_total / L.nblk));
    conv_tc_kernel<NTAPS><<<grid, kThreads, L.smem_bytes, st>>>(L);
}

void launch_conv(const ConvLaunch& L, cudaStream_t st)
{
    k::ProfScope ps(k::F_CONV_TC, L.flops, L.bytes, st);
    switch (L.ntaps) {
    case 1: launch_conv_t<1>(L, st); break;
    case 2: launch_conv_t<2>(L, st); break;
    case 4: launch_conv_t<4>(L, st); break;
    case 9: launch_conv_t<9>(L, st); break;
    default: throw std::logic_error("unsupported tap count");
    }
    k::count_launch();
}

}  // namespace tc
}  // namespace ab
