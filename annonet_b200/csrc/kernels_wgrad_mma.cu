// Filter gradients of con_/cont_ (cpu_dlib.cpp:2327-2355, layers.h:222,547) on the bf16 path:
//
//     G[tap][y][x] = sum_pos DY[pos][y] * X[pos + dy[tap]*P + dx[tap]][x]            (fp32 accumulate)
//
// over the padded flat NHWC tensors of tc.h -- a GEMM whose K dimension is the POSITIONS, so both operands
// are "MN-major" ([pos][channel]).  This version runs on the warp-level tensor-core path (mma.sync m16n8k16
// bf16, operands transposed on the fly by ldmatrix.trans); a tcgen05 version with MN-major shared-memory
// descriptors is the planned replacement (DESIGN.md "next").
//
// A CTA owns one 64-channel tile of X's channels, all cy <= 64 channels of DY, all taps, and a contiguous
// range of position chunks (split-K).  Per chunk the DY runs and the X halo segments are staged with cp.async
// (two stages); the tap shift is a row offset into the staged segment, exactly as in the forward kernel.
// Warp w: x-channels [16*(w&3), +16), taps of half (w>>2).  Partial sums go to scratch[split][tile][tap][64][64];
// a second kernel adds the splits and scatters into the gradient arena in dlib's filter layout.
#include "kcommon.cuh"
#include "tc.h"

#include <stdexcept>

namespace ab {
namespace tc {
namespace {

constexpr int kThreads = 256;
constexpr int XP = 64 * 2 + 16;   // shared-memory pitch of an X row (64 channels + 16 B: conflict-free ldmatrix)
constexpr int kMaxSplitTiles = 148;

struct WgradParams {
    const bf16* dy; int dy_ctot, dy_c0, cy;
    const bf16* x;  int x_ctot, x_c0, cx;
    long long total;     // N*S: DY positions at or beyond it read as zero
    long long x_alloc;   // X positions outside [0, x_alloc) read as zero
    int P;
    int rows_mode, th, kw, ncx;   // a chunk = th runs of kw positions (ROWS: the same kw columns of th consecutive rows)
    int nseg, seg_rows;
    int seg_pos[12];
    int tap_seg[9], tap_off[9];
    long long chunks;
    int nsplit, cxtiles;
    float* scratch;
    float* out; long long sy, sx; int tapmap[9]; int ntaps; int cy_real, cx_real;
};

__device__ __forceinline__ uint32_t smem_u32(const void* p) { return (uint32_t)__cvta_generic_to_shared(p); }
__device__ __forceinline__ void cp_async16(uint32_t dst, const void* src, bool valid)
{
    const int sz = valid ? 16 : 0;
    asm volatile("cp.async.cg.shared.global [%0], [%1], 16, %2;" ::"r"(dst), "l"(src), "r"(sz) : "memory");
}
__device__ __forceinline__ void ldsm_x4_t(uint32_t addr, uint32_t (&r)[4])
{
    asm volatile("ldmatrix.sync.aligned.m8n8.x4.trans.shared.b16 {%0,%1,%2,%3}, [%4];"
                 : "=r"(r[0]), "=r"(r[1]), "=r"(r[2]), "=r"(r[3]) : "r"(addr));
}
__device__ __forceinline__ void mma_bf16(float (&c)[4], const uint32_t (&a)[4], uint32_t b0, uint32_t b1)
{
    asm volatile("mma.sync.aligned.m16n8k16.row.col.f32.bf16.bf16.f32 {%0,%1,%2,%3}, {%4,%5,%6,%7}, {%8,%9}, {%0,%1,%2,%3};"
                 : "+f"(c[0]), "+f"(c[1]), "+f"(c[2]), "+f"(c[3])
                 : "r"(a[0]), "r"(a[1]), "r"(a[2]), "r"(a[3]), "r"(b0), "r"(b1));
}

template <int MT, int NTAPS>
__global__ void __launch_bounds__(kThreads, 1) wgrad_mma_kernel(const __grid_constant__ WgradParams p)
{
    constexpr int TPW = (NTAPS + 1) / 2;     // taps per warp
    constexpr int YP = MT * 32 + 16;         // pitch of a DY row
    extern __shared__ __align__(128) uint8_t smem[];
    const int tid = threadIdx.x, warp = tid >> 5, lane = tid & 31;
    const int split = blockIdx.x, cxt = blockIdx.y, cx0 = cxt * 64;
    const int cxv = min(64, p.cx - cx0);
    const int dy_rows = p.th * p.kw, x_rows = p.nseg * p.seg_rows;
    const uint32_t x_bytes = (uint32_t)(x_rows * XP + 127) & ~127u, stage_bytes = x_bytes + (((uint32_t)(dy_rows * YP) + 127) & ~127u);
    const uint32_t s0 = smem_u32(smem);

    const long long c_begin = p.chunks * split / p.nsplit, c_end = p.chunks * (split + 1) / p.nsplit;

    auto load_chunk = [&](long long chunk, int stage) {
        long long o0;
        int col0 = 0;
        if (p.rows_mode) {
            const long long rg = chunk / p.ncx;
            col0 = (int)(chunk - rg * p.ncx) * p.kw;
            o0 = rg * p.th * p.P + col0;
        } else o0 = chunk * p.kw;
        const uint32_t xs = s0 + (uint32_t)stage * stage_bytes, ys = xs + x_bytes;
        for (int i = tid; i < x_rows * 8; i += kThreads) {
            const int row = i >> 3, v = i & 7;
            const int sg = row / p.seg_rows, j = row - sg * p.seg_rows;
            const long long pos = o0 + p.seg_pos[sg] + j;
            const bool ok = pos >= 0 && pos < p.x_alloc && v * 8 < cxv;
            cp_async16(xs + (uint32_t)(row * XP + v * 16), ok ? (const void*)(p.x + pos * p.x_ctot + p.x_c0 + cx0 + v * 8) : (const void*)p.x, ok);
        }
        for (int i = tid; i < dy_rows * (MT * 2); i += kThreads) {
            const int row = i / (MT * 2), v = i - row * (MT * 2);
            const int m = row / p.kw, j = row - m * p.kw;
            const long long pos = o0 + (p.rows_mode ? (long long)m * p.P : 0) + j;
            const bool ok = pos < p.total && v * 8 < p.cy && (!p.rows_mode || col0 + j < p.P);
            cp_async16(ys + (uint32_t)(row * YP + v * 16), ok ? (const void*)(p.dy + pos * p.dy_ctot + p.dy_c0 + v * 8) : (const void*)p.dy, ok);
        }
        asm volatile("cp.async.commit_group;" ::: "memory");
    };

    float acc[TPW][MT][2][4];
#pragma unroll
    for (int a = 0; a < TPW; ++a)
#pragma unroll
        for (int b = 0; b < MT; ++b)
#pragma unroll
            for (int c = 0; c < 2; ++c)
#pragma unroll
                for (int d = 0; d < 4; ++d) acc[a][b][c][d] = 0.f;

    const int cxg = warp & 3, tbeg = (warp >> 2) * TPW;
    const bool active = cxg * 16 < cxv && tbeg < NTAPS;
    // ldmatrix lane -> row / column offsets (see the fragment layouts of mma.m16n8k16)
    const int a_k = (lane & 7) + ((lane >> 4) & 1) * 8, a_m = ((lane >> 3) & 1) * 8;
    const int b_k = (lane & 7) + ((lane >> 3) & 1) * 8, b_n = ((lane >> 4) & 1) * 8;
    const uint32_t a_lane = (uint32_t)(a_k * YP + a_m * 2), b_lane = (uint32_t)(b_k * XP + (cxg * 16 + b_n) * 2);
    int t_seg[TPW], t_off[TPW];
#pragma unroll
    for (int ti = 0; ti < TPW; ++ti) {
        const int t = min(tbeg + ti, NTAPS - 1);
        t_seg[ti] = p.tap_seg[t];
        t_off[ti] = p.tap_off[t];
    }
    const int ksteps = p.kw / 16;

    if (c_begin < c_end) load_chunk(c_begin, 0);
    for (long long chunk = c_begin; chunk < c_end; ++chunk) {
        const int stage = (int)((chunk - c_begin) & 1);
        if (chunk + 1 < c_end) {
            load_chunk(chunk + 1, stage ^ 1);
            asm volatile("cp.async.wait_group 1;" ::: "memory");
        } else {
            asm volatile("cp.async.wait_group 0;" ::: "memory");
        }
        __syncthreads();
        if (active) {
            const uint32_t xs = s0 + (uint32_t)stage * stage_bytes, ys = xs + x_bytes;
            for (int m = 0; m < p.th; ++m) {
                for (int kk = 0; kk < ksteps; ++kk) {
                    uint32_t a[MT][4];
                    const uint32_t ya = ys + (uint32_t)((m * p.kw + kk * 16) * YP) + a_lane;
#pragma unroll
                    for (int mt = 0; mt < MT; ++mt) ldsm_x4_t(ya + mt * 32, a[mt]);
#pragma unroll
                    for (int ti = 0; ti < TPW; ++ti) {
                        if (tbeg + ti < NTAPS) {
                            uint32_t b[4];
                            ldsm_x4_t(xs + (uint32_t)(((m + t_seg[ti]) * p.seg_rows + t_off[ti] + kk * 16) * XP) + b_lane, b);
#pragma unroll
                            for (int mt = 0; mt < MT; ++mt) {
                                mma_bf16(acc[ti][mt][0], a[mt], b[0], b[1]);
                                mma_bf16(acc[ti][mt][1], a[mt], b[2], b[3]);
                            }
                        }
                    }
                }
            }
        }
        __syncthreads();
    }

    if (active) {
        float* sc = p.scratch + ((size_t)split * p.cxtiles + cxt) * NTAPS * 4096;
#pragma unroll
        for (int ti = 0; ti < TPW; ++ti) {
            if (tbeg + ti >= NTAPS) continue;
            float* st = sc + (size_t)(tbeg + ti) * 4096;
#pragma unroll
            for (int mt = 0; mt < MT; ++mt)
#pragma unroll
                for (int nt = 0; nt < 2; ++nt) {
                    const int y = mt * 16 + (lane >> 2), xx = cxg * 16 + nt * 8 + (lane & 3) * 2;
                    *reinterpret_cast<float2*>(st + y * 64 + xx) = make_float2(acc[ti][mt][nt][0], acc[ti][mt][nt][1]);
                    *reinterpret_cast<float2*>(st + (y + 8) * 64 + xx) = make_float2(acc[ti][mt][nt][2], acc[ti][mt][nt][3]);
                }
        }
    }
}

__global__ void __launch_bounds__(256) wgrad_reduce_kernel(const __grid_constant__ WgradParams p)
{
    const long long total = (long long)p.ntaps * p.cy_real * p.cx_real;
    const long long i = (long long)blockIdx.x * blockDim.x + threadIdx.x;
    if (i >= total) return;
    const int xg = (int)(i % p.cx_real);
    const int y = (int)((i / p.cx_real) % p.cy_real);
    const int t = (int)(i / ((long long)p.cx_real * p.cy_real));
    const int cxt = xg >> 6, x = xg & 63;
    float s = 0.f;
    for (int sp = 0; sp < p.nsplit; ++sp) s += p.scratch[(((size_t)sp * p.cxtiles + cxt) * p.ntaps + t) * 4096 + y * 64 + x];
    p.out[y * p.sy + xg * p.sx + p.tapmap[t]] = s;
}

template <int MT, int NTAPS>
void launch(const WgradParams& p, size_t smem, cudaStream_t st)
{
    static bool attr_set = false;
    if (!attr_set) {
        cudaFuncSetAttribute(wgrad_mma_kernel<MT, NTAPS>, cudaFuncAttributeMaxDynamicSharedMemorySize, 220 * 1024);
        attr_set = true;
    }
    wgrad_mma_kernel<MT, NTAPS><<<dim3(p.nsplit, p.cxtiles), kThreads, smem, st>>>(p);
}

template <int NTAPS>
void launch_mt(int mt, const WgradParams& p, size_t smem, cudaStream_t st)
{
    switch (mt) {
    case 1: launch<1, NTAPS>(p, smem, st); break;
    case 2: launch<2, NTAPS>(p, smem, st); break;
    case 3: launch<3, NTAPS>(p, smem, st); break;
    case 4: launch<4, NTAPS>(p, smem, st); break;
    default: throw std::logic_error("wgrad_mma: cy must be <= 64");
    }
}

}  // namespace

size_t wgrad_mma_scratch_floats() { return (size_t)kMaxSplitTiles * 9 * 4096; }

void wgrad_mma(const WgradJob& job, float* scratch, cudaStream_t st)
{
    if (job.cy % 8 || job.cx % 8 || job.dy_ctot % 8 || job.x_ctot % 8 || job.dy_c0 % 8 || job.x_c0 % 8)
        throw std::invalid_argument("wgrad_mma needs channel counts that are multiples of 8");
    if (job.cy > 64 || job.cy < 8) throw std::invalid_argument("wgrad_mma: 8 <= cy <= 64");
    if (job.ntaps != 1 && job.ntaps != 9) throw std::invalid_argument("wgrad_mma: 1 or 9 taps");
    WgradParams p{};
    p.dy = job.dy; p.dy_ctot = job.dy_ctot; p.dy_c0 = job.dy_c0; p.cy = job.cy;
    p.x = job.x; p.x_ctot = job.x_ctot; p.x_c0 = job.x_c0; p.cx = job.cx;
    const int P = job.g.P();
    p.P = P;
    p.total = (long long)job.g.N * job.g.S();
    p.x_alloc = job.g.positions();
    p.ntaps = job.ntaps;
    int dylo = 0, dyhi = 0, dxlo = 0, dxhi = 0, lo = 0, hi = 0;
    for (int t = 0; t < job.ntaps; ++t) {
        dylo = std::min(dylo, job.tap_dy[t]); dyhi = std::max(dyhi, job.tap_dy[t]);
        dxlo = std::min(dxlo, job.tap_dx[t]); dxhi = std::max(dxhi, job.tap_dx[t]);
        const int off = job.tap_dy[t] * P + job.tap_dx[t];
        lo = std::min(lo, off); hi = std::max(hi, off);
        p.tapmap[t] = job.tapmap[t];
    }
    const int mt = (job.cy + 15) / 16;
    if (P >= 32 && dyhi - dylo <= 2) {
        p.rows_mode = 1; p.kw = 32; p.th = 8;
        p.ncx = (P + p.kw - 1) / p.kw;
        p.nseg = p.th + dyhi - dylo;
        p.seg_rows = p.kw + dxhi - dxlo;
        for (int s = 0; s < p.nseg; ++s) p.seg_pos[s] = (s + dylo) * P + dxlo;
        for (int t = 0; t < job.ntaps; ++t) { p.tap_seg[t] = job.tap_dy[t] - dylo; p.tap_off[t] = job.tap_dx[t] - dxlo; }
        const long long rows_total = (long long)job.g.N * (job.g.H + 1);
        p.chunks = (rows_total + p.th - 1) / p.th * p.ncx;
    } else {
        p.rows_mode = 0; p.th = 1; p.kw = 256; p.ncx = 1;
        p.nseg = 1;
        p.seg_rows = p.kw + hi - lo;
        p.seg_pos[0] = lo;
        for (int t = 0; t < job.ntaps; ++t) { p.tap_seg[t] = 0; p.tap_off[t] = job.tap_dy[t] * P + job.tap_dx[t] - lo; }
        p.chunks = (p.total + p.kw - 1) / p.kw;
    }
    p.cxtiles = (job.cx + 63) / 64;
    p.nsplit = (int)std::max<long long>(1, std::min<long long>(p.chunks, kMaxSplitTiles / p.cxtiles));
    p.scratch = scratch;
    p.out = job.out; p.sy = job.sy; p.sx = job.sx;
    p.cy_real = job.cy_real ? job.cy_real : job.cy; p.cx_real = job.cx_real ? job.cx_real : job.cx;
    const int YP = mt * 32 + 16;
    const size_t x_bytes = ((size_t)p.nseg * p.seg_rows * XP + 127) / 128 * 128, y_bytes = ((size_t)p.th * p.kw * YP + 127) / 128 * 128;
    const size_t smem = 2 * (x_bytes + y_bytes);
    if (smem > 220 * 1024) throw std::logic_error("wgrad_mma: tile does not fit shared memory");
    const double px = (double)job.g.N * job.g.H * job.g.W;
    {
        k::ProfScope ps(k::F_WGRAD_TC, 2.0 * px * job.cx * job.cy * job.ntaps, 2.0 * px * (job.cx + job.cy), st);
        if (job.ntaps == 9) launch_mt<9>(mt, p, smem, st);
        else launch_mt<1>(mt, p, smem, st);
        const long long total = (long long)job.ntaps * p.cy_real * p.cx_real;
        wgrad_reduce_kernel<<<(unsigned)((total + 255) / 256), 256, 0, st>>>(p);
        k::count_launch(2);
    }
}

}  // namespace tc
}  // namespace ab
