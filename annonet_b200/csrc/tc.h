// Tensor-core path (AB_PRECISION_F16): tcgen05 / TMEM / TMA implicit-GEMM convolutions (kind::f16, fp16 operands, fp32
// accumulation in TMEM) and the fp16 NHWC kernels around them.
//
// Why fp16 and not bf16: the net's 42 stacked batch norms amplify a forward rounding error of relative size e into a
// parameter-gradient error of ~100 e at the deepest layers (measured and reproduced on the CPU by
// experiments/bf16_emulation.py; DESIGN.md "Precision").  bf16 storage (e = 2^-9) leaves those gradients 20-35 % off, fp16
// storage alone 3-8 %.  The TRAINING forward pass therefore keeps every activation as a PAIR of fp16 values (hi + lo, 22
// significant bits) and every filter as a pair as well.  The forward convolutions then compute (A_hi + A_lo)(W_hi + W_lo) as a
// GEMM with the filter pair side by side along N ([W_hi | W_lo], N doubled) and the activation pair as two K chunks; the
// epilogue adds the two column halves -- the tensor pipe has the headroom, the layers are HBM / epilogue bound.  The
// backward pass (gradients, filter gradients) and inference use single fp16 values; gradients are multiplied by a
// power-of-two loss scale so that they sit in fp16's normal range.
//
// Activation layout ("padded flat NHWC"): a tensor of N samples, H x W pixels, C channels is stored as
//     h16 [N*S + P + 1 positions][C],   P = W + 1,  S = (H + 1) * P,
//     position(n, y, x) = n*S + (y + 1)*P + (x + 1),
// i.e. every sample carries one zero row above it and every row one zero pixel to its left.  The pad to
// the right of a row is the left pad of the next row, the pad below a sample is the top pad of the next
// sample (or the P + 1 zero positions at the very end).  Producers never write pad positions, so they
// stay zero from allocation on.  With that layout a 3x3 (or any) convolution is a GEMM over runs of 128
// CONSECUTIVE positions whose tap (dy, dx) operand is the same run shifted by dy*P + dx positions: one
// TMA-loaded halo segment in shared memory serves all nine taps through row-offset shared-memory
// descriptors (the swizzle of tcgen05 descriptors is a function of the absolute shared address --
// measured with experiments/umma_probe.cu -- so any row offset is legal).
#pragma once
#include <cuda.h>
#include <cuda_fp16.h>
#include <cuda_runtime.h>
#include <cstdint>

namespace ab {
namespace tc {

typedef __half h16;

struct Geo {            // geometry of one padded flat tensor
    int N = 0, H = 0, W = 0;
    int P() const { return W + 1; }
    long long S() const { return (long long)(H + 1) * (W + 1); }
    long long positions() const { return (long long)N * S() + P() + 1; }  // allocated / tensor-map rows
    long long pos(int n, int y, int x) const { return n * S() + (long long)(y + 1) * P() + (x + 1); }
};

enum OutMode {
    OUT_BF16 = 0,        // out[pos][c0 + co] = act(acc + bias)            (same grid)
    OUT_BF16_ACCUM = 1,  // out[pos][c0 + co] += acc                        (gradient accumulation)
    OUT_F32_NCHW = 2,    // planes[n][co][y][x] = acc + bias                (class head -> logits)
    OUT_BF16_UP = 3,     // out[fine pos (2r + a, 2c + b)][c0 + co] = acc + bias   (cont_ stride 2, one parity class)
    OUT_BF16_UP_ACCUM = 4,
    OUT_BF16_DOWN = 7,       // as OUT_BF16_DOWN_ACCUM, but assigns (first writer of the slice in a backward pass)
    OUT_F32_NHWC = 6,        // out32[pos][c0 + co] = acc + bias    (training: the fp32 conv output z that feeds a batch norm)
    OUT_BF16_DOWN_ACCUM = 5  // out[coarse pos ((y-1)/2, (x-1)/2)][c0 + co] += acc at odd (y, x) of the fine M grid: backward-data of
                             // the stride-2 cont_ (layers.h:542) evaluated as a stride-1 conv on the fine grid
};
enum Act { ACT_NONE = 0, ACT_MISH_AFFINE = 1, ACT_AFFINE_MISH = 2 };

struct ConvLaunch {
    alignas(64) CUtensorMap tmA;  // source positions x channels, box {kc, box_rows}
    alignas(64) CUtensorMap tmB;  // packed filters [tap][npad_total][padded K], box {kc, nblk}
    int a_c0 = 0;                 // first channel of the input slice
    int nchunks = 0, kc = 16;     // K = nchunks * kc channels per tap (filters zero-padded beyond the real Cin)
    // Split operands (training forward): nchunks = 2 * nch1.  Chunk c of part p (0: hi halves, 1: lo halves) reads the A box at
    // channel a_c0 + c*kc, row offset p * a_lo_rows -- the lo halves live a_lo_rows positions behind the hi halves in the same
    // allocation -- and the SAME B box (K coordinate c*kc) of the pack whose N rows are [W_hi (nsum_stride rows) | W_lo]: the
    // accumulator holds (A_hi + A_lo) W_hi in columns [0, nsum_stride) and (A_hi + A_lo) W_lo behind them, the epilogue adds
    // them.  Unsplit: nch1 = nchunks, nsum = 1.
    int nch1 = 0;
    long long a_lo_rows = 0;
    int nsum = 1, nsum_stride = 0;   // column groups of one accumulator tile the epilogue adds; their distance
    int ntaps = 0;
    // Tiling of the positions.  The tensor is a 2-D array of [N*(H+1) rows][P columns]; a CTA owns `th` M tiles of
    // 128 positions: FLAT: consecutive runs (tile m starts at o0 + 128*m), one merged halo segment;
    // ROWS: the same 128 columns of `th` consecutive rows (tile m starts at o0 + m*P), one 130-position
    // segment per input row, shared by the three dy taps of three output rows.
    int rows_mode = 0;
    // ROWS + merged dy taps ("kym", 3x3 only, taps ordered dx-major): one accumulator tile PER INPUT ROW holding the
    // partial sums of its three dy taps side by side (N = 3*npad), three MMAs (one per dx) per input row and K step
    // instead of nine per output row; the epilogue adds the three row partials (same TMEM lane, no data movement).
    int kym = 0;
    int ncx = 1;                  // ROWS: column tiles per row
    int m_pos_stride = 128;       // position step between the M tiles of a CTA
    int m_row_stride = 128;       // shared-memory row step between the M tiles of a CTA
    int tap_row[9] = {0};         // shared-memory row of tap t for M tile 0
    int nseg = 1;                 // segments per stage
    int seg_pos[10] = {0};        // position of a segment's first row relative to o0
    int seg_alloc = 0;            // rows reserved per segment (= nloads * box_rows)
    int box_rows = 0, nloads = 0; // a segment = nloads TMA boxes of box_rows positions
    int npad_total = 0, nblk = 0; // N of the GEMM (multiple of 16) and the part of it one CTA owns (<= 256)
    int th = 1;                   // 128-position M tiles per CTA
    int stages = 1;
    int tmem_cols = 32;           // allocated TMEM columns (power of two)
    int cols_buf = 0, nbuf = 1;   // accumulator columns of one work item; 2 buffers when they fit
    unsigned smem_bytes = 0, a_stage_bytes = 0, stage_bytes = 0;
    // geometry of the M grid (source/positions) and of the output
    long long total = 0;          // N * S of the M grid
    long long S = 0; int P = 0, H = 0, W = 0, Nsamples = 0;
    int mode = OUT_BF16, act = ACT_NONE;
    int cout = 0;                 // real output channels (<= npad_total)
    void* out = nullptr;          // h16 [pos][out_ctot] or f32 planes
    int out_ctot = 0, out_c0 = 0;
    long long out_lo_off = 0;     // OUT_BF16 / OUT_BF16_UP: also store the rounding residual (lo half) this many ELEMENTS behind; 0: none
    long long out_S = 0; int out_P = 0;  // fine grid (OUT_BF16_UP*)
    int par_a = 0, par_b = 0;     // parity class of OUT_BF16_UP*
    const float* bias = nullptr;  // [cout] or null
    const float* scale = nullptr; // affine_/bn gamma-like and beta-like per channel (ACT_*)
    const float* shift = nullptr;
    // fused batch statistics (OUT_F32_NHWC, one N block of 16): stat_sums[c] += sum, stat_sums[cout + c] += sum of squares of the
    // stored values (of mish(value) with stat_mish) over the valid positions -- what bn_stats_nhwc would re-read z for
    double* stat_sums = nullptr;
    int stat_mish = 0;
    int dbg = 0;                  // timing experiments only (AB_TC_DBG): 1 skip the MMAs, 2 skip the TMA loads, 4 skip the stores
    long long grid_x = 0;         // CTAs along the positions
    double flops = 0, bytes = 0;  // algorithmic work (for the profiler)
};

// host: fill tmA/tmB and the derived tiling fields.  src: h16 [src_positions][src_ctot].
struct ConvSetup {
    const h16* src = nullptr; long long src_positions = 0; int src_ctot = 0, a_c0 = 0, cin = 0;
    const h16* packed = nullptr;  // packed filters (see pack_filters_bf16)
    int ntaps = 0; int tap_dy[9] = {0}, tap_dx[9] = {0};  // tap offset = dy*P + dx on the M grid
    int npad_total = 0;
    Geo grid;                      // geometry the M tiles walk over (the source tensor's)
    // split operands: src holds 2 * src_positions rows (hi halves, then lo halves) and `packed` has 2 * npad_total rows per tap:
    // the hi filters, then the lo filters (PackJob::parts = 2)
    int split = 0;
};
int padded_k(int cin);             // multiple of 64
int padded_n(int cout);            // multiple of 16
size_t packed_filter_elems(int cin, int npad_total, int ntaps);
void setup_conv(ConvLaunch& L, const ConvSetup& s);
// rev: the CTAs walk the work items downwards (see bn_stats_nhwc)
void launch_conv(const ConvLaunch& L, cudaStream_t st, int rev = 0);

// packed[tap][n][k] = (n < nreal && k < kreal) ? w[n*sn + k*sk + tapmap[tap]] : 0,   n < npad_total, k < padded_k(kreal)
void pack_filters_bf16(const float* w, h16* packed, int nreal, int npad_total, int kreal, int ntaps, const int* tapmap, long long sn,
                       long long sk, cudaStream_t st);

// ---- h16 NHWC kernels around the convolutions (kernels_nhwc.cu) ---------------------------------
// input_rgb_image::to_tensor (input.h:68-117) [+ tile extraction with clamp-to-edge, annonet_infer.cpp:68-75]:
// u8 HWC images [N][IH][IW][3] -> padded flat NHWC with 8 channels (3 real + 5 zero); out pixel (y, x) of sample n
// reads image pixel (clamp(top + y), clamp(left + x)).
// lo_off != 0: the rounding residual of every value goes to out[lo_off + ...] (the lo half of a split tensor)
void image_to_nhwc(const uint8_t* img, int IH, int IW, int left, int top, h16* out, const Geo& g, float ar, float ag, float ab_,
                   cudaStream_t st, long long lo_off = 0);
// max_pool_<3,3,2,2,0,0> (cpu_dlib.cpp:1888-1943): channels [c0, c0 + C) of `in` (geometry gi, ctot channels) -> dense
// pooled tensor [positions of go][C]
// arg (optional, training): u8 [positions of go][C], the index (r*3 + q) of the window element that won
// in_lo_off / out_lo_off != 0: split tensors -- the maximum is taken over hi + lo and both halves are written
void maxpool_nhwc(const h16* in, int in_ctot, int c0, int C, const Geo& gi, h16* out, const Geo& go, cudaStream_t st, uint8_t* arg = nullptr,
                  long long in_lo_off = 0, long long out_lo_off = 0);

// ---- training-side h16 NHWC kernels (kernels_train_nhwc.cu) ------------------------------------
// A "view" = channels [c0, c0 + C) of a padded flat tensor with ctot channels per position.
struct View {
    h16* p = nullptr;
    int ctot = 0, c0 = 0, C = 0;
    long long lo_off = 0;   // split tensors: element offset of the lo half (0: single fp16 values)
};
struct ViewF {   // same, fp32 elements (the conv outputs z that feed a batch norm stay fp32: normalisation amplifies rounding)
    float* p = nullptr;
    int ctot = 0, c0 = 0, C = 0;
};
// sums[c] = sum x, sums[C + c] = sum x^2 over all positions of x = mish_first ? mish(z) : z (pads are zero and add nothing)
// rev (all bn passes): walk the positions downwards, so that a pass starts where its producer stopped (L2 reuse)
// pre_zeroed: `sums` / `red` already holds zeros (the engine clears every layer's slot with ONE memset per step: a memset node
// in front of each launch would cut the programmatic launch dependency, kcommon.cuh)
void bn_stats_nhwc(const ViewF& z, const Geo& g, int mish_first, double* sums, cudaStream_t st, int rev = 0, int pre_zeroed = 0);
// y = mish_first ? gamma*(mish(z)-mean)*invstd+beta : mish(gamma*(z-mean)*invstd+beta)   [mish optional]; pads are left alone
// Optional fused bn_finalize (kernels_bn.cu): every CTA derives mean / invstd from the layer's sums itself, CTA 0 also
// updates the running statistics, the saved stats and the device-side num_updates counter (layers.h:1392-1397).
struct BnFinalize {
    const double* sums = nullptr;   // [2][C] sum, sum of squares (null: read `stats`)
    double m = 0, eps = 0;
    unsigned long long* num_updates = nullptr;
    unsigned long long window = 0;
    float* rmean = nullptr; float* rvar = nullptr;
    float* stats_out = nullptr;     // [2][C] mean, invstd for the backward pass
};
void bn_apply_nhwc(const ViewF& z, const View& y, const Geo& g, int mish_first, int has_mish, const float* gamma, const float* beta,
                   const float* stats, cudaStream_t st, const BnFinalize* fin = nullptr, int rev = 0);
// backward pass 1 (cpu_dlib.cpp:1228-1250): red[c] = sum dv, red[C + c] = sum dv*xhat
void bn_backward_reduce_nhwc(const ViewF& z, const View& gy, const Geo& g, int mish_first, int has_mish, const float* gamma,
                             const float* beta, const float* stats, double* red, cudaStream_t st, int rev = 0, int pre_zeroed = 0);
// backward pass 2 (cpu_dlib.cpp:1252-1288): dz = gradient at the conv output (h16, pads untouched), dgamma/dbeta from `red`,
// dbias[c] = sum dz (assign_conv_bias_gradient, cpu_dlib.cpp:470-504; dbias must be zeroed by the caller).  gy and dz carry
// the loss scale; the three parameter gradients are multiplied by `unscale` = 1 / loss scale
void bn_backward_apply_nhwc(const ViewF& z, const View& gy, const View& dz, const Geo& g, int mish_first, int has_mish, const float* gamma,
                            const float* beta, const float* stats, const double* red, float* dgamma, float* dbeta, float* dbias,
                            cudaStream_t st, int rev = 0, float unscale = 1.f);
// g[c] += unscale * sum over positions of v (caller zeroes g)
void channel_sum_nhwc(const View& v, const Geo& g, float* out, cudaStream_t st, float unscale = 1.f);
// max_pool_<3,3,2,2,0,0> gradient (cpu_dlib.cpp:1971-2017): the first maximum of a window in row-major order receives it; gin +=
// (`arg` as written by maxpool_nhwc in the forward pass)
void maxpool_backward_nhwc(const uint8_t* arg, int C, const Geo& gi, const h16* gout_dense, const Geo& go, const View& gin, cudaStream_t st);
// f32 planes [N][K][H][W] * scale -> h16 padded flat [positions][K rounded up to 8] (channels >= K are zero), K <= 64
void planes_to_nhwc8(const float* planes, int K, h16* out, const Geo& g, cudaStream_t st, float scale = 1.f);
// concat_prev_'s bilinear resize (layers.h:2494-2521; plan.h resize_axis): `in` (geometry gi, dense [positions][C], split when
// in_lo_off != 0) -> channels of `out` (geometry go; both halves when out.lo_off != 0).  tab: device table as k::ResizeTab.
struct ResizeTabDev { const int* top; const int* bottom; const int* left; const int* right; const float* tb; const float* itb; const float* lr; const float* ilr; };
void resize_bilinear_nhwc(const h16* in, long long in_lo_off, int C, const Geo& gi, const View& out, const Geo& go, const ResizeTabDev& t, cudaStream_t st);
// out (fine geometry, dense [positions][C]) = zero-stuffed in: out[2i][2j] = in[i][j], 0 at the other pixels
void upsample_zero_nhwc(const View& in, const Geo& gc, h16* out, const Geo& gf, cudaStream_t st);

// one filter-pack job of pack_filters_batched: same element mapping as pack_filters_bf16
struct PackJob {
    long long w_off, packed_off;   // floats into the arena, h16 elements into the pack buffer
    long long sn, sk;
    int nreal, npad, kreal, kp, ntaps;
    int tapmap[9];
    int parts;                     // 1: [tap][n][kp] of fp16(w);  2: [tap][2*npad][kp]: rows n < npad hi = fp16(w), rows npad + n lo = fp16(w - hi)
};
void pack_filters_batched(const float* arena, h16* packs, const PackJob* jobs_dev, int njobs, int max_elems, cudaStream_t st);

// ---- filter gradients on the tensor cores (kernels_wgrad_tc.cu) -----------------------------------
//   G[tap][y][x] = sum_pos DY[pos][dy_c0 + y] * X[pos + tap_dy*P + tap_dx][x_c0 + x]      y < cy <= 64, x < cx
// over ALL positions of geometry g (pads of both tensors are zero), scattered to out[y*sy + x*sx + tapmap[tap]].
struct WgradJob {
    const h16* dy = nullptr; int dy_ctot = 0, dy_c0 = 0, cy = 0;   // cy, cx: loaded channels (multiples of 8)
    const h16* x = nullptr;  int x_ctot = 0, x_c0 = 0, cx = 0;
    int cy_real = 0, cx_real = 0;                                    // channels written to `out` (<= cy, cx)
    Geo g;
    int ntaps = 0, tap_dy[9] = {0}, tap_dx[9] = {0}, tapmap[9] = {0};
    float* out = nullptr; long long sy = 0, sx = 0;
    float unscale = 1.f;            // out = unscale * G (1 / loss scale)
};
// h16 [rows][ctot] tensor, box {box_ch (16 / 32 / 64: 32 / 64 / 128-byte swizzle), box_rows}
void make_tensor_map_2d(CUtensorMap* m, const void* base, long long rows, int ctot, int box_ch, int box_rows);
size_t wgrad_tc_scratch_floats();

// tcgen05 (kernels_wgrad_tc.cu): MN-major shared-memory descriptors over TMA-staged [position][64 channel]
// tiles, two taps fused into one M = 128 MMA through the descriptor's leading-dimension offset, accumulators in TMEM.
struct WgradTcLaunch {
    alignas(64) CUtensorMap tmX, tmY;
    int x_c0 = 0, dy_c0 = 0;
    int th = 1, kw = 64;              // a chunk = th runs of kw positions; run m starts at o0 + m*run_stride
    int run_stride = 0;
    int rows_mode = 0, ncx = 1, P = 0;
    int nseg = 1, seg_alloc = 0;      // X segments of a chunk (seg_alloc rows reserved each, multiple of 8)
    int seg_pos[12] = {0};
    int npairs = 0;                   // accumulators: tap pair (t1, t2) in rows 0-63 / 64-127
    int pair_seg[5] = {0}, pair_off[5] = {0}, pair_lbo[5] = {0};   // segment (relative to the run) and row offset of t1; byte distance to t2
    int pair_t1[5] = {0}, pair_t2[5] = {0};                          // tap ids (t2 = -1: none)
    int N = 16;                       // padded cy
    // 3x3, cy <= 16 ("shift" mode): the three dx taps ride on B instead of A -- DY is staged three times, shifted by 0, 1, 2
    // positions, as the three 16-channel atoms of one N = 48 operand, so one MMA against X[p + dy*P + 1] yields dx = +1, 0, -1
    // at once: two MMAs (dy pair, single dy) per 16 positions instead of five.
    int shift = 0, Nmma = 16;
    int sh_tap[5][2][3] = {};         // [accumulator][rows 0-63 / 64-127][shift] -> tap id or -1
    int stages = 2, tmem_cols = 64;
    unsigned x_stage_bytes = 0, stage_bytes = 0, smem_bytes = 0;
    long long chunks = 0;
    int nsplit = 1, cxtiles = 1, ntaps = 0;
    float* scratch = nullptr;
    // reduction + scatter (same as WgradJob)
    float* out = nullptr; long long sy = 0, sx = 0; int tapmap[9] = {0}; int cy_real = 0, cx_real = 0;
    float unscale = 1.f;
    double flops = 0, bytes = 0;
};
void setup_wgrad_tc(WgradTcLaunch& L, const WgradJob& job, float* scratch);
void launch_wgrad_tc(const WgradTcLaunch& L, cudaStream_t st);

}  // namespace tc
}  // namespace ab
