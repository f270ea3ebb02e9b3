// Filter gradients (tensor_conv::get_gradient_for_filters, cpu_dlib.cpp:2327-2355) as a split-K
// implicit GEMM on the FMA pipe:
//     G[a][b][ky][kx] = sum_{n,y,x} A[n,a,y,x] * B[n,b,S*y+ky-P,S*x+kx-P]
// A CTA owns 16 a-channels x 8 b-channels and walks its share of the pixel tiles (split K); warp w
// owns b-channel w, each lane one pixel column, and keeps the full 16 x 9 accumulator tile in
// registers, so the only cross-thread reduction is one warp shuffle tree per CTA at the end.
// Partials of the K splits go to scratch and are summed by a second kernel in a fixed order
// (bitwise reproducible, unlike atomics).
#include "kcommon.cuh"

namespace ab {
namespace k {
namespace {

constexpr int WG_A = 16, WG_B = 8;
constexpr size_t kScratchFloats = 8u << 20;

template <int S, int TH>
__global__ void __launch_bounds__(256) wgrad3x3_kernel(const Slice A, const Slice B, const int N, const int Ha, const int Wa,
                                                       const int Hb, const int Wb, const int P, float* __restrict__ part,
                                                       const int tiles_x, const int tiles_y)
{
    constexpr int BR = S * (TH - 1) + 3, BC = S * 31 + 3;
    __shared__ float As[WG_A][TH][32];
    __shared__ float Bs[WG_B][BR][BC + (BC % 2 == 0 ? 1 : 0)];
    constexpr int BCP = BC + (BC % 2 == 0 ? 1 : 0);
    const int tid = threadIdx.x, lane = tid & 31, warp = tid >> 5;
    const int b0 = blockIdx.y * WG_B, a0 = blockIdx.z * WG_A;
    const long long hwa = (long long)Ha * Wa, hwb = (long long)Hb * Wb;
    const long long total = (long long)N * tiles_x * tiles_y;

    float acc[WG_A][9];
#pragma unroll
    for (int a = 0; a < WG_A; ++a)
#pragma unroll
        for (int t = 0; t < 9; ++t) acc[a][t] = 0.f;

    for (long long tile = blockIdx.x; tile < total; tile += gridDim.x) {
        const int n = (int)(tile / (tiles_x * tiles_y));
        const int tr = (int)(tile % (tiles_x * tiles_y));
        const int y0 = (tr / tiles_x) * TH, x0 = (tr % tiles_x) * 32;
        const float* __restrict__ ab_ = A.p + (long long)n * A.sstride + (long long)A.c0 * hwa;
        const float* __restrict__ bb_ = B.p + (long long)n * B.sstride + (long long)B.c0 * hwb;
        for (int idx = tid; idx < WG_A * TH * 32; idx += 256) {
            const int x = idx & 31, y = (idx >> 5) % TH, a = idx / (32 * TH);
            const int gy = y0 + y, gx = x0 + x;
            float v = 0.f;
            if (a0 + a < A.C && gy < Ha && gx < Wa) v = __ldg(ab_ + (long long)(a0 + a) * hwa + (long long)gy * Wa + gx);
            As[a][y][x] = v;
        }
        const int by0 = S * y0 - P, bx0 = S * x0 - P;
        for (int idx = tid; idx < WG_B * BR * BC; idx += 256) {
            const int x = idx % BC, y = (idx / BC) % BR, b = idx / (BC * BR);
            const int gy = by0 + y, gx = bx0 + x;
            float v = 0.f;
            if (b0 + b < B.C && gy >= 0 && gy < Hb && gx >= 0 && gx < Wb) v = __ldg(bb_ + (long long)(b0 + b) * hwb + (long long)gy * Wb + gx);
            Bs[b][y][x] = v;
        }
        __syncthreads();
#pragma unroll 1
        for (int y = 0; y < TH; ++y) {
            float bv[9];
#pragma unroll
            for (int ky = 0; ky < 3; ++ky)
#pragma unroll
                for (int kx = 0; kx < 3; ++kx) bv[ky * 3 + kx] = Bs[warp][S * y + ky][S * lane + kx];
#pragma unroll
            for (int a = 0; a < WG_A; ++a) {
                const float av = As[a][y][lane];
#pragma unroll
                for (int t = 0; t < 9; ++t) acc[a][t] = fmaf(av, bv[t], acc[a][t]);
            }
        }
        __syncthreads();
    }
    // one shuffle tree per accumulator; lane (i % 32) stores accumulator i
    const int Ca = A.C, Cb = B.C;
    float* __restrict__ dst = part + (long long)blockIdx.x * Ca * Cb * 9;
    const int b = b0 + warp;
#pragma unroll
    for (int a = 0; a < WG_A; ++a)
#pragma unroll
        for (int t = 0; t < 9; ++t) {
            const float s = warp_sum(acc[a][t]);
            if (lane == ((a * 9 + t) & 31) && a0 + a < Ca && b < Cb) dst[((long long)(a0 + a) * Cb + b) * 9 + t] = s;
        }
}

__global__ void __launch_bounds__(256) wgrad1x1_kernel(const Slice A, const Slice B, const int N, const long long hw,
                                                       float* __restrict__ part, const int chunks_per_sample)
{
    const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
    const int b = blockIdx.y * WG_B + warp, a0 = blockIdx.z * WG_A;
    const int Ca = A.C, Cb = B.C;
    float acc[WG_A];
#pragma unroll
    for (int a = 0; a < WG_A; ++a) acc[a] = 0.f;
    const long long total = (long long)N * chunks_per_sample;
    constexpr int CH = 256;  // pixels per chunk
    for (long long ch = blockIdx.x; ch < total; ch += gridDim.x) {
        const int n = (int)(ch / chunks_per_sample);
        const long long p0 = (ch % chunks_per_sample) * CH;
        const float* __restrict__ ab_ = A.p + (long long)n * A.sstride + (long long)A.c0 * hw;
        const float* __restrict__ bb_ = B.p + (long long)n * B.sstride + (long long)(B.c0 + b) * hw;
#pragma unroll 2
        for (int k = 0; k < CH / 32; ++k) {
            const long long p = p0 + k * 32 + lane;
            if (p < hw && b < Cb) {
                const float bv = __ldg(bb_ + p);
#pragma unroll
                for (int a = 0; a < WG_A; ++a)
                    if (a0 + a < Ca) acc[a] = fmaf(__ldg(ab_ + (long long)(a0 + a) * hw + p), bv, acc[a]);
            }
        }
    }
    float* __restrict__ dst = part + (long long)blockIdx.x * Ca * Cb;
#pragma unroll
    for (int a = 0; a < WG_A; ++a) {
        const float s = warp_sum(acc[a]);
        if (lane == a && a0 + a < Ca && b < Cb) dst[(long long)(a0 + a) * Cb + b] = s;
    }
}

__global__ void reduce_splits_kernel(const float* __restrict__ part, float* __restrict__ g, const long long count, const int splits)
{
    const long long i = (long long)blockIdx.x * blockDim.x + threadIdx.x;
    if (i >= count) return;
    float s = 0.f;
    for (int k = 0; k < splits; ++k) s += part[(long long)k * count + i];
    g[i] = s;
}

__global__ void __launch_bounds__(256) channel_sum_kernel(const Slice A, const int N, const long long hw, float* __restrict__ part,
                                                          const int splits)
{
    __shared__ float red[8];
    const int c = blockIdx.y, split = blockIdx.x;
    const long long total = (long long)N * hw;
    float s = 0.f;
    for (long long i = (long long)split * 256 + threadIdx.x; i < total; i += (long long)splits * 256) {
        const int n = (int)(i / hw);
        const long long p = i % hw;
        s += __ldg(A.p + (long long)n * A.sstride + (long long)(A.c0 + c) * hw + p);
    }
    s = warp_sum(s);
    if ((threadIdx.x & 31) == 0) red[threadIdx.x >> 5] = s;
    __syncthreads();
    if (threadIdx.x == 0) {
        float t = 0.f;
        for (int w = 0; w < 8; ++w) t += red[w];
        part[(long long)split * A.C + c] = t;
    }
}

int pick_splits(long long work_items, int cta_per_split)
{
    long long s = (2 * num_sms() + cta_per_split - 1) / cta_per_split;
    if (s > work_items) s = work_items;
    if (s < 1) s = 1;
    return (int)s;
}

}  // namespace

size_t wgrad_scratch_floats(int, int) { return kScratchFloats; }

void wgrad3x3(const Slice& A, const Slice& B, int N, int Ha, int Wa, int Hb, int Wb, int S, int P, float* g, float* scratch,
              cudaStream_t st)
{
    ProfScope ps(F_WGRAD, 18.0 * N * Ha * Wa * A.C * B.C, 4.0 * N * ((double)Ha * Wa * A.C + (double)Hb * Wb * B.C), st);
    const int TH = S == 1 ? 8 : 4;
    const int tiles_x = (Wa + 31) / 32, tiles_y = (Ha + TH - 1) / TH;
    const int chunks = ((B.C + WG_B - 1) / WG_B) * ((A.C + WG_A - 1) / WG_A);
    const long long count = (long long)A.C * B.C * 9;
    int splits = pick_splits((long long)N * tiles_x * tiles_y, chunks);
    while (splits > 1 && (size_t)splits * count > kScratchFloats) --splits;
    dim3 grid(splits, (B.C + WG_B - 1) / WG_B, (A.C + WG_A - 1) / WG_A);
    if (S == 1) wgrad3x3_kernel<1, 8><<<grid, 256, 0, st>>>(A, B, N, Ha, Wa, Hb, Wb, P, scratch, tiles_x, tiles_y);
    else wgrad3x3_kernel<2, 4><<<grid, 256, 0, st>>>(A, B, N, Ha, Wa, Hb, Wb, P, scratch, tiles_x, tiles_y);
    reduce_splits_kernel<<<(unsigned)((count + 255) / 256), 256, 0, st>>>(scratch, g, count, splits);
    count_launch(2);
}

void wgrad1x1(const Slice& A, const Slice& B, int N, int H, int W, float* g, float* scratch, cudaStream_t st)
{
    ProfScope ps(F_WGRAD, 2.0 * N * H * W * A.C * B.C, 4.0 * N * H * W * (A.C + B.C), st);
    const long long hw = (long long)H * W;
    const int cps = (int)((hw + 255) / 256);
    const int chunks = ((B.C + WG_B - 1) / WG_B) * ((A.C + WG_A - 1) / WG_A);
    const long long count = (long long)A.C * B.C;
    const int splits = pick_splits((long long)N * cps, chunks);
    dim3 grid(splits, (B.C + WG_B - 1) / WG_B, (A.C + WG_A - 1) / WG_A);
    wgrad1x1_kernel<<<grid, 256, 0, st>>>(A, B, N, hw, scratch, cps);
    reduce_splits_kernel<<<(unsigned)((count + 255) / 256), 256, 0, st>>>(scratch, g, count, splits);
    count_launch(2);
}

void channel_sum(const Slice& A, int N, int H, int W, float* g, float* scratch, cudaStream_t st)
{
    ProfScope ps(F_BIAS_GRAD, 1.0 * N * H * W * A.C, 4.0 * N * H * W * A.C, st);
    const long long hw = (long long)H * W;
    long long s = ((long long)N * hw + 256 * 16 - 1) / (256 * 16);
    const int splits = (int)(s < 1 ? 1 : (s > 64 ? 64 : s));
    channel_sum_kernel<<<dim3(splits, A.C), 256, 0, st>>>(A, N, hw, scratch, splits);
    reduce_splits_kernel<<<(A.C + 255) / 256, 256, 0, st>>>(scratch, g, A.C, splits);
    count_launch(2);
}

}  // namespace k
}  // namespace ab
