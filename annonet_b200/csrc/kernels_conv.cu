// fp32 implicit-GEMM convolutions on the FMA pipe (the exact-parity path; the h16 tcgen05 path is
// kernels_tc.cu).  Direct form: a CTA owns a TW x TH pixel tile of one sample and CO_T output
// channels; input channels stream through shared memory 8 at a time (patch with a 1-pixel halo +
// the 8x9xCO_T filter slab); each thread keeps a 4-pixel x CO_T register tile.
#include "kcommon.cuh"

namespace ab {
namespace k {

std::atomic<unsigned long long> g_launches{0};

namespace {

constexpr int CI_T = 8;

template <int CO_T, int TQ, int TH>
__global__ void __launch_bounds__(TQ* TH) conv3x3_kernel(const ConvArgs a, const int tiles_x)
{
    constexpr int NT = TQ * TH, TW = TQ * 4, PW = TW + 4, PH = TH + 2;
    __shared__ __align__(16) float xs[CI_T][PH][PW];
    __shared__ __align__(16) float ws[CI_T][9][CO_T];

    const int tid = threadIdx.x, tx = tid % TQ, ty = tid / TQ;
    const int x0 = (blockIdx.x % tiles_x) * TW, y0 = (blockIdx.x / tiles_x) * TH;
    const int co0 = blockIdx.y * CO_T, n = blockIdx.z;
    const int H = a.H, W = a.W, Cin = a.in.C, Cout = a.out.C;
    const long long hw = (long long)H * W;
    const float* __restrict__ inb = a.in.p + (long long)n * a.in.sstride + (long long)a.in.c0 * hw;
    const float* __restrict__ wsrc = a.wv.w;

    float acc[4][CO_T];
#pragma unroll
    for (int p = 0; p < 4; ++p)
#pragma unroll
        for (int c = 0; c < CO_T; ++c) acc[p][c] = 0.f;

    for (int ci0 = 0; ci0 < Cin; ci0 += CI_T) {
        float* xflat = &xs[0][0][0];
        for (int idx = tid; idx < CI_T * PH * PW; idx += NT) {
            const int j = idx % PW, r = (idx / PW) % PH, ci = idx / (PW * PH);
            const int gy = y0 - 1 + r, gx = x0 - 1 + j;
            float v = 0.f;
            if (ci0 + ci < Cin && gy >= 0 && gy < H && gx >= 0 && gx < W)
                v = __ldg(inb + (long long)(ci0 + ci) * hw + (long long)gy * W + gx);
            xflat[idx] = v;
        }
        float* wflat = &ws[0][0][0];
        for (int idx = tid; idx < CI_T * 9 * CO_T; idx += NT) {
            const int c = idx % CO_T, t = (idx / CO_T) % 9, ci = idx / (CO_T * 9);
            float v = 0.f;
            if (ci0 + ci < Cin && co0 + c < Cout)
                v = __ldg(wsrc + (long long)(co0 + c) * a.wv.s_out + (long long)(ci0 + ci) * a.wv.s_in + (a.wv.flip ? 8 - t : t));
            wflat[idx] = v;
        }
        __syncthreads();
#pragma unroll 2
        for (int ci = 0; ci < CI_T; ++ci) {
#pragma unroll
            for (int ky = 0; ky < 3; ++ky) {
                const float* row = &xs[ci][ty + ky][4 * tx];
                const float4 xa = *reinterpret_cast<const float4*>(row);
                const float2 xb = *reinterpret_cast<const float2*>(row + 4);
                const float xv[6] = {xa.x, xa.y, xa.z, xa.w, xb.x, xb.y};
#pragma unroll
                for (int kx = 0; kx < 3; ++kx) {
                    const float* wr = &ws[ci][ky * 3 + kx][0];
#pragma unroll
                    for (int c = 0; c < CO_T; c += 4) {
                        const float4 wv = *reinterpret_cast<const float4*>(wr + c);
#pragma unroll
                        for (int p = 0; p < 4; ++p) {
                            acc[p][c + 0] = fmaf(xv[p + kx], wv.x, acc[p][c + 0]);
                            acc[p][c + 1] = fmaf(xv[p + kx], wv.y, acc[p][c + 1]);
                            acc[p][c + 2] = fmaf(xv[p + kx], wv.z, acc[p][c + 2]);
                            acc[p][c + 3] = fmaf(xv[p + kx], wv.w, acc[p][c + 3]);
                        }
                    }
                }
            }
        }
        __syncthreads();
    }

    const int y = y0 + ty;
    if (y >= H) return;
    float* __restrict__ outb = a.out.p + (long long)n * a.out.sstride + (long long)a.out.c0 * hw + (long long)y * W;
#pragma unroll
    for (int c = 0; c < CO_T; ++c) {
        const int co = co0 + c;
        if (co >= Cout) break;
        const float b = a.bias ? __ldg(a.bias + co) : 0.f;
        float sc = 1.f, sh = 0.f;
        if (a.epi != EPI_NONE) { sc = __ldg(a.scale + co); sh = __ldg(a.shift + co); }
#pragma unroll
        for (int p = 0; p < 4; ++p) {
            const int x = x0 + 4 * tx + p;
            if (x >= W) break;
            float v = acc[p][c] + b;
            if (a.epi == EPI_MISH_AFFINE) v = mish_f(v) * sc + sh;
            else if (a.epi == EPI_AFFINE_MISH) v = mish_f(v * sc + sh);
            float* o = outb + (long long)co * hw + x;
            if (a.accumulate) v += *o;
            *o = v;
        }
    }
}

// 1x1: one thread per output pixel, CO_T output channels in registers, filters broadcast from smem.
template <int CO_T>
__global__ void __launch_bounds__(128) conv1x1_kernel(const ConvArgs a)
{
    extern __shared__ float w1[];  // [Cin][CO_T]
    const int co0 = blockIdx.y * CO_T, n = blockIdx.z;
    const int Cin = a.in.C, Cout = a.out.C;
    for (int idx = threadIdx.x; idx < Cin * CO_T; idx += blockDim.x) {
        const int c = idx % CO_T, ci = idx / CO_T;
        w1[idx] = (co0 + c < Cout) ? __ldg(a.wv.w + (long long)(co0 + c) * a.wv.s_out + (long long)ci * a.wv.s_in) : 0.f;
    }
    __syncthreads();
    const long long hw = (long long)a.H * a.W;
    const long long pix = (long long)blockIdx.x * blockDim.x + threadIdx.x;
    if (pix >= hw) return;
    float acc[CO_T];
#pragma unroll
    for (int c = 0; c < CO_T; ++c) acc[c] = 0.f;
    if (!a.pool) {
        const float* __restrict__ inb = a.in.p + (long long)n * a.in.sstride + (long long)a.in.c0 * hw + pix;
        for (int ci = 0; ci < Cin; ++ci) {
            const float xv = __ldg(inb + (long long)ci * hw);
#pragma unroll
            for (int c = 0; c < CO_T; ++c) acc[c] = fmaf(xv, w1[ci * CO_T + c], acc[c]);
        }
    } else {
        const long long hwin = (long long)a.Hin * a.Win;
        const int y = (int)(pix / a.W), x = (int)(pix % a.W);
        const float* __restrict__ inb = a.in.p + (long long)n * a.in.sstride + (long long)a.in.c0 * hwin + (long long)(2 * y) * a.Win + 2 * x;
        for (int ci = 0; ci < Cin; ++ci) {
            const float* s = inb + (long long)ci * hwin;
            float m = -INFINITY;  // window always inside the image for out = 1 + (in-3)/2
#pragma unroll
            for (int r = 0; r < 3; ++r)
#pragma unroll
                for (int q = 0; q < 3; ++q) m = fmaxf(m, __ldg(s + r * a.Win + q));
#pragma unroll
            for (int c = 0; c < CO_T; ++c) acc[c] = fmaf(m, w1[ci * CO_T + c], acc[c]);
        }
    }
    float* __restrict__ outb = a.out.p + (long long)n * a.out.sstride + (long long)a.out.c0 * hw + pix;
#pragma unroll
    for (int c = 0; c < CO_T; ++c) {
        const int co = co0 + c;
        if (co >= Cout) break;
        float v = acc[c] + (a.bias ? __ldg(a.bias + co) : 0.f);
        if (a.epi == EPI_MISH_AFFINE) v = mish_f(v) * __ldg(a.scale + co) + __ldg(a.shift + co);
        else if (a.epi == EPI_AFFINE_MISH) v = mish_f(v * __ldg(a.scale + co) + __ldg(a.shift + co));
        float* o = outb + (long long)co * hw;
        if (a.accumulate) v += *o;
        *o = v;
    }
}

int pick_co_tile(int cout)
{
    const int pad16 = (cout + 15) / 16 * 16 - cout;
    if (pad16 * 10 <= cout) return 16;
    if (cout % 24 == 0) return 24;
    if (cout <= 8 || cout % 8 == 0) return 8;
    return 16;
}

template <int CO_T>
void launch3x3(const ConvArgs& a, cudaStream_t st)
{
    if (a.W >= 48) {
        constexpr int TQ = 16, TH = 16;
        const int tx = (a.W + TQ * 4 - 1) / (TQ * 4), ty = (a.H + TH - 1) / TH;
        dim3 grid(tx * ty, (a.out.C + CO_T - 1) / CO_T, a.N);
        conv3x3_kernel<CO_T, TQ, TH><<<grid, TQ * TH, 0, st>>>(a, tx);
    } else {
        constexpr int TQ = 4, TH = 16;
        const int tx = (a.W + TQ * 4 - 1) / (TQ * 4), ty = (a.H + TH - 1) / TH;
        dim3 grid(tx * ty, (a.out.C + CO_T - 1) / CO_T, a.N);
        conv3x3_kernel<CO_T, TQ, TH><<<grid, TQ * TH, 0, st>>>(a, tx);
    }
    count_launch();
}

template <int CO_T>
void launch1x1(const ConvArgs& a, cudaStream_t st)
{
    const long long hw = (long long)a.H * a.W;
    dim3 grid((unsigned)((hw + 127) / 128), (a.out.C + CO_T - 1) / CO_T, a.N);
    conv1x1_kernel<CO_T><<<grid, 128, (size_t)a.in.C * CO_T * sizeof(float), st>>>(a);
    count_launch();
}

}  // namespace

void conv3x3(const ConvArgs& a, cudaStream_t st)
{
    const double px = (double)a.N * a.H * a.W;
    ProfScope ps(F_CONV3X3, 18.0 * px * a.in.C * a.out.C, 4.0 * px * (a.in.C + a.out.C * (a.accumulate ? 2 : 1)), st);
    switch (pick_co_tile(a.out.C)) {
    case 8: launch3x3<8>(a, st); break;
    case 24: launch3x3<24>(a, st); break;
    default: launch3x3<16>(a, st); break;
    }
}

void conv1x1(const ConvArgs& a, cudaStream_t st)
{
    const double px = (double)a.N * a.H * a.W, pin = a.pool ? (double)a.N * a.Hin * a.Win : px;
    ProfScope ps(F_CONV1X1, 2.0 * px * a.in.C * a.out.C, 4.0 * (pin * a.in.C + px * a.out.C * (a.accumulate ? 2 : 1)), st);
    if (a.out.C <= 8 || (a.out.C % 16 != 0 && a.out.C % 8 == 0)) launch1x1<8>(a, st);
    else launch1x1<16>(a, st);
}

}  // namespace k
}  // namespace ab
