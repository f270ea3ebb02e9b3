// cont_<N,3,3,2,2,0,0> ("up", NetStructure.h:83): transposed convolution with stride 2, no padding.
// Forward = tensor_conv::get_gradient_for_data of the equivalent con_ (layers.h:531, cpu_dlib.cpp:2298-2325):
//     out[n,co,2y+ky,2x+kx] += in[n,ci,y,x] * w[ci][co][ky][kx]        (out is (2h+1) x (2w+1))
// Written in gather form: one thread owns a 2x2 output quad, which depends on a 2x2 input
// neighbourhood with a fixed tap pattern, so no multiply is wasted on parity tests.
#include "kcommon.cuh"

namespace ab {
namespace k {
namespace {

constexpr int UP_CI_T = 8;
constexpr int UP_CO_T = 16;

__global__ void __launch_bounds__(128) up_forward_kernel(const Slice in, const Slice out, const int h, const int w,
                                                         const float* __restrict__ wt, const float* __restrict__ bias)
{
    __shared__ __align__(16) float ws[UP_CI_T][9][UP_CO_T];
    const int co0 = blockIdx.y * UP_CO_T, n = blockIdx.z;
    const int Cin = in.C, Cout = out.C;
    const int qw = w + 1, qh = h + 1;
    const int q = blockIdx.x * blockDim.x + threadIdx.x;
    const bool active = q < qw * qh;
    const int i = active ? q / qw : 0, j = active ? q % qw : 0;
    const long long hw = (long long)h * w;
    const int H2 = 2 * h + 1, W2 = 2 * w + 1;
    const float* __restrict__ inb = in.p + (long long)n * in.sstride + (long long)in.c0 * hw;
    const bool va = active && i < h && j < w, vb = active && i >= 1 && j < w, vc = active && i < h && j >= 1,
               vd = active && i >= 1 && j >= 1;
    const long long oa = (long long)i * w + j, ob = (long long)(i - 1) * w + j, oc = (long long)i * w + j - 1,
                    od = (long long)(i - 1) * w + j - 1;

    float acc[4][UP_CO_T];
#pragma unroll
    for (int p = 0; p < 4; ++p)
#pragma unroll
        for (int c = 0; c < UP_CO_T; ++c) acc[p][c] = 0.f;

    for (int ci0 = 0; ci0 < Cin; ci0 += UP_CI_T) {
        float* wflat = &ws[0][0][0];
        for (int idx = threadIdx.x; idx < UP_CI_T * 9 * UP_CO_T; idx += blockDim.x) {
            const int c = idx % UP_CO_T, t = (idx / UP_CO_T) % 9, ci = idx / (UP_CO_T * 9);
            wflat[idx] = (ci0 + ci < Cin && co0 + c < Cout) ? __ldg(wt + ((long long)(ci0 + ci) * Cout + co0 + c) * 9 + t) : 0.f;
        }
        __syncthreads();
        const int cmax = min(UP_CI_T, Cin - ci0);
        for (int ci = 0; ci < cmax; ++ci) {
            const float* s = inb + (long long)(ci0 + ci) * hw;
            const float a = va ? __ldg(s + oa) : 0.f, b = vb ? __ldg(s + ob) : 0.f, c_ = vc ? __ldg(s + oc) : 0.f,
                        d = vd ? __ldg(s + od) : 0.f;
#pragma unroll
            for (int c = 0; c < UP_CO_T; ++c) {
                acc[0][c] += a * ws[ci][0][c] + b * ws[ci][6][c] + c_ * ws[ci][2][c] + d * ws[ci][8][c];
                acc[1][c] += a * ws[ci][1][c] + b * ws[ci][7][c];
                acc[2][c] += a * ws[ci][3][c] + c_ * ws[ci][5][c];
                acc[3][c] += a * ws[ci][4][c];
            }
        }
        __syncthreads();
    }
    if (!active) return;
    const long long hw2 = (long long)H2 * W2;
    float* __restrict__ outb = out.p + (long long)n * out.sstride + (long long)out.c0 * hw2;
#pragma unroll
    for (int c = 0; c < UP_CO_T; ++c) {
        const int co = co0 + c;
        if (co >= Cout) break;
        const float b = bias ? __ldg(bias + co) : 0.f;
        float* o = outb + (long long)co * hw2 + (long long)(2 * i) * W2 + 2 * j;
        o[0] = acc[0][c] + b;
        if (j < w) o[1] = acc[1][c] + b;
        if (i < h) {
            o[W2] = acc[2][c] + b;
            if (j < w) o[W2 + 1] = acc[3][c] + b;
        }
    }
}

// backward-data of `up` (layers.h:542: conv of the output gradient with the same filters, stride 2):
//     gin[n,ci,i,j] += sum_{co,ky,kx} gout[n,co,2i+ky,2j+kx] * w[ci][co][ky][kx]
__global__ void __launch_bounds__(128) up_backward_data_kernel(const Slice gout, const Slice gin, const int h, const int w,
                                                               const float* __restrict__ wt)
{
    constexpr int CI_R = 16, CO_S = 8;
    __shared__ __align__(16) float ws[CO_S][9][CI_R];
    const int ci0 = blockIdx.y * CI_R, n = blockIdx.z;
    const int Cin = gin.C, Cout = gout.C;
    const long long hw = (long long)h * w;
    const int W2 = 2 * w + 1;
    const long long hw2 = (long long)(2 * h + 1) * W2;
    const long long pix = (long long)blockIdx.x * blockDim.x + threadIdx.x;
    const bool active = pix < hw;
    const int i = active ? (int)(pix / w) : 0, j = active ? (int)(pix % w) : 0;
    const float* __restrict__ gb = gout.p + (long long)n * gout.sstride + (long long)gout.c0 * hw2 + (long long)(2 * i) * W2 + 2 * j;
    float acc[CI_R];
#pragma unroll
    for (int c = 0; c < CI_R; ++c) acc[c] = 0.f;
    for (int co0 = 0; co0 < Cout; co0 += CO_S) {
        float* wflat = &ws[0][0][0];
        for (int idx = threadIdx.x; idx < CO_S * 9 * CI_R; idx += blockDim.x) {
            const int c = idx % CI_R, t = (idx / CI_R) % 9, co = idx / (CI_R * 9);
            wflat[idx] = (co0 + co < Cout && ci0 + c < Cin) ? __ldg(wt + ((long long)(ci0 + c) * Cout + co0 + co) * 9 + t) : 0.f;
        }
        __syncthreads();
        if (active) {
            const int cmax = min(CO_S, Cout - co0);
            for (int co = 0; co < cmax; ++co) {
                const float* s = gb + (long long)(co0 + co) * hw2;
                float g[9];
#pragma unroll
                for (int ky = 0; ky < 3; ++ky)
#pragma unroll
                    for (int kx = 0; kx < 3; ++kx) g[ky * 3 + kx] = __ldg(s + ky * W2 + kx);
#pragma unroll
                for (int t = 0; t < 9; ++t)
#pragma unroll
                    for (int c = 0; c < CI_R; ++c) acc[c] = fmaf(g[t], ws[co][t][c], acc[c]);
            }
        }
        __syncthreads();
    }
    if (!active) return;
    float* __restrict__ o = gin.p + (long long)n * gin.sstride + (long long)gin.c0 * hw + pix;
#pragma unroll
    for (int c = 0; c < CI_R; ++c)
        if (ci0 + c < Cin) o[(long long)(ci0 + c) * hw] += acc[c];
}

}  // namespace

void up_forward(const Slice& in, const Slice& out, int N, int h, int w, const float* wt, const float* bias, cudaStream_t st)
{
    ProfScope ps(F_UP, 18.0 * N * h * w * in.C * out.C, 4.0 * N * ((double)h * w * in.C + (2.0 * h + 1) * (2.0 * w + 1) * out.C), st);
    const int quads = (h + 1) * (w + 1);
    dim3 grid((quads + 127) / 128, (out.C + UP_CO_T - 1) / UP_CO_T, N);
    up_forward_kernel<<<grid, 128, 0, st>>>(in, out, h, w, wt, bias);
    count_launch();
}

void up_backward_data(const Slice& gout, const Slice& gin, int N, int h, int w, const float* wt, cudaStream_t st)
{
    ProfScope ps(F_UP, 18.0 * N * h * w * gin.C * gout.C, 4.0 * N * (2.0 * h * w * gin.C + (2.0 * h + 1) * (2.0 * w + 1) * gout.C), st);
    dim3 grid((unsigned)(((long long)h * w + 127) / 128), (gin.C + 15) / 16, N);
    up_backward_data_kernel<<<grid, 128, 0, st>>>(gout, gin, h, w, wt);
    count_launch();
}

namespace {
__global__ void __launch_bounds__(256) resize_nchw_kernel(const Slice in, int ih, int iw, const Slice out, int oh, int ow, const ResizeTab t)
{
    const int n = blockIdx.z, c = blockIdx.y;
    const float* s = in.p + (long long)n * in.sstride + (long long)(in.c0 + c) * ih * iw;
    float* d = out.p + (long long)n * out.sstride + (long long)(out.c0 + c) * oh * ow;
    for (int i = blockIdx.x * blockDim.x + threadIdx.x; i < oh * ow; i += gridDim.x * blockDim.x) {
        const int y = i / ow, x = i - y * ow;
        const int top = t.top[y], bottom = t.bottom[y], left = t.left[x], right = t.right[x];
        const float tb = t.tb[y], itb = t.itb[y], lr = t.lr[x], ilr = t.ilr[x];
        const float tl = s[top * iw + left], tr = s[top * iw + right], bl = s[bottom * iw + left], br = s[bottom * iw + right];
        d[i] = __fadd_rn(__fadd_rn(__fadd_rn(__fmul_rn(__fmul_rn(itb, ilr), tl), __fmul_rn(__fmul_rn(itb, lr), tr)), __fmul_rn(__fmul_rn(tb, ilr), bl)),
                         __fmul_rn(__fmul_rn(tb, lr), br));
    }
}
}  // namespace

void resize_bilinear_nchw(const Slice& in, int ih, int iw, const Slice& out, int N, int oh, int ow, const ResizeTab& t, cudaStream_t st)
{
    ProfScope ps(F_UP, 0, 4.0 * N * in.C * ((double)ih * iw + (double)oh * ow), st);
    const unsigned gx = (unsigned)std::min((oh * ow + 255) / 256, 64);
    resize_nchw_kernel<<<dim3(gx, in.C, N), 256, 0, st>>>(in, ih, iw, out, oh, ow, t);
    count_launch();
}

}  // namespace k
}  // namespace ab
