// Training-side bf16 NHWC ("padded flat", tc.h) elementwise / reduction kernels of the tensor-core path:
// batch normalisation in CONV_MODE fused with mish (cpu_dlib.cpp:1073-1289,1510-1558), the max-pool gradient
// (cpu_dlib.cpp:1971-2017), bias gradients (cpu_dlib.cpp:470-504) and the small layout conversions around them.
// All HBM-bound: 16-byte accesses of 8 channels, fp32 arithmetic, per-channel reductions in registers ->
// shared memory -> one double (or float) atomic per CTA and channel.
//
// Thread mapping of the per-channel kernels: blockDim = (C/8, TY); threadIdx.x owns 8 consecutive channels for the
// whole kernel, threadIdx.y strides over positions, so a warp reads consecutive 16-byte vectors of [pos][C].
#include "kcommon.cuh"
#include "tc.h"

#include <algorithm>
#include <stdexcept>

namespace ab {
namespace tc {
namespace {

struct alignas(16) BF8 { __nv_bfloat162 v[4]; };

__device__ __forceinline__ void unpack8(const BF8& b, float f[8])
{
#pragma unroll
    for (int i = 0; i < 4; ++i) { const float2 t = __bfloat1622float2(b.v[i]); f[2 * i] = t.x; f[2 * i + 1] = t.y; }
}
__device__ __forceinline__ BF8 pack8(const float f[8])
{
    BF8 b;
#pragma unroll
    for (int i = 0; i < 4; ++i) b.v[i] = __floats2bfloat162_rn(f[2 * i], f[2 * i + 1]);
    return b;
}
__device__ __forceinline__ BF8 ld8(const bf16* p) { return *reinterpret_cast<const BF8*>(p); }
__device__ __forceinline__ void ldf8(const float* p, float f[8])
{
    const float4 a = *reinterpret_cast<const float4*>(p), b = *reinterpret_cast<const float4*>(p + 4);
    f[0] = a.x; f[1] = a.y; f[2] = a.z; f[3] = a.w; f[4] = b.x; f[5] = b.y; f[6] = b.z; f[7] = b.w;
}
__device__ __forceinline__ void st8(bf16* p, const BF8& v) { *reinterpret_cast<BF8*>(p) = v; }

struct GeoD { int N, H, W, P; unsigned S; long long total; };  // total = N * S  (< 2^31: checked by geod)
GeoD geod(const Geo& g)
{
    if (g.positions() >= (1ll << 31)) throw std::invalid_argument("tensor too large for the bf16 NHWC kernels (positions must fit 31 bits)");
    return GeoD{g.N, g.H, g.W, g.P(), (unsigned)g.S(), (long long)g.N * g.S()};
}

// is `pos` a real pixel (not a pad position)?  32-bit arithmetic: positions fit 31 bits
__device__ __forceinline__ bool is_pixel(const GeoD& g, long long pos)
{
    if (pos >= g.total) return false;
    const unsigned rem = (unsigned)pos % g.S;
    return rem >= (unsigned)g.P && (rem % (unsigned)g.P) != 0;
}

// mish_ (cpu_dlib.cpp:1510-1558) on the fast-math intrinsics: the results feed bf16 tensors or fp32 sums of bf16-rounded
// data, so 2-ulp exp / division are far below the path's rounding noise.  e = inf -> r = 0 -> y = x.
__device__ __forceinline__ float mish_fast(float x)
{
    const float e = __expf(x);
    const float r = __fdividef(1.f, fmaf(e, e + 2.f, 2.f));
    return fmaf(-2.f * x, r, x);
}
// y = mish(x), dy = mish'(x) from ONE exponential; dlib's clamps at |x| >= 8 (cpu_dlib.cpp:1531-1543)
__device__ __forceinline__ void mish_both(float x, float& y, float& dy)
{
    const float e = __expf(x);
    const float r = __fdividef(1.f, fmaf(e, e + 2.f, 2.f));
    y = fmaf(-2.f * x, r, x);
    const float omega = fmaf(4.f, x + 1.f, fmaf(e, fmaf(e, e + 4.f, fmaf(4.f, x, 6.f)), 0.f));
    const float d = e * omega * r * r;
    dy = x >= 8.f ? 1.f : (x <= -8.f ? 0.f : d);
}
__device__ __forceinline__ float mish_grad_fast(float x)
{
    float y, dy;
    mish_both(x, y, dy);
    return dy;
}

constexpr int kMaxT = 256;

// sum the NV per-thread values over threadIdx.y and add them to out[v_base(k) + channel] with one atomic per value
template <int NV, typename OutT>
__device__ __forceinline__ void block_reduce_channels(const float (&v)[NV][8], OutT* const (&dst)[NV])
{
    __shared__ float sh[NV * 8][kMaxT + 1];
    const int C8 = blockDim.x, TY = blockDim.y, tid = threadIdx.y * C8 + threadIdx.x;
#pragma unroll
    for (int k = 0; k < NV; ++k)
#pragma unroll
        for (int i = 0; i < 8; ++i) sh[k * 8 + i][tid] = v[k][i];
    __syncthreads();
    for (int j = tid; j < NV * 8 * C8; j += C8 * TY) {
        const int cg = j / (NV * 8), ki = j - cg * (NV * 8);
        float s = 0.f;
        for (int y = 0; y < TY; ++y) s += sh[ki][y * C8 + cg];
        const int k = ki >> 3, i = ki & 7;
        atomicAdd(dst[k] + cg * 8 + i, (OutT)s);
    }
}

__global__ void __launch_bounds__(kMaxT) bn_stats_kernel(const float* __restrict__ z, int ctot, int c0, long long npos, int mish_first,
                                                         double* __restrict__ sums)
{
    const int C8 = blockDim.x;
    float acc[2][8] = {};
    for (long long pos = (long long)blockIdx.x * blockDim.y + threadIdx.y; pos < npos; pos += (long long)gridDim.x * blockDim.y) {
        float f[8];
        ldf8(z + pos * ctot + c0 + threadIdx.x * 8, f);
#pragma unroll
        for (int i = 0; i < 8; ++i) {
            const float v = mish_first ? mish_fast(f[i]) : f[i];
            acc[0][i] += v;
            acc[1][i] = fmaf(v, v, acc[1][i]);
        }
    }
    double* const dst[2] = {sums, sums + C8 * 8};
    block_reduce_channels<2, double>(acc, dst);
}

__global__ void __launch_bounds__(kMaxT) bn_apply_kernel(const float* __restrict__ z, int z_ctot, int z_c0, bf16* __restrict__ y, int y_ctot,
                                                         int y_c0, GeoD g, int mish_first, int has_mish, const float* __restrict__ gamma,
                                                         const float* __restrict__ beta, const float* __restrict__ stats)
{
    const int C = blockDim.x * 8, c = threadIdx.x * 8;
    float mean[8], sc[8], sh[8];
#pragma unroll
    for (int i = 0; i < 8; ++i) {
        mean[i] = stats[c + i];
        sc[i] = stats[C + c + i] * gamma[c + i];
        sh[i] = beta[c + i];
    }
    for (long long pos = (long long)blockIdx.x * blockDim.y + threadIdx.y; pos < g.total; pos += (long long)gridDim.x * blockDim.y) {
        if (!is_pixel(g, pos)) continue;
        float f[8];
        ldf8(z + pos * z_ctot + z_c0 + c, f);
#pragma unroll
        for (int i = 0; i < 8; ++i) {
            float v = f[i];
            if (has_mish && mish_first) v = mish_fast(v);
            v = fmaf(v - mean[i], sc[i], sh[i]);
            if (has_mish && !mish_first) v = mish_fast(v);
            f[i] = v;
        }
        st8(y + pos * y_ctot + y_c0 + c, pack8(f));
    }
}

__global__ void __launch_bounds__(kMaxT) bn_bwd_reduce_kernel(const float* __restrict__ z, int z_ctot, int z_c0, const bf16* __restrict__ gy,
                                                              int gy_ctot, int gy_c0, long long npos, int mish_first, int has_mish,
                                                              const float* __restrict__ gamma, const float* __restrict__ beta,
                                                              const float* __restrict__ stats, double* __restrict__ red)
{
    const int C = blockDim.x * 8, c = threadIdx.x * 8;
    float mean[8], invstd[8], ga[8], be[8];
#pragma unroll
    for (int i = 0; i < 8; ++i) { mean[i] = stats[c + i]; invstd[i] = stats[C + c + i]; ga[i] = gamma[c + i]; be[i] = beta[c + i]; }
    float acc[2][8] = {};
    for (long long pos = (long long)blockIdx.x * blockDim.y + threadIdx.y; pos < npos; pos += (long long)gridDim.x * blockDim.y) {
        float f[8], d[8];
        ldf8(z + pos * z_ctot + z_c0 + c, f);
        unpack8(ld8(gy + pos * gy_ctot + gy_c0 + c), d);
#pragma unroll
        for (int i = 0; i < 8; ++i) {
            float x = f[i], dv = d[i];
            if (has_mish && mish_first) x = mish_fast(x);
            const float xhat = (x - mean[i]) * invstd[i];
            if (has_mish && !mish_first) dv *= mish_grad_fast(fmaf(xhat, ga[i], be[i]));
            acc[0][i] += dv;
            acc[1][i] = fmaf(dv, xhat, acc[1][i]);
        }
    }
    double* const dst[2] = {red, red + C};
    block_reduce_channels<2, double>(acc, dst);
}

__global__ void __launch_bounds__(kMaxT) bn_bwd_apply_kernel(const float* __restrict__ z, int z_ctot, int z_c0, const bf16* __restrict__ gy, int gy_ctot,
                                                             int gy_c0, bf16* __restrict__ dz, int dz_ctot, int dz_c0, GeoD g, int mish_first, int has_mish,
                                                             const float* __restrict__ gamma, const float* __restrict__ beta,
                                                             const float* __restrict__ stats, const double* __restrict__ red, float invnum,
                                                             float* __restrict__ dgamma, float* __restrict__ dbeta, float* __restrict__ dbias)
{
    const int C = blockDim.x * 8, c = threadIdx.x * 8;
    float mean[8], invstd[8], ga[8], be[8], k1[8], k2[8], k3[8];
#pragma unroll
    for (int i = 0; i < 8; ++i) {
        mean[i] = stats[c + i]; invstd[i] = stats[C + c + i]; ga[i] = gamma[c + i]; be[i] = beta[c + i];
        const float s_dv = (float)red[c + i], s_dvx = (float)red[C + c + i];
        // cpu_dlib.cpp:1228-1288 with the per-channel sums factored out (see kernels_bn.cu)
        const float dvars = -0.5f * ga[i] * invstd[i] * invstd[i] * s_dvx;
        const float dmeans = -ga[i] * invstd[i] * s_dv;
        k1[i] = ga[i] * invstd[i];
        k2[i] = dvars * 2.f * invnum;
        k3[i] = dmeans * invnum;
        if (blockIdx.x == 0 && threadIdx.y == 0) { dgamma[c + i] = s_dvx; dbeta[c + i] = s_dv; }
    }
    float acc[1][8] = {};
    for (long long pos = (long long)blockIdx.x * blockDim.y + threadIdx.y; pos < g.total; pos += (long long)gridDim.x * blockDim.y) {
        if (!is_pixel(g, pos)) continue;
        float f[8], d[8];
        ldf8(z + pos * z_ctot + z_c0 + c, f);
        unpack8(ld8(gy + pos * gy_ctot + gy_c0 + c), d);
#pragma unroll
        for (int i = 0; i < 8; ++i) {
            const float zv = f[i];
            float dv = d[i], x = zv, dm = 1.f;
            if (has_mish && mish_first) mish_both(zv, x, dm);
            const float xc = x - mean[i];
            if (has_mish && !mish_first) dv *= mish_grad_fast(fmaf(xc * invstd[i], ga[i], be[i]));
            float r = fmaf(dv, k1[i], fmaf(k2[i], xc, k3[i]));
            if (has_mish && mish_first) r *= dm;
            f[i] = r;
            acc[0][i] += r;
        }
        st8(dz + pos * dz_ctot + dz_c0 + c, pack8(f));
    }
    float* const dst[1] = {dbias};
    block_reduce_channels<1, float>(acc, dst);
}

__global__ void __launch_bounds__(kMaxT) channel_sum_kernel(const bf16* __restrict__ v, int ctot, int c0, long long npos, float* __restrict__ out)
{
    float acc[1][8] = {};
    for (long long pos = (long long)blockIdx.x * blockDim.y + threadIdx.y; pos < npos; pos += (long long)gridDim.x * blockDim.y) {
        float f[8];
        unpack8(ld8(v + pos * ctot + c0 + threadIdx.x * 8), f);
#pragma unroll
        for (int i = 0; i < 8; ++i) acc[0][i] += f[i];
    }
    float* const dst[1] = {out};
    block_reduce_channels<1, float>(acc, dst);
}

// gather form: an input pixel belongs to at most 2 x 2 windows and receives a window's gradient iff it is that window's
// first maximum in row-major order -- which element won is read from the byte map the forward pass wrote
__global__ void __launch_bounds__(256) maxpool_bwd_kernel(const uint8_t* __restrict__ arg, int C8, int N, int Hi, int Wi,
                                                          const bf16* __restrict__ gout, int Ho, int Wo, bf16* __restrict__ gin, int gin_ctot,
                                                          int gin_c0)
{
    const long long total = (long long)N * Hi * Wi * C8;
    const long long i = (long long)blockIdx.x * blockDim.x + threadIdx.x;
    if (i >= total) return;
    const int cg = (int)((unsigned long long)i % (unsigned)C8);
    unsigned p = (unsigned)((unsigned long long)i / (unsigned)C8);
    const int x = (int)(p % (unsigned)Wi); p /= (unsigned)Wi;
    const int y = (int)(p % (unsigned)Hi);
    const int n = (int)(p / (unsigned)Hi);
    const long long Si = (long long)(Hi + 1) * (Wi + 1), So = (long long)(Ho + 1) * (Wo + 1);
    float acc[8] = {};
    bool any = false;
    const int Y0 = max((y - 1) >> 1, 0), Y1 = min(y >> 1, Ho - 1), X0 = max((x - 1) >> 1, 0), X1 = min(x >> 1, Wo - 1);
    for (int Y = Y0; Y <= Y1; ++Y)
        for (int X = X0; X <= X1; ++X) {
            const long long po = (n * So + (long long)(Y + 1) * (Wo + 1) + (X + 1)) * (C8 * 8) + cg * 8;
            const uint2 w = *reinterpret_cast<const uint2*>(arg + po);
            const uint32_t mine = (uint32_t)((y - 2 * Y) * 3 + (x - 2 * X)) * 0x01010101u;
            // bytes of w equal to `mine`: zero bytes of the xor
            const uint32_t dx = w.x ^ mine, dy = w.y ^ mine;
            if ((((dx - 0x01010101u) & ~dx) | ((dy - 0x01010101u) & ~dy)) & 0x80808080u) {
                float go[8];
                unpack8(ld8(gout + po), go);
#pragma unroll
                for (int k = 0; k < 4; ++k) {
                    if (((dx >> (8 * k)) & 0xffu) == 0) { acc[k] += go[k]; any = true; }
                    if (((dy >> (8 * k)) & 0xffu) == 0) { acc[4 + k] += go[4 + k]; any = true; }
                }
            }
        }
    if (!any) return;
    bf16* gp = gin + (n * Si + (long long)(y + 1) * (Wi + 1) + (x + 1)) * gin_ctot + gin_c0 + cg * 8;
    float old[8];
    unpack8(ld8(gp), old);
#pragma unroll
    for (int k = 0; k < 8; ++k) old[k] += acc[k];
    st8(gp, pack8(old));
}

__global__ void __launch_bounds__(256) planes_to_nhwc8_kernel(const float* __restrict__ planes, int K, bf16* __restrict__ out, int N, int H, int W)
{
    const long long hw = (long long)H * W;
    const long long i = (long long)blockIdx.x * blockDim.x + threadIdx.x;
    if (i >= (long long)N * hw) return;
    const int n = (int)(i / hw);
    const long long rem = i - n * hw;
    const int y = (int)(rem / W), x = (int)(rem - (long long)y * W);
    float f[8] = {0, 0, 0, 0, 0, 0, 0, 0};
    for (int k = 0; k < K && k < 8; ++k) f[k] = planes[((long long)n * K + k) * hw + rem];
    const long long pos = n * (long long)(H + 1) * (W + 1) + (long long)(y + 1) * (W + 1) + (x + 1);
    st8(out + pos * 8, pack8(f));
}

__global__ void __launch_bounds__(256) upsample_zero_kernel(const bf16* __restrict__ in, int in_ctot, int in_c0, int C8, int N, int h, int w,
                                                            bf16* __restrict__ out)
{
    const int Hf = 2 * h + 1, Wf = 2 * w + 1;
    const long long total = (long long)N * Hf * Wf * C8;
    const long long i = (long long)blockIdx.x * blockDim.x + threadIdx.x;
    if (i >= total) return;
    const int cg = (int)(i % C8);
    long long p = i / C8;
    const int x = (int)(p % Wf); p /= Wf;
    const int y = (int)(p % Hf);
    const int n = (int)(p / Hf);
    BF8 v;
#pragma unroll
    for (int k = 0; k < 4; ++k) v.v[k] = __floats2bfloat162_rn(0.f, 0.f);
    if (!(y & 1) && !(x & 1) && (y >> 1) < h && (x >> 1) < w)
        v = ld8(in + (n * (long long)(h + 1) * (w + 1) + (long long)((y >> 1) + 1) * (w + 1) + ((x >> 1) + 1)) * in_ctot + in_c0 + cg * 8);
    st8(out + (n * (long long)(Hf + 1) * (Wf + 1) + (long long)(y + 1) * (Wf + 1) + (x + 1)) * (C8 * 8) + cg * 8, v);
}

__global__ void __launch_bounds__(256) pack_batched_kernel(const float* __restrict__ arena, bf16* __restrict__ packs, const PackJob* __restrict__ jobs)
{
    const PackJob& j = jobs[blockIdx.y];
    const long long total = (long long)j.ntaps * j.npad * j.kp;
    const float* w = arena + j.w_off;
    bf16* packed = packs + j.packed_off;
    for (long long i = (long long)blockIdx.x * blockDim.x + threadIdx.x; i < total; i += (long long)gridDim.x * blockDim.x) {
        const int kidx = (int)(i % j.kp);
        const int n = (int)((i / j.kp) % j.npad);
        const int t = (int)(i / ((long long)j.kp * j.npad));
        float v = 0.f;
        if (n < j.nreal && kidx < j.kreal) v = w[n * j.sn + kidx * j.sk + j.tapmap[t]];
        packed[i] = __float2bfloat16(v);
    }
}

struct Shape { dim3 block; unsigned grid; };
Shape channel_shape(int C, long long npos)
{
    const int C8 = C / 8;
    if (C % 8 || C8 < 1 || C8 > 32) throw std::invalid_argument("bf16 NHWC kernels need 8 <= C <= 256, C % 8 == 0");
    const int TY = kMaxT / C8;
    const long long want = (npos + TY - 1) / TY;
    return Shape{dim3(C8, TY), (unsigned)std::max<long long>(1, std::min<long long>(want, k::kNumSMs * 8))};
}

}  // namespace

void bn_stats_nhwc(const ViewF& z, const Geo& g, int mish_first, double* sums, cudaStream_t st)
{
    const GeoD d = geod(g);
    k::ProfScope ps(k::F_EW_BF16, 3.0 * g.N * g.H * g.W * z.C, 2.0 * g.N * g.H * g.W * z.C, st);
    cudaMemsetAsync(sums, 0, sizeof(double) * 2 * z.C, st);
    const Shape s = channel_shape(z.C, d.total);
    bn_stats_kernel<<<s.grid, s.block, 0, st>>>(z.p, z.ctot, z.c0, d.total, mish_first, sums);
    k::count_launch();
}

void bn_apply_nhwc(const ViewF& z, const View& y, const Geo& g, int mish_first, int has_mish, const float* gamma, const float* beta,
                   const float* stats, cudaStream_t st)
{
    const GeoD d = geod(g);
    k::ProfScope ps(k::F_EW_BF16, 4.0 * g.N * g.H * g.W * z.C, 4.0 * g.N * g.H * g.W * z.C, st);
    const Shape s = channel_shape(z.C, d.total);
    bn_apply_kernel<<<s.grid, s.block, 0, st>>>(z.p, z.ctot, z.c0, y.p, y.ctot, y.c0, d, mish_first, has_mish, gamma, beta, stats);
    k::count_launch();
}

void bn_backward_reduce_nhwc(const ViewF& z, const View& gy, const Geo& g, int mish_first, int has_mish, const float* gamma,
                             const float* beta, const float* stats, double* red, cudaStream_t st)
{
    const GeoD d = geod(g);
    k::ProfScope ps(k::F_EW_BF16, 6.0 * g.N * g.H * g.W * z.C, 4.0 * g.N * g.H * g.W * z.C, st);
    cudaMemsetAsync(red, 0, sizeof(double) * 2 * z.C, st);
    const Shape s = channel_shape(z.C, d.total);
    bn_bwd_reduce_kernel<<<s.grid, s.block, 0, st>>>(z.p, z.ctot, z.c0, gy.p, gy.ctot, gy.c0, d.total, mish_first, has_mish, gamma, beta, stats,
                                                     red);
    k::count_launch();
}

void bn_backward_apply_nhwc(const ViewF& z, const View& gy, const View& dz, const Geo& g, int mish_first, int has_mish, const float* gamma,
                            const float* beta, const float* stats, const double* red, float* dgamma, float* dbeta, float* dbias,
                            cudaStream_t st)
{
    const GeoD d = geod(g);
    k::ProfScope ps(k::F_EW_BF16, 8.0 * g.N * g.H * g.W * z.C, 6.0 * g.N * g.H * g.W * z.C, st);
    const Shape s = channel_shape(z.C, d.total);
    const float invnum = 1.0f / (float)((long long)g.N * g.H * g.W);
    bn_bwd_apply_kernel<<<s.grid, s.block, 0, st>>>(z.p, z.ctot, z.c0, gy.p, gy.ctot, gy.c0, dz.p, dz.ctot, dz.c0, d, mish_first, has_mish, gamma, beta, stats, red,
                                                    invnum, dgamma, dbeta, dbias);
    k::count_launch();
}

void channel_sum_nhwc(const View& v, const Geo& g, float* out, cudaStream_t st)
{
    const GeoD d = geod(g);
    k::ProfScope ps(k::F_EW_BF16, 1.0 * g.N * g.H * g.W * v.C, 2.0 * g.N * g.H * g.W * v.C, st);
    const Shape s = channel_shape(v.C, d.total);
    channel_sum_kernel<<<s.grid, s.block, 0, st>>>(v.p, v.ctot, v.c0, d.total, out);
    k::count_launch();
}

void maxpool_backward_nhwc(const uint8_t* arg, int C, const Geo& gi, const bf16* gout_dense, const Geo& go, const View& gin, cudaStream_t st)
{
    const long long n = (long long)gi.N * gi.H * gi.W * (C / 8);
    if (n >= (1ll << 31) * 8) throw std::invalid_argument("maxpool_backward_nhwc: tensor too large");
    k::ProfScope ps(k::F_EW_BF16, 0, 2.0 * C * (2.0 * gi.N * gi.H * gi.W + 1.5 * (double)go.N * go.H * go.W), st);
    maxpool_bwd_kernel<<<(unsigned)((n + 255) / 256), 256, 0, st>>>(arg, C / 8, gi.N, gi.H, gi.W, gout_dense, go.H, go.W, gin.p, gin.ctot, gin.c0);
    k::count_launch();
}

void planes_to_nhwc8(const float* planes, int K, bf16* out, const Geo& g, cudaStream_t st)
{
    if (K > 8) throw std::invalid_argument("planes_to_nhwc8: at most 8 planes");
    const long long n = (long long)g.N * g.H * g.W;
    k::ProfScope ps(k::F_PACK, 0, (4.0 * K + 16.0) * n, st);
    planes_to_nhwc8_kernel<<<(unsigned)((n + 255) / 256), 256, 0, st>>>(planes, K, out, g.N, g.H, g.W);
    k::count_launch();
}

void upsample_zero_nhwc(const View& in, const Geo& gc, bf16* out, const Geo& gf, cudaStream_t st)
{
    if (gf.H != 2 * gc.H + 1 || gf.W != 2 * gc.W + 1) throw std::logic_error("upsample_zero_nhwc: geometry mismatch");
    const long long n = (long long)gf.N * gf.H * gf.W * (in.C / 8);
    k::ProfScope ps(k::F_PACK, 0, 2.0 * in.C * ((double)gc.N * gc.H * gc.W + (double)gf.N * gf.H * gf.W), st);
    upsample_zero_kernel<<<(unsigned)((n + 255) / 256), 256, 0, st>>>(in.p, in.ctot, in.c0, in.C / 8, gc.N, gc.H, gc.W, out);
    k::count_launch();
}

void pack_filters_batched(const float* arena, bf16* packs, const PackJob* jobs_dev, int njobs, int max_elems, cudaStream_t st)
{
    if (njobs <= 0) return;
    k::ProfScope ps(k::F_PACK, 0, 0, st);
    const unsigned gx = (unsigned)std::max(1, std::min((max_elems + 255) / 256, 64));
    pack_batched_kernel<<<dim3(gx, njobs), 256, 0, st>>>(arena, packs, jobs_dev);
    k::count_launch();
}

}  // namespace tc
}  // namespace ab
