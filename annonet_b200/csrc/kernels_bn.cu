// Batch normalisation (CONV_MODE, cpu_dlib.cpp:1073-1289) fused with mish (cpu_dlib.cpp:1510-1558),
// and the 3x3/2 max-pool (cpu_dlib.cpp:1888-2017).  All HBM-bound: one read (+ one write) of the
// tensor per pass, per-channel statistics reduced with warp shuffles and one double atomic per CTA.
#include "kcommon.cuh"

namespace ab {
namespace k {
namespace {

constexpr int EW_T = 256;   // threads per CTA
constexpr int EW_PER = 8;   // elements per thread
constexpr int EW_CHUNK = EW_T * EW_PER;

__device__ __forceinline__ void block_sum2_atomic(double a, double b, double* da, double* db)
{
    __shared__ double sa[EW_T / 32], sb[EW_T / 32];
    a = warp_sum(a);
    b = warp_sum(b);
    const int lane = threadIdx.x & 31, w = threadIdx.x >> 5;
    if (lane == 0) { sa[w] = a; sb[w] = b; }
    __syncthreads();
    if (threadIdx.x == 0) {
        double ta = 0, tb = 0;
        for (int i = 0; i < EW_T / 32; ++i) { ta += sa[i]; tb += sb[i]; }
        atomicAdd(da, ta);
        atomicAdd(db, tb);
    }
}

// grid: (chunks, C, N)
__global__ void __launch_bounds__(EW_T) bn_stats_kernel(const Slice z, const long long hw, const int mish_first, double* __restrict__ sums)
{
    const int c = blockIdx.y, n = blockIdx.z;
    const float* __restrict__ s = plane(z, n, c, hw);
    const long long base = (long long)blockIdx.x * EW_CHUNK;
    float a = 0.f, b = 0.f;
#pragma unroll
    for (int k = 0; k < EW_PER; ++k) {
        const long long i = base + k * EW_T + threadIdx.x;
        if (i < hw) {
            float v = __ldg(s + i);
            if (mish_first) v = mish_f(v);
            a += v;
            b = fmaf(v, v, b);
        }
    }
    block_sum2_atomic((double)a, (double)b, sums + c, sums + z.C + c);
}

// one CTA; the counter is read by every thread before thread 0 advances it
__global__ void bn_finalize_kernel(const double* __restrict__ sums, const int C, const double m, const double eps,
                                   unsigned long long* __restrict__ num_updates, const unsigned long long window,
                                   float* __restrict__ rmean, float* __restrict__ rvar, float* __restrict__ stats)
{
    const int c = threadIdx.x;
    const unsigned long long nu = *num_updates;
    // layers.h:1392-1397: decay = 1 - num_updates/(num_updates+1); ++num_updates (capped by the window)
    const double decay = 1.0 - (double)nu / ((double)nu + 1.0);
    __syncthreads();
    if (c == 0) *num_updates = nu + 1 > window ? window : nu + 1;
    if (c >= C) return;
    const float mean = (float)(sums[c] / m);
    const float sq = (float)(sums[C + c] / m);
    const float var = sq - mean * mean;
    const double scale = m / (m - 1.0);  // unbiased running variance
    if (decay == 1.0) {
        rvar[c] = (float)(scale * var);
        rmean[c] = mean;
    } else {
        rvar[c] = (float)((1 - decay) * rvar[c] + scale * decay * var);
        rmean[c] = (float)((1 - decay) * rmean[c] + decay * mean);
    }
    stats[c] = mean;
    stats[C + c] = (float)(1.0 / sqrt((double)var + eps));
}

__global__ void __launch_bounds__(EW_T) bn_apply_kernel(const Slice z, const Slice y, const long long hw, const int mish_first,
                                                        const int has_mish, const float* __restrict__ gamma,
                                                        const float* __restrict__ beta, const float* __restrict__ stats)
{
    const int c = blockIdx.y, n = blockIdx.z;
    const float* __restrict__ s = plane(z, n, c, hw);
    float* __restrict__ d = plane_w(y, n, c, hw);
    const float mean = stats[c], invstd = stats[z.C + c], g = gamma[c], b = beta[c];
    const long long base = (long long)blockIdx.x * EW_CHUNK;
#pragma unroll
    for (int k = 0; k < EW_PER; ++k) {
        const long long i = base + k * EW_T + threadIdx.x;
        if (i < hw) {
            float v = __ldg(s + i);
            if (has_mish && mish_first) v = mish_f(v);
            v = (v - mean) * invstd;
            v = v * g + b;
            if (has_mish && !mish_first) v = mish_f(v);
            d[i] = v;
        }
    }
}

__global__ void __launch_bounds__(EW_T) bn_bwd_reduce_kernel(const Slice z, const Slice gy, const long long hw, const int mish_first,
                                                             const int has_mish, const float* __restrict__ gamma,
                                                             const float* __restrict__ beta, const float* __restrict__ stats,
                                                             double* __restrict__ red)
{
    const int c = blockIdx.y, n = blockIdx.z;
    const float* __restrict__ s = plane(z, n, c, hw);
    const float* __restrict__ gp = plane(gy, n, c, hw);
    const float mean = stats[c], invstd = stats[z.C + c], g = gamma[c], b = beta[c];
    const long long base = (long long)blockIdx.x * EW_CHUNK;
    float a0 = 0.f, a1 = 0.f;
#pragma unroll
    for (int k = 0; k < EW_PER; ++k) {
        const long long i = base + k * EW_T + threadIdx.x;
        if (i < hw) {
            float x = __ldg(s + i);
            float dv = __ldg(gp + i);
            if (has_mish && mish_first) x = mish_f(x);
            const float xhat = (x - mean) * invstd;
            if (has_mish && !mish_first) dv *= mish_grad_f(xhat * g + b);
            a0 += dv;
            a1 = fmaf(dv, xhat, a1);
        }
    }
    block_sum2_atomic((double)a0, (double)a1, red + c, red + z.C + c);
}

__global__ void __launch_bounds__(EW_T) bn_bwd_apply_kernel(const Slice z, const Slice gy, float* __restrict__ dz, const long long hw,
                                                            const int mish_first, const int has_mish,
                                                            const float* __restrict__ gamma, const float* __restrict__ beta,
                                                            const float* __restrict__ stats, const double* __restrict__ red,
                                                            const float invnum, float* __restrict__ dgamma, float* __restrict__ dbeta)
{
    const int c = blockIdx.y, n = blockIdx.z, C = z.C;
    const float* __restrict__ s = plane(z, n, c, hw);
    const float* __restrict__ gp = plane(gy, n, c, hw);
    float* __restrict__ d = dz + ((long long)n * C + c) * hw;
    const float mean = stats[c], invstd = stats[C + c], g = gamma[c], b = beta[c];
    const float s_dv = (float)red[c], s_dvx = (float)red[C + c];
    // cpu_dlib.cpp:1228-1288 with the per-channel sums factored out:
    //   dvars  = sum dx*(x-mean) * -0.5*invstd^3,  dx = dv*gamma  ->  -0.5*gamma*invstd^2 * sum(dv*xhat)
    //   dmeans = sum -dx*invstd   (the term with sum(x-mean) vanishes)
    const float dvars = -0.5f * g * invstd * invstd * s_dvx;
    const float dmeans = -g * invstd * s_dv;
    if (blockIdx.x == 0 && n == 0 && threadIdx.x == 0) { dgamma[c] = s_dvx; dbeta[c] = s_dv; }
    const long long base = (long long)blockIdx.x * EW_CHUNK;
#pragma unroll
    for (int k = 0; k < EW_PER; ++k) {
        const long long i = base + k * EW_T + threadIdx.x;
        if (i < hw) {
            const float zv = __ldg(s + i);
            float dv = __ldg(gp + i);
            const float x = (has_mish && mish_first) ? mish_f(zv) : zv;
            const float xc = x - mean;
            if (has_mish && !mish_first) dv *= mish_grad_f(xc * invstd * g + b);
            float r = dv * g * invstd + dvars * 2.f * xc * invnum + dmeans * invnum;
            if (has_mish && mish_first) r *= mish_grad_f(zv);
            d[i] = r;
        }
    }
}

// grid: (chunks over Hout*Wout, C, N)
__global__ void __launch_bounds__(256) maxpool_fwd_kernel(const Slice in, const Slice out, const int Hin, const int Win, const int Ho,
                                                          const int Wo)
{
    const int c = blockIdx.y, n = blockIdx.z;
    const long long i = (long long)blockIdx.x * blockDim.x + threadIdx.x;
    if (i >= (long long)Ho * Wo) return;
    const int y = (int)(i / Wo), x = (int)(i % Wo);
    const float* __restrict__ s = plane(in, n, c, (long long)Hin * Win) + (long long)(2 * y) * Win + 2 * x;
    float m = -INFINITY;
#pragma unroll
    for (int r = 0; r < 3; ++r)
#pragma unroll
        for (int q = 0; q < 3; ++q) m = fmaxf(m, __ldg(s + r * Win + q));
    plane_w(out, n, c, (long long)Ho * Wo)[i] = m;
}

// gather form of the max-pool gradient: an input pixel belongs to at most 2x2 windows; it receives a
// window's gradient iff it is that window's first maximum in row-major order (cpu_dlib.cpp:2003-2014).
__global__ void __launch_bounds__(256) maxpool_bwd_kernel(const Slice in, const Slice gout, const Slice gin, const int Hin, const int Win,
                                                          const int Ho, const int Wo)
{
    const int c = blockIdx.y, n = blockIdx.z;
    const long long i = (long long)blockIdx.x * blockDim.x + threadIdx.x;
    if (i >= (long long)Hin * Win) return;
    const int y = (int)(i / Win), x = (int)(i % Win);
    const float* __restrict__ s = plane(in, n, c, (long long)Hin * Win);
    const float* __restrict__ g = plane(gout, n, c, (long long)Ho * Wo);
    float acc = 0.f;
    const int Y1 = min(y / 2, Ho - 1), Y0 = max((y - 1) / 2, 0), X1 = min(x / 2, Wo - 1), X0 = max((x - 1) / 2, 0);
    for (int Y = (y >= 2 ? Y0 : 0); Y <= Y1; ++Y) {
        if (2 * Y > y || 2 * Y + 2 < y) continue;
        for (int X = (x >= 2 ? X0 : 0); X <= X1; ++X) {
            if (2 * X > x || 2 * X + 2 < x) continue;
            const float* wdw = s + (long long)(2 * Y) * Win + 2 * X;
            float m = -INFINITY;
            int arg = 0;
#pragma unroll
            for (int r = 0; r < 3; ++r)
#pragma unroll
                for (int q = 0; q < 3; ++q) {
                    const float v = __ldg(wdw + r * Win + q);
                    if (v > m) { m = v; arg = r * 3 + q; }
                }
            if (arg == (y - 2 * Y) * 3 + (x - 2 * X)) acc += __ldg(g + (long long)Y * Wo + X);
        }
    }
    if (acc != 0.f) plane_w(gin, n, c, (long long)Hin * Win)[i] += acc;
}

inline dim3 ew_grid(long long hw, int C, int N) { return dim3((unsigned)((hw + EW_CHUNK - 1) / EW_CHUNK), C, N); }

}  // namespace

void bn_stats(const Slice& z, int N, int H, int W, int mish_first, double* sums, cudaStream_t st)
{
    ProfScope ps(F_BN, 3.0 * N * H * W * z.C, 4.0 * N * H * W * z.C, st);
    cudaMemsetAsync(sums, 0, sizeof(double) * 2 * z.C, st);
    const long long hw = (long long)H * W;
    bn_stats_kernel<<<ew_grid(hw, z.C, N), EW_T, 0, st>>>(z, hw, mish_first, sums);
    count_launch();
}

void bn_finalize(const double* sums, int C, long long m, float eps, unsigned long long* num_updates, unsigned long long window,
                 float* running_mean, float* running_var, float* stats, cudaStream_t st)
{
    bn_finalize_kernel<<<1, (C + 31) / 32 * 32, 0, st>>>(sums, C, (double)m, (double)eps, num_updates, window, running_mean, running_var, stats);
    count_launch();
}

void bn_apply(const Slice& z, const Slice& y, int N, int H, int W, int mish_first, int has_mish, const float* gamma,
              const float* beta, const float* stats, cudaStream_t st)
{
    ProfScope ps(F_BN, 4.0 * N * H * W * z.C, 8.0 * N * H * W * z.C, st);
    const long long hw = (long long)H * W;
    bn_apply_kernel<<<ew_grid(hw, z.C, N), EW_T, 0, st>>>(z, y, hw, mish_first, has_mish, gamma, beta, stats);
    count_launch();
}

void bn_backward_reduce(const Slice& z, const Slice& gy, int N, int H, int W, int mish_first, int has_mish, const float* gamma,
                        const float* beta, const float* stats, double* red, cudaStream_t st)
{
    ProfScope ps(F_BN, 6.0 * N * H * W * z.C, 8.0 * N * H * W * z.C, st);
    cudaMemsetAsync(red, 0, sizeof(double) * 2 * z.C, st);
    const long long hw = (long long)H * W;
    bn_bwd_reduce_kernel<<<ew_grid(hw, z.C, N), EW_T, 0, st>>>(z, gy, hw, mish_first, has_mish, gamma, beta, stats, red);
    count_launch();
}

void bn_backward_apply(const Slice& z, const Slice& gy, float* dz, int N, int H, int W, int mish_first, int has_mish,
                       const float* gamma, const float* beta, const float* stats, const double* red, float* dgamma,
                       float* dbeta, cudaStream_t st)
{
    ProfScope ps(F_BN, 8.0 * N * H * W * z.C, 12.0 * N * H * W * z.C, st);
    const long long hw = (long long)H * W;
    const float invnum = 1.0f / (float)((long long)N * hw);
    bn_bwd_apply_kernel<<<ew_grid(hw, z.C, N), EW_T, 0, st>>>(z, gy, dz, hw, mish_first, has_mish, gamma, beta, stats, red, invnum,
                                                            dgamma, dbeta);
    count_launch();
}

void maxpool_forward(const Slice& in, const Slice& out, int N, int Hin, int Win, cudaStream_t st)
{
    ProfScope ps(F_POOL, 0, 4.0 * N * in.C * ((double)Hin * Win + (double)(1 + (Hin - 3) / 2) * (1 + (Win - 3) / 2)), st);
    const int Ho = 1 + (Hin - 3) / 2, Wo = 1 + (Win - 3) / 2;
    maxpool_fwd_kernel<<<dim3((unsigned)(((long long)Ho * Wo + 255) / 256), in.C, N), 256, 0, st>>>(in, out, Hin, Win, Ho, Wo);
    count_launch();
}

void maxpool_backward(const Slice& in, const Slice& gout, const Slice& gin, int N, int Hin, int Win, cudaStream_t st)
{
    ProfScope ps(F_POOL, 0, 4.0 * N * in.C * (3.0 * Hin * Win + (double)(1 + (Hin - 3) / 2) * (1 + (Win - 3) / 2)), st);
    const int Ho = 1 + (Hin - 3) / 2, Wo = 1 + (Win - 3) / 2;
    maxpool_bwd_kernel<<<dim3((unsigned)(((long long)Hin * Win + 255) / 256), in.C, N), 256, 0, st>>>(in, gout, gin, Hin, Win, Ho, Wo);
    count_launch();
}

}  // namespace k
}  // namespace ab
