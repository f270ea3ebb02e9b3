// Training-side fp16 NHWC ("padded flat", tc.h) elementwise / reduction kernels of the tensor-core path:
// batch normalisation in CONV_MODE fused with mish (cpu_dlib.cpp:1073-1289,1510-1558), the max-pool gradient
// (cpu_dlib.cpp:1971-2017), bias gradients (cpu_dlib.cpp:470-504) and the small layout conversions around them.
// All HBM-bound: 16-byte accesses of 8 channels, fp32 arithmetic, per-channel reductions in registers ->
// shared memory -> one double (or float) atomic per CTA and channel.
//
// Thread mapping of the per-channel kernels: blockDim = (C/8, TY); threadIdx.x owns 8 consecutive channels for the
// whole kernel, threadIdx.y strides over positions, so a warp reads consecutive 16-byte vectors of [pos][C].
#include "h16.cuh"
#include "kcommon.cuh"
#include "tc.h"

#include <algorithm>
#include <stdexcept>

namespace ab {
namespace tc {
namespace {

struct alignas(16) BF8 { __half2 v[4]; };

__device__ __forceinline__ void unpack8(const BF8& b, float f[8])
{
#pragma unroll
    for (int i = 0; i < 4; ++i) { const float2 t = __half22float2(b.v[i]); f[2 * i] = t.x; f[2 * i + 1] = t.y; }
}
__device__ __forceinline__ BF8 pack8(const float f[8])
{
    BF8 b;
#pragma unroll
    for (int i = 0; i < 4; ++i) b.v[i] = sat_half2(f[2 * i], f[2 * i + 1]);
    return b;
}
__device__ __forceinline__ void split8(const float f[8], BF8& hi, BF8& lo)
{
#pragma unroll
    for (int i = 0; i < 4; ++i) split_half2(f[2 * i], f[2 * i + 1], hi.v[i], lo.v[i]);
}
// 8 h16 of a channel slice.  The slice is 16..64 bytes of a wider row: ask L2 for 64-byte fetches instead of whole 128-byte
// lines (ncu: the backward bn passes read 269 MB of DRAM for a 67 MB slice, 4 x the algorithmic bytes)
__device__ __forceinline__ BF8 ld8(const h16* p)
{
    BF8 r;
    uint32_t* u = reinterpret_cast<uint32_t*>(&r);
    asm volatile("ld.global.L2::64B.v4.u32 {%0, %1, %2, %3}, [%4];" : "=r"(u[0]), "=r"(u[1]), "=r"(u[2]), "=r"(u[3]) : "l"(p));
    return r;
}
// 8 floats = one 32-byte sector in ONE load (LDG.E.256, sm_100): as two 16-byte loads every sector was requested twice (ncu on
// bn_apply: "16 of the 32 bytes per sector used per thread").  p is 32-byte aligned: rows and slices are multiples of 8 floats.
__device__ __forceinline__ void ldf8(const float* p, float f[8])
{
    asm volatile("ld.global.nc.v8.f32 {%0,%1,%2,%3,%4,%5,%6,%7}, [%8];"
                 : "=f"(f[0]), "=f"(f[1]), "=f"(f[2]), "=f"(f[3]), "=f"(f[4]), "=f"(f[5]), "=f"(f[6]), "=f"(f[7]) : "l"(p));
}
// one 16-byte store (a struct copy of BF8 is emitted member by member: four 4-byte stores -- seen in the SASS of bn_apply)
__device__ __forceinline__ void st8(h16* p, const BF8& v)
{
    const uint32_t* u = reinterpret_cast<const uint32_t*>(&v);
    asm volatile("st.global.v4.u32 [%0], {%1, %2, %3, %4};" ::"l"(p), "r"(u[0]), "r"(u[1]), "r"(u[2]), "r"(u[3]) : "memory");
}

struct GeoD { int N, H, W, P; unsigned S; long long total; };  // total = N * S  (< 2^31: checked by geod)
GeoD geod(const Geo& g)
{
    if (g.positions() >= (1ll << 31)) throw std::invalid_argument("tensor too large for the h16 NHWC kernels (positions must fit 31 bits)");
    return GeoD{g.N, g.H, g.W, g.P(), (unsigned)g.S(), (long long)g.N * g.S()};
}

// is `pos` a real pixel (not a pad position)?  32-bit arithmetic: positions fit 31 bits
__device__ __forceinline__ bool is_pixel(const GeoD& g, long long pos)
{
    if (pos >= g.total) return false;
    const unsigned rem = (unsigned)pos % g.S;
    return rem >= (unsigned)g.P && (rem % (unsigned)g.P) != 0;
}

// mish_ (cpu_dlib.cpp:1510-1558) on the fast-math intrinsics: the results feed h16 tensors or fp32 sums of h16-rounded
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
constexpr int kU = 2;   // positions per thread and loop iteration: their loads are issued together (memory-level parallelism)

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

// Per-channel constants live in SHARED memory (read back as float4 pairs), not in registers: the kernels below are
// bound by memory latency, so what matters is how many threads with loads in flight fit on an SM (B200 needs ~45 KB in
// flight per SM to saturate HBM).  Two positions per thread and iteration, loads issued before the first use.
constexpr int kMaxC = 256;
__device__ __forceinline__ void lds8(const float* p, float f[8])
{
    const float4 a = *reinterpret_cast<const float4*>(p), b = *reinterpret_cast<const float4*>(p + 4);
    f[0] = a.x; f[1] = a.y; f[2] = a.z; f[3] = a.w; f[4] = b.x; f[5] = b.y; f[6] = b.z; f[7] = b.w;
}

__global__ void __launch_bounds__(kMaxT, 4) bn_stats_kernel(const float* __restrict__ z, int ctot, int c0, long long npos, int mish_first,
                                                            double* __restrict__ sums, int rev)
{
    k::pdl_enter();
    const int C8 = blockDim.x;
    float acc[2][8] = {};
    const long long stride = (long long)gridDim.x * blockDim.y;
    for (long long p0 = (long long)blockIdx.x * blockDim.y + threadIdx.y; p0 < npos; p0 += kU * stride) {
        float f[kU][8];
#pragma unroll
        for (int u = 0; u < kU; ++u) {
            // rev: descending positions -- a producer that walked upwards left its last writes in L2
            const long long q = p0 + u * stride, pos = rev ? npos - 1 - q : q;
            if (q < npos) ldf8(z + pos * ctot + c0 + threadIdx.x * 8, f[u]);
            else {
#pragma unroll
                for (int i = 0; i < 8; ++i) f[u][i] = 0.f;   // mish(0) = 0: adds nothing
            }
        }
#pragma unroll
        for (int u = 0; u < kU; ++u)
#pragma unroll
            for (int i = 0; i < 8; ++i) {
                const float v = mish_first ? mish_fast(f[u][i]) : f[u][i];
                acc[0][i] += v;
                acc[1][i] = fmaf(v, v, acc[1][i]);
            }
    }
    double* const dst[2] = {sums, sums + C8 * 8};
    block_reduce_channels<2, double>(acc, dst);
}

// y = mish_first ? mish(z)*sc + sh : mish(z*sc + sh),   sc = gamma*invstd, sh = beta - mean*sc
template <int U>
__global__ void __launch_bounds__(kMaxT, U == 2 ? 4 : 3) bn_apply_kernel(const float* __restrict__ z, int z_ctot, int z_c0, h16* __restrict__ y, int y_ctot,
                                                            int y_c0, long long y_lo, GeoD g, int mish_first, int has_mish, const float* __restrict__ gamma,
                                                            const float* __restrict__ beta, const float* __restrict__ stats,
                                                            const BnFinalize fin, int rev)
{
    k::pdl_enter();
    __shared__ __align__(16) float s_sc[kMaxC], s_sh[kMaxC];
    const int C = blockDim.x * 8, c = threadIdx.x * 8;
    if (fin.sums) {
        // same arithmetic as k::bn_finalize_kernel
        const bool first = blockIdx.x == 0;
        const unsigned long long nu = first ? *fin.num_updates : 0;
        const double decay = 1.0 - (double)nu / ((double)nu + 1.0);
        __syncthreads();   // every thread of CTA 0 has read the counter before it advances
        if (first && threadIdx.x == 0 && threadIdx.y == 0) *fin.num_updates = nu + 1 > fin.window ? fin.window : nu + 1;
        for (int i = threadIdx.y * blockDim.x + threadIdx.x; i < C; i += blockDim.x * blockDim.y) {
            const float mean = (float)(fin.sums[i] / fin.m);
            const float sq = (float)(fin.sums[C + i] / fin.m);
            const float var = sq - mean * mean;
            const float invstd = (float)(1.0 / sqrt((double)var + fin.eps));
            if (first) {
                const double scale = fin.m / (fin.m - 1.0);  // unbiased running variance
                if (decay == 1.0) {
                    fin.rvar[i] = (float)(scale * var);
                    fin.rmean[i] = mean;
                } else {
                    fin.rvar[i] = (float)((1 - decay) * fin.rvar[i] + scale * decay * var);
                    fin.rmean[i] = (float)((1 - decay) * fin.rmean[i] + decay * mean);
                }
                fin.stats_out[i] = mean;
                fin.stats_out[C + i] = invstd;
            }
            const float sc = invstd * gamma[i];
            s_sc[i] = sc;
            s_sh[i] = beta[i] - mean * sc;
        }
    } else {
        for (int i = threadIdx.y * blockDim.x + threadIdx.x; i < C; i += blockDim.x * blockDim.y) {
            const float sc = stats[C + i] * gamma[i];
            s_sc[i] = sc;
            s_sh[i] = beta[i] - stats[i] * sc;
        }
    }
    __syncthreads();
    const long long stride = (long long)gridDim.x * blockDim.y;
    for (long long p0 = (long long)blockIdx.x * blockDim.y + threadIdx.y; p0 < g.total; p0 += U * stride) {
        float f[U][8];
        bool ok[U];
#pragma unroll
        for (int u = 0; u < U; ++u) {
            const long long q = p0 + u * stride, pos = rev ? g.total - 1 - q : q;
            ok[u] = q < g.total && is_pixel(g, pos);
            if (ok[u]) ldf8(z + pos * z_ctot + z_c0 + c, f[u]);
        }
        float sc[8], sh[8];
        lds8(s_sc + c, sc);
        lds8(s_sh + c, sh);
#pragma unroll
        for (int u = 0; u < U; ++u) {
            if (!ok[u]) continue;
#pragma unroll
            for (int i = 0; i < 8; ++i) {
                float v = f[u][i];
                if (has_mish && mish_first) v = mish_fast(v);
                v = fmaf(v, sc[i], sh[i]);
                if (has_mish && !mish_first) v = mish_fast(v);
                f[u][i] = v;
            }
            h16* yp = y + (rev ? g.total - 1 - (p0 + u * stride) : p0 + u * stride) * y_ctot + y_c0 + c;
            BF8 hi, lo;
            split8(f[u], hi, lo);
            st8(yp, hi);
            if (y_lo) st8(yp + y_lo, lo);   // split tensor (training forward): the rounding residual
        }
    }
}

// xhat = x*I + M (I = invstd, M = -mean*invstd);  conv->bn->mish layers: dv *= mish'(xhat*gamma + beta)
__global__ void __launch_bounds__(kMaxT, 3) bn_bwd_reduce_kernel(const float* __restrict__ z, int z_ctot, int z_c0, const h16* __restrict__ gy,
                                                                 int gy_ctot, int gy_c0, long long npos, int mish_first, int has_mish,
                                                                 const float* __restrict__ gamma, const float* __restrict__ beta,
                                                                 const float* __restrict__ stats, double* __restrict__ red, int rev)
{
    k::pdl_enter();
    __shared__ __align__(16) float s_i[kMaxC], s_m[kMaxC], s_g[kMaxC], s_b[kMaxC];
    const int C = blockDim.x * 8, c = threadIdx.x * 8;
    for (int i = threadIdx.y * blockDim.x + threadIdx.x; i < C; i += blockDim.x * blockDim.y) {
        s_i[i] = stats[C + i];
        s_m[i] = -stats[i] * stats[C + i];
        s_g[i] = gamma[i];
        s_b[i] = beta[i];
    }
    __syncthreads();
    const bool act_after = has_mish && !mish_first;
    float acc[2][8] = {};
    const long long stride = (long long)gridDim.x * blockDim.y;
    for (long long p0 = (long long)blockIdx.x * blockDim.y + threadIdx.y; p0 < npos; p0 += kU * stride) {
        float f[kU][8];
        BF8 dq[kU];
#pragma unroll
        for (int u = 0; u < kU; ++u) {
            const long long q = p0 + u * stride, pos = rev ? npos - 1 - q : q;
            if (q < npos) { ldf8(z + pos * z_ctot + z_c0 + c, f[u]); dq[u] = ld8(gy + pos * gy_ctot + gy_c0 + c); }
            else {
#pragma unroll
                for (int i = 0; i < 8; ++i) f[u][i] = 0.f;
#pragma unroll
                for (int i = 0; i < 4; ++i) dq[u].v[i] = __floats2half2_rn(0.f, 0.f);   // dv = 0: adds nothing
            }
        }
        float ci[8], cm[8];
        lds8(s_i + c, ci);
        lds8(s_m + c, cm);
#pragma unroll
        for (int u = 0; u < kU; ++u) {
            float d[8];
            unpack8(dq[u], d);
            if (act_after) {
                float cg[8], cb[8];
                lds8(s_g + c, cg);
                lds8(s_b + c, cb);
#pragma unroll
                for (int i = 0; i < 8; ++i) d[i] *= mish_grad_fast(fmaf(fmaf(f[u][i], ci[i], cm[i]), cg[i], cb[i]));
            }
#pragma unroll
            for (int i = 0; i < 8; ++i) {
                const float x = (has_mish && mish_first) ? mish_fast(f[u][i]) : f[u][i];
                const float xhat = fmaf(x, ci[i], cm[i]);
                acc[0][i] += d[i];
                acc[1][i] = fmaf(d[i], xhat, acc[1][i]);
            }
        }
    }
    double* const dst[2] = {red, red + C};
    block_reduce_channels<2, double>(acc, dst);
}

// cpu_dlib.cpp:1228-1288 with the per-channel sums factored out (see kernels_bn.cu):
//   dvars = -0.5*gamma*invstd^2 * sum(dv*xhat),  dmeans = -gamma*invstd * sum(dv)
//   r = dv*k1 + k2*x + k3,   k1 = gamma*invstd,  k2 = 2*dvars/num,  k3 = dmeans/num - k2*mean
// conv->mish->bn layers: x = mish(z), dz = r*mish'(z);   conv->bn->mish layers: x = z, dv *= mish'(x*A + B) first
__global__ void __launch_bounds__(kMaxT, 3) bn_bwd_apply_kernel(const float* __restrict__ z, int z_ctot, int z_c0, const h16* __restrict__ gy,
                                                                int gy_ctot, int gy_c0, h16* __restrict__ dz, int dz_ctot, int dz_c0, GeoD g,
                                                                int mish_first, int has_mish, const float* __restrict__ gamma,
                                                                const float* __restrict__ beta, const float* __restrict__ stats,
                                                                const double* __restrict__ red, float invnum, float* __restrict__ dgamma,
                                                                float* __restrict__ dbeta, float* __restrict__ dbias, int rev, float unscale)
{
    k::pdl_enter();
    __shared__ __align__(16) float s_k1[kMaxC], s_k2[kMaxC], s_k3[kMaxC], s_a[kMaxC], s_b[kMaxC];
    const int C = blockDim.x * 8, c = threadIdx.x * 8;
    for (int i = threadIdx.y * blockDim.x + threadIdx.x; i < C; i += blockDim.x * blockDim.y) {
        const float mean = stats[i], invstd = stats[C + i], ga = gamma[i], be = beta[i];
        const float s_dv = (float)red[i], s_dvx = (float)red[C + i];
        const float dvars = -0.5f * ga * invstd * invstd * s_dvx, dmeans = -ga * invstd * s_dv;
        const float k2 = dvars * 2.f * invnum;
        s_k1[i] = ga * invstd;
        s_k2[i] = k2;
        s_k3[i] = dmeans * invnum - k2 * mean;
        s_a[i] = invstd * ga;
        s_b[i] = be - mean * invstd * ga;
        if (blockIdx.x == 0) { dgamma[i] = s_dvx * unscale; dbeta[i] = s_dv * unscale; }
    }
    __syncthreads();
    const bool act_after = has_mish && !mish_first, act_before = has_mish && mish_first;
    float acc[1][8] = {};
    const long long stride = (long long)gridDim.x * blockDim.y;
    for (long long p0 = (long long)blockIdx.x * blockDim.y + threadIdx.y; p0 < g.total; p0 += kU * stride) {
        float f[kU][8];
        BF8 dq[kU];
        bool ok[kU];
#pragma unroll
        for (int u = 0; u < kU; ++u) {
            const long long q = p0 + u * stride, pos = rev ? g.total - 1 - q : q;
            ok[u] = q < g.total && is_pixel(g, pos);
            if (ok[u]) { ldf8(z + pos * z_ctot + z_c0 + c, f[u]); dq[u] = ld8(gy + pos * gy_ctot + gy_c0 + c); }
        }
        float k1[8], k2[8], k3[8];
        lds8(s_k1 + c, k1);
        lds8(s_k2 + c, k2);
        lds8(s_k3 + c, k3);
#pragma unroll
        for (int u = 0; u < kU; ++u) {
            if (!ok[u]) continue;
            float d[8];
            unpack8(dq[u], d);
            if (act_after) {
                float ca[8], cb[8];
                lds8(s_a + c, ca);
                lds8(s_b + c, cb);
#pragma unroll
                for (int i = 0; i < 8; ++i) d[i] *= mish_grad_fast(fmaf(f[u][i], ca[i], cb[i]));
            }
#pragma unroll
            for (int i = 0; i < 8; ++i) {
                float x = f[u][i], dm = 1.f;
                if (act_before) mish_both(f[u][i], x, dm);
                const float r = fmaf(d[i], k1[i], fmaf(k2[i], x, k3[i])) * dm;
                d[i] = r;
                acc[0][i] += r;
            }
            st8(dz + (rev ? g.total - 1 - (p0 + u * stride) : p0 + u * stride) * dz_ctot + dz_c0 + c, pack8(d));
        }
    }
#pragma unroll
    for (int i = 0; i < 8; ++i) acc[0][i] *= unscale;
    float* const dst[1] = {dbias};
    block_reduce_channels<1, float>(acc, dst);
}

__global__ void __launch_bounds__(kMaxT) channel_sum_kernel(const h16* __restrict__ v, int ctot, int c0, long long npos, float* __restrict__ out,
                                                            float unscale)
{
    k::pdl_enter();
    float acc[1][8] = {};
    for (long long pos = (long long)blockIdx.x * blockDim.y + threadIdx.y; pos < npos; pos += (long long)gridDim.x * blockDim.y) {
        float f[8];
        unpack8(ld8(v + pos * ctot + c0 + threadIdx.x * 8), f);
#pragma unroll
        for (int i = 0; i < 8; ++i) acc[0][i] += f[i];
    }
#pragma unroll
    for (int i = 0; i < 8; ++i) acc[0][i] *= unscale;
    float* const dst[1] = {out};
    block_reduce_channels<1, float>(acc, dst);
}

// gather form: an input pixel belongs to at most 2 x 2 windows and receives a window's gradient iff it is that window's
// first maximum in row-major order -- which element won is read from the byte map the forward pass wrote
__global__ void __launch_bounds__(256) maxpool_bwd_kernel(const uint8_t* __restrict__ arg, int C8, int N, int Hi, int Wi,
                                                          const h16* __restrict__ gout, int Ho, int Wo, h16* __restrict__ gin, int gin_ctot,
                                                          int gin_c0)
{
    k::pdl_enter();
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
    h16* gp = gin + (n * Si + (long long)(y + 1) * (Wi + 1) + (x + 1)) * gin_ctot + gin_c0 + cg * 8;
    const BF8 a = pack8(acc);   // one 16-byte reduction at the L2 instead of load + add + store (h16.cuh)
    red_add_h8(gp, reinterpret_cast<const uint32_t*>(a.v));
}

__global__ void __launch_bounds__(256) planes_to_nhwc8_kernel(const float* __restrict__ planes, int K, h16* __restrict__ out, int N, int H, int W,
                                                              float scale)
{
    const long long hw = (long long)H * W;
    const long long i = (long long)blockIdx.x * blockDim.x + threadIdx.x;
    if (i >= (long long)N * hw) return;
    const int n = (int)(i / hw);
    const long long rem = i - n * hw;
    const int y = (int)(rem / W), x = (int)(rem - (long long)y * W);
    const long long pos = n * (long long)(H + 1) * (W + 1) + (long long)(y + 1) * (W + 1) + (x + 1);
    const int Kp = (K + 7) / 8 * 8;
    for (int k0 = 0; k0 < Kp; k0 += 8) {
        float f[8] = {0, 0, 0, 0, 0, 0, 0, 0};
        for (int k = k0; k < K && k < k0 + 8; ++k) f[k - k0] = planes[((long long)n * K + k) * hw + rem] * scale;
        st8(out + pos * Kp + k0, pack8(f));
    }
}

__global__ void __launch_bounds__(256) upsample_zero_kernel(const h16* __restrict__ in, int in_ctot, int in_c0, int C8, int N, int h, int w,
                                                            h16* __restrict__ out)
{
    k::pdl_enter();
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
    for (int k = 0; k < 4; ++k) v.v[k] = __floats2half2_rn(0.f, 0.f);
    if (!(y & 1) && !(x & 1) && (y >> 1) < h && (x >> 1) < w)
        v = ld8(in + (n * (long long)(h + 1) * (w + 1) + (long long)((y >> 1) + 1) * (w + 1) + ((x >> 1) + 1)) * in_ctot + in_c0 + cg * 8);
    st8(out + (n * (long long)(Hf + 1) * (Wf + 1) + (long long)(y + 1) * (Wf + 1) + (x + 1)) * (C8 * 8) + cg * 8, v);
}

__global__ void __launch_bounds__(256) resize_nhwc_kernel(const h16* __restrict__ in, long long in_lo, int C8, int N, int ih, int iw, h16* __restrict__ out,
                                                          int out_ctot, int out_c0, long long out_lo, int oh, int ow, const ResizeTabDev t)
{
    const long long total = (long long)N * oh * ow * C8;
    const long long i = (long long)blockIdx.x * blockDim.x + threadIdx.x;
    if (i >= total) return;
    const int cg = (int)(i % C8);
    long long p = i / C8;
    const int x = (int)(p % ow); p /= ow;
    const int y = (int)(p % oh);
    const int n = (int)(p / oh);
    const long long Si = (long long)(ih + 1) * (iw + 1), So = (long long)(oh + 1) * (ow + 1);
    const int top = t.top[y], bottom = t.bottom[y], left = t.left[x], right = t.right[x];
    const float wtl = __fmul_rn(t.itb[y], t.ilr[x]), wtr = __fmul_rn(t.itb[y], t.lr[x]), wbl = __fmul_rn(t.tb[y], t.ilr[x]), wbr = __fmul_rn(t.tb[y], t.lr[x]);
    auto load = [&](int yy, int xx, float f[8]) {
        const long long e = (n * Si + (long long)(yy + 1) * (iw + 1) + (xx + 1)) * (C8 * 8) + cg * 8;
        unpack8(ld8(in + e), f);
        if (in_lo) {
            float l[8];
            unpack8(ld8(in + in_lo + e), l);
#pragma unroll
            for (int k = 0; k < 8; ++k) f[k] += l[k];
        }
    };
    float tl[8], tr[8], bl[8], br[8], o[8];
    load(top, left, tl); load(top, right, tr); load(bottom, left, bl); load(bottom, right, br);
#pragma unroll
    for (int k = 0; k < 8; ++k)
        o[k] = __fadd_rn(__fadd_rn(__fadd_rn(__fmul_rn(wtl, tl[k]), __fmul_rn(wtr, tr[k])), __fmul_rn(wbl, bl[k])), __fmul_rn(wbr, br[k]));
    h16* op = out + (n * So + (long long)(y + 1) * (ow + 1) + (x + 1)) * out_ctot + out_c0 + cg * 8;
    BF8 hi, lo;
    split8(o, hi, lo);
    st8(op, hi);
    if (out_lo) st8(op + out_lo, lo);
}

__global__ void __launch_bounds__(256) pack_batched_kernel(const float* __restrict__ arena, h16* __restrict__ packs, const PackJob* __restrict__ jobs)
{
    const PackJob& j = jobs[blockIdx.y];
    const int rows = j.npad * j.parts;   // N rows per tap: [hi filters | lo filters] when the operands are split
    const long long total = (long long)j.ntaps * rows * j.kp;
    const float* w = arena + j.w_off;
    h16* packed = packs + j.packed_off;
    for (long long i = (long long)blockIdx.x * blockDim.x + threadIdx.x; i < total; i += (long long)gridDim.x * blockDim.x) {
        const int kidx = (int)(i % j.kp);
        const int rr = (int)((i / j.kp) % rows);
        const int part = rr / j.npad, n = rr - part * j.npad;
        const int t = (int)(i / ((long long)j.kp * rows));
        float v = 0.f;
        if (n < j.nreal && kidx < j.kreal) v = w[n * j.sn + kidx * j.sk + j.tapmap[t]];
        const h16 hi = __float2half_rn(fminf(fmaxf(v, -65504.f), 65504.f));
#ifdef AB_ZERO_LO
        packed[i] = part == 1 ? __float2half_rn(0.f) : hi;
#else
        packed[i] = part == 1 ? __float2half_rn(v - __half2float(hi)) : hi;
#endif
    }
}

struct Shape { dim3 block; unsigned grid; };
// ctas_per_sm: the kernel's residency (__launch_bounds__): a grid of exactly the resident CTAs runs as ONE wave of equal
// grid-stride shares -- no partially filled last wave, and the per-CTA prologues (constants, fused bn_finalize) run once per slot
Shape channel_shape(int C, long long npos, int ctas_per_sm = 8)
{
    const int C8 = C / 8;
    if (C % 8 || C8 < 1 || C8 > 32) throw std::invalid_argument("h16 NHWC kernels need 8 <= C <= 256, C % 8 == 0");
    const int TY = kMaxT / C8;
    const long long want = (npos + TY - 1) / TY;
    return Shape{dim3(C8, TY), (unsigned)std::max<long long>(1, std::min<long long>(want, (long long)k::num_sms() * ctas_per_sm))};
}

}  // namespace

void bn_stats_nhwc(const ViewF& z, const Geo& g, int mish_first, double* sums, cudaStream_t st, int rev, int pre_zeroed)
{
    const GeoD d = geod(g);
    k::ProfScope ps(k::F_EW_BF16, 3.0 * g.N * g.H * g.W * z.C, 2.0 * g.N * g.H * g.W * z.C, st);
    if (!pre_zeroed) cudaMemsetAsync(sums, 0, sizeof(double) * 2 * z.C, st);
    const Shape s = channel_shape(z.C, d.total, 4);
    k::launch_pdl(bn_stats_kernel, s.grid, s.block, 0, st, z.p, z.ctot, z.c0, d.total, mish_first, sums, rev);
    k::count_launch();
}

void bn_apply_nhwc(const ViewF& z, const View& y, const Geo& g, int mish_first, int has_mish, const float* gamma, const float* beta,
                   const float* stats, cudaStream_t st, const BnFinalize* fin, int rev)
{
    const GeoD d = geod(g);
    k::ProfScope ps(k::F_EW_BF16, 4.0 * g.N * g.H * g.W * z.C, (y.lo_off ? 8.0 : 6.0) * g.N * g.H * g.W * z.C, st);
    // positions per thread and loop iteration: ncu on the level-0 layers shows the kernel waiting on memory (17 warps stalled on the
    // long scoreboard per issue, 33 % issue activity), yet four positions per thread at 3 CTAs per SM are no faster than two at 4
    static const int unroll = getenv("AB_BN_U") ? atoi(getenv("AB_BN_U")) : 2;   // measured: 2 -> 7.96, 4 -> 7.98 ms per step
    if (unroll == 2) {
        const Shape s = channel_shape(z.C, d.total, 4);
        k::launch_pdl(bn_apply_kernel<2>, s.grid, s.block, 0, st, z.p, z.ctot, z.c0, y.p, y.ctot, y.c0, y.lo_off, d, mish_first, has_mish, gamma, beta, stats, fin ? *fin : BnFinalize{}, rev);
    } else {
        const Shape s = channel_shape(z.C, d.total, 3);
        k::launch_pdl(bn_apply_kernel<4>, s.grid, s.block, 0, st, z.p, z.ctot, z.c0, y.p, y.ctot, y.c0, y.lo_off, d, mish_first, has_mish, gamma, beta, stats, fin ? *fin : BnFinalize{}, rev);
    }
    k::count_launch();
}

void bn_backward_reduce_nhwc(const ViewF& z, const View& gy, const Geo& g, int mish_first, int has_mish, const float* gamma,
                             const float* beta, const float* stats, double* red, cudaStream_t st, int rev, int pre_zeroed)
{
    const GeoD d = geod(g);
    k::ProfScope ps(k::F_EW_BF16, 6.0 * g.N * g.H * g.W * z.C, 4.0 * g.N * g.H * g.W * z.C, st);
    if (!pre_zeroed) cudaMemsetAsync(red, 0, sizeof(double) * 2 * z.C, st);
    const Shape s = channel_shape(z.C, d.total, 3);
    k::launch_pdl(bn_bwd_reduce_kernel, s.grid, s.block, 0, st, z.p, z.ctot, z.c0, gy.p, gy.ctot, gy.c0, d.total, mish_first, has_mish, gamma, beta, stats,
                                                     red, rev);
    k::count_launch();
}

void bn_backward_apply_nhwc(const ViewF& z, const View& gy, const View& dz, const Geo& g, int mish_first, int has_mish, const float* gamma,
                            const float* beta, const float* stats, const double* red, float* dgamma, float* dbeta, float* dbias,
                            cudaStream_t st, int rev, float unscale)
{
    const GeoD d = geod(g);
    k::ProfScope ps(k::F_EW_BF16, 8.0 * g.N * g.H * g.W * z.C, 6.0 * g.N * g.H * g.W * z.C, st);
    const Shape s = channel_shape(z.C, d.total, 3);
    const float invnum = 1.0f / (float)((long long)g.N * g.H * g.W);
    k::launch_pdl(bn_bwd_apply_kernel, s.grid, s.block, 0, st, z.p, z.ctot, z.c0, gy.p, gy.ctot, gy.c0, dz.p, dz.ctot, dz.c0, d, mish_first, has_mish, gamma, beta, stats, red,
                                                    invnum, dgamma, dbeta, dbias, rev, unscale);
    k::count_launch();
}

void channel_sum_nhwc(const View& v, const Geo& g, float* out, cudaStream_t st, float unscale)
{
    const GeoD d = geod(g);
    k::ProfScope ps(k::F_EW_BF16, 1.0 * g.N * g.H * g.W * v.C, 2.0 * g.N * g.H * g.W * v.C, st);
    const Shape s = channel_shape(v.C, d.total);
    k::launch_pdl(channel_sum_kernel, s.grid, s.block, 0, st, v.p, v.ctot, v.c0, d.total, out, unscale);
    k::count_launch();
}

void maxpool_backward_nhwc(const uint8_t* arg, int C, const Geo& gi, const h16* gout_dense, const Geo& go, const View& gin, cudaStream_t st)
{
    const long long n = (long long)gi.N * gi.H * gi.W * (C / 8);
    if (n >= (1ll << 31) * 8) throw std::invalid_argument("maxpool_backward_nhwc: tensor too large");
    k::ProfScope ps(k::F_EW_BF16, 0, 2.0 * C * (2.0 * gi.N * gi.H * gi.W + 1.5 * (double)go.N * go.H * go.W), st);
    k::launch_pdl(maxpool_bwd_kernel, dim3((unsigned)((n + 255) / 256)), dim3(256), 0, st, arg, C / 8, gi.N, gi.H, gi.W, gout_dense, go.H, go.W, gin.p, gin.ctot, gin.c0);
    k::count_launch();
}

void planes_to_nhwc8(const float* planes, int K, h16* out, const Geo& g, cudaStream_t st, float scale)
{
    if (K > 64) throw std::invalid_argument("planes_to_nhwc8: at most 64 planes");
    const long long n = (long long)g.N * g.H * g.W;
    k::ProfScope ps(k::F_PACK, 0, (4.0 * K + 16.0) * n, st);
    planes_to_nhwc8_kernel<<<(unsigned)((n + 255) / 256), 256, 0, st>>>(planes, K, out, g.N, g.H, g.W, scale);
    k::count_launch();
}

void upsample_zero_nhwc(const View& in, const Geo& gc, h16* out, const Geo& gf, cudaStream_t st)
{
    if (gf.H != 2 * gc.H + 1 || gf.W != 2 * gc.W + 1) throw std::logic_error("upsample_zero_nhwc: geometry mismatch");
    const long long n = (long long)gf.N * gf.H * gf.W * (in.C / 8);
    k::ProfScope ps(k::F_PACK, 0, 2.0 * in.C * ((double)gc.N * gc.H * gc.W + (double)gf.N * gf.H * gf.W), st);
    k::launch_pdl(upsample_zero_kernel, dim3((unsigned)((n + 255) / 256)), dim3(256), 0, st, in.p, in.ctot, in.c0, in.C / 8, gc.N, gc.H, gc.W, out);
    k::count_launch();
}

void resize_bilinear_nhwc(const h16* in, long long in_lo_off, int C, const Geo& gi, const View& out, const Geo& go, const ResizeTabDev& t, cudaStream_t st)
{
    const long long n = (long long)go.N * go.H * go.W * (C / 8);
    k::ProfScope ps(k::F_EW_BF16, 0, (in_lo_off ? 4.0 : 2.0) * C * ((double)gi.N * gi.H * gi.W + (double)go.N * go.H * go.W), st);
    resize_nhwc_kernel<<<(unsigned)((n + 255) / 256), 256, 0, st>>>(in, in_lo_off, C / 8, go.N, gi.H, gi.W, out.p, out.ctot, out.c0, out.lo_off, go.H, go.W, t);
    k::count_launch();
}

void pack_filters_batched(const float* arena, h16* packs, const PackJob* jobs_dev, int njobs, int max_elems, cudaStream_t st)
{
    if (njobs <= 0) return;
    k::ProfScope ps(k::F_PACK, 0, 0, st);
    const unsigned gx = (unsigned)std::max(1, std::min((max_elems + 255) / 256, 64));
    pack_batched_kernel<<<dim3(gx, njobs), 256, 0, st>>>(arena, packs, jobs_dev);
    k::count_launch();
}

}  // namespace tc
}  // namespace ab
