// fp16 NHWC ("padded flat", see tc.h) elementwise / reduction kernels of the tensor-core path.
// All are HBM-bound: 16-byte vector accesses of 8 channels, fp32 arithmetic.
#include "h16.cuh"
#include "kcommon.cuh"
#include "tc.h"

namespace ab {
namespace tc {
namespace {

struct alignas(16) BF8 { __half2 v[4]; };

__device__ __forceinline__ void unpack8(const BF8& b, float f[8])
{
#pragma unroll
    for (int i = 0; i < 4; ++i) { const float2 t = __half22float2(b.v[i]); f[2 * i] = t.x; f[2 * i + 1] = t.y; }
}
__device__ __forceinline__ BF8 pack8_unused(const float f[8])
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

__device__ __forceinline__ BF8 ld8(const h16* p)   // one 16-byte load (a BF8 struct read is emitted as four 4-byte loads)
{
    BF8 r;
    uint32_t* u = reinterpret_cast<uint32_t*>(&r);
    asm volatile("ld.global.nc.v4.u32 {%0, %1, %2, %3}, [%4];" : "=r"(u[0]), "=r"(u[1]), "=r"(u[2]), "=r"(u[3]) : "l"(p));
    return r;
}
__device__ __forceinline__ void st8(h16* p, const BF8& v)   // one 16-byte store
{
    const uint32_t* u = reinterpret_cast<const uint32_t*>(&v);
    asm volatile("st.global.v4.u32 [%0], {%1, %2, %3, %4};" ::"l"(p), "r"(u[0]), "r"(u[1]), "r"(u[2]), "r"(u[3]) : "memory");
}

__global__ void __launch_bounds__(256) image_to_nhwc_kernel(const uint8_t* __restrict__ img, int IH, int IW, int left, int top,
                                                            h16* __restrict__ out, int N, int H, int W, float ar, float ag, float ab_,
                                                            long long lo_off)
{
    const long long hw = (long long)H * W;
    const long long i = (long long)blockIdx.x * blockDim.x + threadIdx.x;
    if (i >= (long long)N * hw) return;
    const int n = (int)(i / hw);
    const int rem = (int)(i - n * hw);
    const int y = rem / W, x = rem - y * W;
    const int sy = min(max(top + y, 0), IH - 1), sx = min(max(left + x, 0), IW - 1);
    const uint8_t* p = img + ((long long)n * IH * IW + (long long)sy * IW + sx) * 3;
    // (p - avg)/256 in double like input.h:96-115, then rounded once to h16
    float f[8] = {0, 0, 0, 0, 0, 0, 0, 0};
    f[0] = (float)(((double)p[0] - (double)ar) / 256.0);
    f[1] = (float)(((double)p[1] - (double)ag) / 256.0);
    f[2] = (float)(((double)p[2] - (double)ab_) / 256.0);
    const long long S = (long long)(H + 1) * (W + 1);
    const long long pos = n * S + (long long)(y + 1) * (W + 1) + (x + 1);
    BF8 hi, lo;
    split8(f, hi, lo);
    st8(out + pos * 8, hi);
    if (lo_off) st8(out + lo_off + pos * 8, lo);
}

__global__ void __launch_bounds__(256) maxpool_nhwc_kernel(const h16* __restrict__ in, int in_ctot, int c0, int C8, int N, int Hi, int Wi,
                                                           h16* __restrict__ out, int Ho, int Wo, uint8_t* __restrict__ arg, long long in_lo,
                                                           long long out_lo)
{
    k::pdl_enter();
    const long long total = (long long)N * Ho * Wo * C8;
    const long long i = (long long)blockIdx.x * blockDim.x + threadIdx.x;
    if (i >= total) return;
    const int cg = (int)(i % C8);
    long long p = i / C8;
    const int x = (int)(p % Wo); p /= Wo;
    const int y = (int)(p % Ho);
    const int n = (int)(p / Ho);
    const long long Si = (long long)(Hi + 1) * (Wi + 1), So = (long long)(Ho + 1) * (Wo + 1);
    float m[8];
    uint32_t a[8];
#pragma unroll
    for (int k = 0; k < 8; ++k) { m[k] = -INFINITY; a[k] = 0; }
#pragma unroll
    for (int r = 0; r < 3; ++r)
#pragma unroll
        for (int q = 0; q < 3; ++q) {
            const long long pi = n * Si + (long long)(2 * y + r + 1) * (Wi + 1) + (2 * x + q + 1);
            float f[8];
            unpack8(ld8(in + pi * in_ctot + c0 + cg * 8), f);
            if (in_lo) {   // split tensor: the value is hi + lo (exact in fp32)
                float l[8];
                unpack8(ld8(in + in_lo + pi * in_ctot + c0 + cg * 8), l);
#pragma unroll
                for (int k = 0; k < 8; ++k) f[k] += l[k];
            }
#pragma unroll
            for (int k = 0; k < 8; ++k)
                if (f[k] > m[k]) { m[k] = f[k]; a[k] = r * 3 + q; }   // first maximum in row-major order (cpu_dlib.cpp:2003-2014)
        }
    const long long po = n * So + (long long)(y + 1) * (Wo + 1) + (x + 1);
    BF8 hi, lo;
    split8(m, hi, lo);
    st8(out + po * (C8 * 8) + cg * 8, hi);
    if (out_lo) st8(out + out_lo + po * (C8 * 8) + cg * 8, lo);
    if (arg) {   // training: which window element won, one byte per channel, for the gradient routing
        uint2 w;
        w.x = a[0] | (a[1] << 8) | (a[2] << 16) | (a[3] << 24);
        w.y = a[4] | (a[5] << 8) | (a[6] << 16) | (a[7] << 24);
        *reinterpret_cast<uint2*>(arg + po * (C8 * 8) + cg * 8) = w;
    }
}

}  // namespace

void image_to_nhwc(const uint8_t* img, int IH, int IW, int left, int top, h16* out, const Geo& g, float ar, float ag, float ab_,
                   cudaStream_t st, long long lo_off)
{
    const long long n = (long long)g.N * g.H * g.W;
    k::ProfScope ps(k::F_INPUT, 0, 19.0 * n, st);
    image_to_nhwc_kernel<<<(unsigned)((n + 255) / 256), 256, 0, st>>>(img, IH, IW, left, top, out, g.N, g.H, g.W, ar, ag, ab_, lo_off);
    k::count_launch();
}

void maxpool_nhwc(const h16* in, int in_ctot, int c0, int C, const Geo& gi, h16* out, const Geo& go, cudaStream_t st, uint8_t* arg,
                  long long in_lo_off, long long out_lo_off)
{
    const long long n = (long long)go.N * go.H * go.W * (C / 8);
    k::ProfScope ps(k::F_EW_BF16, 0, (in_lo_off ? 4.0 : 2.0) * C * ((double)gi.N * gi.H * gi.W + (double)go.N * go.H * go.W), st);
    k::launch_pdl(maxpool_nhwc_kernel, dim3((unsigned)((n + 255) / 256)), dim3(256), 0, st, in, in_ctot, c0, C / 8, gi.N, gi.H, gi.W, out, go.H, go.W, arg, in_lo_off, out_lo_off);
    k::count_launch();
}

}  // namespace tc
}  // namespace ab
