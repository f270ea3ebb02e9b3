// Input conversion, loss, label maps, tile blending and the optimizer: coalesced HBM-bound kernels.
#include "kcommon.cuh"

namespace ab {
namespace k {
namespace {

// input_rgb_image::to_tensor (input.h:96-115): (p - avg)/256, planar.  grid: (chunks, N)
__global__ void __launch_bounds__(256) to_tensor_kernel(const uint8_t* __restrict__ img, float* __restrict__ x, const long long hw,
                                                        const float ar, const float ag, const float ab_)
{
    const int n = blockIdx.y;
    const long long i = (long long)blockIdx.x * blockDim.x + threadIdx.x;
    if (i >= hw) return;
    const uint8_t* p = img + ((long long)n * hw + i) * 3;
    float* o = x + (long long)n * 3 * hw + i;
    o[0] = (float)(((double)((float)p[0] - ar)) / 256.0);
    o[hw] = (float)(((double)((float)p[1] - ag)) / 256.0);
    o[2 * hw] = (float)(((double)((float)p[2] - ab_)) / 256.0);
}

__global__ void __launch_bounds__(256) extract_tile_kernel(const uint8_t* __restrict__ img, const int IH, const int IW, const int left,
                                                           const int top, float* __restrict__ x, const int TH, const int TW,
                                                           const float ar, const float ag, const float ab_)
{
    const long long hw = (long long)TH * TW;
    const long long i = (long long)blockIdx.x * blockDim.x + threadIdx.x;
    if (i >= hw) return;
    const int ty = (int)(i / TW), tx = (int)(i % TW);
    const int sy = min(max(top + ty, 0), IH - 1), sx = min(max(left + tx, 0), IW - 1);
    const uint8_t* p = img + ((long long)sy * IW + sx) * 3;
    x[i] = (float)(((double)((float)p[0] - ar)) / 256.0);
    x[hw + i] = (float)(((double)((float)p[1] - ag)) / 256.0);
    x[2 * hw + i] = (float)(((double)((float)p[2] - ab_)) / 256.0);
}

struct WLabel { uint16_t label; uint16_t pad; float weight; };

// loss.h:3155-3190 with cpu::softmax (cpu_dlib.cpp:1326-1383); one thread per pixel, grid (chunks, N)
__global__ void __launch_bounds__(256) loss_kernel(const float* __restrict__ logits, const WLabel* __restrict__ wl, float* __restrict__ grad,
                                                   const int K, const long long hw, const double scale, double* __restrict__ loss_sum)
{
    __shared__ double red[8];
    const int n = blockIdx.y;
    const long long i = (long long)blockIdx.x * blockDim.x + threadIdx.x;
    double loss = 0.0;
    if (i < hw) {
        const float* s = logits + (long long)n * K * hw + i;
        float* g = grad + (long long)n * K * hw + i;
        float mx = -INFINITY;
        for (int k = 0; k < K; ++k) mx = fmaxf(mx, __ldg(s + (long long)k * hw));
        float sum = 0.f;
        for (int k = 0; k < K; ++k) sum += expf(__ldg(s + (long long)k * hw) - mx);
        const WLabel l = wl[(long long)n * hw + i];
        const double ws = (double)l.weight * scale;
        for (int k = 0; k < K; ++k) {
            const float p = expf(__ldg(s + (long long)k * hw) - mx) / sum;
            if (k == (int)l.label) {
                loss += ws * (double)(-logf(fmaxf(p, 1e-10f)));
                g[(long long)k * hw] = (float)(ws * (double)(p - 1.f));
            } else
                g[(long long)k * hw] = (float)(ws * (double)p);
        }
    }
    loss = warp_sum(loss);
    if ((threadIdx.x & 31) == 0) red[threadIdx.x >> 5] = loss;
    __syncthreads();
    if (threadIdx.x == 0) {
        double t = 0;
        for (int w = 0; w < 8; ++w) t += red[w];
        if (t != 0.0) atomicAdd(loss_sum, t);
    }
}

__global__ void __launch_bounds__(256) argmax_kernel(const float* __restrict__ planes, const long long plane_stride, const int K,
                                                     const long long npix, const double* __restrict__ gains, uint16_t* __restrict__ labels,
                                                     const double* __restrict__ det, uint8_t* __restrict__ seeds)
{
    const long long i = (long long)blockIdx.x * blockDim.x + threadIdx.x;
    if (i >= npix) return;
    uint16_t label = 65535;
    float best = -INFINITY;
    for (int k = 0; k < K; ++k) {
        const float raw = __ldg(planes + (long long)k * plane_stride + i);
        const float v = gains ? (float)((double)raw + gains[k]) : raw;
        if (v > best) { best = v; label = (uint16_t)k; }
    }
    labels[i] = label;
    if (seeds) {   // detection seed (annonet_infer.cpp:204-212): the winner beats the clean class by more than its detection level
        bool s = false;
        if (label > 0 && label < K) {
            const float clean = __ldg(planes + i), mine = __ldg(planes + (long long)label * plane_stride + i);
            s = (double)(mine - clean) > det[label] - det[0];
        }
        seeds[i] = s ? 1 : 0;
    }
}

__device__ __forceinline__ double ramp(long long c, long long first_possible, long long first_in, long long last_in, long long last_possible)
{
    if (c < first_in) return (double)(c - first_possible) / (double)(first_in - first_possible);
    if (c > last_in) return (double)(last_possible - c) / (double)(last_possible - last_in);
    return 1.0;
}

// annonet_infer.cpp:116-164; one thread per tile pixel
__global__ void __launch_bounds__(256) blend_kernel(const BlendArgs a)
{
    const long long i = (long long)blockIdx.x * blockDim.x + threadIdx.x;
    if (i >= (long long)a.th * a.tw) return;
    const int y = (int)(i / a.tw), x = (int)(i % a.tw);
    const int by = a.top + y, bx = a.left + x;
    if (by < a.full[1] || by > a.full[3] || bx < a.full[0] || bx > a.full[2]) return;
    if (by < 0 || by >= a.H || bx < 0 || bx >= a.W) return;
    const bool unique = bx >= a.unique[0] && bx <= a.unique[2] && by >= a.unique[1] && by <= a.unique[3];
    const long long hw = (long long)a.H * a.W, thw = (long long)a.th * a.tw;
    float* o = a.canvas + (long long)by * a.W + bx;
    const float* s = a.logits + i;
    if (unique) {
        for (int k = 0; k < a.K; ++k) o[k * hw] = __ldg(s + k * thw);
    } else {
        const double t = ramp(bx, a.full[0], a.unique[0], a.unique[2], a.full[2]) * ramp(by, a.full[1], a.unique[1], a.unique[3], a.full[3]);
        for (int k = 0; k < a.K; ++k) o[k * hw] = (float)((double)o[k * hw] + t * (double)__ldg(s + k * thw));
    }
}

__global__ void __launch_bounds__(256) sgd_kernel(float* __restrict__ w, const float* __restrict__ g, float* __restrict__ v,
                                                  const float* __restrict__ lr_mul, const float* __restrict__ wd_mul, const long long n,
                                                  const float lr, const float wd, const float momentum, const float grad_scale)
{
    const long long i = (long long)blockIdx.x * blockDim.x + threadIdx.x;
    if (i >= n) return;
    const double lrm = (double)lr * (double)lr_mul[i];
    const float B = (float)(-((double)wd * (double)wd_mul[i]) * lrm);
    const float C = (float)(-lrm);
    const float gi = grad_scale == 1.f ? g[i] : g[i] * grad_scale;
    // tt::affine_transform(v, v, params, params_grad, momentum, B, C), solvers.h:177-184
    const float nv = __fadd_rn(__fadd_rn(__fmul_rn(momentum, v[i]), __fmul_rn(B, w[i])), __fmul_rn(C, gi));
    v[i] = nv;
    w[i] = w[i] + nv;
}

__global__ void bump_kernel(float* t) { *t += 1.f; }
__global__ void loss_to_slot_kernel(const double* loss, float* slot) { *slot = (float)*loss; }
__global__ void slot_to_loss_kernel(const float* slot, double* loss, float inv_world) { *loss = (double)(*slot * inv_world); }

__global__ void __launch_bounds__(256) adam_kernel(float* __restrict__ w, const float* __restrict__ g, float* __restrict__ m,
                                                   float* __restrict__ v, const float* __restrict__ lr_mul, const float* __restrict__ wd_mul,
                                                   const long long n, const float* __restrict__ t_dev, const float lr, const float wd,
                                                   const float b1, const float b2, const float grad_scale)
{
    const long long i = (long long)blockIdx.x * blockDim.x + threadIdx.x;
    if (i >= n) return;
    // bias correction from the device-side step counter (a kernel argument would pin the step into a captured CUDA graph)
    const float t = *t_dev;
    const float alpha_base = sqrtf(1.f - powf(b2, t)) / (1.f - powf(b1, t));
    const float lri = (float)((double)lr * (double)lr_mul[i]);
    const float wdi = (float)((double)wd * (double)wd_mul[i]);
    const float alpha = lri * alpha_base;
    const float gg = wdi * w[i] + g[i] * grad_scale;
    const float mi = b1 * m[i] + (1 - b1) * gg;
    const float vi = b2 * v[i] + (1 - b2) * gg * gg;
    m[i] = mi;
    v[i] = vi;
    w[i] += -alpha * mi / (sqrtf(vi) + 1e-8f);
}

}  // namespace

void to_tensor(const uint8_t* img, float* x, int N, int H, int W, float ar, float ag, float ab_, cudaStream_t st)
{
    ProfScope ps(F_INPUT, 0, 15.0 * N * H * W, st);
    const long long hw = (long long)H * W;
    to_tensor_kernel<<<dim3((unsigned)((hw + 255) / 256), N), 256, 0, st>>>(img, x, hw, ar, ag, ab_);
    count_launch();
}

void extract_tile(const uint8_t* img, int IH, int IW, int left, int top, float* x, int TH, int TW, float ar, float ag, float ab_,
                  cudaStream_t st)
{
    ProfScope ps(F_INPUT, 0, 15.0 * TH * TW, st);
    const long long hw = (long long)TH * TW;
    extract_tile_kernel<<<(unsigned)((hw + 255) / 256), 256, 0, st>>>(img, IH, IW, left, top, x, TH, TW, ar, ag, ab_);
    count_launch();
}

void loss_forward_backward(const float* logits, const void* wlabels, float* grad, int N, int K, int H, int W, double* loss_sum,
                           cudaStream_t st)
{
    ProfScope ps(F_LOSS, 0, (8.0 * K + 8.0) * N * H * W, st);
    const long long hw = (long long)H * W;
    cudaMemsetAsync(loss_sum, 0, sizeof(double), st);
    const double scale = 1.0 / ((double)N * (double)H * (double)W);
    loss_kernel<<<dim3((unsigned)((hw + 255) / 256), N), 256, 0, st>>>(logits, (const WLabel*)wlabels, grad, K, hw, scale, loss_sum);
    count_launch();
}

void argmax_labels(const float* planes, long long plane_stride, int K, long long npix, const double* gains_dev, uint16_t* labels,
                   cudaStream_t st, const double* det_dev, uint8_t* seeds)
{
    ProfScope ps(F_ARGMAX, 0, (4.0 * K + 2.0) * npix, st);
    argmax_kernel<<<(unsigned)((npix + 255) / 256), 256, 0, st>>>(planes, plane_stride, K, npix, gains_dev, labels, det_dev, seeds);
    count_launch();
}

void blend_tile(const BlendArgs& a, cudaStream_t st)
{
    ProfScope ps(F_BLEND, 0, 12.0 * a.K * a.th * a.tw, st);
    const long long n = (long long)a.th * a.tw;
    blend_kernel<<<(unsigned)((n + 255) / 256), 256, 0, st>>>(a);
    count_launch();
}

void sgd_step(float* w, const float* g, float* v, const float* lr_mul, const float* wd_mul, long long n, float lr, float wd,
              float momentum, float grad_scale, cudaStream_t st)
{
    ProfScope ps(F_OPTIMIZER, 0, 28.0 * n, st);
    sgd_kernel<<<(unsigned)((n + 255) / 256), 256, 0, st>>>(w, g, v, lr_mul, wd_mul, n, lr, wd, momentum, grad_scale);
    count_launch();
}

void adam_step(float* w, const float* g, float* m, float* v, const float* lr_mul, const float* wd_mul, long long n,
               float* t_dev, float lr,
               float wd, float beta1, float beta2, float grad_scale, cudaStream_t st)
{
    ProfScope ps(F_OPTIMIZER, 0, 36.0 * n, st);
    bump_kernel<<<1, 1, 0, st>>>(t_dev);   // t = t + 1 (solvers.h:262: the first step runs with t = 1)
    adam_kernel<<<(unsigned)((n + 255) / 256), 256, 0, st>>>(w, g, m, v, lr_mul, wd_mul, n, t_dev, lr, wd, beta1, beta2, grad_scale);
    count_launch(2);
}

void fill_zero(void* p, size_t bytes, cudaStream_t st) { cudaMemsetAsync(p, 0, bytes, st); }

__global__ void __launch_bounds__(256) add_inplace_kernel(float* __restrict__ dst, const float* __restrict__ src, long long n)
{
    for (long long i = ((long long)blockIdx.x * blockDim.x + threadIdx.x) * 4; i < n; i += (long long)gridDim.x * blockDim.x * 4) {
        if (i + 3 < n && ((((uintptr_t)dst | (uintptr_t)src) & 15) == 0)) {
            float4 a = *reinterpret_cast<float4*>(dst + i);
            const float4 b = *reinterpret_cast<const float4*>(src + i);
            a.x += b.x; a.y += b.y; a.z += b.z; a.w += b.w;
            *reinterpret_cast<float4*>(dst + i) = a;
        } else
            for (long long j = i; j < n && j < i + 4; ++j) dst[j] += src[j];
    }
}

void add_inplace(float* dst, const float* src, long long n, cudaStream_t st)
{
    if (n <= 0) return;
    ProfScope ps(F_BLEND, 0, 12.0 * n, st);
    add_inplace_kernel<<<(unsigned)std::min<long long>((n / 4 + 255) / 256 + 1, 148 * 16), 256, 0, st>>>(dst, src, n);
    count_launch();
}

void loss_to_slot(const double* loss, float* slot, cudaStream_t st) { loss_to_slot_kernel<<<1, 1, 0, st>>>(loss, slot); count_launch(); }
void slot_to_loss(const float* slot, double* loss, int world, cudaStream_t st) { slot_to_loss_kernel<<<1, 1, 0, st>>>(slot, loss, 1.f / world); count_launch(); }

}  // namespace k
}  // namespace ab
