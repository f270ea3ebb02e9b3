// Launchers of the fp32 (NCHW) CUDA kernels.  All tensors are channel slices of level buffers:
// element (n, c, y, x) of a slice lives at  base + n*sstride + (c0 + c)*H*W + y*W + x.
// Every launcher enqueues on `st` and returns; errors surface through cudaGetLastError in the caller.
#pragma once
#include <cuda_runtime.h>
#include <cstdint>

namespace ab {
namespace k {

struct Slice {
    float* p = nullptr;
    long long sstride = 0;  // floats between samples
    int c0 = 0, C = 0;
};

// weight addressing of the implicit-GEMM kernels: w[a*sa + b*sb + tap], tap = ky*k + kx
// (con_: dlib [Cout][Cin][k][k]; cont_: [Cin][Cout][k][k], layers.h:182-193,512)
struct WView {
    const float* w = nullptr;
    int s_out = 0, s_in = 0;  // stride of the kernel's output-channel / input-channel index
    int flip = 0;             // use tap (k*k-1 - tap): correlation with the filter rotated by 180 degrees
};

enum Epilogue { EPI_NONE = 0, EPI_MISH_AFFINE = 1, EPI_AFFINE_MISH = 2 };

struct ConvArgs {
    Slice in, out;
    int N = 0, H = 0, W = 0;      // output == input spatial size (stride 1, "same" padding)
    WView wv;
    const float* bias = nullptr;  // [out.C] or null
    int accumulate = 0;           // out += result (gradient accumulation, dnn/core.h:958-974 semantics)
    int epi = EPI_NONE;
    const float* scale = nullptr; // affine_ gamma/beta per output channel (inference, layers.h:2199)
    const float* shift = nullptr;
    int pool = 0;                 // 1x1 only: 3x3/2 max-pool of `in` computed on load; in has size Hin x Win
    int Hin = 0, Win = 0;
};

// out[n,co,y,x] = bias[co] + sum_{ci,ky,kx} w(co,ci,ky,kx) * in[n,ci,y+ky-1,x+kx-1]   (zero padding)
// con_ forward (cpu_dlib.cpp:2250-2296), con_ backward-data with flip (:2298-2325), cont_<3,3,1,1,1,1>
// forward (layers.h:531) and its backward-data (layers.h:542).
void conv3x3(const ConvArgs& a, cudaStream_t st);
// 1x1 convolution, optionally with the transition_down max-pool fused into the load
void conv1x1(const ConvArgs& a, cudaStream_t st);

// cont_<3,3,2,2,0,0> ("up", NetStructure.h:83): out (2h+1 x 2w+1) = bias + scatter(in (h x w) * w[ci][co][ky][kx])
void up_forward(const Slice& in, const Slice& out, int N, int h, int w, const float* wt, const float* bias, cudaStream_t st);
// its backward-data (layers.h:542): gin[n,ci,i,j] += sum_{co,ky,kx} gout[n,co,2i+ky,2j+kx] * w[ci][co][ky][kx]
void up_backward_data(const Slice& gout, const Slice& gin, int N, int h, int w, const float* wt, cudaStream_t st);

// G[a][b][ky][kx] = sum_{n,y,x} A[n,a,y,x] * B[n,b,S*y+ky-P,S*x+kx-P]   (A is Ha x Wa, B is Hb x Wb)
// con_ filters gradient: A = dY, B = X (cpu_dlib.cpp:2327-2355); cont_: A = X, B = dY (layers.h:547).
// g is fully overwritten; `scratch` holds the per-split partial sums (>= wgrad_scratch_floats()).
void wgrad3x3(const Slice& A, const Slice& B, int N, int Ha, int Wa, int Hb, int Wb, int S, int P,
              float* g, float* scratch, cudaStream_t st);
// G[co][ci] = sum A[n,co,p] * B[n,ci,p]
void wgrad1x1(const Slice& A, const Slice& B, int N, int H, int W, float* g, float* scratch, cudaStream_t st);
size_t wgrad_scratch_floats(int max_a, int max_b);
// g[c] = sum_{n,y,x} A[n,c,y,x]   (assign_conv_bias_gradient, cpu_dlib.cpp:470-504)
void channel_sum(const Slice& A, int N, int H, int W, float* g, float* scratch, cudaStream_t st);

// max_pool_<3,3,2,2,0,0> (cpu_dlib.cpp:1888-1943) and its gradient (first maximum in row-major window
// order receives it, :1971-2017); gin += .
void maxpool_forward(const Slice& in, const Slice& out, int N, int Hin, int Win, cudaStream_t st);
void maxpool_backward(const Slice& in, const Slice& gout, const Slice& gin, int N, int Hin, int Win, cudaStream_t st);

// ---- batch norm (CONV_MODE) + mish ------------------------------------------------------------
// sums[c] = sum x, sums[C+c] = sum x^2 over (n,y,x) of x = mish_first ? mish(z) : z   (double accumulation)
void bn_stats(const Slice& z, int N, int H, int W, int mish_first, double* sums, cudaStream_t st);
// mean/invstd (biased variance, eps), running stats with dlib's averaging factor (layers.h:1392-1397,
// cpu_dlib.cpp:1127-1182); stats = [mean(C) | invstd(C)]
// num_updates: device counter of this bn_ layer; decay = 1 - n/(n+1), then n = min(n+1, window)
void bn_finalize(const double* sums, int C, long long m, float eps, unsigned long long* num_updates, unsigned long long window,
                 float* running_mean, float* running_var, float* stats, cudaStream_t st);
// y = mish_first ? gamma*(mish(z)-mean)*invstd+beta : mish(gamma*(z-mean)*invstd+beta)   [mish optional]
void bn_apply(const Slice& z, const Slice& y, int N, int H, int W, int mish_first, int has_mish,
              const float* gamma, const float* beta, const float* stats, cudaStream_t st);
// backward, pass 1: red[c] = sum dv, red[C+c] = sum dv*xhat  where dv = gradient at the bn output
// (= gy for conv->mish->bn, gy*mish'(u) for conv->bn->mish) -- cpu_dlib.cpp:1228-1250
void bn_backward_reduce(const Slice& z, const Slice& gy, int N, int H, int W, int mish_first, int has_mish,
                        const float* gamma, const float* beta, const float* stats, double* red, cudaStream_t st);
// pass 2: dz = gradient at the conv output (dense [N][C][H][W]); also writes dgamma/dbeta (cpu_dlib.cpp:1252-1288)
void bn_backward_apply(const Slice& z, const Slice& gy, float* dz, int N, int H, int W, int mish_first, int has_mish,
                       const float* gamma, const float* beta, const float* stats, const double* red,
                       float* dgamma, float* dbeta, cudaStream_t st);

// ---- input / loss / labels --------------------------------------------------------------------
// input_rgb_image::to_tensor (input.h:68-117): u8 HWC -> f32 planar (p - avg)/256
void to_tensor(const uint8_t* img, float* x, int N, int H, int W, float ar, float ag, float ab, cudaStream_t st);
// tile extraction with clamp-to-edge (extract_image_chip + outpaint, annonet_infer.cpp:68-75, annonet.h:74-120) fused with to_tensor
void extract_tile(const uint8_t* img, int IH, int IW, int left, int top, float* x, int TH, int TW,
                  float ar, float ag, float ab, cudaStream_t st);
// loss_multiclass_log_per_pixel_weighted_ (loss.h:3121-3192): grad = w*s*(softmax - onehot); loss_sum += w*s*-log(max(p_y,1e-10))
void loss_forward_backward(const float* logits, const void* wlabels, float* grad, int N, int K, int H, int W,
                           double* loss_sum, cudaStream_t st);
// to_label (loss.h:2921-2969): argmax_k float(logit + gain), lowest index wins ties, all-NaN -> 65535
// det_dev / seeds (optional): seeds[i] = 1 where label > 0 and plane[label] - plane[0] > det[label] - det[0]
// (the detection seeds of annonet_infer.cpp:204-212)
void argmax_labels(const float* planes, long long plane_stride, int K, long long npix, const double* gains_dev,
                   uint16_t* labels, cudaStream_t st, const double* det_dev = nullptr, uint8_t* seeds = nullptr);
// annonet_infer.cpp:214-239: blobs of equal non-zero labels (8-neighbourhood) without a seed are set to 0.
// parent: h*w ints, keep: h*w bytes of scratch.  transposed_seeds: mirror the reference's (r, c) -> point(x = r, y = c) ->
// blobs(y, x) seed lookup (annonet_infer.cpp:210,221) instead of the evidently intended same-pixel lookup.
// ---- mini-batch assembly on the device (kernels_crop.cu; annonet_train_main.cpp:110-232, annonet_train.h:20-83) -----------
struct CropJob {
    const uint8_t* image;       // full image, u8 HWC rgb, resident on the device
    const uint16_t* labels;     // its label map, u16 (65535 = ignore)
    int h, w;                   // size of the full image
    int left, top;              // crop window = [left, left + dim) x [top, top + dim); may leave the image (outpaint / ignore)
    int flip_lr, flip_ud;
    double brightness;          // multiplicative brightness change (1.0: none)
    int noise_level;            // uniform integer noise in [-level, level] per channel (0: none)
    unsigned seed;              // of this crop's noise
    int color_offset[3];        // apply_random_color_offset's per-channel offsets (0: none)
};
size_t crop_scratch_bytes(int n);
// out_img: u8 [n][dim][dim][3]; out_lab: weighted_label<uint16_t> [n][dim][dim] (8 bytes each); overflow_dev: counts labels >= 256
void build_crops(const CropJob* jobs_dev, int n, int dim, double class_weight, double image_weight, void* scratch, uint8_t* out_img, void* out_lab,
                 unsigned int* overflow_dev, cudaStream_t st);

// ---- evaluation and write-back of annonet_infer_main.cpp (confusion matrices :91-272,482-491; result images :408-418) --------
// connected blobs of equal non-zero labels (8-neighbourhood): roots[i] = -1 (background) or the blob's smallest pixel index
void label_blobs(const uint16_t* labels, int h, int w, int* roots, uint8_t* scratch_bytes, cudaStream_t st);
// counts[t * K + p] = number of pixels with truth t < K and prediction p < K (ignored truth 65535 drops out)
void confusion_per_pixel(const uint16_t* truth, const uint16_t* pred, long long n, int K, unsigned long long* counts, cudaStream_t st);
// nearest-neighbour resize of a label map to oh x ow (dlib's rounding) + class colours -> rgba8
void render_labels(const uint16_t* labels, int ih, int iw, int oh, int ow, const uint8_t* palette_rgba, int K, uint8_t* out_rgba, cudaStream_t st);
void filter_detections(uint16_t* labels, const uint8_t* seeds, int h, int w, int* parent, uint8_t* keep, cudaStream_t st, int transposed_seeds = 1);
// blend (annonet_infer.cpp:102-164): canvas[k, Y, X] (+)= t * logits[k, y, x]
struct BlendArgs {
    const float* logits; int K, th, tw;        // tile logits [K][th][tw]
    float* canvas; int H, W;                    // [K][H][W]
    int left, top;                              // canvas position of tile pixel (0,0)
    int full[4], unique[4];                     // inclusive l,t,r,b
};
void blend_tile(const BlendArgs& a, cudaStream_t st);
// dst += src (a neighbour rank's seam strip into the blend canvas)
void add_inplace(float* dst, const float* src, long long n, cudaStream_t st);

// ---- optimizer --------------------------------------------------------------------------------
// sgd::operator() (solvers.h:155-185): v = m*v - (wd*wdm)*(lr*lrm)*w - (lr*lrm)*g ; w += v
void sgd_step(float* w, const float* g, float* v, const float* lr_mul, const float* wd_mul, long long n,
              float lr, float wd, float momentum, float grad_scale, cudaStream_t st);
// adam (solvers.h:353-372, cpu_dlib.cpp:704-743)
// t_dev: device-side step counter, advanced by the launch itself (the step stays a static CUDA graph)
void adam_step(float* w, const float* g, float* m, float* v, const float* lr_mul, const float* wd_mul, long long n,
               float* t_dev, float lr, float wd, float beta1, float beta2, float grad_scale, cudaStream_t st);
// data parallel: the step loss rides in one extra float behind the gradient arena through the last all-reduce
// (dnn_trainer reports the mean of the device losses, trainer.h:1206-1253)
void loss_to_slot(const double* loss, float* slot, cudaStream_t st);
void slot_to_loss(const float* slot, double* loss, int world, cudaStream_t st);
void fill_zero(void* p, size_t bytes, cudaStream_t st);

// concat_prev_'s bilinear resize of the up-sampled tensor to the skip's size (layers.h:2494-2521, cpu_dlib.cpp:1735-1762):
// out[n, c, y, x] = ((wtl*tl + wtr*tr) + wbl*bl) + wbr*br with the source indices / fractions of plan.h resize_axis.
// tab (device): ints top[oh] bottom[oh] left[ow] right[ow], then floats tb[oh] itb[oh] lr[ow] ilr[ow]
struct ResizeTab { const int* top; const int* bottom; const int* left; const int* right; const float* tb; const float* itb; const float* lr; const float* ilr; };
void resize_bilinear_nchw(const Slice& in, int ih, int iw, const Slice& out, int N, int oh, int ow, const ResizeTab& t, cudaStream_t st);

}  // namespace k
}  // namespace ab
