// Mini-batch assembly on the device (annonet_train_main.cpp:110-232 randomly_crop_image + annonet_train.h:20-83 set_weights):
// at B200 step rates (a few thousand crops per second and GPU) the reference's loader threads -- extract_image_chip, outpaint,
// set_weights, flips, brightness, noise, colour offset per crop on the CPU, then a host-to-device copy -- cannot keep one GPU
// busy.  Here the full images and label maps stay resident in HBM; the host only draws the random decisions (which image,
// where, which flips, which factors: a few dozen bytes per crop) and three kernels write the crops and their weighted
// labels straight into the training step's input buffers.
//   crop_hist_kernel    : per crop, the label histogram of the crop window + what set_weights' growing vector needs
//   crop_weights_kernel : per crop, the class weights (double arithmetic as annonet_train.h:52-72)
//   crop_write_kernel   : per output pixel, gather (edge-replicated outside the image = outpaint), flips, brightness,
//                         noise, colour offset -> u8 rgb;  label + weight -> weighted_label
// HBM-bound byte/integer work: coalesced rows, one thread per output pixel.
#include "kcommon.cuh"

#include <stdexcept>

namespace ab {
namespace k {
namespace {

constexpr int kMaxLabels = 256;   // class ids a crop may hold on the device path (the reference's vector grows without bound)

__global__ void __launch_bounds__(256) crop_hist_kernel(const CropJob* __restrict__ jobs, int dim, unsigned int* __restrict__ hist,
                                                        unsigned int* __restrict__ first_idx, unsigned int* __restrict__ overflow)
{
    __shared__ unsigned int sh[kMaxLabels];
    __shared__ unsigned int s_first;
    const CropJob j = jobs[blockIdx.y];
    for (int i = threadIdx.x; i < kMaxLabels; i += blockDim.x) sh[i] = 0;
    if (threadIdx.x == 0) s_first = 0xffffffffu;
    __syncthreads();
    const int n = dim * dim;
    for (int i = blockIdx.x * blockDim.x + threadIdx.x; i < n; i += gridDim.x * blockDim.x) {
        const int y = i / dim, x = i - y * dim;
        const int sy = j.top + y, sx = j.left + x;
        if (sy < 0 || sy >= j.h || sx < 0 || sx >= j.w) continue;   // outside the image: label_to_ignore (set_to_unknown_outside)
        const unsigned l = j.labels[(long long)sy * j.w + sx];
        if (l == 65535u) continue;
        if (l >= (unsigned)kMaxLabels) { atomicAdd(overflow, 1u); continue; }
        atomicAdd(&sh[l], 1u);
        atomicMin(&s_first, (unsigned)i);
    }
    __syncthreads();
    for (int i = threadIdx.x; i < kMaxLabels; i += blockDim.x)
        if (sh[i]) atomicAdd(&hist[blockIdx.y * kMaxLabels + i], sh[i]);
    if (threadIdx.x == 0 && s_first != 0xffffffffu) atomicMin(&first_idx[blockIdx.y], s_first);
}

// annonet_train.h:36-72.  label_counts is a vector that grows to index*2+16 whenever a label does not fit (:31-35), and
// average_count divides by its SIZE: the size depends on the order in which labels appear.  With the first labelled pixel's
// label f the size is 2f+16; a later label >= that size grows it again -- replayed here in scan order only in that rare case.
__global__ void crop_weights_kernel(const CropJob* __restrict__ jobs, int dim, const unsigned int* __restrict__ hist,
                                    const unsigned int* __restrict__ first_idx, double class_weight, double image_weight,
                                    float* __restrict__ weights)
{
    const int c = blockIdx.x;
    const CropJob j = jobs[c];
    const unsigned int* h = hist + c * kMaxLabels;
    float* wout = weights + c * kMaxLabels;
    if (threadIdx.x != 0) return;
    unsigned long long total = 0;
    int maxl = -1;
    for (int l = 0; l < kMaxLabels; ++l) { total += h[l]; if (h[l]) maxl = l; }
    for (int l = 0; l < kMaxLabels; ++l) wout[l] = 0.f;
    if (total == 0) return;
    const unsigned fi = first_idx[c];
    const int fy = (int)(fi / dim), fx = (int)(fi - fy * dim);
    long long size = 2ll * j.labels[(long long)(j.top + fy) * j.w + (j.left + fx)] + 16;
    if (maxl >= size) {   // replay the growth in scan order
        for (int i = (int)fi; i < dim * dim; ++i) {
            const int y = i / dim, x = i - y * dim, sy = j.top + y, sx = j.left + x;
            if (sy < 0 || sy >= j.h || sx < 0 || sx >= j.w) continue;
            const unsigned l = j.labels[(long long)sy * j.w + sx];
            if (l != 65535u && (long long)l >= size) size = 2ll * l + 16;
        }
    }
    const double average_count = (double)total / (double)size;
    double total_unnormalized = 0.0;
    double lw[kMaxLabels];
    for (int l = 0; l <= maxl; ++l) {
        lw[l] = 0.0;
        if (h[l]) { lw[l] = pow(average_count / (double)h[l], class_weight); total_unnormalized += (double)h[l] * lw[l]; }
    }
    const double target = (double)total * pow((double)dim * dim / (double)total, image_weight);
    for (int l = 0; l <= maxl; ++l) wout[l] = (float)(lw[l] * (target / total_unnormalized));
}

// counter-based random integer in [-level, level] (the reference draws dlib::rand::get_integer_in_range per channel in scan
// order from the loader thread's generator; any reproducible stream does -- this one is a function of (seed, pixel, channel))
__device__ __forceinline__ int noise_of(unsigned seed, unsigned idx, unsigned ch, int level)
{
    unsigned x = seed ^ (idx * 0x9E3779B9u) ^ (ch * 0x85EBCA6Bu);
    x ^= x >> 16; x *= 0x7FEB352Du; x ^= x >> 15; x *= 0x846CA68Bu; x ^= x >> 16;
    return (int)(x % (unsigned)(2 * level + 1)) - level;
}

__global__ void __launch_bounds__(256) crop_write_kernel(const CropJob* __restrict__ jobs, int dim, const float* __restrict__ weights,
                                                         uint8_t* __restrict__ out_img, uint2* __restrict__ out_lab)
{
    const CropJob j = jobs[blockIdx.y];
    const int n = dim * dim;
    for (int i = blockIdx.x * blockDim.x + threadIdx.x; i < n; i += gridDim.x * blockDim.x) {
        const int oy = i / dim, ox = i - oy * dim;
        // flips act on the finished crop (annonet_train_main.cpp:187-196): output pixel (oy, ox) shows crop pixel (y, x)
        const int y = j.flip_ud ? dim - 1 - oy : oy, x = j.flip_lr ? dim - 1 - ox : ox;
        const int sy = j.top + y, sx = j.left + x;
        const bool inside = sy >= 0 && sy < j.h && sx >= 0 && sx < j.w;
        // outpaint (annonet.h:74-120) replicates the nearest pixel of the valid rectangle: clamped coordinates
        const int cy = min(max(sy, 0), j.h - 1), cx = min(max(sx, 0), j.w - 1);
        const uint8_t* p = j.image + ((long long)cy * j.w + cx) * 3;
        int v[3] = {p[0], p[1], p[2]};
#pragma unroll
        for (int ch = 0; ch < 3; ++ch) {
            if (j.brightness != 1.0) v[ch] = (int)floor(fmin(fmax((double)v[ch] * j.brightness, 0.0), 255.0) + 0.5);   // :204-206
            if (j.noise_level > 0) v[ch] = min(max(v[ch] + noise_of(j.seed, (unsigned)i, (unsigned)ch, j.noise_level), 0), 255);   // :73-101
            v[ch] = min(max(v[ch] + j.color_offset[ch], 0), 255);   // apply_random_color_offset, random_color_transform.h:98-150
        }
        uint8_t* o = out_img + ((long long)blockIdx.y * n + i) * 3;
        o[0] = (uint8_t)v[0]; o[1] = (uint8_t)v[1]; o[2] = (uint8_t)v[2];
        const unsigned l = inside ? j.labels[(long long)sy * j.w + sx] : 65535u;
        const float wgt = (l < (unsigned)kMaxLabels) ? weights[blockIdx.y * kMaxLabels + l] : 0.f;
        out_lab[(long long)blockIdx.y * n + i] = make_uint2(l, __float_as_uint(wgt));   // weighted_label {u16 label, pad, f32 weight}
    }
}

}  // namespace

size_t crop_scratch_bytes(int n) { return (size_t)n * (kMaxLabels * 4 + 4 + kMaxLabels * 4) + 16; }

void build_crops(const CropJob* jobs_dev, int n, int dim, double class_weight, double image_weight, void* scratch, uint8_t* out_img, void* out_lab,
                 unsigned int* overflow_dev, cudaStream_t st)
{
    ProfScope ps(F_INPUT, 0, 14.0 * n * dim * dim, st);
    unsigned int* hist = (unsigned int*)scratch;
    unsigned int* first = hist + (size_t)n * kMaxLabels;
    float* weights = (float*)(first + n);
    cudaMemsetAsync(hist, 0, (size_t)n * kMaxLabels * 4, st);
    cudaMemsetAsync(first, 0xff, (size_t)n * 4, st);
    cudaMemsetAsync(overflow_dev, 0, 4, st);
    const unsigned gx = (unsigned)std::min((dim * dim + 255) / 256, 64);
    crop_hist_kernel<<<dim3(gx, n), 256, 0, st>>>(jobs_dev, dim, hist, first, overflow_dev);
    crop_weights_kernel<<<n, 32, 0, st>>>(jobs_dev, dim, hist, first, class_weight, image_weight, weights);
    crop_write_kernel<<<dim3(gx, n), 256, 0, st>>>(jobs_dev, dim, weights, out_img, (uint2*)out_lab);
    count_launch(3);
}

}  // namespace k
}  // namespace ab
