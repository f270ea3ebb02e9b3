// Detection-level blob filter of annonet_infer() (annonet_infer.cpp:187-239): connected blobs of equal non-zero labels
// (dlib::label_connected_blobs with zero_pixels_are_background, neighbors_8, connected_if_equal) survive only if they
// contain a "seed" -- a pixel whose winning class beats the clean class 0 by more than its detection level.
//
// Connected components by union-find on the pixel grid (label equivalence, after Komura / Playne): every labelled pixel
// starts as its own root, is united with its W / NW / N / NE neighbours of the same label through atomicMin on the parent
// array (roots only ever decrease, so the structure stays a forest), then every pixel is pointed at its root.  Blob
// NUMBERS differ from dlib's scan order, which the filter never looks at: a blob is kept or zeroed as a whole.
// HBM-bound integer work: 2 B label + 4 B parent per pixel and pass, coalesced rows, grids sized to the image.
#include "kcommon.cuh"

#include <algorithm>
#include <stdexcept>

namespace ab {
namespace k {
namespace {

__device__ __forceinline__ int find_root(const int* parent, int i)
{
    int p = __ldcg(parent + i);
    while (p != i) { i = p; p = __ldcg(parent + i); }
    return i;
}

__device__ __forceinline__ void unite(int* parent, int a, int b)
{
    while (true) {
        a = find_root(parent, a);
        b = find_root(parent, b);
        if (a == b) return;
        if (a < b) { const int t = a; a = b; b = t; }   // a > b: hang the larger root under the smaller
        const int old = atomicMin(parent + a, b);
        if (old == a) return;                            // a was still a root: linked
        a = old;                                         // somebody re-parented a meanwhile: retry from there
    }
}

__global__ void __launch_bounds__(256) blobs_init_kernel(const uint16_t* __restrict__ labels, int* __restrict__ parent, uint8_t* __restrict__ keep, long long n)
{
    const long long i = (long long)blockIdx.x * blockDim.x + threadIdx.x;
    if (i >= n) return;
    parent[i] = labels[i] != 0 ? (int)i : -1;
    keep[i] = 0;
}

__global__ void __launch_bounds__(256) blobs_merge_kernel(const uint16_t* __restrict__ labels, int* __restrict__ parent, int h, int w)
{
    const long long i = (long long)blockIdx.x * blockDim.x + threadIdx.x;
    if (i >= (long long)h * w) return;
    const uint16_t l = labels[i];
    if (l == 0) return;
    const int y = (int)(i / w), x = (int)(i - (long long)y * w);
    if (x > 0 && labels[i - 1] == l) unite(parent, (int)i, (int)i - 1);
    if (y > 0) {
        const long long up = i - w;
        if (labels[up] == l) unite(parent, (int)i, (int)up);
        if (x > 0 && labels[up - 1] == l) unite(parent, (int)i, (int)up - 1);
        if (x + 1 < w && labels[up + 1] == l) unite(parent, (int)i, (int)up + 1);
    }
}

// transposed: the reference stores a seed found at pixel (r, c) as dlib::point(x = r, y = c) and looks the blob up at
// connected_blobs(point.y(), point.x()) = pixel (row c, col r) (annonet_infer.cpp:210,221).  Mirrored here: pixel (y, x) is
// marked by the seed flag of pixel (x, y); lookups that would fall outside the image (the reference reads out of bounds
// there) mark nothing.
__global__ void __launch_bounds__(256) blobs_flatten_kernel(int* __restrict__ parent, const uint8_t* __restrict__ seeds, uint8_t* __restrict__ keep, int h,
                                                            int w, int transposed)
{
    const long long i = (long long)blockIdx.x * blockDim.x + threadIdx.x;
    if (i >= (long long)h * w) return;
    if (__ldcg(parent + i) < 0) return;
    const int r = find_root(parent, (int)i);
    parent[i] = r;          // (writes only shorten paths: concurrent readers still reach the same root)
    bool s;
    if (transposed) {
        const int y = (int)(i / w), x = (int)(i - (long long)y * w);
        s = x < h && y < w && seeds[(long long)x * w + y];
    } else s = seeds[i];
    if (s) keep[r] = 1;
}

__global__ void __launch_bounds__(256) blobs_apply_kernel(uint16_t* __restrict__ labels, const int* __restrict__ parent, const uint8_t* __restrict__ keep, long long n)
{
    const long long i = (long long)blockIdx.x * blockDim.x + threadIdx.x;
    if (i >= n) return;
    const int p = parent[i];
    if (p < 0) return;
    if (!keep[find_root(parent, p)]) labels[i] = 0;
}

__global__ void __launch_bounds__(256) blobs_roots_kernel(int* __restrict__ parent, long long n)
{
    const long long i = (long long)blockIdx.x * blockDim.x + threadIdx.x;
    if (i >= n) return;
    if (__ldcg(parent + i) < 0) return;
    parent[i] = find_root(parent, (int)i);
}

// confusion_matrix_per_pixel (annonet_infer_main.cpp:482-489): counts[truth][predicted] over the labelled pixels
__global__ void __launch_bounds__(256) confusion_kernel(const uint16_t* __restrict__ truth, const uint16_t* __restrict__ pred, long long n, int K,
                                                        unsigned long long* __restrict__ counts)
{
    extern __shared__ unsigned int sh[];   // K * K
    for (int i = threadIdx.x; i < K * K; i += blockDim.x) sh[i] = 0;
    __syncthreads();
    for (long long i = (long long)blockIdx.x * blockDim.x + threadIdx.x; i < n; i += (long long)gridDim.x * blockDim.x) {
        const unsigned t = truth[i], p = pred[i];
        if (t < (unsigned)K && p < (unsigned)K) atomicAdd(&sh[t * K + p], 1u);
    }
    __syncthreads();
    for (int i = threadIdx.x; i < K * K; i += blockDim.x)
        if (sh[i]) atomicAdd(&counts[i], (unsigned long long)sh[i]);
}

// resize_label_image (annonet.cpp:132-141: dlib::resize_image with interpolate_nearest_neighbor -- source pixel
// (round(c * (iw-1)/(ow-1)), round(r * (ih-1)/(oh-1))), dlib rounds a vector<double> to a point with floor(x + 0.5)) fused with
// index_label_image_to_rgba_label_image (annonet_infer_main.cpp:74-86): labels -> the classes' rgba colours
__global__ void __launch_bounds__(256) render_labels_kernel(const uint16_t* __restrict__ labels, int ih, int iw, int oh, int ow, double ys, double xs,
                                                            const uchar4* __restrict__ palette, int K, uchar4* __restrict__ out)
{
    const long long i = (long long)blockIdx.x * blockDim.x + threadIdx.x;
    if (i >= (long long)oh * ow) return;
    const int r = (int)(i / ow), c = (int)(i - (long long)r * ow);
    const long long sy = (long long)floor(r * ys + 0.5), sx = (long long)floor(c * xs + 0.5);
    uchar4 v = make_uchar4(0, 0, 0, 0);   // outside the source: black_background (never for a resize); unknown label: transparent
    if (sy >= 0 && sy < ih && sx >= 0 && sx < iw) {
        const unsigned l = labels[sy * iw + sx];
        if (l < (unsigned)K) v = palette[l];
    }
    out[i] = v;
}

}  // namespace

// dlib::label_connected_blobs(labels, zero_pixels_are_background(), neighbors_8(), connected_if_equal()): roots[i] = -1 for
// background pixels, else the smallest pixel index of pixel i's blob (a canonical blob id; dlib numbers blobs in scan order,
// which nothing downstream depends on)
void label_blobs(const uint16_t* labels, int h, int w, int* roots, uint8_t* scratch_keep, cudaStream_t st)
{
    const long long n = (long long)h * w;
    if (n >= (1ll << 31)) throw std::invalid_argument("label_blobs: image too large (pixel indices must fit 31 bits)");
    ProfScope ps(F_BLEND, 0, 20.0 * n, st);
    const unsigned grid = (unsigned)((n + 255) / 256);
    blobs_init_kernel<<<grid, 256, 0, st>>>(labels, roots, scratch_keep, n);
    blobs_merge_kernel<<<grid, 256, 0, st>>>(labels, roots, h, w);
    blobs_roots_kernel<<<grid, 256, 0, st>>>(roots, n);
    count_launch(3);
}

void confusion_per_pixel(const uint16_t* truth, const uint16_t* pred, long long n, int K, unsigned long long* counts, cudaStream_t st)
{
    if (K < 1 || K > 64) throw std::invalid_argument("confusion_per_pixel: 1..64 classes");
    ProfScope ps(F_ARGMAX, 0, 4.0 * n, st);
    cudaMemsetAsync(counts, 0, sizeof(unsigned long long) * K * K, st);
    const unsigned grid = (unsigned)std::min<long long>((n + 255) / 256, (long long)num_sms() * 8);
    confusion_kernel<<<grid, 256, sizeof(unsigned) * K * K, st>>>(truth, pred, n, K, counts);
    count_launch();
}

void render_labels(const uint16_t* labels, int ih, int iw, int oh, int ow, const uint8_t* palette_rgba, int K, uint8_t* out_rgba, cudaStream_t st)
{
    const long long n = (long long)oh * ow;
    ProfScope ps(F_ARGMAX, 0, 6.0 * n, st);
    const double xs = (iw - 1) / (double)std::max(ow - 1, 1), ys = (ih - 1) / (double)std::max(oh - 1, 1);
    render_labels_kernel<<<(unsigned)((n + 255) / 256), 256, 0, st>>>(labels, ih, iw, oh, ow, ys, xs, (const uchar4*)palette_rgba, K, (uchar4*)out_rgba);
    count_launch();
}

void filter_detections(uint16_t* labels, const uint8_t* seeds, int h, int w, int* parent, uint8_t* keep, cudaStream_t st, int transposed_seeds)
{
    const long long n = (long long)h * w;
    if (n >= (1ll << 31)) throw std::invalid_argument("filter_detections: image too large (pixel indices must fit 31 bits)");
    ProfScope ps(F_BLEND, 0, 30.0 * n, st);
    const unsigned grid = (unsigned)((n + 255) / 256);
    blobs_init_kernel<<<grid, 256, 0, st>>>(labels, parent, keep, n);
    blobs_merge_kernel<<<grid, 256, 0, st>>>(labels, parent, h, w);
    blobs_flatten_kernel<<<grid, 256, 0, st>>>(parent, seeds, keep, h, w, transposed_seeds);
    blobs_apply_kernel<<<grid, 256, 0, st>>>(labels, parent, keep, n);
    count_launch(4);
}

}  // namespace k
}  // namespace ab
