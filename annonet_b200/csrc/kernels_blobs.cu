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

}  // namespace

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
