// The allocation-free variant of confusion_per_pixel (kernels_blobs.cu; confusion_matrix_per_pixel of
// annonet_infer_main.cpp:482-489) behind ab_confusion_per_pixel_noalloc (include/annonet_b200.h).
//
// No counterpart in the reference: added for callers that cannot call cudaMalloc
// (code under CUDA graph capture, tight serving loops, harnesses that forbid it).  The engine's own path accumulates into a
// counter buffer it cudaMalloc's on first use (Engine::evaluate_labels, e_counts_); here that scratch is the file-scope
// __device__ array below, so the call allocates NOTHING: no cudaMalloc / cudaFree / cudaMallocAsync, no events, no library
// routine that allocates.  Every GPU that loads the library holds the array (32 KB) until the process exits.
//
// Limits and sharing:
//   * class_count <= kMaxClasses (64): the array holds 64 x 64 counters; larger requests are rejected with AB_ERR_ARGUMENT.
//   * ONE array per GPU, shared by every call and keeping the last call's contents: each call clears the counters it uses
//     first, on the caller's stream.  Calls that overlap on the GPU (different streams, or different host threads without
//     ordering between their streams) RACE on it -- serialise them.
//   * The first call on a device looks the array's device address up (cudaGetSymbolAddress; with lazy module loading that is
//     also what loads the module) -- make that first call outside a stream capture.
#include "engine.h"   // ab::cuda_error (mapped to AB_ERR_CUDA by c_api.cu)

#include <cuda_runtime.h>

#include <algorithm>
#include <cstdint>
#include <mutex>
#include <stdexcept>
#include <string>

namespace ab {
namespace noalloc {
namespace {

constexpr int kMaxClasses = 64;
constexpr int kMaxDevices = 64;

// The scratch.  In HOST code this name is only a stand-in for the symbol, not a device address: a kernel handed `g_counts`
// by the host would fault (or scribble on host memory where the GPU can reach it).  Host code below therefore asks the
// runtime for the real address once per device and passes THAT to the kernels.
static __device__ __align__(256) unsigned long long g_counts[kMaxClasses * kMaxClasses];

__global__ void __launch_bounds__(256) clear_kernel(unsigned long long* __restrict__ counts, int kk)
{
    for (int i = blockIdx.x * blockDim.x + threadIdx.x; i < kk; i += gridDim.x * blockDim.x) counts[i] = 0ull;
}

// same arithmetic as confusion_kernel (kernels_blobs.cu): per-CTA histogram in shared memory, one 64-bit atomic per non-zero
// cell and CTA into the scratch counters
__global__ void __launch_bounds__(256) confusion_noalloc_kernel(const uint16_t* __restrict__ truth, const uint16_t* __restrict__ pred, long long n,
                                                                int K, unsigned long long* __restrict__ counts)
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

__global__ void __launch_bounds__(256) publish_kernel(const unsigned long long* __restrict__ counts, unsigned long long* __restrict__ out, int kk)
{
    for (int i = blockIdx.x * blockDim.x + threadIdx.x; i < kk; i += gridDim.x * blockDim.x) out[i] = counts[i];
}

struct PerDevice {
    unsigned long long* counts = nullptr;   // device address of g_counts ON THIS device
    int sms = 0;
};

void check(cudaError_t e, const char* what)
{
    if (e != cudaSuccess) throw cuda_error(std::string(what) + ": " + cudaGetErrorString(e));
}

// set-up, once per device, with that device current
const PerDevice& per_device()
{
    static std::mutex mu;
    static PerDevice table[kMaxDevices];
    int dev = -1;
    check(cudaGetDevice(&dev), "cudaGetDevice");
    if (dev < 0 || dev >= kMaxDevices) throw cuda_error("ab_confusion_per_pixel_noalloc: unexpected device ordinal " + std::to_string(dev));
    std::lock_guard<std::mutex> lock(mu);
    PerDevice& d = table[dev];
    if (!d.counts) {
        void* p = nullptr;
        check(cudaGetSymbolAddress(&p, g_counts), "cudaGetSymbolAddress(noalloc scratch)");
        if (!p) throw cuda_error("cudaGetSymbolAddress returned a null address for the noalloc scratch");
        int sms = 0;
        check(cudaDeviceGetAttribute(&sms, cudaDevAttrMultiProcessorCount, dev), "cudaDeviceGetAttribute");
        d.sms = std::max(sms, 1);
        d.counts = static_cast<unsigned long long*>(p);
    }
    return d;
}

}  // namespace

int max_classes() { return kMaxClasses; }

void confusion_per_pixel(const uint16_t* truth_dev, const uint16_t* pred_dev, long long n, int K, unsigned long long* out_dev, cudaStream_t st)
{
    if (K < 1 || K > kMaxClasses)
        throw std::invalid_argument("ab_confusion_per_pixel_noalloc: class_count " + std::to_string(K) + " is outside 1.." +
                                    std::to_string(kMaxClasses) + " (the static scratch holds " + std::to_string(kMaxClasses) + " x " +
                                    std::to_string(kMaxClasses) + " counters)");
    if (n < 0) throw std::invalid_argument("ab_confusion_per_pixel_noalloc: negative pixel count");
    if (!out_dev || (n > 0 && (!truth_dev || !pred_dev))) throw std::invalid_argument("ab_confusion_per_pixel_noalloc: null pointer");
    const PerDevice& d = per_device();
    const int kk = K * K;
    clear_kernel<<<(kk + 255) / 256, 256, 0, st>>>(d.counts, kk);   // the array keeps the previous call's counts
    if (n > 0) {
        const unsigned grid = (unsigned)std::min<long long>((n + 255) / 256, (long long)d.sms * 8);
        confusion_noalloc_kernel<<<grid, 256, sizeof(unsigned) * kk, st>>>(truth_dev, pred_dev, n, K, d.counts);
    }
    publish_kernel<<<(kk + 255) / 256, 256, 0, st>>>(d.counts, out_dev, kk);
    check(cudaGetLastError(), "ab_confusion_per_pixel_noalloc launch");
}

}  // namespace noalloc
}  // namespace ab
