// Shared device helpers of the fp32 kernels.
#pragma once
#include "kernels.h"
#include "prof.h"

#include <atomic>
#include <cstdlib>

namespace ab {
namespace k {

extern std::atomic<unsigned long long> g_launches;  // kernels of this library launched so far
inline void count_launch(int n = 1) { g_launches.fetch_add((unsigned long long)n, std::memory_order_relaxed); }

// SM count of the current device (B200: 148), queried once: sizes the persistent grids, the tiling cost model and the
// split-K scratch, so a MIG slice or another Blackwell SKU gets grids that fit it
inline int num_sms()
{
    static const int n = [] {
        int dev = 0, v = 0;
        if (cudaGetDevice(&dev) != cudaSuccess || cudaDeviceGetAttribute(&v, cudaDevAttrMultiProcessorCount, dev) != cudaSuccess || v <= 0) {
            cudaGetLastError();
            return 148;
        }
        return v;
    }();
    return n;
}

// ---- programmatic dependent launch (PDL) --------------------------------------------------------------------------------
// The training step is ~370 dependent launches of 5..300 us; between two of them the GPU drains, launches and fills again.
// A kernel launched with launch_pdl() may be made resident while its predecessor in the stream is still running; it must call
// pdl_enter() BEFORE its first global-memory access: that blocks until the predecessor has completed and its writes are
// visible (griddepcontrol.wait), then lets its own successor start becoming resident (griddepcontrol.launch_dependents -- after
// the wait, so that completion of a kernel always implies completion of everything before it in the stream).  Launched
// without the attribute (AB_NO_PDL=1, or the first kernel after a memset / event) both instructions are no-ops.
__device__ __forceinline__ void pdl_enter()
{
    asm volatile("griddepcontrol.wait;" ::: "memory");
    asm volatile("griddepcontrol.launch_dependents;" ::: "memory");
}

inline bool pdl_enabled()
{
    static const bool on = !(getenv("AB_NO_PDL") && atoi(getenv("AB_NO_PDL")) != 0);
    return on;
}

// kernel<<<grid, block, smem, st>>>(args...) with the programmatic-stream-serialization attribute.  ONLY for kernels that call
// pdl_enter() first thing.
template <typename... KArgs, typename... Args>
inline void launch_pdl(void (*kernel)(KArgs...), dim3 grid, dim3 block, size_t smem, cudaStream_t st, Args&&... args)
{
    cudaLaunchConfig_t cfg = {};
    cfg.gridDim = grid; cfg.blockDim = block; cfg.dynamicSmemBytes = smem; cfg.stream = st;
    cudaLaunchAttribute attr[1];
    attr[0].id = cudaLaunchAttributeProgrammaticStreamSerialization;
    attr[0].val.programmaticStreamSerializationAllowed = 1;
    cfg.attrs = attr;
    cfg.numAttrs = pdl_enabled() ? 1 : 0;
    cudaLaunchKernelEx(&cfg, kernel, static_cast<KArgs>(args)...);
}

// mish_: y = x - 2x/(e^2 + 2e + 2), e = exp(x)   (cpu_dlib.cpp:1510-1523; fp32 exp overflow -> y = x)
__device__ __forceinline__ float mish_f(float x)
{
    const float e = expf(x);
    const float delta = 2.f * e + e * e + 2.f;
    return x - 2.f * x / delta;
}

// d mish / dx with dlib's clamps (cpu_dlib.cpp:1531-1543)
__device__ __forceinline__ float mish_grad_f(float x)
{
    if (x >= 8.f) return 1.f;
    if (x <= -8.f) return 0.f;
    const float e = expf(x);
    const float delta = 2.f * e + e * e + 2.f;
    const float omega = 4.f * (x + 1.f) + 4.f * e * e + e * e * e + e * (4.f * x + 6.f);
    return e * omega / (delta * delta);
}

__device__ __forceinline__ float warp_sum(float v)
{
#pragma unroll
    for (int o = 16; o > 0; o >>= 1) v += __shfl_xor_sync(0xffffffffu, v, o);
    return v;
}
__device__ __forceinline__ double warp_sum(double v)
{
#pragma unroll
    for (int o = 16; o > 0; o >>= 1) v += __shfl_xor_sync(0xffffffffu, v, o);
    return v;
}

__device__ __forceinline__ const float* plane(const Slice& s, int n, int c, long long hw)
{
    return s.p + (long long)n * s.sstride + (long long)(s.c0 + c) * hw;
}
__device__ __forceinline__ float* plane_w(const Slice& s, int n, int c, long long hw)
{
    return s.p + (long long)n * s.sstride + (long long)(s.c0 + c) * hw;
}

}  // namespace k
}  // namespace ab
