// Per-kernel-family device timing for bench.py's roofline line: when enabled, every launcher brackets
// its launches with a CUDA event pair on the launching stream and records the ALGORITHMIC work of that
// launch (flops and minimum bytes: one read of each input, one write of the output -- SURVEY §8d).
// Off by default; the off path is one relaxed atomic load per launcher.
#pragma once
#include <cuda_runtime.h>

namespace ab {
namespace k {

enum Family {
    F_CONV3X3 = 0,   // con_/cont_ 3x3 s1 forward and backward-data (implicit GEMM)
    F_CONV1X1,       // con_ 1x1 (+ fused max-pool) forward / backward-data
    F_UP,            // cont_ 3x3 s2 forward / backward-data
    F_WGRAD,         // filter gradients (split-K implicit GEMM + split reduction)
    F_BIAS_GRAD,     // per-channel sums
    F_BN,            // bn_ statistics / apply / backward, fused with mish
    F_POOL,          // max_pool forward / backward
    F_LOSS,          // softmax + weighted log loss + gradient
    F_ARGMAX,
    F_BLEND,
    F_INPUT,         // to_tensor / tile extraction
    F_OPTIMIZER,
    F_PACK,          // layout / precision conversion of the h16 tensor-core path
    F_CONV_TC,       // tcgen05 implicit-GEMM convolutions (forward, backward-data)
    F_WGRAD_TC,      // tcgen05 filter gradients
    F_EW_BF16,       // h16 NHWC elementwise / reduction kernels (bn, mish, pool, bias gradient)
    F_COUNT
};

struct FamilyStat {
    const char* name;
    unsigned long long launches;  // event-bracketed launcher calls
    double ms, flops, bytes;
};

void prof_enable(bool on);               // (re)starts collection; drops earlier samples
bool prof_enabled();
void prof_collect(FamilyStat out[F_COUNT]);  // synchronizes the device, sums the samples

struct ProfScope {
    ProfScope(int family, double flops, double bytes, cudaStream_t st);
    ~ProfScope();
    int slot;
    cudaStream_t st;
};

}  // namespace k
}  // namespace ab
