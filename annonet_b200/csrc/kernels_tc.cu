// bf16 tcgen05 / TMEM / TMA implicit-GEMM convolutions (AB_PRECISION_BF16).  Placeholder translation
// unit: the kernels land here next.
#include "kcommon.cuh"
namespace ab { namespace k { } }
