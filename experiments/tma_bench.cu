// Hardware probe (not part of the product): TMA tiled-load throughput per SM as a function of the box row width
// (kc channels of a [rows][64-channel] bf16 tensor) -- cycles per box row and bytes per cycle.
#include <cuda.h>
#include <cuda_runtime.h>
#include <cstdio>
#include <cstdint>
#define CK(x) do { cudaError_t e_ = (x); if (e_ != cudaSuccess) { printf("CUDA error %s at %d\n", cudaGetErrorString(e_), __LINE__); return 1; } } while (0)
typedef CUresult (*EncodeTiled)(CUtensorMap*, CUtensorMapDataType, cuuint32_t, void*, const cuuint64_t*, const cuuint64_t*,
                                const cuuint32_t*, const cuuint32_t*, CUtensorMapInterleave, CUtensorMapSwizzle,
                                CUtensorMapL2promotion, CUtensorMapFloatOOBfill);
__device__ __forceinline__ uint32_t smem_u32(const void* p) { return (uint32_t)__cvta_generic_to_shared(p); }
__device__ __forceinline__ void mbar_wait(uint64_t* b, uint32_t parity)
{
    asm volatile("{\n.reg .pred p;\nW: mbarrier.try_wait.parity.shared::cta.b64 p, [%0], %1;\n@p bra D;\nbra W;\nD:\n}" ::"r"(smem_u32(b)), "r"(parity) : "memory");
}
__global__ void __launch_bounds__(32) bench(const __grid_constant__ CUtensorMap tm, int kc, int box_rows, int boxes, int rounds, int total_rows, long long* cycles)
{
    extern __shared__ __align__(1024) uint8_t smem[];
    __shared__ uint64_t bar;
    const uint32_t base = (smem_u32(smem) + 1023u) & ~1023u;
    if (threadIdx.x == 0) {
        asm volatile("mbarrier.init.shared::cta.b64 [%0], 1;" ::"r"(smem_u32(&bar)));
        asm volatile("fence.mbarrier_init.release.cluster;");
        const uint32_t box_bytes = (uint32_t)(box_rows * kc * 2);
        const long long t0 = clock64();
        for (int r = 0; r < rounds; ++r) {
            asm volatile("mbarrier.arrive.expect_tx.shared::cta.b64 _, [%0], %1;" ::"r"(smem_u32(&bar)), "r"(box_bytes * boxes) : "memory");
            for (int b = 0; b < boxes; ++b) {
                const int row = (int)(((long long)(blockIdx.x * rounds + r) * boxes + b) * box_rows % (total_rows - box_rows));
                asm volatile("cp.async.bulk.tensor.2d.shared::cluster.global.mbarrier::complete_tx::bytes [%0], [%1, {%2, %3}], [%4];"
                             ::"r"(base + b * box_bytes), "l"(&tm), "r"(0), "r"(row), "r"(smem_u32(&bar)) : "memory");
            }
            mbar_wait(&bar, r & 1);
        }
        cycles[blockIdx.x] = clock64() - t0;
    }
}
int main()
{
    EncodeTiled enc = nullptr; cudaDriverEntryPointQueryResult q;
    CK(cudaGetDriverEntryPoint("cuTensorMapEncodeTiled", (void**)&enc, cudaEnableDefault, &q));
    const int total_rows = 1 << 21, ctot = 64;   // 256 MB tensor
    void* d; CK(cudaMalloc(&d, (size_t)total_rows * ctot * 2)); CK(cudaMemset(d, 0, (size_t)total_rows * ctot * 2));
    long long* dc; CK(cudaMalloc(&dc, 148 * 8));
    CK(cudaFuncSetAttribute(bench, cudaFuncAttributeMaxDynamicSharedMemorySize, 200 * 1024));
    for (int grid : {1, 148})
        for (int kc : {16, 32, 64})
            for (int promo = 0; promo < 4; ++promo) { const int box_rows = 256;
                CUtensorMap tm;
                cuuint64_t dims[2] = {(cuuint64_t)ctot, (cuuint64_t)total_rows}, strides[1] = {(cuuint64_t)ctot * 2};
                cuuint32_t box[2] = {(cuuint32_t)kc, (cuuint32_t)box_rows}, es[2] = {1, 1};
                const CUtensorMapSwizzle sw = kc == 64 ? CU_TENSOR_MAP_SWIZZLE_128B : kc == 32 ? CU_TENSOR_MAP_SWIZZLE_64B : CU_TENSOR_MAP_SWIZZLE_32B;
                if (enc(&tm, CU_TENSOR_MAP_DATA_TYPE_BFLOAT16, 2, d, dims, strides, box, es, CU_TENSOR_MAP_INTERLEAVE_NONE, sw,
                        (CUtensorMapL2promotion)promo, CU_TENSOR_MAP_FLOAT_OOB_FILL_NONE) != CUDA_SUCCESS) { printf("encode failed\n"); return 1; }
                const int boxes = 160 * 1024 / (box_rows * kc * 2) > 10 ? 10 : 160 * 1024 / (box_rows * kc * 2), rounds = 50;
                bench<<<grid, 32, 200 * 1024>>>(tm, kc, box_rows, boxes, rounds, total_rows, dc);
                CK(cudaDeviceSynchronize());
                long long h[148]; CK(cudaMemcpy(h, dc, grid * 8, cudaMemcpyDeviceToHost));
                long long mx = 0; for (int i = 0; i < grid; ++i) mx = h[i] > mx ? h[i] : mx;
                const double rows = (double)rounds * boxes * box_rows;
                printf("promo %d grid %3d kc %2d (%3d B rows) box_rows %3d x %2d boxes: %.2f cycles/row, %.1f B/cycle/SM, round latency %.0f cycles\n", promo, grid, kc, kc * 2,
                       box_rows, boxes, mx / rows, rows * kc * 2 / mx, (double)mx / rounds);
            }
    return 0;
}
