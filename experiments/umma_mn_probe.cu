// Hardware probe (not part of the product): MN-major tcgen05 shared-memory descriptors for the filter-gradient GEMM
// (K = positions, so both operands are [position][channel] tiles written by TMA with 128-byte swizzle).
//   T1: M = 128 as two 64-channel TMA boxes, LBO = box stride, SBO = 1024 (8 rows), start address offset by r rows
//   T2: M = 128 as the SAME 64-channel box at two row offsets (two taps fused into one MMA): LBO = delta_rows * 128
//   T3: B from a tensor with only 16 channels (TMA box of 64 channels, out-of-bounds columns zero-filled)
// variant 1 swaps the roles of LBO and SBO.   nvcc -gencode arch=compute_100a,code=sm_100a -o umma_mn_probe umma_mn_probe.cu
#include <cuda.h>
#include <cuda_bf16.h>
#include <cuda_runtime.h>
#include <cstdio>
#include <cstdlib>
#include <vector>

#define CK(x) do { cudaError_t e_ = (x); if (e_ != cudaSuccess) { printf("CUDA error %s at %d\n", cudaGetErrorString(e_), __LINE__); exit(1); } } while (0)

typedef CUresult (*EncodeTiled)(CUtensorMap*, CUtensorMapDataType, cuuint32_t, void*, const cuuint64_t*, const cuuint64_t*,
                                const cuuint32_t*, const cuuint32_t*, CUtensorMapInterleave, CUtensorMapSwizzle,
                                CUtensorMapL2promotion, CUtensorMapFloatOOBfill);

__device__ __forceinline__ uint32_t smem_u32(const void* p) { return (uint32_t)__cvta_generic_to_shared(p); }
__device__ __forceinline__ void mbar_init(uint64_t* b, int n) { asm volatile("mbarrier.init.shared::cta.b64 [%0], %1;" ::"r"(smem_u32(b)), "r"(n)); }
__device__ __forceinline__ void mbar_expect(uint64_t* b, uint32_t tx) { asm volatile("mbarrier.arrive.expect_tx.shared::cta.b64 _, [%0], %1;" ::"r"(smem_u32(b)), "r"(tx) : "memory"); }
__device__ __forceinline__ void mbar_wait(uint64_t* b, uint32_t parity)
{
    asm volatile("{\n.reg .pred p;\nW: mbarrier.try_wait.parity.shared::cta.b64 p, [%0], %1;\n@p bra D;\nbra W;\nD:\n}" ::"r"(smem_u32(b)), "r"(parity) : "memory");
}
__device__ __forceinline__ void tma_2d(void* dst, const CUtensorMap* m, int c0, int c1, uint64_t* bar)
{
    asm volatile("cp.async.bulk.tensor.2d.shared::cluster.global.mbarrier::complete_tx::bytes [%0], [%1, {%2, %3}], [%4];"
                 ::"r"(smem_u32(dst)), "l"(m), "r"(c0), "r"(c1), "r"(smem_u32(bar)) : "memory");
}


constexpr int ROWS = 96;      // positions staged per operand
constexpr int KSTEPS = 3;     // 48 positions of K

// test: 1 / 2 / 3 (see the header); r: row offset of the start address; variant: 0 (LBO = M-block stride, SBO = 8-row stride) or 1 (swapped)
__global__ void __launch_bounds__(128) probe(const __grid_constant__ CUtensorMap tx, const __grid_constant__ CUtensorMap ty,
                                             float* out, int test, int r, int delta, int variant)
{
    extern __shared__ __align__(1024) uint8_t smem[];
    uint8_t* sx = (uint8_t*)(((uintptr_t)smem + 1023) & ~(uintptr_t)1023);   // X: two boxes of ROWS x 64 channels
    uint8_t* sy = sx + 2 * ROWS * 128;
    __shared__ uint64_t bar_full, bar_mma;
    __shared__ uint32_t tmem_base;
    const int warp = threadIdx.x / 32;
    if (threadIdx.x == 0) { mbar_init(&bar_full, 1); mbar_init(&bar_mma, 1); asm volatile("fence.mbarrier_init.release.cluster;"); }
    if (warp == 0) {
        asm volatile("tcgen05.alloc.cta_group::1.sync.aligned.shared::cta.b32 [%0], 64;" ::"r"(smem_u32(&tmem_base)));
        asm volatile("tcgen05.relinquish_alloc_permit.cta_group::1.sync.aligned;");
    }
    asm volatile("tcgen05.fence::before_thread_sync;");
    __syncthreads();
    asm volatile("tcgen05.fence::after_thread_sync;");
    const uint32_t tm = tmem_base;
    if (threadIdx.x == 0) {
        mbar_expect(&bar_full, (uint32_t)(3 * ROWS * 128));
        tma_2d(sx, &tx, 0, 0, &bar_full);
        tma_2d(sx + ROWS * 128, &tx, 64, 0, &bar_full);
        tma_2d(sy, &ty, 0, 0, &bar_full);
        mbar_wait(&bar_full, 0);
        asm volatile("tcgen05.fence::after_thread_sync;");
        const uint32_t N = 64;
        const uint32_t idesc = (1u << 4) | (1u << 7) | (1u << 10) | (1u << 15) | (1u << 16) | ((N >> 3) << 17) | ((128u >> 4) << 24);
        const uint32_t mblock = test == 2 ? (uint32_t)delta * 128u : (uint32_t)ROWS * 128u;   // byte distance of the second 64 rows of M
        const uint32_t lbo = variant == 0 ? mblock : 1024u, sbo = variant == 0 ? 1024u : mblock;
        const uint64_t hi = ((uint64_t)(sbo >> 4) << 32) | (1ull << 46) | (2ull << 61);
        const uint32_t a_addr = smem_u32(sx) + (uint32_t)r * 128, b_addr = smem_u32(sy) + (uint32_t)(r & 7) * 0;   // B starts at row 0
        for (int k = 0; k < KSTEPS; ++k) {
            const uint64_t da = hi | ((uint64_t)(lbo >> 4) << 16) | (uint64_t)(((a_addr + k * 2048) >> 4) & 0x3fff);
            const uint64_t db = hi | ((uint64_t)(lbo >> 4) << 16) | (uint64_t)(((b_addr + k * 2048) >> 4) & 0x3fff);
            asm volatile("{\n.reg .pred p;\nsetp.ne.b32 p, %4, 0;\ntcgen05.mma.cta_group::1.kind::f16 [%0], %1, %2, %3, p;\n}"
                         ::"r"(tm), "l"(da), "l"(db), "r"(idesc), "r"(k > 0 ? 1 : 0) : "memory");
        }
        asm volatile("tcgen05.commit.cta_group::1.mbarrier::arrive::one.shared::cluster.b64 [%0];" ::"r"(smem_u32(&bar_mma)) : "memory");
    }
    mbar_wait(&bar_mma, 0);
    asm volatile("tcgen05.fence::after_thread_sync;");
    for (int j0 = 0; j0 < 64; j0 += 16) {
        uint32_t v[16];
        const uint32_t taddr = tm + ((uint32_t)(warp * 32) << 16) + j0;
        asm volatile("tcgen05.ld.sync.aligned.32x32b.x16.b32 {%0,%1,%2,%3,%4,%5,%6,%7,%8,%9,%10,%11,%12,%13,%14,%15}, [%16];"
                     : "=r"(v[0]), "=r"(v[1]), "=r"(v[2]), "=r"(v[3]), "=r"(v[4]), "=r"(v[5]), "=r"(v[6]), "=r"(v[7]), "=r"(v[8]), "=r"(v[9]),
                       "=r"(v[10]), "=r"(v[11]), "=r"(v[12]), "=r"(v[13]), "=r"(v[14]), "=r"(v[15]) : "r"(taddr));
        asm volatile("tcgen05.wait::ld.sync.aligned;");
        for (int j = 0; j < 16; ++j) out[threadIdx.x * 64 + j0 + j] = __uint_as_float(v[j]);
    }
    asm volatile("tcgen05.fence::before_thread_sync;");
    __syncthreads();
    if (warp == 0) asm volatile("tcgen05.dealloc.cta_group::1.sync.aligned.b32 %0, 64;" ::"r"(tm));
}

int main()
{
    EncodeTiled enc = nullptr;
    cudaDriverEntryPointQueryResult q;
    CK(cudaGetDriverEntryPoint("cuTensorMapEncodeTiled", (void**)&enc, cudaEnableDefault, &q));
    const int CX = 128;
    for (int test = 1; test <= 3; ++test) {
        const int CY = test == 3 ? 16 : 64;
        std::vector<__nv_bfloat16> hx(ROWS * CX), hy(ROWS * CY);
        std::vector<float> fx(ROWS * CX), fy(ROWS * CY);
        srand(7 + test);
        for (size_t i = 0; i < hx.size(); ++i) { fx[i] = (float)(rand() % 17 - 8); hx[i] = __float2bfloat16(fx[i]); }
        for (size_t i = 0; i < hy.size(); ++i) { fy[i] = (float)(rand() % 9 - 4); hy[i] = __float2bfloat16(fy[i]); }
        __nv_bfloat16 *dx, *dy; float* dout;
        CK(cudaMalloc(&dx, hx.size() * 2)); CK(cudaMalloc(&dy, hy.size() * 2)); CK(cudaMalloc(&dout, 128 * 64 * 4));
        CK(cudaMemcpy(dx, hx.data(), hx.size() * 2, cudaMemcpyHostToDevice));
        CK(cudaMemcpy(dy, hy.data(), hy.size() * 2, cudaMemcpyHostToDevice));
        CUtensorMap tx, ty;
        cuuint64_t dimsX[2] = {(cuuint64_t)CX, ROWS}, dimsY[2] = {(cuuint64_t)CY, ROWS}, sX[1] = {(cuuint64_t)CX * 2}, sY[1] = {(cuuint64_t)CY * 2};
        cuuint32_t box[2] = {64, ROWS}, es[2] = {1, 1};
        CUresult r1 = enc(&tx, CU_TENSOR_MAP_DATA_TYPE_BFLOAT16, 2, dx, dimsX, sX, box, es, CU_TENSOR_MAP_INTERLEAVE_NONE, CU_TENSOR_MAP_SWIZZLE_128B,
                          CU_TENSOR_MAP_L2_PROMOTION_L2_128B, CU_TENSOR_MAP_FLOAT_OOB_FILL_NONE);
        CUresult r2 = enc(&ty, CU_TENSOR_MAP_DATA_TYPE_BFLOAT16, 2, dy, dimsY, sY, box, es, CU_TENSOR_MAP_INTERLEAVE_NONE, CU_TENSOR_MAP_SWIZZLE_128B,
                          CU_TENSOR_MAP_L2_PROMOTION_L2_128B, CU_TENSOR_MAP_FLOAT_OOB_FILL_NONE);
        if (r1 != CUDA_SUCCESS || r2 != CUDA_SUCCESS) { printf("test %d: tensor map encode failed %d %d\n", test, (int)r1, (int)r2); continue; }
        CK(cudaFuncSetAttribute(probe, cudaFuncAttributeMaxDynamicSharedMemorySize, 64 * 1024));
        for (int variant = 0; variant < 2; ++variant)
            for (int delta : {1, 3, 34}) {
                if (test != 2 && delta != 1) continue;
                printf("test %d variant %d delta %2d, r = 0..17: ", test, variant, delta);
                for (int r = 0; r <= 17; ++r) {
                    CK(cudaMemset(dout, 0, 128 * 64 * 4));
                    probe<<<1, 128, 64 * 1024>>>(tx, ty, dout, test, r, delta, variant);
                    CK(cudaDeviceSynchronize());
                    std::vector<float> ho(128 * 64);
                    CK(cudaMemcpy(ho.data(), dout, ho.size() * 4, cudaMemcpyDeviceToHost));
                    double worst = 0;
                    for (int m = 0; m < 128; ++m)
                        for (int n = 0; n < 64; ++n) {
                            double ref = 0;
                            for (int k = 0; k < 16 * KSTEPS; ++k) {
                                // T1/T3: row m of M = channel m of X at position r + k;  T2: rows 64.. = channel m - 64 at position r + delta + k
                                const int ch = test == 2 ? (m & 63) : m;
                                const int pos = r + k + (test == 2 && m >= 64 ? delta : 0);
                                const double a = fx[pos * CX + ch];
                                const double b = n < CY ? fy[k * CY + n] : 0.0;
                                ref += a * b;
                            }
                            worst = fmax(worst, fabs(ref - ho[m * 64 + n]));
                        }
                    printf("%s", worst == 0 ? "." : "X");
                }
                printf("\n");
            }
        cudaFree(dx); cudaFree(dy); cudaFree(dout);
    }
    return 0;
}
