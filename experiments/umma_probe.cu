// Hardware probe (not part of the product): does a tcgen05 K-major shared-memory descriptor whose start
// address is offset by r rows (r not a multiple of the 8-row swizzle atom) address a TMA-written swizzled
// tile correctly, and with which base_offset value?  Decides how the conv kernel shares one halo segment
// between the nine taps.     nvcc -gencode arch=compute_100a,code=sm_100a -o umma_probe umma_probe.cu
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

__global__ void __launch_bounds__(128) probe(const __grid_constant__ CUtensorMap ta, const __grid_constant__ CUtensorMap tb,
                                             float* out, int KC, int rows_a, int r, int base_off_mode, int layout_type, int N)
{
    extern __shared__ __align__(1024) uint8_t smem[];
    uint8_t* sa = (uint8_t*)(((uintptr_t)smem + 1023) & ~(uintptr_t)1023);
    const int pitch = KC * 2;
    uint8_t* sb = sa + ((rows_a * pitch + 1023) / 1024) * 1024;
    __shared__ uint64_t bar_full, bar_mma;
    __shared__ uint32_t tmem_base;
    const int warp = threadIdx.x / 32;
    if (threadIdx.x == 0) { mbar_init(&bar_full, 1); mbar_init(&bar_mma, 1); asm volatile("fence.mbarrier_init.release.cluster;"); }
    if (warp == 0) {
        asm volatile("tcgen05.alloc.cta_group::1.sync.aligned.shared::cta.b32 [%0], 32;" ::"r"(smem_u32(&tmem_base)));
        asm volatile("tcgen05.relinquish_alloc_permit.cta_group::1.sync.aligned;");
    }
    asm volatile("tcgen05.fence::before_thread_sync;");
    __syncthreads();
    asm volatile("tcgen05.fence::after_thread_sync;");
    const uint32_t tm = tmem_base;
    if (threadIdx.x == 0) {
        mbar_expect(&bar_full, (uint32_t)(rows_a * pitch + N * pitch));
        tma_2d(sa, &ta, 0, 0, &bar_full);
        tma_2d(sb, &tb, 0, 0, &bar_full);
        mbar_wait(&bar_full, 0);
        asm volatile("tcgen05.fence::after_thread_sync;");
        const uint32_t idesc = (1u << 4) | (1u << 7) | (1u << 10) | ((uint32_t)(N >> 3) << 17) | ((128u >> 4) << 24);
        const uint32_t a_addr = smem_u32(sa) + (uint32_t)r * pitch, b_addr = smem_u32(sb);
        const uint64_t sbo = (uint64_t)(8 * pitch) >> 4;
        uint64_t hi = (sbo << 32) | (1ull << 46) | ((uint64_t)layout_type << 61) | (1ull << 16);
        for (int k = 0; k < KC / 16; ++k) {
            uint64_t bo = 0;
            if (base_off_mode == 1) bo = (uint64_t)(((a_addr + k * 32) >> 7) & 7);
            if (base_off_mode == 2) bo = (uint64_t)(r & 7);
            const uint64_t da = hi | (bo << 49) | (uint64_t)(((a_addr + k * 32) >> 4) & 0x3fff);
            const uint64_t db = hi | (uint64_t)(((b_addr + k * 32) >> 4) & 0x3fff);
            asm volatile("{\n.reg .pred p;\nsetp.ne.b32 p, %4, 0;\ntcgen05.mma.cta_group::1.kind::f16 [%0], %1, %2, %3, p;\n}"
                         ::"r"(tm), "l"(da), "l"(db), "r"(idesc), "r"(k > 0 ? 1 : 0) : "memory");
        }
        asm volatile("tcgen05.commit.cta_group::1.mbarrier::arrive::one.shared::cluster.b64 [%0];" ::"r"(smem_u32(&bar_mma)) : "memory");
    }
    mbar_wait(&bar_mma, 0);
    asm volatile("tcgen05.fence::after_thread_sync;");
    uint32_t v[16];
    const uint32_t taddr = tm + ((uint32_t)(warp * 32) << 16);
    asm volatile("tcgen05.ld.sync.aligned.32x32b.x16.b32 {%0,%1,%2,%3,%4,%5,%6,%7,%8,%9,%10,%11,%12,%13,%14,%15}, [%16];"
                 : "=r"(v[0]), "=r"(v[1]), "=r"(v[2]), "=r"(v[3]), "=r"(v[4]), "=r"(v[5]), "=r"(v[6]), "=r"(v[7]), "=r"(v[8]), "=r"(v[9]),
                   "=r"(v[10]), "=r"(v[11]), "=r"(v[12]), "=r"(v[13]), "=r"(v[14]), "=r"(v[15]) : "r"(taddr));
    asm volatile("tcgen05.wait::ld.sync.aligned;");
    for (int j = 0; j < 16; ++j) out[threadIdx.x * 16 + j] = __uint_as_float(v[j]);
    asm volatile("tcgen05.fence::before_thread_sync;");
    __syncthreads();
    if (warp == 0) asm volatile("tcgen05.dealloc.cta_group::1.sync.aligned.b32 %0, 32;" ::"r"(tm));
}

int main()
{
    EncodeTiled enc = nullptr;
    cudaDriverEntryPointQueryResult q;
    CK(cudaGetDriverEntryPoint("cuTensorMapEncodeTiled", (void**)&enc, cudaEnableDefault, &q));
    const int ROWS = 160, N = 16;
    for (int KC : {64, 32, 16}) {
        std::vector<__nv_bfloat16> ha(ROWS * KC), hb(N * KC);
        std::vector<float> fa(ROWS * KC), fb(N * KC);
        srand(7);
        for (size_t i = 0; i < ha.size(); ++i) { fa[i] = (float)(rand() % 17 - 8); ha[i] = __float2bfloat16(fa[i]); }
        for (size_t i = 0; i < hb.size(); ++i) { fb[i] = (float)(rand() % 9 - 4); hb[i] = __float2bfloat16(fb[i]); }
        __nv_bfloat16 *da, *db; float* dout;
        CK(cudaMalloc(&da, ha.size() * 2)); CK(cudaMalloc(&db, hb.size() * 2)); CK(cudaMalloc(&dout, 128 * 16 * 4));
        CK(cudaMemcpy(da, ha.data(), ha.size() * 2, cudaMemcpyHostToDevice));
        CK(cudaMemcpy(db, hb.data(), hb.size() * 2, cudaMemcpyHostToDevice));
        const CUtensorMapSwizzle sw = KC == 64 ? CU_TENSOR_MAP_SWIZZLE_128B : KC == 32 ? CU_TENSOR_MAP_SWIZZLE_64B : CU_TENSOR_MAP_SWIZZLE_32B;
        const int layout_type = KC == 64 ? 2 : KC == 32 ? 4 : 6;
        CUtensorMap ta, tb;
        cuuint64_t dimsA[2] = {(cuuint64_t)KC, ROWS}, dimsB[2] = {(cuuint64_t)KC, N}, strides[1] = {(cuuint64_t)KC * 2};
        cuuint32_t boxA[2] = {(cuuint32_t)KC, ROWS}, boxB[2] = {(cuuint32_t)KC, N}, es[2] = {1, 1};
        CUresult r1 = enc(&ta, CU_TENSOR_MAP_DATA_TYPE_BFLOAT16, 2, da, dimsA, strides, boxA, es, CU_TENSOR_MAP_INTERLEAVE_NONE, sw,
                          CU_TENSOR_MAP_L2_PROMOTION_L2_128B, CU_TENSOR_MAP_FLOAT_OOB_FILL_NONE);
        CUresult r2 = enc(&tb, CU_TENSOR_MAP_DATA_TYPE_BFLOAT16, 2, db, dimsB, strides, boxB, es, CU_TENSOR_MAP_INTERLEAVE_NONE, sw,
                          CU_TENSOR_MAP_L2_PROMOTION_L2_128B, CU_TENSOR_MAP_FLOAT_OOB_FILL_NONE);
        if (r1 != CUDA_SUCCESS || r2 != CUDA_SUCCESS) { printf("tensor map encode failed %d %d\n", (int)r1, (int)r2); return 1; }
        CK(cudaFuncSetAttribute(probe, cudaFuncAttributeMaxDynamicSharedMemorySize, 64 * 1024));
        for (int mode = 0; mode < 3; ++mode) {
            printf("KC=%d base_offset mode %d (0: zero, 1: (addr>>7)&7, 2: r&7): ", KC, mode);
            for (int r = 0; r <= 17; ++r) {
                CK(cudaMemset(dout, 0, 128 * 16 * 4));
                probe<<<1, 128, 64 * 1024>>>(ta, tb, dout, KC, ROWS, r, mode, layout_type, N);
                CK(cudaDeviceSynchronize());
                std::vector<float> ho(128 * 16);
                CK(cudaMemcpy(ho.data(), dout, ho.size() * 4, cudaMemcpyDeviceToHost));
                double worst = 0;
                for (int i = 0; i < 128; ++i)
                    for (int n = 0; n < N; ++n) {
                        double ref = 0;
                        for (int k = 0; k < KC; ++k) ref += (double)fa[(r + i) * KC + k] * fb[n * KC + k];
                        worst = fmax(worst, fabs(ref - ho[i * 16 + n]));
                    }
                printf("%s", worst == 0 ? "." : "X");
            }
            printf("\n");
        }
        cudaFree(da); cudaFree(db); cudaFree(dout);
    }
    return 0;
}
