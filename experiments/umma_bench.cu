// Hardware probe (not part of the product): cycles per tcgen05.mma (M = 128, K = 16, bf16, SS mode) as a function of N,
// of the shared-memory swizzle mode of the K-major operands and of whether consecutive MMAs hit the same accumulator.
#include <cuda_runtime.h>
#include <cstdio>
#include <cstdint>
#define CK(x) do { cudaError_t e_ = (x); if (e_ != cudaSuccess) { printf("CUDA error %s at %d\n", cudaGetErrorString(e_), __LINE__); return 1; } } while (0)
__device__ __forceinline__ uint32_t smem_u32(const void* p) { return (uint32_t)__cvta_generic_to_shared(p); }
__device__ __forceinline__ void mbar_wait(uint64_t* b, uint32_t parity)
{
    asm volatile("{\n.reg .pred p;\nW: mbarrier.try_wait.parity.shared::cta.b64 p, [%0], %1;\n@p bra D;\nbra W;\nD:\n}" ::"r"(smem_u32(b)), "r"(parity) : "memory");
}
__global__ void __launch_bounds__(128) bench(int N, int pitch, int reps, int rotate, int row_off, long long* cycles)
{
    extern __shared__ __align__(1024) uint8_t smem[];
    __shared__ uint64_t bar;
    __shared__ uint32_t tmem_slot;
    uint8_t* base = (uint8_t*)(((uintptr_t)smem + 1023) & ~(uintptr_t)1023);
    for (int i = threadIdx.x; i < 48 * 1024 / 4; i += 128) ((uint32_t*)base)[i] = 0x3c003c00u;  // bf16 pairs, small values
    if (threadIdx.x == 0) { asm volatile("mbarrier.init.shared::cta.b64 [%0], 1;" ::"r"(smem_u32(&bar))); asm volatile("fence.mbarrier_init.release.cluster;"); }
    if (threadIdx.x < 32) {
        asm volatile("tcgen05.alloc.cta_group::1.sync.aligned.shared::cta.b32 [%0], 512;" ::"r"(smem_u32(&tmem_slot)));
        asm volatile("tcgen05.relinquish_alloc_permit.cta_group::1.sync.aligned;");
    }
    asm volatile("fence.proxy.async.shared::cta;" ::: "memory");
    asm volatile("tcgen05.fence::before_thread_sync;");
    __syncthreads();
    asm volatile("tcgen05.fence::after_thread_sync;");
    const uint32_t tm = tmem_slot;
    uint32_t el; asm volatile("{\n.reg .pred p;\nelect.sync _|p, 0xffffffff;\nselp.u32 %0, 1, 0, p;\n}" : "=r"(el));
    if (threadIdx.x < 32 && el) {
        const uint32_t idesc = (1u << 4) | (1u << 7) | (1u << 10) | ((uint32_t)(N >> 3) << 17) | ((128u >> 4) << 24);
        const uint32_t layout = pitch == 128 ? 2 : pitch == 64 ? 4 : 6;
        const uint32_t hi = (uint32_t)((8 * pitch) >> 4) | (1u << 14) | (layout << 29);
        const uint32_t a0 = ((smem_u32(base) + row_off * pitch) >> 4) | (1u << 16), b0 = ((smem_u32(base) + 32 * 1024) >> 4) | (1u << 16);
        const uint32_t dstep = rotate ? (uint32_t)N : 0u;   // rotate over two accumulators
        const long long t0 = clock64();
#define MMA(dd, aa, acc) asm volatile("{\n.reg .b64 da, db;\n.reg .pred p;\nsetp.ne.b32 p, %5, 0;\nmov.b64 da, {%1, %3};\nmov.b64 db, {%2, %3};\n" \
                         "tcgen05.mma.cta_group::1.kind::f16 [%0], da, db, %4, p;\n}" ::"r"(dd), "r"(aa), "r"(b0), "r"(hi), "r"(idesc), "r"(acc) : "memory")
        for (int r = 0; r < reps; r += 8) {
            const uint32_t acc = r > 0;
            MMA(tm, a0, acc); MMA(tm + dstep, a0 + 64, acc); MMA(tm, a0 + 128, 1); MMA(tm + dstep, a0 + 192, 1);
            MMA(tm, a0 + 256, 1); MMA(tm + dstep, a0 + 320, 1); MMA(tm, a0 + 384, 1); MMA(tm + dstep, a0 + 448, 1);
        }
        asm volatile("tcgen05.commit.cta_group::1.mbarrier::arrive::one.shared::cluster.b64 [%0];" ::"r"(smem_u32(&bar)) : "memory");
        mbar_wait(&bar, 0);
        cycles[blockIdx.x] = clock64() - t0;
    }
    asm volatile("tcgen05.fence::before_thread_sync;");
    __syncthreads();
    if (threadIdx.x < 32) asm volatile("tcgen05.dealloc.cta_group::1.sync.aligned.b32 %0, 512;" ::"r"(tm));
}
int main()
{
    long long* d; CK(cudaMalloc(&d, 148 * 8));
    CK(cudaFuncSetAttribute(bench, cudaFuncAttributeMaxDynamicSharedMemorySize, 64 * 1024));
    const int reps = 4000;
    for (int grid : {1, 148})
        for (int pitch : {32, 64, 128})
            for (int rotate : {0, 1})
                for (int off : {0, 1}) {
                    printf("grid %3d swizzle %3dB rotate %d rowoff %d: cycles/MMA", grid, pitch, rotate, off);
                    for (int N : {16, 32, 48, 64, 96, 128, 256}) {
                        bench<<<grid, 128, 64 * 1024>>>(N, pitch, reps, rotate, off, d);
                        CK(cudaDeviceSynchronize());
                        long long h[148]; CK(cudaMemcpy(h, d, grid * 8, cudaMemcpyDeviceToHost));
                        long long mx = 0; for (int i = 0; i < grid; ++i) mx = h[i] > mx ? h[i] : mx;
                        printf("  N%-3d %6.1f", N, (double)mx / reps);
                    }
                    printf("\n");
                }
    return 0;
}
