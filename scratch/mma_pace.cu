// micro-benchmark: cycles per tcgen05.mma (kind::f16, M = 128, K = 16) by operand source, N and accumulator pattern
#include <cstdio>
#include <cstdint>
#include <cuda_runtime.h>
__device__ __forceinline__ uint32_t s32(const void *p) { return (uint32_t)__cvta_generic_to_shared(p); }
__device__ __forceinline__ uint64_t desc(uint32_t saddr, uint32_t lbo, uint32_t sbo) {
    return (uint64_t)((saddr & 0x3FFFF) >> 4) | ((uint64_t)(lbo >> 4) << 16) | ((uint64_t)(sbo >> 4) << 32) | (1ull << 46);
}
__device__ __forceinline__ void mma_ts(uint32_t d, uint32_t a, uint64_t b, uint32_t idesc, uint32_t acc) {
    asm volatile("{\n.reg .pred p;\nsetp.ne.b32 p, %4, 0;\ntcgen05.mma.cta_group::1.kind::f16 [%0], [%1], %2, %3, p;\n}\n" ::"r"(d), "r"(a), "l"(b), "r"(idesc), "r"(acc) : "memory");
}
__device__ __forceinline__ void mma_ss(uint32_t d, uint64_t a, uint64_t b, uint32_t idesc, uint32_t acc) {
    asm volatile("{\n.reg .pred p;\nsetp.ne.b32 p, %4, 0;\ntcgen05.mma.cta_group::1.kind::f16 [%0], %1, %2, %3, p;\n}\n" ::"r"(d), "l"(a), "l"(b), "r"(idesc), "r"(acc) : "memory");
}
__global__ void __launch_bounds__(128, 1) pace(int mode, int N, int nd, int iters, int M, long long *out) {
    extern __shared__ __align__(1024) uint8_t sm[];
    __shared__ uint64_t bar;
    __shared__ uint32_t slot;
    if (threadIdx.x == 0) { asm volatile("mbarrier.init.shared::cta.b64 [%0], 1;" ::"r"(s32(&bar))); asm volatile("fence.mbarrier_init.release.cluster;" ::: "memory"); }
    for (int i = threadIdx.x; i < 160 * 1024 / 4; i += blockDim.x) ((uint32_t *)sm)[i] = 0x3c003c00u; // fp16 1.0
    if (threadIdx.x < 32) {
        asm volatile("tcgen05.alloc.cta_group::1.sync.aligned.shared::cta.b32 [%0], 512;" ::"r"(s32(&slot)) : "memory");
        asm volatile("tcgen05.relinquish_alloc_permit.cta_group::1.sync.aligned;" ::: "memory");
    }
    asm volatile("fence.proxy.async.shared::cta;" ::: "memory");
    asm volatile("tcgen05.fence::before_thread_sync;" ::: "memory");
    __syncthreads();
    asm volatile("tcgen05.fence::after_thread_sync;" ::: "memory");
    const uint32_t tmem = slot;
    if (threadIdx.x == 0) {
        const uint32_t idesc = (1u << 4) | ((uint32_t)(N >> 3) << 17) | ((uint32_t)(M >> 4) << 24);
        const uint32_t sa = s32(sm), sb = s32(sm + 64 * 1024);
        long long t0 = clock64();
        for (int it = 0; it < iters; it++) {
#pragma unroll
            for (int i = 0; i < 8; i++) {
                const uint32_t d = tmem + ((it * 8 + i) % nd) * 64;
                if (mode == 0) mma_ts(d, tmem + 256 + (it & 3) * 64 + i * 8, desc(sb + i * 256, 128, 4096), idesc, 1);
                else mma_ss(d, desc(sa + (it & 3) * 16384 + i * 256, 128, 1024), desc(sb + i * 256, 128, 4096), idesc, 1);
            }
        }
        long long t1 = clock64();
        asm volatile("tcgen05.commit.cta_group::1.mbarrier::arrive::one.shared::cluster.b64 [%0];" ::"r"(s32(&bar)) : "memory");
        asm volatile("{\n.reg .pred p;\nW:\nmbarrier.try_wait.parity.shared::cta.b64 p, [%0], 0;\n@p bra D;\nbra W;\nD:\n}\n" ::"r"(s32(&bar)) : "memory");
        long long t2 = clock64();
        out[0] = t1 - t0; out[1] = t2 - t0;
    }
    asm volatile("tcgen05.fence::before_thread_sync;" ::: "memory");
    __syncthreads();
    if (threadIdx.x < 32) asm volatile("tcgen05.dealloc.cta_group::1.sync.aligned.b32 %0, 512;" ::"r"(tmem) : "memory");
}
int main() {
    long long *d; cudaMalloc(&d, 16); long long h[2];
    cudaFuncSetAttribute(pace, cudaFuncAttributeMaxDynamicSharedMemorySize, 200 * 1024);
    const int iters = 256;
    struct { int mode, N, nd, M; } cfg[] = {{0, 64, 1, 128}, {0, 64, 2, 128}, {0, 64, 4, 128}, {0, 16, 1, 128}, {0, 128, 1, 128}, {0, 256, 1, 128}, {1, 64, 1, 128}, {1, 64, 2, 128}, {1, 16, 1, 128}, {1, 128, 1, 128},
                                         {1, 256, 1, 128}, {0, 64, 1, 64}, {1, 64, 1, 64}};
    for (auto &c : cfg) {
        for (int rep = 0; rep < 2; rep++) {
            pace<<<1, 128, 200 * 1024>>>(c.mode, c.N, c.nd, iters, c.M, d);
            cudaError_t e = cudaDeviceSynchronize();
            if (e != cudaSuccess) { printf("error %s\n", cudaGetErrorString(e)); return 1; }
        }
        cudaMemcpy(h, d, 16, cudaMemcpyDeviceToHost);
        printf("%s M=%d N=%3d accumulators=%d: issue %.1f cyc/MMA, complete %.1f cyc/MMA\n", c.mode ? "SS" : "TS", c.M, c.N, c.nd, (double)h[0] / (iters * 8), (double)h[1] / (iters * 8));
    }
    return 0;
}
