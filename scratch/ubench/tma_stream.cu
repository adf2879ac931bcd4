// How fast can per-warp cp.async.bulk rings stream HBM -> shared memory?  (no compute; optional smem read-back)
#include <cstdio>
#include <cstdint>
#include <cuda_runtime.h>
__device__ __forceinline__ uint32_t s32(const void *p) { return (uint32_t)__cvta_generic_to_shared(p); }
__device__ __forceinline__ void mbar_init(uint64_t *b, uint32_t c) { asm volatile("mbarrier.init.shared::cta.b64 [%0], %1;" ::"r"(s32(b)), "r"(c)); }
__device__ __forceinline__ void expect_tx(uint32_t b, uint32_t n) { asm volatile("mbarrier.arrive.expect_tx.shared::cta.b64 _, [%0], %1;" ::"r"(b), "r"(n) : "memory"); }
__device__ __forceinline__ void mwait(uint32_t bar, uint32_t parity) {
    asm volatile("{\n.reg .pred p;\nW_%=:\nmbarrier.try_wait.parity.shared::cta.b64 p, [%0], %1;\n@p bra D_%=;\nbra W_%=;\nD_%=:\n}\n" ::"r"(bar), "r"(parity) : "memory");
}
__device__ __forceinline__ void bulk(uint32_t dst, const void *src, uint32_t bytes, uint32_t bar) {
    asm volatile("cp.async.bulk.shared::cluster.global.mbarrier::complete_tx::bytes [%0], [%1], %2, [%3];" ::"r"(dst), "l"(src), "r"(bytes), "r"(bar) : "memory");
}
// mode 0: every warp owns `stages` buffers of `bytes` and re-issues after each wait
// readback: each lane reads `readback` x 16 B per KB from the tile (simulates the consumer)
__global__ void __launch_bounds__(256, 1) k(const uint8_t *src, size_t per_cta, int bytes, int stages, int warps, int readback, long long *out, int *sink) {
    extern __shared__ __align__(128) uint8_t smem[];
    uint64_t *bars = (uint64_t *)smem;
    const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
    if (threadIdx.x == 0) { for (int i = 0; i < 64; i++) mbar_init(bars + i, 1); asm volatile("fence.mbarrier_init.release.cluster;" ::: "memory"); }
    __syncthreads();
    if (warp >= warps) return;
    const uint32_t ring = s32(smem + 1024) + warp * stages * bytes, bar0 = s32(bars + warp * stages);
    const size_t per_warp = per_cta / warps;
    const uint8_t *base = src + (size_t)blockIdx.x * per_cta + (size_t)warp * per_warp;
    const int n_tiles = (int)(per_warp / bytes);
    long long t0 = clock64(), t_issue = 0, t_wait = 0;
    int issued = 0, acc = 0;
    for (; issued < stages && issued < n_tiles; issued++)
        if (lane == 0) { expect_tx(bar0 + issued * 8, bytes); bulk(ring + issued * bytes, base + (size_t)issued * bytes, bytes, bar0 + issued * 8); }
    for (int t = 0; t < n_tiles; t++) {
        const int s = t % stages;
        long long a0 = clock64();
        mwait(bar0 + s * 8, (t / stages) & 1);
        long long a1 = clock64();
        for (int i = 0; i < readback * bytes / 1024; i++) {
            int4 v;
            asm volatile("ld.shared.v4.b32 {%0,%1,%2,%3}, [%4];" : "=r"(v.x), "=r"(v.y), "=r"(v.z), "=r"(v.w) : "r"(ring + s * bytes + ((i * 32 + lane) * 16) % bytes));
            acc += v.x ^ v.y ^ v.z ^ v.w;
        }
        __syncwarp();
        long long a2 = clock64();
        if (issued < n_tiles) {
            if (lane == 0) { asm volatile("fence.proxy.async.shared::cta;" ::: "memory"); expect_tx(bar0 + s * 8, bytes); bulk(ring + s * bytes, base + (size_t)issued * bytes, bytes, bar0 + s * 8); }
            issued++;
        }
        __syncwarp();
        long long a3 = clock64();
        t_wait += a1 - a0; t_issue += a3 - a2;
    }
    if (acc == 0x12345678) *sink = acc;
    if (blockIdx.x == 0 && threadIdx.x == 0) { out[0] = clock64() - t0; out[1] = t_wait; out[2] = t_issue; out[3] = n_tiles; }
}
int main() {
    const size_t total = (size_t)3 << 30;
    uint8_t *src; long long *out; int *sink;
    cudaMalloc(&src, total); cudaMemset(src, 1, total); cudaMallocManaged(&out, 64); cudaMalloc(&sink, 4);
    cudaFuncSetAttribute(k, cudaFuncAttributeMaxDynamicSharedMemorySize, 220 * 1024);
    const size_t per_cta = (total / 148) / (1 << 20) * (1 << 20);
    cudaEvent_t e0, e1; cudaEventCreate(&e0); cudaEventCreate(&e1);
    struct Cfg { int bytes, stages, warps, readback; } cfgs[] = {
        {9216, 2, 8, 0}, {9216, 2, 8, 2}, {12288, 2, 8, 0}, {4608, 4, 8, 0}, {4608, 5, 8, 0}, {2304, 8, 8, 0}, {2304, 10, 8, 0}, {16384, 3, 4, 0}, {32768, 3, 2, 0}, {32768, 6, 1, 0},
        {9216, 1, 8, 0}, {6144, 2, 16, 0}, {1024, 16, 8, 0}, {8192, 3, 8, 0}};
    for (auto c : cfgs) {
        const size_t smem = 1024 + (size_t)c.bytes * c.stages * c.warps;
        if (smem > 220 * 1024) { printf("skip\n"); continue; }
        for (int rep = 0; rep < 2; rep++) {
            cudaEventRecord(e0);
            k<<<148, 256 * (c.warps > 8 ? 2 : 1) , smem>>>(src, per_cta, c.bytes, c.stages, c.warps, c.readback, out, sink);
            cudaEventRecord(e1);
            cudaError_t e = cudaDeviceSynchronize();
            float ms; cudaEventElapsedTime(&ms, e0, e1);
            if (rep == 1) printf("copy %6d B x %2d stages x %2d warps readback %d: %7.1f GB/s  (in flight/SM %4d KB)  warp0: wait %5.1f%% issue %5.1f%% of %lld cycles, %lld tiles  %s\n", c.bytes, c.stages, c.warps, c.readback,
                                 per_cta * 148 / (ms * 1e-3) / 1e9, c.bytes * c.stages * c.warps / 1024, 100.0 * out[1] / out[0], 100.0 * out[2] / out[0], out[0], out[3], cudaGetErrorString(e));
        }
    }
    return 0;
}
