// grid-barrier micro-benchmark: cycles per barrier for several implementations (148 CTAs x 256 threads)
#include <cooperative_groups.h>
#include <cstdio>
#include <cuda_runtime.h>
namespace cg = cooperative_groups;

__device__ __forceinline__ unsigned ld_acquire(const unsigned *p) { unsigned v; asm volatile("ld.acquire.gpu.global.u32 %0, [%1];" : "=r"(v) : "l"(p) : "memory"); return v; }
__device__ __forceinline__ unsigned ld_relaxed(const unsigned *p) { unsigned v; asm volatile("ld.relaxed.gpu.global.u32 %0, [%1];" : "=r"(v) : "l"(p) : "memory"); return v; }
__device__ __forceinline__ unsigned ld_volatile(const unsigned *p) { unsigned v; asm volatile("ld.volatile.global.u32 %0, [%1];" : "=r"(v) : "l"(p) : "memory"); return v; }

// A: atomic counter + generation (v4)
__device__ __forceinline__ void bar_A(unsigned *bar, unsigned n, unsigned &gen) {
    __syncthreads();
    if (threadIdx.x == 0) {
        unsigned prev;
        asm volatile("atom.add.acq_rel.gpu.global.u32 %0, [%1], 1;" : "=r"(prev) : "l"(bar) : "memory");
        if (prev == n - 1) {
            asm volatile("st.relaxed.gpu.global.u32 [%0], 0;" ::"l"(bar) : "memory");
            asm volatile("red.release.gpu.global.add.u32 [%0], 1;" ::"l"(bar + 1) : "memory");
        } else {
            while (ld_acquire(bar + 1) == gen) { }
        }
        gen++;
    }
    __syncthreads();
}
// B: per-CTA flags, release store, relaxed polls by warp 0, trailing fence
__device__ __forceinline__ void bar_B(unsigned *flags, unsigned n, unsigned epoch) {
    __syncthreads();
    if (threadIdx.x == 0) asm volatile("st.release.gpu.global.u32 [%0], %1;" ::"l"(flags + blockIdx.x), "r"(epoch) : "memory");
    if (threadIdx.x < 32) {
        bool done;
        do {
            done = true;
            for (unsigned c = threadIdx.x; c < n; c += 32) done &= (int)(ld_relaxed(flags + c) - epoch) >= 0;
        } while (!__all_sync(0xffffffffu, done));
        asm volatile("fence.acq_rel.gpu;" ::: "memory");
    }
    __syncthreads();
}
// C: like B without the trailing fence (volatile polls)
__device__ __forceinline__ void bar_C(unsigned *flags, unsigned n, unsigned epoch) {
    __syncthreads();
    if (threadIdx.x == 0) { __threadfence(); asm volatile("st.relaxed.gpu.global.u32 [%0], %1;" ::"l"(flags + blockIdx.x), "r"(epoch) : "memory"); }
    if (threadIdx.x < 32) {
        bool done;
        do {
            done = true;
            for (unsigned c = threadIdx.x; c < n; c += 32) done &= (int)(ld_volatile(flags + c) - epoch) >= 0;
        } while (!__all_sync(0xffffffffu, done));
    }
    __syncthreads();
}
// E: monotonic counter: red.release add, poll until >= target
__device__ __forceinline__ void bar_E(unsigned *cnt, unsigned n, unsigned target) {
    __syncthreads();
    if (threadIdx.x == 0) {
        asm volatile("red.release.gpu.global.add.u32 [%0], 1;" ::"l"(cnt) : "memory");
        while ((int)(ld_acquire(cnt) - target) < 0) { }
    }
    __syncthreads();
}
// F: monotonic counter, relaxed red after a fence, volatile polls, no trailing fence
__device__ __forceinline__ void bar_F(unsigned *cnt, unsigned n, unsigned target) {
    __syncthreads();
    if (threadIdx.x == 0) {
        __threadfence();
        asm volatile("red.relaxed.gpu.global.add.u32 [%0], 1;" ::"l"(cnt) : "memory");
        while ((int)(ld_volatile(cnt) - target) < 0) { }
    }
    __syncthreads();
}

// G: counter + nanosleep backoff
__device__ __forceinline__ void bar_G(unsigned *cnt, unsigned n, unsigned target) {
    __syncthreads();
    if (threadIdx.x == 0) {
        asm volatile("red.release.gpu.global.add.u32 [%0], 1;" ::"l"(cnt) : "memory");
        while ((int)(ld_relaxed(cnt) - target) < 0) __nanosleep(32);
        asm volatile("fence.acq_rel.gpu;" ::: "memory");
    }
    __syncthreads();
}
// H: monotonic counter with returned value; the last arriver publishes a separate go word that everyone polls
__device__ __forceinline__ void bar_H(unsigned *cnt, unsigned n, unsigned target, unsigned epoch) {
    __syncthreads();
    if (threadIdx.x == 0) {
        unsigned prev;
        asm volatile("atom.add.acq_rel.gpu.global.u32 %0, [%1], 1;" : "=r"(prev) : "l"(cnt) : "memory");
        if (prev + 1 == target) asm volatile("st.release.gpu.global.u32 [%0], %1;" ::"l"(cnt + 32), "r"(epoch) : "memory");
        else while ((int)(ld_acquire(cnt + 32) - epoch) < 0) { }
    }
    __syncthreads();
}
// I: like E but only lane 0 of warp 0 of every 4th CTA... (skipped) -- instead: E with polls by relaxed loads + one fence
__device__ __forceinline__ void bar_I(unsigned *cnt, unsigned n, unsigned target) {
    __syncthreads();
    if (threadIdx.x == 0) {
        asm volatile("red.release.gpu.global.add.u32 [%0], 1;" ::"l"(cnt) : "memory");
        while ((int)(ld_relaxed(cnt) - target) < 0) { }
        asm volatile("fence.acq_rel.gpu;" ::: "memory");
    }
    __syncthreads();
}

__global__ void k(int which, unsigned *bar, float *data, int iters, long long *out) {
    cg::grid_group g = cg::this_grid();
    unsigned gen = 0;
    const unsigned n = gridDim.x;
    long long t0 = 0;
    for (int it = 0; it < iters + 10; it++) {
        if (it == 10) t0 = clock64();
        // a little traffic so that the release has something to order
        data[(size_t)blockIdx.x * 256 + threadIdx.x] = (float)it;
        switch (which) {
            case 0: bar_A(bar, n, gen); break;
            case 1: bar_B(bar + 64, n, it + 1); break;
            case 2: bar_C(bar + 64, n, it + 1); break;
            case 3: g.sync(); break;
            case 4: bar_E(bar + 32, n, (it + 1) * n); break;
            case 5: bar_F(bar + 32, n, (it + 1) * n); break;
            case 6: bar_G(bar + 32, n, (it + 1) * n); break;
            case 7: bar_H(bar + 128, n, (it + 1) * n, it + 1); break;
            case 8: bar_I(bar + 32, n, (it + 1) * n); break;
            case 9: break; // no barrier: cost of the loop body alone
        }
        // check visibility of a neighbour's write (correctness smoke test)
        const float v = __ldcg(&data[(size_t)((blockIdx.x + 1) % n) * 256 + threadIdx.x]);
        if (v != (float)it) atomicAdd((unsigned long long *)&out[1], 1ull);
        __syncthreads();
    }
    if (blockIdx.x == 0 && threadIdx.x == 0) out[0] = (clock64() - t0) / iters;
}

int main() {
    unsigned *bar; float *data; long long *out;
    cudaMalloc(&bar, 8192); cudaMalloc(&data, 148 * 256 * 4); cudaMallocManaged(&out, 16);
    int n_sm = 0; cudaDeviceGetAttribute(&n_sm, cudaDevAttrMultiProcessorCount, 0);
    const char *names[] = {"A atomic+generation (acq_rel / ld.acquire)", "B flags: st.release, relaxed polls, fence", "C flags: fence+st.relaxed, volatile polls, no fence", "D cooperative_groups grid.sync", "E counter: red.release + ld.acquire polls", "F counter: fence+red.relaxed + volatile polls", "G counter: red.release, relaxed polls + nanosleep(32), fence", "H counter (atom) + separate go word", "I counter: red.release, relaxed polls, fence", "- no barrier (loop body only)"};
    for (int which = 0; which < 10; which++) {
        cudaMemset(bar, 0, 8192); out[0] = out[1] = 0;
        int iters = 2000;
        void *args[] = {&which, &bar, &data, &iters, &out};
        cudaLaunchCooperativeKernel((const void *)k, dim3(n_sm), dim3(256), args, 0, 0);
        cudaError_t e = cudaDeviceSynchronize();
        printf("%-55s %6lld cycles/iter (incl. ~1 L2 write + read)  stale reads %lld  %s\n", names[which], out[0], out[1], cudaGetErrorString(e));
    }
    return 0;
}
