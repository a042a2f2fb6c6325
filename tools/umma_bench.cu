// Micro-benchmark: cycles per tcgen05.mma for the operand layouts / shapes considered for the conv kernels.
// Build: nvcc -gencode arch=compute_100a,code=sm_100a -O3 -I stofnet_b200/csrc -o tools/umma_bench tools/umma_bench.cu
#include <cstdio>
#include <cstdint>
#include <cuda_runtime.h>
#include "tc_ptx.cuh"
using namespace stof::ptx;

struct Cfg { int n; int layout; int kind_tf32; int reps; int a_shift; int cta; int nacc; int pattern; };

__device__ __forceinline__ void umma_tf32(uint32_t d, uint64_t a, uint64_t b, uint32_t idesc, uint32_t acc) {
    asm volatile("{\n\t.reg .pred p;\n\tsetp.ne.b32 p, %4, 0;\n\ttcgen05.mma.cta_group::1.kind::tf32 [%0], %1, %2, %3, p;\n\t}"
                 :: "r"(d), "l"(a), "l"(b), "r"(idesc), "r"(acc) : "memory");
}

__global__ void __launch_bounds__(128, 1) bench(Cfg c, long long *out) {
    extern __shared__ __align__(1024) uint8_t smem[];
    __shared__ uint64_t bar, bar2, bar3, barB;
    __shared__ uint32_t tm;
    const uint32_t sb = smem_u32(smem);
    for (int i = threadIdx.x; i < 160 * 1024 / 4; i += 128) reinterpret_cast<uint32_t *>(smem)[i] = 0x3c003c00u;
    if (threadIdx.x == 0) { mbar_init(smem_u32(&bar), 1); mbar_init(smem_u32(&bar2), 1); mbar_init(smem_u32(&bar3), 1); mbar_init(smem_u32(&barB), 1); fence_mbar_init(); }
    if (threadIdx.x < 32) tmem_alloc(smem_u32(&tm), 512);
    fence_proxy_async_smem();
    tc_fence_before_sync();
    __syncthreads();
    tc_fence_after_sync();
    const uint32_t tmem = tm;
    const int wid = threadIdx.x >> 5;
    if (threadIdx.x < 32 || (c.pattern >= 64 && wid == 1)) {
        uint32_t idesc = (1u << 4) | ((uint32_t)(c.n >> 3) << 17) | (8u << 24);
        if (c.kind_tf32) idesc |= (2u << 7) | (2u << 10);
        uint32_t hi, a_lo, b_lo, kstep;
        if (c.layout == 0) {          // no swizzle: SBO 128, LBO = 4 KB-ish slab
            hi = (128u >> 4) | (1u << 14);
            a_lo = (sb >> 4) | ((4480u >> 4) << 16);
            b_lo = ((sb + 80 * 1024) >> 4) | ((4096u >> 4) << 16);
            kstep = 2 * 4480 / 16;
        } else {                       // 128B swizzle: SBO 1024, start += 32 B per K step
            hi = (1024u >> 4) | (1u << 14) | (2u << 29);
            a_lo = (sb >> 4) | (1u << 16);
            b_lo = ((sb + 80 * 1024) >> 4) | (1u << 16);
            kstep = 2;
        }
        a_lo += c.a_shift;
        long long t0 = 0, t1 = 0;
        for (int rep = 0; rep < 2; ++rep) {
            __syncwarp();
            t0 = clock64();
            if ((c.pattern & 63) >= 6) {
                // realistic tap loop: [wait on a completed barrier] [fence] elect { 4 x (N128 + N64) ; [commit] }
                const uint32_t ia = (1u << 4) | (16u << 17) | (8u << 24), ib = (1u << 4) | (8u << 17) | (8u << 24);
                for (int i = 0; i < c.reps; ++i) {
                    if (c.pattern & 1) { mbar_wait(smem_u32(&bar2), 1); tc_fence_after_sync(); }
                    if (elect_one()) {
                        const uint32_t d = tmem + ((c.pattern & 8) ? (uint32_t)(i & 1) * 128u : 0u) + (uint32_t)wid * 256u;
#pragma unroll
                        for (int ks = 0; ks < 4; ++ks) {
                            const uint64_t da = make_desc(a_lo + ks * kstep + (i & 7), hi), db = make_desc(b_lo + ks * 512, hi);
                            umma_f16(d, da, db, ia, 1u);
                            umma_f16(d + 64, da, db, ib, 1u);
                        }
                        if (c.pattern & 2) umma_commit(smem_u32(&bar3));
                    }
                    __syncwarp();
                }
                if (elect_one()) umma_commit(smem_u32(wid ? &barB : &bar));
            } else
            if (elect_one()) {
                for (int i = 0; i < c.reps; ++i) {
#pragma unroll
                    for (int ks = 0; ks < 4; ++ks) {
                        const uint64_t da = make_desc(a_lo + ks * kstep, hi), db = make_desc(b_lo + ks * (c.layout ? 2 : 512), hi);
                        const uint32_t d = tmem + (uint32_t)((i * 4 + ks) & (c.nacc - 1)) * (uint32_t)c.a_shift;
                        if (c.pattern == 0) {
                            if (c.kind_tf32) umma_tf32(d, da, db, idesc, 1u); else umma_f16(d, da, db, idesc, 1u);
                        } else if (c.pattern == 1) {     // conv pattern: N=128 into D, N=64 into D+64, same accumulator
                            umma_f16(d, da, db, (1u << 4) | (16u << 17) | (8u << 24), 1u);
                            umma_f16(d + 64, da, db, (1u << 4) | (8u << 17) | (8u << 24), 1u);
                        } else if (c.pattern == 3) {     // a0 b0 a1 b1
                            umma_f16(tmem, da, db, (1u << 4) | (16u << 17) | (8u << 24), 1u);
                            umma_f16(tmem + 64, da, db, (1u << 4) | (8u << 17) | (8u << 24), 1u);
                            umma_f16(tmem + 128, da, db, (1u << 4) | (16u << 17) | (8u << 24), 1u);
                            umma_f16(tmem + 192, da, db, (1u << 4) | (8u << 17) | (8u << 24), 1u);
                        } else if (c.pattern == 4) {     // 4 blocks: a0 a1 a2 a3 b0 b1 b2 b3
                            for (int j = 0; j < 4; ++j) umma_f16(tmem + 128 * j, da, db, (1u << 4) | (16u << 17) | (8u << 24), 1u);
                            for (int j = 0; j < 4; ++j) umma_f16(tmem + 128 * j + 64, da, db, (1u << 4) | (8u << 17) | (8u << 24), 1u);
                        } else if (c.pattern == 5) {     // single N=192 MMA per block (hi|lo|extra), 2 blocks
                            umma_f16(tmem, da, db, (1u << 4) | (24u << 17) | (8u << 24), 1u);
                            umma_f16(tmem + 256, da, db, (1u << 4) | (24u << 17) | (8u << 24), 1u);
                        } else {                         // interleave two blocks: a(b0) a(b1) b(b0) b(b1)
                            umma_f16(tmem, da, db, (1u << 4) | (16u << 17) | (8u << 24), 1u);
                            umma_f16(tmem + 128, da, db, (1u << 4) | (16u << 17) | (8u << 24), 1u);
                            umma_f16(tmem + 64, da, db, (1u << 4) | (8u << 17) | (8u << 24), 1u);
                            umma_f16(tmem + 192, da, db, (1u << 4) | (8u << 17) | (8u << 24), 1u);
                        }
                    }
                }
                umma_commit(smem_u32(&bar));
            }
            __syncwarp();
            mbar_wait(smem_u32(wid ? &barB : &bar), rep & 1);
            t1 = clock64();
        }
        if (threadIdx.x == 0 || threadIdx.x == 32) atomicMax((unsigned long long *)&out[blockIdx.x], (unsigned long long)(t1 - t0));
    }
    tc_fence_before_sync();
    __syncthreads();
    if (threadIdx.x < 32) { tc_fence_after_sync(); tmem_dealloc(tmem, 512); }
}

int main() {
    long long *d; cudaMalloc(&d, 148 * 8);
    cudaFuncSetAttribute(bench, cudaFuncAttributeMaxDynamicSharedMemorySize, 200 * 1024);
    auto run = [&](Cfg c, const char *what, double mmas_per_iter, double macs_per_iter) {
        cudaMemset(d, 0, 148 * 8);
        bench<<<c.cta, 128, 200 * 1024>>>(c, d);
        cudaError_t e = cudaDeviceSynchronize();
        long long h[148]; cudaMemcpy(h, d, c.cta * 8, cudaMemcpyDeviceToHost);
        long long mx = 0; for (int i = 0; i < c.cta; ++i) mx = h[i] > mx ? h[i] : mx;
        const double iters = ((c.pattern & 63) >= 6) ? 256.0 : 256.0 * 4;
        printf("%-40s N=%3d nacc=%d : %.1f cycles/iter  %.1f cycles/MMA  %.0f MAC/cyc err=%d\n", what, c.n, c.nacc, mx / iters, mx / iters / mmas_per_iter,
               macs_per_iter * iters / mx, (int)e);
    };
    for (int n : {64, 128, 256})
        for (int nacc : {1, 2, 4}) {
            if (n == 256 && nacc > 2) continue;
            run(Cfg{n, 0, 0, 256, n == 256 ? 256 : 128, 1, nacc, 0}, "f16 none, rotating accumulators", 1, 128.0 * n * 16);
        }
    run(Cfg{128, 0, 0, 256, 128, 1, 1, 1}, "conv N128+N64 same acc", 2, 128.0 * 192 * 16);
    run(Cfg{128, 0, 0, 256, 128, 1, 2, 1}, "conv, 2 accs alternating per ks", 2, 128.0 * 192 * 16);
    run(Cfg{128, 0, 0, 256, 128, 1, 4, 1}, "conv, 4 accs alternating per ks", 2, 128.0 * 192 * 16);
    run(Cfg{128, 0, 0, 256, 0, 1, 1, 2}, "conv, a0 a1 b0 b1", 4, 2 * 128.0 * 192 * 16);
    run(Cfg{128, 0, 0, 256, 0, 1, 1, 3}, "conv, a0 b0 a1 b1", 4, 2 * 128.0 * 192 * 16);
    run(Cfg{128, 0, 0, 256, 0, 1, 1, 4}, "conv, a0..a3 b0..b3", 8, 4 * 128.0 * 192 * 16);
    run(Cfg{192, 0, 0, 256, 0, 1, 1, 5}, "N=192 x 2 accs", 2, 2 * 128.0 * 192 * 16);
    run(Cfg{128, 0, 0, 256, 0, 1, 1, 64 | 6}, "2 warps, tap loop: elect only", 16, 2 * 4 * 128.0 * 192 * 16);
    run(Cfg{128, 0, 0, 256, 0, 1, 1, 64 | 7 | 2}, "2 warps, tap loop: wait + commit", 16, 2 * 4 * 128.0 * 192 * 16);
    run(Cfg{128, 0, 0, 256, 0, 1, 1, 6}, "tap loop: elect only", 8, 4 * 128.0 * 192 * 16);
    run(Cfg{128, 0, 0, 256, 0, 1, 1, 7}, "tap loop: + barrier wait/fence", 8, 4 * 128.0 * 192 * 16);
    run(Cfg{128, 0, 0, 256, 0, 1, 1, 6 | 2 | 16}, "tap loop: + commit per tap", 8, 4 * 128.0 * 192 * 16);
    run(Cfg{128, 0, 0, 256, 0, 1, 1, 7 | 2 | 16}, "tap loop: + wait + commit", 8, 4 * 128.0 * 192 * 16);
    run(Cfg{128, 0, 0, 256, 0, 1, 1, 7 | 2 | 8 | 16}, "tap loop: + wait + commit, alt acc per tap", 8, 4 * 128.0 * 192 * 16);
    return 0;
}
