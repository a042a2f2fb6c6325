// Micro-benchmark: does epilogue traffic (tcgen05.ld, smem stores) slow a concurrent tcgen05.mma stream?
// Build: nvcc -gencode arch=compute_100a,code=sm_100a -O3 -I stofnet_b200/csrc -o tools/umma_interf tools/umma_interf.cu
#include <cstdio>
#include <cstdint>
#include <cuda_runtime.h>
#include "tc_ptx.cuh"
using namespace stof::ptx;

// mode bits: 1 = epilogue warps do tcgen05.ld, 2 = epilogue warps do smem uint4 stores, 4 = smem uint4 loads,
//            8 = second MMA warp (other accumulator), 16 = FMA busy loop
__global__ void __launch_bounds__(352, 1) bench(int mode, int reps, long long *out) {
    extern __shared__ __align__(1024) uint8_t smem[];
    __shared__ uint64_t bar[4];
    __shared__ uint32_t tm;
    __shared__ volatile int stop;
    const uint32_t sb = smem_u32(smem);
    for (int i = threadIdx.x; i < 200 * 1024 / 4; i += blockDim.x) reinterpret_cast<uint32_t *>(smem)[i] = 0x3c003c00u;
    if (threadIdx.x == 0) { for (int i = 0; i < 4; ++i) mbar_init(smem_u32(&bar[i]), 1); stop = 0; fence_mbar_init(); }
    if (threadIdx.x < 32) tmem_alloc(smem_u32(&tm), 512);
    fence_proxy_async_smem();
    tc_fence_before_sync();
    __syncthreads();
    tc_fence_after_sync();
    const uint32_t tmem = tm;
    const int wid = threadIdx.x >> 5, lane = threadIdx.x & 31;
    if (wid == 0 || (wid == 1 && (mode & 8))) {
        const uint32_t hi = (128u >> 4) | (1u << 14);
        const uint32_t a_lo = ((sb + wid * 2048) >> 4) | ((4480u >> 4) << 16);
        const uint32_t b_lo = ((sb + 150 * 1024) >> 4) | ((2048u >> 4) << 16);
        const uint32_t ia = (1u << 4) | (16u << 17) | (8u << 24), ib = (1u << 4) | (8u << 17) | (8u << 24);
        long long t0 = clock64();
        for (int i = 0; i < reps; ++i) {
            if (elect_one()) {
                const uint32_t d = tmem + (uint32_t)wid * 128u;
#pragma unroll
                for (int ks = 0; ks < 4; ++ks) {
                    const uint64_t da = make_desc(a_lo + ks * (2 * 4480 / 16) + (i & 7), hi), db = make_desc(b_lo + ks * 256, hi);
                    umma_f16(d, da, db, ia, 1u);
                    umma_f16(d + 64, da, db, ib, 1u);
                }
                umma_commit(smem_u32(&bar[2 + wid]));
            }
            __syncwarp();
        }
        if (elect_one()) umma_commit(smem_u32(&bar[wid]));
        __syncwarp();
        mbar_wait(smem_u32(&bar[wid]), 0);
        long long t1 = clock64();
        if (lane == 0) { out[2 * blockIdx.x + wid] = t1 - t0; }
        __threadfence_block();
        if (wid == 0 && lane == 0) stop = 1;
    } else if (wid >= 3) {
        const uint32_t lane_addr = (uint32_t)(32 * (wid & 3)) << 16;
        uint32_t acc = 0;
        float f = (float)lane;
        uint8_t *buf = smem + 80 * 1024 + ((wid - 3) * 32 + lane) * 16;
        while (!stop) {
            if (mode & 1) {
                uint32_t d[16];
#pragma unroll
                for (int j = 0; j < 4; ++j) {
                    tmem_ld16(tmem + 256 + lane_addr + 16 * j, d);
                    tmem_ld_wait();
                    acc += d[0] + d[15];
                }
            }
            if (mode & 2) {
#pragma unroll
                for (int j = 0; j < 8; ++j) *reinterpret_cast<uint4 *>(buf + j * 4096) = make_uint4(acc, acc, acc, acc);
            }
            if (mode & 4) {
#pragma unroll
                for (int j = 0; j < 8; ++j) { uint32_t vx, vy, vz, vw; asm volatile("ld.shared.v4.u32 {%0,%1,%2,%3}, [%4];" : "=r"(vx), "=r"(vy), "=r"(vz), "=r"(vw) : "r"(smem_u32(buf + j * 4096))); acc += vx + vw; }
            }
            if (mode & 16) {
#pragma unroll
                for (int j = 0; j < 64; ++j) f = fmaf(f, 1.0001f, 0.5f);
            }
        }
        if (acc == 0x12345 || f == 1.2345f) out[100] = acc;
    }
    tc_fence_before_sync();
    __syncthreads();
    if (threadIdx.x < 32) { tc_fence_after_sync(); tmem_dealloc(tmem, 512); }
}

int main() {
    long long *d; cudaMalloc(&d, 4096);
    cudaFuncSetAttribute(bench, cudaFuncAttributeMaxDynamicSharedMemorySize, 200 * 1024);
    const int reps = 512;
    for (int mode : {0, 1, 2, 4, 16, 3, 7, 23, 8, 9, 10, 12, 15, 31}) {
        cudaMemset(d, 0, 4096);
        bench<<<1, 352, 200 * 1024>>>(mode, reps, d);
        cudaError_t e = cudaDeviceSynchronize();
        long long h[2]; cudaMemcpy(h, d, 16, cudaMemcpyDeviceToHost);
        printf("mode %2d (ld=%d st=%d lds=%d fma=%d mma2=%d): warp0 %.1f cycles/tap  warp1 %.1f  err=%d\n", mode, mode & 1, (mode >> 1) & 1,
               (mode >> 2) & 1, (mode >> 4) & 1, (mode >> 3) & 1, (double)h[0] / reps, (double)h[1] / reps, (int)e);
    }
    return 0;
}
