// Shared-memory-port microbenchmark for the chain kernel (tools/, not part of the library).
// Question (round 2): the chain kernel's layer period (~8,000 cycles per 256-row tile and layer) matches the sum of ALL its
// shared-memory traffic at 128 B/clk (tcgen05 operand reads 784 KB + epilogue LDS/STS ~110-180 KB + weight-ring bulk writes
// 112 KB), not the tensor pipe's 5,376 cycles.  This benchmark measures, on one CTA per SM with the kernel's operand layouts:
//   * the MMA rate of two issuing warps (today's pattern v0, and the tap-PAIR pattern vP: X_hi(t) x [W(t); W(t+1)] N = 256 and
//     X_lo(t) x [W_hi(t); W_hi(t+1)] N = 128, which reads 20 KB instead of 28 KB per tap pair and K-step),
//   * with 8 "epilogue" warps hammering shared memory with LDS.128 / STS.128 at the same time (achieved LSU bytes / clk),
//   * and the TMEM -> register bandwidth of tcgen05.ld (8 warps, 32x32b.x32), alone and under MMA load.
//   nvcc -gencode arch=compute_100a,code=sm_100a -O3 -o tools/umma_bench3 tools/umma_bench3.cu && tools/umma_bench3
#include <cstdio>
#include <cstdint>
#include <cuda_runtime.h>
#include "../stofnet_b200/csrc/tc_ptx.cuh"

using namespace stof::ptx;

constexpr int R = 280, SLAB = R * 16, PLANE = 8 * SLAB, BUF = 2 * PLANE;
constexpr int OFF_W = BUF, W_BYTES = 81920, OFF_X = OFF_W + W_BYTES, X_BYTES = 4096, OFF_BAR = OFF_X + X_BYTES, SMEM = OFF_BAR + 128;

struct Cfg {
    int variant;   // 0: today (N128 + N64 per tap and K-step); 1: tap pairs (N256 + N128 per 2 taps, last tap N128 + N64); 2: no MMAs
    int reps;
    int lsu;       // bit 0: STS.128, bit 1: LDS.128 loops on the 8 epilogue warps (conflict-free 512 B per warp instruction)
    int ldtm;      // 1: the 8 epilogue warps run tcgen05.ld 32x32b.x32 loops
    int pace;      // idle FMAs between LSU instructions (0 = flat out)
    int tma;       // > 0: warp 2 streams 16 KB bulk copies global -> shared into the weight area with this many in flight
    int fence;     // 1: a fence.proxy.async.shared::cta after every 2 STS (as the chain kernel's epilogue does per chunk)
};

__device__ __forceinline__ void tmem_ld32(uint32_t taddr, uint32_t (&r)[32]) {
    asm volatile(
        "tcgen05.ld.sync.aligned.32x32b.x32.b32 {%0, %1, %2, %3, %4, %5, %6, %7, %8, %9, %10, %11, %12, %13, %14, %15, %16, %17, %18, %19, "
        "%20, %21, %22, %23, %24, %25, %26, %27, %28, %29, %30, %31}, [%32];"
        : "=r"(r[0]), "=r"(r[1]), "=r"(r[2]), "=r"(r[3]), "=r"(r[4]), "=r"(r[5]), "=r"(r[6]), "=r"(r[7]), "=r"(r[8]), "=r"(r[9]),
          "=r"(r[10]), "=r"(r[11]), "=r"(r[12]), "=r"(r[13]), "=r"(r[14]), "=r"(r[15]), "=r"(r[16]), "=r"(r[17]), "=r"(r[18]),
          "=r"(r[19]), "=r"(r[20]), "=r"(r[21]), "=r"(r[22]), "=r"(r[23]), "=r"(r[24]), "=r"(r[25]), "=r"(r[26]), "=r"(r[27]),
          "=r"(r[28]), "=r"(r[29]), "=r"(r[30]), "=r"(r[31])
        : "r"(taddr));
}

__global__ void __launch_bounds__(352, 1) bench(Cfg c, long long *out, const uint8_t *gsrc) {
    extern __shared__ __align__(1024) uint8_t smem[];
    const uint32_t sb = smem_u32(smem);
    __shared__ uint32_t tm;
    __shared__ volatile int stop;
    const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
    for (int i = threadIdx.x; i < SMEM / 4; i += 352) {
        uint32_t h = (uint32_t)i * 2654435761u + blockIdx.x * 40503u; h ^= h >> 15; h *= 2246822519u; h ^= h >> 13;
        const uint32_t a = (h & 0x83ffu) | ((11u + ((h >> 10) & 3u)) << 10), b = ((h >> 16) & 0x83ffu) | ((11u + ((h >> 26) & 3u)) << 10);
        reinterpret_cast<uint32_t *>(smem)[i] = a | (b << 16);
    }
    const uint32_t bar = sb + OFF_BAR;
    if (threadIdx.x == 0) { mbar_init(bar, 2); for (int k = 0; k < 4; ++k) mbar_init(bar + 32 + 8 * k, 1); stop = 0; fence_mbar_init(); }
    if (warp == 0) tmem_alloc(smem_u32(&tm), 512);
    fence_proxy_async_smem();
    tc_fence_before_sync();
    __syncthreads();
    tc_fence_after_sync();
    const uint32_t tmem = tm;
    constexpr uint32_t kHi = (128u >> 4) | (1u << 14);
    constexpr uint32_t kALo = (uint32_t)(SLAB >> 4) << 16;
    const uint32_t i128 = umma_idesc_f16(128, 128), i64 = umma_idesc_f16(128, 64), i256 = umma_idesc_f16(128, 256);
    if (warp < 2) {
        const int b = warp;
        long long t0 = clock64();
        if (c.variant != 2) {
            for (int rep = 0; rep < c.reps; ++rep) {
                if (elect_one()) {
                    auto A = [&](int plane, int ks, int t) {
                        return make_desc((((sb) >> 4) | kALo) + (uint32_t)(3 + t + 128 * b) + (uint32_t)(plane * 8 + 2 * ks) * R, kHi);
                    };
                    if (c.variant == 0) {
                        for (int t = 0; t < 7; ++t) {
                            const uint32_t w_lo = ((sb + OFF_W + (t % 5) * 16384) >> 4) | (128u << 16);
#pragma unroll
                            for (int ks = 0; ks < 4; ++ks) {
                                umma_f16(tmem + 256 * b, A(0, ks, t), make_desc(w_lo + 2 * ks * 128, kHi), i128, 1u);
                                umma_f16(tmem + 256 * b + 64, A(1, ks, t), make_desc(w_lo + 2 * ks * 128, kHi), i64, 1u);
                            }
                        }
                    } else if (c.variant == 3) {
                        // today's MMAs, K-group outer / tap inner (the order of the K-group-pipelined kernel)
#pragma unroll 1
                        for (int ks = 0; ks < 4; ++ks)
#pragma unroll
                            for (int t = 0; t < 7; ++t) {
                                const uint32_t w_lo = ((sb + OFF_W + ((ks * 2 + (t >> 2)) % 5) * 16384 + (t & 3) * 4096) >> 4) | (128u << 16);
                                umma_f16(tmem + 256 * b, A(0, ks, t), make_desc(w_lo, kHi), i128, 1u);
                                umma_f16(tmem + 256 * b + 64, A(1, ks, t), make_desc(w_lo, kHi), i64, 1u);
                            }
                    } else {
                        // K-group outer; per K-group 3 tap pairs (N256 hi + N128 lo) and the single last tap (N128 + N64)
#pragma unroll 1
                        for (int ks = 0; ks < 4; ++ks) {
#pragma unroll
                            for (int pr = 0; pr < 3; ++pr) {
                                const uint32_t w_lo = ((sb + OFF_W + ((ks * 4 + pr) % 10) * 8192) >> 4) | (256u << 16);
                                umma_f16(tmem + 256 * b, A(0, ks, 2 * pr), make_desc(w_lo, kHi), i256, 1u);
                                umma_f16(tmem + 256 * b + 128, A(1, ks, 2 * pr), make_desc(w_lo, kHi), i128, 1u);
                            }
                            const uint32_t w_lo = ((sb + OFF_W + ((ks * 4 + 3) % 10) * 8192) >> 4) | (128u << 16);
                            umma_f16(tmem + 256 * b + 64, A(0, ks, 6), make_desc(w_lo, kHi), i128, 1u);
                            umma_f16(tmem + 256 * b + 128, A(1, ks, 6), make_desc(w_lo, kHi), i64, 1u);
                        }
                    }
                }
                __syncwarp();
            }
        }
        if (elect_one()) umma_commit(bar);
        __syncwarp();
        mbar_wait(bar, 0);
        long long t1 = clock64();
        if (blockIdx.x == 0 && lane == 0) out[warp] = t1 - t0;
        if (c.variant != 2 && warp == 0 && lane == 0) stop = 1;
    } else if (warp == 2 && c.tma > 0) {
        if (lane == 0) {
            const uint32_t tb = bar + 32;
            int issued = 0, done = 0;
            uint32_t ph[4] = {0, 0, 0, 0};
            const long long c0 = clock64();
            while (!stop) {
                while (issued - done < c.tma) {
                    const int k = issued % c.tma;
                    mbar_arrive_expect_tx(tb + 8 * k, 16384);
                    bulk_g2s(sb + OFF_W + (issued % 5) * 16384, gsrc + (size_t)(issued % 77) * 16384, 16384, tb + 8 * k);
                    ++issued;
                }
                const int k = done % c.tma;
                mbar_wait(tb + 8 * k, ph[k]);
                ph[k] ^= 1u;
                ++done;
            }
            while (done < issued) { const int k = done % c.tma; mbar_wait(tb + 8 * k, ph[k]); ph[k] ^= 1u; ++done; }
            if (blockIdx.x == 0) { out[2] = clock64() - c0; out[3] = done; }
        }
    } else if (warp >= 3) {
        // "epilogue" warps: shared-memory and / or TMEM traffic until the MMA warps are done (variant 2: fixed iteration count)
        const uint32_t lane_addr = (uint32_t)(32 * (warp & 3)) << 16;
        uint8_t *buf = smem + ((warp - 3) * 32 + lane) * 16;          // 8 warps x 512 B rows, + j * 4096 per instruction
        uint32_t acc = lane;
        float f = (float)lane;
        long long bytes_lsu = 0, bytes_tm = 0;
        const long long t0 = clock64();
        int iter = 0;
        while (c.variant == 2 ? (iter < c.reps) : !stop) {
            ++iter;
            if (c.lsu & 1) {
#pragma unroll
                for (int j = 0; j < 8; ++j) {
                    *reinterpret_cast<uint4 *>(buf + (j & 3) * 4096 + (j >> 2) * 71680) = make_uint4(acc, acc, acc, acc);
                    if (c.fence == 1 && (j & 1)) fence_proxy_async_smem();
                    for (int q = 0; q < c.pace; ++q) f = fmaf(f, 1.0001f, 0.5f);
                    if (c.fence == 2 && (j & 1)) fence_proxy_async_smem();      // fence delayed by the pacing work
                }
                bytes_lsu += 8 * 512;
            }
            if (c.lsu & 2) {
#pragma unroll
                for (int j = 0; j < 8; ++j) {
                    uint32_t vx, vy, vz, vw;
                    asm volatile("ld.shared.v4.u32 {%0,%1,%2,%3}, [%4];" : "=r"(vx), "=r"(vy), "=r"(vz), "=r"(vw) : "r"(smem_u32(buf + (j & 3) * 4096 + (j >> 2) * 71680)));
                    acc += vx + vw;
                    for (int q = 0; q < c.pace; ++q) f = fmaf(f, 1.0001f, 0.5f);
                }
                bytes_lsu += 8 * 512;
            }
            if (c.ldtm) {
                uint32_t d[32];
#pragma unroll
                for (int j = 0; j < 4; ++j) {
                    tmem_ld32(tmem + lane_addr + 32 * j + 256 * (warp >= 7), d);
                    tmem_ld_wait();
                    acc += d[0] + d[31];
                }
                bytes_tm += 4 * 32 * 32 * 4;
            }
        }
        const long long t1 = clock64();
        if (blockIdx.x == 0 && lane == 0) {
            out[4 + (warp - 3) * 3 + 0] = t1 - t0;
            out[4 + (warp - 3) * 3 + 1] = bytes_lsu;
            out[4 + (warp - 3) * 3 + 2] = bytes_tm;
        }
        if (acc == 0x12345 || f == 1.2345f) out[63] = acc;
    }
    tc_fence_before_sync();
    __syncthreads();
    if (warp == 0) { tc_fence_after_sync(); tmem_dealloc(tmem, 512); }
}

int main() {
    long long *d;
    cudaMalloc(&d, 512);
    uint8_t *gsrc;
    cudaMalloc(&gsrc, 77 * 16384);
    cudaMemset(gsrc, 0x3c, 77 * 16384);
    cudaFuncSetAttribute(bench, cudaFuncAttributeMaxDynamicSharedMemorySize, SMEM);
    const char *vn[] = {"v0 today (N128+N64 per tap, tap outer)", "vP tap pairs (N256+N128 per 2 taps)", "no MMAs", "v0k today's MMAs, K-group outer"};
    struct Run { int variant, lsu, ldtm, pace, tma, fence; };
    const Run runs[] = {{2, 1, 0, 16, 0, 0}, {2, 1, 0, 16, 0, 1}, {2, 1, 0, 16, 0, 2}, {2, 1, 0, 64, 0, 1}, {2, 1, 0, 64, 0, 2},
                        {0, 1, 0, 16, 0, 0}, {0, 1, 0, 16, 0, 1}, {0, 1, 0, 16, 0, 2}, {0, 1, 0, 64, 0, 0}, {0, 1, 0, 64, 0, 1}, {0, 1, 0, 64, 0, 2},
                        {0, 1, 0, 128, 0, 1}, {0, 1, 0, 128, 0, 2},
                        {1, 1, 0, 64, 0, 1}, {1, 1, 0, 64, 0, 2}, {1, 1, 0, 128, 0, 1}, {1, 1, 0, 128, 0, 2}};
    for (const Run &r : runs) {
        Cfg c{r.variant, r.variant == 2 ? 4096 : 256, r.lsu, r.ldtm, r.pace, r.tma, r.fence};
        long long h[64] = {0};
        for (int it = 0; it < 2; ++it) {
            cudaMemset(d, 0, 512);
            bench<<<148, 352, SMEM>>>(c, d, gsrc);
            cudaError_t e = cudaDeviceSynchronize();
            if (e != cudaSuccess) { printf("error: %s\n", cudaGetErrorString(e)); return 1; }
        }
        cudaMemcpy(h, d, 512, cudaMemcpyDeviceToHost);
        const long long cyc = h[0] > h[1] ? h[0] : h[1];
        double lsu_b = 0, tm_b = 0, ecyc = 0;
        for (int w = 0; w < 8; ++w) { lsu_b += (double)h[4 + 3 * w + 1]; tm_b += (double)h[4 + 3 * w + 2]; if ((double)h[4 + 3 * w] > ecyc) ecyc = (double)h[4 + 3 * w]; }
        printf("%-40s lsu=%d ldtm=%d pace=%2d tma=%d fence=%d : %7.1f cycles per tap (both blocks) | LSU %6.1f B/clk | TMEM ld %6.1f B/clk | bulk %6.1f B/clk\n",
               vn[r.variant], r.lsu, r.ldtm, r.pace, r.tma, r.fence, r.variant == 2 ? 0.0 : (double)cyc / (c.reps * 7), ecyc > 0 ? lsu_b / ecyc : 0.0,
               ecyc > 0 ? tm_b / ecyc : 0.0, h[2] > 0 ? (double)h[3] * 16384.0 / (double)h[2] : 0.0);
    }
    return 0;
}
