// Issue-pattern microbenchmark for the chain kernel's inner loop (tools/, not part of the library):
// one CTA per SM, operands laid out exactly like conv_tc.cu (split hi/lo planes [plane][c8][280 rows][16 B], weight taps
// [c8][128 rows][16 B] in a 5-stage ring), two 128-row blocks.  Each variant issues  reps x 7 taps x 4 k-steps  and reports
// cycles per tap (both blocks) so that variants can be compared with the kernel's in-situ ~1250 cycles / tap.
//   nvcc -gencode arch=compute_100a,code=sm_100a -O3 -o tools/umma_bench2 tools/umma_bench2.cu && tools/umma_bench2
#include <cstdio>
#include <cstdint>
#include <cuda_runtime.h>
#include "../stofnet_b200/csrc/tc_ptx.cuh"

using namespace stof::ptx;

constexpr int R = 280, SLAB = R * 16, PLANE = 8 * SLAB, BUF = 2 * PLANE;
constexpr int STAGE = 16384, STAGES = 5;
constexpr int OFF_W = BUF, OFF_BAR = OFF_W + STAGES * STAGE, SMEM = OFF_BAR + 128;

struct Cfg {
    int variant;      // see main()
    int reps;
    int data;         // 0: all operands 1.0, 1: pseudo-random fp16 in [-2, 2) (switching activity like real activations)
    int tma;          // > 0: warp 2 streams 16 KB bulk copies into the weight ring with this many in flight (no pacing by the MMAs)
    int spin;         // 0: no pollers, 1: 9 warps spin on try_wait (as the kernel's idle roles do), 2: same with a suspend-time hint,
                      // 3: only lane 0 polls, 4: lane 0 polls with nanosleep backoff
};

__device__ __forceinline__ bool try_wait_hint(uint32_t bar, uint32_t parity, uint32_t ticks) {
    uint32_t ok;
    asm volatile("{\n\t.reg .pred p;\n\tmbarrier.try_wait.parity.shared::cta.b64 p, [%1], %2, %3;\n\tselp.b32 %0, 1, 0, p;\n\t}"
                 : "=r"(ok) : "r"(bar), "r"(parity), "r"(ticks) : "memory");
    return ok != 0;
}

__global__ void __launch_bounds__(416, 1) bench(Cfg c, long long *out, const uint8_t *gsrc) {
    extern __shared__ __align__(1024) uint8_t smem[];
    const uint32_t sb = smem_u32(smem);
    __shared__ uint32_t tm;
    const int warp = threadIdx.x >> 5;
    for (int i = threadIdx.x; i < SMEM / 4; i += 416) {
        uint32_t w = 0x3c003c00u;                                   // fp16 1.0
        if (c.data) {
            uint32_t h = (uint32_t)i * 2654435761u + blockIdx.x * 40503u; h ^= h >> 15; h *= 2246822519u; h ^= h >> 13;
            // sign + exponent in [11, 15] + random mantissa: |v| in [2^-4, 2)
            const uint32_t a = (h & 0x83ffu) | ((11u + ((h >> 10) & 3u)) << 10), b = ((h >> 16) & 0x83ffu) | ((11u + ((h >> 26) & 3u)) << 10);
            w = a | (b << 16);
        }
        reinterpret_cast<uint32_t *>(smem)[i] = w;
    }
    const uint32_t bar = sb + OFF_BAR;
    if (threadIdx.x == 0) { mbar_init(bar, 1); mbar_init(bar + 8, 1); mbar_init(bar + 16, 2); mbar_init(bar + 24, 1); for (int k = 0; k < 4; ++k) mbar_init(bar + 32 + 8 * k, 1); fence_mbar_init(); }
    if (warp == 0) tmem_alloc(smem_u32(&tm), 512);
    fence_proxy_async_smem();
    tc_fence_before_sync();
    __syncthreads();
    tc_fence_after_sync();
    const uint32_t tmem = tm;
    constexpr uint32_t kHi = (128u >> 4) | (1u << 14);
    constexpr uint32_t kALo = (uint32_t)(SLAB >> 4) << 16;
    const uint32_t ia = umma_idesc_f16(128, 128), ib = umma_idesc_f16(128, 64), i192 = umma_idesc_f16(128, 192);
    const int v = c.variant % 30;            // variants 30 + v: variant v issued by ONE warp (block 0 only)
    const bool solo = c.variant >= 30;
    long long t0 = 0, t1 = 0;
    const bool two_warps = (v < 10) && !solo;
    if (warp < 2 && (two_warps || warp == 0)) {
        const int b = warp;
        __syncwarp();
        t0 = clock64();
        int s = 0;
        for (int rep = 0; rep < c.reps; ++rep) {
            for (int t = 0; t < 7; ++t) {
                if (elect_one()) {
                    const uint32_t w_lo = ((sb + OFF_W + s * STAGE) >> 4) | (128u << 16);
                    auto A = [&](int blk, int plane, int ks) {
                        return make_desc((((sb) >> 4) | kALo) + (uint32_t)(3 + t + 128 * blk) + (uint32_t)(plane * 8 + 2 * ks) * R, kHi);
                    };
                    auto B = [&](int ks, int row0) { return make_desc(w_lo + 2 * ks * 128 + row0, kHi); };
                    if (v == 0) {            // the kernel today: N128 (hi) then N64 (lo) into overlapping columns
#pragma unroll
                        for (int ks = 0; ks < 4; ++ks) {
                            umma_f16(tmem + 128 * b, A(b, 0, ks), B(ks, 0), ia, 1u);
                            umma_f16(tmem + 128 * b + 64, A(b, 1, ks), B(ks, 0), ib, 1u);
                        }
                    } else if (v == 1) {     // lo product into its own columns
#pragma unroll
                        for (int ks = 0; ks < 4; ++ks) {
                            umma_f16(tmem + 192 * b, A(b, 0, ks), B(ks, 0), ia, 1u);
                            umma_f16(tmem + 192 * b + 128, A(b, 1, ks), B(ks, 0), ib, 1u);
                        }
                    } else if (v == 2) {     // all hi k-steps, then all lo k-steps (own columns)
#pragma unroll
                        for (int ks = 0; ks < 4; ++ks) umma_f16(tmem + 192 * b, A(b, 0, ks), B(ks, 0), ia, 1u);
#pragma unroll
                        for (int ks = 0; ks < 4; ++ks) umma_f16(tmem + 192 * b + 128, A(b, 1, ks), B(ks, 0), ib, 1u);
                    } else if (v == 3) {     // A reuse test: lo uses the hi descriptor
#pragma unroll
                        for (int ks = 0; ks < 4; ++ks) {
                            umma_f16(tmem + 128 * b, A(b, 0, ks), B(ks, 0), ia, 1u);
                            umma_f16(tmem + 128 * b + 64, A(b, 0, ks), B(ks, 0), ib, 1u);
                        }
                    } else if (v == 4) {     // hi only (N128)
#pragma unroll
                        for (int ks = 0; ks < 4; ++ks) umma_f16(tmem + 128 * b, A(b, 0, ks), B(ks, 0), ia, 1u);
                    } else if (v == 5) {     // lo only (N64)
#pragma unroll
                        for (int ks = 0; ks < 4; ++ks) umma_f16(tmem + 128 * b + 64, A(b, 1, ks), B(ks, 0), ib, 1u);
                    } else if (v == 6) {     // hi only, two accumulators alternating per k-step
#pragma unroll
                        for (int ks = 0; ks < 4; ++ks) umma_f16(tmem + 256 * b + 128 * (ks & 1), A(b, 0, ks), B(ks, 0), ia, 1u);
                    } else if (v == 7) {     // today's pattern without the tap shift (aligned A)
#pragma unroll
                        for (int ks = 0; ks < 4; ++ks) {
                            umma_f16(tmem + 128 * b, make_desc((((sb) >> 4) | kALo) + (uint32_t)(128 * b) + (uint32_t)(2 * ks) * R, kHi), B(ks, 0), ia, 1u);
                            umma_f16(tmem + 128 * b + 64, make_desc((((sb) >> 4) | kALo) + (uint32_t)(128 * b) + (uint32_t)(8 + 2 * ks) * R, kHi), B(ks, 0), ib, 1u);
                        }
                    } else if (v == 9) {     // term split: BOTH warps work on block 0 - warp 0 issues the hi products, warp 1 the lo ones (own columns)
                        if (b == 0) {
#pragma unroll
                            for (int ks = 0; ks < 4; ++ks) umma_f16(tmem, A(0, 0, ks), B(ks, 0), ia, 1u);
                        } else {
#pragma unroll
                            for (int ks = 0; ks < 4; ++ks) umma_f16(tmem + 128, A(0, 1, ks), B(ks, 0), ib, 1u);
                        }
                    } else if (v == 8) {     // N192 single MMA (hi x [Whi;Wlo;Whi-ish]) - operand-traffic probe
#pragma unroll
                        for (int ks = 0; ks < 4; ++ks) umma_f16(tmem + 256 * b, A(b, 0, ks), B(ks, 0), i192, 1u);
                    }
                    if (v == 20 || v == 21) {  // contract conv pattern: one issuer, block = t & 1, per k-step N256 (hi) + N128 (lo)
                        const uint32_t i256 = umma_idesc_f16(128, 256), i128 = umma_idesc_f16(128, 128);
                        const uint32_t wl = ((sb + OFF_W) >> 4) | (256u << 16);
                        const int blk = t & 1;
#pragma unroll
                        for (int tt = 0; tt < 5; ++tt)
#pragma unroll
                            for (int ks = 0; ks < 4; ++ks) {
                                const uint64_t ah = make_desc((((sb) >> 4) | kALo) + (uint32_t)(tt + 128 * blk) + (uint32_t)(2 * ks) * R, kHi);
                                const uint64_t al = make_desc((((sb) >> 4) | kALo) + (uint32_t)(tt + 128 * blk) + (uint32_t)(8 + 2 * ks) * R, kHi);
                                const uint64_t db = make_desc(wl + (tt & 1) * 2048 + 2 * ks * 256, kHi);
                                umma_f16(tmem + 256 * blk, ah, db, i256, 1u);
                                if (v == 20) umma_f16(tmem + 256 * blk + 128, al, db, i128, 1u);
                            }
                    }
                    if (v >= 13 && v <= 16) {  // role swap probe: weights as A, 256 positions (both blocks) as one N = 256 B operand
                        const uint32_t i128x256 = umma_idesc_f16(128, 256), i64x256 = umma_idesc_f16(64, 256), i64x128 = umma_idesc_f16(64, 128);
                        const uint32_t xl = ((sb + OFF_W) >> 4) | (256u << 16);
#pragma unroll
                        for (int ks = 0; ks < 4; ++ks) {
                            const uint64_t aw = make_desc((((sb) >> 4) | kALo) + (uint32_t)(2 * ks) * R, kHi);          // [W_hi; W_lo] 128 rows
                            const uint64_t aw64 = make_desc((((sb) >> 4) | kALo) + (uint32_t)(8 + 2 * ks) * R, kHi);   // W_hi 64 rows
                            const uint64_t bx = make_desc(xl + (t & 1) * 2048 + 2 * ks * 256, kHi);
                            if (v == 13 || v == 15) umma_f16(tmem, aw, bx, i128x256, 1u);
                            if (v == 14 || v == 15) umma_f16(tmem + 256, aw64, bx, i64x256, 1u);
                            if (v == 16) umma_f16(tmem + 256, aw64, bx, i64x128, 1u);
                        }
                    }
                    // single-warp variants (one issuer for both blocks)
                    if (v == 10) {           // a(b0) b(b0) a(b1) b(b1) per k-step
#pragma unroll
                        for (int ks = 0; ks < 4; ++ks)
#pragma unroll
                            for (int bb = 0; bb < 2; ++bb) {
                                umma_f16(tmem + 192 * bb, A(bb, 0, ks), B(ks, 0), ia, 1u);
                                umma_f16(tmem + 192 * bb + 128, A(bb, 1, ks), B(ks, 0), ib, 1u);
                            }
                    } else if (v == 11) {    // a(b0) a(b1) b(b0) b(b1) per k-step
#pragma unroll
                        for (int ks = 0; ks < 4; ++ks) {
                            umma_f16(tmem, A(0, 0, ks), B(ks, 0), ia, 1u);
                            umma_f16(tmem + 192, A(1, 0, ks), B(ks, 0), ia, 1u);
                            umma_f16(tmem + 128, A(0, 1, ks), B(ks, 0), ib, 1u);
                            umma_f16(tmem + 192 + 128, A(1, 1, ks), B(ks, 0), ib, 1u);
                        }
                    } else if (v == 12) {    // same as 11 but overlapping lo columns (today's TMEM map)
#pragma unroll
                        for (int ks = 0; ks < 4; ++ks) {
                            umma_f16(tmem, A(0, 0, ks), B(ks, 0), ia, 1u);
                            umma_f16(tmem + 128, A(1, 0, ks), B(ks, 0), ia, 1u);
                            umma_f16(tmem + 64, A(0, 1, ks), B(ks, 0), ib, 1u);
                            umma_f16(tmem + 192, A(1, 1, ks), B(ks, 0), ib, 1u);
                        }
                    }
                }
                __syncwarp();
                if (++s == STAGES) s = 0;
            }
        }
        if (elect_one()) umma_commit(bar + 16);
        __syncwarp();
        if (!two_warps && elect_one()) mbar_arrive(bar + 16);
        mbar_wait(bar + 16, 0);
        t1 = clock64();
        if (blockIdx.x == 0 && (threadIdx.x & 31) == 0) out[warp] = t1 - t0;
        if (warp == 0 && (threadIdx.x & 31) == 0) mbar_arrive(bar + 24);       // release the pollers
    } else if (warp == 2 && c.tma > 0) {
        // weight-ring traffic: 16 KB bulk copies global -> shared, c.tma in flight, until the MMA warps are done
        volatile uint32_t *flag = &tm + 0;   // unused here; termination by copy count below
        (void)flag;
        const int total = c.reps * 7;
        if ((threadIdx.x & 31) == 0) {
            const uint32_t tb = bar + 32;    // c.tma (<= 4) barriers
            int issued = 0, done = 0;
            uint32_t ph[4] = {0, 0, 0, 0};
            const long long c0 = clock64();
            while (done < total) {
                while (issued < total && issued - done < c.tma) {
                    const int k = issued % c.tma;
                    mbar_arrive_expect_tx(tb + 8 * k, STAGE);
                    bulk_g2s(sb + OFF_W + (issued % STAGES) * STAGE, gsrc + (size_t)(issued % 77) * STAGE, STAGE, tb + 8 * k);
                    ++issued;
                }
                const int k = done % c.tma;
                mbar_wait(tb + 8 * k, ph[k]);
                ph[k] ^= 1u;
                ++done;
            }
            if (blockIdx.x == 0) out[2] = clock64() - c0;
        }
    } else if (warp >= 4 && c.spin) {
        const int lane = threadIdx.x & 31;
        if (c.spin == 1) { while (!mbar_try_wait(bar + 24, 0)) {} }
        else if (c.spin == 2) { while (!try_wait_hint(bar + 24, 0, 10000000u)) {} }
        else if (c.spin == 3) { if (lane == 0) { while (!mbar_try_wait(bar + 24, 0)) {} } __syncwarp(); }
        else { if (lane == 0) { while (!mbar_try_wait(bar + 24, 0)) { __nanosleep(100); } } __syncwarp(); }
    }
    tc_fence_before_sync();
    __syncthreads();
    if (warp == 0) { tc_fence_after_sync(); tmem_dealloc(tmem, 512); }
}

int main() {
    long long *d;
    cudaMalloc(&d, 64);
    cudaFuncSetAttribute(bench, cudaFuncAttributeMaxDynamicSharedMemorySize, SMEM);
    const char *names[] = {"v0 today: hi N128 + lo N64 overlapping cols, 2 warps", "v1 lo in own columns, 2 warps",
                           "v2 4x hi then 4x lo, own columns, 2 warps", "v3 lo reuses hi descriptor (A reuse probe)",
                           "v4 hi only (N128)", "v5 lo only (N64)", "v6 hi only, 2 accumulators alternating", "v7 today without tap shift",
                           "v8 N192 single MMA", "v9 term split: 2 warps on ONE block (hi / lo)", "v10 one issuer: a0 l0 a1 l1", "v11 one issuer: a0 a1 l0 l1 (own cols)",
                           "v12 one issuer: a0 a1 l0 l1 (overlapping cols)", "v13 swap: 4 x (M128 N256) per tap", "v14 swap: 4 x (M64 N256) per tap", "v15 swap: 4 x (M128 N256 + M64 N256) per tap = both blocks", "v16 swap: 4 x (M64 N128) per tap", "", "", "",
                           "v20 contract: one issuer, per 'tap' = one block of 5x4x(N256+N128)", "v21 contract hi only: 5x4xN256"};
    uint8_t *gsrc;
    cudaMalloc(&gsrc, 77 * STAGE);
    cudaMemset(gsrc, 0x3c, 77 * STAGE);
    for (int vv : {0, 13, 14, 15, 16, 21}) {
        const int v = vv % 100 % 30, solo = (vv % 100) >= 30, spin = (vv / 100) % 10, data = (vv / 1000) % 10, reps = 256, tma = vv / 100000;
        Cfg c{vv % 100, reps, data, tma, spin};
        long long h[3] = {0, 0, 0};
        for (int it = 0; it < 2; ++it) {
            cudaMemset(d, 0, 64);
            bench<<<148, 416, SMEM>>>(c, d, gsrc);
            cudaError_t e = cudaDeviceSynchronize();
            if (e != cudaSuccess) { printf("variant %d: %s\n", v, cudaGetErrorString(e)); return 1; }
        }
        cudaMemcpy(h, d, 24, cudaMemcpyDeviceToHost);
        const long long cyc = h[0] > h[1] ? h[0] : h[1];
        printf("%s%-58s tma=%d %8.1f cycles/tap (both blocks); %8.1f cycles per 16 KB copy\n", solo ? "[one warp, one block] " : "", names[v], tma, (double)cyc / (reps * 7), (double)h[2] / (reps * 7));
    }
    return 0;
}
