// Tensor-core kernel of the semi-global block's contracting half (models/stofnet.py:88-91,100-103 of the reference):
//
//     pooled = MaxPool1d(80, 80)( LeakyReLU( contract_conv( relu(conv1(x)) ) ) )        (N, 512, floor(L / 80))
//
// The (N, 512, L) tensor - 34 % of the network's FLOPs and by far its largest intermediate - never exists: the conv is an
// implicit GEMM on tcgen05 whose epilogue max-pools the raw accumulators (LeakyReLU and the per-channel bias are monotone,
// so they are applied once per pooled value).
//
// Weight-stationary schedule: a CTA owns one 128-channel slice of the 512 output channels and keeps that slice's five taps
// (fp16 hi/lo pairs, 160 KB) resident in shared memory for the whole launch; it streams 240-position tiles (3 pooling
// windows = two 128-row MMA blocks) of waveform rows through them.  conv1 (1 -> 64, k9, K = 9: CUDA-core work) is recomputed
// per tile by eight loader warps straight into the UMMA operand layout, so the kernel reads 4 B and writes 6.4 B per
// position and moves no weights after start-up.
//
// fp32 accuracy on fp16 tensor cores as in conv_tc.cu: v = hi + 2^-11 lo;  per K-step
//     [D0 | D1] (+)= X_hi [W_hi ; W_lo]^T   (N = 256),      D1 += X_lo W_hi^T   (N = 128),      result D0 + 2^-11 D1.
//
// Roles: warp 0 = weight loader (bulk copies, once), warp 1 = MMA issuer + TMEM owner, warps 2-9 = conv1 loaders
// (one thread per row and 32-channel half), warps 10-17 = pooling epilogue (warp & 3 = TMEM lane quadrant,
// (warp - 10) >> 2 = 64-channel half).
#include <cuda_fp16.h>
#include <limits.h>

#include "common.cuh"
#include "kernels.cuh"
#include "tc_ptx.cuh"

namespace stof {

using namespace ptx;

constexpr int CT_THREADS = 576;                       // 2 + 8 loader + 8 epilogue warps
constexpr int CT_TILE = 240;                         // positions per tile = 3 pooling windows of 80
constexpr int CT_WINDOW = 80;
constexpr int CT_ROWS0 = 132;                        // operand rows of block 0: positions [-2, 130) of the tile
constexpr int CT_ROWS1 = 116;                        // block 1: positions [126, 242); rows 116..131 are never used
constexpr int CT_TAP_BYTES = 8 * 256 * 16;           // one tap of one 128-channel slice: [c8][W_hi 128 | W_lo 128][8] fp16
constexpr int CT_OFF_W = 0;
constexpr int CT_OFF_A0 = 5 * CT_TAP_BYTES;                      // 163,840
constexpr int CT_OFF_A1 = CT_OFF_A0 + 16 * CT_ROWS0 * 16;        // + 33,792
constexpr int CT_OFF_POOL = CT_OFF_A1 + 16 * CT_ROWS1 * 16;      // + 29,696 ; block 1's MMAs over-read 256 B into here (unused rows)
constexpr int CT_OFF_W1 = CT_OFF_POOL + 3 * 128 * 4;             // conv1 weights [9][64] + bias [64] fp32
constexpr int CT_OFF_BAR = CT_OFF_W1 + (9 * 64 + 64) * 4;
constexpr int CT_NBAR = 9;                                       // w_full, a_full[2], a_empty[2], acc_full[2], acc_empty[2]
constexpr int CT_OFF_TMEM = CT_OFF_BAR + CT_NBAR * 8;
constexpr int CT_SMEM_BYTES = CT_OFF_TMEM + 16;
static_assert(CT_SMEM_BYTES <= 232448, "contract kernel exceeds the 227 KB shared-memory limit");

__device__ __forceinline__ uint32_t ct_cvt_sat_half2(float a, float b) {
    uint32_t r;
    asm("cvt.rn.satfinite.f16x2.f32 %0, %1, %2;" : "=r"(r) : "f"(b), "f"(a));
    return r;
}
__device__ __forceinline__ void ct_split_pair(float v0, float v1, uint32_t &hi, uint32_t &lo) {
    hi = ct_cvt_sat_half2(v0, v1);
    const float2 h = __half22float2(*reinterpret_cast<const __half2 *>(&hi));
    lo = ct_cvt_sat_half2((v0 - h.x) * 2048.0f, (v1 - h.y) * 2048.0f);
}
// order-preserving float <-> int32 map (so the pooling can use integer warp reductions and atomics)
__device__ __forceinline__ int float_key(float v) {
    const int b = __float_as_int(v);
    return b ^ ((b >> 31) & 0x7fffffff);
}
__device__ __forceinline__ float key_float(int k) { return __int_as_float(k ^ ((k >> 31) & 0x7fffffff)); }

// Warp "transpose-reduce": every lane holds 32 column values of its own row; afterwards lane j holds the maximum over the
// rows (lanes) of its segment for column j (FULL: one 32-lane segment; otherwise two 16-lane segments and lane j holds
// columns 2 * (j & 15) + {0, 1} in v[0], v[1]).  log-step butterfly: 31 (30) shuffles instead of 32 warp reductions.
template <bool FULL>
__device__ __forceinline__ void transpose_max(float (&v)[32], int lane) {
    if (FULL) {
        const bool up = lane & 16;
#pragma unroll
        for (int i = 0; i < 16; ++i) {
            const float send = up ? v[i] : v[i + 16], keep = up ? v[i + 16] : v[i];
            v[i] = fmaxf(keep, __shfl_xor_sync(0xffffffffu, send, 16));
        }
    }
    constexpr int N0 = FULL ? 16 : 32;
    {
        const bool up = lane & 8;
#pragma unroll
        for (int i = 0; i < N0 / 2; ++i) {
            const float send = up ? v[i] : v[i + N0 / 2], keep = up ? v[i + N0 / 2] : v[i];
            v[i] = fmaxf(keep, __shfl_xor_sync(0xffffffffu, send, 8));
        }
    }
    {
        const bool up = lane & 4;
#pragma unroll
        for (int i = 0; i < N0 / 4; ++i) {
            const float send = up ? v[i] : v[i + N0 / 4], keep = up ? v[i + N0 / 4] : v[i];
            v[i] = fmaxf(keep, __shfl_xor_sync(0xffffffffu, send, 4));
        }
    }
    {
        const bool up = lane & 2;
#pragma unroll
        for (int i = 0; i < N0 / 8; ++i) {
            const float send = up ? v[i] : v[i + N0 / 8], keep = up ? v[i + N0 / 8] : v[i];
            v[i] = fmaxf(keep, __shfl_xor_sync(0xffffffffu, send, 2));
        }
    }
    {
        const bool up = lane & 1;
#pragma unroll
        for (int i = 0; i < N0 / 16; ++i) {
            const float send = up ? v[i] : v[i + N0 / 16], keep = up ? v[i + N0 / 16] : v[i];
            v[i] = fmaxf(keep, __shfl_xor_sync(0xffffffffu, send, 1));
        }
    }
}

struct ContractParams {
    const float *x;          // (N, L) fp32
    const void *w;           // [4 slices][5 taps][8 c8][256][8] fp16
    const float *bias;       // [512]
    const float *w1;         // conv1 weights [9][64] fp32
    const float *b1;         // [64]
    float *pooled;           // (N, 512, P)
    int N, L, P;
    int tiles_per_row, total_tiles;
    int fast;                // 1: one-pass fp16 mode (X_hi W_hi only)
};

__global__ void __launch_bounds__(CT_THREADS, 1) contract_tc_kernel(const ContractParams p) {
    extern __shared__ __align__(1024) uint8_t smem[];
    const uint32_t sbase = smem_u32(smem);
    const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
    const uint32_t bar0 = sbase + CT_OFF_BAR;
    const uint32_t BAR_W_FULL = bar0;
    auto BAR_A_FULL = [&](int b) { return bar0 + 8 * (1 + b); };
    auto BAR_A_EMPTY = [&](int b) { return bar0 + 8 * (3 + b); };
    auto BAR_ACC_FULL = [&](int b) { return bar0 + 8 * (5 + b); };
    auto BAR_ACC_EMPTY = [&](int b) { return bar0 + 8 * (7 + b); };
    int *s_pool = reinterpret_cast<int *>(smem + CT_OFF_POOL);            // [3 windows][128 channels] keys
    float *s_w1 = reinterpret_cast<float *>(smem + CT_OFF_W1);            // [9][64] then bias [64]
    volatile uint32_t *s_tmem = reinterpret_cast<volatile uint32_t *>(smem + CT_OFF_TMEM);

    const int slice = blockIdx.x & 3;                 // 128-channel slice of the 512 output channels
    const int worker = blockIdx.x >> 2;               // index among the CTAs that share this slice
    const int workers = (gridDim.x - slice + 3) >> 2;

    for (int i = threadIdx.x; i < 9 * 64; i += CT_THREADS) s_w1[i] = __ldg(p.w1 + i);
    for (int i = threadIdx.x; i < 64; i += CT_THREADS) s_w1[9 * 64 + i] = __ldg(p.b1 + i);
    for (int i = threadIdx.x; i < 3 * 128; i += CT_THREADS) s_pool[i] = INT_MIN;
    if (threadIdx.x == 0) {
        mbar_init(BAR_W_FULL, 1);
        for (int b = 0; b < 2; ++b) {
            mbar_init(BAR_A_FULL(b), 8);
            mbar_init(BAR_A_EMPTY(b), 1);
            mbar_init(BAR_ACC_FULL(b), 1);
            mbar_init(BAR_ACC_EMPTY(b), 8);
        }
        fence_mbar_init();
    }
    if (warp == 1) tmem_alloc(sbase + CT_OFF_TMEM, 512);
    tc_fence_before_sync();
    __syncthreads();
    tc_fence_after_sync();
    const uint32_t tmem0 = *s_tmem;

    if (warp == 0) {
        // =================================== weight loader (once) ======================================
        if (lane == 0) {
            mbar_arrive_expect_tx(BAR_W_FULL, 5u * CT_TAP_BYTES);
            const uint8_t *src = reinterpret_cast<const uint8_t *>(p.w) + (size_t)slice * 5 * CT_TAP_BYTES;
            for (int t = 0; t < 5; ++t) bulk_g2s(sbase + CT_OFF_W + t * CT_TAP_BYTES, src + (size_t)t * CT_TAP_BYTES, CT_TAP_BYTES, BAR_W_FULL);
        }
    } else if (warp == 1) {
        // =================================== MMA issuer ==================================================
        constexpr uint32_t kDescHi = (128u >> 4) | (1u << 14);                  // SBO = 128 B, descriptor version 1
        constexpr uint32_t kIdescA = umma_idesc_f16(128, 256), kIdescB = umma_idesc_f16(128, 128);
        const uint32_t w_lo0 = ((sbase + CT_OFF_W) >> 4) | (256u << 16);        // LBO = 256 rows * 16 B
        mbar_wait(BAR_W_FULL, 0);
        int it = 0;
        for (int tile = worker; tile < p.total_tiles; tile += workers, ++it) {
#pragma unroll 1
            for (int b = 0; b < 2; ++b) {
                const uint32_t rows = b ? CT_ROWS1 : CT_ROWS0;
                const uint32_t a_lo0 = ((sbase + (b ? CT_OFF_A1 : CT_OFF_A0)) >> 4) | (rows << 16);
                const uint32_t d_tmem = tmem0 + 256u * b;
                mbar_wait(BAR_A_FULL(b), (uint32_t)(it & 1));
                mbar_wait(BAR_ACC_EMPTY(b), (uint32_t)(it & 1) ^ 1u);
                tc_fence_after_sync();
                if (elect_one()) {
#pragma unroll
                    for (int t = 0; t < 5; ++t) {
#pragma unroll
                        for (int ks = 0; ks < 4; ++ks) {
                            const uint64_t da_hi = make_desc(a_lo0 + t + 2 * ks * rows, kDescHi);
                            const uint64_t da_lo = make_desc(a_lo0 + t + (8 + 2 * ks) * rows, kDescHi);
                            const uint64_t db = make_desc(w_lo0 + t * (CT_TAP_BYTES >> 4) + 2 * ks * 256, kDescHi);
                            if (p.fast) {
                                umma_f16(d_tmem, da_hi, db, kIdescB, (t > 0 || ks > 0) ? 1u : 0u);      // N = 128: W_hi rows only
                            } else {
                                umma_f16(d_tmem, da_hi, db, kIdescA, (t > 0 || ks > 0) ? 1u : 0u);
                                umma_f16(d_tmem + 128, da_lo, db, kIdescB, 1u);
                            }
                        }
                    }
                    umma_commit(BAR_A_EMPTY(b));
                    umma_commit(BAR_ACC_FULL(b));
                }
                __syncwarp();
            }
        }
    } else if (warp < 10) {
        // =================================== conv1 loaders ===============================================
        // x0[l][c] = relu(b1[c] + sum_t w1[c][t] x[l + t - 4]) for 0 <= l < L, 0 outside (the contract conv's zero padding).
        // Work item = (row, 32-channel half); 256 threads.
        const int tid = threadIdx.x - 64;               // 0..255
        const float4 *w4 = reinterpret_cast<const float4 *>(s_w1);
        int it = 0;
        for (int tile = worker; tile < p.total_tiles; tile += workers, ++it) {
            const int n = tile / p.tiles_per_row;
            const int l_tile = (tile - n * p.tiles_per_row) * CT_TILE;
            const float *xr = p.x + (size_t)n * p.L;
#pragma unroll 1
            for (int b = 0; b < 2; ++b) {
                const int rows = b ? CT_ROWS1 : CT_ROWS0;
                uint8_t *abuf = smem + (b ? CT_OFF_A1 : CT_OFF_A0);
                mbar_wait(BAR_A_EMPTY(b), (uint32_t)(it & 1) ^ 1u);
                for (int item = tid; item < 2 * rows; item += 256) {
                    const int half = item >= rows ? 1 : 0;
                    const int row = item - half * rows;
                    const int l = l_tile + 128 * b - 2 + row;
                    const bool inside = (l >= 0) && (l < p.L);
                    float xv[9];
#pragma unroll
                    for (int t = 0; t < 9; ++t) {
                        const int g = l + t - 4;
                        xv[t] = (inside && g >= 0 && g < p.L) ? __ldg(xr + g) : 0.0f;
                    }
#pragma unroll 2
                    for (int j = 0; j < 4; ++j) {
                        const int c8 = 4 * half + j;
                        float v[8];
#pragma unroll
                        for (int h = 0; h < 2; ++h) {
                            float4 a = w4[9 * 16 + 2 * c8 + h];
#pragma unroll
                            for (int t = 0; t < 9; ++t) {
                                const float4 wv = w4[t * 16 + 2 * c8 + h];
                                a.x = fmaf(wv.x, xv[t], a.x); a.y = fmaf(wv.y, xv[t], a.y);
                                a.z = fmaf(wv.z, xv[t], a.z); a.w = fmaf(wv.w, xv[t], a.w);
                            }
                            v[4 * h + 0] = inside ? fmaxf(a.x, 0.f) : 0.f; v[4 * h + 1] = inside ? fmaxf(a.y, 0.f) : 0.f;
                            v[4 * h + 2] = inside ? fmaxf(a.z, 0.f) : 0.f; v[4 * h + 3] = inside ? fmaxf(a.w, 0.f) : 0.f;
                        }
                        uint4 uh, ul;
                        ct_split_pair(v[0], v[1], uh.x, ul.x);
                        ct_split_pair(v[2], v[3], uh.y, ul.y);
                        ct_split_pair(v[4], v[5], uh.z, ul.z);
                        ct_split_pair(v[6], v[7], uh.w, ul.w);
                        *reinterpret_cast<uint4 *>(abuf + (c8 * rows + row) * 16) = uh;
                        if (!p.fast) *reinterpret_cast<uint4 *>(abuf + ((8 + c8) * rows + row) * 16) = ul;
                    }
                }
                fence_proxy_async_smem();
                __syncwarp();
                if (lane == 0) mbar_arrive(BAR_A_FULL(b));
            }
        }
    } else {
        // =================================== pooling epilogue ============================================
        const int quad = warp & 3;
        const int grp = (warp - 10) >> 2;                // channel half: co [64 grp, +64) of the slice
        const int etid = threadIdx.x - 320;              // 0..255
        const uint32_t lane_addr = (uint32_t)(32 * quad) << 16;
        int it = 0;
        for (int tile = worker; tile < p.total_tiles; tile += workers, ++it) {
            const int n = tile / p.tiles_per_row;
            const int ti = tile - n * p.tiles_per_row;
            const int p0 = 3 * ti;                        // first pooling window of the tile
#pragma unroll 1
            for (int b = 0; b < 2; ++b) {
                const int row = 128 * b + 32 * quad + lane;             // row within the tile
                const int win = row / CT_WINDOW;                        // 0..3 (3 = beyond the tile)
                const bool valid = (row < CT_TILE) && (p0 + win < p.P);
                // the only warp whose rows belong to two pooling windows of the tile: rows 64..95 (split at lane 16)
                const bool split = (b == 0) && (quad == 2);
                mbar_wait(BAR_ACC_FULL(b), (uint32_t)(it & 1));
                tc_fence_after_sync();
                const uint32_t d_tmem = tmem0 + 256u * b + lane_addr + 64u * grp;
#pragma unroll 1
                for (int c32 = 0; c32 < 2; ++c32) {
                    uint32_t d0[32], d1[32];
                    tmem_ld16(d_tmem + 32 * c32, *reinterpret_cast<uint32_t(*)[16]>(&d0[0]));
                    tmem_ld16(d_tmem + 32 * c32 + 16, *reinterpret_cast<uint32_t(*)[16]>(&d0[16]));
                    if (!p.fast) {
                        tmem_ld16(d_tmem + 128 + 32 * c32, *reinterpret_cast<uint32_t(*)[16]>(&d1[0]));
                        tmem_ld16(d_tmem + 128 + 32 * c32 + 16, *reinterpret_cast<uint32_t(*)[16]>(&d1[16]));
                    } else {
#pragma unroll
                        for (int i = 0; i < 32; ++i) d1[i] = 0u;
                    }
                    tmem_ld_wait();
                    if (c32 == 1) {
                        tc_fence_before_sync();
                        __syncwarp();
                        if (lane == 0) mbar_arrive(BAR_ACC_EMPTY(b));
                    }
                    float v[32];
#pragma unroll
                    for (int i = 0; i < 32; ++i) {
                        const float t = __uint_as_float(d0[i]) + (1.0f / 2048.0f) * __uint_as_float(d1[i]);
                        v[i] = valid ? t : -INFINITY;
                    }
                    const int co0 = 64 * grp + 32 * c32;
                    if (!split) {
                        transpose_max<true>(v, lane);                   // lane j: max over the warp's rows of column j
                        const int w = (128 * b + 32 * quad) / CT_WINDOW;
                        if (v[0] != -INFINITY) atomicMax(&s_pool[w * 128 + co0 + lane], float_key(v[0]));
                    } else {
                        transpose_max<false>(v, lane);                  // lane j: columns 2 (j & 15) + {0, 1} of its 16-row half
                        const int w = lane >> 4;                        // rows 64..79 -> window 0, rows 80..95 -> window 1
                        const int c = 2 * (lane & 15);
                        if (v[0] != -INFINITY) atomicMax(&s_pool[w * 128 + co0 + c], float_key(v[0]));
                        if (v[1] != -INFINITY) atomicMax(&s_pool[w * 128 + co0 + c + 1], float_key(v[1]));
                    }
                }
            }
            // all 8 epilogue warps have folded both blocks: publish bias + LeakyReLU of the window maxima, reset the pool
            asm volatile("bar.sync 1, 256;" ::: "memory");
            for (int idx = etid; idx < 3 * 128; idx += 256) {
                const int w = idx >> 7, co = idx & 127;
                const int key = s_pool[idx];
                s_pool[idx] = INT_MIN;
                if (p0 + w < p.P) {
                    const float v = key_float(key) + __ldg(p.bias + slice * 128 + co);
                    p.pooled[((size_t)n * 512 + slice * 128 + co) * p.P + p0 + w] = leaky(v);
                }
            }
            asm volatile("bar.sync 1, 256;" ::: "memory");
        }
    }

    tc_fence_before_sync();
    __syncthreads();
    if (warp == 1) {
        tc_fence_after_sync();
        tmem_dealloc(tmem0, 512);
    }
}

int launch_contract_pool_tc(const float *x, const void *w, const float *bias, const float *w1, const float *b1, float *pooled,
                            int N, int L, int P, cudaStream_t st, int fast) {
    static bool attr = false;
    if (!attr) {
        STOF_CUDA(cudaFuncSetAttribute(contract_tc_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, CT_SMEM_BYTES));
        attr = true;
    }
    ContractParams p;
    p.x = x; p.w = w; p.bias = bias; p.w1 = w1; p.b1 = b1; p.pooled = pooled;
    p.N = N; p.L = L; p.P = P; p.fast = fast;
    p.tiles_per_row = (P + 2) / 3;
    const long long total = (long long)N * p.tiles_per_row;
    STOF_REQUIRE(total < (1ll << 31), STOF_E_INVALID, "contract_tc: too many tiles");
    p.total_tiles = (int)total;
    // 4 slices x workers; every slice needs at least one CTA
    long long grid = 4 * total;
    if (grid > kSMs) grid = kSMs;
    contract_tc_kernel<<<(int)grid, CT_THREADS, CT_SMEM_BYTES, st>>>(p);
    STOF_LAUNCH_CHECK();
    return STOF_OK;
}

}  // namespace stof
