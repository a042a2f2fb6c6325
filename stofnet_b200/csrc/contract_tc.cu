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
// Orientation: the WEIGHTS are the MMA's A operand (M = the slice's 128 output channels = TMEM lanes) and the positions are
// its N dimension (TMEM columns), so that max-pooling over positions is a running maximum over the columns a thread reads
// from its own lane - no cross-lane reduction, no shared-memory atomics (the positions-on-lanes form spent more issue slots
// on its butterfly transpose-max than the tensor pipe needed for the MMAs: the epilogue warps paced the kernel).
//
// fp32 accuracy on fp16 tensor cores as in conv_tc.cu: v = hi + 2^-11 lo;  per K-step three M = 128 MMAs
//     D0 (+)= W_hi X_hi^T,      D1 (+)= W_lo X_hi^T,      D1 += W_hi X_lo^T,      result D0 + 2^-11 D1.
// A tile of 240 positions is two accumulator halves: positions [0, 128) (N = 128) and [128, 240) (N = 112); pooling window 1
// (positions 80..159) straddles them and is carried in a register.
//
// Roles: warp 0 = weight loader (bulk copies, once), warp 1 = MMA issuer + TMEM owner, warps 2-9 = conv1 loaders
// (one thread per row and 32-channel half), warps 10-13 = pooling epilogue (warp & 3 = TMEM lane quadrant = 32 channels).
#include <cuda_fp16.h>
#include <limits.h>

#include "common.cuh"
#include "kernels.cuh"
#include "tc_ptx.cuh"

namespace stof {

using namespace ptx;

constexpr int CT_THREADS = 448;                       // 2 + 8 loader + 4 epilogue warps
constexpr int CT_TILE = 240;                         // positions per tile = 3 pooling windows of 80
constexpr int CT_WINDOW = 80;
constexpr int CT_ROWS0 = 132;                        // operand rows of block 0: positions [-2, 130) of the tile
constexpr int CT_ROWS1 = 116;                        // block 1: positions [126, 242); rows 116..131 are never used
constexpr int CT_TAP_BYTES = 8 * 256 * 16;           // one tap of one 128-channel slice: [c8][W_hi 128 | W_lo 128][8] fp16
constexpr int CT_OFF_W = 0;
constexpr int CT_OFF_A0 = 5 * CT_TAP_BYTES;                      // 163,840
constexpr int CT_OFF_A1 = CT_OFF_A0 + 16 * CT_ROWS0 * 16;        // + 33,792
constexpr int CT_OFF_POOL = CT_OFF_A1 + 16 * CT_ROWS1 * 16;      // + 29,696 ; 1.5 KB spare (guard behind the operand buffers)
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
    // (v - h) * 2^11 with packed fp32 ops; both steps are exact, so this equals the scalar form bit for bit
    const float2 d = __fmul2_rn(__ffma2_rn(h, make_float2(-1.0f, -1.0f), make_float2(v0, v1)), make_float2(2048.0f, 2048.0f));
    lo = ct_cvt_sat_half2(d.x, d.y);
}
struct ContractParams {
    const float *x;          // (N, L) fp32
    const void *w;           // [4 slices][5 taps][8 c8][256][8] fp16
    const float *bias;       // [512]
    const float *w1;         // conv1 weights [9][64] fp32
    const float *b1;         // [64]
    float *pooled;           // (N, 512, P)
    int *argmax;             // (N, 512, P) position of each pooled maximum (training forward), or NULL
    int N, L, P;
    int tiles_per_row, total_tiles;
    int fast;                // 1: one-pass fp16 mode (X_hi W_hi only)
};

// ARGMAX: also record the position of every pooled maximum (first one on ties, as ATen's max_pool1d) for the backward pass
template <bool ARGMAX>
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
    float *s_w1 = reinterpret_cast<float *>(smem + CT_OFF_W1);            // [9][64] then bias [64]
    volatile uint32_t *s_tmem = reinterpret_cast<volatile uint32_t *>(smem + CT_OFF_TMEM);

    const int slice = blockIdx.x & 3;                 // 128-channel slice of the 512 output channels
    const int worker = blockIdx.x >> 2;               // index among the CTAs that share this slice
    const int workers = (gridDim.x - slice + 3) >> 2;

    for (int i = threadIdx.x; i < 9 * 64; i += CT_THREADS) s_w1[i] = __ldg(p.w1 + i);
    for (int i = threadIdx.x; i < 64; i += CT_THREADS) s_w1[9 * 64 + i] = __ldg(p.b1 + i);
    if (threadIdx.x == 0) {
        mbar_init(BAR_W_FULL, 1);
        for (int b = 0; b < 2; ++b) {
            mbar_init(BAR_A_FULL(b), 8);
            mbar_init(BAR_A_EMPTY(b), 1);
            mbar_init(BAR_ACC_FULL(b), 1);
            mbar_init(BAR_ACC_EMPTY(b), 4);
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
        const uint32_t w_lo0 = ((sbase + CT_OFF_W) >> 4) | (256u << 16);        // weights (A): LBO = 256 rows * 16 B; W_hi rows 0-127, W_lo 128-255
        mbar_wait(BAR_W_FULL, 0);
        int it = 0;
        for (int tile = worker; tile < p.total_tiles; tile += workers, ++it) {
#pragma unroll 1
            for (int b = 0; b < 2; ++b) {
                const uint32_t rows = b ? CT_ROWS1 : CT_ROWS0;
                const uint32_t idesc = b ? umma_idesc_f16(128, 112) : umma_idesc_f16(128, 128);   // N = positions of this half
                const uint32_t x_lo0 = ((sbase + (b ? CT_OFF_A1 : CT_OFF_A0)) >> 4) | (rows << 16);
                const uint32_t d0 = tmem0 + 256u * b, d1 = d0 + 128u;
                mbar_wait(BAR_A_FULL(b), (uint32_t)(it & 1));
                mbar_wait(BAR_ACC_EMPTY(b), (uint32_t)(it & 1) ^ 1u);
                tc_fence_after_sync();
                if (elect_one()) {
#pragma unroll
                    for (int t = 0; t < 5; ++t) {
#pragma unroll
                        for (int ks = 0; ks < 4; ++ks) {
                            const uint64_t dx_hi = make_desc(x_lo0 + t + 2 * ks * rows, kDescHi);
                            const uint64_t dx_lo = make_desc(x_lo0 + t + (8 + 2 * ks) * rows, kDescHi);
                            const uint64_t dw_hi = make_desc(w_lo0 + t * (CT_TAP_BYTES >> 4) + 2 * ks * 256, kDescHi);
                            const uint64_t dw_lo = make_desc(w_lo0 + t * (CT_TAP_BYTES >> 4) + 2 * ks * 256 + 128, kDescHi);
                            const uint32_t acc = (t > 0 || ks > 0) ? 1u : 0u;
                            umma_f16(d0, dw_hi, dx_hi, idesc, acc);
                            if (!p.fast) {
                                umma_f16(d1, dw_lo, dx_hi, idesc, acc);
                                umma_f16(d1, dw_hi, dx_lo, idesc, 1u);
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
        // Warp w owns the 8-channel chunk c8 = w and keeps its 72 conv1 weights + 8 biases in registers for the whole launch;
        // lane = row within a group of 32 (conflict-free 16-byte operand stores).  A block's waveform samples are fetched once
        // per lane up front (6 coalesced loads issued before the buffer wait: one memory latency per block, overlapped with
        // the MMAs still reading the buffer) and the 9-tap windows are assembled with warp shuffles; the FMAs are packed
        // (fma.rn.f32x2, bit-identical to fmaf).  No shared-memory reads at all in this role.
        const int c8 = warp - 2;                        // 0..7
        float2 wr[9][4], br[4];
        {
            const float4 *w4 = reinterpret_cast<const float4 *>(s_w1) + 2 * c8;
#pragma unroll
            for (int h = 0; h < 2; ++h) {
                const float4 bq = w4[9 * 16 + h];
                br[2 * h] = make_float2(bq.x, bq.y); br[2 * h + 1] = make_float2(bq.z, bq.w);
#pragma unroll
                for (int t = 0; t < 9; ++t) {
                    const float4 wv = w4[t * 16 + h];
                    wr[t][2 * h] = make_float2(wv.x, wv.y); wr[t][2 * h + 1] = make_float2(wv.z, wv.w);
                }
            }
        }
        int it = 0;
        for (int tile = worker; tile < p.total_tiles; tile += workers, ++it) {
            const int n = tile / p.tiles_per_row;
            const int l_tile = (tile - n * p.tiles_per_row) * CT_TILE;
            const float *xr = p.x + (size_t)n * p.L;
#pragma unroll 1
            for (int b = 0; b < 2; ++b) {
                const int rows = b ? CT_ROWS1 : CT_ROWS0;
                uint8_t *abuf = smem + (b ? CT_OFF_A1 : CT_OFF_A0) + (size_t)c8 * rows * 16;
                const int l_base = l_tile + 128 * b - 2;                 // position of operand row 0
                float xg[6];                                             // xg[k] = x[l_base - 4 + 32 k + lane], 0 outside the row
#pragma unroll
                for (int k = 0; k < 6; ++k) {
                    const int g = l_base - 4 + 32 * k + lane;
                    xg[k] = (g >= 0 && g < p.L) ? __ldg(xr + g) : 0.0f;
                }
                mbar_wait(BAR_A_EMPTY(b), (uint32_t)(it & 1) ^ 1u);
#pragma unroll
                for (int k = 0; k < 5; ++k) {
                    const int row = 32 * k + lane;
                    const int l = l_base + row;
                    const bool inside = (l >= 0) && (l < p.L);
                    float2 acc[4] = {br[0], br[1], br[2], br[3]};
#pragma unroll
                    for (int t = 0; t < 9; ++t) {
                        const float lo = __shfl_sync(0xffffffffu, xg[k], (lane + t) & 31);
                        const float hi = __shfl_sync(0xffffffffu, xg[k + 1], (lane + t) & 31);
                        const float xv = (lane + t < 32) ? lo : hi;
                        const float2 xx = make_float2(xv, xv);
#pragma unroll
                        for (int q = 0; q < 4; ++q) acc[q] = __ffma2_rn(wr[t][q], xx, acc[q]);
                    }
                    if (row < rows) {
                        float v[8];
#pragma unroll
                        for (int q = 0; q < 4; ++q) {
                            v[2 * q] = inside ? fmaxf(acc[q].x, 0.f) : 0.f;
                            v[2 * q + 1] = inside ? fmaxf(acc[q].y, 0.f) : 0.f;
                        }
                        uint4 uh, ul;
                        ct_split_pair(v[0], v[1], uh.x, ul.x);
                        ct_split_pair(v[2], v[3], uh.y, ul.y);
                        ct_split_pair(v[4], v[5], uh.z, ul.z);
                        ct_split_pair(v[6], v[7], uh.w, ul.w);
                        *reinterpret_cast<uint4 *>(abuf + row * 16) = uh;
                        if (!p.fast) *reinterpret_cast<uint4 *>(abuf + (size_t)8 * rows * 16 + row * 16) = ul;
                    }
                }
                fence_proxy_async_smem();
                __syncwarp();
                if (lane == 0) mbar_arrive(BAR_A_FULL(b));
            }
        }
    } else {
        // =================================== pooling epilogue ============================================
        // thread = output channel (TMEM lane) of the slice; a tile's three pooling windows are running maxima over the columns
        const int quad = warp & 3;
        const int co = slice * 128 + 32 * quad + lane;
        const uint32_t lane_addr = (uint32_t)(32 * quad) << 16;
        const float bias = __ldg(p.bias + co);
        const float li = p.fast ? 0.0f : (1.0f / 2048.0f);
        int it = 0;
        for (int tile = worker; tile < p.total_tiles; tile += workers, ++it) {
            const int n = tile / p.tiles_per_row;
            const int ti = tile - n * p.tiles_per_row;
            const int p0 = 3 * ti;                        // first pooling window of the tile
            float m[3] = {-INFINITY, -INFINITY, -INFINITY};
            int mi[3] = {0, 0, 0};
#pragma unroll
            for (int b = 0; b < 2; ++b) {
                mbar_wait(BAR_ACC_FULL(b), (uint32_t)(it & 1));
                tc_fence_after_sync();
                const uint32_t d_tmem = tmem0 + 256u * b + lane_addr;
                const int nch = (b == 0) ? 8 : 7;         // 16-column chunks: positions [0, 128) and [128, 240)
#pragma unroll
                for (int c = 0; c < 8; c += 2) {
                    uint32_t d0[32], d1[32];
                    const bool two = (c + 1 < nch);
                    tmem_ld16(d_tmem + 16 * c, *reinterpret_cast<uint32_t(*)[16]>(&d0[0]));
                    if (two) tmem_ld16(d_tmem + 16 * c + 16, *reinterpret_cast<uint32_t(*)[16]>(&d0[16]));
                    if (!p.fast) {
                        tmem_ld16(d_tmem + 128 + 16 * c, *reinterpret_cast<uint32_t(*)[16]>(&d1[0]));
                        if (two) tmem_ld16(d_tmem + 128 + 16 * c + 16, *reinterpret_cast<uint32_t(*)[16]>(&d1[16]));
                    } else {
#pragma unroll
                        for (int i = 0; i < 32; ++i) d1[i] = 0u;
                    }
                    tmem_ld_wait();
#pragma unroll
                    for (int h = 0; h < 2; ++h) {
                        if (h == 0 || two) {
                            const int w = (128 * b + 16 * (c + h)) / CT_WINDOW;  // compile-time: window of this 16-column chunk
                            float mm = m[w];
                            if (ARGMAX) {
                                int ii = mi[w];
#pragma unroll
                                for (int i = 0; i < 16; ++i) {
                                    const float v = fmaf(li, __uint_as_float(d1[16 * h + i]), __uint_as_float(d0[16 * h + i]));
                                    if (v > mm) { mm = v; ii = 128 * b + 16 * (c + h) + i; }
                                }
                                mi[w] = ii;
                            } else {
#pragma unroll
                                for (int i = 0; i < 16; ++i)
                                    mm = fmaxf(mm, fmaf(li, __uint_as_float(d1[16 * h + i]), __uint_as_float(d0[16 * h + i])));
                            }
                            m[w] = mm;
                        }
                    }
                }
                tc_fence_before_sync();
                __syncwarp();
                if (lane == 0) mbar_arrive(BAR_ACC_EMPTY(b));
            }
            // bias and LeakyReLU are monotone: applied once per pooled value (models/stofnet.py:89-91)
            float *dst = p.pooled + ((size_t)n * 512 + co) * p.P + p0;
#pragma unroll
            for (int w = 0; w < 3; ++w)
                if (p0 + w < p.P) {
                    dst[w] = leaky(m[w] + bias);
                    if (ARGMAX) p.argmax[((size_t)n * 512 + co) * p.P + p0 + w] = CT_TILE * ti + mi[w];
                }
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
                            int N, int L, int P, cudaStream_t st, int fast, int *argmax) {
    static SmemAttrOnce attr, attr_arg;
    { const int rc = argmax ? ensure_dyn_smem(contract_tc_kernel<true>, CT_SMEM_BYTES, attr_arg)
                            : ensure_dyn_smem(contract_tc_kernel<false>, CT_SMEM_BYTES, attr); if (rc) return rc; }
    ContractParams p;
    p.x = x; p.w = w; p.bias = bias; p.w1 = w1; p.b1 = b1; p.pooled = pooled; p.argmax = argmax;
    p.N = N; p.L = L; p.P = P; p.fast = fast;
    p.tiles_per_row = (P + 2) / 3;
    const long long total = (long long)N * p.tiles_per_row;
    STOF_REQUIRE(total < (1ll << 31), STOF_E_INVALID, "contract_tc: too many tiles");
    p.total_tiles = (int)total;
    // 4 slices x workers; every slice needs at least one CTA
    long long grid = 4 * total;
    if (grid > kSMs) grid = kSMs;
    if (argmax) contract_tc_kernel<true><<<(int)grid, CT_THREADS, CT_SMEM_BYTES, st>>>(p);
    else contract_tc_kernel<false><<<(int)grid, CT_THREADS, CT_SMEM_BYTES, st>>>(p);
    STOF_LAUNCH_CHECK();
    return STOF_OK;
}

}  // namespace stof
