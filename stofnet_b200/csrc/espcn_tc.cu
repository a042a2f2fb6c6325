// ESPCN_1D forward (models/espcn_1d.py:8-36; SURVEY.md 8f row 1, BASELINE config 2's wobbly-sponge) with conv2 on tcgen05:
//   conv1 (1 -> 64, k5, pad 2) + tanh  ->  conv2 (64 -> 32, k3, pad 1) + tanh  ->  conv3 (32 -> 4, k3, pad 1)
//   -> SampleShuffle1D(4) -> sigmoid
// conv2 is 6,144 of the model's 6,848 MACs per position: a dense contraction (K = 3 taps x 64 channels, N = 32) that the fused
// fp32 kernel (espcn_fused.cu) runs on the FMA pipe.  Here it is an implicit GEMM on the tensor cores, exactly like the chain
// kernel's layers (conv_tc.cu): M = 128 positions (TMEM lanes), conv1's output written by the producer threads straight into
// shared memory as split fp16 planes [plane hi/lo][c8][row][8 channels] (the no-swizzle K-major UMMA operand layout, a tap
// shift = a 16-byte shift of the descriptor), fp32-faithful through the hi/lo split
//          [D0 | D1] (+)= H1_hi * [W_hi ; W_lo]^T   (N = 64)          D1 += H1_lo * W_hi^T   (N = 32)
// with conv2's weights (24 KB as packed fp16 taps) resident in shared memory for the whole launch.  conv3 (32 -> 4, k3) follows
// the same way (M = 128, N = 16 like the chain kernel's conv_last: rows 0-3 W_hi, 8-11 W_lo, 16-19 W_hi of a 32-row tap), fed by
// the conv2 epilogue's split planes of h2.  conv1 + tanh (producer), bias + tanh (conv2 epilogue), bias + shuffle + sigmoid
// (conv3 epilogue) stay on the CUDA cores; HBM traffic is 4 B in + 16 B out per position.
//
// One persistent CTA per SM: 16 worker warps + 3 MMA-issuing warps, synchronised by mbarriers only.  Per tile of 124 output
// positions the workers
//   (1) drain tile k's conv2 accumulator (TMEM -> + bias, tanh -> h2 split planes in shared memory),
//   (2) produce tile k+2's conv1 operand (two (row, 8-channel) items per thread) into the A buffer tile k used, and publish both
//       with one proxy fence (conv3 of tile k and conv2 of tile k+2 may start),
//   (3) one warp per TMEM quadrant (rotating per tile) drains tile k-1's conv3 accumulator: sigmoid + the shuffled 16-byte stores,
// while the tensor pipe works on conv2 of tile k+1 and conv3 of tile k (everything double-buffered: A, h2, both accumulators).
// tanh / sigmoid use ex2.approx + rcp.approx forms on pre-scaled arguments (absolute error < 2e-7); the fp32 FFMA kernels
// (STOF_PATH_SIMT) stay as the cross-check (tests/test_gpu_parity.py).
#include <cuda_fp16.h>

#include "conv_tc_shared.cuh"

namespace stof {

using namespace ptx;

constexpr int ET_WORKER_WARPS = 16;
constexpr int ET_WORKERS = ET_WORKER_WARPS * 32;       // 512
constexpr int ET_THREADS = ET_WORKERS + 3 * 32;        // + three MMA-issuing warps (conv2 even / odd K-steps, conv3)
constexpr int ET_OUT = 124;                            // output positions per tile
constexpr int ET_R = 136;                              // rows per (plane, c8) slab of an A buffer (128 produced + tap overhang)
constexpr int ET_SLAB = ET_R * 16;
constexpr int ET_PLANE = 8 * ET_SLAB;
constexpr int ET_ABUF = 2 * ET_PLANE;                  // 34,816 B
constexpr int ET_W2_TAP = 8 * 64 * 16;                 // one tap: [8 c8][W_hi 32 | W_lo 32][8] fp16 = 8 KB
constexpr int ET_H2BUF = 2 * 4 * ET_SLAB;               // h2 split planes [plane][4 c8][row][8]: 17,408 B
constexpr int ET_W3_TAP = 4 * 32 * 16;                 // one conv3 tap: [4 c8][32 rows][8] fp16 = 2 KB

constexpr int ET_OFF_W2 = 0;
constexpr int ET_OFF_A = ET_OFF_W2 + 3 * ET_W2_TAP;                    // 24,576
constexpr int ET_OFF_H2 = ET_OFF_A + 2 * ET_ABUF;                      // 94,208
constexpr int ET_OFF_W3 = ET_OFF_H2 + 2 * ET_H2BUF;                    // (two h2 buffers) packed conv3 taps, 6 KB
constexpr int ET_OFF_W1 = ET_OFF_W3 + 3 * ET_W3_TAP;                   // [5][64] fp32
constexpr int ET_OFF_B1 = ET_OFF_W1 + 5 * 64 * 4;
constexpr int ET_OFF_B2 = ET_OFF_B1 + 64 * 4;
constexpr int ET_OFF_B3 = ET_OFF_B2 + 32 * 4;
constexpr int ET_OFF_BAR = ET_OFF_B3 + 16;                             // a_full[2] acc_full[2] h2_full[2] out_full[2]
constexpr int ET_OFF_TMEM = ET_OFF_BAR + 8 * 8;
// TMEM: conv2 accumulator of A buffer b, issuer j at column 64 (2 b + j); conv3's accumulators at 256 / 288
constexpr uint32_t ET_ACC3_COL = 256;
constexpr int ET_SMEM_BYTES = ET_OFF_TMEM + 16;
static_assert(ET_OFF_A % 1024 == 0 && ET_OFF_W3 % 16 == 0 && ET_SMEM_BYTES <= 232448, "espcn_tc shared-memory map");

// The activations take PRE-SCALED arguments: the weights and biases of conv1 / conv2 carry the factor 2 log2(e), conv3's the
// factor -log2(e) (folded in when the operands are staged / packed), so that
//   tanh(v)    = 1 - 2 / (1 + 2^(2 log2(e) v))      sigmoid(v) = 1 / (1 + 2^(-log2(e) v))
// cost one ex2.approx + one rcp.approx (+ one FFMA) each; both saturate correctly (2^a -> inf: 1 / 0, 2^a -> 0: -1 / 1).
constexpr float ET_TANH_SCALE = 2.885390081777927f;        // 2 log2(e)
constexpr float ET_SIGM_SCALE = -1.4426950408889634f;      // -log2(e)
// (A tanhf-like variant - polynomial a P(a^2) below |v| = 0.55 for relative accuracy near zero, selected against the ex2 form -
// was measured: 12 instead of 5 instructions per tanh, forward 1.96 -> 2.53 ms, and NO change of the final mask error, which is
// set by the 22-bit hi/lo operands of the two tensor-core convolutions, not by the activations.  Sharing one reciprocal between
// two tanh - 1 / (1 + ea) = (1 + eb) / ((1 + ea)(1 + eb)): three MUFU operations instead of four, four more FP32 instructions per
// pair - was measured too: 1.815 -> 1.795 ms, within noise of a wash between the MUFU and the issue floors; not kept.
// Also measured (r11e, with an ncu opcode histogram: 96 FSEL per position for the zero padding, 16 FADD + two TMEM loads per
// conv2 item for the second issuer's partial sums): zero rows by a branch around the item instead of selects, and ONE conv2 issuer
// with one accumulator - 9 % fewer instructions, 1.815 -> 1.836 ms: the kernel is not short of issue slots; not kept.)
__device__ __forceinline__ float tanh_scaled(float a) {
    float e, r;
    asm volatile("ex2.approx.ftz.f32 %0, %1;" : "=f"(e) : "f"(a));
    asm("rcp.approx.ftz.f32 %0, %1;" : "=f"(r) : "f"(e + 1.0f));
    return fmaf(-2.0f, r, 1.0f);
}
__device__ __forceinline__ float sigmoid_scaled(float a) {
    float e, r;
    asm volatile("ex2.approx.ftz.f32 %0, %1;" : "=f"(e) : "f"(a));
    asm("rcp.approx.ftz.f32 %0, %1;" : "=f"(r) : "f"(e + 1.0f));
    return r;
}

struct EtParams {
    const float *x;
    float *mask;
    const void *w2p;           // packed conv2 taps, then packed conv3 taps (set_weights_locked in stof_api.cu)
    const float *w1, *b1, *b2, *b3;
    int N, L, tiles_per_row, total_tiles;
};

__global__ void __launch_bounds__(ET_THREADS, 1) espcn_tc_kernel(const EtParams p) {
    extern __shared__ __align__(1024) uint8_t smem[];
    const uint32_t sbase = smem_u32(smem);
    const int tid = threadIdx.x, warp = tid >> 5, lane = tid & 31;
    const uint32_t bar0 = sbase + ET_OFF_BAR;
    auto BAR_A_FULL = [&](int b) { return bar0 + 8 * b; };
    auto BAR_ACC_FULL = [&](int b) { return bar0 + 16 + 8 * b; };
    // per-buffer barriers: a producer may run one tile ahead of its consumer, and a single barrier two phases ahead of a parity wait aliases
    auto BAR_H2_FULL = [&](int b) { return bar0 + 32 + 8 * b; };
    auto BAR_OUT_FULL = [&](int b) { return bar0 + 48 + 8 * b; };
    volatile uint32_t *s_tmem = reinterpret_cast<volatile uint32_t *>(smem + ET_OFF_TMEM);
    const float4 *s_w1 = reinterpret_cast<const float4 *>(smem + ET_OFF_W1);
    const float4 *s_b1 = reinterpret_cast<const float4 *>(smem + ET_OFF_B1);
    const float4 *s_b2 = reinterpret_cast<const float4 *>(smem + ET_OFF_B2);
    const float4 *s_b3 = reinterpret_cast<const float4 *>(smem + ET_OFF_B3);

    // ---- one-time setup: weights to shared memory, zero rows above the produced 128 of both A buffers, barriers, TMEM ----
    for (int i = tid; i < 3 * ET_W2_TAP / 16; i += ET_THREADS)
        reinterpret_cast<uint4 *>(smem + ET_OFF_W2)[i] = __ldg(reinterpret_cast<const uint4 *>(p.w2p) + i);
    for (int i = tid; i < 3 * ET_W3_TAP / 16; i += ET_THREADS)
        reinterpret_cast<uint4 *>(smem + ET_OFF_W3)[i] = __ldg(reinterpret_cast<const uint4 *>(p.w2p) + 3 * ET_W2_TAP / 16 + i);
    for (int i = tid; i < 5 * 64; i += ET_THREADS) reinterpret_cast<float *>(smem + ET_OFF_W1)[i] = ET_TANH_SCALE * __ldg(p.w1 + i);
    for (int i = tid; i < 64; i += ET_THREADS) reinterpret_cast<float *>(smem + ET_OFF_B1)[i] = ET_TANH_SCALE * __ldg(p.b1 + i);
    for (int i = tid; i < 32; i += ET_THREADS) reinterpret_cast<float *>(smem + ET_OFF_B2)[i] = ET_TANH_SCALE * __ldg(p.b2 + i);
    if (tid < 4) reinterpret_cast<float *>(smem + ET_OFF_B3)[tid] = ET_SIGM_SCALE * __ldg(p.b3 + tid);
    for (int i = tid; i < (2 * 16 + 2 * 8) * (ET_R - 128); i += ET_THREADS) {
        // slab = buffer * 16 + plane * 8 + c8 of the A buffers, then buffer * 8 + plane * 4 + c8 of the h2 planes (contiguous in the map)
        const int slab = i / (ET_R - 128), r = 128 + i % (ET_R - 128);
        *reinterpret_cast<uint4 *>(smem + ET_OFF_A + slab * ET_SLAB + r * 16) = make_uint4(0, 0, 0, 0);
    }
    if (tid == 0) {
        for (int b = 0; b < 2; ++b) {
            mbar_init(BAR_A_FULL(b), ET_WORKER_WARPS);
            mbar_init(BAR_ACC_FULL(b), 2);                 // one commit per conv2 issuer
        }
        for (int b = 0; b < 2; ++b) {
            mbar_init(BAR_H2_FULL(b), ET_WORKER_WARPS);
            mbar_init(BAR_OUT_FULL(b), 1);
        }
        fence_mbar_init();
    }
    if (warp == ET_WORKER_WARPS) tmem_alloc(sbase + ET_OFF_TMEM, 512);
    fence_proxy_async_smem();
    tc_fence_before_sync();
    __syncthreads();
    tc_fence_after_sync();
    const uint32_t tmem0 = *s_tmem;

    const int n_my = (p.total_tiles - (int)blockIdx.x + (int)gridDim.x - 1) / (int)gridDim.x;     // tiles of this CTA

    if (warp >= ET_WORKER_WARPS) {
        // =================================== MMA issuers =================================================
        // A single thread issues these small MMAs (N = 64 / 32 / 16) at ~90 cycles each, far below the pipe's rate, and 36 of
        // them per tile would bound the kernel: three issuing warps work in parallel on independent accumulators - conv2's
        // twelve (tap, K-step) products split even / odd between two of them (the epilogue adds the two partial sums), conv3 on
        // the third.
        constexpr uint32_t kDescHi = (128u >> 4) | (1u << 14);
        constexpr uint32_t idesc_a = umma_idesc_f16(128, 64), idesc_b = umma_idesc_f16(128, 32), idesc_c = umma_idesc_f16(128, 16);
        const int role = warp - ET_WORKER_WARPS;            // 0 / 1: conv2 even / odd products, 2: conv3
        auto conv2 = [&](int k) {
            const int buf = k & 1;
            mbar_wait(BAR_A_FULL(buf), (uint32_t)(k >> 1) & 1u);
            tc_fence_after_sync();
            if (elect_one()) {
                const uint32_t a0 = ((sbase + ET_OFF_A + buf * ET_ABUF) >> 4) | ((uint32_t)(ET_SLAB >> 4) << 16);
                const uint32_t w0 = ((sbase + ET_OFF_W2) >> 4) | (64u << 16);         // LBO = 64 rows of 16 B between c8 chunks
                const uint32_t d = tmem0 + 64u * (uint32_t)(2 * buf + role);
#pragma unroll
                for (int i = 0; i < 6; ++i) {
                    const int q = 2 * i + role, t = q >> 2, ks = q & 3;
                    const uint64_t bw = make_desc(w0 + (uint32_t)(t * (ET_W2_TAP >> 4) + 2 * ks * 64), kDescHi);
                    umma_f16(d, make_desc(a0 + (uint32_t)(t + 2 * ks * ET_R), kDescHi), bw, idesc_a, i > 0 ? 1u : 0u);
                    umma_f16(d + 32u, make_desc(a0 + (uint32_t)(t + (8 + 2 * ks) * ET_R), kDescHi), bw, idesc_b, 1u);
                }
                umma_commit(BAR_ACC_FULL(buf));
            }
            __syncwarp();
        };
        // conv3 (espcn_1d.py:14,34): D0 = columns 0-3, D1 = 8-11 of the 16-column product; the X_lo product lands at + 8
        auto conv3 = [&](int k) {
            mbar_wait(BAR_H2_FULL(k & 1), (uint32_t)(k >> 1) & 1u);
            tc_fence_after_sync();
            if (elect_one()) {
                const uint32_t a0 = ((sbase + ET_OFF_H2 + (k & 1) * ET_H2BUF) >> 4) | ((uint32_t)(ET_SLAB >> 4) << 16);
                const uint32_t w0 = ((sbase + ET_OFF_W3) >> 4) | (32u << 16);
                const uint32_t d = tmem0 + ET_ACC3_COL + 32u * (uint32_t)(k & 1);
#pragma unroll
                for (int t = 0; t < 3; ++t) {
#pragma unroll
                    for (int ks = 0; ks < 2; ++ks) {
                        const uint32_t bw = w0 + (uint32_t)(t * (ET_W3_TAP >> 4) + 2 * ks * 32);
                        umma_f16(d, make_desc(a0 + (uint32_t)(t + 2 * ks * ET_R), kDescHi), make_desc(bw, kDescHi), idesc_c,
                                 (t > 0 || ks > 0) ? 1u : 0u);
                        umma_f16(d + 8u, make_desc(a0 + (uint32_t)(t + (4 + 2 * ks) * ET_R), kDescHi), make_desc(bw + 16u, kDescHi), idesc_c, 1u);
                    }
                }
                umma_commit(BAR_OUT_FULL(k & 1));
            }
            __syncwarp();
        };
        if (role < 2) {
            for (int k = 0; k < n_my; ++k) conv2(k);
        } else {
            for (int k = 0; k < n_my; ++k) conv3(k);
        }
    } else {
        // =================================== workers =====================================================
        const int prow = tid & 127, pcq = tid >> 7;            // producer item: row of the tile, c8 in {pcq, pcq + 4}
        const int quad = warp & 3, cg = warp >> 2;             // epilogue: TMEM lanes of this warp, channels [8 cg, 8 cg + 8)
        const int qrow = 32 * quad + lane;
        const uint32_t lane_addr = (uint32_t)(32 * quad) << 16;

        // tile k of this CTA = tile blockIdx.x + k gridDim.x of the launch = (row n, tile j of the row); the workers walk two
        // cursors (tile k for the epilogues, tile k + 2 for the producer) without dividing
        const int step_n = (int)gridDim.x / p.tiles_per_row, step_j = (int)gridDim.x % p.tiles_per_row;
        struct Cursor { int n, j; };
        auto advance = [&](Cursor &c) {
            c.n += step_n; c.j += step_j;
            if (c.j >= p.tiles_per_row) { c.j -= p.tiles_per_row; ++c.n; }
        };
        // the five input samples of this thread's conv1 row: positions l0 - 4 + prow + t (zero outside the waveform)
        auto load_x = [&](const Cursor &c, float (&xv)[5]) {
            const float *xr = p.x + (size_t)c.n * p.L;
            const int g0 = c.j * ET_OUT - 4 + prow;
#pragma unroll
            for (int t = 0; t < 5; ++t) xv[t] = ((unsigned)(g0 + t) < (unsigned)p.L) ? __ldg(xr + g0 + t) : 0.0f;
        };
        // conv1 + tanh (espcn_1d.py:12,32) of row prow, 2 x 8 channels -> split planes of A buffer `buf`; bias first, taps ascending
        // (the sum order of conv_simt.cu / espcn_fused.cu).  Rows outside [0, L) are conv2's zero padding.
        auto produce = [&](int k, const Cursor &c, const float (&xv)[5]) {
            const bool inside = (unsigned)(c.j * ET_OUT - 2 + prow) < (unsigned)p.L;
            uint8_t *abuf = smem + ET_OFF_A + (k & 1) * ET_ABUF + prow * 16;
#pragma unroll
            for (int half = 0; half < 2; ++half) {
                const int c8 = pcq + 4 * half;
                float v[8];
#pragma unroll
                for (int h = 0; h < 2; ++h) {
                    const float4 bq = s_b1[2 * c8 + h];
                    float2 a01 = make_float2(bq.x, bq.y), a23 = make_float2(bq.z, bq.w);
#pragma unroll
                    for (int t = 0; t < 5; ++t) {
                        const float4 wv = s_w1[t * 16 + 2 * c8 + h];
                        const float2 xx = make_float2(xv[t], xv[t]);
                        a01 = __ffma2_rn(make_float2(wv.x, wv.y), xx, a01);
                        a23 = __ffma2_rn(make_float2(wv.z, wv.w), xx, a23);
                    }
                    // evaluated unconditionally, then selected: no branch around the MUFU chains
                    const float t0 = tanh_scaled(a01.x), t1 = tanh_scaled(a01.y), t2 = tanh_scaled(a23.x), t3 = tanh_scaled(a23.y);
                    v[4 * h + 0] = inside ? t0 : 0.0f;
                    v[4 * h + 1] = inside ? t1 : 0.0f;
                    v[4 * h + 2] = inside ? t2 : 0.0f;
                    v[4 * h + 3] = inside ? t3 : 0.0f;
                }
                uint4 uh, ul;
                split_pair(v[0], v[1], uh.x, ul.x);
                split_pair(v[2], v[3], uh.y, ul.y);
                split_pair(v[4], v[5], uh.z, ul.z);
                split_pair(v[6], v[7], uh.w, ul.w);
                *reinterpret_cast<uint4 *>(abuf + c8 * ET_SLAB) = uh;
                *reinterpret_cast<uint4 *>(abuf + ET_PLANE + c8 * ET_SLAB) = ul;
            }
        };
        // generic-proxy stores of this warp (A buffer and / or h2 planes) -> visible to the MMAs, then the barrier arrivals
        auto publish = [&](int a_buf, int h2_buf) {
            fence_proxy_async_smem();
            tc_fence_before_sync();                       // also orders this thread's earlier tcgen05.ld of accumulators the MMAs will reuse
            __syncwarp();
            if (lane == 0) {
                if (h2_buf >= 0) mbar_arrive(BAR_H2_FULL(h2_buf));
                if (a_buf >= 0) mbar_arrive(BAR_A_FULL(a_buf));
            }
        };
        // conv3 accumulator of tile kk -> SampleShuffle1D + sigmoid (espcn_1d.py:34-36; utils/sample_shuffle.py:20-26): every worker
        // waits (the tile's h2 buffer and accumulator are reused two tiles later), one warp per quadrant (rotating) stores
        auto finish = [&](int kk, int n, int l0) {
            mbar_wait(BAR_OUT_FULL(kk & 1), (uint32_t)(kk >> 1) & 1u);
            tc_fence_after_sync();
            if (cg == (kk & 3)) {
                uint32_t d[16];
                tmem_ld16(tmem0 + ET_ACC3_COL + 32u * (uint32_t)(kk & 1) + lane_addr, d);
                tmem_ld_wait();
                const int l = l0 + qrow;
                if (qrow < ET_OUT && l < p.L) {
                    const float4 b3 = s_b3[0];
                    float4 out;
                    out.x = sigmoid_scaled((__uint_as_float(d[0]) + TC_LO_INV * __uint_as_float(d[8])) + b3.x);
                    out.y = sigmoid_scaled((__uint_as_float(d[1]) + TC_LO_INV * __uint_as_float(d[9])) + b3.y);
                    out.z = sigmoid_scaled((__uint_as_float(d[2]) + TC_LO_INV * __uint_as_float(d[10])) + b3.z);
                    out.w = sigmoid_scaled((__uint_as_float(d[3]) + TC_LO_INV * __uint_as_float(d[11])) + b3.w);
                    reinterpret_cast<float4 *>(p.mask)[(size_t)n * p.L + l] = out;
                }
            }
        };

        float xv[5];
        Cursor cur, nxt;                                  // tile k, tile k + 2
        cur.n = (int)blockIdx.x / p.tiles_per_row;
        cur.j = (int)blockIdx.x - cur.n * p.tiles_per_row;
        nxt = cur;
        for (int k = 0; k < 2 && k < n_my; ++k) {
            load_x(nxt, xv);
            produce(k, nxt, xv);
            publish(k, -1);
            advance(nxt);
        }
        int prev_n = 0, prev_l0 = 0;
        for (int k = 0; k < n_my; ++k, advance(cur)) {
            const int n = cur.n, l0 = cur.j * ET_OUT;
            if (k + 2 < n_my) load_x(nxt, xv);            // in flight behind the epilogue of tile k
            // ---- (1) conv2 accumulator -> h2 = tanh(conv2 + bias), zero outside the waveform (espcn_1d.py:13,33), as split planes:
            //          this thread's 8 channels of one row are exactly one 16-byte operand vector per plane
            mbar_wait(BAR_ACC_FULL(k & 1), (uint32_t)(k >> 1) & 1u);
            tc_fence_after_sync();
            {
                const uint32_t d = tmem0 + 128u * (uint32_t)(k & 1) + lane_addr + 8u * (uint32_t)cg;
                uint32_t d0[8], d1[8], e0[8], e1[8];                  // the two issuers' partial sums
                tmem_ld8(d, d0);
                tmem_ld8(d + 32u, d1);
                tmem_ld8(d + 64u, e0);
                tmem_ld8(d + 96u, e1);
                tmem_ld_wait();
                const bool inside = (unsigned)(l0 - 1 + qrow) < (unsigned)p.L;
                const float4 ba = s_b2[2 * cg], bb = s_b2[2 * cg + 1];
                const float bv[8] = {ba.x, ba.y, ba.z, ba.w, bb.x, bb.y, bb.z, bb.w};
                float v[8];
#pragma unroll
                for (int i = 0; i < 8; ++i) {
                    const float a = ((__uint_as_float(d0[i]) + __uint_as_float(e0[i])) +
                                     TC_LO_INV * (__uint_as_float(d1[i]) + __uint_as_float(e1[i]))) + bv[i];
                    const float th = tanh_scaled(a);
                    v[i] = inside ? th : 0.0f;
                }
                uint4 uh, ul;
                split_pair(v[0], v[1], uh.x, ul.x);
                split_pair(v[2], v[3], uh.y, ul.y);
                split_pair(v[4], v[5], uh.z, ul.z);
                split_pair(v[6], v[7], uh.w, ul.w);
                uint8_t *hb = smem + ET_OFF_H2 + (k & 1) * ET_H2BUF + cg * ET_SLAB + qrow * 16;
                *reinterpret_cast<uint4 *>(hb) = uh;
                *reinterpret_cast<uint4 *>(hb + 4 * ET_SLAB) = ul;
            }
            // ---- (2) conv1 operand of tile k + 2 into the A buffer tile k's MMAs have released; ONE proxy fence then publishes
            //          both the h2 planes of tile k (conv3 may start) and that operand (conv2 of tile k + 2 may start)
            if (k + 2 < n_my) {
                produce(k + 2, nxt, xv);
                advance(nxt);
                publish(k & 1, k & 1);
            } else {
                publish(-1, k & 1);
            }
            // ---- (3) the PREVIOUS tile's conv3 accumulator (issued a whole iteration ago: no wait in the steady state) ------------
            if (k > 0) finish(k - 1, prev_n, prev_l0);
            prev_n = n; prev_l0 = l0;
        }
        finish(n_my - 1, prev_n, prev_l0);
    }

    // ---- teardown --------------------------------------------------------------------------------------
    tc_fence_before_sync();
    __syncthreads();
    if (warp == ET_WORKER_WARPS) {
        tc_fence_after_sync();
        tmem_dealloc(tmem0, 512);
    }
}

int launch_espcn_tc(const float *x, float *mask, int N, int L, const float *w1, const float *b1, const void *w23_packed, const float *b2,
                    const float *b3, cudaStream_t st) {
    static SmemAttrOnce attr;
    { const int rc = ensure_dyn_smem(espcn_tc_kernel, ET_SMEM_BYTES, attr); if (rc) return rc; }
    EtParams p;
    p.x = x; p.mask = mask; p.w2p = w23_packed; p.w1 = w1; p.b1 = b1; p.b2 = b2; p.b3 = b3;
    p.N = N; p.L = L;
    p.tiles_per_row = (L + ET_OUT - 1) / ET_OUT;
    const long long total = (long long)N * p.tiles_per_row;
    STOF_REQUIRE(total < (1ll << 31), STOF_E_INVALID, "espcn_tc: too many tiles");
    p.total_tiles = (int)total;
    const int grid = (int)(total < kSMs ? total : kSMs);
    espcn_tc_kernel<<<grid, ET_THREADS, ET_SMEM_BYTES, st>>>(p);
    STOF_LAUNCH_CHECK();
    return STOF_OK;
}

}  // namespace stof
