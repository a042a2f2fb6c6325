// EXPERIMENT (round 2, not built into the library): K-group-pipelined variant of the chain kernel.  Drop-in replacement for
// stofnet_b200/csrc/conv_tc.cu (needs `int in_ks;` in TcParams); results verified (tests/test_gpu_tc.py, bench parity check),
// measured SLOWER than the shipped tap-outer kernel on the same box: mid segment 1.96 vs 1.84 ms, tail 1.77 vs 1.55 ms
// (profiles/r4h_kgroup_vs_base_*.txt, timeline profiles/r4j_kgroup_timeline_mid.txt, ncu profiles/r4i_kgroup_ncu_summary.txt:
// tensor pipe active 59 % vs 65 %).  Why: a block's next layer can start after the epilogue's first
// 8-channel chunk, but that chunk's LATENCY (TMEM load 330 + dependent math / split / STS / proxy fence ~570 + barrier
// hand-over) is ~1,600 cycles from "accumulator full" to the first MMA of the next layer - as long as the whole epilogue of
// the tap-outer kernel (1,550) - while its extra cost (4 fences + arrivals per block and layer, 4 KB weight slices, the
// tcgen05.cp short-skip staging at 64 B/clk) is paid in full.  Kept for the record of what was built and measured.
//
// Tensor-core (tcgen05 / TMEM) kernels of the StofNet conv stack (models/stofnet.py:30-31,51-65).
//
// chain_tc_kernel runs a SEGMENT of consecutive layers of the 64->64 k7 stack (conv2..conv12) and,
// optionally, conv_last + SampleShuffle1D, on one tile of a row without leaving the SM:
//
//   * a Conv1d layer is an implicit GEMM  D[position, co] = sum_tap sum_ci X[position + tap - pad, ci] * W_tap[co, ci]
//     with M = 128 positions (TMEM lanes), N = output channels (TMEM columns), K = 64 input channels per tap.
//     Activations live in shared memory as "rows x 8-channel" 16-byte vectors (layout [plane][c8][row][8] fp16), which
//     is exactly the no-swizzle K-major UMMA operand layout with SBO = 128 B: a one-sample tap shift is a 16-byte
//     shift of the descriptor start address, so the 7 taps are 7 descriptors into the same tile (no im2col copy).
//   * fp32 accuracy on fp16 tensor cores: every fp32 value v is carried as two fp16 planes, hi = fp16(v) and
//     lo = fp16((v - hi) * 2^11) (22 significand bits); weights likewise.  Per K-step two MMAs are issued,
//          [D0 | D1] (+)= X_hi * [W_hi ; W_lo]^T      (N = 128)
//               D1    += X_lo * W_hi^T                 (N = 64)
//     and the epilogue forms D0 + 2^-11 * D1 (the dropped lo*lo term is 2^-22 relative).  Stacking W_hi and W_lo along N
//     halves the shared-memory operand traffic per MAC, which is what bounds N = 64 MMAs.
//   * the epilogue (8 warps, one TMEM lane = one position per thread) adds bias / residual, applies LeakyReLU, zeroes
//     positions outside the row (per-layer zero padding), re-splits to hi/lo fp16 and writes the next layer's operand
//     straight back to shared memory; block grids of successive layers are skewed by the layer's padding so block b of
//     layer k+1 only depends on blocks <= b of layer k.
//   * K-GROUP PIPELINING between a block's epilogue and its next layer: the MMAs of a layer are issued K-group-outer
//     (ks = 16 input channels: all taps of ks 0, then of ks 1, ...), and the epilogue produces the output channels in the same
//     order (warp group g handles the 8-channel chunks 2h + g, h = 0..3), signalling "channels of K-group h are in shared
//     memory" after each chunk.  The next layer's first MMAs therefore start after a QUARTER of the epilogue instead of after
//     all of it; they write a different accumulator (three 128-column accumulators rotate over the two blocks' layers), so
//     the epilogue keeps draining TMEM chunk by chunk while the tensor pipe is already busy with the same block.
//   * weights stream through a ring of 16 KB stages filled by bulk async copies (TMA engine) on mbarriers; one stage =
//     up to 4 taps of one K-group (4 KB per tap and K-group), both blocks' streams consume every stage.
//   * the short skip of a segment's LAST layer is parked as fp32 in spare TMEM columns by the previous layer's epilogue, so
//     the input buffer is free one layer before the tile ends and the next tile's input load hides behind that layer.
//
// Roles: warp 0 = producer (input tile + weight ring), warps 1-2 = MMA issuers (one per 128-row block; warp 1 owns
// TMEM), warps 3-10 = epilogue.
#include <cuda_fp16.h>
#include <string.h>

#include "conv_tc_shared.cuh"

namespace stof {

using namespace ptx;

// shared memory map (dynamic, 1024-aligned base)
constexpr int TC_OFF_P = 0;
constexpr int TC_OFF_Q = TC_BUF;
constexpr int TC_OFF_W = 2 * TC_BUF;
constexpr int TC_OFF_BIAS = TC_OFF_W + TC_STAGES * TC_STAGE_BYTES;          // [TC_MAX_LAYERS][64] fp32
constexpr int TC_OFF_BAR = TC_OFF_BIAS + TC_MAX_LAYERS * 64 * 4;
// barriers (8 B each): w_full[STAGES] w_empty[STAGES] acc_full[2] acc_free[3] act[TC_MAX_LAYERS][2][4] in_full p_free
constexpr int TC_NBAR = 2 * TC_MAX_STAGES + 2 + 3 + 8 * TC_MAX_LAYERS + 2;
constexpr int TC_OFF_TMEM = TC_OFF_BAR + TC_NBAR * 8;
constexpr int TC_SMEM_BYTES = TC_OFF_TMEM + 16;
constexpr int TC_GRP_MAX = 4;                                               // taps per weight stage (4 KB each in exact mode)
constexpr uint32_t TC_STASH_COL = 384;                                      // TMEM: 3 accumulators x 128 columns, then the stash

static_assert(TC_SMEM_BYTES <= 232448, "chain kernel exceeds the 227 KB shared-memory limit");
static_assert(TC_GRP_MAX * (TC_STAGE_BYTES / 4) <= TC_STAGE_BYTES, "a tap group must fit one ring stage");

// optional timeline (stof_debug_set_timeline): roles 0..3 (producer, MMA issuer 0 / 1, epilogue warp 3) of CTA 0 each append
// (code, clock64) pairs to a private quarter of the buffer with a register counter (no atomics: the probe must not stall
// the warp it observes); code = role<<24 | k<<16 | b<<8 | what.  tl[role] receives the role's event count at kernel end.
__device__ __forceinline__ void tl_event(unsigned long long *tl, int cap, uint32_t role, uint32_t code, int &cnt) {
    if (tl == nullptr || blockIdx.x != 0) return;
    const int capr = cap >> 2;
    if (cnt < capr) {
        unsigned long long *e = tl + 4 + 2 * ((size_t)role * capr + cnt);
        e[0] = code;
        e[1] = (unsigned long long)clock64();
    }
    ++cnt;
}
#define TL(role, k, b, what) tl_event(p.timeline, p.timeline_cap, (role), ((role) << 24) | ((uint32_t)(k) << 16) | ((uint32_t)(b) << 8) | (what), tl_cnt)
#define TL_FLUSH(role) do { if (p.timeline != nullptr && blockIdx.x == 0) p.timeline[(role)] = (unsigned long long)tl_cnt; } while (0)

// taps of a layer are split into ceil(taps / 4) groups of equal size (7 -> 4 + 3, 3 -> 3): one weight-ring stage per (K-group, tap group)
__device__ __forceinline__ int tap_group_size(int taps) {
    const int ngrp = (taps + TC_GRP_MAX - 1) / TC_GRP_MAX;
    return (taps + ngrp - 1) / ngrp;
}

// tcgen05.wait::ld that also "touches" the registers the loads wrote, so that no use of them can be scheduled above the wait
__device__ __forceinline__ void tmem_ld_wait_dep(uint32_t (&a)[8], uint32_t (&b)[8]) {
    asm volatile("tcgen05.wait::ld.sync.aligned;"
                 : "+r"(a[0]), "+r"(a[1]), "+r"(a[2]), "+r"(a[3]), "+r"(a[4]), "+r"(a[5]), "+r"(a[6]), "+r"(a[7]), "+r"(b[0]), "+r"(b[1]),
                   "+r"(b[2]), "+r"(b[3]), "+r"(b[4]), "+r"(b[5]), "+r"(b[6]), "+r"(b[7])
                 :
                 : "memory");
}

// All MMAs of one layer for one 128-row block, issued by a single thread, K-group outer / tap inner.
// Layer kinds (upload_tc in stof_api.cu):
//   64 -> 64 conv : taps [c8][W_hi 64 | W_lo 64][8] (one-pass mode: [c8][W_hi 64][8]); D0 = columns 0-63, D1 = 64-127
//   conv_last     : taps [c8][32 rows][8]: rows 0-3 W_hi, 8-11 W_lo, 16-19 W_hi (one-pass: [c8][16][8]); D0 = 0-3, D1 = 8-11
// A ring stage holds, for one K-group, the (tap, K-group) slices [2 c8][W_ROWS][8] of up to 4 consecutive taps.
// Descriptor low word = (addr >> 4) | (LBO >> 4) << 16; the high word (SBO = 128 B, version 1) is constant.
// bar_act0 / bar_act1: "K-group 0 of the previous layer's output is in shared memory" barriers of block 0 / of this block when it
// is block 1 (0 = no dependency: first layer of a tile); the barriers of K-group ks follow at + 8 ks.
template <bool FAST, bool LAST>
__device__ __forceinline__ void issue_layer(uint32_t a_base, uint32_t w_base, uint32_t d_tmem, int taps, int s, uint32_t ph,
                                            uint32_t bar_w_full0, uint32_t bar_w_empty0, uint32_t bar_acc_full, uint32_t bar_act0,
                                            uint32_t bar_act1, uint32_t act_par, uint32_t res_src, uint32_t res_tmem, uint32_t bar_stash, uint32_t stash_par,
                                            unsigned long long *tl, int tl_cap, int &tl_cnt, uint32_t tl_code) {
    constexpr uint32_t kDescHi = (128u >> 4) | (1u << 14);
    constexpr uint32_t W_ROWS = LAST ? (FAST ? 16u : 32u) : (FAST ? 64u : 128u);    // rows per c8 chunk of a tap
    constexpr uint32_t SLICE16 = 2u * W_ROWS;                                        // one (tap, K-group) slice / 16 B
    constexpr uint32_t LO_ROW = LAST ? 16u : 0u;                                     // W_hi rows of the X_lo product
    constexpr uint32_t LO_COL = LAST ? 8u : 64u;                                     // its accumulator columns
    constexpr uint32_t STRIDE16 = (uint32_t)(FAST ? TC_STAGE_BYTES / 2 : TC_STAGE_BYTES) >> 4;
    constexpr int NST = FAST ? TC_MAX_STAGES : TC_STAGES;
    constexpr uint32_t idesc_a = umma_idesc_f16(128, LAST ? 16 : (FAST ? 64 : 128));
    constexpr uint32_t idesc_b = umma_idesc_f16(128, LAST ? 16 : 64);
    const int grp = tap_group_size(taps);
    // timeline (debug): per K-group a time stamp after the activation wait, per layer the cycles spent waiting for activations /
    // for weight stages (the value replaces the time stamp in events 40 / 41)
    const bool tl_on = (tl != nullptr) && blockIdx.x == 0;
    long long wait_act = 0, wait_w = 0;
#pragma unroll 1
    for (int ks = 0; ks < 4; ++ks) {
        if (bar_act0 != 0u) {
            const long long c0 = tl_on ? clock64() : 0;
            mbar_wait(bar_act0 + 8u * (uint32_t)ks, act_par);
            if (bar_act1 != 0u) mbar_wait(bar_act1 + 8u * (uint32_t)ks, act_par);
            tc_fence_after_sync();
            if (tl_on) wait_act += clock64() - c0;
        }
        if (tl_on) tl_event(tl, tl_cap, tl_code >> 24, tl_code | (16u + (uint32_t)ks), tl_cnt);
        if (ks == 3 && res_src != 0u) {
            // short skip of the NEXT layer (x two layers back from it = this layer's input): the block's 128 rows (hi / lo planes)
            // go from this layer's input buffer to the TMEM stash through the tensor pipe's own shared-memory path - 4 columns per
            // 8-channel slab, packed fp16 pairs - so the epilogue never reads them with LDS (which starves while the MMA streams own
            // the read port), and the buffer is no longer needed for it once this layer's MMAs have completed.  The stash was last
            // read by this block's epilogue two steps back (bar_stash).
            if (bar_stash != 0u) {
                mbar_wait(bar_stash, stash_par);
                tc_fence_after_sync();
            }
#pragma unroll
            for (int slab = 0; slab < (FAST ? 8 : 16); ++slab)
                tmem_cp_128x128b(res_tmem + 4u * (uint32_t)slab, make_desc(res_src + (uint32_t)(slab * TC_R), kDescHi));
        }
        const uint32_t a_hi = a_base + (uint32_t)(2 * ks * TC_R);
        const uint32_t a_lo = a_base + (uint32_t)((8 + 2 * ks) * TC_R);
#pragma unroll 1
        for (int t0 = 0; t0 < taps; t0 += grp) {
            const int nt = min(grp, taps - t0);
            const long long c1 = tl_on ? clock64() : 0;
            mbar_wait(bar_w_full0 + 8u * (uint32_t)s, ph);
            tc_fence_after_sync();
            if (tl_on) wait_w += clock64() - c1;
            const uint32_t w0 = (w_base + (uint32_t)s * STRIDE16) | (W_ROWS << 16);
#pragma unroll
            for (int j = 0; j < TC_GRP_MAX; ++j) {
                if (j < nt) {
                    const uint32_t t = (uint32_t)(t0 + j);
                    umma_f16(d_tmem, make_desc(a_hi + t, kDescHi), make_desc(w0 + (uint32_t)j * SLICE16, kDescHi), idesc_a,
                             (ks > 0 || t > 0u) ? 1u : 0u);
                    if (!FAST)
                        umma_f16(d_tmem + LO_COL, make_desc(a_lo + t, kDescHi), make_desc(w0 + (uint32_t)j * SLICE16 + LO_ROW, kDescHi),
                                 idesc_b, 1u);
                }
            }
            umma_commit(bar_w_empty0 + 8u * (uint32_t)s);                        // this stream is done with the stage
            if (++s == NST) { s = 0; ph ^= 1u; }
        }
    }
    umma_commit(bar_acc_full);
    if (tl_on) {
        const int capr = tl_cap >> 2;
        const uint32_t role = tl_code >> 24;
        if (tl_cnt + 1 < capr) {
            unsigned long long *e = tl + 4 + 2 * ((size_t)role * capr + tl_cnt);
            e[0] = tl_code | 40u; e[1] = (unsigned long long)wait_act;
            e[2] = tl_code | 41u; e[3] = (unsigned long long)wait_w;
        }
        tl_cnt += 2;
    }
}

template <bool FAST>
__global__ void __launch_bounds__(TC_THREADS, 1) chain_tc_kernel(const TcParams p) {
    extern __shared__ __align__(1024) uint8_t smem[];
    const uint32_t sbase = smem_u32(smem);
    const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
    int tl_cnt = 0;
    const uint32_t bar0 = sbase + TC_OFF_BAR;
    auto BAR_W_FULL = [&](int s) { return bar0 + 8 * s; };
    auto BAR_W_EMPTY = [&](int s) { return bar0 + 8 * (TC_MAX_STAGES + s); };
    auto BAR_ACC_FULL = [&](int b) { return bar0 + 8 * (2 * TC_MAX_STAGES + b); };
    auto BAR_ACC_FREE = [&](int a) { return bar0 + 8 * (2 * TC_MAX_STAGES + 2 + a); };
    // "output channels of K-group h of layer k, block b are in shared memory"
    auto BAR_ACT = [&](int k, int b, int h) { return bar0 + 8 * (2 * TC_MAX_STAGES + 5 + 8 * k + 4 * b + h); };
    const uint32_t BAR_IN_FULL = bar0 + 8 * (2 * TC_MAX_STAGES + 5 + 8 * TC_MAX_LAYERS);
    constexpr int n_stages = FAST ? TC_MAX_STAGES : TC_STAGES;          // weight ring depth / stage stride
    constexpr uint32_t stage_stride = FAST ? TC_STAGE_BYTES / 2 : TC_STAGE_BYTES;
    const uint32_t BAR_P_FREE = BAR_IN_FULL + 8;
    float *s_bias = reinterpret_cast<float *>(smem + TC_OFF_BIAS);
    volatile uint32_t *s_tmem = reinterpret_cast<volatile uint32_t *>(smem + TC_OFF_TMEM);

    // ---- one-time setup ------------------------------------------------------------------------------
    for (int i = threadIdx.x; i < p.n_layers * 64; i += TC_THREADS) {
        const int k = i >> 6, c = i & 63;
        s_bias[i] = (c < p.layer[k].n_out) ? __ldg(p.layer[k].bias + c) : 0.0f;
    }
    if (threadIdx.x == 0) {
        for (int s = 0; s < TC_MAX_STAGES; ++s) {
            mbar_init(BAR_W_FULL(s), 1);
            mbar_init(BAR_W_EMPTY(s), 2);                 // one issuing warp per block consumes a stage
        }
        for (int b = 0; b < 2; ++b) mbar_init(BAR_ACC_FULL(b), 1);
        for (int a = 0; a < 3; ++a) mbar_init(BAR_ACC_FREE(a), TC_EPI_WARPS);
        for (int k = 0; k < TC_MAX_LAYERS; ++k)
            for (int b = 0; b < 2; ++b)
                for (int h = 0; h < 4; ++h) mbar_init(BAR_ACT(k, b, h), TC_EPI_WARPS);
        mbar_init(BAR_IN_FULL, 1);
        mbar_init(BAR_P_FREE, TC_EPI_WARPS);
        fence_mbar_init();
    }
    if (warp == 1) tmem_alloc(sbase + TC_OFF_TMEM, 512);
    tc_fence_before_sync();
    __syncthreads();
    tc_fence_after_sync();
    const uint32_t tmem0 = *s_tmem;

    const int n_layers = p.n_layers;
    const int r_in = p.r_in;                 // input rows staged per tile

    if (warp == 0) {
        // =================================== producer ===================================================
        // Input tile j of this CTA goes into buffer P as soon as tile j-1 no longer touches P (BAR_P_FREE, arrived by the
        // epilogue after layer p.free_layer); with an even number of layers that is one whole layer before the tile
        // ends, so the load of the next tile hides behind the last layer.  Weights: every stage of every layer in issue order.
        auto issue_input = [&](int tile, int j) {
            const int n = tile / p.tiles_per_row;
            const int l_org = (tile - n * p.tiles_per_row) * p.t_out - p.l_shift;
            if (j > 0) mbar_wait(BAR_P_FREE, (uint32_t)(j - 1) & 1u);
            if (lane == 0) TL(0u, 0, 0, 1);
            const int r_lo = max(0, -l_org), r_hi = min(r_in, p.L - l_org);
            const int nslab = FAST ? 8 : 16;               // one-pass fp16 mode: hi plane only
            // rows outside the waveform are the conv's zero padding (only tiles at the ends of a row have any)
            if (r_lo > 0 || r_hi < r_in) {
                for (int slab = 0; slab < nslab; ++slab) {
                    uint8_t *sl = smem + TC_OFF_P + slab * TC_SLAB;
                    for (int r = lane; r < r_lo; r += 32) *reinterpret_cast<uint4 *>(sl + r * 16) = make_uint4(0, 0, 0, 0);
                    for (int r = r_hi + lane; r < r_in; r += 32) *reinterpret_cast<uint4 *>(sl + r * 16) = make_uint4(0, 0, 0, 0);
                }
            }
            fence_proxy_async_smem();
            __syncwarp();
            if (lane == 0) {
                const uint32_t bytes = (uint32_t)(r_hi - r_lo) * 16u;
                mbar_arrive_expect_tx(BAR_IN_FULL, bytes * (uint32_t)nslab);
                const __half *src = reinterpret_cast<const __half *>(p.in) + (size_t)n * 16 * p.L * 8;
                for (int slab = 0; slab < nslab; ++slab)
                    bulk_g2s(sbase + TC_OFF_P + slab * TC_SLAB + r_lo * 16, src + ((size_t)slab * p.L + l_org + r_lo) * 8, bytes,
                             BAR_IN_FULL);
            }
            __syncwarp();
        };
        int s = 0;
        uint32_t ph = 0;
        int it = 0;
        if ((int)blockIdx.x < p.total_tiles) issue_input(blockIdx.x, 0);
        for (int tile = blockIdx.x; tile < p.total_tiles; tile += gridDim.x, ++it) {
            const int next = tile + gridDim.x;
            for (int k = 0; k < n_layers; ++k) {
                const TcLayer &ly = p.layer[k];
                if (ly.type == TC_LONGSKIP && k > 0 && lane == 0) {
                    // pull this tile's rows of the long-skip tensor into L2 one layer before the epilogue reads them
                    const int n = tile / p.tiles_per_row;
                    const int l0 = (tile - n * p.tiles_per_row) * p.t_out;
                    const int a0 = max(0, l0 - 16), a1 = min(p.L, l0 + p.t_out + 16);
                    const __half *src = reinterpret_cast<const __half *>(p.x0) + (size_t)n * 16 * p.L * 8;
                    for (int slab = 0; slab < (FAST ? 8 : 16); ++slab)
                        bulk_prefetch_l2(src + ((size_t)slab * p.L + a0) * 8, (uint32_t)(a1 - a0) * 16u);
                }
                const int taps = ly.taps;
                const int grp = tap_group_size(taps);
                const uint32_t tap_bytes = (uint32_t)(FAST ? ly.stage_bytes_fast : ly.stage_bytes);
                const uint32_t slice = tap_bytes >> 2;                       // one tap, one K-group: [2 c8][rows][8]
                const uint8_t *wsrc = reinterpret_cast<const uint8_t *>(FAST ? ly.w_fast : ly.w);
                for (int ks = 0; ks < 4; ++ks) {
                    for (int t0 = 0; t0 < taps; t0 += grp) {
                        const int nt = min(grp, taps - t0);
                        if (lane == 0) {
                            mbar_wait(BAR_W_EMPTY(s), ph ^ 1u);
                            mbar_arrive_expect_tx(BAR_W_FULL(s), slice * (uint32_t)nt);
                        }
                        __syncwarp();                                        // the stage is free: lanes 0..nt-1 fetch one tap slice each
                        if (lane < nt)
                            bulk_g2s(sbase + TC_OFF_W + s * stage_stride + (uint32_t)lane * slice,
                                     wsrc + (size_t)(t0 + lane) * tap_bytes + (size_t)ks * slice, slice, BAR_W_FULL(s));
                        if (++s == n_stages) { s = 0; ph ^= 1u; }
                    }
                    if (k == p.in_layer && ks == p.in_ks && next < p.total_tiles) {
                        __syncwarp();
                        issue_input(next, it + 1);
                    }
                }
            }
        }
        if (lane == 0) TL_FLUSH(0u);
    } else if (warp <= 2) {
        // =================================== MMA issuers =================================================
        // Two issuing warps, one per 128-row block of the tile (warp 1 -> block 0, warp 2 -> block 1); each block's MMAs are
        // issued by one thread in a fixed order, so the accumulation order - hence the result - is deterministic.  The two
        // streams (different accumulators) hide each other's issue gaps; the weight ring bounds the lag between the blocks.
        // Accumulators rotate: the q-th (layer, block) step of this CTA, q = 2 * layer sequence + block, uses accumulator
        // q mod 3, which was last read by the epilogue of step q - 3 (BAR_ACC_FREE) - always complete before step q can start,
        // because the epilogue handles the steps in order and step q waits for the first chunk of step q - 2.
        // Each warp runs warp-uniform control flow (descriptors stay in uniform registers); one elected lane issues.
        constexpr uint32_t kALoBase = (uint32_t)(TC_SLAB >> 4) << 16;            // A: LBO = one slab
        const int b = warp - 1;
        const uint32_t role = 1u + (uint32_t)b;
        const bool tl_on = (lane == 0);
        int s = 0;
        uint32_t ph = 0, lseq = 0;
        int accsel = b;                      // q mod 3
        uint32_t accph = 0;                  // (q / 3) & 1
        int it = 0;
        for (int tile = blockIdx.x; tile < p.total_tiles; tile += gridDim.x, ++it) {
            if (tl_on) TL(role, 0, b, 1);
            mbar_wait(BAR_IN_FULL, (uint32_t)(it & 1));
            tc_fence_after_sync();
            if (tl_on) TL(role, 0, b, 2);
            for (int k = 0; k < n_layers; ++k, ++lseq) {
                const TcLayer &ly = p.layer[k];
                const int taps = ly.taps;
                const uint32_t a_base = (((sbase + ((k & 1) ? TC_OFF_Q : TC_OFF_P)) >> 4) | kALoBase) +
                                        (uint32_t)(p.b_row[k] - ly.pad + 128 * b);
                const uint32_t d_tmem = tmem0 + 128u * (uint32_t)accsel;
                if (tl_on) TL(role, k, b, 3);
                mbar_wait(BAR_ACC_FREE(accsel), accph ^ 1u);               // the epilogue of step q - 3 has drained this accumulator
                tc_fence_after_sync();
                // the layer's MMAs are issued by ONE elected lane in a single region (it also polls the activation and weight
                // barriers), with every per-layer quantity a compile-time constant of the layer kind
                if (elect_one()) {
                    const uint32_t w_base = (sbase + TC_OFF_W) >> 4;
                    const uint32_t act0 = (k > 0) ? BAR_ACT(k - 1, 0, 0) : 0u;
                    const uint32_t act1 = (k > 0 && b == 1) ? BAR_ACT(k - 1, 1, 0) : 0u;
                    const uint32_t tlc = (role << 24) | ((uint32_t)k << 16) | ((uint32_t)b << 8);
                    // short skip of the next layer: rows of its block b in ITS output buffer = this layer's input buffer
                    const bool next_res = (k + 1 < n_layers) && (p.layer[k + 1].type == TC_RES);
                    const uint32_t res_src = next_res ? (((sbase + ((k & 1) ? TC_OFF_Q : TC_OFF_P)) >> 4) | kALoBase) +
                                                            (uint32_t)(p.b_row[k + 1] + 128 * b)
                                                      : 0u;
                    // the stash's last reader: this block's epilogue two steps back (q - 2), whose "accumulator drained" arrival
                    // comes after its last stash load
                    const uint32_t bar_stash = (lseq > 0u) ? BAR_ACC_FREE(accsel == 2 ? 0 : accsel + 1) : 0u;
                    const uint32_t stash_par = (accsel == 2) ? accph : (accph ^ 1u);
                    const uint32_t res_tmem = tmem0 + TC_STASH_COL + 64u * (uint32_t)b;
                    if (ly.type == TC_LAST)
                        issue_layer<FAST, true>(a_base, w_base, d_tmem, taps, s, ph, BAR_W_FULL(0), BAR_W_EMPTY(0), BAR_ACC_FULL(b), act0,
                                                act1, (uint32_t)(it & 1), res_src, res_tmem, bar_stash, stash_par, p.timeline, p.timeline_cap, tl_cnt, tlc);
                    else
                        issue_layer<FAST, false>(a_base, w_base, d_tmem, taps, s, ph, BAR_W_FULL(0), BAR_W_EMPTY(0), BAR_ACC_FULL(b), act0,
                                                 act1, (uint32_t)(it & 1), res_src, res_tmem, bar_stash, stash_par, p.timeline, p.timeline_cap, tl_cnt, tlc);
                }
                __syncwarp();
                {
                    const int grp = tap_group_size(taps);
                    s += 4 * ((taps + grp - 1) / grp);
                    while (s >= n_stages) { s -= n_stages; ph ^= 1u; }
                }
                accsel += 2;
                if (accsel >= 3) { accsel -= 3; accph ^= 1u; }
                if (tl_on) TL(role, k, b, 4);
            }
        }
        if (tl_on) TL_FLUSH(role);
    } else {
        // =================================== epilogue ====================================================
        // 8 warps: warp quadrant (warp & 3) selects the 32 TMEM lanes, group g (0/1) the 8-channel chunks 2h + g, h = 0..3.
        // The blocks are processed one after the other; within a block the chunks are produced in K-group order and signalled
        // one by one, so the block's next layer starts after the first chunk while the rest of the epilogue still runs.
        const int quad = warp & 3;
        const int egrp = (warp - 3) >> 2;
        const int trow = 32 * quad + lane;                       // TMEM lane = row within the 128-block
        const uint32_t lane_addr = (uint32_t)(32 * quad) << 16;
        const bool tl_on = (warp == 3) && (lane == 0);
        uint32_t lseq = 0;
        int accsel = 0;
        int it = 0;
        for (int tile = blockIdx.x; tile < p.total_tiles; tile += gridDim.x, ++it) {
            const int n = tile / p.tiles_per_row;
            const int l0 = (tile - n * p.tiles_per_row) * p.t_out;
            const int l_org = l0 - p.l_shift;
            const int l_end = min(l0 + p.t_out, p.L);
            for (int k = 0; k < n_layers; ++k, ++lseq) {
                const TcLayer &ly = p.layer[k];
                const int type = ly.type;
                const bool last_in_seg = (k == n_layers - 1);
                uint8_t *outbuf = smem + ((k & 1) ? TC_OFF_P : TC_OFF_Q);
                const float4 *bias4 = reinterpret_cast<const float4 *>(s_bias + 64 * k);
                // short skip (x two layers back): the block's MMA issuer copies those rows (packed fp16 hi / lo pairs) from this
                // layer's output buffer into the TMEM stash with tcgen05.cp, behind the layer's first K-group
                const bool from_stash = (type == TC_RES);
                for (int b = 0; b < 2; ++b) {
                    const int r = p.b_row[k] + 128 * b + trow;
                    const int l = l_org + r;
                    const bool inside = (l >= 0) && (l < p.L);
                    // Everything the epilogue reads from shared / global memory is requested BEFORE waiting for the accumulator:
                    // while the MMA streams run, their operand reads own the shared-memory read port and an LDS issued on the
                    // critical path takes hundreds of cycles (tools/umma_bench3: 8 B/clk left for LDS under MMA load).
                    //   long skip (res1 = x0 planes in global memory; the producer has pulled the rows into L2)
                    uint4 xs_h[4], xs_l[4];
                    if (type == TC_LONGSKIP) {
                        const uint4 *x0 = reinterpret_cast<const uint4 *>(p.x0) + (size_t)n * 16 * p.L;
#pragma unroll
                        for (int h = 0; h < 4; ++h) {
                            xs_h[h] = inside ? __ldg(x0 + (size_t)(2 * h + egrp) * p.L + l) : make_uint4(0, 0, 0, 0);
                            xs_l[h] = (inside && !FAST) ? __ldg(x0 + (size_t)(8 + 2 * h + egrp) * p.L + l) : make_uint4(0, 0, 0, 0);
                        }
                    }
                    //   bias of the first chunk (the following chunks' are fetched one chunk ahead)
                    float4 bn0 = bias4[2 * egrp], bn1 = bias4[2 * egrp + 1];
                    mbar_wait(BAR_ACC_FULL(b), lseq & 1u);
                    tc_fence_after_sync();
                    if (tl_on) TL(3u, k, b, 1);
                    if (b == 1 && k == p.free_layer && !(k & 1)) {
                        // both blocks' MMAs of the last layer that reads buffer P have completed, and this layer's epilogue does
                        // not touch P (even index: its output goes to Q / global, its short skip comes from the stash): hand P
                        // to the producer before the epilogue math, so the next tile's input load overlaps it
                        if (lane == 0) mbar_arrive(BAR_P_FREE);
                    }
                    const uint32_t d_tmem = tmem0 + 128u * (uint32_t)accsel + lane_addr;
                    const uint32_t bar_free = BAR_ACC_FREE(accsel);
                    if (++accsel == 3) accsel = 0;
                    // stash: hi plane of slab c8 in columns 4 c8 .. 4 c8 + 3, lo plane at + 32
                    const uint32_t stash = tmem0 + TC_STASH_COL + 64u * (uint32_t)b + 4u * (uint32_t)egrp + lane_addr;
                    if (type == TC_LAST) {
                        // conv_last + SampleShuffle1D: mask[n][4 l + u] = D0[u] + 2^-11 D1[u] + bias[u]   (group 0 only)
                        uint32_t d[16];
                        if (egrp == 0) {
                            tmem_ld16(d_tmem, d);
                            tmem_ld_wait();
                        }
                        tc_fence_before_sync();
                        __syncwarp();
                        if (lane == 0) mbar_arrive(bar_free);
                        if (egrp == 0 && l >= l0 && l < l_end) {
                            float4 o;
                            const float li = FAST ? 0.0f : TC_LO_INV;      // one-pass mode: columns 8..11 hold X_hi W_lo only
                            o.x = (__uint_as_float(d[0]) + li * __uint_as_float(d[8])) + bn0.x;
                            o.y = (__uint_as_float(d[1]) + li * __uint_as_float(d[9])) + bn0.y;
                            o.z = (__uint_as_float(d[2]) + li * __uint_as_float(d[10])) + bn0.z;
                            o.w = (__uint_as_float(d[3]) + li * __uint_as_float(d[11])) + bn0.w;
                            reinterpret_cast<float4 *>(p.out_mask)[(size_t)n * p.L + l] = o;
                        }
                    } else {
                        const bool store_global = last_in_seg && (l >= l0) && (l < l_end);
                        const bool fast = FAST;
                        // chunk h = channels [8 c8, 8 c8 + 8), c8 = 2 h + egrp: columns 8 c8 of D0 and of D1 (at +64); the TMEM
                        // loads of chunk h + 1 are in flight while chunk h is processed
                        uint32_t n0[8], n1[8], ns[8];
                        tmem_ld8(d_tmem + 8u * (uint32_t)egrp, n0);
                        if (!fast) tmem_ld8(d_tmem + 64u + 8u * (uint32_t)egrp, n1);
                        auto ld_stash = [&](int h) {
                            uint32_t a[4], c[4] = {0u, 0u, 0u, 0u};
                            tmem_ld4(stash + 8u * (uint32_t)h, a);
                            if (!fast) tmem_ld4(stash + 32u + 8u * (uint32_t)h, c);
#pragma unroll
                            for (int i = 0; i < 4; ++i) { ns[i] = a[i]; ns[4 + i] = c[i]; }
                        };
                        if (from_stash) ld_stash(0);
#pragma unroll
                        for (int h = 0; h < 4; ++h) {                                  // 4 chunks of 8 channels
                            const int c8 = 2 * h + egrp;
                            uint32_t d0[8], d1[8], rs[8];
                            if (fast) {
#pragma unroll
                                for (int i = 0; i < 8; ++i) n1[i] = 0u;
                            }
                            if (!from_stash) {
#pragma unroll
                                for (int i = 0; i < 8; ++i) ns[i] = 0u;
                            }
                            tmem_ld_wait_dep(n0, n1);
                            tmem_ld_wait_dep(ns, n0);
                            if (tl_on && h == 0) TL(3u, k, b, 30);
#pragma unroll
                            for (int i = 0; i < 8; ++i) { d0[i] = n0[i]; d1[i] = n1[i]; rs[i] = ns[i]; }
                            const float4 b0 = bn0, b1 = bn1;
                            if (h < 3) {
                                tmem_ld8(d_tmem + 8u * (uint32_t)(c8 + 2), n0);
                                if (!fast) tmem_ld8(d_tmem + 64u + 8u * (uint32_t)(c8 + 2), n1);
                                if (from_stash) ld_stash(h + 1);
                                bn0 = bias4[2 * (c8 + 2)];
                                bn1 = bias4[2 * (c8 + 2) + 1];
                            } else {
                                tc_fence_before_sync();
                                __syncwarp();
                                if (lane == 0) mbar_arrive(bar_free);                  // accumulator read: hand it back
                            }
                            const float bv[8] = {b0.x, b0.y, b0.z, b0.w, b1.x, b1.y, b1.z, b1.w};
                            float v[8];
#pragma unroll
                            for (int i = 0; i < 8; ++i)
                                v[i] = (__uint_as_float(d0[i]) + TC_LO_INV * __uint_as_float(d1[i])) + bv[i];
                            if (type == TC_LEAKY) {
#pragma unroll
                                for (int i = 0; i < 8; ++i) v[i] = fmaxf(v[i], kLeakySlope * v[i]);
                            } else if (type == TC_RELU) {
#pragma unroll
                                for (int i = 0; i < 8; ++i) v[i] = fmaxf(v[i], 0.0f);
                            } else {
                                // short skip (packed fp16 pairs from the TMEM stash) or long skip (res1 = x0, vectors loaded above)
                                float fh[8], fl[8];
                                unpack8(from_stash ? make_uint4(rs[0], rs[1], rs[2], rs[3]) : xs_h[h], fh);
                                unpack8(from_stash ? make_uint4(rs[4], rs[5], rs[6], rs[7]) : xs_l[h], fl);
#pragma unroll
                                for (int i = 0; i < 8; ++i) v[i] += fh[i] + TC_LO_INV * fl[i];
                            }
                            if (!inside) {
#pragma unroll
                                for (int i = 0; i < 8; ++i) v[i] = 0.0f;               // per-layer zero padding outside the row
                            }
                            uint4 uh, ul;
                            split_pair(v[0], v[1], uh.x, ul.x);
                            split_pair(v[2], v[3], uh.y, ul.y);
                            split_pair(v[4], v[5], uh.z, ul.z);
                            split_pair(v[6], v[7], uh.w, ul.w);
                            if (!last_in_seg) {
                                *reinterpret_cast<uint4 *>(outbuf + c8 * TC_SLAB + r * 16) = uh;
                                if (!fast) *reinterpret_cast<uint4 *>(outbuf + TC_PLANE + c8 * TC_SLAB + r * 16) = ul;
                                // K-group h of this block's output is complete once both warp groups of all four quadrants arrive
                                fence_proxy_async_smem();
                                __syncwarp();
                                if (lane == 0) mbar_arrive(BAR_ACT(k, b, h));
                                if (tl_on) TL(3u, k, b, 32 + h);
                            } else if (store_global) {
                                uint4 *dst = reinterpret_cast<uint4 *>(p.out_planes) + (size_t)n * 16 * p.L;
                                dst[(size_t)c8 * p.L + l] = uh;
                                if (!fast) dst[(size_t)(8 + c8) * p.L + l] = ul;
                            }
                        }
                    }
                    if (b == 1 && k == p.free_layer && (k & 1)) {
                        // two-layer segment whose layer 1 takes its short skip from the input tile: P is free once those
                        // vectors have been consumed
                        __syncwarp();
                        if (lane == 0) mbar_arrive(BAR_P_FREE);
                    }
                    if (tl_on) TL(3u, k, b, 3);
                }
            }
        }
        if (tl_on) TL_FLUSH(3u);
    }

    // ---- teardown --------------------------------------------------------------------------------------
    tc_fence_before_sync();
    __syncthreads();
    if (warp == 1) {
        tc_fence_after_sync();
        tmem_dealloc(tmem0, 512);
    }
}

// ------------------------------------------------------------------------------------------------
// x0 = act(conv_first(x)) [+ E upsampled x scale, centred, zero outside]  ->  split planes.
// conv1 (1 -> 64, k9) has K = 9: CUDA-core FMAs, HBM-bound (4 B in, 256 B out per position).
// (models/stofnet.py:23,45 and :96,108-115; K = 3 without E is EDSR_1D's relu(conv_input(x)), models/edsr_1d.py:33)
// ------------------------------------------------------------------------------------------------
template <int K>
__global__ void __launch_bounds__(256)
x0_split_kernel(const float *__restrict__ x, const float *__restrict__ e, int N, int L, int P, int scale, int off,
                const float *__restrict__ w /*[K][64]*/, const float *__restrict__ b, uint4 *__restrict__ planes, int fast) {
    __shared__ float4 ws[K * 16];
    __shared__ float4 bs[16];
    // E of the (at most 5) pooled positions this block's 256 samples fall into, [slot][64 channels]: one shared-memory
    // broadcast read per 4 channels instead of 64 scalar global loads per thread
    __shared__ float4 es[6 * 16];
    const int n = blockIdx.y;
    const int l0 = blockIdx.x * 256;
    for (int i = threadIdx.x; i < K * 16; i += 256) ws[i] = reinterpret_cast<const float4 *>(w)[i];
    if (threadIdx.x < 16) bs[threadIdx.x] = reinterpret_cast<const float4 *>(b)[threadIdx.x];
    const int j0 = (l0 - off >= 0) ? (l0 - off) / scale : 0;                 // first pooled slot touched by this block
    if (e != nullptr) {
        for (int i = threadIdx.x; i < 6 * 64; i += 256) {
            const int slot = i >> 6, c = i & 63;
            const int j = j0 + slot;
            reinterpret_cast<float *>(es)[slot * 64 + c] = (j < P) ? __ldg(e + ((size_t)n * 64 + c) * P + j) : 0.0f;
        }
    }
    __syncthreads();
    const int l = l0 + threadIdx.x;
    if (l >= L) return;
    const float *xr = x + (size_t)n * L;
    float xv[K];
#pragma unroll
    for (int t = 0; t < K; ++t) {
        const int g = l + t - K / 2;
        xv[t] = (g >= 0 && g < L) ? __ldg(xr + g) : 0.0f;
    }
    const int j = l - off;
    const bool has_e = (e != nullptr) && j >= 0 && j < scale * P;
    const float4 *er = es + (has_e ? (j / scale - j0) : 0) * 16;
    uint4 *dst = planes + (size_t)n * 16 * L + l;
#pragma unroll 2
    for (int c8 = 0; c8 < 8; ++c8) {
        float v[8];
#pragma unroll
        for (int h = 0; h < 2; ++h) {
            // packed fp32 FMAs (fma.rn.f32x2, bit-identical to fmaf): half the FMA-pipe instructions of this FMA-bound loop
            const float4 bq = bs[2 * c8 + h];
            float2 a01 = make_float2(bq.x, bq.y), a23 = make_float2(bq.z, bq.w);
#pragma unroll
            for (int t = 0; t < K; ++t) {
                const float4 wv = ws[t * 16 + 2 * c8 + h];
                const float2 xx = make_float2(xv[t], xv[t]);
                a01 = __ffma2_rn(make_float2(wv.x, wv.y), xx, a01);
                a23 = __ffma2_rn(make_float2(wv.z, wv.w), xx, a23);
            }
            v[4 * h + 0] = fmaxf(a01.x, 0.f); v[4 * h + 1] = fmaxf(a01.y, 0.f);
            v[4 * h + 2] = fmaxf(a23.x, 0.f); v[4 * h + 3] = fmaxf(a23.y, 0.f);
            if (has_e) {
                const float4 ev = er[2 * c8 + h];
                v[4 * h + 0] += ev.x; v[4 * h + 1] += ev.y; v[4 * h + 2] += ev.z; v[4 * h + 3] += ev.w;
            }
        }
        // (hi, lo) planes with the packed conversions of the chain kernel's epilogue (same rounding and saturation as split_hi_lo)
        uint4 uh, ul;
        split_pair(v[0], v[1], uh.x, ul.x);
        split_pair(v[2], v[3], uh.y, ul.y);
        split_pair(v[4], v[5], uh.z, ul.z);
        split_pair(v[6], v[7], uh.w, ul.w);
        dst[(size_t)c8 * L] = uh;
        if (!fast) dst[(size_t)(8 + c8) * L] = ul;
    }
}

// fp32 C4L4 [n][c/4][l][4]  <->  split planes [n][plane][c8][l][8]
__global__ void __launch_bounds__(256)
c4l4_to_planes_kernel(const float4 *__restrict__ act, uint4 *__restrict__ planes, int N, int L) {
    const int n = blockIdx.z, c8 = blockIdx.y;
    const int l = blockIdx.x * 256 + threadIdx.x;
    if (l >= L) return;
    const float4 a0 = act[((size_t)n * 16 + 2 * c8) * L + l], a1 = act[((size_t)n * 16 + 2 * c8 + 1) * L + l];
    const float v[8] = {a0.x, a0.y, a0.z, a0.w, a1.x, a1.y, a1.z, a1.w};
    float hi[8], lo[8];
#pragma unroll
    for (int i = 0; i < 8; ++i) split_hi_lo(v[i], hi[i], lo[i]);
    uint4 uh, ul;
    uh.x = pack_half2(hi[0], hi[1]); uh.y = pack_half2(hi[2], hi[3]); uh.z = pack_half2(hi[4], hi[5]); uh.w = pack_half2(hi[6], hi[7]);
    ul.x = pack_half2(lo[0], lo[1]); ul.y = pack_half2(lo[2], lo[3]); ul.z = pack_half2(lo[4], lo[5]); ul.w = pack_half2(lo[6], lo[7]);
    planes[((size_t)n * 16 + c8) * L + l] = uh;
    planes[((size_t)n * 16 + 8 + c8) * L + l] = ul;
}

__global__ void __launch_bounds__(256)
planes_to_c4l4_kernel(const uint4 *__restrict__ planes, float4 *__restrict__ act, int N, int L) {
    const int n = blockIdx.z, c8 = blockIdx.y;
    const int l = blockIdx.x * 256 + threadIdx.x;
    if (l >= L) return;
    float fh[8], fl[8];
    unpack8(planes[((size_t)n * 16 + c8) * L + l], fh);
    unpack8(planes[((size_t)n * 16 + 8 + c8) * L + l], fl);
    float v[8];
#pragma unroll
    for (int i = 0; i < 8; ++i) v[i] = fh[i] + TC_LO_INV * fl[i];
    act[((size_t)n * 16 + 2 * c8) * L + l] = make_float4(v[0], v[1], v[2], v[3]);
    act[((size_t)n * 16 + 2 * c8 + 1) * L + l] = make_float4(v[4], v[5], v[6], v[7]);
}

// ------------------------------------------------------------------------------------------------ launchers
static unsigned long long *g_timeline = nullptr;
static int g_timeline_cap = 0;
void set_chain_timeline(void *dev_buf, int cap) { g_timeline = reinterpret_cast<unsigned long long *>(dev_buf); g_timeline_cap = cap; }

int launch_chain_tc(const TcLayer *layers, int count, const void *in, const void *x0, void *out_planes, float *out_mask,
                    int N, int L, cudaStream_t st, int fast) {
    STOF_REQUIRE(count >= 1 && count <= TC_MAX_LAYERS, STOF_E_INVALID, "chain_tc: %d layers per segment (max %d)", count, TC_MAX_LAYERS);
    static SmemAttrOnce attr_exact, attr_fast;
    { const int rc = fast ? ensure_dyn_smem(chain_tc_kernel<true>, TC_SMEM_BYTES, attr_fast)
                          : ensure_dyn_smem(chain_tc_kernel<false>, TC_SMEM_BYTES, attr_exact); if (rc) return rc; }
    TcParams p;
    memset(&p, 0, sizeof(p));
    int S = 0, Sp = 0;
    for (int k = 0; k < count; ++k) {
        p.layer[k] = layers[k];
        S += layers[k].pad;
        if (k > 0) Sp += layers[k].pad;
        STOF_REQUIRE(layers[k].type != TC_LAST || k == count - 1, STOF_E_INVALID, "chain_tc: conv_last must end a segment");
        STOF_REQUIRE(layers[k].type != TC_RES || k >= 1, STOF_E_INVALID, "chain_tc: a short-skip layer cannot start a segment");
        STOF_REQUIRE(layers[k].type != TC_RES || layers[k - 1].type != TC_RES, STOF_E_INVALID,
                     "chain_tc: two consecutive short-skip layers (one TMEM stash per block)");
        STOF_REQUIRE(layers[k].type != TC_LONGSKIP || x0 != nullptr, STOF_E_INVALID, "chain_tc: long skip needs x0");
    }
    const int OFF = S + 3;
    p.n_layers = count;
    int acc = 0;
    for (int k = 0; k < count; ++k) {
        acc += layers[k].pad;
        p.b_row[k] = OFF - acc;
    }
    p.r_in = OFF + 256;
    STOF_REQUIRE(p.r_in <= TC_R, STOF_E_INVALID, "chain_tc: segment too deep for the smem tile (%d rows)", p.r_in);
    p.l_shift = p.b_row[0] + Sp;
    p.t_out = 256 - 2 * Sp;
    p.tiles_per_row = (L + p.t_out - 1) / p.t_out;
    const long long total = (long long)N * p.tiles_per_row;
    STOF_REQUIRE(total < (1ll << 31), STOF_E_INVALID, "chain_tc: too many tiles");
    p.total_tiles = (int)total;
    p.N = N; p.L = L; p.fast = fast;
    p.n_stages = fast ? TC_MAX_STAGES : TC_STAGES;
    p.stage_stride = fast ? TC_STAGE_BYTES / 2 : TC_STAGE_BYTES;
    p.in = in; p.x0 = x0; p.out_planes = out_planes; p.out_mask = out_mask;
    p.timeline = g_timeline; p.timeline_cap = g_timeline_cap;
    // Buffer P holds the input tile; it is the operand of even layers, the output of odd layers that do not end the segment,
    // and the short-skip source of odd layers, copied to the TMEM stash by the MMA stream of the even layer before.
    // free_layer = last layer touching it (always even).
    const bool even = (count % 2) == 0;
    p.stash = 1;
    p.free_layer = even ? count - 2 : count - 1;
    // the producer requests the next tile's input after issuing this K-group of the last layer's weights: with an even layer
    // count buffer P is handed over by the epilogue of layer count-2, so the request can go out early in the last layer
    // (tuned on the timeline); otherwise (and in the one-pass mode with its deep ring) after the tile's last stage
    p.in_layer = count - 1;
    p.in_ks = (!fast && even && !(p.free_layer & 1)) ? 1 : 3;
    for (int k = 0; k < count; ++k)
        STOF_REQUIRE(layers[k].taps >= 1, STOF_E_INVALID, "chain_tc: layer without taps");
    const bool ends_last = layers[count - 1].type == TC_LAST;
    STOF_REQUIRE(ends_last ? out_mask != nullptr : out_planes != nullptr, STOF_E_INVALID, "chain_tc: missing output buffer");
    const int grid = (int)(total < kSMs ? total : kSMs);
    if (fast) chain_tc_kernel<true><<<grid, TC_THREADS, TC_SMEM_BYTES, st>>>(p);
    else chain_tc_kernel<false><<<grid, TC_THREADS, TC_SMEM_BYTES, st>>>(p);
    STOF_LAUNCH_CHECK();
    return STOF_OK;
}

int launch_x0_split(const float *x, const float *e, int N, int L, int P, int scale, int off, const float *w, const float *b,
                    void *planes, cudaStream_t st, int fast, int K) {
    dim3 grid((L + 255) / 256, N);
    STOF_REQUIRE(K == 9 || K == 3, STOF_E_UNSUPPORTED, "x0_split: first conv kernel size %d (built for 9 and 3)", K);
    if (K == 9) x0_split_kernel<9><<<grid, 256, 0, st>>>(x, e, N, L, P, scale, off, w, b, reinterpret_cast<uint4 *>(planes), fast);
    else x0_split_kernel<3><<<grid, 256, 0, st>>>(x, e, N, L, P, scale, off, w, b, reinterpret_cast<uint4 *>(planes), fast);
    STOF_LAUNCH_CHECK();
    return STOF_OK;
}

int launch_c4l4_to_planes(const float *act, void *planes, int N, int L, cudaStream_t st) {
    dim3 grid((L + 255) / 256, 8, N);
    c4l4_to_planes_kernel<<<grid, 256, 0, st>>>(reinterpret_cast<const float4 *>(act), reinterpret_cast<uint4 *>(planes), N, L);
    STOF_LAUNCH_CHECK();
    return STOF_OK;
}

int launch_planes_to_c4l4(const void *planes, float *act, int N, int L, cudaStream_t st) {
    dim3 grid((L + 255) / 256, 8, N);
    planes_to_c4l4_kernel<<<grid, 256, 0, st>>>(reinterpret_cast<const uint4 *>(planes), reinterpret_cast<float4 *>(act), N, L);
    STOF_LAUNCH_CHECK();
    return STOF_OK;
}

}  // namespace stof
