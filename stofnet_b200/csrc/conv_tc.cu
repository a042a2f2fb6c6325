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
//     layer k+1 only depends on blocks <= b of layer k and the tensor pipe never waits for a whole layer.
//   * weights stream through a 5-stage ring of 16 KB taps filled by bulk async copies (TMA engine) on mbarriers.
//
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
// barriers (8 B each): w_full[STAGES] w_empty[STAGES] acc_full[2] acc_empty[2] act_ready[TC_MAX_LAYERS][2] in_full p_free
constexpr int TC_NBAR = 2 * TC_MAX_STAGES + 4 + 2 * TC_MAX_LAYERS + 4;       // ... in_full[2] p_free[2]
constexpr int TC_OFF_TMEM = TC_OFF_BAR + TC_NBAR * 8;
constexpr int TC_SMEM_BYTES = TC_OFF_TMEM + 16;

static_assert(TC_SMEM_BYTES <= 232448, "chain kernel exceeds the 227 KB shared-memory limit");

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

// All taps of one layer for one 128-row block, issued by a single thread.  Layer kinds (upload_tc in stof_api.cu):
//   64 -> 64 conv : taps [c8][W_hi 64 | W_lo 64][8] (one-pass mode: [c8][W_hi 64][8]); D0 = columns 0-63, D1 = 64-127
//   conv_last     : taps [c8][32 rows][8]: rows 0-3 W_hi, 8-11 W_lo, 16-19 W_hi (one-pass: [c8][16][8]); D0 = 0-3, D1 = 8-11
// Descriptor low word = (addr >> 4) | (LBO >> 4) << 16; the high word (SBO = 128 B, version 1) is constant.
template <bool FAST, bool LAST>
__device__ __forceinline__ void issue_layer_taps(uint32_t a_base, uint32_t w_base, uint32_t d_tmem, int taps, int s, uint32_t ph,
                                                 uint32_t bar_w_full0, uint32_t bar_w_empty0, uint32_t bar_acc_full) {
    constexpr uint32_t kDescHi = (128u >> 4) | (1u << 14);
    constexpr uint32_t W_ROWS = LAST ? (FAST ? 16u : 32u) : (FAST ? 64u : 128u);    // K-chunk stride of a weight stage / 16
    constexpr uint32_t LO_ROW = LAST ? 16u : 0u;                                     // W_hi rows of the X_lo product
    constexpr uint32_t LO_COL = LAST ? 8u : 64u;                                     // its accumulator columns
    constexpr uint32_t STRIDE16 = (uint32_t)(FAST ? TC_STAGE_BYTES / 2 : TC_STAGE_BYTES) >> 4;
    constexpr int NST = FAST ? TC_MAX_STAGES : TC_STAGES;
    constexpr uint32_t idesc_a = umma_idesc_f16(128, LAST ? 16 : (FAST ? 64 : 128));
    constexpr uint32_t idesc_b = umma_idesc_f16(128, LAST ? 16 : 64);
    for (int t = 0; t < taps; ++t) {
        mbar_wait(bar_w_full0 + 8u * (uint32_t)s, ph);
        tc_fence_after_sync();
        const uint32_t a_lo = a_base + (uint32_t)t;
        const uint32_t w_lo = (w_base + (uint32_t)s * STRIDE16) | (W_ROWS << 16);
#pragma unroll
        for (int ks = 0; ks < 4; ++ks) {
            umma_f16(d_tmem, make_desc(a_lo + 2 * ks * TC_R, kDescHi), make_desc(w_lo + 2 * ks * W_ROWS, kDescHi), idesc_a,
                     (t > 0 || ks > 0) ? 1u : 0u);
            if (!FAST)
                umma_f16(d_tmem + LO_COL, make_desc(a_lo + (8 + 2 * ks) * TC_R, kDescHi),
                         make_desc(w_lo + 2 * ks * W_ROWS + LO_ROW, kDescHi), idesc_b, 1u);
        }
        umma_commit(bar_w_empty0 + 8u * (uint32_t)s);                            // this stream is done with the stage
        if (++s == NST) { s = 0; ph ^= 1u; }
    }
    umma_commit(bar_acc_full);
}

template <bool FAST, bool GRAD>
__global__ void __launch_bounds__(TC_THREADS, 1) chain_tc_kernel(const TcParams p) {
    extern __shared__ __align__(1024) uint8_t smem[];
    const uint32_t sbase = smem_u32(smem);
    const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
    int tl_cnt = 0;
    const uint32_t bar0 = sbase + TC_OFF_BAR;
    auto BAR_W_FULL = [&](int s) { return bar0 + 8 * s; };
    auto BAR_W_EMPTY = [&](int s) { return bar0 + 8 * (TC_MAX_STAGES + s); };
    auto BAR_ACC_FULL = [&](int b) { return bar0 + 8 * (2 * TC_MAX_STAGES + b); };
    auto BAR_ACC_EMPTY = [&](int b) { return bar0 + 8 * (2 * TC_MAX_STAGES + 2 + b); };
    auto BAR_ACT = [&](int k, int b) { return bar0 + 8 * (2 * TC_MAX_STAGES + 4 + 2 * k + b); };
    // Input buffer(s): a one-layer segment never touches buffer Q, so it alternates its tiles' inputs between P and Q and the next
    // tile's load runs behind the whole current tile (p.dbl; barriers per buffer: a single one could run two phases ahead of a
    // parity wait).  Otherwise buffer P only, barrier set 0.
    const uint32_t BAR_IN_FULL0 = bar0 + 8 * (2 * TC_MAX_STAGES + 4 + 2 * TC_MAX_LAYERS);
    const bool dbl = (p.dbl != 0);
    auto in_sel = [&](int j) { return dbl ? (j & 1) : 0; };                         // input buffer / barrier set of this CTA's j-th tile
    auto in_phase = [&](int j) { return (uint32_t)(dbl ? (j >> 1) : j) & 1u; };     // parity of that buffer's phase for tile j
    auto BAR_IN_FULL_OF = [&](int sel) { return BAR_IN_FULL0 + 16u * (uint32_t)sel; };
    auto BAR_P_FREE_OF = [&](int sel) { return BAR_IN_FULL0 + 8u + 16u * (uint32_t)sel; };
    const uint32_t BAR_IN_FULL = BAR_IN_FULL0;
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
            mbar_init(BAR_W_EMPTY(s), 2);                 // one issuing warp per block consumes a tap
        }
        for (int b = 0; b < 2; ++b) {
            mbar_init(BAR_ACC_FULL(b), 1);
            mbar_init(BAR_ACC_EMPTY(b), TC_EPI_WARPS);
        }
        for (int k = 0; k < TC_MAX_LAYERS; ++k) {
            mbar_init(BAR_ACT(k, 0), TC_EPI_WARPS);
            mbar_init(BAR_ACT(k, 1), TC_EPI_WARPS);
        }
        for (int sel = 0; sel < 2; ++sel) {
            mbar_init(BAR_IN_FULL_OF(sel), 1);
            mbar_init(BAR_P_FREE_OF(sel), TC_EPI_WARPS);
        }
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
        // ends, so the load of the next tile hides behind the last layer.  Weights: every tap of every layer in issue order.
        auto issue_input = [&](int tile, int j) {
            const int n = tile / p.tiles_per_row;
            const int l_org = (tile - n * p.tiles_per_row) * p.t_out - p.l_shift;
            const int sel = in_sel(j);
            const uint32_t in_off = sel ? TC_OFF_Q : TC_OFF_P;
            const uint32_t bar_in = BAR_IN_FULL_OF(sel);
            // the buffer's previous tile (j - 1, or j - 2 when the tiles alternate) no longer touches it
            if (dbl) { if (j > 1) mbar_wait(BAR_P_FREE_OF(sel), in_phase(j - 2)); }
            else if (j > 0) mbar_wait(BAR_P_FREE, (uint32_t)(j - 1) & 1u);
            if (lane == 0) TL(0u, 0, 0, 1);
            const int r_lo = max(0, -l_org), r_hi = min(r_in, p.L - l_org);
            const int nslab = FAST ? 8 : 16;               // one-pass fp16 mode: hi plane only
            // rows outside the waveform are the conv's zero padding (only tiles at the ends of a row have any)
            if (r_lo > 0 || r_hi < r_in) {
                for (int slab = 0; slab < nslab; ++slab) {
                    uint8_t *sl = smem + in_off + slab * TC_SLAB;
                    for (int r = lane; r < r_lo; r += 32) *reinterpret_cast<uint4 *>(sl + r * 16) = make_uint4(0, 0, 0, 0);
                    for (int r = r_hi + lane; r < r_in; r += 32) *reinterpret_cast<uint4 *>(sl + r * 16) = make_uint4(0, 0, 0, 0);
                }
            }
            fence_proxy_async_smem();
            __syncwarp();
            if (lane == 0) {
                const uint32_t bytes = (uint32_t)(r_hi - r_lo) * 16u;
                mbar_arrive_expect_tx(bar_in, bytes * (uint32_t)nslab);
                const __half *src = reinterpret_cast<const __half *>(p.in) + (size_t)n * 16 * p.L * 8;
                for (int slab = 0; slab < nslab; ++slab)
                    bulk_g2s(sbase + in_off + slab * TC_SLAB + r_lo * 16, src + ((size_t)slab * p.L + l_org + r_lo) * 8, bytes, bar_in);
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
                for (int t = 0; t < ly.taps; ++t) {
                    if (lane == 0) {
                        mbar_wait(BAR_W_EMPTY(s), ph ^ 1u);
                        const uint32_t sbytes = (uint32_t)(FAST ? ly.stage_bytes_fast : ly.stage_bytes);
                        mbar_arrive_expect_tx(BAR_W_FULL(s), sbytes);
                        bulk_g2s(sbase + TC_OFF_W + s * stage_stride,
                                 reinterpret_cast<const uint8_t *>(FAST ? ly.w_fast : ly.w) + (size_t)t * sbytes, sbytes, BAR_W_FULL(s));
                    }
                    if (++s == n_stages) { s = 0; ph ^= 1u; }
                    if (k == p.in_layer && t == p.in_tap && next < p.total_tiles) {
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
        // streams (different accumulators) hide each other's issue gaps, and one block's MMAs cover the other block's
        // epilogue; the weight ring bounds the lag between the blocks to TC_STAGES taps.
        // Each warp runs warp-uniform control flow (descriptors stay in uniform registers); one elected lane issues.
        // Descriptor low word = (addr >> 4) | (LBO >> 4) << 16; the high word (SBO = 128 B, version 1) is constant.
        constexpr uint32_t kALoBase = (uint32_t)(TC_SLAB >> 4) << 16;            // A: LBO = one slab
        const int b = warp - 1;
        const uint32_t role = 1u + (uint32_t)b;
        const bool tl_on = (lane == 0);
        const uint32_t d_tmem = tmem0 + (uint32_t)b * 128u;
        int s = 0;
        uint32_t ph = 0, lseq = 0;
        int it = 0;
        for (int tile = blockIdx.x; tile < p.total_tiles; tile += gridDim.x, ++it) {
            if (tl_on) TL(role, 0, b, 1);
            mbar_wait(BAR_IN_FULL_OF(in_sel(it)), in_phase(it));
            tc_fence_after_sync();
            if (tl_on) TL(role, 0, b, 2);
            for (int k = 0; k < n_layers; ++k, ++lseq) {
                const TcLayer &ly = p.layer[k];
                const int taps = ly.taps;
                const uint32_t a_base = (((sbase + (((k & 1) ^ in_sel(it)) ? TC_OFF_Q : TC_OFF_P)) >> 4) | kALoBase) +
                                        (uint32_t)(p.b_row[k] - ly.pad + 128 * b);
                if (tl_on) TL(role, k, b, 3);
                mbar_wait(BAR_ACC_EMPTY(b), (lseq & 1u) ^ 1u);
                if (k > 0) {
                    mbar_wait(BAR_ACT(k - 1, 0), (uint32_t)(it & 1));
                    if (b == 1) mbar_wait(BAR_ACT(k - 1, 1), (uint32_t)(it & 1));
                }
                tc_fence_after_sync();
                // the layer's taps are issued by ONE elected lane in a single region (it also polls the weight ring), with every
                // per-layer quantity a compile-time constant of the layer kind: per tap only the A row shift and the ring stage change
                if (elect_one()) {
                    const uint32_t w_base = (sbase + TC_OFF_W) >> 4;
                    if (ly.type == TC_LAST)
                        issue_layer_taps<FAST, true>(a_base, w_base, d_tmem, taps, s, ph, BAR_W_FULL(0), BAR_W_EMPTY(0), BAR_ACC_FULL(b));
                    else
                        issue_layer_taps<FAST, false>(a_base, w_base, d_tmem, taps, s, ph, BAR_W_FULL(0), BAR_W_EMPTY(0), BAR_ACC_FULL(b));
                }
                __syncwarp();
                s += taps;
                while (s >= n_stages) { s -= n_stages; ph ^= 1u; }
                if (tl_on) TL(role, k, b, 4);
            }
        }
        if (tl_on) TL_FLUSH(role);
    } else {
        // =================================== epilogue ====================================================
        // 8 warps: warp quadrant (warp & 3) selects the 32 TMEM lanes, group (0/1) the 32-channel half of the block.
        // The blocks are processed one after the other, which staggers their layer pipelines by one epilogue: while block
        // 0's epilogue runs, block 1's MMAs keep the tensor pipe and the shared-memory ports busy, and vice versa.
        const int quad = warp & 3;
        const int egrp = (warp - 3) >> 2;
        const int trow = 32 * quad + lane;                       // TMEM lane = row within the 128-block
        const uint32_t lane_addr = (uint32_t)(32 * quad) << 16;
        const bool tl_on = (warp == 3) && (lane == 0);
        const bool free_early = (p.stash == 0);     // free_layer has an even index: it reads P as operand only
        // gradient mode (training step, one TC_NONE layer per launch): out_stored = (conv(in_stored) * g_sk [+ aux_stored * g_ss])
        // [* lrelu'(aux)], the powers of two chosen so that a bound of max |out| lands in [2^10, 2^11)
        float g_sk = 1.0f, g_ss = 1.0f, g_amax = 0.0f;
        float g_bsum[32];
        int g_mode = 0;
        if (GRAD) {
            g_mode = p.grad.mode;
#pragma unroll
            for (int i = 0; i < 32; ++i) g_bsum[i] = 0.0f;
            const unsigned in_bits = __ldg(&p.grad.g_in->amax_bits);
            const int e_in = __ldg(&p.grad.g_in->E);
            auto pow2 = [](int d) { d = d < -126 ? -126 : (d > 127 ? 127 : d); return __uint_as_float((unsigned)(d + 127) << 23); };
            float bound = __uint_as_float(in_bits) * pow2(-e_in);
            int e_aux = 0;
            if (g_mode == 2) {
                e_aux = __ldg(&p.grad.g_aux->E);
                bound += __uint_as_float(__ldg(&p.grad.g_aux->amax_bits)) * pow2(-e_aux);
            }
            const int eb = (int)((__float_as_uint(bound) >> 23) & 255u);
            const int e_out = (eb == 0 || eb == 255) ? e_in : 10 - (eb - 127);
            g_sk = pow2(e_out - e_in);
            g_ss = pow2(e_out - e_aux);
            if (blockIdx.x == 0 && warp == 3 && lane == 0) p.grad.g_out->E = e_out;
        }
        uint32_t lseq = 0;
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
                const float *bias = s_bias + 64 * k;
                for (int b = 0; b < 2; ++b) {
                    const int r = p.b_row[k] + 128 * b + trow;
                    const int l = l_org + r;
                    const bool inside = (l >= 0) && (l < p.L);
                    // long skip (res1 = x0 planes in global memory): issue the loads before waiting for the accumulator so
                    // that their latency hides behind the block's MMAs (the producer has pulled the rows into L2)
                    uint4 xs_h[4], xs_l[4];
                    if (type == TC_LONGSKIP || (GRAD && g_mode >= 2)) {
                        const uint4 *x0 = reinterpret_cast<const uint4 *>(GRAD ? p.grad.aux : p.x0) + (size_t)n * 16 * p.L;
#pragma unroll
                        for (int h = 0; h < 4; ++h) {
                            xs_h[h] = inside ? __ldg(x0 + (size_t)(4 * egrp + h) * p.L + l) : make_uint4(0, 0, 0, 0);
                            xs_l[h] = (inside && !FAST) ? __ldg(x0 + (size_t)(8 + 4 * egrp + h) * p.L + l) : make_uint4(0, 0, 0, 0);
                        }
                    }
                    mbar_wait(BAR_ACC_FULL(b), lseq & 1u);
                    tc_fence_after_sync();
                    if (tl_on) TL(3u, k, b, 1);
                    if (free_early && b == 1 && k == p.free_layer) {
                        // both blocks' MMAs of the last layer that reads buffer P have completed, and this layer's epilogue does
                        // not touch P (its output goes to Q / global, no TMEM stash): hand P to the producer before the epilogue
                        // math, so the next tile's input load overlaps it
                        if (lane == 0) mbar_arrive(BAR_P_FREE_OF(in_sel(it)));
                    }
                    const uint32_t d_tmem = tmem0 + (uint32_t)b * 128u + lane_addr;
                    const uint32_t stash = tmem0 + 256u + 64u * (uint32_t)b + 32u * (uint32_t)egrp + lane_addr;
                    if (type == TC_LAST) {
                        // conv_last + SampleShuffle1D: mask[n][4 l + u] = D0[u] + 2^-11 D1[u] + bias[u]   (group 0 only)
                        uint32_t d[16];
                        if (egrp == 0) {
                            tmem_ld16(d_tmem, d);
                            tmem_ld_wait();
                        }
                        tc_fence_before_sync();
                        __syncwarp();
                        if (lane == 0) mbar_arrive(BAR_ACC_EMPTY(b));
                        if (egrp == 0 && l >= l0 && l < l_end) {
                            float4 o;
                            const float li = FAST ? 0.0f : TC_LO_INV;      // one-pass mode: columns 8..11 hold X_hi W_lo only
                            o.x = (__uint_as_float(d[0]) + li * __uint_as_float(d[8])) + bias[0];
                            o.y = (__uint_as_float(d[1]) + li * __uint_as_float(d[9])) + bias[1];
                            o.z = (__uint_as_float(d[2]) + li * __uint_as_float(d[10])) + bias[2];
                            o.w = (__uint_as_float(d[3]) + li * __uint_as_float(d[11])) + bias[3];
                            reinterpret_cast<float4 *>(p.out_mask)[(size_t)n * p.L + l] = o;
                        }
                    } else {
                        const bool store_global = last_in_seg && (l >= l0) && (l < l_end);
                        const bool from_stash = (type == TC_RES) && last_in_seg && (p.stash != 0);
                        // this group's 32 channels: columns [32 egrp, +32) of D0 and of D1 (at +64)
                        uint32_t d0[32], d1[32];
                        tmem_ld16(d_tmem + 32 * egrp, *reinterpret_cast<uint32_t(*)[16]>(&d0[0]));
                        tmem_ld16(d_tmem + 32 * egrp + 16, *reinterpret_cast<uint32_t(*)[16]>(&d0[16]));
                        const bool fast = FAST;
                        if (!fast) {
                            tmem_ld16(d_tmem + 64 + 32 * egrp, *reinterpret_cast<uint32_t(*)[16]>(&d1[0]));
                            tmem_ld16(d_tmem + 64 + 32 * egrp + 16, *reinterpret_cast<uint32_t(*)[16]>(&d1[16]));
                        } else {
#pragma unroll
                            for (int i = 0; i < 32; ++i) d1[i] = 0u;
                        }
                        tmem_ld_wait();
                        tc_fence_before_sync();
                        __syncwarp();
                        if (lane == 0) mbar_arrive(BAR_ACC_EMPTY(b));              // accumulator read: hand it back
#pragma unroll
                        for (int h = 0; h < 4; ++h) {                                  // 4 chunks of 8 channels
                            const int c8 = 4 * egrp + h;
                            float v[8];
#pragma unroll
                            for (int i = 0; i < 8; ++i)
                                v[i] = (__uint_as_float(d0[8 * h + i]) + TC_LO_INV * __uint_as_float(d1[8 * h + i])) + bias[8 * c8 + i];
                            if (type == TC_LEAKY) {
#pragma unroll
                                for (int i = 0; i < 8; ++i) v[i] = fmaxf(v[i], kLeakySlope * v[i]);
                            } else if (type == TC_RELU) {
#pragma unroll
                                for (int i = 0; i < 8; ++i) v[i] = fmaxf(v[i], 0.0f);
                            } else if (type == TC_NONE) {
                                // conv + bias only (data gradient of a conv: no activation, no skip)
                                if (GRAD) {
#pragma unroll
                                    for (int i = 0; i < 8; ++i) v[i] *= g_sk;
                                    if (g_mode == 2) {
                                        float fh[8], fl[8];
                                        unpack8(xs_h[h], fh);
                                        unpack8(xs_l[h], fl);
#pragma unroll
                                        for (int i = 0; i < 8; ++i) v[i] = fmaf(fh[i] + TC_LO_INV * fl[i], g_ss, v[i]);
                                    } else if (g_mode == 3) {
                                        const uint32_t hw[4] = {xs_h[h].x, xs_h[h].y, xs_h[h].z, xs_h[h].w};
                                        const uint32_t lw[4] = {xs_l[h].x, xs_l[h].y, xs_l[h].z, xs_l[h].w};
#pragma unroll
                                        for (int i = 0; i < 8; ++i) {
                                            const uint32_t hv = (hw[i >> 1] >> (16 * (i & 1))) & 0xffffu, lv = (lw[i >> 1] >> (16 * (i & 1))) & 0xffffu;
                                            // a > 0 <=> its hi plane is positive, or zero with a positive lo plane
                                            const bool pos = ((hv & 0x7fffu) != 0u) ? ((hv & 0x8000u) == 0u)
                                                                                    : (((lv & 0x7fffu) != 0u) && ((lv & 0x8000u) == 0u));
                                            v[i] *= pos ? 1.0f : kLeakySlope;
                                        }
                                    }
                                    if (store_global) {
#pragma unroll
                                        for (int i = 0; i < 8; ++i) { g_amax = fmaxf(g_amax, fabsf(v[i])); g_bsum[8 * h + i] += v[i]; }
                                    }
                                }
                            } else if (from_stash) {
                                // short skip of the segment's last layer: fp32 copy parked in TMEM by the previous layer's epilogue
                                uint32_t rs[8];
                                tmem_ld8(stash + 8 * h, rs);
                                tmem_ld_wait();
#pragma unroll
                                for (int i = 0; i < 8; ++i) v[i] += __uint_as_float(rs[i]);
                            } else {
                                // residual: short skip = value stored at this row of the output buffer (x two layers back);
                                // long skip (res1) = the x0 vectors loaded above
                                uint4 rh, rl;
                                if (type == TC_RES) {
                                    rh = *reinterpret_cast<const uint4 *>(outbuf + c8 * TC_SLAB + r * 16);
                                    rl = fast ? make_uint4(0, 0, 0, 0) : *reinterpret_cast<const uint4 *>(outbuf + TC_PLANE + c8 * TC_SLAB + r * 16);
                                } else {
                                    rh = xs_h[h];
                                    rl = xs_l[h];
                                }
                                float fh[8], fl[8];
                                unpack8(rh, fh);
                                unpack8(rl, fl);
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
                            } else if (store_global) {
                                uint4 *dst = reinterpret_cast<uint4 *>(p.out_planes) + (size_t)n * 16 * p.L;
                                dst[(size_t)c8 * p.L + l] = uh;
                                if (!fast) dst[(size_t)(8 + c8) * p.L + l] = ul;
                            }
                        }
                        if (!last_in_seg) {
                            fence_proxy_async_smem();
                            __syncwarp();
                            if (lane == 0) mbar_arrive(BAR_ACT(k, b));
                        }
                        if (p.stash != 0 && k == n_layers - 2) {
                            // park the last layer's short skip (rows of buffer P, the output of layer k - 1) as fp32 in spare
                            // TMEM columns, so that P can take the next tile's input while the last layer still runs
                            const int rr = p.b_row[n_layers - 1] + 128 * b + trow;
                            const uint8_t *pb = smem + TC_OFF_P + rr * 16;
#pragma unroll
                            for (int h2 = 0; h2 < 2; ++h2) {
                                uint32_t rs[16];
#pragma unroll
                                for (int h = 0; h < 2; ++h) {
                                    const int c8 = 4 * egrp + 2 * h2 + h;
                                    float fh[8], fl[8];
                                    unpack8(*reinterpret_cast<const uint4 *>(pb + c8 * TC_SLAB), fh);
                                    unpack8(FAST ? make_uint4(0, 0, 0, 0) : *reinterpret_cast<const uint4 *>(pb + TC_PLANE + c8 * TC_SLAB), fl);
#pragma unroll
                                    for (int i = 0; i < 8; ++i) rs[8 * h + i] = __float_as_uint(fh[i] + TC_LO_INV * fl[i]);
                                }
                                tmem_st16(stash + 16 * h2, rs);
                            }
                            tmem_st_wait();
                        }
                    }
                    if (tl_on) TL(3u, k, b, 3);
                }
                if (k == p.free_layer && !free_early) {
                    // every read of buffer P for this tile is done (this warp's share)
                    __syncwarp();
                    if (lane == 0) mbar_arrive(BAR_P_FREE);
                }
            }
        }
        if (tl_on) TL_FLUSH(3u);
        if (GRAD) {
            // max |stored output| of this launch; bias-gradient sums of this CTA: lanes (rows) by shuffle tree, then the four
            // quadrant warps of each channel half in order (after the CTA-wide barrier below: the operand buffers are free then)
#pragma unroll
            for (int o = 16; o > 0; o >>= 1) g_amax = fmaxf(g_amax, __shfl_xor_sync(0xffffffffu, g_amax, o));
            if (lane == 0 && g_amax > 0.0f) atomicMax(&p.grad.g_out->amax_bits, __float_as_uint(g_amax));
#pragma unroll
            for (int i = 0; i < 32; ++i)
#pragma unroll
                for (int o = 16; o > 0; o >>= 1) g_bsum[i] += __shfl_xor_sync(0xffffffffu, g_bsum[i], o);
            tc_fence_before_sync();
            __syncthreads();
            float *red = reinterpret_cast<float *>(smem + TC_OFF_P);                   // [8 epilogue warps][32]
            float mine = 0.0f;
#pragma unroll
            for (int i = 0; i < 32; ++i) if (lane == i) mine = g_bsum[i];
            red[(warp - 3) * 32 + lane] = mine;
        }
    }
    if (GRAD) {
        if (warp < 3) { tc_fence_before_sync(); __syncthreads(); }
        __syncthreads();
        if (p.grad.bias_partial != nullptr && threadIdx.x < 64) {
            const int egrp = threadIdx.x >> 5, c = threadIdx.x & 31;
            const float *red = reinterpret_cast<const float *>(smem + TC_OFF_P) + (4 * egrp) * 32 + c;
            p.grad.bias_partial[(size_t)blockIdx.x * 64 + threadIdx.x] = ((red[0] + red[32]) + red[64]) + red[96];
        }
        if (warp == 1) {
            tc_fence_after_sync();
            tmem_dealloc(tmem0, 512);
        }
        return;
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
// Thread = X0_Q positions (256 apart, so that a warp's 16-byte stores stay contiguous) x all 64 channels: every weight vector read
// from shared memory feeds X0_Q packed FMAs (with one position per thread the kernel waited on shared-memory loads: short-scoreboard
// stalls 8.9 per issued instruction, 3.6 TB/s of writes; ncu profiles/r8c).
constexpr int X0_Q = 4;
template <int K>
__global__ void __launch_bounds__(256)
x0_split_kernel(const float *__restrict__ x, const float *__restrict__ e, int N, int L, int P, int scale, int off,
                const float *__restrict__ w /*[K][64]*/, const float *__restrict__ b, uint4 *__restrict__ planes, int fast) {
    __shared__ float4 ws[K * 16];
    __shared__ float4 bs[16];
    // E of the pooled positions this block's 1,024 samples fall into (at most 14 windows of 80 + 1), [slot][64 channels]: one
    // shared-memory broadcast read per 4 channels instead of 64 scalar global loads per thread
    constexpr int SLOTS = 16;
    __shared__ float4 es[SLOTS * 16];
    const int n = blockIdx.y;
    const int l0 = blockIdx.x * (256 * X0_Q);
    for (int i = threadIdx.x; i < K * 16; i += 256) ws[i] = reinterpret_cast<const float4 *>(w)[i];
    if (threadIdx.x < 16) bs[threadIdx.x] = reinterpret_cast<const float4 *>(b)[threadIdx.x];
    const int j0 = (l0 - off >= 0) ? (l0 - off) / scale : 0;                 // first pooled slot touched by this block
    if (e != nullptr) {
        for (int i = threadIdx.x; i < SLOTS * 64; i += 256) {
            const int slot = i >> 6, c = i & 63;
            const int j = j0 + slot;
            reinterpret_cast<float *>(es)[slot * 64 + c] = (j < P) ? __ldg(e + ((size_t)n * 64 + c) * P + j) : 0.0f;
        }
    }
    __syncthreads();
    const float *xr = x + (size_t)n * L;
    float xv[X0_Q][K];
    bool has_e[X0_Q];
    const float4 *er[X0_Q];
#pragma unroll
    for (int q = 0; q < X0_Q; ++q) {
        const int l = l0 + 256 * q + threadIdx.x;
#pragma unroll
        for (int t = 0; t < K; ++t) {
            const int g = l + t - K / 2;
            xv[q][t] = (g >= 0 && g < L) ? __ldg(xr + g) : 0.0f;
        }
        const int j = l - off;
        has_e[q] = (e != nullptr) && j >= 0 && j < scale * P;
        er[q] = es + (has_e[q] ? (j / scale - j0) : 0) * 16;
    }
    uint4 *dst = planes + (size_t)n * 16 * L + l0 + threadIdx.x;
#pragma unroll 1
    for (int c8 = 0; c8 < 8; ++c8) {
        float v[X0_Q][8];
#pragma unroll
        for (int h = 0; h < 2; ++h) {
            // packed fp32 FMAs (fma.rn.f32x2, bit-identical to fmaf): half the FMA-pipe instructions of this FMA-bound loop
            const float4 bq = bs[2 * c8 + h];
            float2 a01[X0_Q], a23[X0_Q];
#pragma unroll
            for (int q = 0; q < X0_Q; ++q) { a01[q] = make_float2(bq.x, bq.y); a23[q] = make_float2(bq.z, bq.w); }
#pragma unroll
            for (int t = 0; t < K; ++t) {
                const float4 wv = ws[t * 16 + 2 * c8 + h];
#pragma unroll
                for (int q = 0; q < X0_Q; ++q) {
                    const float2 xx = make_float2(xv[q][t], xv[q][t]);
                    a01[q] = __ffma2_rn(make_float2(wv.x, wv.y), xx, a01[q]);
                    a23[q] = __ffma2_rn(make_float2(wv.z, wv.w), xx, a23[q]);
                }
            }
#pragma unroll
            for (int q = 0; q < X0_Q; ++q) {
                v[q][4 * h + 0] = fmaxf(a01[q].x, 0.f); v[q][4 * h + 1] = fmaxf(a01[q].y, 0.f);
                v[q][4 * h + 2] = fmaxf(a23[q].x, 0.f); v[q][4 * h + 3] = fmaxf(a23[q].y, 0.f);
                if (has_e[q]) {
                    const float4 ev = er[q][2 * c8 + h];
                    v[q][4 * h + 0] += ev.x; v[q][4 * h + 1] += ev.y; v[q][4 * h + 2] += ev.z; v[q][4 * h + 3] += ev.w;
                }
            }
        }
#pragma unroll
        for (int q = 0; q < X0_Q; ++q) {
            if (l0 + 256 * q + (int)threadIdx.x >= L) continue;
            // (hi, lo) planes with the packed conversions of the chain kernel's epilogue (same rounding and saturation as split_hi_lo)
            uint4 uh, ul;
            split_pair(v[q][0], v[q][1], uh.x, ul.x);
            split_pair(v[q][2], v[q][3], uh.y, ul.y);
            split_pair(v[q][4], v[q][5], uh.z, ul.z);
            split_pair(v[q][6], v[q][7], uh.w, ul.w);
            // streaming stores: 256 B per position that nothing re-reads before the tensor has passed through L2 many times over
            __stcs(dst + (size_t)c8 * L + 256 * q, uh);
            if (!fast) __stcs(dst + (size_t)(8 + c8) * L + 256 * q, ul);
        }
    }
}

// fp32 C4L4 [n][c/4][l][4]  <->  split planes [n][plane][c8][l][8], optionally scaled by a power of two derived on the device
// from the tensor's max |value| (gradients: see kernels.cuh)
__device__ __forceinline__ float plane_scale(const unsigned *absmax_bits) {
    if (absmax_bits == nullptr) return 1.0f;
    const unsigned b = __ldg(absmax_bits);
    const int e = (int)((b >> 23) & 255u);                       // biased exponent of max |v|: max in [2^(e-127), 2^(e-126))
    if (e == 0 || e == 255) return 1.0f;                         // all zero (or subnormal) / inf, nan: leave as is
    return __uint_as_float((unsigned)(127 + 10 - (e - 127)) << 23);    // 2^(10 - (e - 127)): scaled max in [2^10, 2^11)
}

__global__ void __launch_bounds__(256)
c4l4_to_planes_kernel(const float4 *__restrict__ act, uint4 *__restrict__ planes, int N, int L, const unsigned *__restrict__ absmax_bits) {
    const int n = blockIdx.z, c8 = blockIdx.y;
    const int l = blockIdx.x * 256 + threadIdx.x;
    if (l >= L) return;
    const float sc = plane_scale(absmax_bits);
    const float4 a0 = act[((size_t)n * 16 + 2 * c8) * L + l], a1 = act[((size_t)n * 16 + 2 * c8 + 1) * L + l];
    const float v[8] = {a0.x * sc, a0.y * sc, a0.z * sc, a0.w * sc, a1.x * sc, a1.y * sc, a1.z * sc, a1.w * sc};
    float hi[8], lo[8];
#pragma unroll
    for (int i = 0; i < 8; ++i) split_hi_lo(v[i], hi[i], lo[i]);
    uint4 uh, ul;
    uh.x = pack_half2(hi[0], hi[1]); uh.y = pack_half2(hi[2], hi[3]); uh.z = pack_half2(hi[4], hi[5]); uh.w = pack_half2(hi[6], hi[7]);
    ul.x = pack_half2(lo[0], lo[1]); ul.y = pack_half2(lo[2], lo[3]); ul.z = pack_half2(lo[4], lo[5]); ul.w = pack_half2(lo[6], lo[7]);
    planes[((size_t)n * 16 + c8) * L + l] = uh;
    planes[((size_t)n * 16 + 8 + c8) * L + l] = ul;
}

// EX: the end of the training step's gradient chain (train.cu) - planes written by the chain kernel's gradient mode (g * 2^E, E
// in the tensor's slot) back to fp32, plus a second addend (the long skip's gradient)
template <bool EX>
__global__ void __launch_bounds__(256)
planes_to_c4l4_kernel(const uint4 *__restrict__ planes, float4 *__restrict__ act, int N, int L, const unsigned *__restrict__ absmax_bits,
                      int add, const float4 *__restrict__ add2, const GradSlot *__restrict__ slot) {
    const int n = blockIdx.z, c8 = blockIdx.y;
    const int l = blockIdx.x * 256 + threadIdx.x;
    if (l >= L) return;
    float inv = 1.0f / plane_scale(absmax_bits);                     // exact: a power of two
    if (EX && slot != nullptr) {
        int d = -__ldg(&slot->E);
        d = d < -126 ? -126 : (d > 127 ? 127 : d);
        inv = __uint_as_float((unsigned)(d + 127) << 23);
    }
    float fh[8], fl[8];
    unpack8(planes[((size_t)n * 16 + c8) * L + l], fh);
    unpack8(planes[((size_t)n * 16 + 8 + c8) * L + l], fl);
    float v[8];
#pragma unroll
    for (int i = 0; i < 8; ++i) v[i] = (fh[i] + TC_LO_INV * fl[i]) * inv;
    float4 *o0 = act + ((size_t)n * 16 + 2 * c8) * L + l, *o1 = act + ((size_t)n * 16 + 2 * c8 + 1) * L + l;
    float4 r0 = make_float4(v[0], v[1], v[2], v[3]), r1 = make_float4(v[4], v[5], v[6], v[7]);
    if (add) {
        const float4 p0 = *o0, p1 = *o1;
        r0.x += p0.x; r0.y += p0.y; r0.z += p0.z; r0.w += p0.w;
        r1.x += p1.x; r1.y += p1.y; r1.z += p1.z; r1.w += p1.w;
    }
    if (EX && add2 != nullptr) {
        const float4 q0 = __ldg(add2 + ((size_t)n * 16 + 2 * c8) * L + l), q1 = __ldg(add2 + ((size_t)n * 16 + 2 * c8 + 1) * L + l);
        r0.x += q0.x; r0.y += q0.y; r0.z += q0.z; r0.w += q0.w;
        r1.x += q1.x; r1.y += q1.y; r1.z += q1.z; r1.w += q1.w;
    }
    *o0 = r0;
    *o1 = r1;
}

// fp32 C4L4 gradient -> scaled split planes (as c4l4_to_planes_kernel with absmax_bits), 4 positions per thread, and the bias
// gradient's partial sums on the way: bias_partial[(n * gridDim.x + blockIdx.x) * 64 + c] = sum of channel c over this block's
// (up to) 1024 positions, summed in a fixed order (shuffle tree, then the 8 warps in order)
constexpr int G2P_Q = 4;
__global__ void __launch_bounds__(256)
grad_to_planes_kernel(const float4 *__restrict__ act, uint4 *__restrict__ planes, int N, int L, const unsigned *__restrict__ absmax_bits,
                      float *__restrict__ bias_partial, GradSlot *__restrict__ slot_out) {
    const int n = blockIdx.z, c8 = blockIdx.y;
    const float sc = plane_scale(absmax_bits);
    if (slot_out != nullptr && blockIdx.x == 0 && c8 == 0 && n == 0 && threadIdx.x == 0) {
        // the planes' slot for the chain kernel's gradient mode: E = log2 of the scale, max |stored| = max |g| * scale
        slot_out->E = (int)((__float_as_uint(sc) >> 23) & 255u) - 127;
        slot_out->amax_bits = __float_as_uint(__uint_as_float(__ldg(absmax_bits)) * sc);
    }
    float s[8] = {0.f, 0.f, 0.f, 0.f, 0.f, 0.f, 0.f, 0.f};
#pragma unroll
    for (int q = 0; q < G2P_Q; ++q) {
        const int l = blockIdx.x * (256 * G2P_Q) + q * 256 + threadIdx.x;
        if (l < L) {
            const float4 a0 = act[((size_t)n * 16 + 2 * c8) * L + l], a1 = act[((size_t)n * 16 + 2 * c8 + 1) * L + l];
            const float r[8] = {a0.x, a0.y, a0.z, a0.w, a1.x, a1.y, a1.z, a1.w};
            float hi[8], lo[8];
#pragma unroll
            for (int i = 0; i < 8; ++i) { s[i] += r[i]; split_hi_lo(r[i] * sc, hi[i], lo[i]); }
            uint4 uh, ul;
            uh.x = pack_half2(hi[0], hi[1]); uh.y = pack_half2(hi[2], hi[3]); uh.z = pack_half2(hi[4], hi[5]); uh.w = pack_half2(hi[6], hi[7]);
            ul.x = pack_half2(lo[0], lo[1]); ul.y = pack_half2(lo[2], lo[3]); ul.z = pack_half2(lo[4], lo[5]); ul.w = pack_half2(lo[6], lo[7]);
            planes[((size_t)n * 16 + c8) * L + l] = uh;
            planes[((size_t)n * 16 + 8 + c8) * L + l] = ul;
        }
    }
#pragma unroll
    for (int i = 0; i < 8; ++i)
#pragma unroll
        for (int o = 16; o > 0; o >>= 1) s[i] += __shfl_xor_sync(0xffffffffu, s[i], o);
    __shared__ float ws[8][8];
    if ((threadIdx.x & 31) == 0) {
#pragma unroll
        for (int i = 0; i < 8; ++i) ws[threadIdx.x >> 5][i] = s[i];
    }
    __syncthreads();
    if (threadIdx.x < 8) {
        float t = 0.0f;
#pragma unroll
        for (int w = 0; w < 8; ++w) t += ws[w][threadIdx.x];
        bias_partial[((size_t)n * gridDim.x + blockIdx.x) * 64 + 8 * c8 + threadIdx.x] = t;
    }
}

__global__ void __launch_bounds__(256) absmax_kernel(const float4 *__restrict__ v, size_t n4, unsigned *__restrict__ out_bits) {
    float m = 0.0f;
    for (size_t i = (size_t)blockIdx.x * 256 + threadIdx.x; i < n4; i += (size_t)gridDim.x * 256) {
        const float4 a = __ldg(v + i);
        m = fmaxf(m, fmaxf(fmaxf(fabsf(a.x), fabsf(a.y)), fmaxf(fabsf(a.z), fabsf(a.w))));
    }
#pragma unroll
    for (int o = 16; o > 0; o >>= 1) m = fmaxf(m, __shfl_xor_sync(0xffffffffu, m, o));
    __shared__ float wm[8];
    if ((threadIdx.x & 31) == 0) wm[threadIdx.x >> 5] = m;
    __syncthreads();
    if (threadIdx.x == 0) {
#pragma unroll
        for (int i = 1; i < 8; ++i) m = fmaxf(m, wm[i]);
        atomicMax(out_bits, __float_as_uint(m));
    }
}

// ------------------------------------------------------------------------------------------------ launchers
static unsigned long long *g_timeline = nullptr;
static int g_timeline_cap = 0;
void set_chain_timeline(void *dev_buf, int cap) { g_timeline = reinterpret_cast<unsigned long long *>(dev_buf); g_timeline_cap = cap; }

int chain_tc_grid(int count, const TcLayer *layers, int N, int L) {
    int Sp = 0;
    for (int k = 1; k < count; ++k) Sp += layers[k].pad;
    const int t_out = 256 - 2 * Sp;
    const long long total = (long long)N * ((L + t_out - 1) / t_out);
    return (int)(total < kSMs ? total : kSMs);
}

static int launch_chain_impl(const TcLayer *layers, int count, const void *in, const void *x0, void *out_planes, float *out_mask,
                             int N, int L, cudaStream_t st, int fast, const TcGrad *grad) {
    STOF_REQUIRE(count >= 1 && count <= TC_MAX_LAYERS, STOF_E_INVALID, "chain_tc: %d layers per segment (max %d)", count, TC_MAX_LAYERS);
    static SmemAttrOnce attr_exact, attr_fast;
    static SmemAttrOnce attr_grad;
    STOF_REQUIRE(grad == nullptr || (!fast && count == 1 && layers[0].type == TC_NONE && grad->mode >= 1 && grad->mode <= 3 && grad->g_in &&
                                     grad->g_out && (grad->mode == 1 || grad->aux) && (grad->mode != 2 || grad->g_aux)),
                 STOF_E_INVALID, "chain_tc: gradient mode needs one TC_NONE layer, exact arithmetic and its slots");
    { const int rc = grad ? ensure_dyn_smem(chain_tc_kernel<false, true>, TC_SMEM_BYTES, attr_grad)
                   : fast ? ensure_dyn_smem(chain_tc_kernel<true, false>, TC_SMEM_BYTES, attr_fast)
                          : ensure_dyn_smem(chain_tc_kernel<false, false>, TC_SMEM_BYTES, attr_exact); if (rc) return rc; }
    TcParams p;
    memset(&p, 0, sizeof(p));
    int S = 0, Sp = 0;
    for (int k = 0; k < count; ++k) {
        p.layer[k] = layers[k];
        S += layers[k].pad;
        if (k > 0) Sp += layers[k].pad;
        STOF_REQUIRE(layers[k].type != TC_LAST || k == count - 1, STOF_E_INVALID, "chain_tc: conv_last must end a segment");
        STOF_REQUIRE(layers[k].type != TC_RES || k >= 1, STOF_E_INVALID, "chain_tc: a short-skip layer cannot start a segment");
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
    // Buffer P (input tile; operand of even layers, output / short-skip source of odd layers) is last touched by layer
    // count-2 when the count is even, provided the last layer's short skip is parked in TMEM (stash); else by the last layer.
    const bool even = (count % 2) == 0;
    p.stash = (even && layers[count - 1].type == TC_RES) ? 1 : 0;
    p.free_layer = even ? count - 2 : count - 1;
    // the producer requests the next tile's input after this tap of the last layer (the ring holds that many taps of the
    // last layer while the epilogue of free_layer is still running, so waiting for P there delays no weights)
    p.in_layer = count - 1;
    // exact mode: the producer asks for the next tile's input after tap STAGES-2 of the last layer (tuned on the timeline);
    // one-pass mode (deep ring): after the tile's last tap
    p.in_tap = (!fast && even && layers[count - 1].taps > TC_STAGES - 1) ? TC_STAGES - 2 : layers[count - 1].taps - 1;
    // one layer per launch (the training step keeps every activation): buffer Q is idle, so the tiles' inputs alternate between P
    // and Q; the next tile's load is requested after the current tile's last tap (in_tap above: the producer is then up to a ring
    // of taps ahead of the MMAs, so the load runs behind the rest of the tile's MMAs and its epilogue)
    p.dbl = (count == 1) ? 1 : 0;
    for (int k = 0; k < count; ++k)
        STOF_REQUIRE(layers[k].taps >= 1, STOF_E_INVALID, "chain_tc: layer without taps");
    const bool ends_last = layers[count - 1].type == TC_LAST;
    STOF_REQUIRE(ends_last ? out_mask != nullptr : out_planes != nullptr, STOF_E_INVALID, "chain_tc: missing output buffer");
    const int grid = (int)(total < kSMs ? total : kSMs);
    if (grad) { p.grad = *grad; chain_tc_kernel<false, true><<<grid, TC_THREADS, TC_SMEM_BYTES, st>>>(p); }
    else if (fast) chain_tc_kernel<true, false><<<grid, TC_THREADS, TC_SMEM_BYTES, st>>>(p);
    else chain_tc_kernel<false, false><<<grid, TC_THREADS, TC_SMEM_BYTES, st>>>(p);
    STOF_LAUNCH_CHECK();
    return STOF_OK;
}

int launch_chain_tc(const TcLayer *layers, int count, const void *in, const void *x0, void *out_planes, float *out_mask,
                    int N, int L, cudaStream_t st, int fast) {
    return launch_chain_impl(layers, count, in, x0, out_planes, out_mask, N, L, st, fast, nullptr);
}
int launch_chain_tc_grad(const TcLayer *layer, const void *in, void *out_planes, int N, int L, cudaStream_t st, const TcGrad &grad) {
    return launch_chain_impl(layer, 1, in, nullptr, out_planes, nullptr, N, L, st, 0, &grad);
}

int launch_x0_split(const float *x, const float *e, int N, int L, int P, int scale, int off, const float *w, const float *b,
                    void *planes, cudaStream_t st, int fast, int K) {
    dim3 grid((L + 256 * X0_Q - 1) / (256 * X0_Q), N);
    STOF_REQUIRE(e == nullptr || scale >= 72, STOF_E_UNSUPPORTED, "x0_split: upsampling factor %d (the shared-memory slot table is sized for >= 72)", scale);
    STOF_REQUIRE(K == 9 || K == 3, STOF_E_UNSUPPORTED, "x0_split: first conv kernel size %d (built for 9 and 3)", K);
    if (K == 9) x0_split_kernel<9><<<grid, 256, 0, st>>>(x, e, N, L, P, scale, off, w, b, reinterpret_cast<uint4 *>(planes), fast);
    else x0_split_kernel<3><<<grid, 256, 0, st>>>(x, e, N, L, P, scale, off, w, b, reinterpret_cast<uint4 *>(planes), fast);
    STOF_LAUNCH_CHECK();
    return STOF_OK;
}

int launch_c4l4_to_planes(const float *act, void *planes, int N, int L, cudaStream_t st, const unsigned *absmax_bits) {
    STOF_REQUIRE(N <= 65535, STOF_E_INVALID, "c4l4_to_planes: at most 65535 rows per call");
    dim3 grid((L + 255) / 256, 8, N);
    c4l4_to_planes_kernel<<<grid, 256, 0, st>>>(reinterpret_cast<const float4 *>(act), reinterpret_cast<uint4 *>(planes), N, L, absmax_bits);
    STOF_LAUNCH_CHECK();
    return STOF_OK;
}

int launch_planes_to_c4l4(const void *planes, float *act, int N, int L, cudaStream_t st, const unsigned *absmax_bits, int add) {
    STOF_REQUIRE(N <= 65535, STOF_E_INVALID, "planes_to_c4l4: at most 65535 rows per call");
    dim3 grid((L + 255) / 256, 8, N);
    planes_to_c4l4_kernel<false><<<grid, 256, 0, st>>>(reinterpret_cast<const uint4 *>(planes), reinterpret_cast<float4 *>(act), N, L, absmax_bits, add,
                                                       nullptr, nullptr);
    STOF_LAUNCH_CHECK();
    return STOF_OK;
}

int launch_grad_planes_to_c4l4(const void *planes, float *act, int N, int L, cudaStream_t st, const GradSlot *slot, const float *add2) {
    STOF_REQUIRE(N <= 65535, STOF_E_INVALID, "planes_to_c4l4: at most 65535 rows per call");
    dim3 grid((L + 255) / 256, 8, N);
    planes_to_c4l4_kernel<true><<<grid, 256, 0, st>>>(reinterpret_cast<const uint4 *>(planes), reinterpret_cast<float4 *>(act), N, L, nullptr, 0,
                                                      reinterpret_cast<const float4 *>(add2), slot);
    STOF_LAUNCH_CHECK();
    return STOF_OK;
}

int grad_to_planes_parts(int N, int L) { return N * ((L + 256 * G2P_Q - 1) / (256 * G2P_Q)); }

int launch_grad_to_planes(const float *g, void *planes, int N, int L, cudaStream_t st, const unsigned *absmax_bits, float *bias_partial,
                          GradSlot *slot_out) {
    STOF_REQUIRE(N <= 65535, STOF_E_INVALID, "grad_to_planes: at most 65535 rows per call");
    dim3 grid((L + 256 * G2P_Q - 1) / (256 * G2P_Q), 8, N);
    grad_to_planes_kernel<<<grid, 256, 0, st>>>(reinterpret_cast<const float4 *>(g), reinterpret_cast<uint4 *>(planes), N, L, absmax_bits, bias_partial,
                                                slot_out);
    STOF_LAUNCH_CHECK();
    return STOF_OK;
}

int launch_absmax(const float *v, size_t n4, unsigned *out_bits, cudaStream_t st) {
    const size_t g = (n4 + 255) / 256;
    absmax_kernel<<<(unsigned)(g < (size_t)kSMs * 8 ? g : (size_t)kSMs * 8), 256, 0, st>>>(reinterpret_cast<const float4 *>(v), n4, out_bits);
    STOF_LAUNCH_CHECK();
    return STOF_OK;
}

}  // namespace stof
