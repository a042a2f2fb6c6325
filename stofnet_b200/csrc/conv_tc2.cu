// The chain kernel on CTA PAIRS: the same segment of 64 -> 64 conv layers (+ conv_last / SampleShuffle1D) as conv_tc.cu
// (models/stofnet.py:51-65), with every tcgen05.mma issued as cta_group::2 over a 2-CTA cluster.
//
//   * each CTA of the pair owns its own 256-row tile (two 128-row blocks), its own activation buffers, TMEM accumulators and
//     epilogue - exactly as in the one-CTA kernel; block b of BOTH tiles forms one M = 256 MMA stream issued by one thread of
//     the leader CTA (cluster rank 0), so an instruction carries twice the rows;
//   * the B operand (the weight tap) is split between the two CTAs: an N = 128 MMA reads 64 rows from each CTA's shared
//     memory, the N = 64 X_lo product 32 rows.  Per SM and K-step the operand traffic drops from 14 KB to 11 KB (the N = 64 MMA
//     was shared-memory bound: 48 cycles for 32 cycles of tensor work; now 40), and each CTA streams 12 KB per tap from L2
//     instead of 16 KB:   rank 0 holds [W_hi (64 rows) | W_hi rows 0-31],   rank 1 holds [W_lo (64 rows) | W_hi rows 32-63];
//     the stacked MMA reads rows 0-63 of both (D0 = X_hi W_hi in columns 0-63, D1 += X_hi W_lo in 64-127), the X_lo product rows
//     64-95 of both (D1 columns 64-95 from rank 0's rows, 96-127 from rank 1's) - the accumulator layout, the accumulation order
//     and therefore the results are those of the one-CTA kernel, bit for bit.
//   * synchronisation: weight / input arrival of the peer is relayed to the leader by the peer's idle issuer warps (a bulk copy
//     may only signal a barrier of its own CTA); tcgen05.commit multicasts "stage free" / "accumulator full" to both CTAs;
//     the epilogue warps of both CTAs arrive on the LEADER's "accumulator empty" / "activations ready" barriers (remote
//     mbarrier arrive with release at cluster scope after fence.proxy.async, acquired at cluster scope by the issuer).
#include <string.h>

#include "conv_tc_shared.cuh"

namespace stof {

using namespace ptx;

constexpr int T2_STAGES = 6;                          // exact mode: 6 x 12 KB half-taps
constexpr int T2_STAGE_BYTES = 12288;                  // [8 c8][96 rows][8] fp16
constexpr int T2_FAST_STAGES = 12;                    // one-pass mode: 12 x 4 KB ([8 c8][32 rows][8])
constexpr int T2_FAST_STAGE_BYTES = 4096;
constexpr int T2_MAX_STAGES = 12;

constexpr int T2_OFF_P = 0;
constexpr int T2_OFF_Q = TC_BUF;
constexpr int T2_OFF_W = 2 * TC_BUF;
constexpr int T2_OFF_BIAS = T2_OFF_W + T2_STAGES * T2_STAGE_BYTES;           // [TC_MAX_LAYERS][64] fp32
constexpr int T2_OFF_BAR = T2_OFF_BIAS + TC_MAX_LAYERS * 64 * 4;
// barriers (8 B each): w_full[S] w_peer[S] w_empty[S] acc_full[2] acc_empty[2] act[TC_MAX_LAYERS][2] in_full in_peer p_free
constexpr int T2_NBAR = 3 * T2_MAX_STAGES + 4 + 2 * TC_MAX_LAYERS + 3;
constexpr int T2_OFF_TMEM = T2_OFF_BAR + T2_NBAR * 8;
constexpr int T2_SMEM_BYTES = T2_OFF_TMEM + 16;
static_assert(T2_SMEM_BYTES <= 232448, "pair chain kernel exceeds the 227 KB shared-memory limit");
static_assert(T2_FAST_STAGES * T2_FAST_STAGE_BYTES <= T2_STAGES * T2_STAGE_BYTES, "one-pass ring must fit the exact ring");

// All taps of one layer for block b of BOTH CTAs (M = 256), issued by a single thread of the leader.
//   64 -> 64 conv : half-taps [c8][96 rows][8]; stacked MMA N = 128 on rows 0-63 of both CTAs, X_lo product N = 64 on rows 64-95
//   conv_last     : half-taps [c8][16 rows][8]; N = 16 on rows 0-7 (rank 0: W_hi, rank 1: W_lo -> columns 0-3 / 8-11), X_lo product
//                   N = 16 on rows 8-15 (rank 0: W_hi, rank 1: zeros) into columns 8-23
//   one-pass mode : [c8][32 rows][8], one N = 64 MMA; conv_last [c8][8 rows][8], N = 16
template <bool FAST, bool LAST>
__device__ __forceinline__ void issue_layer_taps_pair(uint32_t a_base, uint32_t w_base, uint32_t d_tmem, int taps, int s, uint32_t ph,
                                                      uint32_t bar_w_full0, uint32_t bar_w_peer0, uint32_t bar_w_empty0,
                                                      uint32_t bar_acc_full) {
    constexpr uint32_t kDescHi = (128u >> 4) | (1u << 14);
    constexpr uint32_t W_ROWS = LAST ? (FAST ? 8u : 16u) : (FAST ? 32u : 96u);      // rows per c8 chunk of a half-tap
    constexpr uint32_t LO_ROW = LAST ? 8u : 64u;                                    // first row of the X_lo product's B rows
    constexpr uint32_t LO_COL = LAST ? 8u : 64u;                                    // its first accumulator column
    constexpr uint32_t STRIDE16 = (uint32_t)(FAST ? T2_FAST_STAGE_BYTES : T2_STAGE_BYTES) >> 4;
    constexpr int NST = FAST ? T2_FAST_STAGES : T2_STAGES;
    constexpr uint32_t idesc_a = umma_idesc_f16(256, LAST ? 16 : (FAST ? 64 : 128));
    constexpr uint32_t idesc_b = umma_idesc_f16(256, LAST ? 16 : 64);
    for (int t = 0; t < taps; ++t) {
        mbar_wait(bar_w_full0 + 8u * (uint32_t)s, ph);                  // this CTA's half of the tap
        mbar_wait_cluster(bar_w_peer0 + 8u * (uint32_t)s, ph);         // the peer's half (relayed)
        tc_fence_after_sync();
        const uint32_t a_lo = a_base + (uint32_t)t;
        const uint32_t w_lo = (w_base + (uint32_t)s * STRIDE16) | (W_ROWS << 16);
#pragma unroll
        for (int ks = 0; ks < 4; ++ks) {
            umma_f16_pair(d_tmem, make_desc(a_lo + 2 * ks * TC_R, kDescHi), make_desc(w_lo + 2 * ks * W_ROWS, kDescHi), idesc_a,
                          (t > 0 || ks > 0) ? 1u : 0u);
            if (!FAST)
                umma_f16_pair(d_tmem + LO_COL, make_desc(a_lo + (8 + 2 * ks) * TC_R, kDescHi),
                              make_desc(w_lo + 2 * ks * W_ROWS + LO_ROW, kDescHi), idesc_b, 1u);
        }
        umma_commit_pair(bar_w_empty0 + 8u * (uint32_t)s, 3);            // both CTAs' producers: this stream is done with the stage
        if (++s == NST) { s = 0; ph ^= 1u; }
    }
    umma_commit_pair(bar_acc_full, 3);                                   // both CTAs' epilogues
}

template <bool FAST>
__global__ void __cluster_dims__(2, 1, 1) __launch_bounds__(TC_THREADS, 1) chain_tc2_kernel(const TcParams p) {
    extern __shared__ __align__(1024) uint8_t smem[];
    const uint32_t sbase = smem_u32(smem);
    const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
    const uint32_t rank = cluster_ctarank();
    const bool leader = (rank == 0);
    const uint32_t bar0 = sbase + T2_OFF_BAR;
    auto BAR_W_FULL = [&](int s) { return bar0 + 8 * s; };
    auto BAR_W_PEER = [&](int s) { return bar0 + 8 * (T2_MAX_STAGES + s); };
    auto BAR_W_EMPTY = [&](int s) { return bar0 + 8 * (2 * T2_MAX_STAGES + s); };
    auto BAR_ACC_FULL = [&](int b) { return bar0 + 8 * (3 * T2_MAX_STAGES + b); };
    auto BAR_ACC_EMPTY = [&](int b) { return bar0 + 8 * (3 * T2_MAX_STAGES + 2 + b); };
    auto BAR_ACT = [&](int k, int b) { return bar0 + 8 * (3 * T2_MAX_STAGES + 4 + 2 * k + b); };
    const uint32_t BAR_IN_FULL = bar0 + 8 * (3 * T2_MAX_STAGES + 4 + 2 * TC_MAX_LAYERS);
    const uint32_t BAR_IN_PEER = BAR_IN_FULL + 8;
    const uint32_t BAR_P_FREE = BAR_IN_FULL + 16;
    constexpr int n_stages = FAST ? T2_FAST_STAGES : T2_STAGES;
    constexpr uint32_t stage_stride = FAST ? T2_FAST_STAGE_BYTES : T2_STAGE_BYTES;
    float *s_bias = reinterpret_cast<float *>(smem + T2_OFF_BIAS);
    volatile uint32_t *s_tmem = reinterpret_cast<volatile uint32_t *>(smem + T2_OFF_TMEM);
    // arrive on a barrier that lives in the LEADER's shared memory (release at cluster scope; the issuer acquires at cluster scope)
    auto arrive_leader = [&](uint32_t bar) { mbar_arrive_cluster(mapa_shared(bar, 0)); };

    // ---- one-time setup ------------------------------------------------------------------------------
    for (int i = threadIdx.x; i < p.n_layers * 64; i += TC_THREADS) {
        const int k = i >> 6, c = i & 63;
        s_bias[i] = (c < p.layer[k].n_out) ? __ldg(p.layer[k].bias + c) : 0.0f;
    }
    if (threadIdx.x == 0) {
        for (int s = 0; s < T2_MAX_STAGES; ++s) {
            mbar_init(BAR_W_FULL(s), 1);
            mbar_init(BAR_W_PEER(s), 1);
            mbar_init(BAR_W_EMPTY(s), 2);                 // one commit per issuing stream (block 0 / block 1)
        }
        for (int b = 0; b < 2; ++b) {
            mbar_init(BAR_ACC_FULL(b), 1);
            mbar_init(BAR_ACC_EMPTY(b), 2 * TC_EPI_WARPS);          // the epilogue warps of both CTAs
        }
        for (int k = 0; k < TC_MAX_LAYERS; ++k) {
            mbar_init(BAR_ACT(k, 0), 2 * TC_EPI_WARPS);
            mbar_init(BAR_ACT(k, 1), 2 * TC_EPI_WARPS);
        }
        mbar_init(BAR_IN_FULL, 1);
        mbar_init(BAR_IN_PEER, 1);
        mbar_init(BAR_P_FREE, TC_EPI_WARPS);
        fence_mbar_init();
    }
    if (warp == 1) tmem_alloc_pair(sbase + T2_OFF_TMEM, 512);
    tc_fence_before_sync();
    cluster_sync_all();                                   // barriers of BOTH CTAs initialised before any remote arrive
    tc_fence_after_sync();
    const uint32_t tmem0 = *s_tmem;

    const int n_layers = p.n_layers;
    const int r_in = p.r_in;                 // input rows staged per tile
    const int n_pairs = (p.total_tiles + 1) >> 1;
    const int cl = (int)(blockIdx.x >> 1), ncl = (int)(gridDim.x >> 1);

    if (warp == 0) {
        // =================================== producer (both CTAs) =======================================
        // own input tile into buffer P, own half of every weight tap into the ring (see conv_tc.cu for the schedule)
        auto issue_input = [&](int pair, int j) {
            const int tile = 2 * pair + (int)rank;
            const bool valid = tile < p.total_tiles;
            if (j > 0) mbar_wait(BAR_P_FREE, (uint32_t)(j - 1) & 1u);
            const int nslab = FAST ? 8 : 16;               // one-pass fp16 mode: hi plane only
            if (!valid) {
                // phantom tile: all-zero input, nothing loaded
                for (int slab = 0; slab < nslab; ++slab) {
                    uint8_t *sl = smem + T2_OFF_P + slab * TC_SLAB;
                    for (int r = lane; r < r_in; r += 32) *reinterpret_cast<uint4 *>(sl + r * 16) = make_uint4(0, 0, 0, 0);
                }
                fence_proxy_async_smem();
                __syncwarp();
                if (lane == 0) mbar_arrive(BAR_IN_FULL);
                __syncwarp();
                return;
            }
            const int n = tile / p.tiles_per_row;
            const int l_org = (tile - n * p.tiles_per_row) * p.t_out - p.l_shift;
            const int r_lo = max(0, -l_org), r_hi = min(r_in, p.L - l_org);
            // rows outside the waveform are the conv's zero padding (only tiles at the ends of a row have any)
            if (r_lo > 0 || r_hi < r_in) {
                for (int slab = 0; slab < nslab; ++slab) {
                    uint8_t *sl = smem + T2_OFF_P + slab * TC_SLAB;
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
                    bulk_g2s(sbase + T2_OFF_P + slab * TC_SLAB + r_lo * 16, src + ((size_t)slab * p.L + l_org + r_lo) * 8, bytes,
                             BAR_IN_FULL);
            }
            __syncwarp();
        };
        int s = 0;
        uint32_t ph = 0;
        int it = 0;
        if (cl < n_pairs) issue_input(cl, 0);
        for (int pair = cl; pair < n_pairs; pair += ncl, ++it) {
            const int next = pair + ncl;
            const int tile = 2 * pair + (int)rank;
            for (int k = 0; k < n_layers; ++k) {
                const TcLayer &ly = p.layer[k];
                if (ly.type == TC_LONGSKIP && k > 0 && lane == 0 && tile < p.total_tiles) {
                    // pull this tile's rows of the long-skip tensor into L2 one layer before the epilogue reads them
                    const int n = tile / p.tiles_per_row;
                    const int l0 = (tile - n * p.tiles_per_row) * p.t_out;
                    const int a0 = max(0, l0 - 16), a1 = min(p.L, l0 + p.t_out + 16);
                    const __half *src = reinterpret_cast<const __half *>(p.x0) + (size_t)n * 16 * p.L * 8;
                    for (int slab = 0; slab < (FAST ? 8 : 16); ++slab)
                        bulk_prefetch_l2(src + ((size_t)slab * p.L + a0) * 8, (uint32_t)(a1 - a0) * 16u);
                }
                const uint32_t sbytes = (uint32_t)(FAST ? ly.stage_bytes2_fast : ly.stage_bytes2);
                const uint8_t *wsrc = reinterpret_cast<const uint8_t *>(FAST ? ly.w2_fast[rank] : ly.w2[rank]);
                for (int t = 0; t < ly.taps; ++t) {
                    if (lane == 0) {
                        mbar_wait(BAR_W_EMPTY(s), ph ^ 1u);
                        mbar_arrive_expect_tx(BAR_W_FULL(s), sbytes);
                        bulk_g2s(sbase + T2_OFF_W + s * stage_stride, wsrc + (size_t)t * sbytes, sbytes, BAR_W_FULL(s));
                    }
                    if (++s == n_stages) { s = 0; ph ^= 1u; }
                    if (k == p.in_layer && t == p.in_tap && next < n_pairs) {
                        __syncwarp();
                        issue_input(next, it + 1);
                    }
                }
            }
        }
    } else if (warp <= 2) {
        if (leader) {
            // =================================== MMA issuers (leader CTA) ====================================
            // warp 1 -> block 0 of both tiles, warp 2 -> block 1 of both tiles; one elected lane issues, in a fixed order
            constexpr uint32_t kALoBase = (uint32_t)(TC_SLAB >> 4) << 16;            // A: LBO = one slab
            const int b = warp - 1;
            const uint32_t d_tmem = tmem0 + (uint32_t)b * 128u;
            int s = 0;
            uint32_t ph = 0, lseq = 0;
            int it = 0;
            for (int pair = cl; pair < n_pairs; pair += ncl, ++it) {
                mbar_wait(BAR_IN_FULL, (uint32_t)(it & 1));
                mbar_wait_cluster(BAR_IN_PEER, (uint32_t)(it & 1));
                tc_fence_after_sync();
                for (int k = 0; k < n_layers; ++k, ++lseq) {
                    const TcLayer &ly = p.layer[k];
                    const int taps = ly.taps;
                    const uint32_t a_base = (((sbase + ((k & 1) ? T2_OFF_Q : T2_OFF_P)) >> 4) | kALoBase) +
                                            (uint32_t)(p.b_row[k] - ly.pad + 128 * b);
                    mbar_wait_cluster(BAR_ACC_EMPTY(b), (lseq & 1u) ^ 1u);
                    if (k > 0) {
                        mbar_wait_cluster(BAR_ACT(k - 1, 0), (uint32_t)(it & 1));
                        if (b == 1) mbar_wait_cluster(BAR_ACT(k - 1, 1), (uint32_t)(it & 1));
                    }
                    tc_fence_after_sync();
                    if (elect_one()) {
                        const uint32_t w_base = (sbase + T2_OFF_W) >> 4;
                        if (ly.type == TC_LAST)
                            issue_layer_taps_pair<FAST, true>(a_base, w_base, d_tmem, taps, s, ph, BAR_W_FULL(0), BAR_W_PEER(0),
                                                              BAR_W_EMPTY(0), BAR_ACC_FULL(b));
                        else
                            issue_layer_taps_pair<FAST, false>(a_base, w_base, d_tmem, taps, s, ph, BAR_W_FULL(0), BAR_W_PEER(0),
                                                               BAR_W_EMPTY(0), BAR_ACC_FULL(b));
                    }
                    __syncwarp();
                    s += taps;
                    while (s >= n_stages) { s -= n_stages; ph ^= 1u; }
                }
            }
        } else if (lane == 0) {
            // =================================== relays (peer CTA) ===========================================
            // a bulk copy can only signal a barrier of its own CTA: the peer's otherwise idle issuer warps forward "my half of
            // the tap has landed" (warp 1) and "my input tile has landed" (warp 2) to the leader's barriers
            if (warp == 1) {
                int s = 0;
                uint32_t ph = 0;
                for (int pair = cl; pair < n_pairs; pair += ncl)
                    for (int k = 0; k < n_layers; ++k)
                        for (int t = 0; t < p.layer[k].taps; ++t) {
                            mbar_wait(BAR_W_FULL(s), ph);
                            mbar_arrive_cluster(mapa_shared(BAR_W_PEER(s), 0));
                            if (++s == n_stages) { s = 0; ph ^= 1u; }
                        }
            } else {
                int it = 0;
                for (int pair = cl; pair < n_pairs; pair += ncl, ++it) {
                    mbar_wait(BAR_IN_FULL, (uint32_t)(it & 1));
                    mbar_arrive_cluster(mapa_shared(BAR_IN_PEER, 0));
                }
            }
        }
    } else {
        // =================================== epilogue (both CTAs) ========================================
        // 8 warps: warp quadrant (warp & 3) selects the 32 TMEM lanes, group (0/1) the 32-channel half of the block.
        // The blocks are processed one after the other, which staggers their layer pipelines by one epilogue: while block
        // 0's epilogue runs, block 1's MMAs keep the tensor pipe and the shared-memory ports busy, and vice versa.
        const int quad = warp & 3;
        const int egrp = (warp - 3) >> 2;
        const int trow = 32 * quad + lane;                       // TMEM lane = row within the 128-block
        const uint32_t lane_addr = (uint32_t)(32 * quad) << 16;
        const bool free_early = (p.stash == 0);     // free_layer has an even index: it reads P as operand only
        uint32_t lseq = 0;
        int it = 0;
        for (int pair = cl; pair < n_pairs; pair += ncl, ++it) {
            const int tile_raw = 2 * pair + (int)rank;
            const bool valid = tile_raw < p.total_tiles;             // an odd tile count leaves the last pair's second CTA a phantom tile
            const int tile = valid ? tile_raw : p.total_tiles - 1;
            const int n = tile / p.tiles_per_row;
            const int l0 = (tile - n * p.tiles_per_row) * p.t_out;
            const int l_org = l0 - p.l_shift;
            const int l_end = min(l0 + p.t_out, p.L);
            for (int k = 0; k < n_layers; ++k, ++lseq) {
                const TcLayer &ly = p.layer[k];
                const int type = ly.type;
                const bool last_in_seg = (k == n_layers - 1);
                uint8_t *outbuf = smem + ((k & 1) ? T2_OFF_P : T2_OFF_Q);
                const float *bias = s_bias + 64 * k;
                for (int b = 0; b < 2; ++b) {
                    const int r = p.b_row[k] + 128 * b + trow;
                    const int l = l_org + r;
                    const bool inside = valid && (l >= 0) && (l < p.L);
                    // long skip (res1 = x0 planes in global memory): issue the loads before waiting for the accumulator so
                    // that their latency hides behind the block's MMAs (the producer has pulled the rows into L2)
                    uint4 xs_h[4], xs_l[4];
                    if (type == TC_LONGSKIP) {
                        const uint4 *x0 = reinterpret_cast<const uint4 *>(p.x0) + (size_t)n * 16 * p.L;
#pragma unroll
                        for (int h = 0; h < 4; ++h) {
                            xs_h[h] = inside ? __ldg(x0 + (size_t)(4 * egrp + h) * p.L + l) : make_uint4(0, 0, 0, 0);
                            xs_l[h] = (inside && !FAST) ? __ldg(x0 + (size_t)(8 + 4 * egrp + h) * p.L + l) : make_uint4(0, 0, 0, 0);
                        }
                    }
                    mbar_wait(BAR_ACC_FULL(b), lseq & 1u);
                    tc_fence_after_sync();
                    if (free_early && b == 1 && k == p.free_layer) {
                        // both blocks' MMAs of the last layer that reads buffer P have completed, and this layer's epilogue does
                        // not touch P (its output goes to Q / global, no TMEM stash): hand P to the producer before the epilogue
                        // math, so the next tile's input load overlaps it
                        if (lane == 0) mbar_arrive(BAR_P_FREE);
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
                        if (lane == 0) arrive_leader(BAR_ACC_EMPTY(b));
                        if (egrp == 0 && valid && l >= l0 && l < l_end) {
                            float4 o;
                            const float li = FAST ? 0.0f : TC_LO_INV;      // one-pass mode: columns 8..11 hold X_hi W_lo only
                            o.x = (__uint_as_float(d[0]) + li * __uint_as_float(d[8])) + bias[0];
                            o.y = (__uint_as_float(d[1]) + li * __uint_as_float(d[9])) + bias[1];
                            o.z = (__uint_as_float(d[2]) + li * __uint_as_float(d[10])) + bias[2];
                            o.w = (__uint_as_float(d[3]) + li * __uint_as_float(d[11])) + bias[3];
                            reinterpret_cast<float4 *>(p.out_mask)[(size_t)n * p.L + l] = o;
                        }
                    } else {
                        const bool store_global = last_in_seg && valid && (l >= l0) && (l < l_end);
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
                        if (lane == 0) arrive_leader(BAR_ACC_EMPTY(b));              // accumulator read: hand it back
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
                            if (lane == 0) arrive_leader(BAR_ACT(k, b));
                        }
                        if (p.stash != 0 && k == n_layers - 2) {
                            // park the last layer's short skip (rows of buffer P, the output of layer k - 1) as fp32 in spare
                            // TMEM columns, so that P can take the next tile's input while the last layer still runs
                            const int rr = p.b_row[n_layers - 1] + 128 * b + trow;
                            const uint8_t *pb = smem + T2_OFF_P + rr * 16;
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
                }
                if (k == p.free_layer && !free_early) {
                    // every read of buffer P for this tile is done (this warp's share)
                    __syncwarp();
                    if (lane == 0) mbar_arrive(BAR_P_FREE);
                }
            }
        }
    }

    // ---- teardown --------------------------------------------------------------------------------------
    // the leader's MMAs read the peer's shared memory and the peer arrives on the leader's barriers: nobody leaves early
    tc_fence_before_sync();
    cluster_sync_all();
    if (warp == 1) {
        tc_fence_after_sync();
        tmem_dealloc_pair(tmem0, 512);
    }
}

// ------------------------------------------------------------------------------------------------ launcher
int launch_chain_tc2(const TcLayer *layers, int count, const void *in, const void *x0, void *out_planes, float *out_mask,
                     int N, int L, cudaStream_t st, int fast) {
    STOF_REQUIRE(count >= 1 && count <= TC_MAX_LAYERS, STOF_E_INVALID, "chain_tc2: %d layers per segment (max %d)", count, TC_MAX_LAYERS);
    static SmemAttrOnce attr_exact, attr_fast;
    { const int rc = fast ? ensure_dyn_smem(chain_tc2_kernel<true>, T2_SMEM_BYTES, attr_fast)
                          : ensure_dyn_smem(chain_tc2_kernel<false>, T2_SMEM_BYTES, attr_exact); if (rc) return rc; }
    TcParams p;
    memset(&p, 0, sizeof(p));
    int S = 0, Sp = 0;
    for (int k = 0; k < count; ++k) {
        p.layer[k] = layers[k];
        S += layers[k].pad;
        if (k > 0) Sp += layers[k].pad;
        STOF_REQUIRE(layers[k].type != TC_LAST || k == count - 1, STOF_E_INVALID, "chain_tc2: conv_last must end a segment");
        STOF_REQUIRE(layers[k].type != TC_RES || k >= 1, STOF_E_INVALID, "chain_tc2: a short-skip layer cannot start a segment");
        STOF_REQUIRE(layers[k].type != TC_LONGSKIP || x0 != nullptr, STOF_E_INVALID, "chain_tc2: long skip needs x0");
        STOF_REQUIRE(layers[k].taps >= 1 && layers[k].w2[0] && layers[k].w2[1], STOF_E_INVALID, "chain_tc2: layer without pair packs");
    }
    const int OFF = S + 3;
    p.n_layers = count;
    int acc = 0;
    for (int k = 0; k < count; ++k) {
        acc += layers[k].pad;
        p.b_row[k] = OFF - acc;
    }
    p.r_in = OFF + 256;
    STOF_REQUIRE(p.r_in <= TC_R, STOF_E_INVALID, "chain_tc2: segment too deep for the smem tile (%d rows)", p.r_in);
    p.l_shift = p.b_row[0] + Sp;
    p.t_out = 256 - 2 * Sp;
    p.tiles_per_row = (L + p.t_out - 1) / p.t_out;
    const long long total = (long long)N * p.tiles_per_row;
    STOF_REQUIRE(total < (1ll << 31) - 2, STOF_E_INVALID, "chain_tc2: too many tiles");
    p.total_tiles = (int)total;
    p.N = N; p.L = L; p.fast = fast;
    p.n_stages = fast ? T2_FAST_STAGES : T2_STAGES;
    p.stage_stride = fast ? T2_FAST_STAGE_BYTES : T2_STAGE_BYTES;
    p.in = in; p.x0 = x0; p.out_planes = out_planes; p.out_mask = out_mask;
    const bool even = (count % 2) == 0;
    p.stash = (even && layers[count - 1].type == TC_RES) ? 1 : 0;
    p.free_layer = even ? count - 2 : count - 1;
    p.in_layer = count - 1;
    p.in_tap = (!fast && even && layers[count - 1].taps > T2_STAGES - 1) ? T2_STAGES - 2 : layers[count - 1].taps - 1;
    const bool ends_last = layers[count - 1].type == TC_LAST;
    STOF_REQUIRE(ends_last ? out_mask != nullptr : out_planes != nullptr, STOF_E_INVALID, "chain_tc2: missing output buffer");
    const long long pairs = (total + 1) / 2;
    const int grid = 2 * (int)(pairs < kSMs / 2 ? pairs : kSMs / 2);
    if (fast) chain_tc2_kernel<true><<<grid, TC_THREADS, T2_SMEM_BYTES, st>>>(p);
    else chain_tc2_kernel<false><<<grid, TC_THREADS, T2_SMEM_BYTES, st>>>(p);
    STOF_LAUNCH_CHECK();
    return STOF_OK;
}

}  // namespace stof
