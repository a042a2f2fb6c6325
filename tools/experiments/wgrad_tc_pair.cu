// EXPERIMENT (round 2, verified, not shipped): wgrad_tc on CTA pairs.  This is the kernel as it was built into
// stofnet_b200/csrc/wgrad_tc.cu behind STOF_WGRAD_PAIR=1 (it uses that file's constants and WtParams; the launcher set
// tiles_per_row = (L + 128) / 128, grid = 2 x pairs, and the reduce kernel read tap t from CTA (t & 1), accumulator (t >> 1)).
// Result at 256 x 2560: all training tests green (gradients vs the reference's autograd, bit-identical run to run), 154 us per
// layer against 139 us for the shipped kernel (profiles/r11d_wgrad_pair_ncu_summary.txt): each CTA takes in 49 KB per tile
// instead of 66.5 KB and every tile is fetched once, but a tile now costs ~3,540 cycles instead of ~3,360 - the 32
// cta_group::2 MMAs of a tile take ~110 cycles each with MN-major operands (each SM needs the other half of B from its peer for
// every MMA), where the single-CTA M = 128, N = 128 MMA takes ~63.  The step got 0.2 ms slower (7.74 vs 7.54 ms).
// ---------------------------------------------------------------------------------------------------------------------------
// The same product on CTA PAIRS (2-CTA clusters, tcgen05 cta_group::2), built to cut the bytes an SM takes in per MMA - the bound
// of the kernel above.  One M = 256 MMA spans both CTAs: each contributes its own 128 A-operand rows and HALF of the B operand.
//   * B = [A_hi (64 ci); A_lo (64 ci)] at tap shift t_B: the leader holds the hi plane's 8 slabs, the peer the lo plane's - 17 KB
//     per tile each instead of 34 KB;
//   * A-operand rows: both CTAs hold [G_hi; G_lo], but the peer's copy starts ONE POSITION EARLIER.  With the same B,
//       leader:  sum_k G[l0 + k]     A[l0 + k + t_B - 3]  = tap t_B       peer:  sum_k G[l0 - 1 + k] A[l0 + k + t_B - 3]  = tap t_B + 1
//     so four MMAs per 16 positions (t_B = 0, 2, 4, 6) give all 7 taps (the peer's "tap 7" is discarded), 4 x 128 TMEM columns in
//     each CTA, and every tile is fetched once per pair: 49 KB per CTA and tile for 32 MMAs (66.5 KB before, and every tile
//     went through two CTAs).  The peer's position window [l0 - 1, l0 + 126] needs one more tile per row to reach position L - 1.
// Barriers: a bulk copy signals its own CTA only, so the peer's warp 1 relays "my stage has landed" to the leader (release /
// acquire at cluster scope); the leader's commits multicast "stage free" / "accumulators complete" to both CTAs.
constexpr int WP_ABUF = 8 * WT_ASLAB;            // this CTA's half of B: one plane's 8 slabs, 17,408 B
constexpr int WP_STAGE = WT_GBUF + WP_ABUF;      // 50,176 B
constexpr int WP_STAGES = 3;
constexpr int WP_OFF_BAR = WP_STAGES * WP_STAGE; // full[S] peer_full[S] empty[S] acc_full
constexpr int WP_OFF_TMEM = WP_OFF_BAR + (3 * WP_STAGES + 1) * 8;
constexpr int WP_SMEM_BYTES = WP_OFF_TMEM + 16;

__global__ void __cluster_dims__(2, 1, 1) __launch_bounds__(WT_THREADS, 1) wgrad_tc_pair_kernel(const WtParams p) {
    extern __shared__ __align__(1024) uint8_t smem[];
    const uint32_t sbase = smem_u32(smem);
    const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
    const int rank = (int)cluster_ctarank();
    const bool leader = (rank == 0);
    const uint32_t bar0 = sbase + WP_OFF_BAR;
    auto BAR_FULL = [&](int s) { return bar0 + 8 * s; };
    auto BAR_PEER = [&](int s) { return bar0 + 8 * (WP_STAGES + s); };
    auto BAR_EMPTY = [&](int s) { return bar0 + 8 * (2 * WP_STAGES + s); };
    const uint32_t BAR_ACC = bar0 + 8 * (3 * WP_STAGES);
    volatile uint32_t *s_tmem = reinterpret_cast<volatile uint32_t *>(smem + WP_OFF_TMEM);
    if (threadIdx.x == 0) {
        for (int s = 0; s < WP_STAGES; ++s) { mbar_init(BAR_FULL(s), 1); mbar_init(BAR_PEER(s), 1); mbar_init(BAR_EMPTY(s), 1); }
        mbar_init(BAR_ACC, 1);
        fence_mbar_init();
    }
    if (warp == 1) tmem_alloc_pair(sbase + WP_OFF_TMEM, 512);
    tc_fence_before_sync();
    cluster_sync_all();                                   // barriers of BOTH CTAs initialised before any remote arrive / commit
    tc_fence_after_sync();
    const uint32_t tmem0 = *s_tmem;

    const int pair = blockIdx.x >> 1;
    const int n_my = (p.total_tiles - pair + p.shares - 1) / p.shares;          // tiles pair, pair + shares, ...

    if (warp == 0) {
        // =================================== producer (both CTAs) =======================================
        for (int k = 0; k < n_my; ++k) {
            const int s = k % WP_STAGES;
            const int tile = pair + k * p.shares;
            const int n = tile / p.tiles_per_row, l0 = (tile - n * p.tiles_per_row) * WT_TK;
            mbar_wait(BAR_EMPTY(s), ((uint32_t)(k / WP_STAGES) & 1u) ^ 1u);
            uint8_t *gb = smem + s * WP_STAGE, *ab = gb + WT_GBUF;
            const int g0 = l0 - rank;                                        // position of G row 0 (the peer: one earlier)
            const int g_lo = min(WT_TK, max(0, -g0)), g_hi = max(g_lo, min(WT_TK, p.L - g0));      // valid G rows [g_lo, g_hi)
            const int a_lo = min(WT_TK + 6, max(0, 3 - l0)), a_hi = max(a_lo, min(WT_TK + 6, p.L - l0 + 3));   // valid A rows
            if (g_lo > 0 || g_hi < WT_TK || a_lo > 0 || a_hi < WT_TK + 6) {
                for (int slab = 0; slab < 16; ++slab) {
                    for (int r = lane; r < g_lo; r += 32) *reinterpret_cast<uint4 *>(gb + slab * WT_GSLAB + r * 16) = make_uint4(0, 0, 0, 0);
                    for (int r = g_hi + lane; r < WT_TK; r += 32) *reinterpret_cast<uint4 *>(gb + slab * WT_GSLAB + r * 16) = make_uint4(0, 0, 0, 0);
                }
                for (int slab = 0; slab < 8; ++slab) {
                    for (int r = lane; r < a_lo; r += 32) *reinterpret_cast<uint4 *>(ab + slab * WT_ASLAB + r * 16) = make_uint4(0, 0, 0, 0);
                    for (int r = a_hi + lane; r < WT_TK + 6; r += 32) *reinterpret_cast<uint4 *>(ab + slab * WT_ASLAB + r * 16) = make_uint4(0, 0, 0, 0);
                }
                fence_proxy_async_smem();
            }
            __syncwarp();
            if (lane == 0) {
                const uint32_t gbytes = (uint32_t)(g_hi - g_lo) * 16u, abytes = (uint32_t)(a_hi - a_lo) * 16u;
                mbar_arrive_expect_tx(BAR_FULL(s), 16u * gbytes + 8u * abytes);
                const __half *gsrc = reinterpret_cast<const __half *>(p.g_planes) + (size_t)n * 16 * p.L * 8;
                const __half *asrc = reinterpret_cast<const __half *>(p.a_planes) + (size_t)n * 16 * p.L * 8;
                if (gbytes)
                    for (int slab = 0; slab < 16; ++slab)
                        bulk_g2s(sbase + s * WP_STAGE + slab * WT_GSLAB + g_lo * 16, gsrc + ((size_t)slab * p.L + g0 + g_lo) * 8, gbytes, BAR_FULL(s));
                if (abytes)
                    for (int i = 0; i < 8; ++i)                              // this CTA's plane of the conv's input: slabs 8 rank + i
                        bulk_g2s(sbase + s * WP_STAGE + WT_GBUF + i * WT_ASLAB + a_lo * 16,
                                 asrc + ((size_t)(8 * rank + i) * p.L + l0 - 3 + a_lo) * 8, abytes, BAR_FULL(s));
            }
            __syncwarp();
        }
    } else if (warp == 1) {
        if (leader) {
            // =================================== MMA issuer (leader) ====================================
            constexpr uint32_t idesc = umma_idesc_f16_mn(256, 128);
            constexpr uint32_t kLbo = (128u >> 4) << 16;
            constexpr uint32_t kGHi = ((uint32_t)WT_GSLAB >> 4) | (1u << 14), kAHi = ((uint32_t)WT_ASLAB >> 4) | (1u << 14);
            for (int k = 0; k < n_my; ++k) {
                const int s = k % WP_STAGES;
                const uint32_t ph = (uint32_t)(k / WP_STAGES) & 1u;
                mbar_wait(BAR_FULL(s), ph);
                mbar_wait_cluster(BAR_PEER(s), ph);
                tc_fence_after_sync();
                if (elect_one()) {
                    const uint32_t g0 = ((sbase + s * WP_STAGE) >> 4) | kLbo;
                    const uint32_t a0 = ((sbase + s * WP_STAGE + WT_GBUF) >> 4) | kLbo;
                    for (int pp = 0; pp < 4; ++pp) {
#pragma unroll
                        for (int ks = 0; ks < WT_TK / 16; ++ks)
                            umma_f16_pair(tmem0 + 128u * (uint32_t)pp, make_desc(g0 + (uint32_t)(16 * ks), kGHi),
                                          make_desc(a0 + (uint32_t)(2 * pp + 16 * ks), kAHi), idesc, (k > 0 || ks > 0) ? 1u : 0u);
                    }
                    umma_commit_pair(BAR_EMPTY(s), 3);
                    if (k == n_my - 1) umma_commit_pair(BAR_ACC, 3);
                }
                __syncwarp();
            }
        } else if (lane == 0) {
            // =================================== relay (peer) ===========================================
            for (int k = 0; k < n_my; ++k) {
                const int s = k % WP_STAGES;
                mbar_wait(BAR_FULL(s), (uint32_t)(k / WP_STAGES) & 1u);
                mbar_arrive_cluster(mapa_shared(BAR_PEER(s), 0));
            }
        }
    } else {
        // =================================== epilogue (both CTAs, once) ==================================
        // lane = row of [G_hi (64 co); G_lo (64 co)], column = [A_hi (64 ci) | A_lo (64 ci)]; the lo x lo quadrant is not stored;
        // accumulator pp of rank r is tap 2 pp + r
        const int quad = warp & 3;
        const int row = 32 * quad + lane;
        const uint32_t lane_addr = (uint32_t)(32 * quad) << 16;
        float *out = p.partial + (size_t)blockIdx.x * WT_PART;
        if (n_my > 0) {
            mbar_wait_cluster(BAR_ACC, 0u);
            tc_fence_after_sync();
            const int ncol = (row < 64) ? 128 : 64;
            for (int pp = 0; pp < 4; ++pp) {
                if (2 * pp + rank >= 7) continue;
                float4 *dst = reinterpret_cast<float4 *>(out + ((size_t)pp * 128 + row) * 128);
#pragma unroll 1
                for (int c = 0; c < ncol; c += 16) {
                    uint32_t d[16];
                    tmem_ld16(tmem0 + 128u * (uint32_t)pp + lane_addr + (uint32_t)c, d);
                    tmem_ld_wait();
#pragma unroll
                    for (int i = 0; i < 4; ++i)
                        dst[(c >> 2) + i] = make_float4(__uint_as_float(d[4 * i]), __uint_as_float(d[4 * i + 1]), __uint_as_float(d[4 * i + 2]),
                                                        __uint_as_float(d[4 * i + 3]));
                }
            }
        }
    }
    // the leader's MMAs read the peer's shared memory and the peer arrives on the leader's barriers: nobody leaves early
    tc_fence_before_sync();
    cluster_sync_all();
    if (warp == 1) {
        tc_fence_after_sync();
        tmem_dealloc_pair(tmem0, 512);
    }
}

