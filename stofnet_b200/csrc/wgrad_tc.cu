// Weight gradient of a 64 -> 64 k7 conv on tcgen05 (training step, reference main.py:247 loss.backward(); models/stofnet.py:30-31):
//
//     dW[co][ci][t] = sum over rows n and positions l of   G[n][co][l] * A[n][ci][l + t - 3]
//
// is, per tap, a GEMM whose reduction dimension is the POSITION: D_t (64 x 64) = G (64 x K) * A_t (64 x K)^T with K = N * L
// (655,360 for the 256 x 2560 batch).  Both operands already exist as split fp16 planes [plane hi/lo][c8][position][8 channels]
// (conv_tc.cu): 16-byte vectors of 8 channels, positions 16 bytes apart - which is exactly the no-swizzle MN-MAJOR UMMA operand
// layout (core matrix = 8 positions x 8 channels = 128 contiguous bytes; next 8 channels = next slab (SBO), next 8 positions =
// + 128 bytes (LBO)); a tap shift of the activation operand is a 16-byte shift of its descriptor.  The hi and lo planes are
// stacked along M and N, so ONE M = 128, N = 128, K = 16 MMA per (tap, 16 positions) yields all four hi/lo products
//     [G_hi; G_lo] * [A_hi; A_lo]^T = [[hh, hl], [lh, ll]]        dW = hh + 2^-11 (hl + lh)      (ll, 2^-22, is dropped)
// with fp32 accumulation in TMEM over the whole launch: 128 columns per tap, so a CTA owns 4 taps (taps 0-3 or 4-6) and the
// grid is two tap groups x 74 position shares.  Each CTA streams its tiles (128 positions of one row: 32 KB of G planes, 34 KB
// of A planes incl. the 6-position halo, zero outside the row) through a two-stage bulk-copy ring, issues 8 MMAs per tap and
// tile, and writes its accumulators once at the end; a second kernel adds the shares in a fixed order (bit-reproducible, no
// float atomics), folds the hi/lo terms and removes the power-of-two scale the gradient planes carry.
#include <cuda_fp16.h>
#include <string.h>

#include "conv_tc_shared.cuh"

namespace stof {

using namespace ptx;

constexpr int WT_TK = 128;                       // positions per tile
constexpr int WT_AR = 136;                       // rows of an activation slab: 128 + 6 halo (+ 2 pad)
constexpr int WT_GSLAB = WT_TK * 16;
constexpr int WT_ASLAB = WT_AR * 16;
constexpr int WT_GBUF = 16 * WT_GSLAB;           // 32,768 B
constexpr int WT_ABUF = 16 * WT_ASLAB;           // 34,816 B
constexpr int WT_STAGE = WT_GBUF + WT_ABUF;
constexpr int WT_STAGES = 2;                       // 3 stages measured no faster (1.52 ms per 11 layers either way): ingest-bound, not latency-bound
// (Also measured: a CTA owning ALL 7 taps of HALF of the input channels - B = [A_hi (32 ci); A_lo (32 ci)], N = 64, 448 TMEM
// columns, 49 KB per tile for 56 MMAs instead of 66.5 KB for 32.  147 us per layer instead of 139: an M = 128, K = 16 MMA with
// MN-major operands takes ~63 cycles whether N is 64 or 128 - the transposed A-operand read sets its time - so halving N halves
// the work per MMA for nothing; profiles/r11b_wgrad_nhalf_ncu_summary.txt.  And CTA pairs - one cta_group::2 MMA computing two
// taps, the peer's G operand one position earlier, each CTA holding half of B: 49 KB per tile, every tile fetched once - verified
// against all training tests and slower still, 154 us: tools/experiments/wgrad_tc_pair.cu, profiles/r11d_wgrad_pair_ncu_summary.txt.
// Slab strides of 16 bytes mod 128 instead of 0 mod 128 (in case one position's vectors of the 16 slabs collide in one bank
// group): 138 vs 139.5 us - no.)
constexpr int WT_OFF_BAR = WT_STAGES * WT_STAGE;  // full[stages] empty[stages] acc_full
constexpr int WT_OFF_TMEM = WT_OFF_BAR + (2 * WT_STAGES + 1) * 8;
constexpr int WT_SMEM_BYTES = WT_OFF_TMEM + 16;
constexpr int WT_THREADS = 6 * 32;               // producer, MMA issuer, 4 epilogue warps
constexpr int WT_SHARES = kSMs / 2;              // position shares per tap group
constexpr size_t WT_PART = (size_t)4 * 128 * 128; // floats of one CTA's partial: [tap of the group][lane][column]

struct WtParams {
    const void *g_planes;      // [N][2][8][L][8] fp16, the gradient times 2^E (E in its GradSlot)
    const void *a_planes;      // [N][2][8][L][8] fp16, the conv's input
    float *partial;            // [gridDim.x][4][128][128]
    int N, L, tiles_per_row, total_tiles, shares;
};

// instruction descriptor: kind::f16, fp16 A / B, fp32 accumulate, A and B MN-major (bits 15 / 16)
__host__ __device__ constexpr uint32_t umma_idesc_f16_mn(int M, int N) { return umma_idesc_f16(M, N) | (1u << 15) | (1u << 16); }

__global__ void __launch_bounds__(WT_THREADS, 1) wgrad_tc_kernel(const WtParams p) {
    extern __shared__ __align__(1024) uint8_t smem[];
    const uint32_t sbase = smem_u32(smem);
    const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
    const uint32_t bar0 = sbase + WT_OFF_BAR;
    auto BAR_FULL = [&](int s) { return bar0 + 8 * s; };
    auto BAR_EMPTY = [&](int s) { return bar0 + 8 * WT_STAGES + 8 * s; };
    const uint32_t BAR_ACC = bar0 + 16 * WT_STAGES;
    volatile uint32_t *s_tmem = reinterpret_cast<volatile uint32_t *>(smem + WT_OFF_TMEM);
    if (threadIdx.x == 0) {
        for (int s = 0; s < WT_STAGES; ++s) { mbar_init(BAR_FULL(s), 1); mbar_init(BAR_EMPTY(s), 1); }
        mbar_init(BAR_ACC, 1);
        fence_mbar_init();
    }
    if (warp == 1) tmem_alloc(sbase + WT_OFF_TMEM, 512);
    tc_fence_before_sync();
    __syncthreads();
    tc_fence_after_sync();
    const uint32_t tmem0 = *s_tmem;

    const int group = blockIdx.x & 1, share = blockIdx.x >> 1;
    const int t0 = 4 * group, ntaps = group ? 3 : 4;
    const int n_my = (p.total_tiles - share + p.shares - 1) / p.shares;        // tiles share, share + shares, ...

    if (warp == 0) {
        // =================================== producer ===================================================
        for (int k = 0; k < n_my; ++k) {
            const int s = k % WT_STAGES;
            const int tile = share + k * p.shares;
            const int n = tile / p.tiles_per_row, l0 = (tile - n * p.tiles_per_row) * WT_TK;
            mbar_wait(BAR_EMPTY(s), ((uint32_t)(k / WT_STAGES) & 1u) ^ 1u);
            uint8_t *gb = smem + s * WT_STAGE, *ab = gb + WT_GBUF;
            const int gv = min(WT_TK, p.L - l0);                            // valid G rows [0, gv)
            const int a_lo = max(0, 3 - l0), a_hi = min(WT_TK + 6, p.L - l0 + 3);   // valid A rows: position l0 - 3 + row in [0, L)
            if (gv < WT_TK || a_lo > 0 || a_hi < WT_TK + 6) {
                // rows outside the waveform: zero (the conv's padding for A; no contribution for G)
                for (int slab = 0; slab < 16; ++slab) {
                    for (int r = gv + lane; r < WT_TK; r += 32) *reinterpret_cast<uint4 *>(gb + slab * WT_GSLAB + r * 16) = make_uint4(0, 0, 0, 0);
                    for (int r = lane; r < a_lo; r += 32) *reinterpret_cast<uint4 *>(ab + slab * WT_ASLAB + r * 16) = make_uint4(0, 0, 0, 0);
                    for (int r = a_hi + lane; r < WT_TK + 6; r += 32) *reinterpret_cast<uint4 *>(ab + slab * WT_ASLAB + r * 16) = make_uint4(0, 0, 0, 0);
                }
                fence_proxy_async_smem();
            }
            __syncwarp();
            if (lane == 0) {
                const uint32_t gbytes = (uint32_t)gv * 16u, abytes = (uint32_t)(a_hi - a_lo) * 16u;
                mbar_arrive_expect_tx(BAR_FULL(s), 16u * (gbytes + abytes));
                const __half *gsrc = reinterpret_cast<const __half *>(p.g_planes) + (size_t)n * 16 * p.L * 8;
                const __half *asrc = reinterpret_cast<const __half *>(p.a_planes) + (size_t)n * 16 * p.L * 8;
                for (int slab = 0; slab < 16; ++slab) {
                    bulk_g2s(sbase + s * WT_STAGE + slab * WT_GSLAB, gsrc + ((size_t)slab * p.L + l0) * 8, gbytes, BAR_FULL(s));
                    bulk_g2s(sbase + s * WT_STAGE + WT_GBUF + slab * WT_ASLAB + a_lo * 16, asrc + ((size_t)slab * p.L + l0 - 3 + a_lo) * 8,
                             abytes, BAR_FULL(s));
                }
            }
            __syncwarp();
        }
    } else if (warp == 1) {
        // =================================== MMA issuer ==================================================
        // descriptor: start address, LBO = 128 B (next 8 positions = next K core block), SBO = slab (next 8 channels)
        constexpr uint32_t idesc = umma_idesc_f16_mn(128, 128);
        constexpr uint32_t kLbo = (128u >> 4) << 16;
        constexpr uint32_t kGHi = ((uint32_t)WT_GSLAB >> 4) | (1u << 14), kAHi = ((uint32_t)WT_ASLAB >> 4) | (1u << 14);
        for (int k = 0; k < n_my; ++k) {
            const int s = k % WT_STAGES;
            mbar_wait(BAR_FULL(s), (uint32_t)(k / WT_STAGES) & 1u);
            tc_fence_after_sync();
            if (elect_one()) {
                const uint32_t g0 = ((sbase + s * WT_STAGE) >> 4) | kLbo;
                const uint32_t a0 = ((sbase + s * WT_STAGE + WT_GBUF) >> 4) | kLbo;
                for (int tl = 0; tl < ntaps; ++tl) {
#pragma unroll
                    for (int ks = 0; ks < WT_TK / 16; ++ks)
                        umma_f16(tmem0 + 128u * (uint32_t)tl, make_desc(g0 + (uint32_t)(16 * ks), kGHi),
                                 make_desc(a0 + (uint32_t)(t0 + tl + 16 * ks), kAHi), idesc, (k > 0 || ks > 0) ? 1u : 0u);
                }
                umma_commit(BAR_EMPTY(s));
                if (k == n_my - 1) umma_commit(BAR_ACC);
            }
            __syncwarp();
        }
    } else {
        // =================================== epilogue (once) =============================================
        // lane = row of [G_hi (64 co); G_lo (64 co)], column = [A_hi (64 ci) | A_lo (64 ci)]; the lo x lo quadrant is not stored
        const int quad = warp & 3;
        const int row = 32 * quad + lane;
        const uint32_t lane_addr = (uint32_t)(32 * quad) << 16;
        float *out = p.partial + (size_t)blockIdx.x * WT_PART;
        if (n_my > 0) {
            mbar_wait(BAR_ACC, 0u);
            tc_fence_after_sync();
            const int ncol = (row < 64) ? 128 : 64;
            for (int tl = 0; tl < ntaps; ++tl) {
                float4 *dst = reinterpret_cast<float4 *>(out + ((size_t)tl * 128 + row) * 128);
#pragma unroll 1
                for (int c = 0; c < ncol; c += 16) {
                    uint32_t d[16];
                    tmem_ld16(tmem0 + 128u * (uint32_t)tl + lane_addr + (uint32_t)c, d);
                    tmem_ld_wait();
#pragma unroll
                    for (int i = 0; i < 4; ++i)
                        dst[(c >> 2) + i] = make_float4(__uint_as_float(d[4 * i]), __uint_as_float(d[4 * i + 1]), __uint_as_float(d[4 * i + 2]),
                                                        __uint_as_float(d[4 * i + 3]));
                }
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

// dW[co][ci][t] (PyTorch layout) = 2^-k * sum over shares (fixed order) of  hh + 2^-11 (hl + lh).
// One block per (tap, co, half of the ci): lane = ci (128-byte coalesced reads of the partials), warp w adds the shares
// w, w + 8, ... in order, and the eight sub-sums are added in order: as deterministic as a single chain, with an eighth of its
// dependent-load latency (a lone thread per output walking 74 shares took 42 us per layer: 0.47 ms of the 256-row step)
__global__ void __launch_bounds__(256)
wgrad_tc_reduce_kernel(const float *__restrict__ partial, int shares, int tiles, const GradSlot *__restrict__ slot,
                       float *__restrict__ dW) {
    const int half = blockIdx.x & 1, co = (blockIdx.x >> 1) & 63, t = blockIdx.x >> 7;
    const int warp = threadIdx.x >> 5, ci = 32 * half + (threadIdx.x & 31);
    const int group = t >> 2, tl = t & 3;
    float hh = 0.0f, x = 0.0f;
    const int used = shares < tiles ? shares : tiles;            // shares beyond the tile count wrote nothing
    for (int sh = warp; sh < used; sh += 8) {
        const float *P = partial + ((size_t)(2 * sh + group) * 4 + tl) * 128 * 128;
        hh += __ldg(P + (size_t)co * 128 + ci);
        x += __ldg(P + (size_t)co * 128 + 64 + ci) + __ldg(P + (size_t)(64 + co) * 128 + ci);
    }
    __shared__ float red[2][8][32];
    red[0][warp][threadIdx.x & 31] = hh;
    red[1][warp][threadIdx.x & 31] = x;
    __syncthreads();
    if (warp != 0) return;
    hh = 0.0f; x = 0.0f;
#pragma unroll
    for (int w = 0; w < 8; ++w) { hh += red[0][w][threadIdx.x]; x += red[1][w][threadIdx.x]; }
    float inv = 1.0f;
    if (slot != nullptr) {                                           // the gradient planes hold g * 2^E
        int d = -__ldg(&slot->E);
        d = d < -126 ? -126 : (d > 127 ? 127 : d);
        inv = __uint_as_float((unsigned)(d + 127) << 23);
    }
    dW[((size_t)co * 64 + ci) * 7 + t] = (hh + TC_LO_INV * x) * inv;
}

// db[c] = sum of bias_partial[q][c] over the parts q (grad_to_planes_kernel): thread t adds q = t, t + 256, ... in order, then a
// fixed tree
__global__ void __launch_bounds__(256) bias_parts_reduce_kernel(const float *__restrict__ partial, int parts, float *__restrict__ db,
                                                                const GradSlot *__restrict__ slot) {
    const int c = blockIdx.x;
    float s = 0.0f;
    for (int q = threadIdx.x; q < parts; q += 256) s += __ldg(partial + (size_t)q * 64 + c);
    __shared__ float red[256];
    red[threadIdx.x] = s;
    __syncthreads();
    for (int o = 128; o > 0; o >>= 1) {
        if (threadIdx.x < o) red[threadIdx.x] += red[threadIdx.x + o];
        __syncthreads();
    }
    if (threadIdx.x == 0) {
        float inv = 1.0f;
        if (slot != nullptr) {
            int d = -__ldg(&slot->E);
            d = d < -126 ? -126 : (d > 127 ? 127 : d);
            inv = __uint_as_float((unsigned)(d + 127) << 23);
        }
        db[c] = red[0] * inv;
    }
}

size_t wgrad_tc_scratch_floats() { return (size_t)2 * WT_SHARES * WT_PART; }

// g_planes / a_planes: split planes of the layer's output gradient (g * 2^E, slot) and of its input; scratch:
// wgrad_tc_scratch_floats() floats.  (The bias gradient comes from the partial sums of launch_grad_to_planes, conv_tc.cu.)
int launch_wgrad_tc(const void *g_planes, const void *a_planes, const GradSlot *slot, int N, int L, float *scratch, float *dW,
                    cudaStream_t st) {
    static SmemAttrOnce attr;
    { const int rc = ensure_dyn_smem(wgrad_tc_kernel, WT_SMEM_BYTES, attr); if (rc) return rc; }
    WtParams p;
    p.g_planes = g_planes; p.a_planes = a_planes; p.partial = scratch;
    p.N = N; p.L = L;
    p.tiles_per_row = (L + WT_TK - 1) / WT_TK;
    const long long total = (long long)N * p.tiles_per_row;
    STOF_REQUIRE(total < (1ll << 31), STOF_E_INVALID, "wgrad_tc: too many tiles");
    p.total_tiles = (int)total;
    p.shares = (int)(total < WT_SHARES ? total : WT_SHARES);
    wgrad_tc_kernel<<<2 * p.shares, WT_THREADS, WT_SMEM_BYTES, st>>>(p);
    STOF_LAUNCH_CHECK();
    wgrad_tc_reduce_kernel<<<7 * 64 * 2, 256, 0, st>>>(scratch, p.shares, p.total_tiles, slot, dW);
    STOF_LAUNCH_CHECK();
    return STOF_OK;
}

int launch_bias_parts_reduce(const float *bias_partial, int parts, float *db, cudaStream_t st, const GradSlot *slot) {
    bias_parts_reduce_kernel<<<64, 256, 0, st>>>(bias_partial, parts, db, slot);
    STOF_LAUNCH_CHECK();
    return STOF_OK;
}

}  // namespace stof
