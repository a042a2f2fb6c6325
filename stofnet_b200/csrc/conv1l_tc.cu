// ONE 64 -> 64 k7 conv layer per launch on tcgen05, weight-stationary: the training step's form of the chain kernel's layer
// (reference models/stofnet.py:51-62 forward; main.py:247 loss.backward() for the data gradients).
//
// The training step keeps every layer's activation, so it cannot fuse layers the way inference does (conv_tc.cu), and a
// one-layer launch of the chain kernel takes in 176 KB per 256-position tile - 64 KB of input planes and the layer's seven
// 16 KB weight taps again for every tile - for 6,272 cycles of MMAs: at the ~22 B/clk an SM's bulk copies sustain under MMA
// load that is 8,000 cycles, and the measured tile period was 12,000.  Here the layer's 112 KB of packed taps are loaded ONCE
// per CTA and stay in shared memory; the input streams through three 34 KB half-tile buffers (128 positions + the 6-position
// halo of one waveform row, zero outside it), so a tile costs 68 KB of bulk copies.
//
// Work unit = half-tile of 128 positions (= one M = 128 MMA block).  Per CTA the half-tiles alternate between two parity
// classes; each class has its own MMA-issuing warp, its own group of four epilogue warps (TMEM lane quadrants) and two of the
// four 128-column accumulators [D0 64 | D1 64], so both issue streams keep the tensor pipe busy (as the chain kernel's two
// blocks do) while each epilogue group drains the class's previous accumulator.  Arithmetic, operand layouts and the order of the
// accumulation are those of conv_tc.cu (fp16 hi/lo split, [D0 | D1] += X_hi [W_hi; W_lo]^T, D1 += X_lo W_hi^T; result D0 +
// 2^-11 D1), so the results equal the chain kernel's one-layer launch bit for bit.
//
// (Measured and not kept: an L2 prefetch of the half-tile that takes a buffer next - 92 vs 90 us per forward layer at 256 x 2560:
// the input copies are not what the MMAs wait for; fetching the activation's lo plane only when a hi value is zero in the lrelu'
// mode - within noise.)
//
// Epilogue variants: bias + LeakyReLU, bias + skip (split planes from global memory), and the gradient mode of kernels.cuh
// (TcGrad: rescale by a power of two chosen from the operands' slots, + skip gradient or * lrelu'(activation), max |out| and
// per-CTA bias-gradient sums recorded on the way).
#include <cuda_fp16.h>
#include <string.h>

#include "conv_tc_shared.cuh"

namespace stof {

using namespace ptx;

constexpr int C1_R = 136;                                   // rows per (plane, c8) slab of a half-tile buffer: 128 + 6 halo (+ 2 pad)
constexpr int C1_SLAB = C1_R * 16;
constexpr int C1_HBUF = 16 * C1_SLAB;                       // 34,816 B: [hi 8 slabs; lo 8 slabs]
constexpr int C1_NBUF = 3;
constexpr int C1_OFF_W = 0;                                 // 7 taps x 16 KB, resident
constexpr int C1_OFF_IN = 7 * TC_STAGE_BYTES;               // 114,688
constexpr int C1_OFF_BIAS = C1_OFF_IN + C1_NBUF * C1_HBUF;  // 219,136
constexpr int C1_OFF_BAR = C1_OFF_BIAS + 64 * 4;            // w_full[7], in_full[3], in_empty[3], acc_full[4], acc_empty[4]
constexpr int C1_NBAR = 7 + 2 * C1_NBUF + 8;
constexpr int C1_OFF_TMEM = C1_OFF_BAR + C1_NBAR * 8;
constexpr int C1_SMEM_BYTES = C1_OFF_TMEM + 16;
static_assert(C1_SMEM_BYTES <= 232448, "conv1l kernel exceeds the 227 KB shared-memory limit");

struct C1Params {
    const void *in;            // split planes [N][2][8][L][8] fp16
    void *out;                 // the same layout
    const void *w;             // packed taps [7][8 c8][W_hi 64 | W_lo 64][8] fp16 (TcLayer::w)
    const float *bias;         // [64]
    const void *aux;           // TC_LONGSKIP: skip planes; gradient modes 2 / 3: skip gradient / activation planes
    int type;                  // TC_LEAKY, TC_LONGSKIP or TC_NONE
    int N, L, halves_per_row, total_halves;
    TcGrad grad;               // mode 0: off
};

template <bool GRAD>
__global__ void __launch_bounds__(TC_THREADS, 1) conv1l_tc_kernel(const C1Params p) {
    extern __shared__ __align__(1024) uint8_t smem[];
    const uint32_t sbase = smem_u32(smem);
    const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
    const uint32_t bar0 = sbase + C1_OFF_BAR;
    auto BAR_W_FULL = [&](int t) { return bar0 + 8 * t; };           // one per tap: the first MMAs start behind tap 0, not behind all 112 KB
    auto BAR_IN_FULL = [&](int b) { return bar0 + 8 * (7 + b); };
    auto BAR_IN_EMPTY = [&](int b) { return bar0 + 8 * (7 + C1_NBUF + b); };
    auto BAR_ACC_FULL = [&](int a) { return bar0 + 8 * (7 + 2 * C1_NBUF + a); };
    auto BAR_ACC_EMPTY = [&](int a) { return bar0 + 8 * (7 + 2 * C1_NBUF + 4 + a); };
    float *s_bias = reinterpret_cast<float *>(smem + C1_OFF_BIAS);
    volatile uint32_t *s_tmem = reinterpret_cast<volatile uint32_t *>(smem + C1_OFF_TMEM);

    if (threadIdx.x < 64) s_bias[threadIdx.x] = __ldg(p.bias + threadIdx.x);
    if (threadIdx.x == 0) {
        for (int t = 0; t < 7; ++t) mbar_init(BAR_W_FULL(t), 1);
        for (int b = 0; b < C1_NBUF; ++b) { mbar_init(BAR_IN_FULL(b), 1); mbar_init(BAR_IN_EMPTY(b), 1); }
        for (int a = 0; a < 4; ++a) { mbar_init(BAR_ACC_FULL(a), 1); mbar_init(BAR_ACC_EMPTY(a), 4); }   // four epilogue warps drain it
        fence_mbar_init();
    }
    if (warp == 1) tmem_alloc(sbase + C1_OFF_TMEM, 512);
    tc_fence_before_sync();
    __syncthreads();
    tc_fence_after_sync();
    const uint32_t tmem0 = *s_tmem;

    // this CTA's half-tiles: blockIdx.x, + gridDim.x, ...; the i-th of them uses input buffer i % 3, parity class i & 1 and that
    // class's accumulator (i >> 1) & 1
    const int n_my = (p.total_halves - (int)blockIdx.x + (int)gridDim.x - 1) / (int)gridDim.x;
    auto acc_of = [](int i) { return 2 * (i & 1) + ((i >> 1) & 1); };

    if (warp == 0) {
        // =================================== producer ===================================================
        if (lane == 0) {
            for (int t = 0; t < 7; ++t) {
                mbar_arrive_expect_tx(BAR_W_FULL(t), TC_STAGE_BYTES);
                bulk_g2s(sbase + C1_OFF_W + t * TC_STAGE_BYTES, reinterpret_cast<const uint8_t *>(p.w) + (size_t)t * TC_STAGE_BYTES,
                         TC_STAGE_BYTES, BAR_W_FULL(t));
            }
        }
        for (int i = 0; i < n_my; ++i) {
            const int b = i % C1_NBUF;
            const int h = (int)blockIdx.x + i * (int)gridDim.x;
            const int n = h / p.halves_per_row, l0 = (h - n * p.halves_per_row) * 128;
            mbar_wait(BAR_IN_EMPTY(b), ((uint32_t)(i / C1_NBUF) & 1u) ^ 1u);
            uint8_t *hb = smem + C1_OFF_IN + b * C1_HBUF;
            // buffer row r <-> position l0 - 3 + r; rows outside the waveform are the conv's zero padding
            const int r_lo = max(0, 3 - l0), r_hi = min(134, p.L - l0 + 3);
            if (r_lo > 0 || r_hi < 134) {
                for (int slab = 0; slab < 16; ++slab) {
                    for (int r = lane; r < r_lo; r += 32) *reinterpret_cast<uint4 *>(hb + slab * C1_SLAB + r * 16) = make_uint4(0, 0, 0, 0);
                    for (int r = r_hi + lane; r < 134; r += 32) *reinterpret_cast<uint4 *>(hb + slab * C1_SLAB + r * 16) = make_uint4(0, 0, 0, 0);
                }
                fence_proxy_async_smem();
            }
            __syncwarp();
            if (lane == 0) {
                const uint32_t bytes = (uint32_t)(r_hi - r_lo) * 16u;
                mbar_arrive_expect_tx(BAR_IN_FULL(b), 16u * bytes);
                const __half *src = reinterpret_cast<const __half *>(p.in) + (size_t)n * 16 * p.L * 8;
                for (int slab = 0; slab < 16; ++slab)
                    bulk_g2s(sbase + C1_OFF_IN + b * C1_HBUF + slab * C1_SLAB + r_lo * 16, src + ((size_t)slab * p.L + l0 - 3 + r_lo) * 8, bytes,
                             BAR_IN_FULL(b));
            }
            __syncwarp();
        }
    } else if (warp < 3) {
        // =================================== MMA issuers (one per parity class) ==========================
        const int cls = warp - 1;
        constexpr uint32_t kDescHi = (128u >> 4) | (1u << 14);
        constexpr uint32_t idesc_a = umma_idesc_f16(128, 128), idesc_b = umma_idesc_f16(128, 64);
        for (int i = cls; i < n_my; i += 2) {
            const int b = i % C1_NBUF, a = acc_of(i);
            mbar_wait(BAR_IN_FULL(b), (uint32_t)(i / C1_NBUF) & 1u);
            mbar_wait(BAR_ACC_EMPTY(a), ((uint32_t)(i >> 2) & 1u) ^ 1u);
            tc_fence_after_sync();
            if (elect_one()) {
                const uint32_t a_base = ((sbase + C1_OFF_IN + b * C1_HBUF) >> 4) | ((uint32_t)(C1_SLAB >> 4) << 16);
                const uint32_t w_base = ((sbase + C1_OFF_W) >> 4) | (128u << 16);
                const uint32_t d = tmem0 + 128u * (uint32_t)a;
                for (int t = 0; t < 7; ++t) {
                    if (i < 2) { mbar_wait(BAR_W_FULL(t), 0u); tc_fence_after_sync(); }     // this issuer's first half-tile: tap by tap
                    const uint32_t a_lo = a_base + (uint32_t)t;
                    const uint32_t w_lo = w_base + (uint32_t)t * (TC_STAGE_BYTES >> 4);
#pragma unroll
                    for (int ks = 0; ks < 4; ++ks) {
                        umma_f16(d, make_desc(a_lo + 2 * ks * C1_R, kDescHi), make_desc(w_lo + 2 * ks * 128, kDescHi), idesc_a,
                                 (t > 0 || ks > 0) ? 1u : 0u);
                        umma_f16(d + 64u, make_desc(a_lo + (8 + 2 * ks) * C1_R, kDescHi), make_desc(w_lo + 2 * ks * 128, kDescHi), idesc_b, 1u);
                    }
                }
                umma_commit(BAR_IN_EMPTY(b));
                umma_commit(BAR_ACC_FULL(a));
            }
            __syncwarp();
        }
    } else {
        // =================================== epilogue (four warps per parity class) ======================
        const int quad = warp & 3;                               // TMEM lane quadrant of this warp
        const int cls = (warp - 3) >> 2;
        const int trow = 32 * quad + lane;
        const uint32_t lane_addr = (uint32_t)(32 * quad) << 16;
        const int type = p.type;
        // gradient mode (kernels.cuh, TcGrad): out_stored = (conv(in_stored) * g_sk [+ aux_stored * g_ss]) [* lrelu'(aux)], the powers
        // of two chosen so that a bound of max |out| lands in [2^10, 2^11)
        float g_sk = 1.0f, g_ss = 1.0f, g_amax = 0.0f;
        float g_bs0[32], g_bs1[32];                              // per pass (channels 0-31 / 32-63): static register indices
        int g_mode = 0;
        if (GRAD) {
            g_mode = p.grad.mode;
#pragma unroll
            for (int i = 0; i < 32; ++i) { g_bs0[i] = 0.0f; g_bs1[i] = 0.0f; }
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
        const bool use_aux = (type == TC_LONGSKIP) || (GRAD && g_mode >= 2);
        for (int i = cls; i < n_my; i += 2) {
            const int a = 2 * (i & 1) + ((i >> 1) & 1);
            const int h = (int)blockIdx.x + i * (int)gridDim.x;
            const int n = h / p.halves_per_row, l0 = (h - n * p.halves_per_row) * 128;
            const int l = l0 + trow;
            const bool inside = l < p.L;
            const uint4 *aux = reinterpret_cast<const uint4 *>(p.aux) + (size_t)n * 16 * p.L + l;
            uint4 *dst = reinterpret_cast<uint4 *>(p.out) + (size_t)n * 16 * p.L + l;
            // the first pass's skip / activation vectors: requested before the accumulator wait, behind the block's MMAs
            uint4 xs_h[4], xs_l[4];
            if (use_aux) {
#pragma unroll
                for (int q = 0; q < 4; ++q) {
                    xs_h[q] = inside ? aux[(size_t)q * p.L] : make_uint4(0, 0, 0, 0);     // plain loads: the skip may be updated in place (out == aux)
                    xs_l[q] = inside ? aux[(size_t)(8 + q) * p.L] : make_uint4(0, 0, 0, 0);
                }
            }
            mbar_wait(BAR_ACC_FULL(a), (uint32_t)(i >> 2) & 1u);
            tc_fence_after_sync();
            const uint32_t d_tmem = tmem0 + 128u * (uint32_t)a + lane_addr;
#pragma unroll 1
            for (int ps = 0; ps < 2; ++ps) {
                // this pass's 32 channels: columns [32 ps, +32) of D0 and of D1 (at +64)
                uint32_t d0[32], d1[32];
                tmem_ld16(d_tmem + 32 * ps, *reinterpret_cast<uint32_t(*)[16]>(&d0[0]));
                tmem_ld16(d_tmem + 32 * ps + 16, *reinterpret_cast<uint32_t(*)[16]>(&d0[16]));
                tmem_ld16(d_tmem + 64 + 32 * ps, *reinterpret_cast<uint32_t(*)[16]>(&d1[0]));
                tmem_ld16(d_tmem + 64 + 32 * ps + 16, *reinterpret_cast<uint32_t(*)[16]>(&d1[16]));
                tmem_ld_wait();
                if (ps == 1) {
                    tc_fence_before_sync();
                    __syncwarp();
                    if (lane == 0) mbar_arrive(BAR_ACC_EMPTY(a));              // accumulator read: hand it back
                }
#pragma unroll
                for (int q = 0; q < 4; ++q) {
                    const int c8 = 4 * ps + q;
                    float v[8];
#pragma unroll
                    for (int k = 0; k < 8; ++k)
                        v[k] = (__uint_as_float(d0[8 * q + k]) + TC_LO_INV * __uint_as_float(d1[8 * q + k])) + s_bias[8 * c8 + k];
                    if (type == TC_LEAKY) {
#pragma unroll
                        for (int k = 0; k < 8; ++k) v[k] = fmaxf(v[k], kLeakySlope * v[k]);
                    } else if (type == TC_LONGSKIP) {
                        float fh[8], fl[8];
                        unpack8(xs_h[q], fh);
                        unpack8(xs_l[q], fl);
#pragma unroll
                        for (int k = 0; k < 8; ++k) v[k] += fh[k] + TC_LO_INV * fl[k];
                    } else if (GRAD) {
#pragma unroll
                        for (int k = 0; k < 8; ++k) v[k] *= g_sk;
                        if (g_mode == 2) {
                            float fh[8], fl[8];
                            unpack8(xs_h[q], fh);
                            unpack8(xs_l[q], fl);
#pragma unroll
                            for (int k = 0; k < 8; ++k) v[k] = fmaf(fh[k] + TC_LO_INV * fl[k], g_ss, v[k]);
                        } else if (g_mode == 3) {
                            const uint32_t hw[4] = {xs_h[q].x, xs_h[q].y, xs_h[q].z, xs_h[q].w};
                            const uint32_t lw[4] = {xs_l[q].x, xs_l[q].y, xs_l[q].z, xs_l[q].w};
#pragma unroll
                            for (int k = 0; k < 8; ++k) {
                                const uint32_t hv = (hw[k >> 1] >> (16 * (k & 1))) & 0xffffu, lv = (lw[k >> 1] >> (16 * (k & 1))) & 0xffffu;
                                // a > 0 <=> its hi plane is positive, or zero with a positive lo plane
                                const bool pos = ((hv & 0x7fffu) != 0u) ? ((hv & 0x8000u) == 0u)
                                                                        : (((lv & 0x7fffu) != 0u) && ((lv & 0x8000u) == 0u));
                                v[k] *= pos ? 1.0f : kLeakySlope;
                            }
                        }
                        if (inside) {
#pragma unroll
                            for (int k = 0; k < 8; ++k) {
                                g_amax = fmaxf(g_amax, fabsf(v[k]));
                                if (ps == 0) g_bs0[8 * q + k] += v[k]; else g_bs1[8 * q + k] += v[k];
                            }
                        }
                    }
                    uint4 uh, ul;
                    split_pair(v[0], v[1], uh.x, ul.x);
                    split_pair(v[2], v[3], uh.y, ul.y);
                    split_pair(v[4], v[5], uh.z, ul.z);
                    split_pair(v[6], v[7], uh.w, ul.w);
                    if (inside) {
                        dst[(size_t)c8 * p.L] = uh;
                        dst[(size_t)(8 + c8) * p.L] = ul;
                    }
                }
                if (ps == 0 && use_aux) {
                    // the second pass's vectors (their latency hides behind the TMEM loads and the first chunk's math of that pass)
#pragma unroll
                    for (int q = 0; q < 4; ++q) {
                        xs_h[q] = inside ? aux[(size_t)(4 + q) * p.L] : make_uint4(0, 0, 0, 0);
                        xs_l[q] = inside ? aux[(size_t)(12 + q) * p.L] : make_uint4(0, 0, 0, 0);
                    }
                }
            }
        }
        if (GRAD) {
            // max |stored output| of this launch; bias-gradient sums of this CTA: lanes (rows) by shuffle tree here, the eight
            // epilogue warps in order after the CTA-wide barrier below (the input buffers are free then)
#pragma unroll
            for (int o = 16; o > 0; o >>= 1) g_amax = fmaxf(g_amax, __shfl_xor_sync(0xffffffffu, g_amax, o));
            if (lane == 0 && g_amax > 0.0f) atomicMax(&p.grad.g_out->amax_bits, __float_as_uint(g_amax));
#pragma unroll
            for (int k = 0; k < 32; ++k)
#pragma unroll
                for (int o = 16; o > 0; o >>= 1) {
                    g_bs0[k] += __shfl_xor_sync(0xffffffffu, g_bs0[k], o);
                    g_bs1[k] += __shfl_xor_sync(0xffffffffu, g_bs1[k], o);
                }
        }
        tc_fence_before_sync();
        __syncthreads();                                          // every MMA has completed, every half-tile is stored
        if (GRAD) {
            float *red = reinterpret_cast<float *>(smem + C1_OFF_IN);            // [8 epilogue warps][64]
            float m0 = 0.0f, m1 = 0.0f;
#pragma unroll
            for (int k = 0; k < 32; ++k) if (lane == k) { m0 = g_bs0[k]; m1 = g_bs1[k]; }
            red[(warp - 3) * 64 + lane] = m0;
            red[(warp - 3) * 64 + 32 + lane] = m1;
        }
    }
    if (warp < 3) { tc_fence_before_sync(); __syncthreads(); }
    if (GRAD) {
        __syncthreads();
        if (p.grad.bias_partial != nullptr && threadIdx.x < 64) {
            const float *red = reinterpret_cast<const float *>(smem + C1_OFF_IN) + threadIdx.x;
            float s = 0.0f;
#pragma unroll
            for (int w = 0; w < 8; ++w) s += red[w * 64];
            p.grad.bias_partial[(size_t)blockIdx.x * 64 + threadIdx.x] = s;
        }
    }
    if (warp == 1) {
        tc_fence_after_sync();
        tmem_dealloc(tmem0, 512);
    }
}

int conv1l_tc_grid(int N, int L) {
    const long long total = (long long)N * ((L + 127) / 128);
    return (int)(total < kSMs ? total : kSMs);
}

// layer: a 64 -> 64 k7 TcLayer (its packed taps and bias); type: TC_LEAKY, TC_LONGSKIP (aux = skip planes) or TC_NONE; grad (or
// NULL): gradient mode (type TC_NONE; aux inside it).  bias_partial of the gradient mode has conv1l_tc_grid(N, L) parts.
int launch_conv1l_tc(const TcLayer *layer, int type, const void *in, const void *aux, void *out, int N, int L, cudaStream_t st,
                     const TcGrad *grad) {
    static SmemAttrOnce attr_fwd, attr_grad;
    STOF_REQUIRE(layer && layer->taps == 7 && layer->stage_bytes == TC_STAGE_BYTES && layer->n_out == 64, STOF_E_INVALID,
                 "conv1l_tc: a 64 -> 64 k7 layer is required");
    STOF_REQUIRE(in && out && (type == TC_LEAKY || type == TC_NONE || (type == TC_LONGSKIP && aux)), STOF_E_INVALID,
                 "conv1l_tc: bad arguments (type %d)", type);
    STOF_REQUIRE(grad == nullptr || (type == TC_NONE && grad->mode >= 1 && grad->mode <= 3 && grad->g_in && grad->g_out &&
                                     (grad->mode == 1 || grad->aux) && (grad->mode != 2 || grad->g_aux)),
                 STOF_E_INVALID, "conv1l_tc: gradient mode needs type TC_NONE and its slots");
    { const int rc = grad ? ensure_dyn_smem(conv1l_tc_kernel<true>, C1_SMEM_BYTES, attr_grad)
                          : ensure_dyn_smem(conv1l_tc_kernel<false>, C1_SMEM_BYTES, attr_fwd); if (rc) return rc; }
    C1Params p;
    memset(&p, 0, sizeof(p));
    p.in = in; p.out = out; p.w = layer->w; p.bias = layer->bias; p.aux = grad ? grad->aux : aux; p.type = type;
    p.N = N; p.L = L;
    p.halves_per_row = (L + 127) / 128;
    const long long total = (long long)N * p.halves_per_row;
    STOF_REQUIRE(total < (1ll << 31), STOF_E_INVALID, "conv1l_tc: too many tiles");
    p.total_halves = (int)total;
    const int grid = conv1l_tc_grid(N, L);
    if (grad) { p.grad = *grad; conv1l_tc_kernel<true><<<grid, TC_THREADS, C1_SMEM_BYTES, st>>>(p); }
    else conv1l_tc_kernel<false><<<grid, TC_THREADS, C1_SMEM_BYTES, st>>>(p);
    STOF_LAUNCH_CHECK();
    return STOF_OK;
}

}  // namespace stof
