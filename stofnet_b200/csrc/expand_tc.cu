// Tensor-core kernel of the semi-global block's expanding half (models/stofnet.py:94-95,104-105 of the reference):
//
//     E = LeakyReLU( expand_conv( pooled ) )          pooled (N, 512, P) fp32  ->  E (N, 64, P) fp32,  Conv1d(512 -> 64, k5, pad 2)
//
// Implicit GEMM on tcgen05: M = 128 consecutive "slots" of the padded slot space (every waveform row owns P + 4 slots, two
// zero slots on either side, so a tap shift across a row boundary reads the conv's zero padding), N = 64 output channels,
// K = 512 channels x 5 taps, walked as 8 groups of 64 channels.  fp32 accuracy on fp16 tensor cores exactly as in conv_tc.cu:
//     [D0 | D1] (+)= X_hi [W_hi ; W_lo]^T   (N = 128),      D1 += X_lo W_hi^T   (N = 64),      result D0 + 2^-11 D1.
// The fp32 input is converted to the split fp16 operand layout ([plane][c8][row][8], no-swizzle K-major) on the fly by nine
// loader warps (coalesced 4-byte reads along p), double-buffered per channel group; the 40 weight taps (16 KB each) stream
// through a 5-stage bulk-copy ring; two accumulators alternate per tile so the epilogue (bias + LeakyReLU, coalesced fp32
// stores) overlaps the next tile's MMAs.
//
// Roles: warp 0 = weight producer, warp 1 = MMA issuer + TMEM owner, warps 2-10 = loaders, of which warps 3-10 are also the
// epilogue (warp & 3 = TMEM lane quadrant, (warp - 3) >> 2 = 32-channel half).
#include <cuda_fp16.h>

#include "common.cuh"
#include "kernels.cuh"
#include "tc_ptx.cuh"

namespace stof {

using namespace ptx;

constexpr int XT_THREADS = 352;
constexpr int XT_M = 128;                              // output slots per tile
constexpr int XT_R = XT_M + 4;                         // operand rows per tile (halo 2 + 2)
constexpr int XT_SLAB = XT_R * 16;                     // one (plane, c8) slab
constexpr int XT_PLANE = 8 * XT_SLAB;
constexpr int XT_BUF = 2 * XT_PLANE;                   // hi + lo planes of one 64-channel group: 33,792 B
constexpr int XT_STAGES = 5;
constexpr int XT_TAP_BYTES = 16384;                    // [c8][W_hi 64 | W_lo 64][8] fp16
constexpr int XT_GROUPS = 8, XT_TAPS = 5;
constexpr int XT_OFF_A = 0;
constexpr int XT_OFF_W = 2 * XT_BUF;
constexpr int XT_OFF_BAR = XT_OFF_W + XT_STAGES * XT_TAP_BYTES;
// barriers: w_full[5] w_empty[5] a_full[2] a_empty[2] acc_full[2] acc_empty[2]
constexpr int XT_NBAR = 2 * XT_STAGES + 8;
constexpr int XT_OFF_TMEM = XT_OFF_BAR + XT_NBAR * 8;
constexpr int XT_SMEM_BYTES = XT_OFF_TMEM + 16;
constexpr int XT_LOADER_WARPS = 9, XT_EPI_WARPS = 8;

struct ExpandParams {
    const float *pooled;      // (N, 512, P)
    float *e;                 // (N, 64, P)
    const void *w;            // packed taps [group 8][tap 5][c8 8][128][8] fp16
    const float *bias;        // (64)
    int N, P;
    long long slots;          // N * (P + 4)
    int tiles;
};

__device__ __forceinline__ uint32_t xt_cvt_sat_half2(float a, float b) {
    uint32_t r;
    asm("cvt.rn.satfinite.f16x2.f32 %0, %1, %2;" : "=r"(r) : "f"(b), "f"(a));
    return r;
}
__device__ __forceinline__ void xt_split_pair(float v0, float v1, uint32_t &hi, uint32_t &lo) {
    hi = xt_cvt_sat_half2(v0, v1);
    const float2 h = __half22float2(*reinterpret_cast<const __half2 *>(&hi));
    lo = xt_cvt_sat_half2((v0 - h.x) * 2048.0f, (v1 - h.y) * 2048.0f);
}

__global__ void __launch_bounds__(XT_THREADS, 1) expand_tc_kernel(const ExpandParams p) {
    extern __shared__ __align__(1024) uint8_t smem[];
    const uint32_t sbase = smem_u32(smem);
    const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
    const uint32_t bar0 = sbase + XT_OFF_BAR;
    auto BAR_W_FULL = [&](int s) { return bar0 + 8 * s; };
    auto BAR_W_EMPTY = [&](int s) { return bar0 + 8 * (XT_STAGES + s); };
    auto BAR_A_FULL = [&](int b) { return bar0 + 8 * (2 * XT_STAGES + b); };
    auto BAR_A_EMPTY = [&](int b) { return bar0 + 8 * (2 * XT_STAGES + 2 + b); };
    auto BAR_ACC_FULL = [&](int b) { return bar0 + 8 * (2 * XT_STAGES + 4 + b); };
    auto BAR_ACC_EMPTY = [&](int b) { return bar0 + 8 * (2 * XT_STAGES + 6 + b); };
    volatile uint32_t *s_tmem = reinterpret_cast<volatile uint32_t *>(smem + XT_OFF_TMEM);

    if (threadIdx.x == 0) {
        for (int s = 0; s < XT_STAGES; ++s) { mbar_init(BAR_W_FULL(s), 1); mbar_init(BAR_W_EMPTY(s), 1); }
        for (int b = 0; b < 2; ++b) {
            mbar_init(BAR_A_FULL(b), XT_LOADER_WARPS);
            mbar_init(BAR_A_EMPTY(b), 1);
            mbar_init(BAR_ACC_FULL(b), 1);
            mbar_init(BAR_ACC_EMPTY(b), XT_EPI_WARPS);
        }
        fence_mbar_init();
    }
    if (warp == 1) tmem_alloc(sbase + XT_OFF_TMEM, 256);
    tc_fence_before_sync();
    __syncthreads();
    tc_fence_after_sync();
    const uint32_t tmem0 = *s_tmem;
    const int PP = p.P + 4;
    int n_local = 0;
    if ((int)blockIdx.x < p.tiles) n_local = (p.tiles - 1 - (int)blockIdx.x) / (int)gridDim.x + 1;

    if (warp == 0) {
        // =================================== weight producer =============================================
        if (lane == 0) {
            int s = 0;
            uint32_t ph = 0;
            for (int it = 0; it < n_local; ++it)
                for (int gt = 0; gt < XT_GROUPS * XT_TAPS; ++gt) {
                    mbar_wait(BAR_W_EMPTY(s), ph ^ 1u);
                    mbar_arrive_expect_tx(BAR_W_FULL(s), XT_TAP_BYTES);
                    bulk_g2s(sbase + XT_OFF_W + s * XT_TAP_BYTES, reinterpret_cast<const uint8_t *>(p.w) + (size_t)gt * XT_TAP_BYTES,
                             XT_TAP_BYTES, BAR_W_FULL(s));
                    if (++s == XT_STAGES) { s = 0; ph ^= 1u; }
                }
        }
    } else if (warp == 1) {
        // =================================== MMA issuer ==================================================
        if (elect_one()) {
            constexpr uint32_t kDescHi = (128u >> 4) | (1u << 14);                   // SBO = 128 B, descriptor version 1
            constexpr uint32_t idesc_a = umma_idesc_f16(128, 128), idesc_b = umma_idesc_f16(128, 64);
            int s = 0;
            uint32_t ph = 0;
            for (int it = 0; it < n_local; ++it) {
                const int ab = it & 1;
                const uint32_t d_tmem = tmem0 + 128u * (uint32_t)ab;
                mbar_wait(BAR_ACC_EMPTY(ab), (((uint32_t)it >> 1) & 1u) ^ 1u);
                for (int g = 0; g < XT_GROUPS; ++g) {
                    const int c = it * XT_GROUPS + g, buf = c & 1;
                    mbar_wait(BAR_A_FULL(buf), ((uint32_t)c >> 1) & 1u);
                    tc_fence_after_sync();
                    const uint32_t a_base = ((sbase + XT_OFF_A + buf * XT_BUF) >> 4) | ((uint32_t)(XT_SLAB >> 4) << 16);
                    for (int t = 0; t < XT_TAPS; ++t) {
                        mbar_wait(BAR_W_FULL(s), ph);
                        tc_fence_after_sync();
                        const uint32_t a_lo = a_base + (uint32_t)t;
                        const uint32_t w_lo = ((sbase + XT_OFF_W + s * XT_TAP_BYTES) >> 4) | (128u << 16);
#pragma unroll
                        for (int ks = 0; ks < 4; ++ks) {
                            umma_f16(d_tmem, make_desc(a_lo + 2 * ks * XT_R, kDescHi), make_desc(w_lo + 2 * ks * 128, kDescHi), idesc_a,
                                     (g > 0 || t > 0 || ks > 0) ? 1u : 0u);
                            umma_f16(d_tmem + 64u, make_desc(a_lo + (8 + 2 * ks) * XT_R, kDescHi), make_desc(w_lo + 2 * ks * 128, kDescHi),
                                     idesc_b, 1u);
                        }
                        umma_commit(BAR_W_EMPTY(s));
                        if (++s == XT_STAGES) { s = 0; ph ^= 1u; }
                    }
                    umma_commit(BAR_A_EMPTY(buf));
                }
                umma_commit(BAR_ACC_FULL(ab));
            }
        }
        __syncwarp();
    } else {
        // =================================== loaders + epilogue ==========================================
        const int lt = threadIdx.x - 64;                      // 0 .. 287
        const bool ld_on = lt < 2 * XT_R;
        const int half = lt / XT_R, r = lt - half * XT_R;     // operand row r, channel chunks c8 in [4 half, +4)
        const int quad = warp & 3, egrp = (warp - 3) >> 2;    // epilogue: TMEM lane quadrant, 32-channel half (warps 3-10)
        const uint32_t lane_addr = (uint32_t)(32 * quad) << 16;
        const float4 *bias4 = reinterpret_cast<const float4 *>(p.bias) + 8 * egrp;

        auto epilogue = [&](int it) {
            const int ab = it & 1;
            const long long q = ((long long)blockIdx.x + (long long)it * gridDim.x) * XT_M + 32 * quad + lane + 2;   // slot of row m + 2
            mbar_wait(BAR_ACC_FULL(ab), ((uint32_t)it >> 1) & 1u);
            tc_fence_after_sync();
            const uint32_t d_tmem = tmem0 + 128u * (uint32_t)ab + lane_addr + 32u * (uint32_t)egrp;
            uint32_t d0[32], d1[32];
            tmem_ld16(d_tmem, *reinterpret_cast<uint32_t(*)[16]>(&d0[0]));
            tmem_ld16(d_tmem + 16, *reinterpret_cast<uint32_t(*)[16]>(&d0[16]));
            tmem_ld16(d_tmem + 64, *reinterpret_cast<uint32_t(*)[16]>(&d1[0]));
            tmem_ld16(d_tmem + 64 + 16, *reinterpret_cast<uint32_t(*)[16]>(&d1[16]));
            tmem_ld_wait();
            tc_fence_before_sync();
            __syncwarp();
            if (lane == 0) mbar_arrive(BAR_ACC_EMPTY(ab));
            if (q < p.slots) {
                const long long row = q / PP;
                const int j = (int)(q - row * PP);
                if (j >= 2 && j < p.P + 2) {
                    float *dst = p.e + ((size_t)row * 64 + 32 * egrp) * p.P + (j - 2);
#pragma unroll
                    for (int h = 0; h < 8; ++h) {
                        const float4 b4 = __ldg(bias4 + h);
                        const float bb[4] = {b4.x, b4.y, b4.z, b4.w};
#pragma unroll
                        for (int i = 0; i < 4; ++i) {
                            const float v = (__uint_as_float(d0[4 * h + i]) + (1.0f / 2048.0f) * __uint_as_float(d1[4 * h + i])) + bb[i];
                            dst[(size_t)(4 * h + i) * p.P] = fmaxf(v, kLeakySlope * v);
                        }
                    }
                }
            }
        };

        for (int c = 0; c < n_local * XT_GROUPS; ++c) {
            const int it = c >> 3, g = c & 7, buf = c & 1;
            // this thread's operand row: slot q of the padded slot space -> (waveform row, pooled position) or zero padding
            const long long q = ((long long)blockIdx.x + (long long)it * gridDim.x) * XT_M + r;
            const float *src = nullptr;
            if (ld_on && q < p.slots) {
                const long long row = q / PP;
                const int j = (int)(q - row * PP);
                if (j >= 2 && j < p.P + 2) src = p.pooled + ((size_t)row * 512 + 64 * g + 32 * half) * p.P + (j - 2);
            }
            float v[32];
#pragma unroll
            for (int i = 0; i < 32; ++i) v[i] = (src != nullptr) ? __ldg(src + (size_t)i * p.P) : 0.0f;
            mbar_wait(BAR_A_EMPTY(buf), (((uint32_t)c >> 1) & 1u) ^ 1u);
            if (ld_on) {
                uint8_t *dst = smem + XT_OFF_A + buf * XT_BUF + (4 * half) * XT_SLAB + r * 16;
#pragma unroll
                for (int h = 0; h < 4; ++h) {
                    uint4 uh, ul;
                    xt_split_pair(v[8 * h + 0], v[8 * h + 1], uh.x, ul.x);
                    xt_split_pair(v[8 * h + 2], v[8 * h + 3], uh.y, ul.y);
                    xt_split_pair(v[8 * h + 4], v[8 * h + 5], uh.z, ul.z);
                    xt_split_pair(v[8 * h + 6], v[8 * h + 7], uh.w, ul.w);
                    *reinterpret_cast<uint4 *>(dst + h * XT_SLAB) = uh;
                    *reinterpret_cast<uint4 *>(dst + XT_PLANE + h * XT_SLAB) = ul;
                }
            }
            fence_proxy_async_smem();
            __syncwarp();
            if (lane == 0) mbar_arrive(BAR_A_FULL(buf));
            // the previous tile's epilogue runs two channel groups into this tile, so the MMAs never wait for it
            if (warp >= 3 && g == 1 && it >= 1) epilogue(it - 1);
        }
        if (warp >= 3 && n_local > 0) epilogue(n_local - 1);
    }

    tc_fence_before_sync();
    __syncthreads();
    if (warp == 1) {
        tc_fence_after_sync();
        tmem_dealloc(tmem0, 256);
    }
}

int launch_expand_tc(const float *pooled, float *e, const void *w_packed, const float *bias, int N, int P, cudaStream_t st) {
    static SmemAttrOnce attr;
    { const int rc = ensure_dyn_smem(expand_tc_kernel, XT_SMEM_BYTES, attr); if (rc) return rc; }
    ExpandParams p;
    p.pooled = pooled; p.e = e; p.w = w_packed; p.bias = bias; p.N = N; p.P = P;
    p.slots = (long long)N * (P + 4);
    const long long tiles = (p.slots + XT_M - 1) / XT_M;
    STOF_REQUIRE(tiles < (1ll << 31), STOF_E_INVALID, "expand_tc: too many tiles");
    p.tiles = (int)tiles;
    const int grid = (int)(tiles < kSMs ? tiles : kSMs);
    expand_tc_kernel<<<grid, XT_THREADS, XT_SMEM_BYTES, st>>>(p);
    STOF_LAUNCH_CHECK();
    return STOF_OK;
}

}  // namespace stof
