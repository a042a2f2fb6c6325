// ESPCN_1D forward in ONE kernel (models/espcn_1d.py:8-36; SURVEY.md 8f row 1, BASELINE config 2's wobbly-sponge):
//   conv1 (1 -> 64, k5, pad 2) + tanh  ->  conv2 (64 -> 32, k3, pad 1) + tanh  ->  conv3 (32 -> 4, k3, pad 1)
//   -> SampleShuffle1D(4) -> sigmoid
// Layer by layer the model moves 784 B per position through HBM for 6,848 MACs; fused it reads 4 B and writes 16 B, and
// the 64- and 32-channel activations only ever exist in shared memory.  What is left is fp32 FMA work (conv2 is 90 % of
// it) plus 96 tanh per position, so this is a CUDA-core kernel: K = 64 x 3 with N = 32 is far too thin a contraction to
// pay for operand splitting on the tensor pipe.
//
// Persistent CTAs (3 per SM), each looping over tiles of EF_T = 126 output positions of one row:
//   phase A  x[l0-4, l0+130) -> h1[32][130]  (positions l0-2 .. l0+128)   one (position, 4 channels) item per thread step;
//            32 of the 64 channels at a time (A, B, A, B), conv2's accumulators live in registers across the halves
//   phase B  h1 -> h2[32][128]               (positions l0-1 .. l0+127)   thread = 4 positions x 8 channels, packed FMAs
//   phase C  h2 -> mask[n][4 l .. 4 l + 3]                                 thread = (position, sub-pixel pair), 8-byte store
// Activations outside [0, L) are the next conv's zero padding, so they are stored as exact zeros (not tanh(bias)).
// The accumulation order of every sum is the one of the layer-by-layer kernels in conv_simt.cu (bias first, then
// (channel group of 4, tap, channel) order), so both paths give bit-identical masks (tests/test_gpu_parity.py).
#include "common.cuh"
#include "kernels.cuh"

namespace stof {

constexpr int EF_T = 126;                  // output positions per tile
constexpr int EF_THREADS = 256;
constexpr int EF_X = EF_T + 8;             // staged input samples
constexpr int EF_H1 = EF_T + 4;            // conv1 outputs per tile (130)
constexpr int EF_H1P = 132;                // row pitch of h1 (floats)
constexpr int EF_H2 = EF_T + 2;            // conv2 outputs per tile (128)
constexpr int EF_H2P = 128;

struct EfSmem {
    float h1[32 * EF_H1P];                 // one half (32 channels) of conv1's output at a time
    float h2[32 * EF_H2P];
    float4 w2[64 * 3 * 8];                 // [ci][t][32 oc]
    float4 w1[5 * 16];                     // [t][64 oc]
    float4 w3[32 * 3];                     // [ci][t][4 u]
    float4 b1[16];
    float4 b2[8];
    float4 b3;
    float x[EF_X + 2];
};

__global__ void __launch_bounds__(EF_THREADS, 3)
espcn_fused_kernel(const float *__restrict__ x, float *__restrict__ mask, int N, int L, const float *__restrict__ w1,
                   const float *__restrict__ b1, const float *__restrict__ w2, const float *__restrict__ b2,
                   const float *__restrict__ w3, const float *__restrict__ b3, int tiles_per_row, int total_tiles) {
    extern __shared__ __align__(16) uint8_t ef_raw[];
    EfSmem &s = *reinterpret_cast<EfSmem *>(ef_raw);
    const int tid = threadIdx.x;
    for (int i = tid; i < 64 * 3 * 8; i += EF_THREADS) s.w2[i] = __ldg(reinterpret_cast<const float4 *>(w2) + i);
    for (int i = tid; i < 5 * 16; i += EF_THREADS) s.w1[i] = __ldg(reinterpret_cast<const float4 *>(w1) + i);
    for (int i = tid; i < 32 * 3; i += EF_THREADS) s.w3[i] = __ldg(reinterpret_cast<const float4 *>(w3) + i);
    if (tid < 16) s.b1[tid] = __ldg(reinterpret_cast<const float4 *>(b1) + tid);
    if (tid < 8) s.b2[tid] = __ldg(reinterpret_cast<const float4 *>(b2) + tid);
    if (tid == 0) s.b3 = __ldg(reinterpret_cast<const float4 *>(b3));

    for (int tile = blockIdx.x; tile < total_tiles; tile += gridDim.x) {
        const int n = tile / tiles_per_row;
        const int l0 = (tile - n * tiles_per_row) * EF_T;
        const float *xr = x + (size_t)n * L;
        __syncthreads();                                   // previous tile's phase C is done with h2 / weights are staged
        for (int j = tid; j < EF_X; j += EF_THREADS) {
            const int g = l0 - 4 + j;
            s.x[j] = (g >= 0 && g < L) ? __ldg(xr + g) : 0.0f;
        }
        __syncthreads();

        // ---- phases A + B, in two halves of conv1's 64 output channels so that h1 needs 17 KB instead of 34 KB (3 CTAs per
        //      SM); conv2's accumulators stay in registers across the halves, its sum order (channel ascending) is unchanged
        const int og = (tid >> 5) & 3, pp = tid & 31;      // phase B: warps 0-3, og = warp (weight loads are broadcasts)
        float2 acc[4][4];
#pragma unroll
        for (int q = 0; q < 4; ++q) {
            const float4 ba = s.b2[2 * og], bb = s.b2[2 * og + 1];
            acc[q][0] = make_float2(ba.x, ba.y); acc[q][1] = make_float2(ba.z, ba.w);
            acc[q][2] = make_float2(bb.x, bb.y); acc[q][3] = make_float2(bb.z, bb.w);
        }
#pragma unroll 1
        for (int half = 0; half < 2; ++half) {
            if (half) __syncthreads();                     // phase B of the first half is done with h1
            // ---- phase A: conv1 + tanh (espcn_1d.py:12,32); item = (position, group of 4 output channels) -------------
            for (int it = tid; it < EF_H1 * 8; it += EF_THREADS) {
                const int cl = it / EF_H1, p = it - cl * EF_H1;
                const int c4 = 8 * half + cl;
                const int g = l0 - 2 + p;
                float4 a = s.b1[c4];
#pragma unroll
                for (int t = 0; t < 5; ++t) {
                    const float4 wv = s.w1[t * 16 + c4];
                    const float xv = s.x[p + t];
                    a.x = fmaf(wv.x, xv, a.x); a.y = fmaf(wv.y, xv, a.y);
                    a.z = fmaf(wv.z, xv, a.z); a.w = fmaf(wv.w, xv, a.w);
                }
                const bool inside = (g >= 0) && (g < L);
                s.h1[(4 * cl + 0) * EF_H1P + p] = inside ? tanhf(a.x) : 0.0f;
                s.h1[(4 * cl + 1) * EF_H1P + p] = inside ? tanhf(a.y) : 0.0f;
                s.h1[(4 * cl + 2) * EF_H1P + p] = inside ? tanhf(a.z) : 0.0f;
                s.h1[(4 * cl + 3) * EF_H1P + p] = inside ? tanhf(a.w) : 0.0f;
            }
            __syncthreads();
            // ---- phase B: conv2 (espcn_1d.py:13,33); thread = positions 4 pp .. 4 pp + 3 x channels [8 og, +8) ---------
            // A 4-position register tile gets 8 packed FMAs out of every 16-byte weight wavefront - with 2 positions the
            // shared-memory pipe, not the FMA pipe, bounded this phase.  Warps 4-7 go on to the barrier; the SM's other
            // CTAs fill their issue slots.
            if (tid < 128) {
#pragma unroll 1
                for (int cl = 0; cl < 8; ++cl) {
                    const int c4 = 8 * half + cl;
                    float hv[4][6];                        // [channel of the group][position 4 pp + 0..5]
#pragma unroll
                    for (int i = 0; i < 4; ++i) {
                        const float4 lo = *reinterpret_cast<const float4 *>(&s.h1[(4 * cl + i) * EF_H1P + 4 * pp]);
                        const float2 hi = *reinterpret_cast<const float2 *>(&s.h1[(4 * cl + i) * EF_H1P + 4 * pp + 4]);
                        hv[i][0] = lo.x; hv[i][1] = lo.y; hv[i][2] = lo.z; hv[i][3] = lo.w; hv[i][4] = hi.x; hv[i][5] = hi.y;
                    }
#pragma unroll
                    for (int t = 0; t < 3; ++t) {
#pragma unroll
                        for (int i = 0; i < 4; ++i) {
                            const float4 wa = s.w2[((4 * c4 + i) * 3 + t) * 8 + 2 * og];
                            const float4 wb = s.w2[((4 * c4 + i) * 3 + t) * 8 + 2 * og + 1];
#pragma unroll
                            for (int q = 0; q < 4; ++q) {
                                const float2 xx = make_float2(hv[i][q + t], hv[i][q + t]);
                                acc[q][0] = __ffma2_rn(make_float2(wa.x, wa.y), xx, acc[q][0]);
                                acc[q][1] = __ffma2_rn(make_float2(wa.z, wa.w), xx, acc[q][1]);
                                acc[q][2] = __ffma2_rn(make_float2(wb.x, wb.y), xx, acc[q][2]);
                                acc[q][3] = __ffma2_rn(make_float2(wb.z, wb.w), xx, acc[q][3]);
                            }
                        }
                    }
                }
            }
        }
        if (tid < 128) {                                   // tanh + zero padding outside the row -> h2
#pragma unroll
            for (int k = 0; k < 4; ++k) {
                float4 v0, v1;                             // channel 8 og + 2 k (+1), positions 4 pp .. 4 pp + 3
                v0.x = tanhf(acc[0][k].x); v0.y = tanhf(acc[1][k].x); v0.z = tanhf(acc[2][k].x); v0.w = tanhf(acc[3][k].x);
                v1.x = tanhf(acc[0][k].y); v1.y = tanhf(acc[1][k].y); v1.z = tanhf(acc[2][k].y); v1.w = tanhf(acc[3][k].y);
                const int g = l0 - 1 + 4 * pp;
                if (g < 0 || g >= L) { v0.x = 0.0f; v1.x = 0.0f; }
                if (g + 1 < 0 || g + 1 >= L) { v0.y = 0.0f; v1.y = 0.0f; }
                if (g + 2 < 0 || g + 2 >= L) { v0.z = 0.0f; v1.z = 0.0f; }
                if (g + 3 < 0 || g + 3 >= L) { v0.w = 0.0f; v1.w = 0.0f; }
                *reinterpret_cast<float4 *>(&s.h2[(8 * og + 2 * k) * EF_H2P + 4 * pp]) = v0;
                *reinterpret_cast<float4 *>(&s.h2[(8 * og + 2 * k + 1) * EF_H2P + 4 * pp]) = v1;
            }
        }
        __syncthreads();

        // ---- phase C: conv3 + SampleShuffle1D + sigmoid (espcn_1d.py:14,34-36; utils/sample_shuffle.py:20-26) ----------
        // thread = (position, output pair): the four sub-pixel sums are independent, so splitting them keeps every
        // sum's order and gives all 252 of 256 threads work
        {
            const int p = tid >> 1, half = tid & 1;
            const int l = l0 + p;
            if (p < EF_T && l < L) {
                float2 acc = half ? make_float2(s.b3.z, s.b3.w) : make_float2(s.b3.x, s.b3.y);
#pragma unroll 2
                for (int c4 = 0; c4 < 8; ++c4) {
#pragma unroll
                    for (int t = 0; t < 3; ++t) {
#pragma unroll
                        for (int i = 0; i < 4; ++i) {
                            const float xv = s.h2[(4 * c4 + i) * EF_H2P + p + t];
                            const float2 wv = reinterpret_cast<const float2 *>(&s.w3[(4 * c4 + i) * 3 + t])[half];
                            acc = __ffma2_rn(wv, make_float2(xv, xv), acc);
                        }
                    }
                }
                acc.x = 1.0f / (1.0f + expf(-acc.x));
                acc.y = 1.0f / (1.0f + expf(-acc.y));
                reinterpret_cast<float2 *>(mask)[((size_t)n * L + l) * 2 + half] = acc;
            }
        }
    }
}

int launch_espcn_fused(const float *x, float *mask, int N, int L, const float *w1, const float *b1, const float *w2,
                       const float *b2, const float *w3, const float *b3, cudaStream_t st) {
    static SmemAttrOnce attr;
    { const int rc = ensure_dyn_smem(espcn_fused_kernel, sizeof(EfSmem), attr); if (rc) return rc; }
    const int tiles_per_row = (L + EF_T - 1) / EF_T;
    const long long total = (long long)N * tiles_per_row;
    STOF_REQUIRE(total < (1ll << 31), STOF_E_INVALID, "espcn_fused: too many tiles");
    const int grid = (int)(total < 3 * kSMs ? total : 3 * kSMs);
    espcn_fused_kernel<<<grid, EF_THREADS, sizeof(EfSmem), st>>>(x, mask, N, L, w1, b1, w2, b2, w3, b3, tiles_per_row, (int)total);
    STOF_LAUNCH_CHECK();
    return STOF_OK;
}

}  // namespace stof
