// fp32 CUDA-core kernels of the StofNet forward (models/stofnet.py:42-117 of the reference).
//
// These cover the stages that are NOT dense 64x64 contractions (conv1 with C_in = 1, the pooled
// expand conv, the nearest-upsample/pad/add, conv_last + sub-pixel shuffle) and provide an exact
// fp32 FFMA implementation of the dense convs (conv2..conv12, contract conv + fused max-pool) that
// the tensor-core path (conv_tc.cu) is validated against on the device.
//
// Activation layout between kernels: C4L4 = act[n][c/4][l][c%4] (see common.cuh).
#include "common.cuh"
#include "kernels.cuh"

namespace stof {

// ------------------------------------------------------------------------------------------------
// conv1: Conv1d(1 -> 64, k=9, pad 4) + ReLU   (models/stofnet.py:23,45)
// HBM-bound: reads 4 B, writes 256 B per position.  One thread per position, 16 float4 stores that
// are contiguous across the warp (512 B per warp per store).
// ------------------------------------------------------------------------------------------------
constexpr int C1_THREADS = 256;

template <int K, int ACT>   // ACT 0: relu, 1: tanh
__global__ void __launch_bounds__(C1_THREADS)
conv_first_kernel(const float *__restrict__ x, int N, int L, const float *__restrict__ w /*[K][64]*/,
                  const float *__restrict__ b, float *__restrict__ out /*C4L4*/) {
    __shared__ float4 ws[K * 16];
    __shared__ float4 bs[16];
    for (int i = threadIdx.x; i < K * 16; i += C1_THREADS) ws[i] = reinterpret_cast<const float4 *>(w)[i];
    if (threadIdx.x < 16) bs[threadIdx.x] = reinterpret_cast<const float4 *>(b)[threadIdx.x];
    __syncthreads();
    const int n = blockIdx.y;
    const int l = blockIdx.x * C1_THREADS + threadIdx.x;
    if (l >= L) return;
    const float *xr = x + (size_t)n * L;
    float xv[K];
#pragma unroll
    for (int t = 0; t < K; ++t) {
        const int g = l + t - K / 2;
        xv[t] = (g >= 0 && g < L) ? __ldg(xr + g) : 0.0f;
    }
    float4 *o = reinterpret_cast<float4 *>(out) + (size_t)n * 16 * L + l;
#pragma unroll 4
    for (int c4 = 0; c4 < 16; ++c4) {
        float4 a = bs[c4];
#pragma unroll
        for (int t = 0; t < K; ++t) {
            const float4 wv = ws[t * 16 + c4];
            a.x = fmaf(wv.x, xv[t], a.x); a.y = fmaf(wv.y, xv[t], a.y);
            a.z = fmaf(wv.z, xv[t], a.z); a.w = fmaf(wv.w, xv[t], a.w);
        }
        if (ACT == 0) { a.x = fmaxf(a.x, 0.f); a.y = fmaxf(a.y, 0.f); a.z = fmaxf(a.z, 0.f); a.w = fmaxf(a.w, 0.f); }
        else { a.x = tanhf(a.x); a.y = tanhf(a.y); a.z = tanhf(a.z); a.w = tanhf(a.w); }
        o[(size_t)c4 * L] = a;
    }
}

// ------------------------------------------------------------------------------------------------
// Dense conv 64 -> 64 * NT (k = K) as register-tiled FFMA: each CTA owns 256 (or 240) positions,
// each warp 8 output channels, each lane 8 consecutive positions (64 accumulators per thread).
//   MODE_MID      conv2..conv12: epilogue bias (+LeakyReLU | +residual), C4L4 store
//   MODE_CONTRACT SemiGlobalBlock.contract_conv + LeakyReLU + MaxPool1d(80,80) fused: the (N,512,L)
//                 tensor is never stored (models/stofnet.py:88-91,100-103)
// ------------------------------------------------------------------------------------------------
constexpr int DC_THREADS = 256;
constexpr int DC_PITCH = 264;     // smem row pitch of the input tile (>= 256 + K - 1, multiple of 4)

template <int K>
__device__ __forceinline__ void dc_load_x(float *xs, const float *__restrict__ in, int n, int L, int l0) {
    // xs[ci][j] <- in[n][ci][l0 - K/2 + j], zero outside the row (per-layer zero padding)
    constexpr int W = 256 + K - 1;
    const float4 *src = reinterpret_cast<const float4 *>(in) + (size_t)n * 16 * L;
    for (int idx = threadIdx.x; idx < 16 * W; idx += DC_THREADS) {
        const int c4 = idx / W, j = idx - c4 * W;
        const int l = l0 - K / 2 + j;
        float4 v = make_float4(0.f, 0.f, 0.f, 0.f);
        if (l >= 0 && l < L) v = __ldg(src + (size_t)c4 * L + l);
        float *d = xs + (4 * c4) * DC_PITCH + j;
        d[0] = v.x; d[DC_PITCH] = v.y; d[2 * DC_PITCH] = v.z; d[3 * DC_PITCH] = v.w;
    }
}

template <int K>
__device__ __forceinline__ void dc_compute(const float *xs, const float *ws, float (&acc)[8][8]) {
    const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
#pragma unroll
    for (int c = 0; c < 8; ++c)
#pragma unroll
        for (int j = 0; j < 8; ++j) acc[c][j] = 0.0f;
#pragma unroll 1
    for (int ci = 0; ci < 64; ++ci) {
        float xv[16];
        const float4 *xp = reinterpret_cast<const float4 *>(xs + ci * DC_PITCH + 8 * lane);
#pragma unroll
        for (int q = 0; q < 4; ++q) {
            const float4 t = xp[q];
            xv[4 * q] = t.x; xv[4 * q + 1] = t.y; xv[4 * q + 2] = t.z; xv[4 * q + 3] = t.w;
        }
#pragma unroll
        for (int t = 0; t < K; ++t) {
            const float4 *wp = reinterpret_cast<const float4 *>(ws + (ci * K + t) * 64 + 8 * warp);
            const float4 w0 = wp[0], w1 = wp[1];
            const float wv[8] = {w0.x, w0.y, w0.z, w0.w, w1.x, w1.y, w1.z, w1.w};
#pragma unroll
            for (int c = 0; c < 8; ++c)
#pragma unroll
                for (int j = 0; j < 8; ++j) acc[c][j] = fmaf(wv[c], xv[j + t], acc[c][j]);
        }
    }
}

// EPI: 0 = LeakyReLU, 1 = + residual (no activation), 2 = plain (bias only; the dgrad convs of train.cu)
template <int K, int EPI>
__global__ void __launch_bounds__(DC_THREADS, 1)
conv_mid_simt_kernel(const float *__restrict__ in, const float *res, float *out,   // res may alias out (in-place skip add)
                     const float *__restrict__ w /*[64][K][64]*/, const float *__restrict__ b,
                     int N, int L, int tiles_per_row) {
    extern __shared__ __align__(16) float dc_smem[];
    float *xs = dc_smem;                          // [64][DC_PITCH]  (reused as the output stage)
    float *ws = dc_smem + 64 * DC_PITCH;          // [64][K][64]
    for (int i = threadIdx.x; i < 64 * K * 16; i += DC_THREADS)
        reinterpret_cast<float4 *>(ws)[i] = __ldg(reinterpret_cast<const float4 *>(w) + i);
    const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
    float bias[8];
#pragma unroll
    for (int c = 0; c < 8; ++c) bias[c] = __ldg(b + 8 * warp + c);
    const int total = N * tiles_per_row;
    for (int tile = blockIdx.x; tile < total; tile += gridDim.x) {
        const int n = tile / tiles_per_row, l0 = (tile - n * tiles_per_row) * 256;
        __syncthreads();
        dc_load_x<K>(xs, in, n, L, l0);
        __syncthreads();
        float acc[8][8];
        dc_compute<K>(xs, ws, acc);
        __syncthreads();
        float4 *stage = reinterpret_cast<float4 *>(xs);   // [16][256] float4
#pragma unroll
        for (int h = 0; h < 2; ++h)
#pragma unroll
            for (int j = 0; j < 8; ++j) {
                float4 v;
                v.x = acc[4 * h + 0][j] + bias[4 * h + 0]; v.y = acc[4 * h + 1][j] + bias[4 * h + 1];
                v.z = acc[4 * h + 2][j] + bias[4 * h + 2]; v.w = acc[4 * h + 3][j] + bias[4 * h + 3];
                if (EPI == 0) { v.x = leaky(v.x); v.y = leaky(v.y); v.z = leaky(v.z); v.w = leaky(v.w); }
                stage[(2 * warp + h) * 256 + 8 * lane + j] = v;
            }
        __syncthreads();
        const size_t rowbase = (size_t)n * 16 * L;
        for (int idx = threadIdx.x; idx < 16 * 256; idx += DC_THREADS) {
            const int c4 = idx >> 8, p = idx & 255;
            const int l = l0 + p;
            if (l < L) {
                float4 v = stage[idx];
                const size_t g = rowbase + (size_t)c4 * L + l;
                if (EPI == 1) {
                    const float4 r = reinterpret_cast<const float4 *>(res)[g];
                    v.x += r.x; v.y += r.y; v.z += r.z; v.w += r.w;
                }
                reinterpret_cast<float4 *>(out)[g] = v;
            }
        }
    }
}

// contract conv (64 -> 512, k5) + LeakyReLU + MaxPool1d(80,80): tile = 3 pooling windows = 240 positions
constexpr int CT_WIN = 80;

template <int K>
__global__ void __launch_bounds__(DC_THREADS, 1)
contract_pool_simt_kernel(const float *__restrict__ in, float *__restrict__ pooled /*[N][512][P]*/,
                          const float *__restrict__ w /*[8][64][K][64]*/, const float *__restrict__ b /*[512]*/,
                          int N, int L, int P, int tiles_per_row, int *__restrict__ argmax /*[N][512][P] or NULL*/) {
    extern __shared__ __align__(16) float dc_smem[];
    float *xs = dc_smem;
    float *ws = dc_smem + 64 * DC_PITCH;          // [64][K][64] of the current cout tile
    float *pm = ws + 64 * K * 64;                 // [64 co][32 lanes]
    int *pi = reinterpret_cast<int *>(pm + 64 * 32);   // [64 co][32 lanes] position of the lane's maximum (training only)
    const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
    const int total = N * tiles_per_row;
    for (int tile = blockIdx.x; tile < total; tile += gridDim.x) {
        const int n = tile / tiles_per_row, p0 = (tile - n * tiles_per_row) * 3, l0 = p0 * CT_WIN;
        __syncthreads();
        dc_load_x<K>(xs, in, n, L, l0);
        for (int ct = 0; ct < 8; ++ct) {
            __syncthreads();
            for (int i = threadIdx.x; i < 64 * K * 16; i += DC_THREADS)
                reinterpret_cast<float4 *>(ws)[i] = __ldg(reinterpret_cast<const float4 *>(w) + (size_t)ct * 64 * K * 16 + i);
            __syncthreads();
            float acc[8][8];
            dc_compute<K>(xs, ws, acc);
#pragma unroll
            for (int c = 0; c < 8; ++c) {
                const float bc = __ldg(b + ct * 64 + 8 * warp + c);
                float m = leaky(acc[c][0] + bc);
                int mj = 0;
#pragma unroll
                for (int j = 1; j < 8; ++j) {
                    const float v = leaky(acc[c][j] + bc);
                    if (v > m) { m = v; mj = j; }             // first maximum wins, like torch's max_pool1d
                }
                pm[(8 * warp + c) * 32 + lane] = m;
                if (argmax != nullptr) pi[(8 * warp + c) * 32 + lane] = l0 + 8 * lane + mj;
            }
            __syncthreads();
            if (threadIdx.x < 192) {
                const int co = threadIdx.x & 63, wdw = threadIdx.x >> 6;
                if (p0 + wdw < P) {
                    float m = pm[co * 32 + 10 * wdw];
                    int mi = 0;
#pragma unroll
                    for (int i = 1; i < 10; ++i) {
                        const float v = pm[co * 32 + 10 * wdw + i];
                        if (v > m) { m = v; mi = i; }
                    }
                    pooled[((size_t)n * 512 + ct * 64 + co) * P + p0 + wdw] = m;
                    if (argmax != nullptr) argmax[((size_t)n * 512 + ct * 64 + co) * P + p0 + wdw] = pi[co * 32 + 10 * wdw + mi];
                }
            }
        }
    }
}

// ------------------------------------------------------------------------------------------------
// expand conv: Conv1d(512 -> 64, k5, pad 2) + LeakyReLU on the pooled tensor (N,512,P)
// (models/stofnet.py:94-95,106-107).  0.4 % of the FLOPs, on CUDA cores as a register-tiled fp32 GEMM:
// the rows of a chunk are laid end to end in "slots", each row padded by 2 zero slots on both sides (the conv's zero
// padding), so the 5 taps are plain slot shifts and a 128-slot CTA tile may span several short rows.  Thread tile = 8 slots
// x 4 output channels; the 512 input channels stream through shared memory in double-buffered chunks of 8 (register
// prefetch of the next chunk while the current one is multiplied).
// ------------------------------------------------------------------------------------------------
constexpr int EX_THREADS = 256;
constexpr int EX_TS = 128;                 // output slots per CTA
constexpr int EX_KC = 8;                   // input channels per shared-memory chunk
constexpr int EX_AP = 136;                 // pitch of one channel's slot line (EX_TS + 4 used, 16-byte aligned)
constexpr int EX_A_ELEMS = EX_KC * (EX_TS + 4);          // 1056
constexpr int EX_A_PER_THREAD = (EX_A_ELEMS + EX_THREADS - 1) / EX_THREADS;   // 5
constexpr int EX_W_VEC = EX_KC * 5 * 64 / 4;             // 640 float4
constexpr int EX_W_PER_THREAD = (EX_W_VEC + EX_THREADS - 1) / EX_THREADS;     // 3

__global__ void __launch_bounds__(EX_THREADS, 3)
expand_simt_kernel(const float *__restrict__ pooled /*[N][512][P]*/, float *__restrict__ e /*[N][64][P]*/,
                   const float *__restrict__ w /*[512][5][64]*/, const float *__restrict__ b, int N, int P) {
    __shared__ __align__(16) float As[2][EX_KC][EX_AP];
    __shared__ __align__(16) float Ws[2][EX_KC * 5 * 64];
    const int PP = P + 4;
    const long long S = (long long)N * PP;                     // slots in this launch
    const long long q0 = (long long)blockIdx.x * EX_TS;
    const int tid = threadIdx.x;
    const int sg = tid >> 4, cg = tid & 15;                    // slots [8 sg, +8), channels [4 cg, +4)

    // fixed (slot, channel-in-chunk) -> global offset of this thread's share of an A chunk (-1: zero padding)
    int a_off[EX_A_PER_THREAD];               // < 2^31: checked by the launcher
    int a_dst[EX_A_PER_THREAD];
#pragma unroll
    for (int i = 0; i < EX_A_PER_THREAD; ++i) {
        const int el = tid + i * EX_THREADS;
        a_off[i] = -1;
        a_dst[i] = -1;
        if (el < EX_A_ELEMS) {
            const int ci = el / (EX_TS + 4), sl = el - ci * (EX_TS + 4);
            a_dst[i] = ci * EX_AP + sl;
            const long long q = q0 - 2 + sl;
            if (q >= 0 && q < S) {
                const long long row = q / PP;
                const int j = (int)(q - row * PP);
                if (j >= 2 && j < P + 2) a_off[i] = (int)((row * 512 + ci) * P + (j - 2));
            }
        }
    }
    float a_reg[EX_A_PER_THREAD];
    float4 w_reg[EX_W_PER_THREAD];
    auto fetch = [&](int ci0) {
#pragma unroll
        for (int i = 0; i < EX_A_PER_THREAD; ++i)
            a_reg[i] = (a_off[i] >= 0) ? __ldg(pooled + a_off[i] + ci0 * P) : 0.0f;
        const float4 *wsrc = reinterpret_cast<const float4 *>(w + (size_t)ci0 * 320);
#pragma unroll
        for (int i = 0; i < EX_W_PER_THREAD; ++i) {
            const int v = tid + i * EX_THREADS;
            if (v < EX_W_VEC) w_reg[i] = __ldg(wsrc + v);
        }
    };
    auto stage = [&](int buf) {
        float *ad = &As[buf][0][0];
#pragma unroll
        for (int i = 0; i < EX_A_PER_THREAD; ++i)
            if (a_dst[i] >= 0) ad[a_dst[i]] = a_reg[i];
        float4 *wd = reinterpret_cast<float4 *>(&Ws[buf][0]);
#pragma unroll
        for (int i = 0; i < EX_W_PER_THREAD; ++i) {
            const int v = tid + i * EX_THREADS;
            if (v < EX_W_VEC) wd[v] = w_reg[i];
        }
    };

    float acc[8][4];
    {
        const float4 bc = __ldg(reinterpret_cast<const float4 *>(b) + cg);
#pragma unroll
        for (int j = 0; j < 8; ++j) { acc[j][0] = bc.x; acc[j][1] = bc.y; acc[j][2] = bc.z; acc[j][3] = bc.w; }
    }
    fetch(0);
    stage(0);
    __syncthreads();
    for (int c = 0; c < 512 / EX_KC; ++c) {
        const int buf = c & 1;
        if (c + 1 < 512 / EX_KC) fetch((c + 1) * EX_KC);
#pragma unroll
        for (int ci = 0; ci < EX_KC; ++ci) {
            float xv[12];
            const float4 *xp = reinterpret_cast<const float4 *>(&As[buf][ci][8 * sg]);
#pragma unroll
            for (int q = 0; q < 3; ++q) {
                const float4 t = xp[q];
                xv[4 * q] = t.x; xv[4 * q + 1] = t.y; xv[4 * q + 2] = t.z; xv[4 * q + 3] = t.w;
            }
#pragma unroll
            for (int t = 0; t < 5; ++t) {
                const float4 wv = *reinterpret_cast<const float4 *>(&Ws[buf][(ci * 5 + t) * 64 + 4 * cg]);
#pragma unroll
                for (int j = 0; j < 8; ++j) {
                    acc[j][0] = fmaf(wv.x, xv[j + t], acc[j][0]);
                    acc[j][1] = fmaf(wv.y, xv[j + t], acc[j][1]);
                    acc[j][2] = fmaf(wv.z, xv[j + t], acc[j][2]);
                    acc[j][3] = fmaf(wv.w, xv[j + t], acc[j][3]);
                }
            }
        }
        if (c + 1 < 512 / EX_KC) stage(buf ^ 1);
        __syncthreads();
    }
#pragma unroll
    for (int j = 0; j < 8; ++j) {
        const long long q = q0 + 8 * sg + j;
        if (q >= S) continue;
        const long long row = q / PP;
        const int jj = (int)(q - row * PP);
        if (jj < 2 || jj >= P + 2) continue;
        float *dst = e + ((size_t)row * 64 + 4 * cg) * P + (jj - 2);
        dst[0] = leaky(acc[j][0]);
        dst[(size_t)P] = leaky(acc[j][1]);
        dst[2 * (size_t)P] = leaky(acc[j][2]);
        dst[3 * (size_t)P] = leaky(acc[j][3]);
    }
}

// x0[n][c][l] += E[n][c][(l - off) / scale] for off <= l < off + scale*P   (nearest Upsample(x80),
// centred zero pad, skip add; models/stofnet.py:96,108-115)
__global__ void __launch_bounds__(256)
sgb_add_kernel(float *__restrict__ act /*C4L4*/, const float *__restrict__ e /*[N][64][P]*/, int N, int L, int P,
               int scale, int off) {
    const int n = blockIdx.z, c4 = blockIdx.y;
    const int l = blockIdx.x * 256 + threadIdx.x;
    if (l >= L) return;
    const int j = l - off;
    if (j < 0 || j >= scale * P) return;
    const int p = j / scale;
    const float *er = e + ((size_t)n * 64 + 4 * c4) * P + p;
    float4 *a = reinterpret_cast<float4 *>(act) + ((size_t)n * 16 + c4) * L + l;
    float4 v = *a;
    v.x += __ldg(er); v.y += __ldg(er + P); v.z += __ldg(er + 2 * P); v.w += __ldg(er + 3 * P);
    *a = v;
}

// ------------------------------------------------------------------------------------------------
// conv_last (CIN -> r = 4, k3, pad 1) with SampleShuffle1D fused into the store:
//   mask[n][4 l + u] = b[u] + sum_ci sum_t w[u][ci][t] * in[n][ci][l + t - 1]
// (models/stofnet.py:24,65; utils/sample_shuffle.py:20-26 with C_long = 1).  One thread per position,
// one 16-byte store per position (fully coalesced).  ACT: 0 none, 1 sigmoid (ESPCN).
// ------------------------------------------------------------------------------------------------
template <int CIN, int ACT>
__global__ void __launch_bounds__(256)
conv_last_shuffle_kernel(const float *__restrict__ in /*C4L4, CIN channels*/, float *__restrict__ mask,
                         const float *__restrict__ w /*[CIN][3][4]*/, const float *__restrict__ b, int N, int L) {
    __shared__ float4 ws[CIN * 3];
    for (int i = threadIdx.x; i < CIN * 3; i += 256) ws[i] = reinterpret_cast<const float4 *>(w)[i];
    __syncthreads();
    const int n = blockIdx.y;
    const int l = blockIdx.x * 256 + threadIdx.x;
    if (l >= L) return;
    const float4 *src = reinterpret_cast<const float4 *>(in) + (size_t)n * (CIN / 4) * L;
    float4 acc = __ldg(reinterpret_cast<const float4 *>(b));
#pragma unroll 4
    for (int c4 = 0; c4 < CIN / 4; ++c4) {
#pragma unroll
        for (int t = 0; t < 3; ++t) {
            const int g = l + t - 1;
            if (g < 0 || g >= L) continue;
            const float4 xv = __ldg(src + (size_t)c4 * L + g);
            const float xs[4] = {xv.x, xv.y, xv.z, xv.w};
#pragma unroll
            for (int i = 0; i < 4; ++i) {
                const float4 wv = ws[(4 * c4 + i) * 3 + t];
                acc.x = fmaf(wv.x, xs[i], acc.x); acc.y = fmaf(wv.y, xs[i], acc.y);
                acc.z = fmaf(wv.z, xs[i], acc.z); acc.w = fmaf(wv.w, xs[i], acc.w);
            }
        }
    }
    if (ACT == 1) {
        acc.x = 1.0f / (1.0f + expf(-acc.x)); acc.y = 1.0f / (1.0f + expf(-acc.y));
        acc.z = 1.0f / (1.0f + expf(-acc.z)); acc.w = 1.0f / (1.0f + expf(-acc.w));
    }
    reinterpret_cast<float4 *>(mask)[(size_t)n * L + l] = acc;
}

// ESPCN conv2: Conv1d(64 -> 32, k3, pad 1) + tanh (models/espcn_1d.py:13,33); HBM-bound, one thread
// per (position, 8 output channels)
__global__ void __launch_bounds__(256)
espcn_conv2_kernel(const float *__restrict__ in /*C4L4 64ch*/, float *__restrict__ out /*C4L4 32ch*/,
                   const float *__restrict__ w /*[64][3][32]*/, const float *__restrict__ b, int N, int L) {
    __shared__ float4 ws[64 * 3 * 8];
    for (int i = threadIdx.x; i < 64 * 3 * 8; i += 256) ws[i] = reinterpret_cast<const float4 *>(w)[i];
    __syncthreads();
    const int n = blockIdx.y;
    const int l = blockIdx.x * 256 + threadIdx.x;
    if (l >= L) return;
    const float4 *src = reinterpret_cast<const float4 *>(in) + (size_t)n * 16 * L;
    float4 acc[8];
#pragma unroll
    for (int o = 0; o < 8; ++o) acc[o] = __ldg(reinterpret_cast<const float4 *>(b) + o);
#pragma unroll 2
    for (int c4 = 0; c4 < 16; ++c4) {
#pragma unroll
        for (int t = 0; t < 3; ++t) {
            const int g = l + t - 1;
            if (g < 0 || g >= L) continue;
            const float4 xv = __ldg(src + (size_t)c4 * L + g);
            const float xs[4] = {xv.x, xv.y, xv.z, xv.w};
#pragma unroll
            for (int i = 0; i < 4; ++i)
#pragma unroll
                for (int o = 0; o < 8; ++o) {
                    const float4 wv = ws[((4 * c4 + i) * 3 + t) * 8 + o];
                    acc[o].x = fmaf(wv.x, xs[i], acc[o].x); acc[o].y = fmaf(wv.y, xs[i], acc[o].y);
                    acc[o].z = fmaf(wv.z, xs[i], acc[o].z); acc[o].w = fmaf(wv.w, xs[i], acc[o].w);
                }
        }
    }
    float4 *dst = reinterpret_cast<float4 *>(out) + (size_t)n * 8 * L + l;
#pragma unroll
    for (int o = 0; o < 8; ++o) {
        float4 v = acc[o];
        v.x = tanhf(v.x); v.y = tanhf(v.y); v.z = tanhf(v.z); v.w = tanhf(v.w);
        dst[(size_t)o * L] = v;
    }
}

// standalone SampleShuffle1D (utils/sample_shuffle.py:10-28)
__global__ void __launch_bounds__(256)
sample_shuffle_kernel(const float *__restrict__ x, int B, int Cs, int W, int r, float *__restrict__ out) {
    const int Cl = Cs / r;
    const size_t total = (size_t)B * Cl * W * r;
    for (size_t i = (size_t)blockIdx.x * 256 + threadIdx.x; i < total; i += (size_t)gridDim.x * 256) {
        const int u = (int)(i % r);
        const size_t t = i / r;
        const int wv = (int)(t % W);
        const size_t t2 = t / W;
        const int c = (int)(t2 % Cl);
        const int bb = (int)(t2 / Cl);
        out[i] = __ldg(x + ((size_t)bb * Cs + (size_t)u * Cl + c) * W + wv);
    }
}

// ------------------------------------------------------------------------------------------------ launchers
static size_t mid_smem(int K) { return sizeof(float) * (64 * DC_PITCH + 64 * K * 64); }

int launch_conv_first(const float *x, int N, int L, const float *w, const float *b, float *out, int K, int act,
                      cudaStream_t st) {
    dim3 grid((L + C1_THREADS - 1) / C1_THREADS, N);
    if (K == 9 && act == 0) conv_first_kernel<9, 0><<<grid, C1_THREADS, 0, st>>>(x, N, L, w, b, out);
    else if (K == 5 && act == 1) conv_first_kernel<5, 1><<<grid, C1_THREADS, 0, st>>>(x, N, L, w, b, out);
    else { set_error("conv_first: unsupported K=%d act=%d", K, act); return STOF_E_UNSUPPORTED; }
    STOF_LAUNCH_CHECK();
    return STOF_OK;
}

int launch_conv_mid_simt(const float *in, const float *res, float *out, const float *w, const float *b, int N, int L,
                         int epi, cudaStream_t st) {
    static SmemAttrOnce attr0, attr1, attr2;
    const size_t smem = mid_smem(7);
    {
        int rc = ensure_dyn_smem(conv_mid_simt_kernel<7, 0>, (int)smem, attr0);
        if (!rc) rc = ensure_dyn_smem(conv_mid_simt_kernel<7, 1>, (int)smem, attr1);
        if (!rc) rc = ensure_dyn_smem(conv_mid_simt_kernel<7, 2>, (int)smem, attr2);
        if (rc) return rc;
    }
    const int tpr = (L + 255) / 256;
    const long long total = (long long)N * tpr;
    const int grid = (int)(total < kSMs ? total : kSMs);
    if (epi == 0) conv_mid_simt_kernel<7, 0><<<grid, DC_THREADS, smem, st>>>(in, res, out, w, b, N, L, tpr);
    else if (epi == 1) conv_mid_simt_kernel<7, 1><<<grid, DC_THREADS, smem, st>>>(in, res, out, w, b, N, L, tpr);
    else conv_mid_simt_kernel<7, 2><<<grid, DC_THREADS, smem, st>>>(in, res, out, w, b, N, L, tpr);
    STOF_LAUNCH_CHECK();
    return STOF_OK;
}

int launch_contract_pool_simt(const float *in, float *pooled, const float *w, const float *b, int N, int L, int P,
                              cudaStream_t st, int *argmax) {
    static SmemAttrOnce attr;
    const size_t smem = mid_smem(5) + sizeof(float) * 64 * 32 * 2;
    { const int rc = ensure_dyn_smem(contract_pool_simt_kernel<5>, (int)smem, attr); if (rc) return rc; }
    const int tpr = (P + 2) / 3;
    const long long total = (long long)N * tpr;
    const int grid = (int)(total < kSMs ? total : kSMs);
    contract_pool_simt_kernel<5><<<grid, DC_THREADS, smem, st>>>(in, pooled, w, b, N, L, P, tpr, argmax);
    STOF_LAUNCH_CHECK();
    return STOF_OK;
}

int launch_expand(const float *pooled, float *e, const float *w, const float *b, int N, int P, cudaStream_t st) {
    const long long slots = (long long)N * (P + 4);
    const long long grid = (slots + EX_TS - 1) / EX_TS;
    STOF_REQUIRE(grid < (1ll << 31) && (long long)N * 512 * P < (1ll << 31), STOF_E_INVALID,
                 "expand: chunk of %d rows x %d pooled positions too large (use a smaller chunk_rows)", N, P);
    expand_simt_kernel<<<(unsigned)grid, EX_THREADS, 0, st>>>(pooled, e, w, b, N, P);
    STOF_LAUNCH_CHECK();
    return STOF_OK;
}

int launch_sgb_add(float *act, const float *e, int N, int L, int P, int scale, int off, cudaStream_t st) {
    dim3 grid((L + 255) / 256, 16, N);
    sgb_add_kernel<<<grid, 256, 0, st>>>(act, e, N, L, P, scale, off);
    STOF_LAUNCH_CHECK();
    return STOF_OK;
}

int launch_conv_last(const float *in, float *mask, const float *w, const float *b, int N, int L, int cin, int act,
                     cudaStream_t st) {
    dim3 grid((L + 255) / 256, N);
    if (cin == 64 && act == 0) conv_last_shuffle_kernel<64, 0><<<grid, 256, 0, st>>>(in, mask, w, b, N, L);
    else if (cin == 32 && act == 1) conv_last_shuffle_kernel<32, 1><<<grid, 256, 0, st>>>(in, mask, w, b, N, L);
    else { set_error("conv_last: unsupported cin=%d act=%d", cin, act); return STOF_E_UNSUPPORTED; }
    STOF_LAUNCH_CHECK();
    return STOF_OK;
}

int launch_espcn_conv2(const float *in, float *out, const float *w, const float *b, int N, int L, cudaStream_t st) {
    dim3 grid((L + 255) / 256, N);
    espcn_conv2_kernel<<<grid, 256, 0, st>>>(in, out, w, b, N, L);
    STOF_LAUNCH_CHECK();
    return STOF_OK;
}

}  // namespace stof

extern "C" int stof_sample_shuffle(const float *x_dev, int B, int Cs, int W, int r, float *out_dev, void *stream) {
    STOF_REQUIRE(x_dev && out_dev && B > 0 && Cs > 0 && W > 0 && r > 0 && Cs % r == 0, STOF_E_INVALID,
                 "stof_sample_shuffle: bad arguments (B=%d Cs=%d W=%d r=%d)", B, Cs, W, r);
    const size_t total = (size_t)B * Cs * W;
    int grid = (int)((total + 255) / 256);
    if (grid > stof::kSMs * 16) grid = stof::kSMs * 16;
    stof::sample_shuffle_kernel<<<grid, 256, 0, (cudaStream_t)stream>>>(x_dev, B, Cs, W, r, out_dev);
    STOF_LAUNCH_CHECK();
    return STOF_OK;
}
