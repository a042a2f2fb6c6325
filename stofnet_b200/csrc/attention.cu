// stof_attention_block: AttentionBlock.forward of utils/attention_block.py:13-25 (reference file:line, relative to the
// reference root).  The class is dead code in the reference (only mentioned in comments at models/stofnet.py:90,102, and
// it raises NameError as shipped because `F` is never imported), no checkpoint holds weights for it: parity UNPINNED.
//
//   q, k, v = x Wq^T + bq, x Wk^T + bk, x Wv^T + bv          (B, S, H)
//   out     = dropout(softmax(q k^T / sqrt(H)), p) v         F.scaled_dot_product_attention(q, k, v, dropout_p=0.5)
//
// The intended shapes are tiny (S = 512 channels of the semi-global block, D = H = pooled length), so this is plain fp32
// CUDA-core work: one batched register-tiled GEMM kernel (C = A B^T or A B, fixed k order -> deterministic), and a
// row-softmax kernel with a counter-based dropout mask.
#include <math.h>

#include "common.cuh"
#include "kernels.cuh"

namespace stof {

constexpr int AG_TILE = 64;     // C tile 64 x 64, 256 threads, 4 x 4 per thread
constexpr int AG_K = 16;

// C[b] (M x N) = alpha * A[b] (M x K) * op(B[b]) + bias;  TRANS_B: B is (N x K) row-major (C = A B^T), else (K x N)
template <bool TRANS_B>
__global__ void __launch_bounds__(256)
batched_gemm_kernel(const float *__restrict__ A, const float *__restrict__ Bm, const float *__restrict__ bias, float *__restrict__ C,
                    int M, int N, int K, long long sA, long long sB, long long sC, float alpha) {
    __shared__ float As[AG_K][AG_TILE + 1];
    __shared__ float Bs[AG_K][AG_TILE + 1];
    const int b = blockIdx.z;
    const float *a = A + (size_t)b * sA;
    const float *bm = Bm + (size_t)b * sB;
    float *c = C + (size_t)b * sC;
    const int m0 = blockIdx.y * AG_TILE, n0 = blockIdx.x * AG_TILE;
    const int tx = threadIdx.x & 15, ty = threadIdx.x >> 4;
    float acc[4][4] = {};
    for (int k0 = 0; k0 < K; k0 += AG_K) {
        for (int i = threadIdx.x; i < AG_TILE * AG_K; i += 256) {
            const int r = i / AG_K, kk = i % AG_K;
            const int gm = m0 + r, gk = k0 + kk;
            As[kk][r] = (gm < M && gk < K) ? __ldg(a + (size_t)gm * K + gk) : 0.0f;
        }
        if (TRANS_B) {
            for (int i = threadIdx.x; i < AG_TILE * AG_K; i += 256) {
                const int r = i / AG_K, kk = i % AG_K;
                const int gn = n0 + r, gk = k0 + kk;
                Bs[kk][r] = (gn < N && gk < K) ? __ldg(bm + (size_t)gn * K + gk) : 0.0f;
            }
        } else {
            for (int i = threadIdx.x; i < AG_TILE * AG_K; i += 256) {
                const int kk = i / AG_TILE, r = i % AG_TILE;
                const int gn = n0 + r, gk = k0 + kk;
                Bs[kk][r] = (gn < N && gk < K) ? __ldg(bm + (size_t)gk * N + gn) : 0.0f;
            }
        }
        __syncthreads();
#pragma unroll
        for (int kk = 0; kk < AG_K; ++kk) {
            float av[4], bv[4];
#pragma unroll
            for (int i = 0; i < 4; ++i) { av[i] = As[kk][ty * 4 + i]; bv[i] = Bs[kk][tx * 4 + i]; }
#pragma unroll
            for (int i = 0; i < 4; ++i)
#pragma unroll
                for (int j = 0; j < 4; ++j) acc[i][j] = fmaf(av[i], bv[j], acc[i][j]);
        }
        __syncthreads();
    }
#pragma unroll
    for (int i = 0; i < 4; ++i)
#pragma unroll
        for (int j = 0; j < 4; ++j) {
            const int gm = m0 + ty * 4 + i, gn = n0 + tx * 4 + j;
            if (gm < M && gn < N) c[(size_t)gm * N + gn] = alpha * acc[i][j] + (bias ? __ldg(bias + gn) : 0.0f);
        }
}

__device__ __forceinline__ unsigned long long mix64(unsigned long long z) {     // splitmix64 finaliser
    z += 0x9e3779b97f4a7c15ull;
    z = (z ^ (z >> 30)) * 0xbf58476d1ce4e5b9ull;
    z = (z ^ (z >> 27)) * 0x94d049bb133111ebull;
    return z ^ (z >> 31);
}

// one warp per score row: softmax over S keys, then dropout (keep with probability 1 - p, scale 1 / (1 - p))
__global__ void __launch_bounds__(256)
softmax_dropout_kernel(float *__restrict__ scores, long long rows, int S, float p, unsigned long long seed) {
    const long long row = (long long)blockIdx.x * 8 + (threadIdx.x >> 5);
    const int lane = threadIdx.x & 31;
    if (row >= rows) return;
    float *s = scores + (size_t)row * S;
    float mx = -INFINITY;
    for (int j = lane; j < S; j += 32) mx = fmaxf(mx, s[j]);
#pragma unroll
    for (int o = 16; o; o >>= 1) mx = fmaxf(mx, __shfl_xor_sync(0xffffffffu, mx, o));
    float sum = 0.0f;
    for (int j = lane; j < S; j += 32) { const float e = expf(s[j] - mx); s[j] = e; sum += e; }
#pragma unroll
    for (int o = 16; o; o >>= 1) sum += __shfl_xor_sync(0xffffffffu, sum, o);
    const float inv = 1.0f / sum;
    const float keep_scale = (p > 0.0f) ? 1.0f / (1.0f - p) : 1.0f;
    for (int j = lane; j < S; j += 32) {
        float v = s[j] * inv;
        if (p > 0.0f) {
            const unsigned long long h = mix64(seed ^ mix64((unsigned long long)row * (unsigned long long)S + j));
            const float u = (float)(h >> 40) * (1.0f / 16777216.0f);       // uniform [0, 1)
            v = (u >= p) ? v * keep_scale : 0.0f;
        }
        s[j] = v;
    }
}

static int gemm(bool trans_b, const float *A, const float *B, const float *bias, float *C, int batch, int M, int N, int K,
                long long sA, long long sB, long long sC, float alpha, cudaStream_t st) {
    dim3 grid((N + AG_TILE - 1) / AG_TILE, (M + AG_TILE - 1) / AG_TILE, batch);
    if (trans_b) batched_gemm_kernel<true><<<grid, 256, 0, st>>>(A, B, bias, C, M, N, K, sA, sB, sC, alpha);
    else batched_gemm_kernel<false><<<grid, 256, 0, st>>>(A, B, bias, C, M, N, K, sA, sB, sC, alpha);
    STOF_LAUNCH_CHECK();
    return STOF_OK;
}

}  // namespace stof

using namespace stof;

extern "C" size_t stof_attention_scratch_bytes(int B, int S, int H) {
    if (B <= 0 || S <= 0 || H <= 0) return 0;
    return ((size_t)3 * B * S * H + (size_t)B * S * S) * sizeof(float);
}

extern "C" int stof_attention_block(const float *x_dev, int B, int S, int D, int H, const float *wq_dev, const float *bq_dev,
                                    const float *wk_dev, const float *bk_dev, const float *wv_dev, const float *bv_dev,
                                    float dropout_p, uint64_t seed, float *out_dev, void *scratch_dev, void *stream) {
    STOF_REQUIRE(x_dev && wq_dev && wk_dev && wv_dev && bq_dev && bk_dev && bv_dev && out_dev && scratch_dev, STOF_E_INVALID,
                 "stof_attention_block: NULL buffer");
    STOF_REQUIRE(B > 0 && S > 0 && D > 0 && H > 0 && B <= 65535, STOF_E_INVALID, "stof_attention_block: B=%d S=%d D=%d H=%d", B, S, D, H);
    STOF_REQUIRE(dropout_p >= 0.0f && dropout_p < 1.0f, STOF_E_INVALID, "stof_attention_block: dropout_p=%g not in [0, 1)", dropout_p);
    cudaStream_t st = (cudaStream_t)stream;
    float *q = reinterpret_cast<float *>(scratch_dev);
    float *k = q + (size_t)B * S * H;
    float *v = k + (size_t)B * S * H;
    float *sc = v + (size_t)B * S * H;
    int rc;
    // nn.Linear: y = x W^T + b   (attention_block.py:18-20); all batches as one (B*S, D) x (H, D)^T product
    if ((rc = gemm(true, x_dev, wq_dev, bq_dev, q, 1, B * S, H, D, 0, 0, 0, 1.0f, st))) return rc;
    if ((rc = gemm(true, x_dev, wk_dev, bk_dev, k, 1, B * S, H, D, 0, 0, 0, 1.0f, st))) return rc;
    if ((rc = gemm(true, x_dev, wv_dev, bv_dev, v, 1, B * S, H, D, 0, 0, 0, 1.0f, st))) return rc;
    // scores = q k^T / sqrt(H); softmax over keys; dropout; out = P v   (:23)
    if ((rc = gemm(true, q, k, nullptr, sc, B, S, S, H, (long long)S * H, (long long)S * H, (long long)S * S, 1.0f / sqrtf((float)H), st)))
        return rc;
    const long long rows = (long long)B * S;
    softmax_dropout_kernel<<<(unsigned)((rows + 7) / 8), 256, 0, st>>>(sc, rows, S, dropout_p, (unsigned long long)seed);
    STOF_LAUNCH_CHECK();
    return gemm(false, sc, v, nullptr, out_dev, B, S, H, S, (long long)S * S, (long long)S * H, (long long)S * H, 1.0f, st);
}
