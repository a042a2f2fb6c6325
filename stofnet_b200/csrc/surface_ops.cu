// Standalone operators of the reference's module surface that the fused hot path does not need, but a caller of the
// reference may use on their own (SURVEY.md 8b "Operator surface").  Citations relative to the reference repository root.
//
//   stof_thresholding         utils/mask2samples.py:14-23    float threshold in place, or keep entries equal to the row max
//   stof_get_amplitudes       utils/mask2samples.py:135-136  gather(frames, round(samples))
//   stof_reduce_echoes_pairs  utils/mask2samples.py:117-132  echo_max largest amplitudes, re-ordered by time of arrival
//   stof_semi_global_block    models/stofnet.py:98-117       SemiGlobalBlock.forward on an (N, 64, L) fp32 tensor
//
// All but the last are HBM-bound byte / index work: one pass, coalesced, warp-shuffle reductions.
#include <math.h>

#include "model.cuh"

namespace stof {

// ------------------------------------------------------------------------------------------------
// thresholding (:14-23).  Float threshold: scores[scores < th] = 0 (NaN < th is false: NaN stays).
// None / 0.0: masked_fill(scores < rowmax, 0) with torch.max's NaN propagation (a NaN anywhere makes the row max NaN,
// `s < NaN` is false everywhere, nothing is filled).  One CTA per row; the row-max mode reads the row twice (second
// time from L2).
// ------------------------------------------------------------------------------------------------
__global__ void __launch_bounds__(256)
thresholding_kernel(const float *__restrict__ in, float *__restrict__ out, int M, int th_is_none, float th) {
    __shared__ float s_mx[8];
    __shared__ int s_nan[8];
    const float *row = in + (size_t)blockIdx.x * M;
    float *orow = out + (size_t)blockIdx.x * M;
    float limit = th;
    bool keep_all = false;
    if (th_is_none) {
        float mx = -INFINITY;
        int nan = 0;
        for (int i = threadIdx.x; i < M; i += 256) {
            const float v = __ldg(row + i);
            nan |= (v != v);
            mx = fmaxf(mx, v);
        }
#pragma unroll
        for (int o = 16; o; o >>= 1) {
            mx = fmaxf(mx, __shfl_xor_sync(0xffffffffu, mx, o));
            nan |= __shfl_xor_sync(0xffffffffu, nan, o);
        }
        if ((threadIdx.x & 31) == 0) { s_mx[threadIdx.x >> 5] = mx; s_nan[threadIdx.x >> 5] = nan; }
        __syncthreads();
        mx = s_mx[0]; nan = s_nan[0];
#pragma unroll
        for (int w = 1; w < 8; ++w) { mx = fmaxf(mx, s_mx[w]); nan |= s_nan[w]; }
        limit = mx;
        keep_all = nan != 0;
    }
    for (int i = threadIdx.x; i < M; i += 256) {
        const float v = row[i];
        orow[i] = (!keep_all && v < limit) ? 0.0f : v;
    }
}

// get_amplitudes (:135-136): out[n][k] = frames[n][round(samples[n][k])], torch.round = half to even
__global__ void __launch_bounds__(256)
get_amplitudes_kernel(const float *__restrict__ frames, const float *__restrict__ samples, int N, int M, int K,
                      float *__restrict__ out, int *__restrict__ oob) {
    const size_t i = (size_t)blockIdx.x * 256 + threadIdx.x;
    if (i >= (size_t)N * K) return;
    const int n = (int)(i / K);
    const long long p = (long long)rintf(__ldg(samples + i));
    if (p < 0 || p >= M) {           // torch.gather raises on an out-of-range index
        atomicExch(oob, 1);
        out[i] = 0.0f;
        return;
    }
    out[i] = __ldg(frames + (size_t)n * M + p);
}

// reduce_echoes (:117-132) on (N, K, C) entries, column 0 = time sample, column 1 = amplitude: keep the echo_max entries
// with the largest amplitude (descending, NaN first like torch.sort, equal amplitudes in input order), then order the kept
// entries by column 0 ascending (NaN last, ties in kept order).  One block per row, rank counting in shared memory.
__device__ __forceinline__ bool desc_before(float a, int ia, float b, int ib) {
    const bool an = isnan(a), bn = isnan(b);
    if (an || bn) return (an && !bn) || (an && bn && ia < ib);
    return a > b || (a == b && ia < ib);
}
__device__ __forceinline__ bool asc_before(float a, int ia, float b, int ib) {
    const bool an = isnan(a), bn = isnan(b);
    if (an || bn) return (!an && bn) || (an && bn && ia < ib);
    return a < b || (a == b && ia < ib);
}

__global__ void __launch_bounds__(128)
reduce_pairs_kernel(const float *__restrict__ in, int K, int C, int echo_max, float *__restrict__ out) {
    extern __shared__ float rp_sm[];
    float *t = rp_sm, *a = rp_sm + K;
    int *rank = reinterpret_cast<int *>(rp_sm + 2 * K);
    const float *src = in + (size_t)blockIdx.x * K * C;
    float *dst = out + (size_t)blockIdx.x * echo_max * C;
    for (int j = threadIdx.x; j < K; j += blockDim.x) { t[j] = __ldg(src + (size_t)j * C); a[j] = __ldg(src + (size_t)j * C + 1); }
    __syncthreads();
    for (int j = threadIdx.x; j < K; j += blockDim.x) {
        int r = 0;
        for (int i = 0; i < K; ++i) r += (i != j && desc_before(a[i], i, a[j], j)) ? 1 : 0;
        rank[j] = r;                                    // position after the amplitude sort
    }
    __syncthreads();
    for (int j = threadIdx.x; j < K; j += blockDim.x) {
        if (rank[j] >= echo_max) continue;
        int slot = 0;                                   // stable ascending sort of the kept entries (in amplitude order)
        for (int i = 0; i < K; ++i) slot += (i != j && rank[i] < echo_max && asc_before(t[i], rank[i], t[j], rank[j])) ? 1 : 0;
        for (int c = 0; c < C; ++c) dst[(size_t)slot * C + c] = __ldg(src + (size_t)j * C + c);
    }
}

// fp32 NCL (N, 64, L) <-> C4L4 [n][c/4][l][4] (the layout of the fp32 FFMA kernels, common.cuh)
__global__ void __launch_bounds__(256)
ncl_to_c4l4_kernel(const float *__restrict__ x, float4 *__restrict__ act, int L) {
    const int n = blockIdx.z, c4 = blockIdx.y;
    const int l = blockIdx.x * 256 + threadIdx.x;
    if (l >= L) return;
    const float *p = x + ((size_t)n * 64 + 4 * c4) * L + l;
    act[((size_t)n * 16 + c4) * L + l] = make_float4(p[0], p[L], p[2 * (size_t)L], p[3 * (size_t)L]);
}
__global__ void __launch_bounds__(256)
c4l4_to_ncl_kernel(const float4 *__restrict__ act, float *__restrict__ x, int L) {
    const int n = blockIdx.z, c4 = blockIdx.y;
    const int l = blockIdx.x * 256 + threadIdx.x;
    if (l >= L) return;
    const float4 v = act[((size_t)n * 16 + c4) * L + l];
    float *p = x + ((size_t)n * 64 + 4 * c4) * L + l;
    p[0] = v.x; p[L] = v.y; p[2 * (size_t)L] = v.z; p[3 * (size_t)L] = v.w;
}

static size_t sgb_align(size_t b) { return (b + 255) & ~(size_t)255; }

}  // namespace stof

extern "C" int stof_thresholding(const float *scores_dev, int N, int M, int th_is_none, float th, float *out_dev, void *stream) {
    STOF_REQUIRE(scores_dev && out_dev && N > 0 && M > 0, STOF_E_INVALID, "stof_thresholding: bad arguments (N=%d M=%d)", N, M);
    stof::thresholding_kernel<<<N, 256, 0, (cudaStream_t)stream>>>(scores_dev, out_dev, M, th_is_none, th);
    STOF_LAUNCH_CHECK();
    return STOF_OK;
}

extern "C" int stof_get_amplitudes(const float *frames_dev, const float *samples_dev, int N, int M, int K, float *out_dev,
                                   int *oob_flag_dev, void *stream) {
    STOF_REQUIRE(frames_dev && samples_dev && out_dev && oob_flag_dev && N > 0 && M > 0 && K > 0, STOF_E_INVALID,
                 "stof_get_amplitudes: bad arguments (N=%d M=%d K=%d)", N, M, K);
    const size_t total = (size_t)N * K;
    stof::get_amplitudes_kernel<<<(unsigned)((total + 255) / 256), 256, 0, (cudaStream_t)stream>>>(frames_dev, samples_dev, N, M, K,
                                                                                                out_dev, oob_flag_dev);
    STOF_LAUNCH_CHECK();
    return STOF_OK;
}

extern "C" int stof_reduce_echoes_pairs(const float *pairs_dev, int N, int K, int C, int echo_max, float *out_dev, void *stream) {
    STOF_REQUIRE(pairs_dev && out_dev && N > 0 && K > 0 && C >= 2 && echo_max > 0 && echo_max < K, STOF_E_INVALID,
                 "stof_reduce_echoes_pairs: bad arguments (N=%d K=%d C=%d echo_max=%d; needs 0 < echo_max < K, C >= 2)", N, K, C, echo_max);
    STOF_REQUIRE(K <= 4096, STOF_E_UNSUPPORTED, "stof_reduce_echoes_pairs: at most 4096 entries per row (got %d)", K);
    stof::reduce_pairs_kernel<<<N, 128, (size_t)K * 12, (cudaStream_t)stream>>>(pairs_dev, K, C, echo_max, out_dev);
    STOF_LAUNCH_CHECK();
    return STOF_OK;
}

extern "C" int stof_sgb_workspace_bytes(const stof_model *m, int N, int L, size_t *bytes) {
    STOF_REQUIRE(m && bytes && N > 0 && L > 0, STOF_E_INVALID, "stof_sgb_workspace_bytes: bad arguments");
    STOF_REQUIRE(m->kind == STOF_KIND_STOFNET && m->sgb_scale != 1, STOF_E_INVALID, "stof_sgb_workspace_bytes: model has no semi-global block");
    const int P = L / m->sgb_scale;
    *bytes = stof::sgb_align((size_t)N * 64 * L * 4) + stof::sgb_align((size_t)N * 512 * (P > 0 ? P : 1) * 4) +
             stof::sgb_align((size_t)N * 64 * (P > 0 ? P : 1) * 4);
    return STOF_OK;
}

extern "C" int stof_semi_global_block(stof_model *m, const float *x_dev, int N, int L, float *out_dev, void *ws_dev, size_t ws_bytes,
                                      void *stream) {
    using namespace stof;
    STOF_REQUIRE(m && m->kind == STOF_KIND_STOFNET && m->sgb_scale != 1, STOF_E_INVALID,
                 "stof_semi_global_block: needs a StofNet model with a semi-global block");
    STOF_REQUIRE(x_dev && out_dev && ws_dev && N > 0 && L > 0, STOF_E_INVALID, "stof_semi_global_block: bad arguments");
    STOF_REQUIRE(L >= m->sgb_scale, STOF_E_LENGTH,
                 "semi-global block: L=%d shorter than the pooling window %d (reference: max_pool1d output size 0)", L, m->sgb_scale);
    STOF_REQUIRE(((L % m->sgb_scale) & 1) == 0, STOF_E_LENGTH_ODD,
                 "semi-global block: (L mod %d) = %d is odd; the reference pads (p//2, p//2) and fails at torch.add", m->sgb_scale,
                 L % m->sgb_scale);
    size_t need = 0;
    int rc = stof_sgb_workspace_bytes(m, N, L, &need);
    if (rc) return rc;
    STOF_REQUIRE(ws_bytes >= need, STOF_E_WORKSPACE, "stof_semi_global_block: workspace %zu < %zu bytes", ws_bytes, need);
    STOF_REQUIRE((reinterpret_cast<uintptr_t>(ws_dev) & 255) == 0, STOF_E_INVALID, "stof_semi_global_block: workspace must be 256-byte aligned");
    cudaStream_t st = (cudaStream_t)stream;
    const int P = L / m->sgb_scale;
    char *ws = reinterpret_cast<char *>(ws_dev);
    float *act = reinterpret_cast<float *>(ws);
    float *pooled = reinterpret_cast<float *>(ws + sgb_align((size_t)N * 64 * L * 4));
    float *e = reinterpret_cast<float *>(ws + sgb_align((size_t)N * 64 * L * 4) + sgb_align((size_t)N * 512 * P * 4));
    dim3 grid((L + 255) / 256, 16, N);
    ncl_to_c4l4_kernel<<<grid, 256, 0, st>>>(x_dev, reinterpret_cast<float4 *>(act), L);
    STOF_LAUNCH_CHECK();
    // models/stofnet.py:100-103 contract conv + LeakyReLU + MaxPool1d fused; :106-107 expand conv + LeakyReLU;
    // :108-115 nearest upsample, centred pad, skip add
    if ((rc = launch_contract_pool_simt(act, pooled, m->contract.w, m->contract.b, N, L, P, st))) return rc;
    if ((rc = launch_expand(pooled, e, m->expand.w, m->expand.b, N, P, st))) return rc;
    if ((rc = launch_sgb_add(act, e, N, L, P, m->sgb_scale, (L - P * m->sgb_scale) / 2, st))) return rc;
    c4l4_to_ncl_kernel<<<grid, 256, 0, st>>>(reinterpret_cast<const float4 *>(act), out_dev, L);
    STOF_LAUNCH_CHECK();
    return STOF_OK;
}
