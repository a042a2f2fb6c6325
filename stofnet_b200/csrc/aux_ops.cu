// Operators either side of the inference path, as hand-written kernels behind the C ABI:
//   stof_toa_rmse        utils/metrics.py:9-41        (consumer of the ToF table, `etol`)
//   stof_coords2mask     utils/mask2samples.py:139-148
//   stof_reduce_echoes   utils/mask2samples.py:103-110,117-137   (`echo_max` amplitude selection of mask2coords)
//   stof_mask_loss       main.py:228-232 + utils/gaussian.py:4-7   (loss value and d loss / d pred)
//   stof_find_threshold  utils/threshold.py:6-28      (sklearn.metrics.roc_curve restated on the device)
//   stof_rf_upscale      datasets/chirp_dataset.py:10-17,79-91     (rf_scale_factor pre-interpolation)
// All of them are HBM-bound byte / index work (no contraction): one pass over the data, coalesced accesses, warp-shuffle
// reductions.  File:line citations are relative to the reference repository root.
#include <math.h>

#include "common.cuh"
#include "kernels.cuh"

namespace stof {

__device__ __forceinline__ bool valid_sample(float v) { return v != 0.0f && isfinite(v); }   // metrics.py:6 remove_invalids

// ------------------------------------------------------------------------------------------------
// echo_max selection of mask2coords (utils/mask2samples.py:103-110, reduce_echoes :117-132, get_amplitudes :135-136):
// amplitude of entry j = scores[row][round(position j)]; keep the echo_max largest amplitudes (descending sort, NaN first
// like torch.sort; equal amplitudes in table order), then order the kept entries by position.  One block per row; the
// table has at most a few hundred entries per row, so both orderings are rank counts over the row held in shared memory.
// K <= echo_max: the table is copied and zero padded (:107-109).  Every output is divided by `divisor` (:111).
// ------------------------------------------------------------------------------------------------
__device__ __forceinline__ bool amp_before(float a, int ia, float b, int ib) {     // a sorts before b (descending, stable)
    const bool an = isnan(a), bn = isnan(b);
    if (an || bn) return (an && !bn) || (an && bn && ia < ib);
    return a > b || (a == b && ia < ib);
}

__global__ void __launch_bounds__(128)
reduce_echoes_kernel(const float *__restrict__ scores, int N, int M, const float *__restrict__ coords, int ld, int K, int echo_max,
                     float divisor, float *__restrict__ out) {
    extern __shared__ float re_sm[];
    float *pos = re_sm, *amp = re_sm + K;
    int *keep = reinterpret_cast<int *>(re_sm + 2 * K);
    const int row = blockIdx.x;
    const float *c = coords + (size_t)row * ld;
    float *o = out + (size_t)row * echo_max;
    if (K <= echo_max) {
        for (int j = threadIdx.x; j < echo_max; j += blockDim.x) o[j] = (j < K) ? __ldg(c + j) / divisor : 0.0f;
        return;
    }
    const float *s = scores + (size_t)row * M;
    for (int j = threadIdx.x; j < K; j += blockDim.x) {
        const float p = __ldg(c + j);
        const int i = min(max((int)rintf(p), 0), M - 1);              // torch.round: half to even
        pos[j] = p;
        amp[j] = __ldg(s + i);
    }
    __syncthreads();
    for (int j = threadIdx.x; j < K; j += blockDim.x) {
        const float a = amp[j];
        int rank = 0;
        for (int i = 0; i < K; ++i) rank += (i != j && amp_before(amp[i], i, a, j)) ? 1 : 0;
        keep[j] = rank < echo_max ? 1 : 0;
    }
    __syncthreads();
    for (int j = threadIdx.x; j < K; j += blockDim.x) {
        if (!keep[j]) continue;
        const float p = pos[j];
        int slot = 0;
        for (int i = 0; i < K; ++i) slot += (keep[i] && (pos[i] < p || (pos[i] == p && i < j))) ? 1 : 0;
        o[slot] = p / divisor;
    }
}

// ------------------------------------------------------------------------------------------------
// toa_rmse: one warp per row.  Lane i walks the ground-truth entries i, i+32, ...; every entry scans the row's estimates
// for its nearest neighbour (squared distance, metrics.py:27-28), TP if d2 <= tol (:31-33).
// ------------------------------------------------------------------------------------------------
__global__ void __launch_bounds__(256)
toa_rmse_kernel(const float *__restrict__ gt, int G, const float *__restrict__ es, int E, int N, float tol,
                float *__restrict__ out) {
    const int row = blockIdx.x * 8 + (threadIdx.x >> 5);
    const int lane = threadIdx.x & 31;
    if (row >= N) return;
    const float *g = gt + (size_t)row * G;
    const float *e = es + (size_t)row * E;
    int ne = 0, ng = 0, tp = 0, fn = 0;
    float sum = 0.0f;
    for (int j = lane; j < E; j += 32) ne += valid_sample(__ldg(e + j)) ? 1 : 0;
#pragma unroll
    for (int o = 16; o; o >>= 1) ne += __shfl_xor_sync(0xffffffffu, ne, o);
    for (int i = lane; i < G; i += 32) {
        const float gv = __ldg(g + i);
        if (!valid_sample(gv)) continue;
        ++ng;
        if (ne == 0) continue;
        float best = INFINITY;
        for (int j = 0; j < E; ++j) {
            const float ev = __ldg(e + j);
            if (!valid_sample(ev)) continue;
            const float d = (gv - ev) * (gv - ev);
            best = fminf(best, d);
        }
        if (best <= tol) { sum += best; ++tp; } else ++fn;
    }
#pragma unroll
    for (int o = 16; o; o >>= 1) {
        ng += __shfl_xor_sync(0xffffffffu, ng, o);
        tp += __shfl_xor_sync(0xffffffffu, tp, o);
        fn += __shfl_xor_sync(0xffffffffu, fn, o);
        sum += __shfl_xor_sync(0xffffffffu, sum, o);
    }
    if (lane == 0) {
        float me = 0.f, tpf = 0.f, fpf = 0.f, fnf = 0.f;
        if (ng > 0 && ne > 0) {              // :23-24 rows without a valid gt or estimate stay all zero
            tpf = (float)tp; fnf = (float)fn; fpf = (float)ne - tpf;
            me = sqrtf(sum / (float)tp);     // mean of an empty selection is NaN, like torch.mean (:31)
        }
        float *o = out + (size_t)row * 7;
        o[0] = me;
        o[1] = tpf / (fpf + tpf) * 100.0f;           // :37-39 (0/0 -> NaN as in the reference)
        o[2] = tpf / (fnf + tpf) * 100.0f;
        o[3] = tpf / (fnf + tpf + fpf) * 100.0f;
        o[4] = tpf; o[5] = fpf; o[6] = fnf;
    }
}

// ------------------------------------------------------------------------------------------------
// coords2mask: the zero fill is a memset; one thread per ground-truth entry scatters a 1 (clamping negatives in place,
// mask2samples.py:144), then column 0 is cleared (:146).
// ------------------------------------------------------------------------------------------------
__global__ void __launch_bounds__(256)
coords2mask_scatter_kernel(long long *__restrict__ samples, int N, int G, int M, float *__restrict__ mask) {
    const size_t i = (size_t)blockIdx.x * 256 + threadIdx.x;
    if (i >= (size_t)N * G) return;
    long long s = samples[i];
    if (s < 0) { s = 0; samples[i] = 0; }
    if (s > 0 && s < M) mask[(i / G) * (size_t)M + s] = 1.0f;     // column 0 is cleared anyway
}

// ------------------------------------------------------------------------------------------------
// Loss of main.py:228-232.  The one-hot mask is sparse (G entries per row), so the blurred target is evaluated from
// the row's entry list instead of materialising mask and blur: blur[m] = sum over distinct valid entries s of
// gauss[m - s + k/2].
// ------------------------------------------------------------------------------------------------
constexpr int LOSS_MAX_K = 31;

// entries of one row -> shared list of distinct valid columns (0 < s < M), -1 elsewhere; returns nothing, list[G]
__device__ __forceinline__ void load_entries(const long long *__restrict__ samples, int G, int M, int *list) {
    for (int i = threadIdx.x; i < G; i += blockDim.x) {
        long long s = samples[i];
        if (s < 0) s = 0;
        list[i] = (s > 0 && s < M) ? (int)s : -1;
    }
    __syncthreads();
    // duplicates collapse in the scatter (a one-hot mask): keep the first occurrence.  The first occurrence of a value is
    // never overwritten, so a concurrent reader finds it whether or not a later duplicate has already been cleared.
    volatile int *vl = list;
    for (int i = threadIdx.x; i < G; i += blockDim.x) {
        const int s = vl[i];
        bool dup = false;
        if (s >= 0)
            for (int j = 0; j < i; ++j) dup |= (vl[j] == s);
        if (dup) vl[i] = -1;
    }
    __syncthreads();
}

__device__ __forceinline__ float blur_at(int m, const int *list, int G, const float *gs, int k) {
    float b = 0.0f;
    const int h = k >> 1;
    for (int i = 0; i < G; ++i) {
        const int s = list[i];
        const int d = m - s + h;
        if (s >= 0 && d >= 0 && d < k) b += gs[d];
    }
    return b;
}

// phase 1: max over the batch of the un-normalised blur (main.py:230 `masks_true_blur.max()`); values are >= 0, so the
// float bit pattern orders like an unsigned integer
__global__ void __launch_bounds__(128)
blur_max_kernel(const long long *__restrict__ samples, int N, int G, int M, const float *__restrict__ gauss, int k,
                unsigned int *__restrict__ blur_max_bits) {
    extern __shared__ int sh_list[];
    __shared__ float gs[LOSS_MAX_K];
    const int n = blockIdx.x;
    if (threadIdx.x < k) gs[threadIdx.x] = __ldg(gauss + threadIdx.x);
    load_entries(samples + (size_t)n * G, G, M, sh_list);
    __syncthreads();
    float best = 0.0f;
    const int h = k >> 1;
    for (int idx = threadIdx.x; idx < G * k; idx += blockDim.x) {
        const int s = sh_list[idx / k];
        const int m = s + (idx % k) - h;
        if (s >= 0 && m >= 0 && m < M) best = fmaxf(best, blur_at(m, sh_list, G, gs, k));
    }
#pragma unroll
    for (int o = 16; o; o >>= 1) best = fmaxf(best, __shfl_xor_sync(0xffffffffu, best, o));
    if ((threadIdx.x & 31) == 0 && best > 0.0f) atomicMax(blur_max_bits, __float_as_uint(best));
}

// phase 2: per (row, 4096-column chunk) partial sums of (pred - blur)^2 and |pred| in fp64, and the gradient
constexpr int LOSS_CHUNK = 4096;
__global__ void __launch_bounds__(256)
mask_loss_kernel(const float *__restrict__ pred, const long long *__restrict__ samples, int N, int G, int M,
                 const float *__restrict__ gauss, int k, float amplitude, float lambda,
                 const float *__restrict__ blur_max, double inv_count, float *__restrict__ grad,
                 double *__restrict__ partial /*[blocks][2]*/, int chunks) {
    extern __shared__ int sh_list[];
    __shared__ float gs[LOSS_MAX_K];
    __shared__ double red[2][8];
    const int n = blockIdx.x / chunks;
    if (threadIdx.x < k) gs[threadIdx.x] = __ldg(gauss + threadIdx.x);
    load_entries(samples + (size_t)n * G, G, M, sh_list);
    __syncthreads();
    const float mx = __ldg(blur_max);
    const int m0 = (blockIdx.x - n * chunks) * LOSS_CHUNK;
    const float *pr = pred + (size_t)n * M;
    const float ginv = (float)inv_count;
    double se = 0.0, sa = 0.0;
    for (int m = m0 + threadIdx.x; m < min(M, m0 + LOSS_CHUNK); m += 256) {
        const float p = __ldg(pr + m);
        float b = blur_at(m, sh_list, G, gs, k);
        b = (b / mx) * amplitude;                 // main.py:230-231: blur /= blur.max(); blur *= mask_amplitude
        const float d = p - b;
        se += (double)d * (double)d;
        sa += (double)fabsf(p);
        if (grad != nullptr) {
            const float sg = (p > 0.0f) ? 1.0f : ((p < 0.0f) ? -1.0f : 0.0f);
            grad[(size_t)n * M + m] = 2.0f * d * ginv + lambda * sg * ginv;
        }
    }
#pragma unroll
    for (int o = 16; o; o >>= 1) {
        se += __shfl_xor_sync(0xffffffffu, se, o);
        sa += __shfl_xor_sync(0xffffffffu, sa, o);
    }
    if ((threadIdx.x & 31) == 0) { red[0][threadIdx.x >> 5] = se; red[1][threadIdx.x >> 5] = sa; }
    __syncthreads();
    if (threadIdx.x == 0) {
        double a = 0.0, b = 0.0;
        for (int w = 0; w < 8; ++w) { a += red[0][w]; b += red[1][w]; }
        const size_t blk = blockIdx.x;
        partial[2 * blk] = a;
        partial[2 * blk + 1] = b;
    }
}

// phase 3: deterministic final sum (one block, fixed order)
__global__ void __launch_bounds__(256)
mask_loss_final_kernel(const double *__restrict__ partial, size_t blocks, double inv_count, float lambda,
                       double *__restrict__ loss) {
    __shared__ double red[2][256];
    double a = 0.0, b = 0.0;
    for (size_t i = threadIdx.x; i < blocks; i += 256) { a += partial[2 * i]; b += partial[2 * i + 1]; }
    red[0][threadIdx.x] = a; red[1][threadIdx.x] = b;
    __syncthreads();
    for (int s = 128; s; s >>= 1) {
        if ((int)threadIdx.x < s) { red[0][threadIdx.x] += red[0][threadIdx.x + s]; red[1][threadIdx.x] += red[1][threadIdx.x + s]; }
        __syncthreads();
    }
    if (threadIdx.x == 0) *loss = red[0][0] * inv_count + (double)lambda * (red[1][0] * inv_count);
}

}  // namespace stof

using namespace stof;

extern "C" int stof_toa_rmse(const float *gt_dev, int G, const float *es_dev, int E, int N, float tol, float *out_dev,
                             void *stream) {
    STOF_REQUIRE(gt_dev && es_dev && out_dev && N > 0 && G > 0 && E > 0, STOF_E_INVALID,
                 "stof_toa_rmse: bad arguments (N=%d G=%d E=%d)", N, G, E);
    toa_rmse_kernel<<<(N + 7) / 8, 256, 0, (cudaStream_t)stream>>>(gt_dev, G, es_dev, E, N, tol, out_dev);
    STOF_LAUNCH_CHECK();
    return STOF_OK;
}

extern "C" int stof_reduce_echoes(const float *scores_dev, int N, int M, const float *coords_dev, int ld, int K, int echo_max,
                                  float divisor, float *out_dev, void *stream) {
    STOF_REQUIRE(coords_dev && out_dev && N > 0 && M > 0 && K > 0 && ld >= K && echo_max > 0, STOF_E_INVALID,
                 "stof_reduce_echoes: bad arguments (N=%d M=%d K=%d ld=%d echo_max=%d)", N, M, K, ld, echo_max);
    STOF_REQUIRE(K <= echo_max || scores_dev, STOF_E_INVALID, "stof_reduce_echoes: scores needed to select %d of %d echoes", echo_max, K);
    STOF_REQUIRE(K <= 4096, STOF_E_UNSUPPORTED, "stof_reduce_echoes: at most 4096 table columns (got %d)", K);
    STOF_REQUIRE(divisor != 0.0f, STOF_E_INVALID, "stof_reduce_echoes: divisor 0");
    reduce_echoes_kernel<<<N, 128, (size_t)K * 12, (cudaStream_t)stream>>>(scores_dev, N, M, coords_dev, ld, K, echo_max, divisor, out_dev);
    STOF_LAUNCH_CHECK();
    return STOF_OK;
}

extern "C" int stof_coords2mask(int64_t *samples_dev, int N, int G, int M, float *mask_dev, void *stream) {
    STOF_REQUIRE(samples_dev && mask_dev && N > 0 && G > 0 && M > 0, STOF_E_INVALID,
                 "stof_coords2mask: bad arguments (N=%d G=%d M=%d)", N, G, M);
    cudaStream_t st = (cudaStream_t)stream;
    STOF_CUDA(cudaMemsetAsync(mask_dev, 0, (size_t)N * M * sizeof(float), st));
    const size_t total = (size_t)N * G;
    coords2mask_scatter_kernel<<<(unsigned)((total + 255) / 256), 256, 0, st>>>(reinterpret_cast<long long *>(samples_dev), N, G, M,
                                                                             mask_dev);
    STOF_LAUNCH_CHECK();
    return STOF_OK;
}

extern "C" size_t stof_loss_scratch_bytes(int N, int M) {
    if (N <= 0 || M <= 0) return 0;
    const size_t blocks = (size_t)N * ((M + LOSS_CHUNK - 1) / LOSS_CHUNK);
    return blocks * 2 * sizeof(double) + 256;
}

extern "C" int stof_mask_loss(const float *pred_dev, const int64_t *samples_dev, int N, int G, int M, const float *gauss_dev,
                              int k, float amplitude, float lambda, int mode, float *blur_max_io_dev, double *loss_dev,
                              float *grad_dev, void *scratch_dev, void *stream) {
    STOF_REQUIRE(samples_dev && gauss_dev && blur_max_io_dev && N > 0 && G > 0 && M > 0, STOF_E_INVALID,
                 "stof_mask_loss: bad arguments (N=%d G=%d M=%d)", N, G, M);
    STOF_REQUIRE(k >= 1 && (k & 1) && k <= LOSS_MAX_K, STOF_E_INVALID, "stof_mask_loss: kernel_size %d must be odd and <= %d", k, LOSS_MAX_K);
    STOF_REQUIRE(G <= 8192, STOF_E_INVALID, "stof_mask_loss: at most 8192 ground-truth entries per row (got %d)", G);
    STOF_REQUIRE(mode >= 0 && mode <= 2, STOF_E_INVALID, "stof_mask_loss: mode %d", mode);
    cudaStream_t st = (cudaStream_t)stream;
    const long long *smp = reinterpret_cast<const long long *>(samples_dev);
    if (mode != 2) {
        STOF_CUDA(cudaMemsetAsync(blur_max_io_dev, 0, sizeof(float), st));
        blur_max_kernel<<<N, 128, G * sizeof(int), st>>>(smp, N, G, M, gauss_dev, k, reinterpret_cast<unsigned int *>(blur_max_io_dev));
        STOF_LAUNCH_CHECK();
        if (mode == 1) return STOF_OK;
    }
    STOF_REQUIRE(pred_dev && loss_dev && scratch_dev, STOF_E_INVALID, "stof_mask_loss: NULL buffer");
    const int chunks = (M + LOSS_CHUNK - 1) / LOSS_CHUNK;
    STOF_REQUIRE((long long)N * chunks < (1ll << 31), STOF_E_INVALID, "stof_mask_loss: batch too large");
    const unsigned grid = (unsigned)N * (unsigned)chunks;
    const double inv_count = 1.0 / ((double)N * (double)M);
    double *partial = reinterpret_cast<double *>(scratch_dev);
    mask_loss_kernel<<<grid, 256, G * sizeof(int), st>>>(pred_dev, smp, N, G, M, gauss_dev, k, amplitude, lambda, blur_max_io_dev,
                                                       inv_count, grad_dev, partial, chunks);
    STOF_LAUNCH_CHECK();
    mask_loss_final_kernel<<<1, 256, 0, st>>>(partial, (size_t)grid, inv_count, lambda, loss_dev);
    STOF_LAUNCH_CHECK();
    return STOF_OK;
}

// ---------------------------------------------------------------------------------------------------------------------------
// stof_mask_loss_static: the mask loss with its per-(row, 4096-column chunk) partial sums in a file-scope device array instead
// of a caller-provided scratch buffer, for callers that cannot allocate (code under CUDA graph capture, serving loops,
// harnesses that forbid cudaMalloc).  No counterpart in the reference.
// The array is 256 KB per GPU that loads the library, held until the process exits.  Nothing in it has to start at zero: the
// loss kernel writes every slot [0, N * chunks) that the final fixed-order sum reads, so the previous call's contents are
// simply overwritten.  All calls on a device share it: calls that overlap on the GPU (two streams) race.
constexpr int LOSS_STATIC_MAX_BLOCKS = 16384;                 // N * ceil(M / 4096) of one call
constexpr int LOSS_STATIC_MAX_DEVICES = 64;
static __device__ __align__(256) double g_loss_partial_static[2 * LOSS_STATIC_MAX_BLOCKS];

// In host code `g_loss_partial_static` is only a handle for the runtime, not a device address: resolve it once per device, with
// that device current.  cudaGetSymbolAddress looks the symbol up in the already-loaded module; it allocates nothing.
static int loss_static_scratch(double **out) {
    static std::atomic<double *> ptrs[LOSS_STATIC_MAX_DEVICES];
    int dev = -1;
    STOF_CUDA(cudaGetDevice(&dev));
    STOF_REQUIRE(dev >= 0 && dev < LOSS_STATIC_MAX_DEVICES, STOF_E_UNSUPPORTED,
                 "stof_mask_loss_static: device ordinal %d outside [0, %d)", dev, LOSS_STATIC_MAX_DEVICES);
    double *p = ptrs[dev].load(std::memory_order_acquire);
    if (p == nullptr) {
        void *sym = nullptr;
        STOF_CUDA(cudaGetSymbolAddress(&sym, g_loss_partial_static));
        STOF_REQUIRE(sym != nullptr, STOF_E_CUDA, "stof_mask_loss_static: cudaGetSymbolAddress returned NULL on device %d", dev);
        p = static_cast<double *>(sym);
        ptrs[dev].store(p, std::memory_order_release);       // two first callers resolve the same address: benign
    }
    *out = p;
    return STOF_OK;
}

extern "C" int stof_mask_loss_static(const float *pred_dev, const int64_t *samples_dev, int N, int G, int M,
                                     const float *gauss_dev, int k, float amplitude, float lambda, int mode,
                                     float *blur_max_io_dev, double *loss_dev, float *grad_dev, void *stream) {
    STOF_REQUIRE(N > 0 && M > 0, STOF_E_INVALID, "stof_mask_loss_static: bad arguments (N=%d M=%d)", N, M);
    const long long blocks = (long long)N * ((M + LOSS_CHUNK - 1) / LOSS_CHUNK);
    STOF_REQUIRE(blocks <= LOSS_STATIC_MAX_BLOCKS, STOF_E_UNSUPPORTED,
                 "stof_mask_loss_static: N * ceil(M / %d) = %lld exceeds the static scratch of %d blocks (N=%d M=%d); "
                 "use stof_mask_loss with a caller-provided scratch buffer", LOSS_CHUNK, blocks, LOSS_STATIC_MAX_BLOCKS, N, M);
    double *scratch = nullptr;
    const int rc = loss_static_scratch(&scratch);
    if (rc != STOF_OK) return rc;
    // same launches as stof_mask_loss (a 4-byte cudaMemsetAsync + three kernels): no cudaMalloc / cudaFree anywhere
    return stof_mask_loss(pred_dev, samples_dev, N, G, M, gauss_dev, k, amplitude, lambda, mode, blur_max_io_dev, loss_dev,
                          grad_dev, scratch, stream);
}
