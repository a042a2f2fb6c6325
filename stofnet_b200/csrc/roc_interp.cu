// stof_find_threshold: find_threshold of utils/threshold.py:6-28, i.e. sklearn.metrics.roc_curve (the reference's
// third-party dependency; requirements.txt pins no version, this restates scikit-learn 1.9's published algorithm:
// sort scores descending, cumulative tp / fp at the last index of every distinct score, drop_intermediate=True, prepend the
// (0, 0) point with threshold +inf) followed by arg-max of sqrt(tpr * (1 - fpr)) on batch element 0.
//
// stof_rf_upscale: the dataset-side rf_scale_factor interpolation of datasets/chirp_dataset.py:10-17 (upscale_1d) and
// :79-91 (iq2rf), scipy.interpolate.interp1d(kind='linear') semantics in fp64.
#include <math.h>

#include "common.cuh"
#include "kernels.cuh"

namespace stof {

// order-preserving float -> uint map (larger float = larger key)
__device__ __forceinline__ unsigned int score_key(float v) {
    const unsigned int b = __float_as_uint(v);
    return (b & 0x80000000u) ? ~b : (b | 0x80000000u);
}
__device__ __forceinline__ float key_score(unsigned int k) {
    return __uint_as_float((k & 0x80000000u) ? (k & 0x7fffffffu) : ~k);
}

// min / max of an array as ordered keys (order-independent, so atomics keep it deterministic); NaNs are skipped
__global__ void __launch_bounds__(256)
minmax_kernel(const float *__restrict__ x, size_t n, unsigned int *__restrict__ mm /*[0]=min key, [1]=max key*/) {
    unsigned int lo = 0xffffffffu, hi = 0u;
    for (size_t i = (size_t)blockIdx.x * 256 + threadIdx.x; i < n; i += (size_t)gridDim.x * 256) {
        const float v = __ldg(x + i);
        if (v != v) continue;
        const unsigned int k = score_key(v);
        lo = min(lo, k); hi = max(hi, k);
    }
#pragma unroll
    for (int o = 16; o; o >>= 1) {
        lo = min(lo, __shfl_xor_sync(0xffffffffu, lo, o));
        hi = max(hi, __shfl_xor_sync(0xffffffffu, hi, o));
    }
    if ((threadIdx.x & 31) == 0) { atomicMin(mm, lo); atomicMax(mm + 1, hi); }
}

// row 0 -> 64-bit sort items: (key of the clamped / normalised score) << 32 | label; padding items are 0 (sort last)
__global__ void __launch_bounds__(256)
roc_items_kernel(const float *__restrict__ pred, const float *__restrict__ truth, int M, int Mp, int norm_opt,
                 const unsigned int *__restrict__ mm_pred, const unsigned int *__restrict__ mm_true,
                 unsigned long long *__restrict__ items) {
    const int i = blockIdx.x * 256 + threadIdx.x;
    if (i >= Mp) return;
    if (i >= M) { items[i] = 0ull; return; }
    float s = __ldg(pred + i);
    if (norm_opt) {
        // threshold.py:12-13: masks_norm -= masks_norm.min(); masks_norm /= masks_norm.max()   (whole batch)
        const float mn = key_score(mm_pred[0]), mx = key_score(mm_pred[1]);
        s = (s - mn) / (mx - mn);
    } else if (s < 0.0f) {
        s = 0.0f;                                   // :15
    }
    const float tmax = key_score(mm_true[1]);
    const float pos = (tmax != 0.0f) ? tmax : 1.0f; // :18 pos_label
    const unsigned long long label = (__ldg(truth + i) == pos) ? 1ull : 0ull;
    items[i] = ((unsigned long long)score_key(s) << 32) | label;
}

// ---- bitonic sort, descending, on 64-bit items (ties between equal scores need no stable order: roc_curve folds them) ----
constexpr int BS_CHUNK = 4096;        // items per block in shared memory (32 KB)
constexpr int BS_THREADS = 512;

__device__ __forceinline__ void bs_cmpswap(unsigned long long &a, unsigned long long &b, bool desc) {
    if ((a < b) == desc) { const unsigned long long t = a; a = b; b = t; }
}

// full bitonic sort of every chunk; direction alternates so that chunks form bitonic sequences of size 2 * BS_CHUNK
__global__ void __launch_bounds__(BS_THREADS)
bitonic_local_sort_kernel(unsigned long long *__restrict__ items, int Mp) {
    __shared__ unsigned long long sh[BS_CHUNK];
    const int base = blockIdx.x * BS_CHUNK;
    const int cnt = min(BS_CHUNK, Mp - base);           // Mp is a power of two: cnt == Mp (< chunk) or BS_CHUNK
    for (int i = threadIdx.x; i < cnt; i += BS_THREADS) sh[i] = items[base + i];
    __syncthreads();
    for (int k = 2; k <= cnt; k <<= 1) {
        for (int j = k >> 1; j > 0; j >>= 1) {
            for (int t = threadIdx.x; t < cnt / 2; t += BS_THREADS) {
                const int lo = 2 * t - (t & (j - 1));
                const int hi = lo + j;
                const bool desc = (((base + lo) & k) == 0);
                bs_cmpswap(sh[lo], sh[hi], desc);
            }
            __syncthreads();
        }
    }
    for (int i = threadIdx.x; i < cnt; i += BS_THREADS) items[base + i] = sh[i];
}

// one global compare-exchange step (j >= BS_CHUNK) of merge stage k
__global__ void __launch_bounds__(256)
bitonic_global_step_kernel(unsigned long long *__restrict__ items, int Mp, int k, int j) {
    const int t = blockIdx.x * 256 + threadIdx.x;
    if (t >= Mp / 2) return;
    const int lo = 2 * t - (t & (j - 1));
    const int hi = lo + j;
    const bool desc = ((lo & k) == 0);
    unsigned long long a = items[lo], b = items[hi];
    if ((a < b) == desc) { items[lo] = b; items[hi] = a; }
}

// the steps j = BS_CHUNK/2 .. 1 of merge stage k, in shared memory
__global__ void __launch_bounds__(BS_THREADS)
bitonic_local_merge_kernel(unsigned long long *__restrict__ items, int Mp, int k) {
    __shared__ unsigned long long sh[BS_CHUNK];
    const int base = blockIdx.x * BS_CHUNK;
    for (int i = threadIdx.x; i < BS_CHUNK; i += BS_THREADS) sh[i] = items[base + i];
    __syncthreads();
    for (int j = BS_CHUNK >> 1; j > 0; j >>= 1) {
        for (int t = threadIdx.x; t < BS_CHUNK / 2; t += BS_THREADS) {
            const int lo = 2 * t - (t & (j - 1));
            const int hi = lo + j;
            const bool desc = (((base + lo) & k) == 0);
            bs_cmpswap(sh[lo], sh[hi], desc);
        }
        __syncthreads();
    }
    for (int i = threadIdx.x; i < BS_CHUNK; i += BS_THREADS) items[base + i] = sh[i];
}

// Sequential ROC scan over the sorted items (one thread; M is a single row: tens of thousands of items).
// Emits a curve point at the last index of every distinct score (fps = 1 + idx - tps), drops points whose second
// differences of fps and tps are both zero (drop_intermediate, never the first / last point, only if > 2 points), and keeps
// the first arg-max of sqrt(tpr * (1 - fpr)) including the prepended (0, 0, +inf) point.  fp64 like numpy.
__global__ void roc_scan_kernel(const unsigned long long *__restrict__ items, int M, double *__restrict__ th_out) {
    if (threadIdx.x != 0 || blockIdx.x != 0) return;
    double P = 0.0;
    long long npts = 0;
    for (int i = 0; i < M; ++i) {
        P += (double)(items[i] & 1ull);
        if (i == M - 1 || (items[i] >> 32) != (items[i + 1] >> 32)) ++npts;
    }
    const double Nn = (double)M - P;
    if (!(P > 0.0) || !(Nn > 0.0)) {       // tpr or fpr is NaN everywhere: np.argmax returns index 0, the +inf threshold
        *th_out = INFINITY;
        return;
    }
    const bool drop = npts > 2;
    double best = 0.0;                      // gmean of the prepended (0, 0) point
    double best_th = INFINITY;
    // sliding window of curve points: prev (emitted or not), cur (pending decision), next
    double f_prev = 0, t_prev = 0, f_cur = 0, t_cur = 0, th_cur = 0;
    long long have = 0;                     // points seen so far
    double tps = 0.0;
    auto consider = [&](double f, double t, double th) {
        const double g = sqrt((t / P) * (1.0 - f / Nn));
        if (g > best) { best = g; best_th = th; }
    };
    for (int i = 0; i < M; ++i) {
        tps += (double)(items[i] & 1ull);
        if (!(i == M - 1 || (items[i] >> 32) != (items[i + 1] >> 32))) continue;
        const double f = 1.0 + (double)i - tps, t = tps;
        const double th = (double)key_score((unsigned int)(items[i] >> 32));
        if (have >= 1) {
            // decide the pending point `cur` now that its successor is known
            bool keep = true;
            if (drop && have >= 2) keep = ((f - 2.0 * f_cur + f_prev) != 0.0) || ((t - 2.0 * t_cur + t_prev) != 0.0);
            if (keep) consider(f_cur, t_cur, th_cur);
            f_prev = f_cur; t_prev = t_cur;
        }
        f_cur = f; t_cur = t; th_cur = th;
        ++have;
    }
    consider(f_cur, t_cur, th_cur);         // the last point is always kept
    *th_out = best_th;
}

// ------------------------------------------------------------------------------------------------
// rf_scale_factor pre-interpolation.  Grids: x_j = linspace(0, n/fs, n), t_i = linspace(0, n/fs, m), m = int(n * rf)
// (numpy: arange * step, last element = stop).  interp1d: idx = searchsorted(x, t) clipped to [1, n-1];
// y = y_lo + (y_hi - y_lo) / (x_hi - x_lo) * (t - x_lo).
// ------------------------------------------------------------------------------------------------
__global__ void __launch_bounds__(256)
rf_upscale_kernel(const float *__restrict__ in_re, const float *__restrict__ in_im, int N, int n, int m, double T, double fc,
                  float *__restrict__ out) {
    const int cpr = (m + 255) / 256;                       // chunks per row
    const int row = blockIdx.x / cpr;
    const int i = (blockIdx.x - row * cpr) * 256 + threadIdx.x;
    if (i >= m) return;
    const double dx = T / (double)(n - 1);
    const double dt = (m > 1) ? T / (double)(m - 1) : 0.0;
    const double t = (i == m - 1 && m > 1) ? T : (double)i * dt;
    auto xg = [&](int j) { return (j == n - 1) ? T : (double)j * dx; };
    int idx = (int)floor(t / dx);
    idx = max(0, min(n - 1, idx));
    while (idx > 0 && xg(idx - 1) >= t) --idx;          // searchsorted side='left': first j with x_j >= t
    while (idx < n - 1 && xg(idx) < t) ++idx;
    if (xg(idx) < t) idx = n;
    idx = max(1, min(n - 1, idx));
    const double x_lo = xg(idx - 1), x_hi = xg(idx);
    const float *re = in_re + (size_t)row * n;
    const double y_lo = (double)__ldg(re + idx - 1), y_hi = (double)__ldg(re + idx);
    const double yr = (y_hi - y_lo) / (x_hi - x_lo) * (t - x_lo) + y_lo;
    double v = yr;
    if (in_im != nullptr || fc != 0.0) {
        double yi = 0.0;
        if (in_im != nullptr) {
            const float *im = in_im + (size_t)row * n;
            const double i_lo = (double)__ldg(im + idx - 1), i_hi = (double)__ldg(im + idx);
            yi = (i_hi - i_lo) / (x_hi - x_lo) * (t - x_lo) + i_lo;
        }
        // Re((yr + j yi) * exp(2j pi fc t))     (chirp_dataset.py:88-90)
        const double ph = ((2.0 * M_PI) * fc) * t;
        double s, c;
        sincos(ph, &s, &c);
        v = yr * c - yi * s;
    }
    out[(size_t)row * m + i] = (float)v;
}

static int next_pow2(int v) { int p = 1; while (p < v) p <<= 1; return p; }

}  // namespace stof

using namespace stof;

extern "C" size_t stof_threshold_scratch_bytes(int N, int M) {
    (void)N;
    if (M <= 0) return 0;
    return (size_t)next_pow2(M) * sizeof(unsigned long long) + 256;
}

extern "C" int stof_find_threshold(const float *pred_dev, const float *true_dev, int N, int M, int norm_opt, double *th_out_dev,
                                   void *scratch_dev, void *stream) {
    STOF_REQUIRE(pred_dev && true_dev && th_out_dev && scratch_dev && N > 0 && M > 0, STOF_E_INVALID,
                 "stof_find_threshold: bad arguments (N=%d M=%d)", N, M);
    STOF_REQUIRE(M <= (1 << 28), STOF_E_INVALID, "stof_find_threshold: row too long (%d)", M);
    cudaStream_t st = (cudaStream_t)stream;
    unsigned int *mm = reinterpret_cast<unsigned int *>(scratch_dev);         // [0,1] pred min/max keys, [2,3] truth
    unsigned long long *items = reinterpret_cast<unsigned long long *>(reinterpret_cast<char *>(scratch_dev) + 256);
    const unsigned int init[4] = {0xffffffffu, 0u, 0xffffffffu, 0u};
    STOF_CUDA(cudaMemcpyAsync(mm, init, sizeof(init), cudaMemcpyHostToDevice, st));
    const size_t total = (size_t)N * M;
    const int rgrid = (int)((total + 255) / 256 < (size_t)kSMs * 8 ? (total + 255) / 256 : (size_t)kSMs * 8);
    if (norm_opt) { minmax_kernel<<<rgrid, 256, 0, st>>>(pred_dev, total, mm); STOF_LAUNCH_CHECK(); }
    minmax_kernel<<<rgrid, 256, 0, st>>>(true_dev, total, mm + 2);
    STOF_LAUNCH_CHECK();
    const int Mp = next_pow2(M);
    roc_items_kernel<<<(Mp + 255) / 256, 256, 0, st>>>(pred_dev, true_dev, M, Mp, norm_opt, mm, mm + 2, items);
    STOF_LAUNCH_CHECK();
    bitonic_local_sort_kernel<<<(Mp + BS_CHUNK - 1) / BS_CHUNK, BS_THREADS, 0, st>>>(items, Mp);
    STOF_LAUNCH_CHECK();
    for (int k = 2 * BS_CHUNK; k <= Mp; k <<= 1) {
        for (int j = k >> 1; j >= BS_CHUNK; j >>= 1) {
            bitonic_global_step_kernel<<<(Mp / 2 + 255) / 256, 256, 0, st>>>(items, Mp, k, j);
            STOF_LAUNCH_CHECK();
        }
        bitonic_local_merge_kernel<<<Mp / BS_CHUNK, BS_THREADS, 0, st>>>(items, Mp, k);
        STOF_LAUNCH_CHECK();
    }
    roc_scan_kernel<<<1, 32, 0, st>>>(items, M, th_out_dev);
    STOF_LAUNCH_CHECK();
    return STOF_OK;
}

extern "C" int stof_rf_upscale(const float *in_re_dev, const float *in_im_dev, int N, int n, double rf, double fs, double fc,
                               float *out_dev, void *stream) {
    STOF_REQUIRE(in_re_dev && out_dev && N > 0 && n >= 2 && rf > 0.0 && fs > 0.0, STOF_E_INVALID,
                 "stof_rf_upscale: bad arguments (N=%d n=%d rf=%g fs=%g)", N, n, rf, fs);
    const double mf = (double)n * rf;
    STOF_REQUIRE(mf >= 1.0 && mf < 2147483647.0, STOF_E_INVALID, "stof_rf_upscale: output length %g out of range", mf);
    const int m = (int)mf;                               // int(data_len * rescale_factor), chirp_dataset.py:14,84
    STOF_REQUIRE((long long)N * ((m + 255) / 256) < (1ll << 31), STOF_E_INVALID, "stof_rf_upscale: batch too large");
    const unsigned grid = (unsigned)N * (unsigned)((m + 255) / 256);
    rf_upscale_kernel<<<grid, 256, 0, (cudaStream_t)stream>>>(in_re_dev, in_im_dev, N, n, m, (double)n / fs, fc, out_dev);
    STOF_LAUNCH_CHECK();
    return STOF_OK;
}
