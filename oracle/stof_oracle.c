/*
 * stof_oracle.c - CPU restatement of hahnec/stofnet's batched echo-waveform inference path.
 *
 * TEST INFRASTRUCTURE ONLY.  This file is the checker that tests/, __graft_entry__.smoke()
 * and bench.py's cpu_baseline leg hold the CUDA path against.  Nothing under stofnet_b200/
 * may import, link or execute it; the product has no CPU path.
 *
 * Where the arithmetic lives: the reference is pure Python that calls PyTorch (unpinned in
 * /root/reference/requirements.txt:2; this environment: torch 2.11.0+cu128, CPU = oneDNN).
 * The functions below restate the published semantics of the ATen ops the reference calls,
 * at the reference's own call sites (cited per function, paths relative to /root/reference).
 * Parity pin: tests/test_oracle.py checks every function here against tests/golden/ (.npz),
 * which tests/golden/make_golden.py produced by running the unmodified reference in the
 * build container (the reference ships no tests or golden vectors of its own).
 *
 * Plain C99; accumulation is double (acc64=1, the tie-breaker) or float (acc64=0).
 */
#include <math.h>
#include <stdint.h>
#include <stdlib.h>
#include <string.h>

#define STOF_MAX_CONVS 16

/* ------------------------------------------------------------------ conv1d */
/* nn.Conv1d(Cin, Cout, K, stride 1, padding pad), cross-correlation with zero padding:
 *   y[co][l] = b[co] + sum_ci sum_k w[co][ci][k] * x[ci][l + k - pad]
 * call sites: models/stofnet.py:23,24,31,88,94 ; models/espcn_1d.py:12-14            */
void stof_oracle_conv1d(const float *x, int Cin, int L, const float *w, const float *b,
                        int Cout, int K, int pad, float *y, int acc64)
{
#pragma omp parallel for schedule(static)
    for (int co = 0; co < Cout; ++co) {
        if (acc64) {
            double *acc = (double *)malloc(sizeof(double) * (size_t)L);
            for (int l = 0; l < L; ++l) acc[l] = 0.0;
            for (int ci = 0; ci < Cin; ++ci) {
                const float *xr = x + (size_t)ci * L;
                for (int k = 0; k < K; ++k) {
                    const double wv = (double)w[((size_t)co * Cin + ci) * K + k];
                    const int s = k - pad;
                    const int l0 = s < 0 ? -s : 0;
                    const int l1 = s > 0 ? L - s : L;
                    for (int l = l0; l < l1; ++l) acc[l] += wv * (double)xr[l + s];
                }
            }
            for (int l = 0; l < L; ++l) y[(size_t)co * L + l] = (float)(acc[l] + (double)b[co]);
            free(acc);
        } else {
            float *acc = y + (size_t)co * L;
            for (int l = 0; l < L; ++l) acc[l] = 0.0f;
            for (int ci = 0; ci < Cin; ++ci) {
                const float *xr = x + (size_t)ci * L;
                for (int k = 0; k < K; ++k) {
                    const float wv = w[((size_t)co * Cin + ci) * K + k];
                    const int s = k - pad;
                    const int l0 = s < 0 ? -s : 0;
                    const int l1 = s > 0 ? L - s : L;
                    for (int l = l0; l < l1; ++l) acc[l] += wv * xr[l + s];
                }
            }
            for (int l = 0; l < L; ++l) acc[l] += b[co];
        }
    }
}

static void relu_(float *x, size_t n)
{
    for (size_t i = 0; i < n; ++i) x[i] = x[i] > 0.0f ? x[i] : 0.0f;
}

/* F.leaky_relu / nn.LeakyReLU(), negative_slope 0.01 (models/stofnet.py:56,89,95) */
static void leaky_relu_(float *x, size_t n)
{
    for (size_t i = 0; i < n; ++i) x[i] = x[i] > 0.0f ? x[i] : x[i] * 0.01f;
}

/* ------------------------------------------------------------------ weights */
typedef struct {
    /* all tensors fp32, PyTorch (Cout, Cin, K) layout, as stored in the .pth state_dict */
    const float *conv1_w, *conv1_b;       /* (F,1,9)  models/stofnet.py:23 */
    const float *convs_w[STOF_MAX_CONVS]; /* conv2..conv{num_blocks-1}: (F,F,7) :30-31 */
    const float *convs_b[STOF_MAX_CONVS];
    const float *last_w, *last_b;         /* (r,F,3)  :24 */
    const float *contract_w, *contract_b; /* (8F,F,5) :88 ; NULL when no semi-global block */
    const float *expand_w, *expand_b;     /* (F,8F,5) :94 */
    int num_features;                     /* F = 64 */
    int num_blocks;                       /* 13 */
    int upsample_factor;                  /* r = conv_last.out_channels (4) */
    int k_first, k_mid, k_last;           /* 9, 7, 3 */
    int sgb_scale;                        /* 80, or 1 = no semi-global block (:27) */
    int sgb_feat;                         /* feat_scale*F = 512 (:85) */
    int sgb_k;                            /* 5 */
} stof_oracle_weights;

/* SemiGlobalBlock.forward (models/stofnet.py:98-117): x (F,L) -> x + pad(up(act(expand(pool(act(contract x)))))) */
static int sgb_forward(const stof_oracle_weights *W, float *x, int L, int acc64)
{
    const int F = W->num_features, C = W->sgb_feat, S = W->sgb_scale, K = W->sgb_k;
    const int P = L / S;                    /* nn.MaxPool1d(S, S), floor mode (:91) */
    if (P < 1) return -2;                   /* torch raises: output size too small */
    const int pad_total = L - P * S;        /* max(0, L - up_len) (:111) */
    if (pad_total % 2) return -3;           /* F.pad(p//2, p//2) then add: shape mismatch (:112-115) */
    float *t = (float *)malloc(sizeof(float) * (size_t)C * L);
    float *pool = (float *)malloc(sizeof(float) * (size_t)C * P);
    float *e = (float *)malloc(sizeof(float) * (size_t)F * P);
    stof_oracle_conv1d(x, F, L, W->contract_w, W->contract_b, C, K, K / 2, t, acc64);
    leaky_relu_(t, (size_t)C * L);
    for (int c = 0; c < C; ++c)
        for (int p = 0; p < P; ++p) {
            float m = -INFINITY;
            for (int j = 0; j < S; ++j) {
                float v = t[(size_t)c * L + (size_t)p * S + j];
                if (v > m || v != v) m = v;     /* ATen max-pool propagates NaN */
            }
            pool[(size_t)c * P + p] = m;
        }
    stof_oracle_conv1d(pool, C, P, W->expand_w, W->expand_b, F, K, K / 2, e, acc64);
    leaky_relu_(e, (size_t)F * P);
    /* nn.Upsample(scale_factor=S, 'nearest') (:96) then centred zero pad (:112) then add (:115) */
    const int off = pad_total / 2;
    for (int c = 0; c < F; ++c)
        for (int j = 0; j < P * S; ++j)
            x[(size_t)c * L + off + j] += e[(size_t)c * P + j / S];
    free(t); free(pool); free(e);
    return 0;
}

/* StofNet.forward for one row (models/stofnet.py:42-67); x (L) -> mask (r*L) */
static int forward_row(const stof_oracle_weights *W, const float *x, int L, float *mask, int acc64)
{
    const int F = W->num_features, r = W->upsample_factor, nb = W->num_blocks;
    const size_t n = (size_t)F * L;
    float *a = (float *)malloc(sizeof(float) * n);     /* running x */
    float *t = (float *)malloc(sizeof(float) * n);     /* conv output */
    float *res = (float *)malloc(sizeof(float) * n);   /* short skip  */
    float *res1 = (float *)malloc(sizeof(float) * n);  /* long skip   */
    float *y = (float *)malloc(sizeof(float) * (size_t)r * L);
    int rc = 0;
    stof_oracle_conv1d(x, 1, L, W->conv1_w, W->conv1_b, F, W->k_first, 4, a, acc64); /* :23 pad 4 */
    relu_(a, n);                                                                       /* :45 */
    if (W->contract_w) rc = sgb_forward(W, a, L, acc64);                               /* :48 */
    if (rc == 0) {
        memcpy(res, a, sizeof(float) * n);
        memcpy(res1, a, sizeof(float) * n);                                            /* :51 */
        for (int i = 2; i <= nb - 2; ++i) {                                            /* :52 */
            /* residual_layers = 3,5,..,nb-2 (+ nb-1, nb) (:37) */
            const int is_res = (i >= 3) && (i % 2 == 1);
            stof_oracle_conv1d(a, F, L, W->convs_w[i - 2], W->convs_b[i - 2], F, W->k_mid,
                               W->k_mid / 2, t, acc64);                                /* 'same' */
            if (is_res) {
                for (size_t j = 0; j < n; ++j) a[j] = res[j] + t[j];                   /* :56 */
                memcpy(res, a, sizeof(float) * n);                                     /* :58 */
            } else {
                leaky_relu_(t, n);
                memcpy(a, t, sizeof(float) * n);
            }
        }
        stof_oracle_conv1d(a, F, L, W->convs_w[nb - 3], W->convs_b[nb - 3], F, W->k_mid,
                           W->k_mid / 2, t, acc64);                                    /* conv{nb-1} :61 */
        for (size_t j = 0; j < n; ++j) a[j] = res1[j] + t[j];                          /* :62 */
        stof_oracle_conv1d(a, F, L, W->last_w, W->last_b, r, W->k_last, 1, y, acc64);  /* :24 pad 1 */
        /* SampleShuffle1D with C_long = 1 (utils/sample_shuffle.py:20-26): out[w*r+u] = y[u][w] */
        for (int u = 0; u < r; ++u)
            for (int l = 0; l < L; ++l) mask[(size_t)l * r + u] = y[(size_t)u * L + l];
    }
    free(a); free(t); free(res); free(res1); free(y);
    return rc;
}

/* x (N,L) fp32 -> mask (N, r*L) fp32.  0 ok; -2 L too short; -3 (L mod scale) odd. */
int stof_oracle_forward(const stof_oracle_weights *W, const float *x, int N, int L, float *mask,
                        int acc64)
{
    for (int n = 0; n < N; ++n) {
        int rc = forward_row(W, x + (size_t)n * L, L, mask + (size_t)n * L * W->upsample_factor, acc64);
        if (rc) return rc;
    }
    return 0;
}

/* ESPCN_1D.forward (models/espcn_1d.py:31-36): tanh(conv1) -> tanh(conv2) -> sigmoid(shuffle(conv3)) */
int stof_oracle_espcn_forward(const float *w1, const float *b1, const float *w2, const float *b2,
                              const float *w3, const float *b3, int r, const float *x, int N, int L,
                              float *mask, int acc64)
{
    float *a = (float *)malloc(sizeof(float) * 64 * (size_t)L);
    float *c = (float *)malloc(sizeof(float) * 32 * (size_t)L);
    float *y = (float *)malloc(sizeof(float) * (size_t)r * L);
    for (int n = 0; n < N; ++n) {
        stof_oracle_conv1d(x + (size_t)n * L, 1, L, w1, b1, 64, 5, 2, a, acc64);
        for (size_t j = 0; j < 64 * (size_t)L; ++j) a[j] = tanhf(a[j]);
        stof_oracle_conv1d(a, 64, L, w2, b2, 32, 3, 1, c, acc64);
        for (size_t j = 0; j < 32 * (size_t)L; ++j) c[j] = tanhf(c[j]);
        stof_oracle_conv1d(c, 32, L, w3, b3, r, 3, 1, y, acc64);
        float *m = mask + (size_t)n * L * r;
        for (int u = 0; u < r; ++u)
            for (int l = 0; l < L; ++l) m[(size_t)l * r + u] = 1.0f / (1.0f + expf(-y[(size_t)u * L + l]));
    }
    free(a); free(c); free(y);
    return 0;
}

/* SampleShuffle1D.forward (utils/sample_shuffle.py:10-28): x (B, r*Cl, W) -> (B, Cl, r*W),
 * out[b][c][w*r+u] = x[b][u*Cl+c][w] */
void stof_oracle_sample_shuffle(const float *x, int B, int Cs, int Wd, int r, float *out)
{
    const int Cl = Cs / r;
    for (int b = 0; b < B; ++b)
        for (int c = 0; c < Cl; ++c)
            for (int w = 0; w < Wd; ++w)
                for (int u = 0; u < r; ++u)
                    out[((size_t)b * Cl + c) * Wd * r + (size_t)w * r + u] =
                        x[((size_t)b * Cs + (size_t)u * Cl + c) * Wd + w];
}

/* ------------------------------------------------------------ stage 4: peaks */
/* nms_1d (utils/mask2samples.py:5-11): window -> odd; max_pool1d(win, stride 1, pad (win-1)/2)
 * with implicit -inf padding and NaN propagation; suppressed = (s == max).float() * s       */
void stof_oracle_nms_1d(const float *s, int M, int window, float *out)
{
    const int win = window / 2 * 2 + 1, h = (win - 1) / 2;
    for (int i = 0; i < M; ++i) {
        float m = -INFINITY;
        const int j0 = i - h < 0 ? 0 : i - h, j1 = i + h >= M ? M - 1 : i + h;
        for (int j = j0; j <= j1; ++j) {
            const float v = s[j];
            if (v > m || v != v) m = v;
        }
        out[i] = ((s[i] == m) ? 1.0f : 0.0f) * s[i];
    }
}

/* thresholding (utils/mask2samples.py:14-23) on one row, in place.
 * th_is_none also covers th == 0.0 ("if threshold:" is falsy for both).                     */
static void thresholding_row(float *s, int M, int th_is_none, float th)
{
    if (!th_is_none) {
        for (int i = 0; i < M; ++i) if (s[i] < th) s[i] = 0.0f;
    } else {
        float m = s[0];                       /* torch.max(dim) propagates NaN */
        for (int i = 1; i < M; ++i) if (s[i] > m || s[i] != s[i]) { m = s[i]; if (m != m) break; }
        for (int i = 0; i < M; ++i) if (s[i] < m) s[i] = 0.0f;
    }
}

/* get_maxima_positions (utils/mask2samples.py:26-34) per row: ascending positions of non-zero
 * entries after nms + thresholding.  Returns total K; idx (cap,2) int64 row-major [row, pos];
 * counts[n] = detections in row n (bincount, :92).                                          */
int64_t stof_oracle_maxima(const float *scores, int N, int M, int window, int th_is_none, float th,
                           int64_t *idx, int64_t cap, int32_t *counts)
{
    int64_t K = 0;
    float *t = (float *)malloc(sizeof(float) * (size_t)M);
    for (int n = 0; n < N; ++n) {
        stof_oracle_nms_1d(scores + (size_t)n * M, M, window, t);
        thresholding_row(t, M, th_is_none, th);
        int32_t c = 0;
        for (int i = 0; i < M; ++i)
            if (t[i] != 0.0f) {               /* torch.nonzero: NaN counts as non-zero */
                if (idx && K < cap) { idx[2 * K] = n; idx[2 * K + 1] = i; }
                ++K; ++c;
            }
        if (counts) counts[n] = c;
    }
    free(t);
    return K;
}

/* mask2coords (utils/mask2samples.py:81-114) without echo_max: coords (N, kmax) fp32,
 * ascending positions / upsample_factor, zero padded on the right.  Returns kmax
 * (0 = the reference's "no maxima anywhere" branch, which returns zeros((N,1,1)), :87-88).
 * coords may be NULL to query kmax; stride = row pitch of coords in floats (>= kmax).        */
int stof_oracle_mask2coords(const float *scores, int N, int M, int window, int th_is_none, float th,
                            int upsample_factor, float *coords, int stride, int32_t *counts)
{
    int kmax = 0;
    float *t = (float *)malloc(sizeof(float) * (size_t)M);
    for (int n = 0; n < N; ++n) {
        stof_oracle_nms_1d(scores + (size_t)n * M, M, window, t);
        thresholding_row(t, M, th_is_none, th);
        int c = 0;
        for (int i = 0; i < M; ++i)
            if (t[i] != 0.0f) {
                if (coords && c < stride)
                    coords[(size_t)n * stride + c] = (float)i / (float)upsample_factor; /* :102,112 */
                ++c;
            }
        if (coords) for (int j = c; j < stride; ++j) coords[(size_t)n * stride + j] = 0.0f;
        if (counts) counts[n] = c;
        if (c > kmax) kmax = c;
    }
    free(t);
    return kmax;
}

/* coords2mask (utils/mask2samples.py:139-148): scatter ones at (clamped) sample indices, then
 * zero column 0 (placeholder for absent ground truth).  samples (N,G) int64, mask (N,M).     */
void stof_oracle_coords2mask(const int64_t *samples, int N, int G, int M, float *mask)
{
    memset(mask, 0, sizeof(float) * (size_t)N * M);
    for (int n = 0; n < N; ++n) {
        for (int g = 0; g < G; ++g) {
            int64_t s = samples[(size_t)n * G + g];
            if (s < 0) s = 0;
            if (s < M) mask[(size_t)n * M + s] = 1.0f;
        }
        mask[(size_t)n * M] = 0.0f;
    }
}

/* toa_rmse (utils/metrics.py:9-41): per row, valid = non-zero finite; for each gt the squared
 * distance to the nearest estimate; TP if d2 <= tol.  out (N,7) = rmse, precision, recall,
 * jaccard, tp, fp, fn (percentages; 0/0 -> NaN as in the reference).                         */
void stof_oracle_toa_rmse(const float *gt, int G, const float *es, int E, int N, float tol, float *out)
{
    for (int n = 0; n < N; ++n) {
        float me = 0.0f, tp = 0.0f, fp = 0.0f, fn = 0.0f;
        int ng = 0, ne = 0;
        for (int i = 0; i < G; ++i) { float v = gt[(size_t)n * G + i]; if (v != 0.0f && isfinite(v)) ++ng; }
        for (int j = 0; j < E; ++j) { float v = es[(size_t)n * E + j]; if (v != 0.0f && isfinite(v)) ++ne; }
        if (ng && ne) {
            float sum = 0.0f; int cnt = 0;
            for (int i = 0; i < G; ++i) {
                const float g = gt[(size_t)n * G + i];
                if (!(g != 0.0f && isfinite(g))) continue;
                float best = INFINITY;
                for (int j = 0; j < E; ++j) {
                    const float e = es[(size_t)n * E + j];
                    if (!(e != 0.0f && isfinite(e))) continue;
                    const float d = (g - e) * (g - e);
                    if (d < best) best = d;
                }
                if (best <= tol) { sum += best; ++cnt; tp += 1.0f; } else fn += 1.0f;
            }
            me = sqrtf(sum / (float)cnt);     /* mean of empty -> NaN, as torch.mean */
            fp = (float)ne - tp;
        }
        float *o = out + (size_t)n * 7;
        o[0] = me;
        o[1] = tp / (fp + tp) * 100.0f;
        o[2] = tp / (fn + tp) * 100.0f;
        o[3] = tp / (fn + tp + fp) * 100.0f;
        o[4] = tp; o[5] = fp; o[6] = fn;
    }
}

/* ------------------------------------------------------------------ loss (training / validation step)
 * main.py:228-232 with utils/gaussian.py:4-7 and utils/mask2samples.py:139-148:
 *   true = coords2mask(gt);  blur = conv1d(true, gauss, padding k/2);  blur /= blur.max();  blur *= amp
 *   loss = mean((pred - blur)^2) + lambda * mean(|pred|)
 * Dense restatement (mask and blur are materialised, unlike the CUDA kernel).  grad (N,M) may be NULL:
 * d loss / d pred = (2 (pred - blur) + lambda sign(pred)) / (N M).  Returns the loss; *blur_max_out = blur.max().     */
double stof_oracle_mask_loss(const float *pred, const int64_t *samples, int N, int G, int M, const float *gauss, int k,
                             float amp, float lambda, float *grad, float *blur_max_out)
{
    const size_t tot = (size_t)N * M;
    float *mask = (float *)malloc(sizeof(float) * tot);
    float *blur = (float *)malloc(sizeof(float) * tot);
    stof_oracle_coords2mask(samples, N, G, M, mask);
    const int h = k / 2;
    float mx = -INFINITY;
    for (int n = 0; n < N; ++n)
        for (int m = 0; m < M; ++m) {
            float a = 0.0f;
            for (int t = 0; t < k; ++t) {                   /* cross-correlation, zero padding */
                const int q = m + t - h;
                if (q >= 0 && q < M) a += gauss[t] * mask[(size_t)n * M + q];
            }
            blur[(size_t)n * M + m] = a;
            if (a > mx) mx = a;
        }
    if (blur_max_out) *blur_max_out = mx;
    double se = 0.0, sa = 0.0;
    const float inv = (float)(1.0 / ((double)N * (double)M));
    for (size_t i = 0; i < tot; ++i) {
        const float b = (blur[i] / mx) * amp;
        const float d = pred[i] - b;
        se += (double)d * (double)d;
        sa += fabs((double)pred[i]);
        if (grad) {
            const float sg = pred[i] > 0.0f ? 1.0f : (pred[i] < 0.0f ? -1.0f : 0.0f);
            grad[i] = 2.0f * d * inv + lambda * sg * inv;
        }
    }
    free(mask); free(blur);
    return se / (double)tot + (double)lambda * (sa / (double)tot);
}

/* ------------------------------------------------------------------ find_threshold
 * utils/threshold.py:6-28.  The ROC itself is scikit-learn's (third-party dependency, unpinned in requirements.txt;
 * restated from scikit-learn 1.9 sklearn/metrics/_ranking.py roc_curve + confusion_matrix_at_thresholds): stable sort
 * by descending score, cumulative tp at the last index of each distinct score, fps = 1 + idx - tps, drop_intermediate
 * (second differences of fps and tps both zero), prepend (0,0) at threshold +inf, fpr/tpr = NaN when a class is empty;
 * then threshold.py:24-26: first arg-max of sqrt(tpr (1 - fpr)) with numpy's NaN-is-max rule.                       */
typedef struct { float s; int y; } roc_item;
static int roc_cmp(const void *a, const void *b)
{
    const float x = ((const roc_item *)a)->s, y = ((const roc_item *)b)->s;
    return (x < y) - (x > y);                               /* descending */
}
double stof_oracle_find_threshold(const float *pred, const float *truth, int N, int M, int norm_opt)
{
    const size_t tot = (size_t)N * M;
    float tmax = -INFINITY, pmin = INFINITY, pmax = -INFINITY;
    for (size_t i = 0; i < tot; ++i) {
        if (truth[i] > tmax) tmax = truth[i];
        if (pred[i] < pmin) pmin = pred[i];
        if (pred[i] > pmax) pmax = pred[i];
    }
    const float pos = (tmax != 0.0f) ? tmax : 1.0f;         /* :18 */
    roc_item *it = (roc_item *)malloc(sizeof(roc_item) * (size_t)M);
    for (int i = 0; i < M; ++i) {
        float s = pred[i];
        if (norm_opt) { s = s - pmin; s = s / (pmax - pmin); }     /* :12-13 (max of the shifted tensor = pmax - pmin) */
        else if (s < 0.0f) s = 0.0f;                               /* :15 */
        it[i].s = s;
        it[i].y = (truth[i] == pos);
    }
    qsort(it, (size_t)M, sizeof(roc_item), roc_cmp);
    double *fps = (double *)malloc(sizeof(double) * (size_t)(M + 1));
    double *tps = (double *)malloc(sizeof(double) * (size_t)(M + 1));
    double *thr = (double *)malloc(sizeof(double) * (size_t)(M + 1));
    int np_ = 0;
    double c = 0.0;
    for (int i = 0; i < M; ++i) {
        c += it[i].y;
        if (i == M - 1 || it[i].s != it[i + 1].s) { tps[np_] = c; fps[np_] = 1.0 + i - c; thr[np_] = it[i].s; ++np_; }
    }
    double *f2 = (double *)malloc(sizeof(double) * (size_t)(M + 2));
    double *t2 = (double *)malloc(sizeof(double) * (size_t)(M + 2));
    double *h2 = (double *)malloc(sizeof(double) * (size_t)(M + 2));
    /* drop_intermediate decides on the uncompacted neighbours, then the (0, 0) point is prepended */
    int n2 = 0;
    f2[n2] = 0.0; t2[n2] = 0.0; h2[n2] = INFINITY; ++n2;     /* prepended (0, 0) point */
    for (int p = 0; p < np_; ++p) {
        int keep = 1;
        if (np_ > 2 && p > 0 && p < np_ - 1)
            keep = (fps[p + 1] - 2 * fps[p] + fps[p - 1] != 0.0) || (tps[p + 1] - 2 * tps[p] + tps[p - 1] != 0.0);
        if (keep) { f2[n2] = fps[p]; t2[n2] = tps[p]; h2[n2] = thr[p]; ++n2; }
    }
    const double P = t2[n2 - 1], Nn = f2[n2 - 1];
    double best = -1.0, best_th = INFINITY;
    int have_nan = 0;
    for (int p = 0; p < n2 && !have_nan; ++p) {
        const double tpr = (P <= 0.0) ? NAN : t2[p] / P;
        const double fpr = (Nn <= 0.0) ? NAN : f2[p] / Nn;
        const double g = sqrt(tpr * (1.0 - fpr));
        if (g != g) { best_th = h2[p]; have_nan = 1; break; }     /* np.argmax: the first NaN wins */
        if (g > best) { best = g; best_th = h2[p]; }
    }
    free(it); free(fps); free(tps); free(thr); free(f2); free(t2); free(h2);
    return best_th;
}

/* ------------------------------------------------------------------ rf_scale_factor pre-interpolation
 * datasets/chirp_dataset.py:10-17 (upscale_1d) and :79-91 (iq2rf).  The interpolation is scipy.interpolate.interp1d
 * (kind='linear'; third-party, unpinned; restated from scipy 1.x _interpolate.py _call_linear): idx =
 * searchsorted(x, t) clipped to [1, n-1]; y = slope * (t - x_lo) + y_lo.  Grids are np.linspace(0, n/fs, num,
 * endpoint=True) = arange(num) * step with the last sample set to the stop value.  in_im may be NULL (upscale_1d).     */
void stof_oracle_rf_upscale(const float *in_re, const float *in_im, int N, int n, double rf, double fs, double fc, float *out)
{
    const int m = (int)((double)n * rf);
    const double T = (double)n / fs;
    const double dx = T / (double)(n - 1), dt = (m > 1) ? T / (double)(m - 1) : 0.0;
    for (int r = 0; r < N; ++r)
        for (int i = 0; i < m; ++i) {
            const double t = (i == m - 1 && m > 1) ? T : (double)i * dt;
            int idx = 0;                                       /* searchsorted(x, t, side='left'), linear scan */
            while (idx < n && ((idx == n - 1) ? T : (double)idx * dx) < t) ++idx;
            if (idx < 1) idx = 1;
            if (idx > n - 1) idx = n - 1;
            const double x_lo = (double)(idx - 1) * dx, x_hi = (idx == n - 1) ? T : (double)idx * dx;
            const double a = in_re[(size_t)r * n + idx - 1], b = in_re[(size_t)r * n + idx];
            const double yr = (b - a) / (x_hi - x_lo) * (t - x_lo) + a;
            double v = yr;
            if (in_im || fc != 0.0) {
                double yi = 0.0;
                if (in_im) {
                    const double c = in_im[(size_t)r * n + idx - 1], d = in_im[(size_t)r * n + idx];
                    yi = (d - c) / (x_hi - x_lo) * (t - x_lo) + c;
                }
                const double ph = ((2.0 * M_PI) * fc) * t;     /* exp(2j*np.pi*fc*t), :88 */
                v = yr * cos(ph) - yi * sin(ph);
            }
            out[(size_t)r * m + i] = (float)v;
        }
}

/* ------------------------------------------------------------------ AttentionBlock (parity UNPINNED)
 * utils/attention_block.py:13-25 with dropout disabled: out = softmax(q k^T / sqrt(H)) v, q/k/v = x W^T + b.
 * The shipped class raises NameError (F is never imported) and no checkpoint has its weights.                     */
void stof_oracle_attention(const float *x, int B, int S, int D, int H, const float *wq, const float *bq, const float *wk,
                           const float *bk, const float *wv, const float *bv, float *out)
{
    double *q = (double *)malloc(sizeof(double) * (size_t)S * H * 3), *k = q + (size_t)S * H, *v = k + (size_t)S * H;
    double *p = (double *)malloc(sizeof(double) * (size_t)S);
    for (int b = 0; b < B; ++b) {
        const float *xb = x + (size_t)b * S * D;
        for (int s = 0; s < S; ++s)
            for (int h = 0; h < H; ++h) {
                double aq = bq[h], ak = bk[h], av = bv[h];
                for (int d = 0; d < D; ++d) {
                    const double xv = xb[(size_t)s * D + d];
                    aq += xv * wq[(size_t)h * D + d]; ak += xv * wk[(size_t)h * D + d]; av += xv * wv[(size_t)h * D + d];
                }
                q[(size_t)s * H + h] = aq; k[(size_t)s * H + h] = ak; v[(size_t)s * H + h] = av;
            }
        const double sc = 1.0 / sqrt((double)H);
        for (int s = 0; s < S; ++s) {
            double mx = -INFINITY, sum = 0.0;
            for (int j = 0; j < S; ++j) {
                double a = 0.0;
                for (int h = 0; h < H; ++h) a += q[(size_t)s * H + h] * k[(size_t)j * H + h];
                p[j] = a * sc;
                if (p[j] > mx) mx = p[j];
            }
            for (int j = 0; j < S; ++j) { p[j] = exp(p[j] - mx); sum += p[j]; }
            for (int h = 0; h < H; ++h) {
                double a = 0.0;
                for (int j = 0; j < S; ++j) a += p[j] * v[(size_t)j * H + h];
                out[((size_t)b * S + s) * H + h] = (float)(a / sum);
            }
        }
    }
    free(q); free(p);
}
