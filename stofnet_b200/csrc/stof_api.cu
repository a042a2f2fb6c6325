// C ABI of libstofnet_b200.so: model handle, weight repack, forward orchestration, host-buffer
// end-to-end call.  See include/stofnet_b200.h for the contract of every entry point.
#include <stdarg.h>
#include <stdlib.h>
#include <string.h>
#include <vector>

#include <cuda_fp16.h>

#include "common.cuh"
#include "kernels.cuh"

namespace stof {

static thread_local char g_err[512] = "";
std::atomic<unsigned long long> g_launches{0};

void set_error(const char *fmt, ...) {
    va_list ap;
    va_start(ap, fmt);
    vsnprintf(g_err, sizeof(g_err), fmt, ap);
    va_end(ap);
}

int cuda_fail(cudaError_t e, const char *what, const char *file, int line) {
    set_error("CUDA error %d (%s) at %s:%d: %s", (int)e, cudaGetErrorString(e), file, line, what);
    return STOF_E_CUDA;
}

}  // namespace stof

#include "model.cuh"

namespace stof {

static int upload(DevConv &c, const std::vector<float> &w, const float *b, int nb) {
    if (!c.w) STOF_CUDA(cudaMalloc(&c.w, w.size() * sizeof(float)));
    if (!c.b) STOF_CUDA(cudaMalloc(&c.b, (size_t)nb * sizeof(float)));
    STOF_CUDA(cudaMemcpy(c.w, w.data(), w.size() * sizeof(float), cudaMemcpyHostToDevice));
    STOF_CUDA(cudaMemcpy(c.b, b, (size_t)nb * sizeof(float), cudaMemcpyHostToDevice));
    return STOF_OK;
}

// (Cout, Cin, K) -> [Cin][K][Cout]
static std::vector<float> repack_cikco(const float *w, int Cout, int Cin, int K) {
    std::vector<float> o((size_t)Cout * Cin * K);
    for (int co = 0; co < Cout; ++co)
        for (int ci = 0; ci < Cin; ++ci)
            for (int t = 0; t < K; ++t) o[((size_t)ci * K + t) * Cout + co] = w[((size_t)co * Cin + ci) * K + t];
    return o;
}


// ---- tensor-core weight pack ------------------------------------------------------------------------
// fp32 w -> hi = fp16(w), lo = fp16((w - hi) * 2^11); one tap = [c8][rows][8] fp16 (K-major, no-swizzle core matrices)
static void split_w(float w, __half &hi, __half &lo) {
    hi = __float2half_rn(w);
    lo = __float2half_rn((w - __half2float(hi)) * 2048.0f);
}

// EDSR_1D's tail (models/edsr_1d.py:43-44): SampleShuffle1D(4) of the 64-channel tensor Y, then conv_output (16 -> 1, k3,
// pad 1) on the 4x longer signal S[c][4 w + u] = Y[16 u + c][w].  Output sample 4 l + u reads S at 4 l + u + t - 1
// (t = 0..2), i.e. Y channel 16 ((u + t - 1) mod 4) + c at position l + floor((u + t - 1) / 4): a (4, 64, 3) conv on Y
// followed by the same shuffle store as StofNet's conv_last - which is what the chain kernel's TC_LAST layer computes.
static std::vector<float> edsr_tail_weights(const float *wo /*(1,16,3)*/) {
    std::vector<float> e((size_t)4 * 64 * 3, 0.0f);
    for (int u = 0; u < 4; ++u)
        for (int t = 0; t < 3; ++t) {
            const int q = u + t - 1;
            const int dw = (q < 0) ? -1 : (q > 3 ? 1 : 0), up = (q + 4) % 4;
            for (int c = 0; c < 16; ++c) e[((size_t)u * 64 + 16 * up + c) * 3 + (dw + 1)] = wo[c * 3 + t];
        }
    return e;
}

static int upload_tc(stof_model *m, const stof_weights *w, const float *last_w /*(4,64,3)*/) {
    const int n_mid = w->n_mid;
    const int KM = w->k_mid;                                             // taps of the 64 -> 64 convs (7, or 3 for EDSR_1D)
    const bool edsr = w->kind == STOF_KIND_EDSR;
    const size_t mid_tap = 8 * 128 * 8, last_tap = 8 * 32 * 8;          // halves per tap
    std::vector<__half> buf((size_t)n_mid * KM * mid_tap + 3 * last_tap, __float2half_rn(0.0f));
    for (int j = 0; j < n_mid; ++j)
        for (int t = 0; t < KM; ++t) {
            __half *tap = buf.data() + ((size_t)j * KM + t) * mid_tap;
            for (int co = 0; co < 64; ++co)
                for (int ci = 0; ci < 64; ++ci) {
                    __half hi, lo;
                    split_w(w->mid_w[j][((size_t)co * 64 + ci) * KM + t], hi, lo);
                    tap[((size_t)(ci >> 3) * 128 + co) * 8 + (ci & 7)] = hi;
                    tap[((size_t)(ci >> 3) * 128 + 64 + co) * 8 + (ci & 7)] = lo;
                }
        }
    __half *lastp = buf.data() + (size_t)n_mid * KM * mid_tap;
    for (int t = 0; t < 3; ++t)
        for (int u = 0; u < 4; ++u)
            for (int ci = 0; ci < 64; ++ci) {
                __half hi, lo;
                split_w(last_w[((size_t)u * 64 + ci) * 3 + t], hi, lo);
                __half *tap = lastp + (size_t)t * last_tap;
                tap[((size_t)(ci >> 3) * 32 + u) * 8 + (ci & 7)] = hi;          // rows 0-3   : W_hi   (first MMA, D0)
                tap[((size_t)(ci >> 3) * 32 + 8 + u) * 8 + (ci & 7)] = lo;      // rows 8-11  : W_lo   (first MMA, D1)
                tap[((size_t)(ci >> 3) * 32 + 16 + u) * 8 + (ci & 7)] = hi;     // rows 16-19 : W_hi   (second MMA, D1)
            }
    if (!m->tc_w) STOF_CUDA(cudaMalloc(&m->tc_w, buf.size() * sizeof(__half)));
    STOF_CUDA(cudaMemcpy(m->tc_w, buf.data(), buf.size() * sizeof(__half), cudaMemcpyHostToDevice));
    // one-pass fp16 mode: W_hi only, [tap][c8][64][8] (8 KB per k7 tap); conv_last keeps its 16-row tap (rows 0-3 = W_hi)
    const size_t mid_tap_f = 8 * 64 * 8, last_tap_f = 8 * 16 * 8;
    std::vector<__half> bf((size_t)n_mid * KM * mid_tap_f + 3 * last_tap_f, __float2half_rn(0.0f));
    for (int j = 0; j < n_mid; ++j)
        for (int t = 0; t < KM; ++t) {
            __half *tap = bf.data() + ((size_t)j * KM + t) * mid_tap_f;
            for (int co = 0; co < 64; ++co)
                for (int ci = 0; ci < 64; ++ci)
                    tap[((size_t)(ci >> 3) * 64 + co) * 8 + (ci & 7)] = __float2half_rn(w->mid_w[j][((size_t)co * 64 + ci) * KM + t]);
        }
    for (int t = 0; t < 3; ++t)
        for (int u = 0; u < 4; ++u)
            for (int ci = 0; ci < 64; ++ci)
                bf[(size_t)n_mid * KM * mid_tap_f + (size_t)t * last_tap_f + ((size_t)(ci >> 3) * 16 + u) * 8 + (ci & 7)] =
                    __float2half_rn(last_w[((size_t)u * 64 + ci) * 3 + t]);
    if (!m->tc_w_fast) STOF_CUDA(cudaMalloc(&m->tc_w_fast, bf.size() * sizeof(__half)));
    STOF_CUDA(cudaMemcpy(m->tc_w_fast, bf.data(), bf.size() * sizeof(__half), cudaMemcpyHostToDevice));
    // CTA-pair kernel (conv_tc2.cu): per cluster rank its half of every tap.
    //   exact, 64 -> 64: [tap][c8][96][8]   rank 0: rows 0-63 W_hi, rows 64-95 W_hi[co 0-31];  rank 1: rows 0-63 W_lo, rows 64-95 W_hi[co 32-63]
    //   exact, conv_last: [tap][c8][16][8]  rank 0: rows 0-3 W_hi, rows 8-11 W_hi;            rank 1: rows 0-3 W_lo
    //   one-pass, 64 -> 64: [tap][c8][32][8] rank r: W_hi[co 32 r ..];   conv_last: [tap][c8][8][8]  rank 0: rows 0-3 W_hi, rank 1: zeros
    const size_t mid2 = 8 * 96 * 8, last2 = 8 * 16 * 8, mid2f = 8 * 32 * 8, last2f = 8 * 8 * 8;       // halves per tap and rank
    const size_t rank2 = (size_t)n_mid * KM * mid2 + 3 * last2, rank2f = (size_t)n_mid * KM * mid2f + 3 * last2f;
    std::vector<__half> b2(2 * rank2, __float2half_rn(0.0f)), b2f(2 * rank2f, __float2half_rn(0.0f));
    for (int j = 0; j < n_mid; ++j)
        for (int t = 0; t < KM; ++t)
            for (int co = 0; co < 64; ++co)
                for (int ci = 0; ci < 64; ++ci) {
                    __half hi, lo;
                    split_w(w->mid_w[j][((size_t)co * 64 + ci) * KM + t], hi, lo);
                    const size_t c8o = (size_t)(ci >> 3), e = ci & 7;
                    __half *t0 = b2.data() + ((size_t)j * KM + t) * mid2, *t1 = b2.data() + rank2 + ((size_t)j * KM + t) * mid2;
                    t0[(c8o * 96 + co) * 8 + e] = hi;
                    t1[(c8o * 96 + co) * 8 + e] = lo;
                    if (co < 32) t0[(c8o * 96 + 64 + co) * 8 + e] = hi;
                    else t1[(c8o * 96 + 64 + (co - 32)) * 8 + e] = hi;
                    __half *f = b2f.data() + (size_t)(co >> 5) * rank2f + ((size_t)j * KM + t) * mid2f;
                    f[(c8o * 32 + (co & 31)) * 8 + e] = hi;
                }
    for (int t = 0; t < 3; ++t)
        for (int u = 0; u < 4; ++u)
            for (int ci = 0; ci < 64; ++ci) {
                __half hi, lo;
                split_w(last_w[((size_t)u * 64 + ci) * 3 + t], hi, lo);
                const size_t c8o = (size_t)(ci >> 3), e = ci & 7;
                __half *t0 = b2.data() + (size_t)n_mid * KM * mid2 + (size_t)t * last2, *t1 = t0 + rank2;
                t0[(c8o * 16 + u) * 8 + e] = hi;
                t0[(c8o * 16 + 8 + u) * 8 + e] = hi;
                t1[(c8o * 16 + u) * 8 + e] = lo;
                b2f[(size_t)n_mid * KM * mid2f + (size_t)t * last2f + (c8o * 8 + u) * 8 + e] = hi;
            }
    if (!m->tc_w2) STOF_CUDA(cudaMalloc(&m->tc_w2, b2.size() * sizeof(__half)));
    STOF_CUDA(cudaMemcpy(m->tc_w2, b2.data(), b2.size() * sizeof(__half), cudaMemcpyHostToDevice));
    if (!m->tc_w2_fast) STOF_CUDA(cudaMalloc(&m->tc_w2_fast, b2f.size() * sizeof(__half)));
    STOF_CUDA(cudaMemcpy(m->tc_w2_fast, b2f.data(), b2f.size() * sizeof(__half), cudaMemcpyHostToDevice));
    const __half *base2 = reinterpret_cast<const __half *>(m->tc_w2), *base2f = reinterpret_cast<const __half *>(m->tc_w2_fast);
    const __half *fbase = reinterpret_cast<const __half *>(m->tc_w_fast);
    const __half *base = reinterpret_cast<const __half *>(m->tc_w);
    for (int j = 0; j < n_mid; ++j) {
        TcLayer &l = m->tc_layer[j];
        l.w = base + (size_t)j * KM * mid_tap;
        l.bias = m->mid[j].b;
        l.taps = KM; l.pad = KM / 2;
        // models/stofnet.py:51-62: conv index i = j + 2; even i -> LeakyReLU, odd i -> res + conv, last -> res1 + conv
        // models/edsr_1d.py:14-19,36-40: blocks of relu(conv1) -> conv2 + block input; conv_mid + first activation
        l.type = (j == n_mid - 1) ? TC_LONGSKIP : ((j & 1) ? TC_RES : (edsr ? TC_RELU : TC_LEAKY));
        l.stage_bytes = 16384; l.n_out = 64; l.n_stack = 128; l.n_lo = 64; l.lo_row = 0; l.lo_col = 64; l.w_rows = 128; l.n_fast = 64;
        l.w_fast = fbase + (size_t)j * KM * mid_tap_f; l.stage_bytes_fast = 8192; l.w_rows_fast = 64;
        for (int r = 0; r < 2; ++r) {
            l.w2[r] = base2 + (size_t)r * rank2 + (size_t)j * KM * mid2;
            l.w2_fast[r] = base2f + (size_t)r * rank2f + (size_t)j * KM * mid2f;
        }
        l.stage_bytes2 = 12288; l.rows2 = 96; l.lo_row2 = 64; l.stage_bytes2_fast = 4096; l.rows2_fast = 32;
    }
    TcLayer &l = m->tc_layer[n_mid];
    l.w = base + (size_t)n_mid * KM * mid_tap;
    l.bias = m->last.b;
    l.taps = 3; l.pad = 1; l.type = TC_LAST;
    l.stage_bytes = 4096; l.n_out = 4; l.n_stack = 16; l.n_lo = 16; l.lo_row = 16; l.lo_col = 8; l.w_rows = 32; l.n_fast = 16;
    l.w_fast = fbase + (size_t)n_mid * KM * mid_tap_f; l.stage_bytes_fast = 2048; l.w_rows_fast = 16;
    for (int r = 0; r < 2; ++r) {
        l.w2[r] = base2 + (size_t)r * rank2 + (size_t)n_mid * KM * mid2;
        l.w2_fast[r] = base2f + (size_t)r * rank2f + (size_t)n_mid * KM * mid2f;
    }
    l.stage_bytes2 = 2048; l.rows2 = 16; l.lo_row2 = 8; l.stage_bytes2_fast = 1024; l.rows2_fast = 8;
    m->n_tc = n_mid + 1;
    return STOF_OK;
}

static int upload_raw(float **dst, const float *src, size_t n) {
    if (!*dst) STOF_CUDA(cudaMalloc(dst, n * sizeof(float)));
    STOF_CUDA(cudaMemcpy(*dst, src, n * sizeof(float), cudaMemcpyHostToDevice));
    return STOF_OK;
}

// weights in the layouts the backward kernels read (train.cu)
static int upload_train(stof_model *m, const stof_weights *w) {
    int rc;
    if (!m->zero_bias) {
        STOF_CUDA(cudaMalloc(&m->zero_bias, 512 * sizeof(float)));
        STOF_CUDA(cudaMemset(m->zero_bias, 0, 512 * sizeof(float)));
    }
    std::vector<float> t((size_t)64 * 7 * 64);
    for (int j = 0; j < w->n_mid; ++j) {
        // dX[ci][l] = sum_co sum_t W[co][ci][t] dY[co][l - t + 3] = sum_co sum_t' W[co][ci][6 - t'] dY[co][l + t' - 3]:
        // the forward kernel's [in][tap][out] layout with in = co, out = ci, tap flipped
        for (int co = 0; co < 64; ++co)
            for (int ci = 0; ci < 64; ++ci)
                for (int tp = 0; tp < 7; ++tp) t[((size_t)co * 7 + tp) * 64 + ci] = w->mid_w[j][((size_t)co * 64 + ci) * 7 + (6 - tp)];
        if ((rc = upload_raw(&m->bw_mid[j], t.data(), t.size()))) return rc;
    }
    if ((rc = upload_raw(&m->raw_last_w, w->last_w, (size_t)4 * 64 * 3))) return rc;
    if (w->contract_w) {
        if ((rc = upload_raw(&m->raw_contract_w, w->contract_w, (size_t)512 * 64 * 5))) return rc;
        if ((rc = upload_raw(&m->raw_expand_w, w->expand_w, (size_t)64 * 512 * 5))) return rc;
    }
    return STOF_OK;
}

static int validate(const stof_weights *w) {
    STOF_REQUIRE(w != nullptr, STOF_E_INVALID, "weights: NULL");
    STOF_REQUIRE(w->conv1_w && w->conv1_b && w->last_w && w->last_b, STOF_E_INVALID, "weights: missing conv1/conv_last");
    if (w->kind == STOF_KIND_STOFNET) {
        STOF_REQUIRE(w->num_features == 64 && w->upsample_factor == 4 && w->k_first == 9 && w->k_mid == 7 && w->k_last == 3,
                     STOF_E_UNSUPPORTED, "StofNet kernels are built for F=64, r=4, kernel_sizes=[9,7,3] (got F=%d r=%d k=%d,%d,%d)",
                     w->num_features, w->upsample_factor, w->k_first, w->k_mid, w->k_last);
        STOF_REQUIRE(w->n_mid >= 3 && w->n_mid <= STOF_MAX_MID_CONVS && (w->n_mid % 2) == 1, STOF_E_UNSUPPORTED,
                     "StofNet: n_mid=%d (num_blocks-2) must be odd and in [3,%d]", w->n_mid, STOF_MAX_MID_CONVS);
        for (int i = 0; i < w->n_mid; ++i)
            STOF_REQUIRE(w->mid_w[i] && w->mid_b[i], STOF_E_INVALID, "weights: conv%d missing", i + 2);
        if (w->contract_w) {
            STOF_REQUIRE(w->contract_b && w->expand_w && w->expand_b, STOF_E_INVALID, "weights: incomplete semi-global block");
            STOF_REQUIRE(w->sgb_scale == 80 && w->sgb_feat == 512 && w->sgb_k == 5, STOF_E_UNSUPPORTED,
                         "semi-global block kernels are built for scale=80, 512 features, k=5 (got %d, %d, %d)",
                         w->sgb_scale, w->sgb_feat, w->sgb_k);
        }
    } else if (w->kind == STOF_KIND_ESPCN) {
        STOF_REQUIRE(w->upsample_factor == 4 && w->mid_w[0] && w->mid_b[0], STOF_E_UNSUPPORTED,
                     "ESPCN_1D kernels are built for upscale_factor=4");
    } else if (w->kind == STOF_KIND_EDSR) {
        STOF_REQUIRE(w->num_features == 64 && w->upsample_factor == 4 && w->k_first == 3 && w->k_mid == 3 && w->k_last == 3,
                     STOF_E_UNSUPPORTED, "EDSR_1D kernels are built for 64 features, upscale_factor=4, k=3 (got F=%d r=%d k=%d,%d,%d)",
                     w->num_features, w->upsample_factor, w->k_first, w->k_mid, w->k_last);
        STOF_REQUIRE(w->n_mid >= 3 && w->n_mid <= STOF_MAX_MID_CONVS && (w->n_mid % 2) == 1, STOF_E_UNSUPPORTED,
                     "EDSR_1D: n_mid=%d (2*num_blocks+1) must be odd and in [3,%d]", w->n_mid, STOF_MAX_MID_CONVS);
        for (int i = 0; i < w->n_mid; ++i)
            STOF_REQUIRE(w->mid_w[i] && w->mid_b[i], STOF_E_INVALID, "weights: EDSR conv %d missing", i);
    } else {
        STOF_REQUIRE(false, STOF_E_INVALID, "weights: unknown kind %d", w->kind);
    }
    return STOF_OK;
}

static int set_weights_locked(stof_model *m, const stof_weights *w);

// Weight uploads are blocking cudaMemcpy calls from pageable host vectors on the legacy default stream, while the kernels that
// read the weights run on the caller's (possibly non-blocking) streams.  Ordering contract: every call that overwrites device
// weights first waits for ALL work on the device (an earlier forward may still be reading the old weights) and returns only
// when the last byte has landed (a pageable H2D copy may return before its DMA has finished).  stof_model_set_weights must
// not be called concurrently with a forward from another thread.
static int set_weights(stof_model *m, const stof_weights *w) {
    int rc = validate(w);
    if (rc) return rc;
    DeviceGuard guard(m->device);
    STOF_REQUIRE(guard.ok, STOF_E_CUDA, "cudaSetDevice(%d) failed", m->device);
    STOF_CUDA(cudaDeviceSynchronize());
    rc = set_weights_locked(m, w);
    if (rc) return rc;
    STOF_CUDA(cudaDeviceSynchronize());
    return STOF_OK;
}

static int set_weights_locked(stof_model *m, const stof_weights *w) {
    int rc;
    m->kind = w->kind;
    m->r = w->upsample_factor;
    if (w->kind == STOF_KIND_ESPCN) {
        m->F = 64; m->n_mid = 1; m->k_first = 5; m->k_mid = 3; m->k_last = 3; m->sgb_scale = 1;
        if ((rc = upload(m->first, repack_cikco(w->conv1_w, 64, 1, 5), w->conv1_b, 64))) return rc;
        if ((rc = upload(m->mid[0], repack_cikco(w->mid_w[0], 32, 64, 3), w->mid_b[0], 32))) return rc;
        if ((rc = upload(m->last, repack_cikco(w->last_w, 4, 32, 3), w->last_b, 4))) return rc;
        // tensor-core path (espcn_tc.cu): conv2 (32, 64, 3) as packed fp16 taps [tap][c8][W_hi 32 | W_lo 32][8] (8 KB per tap),
        // then conv3 (4, 32, 3) as [tap][4 c8][32 rows][8] with rows 0-3 W_hi, 8-11 W_lo, 16-19 W_hi (2 KB per tap)
        std::vector<__half> pk((size_t)3 * 8 * 64 * 8 + (size_t)3 * 4 * 32 * 8, __float2half_rn(0.0f));
        for (int t = 0; t < 3; ++t)
            for (int oc = 0; oc < 32; ++oc)
                for (int ci = 0; ci < 64; ++ci) {
                    __half hi, lo;
                    // the factor 2 log2(e) of tanh's ex2 form rides on the weights (espcn_tc.cu: tanh_scaled)
                    split_w(2.885390081777927f * w->mid_w[0][((size_t)oc * 64 + ci) * 3 + t], hi, lo);
                    __half *tap = pk.data() + (size_t)t * 8 * 64 * 8;
                    tap[((size_t)(ci >> 3) * 64 + oc) * 8 + (ci & 7)] = hi;
                    tap[((size_t)(ci >> 3) * 64 + 32 + oc) * 8 + (ci & 7)] = lo;
                }
        for (int t = 0; t < 3; ++t)
            for (int u = 0; u < 4; ++u)
                for (int ci = 0; ci < 32; ++ci) {
                    __half hi, lo;
                    split_w(-1.4426950408889634f * w->last_w[((size_t)u * 32 + ci) * 3 + t], hi, lo);     // sigmoid_scaled
                    __half *tap = pk.data() + (size_t)3 * 8 * 64 * 8 + (size_t)t * 4 * 32 * 8;
                    tap[((size_t)(ci >> 3) * 32 + u) * 8 + (ci & 7)] = hi;
                    tap[((size_t)(ci >> 3) * 32 + 8 + u) * 8 + (ci & 7)] = lo;
                    tap[((size_t)(ci >> 3) * 32 + 16 + u) * 8 + (ci & 7)] = hi;
                }
        if (!m->tc_w) STOF_CUDA(cudaMalloc(&m->tc_w, pk.size() * sizeof(__half)));
        STOF_CUDA(cudaMemcpy(m->tc_w, pk.data(), pk.size() * sizeof(__half), cudaMemcpyHostToDevice));
        return STOF_OK;
    }
    m->F = w->num_features; m->n_mid = w->n_mid;
    m->k_first = w->k_first; m->k_mid = w->k_mid; m->k_last = w->k_last;
    if (w->kind == STOF_KIND_EDSR) {
        // models/edsr_1d.py:22-45 on the tensor-core chain: conv_input -> x0 kernel (K = 3), the 2*num_blocks + 1 k3 convs ->
        // chain layers, shuffle + conv_output -> an equivalent (4, 64, 3) conv_last (edsr_tail_weights)
        m->sgb_scale = 1; m->sgb_feat = 0; m->sgb_k = 0;
        m->seg_len = 6;                                   // k3 layers: 6 per launch fit the smem tile
        if ((rc = upload(m->first, repack_cikco(w->conv1_w, 64, 1, 3), w->conv1_b, 64))) return rc;
        for (int i = 0; i < w->n_mid; ++i)
            if ((rc = upload(m->mid[i], repack_cikco(w->mid_w[i], 64, 64, 3), w->mid_b[i], 64))) return rc;
        const std::vector<float> tail = edsr_tail_weights(w->last_w);
        const float b4[4] = {w->last_b[0], w->last_b[0], w->last_b[0], w->last_b[0]};
        if ((rc = upload(m->last, repack_cikco(tail.data(), 4, 64, 3), b4, 4))) return rc;
        return upload_tc(m, w, tail.data());
    }
    if ((rc = upload(m->first, repack_cikco(w->conv1_w, 64, 1, 9), w->conv1_b, 64))) return rc;
    for (int i = 0; i < w->n_mid; ++i)
        if ((rc = upload(m->mid[i], repack_cikco(w->mid_w[i], 64, 64, 7), w->mid_b[i], 64))) return rc;
    if ((rc = upload(m->last, repack_cikco(w->last_w, 4, 64, 3), w->last_b, 4))) return rc;
    if ((rc = upload_train(m, w))) return rc;
    // tensor-core operand packs: deferred to ensure_tc() (first inference call that needs them)
    for (int i = 0; i < w->n_mid; ++i) m->h_mid[i].assign(w->mid_w[i], w->mid_w[i] + (size_t)64 * 64 * 7);
    m->h_last.assign(w->last_w, w->last_w + (size_t)4 * 64 * 3);
    m->tc_dirty = true;
    m->tr_tc_dirty = true;
    m->host_stale = false;
    if (w->contract_w) {
        m->sgb_scale = w->sgb_scale; m->sgb_feat = w->sgb_feat; m->sgb_k = w->sgb_k;
        // 8 cout tiles of 64, each [ci][t][co]
        std::vector<float> cw((size_t)512 * 64 * 5);
        for (int ct = 0; ct < 8; ++ct) {
            std::vector<float> t = repack_cikco(w->contract_w + (size_t)ct * 64 * 64 * 5, 64, 64, 5);
            memcpy(cw.data() + (size_t)ct * 64 * 64 * 5, t.data(), t.size() * sizeof(float));
        }
        if ((rc = upload(m->contract, cw, w->contract_b, 512))) return rc;
        if ((rc = upload(m->expand, repack_cikco(w->expand_w, 64, 512, 5), w->expand_b, 64))) return rc;
        m->h_contract.assign(w->contract_w, w->contract_w + (size_t)512 * 64 * 5);
        m->h_expand.assign(w->expand_w, w->expand_w + (size_t)64 * 512 * 5);
    } else {
        m->sgb_scale = 1; m->sgb_feat = 0; m->sgb_k = 0;
    }
    return STOF_OK;
}

// Builds the fp16 hi/lo operand packs of the tensor-core kernels from the host copies kept by set_weights (StofNet; the EDSR
// packs are built eagerly).  Called by every entry point that launches a tensor-core kernel.
static int ensure_tc(stof_model *m) {
    if (!m->tc_dirty) return STOF_OK;
    int rc;
    DeviceGuard guard(m->device);
    STOF_REQUIRE(guard.ok, STOF_E_CUDA, "cudaSetDevice(%d) failed", m->device);
    STOF_CUDA(cudaDeviceSynchronize());               // nothing in flight may still read the packs that are overwritten below
    if (m->host_stale) {
        // the weights were last set on the device (stof_model_set_weights_device): fetch the host copies the packs are built
        // from - conv_last / contract / expand are kept in PyTorch's layout, the mid convs come back from [ci][t][co]
        std::vector<float> t((size_t)64 * 7 * 64);
        for (int i = 0; i < m->n_mid; ++i) {
            STOF_CUDA(cudaMemcpy(t.data(), m->mid[i].w, t.size() * sizeof(float), cudaMemcpyDeviceToHost));
            m->h_mid[i].resize((size_t)64 * 64 * 7);
            for (int co = 0; co < 64; ++co)
                for (int ci = 0; ci < 64; ++ci)
                    for (int tp = 0; tp < 7; ++tp) m->h_mid[i][((size_t)co * 64 + ci) * 7 + tp] = t[((size_t)ci * 7 + tp) * 64 + co];
        }
        m->h_last.resize((size_t)4 * 64 * 3);
        STOF_CUDA(cudaMemcpy(m->h_last.data(), m->raw_last_w, m->h_last.size() * sizeof(float), cudaMemcpyDeviceToHost));
        if (m->sgb_scale != 1) {
            m->h_contract.resize((size_t)512 * 64 * 5);
            m->h_expand.resize((size_t)64 * 512 * 5);
            STOF_CUDA(cudaMemcpy(m->h_contract.data(), m->raw_contract_w, m->h_contract.size() * sizeof(float), cudaMemcpyDeviceToHost));
            STOF_CUDA(cudaMemcpy(m->h_expand.data(), m->raw_expand_w, m->h_expand.size() * sizeof(float), cudaMemcpyDeviceToHost));
        }
        m->host_stale = false;
    }
    stof_weights v;
    memset(&v, 0, sizeof(v));
    v.kind = m->kind; v.n_mid = m->n_mid; v.num_features = m->F; v.upsample_factor = m->r;
    v.k_first = m->k_first; v.k_mid = m->k_mid; v.k_last = m->k_last;
    for (int i = 0; i < m->n_mid; ++i) v.mid_w[i] = m->h_mid[i].data();
    v.last_w = m->h_last.data();
    if ((rc = upload_tc(m, &v, v.last_w))) return rc;
    if (m->sgb_scale != 1) {
        const float *contract_w = m->h_contract.data(), *expand_w = m->h_expand.data();
        {
            std::vector<__half> buf((size_t)4 * 5 * 8 * 256 * 8);
            for (int sl = 0; sl < 4; ++sl)
                for (int t = 0; t < 5; ++t) {
                    __half *tap = buf.data() + ((size_t)sl * 5 + t) * 8 * 256 * 8;
                    for (int r = 0; r < 128; ++r)
                        for (int ci = 0; ci < 64; ++ci) {
                            __half hi, lo;
                            split_w(contract_w[((size_t)(sl * 128 + r) * 64 + ci) * 5 + t], hi, lo);
                            tap[((size_t)(ci >> 3) * 256 + r) * 8 + (ci & 7)] = hi;
                            tap[((size_t)(ci >> 3) * 256 + 128 + r) * 8 + (ci & 7)] = lo;
                        }
                }
            if (!m->tc_contract_w) STOF_CUDA(cudaMalloc(&m->tc_contract_w, buf.size() * sizeof(__half)));
            STOF_CUDA(cudaMemcpy(m->tc_contract_w, buf.data(), buf.size() * sizeof(__half), cudaMemcpyHostToDevice));
        }
        {
            // expand conv (64, 512, 5) -> [group g][tap t][c8][W_hi 64 | W_lo 64][8]: input channel 64 g + 8 c8 + (ci & 7)
            std::vector<__half> buf((size_t)8 * 5 * 8 * 128 * 8);
            for (int g = 0; g < 8; ++g)
                for (int t = 0; t < 5; ++t) {
                    __half *tap = buf.data() + ((size_t)g * 5 + t) * 8 * 128 * 8;
                    for (int co = 0; co < 64; ++co)
                        for (int ci = 0; ci < 64; ++ci) {
                            __half hi, lo;
                            split_w(expand_w[((size_t)co * 512 + 64 * g + ci) * 5 + t], hi, lo);
                            tap[((size_t)(ci >> 3) * 128 + co) * 8 + (ci & 7)] = hi;
                            tap[((size_t)(ci >> 3) * 128 + 64 + co) * 8 + (ci & 7)] = lo;
                        }
                }
            if (!m->tc_expand_w) STOF_CUDA(cudaMalloc(&m->tc_expand_w, buf.size() * sizeof(__half)));
            STOF_CUDA(cudaMemcpy(m->tc_expand_w, buf.data(), buf.size() * sizeof(__half), cudaMemcpyHostToDevice));
        }
    }
    STOF_CUDA(cudaDeviceSynchronize());               // the packs are complete before any stream's kernel can read them
    m->tc_dirty = false;
    return STOF_OK;
}

// ---- device-side weight sync (training loop) -----------------------------------------------------------------------------
// One launch re-derives every fp32 device layout from PyTorch-layout device tensors: blockIdx.y = job.
//   mode 0: copy;   mode 1: (Cout, Cin, K) -> [Cin][K][Cout];   mode 2: (Cout, Cin, K) -> [Cout][K-1-t][Cin] (dgrad weights)
struct RepackJob {
    const float *src;
    float *dst;
    int cout, cin, K, mode;
};
constexpr int kMaxRepackJobs = 64;
struct RepackJobs {
    RepackJob j[kMaxRepackJobs];
};

__global__ void __launch_bounds__(256) repack_weights_kernel(const RepackJobs jobs) {
    const RepackJob jb = jobs.j[blockIdx.y];
    const int n = jb.cout * jb.cin * jb.K;
    for (int i = blockIdx.x * 256 + threadIdx.x; i < n; i += gridDim.x * 256) {
        int s = i;                                                     // mode 0
        if (jb.mode == 1) {                                            // i = (ci K + t) Cout + co
            const int co = i % jb.cout, r = i / jb.cout, t = r % jb.K, ci = r / jb.K;
            s = (co * jb.cin + ci) * jb.K + t;
        } else if (jb.mode == 2) {                                     // i = (co K + tp) Cin + ci,  t = K - 1 - tp
            const int ci = i % jb.cin, r = i / jb.cin, tp = r % jb.K, co = r / jb.K;
            s = (co * jb.cin + ci) * jb.K + (jb.K - 1 - tp);
        }
        jb.dst[i] = __ldg(jb.src + s);
    }
}

static int set_weights_device(stof_model *m, const stof_weights *w, cudaStream_t st) {
    STOF_REQUIRE(m != nullptr && w != nullptr, STOF_E_INVALID, "stof_model_set_weights_device: NULL argument");
    STOF_REQUIRE(m->kind == STOF_KIND_STOFNET && w->kind == STOF_KIND_STOFNET, STOF_E_UNSUPPORTED,
                 "stof_model_set_weights_device: StofNet models only (the training step, main.py:217-248)");
    STOF_REQUIRE(w->n_mid == m->n_mid && (w->contract_w != nullptr) == (m->sgb_scale != 1), STOF_E_INVALID,
                 "stof_model_set_weights_device: the tensors do not match the model's architecture");
    STOF_REQUIRE(w->conv1_w && w->conv1_b && w->last_w && w->last_b, STOF_E_INVALID, "stof_model_set_weights_device: missing conv1 / conv_last");
    STOF_REQUIRE(m->first.w && m->last.w && m->raw_last_w && m->bw_mid[0], STOF_E_INVALID,
                 "stof_model_set_weights_device: the handle has no device layouts yet (create it from host weights first)");
    RepackJobs jobs;
    int n = 0, biggest = 0;
    auto add = [&](const float *src, float *dst, int cout, int cin, int K, int mode) {
        jobs.j[n++] = RepackJob{src, dst, cout, cin, K, mode};
        if (cout * cin * K > biggest) biggest = cout * cin * K;
    };
    add(w->conv1_w, m->first.w, 64, 1, 9, 1);
    add(w->conv1_b, m->first.b, 64, 1, 1, 0);
    for (int i = 0; i < m->n_mid; ++i) {
        STOF_REQUIRE(w->mid_w[i] && w->mid_b[i], STOF_E_INVALID, "stof_model_set_weights_device: conv%d missing", i + 2);
        add(w->mid_w[i], m->mid[i].w, 64, 64, 7, 1);
        add(w->mid_b[i], m->mid[i].b, 64, 1, 1, 0);
        add(w->mid_w[i], m->bw_mid[i], 64, 64, 7, 2);
    }
    add(w->last_w, m->last.w, 4, 64, 3, 1);
    add(w->last_b, m->last.b, 4, 1, 1, 0);
    add(w->last_w, m->raw_last_w, 4, 64, 3, 0);
    if (m->sgb_scale != 1) {
        STOF_REQUIRE(w->contract_b && w->expand_w && w->expand_b, STOF_E_INVALID, "stof_model_set_weights_device: incomplete semi-global block");
        for (int ct = 0; ct < 8; ++ct)                    // 8 cout tiles of 64, each [ci][t][co]
            add(w->contract_w + (size_t)ct * 64 * 64 * 5, m->contract.w + (size_t)ct * 64 * 64 * 5, 64, 64, 5, 1);
        add(w->contract_b, m->contract.b, 512, 1, 1, 0);
        add(w->contract_w, m->raw_contract_w, 512, 64, 5, 0);
        add(w->expand_w, m->expand.w, 64, 512, 5, 1);
        add(w->expand_b, m->expand.b, 64, 1, 1, 0);
        add(w->expand_w, m->raw_expand_w, 64, 512, 5, 0);
    }
    STOF_REQUIRE(n <= kMaxRepackJobs, STOF_E_INVALID, "stof_model_set_weights_device: too many tensors");
    DeviceGuard guard(m->device);
    STOF_REQUIRE(guard.ok, STOF_E_CUDA, "cudaSetDevice(%d) failed", m->device);
    dim3 grid((unsigned)((biggest + 255) / 256 < 64 ? (biggest + 255) / 256 : 64), (unsigned)n);
    repack_weights_kernel<<<grid, 256, 0, st>>>(jobs);
    STOF_LAUNCH_CHECK();
    m->tc_dirty = true;
    m->tr_tc_dirty = true;
    m->host_stale = true;
    return STOF_OK;
}

static void free_conv(DevConv &c) {
    if (c.w) cudaFree(c.w);
    if (c.b) cudaFree(c.b);
    c.w = c.b = nullptr;
}

// ---- workspace layout of one chunk of rows -----------------------------------------------------
struct Plan {
    int chunk;          // rows per chunk
    int P;              // pooled length
    size_t act;         // floats per activation buffer (chunk * 64 * L)
    size_t pooled, e;   // floats
    size_t total_bytes;
};

static size_t align256(size_t b) { return (b + 255) & ~(size_t)255; }

static int make_plan(const stof_model *m, int N, int L, int chunk_rows, Plan *p) {
    STOF_REQUIRE(m != nullptr, STOF_E_INVALID, "model: NULL");
    STOF_REQUIRE(N > 0 && L > 0, STOF_E_INVALID, "forward: N=%d L=%d must be positive", N, L);
    const bool sgb = m->kind == STOF_KIND_STOFNET && m->sgb_scale != 1;
    if (sgb) {
        STOF_REQUIRE(L >= m->sgb_scale, STOF_E_LENGTH,
                     "forward: L=%d shorter than the semi-global pooling window %d (reference: max_pool1d output size 0)",
                     L, m->sgb_scale);
        STOF_REQUIRE(((L % m->sgb_scale) & 1) == 0, STOF_E_LENGTH_ODD,
                     "forward: (L mod %d) = %d is odd; the reference pads (p//2, p//2) and fails at torch.add",
                     m->sgb_scale, L % m->sgb_scale);
    }
    int chunk = chunk_rows;
    if (chunk <= 0) {
        // default: ~4 Mi positions per chunk (>= 100 CTA tiles per SM-wave several times over, and
        // 3 activation buffers of 1 GiB each)
        long long c = (4ll << 20) / L;
        if (c < 1) c = 1;
        chunk = (int)c;
    }
    if (chunk > N) chunk = N;
    if (chunk_rows <= 0) {
        // balance: 4096 rows at a default of 819 are 6 chunks of 683, not 5 x 819 + a 1-row tail that runs every kernel of
        // the chunk on a fraction of the SMs
        const int n_chunks = (N + chunk - 1) / chunk;
        chunk = (N + n_chunks - 1) / n_chunks;
    }
    p->chunk = chunk;
    p->P = sgb ? L / m->sgb_scale : 0;
    p->act = (size_t)chunk * 64 * L;
    p->pooled = sgb ? (size_t)chunk * 512 * p->P : 0;
    p->e = sgb ? (size_t)chunk * 64 * p->P : 0;
    p->total_bytes = 3 * align256(p->act * 4) + align256(p->pooled * 4) + align256(p->e * 4);
    return STOF_OK;
}

// RAII: records an event pair around the launches issued while it is alive (profiling only)
struct ProfScope {
    stof_model *m; cudaStream_t st; size_t slot; bool on;
    ProfScope(stof_model *m_, int cls, cudaStream_t st_) : m(m_), st(st_), slot(0), on(m_->profiling) {
        if (!on) return;
        if (m->ev_used + 2 > m->ev_pool.size()) {
            cudaEvent_t a, b;
            if (cudaEventCreate(&a) != cudaSuccess || cudaEventCreate(&b) != cudaSuccess) { on = false; return; }
            m->ev_pool.push_back(a); m->ev_pool.push_back(b); m->ev_class.push_back(cls);
        } else {
            m->ev_class[m->ev_used / 2] = cls;
        }
        slot = m->ev_used; m->ev_used += 2;
        cudaEventRecord(m->ev_pool[slot], st);
    }
    ~ProfScope() { if (on) cudaEventRecord(m->ev_pool[slot + 1], st); }
};

// Tensor-core path: conv1 (+ semi-global block) on CUDA cores / SIMT, then the 64->64 stack and conv_last +
// shuffle as fused tcgen05 chain segments (conv_tc.cu).  B0/B1/B2 hold split fp16 planes (same bytes as fp32 C4L4).
static int forward_chunk_tc(stof_model *m, const float *x, int n, int L, float *mask, float *B0, float *B1, float *B2,
                            float *pooled, float *e, const Plan &pl, cudaStream_t st) {
    int rc;
    if ((rc = ensure_tc(m))) return rc;               // operand packs are built on first use after set_weights
    const bool sgb = m->sgb_scale != 1;
    const int fast = (m->path == STOF_PATH_TC_FP16) ? 1 : 0;
    int off = 0;
    if (sgb) {
        // models/stofnet.py:45,98-107: E = LeakyReLU(expand(maxpool80(LeakyReLU(contract(relu(conv1 x))))))
        { ProfScope p(m, STOF_PROF_CONTRACT, st);
          if ((rc = launch_contract_pool_tc(x, m->tc_contract_w, m->contract.b, m->first.w, m->first.b, pooled, n, L, pl.P, st, fast))) return rc; }
        { ProfScope p(m, STOF_PROF_EXPAND, st); if ((rc = launch_expand_tc(pooled, e, m->tc_expand_w, m->expand.b, n, pl.P, st))) return rc; }
        off = (L - pl.P * m->sgb_scale) / 2;
    }
    // x0 = relu(conv1 x) + upsample(E) as split planes (:45,108-115); also the long skip res1 (:51,62)
    { ProfScope p(m, STOF_PROF_SGB_ADD, st);
      if ((rc = launch_x0_split(x, sgb ? e : nullptr, n, L, pl.P, m->sgb_scale, off, m->first.w, m->first.b, B1, st, fast, m->k_first))) return rc; }
    // EXPERIMENT (STOF_INFER_CONV1L=1): the mid convs layer by layer on the weight-stationary kernel (conv1l_tc.cu) instead of
    // fused segments: no halo recompute and no weight re-streaming, but 512 B of HBM traffic per position and layer
    static const bool by_layer = [] { const char *e = getenv("STOF_INFER_CONV1L"); return e != nullptr && e[0] == '1'; }();
    if (by_layer && !fast && m->kind == STOF_KIND_STOFNET && m->tc_layer[0].taps == 7) {
        // a_j: LEAKY layers write the free buffer; short-skip layers add onto their skip IN PLACE (a thread reads its skip vector
        // before it writes the same vector, nobody else reads that buffer); B1 = x0 stays for the long skip
        const void *in = B1, *prev = B1;               // a_{j-1}, a_{j-2}
        void *fr = B2, *other = B0;
        for (int j = 0; j + 1 < m->n_tc; ++j) {
            const int type = m->tc_layer[j].type;
            ProfScope p(m, STOF_PROF_MID, st);
            if (type == TC_LEAKY) {
                void *out = (in == fr) ? other : fr;
                if ((rc = launch_conv1l_tc(m->tc_layer + j, TC_LEAKY, in, nullptr, out, n, L, st))) return rc;
                prev = in; in = out;
            } else if (type == TC_RES) {
                if (prev == B1) {                          // the first short skip is x0 itself: keep it, write the third buffer
                    void *out = (in == fr) ? other : fr;
                    if ((rc = launch_conv1l_tc(m->tc_layer + j, TC_LONGSKIP, in, prev, out, n, L, st))) return rc;
                    prev = in; in = out;
                } else {
                    if ((rc = launch_conv1l_tc(m->tc_layer + j, TC_LONGSKIP, in, prev, const_cast<void *>(prev), n, L, st))) return rc;
                    const void *t = prev; prev = in; in = t;
                }
            } else {                                       // long skip (x0)
                void *out = (in == fr) ? other : fr;
                if ((rc = launch_conv1l_tc(m->tc_layer + j, TC_LONGSKIP, in, B1, out, n, L, st))) return rc;
                prev = in; in = out;
            }
        }
        ProfScope p(m, STOF_PROF_LAST, st);
        return launch_chain_tc(m->tc_layer + m->n_tc - 1, 1, in, B1, nullptr, mask, n, L, st, 0);
    }
    const void *cur = B1;
    void *pp[2] = {B2, B0};
    int which = 0;
    int j = 0;
    while (j < m->n_tc) {
        int cnt = m->seg_len < m->n_tc - j ? m->seg_len : m->n_tc - j;
        // a segment may not start on a short-skip layer: its residual lives in the previous segment's tile
        if (j + cnt < m->n_tc && m->tc_layer[j + cnt].type == TC_RES) cnt = (cnt > 1) ? cnt - 1 : cnt + 1;
        const bool ends = (j + cnt == m->n_tc);
        ProfScope p(m, ends ? STOF_PROF_LAST : STOF_PROF_MID, st);
        if ((rc = (m->chain_pair ? launch_chain_tc2 : launch_chain_tc)(m->tc_layer + j, cnt, cur, B1, ends ? nullptr : pp[which],
                                                                       ends ? mask : nullptr, n, L, st, fast))) return rc;
        cur = pp[which];
        which ^= 1;
        j += cnt;
    }
    return STOF_OK;
}

static int forward_chunk(stof_model *m, const float *x, int n, int L, float *mask, char *ws, const Plan &pl,
                         cudaStream_t st) {
    float *B0 = reinterpret_cast<float *>(ws);
    float *B1 = reinterpret_cast<float *>(ws + align256(pl.act * 4));
    float *B2 = reinterpret_cast<float *>(ws + 2 * align256(pl.act * 4));
    float *pooled = reinterpret_cast<float *>(ws + 3 * align256(pl.act * 4));
    float *e = reinterpret_cast<float *>(ws + 3 * align256(pl.act * 4) + align256(pl.pooled * 4));
    int rc;
    if (m->kind == STOF_KIND_ESPCN) {
        // models/espcn_1d.py:31-36: one fused kernel with conv2 on tcgen05 (default), or - STOF_PATH_SIMT - the fused fp32 FFMA
        // kernel; there, layers_per_launch == 1 (stof_model_set_path) selects the three layer-by-layer kernels instead, which
        // give masks bit-identical to the fused FFMA kernel (the cross-checks of the tensor-core kernel)
        if (m->path != STOF_PATH_SIMT) {
            ProfScope p(m, STOF_PROF_MID, st);
            return launch_espcn_tc(x, mask, n, L, m->first.w, m->first.b, m->tc_w, m->mid[0].b, m->last.b, st);
        }
        if (m->seg_len != 1) {
            ProfScope p(m, STOF_PROF_MID, st);
            return launch_espcn_fused(x, mask, n, L, m->first.w, m->first.b, m->mid[0].w, m->mid[0].b, m->last.w, m->last.b, st);
        }
        { ProfScope p(m, STOF_PROF_FIRST, st); if ((rc = launch_conv_first(x, n, L, m->first.w, m->first.b, B0, 5, 1, st))) return rc; }
        { ProfScope p(m, STOF_PROF_MID, st); if ((rc = launch_espcn_conv2(B0, B1, m->mid[0].w, m->mid[0].b, n, L, st))) return rc; }
        ProfScope p(m, STOF_PROF_LAST, st);
        return launch_conv_last(B1, mask, m->last.w, m->last.b, n, L, 32, 1, st);
    }
    if (m->path != STOF_PATH_SIMT || m->kind == STOF_KIND_EDSR) return forward_chunk_tc(m, x, n, L, mask, B0, B1, B2, pooled, e, pl, st);
    // models/stofnet.py:45
    { ProfScope p(m, STOF_PROF_FIRST, st); if ((rc = launch_conv_first(x, n, L, m->first.w, m->first.b, B0, 9, 0, st))) return rc; }
    // :48 semi-global block
    if (m->sgb_scale != 1) {
        { ProfScope p(m, STOF_PROF_CONTRACT, st); if ((rc = launch_contract_pool_simt(B0, pooled, m->contract.w, m->contract.b, n, L, pl.P, st))) return rc; }
        { ProfScope p(m, STOF_PROF_EXPAND, st); if ((rc = launch_expand(pooled, e, m->expand.w, m->expand.b, n, pl.P, st))) return rc; }
        const int off = (L - pl.P * m->sgb_scale) / 2;
        { ProfScope p(m, STOF_PROF_SGB_ADD, st); if ((rc = launch_sgb_add(B0, e, n, L, pl.P, m->sgb_scale, off, st))) return rc; }
    }
    // :51-58  conv2 .. conv{nb-2}: even index -> LeakyReLU, odd index (>= 3) -> res + conv, res = x
    const float *cur = B0;     // x
    const float *res = B0;     // short skip (B0 stays intact as res1, the long skip)
    for (int i = 2; i <= m->n_mid; ++i) {          // layer index i uses mid[i-2]; last one (i = n_mid+1) below
        const DevConv &c = m->mid[i - 2];
        ProfScope p(m, STOF_PROF_MID, st);
        if (i % 2 == 0) {
            if ((rc = launch_conv_mid_simt(cur, nullptr, B1, c.w, c.b, n, L, 0, st))) return rc;
            cur = B1;
        } else {
            if ((rc = launch_conv_mid_simt(cur, res, B2, c.w, c.b, n, L, 1, st))) return rc;   // in place on B2 after the first
            cur = B2; res = B2;
        }
    }
    // :61-62  x = res1 + conv{nb-1}(x)
    {
        const DevConv &c = m->mid[m->n_mid - 1];
        float *dst = (cur == B1) ? B2 : B1;
        ProfScope p(m, STOF_PROF_MID, st);
        if ((rc = launch_conv_mid_simt(cur, B0, dst, c.w, c.b, n, L, 1, st))) return rc;
        cur = dst;
    }
    // :65 conv_last + sample shuffle
    ProfScope p(m, STOF_PROF_LAST, st);
    return launch_conv_last(cur, mask, m->last.w, m->last.b, n, L, 64, 0, st);
}

}  // namespace stof

#ifndef STOF_SRC_HASH
#define STOF_SRC_HASH "unknown"
#endif
// the hash of the sources this library was built from (csrc/Makefile): tests/test_host.py compares it with the tree, so a
// stale pre-built .so cannot pass silently
extern "C" const char *stof_version(void) { return "stofnet_b200 0.2 (sm_100a) src:" STOF_SRC_HASH; }
extern "C" const char *stof_last_error(void) { return stof::g_err; }
extern "C" uint64_t stof_launch_count(void) { return stof::g_launches.load(); }

extern "C" int stof_device_count(void) {
    int n = 0;
    if (cudaGetDeviceCount(&n) != cudaSuccess) { cudaGetLastError(); return 0; }
    return n;
}

extern "C" int stof_model_create(const stof_weights *w, int device, stof_model **out) {
    STOF_REQUIRE(out != nullptr, STOF_E_INVALID, "stof_model_create: out is NULL");
    *out = nullptr;
    int rc = stof::validate(w);
    if (rc) return rc;
    int ndev = 0;
    STOF_CUDA(cudaGetDeviceCount(&ndev));
    STOF_REQUIRE(device >= 0 && device < ndev, STOF_E_INVALID, "stof_model_create: device %d of %d", device, ndev);
    stof_model *m = new stof_model();
    m->device = device;
    rc = stof::set_weights(m, w);
    if (rc) { stof_model_destroy(m); return rc; }
    *out = m;
    return STOF_OK;
}

extern "C" int stof_model_set_weights(stof_model *m, const stof_weights *w) {
    STOF_REQUIRE(m != nullptr, STOF_E_INVALID, "stof_model_set_weights: model is NULL");
    STOF_REQUIRE(w != nullptr && w->kind == m->kind, STOF_E_INVALID, "stof_model_set_weights: kind mismatch");
    STOF_REQUIRE((w->contract_w != nullptr) == (m->sgb_scale != 1) || m->kind != STOF_KIND_STOFNET, STOF_E_INVALID,
                 "stof_model_set_weights: semi-global block presence differs from the handle");
    return stof::set_weights(m, w);
}

extern "C" void stof_model_destroy(stof_model *m) {
    if (!m) return;
    stof::DeviceGuard guard(m->device);
    stof::free_conv(m->first); stof::free_conv(m->last); stof::free_conv(m->contract); stof::free_conv(m->expand);
    for (int i = 0; i < STOF_MAX_MID_CONVS; ++i) stof::free_conv(m->mid[i]);
    for (int i = 0; i < stof_model::kPipe; ++i) {
        if (m->scratch[i]) cudaFree(m->scratch[i]);
        if (m->streams[i]) cudaStreamDestroy(m->streams[i]);
    }
    for (int i = 0; i < STOF_MAX_MID_CONVS; ++i) if (m->bw_mid[i]) cudaFree(m->bw_mid[i]);
    if (m->zero_bias) cudaFree(m->zero_bias);
    if (m->raw_last_w) cudaFree(m->raw_last_w);
    if (m->raw_contract_w) cudaFree(m->raw_contract_w);
    if (m->raw_expand_w) cudaFree(m->raw_expand_w);
    if (m->tc_w) cudaFree(m->tc_w);
    if (m->tc_w_fast) cudaFree(m->tc_w_fast);
    if (m->tc_w2) cudaFree(m->tc_w2);
    if (m->tc_w2_fast) cudaFree(m->tc_w2_fast);
    if (m->tr_tc_w) cudaFree(m->tr_tc_w);
    if (m->tr_side) cudaStreamDestroy(m->tr_side);
    for (cudaEvent_t e : m->tr_ev) if (e) cudaEventDestroy(e);
    if (m->tc_contract_w) cudaFree(m->tc_contract_w);
    if (m->tc_expand_w) cudaFree(m->tc_expand_w);
    if (m->h_stats) cudaFreeHost(m->h_stats);
    for (cudaEvent_t e : m->ev_pool) cudaEventDestroy(e);
    delete m;
}

extern "C" int stof_forward_workspace_bytes(const stof_model *m, int N, int L, int chunk_rows, size_t *bytes) {
    STOF_REQUIRE(bytes != nullptr, STOF_E_INVALID, "stof_forward_workspace_bytes: bytes is NULL");
    stof::Plan pl;
    int rc = stof::make_plan(m, N, L, chunk_rows, &pl);
    if (rc) return rc;
    *bytes = pl.total_bytes;
    return STOF_OK;
}

extern "C" int stof_forward(stof_model *m, const float *x_dev, int N, int L, float *mask_dev, void *ws_dev,
                            size_t ws_bytes, int chunk_rows, void *stream) {
    stof::Plan pl;
    int rc = stof::make_plan(m, N, L, chunk_rows, &pl);
    if (rc) return rc;
    STOF_REQUIRE(x_dev && mask_dev && ws_dev, STOF_E_INVALID, "stof_forward: NULL buffer");
    STOF_REQUIRE(ws_bytes >= pl.total_bytes, STOF_E_WORKSPACE, "stof_forward: workspace %zu < %zu bytes", ws_bytes, pl.total_bytes);
    STOF_REQUIRE((reinterpret_cast<uintptr_t>(ws_dev) & 255) == 0 && (reinterpret_cast<uintptr_t>(mask_dev) & 15) == 0,
                 STOF_E_INVALID, "stof_forward: workspace must be 256-byte and mask 16-byte aligned");
    cudaStream_t st = (cudaStream_t)stream;
    for (int n0 = 0; n0 < N; n0 += pl.chunk) {
        const int n = (N - n0 < pl.chunk) ? N - n0 : pl.chunk;
        rc = stof::forward_chunk(m, x_dev + (size_t)n0 * L, n, L, mask_dev + (size_t)n0 * L * m->r, (char *)ws_dev, pl, st);
        if (rc) return rc;
    }
    return STOF_OK;
}

extern "C" int stof_infer_host(stof_model *m, const float *x_host, int N, int L, int window, int th_is_none, float th,
                               float *coords_host, int cap, int32_t *counts_host, float *mask_host, int *kmax_out) {
    STOF_REQUIRE(m && x_host && coords_host && counts_host && cap > 0, STOF_E_INVALID, "stof_infer_host: bad arguments");
    stof::Plan pl;
    int rc = stof::make_plan(m, N, L, 0, &pl);
    if (rc) return rc;
    stof::DeviceGuard guard(m->device);
    STOF_REQUIRE(guard.ok, STOF_E_CUDA, "cudaSetDevice(%d) failed", m->device);
    constexpr int kPipe = stof_model::kPipe;
    // split into at least kPipe balanced chunks so copies overlap compute (host_chunk_rows > 0: caller's choice, test hook)
    int chunk = pl.chunk;
    if (m->host_chunk_rows > 0) {
        chunk = m->host_chunk_rows < N ? m->host_chunk_rows : N;
    } else {
        int n_chunks = (N + chunk - 1) / chunk;
        if (N >= 2 * kPipe && n_chunks < kPipe) n_chunks = kPipe;
        chunk = (N + n_chunks - 1) / n_chunks;
    }
    rc = stof::make_plan(m, N, L, chunk, &pl);
    if (rc) return rc;
    const size_t Mr = (size_t)L * m->r;
    const size_t x_b = stof::align256((size_t)pl.chunk * L * 4), mask_b = stof::align256((size_t)pl.chunk * Mr * 4);
    const size_t coords_b = stof::align256((size_t)pl.chunk * cap * 4), counts_b = stof::align256((size_t)pl.chunk * 4);
    const size_t need = pl.total_bytes + x_b + mask_b + coords_b + counts_b + 256;
    if (!m->h_stats) STOF_CUDA(cudaMallocHost(&m->h_stats, sizeof(int64_t) * 2 * kPipe));
    for (int i = 0; i < kPipe; ++i) {
        if (!m->streams[i]) STOF_CUDA(cudaStreamCreateWithFlags(&m->streams[i], cudaStreamNonBlocking));
        if (m->scratch_bytes[i] < need) {
            if (m->scratch[i]) STOF_CUDA(cudaFree(m->scratch[i]));
            m->scratch[i] = nullptr; m->scratch_bytes[i] = 0;
            STOF_CUDA(cudaMalloc(&m->scratch[i], need));
            m->scratch_bytes[i] = need;
        }
    }
    long long kmax = 0;
    int it = 0;
    for (int n0 = 0; n0 < N; n0 += pl.chunk, ++it) {
        const int s = it % kPipe;
        const int n = (N - n0 < pl.chunk) ? N - n0 : pl.chunk;
        cudaStream_t st = m->streams[s];
        if (it >= kPipe) {      // the slot's previous chunk must be done before its stats are read / buffers reused
            STOF_CUDA(cudaStreamSynchronize(st));
            if (m->h_stats[2 * s] > kmax) kmax = m->h_stats[2 * s];
        }
        char *base = (char *)m->scratch[s];
        char *ws = base;
        float *x_d = (float *)(base + pl.total_bytes);
        float *mask_d = (float *)(base + pl.total_bytes + x_b);
        float *coords_d = (float *)(base + pl.total_bytes + x_b + mask_b);
        int32_t *counts_d = (int32_t *)(base + pl.total_bytes + x_b + mask_b + coords_b);
        int64_t *stats_d = (int64_t *)(base + pl.total_bytes + x_b + mask_b + coords_b + counts_b);
        STOF_CUDA(cudaMemcpyAsync(x_d, x_host + (size_t)n0 * L, (size_t)n * L * 4, cudaMemcpyHostToDevice, st));
        STOF_CUDA(cudaMemsetAsync(stats_d, 0, 16, st));
        if ((rc = stof::forward_chunk(m, x_d, n, L, mask_d, ws, pl, st))) return rc;
        {
            stof::ProfScope p(m, STOF_PROF_PEAKS, st);
            if ((rc = stof_peaks(mask_d, n, (int)Mr, window, th_is_none, th, m->r, coords_d, nullptr, cap, counts_d, stats_d, st)))
                return rc;
        }
        STOF_CUDA(cudaMemcpyAsync(coords_host + (size_t)n0 * cap, coords_d, (size_t)n * cap * 4, cudaMemcpyDeviceToHost, st));
        STOF_CUDA(cudaMemcpyAsync(counts_host + n0, counts_d, (size_t)n * 4, cudaMemcpyDeviceToHost, st));
        STOF_CUDA(cudaMemcpyAsync(m->h_stats + 2 * s, stats_d, 16, cudaMemcpyDeviceToHost, st));
        if (mask_host)
            STOF_CUDA(cudaMemcpyAsync(mask_host + (size_t)n0 * Mr, mask_d, (size_t)n * Mr * 4, cudaMemcpyDeviceToHost, st));
    }
    for (int s = 0; s < kPipe && s < it; ++s) {
        STOF_CUDA(cudaStreamSynchronize(m->streams[s]));
        if (m->h_stats[2 * s] > kmax) kmax = m->h_stats[2 * s];
    }
    if (kmax_out) *kmax_out = (int)kmax;
    return STOF_OK;
}

extern "C" int stof_profile_enable(stof_model *m, int on) {
    STOF_REQUIRE(m != nullptr, STOF_E_INVALID, "stof_profile_enable: model is NULL");
    m->profiling = on != 0;
    m->ev_used = 0;
    return STOF_OK;
}

extern "C" int stof_profile_collect(stof_model *m, double *ms_by_class, int64_t *launches_by_class) {
    STOF_REQUIRE(m && ms_by_class && launches_by_class, STOF_E_INVALID, "stof_profile_collect: NULL argument");
    for (int c = 0; c < STOF_PROF_CLASSES; ++c) { ms_by_class[c] = 0.0; launches_by_class[c] = 0; }
    for (size_t i = 0; i + 1 < m->ev_used; i += 2) {
        STOF_CUDA(cudaEventSynchronize(m->ev_pool[i + 1]));
        float ms = 0.f;
        STOF_CUDA(cudaEventElapsedTime(&ms, m->ev_pool[i], m->ev_pool[i + 1]));
        const int c = m->ev_class[i / 2];
        ms_by_class[c] += ms;
        launches_by_class[c] += 1;
    }
    m->ev_used = 0;
    return STOF_OK;
}

extern "C" int stof_model_set_weights_device(stof_model *m, const stof_weights *w_dev, void *stream) {
    return stof::set_weights_device(m, w_dev, (cudaStream_t)stream);
}

extern "C" int stof_model_mark_weights_changed(stof_model *m) {
    STOF_REQUIRE(m != nullptr, STOF_E_INVALID, "stof_model_mark_weights_changed: model is NULL");
    m->tc_dirty = true;
    m->tr_tc_dirty = true;
    m->host_stale = true;
    return STOF_OK;
}

extern "C" int stof_model_set_path(stof_model *m, int path, int layers_per_launch) {
    STOF_REQUIRE(m != nullptr, STOF_E_INVALID, "stof_model_set_path: model is NULL");
    STOF_REQUIRE(path == STOF_PATH_SIMT || path == STOF_PATH_TC || path == STOF_PATH_TC_FP16, STOF_E_INVALID,
                 "stof_model_set_path: path %d", path);
    STOF_REQUIRE(path != STOF_PATH_SIMT || m->kind != STOF_KIND_EDSR, STOF_E_UNSUPPORTED,
                 "stof_model_set_path: EDSR_1D runs on the tensor-core path only");
    STOF_REQUIRE(path != STOF_PATH_TC_FP16 || m->kind != STOF_KIND_ESPCN, STOF_E_UNSUPPORTED,
                 "stof_model_set_path: ESPCN_1D has no one-pass fp16 mode (fp32-faithful tensor-core path or fp32 FFMA path)");
    if (layers_per_launch > 0) {
        STOF_REQUIRE(layers_per_launch <= stof::TC_MAX_LAYERS, STOF_E_INVALID, "stof_model_set_path: at most %d layers per launch",
                     stof::TC_MAX_LAYERS);
        m->seg_len = layers_per_launch;
    }
    m->path = path;
    return STOF_OK;
}

extern "C" int stof_model_set_chain_kernel(stof_model *m, int cta_pairs) {
    STOF_REQUIRE(m != nullptr, STOF_E_INVALID, "stof_model_set_chain_kernel: model is NULL");
    m->chain_pair = cta_pairs != 0;
    return STOF_OK;
}

extern "C" int stof_model_set_host_chunk_rows(stof_model *m, int rows) {
    STOF_REQUIRE(m != nullptr && rows >= 0, STOF_E_INVALID, "stof_model_set_host_chunk_rows: bad arguments");
    m->host_chunk_rows = rows;
    return STOF_OK;
}

extern "C" int stof_chain_segment(stof_model *m, int first, int count, const void *in_planes_dev, const void *x0_planes_dev,
                                  void *out_dev, int N, int L, void *stream) {
    STOF_REQUIRE(m != nullptr && m->kind != STOF_KIND_ESPCN, STOF_E_INVALID, "stof_chain_segment: needs a StofNet or EDSR_1D model");
    { const int rc = stof::ensure_tc(m); if (rc) return rc; }
    STOF_REQUIRE(first >= 0 && count >= 1 && first + count <= m->n_tc, STOF_E_INVALID,
                 "stof_chain_segment: layers [%d, %d) of %d", first, first + count, m->n_tc);
    STOF_REQUIRE(in_planes_dev && out_dev && N > 0 && L > 0, STOF_E_INVALID, "stof_chain_segment: bad arguments");
    const bool ends = m->tc_layer[first + count - 1].type == stof::TC_LAST;
    return (m->chain_pair ? stof::launch_chain_tc2 : stof::launch_chain_tc)(
        m->tc_layer + first, count, in_planes_dev, x0_planes_dev, ends ? nullptr : out_dev,
        ends ? reinterpret_cast<float *>(out_dev) : nullptr, N, L, (cudaStream_t)stream, m->path == STOF_PATH_TC_FP16 ? 1 : 0);
}

extern "C" int stof_sgb_expand(stof_model *m, const float *pooled_dev, int N, int P, float *e_dev, int path, void *stream) {
    STOF_REQUIRE(m != nullptr && m->kind == STOF_KIND_STOFNET && m->sgb_scale != 1, STOF_E_INVALID,
                 "stof_sgb_expand: needs a StofNet model with a semi-global block");
    STOF_REQUIRE(pooled_dev && e_dev && N > 0 && P > 0, STOF_E_INVALID, "stof_sgb_expand: bad arguments");
    stof::DeviceGuard guard(m->device);
    STOF_REQUIRE(guard.ok, STOF_E_CUDA, "cudaSetDevice(%d) failed", m->device);
    if (path == STOF_PATH_SIMT) return stof::launch_expand(pooled_dev, e_dev, m->expand.w, m->expand.b, N, P, (cudaStream_t)stream);
    { const int rc = stof::ensure_tc(m); if (rc) return rc; }
    return stof::launch_expand_tc(pooled_dev, e_dev, m->tc_expand_w, m->expand.b, N, P, (cudaStream_t)stream);
}

extern "C" int stof_debug_set_timeline(void *dev_buf, int capacity_events) {
    stof::set_chain_timeline(dev_buf, dev_buf ? capacity_events : 0);
    return STOF_OK;
}
