// Training step of the StofNet stack (reference main.py:217-248; models/stofnet.py:42-117): a forward pass that keeps every
// layer's activations, the backward pass (what torch autograd derives from the reference's forward), and a fused AdamW.
//
// fp32 on CUDA cores, activations in the C4L4 layout of conv_simt.cu; data gradients of the 64->64 k7 convs reuse the forward
// conv with channel-swapped, tap-flipped weights (a transposed conv with 'same' padding is a conv); weight gradients are
// register-tiled outer-product reductions over positions with per-CTA partial sums that a second kernel adds in a fixed order,
// so gradients are bit-reproducible run to run (no float atomics).
// On the tensor-core path (default; STOF_PATH_SIMT keeps the all-fp32 step) the 22 big contractions of a step - forward and
// data gradient of conv2..conv12, 75 % of its FLOPs - run on the tcgen05 chain kernel (conv_tc.cu), one layer per launch on
// split fp16 planes: operand packs built on the device from the current weights, gradient planes scaled by a power of two
// taken from the tensor's max |g| on the device (no host synchronisation), fp32 C4L4 copies kept for the weight gradients.
//
//   forward  (saved)   x0r = relu(conv1 x);  pooled, argmax = maxpool80(lrelu(contract x0r));  E = lrelu(expand pooled);
//                      x0 = x0r + up80(E);  a2 = lrelu(c2 x0);  a3 = x0 + c3 a2;  ...  a12 = x0 + c12 a11;  y = c_last a12
//   backward           g_y = g_mask (the shuffle is a relabelling);  conv_last, conv12..conv2 (pairs: skip + lrelu'),
//                      semi-global block (window sum, expand, arg-max scatter through the pool, contract), conv1
#include <cuda_fp16.h>
#include <math.h>
#include <stdlib.h>
#include <string.h>

#include "model.cuh"

namespace stof {

// ------------------------------------------------------------------------------------------------ workspace plan
struct TrainPlan {
    int P;
    bool sgb;
    size_t act;                    // floats of one (N, 64, L) tensor
    size_t off_x0r, off_x0, off_a[STOF_MAX_MID_CONVS], off_g[3];     // float offsets
    size_t off_pooled, off_argmax, off_e, off_ge, off_gpool;
    size_t off_partial;            // float offset of the partial-sum scratch
    size_t partial_floats;
    // tensor-core path: split planes (same size as an fp32 tensor) of x0 and of every mid conv's output, two gradient planes,
    // and the max |g| slots (one unsigned per data-gradient call)
    bool tc;
    size_t off_px0, off_pa[STOF_MAX_MID_CONVS], off_gp[2], off_absmax, off_bpart, off_bparts, off_wpart;
    size_t total_bytes;
};

constexpr int WG_MAX_PARTS = kSMs;                 // persistent wgrad CTAs
constexpr int CWG_MAX_PARTS = 37;                  // parts of the semi-global block's weight gradients: 8 channel blocks x 37 = 2 x 148 CTAs
constexpr size_t WG_MID_LEN = 64 * 64 * 7 + 64;    // dW + db of one mid conv

static size_t a64(size_t floats) { return (floats + 63) & ~(size_t)63; }

static int make_train_plan(const stof_model *m, int N, int L, TrainPlan *p) {
    STOF_REQUIRE(m != nullptr && m->kind == STOF_KIND_STOFNET, STOF_E_INVALID, "training: needs a StofNet model");
    STOF_REQUIRE(N > 0 && L > 0, STOF_E_INVALID, "training: N=%d L=%d must be positive", N, L);
    p->sgb = m->sgb_scale != 1;
    if (p->sgb) {
        STOF_REQUIRE(L >= m->sgb_scale, STOF_E_LENGTH, "training: L=%d shorter than the semi-global pooling window %d", L, m->sgb_scale);
        STOF_REQUIRE(((L % m->sgb_scale) & 1) == 0, STOF_E_LENGTH_ODD, "training: (L mod %d) = %d is odd", m->sgb_scale, L % m->sgb_scale);
    }
    p->P = p->sgb ? L / m->sgb_scale : 0;
    p->act = a64((size_t)N * 64 * L);
    size_t o = 0;
    p->tc = (m->path != STOF_PATH_SIMT);
    p->off_x0r = o; o += p->act;
    // the tensor-core path keeps x0 and the mid activations as planes (below); in fp32 it needs x0r and the last activation only
    if (p->sgb && !p->tc) { p->off_x0 = o; o += p->act; } else p->off_x0 = p->off_x0r;
    for (int j = 0; j < m->n_mid; ++j) {
        p->off_a[j] = o;
        if (!p->tc || j == m->n_mid - 1) o += p->act;
    }
    if (p->tc) for (int j = 0; j < m->n_mid - 1; ++j) p->off_a[j] = p->off_a[m->n_mid - 1];     // never read on that path
    for (int i = 0; i < 3; ++i) { p->off_g[i] = o; o += p->act; }
    const size_t np = a64((size_t)N * 512 * p->P), ne = a64((size_t)N * 64 * p->P);
    p->off_pooled = o; o += p->sgb ? np : 0;
    p->off_argmax = o; o += p->sgb ? np : 0;
    p->off_gpool = o; o += p->sgb ? np : 0;
    p->off_e = o; o += p->sgb ? ne : 0;
    p->off_ge = o; o += p->sgb ? ne : 0;
    // partial sums: the largest user is the contract wgrad (one partial per row chunk is avoided: it reduces in-kernel),
    // then mid wgrad (WG_MAX_PARTS x WG_MID_LEN)
    p->partial_floats = a64((size_t)WG_MAX_PARTS * WG_MID_LEN);
    if (m->path != STOF_PATH_SIMT && a64(wgrad_tc_scratch_floats()) > p->partial_floats) p->partial_floats = a64(wgrad_tc_scratch_floats());
    if (p->sgb && a64((size_t)CWG_MAX_PARTS * (512 * 64 * 5)) > p->partial_floats) p->partial_floats = a64((size_t)CWG_MAX_PARTS * (512 * 64 * 5));
    p->off_partial = o; o += p->partial_floats;
    if (p->tc) {
        p->off_px0 = o; o += p->act;
        for (int j = 0; j < m->n_mid; ++j) { p->off_pa[j] = o; o += p->act; }
        for (int i = 0; i < 2; ++i) { p->off_gp[i] = o; o += p->act; }
        p->off_absmax = o; o += 64;
        p->off_bpart = o; o += a64((size_t)grad_to_planes_parts(N, L) * 64);          // bias partials of the first gradient tensor
        p->off_bparts = o; o += a64((size_t)kSMs * 64) * m->n_mid;                     // ... of each chain launch (read on the side stream)
        p->off_wpart = o; o += a64(wgrad_tc_scratch_floats());                         // wgrad_tc's partials (side stream)
    }
    p->total_bytes = o * sizeof(float);
    return STOF_OK;
}

// ------------------------------------------------------------------------------------------------ elementwise helpers
// g[i] *= lrelu'(a[i])   (a = LeakyReLU output: same sign as its input; derivative at 0 is the negative slope, as in ATen)
__global__ void __launch_bounds__(256)
lrelu_bwd_kernel(float4 *__restrict__ g, const float4 *__restrict__ a, size_t n4) {
    for (size_t i = (size_t)blockIdx.x * 256 + threadIdx.x; i < n4; i += (size_t)gridDim.x * 256) {
        float4 gv = g[i];
        const float4 av = __ldg(a + i);
        gv.x *= av.x > 0.f ? 1.f : kLeakySlope; gv.y *= av.y > 0.f ? 1.f : kLeakySlope;
        gv.z *= av.z > 0.f ? 1.f : kLeakySlope; gv.w *= av.w > 0.f ? 1.f : kLeakySlope;
        g[i] = gv;
    }
}
// g[i] = (a[i] > 0) ? g[i] : 0     (ReLU backward)
__global__ void __launch_bounds__(256)
relu_bwd_kernel(float4 *__restrict__ g, const float4 *__restrict__ a, size_t n4) {
    for (size_t i = (size_t)blockIdx.x * 256 + threadIdx.x; i < n4; i += (size_t)gridDim.x * 256) {
        float4 gv = g[i];
        const float4 av = __ldg(a + i);
        gv.x = av.x > 0.f ? gv.x : 0.f; gv.y = av.y > 0.f ? gv.y : 0.f;
        gv.z = av.z > 0.f ? gv.z : 0.f; gv.w = av.w > 0.f ? gv.w : 0.f;
        g[i] = gv;
    }
}
// d[i] += s[i]
__global__ void __launch_bounds__(256)
add_kernel(float4 *__restrict__ d, const float4 *__restrict__ s, size_t n4) {
    for (size_t i = (size_t)blockIdx.x * 256 + threadIdx.x; i < n4; i += (size_t)gridDim.x * 256) {
        float4 dv = d[i];
        const float4 sv = __ldg(s + i);
        dv.x += sv.x; dv.y += sv.y; dv.z += sv.z; dv.w += sv.w;
        d[i] = dv;
    }
}
static int ew_grid(size_t n4) { size_t g = (n4 + 255) / 256; return (int)(g < (size_t)kSMs * 8 ? g : (size_t)kSMs * 8); }

// out[j] = sum_p partial[p][j]   (fixed order)
__global__ void __launch_bounds__(256)
reduce_parts_kernel(const float *__restrict__ partial, int parts, size_t len, float *__restrict__ out_w, size_t len_w,
                    float *__restrict__ out_b) {
    const size_t j = (size_t)blockIdx.x * 256 + threadIdx.x;
    if (j >= len) return;
    float s = 0.0f;
    for (int p = 0; p < parts; ++p) s += partial[(size_t)p * len + j];
    if (j < len_w) out_w[j] = s; else out_b[j - len_w] = s;
}

// ------------------------------------------------------------------------------------------------ conv_last backward
// y[n][u][l] = b[u] + sum_ci sum_t W[u][ci][t] a[n][ci][l + t - 1];  mask[n][4l + u] = y[n][u][l]  => g_y[n][l][u] = g_mask[n][4l+u]
// dgrad: g_a[n][ci][l] = sum_u sum_t W[u][ci][t] g_y[n][u][l - t + 1]     (C4L4 output)
__global__ void __launch_bounds__(256)
last_dgrad_kernel(const float4 *__restrict__ gy /*[N][L] (4 u)*/, const float *__restrict__ w /*(4,64,3)*/, int N, int L,
                  float4 *__restrict__ ga /*C4L4*/) {
    __shared__ float ws[4 * 64 * 3];
    for (int i = threadIdx.x; i < 768; i += 256) ws[i] = __ldg(w + i);
    __syncthreads();
    const int n = blockIdx.y;
    const int l = blockIdx.x * 256 + threadIdx.x;
    if (l >= L) return;
    float4 g[3];                                   // g_y at l + 1, l, l - 1  (t = 0, 1, 2)
#pragma unroll
    for (int t = 0; t < 3; ++t) {
        const int q = l - t + 1;
        g[t] = (q >= 0 && q < L) ? __ldg(gy + (size_t)n * L + q) : make_float4(0.f, 0.f, 0.f, 0.f);
    }
    for (int c4 = 0; c4 < 16; ++c4) {
        float o[4];
#pragma unroll
        for (int i = 0; i < 4; ++i) {
            const int ci = 4 * c4 + i;
            float a = 0.0f;
#pragma unroll
            for (int t = 0; t < 3; ++t) {
                a = fmaf(ws[(0 * 64 + ci) * 3 + t], g[t].x, a);
                a = fmaf(ws[(1 * 64 + ci) * 3 + t], g[t].y, a);
                a = fmaf(ws[(2 * 64 + ci) * 3 + t], g[t].z, a);
                a = fmaf(ws[(3 * 64 + ci) * 3 + t], g[t].w, a);
            }
            o[i] = a;
        }
        ga[((size_t)n * 16 + c4) * L + l] = make_float4(o[0], o[1], o[2], o[3]);
    }
}

// wgrad: dW[u][ci][t] = sum_{n,l} g_y[n][u][l] a[n][ci][l + t - 1];  db[u] = sum g_y.   Persistent CTAs over (row, 128-position)
// tiles staged in shared memory with coalesced 16-byte loads (a: 16 c4 slabs x 130 positions incl. the +-1 halo, zero outside the
// row; g_y: 128 float4); thread = (ci, t) (192 threads) keeps 4 accumulators (u) and reads one broadcast float4 of g_y and one
// word of a per position.  (The first version walked global memory position by position from every thread: latency-bound,
// 0.57 ms of the 256 x 2560 step for 0.5 GMAC.)  partial [768 + 4] per CTA, added in CTA order by reduce_parts_kernel.
constexpr int LW_TL = 128;
constexpr int LW_AP = 131;                      // float4 pitch of an a slab: 524 words = 12 mod 32 -> a warp's three c4 groups miss each other
constexpr int LW_MAX_CTAS = 4 * kSMs;
__global__ void __launch_bounds__(192)
last_wgrad_kernel(const float4 *__restrict__ gy, const float4 *__restrict__ a /*C4L4*/, int N, int L, int tiles_per_row,
                  float *__restrict__ partial /*[gridDim.x][772]*/) {
    __shared__ float4 As[16 * LW_AP];
    __shared__ float4 Gs[LW_TL];
    __shared__ float4 red[LW_TL];
    const int ci = threadIdx.x / 3, t = threadIdx.x - 3 * ci;
    float acc[4] = {0.f, 0.f, 0.f, 0.f};
    float4 bs = make_float4(0.f, 0.f, 0.f, 0.f);            // thread i < 128: sum of g_y over the positions l0 + i of its tiles
    const int total = N * tiles_per_row;
    for (int tile = blockIdx.x; tile < total; tile += gridDim.x) {
        const int n = tile / tiles_per_row, l0 = (tile - n * tiles_per_row) * LW_TL;
        __syncthreads();
        for (int i = threadIdx.x; i < 16 * (LW_TL + 2); i += 192) {
            const int c4 = i / (LW_TL + 2), j = i - c4 * (LW_TL + 2);
            const int q = l0 + j - 1;
            As[c4 * LW_AP + j] = (q >= 0 && q < L) ? __ldg(a + ((size_t)n * 16 + c4) * L + q) : make_float4(0.f, 0.f, 0.f, 0.f);
        }
        if (threadIdx.x < LW_TL) {
            const int l = l0 + threadIdx.x;
            const float4 g = (l < L) ? __ldg(gy + (size_t)n * L + l) : make_float4(0.f, 0.f, 0.f, 0.f);
            Gs[threadIdx.x] = g;
            bs.x += g.x; bs.y += g.y; bs.z += g.z; bs.w += g.w;
        }
        __syncthreads();
        const float *ar = reinterpret_cast<const float *>(As + (ci >> 2) * LW_AP + t) + (ci & 3);
#pragma unroll 8
        for (int j = 0; j < LW_TL; ++j) {
            const float4 g = Gs[j];
            const float av = ar[4 * j];
            acc[0] = fmaf(g.x, av, acc[0]); acc[1] = fmaf(g.y, av, acc[1]);
            acc[2] = fmaf(g.z, av, acc[2]); acc[3] = fmaf(g.w, av, acc[3]);
        }
    }
    float *o = partial + (size_t)blockIdx.x * 772;
#pragma unroll
    for (int u = 0; u < 4; ++u) o[(u * 64 + ci) * 3 + t] = acc[u];
    __syncthreads();
    if (threadIdx.x < LW_TL) red[threadIdx.x] = bs;
    __syncthreads();
    for (int k = LW_TL / 2; k > 0; k >>= 1) {
        if ((int)threadIdx.x < k) {
            const float4 y = red[threadIdx.x + k];
            float4 x = red[threadIdx.x];
            x.x += y.x; x.y += y.y; x.z += y.z; x.w += y.w;
            red[threadIdx.x] = x;
        }
        __syncthreads();
    }
    if (threadIdx.x == 0) { o[768] = red[0].x; o[769] = red[0].y; o[770] = red[0].z; o[771] = red[0].w; }
}

// ------------------------------------------------------------------------------------------------ mid conv wgrad
// dW[co][ci][t] = sum_{n,l} G[n][co][l] A[n][ci][l + t - 3],  db[co] = sum G[n][co][l]        (64 x 64 x 7 + 64 outputs)
// Persistent CTAs; a tile = 128 positions of one row staged in shared memory (G: 16 x 128 float4, A: 16 x 134 float4 with the
// conv's zero padding).  Thread (cog, cig) owns 4 co x 4 ci x 7 taps = 112 accumulators; per position it reads one float4 of G
// and one new float4 of A (the 7-tap window slides through registers).  Each CTA writes one partial [64][64][7] (+ bias).
constexpr int WG_TL = 128;
constexpr int WG_AP = 135;          // float4 pitch of an A slab (>= WG_TL + 6, odd multiple keeps 8 consecutive slabs conflict-free)
constexpr int WG_GP = 129;

__global__ void __launch_bounds__(256, 1)
mid_wgrad_kernel(const float4 *__restrict__ G /*C4L4*/, const float4 *__restrict__ A /*C4L4*/, int N, int L, int tiles_per_row,
                 float *__restrict__ partial /*[gridDim.x][WG_MID_LEN]*/) {
    extern __shared__ __align__(16) float4 wg_smem[];
    float4 *Gs = wg_smem;                     // [16][WG_GP]
    float4 *As = wg_smem + 16 * WG_GP;        // [16][WG_AP]
    const int cog = threadIdx.x >> 4, cig = threadIdx.x & 15;
    float acc[7][4][4];
#pragma unroll
    for (int t = 0; t < 7; ++t)
#pragma unroll
        for (int i = 0; i < 4; ++i)
#pragma unroll
            for (int j = 0; j < 4; ++j) acc[t][i][j] = 0.0f;
    float bsum[4] = {0.f, 0.f, 0.f, 0.f};
    const int total = N * tiles_per_row;
    for (int tile = blockIdx.x; tile < total; tile += gridDim.x) {
        const int n = tile / tiles_per_row, l0 = (tile - n * tiles_per_row) * WG_TL;
        __syncthreads();
        for (int idx = threadIdx.x; idx < 16 * WG_TL; idx += 256) {
            const int c4 = idx / WG_TL, j = idx - c4 * WG_TL;
            const int l = l0 + j;
            Gs[c4 * WG_GP + j] = (l < L) ? __ldg(G + ((size_t)n * 16 + c4) * L + l) : make_float4(0.f, 0.f, 0.f, 0.f);
        }
        for (int idx = threadIdx.x; idx < 16 * (WG_TL + 6); idx += 256) {
            const int c4 = idx / (WG_TL + 6), j = idx - c4 * (WG_TL + 6);
            const int l = l0 + j - 3;
            As[c4 * WG_AP + j] = (l >= 0 && l < L) ? __ldg(A + ((size_t)n * 16 + c4) * L + l) : make_float4(0.f, 0.f, 0.f, 0.f);
        }
        __syncthreads();
        const float4 *gp = Gs + cog * WG_GP;
        const float4 *ap = As + cig * WG_AP;
        float4 win[7];                       // A at positions j .. j + 6  (tap t reads win[t])
#pragma unroll
        for (int t = 0; t < 6; ++t) win[t + 1] = ap[t];
#pragma unroll 2
        for (int j = 0; j < WG_TL; ++j) {
#pragma unroll
            for (int t = 0; t < 6; ++t) win[t] = win[t + 1];
            win[6] = ap[j + 6];
            const float4 g = gp[j];
            const float gv[4] = {g.x, g.y, g.z, g.w};
#pragma unroll
            for (int t = 0; t < 7; ++t) {
                const float av[4] = {win[t].x, win[t].y, win[t].z, win[t].w};
#pragma unroll
                for (int i = 0; i < 4; ++i)
#pragma unroll
                    for (int jj = 0; jj < 4; ++jj) acc[t][i][jj] = fmaf(gv[i], av[jj], acc[t][i][jj]);
            }
            if (cig == 0) { bsum[0] += g.x; bsum[1] += g.y; bsum[2] += g.z; bsum[3] += g.w; }
        }
    }
    float *o = partial + (size_t)blockIdx.x * WG_MID_LEN;
#pragma unroll
    for (int i = 0; i < 4; ++i)
#pragma unroll
        for (int jj = 0; jj < 4; ++jj)
#pragma unroll
            for (int t = 0; t < 7; ++t) o[((size_t)(4 * cog + i) * 64 + 4 * cig + jj) * 7 + t] = acc[t][i][jj];
    if (cig == 0) {
#pragma unroll
        for (int i = 0; i < 4; ++i) o[64 * 64 * 7 + 4 * cog + i] = bsum[i];
    }
}

// ------------------------------------------------------------------------------------------------ semi-global block backward
// x0 = x0r + pad(up80(E)):  g_E[n][c][p] = sum_{l = off + 80 p}^{+79} g_x0[n][c][l]      (one warp per (n, c4, p))
__global__ void __launch_bounds__(256)
window_sum_kernel(const float4 *__restrict__ gx0 /*C4L4*/, int N, int L, int P, int scale, int off, float *__restrict__ ge /*[N][64][P]*/) {
    const long long wid = (long long)blockIdx.x * 8 + (threadIdx.x >> 5);
    const int lane = threadIdx.x & 31;
    if (wid >= (long long)N * 16 * P) return;
    const int p = (int)(wid % P);
    const int c4 = (int)((wid / P) % 16);
    const int n = (int)(wid / ((long long)P * 16));
    const float4 *src = gx0 + ((size_t)n * 16 + c4) * L + off + (size_t)p * scale;
    float4 s = make_float4(0.f, 0.f, 0.f, 0.f);
    for (int j = lane; j < scale; j += 32) {
        const float4 v = __ldg(src + j);
        s.x += v.x; s.y += v.y; s.z += v.z; s.w += v.w;
    }
#pragma unroll
    for (int o = 16; o; o >>= 1) {
        s.x += __shfl_xor_sync(0xffffffffu, s.x, o); s.y += __shfl_xor_sync(0xffffffffu, s.y, o);
        s.z += __shfl_xor_sync(0xffffffffu, s.z, o); s.w += __shfl_xor_sync(0xffffffffu, s.w, o);
    }
    if (lane == 0) {
        float *d = ge + ((size_t)n * 64 + 4 * c4) * P + p;
        d[0] = s.x; d[P] = s.y; d[2 * (size_t)P] = s.z; d[3 * (size_t)P] = s.w;
    }
}

// plain elementwise lrelu' on NCL tensors (E, pooled): g *= a > 0 ? 1 : slope
__global__ void __launch_bounds__(256)
lrelu_bwd1_kernel(float *__restrict__ g, const float *__restrict__ a, size_t n) {
    for (size_t i = (size_t)blockIdx.x * 256 + threadIdx.x; i < n; i += (size_t)gridDim.x * 256)
        g[i] *= (__ldg(a + i) > 0.f) ? 1.f : kLeakySlope;
}

// expand conv (512 -> 64, k5, pad 2 on the pooled axis):  e_pre[n][co][p] = b + sum_ci sum_t We[co][ci][t] pooled[n][ci][p+t-2]
// wgrad: dWe[co][ci][t] = sum_{n,p} g[n][co][p] pooled[n][ci][p + t - 2];  CTA = (block of 64 ci, part of the rows); thread =
// (co group, ci group) owns 4 co x 4 ci x 5 taps = 80 accumulators (co = cog + 16 a, ci = cig + 16 b: a warp's 16 ci rows sit in
// distinct banks, its two co rows are broadcasts).  Per row and 32-position chunk the CTA stages g (64 x 32) and the block's 64
// pooled rows with their +-2 halo (zero outside [0, P)); per position a thread reads 4 g values and the 4 x 5 window (the
// unrolled loop keeps the overlapping window values in registers): ~0.2 shared-memory reads per FMA.
// (History at 256 x 2560: both operands from global memory in every thread 1.49 ms; thread = (co, t) with 16 ci accumulators and
// one shared-memory read per FMA 0.23 ms.)  Fixed (n, p) order; partial [part][(64,512,5)], added in part order by reduce_parts_kernel.
constexpr int EW_CB = 64, EW_PC = 32;
constexpr int EW_GP = EW_PC + 1, EW_PP = EW_PC + 5;
__global__ void __launch_bounds__(256, 2)
expand_wgrad_kernel(const float *__restrict__ g /*[N][64][P]*/, const float *__restrict__ pooled /*[N][512][P]*/, int N, int P,
                    float *__restrict__ partial) {
    __shared__ float gs[64][EW_GP];
    __shared__ float ps[EW_CB][EW_PP];
    const int ci0 = blockIdx.x * EW_CB;
    const int cog = threadIdx.x >> 4, cig = threadIdx.x & 15;
    float acc[4][4][5];
#pragma unroll
    for (int a = 0; a < 4; ++a)
#pragma unroll
        for (int b = 0; b < 4; ++b)
#pragma unroll
            for (int t = 0; t < 5; ++t) acc[a][b][t] = 0.0f;
    for (int n = blockIdx.y; n < N; n += gridDim.y) {
        for (int p0 = 0; p0 < P; p0 += EW_PC) {
            __syncthreads();
            for (int i = threadIdx.x; i < 64 * EW_PC; i += 256) {
                const int c = i / EW_PC, j = i - c * EW_PC;
                gs[c][j] = (p0 + j < P) ? __ldg(g + ((size_t)n * 64 + c) * P + p0 + j) : 0.0f;
            }
            for (int i = threadIdx.x; i < EW_CB * (EW_PC + 4); i += 256) {
                const int c = i / (EW_PC + 4), j = i - c * (EW_PC + 4);
                const int q = p0 + j - 2;
                ps[c][j] = (q >= 0 && q < P) ? __ldg(pooled + ((size_t)n * 512 + ci0 + c) * P + q) : 0.0f;
            }
            __syncthreads();
#pragma unroll 8
            for (int j = 0; j < EW_PC; ++j) {
                float gv[4];
#pragma unroll
                for (int a = 0; a < 4; ++a) gv[a] = gs[cog + 16 * a][j];
#pragma unroll
                for (int b = 0; b < 4; ++b)
#pragma unroll
                    for (int t = 0; t < 5; ++t) {
                        const float x = ps[cig + 16 * b][j + t];
#pragma unroll
                        for (int a = 0; a < 4; ++a) acc[a][b][t] = fmaf(gv[a], x, acc[a][b][t]);
                    }
            }
        }
    }
    float *dW = partial + (size_t)blockIdx.y * (64 * 512 * 5);
#pragma unroll
    for (int a = 0; a < 4; ++a)
#pragma unroll
        for (int b = 0; b < 4; ++b)
#pragma unroll
            for (int t = 0; t < 5; ++t) dW[((size_t)(cog + 16 * a) * 512 + ci0 + cig + 16 * b) * 5 + t] = acc[a][b][t];
}
// db[c] = sum_{n,p} g[n][c][p]     (NCL tensor with C channels; one CTA per channel, fixed-order tree)
__global__ void __launch_bounds__(256)
bias_grad_ncl_kernel(const float *__restrict__ g, int N, int Cn, int P, float *__restrict__ db) {
    __shared__ float red[256];
    const int c = blockIdx.x;
    float s = 0.0f;
    for (long long i = threadIdx.x; i < (long long)N * P; i += 256) {
        const int n = (int)(i / P), p = (int)(i - (long long)n * P);
        s += __ldg(g + ((size_t)n * Cn + c) * P + p);
    }
    red[threadIdx.x] = s;
    __syncthreads();
    for (int k = 128; k; k >>= 1) { if ((int)threadIdx.x < k) red[threadIdx.x] += red[threadIdx.x + k]; __syncthreads(); }
    if (threadIdx.x == 0) db[c] = red[0];
}
// dgrad: g_pooled[n][ci][p] = sum_co sum_t We[co][ci][t] g[n][co][p - t + 2];  then through LeakyReLU of the pooled value.
// CTA = (block of 32 ci, part of the rows): the block's weights (64 co x 32 ci x 5 = 40 KB) stay in shared memory for the whole
// launch, each row's g (64 co x P with a zero halo of 2 + 4) is staged once; lane = ci, a warp takes 4 consecutive p at a time:
// per co it reads 5 weights (stride 5: conflict-free) and 8 broadcast g values for 20 FMAs.  (One thread per output reading
// weights and g from global memory per FMA: 0.41 ms of the 256 x 2560 step.)
constexpr int ED_CI = 32;
constexpr int ED_MAXP = 128;                                  // longer pooled rows: chunks of ED_MAXP positions
constexpr int ED_GP = ED_MAXP + 8;                            // g row pitch: 2 halo words left, >= 6 right, multiple of 4
constexpr int ED_SMEM = (64 * ED_CI * 5 + 64 * ED_GP) * 4;
__global__ void __launch_bounds__(256)
expand_dgrad_kernel(const float *__restrict__ g /*[N][64][P]*/, const float *__restrict__ We /*(64,512,5)*/,
                    const float *__restrict__ pooled, int N, int P, float *__restrict__ gpool /*[N][512][P]*/) {
    extern __shared__ __align__(16) float ed_smem[];
    float *s_w = ed_smem;                                     // [64 co][32 ci][5]
    float *s_g = ed_smem + 64 * ED_CI * 5;                    // [64 co][ED_GP]: s_g[co][2 + j] = g[co][p0 + j]
    const int ci0 = blockIdx.x * ED_CI;
    const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
    for (int i = threadIdx.x; i < 64 * ED_CI * 5; i += 256) {
        const int co = i / (ED_CI * 5), r = i - co * (ED_CI * 5);
        s_w[i] = __ldg(We + ((size_t)co * 512 + ci0) * 5 + r);
    }
    for (int n = blockIdx.y; n < N; n += gridDim.y) {
        for (int p0 = 0; p0 < P; p0 += ED_MAXP) {
            const int np = min(ED_MAXP, P - p0);
            __syncthreads();
            for (int i = threadIdx.x; i < 64 * ED_GP; i += 256) {
                const int co = i / ED_GP, j = i - co * ED_GP;
                const int q = p0 + j - 2;
                s_g[i] = (q >= 0 && q < P) ? __ldg(g + ((size_t)n * 64 + co) * P + q) : 0.0f;
            }
            __syncthreads();
            for (int pq = warp; 4 * pq < np; pq += 8) {
                float acc[4] = {0.f, 0.f, 0.f, 0.f};
                const float *wr = s_w + lane * 5;
                const float *gr = s_g + 4 * pq;               // gr[k] = g[p0 + 4 pq + k - 2]
#pragma unroll 4
                for (int co = 0; co < 64; ++co) {
                    const float4 ga = *reinterpret_cast<const float4 *>(gr + co * ED_GP), gb = *reinterpret_cast<const float4 *>(gr + co * ED_GP + 4);
                    const float gv[8] = {ga.x, ga.y, ga.z, ga.w, gb.x, gb.y, gb.z, gb.w};
#pragma unroll
                    for (int t = 0; t < 5; ++t) {
                        const float w = wr[co * (ED_CI * 5) + t];
#pragma unroll
                        for (int k = 0; k < 4; ++k) acc[k] = fmaf(w, gv[k + 4 - t], acc[k]);      // g[p - t + 2]
                    }
                }
                // pooled = lrelu(max + bias): the max-pool routes the gradient to the arg-max position, scaled by lrelu'
                const size_t o = ((size_t)n * 512 + ci0 + lane) * P + p0 + 4 * pq;
#pragma unroll
                for (int k = 0; k < 4; ++k)
                    if (4 * pq + k < np) gpool[o + k] = acc[k] * ((__ldg(pooled + o + k) > 0.f) ? 1.f : kLeakySlope);
            }
        }
    }
}

// contract conv (64 -> 512, k5, pad 2) behind the max-pool: the gradient of c[n][co][.] is gpool[n][co][p] at l* = argmax only.
// wgrad: dWc[co][ci][t] = sum_{n,p} gpool[n][co][p] x0r[n][ci][l* + t - 2]  (dbc[co] = sum gpool: bias_grad_ncl_kernel).
// Every arg-max of pooling window p lies inside that window, so all the x0r values its 512 channels gather come from one
// 64 x 84 tile of row n (80 positions + 2 either side, zero outside the row).  CTA = (block of 64 co, part of the (n, p) tiles):
// the tile is staged in shared memory once (C4L4 order kept: float4 per (c4, position), pitch 85 -> the 8 c4 groups of a warp hit
// disjoint banks) by 16-byte cp.async through a 4-stage ring (three tiles in flight behind the current one: the work on a
// tile is much shorter than a trip to L2 / HBM); thread = (ci, one of 8 co groups) keeps 8 x 5 accumulators and reads, per entry, one broadcast (column, g) pair and the 5 tap values of its ci.
// (One CTA per co gathering from global memory per entry: 6.4 GB of L2 sector reads, 0.79 ms of the 256 x 2560 step.)
// Fixed tile order per part; partial [part][(512,64,5)], added in part order by reduce_parts_kernel.
constexpr int CWG_COB = 64;
constexpr int CWG_AP = 85;
constexpr int CWG_THREADS = 512;
constexpr int CWG_K = CWG_COB / (CWG_THREADS / 64);                   // entries per thread and tile (8)
constexpr int CWG_STAGES = 4;                                         // tiles in the cp.async ring (3 in flight behind the current one)
constexpr int CWG_STAGE_BYTES = 16 * CWG_AP * 16 + 2 * CWG_COB * 4;   // tile + the block's (arg-max, g) entries
constexpr int CWG_SMEM = CWG_STAGES * CWG_STAGE_BYTES;
__global__ void __launch_bounds__(CWG_THREADS, 1)
contract_wgrad_kernel(const float *__restrict__ gpool, const int *__restrict__ argmax, const float4 *__restrict__ x0r /*C4L4*/,
                      int N, int L, int P, float *__restrict__ partial) {
    extern __shared__ __align__(16) uint8_t cw_raw[];
    const uint32_t sbase = (uint32_t)__cvta_generic_to_shared(cw_raw);
    const int co0 = blockIdx.x * CWG_COB;
    const int ci = threadIdx.x & 63, cg = threadIdx.x >> 6;           // co = co0 + cg + 8 k
    float acc[CWG_K][5];
#pragma unroll
    for (int k = 0; k < CWG_K; ++k)
#pragma unroll
        for (int t = 0; t < 5; ++t) acc[k][t] = 0.0f;
    constexpr int PRE = (16 * 84 + CWG_THREADS - 1) / CWG_THREADS;    // 3 16-byte copies of a tile per thread
    const int total = N * P;
    // this thread's share of a tile copy does not depend on the tile: (c4, column) of its PRE chunks, as source / destination offsets
    int src_off[PRE], col_of[PRE];
    uint32_t dst_off[PRE];
#pragma unroll
    for (int k = 0; k < PRE; ++k) {
        const int i = threadIdx.x + k * CWG_THREADS;
        const int c4 = i / 84, c = i - c4 * 84;
        col_of[k] = (i < 16 * 84) ? c - 2 : (1 << 30);                  // position relative to l0; beyond the tile: never copied
        src_off[k] = c4 * L;
        dst_off[k] = (uint32_t)(c4 * CWG_AP + c) * 16u;
    }
    const int step_n = gridDim.y / P, step_w = gridDim.y - step_n * P;  // tile cursor increments (tile = n * P + w)
    int in_ = blockIdx.y / P, iw = blockIdx.y - in_ * P, itl = blockIdx.y;     // cursor of the copies
    // one cp.async group per tile (empty past the end): x0r positions l0 - 2 .. l0 + 81 of the 16 c4 slabs, zero-filled outside the row
    auto issue = [&](int stage) {
        if (itl < total) {
            const int l0 = iw * 80;
            const uint32_t dst0 = sbase + stage * CWG_STAGE_BYTES;
            const float4 *row = x0r + (size_t)in_ * 16 * L;
#pragma unroll
            for (int k = 0; k < PRE; ++k) {
                const int pos = l0 + col_of[k];
                if (col_of[k] < 84) {
                    const bool ok = (pos >= 0 && pos < L);
                    asm volatile("cp.async.cg.shared.global [%0], [%1], 16, %2;" ::"r"(dst0 + dst_off[k]), "l"(row + src_off[k] + (ok ? pos : 0)),
                                 "r"(ok ? 16u : 0u) : "memory");
                }
            }
            if (threadIdx.x < CWG_COB) {
                const size_t e = ((size_t)in_ * 512 + co0 + threadIdx.x) * P + iw;
                const uint32_t d = dst0 + 16 * CWG_AP * 16 + threadIdx.x * 4;
                asm volatile("cp.async.ca.shared.global [%0], [%1], 4;" ::"r"(d), "l"(argmax + e) : "memory");
                asm volatile("cp.async.ca.shared.global [%0], [%1], 4;" ::"r"(d + CWG_COB * 4), "l"(gpool + e) : "memory");
            }
        }
        asm volatile("cp.async.commit_group;" ::: "memory");
        itl += gridDim.y; in_ += step_n; iw += step_w;
        if (iw >= P) { iw -= P; ++in_; }
    };
#pragma unroll
    for (int s = 0; s < CWG_STAGES - 1; ++s) issue(s);
    int stage = 0, w = blockIdx.y % P;                                  // cursor of the work
    for (int tl = blockIdx.y; tl < total; tl += gridDim.y) {
        asm volatile("cp.async.wait_group %0;" ::"n"(CWG_STAGES - 2) : "memory");
        __syncthreads();                  // this tile has landed for everyone; everyone is done with the previous one, whose stage is refilled now
        issue(stage == 0 ? CWG_STAGES - 1 : stage - 1);
        const uint8_t *st = cw_raw + stage * CWG_STAGE_BYTES;
        const float *xr = reinterpret_cast<const float *>(st) + ((ci >> 2) * CWG_AP) * 4 + (ci & 3);
        const int *s_l = reinterpret_cast<const int *>(st + 16 * CWG_AP * 16);
        const float *s_g = reinterpret_cast<const float *>(s_l + CWG_COB);
        const int l0 = w * 80;
#pragma unroll
        for (int k = 0; k < CWG_K; ++k) {
            const int kk = cg + (CWG_THREADS / 64) * k;
            const float g = s_g[kk];
            const float *x = xr + 4 * (s_l[kk] - l0);                    // x[4 t] = x0r[ci][l* + t - 2]
#pragma unroll
            for (int t = 0; t < 5; ++t) acc[k][t] = fmaf(g, x[4 * t], acc[k][t]);
        }
        stage = (stage + 1 == CWG_STAGES) ? 0 : stage + 1;
        w += step_w;
        if (w >= P) w -= P;
    }
    asm volatile("cp.async.wait_group 0;" ::: "memory");
    float *dW = partial + (size_t)blockIdx.y * (512 * 64 * 5);
#pragma unroll
    for (int k = 0; k < CWG_K; ++k)
#pragma unroll
        for (int t = 0; t < 5; ++t) dW[((size_t)(co0 + cg + (CWG_THREADS / 64) * k) * 64 + ci) * 5 + t] = acc[k][t];
}
// dgrad (deterministic): g_x0r[n][ci][l] += sum over (co, p) with |l - l*| <= 2 of Wc[co][ci][l - l* + 2] gpool[n][co][p].
// One CTA per (pooling window = 80-position tile, group of 8 rows): an entry (n, co, p) scatters a 64 x 5 patch g * Wc[co] around
// its arg-max, so the rows of a group SHARE every channel's weights - the 512 channels pass through shared memory in 16 blocks of
// 32 (40 KB, read from L2 once per CTA, the next block prefetched into registers behind the current one's work).  Thread = (ci,
// row) keeps its row's 64 x 84 output tile in shared memory (accs[row][ci][85]: lanes = consecutive ci, stride 85 -> conflict-free;
// 2 halo columns either side) and walks the channels in order.  The window's own entry is branch-free: its 5 taps always land
// inside the 84 columns (what falls into the halo belongs to the neighbouring tile's CTA and is dropped here).  The neighbouring
// windows' entries whose arg-max lies within 2 positions of the tile (~2.5 % of them) are compacted in channel order, per row,
// by the warp that stages that row's entries (ballot + prefix: no atomics), and applied after the block's own entries.
// One accumulation order per output, no atomics.
// History at 256 x 2560: every thread scanning all candidates 3.1 ms; compacted entries, one CTA per (row, tile), weights fetched
// per entry from L2 (5.4 GB of 20-byte reads per step) 2.0 ms; a weight-stationary variant with 4 windows per CTA 2.7 ms; this
// layout with the three windows' entries tested inline (six divergence-managed branches per channel) 1.46 ms.
constexpr int CD_T = 80;
constexpr int CD_R = 8;                        // rows per CTA
constexpr int CD_CO = 32;                      // channels per staged block
constexpr int CD_AP = CD_T + 5;                // accumulator pitch: columns = positions l0 - 2 .. l0 + 81, + 1 pad
constexpr int CD_THREADS = 64 * CD_R;
constexpr int CD_SMEM = (CD_R * 64 * CD_AP + CD_CO * 320 + 2 * CD_R * CD_CO + 4 * CD_R * 2 * CD_CO + 16) * 4;
__global__ void __launch_bounds__(CD_THREADS, 1)
contract_dgrad_kernel(const float *__restrict__ gpool, const int *__restrict__ argmax, const float *__restrict__ Wc /*(512,64,5)*/,
                      int N, int L, int P, int tiles, float *__restrict__ gx /*C4L4, accumulated into*/,
                      const float4 *__restrict__ relu_src /*C4L4 or NULL: gx = (gx + ...) * (relu_src > 0), conv1's ReLU backward*/) {
    extern __shared__ __align__(16) uint8_t cd_raw[];
    float *accs = reinterpret_cast<float *>(cd_raw);                       // [8 rows][64 ci][CD_AP]
    float *s_w = accs + CD_R * 64 * CD_AP;                                 // [32 co][64 ci][5]
    int2 *s_e = reinterpret_cast<int2 *>(s_w + CD_CO * 320);               // [8 rows][32 co]: (arg-max - l0, bits of g) of the own window
    int4 *s_edge = reinterpret_cast<int4 *>(s_e + CD_R * CD_CO);           // [8 rows][64]: (arg-max - l0, bits of g, co in block, -)
    int *s_cnt = reinterpret_cast<int *>(s_edge + CD_R * 2 * CD_CO);       // [8 rows] edge entries of the block
    const int group = blockIdx.x / tiles, tile = blockIdx.x - group * tiles;
    const int n0 = group * CD_R, l0 = tile * CD_T;
    const int ci = threadIdx.x & 63, row = threadIdx.x >> 6;               // compute role (row is warp-uniform)
    const int lane = threadIdx.x & 31;
    const int sr = threadIdx.x >> 5, sn = n0 + sr;                         // staging role of warps 0-7: row sr, lane = co of the block
    const bool stager = threadIdx.x < 32 * CD_R;
    for (int i = threadIdx.x; i < CD_R * 64 * CD_AP; i += CD_THREADS) accs[i] = 0.0f;
    float *mine = accs + ((size_t)row * 64 + ci) * CD_AP;
    constexpr int W4 = CD_CO * 320 / 4 / CD_THREADS;                       // float4 of a weight block per thread (5)
    static_assert(W4 * CD_THREADS * 4 == CD_CO * 320, "weight block / thread count");
    float4 wreg[W4];
    int pl[3] = {0, 0, 0};
    float pg[3] = {0.f, 0.f, 0.f};
    auto prefetch = [&](int c0) {
#pragma unroll
        for (int k = 0; k < W4; ++k) wreg[k] = __ldg(reinterpret_cast<const float4 *>(Wc + (size_t)c0 * 320) + threadIdx.x + k * CD_THREADS);
        if (stager && sn < N) {
            const size_t base = ((size_t)sn * 512 + c0 + lane) * P;
#pragma unroll
            for (int d = 0; d < 3; ++d) {
                const int w = tile - 1 + d;
                const bool ok = (w >= 0 && w < P);
                pl[d] = ok ? __ldg(argmax + base + w) : (1 << 28);
                pg[d] = ok ? __ldg(gpool + base + w) : 0.0f;
            }
        }
    };
    prefetch(0);
    for (int c0 = 0; c0 < 512; c0 += CD_CO) {
        __syncthreads();                                                    // the previous block's weights / entries are consumed
#pragma unroll
        for (int k = 0; k < W4; ++k) reinterpret_cast<float4 *>(s_w)[threadIdx.x + k * CD_THREADS] = wreg[k];
        if (stager) {
            int cnt = 0;
            const bool have = (sn < N);
            // own window: always inside the tile (col 0..79); absent row / window: a zero entry
            const bool own = have && tile < P;
            s_e[sr * CD_CO + lane] = make_int2(own ? pl[1] - l0 : 0, own ? __float_as_int(pg[1]) : 0);
#pragma unroll
            for (int d = 0; d < 3; d += 2) {
                const int rel = pl[d] - l0;                                 // absent: far away
                const bool ok = have && (d == 0 ? (rel >= -2 && rel < 0) : (rel >= CD_T && rel < CD_T + 2));
                const unsigned bal = __ballot_sync(0xffffffffu, ok);
                if (ok) s_edge[sr * 2 * CD_CO + cnt + __popc(bal & ((1u << lane) - 1u))] = make_int4(rel, __float_as_int(pg[d]), lane, 0);
                cnt += __popc(bal);
            }
            if (lane == 0) s_cnt[sr] = cnt;
        }
        __syncthreads();
        if (c0 + CD_CO < 512) prefetch(c0 + CD_CO);                         // in flight behind this block's work
#pragma unroll 4
        for (int co = 0; co < CD_CO; ++co) {
            const float *wr = s_w + ((size_t)co * 64 + ci) * 5;
            const int2 e = s_e[row * CD_CO + co];
            const float g = __int_as_float(e.y);
            float *m = mine + e.x;                                          // columns e.x .. e.x + 4  <->  positions l* - 2 .. l* + 2
            const float a0 = m[0], a1 = m[1], a2 = m[2], a3 = m[3], a4 = m[4];
            m[0] = fmaf(wr[0], g, a0); m[1] = fmaf(wr[1], g, a1); m[2] = fmaf(wr[2], g, a2);
            m[3] = fmaf(wr[3], g, a3); m[4] = fmaf(wr[4], g, a4);
        }
        const int ne = s_cnt[row];
        for (int k = 0; k < ne; ++k) {
            const int4 e = s_edge[row * 2 * CD_CO + k];
            const float *wr = s_w + ((size_t)e.z * 64 + ci) * 5;
            const float g = __int_as_float(e.y);
#pragma unroll
            for (int t = 0; t < 5; ++t) {
                const int c = e.x + t;
                if (c >= 0 && c < CD_T + 4) mine[c] = fmaf(wr[t], g, mine[c]);
            }
        }
    }
    __syncthreads();
    // tile position j <-> column j + 2; four independent read-modify-writes in flight per thread
    constexpr int WO = CD_R * 16 * CD_T;
    for (int i0 = threadIdx.x; i0 < WO; i0 += 4 * CD_THREADS) {
        float4 *d[4];
        float4 v[4];
#pragma unroll
        for (int u = 0; u < 4; ++u) {
            const int i = i0 + u * CD_THREADS;
            d[u] = nullptr;
            if (i < WO) {
                const int r = i / (16 * CD_T), rem = i - r * (16 * CD_T);
                const int c4 = rem / CD_T, j = rem - c4 * CD_T;
                const int l = l0 + j, n = n0 + r;
                if (l < L && n < N) {
                    d[u] = reinterpret_cast<float4 *>(gx) + ((size_t)n * 16 + c4) * L + l;
                    const float *a = accs + ((size_t)r * 64 + 4 * c4) * CD_AP + j + 2;
                    v[u] = make_float4(a[0], a[CD_AP], a[2 * CD_AP], a[3 * CD_AP]);
                }
            }
        }
        float4 o[4], a[4];
#pragma unroll
        for (int u = 0; u < 4; ++u) {
            if (d[u]) o[u] = *d[u];
            if (d[u] && relu_src) a[u] = __ldg(relu_src + (d[u] - reinterpret_cast<float4 *>(gx)));
        }
#pragma unroll
        for (int u = 0; u < 4; ++u)
            if (d[u]) {
                o[u].x += v[u].x; o[u].y += v[u].y; o[u].z += v[u].z; o[u].w += v[u].w;
                if (relu_src) {
                    o[u].x = a[u].x > 0.f ? o[u].x : 0.f; o[u].y = a[u].y > 0.f ? o[u].y : 0.f;
                    o[u].z = a[u].z > 0.f ? o[u].z : 0.f; o[u].w = a[u].w > 0.f ? o[u].w : 0.f;
                }
                *d[u] = o[u];
            }
    }
}

// ------------------------------------------------------------------------------------------------ conv1 backward
// x0r = relu(conv1 x), g already masked by relu':  dW1[c][t] = sum_{n,l} g[n][c][l] x[n][l + t - 4];  db1[c] = sum g.
// Persistent CTAs over (row, 1024-position) tiles; thread = (c4, lane16): 4 channels x 9 taps + 4 bias accumulators over the
// positions l = lane16 mod 16; block tree over the 16 lanes; one partial [64*9 + 64] per CTA.
constexpr int C1_CHUNK = 1024;
constexpr int C1_LEN = 64 * 9 + 64;
__global__ void __launch_bounds__(256)
first_wgrad_kernel(const float4 *__restrict__ g /*C4L4*/, const float *__restrict__ x, int N, int L, int chunks,
                   float *__restrict__ partial /*[gridDim.x][C1_LEN]*/) {
    __shared__ float red[16][16][40];
    const int c4 = threadIdx.x >> 4, ln = threadIdx.x & 15;
    float acc[4][9], bs[4];
#pragma unroll
    for (int i = 0; i < 4; ++i) { bs[i] = 0.f;
#pragma unroll
        for (int t = 0; t < 9; ++t) acc[i][t] = 0.f; }
    const int total = N * chunks;
    for (int tile = blockIdx.x; tile < total; tile += gridDim.x) {
        const int n = tile / chunks, l0 = (tile - n * chunks) * C1_CHUNK;
        const int l1 = min(L, l0 + C1_CHUNK);
        const float *xr = x + (size_t)n * L;
        for (int l = l0 + ln; l < l1; l += 16) {
            const float4 gv = __ldg(g + ((size_t)n * 16 + c4) * L + l);
            const float ga[4] = {gv.x, gv.y, gv.z, gv.w};
#pragma unroll
            for (int t = 0; t < 9; ++t) {
                const int q = l + t - 4;
                const float xv = (q >= 0 && q < L) ? __ldg(xr + q) : 0.0f;
#pragma unroll
                for (int i = 0; i < 4; ++i) acc[i][t] = fmaf(ga[i], xv, acc[i][t]);
            }
#pragma unroll
            for (int i = 0; i < 4; ++i) bs[i] += ga[i];
        }
    }
#pragma unroll
    for (int i = 0; i < 4; ++i) {
#pragma unroll
        for (int t = 0; t < 9; ++t) red[c4][ln][i * 9 + t] = acc[i][t];
        red[c4][ln][36 + i] = bs[i];
    }
    __syncthreads();
    for (int idx = threadIdx.x; idx < 16 * 40; idx += 256) {
        const int cc = idx / 40, k = idx - cc * 40;
        float s = 0.0f;
        for (int j = 0; j < 16; ++j) s += red[cc][j][k];
        float *o = partial + (size_t)blockIdx.x * C1_LEN;
        if (k < 36) o[(4 * cc + k / 9) * 9 + (k % 9)] = s;          // (64, 1, 9) layout
        else o[64 * 9 + 4 * cc + (k - 36)] = s;
    }
}

// ------------------------------------------------------------------------------------------------ AdamW
__global__ void __launch_bounds__(256)
adamw_kernel(float *__restrict__ p, const float *__restrict__ g, float *__restrict__ m, float *__restrict__ v, size_t n,
             float lr, float b1, float b2, float eps, float wd, float bc1, float bc2_sqrt) {
    for (size_t i = (size_t)blockIdx.x * 256 + threadIdx.x; i < n; i += (size_t)gridDim.x * 256) {
        const float gi = g[i];
        float pi = p[i] * (1.0f - lr * wd);                       // decoupled weight decay
        const float mi = m[i] + (1.0f - b1) * (gi - m[i]);        // torch: exp_avg.lerp_(grad, 1 - beta1)
        const float vi = b2 * v[i] + (1.0f - b2) * gi * gi;
        m[i] = mi; v[i] = vi;
        const float denom = sqrtf(vi) / bc2_sqrt + eps;
        pi -= (lr / bc1) * (mi / denom);
        p[i] = pi;
    }
}

// the same with the step-dependent scalars in device memory: h = {lr, beta1, beta2, eps, weight_decay, 1 - beta1^step, sqrt(1 - beta2^step)}
__global__ void __launch_bounds__(256)
adamw_dev_kernel(float *__restrict__ p, const float *__restrict__ g, float *__restrict__ m, float *__restrict__ v, size_t n,
                 const float *__restrict__ h) {
    const float lr = __ldg(h), b1 = __ldg(h + 1), b2 = __ldg(h + 2), eps = __ldg(h + 3), wd = __ldg(h + 4), bc1 = __ldg(h + 5),
                bc2_sqrt = __ldg(h + 6);
    for (size_t i = (size_t)blockIdx.x * 256 + threadIdx.x; i < n; i += (size_t)gridDim.x * 256) {
        const float gi = g[i];
        float pi = p[i] * (1.0f - lr * wd);
        const float mi = m[i] + (1.0f - b1) * (gi - m[i]);
        const float vi = b2 * v[i] + (1.0f - b2) * gi * gi;
        m[i] = mi; v[i] = vi;
        const float denom = sqrtf(vi) / bc2_sqrt + eps;
        pi -= (lr / bc1) * (mi / denom);
        p[i] = pi;
    }
}

}  // namespace stof

using namespace stof;

// ------------------------------------------------------------------------------------------------ tensor-core operand packs
// One tap of a 64 -> 64 conv for the chain kernel: [c8 = k/8][W_hi 64 rows | W_lo 64 rows][k%8] fp16 (upload_tc in stof_api.cu).
// Built here on the device from mid[j].w = [ci][t][co] fp32 (kept current by set_weights / set_weights_device):
//   forward   tap t : k = ci, row = co              <- W[co][ci][t]
//   backward  tap t': k = co, row = ci              <- W[co][ci][6 - t']     (dX = conv of dY with swapped channels, flipped taps)
struct TrainPackJobs {
    const float *w[STOF_MAX_MID_CONVS];
    __half *fwd[STOF_MAX_MID_CONVS], *bwd[STOF_MAX_MID_CONVS];
};

__global__ void __launch_bounds__(256) train_pack_kernel(const TrainPackJobs jobs) {
    const int j = blockIdx.y;
    const float *w = jobs.w[j];
    const int i = blockIdx.x * 256 + threadIdx.x;            // (t, k, row) with row fastest: 7 * 64 * 64
    if (i >= 7 * 64 * 64) return;
    const int row = i & 63, k = (i >> 6) & 63, t = i >> 12;
    {
        const float v = __ldg(w + ((size_t)k * 7 + t) * 64 + row);                  // W[co = row][ci = k][t]
        const float c = fminf(fmaxf(v, -65504.0f), 65504.0f);
        const __half hi = __float2half_rn(c);
        const __half lo = __float2half_rn((c - __half2float(hi)) * 2048.0f);
        __half *tap = jobs.fwd[j] + (size_t)t * 8 * 128 * 8;
        tap[((size_t)(k >> 3) * 128 + row) * 8 + (k & 7)] = hi;
        tap[((size_t)(k >> 3) * 128 + 64 + row) * 8 + (k & 7)] = lo;
    }
    {
        const float v = __ldg(w + ((size_t)row * 7 + (6 - t)) * 64 + k);            // W[co = k][ci = row][6 - t]
        const float c = fminf(fmaxf(v, -65504.0f), 65504.0f);
        const __half hi = __float2half_rn(c);
        const __half lo = __float2half_rn((c - __half2float(hi)) * 2048.0f);
        __half *tap = jobs.bwd[j] + (size_t)t * 8 * 128 * 8;
        tap[((size_t)(k >> 3) * 128 + row) * 8 + (k & 7)] = hi;
        tap[((size_t)(k >> 3) * 128 + 64 + row) * 8 + (k & 7)] = lo;
    }
}

// semi-global block packs (the layouts of ensure_tc in stof_api.cu) from the PyTorch-layout device copies:
//   contract (512, 64, 5) -> [4 slices][5 taps][8 c8][W_hi 128 | W_lo 128][8];   expand (64, 512, 5) -> [8 groups][5 taps][8 c8][W_hi 64 | W_lo 64][8]
__global__ void __launch_bounds__(256)
train_pack_sgb_kernel(const float *__restrict__ contract_w, const float *__restrict__ expand_w, __half *__restrict__ cpk, __half *__restrict__ epk) {
    const int i = blockIdx.x * 256 + threadIdx.x;
    if (i >= 512 * 64 * 5) return;
    const float v = __ldg((blockIdx.y ? expand_w : contract_w) + i);
    const float c = fminf(fmaxf(v, -65504.0f), 65504.0f);
    const __half hi = __float2half_rn(c);
    const __half lo = __float2half_rn((c - __half2float(hi)) * 2048.0f);
    const int t = i % 5;
    if (blockIdx.y == 0) {
        const int ci = (i / 5) & 63, co = i / (5 * 64);
        const int sl = co >> 7, r = co & 127;
        __half *tap = cpk + ((size_t)sl * 5 + t) * 8 * 256 * 8;
        tap[((size_t)(ci >> 3) * 256 + r) * 8 + (ci & 7)] = hi;
        tap[((size_t)(ci >> 3) * 256 + 128 + r) * 8 + (ci & 7)] = lo;
    } else {
        const int cin = (i / 5) & 511, co = i / (5 * 512);
        const int g = cin >> 6, ci = cin & 63;
        __half *tap = epk + ((size_t)g * 5 + t) * 8 * 128 * 8;
        tap[((size_t)(ci >> 3) * 128 + co) * 8 + (ci & 7)] = hi;
        tap[((size_t)(ci >> 3) * 128 + 64 + co) * 8 + (ci & 7)] = lo;
    }
}

constexpr size_t TR_PER_LAYER = (size_t)7 * 8 * 128 * 8;       // halfs of one mid conv pack
constexpr size_t TR_SGB_PACK = (size_t)512 * 64 * 5 * 2;       // halfs of the contract (and of the expand) pack
static __half *train_contract_pack(const stof_model *m) { return reinterpret_cast<__half *>(m->tr_tc_w) + 2 * (size_t)m->n_mid * TR_PER_LAYER; }
static __half *train_expand_pack(const stof_model *m) { return train_contract_pack(m) + TR_SGB_PACK; }

// (re)build the training packs on `st` when the weights changed since the last build
// The mid convs of the training step: the weight-stationary one-layer kernel (conv1l_tc.cu), or - STOF_TRAIN_CONV1L=0, kept for
// A/B measurements - one-layer launches of the inference chain kernel; bit-identical results
static bool train_conv1l() {
    static const bool on = [] { const char *e = getenv("STOF_TRAIN_CONV1L"); return !(e != nullptr && e[0] == '0'); }();
    return on;
}

static int ensure_train_tc(stof_model *m, cudaStream_t st) {
    const size_t per_layer = TR_PER_LAYER;
    if (!m->tr_tc_w) STOF_CUDA(cudaMalloc(&m->tr_tc_w, (2 * (size_t)m->n_mid * per_layer + 2 * TR_SGB_PACK) * sizeof(__half)));
    if (!m->tr_side) {
        STOF_CUDA(cudaStreamCreateWithFlags(&m->tr_side, cudaStreamNonBlocking));
        for (cudaEvent_t &e : m->tr_ev) STOF_CUDA(cudaEventCreateWithFlags(&e, cudaEventDisableTiming));
    }
    if (!m->tr_tc_dirty) return STOF_OK;
    if (m->sgb_scale != 1) {
        train_pack_sgb_kernel<<<dim3((512 * 64 * 5 + 255) / 256, 2), 256, 0, st>>>(m->raw_contract_w, m->raw_expand_w, train_contract_pack(m),
                                                                                  train_expand_pack(m));
        STOF_LAUNCH_CHECK();
    }
    TrainPackJobs jobs;
    __half *base = reinterpret_cast<__half *>(m->tr_tc_w);
    for (int j = 0; j < m->n_mid; ++j) {
        jobs.w[j] = m->mid[j].w;
        jobs.fwd[j] = base + (size_t)j * per_layer;
        jobs.bwd[j] = base + (size_t)(m->n_mid + j) * per_layer;
        for (int dir = 0; dir < 2; ++dir) {
            TcLayer &l = dir ? m->tr_bwd[j] : m->tr_fwd[j];
            memset(&l, 0, sizeof(l));
            l.w = dir ? jobs.bwd[j] : jobs.fwd[j];
            l.bias = dir ? m->zero_bias : m->mid[j].b;
            l.taps = 7; l.pad = 3;
            l.stage_bytes = 16384; l.n_out = 64; l.n_stack = 128; l.n_lo = 64; l.lo_row = 0; l.lo_col = 64; l.w_rows = 128; l.n_fast = 64;
            // forward (models/stofnet.py:51-62): even index -> LeakyReLU; odd -> skip + conv with the skip read from global
            // planes (one layer per launch: every output is kept for the backward pass).  backward: set per call.
            l.type = dir ? TC_NONE : (((j & 1) || j == m->n_mid - 1) ? TC_LONGSKIP : TC_LEAKY);
        }
    }
    dim3 grid((7 * 64 * 64 + 255) / 256, (unsigned)m->n_mid);
    train_pack_kernel<<<grid, 256, 0, st>>>(jobs);
    STOF_LAUNCH_CHECK();
    m->tr_tc_dirty = false;
    return STOF_OK;
}

extern "C" int stof_train_workspace_bytes(const stof_model *m, int N, int L, size_t *bytes) {
    STOF_REQUIRE(bytes != nullptr, STOF_E_INVALID, "stof_train_workspace_bytes: bytes is NULL");
    TrainPlan pl;
    int rc = make_train_plan(m, N, L, &pl);
    if (rc) return rc;
    *bytes = pl.total_bytes;
    return STOF_OK;
}

extern "C" int stof_forward_train(stof_model *m, const float *x, int N, int L, float *mask, void *ws_dev, size_t ws_bytes,
                                  void *stream) {
    TrainPlan pl;
    int rc = make_train_plan(m, N, L, &pl);
    if (rc) return rc;
    STOF_REQUIRE(x && mask && ws_dev, STOF_E_INVALID, "stof_forward_train: NULL buffer");
    STOF_REQUIRE(ws_bytes >= pl.total_bytes, STOF_E_WORKSPACE, "stof_forward_train: workspace %zu < %zu bytes", ws_bytes, pl.total_bytes);
    STOF_REQUIRE((reinterpret_cast<uintptr_t>(ws_dev) & 255) == 0 && (reinterpret_cast<uintptr_t>(mask) & 15) == 0, STOF_E_INVALID,
                 "stof_forward_train: workspace must be 256-byte and mask 16-byte aligned");
    cudaStream_t st = (cudaStream_t)stream;
    float *ws = reinterpret_cast<float *>(ws_dev);
    float *x0r = ws + pl.off_x0r, *x0 = ws + pl.off_x0;
    // models/stofnet.py:45
    if ((rc = launch_conv_first(x, N, L, m->first.w, m->first.b, x0r, 9, 0, st))) return rc;
    if (pl.tc && pl.sgb) {
        // :98-117 on the tensor cores (contract_tc.cu with the arg-max output the backward pass needs, expand_tc.cu); x0 comes
        // out of the fused conv1 + upsample-add kernel directly as split planes.  x0r (fp32) stays for relu' and the contract wgrad.
        if ((rc = ensure_train_tc(m, st))) return rc;
        float *pooled = ws + pl.off_pooled, *e = ws + pl.off_e;
        int *argmax = reinterpret_cast<int *>(ws + pl.off_argmax);
        if ((rc = launch_contract_pool_tc(x, train_contract_pack(m), m->contract.b, m->first.w, m->first.b, pooled, N, L, pl.P, st, 0, argmax))) return rc;
        if ((rc = launch_expand_tc(pooled, e, train_expand_pack(m), m->expand.b, N, pl.P, st))) return rc;
    } else if (pl.sgb) {
        // :98-117
        float *pooled = ws + pl.off_pooled, *e = ws + pl.off_e;
        int *argmax = reinterpret_cast<int *>(ws + pl.off_argmax);
        if ((rc = launch_contract_pool_simt(x0r, pooled, m->contract.w, m->contract.b, N, L, pl.P, st, argmax))) return rc;
        if ((rc = launch_expand(pooled, e, m->expand.w, m->expand.b, N, pl.P, st))) return rc;
        STOF_CUDA(cudaMemcpyAsync(x0, x0r, (size_t)N * 64 * L * sizeof(float), cudaMemcpyDeviceToDevice, st));
        if ((rc = launch_sgb_add(x0, e, N, L, pl.P, m->sgb_scale, (L - pl.P * m->sgb_scale) / 2, st))) return rc;
    }
    if (pl.tc) {
        // :51-62 on the tensor cores: one chain launch per conv, planes in / planes out (kept), fp32 C4L4 copy for the backward
        if ((rc = ensure_train_tc(m, st))) return rc;
        void *px0 = ws + pl.off_px0;
        if (pl.sgb) {
            if ((rc = launch_x0_split(x, ws + pl.off_e, N, L, pl.P, m->sgb_scale, (L - pl.P * m->sgb_scale) / 2, m->first.w, m->first.b, px0, st, 0,
                                      m->k_first))) return rc;
        } else if ((rc = launch_c4l4_to_planes(x0, px0, N, L, st))) return rc;
        const void *curp = px0;
        for (int j = 0; j < m->n_mid; ++j) {
            void *outp = ws + pl.off_pa[j];
            const bool longskip = (j == m->n_mid - 1);
            const void *skip = (longskip || j < 2) ? px0 : (const void *)(ws + pl.off_pa[j - 2]);
            const bool add_skip = (j & 1) || longskip;
            if (train_conv1l())
                rc = launch_conv1l_tc(&m->tr_fwd[j], add_skip ? TC_LONGSKIP : TC_LEAKY, curp, add_skip ? skip : nullptr, outp, N, L, st);
            else
                rc = launch_chain_tc(&m->tr_fwd[j], 1, curp, add_skip ? skip : nullptr, outp, nullptr, N, L, st);
            if (rc) return rc;
            // fp32 copy only of the last activation (conv_last forward and weight gradient); lrelu' reads the planes
            if (longskip && (rc = launch_planes_to_c4l4(outp, ws + pl.off_a[j], N, L, st))) return rc;
            curp = outp;
        }
        return launch_conv_last(ws + pl.off_a[m->n_mid - 1], mask, m->last.w, m->last.b, N, L, 64, 0, st);
    }
    // :51-62  conv index i = j + 2
    const float *cur = x0;
    for (int j = 0; j < m->n_mid; ++j) {
        float *out = ws + pl.off_a[j];
        const bool longskip = (j == m->n_mid - 1), res = (j & 1) && !longskip;
        if (longskip) rc = launch_conv_mid_simt(cur, x0, out, m->mid[j].w, m->mid[j].b, N, L, 1, st);
        else if (res) rc = launch_conv_mid_simt(cur, (j >= 2) ? ws + pl.off_a[j - 2] : x0, out, m->mid[j].w, m->mid[j].b, N, L, 1, st);
        else rc = launch_conv_mid_simt(cur, nullptr, out, m->mid[j].w, m->mid[j].b, N, L, 0, st);
        if (rc) return rc;
        cur = out;
    }
    // :65
    return launch_conv_last(cur, mask, m->last.w, m->last.b, N, L, 64, 0, st);
}

static int mid_wgrad(const float *G, const float *A, int N, int L, float *partial, float *dW, float *db, cudaStream_t st) {
    const int tpr = (L + WG_TL - 1) / WG_TL;
    const long long total = (long long)N * tpr;
    const int grid = (int)(total < WG_MAX_PARTS ? total : WG_MAX_PARTS);
    constexpr int smem = 16 * (WG_GP + WG_AP) * sizeof(float4);
    static SmemAttrOnce attr;
    { const int rc = ensure_dyn_smem(mid_wgrad_kernel, (int)smem, attr); if (rc) return rc; }
    mid_wgrad_kernel<<<grid, 256, smem, st>>>(reinterpret_cast<const float4 *>(G), reinterpret_cast<const float4 *>(A), N, L, tpr, partial);
    STOF_LAUNCH_CHECK();
    reduce_parts_kernel<<<(unsigned)((WG_MID_LEN + 255) / 256), 256, 0, st>>>(partial, grid, WG_MID_LEN, dW, 64 * 64 * 7, db);
    STOF_LAUNCH_CHECK();
    return STOF_OK;
}

extern "C" int stof_backward(stof_model *m, const float *x, const float *g_mask, int N, int L, void *ws_dev, size_t ws_bytes,
                             const stof_grads *gr, void *stream) {
    TrainPlan pl;
    int rc = make_train_plan(m, N, L, &pl);
    if (rc) return rc;
    STOF_REQUIRE(x && g_mask && ws_dev && gr, STOF_E_INVALID, "stof_backward: NULL buffer");
    STOF_REQUIRE(ws_bytes >= pl.total_bytes, STOF_E_WORKSPACE, "stof_backward: workspace %zu < %zu bytes", ws_bytes, pl.total_bytes);
    STOF_REQUIRE(gr->conv1_w && gr->conv1_b && gr->last_w && gr->last_b, STOF_E_INVALID, "stof_backward: missing gradient buffer");
    for (int j = 0; j < m->n_mid; ++j) STOF_REQUIRE(gr->mid_w[j] && gr->mid_b[j], STOF_E_INVALID, "stof_backward: conv%d gradient buffer missing", j + 2);
    if (pl.sgb) STOF_REQUIRE(gr->contract_w && gr->contract_b && gr->expand_w && gr->expand_b, STOF_E_INVALID, "stof_backward: semi-global block gradient buffers missing");
    STOF_REQUIRE((reinterpret_cast<uintptr_t>(g_mask) & 15) == 0, STOF_E_INVALID, "stof_backward: g_mask must be 16-byte aligned");
    cudaStream_t st = (cudaStream_t)stream;
    float *ws = reinterpret_cast<float *>(ws_dev);
    float *x0r = ws + pl.off_x0r, *x0 = ws + pl.off_x0;
    float *X = ws + pl.off_g[0], *S = ws + pl.off_g[1], *T = ws + pl.off_g[2];
    float *partial = ws + pl.off_partial;
    const size_t n4 = (size_t)N * 16 * L;
    const float4 *gy = reinterpret_cast<const float4 *>(g_mask);
    auto act = [&](int j) -> const float * { return j < 0 ? x0 : ws + pl.off_a[j]; };     // output of mid conv j (x0 for j = -1)
    const int nm = m->n_mid;

    // ---- conv_last: g_a12 -> X
    {
        dim3 grid((L + 255) / 256, N);
        STOF_REQUIRE(N <= 65535, STOF_E_INVALID, "stof_backward: at most 65535 rows per call");
        last_dgrad_kernel<<<grid, 256, 0, st>>>(gy, m->raw_last_w, N, L, reinterpret_cast<float4 *>(X));
        STOF_LAUNCH_CHECK();
        const int tpr = (L + LW_TL - 1) / LW_TL;
        const long long total = (long long)N * tpr;
        const int g2 = (int)(total < LW_MAX_CTAS ? total : LW_MAX_CTAS);
        last_wgrad_kernel<<<g2, 192, 0, st>>>(gy, reinterpret_cast<const float4 *>(act(nm - 1)), N, L, tpr, partial);
        STOF_LAUNCH_CHECK();
        reduce_parts_kernel<<<4, 256, 0, st>>>(partial, g2, 772, gr->last_w, 768, gr->last_b);
        STOF_LAUNCH_CHECK();
    }
    auto pact = [&](int j) -> const void * { return j < 0 ? (const void *)(ws + pl.off_px0) : (const void *)(ws + pl.off_pa[j]); };
    if (pl.tc) {
        // ---- mid convs, tensor-core path: the gradients stay in split planes from X to the gradient of x0.
        // A planes tensor holds g * 2^E with its GradSlot {max |stored|, E} in device memory.  X is converted once (max |X| ->
        // E, bias partial sums on the way); after that every data gradient is ONE launch of the chain kernel in gradient mode
        // (conv_tc.cu: the tap-flipped channel-swapped pack, then * lrelu' of the activation planes or + the skip gradient's
        // planes in the epilogue, the output's E chosen on the device from the operands' slots, max |out| and the bias
        // gradient's per-CTA sums recorded on the way), and every weight gradient one launch of wgrad_tc on the same planes.
        if ((rc = ensure_train_tc(m, st))) return rc;
        STOF_REQUIRE(nm + 2 <= 31, STOF_E_INVALID, "stof_backward: too many mid convs");
        STOF_CUDA(cudaMemsetAsync(ws + pl.off_absmax, 0, 64 * sizeof(float), st));
        unsigned *raw_amax = reinterpret_cast<unsigned *>(ws + pl.off_absmax);
        GradSlot *slots = reinterpret_cast<GradSlot *>(ws + pl.off_absmax + 2);
        void *buf[3] = {ws + pl.off_gp[0], ws + pl.off_gp[1], ws + pl.off_g[2]};       // planes (the fp32 T buffer is not used on this path)
        float *bpart = ws + pl.off_bpart;
        int n_slot = 0;
        const int cgrid = train_conv1l() ? conv1l_tc_grid(N, L) : chain_tc_grid(1, &m->tr_bwd[0], N, L);
        const size_t bp_stride = ((size_t)kSMs * 64 + 63) / 64 * 64;
        auto bparts = [&](int slot) { return ws + pl.off_bparts + (size_t)(slot - 1) * bp_stride; };     // of the chain launch that wrote `slot`
        // The weight and bias gradients only READ what the data-gradient chain produces, so they run on a side stream next to
        // it (forked and joined with events - also under CUDA-graph capture): at the reference's batch size the backward is a
        // chain of ~50 tiny dependent kernels, and this halves its critical path.
        cudaStream_t side = m->tr_side;
        float *wpart = ws + pl.off_wpart;
        int ev_i = 0;
        cudaEvent_t buf_free[3] = {nullptr, nullptr, nullptr}, last_side = nullptr;      // side-stream work that last read buf[i]
        // dW_j, db_j from g = buf[gi] (gradient wrt the output of conv j, just produced on the main stream with its bias partials)
        auto side_grads = [&](int j, int gi, int gslot, const float *bp, int parts, const GradSlot *bslot) -> int {
            int r;
            cudaEvent_t ready = m->tr_ev[ev_i++], done = m->tr_ev[ev_i++];
            STOF_CUDA(cudaEventRecord(ready, st));
            STOF_CUDA(cudaStreamWaitEvent(side, ready, 0));
            if ((r = launch_wgrad_tc(buf[gi], pact(j - 1), &slots[gslot], N, L, wpart, gr->mid_w[j], side))) return r;
            if ((r = launch_bias_parts_reduce(bp, parts, gr->mid_b[j], side, bslot))) return r;
            STOF_CUDA(cudaEventRecord(done, side));
            buf_free[gi] = done; last_side = done;
            return STOF_OK;
        };
        // main stream: buf[oi] = dgrad_j(buf[gi]) [* lrelu'(aux) | + aux], with the new tensor's slot and bias partial sums
        auto step = [&](int j, int gi, int gslot, int oi, int mode, const void *aux, int aux_slot, bool want_bias) -> int {
            if (buf_free[oi]) STOF_CUDA(cudaStreamWaitEvent(st, buf_free[oi], 0));      // its last reader on the side stream is done
            TcGrad g;
            g.mode = mode; g.aux = aux;
            g.g_in = &slots[gslot]; g.g_aux = aux_slot >= 0 ? &slots[aux_slot] : nullptr;
            g.g_out = &slots[++n_slot];
            g.bias_partial = want_bias ? bparts(n_slot) : nullptr;
            if (train_conv1l()) return launch_conv1l_tc(&m->tr_bwd[j], TC_NONE, buf[gi], nullptr, buf[oi], N, L, st, &g);
            return launch_chain_tc_grad(&m->tr_bwd[j], buf[gi], buf[oi], N, L, st, g);
        };
        // X -> planes
        if ((rc = launch_absmax(X, n4, raw_amax, st))) return rc;
        if ((rc = launch_grad_to_planes(X, buf[0], N, L, st, raw_amax, bpart, &slots[0]))) return rc;
        // conv{nm-1} (long skip): a = x0 + conv(a_prev): g_conv = X;  S = dgrad(X) = gradient wrt a_{nm-2}
        int si = 1, s_slot;
        if ((rc = side_grads(nm - 1, 0, 0, bpart, grad_to_planes_parts(N, L), nullptr))) return rc;
        if ((rc = step(nm - 1, 0, 0, si, 1, nullptr, -1, nm >= 2))) return rc;
        s_slot = n_slot;
        // pairs (RES j, LEAKY j - 1), j = nm-2, nm-4, ..., 1
        for (int j = nm - 2; j >= 1; j -= 2) {
            // a_j = a_{j-2} + conv_j(a_{j-1}):  T = dgrad_j(S) * lrelu'(a_{j-1}) = gradient wrt the output of conv j-1; the skip keeps S
            const int ti = (si + 1) % 3, s2 = (si + 2) % 3;
            if ((rc = side_grads(j, si, s_slot, bparts(s_slot), cgrid, &slots[s_slot]))) return rc;
            if ((rc = step(j, si, s_slot, ti, 3, pact(j - 1), -1, true))) return rc;
            const int t_slot = n_slot;
            // a_{j-1} = lrelu(conv_{j-1}(a_{j-2})):  S' = S + dgrad_{j-1}(T) = gradient wrt a_{j-2} (the output of conv j-2, or x0)
            if ((rc = side_grads(j - 1, ti, t_slot, bparts(t_slot), cgrid, &slots[t_slot]))) return rc;
            if ((rc = step(j - 1, ti, t_slot, s2, 2, buf[si], s_slot, j - 2 >= 0))) return rc;
            si = s2; s_slot = n_slot;
        }
        // S (planes) = gradient wrt x0 through the chain: back to fp32, + the long skip's X
        if ((rc = launch_grad_planes_to_c4l4(buf[si], S, N, L, st, &slots[s_slot], X))) return rc;
        if (last_side) STOF_CUDA(cudaStreamWaitEvent(st, last_side, 0));                // join: everything after sees dW / db
    } else {
        // ---- mid convs, fp32 path: weight gradient with the conv's input (the output of conv j - 1, or x0), then out (+)= data gradient
        auto mid_backward = [&](int j, const float *g, float *out, bool accumulate) -> int {
            int r;
            if ((r = mid_wgrad(g, act(j - 1), N, L, partial, gr->mid_w[j], gr->mid_b[j], st))) return r;
            return launch_conv_mid_simt(g, accumulate ? out : nullptr, out, m->bw_mid[j], m->zero_bias, N, L, accumulate ? 1 : 2, st);
        };
        // conv{nb-1} (long skip): a = x0 + conv(a_prev): g_conv = X;  S = dgrad(X)
        if ((rc = mid_backward(nm - 1, X, S, false))) return rc;
        // pairs (RES j, LEAKY j - 1), j = nm-2, nm-4, ..., 1:   S = g wrt a_j
        for (int j = nm - 2; j >= 1; j -= 2) {
            // a_j = a_{j-2} + conv_j(a_{j-1}):  wgrad with input a_{j-1};  T = dgrad_j(S);  skip keeps S
            if ((rc = mid_backward(j, S, T, false))) return rc;
            // a_{j-1} = lrelu(conv_{j-1}(a_{j-2})):  T *= lrelu'(a_{j-1});  wgrad with input a_{j-2};  S += dgrad_{j-1}(T)
            lrelu_bwd_kernel<<<ew_grid(n4), 256, 0, st>>>(reinterpret_cast<float4 *>(T), reinterpret_cast<const float4 *>(act(j - 1)), n4);
            STOF_LAUNCH_CHECK();
            if ((rc = mid_backward(j - 1, T, S, true))) return rc;
        }
        // S = g wrt x0 through the chain; add the long skip
        add_kernel<<<ew_grid(n4), 256, 0, st>>>(reinterpret_cast<float4 *>(S), reinterpret_cast<const float4 *>(X), n4);
        STOF_LAUNCH_CHECK();
    }
    // ---- semi-global block: x0 = x0r + pad(up(E))
    if (pl.sgb) {
        float *pooled = ws + pl.off_pooled, *e = ws + pl.off_e, *ge = ws + pl.off_ge, *gpool = ws + pl.off_gpool;
        const int *argmax = reinterpret_cast<const int *>(ws + pl.off_argmax);
        const int P = pl.P, off = (L - P * m->sgb_scale) / 2;
        const long long warps = (long long)N * 16 * P;
        window_sum_kernel<<<(unsigned)((warps + 7) / 8), 256, 0, st>>>(reinterpret_cast<const float4 *>(S), N, L, P, m->sgb_scale, off, ge);
        STOF_LAUNCH_CHECK();
        const size_t ne = (size_t)N * 64 * P;
        lrelu_bwd1_kernel<<<ew_grid(ne), 256, 0, st>>>(ge, e, ne);
        STOF_LAUNCH_CHECK();
        const int parts = N < CWG_MAX_PARTS ? N : CWG_MAX_PARTS;
        expand_wgrad_kernel<<<dim3(512 / EW_CB, (unsigned)parts), 256, 0, st>>>(ge, pooled, N, P, partial);
        STOF_LAUNCH_CHECK();
        reduce_parts_kernel<<<(64 * 512 * 5 + 255) / 256, 256, 0, st>>>(partial, parts, (size_t)64 * 512 * 5, gr->expand_w, (size_t)64 * 512 * 5, nullptr);
        STOF_LAUNCH_CHECK();
        bias_grad_ncl_kernel<<<64, 256, 0, st>>>(ge, N, 64, P, gr->expand_b);
        STOF_LAUNCH_CHECK();
        static SmemAttrOnce ed_attr;
        if ((rc = ensure_dyn_smem(expand_dgrad_kernel, ED_SMEM, ed_attr))) return rc;
        expand_dgrad_kernel<<<dim3(512 / ED_CI, (unsigned)(N < 32 ? N : 32)), 256, ED_SMEM, st>>>(ge, m->raw_expand_w, pooled, N, P, gpool);
        STOF_LAUNCH_CHECK();
        const int cparts = (long long)N * P < CWG_MAX_PARTS ? N * P : CWG_MAX_PARTS;
        static SmemAttrOnce cw_attr;
        if ((rc = ensure_dyn_smem(contract_wgrad_kernel, CWG_SMEM, cw_attr))) return rc;
        contract_wgrad_kernel<<<dim3(512 / CWG_COB, (unsigned)cparts), CWG_THREADS, CWG_SMEM, st>>>(gpool, argmax, reinterpret_cast<const float4 *>(x0r), N, L, P, partial);
        STOF_LAUNCH_CHECK();
        reduce_parts_kernel<<<(512 * 64 * 5 + 255) / 256, 256, 0, st>>>(partial, cparts, (size_t)512 * 64 * 5, gr->contract_w, (size_t)512 * 64 * 5, nullptr);
        STOF_LAUNCH_CHECK();
        bias_grad_ncl_kernel<<<512, 256, 0, st>>>(gpool, N, 512, P, gr->contract_b);
        STOF_LAUNCH_CHECK();
        const int tiles = (L + CD_T - 1) / CD_T;
        STOF_REQUIRE((long long)N * tiles < (1ll << 31), STOF_E_INVALID, "stof_backward: batch too large");
        static SmemAttrOnce cd_attr;
        if ((rc = ensure_dyn_smem(contract_dgrad_kernel, CD_SMEM, cd_attr))) return rc;
        contract_dgrad_kernel<<<(unsigned)(((N + CD_R - 1) / CD_R) * tiles), CD_THREADS, CD_SMEM, st>>>(gpool, argmax, m->raw_contract_w, N, L, P, tiles, S,
                                                                                                        reinterpret_cast<const float4 *>(x0r));
        STOF_LAUNCH_CHECK();
    }
    // ---- conv1: x0r = relu(conv1 x)
    if (!pl.sgb) {                                    // with the semi-global block: applied by contract_dgrad's write-out
        relu_bwd_kernel<<<ew_grid(n4), 256, 0, st>>>(reinterpret_cast<float4 *>(S), reinterpret_cast<const float4 *>(x0r), n4);
        STOF_LAUNCH_CHECK();
    }
    {
        const int chunks = (L + C1_CHUNK - 1) / C1_CHUNK;
        const long long total = (long long)N * chunks;
        const int g1 = (int)(total < 4 * kSMs ? total : 4 * kSMs);
        first_wgrad_kernel<<<g1, 256, 0, st>>>(reinterpret_cast<const float4 *>(S), x, N, L, chunks, partial);
        STOF_LAUNCH_CHECK();
        reduce_parts_kernel<<<(C1_LEN + 255) / 256, 256, 0, st>>>(partial, g1, C1_LEN, gr->conv1_w, 64 * 9, gr->conv1_b);
        STOF_LAUNCH_CHECK();
    }
    return STOF_OK;
}

extern "C" int stof_adamw_step(float *params_dev, const float *grads_dev, float *exp_avg_dev, float *exp_avg_sq_dev, size_t n,
                               int step, float lr, float beta1, float beta2, float eps, float weight_decay, void *stream) {
    STOF_REQUIRE(params_dev && grads_dev && exp_avg_dev && exp_avg_sq_dev && n > 0 && step >= 1, STOF_E_INVALID,
                 "stof_adamw_step: bad arguments (n=%zu step=%d)", n, step);
    const double bc1 = 1.0 - pow((double)beta1, (double)step);
    const double bc2 = 1.0 - pow((double)beta2, (double)step);
    adamw_kernel<<<ew_grid(n), 256, 0, (cudaStream_t)stream>>>(params_dev, grads_dev, exp_avg_dev, exp_avg_sq_dev, n, lr, beta1, beta2, eps,
                                                              weight_decay, (float)bc1, (float)sqrt(bc2));
    STOF_LAUNCH_CHECK();
    return STOF_OK;
}

extern "C" int stof_adamw_step_dev(float *params_dev, const float *grads_dev, float *exp_avg_dev, float *exp_avg_sq_dev, size_t n,
                                   const float *hyper_dev, void *stream) {
    STOF_REQUIRE(params_dev && grads_dev && exp_avg_dev && exp_avg_sq_dev && hyper_dev && n > 0, STOF_E_INVALID,
                 "stof_adamw_step_dev: bad arguments (n=%zu)", n);
    adamw_dev_kernel<<<ew_grid(n), 256, 0, (cudaStream_t)stream>>>(params_dev, grads_dev, exp_avg_dev, exp_avg_sq_dev, n, hyper_dev);
    STOF_LAUNCH_CHECK();
    return STOF_OK;
}
