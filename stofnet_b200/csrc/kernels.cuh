// Internal launcher declarations shared by the translation units of libstofnet_b200.so.
#pragma once
#include <cuda_runtime.h>
#include <stdint.h>

namespace stof {

// conv_simt.cu ------------------------------------------------------------------------------
int launch_conv_first(const float *x, int N, int L, const float *w, const float *b, float *out, int K, int act,
                      cudaStream_t st);
int launch_conv_mid_simt(const float *in, const float *res, float *out, const float *w, const float *b, int N, int L,
                         int epi, cudaStream_t st);
int launch_contract_pool_simt(const float *in, float *pooled, const float *w, const float *b, int N, int L, int P,
                              cudaStream_t st, int *argmax = nullptr);
int launch_expand(const float *pooled, float *e, const float *w, const float *b, int N, int P, cudaStream_t st);
int launch_sgb_add(float *act, const float *e, int N, int L, int P, int scale, int off, cudaStream_t st);
int launch_conv_last(const float *in, float *mask, const float *w, const float *b, int N, int L, int cin, int act,
                     cudaStream_t st);
int launch_espcn_conv2(const float *in, float *out, const float *w, const float *b, int N, int L, cudaStream_t st);
// ESPCN_1D forward in one kernel (espcn_fused.cu): x (N, L) -> mask (N, 4 L); weights in the layouts of the three kernels above
int launch_espcn_fused(const float *x, float *mask, int N, int L, const float *w1, const float *b1, const float *w2,
                       const float *b2, const float *w3, const float *b3, cudaStream_t st);

// espcn_tc.cu: the same forward with conv2 (64 -> 32, k3) and conv3 (32 -> 4, k3) on tcgen05; w23_packed = fp16 taps of conv2
// [tap][8 c8][W_hi 32 | W_lo 32][8] followed by conv3's [tap][4 c8][32 rows: 0-3 W_hi, 8-11 W_lo, 16-19 W_hi][8]
int launch_espcn_tc(const float *x, float *mask, int N, int L, const float *w1, const float *b1, const void *w23_packed, const float *b2,
                    const float *b3, cudaStream_t st);

// conv_tc.cu (tcgen05 / TMEM path) ---------------------------------------------------------------
constexpr int TC_MAX_LAYERS = 6;
enum TcLayerType { TC_LEAKY = 0, TC_RES = 1, TC_LONGSKIP = 2, TC_LAST = 3, TC_RELU = 4, TC_NONE = 5 /* conv + bias only (data gradients) */ };

struct TcLayer {
    const void *w;        // packed fp16 taps: [tap][c8][w_rows][8]
    const float *bias;
    int taps, pad, type;
    int stage_bytes;      // bytes of one tap
    int n_out;            // real output channels (64, or r = 4 for conv_last)
    int n_stack, n_lo;    // N of the stacked [W_hi; W_lo] MMA and of the X_lo * W_hi MMA
    int lo_row, lo_col;   // first weight row of the W_hi block used by the second MMA; its first accumulator column
    int w_rows;           // rows per c8 chunk of a tap
    int n_fast;           // N of the single X_hi * W_hi MMA of the one-pass fp16 mode (64, or 16 for conv_last)
    const void *w_fast;   // one-pass mode: W_hi-only taps [tap][c8][w_rows_fast][8]
    int stage_bytes_fast, w_rows_fast;
    // CTA-pair kernel (conv_tc2.cu): each CTA of the pair streams its own half of every tap, [tap][c8][rows2][8];
    // rank 0: W_hi | first half of W_hi again, rank 1: W_lo | second half of W_hi (see upload_tc)
    const void *w2[2], *w2_fast[2];
    int stage_bytes2, rows2, lo_row2;          // exact mode: bytes per tap and CTA, rows per c8 chunk, first row of the X_lo product's B
    int stage_bytes2_fast, rows2_fast;
};

// Training step: a gradient tensor kept as split planes holds g * 2^E; its slot records E and the bit pattern of max |stored value|
// (non-negative floats order like their bit patterns: atomicMax).  The chain kernel's gradient mode (TcGrad) reads the slots of
// its operands, chooses the output's E from them on the device, and records the output's slot - no pass over a gradient tensor
// just to find its scale.
struct GradSlot { unsigned amax_bits; int E; };
struct TcGrad {
    int mode;                   // 1: out = conv(in);  2: out = conv(in) + aux (aux = gradient planes with their own slot);
                                // 3: out = conv(in) * lrelu'(aux) (aux = activation planes)
    const void *aux;
    const GradSlot *g_in, *g_aux;
    GradSlot *g_out;            // amax_bits zeroed by the caller
    float *bias_partial;        // [gridDim][64]: per-CTA sums of the stored output over its valid rows (bias gradient of the conv
                                // whose output gradient this is), or NULL
};

struct TcParams {
    TcLayer layer[TC_MAX_LAYERS];
    int n_layers;
    int b_row[TC_MAX_LAYERS];   // first smem row computed by layer k
    int r_in;                   // input rows staged per tile
    int l_shift;                // position of smem row 0 relative to the tile's first output position
    int t_out;                  // valid output positions per tile
    int free_layer;             // layer after whose epilogue the input buffer P is no longer touched by the tile
    int stash;                  // 1: the last layer's short skip is parked in TMEM by the epilogue of layer n_layers-2
    int dbl;                    // 1: one-layer segment, the tiles' inputs alternate between buffers P and Q
    int in_layer, in_tap;       // the producer requests the next tile's input after issuing this weight tap
    int tiles_per_row, total_tiles;
    int N, L;
    int n_stages, stage_stride; // weight ring depth / bytes between stages
    int fast;                   // 1: one-pass fp16 mode (hi planes only, one MMA per k-step); 0: fp32-faithful hi/lo split
    const void *in;             // split planes [N][2][8][L][8] fp16
    const void *x0;             // split planes of the long-skip tensor (TC_LONGSKIP), or NULL
    void *out_planes;           // segment output (split planes) unless the segment ends with TC_LAST
    float *out_mask;            // (N, 4L) when the segment ends with TC_LAST
    unsigned long long *timeline;   // debug: [0] = event count, then (code, clock64) pairs written by CTA 0; or NULL
    int timeline_cap;
    TcGrad grad;                // gradient mode of the training step (chain_tc_kernel<false, true> only; mode 0 otherwise)
};
void set_chain_timeline(void *dev_buf, int cap);

// layers[first .. first+count) of the spec list; returns STOF_E_* on error
int launch_chain_tc(const TcLayer *layers, int count, const void *in, const void *x0, void *out_planes, float *out_mask,
                    int N, int L, cudaStream_t st, int fast = 0);
// one TC_NONE layer in gradient mode (the training step's data gradients; see TcGrad)
int launch_chain_tc_grad(const TcLayer *layer, const void *in, void *out_planes, int N, int L, cudaStream_t st, const TcGrad &grad);
int chain_tc_grid(int count, const TcLayer *layers, int N, int L);       // CTAs of that launch (= parts of TcGrad::bias_partial)
// conv1l_tc.cu: ONE 64 -> 64 k7 layer per launch with the layer's taps resident in shared memory (the training step keeps every
// activation and cannot fuse layers).  type: TC_LEAKY, TC_LONGSKIP (aux = skip planes) or TC_NONE; grad (or NULL): gradient mode,
// TcGrad::bias_partial with conv1l_tc_grid(N, L) parts.  Results equal launch_chain_tc's one-layer launch bit for bit.
int conv1l_tc_grid(int N, int L);
int launch_conv1l_tc(const TcLayer *layer, int type, const void *in, const void *aux, void *out, int N, int L, cudaStream_t st,
                     const TcGrad *grad = nullptr);
// conv_tc2.cu: the same segment on CTA pairs (2-CTA clusters, tcgen05 cta_group::2: M = 256 per MMA, each CTA reads half of B)
int launch_chain_tc2(const TcLayer *layers, int count, const void *in, const void *x0, void *out_planes, float *out_mask,
                     int N, int L, cudaStream_t st, int fast = 0);
// wgrad_tc.cu: weight gradient of a 64 -> 64 k7 conv on tcgen05 from the split planes of the output gradient (g * 2^E, E in its
// slot) and of the conv's input; dW in PyTorch's (64, 64, 7) layout; scratch: wgrad_tc_scratch_floats()
// floats.  Deterministic (fixed-order reduction).  The bias gradient: launch_grad_to_planes + launch_bias_parts_reduce below.
size_t wgrad_tc_scratch_floats();
int launch_wgrad_tc(const void *g_planes, const void *a_planes, const GradSlot *slot, int N, int L, float *scratch, float *dW,
                    cudaStream_t st);
// x (N, L) fp32 -> split planes of x0 = act(conv_first(x)) [+ nearest-upsampled E, centred]  (models/stofnet.py:45,96-115)
int launch_x0_split(const float *x, const float *e, int N, int L, int P, int scale, int off, const float *w, const float *b,
                    void *planes, cudaStream_t st, int fast = 0, int K = 9);
// contract_tc.cu: pooled = maxpool80(LeakyReLU(contract_conv(relu(conv1 x)))) on tcgen05, conv1 recomputed per tile
// argmax (or NULL): (N, 512, P) positions of the pooled maxima, for the training step's backward pass
int launch_contract_pool_tc(const float *x, const void *w, const float *bias, const float *w1, const float *b1, float *pooled,
                            int N, int L, int P, cudaStream_t st, int fast = 0, int *argmax = nullptr);
// expand_tc.cu: E = LeakyReLU(expand_conv(pooled)) on tcgen05 (fp32 in / out, operands split on the fly)
int launch_expand_tc(const float *pooled, float *e, const void *w_packed, const float *bias, int N, int P, cudaStream_t st);
// fp32 C4L4 activations <-> split planes (hand-over between the fp32 and tensor paths).  absmax_bits (device, or NULL): bit
// pattern of max |act| of the tensor; the planes then carry act * 2^k with 2^k max|act| in [2^10, 2^11) - gradients are ~1e-6
// and would sit in fp16's subnormal range unscaled - and the way back divides by 2^k.  add: act (+)= instead of =.
int launch_c4l4_to_planes(const float *act, void *planes, int N, int L, cudaStream_t st, const unsigned *absmax_bits = nullptr);
int launch_planes_to_c4l4(const void *planes, float *act, int N, int L, cudaStream_t st, const unsigned *absmax_bits = nullptr,
                          int add = 0);
// The training step's gradient hand-overs (train.cu):
//   fp32 -> scaled planes with the bias gradient's partial sums on the way (bias_partial[part][64], grad_to_planes_parts(N, L)
//   parts, added in order by launch_bias_parts_reduce, wgrad_tc.cu) and the planes' slot for the chain kernel's gradient mode;
//   planes of a slot (g * 2^E) -> fp32, plus a second fp32 addend (or NULL)
int launch_grad_planes_to_c4l4(const void *planes, float *act, int N, int L, cudaStream_t st, const GradSlot *slot, const float *add2);
int grad_to_planes_parts(int N, int L);
int launch_grad_to_planes(const float *g, void *planes, int N, int L, cudaStream_t st, const unsigned *absmax_bits, float *bias_partial,
                          GradSlot *slot_out = nullptr);
// slot (or NULL): the partial sums are of stored values g * 2^E - db = 2^-E * sum
int launch_bias_parts_reduce(const float *bias_partial, int parts, float *db, cudaStream_t st, const GradSlot *slot = nullptr);
// *out_bits = max(*out_bits, bits of max |v|) over n4 float4 (non-negative floats order like their bit patterns); zero the slot first
int launch_absmax(const float *v, size_t n4, unsigned *out_bits, cudaStream_t st);

}  // namespace stof
