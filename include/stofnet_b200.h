/*
 * stofnet_b200.h - C ABI of the B200-native StofNet echo-waveform inference path.
 *
 * Drop-in boundary for ONE hot path of hahnec/stofnet: StofNet.forward -> sub-pixel shuffle ->
 * NMS / threshold / time-of-flight extraction.  The reference is pure Python on PyTorch and has
 * no FFI of its own; each entry point below names the reference call site it replaces (paths
 * relative to the reference repository root).  INTEGRATION.md shows the ctypes stub a maintainer
 * of the reference would add.
 *
 * Conventions
 *   - plain pointers and sizes only; `stream` is a cudaStream_t passed as void* (NULL = default)
 *   - *_dev pointers are device memory owned by the caller (e.g. torch's caching allocator);
 *     *_host pointers are host memory (pinned recommended)
 *   - every call is asynchronous on `stream` unless stated otherwise
 *   - returns 0 on success, a negative STOF_E_* code otherwise; stof_last_error() gives the text
 *   - there is no CPU fallback: without a CUDA device every compute call returns STOF_E_CUDA
 *   - a model handle is not thread-safe; use one handle per device / stream
 */
#ifndef STOFNET_B200_H
#define STOFNET_B200_H

#include <stddef.h>
#include <stdint.h>

#ifdef __cplusplus
extern "C" {
#endif

#define STOF_OK            0
#define STOF_E_INVALID    -1   /* bad argument (null pointer, non-positive size, unsupported shape) */
#define STOF_E_LENGTH     -2   /* L < scale: pooled length would be 0 (torch raises in max_pool1d; models/stofnet.py:91) */
#define STOF_E_LENGTH_ODD -3   /* (L mod scale) odd: reference fails at torch.add (models/stofnet.py:111-115) */
#define STOF_E_WORKSPACE  -4   /* workspace too small */
#define STOF_E_CUDA       -5   /* CUDA runtime error (text in stof_last_error) */
#define STOF_E_NCCL       -6   /* reserved: collective error */
#define STOF_E_UNSUPPORTED -7  /* architecture of the checkpoint not supported by the kernels */

#define STOF_MAX_MID_CONVS 24

/* Model kinds */
#define STOF_KIND_STOFNET 0    /* models/stofnet.py  */
#define STOF_KIND_ESPCN   1    /* models/espcn_1d.py */
#define STOF_KIND_EDSR    2    /* models/edsr_1d.py  */

/* Host-side description of a checkpoint: fp32 tensors in PyTorch's (Cout, Cin, K) layout exactly
 * as stored in the reference's .pth state_dict (main.py:174-177).  Pointers are host memory and
 * are only read during stof_model_create / stof_model_set_weights.                              */
typedef struct stof_weights {
    int kind;                   /* STOF_KIND_* */
    /* StofNet (models/stofnet.py:11-40) */
    const float *conv1_w, *conv1_b;                 /* (F,1,9), (F)            :23 */
    int n_mid;                                      /* num_blocks-2 = 11 convs: conv2..conv12 */
    const float *mid_w[STOF_MAX_MID_CONVS];         /* (F,F,7)                 :30-31 */
    const float *mid_b[STOF_MAX_MID_CONVS];
    const float *last_w, *last_b;                   /* (r,F,3), (r)            :24 */
    const float *contract_w, *contract_b;           /* (8F,F,5) or NULL        :88 */
    const float *expand_w, *expand_b;               /* (F,8F,5) or NULL        :94 */
    int num_features;                               /* F = 64 */
    int upsample_factor;                            /* r = conv_last.out_channels (4); NOT rf_scale_factor */
    int k_first, k_mid, k_last;                     /* 9, 7, 3 */
    int sgb_scale, sgb_feat, sgb_k;                 /* 80, 512, 5 ; sgb_scale = 1 when no semi-global block (:27) */
    /* ESPCN_1D (models/espcn_1d.py:12-14): conv1 (64,1,5) -> conv1_w/b, conv2 (32,64,3) -> mid_w[0]/mid_b[0],
     * conv3 (r,32,3) -> last_w/b
     * EDSR_1D (models/edsr_1d.py:22-30): conv_input (64,1,3) -> conv1_w/b; residual_blocks[i].conv1 / .conv2 (64,64,3) ->
     * mid_w[2i] / mid_w[2i+1]; conv_mid -> mid_w[2*num_blocks]; conv_output (1,16,3) -> last_w/b; n_mid = 2*num_blocks+1,
     * k_first = k_mid = k_last = 3 */
} stof_weights;

typedef struct stof_model stof_model;   /* opaque: device-resident repacked weights */

/* Library / device probes (no GPU work) */
const char *stof_version(void);
const char *stof_last_error(void);
int stof_device_count(void);            /* 0 when no CUDA device is visible */

/* Replaces StofNet(...).to(device) + load_state_dict (main.py:133-134,173-177).  Synchronous. */
int stof_model_create(const stof_weights *w, int device, stof_model **out);
/* Re-upload weights into an existing handle (after an optimizer step).  Synchronous. */
int stof_model_set_weights(stof_model *m, const stof_weights *w);
/* The same for a training loop whose optimizer updates the parameters IN DEVICE MEMORY (main.py:246-248 optimizer.step(),
 * :179 optim.AdamW): `w_dev` holds DEVICE pointers to fp32 tensors in PyTorch's layouts on the model's device; ONE kernel on
 * `stream` re-derives every device-side layout the fp32 forward / backward kernels read (no host round trip, no
 * synchronisation; ordered after the optimizer kernel on the same stream).  StofNet only.  The fp16 operand packs of the
 * tensor-core inference kernels are rebuilt on the next inference call (one device-to-host copy there).                  */
int stof_model_set_weights_device(stof_model *m, const stof_weights *w_dev, void *stream);
/* Host-side bookkeeping of the above without a launch: for a training step replayed from a CUDA graph (the graph holds the
 * repack kernel; the handle must still learn that the tensor-core inference packs are stale).                             */
int stof_model_mark_weights_changed(stof_model *m);
void stof_model_destroy(stof_model *m);

/* Rows processed per internal chunk and the workspace this needs.  chunk_rows <= 0 picks a default. */
int stof_forward_workspace_bytes(const stof_model *m, int N, int L, int chunk_rows, size_t *bytes);

/* Replaces `masks_pred = model(frame)` (main.py:221,314; models/stofnet.py:42-67 incl. the
 * SampleShuffle1D of utils/sample_shuffle.py:10-28 fused into the last conv's store).
 *   x_dev    (N, 1, L) fp32 contiguous
 *   mask_dev (N, 1, r*L) fp32 contiguous
 *   ws_dev   workspace of at least stof_forward_workspace_bytes(...) bytes                     */
int stof_forward(stof_model *m, const float *x_dev, int N, int L, float *mask_dev,
                 void *ws_dev, size_t ws_bytes, int chunk_rows, void *stream);

/* Replaces utils/mask2samples.py: nms_1d (:5-11) + thresholding (:14-23) + torch.nonzero (:32) +
 * bincount / ragged pack / divide by upsample_factor of mask2coords (:92-112).
 *   scores_dev (N, M) fp32 contiguous (M = r*L)
 *   window      nms window (made odd like the reference: window/2*2+1)
 *   th_is_none  1 = "threshold is None or 0.0" branch (keep entries equal to the row max)
 *   coords_dev (N, cap) fp32 or NULL: ascending position / upsample_factor, zero padded
 *   pos_dev    (N, cap) int32 or NULL: ascending integer positions, -1 padded
 *   counts_dev (N) int32: detections per row (may exceed cap: caller re-runs with a larger cap)
 *   stats_dev  (2) int64: [0] = max_k counts_k (K_max, :93), [1] = total detections K; must be
 *               zeroed by the caller before the call                                            */
int stof_peaks(const float *scores_dev, int N, int M, int window, int th_is_none, float th,
               int upsample_factor, float *coords_dev, int32_t *pos_dev, int cap,
               int32_t *counts_dev, int64_t *stats_dev, void *stream);

/* (K,2) int64 [row, position] list of get_maxima_positions (utils/mask2samples.py:26-34) from the
 * per-row output of stof_peaks.  offsets_dev (N) int64 scratch.  K = stats[1].                   */
int stof_peaks_indices(const int32_t *pos_dev, const int32_t *counts_dev, int N, int cap,
                       int64_t *offsets_dev, int64_t *idx_dev, void *stream);

/* Replaces nms_1d alone (utils/mask2samples.py:5-11): out = (s == window_max(s)) * s. */
int stof_nms_1d(const float *scores_dev, int N, int M, int window, float *out_dev, void *stream);

/* End-to-end call with HOST buffers: the reference's eval step `model(frame)` + `mask2coords`
 * (main.py:307-320) including the host->device copy of the waveforms and the device->host copy
 * of the time-of-flight table.  Synchronous; rows are pipelined in chunks over internal streams.
 *   x_host (N, L) fp32; coords_host (N, cap) fp32; counts_host (N) int32; mask_host (N, r*L) or NULL
 *   kmax_out: max detections in any row (if > cap the table is truncated: re-run with cap >= kmax) */
int stof_infer_host(stof_model *m, const float *x_host, int N, int L, int window, int th_is_none,
                    float th, float *coords_host, int cap, int32_t *counts_host, float *mask_host,
                    int *kmax_out);

/* Rows per pipelined chunk of stof_infer_host (0 = automatic: balanced chunks of about 4 Mi positions, at least two). */
int stof_model_set_host_chunk_rows(stof_model *m, int rows);

/* SampleShuffle1D.forward (utils/sample_shuffle.py:10-28) as a standalone operator:
 * x (B, r*Cl, W) -> out (B, Cl, r*W), out[b,c,w*r+u] = x[b,u*Cl+c,w].                           */
int stof_sample_shuffle(const float *x_dev, int B, int Cs, int W, int r, float *out_dev, void *stream);

/* Per-kernel-class device timing for bench.py's roofline: when enabled, stof_forward /
 * stof_infer_host record a CUDA-event pair on the launch stream around every kernel launch;
 * stof_profile_collect waits for them and returns summed milliseconds and launch counts per class.
 * (The reference has only a host process_time() pair around model(frame), main.py:313-315.)      */
#define STOF_PROF_FIRST    0   /* conv1 + ReLU */
#define STOF_PROF_CONTRACT 1   /* contract conv + LeakyReLU + max-pool */
#define STOF_PROF_EXPAND   2   /* expand conv + LeakyReLU */
#define STOF_PROF_SGB_ADD  3   /* nearest upsample + pad + skip add */
#define STOF_PROF_MID      4   /* conv2..conv12 */
#define STOF_PROF_LAST     5   /* conv_last + sample shuffle */
#define STOF_PROF_PEAKS    6   /* nms + threshold + compaction (stof_infer_host only) */
#define STOF_PROF_CLASSES  8
int stof_profile_enable(stof_model *m, int on);
int stof_profile_collect(stof_model *m, double *ms_by_class, int64_t *launches_by_class);

/* Arithmetic path of the 64->64 conv stack (models/stofnet.py:51-65).
 *   STOF_PATH_TC    tcgen05/TMEM implicit GEMM; every fp32 operand is carried as an fp16 hi/lo pair (22 significand
 *                   bits, fp32 accumulation): fp32-faithful results at fp16 tensor-core rate.  Default for StofNet.
 *   STOF_PATH_SIMT  plain fp32 FFMA kernels (bring-up / cross-check path; also what ESPCN_1D uses).
 * layers_per_launch > 0 sets how many conv layers one fused chain launch covers (1..6, default 4).     */
#define STOF_PATH_SIMT 0
#define STOF_PATH_TC   1
/*   STOF_PATH_TC_FP16  same kernels, ONE tensor-core pass on fp16-rounded operands (fp32 accumulation): a third of the MMAs and
 *                   half of the activation traffic.  Not index-exact: it is the "stated tolerance" mode of the reduced-precision
 *                   path (a TF32-like 11-bit significand; see DESIGN.md for the measured detection agreement).  Opt-in only. */
#define STOF_PATH_TC_FP16 2
int stof_model_set_path(stof_model *m, int path, int layers_per_launch);

/* Variant of the fused chain kernel: cta_pairs = 0 (default) runs one CTA per tile; 1 runs every tile pair on a 2-CTA cluster
 * with tcgen05 cta_group::2 MMAs (each SM reads half of the weight operand).  Same arithmetic in the same order: the masks are
 * bit-identical (tests/test_gpu_tc.py); the pair form measured slower (cross-CTA barrier latency on the per-layer critical path,
 * DESIGN.md).  STOF_CHAIN_PAIR=1 in the environment sets the default to 1.                                              */
int stof_model_set_chain_kernel(stof_model *m, int cta_pairs);

/* One fused chain launch on caller-provided activations (test hook for single layers / segments):
 * layers [first, first+count) of the list conv2 .. conv{nb-1}, conv_last.
 *   in_planes_dev  (N, 2, 8, L, 8) fp16: plane 0 = fp16(v), plane 1 = fp16((v - plane0) * 2^11), channel c at [c/8][l][c%8]
 *   x0_planes_dev  same layout: the long-skip tensor res1 (needed when the range contains conv{nb-1}); may be NULL otherwise
 *   out_dev        same layout, or the (N, r*L) fp32 mask when the range ends with conv_last                 */
int stof_chain_segment(stof_model *m, int first, int count, const void *in_planes_dev, const void *x0_planes_dev,
                       void *out_dev, int N, int L, void *stream);

/* The expanding half of the semi-global block alone (test hook): E = LeakyReLU(expand_conv(pooled)), models/stofnet.py:94-95,
 * 104-105.  pooled_dev (N, 512, P) fp32 -> e_dev (N, 64, P) fp32.  path: STOF_PATH_TC = tcgen05 kernel (what stof_forward
 * runs), STOF_PATH_SIMT = fp32 FFMA kernel (what the training forward runs).                                  */
int stof_sgb_expand(stof_model *m, const float *pooled_dev, int N, int P, float *e_dev, int path, void *stream);

/* Debug: chain launches issued after this call make four warps of CTA 0 (producer, the two MMA issuers, one epilogue
 * warp) append (code, clock64) event pairs to dev_buf (uint64[4 + 2*capacity]: [role] = event count of that role, then
 * four regions of capacity/4 pairs; must be zeroed by the caller).  NULL switches it off.                  */
int stof_debug_set_timeline(void *dev_buf, int capacity_events);

/* Counter of kernels launched by this library since load (bench.py's gpu_launches). */
uint64_t stof_launch_count(void);

/* ------------------------------------------------------------------------------------------------
 * Operators either side of the path (SURVEY.md 8f "next" rows and 8a rows a9-a13)
 * ------------------------------------------------------------------------------------------------ */

/* Replaces toa_rmse (utils/metrics.py:9-41), the consumer of the ToF table with `etol` (main.py:347).
 *   gt_dev (N, G) fp32, es_dev (N, E) fp32: zeros / NaN / inf are "no echo" and dropped (:6,19-20)
 *   tol compares with the SQUARED distance (:27-33)
 *   out_dev (N, 7) fp32: rmse, precision, recall, jaccard [percent], tp, fp, fn; rows without a valid gt or
 *   estimate give 0, NaN, NaN, NaN, 0, 0, 0 exactly like the reference's 0/0 divisions                  */
int stof_toa_rmse(const float *gt_dev, int G, const float *es_dev, int E, int N, float tol, float *out_dev, void *stream);

/* Replaces the `echo_max` branch of mask2coords (utils/mask2samples.py:103-111) with reduce_echoes (:117-132) and
 * get_amplitudes (:135-136): per row keep the echo_max entries of the position table with the largest mask amplitude
 * scores[row][round(position)] (descending, NaN first, ties in table order), re-ordered by position, / divisor.
 *   scores_dev (N, M) fp32 (may be NULL when K <= echo_max), coords_dev (N, K) fp32 positions (undivided, zero padded,
 *   row stride ld >= K), out_dev (N, echo_max) fp32; K <= echo_max copies the table and pads it with zeros (:107-109).  */
int stof_reduce_echoes(const float *scores_dev, int N, int M, const float *coords_dev, int ld, int K, int echo_max,
                       float divisor, float *out_dev, void *stream);

/* Replaces coords2mask (utils/mask2samples.py:139-148): mask = 0; mask[n, max(s,0)] = 1; mask[n, 0] = 0.
 *   samples_dev (N, G) int64 (negative entries are clamped to 0 IN PLACE like the reference; entries >= M raise in
 *   torch's scatter_ and are ignored here), mask_dev (N, M) fp32                                        */
int stof_coords2mask(int64_t *samples_dev, int N, int G, int M, float *mask_dev, void *stream);

/* Replaces the training / validation loss of main.py:228-232 in one pass over the mask:
 *   true  = coords2mask(gt_true);  blur = conv1d(true, gauss(k, sigma), pad k/2);  blur *= amplitude / max(blur)
 *   loss  = mean((pred - blur)^2) + lambda * mean(|pred|)
 *   gauss_dev (k) fp32 taps (utils/gaussian.py:4-7), k odd <= 31
 *   blur_max_io_dev (1) fp32: global max of the un-normalised blur.  mode 0: computed here (single GPU, main.py:230);
 *       mode 1: only compute it (first phase of the sharded loss: MAX-all-reduce it, then call with mode 2);
 *       mode 2: use the value already in blur_max_io_dev
 *   loss_dev (1) fp64: the loss over these N rows (mean over N*M elements); grad_dev (N, M) fp32 or NULL: d loss / d pred
 *   scratch_dev: at least stof_loss_scratch_bytes(N, M) bytes                                              */
size_t stof_loss_scratch_bytes(int N, int M);
int stof_mask_loss(const float *pred_dev, const int64_t *samples_dev, int N, int G, int M, const float *gauss_dev, int k,
                   float amplitude, float lambda, int mode, float *blur_max_io_dev, double *loss_dev, float *grad_dev,
                   void *scratch_dev, void *stream);
/* The same loss for callers that cannot allocate (code under CUDA graph capture, tight serving loops, harnesses that forbid
 * cudaMalloc): no scratch argument - the partial sums live in a 256 KB file-scope device array inside the library (held by
 * every GPU that loads it until the process exits), and the call performs no cudaMalloc / cudaFree or any allocating
 * routine.  No counterpart in the reference.
 *   limit: N * ceil(M / 4096) <= 16384, else STOF_E_UNSUPPORTED (use stof_mask_loss with your own scratch buffer)
 *   the device that owns the buffers must be current (as for every entry point here); its ordinal must be < 64
 *   every call on a device shares the array: calls that overlap on the GPU (different streams) race - serialise them.
 *   Nothing in it needs clearing between calls (every slot the final sum reads is rewritten first).
 *   Results are bit-identical to stof_mask_loss; all other arguments and the modes are the same.                       */
int stof_mask_loss_static(const float *pred_dev, const int64_t *samples_dev, int N, int G, int M, const float *gauss_dev,
                          int k, float amplitude, float lambda, int mode, float *blur_max_io_dev, double *loss_dev,
                          float *grad_dev, void *stream);

/* Replaces find_threshold (utils/threshold.py:6-28; sklearn.metrics.roc_curve on batch element 0, main.py:330):
 * clamp negatives (or min-max normalise over the WHOLE batch when norm_opt), ROC of row 0 with drop_intermediate,
 * threshold at the arg-max of sqrt(tpr * (1 - fpr)).
 *   pred_dev, true_dev (N, M) fp32; th_out_dev (1) fp64 (+inf when row 0 has no positive, as sklearn's first point)
 *   scratch_dev: at least stof_threshold_scratch_bytes(N, M) bytes                                       */
size_t stof_threshold_scratch_bytes(int N, int M);
int stof_find_threshold(const float *pred_dev, const float *true_dev, int N, int M, int norm_opt, double *th_out_dev,
                        void *scratch_dev, void *stream);

/* Replaces the dataset-side rf_scale_factor interpolation (datasets/chirp_dataset.py:10-17 upscale_1d and :79-91 iq2rf):
 * linear interpolation of each row from n to floor(n * rf) samples on np.linspace(0, n/fs, ., endpoint=True) grids,
 * optionally modulated onto the carrier: out = Re((I + jQ)(t) * exp(2j pi fc t)).  Computed in fp64 like scipy, stored fp32.
 *   in_re_dev (N, n) fp32 (envelope / I), in_im_dev (N, n) fp32 or NULL (Q); out_dev (N, floor(n * rf)) fp32
 *   fc = 0 and in_im_dev = NULL gives upscale_1d                                                          */
int stof_rf_upscale(const float *in_re_dev, const float *in_im_dev, int N, int n, double rf, double fs, double fc,
                    float *out_dev, void *stream);

/* Replaces AttentionBlock.forward (utils/attention_block.py:13-25; dead code in the reference, parity unpinned):
 *   q, k, v = x Wq^T + bq, ...;  out = dropout(softmax(q k^T / sqrt(H)), p) v
 *   x_dev (B, S, D) fp32; w*_dev (H, D), b*_dev (H) fp32 (nn.Linear layout); out_dev (B, S, H)
 *   dropout_p in [0, 1): 0 disables it (the reference hard-codes 0.5, also in eval); seed drives a counter-based RNG
 *   scratch_dev: at least stof_attention_scratch_bytes(B, S, H) bytes                                     */
size_t stof_attention_scratch_bytes(int B, int S, int H);
int stof_attention_block(const float *x_dev, int B, int S, int D, int H, const float *wq_dev, const float *bq_dev,
                         const float *wk_dev, const float *bk_dev, const float *wv_dev, const float *bv_dev,
                         float dropout_p, uint64_t seed, float *out_dev, void *scratch_dev, void *stream);

/* Standalone operators of the reference's module surface (the fused hot path does not need them; callers of the
 * reference may).  thresholding (utils/mask2samples.py:14-23): th_is_none = 0 -> out = scores with entries < th zeroed
 * (the reference does this in place: pass out_dev = scores_dev); th_is_none = 1 ("threshold is None or 0.0") -> entries
 * below the row maximum zeroed, torch.max's NaN propagation kept.  scores_dev, out_dev (N, M) fp32.                      */
int stof_thresholding(const float *scores_dev, int N, int M, int th_is_none, float th, float *out_dev, void *stream);

/* get_amplitudes (utils/mask2samples.py:135-136): out[n][k] = frames[n][round_half_even(samples[n][k])].
 *   frames_dev (N, M), samples_dev / out_dev (N, K) fp32; *oob_flag_dev (int32, zeroed by the caller) is set to 1 when an
 *   index falls outside [0, M) (torch.gather raises there; the wrapper turns the flag into the same RuntimeError).       */
int stof_get_amplitudes(const float *frames_dev, const float *samples_dev, int N, int M, int K, float *out_dev,
                        int *oob_flag_dev, void *stream);

/* reduce_echoes (utils/mask2samples.py:117-132) on (N, K, C) fp32 entries, column 0 = time sample, column 1 = amplitude,
 * for echo_max < K: the echo_max entries with the largest amplitude, re-ordered by time sample -> out_dev (N, echo_max, C).
 * Equal amplitudes keep input order (torch.argsort without stable=True leaves that order unspecified).                   */
int stof_reduce_echoes_pairs(const float *pairs_dev, int N, int K, int C, int echo_max, float *out_dev, void *stream);

/* SemiGlobalBlock.forward (models/stofnet.py:98-117) as a standalone operator on an fp32 (N, 64, L) NCL tensor:
 * out = x + pad(upsample_nearest(LeakyReLU(expand_conv(maxpool(LeakyReLU(contract_conv(x))))))), fp32 FFMA kernels
 * (inside stof_forward the block runs fused on the tensor cores instead).  Same length constraints as stof_forward.     */
int stof_sgb_workspace_bytes(const stof_model *m, int N, int L, size_t *bytes);
int stof_semi_global_block(stof_model *m, const float *x_dev, int N, int L, float *out_dev, void *ws_dev, size_t ws_bytes,
                           void *stream);

/* ------------------------------------------------------------------------------------------------
 * Training step (main.py:217-248): forward that keeps the activations, backward of the whole stack, fused AdamW.
 * fp32 CUDA-core kernels; gradients are deterministic (fixed reduction order, no float atomics).
 * ------------------------------------------------------------------------------------------------ */

/* Device-side gradient buffers, one per parameter tensor, fp32 in PyTorch's (Cout, Cin, K) / (Cout) layout - i.e. what
 * `param.grad` holds after loss.backward() (main.py:247).  The caller typically carves them out of ONE flat buffer so that the
 * data-parallel gradient all-reduce is a single collective.  Pointers are device memory.                                */
typedef struct stof_grads {
    float *conv1_w, *conv1_b;
    float *mid_w[STOF_MAX_MID_CONVS];
    float *mid_b[STOF_MAX_MID_CONVS];
    float *last_w, *last_b;
    float *contract_w, *contract_b;     /* NULL when the model has no semi-global block */
    float *expand_w, *expand_b;
} stof_grads;

/* Bytes of the activation / scratch workspace one training step of N rows of length L needs (forward + backward). */
int stof_train_workspace_bytes(const stof_model *m, int N, int L, size_t *bytes);

/* `masks_pred = model(frame)` in training mode (main.py:221): same result as stof_forward's fp32 path, and every layer's
 * activations (plus the max-pool arg-max of the semi-global block) stay in ws_dev for stof_backward.                    */
int stof_forward_train(stof_model *m, const float *x_dev, int N, int L, float *mask_dev, void *ws_dev, size_t ws_bytes,
                       void *stream);

/* `loss.backward()` (main.py:247) for the StofNet stack: g_mask_dev (N, r*L) = d loss / d masks_pred -> every parameter's
 * gradient (overwritten, not accumulated: the reference calls optimizer.zero_grad() before every backward, main.py:246).
 * Must follow stof_forward_train on the same x_dev / ws_dev with unchanged weights.
 * Stream-ordered on `stream`: on the tensor-core path the weight / bias gradients of the conv stack run on a side stream of
 * the model, forked from and joined back into `stream` with events before the call returns (safe under CUDA-graph capture
 * once a first, uncaptured call has created that stream).  Deterministic: bit-identical gradients run to run.            */
int stof_backward(stof_model *m, const float *x_dev, const float *g_mask_dev, int N, int L, void *ws_dev, size_t ws_bytes,
                  const stof_grads *grads, void *stream);

/* torch.optim.AdamW step (main.py:179,248; amsgrad off, maximize off) on flat fp32 buffers of n elements:
 *   p *= 1 - lr * wd;  m = b1 m + (1 - b1) g;  v = b2 v + (1 - b2) g^2;
 *   p -= lr / (1 - b1^step) * m / (sqrt(v) / sqrt(1 - b2^step) + eps)            step counts from 1                    */
int stof_adamw_step(float *params_dev, const float *grads_dev, float *exp_avg_dev, float *exp_avg_sq_dev, size_t n, int step,
                    float lr, float beta1, float beta2, float eps, float weight_decay, void *stream);
/* The same with the scalars in DEVICE memory, hyper_dev[7] = {lr, beta1, beta2, eps, weight_decay, 1 - beta1^step,
 * sqrt(1 - beta2^step)}: nothing step-dependent is baked into the launch, so the whole training step can be captured once in
 * a CUDA graph and replayed (copy the seven floats in before each replay).                                                */
int stof_adamw_step_dev(float *params_dev, const float *grads_dev, float *exp_avg_dev, float *exp_avg_sq_dev, size_t n,
                        const float *hyper_dev, void *stream);

#ifdef __cplusplus
}
#endif
#endif /* STOFNET_B200_H */
