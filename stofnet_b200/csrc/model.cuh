// Internal definition of the opaque stof_model handle (shared by stof_api.cu and train.cu).
#pragma once
#include <stdlib.h>
#include <vector>

#include "common.cuh"
#include "kernels.cuh"

namespace stof {

// Device-resident weights in the layouts the kernels read (all fp32):
//   first : [K][64]            w[t][co]        = W[co][0][t]
//   mid   : [64][K][64]        w[ci][t][co]    = W[co][ci][t]          (conv2..conv12, contract tiles)
//   expand: [512][5][64]       w[ci][t][co]
//   last  : [CIN][3][4]        w[ci][t][u]     = W[u][ci][t]
struct DevConv {
    float *w = nullptr;
    float *b = nullptr;
};

}  // namespace stof

struct stof_model {
    int device = 0;
    int kind = STOF_KIND_STOFNET;
    int F = 64, r = 4, n_mid = 0;
    int k_first = 9, k_mid = 7, k_last = 3;
    int sgb_scale = 1, sgb_feat = 0, sgb_k = 0;
    stof::DevConv first, last, contract, expand;
    stof::DevConv mid[STOF_MAX_MID_CONVS];
    // tensor-core path (conv_tc.cu): packed fp16 hi/lo taps of conv2..conv{nb-1} + conv_last, one allocation
    int path = STOF_PATH_TC;
    int seg_len = 4;                                  // layers fused per chain launch
    void *tc_w = nullptr;
    void *tc_w_fast = nullptr;                        // W_hi-only taps of the one-pass fp16 mode
    void *tc_w2 = nullptr, *tc_w2_fast = nullptr;     // per-rank half-taps of the CTA-pair chain kernel (conv_tc2.cu)
    // chain kernel variant: one CTA per tile (default), or CTA pairs with cta_group::2 MMAs (STOF_CHAIN_PAIR=1 in the
    // environment, or stof_model_set_chain_kernel; bit-identical results, measured 13-21 % slower: see DESIGN.md)
    bool chain_pair = getenv("STOF_CHAIN_PAIR") && getenv("STOF_CHAIN_PAIR")[0] == '1';
    void *tc_expand_w = nullptr;                      // expand conv: [8 groups][5 taps][8 c8][W_hi 64 | W_lo 64][8] fp16
    void *tc_contract_w = nullptr;                    // contract conv: [4 slices][5 taps][8 c8][W_hi 128 | W_lo 128][8] fp16
    stof::TcLayer tc_layer[STOF_MAX_MID_CONVS + 1];   // spec list: mid convs then conv_last
    int n_tc = 0;
    // The fp16 hi/lo operand packs above are built on first use after stof_model_set_weights (an optimizer step re-uploads
    // the weights every iteration and its fp32 training kernels never read them): host copies of the tensors they need
    bool tc_dirty = false;
    // stof_model_set_weights_device: the device layouts are newer than the host copies below (refreshed by ensure_tc)
    bool host_stale = false;
    std::vector<float> h_mid[STOF_MAX_MID_CONVS], h_last, h_contract, h_expand;
    // training step (train.cu): dgrad weights of the mid convs [co][K-1-t][ci] (a conv with channels swapped and taps
    // flipped), a zero bias vector, and PyTorch-layout copies of conv_last / contract / expand for the backward kernels
    float *bw_mid[STOF_MAX_MID_CONVS] = {};
    float *zero_bias = nullptr;                       // 512 zeros
    float *raw_last_w = nullptr;                      // (4, 64, 3)
    float *raw_contract_w = nullptr;                  // (512, 64, 5)
    float *raw_expand_w = nullptr;                    // (64, 512, 5)
    // tensor-core training step (train.cu): fp16 hi/lo taps of the mid convs in the forward and the data-gradient orientation,
    // packed ON THE DEVICE from mid[].w by the first training call after a weight change (an optimizer step changes them all)
    void *tr_tc_w = nullptr;
    stof::TcLayer tr_fwd[STOF_MAX_MID_CONVS], tr_bwd[STOF_MAX_MID_CONVS];
    bool tr_tc_dirty = true;
    // backward pass: the weight / bias gradients of the mid convs run on a side stream next to the data-gradient chain (they only
    // read what the chain has produced); events fork and join it, also under CUDA-graph capture
    cudaStream_t tr_side = nullptr;
    static constexpr int kTrainEvents = 2 * STOF_MAX_MID_CONVS + 4;
    cudaEvent_t tr_ev[kTrainEvents] = {};
    // scratch of stof_infer_host (grow-only)
    static constexpr int kPipe = 2;
    cudaStream_t streams[kPipe] = {nullptr, nullptr};
    void *scratch[kPipe] = {nullptr, nullptr};
    size_t scratch_bytes[kPipe] = {0, 0};
    int64_t *h_stats = nullptr;   // pinned, kPipe * 2
    int host_chunk_rows = 0;      // stof_model_set_host_chunk_rows: rows per pipelined chunk of stof_infer_host (0 = automatic)
    // optional per-kernel-class CUDA-event timing (stof_profile_*): pairs recorded on the launch stream
    bool profiling = false;
    std::vector<cudaEvent_t> ev_pool;                 // start/stop pairs
    std::vector<int> ev_class;                        // class of pair i
    size_t ev_used = 0;
};

