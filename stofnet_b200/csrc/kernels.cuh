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
                              cudaStream_t st);
int launch_expand(const float *pooled, float *e, const float *w, const float *b, int N, int P, cudaStream_t st);
int launch_sgb_add(float *act, const float *e, int N, int L, int P, int scale, int off, cudaStream_t st);
int launch_conv_last(const float *in, float *mask, const float *w, const float *b, int N, int L, int cin, int act,
                     cudaStream_t st);
int launch_espcn_conv2(const float *in, float *out, const float *w, const float *b, int N, int L, cudaStream_t st);

}  // namespace stof
