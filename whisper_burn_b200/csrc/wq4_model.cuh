// wq4_model.cuh -- launchers of the non-GEMM kernels used by the encoder/decoder host code.
#pragma once
#include <cuda_runtime.h>
#include <stdint.h>

int32_t wq4m_layernorm(const float *x, const float *w, const float *b, float *y32, void *y16, int64_t rows, int64_t D,
                       cudaStream_t st);
int32_t wq4m_enc_attention(const void *qkv16, void *out16, int64_t B, int64_t T, int64_t D, int64_t H, cudaStream_t st);
int32_t wq4m_dec_attention(const float *q, int64_t ldq, const void *kc, const void *vc, int64_t ldkv, int64_t kv_batch_stride,
                           float *out, int64_t ldo, int64_t B, int64_t H, int64_t q_len, int64_t kv_len, int causal_base,
                           cudaStream_t st);
int32_t wq4m_mel_transpose(const float *x, void *xt16, int64_t B, int64_t C, int64_t T, int64_t Cp, int64_t Tp, int64_t pad,
                           cudaStream_t st);
int32_t wq4m_embed(const int *tokens, const float *tok_emb, const float *pos_emb, float *x, int64_t B, int64_t q_len, int64_t D,
                   int64_t pos0, cudaStream_t st);
int32_t wq4m_logits(const float *x, const float *emb, float *logits, int64_t nb, int64_t D, int64_t V, cudaStream_t st);
int32_t wq4m_argmax(const float *logits, int64_t nb, int64_t V, int64_t lo, int64_t hi, const int *suppress, int *out,
                    cudaStream_t st);
int32_t wq4m_kv_append(const float *qkv, int64_t ld_src, void *kc, void *vc, int64_t ldkv, int64_t kv_batch_stride, int64_t B,
                       int64_t D, int64_t q_len, int64_t pos0, cudaStream_t st);
