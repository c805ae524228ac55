// wq4_model.cuh -- launchers of the non-GEMM kernels used by the encoder/decoder host code.
#pragma once
#include <cuda_runtime.h>
#include <stdint.h>
#include <cuda_fp16.h>
struct wq4_tensor;

int32_t wq4m_layernorm(const float *x, const float *w, const float *b, float *y32, void *y16, int64_t rows, int64_t D,
                       cudaStream_t st);
// K/V addressing of the decode attention kernels: element (clip b, head h, key t, dim d) sits at
//   base + b*kv_batch_stride + h*hstride + t*ldkv + d.   Self-attention cache [B][ctx][D]: ldkv = D, hstride = 64.
//   Cross-attention K/V, head-major [B][H][T][128] (64 K | 64 V per key, wq4m_cross_kv_pack): ldkv = 128, hstride = T*128, V = K + 64.
int32_t wq4m_cross_kv_pack(const void *src16, void *dst16, int64_t B, int64_t T, int64_t H, cudaStream_t st);
int32_t wq4m_dec_attention_split(const float *q, int64_t ldq, const void *kc, const void *vc, int64_t ldkv, int64_t kv_batch_stride,
                                 int64_t hstride, float *out, void *out16, int64_t ldo, int64_t B, int64_t H, int64_t kv_len, float *part,
                                 int64_t part_floats, int *cnt, int device, cudaStream_t st, int64_t append_off = 0, const int *slot = nullptr);
// slot: device table, row b of q / out reads and appends the caches of clip slot[b] (active-set compaction); null = b.
// append_off < 0: K/V are static (cross-attention): their first ring stages are requested before the programmatic-launch wait.
// append_off > 0: q points into a fused q|k|v row and the kernel first stores the new key / value (q + append_off,
// q + 2*append_off) of every (clip, head) at cache row kv_len - 1 -- the cache append of a decode step without its own launch.
int32_t wq4m_enc_attention(const void *qkv16, void *out16, int64_t B, int64_t T, int64_t D, int64_t H, cudaStream_t st);
// tcgen05 implementation (wq4_attn.cu); wq4m_enc_attention dispatches to it unless WQ4_ATTN_IMPL=1 (A/B runs).
int32_t wq4m_enc_attention_tc(const void *qkv16, void *out16, int64_t B, int64_t T, int64_t D, int64_t H, cudaStream_t st);
int32_t wq4m_dec_attention(const float *q, int64_t ldq, const void *kc, const void *vc, int64_t ldkv, int64_t kv_batch_stride,
                           int64_t hstride, float *out, int64_t ldo, int64_t B, int64_t H, int64_t q_len, int64_t kv_len, int causal_base,
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

// Greedy-loop bookkeeping on the device (batches <= 1024 clips): see greedy_bookkeep_kernel.
int32_t wq4m_greedy_init(int *slot, int *n_active, int *count, int64_t B, cudaStream_t st);
int32_t wq4m_greedy_bookkeep(const int *next, int *tokens_next, int *slot, int *n_active, int *count, const int *budget,
                             int *tokens_out, int64_t max_tokens, int eot, cudaStream_t st);

// ---- decode-step megakernel (wq4_decode.cu) ------------------------------------------------------------
// Decode-layout Q4 weight: rows are dealt to the CTAs of the persistent decode kernel in contiguous ranges of
// `rpc` rows; CTA c's chunk is nib[c][kb][rpc] (uint4 = one 32-weight block, same nibble interleave as the
// panel layout) and scl[c][kb][rpc] (f16, bit-identical GGUF scales).  Rows >= n inside the last chunk carry
// scale 0.  Same 18 bytes per block as GGUF.
struct DecWeight {
    const uint4 *nib;
    const __half *scl;
    const float *bias;   // [n] or null
    int n, k, kb, rpc;
};
struct DecLayerW {
    const float *ln1w, *ln1b, *ln2w, *ln2b, *ln3w, *ln3b;
    DecWeight qkv, out, cq, cout, fc1, fc2;
};
constexpr int kDecMaxLayers = 32;
constexpr int kDecTimelineSlots = 512;
struct DecodeStepArgs {           // passed BY VALUE: the per-layer table sits in the kernel-parameter constant bank
    DecLayerW layers[kDecMaxLayers];
    int n_layers, B, D, F, H, V, ctx, T_enc;
    const int *tokens;            // [B] device
    const float *tok_emb;         // [V][D] f32 (embedding lookup)
    const uint4 *tok_emb16t;      // logits matrix, f16, transposed in 16-byte chunks: [D/8][Vp] (Vp = V rounded up to 32)
    int Vp;
    const float *pos_emb;         // [ctx][D]
    int emb_pos;                  // positional row of this step
    int kv_len;                   // self-attention cache rows valid BEFORE this step (the new row is written at kv_len)
    float *x, *q, *h, *part;      // [B][D], [B][D], [B][F], attention partials
    __half *self_k, *self_v;      // [L][B][ctx][D]
    const __half *cross_kv;       // [L][kv_clips][H][T_enc][128] head-major, 64 K | 64 V per key (wq4m_cross_kv_pack)
    const float *lnfw, *lnfb;
    float *logits;                // [B][V]
    int arg_lo, arg_hi;           // greedy pick over [lo, hi); arg_hi <= arg_lo: no pick
    const int *suppress;          // [B] token to exclude (-1 none) or null
    int *next_out;                // [B]
    float *arg_scratch;           // [grid][B][2]
    unsigned *barrier;            // grid barrier counter (zeroed before every launch)
    int *dev_err;
    const int *slot;       // [B] device: cache slot of batch row b (active-set compaction) or null = b
    int kv_clips;          // clips the KV caches were laid out for ([layer][kv_clips][ctx|T_enc][D]); >= B
    long long *timeline;   // debug: kDecTimelineSlots x int64 barrier time stamps, normally NULL
};
// Builds the decode layout of a panel-layout tensor (device to device).
int32_t wq4m_decw_create(const wq4_tensor *t, const float *bias_dev, int n_cta, DecWeight *out, void **alloc_nib, void **alloc_scl,
                         cudaStream_t st);
int wq4m_decode_grid(int device);
int32_t wq4m_decode_step(const DecodeStepArgs &a, int device, cudaStream_t st);
// f32 [V][D] -> f16 chunks [D/8][Vp] (rows >= V zero)
int32_t wq4m_logits_matrix(const float *emb, void *dst, int64_t V, int64_t Vp, int64_t D, cudaStream_t st);

// ---- decode step, second generation (wq4_decode2.cu): one cluster per attention head, three grid-wide exchanges per layer ----
struct Dec2Layer {
    const uint8_t *wA, *wB, *wC;                                  // [n_cta][sA | sB | sC] weight slabs (see wq4_decode2.cu)
    const float *ln1w, *ln1b, *ln2w, *ln2b, *ln3w, *ln3b;
    const float *b_qkv, *b_out, *b_cq, *b_cout, *b_fc1, *b_fc2;    // [N] or null
};
struct Dec2Args {                                                // passed BY VALUE (kernel-parameter constant bank)
    Dec2Layer layers[kDecMaxLayers];
    int n_layers, D, F, H, V, Vp, ctx, T_enc, CS, kbD;
    int rq, ro, rc, rf, kpc;                                      // rows per CTA of the q|k|v, out/fc2 slices, cross-q and fc1 jobs; cross keys per CTA
    uint32_t sA, sB, sC;                                          // slab bytes (multiples of 128)
    uint32_t a_scl1, a_nib2, a_scl2, b_scl1, b_nib2, b_scl2, c_scl1, c_nib2, c_scl2;   // offsets inside the slabs (nib1 at 0)
    uint32_t off_bufA;                                            // shared-memory offset of the first weight buffer
    const int *tokens;
    const float *tok_emb;
    const uint4 *tok_emb16t;
    const float *pos_emb;
    int emb_pos, kv_len;
    __half *self_k, *self_v;
    const __half *cross_kv;
    int kv_clips;
    const int *slot;                                              // [1] device: cache slot of the clip, or null = 0
    const float *lnfw, *lnfb;
    float *logits;
    int arg_lo, arg_hi;
    const int *suppress;
    int *next_out;
    float *arg_scratch;                                           // [n_cta][2]
    float *partial;                                               // [2][H][D]
    unsigned *barrier;                                            // [0] grid counter, [1] arg-max arrivals (zeroed before every launch)
    int *dev_err;
    long long *timeline;
};
bool wq4m_dec2_supported(int64_t D, int64_t F, int64_t H);
int32_t wq4m_dec2_plan(int device, int64_t D, int64_t F, int64_t H, int64_t T_enc, int *cluster_size);
void wq4m_dec2_geometry(Dec2Args &a, int D, int F, int H, int T_enc, int CS);
int32_t wq4m_dec2_layer_create(const Dec2Args &geo, const wq4_tensor *qkv, const wq4_tensor *out, const wq4_tensor *cq, const wq4_tensor *cout,
                               const wq4_tensor *fc1, const wq4_tensor *fc2, Dec2Layer *L, cudaStream_t st);
int32_t wq4m_decode2_step(const Dec2Args &a, int device, cudaStream_t st);
