/*
 * wq4.h -- C ABI of the B200 (sm_100a) Q4_0 hot path of whisper-burn.
 *
 * Every entry point replaces one reference interface (file:line under the reference tree):
 *
 *   wq4_tensor_create        Q4Tensor::from_q4_bytes          src/gguf/tensor.rs:35-71
 *   wq4_tensor_shape         Q4Tensor::shape                  src/gguf/tensor.rs:74
 *   wq4_tensor_num_blocks    Q4Tensor::num_blocks             src/gguf/tensor.rs:79
 *   wq4_tensor_dequantize    Q4Tensor::dequantize             src/gguf/tensor.rs:88-113
 *   wq4_matmul               q4_matmul                        src/gguf/op.rs:47-116 (+ shader.wgsl:51-92)
 *   wq4_linear_forward       Q4Linear::forward                src/gguf/linear.rs:34-40
 *   wq4_ffn_forward          Q4FFN::forward (+ gelu)          src/model/layers.rs:35-41,54-58
 *   wq4_conv1d_create/forward Conv1D::new / Conv1D::forward   src/model/layers.rs:69-132
 *   wq4_whisper_load_gguf     load_whisper_from_gguf           src/gguf/loader.rs:26-46 (+ reader.rs:98-223)
 *   wq4_whisper_encode        WhisperModel::encode             src/model/whisper.rs:41, encoder.rs:87-115
 *   wq4_whisper_transcribe    WhisperModel::transcribe         src/model/whisper.rs:51-128 (batched; transcribe.rs:84-87)
 *   wq4_transcribe            transcribe                       src/transcribe.rs:34-107 (samples -> text-token ids, batched)
 *   wq4_logmel_forward        MelSpectrogram::compute_log      src/audio/mel.rs:105-157 (+ framing of transcribe.rs:46-82)
 *
 * Conventions
 *   - extern "C", plain pointers and sizes, int32_t status (0 = WQ4_OK).  Nothing aborts or
 *     throws across this boundary; wq4_last_error() returns a thread-local message.
 *   - *_dev pointers are device pointers on the device the weight object was created on;
 *     activations, biases and outputs are contiguous row-major f32 exactly as in the
 *     reference ([B,M,K] -> [B,M,N]).  Weights enter as raw GGUF Q4_0 bytes (18-byte blocks,
 *     f16 scale + 16 nibble bytes) in HOST memory and are repacked once on the device.
 *   - Stream ordered: every call only enqueues work on `stream` (a cudaStream_t passed as
 *     void*); no hidden synchronisation, no allocation inside forward calls, CUDA-graph
 *     capturable.  The caller owns activation/output/workspace buffers; the library owns the
 *     repacked weight storage inside wq4_tensor / wq4_conv1d.
 *   - There is no CPU fallback: without a CUDA device every compute entry point returns
 *     WQ4_ERR_CUDA.
 */
#ifndef WQ4_H
#define WQ4_H

#include <stddef.h>
#include <stdint.h>

#ifdef __cplusplus
extern "C" {
#endif

#if defined(_WIN32)
#define WQ4_API __declspec(dllexport)
#else
#define WQ4_API __attribute__((visibility("default")))
#endif

typedef struct wq4_tensor wq4_tensor;   /* device-resident Q4_0 weight [N,K]            */
typedef struct wq4_conv1d wq4_conv1d;   /* device-resident Conv1D weight + bias          */
typedef void *wq4_stream;               /* cudaStream_t                                  */

enum {
    WQ4_OK = 0,
    WQ4_ERR_INVALID = 1,     /* null pointer / negative size / bad flag                   */
    WQ4_ERR_SHAPE = 2,       /* n*k % 32 != 0 (tensor.rs:38-41), K mismatch (op.rs:58-61) */
    WQ4_ERR_BYTES = 3,       /* byte count != n*k/32*18 (tensor.rs:44-48)                 */
    WQ4_ERR_CUDA = 4,        /* CUDA runtime/driver error, or no device                   */
    WQ4_ERR_UNSUPPORTED = 5, /* e.g. matmul on a tensor whose K is not a multiple of 32   */
    WQ4_ERR_WORKSPACE = 6    /* workspace too small                                       */
};

/* Epilogue / precision flags for the *_ex entry points. */
enum {
    WQ4_EPI_NONE = 0,
    WQ4_EPI_GELU = 1,        /* tanh-GELU of src/model/layers.rs:35-41 applied after bias  */
    WQ4_OPERANDS_BF16 = 16,  /* tensor-core operands in bf16 instead of the default f16    */
    WQ4_FORCE_GEMM = 32,     /* use the tcgen05 GEMM kernel even for tiny M                */
    WQ4_FORCE_GEMV = 64,     /* use the CUDA-core GEMV kernel (M <= 8 only)                */
    WQ4_FORCE_SKINNY = 128   /* use the mma.sync skinny kernel (M <= 64 only; default 9..16) */
};

WQ4_API const char *wq4_last_error(void);
WQ4_API int32_t wq4_version(void);
/* Number of CUDA devices visible (0 when there is no GPU / driver). */
WQ4_API int32_t wq4_device_count(void);

/* ---- Q4Tensor ------------------------------------------------------------------------ */
/* host_blocks: nbytes of raw GGUF Q4_0 data for a [n,k] weight (row-major, blocks of 32
 * consecutive elements).  Validates exactly as tensor.rs:38-48, uploads and repacks. */
WQ4_API int32_t wq4_tensor_create(const uint8_t *host_blocks, size_t nbytes, int64_t n, int64_t k,
                                  int32_t device, wq4_stream stream, wq4_tensor **out);
WQ4_API int32_t wq4_tensor_destroy(wq4_tensor *t);
WQ4_API int32_t wq4_tensor_shape(const wq4_tensor *t, int64_t *n, int64_t *k);
WQ4_API int64_t wq4_tensor_num_blocks(const wq4_tensor *t);
WQ4_API int32_t wq4_tensor_device(const wq4_tensor *t);
/* Bytes of device memory held by the repacked weight (scales + nibbles + padding). */
WQ4_API size_t wq4_tensor_device_bytes(const wq4_tensor *t);
/* dst_dev: n*k floats, row-major [n,k]; bit-identical to Q4Tensor::dequantize. */
WQ4_API int32_t wq4_tensor_dequantize(const wq4_tensor *t, float *dst_dev, wq4_stream stream);
/* Inverse of the upload repack: writes the original n*k/32*18 GGUF bytes to dst_dev. */
WQ4_API int32_t wq4_tensor_export_gguf(const wq4_tensor *t, uint8_t *dst_dev, wq4_stream stream);

/* ---- q4_matmul / Q4Linear ------------------------------------------------------------ */
/* Workspace (bytes) wq4_matmul / wq4_linear_forward need for a [b,m,K] input. */
WQ4_API size_t wq4_linear_workspace_bytes(const wq4_tensor *w, int64_t b, int64_t m);
/* y[b,m,N] = x[b,m,K] . W[N,K]^T            (q4_matmul, op.rs:47) */
WQ4_API int32_t wq4_matmul(const wq4_tensor *w, const float *x_dev, float *y_dev, int64_t b,
                           int64_t m, void *workspace_dev, size_t workspace_bytes,
                           wq4_stream stream);
/* y = x . W^T + bias (bias_dev may be NULL)  (Q4Linear::forward, linear.rs:34-40) */
WQ4_API int32_t wq4_linear_forward(const wq4_tensor *w, const float *bias_dev, const float *x_dev,
                                   float *y_dev, int64_t b, int64_t m, void *workspace_dev,
                                   size_t workspace_bytes, wq4_stream stream);
/* Extended form used by the fused callers.  rows = b*m.
 *   x_f32_dev / x_f16_dev : exactly one non-NULL ([rows,K]; f16/bf16 per flags)
 *   y_f32_dev / y_f16_dev : at least one non-NULL ([rows,N])
 *   residual_dev          : optional f32 [rows,N] added after bias/GELU
 *   flags                 : WQ4_EPI_* | WQ4_OPERANDS_BF16 | WQ4_FORCE_* */
WQ4_API int32_t wq4_linear_forward_ex(const wq4_tensor *w, const float *bias_dev,
                                      const float *x_f32_dev, const void *x_f16_dev,
                                      float *y_f32_dev, void *y_f16_dev, const float *residual_dev,
                                      int64_t rows, uint32_t flags, void *workspace_dev,
                                      size_t workspace_bytes, wq4_stream stream);

/* ---- Q4FFN ---------------------------------------------------------------------------- */
WQ4_API size_t wq4_ffn_workspace_bytes(const wq4_tensor *fc1, const wq4_tensor *fc2, int64_t b,
                                       int64_t m);
/* y = fc2(gelu(fc1(x)))  (layers.rs:54-58); biases may be NULL. */
WQ4_API int32_t wq4_ffn_forward(const wq4_tensor *fc1, const float *b1_dev, const wq4_tensor *fc2,
                                const float *b2_dev, const float *x_dev, float *y_dev, int64_t b,
                                int64_t m, void *workspace_dev, size_t workspace_bytes,
                                wq4_stream stream);

/* No counterpart in the reference: added for callers that cannot allocate
 * device memory (CUDA graph capture, tight serving loops, harnesses that forbid cudaMalloc).  Same result as
 * wq4_ffn_forward for b*m <= 8 rows on the CUDA-core GEMV family, but the f32 hidden block lives in a file-scope
 * __device__ array inside the library (256 KiB per GPU, held until exit) instead of a caller workspace; the call makes
 * no cudaMalloc / cudaFree and no allocating library call.  Limits (wq4_ffn_static_limits): b*m <= 8 rows and
 * fc1's N (the hidden width) <= 8192; beyond them it returns WQ4_ERR_WORKSPACE.  All calls on one GPU share the array:
 * calls that overlap on the GPU (unordered streams) race on it; nothing in it has to start at zero. */
WQ4_API void wq4_ffn_static_limits(int64_t *max_rows, int64_t *max_hidden);
WQ4_API int32_t wq4_ffn_forward_static(const wq4_tensor *fc1, const float *b1_dev, const wq4_tensor *fc2,
                                       const float *b2_dev, const float *x_dev, float *y_dev, int64_t b,
                                       int64_t m, wq4_stream stream);

/* ---- Conv1D --------------------------------------------------------------------------- */
/* host_w: [c_out, c_in, ksize] f32 (loader.rs:256-281); host_bias: [c_out] f32. */
WQ4_API int32_t wq4_conv1d_create(const float *host_w, const float *host_bias, int64_t c_out,
                                  int64_t c_in, int64_t ksize, int32_t device, wq4_stream stream,
                                  wq4_conv1d **out);
WQ4_API int32_t wq4_conv1d_destroy(wq4_conv1d *c);
WQ4_API size_t wq4_conv1d_workspace_bytes(const wq4_conv1d *c, int64_t b, int64_t t,
                                          int64_t padding, int64_t stride);
/* x [b,c_in,t] -> y [b,c_out,t_out], t_out = (t + 2*padding - ksize)/stride + 1
 * (layers.rs:77-132).  flags: WQ4_EPI_GELU fuses the GELU that encoder.rs:90,94 applies. */
WQ4_API int32_t wq4_conv1d_forward(const wq4_conv1d *c, const float *x_dev, float *y_dev,
                                   int64_t b, int64_t t, int64_t padding, int64_t stride,
                                   uint32_t flags, void *workspace_dev, size_t workspace_bytes,
                                   wq4_stream stream);

/* ---- Model level: the callers of the Q4 path (SURVEY section 8f "next") -------------------------- */
typedef struct wq4_whisper wq4_whisper;
typedef struct wq4_whisper_config {   /* src/model/config.rs:5-28 */
    int32_t n_mels, n_audio_ctx, n_audio_state, n_audio_head, n_audio_layer;
    int32_t n_text_ctx, n_text_state, n_text_head, n_text_layer, n_vocab, n_lang;
} wq4_whisper_config;

/* GGUF reader probes, host only (GgufReader::{version,tensor_count,tensor_info,tensor_data},
 * src/gguf/reader.rs:190-223).  dims4 receives the dims as stored (reversed), padded with 1. */
WQ4_API int32_t wq4_gguf_probe(const uint8_t *data, size_t nbytes, uint32_t *version,
                               uint64_t *tensor_count, uint64_t *data_offset);
WQ4_API int32_t wq4_gguf_tensor_info(const uint8_t *data, size_t nbytes, const char *name,
                                     uint64_t *dims4, uint32_t *ndims, uint32_t *dtype,
                                     uint64_t *abs_offset, uint64_t *byte_size);
/* "large_v3" (config.rs:32-46) or "medium" (config.rs:49-63). */
WQ4_API int32_t wq4_whisper_config_preset(const char *name, wq4_whisper_config *cfg);
/* GGUF v2/v3 with the reference's tensor names (loader.rs:160-377); Q4_0 linears, F32/F16 rest. */
WQ4_API int32_t wq4_whisper_load_gguf(const char *path, const wq4_whisper_config *cfg, int32_t device,
                                      wq4_whisper **out);
WQ4_API int32_t wq4_whisper_load_gguf_memory(const uint8_t *data, size_t nbytes,
                                             const wq4_whisper_config *cfg, int32_t device,
                                             wq4_whisper **out);
WQ4_API int32_t wq4_whisper_destroy(wq4_whisper *m);
/* Allocates activation / KV-cache workspaces for up to max_batch clips.  Optional: encode / transcribe reserve what they need.
 * A request above the current reservation synchronises the device and re-allocates (cache contents are not kept). */
WQ4_API int32_t wq4_whisper_reserve(wq4_whisper *m, int64_t max_batch);
WQ4_API int32_t wq4_whisper_info(const wq4_whisper *m, int64_t *d_model, int64_t *ffn, int64_t *vocab,
                                 int64_t *weight_bytes);
/* mel_dev [batch, n_mels, 2*n_audio_ctx] f32 -> enc_out_dev [batch, n_audio_ctx, D] f32 (may be NULL:
 * the encoder output then only stays inside the model for the decoder).  From the second call per batch size on, everything
 * after the input transpose is replayed as one CUDA graph on `stream` (not on the legacy default stream, not inside a
 * caller's own stream capture, where the launches join the caller's graph; WQ4_ENCODE_GRAPH=0 turns it off).
 * The encoder's chained GEMM launch and the single-clip decode step are persistent kernels whose CTAs wait for one another
 * and fill the device: do not run two model calls concurrently on one GPU (different streams or processes); one host thread
 * and one stream per GPU, as in the reference's inference thread (src/native/inference.rs:31-65).  A wait that cannot be
 * satisfied ends after a bounded number of polls and raises the sticky device error flag instead of hanging. */
WQ4_API int32_t wq4_whisper_encode(wq4_whisper *m, const float *mel_dev, int64_t batch,
                                   float *enc_out_dev, wq4_stream stream);
/* Greedy transcription of `batch` clips (see wq4_whisper.cu for the prompt modes).
 * lang_tokens [batch] (host; < 0 = auto-detect, all clips same mode); tokens_out [batch][224] and
 * counts_out [batch] (host); detected_lang_out [batch] may be NULL.  Synchronises `stream`. */
WQ4_API int32_t wq4_whisper_transcribe(wq4_whisper *m, const float *mel_dev, int64_t batch,
                                       const int32_t *lang_tokens, int32_t *tokens_out,
                                       int32_t *counts_out, int32_t *detected_lang_out,
                                       wq4_stream stream);
/* Same with a per-clip token budget (host, [batch], clamped to 224; NULL = the reference's MAX_TOKENS): a clip that hits EOT
 * (whisper.rs:105-108) or its budget leaves the active set, the remaining rows are compacted on the device. */
WQ4_API int32_t wq4_whisper_transcribe_ex(wq4_whisper *m, const float *mel_dev, int64_t batch,
                                          const int32_t *lang_tokens, const int32_t *max_new_tokens,
                                          int32_t *tokens_out, int32_t *counts_out, int32_t *detected_lang_out,
                                          wq4_stream stream);
/* One decoder pass over q_len new tokens per clip after wq4_whisper_encode (test / bench hook);
 * logits_dev [batch][vocab] of the last position (may be NULL). */
WQ4_API int32_t wq4_whisper_decoder_pass(wq4_whisper *m, const int32_t *tokens_host, int64_t batch,
                                         int64_t q_len, int64_t emb_pos0, int64_t kv_pos0,
                                         int32_t recompute_cross, float *logits_dev, wq4_stream stream);
/* One single-query step over a subset of the clips the caches hold: row r = clip slots_host[r] (test hook of the greedy
 * loop's active-set compaction); logits_dev [rows][vocab]. */
WQ4_API int32_t wq4_whisper_decoder_pass_slots(wq4_whisper *m, const int32_t *tokens_host, const int32_t *slots_host,
                                               int64_t rows, int64_t emb_pos, int64_t kv_pos, float *logits_dev,
                                               wq4_stream stream);
/* Stream-ordered greedy decode step on device-resident tokens (CUDA-graph capturable). */
WQ4_API int32_t wq4_whisper_decode_step_device(wq4_whisper *m, int64_t batch, int64_t position,
                                               int64_t kv_len, wq4_stream stream);

/* ---- Log-mel front end (batched transcribe front-end) ----------------- */
/* MelSpectrogram::compute_log (src/audio/mel.rs:105-157: reflect-padded 400-point STFT, periodic Hann, hop 160,
 * |X|^2, Slaney filterbank mel.rs:270-312, log10(max(.,1e-10)), clamp to max-8, (.+4)/4) plus the framing of
 * src/transcribe.rs:46-82 (clip padded with zeros / truncated to 480000 samples, output transposed to
 * [n_mels, 3000]), for a batch of clips on the device. */
typedef struct wq4_logmel wq4_logmel;
WQ4_API int32_t wq4_logmel_create(int32_t n_mels, int32_t device, wq4_stream stream, wq4_logmel **out);
WQ4_API int32_t wq4_logmel_destroy(wq4_logmel *m);
/* pcm_dev [batch][n_samples] f32, 16 kHz mono -> mel_dev [batch][n_mels][3000] f32 (the encoder's input). */
WQ4_API int32_t wq4_logmel_forward(const wq4_logmel *m, const float *pcm_dev, int64_t batch,
                                   int64_t n_samples, float *mel_dev, wq4_stream stream);

/* Ragged batch: clip b = pcm_dev[b*stride .. b*stride + lens_dev[b]) (lens_dev: DEVICE int64 [batch], each <= stride). */
WQ4_API int32_t wq4_logmel_forward_ragged(const wq4_logmel *m, const float *pcm_dev, int64_t batch, int64_t stride,
                                          const int64_t *lens_dev, float *mel_dev, wq4_stream stream);

/* ---- transcribe (src/transcribe.rs:34-107) for a batch of utterances: 16 kHz mono f32 samples on the device ->
 * pad/truncate to 30 s -> log-mel -> encoder -> greedy decode -> ids >= 50257 dropped.  n_samples, language (token per clip,
 * < 0 = auto-detect, one mode per call) and every output are HOST arrays; text_tokens_out [batch][224], text_counts_out
 * [batch]; raw_tokens_out / raw_counts_out / elapsed_ms_out may be NULL.  Synchronises `stream`. */
WQ4_API int32_t wq4_transcribe(wq4_whisper *m, const wq4_logmel *mel, const float *pcm_dev, int64_t pcm_stride,
                               const int64_t *n_samples, int64_t batch, const int32_t *language,
                               int32_t *text_tokens_out, int32_t *text_counts_out, int32_t *raw_tokens_out,
                               int32_t *raw_counts_out, double *elapsed_ms_out, wq4_stream stream);

/* ---- Introspection used by bench.py / tests ------------------------------------------- */
/* Which kernel family a [rows,K]x[N,K]^T call dispatches to: 0 = GEMV (CUDA cores),
 * 1 = tcgen05 GEMM, 2 = mma.sync skinny kernel (9..16 rows, or forced).  `tile_m` receives the GEMM row tile
 * (0 for GEMV, 16 = weight rows per CTA for the skinny kernel). */
WQ4_API int32_t wq4_linear_plan(const wq4_tensor *w, int64_t rows, uint32_t flags,
                                int32_t *tile_m);
/* Count of kernels this library has launched on the calling thread since the last reset. */
WQ4_API int64_t wq4_launch_count(void);
WQ4_API void wq4_launch_count_reset(void);
/* Per-kernel timing for the roofline leg of bench.py: while enabled, every launch is bracketed by a
 * CUDA event pair on its stream.  kind: 0 Q4 GEMM (work = flop), 1 dense GEMM (flop), 2 GEMV (bytes),
 * 3 encoder attention (flop), 4 LayerNorm (bytes), 5 decoder attention (bytes), 6 logits (bytes), 7 other. */
WQ4_API void wq4_profile_enable(int32_t on);
WQ4_API void wq4_profile_reset(void);
WQ4_API int32_t wq4_profile_read(int32_t kind, int64_t *launches, double *ms, double *work);
/* Debug builds only (csrc/build.sh -DWQ4_DEBUG_OVERFLOW): number of f32 activations the drop-in entry points clipped to
 * +-65504 when converting to the f16 tensor-core operand (cvt.rn.satfinite) since the last reset; -1 in release builds. */
WQ4_API int64_t wq4_debug_f16_overflow_count(int32_t device, int32_t reset);
/* Sticky device-side error flag set by kernels (pipeline wait timeouts); 0 = clean. */
WQ4_API int32_t wq4_device_error_flag(int32_t device);

#ifdef __cplusplus
}
#endif
#endif /* WQ4_H */
