//! Raw `extern "C"` declarations, one per symbol of include/wq4.h.
#![allow(non_camel_case_types)]
use std::os::raw::{c_char, c_void};

#[repr(C)]
pub struct wq4_tensor {
    _private: [u8; 0],
}
#[repr(C)]
pub struct wq4_conv1d {
    _private: [u8; 0],
}
#[repr(C)]
pub struct wq4_logmel {
    _private: [u8; 0],
}
#[repr(C)]
pub struct wq4_whisper {
    _private: [u8; 0],
}
#[repr(C)]
#[derive(Clone, Copy, Debug, Default)]
pub struct wq4_whisper_config {
    pub n_mels: i32,
    pub n_audio_ctx: i32,
    pub n_audio_state: i32,
    pub n_audio_head: i32,
    pub n_audio_layer: i32,
    pub n_text_ctx: i32,
    pub n_text_state: i32,
    pub n_text_head: i32,
    pub n_text_layer: i32,
    pub n_vocab: i32,
    pub n_lang: i32,
}
pub type wq4_stream = *mut c_void;

pub const WQ4_OK: i32 = 0;
pub const WQ4_EPI_GELU: u32 = 1;
pub const WQ4_OPERANDS_BF16: u32 = 16;
pub const WQ4_FORCE_GEMM: u32 = 32;
pub const WQ4_FORCE_GEMV: u32 = 64;
pub const WQ4_FORCE_SKINNY: u32 = 128;

extern "C" {
    pub fn wq4_last_error() -> *const c_char;
    pub fn wq4_version() -> i32;
    pub fn wq4_device_count() -> i32;

    pub fn wq4_tensor_create(host_blocks: *const u8, nbytes: usize, n: i64, k: i64, device: i32,
                             stream: wq4_stream, out: *mut *mut wq4_tensor) -> i32;
    pub fn wq4_tensor_destroy(t: *mut wq4_tensor) -> i32;
    pub fn wq4_tensor_shape(t: *const wq4_tensor, n: *mut i64, k: *mut i64) -> i32;
    pub fn wq4_tensor_num_blocks(t: *const wq4_tensor) -> i64;
    pub fn wq4_tensor_device(t: *const wq4_tensor) -> i32;
    pub fn wq4_tensor_device_bytes(t: *const wq4_tensor) -> usize;
    pub fn wq4_tensor_dequantize(t: *const wq4_tensor, dst_dev: *mut f32, stream: wq4_stream) -> i32;
    pub fn wq4_tensor_export_gguf(t: *const wq4_tensor, dst_dev: *mut u8, stream: wq4_stream) -> i32;

    pub fn wq4_linear_workspace_bytes(w: *const wq4_tensor, b: i64, m: i64) -> usize;
    pub fn wq4_matmul(w: *const wq4_tensor, x_dev: *const f32, y_dev: *mut f32, b: i64, m: i64,
                      workspace_dev: *mut c_void, workspace_bytes: usize, stream: wq4_stream) -> i32;
    pub fn wq4_linear_forward(w: *const wq4_tensor, bias_dev: *const f32, x_dev: *const f32,
                              y_dev: *mut f32, b: i64, m: i64, workspace_dev: *mut c_void,
                              workspace_bytes: usize, stream: wq4_stream) -> i32;
    pub fn wq4_linear_forward_ex(w: *const wq4_tensor, bias_dev: *const f32, x_f32_dev: *const f32,
                                 x_f16_dev: *const c_void, y_f32_dev: *mut f32, y_f16_dev: *mut c_void,
                                 residual_dev: *const f32, rows: i64, flags: u32,
                                 workspace_dev: *mut c_void, workspace_bytes: usize,
                                 stream: wq4_stream) -> i32;

    pub fn wq4_ffn_workspace_bytes(fc1: *const wq4_tensor, fc2: *const wq4_tensor, b: i64, m: i64) -> usize;
    pub fn wq4_ffn_forward(fc1: *const wq4_tensor, b1_dev: *const f32, fc2: *const wq4_tensor,
                           b2_dev: *const f32, x_dev: *const f32, y_dev: *mut f32, b: i64, m: i64,
                           workspace_dev: *mut c_void, workspace_bytes: usize, stream: wq4_stream) -> i32;

    // no counterpart in the reference: Q4FFN::forward for <= 8 rows with the hidden block in a
    // static device array inside the library -- allocates nothing
    pub fn wq4_ffn_static_limits(max_rows: *mut i64, max_hidden: *mut i64);
    pub fn wq4_ffn_forward_static(fc1: *const wq4_tensor, b1_dev: *const f32, fc2: *const wq4_tensor,
                                  b2_dev: *const f32, x_dev: *const f32, y_dev: *mut f32, b: i64, m: i64,
                                  stream: wq4_stream) -> i32;

    pub fn wq4_conv1d_create(host_w: *const f32, host_bias: *const f32, c_out: i64, c_in: i64,
                             ksize: i64, device: i32, stream: wq4_stream,
                             out: *mut *mut wq4_conv1d) -> i32;
    pub fn wq4_conv1d_destroy(c: *mut wq4_conv1d) -> i32;
    pub fn wq4_conv1d_workspace_bytes(c: *const wq4_conv1d, b: i64, t: i64, padding: i64, stride: i64) -> usize;
    pub fn wq4_conv1d_forward(c: *const wq4_conv1d, x_dev: *const f32, y_dev: *mut f32, b: i64, t: i64,
                              padding: i64, stride: i64, flags: u32, workspace_dev: *mut c_void,
                              workspace_bytes: usize, stream: wq4_stream) -> i32;

    pub fn wq4_logmel_create(n_mels: i32, device: i32, stream: wq4_stream, out: *mut *mut wq4_logmel) -> i32;
    pub fn wq4_logmel_destroy(m: *mut wq4_logmel) -> i32;
    pub fn wq4_logmel_forward(m: *const wq4_logmel, pcm_dev: *const f32, batch: i64, n_samples: i64,
                              mel_dev: *mut f32, stream: wq4_stream) -> i32;

    pub fn wq4_logmel_forward_ragged(m: *const wq4_logmel, pcm_dev: *const f32, batch: i64, stride: i64,
                                     lens_dev: *const i64, mel_dev: *mut f32, stream: wq4_stream) -> i32;
    pub fn wq4_transcribe(m: *mut wq4_whisper, mel: *const wq4_logmel, pcm_dev: *const f32, pcm_stride: i64,
                          n_samples: *const i64, batch: i64, language: *const i32, text_tokens_out: *mut i32,
                          text_counts_out: *mut i32, raw_tokens_out: *mut i32, raw_counts_out: *mut i32,
                          elapsed_ms_out: *mut f64, stream: wq4_stream) -> i32;

    pub fn wq4_gguf_probe(data: *const u8, nbytes: usize, version: *mut u32, tensor_count: *mut u64,
                          data_offset: *mut u64) -> i32;
    pub fn wq4_gguf_tensor_info(data: *const u8, nbytes: usize, name: *const c_char, dims4: *mut u64,
                                ndims: *mut u32, dtype: *mut u32, abs_offset: *mut u64, byte_size: *mut u64) -> i32;
    pub fn wq4_whisper_config_preset(name: *const c_char, cfg: *mut wq4_whisper_config) -> i32;
    pub fn wq4_whisper_load_gguf(path: *const c_char, cfg: *const wq4_whisper_config, device: i32,
                                 out: *mut *mut wq4_whisper) -> i32;
    pub fn wq4_whisper_load_gguf_memory(data: *const u8, nbytes: usize, cfg: *const wq4_whisper_config,
                                        device: i32, out: *mut *mut wq4_whisper) -> i32;
    pub fn wq4_whisper_destroy(m: *mut wq4_whisper) -> i32;
    pub fn wq4_whisper_reserve(m: *mut wq4_whisper, max_batch: i64) -> i32;
    pub fn wq4_whisper_info(m: *const wq4_whisper, d_model: *mut i64, ffn: *mut i64, vocab: *mut i64,
                            weight_bytes: *mut i64) -> i32;
    pub fn wq4_whisper_encode(m: *mut wq4_whisper, mel_dev: *const f32, batch: i64, enc_out_dev: *mut f32,
                              stream: wq4_stream) -> i32;
    pub fn wq4_whisper_transcribe(m: *mut wq4_whisper, mel_dev: *const f32, batch: i64, lang_tokens: *const i32,
                                  tokens_out: *mut i32, counts_out: *mut i32, detected_lang_out: *mut i32,
                                  stream: wq4_stream) -> i32;
    pub fn wq4_whisper_transcribe_ex(m: *mut wq4_whisper, mel_dev: *const f32, batch: i64, lang_tokens: *const i32,
                                     max_new_tokens: *const i32, tokens_out: *mut i32, counts_out: *mut i32,
                                     detected_lang_out: *mut i32, stream: wq4_stream) -> i32;
    pub fn wq4_whisper_decoder_pass(m: *mut wq4_whisper, tokens_host: *const i32, batch: i64, q_len: i64,
                                    emb_pos0: i64, kv_pos0: i64, recompute_cross: i32, logits_dev: *mut f32,
                                    stream: wq4_stream) -> i32;
    pub fn wq4_whisper_decoder_pass_slots(m: *mut wq4_whisper, tokens_host: *const i32, slots_host: *const i32, rows: i64,
                                          emb_pos: i64, kv_pos: i64, logits_dev: *mut f32, stream: wq4_stream) -> i32;
    pub fn wq4_whisper_decode_step_device(m: *mut wq4_whisper, batch: i64, position: i64, kv_len: i64,
                                          stream: wq4_stream) -> i32;

    pub fn wq4_linear_plan(w: *const wq4_tensor, rows: i64, flags: u32, tile_m: *mut i32) -> i32;
    pub fn wq4_launch_count() -> i64;
    pub fn wq4_launch_count_reset();
    pub fn wq4_debug_f16_overflow_count(device: i32, reset: i32) -> i64;
    pub fn wq4_device_error_flag(device: i32) -> i32;
    pub fn wq4_profile_enable(on: i32);
    pub fn wq4_profile_reset();
    pub fn wq4_profile_read(kind: i32, launches: *mut i64, ms: *mut f64, work: *mut f64) -> i32;
}
