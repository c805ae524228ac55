//! Safe wrappers with the reference's names and error behaviour.
//!
//! `Q4Tensor::from_q4_bytes` returns `anyhow::Result` like src/gguf/tensor.rs:35; the forward
//! calls panic on a non-zero status like the `assert_eq!`/`expect` of src/gguf/op.rs:53,58,106.
//! Device buffers are plain `*mut f32` device pointers here; INTEGRATION.md shows how the
//! reference's `Tensor<Wgpu, 3>` call sites map onto them.
pub mod sys;

use anyhow::{bail, Result};
use std::ffi::CStr;
use std::os::raw::c_void;
use std::ptr;

fn last_error() -> String {
    unsafe { CStr::from_ptr(sys::wq4_last_error()).to_string_lossy().into_owned() }
}

fn expect_ok(status: i32, what: &str) {
    assert!(status == sys::WQ4_OK, "{what} failed: {}", last_error());
}

/// Caller-owned device scratch (any allocator: cudaMalloc, a pool, a framework tensor).
#[derive(Clone, Copy)]
pub struct Workspace {
    pub ptr: *mut c_void,
    pub bytes: usize,
}

/// A Q4_0 quantised weight `[N, K]` living on the GPU (src/gguf/tensor.rs:21-27).
pub struct Q4Tensor {
    raw: *mut sys::wq4_tensor,
    shape: [usize; 2],
}
unsafe impl Send for Q4Tensor {}

impl Q4Tensor {
    pub fn from_q4_bytes(raw_bytes: &[u8], shape: [usize; 2], device: i32, stream: sys::wq4_stream) -> Result<Self> {
        let mut out = ptr::null_mut();
        let st = unsafe {
            sys::wq4_tensor_create(raw_bytes.as_ptr(), raw_bytes.len(), shape[0] as i64, shape[1] as i64,
                                   device, stream, &mut out)
        };
        if st != sys::WQ4_OK {
            bail!("{}", last_error());
        }
        Ok(Self { raw: out, shape })
    }
    pub fn shape(&self) -> [usize; 2] { self.shape }
    pub fn num_blocks(&self) -> usize { unsafe { sys::wq4_tensor_num_blocks(self.raw) as usize } }
    /// `dst_dev`: `n*k` f32 on the device; bit-identical to the reference's CPU dequantiser.
    pub fn dequantize(&self, dst_dev: *mut f32, stream: sys::wq4_stream) {
        expect_ok(unsafe { sys::wq4_tensor_dequantize(self.raw, dst_dev, stream) }, "Q4 dequantize");
    }
    pub fn as_ptr(&self) -> *const sys::wq4_tensor { self.raw }
}
impl Drop for Q4Tensor {
    fn drop(&mut self) { unsafe { sys::wq4_tensor_destroy(self.raw); } }
}

/// `output[B,M,N] = input[B,M,K] x weights[N,K]^T` (src/gguf/op.rs:47).
pub fn q4_matmul(x_dev: *const f32, dims: [usize; 3], weights: &Q4Tensor, y_dev: *mut f32, ws: Workspace,
                 stream: sys::wq4_stream) {
    let [b, m, k] = dims;
    let [_n, wk] = weights.shape();
    assert_eq!(k, wk, "K dimension mismatch: input has {k}, weights have {wk}");
    expect_ok(unsafe { sys::wq4_matmul(weights.raw, x_dev, y_dev, b as i64, m as i64, ws.ptr, ws.bytes, stream) },
              "Q4 matmul kernel launch");
}

/// src/gguf/linear.rs:17-40; the bias is a device pointer to `[N]` f32 and is fused into the kernel.
pub struct Q4Linear {
    weights: Q4Tensor,
    bias: Option<*const f32>,
}
impl Q4Linear {
    pub fn new(weights: Q4Tensor, bias: Option<*const f32>) -> Self { Self { weights, bias } }
    pub fn forward(&self, x_dev: *const f32, dims: [usize; 3], y_dev: *mut f32, ws: Workspace, stream: sys::wq4_stream) {
        let [b, m, k] = dims;
        let [_n, wk] = self.weights.shape();
        assert_eq!(k, wk, "K dimension mismatch: input has {k}, weights have {wk}");
        expect_ok(unsafe {
            sys::wq4_linear_forward(self.weights.raw, self.bias.unwrap_or(ptr::null()), x_dev, y_dev, b as i64,
                                    m as i64, ws.ptr, ws.bytes, stream)
        }, "Q4Linear forward");
    }
    pub fn workspace_bytes(&self, b: usize, m: usize) -> usize {
        unsafe { sys::wq4_linear_workspace_bytes(self.weights.raw, b as i64, m as i64) }
    }
    pub fn weights(&self) -> &Q4Tensor { &self.weights }
    pub fn bias(&self) -> Option<*const f32> { self.bias }
}

/// src/model/layers.rs:44-59: fc1 -> tanh-GELU -> fc2 in one call.
pub struct Q4FFN {
    fc1: Q4Linear,
    fc2: Q4Linear,
}
impl Q4FFN {
    pub fn new(fc1: Q4Linear, fc2: Q4Linear) -> Self { Self { fc1, fc2 } }
    pub fn forward(&self, x_dev: *const f32, dims: [usize; 3], y_dev: *mut f32, ws: Workspace, stream: sys::wq4_stream) {
        let [b, m, _k] = dims;
        expect_ok(unsafe {
            sys::wq4_ffn_forward(self.fc1.weights.raw, self.fc1.bias.unwrap_or(ptr::null()), self.fc2.weights.raw,
                                 self.fc2.bias.unwrap_or(ptr::null()), x_dev, y_dev, b as i64, m as i64, ws.ptr,
                                 ws.bytes, stream)
        }, "Q4FFN forward");
    }
    pub fn workspace_bytes(&self, b: usize, m: usize) -> usize {
        unsafe { sys::wq4_ffn_workspace_bytes(self.fc1.weights.raw, self.fc2.weights.raw, b as i64, m as i64) }
    }
    /// No counterpart in the reference: `forward` for b*m <= 8 rows and a hidden width <= 8192 with the hidden block in a
    /// static device array of the library -- no workspace, no device allocation (usable under CUDA graph capture).
    /// Calls on one GPU share that array: do not overlap two of them on unordered streams.
    pub fn forward_static(&self, x_dev: *const f32, dims: [usize; 3], y_dev: *mut f32, stream: sys::wq4_stream) {
        let [b, m, _k] = dims;
        expect_ok(unsafe {
            sys::wq4_ffn_forward_static(self.fc1.weights.raw, self.fc1.bias.unwrap_or(ptr::null()), self.fc2.weights.raw,
                                        self.fc2.bias.unwrap_or(ptr::null()), x_dev, y_dev, b as i64, m as i64, stream)
        }, "Q4FFN forward_static");
    }
}

/// src/model/layers.rs:61-133: weight `[out, in, k]` f32 (host), bias `[out]` f32 (host).
pub struct Conv1D {
    raw: *mut sys::wq4_conv1d,
    dims: [usize; 3],
}
unsafe impl Send for Conv1D {}
impl Conv1D {
    pub fn new(weight: &[f32], bias: &[f32], dims: [usize; 3], device: i32, stream: sys::wq4_stream) -> Result<Self> {
        let [c_out, c_in, k] = dims;
        if weight.len() != c_out * c_in * k || bias.len() != c_out {
            bail!("Conv1D weight/bias size mismatch");
        }
        let mut out = ptr::null_mut();
        let st = unsafe {
            sys::wq4_conv1d_create(weight.as_ptr(), bias.as_ptr(), c_out as i64, c_in as i64, k as i64, device, stream, &mut out)
        };
        if st != sys::WQ4_OK {
            bail!("{}", last_error());
        }
        Ok(Self { raw: out, dims })
    }
    /// x `[B, C_in, T]` -> y `[B, C_out, T_out]`.
    pub fn forward(&self, x_dev: *const f32, b: usize, t: usize, padding: usize, stride: usize, y_dev: *mut f32,
                   ws: Workspace, stream: sys::wq4_stream) {
        expect_ok(unsafe {
            sys::wq4_conv1d_forward(self.raw, x_dev, y_dev, b as i64, t as i64, padding as i64, stride as i64, 0,
                                    ws.ptr, ws.bytes, stream)
        }, "Conv1D forward");
    }
    pub fn dims(&self) -> [usize; 3] { self.dims }
}
impl Drop for Conv1D {
    fn drop(&mut self) { unsafe { sys::wq4_conv1d_destroy(self.raw); } }
}

// ---------------------------------------------------------------------------------------------
// Model level: the reference's names for the callers of the Q4 path, over the batched C ABI.
// ---------------------------------------------------------------------------------------------

/// Whisper token budget of one transcription (src/model/whisper.rs:20).
pub const MAX_TOKENS: usize = 224;

/// `WhisperConfig` (src/model/config.rs:5-28); `large_v3()` / `medium()` as config.rs:32-63.
pub type WhisperConfig = sys::wq4_whisper_config;

pub fn whisper_config(preset: &str) -> Result<WhisperConfig> {
    let name = std::ffi::CString::new(preset)?;
    let mut cfg: WhisperConfig = unsafe { std::mem::zeroed() };
    if unsafe { sys::wq4_whisper_config_preset(name.as_ptr(), &mut cfg) } != sys::WQ4_OK {
        bail!("{}", last_error());
    }
    Ok(cfg)
}

/// `WhisperModel` (src/model/whisper.rs:22-128): encoder + decoder with in-place f16 KV caches, batched.
pub struct WhisperModel {
    raw: *mut sys::wq4_whisper,
}
unsafe impl Send for WhisperModel {}

impl WhisperModel {
    /// `load_whisper_from_gguf` (src/gguf/loader.rs:26-46): the file is mapped and uploaded tensor by tensor.
    pub fn load_from_gguf(path: &std::path::Path, config: &WhisperConfig, device: i32) -> Result<Self> {
        let p = std::ffi::CString::new(path.to_string_lossy().as_bytes())?;
        let mut out = ptr::null_mut();
        if unsafe { sys::wq4_whisper_load_gguf(p.as_ptr(), config, device, &mut out) } != sys::WQ4_OK {
            bail!("Failed to load GGUF model: {}", last_error());
        }
        Ok(Self { raw: out })
    }
    /// `encode` (whisper.rs:41-43): mel `[B, n_mels, 3000]` on the device -> `[B, 1500, D]` (optional copy out).
    pub fn encode(&self, mel_dev: *const f32, batch: usize, enc_out_dev: *mut f32, stream: sys::wq4_stream) {
        expect_ok(unsafe { sys::wq4_whisper_encode(self.raw, mel_dev, batch as i64, enc_out_dev, stream) }, "encode");
    }
    /// `transcribe` (whisper.rs:51-128) for a batch of clips; `lang` = language token per clip, `None` = auto-detect
    /// (one mode per call).  Returns the generated ids per clip.
    pub fn transcribe(&self, mel_dev: *const f32, batch: usize, lang: Option<&[i32]>, stream: sys::wq4_stream) -> Vec<Vec<i32>> {
        let langs: Vec<i32> = match lang { Some(l) => l.to_vec(), None => vec![-1; batch] };
        let mut toks = vec![0i32; batch * MAX_TOKENS];
        let mut counts = vec![0i32; batch];
        expect_ok(unsafe {
            sys::wq4_whisper_transcribe(self.raw, mel_dev, batch as i64, langs.as_ptr(), toks.as_mut_ptr(),
                                        counts.as_mut_ptr(), ptr::null_mut(), stream)
        }, "transcribe");
        (0..batch).map(|b| toks[b * MAX_TOKENS..b * MAX_TOKENS + counts[b] as usize].to_vec()).collect()
    }
    pub fn as_ptr(&self) -> *mut sys::wq4_whisper { self.raw }
}
impl Drop for WhisperModel {
    fn drop(&mut self) { unsafe { sys::wq4_whisper_destroy(self.raw); } }
}

/// `MelSpectrogram` (src/audio/mel.rs:45-157) on the device, batched.
pub struct MelSpectrogram {
    raw: *mut sys::wq4_logmel,
}
unsafe impl Send for MelSpectrogram {}
impl MelSpectrogram {
    pub fn new(n_mels: usize, device: i32, stream: sys::wq4_stream) -> Result<Self> {
        let mut out = ptr::null_mut();
        if unsafe { sys::wq4_logmel_create(n_mels as i32, device, stream, &mut out) } != sys::WQ4_OK {
            bail!("{}", last_error());
        }
        Ok(Self { raw: out })
    }
    pub fn as_ptr(&self) -> *const sys::wq4_logmel { self.raw }
}
impl Drop for MelSpectrogram {
    fn drop(&mut self) { unsafe { sys::wq4_logmel_destroy(self.raw); } }
}

/// `InferenceState` (src/transcribe.rs:26-31); the tokenizer stays with the caller.
pub struct InferenceState {
    pub model: WhisperModel,
    pub mel: MelSpectrogram,
    pub device: i32,
    pub n_mels: usize,
}

/// `transcribe` (src/transcribe.rs:34-107) for a batch of utterances already resampled to 16 kHz and resident on the
/// device: clip `b` = `pcm_dev[b*pcm_stride ..][..n_samples[b]]`.  Returns the text-token ids per clip (ids >= 50257
/// dropped, :96-100) and the inference time in ms (:84-87); the caller runs `tokenizer.decode` (:102-105).
pub fn transcribe(state: &InferenceState, pcm_dev: *const f32, pcm_stride: usize, n_samples: &[i64], language: Option<&[i32]>,
                  stream: sys::wq4_stream) -> Result<(Vec<Vec<u32>>, u128)> {
    let batch = n_samples.len();
    let langs: Vec<i32> = match language { Some(l) => l.to_vec(), None => vec![-1; batch] };
    let mut text = vec![0i32; batch * MAX_TOKENS];
    let mut counts = vec![0i32; batch];
    let mut ms = 0f64;
    let st = unsafe {
        sys::wq4_transcribe(state.model.as_ptr(), state.mel.as_ptr(), pcm_dev, pcm_stride as i64, n_samples.as_ptr(),
                            batch as i64, langs.as_ptr(), text.as_mut_ptr(), counts.as_mut_ptr(), ptr::null_mut(),
                            ptr::null_mut(), &mut ms, stream)
    };
    if st != sys::WQ4_OK {
        bail!("{}", last_error());
    }
    let ids = (0..batch)
        .map(|b| text[b * MAX_TOKENS..b * MAX_TOKENS + counts[b] as usize].iter().map(|&t| t as u32).collect())
        .collect();
    Ok((ids, ms as u128))
}
