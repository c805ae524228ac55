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
