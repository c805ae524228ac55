// wq4_common.cuh -- shared declarations for the sm_100a Q4_0 kernels.
//
// Device weight layout ("panels").  The GGUF layout (18-byte blocks, f16 scale interleaved
// with nibbles, src/gguf/shader.wgsl:7-21) is only 2-byte aligned and hostile to 16-byte
// loads, bulk copies and TMA, so wq4_tensor_create repacks once on the device into
//
//   nib[p][kb][r] : uint4   16 bytes = the 32 nibbles of block (row = 128*p + r, kb)
//   scl[p][kb][r] : __half  the block's f16 scale, bit-identical to the GGUF scale
//
// with p < NP = ceil(rows/128), kb < KBP = round_up(K/32, 4), r < 128.  A (panel, kb) slab of
// 128 blocks is 2 KiB contiguous, so one warp reads 512 contiguous bytes and one K-stage of
// the GEMM (two kb) is a single 4 KiB bulk copy.  Padding rows / padding kb carry scale 0 and
// nibble 8.  Within a block, 32-bit word c (0..3) holds elements 8c..8c+7; element 8c+j sits
// at nibble position (j>>1) + 4*(j&1), which makes the in-place `lop3` extraction
//   (w      & 0x000F000F) | 0x64006400 -> half2(1024+e0,    1024+e1)
//   (w      & 0x00F000F0) | 0x64006400 -> half2(1024+16*e2, 1024+16*e3)
//   (w >> 8 & 0x000F000F) | 0x64006400 -> half2(1024+e4,    1024+e5)
//   (w >> 8 & 0x00F000F0) | 0x64006400 -> half2(1024+16*e6, 1024+16*e7)
// produce K-consecutive pairs.  The same 18 bytes per block move as in GGUF; nothing is lost
// (wq4_tensor_export_gguf inverts the repack and is tested byte-exact).
#pragma once

#include <cuda_fp16.h>
#include <cuda_bf16.h>
#include <cuda_runtime.h>
#include <stdint.h>

#include "../../include/wq4.h"

#define WQ4_PANEL_ROWS 128
#define WQ4_BLOCK_ELEMS 32
#define WQ4_GGUF_BLOCK_BYTES 18

struct wq4_tensor {
    int64_t n, k;          // logical [N,K] (tensor.rs:22-24)
    int64_t num_blocks;    // n*k/32
    int64_t rows;          // matrix mode: n ; flat mode (k % 32 != 0): num_blocks
    int64_t kb;            // blocks per row: k/32 ; flat mode: 1
    int64_t kbp;           // kb rounded up to a multiple of 4
    int64_t np;            // ceil(rows/128)
    int64_t np_alloc;      // np rounded up to even: the CTA-pair GEMM always works on two panels (padding: scale 0)
    int matrix_mode;       // 1 when k % 32 == 0 (matmul allowed)
    int device;
    uint4 *nib;            // [np][kbp][128]
    __half *scl;           // [np][kbp][128]
    size_t device_bytes;
};

struct wq4_conv1d {
    int64_t c_out, c_in, ksize;
    int64_t kdim;          // ksize*c_in, kernel-major (layers.rs:114,120-121)
    int64_t kdim_pad;      // rounded up to a multiple of 64 (one GEMM K-stage)
    int64_t rows_pad;      // c_out rounded up to 128
    int device;
    __half *w16;           // [rows_pad][kdim_pad] f16, zero padded
    __nv_bfloat16 *w16b;   // same in bf16 (built lazily is not graph-safe, so built at create)
    float *bias;           // [c_out]
    size_t device_bytes;
};

// ---- error plumbing (host) -----------------------------------------------------------------
void wq4_set_error(const char *fmt, ...);
int32_t wq4_cuda_fail(cudaError_t e, const char *what);
void wq4_count_launch(int n = 1);
bool wq4_profile_active();              // per-kernel profiler on (bench.py roofline leg): launches must stay individual
int *wq4_error_flag_ptr(int device);   // device int, sticky, lazily allocated per device
int wq4_sm_count(int device);

// Profiler scope: when wq4_profile_enable(1) is set, brackets the launches made inside its lifetime
// with a CUDA event pair on `st` and books `work` (flops or bytes) under `kind`.
enum { WQ4_PROF_GEMM_Q4 = 0, WQ4_PROF_GEMM_DENSE = 1, WQ4_PROF_GEMV = 2, WQ4_PROF_ATTN_ENC = 3, WQ4_PROF_LAYERNORM = 4,
       WQ4_PROF_ATTN_DEC = 5, WQ4_PROF_LOGITS = 6, WQ4_PROF_OTHER = 7 };
struct WqProfScope {
    bool active;
    cudaStream_t stream;
    WqProfScope(int kind, double work, cudaStream_t st);
    ~WqProfScope();
};

#define WQ4_CUDA(call)                                            \
    do {                                                          \
        cudaError_t e__ = (call);                                 \
        if (e__ != cudaSuccess) return wq4_cuda_fail(e__, #call); \
    } while (0)

struct DeviceGuard {
    int prev = -1;
    bool ok = true;
    explicit DeviceGuard(int dev)
    {
        if (cudaGetDevice(&prev) != cudaSuccess) { ok = false; return; }
        if (prev != dev && cudaSetDevice(dev) != cudaSuccess) ok = false;
        target = dev;
    }
    ~DeviceGuard()
    {
        if (prev >= 0 && prev != target) cudaSetDevice(prev);
    }
    int target = -1;
};

// ---- kernel launch entry points (implemented in the .cu files) -------------------------------
struct LinearArgs {
    const wq4_tensor *w;
    const float *bias;       // [N] or null
    const float *x_f32;      // [rows,K] or null
    const void *x_f16;       // [rows,K] f16/bf16 or null
    float *y_f32;            // [rows,N] or null
    void *y_f16;             // [rows,N] or null
    const float *residual;   // [rows,N] or null
    int64_t rows;
    uint32_t flags;
    void *workspace;
    size_t workspace_bytes;
    cudaStream_t stream;
};

int32_t wq4_launch_gemv(const LinearArgs &a);
int32_t wq4_launch_gemm_q4(const LinearArgs &a, int tile_m);
// Skinny linear (wq4_skinny.cu): 1..64 activation rows on mma.sync, 16 weight rows per CTA (default dispatch: 9..16 rows).
// Optional extras: LayerNorm fused on load (f32 x only) and the K/V thirds of a fused q|k|v projection appended to the f16
// self-attention cache.
#define WQ4_SKINNY_MAX_ROWS 64
#define WQ4_SKINNY_DEFAULT_ROWS 16
struct SkinnyExtra {
    const float *ln_w, *ln_b;          // LayerNorm weight / bias over K, or null
    void *kc, *vc;                     // f16 caches [clip][ctx][kv_ld] or null; row m of x belongs to clip m
    int64_t kv_batch_stride, kv_ld, kv_pos, kv_d;
    const int *kv_slot;                // device: cache slot of row m (active-set compaction of a decode batch), or null = m
};
int32_t wq4_launch_skinny(const LinearArgs &a, const SkinnyExtra *ex);
bool wq4_skinny_supported(const wq4_tensor *w, int64_t rows);   // matrix-mode weight, 1..64 rows
int wq4_pick_tile_m(int64_t rows, int64_t np);
// second-generation Q4 GEMM (CTA pair, A operand in tensor memory; wq4_gemm2.cu)
int32_t wq4_launch_gemm2_q4(const LinearArgs &a, int tile_m);
int wq4_pick_tile_m2(int64_t rows, int64_t npairs, int device);
int wq4_gemm_impl();   // 2 (default) or 1 (WQ4_GEMM_IMPL=1: first-generation kernel, kept for A/B runs)
int wq4_gemm_cluster_size(int64_t np);

// Chain of Q4 GEMMs in one persistent launch with row-block dependencies and the LayerNorm between two of them folded
// into the consumer's epilogue (wq4_chain.cu).  Problem i reads x16 [rows][K_i] f16 and may depend on problem dep < i
// (its x16 / residual / stats_in are that problem's outputs); dep = -1: inputs come from the predecessor kernel.
#define WQ4_CHAIN_MAX 8
struct ChainGemm {
    const wq4_tensor *w;
    const float *bias;        // [N]: bias, or b' = bias + W.beta when cvec is given
    const float *cvec;        // [N]: W.gamma -> the epilogue applies r*(acc - mu*c) + b' with (mu, r) from stats_in; null = plain
    const void *x16;          // f16 [rows][K]
    const float *residual;    // f32 [rows][N] or null (may alias y32)
    float *y32;               // f32 [rows][N] or null
    void *y16;                // f16 [rows][N] or null
    const float *y16_scale;   // [N] multiplied into y16 (gamma of the LayerNorm the consumer folds) or null
    void *stats_out;          // float2 [N/128][rows]: per 128-column group (sum, sum of squares) of the f32 result rows, or null
    const void *stats_in;     // float2 [stats_groups][rows] written by problem `dep`
    int stats_groups;
    int dep;
    uint32_t flags;           // WQ4_EPI_GELU
    int k_stage0, k_stages;   // K range in stages of 128 (k_stages = 0: all of K): the two halves of a split-K weight are two problems
    int edep_plus1;           // 1 + problem whose row block must be complete before this problem's epilogue reads `residual` (0: none)
};
size_t wq4_chain_counter_words(int64_t rows);      // unsigned words of the zero-initialised counter buffer a launch needs
int wq4_chain_pick_tile_m(int64_t rows);
int32_t wq4_launch_chain(const ChainGemm *g, int n, int64_t rows, unsigned *counters, int tile_m, cudaStream_t stream);

// Tensor-core GEMM on a dense 16-bit weight (the Conv1D stem).  The activation operand is a
// strided 3-D view [x_batch][x_rows][x_k] (element strides x_batch_stride / x_row_stride / 1), which
// is how the im2col matrix of src/model/layers.rs:93-116 is expressed without materialising it:
// with the input stored time-major [B][T+2p][C], row (b,t) of the im2col matrix is the contiguous
// window of ksize*C elements starting at (t*stride)*C.
struct DenseGemmArgs {
    const void *w16;         // [n_pad][kdim_pad] f16/bf16, K-major, zero padded
    const void *x16;         // base of the activation view
    const float *bias;       // [n] or null
    const float *residual;   // optional, element strides (rb, ot, on)
    float *y_f32;            // element strides (ob, ot, on)
    void *y_f16;
    int64_t n, n_pad, kdim_pad;
    int64_t x_k, x_rows, x_batch, x_row_stride, x_batch_stride;
    int64_t ob, ot, on, rb;
    uint32_t flags;
    int device;
    cudaStream_t stream;
};
int32_t wq4_launch_gemm_dense(const DenseGemmArgs &a, int tile_m);

// PDL on the encoder chain (GEMM, LayerNorm, attention); WQ4_PDL=0 turns the launch attribute off (A/B runs).
bool wq4_pdl_enabled();

int32_t wq4_launch_cvt_f32_to_16(const float *src, void *dst, int64_t n_elems, bool bf16,
                                 cudaStream_t stream);

// ---- small device helpers ----------------------------------------------------------------------
__device__ __forceinline__ float wq4_gelu_tanh(float h)
{
    // src/model/layers.rs:35-41: 0.5*h*(1 + tanh(u)), u = sqrt(2/pi)*(h + 0.044715 h^3).
    // 0.5*(1 + tanh(u)) == 1/(1 + exp(-2u)) == 1/(1 + 2^(h*(k0 + k1*h^2))), k0 = -2*log2(e)*sqrt(2/pi):
    // 3 FMA-pipe ops + ex2 + add + rcp + mul, relative error ~1e-6, exact limits (0 for h -> -inf, h for h -> +inf).
    const float k0 = -2.3022081985f, k1 = k0 * 0.044715f;
    const float a = h * fmaf(h * h, k1, k0);
    float e, r;
    asm("ex2.approx.ftz.f32 %0, %1;" : "=f"(e) : "f"(a));
    asm("rcp.approx.ftz.f32 %0, %1;" : "=f"(r) : "f"(e + 1.0f));
    return h * r;
}

// Undo the nibble interleave of one 32-bit word: returns the 8 element nibbles e0..e7
// (4 bits each, e0 in bits 0-3) given the device word.
__host__ __device__ __forceinline__ uint32_t wq4_word_deinterleave(uint32_t w)
{
    uint32_t out = 0;
#pragma unroll
    for (int j = 0; j < 8; j++) {
        int pos = (j >> 1) + 4 * (j & 1);
        out |= ((w >> (4 * pos)) & 0xFu) << (4 * j);
    }
    return out;
}
__host__ __device__ __forceinline__ uint32_t wq4_word_interleave(uint32_t e)
{
    uint32_t out = 0;
#pragma unroll
    for (int j = 0; j < 8; j++) {
        int pos = (j >> 1) + 4 * (j & 1);
        out |= ((e >> (4 * j)) & 0xFu) << (4 * pos);
    }
    return out;
}
