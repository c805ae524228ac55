// wq4_tensor.cu -- Q4Tensor on the device: upload + repack, dequantise, export.
// Replaces src/gguf/tensor.rs:35-113 of the reference.
#include <vector>

#include "wq4_common.cuh"

namespace {

// One thread per device entry (p, kb, r), padding included.
__global__ void repack_kernel(const uint16_t *__restrict__ raw /* GGUF bytes, 2-byte units */,
                              uint4 *__restrict__ nib, __half *__restrict__ scl, int64_t rows,
                              int64_t kb_valid, int64_t kbp, int64_t entries)
{
    int64_t i = (int64_t)blockIdx.x * blockDim.x + threadIdx.x;
    if (i >= entries) return;
    int64_t r = i % WQ4_PANEL_ROWS;
    int64_t kb = (i / WQ4_PANEL_ROWS) % kbp;
    int64_t p = i / (WQ4_PANEL_ROWS * kbp);
    int64_t row = p * WQ4_PANEL_ROWS + r;
    if (row >= rows || kb >= kb_valid) {
        nib[i] = make_uint4(0x88888888u, 0x88888888u, 0x88888888u, 0x88888888u);
        scl[i] = __ushort_as_half((unsigned short)0);
        return;
    }
    const uint16_t *src = raw + (row * kb_valid + kb) * (WQ4_GGUF_BLOCK_BYTES / 2);
    uint16_t h[9];
#pragma unroll
    for (int j = 0; j < 9; j++) h[j] = src[j];
    // bytes 2..17 of the block = h[1..8]; byte (2+e) low nibble = element e, high = e+16
    uint32_t lo01 = 0, lo23 = 0, hi01 = 0, hi23 = 0;  // natural-order words (element j at bits 4j)
#pragma unroll
    for (int j = 0; j < 8; j++) {
        uint32_t b0 = (h[1 + (j >> 1)] >> (8 * (j & 1))) & 0xFFu;        // byte 2+j
        uint32_t b1 = (h[5 + (j >> 1)] >> (8 * (j & 1))) & 0xFFu;        // byte 10+j
        lo01 |= (b0 & 0xFu) << (4 * j);
        lo23 |= (b1 & 0xFu) << (4 * j);
        hi01 |= (b0 >> 4) << (4 * j);
        hi23 |= (b1 >> 4) << (4 * j);
    }
    nib[i] = make_uint4(wq4_word_interleave(lo01), wq4_word_interleave(lo23),
                        wq4_word_interleave(hi01), wq4_word_interleave(hi23));
    scl[i] = __ushort_as_half(h[0]);
}

// One thread per (entry, word): 8 consecutive f32 outputs (32 bytes).
__global__ void dequant_kernel(const uint4 *__restrict__ nib, const __half *__restrict__ scl,
                               float *__restrict__ out, int64_t rows, int64_t kb_valid,
                               int64_t kbp, int64_t entries)
{
    int64_t t = (int64_t)blockIdx.x * blockDim.x + threadIdx.x;
    int64_t i = t >> 2;
    int c = (int)(t & 3);
    if (i >= entries) return;
    int64_t r = i % WQ4_PANEL_ROWS;
    int64_t kb = (i / WQ4_PANEL_ROWS) % kbp;
    int64_t p = i / (WQ4_PANEL_ROWS * kbp);
    int64_t row = p * WQ4_PANEL_ROWS + r;
    if (row >= rows || kb >= kb_valid) return;
    const uint32_t *wp = reinterpret_cast<const uint32_t *>(nib + i);
    uint32_t e = wq4_word_deinterleave(wp[c]);
    float d = __half2float(scl[i]);  // exact (half 2.7.1 f16::to_f32, tensor.rs:98-99)
    float v[8];
#pragma unroll
    for (int j = 0; j < 8; j++) {
        float q = (float)((e >> (4 * j)) & 0xFu) - 8.0f;  // tensor.rs:104-105
        v[j] = q * d;                                       // tensor.rs:106-107
    }
    float4 *dst = reinterpret_cast<float4 *>(out + ((row * kb_valid + kb) * 32 + 8 * c));
    dst[0] = make_float4(v[0], v[1], v[2], v[3]);
    dst[1] = make_float4(v[4], v[5], v[6], v[7]);
}

__global__ void export_kernel(const uint4 *__restrict__ nib, const __half *__restrict__ scl,
                              uint16_t *__restrict__ raw, int64_t rows, int64_t kb_valid,
                              int64_t kbp, int64_t entries)
{
    int64_t i = (int64_t)blockIdx.x * blockDim.x + threadIdx.x;
    if (i >= entries) return;
    int64_t r = i % WQ4_PANEL_ROWS;
    int64_t kb = (i / WQ4_PANEL_ROWS) % kbp;
    int64_t p = i / (WQ4_PANEL_ROWS * kbp);
    int64_t row = p * WQ4_PANEL_ROWS + r;
    if (row >= rows || kb >= kb_valid) return;
    uint4 q = nib[i];
    uint32_t lo01 = wq4_word_deinterleave(q.x), lo23 = wq4_word_deinterleave(q.y);
    uint32_t hi01 = wq4_word_deinterleave(q.z), hi23 = wq4_word_deinterleave(q.w);
    uint16_t *dst = raw + (row * kb_valid + kb) * (WQ4_GGUF_BLOCK_BYTES / 2);
    dst[0] = __half_as_ushort(scl[i]);
#pragma unroll
    for (int j = 0; j < 4; j++) {
        uint32_t b0 = ((lo01 >> (8 * j)) & 0xFu) | (((hi01 >> (8 * j)) & 0xFu) << 4);          // byte 2+2j
        uint32_t b1 = ((lo01 >> (8 * j + 4)) & 0xFu) | (((hi01 >> (8 * j + 4)) & 0xFu) << 4);  // byte 3+2j
        dst[1 + j] = (uint16_t)(b0 | (b1 << 8));
        uint32_t c0 = ((lo23 >> (8 * j)) & 0xFu) | (((hi23 >> (8 * j)) & 0xFu) << 4);
        uint32_t c1 = ((lo23 >> (8 * j + 4)) & 0xFu) | (((hi23 >> (8 * j + 4)) & 0xFu) << 4);
        dst[5 + j] = (uint16_t)(c0 | (c1 << 8));
    }
}

}  // namespace

extern "C" {

int32_t wq4_tensor_create(const uint8_t *host_blocks, size_t nbytes, int64_t n, int64_t k,
                          int32_t device, wq4_stream stream_, wq4_tensor **out)
{
    if (!out) { wq4_set_error("wq4_tensor_create: out is null"); return WQ4_ERR_INVALID; }
    *out = nullptr;
    if (!host_blocks && nbytes) { wq4_set_error("wq4_tensor_create: host_blocks is null"); return WQ4_ERR_INVALID; }
    if (n <= 0 || k <= 0) { wq4_set_error("wq4_tensor_create: shape [%lld,%lld] must be positive", (long long)n, (long long)k); return WQ4_ERR_INVALID; }
    int64_t num_elements;
    if (__builtin_mul_overflow(n, k, &num_elements) || num_elements > (int64_t)1 << 56) {
        wq4_set_error("wq4_tensor_create: shape [%lld,%lld] overflows", (long long)n, (long long)k);
        return WQ4_ERR_INVALID;
    }
    if (num_elements % 32 != 0) {  // tensor.rs:38-41
        wq4_set_error("Q4_0 requires element count divisible by 32, got %lld", (long long)num_elements);
        return WQ4_ERR_SHAPE;
    }
    int64_t num_blocks = num_elements / 32;
    size_t expected = (size_t)num_blocks * WQ4_GGUF_BLOCK_BYTES;
    if (nbytes != expected) {      // tensor.rs:44-48
        wq4_set_error("Q4_0 byte count mismatch: expected %zu for %lld blocks, got %zu", expected,
                      (long long)num_blocks, nbytes);
        return WQ4_ERR_BYTES;
    }
    int ndev = 0;
    if (cudaGetDeviceCount(&ndev) != cudaSuccess || device < 0 || device >= ndev) {
        cudaGetLastError();
        wq4_set_error("wq4_tensor_create: CUDA device %d not available (%d visible); there is no CPU fallback", device, ndev);
        return WQ4_ERR_CUDA;
    }
    DeviceGuard g(device);
    if (!g.ok) return wq4_cuda_fail(cudaGetLastError(), "cudaSetDevice");
    cudaStream_t stream = (cudaStream_t)stream_;

    wq4_tensor *t = new wq4_tensor();
    t->n = n; t->k = k; t->num_blocks = num_blocks; t->device = device;
    t->matrix_mode = (k % 32 == 0) ? 1 : 0;
    t->rows = t->matrix_mode ? n : num_blocks;
    t->kb = t->matrix_mode ? k / 32 : 1;
    t->kbp = (t->kb + 3) / 4 * 4;   // one K stage of the CTA-pair GEMM is four k-blocks (128 elements)
    t->np = (t->rows + WQ4_PANEL_ROWS - 1) / WQ4_PANEL_ROWS;
    t->np_alloc = (t->np + 1) / 2 * 2;
    int64_t entries = t->np_alloc * t->kbp * WQ4_PANEL_ROWS;
    t->nib = nullptr; t->scl = nullptr;
    uint8_t *staging = nullptr;
    cudaError_t e;
    if ((e = cudaMalloc(&t->nib, (size_t)entries * sizeof(uint4))) != cudaSuccess ||
        (e = cudaMalloc(&t->scl, (size_t)entries * sizeof(__half))) != cudaSuccess ||
        (e = cudaMalloc(&staging, expected)) != cudaSuccess ||
        (e = cudaMemcpyAsync(staging, host_blocks, expected, cudaMemcpyHostToDevice, stream)) != cudaSuccess) {
        cudaFree(t->nib); cudaFree(t->scl); cudaFree(staging); delete t;
        return wq4_cuda_fail(e, "wq4_tensor_create alloc/upload");
    }
    t->device_bytes = (size_t)entries * (sizeof(uint4) + sizeof(__half));
    int threads = 256;
    int64_t blocks = (entries + threads - 1) / threads;
    repack_kernel<<<(unsigned)blocks, threads, 0, stream>>>(
        reinterpret_cast<const uint16_t *>(staging), t->nib, t->scl, t->rows, t->kb, t->kbp, entries);
    wq4_count_launch();
    e = cudaGetLastError();
    if (e == cudaSuccess) e = cudaStreamSynchronize(stream);   // load-time only; staging is freed below
    cudaFree(staging);
    if (e != cudaSuccess) {
        cudaFree(t->nib); cudaFree(t->scl); delete t;
        return wq4_cuda_fail(e, "wq4_tensor_create repack");
    }
    *out = t;
    return WQ4_OK;
}

int32_t wq4_tensor_destroy(wq4_tensor *t)
{
    if (!t) return WQ4_OK;
    DeviceGuard g(t->device);
    cudaFree(t->nib);
    cudaFree(t->scl);
    delete t;
    return WQ4_OK;
}

int32_t wq4_tensor_shape(const wq4_tensor *t, int64_t *n, int64_t *k)
{
    if (!t || !n || !k) { wq4_set_error("wq4_tensor_shape: null argument"); return WQ4_ERR_INVALID; }
    *n = t->n; *k = t->k;
    return WQ4_OK;
}

int64_t wq4_tensor_num_blocks(const wq4_tensor *t) { return t ? t->num_blocks : -1; }
int32_t wq4_tensor_device(const wq4_tensor *t) { return t ? t->device : -1; }
size_t wq4_tensor_device_bytes(const wq4_tensor *t) { return t ? t->device_bytes : 0; }

int32_t wq4_tensor_dequantize(const wq4_tensor *t, float *dst_dev, wq4_stream stream_)
{
    if (!t || !dst_dev) { wq4_set_error("wq4_tensor_dequantize: null argument"); return WQ4_ERR_INVALID; }
    DeviceGuard g(t->device);
    if (!g.ok) return wq4_cuda_fail(cudaGetLastError(), "cudaSetDevice");
    int64_t entries = t->np * t->kbp * WQ4_PANEL_ROWS;
    int threads = 256;
    int64_t blocks = (entries * 4 + threads - 1) / threads;
    dequant_kernel<<<(unsigned)blocks, threads, 0, (cudaStream_t)stream_>>>(t->nib, t->scl, dst_dev, t->rows,
                                                                           t->kb, t->kbp, entries);
    wq4_count_launch();
    WQ4_CUDA(cudaGetLastError());
    return WQ4_OK;
}

int32_t wq4_tensor_export_gguf(const wq4_tensor *t, uint8_t *dst_dev, wq4_stream stream_)
{
    if (!t || !dst_dev) { wq4_set_error("wq4_tensor_export_gguf: null argument"); return WQ4_ERR_INVALID; }
    DeviceGuard g(t->device);
    if (!g.ok) return wq4_cuda_fail(cudaGetLastError(), "cudaSetDevice");
    int64_t entries = t->np * t->kbp * WQ4_PANEL_ROWS;
    int threads = 256;
    int64_t blocks = (entries + threads - 1) / threads;
    export_kernel<<<(unsigned)blocks, threads, 0, (cudaStream_t)stream_>>>(
        t->nib, t->scl, reinterpret_cast<uint16_t *>(dst_dev), t->rows, t->kb, t->kbp, entries);
    wq4_count_launch();
    WQ4_CUDA(cudaGetLastError());
    return WQ4_OK;
}

}  // extern "C"
