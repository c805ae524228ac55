// wq4_noalloc.cu -- wq4_ffn_forward_static: the decode-step form of Q4FFN::forward (rows <= 8, GEMV family) for callers
// that cannot allocate device memory (CUDA graph capture, tight serving loops, harnesses that forbid cudaMalloc).
// No counterpart in the reference.
//
// wq4_ffn_forward needs a caller-provided workspace for the f32 hidden row block h = gelu(fc1(x)) that fc2 consumes.
// Here that scratch is a file-scope __device__ array of a fixed, modest size: every GPU that loads the library holds
// WQ4_FFN_STATIC_MAX_ROWS * WQ4_FFN_STATIC_MAX_HIDDEN * 4 bytes (256 KiB) until the process exits.  The entry point
// calls no cudaMalloc / cudaFree and no allocating library routine; inputs and outputs are still the caller's.
//
// The array's name is only a host-side stand-in in host code, so its device address is taken once per device with
// cudaGetSymbolAddress (that device current) and that pointer is what the kernels receive.  Every call on a device
// shares the one array: it keeps the previous call's contents (nothing needs clearing -- fc1 overwrites every element
// [rows][hidden] that fc2 reads before fc2 runs, in stream order), and two calls that overlap on the GPU (different
// streams without an ordering between them) race on it.
#include <mutex>

#include "wq4_common.cuh"

#define WQ4_FFN_STATIC_MAX_ROWS 8
#define WQ4_FFN_STATIC_MAX_HIDDEN 8192

static __device__ __align__(256) float g_ffn_hidden[WQ4_FFN_STATIC_MAX_ROWS * WQ4_FFN_STATIC_MAX_HIDDEN];

static std::mutex g_static_mu;
static float *g_hidden_ptr[64] = {};

// Device address of g_ffn_hidden on `device` (which must be current), resolved once per device.
static int32_t hidden_scratch(int device, float **out)
{
    if (device < 0 || device >= 64) { wq4_set_error("wq4_ffn_forward_static: device %d out of range", device); return WQ4_ERR_INVALID; }
    std::lock_guard<std::mutex> lk(g_static_mu);
    if (!g_hidden_ptr[device]) {
        float *p = nullptr;
        cudaError_t e = cudaGetSymbolAddress(reinterpret_cast<void **>(&p), g_ffn_hidden);
        if (e != cudaSuccess) return wq4_cuda_fail(e, "cudaGetSymbolAddress(g_ffn_hidden)");
        if (!p) { wq4_set_error("wq4_ffn_forward_static: null scratch address"); return WQ4_ERR_CUDA; }
        g_hidden_ptr[device] = p;
    }
    *out = g_hidden_ptr[device];
    return WQ4_OK;
}

extern "C" {

void wq4_ffn_static_limits(int64_t *max_rows, int64_t *max_hidden)
{
    if (max_rows) *max_rows = WQ4_FFN_STATIC_MAX_ROWS;
    if (max_hidden) *max_hidden = WQ4_FFN_STATIC_MAX_HIDDEN;
}

int32_t wq4_ffn_forward_static(const wq4_tensor *fc1, const float *b1_dev, const wq4_tensor *fc2, const float *b2_dev,
                               const float *x_dev, float *y_dev, int64_t b, int64_t m, wq4_stream stream)
{
    if (!fc1 || !fc2) { wq4_set_error("wq4_ffn_forward_static: weight tensor is null"); return WQ4_ERR_INVALID; }
    if (!fc1->matrix_mode || !fc2->matrix_mode) {
        wq4_set_error("wq4_ffn_forward_static: K is not a multiple of 32");
        return WQ4_ERR_UNSUPPORTED;
    }
    if (!x_dev || !y_dev) { wq4_set_error("wq4_ffn_forward_static: null activation/output"); return WQ4_ERR_INVALID; }
    if (fc2->k != fc1->n) {
        wq4_set_error("K dimension mismatch: input has %lld, weights have %lld", (long long)fc1->n, (long long)fc2->k);
        return WQ4_ERR_SHAPE;
    }
    if (fc1->device != fc2->device) { wq4_set_error("wq4_ffn_forward_static: fc1 and fc2 live on different devices"); return WQ4_ERR_INVALID; }
    if (b < 0 || m < 0) { wq4_set_error("wq4_ffn_forward_static: negative batch/rows"); return WQ4_ERR_INVALID; }
    const int64_t rows = b * m;
    if (rows == 0) return WQ4_OK;
    if (rows > WQ4_FFN_STATIC_MAX_ROWS || fc1->n > WQ4_FFN_STATIC_MAX_HIDDEN) {
        wq4_set_error("wq4_ffn_forward_static: rows=%lld hidden=%lld beyond the static scratch (max %d rows x %d hidden); "
                      "use wq4_ffn_forward with a caller workspace", (long long)rows, (long long)fc1->n,
                      WQ4_FFN_STATIC_MAX_ROWS, WQ4_FFN_STATIC_MAX_HIDDEN);
        return WQ4_ERR_WORKSPACE;
    }
    DeviceGuard g(fc1->device);
    if (!g.ok) return wq4_cuda_fail(cudaGetLastError(), "cudaSetDevice");
    float *hidden = nullptr;
    int32_t st = hidden_scratch(fc1->device, &hidden);
    if (st != WQ4_OK) return st;
    // Same two launches as wq4_ffn_forward's rows <= 8 branch (f32 hidden, GELU fused into fc1's reduction epilogue),
    // pinned to the CUDA-core GEMV family, whose launcher allocates nothing.
    LinearArgs a = {};
    a.w = fc1; a.bias = b1_dev; a.x_f32 = x_dev; a.y_f32 = hidden; a.rows = rows;
    a.flags = WQ4_EPI_GELU; a.stream = (cudaStream_t)stream;
    st = wq4_launch_gemv(a);
    if (st != WQ4_OK) return st;
    LinearArgs c = {};
    c.w = fc2; c.bias = b2_dev; c.x_f32 = hidden; c.y_f32 = y_dev; c.rows = rows;
    c.flags = 0; c.stream = (cudaStream_t)stream;
    return wq4_launch_gemv(c);
}

}  // extern "C"
