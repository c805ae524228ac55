// wq4_api.cu -- C ABI dispatch (include/wq4.h): error plumbing, q4_matmul / Q4Linear / Q4FFN /
// Conv1D entry points, f32->16-bit conversion and the Conv1D input transpose.
#include <stdarg.h>
#include <stdio.h>
#include <stdlib.h>

#include <mutex>
#include <vector>

#include "wq4_common.cuh"

// ---- error plumbing ----------------------------------------------------------------------------
static thread_local char g_err[512] = "";
static thread_local int64_t g_launches = 0;

void wq4_set_error(const char *fmt, ...)
{
    va_list ap;
    va_start(ap, fmt);
    vsnprintf(g_err, sizeof(g_err), fmt, ap);
    va_end(ap);
}

int32_t wq4_cuda_fail(cudaError_t e, const char *what)
{
    wq4_set_error("CUDA error %d (%s) in %s", (int)e, cudaGetErrorString(e), what);
    return WQ4_ERR_CUDA;
}

void wq4_count_launch(int n) { g_launches += n; }

// ---- per-kernel profiler (bench.py's roofline leg): CUDA event pairs around launches ----------------
struct ProfRec { int kind; double work; cudaEvent_t e0, e1; };
static thread_local bool g_prof_on = false;
static thread_local std::vector<ProfRec> *g_prof = nullptr;
bool wq4_profile_active() { return g_prof_on; }

WqProfScope::WqProfScope(int kind, double work, cudaStream_t st) : active(false), stream(st)
{
    if (!g_prof_on) return;
    if (!g_prof) g_prof = new std::vector<ProfRec>();
    ProfRec r;
    r.kind = kind; r.work = work;
    if (cudaEventCreate(&r.e0) != cudaSuccess || cudaEventCreate(&r.e1) != cudaSuccess) return;
    cudaEventRecord(r.e0, st);
    g_prof->push_back(r);
    active = true;
}
WqProfScope::~WqProfScope()
{
    if (active) cudaEventRecord(g_prof->back().e1, stream);
}

static std::mutex g_dev_mu;
static int *g_err_flag[64] = {};
static int g_sms[64] = {};

int *wq4_error_flag_ptr(int device)
{
    if (device < 0 || device >= 64) return nullptr;
    std::lock_guard<std::mutex> lk(g_dev_mu);
    if (!g_err_flag[device]) {
        DeviceGuard g(device);
        int *p = nullptr;
        if (cudaMalloc(&p, sizeof(int)) != cudaSuccess || cudaMemset(p, 0, sizeof(int)) != cudaSuccess) {
            wq4_cuda_fail(cudaGetLastError(), "error flag alloc");
            return nullptr;
        }
        g_err_flag[device] = p;
    }
    return g_err_flag[device];
}

int wq4_sm_count(int device)
{
    if (device < 0 || device >= 64) return 148;
    std::lock_guard<std::mutex> lk(g_dev_mu);
    if (!g_sms[device]) {
        int v = 0;
        if (cudaDeviceGetAttribute(&v, cudaDevAttrMultiProcessorCount, device) != cudaSuccess || v <= 0) v = 148;
        g_sms[device] = v;
    }
    return g_sms[device];
}

// ---- small kernels -----------------------------------------------------------------------------
// Debug builds (build.sh -DWQ4_DEBUG_OVERFLOW): the f32 -> f16 conversion of the drop-in entry points counts the
// activations it clips to +-65504 (cvt.rn.satfinite), SURVEY 7-B.  Release builds carry no counter code.
__device__ unsigned long long g_f16_overflows = 0ull;

namespace {

__device__ __forceinline__ void count_overflow(float v)
{
#ifdef WQ4_DEBUG_OVERFLOW
    if (fabsf(v) > 65504.0f) atomicAdd(&g_f16_overflows, 1ull);
#else
    (void)v;
#endif
}

template <bool kBF16>
__global__ void cvt_f32_to_16_kernel(const float4 *__restrict__ src, uint2 *__restrict__ dst, int64_t n4,
                                     const float *__restrict__ src_tail, uint16_t *__restrict__ dst_tail, int tail)
{
    int64_t i = (int64_t)blockIdx.x * blockDim.x + threadIdx.x;
    const int64_t stride = (int64_t)gridDim.x * blockDim.x;
    for (; i < n4; i += stride) {
        float4 v = src[i];
        uint2 o;
        if constexpr (kBF16) {
            __nv_bfloat162 a = __floats2bfloat162_rn(v.x, v.y), b = __floats2bfloat162_rn(v.z, v.w);
            o.x = *reinterpret_cast<uint32_t *>(&a);
            o.y = *reinterpret_cast<uint32_t *>(&b);
        } else {
            // saturating: activations beyond the f16 range clamp to +-65504 instead of inf
            count_overflow(v.x); count_overflow(v.y); count_overflow(v.z); count_overflow(v.w);
            asm("{.reg .b16 l, h; cvt.rn.satfinite.f16.f32 l, %1; cvt.rn.satfinite.f16.f32 h, %2; mov.b32 %0, {l, h};}"
                : "=r"(o.x) : "f"(v.x), "f"(v.y));
            asm("{.reg .b16 l, h; cvt.rn.satfinite.f16.f32 l, %1; cvt.rn.satfinite.f16.f32 h, %2; mov.b32 %0, {l, h};}"
                : "=r"(o.y) : "f"(v.z), "f"(v.w));
        }
        dst[i] = o;
    }
    if (blockIdx.x == 0 && (int)threadIdx.x < tail) {
        float v = src_tail[threadIdx.x];
        if constexpr (kBF16) {
            __nv_bfloat16 b = __float2bfloat16_rn(v);
            dst_tail[threadIdx.x] = *reinterpret_cast<uint16_t *>(&b);
        } else {
            uint16_t h;
            asm("cvt.rn.satfinite.f16.f32 %0, %1;" : "=h"(h) : "f"(v));
            dst_tail[threadIdx.x] = h;
        }
    }
}

// x [B][C][T] f32 -> xt [B][Tp][Cp] 16-bit with `pad` zero rows on both ends of T and zero
// channels beyond C (src/model/layers.rs:83-90 builds the same padding with Tensor::cat).
template <bool kBF16>
__global__ void conv_transpose_kernel(const float *__restrict__ x, uint16_t *__restrict__ xt, int C, int T,
                                      int Cp, int Tp, int pad)
{
    __shared__ float tile[32][33];
    const int b = blockIdx.z;
    const int tp0 = blockIdx.x * 32, c0 = blockIdx.y * 32;
    for (int i = threadIdx.y; i < 32; i += blockDim.y) {
        int c = c0 + i, t = tp0 + threadIdx.x - pad;
        float v = 0.0f;
        if (c < C && t >= 0 && t < T) v = x[((size_t)b * C + c) * T + t];
        tile[i][threadIdx.x] = v;
    }
    __syncthreads();
    for (int i = threadIdx.y; i < 32; i += blockDim.y) {
        int tp = tp0 + i, c = c0 + threadIdx.x;
        if (tp < Tp && c < Cp) {
            float v = tile[threadIdx.x][i];
            uint16_t h;
            if constexpr (kBF16) {
                __nv_bfloat16 bb = __float2bfloat16_rn(v);
                h = *reinterpret_cast<uint16_t *>(&bb);
            } else {
                asm("cvt.rn.satfinite.f16.f32 %0, %1;" : "=h"(h) : "f"(v));
            }
            xt[((size_t)b * Tp + tp) * Cp + c] = h;
        }
    }
}

inline size_t align_up(size_t v, size_t a) { return (v + a - 1) / a * a; }

}  // namespace

int32_t wq4_launch_cvt_f32_to_16(const float *src, void *dst, int64_t n, bool bf16, cudaStream_t stream)
{
    if (n <= 0) return WQ4_OK;
    int64_t n4 = n / 4;
    int tail = (int)(n - n4 * 4);
    int threads = 256;
    int64_t blocks = (n4 + threads - 1) / threads;
    if (blocks < 1) blocks = 1;
    if (blocks > 148 * 16) blocks = 148 * 16;
    if (bf16)
        cvt_f32_to_16_kernel<true><<<(unsigned)blocks, threads, 0, stream>>>(
            reinterpret_cast<const float4 *>(src), reinterpret_cast<uint2 *>(dst), n4, src + n4 * 4,
            reinterpret_cast<uint16_t *>(dst) + n4 * 4, tail);
    else
        cvt_f32_to_16_kernel<false><<<(unsigned)blocks, threads, 0, stream>>>(
            reinterpret_cast<const float4 *>(src), reinterpret_cast<uint2 *>(dst), n4, src + n4 * 4,
            reinterpret_cast<uint16_t *>(dst) + n4 * 4, tail);
    wq4_count_launch();
    WQ4_CUDA(cudaGetLastError());
    return WQ4_OK;
}

// ---- dispatch -----------------------------------------------------------------------------------
bool wq4_pdl_enabled()
{
    static const bool v = [] { const char *e = getenv("WQ4_PDL"); return !(e && atoi(e) == 0); }();   // thread-safe one-time read
    return v;
}

int wq4_gemm_impl()
{
    static const int impl = [] { const char *e = getenv("WQ4_GEMM_IMPL"); return (e && atoi(e) == 1) ? 1 : 2; }();
    return impl;
}

static bool use_gemv(int64_t rows, uint32_t flags, bool has_f32_x)
{
    if (flags & WQ4_FORCE_GEMM) return false;
    if (!has_f32_x) return false;
    // 1-2 rows: f32 CUDA-core GEMV.  From 3 rows the mma.sync skinny kernel is faster (8 rows 1280 -> 5120: 15.9 us on the
    // GEMV, whose cost grows with the row count, against ~5 us; profiles/r01_microbench_*.txt); WQ4_FORCE_GEMV keeps <= 8 rows here.
    return rows <= (flags & WQ4_FORCE_SKINNY ? 0 : 2);
}

// 3..16 rows (up to 64 with WQ4_FORCE_SKINNY): the mma.sync skinny kernel (16 weight rows per CTA) instead of 256-row
// tcgen05 tiles on a handful of SMs; beyond 16 rows the tcgen05 kernel is the faster one (wq4_skinny.cu header).
// WQ4_SKINNY=0 in the environment restores the tcgen05 path for these row counts (A/B runs).
static bool use_skinny(const wq4_tensor *w, int64_t rows, uint32_t flags)
{
    static const bool on = [] { const char *e = getenv("WQ4_SKINNY"); return !(e && atoi(e) == 0); }();
    if (flags & (WQ4_FORCE_GEMM | WQ4_OPERANDS_BF16)) return false;
    if (!wq4_skinny_supported(w, rows)) return false;
    if (flags & WQ4_FORCE_SKINNY) return true;
    return on && rows > 2 && rows <= WQ4_SKINNY_DEFAULT_ROWS;
}

static int32_t check_linear(const wq4_tensor *w, const char *who)
{
    if (!w) { wq4_set_error("%s: weight tensor is null", who); return WQ4_ERR_INVALID; }
    if (!w->matrix_mode) {
        wq4_set_error("%s: K=%lld is not a multiple of 32; the reference kernel truncates K/32 "
                      "(shader.wgsl:70) and is undefined here", who, (long long)w->k);
        return WQ4_ERR_UNSUPPORTED;
    }
    return WQ4_OK;
}

extern "C" {

const char *wq4_last_error(void) { return g_err; }
int32_t wq4_version(void) { return 100; }

int32_t wq4_device_count(void)
{
    int n = 0;
    if (cudaGetDeviceCount(&n) != cudaSuccess) { cudaGetLastError(); return 0; }
    return n;
}

int64_t wq4_launch_count(void) { return g_launches; }
void wq4_launch_count_reset(void) { g_launches = 0; }

void wq4_profile_enable(int32_t on)
{
    g_prof_on = on != 0;
}

void wq4_profile_reset(void)
{
    if (!g_prof) return;
    for (auto &r : *g_prof) { cudaEventDestroy(r.e0); cudaEventDestroy(r.e1); }
    g_prof->clear();
}

int32_t wq4_profile_read(int32_t kind, int64_t *launches, double *ms, double *work)
{
    int64_t n = 0;
    double t = 0.0, w = 0.0;
    if (g_prof) {
        for (auto &r : *g_prof) {
            if (r.kind != kind) continue;
            float e = 0.0f;
            cudaError_t st = cudaEventSynchronize(r.e1);
            if (st == cudaSuccess) st = cudaEventElapsedTime(&e, r.e0, r.e1);
            if (st != cudaSuccess) return wq4_cuda_fail(st, "wq4_profile_read");
            n++; t += e; w += r.work;
        }
    }
    if (launches) *launches = n;
    if (ms) *ms = t;
    if (work) *work = w;
    return WQ4_OK;
}

int64_t wq4_debug_f16_overflow_count(int32_t device, int32_t reset)
{
#ifdef WQ4_DEBUG_OVERFLOW
    DeviceGuard g(device);
    unsigned long long v = 0ull, z = 0ull;
    if (!g.ok || cudaMemcpyFromSymbol(&v, g_f16_overflows, sizeof(v)) != cudaSuccess) { cudaGetLastError(); return -2; }
    if (reset) cudaMemcpyToSymbol(g_f16_overflows, &z, sizeof(z));
    return (int64_t)v;
#else
    (void)device; (void)reset;
    return -1;      // not a debug build
#endif
}

int32_t wq4_device_error_flag(int32_t device)
{
    int *p = wq4_error_flag_ptr(device);
    if (!p) return -1;
    DeviceGuard g(device);
    int v = 0;
    if (cudaMemcpy(&v, p, sizeof(int), cudaMemcpyDeviceToHost) != cudaSuccess) { cudaGetLastError(); return -1; }
    return v;
}

int32_t wq4_linear_plan(const wq4_tensor *w, int64_t rows, uint32_t flags, int32_t *tile_m)
{
    if (!w) return -1;
    if (use_gemv(rows, flags, true) || ((flags & WQ4_FORCE_GEMV) && rows <= 8)) {
        if (tile_m) *tile_m = 0;
        return 0;
    }
    if (use_skinny(w, rows, flags)) {
        if (tile_m) *tile_m = 16;      /* weight rows per CTA of the skinny kernel */
        return 2;
    }
    if (tile_m) *tile_m = wq4_gemm_impl() == 2 ? wq4_pick_tile_m2(rows, w->np_alloc / 2, w->device) : wq4_pick_tile_m(rows, w->np);
    return 1;
}

size_t wq4_linear_workspace_bytes(const wq4_tensor *w, int64_t b, int64_t m)
{
    if (!w || b <= 0 || m <= 0) return 0;
    return align_up((size_t)(b * m) * (size_t)w->k * 2, 256);   // 16-bit copy of x for the GEMM path
}

int32_t wq4_linear_forward_ex(const wq4_tensor *w, const float *bias_dev, const float *x_f32_dev,
                              const void *x_f16_dev, float *y_f32_dev, void *y_f16_dev,
                              const float *residual_dev, int64_t rows, uint32_t flags, void *workspace_dev,
                              size_t workspace_bytes, wq4_stream stream)
{
    int32_t st = check_linear(w, "wq4_linear_forward");
    if (st != WQ4_OK) return st;
    if ((x_f32_dev == nullptr) == (x_f16_dev == nullptr)) {
        wq4_set_error("wq4_linear_forward: exactly one of x_f32 / x_f16 must be given");
        return WQ4_ERR_INVALID;
    }
    if (!y_f32_dev && !y_f16_dev) { wq4_set_error("wq4_linear_forward: no output buffer"); return WQ4_ERR_INVALID; }
    if (rows < 0) { wq4_set_error("wq4_linear_forward: negative rows"); return WQ4_ERR_INVALID; }
    if (rows == 0) return WQ4_OK;
    DeviceGuard g(w->device);
    if (!g.ok) return wq4_cuda_fail(cudaGetLastError(), "cudaSetDevice");
    LinearArgs a;
    a.w = w; a.bias = bias_dev; a.x_f32 = x_f32_dev; a.x_f16 = x_f16_dev; a.y_f32 = y_f32_dev;
    a.y_f16 = y_f16_dev; a.residual = residual_dev; a.rows = rows; a.flags = flags;
    a.workspace = workspace_dev; a.workspace_bytes = workspace_bytes; a.stream = (cudaStream_t)stream;
    bool gemv = use_gemv(rows, flags, x_f32_dev != nullptr);
    if ((flags & WQ4_FORCE_GEMV) && rows <= 8 && x_f32_dev) gemv = true;
    if (gemv) return wq4_launch_gemv(a);
    if (use_skinny(w, rows, flags)) return wq4_launch_skinny(a, nullptr);
    if (wq4_gemm_impl() == 2) return wq4_launch_gemm2_q4(a, wq4_pick_tile_m2(rows, w->np_alloc / 2, w->device));
    return wq4_launch_gemm_q4(a, wq4_pick_tile_m(rows, w->np));
}

int32_t wq4_linear_forward(const wq4_tensor *w, const float *bias_dev, const float *x_dev, float *y_dev,
                           int64_t b, int64_t m, void *workspace_dev, size_t workspace_bytes, wq4_stream stream)
{
    if (!x_dev || !y_dev) { wq4_set_error("wq4_linear_forward: null activation/output"); return WQ4_ERR_INVALID; }
    if (b < 0 || m < 0) { wq4_set_error("wq4_linear_forward: negative batch/rows"); return WQ4_ERR_INVALID; }
    return wq4_linear_forward_ex(w, bias_dev, x_dev, nullptr, y_dev, nullptr, nullptr, b * m, 0, workspace_dev,
                                 workspace_bytes, stream);
}

int32_t wq4_matmul(const wq4_tensor *w, const float *x_dev, float *y_dev, int64_t b, int64_t m,
                   void *workspace_dev, size_t workspace_bytes, wq4_stream stream)
{
    return wq4_linear_forward(w, nullptr, x_dev, y_dev, b, m, workspace_dev, workspace_bytes, stream);
}

size_t wq4_ffn_workspace_bytes(const wq4_tensor *fc1, const wq4_tensor *fc2, int64_t b, int64_t m)
{
    if (!fc1 || !fc2 || b <= 0 || m <= 0) return 0;
    size_t rows = (size_t)(b * m);
    size_t gemm = align_up(rows * (size_t)fc1->k * 2, 256) + align_up(rows * (size_t)fc1->n * 2, 256);
    size_t gemv = align_up(rows * (size_t)fc1->n * 4, 256);
    return gemm > gemv ? gemm : gemv;
}

int32_t wq4_ffn_forward(const wq4_tensor *fc1, const float *b1_dev, const wq4_tensor *fc2, const float *b2_dev,
                        const float *x_dev, float *y_dev, int64_t b, int64_t m, void *workspace_dev,
                        size_t workspace_bytes, wq4_stream stream)
{
    int32_t st = check_linear(fc1, "wq4_ffn_forward(fc1)");
    if (st != WQ4_OK) return st;
    st = check_linear(fc2, "wq4_ffn_forward(fc2)");
    if (st != WQ4_OK) return st;
    if (!x_dev || !y_dev) { wq4_set_error("wq4_ffn_forward: null activation/output"); return WQ4_ERR_INVALID; }
    if (fc2->k != fc1->n) {
        wq4_set_error("K dimension mismatch: input has %lld, weights have %lld", (long long)fc1->n, (long long)fc2->k);
        return WQ4_ERR_SHAPE;
    }
    if (b < 0 || m < 0) { wq4_set_error("wq4_ffn_forward: negative batch/rows"); return WQ4_ERR_INVALID; }
    int64_t rows = b * m;
    if (rows == 0) return WQ4_OK;
    size_t need = wq4_ffn_workspace_bytes(fc1, fc2, b, m);
    if (!workspace_dev || workspace_bytes < need) {
        wq4_set_error("wq4_ffn_forward: workspace too small (%zu < %zu)", workspace_bytes, need);
        return WQ4_ERR_WORKSPACE;
    }
    uint8_t *ws = reinterpret_cast<uint8_t *>(workspace_dev);
    if (rows <= 8) {
        // decoder step: f32 hidden, GELU fused into fc1's reduction epilogue
        float *hidden = reinterpret_cast<float *>(ws);
        st = wq4_linear_forward_ex(fc1, b1_dev, x_dev, nullptr, hidden, nullptr, nullptr, rows, WQ4_EPI_GELU,
                                   nullptr, 0, stream);
        if (st != WQ4_OK) return st;
        return wq4_linear_forward_ex(fc2, b2_dev, hidden, nullptr, y_dev, nullptr, nullptr, rows, 0, nullptr, 0, stream);
    }
    // encoder: x -> f16, fc1 epilogue emits GELU(h) directly as the f16 operand of fc2
    size_t x16_bytes = align_up((size_t)rows * (size_t)fc1->k * 2, 256);
    void *x16 = ws;
    void *h16 = ws + x16_bytes;
    DeviceGuard g(fc1->device);
    st = wq4_launch_cvt_f32_to_16(x_dev, x16, rows * fc1->k, false, (cudaStream_t)stream);
    if (st != WQ4_OK) return st;
    st = wq4_linear_forward_ex(fc1, b1_dev, nullptr, x16, nullptr, h16, nullptr, rows, WQ4_EPI_GELU, nullptr, 0, stream);
    if (st != WQ4_OK) return st;
    return wq4_linear_forward_ex(fc2, b2_dev, nullptr, h16, y_dev, nullptr, nullptr, rows, 0, nullptr, 0, stream);
}

// ---- Conv1D -------------------------------------------------------------------------------------
int32_t wq4_conv1d_create(const float *host_w, const float *host_bias, int64_t c_out, int64_t c_in, int64_t ksize,
                          int32_t device, wq4_stream stream_, wq4_conv1d **out)
{
    if (!out) { wq4_set_error("wq4_conv1d_create: out is null"); return WQ4_ERR_INVALID; }
    *out = nullptr;
    if (!host_w || !host_bias || c_out <= 0 || c_in <= 0 || ksize <= 0) {
        wq4_set_error("wq4_conv1d_create: invalid argument");
        return WQ4_ERR_INVALID;
    }
    int ndev = 0;
    if (cudaGetDeviceCount(&ndev) != cudaSuccess || device < 0 || device >= ndev) {
        cudaGetLastError();
        wq4_set_error("wq4_conv1d_create: CUDA device %d not available (%d visible); there is no CPU fallback", device, ndev);
        return WQ4_ERR_CUDA;
    }
    DeviceGuard g(device);
    cudaStream_t stream = (cudaStream_t)stream_;
    wq4_conv1d *c = new wq4_conv1d();
    c->c_out = c_out; c->c_in = c_in; c->ksize = ksize; c->device = device;
    int64_t cp = (c_in + 7) / 8 * 8;                 // channel stride of the transposed input (16-byte rows)
    c->kdim = ksize * cp;
    c->kdim_pad = (c->kdim + 63) / 64 * 64;
    c->rows_pad = (c_out + WQ4_PANEL_ROWS - 1) / WQ4_PANEL_ROWS * WQ4_PANEL_ROWS;
    size_t nel = (size_t)c->rows_pad * c->kdim_pad;
    std::vector<__half> h16(nel, __float2half(0.0f));
    std::vector<__nv_bfloat16> b16(nel, __float2bfloat16(0.0f));
    // [c_out, c_in, ks] -> [c_out, ks, c_in] kernel-major (layers.rs:120-121), channel-padded
    for (int64_t co = 0; co < c_out; co++)
        for (int64_t ci = 0; ci < c_in; ci++)
            for (int64_t k = 0; k < ksize; k++) {
                float v = host_w[(co * c_in + ci) * ksize + k];
                size_t o = (size_t)co * c->kdim_pad + (size_t)k * cp + ci;
                h16[o] = __float2half_rn(v);
                b16[o] = __float2bfloat16_rn(v);
            }
    cudaError_t e;
    c->w16 = nullptr; c->w16b = nullptr; c->bias = nullptr;
    if ((e = cudaMalloc(&c->w16, nel * 2)) != cudaSuccess || (e = cudaMalloc(&c->w16b, nel * 2)) != cudaSuccess ||
        (e = cudaMalloc(&c->bias, (size_t)c_out * 4)) != cudaSuccess ||
        (e = cudaMemcpyAsync(c->w16, h16.data(), nel * 2, cudaMemcpyHostToDevice, stream)) != cudaSuccess ||
        (e = cudaMemcpyAsync(c->w16b, b16.data(), nel * 2, cudaMemcpyHostToDevice, stream)) != cudaSuccess ||
        (e = cudaMemcpyAsync(c->bias, host_bias, (size_t)c_out * 4, cudaMemcpyHostToDevice, stream)) != cudaSuccess ||
        (e = cudaStreamSynchronize(stream)) != cudaSuccess) {
        cudaFree(c->w16); cudaFree(c->w16b); cudaFree(c->bias); delete c;
        return wq4_cuda_fail(e, "wq4_conv1d_create");
    }
    c->device_bytes = nel * 4 + (size_t)c_out * 4;
    *out = c;
    return WQ4_OK;
}

int32_t wq4_conv1d_destroy(wq4_conv1d *c)
{
    if (!c) return WQ4_OK;
    DeviceGuard g(c->device);
    cudaFree(c->w16); cudaFree(c->w16b); cudaFree(c->bias);
    delete c;
    return WQ4_OK;
}

size_t wq4_conv1d_workspace_bytes(const wq4_conv1d *c, int64_t b, int64_t t, int64_t padding, int64_t stride)
{
    (void)stride;
    if (!c || b <= 0 || t <= 0 || padding < 0) return 0;
    int64_t cp = (c->c_in + 7) / 8 * 8;
    // one extra ksize rows of slack so the last (partial) TMA box of a batch item stays inside the buffer
    return align_up((size_t)(b * (t + 2 * padding) + c->ksize) * (size_t)cp * 2, 256);
}

int32_t wq4_conv1d_forward(const wq4_conv1d *c, const float *x_dev, float *y_dev, int64_t b, int64_t t,
                           int64_t padding, int64_t stride, uint32_t flags, void *workspace_dev,
                           size_t workspace_bytes, wq4_stream stream_)
{
    if (!c || !x_dev || !y_dev) { wq4_set_error("wq4_conv1d_forward: null argument"); return WQ4_ERR_INVALID; }
    if (b < 0 || t <= 0 || padding < 0 || stride <= 0 || t + 2 * padding < c->ksize) {
        wq4_set_error("wq4_conv1d_forward: invalid geometry b=%lld t=%lld padding=%lld stride=%lld", (long long)b,
                      (long long)t, (long long)padding, (long long)stride);
        return WQ4_ERR_INVALID;
    }
    if (b == 0) return WQ4_OK;
    size_t need = wq4_conv1d_workspace_bytes(c, b, t, padding, stride);
    if (!workspace_dev || workspace_bytes < need) {
        wq4_set_error("wq4_conv1d_forward: workspace too small (%zu < %zu)", workspace_bytes, need);
        return WQ4_ERR_WORKSPACE;
    }
    DeviceGuard g(c->device);
    if (!g.ok) return wq4_cuda_fail(cudaGetLastError(), "cudaSetDevice");
    cudaStream_t stream = (cudaStream_t)stream_;
    const bool bf16 = (flags & WQ4_OPERANDS_BF16) != 0;
    const int64_t cp = (c->c_in + 7) / 8 * 8;
    const int64_t tp = t + 2 * padding;
    const int64_t t_out = (t + 2 * padding - c->ksize) / stride + 1;   // layers.rs:80
    dim3 grid((unsigned)((tp + 31) / 32), (unsigned)((cp + 31) / 32), (unsigned)b), block(32, 8);
    if (bf16)
        conv_transpose_kernel<true><<<grid, block, 0, stream>>>(x_dev, reinterpret_cast<uint16_t *>(workspace_dev),
                                                               (int)c->c_in, (int)t, (int)cp, (int)tp, (int)padding);
    else
        conv_transpose_kernel<false><<<grid, block, 0, stream>>>(x_dev, reinterpret_cast<uint16_t *>(workspace_dev),
                                                                (int)c->c_in, (int)t, (int)cp, (int)tp, (int)padding);
    wq4_count_launch();
    WQ4_CUDA(cudaGetLastError());
    DenseGemmArgs a = {};
    a.w16 = bf16 ? (const void *)c->w16b : (const void *)c->w16;
    a.x16 = workspace_dev;
    a.bias = c->bias;
    a.y_f32 = y_dev; a.y_f16 = nullptr;
    a.n = c->c_out; a.n_pad = c->rows_pad; a.kdim_pad = c->kdim_pad;
    a.x_k = c->kdim; a.x_rows = t_out; a.x_batch = b;
    a.x_row_stride = stride * cp; a.x_batch_stride = tp * cp;
    a.ob = c->c_out * t_out; a.ot = 1; a.on = t_out; a.rb = 0; a.residual = nullptr;
    a.flags = flags; a.device = c->device; a.stream = stream;
    return wq4_launch_gemm_dense(a, wq4_pick_tile_m(t_out, a.n_pad / WQ4_PANEL_ROWS * b));
}

}  // extern "C"
