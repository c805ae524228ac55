// wq4_gemv.cu -- kernel family (2): decoder-step Q4_0 GEMV for rows (= B*M) <= 8.
//
// Replaces the M=1 use of src/gguf/shader.wgsl:51-92 (where 240 of 256 threads of every
// workgroup idle, op.rs:96-97) and folds in the bias add of src/gguf/linear.rs:36-39.
//
// HBM-bound: algorithmic bytes = N*K/32*18 (weights) + 4*K*rows (x) + 4*N (bias) + 4*N*rows (y).
// Mapping: one CTA = 64 weight rows x one K-slice; 256 threads = 64 rows x 4 k-lanes, so a warp
// reads 512 contiguous bytes of nibbles per load (16-byte ld.global.nc per lane).  The K
// dimension is split over a thread-block cluster (<= 8 CTAs); partial sums are reduced through
// distributed shared memory in a fixed order (deterministic, no atomics, no workspace), then
// bias / GELU / residual are applied by cluster rank 0.  All arithmetic is f32: (nib-8) is
// exact, the per-block scale is applied once per 32 products.
#include <cooperative_groups.h>

#include "wq4_common.cuh"

namespace cg = cooperative_groups;

namespace {

constexpr int kRowsPerCta = 64;
constexpr int kLanesK = 4;
constexpr int kThreads = kRowsPerCta * kLanesK;

struct GemvParams {
    const uint4 *nib;
    const __half *scl;
    const float *x;         // [rows][K]
    const float *bias;      // [N] or null
    const float *residual;  // [rows][N] or null
    float *y;               // [rows][N] or null
    __half *y16;            // [rows][N] or null (f16 or bf16 bits per flags)
    int rows;               // actual rows (<= MV)
    int n;                  // N
    int k;                  // K
    int kb;                 // K/32
    int kbp;
    int per;                // kb per cluster rank
    uint32_t flags;
};

__device__ __forceinline__ uint4 ldg_stream_v4(const uint4 *p)
{
    uint4 r;
    asm volatile("ld.global.nc.L1::no_allocate.v4.u32 {%0,%1,%2,%3}, [%4];"
                 : "=r"(r.x), "=r"(r.y), "=r"(r.z), "=r"(r.w)
                 : "l"(p));
    return r;
}

// 8 weights of one device word against 8 activations, for MV rows.
template <int MV>
__device__ __forceinline__ void word_dot(uint32_t w, const float *__restrict__ xs, int xstride,
                                         float (&sum)[MV])
{
    const __half2 k1032 = __floats2half2_rn(1032.0f, 1032.0f);
    const __half2 k16th = __floats2half2_rn(0.0625f, 0.0625f);
    const __half2 km72 = __floats2half2_rn(-72.0f, -72.0f);
    uint32_t t0 = (w & 0x000F000Fu) | 0x64006400u;
    uint32_t t1 = (w & 0x00F000F0u) | 0x64006400u;
    uint32_t w8 = w >> 8;
    uint32_t t2 = (w8 & 0x000F000Fu) | 0x64006400u;
    uint32_t t3 = (w8 & 0x00F000F0u) | 0x64006400u;
    // exact: (1024 + e) - 1032 = e - 8 ; (1024 + 16 e)/16 - 72 = e - 8
    float2 q01 = __half22float2(__hsub2(*reinterpret_cast<__half2 *>(&t0), k1032));
    float2 q23 = __half22float2(__hfma2(*reinterpret_cast<__half2 *>(&t1), k16th, km72));
    float2 q45 = __half22float2(__hsub2(*reinterpret_cast<__half2 *>(&t2), k1032));
    float2 q67 = __half22float2(__hfma2(*reinterpret_cast<__half2 *>(&t3), k16th, km72));
#pragma unroll
    for (int m = 0; m < MV; m++) {
        const float4 xa = *reinterpret_cast<const float4 *>(xs + m * xstride);
        const float4 xb = *reinterpret_cast<const float4 *>(xs + m * xstride + 4);
        float s = sum[m];
        s = fmaf(q01.x, xa.x, s);
        s = fmaf(q01.y, xa.y, s);
        s = fmaf(q23.x, xa.z, s);
        s = fmaf(q23.y, xa.w, s);
        s = fmaf(q45.x, xb.x, s);
        s = fmaf(q45.y, xb.y, s);
        s = fmaf(q67.x, xb.z, s);
        s = fmaf(q67.y, xb.w, s);
        sum[m] = s;
    }
}

template <int MV>
__global__ void __launch_bounds__(kThreads) q4_gemv_kernel(const GemvParams p)
{
    extern __shared__ __align__(16) float smem_f[];
    cg::cluster_group cluster = cg::this_cluster();
    const int rank = (int)cluster.block_rank();
    const int nranks = (int)cluster.num_blocks();

    const int tid = threadIdx.x;
    const int r = tid & (kRowsPerCta - 1);
    const int kl = tid >> 6;
    const int panel = blockIdx.x >> 1;
    const int r0 = (blockIdx.x & 1) * kRowsPerCta;

    const int kb0 = rank * p.per;
    const int kb1 = min(p.kb, kb0 + p.per);
    const int xlen = p.per * 32;                 // floats per row held in smem
    float *xs = smem_f;                          // [MV][xlen]
    float *red = xs + MV * xlen;                 // [kLanesK][MV][64]
    float *part = red + kLanesK * MV * kRowsPerCta;  // [MV][64]  (read by rank 0 over DSMEM)

    // stage this K-slice of x (zeros beyond the valid range / rows)
    const int nvalid = max(0, (kb1 - kb0) * 32);
    for (int idx = tid; idx < MV * xlen; idx += kThreads) {
        int m = idx / xlen, j = idx - m * xlen;
        float v = 0.0f;
        if (m < p.rows && j < nvalid) v = p.x[(size_t)m * p.k + (size_t)kb0 * 32 + j];
        xs[idx] = v;
    }
    __syncthreads();

    float acc[MV];
#pragma unroll
    for (int m = 0; m < MV; m++) acc[m] = 0.0f;

    const size_t panel_base = (size_t)panel * p.kbp * WQ4_PANEL_ROWS + r0 + r;
    constexpr int U = 4;  // entries in flight per thread
    for (int kb = kb0 + kl; kb < kb1; kb += kLanesK * U) {
        uint4 q[U];
        float d[U];
#pragma unroll
        for (int u = 0; u < U; u++) {
            int kbu = kb + u * kLanesK;
            if (kbu < kb1) {
                size_t e = panel_base + (size_t)kbu * WQ4_PANEL_ROWS;
                q[u] = ldg_stream_v4(p.nib + e);
                d[u] = __half2float(__ldg(p.scl + e));
            } else {
                q[u] = make_uint4(0x88888888u, 0x88888888u, 0x88888888u, 0x88888888u);
                d[u] = 0.0f;
            }
        }
#pragma unroll
        for (int u = 0; u < U; u++) {
            int kbu = kb + u * kLanesK;
            if (kbu >= kb1) break;
            const float *xb = xs + (kbu - kb0) * 32;
            float bs[MV];
#pragma unroll
            for (int m = 0; m < MV; m++) bs[m] = 0.0f;
            word_dot<MV>(q[u].x, xb + 0, xlen, bs);
            word_dot<MV>(q[u].y, xb + 8, xlen, bs);
            word_dot<MV>(q[u].z, xb + 16, xlen, bs);
            word_dot<MV>(q[u].w, xb + 24, xlen, bs);
#pragma unroll
            for (int m = 0; m < MV; m++) acc[m] = fmaf(d[u], bs[m], acc[m]);
        }
    }

    // reduce the 4 k-lanes inside the CTA
#pragma unroll
    for (int m = 0; m < MV; m++) red[(kl * MV + m) * kRowsPerCta + r] = acc[m];
    __syncthreads();
    for (int idx = tid; idx < MV * kRowsPerCta; idx += kThreads) {
        int m = idx / kRowsPerCta, rr = idx - m * kRowsPerCta;
        float s = 0.0f;
#pragma unroll
        for (int l = 0; l < kLanesK; l++) s += red[(l * MV + m) * kRowsPerCta + rr];
        part[idx] = s;
    }
    cluster.sync();

    if (rank == 0) {
        for (int idx = tid; idx < MV * kRowsPerCta; idx += kThreads) {
            int m = idx / kRowsPerCta, rr = idx - m * kRowsPerCta;
            int nrow = panel * WQ4_PANEL_ROWS + r0 + rr;
            float s = 0.0f;
            for (int rk = 0; rk < nranks; rk++) {   // fixed order: deterministic
                const float *remote = cluster.map_shared_rank(part, rk);
                s += remote[idx];
            }
            if (m < p.rows && nrow < p.n) {
                if (p.bias) s += p.bias[nrow];
                if (p.flags & WQ4_EPI_GELU) s = wq4_gelu_tanh(s);
                size_t o = (size_t)m * p.n + nrow;
                if (p.residual) s += p.residual[o];
                if (p.y) p.y[o] = s;
                if (p.y16) {
                    if (p.flags & WQ4_OPERANDS_BF16)
                        reinterpret_cast<__nv_bfloat16 *>(p.y16)[o] = __float2bfloat16_rn(s);
                    else
                        p.y16[o] = __float2half_rn(s);
                }
            }
        }
    }
    cluster.sync();  // keep every rank's shared memory alive until rank 0 has read it
}

template <int MV>
int32_t launch(const GemvParams &p, int np, int ks, cudaStream_t stream)
{
    size_t smem = (size_t)(MV * p.per * 32 + kLanesK * MV * kRowsPerCta + MV * kRowsPerCta) * sizeof(float);
    if (smem > 48 * 1024)
        WQ4_CUDA(cudaFuncSetAttribute(q4_gemv_kernel<MV>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem));
    cudaLaunchConfig_t cfg = {};
    cfg.gridDim = dim3((unsigned)(np * 2), (unsigned)ks, 1);
    cfg.blockDim = dim3(kThreads, 1, 1);
    cfg.dynamicSmemBytes = smem;
    cfg.stream = stream;
    cudaLaunchAttribute attr[1];
    attr[0].id = cudaLaunchAttributeClusterDimension;
    attr[0].val.clusterDim.x = 1;
    attr[0].val.clusterDim.y = (unsigned)ks;
    attr[0].val.clusterDim.z = 1;
    cfg.attrs = attr;
    cfg.numAttrs = 1;
    WQ4_CUDA(cudaLaunchKernelEx(&cfg, q4_gemv_kernel<MV>, p));
    wq4_count_launch();
    return WQ4_OK;
}

}  // namespace

int32_t wq4_launch_gemv(const LinearArgs &a)
{
    const wq4_tensor *w = a.w;
    if (a.rows < 1 || a.rows > 8) { wq4_set_error("gemv: rows=%lld outside [1,8]", (long long)a.rows); return WQ4_ERR_INVALID; }
    if (!a.x_f32) { wq4_set_error("gemv: needs f32 activations"); return WQ4_ERR_INVALID; }
    GemvParams p;
    p.nib = w->nib; p.scl = w->scl; p.x = a.x_f32; p.bias = a.bias; p.residual = a.residual;
    p.y = a.y_f32; p.y16 = reinterpret_cast<__half *>(a.y_f16);
    p.rows = (int)a.rows; p.n = (int)w->n; p.k = (int)w->k; p.kb = (int)w->kb; p.kbp = (int)w->kbp;
    p.flags = a.flags;
    // K split over the cluster: as many ranks (<= 8) as leave every k-lane at least one block
    int ks = 8;
    while (ks > 1 && (p.kb + ks - 1) / ks < kLanesK) ks >>= 1;
    p.per = (p.kb + ks - 1) / ks;
    int mv = a.rows <= 1 ? 1 : a.rows <= 2 ? 2 : a.rows <= 4 ? 4 : 8;
    // algorithmic bytes: Q4 weights + x + bias + y
    const double bytes = (double)w->n * (double)w->k / 32.0 * 18.0 + 4.0 * (double)w->k * (double)a.rows +
                         (a.bias ? 4.0 * (double)w->n : 0.0) + 4.0 * (double)w->n * (double)a.rows;
    WqProfScope prof(WQ4_PROF_GEMV, bytes, a.stream);
    switch (mv) {
    case 1: return launch<1>(p, (int)w->np, ks, a.stream);
    case 2: return launch<2>(p, (int)w->np, ks, a.stream);
    case 4: return launch<4>(p, (int)w->np, ks, a.stream);
    default: return launch<8>(p, (int)w->np, ks, a.stream);
    }
}
