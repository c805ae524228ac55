// wq4_skinny.cu -- Q4_0 linear for a handful of activation rows (default 9..16, up to 64 on request): decode steps of batches
// just beyond the single-kernel step and short decoder prompt passes (Q4Linear::forward at M = batch,
// src/gguf/linear.rs:34-40; shader.wgsl:51-92).
//
// The CTA-pair tcgen05 GEMM tiles N in units of 256 weight rows and walks K serially, so at N = 1280 it runs on 10 of the
// 148 SMs whatever the row count (5.5 us at K = 1280, 16.4 us at K = 5120, back to back).  Here the weight rows are dealt out
// 16 per CTA (80..320 CTAs), the warps of a CTA split K, and the product runs on mma.sync m16n8k16 with the WEIGHT rows
// as the MMA M dimension and the activation rows as N (8 per n-tile).  Measured at 16 rows: 3.4 / 4.8 / 5.0 us for
// N = 1280 / 3840 / 5120 at K = 1280 (profiles/r01_microbench_skinny.txt).  Every CTA re-reads all activation rows, so the
// advantage is gone by 32 rows (x traffic = N/16 * rows * K * 2 bytes out of L2) and the default dispatch stops at 16.  A
// variant with 128-row CTAs, bulk-copy staging and the K split over a thread-block cluster was measured slower at every
// size (cluster launch and the DSMEM exchange cost more than they saved) and dropped.
//   * A fragment straight from the panel layout, no shared memory: lane (g = lane/4, t = lane%4) loads word t of the
//     32-nibble block of rows g and g+8 -- a warp reads 2 x 128 contiguous bytes per block -- and the four half2 pairs of
//     the word (lop3 + 0x6400 trick, exact nib-8, times the f16 block scale) are the a0..a3 registers of two k16 steps.
//     That assigns block element 8t+2i+{0,1} to MMA k-slot {2t, 2t+1} (+8 for odd i), a permutation of K inside the block;
//   * the B fragment follows the same permutation for free: lane (g, t) loads the 16 contiguous bytes x[row g][32 kb + 8t ..
//     +7] and uses its four half2 as b0,b1 of the two steps.  A dot product does not care about the order of k.
//   * cross-warp reduction of the K split in shared memory in a fixed order (deterministic), then bias -> GELU -> +residual
//     -> f32 / f16 stores, and optionally the K/V rows of a fused q|k|v projection appended to the f16 self-attention
//     cache in place (attention.rs:106-113 without the realloc), which removes the separate append launch.
//   * launch chain: griddepcontrol.launch_dependents first, then the warp's weight blocks are prefetched into L2 BEFORE
//     griddepcontrol.wait (weights do not depend on the previous kernel), bias / residual are requested before the weight loop.
// Operand precision is that of the tensor-core GEMM (f16 operands, f32 accumulate).
// XMODE 0: x is f16 [rows][K];  1: x is f32 (converted on load);  2: x is f32 and LayerNorm (layers.rs:23-31: biased
// variance, eps 1e-5) is applied on load -- every CTA derives the row statistics itself (rows*K*4 bytes out of L2), which
// removes the LayerNorm launch and its dependency edge in front of the q|k|v, cross-q and fc1 projections.
#include "wq4_common.cuh"
#include "wq4_ptx.cuh"

namespace {

constexpr int kSRows = 16;          // weight rows per CTA
__host__ __device__ constexpr int sk_max_warps(int nt) { return nt <= 4 ? 16 : 8; }   // K split; 16 warps when K >= 2048

struct SkinnyParams {
    const uint4 *nib;
    const __half *scl;
    int kb, kbp, n, k, rows;
    const __half *x16;
    const float *x32;
    const float *ln_w, *ln_b;
    const float *bias, *residual;
    float *y32;
    __half *y16;
    int ldy;
    uint32_t flags;
    // optional K/V append (n in [kvD, 2kvD) -> kc, [2kvD, 3kvD) -> vc), row m of x = clip m
    __half *kc, *vc;
    long long kv_bstride;
    int kv_ld, kv_pos, kvD;
    const int *kv_slot;        // row m of x belongs to cache slot kv_slot[m] (active-set compaction); null = m
};

__device__ __forceinline__ uint32_t ldg_nc32(const uint32_t *p)
{
    uint32_t r;
    asm volatile("ld.global.nc.L1::no_allocate.u32 %0, [%1];" : "=r"(r) : "l"(p));
    return r;
}
__device__ __forceinline__ uint16_t f2h_sat(float v)
{
    uint16_t h;
    asm("cvt.rn.satfinite.f16.f32 %0, %1;" : "=h"(h) : "f"(v));
    return h;
}
__device__ __forceinline__ uint32_t pack_h2_sat(float a, float b) { return (uint32_t)f2h_sat(a) | ((uint32_t)f2h_sat(b) << 16); }

__device__ __forceinline__ void mma_16816(float (&c)[4], uint32_t a0, uint32_t a1, uint32_t a2, uint32_t a3, uint32_t b0, uint32_t b1)
{
    asm volatile("mma.sync.aligned.m16n8k16.row.col.f32.f16.f16.f32 {%0,%1,%2,%3}, {%4,%5,%6,%7}, {%8,%9}, {%0,%1,%2,%3};"
                 : "+f"(c[0]), "+f"(c[1]), "+f"(c[2]), "+f"(c[3])
                 : "r"(a0), "r"(a1), "r"(a2), "r"(a3), "r"(b0), "r"(b1));
}

// the four half2 pairs of one device word, exact (nib - 8), times the block scale
__device__ __forceinline__ void word_pairs(uint32_t w, __half2 d2, uint32_t (&o)[4])
{
    const __half2 k1032 = __floats2half2_rn(1032.0f, 1032.0f);
    const __half2 k16th = __floats2half2_rn(0.0625f, 0.0625f);
    const __half2 km72 = __floats2half2_rn(-72.0f, -72.0f);
    uint32_t t0 = (w & 0x000F000Fu) | 0x64006400u;
    uint32_t t1 = (w & 0x00F000F0u) | 0x64006400u;
    const uint32_t w8 = w >> 8;
    uint32_t t2 = (w8 & 0x000F000Fu) | 0x64006400u;
    uint32_t t3 = (w8 & 0x00F000F0u) | 0x64006400u;
    __half2 p0 = __hmul2(__hsub2(*reinterpret_cast<__half2 *>(&t0), k1032), d2);
    __half2 p1 = __hmul2(__hfma2(*reinterpret_cast<__half2 *>(&t1), k16th, km72), d2);
    __half2 p2 = __hmul2(__hsub2(*reinterpret_cast<__half2 *>(&t2), k1032), d2);
    __half2 p3 = __hmul2(__hfma2(*reinterpret_cast<__half2 *>(&t3), k16th, km72), d2);
    o[0] = *reinterpret_cast<uint32_t *>(&p0);
    o[1] = *reinterpret_cast<uint32_t *>(&p1);
    o[2] = *reinterpret_cast<uint32_t *>(&p2);
    o[3] = *reinterpret_cast<uint32_t *>(&p3);
}

template <int NT, int XMODE>
__global__ void __launch_bounds__(32 * sk_max_warps(NT)) q4_skinny_kernel(const SkinnyParams p)
{
    constexpr int RP = 8 * NT + 1;                       // padded row pitch of the reduction buffer (floats)
    __shared__ float red[sk_max_warps(NT) * kSRows * RP];
    const int kST = blockDim.x, kSW = kST >> 5;
    __shared__ float s_mean[XMODE == 2 ? 8 * NT : 1], s_rstd[XMODE == 2 ? 8 * NT : 1];
    const int tid = threadIdx.x, warp = tid >> 5, lane = tid & 31;
    const int g = lane >> 2, t = lane & 3;
    const int row0 = blockIdx.x * kSRows;                // first weight row of this CTA
    const int panel = row0 >> 7, rin = row0 & 127;
    const int K = p.k;
    ptx::grid_dep_launch();     // PDL: the next kernel of the decode chain may be scheduled
    // The weights do not depend on the previous kernel: pull this warp's blocks (two 128-byte nibble lines and the scale line
    // per block) into L2 while the predecessor is still running, so that the loads after the wait do not pay the HBM latency.
    if (lane < 3) {
        const size_t slab0 = (size_t)panel * p.kbp;
        for (int kbi = warp; kbi < p.kb; kbi += kSW) {
            const char *base = lane < 2 ? reinterpret_cast<const char *>(p.nib + (slab0 + kbi) * WQ4_PANEL_ROWS + rin) + lane * 128
                                        : reinterpret_cast<const char *>(p.scl + (slab0 + kbi) * WQ4_PANEL_ROWS + rin);
            asm volatile("prefetch.global.L2 [%0];" ::"l"(base));
        }
    }
    ptx::grid_dep_wait();       // x / residual come from the previous kernel

    if (XMODE == 2) {
        // LayerNorm statistics of every activation row (same summation order as layernorm_kernel)
        const bool in_regs = (K & 127) == 0 && K <= 2048;     // the row fits 16 float4 per lane: one pass over global memory
        for (int m = warp; m < p.rows; m += kSW) {
            const float4 *xr = reinterpret_cast<const float4 *>(p.x32 + (size_t)m * K);
            if (in_regs) {
                float4 v[16];
                const int ni = K >> 7;
                float s = 0.0f;
#pragma unroll
                for (int i = 0; i < 16; i++)
                    if (i < ni) v[i] = xr[i * 32 + lane];
#pragma unroll
                for (int i = 0; i < 16; i++)
                    if (i < ni) s += v[i].x + v[i].y + v[i].z + v[i].w;
#pragma unroll
                for (int o = 16; o > 0; o >>= 1) s += __shfl_xor_sync(0xffffffffu, s, o);
                const float mean = s / (float)K;
                float q = 0.0f;
#pragma unroll
                for (int i = 0; i < 16; i++)
                    if (i < ni) {
                        const float a = v[i].x - mean, b = v[i].y - mean, c = v[i].z - mean, d = v[i].w - mean;
                        q += a * a + b * b + c * c + d * d;
                    }
#pragma unroll
                for (int o = 16; o > 0; o >>= 1) q += __shfl_xor_sync(0xffffffffu, q, o);
                if (lane == 0) { s_mean[m] = mean; s_rstd[m] = 1.0f / sqrtf(q / (float)K + 1e-5f); }
                continue;
            }
            float s = 0.0f;
            for (int i = lane; i < K / 4; i += 32) { const float4 v = xr[i]; s += v.x + v.y + v.z + v.w; }
#pragma unroll
            for (int o = 16; o > 0; o >>= 1) s += __shfl_xor_sync(0xffffffffu, s, o);
            const float mean = s / (float)K;
            float q = 0.0f;
            for (int i = lane; i < K / 4; i += 32) {
                float4 v = xr[i];
                v.x -= mean; v.y -= mean; v.z -= mean; v.w -= mean;
                q += v.x * v.x + v.y * v.y + v.z * v.z + v.w * v.w;
            }
#pragma unroll
            for (int o = 16; o > 0; o >>= 1) q += __shfl_xor_sync(0xffffffffu, q, o);
            if (lane == 0) { s_mean[m] = mean; s_rstd[m] = 1.0f / sqrtf(q / (float)K + 1e-5f); }
        }
        __syncthreads();
    }

    // bias and residual of the first output this thread will finish (the only one up to 16 rows), requested before the
    // weight stream so that the epilogue has no dependent global load
    float pre_b = 0.0f, pre_r = 0.0f;
    if (tid < kSRows * p.rows) {
        const int nn = row0 + (tid & (kSRows - 1));
        if (nn < p.n) {
            if (p.bias) pre_b = __ldg(p.bias + nn);
            if (p.residual) pre_r = p.residual[(size_t)(tid >> 4) * p.ldy + nn];
        }
    }

    float acc[NT][4];
#pragma unroll
    for (int j = 0; j < NT; j++) acc[j][0] = acc[j][1] = acc[j][2] = acc[j][3] = 0.0f;

    float mu[NT], rs[NT];
    if (XMODE == 2) {
#pragma unroll
        for (int j = 0; j < NT; j++) {
            const int m = j * 8 + g;
            mu[j] = m < p.rows ? s_mean[m] : 0.0f;
            rs[j] = m < p.rows ? s_rstd[m] : 0.0f;
        }
    }

    const size_t slab = (size_t)panel * p.kbp;
#pragma unroll(NT <= 2 ? 4 : (NT == 4 ? 2 : 1))
    for (int kbi = warp; kbi < p.kb; kbi += kSW) {
        const uint32_t *wp = reinterpret_cast<const uint32_t *>(p.nib + (slab + kbi) * WQ4_PANEL_ROWS + rin);
        const uint32_t w_lo = ldg_nc32(wp + lane);            // rows g     (lane = 4g + t: word t of row g)
        const uint32_t w_hi = ldg_nc32(wp + 32 + lane);       // rows g + 8
        const __half *sp = p.scl + (slab + kbi) * WQ4_PANEL_ROWS + rin;
        const __half d_lo = __ldg(sp + g), d_hi = __ldg(sp + g + 8);
        uint32_t xb[NT][4];
        const int kofs = kbi * 32 + t * 8;
        if (XMODE == 0) {
#pragma unroll
            for (int j = 0; j < NT; j++) {
                const int m = j * 8 + g;
                uint4 v = make_uint4(0u, 0u, 0u, 0u);
                if (m < p.rows) v = __ldg(reinterpret_cast<const uint4 *>(p.x16 + (size_t)m * K + kofs));
                xb[j][0] = v.x; xb[j][1] = v.y; xb[j][2] = v.z; xb[j][3] = v.w;
            }
        } else {
            float4 lw0, lw1, lb0, lb1;
            if (XMODE == 2) {
                lw0 = __ldg(reinterpret_cast<const float4 *>(p.ln_w + kofs));
                lw1 = __ldg(reinterpret_cast<const float4 *>(p.ln_w + kofs) + 1);
                lb0 = __ldg(reinterpret_cast<const float4 *>(p.ln_b + kofs));
                lb1 = __ldg(reinterpret_cast<const float4 *>(p.ln_b + kofs) + 1);
            }
#pragma unroll
            for (int j = 0; j < NT; j++) {
                const int m = j * 8 + g;
                float4 a = make_float4(0.f, 0.f, 0.f, 0.f), b = a;
                if (m < p.rows) {
                    const float4 *xp = reinterpret_cast<const float4 *>(p.x32 + (size_t)m * K + kofs);
                    a = __ldg(xp);
                    b = __ldg(xp + 1);
                }
                if (XMODE == 2) {
                    a.x = (a.x - mu[j]) * rs[j] * lw0.x + lb0.x; a.y = (a.y - mu[j]) * rs[j] * lw0.y + lb0.y;
                    a.z = (a.z - mu[j]) * rs[j] * lw0.z + lb0.z; a.w = (a.w - mu[j]) * rs[j] * lw0.w + lb0.w;
                    b.x = (b.x - mu[j]) * rs[j] * lw1.x + lb1.x; b.y = (b.y - mu[j]) * rs[j] * lw1.y + lb1.y;
                    b.z = (b.z - mu[j]) * rs[j] * lw1.z + lb1.z; b.w = (b.w - mu[j]) * rs[j] * lw1.w + lb1.w;
                    if (m >= p.rows) { a = make_float4(0.f, 0.f, 0.f, 0.f); b = a; }
                }
                xb[j][0] = pack_h2_sat(a.x, a.y); xb[j][1] = pack_h2_sat(a.z, a.w);
                xb[j][2] = pack_h2_sat(b.x, b.y); xb[j][3] = pack_h2_sat(b.z, b.w);
            }
        }
        uint32_t alo[4], ahi[4];
        word_pairs(w_lo, __half2half2(d_lo), alo);
        word_pairs(w_hi, __half2half2(d_hi), ahi);
#pragma unroll
        for (int j = 0; j < NT; j++) {
            mma_16816(acc[j], alo[0], ahi[0], alo[1], ahi[1], xb[j][0], xb[j][1]);
            mma_16816(acc[j], alo[2], ahi[2], alo[3], ahi[3], xb[j][2], xb[j][3]);
        }
    }

    // C fragment: c0,c1 -> (weight row g, x rows 8j + 2t, +1); c2,c3 -> (weight row g + 8, same x rows)
    float *rw = red + warp * (kSRows * RP);
#pragma unroll
    for (int j = 0; j < NT; j++) {
        rw[g * RP + j * 8 + 2 * t] = acc[j][0];
        rw[g * RP + j * 8 + 2 * t + 1] = acc[j][1];
        rw[(g + 8) * RP + j * 8 + 2 * t] = acc[j][2];
        rw[(g + 8) * RP + j * 8 + 2 * t + 1] = acc[j][3];
    }
    __syncthreads();
    for (int o = tid; o < kSRows * p.rows; o += kST) {
        const int r = o & (kSRows - 1), m = o >> 4, nn = row0 + r;
        if (nn >= p.n) continue;
        float s = 0.0f;
        for (int w2 = 0; w2 < kSW; w2++) s += red[w2 * (kSRows * RP) + r * RP + m];      // fixed order
        const bool first = o == tid;
        if (p.bias) s += first ? pre_b : __ldg(p.bias + nn);
        if (p.flags & WQ4_EPI_GELU) s = wq4_gelu_tanh(s);
        const size_t oi = (size_t)m * p.ldy + nn;
        if (p.residual) s += first ? pre_r : p.residual[oi];
        if (p.y32) p.y32[oi] = s;
        if (p.y16) p.y16[oi] = __ushort_as_half(f2h_sat(s));
        if (p.kc && nn >= p.kvD) {
            const bool isv = nn >= 2 * p.kvD;
            const int clip = p.kv_slot ? __ldg(p.kv_slot + m) : m;
            __half *dst = (isv ? p.vc : p.kc) + (size_t)clip * p.kv_bstride + (size_t)p.kv_pos * p.kv_ld + (nn - (isv ? 2 : 1) * p.kvD);
            *dst = __ushort_as_half(f2h_sat(s));
        }
    }
}

template <int NT>
int32_t launch_nt(const SkinnyParams &p, int mode, unsigned grid, cudaStream_t st)
{
    cudaLaunchConfig_t cfg = {};
    cfg.gridDim = dim3(grid, 1, 1);
    cfg.blockDim = dim3((p.kb >= 64 ? sk_max_warps(NT) : 8) * 32, 1, 1);
    cfg.stream = st;
    cudaLaunchAttribute attr[1];
    attr[0].id = cudaLaunchAttributeProgrammaticStreamSerialization;
    attr[0].val.programmaticStreamSerializationAllowed = 1;
    cfg.attrs = attr;
    cfg.numAttrs = wq4_pdl_enabled() ? 1 : 0;
    switch (mode) {
    case 0: WQ4_CUDA(cudaLaunchKernelEx(&cfg, q4_skinny_kernel<NT, 0>, p)); break;
    case 1: WQ4_CUDA(cudaLaunchKernelEx(&cfg, q4_skinny_kernel<NT, 1>, p)); break;
    default: WQ4_CUDA(cudaLaunchKernelEx(&cfg, q4_skinny_kernel<NT, 2>, p)); break;
    }
    return WQ4_OK;
}

}  // namespace

bool wq4_skinny_supported(const wq4_tensor *w, int64_t rows) { return w && w->matrix_mode && rows >= 1 && rows <= WQ4_SKINNY_MAX_ROWS; }

int32_t wq4_launch_skinny(const LinearArgs &a, const SkinnyExtra *ex)
{
    const wq4_tensor *w = a.w;
    if (a.rows < 1 || a.rows > WQ4_SKINNY_MAX_ROWS) { wq4_set_error("skinny linear: rows %lld outside 1..%d", (long long)a.rows, WQ4_SKINNY_MAX_ROWS); return WQ4_ERR_INVALID; }
    if (a.flags & WQ4_OPERANDS_BF16) { wq4_set_error("skinny linear: f16 operands only"); return WQ4_ERR_UNSUPPORTED; }
    const bool ln = ex && ex->ln_w;
    if (ln && (!a.x_f32 || !ex->ln_b || (w->k & 3))) { wq4_set_error("skinny linear: fused LayerNorm needs f32 x, both parameters and K %% 4 == 0"); return WQ4_ERR_INVALID; }
    SkinnyParams p = {};
    p.nib = w->nib; p.scl = w->scl; p.kb = (int)w->kb; p.kbp = (int)w->kbp; p.n = (int)w->n; p.k = (int)w->k; p.rows = (int)a.rows;
    p.x16 = reinterpret_cast<const __half *>(a.x_f16); p.x32 = a.x_f32;
    p.ln_w = ln ? ex->ln_w : nullptr; p.ln_b = ln ? ex->ln_b : nullptr;
    p.bias = a.bias; p.residual = a.residual; p.y32 = a.y_f32; p.y16 = reinterpret_cast<__half *>(a.y_f16);
    p.ldy = (int)w->n; p.flags = a.flags;
    if (ex && ex->kc) {
        p.kc = reinterpret_cast<__half *>(ex->kc); p.vc = reinterpret_cast<__half *>(ex->vc);
        p.kv_bstride = ex->kv_batch_stride; p.kv_ld = (int)ex->kv_ld; p.kv_pos = (int)ex->kv_pos; p.kvD = (int)ex->kv_d;
        p.kv_slot = ex->kv_slot;
    }
    const int mode = a.x_f16 ? 0 : (ln ? 2 : 1);
    const unsigned grid = (unsigned)((w->n + kSRows - 1) / kSRows);
    WqProfScope prof(WQ4_PROF_GEMM_Q4, 2.0 * (double)a.rows * (double)w->n * (double)w->k, a.stream);
    int32_t st;
    if (a.rows <= 8) st = launch_nt<1>(p, mode, grid, a.stream);
    else if (a.rows <= 16) st = launch_nt<2>(p, mode, grid, a.stream);
    else if (a.rows <= 32) st = launch_nt<4>(p, mode, grid, a.stream);
    else st = launch_nt<8>(p, mode, grid, a.stream);
    if (st != WQ4_OK) return st;
    wq4_count_launch();
    WQ4_CUDA(cudaGetLastError());
    return WQ4_OK;
}
