// wq4_decode.cu -- kernel family (2) as ONE persistent kernel: the whole greedy decode step
// (WhisperDecoder::decode_step, src/model/decoder.rs:306-348, with forward_with_cache of
// src/model/attention.rs:93-125,208-236 and Q4FFN of src/model/layers.rs:54-58) for up to 8 clips.
//
// Why one kernel: at M = 1 a Large-V3 layer streams only 12.9 MB of Q4 weights (0.09 MB per SM); as separate
// launches the step was 388 kernels and 8.0 ms (profiles/r01_decode_launches.txt), i.e. 50 GB/s.  Here one
// cooperative grid of one CTA per SM walks embed -> 32 x [LN+QKV | self-attn | out+res | LN+q | cross-attn |
// out+res | LN+fc1+GELU | fc2+res] -> LN -> logits -> argmax with a grid barrier between phases; every
// activation vector lives in L2 between phases, weights are read exactly once per step with 16-byte loads.
//
// GEMV mapping: the rows of every weight are dealt to the CTAs in contiguous ranges (decode layout, see
// wq4_model.cuh); inside a CTA thread (kl, slot) owns row `slot` and the k-blocks kl, kl+KL, ...; a warp's
// lanes read consecutive 16-byte blocks (coalesced), the activation vector sits in shared memory (f32) and is
// read as broadcasts; partial sums are reduced through shared memory in a fixed order (deterministic).
// LayerNorm is recomputed by every CTA in the prologue of the GEMV that consumes it; bias, GELU, residual add
// and the f16 KV-cache append are fused into the GEMV epilogues.  All arithmetic is f32 ((nib-8) exact, one
// scale multiply per 32 products), accumulation order fixed.
//
// Attention: unit = (clip, head, key split); 8 lanes per key (16-byte f16 chunks), online softmax per lane
// group, merged through shuffles and shared memory; the splits are merged by the consumer GEMV's prologue.
//
// Roofline: HBM.  Algorithmic bytes per step = Q4 bytes of the six decoder linears per layer (412.9 MB at
// Large-V3) + f16 cross K/V + f16 self K/V + the f16 logits matrix; bench: scripts/bench_decode.py.
#include <cooperative_groups.h>
#include <float.h>
#include <stdio.h>
#include <stdlib.h>

#include <atomic>

#include "wq4_common.cuh"
#include "wq4_model.cuh"
#include "wq4_ptx.cuh"

namespace {

constexpr int kDT = 512;          // threads per CTA
constexpr int kMaxR = 8;          // clips per launch
constexpr int kPartStride = 68;   // o[64], m, l, pad (16-byte aligned rows) per attention partial
constexpr int kMaxSeg = 4;        // units a CTA's key range may touch
constexpr int kMaxParts = 10;     // partials per unit

__device__ __forceinline__ uint4 ldg_stream16(const uint4 *p)
{
    uint4 r;
    asm volatile("ld.global.nc.L1::no_allocate.v4.u32 {%0,%1,%2,%3}, [%4];"
                 : "=r"(r.x), "=r"(r.y), "=r"(r.z), "=r"(r.w) : "l"(p));
    return r;
}
__device__ __forceinline__ uint4 ldcg16(const void *p)
{
    uint4 r;
    asm volatile("ld.global.cg.v4.u32 {%0,%1,%2,%3}, [%4];" : "=r"(r.x), "=r"(r.y), "=r"(r.z), "=r"(r.w) : "l"(p));
    return r;
}
__device__ __forceinline__ unsigned ld_acquire_u32(const unsigned *p)
{
    unsigned v;
    asm volatile("ld.acquire.gpu.global.u32 %0, [%1];" : "=r"(v) : "l"(p) : "memory");
    return v;
}
__device__ __forceinline__ float warp_sum_d(float v)
{
#pragma unroll
    for (int o = 16; o > 0; o >>= 1) v += __shfl_xor_sync(0xffffffffu, v, o);
    return v;
}
__device__ __forceinline__ uint16_t f32_to_f16_sat_d(float v)
{
    uint16_t h;
    asm("cvt.rn.satfinite.f16.f32 %0, %1;" : "=h"(h) : "f"(v));
    return h;
}

// Grid-wide barrier over a monotonically increasing counter (zeroed by the host before each launch).
// Bounded: a co-residency or logic bug raises the device error flag instead of hanging the GPU.
// `timeline` (normally NULL) receives CTA 0's %globaltimer after barrier i at slot i: scripts/decode_timeline.py.
__device__ __forceinline__ void grid_barrier(unsigned *counter, unsigned &target, int *dev_err, long long *timeline)
{
    __syncthreads();
    if (threadIdx.x == 0) {
        target += gridDim.x;
        __threadfence();
        atomicAdd(counter, 1u);
        unsigned spins = 0;
        while (ld_acquire_u32(counter) < target) {
            if (++spins > (1u << 24)) { atomicCAS(dev_err, 0, 301); break; }
        }
        if (timeline && blockIdx.x == 0) {
            const unsigned slot = target / gridDim.x - 1;
            long long t;
            asm volatile("mov.u64 %0, %%globaltimer;" : "=l"(t));
            if (slot < (unsigned)kDecTimelineSlots) timeline[slot] = t;
        }
    }
    __syncthreads();
}

// 8 weights of one device word -> 8 f32 values (nib - 8), exact.
__device__ __forceinline__ void word_to_f32(uint32_t w, float (&f)[8])
{
    const __half2 k1032 = __floats2half2_rn(1032.0f, 1032.0f);
    const __half2 k16th = __floats2half2_rn(0.0625f, 0.0625f);
    const __half2 km72 = __floats2half2_rn(-72.0f, -72.0f);
    uint32_t t0 = (w & 0x000F000Fu) | 0x64006400u;
    uint32_t t1 = (w & 0x00F000F0u) | 0x64006400u;
    uint32_t w8 = w >> 8;
    uint32_t t2 = (w8 & 0x000F000Fu) | 0x64006400u;
    uint32_t t3 = (w8 & 0x00F000F0u) | 0x64006400u;
    float2 a = __half22float2(__hsub2(*reinterpret_cast<__half2 *>(&t0), k1032));
    float2 b = __half22float2(__hfma2(*reinterpret_cast<__half2 *>(&t1), k16th, km72));
    float2 c = __half22float2(__hsub2(*reinterpret_cast<__half2 *>(&t2), k1032));
    float2 d = __half22float2(__hfma2(*reinterpret_cast<__half2 *>(&t3), k16th, km72));
    f[0] = a.x; f[1] = a.y; f[2] = b.x; f[3] = b.y; f[4] = c.x; f[5] = c.y; f[6] = d.x; f[7] = d.y;
}

// acc[m] += d * sum_k (nib - 8) * xs[m][kbi*32 + k]   for one block.
// The word loop is deliberately NOT unrolled: every phase of the persistent kernel executes its code once per
// layer, so straight-line code is fetched cold from L2 each time; a compact loop body keeps the instruction
// stream of a phase within the instruction cache (the fully unrolled R = 8 body was 24 KB and fetch bound).
template <int R>
__device__ __forceinline__ void block_dot(uint4 q, float d, const float *xs, int K, int kbi, float (&acc)[R])
{
    float bs[R];
#pragma unroll
    for (int m = 0; m < R; m++) bs[m] = 0.0f;
    const float *xp = xs + kbi * 32;
#pragma unroll 1
    for (int c = 0; c < 4; c++) {
        float f[8];
        word_to_f32(q.x, f);
        q.x = q.y; q.y = q.z; q.z = q.w;
#pragma unroll
        for (int m = 0; m < R; m++) {
            const float4 xa = *reinterpret_cast<const float4 *>(xp + (size_t)m * K + c * 8);
            const float4 xb = *reinterpret_cast<const float4 *>(xp + (size_t)m * K + c * 8 + 4);
            float s = bs[m];
            s = fmaf(f[0], xa.x, s); s = fmaf(f[1], xa.y, s); s = fmaf(f[2], xa.z, s); s = fmaf(f[3], xa.w, s);
            s = fmaf(f[4], xb.x, s); s = fmaf(f[5], xb.y, s); s = fmaf(f[6], xb.z, s); s = fmaf(f[7], xb.w, s);
            bs[m] = s;
        }
    }
#pragma unroll
    for (int m = 0; m < R; m++) acc[m] = fmaf(d, bs[m], acc[m]);
}

// One GEMV phase: y[m][n] = epi(sum_k W[n][k] xs[m][k] + bias[n]) for this CTA's rows.
//   xs  : shared memory, f32 [R][K]
//   red : shared memory scratch, >= kDT * R floats
// Epilogue of a GEMV phase (selected at run time so one copy of the GEMV code serves all six linears).
enum { EPI_STORE = 0, EPI_RESID = 1, EPI_GELU = 2, EPI_QKV = 3 };
struct EpiCtx {
    int kind;
    float *dst;        // STORE/GELU/QKV(q): dst[m*ld + n] ; RESID: dst[m*ld + n] += v
    int ld;
    __half *sk, *sv;   // QKV: f16 cache rows (clip stride ctx*D), row kvrow
    int ctx, kvrow, D;
    const int *slot;   // QKV: cache slot of batch row m (active-set compaction) or null = m
};
__device__ __forceinline__ void epi_apply(const EpiCtx &e, int m, int n, float v)
{
    if (e.kind == EPI_STORE) e.dst[(size_t)m * e.ld + n] = v;
    else if (e.kind == EPI_RESID) e.dst[(size_t)m * e.ld + n] = v;   // residual already added by the caller
    else if (e.kind == EPI_GELU) e.dst[(size_t)m * e.ld + n] = wq4_gelu_tanh(v);
    else {
        const int D = e.D;
        const int cm = e.slot ? __ldg(e.slot + m) : m;
        if (n < D) e.dst[(size_t)m * e.ld + n] = v;
        else if (n < 2 * D) e.sk[((size_t)cm * e.ctx + e.kvrow) * D + (n - D)] = __ushort_as_half(f32_to_f16_sat_d(v));
        else e.sv[((size_t)cm * e.ctx + e.kvrow) * D + (n - 2 * D)] = __ushort_as_half(f32_to_f16_sat_d(v));
    }
}

template <int R>
__device__ __noinline__ void gemv_phase_fma(const DecWeight w, const float *xs, float *red, int nrows, const EpiCtx epi)
{
    const int tid = threadIdx.x;
    const int rpc = w.rpc, kb = w.kb, K = w.k;
    const int row0 = blockIdx.x * rpc;
    int KL = kDT / rpc;
    if (KL > kb) KL = kb;
    const int kl = tid / rpc, slot = tid - kl * rpc;
    const bool active = kl < KL && row0 < w.n;
    // the bias of the outputs this thread will finish (four lanes per output, see below), loaded before the
    // weight stream so its latency is hidden
    float bias_pre[3];       // + the residual value for EPI_RESID (x += ...), so the epilogue has no dependent load
    {
        const int g4 = tid >> 2;
#pragma unroll
        for (int i = 0; i < 3; i++) {
            const int o = i * (kDT / 4) + g4;
            const int s2 = o / R, m = o - s2 * R, n = row0 + s2;
            const bool ok = o < rpc * R && n < w.n && m < nrows;
            float v = (w.bias && ok) ? __ldg(w.bias + n) : 0.0f;
            if (epi.kind == EPI_RESID && ok) v += __ldcg(epi.dst + (size_t)m * epi.ld + n);
            bias_pre[i] = v;
        }
    }
    float acc[R];
#pragma unroll
    for (int m = 0; m < R; m++) acc[m] = 0.0f;
    if (active) {
        const size_t cbase = (size_t)blockIdx.x * kb * rpc + slot;
        for (int kb0 = kl; kb0 < kb; kb0 += 4 * KL) {
            uint4 q[4];
            float d[4];
#pragma unroll
            for (int u = 0; u < 4; u++) {
                const int kbi = kb0 + u * KL;
                q[u] = make_uint4(0, 0, 0, 0); d[u] = 0.0f;
                if (kbi < kb) {
                    q[u] = ldg_stream16(w.nib + cbase + (size_t)kbi * rpc);
                    d[u] = __half2float(__ldg(w.scl + cbase + (size_t)kbi * rpc));
                }
            }
#pragma unroll 1
            for (int u = 0; u < 4; u++) {                 // compact loop: rotate the four prefetched blocks
                const int kbi = kb0 + u * KL;
                if (kbi < kb) block_dot<R>(q[0], d[0], xs, K, kbi, acc);
                q[0] = q[1]; q[1] = q[2]; q[2] = q[3];
                d[0] = d[1]; d[1] = d[2]; d[2] = d[3];
            }
        }
#pragma unroll
        for (int m = 0; m < R; m++) red[(kl * rpc + slot) * R + m] = acc[m];
    }
    __syncthreads();
    // four lanes per output: strided partial sums, then two shuffles (fixed order: deterministic)
    const int g4 = tid >> 2, sub = tid & 3;
    int it = 0;
    for (int o0 = 0; o0 < rpc * R; o0 += kDT / 4, it++) {      // uniform trip count (shuffles inside)
        const int o = o0 + g4;
        const bool ok = o < rpc * R;
        const int s2 = ok ? o / R : 0, m = ok ? o - s2 * R : 0;
        float s = 0.0f;
        if (ok) for (int l = sub; l < KL; l += 4) s += red[(l * rpc + s2) * R + m];
        s += __shfl_xor_sync(0xffffffffu, s, 1);
        s += __shfl_xor_sync(0xffffffffu, s, 2);
        const int n = row0 + s2;
        if (ok && sub == 0 && n < w.n && m < nrows) {
            if (it < 3) s += bias_pre[it];
            else { if (w.bias) s += w.bias[n]; if (epi.kind == EPI_RESID) s += __ldcg(epi.dst + (size_t)m * epi.ld + n); }
            epi_apply(epi, m, n, s);
        }
    }
    __syncthreads();
}

// Tensor-core form of the GEMV phase for R >= 4 rows (mma.sync.m16n8k16, f16 operands, f32 accumulate): the batch rows
// are the MMA's N = 8 columns, 16 weight rows its M.  On the CUDA cores the phase costs R FMAs per weight and is
// FMA/LDS bound from R = 4 up (~15-20 us per phase at R = 8 against ~5 us at R = 1); here a weight costs one half2 op
// to dequantise and the phase is back to being bound by the weight stream and the barrier.
//   * warp w owns the k-blocks w, w+16, ... of ALL of this CTA's rows (groups of 16 rows, G <= 3);
//   * A fragment: the device word layout (wq4_common.cuh) already holds K-consecutive pairs (e0|e1, e2|e3, ...) in
//     half2-ready bit positions, so lane%4 selects its pair by a shift and a mask: rows lane/4 and lane/4+8,
//     k = 2*(lane%4)+{0,1} (+8) of the 16-wide step; (nib - 8) exact in f16, then one rounding by the block scale;
//   * B fragment: x[n = lane/4][k] from shared memory (f32 -> f16x2);
//   * partial 16 x 8 tiles of the 16 warps are summed through `red` in a fixed order (deterministic).
__device__ __forceinline__ void mma_16816(float (&c)[4], uint32_t a0, uint32_t a1, uint32_t a2, uint32_t a3, uint32_t b0, uint32_t b1)
{
    asm volatile("mma.sync.aligned.m16n8k16.row.col.f32.f16.f16.f32 {%0,%1,%2,%3}, {%4,%5,%6,%7}, {%8,%9}, {%0,%1,%2,%3};"
                 : "+f"(c[0]), "+f"(c[1]), "+f"(c[2]), "+f"(c[3])
                 : "r"(a0), "r"(a1), "r"(a2), "r"(a3), "r"(b0), "r"(b1));
}

constexpr int kMmaMaxGroups = 3;
constexpr int kMmaRedFloats = (kDT / 32) * kMmaMaxGroups * 16 * 8;

template <int R>
__device__ __noinline__ void gemv_phase_mma(const DecWeight w, const float *xs, float *red, int nrows, const EpiCtx epi)
{
    static_assert(R == 4 || R == 8, "mma GEMV: 4 or 8 batch rows");
    const int tid = threadIdx.x, warp = tid >> 5, lane = tid & 31;
    const int rpc = w.rpc, kb = w.kb, K = w.k;
    const int row0 = blockIdx.x * rpc;
    const int G = (rpc + 15) >> 4;
    const int j = lane & 3, n = lane >> 2;                 // pair index inside the word / batch row (B column)
    // bias (+ residual) of the output this thread will finish, fetched before the weight stream
    float pre[2] = {0.0f, 0.0f};
#pragma unroll
    for (int i = 0; i < 2; i++) {
        const int o = tid + i * kDT;
        const int s2 = o / R, m = o - s2 * R, nn = row0 + s2;
        if (o < rpc * R && nn < w.n && m < nrows) {
            float v = w.bias ? __ldg(w.bias + nn) : 0.0f;
            if (epi.kind == EPI_RESID) v += __ldcg(epi.dst + (size_t)m * epi.ld + nn);
            pre[i] = v;
        }
    }
    float acc[kMmaMaxGroups][4];
#pragma unroll
    for (int g = 0; g < kMmaMaxGroups; g++) acc[g][0] = acc[g][1] = acc[g][2] = acc[g][3] = 0.0f;
    if (row0 < w.n) {
        const uint32_t sh = (uint32_t)(j >> 1) * 8u;
        const uint32_t mask = (j & 1) ? 0x00F000F0u : 0x000F000Fu;
        const __half2 mul = (j & 1) ? __floats2half2_rn(0.0625f, 0.0625f) : __floats2half2_rn(1.0f, 1.0f);
        const __half2 add = (j & 1) ? __floats2half2_rn(-72.0f, -72.0f) : __floats2half2_rn(-1032.0f, -1032.0f);
        const size_t cbase = (size_t)blockIdx.x * kb * rpc;
        const float *xrow = xs + (size_t)(n < R ? n : 0) * K;
        auto frag = [&](uint32_t word, __half2 d2) -> uint32_t {
            uint32_t t = ((word >> sh) & mask) | 0x64006400u;
            __half2 v = __hmul2(__hfma2(*reinterpret_cast<__half2 *>(&t), mul, add), d2);
            return *reinterpret_cast<uint32_t *>(&v);
        };
        for (int kbi = warp; kbi < kb; kbi += kDT / 32) {
            // B fragments of the two 16-wide steps of this block
            uint32_t bfrag[2][2];
#pragma unroll
            for (int s2 = 0; s2 < 2; s2++) {
#pragma unroll
                for (int h2 = 0; h2 < 2; h2++) {
                    float2 xv = make_float2(0.0f, 0.0f);
                    if (n < R) xv = *reinterpret_cast<const float2 *>(xrow + kbi * 32 + s2 * 16 + h2 * 8 + 2 * j);
                    const __half2 hv = __floats2half2_rn(xv.x, xv.y);
                    bfrag[s2][h2] = *reinterpret_cast<const uint32_t *>(&hv);
                }
            }
#pragma unroll
            for (int g = 0; g < kMmaMaxGroups; g++) {
                if (g < G) {
                    const int r0 = g * 16 + n, r1 = r0 + 8;
                    uint4 q0 = make_uint4(0x88888888u, 0x88888888u, 0x88888888u, 0x88888888u), q1 = q0;
                    __half d0 = __ushort_as_half((unsigned short)0), d1 = d0;
                    if (r0 < rpc) { q0 = ldg_stream16(w.nib + cbase + (size_t)kbi * rpc + r0); d0 = __ldg(w.scl + cbase + (size_t)kbi * rpc + r0); }
                    if (r1 < rpc) { q1 = ldg_stream16(w.nib + cbase + (size_t)kbi * rpc + r1); d1 = __ldg(w.scl + cbase + (size_t)kbi * rpc + r1); }
                    const __half2 d02 = __half2half2(d0), d12 = __half2half2(d1);
                    mma_16816(acc[g], frag(q0.x, d02), frag(q1.x, d12), frag(q0.y, d02), frag(q1.y, d12), bfrag[0][0], bfrag[0][1]);
                    mma_16816(acc[g], frag(q0.z, d02), frag(q1.z, d12), frag(q0.w, d02), frag(q1.w, d12), bfrag[1][0], bfrag[1][1]);
                }
            }
        }
    }
    // C fragment: c0,c1 -> (row n, batch 2j, 2j+1); c2,c3 -> (row n + 8, same batch columns)
#pragma unroll
    for (int g = 0; g < kMmaMaxGroups; g++) {
        if (g < G) {
            float *t = red + ((size_t)(warp * kMmaMaxGroups + g) * 16) * 8;
            *reinterpret_cast<float2 *>(t + n * 8 + 2 * j) = make_float2(acc[g][0], acc[g][1]);
            *reinterpret_cast<float2 *>(t + (n + 8) * 8 + 2 * j) = make_float2(acc[g][2], acc[g][3]);
        }
    }
    __syncthreads();
#pragma unroll
    for (int i = 0; i < 2; i++) {
        const int o = tid + i * kDT;
        const int s2 = o / R, m = o - s2 * R, nn = row0 + s2;
        if (o < rpc * R && nn < w.n && m < nrows) {
            const int g = s2 >> 4, rr = s2 & 15;
            float s = 0.0f;
#pragma unroll
            for (int w2 = 0; w2 < kDT / 32; w2++) s += red[((size_t)(w2 * kMmaMaxGroups + g) * 16 + rr) * 8 + m];
            epi_apply(epi, m, nn, s + pre[i]);
        }
    }
    __syncthreads();
}

template <int R>
__device__ __forceinline__ void gemv_phase(const DecWeight w, const float *xs, float *red, int nrows, const EpiCtx epi)
{
    if constexpr (R >= 4) {
        if (w.rpc <= 16 * kMmaMaxGroups && w.rpc * R <= 2 * kDT) { gemv_phase_mma<R>(w, xs, red, nrows, epi); return; }
    }
    gemv_phase_fma<R>(w, xs, red, nrows, epi);
}

// LayerNorm::forward (src/model/layers.rs:23-31) of R rows of a global [R][D] f32 matrix into shared memory.
// kDT/R threads per row issue all their loads at once (one L2 round trip); two block-wide reductions through
// `red`.  The rows were written by other CTAs in an earlier phase: read through L2 (ld.cg).  D <= 1280.
template <int R>
__device__ __noinline__ void ln_rows(const float *x, const float *w, const float *b, float *xs, float *red, int D)
{
    constexpr int TPR = kDT / R, WPR = TPR / 32, NV = (320 + TPR - 1) / TPR;
    const int tid = threadIdx.x, m = tid / TPR, t = tid - m * TPR, warp = tid >> 5, lane = tid & 31;
    const int cnt = D >> 2;
    const float4 *xr = reinterpret_cast<const float4 *>(x + (size_t)m * D);
    float4 v[NV], ww[NV], bb[NV];
    float s = 0.0f;
#pragma unroll
    for (int j = 0; j < NV; j++) {                       // all loads (x through L2, gamma/beta) in flight together
        const int i = t + j * TPR;
        const bool ok = i < cnt;
        v[j] = ok ? __ldcg(xr + i) : make_float4(0.f, 0.f, 0.f, 0.f);
        ww[j] = ok ? __ldg(reinterpret_cast<const float4 *>(w) + i) : make_float4(0.f, 0.f, 0.f, 0.f);
        bb[j] = ok ? __ldg(reinterpret_cast<const float4 *>(b) + i) : make_float4(0.f, 0.f, 0.f, 0.f);
    }
#pragma unroll
    for (int j = 0; j < NV; j++) s += (v[j].x + v[j].y) + (v[j].z + v[j].w);
    s = warp_sum_d(s);
    if (lane == 0) red[warp] = s;
    __syncthreads();
    float tot = 0.0f;
#pragma unroll
    for (int i = 0; i < WPR; i++) tot += red[m * WPR + i];
    const float mean = tot / (float)D;
    float qv = 0.0f;
#pragma unroll
    for (int j = 0; j < NV; j++) {
        const int i = t + j * TPR;
        if (i < cnt) {
            v[j].x -= mean; v[j].y -= mean; v[j].z -= mean; v[j].w -= mean;
            qv += (v[j].x * v[j].x + v[j].y * v[j].y) + (v[j].z * v[j].z + v[j].w * v[j].w);
        }
    }
    qv = warp_sum_d(qv);
    if (lane == 0) red[32 + warp] = qv;
    __syncthreads();
    float vt = 0.0f;
#pragma unroll
    for (int i = 0; i < WPR; i++) vt += red[32 + m * WPR + i];
    const float inv = 1.0f / sqrtf(vt / (float)D + 1e-5f);
    float4 *row = reinterpret_cast<float4 *>(xs + (size_t)m * D);
#pragma unroll
    for (int j = 0; j < NV; j++) {
        const int i = t + j * TPR;
        if (i < cnt)
            row[i] = make_float4(v[j].x * inv * ww[j].x + bb[j].x, v[j].y * inv * ww[j].y + bb[j].y,
                                 v[j].z * inv * ww[j].z + bb[j].z, v[j].w * inv * ww[j].w + bb[j].w);
    }
    __syncthreads();
}

// L2 prefetch of this CTA's chunk of a weight that a LATER phase will stream (weights do not depend on the
// activations, so the HBM latency hides behind the current phase and its grid barrier).
__device__ __noinline__ void prefetch_weight(const DecWeight w, const float *ln_w, const float *ln_b, int D)
{
    if (ln_w) {                                          // LayerNorm parameters of the consuming phase (every CTA needs all)
        const int lines = (D * 4 + 127) >> 7;
        for (int i = threadIdx.x; i < 2 * lines; i += kDT) {
            const char *p = i < lines ? reinterpret_cast<const char *>(ln_w) + ((size_t)i << 7)
                                      : reinterpret_cast<const char *>(ln_b) + ((size_t)(i - lines) << 7);
            asm volatile("prefetch.global.L2 [%0];" ::"l"(p));
        }
    }
    if (blockIdx.x * w.rpc >= w.n) return;
    if (w.bias && threadIdx.x < 2) {
        const char *p = reinterpret_cast<const char *>(w.bias + (size_t)blockIdx.x * w.rpc) + ((size_t)threadIdx.x << 7);
        asm volatile("prefetch.global.L2 [%0];" ::"l"(p));
    }
    const size_t blocks = (size_t)w.kb * w.rpc;
    const char *pn = reinterpret_cast<const char *>(w.nib + (size_t)blockIdx.x * blocks);
    const char *ps = reinterpret_cast<const char *>(w.scl + (size_t)blockIdx.x * blocks);
    const int ln = (int)((blocks * 16 + 127) >> 7), ls = (int)((blocks * 2 + 127) >> 7);
    for (int i = threadIdx.x; i < ln + ls; i += kDT) {
        const char *p = i < ln ? pn + ((size_t)i << 7) : ps + ((size_t)(i - ln) << 7);
        asm volatile("prefetch.global.L2 [%0];" ::"l"(p));
    }
}

// ---- attention over an f16 KV cache, query length 1 (attention.rs:93-125 self, :208-236 cross) ----------
struct Soft { float m, l, o[8]; };

__device__ __forceinline__ void soft_merge(Soft &a, float m2, float l2, const float (&o2)[8])
{
    const float mn = fmaxf(a.m, m2);
    const float sa = (a.m == -INFINITY) ? 0.0f : __expf(a.m - mn);
    const float sb = (m2 == -INFINITY) ? 0.0f : __expf(m2 - mn);
    a.l = a.l * sa + l2 * sb;
#pragma unroll
    for (int j = 0; j < 8; j++) a.o[j] = a.o[j] * sa + o2[j] * sb;
    a.m = mn;
}

// The keys of all (clip, head) units form one flattened sequence of units*kv keys that is cut into equal ranges
// of `per` keys, one per CTA (perfect balance for any batch size); a CTA's range touches at most kMaxSeg
// consecutive units and it writes one partial per touched unit:
//   part[(cta*kMaxSeg + seg)*68] = unnormalised o[64], running max, sum.
// q: global f32 [B][D] (written in the previous phase); K/V: f16 rows of D (head h at column h*64), row stride
// ldkv, clip stride bstride.
// Few units (units*2 <= grid, e.g. one clip): every unit is cut into ns aligned splits, one per CTA (one segment,
// one block-wide merge per CTA); kvp = per*ns is the padded unit length.  Many units: kvp = kv, ranges of `per`
// keys straddle unit boundaries (<= kMaxSeg segments per CTA) for perfect balance.
struct AttnPlan { int per, kvp; };
__device__ __forceinline__ AttnPlan attn_plan(int units, int kv, int grid)
{
    AttnPlan p;
    if (units * 2 <= grid) {
        int ns = min(grid / units, 8);
        int per = (kv + ns - 1) / ns;
        if (per < 16) { per = min(kv, 16); ns = (kv + per - 1) / per; }
        p.per = per; p.kvp = per * ns;
    } else {
        const int mn = max((kv + 7) / 8, min(kv, 16));
        const int per = (int)(((long long)units * kv + grid - 1) / grid);
        p.per = per < mn ? mn : per; p.kvp = kv;
    }
    return p;
}

template <bool kFreshKV>
__device__ __noinline__ void attn_phase(const float *q, int D, const __half *kc, const __half *vc, long long ldkv,
                                        long long bstride, long long hstride, int kv_len, int B, int H, AttnPlan pl, float *part, float *sm,
                                        const int *slot)
{
    const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
    const int g = lane >> 3, dch = lane & 7;
    const int per = pl.per, kvp = pl.kvp;
    const int total = B * H * kvp;
    int start = (int)blockIdx.x * per;
    const int end = min(total, start + per);
    for (int seg = 0; start < end && seg < kMaxSeg; seg++) {
        const int u = start / kvp;
        const int k0 = start - u * kvp;
        const int k1p = min(kvp, k0 + (end - start));
        start += k1p - k0;
        const int k1 = min(k1p, kv_len);                   // keys beyond kv_len are padding of the aligned plan
        const int h = u % H, b = u / H;
        float qv[8];
        {
            const float *qp = q + (size_t)b * D + h * 64 + dch * 8;
            const float4 a = __ldcg(reinterpret_cast<const float4 *>(qp));
            const float4 c = __ldcg(reinterpret_cast<const float4 *>(qp) + 1);
            // (q . k) / sqrt(64) (attention.rs:266-267): scaling q by 1/8 first is exact (power of two)
            qv[0] = a.x * 0.125f; qv[1] = a.y * 0.125f; qv[2] = a.z * 0.125f; qv[3] = a.w * 0.125f;
            qv[4] = c.x * 0.125f; qv[5] = c.y * 0.125f; qv[6] = c.z * 0.125f; qv[7] = c.w * 0.125f;
        }
        const int cb = slot ? __ldg(slot + b) : b;
        const __half *kb_ = kc + (size_t)cb * bstride + (size_t)h * hstride + dch * 8;
        const __half *vb_ = vc + (size_t)cb * bstride + (size_t)h * hstride + dch * 8;
        Soft st;
        st.m = -INFINITY; st.l = 0.0f;
#pragma unroll
        for (int j = 0; j < 8; j++) st.o[j] = 0.0f;
        for (int base0 = k0 + warp * 4; base0 < k1; base0 += 256) {   // warp-uniform trip count (shuffles inside)
            const int key0 = base0 + g;
            constexpr int NT = 4;                                      // four keys per lane group in flight
            uint4 kq[NT], vq[NT];
            bool ok[NT];
#pragma unroll
            for (int t = 0; t < NT; t++) {
                const int key = key0 + t * 64;
                ok[t] = key < k1;
                if (ok[t]) {
                    const __half *kp = kb_ + (size_t)key * ldkv, *vp = vb_ + (size_t)key * ldkv;
                    if (kFreshKV) { kq[t] = ldcg16(kp); vq[t] = ldcg16(vp); }
                    else { kq[t] = ldg_stream16(reinterpret_cast<const uint4 *>(kp)); vq[t] = ldg_stream16(reinterpret_cast<const uint4 *>(vp)); }
                }
            }
#pragma unroll
            for (int t = 0; t < NT; t++) {
                float sc = 0.0f;
                if (ok[t]) {
                    const __half2 *hp = reinterpret_cast<const __half2 *>(&kq[t]);
#pragma unroll
                    for (int e = 0; e < 4; e++) {
                        const float2 f = __half22float2(hp[e]);
                        sc = fmaf(qv[2 * e], f.x, sc);
                        sc = fmaf(qv[2 * e + 1], f.y, sc);
                    }
                }
                sc += __shfl_xor_sync(0xffffffffu, sc, 1);
                sc += __shfl_xor_sync(0xffffffffu, sc, 2);
                sc += __shfl_xor_sync(0xffffffffu, sc, 4);
                if (ok[t]) {
                    const float mn = fmaxf(st.m, sc);
                    const float alpha = (st.m == -INFINITY) ? 0.0f : __expf(st.m - mn);
                    const float pw = __expf(sc - mn);
                    st.l = st.l * alpha + pw;
                    const __half2 *vp2 = reinterpret_cast<const __half2 *>(&vq[t]);
#pragma unroll
                    for (int e = 0; e < 4; e++) {
                        const float2 f = __half22float2(vp2[e]);
                        st.o[2 * e] = st.o[2 * e] * alpha + pw * f.x;
                        st.o[2 * e + 1] = st.o[2 * e + 1] * alpha + pw * f.y;
                    }
                    st.m = mn;
                }
            }
        }
        // merge the four key groups of the warp (lanes with equal dch)
#pragma unroll
        for (int off = 8; off <= 16; off <<= 1) {
            const float m2 = __shfl_xor_sync(0xffffffffu, st.m, off);
            const float l2 = __shfl_xor_sync(0xffffffffu, st.l, off);
            float o2[8];
#pragma unroll
            for (int j = 0; j < 8; j++) o2[j] = __shfl_xor_sync(0xffffffffu, st.o[j], off);
            soft_merge(st, m2, l2, o2);
        }
        if (g == 0) {
            float *dst = sm + warp * kPartStride;
            if (dch == 0) { dst[64] = st.m; dst[65] = st.l; }
#pragma unroll
            for (int j = 0; j < 8; j++) dst[dch * 8 + j] = st.o[j];
        }
        __syncthreads();
        if (warp == 0) {
            float M = -INFINITY;
            for (int w2 = 0; w2 < kDT / 32; w2++) M = fmaxf(M, sm[w2 * kPartStride + 64]);
            float Lsum = 0.0f, o0 = 0.0f, o1 = 0.0f;
            for (int w2 = 0; w2 < kDT / 32; w2++) {      // fixed order
                const float mw = sm[w2 * kPartStride + 64];
                const float sc = (mw == -INFINITY) ? 0.0f : __expf(mw - M);
                Lsum += sc * sm[w2 * kPartStride + 65];
                o0 += sc * sm[w2 * kPartStride + lane];
                o1 += sc * sm[w2 * kPartStride + 32 + lane];
            }
            float *pd = part + (size_t)(blockIdx.x * kMaxSeg + seg) * kPartStride;
            if (lane == 0) { pd[64] = M; pd[65] = Lsum; }
            pd[lane] = o0;
            pd[32 + lane] = o1;
        }
        __syncthreads();
    }
}

// ---- stand-alone form for the per-op decoder path (batches beyond the grouped kernel's reach) -------------------------
// One query row per clip, f16 KV cache, non-causal: grid (clip*head, split).  Same inner loop as attn_phase (8 lanes per
// 128-byte key row, 4 keys per lane group in flight, online softmax); with more than one split per unit the last CTA to
// finish a unit merges its partials ("fix-up" by an arrival counter that it resets), so there is no second launch.
// sm_100 mixed-precision FMA (PTX fma.rn.f32.f16, SASS FHFMA): acc += a.h * b.h with f16 operands taken as register halves.
__device__ __forceinline__ void fhfma_pair(float &acc_lo, float &acc_hi, uint32_t a2, uint32_t b2)      // (a.lo*b.lo, a.hi*b.hi)
{
    asm("{\n\t.reg .b16 al, ah, bl, bh;\n\tmov.b32 {al, ah}, %2;\n\tmov.b32 {bl, bh}, %3;\n\t"
        "fma.rn.f32.f16 %0, al, bl, %0;\n\tfma.rn.f32.f16 %1, ah, bh, %1;\n\t}"
        : "+f"(acc_lo), "+f"(acc_hi) : "r"(a2), "r"(b2));
}
__device__ __forceinline__ void fhfma_scalar(float &acc_lo, float &acc_hi, uint32_t a2, unsigned short s)   // (a.lo*s, a.hi*s)
{
    asm("{\n\t.reg .b16 al, ah;\n\tmov.b32 {al, ah}, %2;\n\t"
        "fma.rn.f32.f16 %0, al, %3, %0;\n\tfma.rn.f32.f16 %1, ah, %3, %1;\n\t}"
        : "+f"(acc_lo), "+f"(acc_hi) : "r"(a2), "h"(s));
}

constexpr int kDaStages = 3;
constexpr size_t da_smem(int warps) { return (size_t)warps * kDaStages * 2 * 32 * 128; }   // warp-private rings: 24 KB per warp
// kDaW warps per CTA: 8 for long key ranges (cross-attention: 192 KB of ring, one CTA per SM, 128 KB in flight), 2 for
// short ones (self-attention over <= 448 keys: 48 KB, four CTAs per SM, so that a 640-unit launch is about one wave).
template <int kDaW>
__global__ void __launch_bounds__(kDaW * 32, kDaW == 8 ? 1 : 4)
dec_attention_split_kernel(const float *__restrict__ q, int ldq, const __half *__restrict__ kc, const __half *__restrict__ vc,
                           long long ldkv, long long bstride, long long hstride, float *__restrict__ out, __half *__restrict__ out16, int ldo, int kv_len,
                           int H, int nsplit, int append_off,
                           float *__restrict__ part, int *__restrict__ cnt, const int *__restrict__ slot)
{
    __shared__ float sm[(kDaW) * kPartStride];
    __shared__ int s_last;
    const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
    const int u = blockIdx.x, sp = blockIdx.y, h = u % H, b = u / H;
    // Active-set compaction (whisper.rs:105-108 per clip): query / output rows are dense (row b), the K/V caches stay where
    // the clip's slot put them.  The slot table is written by the bookkeeping kernel of an EARLIER step, long before this
    // kernel's predecessor, so it may be read before the programmatic-launch wait.
    const int cb = slot ? __ldg(slot + b) : b;
    const int per = (kv_len + nsplit - 1) / nsplit;
    const int k0 = sp * per, k1 = min(kv_len, k0 + per);
    ptx::grid_dep_launch();     // PDL: the following projection may be scheduled and prefetch its weights
    // append_off < 0: the K/V rows are static (cross-attention over the encoder K/V), so the first ring stages are requested
    // BEFORE waiting for the previous kernel and only q is read after the wait
    const bool static_kv = append_off < 0;
    if (!static_kv) ptx::grid_dep_wait();       // q and the new K/V row come from the previous kernel

    // Optional cache append (self-attention of a decode step, attention.rs:106-113 without the realloc): q points into the
    // fused q|k|v projection output and the new key / value of this (clip, head) sit append_off / 2*append_off floats
    // further.  The split that owns the newest key stores them (f16, saturating, as kv_append_kernel does) before it
    // starts streaming, which removes the separate append launch.
    if (append_off > 0 && k0 < k1 && k1 == kv_len) {
        if (threadIdx.x < 16) {
            const int isv = threadIdx.x >> 3, c = threadIdx.x & 7;
            const float *src = q + (size_t)b * ldq + (size_t)(1 + isv) * append_off + h * 64 + c * 8;
            const float4 a = __ldcg(reinterpret_cast<const float4 *>(src)), d = __ldcg(reinterpret_cast<const float4 *>(src) + 1);
            uint4 pk;
            pk.x = (uint32_t)f32_to_f16_sat_d(a.x) | ((uint32_t)f32_to_f16_sat_d(a.y) << 16);
            pk.y = (uint32_t)f32_to_f16_sat_d(a.z) | ((uint32_t)f32_to_f16_sat_d(a.w) << 16);
            pk.z = (uint32_t)f32_to_f16_sat_d(d.x) | ((uint32_t)f32_to_f16_sat_d(d.y) << 16);
            pk.w = (uint32_t)f32_to_f16_sat_d(d.z) | ((uint32_t)f32_to_f16_sat_d(d.w) << 16);
            __half *dst = const_cast<__half *>(isv ? vc : kc) + (size_t)cb * bstride + (size_t)h * hstride + (size_t)(kv_len - 1) * ldkv + c * 8;
            *reinterpret_cast<uint4 *>(dst) = pk;
        }
        __syncthreads();
    }

    // ONE KEY PER LANE.  The first form of this kernel spread a key over 8 lanes (8 dims each): three shuffles, two
    // exponentials and the loop overhead per key quad made it issue bound -- 24 M warp instructions, 52 % issue-active with
    // 16 resident warps, 54 us for the 246 MB cross-attention read of a 32-clip step (4.5 TB/s; ncu profiles/r01).  Here a lane
    // owns a whole 64-dim key row: no shuffles in the loop, one exponential per key, 64 independent accumulators of ILP, and
    // a single butterfly reduce-scatter over the warp at the end.  Rows are staged by cp.async into a warp-private ring
    // (kDaStages x 32 keys x (128 B K + 128 B V)), copied as coalesced 16-byte chunks and swizzled (chunk ^ key&7) so that
    // the row-per-lane reads are bank-conflict free.
    extern __shared__ __align__(128) uint8_t da_ring[];    // [warp][stage][K|V][32 keys][128 B]
    constexpr uint32_t kHalf = 32 * 128, kStage = 2 * kHalf, kWarpRing = kDaStages * kStage;
    const uint32_t ring = ptx::smem_u32(da_ring) + (uint32_t)warp * kWarpRing;
    const uint8_t *ring_p = da_ring + (size_t)warp * kWarpRing;
    const __half *kb_ = kc + (size_t)cb * bstride + (size_t)h * hstride;
    const __half *vb_ = vc + (size_t)cb * bstride + (size_t)h * hstride;
    const int nblocks = (k1 - k0 + 31) >> 5;                                   // 32-key blocks of this CTA's range
    const int n_it = nblocks > warp ? (nblocks - warp + kDaW - 1) / (kDaW) : 0;   // blocks warp, warp+8, ...
    // chunk id = i*32 + lane of a 32-key block: key kl = 4i + (lane >> 3), 16-byte chunk c = lane & 7 (8 lanes = one 128-byte
    // row: coalesced); swizzled slot (c ^ (kl & 7)) with kl & 7 = 4*(i & 1) + (lane >> 3).  All per-lane parts are hoisted.
    const int kl0 = lane >> 3, cch = lane & 7;
    const char *gk0 = reinterpret_cast<const char *>(kb_ + (size_t)(k0 + warp * 32 + kl0) * ldkv + cch * 8);
    const char *gv0 = reinterpret_cast<const char *>(vb_ + (size_t)(k0 + warp * 32 + kl0) * ldkv + cch * 8);
    const size_t step_i = (size_t)ldkv * 8, step_it = (size_t)ldkv * (kDaW) * 64;      // bytes: 4 keys / 256 keys
    const uint32_t sw_even = (uint32_t)(kl0 * 128 + ((cch ^ kl0) << 4)), sw_odd = (uint32_t)(kl0 * 128 + ((cch ^ (4 | kl0)) << 4));
    auto issue = [&](int it) {
        if (it < n_it) {
            const int rem = k1 - (k0 + (warp + it * (kDaW)) * 32) - kl0;     // key 4i + kl0 of the block is valid iff 4i < rem
            const uint32_t dst = ring + (uint32_t)(it % kDaStages) * kStage;
            const char *gk = gk0 + (size_t)it * step_it, *gv = gv0 + (size_t)it * step_it;
#pragma unroll
            for (int i = 0; i < 8; i++) {
                if (4 * i < rem) {
                    const uint32_t d = dst + (uint32_t)(i * 512) + ((i & 1) ? sw_odd : sw_even);
                    asm volatile("cp.async.cg.shared.global [%0], [%1], 16;" ::"r"(d), "l"(gk + i * step_i) : "memory");
                    asm volatile("cp.async.cg.shared.global [%0], [%1], 16;" ::"r"(d + kHalf), "l"(gv + i * step_i) : "memory");
                }
            }
        }
        asm volatile("cp.async.commit_group;" ::: "memory");
    };
#pragma unroll
    for (int s2 = 0; s2 < kDaStages - 1; s2++) issue(s2);
    if (static_kv) ptx::grid_dep_wait();        // q comes from the previous kernel

    // q as packed f16 pairs: the products run on the mixed-precision FMA of sm_100 (fma.rn.f32.f16 -> FHFMA: f16 x f16 + f32,
    // operand halves addressed in place), which removes the 128 half->float conversions per key that made up a fifth of
    // the instruction stream.  q is rounded to f16 like every tensor-core operand of this library; the
    // product and the accumulation are exact / f32.
    uint32_t qh[32];
    {
        const float4 *qp = reinterpret_cast<const float4 *>(q + (size_t)b * ldq + h * 64);
#pragma unroll
        for (int i = 0; i < 16; i++) {
            const float4 a = __ldg(qp + i);
            // (q . k) / sqrt(64) (attention.rs:266-267): scaling q by 1/8 first is exact (power of two)
            const __half2 lo = __floats2half2_rn(a.x * 0.125f, a.y * 0.125f), hi = __floats2half2_rn(a.z * 0.125f, a.w * 0.125f);
            qh[2 * i] = *reinterpret_cast<const uint32_t *>(&lo);
            qh[2 * i + 1] = *reinterpret_cast<const uint32_t *>(&hi);
        }
    }
    float m_run = -INFINITY, l_run = 0.0f, o[64];
#pragma unroll
    for (int j = 0; j < 64; j++) o[j] = 0.0f;

    for (int it = 0; it < n_it; it++) {
        issue(it + kDaStages - 1);
        asm volatile("cp.async.wait_group %0;" ::"n"(kDaStages - 1) : "memory");
        __syncwarp();                                                          // the other lanes' chunks of this stage
        const int key = k0 + (warp + it * (kDaW)) * 32 + lane;
        if (key < k1) {
            const uint8_t *kr = ring_p + (size_t)(it % kDaStages) * kStage + lane * 128;
            float sc0 = 0.0f, sc1 = 0.0f, sc2 = 0.0f, sc3 = 0.0f;
#pragma unroll
            for (int c = 0; c < 8; c++) {
                const uint4 kq = *reinterpret_cast<const uint4 *>(kr + ((c ^ (lane & 7)) << 4));
                fhfma_pair(sc0, sc1, kq.x, qh[4 * c]);
                fhfma_pair(sc2, sc3, kq.y, qh[4 * c + 1]);
                fhfma_pair(sc0, sc1, kq.z, qh[4 * c + 2]);
                fhfma_pair(sc2, sc3, kq.w, qh[4 * c + 3]);
            }
            const float sc = (sc0 + sc1) + (sc2 + sc3);
            if (sc > m_run) {                                                  // new running maximum: rescale (rare after the first keys)
                const float alpha = (m_run == -INFINITY) ? 0.0f : __expf(m_run - sc);
                l_run *= alpha;
#pragma unroll
                for (int j = 0; j < 64; j++) o[j] *= alpha;
                m_run = sc;
            }
            // the weight as f16 (<= 1, relative rounding 2^-11) for the mixed-precision FMA; the running sum uses the same
            // rounded value so that numerator and denominator stay consistent
            const __half pw16 = __float2half_rn(__expf(sc - m_run));
            l_run += __half2float(pw16);
            const unsigned short pws = __half_as_ushort(pw16);
            const uint8_t *vr = kr + kHalf;
#pragma unroll
            for (int c = 0; c < 8; c++) {
                const uint4 vq = *reinterpret_cast<const uint4 *>(vr + ((c ^ (lane & 7)) << 4));
                fhfma_scalar(o[8 * c], o[8 * c + 1], vq.x, pws);
                fhfma_scalar(o[8 * c + 2], o[8 * c + 3], vq.y, pws);
                fhfma_scalar(o[8 * c + 4], o[8 * c + 5], vq.z, pws);
                fhfma_scalar(o[8 * c + 6], o[8 * c + 7], vq.w, pws);
            }
        }
        __syncwarp();                                                          // stage consumed by every lane before it is refilled
    }

    // merge the 32 lanes of the warp: common maximum, then a butterfly reduce-scatter of the 64 outputs (62 shuffles);
    // lane l ends up with dims 2l and 2l+1
    float m_w = m_run;
#pragma unroll
    for (int off = 16; off > 0; off >>= 1) m_w = fmaxf(m_w, __shfl_xor_sync(0xffffffffu, m_w, off));
    const float scl = (m_run == -INFINITY) ? 0.0f : __expf(m_run - m_w);
    float l_w = l_run * scl;
#pragma unroll
    for (int off = 16; off > 0; off >>= 1) l_w += __shfl_xor_sync(0xffffffffu, l_w, off);
#pragma unroll
    for (int j = 0; j < 64; j++) o[j] *= scl;
    float r32[32], r16[16], r8[8], r4[4], r2[2];
    {
        const bool up = (lane & 16) != 0;
#pragma unroll
        for (int j = 0; j < 32; j++) { const float mine = up ? o[32 + j] : o[j], send = up ? o[j] : o[32 + j]; r32[j] = mine + __shfl_xor_sync(0xffffffffu, send, 16); }
    }
    {
        const bool up = (lane & 8) != 0;
#pragma unroll
        for (int j = 0; j < 16; j++) { const float mine = up ? r32[16 + j] : r32[j], send = up ? r32[j] : r32[16 + j]; r16[j] = mine + __shfl_xor_sync(0xffffffffu, send, 8); }
    }
    {
        const bool up = (lane & 4) != 0;
#pragma unroll
        for (int j = 0; j < 8; j++) { const float mine = up ? r16[8 + j] : r16[j], send = up ? r16[j] : r16[8 + j]; r8[j] = mine + __shfl_xor_sync(0xffffffffu, send, 4); }
    }
    {
        const bool up = (lane & 2) != 0;
#pragma unroll
        for (int j = 0; j < 4; j++) { const float mine = up ? r8[4 + j] : r8[j], send = up ? r8[j] : r8[4 + j]; r4[j] = mine + __shfl_xor_sync(0xffffffffu, send, 2); }
    }
    {
        const bool up = (lane & 1) != 0;
#pragma unroll
        for (int j = 0; j < 2; j++) { const float mine = up ? r4[2 + j] : r4[j], send = up ? r4[j] : r4[2 + j]; r2[j] = mine + __shfl_xor_sync(0xffffffffu, send, 1); }
    }
    {
        float *dst = sm + warp * kPartStride;
        if (lane == 0) { dst[64] = m_w; dst[65] = l_w; }
        *reinterpret_cast<float2 *>(dst + 2 * lane) = make_float2(r2[0], r2[1]);
    }
    __syncthreads();
    if (warp != 0) return;
    // warp 0: merge the warps' partials (fixed order), two output dims per lane
    float M = -INFINITY;
    for (int w2 = 0; w2 < kDaW; w2++) M = fmaxf(M, sm[w2 * kPartStride + 64]);
    float Lsum = 0.0f, o0 = 0.0f, o1 = 0.0f;
    for (int w2 = 0; w2 < kDaW; w2++) {
        const float mw = sm[w2 * kPartStride + 64];
        const float sc = (mw == -INFINITY) ? 0.0f : __expf(mw - M);
        Lsum += sc * sm[w2 * kPartStride + 65];
        o0 += sc * sm[w2 * kPartStride + lane];
        o1 += sc * sm[w2 * kPartStride + 32 + lane];
    }
    float *orow = out + (size_t)b * ldo + h * 64;
    __half *orow16 = out16 ? out16 + (size_t)b * ldo + h * 64 : nullptr;
    if (nsplit == 1) {
        const float inv = 1.0f / Lsum;
        orow[lane] = o0 * inv;
        orow[32 + lane] = o1 * inv;
        if (orow16) { orow16[lane] = __float2half_rn(o0 * inv); orow16[32 + lane] = __float2half_rn(o1 * inv); }
        return;
    }
    float *pd = part + ((size_t)u * nsplit + sp) * kPartStride;
    if (lane == 0) { pd[64] = M; pd[65] = Lsum; }
    pd[lane] = o0;
    pd[32 + lane] = o1;
    __threadfence();
    __syncwarp();
    if (lane == 0) s_last = (atomicAdd(cnt + u, 1) == nsplit - 1) ? 1 : 0;
    __syncwarp();
    if (!s_last) return;
    __threadfence();
    float Mg = -INFINITY;
    for (int s2 = 0; s2 < nsplit; s2++) Mg = fmaxf(Mg, __ldcg(part + ((size_t)u * nsplit + s2) * kPartStride + 64));
    float Lg = 0.0f, a0 = 0.0f, a1 = 0.0f;
    for (int s2 = 0; s2 < nsplit; s2++) {                   // fixed order: deterministic whichever CTA arrives last
        const float *ps = part + ((size_t)u * nsplit + s2) * kPartStride;
        const float mw = __ldcg(ps + 64);
        const float sc = (mw == -INFINITY) ? 0.0f : __expf(mw - Mg);
        Lg += sc * __ldcg(ps + 65);
        a0 += sc * __ldcg(ps + lane);
        a1 += sc * __ldcg(ps + 32 + lane);
    }
    const float inv = 1.0f / Lg;
    orow[lane] = a0 * inv;
    orow[32 + lane] = a1 * inv;
    if (orow16) { orow16[lane] = __float2half_rn(a0 * inv); orow16[32 + lane] = __float2half_rn(a1 * inv); }
    if (lane == 0) cnt[u] = 0;
}

// Merge the partials of every (clip, head) into the attention output att[R][D] in shared memory.
//   stage A: one thread per unit loads its (max, sum) pairs, derives the merge coefficients -> shared memory
//   stage B: one thread per 4 output columns: coefficient-weighted sum of the partial o vectors (16-byte loads)
template <int R>
__device__ __noinline__ void attn_combine(const float *part, int D, int H, int kv_len, AttnPlan pl, int nrows, float *att,
                                          float *sc_s /* >= nrows*H*24 floats */)
{
    const int units = nrows * H, per = pl.per, kvp = pl.kvp;
    if (R <= 2) {
        // one clip or two: a single pass, every thread fetches the (max, sum, o) of all partials of its 4 columns at
        // once (one L2 round trip)
        const int d4 = D >> 2;
        for (int idx = threadIdx.x; idx < R * d4; idx += kDT) {
            const int m = idx / d4, c4 = idx - m * d4;
            float4 acc = make_float4(0.f, 0.f, 0.f, 0.f);
            if (m < nrows) {
                const int h = c4 >> 4, dd = (c4 & 15) * 4, u = m * H + h;
                const int f0 = u * kvp, f1 = f0 + kv_len - 1;     // < 2^31: units * kvp <= 160 * 1500
                const int c_lo = f0 / per, np = min(f1 / per - c_lo + 1, kMaxParts);
                float ms[kMaxParts], ls[kMaxParts];
                float4 v[kMaxParts];
#pragma unroll
                for (int p2 = 0; p2 < kMaxParts; p2++) {
                    const int c = c_lo + p2;
                    const int pi = c * kMaxSeg + (u - (c * per) / kvp);
                    const bool ok = p2 < np;
                    const float *pp = part + (size_t)(ok ? pi : 0) * kPartStride;
                    ms[p2] = ok ? __ldcg(pp + 64) : -INFINITY;
                    ls[p2] = ok ? __ldcg(pp + 65) : 0.0f;
                    v[p2] = ok ? __ldcg(reinterpret_cast<const float4 *>(pp + dd)) : make_float4(0.f, 0.f, 0.f, 0.f);
                }
                float M = ms[0];
#pragma unroll
                for (int p2 = 1; p2 < kMaxParts; p2++) M = fmaxf(M, ms[p2]);
                float den = 0.0f;
#pragma unroll
                for (int p2 = 0; p2 < kMaxParts; p2++) {
                    const float cfp = (ms[p2] == -INFINITY) ? 0.0f : __expf(ms[p2] - M);
                    den += cfp * ls[p2];
                    acc.x += cfp * v[p2].x; acc.y += cfp * v[p2].y; acc.z += cfp * v[p2].z; acc.w += cfp * v[p2].w;
                }
                const float inv = 1.0f / den;
                acc.x *= inv; acc.y *= inv; acc.z *= inv; acc.w *= inv;
            }
            reinterpret_cast<float4 *>(att)[idx] = acc;
        }
        __syncthreads();
        return;
    }
    for (int u = threadIdx.x; u < units; u += kDT) {
        const int f0 = u * kvp, f1 = f0 + kv_len - 1;
        const int c_lo = f0 / per, c_hi = f1 / per;
        const int np = min(c_hi - c_lo + 1, kMaxParts);
        float ms[kMaxParts], ls[kMaxParts];
        int pi[kMaxParts];
#pragma unroll
        for (int p2 = 0; p2 < kMaxParts; p2++) {
            const int c = c_lo + p2;
            const int u0 = (c * per) / kvp;                               // first unit of CTA c
            pi[p2] = c * kMaxSeg + (u - u0);
            const bool ok = p2 < np;
            ms[p2] = ok ? __ldcg(part + (size_t)pi[p2] * kPartStride + 64) : -INFINITY;
            ls[p2] = ok ? __ldcg(part + (size_t)pi[p2] * kPartStride + 65) : 0.0f;
        }
        float M = ms[0];
#pragma unroll
        for (int p2 = 1; p2 < kMaxParts; p2++) M = fmaxf(M, ms[p2]);
        float den = 0.0f, cf[kMaxParts];
#pragma unroll
        for (int p2 = 0; p2 < kMaxParts; p2++) {
            cf[p2] = (ms[p2] == -INFINITY) ? 0.0f : __expf(ms[p2] - M);
            den += cf[p2] * ls[p2];
        }
        const float inv = 1.0f / den;
        float *o = sc_s + u * 24;
        o[0] = __int_as_float(np);
#pragma unroll
        for (int p2 = 0; p2 < kMaxParts; p2++) { o[1 + p2] = cf[p2] * inv; o[12 + p2] = __int_as_float(pi[p2]); }
    }
    __syncthreads();
    const int d4 = D >> 2;
    for (int idx = threadIdx.x; idx < R * d4; idx += kDT) {
        const int m = idx / d4, c4 = idx - m * d4;
        float4 acc = make_float4(0.f, 0.f, 0.f, 0.f);
        if (m < nrows) {
            const int h = c4 >> 4, dd = (c4 & 15) * 4;
            const float *o = sc_s + (m * H + h) * 24;
            const int np = __float_as_int(o[0]);
            float4 v[kMaxParts];
#pragma unroll
            for (int p2 = 0; p2 < kMaxParts; p2++)
                if (p2 < np) v[p2] = __ldcg(reinterpret_cast<const float4 *>(part + (size_t)__float_as_int(o[12 + p2]) * kPartStride + dd));
#pragma unroll
            for (int p2 = 0; p2 < kMaxParts; p2++) {
                if (p2 < np) {
                    const float cfp = o[1 + p2];
                    acc.x += cfp * v[p2].x; acc.y += cfp * v[p2].y; acc.z += cfp * v[p2].z; acc.w += cfp * v[p2].w;
                }
            }
        }
        reinterpret_cast<float4 *>(att)[idx] = acc;
    }
    __syncthreads();
}

// ---- the decode step ---------------------------------------------------------------------------------------
template <int R>
__global__ void __launch_bounds__(kDT, 1) decode_step_kernel(const DecodeStepArgs a)
{
    extern __shared__ __align__(16) float dsm[];
    const int D = a.D, F = a.F, H = a.H, B = a.B, ctx = a.ctx;
    float *const gx = a.x, *const gq = a.q, *const gh = a.h;
    float *xs = dsm;                                   // [R][max(D, F)] activation of the current GEMV
    float *red = xs + (size_t)R * (F > D ? F : D);     // [kDT * R + 1088] partial sums / attention scratch
    const int tid = threadIdx.x;
    unsigned bar_target = 0;
    const int kv_len = a.kv_len + 1;                   // including the row appended in this step
    const AttnPlan per_self = attn_plan(B * H, kv_len, (int)gridDim.x);
    const AttnPlan per_cross = attn_plan(B * H, a.T_enc, (int)gridDim.x);
    prefetch_weight(a.layers[0].qkv, a.layers[0].ln1w, a.layers[0].ln1b, D);

    // ---- phase 0: x[b] = token_embedding[tok[b]] + positional_embedding[pos]  (decoder.rs:317-329)
    for (int b = blockIdx.x; b < B; b += gridDim.x) {
        const int tok = a.tokens[b];
        const float *te = a.tok_emb + (size_t)tok * D, *pe = a.pos_emb + (size_t)a.emb_pos * D;
        for (int j = tid; j < D; j += kDT) a.x[(size_t)b * D + j] = te[j] + pe[j];
    }
    if (blockIdx.x == 0)
        for (int j = tid; j < (R - B) * D; j += kDT) a.x[(size_t)B * D + j] = 0.0f;   // unused rows of the R-row tile
    grid_barrier(a.barrier, bar_target, a.dev_err, a.timeline);

    for (int l = 0; l < a.n_layers; l++) {
        const DecLayerW &L = a.layers[l];   // kernel-parameter constant bank
        __half *sk = a.self_k + (size_t)l * a.kv_clips * a.ctx * D, *sv = a.self_v + (size_t)l * a.kv_clips * a.ctx * D;
        const __half *ck = a.cross_kv + (size_t)l * a.kv_clips * a.T_enc * 2 * D;

        // P1: LN1 -> fused q|k|v projection; q -> a.q, k/v -> f16 cache row kv_len-1 (decoder.rs:84-92, attention.rs:99-113)
        prefetch_weight(L.out, nullptr, nullptr, D);
        ln_rows<R>(a.x, L.ln1w, L.ln1b, xs, red, D);
        gemv_phase<R>(L.qkv, xs, red, B, EpiCtx{EPI_QKV, gq, D, sk, sv, ctx, kv_len - 1, D, a.slot});
        grid_barrier(a.barrier, bar_target, a.dev_err, a.timeline);

        // P2: self-attention over the cache (attention.rs:115-119)
        attn_phase<true>(a.q, D, sk, sv, D, (long long)a.ctx * D, 64, kv_len, B, H, per_self, a.part, red, a.slot);
        grid_barrier(a.barrier, bar_target, a.dev_err, a.timeline);

        // P3: out projection + residual (decoder.rs:91-92)
        prefetch_weight(L.cq, L.ln2w, L.ln2b, D);
        attn_combine<R>(a.part, D, H, kv_len, per_self, B, xs, red);
        gemv_phase<R>(L.out, xs, red, B, EpiCtx{EPI_RESID, gx, D, nullptr, nullptr, 0, 0, D, nullptr});
        grid_barrier(a.barrier, bar_target, a.dev_err, a.timeline);

        // P4: LN2 -> cross-attention query (decoder.rs:95-99, attention.rs:214-217)
        prefetch_weight(L.cout, nullptr, nullptr, D);
        ln_rows<R>(a.x, L.ln2w, L.ln2b, xs, red, D);
        gemv_phase<R>(L.cq, xs, red, B, EpiCtx{EPI_STORE, gq, D, nullptr, nullptr, 0, 0, D, nullptr});
        grid_barrier(a.barrier, bar_target, a.dev_err, a.timeline);

        // P5: cross-attention over the cached encoder K/V (attention.rs:219-233)
        attn_phase<false>(a.q, D, ck, ck + 64, 128, (long long)a.T_enc * 2 * D, (long long)a.T_enc * 128, a.T_enc, B, H, per_cross, a.part, red, a.slot);
        grid_barrier(a.barrier, bar_target, a.dev_err, a.timeline);

        // P6: cross out projection + residual (decoder.rs:103-104)
        prefetch_weight(L.fc1, L.ln3w, L.ln3b, D);
        attn_combine<R>(a.part, D, H, a.T_enc, per_cross, B, xs, red);
        gemv_phase<R>(L.cout, xs, red, B, EpiCtx{EPI_RESID, gx, D, nullptr, nullptr, 0, 0, D, nullptr});
        grid_barrier(a.barrier, bar_target, a.dev_err, a.timeline);

        // P7: LN3 -> fc1 + tanh-GELU (layers.rs:35-41,55-56)
        prefetch_weight(L.fc2, nullptr, nullptr, D);
        ln_rows<R>(a.x, L.ln3w, L.ln3b, xs, red, D);
        gemv_phase<R>(L.fc1, xs, red, B, EpiCtx{EPI_GELU, gh, F, nullptr, nullptr, 0, 0, D, nullptr});
        grid_barrier(a.barrier, bar_target, a.dev_err, a.timeline);

        // P8: fc2 + residual (layers.rs:57, decoder.rs:108-110)
        if (l + 1 < a.n_layers) prefetch_weight(a.layers[l + 1].qkv, a.layers[l + 1].ln1w, a.layers[l + 1].ln1b, D);
        else prefetch_weight(a.layers[l].fc2, a.lnfw, a.lnfb, D);   // final LayerNorm parameters
        for (int i = tid; i < R * F / 4; i += kDT)
            reinterpret_cast<float4 *>(xs)[i] = __ldcg(reinterpret_cast<const float4 *>(a.h) + i);
        __syncthreads();
        gemv_phase<R>(L.fc2, xs, red, B, EpiCtx{EPI_RESID, gx, D, nullptr, nullptr, 0, 0, D, nullptr});
        grid_barrier(a.barrier, bar_target, a.dev_err, a.timeline);
    }

    // ---- final LayerNorm + logits = x . token_embedding^T (decoder.rs:338-343).  The matrix is an f16 copy stored
    // transposed in 16-byte chunks [D/8][Vp]: lane = vocabulary row, so loads are coalesced, x is a shared-memory
    // broadcast and no cross-lane reduction is needed.
    ln_rows<R>(a.x, a.lnfw, a.lnfb, xs, red, D);
    const int warp = tid >> 5, lane = tid & 31;
    const int V = a.V, Vp = a.Vp;
    const int chunks = D >> 3;
    const bool pick = a.arg_hi > a.arg_lo && a.next_out != nullptr;
    float best[R];
    int besti[R];
#pragma unroll
    for (int m = 0; m < R; m++) { best[m] = -INFINITY; besti[m] = -1; }
    const int ngroups = Vp >> 5;                        // groups of 32 vocabulary rows
    for (int gidx = blockIdx.x * (kDT / 32) + warp; gidx < ngroups; gidx += (int)gridDim.x * (kDT / 32)) {
        const int v = gidx * 32 + lane;
        const uint4 *ep = a.tok_emb16t + v;
        float acc[R];
#pragma unroll
        for (int m = 0; m < R; m++) acc[m] = 0.0f;
        for (int c0 = 0; c0 < chunks; c0 += 8) {
            uint4 e[8];
#pragma unroll
            for (int u = 0; u < 8; u++)
                if (c0 + u < chunks) e[u] = ldg_stream16(ep + (size_t)(c0 + u) * Vp);
#pragma unroll
            for (int u = 0; u < 8; u++) {
                if (c0 + u < chunks) {
                    const __half2 *hp = reinterpret_cast<const __half2 *>(&e[u]);
                    const float2 f0 = __half22float2(hp[0]), f1 = __half22float2(hp[1]);
                    const float2 f2 = __half22float2(hp[2]), f3 = __half22float2(hp[3]);
#pragma unroll
                    for (int m = 0; m < R; m++) {
                        const float4 xa = *reinterpret_cast<const float4 *>(xs + (size_t)m * D + (c0 + u) * 8);
                        const float4 xb = *reinterpret_cast<const float4 *>(xs + (size_t)m * D + (c0 + u) * 8 + 4);
                        float s = acc[m];
                        s = fmaf(f0.x, xa.x, s); s = fmaf(f0.y, xa.y, s); s = fmaf(f1.x, xa.z, s); s = fmaf(f1.y, xa.w, s);
                        s = fmaf(f2.x, xb.x, s); s = fmaf(f2.y, xb.y, s); s = fmaf(f3.x, xb.z, s); s = fmaf(f3.y, xb.w, s);
                        acc[m] = s;
                    }
                }
            }
        }
        if (v < V) {
#pragma unroll
            for (int m = 0; m < R; m++) {
                if (m < B) {
                    a.logits[(size_t)m * V + v] = acc[m];
                    const int sup = a.suppress ? a.suppress[m] : -1;
                    // whisper.rs:131-138 max_by: the LAST maximum wins; a lane's rows ascend, so ">=" keeps the later one
                    if (pick && v >= a.arg_lo && v < a.arg_hi && v != sup && (besti[m] < 0 || acc[m] >= best[m])) { best[m] = acc[m]; besti[m] = v; }
                }
            }
        }
    }
    // warp-level merge of the 32 lanes' candidates (higher index wins ties)
#pragma unroll
    for (int m = 0; m < R; m++) {
#pragma unroll
        for (int off = 16; off > 0; off >>= 1) {
            const float ov = __shfl_xor_sync(0xffffffffu, best[m], off);
            const int oi = __shfl_xor_sync(0xffffffffu, besti[m], off);
            if (oi >= 0 && (besti[m] < 0 || ov > best[m] || (ov == best[m] && oi > besti[m]))) { best[m] = ov; besti[m] = oi; }
        }
    }
    if (!pick) return;                                  // uniform over the grid
    // CTA-level merge (warps ascend in v), then grid-level merge by CTA 0
    __syncthreads();
    if (lane == 0) {
#pragma unroll
        for (int m = 0; m < R; m++) { red[(warp * R + m) * 2] = best[m]; red[(warp * R + m) * 2 + 1] = __int_as_float(besti[m]); }
    }
    __syncthreads();
    if (tid < B) {
        float bv = -INFINITY; int bi = -1;
        for (int w2 = 0; w2 < kDT / 32; w2++) {
            const float v = red[(w2 * R + tid) * 2]; const int i = __float_as_int(red[(w2 * R + tid) * 2 + 1]);
            if (i >= 0 && (bi < 0 || v > bv || (v == bv && i > bi))) { bv = v; bi = i; }
        }
        a.arg_scratch[((size_t)blockIdx.x * kMaxR + tid) * 2] = bv;
        a.arg_scratch[((size_t)blockIdx.x * kMaxR + tid) * 2 + 1] = __int_as_float(bi);
    }
    grid_barrier(a.barrier, bar_target, a.dev_err, a.timeline);
    if (blockIdx.x == 0 && tid < B) {
        float bv = -INFINITY; int bi = -1;
        for (int c = 0; c < (int)gridDim.x; c++) {
            const float v = __ldcg(a.arg_scratch + ((size_t)c * kMaxR + tid) * 2);
            const int i = __float_as_int(__ldcg(a.arg_scratch + ((size_t)c * kMaxR + tid) * 2 + 1));
            if (i >= 0 && (bi < 0 || v > bv || (v == bv && i > bi))) { bv = v; bi = i; }
        }
        a.next_out[tid] = bi;
    }
}

// panel layout -> decode layout (one thread per destination block)
__global__ void decw_build_kernel(const uint4 *__restrict__ pnib, const __half *__restrict__ pscl, uint4 *__restrict__ dnib,
                                  __half *__restrict__ dscl, int n, int kb, int kbp, int rpc, int n_cta)
{
    const long long total = (long long)n_cta * kb * rpc;
    for (long long i = (long long)blockIdx.x * blockDim.x + threadIdx.x; i < total; i += (long long)gridDim.x * blockDim.x) {
        const int slot = (int)(i % rpc);
        const int kbi = (int)((i / rpc) % kb);
        const int cta = (int)(i / ((long long)rpc * kb));
        const int row = cta * rpc + slot;
        if (row < n) {
            const size_t src = ((size_t)(row / WQ4_PANEL_ROWS) * kbp + kbi) * WQ4_PANEL_ROWS + (row % WQ4_PANEL_ROWS);
            dnib[i] = pnib[src];
            dscl[i] = pscl[src];
        } else {
            dnib[i] = make_uint4(0x88888888u, 0x88888888u, 0x88888888u, 0x88888888u);
            dscl[i] = __ushort_as_half((unsigned short)0);
        }
    }
}

// f32 [V][D] -> f16 16-byte chunks [D/8][Vp]; rows >= V are zero
__global__ void logits_matrix_kernel(const float *__restrict__ emb, uint4 *__restrict__ dst, int V, int Vp, int D)
{
    const long long total = (long long)(D >> 3) * Vp;
    for (long long i = (long long)blockIdx.x * blockDim.x + threadIdx.x; i < total; i += (long long)gridDim.x * blockDim.x) {
        const int v = (int)(i % Vp), c = (int)(i / Vp);
        uint4 o = make_uint4(0, 0, 0, 0);
        if (v < V) {
            const float *src = emb + (size_t)v * D + c * 8;
            uint32_t h[8];
#pragma unroll
            for (int j = 0; j < 8; j++) h[j] = f32_to_f16_sat_d(src[j]);
            o = make_uint4(h[0] | (h[1] << 16), h[2] | (h[3] << 16), h[4] | (h[5] << 16), h[6] | (h[7] << 16));
        }
        dst[i] = o;
    }
}

template <int R>
int32_t launch_decode(const DecodeStepArgs &a, int device, cudaStream_t st)
{
    const size_t red_floats = (size_t)kDT * R > (size_t)kMmaRedFloats ? (size_t)kDT * R : (size_t)kMmaRedFloats;
    const size_t smem = ((size_t)R * (a.F > a.D ? a.F : a.D) + red_floats + 1088) * sizeof(float);
    auto kern = decode_step_kernel<R>;
    static std::atomic<bool> attr_done[64];
    if (device < 0 || device >= 64 || !attr_done[device]) {
        WQ4_CUDA(cudaFuncSetAttribute(kern, cudaFuncAttributeMaxDynamicSharedMemorySize, 200 * 1024));
        if (device >= 0 && device < 64) attr_done[device] = true;
    }
    if (smem > 200 * 1024) { wq4_set_error("decode step: %zu bytes of shared memory needed (R=%d, F=%d)", smem, R, a.F); return WQ4_ERR_UNSUPPORTED; }
    WQ4_CUDA(cudaMemsetAsync(a.barrier, 0, sizeof(unsigned), st));
    cudaLaunchConfig_t cfg = {};
    cfg.gridDim = dim3((unsigned)wq4m_decode_grid(device), 1, 1);
    cfg.blockDim = dim3(kDT, 1, 1);
    cfg.dynamicSmemBytes = smem;
    cfg.stream = st;
    cudaLaunchAttribute attr[1];
    attr[0].id = cudaLaunchAttributeCooperative;       // co-residency of all CTAs is required by the grid barrier
    attr[0].val.cooperative = 1;
    cfg.attrs = attr;
    cfg.numAttrs = 1;
    WQ4_CUDA(cudaLaunchKernelEx(&cfg, kern, a));
    wq4_count_launch();
    return WQ4_OK;
}

}  // namespace

int wq4m_decode_grid(int device) { return wq4_sm_count(device); }

// Debug hook (not part of the drop-in surface): dev_ptr = kDecTimelineSlots x int64 receiving CTA 0's globaltimer (ns)
// after every grid barrier of the following decode steps; NULL disables.
static long long *g_dec_timeline_host = nullptr;
extern "C" WQ4_API void wq4_debug_decode_timeline(void *dev_ptr) { g_dec_timeline_host = reinterpret_cast<long long *>(dev_ptr); }

int32_t wq4m_decw_create(const wq4_tensor *t, const float *bias_dev, int n_cta, DecWeight *out, void **alloc_nib, void **alloc_scl,
                         cudaStream_t st)
{
    if (!t || !t->matrix_mode) { wq4_set_error("decode layout needs a [N,K] tensor with K %% 32 == 0"); return WQ4_ERR_UNSUPPORTED; }
    const int n = (int)t->n, kb = (int)t->kb;
    const int rpc = (n + n_cta - 1) / n_cta;
    if (rpc > kDT) { wq4_set_error("decode layout: %d rows per CTA exceed %d", rpc, kDT); return WQ4_ERR_UNSUPPORTED; }
    const size_t total = (size_t)n_cta * kb * rpc;
    uint4 *dn = nullptr;
    __half *ds = nullptr;
    WQ4_CUDA(cudaMalloc(&dn, total * sizeof(uint4)));
    WQ4_CUDA(cudaMalloc(&ds, total * sizeof(__half) + 16));
    decw_build_kernel<<<(unsigned)((total + 255) / 256 > 4096 ? 4096 : (total + 255) / 256), 256, 0, st>>>(
        t->nib, t->scl, dn, ds, n, kb, (int)t->kbp, rpc, n_cta);
    wq4_count_launch();
    WQ4_CUDA(cudaGetLastError());
    out->nib = dn; out->scl = ds; out->bias = bias_dev; out->n = n; out->k = (int)t->k; out->kb = kb; out->rpc = rpc;
    *alloc_nib = dn; *alloc_scl = ds;
    return WQ4_OK;
}

int32_t wq4m_logits_matrix(const float *emb, void *dst, int64_t V, int64_t Vp, int64_t D, cudaStream_t st)
{
    logits_matrix_kernel<<<4096, 256, 0, st>>>(emb, reinterpret_cast<uint4 *>(dst), (int)V, (int)Vp, (int)D);
    wq4_count_launch();
    WQ4_CUDA(cudaGetLastError());
    return WQ4_OK;
}

// q [B][ldq] f32, K/V f16 caches (row stride ldkv, clip stride kv_batch_stride elements), out [B][ldo] f32 (+ optional f16
// copy out16 for the next GEMM); one query row
// per clip.  part: >= B*H*nsplit*68 floats, cnt: >= B*H ints (zero; left zero).  nsplit is chosen here.
int32_t wq4m_dec_attention_split(const float *q, int64_t ldq, const void *kc, const void *vc, int64_t ldkv, int64_t kv_batch_stride,
                                 int64_t hstride, float *out, void *out16, int64_t ldo, int64_t B, int64_t H, int64_t kv_len, float *part,
                                 int64_t part_floats, int *cnt, int device, cudaStream_t st, int64_t append_off, const int *slot)
{
    const int64_t units = B * H;
    if (units < 1 || kv_len < 1) return WQ4_OK;
    const int sms = wq4_sm_count(device);
    // Split the keys only when there are fewer units than half the SMs: a split costs the partial write, a fence, an atomic
    // and the merge by the last arriver.  Measured (scripts/sweep_da.sh, Large-V3 step): 80 units x 1500 keys unsplit on the
    // 8-warp kernel 1.75 ms per step against 1.82 ms split in four and 1.91 ms in eight; 320 units unsplit 2.58 ms against
    // 2.76-2.91 ms split; a 225-key self-attention row split in two costs 15 -> 23 us.
    int64_t ns = ((int64_t)sms / 2 + units - 1) / units;
    if (ns > 8) ns = 8;
    while (ns > 1 && (kv_len + ns - 1) / ns < 128) ns--;
    while (ns > 1 && units * ns * kPartStride > part_floats) ns--;
    int64_t keys_per_cta = (kv_len + ns - 1) / ns;
    bool wide = keys_per_cta > 512 || units * ns <= sms;      // few units: eight warps per unit even on short rows
    {   // tuning overrides: WQ4_DA_CROSS / WQ4_DA_SELF = "<splits>,<0|1 wide>" (kv_len > 512 counts as cross)
        struct Ov { int ns[2] = {-1, -1}, w[2] = {-1, -1}; };
        static const Ov ov = [] {              // read once, thread-safely
            Ov o;
            const char *e0 = getenv("WQ4_DA_SELF"), *e1 = getenv("WQ4_DA_CROSS");
            if (e0) sscanf(e0, "%d,%d", &o.ns[0], &o.w[0]);
            if (e1) sscanf(e1, "%d,%d", &o.ns[1], &o.w[1]);
            return o;
        }();
        const int *o_ns = ov.ns, *o_w = ov.w;
        const int c = kv_len > 512 ? 1 : 0;
        if (o_ns[c] > 0) {
            ns = o_ns[c];
            while (ns > 1 && units * ns * kPartStride > part_floats) ns--;
            keys_per_cta = (kv_len + ns - 1) / ns;
        }
        if (o_w[c] >= 0) wide = o_w[c] != 0;
    }
    static std::atomic<bool> configured[64];
    if (!configured[device & 63]) {
        WQ4_CUDA(cudaFuncSetAttribute(dec_attention_split_kernel<8>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)da_smem(8)));
        WQ4_CUDA(cudaFuncSetAttribute(dec_attention_split_kernel<2>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)da_smem(2)));
        configured[device & 63] = true;
    }
    cudaLaunchConfig_t cfg = {};
    cfg.gridDim = dim3((unsigned)units, (unsigned)ns, 1);
    cfg.blockDim = dim3(wide ? 256 : 64, 1, 1);
    cfg.dynamicSmemBytes = da_smem(wide ? 8 : 2);
    cfg.stream = st;
    cudaLaunchAttribute attr[1];
    attr[0].id = cudaLaunchAttributeProgrammaticStreamSerialization;
    attr[0].val.programmaticStreamSerializationAllowed = 1;
    cfg.attrs = attr;
    cfg.numAttrs = wq4_pdl_enabled() ? 1 : 0;
    if (wide)
        WQ4_CUDA(cudaLaunchKernelEx(&cfg, dec_attention_split_kernel<8>, q, (int)ldq, reinterpret_cast<const __half *>(kc),
                                    reinterpret_cast<const __half *>(vc), (long long)ldkv, (long long)kv_batch_stride, (long long)hstride, out,
                                    reinterpret_cast<__half *>(out16), (int)ldo, (int)kv_len, (int)H, (int)ns, (int)append_off, part, cnt, slot));
    else
        WQ4_CUDA(cudaLaunchKernelEx(&cfg, dec_attention_split_kernel<2>, q, (int)ldq, reinterpret_cast<const __half *>(kc),
                                    reinterpret_cast<const __half *>(vc), (long long)ldkv, (long long)kv_batch_stride, (long long)hstride, out,
                                    reinterpret_cast<__half *>(out16), (int)ldo, (int)kv_len, (int)H, (int)ns, (int)append_off, part, cnt, slot));
    wq4_count_launch();
    WQ4_CUDA(cudaGetLastError());
    return WQ4_OK;
}

int32_t wq4m_decode_step(const DecodeStepArgs &a_in, int device, cudaStream_t st)
{
    DecodeStepArgs a = a_in;
    a.timeline = g_dec_timeline_host;
    if (a.B < 1 || a.B > kMaxR) { wq4_set_error("decode step kernel handles 1..8 clips, got %d", a.B); return WQ4_ERR_INVALID; }
    if (a.D % 64 != 0 || a.D > 1280 || a.D != a.H * 64) { wq4_set_error("decode step kernel: unsupported d_model %d / heads %d", a.D, a.H); return WQ4_ERR_UNSUPPORTED; }
    WqProfScope prof(WQ4_PROF_GEMV, 0.0, st);
    if (a.B == 1) return launch_decode<1>(a, device, st);
    if (a.B == 2) return launch_decode<2>(a, device, st);
    if (a.B <= 4) return launch_decode<4>(a, device, st);
    return launch_decode<8>(a, device, st);
}
