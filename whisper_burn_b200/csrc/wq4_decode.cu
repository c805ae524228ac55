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

#include "wq4_common.cuh"
#include "wq4_model.cuh"

namespace {

constexpr int kDT = 512;          // threads per CTA
constexpr int kMaxR = 8;          // clips per launch
constexpr int kPartStride = 66;   // (m, l, o[64]) per attention partial
constexpr int kMaxSplit = 8;

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
__device__ __forceinline__ void grid_barrier(unsigned *counter, unsigned &target, int *dev_err)
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

// acc[m] += d * sum_k (nib - 8) * xs[m][kbi*32 + k]   for one block
template <int R>
__device__ __forceinline__ void block_dot(const uint4 &q, float d, const float *xs, int K, int kbi, float (&acc)[R])
{
    float bs[R];
#pragma unroll
    for (int m = 0; m < R; m++) bs[m] = 0.0f;
    const uint32_t w[4] = {q.x, q.y, q.z, q.w};
#pragma unroll
    for (int c = 0; c < 4; c++) {
        float f[8];
        word_to_f32(w[c], f);
#pragma unroll
        for (int m = 0; m < R; m++) {
            const float4 xa = *reinterpret_cast<const float4 *>(xs + (size_t)m * K + kbi * 32 + c * 8);
            const float4 xb = *reinterpret_cast<const float4 *>(xs + (size_t)m * K + kbi * 32 + c * 8 + 4);
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
template <int R, class Epi>
__device__ __forceinline__ void gemv_phase(const DecWeight &w, const float *xs, float *red, int nrows, Epi epi)
{
    const int tid = threadIdx.x;
    const int rpc = w.rpc, kb = w.kb, K = w.k;
    const int row0 = blockIdx.x * rpc;
    int KL = kDT / rpc;
    if (KL > kb) KL = kb;
    const int kl = tid / rpc, slot = tid - kl * rpc;
    const bool active = kl < KL && row0 < w.n;
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
                if (kbi < kb) {
                    q[u] = ldg_stream16(w.nib + cbase + (size_t)kbi * rpc);
                    d[u] = __half2float(__ldg(w.scl + cbase + (size_t)kbi * rpc));
                }
            }
#pragma unroll
            for (int u = 0; u < 4; u++) {
                const int kbi = kb0 + u * KL;
                if (kbi < kb) block_dot<R>(q[u], d[u], xs, K, kbi, acc);
            }
        }
#pragma unroll
        for (int m = 0; m < R; m++) red[(kl * rpc + slot) * R + m] = acc[m];
    }
    __syncthreads();
    for (int t = tid; t < rpc * R; t += kDT) {
        const int s2 = t / R, m = t - s2 * R;
        const int n = row0 + s2;
        if (n < w.n && m < nrows) {
            float s = 0.0f;
            for (int l = 0; l < KL; l++) s += red[(l * rpc + s2) * R + m];   // fixed order: deterministic
            if (w.bias) s += w.bias[n];
            epi(m, n, s);
        }
    }
    __syncthreads();
}

// LayerNorm::forward (src/model/layers.rs:23-31) of R rows of a global [R][D] f32 matrix into shared memory.
// Warp m handles row m.  The rows were written by other CTAs in an earlier phase: read through L2 (ld.cg).
template <int R>
__device__ __forceinline__ void ln_rows(const float *x, const float *w, const float *b, float *xs, int D)
{
    const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
    if (warp < R) {
        const float4 *xr = reinterpret_cast<const float4 *>(x + (size_t)warp * D);
        float4 *row = reinterpret_cast<float4 *>(xs + (size_t)warp * D);
        const int cnt = D >> 2;            // float4 per row (D is a multiple of 64)
        float s = 0.0f;
        for (int i = lane; i < cnt; i += 32) {
            const float4 v = __ldcg(xr + i);
            row[i] = v;
            s += (v.x + v.y) + (v.z + v.w);
        }
        const float mean = warp_sum_d(s) / (float)D;
        float qv = 0.0f;
        for (int i = lane; i < cnt; i += 32) {
            const float4 v = row[i];
            const float a0 = v.x - mean, a1 = v.y - mean, a2 = v.z - mean, a3 = v.w - mean;
            qv += (a0 * a0 + a1 * a1) + (a2 * a2 + a3 * a3);
        }
        const float inv = 1.0f / sqrtf(warp_sum_d(qv) / (float)D + 1e-5f);
        for (int i = lane; i < cnt; i += 32) {
            const float4 v = row[i];
            const float4 ww = reinterpret_cast<const float4 *>(w)[i], bb = reinterpret_cast<const float4 *>(b)[i];
            float4 o;
            o.x = (v.x - mean) * inv * ww.x + bb.x;
            o.y = (v.y - mean) * inv * ww.y + bb.y;
            o.z = (v.z - mean) * inv * ww.z + bb.z;
            o.w = (v.w - mean) * inv * ww.w + bb.w;
            row[i] = o;
        }
    }
    __syncthreads();
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

// Units (clip b, head h, split s) are dealt round-robin to the CTAs.  q: global f32 [B][D] (written in the
// previous phase); K/V: f16 rows of D (head h at column h*64), row stride ldkv, clip stride bstride.
// part[((b*H + h)*nsplit + s)*66] = (running max, sum, unnormalised o[64]).
template <bool kFreshKV>
__device__ __forceinline__ void attn_phase(const float *q, int D, const __half *kc, const __half *vc, long long ldkv,
                                           long long bstride, int kv_len, int B, int H, int nsplit, float *part, float *sm)
{
    const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
    const int g = lane >> 3, dch = lane & 7;
    const int units = B * H * nsplit;
    const int per = (kv_len + nsplit - 1) / nsplit;
    for (int u = blockIdx.x; u < units; u += gridDim.x) {
        const int s = u % nsplit, h = (u / nsplit) % H, b = u / (nsplit * H);
        const int k0 = s * per, k1 = min(kv_len, k0 + per);
        float qv[8];
        {
            const float *qp = q + (size_t)b * D + h * 64 + dch * 8;
            const float4 a = __ldcg(reinterpret_cast<const float4 *>(qp));
            const float4 c = __ldcg(reinterpret_cast<const float4 *>(qp) + 1);
            // (q . k) / sqrt(64) (attention.rs:266-267): scaling q by 1/8 first is exact (power of two)
            qv[0] = a.x * 0.125f; qv[1] = a.y * 0.125f; qv[2] = a.z * 0.125f; qv[3] = a.w * 0.125f;
            qv[4] = c.x * 0.125f; qv[5] = c.y * 0.125f; qv[6] = c.z * 0.125f; qv[7] = c.w * 0.125f;
        }
        const __half *kb_ = kc + (size_t)b * bstride + h * 64 + dch * 8;
        const __half *vb_ = vc + (size_t)b * bstride + h * 64 + dch * 8;
        Soft st;
        st.m = -INFINITY; st.l = 0.0f;
#pragma unroll
        for (int j = 0; j < 8; j++) st.o[j] = 0.0f;
        for (int base0 = k0 + warp * 4; base0 < k1; base0 += 128) {   // warp-uniform trip count (shuffles inside)
            const int key0 = base0 + g;
            // two keys per lane group in flight
            uint4 kq[2], vq[2];
            bool ok[2];
#pragma unroll
            for (int t = 0; t < 2; t++) {
                const int key = key0 + t * 64;
                ok[t] = key < k1;
                if (ok[t]) {
                    const __half *kp = kb_ + (size_t)key * ldkv, *vp = vb_ + (size_t)key * ldkv;
                    if (kFreshKV) { kq[t] = ldcg16(kp); vq[t] = ldcg16(vp); }
                    else { kq[t] = ldg_stream16(reinterpret_cast<const uint4 *>(kp)); vq[t] = ldg_stream16(reinterpret_cast<const uint4 *>(vp)); }
                }
            }
#pragma unroll
            for (int t = 0; t < 2; t++) {
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
            if (dch == 0) { dst[0] = st.m; dst[1] = st.l; }
#pragma unroll
            for (int j = 0; j < 8; j++) dst[2 + dch * 8 + j] = st.o[j];
        }
        __syncthreads();
        if (warp == 0) {
            float M = -INFINITY;
            for (int w2 = 0; w2 < kDT / 32; w2++) M = fmaxf(M, sm[w2 * kPartStride]);
            float Lsum = 0.0f, o0 = 0.0f, o1 = 0.0f;
            for (int w2 = 0; w2 < kDT / 32; w2++) {      // fixed order
                const float mw = sm[w2 * kPartStride];
                const float sc = (mw == -INFINITY) ? 0.0f : __expf(mw - M);
                Lsum += sc * sm[w2 * kPartStride + 1];
                o0 += sc * sm[w2 * kPartStride + 2 + lane];
                o1 += sc * sm[w2 * kPartStride + 2 + 32 + lane];
            }
            float *pd = part + (size_t)u * kPartStride;
            if (lane == 0) { pd[0] = M; pd[1] = Lsum; }
            pd[2 + lane] = o0;
            pd[2 + 32 + lane] = o1;
        }
        __syncthreads();
    }
}

// Merge the key splits of every (clip, head) into the attention output att[R][D] in shared memory.
template <int R>
__device__ __forceinline__ void attn_combine(const float *part, int D, int H, int nsplit, int nrows, float *att)
{
    for (int idx = threadIdx.x; idx < R * D; idx += kDT) {
        const int m = idx / D, col = idx - m * D;
        if (m >= nrows) { att[idx] = 0.0f; continue; }
        const int h = col >> 6, d = col & 63;
        const float *pp = part + (size_t)((m * H + h) * nsplit) * kPartStride;
        float M = -INFINITY;
        for (int s = 0; s < nsplit; s++) M = fmaxf(M, __ldcg(pp + s * kPartStride));
        float num = 0.0f, den = 0.0f;
        for (int s = 0; s < nsplit; s++) {
            const float ms = __ldcg(pp + s * kPartStride);
            const float sc = (ms == -INFINITY) ? 0.0f : __expf(ms - M);
            den += sc * __ldcg(pp + s * kPartStride + 1);
            num += sc * __ldcg(pp + s * kPartStride + 2 + d);
        }
        att[idx] = num / den;
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
    int nsplit_self, nsplit_cross;
    {
        int ns = (int)gridDim.x / (B * H);
        if (ns < 1) ns = 1;
        if (ns > kMaxSplit) ns = kMaxSplit;
        nsplit_cross = ns;
        nsplit_self = ns;
        const int kvl = a.kv_len + 1;
        if (nsplit_self > (kvl + 63) / 64) nsplit_self = (kvl + 63) / 64;   // at least 64 keys per split
    }
    const int kv_len = a.kv_len + 1;                   // including the row appended in this step

    // ---- phase 0: x[b] = token_embedding[tok[b]] + positional_embedding[pos]  (decoder.rs:317-329)
    for (int b = blockIdx.x; b < B; b += gridDim.x) {
        const int tok = a.tokens[b];
        const float *te = a.tok_emb + (size_t)tok * D, *pe = a.pos_emb + (size_t)a.emb_pos * D;
        for (int j = tid; j < D; j += kDT) a.x[(size_t)b * D + j] = te[j] + pe[j];
    }
    if (blockIdx.x == 0)
        for (int j = tid; j < (R - B) * D; j += kDT) a.x[(size_t)B * D + j] = 0.0f;   // unused rows of the R-row tile
    grid_barrier(a.barrier, bar_target, a.dev_err);

    for (int l = 0; l < a.n_layers; l++) {
        const DecLayerW &L = a.layers[l];
        __half *sk = a.self_k + (size_t)l * B * a.ctx * D, *sv = a.self_v + (size_t)l * B * a.ctx * D;
        const __half *ck = a.cross_kv + (size_t)l * B * a.T_enc * 2 * D;

        // P1: LN1 -> fused q|k|v projection; q -> a.q, k/v -> f16 cache row kv_len-1 (decoder.rs:84-92, attention.rs:99-113)
        ln_rows<R>(a.x, L.ln1w, L.ln1b, xs, D);
        gemv_phase<R>(L.qkv, xs, red, B, [=](int m, int n, float v) {
            if (n < D) gq[(size_t)m * D + n] = v;
            else if (n < 2 * D) sk[((size_t)m * ctx + (kv_len - 1)) * D + (n - D)] = __ushort_as_half(f32_to_f16_sat_d(v));
            else sv[((size_t)m * ctx + (kv_len - 1)) * D + (n - 2 * D)] = __ushort_as_half(f32_to_f16_sat_d(v));
        });
        grid_barrier(a.barrier, bar_target, a.dev_err);

        // P2: self-attention over the cache (attention.rs:115-119)
        attn_phase<true>(a.q, D, sk, sv, D, (long long)a.ctx * D, kv_len, B, H, nsplit_self, a.part, red);
        grid_barrier(a.barrier, bar_target, a.dev_err);

        // P3: out projection + residual (decoder.rs:91-92)
        attn_combine<R>(a.part, D, H, nsplit_self, B, xs);
        gemv_phase<R>(L.out, xs, red, B, [=](int m, int n, float v) {
            float *p = gx + (size_t)m * D + n;
            *p = __ldcg(p) + v;
        });
        grid_barrier(a.barrier, bar_target, a.dev_err);

        // P4: LN2 -> cross-attention query (decoder.rs:95-99, attention.rs:214-217)
        ln_rows<R>(a.x, L.ln2w, L.ln2b, xs, D);
        gemv_phase<R>(L.cq, xs, red, B, [=](int m, int n, float v) { gq[(size_t)m * D + n] = v; });
        grid_barrier(a.barrier, bar_target, a.dev_err);

        // P5: cross-attention over the cached encoder K/V (attention.rs:219-233)
        attn_phase<false>(a.q, D, ck, ck + D, 2 * D, (long long)a.T_enc * 2 * D, a.T_enc, B, H, nsplit_cross, a.part, red);
        grid_barrier(a.barrier, bar_target, a.dev_err);

        // P6: cross out projection + residual (decoder.rs:103-104)
        attn_combine<R>(a.part, D, H, nsplit_cross, B, xs);
        gemv_phase<R>(L.cout, xs, red, B, [=](int m, int n, float v) {
            float *p = gx + (size_t)m * D + n;
            *p = __ldcg(p) + v;
        });
        grid_barrier(a.barrier, bar_target, a.dev_err);

        // P7: LN3 -> fc1 + tanh-GELU (layers.rs:35-41,55-56)
        ln_rows<R>(a.x, L.ln3w, L.ln3b, xs, D);
        gemv_phase<R>(L.fc1, xs, red, B, [=](int m, int n, float v) { gh[(size_t)m * F + n] = wq4_gelu_tanh(v); });
        grid_barrier(a.barrier, bar_target, a.dev_err);

        // P8: fc2 + residual (layers.rs:57, decoder.rs:108-110)
        for (int i = tid; i < R * F / 4; i += kDT)
            reinterpret_cast<float4 *>(xs)[i] = __ldcg(reinterpret_cast<const float4 *>(a.h) + i);
        __syncthreads();
        gemv_phase<R>(L.fc2, xs, red, B, [=](int m, int n, float v) {
            float *p = gx + (size_t)m * D + n;
            *p = __ldcg(p) + v;
        });
        grid_barrier(a.barrier, bar_target, a.dev_err);
    }

    // ---- final LayerNorm + logits = x . token_embedding^T (decoder.rs:338-343), f16 copy of the matrix
    ln_rows<R>(a.x, a.lnfw, a.lnfb, xs, D);
    const int warp = tid >> 5, lane = tid & 31;
    const int V = a.V;
    const int nwarps = (int)gridDim.x * (kDT / 32);
    const int rpw = (V + nwarps - 1) / nwarps;
    const int gw = blockIdx.x * (kDT / 32) + warp;
    const int v0 = gw * rpw, v1 = min(V, v0 + rpw);
    const int chunks = D >> 3;                          // 16-byte chunks per row
    const bool pick = a.arg_hi > a.arg_lo && a.next_out != nullptr;
    float best[R];
    int besti[R];
    int sup[R];
#pragma unroll
    for (int m = 0; m < R; m++) { best[m] = -INFINITY; besti[m] = -1; sup[m] = (a.suppress && m < B) ? a.suppress[m] : -1; }
    for (int v = v0; v < v1; v++) {
        const uint4 *er = reinterpret_cast<const uint4 *>(a.tok_emb16 + (size_t)v * D);
        float acc[R];
#pragma unroll
        for (int m = 0; m < R; m++) acc[m] = 0.0f;
        for (int c0 = lane; c0 < chunks; c0 += 128) {
            uint4 e[4];
#pragma unroll
            for (int u = 0; u < 4; u++)
                if (c0 + 32 * u < chunks) e[u] = ldg_stream16(er + c0 + 32 * u);
#pragma unroll
            for (int u = 0; u < 4; u++) {
                const int c = c0 + 32 * u;
                if (c < chunks) {
                    const __half2 *hp = reinterpret_cast<const __half2 *>(&e[u]);
                    const float2 f0 = __half22float2(hp[0]), f1 = __half22float2(hp[1]);
                    const float2 f2 = __half22float2(hp[2]), f3 = __half22float2(hp[3]);
#pragma unroll
                    for (int m = 0; m < R; m++) {
                        const float4 xa = *reinterpret_cast<const float4 *>(xs + (size_t)m * D + c * 8);
                        const float4 xb = *reinterpret_cast<const float4 *>(xs + (size_t)m * D + c * 8 + 4);
                        float s = acc[m];
                        s = fmaf(f0.x, xa.x, s); s = fmaf(f0.y, xa.y, s); s = fmaf(f1.x, xa.z, s); s = fmaf(f1.y, xa.w, s);
                        s = fmaf(f2.x, xb.x, s); s = fmaf(f2.y, xb.y, s); s = fmaf(f3.x, xb.z, s); s = fmaf(f3.y, xb.w, s);
                        acc[m] = s;
                    }
                }
            }
        }
#pragma unroll
        for (int m = 0; m < R; m++) {
            const float s = warp_sum_d(acc[m]);
            if (lane == 0 && m < B) {
                a.logits[(size_t)m * V + v] = s;
                // whisper.rs:131-138 max_by: the LAST maximum wins; rows ascend, so ">=" keeps the later one
                if (pick && v >= a.arg_lo && v < a.arg_hi && v != sup[m] && (besti[m] < 0 || s >= best[m])) { best[m] = s; besti[m] = v; }
            }
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
            if (i >= 0 && (bi < 0 || v >= bv)) { bv = v; bi = i; }
        }
        a.arg_scratch[((size_t)blockIdx.x * kMaxR + tid) * 2] = bv;
        a.arg_scratch[((size_t)blockIdx.x * kMaxR + tid) * 2 + 1] = __int_as_float(bi);
    }
    grid_barrier(a.barrier, bar_target, a.dev_err);
    if (blockIdx.x == 0 && tid < B) {
        float bv = -INFINITY; int bi = -1;
        for (int c = 0; c < (int)gridDim.x; c++) {
            const float v = __ldcg(a.arg_scratch + ((size_t)c * kMaxR + tid) * 2);
            const int i = __float_as_int(__ldcg(a.arg_scratch + ((size_t)c * kMaxR + tid) * 2 + 1));
            if (i >= 0 && (bi < 0 || v >= bv)) { bv = v; bi = i; }
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

__global__ void cvt_f32_f16_kernel(const float *__restrict__ src, __half *__restrict__ dst, long long n)
{
    for (long long i = (long long)blockIdx.x * blockDim.x + threadIdx.x; i < n; i += (long long)gridDim.x * blockDim.x)
        dst[i] = __ushort_as_half(f32_to_f16_sat_d(src[i]));
}

template <int R>
int32_t launch_decode(const DecodeStepArgs &a, int device, cudaStream_t st)
{
    const size_t smem = ((size_t)R * (a.F > a.D ? a.F : a.D) + (size_t)kDT * R + 1088) * sizeof(float);
    auto kern = decode_step_kernel<R>;
    static bool attr_done[64] = {};
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

int32_t wq4m_cvt_f32_f16(const float *src, void *dst, int64_t n, cudaStream_t st)
{
    cvt_f32_f16_kernel<<<2048, 256, 0, st>>>(src, reinterpret_cast<__half *>(dst), (long long)n);
    wq4_count_launch();
    WQ4_CUDA(cudaGetLastError());
    return WQ4_OK;
}

int32_t wq4m_decode_step(const DecodeStepArgs &a, int device, cudaStream_t st)
{
    if (a.B < 1 || a.B > kMaxR) { wq4_set_error("decode step kernel handles 1..8 clips, got %d", a.B); return WQ4_ERR_INVALID; }
    if (a.D % 64 != 0 || a.D > 1280 || a.D != a.H * 64) { wq4_set_error("decode step kernel: unsupported d_model %d / heads %d", a.D, a.H); return WQ4_ERR_UNSUPPORTED; }
    WqProfScope prof(WQ4_PROF_GEMV, 0.0, st);
    if (a.B == 1) return launch_decode<1>(a, device, st);
    if (a.B == 2) return launch_decode<2>(a, device, st);
    if (a.B <= 4) return launch_decode<4>(a, device, st);
    return launch_decode<8>(a, device, st);
}
