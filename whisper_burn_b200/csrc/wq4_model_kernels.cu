// wq4_model_kernels.cu -- the non-GEMM kernels the encoder / decoder callers of the Q4 path need
// (SURVEY 8f "next" rows): LayerNorm, encoder self-attention, decoder attention over the KV caches,
// embedding, log-mel transpose, argmax.  All emit the 16-bit operand the next GEMM consumes so no
// separate conversion pass exists.
#include <cooperative_groups.h>
#include <float.h>

#include <cstdlib>

#include <atomic>

#include "wq4_common.cuh"
#include "wq4_ptx.cuh"
#include "wq4_model.cuh"

namespace {

__device__ __forceinline__ float warp_sum(float v)
{
#pragma unroll
    for (int o = 16; o > 0; o >>= 1) v += __shfl_xor_sync(0xffffffffu, v, o);
    return v;
}
__device__ __forceinline__ float warp_max(float v)
{
#pragma unroll
    for (int o = 16; o > 0; o >>= 1) v = fmaxf(v, __shfl_xor_sync(0xffffffffu, v, o));
    return v;
}
__device__ __forceinline__ uint16_t f32_to_f16_sat(float v)
{
    uint16_t h;
    asm("cvt.rn.satfinite.f16.f32 %0, %1;" : "=h"(h) : "f"(v));
    return h;
}

// ---------------------------------------------------------------------------------------------
// LayerNorm::forward (src/model/layers.rs:23-31): biased variance, eps 1e-5, two-pass.
// One warp per row; the row lives in registers (NI float4 per lane, D = 128*NI).
// ---------------------------------------------------------------------------------------------
template <int NI>
__global__ void __launch_bounds__(128) layernorm_kernel(const float *__restrict__ x, const float *__restrict__ w,
                                                        const float *__restrict__ b, float *__restrict__ y32,
                                                        __half *__restrict__ y16, int rows)
{
    constexpr int D = NI * 128;
    const int row = blockIdx.x * 4 + (threadIdx.x >> 5);
    const int lane = threadIdx.x & 31;
    ptx::grid_dep_launch();     // PDL (wq4_ptx.cuh): successor may be scheduled; we wait for our own predecessor
    ptx::grid_dep_wait();
    if (row >= rows) return;
    const float4 *xr = reinterpret_cast<const float4 *>(x + (size_t)row * D);
    float4 v[NI];
    float s = 0.0f;
#pragma unroll
    for (int i = 0; i < NI; i++) {
        v[i] = xr[i * 32 + lane];
        s += v[i].x + v[i].y + v[i].z + v[i].w;
    }
    const float mean = warp_sum(s) / (float)D;
    float q = 0.0f;
#pragma unroll
    for (int i = 0; i < NI; i++) {
        v[i].x -= mean; v[i].y -= mean; v[i].z -= mean; v[i].w -= mean;
        q += v[i].x * v[i].x + v[i].y * v[i].y + v[i].z * v[i].z + v[i].w * v[i].w;
    }
    const float var = warp_sum(q) / (float)D;
    const float inv = 1.0f / sqrtf(var + 1e-5f);
#pragma unroll
    for (int i = 0; i < NI; i++) {
        const float4 ww = reinterpret_cast<const float4 *>(w)[i * 32 + lane];
        const float4 bb = reinterpret_cast<const float4 *>(b)[i * 32 + lane];
        float4 o;
        o.x = v[i].x * inv * ww.x + bb.x;
        o.y = v[i].y * inv * ww.y + bb.y;
        o.z = v[i].z * inv * ww.z + bb.z;
        o.w = v[i].w * inv * ww.w + bb.w;
        if (y32) reinterpret_cast<float4 *>(y32 + (size_t)row * D)[i * 32 + lane] = o;
        if (y16) {
            uint2 pk;
            pk.x = (uint32_t)f32_to_f16_sat(o.x) | ((uint32_t)f32_to_f16_sat(o.y) << 16);
            pk.y = (uint32_t)f32_to_f16_sat(o.z) | ((uint32_t)f32_to_f16_sat(o.w) << 16);
            reinterpret_cast<uint2 *>(y16 + (size_t)row * D)[i * 32 + lane] = pk;
        }
    }
}

// Generic D (any multiple of 4): one warp per row, re-reads the row from L1/L2.
__global__ void __launch_bounds__(128) layernorm_generic_kernel(const float *__restrict__ x, const float *__restrict__ w,
                                                                const float *__restrict__ b, float *__restrict__ y32,
                                                                __half *__restrict__ y16, int rows, int D)
{
    const int row = blockIdx.x * 4 + (threadIdx.x >> 5);
    const int lane = threadIdx.x & 31;
    if (row >= rows) return;
    const float *xr = x + (size_t)row * D;
    float s = 0.0f;
    for (int i = lane; i < D; i += 32) s += xr[i];
    const float mean = warp_sum(s) / (float)D;
    float q = 0.0f;
    for (int i = lane; i < D; i += 32) { float d = xr[i] - mean; q += d * d; }
    const float inv = 1.0f / sqrtf(warp_sum(q) / (float)D + 1e-5f);
    for (int i = lane; i < D; i += 32) {
        float o = (xr[i] - mean) * inv * w[i] + b[i];
        if (y32) y32[(size_t)row * D + i] = o;
        if (y16) y16[(size_t)row * D + i] = __ushort_as_half(f32_to_f16_sat(o));
    }
}

// ---------------------------------------------------------------------------------------------
// Encoder self-attention (scaled_dot_product_attention, src/model/attention.rs:243-298, non-causal):
// softmax(Q K^T / sqrt(64)) V per (batch, head), flash-style so the [H,T,T] score tensor the
// reference materialises (:267) never exists.  head_dim is 64 as hard-coded in the reference (:30).
// qkv: [B*T][3*D] f16 (q | k | v), out: [B*T][D] f16.  64 queries x 64 keys per step, 4 warps,
// mma.sync m16n8k16 (legacy tensor path; the tcgen05 version of this kernel is future work and
// noted as such in DESIGN.md -- SDPA is outside the Q4 GEMM/GEMV coverage rows).
// ---------------------------------------------------------------------------------------------
constexpr int kHD = 64;
constexpr int kBQ = 64;
constexpr int kBKV = 64;

__device__ __forceinline__ void cp_async_16(uint32_t dst, const void *src, bool valid)
{
    const int sz = valid ? 16 : 0;
    asm volatile("cp.async.cg.shared.global [%0], [%1], 16, %2;" ::"r"(dst), "l"(src), "r"(sz) : "memory");
}
__device__ __forceinline__ void cp_async_commit() { asm volatile("cp.async.commit_group;" ::: "memory"); }
template <int N>
__device__ __forceinline__ void cp_async_wait() { asm volatile("cp.async.wait_group %0;" ::"n"(N) : "memory"); }

__device__ __forceinline__ void ldsm_x4(uint32_t addr, uint32_t &r0, uint32_t &r1, uint32_t &r2, uint32_t &r3)
{
    asm volatile("ldmatrix.sync.aligned.m8n8.x4.shared.b16 {%0,%1,%2,%3}, [%4];"
                 : "=r"(r0), "=r"(r1), "=r"(r2), "=r"(r3) : "r"(addr));
}
__device__ __forceinline__ void ldsm_x4_t(uint32_t addr, uint32_t &r0, uint32_t &r1, uint32_t &r2, uint32_t &r3)
{
    asm volatile("ldmatrix.sync.aligned.m8n8.x4.trans.shared.b16 {%0,%1,%2,%3}, [%4];"
                 : "=r"(r0), "=r"(r1), "=r"(r2), "=r"(r3) : "r"(addr));
}
__device__ __forceinline__ void mma_16816(float (&c)[4], uint32_t a0, uint32_t a1, uint32_t a2, uint32_t a3,
                                          uint32_t b0, uint32_t b1)
{
    asm volatile("mma.sync.aligned.m16n8k16.row.col.f32.f16.f16.f32 {%0,%1,%2,%3}, {%4,%5,%6,%7}, {%8,%9}, {%0,%1,%2,%3};"
                 : "+f"(c[0]), "+f"(c[1]), "+f"(c[2]), "+f"(c[3])
                 : "r"(a0), "r"(a1), "r"(a2), "r"(a3), "r"(b0), "r"(b1));
}
__device__ __forceinline__ uint32_t pack_half2(float a, float b)
{
    __half2 h = __floats2half2_rn(a, b);
    return *reinterpret_cast<uint32_t *>(&h);
}

// tile [64 rows][64 f16] in shared memory, 16-byte chunk c of row r stored at chunk c ^ (r & 7)
__device__ __forceinline__ uint32_t tile_addr(uint32_t base, int r, int c) { return base + r * 128 + ((c ^ (r & 7)) << 4); }

__device__ __forceinline__ void load_tile_async(uint32_t smem, const __half *g, int ld, int row0, int nrows_valid,
                                                int tid)
{
    // 64 rows x 8 chunks = 512 chunks, 128 threads x 4
#pragma unroll
    for (int i = 0; i < 4; i++) {
        int idx = tid + i * 128;
        int r = idx >> 3, c = idx & 7;
        bool ok = (row0 + r) < nrows_valid;
        const __half *src = g + (size_t)(ok ? (row0 + r) : 0) * ld + c * 8;
        cp_async_16(tile_addr(smem, r, c), src, ok);
    }
}

__global__ void __launch_bounds__(128) enc_attention_kernel(const __half *__restrict__ qkv, __half *__restrict__ out,
                                                            int T, int D, float scale_log2)
{
    __shared__ __align__(128) uint8_t smem[5 * kBQ * 128];   // Q, K0, K1, V0, V1 : 40 KiB
    const uint32_t sQ = (uint32_t)__cvta_generic_to_shared(smem);
    const uint32_t sK[2] = {sQ + 8192, sQ + 16384};
    const uint32_t sV[2] = {sQ + 24576, sQ + 32768};
    const int tid = threadIdx.x, warp = tid >> 5, lane = tid & 31;
    const int q0 = blockIdx.x * kBQ, h = blockIdx.y, b = blockIdx.z;
    const int ld = 3 * D;
    const __half *base = qkv + (size_t)b * T * ld + h * kHD;
    const __half *gQ = base, *gK = base + D, *gV = base + 2 * D;

    load_tile_async(sQ, gQ, ld, q0, T, tid);
    load_tile_async(sK[0], gK, ld, 0, T, tid);
    load_tile_async(sV[0], gV, ld, 0, T, tid);
    cp_async_commit();

    const int nkv = (T + kBKV - 1) / kBKV;
    float o[8][4];
#pragma unroll
    for (int j = 0; j < 8; j++) { o[j][0] = o[j][1] = o[j][2] = o[j][3] = 0.0f; }
    float m_run[2] = {-INFINITY, -INFINITY}, l_run[2] = {0.0f, 0.0f};
    uint32_t qa[4][4];

    for (int it = 0; it < nkv; it++) {
        const int buf = it & 1;
        if (it + 1 < nkv) {
            load_tile_async(sK[buf ^ 1], gK, ld, (it + 1) * kBKV, T, tid);
            load_tile_async(sV[buf ^ 1], gV, ld, (it + 1) * kBKV, T, tid);
            cp_async_commit();
            cp_async_wait<1>();
        } else {
            cp_async_wait<0>();
        }
        __syncthreads();
        if (it == 0) {
            // Q fragments: rows warp*16 + ..., 4 k-steps of 16
#pragma unroll
            for (int kk = 0; kk < 4; kk++) {
                int r = warp * 16 + (lane & 7) + ((lane >> 3) & 1) * 8;
                int c = kk * 2 + (lane >> 4);
                ldsm_x4(tile_addr(sQ, r, c), qa[kk][0], qa[kk][1], qa[kk][2], qa[kk][3]);
            }
        }
        // S = Q K^T  (16 x 64 per warp)
        float s[8][4];
#pragma unroll
        for (int j = 0; j < 8; j++) { s[j][0] = s[j][1] = s[j][2] = s[j][3] = 0.0f; }
#pragma unroll
        for (int kk = 0; kk < 4; kk++) {
#pragma unroll
            for (int jp = 0; jp < 4; jp++) {   // pairs of 8-key tiles
                uint32_t b0, b1, b2, b3;
                int mi = lane >> 3;
                int r = jp * 16 + (mi >> 1) * 8 + (lane & 7);
                int c = kk * 2 + (mi & 1);
                ldsm_x4(tile_addr(sK[buf], r, c), b0, b1, b2, b3);
                mma_16816(s[2 * jp], qa[kk][0], qa[kk][1], qa[kk][2], qa[kk][3], b0, b1);
                mma_16816(s[2 * jp + 1], qa[kk][0], qa[kk][1], qa[kk][2], qa[kk][3], b2, b3);
            }
        }
        // mask keys >= T (last tile only), online softmax in the exp2 domain
        const int kbase = it * kBKV;
        if (kbase + kBKV > T) {
#pragma unroll
            for (int j = 0; j < 8; j++) {
                int key = kbase + j * 8 + (lane & 3) * 2;
                if (key >= T) { s[j][0] = -INFINITY; s[j][2] = -INFINITY; }
                if (key + 1 >= T) { s[j][1] = -INFINITY; s[j][3] = -INFINITY; }
            }
        }
        float mx[2] = {-INFINITY, -INFINITY};
#pragma unroll
        for (int j = 0; j < 8; j++) {
            mx[0] = fmaxf(mx[0], fmaxf(s[j][0], s[j][1]));
            mx[1] = fmaxf(mx[1], fmaxf(s[j][2], s[j][3]));
        }
#pragma unroll
        for (int i = 0; i < 2; i++) {
            mx[i] = fmaxf(mx[i], __shfl_xor_sync(0xffffffffu, mx[i], 1));
            mx[i] = fmaxf(mx[i], __shfl_xor_sync(0xffffffffu, mx[i], 2));
        }
        float alpha[2], msc[2];
#pragma unroll
        for (int i = 0; i < 2; i++) {
            float m_new = fmaxf(m_run[i], mx[i]);
            alpha[i] = exp2f((m_run[i] - m_new) * scale_log2);   // m_run = -inf on the first tile -> 0
            m_run[i] = m_new;
            msc[i] = m_new * scale_log2;
        }
        float rs[2] = {0.0f, 0.0f};
        uint32_t pa[4][4];
#pragma unroll
        for (int j = 0; j < 8; j++) {
            float p0 = exp2f(s[j][0] * scale_log2 - msc[0]);
            float p1 = exp2f(s[j][1] * scale_log2 - msc[0]);
            float p2 = exp2f(s[j][2] * scale_log2 - msc[1]);
            float p3 = exp2f(s[j][3] * scale_log2 - msc[1]);
            rs[0] += p0 + p1;
            rs[1] += p2 + p3;
            pa[j >> 1][(j & 1) * 2 + 0] = pack_half2(p0, p1);
            pa[j >> 1][(j & 1) * 2 + 1] = pack_half2(p2, p3);
        }
#pragma unroll
        for (int i = 0; i < 2; i++) l_run[i] = l_run[i] * alpha[i] + rs[i];
#pragma unroll
        for (int j = 0; j < 8; j++) {
            o[j][0] *= alpha[0]; o[j][1] *= alpha[0];
            o[j][2] *= alpha[1]; o[j][3] *= alpha[1];
        }
        // O += P V   (k = keys, n = d)
#pragma unroll
        for (int kk = 0; kk < 4; kk++) {
#pragma unroll
            for (int jp = 0; jp < 4; jp++) {   // pairs of 8-wide d tiles
                uint32_t b0, b1, b2, b3;
                int mi = lane >> 3;
                int r = kk * 16 + (mi & 1) * 8 + (lane & 7);
                int c = jp * 2 + (mi >> 1);
                ldsm_x4_t(tile_addr(sV[buf], r, c), b0, b1, b2, b3);
                mma_16816(o[2 * jp], pa[kk][0], pa[kk][1], pa[kk][2], pa[kk][3], b0, b1);
                mma_16816(o[2 * jp + 1], pa[kk][0], pa[kk][1], pa[kk][2], pa[kk][3], b2, b3);
            }
        }
        __syncthreads();   // everyone is done with buf before it is refilled
    }
    // finalise: divide by the row sum (quad reduce), stage through sQ for 16-byte stores
#pragma unroll
    for (int i = 0; i < 2; i++) {
        l_run[i] += __shfl_xor_sync(0xffffffffu, l_run[i], 1);
        l_run[i] += __shfl_xor_sync(0xffffffffu, l_run[i], 2);
    }
    const float inv0 = 1.0f / l_run[0], inv1 = 1.0f / l_run[1];
    uint8_t *sO = smem;   // Q tile is dead (fragments are in registers)
    {
        const int g = lane >> 2, t = lane & 3;
        const int r0 = warp * 16 + g, r1 = r0 + 8;
#pragma unroll
        for (int j = 0; j < 8; j++) {
            // column 8j + 2t -> chunk j, byte offset 4t inside the chunk
            *reinterpret_cast<uint32_t *>(sO + r0 * 128 + ((j ^ (r0 & 7)) << 4) + t * 4) = pack_half2(o[j][0] * inv0, o[j][1] * inv0);
            *reinterpret_cast<uint32_t *>(sO + r1 * 128 + ((j ^ (r1 & 7)) << 4) + t * 4) = pack_half2(o[j][2] * inv1, o[j][3] * inv1);
        }
    }
    __syncthreads();
    __half *gO = out + (size_t)b * T * D + h * kHD;
#pragma unroll
    for (int i = 0; i < 4; i++) {
        int idx = tid + i * 128;
        int r = idx >> 3, c = idx & 7;
        if (q0 + r < T) {
            uint4 v = *reinterpret_cast<const uint4 *>(sO + r * 128 + ((c ^ (r & 7)) << 4));
            *reinterpret_cast<uint4 *>(gO + (size_t)(q0 + r) * D + c * 8) = v;
        }
    }
}

// ---------------------------------------------------------------------------------------------
// Decoder attention over a KV cache (src/model/attention.rs:93-125 self, :208-236 cross, and the
// causal prompt pass :62-85): one block per (query row, head, batch item).  q is f32 [.., D],
// K/V are f16 caches with row stride ldkv.  kv_len keys are visible to query row i when
// i_limit = causal ? (kv_base + i + 1) : kv_len.   Scores: q.k / sqrt(64) (:266-267), softmax, .V.
// ---------------------------------------------------------------------------------------------
__global__ void __launch_bounds__(256) dec_attention_kernel(const float *__restrict__ q, int ldq, const __half *__restrict__ kc,
                                                            const __half *__restrict__ vc, int ldkv, long long kv_batch_stride,
                                                            long long hstride, float *__restrict__ out, int ldo, int q_len, int kv_len,
                                                            int causal_base /* <0: not causal */)
{
    extern __shared__ float sm_scores[];   // [kv_len] + 64 q + 256*... partials
    const int qi = blockIdx.x, h = blockIdx.y, b = blockIdx.z;
    const int tid = threadIdx.x;
    float *sq = sm_scores + ((kv_len + 3) & ~3);
    float *sred = sq + kHD;                       // [4][64] partial outputs + 8 scratch
    const int visible = causal_base >= 0 ? min(kv_len, causal_base + qi + 1) : kv_len;
    const float *qrow = q + ((size_t)b * q_len + qi) * ldq + h * kHD;
    if (tid < kHD) sq[tid] = qrow[tid];
    __syncthreads();
    const __half *kb = kc + (size_t)b * kv_batch_stride + (size_t)h * hstride;      // hstride: 64 inside a [ctx][D] row, T*128 head-major
    const __half *vb = vc + (size_t)b * kv_batch_stride + (size_t)h * hstride;
    // scores: one thread per key (128-byte contiguous row read)
    float lmax = -INFINITY;
    for (int key = tid; key < visible; key += 256) {
        const uint4 *kr = reinterpret_cast<const uint4 *>(kb + (size_t)key * ldkv);
        float acc = 0.0f;
#pragma unroll
        for (int c = 0; c < 8; c++) {
            uint4 v = kr[c];
            const __half2 *hp = reinterpret_cast<const __half2 *>(&v);
#pragma unroll
            for (int e = 0; e < 4; e++) {
                float2 f = __half22float2(hp[e]);
                acc = fmaf(sq[c * 8 + 2 * e], f.x, acc);
                acc = fmaf(sq[c * 8 + 2 * e + 1], f.y, acc);
            }
        }
        acc = acc / 8.0f;   // / sqrt(head_dim), attention.rs:266-267
        sm_scores[key] = acc;
        lmax = fmaxf(lmax, acc);
    }
    __shared__ float s_wred[8];
    __shared__ float s_bcast[2];
    lmax = warp_max(lmax);
    if ((tid & 31) == 0) s_wred[tid >> 5] = lmax;
    __syncthreads();
    if (tid == 0) {
        float m = s_wred[0];
        for (int i = 1; i < 8; i++) m = fmaxf(m, s_wred[i]);
        s_bcast[0] = m;
    }
    __syncthreads();
    const float gmax = s_bcast[0];
    float lsum = 0.0f;
    for (int key = tid; key < visible; key += 256) {
        float e = __expf(sm_scores[key] - gmax);
        sm_scores[key] = e;
        lsum += e;
    }
    lsum = warp_sum(lsum);
    __syncthreads();
    if ((tid & 31) == 0) s_wred[tid >> 5] = lsum;
    __syncthreads();
    if (tid == 0) {
        float s = 0.0f;
        for (int i = 0; i < 8; i++) s += s_wred[i];
        s_bcast[1] = s;
    }
    __syncthreads();
    const float inv = 1.0f / s_bcast[1];
    // out[d] = sum_key p[key] * V[key][d]: 4 key groups x 64 d
    const int d = tid & 63, grp = tid >> 6;
    float acc = 0.0f;
    for (int key = grp; key < visible; key += 4) acc = fmaf(sm_scores[key], __half2float(vb[(size_t)key * ldkv + d]), acc);
    sred[grp * 64 + d] = acc;
    __syncthreads();
    if (tid < 64) {
        float s = sred[tid] + sred[64 + tid] + sred[128 + tid] + sred[192 + tid];
        out[((size_t)b * q_len + qi) * ldo + h * kHD + tid] = s * inv;
    }
}

// x [B][C][T] f32 -> xt [B][Tp][Cp] f16 with `pad` zero rows at both ends (Tp = T + 2*pad rounded up
// by the caller) -- the time-major layout the implicit-im2col GEMM view needs.
__global__ void mel_transpose_kernel(const float *__restrict__ x, __half *__restrict__ xt, int C, int T, int Cp, int Tp,
                                     int pad)
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
        if (tp < Tp && c < Cp) xt[((size_t)b * Tp + tp) * Cp + c] = __ushort_as_half(f32_to_f16_sat(tile[threadIdx.x][i]));
    }
}

// x[b][i][:] = token_embedding[tok[b][i]][:] + positional_embedding[pos0 + i][:]  (decoder.rs:262-277,317-329)
__global__ void embed_kernel(const int *__restrict__ tokens, const float *__restrict__ tok_emb, const float *__restrict__ pos_emb,
                             float *__restrict__ x, int D, int q_len, int pos0)
{
    const int i = blockIdx.x, b = blockIdx.y;
    const int tok = tokens[b * q_len + i];
    const float4 *te = reinterpret_cast<const float4 *>(tok_emb + (size_t)tok * D);
    const float4 *pe = reinterpret_cast<const float4 *>(pos_emb + (size_t)(pos0 + i) * D);
    float4 *xo = reinterpret_cast<float4 *>(x + ((size_t)b * q_len + i) * D);
    for (int j = threadIdx.x; j < D / 4; j += blockDim.x) {
        float4 a = te[j], p = pe[j];
        xo[j] = make_float4(a.x + p.x, a.y + p.y, a.z + p.z, a.w + p.w);
    }
}

// logits[b][v] = x[b][:] . token_embedding[v][:]  (decoder.rs:292,343 -- an f32 matmul in the
// reference, kept f32 here).  One warp per vocabulary row; x rows staged in shared memory.
template <int MB>
__global__ void __launch_bounds__(256) logits_kernel(const float *__restrict__ x, const float *__restrict__ emb,
                                                     float *__restrict__ logits, int D, int V, int nb)
{
    extern __shared__ float sx[];   // [MB][D]
    for (int i = threadIdx.x; i < MB * D; i += 256) {
        int m = i / D;
        sx[i] = m < nb ? x[i] : 0.0f;
    }
    __syncthreads();
    const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
    const int v = blockIdx.x * 8 + warp;
    if (v >= V) return;
    const float4 *er = reinterpret_cast<const float4 *>(emb + (size_t)v * D);
    float acc[MB];
#pragma unroll
    for (int m = 0; m < MB; m++) acc[m] = 0.0f;
    for (int j = lane; j < D / 4; j += 32) {
        float4 e = __ldg(er + j);
#pragma unroll
        for (int m = 0; m < MB; m++) {
            const float4 xv = reinterpret_cast<const float4 *>(sx + m * D)[j];
            acc[m] = fmaf(e.x, xv.x, acc[m]);
            acc[m] = fmaf(e.y, xv.y, acc[m]);
            acc[m] = fmaf(e.z, xv.z, acc[m]);
            acc[m] = fmaf(e.w, xv.w, acc[m]);
        }
    }
#pragma unroll
    for (int m = 0; m < MB; m++) {
        float s = warp_sum(acc[m]);
        if (lane == 0 && m < nb) logits[(size_t)m * V + v] = s;
    }
}

// Greedy pick (whisper.rs:98-99,120-124,131-138): optional EOT suppression, optional restriction to a
// token range (language detection :77-83), argmax with Rust `max_by` tie-breaking (LAST maximum wins).
__global__ void __launch_bounds__(1024) argmax_kernel(const float *__restrict__ logits, int V, int lo, int hi,
                                                      const int *__restrict__ suppress /* per batch, -1 = none */,
                                                      int *__restrict__ out)
{
    const int b = blockIdx.x;
    const float *lg = logits + (size_t)b * V;
    const int sup = suppress ? suppress[b] : -1;
    float best = -INFINITY;
    int bi = -1;
    for (int i = lo + threadIdx.x; i < hi; i += blockDim.x) {
        float v = (i == sup) ? -INFINITY : lg[i];
        if (bi < 0 || v >= best) { best = v; bi = i; }   // ascending i, ">=": the last maximum wins
    }
    const bool any = bi >= 0;
    __shared__ float sv[1024];
    __shared__ int si[1024];
    sv[threadIdx.x] = any ? best : -INFINITY;
    si[threadIdx.x] = any ? bi : -1;
    __syncthreads();
    for (int s = blockDim.x / 2; s > 0; s >>= 1) {
        if ((int)threadIdx.x < s) {
            float ov = sv[threadIdx.x + s];
            int oi = si[threadIdx.x + s];
            float mv = sv[threadIdx.x];
            int mi = si[threadIdx.x];
            if (oi >= 0 && (mi < 0 || ov > mv || (ov == mv && oi > mi))) { sv[threadIdx.x] = ov; si[threadIdx.x] = oi; }
        }
        __syncthreads();
    }
    if (threadIdx.x == 0) out[b] = si[0];
}

// The same pick (whisper.rs:131-138: the LAST maximum wins, `suppress` excluded) by a cluster of kArgCluster CTAs per row:
// every CTA scans a contiguous eighth of [lo, hi), rank 0 merges the eight (value, index) partials through distributed
// shared memory in rank order (ascending index ranges, ">=": the later one wins ties).
constexpr int kArgCluster = 8;
__global__ void __launch_bounds__(256) argmax_cluster_kernel(const float *__restrict__ logits, int V, int lo, int hi,
                                                             const int *__restrict__ suppress, int *__restrict__ out)
{
    namespace cg = cooperative_groups;
    cg::cluster_group cluster = cg::this_cluster();
    const int b = blockIdx.y, rank = (int)cluster.block_rank();
    const float *lg = logits + (size_t)b * V;
    const int sup = suppress ? suppress[b] : -1;
    const int per = (hi - lo + kArgCluster - 1) / kArgCluster;
    const int c0 = lo + rank * per, c1 = min(hi, c0 + per);
    float best = -INFINITY;
    int bi = -1;
    for (int i = c0 + threadIdx.x; i < c1; i += 256) {
        const float v = (i == sup) ? -INFINITY : lg[i];
        if (bi < 0 || v >= best) { best = v; bi = i; }
    }
    // warp: keep the larger value, the larger index on ties
#pragma unroll
    for (int o = 16; o > 0; o >>= 1) {
        const float ov = __shfl_xor_sync(0xffffffffu, best, o);
        const int oi = __shfl_xor_sync(0xffffffffu, bi, o);
        if (oi >= 0 && (bi < 0 || ov > best || (ov == best && oi > bi))) { best = ov; bi = oi; }
    }
    __shared__ float sv[8];
    __shared__ int si[8];
    __shared__ float cv;
    __shared__ int ci;
    if ((threadIdx.x & 31) == 0) { sv[threadIdx.x >> 5] = best; si[threadIdx.x >> 5] = bi; }
    __syncthreads();
    if (threadIdx.x == 0) {
        float mv = -INFINITY;
        int mi = -1;
        for (int w = 0; w < 8; w++)
            if (si[w] >= 0 && (mi < 0 || sv[w] > mv || (sv[w] == mv && si[w] > mi))) { mv = sv[w]; mi = si[w]; }
        cv = mv; ci = mi;
    }
    cluster.sync();
    if (rank == 0 && threadIdx.x == 0) {
        float mv = -INFINITY;
        int mi = -1;
        for (int r = 0; r < kArgCluster; r++) {
            const float rv = *cluster.map_shared_rank(&cv, r);
            const int ri = *cluster.map_shared_rank(&ci, r);
            if (ri >= 0 && (mi < 0 || rv > mv || (rv == mv && ri > mi))) { mv = rv; mi = ri; }
        }
        out[b] = mi;
    }
    cluster.sync();      // keep every rank's shared memory alive until rank 0 has read it
}

// f32 [rows][D] -> f16 K/V cache rows (self-attention cache append, attention.rs:106-113 without the realloc)
__global__ void kv_append_kernel(const float *__restrict__ kv, int ld_src, __half *__restrict__ kc, __half *__restrict__ vc,
                                 int ldkv, long long kv_batch_stride, int D, int q_len, int pos0)
{
    const int i = blockIdx.x, b = blockIdx.y;
    const float *src = kv + ((size_t)b * q_len + i) * ld_src;   // [q | k | v] fused projection output
    __half *kd = kc + (size_t)b * kv_batch_stride + (size_t)(pos0 + i) * ldkv;
    __half *vd = vc + (size_t)b * kv_batch_stride + (size_t)(pos0 + i) * ldkv;
    for (int j = threadIdx.x; j < D; j += blockDim.x) {
        kd[j] = __ushort_as_half(f32_to_f16_sat(src[D + j]));
        vd[j] = __ushort_as_half(f32_to_f16_sat(src[2 * D + j]));
    }
}

// Cross-attention K/V of one decoder layer, [B*T][2D] f16 as the fused key|value projection writes it (attention.rs:185-186)
// -> head-major [B][H][T][128] (64 K | 64 V per key): a (clip, head) unit of the decode-step attention then reads ONE
// contiguous 384 KB stream instead of 128-byte chunks 5 KB apart (HBM ran at ~4.5 TB/s on the strided form).  One thread moves
// one 16-byte chunk; reads are fully coalesced rows, writes are whole 128-byte lines.
__global__ void __launch_bounds__(256) cross_kv_pack_kernel(const uint4 *__restrict__ src, uint4 *__restrict__ dst, int T, int H,
                                                            long long total_chunks)
{
    const long long i = (long long)blockIdx.x * 256 + threadIdx.x;      // chunk index in the source: [b*T + t][2][H][8]
    if (i >= total_chunks) return;
    const int c = (int)(i & 7);
    long long r = i >> 3;
    const int h = (int)(r % H); r /= H;
    const int kv = (int)(r & 1); r >>= 1;                               // 0 = key half of the row, 1 = value half
    const int t = (int)(r % T);
    const long long b = r / T;
    dst[(((b * H + h) * T + t) * 2 + kv) * 8 + c] = src[i];
}


// Greedy-loop bookkeeping of a batch of clips on the device (WhisperModel::transcribe, whisper.rs:104-125, per clip):
// consumes the arg-max of the rows still active, appends the token to its clip's output (or retires the clip on EOT /
// when its token budget is used up), COMPACTS the active set -- surviving rows keep their order and move to the front,
// the retired clips' cache slots fill the tail -- and writes the tokens the next step embeds.  One CTA; the host never
// has to look at the tokens between steps (it reads *n_active every few steps to shrink its launches).
// Rows >= *n_active of a launch that still uses an older, larger row count point at retired clips' slots (their caches
// are dead storage), so whatever they compute is harmless.
__global__ void __launch_bounds__(1024) greedy_bookkeep_kernel(const int *__restrict__ next, int *__restrict__ tokens_next,
                                                               int *__restrict__ slot, int *__restrict__ n_active,
                                                               int *__restrict__ count, const int *__restrict__ budget,
                                                               int *__restrict__ tokens_out, int max_tokens, int eot)
{
    __shared__ int warp_keep[32];
    __shared__ int s_total;
    const int r = threadIdx.x, lane = r & 31, warp = r >> 5;
    const int n = *n_active;
    int tok = 0, c = 0;
    bool keep = false;
    const bool mine = r < n;
    if (mine) {
        tok = next[r];
        c = slot[r];
        if (tok != eot) {
            const int k = count[c];
            if (k < budget[c]) {
                tokens_out[(size_t)c * max_tokens + k] = tok;
                count[c] = k + 1;
                keep = k + 1 < budget[c];
            }
        }
    }
    const unsigned bal = __ballot_sync(0xffffffffu, keep);
    if (lane == 0) warp_keep[warp] = __popc(bal);
    __syncthreads();                                   // also: every read of next / slot above is done
    if (warp == 0) {
        int v = warp_keep[lane];
        int incl = v;
#pragma unroll
        for (int o = 1; o < 32; o <<= 1) { const int t = __shfl_up_sync(0xffffffffu, incl, o); if (lane >= o) incl += t; }
        warp_keep[lane] = incl - v;                    // exclusive prefix over warps
        if (lane == 31) s_total = incl;
    }
    __syncthreads();
    const int n_new = s_total;
    if (mine) {
        const int before = warp_keep[warp] + __popc(bal & ((1u << lane) - 1u));     // kept rows in front of r
        const int dst = keep ? before : n_new + (r - before);                       // retired rows fill the tail, in order
        slot[dst] = c;
        tokens_next[dst] = keep ? tok : 0;
    }
    if (r == 0) *n_active = n_new;
}

__global__ void greedy_init_kernel(int *slot, int *n_active, int *count, int B)
{
    const int r = blockIdx.x * blockDim.x + threadIdx.x;
    if (r < B) { slot[r] = r; count[r] = 0; }
    if (r == 0) *n_active = B;
}

}  // namespace

// ---- host launchers ------------------------------------------------------------------------------
int32_t wq4m_cross_kv_pack(const void *src16, void *dst16, int64_t B, int64_t T, int64_t H, cudaStream_t st)
{
    const long long chunks = (long long)B * T * 2 * H * 8;
    cross_kv_pack_kernel<<<(unsigned)((chunks + 255) / 256), 256, 0, st>>>(reinterpret_cast<const uint4 *>(src16),
                                                                          reinterpret_cast<uint4 *>(dst16), (int)T, (int)H, chunks);
    wq4_count_launch();
    WQ4_CUDA(cudaGetLastError());
    return WQ4_OK;
}

int32_t wq4m_layernorm(const float *x, const float *w, const float *b, float *y32, void *y16, int64_t rows, int64_t D,
                       cudaStream_t st)
{
    if (rows <= 0) return WQ4_OK;
    dim3 grid((unsigned)((rows + 3) / 4)), block(128);
    __half *h = reinterpret_cast<__half *>(y16);
    WqProfScope prof(WQ4_PROF_LAYERNORM, (double)rows * (double)D * (4.0 + (y32 ? 4.0 : 0.0) + (y16 ? 2.0 : 0.0)), st);
    cudaLaunchConfig_t cfg = {};
    cfg.gridDim = grid;
    cfg.blockDim = block;
    cfg.stream = st;
    cudaLaunchAttribute attr[1];
    attr[0].id = cudaLaunchAttributeProgrammaticStreamSerialization;
    attr[0].val.programmaticStreamSerializationAllowed = 1;
    cfg.attrs = attr;
    cfg.numAttrs = wq4_pdl_enabled() ? 1 : 0;
    const int irows = (int)rows;
    switch (D) {
    case 256: WQ4_CUDA(cudaLaunchKernelEx(&cfg, layernorm_kernel<2>, x, w, b, y32, h, irows)); break;
    case 384: WQ4_CUDA(cudaLaunchKernelEx(&cfg, layernorm_kernel<3>, x, w, b, y32, h, irows)); break;
    case 512: WQ4_CUDA(cudaLaunchKernelEx(&cfg, layernorm_kernel<4>, x, w, b, y32, h, irows)); break;
    case 768: WQ4_CUDA(cudaLaunchKernelEx(&cfg, layernorm_kernel<6>, x, w, b, y32, h, irows)); break;
    case 1024: WQ4_CUDA(cudaLaunchKernelEx(&cfg, layernorm_kernel<8>, x, w, b, y32, h, irows)); break;
    case 1280: WQ4_CUDA(cudaLaunchKernelEx(&cfg, layernorm_kernel<10>, x, w, b, y32, h, irows)); break;
    default: layernorm_generic_kernel<<<grid, block, 0, st>>>(x, w, b, y32, h, (int)rows, (int)D); break;
    }
    wq4_count_launch();
    WQ4_CUDA(cudaGetLastError());
    return WQ4_OK;
}

int32_t wq4m_enc_attention(const void *qkv16, void *out16, int64_t B, int64_t T, int64_t D, int64_t H, cudaStream_t st)
{
    if (D != H * kHD) { wq4_set_error("attention: head_dim must be 64 (attention.rs:30), got D=%lld H=%lld", (long long)D, (long long)H); return WQ4_ERR_SHAPE; }
    static int impl = -1;
    if (impl < 0) { const char *e = getenv("WQ4_ATTN_IMPL"); impl = (e && atoi(e) == 1) ? 1 : 0; }   // 1 = mma.sync kernel below
    if (impl == 0) return wq4m_enc_attention_tc(qkv16, out16, B, T, D, H, st);
    dim3 grid((unsigned)((T + kBQ - 1) / kBQ), (unsigned)H, (unsigned)B);
    const float scale_log2 = 1.4426950408889634f / 8.0f;   // log2(e) / sqrt(64)
    WqProfScope prof(WQ4_PROF_ATTN_ENC, 4.0 * (double)B * (double)H * (double)T * (double)T * 64.0, st);
    enc_attention_kernel<<<grid, 128, 0, st>>>(reinterpret_cast<const __half *>(qkv16), reinterpret_cast<__half *>(out16),
                                               (int)T, (int)D, scale_log2);
    wq4_count_launch();
    WQ4_CUDA(cudaGetLastError());
    return WQ4_OK;
}

int32_t wq4m_dec_attention(const float *q, int64_t ldq, const void *kc, const void *vc, int64_t ldkv, int64_t kv_batch_stride,
                           int64_t hstride, float *out, int64_t ldo, int64_t B, int64_t H, int64_t q_len, int64_t kv_len,
                           int causal_base, cudaStream_t st)
{
    dim3 grid((unsigned)q_len, (unsigned)H, (unsigned)B);
    size_t smem = (size_t)(((kv_len + 3) & ~3) + kHD + 4 * 64 + 8) * sizeof(float);
    WqProfScope prof(WQ4_PROF_ATTN_DEC, (double)B * (double)H * (double)kv_len * 64.0 * 2.0 * 2.0, st);
    dec_attention_kernel<<<grid, 256, smem, st>>>(q, (int)ldq, reinterpret_cast<const __half *>(kc),
                                                  reinterpret_cast<const __half *>(vc), (int)ldkv, (long long)kv_batch_stride,
                                                  (long long)hstride, out, (int)ldo, (int)q_len, (int)kv_len, causal_base);
    wq4_count_launch();
    WQ4_CUDA(cudaGetLastError());
    return WQ4_OK;
}

int32_t wq4m_mel_transpose(const float *x, void *xt16, int64_t B, int64_t C, int64_t T, int64_t Cp, int64_t Tp, int64_t pad,
                           cudaStream_t st)
{
    dim3 grid((unsigned)((Tp + 31) / 32), (unsigned)((Cp + 31) / 32), (unsigned)B), block(32, 8);
    mel_transpose_kernel<<<grid, block, 0, st>>>(x, reinterpret_cast<__half *>(xt16), (int)C, (int)T, (int)Cp, (int)Tp, (int)pad);
    wq4_count_launch();
    WQ4_CUDA(cudaGetLastError());
    return WQ4_OK;
}

int32_t wq4m_embed(const int *tokens, const float *tok_emb, const float *pos_emb, float *x, int64_t B, int64_t q_len, int64_t D,
                   int64_t pos0, cudaStream_t st)
{
    dim3 grid((unsigned)q_len, (unsigned)B);
    embed_kernel<<<grid, 256, 0, st>>>(tokens, tok_emb, pos_emb, x, (int)D, (int)q_len, (int)pos0);
    wq4_count_launch();
    WQ4_CUDA(cudaGetLastError());
    return WQ4_OK;
}

int32_t wq4m_logits(const float *x, const float *emb, float *logits, int64_t nb, int64_t D, int64_t V, cudaStream_t st)
{
    WqProfScope prof(WQ4_PROF_LOGITS, (double)((nb + 7) / 8) * (double)V * (double)D * 4.0, st);
    // batches of up to 8 rows share one pass over the embedding matrix
    for (int64_t b0 = 0; b0 < nb; b0 += 8) {
        int64_t cnt = nb - b0 < 8 ? nb - b0 : 8;
        dim3 grid((unsigned)((V + 7) / 8));
        const float *xs = x + b0 * D;
        float *lg = logits + b0 * V;
        if (cnt == 1) {
            logits_kernel<1><<<grid, 256, D * sizeof(float), st>>>(xs, emb, lg, (int)D, (int)V, 1);
        } else {
            size_t smem = 8 * D * sizeof(float);
            static std::atomic<bool> attr{false};
            if (!attr && smem > 48 * 1024) { cudaFuncSetAttribute(logits_kernel<8>, cudaFuncAttributeMaxDynamicSharedMemorySize, 96 * 1024); attr = true; }
            logits_kernel<8><<<grid, 256, smem, st>>>(xs, emb, lg, (int)D, (int)V, (int)cnt);
        }
        wq4_count_launch();
    }
    WQ4_CUDA(cudaGetLastError());
    return WQ4_OK;
}

int32_t wq4m_argmax(const float *logits, int64_t nb, int64_t V, int64_t lo, int64_t hi, const int *suppress, int *out,
                    cudaStream_t st)
{
    if (hi - lo >= 8192) {
        // whole-vocabulary pick: a cluster of eight CTAs per row (one CTA per row spends ~30 us on 51 866 logits)
        cudaLaunchConfig_t cfg = {};
        cfg.gridDim = dim3(kArgCluster, (unsigned)nb, 1);
        cfg.blockDim = dim3(256, 1, 1);
        cfg.stream = st;
        cudaLaunchAttribute attr[1];
        attr[0].id = cudaLaunchAttributeClusterDimension;
        attr[0].val.clusterDim.x = kArgCluster; attr[0].val.clusterDim.y = 1; attr[0].val.clusterDim.z = 1;
        cfg.attrs = attr;
        cfg.numAttrs = 1;
        WQ4_CUDA(cudaLaunchKernelEx(&cfg, argmax_cluster_kernel, logits, (int)V, (int)lo, (int)hi, suppress, out));
        wq4_count_launch();
        return WQ4_OK;
    }
    argmax_kernel<<<(unsigned)nb, 1024, 0, st>>>(logits, (int)V, (int)lo, (int)hi, suppress, out);
    wq4_count_launch();
    WQ4_CUDA(cudaGetLastError());
    return WQ4_OK;
}

int32_t wq4m_kv_append(const float *qkv, int64_t ld_src, void *kc, void *vc, int64_t ldkv, int64_t kv_batch_stride, int64_t B,
                       int64_t D, int64_t q_len, int64_t pos0, cudaStream_t st)
{
    dim3 grid((unsigned)q_len, (unsigned)B);
    kv_append_kernel<<<grid, 256, 0, st>>>(qkv, (int)ld_src, reinterpret_cast<__half *>(kc), reinterpret_cast<__half *>(vc),
                                           (int)ldkv, (long long)kv_batch_stride, (int)D, (int)q_len, (int)pos0);
    wq4_count_launch();
    WQ4_CUDA(cudaGetLastError());
    return WQ4_OK;
}

int32_t wq4m_greedy_init(int *slot, int *n_active, int *count, int64_t B, cudaStream_t st)
{
    greedy_init_kernel<<<(unsigned)((B + 255) / 256), 256, 0, st>>>(slot, n_active, count, (int)B);
    wq4_count_launch();
    WQ4_CUDA(cudaGetLastError());
    return WQ4_OK;
}

int32_t wq4m_greedy_bookkeep(const int *next, int *tokens_next, int *slot, int *n_active, int *count, const int *budget,
                             int *tokens_out, int64_t max_tokens, int eot, cudaStream_t st)
{
    greedy_bookkeep_kernel<<<1, 1024, 0, st>>>(next, tokens_next, slot, n_active, count, budget, tokens_out, (int)max_tokens, eot);
    wq4_count_launch();
    WQ4_CUDA(cudaGetLastError());
    return WQ4_OK;
}
