// wq4_decode2.cu -- second generation of the single-clip greedy decode step (kernel family (2), config 3 of BASELINE.json):
// WhisperDecoder::decode_step (src/model/decoder.rs:306-348) with forward_with_cache of src/model/attention.rs:93-125,
// 208-236 and Q4FFN of src/model/layers.rs:54-58, as ONE persistent kernel with THREE grid-wide synchronisations per
// decoder layer instead of the eight of wq4_decode.cu.
//
// Why: at M = 1 the step is a chain of ~260 dependent vector operations; 412.9 MB of Q4 weights stream in 63 us at the
// measured HBM rate, so the step time is (number of grid-wide exchanges) x (cost of one exchange + the dependent work
// between two of them).  The first generation split every linear by OUTPUT rows over all CTAs, which needs the whole
// input vector and therefore a grid barrier in front of each of the 8 operations of a layer (258 barriers, 1486 us).
//
// Here one thread-block CLUSTER owns one attention head (and one 256-wide slice of the FFN hidden layer):
//   phase A  x -> LN1 -> q_h|k_h|v_h (192 weight rows, split over the cluster's CTAs)      [cluster exchange]
//            -> self-attention of head h over the cache, keys split over the CTAs          [cluster exchange]
//            -> out-projection restricted to the head's 64 input columns: a PARTIAL D-vector per head   -> global
//   phase B  (sum of the partials + bias + x) -> LN2 -> cross-q_h -> cross-attention of head h -> partial of cross-out
//   phase C  (sum ...) -> LN3 -> fc1 rows of the cluster's F-slice + GELU [cluster exchange] -> fc2 restricted to those
//            256 input columns: a partial D-vector per cluster
// so the only grid-wide exchange is "H partial vectors -> everyone", three times per layer.  The partial vectors are summed by
// the memory system: every CTA adds its slice of its cluster's partial into one vector of 64-bit fixed-point accumulators
// (red.global.add.u64 of value * 2^28: integer addition is associative, so the result is bit-identical run to run), and
// after the grid barrier every CTA reads the D accumulators (10 KB out of L2, one round trip) and keeps the whole residual
// vector x in shared memory for the life of the kernel.  (First form of this kernel: H f32 partial vectors in global
// memory, each CTA summing 1/CS of the columns and broadcasting them through distributed shared memory: +2 us per phase.)
//
// Weights never wait for the activations: a layer's three weight slabs of a CTA (31 + 16 + 62 KB at Large-V3) and its
// share of the head's cross K/V (64 KB) are fetched by cp.async.bulk into dedicated shared-memory buffers ONE LAYER AHEAD
// (each buffer is refilled as soon as its phase has consumed it), so by the time a phase needs them they are resident
// and the dependent chain of a phase is: poll -> 10 KB of L2 reads -> LayerNorm -> GEMV from shared memory -> DSMEM
// exchange -> attention from shared memory / registers -> DSMEM exchange -> GEMV -> fixed-point reductions + release.
//
// Decode layout (built at load, wq4m_dec2_create): per (layer, phase, CTA) one contiguous slab
//   [nib1: rows1 x kb1 uint4, row-major][scl1: rows1 x kb1 f16][nib2: rows2 x kb2 uint4][scl2]
// same 18 bytes per 32 weights as GGUF, same nibble interleave as the panel layout (wq4_common.cuh).
// GEMV mapping: G lanes of a warp share one weight row and split its k-blocks (consecutive lanes read consecutive
// 16-byte blocks: conflict-free), the activation vector sits in shared memory padded by 4 floats per 32 so that the
// eight lanes of a quarter-warp read eight different bank groups; the row sum is a shuffle reduction (no shared-memory
// round trip, no __syncthreads inside a GEMV).  Activations are f16 in shared memory and the products run on the
// mixed-precision FMA (fma.rn.f32.f16: (nib - 8) exact in f16, f32 accumulation, one f32 scale multiply per block).
#include <atomic>
#include <float.h>
#include <stdio.h>
#include <stdlib.h>

#include "wq4_common.cuh"
#include "wq4_model.cuh"
#include "wq4_ptx.cuh"

namespace {

constexpr int kT = 512;            // threads per CTA
constexpr int kMaxCS = 8;          // portable cluster size limit
constexpr int kPS = 68;            // floats per attention partial: o[64], m, l, pad
constexpr int kHC = 24;            // partial vectors summed per round trip (H <= 24: one trip)

__device__ __forceinline__ int pidx(int i) { return i + ((i >> 5) << 2); }     // padded activation index: 36 floats per 32

__device__ __forceinline__ float warp_sum2(float v)
{
#pragma unroll
    for (int o = 16; o > 0; o >>= 1) v += __shfl_xor_sync(0xffffffffu, v, o);
    return v;
}
__device__ __forceinline__ uint16_t f2h_sat2(float v)
{
    uint16_t h;
    asm("cvt.rn.satfinite.f16.f32 %0, %1;" : "=h"(h) : "f"(v));
    return h;
}
__device__ __forceinline__ float h_round(float v) { return __half2float(__ushort_as_half(f2h_sat2(v))); }   // what the f16 cache will hold
__device__ __forceinline__ void st_cluster_f32(uint32_t addr, float v)
{
    asm volatile("st.shared::cluster.f32 [%0], %1;" ::"r"(addr), "f"(v) : "memory");
}
__device__ __forceinline__ void st_cluster_b16(uint32_t addr, uint16_t v)
{
    asm volatile("st.shared::cluster.b16 [%0], %1;" ::"r"(addr), "h"(v) : "memory");
}
__device__ __forceinline__ unsigned ld_acquire_gpu(const unsigned *p)
{
    unsigned v;
    asm volatile("ld.acquire.gpu.global.u32 %0, [%1];" : "=r"(v) : "l"(p) : "memory");
    return v;
}
__device__ __forceinline__ void red_release_gpu_add(unsigned *p, unsigned v)
{
    asm volatile("red.release.gpu.global.add.u32 [%0], %1;" ::"l"(p), "r"(v) : "memory");
}
__device__ __forceinline__ uint4 ldg_nc16(const void *p)
{
    uint4 r;
    asm volatile("ld.global.nc.L1::no_allocate.v4.u32 {%0,%1,%2,%3}, [%4];" : "=r"(r.x), "=r"(r.y), "=r"(r.z), "=r"(r.w) : "l"(p));
    return r;
}

// 8 weights of one device word -> (nib - 8) as f32, exact (same extraction as wq4_decode.cu)
__device__ __forceinline__ void word_f32(uint32_t w, float (&f)[8])
{
    const __half2 k1032 = __floats2half2_rn(1032.0f, 1032.0f);
    const __half2 k16th = __floats2half2_rn(0.0625f, 0.0625f);
    const __half2 km72 = __floats2half2_rn(-72.0f, -72.0f);
    uint32_t t0 = (w & 0x000F000Fu) | 0x64006400u;
    uint32_t t1 = (w & 0x00F000F0u) | 0x64006400u;
    const uint32_t w8 = w >> 8;
    uint32_t t2 = (w8 & 0x000F000Fu) | 0x64006400u;
    uint32_t t3 = (w8 & 0x00F000F0u) | 0x64006400u;
    const float2 a = __half22float2(__hsub2(*reinterpret_cast<__half2 *>(&t0), k1032));
    const float2 b = __half22float2(__hfma2(*reinterpret_cast<__half2 *>(&t1), k16th, km72));
    const float2 c = __half22float2(__hsub2(*reinterpret_cast<__half2 *>(&t2), k1032));
    const float2 d = __half22float2(__hfma2(*reinterpret_cast<__half2 *>(&t3), k16th, km72));
    f[0] = a.x; f[1] = a.y; f[2] = b.x; f[3] = b.y; f[4] = c.x; f[5] = c.y; f[6] = d.x; f[7] = d.y;
}

// Activation vectors that feed a GEMV are kept in shared memory as f16 (like every tensor-core operand of this library;
// the weights (nib - 8) are exact in f16) so that the products run on the sm_100 mixed-precision FMA (fma.rn.f32.f16 ->
// SASS FHFMA: f16 x f16 + f32, exact product, f32 accumulation) with no half -> float conversion: 17 instructions per
// 8 weights instead of 25.  Layout: 40 halfs (80 bytes) per block of 32, so that the eight lanes of a quarter-warp, which
// read eight different blocks, hit eight different bank groups.
__device__ __forceinline__ int hidx(int i) { return i + ((i >> 5) << 3); }
__device__ __forceinline__ void fhfma2(float &acc_lo, float &acc_hi, uint32_t a2, uint32_t b2)      // (a.lo*b.lo, a.hi*b.hi)
{
    asm("{\n\t.reg .b16 al, ah, bl, bh;\n\tmov.b32 {al, ah}, %2;\n\tmov.b32 {bl, bh}, %3;\n\t"
        "fma.rn.f32.f16 %0, al, bl, %0;\n\tfma.rn.f32.f16 %1, ah, bh, %1;\n\t}"
        : "+f"(acc_lo), "+f"(acc_hi) : "r"(a2), "r"(b2));
}
// sum_k (nib - 8) * x[k] over one 32-weight block; xh = the block's 32 activations as f16 (16-byte aligned)
__device__ __forceinline__ float block_dot32_h(uint4 q, const __half *xh)
{
    const __half2 k1032 = __floats2half2_rn(1032.0f, 1032.0f);
    const __half2 k16th = __floats2half2_rn(0.0625f, 0.0625f);
    const __half2 km72 = __floats2half2_rn(-72.0f, -72.0f);
    float s0 = 0.0f, s1 = 0.0f, s2 = 0.0f, s3 = 0.0f;
    const uint32_t w[4] = {q.x, q.y, q.z, q.w};
#pragma unroll
    for (int c = 0; c < 4; c++) {
        const uint4 xq = *reinterpret_cast<const uint4 *>(xh + c * 8);
        uint32_t t0 = (w[c] & 0x000F000Fu) | 0x64006400u;
        uint32_t t1 = (w[c] & 0x00F000F0u) | 0x64006400u;
        const uint32_t w8 = w[c] >> 8;
        uint32_t t2 = (w8 & 0x000F000Fu) | 0x64006400u;
        uint32_t t3 = (w8 & 0x00F000F0u) | 0x64006400u;
        const __half2 h0 = __hsub2(*reinterpret_cast<__half2 *>(&t0), k1032);               // (e0 - 8, e1 - 8), exact
        const __half2 h1 = __hfma2(*reinterpret_cast<__half2 *>(&t1), k16th, km72);        // (e2 - 8, e3 - 8)
        const __half2 h2 = __hsub2(*reinterpret_cast<__half2 *>(&t2), k1032);
        const __half2 h3 = __hfma2(*reinterpret_cast<__half2 *>(&t3), k16th, km72);
        fhfma2(s0, s1, *reinterpret_cast<const uint32_t *>(&h0), xq.x);
        fhfma2(s2, s3, *reinterpret_cast<const uint32_t *>(&h1), xq.y);
        fhfma2(s0, s1, *reinterpret_cast<const uint32_t *>(&h2), xq.z);
        fhfma2(s2, s3, *reinterpret_cast<const uint32_t *>(&h3), xq.w);
    }
    return (s0 + s1) + (s2 + s3);
}

// One GEMV job of this CTA out of shared memory: y[r] = sum_k W[r][k] x[k], r < rows.  G lanes per row (power of two).
//   nib : [rows_alloc][kb] uint4 row-major, scl : [rows_alloc][kb] f16;   xp : padded f16 activation (hidx);   epi(r, y)
template <int G, typename Epi>
__device__ __forceinline__ void gemv_job(const uint4 *nib, const __half *scl, int rows, int kb, const __half *xp, Epi epi)
{
    constexpr int RPP = kT / G;
    const int g = threadIdx.x & (G - 1), rl = threadIdx.x / G;
    for (int r0 = 0; r0 < rows; r0 += RPP) {            // trip count uniform over the CTA (shuffles inside)
        const int r = r0 + rl;
        float acc = 0.0f;
        if (r < rows) {
            const uint4 *nr = nib + (size_t)r * kb;
            const __half *sr = scl + (size_t)r * kb;
#pragma unroll 2
            for (int k = g; k < kb; k += G) acc = fmaf(__half2float(sr[k]), block_dot32_h(nr[k], xp + k * 40), acc);
        }
#pragma unroll
        for (int o = G >> 1; o > 0; o >>= 1) acc += __shfl_xor_sync(0xffffffffu, acc, o);
        if (r < rows && g == 0) epi(r, acc);
    }
}

struct Soft2 { float m, l, o[8]; };
__device__ __forceinline__ void soft_merge2(Soft2 &a, float m2, float l2, const float (&o2)[8])
{
    const float mn = fmaxf(a.m, m2);
    const float sa = (a.m == -INFINITY) ? 0.0f : __expf(a.m - mn);
    const float sb = (m2 == -INFINITY) ? 0.0f : __expf(m2 - mn);
    a.l = a.l * sa + l2 * sb;
#pragma unroll
    for (int j = 0; j < 8; j++) a.o[j] = a.o[j] * sa + o2[j] * sb;
    a.m = mn;
}
// one key into the running softmax of this 8-lane group: kq / vq = the lane's 8 dims of the key / value row (f16)
__device__ __forceinline__ void soft_key(Soft2 &st, const float (&qv)[8], uint4 kq, uint4 vq, bool valid)
{
    float sc = 0.0f;
    const __half2 *hp = reinterpret_cast<const __half2 *>(&kq);
#pragma unroll
    for (int e = 0; e < 4; e++) {
        const float2 f = __half22float2(hp[e]);
        sc = fmaf(qv[2 * e], f.x, sc);
        sc = fmaf(qv[2 * e + 1], f.y, sc);
    }
    sc += __shfl_xor_sync(0xffffffffu, sc, 1);
    sc += __shfl_xor_sync(0xffffffffu, sc, 2);
    sc += __shfl_xor_sync(0xffffffffu, sc, 4);
    if (valid) {
        const float mn = fmaxf(st.m, sc);
        const float alpha = (st.m == -INFINITY) ? 0.0f : __expf(st.m - mn);
        const float pw = __expf(sc - mn);
        st.l = st.l * alpha + pw;
        const __half2 *vp2 = reinterpret_cast<const __half2 *>(&vq);
#pragma unroll
        for (int e = 0; e < 4; e++) {
            const float2 f = __half22float2(vp2[e]);
            st.o[2 * e] = st.o[2 * e] * alpha + pw * f.x;
            st.o[2 * e + 1] = st.o[2 * e + 1] * alpha + pw * f.y;
        }
        st.m = mn;
    }
}

struct Smem2 {
    float *xres, *xn, *qkvh, *clp, *wpart, *red;
    __half *xn16, *atth16, *hsl16;      // f16 operands of the GEMVs (hidx layout)
    uint64_t *bars;
    uint8_t *bufA, *bufB, *bufC, *bufKV;
};

// merge the CTA's 64 key groups -> one partial, push it to every CTA of the cluster, merge the cluster's partials into
// atth (padded).  Contains two __syncthreads and one cluster barrier; every thread of every CTA of the cluster calls it.
__device__ __forceinline__ void attn_exchange(Soft2 st, const Smem2 &s, int rank, int CS)
{
    const int tid = threadIdx.x, warp = tid >> 5, lane = tid & 31, dch = lane & 7;
#pragma unroll
    for (int off = 8; off <= 16; off <<= 1) {                // the four key groups of a warp (lanes with equal dch)
        const float m2 = __shfl_xor_sync(0xffffffffu, st.m, off);
        const float l2 = __shfl_xor_sync(0xffffffffu, st.l, off);
        float o2[8];
#pragma unroll
        for (int j = 0; j < 8; j++) o2[j] = __shfl_xor_sync(0xffffffffu, st.o[j], off);
        soft_merge2(st, m2, l2, o2);
    }
    if (lane < 8) {
        float *dst = s.wpart + warp * kPS;
        if (dch == 0) { dst[64] = st.m; dst[65] = st.l; }
#pragma unroll
        for (int j = 0; j < 8; j++) dst[dch * 8 + j] = st.o[j];
    }
    __syncthreads();
    if (tid < 64) {                                          // the 16 warps' partials, fixed order; thread = output dim
        float M = -INFINITY;
#pragma unroll
        for (int w2 = 0; w2 < kT / 32; w2++) M = fmaxf(M, s.wpart[w2 * kPS + 64]);
        float L = 0.0f, o = 0.0f;
#pragma unroll
        for (int w2 = 0; w2 < kT / 32; w2++) {
            const float mw = s.wpart[w2 * kPS + 64];
            const float sc = (mw == -INFINITY) ? 0.0f : __expf(mw - M);
            L += sc * s.wpart[w2 * kPS + 65];
            o += sc * s.wpart[w2 * kPS + tid];
        }
        const uint32_t base = ptx::smem_u32(s.clp + rank * kPS);
        for (int p = 0; p < CS; p++) {
            st_cluster_f32(ptx::mapa_u32(base + tid * 4, p), o);
            if (tid == 0) { st_cluster_f32(ptx::mapa_u32(base + 64 * 4, p), M); st_cluster_f32(ptx::mapa_u32(base + 65 * 4, p), L); }
        }
    }
    ptx::cluster_sync_all();
    if (tid < 64) {
        float M = -INFINITY;
        for (int p = 0; p < CS; p++) M = fmaxf(M, s.clp[p * kPS + 64]);
        float L = 0.0f, o = 0.0f;
        for (int p = 0; p < CS; p++) {                       // fixed order: deterministic
            const float mp = s.clp[p * kPS + 64];
            const float sc = (mp == -INFINITY) ? 0.0f : __expf(mp - M);
            L += sc * s.clp[p * kPS + 65];
            o += sc * s.clp[p * kPS + tid];
        }
        s.atth16[hidx(tid)] = __ushort_as_half(f2h_sat2(o / L));
    }
    __syncthreads();
}

// The partial D-vectors of a phase (one per cluster) are summed by the memory system: every CTA adds its slice of its
// cluster's partial into ONE vector of 64-bit fixed-point accumulators (value * 2^28, red.global.add.u64).  Integer
// addition is associative, so the sum is bit-identical run to run whatever the arrival order, and it is at least as
// accurate as an f32 sum (resolution 3.7e-9 absolute, range +-3.4e10).  The next phase reads D accumulators (10 KB per
// CTA out of L2, one round trip) instead of exchanging H partial vectors through distributed shared memory.
// Three accumulator vectors rotate: phase p adds into p % 3, reads (p - 1) % 3 and clears (p + 1) % 3 (last read in
// phase p - 1, next written in phase p + 1; both are a grid barrier away).  3 * n_layers phases end on vector 2, which
// the final stage reads while it clears vector 1; vectors 0 and 1 are clear again when the kernel ends and vector 2 is
// cleared in phase 1 of the next launch before phase 2 adds into it: no host-side memset per step.
constexpr float kFixScale = 268435456.0f;            // 2^28
constexpr float kFixInv = 1.0f / 268435456.0f;
struct PhasePre { float4 bias, w, b; };              // per-thread values a phase needs right after the grid barrier, requested before it

__device__ __forceinline__ void phase_prefetch(PhasePre &pp, const float *bias, const float *lnw, const float *lnb, int D)
{
    const int tid = threadIdx.x;
    pp.bias = make_float4(0.f, 0.f, 0.f, 0.f); pp.w = pp.bias; pp.b = pp.bias;
    if (tid < (D >> 2)) {
        if (bias) pp.bias = __ldg(reinterpret_cast<const float4 *>(bias) + tid);
        pp.w = __ldg(reinterpret_cast<const float4 *>(lnw) + tid);
        pp.b = __ldg(reinterpret_cast<const float4 *>(lnb) + tid);
    }
}

// x <- x + bias + (accumulated partial sums of the previous phase; acc == null: none), then LayerNorm::forward
// (src/model/layers.rs:23-31, two-pass as the reference) of x -> xn (f32, padded) and xn16 (f16 operand of the GEMVs).
// Every CTA does the whole vector: thread t owns columns 4t..4t+3 from the accumulator read to the normalised store, so
// the only block-wide synchronisations are the two statistics and the final hand-over.  D % 4 == 0, D <= 4 * kT.
// `clr` = the accumulator vector to clear for the phase after next (this CTA's share), or null.
__device__ __forceinline__ void reduce_ln(const long long *acc, const PhasePre &pp, const Smem2 &s, int D, long long *clr, int n_cta)
{
    const int tid = threadIdx.x, warp = tid >> 5, lane = tid & 31;
    const bool on = tid < (D >> 2);
    float4 v = make_float4(0.f, 0.f, 0.f, 0.f);
    if (on) {
        v = reinterpret_cast<const float4 *>(s.xres)[tid];
        if (acc) {
            const longlong2 a0 = __ldcg(reinterpret_cast<const longlong2 *>(acc) + 2 * tid);
            const longlong2 a1 = __ldcg(reinterpret_cast<const longlong2 *>(acc) + 2 * tid + 1);
            v.x += pp.bias.x + __ll2float_rn(a0.x) * kFixInv;
            v.y += pp.bias.y + __ll2float_rn(a0.y) * kFixInv;
            v.z += pp.bias.z + __ll2float_rn(a1.x) * kFixInv;
            v.w += pp.bias.w + __ll2float_rn(a1.y) * kFixInv;
            reinterpret_cast<float4 *>(s.xres)[tid] = v;
        }
    }
    if (clr)
        for (int c = (int)blockIdx.x * kT + tid; c < D; c += n_cta * kT) clr[c] = 0ll;
    const float sum = warp_sum2((v.x + v.y) + (v.z + v.w));
    if (lane == 0) s.red[warp] = sum;
    __syncthreads();
    float tot = 0.0f;
#pragma unroll
    for (int i = 0; i < kT / 32; i += 4) { const float4 t = *reinterpret_cast<const float4 *>(s.red + i); tot += (t.x + t.y) + (t.z + t.w); }
    const float mean = tot / (float)D;
    float q = 0.0f;
    if (on) { v.x -= mean; v.y -= mean; v.z -= mean; v.w -= mean; q = (v.x * v.x + v.y * v.y) + (v.z * v.z + v.w * v.w); }
    q = warp_sum2(q);
    if (lane == 0) s.red[32 + warp] = q;
    __syncthreads();
    float vt = 0.0f;
#pragma unroll
    for (int i = 0; i < kT / 32; i += 4) { const float4 t = *reinterpret_cast<const float4 *>(s.red + 32 + i); vt += (t.x + t.y) + (t.z + t.w); }
    const float inv = 1.0f / sqrtf(vt / (float)D + 1e-5f);
    if (on) {
        const float4 o = make_float4(v.x * inv * pp.w.x + pp.b.x, v.y * inv * pp.w.y + pp.b.y, v.z * inv * pp.w.z + pp.b.z, v.w * inv * pp.w.w + pp.b.w);
        *reinterpret_cast<float4 *>(s.xn + pidx(tid * 4)) = o;                              // f32 copy: the logits stage
        uint2 pk;
        pk.x = (uint32_t)f2h_sat2(o.x) | ((uint32_t)f2h_sat2(o.y) << 16);
        pk.y = (uint32_t)f2h_sat2(o.z) | ((uint32_t)f2h_sat2(o.w) << 16);
        *reinterpret_cast<uint2 *>(s.xn16 + hidx(tid * 4)) = pk;                            // f16 copy: the GEMVs
    }
    __syncthreads();
}
__device__ __forceinline__ void acc_add(long long *acc, float y)
{
    atomicAdd(reinterpret_cast<unsigned long long *>(acc), (unsigned long long)__float2ll_rn(y * kFixScale));
}

__device__ __forceinline__ void grid_arrive(unsigned *counter)
{
    __syncthreads();                      // every thread's partial sums are issued
    if (threadIdx.x == 0) red_release_gpu_add(counter, 1u);    // release: cumulative over the CTA's reductions (bar.sync above)
}
__device__ __forceinline__ void grid_wait(const unsigned *counter, unsigned target, int *dev_err)
{
    if (threadIdx.x == 0) {
        unsigned spins = 0;
        while (ld_acquire_gpu(counter) < target) {
            if (++spins > (1u << 24)) { atomicCAS(dev_err, 0, 302); break; }    // bounded: an error flag, never a hang
        }
    }
    __syncthreads();
}

__global__ void __launch_bounds__(kT, 1) decode2_kernel(const Dec2Args a)
{
    extern __shared__ __align__(128) uint8_t sm_raw[];
    const int D = a.D, H = a.H, CS = a.CS, kbD = a.kbD;
    const int tid = threadIdx.x, warp = tid >> 5, lane = tid & 31;
    const int rank = (int)ptx::cluster_ctarank();
    const int cl = (int)blockIdx.x / CS;                 // cluster = attention head = FFN slice
    const int Dp = (D >> 5) * 36;
    Smem2 s;
    s.xres = reinterpret_cast<float *>(sm_raw);
    s.xn = s.xres + D;
    s.qkvh = s.xn + Dp;
    s.clp = s.qkvh + 192;
    s.wpart = s.clp + kMaxCS * kPS;
    s.red = s.wpart + (kT / 32) * kPS;
    s.bars = reinterpret_cast<uint64_t *>(s.red + 64);
    s.xn16 = reinterpret_cast<__half *>(s.bars + 4);
    s.atth16 = s.xn16 + (D >> 5) * 40;
    s.hsl16 = s.atth16 + 2 * 40;
    s.bufA = sm_raw + a.off_bufA;
    s.bufB = s.bufA + a.sA;
    s.bufC = s.bufB + a.sB;
    s.bufKV = s.bufC + a.sC;
    const uint32_t barA = ptx::smem_u32(s.bars), barB = barA + 8, barC = barA + 16, barKV = barA + 24;
    const int clip = a.slot ? __ldg(a.slot) : 0;
    const int kv_len = a.kv_len + 1;                     // including the row appended in this step
    const int kpc = a.kpc;                               // cross keys per CTA
    const int ck0 = rank * kpc, ck1 = min(a.T_enc, ck0 + kpc);
    const uint32_t kv_bytes = ck1 > ck0 ? (uint32_t)(ck1 - ck0) * 256u : 0u;
    long long *tl = (a.timeline && blockIdx.x == 0 && tid == 0) ? a.timeline : nullptr;
    int tl_n = 0;
#define STAMP()                                                                          \
    do {                                                                                 \
        if (tl && tl_n < kDecTimelineSlots) { long long t_; asm volatile("mov.u64 %0, %%globaltimer;" : "=l"(t_)); tl[tl_n++] = t_; } \
    } while (0)

#define STAMP2(id_)                                                                      \
    do {                                                                                 \
        if (tl && l == 2) { long long t_; asm volatile("mov.u64 %0, %%globaltimer;" : "=l"(t_)); tl[480 + (id_)] = t_; } \
    } while (0)
    if (tid == 0) {
        ptx::mbar_init(barA, 1); ptx::mbar_init(barB, 1); ptx::mbar_init(barC, 1); ptx::mbar_init(barKV, 1);
        ptx::fence_mbar_init();
    }
    __syncthreads();
    auto issue_A = [&](int l) { ptx::mbar_arrive_expect_tx(barA, a.sA); ptx::bulk_load(ptx::smem_u32(s.bufA), a.layers[l].wA + (size_t)blockIdx.x * a.sA, a.sA, barA); };
    auto issue_B = [&](int l) { ptx::mbar_arrive_expect_tx(barB, a.sB); ptx::bulk_load(ptx::smem_u32(s.bufB), a.layers[l].wB + (size_t)blockIdx.x * a.sB, a.sB, barB); };
    auto issue_C = [&](int l) { ptx::mbar_arrive_expect_tx(barC, a.sC); ptx::bulk_load(ptx::smem_u32(s.bufC), a.layers[l].wC + (size_t)blockIdx.x * a.sC, a.sC, barC); };
    auto issue_KV = [&](int l) {
        if (kv_bytes) {
            const __half *src = a.cross_kv + (((size_t)l * a.kv_clips + clip) * H + cl) * (size_t)a.T_enc * 128 + (size_t)ck0 * 128;
            ptx::mbar_arrive_expect_tx(barKV, kv_bytes);
            ptx::bulk_load(ptx::smem_u32(s.bufKV), src, kv_bytes, barKV);
        }
    };
    auto wait_bar = [&](uint32_t bar, uint32_t parity) {
        unsigned spins = 0;
        while (!ptx::mbar_try_wait(bar, parity)) {
            if (++spins > (1u << 22)) { atomicCAS(a.dev_err, 0, 303); break; }
        }
    };
    if (tid == 0) { issue_A(0); issue_B(0); issue_C(0); issue_KV(0); }

    // x = token_embedding[tok] + positional_embedding[pos] (decoder.rs:317-329), replicated in every CTA
    {
        const int tok = a.tokens[0];
        const float *te = a.tok_emb + (size_t)tok * D, *pe = a.pos_emb + (size_t)a.emb_pos * D;
        for (int j = tid; j < D; j += kT) s.xres[j] = __ldg(te + j) + __ldg(pe + j);
    }
    __syncthreads();
    // every CTA of the cluster must have started before anyone writes into its shared memory
    ptx::cluster_sync_all();
    STAMP();

    const int n_cta = (int)gridDim.x;
    unsigned n_sync = 0;
    unsigned ph = 0;                                     // phase counter: phase p writes partial buffer p & 1, phase p + 1 reads it
    const int g8 = lane >> 3, dch = lane & 7;             // attention: key group inside the warp, 16-byte chunk of the row
    const int ks = tid >> 3;                             // key slot 0..63
    const int per_self = (kv_len + CS - 1) / CS;
    const int sk0 = rank * per_self, sk1 = min(kv_len, sk0 + per_self);

    for (int l = 0; l < a.n_layers; l++) {
        const Dec2Layer &L = a.layers[l];
        const uint32_t par = (uint32_t)(l & 1);
        const size_t self_base = ((size_t)l * a.kv_clips + clip) * (size_t)a.ctx * D + (size_t)cl * 64 + dch * 8;
        long long *const accv = reinterpret_cast<long long *>(a.partial);   // [3][D] fixed-point accumulators
        PhasePre pp;

        // ================= phase A: self-attention block (decoder.rs:84-92) =================
        // cached K/V rows of this CTA's key range: requested now, consumed after q is known (they do not depend on this step)
        uint4 kq[2], vq[2];
#pragma unroll
        for (int p = 0; p < 2; p++) {
            const int key = sk0 + ks + 64 * p;
            kq[p] = make_uint4(0, 0, 0, 0); vq[p] = kq[p];
            if (key < sk1 && key < kv_len - 1) {
                kq[p] = ldg_nc16(a.self_k + self_base + (size_t)key * D);
                vq[p] = ldg_nc16(a.self_v + self_base + (size_t)key * D);
            }
        }
        STAMP2(0);
        phase_prefetch(pp, l > 0 ? a.layers[l - 1].b_fc2 : nullptr, L.ln1w, L.ln1b, D);
        if (l > 0) {
            grid_wait(a.barrier, ++n_sync * n_cta, a.dev_err);
            STAMP();
            STAMP2(1);
        }
        STAMP2(2);
        reduce_ln(l > 0 ? accv + (size_t)((ph + 2) % 3) * D : nullptr, pp, s, D, l > 0 ? accv + (size_t)((ph + 1) % 3) * D : nullptr, n_cta);
        STAMP2(3);
        wait_bar(barA, par);
        STAMP2(4);
        {   // q_h | k_h | v_h rows of this CTA (attention.rs:99-104): virtual row vr = 64*kind + dim
            const uint4 *nib = reinterpret_cast<const uint4 *>(s.bufA);
            const __half *scl = reinterpret_cast<const __half *>(s.bufA + a.a_scl1);
            const int vr0 = rank * a.rq, rows = max(0, min(192, vr0 + a.rq) - vr0);
            const uint32_t qb = ptx::smem_u32(s.qkvh);
            // the row's bias is requested before the weight loop (one pass when rows <= kT / 16: every CTA of a Large-V3 cluster)
            const int pr = vr0 + tid / 16;
            const float pre_b = (L.b_qkv && pr < 192) ? __ldg(L.b_qkv + (pr >> 6) * D + cl * 64 + (pr & 63)) : 0.0f;
            gemv_job<16>(nib, scl, rows, kbD, s.xn16, [&](int r, float y) {
                const int vr = vr0 + r, kind = vr >> 6;
                if (L.b_qkv) y += (r < kT / 16) ? pre_b : __ldg(L.b_qkv + kind * D + cl * 64 + (vr & 63));
                for (int p = 0; p < CS; p++) st_cluster_f32(ptx::mapa_u32(qb + vr * 4, p), y);
            });
        }
        STAMP2(5);
        ptx::cluster_sync_all();
        STAMP();
        STAMP2(6);
        {   // self-attention of head `cl` over keys [sk0, sk1) (attention.rs:106-119); key kv_len-1 is the new row
            if (rank == 0 && tid < 128) {                // append the new K/V row to the f16 cache in place (no realloc)
                __half *dst = (tid < 64 ? a.self_k : a.self_v) + ((size_t)l * a.kv_clips + clip) * (size_t)a.ctx * D + (size_t)(kv_len - 1) * D + cl * 64 + (tid & 63);
                *dst = __ushort_as_half(f2h_sat2(s.qkvh[64 + tid]));
            }
            float qv[8];
#pragma unroll
            for (int i = 0; i < 8; i++) qv[i] = s.qkvh[dch * 8 + i] * 0.125f;      // (q.k)/sqrt(64): scaling q first is exact
            Soft2 st;
            st.m = -INFINITY; st.l = 0.0f;
#pragma unroll
            for (int j = 0; j < 8; j++) st.o[j] = 0.0f;
            const int n_pass = (sk1 - sk0 + 63) >> 6;                             // uniform over the CTA
            for (int p = 0; p < n_pass; p++) {
                const int key = sk0 + ks + 64 * p;
                const bool valid = key < sk1;
                uint4 kk = make_uint4(0, 0, 0, 0), vv = kk;
                if (valid && key == kv_len - 1) {                                   // the new row: what the cache now holds (f16)
                    uint32_t kw[4], vw[4];
#pragma unroll
                    for (int e = 0; e < 4; e++) {
                        kw[e] = (uint32_t)f2h_sat2(s.qkvh[64 + dch * 8 + 2 * e]) | ((uint32_t)f2h_sat2(s.qkvh[64 + dch * 8 + 2 * e + 1]) << 16);
                        vw[e] = (uint32_t)f2h_sat2(s.qkvh[128 + dch * 8 + 2 * e]) | ((uint32_t)f2h_sat2(s.qkvh[128 + dch * 8 + 2 * e + 1]) << 16);
                    }
                    kk = make_uint4(kw[0], kw[1], kw[2], kw[3]); vv = make_uint4(vw[0], vw[1], vw[2], vw[3]);
                } else if (valid) {
                    if (p < 2) { kk = kq[p]; vv = vq[p]; }
                    else { kk = ldg_nc16(a.self_k + self_base + (size_t)key * D); vv = ldg_nc16(a.self_v + self_base + (size_t)key * D); }
                }
                soft_key(st, qv, kk, vv, valid);
            }
            STAMP2(7);
            attn_exchange(st, s, rank, CS);
        }
        STAMP();
        STAMP2(8);
        {   // out-projection restricted to the head's 64 input columns, added into the phase's accumulator (bias added by the reader)
            const uint4 *nib = reinterpret_cast<const uint4 *>(s.bufA + a.a_nib2);
            const __half *scl = reinterpret_cast<const __half *>(s.bufA + a.a_scl2);
            const int n0 = rank * a.ro, rows = max(0, min(D, n0 + a.ro) - n0);
            long long *dst = accv + (size_t)(ph++ % 3) * D + n0;
            gemv_job<2>(nib, scl, rows, 2, s.atth16, [&](int r, float y) { acc_add(dst + r, y); });
        }
        STAMP2(9);
        grid_arrive(a.barrier);
        if (tid == 0 && l + 1 < a.n_layers) { ptx::fence_proxy_async_smem(); issue_A(l + 1); }
        STAMP();
        STAMP2(10);

        // ================= phase B: cross-attention block (decoder.rs:95-104) =================
        phase_prefetch(pp, L.b_out, L.ln2w, L.ln2b, D);
        grid_wait(a.barrier, ++n_sync * n_cta, a.dev_err);
        STAMP();
        STAMP2(11);
        STAMP2(12);
        reduce_ln(accv + (size_t)((ph + 2) % 3) * D, pp, s, D, accv + (size_t)((ph + 1) % 3) * D, n_cta);
        STAMP2(13);
        wait_bar(barB, par);
        STAMP2(14);
        {   // cross-attention query rows of this CTA (attention.rs:214-217)
            const uint4 *nib = reinterpret_cast<const uint4 *>(s.bufB);
            const __half *scl = reinterpret_cast<const __half *>(s.bufB + a.b_scl1);
            const int vr0 = rank * a.rc, rows = max(0, min(64, vr0 + a.rc) - vr0);
            const uint32_t qb = ptx::smem_u32(s.qkvh);
            const int pr = vr0 + tid / 32;
            const float pre_b = (L.b_cq && pr < 64) ? __ldg(L.b_cq + cl * 64 + pr) : 0.0f;
            gemv_job<32>(nib, scl, rows, kbD, s.xn16, [&](int r, float y) {
                const int vr = vr0 + r;
                if (L.b_cq) y += (r < kT / 32) ? pre_b : __ldg(L.b_cq + cl * 64 + vr);
                for (int p = 0; p < CS; p++) st_cluster_f32(ptx::mapa_u32(qb + vr * 4, p), y);
            });
        }
        STAMP2(15);
        ptx::cluster_sync_all();
        STAMP();
        STAMP2(16);
        {   // cross-attention of head `cl` over this CTA's keys, K/V rows resident in shared memory (attention.rs:219-233)
            if (kv_bytes) wait_bar(barKV, par);
            float qv[8];
#pragma unroll
            for (int i = 0; i < 8; i++) qv[i] = s.qkvh[dch * 8 + i] * 0.125f;
            Soft2 st;
            st.m = -INFINITY; st.l = 0.0f;
#pragma unroll
            for (int j = 0; j < 8; j++) st.o[j] = 0.0f;
            const int nk = ck1 - ck0, n_pass = (max(nk, 0) + 63) >> 6;
            for (int p = 0; p < n_pass; p++) {
                const int r = ks + 64 * p;
                const bool valid = r < nk;
                uint4 kk = make_uint4(0, 0, 0, 0), vv = kk;
                if (valid) {
                    kk = *reinterpret_cast<const uint4 *>(s.bufKV + (size_t)r * 256 + dch * 16);
                    vv = *reinterpret_cast<const uint4 *>(s.bufKV + (size_t)r * 256 + 128 + dch * 16);
                }
                soft_key(st, qv, kk, vv, valid);
            }
            STAMP2(17);
            attn_exchange(st, s, rank, CS);        // its first __syncthreads also ends every read of bufKV
            if (tid == 0 && l + 1 < a.n_layers) { ptx::fence_proxy_async_smem(); issue_KV(l + 1); }
        }
        STAMP();
        STAMP2(18);
        {
            const uint4 *nib = reinterpret_cast<const uint4 *>(s.bufB + a.b_nib2);
            const __half *scl = reinterpret_cast<const __half *>(s.bufB + a.b_scl2);
            const int n0 = rank * a.ro, rows = max(0, min(D, n0 + a.ro) - n0);
            long long *dst = accv + (size_t)(ph++ % 3) * D + n0;
            gemv_job<2>(nib, scl, rows, 2, s.atth16, [&](int r, float y) { acc_add(dst + r, y); });
        }
        STAMP2(19);
        grid_arrive(a.barrier);
        if (tid == 0 && l + 1 < a.n_layers) { ptx::fence_proxy_async_smem(); issue_B(l + 1); }
        STAMP();
        STAMP2(20);

        // ================= phase C: feed-forward block (layers.rs:54-58, decoder.rs:107-110) =================
        phase_prefetch(pp, L.b_cout, L.ln3w, L.ln3b, D);
        grid_wait(a.barrier, ++n_sync * n_cta, a.dev_err);
        STAMP();
        STAMP2(21);
        STAMP2(22);
        reduce_ln(accv + (size_t)((ph + 2) % 3) * D, pp, s, D, accv + (size_t)((ph + 1) % 3) * D, n_cta);
        STAMP2(23);
        wait_bar(barC, par);
        STAMP2(24);
        {   // fc1 rows of the cluster's 256-wide slice of the hidden layer + tanh-GELU (layers.rs:35-41,55-56)
            const uint4 *nib = reinterpret_cast<const uint4 *>(s.bufC);
            const __half *scl = reinterpret_cast<const __half *>(s.bufC + a.c_scl1);
            const int vr0 = rank * a.rf, rows = max(0, min(256, vr0 + a.rf) - vr0);
            const uint32_t hb = ptx::smem_u32(s.hsl16);
            const int pr = vr0 + tid / 8;
            const float pre_b = (L.b_fc1 && pr < 256) ? __ldg(L.b_fc1 + cl * 256 + pr) : 0.0f;
            gemv_job<8>(nib, scl, rows, kbD, s.xn16, [&](int r, float y) {
                const int vr = vr0 + r;
                if (L.b_fc1) y += (r < kT / 8) ? pre_b : __ldg(L.b_fc1 + cl * 256 + vr);
                y = wq4_gelu_tanh(y);
                const uint32_t off = (uint32_t)hidx(vr) * 2u;
                const uint16_t yh = f2h_sat2(y);
                for (int p = 0; p < CS; p++) st_cluster_b16(ptx::mapa_u32(hb + off, p), yh);
            });
        }
        STAMP2(25);
        ptx::cluster_sync_all();
        STAMP();
        STAMP2(26);
        {   // fc2 restricted to the slice's 256 input columns, added into the phase's accumulator
            const uint4 *nib = reinterpret_cast<const uint4 *>(s.bufC + a.c_nib2);
            const __half *scl = reinterpret_cast<const __half *>(s.bufC + a.c_scl2);
            const int n0 = rank * a.ro, rows = max(0, min(D, n0 + a.ro) - n0);
            long long *dst = accv + (size_t)(ph++ % 3) * D + n0;
            gemv_job<8>(nib, scl, rows, 8, s.hsl16, [&](int r, float y) { acc_add(dst + r, y); });
        }
        STAMP2(27);
        grid_arrive(a.barrier);
        if (tid == 0 && l + 1 < a.n_layers) { ptx::fence_proxy_async_smem(); issue_C(l + 1); }
        STAMP();
        STAMP2(28);
    }

    // ---- final LayerNorm + logits = x . token_embedding^T (decoder.rs:338-343) against the f16 transposed matrix
    // ([D/8][Vp] 16-byte chunks: lane = vocabulary row, coalesced, no cross-lane reduction), greedy pick (whisper.rs:131-138)
    {
        PhasePre pp;
        phase_prefetch(pp, a.layers[a.n_layers - 1].b_fc2, a.lnfw, a.lnfb, D);
        grid_wait(a.barrier, ++n_sync * n_cta, a.dev_err);
        STAMP();
        // reads vector 2 and clears vector 1 (read in the last phase); vector 0 was cleared in the last phase, vector 2 is
        // cleared in phase 1 of the next launch, before phase 2 adds into it
        long long *const accv = reinterpret_cast<long long *>(a.partial);
        reduce_ln(accv + (size_t)((ph + 2) % 3) * D, pp, s, D, accv + (size_t)((ph + 1) % 3) * D, n_cta);
    }
    const int V = a.V, Vp = a.Vp, chunks = D >> 3;
    const bool pick = a.arg_hi > a.arg_lo && a.next_out != nullptr;
    float best = -INFINITY;
    int besti = -1;
    const int sup = a.suppress ? a.suppress[0] : -1;
    const int ngroups = Vp >> 5;
    for (int gidx = (int)blockIdx.x * (kT / 32) + warp; gidx < ngroups; gidx += n_cta * (kT / 32)) {
        const int v = gidx * 32 + lane;
        const uint4 *ep = a.tok_emb16t + v;
        float acc0 = 0.0f, acc1 = 0.0f;
        for (int c0 = 0; c0 < chunks; c0 += 8) {
            uint4 e[8];
#pragma unroll
            for (int u = 0; u < 8; u++)
                if (c0 + u < chunks) e[u] = ldg_nc16(ep + (size_t)(c0 + u) * Vp);
#pragma unroll
            for (int u = 0; u < 8; u++) {
                if (c0 + u < chunks) {
                    const __half2 *hp = reinterpret_cast<const __half2 *>(&e[u]);
                    const float2 f0 = __half22float2(hp[0]), f1 = __half22float2(hp[1]);
                    const float2 f2 = __half22float2(hp[2]), f3 = __half22float2(hp[3]);
                    const float *xp = s.xn + pidx((c0 + u) * 8);
                    const float4 xa = *reinterpret_cast<const float4 *>(xp);
                    const float4 xb = *reinterpret_cast<const float4 *>(xp + 4);
                    acc0 = fmaf(f0.x, xa.x, acc0); acc0 = fmaf(f0.y, xa.y, acc0); acc0 = fmaf(f1.x, xa.z, acc0); acc0 = fmaf(f1.y, xa.w, acc0);
                    acc1 = fmaf(f2.x, xb.x, acc1); acc1 = fmaf(f2.y, xb.y, acc1); acc1 = fmaf(f3.x, xb.z, acc1); acc1 = fmaf(f3.y, xb.w, acc1);
                }
            }
        }
        const float acc = acc0 + acc1;
        if (v < V) {
            a.logits[v] = acc;
            // whisper.rs:131-138 max_by: the LAST maximum wins; a lane's rows ascend, so ">=" keeps the later one
            if (pick && v >= a.arg_lo && v < a.arg_hi && v != sup && (besti < 0 || acc >= best)) { best = acc; besti = v; }
        }
    }
    STAMP();
    if (!pick) return;                                   // uniform over the grid
#pragma unroll
    for (int off = 16; off > 0; off >>= 1) {             // higher index wins ties
        const float ov = __shfl_xor_sync(0xffffffffu, best, off);
        const int oi = __shfl_xor_sync(0xffffffffu, besti, off);
        if (oi >= 0 && (besti < 0 || ov > best || (ov == best && oi > besti))) { best = ov; besti = oi; }
    }
    if (lane == 0) { s.red[warp * 2] = best; s.red[warp * 2 + 1] = __int_as_float(besti); }
    __syncthreads();
    if (tid == 0) {
        float bv = -INFINITY; int bi = -1;
        for (int w2 = 0; w2 < kT / 32; w2++) {
            const float v = s.red[w2 * 2]; const int i = __float_as_int(s.red[w2 * 2 + 1]);
            if (i >= 0 && (bi < 0 || v > bv || (v == bv && i > bi))) { bv = v; bi = i; }
        }
        a.arg_scratch[(size_t)blockIdx.x * 2] = bv;
        a.arg_scratch[(size_t)blockIdx.x * 2 + 1] = __int_as_float(bi);
        __threadfence();
        const unsigned prev = atomicAdd(a.barrier + 1, 1u);
        if (prev == (unsigned)n_cta - 1) {               // last arriver: merge the CTAs' candidates (order-independent rule)
            __threadfence();
            bv = -INFINITY; bi = -1;
            for (int c = 0; c < n_cta; c++) {
                const float v = __ldcg(a.arg_scratch + (size_t)c * 2);
                const int i = __float_as_int(__ldcg(a.arg_scratch + (size_t)c * 2 + 1));
                if (i >= 0 && (bi < 0 || v > bv || (v == bv && i > bi))) { bv = v; bi = i; }
            }
            a.next_out[0] = bi;
        }
    }
    STAMP();
#undef STAMP
#undef STAMP2
}

// ---- decode layout builder: panel layout -> per-CTA slabs -----------------------------------------------------------
// mode 0: q|k|v   virtual row vr = j*rpc + r < 192: n = (vr/64)*D + h*64 + vr%64, k-block = k
// mode 1: slice   n = j*rpc + r < n_total, k-block = kb0_per_cluster*h + k      (out / cross-out: 2 per head; fc2: 8 per slice)
// mode 2: rows    vr = j*rpc + r < rows_per_cluster: n = h*rows_per_cluster + vr, k-block = k   (cross-q: 64; fc1: 256)
__global__ void dec2_pack_kernel(const uint4 *__restrict__ pnib, const __half *__restrict__ pscl, int kbp, uint8_t *__restrict__ slabs,
                                 size_t slab_stride, uint32_t nib_off, uint32_t scl_off, int mode, int CS, int n_cta, int rpc, int kbj,
                                 int D, int rows_per_cluster, int n_total)
{
    const long long total = (long long)n_cta * rpc * kbj;
    for (long long i = (long long)blockIdx.x * blockDim.x + threadIdx.x; i < total; i += (long long)gridDim.x * blockDim.x) {
        const int k = (int)(i % kbj);
        const int r = (int)((i / kbj) % rpc);
        const int cta = (int)(i / ((long long)kbj * rpc));
        const int h = cta / CS, j = cta % CS;
        int n = -1, kbi = k;
        if (mode == 0) { const int vr = j * rpc + r; if (vr < 192) n = (vr >> 6) * D + h * 64 + (vr & 63); }
        else if (mode == 1) { const int nn = j * rpc + r; if (nn < n_total) n = nn; kbi = kbj * h + k; }
        else { const int vr = j * rpc + r; if (vr < rows_per_cluster) n = h * rows_per_cluster + vr; }
        uint4 q = make_uint4(0x88888888u, 0x88888888u, 0x88888888u, 0x88888888u);
        __half d = __ushort_as_half((unsigned short)0);
        if (n >= 0) {
            const size_t src = ((size_t)(n / WQ4_PANEL_ROWS) * kbp + kbi) * WQ4_PANEL_ROWS + (n % WQ4_PANEL_ROWS);
            q = pnib[src];
            d = pscl[src];
        }
        uint8_t *slab = slabs + (size_t)cta * slab_stride;
        reinterpret_cast<uint4 *>(slab + nib_off)[(size_t)r * kbj + k] = q;
        reinterpret_cast<__half *>(slab + scl_off)[(size_t)r * kbj + k] = d;
    }
}

inline uint32_t up16(size_t v) { return (uint32_t)((v + 15) / 16 * 16); }
inline uint32_t up128(size_t v) { return (uint32_t)((v + 127) / 128 * 128); }

}  // namespace

// Geometry of the cluster decomposition for a model (D, F, H) and cluster size CS.
static void dec2_geometry(Dec2Args &a, int D, int F, int H, int T_enc, int CS)
{
    a.D = D; a.F = F; a.H = H; a.T_enc = T_enc; a.CS = CS; a.kbD = D / 32;
    a.rq = (192 + CS - 1) / CS;
    a.ro = ((D + CS - 1) / CS + 3) / 4 * 4;
    a.rc = (64 + CS - 1) / CS;
    a.rf = (256 + CS - 1) / CS;
    a.kpc = (T_enc + CS - 1) / CS;
    const int kb = a.kbD;
    a.a_scl1 = up16((size_t)a.rq * kb * 16);
    a.a_nib2 = a.a_scl1 + up16((size_t)a.rq * kb * 2);
    a.a_scl2 = a.a_nib2 + up16((size_t)a.ro * 2 * 16);
    a.sA = up128(a.a_scl2 + up16((size_t)a.ro * 2 * 2));
    a.b_scl1 = up16((size_t)a.rc * kb * 16);
    a.b_nib2 = a.b_scl1 + up16((size_t)a.rc * kb * 2);
    a.b_scl2 = a.b_nib2 + up16((size_t)a.ro * 2 * 16);
    a.sB = up128(a.b_scl2 + up16((size_t)a.ro * 2 * 2));
    a.c_scl1 = up16((size_t)a.rf * kb * 16);
    a.c_nib2 = a.c_scl1 + up16((size_t)a.rf * kb * 2);
    a.c_scl2 = a.c_nib2 + up16((size_t)a.ro * 8 * 16);
    a.sC = up128(a.c_scl2 + up16((size_t)a.ro * 8 * 2));
    const int Dp = D / 32 * 36;
    const size_t floats = (size_t)D + Dp + 192 + kMaxCS * kPS + (kT / 32) * kPS + 64;
    const size_t halfs = (size_t)(D / 32 + 2 + 8) * 40;
    a.off_bufA = up128(floats * 4 + 4 * 8 + halfs * 2);
}

static size_t dec2_smem_bytes(const Dec2Args &a) { return (size_t)a.off_bufA + a.sA + a.sB + a.sC + up128((size_t)a.kpc * 256); }

// The kernel's dynamic shared-memory limit is a per-device function attribute that both the planner (one query per
// candidate cluster size, per model) and the launcher change: one cache of the value last set, so that a model planned
// later with a smaller footprint cannot leave the limit below what an earlier model's step needs.
static std::atomic<int> g_dec2_smem_attr[64];
static cudaError_t dec2_set_smem_limit(int device, size_t smem)
{
    if (device >= 0 && device < 64 && g_dec2_smem_attr[device].load() == (int)smem) return cudaSuccess;
    const cudaError_t e = cudaFuncSetAttribute(decode2_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem);
    if (e == cudaSuccess && device >= 0 && device < 64) g_dec2_smem_attr[device].store((int)smem);
    return e;
}

bool wq4m_dec2_supported(int64_t D, int64_t F, int64_t H)
{
    return D % 64 == 0 && D == H * 64 && F == 4 * D && D <= 4 * kT && H >= 1 && H <= 148;
}

// Largest cluster size (<= 8) for which H clusters are co-resident on `device` with the kernel's shared-memory footprint.
int32_t wq4m_dec2_plan(int device, int64_t D, int64_t F, int64_t H, int64_t T_enc, int *cluster_size)
{
    *cluster_size = 0;
    if (!wq4m_dec2_supported(D, F, H)) return WQ4_OK;
    static const int forced = [] { const char *e = getenv("WQ4_DECODE2_CS"); return e ? atoi(e) : 0; }();
    const int sms = wq4_sm_count(device);
    for (int cs = kMaxCS; cs >= 1; cs--) {
        if (forced > 0 && cs != forced) continue;
        if ((int64_t)cs * H > sms) continue;
        Dec2Args g = {};
        dec2_geometry(g, (int)D, (int)F, (int)H, (int)T_enc, cs);
        const size_t smem = dec2_smem_bytes(g);
        if (smem > 227 * 1024) continue;
        if (dec2_set_smem_limit(device, smem) != cudaSuccess) { cudaGetLastError(); continue; }
        cudaLaunchConfig_t cfg = {};
        cfg.gridDim = dim3((unsigned)(cs * H), 1, 1);
        cfg.blockDim = dim3(kT, 1, 1);
        cfg.dynamicSmemBytes = smem;
        cudaLaunchAttribute attr[1];
        attr[0].id = cudaLaunchAttributeClusterDimension;
        attr[0].val.clusterDim.x = (unsigned)cs; attr[0].val.clusterDim.y = 1; attr[0].val.clusterDim.z = 1;
        cfg.attrs = attr;
        cfg.numAttrs = 1;
        int n_clusters = 0;
        if (cudaOccupancyMaxActiveClusters(&n_clusters, decode2_kernel, &cfg) != cudaSuccess) { cudaGetLastError(); continue; }
        if (n_clusters >= H) { *cluster_size = cs; return WQ4_OK; }
    }
    return WQ4_OK;
}

// Builds the slabs of one decoder layer.  Returns the three device allocations through out->wA/wB/wC (caller frees).
int32_t wq4m_dec2_layer_create(const Dec2Args &geo, const wq4_tensor *qkv, const wq4_tensor *out, const wq4_tensor *cq, const wq4_tensor *cout,
                               const wq4_tensor *fc1, const wq4_tensor *fc2, Dec2Layer *L, cudaStream_t st)
{
    const int n_cta = geo.CS * geo.H, D = geo.D, kb = geo.kbD;
    const wq4_tensor *all[6] = {qkv, out, cq, cout, fc1, fc2};
    for (auto *t : all)
        if (!t || !t->matrix_mode) { wq4_set_error("decode layout (v2) needs [N,K] tensors with K %% 32 == 0"); return WQ4_ERR_UNSUPPORTED; }
    if (qkv->n != 3 * D || qkv->k != D || out->n != D || out->k != D || cq->n != D || cq->k != D || cout->n != D || cout->k != D ||
        fc1->n != geo.F || fc1->k != D || fc2->n != D || fc2->k != geo.F) { wq4_set_error("decode layout (v2): unexpected decoder weight shapes"); return WQ4_ERR_SHAPE; }
    uint8_t *wA = nullptr, *wB = nullptr, *wC = nullptr;
    WQ4_CUDA(cudaMalloc(&wA, (size_t)n_cta * geo.sA));
    WQ4_CUDA(cudaMalloc(&wB, (size_t)n_cta * geo.sB));
    WQ4_CUDA(cudaMalloc(&wC, (size_t)n_cta * geo.sC));
    WQ4_CUDA(cudaMemsetAsync(wA, 0, (size_t)n_cta * geo.sA, st));
    WQ4_CUDA(cudaMemsetAsync(wB, 0, (size_t)n_cta * geo.sB, st));
    WQ4_CUDA(cudaMemsetAsync(wC, 0, (size_t)n_cta * geo.sC, st));
    auto pack = [&](const wq4_tensor *t, uint8_t *slabs, size_t stride, uint32_t nib_off, uint32_t scl_off, int mode, int rpc, int kbj,
                    int rows_per_cluster, int n_total) {
        const long long total = (long long)n_cta * rpc * kbj;
        const unsigned grid = (unsigned)std::min<long long>((total + 255) / 256, 4096);
        dec2_pack_kernel<<<grid, 256, 0, st>>>(t->nib, t->scl, (int)t->kbp, slabs, stride, nib_off, scl_off, mode, geo.CS, n_cta, rpc, kbj, D,
                                               rows_per_cluster, n_total);
        wq4_count_launch();
    };
    pack(qkv, wA, geo.sA, 0, geo.a_scl1, 0, geo.rq, kb, 192, 0);
    pack(out, wA, geo.sA, geo.a_nib2, geo.a_scl2, 1, geo.ro, 2, 0, D);
    pack(cq, wB, geo.sB, 0, geo.b_scl1, 2, geo.rc, kb, 64, 0);
    pack(cout, wB, geo.sB, geo.b_nib2, geo.b_scl2, 1, geo.ro, 2, 0, D);
    pack(fc1, wC, geo.sC, 0, geo.c_scl1, 2, geo.rf, kb, 256, 0);
    pack(fc2, wC, geo.sC, geo.c_nib2, geo.c_scl2, 1, geo.ro, 8, 0, D);
    WQ4_CUDA(cudaGetLastError());
    L->wA = wA; L->wB = wB; L->wC = wC;
    return WQ4_OK;
}

void wq4m_dec2_geometry(Dec2Args &a, int D, int F, int H, int T_enc, int CS) { dec2_geometry(a, D, F, H, T_enc, CS); }

static long long *g_dec2_timeline = nullptr;
extern "C" WQ4_API void wq4_debug_decode2_timeline(void *dev_ptr) { g_dec2_timeline = reinterpret_cast<long long *>(dev_ptr); }

int32_t wq4m_decode2_step(const Dec2Args &a_in, int device, cudaStream_t st)
{
    Dec2Args a = a_in;
    a.timeline = g_dec2_timeline;
    const size_t smem = dec2_smem_bytes(a);
    WQ4_CUDA(dec2_set_smem_limit(device, smem));
    WqProfScope prof(WQ4_PROF_GEMV, 0.0, st);
    WQ4_CUDA(cudaMemsetAsync(a.barrier, 0, 2 * sizeof(unsigned), st));
    cudaLaunchConfig_t cfg = {};
    cfg.gridDim = dim3((unsigned)(a.CS * a.H), 1, 1);
    cfg.blockDim = dim3(kT, 1, 1);
    cfg.dynamicSmemBytes = smem;
    cfg.stream = st;
    cudaLaunchAttribute attr[2];
    attr[0].id = cudaLaunchAttributeClusterDimension;
    attr[0].val.clusterDim.x = (unsigned)a.CS; attr[0].val.clusterDim.y = 1; attr[0].val.clusterDim.z = 1;
    attr[1].id = cudaLaunchAttributeCooperative;       // co-residency of all CTAs is required by the grid-wide exchanges
    attr[1].val.cooperative = 1;
    cfg.attrs = attr;
    // Cooperative + cluster launch; if this driver rejects the combination, launch the clusters plainly: wq4m_dec2_plan has
    // checked with cudaOccupancyMaxActiveClusters that all H clusters are co-resident on an otherwise idle device.
    static std::atomic<int> coop_ok[64];          // 0 unknown, 1 yes, -1 no
    const int di = device >= 0 && device < 64 ? device : 0;
    cudaError_t e = cudaErrorNotSupported;
    if (coop_ok[di].load() >= 0) {
        cfg.numAttrs = 2;
        e = cudaLaunchKernelEx(&cfg, decode2_kernel, a);
        if (e == cudaSuccess) coop_ok[di].store(1);
        else if (coop_ok[di].load() == 0) { cudaGetLastError(); coop_ok[di].store(-1); }
    }
    if (e != cudaSuccess && coop_ok[di].load() < 0) {
        cfg.numAttrs = 1;
        e = cudaLaunchKernelEx(&cfg, decode2_kernel, a);
    }
    if (e != cudaSuccess) return wq4_cuda_fail(e, "decode step (v2) launch");
    wq4_count_launch();
    return WQ4_OK;
}
