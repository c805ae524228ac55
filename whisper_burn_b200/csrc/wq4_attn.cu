// wq4_attn.cu -- encoder self-attention (non-causal, head_dim 64) on tcgen05 tensor cores.
//
// Replaces the SDPA of src/model/attention.rs:30-58 for the audio encoder (encoder.rs:86-88): per (clip, head),
// softmax(Q K^T / 8) V over T = 1500 positions, reading the fused f16 QKV rows the Q4 GEMM epilogue wrote
// ([B*T, 3D]: q | k | v) and writing the f16 context rows [B*T, D] the out-projection GEMM consumes.
//
// One CTA = one 128-row query tile of one (clip, head); two CTAs are co-resident per SM (256 TMEM columns and
// ~83 KB of shared memory each), so one CTA's softmax covers the other's MMA/TMEM round trips.
//   warp 0      TMA producer: Q tile once, then K and V tiles of 128 keys through a 2-stage ring
//   warp 1      TMEM allocator + MMA issuer (one elected lane):
//                 S = Q K^T     SS form, M 128 x N 128 x K 64  -> TMEM columns [0,128)
//                 O_j = P V     TS form, A = P from TMEM (f16 pairs, aliasing S), B = V tile MN-major straight from
//                               the TMA box (no transpose pass), M 128 x N 64 x K 128 -> TMEM columns [128,192)
//   warps 2-5   softmax: thread = query row = TMEM lane, so row max / row sum need no shuffles.  Two passes over the
//               S row (max, then exp2 + f16 pack written back over S), running (m, l) and the 64-wide output row in
//               registers: o = o * alpha + O_j per key tile -- the partial product is never rescaled in TMEM.
// Bound: the 16 ex2/clk/SM special-function rate (128 x 128 exponentials per tile against 2 x 256 clk of MMA).
#include <cuda.h>
#include <cuda_fp16.h>

#include <cstdlib>
#include <mutex>

#include <atomic>

#include "wq4_common.cuh"
#include "wq4_model.cuh"
#include "wq4_ptx.cuh"

namespace {

using namespace ptx;

constexpr int kAttThreads = 192;
constexpr int kBQ = 128;            // query rows per CTA
constexpr int kBK = 128;            // keys per tile
constexpr int kHD = 64;             // head dim (attention.rs:30)
constexpr int kTile = kBQ * kHD * 2;            // 16 KiB: 128 rows x 128 bytes
constexpr int kKVStages = 2;
constexpr int kSmemQ = 0;
constexpr int kSmemK = kSmemQ + kTile;
constexpr int kSmemV = kSmemK + kKVStages * kTile;
constexpr int kSmemBar = kSmemV + kKVStages * kTile;
constexpr int kNumBars = 1 + 2 * kKVStages + 3;
constexpr int kSmemMisc = kSmemBar + kNumBars * 8;
constexpr int kSmemTotal = kSmemMisc + 16;
constexpr int kSmemAlloc = kSmemTotal + 1024;
constexpr uint32_t kTmemCols = 256;             // S/P [0,128)  O_j [128,192)
constexpr uint32_t kColO = 128;

struct AttnParams {
    __half *out;        // [B*T, D]
    int T, D;
    float scale_log2;   // log2(e) / sqrt(head_dim)
    int *dev_err;
};

__device__ __forceinline__ void named_bar_sync(int id, int nthreads)
{
    asm volatile("bar.sync %0, %1;" ::"r"(id), "r"(nthreads) : "memory");
}
__device__ __forceinline__ void named_bar_arrive(int id, int nthreads)
{
    asm volatile("bar.arrive %0, %1;" ::"r"(id), "r"(nthreads) : "memory");
}
__device__ __forceinline__ float ex2f(float x)
{
    float y;
    asm("ex2.approx.ftz.f32 %0, %1;" : "=f"(y) : "f"(x));
    return y;
}

__global__ void __launch_bounds__(kAttThreads, 2)
enc_attention_tc_kernel(const __grid_constant__ CUtensorMap tm_qkv, const AttnParams p)
{
    extern __shared__ uint8_t smem_raw_att[];
    const uint32_t base = (smem_u32(smem_raw_att) + 1023u) & ~1023u;
    uint8_t *base_ptr = smem_raw_att + (base - smem_u32(smem_raw_att));
    const uint32_t bars = base + kSmemBar;
    const uint32_t bar_q = bars;
    auto bar_kv_full = [&](int s) { return bars + 8u * (1 + s); };
    auto bar_kv_empty = [&](int s) { return bars + 8u * (1 + kKVStages + s); };
    const uint32_t bar_s = bars + 8u * (1 + 2 * kKVStages);
    const uint32_t bar_p = bar_s + 8u;
    const uint32_t bar_o = bar_s + 16u;
    volatile uint32_t *tmem_slot = reinterpret_cast<volatile uint32_t *>(base_ptr + kSmemMisc);
    volatile int *cta_abort = reinterpret_cast<volatile int *>(base_ptr + kSmemMisc + 4);

    const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
    const int q0 = blockIdx.x * kBQ, h = blockIdx.y, b = blockIdx.z;
    const int nkv = (p.T + kBK - 1) / kBK;

    if (threadIdx.x == 0) {
        *cta_abort = 0;
        mbar_init(bar_q, 1u);
        for (int s = 0; s < kKVStages; s++) { mbar_init(bar_kv_full(s), 1u); mbar_init(bar_kv_empty(s), 1u); }
        mbar_init(bar_s, 1u);
        mbar_init(bar_p, 4u);
        mbar_init(bar_o, 1u);
        fence_mbar_init();
        prefetch_tensormap(&tm_qkv);
    }
    if (warp == 1) {
        tmem_alloc(smem_u32(const_cast<uint32_t *>(tmem_slot)), kTmemCols);
        tmem_relinquish();
    }
    tc_fence_before();
    __syncthreads();
    tc_fence_after();
    const uint32_t tmem_base = *tmem_slot;

    if (warp == 0) {
        // ================================ TMA producer ================================
        if (elect_one()) {
            mbar_arrive_expect_tx(bar_q, kTile);
            tma_load_3d(base + kSmemQ, &tm_qkv, bar_q, h * kHD, q0, b);
        }
        __syncwarp();
        for (int j = 0; j < nkv; j++) {
            const int s = j % kKVStages;
            mbar_wait(bar_kv_empty(s), ((j / kKVStages) & 1u) ^ 1u, cta_abort, p.dev_err, 401);
            if (elect_one()) {
                mbar_arrive_expect_tx(bar_kv_full(s), 2u * kTile);
                tma_load_3d(base + kSmemK + s * kTile, &tm_qkv, bar_kv_full(s), p.D + h * kHD, j * kBK, b);
                tma_load_3d(base + kSmemV + s * kTile, &tm_qkv, bar_kv_full(s), 2 * p.D + h * kHD, j * kBK, b);
            }
            __syncwarp();
        }
    } else if (warp == 1) {
        // ================================ MMA issuer ==================================
        constexpr uint32_t idesc_s = umma_idesc_f16(kBQ, kBK, false);
        constexpr uint32_t idesc_o = umma_idesc_f16_bmn(kBQ, kHD);
        const uint64_t qdesc = umma_smem_desc_sw128(base + kSmemQ);
        mbar_wait(bar_q, 0u, cta_abort, p.dev_err, 402);
        for (int j = 0; j < nkv; j++) {
            const int s = j % kKVStages;
            mbar_wait(bar_kv_full(s), (j / kKVStages) & 1u, cta_abort, p.dev_err, 403);
            tc_fence_after();
            const uint64_t kdesc = umma_smem_desc_sw128(base + kSmemK + s * kTile);
            if (elect_one()) {
#pragma unroll
                for (int k = 0; k < kHD / 16; k++)       // +32 bytes along the 128-byte row per K step
                    umma_f16(tmem_base, qdesc + 2u * k, kdesc + 2u * k, idesc_s, k != 0 ? 1u : 0u);
                umma_commit(bar_s);
            }
            __syncwarp();
            mbar_wait(bar_p, j & 1u, cta_abort, p.dev_err, 404);
            tc_fence_after();
            const uint64_t vdesc = umma_smem_desc_mn_sw128(base + kSmemV + s * kTile);
            if (elect_one()) {
#pragma unroll
                for (int k = 0; k < kBK / 16; k++)       // 16 keys = 8 TMEM columns of P = 2048 bytes of V rows
                    umma_f16_ts(tmem_base + kColO, tmem_base + 8u * k, vdesc + 128u * k, idesc_o, k != 0 ? 1u : 0u);
                umma_commit(bar_o);
                umma_commit(bar_kv_empty(s));
            }
            __syncwarp();
        }
    } else {
        // ================================ softmax (4 warps) ===========================
        const int quarter = warp & 3;                          // TMEM lane quarter this warp may access
        const int row = quarter * 32 + lane;                   // query row inside the tile
        const uint32_t lane_base = tmem_base + ((uint32_t)(quarter * 32) << 16);
        const float c = p.scale_log2;
        float m_run = -INFINITY, l_run = 0.0f;
        float o[kHD];
#pragma unroll
        for (int i = 0; i < kHD; i++) o[i] = 0.0f;
        for (int j = 0; j < nkv; j++) {
            const int kvalid = p.T - j * kBK;                  // < kBK only in the last tile
            const bool ragged = kvalid < kBK;
            mbar_wait(bar_s, j & 1u, cta_abort, p.dev_err, 405);
            tc_fence_after();
            uint32_t v[32];
            // pass 1: row maximum
            float mx = m_run;
#pragma unroll 1
            for (int cc = 0; cc < kBK / 32; cc++) {
                tmem_ld_32x32b_x32(lane_base + cc * 32, v);
                tmem_ld_wait();
                if (ragged) {
#pragma unroll
                    for (int i = 0; i < 32; i++)
                        if (cc * 32 + i >= kvalid) v[i] = 0xff800000u;      // -inf: key beyond the sequence
                }
                float m0 = mx, m1 = -INFINITY, m2 = -INFINITY, m3 = -INFINITY;
#pragma unroll
                for (int i = 0; i < 32; i += 4) {
                    m0 = fmaxf(m0, __uint_as_float(v[i]));
                    m1 = fmaxf(m1, __uint_as_float(v[i + 1]));
                    m2 = fmaxf(m2, __uint_as_float(v[i + 2]));
                    m3 = fmaxf(m3, __uint_as_float(v[i + 3]));
                }
                mx = fmaxf(fmaxf(m0, m1), fmaxf(m2, m3));
            }
            const float mc = mx * c;
            const float alpha = ex2f(m_run * c - mc);          // first tile: exp2(-inf) = 0
            m_run = mx;
            // pass 2: p = exp2(s*c - m*c), row sum, f16 pairs written back over the S columns already consumed
            float sum0 = 0.0f, sum1 = 0.0f;
#pragma unroll 1
            for (int cc = 0; cc < kBK / 32; cc++) {
                tmem_ld_32x32b_x32(lane_base + cc * 32, v);
                tmem_ld_wait();
                if (ragged) {
#pragma unroll
                    for (int i = 0; i < 32; i++)
                        if (cc * 32 + i >= kvalid) v[i] = 0xff800000u;
                }
                uint32_t pk[16];
#pragma unroll
                for (int i = 0; i < 16; i++) {
                    const float p0 = ex2f(fmaf(__uint_as_float(v[2 * i]), c, -mc));
                    const float p1 = ex2f(fmaf(__uint_as_float(v[2 * i + 1]), c, -mc));
                    sum0 += p0;
                    sum1 += p1;
                    const __half2 hh = __floats2half2_rn(p0, p1);
                    pk[i] = *reinterpret_cast<const uint32_t *>(&hh);
                }
                tmem_st_32x32b_x16(lane_base + cc * 16, pk);
            }
            tmem_st_wait();
            tc_fence_before();
            __syncwarp();
            if (elect_one()) mbar_arrive(bar_p);
            l_run = l_run * alpha + (sum0 + sum1);
            // partial product of this key tile
            mbar_wait(bar_o, j & 1u, cta_abort, p.dev_err, 406);
            tc_fence_after();
#pragma unroll
            for (int hc = 0; hc < kHD / 32; hc++) {
                tmem_ld_32x32b_x32(lane_base + kColO + hc * 32, v);
                tmem_ld_wait();
#pragma unroll
                for (int i = 0; i < 32; i++) o[hc * 32 + i] = fmaf(o[hc * 32 + i], alpha, __uint_as_float(v[i]));
            }
            tc_fence_before();
        }
        if (q0 + row < p.T) {
            const float inv = 1.0f / l_run;
            uint4 *dst = reinterpret_cast<uint4 *>(p.out + ((size_t)b * p.T + q0 + row) * p.D + h * kHD);
#pragma unroll
            for (int i = 0; i < kHD / 8; i++) {
                uint4 w;
                __half2 h0 = __floats2half2_rn(o[8 * i] * inv, o[8 * i + 1] * inv);
                __half2 h1 = __floats2half2_rn(o[8 * i + 2] * inv, o[8 * i + 3] * inv);
                __half2 h2 = __floats2half2_rn(o[8 * i + 4] * inv, o[8 * i + 5] * inv);
                __half2 h3 = __floats2half2_rn(o[8 * i + 6] * inv, o[8 * i + 7] * inv);
                w.x = *reinterpret_cast<uint32_t *>(&h0);
                w.y = *reinterpret_cast<uint32_t *>(&h1);
                w.z = *reinterpret_cast<uint32_t *>(&h2);
                w.w = *reinterpret_cast<uint32_t *>(&h3);
                dst[i] = w;
            }
        }
    }

    tc_fence_before();
    __syncthreads();
    if (warp == 1) {
        tc_fence_after();
        tmem_dealloc(tmem_base, kTmemCols);
    }
}

// ---------------------------------------------------------------------------------------------------------
// Second arrangement: one CTA per SM, TWO query tiles (256 rows) per CTA sharing every K/V tile, one softmax
// warpgroup per query tile, all 512 TMEM columns:
//   S0 [0,128)  S1 [128,256)  P0 [256,320)  P1 [320,384)  O0 [384,448)  O1 [448,512)
// P no longer aliases S, so S(j+1) = Q K(j+1)^T is issued the moment a warpgroup has pulled S(j) into registers
// (the whole 128-column row: one pass, max and exp2 from registers), and O accumulates in TMEM across key tiles
// (rescaled lazily), so neither MMA's latency nor a per-tile read of O is on the softmax path.
// Registers: setmaxnreg moves the producer warpgroup's allocation to the two softmax warpgroups (40 / 232).
constexpr int kA2Threads = 384;
constexpr int kA2Stages = 3;
constexpr int kA2SmemQ = 0;                                   // two Q tiles
constexpr int kA2SmemK = kA2SmemQ + 2 * kTile;
constexpr int kA2SmemV = kA2SmemK + kA2Stages * kTile;
constexpr int kA2SmemBar = kA2SmemV + kA2Stages * kTile;
constexpr int kA2NumBars = 1 + 2 * kA2Stages + 8;             // q, kv_full[], kv_empty[], {s_full,s_free,p_full,o_full}[2]
constexpr int kA2SmemMisc = kA2SmemBar + kA2NumBars * 8;
constexpr int kA2SmemAlloc = kA2SmemMisc + 16 + 1024;

__global__ void __launch_bounds__(kA2Threads, 1)
enc_attention_tc2_kernel(const __grid_constant__ CUtensorMap tm_qkv, const AttnParams p)
{
    extern __shared__ uint8_t smem_raw_att[];
    const uint32_t base = (smem_u32(smem_raw_att) + 1023u) & ~1023u;
    uint8_t *base_ptr = smem_raw_att + (base - smem_u32(smem_raw_att));
    const uint32_t bars = base + kA2SmemBar;
    const uint32_t bar_q = bars;
    auto bar_kv_full = [&](int s) { return bars + 8u * (1 + s); };
    auto bar_kv_empty = [&](int s) { return bars + 8u * (1 + kA2Stages + s); };
    const uint32_t bar_t0 = bars + 8u * (1 + 2 * kA2Stages);
    auto bar_s_full = [&](int t) { return bar_t0 + 8u * (4 * t); };
    auto bar_s_free = [&](int t) { return bar_t0 + 8u * (4 * t + 1); };
    auto bar_p_full = [&](int t) { return bar_t0 + 8u * (4 * t + 2); };
    auto bar_o_full = [&](int t) { return bar_t0 + 8u * (4 * t + 3); };
    volatile uint32_t *tmem_slot = reinterpret_cast<volatile uint32_t *>(base_ptr + kA2SmemMisc);
    volatile int *cta_abort = reinterpret_cast<volatile int *>(base_ptr + kA2SmemMisc + 4);

    const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
    const int q0 = blockIdx.x * 2 * kBQ, h = blockIdx.y, b = blockIdx.z;
    const int nkv = (p.T + kBK - 1) / kBK;

    if (threadIdx.x == 0) {
        *cta_abort = 0;
        mbar_init(bar_q, 1u);
        for (int s = 0; s < kA2Stages; s++) { mbar_init(bar_kv_full(s), 1u); mbar_init(bar_kv_empty(s), 1u); }
        for (int t = 0; t < 2; t++) {
            mbar_init(bar_s_full(t), 1u);
            mbar_init(bar_s_free(t), 4u);
            mbar_init(bar_p_full(t), 4u);
            mbar_init(bar_o_full(t), 1u);
        }
        fence_mbar_init();
        prefetch_tensormap(&tm_qkv);
    }
    if (warp == 1) {
        tmem_alloc(smem_u32(const_cast<uint32_t *>(tmem_slot)), 512u);
        tmem_relinquish();
    }
    tc_fence_before();
    __syncthreads();
    tc_fence_after();
    const uint32_t tmem_base = *tmem_slot;
    grid_dep_launch();      // PDL: see wq4_gemm2.cu; everything this kernel reads comes through the TMA producer

    if (warp < 4) {
        setmaxnreg_dec<40>();
        if (warp == 0) {
            // ================================ TMA producer ================================
            grid_dep_wait();
            if (elect_one()) {
                mbar_arrive_expect_tx(bar_q, 2u * kTile);
                tma_load_3d(base + kA2SmemQ, &tm_qkv, bar_q, h * kHD, q0, b);
                tma_load_3d(base + kA2SmemQ + kTile, &tm_qkv, bar_q, h * kHD, q0 + kBQ, b);
            }
            __syncwarp();
            for (int j = 0; j < nkv; j++) {
                const int s = j % kA2Stages;
                mbar_wait(bar_kv_empty(s), ((j / kA2Stages) & 1u) ^ 1u, cta_abort, p.dev_err, 411);
                if (elect_one()) {
                    mbar_arrive_expect_tx(bar_kv_full(s), 2u * kTile);
                    tma_load_3d(base + kA2SmemK + s * kTile, &tm_qkv, bar_kv_full(s), p.D + h * kHD, j * kBK, b);
                    tma_load_3d(base + kA2SmemV + s * kTile, &tm_qkv, bar_kv_full(s), 2 * p.D + h * kHD, j * kBK, b);
                }
                __syncwarp();
            }
        } else if (warp == 1) {
            // ================================ MMA issuer ==================================
            constexpr uint32_t idesc_s = umma_idesc_f16(kBQ, kBK, false);
            constexpr uint32_t idesc_o = umma_idesc_f16_bmn(kBQ, kHD);
            auto issue_s = [&](int t, int j) {
                const uint64_t qdesc = umma_smem_desc_sw128(base + kA2SmemQ + t * kTile);
                const uint64_t kdesc = umma_smem_desc_sw128(base + kA2SmemK + (j % kA2Stages) * kTile);
                if (elect_one()) {
#pragma unroll
                    for (int k = 0; k < kHD / 16; k++)
                        umma_f16(tmem_base + t * 128u, qdesc + 2u * k, kdesc + 2u * k, idesc_s, k != 0 ? 1u : 0u);
                    umma_commit(bar_s_full(t));
                }
                __syncwarp();
            };
            mbar_wait(bar_q, 0u, cta_abort, p.dev_err, 412);
            mbar_wait(bar_kv_full(0), 0u, cta_abort, p.dev_err, 413);
            tc_fence_after();
            issue_s(0, 0);
            issue_s(1, 0);
            for (int j = 0; j < nkv; j++) {
                if (j + 1 < nkv) {
                    const int sn = (j + 1) % kA2Stages;
                    mbar_wait(bar_kv_full(sn), ((j + 1) / kA2Stages) & 1u, cta_abort, p.dev_err, 413);
                    for (int t = 0; t < 2; t++) {
                        mbar_wait(bar_s_free(t), j & 1u, cta_abort, p.dev_err, 414);   // S(j) is in registers
                        tc_fence_after();
                        issue_s(t, j + 1);
                    }
                }
                const int s = j % kA2Stages;
                const uint64_t vdesc = umma_smem_desc_mn_sw128(base + kA2SmemV + s * kTile);
                for (int t = 0; t < 2; t++) {
                    mbar_wait(bar_p_full(t), j & 1u, cta_abort, p.dev_err, 415);
                    tc_fence_after();
                    if (elect_one()) {
#pragma unroll
                        for (int k = 0; k < kBK / 16; k++)
                            umma_f16_ts(tmem_base + 384u + t * 64u, tmem_base + 256u + t * 64u + 8u * k, vdesc + 128u * k,
                                        idesc_o, (j | k) != 0 ? 1u : 0u);
                        umma_commit(bar_o_full(t));
                        if (t == 1) umma_commit(bar_kv_empty(s));
                    }
                    __syncwarp();
                }
            }
        }
    } else {
        // ================================ softmax warpgroups ==========================
        setmaxnreg_inc<232>();
        const int t = (warp >> 2) - 1;                         // query tile of this warpgroup
        const int quarter = warp & 3;
        const int row = quarter * 32 + lane;
        const uint32_t lane_sel = (uint32_t)(quarter * 32) << 16;
        const uint32_t tm_s = tmem_base + lane_sel + t * 128u;
        const uint32_t tm_p = tmem_base + lane_sel + 256u + t * 64u;
        const uint32_t tm_o = tmem_base + lane_sel + 384u + t * 64u;
        const float c = p.scale_log2;
        // O_t accumulates in TMEM over all key tiles (TMEM reads are 64 B/clk/SM: re-reading a 32 KB partial product
        // per tile cost as much as the exponentials).  The exponent offset m_used is only raised -- and O and l rescaled,
        // a TMEM read-modify-write -- when a row's maximum outgrows it by more than 2^8; until then p <= 256, which f16
        // holds, and the final o / l is exact for any common offset.
        float m_used = -INFINITY, l_run = 0.0f;
        for (int j = 0; j < nkv; j++) {
            const int kvalid = p.T - j * kBK;
            uint32_t sreg[kBK];
            mbar_wait(bar_s_full(t), j & 1u, cta_abort, p.dev_err, 416);
            tc_fence_after();
            tmem_ld_32x32b_x64(tm_s, sreg);
            tmem_ld_32x32b_x64(tm_s + 64, sreg + 64);
            tmem_ld_wait();
            tc_fence_before();
            __syncwarp();
            if (elect_one()) mbar_arrive(bar_s_free(t));
            if (kvalid < kBK) {
#pragma unroll
                for (int i = 0; i < kBK; i++)
                    if (i >= kvalid) sreg[i] = 0xff800000u;    // -inf: key beyond the sequence
            }
            float m0 = -INFINITY, m1 = -INFINITY, m2 = -INFINITY, m3 = -INFINITY;
#pragma unroll
            for (int i = 0; i < kBK; i += 4) {
                m0 = fmaxf(m0, __uint_as_float(sreg[i]));
                m1 = fmaxf(m1, __uint_as_float(sreg[i + 1]));
                m2 = fmaxf(m2, __uint_as_float(sreg[i + 2]));
                m3 = fmaxf(m3, __uint_as_float(sreg[i + 3]));
            }
            const float mx = fmaxf(fmaxf(m0, m1), fmaxf(m2, m3));
            const bool grow = (mx - m_used) * c > 8.0f;        // always true on the first tile (m_used = -inf)
            const bool warp_grow = __any_sync(0xffffffffu, grow);
            if (j > 0) {
                // PV(j-1) complete: O holds tiles < j and the P region may be overwritten
                mbar_wait(bar_o_full(t), (j - 1) & 1u, cta_abort, p.dev_err, 417);
                tc_fence_after();
            }
            if (warp_grow) {
                const float m_new = fmaxf(m_used, mx);
                const float alpha = ex2f((m_used - m_new) * c);    // 1 for lanes whose offset stays; 0 on the first tile
                m_used = m_new;
                l_run *= alpha;
                if (j > 0) {
                    uint32_t ov[kHD];
                    tmem_ld_32x32b_x64(tm_o, ov);
                    tmem_ld_wait();
#pragma unroll
                    for (int i = 0; i < kHD; i++) ov[i] = __float_as_uint(__uint_as_float(ov[i]) * alpha);
                    tmem_st_32x32b_x64(tm_o, ov);
                }
            }
            const float mc = m_used * c;
            uint32_t pk[kBK / 2];
            float sum0 = 0.0f, sum1 = 0.0f;
            // The exponential phase is bound by the SFU pipe (128 ex2 + 64 f32->f16x2 packs per row, 16 lanes/clk/SM)
            // and is as long as the rest of a tile (TMEM read at 64 B/clk/SM, max, P store).  The two warpgroups take
            // turns in it so the SFU never idles; left alone they fall into lockstep and share it at half speed each.
            if (t == 0) { if (j > 0) named_bar_sync(2, 256); } else named_bar_sync(1, 256);
#pragma unroll
            for (int i = 0; i < kBK / 2; i++) {
                const float p0 = ex2f(fmaf(__uint_as_float(sreg[2 * i]), c, -mc));
                const float p1 = ex2f(fmaf(__uint_as_float(sreg[2 * i + 1]), c, -mc));
                sum0 += p0;
                sum1 += p1;
                const __half2 hh = __floats2half2_rn(p0, p1);
                pk[i] = *reinterpret_cast<const uint32_t *>(&hh);
            }
            if (t == 0) named_bar_arrive(1, 256); else if (j + 1 < nkv) named_bar_arrive(2, 256);
            l_run += sum0 + sum1;
            tmem_st_32x32b_x64(tm_p, pk);
            tmem_st_wait();
            tc_fence_before();
            __syncwarp();
            if (elect_one()) mbar_arrive(bar_p_full(t));
        }
        float o[kHD];
        {
            mbar_wait(bar_o_full(t), (nkv - 1) & 1u, cta_abort, p.dev_err, 418);
            tc_fence_after();
            uint32_t ov[kHD];
            tmem_ld_32x32b_x64(tm_o, ov);
            tmem_ld_wait();
#pragma unroll
            for (int i = 0; i < kHD; i++) o[i] = __uint_as_float(ov[i]);
        }
        const int qrow = q0 + t * kBQ + row;
        if (qrow < p.T) {
            const float inv = 1.0f / l_run;
            uint4 *dst = reinterpret_cast<uint4 *>(p.out + ((size_t)b * p.T + qrow) * p.D + h * kHD);
#pragma unroll
            for (int i = 0; i < kHD / 8; i++) {
                uint4 w;
                __half2 h0 = __floats2half2_rn(o[8 * i] * inv, o[8 * i + 1] * inv);
                __half2 h1 = __floats2half2_rn(o[8 * i + 2] * inv, o[8 * i + 3] * inv);
                __half2 h2 = __floats2half2_rn(o[8 * i + 4] * inv, o[8 * i + 5] * inv);
                __half2 h3 = __floats2half2_rn(o[8 * i + 6] * inv, o[8 * i + 7] * inv);
                w.x = *reinterpret_cast<uint32_t *>(&h0);
                w.y = *reinterpret_cast<uint32_t *>(&h1);
                w.z = *reinterpret_cast<uint32_t *>(&h2);
                w.w = *reinterpret_cast<uint32_t *>(&h3);
                dst[i] = w;
            }
        }
    }

    tc_fence_before();
    __syncthreads();
    if (warp == 1) {
        tc_fence_after();
        tmem_dealloc(tmem_base, 512u);
    }
}

typedef CUresult (*EncodeTiledFnA)(CUtensorMap *, CUtensorMapDataType, cuuint32_t, void *, const cuuint64_t *,
                                   const cuuint64_t *, const cuuint32_t *, const cuuint32_t *,
                                   CUtensorMapInterleave, CUtensorMapSwizzle, CUtensorMapL2promotion,
                                   CUtensorMapFloatOOBfill);

EncodeTiledFnA get_encode_fn_att()
{
    static EncodeTiledFnA fn = nullptr;
    static std::once_flag once;
    std::call_once(once, [] {
        void *p = nullptr;
        cudaDriverEntryPointQueryResult qres;
        if (cudaGetDriverEntryPoint("cuTensorMapEncodeTiled", &p, cudaEnableDefault, &qres) == cudaSuccess &&
            qres == cudaDriverEntryPointSuccess)
            fn = reinterpret_cast<EncodeTiledFnA>(p);
    });
    return fn;
}

}  // namespace

// qkv16: [B*T, 3D] f16 (q | k | v per row), out16: [B*T, D] f16.
int32_t wq4m_enc_attention_tc(const void *qkv16, void *out16, int64_t B, int64_t T, int64_t D, int64_t H, cudaStream_t st)
{
    if (D != H * kHD) { wq4_set_error("attention: head_dim must be 64 (attention.rs:30), got D=%lld H=%lld", (long long)D, (long long)H); return WQ4_ERR_SHAPE; }
    if (B < 1 || T < 1 || B > 65535) { wq4_set_error("attention: bad shape B=%lld T=%lld", (long long)B, (long long)T); return WQ4_ERR_SHAPE; }
    EncodeTiledFnA fn = get_encode_fn_att();
    if (!fn) { wq4_set_error("cuTensorMapEncodeTiled entry point not available"); return WQ4_ERR_CUDA; }
    int device = 0;
    WQ4_CUDA(cudaGetDevice(&device));
    CUtensorMap tm;
    cuuint64_t dims[3] = {(cuuint64_t)(3 * D), (cuuint64_t)T, (cuuint64_t)B};
    cuuint64_t strides[2] = {(cuuint64_t)(3 * D) * 2, (cuuint64_t)(3 * D) * 2 * (cuuint64_t)T};
    cuuint32_t box[3] = {(cuuint32_t)kHD, (cuuint32_t)kBQ, 1};
    cuuint32_t estr[3] = {1, 1, 1};
    CUresult r = fn(&tm, CU_TENSOR_MAP_DATA_TYPE_FLOAT16, 3, const_cast<void *>(qkv16), dims, strides, box, estr,
                    CU_TENSOR_MAP_INTERLEAVE_NONE, CU_TENSOR_MAP_SWIZZLE_128B, CU_TENSOR_MAP_L2_PROMOTION_L2_256B,
                    CU_TENSOR_MAP_FLOAT_OOB_FILL_NONE);
    if (r != CUDA_SUCCESS) {
        wq4_set_error("attention: cuTensorMapEncodeTiled failed (%d): B %lld T %lld D %lld ptr %p", (int)r, (long long)B,
                      (long long)T, (long long)D, qkv16);
        return WQ4_ERR_CUDA;
    }
    static std::atomic<bool> attr_done[64];
    if (device < 0 || device >= 64 || !attr_done[device]) {
        WQ4_CUDA(cudaFuncSetAttribute(enc_attention_tc_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, kSmemAlloc));
        WQ4_CUDA(cudaFuncSetAttribute(enc_attention_tc2_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, kA2SmemAlloc));
        if (device >= 0 && device < 64) attr_done[device] = true;
    }
    static int impl = -1;
    if (impl < 0) { const char *e = getenv("WQ4_ATTN_IMPL"); impl = (e && atoi(e) == 2) ? 2 : 0; }   // 2 = one query tile per CTA
    AttnParams p = {};
    p.out = reinterpret_cast<__half *>(out16);
    p.T = (int)T;
    p.D = (int)D;
    p.scale_log2 = 1.4426950408889634f / 8.0f;   // log2(e) / sqrt(64)
    p.dev_err = wq4_error_flag_ptr(device);
    if (!p.dev_err) return WQ4_ERR_CUDA;
    WqProfScope prof(WQ4_PROF_ATTN_ENC, 4.0 * (double)B * (double)H * (double)T * (double)T * 64.0, st);
    if (impl == 2) {
        dim3 grid((unsigned)((T + kBQ - 1) / kBQ), (unsigned)H, (unsigned)B);
        enc_attention_tc_kernel<<<grid, kAttThreads, kSmemAlloc, st>>>(tm, p);
    } else {
        cudaLaunchConfig_t cfg = {};
        cfg.gridDim = dim3((unsigned)((T + 2 * kBQ - 1) / (2 * kBQ)), (unsigned)H, (unsigned)B);
        cfg.blockDim = dim3(kA2Threads, 1, 1);
        cfg.dynamicSmemBytes = kA2SmemAlloc;
        cfg.stream = st;
        cudaLaunchAttribute attr[1];
        attr[0].id = cudaLaunchAttributeProgrammaticStreamSerialization;
        attr[0].val.programmaticStreamSerializationAllowed = 1;
        cfg.attrs = attr;
        cfg.numAttrs = wq4_pdl_enabled() ? 1 : 0;
        WQ4_CUDA(cudaLaunchKernelEx(&cfg, enc_attention_tc2_kernel, tm, p));
    }
    wq4_count_launch();
    WQ4_CUDA(cudaGetLastError());
    return WQ4_OK;
}

// Test / microbenchmark hook (not part of the drop-in surface): the encoder attention alone on device buffers.
extern "C" WQ4_API int32_t wq4_debug_enc_attention(const void *qkv16_dev, void *out16_dev, int64_t B, int64_t T, int64_t D,
                                                   int64_t H, void *stream)
{
    return wq4m_enc_attention(qkv16_dev, out16_dev, B, T, D, H, reinterpret_cast<cudaStream_t>(stream));
}
