// wq4_chain.cu -- kernel family (1), encoder form: a CHAIN of Q4_0 GEMMs of one encoder block in ONE persistent launch.
//
// EncoderBlock::forward (src/model/encoder.rs:37-49) is  x += out(attn(qkv(ln1(x)))) ; x += fc2(gelu(fc1(ln2(x)))).
// Launched one GEMM at a time (wq4_gemm2.cu) a block at one clip (rows = 1500) is 2 LayerNorm + 4 GEMM + 1 attention
// launches of 5-25 us each, and ~40 % of the block is launch boundaries: every GEMM is one partial wave of tiles plus
// 4-5 us of pipeline fill / drain, every LayerNorm a separate pass over x (profiles/r01_ncu_launches_medium_b1_summary.txt).
// Here the four GEMMs between two attention kernels
//       out-proj (+x)  ->  [LN2]  fc1 + GELU  ->  fc2 (+x)  ->  [LN1 of the next block]  q|k|v
// run as one launch of the CTA-pair kernel (same mainloop as wq4_gemm2.cu: TMA X tiles, Q4 slabs dequantised into TENSOR
// MEMORY, tcgen05.mma.cta_group::2 with the A operand in TMEM, f32 accumulators in TMEM):
//   * the tile list of all problems is walked by the persistent CTA pairs in one fixed order (problem-major, row-block-major);
//     a tile waits until the row block of its input is complete -- a per-(problem, row block) arrival counter in global
//     memory, released by the producer's epilogue warps (st -> fence -> red) and acquired by the consumer's X producer
//     (ld.acquire -> fence.proxy.async -> TMA).  Tiles only depend on tiles earlier in the list and every pair walks the
//     list in order, so the smallest unfinished tile can always run: no deadlock (all 148 CTAs are co-resident, 1 CTA/SM).
//     While a pair waits, its raw-slab producer and dequantisers keep filling the TMEM weight ring of the waiting tile,
//     so the fill of every GEMM but the first is hidden.
//   * LayerNorm needs whole rows, which no tile owns.  It is folded into the consumer GEMM algebraically:
//         LN(x) . W^T = r * ( (gamma o x) . W^T  -  mu * c )  +  b',      c[n] = sum_k W[n,k] gamma[k],
//                                                                         b'[n] = bias[n] + sum_k W[n,k] beta[k]
//     (mu, r = row mean and 1/sqrt(var + 1e-5), src/model/layers.rs:23-31).  The producer GEMM (out-proj / fc2) writes the
//     f32 residual stream x, the f16 operand gamma o x of the consumer, and per 128-column group (one CTA's columns; its
//     four quarter warps add their 32-column sums in a fixed order through shared memory) the row partials (sum, sum of
//     squares).  In the consumer CTA the otherwise idle TMEM-allocator warp turns the D/128 partials of a tile's rows into
//     (mu, r) -- sum (x-mu)^2 = sum_g [q_g - s_g^2/128] + 128 sum_g (s_g/128 - mu)^2: independent terms, deterministic, no
//     cancellation of a large row mean -- into a double-buffered shared-memory table while the tile's MMAs run, and the
//     epilogue applies r, mu, c, b' to the accumulator.  c and b' are computed once at load with the f32 GEMV kernel.
//     No LayerNorm launch and no normalised copy of x exist inside an encoder block any more.
//   * The epilogue runs on eight warps (two per TMEM lane quarter) in the transposed domain: a 32 x 32 chunk goes through a
//     per-warp shared-memory scratch so that fold, bias, GELU, residual, statistics and stores work on float4 of four
//     consecutive columns (128-byte row segments per store instead of 32 scalar stores per lane).
//   * A problem may cover a K range of its weight (k_stage0, k_stages) and wait, in its EPILOGUE only, for another problem's
//     row block (edep): split-K as two problems on different pairs, the second adding to what the first wrote in place.
// Launches per encoder block: 2 (chain + attention) instead of 7.  Measured numbers and what bounds it: DESIGN.md 4.1b.
//
// Roofline: tensor cores; algorithmic work = sum over the problems of 2*rows*N*K flop per launch.
#include <cuda.h>

#include <atomic>
#include <mutex>

#include "wq4_common.cuh"
#include "wq4_ptx.cuh"

namespace {

using namespace ptx;

constexpr int kThreadsC = 640;                            // 4 producer / statistics warps, 8 dequantisers, 8 epilogue warps
constexpr int kEpiWarps = 8;
constexpr int kBKC = 128;                                // K elements per stage
constexpr int kRawNibC = 4 * WQ4_PANEL_ROWS * 16;        // 8 KiB
constexpr int kRawSclC = 4 * WQ4_PANEL_ROWS * 2;         // 1 KiB
constexpr int kRawStageC = kRawNibC + kRawSclC;
constexpr int kRawStagesC = 8;
constexpr int kAColsC = kBKC / 2;                        // TMEM columns per weight-ring slot
constexpr int kMaxProb = WQ4_CHAIN_MAX;
constexpr int kCtrStride = 32;                           // unsigned words between two counters: one 128-byte line (one L2 slice) each
constexpr int kMaxGroups = 16;                           // 128-column statistic groups per row: D <= 2048

template <int TILE_M>
struct CfgC {
    static_assert(TILE_M == 128 || TILE_M == 192, "TILE_M");
    static constexpr int HM = TILE_M / 2;
    static constexpr int x_box = HM * 64 * 2;
    static constexpr int x_stage = 2 * x_box;
    static constexpr int ACC = 2;
    static constexpr int a_col0 = ACC * TILE_M;
    static constexpr int AST = (512 - a_col0) / kAColsC < 4 ? (512 - a_col0) / kAColsC : 4;
    static constexpr int S = 4;
    static_assert(AST >= 2 && a_col0 + AST * kAColsC <= 512, "TMEM budget");
    static constexpr int x_off = 0;
    static constexpr int raw_off = x_off + S * x_stage;
    static constexpr int tr_off = raw_off + kRawStagesC * kRawStageC;       // 8 warps x 32 x 36 floats: transposition scratch
    static constexpr int stat_off = tr_off + kEpiWarps * 32 * 36 * 4;       // 2 x TILE_M x (mean, rstd)
    static constexpr int sp_off = stat_off + 2 * TILE_M * 8;                // [2 halves][2 parities][4 quarters][32 rows] (sum, sumsq)
    static constexpr int bar_off = sp_off + 2 * 2 * 4 * 32 * 8;
    static constexpr int n_bars = 2 * S + 2 * kRawStagesC + 2 * ACC + 4;   // + stat_full[2] stat_empty[2]
    static constexpr int misc_off = bar_off + n_bars * 8;
    static constexpr int total = misc_off + 16;
    static constexpr int alloc = total + 1024;
    static_assert(alloc <= 227 * 1024, "shared memory budget");
};

struct ChainProb {
    const uint4 *nib;
    const __half *scl;
    const float *bias;        // [n]: plain bias, or b' of the LayerNorm fold; may be null
    const float *cvec;        // [n]: c of the fold; null = no fold
    const float *residual;    // f32 [rows][ld] or null
    float *y32;               // f32 [rows][ld] or null
    __half *y16;              // f16 [rows][ld] or null
    const float *y16_scale;   // [n]: gamma of the consumer's LayerNorm, multiplied into y16; null = 1
    float2 *stats_out;        // [n/128][rows] (sum, sum of squares) or null
    const float2 *stats_in;   // fold: partials of the rows of X
    int stats_groups;
    int kbp, n, npairs, kst, ks0, ld;     // kst K stages starting at stage ks0 (split-K halves of one weight are two problems)
    int dep, dep_count;       // producer problem (-1: the predecessor grid) and arrivals that complete one of its row blocks
    int edep, edep_count;     // problem whose row block must be complete before the EPILOGUE reads the residual (-1 none): the
                              // other K half of a split weight, which has written x + bias + its partial sum in place
    int tile0;                // first tile of this problem in the launch's tile list
    int gelu;
};

struct ChainParams {
    ChainProb prob[kMaxProb];
    int nprob, rows, mtiles, total_tiles;
    unsigned *counters;       // [kMaxProb][mtiles] arrivals, then [1] finished CTAs; all zero between launches
    int *dev_err;
    long long *timeline;      // debug (wq4_debug_chain_timeline): [pair][16 tiles][12] %globaltimer stamps of the leader CTA, normally NULL
};

struct ChainMaps { CUtensorMap m[kMaxProb]; };

__device__ __forceinline__ uint32_t lop3_and_or(uint32_t a, uint32_t b, uint32_t c)
{
    uint32_t d;
    asm("lop3.b32 %0, %1, %2, %3, 0xEA;" : "=r"(d) : "r"(a), "r"(b), "r"(c));
    return d;
}
// One device word (8 nibbles, wq4_common.cuh) -> 4 x half2 in K order, exact (nib - 8) times the block scale.
__device__ __forceinline__ void dequant_word(uint32_t w, uint32_t d2, uint32_t *out)
{
    const __half2 k1032 = __floats2half2_rn(1032.0f, 1032.0f);
    const __half2 k16th = __floats2half2_rn(0.0625f, 0.0625f);
    const __half2 km72 = __floats2half2_rn(-72.0f, -72.0f);
    const __half2 d = *reinterpret_cast<const __half2 *>(&d2);
    uint32_t t0 = lop3_and_or(w, 0x000F000Fu, 0x64006400u);
    uint32_t t1 = lop3_and_or(w, 0x00F000F0u, 0x64006400u);
    uint32_t w8 = w >> 8;
    uint32_t t2 = lop3_and_or(w8, 0x000F000Fu, 0x64006400u);
    uint32_t t3 = lop3_and_or(w8, 0x00F000F0u, 0x64006400u);
    __half2 q0 = __hmul2(__hsub2(*reinterpret_cast<__half2 *>(&t0), k1032), d);
    __half2 q1 = __hmul2(__hfma2(*reinterpret_cast<__half2 *>(&t1), k16th, km72), d);
    __half2 q2 = __hmul2(__hsub2(*reinterpret_cast<__half2 *>(&t2), k1032), d);
    __half2 q3 = __hmul2(__hfma2(*reinterpret_cast<__half2 *>(&t3), k16th, km72), d);
    out[0] = *reinterpret_cast<uint32_t *>(&q0);
    out[1] = *reinterpret_cast<uint32_t *>(&q1);
    out[2] = *reinterpret_cast<uint32_t *>(&q2);
    out[3] = *reinterpret_cast<uint32_t *>(&q3);
}

__device__ __forceinline__ unsigned ld_acquire_gpu_u32(const unsigned *p)
{
    unsigned v;
    asm volatile("ld.acquire.gpu.global.u32 %0, [%1];" : "=r"(v) : "l"(p) : "memory");
    return v;
}
__device__ __forceinline__ unsigned ld_relaxed_gpu_u32(const unsigned *p)
{
    unsigned v;
    asm volatile("ld.relaxed.gpu.global.u32 %0, [%1];" : "=r"(v) : "l"(p) : "memory");
    return v;
}
__device__ __forceinline__ void red_release_gpu_inc(unsigned *p)
{
    asm volatile("red.release.gpu.global.add.u32 [%0], 1;" ::"l"(p) : "memory");
}
__device__ __forceinline__ void fence_proxy_async_all() { asm volatile("fence.proxy.async;" ::: "memory"); }

// Wait until `*ctr >= want` (one lane polls, the warp follows).  Bounded like every other wait of this library.
__device__ __forceinline__ void dep_wait(const unsigned *ctr, unsigned want, int lane, volatile int *cta_abort, int *dev_err, int code)
{
    if (lane == 0) {
        // relaxed polls (an acquiring load invalidates the SM's L1 every time), one acquiring load once the count is there
        unsigned spins = 0;
        while (ld_relaxed_gpu_u32(ctr) < want) {
            __nanosleep(100);
            if (++spins > (1u << 22) || *cta_abort) { *cta_abort = 1; atomicCAS(dev_err, 0, code); break; }
        }
        (void)ld_acquire_gpu_u32(ctr);
    }
    __syncwarp();
}

// The same for a count that is known to be there already: straight to the acquiring load (loops only if it is not).
__device__ __forceinline__ void dep_acquire(const unsigned *ctr, unsigned want, int lane, volatile int *cta_abort, int *dev_err, int code)
{
    if (lane == 0) {
        unsigned spins = 0;
        while (ld_acquire_gpu_u32(ctr) < want) {
            __nanosleep(100);
            if (++spins > (1u << 22) || *cta_abort) { *cta_abort = 1; atomicCAS(dev_err, 0, code); break; }
        }
    }
    __syncwarp();
}

// which problem a tile of the launch's list belongs to
__device__ __forceinline__ int prob_of(const ChainParams &p, int g)
{
    int pi = 0;
#pragma unroll
    for (int i = 1; i < kMaxProb; i++)
        if (i < p.nprob && g >= p.prob[i].tile0) pi = i;
    return pi;
}

// Per-tile constants of one epilogue lane (registers; the problem table itself lives in the kernel-parameter bank).
struct EpiTile {
    float4 bias4, cn4, gs4;
    float *y32;
    __half *y16;
    float2 *stats_row;        // stats_out + (128-column group) * rows, or null
    long long ld;
    int nc;                   // first of the lane's four columns
    bool n_ok;
};

// One 32 x 32 chunk in the transposed domain: lane = rows (i*4 + lane/8), columns nc..nc+3.  Branch-free for a given mode so
// that the eight row groups overlap (loads first, then arithmetic, predicated stores).
template <bool FOLD, bool GELU, bool RES, bool Y32, bool Y16, bool STATS, bool SCALE16 = STATS>
__device__ __forceinline__ void epi_chunk_t(const float *tr, const float2 *stat, const EpiTile &T, const float4 (&r4)[8],
                                            int tb, int nvalid, int lane, float2 *sp, int qq, int bar_id)
{
    const int rsub = lane >> 3, col4 = (lane & 7) * 4;
    float4 a[8];
#pragma unroll
    for (int i = 0; i < 8; i++) a[i] = *reinterpret_cast<const float4 *>(tr + (i * 4 + rsub) * 36 + col4);
    if constexpr (FOLD) {
        float2 st[8];
#pragma unroll
        for (int i = 0; i < 8; i++) st[i] = stat[i * 4 + rsub];
#pragma unroll
        for (int i = 0; i < 8; i++) {
            a[i].x = fmaf(st[i].y, fmaf(-st[i].x, T.cn4.x, a[i].x), T.bias4.x);
            a[i].y = fmaf(st[i].y, fmaf(-st[i].x, T.cn4.y, a[i].y), T.bias4.y);
            a[i].z = fmaf(st[i].y, fmaf(-st[i].x, T.cn4.z, a[i].z), T.bias4.z);
            a[i].w = fmaf(st[i].y, fmaf(-st[i].x, T.cn4.w, a[i].w), T.bias4.w);
        }
    } else {
#pragma unroll
        for (int i = 0; i < 8; i++) { a[i].x += T.bias4.x; a[i].y += T.bias4.y; a[i].z += T.bias4.z; a[i].w += T.bias4.w; }
    }
    if constexpr (GELU) {
#pragma unroll
        for (int i = 0; i < 8; i++) { a[i].x = wq4_gelu_tanh(a[i].x); a[i].y = wq4_gelu_tanh(a[i].y); a[i].z = wq4_gelu_tanh(a[i].z); a[i].w = wq4_gelu_tanh(a[i].w); }
    }
    if constexpr (RES) {
#pragma unroll
        for (int i = 0; i < 8; i++) { a[i].x += r4[i].x; a[i].y += r4[i].y; a[i].z += r4[i].z; a[i].w += r4[i].w; }
    }
#pragma unroll
    for (int i = 0; i < 8; i++) {
        const int rI = i * 4 + rsub;
        const bool ok = T.n_ok && rI < nvalid;
        const long long o = (long long)(tb + rI) * T.ld + T.nc;
        if constexpr (Y32) { if (ok) *reinterpret_cast<float4 *>(T.y32 + o) = a[i]; }
        if constexpr (Y16) {
            // the f16 copy carries the consumer LayerNorm's gamma only in the residual + statistics mode
            const __half2 h0 = SCALE16 ? __floats2half2_rn(a[i].x * T.gs4.x, a[i].y * T.gs4.y) : __floats2half2_rn(a[i].x, a[i].y);
            const __half2 h1 = SCALE16 ? __floats2half2_rn(a[i].z * T.gs4.z, a[i].w * T.gs4.w) : __floats2half2_rn(a[i].z, a[i].w);
            uint2 pk;
            pk.x = *reinterpret_cast<const uint32_t *>(&h0);
            pk.y = *reinterpret_cast<const uint32_t *>(&h1);
            if (ok) *reinterpret_cast<uint2 *>(T.y16 + o) = pk;
        }
    }
    if constexpr (STATS) {
        // (sum, sum of squares) of each row over this warp's 32 columns (the 8 lanes that share the row), then over the
        // CTA's 128 columns: the four warps of this half exchange their values through shared memory (buffer `sp`, double
        // buffered by chunk parity: one named barrier per chunk) and quarter 0 adds them in a fixed order.
        float s[8], sq[8];
#pragma unroll
        for (int i = 0; i < 8; i++) {
            s[i] = T.n_ok ? (a[i].x + a[i].y) + (a[i].z + a[i].w) : 0.0f;
            sq[i] = T.n_ok ? fmaf(a[i].x, a[i].x, a[i].y * a[i].y) + fmaf(a[i].z, a[i].z, a[i].w * a[i].w) : 0.0f;
        }
#pragma unroll
        for (int off = 1; off < 8; off <<= 1) {
#pragma unroll
            for (int i = 0; i < 8; i++) {
                s[i] += __shfl_xor_sync(0xffffffffu, s[i], off);
                sq[i] += __shfl_xor_sync(0xffffffffu, sq[i], off);
            }
        }
        if ((lane & 7) == 0) {
#pragma unroll
            for (int i = 0; i < 8; i++) sp[qq * 32 + i * 4 + rsub] = make_float2(s[i], sq[i]);
        }
        asm volatile("bar.sync %0, 128;" ::"r"(bar_id) : "memory");
        if (qq == 0 && T.stats_row && lane < nvalid) {
            const float2 p0 = sp[lane], p1 = sp[32 + lane], p2 = sp[64 + lane], p3 = sp[96 + lane];
            T.stats_row[tb + lane] = make_float2((p0.x + p1.x) + (p2.x + p3.x), (p0.y + p1.y) + (p2.y + p3.y));
        }
    }
}

template <int TILE_M>
__global__ void __launch_bounds__(kThreadsC, 1)
q4_chain_kernel(const __grid_constant__ ChainMaps maps, const __grid_constant__ ChainParams p)
{
    using L = CfgC<TILE_M>;
    constexpr int S = L::S, RS = kRawStagesC, ACC = L::ACC;
    extern __shared__ uint8_t smem_rawc[];
    const uint32_t base = (smem_u32(smem_rawc) + 1023u) & ~1023u;
    uint8_t *base_ptr = smem_rawc + (base - smem_u32(smem_rawc));

    const uint32_t x_smem = base + L::x_off;
    const uint32_t raw_smem = base + L::raw_off;
    const uint32_t bars = base + L::bar_off;
    auto bar_full = [&](int s) { return bars + 8u * s; };
    auto bar_empty = [&](int s) { return bars + 8u * (S + s); };
    auto bar_raw_full = [&](int s) { return bars + 8u * (2 * S + s); };
    auto bar_raw_empty = [&](int s) { return bars + 8u * (2 * S + RS + s); };
    auto bar_acc_full = [&](int b) { return bars + 8u * (2 * S + 2 * RS + b); };
    auto bar_acc_empty = [&](int b) { return bars + 8u * (2 * S + 2 * RS + ACC + b); };
    auto bar_stat_full = [&](int b) { return bars + 8u * (2 * S + 2 * RS + 2 * ACC + b); };        // 32 lanes of the statistics warp
    auto bar_stat_empty = [&](int b) { return bars + 8u * (2 * S + 2 * RS + 2 * ACC + 2 + b); };   // 4 epilogue warps
    volatile uint32_t *tmem_base_slot = reinterpret_cast<volatile uint32_t *>(base_ptr + L::misc_off);
    volatile int *cta_abort = reinterpret_cast<volatile int *>(base_ptr + L::misc_off + 4);

    const int warp = threadIdx.x >> 5;
    const int lane = threadIdx.x & 31;
    const int crank = (int)cluster_ctarank();
    const int num_tiles = p.total_tiles;
    const int slot = (int)blockIdx.x >> 1;
    const int nslots = (int)gridDim.x >> 1;

    if (threadIdx.x == 0) {
        *cta_abort = 0;
        for (int s = 0; s < S; s++) { mbar_init(bar_full(s), 17u); mbar_init(bar_empty(s), 1u); }
        for (int s = 0; s < RS; s++) { mbar_init(bar_raw_full(s), 1u); mbar_init(bar_raw_empty(s), 8u); }
        for (int b = 0; b < ACC; b++) { mbar_init(bar_acc_full(b), 1u); mbar_init(bar_acc_empty(b), 2u * kEpiWarps); }
        for (int b = 0; b < 2; b++) { mbar_init(bar_stat_full(b), 32u); mbar_init(bar_stat_empty(b), (uint32_t)kEpiWarps); }
        fence_mbar_init();
        for (int i = 0; i < p.nprob; i++) prefetch_tensormap(&maps.m[i]);
    }
    if (warp == 3) {
        tmem_alloc2(smem_u32(const_cast<uint32_t *>(tmem_base_slot)), 512u);
        tmem_relinquish2();
    }
    tc_fence_before();
    __syncthreads();
    cluster_sync_all();
    tc_fence_after();
    const uint32_t tmem_base = *tmem_base_slot;
    grid_dep_launch();
    long long *const tl = (p.timeline && crank == 0) ? p.timeline + (size_t)slot * 16 * 12 : nullptr;
#define CH_TL(it_, k_)                                                                                           \
    do {                                                                                                         \
        if (tl && (it_) < 16u && lane == 0) { long long t_; asm volatile("mov.u64 %0, %%globaltimer;" : "=l"(t_)); tl[(it_) * 12 + (k_)] = t_; } \
    } while (0)

    if (warp < 12) setmaxnreg_dec<72>();     // producers / statistics / dequantisers: the epilogue warpgroups take the rest (128)
    if (warp == 0) {
        // ================================ X producer (both CTAs) ================================
        grid_dep_wait();        // also orders every counter poll below after the previous chain launch's counter reset
        const uint32_t full_leader0 = mapa_u32(bar_full(0), 0u);
        uint32_t c = 0, xit = 0;
        for (int g = slot; g < num_tiles; g += nslots, xit++) {
            const int pi = prob_of(p, g);
            const ChainProb &P = p.prob[pi];
            const int rb = (g - P.tile0) / P.npairs;
            const int t0 = rb * TILE_M + crank * L::HM;
            CH_TL(xit, 0);
            if (P.dep >= 0) {
                dep_wait(p.counters + (size_t)(P.dep * p.mtiles + rb) * kCtrStride, (unsigned)P.dep_count, lane, cta_abort, p.dev_err, 211);
                fence_proxy_async_all();      // the producer's generic-proxy stores are read by the async proxy (TMA) below
            }
            CH_TL(xit, 1);
            const CUtensorMap *tm = &maps.m[pi];
            for (int ks = 0; ks < P.kst; ks++, c++) {
                const int s = c % S;
                mbar_wait(bar_empty(s), ((c / S) & 1u) ^ 1u, cta_abort, p.dev_err, 201);
                if (elect_one()) {
                    if (crank == 0) mbar_arrive_expect_tx(bar_full(s), 2u * L::x_stage);
                    const uint32_t dst = x_smem + s * L::x_stage;
                    tma_load_3d_2cta(dst, tm, full_leader0 + 8u * s, (P.ks0 + ks) * kBKC, t0, 0);
                    tma_load_3d_2cta(dst + L::x_box, tm, full_leader0 + 8u * s, (P.ks0 + ks) * kBKC + 64, t0, 0);
                }
                __syncwarp();
            }
            CH_TL(xit, 2);
        }
    } else if (warp == 1) {
        // ================================ MMA issuer (leader only) ==============================
        if (crank == 0) {
            constexpr uint32_t idesc = umma_idesc_f16(256, TILE_M, false);
            uint32_t c = 0, it = 0;
            uint32_t full_ok = 0u;
            for (int g = slot; g < num_tiles; g += nslots, it++) {
                const int kst = p.prob[prob_of(p, g)].kst;
                const uint32_t ab = it % ACC;
                mbar_wait(bar_acc_empty(ab), ((it / ACC) & 1u) ^ 1u, cta_abort, p.dev_err, 202);
                tc_fence_after();
                CH_TL(it, 3);
                const uint32_t tmem_d = tmem_base + ab * TILE_M;
                for (int ks = 0; ks < kst; ks++, c++) {
                    const int s = c % S;
                    if (!full_ok) mbar_wait(bar_full(s), (c / S) & 1u, cta_abort, p.dev_err, 203);
                    tc_fence_after();
                    full_ok = mbar_test_wait(bar_full((c + 1) % S), ((c + 1) / S) & 1u);
                    const uint64_t bdesc0 = umma_smem_desc_sw128(x_smem + s * L::x_stage);
                    const uint64_t bdesc1 = umma_smem_desc_sw128(x_smem + s * L::x_stage + L::x_box);
                    const uint32_t a_tmem = tmem_base + L::a_col0 + (c % L::AST) * kAColsC;
                    if (elect_one()) {
#pragma unroll
                        for (int k = 0; k < 4; k++)
                            umma_f16_ts_2cta(tmem_d, a_tmem + 8u * k, bdesc0 + 2u * k, idesc, (ks | k) != 0 ? 1u : 0u);
#pragma unroll
                        for (int k = 0; k < 4; k++)
                            umma_f16_ts_2cta(tmem_d, a_tmem + 32u + 8u * k, bdesc1 + 2u * k, idesc, 1u);
                        umma_commit2_mcast(bar_empty(s), (uint16_t)0x3);
                        if (ks == kst - 1) umma_commit2_mcast(bar_acc_full(ab), (uint16_t)0x3);
                    }
                    __syncwarp();
                }
                CH_TL(it, 4);
            }
        }
    } else if (warp == 2) {
        // ================================ raw Q4 slab producer (both CTAs) ======================
        // Weights do not depend on any other tile: this role (and the dequantisers) run ahead of the dependency waits.
        uint32_t c = 0;
        for (int g = slot; g < num_tiles; g += nslots) {
            const ChainProb &P = p.prob[prob_of(p, g)];
            const int panel = ((g - P.tile0) % P.npairs) * 2 + crank;
            const size_t panel_entry = (size_t)panel * P.kbp * WQ4_PANEL_ROWS;
            for (int ks = 0; ks < P.kst; ks++, c++) {
                const int s = c % RS;
                mbar_wait(bar_raw_empty(s), ((c / RS) & 1u) ^ 1u, cta_abort, p.dev_err, 205);
                if (elect_one()) {
                    mbar_arrive_expect_tx(bar_raw_full(s), kRawStageC);
                    const size_t e = panel_entry + (size_t)(4 * (P.ks0 + ks)) * WQ4_PANEL_ROWS;
                    const uint32_t dst = raw_smem + s * kRawStageC;
                    bulk_load(dst, P.nib + e, kRawNibC, bar_raw_full(s));
                    bulk_load(dst + kRawNibC, P.scl + e, kRawSclC, bar_raw_full(s));
                }
                __syncwarp();
            }
        }
    } else if (warp >= 4 && warp < 12) {
        // ================================ dequantisers (both CTAs) ==============================
        const int dw = warp - 4;
        const int q = dw & 3;
        const int half = dw >> 2;
        const int r = q * 32 + lane;
        const uint32_t my_tmem = tmem_base + L::a_col0 + half * 32 + ((uint32_t)(q * 32) << 16);
        const uint32_t full_leader0 = mapa_u32(bar_full(0), 0u);
        uint32_t total = 0;
        for (int g = slot; g < num_tiles; g += nslots) total += (uint32_t)p.prob[prob_of(p, g)].kst;
        uint32_t raw_ok = 0u;
        for (uint32_t c = 0; c < total; c++) {
            const int rs = c % RS;
            if (!raw_ok) mbar_wait(bar_raw_full(rs), (c / RS) & 1u, cta_abort, p.dev_err, 206);
            const uint8_t *raw = base_ptr + L::raw_off + rs * kRawStageC;
            const int e0 = (2 * half) * WQ4_PANEL_ROWS + r, e1 = e0 + WQ4_PANEL_ROWS;
            const uint4 q0 = *reinterpret_cast<const uint4 *>(raw + e0 * 16);
            const uint4 q1 = *reinterpret_cast<const uint4 *>(raw + e1 * 16);
            const uint32_t s0 = *reinterpret_cast<const uint16_t *>(raw + kRawNibC + e0 * 2);
            const uint32_t s1 = *reinterpret_cast<const uint16_t *>(raw + kRawNibC + e1 * 2);
            const uint32_t d0 = s0 | (s0 << 16), d1 = s1 | (s1 << 16);
            raw_ok = (c + 1 < total) ? mbar_test_wait(bar_raw_full((c + 1) % RS), ((c + 1) / RS) & 1u) : 1u;
            uint32_t slot_ok = 1u;
            const uint32_t pc = c - L::AST;
            if (c >= (uint32_t)L::AST) slot_ok = mbar_test_wait(bar_empty(pc % S), (pc / S) & 1u);
            uint32_t v[32];
            dequant_word(q0.x, d0, v + 0);
            dequant_word(q0.y, d0, v + 4);
            dequant_word(q0.z, d0, v + 8);
            dequant_word(q0.w, d0, v + 12);
            dequant_word(q1.x, d1, v + 16);
            dequant_word(q1.y, d1, v + 20);
            dequant_word(q1.z, d1, v + 24);
            dequant_word(q1.w, d1, v + 28);
            __syncwarp();
            if (elect_one()) mbar_arrive(bar_raw_empty(rs));
            const int s = c % S;
            if (!slot_ok) mbar_wait(bar_empty(pc % S), (pc / S) & 1u, cta_abort, p.dev_err, 207);
            __syncwarp();
            tc_fence_after();
            tmem_st_32x32b_x32(my_tmem + (c % L::AST) * kAColsC, v);
            tmem_st_wait();
            tc_fence_before();
            __syncwarp();
            if (elect_one()) mbar_arrive_cluster(full_leader0 + 8u * s);
        }
    } else if (warp == 3) {
        // ================================ row statistics of the folded LayerNorm (both CTAs) ====
        // For every tile whose epilogue folds a LayerNorm: (mean, 1/sqrt(var + eps)) of the tile's rows from the producer's
        // per-32-column partials, combined in a fixed order (Chan's update: deterministic, no cancellation of a large
        // row mean), into one of two shared-memory buffers the epilogue warps consume.  Runs ahead of the epilogue.
        grid_dep_wait();
        float2 *stat_all = reinterpret_cast<float2 *>(base_ptr + L::stat_off);
        uint32_t ft = 0, tile_it = 0;
        for (int g = slot; g < num_tiles; g += nslots, tile_it++) {
            const ChainProb &P = p.prob[prob_of(p, g)];
            if (!P.cvec) continue;
            const int rb = (g - P.tile0) / P.npairs;
            const int t0 = rb * TILE_M;
            const uint32_t buf = ft & 1u;
            mbar_wait(bar_stat_empty(buf), ((ft >> 1) & 1u) ^ 1u, cta_abort, p.dev_err, 213);
            if (P.dep >= 0) dep_wait(p.counters + (size_t)(P.dep * p.mtiles + rb) * kCtrStride, (unsigned)P.dep_count, lane, cta_abort, p.dev_err, 214);
            const int G = P.stats_groups;
            // tiles of this role are numbered like the epilogue's: count every tile
            CH_TL(tile_it, 10);
            const float inv_d = 1.0f / (128.0f * (float)G);
            for (int rr = lane; rr < TILE_M; rr += 32) {
                const int row = t0 + rr;
                float mean = 0.0f, var = 0.0f;
                if (row < p.rows) {
                    // all partials of the row in registers (one round trip); mean first, then the exact decomposition
                    //   sum (x - mean)^2 = sum_g [ q_g - s_g^2/128 ] + 128 * sum_g (s_g/128 - mean)^2
                    // whose terms are independent and free of the cancellation of a large row mean
                    const float2 *sp = P.stats_in + row;
                    float2 pp[kMaxGroups];
#pragma unroll
                    for (int u = 0; u < kMaxGroups; u++)
                        pp[u] = (u < G) ? __ldcg(sp + (size_t)u * p.rows) : make_float2(0.0f, 0.0f);
                    float s0 = 0.0f, s1 = 0.0f;
#pragma unroll
                    for (int u = 0; u < kMaxGroups; u += 2) { s0 += pp[u].x; s1 += pp[u + 1].x; }
                    mean = (s0 + s1) * inv_d;
                    float m0 = 0.0f, m1 = 0.0f;
#pragma unroll
                    for (int u = 0; u < kMaxGroups; u++) {
                        const float mg = pp[u].x * (1.0f / 128.0f);
                        const float within = fmaxf(fmaf(-pp[u].x, mg, pp[u].y), 0.0f);
                        const float d = mg - mean;
                        const float t = (u < G) ? fmaf(128.0f * d, d, within) : 0.0f;
                        if (u & 1) m1 += t; else m0 += t;
                    }
                    var = (m0 + m1) * inv_d;
                }
                stat_all[buf * TILE_M + rr] = make_float2(mean, 1.0f / sqrtf(var + 1e-5f));
            }
            CH_TL(tile_it, 9);
            mbar_arrive(bar_stat_full(buf));          // all 32 lanes arrive (count 32): each releases its own writes
            ft++;
        }
    } else if (warp >= 12) {
        // ================================ epilogue (both CTAs): 8 warps ==========================
        // Two warps per TMEM lane quarter (a warp may only read the lanes 32*(warp % 4)..+31): half `eh` takes the 32-row
        // chunks eh, eh + 2, ... of the tile.
        // tcgen05.ld delivers lane = output column n, register j = activation row; a 32 x 32 chunk is transposed through
        // this warp's shared-memory scratch (pitch 36 floats: the column-wise STS.32 and the row-wise LDS.128 are both
        // conflict-free) so that everything after it -- fold, bias, GELU, residual, statistics, stores -- works on float4
        // of four consecutive columns: 128-bit residual loads and stores of 128-byte row segments, 4 rows per instruction.
        setmaxnreg_inc<128>();
        const int q = warp & 3;
        const int eh = (warp - 12) >> 2;
        float *tr = reinterpret_cast<float *>(base_ptr + L::tr_off) + (warp - 12) * (32 * 36);
        const float2 *stat_all = reinterpret_cast<const float2 *>(base_ptr + L::stat_off);
        float2 *sp_all = reinterpret_cast<float2 *>(base_ptr + L::sp_off) + eh * (2 * 4 * 32);
        const uint32_t acc_empty_leader0 = mapa_u32(bar_acc_empty(0), 0u);
        const int rsub = lane >> 3;                           // row inside a group of four
        const int col4 = (lane & 7) * 4;                      // first of this lane's four columns inside the warp's 32
        grid_dep_wait();
        uint32_t it = 0, ft = 0;
        for (int g = slot; g < num_tiles; g += nslots, it++) {
            const int pi = prob_of(p, g);
            const ChainProb &P = p.prob[pi];
            const int local = g - P.tile0;
            const int rb = local / P.npairs;
            const int panel = (local % P.npairs) * 2 + crank;
            const int t0 = rb * TILE_M;
            const uint32_t ab = it % ACC;
            const int n0 = panel * WQ4_PANEL_ROWS + q * 32;   // first column of this warp
            EpiTile T;
            T.nc = n0 + col4;                                 // N % 4 == 0: the lane's four columns are all valid or all not
            T.n_ok = T.nc < P.n;
            const float4 zero4 = make_float4(0.0f, 0.0f, 0.0f, 0.0f);
            const bool fold = P.cvec != nullptr;
            T.bias4 = (P.bias && T.n_ok) ? __ldg(reinterpret_cast<const float4 *>(P.bias + T.nc)) : zero4;
            T.cn4 = (fold && T.n_ok) ? __ldg(reinterpret_cast<const float4 *>(P.cvec + T.nc)) : zero4;
            T.gs4 = (P.y16_scale && T.n_ok) ? __ldg(reinterpret_cast<const float4 *>(P.y16_scale + T.nc)) : make_float4(1.0f, 1.0f, 1.0f, 1.0f);
            T.y32 = P.y32; T.y16 = P.y16; T.ld = P.ld;
            T.stats_row = P.stats_out ? P.stats_out + (size_t)panel * p.rows : nullptr;
            const float *res = P.residual;
            const int rows = p.rows;
            // epilogue mode of the problem (warp-uniform): 0 residual + f32 + scaled f16 + statistics (out-proj, fc2),
            // 1 residual + f32 (last fc2), 2 fold + GELU + f16 (fc1), 3 fold + f16 (q|k|v), 4 f16, 6 fold + f32, 5 generic
            int mode = 5;
            if (!fold && !P.gelu && res && P.y32 && P.y16 && P.stats_out) mode = 0;
            else if (!fold && !P.gelu && res && P.y32 && !P.y16 && !P.stats_out) mode = 1;
            else if (fold && P.gelu && !res && !P.y32 && P.y16 && !P.stats_out && !P.y16_scale) mode = 2;
            else if (fold && !P.gelu && !res && !P.y32 && P.y16 && !P.stats_out && !P.y16_scale) mode = 3;
            else if (!fold && !P.gelu && !res && !P.y32 && P.y16 && !P.stats_out && !P.y16_scale) mode = 4;
            else if (fold && !P.gelu && !res && P.y32 && !P.y16 && !P.stats_out) mode = 6;      // decoder q|k|v / cross-q: f32 for the attention kernel
            // the producer's stores (residual stream) are acquired by this warp itself, not only through the
            // X producer -> TMA -> MMA -> accumulator barrier chain
            float4 r4[8];
            auto load_res = [&](int c0) {
                const int tb = t0 + c0;
#pragma unroll
                for (int i = 0; i < 8; i++) {
                    const int row = tb + i * 4 + rsub;
                    r4[i] = (T.n_ok && row < rows) ? __ldcg(reinterpret_cast<const float4 *>(res + (long long)row * T.ld + T.nc)) : zero4;
                }
            };
            // A tile with a producer: its X producer warp has seen the row block complete (that is why the accumulator is
            // ready); one acquiring load of the same counter makes the producer's residual-stream stores visible to THIS warp
            // (no polling loop here: 16 polling warps per pair tile would saturate the counters' L2 slices)
            const bool early_res = res && P.dep < 0 && P.edep < 0;
            if (early_res) load_res(eh * 32);
            const float2 *stat = stat_all + (ft & 1u) * TILE_M;
            // This warp's own acquire of the producer's stores (one acquiring load of a counter that is already complete: no
            // polling).  A folding tile takes it as soon as its statistics are ready -- the statistics warp has seen the row
            // block complete -- so the L2 round trip runs under the tile's MMAs; other tiles take it once the accumulator is
            // ready (their X producer has seen the row block complete).
            bool acquired = P.dep < 0;
            if (fold) {
                mbar_wait(bar_stat_full(ft & 1u), (ft >> 1) & 1u, cta_abort, p.dev_err, 215);
                if (P.dep >= 0) { dep_acquire(p.counters + (size_t)(P.dep * p.mtiles + rb) * kCtrStride, (unsigned)P.dep_count, lane, cta_abort, p.dev_err, 212); acquired = true; }
            }
            if (warp == 12) CH_TL(it, 5);
            mbar_wait(bar_acc_full(ab), (it / ACC) & 1u, cta_abort, p.dev_err, 208);
            tc_fence_after();
            if (!acquired) dep_acquire(p.counters + (size_t)(P.dep * p.mtiles + rb) * kCtrStride, (unsigned)P.dep_count, lane, cta_abort, p.dev_err, 212);
            if (P.edep >= 0) dep_wait(p.counters + (size_t)(P.edep * p.mtiles + rb) * kCtrStride, (unsigned)P.edep_count, lane, cta_abort, p.dev_err, 216);
            if (res && !early_res) load_res(eh * 32);
            if (warp == 12) CH_TL(it, 6);
            uint32_t chunk_i = 0;
#pragma unroll 1
            for (int c0 = eh * 32; c0 < TILE_M; c0 += 64, chunk_i++) {
                const int tb = t0 + c0;
                if (tb >= rows) break;                     // uniform over the four warps of this half
                float2 *sp = sp_all + (chunk_i & 1u) * (4 * 32);
                {
                    uint32_t v[32];
                    tmem_ld_32x32b_x32(tmem_base + ab * TILE_M + c0 + ((uint32_t)(q * 32) << 16), v);
                    tmem_ld_wait();
#pragma unroll
                    for (int j = 0; j < 32; j++) tr[j * 36 + lane] = __uint_as_float(v[j]);
                }
                __syncwarp();
                const int nvalid = min(32, rows - tb);
                switch (mode) {
                case 0: epi_chunk_t<false, false, true, true, true, true>(tr, stat + c0, T, r4, tb, nvalid, lane, sp, q, 1 + eh); break;
                case 1: epi_chunk_t<false, false, true, true, false, false>(tr, stat + c0, T, r4, tb, nvalid, lane, sp, q, 1 + eh); break;
                case 2: epi_chunk_t<true, true, false, false, true, false>(tr, stat + c0, T, r4, tb, nvalid, lane, sp, q, 1 + eh); break;
                case 3: epi_chunk_t<true, false, false, false, true, false>(tr, stat + c0, T, r4, tb, nvalid, lane, sp, q, 1 + eh); break;
                case 4: epi_chunk_t<false, false, false, false, true, false>(tr, stat + c0, T, r4, tb, nvalid, lane, sp, q, 1 + eh); break;
                case 6: epi_chunk_t<true, false, false, true, false, false>(tr, stat + c0, T, r4, tb, nvalid, lane, sp, q, 1 + eh); break;
                default: {
                    // any other combination: the general form, one row group at a time
                    const bool f_gelu = P.gelu != 0;
#pragma unroll 1
                    for (int i = 0; i < 8; i++) {
                        const int rI = i * 4 + rsub;
                        float4 a = *reinterpret_cast<const float4 *>(tr + rI * 36 + col4);
                        if (fold) {
                            const float2 st = stat[c0 + rI];
                            a.x = fmaf(st.y, fmaf(-st.x, T.cn4.x, a.x), T.bias4.x); a.y = fmaf(st.y, fmaf(-st.x, T.cn4.y, a.y), T.bias4.y);
                            a.z = fmaf(st.y, fmaf(-st.x, T.cn4.z, a.z), T.bias4.z); a.w = fmaf(st.y, fmaf(-st.x, T.cn4.w, a.w), T.bias4.w);
                        } else { a.x += T.bias4.x; a.y += T.bias4.y; a.z += T.bias4.z; a.w += T.bias4.w; }
                        if (f_gelu) { a.x = wq4_gelu_tanh(a.x); a.y = wq4_gelu_tanh(a.y); a.z = wq4_gelu_tanh(a.z); a.w = wq4_gelu_tanh(a.w); }
                        if (res) {
                            const float4 rr = (T.n_ok && tb + rI < rows) ? __ldcg(reinterpret_cast<const float4 *>(res + (long long)(tb + rI) * T.ld + T.nc)) : zero4;
                            a.x += rr.x; a.y += rr.y; a.z += rr.z; a.w += rr.w;
                        }
                        const bool ok = T.n_ok && rI < nvalid;
                        const long long o = (long long)(tb + rI) * T.ld + T.nc;
                        if (T.y32 && ok) *reinterpret_cast<float4 *>(T.y32 + o) = a;
                        if (T.y16 && ok) {
                            const __half2 h0 = __floats2half2_rn(a.x * T.gs4.x, a.y * T.gs4.y), h1 = __floats2half2_rn(a.z * T.gs4.z, a.w * T.gs4.w);
                            uint2 pk;
                            pk.x = *reinterpret_cast<const uint32_t *>(&h0);
                            pk.y = *reinterpret_cast<const uint32_t *>(&h1);
                            *reinterpret_cast<uint2 *>(T.y16 + o) = pk;
                        }
                    }
                } break;
                }
                if (res && mode != 5 && c0 + 64 < TILE_M && tb + 64 < rows) load_res(c0 + 64);
                __syncwarp();                              // the scratch is free for the next chunk
            }
            if (warp == 12) CH_TL(it, 8);
            tc_fence_before();
            __syncwarp();
            if (elect_one()) mbar_arrive_cluster(acc_empty_leader0 + 8u * ab);
            if (fold) {
                if (lane == 0) mbar_arrive(bar_stat_empty(ft & 1u));
                ft++;
            }
            // this warp's part of row block rb of problem pi is in global memory: release it to the consumers
            if (lane == 0) red_release_gpu_inc(p.counters + (size_t)(pi * p.mtiles + rb) * kCtrStride);     // release: orders this warp's stores (via __syncwarp) before it
            if (warp == 12) CH_TL(it, 7);
        }
    }

    tc_fence_before();
    __syncthreads();
    cluster_sync_all();
    if (warp == 3) {
        tc_fence_after();
        tmem_dealloc2(tmem_base, 512u);
    }
    // The last CTA to get here zeroes the counters for the next launch: every other CTA has finished all its polls.
    if (threadIdx.x == 0) {
        unsigned *done = p.counters + (size_t)kMaxProb * p.mtiles * kCtrStride;
        __threadfence();
        const unsigned prev = atomicAdd(done, 1u);
        if (prev == gridDim.x - 1) {
            for (int i = 0; i < kMaxProb * p.mtiles; i++) p.counters[(size_t)i * kCtrStride] = 0u;
            *done = 0u;
            __threadfence();
        }
    }
}

typedef CUresult (*EncodeTiledFnC)(CUtensorMap *, CUtensorMapDataType, cuuint32_t, void *, const cuuint64_t *,
                                   const cuuint64_t *, const cuuint32_t *, const cuuint32_t *,
                                   CUtensorMapInterleave, CUtensorMapSwizzle, CUtensorMapL2promotion,
                                   CUtensorMapFloatOOBfill);

EncodeTiledFnC get_encode_fn_c()
{
    static EncodeTiledFnC fn = nullptr;
    static std::once_flag once;
    std::call_once(once, [] {
        void *p = nullptr;
        cudaDriverEntryPointQueryResult qres;
        if (cudaGetDriverEntryPoint("cuTensorMapEncodeTiled", &p, cudaEnableDefault, &qres) == cudaSuccess &&
            qres == cudaDriverEntryPointSuccess)
            fn = reinterpret_cast<EncodeTiledFnC>(p);
    });
    return fn;
}

template <int TILE_M>
int32_t launch_chain(const ChainMaps &maps, const ChainParams &p, int device, cudaStream_t stream)
{
    using L = CfgC<TILE_M>;
    auto kern = q4_chain_kernel<TILE_M>;
    static std::atomic<bool> attr_done[64];
    if (device < 0 || device >= 64 || !attr_done[device]) {
        WQ4_CUDA(cudaFuncSetAttribute(kern, cudaFuncAttributeMaxDynamicSharedMemorySize, L::alloc));
        if (device >= 0 && device < 64) attr_done[device] = true;
    }
    int pairs = wq4_sm_count(device) / 2;
    cudaLaunchConfig_t cfg = {};
    cfg.blockDim = dim3(kThreadsC, 1, 1);
    cfg.dynamicSmemBytes = L::alloc;
    cfg.stream = stream;
    cudaLaunchAttribute attr[2];
    attr[0].id = cudaLaunchAttributeClusterDimension;
    attr[0].val.clusterDim.x = 2;
    attr[0].val.clusterDim.y = 1;
    attr[0].val.clusterDim.z = 1;
    // Tiles wait for tiles of other CTA pairs, so every launched pair must be resident at once: never launch more pairs
    // than the device can hold of this kernel (asked once per device; an idle B200 holds one pair per two SMs)
    static std::atomic<int> max_pairs[64];
    if (device >= 0 && device < 64) {
        int mp = max_pairs[device].load();
        if (mp == 0) {
            cfg.gridDim = dim3((unsigned)(2 * pairs), 1, 1);
            cfg.attrs = attr;
            cfg.numAttrs = 1;
            int n = 0;
            if (cudaOccupancyMaxActiveClusters(&n, kern, &cfg) != cudaSuccess || n < 1) { cudaGetLastError(); n = pairs; }
            mp = n;
            max_pairs[device].store(mp);
        }
        if (pairs > mp) pairs = mp;
    }
    if (p.total_tiles < pairs) pairs = p.total_tiles;
    if (pairs < 1) return WQ4_OK;
    cfg.gridDim = dim3((unsigned)(2 * pairs), 1, 1);
    attr[1].id = cudaLaunchAttributeProgrammaticStreamSerialization;
    attr[1].val.programmaticStreamSerializationAllowed = 1;
    cfg.attrs = attr;
    cfg.numAttrs = wq4_pdl_enabled() ? 2 : 1;
    WQ4_CUDA(cudaLaunchKernelEx(&cfg, kern, maps, p));
    wq4_count_launch();
    return WQ4_OK;
}

}  // namespace

// Debug hook (not part of the drop-in surface): the NEXT chain launch stamps its leader CTAs' tile times into dev_ptr
// ([pairs][16][12] int64, zero-initialised by the caller); scripts/chain_timeline.py.
static long long *g_chain_timeline = nullptr;
extern "C" WQ4_API void wq4_debug_chain_timeline(void *dev_ptr) { g_chain_timeline = reinterpret_cast<long long *>(dev_ptr); }

size_t wq4_chain_counter_words(int64_t rows)
{
    const int64_t mt = (rows + 127) / 128;             // the smallest row tile gives the most row blocks
    return (size_t)(kMaxProb * mt + 1) * kCtrStride;
}

int wq4_chain_pick_tile_m(int64_t rows)
{
    static const int forced = [] { const char *e = getenv("WQ4_CHAIN_TILE_M"); return e ? atoi(e) : 0; }();
    if (forced == 128 || forced == 192) return forced;
    // one clip (1500 rows): 128-row blocks keep the K = 4F problem (few column pairs) on more CTA pairs; large batches
    // take the wider tile (fewer MMA instructions per flop)
    return rows <= 3000 ? 128 : 192;
}

int32_t wq4_launch_chain(const ChainGemm *g, int n, int64_t rows, unsigned *counters, int tile_m, cudaStream_t stream)
{
    if (!g || n < 1 || n > kMaxProb || rows < 1 || !counters) { wq4_set_error("chain: invalid argument"); return WQ4_ERR_INVALID; }
    if (tile_m != 128 && tile_m != 192) { wq4_set_error("chain: unsupported tile_m %d", tile_m); return WQ4_ERR_INVALID; }
    EncodeTiledFnC fn = get_encode_fn_c();
    if (!fn) { wq4_set_error("cuTensorMapEncodeTiled entry point not available"); return WQ4_ERR_CUDA; }
    ChainMaps maps;
    ChainParams p = {};
    p.nprob = n; p.rows = (int)rows; p.mtiles = (int)((rows + tile_m - 1) / tile_m);
    p.counters = counters;
    const int device = g[0].w->device;
    double flops = 0.0;
    int tile0 = 0;
    for (int i = 0; i < n; i++) {
        const wq4_tensor *w = g[i].w;
        if (!w || !w->matrix_mode || w->device != device || !g[i].x16) { wq4_set_error("chain: problem %d malformed", i); return WQ4_ERR_INVALID; }
        if (g[i].dep >= i) { wq4_set_error("chain: problem %d depends on a later problem", i); return WQ4_ERR_INVALID; }
        if (w->n % 4 != 0 || (g[i].stats_out && w->n % 128 != 0)) { wq4_set_error("chain: N must be a multiple of 4 (128 with statistics)"); return WQ4_ERR_SHAPE; }
        if (g[i].stats_out && !(g[i].residual && g[i].y32 && g[i].y16 && !g[i].cvec && !(g[i].flags & WQ4_EPI_GELU))) { wq4_set_error("chain: statistics are produced by the residual epilogue only"); return WQ4_ERR_INVALID; }
        if (g[i].cvec && (!g[i].stats_in || g[i].stats_groups < 1 || g[i].stats_groups > kMaxGroups)) { wq4_set_error("chain: problem %d folds a LayerNorm without a producer", i); return WQ4_ERR_INVALID; }
        cuuint64_t dims[3] = {(cuuint64_t)w->k, (cuuint64_t)rows, 1};
        cuuint64_t strides[2] = {(cuuint64_t)w->k * 2, (cuuint64_t)w->k * 2 * (cuuint64_t)rows};
        cuuint32_t box[3] = {64u, (cuuint32_t)(tile_m / 2), 1};
        cuuint32_t estr[3] = {1, 1, 1};
        CUresult r = fn(&maps.m[i], CU_TENSOR_MAP_DATA_TYPE_FLOAT16, 3, const_cast<void *>(g[i].x16), dims, strides, box, estr,
                        CU_TENSOR_MAP_INTERLEAVE_NONE, CU_TENSOR_MAP_SWIZZLE_128B, CU_TENSOR_MAP_L2_PROMOTION_L2_256B,
                        CU_TENSOR_MAP_FLOAT_OOB_FILL_NONE);
        if (r != CUDA_SUCCESS) { wq4_set_error("chain: cuTensorMapEncodeTiled failed (%d) for problem %d", (int)r, i); return WQ4_ERR_CUDA; }
        ChainProb &P = p.prob[i];
        P.nib = w->nib; P.scl = w->scl; P.bias = g[i].bias; P.cvec = g[i].cvec; P.residual = g[i].residual;
        P.y32 = g[i].y32; P.y16 = reinterpret_cast<__half *>(g[i].y16); P.y16_scale = g[i].y16_scale;
        P.stats_out = reinterpret_cast<float2 *>(g[i].stats_out); P.stats_in = reinterpret_cast<const float2 *>(g[i].stats_in);
        P.stats_groups = g[i].stats_groups;
        P.kbp = (int)w->kbp; P.n = (int)w->n; P.npairs = (int)(w->np_alloc / 2); P.kst = (int)(w->kbp / 4); P.ks0 = 0; P.ld = (int)w->n;
        if (g[i].k_stages > 0) { P.ks0 = g[i].k_stage0; P.kst = g[i].k_stages; }
        if (P.ks0 < 0 || P.ks0 + P.kst > (int)(w->kbp / 4)) { wq4_set_error("chain: K range of problem %d out of bounds", i); return WQ4_ERR_INVALID; }
        P.dep = g[i].dep;
        P.dep_count = g[i].dep >= 0 ? 2 * kEpiWarps * p.prob[g[i].dep].npairs : 0;     // 2 CTAs x 8 epilogue warps per pair tile
        P.edep = g[i].edep_plus1 - 1;
        if (P.edep >= i) { wq4_set_error("chain: problem %d has a later epilogue dependency", i); return WQ4_ERR_INVALID; }
        P.edep_count = P.edep >= 0 ? 2 * kEpiWarps * p.prob[P.edep].npairs : 0;
        P.tile0 = tile0;
        P.gelu = (g[i].flags & WQ4_EPI_GELU) ? 1 : 0;
        tile0 += P.npairs * p.mtiles;
        flops += 2.0 * (double)rows * (double)w->n * (double)P.kst * kBKC;
    }
    for (int i = n; i < kMaxProb; i++) p.prob[i].tile0 = 0x7fffffff;
    p.total_tiles = tile0;
    p.dev_err = wq4_error_flag_ptr(device);
    if (!p.dev_err) return WQ4_ERR_CUDA;
    p.timeline = g_chain_timeline;
    g_chain_timeline = nullptr;       // one shot
    WqProfScope prof(WQ4_PROF_GEMM_Q4, flops, stream);
    return tile_m == 128 ? launch_chain<128>(maps, p, device, stream) : launch_chain<192>(maps, p, device, stream);
}
