// wq4_gemm2.cu -- kernel family (1), second generation: Q4_0 dequant + GEMM over a CTA PAIR with the
// dequantised weights held in TENSOR MEMORY.
//
// Replaces src/gguf/shader.wgsl:51-92 / src/gguf/op.rs:47-116 for rows (= B*M) > 8.
//
// Why a second kernel: profiles/r01 showed the first one (wq4_gemm.cu, both operands in shared memory,
// cta_group::1) bound by SHARED-MEMORY BANDWIDTH, not by the tensor pipe: per 64-wide K stage the MMA
// read A+B from shared memory (32-48 KB), the dequantisers stored A (16 KB) and TMA wrote X (16-32 KB)
// through the same 128 B/clk port -> ~570-820 clk per stage against 256-512 clk of MMA.  Here
//   * the dequantised weight tile never touches shared memory: dequant warps write f16 pairs straight
//     into TMEM with tcgen05.st (one thread = one weight row = one TMEM lane) and the MMA takes its A
//     operand from TMEM (tcgen05.mma ... [a_tmem], b_desc);
//   * the activation tile (B operand, UMMA N = TILE_M rows) is split over the two CTAs of a pair
//     (cta_group::2): each CTA TMA-loads and keeps only TILE_M/2 rows, so X costs half the shared-memory
//     writes, half the capacity (deeper ring) and half the L2->SM traffic;
// leaving ~64 B/clk of shared-memory traffic per SM at full tensor rate.
//
// acc[n, m] = sum_k W[n,k] * X[m,k]: UMMA M = 256 = two 128-row weight panels (one per CTA, TMEM lanes),
// UMMA N = TILE_M activation rows (TMEM columns).  TMEM lane = output column n, so the epilogue's 32 lanes
// store 32 consecutive floats of one output row.
//
// Warp roles per CTA (512 threads, persistent over tiles, 1 CTA / SM):
//   warp 0      X producer    TMA of this CTA's half of the activation tile; completes on the LEADER's barrier
//   warp 1      MMA issuer    leader CTA only, one thread: tcgen05.mma.cta_group::2.kind::f16, commits multicast
//   warp 2      raw producer  cp.async.bulk of this CTA's panel slab (4 KiB nibbles + 512 B scales per stage)
//   warp 3      TMEM allocator (cta_group::2, 512 columns)
//   warps 4-11  dequantisers  two sets of four warps alternate over K stages; thread r of a set owns weight
//                             row r: 2 blocks -> 32 x half2 -> one tcgen05.st.32x32b.x32
//   warps 12-15 epilogue      tcgen05.ld -> +bias -> GELU -> +residual -> f32 / f16 stores; two accumulator buffers
//                             (TILE_M <= 192) so it runs under the next tile's MMAs
// Barriers: "full"-type barriers (x_full, a_full, acc_empty) live in the leader CTA and are arrived on by
// both CTAs (remote arrive / TMA complete_tx on a mapa address); "empty"-type barriers (x_empty, a_empty,
// acc_full) live in each CTA and are signalled by the leader's tcgen05.commit multicast.
//
// Roofline: tensor cores.  Algorithmic work = 2*rows*N*K flop per launch.
#include <cuda.h>

#include <mutex>

#include <atomic>

#include "wq4_common.cuh"
#include "wq4_ptx.cuh"

namespace {

using namespace ptx;

constexpr int kThreads2 = 512;
constexpr int kBK2 = 128;                                // K elements per stage = 4 k-blocks = 2 TMA boxes of 64
constexpr int kRawNib = 4 * WQ4_PANEL_ROWS * 16;         // 8 KiB
constexpr int kRawScl = 4 * WQ4_PANEL_ROWS * 2;          // 1 KiB
constexpr int kRawStage = kRawNib + kRawScl;             // 9216 B
constexpr int kRawStages2 = 8;
constexpr int kAColsPerStage = kBK2 / 2;                 // 64 TMEM columns: 128 f16 per lane, two per column
constexpr int kAStagesMax = 4;                           // deepest TMEM weight ring (64 columns per slot)

template <int TILE_M>
struct Cfg2 {
    static_assert(TILE_M == 64 || TILE_M == 128 || TILE_M == 192 || TILE_M == 256, "TILE_M");
    static constexpr int HM = TILE_M / 2;                 // activation rows this CTA keeps
    static constexpr int x_box = HM * 64 * 2;             // bytes of one 64-wide TMA box
    static constexpr int x_stage = 2 * x_box;
    static constexpr int ACC = TILE_M >= 256 ? 1 : 2;     // accumulator buffers in TMEM (2: epilogue overlaps the next tile)
    static constexpr int a_col0 = ACC * TILE_M;           // first TMEM column of the A ring
    // A-ring depth: what the 512 columns leave after the accumulators (192 rows: 2 slots; the dequantisers hold the
    // next stage in registers while they wait for a slot, so two slots already keep the tensor pipe fed)
    static constexpr int AST = (512 - a_col0) / kAColsPerStage < kAStagesMax ? (512 - a_col0) / kAColsPerStage : kAStagesMax;
    // X ring in shared memory (decoupled from the TMEM weight ring; depth beyond 4 measured neutral)
    static constexpr int S = TILE_M == 192 ? 6 : 4;
    static_assert(AST >= 2 && a_col0 + AST * kAColsPerStage <= 512, "TMEM budget");
    static constexpr int x_off = 0;
    static constexpr int raw_off = x_off + S * x_stage;
    static constexpr int bar_off = raw_off + kRawStages2 * kRawStage;
    // full[S] empty[S] raw_full[R] raw_empty[R] acc_full[ACC] acc_empty[ACC]
    static constexpr int n_bars = 2 * S + 2 * kRawStages2 + 2 * ACC;
    static constexpr int misc_off = bar_off + n_bars * 8;
    static constexpr int total = misc_off + 16;
    static constexpr int alloc = total + 1024;
    static_assert(alloc <= 227 * 1024, "shared memory budget");
};

struct Gemm2Params {
    const uint4 *nib;
    const __half *scl;
    int kbp;
    int n;          // valid output columns (weight rows)
    int npairs;     // weight panel pairs
    int kst;        // K stages of 64
    int rows;       // activation rows
    int mtiles;     // ceil(rows / TILE_M)
    const float *bias;
    const float *residual;
    float *y32;
    void *y16;
    long long ldy;  // output row stride (elements)
    uint32_t flags;
    int *dev_err;
    long long *timeline;   // debug (wq4_debug_gemm_timeline): clock64 stamps of CTA 0's roles, normally NULL
};

__device__ __forceinline__ uint32_t lop3_and_or2(uint32_t a, uint32_t b, uint32_t c)
{
    uint32_t d;
    asm("lop3.b32 %0, %1, %2, %3, 0xEA;" : "=r"(d) : "r"(a), "r"(b), "r"(c));
    return d;
}

// One device word (8 nibbles, interleaved as described in wq4_common.cuh) -> 4 x (2 x 16-bit) in K order.
template <bool kBF16>
__device__ __forceinline__ void dequant_word2(uint32_t w, uint32_t d2, uint32_t *out)
{
    if constexpr (!kBF16) {
        const __half2 k1032 = __floats2half2_rn(1032.0f, 1032.0f);
        const __half2 k16th = __floats2half2_rn(0.0625f, 0.0625f);
        const __half2 km72 = __floats2half2_rn(-72.0f, -72.0f);
        const __half2 d = *reinterpret_cast<const __half2 *>(&d2);
        uint32_t t0 = lop3_and_or2(w, 0x000F000Fu, 0x64006400u);
        uint32_t t1 = lop3_and_or2(w, 0x00F000F0u, 0x64006400u);
        uint32_t w8 = w >> 8;
        uint32_t t2 = lop3_and_or2(w8, 0x000F000Fu, 0x64006400u);
        uint32_t t3 = lop3_and_or2(w8, 0x00F000F0u, 0x64006400u);
        // exact: (1024 + e) - 1032 = e - 8 ; (1024 + 16 e)/16 - 72 = e - 8 ; one rounding in the product
        __half2 q0 = __hmul2(__hsub2(*reinterpret_cast<__half2 *>(&t0), k1032), d);
        __half2 q1 = __hmul2(__hfma2(*reinterpret_cast<__half2 *>(&t1), k16th, km72), d);
        __half2 q2 = __hmul2(__hsub2(*reinterpret_cast<__half2 *>(&t2), k1032), d);
        __half2 q3 = __hmul2(__hfma2(*reinterpret_cast<__half2 *>(&t3), k16th, km72), d);
        out[0] = *reinterpret_cast<uint32_t *>(&q0);
        out[1] = *reinterpret_cast<uint32_t *>(&q1);
        out[2] = *reinterpret_cast<uint32_t *>(&q2);
        out[3] = *reinterpret_cast<uint32_t *>(&q3);
    } else {
        const __nv_bfloat162 k136 = __floats2bfloat162_rn(136.0f, 136.0f);
        const __nv_bfloat162 d = *reinterpret_cast<const __nv_bfloat162 *>(&d2);
        uint32_t t0 = lop3_and_or2(w, 0x000F000Fu, 0x43004300u);
        uint32_t t1 = lop3_and_or2(w >> 4, 0x000F000Fu, 0x43004300u);
        uint32_t t2 = lop3_and_or2(w >> 8, 0x000F000Fu, 0x43004300u);
        uint32_t t3 = lop3_and_or2(w >> 12, 0x000F000Fu, 0x43004300u);
        __nv_bfloat162 q0 = __hmul2(__hsub2(*reinterpret_cast<__nv_bfloat162 *>(&t0), k136), d);
        __nv_bfloat162 q1 = __hmul2(__hsub2(*reinterpret_cast<__nv_bfloat162 *>(&t1), k136), d);
        __nv_bfloat162 q2 = __hmul2(__hsub2(*reinterpret_cast<__nv_bfloat162 *>(&t2), k136), d);
        __nv_bfloat162 q3 = __hmul2(__hsub2(*reinterpret_cast<__nv_bfloat162 *>(&t3), k136), d);
        out[0] = *reinterpret_cast<uint32_t *>(&q0);
        out[1] = *reinterpret_cast<uint32_t *>(&q1);
        out[2] = *reinterpret_cast<uint32_t *>(&q2);
        out[3] = *reinterpret_cast<uint32_t *>(&q3);
    }
}

template <bool kBF16>
__device__ __forceinline__ uint32_t scale_pair(uint16_t sb)
{
    if constexpr (!kBF16) {
        return (uint32_t)sb | ((uint32_t)sb << 16);
    } else {
        __nv_bfloat16 db = __float2bfloat16_rn(__half2float(__ushort_as_half(sb)));
        uint16_t bb = *reinterpret_cast<uint16_t *>(&db);
        return (uint32_t)bb | ((uint32_t)bb << 16);
    }
}

// Epilogue of one 32-column chunk: lane = output column n, v[j] = accumulator of activation row tb + j,
// r[j] = residual value (RES only, loaded by the caller one chunk ahead).  All 32 values are computed
// unconditionally so the 32 chains overlap (guarding each element serialises bias/GELU behind a branch);
// only the stores of a ragged last chunk are predicated.
template <bool GELU, bool RES, bool Y32, bool Y16, bool kBF16, typename Prefetch>
__device__ __forceinline__ void epi_chunk(const uint32_t (&v)[32], float (&r)[32], float bias, int nvalid,
                                          float *y32, void *y16, long long ld, Prefetch &&prefetch_next)
{
    float val[32];
#pragma unroll
    for (int j = 0; j < 32; j++) {
        val[j] = __uint_as_float(v[j]) + bias;
        if constexpr (GELU) val[j] = wq4_gelu_tanh(val[j]);
        if constexpr (RES) val[j] += r[j];
    }
    if constexpr (RES) prefetch_next();       // r[] is free again: next chunk's residual loads fly under the stores
    auto store = [&](int j) {
        if constexpr (Y32) y32[(long long)j * ld] = val[j];
        if constexpr (Y16) {
            if constexpr (kBF16) reinterpret_cast<__nv_bfloat16 *>(y16)[(long long)j * ld] = __float2bfloat16_rn(val[j]);
            else reinterpret_cast<__half *>(y16)[(long long)j * ld] = __float2half_rn(val[j]);
        }
    };
    if (nvalid == 32) {
#pragma unroll
        for (int j = 0; j < 32; j++) store(j);
    } else {
#pragma unroll
        for (int j = 0; j < 32; j++)
            if (j < nvalid) store(j);
    }
}

template <int TILE_M, bool kBF16>
__global__ void __launch_bounds__(kThreads2, 1)
q4_gemm2_kernel(const __grid_constant__ CUtensorMap tm_x, const Gemm2Params p)
{
    using L = Cfg2<TILE_M>;
    constexpr int S = L::S, RS = kRawStages2, ACC = L::ACC;
    extern __shared__ uint8_t smem_raw2[];
    const uint32_t base = (smem_u32(smem_raw2) + 1023u) & ~1023u;
    uint8_t *base_ptr = smem_raw2 + (base - smem_u32(smem_raw2));

    const uint32_t x_smem = base + L::x_off;
    const uint32_t raw_smem = base + L::raw_off;
    const uint32_t bars = base + L::bar_off;
    auto bar_full = [&](int s) { return bars + 8u * s; };                       // leader: X bytes + 16 dequant warps
    auto bar_empty = [&](int s) { return bars + 8u * (S + s); };                // per CTA: MMA commit (multicast)
    auto bar_raw_full = [&](int s) { return bars + 8u * (2 * S + s); };
    auto bar_raw_empty = [&](int s) { return bars + 8u * (2 * S + RS + s); };
    auto bar_acc_full = [&](int b) { return bars + 8u * (2 * S + 2 * RS + b); };         // per CTA: commit
    auto bar_acc_empty = [&](int b) { return bars + 8u * (2 * S + 2 * RS + ACC + b); };  // leader: 8 epilogue warps
    volatile uint32_t *tmem_base_slot = reinterpret_cast<volatile uint32_t *>(base_ptr + L::misc_off);
    volatile int *cta_abort = reinterpret_cast<volatile int *>(base_ptr + L::misc_off + 4);

    const int warp = threadIdx.x >> 5;
    const int lane = threadIdx.x & 31;
    const int crank = (int)cluster_ctarank();          // 0 = leader
    const int num_tiles = p.npairs * p.mtiles;
    const int slot = (int)blockIdx.x >> 1;
    const int nslots = (int)gridDim.x >> 1;

    if (threadIdx.x == 0) {
        *cta_abort = 0;
        for (int s = 0; s < S; s++) { mbar_init(bar_full(s), 17u); mbar_init(bar_empty(s), 1u); }
        for (int s = 0; s < RS; s++) { mbar_init(bar_raw_full(s), 1u); mbar_init(bar_raw_empty(s), 8u); }
        for (int b = 0; b < ACC; b++) { mbar_init(bar_acc_full(b), 1u); mbar_init(bar_acc_empty(b), 8u); }
        fence_mbar_init();
        prefetch_tensormap(&tm_x);
    }
    if (warp == 3) {
        tmem_alloc2(smem_u32(const_cast<uint32_t *>(tmem_base_slot)), 512u);
        tmem_relinquish2();
    }
    tc_fence_before();
    __syncthreads();
    cluster_sync_all();     // both CTAs' barriers exist before anything remote can land
    tc_fence_after();
    const uint32_t tmem_base = *tmem_base_slot;
#ifdef WQ4_GEMM_TIMELINE      // build.sh -DWQ4_GEMM_TIMELINE: clock64 stamps of CTA 0's role loops (scripts/gemm_timeline.py)
    long long *const tl = blockIdx.x == 0 ? p.timeline : nullptr;
#define WQ4_TL(role, c, k) do { if (tl && (c) < 64u && lane == 0) tl[((role) * 64 + (c)) * 4 + (k)] = clock64(); } while (0)
#else
#define WQ4_TL(role, c, k) do { } while (0)
#endif
    // PDL: the next kernel in the stream may be scheduled from here on (it parks in its own grid_dep_wait()).  In this
    // kernel only the warps that touch the predecessor's output wait for it -- the X producer and the epilogue; the
    // weight stream (raw producer, dequantisers -> TMEM) starts under the predecessor's tail.
    grid_dep_launch();

    // Every role loop below is executed by all 32 lanes of its warp (warp-uniform control flow and operands);
    // only the issuing instruction is predicated on an elected lane.
    if (warp == 0) {
        // ================================ X producer (both CTAs) ================================
        grid_dep_wait();
        const uint32_t full_leader0 = mapa_u32(bar_full(0), 0u);
        uint32_t c = 0;
        for (int tile = slot; tile < num_tiles; tile += nslots) {
            const int t0 = (tile / p.npairs) * TILE_M + crank * L::HM;
            for (int ks = 0; ks < p.kst; ks++, c++) {
                const int s = c % S;
                mbar_wait(bar_empty(s), ((c / S) & 1u) ^ 1u, cta_abort, p.dev_err, 201);
                WQ4_TL(0, c, 0);
                if (elect_one()) {
                    if (crank == 0) mbar_arrive_expect_tx(bar_full(s), 2u * L::x_stage);
                    const uint32_t dst = x_smem + s * L::x_stage;
                    tma_load_3d_2cta(dst, &tm_x, full_leader0 + 8u * s, ks * kBK2, t0, 0);
                    tma_load_3d_2cta(dst + L::x_box, &tm_x, full_leader0 + 8u * s, ks * kBK2 + 64, t0, 0);
                }
                __syncwarp();
            }
        }
    } else if (warp == 1) {
        // ================================ MMA issuer (leader only) ==============================
        if (crank == 0) {
            constexpr uint32_t idesc = umma_idesc_f16(256, TILE_M, kBF16);
            uint32_t c = 0, it = 0;
            uint32_t full_ok = 0u;
            for (int tile = slot; tile < num_tiles; tile += nslots, it++) {
                const uint32_t ab = it % ACC;
                mbar_wait(bar_acc_empty(ab), ((it / ACC) & 1u) ^ 1u, cta_abort, p.dev_err, 202);
                tc_fence_after();
                const uint32_t tmem_d = tmem_base + ab * TILE_M;
                for (int ks = 0; ks < p.kst; ks++, c++) {
                    const int s = c % S;
                    WQ4_TL(1, c, 0);
                    if (!full_ok) mbar_wait(bar_full(s), (c / S) & 1u, cta_abort, p.dev_err, 203);
                    WQ4_TL(1, c, 1);
                    tc_fence_after();
                    // probe the next stage's barrier now: the ~200 clk probe latency runs under this stage's MMA issue
                    full_ok = mbar_test_wait(bar_full((c + 1) % S), ((c + 1) / S) & 1u);
                    const uint64_t bdesc0 = umma_smem_desc_sw128(x_smem + s * L::x_stage);
                    const uint64_t bdesc1 = umma_smem_desc_sw128(x_smem + s * L::x_stage + L::x_box);
                    const uint32_t a_tmem = tmem_base + L::a_col0 + (c % L::AST) * kAColsPerStage;
                    if (elect_one()) {
#pragma unroll
                        for (int k = 0; k < 4; k++)
                            umma_f16_ts_2cta(tmem_d, a_tmem + 8u * k, bdesc0 + 2u * k, idesc, (ks | k) != 0 ? 1u : 0u);
#pragma unroll
                        for (int k = 0; k < 4; k++)
                            umma_f16_ts_2cta(tmem_d, a_tmem + 32u + 8u * k, bdesc1 + 2u * k, idesc, 1u);
                        umma_commit2_mcast(bar_empty(s), (uint16_t)0x3);
                        if (ks == p.kst - 1) umma_commit2_mcast(bar_acc_full(ab), (uint16_t)0x3);
                    }
                    __syncwarp();
                    WQ4_TL(1, c, 2);
                }
            }
        }
    } else if (warp == 2) {
        // ================================ raw Q4 slab producer (both CTAs) ======================
        uint32_t c = 0;
        for (int tile = slot; tile < num_tiles; tile += nslots) {
            const int panel = (tile % p.npairs) * 2 + crank;
            const size_t panel_entry = (size_t)panel * p.kbp * WQ4_PANEL_ROWS;
            for (int ks = 0; ks < p.kst; ks++, c++) {
                const int s = c % RS;
                mbar_wait(bar_raw_empty(s), ((c / RS) & 1u) ^ 1u, cta_abort, p.dev_err, 205);
                if (elect_one()) {
                    mbar_arrive_expect_tx(bar_raw_full(s), kRawStage);
                    const size_t e = panel_entry + (size_t)(4 * ks) * WQ4_PANEL_ROWS;
                    const uint32_t dst = raw_smem + s * kRawStage;
                    bulk_load(dst, p.nib + e, kRawNib, bar_raw_full(s));
                    bulk_load(dst + kRawNib, p.scl + e, kRawScl, bar_raw_full(s));
                }
                __syncwarp();
            }
        }
    } else if (warp >= 4 && warp < 12) {
        // ================================ dequantisers (both CTAs) ==============================
        // All 8 warps work on every stage: warp (q, half) dequantises the 64 K-elements `half` of the 32 weight
        // rows of TMEM lane quarter q: 2 blocks -> 32 x (2 x f16) -> one tcgen05.st.32x32b.x32.
        const int dw = warp - 4;
        const int q = dw & 3;                         // == warp % 4: the TMEM lane quarter this warp may access
        const int half = dw >> 2;
        const int r = q * 32 + lane;                  // weight row inside the panel == TMEM lane
        const uint32_t my_tmem = tmem_base + L::a_col0 + half * 32 + ((uint32_t)(q * 32) << 16);
        const uint32_t full_leader0 = mapa_u32(bar_full(0), 0u);
        const int my_tiles = (num_tiles - slot + nslots - 1) / nslots;
        const uint32_t total = (uint32_t)(my_tiles > 0 ? my_tiles : 0) * (uint32_t)p.kst;
        // An mbarrier probe costs ~200 clk even when the phase completed long ago (timeline in profiles/), and this loop
        // has two of them per stage.  Both are therefore issued one step early -- the raw slab of stage c+1 and the TMEM
        // slot of stage c are probed (non-blocking test_wait) before the dequant arithmetic of stage c -- and only waited
        // on if that early probe failed, which in steady state it does not.
        uint32_t raw_ok = 0u;
        for (uint32_t c = 0; c < total; c++) {
            const int rs = c % RS;
            if (dw == 0) WQ4_TL(2, c, 0);
            if (!raw_ok) mbar_wait(bar_raw_full(rs), (c / RS) & 1u, cta_abort, p.dev_err, 206);
            if (dw == 0) WQ4_TL(2, c, 1);
            const uint8_t *raw = base_ptr + L::raw_off + rs * kRawStage;
            const int e0 = (2 * half) * WQ4_PANEL_ROWS + r, e1 = e0 + WQ4_PANEL_ROWS;
            const uint4 q0 = *reinterpret_cast<const uint4 *>(raw + e0 * 16);
            const uint4 q1 = *reinterpret_cast<const uint4 *>(raw + e1 * 16);
            const uint32_t d0 = scale_pair<kBF16>(*reinterpret_cast<const uint16_t *>(raw + kRawNib + e0 * 2));
            const uint32_t d1 = scale_pair<kBF16>(*reinterpret_cast<const uint16_t *>(raw + kRawNib + e1 * 2));
            raw_ok = (c + 1 < total) ? mbar_test_wait(bar_raw_full((c + 1) % RS), ((c + 1) / RS) & 1u) : 1u;
            uint32_t slot_ok = 1u;
            const uint32_t pc = c - L::AST;           // slot c % AST is free once the MMA of stage c - AST has completed
            if (c >= (uint32_t)L::AST) slot_ok = mbar_test_wait(bar_empty(pc % S), (pc / S) & 1u);
            uint32_t v[32];
            dequant_word2<kBF16>(q0.x, d0, v + 0);
            dequant_word2<kBF16>(q0.y, d0, v + 4);
            dequant_word2<kBF16>(q0.z, d0, v + 8);
            dequant_word2<kBF16>(q0.w, d0, v + 12);
            dequant_word2<kBF16>(q1.x, d1, v + 16);
            dequant_word2<kBF16>(q1.y, d1, v + 20);
            dequant_word2<kBF16>(q1.z, d1, v + 24);
            dequant_word2<kBF16>(q1.w, d1, v + 28);
            __syncwarp();                             // every lane has consumed its raw bytes
            if (elect_one()) mbar_arrive(bar_raw_empty(rs));
            const int s = c % S;
            if (dw == 0) WQ4_TL(2, c, 2);
            if (!slot_ok) mbar_wait(bar_empty(pc % S), (pc / S) & 1u, cta_abort, p.dev_err, 207);
            __syncwarp();
            if (dw == 0) WQ4_TL(2, c, 3);
            tc_fence_after();
            tmem_st_32x32b_x32(my_tmem + (c % L::AST) * kAColsPerStage, v);
            tmem_st_wait();
            if (dw == 0) WQ4_TL(3, c, 0);
            tc_fence_before();
            __syncwarp();
            if (elect_one()) mbar_arrive_cluster(full_leader0 + 8u * s);
        }
    } else if (warp >= 12) {
        // ================================ epilogue (both CTAs) ==================================
        const int q = warp & 3;
        const bool f_gelu = (p.flags & WQ4_EPI_GELU) != 0, f_res = p.residual != nullptr;
        const bool f_y32 = p.y32 != nullptr, f_y16 = p.y16 != nullptr;
        int mode = 7;
        if (!f_gelu && !f_res && f_y32 && !f_y16) mode = 0;        // q4_matmul / Q4Linear drop-in
        else if (!f_gelu && !f_res && !f_y32 && f_y16) mode = 1;   // fused QKV -> f16
        else if (f_gelu && !f_res && !f_y32 && f_y16) mode = 2;    // fc1 + GELU -> f16
        else if (!f_gelu && f_res && f_y32 && !f_y16) mode = 3;    // out-proj / fc2 + residual
        const uint32_t acc_empty_leader0 = mapa_u32(bar_acc_empty(0), 0u);
        grid_dep_wait();       // residual reads and output writes must follow the predecessor grid
        uint32_t it = 0;
        for (int tile = slot; tile < num_tiles; tile += nslots, it++) {
            const int panel = (tile % p.npairs) * 2 + crank;
            const int t0 = (tile / p.npairs) * TILE_M;
            const uint32_t ab = it % ACC;
            const int n = panel * WQ4_PANEL_ROWS + q * 32 + lane;
            const bool n_ok = n < p.n;
            const float bias = (p.bias && n_ok) ? p.bias[n] : 0.0f;
            // residual values are fetched one chunk ahead (first chunk: before the accumulator is even ready)
            float r[32];
            auto load_res = [&](int c0) {
                const int tb = t0 + c0;
                const float *res = p.residual + (long long)tb * p.ldy + n;
                const int nv = n_ok ? p.rows - tb : 0;
#pragma unroll
                for (int j = 0; j < 32; j++) r[j] = (j < nv) ? __ldg(res + (long long)j * p.ldy) : 0.0f;
            };
            if (mode == 3) load_res(0);
            mbar_wait(bar_acc_full(ab), (it / ACC) & 1u, cta_abort, p.dev_err, 208);
            tc_fence_after();
#pragma unroll 1
            for (int c0 = 0; c0 < TILE_M; c0 += 32) {
                const int tb = t0 + c0;
                if (tb >= p.rows) break;                   // warp-uniform
                uint32_t v[32];
                tmem_ld_32x32b_x32(tmem_base + ab * TILE_M + c0 + ((uint32_t)(q * 32) << 16), v);
                tmem_ld_wait();
                auto prefetch = [&] { if (c0 + 32 < TILE_M && tb + 32 < p.rows) load_res(c0 + 32); };
                if (!n_ok) continue;
                const int nvalid = min(32, p.rows - tb);
                const long long o = (long long)tb * p.ldy + n;
                float *y32 = p.y32 ? p.y32 + o : nullptr;
                void *y16 = nullptr;
                if (p.y16) y16 = kBF16 ? (void *)(reinterpret_cast<__nv_bfloat16 *>(p.y16) + o)
                                       : (void *)(reinterpret_cast<__half *>(p.y16) + o);
                switch (mode) {
                case 0: epi_chunk<false, false, true, false, kBF16>(v, r, bias, nvalid, y32, y16, p.ldy, prefetch); break;
                case 1: epi_chunk<false, false, false, true, kBF16>(v, r, bias, nvalid, y32, y16, p.ldy, prefetch); break;
                case 2: epi_chunk<true, false, false, true, kBF16>(v, r, bias, nvalid, y32, y16, p.ldy, prefetch); break;
                case 3: epi_chunk<false, true, true, false, kBF16>(v, r, bias, nvalid, y32, y16, p.ldy, prefetch); break;
                default:
#pragma unroll
                    for (int j = 0; j < 32; j++) {
                        float val = __uint_as_float(v[j]) + bias;
                        if (f_gelu) val = wq4_gelu_tanh(val);
                        if (f_res) val += __ldg(p.residual + o + (long long)min(j, nvalid - 1) * p.ldy);
                        if (j < nvalid) {
                            if (f_y32) y32[(long long)j * p.ldy] = val;
                            if (f_y16) {
                                if constexpr (kBF16) reinterpret_cast<__nv_bfloat16 *>(y16)[(long long)j * p.ldy] = __float2bfloat16_rn(val);
                                else reinterpret_cast<__half *>(y16)[(long long)j * p.ldy] = __float2half_rn(val);
                            }
                        }
                    }
                }
            }
            tc_fence_before();
            __syncwarp();
            if (elect_one()) mbar_arrive_cluster(acc_empty_leader0 + 8u * ab);
        }
    }

    tc_fence_before();
    __syncthreads();
    cluster_sync_all();     // no CTA may exit (or free TMEM) while its peer can still signal / read it
    if (warp == 3) {
        tc_fence_after();
        tmem_dealloc2(tmem_base, 512u);
    }
}

// ---- host side -------------------------------------------------------------------------------
typedef CUresult (*EncodeTiledFn2)(CUtensorMap *, CUtensorMapDataType, cuuint32_t, void *, const cuuint64_t *,
                                   const cuuint64_t *, const cuuint32_t *, const cuuint32_t *,
                                   CUtensorMapInterleave, CUtensorMapSwizzle, CUtensorMapL2promotion,
                                   CUtensorMapFloatOOBfill);

EncodeTiledFn2 get_encode_fn2()
{
    static EncodeTiledFn2 fn = nullptr;
    static std::once_flag once;
    std::call_once(once, [] {
        void *p = nullptr;
        cudaDriverEntryPointQueryResult qres;
        if (cudaGetDriverEntryPoint("cuTensorMapEncodeTiled", &p, cudaEnableDefault, &qres) == cudaSuccess &&
            qres == cudaDriverEntryPointSuccess)
            fn = reinterpret_cast<EncodeTiledFn2>(p);
    });
    return fn;
}

template <int TILE_M, bool kBF16>
int32_t launch2(const CUtensorMap &tx, const Gemm2Params &p, int device, cudaStream_t stream)
{
    using L = Cfg2<TILE_M>;
    auto kern = q4_gemm2_kernel<TILE_M, kBF16>;
    static std::atomic<bool> attr_done[64];
    if (device < 0 || device >= 64 || !attr_done[device]) {
        WQ4_CUDA(cudaFuncSetAttribute(kern, cudaFuncAttributeMaxDynamicSharedMemorySize, L::alloc));
        if (device >= 0 && device < 64) attr_done[device] = true;
    }
    const int tiles = p.npairs * p.mtiles;              // pair tiles
    int pairs = wq4_sm_count(device) / 2;
    if (tiles < pairs) pairs = tiles;
    if (pairs < 1) return WQ4_OK;
    cudaLaunchConfig_t cfg = {};
    cfg.gridDim = dim3((unsigned)(2 * pairs), 1, 1);
    cfg.blockDim = dim3(kThreads2, 1, 1);
    cfg.dynamicSmemBytes = L::alloc;
    cfg.stream = stream;
    cudaLaunchAttribute attr[2];
    attr[0].id = cudaLaunchAttributeClusterDimension;
    attr[0].val.clusterDim.x = 2;
    attr[0].val.clusterDim.y = 1;
    attr[0].val.clusterDim.z = 1;
    attr[1].id = cudaLaunchAttributeProgrammaticStreamSerialization;
    attr[1].val.programmaticStreamSerializationAllowed = 1;
    cfg.attrs = attr;
    cfg.numAttrs = wq4_pdl_enabled() ? 2 : 1;
    WQ4_CUDA(cudaLaunchKernelEx(&cfg, kern, tx, p));
    wq4_count_launch();
    return WQ4_OK;
}

}  // namespace

// Debug hook (not part of the drop-in surface): dev_ptr = 4 roles x 64 stages x 4 int64 clock64 stamps of CTA 0 of the
// following GEMM launches (role 0 X producer, 1 MMA issuer, 2/3 dequant warp 0); NULL disables.
static long long *g_gemm_timeline_host = nullptr;
extern "C" WQ4_API void wq4_debug_gemm_timeline(void *dev_ptr) { g_gemm_timeline_host = reinterpret_cast<long long *>(dev_ptr); }

// Row-tile heuristic for the CTA-pair kernel.  Per pair tile the time is ~ max(MMA, dequant) per K stage plus a
// fixed fill/drain part; cost model in units of "activation rows": waves * (max(tile_m, 150) + 48), fitted to
// profiles/r01 (per pair tile at K = 1280: 6.1 / 7.3 / 10.6 us for 128 / 192 / 256 rows).
int wq4_pick_tile_m2(int64_t rows, int64_t npairs, int device)
{
    static int forced = -1;
    if (forced < 0) {
        const char *e = getenv("WQ4_GEMM_TILE_M");
        forced = e ? atoi(e) : 0;
    }
    if (forced == 64 || forced == 128 || forced == 192 || forced == 256) return forced;
    // One row tile or less: such launches are a single serial walk over K per CTA pair, and a K stage costs
    // 8 x (36 + TILE_M/2) clk of MMA issue, so the narrowest tile is the fastest (decode steps of 17..64 clips)
    if (rows <= 64) return 64;
    const int pairs = wq4_sm_count(device) / 2 > 0 ? wq4_sm_count(device) / 2 : 74;   // CTA pairs that run at once
    int best = 64;
    double best_cost = 1e30;
    const int cands[4] = {256, 192, 128, 64};
    for (int i = 0; i < 4; i++) {
        const int tm = cands[i];
        const int64_t tiles = npairs * ((rows + tm - 1) / tm);
        const int64_t waves = (tiles + pairs - 1) / pairs;
        const double per = (tm > 150 ? tm : 150) + 48.0 + (tm == 256 ? 46.0 : 0.0);   // 256: epilogue not overlapped
        const double cost = (double)waves * per;
        if (cost < best_cost) { best_cost = cost; best = tm; }
    }
    return best;
}

int32_t wq4_launch_gemm2_q4(const LinearArgs &a, int tile_m)
{
    const wq4_tensor *w = a.w;
    const bool bf16 = (a.flags & WQ4_OPERANDS_BF16) != 0;
    const void *x16 = a.x_f16;
    if (!x16) {
        size_t need = (size_t)a.rows * w->k * 2;
        if (!a.workspace || a.workspace_bytes < need) {
            wq4_set_error("gemm: workspace too small (%zu < %zu)", a.workspace_bytes, need);
            return WQ4_ERR_WORKSPACE;
        }
        int32_t st = wq4_launch_cvt_f32_to_16(a.x_f32, a.workspace, a.rows * w->k, bf16, a.stream);
        if (st != WQ4_OK) return st;
        x16 = a.workspace;
    }
    EncodeTiledFn2 fn = get_encode_fn2();
    if (!fn) { wq4_set_error("cuTensorMapEncodeTiled entry point not available"); return WQ4_ERR_CUDA; }
    CUtensorMap tx;
    cuuint64_t dims[3] = {(cuuint64_t)w->k, (cuuint64_t)a.rows, 1};
    cuuint64_t strides[2] = {(cuuint64_t)w->k * 2, (cuuint64_t)w->k * 2 * (cuuint64_t)a.rows};
    cuuint32_t box[3] = {64u, (cuuint32_t)(tile_m / 2), 1};
    cuuint32_t estr[3] = {1, 1, 1};
    CUresult r = fn(&tx, bf16 ? CU_TENSOR_MAP_DATA_TYPE_BFLOAT16 : CU_TENSOR_MAP_DATA_TYPE_FLOAT16, 3,
                    const_cast<void *>(x16), dims, strides, box, estr, CU_TENSOR_MAP_INTERLEAVE_NONE,
                    CU_TENSOR_MAP_SWIZZLE_128B, CU_TENSOR_MAP_L2_PROMOTION_L2_256B, CU_TENSOR_MAP_FLOAT_OOB_FILL_NONE);
    if (r != CUDA_SUCCESS) {
        wq4_set_error("cuTensorMapEncodeTiled failed (%d): K %lld rows %lld tile_m %d ptr %p", (int)r, (long long)w->k,
                      (long long)a.rows, tile_m, x16);
        return WQ4_ERR_CUDA;
    }
    Gemm2Params p = {};
    p.nib = w->nib; p.scl = w->scl; p.kbp = (int)w->kbp;
    p.n = (int)w->n; p.npairs = (int)(w->np_alloc / 2); p.kst = (int)(w->kbp / 4);
    p.rows = (int)a.rows; p.mtiles = (int)((a.rows + tile_m - 1) / tile_m);
    p.bias = a.bias; p.residual = a.residual; p.y32 = a.y_f32; p.y16 = a.y_f16;
    p.ldy = w->n;
    p.flags = a.flags; p.dev_err = wq4_error_flag_ptr(w->device);
    p.timeline = g_gemm_timeline_host;
    if (!p.dev_err) return WQ4_ERR_CUDA;
    WqProfScope prof(WQ4_PROF_GEMM_Q4, 2.0 * (double)a.rows * (double)w->n * (double)w->k, a.stream);
#define WQ4_CASE2(T)                                                                             \
    case T:                                                                                      \
        return bf16 ? launch2<T, true>(tx, p, w->device, a.stream) : launch2<T, false>(tx, p, w->device, a.stream);
    switch (tile_m) {
        WQ4_CASE2(64)
        WQ4_CASE2(128)
        WQ4_CASE2(192)
        WQ4_CASE2(256)
    default:
        wq4_set_error("gemm2: unsupported tile_m %d", tile_m);
        return WQ4_ERR_INVALID;
    }
#undef WQ4_CASE2
}
