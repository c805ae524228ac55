// wq4_gemm.cu -- kernel family (1): Q4_0 dequant + GEMM on the 5th-generation tensor cores.
//
// Replaces src/gguf/shader.wgsl:51-92 / src/gguf/op.rs:47-116 for rows (= B*M) > 8 and, with a
// dense f16 weight operand, the Burn f32 matmul inside Conv1D::forward (src/model/layers.rs:124-125).
//
// Computes D^T:  acc[n, m] = sum_k W[n,k] * X[m,k]   (W is the UMMA "A" operand: 128 weight rows =
// 128 TMEM lanes; X is the "B" operand: TILE_M activation rows = TMEM columns).  This orientation
//   - dequantises every weight once per CTA tile and reuses it for TILE_M (<= 256) rows,
//   - makes the epilogue store coalesced without a shared-memory transpose: TMEM lane = output
//     column n, so the 32 lanes of a warp write 32 consecutive floats of one output row.
//
// Warp roles (512 threads, 1 CTA / SM, persistent over tiles):
//   warp 0      X producer     TMA (cp.async.bulk.tensor.3d, 128B swizzle) of the activation tile
//   warp 1      MMA issuer     one elected thread, tcgen05.mma.cta_group::1.kind::f16, f32 in TMEM
//   warp 2      raw producer   cp.async.bulk of the raw Q4 slab (4 KiB nibbles + 512 B scales / stage)
//   warp 3      TMEM allocator
//   warps 4-11  dequantisers   nibbles -> f16 (lop3 magic 0x6400, exact (nib-8), one rounding in
//                              the product with the f16 scale) -> 128B-swizzled K-major A tile;
//                              4 groups x 2 warps, group g owns ring slot g (4 stages in flight)
//   warps 12-15 epilogue       tcgen05.ld -> +bias -> GELU -> +residual -> f32 / f16 stores
// Pipelines: raw ring (6 stages) -> A/X ring (4 stages, mbarrier full/empty) -> double-buffered
// TMEM accumulator (full/empty) so the epilogue of tile i overlaps the main loop of tile i+1.
//
// CTA pairs: the kernel is launched as clusters of 2 CTAs that work on the same activation tile and
// on neighbouring weight panels.  Each CTA TMA-loads half of the X tile and multicasts it into both
// CTAs' shared memory, which halves the L2->SM traffic of the dominant operand (at 128x256 tiles a
// lone CTA needs 36.6 KB per 512 tensor-pipe cycles, more than the L2 can deliver to 148 SMs).
//
// Roofline: tensor cores.  Algorithmic work = 2*rows*N*K flop per launch.
#include <cuda.h>

#include <mutex>
#include <type_traits>

#include <atomic>

#include "wq4_common.cuh"
#include "wq4_ptx.cuh"

namespace {

using namespace ptx;

constexpr int kStages = 4;   // == number of dequant groups (group g owns slot g)
constexpr int kRawStages = 6;
constexpr int kThreads = 512;
constexpr int kBK = 64;                                 // K elements per stage
constexpr int kAStageBytes = WQ4_PANEL_ROWS * kBK * 2;  // 16 KiB
constexpr int kRawNibBytes = 2 * WQ4_PANEL_ROWS * 16;   // 4 KiB
constexpr int kRawSclBytes = 2 * WQ4_PANEL_ROWS * 2;    // 512 B
constexpr int kRawStageBytes = kRawNibBytes + kRawSclBytes;
constexpr int kNumDequantThreads = 256;
constexpr int kDequantGroupThreads = 64;   // 4 groups, one per A-ring slot
constexpr int kNumEpilogueThreads = 128;

template <int TILE_M>
struct SmemLayout {
    static constexpr int x_stage = TILE_M * kBK * 2;
    static constexpr int x_off = 0;
    static constexpr int a_off = x_off + kStages * x_stage;
    static constexpr int raw_off = a_off + kStages * kAStageBytes;
    static constexpr int bar_off = raw_off + kRawStages * kRawStageBytes;
    // barriers: full[S], empty[S], raw_full[R], raw_empty[R], acc_full[2], acc_empty[2]
    static constexpr int n_bars = 2 * kStages + 2 * kRawStages + 4;
    static constexpr int misc_off = bar_off + n_bars * 8;   // tmem base (u32), abort flag (int)
    static constexpr int total = misc_off + 16;
    static constexpr int alloc = total + 1024;              // slack for manual 1 KiB alignment
};

struct GemmParams {
    // Q4 weight operand
    const uint4 *nib;
    const __half *scl;
    int kbp;
    // problem
    int n;          // valid output columns (weight rows)
    int np;         // weight panels (n tiles of 128)
    int kst;        // K stages of 64
    int t_len;      // rows per batch item (valid extent of X dim 1)
    int nbatch;     // X dim 2
    int mt_per_b;   // ceil(t_len / TILE_M)
    // epilogue
    const float *bias;
    const float *residual;
    float *y32;
    void *y16;
    long long ob, ot, on;   // output element strides: batch, row, column
    long long rb;           // residual batch stride (rows/cols share ot/on)
    uint32_t flags;
    int *dev_err;
};

// (a & b) | c in one LOP3 (immLut 0xEA); b and c live in registers across the loop.
__device__ __forceinline__ uint32_t lop3_and_or(uint32_t a, uint32_t b, uint32_t c)
{
    uint32_t d;
    asm("lop3.b32 %0, %1, %2, %3, 0xEA;" : "=r"(d) : "r"(a), "r"(b), "r"(c));
    return d;
}

template <bool kBF16>
__device__ __forceinline__ uint4 dequant_word(uint32_t w, uint32_t d2)
{
    uint4 out;
    if constexpr (!kBF16) {
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
        out.x = *reinterpret_cast<uint32_t *>(&q0);
        out.y = *reinterpret_cast<uint32_t *>(&q1);
        out.z = *reinterpret_cast<uint32_t *>(&q2);
        out.w = *reinterpret_cast<uint32_t *>(&q3);
    } else {
        // bf16 has 8 significant bits: 128 + nib is exact, 128 + 16*nib is not -> shift first.
        const __nv_bfloat162 k136 = __floats2bfloat162_rn(136.0f, 136.0f);
        const __nv_bfloat162 d = *reinterpret_cast<const __nv_bfloat162 *>(&d2);
        uint32_t t0 = lop3_and_or(w, 0x000F000Fu, 0x43004300u);
        uint32_t t1 = lop3_and_or(w >> 4, 0x000F000Fu, 0x43004300u);
        uint32_t t2 = lop3_and_or(w >> 8, 0x000F000Fu, 0x43004300u);
        uint32_t t3 = lop3_and_or(w >> 12, 0x000F000Fu, 0x43004300u);
        __nv_bfloat162 q0 = __hmul2(__hsub2(*reinterpret_cast<__nv_bfloat162 *>(&t0), k136), d);
        __nv_bfloat162 q1 = __hmul2(__hsub2(*reinterpret_cast<__nv_bfloat162 *>(&t1), k136), d);
        __nv_bfloat162 q2 = __hmul2(__hsub2(*reinterpret_cast<__nv_bfloat162 *>(&t2), k136), d);
        __nv_bfloat162 q3 = __hmul2(__hsub2(*reinterpret_cast<__nv_bfloat162 *>(&t3), k136), d);
        out.x = *reinterpret_cast<uint32_t *>(&q0);
        out.y = *reinterpret_cast<uint32_t *>(&q1);
        out.z = *reinterpret_cast<uint32_t *>(&q2);
        out.w = *reinterpret_cast<uint32_t *>(&q3);
    }
    return out;
}

template <int TILE_M, bool kDenseA, bool kBF16>
__global__ void __launch_bounds__(kThreads, 1)
q4_gemm_kernel(const __grid_constant__ CUtensorMap tm_x, const __grid_constant__ CUtensorMap tm_a,
               const GemmParams p)
{
    using L = SmemLayout<TILE_M>;
    extern __shared__ uint8_t smem_raw[];
    const uint32_t base = (smem_u32(smem_raw) + 1023u) & ~1023u;
    uint8_t *base_ptr = smem_raw + (base - smem_u32(smem_raw));

    const uint32_t x_smem = base + L::x_off;
    const uint32_t a_smem = base + L::a_off;
    const uint32_t raw_smem = base + L::raw_off;
    const uint32_t bars = base + L::bar_off;
    auto bar_full = [&](int s) { return bars + 8u * s; };
    auto bar_empty = [&](int s) { return bars + 8u * (kStages + s); };
    auto bar_raw_full = [&](int s) { return bars + 8u * (2 * kStages + s); };
    auto bar_raw_empty = [&](int s) { return bars + 8u * (2 * kStages + kRawStages + s); };
    auto bar_acc_full = [&](int b) { return bars + 8u * (2 * kStages + 2 * kRawStages + b); };
    auto bar_acc_empty = [&](int b) { return bars + 8u * (2 * kStages + 2 * kRawStages + 2 + b); };
    volatile uint32_t *tmem_base_slot = reinterpret_cast<volatile uint32_t *>(base_ptr + L::misc_off);
    volatile int *cta_abort = reinterpret_cast<volatile int *>(base_ptr + L::misc_off + 4);

    const int warp = threadIdx.x >> 5;
    const int lane = threadIdx.x & 31;
    constexpr uint32_t kTmemCols = 2 * TILE_M;   // 64..512, power of two
    const int csize = (int)cluster_nctarank();   // 1 or 2 CTAs share each activation tile
    const int crank = (int)cluster_ctarank();
    const int npc = p.np / csize;                // panel groups (np is even when csize == 2)
    const int num_tiles = npc * p.mt_per_b * p.nbatch;      // tiles per cluster-slot
    const int slot = (int)blockIdx.x / csize;               // this CTA's cluster index
    const int nslots = (int)gridDim.x / csize;

    if (threadIdx.x == 0) {
        *cta_abort = 0;
        for (int s = 0; s < kStages; s++) {
            mbar_init(bar_full(s), kDenseA ? 1u : 1u + kDequantGroupThreads);
            mbar_init(bar_empty(s), (uint32_t)csize);
        }
        for (int s = 0; s < kRawStages; s++) {
            mbar_init(bar_raw_full(s), 1u);
            mbar_init(bar_raw_empty(s), kDequantGroupThreads);
        }
        for (int b = 0; b < 2; b++) {
            mbar_init(bar_acc_full(b), 1u);
            mbar_init(bar_acc_empty(b), kNumEpilogueThreads);
        }
        fence_mbar_init();
        prefetch_tensormap(&tm_x);
        if (kDenseA) prefetch_tensormap(&tm_a);
    }
    if (warp == 3) {
        tmem_alloc(smem_u32(const_cast<uint32_t *>(tmem_base_slot)), kTmemCols);
        tmem_relinquish();
    }
    tc_fence_before();
    __syncthreads();
    if (csize > 1) cluster_sync_all();   // peer barriers must exist before any multicast lands
    tc_fence_after();
    const uint32_t tmem_base = *tmem_base_slot;

    // tile -> (batch, m tile, weight panel); consecutive tiles share the activation tile
    auto decode_tile = [&](int tile, int &b, int &t0, int &panel) {
        panel = (tile % npc) * csize + crank;
        int mt = tile / npc;
        b = mt / p.mt_per_b;
        t0 = (mt - b * p.mt_per_b) * TILE_M;
    };

    if (warp == 0) {
        // ============================ X (and dense A) TMA producer ============================
        if (lane == 0) {
            uint32_t c = 0;
            for (int tile = slot; tile < num_tiles; tile += nslots) {
                int b, t0, panel;
                decode_tile(tile, b, t0, panel);
                for (int ks = 0; ks < p.kst; ks++, c++) {
                    const int s = c % kStages;
                    const uint32_t ph = (c / kStages) & 1u;
                    mbar_wait(bar_empty(s), ph ^ 1u, cta_abort, p.dev_err, 101);
                    mbar_arrive_expect_tx(bar_full(s), L::x_stage + (kDenseA ? kAStageBytes : 0));
                    if (csize == 1) {
                        tma_load_3d(x_smem + s * L::x_stage, &tm_x, bar_full(s), ks * kBK, t0, b);
                    } else {
                        // my half of the rows, delivered to both CTAs of the pair
                        tma_load_3d_mcast(x_smem + s * L::x_stage + crank * (L::x_stage / 2), &tm_x, bar_full(s),
                                          ks * kBK, t0 + crank * (TILE_M / 2), b, (uint16_t)0x3);
                    }
                    if (kDenseA) tma_load_2d(a_smem + s * kAStageBytes, &tm_a, bar_full(s), ks * kBK, panel * WQ4_PANEL_ROWS);
                }
            }
        }
    } else if (warp == 1) {
        // ===================================== MMA issuer =====================================
        if (lane == 0) {
            constexpr uint32_t idesc = umma_idesc_f16(128, TILE_M, kBF16);
            uint32_t c = 0, it = 0;
            for (int tile = slot; tile < num_tiles; tile += nslots, it++) {
                const uint32_t ab = it & 1u;
                const uint32_t aph = (it >> 1) & 1u;
                mbar_wait(bar_acc_empty(ab), aph ^ 1u, cta_abort, p.dev_err, 102);
                tc_fence_after();
                const uint32_t tmem_d = tmem_base + ab * TILE_M;
                for (int ks = 0; ks < p.kst; ks++, c++) {
                    const int s = c % kStages;
                    const uint32_t ph = (c / kStages) & 1u;
                    mbar_wait(bar_full(s), ph, cta_abort, p.dev_err, 103);
                    tc_fence_after();
                    const uint64_t adesc = umma_smem_desc_sw128(a_smem + s * kAStageBytes);
                    const uint64_t bdesc = umma_smem_desc_sw128(x_smem + s * L::x_stage);
#pragma unroll
                    for (int k = 0; k < kBK / 16; k++) {
                        // +32 bytes (16 K elements) inside the 128-byte swizzle atom: +2 in >>4 units
                        umma_f16(tmem_d, adesc + 2u * k, bdesc + 2u * k, idesc, (ks | k) != 0 ? 1u : 0u);
                    }
                    // frees the A/X slot when these MMAs retire (in both CTAs: the peer's producer
                    // multicasts into this CTA's slot as well)
                    if (csize == 1) umma_commit(bar_empty(s));
                    else umma_commit_mcast(bar_empty(s), (uint16_t)0x3);
                }
                umma_commit(bar_acc_full(ab));          // accumulator ready for the epilogue
            }
        }
    } else if (warp == 2) {
        // ================================= raw Q4 slab producer ================================
        if (!kDenseA && lane == 0) {
            uint32_t c = 0;
            for (int tile = slot; tile < num_tiles; tile += nslots) {
                int b, t0, panel;
                decode_tile(tile, b, t0, panel);
                const size_t panel_entry = (size_t)panel * p.kbp * WQ4_PANEL_ROWS;
                for (int ks = 0; ks < p.kst; ks++, c++) {
                    const int s = c % kRawStages;
                    const uint32_t ph = (c / kRawStages) & 1u;
                    mbar_wait(bar_raw_empty(s), ph ^ 1u, cta_abort, p.dev_err, 104);
                    mbar_arrive_expect_tx(bar_raw_full(s), kRawStageBytes);
                    const size_t e = panel_entry + (size_t)(2 * ks) * WQ4_PANEL_ROWS;
                    const uint32_t dst = raw_smem + s * kRawStageBytes;
                    bulk_load(dst, p.nib + e, kRawNibBytes, bar_raw_full(s));
                    bulk_load(dst + kRawNibBytes, p.scl + e, kRawSclBytes, bar_raw_full(s));
                }
            }
        }
    } else if (warp >= 4 && warp < 12) {
        // ===================================== dequantisers ====================================
        // Four groups of two warps; group g owns ring slot g, so four K-stages are dequantised
        // concurrently (the per-stage chain wait -> LDS -> ALU -> STS -> fence -> arrive is latency
        // bound; running it four-wide is what keeps the tensor pipe fed).
        if (!kDenseA) {
            const int g = (warp - 4) >> 1;                       // 0..3 == A-ring slot
            const int t64 = threadIdx.x - 128 - g * kDequantGroupThreads;   // 0..63
            const int my_tiles = (num_tiles - slot + nslots - 1) / nslots;
            const uint32_t total = (uint32_t)my_tiles * (uint32_t)p.kst;
            uint8_t *a_slot = base_ptr + L::a_off + g * kAStageBytes;
            for (uint32_t c = g; c < total; c += kStages) {
                const int rs = c % kRawStages;
                const uint32_t rph = (c / kRawStages) & 1u;
                const uint32_t ph = (c / kStages) & 1u;
                mbar_wait(bar_raw_full(rs), rph, cta_abort, p.dev_err, 105);
                const uint8_t *raw = base_ptr + L::raw_off + rs * kRawStageBytes;
                uint4 q[4];
                uint32_t d2[4];
#pragma unroll
                for (int j = 0; j < 4; j++) {                    // task j: rows (j&1)*64 + t64, block j>>1
                    const int e = (j >> 1) * 128 + (j & 1) * 64 + t64;
                    q[j] = *reinterpret_cast<const uint4 *>(raw + e * 16);
                    const uint16_t sb = *reinterpret_cast<const uint16_t *>(raw + kRawNibBytes + e * 2);
                    if constexpr (!kBF16) {
                        d2[j] = (uint32_t)sb | ((uint32_t)sb << 16);
                    } else {
                        __nv_bfloat16 db = __float2bfloat16_rn(__half2float(__ushort_as_half(sb)));
                        uint16_t bb = *reinterpret_cast<uint16_t *>(&db);
                        d2[j] = (uint32_t)bb | ((uint32_t)bb << 16);
                    }
                }
                mbar_arrive(bar_raw_empty(rs));                  // raw slot may be refilled
                mbar_wait(bar_empty(g), ph ^ 1u, cta_abort, p.dev_err, 106);
#pragma unroll
                for (int j = 0; j < 4; j++) {
                    const int r = (j & 1) * 64 + t64;
                    const int h = j >> 1;
                    // row r of the K-major tile: 128 bytes, 16-byte chunk i stored at i ^ (r & 7)
                    uint8_t *row = a_slot + r * 128;
                    const int x7 = r & 7;
                    *reinterpret_cast<uint4 *>(row + (((4 * h + 0) ^ x7) << 4)) = dequant_word<kBF16>(q[j].x, d2[j]);
                    *reinterpret_cast<uint4 *>(row + (((4 * h + 1) ^ x7) << 4)) = dequant_word<kBF16>(q[j].y, d2[j]);
                    *reinterpret_cast<uint4 *>(row + (((4 * h + 2) ^ x7) << 4)) = dequant_word<kBF16>(q[j].z, d2[j]);
                    *reinterpret_cast<uint4 *>(row + (((4 * h + 3) ^ x7) << 4)) = dequant_word<kBF16>(q[j].w, d2[j]);
                }
                fence_proxy_async_smem();                        // generic-proxy writes -> visible to tcgen05
                mbar_arrive(bar_full(g));
            }
        }
    } else if (warp >= 12) {
        // ======================================= epilogue =======================================
        const int q = warp & 3;                    // TMEM lane quarter this warp may access
        // The flag combination is warp-uniform and fixed for the launch: pick a specialised
        // straight-line body once (no per-element branches), generic body as the fallback.
        const bool f_gelu = (p.flags & WQ4_EPI_GELU) != 0, f_res = p.residual != nullptr;
        const bool f_y32 = p.y32 != nullptr, f_y16 = p.y16 != nullptr;
        int mode = 7;
        if (!f_gelu && !f_res && f_y32 && !f_y16) mode = 0;        // q4_matmul / Q4Linear drop-in
        else if (!f_gelu && !f_res && !f_y32 && f_y16) mode = 1;   // fused QKV -> f16
        else if (f_gelu && !f_res && !f_y32 && f_y16) mode = 2;    // fc1 / conv1 + GELU -> f16
        else if (!f_gelu && f_res && f_y32 && !f_y16) mode = 3;    // out-proj / fc2 + residual
        else if (f_gelu && !f_res && f_y32 && !f_y16) mode = 4;    // conv + GELU drop-in
        else if (f_gelu && f_res && f_y32 && !f_y16) mode = 5;     // conv2 + GELU + positional emb.
        uint32_t it = 0;
        for (int tile = slot; tile < num_tiles; tile += nslots, it++) {
            int b, t0, panel;
            decode_tile(tile, b, t0, panel);
            const uint32_t ab = it & 1u;
            const uint32_t aph = (it >> 1) & 1u;
            mbar_wait(bar_acc_full(ab), aph, cta_abort, p.dev_err, 107);
            tc_fence_after();
            const int n = panel * WQ4_PANEL_ROWS + q * 32 + lane;
            const bool n_ok = n < p.n;
            const float bias = (p.bias && n_ok) ? p.bias[n] : 0.0f;
            const long long obase = (long long)b * p.ob + (long long)n * p.on;
            const long long rbase = (long long)b * p.rb + (long long)n * p.on;
#pragma unroll 1
            for (int c0 = 0; c0 < TILE_M; c0 += 32) {
                if (t0 + c0 >= p.t_len) break;     // warp-uniform
                uint32_t v[32];
                tmem_ld_32x32b_x32(tmem_base + ab * TILE_M + c0 + ((uint32_t)(q * 32) << 16), v);
                tmem_ld_wait();
                if (!n_ok) continue;
                const int tb = t0 + c0;
                const int nvalid = min(32, p.t_len - tb);
                float *y32 = p.y32 ? p.y32 + obase + (long long)tb * p.ot : nullptr;
                const float *res = p.residual ? p.residual + rbase + (long long)tb * p.ot : nullptr;
                const long long ot = p.ot;
                auto body = [&](auto GELU, auto RES, auto Y32, auto Y16, auto FULL) {
                    [[maybe_unused]] auto *y16h = reinterpret_cast<__half *>(p.y16) + (Y16.value ? obase + (long long)tb * ot : 0);
                    [[maybe_unused]] auto *y16b = reinterpret_cast<__nv_bfloat16 *>(p.y16) + (Y16.value ? obase + (long long)tb * ot : 0);
                    [[maybe_unused]] float r[32];
                    if constexpr (RES.value) {
#pragma unroll
                        for (int j = 0; j < 32; j++) r[j] = (FULL.value || j < nvalid) ? res[(long long)j * ot] : 0.0f;
                    }
#pragma unroll
                    for (int j = 0; j < 32; j++) {
                        float val = __uint_as_float(v[j]) + bias;
                        if constexpr (GELU.value) val = wq4_gelu_tanh(val);
                        if constexpr (RES.value) val += r[j];
                        if (FULL.value || j < nvalid) {
                            if constexpr (Y32.value) y32[(long long)j * ot] = val;
                            if constexpr (Y16.value) {
                                if constexpr (kBF16) y16b[(long long)j * ot] = __float2bfloat16_rn(val);
                                else y16h[(long long)j * ot] = __float2half_rn(val);
                            }
                        }
                    }
                };
                using T_ = std::true_type;
                using F_ = std::false_type;
                const bool full = nvalid == 32;
#define WQ4_EPI(G, R, A, B)                                        \
    do {                                                           \
        if (full) body(G{}, R{}, A{}, B{}, T_{});                  \
        else body(G{}, R{}, A{}, B{}, F_{});                       \
    } while (0)
                switch (mode) {
                case 0: WQ4_EPI(F_, F_, T_, F_); break;
                case 1: WQ4_EPI(F_, F_, F_, T_); break;
                case 2: WQ4_EPI(T_, F_, F_, T_); break;
                case 3: WQ4_EPI(F_, T_, T_, F_); break;
                case 4: WQ4_EPI(T_, F_, T_, F_); break;
                case 5: WQ4_EPI(T_, T_, T_, F_); break;
                default:
                    // generic (rare flag combinations): runtime flags, per-element checks
#pragma unroll 4
                    for (int j = 0; j < 32; j++) {
                        if (j < nvalid) {
                            float val = __uint_as_float(v[j]) + bias;
                            if (f_gelu) val = wq4_gelu_tanh(val);
                            const long long o = obase + (long long)(tb + j) * ot;
                            if (f_res) val += p.residual[rbase + (long long)(tb + j) * ot];
                            if (f_y32) p.y32[o] = val;
                            if (f_y16) {
                                if constexpr (kBF16) reinterpret_cast<__nv_bfloat16 *>(p.y16)[o] = __float2bfloat16_rn(val);
                                else reinterpret_cast<__half *>(p.y16)[o] = __float2half_rn(val);
                            }
                        }
                    }
                }
#undef WQ4_EPI
            }
            tc_fence_before();
            mbar_arrive(bar_acc_empty(ab));
        }
    }

    tc_fence_before();
    __syncthreads();
    if (csize > 1) cluster_sync_all();   // no CTA may exit while its peer can still signal it
    if (warp == 3) {
        tc_fence_after();
        tmem_dealloc(tmem_base, kTmemCols);
    }
}

// ---- host side -------------------------------------------------------------------------------
typedef CUresult (*EncodeTiledFn)(CUtensorMap *, CUtensorMapDataType, cuuint32_t, void *, const cuuint64_t *,
                                  const cuuint64_t *, const cuuint32_t *, const cuuint32_t *,
                                  CUtensorMapInterleave, CUtensorMapSwizzle, CUtensorMapL2promotion,
                                  CUtensorMapFloatOOBfill);

EncodeTiledFn get_encode_fn()
{
    static EncodeTiledFn fn = nullptr;
    static std::once_flag once;
    std::call_once(once, [] {
        void *p = nullptr;
        cudaDriverEntryPointQueryResult qres;
        if (cudaGetDriverEntryPoint("cuTensorMapEncodeTiled", &p, cudaEnableDefault, &qres) == cudaSuccess &&
            qres == cudaDriverEntryPointSuccess)
            fn = reinterpret_cast<EncodeTiledFn>(p);
    });
    return fn;
}

// K-major 16-bit tensor [d2][d1][d0 = K]; box {64, box_rows, 1}; 128-byte swizzle; OOB -> 0.
int32_t make_tmap(CUtensorMap *m, const void *ptr, uint64_t d0, uint64_t d1, uint64_t d2,
                  uint64_t stride1_bytes, uint64_t stride2_bytes, uint32_t box_rows, bool bf16, int rank)
{
    EncodeTiledFn fn = get_encode_fn();
    if (!fn) { wq4_set_error("cuTensorMapEncodeTiled entry point not available"); return WQ4_ERR_CUDA; }
    cuuint64_t dims[3] = {d0, d1, d2};
    cuuint64_t strides[2] = {stride1_bytes, stride2_bytes};
    cuuint32_t box[3] = {(cuuint32_t)kBK, box_rows, 1};
    cuuint32_t estr[3] = {1, 1, 1};
    CUresult r = fn(m, bf16 ? CU_TENSOR_MAP_DATA_TYPE_BFLOAT16 : CU_TENSOR_MAP_DATA_TYPE_FLOAT16, (cuuint32_t)rank,
                    const_cast<void *>(ptr), dims, strides, box, estr, CU_TENSOR_MAP_INTERLEAVE_NONE,
                    CU_TENSOR_MAP_SWIZZLE_128B, CU_TENSOR_MAP_L2_PROMOTION_L2_256B,
                    CU_TENSOR_MAP_FLOAT_OOB_FILL_NONE);
    if (r != CUDA_SUCCESS) {
        wq4_set_error("cuTensorMapEncodeTiled failed (%d): dims {%llu,%llu,%llu} strides {%llu,%llu} box_rows %u ptr %p",
                      (int)r, (unsigned long long)d0, (unsigned long long)d1, (unsigned long long)d2,
                      (unsigned long long)stride1_bytes, (unsigned long long)stride2_bytes, box_rows, ptr);
        return WQ4_ERR_CUDA;
    }
    return WQ4_OK;
}

template <int TILE_M, bool kDenseA, bool kBF16>
int32_t launch_tile(const CUtensorMap &tx, const CUtensorMap &ta, const GemmParams &p, int device,
                    cudaStream_t stream)
{
    using L = SmemLayout<TILE_M>;
    auto kern = q4_gemm_kernel<TILE_M, kDenseA, kBF16>;
    static std::atomic<bool> attr_done[16];   // per device
    if (device < 16 && !attr_done[device]) {
        WQ4_CUDA(cudaFuncSetAttribute(kern, cudaFuncAttributeMaxDynamicSharedMemorySize, L::alloc));
        attr_done[device] = true;
    } else if (device >= 16) {
        WQ4_CUDA(cudaFuncSetAttribute(kern, cudaFuncAttributeMaxDynamicSharedMemorySize, L::alloc));
    }
    const int csize = wq4_gemm_cluster_size(p.np);
    const int tiles = p.np * p.mt_per_b * p.nbatch;     // CTA tiles (a pair takes two per step)
    int grid = wq4_sm_count(device) / csize * csize;
    if (tiles < grid) grid = tiles;
    if (grid < 1) return WQ4_OK;
    cudaLaunchConfig_t cfg = {};
    cfg.gridDim = dim3((unsigned)grid, 1, 1);
    cfg.blockDim = dim3(kThreads, 1, 1);
    cfg.dynamicSmemBytes = L::alloc;
    cfg.stream = stream;
    cudaLaunchAttribute attr[1];
    attr[0].id = cudaLaunchAttributeClusterDimension;
    attr[0].val.clusterDim.x = (unsigned)csize;
    attr[0].val.clusterDim.y = 1;
    attr[0].val.clusterDim.z = 1;
    cfg.attrs = attr;
    cfg.numAttrs = 1;
    WQ4_CUDA(cudaLaunchKernelEx(&cfg, kern, tx, ta, p));
    wq4_count_launch();
    return WQ4_OK;
}

template <bool kDenseA>
int32_t launch_any(int tile_m, bool bf16, const CUtensorMap &tx, const CUtensorMap &ta, const GemmParams &p,
                   int device, cudaStream_t stream)
{
#define WQ4_CASE(T)                                                                         \
    case T:                                                                                 \
        return bf16 ? launch_tile<T, kDenseA, true>(tx, ta, p, device, stream)              \
                    : launch_tile<T, kDenseA, false>(tx, ta, p, device, stream);
    switch (tile_m) {
        WQ4_CASE(32)
        WQ4_CASE(64)
        WQ4_CASE(128)
        WQ4_CASE(256)
    default:
        wq4_set_error("gemm: unsupported tile_m %d", tile_m);
        return WQ4_ERR_INVALID;
    }
#undef WQ4_CASE
}

}  // namespace

// CTA pairs need an even number of weight panels; WQ4_GEMM_CLUSTER=1 disables pairing.
int wq4_gemm_cluster_size(int64_t np)
{
    static int forced = -1;
    if (forced < 0) {
        const char *e = getenv("WQ4_GEMM_CLUSTER");
        forced = e ? atoi(e) : 0;
    }
    if (forced == 1) return 1;
    return (np % 2 == 0) ? 2 : 1;
}

// Row-tile heuristic.  cost = waves * (tile_m + c): per-tile time grows with the MMA work
// (proportional to tile_m) plus a fixed part (dequantisation of the 128 x K panel, pipeline
// fill, epilogue drain) worth about 96 rows.  Override with WQ4_GEMM_TILE_M for experiments.
int wq4_pick_tile_m(int64_t rows, int64_t np)
{
    static int forced = -1;
    if (forced < 0) {
        const char *e = getenv("WQ4_GEMM_TILE_M");
        forced = e ? atoi(e) : 0;
    }
    if (forced == 32 || forced == 64 || forced == 128 || forced == 256) return forced;
    const int sms = 148;
    int best = 32;
    double best_cost = 1e30;
    const int cands[4] = {256, 128, 64, 32};
    for (int i = 0; i < 4; i++) {
        int tm = cands[i];
        int64_t tiles = np * ((rows + tm - 1) / tm);
        int64_t waves = (tiles + sms - 1) / sms;
        double cost = (double)waves * (tm + 96.0);
        if (cost < best_cost) { best_cost = cost; best = tm; }
    }
    return best;
}

int32_t wq4_launch_gemm_q4(const LinearArgs &a, int tile_m)
{
    const wq4_tensor *w = a.w;
    const bool bf16 = (a.flags & WQ4_OPERANDS_BF16) != 0;
    const void *x16 = a.x_f16;
    if (!x16) {
        // drop-in f32 activations: convert once into the caller's workspace
        size_t need = (size_t)a.rows * w->k * 2;
        if (!a.workspace || a.workspace_bytes < need) {
            wq4_set_error("gemm: workspace too small (%zu < %zu)", a.workspace_bytes, need);
            return WQ4_ERR_WORKSPACE;
        }
        int32_t st = wq4_launch_cvt_f32_to_16(a.x_f32, a.workspace, a.rows * w->k, bf16, a.stream);
        if (st != WQ4_OK) return st;
        x16 = a.workspace;
    }
    CUtensorMap tx, ta;
    const int csize = wq4_gemm_cluster_size(w->np);
    int32_t st = make_tmap(&tx, x16, (uint64_t)w->k, (uint64_t)a.rows, 1, (uint64_t)w->k * 2,
                           (uint64_t)w->k * 2 * (uint64_t)a.rows, (uint32_t)(tile_m / csize), bf16, 3);
    if (st != WQ4_OK) return st;
    ta = tx;  // unused by the Q4 variant
    GemmParams p = {};
    p.nib = w->nib; p.scl = w->scl; p.kbp = (int)w->kbp;
    p.n = (int)w->n; p.np = (int)w->np; p.kst = (int)(w->kbp / 2);
    p.t_len = (int)a.rows; p.nbatch = 1; p.mt_per_b = (int)((a.rows + tile_m - 1) / tile_m);
    p.bias = a.bias; p.residual = a.residual; p.y32 = a.y_f32; p.y16 = a.y_f16;
    p.ob = 0; p.ot = w->n; p.on = 1; p.rb = 0;
    p.flags = a.flags; p.dev_err = wq4_error_flag_ptr(w->device);
    if (!p.dev_err) return WQ4_ERR_CUDA;
    WqProfScope prof(WQ4_PROF_GEMM_Q4, 2.0 * (double)a.rows * (double)w->n * (double)w->k, a.stream);
    return launch_any<false>(tile_m, bf16, tx, ta, p, w->device, a.stream);
}

int32_t wq4_launch_gemm_dense(const DenseGemmArgs &a, int tile_m)
{
    const bool bf16 = (a.flags & WQ4_OPERANDS_BF16) != 0;
    CUtensorMap tx, ta;
    const int csize = wq4_gemm_cluster_size(a.n_pad / WQ4_PANEL_ROWS);
    int32_t st = make_tmap(&tx, a.x16, (uint64_t)a.x_k, (uint64_t)a.x_rows, (uint64_t)a.x_batch,
                           (uint64_t)a.x_row_stride * 2, (uint64_t)a.x_batch_stride * 2, (uint32_t)(tile_m / csize), bf16, 3);
    if (st != WQ4_OK) return st;
    st = make_tmap(&ta, a.w16, (uint64_t)a.kdim_pad, (uint64_t)a.n_pad, 1, (uint64_t)a.kdim_pad * 2, 0,
                   WQ4_PANEL_ROWS, bf16, 2);
    if (st != WQ4_OK) return st;
    GemmParams p = {};
    p.n = (int)a.n; p.np = (int)(a.n_pad / WQ4_PANEL_ROWS); p.kst = (int)(a.kdim_pad / kBK);
    p.t_len = (int)a.x_rows; p.nbatch = (int)a.x_batch; p.mt_per_b = (int)((a.x_rows + tile_m - 1) / tile_m);
    p.bias = a.bias; p.residual = a.residual; p.y32 = a.y_f32; p.y16 = a.y_f16;
    p.ob = a.ob; p.ot = a.ot; p.on = a.on; p.rb = a.rb;
    p.flags = a.flags; p.dev_err = wq4_error_flag_ptr(a.device);
    if (!p.dev_err) return WQ4_ERR_CUDA;
    WqProfScope prof(WQ4_PROF_GEMM_DENSE, 2.0 * (double)a.x_rows * (double)a.x_batch * (double)a.n * (double)a.x_k, a.stream);
    return launch_any<true>(tile_m, bf16, tx, ta, p, a.device, a.stream);
}
