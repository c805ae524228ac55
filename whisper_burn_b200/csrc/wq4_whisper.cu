// wq4_whisper.cu -- host-side C++ of the callers around the Q4 hot path: GGUF reader, model
// loader, encoder, decoder with in-place KV caches, greedy loop, batched transcribe.
//
// Mirrors (file:line under the reference tree):
//   GgufReader                      src/gguf/reader.rs:98-223
//   load_whisper_from_gguf & co     src/gguf/loader.rs:26-377 (tensor names, [k,n] -> [n,k], optional biases)
//   WhisperEncoder::forward         src/model/encoder.rs:37-49,87-115
//   WhisperDecoder::{forward_prompt,decode_step}   src/model/decoder.rs:77-167,251-348
//   WhisperModel::transcribe        src/model/whisper.rs:51-138 (prompt modes, EOT suppression, MAX_TOKENS)
//   transcribe (batching is new)    src/transcribe.rs:34-107
// Differences by design: q/k/v (and cross k/v) projections are fused into one Q4 weight at load; the
// KV caches are preallocated f16 buffers written in place (no Tensor::cat realloc, attention.rs:106-113);
// argmax runs on the device; utterances are batched (M = 1500*B in the encoder, M = B in the decoder).
#include <stdio.h>
#include <string.h>

#include <chrono>
#include <map>
#include <string>
#include <vector>

#include <fcntl.h>
#include <sys/mman.h>
#include <sys/stat.h>
#include <unistd.h>

#include "wq4_common.cuh"
#include "wq4_model.cuh"

namespace {

// ------------------------------------------------------------------------------------ GGUF reader
struct GgufTensor {
    std::vector<uint64_t> dims;   // as stored (reversed w.r.t. row-major shape)
    uint32_t dtype;               // 0 F32, 1 F16, 2 Q4_0 (reader.rs:27-35)
    uint64_t offset;
    // Every size derived from file-controlled values is computed with overflow checks (a crafted file must fail cleanly,
    // as the reference's reader does on short reads); `bad` marks a tensor whose element or byte count does not fit 64 bits.
    bool bad = false;
    uint64_t elems = 0, bytes = 0;
    void finalize()
    {
        uint64_t n = 1;
        for (auto d : dims) if (__builtin_mul_overflow(n, d, &n)) { bad = true; return; }
        elems = n;
        if (dtype == 0) { if (__builtin_mul_overflow(n, (uint64_t)4, &bytes)) bad = true; }
        else if (dtype == 1) { if (__builtin_mul_overflow(n, (uint64_t)2, &bytes)) bad = true; }
        else bytes = n / 32 * 18;             // reader.rs:38-48
    }
    uint64_t num_elements() const { return elems; }
    uint64_t byte_size() const { return bytes; }
};

struct Gguf {
    const uint8_t *data = nullptr;
    size_t size = 0;
    uint32_t version = 0;
    uint64_t tensor_count = 0;
    uint64_t data_offset = 0;
    std::map<std::string, GgufTensor> tensors;
    std::string err;

    bool rd(size_t &pos, void *dst, size_t n)
    {
        if (pos > size || n > size - pos) { err = "unexpected end of GGUF data"; return false; }
        memcpy(dst, data + pos, n);
        pos += n;
        return true;
    }
    bool rd_str(size_t &pos, std::string &s)
    {
        uint64_t n;
        if (!rd(pos, &n, 8)) return false;
        if (n > size - pos) { err = "unexpected end of GGUF string"; return false; }   // pos <= size after rd()
        s.assign(reinterpret_cast<const char *>(data + pos), (size_t)n);
        pos += (size_t)n;
        return true;
    }
    bool skip_value(size_t &pos, uint32_t t, int depth = 0)   // reader.rs:237-285
    {
        static const int sz[13] = {1, 1, 2, 2, 4, 4, 4, 1, -1, -2, 8, 8, 8};
        if (t > 12) { err = "Unknown GGUF metadata value type"; return false; }
        if (sz[t] > 0) { if ((size_t)sz[t] > size - pos) { err = "truncated metadata"; return false; } pos += sz[t]; return true; }
        if (t == 8) { std::string s; return rd_str(pos, s); }
        uint32_t et; uint64_t cnt;
        if (!rd(pos, &et, 4) || !rd(pos, &cnt, 8)) return false;
        if (depth >= 16) { err = "GGUF metadata arrays nested too deeply"; return false; }   // bounds the recursion on crafted files
        if (et <= 12 && sz[et] > 0) {                                                         // fixed-size elements: one checked skip
            uint64_t nb;
            if (__builtin_mul_overflow(cnt, (uint64_t)sz[et], &nb) || nb > size - pos) { err = "truncated metadata array"; return false; }
            pos += (size_t)nb;
            return true;
        }
        for (uint64_t i = 0; i < cnt; i++) if (!skip_value(pos, et, depth + 1)) return false;   // each element consumes >= 4 bytes or fails
        return true;
    }
    bool parse(const uint8_t *d, size_t n)     // reader.rs:105-188
    {
        data = d; size = n;
        size_t pos = 0;
        uint32_t magic;
        if (!rd(pos, &magic, 4)) return false;
        if (magic != 0x46554747u) { err = "Invalid GGUF magic"; return false; }
        if (!rd(pos, &version, 4)) return false;
        if (version != 2 && version != 3) { err = "Unsupported GGUF version (expected 2 or 3)"; return false; }
        uint64_t kv;
        if (!rd(pos, &tensor_count, 8) || !rd(pos, &kv, 8)) return false;
        for (uint64_t i = 0; i < kv; i++) {
            std::string key; uint32_t vt;
            if (!rd_str(pos, key) || !rd(pos, &vt, 4) || !skip_value(pos, vt)) return false;
        }
        for (uint64_t i = 0; i < tensor_count; i++) {
            std::string name; uint32_t nd;
            GgufTensor t;
            if (!rd_str(pos, name) || !rd(pos, &nd, 4)) return false;
            if (nd > 8) { err = "tensor rank too large"; return false; }
            t.dims.resize(nd);
            for (uint32_t j = 0; j < nd; j++) if (!rd(pos, &t.dims[j], 8)) return false;
            if (!rd(pos, &t.dtype, 4) || !rd(pos, &t.offset, 8)) return false;
            if (t.dtype > 2) { err = "Unsupported GGML dtype code"; return false; }
            t.finalize();
            if (t.bad) { err = "tensor '" + name + "' size overflows 64 bits"; return false; }
            tensors[name] = t;
        }
        if (pos > size || size - pos < 31) { data_offset = size; return true; }   // no room for a data section: every tensor_data() fails
        data_offset = (pos + 31) / 32 * 32;    // reader.rs:178-179
        return true;
    }
    const GgufTensor *find(const std::string &name) const
    {
        auto it = tensors.find(name);
        return it == tensors.end() ? nullptr : &it->second;
    }
    const uint8_t *tensor_data(const GgufTensor &t) const
    {
        uint64_t a;
        if (t.bad || __builtin_add_overflow(data_offset, t.offset, &a)) return nullptr;
        if (a > size || t.byte_size() > size - a) return nullptr;
        return data + a;
    }
};

static float f16_bits_to_f32(uint16_t h)
{
    __half hv = __ushort_as_half(h);
    return __half2float(hv);
}

// ------------------------------------------------------------------------------------------ model
struct Lin { wq4_tensor *w = nullptr; float *bias = nullptr; };
struct LNp { float *w = nullptr; float *b = nullptr; };
struct EncLayer {
    LNp attn_ln; Lin qkv, out; LNp mlp_ln; Lin fc1, fc2;
    // LayerNorm folded into the consumer GEMM (wq4_chain.cu): c = W.gamma, b' = bias + W.beta for q|k|v (attn_ln) and fc1 (mlp_ln)
    float *qkv_c = nullptr, *qkv_b2 = nullptr, *fc1_c = nullptr, *fc1_b2 = nullptr;
};
struct DecLayer {
    LNp attn_ln; Lin qkv, out; LNp cross_ln; Lin cq, ckv, cout; LNp mlp_ln; Lin fc1, fc2;
    // LayerNorm fold vectors of the chained decode step (wq4_chain.cu): c = W.gamma, b' = bias + W.beta
    float *qkv_c = nullptr, *qkv_b2 = nullptr, *cq_c = nullptr, *cq_b2 = nullptr, *fc1_c = nullptr, *fc1_b2 = nullptr;
};

}  // namespace

struct wq4_whisper {
    wq4_whisper_config cfg;
    int device = 0;
    int64_t D = 0, F = 0, V = 0, T_mel = 3000, T_enc = 1500;
    wq4_conv1d *conv1 = nullptr, *conv2 = nullptr;
    float *enc_pos = nullptr;        // [n_audio_ctx, D]
    std::vector<EncLayer> enc;
    LNp ln_post;
    float *tok_emb = nullptr;        // [V, D] f32 (decoder.rs:171, logits matmul :292,343)
    float *dec_pos = nullptr;        // [n_text_ctx, D]
    std::vector<DecLayer> dec;
    LNp ln;
    std::vector<void *> owned;       // device allocations that live as long as the model (weights and what is derived from them)
    std::vector<void *> ws_owned;    // per-batch workspaces: freed and re-allocated when reserve() grows
    size_t weight_bytes = 0;
    // workspaces (reserved for max_batch)
    int64_t max_batch = 0;
    int64_t kv_batch = 0;            // clips the KV caches are currently laid out for: [layer][kv_batch][...]; set by cross_kv()
    __half *mel_t16 = nullptr, *h1_16 = nullptr, *xn16 = nullptr, *qkv16 = nullptr, *att16 = nullptr, *ffn16 = nullptr,
           *enc16 = nullptr;
    float *x = nullptr, *enc32 = nullptr;
    float2 *ln_stats_a = nullptr, *ln_stats_b = nullptr;   // [D/128][rows] row partials of the residual stream (wq4_chain.cu)
    unsigned *chain_ctr = nullptr;                         // row-block arrival counters of the chain launches, zero between launches
    bool chain_ready = false;
    // wq4_whisper_encode replays a CUDA graph of the encoder per batch size (first call eager, second call captures): 54
    // launches per Medium pass become one graph launch on the caller's stream.  call count and exec per batch.
    struct EncGraph { int calls = 0; cudaGraphExec_t exec = nullptr; int64_t launches = 0; bool failed = false; };
    std::map<int64_t, EncGraph> enc_graphs;
    float2 *dln_stats_a = nullptr, *dln_stats_b = nullptr; // the same for the decoder's chained steps: [D/128][max_batch * max_q]
    unsigned *dchain_ctr = nullptr;
    int64_t Cp1 = 0, Tp1 = 0, Tp2 = 0;
    // decoder state
    __half *cross_kv = nullptr;      // [L][B][H][T_enc][128]: head-major, 64 K | 64 V per key
    __half *self_k = nullptr, *self_v = nullptr;   // [L][B][n_text_ctx][D]
    float *dx = nullptr, *dxn = nullptr, *dqkv = nullptr, *datt = nullptr, *dq = nullptr, *dh = nullptr, *logits = nullptr;
    __half *dxn16 = nullptr, *datt16 = nullptr, *dh16 = nullptr;   // f16 operands of the decoder GEMMs (rows > 8)
    int *d_tokens = nullptr, *d_next = nullptr, *d_sup_eot = nullptr, *d_sup_none = nullptr;
    int *h_next = nullptr;           // pinned [max_batch + 1]: language picks; [max_batch] = active-clip count read-back
    int *d_slot = nullptr, *d_nactive = nullptr, *d_count = nullptr, *d_budget = nullptr, *d_tokens_out = nullptr;   // greedy-loop bookkeeping
    void *lin_ws = nullptr;
    size_t lin_ws_bytes = 0;
    int64_t max_q = 8;               // rows per batch item in one decoder call (prompt <= 4 tokens)
    // decode-step megakernel state (wq4_decode.cu): decode-layout copies of the decoder linears, f16 logits matrix
    std::vector<DecLayerW> dec_w_host;
    bool dec1_ready = false;         // first-generation decode layout built (only when the second generation cannot run, or for A/B runs)
    Dec2Args dec2 = {};              // second-generation decode step (wq4_decode2.cu): geometry + per-layer slabs
    bool dec2_ready = false;
    float *dec2_partial = nullptr;   // [3][D] int64 fixed-point accumulators
    uint4 *tok_emb16t = nullptr;     // [D/8][Vp] f16 chunks
    __half *tok_emb16 = nullptr;     // [Vp256][D] f16 rows, zero padded: the dense-GEMM logits of batches above one decode group
    int64_t Vp256 = 0;
    int64_t Vp = 0;
    float *att_part = nullptr, *arg_scratch = nullptr;
    float *da_part = nullptr;      // split-KV partials of the per-op decode attention: [B*H][8][68]
    int *da_cnt = nullptr;         // arrival counters [B*H], zero between launches
    int64_t da_part_floats = 0;
    unsigned *grid_bar = nullptr;
    bool decode_kernel_ready = false;
};

namespace {

#define CHK(st_)                          \
    do {                                  \
        int32_t s__ = (st_);              \
        if (s__ != WQ4_OK) return s__;    \
    } while (0)

static int decode_chain_mode();
static bool decode_chain_enabled() { return decode_chain_mode() != 0; }
static int decode_chain_mode()
{
    // The chain kernel on decode steps of 17+ clips, measured on Large-V3 at 32 clips (3.50 ms with the tcgen05 GEMM + LayerNorm
    // launches, the default, mode 0):  2 = the linears between two attention kernels in ONE launch: 3.87 ms;  1 = one launch
    // per linear with the LayerNorm folded into the consumer's epilogue (no LayerNorm launches): 4.00 ms.  With <= 64 rows
    // every tile is a serial walk over K, so a dependent hop inside a launch costs what a PDL launch boundary costs, and the
    // 640-thread, 128-row-tile kernel has ~5 us more fixed cost per launch than q4_gemm2_kernel<64>.  Kept for A/B runs.
    static const int v = [] { const char *e = getenv("WQ4_DECODE_CHAIN"); return e ? atoi(e) : 0; }();
    return v;
}
static int decode_chain_fc2_pieces()
{
    static const int v = [] { const char *e = getenv("WQ4_DECODE_CHAIN_FC2"); int r = e ? atoi(e) : 1; return r < 1 ? 1 : r > 4 ? 4 : r; }();
    return decode_chain_mode() == 2 ? v : 1;
}
// mode 2: one launch; mode 1: one launch per problem (inputs then come from the predecessor kernel: dep = -1)
static int32_t launch_chain_maybe_split(ChainGemm *g, int n, int64_t rows, unsigned *ctr, int tm, cudaStream_t st)
{
    if (decode_chain_mode() == 2) return wq4_launch_chain(g, n, rows, ctr, tm, st);
    for (int i = 0; i < n; i++) {
        g[i].dep = -1; g[i].edep_plus1 = 0;
        int32_t s_ = wq4_launch_chain(g + i, 1, rows, ctr, tm, st);
        if (s_ != WQ4_OK) return s_;
    }
    return WQ4_OK;
}
static bool chain_split_k()
{
    // off by default: measured neutral at one Medium clip (2.166 against 2.148 ms per encoder pass): the extra tile epilogues cost what the shorter critical path saves
    static const bool v = [] { const char *e = getenv("WQ4_CHAIN_SPLITK"); return e && atoi(e) == 1; }();
    return v;
}
// WQ4_CHAIN=0 restores one launch per GEMM and LayerNorm in the encoder blocks (A/B runs)
static bool chain_enabled()
{
    static const bool v = [] { const char *e = getenv("WQ4_CHAIN"); return !(e && atoi(e) == 0); }();
    return v;
}

template <typename T>
int32_t dev_alloc(wq4_whisper *m, T **p, size_t count, bool workspace = false)
{
    void *q = nullptr;
    WQ4_CUDA(cudaMalloc(&q, count * sizeof(T) > 0 ? count * sizeof(T) : 16));
    (workspace ? m->ws_owned : m->owned).push_back(q);
    *p = reinterpret_cast<T *>(q);
    return WQ4_OK;
}
template <typename T>
int32_t ws_alloc(wq4_whisper *m, T **p, size_t count) { return dev_alloc(m, p, count, true); }

int32_t upload_f32(wq4_whisper *m, const Gguf &g, const std::string &name, float **out, uint64_t expect_elems,
                   std::vector<float> *host_copy = nullptr)
{
    const GgufTensor *t = g.find(name);
    if (!t) { wq4_set_error("Tensor '%s' not found", name.c_str()); return WQ4_ERR_INVALID; }
    const uint8_t *raw = g.tensor_data(*t);
    if (!raw) { wq4_set_error("Tensor '%s' data out of range", name.c_str()); return WQ4_ERR_BYTES; }
    uint64_t n = t->num_elements();
    if (expect_elems && n != expect_elems) {
        wq4_set_error("Tensor '%s' has %llu elements, expected %llu", name.c_str(), (unsigned long long)n, (unsigned long long)expect_elems);
        return WQ4_ERR_SHAPE;
    }
    std::vector<float> host(n);
    if (t->dtype == 0) memcpy(host.data(), raw, n * 4);
    else if (t->dtype == 1) { const uint16_t *h = reinterpret_cast<const uint16_t *>(raw); for (uint64_t i = 0; i < n; i++) host[i] = f16_bits_to_f32(h[i]); }
    else { wq4_set_error("Expected F32/F16 for '%s'", name.c_str()); return WQ4_ERR_INVALID; }   // loader.rs:75,108
    CHK(dev_alloc(m, out, n));
    WQ4_CUDA(cudaMemcpy(*out, host.data(), n * 4, cudaMemcpyHostToDevice));
    m->weight_bytes += n * 4;
    if (host_copy) host_copy->swap(host);
    return WQ4_OK;
}

// load_q4_linear (loader.rs:112-148) for one or several projections fused along N.
int32_t load_q4_fused(wq4_whisper *m, const Gguf &g, const std::vector<std::string> &prefixes, Lin *out)
{
    std::vector<uint8_t> bytes;
    std::vector<float> bias;
    bool any_bias = false;
    int64_t k = -1, n_total = 0;
    for (const auto &pfx : prefixes) {
        const std::string wn = pfx + ".weight", bn = pfx + ".bias";
        const GgufTensor *t = g.find(wn);
        if (!t) { wq4_set_error("Tensor '%s' not found", wn.c_str()); return WQ4_ERR_INVALID; }
        if (t->dtype != 2) { wq4_set_error("Expected Q4_0 for weight '%s'. Use the conversion script.", wn.c_str()); return WQ4_ERR_INVALID; }  // loader.rs:132-134
        if (t->dims.size() != 2) { wq4_set_error("Weight '%s' is not 2-D", wn.c_str()); return WQ4_ERR_SHAPE; }
        int64_t kk = (int64_t)t->dims[0], nn = (int64_t)t->dims[1];   // GGUF [k, n] -> [n, k] (loader.rs:126-131)
        if (k >= 0 && kk != k) { wq4_set_error("fused projections disagree on K"); return WQ4_ERR_SHAPE; }
        k = kk;
        const uint8_t *raw = g.tensor_data(*t);
        if (!raw) { wq4_set_error("Tensor '%s' data out of range", wn.c_str()); return WQ4_ERR_BYTES; }
        bytes.insert(bytes.end(), raw, raw + t->byte_size());
        bias.resize(n_total + nn, 0.0f);
        if (const GgufTensor *bt = g.find(bn)) {               // bias optional (loader.rs:137-145)
            const uint8_t *braw = g.tensor_data(*bt);
            if (!braw || bt->num_elements() != (uint64_t)nn) { wq4_set_error("Bias '%s' malformed", bn.c_str()); return WQ4_ERR_SHAPE; }
            if (bt->dtype == 0) memcpy(bias.data() + n_total, braw, nn * 4);
            else if (bt->dtype == 1) { const uint16_t *h = reinterpret_cast<const uint16_t *>(braw); for (int64_t i = 0; i < nn; i++) bias[n_total + i] = f16_bits_to_f32(h[i]); }
            else { wq4_set_error("Expected F32/F16 for '%s'", bn.c_str()); return WQ4_ERR_INVALID; }
            any_bias = true;
        }
        n_total += nn;
    }
    CHK(wq4_tensor_create(bytes.data(), bytes.size(), n_total, k, m->device, nullptr, &out->w));
    m->weight_bytes += wq4_tensor_device_bytes(out->w);
    out->bias = nullptr;
    if (any_bias) {
        CHK(dev_alloc(m, &out->bias, (size_t)n_total));
        WQ4_CUDA(cudaMemcpy(out->bias, bias.data(), n_total * 4, cudaMemcpyHostToDevice));
    }
    return WQ4_OK;
}

int32_t load_ln(wq4_whisper *m, const Gguf &g, const std::string &pfx, LNp *ln, uint64_t D)   // loader.rs:150-158
{
    CHK(upload_f32(m, g, pfx + ".weight", &ln->w, D));
    return upload_f32(m, g, pfx + ".bias", &ln->b, D);
}

int32_t load_conv(wq4_whisper *m, const Gguf &g, const std::string &pfx, wq4_conv1d **out, int64_t *c_out, int64_t *c_in)
{
    const GgufTensor *t = g.find(pfx + ".weight");
    if (!t || t->dims.size() != 3) { wq4_set_error("Tensor '%s.weight' not found or not 3-D", pfx.c_str()); return WQ4_ERR_INVALID; }
    int64_t ks = (int64_t)t->dims[0], ci = (int64_t)t->dims[1], co = (int64_t)t->dims[2];   // loader.rs:256-258
    std::vector<float> w, b;
    float *tmp = nullptr;
    CHK(upload_f32(m, g, pfx + ".weight", &tmp, (uint64_t)(ks * ci * co), &w));   // row-major [out, in, k] (loader.rs:273-276)
    CHK(upload_f32(m, g, pfx + ".bias", &tmp, (uint64_t)co, &b));
    CHK(wq4_conv1d_create(w.data(), b.data(), co, ci, ks, m->device, nullptr, out));
    *c_out = co; *c_in = ci;
    return WQ4_OK;
}

// c = W.gamma and b' = bias + W.beta of a LayerNorm folded into the linear that consumes it (wq4_chain.cu), by the f32 GEMV
int32_t ln_fold_vectors(wq4_whisper *m, const Lin &l, const LNp &ln, float **c_out, float **b_out)
{
    int64_t n, k;
    wq4_tensor_shape(l.w, &n, &k);
    CHK(dev_alloc(m, c_out, (size_t)n));
    CHK(dev_alloc(m, b_out, (size_t)n));
    LinearArgs a = {};
    a.w = l.w; a.bias = nullptr; a.x_f32 = ln.w; a.y_f32 = *c_out; a.rows = 1; a.stream = nullptr;
    CHK(wq4_launch_gemv(a));
    a.bias = l.bias; a.x_f32 = ln.b; a.y_f32 = *b_out;
    return wq4_launch_gemv(a);
}

int32_t load_model(wq4_whisper *m, const Gguf &g)
{
    const wq4_whisper_config &c = m->cfg;
    int64_t co, ci;
    CHK(load_conv(m, g, "encoder.conv1", &m->conv1, &co, &ci));           // loader.rs:291-296
    if (ci != c.n_mels) { wq4_set_error("conv1 expects %lld mel bins, config says %d", (long long)ci, c.n_mels); return WQ4_ERR_SHAPE; }
    m->D = co;
    CHK(load_conv(m, g, "encoder.conv2", &m->conv2, &co, &ci));           // loader.rs:297-302
    if (co != m->D || ci != m->D) { wq4_set_error("conv2 must be D x D"); return WQ4_ERR_SHAPE; }
    if (m->D != c.n_audio_state || m->D != c.n_text_state) { wq4_set_error("state size mismatch: tensors %lld, config %d/%d", (long long)m->D, c.n_audio_state, c.n_text_state); return WQ4_ERR_SHAPE; }
    if (m->D != (int64_t)c.n_audio_head * 64 || m->D != (int64_t)c.n_text_head * 64) { wq4_set_error("head_dim must be 64 (attention.rs:30)"); return WQ4_ERR_SHAPE; }
    const uint64_t D = (uint64_t)m->D;
    CHK(upload_f32(m, g, "encoder.positional_embedding", &m->enc_pos, 0));   // loader.rs:304
    {
        const GgufTensor *t = g.find("encoder.positional_embedding");
        if (t->dims.size() != 2 || t->dims[0] != D || (int64_t)t->dims[1] < m->T_enc) { wq4_set_error("encoder.positional_embedding must be [>=1500, D]"); return WQ4_ERR_SHAPE; }
    }
    m->enc.resize(c.n_audio_layer);
    for (int i = 0; i < c.n_audio_layer; i++) {                            // loader.rs:306-323
        const std::string p = "encoder.blocks." + std::to_string(i);
        EncLayer &L = m->enc[i];
        CHK(load_ln(m, g, p + ".attn_ln", &L.attn_ln, D));
        CHK(load_q4_fused(m, g, {p + ".attn.query", p + ".attn.key", p + ".attn.value"}, &L.qkv));   // loader.rs:160-191
        CHK(load_q4_fused(m, g, {p + ".attn.out"}, &L.out));
        CHK(load_ln(m, g, p + ".mlp_ln", &L.mlp_ln, D));
        CHK(load_q4_fused(m, g, {p + ".mlp.0"}, &L.fc1));                  // loader.rs:226-244
        CHK(load_q4_fused(m, g, {p + ".mlp.2"}, &L.fc2));
        int64_t n, k;
        wq4_tensor_shape(L.fc1.w, &n, &k);
        m->F = n;
    }
    CHK(load_ln(m, g, "encoder.ln_post", &m->ln_post, D));                 // loader.rs:325
    // LayerNorm fold vectors of the chained encoder block (wq4_chain.cu), f32 GEMV over the Q4 weights
    m->chain_ready = chain_enabled() && D % 128 == 0 && D <= 2048;
    if (m->chain_ready) {
        for (auto &L : m->enc) {
            CHK(ln_fold_vectors(m, L.qkv, L.attn_ln, &L.qkv_c, &L.qkv_b2));
            CHK(ln_fold_vectors(m, L.fc1, L.mlp_ln, &L.fc1_c, &L.fc1_b2));
        }
        WQ4_CUDA(cudaDeviceSynchronize());
    }
    CHK(upload_f32(m, g, "decoder.token_embedding.weight", &m->tok_emb, 0));   // loader.rs:338-339
    {
        const GgufTensor *t = g.find("decoder.token_embedding.weight");
        if (t->dims.size() != 2 || t->dims[0] != D) { wq4_set_error("token embedding must be [V, D]"); return WQ4_ERR_SHAPE; }
        m->V = (int64_t)t->dims[1];   // vocabulary size comes from the tensor (config.n_vocab is never read by the reference)
    }
    CHK(upload_f32(m, g, "decoder.positional_embedding", &m->dec_pos, 0));     // loader.rs:340-341
    {
        const GgufTensor *t = g.find("decoder.positional_embedding");
        if (t->dims.size() != 2 || t->dims[0] != D || (int64_t)t->dims[1] < c.n_text_ctx) { wq4_set_error("decoder.positional_embedding must be [>=n_text_ctx, D]"); return WQ4_ERR_SHAPE; }
    }
    m->dec.resize(c.n_text_layer);
    for (int i = 0; i < c.n_text_layer; i++) {                             // loader.rs:343-367
        const std::string p = "decoder.blocks." + std::to_string(i);
        DecLayer &L = m->dec[i];
        CHK(load_ln(m, g, p + ".attn_ln", &L.attn_ln, D));
        CHK(load_q4_fused(m, g, {p + ".attn.query", p + ".attn.key", p + ".attn.value"}, &L.qkv));
        CHK(load_q4_fused(m, g, {p + ".attn.out"}, &L.out));
        CHK(load_ln(m, g, p + ".cross_attn_ln", &L.cross_ln, D));
        CHK(load_q4_fused(m, g, {p + ".cross_attn.query"}, &L.cq));        // loader.rs:193-224 (key has no bias :208)
        CHK(load_q4_fused(m, g, {p + ".cross_attn.key", p + ".cross_attn.value"}, &L.ckv));
        CHK(load_q4_fused(m, g, {p + ".cross_attn.out"}, &L.cout));
        CHK(load_ln(m, g, p + ".mlp_ln", &L.mlp_ln, D));
        CHK(load_q4_fused(m, g, {p + ".mlp.0"}, &L.fc1));
        CHK(load_q4_fused(m, g, {p + ".mlp.2"}, &L.fc2));
    }
    CHK(load_ln(m, g, "decoder.ln", &m->ln, D));                           // loader.rs:369
    if (m->chain_ready) {
        for (auto &L : m->dec) {
            CHK(ln_fold_vectors(m, L.qkv, L.attn_ln, &L.qkv_c, &L.qkv_b2));
            CHK(ln_fold_vectors(m, L.cq, L.cross_ln, &L.cq_c, &L.cq_b2));
            CHK(ln_fold_vectors(m, L.fc1, L.mlp_ln, &L.fc1_c, &L.fc1_b2));
        }
        WQ4_CUDA(cudaDeviceSynchronize());
    }
    return WQ4_OK;
}

// Weight-derived decoder buffers, built once (the first reserve): decode-layout copies of the decoder linears for the
// one-kernel decode step, the f16 logits matrices.
static int64_t decode_grouped_max();
static bool force_decode_v1()
{
    static const bool v = [] { const char *e = getenv("WQ4_DECODE_V1"); return e && atoi(e) == 1; }();   // A/B runs: first-generation kernel
    return v;
}

int32_t prepare_decoder(wq4_whisper *m)
{
    if (m->decode_kernel_ready || m->dec.empty()) return WQ4_OK;
    const int64_t D = m->D;
    const int n_cta = wq4m_decode_grid(m->device);
    // second generation (one cluster per head): whenever the model's shape and the device allow it
    int cs = 0;
    if (!force_decode_v1() && m->dec.size() <= (size_t)kDecMaxLayers)
        CHK(wq4m_dec2_plan(m->device, D, m->F, m->cfg.n_text_head, m->T_enc, &cs));
    if (cs > 0) {
        wq4m_dec2_geometry(m->dec2, (int)D, (int)m->F, m->cfg.n_text_head, (int)m->T_enc, cs);
        for (size_t l = 0; l < m->dec.size(); l++) {
            DecLayer &L = m->dec[l];
            Dec2Layer &W = m->dec2.layers[l];
            CHK(wq4m_dec2_layer_create(m->dec2, L.qkv.w, L.out.w, L.cq.w, L.cout.w, L.fc1.w, L.fc2.w, &W, nullptr));
            m->owned.push_back(const_cast<uint8_t *>(W.wA)); m->owned.push_back(const_cast<uint8_t *>(W.wB)); m->owned.push_back(const_cast<uint8_t *>(W.wC));
            W.ln1w = L.attn_ln.w; W.ln1b = L.attn_ln.b; W.ln2w = L.cross_ln.w; W.ln2b = L.cross_ln.b; W.ln3w = L.mlp_ln.w; W.ln3b = L.mlp_ln.b;
            W.b_qkv = L.qkv.bias; W.b_out = L.out.bias; W.b_cq = L.cq.bias; W.b_cout = L.cout.bias; W.b_fc1 = L.fc1.bias; W.b_fc2 = L.fc2.bias;
        }
        m->dec2.n_layers = (int)m->dec.size();
        // [3][D] 64-bit fixed-point accumulators of the phase exchanges (wq4_decode2.cu), zero between launches
        CHK(dev_alloc(m, &m->dec2_partial, (size_t)3 * D * 2));
        WQ4_CUDA(cudaMemset(m->dec2_partial, 0, (size_t)3 * D * 8));
        m->dec2_ready = true;
    }
    // first generation: only when it can be reached (second generation unavailable, or grouped batches asked for by an override)
    if (!m->dec2_ready || decode_grouped_max() > 1) {
        m->dec_w_host.resize(m->dec.size());
        auto mk = [&](const Lin &l, DecWeight *out) -> int32_t {
            void *a1 = nullptr, *a2 = nullptr;
            CHK(wq4m_decw_create(l.w, l.bias, n_cta, out, &a1, &a2, nullptr));
            m->owned.push_back(a1); m->owned.push_back(a2);
            return WQ4_OK;
        };
        for (size_t l = 0; l < m->dec.size(); l++) {
            DecLayer &L = m->dec[l];
            DecLayerW &W = m->dec_w_host[l];
            W.ln1w = L.attn_ln.w; W.ln1b = L.attn_ln.b; W.ln2w = L.cross_ln.w; W.ln2b = L.cross_ln.b;
            W.ln3w = L.mlp_ln.w; W.ln3b = L.mlp_ln.b;
            CHK(mk(L.qkv, &W.qkv)); CHK(mk(L.out, &W.out)); CHK(mk(L.cq, &W.cq)); CHK(mk(L.cout, &W.cout));
            CHK(mk(L.fc1, &W.fc1)); CHK(mk(L.fc2, &W.fc2));
        }
        m->dec1_ready = true;
    }
    m->Vp = (m->V + 31) / 32 * 32;
    CHK(dev_alloc(m, &m->tok_emb16t, (size_t)m->Vp * (D / 8)));
    CHK(wq4m_logits_matrix(m->tok_emb, m->tok_emb16t, m->V, m->Vp, D, nullptr));
    CHK(dev_alloc(m, &m->att_part, (size_t)n_cta * 4 * 68));   // [cta][kMaxSeg][68]
    CHK(dev_alloc(m, &m->arg_scratch, (size_t)n_cta * 8 * 2));
    CHK(dev_alloc(m, &m->grid_bar, 4));
    WQ4_CUDA(cudaDeviceSynchronize());
    m->decode_kernel_ready = true;
    return WQ4_OK;
}

// Activation / KV-cache workspaces for up to B clips.  Grows on demand (the reference's WhisperModel has no batch
// ceiling): a larger request synchronises the device, frees the per-batch buffers and re-allocates them; the contents
// of the caches are NOT preserved across a growth (a transcription never spans one).  Weight-derived buffers stay.
int32_t reserve(wq4_whisper *m, int64_t B)
{
    if (B <= m->max_batch) return WQ4_OK;
    const int64_t D = m->D, F = m->F, T = m->T_enc, Ld = (int64_t)m->dec.size(), ctx = m->cfg.n_text_ctx;
    if (m->max_batch != 0) {
        WQ4_CUDA(cudaDeviceSynchronize());
        for (void *q : m->ws_owned) cudaFree(q);
        m->ws_owned.clear();
        if (m->h_next) { cudaFreeHost(m->h_next); m->h_next = nullptr; }
        m->max_batch = 0;
        m->kv_batch = 0;
        for (auto &kv : m->enc_graphs) if (kv.second.exec) cudaGraphExecDestroy(kv.second.exec);
        m->enc_graphs.clear();            // the graphs hold the old workspace pointers
    }
    m->Cp1 = (m->cfg.n_mels + 7) / 8 * 8;
    m->Tp1 = m->T_mel + 2;
    m->Tp2 = m->T_mel + 2;   // even, as the stride-2 view requires
    const int64_t M = B * T;
    CHK(ws_alloc(m, &m->mel_t16, (size_t)(B * m->Tp1 + 4) * m->Cp1));
    CHK(ws_alloc(m, &m->h1_16, (size_t)(B * m->Tp2 + 4) * D));
    WQ4_CUDA(cudaMemset(m->h1_16, 0, (size_t)(B * m->Tp2 + 4) * D * 2));   // pad rows stay zero forever
    CHK(ws_alloc(m, &m->x, (size_t)M * D));
    CHK(ws_alloc(m, &m->xn16, (size_t)M * D));
    CHK(ws_alloc(m, &m->qkv16, (size_t)M * 3 * D));
    CHK(ws_alloc(m, &m->att16, (size_t)M * D));
    CHK(ws_alloc(m, &m->ffn16, (size_t)M * F));
    CHK(ws_alloc(m, &m->enc16, (size_t)M * D));
    CHK(ws_alloc(m, &m->enc32, (size_t)M * D));
    if (m->chain_ready) {
        CHK(ws_alloc(m, &m->ln_stats_a, (size_t)(D / 128) * M));
        CHK(ws_alloc(m, &m->ln_stats_b, (size_t)(D / 128) * M));
        const size_t words = wq4_chain_counter_words(M);
        CHK(ws_alloc(m, &m->chain_ctr, words));
        WQ4_CUDA(cudaMemset(m->chain_ctr, 0, words * sizeof(unsigned)));
    }
    CHK(ws_alloc(m, &m->cross_kv, (size_t)Ld * M * 2 * D));
    CHK(ws_alloc(m, &m->self_k, (size_t)Ld * B * ctx * D));
    CHK(ws_alloc(m, &m->self_v, (size_t)Ld * B * ctx * D));
    const int64_t R = B * m->max_q;
    CHK(ws_alloc(m, &m->dx, (size_t)R * D));
    CHK(ws_alloc(m, &m->dxn, (size_t)R * D));
    CHK(ws_alloc(m, &m->dqkv, (size_t)R * 3 * D));
    CHK(ws_alloc(m, &m->datt, (size_t)R * D));
    CHK(ws_alloc(m, &m->dq, (size_t)R * D));
    CHK(ws_alloc(m, &m->dh, (size_t)R * F));
    CHK(ws_alloc(m, &m->dxn16, (size_t)R * D));
    CHK(ws_alloc(m, &m->datt16, (size_t)R * D));
    CHK(ws_alloc(m, &m->dh16, (size_t)R * F));
    if (m->chain_ready) {
        CHK(ws_alloc(m, &m->dln_stats_a, (size_t)(D / 128) * R));
        CHK(ws_alloc(m, &m->dln_stats_b, (size_t)(D / 128) * R));
        const size_t words = wq4_chain_counter_words(R);
        CHK(ws_alloc(m, &m->dchain_ctr, words));
        WQ4_CUDA(cudaMemset(m->dchain_ctr, 0, words * sizeof(unsigned)));
    }
    CHK(ws_alloc(m, &m->logits, (size_t)B * m->V));
    CHK(ws_alloc(m, &m->d_tokens, (size_t)R));
    CHK(ws_alloc(m, &m->d_next, (size_t)B));
    CHK(ws_alloc(m, &m->d_sup_eot, (size_t)B));
    CHK(ws_alloc(m, &m->d_sup_none, (size_t)B));
    std::vector<int> e((size_t)B, 50257), n((size_t)B, -1);
    WQ4_CUDA(cudaMemcpy(m->d_sup_eot, e.data(), B * 4, cudaMemcpyHostToDevice));
    WQ4_CUDA(cudaMemcpy(m->d_sup_none, n.data(), B * 4, cudaMemcpyHostToDevice));
    WQ4_CUDA(cudaMallocHost(&m->h_next, (size_t)(B + 1) * 4));
    CHK(ws_alloc(m, &m->d_slot, (size_t)B));
    CHK(ws_alloc(m, &m->d_nactive, (size_t)4));
    CHK(ws_alloc(m, &m->d_count, (size_t)B));
    CHK(ws_alloc(m, &m->d_budget, (size_t)B));
    CHK(ws_alloc(m, &m->d_tokens_out, (size_t)B * 224));
    m->lin_ws_bytes = (size_t)R * (size_t)F * 2 + 4096;   // f32 -> f16 copy of the widest decoder activation
    CHK(ws_alloc(m, reinterpret_cast<uint8_t **>(&m->lin_ws), m->lin_ws_bytes));
    if (!m->dec.empty()) {
        CHK(prepare_decoder(m));
        if (B > 1 && !m->tok_emb16) {   // batches beyond the single-kernel step: logits as one dense tensor-core GEMM on f16 rows
            m->Vp256 = (m->V + 255) / 256 * 256;
            CHK(dev_alloc(m, &m->tok_emb16, (size_t)m->Vp256 * D));
            WQ4_CUDA(cudaMemset(m->tok_emb16, 0, (size_t)m->Vp256 * D * 2));
            CHK(wq4_launch_cvt_f32_to_16(m->tok_emb, m->tok_emb16, m->V * D, false, nullptr));
        }
        m->da_part_floats = B * (int64_t)m->cfg.n_text_head * 8 * 68;
        CHK(ws_alloc(m, &m->da_part, (size_t)m->da_part_floats));
        CHK(ws_alloc(m, &m->da_cnt, (size_t)(B * m->cfg.n_text_head)));
        WQ4_CUDA(cudaMemset(m->da_cnt, 0, (size_t)(B * m->cfg.n_text_head) * sizeof(int)));
        WQ4_CUDA(cudaDeviceSynchronize());
    }
    m->max_batch = B;
    return WQ4_OK;
}

int32_t lin16(const Lin &l, const void *x16, float *y32, void *y16, const float *res, int64_t rows, uint32_t flags, cudaStream_t st)
{
    return wq4_linear_forward_ex(l.w, l.bias, nullptr, x16, y32, y16, res, rows, flags, nullptr, 0, st);
}
int32_t lin32(wq4_whisper *m, const Lin &l, const float *x, float *y, const float *res, int64_t rows, uint32_t flags, cudaStream_t st)
{
    return wq4_linear_forward_ex(l.w, l.bias, x, nullptr, y, nullptr, res, rows, flags, m->lin_ws, m->lin_ws_bytes, st);
}

// Skinny linear (wq4_skinny.cu) with its optional extras: LayerNorm fused on load, K/V append.
int32_t lin_sk(const Lin &l, const float *x32, const void *x16, const LNp *ln, float *y32, void *y16, const float *res, int64_t rows,
               uint32_t flags, SkinnyExtra ex, cudaStream_t st)
{
    LinearArgs a = {};
    a.w = l.w; a.bias = l.bias; a.x_f32 = x32; a.x_f16 = x16; a.y_f32 = y32; a.y_f16 = y16; a.residual = res;
    a.rows = rows; a.flags = flags; a.stream = st;
    if (ln) { ex.ln_w = ln->w; ex.ln_b = ln->b; }
    return wq4_launch_skinny(a, &ex);
}
// Largest batch whose decode step runs entirely on the skinny linear; WQ4_SKINNY_STEP_ROWS overrides it (tuning runs).
static int64_t skinny_step_rows()
{
    // function-local statics with an initialiser are initialised once, thread-safely (C++11)
    static const int64_t v = [] { const char *e = getenv("WQ4_SKINNY_STEP_ROWS"); int64_t r = e ? atoll(e) : WQ4_SKINNY_DEFAULT_ROWS; return r > WQ4_SKINNY_MAX_ROWS ? (int64_t)WQ4_SKINNY_MAX_ROWS : r; }();
    return v;
}
static bool use_skinny_step()
{
    static const bool v = [] { const char *e = getenv("WQ4_SKINNY"); return !(e && atoi(e) == 0); }();   // 0 = tcgen05 GEMM + separate LN / append (A/B runs)
    return v;
}

// WhisperEncoder::forward (encoder.rs:87-115) for B clips; mel_dev [B, n_mels, 3000] f32.
int32_t encode_front(wq4_whisper *m, const float *mel_dev, int64_t B, cudaStream_t st)
{
    // mel -> time-major f16 with one zero row before/after each clip (padding 1, layers.rs:83-90): the only kernel that
    // reads a caller pointer
    return wq4m_mel_transpose(mel_dev, m->mel_t16, B, m->cfg.n_mels, m->T_mel, m->Cp1, m->Tp1, 1, st);
}

int32_t encode_body(wq4_whisper *m, int64_t B, cudaStream_t st)
{
    const int64_t D = m->D, T = m->T_enc, M = B * T;
    const wq4_whisper_config &c = m->cfg;
    // conv1 (k=3, p=1, s=1) + GELU -> h1 [B][3002][D] f16, rows 1..3000 (encoder.rs:89-90)
    DenseGemmArgs a = {};
    a.w16 = m->conv1->w16; a.x16 = m->mel_t16; a.bias = m->conv1->bias; a.residual = nullptr;
    a.y_f32 = nullptr; a.y_f16 = m->h1_16 + D;
    a.n = D; a.n_pad = m->conv1->rows_pad; a.kdim_pad = m->conv1->kdim_pad;
    a.x_k = 3 * m->Cp1; a.x_rows = m->T_mel; a.x_batch = B; a.x_row_stride = m->Cp1; a.x_batch_stride = m->Tp1 * m->Cp1;
    a.ob = m->Tp2 * D; a.ot = D; a.on = 1; a.rb = 0;
    a.flags = WQ4_EPI_GELU; a.device = m->device; a.stream = st;
    CHK(wq4_launch_gemm_dense(a, wq4_pick_tile_m(m->T_mel, a.n_pad / WQ4_PANEL_ROWS * B)));
    // conv2 (k=3, p=1, s=2) + GELU + positional embedding -> x [B*1500][D] f32 (encoder.rs:93-106)
    DenseGemmArgs b2 = {};
    b2.w16 = m->conv2->w16; b2.x16 = m->h1_16; b2.bias = m->conv2->bias; b2.residual = m->enc_pos;
    b2.y_f32 = m->x; b2.y_f16 = nullptr;
    b2.n = D; b2.n_pad = m->conv2->rows_pad; b2.kdim_pad = m->conv2->kdim_pad;
    b2.x_k = 3 * D; b2.x_rows = T; b2.x_batch = B; b2.x_row_stride = 2 * D; b2.x_batch_stride = m->Tp2 * D;
    b2.ob = T * D; b2.ot = D; b2.on = 1; b2.rb = 0;
    b2.flags = WQ4_EPI_GELU; b2.device = m->device; b2.stream = st;
    CHK(wq4_launch_gemm_dense(b2, wq4_pick_tile_m(T, b2.n_pad / WQ4_PANEL_ROWS * B)));
    if (m->chain_ready && !m->enc.empty() && M >= 128) {
        // EncoderBlock::forward (encoder.rs:37-49) as two launches per block: attention, then ONE chained launch of
        // out-proj(+x) -> [LN2 folded] fc1+GELU -> fc2(+x) -> [next block's LN1 folded] q|k|v (wq4_chain.cu)
        const int tm = wq4_chain_pick_tile_m(M);
        const int groups = (int)(D / 128);
        CHK(wq4m_layernorm(m->x, m->enc[0].attn_ln.w, m->enc[0].attn_ln.b, nullptr, m->xn16, M, D, st));
        CHK(lin16(m->enc[0].qkv, m->xn16, nullptr, m->qkv16, nullptr, M, 0, st));
        for (size_t l = 0; l < m->enc.size(); l++) {
            const EncLayer &L = m->enc[l];
            const EncLayer *N = l + 1 < m->enc.size() ? &m->enc[l + 1] : nullptr;
            CHK(wq4m_enc_attention(m->qkv16, m->att16, B, T, D, c.n_audio_head, st));
            ChainGemm g[WQ4_CHAIN_MAX] = {};
            int n = 0;
            g[n].w = L.out.w; g[n].bias = L.out.bias; g[n].x16 = m->att16; g[n].residual = m->x; g[n].y32 = m->x;
            g[n].y16 = m->xn16; g[n].y16_scale = L.mlp_ln.w; g[n].stats_out = m->ln_stats_a; g[n].dep = -1;
            const int i_out = n++;
            g[n].w = L.fc1.w; g[n].bias = L.fc1_b2; g[n].cvec = L.fc1_c; g[n].x16 = m->xn16; g[n].y16 = m->ffn16;
            g[n].stats_in = m->ln_stats_a; g[n].stats_groups = groups; g[n].dep = i_out; g[n].flags = WQ4_EPI_GELU;
            const int i_fc1 = n++;
            // fc2: few column pairs and a long K.  When its tiles cannot fill the CTA pairs it runs as two K halves on
            // different pairs: the first writes x + bias + partial in place, the second adds its partial to that
            int64_t fn, fk;
            wq4_tensor_shape(L.fc2.w, &fn, &fk);
            const int64_t fc2_tiles = ((fn + 255) / 256) * ((M + tm - 1) / tm), kst = fk / 128;
            const bool split = chain_split_k() && fc2_tiles * 3 <= (int64_t)wq4_sm_count(m->device) && kst >= 16 && kst % 2 == 0;
            int i_fc2 = -1;
            if (split) {
                g[n].w = L.fc2.w; g[n].bias = L.fc2.bias; g[n].x16 = m->ffn16; g[n].residual = m->x; g[n].y32 = m->x; g[n].dep = i_fc1;
                g[n].k_stage0 = 0; g[n].k_stages = (int)(kst / 2);
                const int i_a = n++;
                g[n].w = L.fc2.w; g[n].bias = nullptr; g[n].x16 = m->ffn16; g[n].residual = m->x; g[n].y32 = m->x; g[n].dep = i_fc1;
                g[n].k_stage0 = (int)(kst / 2); g[n].k_stages = (int)(kst / 2); g[n].edep_plus1 = i_a + 1;
                i_fc2 = n++;
            } else {
                g[n].w = L.fc2.w; g[n].bias = L.fc2.bias; g[n].x16 = m->ffn16; g[n].residual = m->x; g[n].y32 = m->x; g[n].dep = i_fc1;
                i_fc2 = n++;
            }
            if (N) {
                g[i_fc2].y16 = m->xn16; g[i_fc2].y16_scale = N->attn_ln.w; g[i_fc2].stats_out = m->ln_stats_b;
                g[n].w = N->qkv.w; g[n].bias = N->qkv_b2; g[n].cvec = N->qkv_c; g[n].x16 = m->xn16; g[n].y16 = m->qkv16;
                g[n].stats_in = m->ln_stats_b; g[n].stats_groups = groups; g[n].dep = i_fc2;
                n++;
            }
            CHK(wq4_launch_chain(g, n, M, m->chain_ctr, tm, st));
        }
        return wq4m_layernorm(m->x, m->ln_post.w, m->ln_post.b, m->enc32, m->enc16, M, D, st);   // encoder.rs:114
    }
    for (auto &L : m->enc) {                                               // EncoderBlock::forward, encoder.rs:37-49
        CHK(wq4m_layernorm(m->x, L.attn_ln.w, L.attn_ln.b, nullptr, m->xn16, M, D, st));
        CHK(lin16(L.qkv, m->xn16, nullptr, m->qkv16, nullptr, M, 0, st));
        CHK(wq4m_enc_attention(m->qkv16, m->att16, B, T, D, c.n_audio_head, st));
        CHK(lin16(L.out, m->att16, m->x, nullptr, m->x, M, 0, st));        // residual + attn
        CHK(wq4m_layernorm(m->x, L.mlp_ln.w, L.mlp_ln.b, nullptr, m->xn16, M, D, st));
        CHK(lin16(L.fc1, m->xn16, nullptr, m->ffn16, nullptr, M, WQ4_EPI_GELU, st));
        CHK(lin16(L.fc2, m->ffn16, m->x, nullptr, m->x, M, 0, st));        // residual + ffn
    }
    return wq4m_layernorm(m->x, m->ln_post.w, m->ln_post.b, m->enc32, m->enc16, M, D, st);   // encoder.rs:114
}

int32_t encode(wq4_whisper *m, const float *mel_dev, int64_t B, cudaStream_t st)
{
    CHK(encode_front(m, mel_dev, B, st));
    return encode_body(m, B, st);
}

// WhisperModel::encode through the C ABI: everything after the mel transpose only touches model-owned buffers, so it is
// captured once per batch size into a CUDA graph (on the second call: the first one runs eagerly and sets up function
// attributes / occupancy queries / the error flag) and replayed -- one graph launch instead of ~54 kernel launches per
// Medium pass from the host.  Not used on the legacy default stream, inside a caller's own capture (the launches then go
// into the caller's graph), while the per-kernel profiler is on, or with WQ4_ENCODE_GRAPH=0.
int32_t encode_graphed(wq4_whisper *m, const float *mel_dev, int64_t B, cudaStream_t st)
{
    static const bool on = [] { const char *e = getenv("WQ4_ENCODE_GRAPH"); return !(e && atoi(e) == 0); }();
    cudaStreamCaptureStatus cs = cudaStreamCaptureStatusNone;
    bool plain = !on || st == nullptr || st == cudaStreamLegacy || st == cudaStreamPerThread || wq4_profile_active();
    if (!plain && (cudaStreamIsCapturing(st, &cs) != cudaSuccess || cs != cudaStreamCaptureStatusNone)) { cudaGetLastError(); plain = true; }
    if (plain) return encode(m, mel_dev, B, st);
    wq4_whisper::EncGraph &g = m->enc_graphs[B];
    if (g.failed) return encode(m, mel_dev, B, st);
    if (g.calls++ == 0) return encode(m, mel_dev, B, st);
    if (!g.exec) {
        const int64_t before = wq4_launch_count();
        cudaGraph_t graph = nullptr;
        if (cudaStreamBeginCapture(st, cudaStreamCaptureModeThreadLocal) != cudaSuccess) { cudaGetLastError(); g.failed = true; return encode(m, mel_dev, B, st); }
        const int32_t rc = encode_body(m, B, st);
        const cudaError_t ce = cudaStreamEndCapture(st, &graph);
        g.launches = wq4_launch_count() - before;
        wq4_count_launch(-(int)g.launches);                         // recorded into the graph, not launched
        if (rc != WQ4_OK || ce != cudaSuccess || !graph || cudaGraphInstantiate(&g.exec, graph, 0) != cudaSuccess) {
            cudaGetLastError();
            if (graph) cudaGraphDestroy(graph);
            g.exec = nullptr; g.failed = true;
            return rc != WQ4_OK ? rc : encode(m, mel_dev, B, st);
        }
        cudaGraphDestroy(graph);
    }
    CHK(encode_front(m, mel_dev, B, st));
    WQ4_CUDA(cudaGraphLaunch(g.exec, st));
    wq4_count_launch((int)g.launches);
    return WQ4_OK;
}

// Cross-attention K/V for every decoder layer from the encoder output (attention.rs:185-186):
// one fused [2D, D] GEMM at M = 1500*B per layer, stored f16.
int32_t cross_kv(wq4_whisper *m, int64_t B, cudaStream_t st)
{
    const int64_t D = m->D, M = B * m->T_enc;
    m->kv_batch = B;            // the caches of this transcription are [layer][B][..]; later steps may run fewer rows (compaction)
    for (size_t l = 0; l < m->dec.size(); l++)
    {
        // projection into the (now idle) encoder q|k|v buffer, then repacked head-major [B][H][T][128] for the decode steps
        CHK(lin16(m->dec[l].ckv, m->enc16, nullptr, m->qkv16, nullptr, M, 0, st));
        CHK(wq4m_cross_kv_pack(m->qkv16, m->cross_kv + l * (size_t)M * 2 * D, B, m->T_enc, m->cfg.n_text_head, st));
    }
    return WQ4_OK;
}

// One decoder pass over q_len new positions per clip (forward_prompt / decode_step, decoder.rs:251-348).
//   emb_pos0 : first positional-embedding row;  kv_pos0 : self-attention cache rows already valid
//   (the prompt pass overwrites from 0, decoder.rs:149-151).  Leaves logits of the LAST position.
struct ArgReq { int lo, hi; const int *suppress; int *out; };

static bool use_decode_kernel()
{
    static const bool v = [] { const char *e = getenv("WQ4_DECODE_IMPL"); return !(e && atoi(e) == 1); }();   // 1 = per-op launches (A/B runs)
    return v;
}

// One decode step as a single persistent kernel per group of up to 8 clips (wq4_decode.cu).  Larger batches run the
// kernel once per group, back to back on the stream: the caches are laid out [layer][B][..] so a group is a pointer
// offset, and the kernel takes the layer stride (kv_clips) separately from its own batch.
constexpr int64_t kDecodeGroup = 8;
// Largest batch that runs through the single-kernel step (in groups of 8 when an override allows more).  Measured
// (profiles/r01_decode_step.txt): the persistent kernel wins for a single clip (1.48 ms against ~1.7 ms); from 2 clips on
// the per-op path -- skinny linear with fused LayerNorm and cache append, weights prefetched under the previous kernel,
// key-per-lane attention kernel, 8 launches per layer -- is faster (1.76 against 1.84 ms at 2, 1.82 against 2.29 ms at 4,
// 2.17 against 3.04 ms at 8), so the default is 1; WQ4_DECODE_GROUP_MAX overrides it (tests).
// Smallest row count of the per-op path that runs on f16 operands (skinny linear / tcgen05 GEMM); below it the f32 GEMV.
// WQ4_DECODE_F16_MIN overrides it (tuning runs together with WQ4_DECODE_GROUP_MAX).
static int64_t decode_f16_min_rows()
{
    static const int64_t v = [] { const char *e = getenv("WQ4_DECODE_F16_MIN"); int64_t r = e ? atoll(e) : 2; return r < 1 ? (int64_t)1 : r; }();
    return v;
}

static int64_t decode_grouped_max()
{
    static const int64_t v = [] { const char *e = getenv("WQ4_DECODE_GROUP_MAX"); int64_t r = e ? atoll(e) : 1; return r < 1 ? (int64_t)1 : r; }();
    return v;
}

int32_t decode_step_fused(wq4_whisper *m, int64_t B, int64_t emb_pos, int64_t kv_pos0, const ArgReq *arg, cudaStream_t st, const int *slot)
{
    if (m->dec.size() > (size_t)kDecMaxLayers) { wq4_set_error("decode step kernel: more than %d decoder layers", kDecMaxLayers); return WQ4_ERR_UNSUPPORTED; }
    const int64_t D = m->D, F = m->F, ctx = m->cfg.n_text_ctx;
    if (B == 1 && m->dec2_ready) {           // second generation: one cluster per head, three grid-wide exchanges per layer
        Dec2Args a = m->dec2;
        a.V = (int)m->V; a.Vp = (int)m->Vp; a.ctx = (int)ctx;
        a.tokens = m->d_tokens; a.tok_emb = m->tok_emb; a.tok_emb16t = m->tok_emb16t; a.pos_emb = m->dec_pos;
        a.emb_pos = (int)emb_pos; a.kv_len = (int)kv_pos0;
        a.self_k = m->self_k; a.self_v = m->self_v; a.cross_kv = m->cross_kv; a.kv_clips = (int)m->kv_batch; a.slot = slot;
        a.lnfw = m->ln.w; a.lnfb = m->ln.b; a.logits = m->logits;
        a.arg_lo = arg ? arg->lo : 0; a.arg_hi = arg ? arg->hi : 0; a.suppress = (arg && arg->suppress) ? arg->suppress : nullptr;
        a.next_out = (arg && arg->out) ? arg->out : nullptr; a.arg_scratch = m->arg_scratch; a.partial = m->dec2_partial;
        a.barrier = m->grid_bar;
        a.dev_err = wq4_error_flag_ptr(m->device);
        if (!a.dev_err) return WQ4_ERR_CUDA;
        return wq4m_decode2_step(a, m->device, st);
    }
    if (!m->dec1_ready) { wq4_set_error("decode step kernel: first-generation layout not built"); return WQ4_ERR_UNSUPPORTED; }
    for (int64_t g0 = 0; g0 < B; g0 += kDecodeGroup) {
        const int64_t gb = B - g0 < kDecodeGroup ? B - g0 : kDecodeGroup;
        DecodeStepArgs a = {};
        for (size_t l = 0; l < m->dec.size(); l++) a.layers[l] = m->dec_w_host[l];
        a.n_layers = (int)m->dec.size(); a.B = (int)gb; a.kv_clips = (int)m->kv_batch; a.D = (int)D; a.F = (int)F;
        a.H = m->cfg.n_text_head; a.V = (int)m->V; a.ctx = (int)ctx; a.T_enc = (int)m->T_enc;
        a.tokens = m->d_tokens + g0; a.tok_emb = m->tok_emb; a.tok_emb16t = m->tok_emb16t; a.Vp = (int)m->Vp; a.pos_emb = m->dec_pos;
        a.emb_pos = (int)emb_pos; a.kv_len = (int)kv_pos0;
        a.x = m->dx + g0 * D; a.q = m->dq + g0 * D; a.h = m->dh + g0 * F; a.part = m->att_part;
        // with a slot table the kernel indexes the caches by clip itself; without one a group is a pointer offset
        const int64_t c0 = slot ? 0 : g0;
        a.self_k = m->self_k + c0 * ctx * D; a.self_v = m->self_v + c0 * ctx * D;
        a.cross_kv = m->cross_kv + c0 * m->T_enc * 2 * D;
        a.lnfw = m->ln.w; a.lnfb = m->ln.b; a.logits = m->logits + g0 * m->V;
        a.arg_lo = arg ? arg->lo : 0; a.arg_hi = arg ? arg->hi : 0; a.suppress = (arg && arg->suppress) ? arg->suppress + g0 : nullptr;
        a.next_out = (arg && arg->out) ? arg->out + g0 : nullptr; a.arg_scratch = m->arg_scratch; a.barrier = m->grid_bar;
        a.dev_err = wq4_error_flag_ptr(m->device);
        a.slot = slot ? slot + g0 : nullptr;
        if (!a.dev_err) return WQ4_ERR_CUDA;
        CHK(wq4m_decode_step(a, m->device, st));
    }
    return WQ4_OK;
}

// slot (device, or null): cache slot of batch row b, single-query steps only (active-set compaction of the greedy loop)
int32_t decoder_pass(wq4_whisper *m, int64_t B, int64_t q_len, int64_t emb_pos0, int64_t kv_pos0, const ArgReq *arg, cudaStream_t st,
                     const int *slot = nullptr)
{
    if (slot && q_len != 1) { wq4_set_error("decoder_pass: a slot table needs q_len == 1"); return WQ4_ERR_INVALID; }
    if (B > m->kv_batch) { wq4_set_error("decoder_pass: batch %lld exceeds the %lld clips the caches hold (encode / cross K/V first)", (long long)B, (long long)m->kv_batch); return WQ4_ERR_INVALID; }
    if (q_len == 1 && B <= decode_grouped_max() && m->decode_kernel_ready && use_decode_kernel() && m->dec.size() <= (size_t)kDecMaxLayers) {
        if (kv_pos0 + 1 > m->cfg.n_text_ctx || emb_pos0 + 1 > m->cfg.n_text_ctx) { wq4_set_error("decoder_pass: context overflow"); return WQ4_ERR_INVALID; }
        return decode_step_fused(m, B, emb_pos0, kv_pos0, arg, st, slot);
    }
    const int64_t D = m->D, F = m->F, R = B * q_len, ctx = m->cfg.n_text_ctx, H = m->cfg.n_text_head;
    const int64_t KB = m->kv_batch, M = KB * m->T_enc;          // layer strides of the caches
    if (q_len > m->max_q || kv_pos0 + q_len > ctx || emb_pos0 + q_len > ctx) { wq4_set_error("decoder_pass: context overflow"); return WQ4_ERR_INVALID; }
    CHK(wq4m_embed(m->d_tokens, m->tok_emb, m->dec_pos, m->dx, B, q_len, D, emb_pos0, st));
    // Single-query steps beyond the skinny kernel's reach (17+ clips): two chained launches per layer (wq4_chain.cu) --
    // out-proj(+x) -> [cross LN folded] cross-q, and cross-out(+x) -> [mlp LN] fc1+GELU -> fc2(+x) -> [next attn LN] q|k|v --
    // instead of six GEMM and three LayerNorm launches that each walk K serially on 10-40 SMs.
    const bool chained = m->chain_ready && m->dchain_ctr && q_len == 1 && m->da_part && R >= decode_f16_min_rows() &&
                         !(R <= skinny_step_rows() && use_skinny_step()) && decode_chain_enabled() && !m->dec.empty();
    if (chained) {
        const int groups = (int)(D / 128), tm = 128;
        CHK(wq4m_layernorm(m->dx, m->dec[0].attn_ln.w, m->dec[0].attn_ln.b, nullptr, m->dxn16, R, D, st));
        CHK(lin16(m->dec[0].qkv, m->dxn16, m->dqkv, nullptr, nullptr, R, 0, st));
        for (size_t l = 0; l < m->dec.size(); l++) {
            const DecLayer &L = m->dec[l];
            const DecLayer *N = l + 1 < m->dec.size() ? &m->dec[l + 1] : nullptr;
            __half *sk = m->self_k + l * (size_t)KB * ctx * D, *sv = m->self_v + l * (size_t)KB * ctx * D;
            const __half *ck = m->cross_kv + l * (size_t)M * 2 * D;
            CHK(wq4m_dec_attention_split(m->dqkv, 3 * D, sk, sv, D, ctx * D, 64, m->datt, m->datt16, D, B, H, kv_pos0 + 1,
                                         m->da_part, m->da_part_floats, m->da_cnt, m->device, st, D, slot));
            {
                ChainGemm g[2] = {};
                g[0].w = L.out.w; g[0].bias = L.out.bias; g[0].x16 = m->datt16; g[0].residual = m->dx; g[0].y32 = m->dx;
                g[0].y16 = m->dxn16; g[0].y16_scale = L.cross_ln.w; g[0].stats_out = m->dln_stats_a; g[0].dep = -1;
                g[1].w = L.cq.w; g[1].bias = L.cq_b2; g[1].cvec = L.cq_c; g[1].x16 = m->dxn16; g[1].y32 = m->dq;
                g[1].stats_in = m->dln_stats_a; g[1].stats_groups = groups; g[1].dep = 0;
                CHK(launch_chain_maybe_split(g, 2, R, m->dchain_ctr, tm, st));
            }
            CHK(wq4m_dec_attention_split(m->dq, D, ck, ck + 64, 128, m->T_enc * 2 * D, m->T_enc * 128, m->datt, m->datt16, D, B, H,
                                         m->T_enc, m->da_part, m->da_part_floats, m->da_cnt, m->device, st, -1, slot));
            {
                ChainGemm g[WQ4_CHAIN_MAX] = {};
                int n = 0;
                g[n].w = L.cout.w; g[n].bias = L.cout.bias; g[n].x16 = m->datt16; g[n].residual = m->dx; g[n].y32 = m->dx;
                g[n].y16 = m->dxn16; g[n].y16_scale = L.mlp_ln.w; g[n].stats_out = m->dln_stats_a; g[n].dep = -1;
                const int i_cout = n++;
                g[n].w = L.fc1.w; g[n].bias = L.fc1_b2; g[n].cvec = L.fc1_c; g[n].x16 = m->dxn16; g[n].y16 = m->dh16;
                g[n].stats_in = m->dln_stats_a; g[n].stats_groups = groups; g[n].dep = i_cout; g[n].flags = WQ4_EPI_GELU;
                const int i_fc1 = n++;
                // fc2 walks K = 4D: split into K pieces on different CTA pairs, chained through their epilogues
                int64_t fn, fk;
                wq4_tensor_shape(L.fc2.w, &fn, &fk);
                const int kst = (int)(fk / 128);
                int pieces = decode_chain_fc2_pieces();
                while (pieces > 1 && kst % pieces != 0) pieces--;
                int i_prev = -1;
                for (int pc = 0; pc < pieces; pc++) {
                    g[n].w = L.fc2.w; g[n].bias = pc == 0 ? L.fc2.bias : nullptr; g[n].x16 = m->dh16; g[n].residual = m->dx; g[n].y32 = m->dx;
                    g[n].dep = i_fc1; g[n].k_stage0 = pc * (kst / pieces); g[n].k_stages = kst / pieces; g[n].edep_plus1 = i_prev + 1;
                    i_prev = n++;
                }
                if (N) {
                    g[i_prev].y16 = m->dxn16; g[i_prev].y16_scale = N->attn_ln.w; g[i_prev].stats_out = m->dln_stats_b;
                    g[n].w = N->qkv.w; g[n].bias = N->qkv_b2; g[n].cvec = N->qkv_c; g[n].x16 = m->dxn16; g[n].y32 = m->dqkv;
                    g[n].stats_in = m->dln_stats_b; g[n].stats_groups = groups; g[n].dep = i_prev;
                    n++;
                }
                CHK(launch_chain_maybe_split(g, n, R, m->dchain_ctr, tm, st));
            }
        }
    }
    for (size_t l = 0; l < (chained ? (size_t)0 : m->dec.size()); l++) {
        DecLayer &L = m->dec[l];
        __half *sk = m->self_k + l * (size_t)KB * ctx * D, *sv = m->self_v + l * (size_t)KB * ctx * D;
        const __half *ck = m->cross_kv + l * (size_t)M * 2 * D;
        // Rows > 8 run on the tensor-core GEMM, whose operand is f16: the producers (LayerNorm, split attention, fc1's
        // epilogue) emit it directly; rows <= 8 use the f32 GEMV.  (f16p = single-query step on the GEMM path.)
        const bool g16 = R >= decode_f16_min_rows(), f16p = g16 && q_len == 1 && m->da_part;
        auto ln = [&](const LNp &w) { return wq4m_layernorm(m->dx, w.w, w.b, g16 ? nullptr : m->dxn, g16 ? m->dxn16 : nullptr, R, D, st); };
        auto lin_n = [&](const Lin &l, float *y32, void *y16, uint32_t flags) {       // input: the LayerNorm output
            return g16 ? lin16(l, m->dxn16, y32, y16, nullptr, R, flags, st) : lin32(m, l, m->dxn, y32, nullptr, R, flags, st);
        };
        auto lin_att = [&](const Lin &l) {                                             // input: attention output, + residual
            return f16p ? lin16(l, m->datt16, m->dx, nullptr, m->dx, R, 0, st) : lin32(m, l, m->datt, m->dx, m->dx, R, 0, st);
        };
        if (f16p && R <= skinny_step_rows() && use_skinny_step()) {
            // Single-query step of 9..16 clips: eight launches per layer (beyond 16 rows the tcgen05 GEMM is faster).  The skinny linear applies LayerNorm on load
            // (q|k|v, cross-q, fc1), appends the new K/V rows to the f16 cache from the q|k|v epilogue and adds the
            // residual in place; the split-KV attention emits the f16 operand of the following projection.
            SkinnyExtra none = {}, kv = {};
            kv.kc = sk; kv.vc = sv; kv.kv_batch_stride = ctx * D; kv.kv_ld = D; kv.kv_pos = kv_pos0; kv.kv_d = D; kv.kv_slot = slot;
            CHK(lin_sk(L.qkv, m->dx, nullptr, &L.attn_ln, m->dqkv, nullptr, nullptr, R, 0, kv, st));
            CHK(wq4m_dec_attention_split(m->dqkv, 3 * D, sk, sv, D, ctx * D, 64, m->datt, m->datt16, D, B, H, kv_pos0 + 1,
                                         m->da_part, m->da_part_floats, m->da_cnt, m->device, st, 0, slot));
            CHK(lin_sk(L.out, nullptr, m->datt16, nullptr, m->dx, nullptr, m->dx, R, 0, none, st));
            CHK(lin_sk(L.cq, m->dx, nullptr, &L.cross_ln, m->dq, nullptr, nullptr, R, 0, none, st));
            CHK(wq4m_dec_attention_split(m->dq, D, ck, ck + 64, 128, m->T_enc * 2 * D, m->T_enc * 128, m->datt, m->datt16, D, B, H,
                                         m->T_enc, m->da_part, m->da_part_floats, m->da_cnt, m->device, st, -1, slot));
            CHK(lin_sk(L.cout, nullptr, m->datt16, nullptr, m->dx, nullptr, m->dx, R, 0, none, st));
            CHK(lin_sk(L.fc1, m->dx, nullptr, &L.mlp_ln, nullptr, m->dh16, nullptr, R, WQ4_EPI_GELU, none, st));
            CHK(lin_sk(L.fc2, nullptr, m->dh16, nullptr, m->dx, nullptr, m->dx, R, 0, none, st));
            continue;
        }
        // self-attention with cache (decoder.rs:84-92 / 147-152)
        CHK(ln(L.attn_ln));
        CHK(lin_n(L.qkv, m->dqkv, nullptr, 0));
        if (q_len == 1 && m->da_part)           // the attention kernel appends the new K/V row itself
            CHK(wq4m_dec_attention_split(m->dqkv, 3 * D, sk, sv, D, ctx * D, 64, m->datt, f16p ? m->datt16 : nullptr, D, B, H, kv_pos0 + 1,
                                         m->da_part, m->da_part_floats, m->da_cnt, m->device, st, D, slot));
        else {
            CHK(wq4m_kv_append(m->dqkv, 3 * D, sk, sv, D, ctx * D, B, D, q_len, kv_pos0, st));
            CHK(wq4m_dec_attention(m->dqkv, 3 * D, sk, sv, D, ctx * D, 64, m->datt, D, B, H, q_len, kv_pos0 + q_len,
                                   q_len > 1 ? (int)kv_pos0 : -1, st));
        }
        CHK(lin_att(L.out));
        // cross-attention over the cached encoder K/V (decoder.rs:95-104)
        CHK(ln(L.cross_ln));
        CHK(lin_n(L.cq, m->dq, nullptr, 0));
        if (q_len == 1 && m->da_part)
            CHK(wq4m_dec_attention_split(m->dq, D, ck, ck + 64, 128, m->T_enc * 2 * D, m->T_enc * 128, m->datt, f16p ? m->datt16 : nullptr, D, B, H,
                                         m->T_enc, m->da_part, m->da_part_floats, m->da_cnt, m->device, st, -1, slot));
        else
            CHK(wq4m_dec_attention(m->dq, D, ck, ck + 64, 128, m->T_enc * 2 * D, m->T_enc * 128, m->datt, D, B, H, q_len, m->T_enc, -1, st));
        CHK(lin_att(L.cout));
        // FFN (decoder.rs:107-110)
        CHK(ln(L.mlp_ln));
        if (g16) {
            CHK(lin16(L.fc1, m->dxn16, nullptr, m->dh16, nullptr, R, WQ4_EPI_GELU, st));
            // fc2 (K = 4D) at 17..32 rows: the skinny kernel's 16-warp K split (11.6 us) beats the serial K walk of the
            // tcgen05 kernel on 10 SMs (16.4 us, profiles/r01_microbench_skinny.txt)
            if (R <= 32 && L.fc2.w->k >= 4096 && use_skinny_step())
                CHK(lin_sk(L.fc2, nullptr, m->dh16, nullptr, m->dx, nullptr, m->dx, R, 0, SkinnyExtra{}, st));
            else
                CHK(lin16(L.fc2, m->dh16, m->dx, nullptr, m->dx, R, 0, st));
        } else {
            CHK(lin32(m, L.fc1, m->dxn, m->dh, nullptr, R, WQ4_EPI_GELU, st));
            CHK(lin32(m, L.fc2, m->dh, m->dx, m->dx, R, 0, st));
        }
    }
    // final LN on the last position of every clip, then logits = x . token_embedding^T (decoder.rs:286-292)
    if (R >= decode_f16_min_rows() && m->tok_emb16) {
        // batches above one decode group: the logits are one tensor-core GEMM on the f16 embedding rows
        const void *last16 = m->dxn16;
        if (q_len == 1) {
            CHK(wq4m_layernorm(m->dx, m->ln.w, m->ln.b, nullptr, m->dxn16, B, D, st));
        } else {
            CHK(wq4m_layernorm(m->dx, m->ln.w, m->ln.b, nullptr, m->dxn16, R, D, st));
            WQ4_CUDA(cudaMemcpy2DAsync(m->datt16, D * 2, m->dxn16 + (q_len - 1) * D, q_len * D * 2, D * 2, B, cudaMemcpyDeviceToDevice, st));
            last16 = m->datt16;
        }
        DenseGemmArgs ga = {};
        ga.w16 = m->tok_emb16; ga.x16 = last16; ga.y_f32 = m->logits;
        ga.n = m->V; ga.n_pad = m->Vp256; ga.kdim_pad = D;
        ga.x_k = D; ga.x_rows = B; ga.x_batch = 1; ga.x_row_stride = D; ga.x_batch_stride = B * D;
        ga.ob = B * m->V; ga.ot = m->V; ga.on = 1;
        ga.device = m->device; ga.stream = st;
        CHK(wq4_launch_gemm_dense(ga, B <= 32 ? 32 : 64));   // row tile of the first-generation kernel
    } else {
        if (q_len == 1) {
            CHK(wq4m_layernorm(m->dx, m->ln.w, m->ln.b, m->dxn, nullptr, B, D, st));
        } else {
            CHK(wq4m_layernorm(m->dx, m->ln.w, m->ln.b, m->dxn, nullptr, R, D, st));
            // gather last rows: dxn[b*q_len + q_len-1] -> dq[b]
            WQ4_CUDA(cudaMemcpy2DAsync(m->dq, D * 4, m->dxn + (q_len - 1) * D, q_len * D * 4, D * 4, B, cudaMemcpyDeviceToDevice, st));
        }
        const float *last = q_len == 1 ? m->dxn : m->dq;
        CHK(wq4m_logits(last, m->tok_emb, m->logits, B, D, m->V, st));
    }
    (void)F;
    if (arg && arg->hi > arg->lo) CHK(wq4m_argmax(m->logits, B, m->V, arg->lo, arg->hi, arg->suppress, arg->out, st));
    return WQ4_OK;
}

}  // namespace

extern "C" {

int32_t wq4_whisper_config_preset(const char *name, wq4_whisper_config *cfg)
{
    if (!name || !cfg) { wq4_set_error("wq4_whisper_config_preset: null argument"); return WQ4_ERR_INVALID; }
    wq4_whisper_config c = {};
    c.n_audio_ctx = 1500; c.n_text_ctx = 448; c.n_vocab = 51865;
    if (!strcmp(name, "large_v3")) {          // config.rs:32-46
        c.n_mels = 128; c.n_audio_state = 1280; c.n_audio_head = 20; c.n_audio_layer = 32;
        c.n_text_state = 1280; c.n_text_head = 20; c.n_text_layer = 32; c.n_lang = 100;
    } else if (!strcmp(name, "medium")) {     // config.rs:49-63
        c.n_mels = 80; c.n_audio_state = 1024; c.n_audio_head = 16; c.n_audio_layer = 24;
        c.n_text_state = 1024; c.n_text_head = 16; c.n_text_layer = 24; c.n_lang = 99;
    } else {
        wq4_set_error("unknown preset '%s' (large_v3 | medium)", name);
        return WQ4_ERR_INVALID;
    }
    *cfg = c;
    return WQ4_OK;
}

int32_t wq4_whisper_load_gguf_memory(const uint8_t *data, size_t nbytes, const wq4_whisper_config *cfg, int32_t device,
                                     wq4_whisper **out)
{
    if (!data || !cfg || !out) { wq4_set_error("wq4_whisper_load_gguf_memory: null argument"); return WQ4_ERR_INVALID; }
    *out = nullptr;
    Gguf g;
    if (!g.parse(data, nbytes)) { wq4_set_error("Failed to parse GGUF: %s", g.err.c_str()); return WQ4_ERR_INVALID; }
    int ndev = 0;
    if (cudaGetDeviceCount(&ndev) != cudaSuccess || device < 0 || device >= ndev) {
        cudaGetLastError();
        wq4_set_error("wq4_whisper_load: CUDA device %d not available (%d visible); there is no CPU fallback", device, ndev);
        return WQ4_ERR_CUDA;
    }
    DeviceGuard guard(device);
    wq4_whisper *m = new wq4_whisper();
    m->cfg = *cfg;
    m->device = device;
    m->T_enc = cfg->n_audio_ctx;
    m->T_mel = 2 * (int64_t)cfg->n_audio_ctx;
    int32_t st = load_model(m, g);
    if (st != WQ4_OK) { wq4_whisper_destroy(m); return st; }
    *out = m;
    return WQ4_OK;
}

int32_t wq4_whisper_load_gguf(const char *path, const wq4_whisper_config *cfg, int32_t device, wq4_whisper **out)
{
    if (!path || !cfg || !out) { wq4_set_error("wq4_whisper_load_gguf: null argument"); return WQ4_ERR_INVALID; }
    // The file is mapped, not read: tensor bytes go from the page cache straight into the driver's staging copies
    // (the reference copies every tensor twice on the host first: reader.rs:220-221, tensor.rs:54-61).
    const int fd = open(path, O_RDONLY);
    if (fd < 0) { wq4_set_error("Failed to open GGUF file: %s", path); return WQ4_ERR_INVALID; }   // loader.rs:31-32
    struct stat sb;
    if (fstat(fd, &sb) != 0 || sb.st_size <= 0) { close(fd); wq4_set_error("cannot stat %s", path); return WQ4_ERR_INVALID; }
    const size_t n = (size_t)sb.st_size;
    void *map = mmap(nullptr, n, PROT_READ, MAP_PRIVATE, fd, 0);
    close(fd);
    if (map == MAP_FAILED) { wq4_set_error("mmap failed on %s", path); return WQ4_ERR_INVALID; }
    madvise(map, n, MADV_SEQUENTIAL);
    madvise(map, n, MADV_WILLNEED);
    const int32_t st = wq4_whisper_load_gguf_memory(static_cast<const uint8_t *>(map), n, cfg, device, out);
    munmap(map, n);
    return st;
}

int32_t wq4_whisper_destroy(wq4_whisper *m)
{
    if (!m) return WQ4_OK;
    DeviceGuard g(m->device);
    for (auto &L : m->enc) { wq4_tensor_destroy(L.qkv.w); wq4_tensor_destroy(L.out.w); wq4_tensor_destroy(L.fc1.w); wq4_tensor_destroy(L.fc2.w); }
    for (auto &L : m->dec) {
        wq4_tensor_destroy(L.qkv.w); wq4_tensor_destroy(L.out.w); wq4_tensor_destroy(L.cq.w); wq4_tensor_destroy(L.ckv.w);
        wq4_tensor_destroy(L.cout.w); wq4_tensor_destroy(L.fc1.w); wq4_tensor_destroy(L.fc2.w);
    }
    wq4_conv1d_destroy(m->conv1);
    wq4_conv1d_destroy(m->conv2);
    for (auto &kv : m->enc_graphs) if (kv.second.exec) cudaGraphExecDestroy(kv.second.exec);
    for (void *p : m->owned) cudaFree(p);
    for (void *p : m->ws_owned) cudaFree(p);
    if (m->h_next) cudaFreeHost(m->h_next);
    delete m;
    return WQ4_OK;
}

int32_t wq4_whisper_reserve(wq4_whisper *m, int64_t max_batch)
{
    if (!m || max_batch <= 0) { wq4_set_error("wq4_whisper_reserve: invalid argument"); return WQ4_ERR_INVALID; }
    DeviceGuard g(m->device);
    return reserve(m, max_batch);
}

int32_t wq4_whisper_info(const wq4_whisper *m, int64_t *d_model, int64_t *ffn, int64_t *vocab, int64_t *weight_bytes)
{
    if (!m) { wq4_set_error("wq4_whisper_info: null model"); return WQ4_ERR_INVALID; }
    if (d_model) *d_model = m->D;
    if (ffn) *ffn = m->F;
    if (vocab) *vocab = m->V;
    if (weight_bytes) *weight_bytes = (int64_t)m->weight_bytes;
    return WQ4_OK;
}

int32_t wq4_whisper_encode(wq4_whisper *m, const float *mel_dev, int64_t batch, float *enc_out_dev, wq4_stream stream)
{
    if (!m || !mel_dev || batch <= 0) { wq4_set_error("wq4_whisper_encode: invalid argument"); return WQ4_ERR_INVALID; }
    DeviceGuard g(m->device);
    CHK(reserve(m, batch));
    if (batch > m->max_batch) { wq4_set_error("wq4_whisper_encode: batch %lld exceeds the reserved %lld", (long long)batch, (long long)m->max_batch); return WQ4_ERR_INVALID; }
    cudaStream_t st = (cudaStream_t)stream;
    CHK(encode_graphed(m, mel_dev, batch, st));
    if (enc_out_dev)
        WQ4_CUDA(cudaMemcpyAsync(enc_out_dev, m->enc32, (size_t)batch * m->T_enc * m->D * 4, cudaMemcpyDeviceToDevice, st));
    return WQ4_OK;
}

// Reads the sticky device error flag (pipeline / grid-barrier timeouts raise it and the kernel carries on with unsynchronised
// data): called at the synchronisation points of transcribe so that a timeout is an error, never silently wrong tokens.
static int32_t check_device_flag(wq4_whisper *m, const char *who)
{
    int *p = wq4_error_flag_ptr(m->device);
    if (!p) return WQ4_ERR_CUDA;
    int v = 0;
    WQ4_CUDA(cudaMemcpy(&v, p, sizeof(int), cudaMemcpyDeviceToHost));
    if (v != 0) { wq4_set_error("%s: device error flag %d (a kernel-side wait timed out); results are invalid", who, v); return WQ4_ERR_CUDA; }
    return WQ4_OK;
}

// WhisperModel::transcribe (whisper.rs:51-128) for a batch of clips that share the prompt mode.
//   lang_tokens[b] >= 0 : explicit language -> prompt [SOT, lang, TRANSCRIBE, NO_TIMESTAMPS] (whisper.rs:67-71)
//   lang_tokens[b] <  0 : auto-detect       -> decode_step(SOT, 0), argmax over the language range, then
//                         prompt [lang, TRANSCRIBE, NO_TIMESTAMPS] at positions 0..3 with the self-attention
//                         cache overwritten (whisper.rs:72-88, decoder.rs:149-151,272-275) -- reproduced as is.
// All clips of one call must use the same mode.  tokens_out: [batch][224] (host), counts_out: [batch] (host).
// max_new_tokens (host, may be NULL): per-clip token budget, clamped to MAX_TOKENS = 224 (whisper.rs:20) -- a batching
// extension; NULL is the reference's behaviour.
//
// The loop is device-driven: the greedy pick, the EOT / budget test (whisper.rs:105-108), the token append and the
// compaction of the active set all run in one small bookkeeping kernel per step (greedy_bookkeep_kernel), so the host
// enqueues steps without reading anything back.  Every kSyncEvery steps it reads the number of clips still active and
// shrinks the row count of the following launches to it: a retired clip stops occupying GEMV rows and stops streaming
// its 7.7 MB x layers of cross K/V.  Tokens come back in one copy at the end.
int32_t wq4_whisper_transcribe_ex(wq4_whisper *m, const float *mel_dev, int64_t batch, const int32_t *lang_tokens,
                                  const int32_t *max_new_tokens, int32_t *tokens_out, int32_t *counts_out,
                                  int32_t *detected_lang_out, wq4_stream stream)
{
    if (!m || !mel_dev || !lang_tokens || !tokens_out || !counts_out || batch <= 0) { wq4_set_error("wq4_whisper_transcribe: invalid argument"); return WQ4_ERR_INVALID; }
    if (batch > 1024) { wq4_set_error("wq4_whisper_transcribe: at most 1024 clips per call (got %lld); split the batch", (long long)batch); return WQ4_ERR_UNSUPPORTED; }
    if (m->dec.empty()) { wq4_set_error("wq4_whisper_transcribe: the model was loaded without decoder layers"); return WQ4_ERR_INVALID; }
    DeviceGuard g(m->device);
    CHK(reserve(m, batch));
    const int64_t B = batch;
    cudaStream_t st = (cudaStream_t)stream;
    const wq4_whisper_config &c = m->cfg;
    const int SOT = 50258, EOT = 50257, MAX_TOKENS = 224, MIN_TOKENS = 3;   // whisper.rs:15-20,97
    const int kSyncEvery = 8;
    const int transcribe_tok = 50260 + c.n_lang, no_ts_tok = transcribe_tok + 4;   // config.rs:66-75
    const bool autodetect = lang_tokens[0] < 0;
    for (int64_t b = 0; b < B; b++)
        if ((lang_tokens[b] < 0) != autodetect) { wq4_set_error("wq4_whisper_transcribe: mixed prompt modes in one batch"); return WQ4_ERR_INVALID; }
    std::vector<int> budget((size_t)B, MAX_TOKENS);
    if (max_new_tokens)
        for (int64_t b = 0; b < B; b++) budget[b] = max_new_tokens[b] < 0 ? 0 : (max_new_tokens[b] > MAX_TOKENS ? MAX_TOKENS : max_new_tokens[b]);

    CHK(encode(m, mel_dev, B, st));
    CHK(cross_kv(m, B, st));

    std::vector<int> prompt;   // [B][q_len]
    int64_t position = 0, q_len;
    std::vector<int> langs((size_t)B);
    if (!autodetect) {
        q_len = 4;
        prompt.resize((size_t)B * 4);
        for (int64_t b = 0; b < B; b++) {
            langs[b] = lang_tokens[b];
            int *p = &prompt[(size_t)b * 4];
            p[0] = SOT; p[1] = lang_tokens[b]; p[2] = transcribe_tok; p[3] = no_ts_tok;
        }
    } else {
        std::vector<int> sot((size_t)B, SOT);
        WQ4_CUDA(cudaMemcpyAsync(m->d_tokens, sot.data(), B * 4, cudaMemcpyHostToDevice, st));
        ArgReq lang_req = {50259, 50259 + c.n_lang, nullptr, m->d_next};      // whisper.rs:77-83
        CHK(decoder_pass(m, B, 1, 0, 0, &lang_req, st));                   // whisper.rs:73
        position = 1;
        WQ4_CUDA(cudaMemcpyAsync(m->h_next, m->d_next, B * 4, cudaMemcpyDeviceToHost, st));
        WQ4_CUDA(cudaStreamSynchronize(st));
        q_len = 3;
        prompt.resize((size_t)B * 3);
        for (int64_t b = 0; b < B; b++) {
            langs[b] = m->h_next[b];
            int *p = &prompt[(size_t)b * 3];
            p[0] = langs[b]; p[1] = transcribe_tok; p[2] = no_ts_tok;
        }
    }
    if (detected_lang_out) for (int64_t b = 0; b < B; b++) detected_lang_out[b] = langs[b];
    WQ4_CUDA(cudaMemcpyAsync(m->d_tokens, prompt.data(), prompt.size() * 4, cudaMemcpyHostToDevice, st));
    WQ4_CUDA(cudaMemcpyAsync(m->d_budget, budget.data(), (size_t)B * 4, cudaMemcpyHostToDevice, st));
    CHK(wq4m_greedy_init(m->d_slot, m->d_nactive, m->d_count, B, st));
    ArgReq first_req = {0, (int)m->V, m->d_sup_eot, m->d_next};            // whisper.rs:98-99
    CHK(decoder_pass(m, B, q_len, 0, 0, &first_req, st));                  // forward_prompt: positions restart at 0
    position += q_len;                                                     // whisper.rs:93
    int64_t kv_len = q_len;                                                // cache now holds the prompt only

    int64_t rows = B;                                                      // rows the next launches run (>= clips still active)
    int *h_nactive = m->h_next + m->max_batch;
    for (int step = 0; step < MAX_TOKENS; step++) {                        // whisper.rs:104-125
        // consume the pick of the previous pass: append / retire / compact, and stage the next step's input tokens
        CHK(wq4m_greedy_bookkeep(m->d_next, m->d_tokens, m->d_slot, m->d_nactive, m->d_count, m->d_budget, m->d_tokens_out,
                                 MAX_TOKENS, EOT, st));
        if (step == MAX_TOKENS - 1) break;   // the reference computes one more step whose result it never uses
        if (position >= c.n_text_ctx || kv_len >= c.n_text_ctx) break;
        bool look = step % kSyncEvery == kSyncEvery - 1;
        if (max_new_tokens)                     // a clip that has just used up its budget retires now: shrink at once
            for (int64_t b = 0; b < B && !look; b++) look = budget[b] == step + 1 || (step == 0 && budget[b] == 0);
        if (look) {
            WQ4_CUDA(cudaMemcpyAsync(h_nactive, m->d_nactive, 4, cudaMemcpyDeviceToHost, st));
            WQ4_CUDA(cudaStreamSynchronize(st));
            rows = *h_nactive;
            if (rows <= 0) break;
        }
        const int *sup = (step + 1 < MIN_TOKENS) ? m->d_sup_eot : m->d_sup_none;   // whisper.rs:120-122
        ArgReq req = {0, (int)m->V, sup, m->d_next};
        CHK(decoder_pass(m, rows, 1, position, kv_len, &req, st, m->d_slot));       // whisper.rs:116
        position += 1;
        kv_len += 1;
    }
    std::vector<int> toks((size_t)B * MAX_TOKENS), count((size_t)B);
    WQ4_CUDA(cudaMemcpyAsync(toks.data(), m->d_tokens_out, toks.size() * 4, cudaMemcpyDeviceToHost, st));
    WQ4_CUDA(cudaMemcpyAsync(count.data(), m->d_count, (size_t)B * 4, cudaMemcpyDeviceToHost, st));
    WQ4_CUDA(cudaStreamSynchronize(st));
    CHK(check_device_flag(m, "wq4_whisper_transcribe"));
    for (int64_t b = 0; b < B; b++) {
        counts_out[b] = count[b];
        memcpy(tokens_out + (size_t)b * MAX_TOKENS, toks.data() + (size_t)b * MAX_TOKENS, (size_t)count[b] * 4);
    }
    return WQ4_OK;
}

int32_t wq4_whisper_transcribe(wq4_whisper *m, const float *mel_dev, int64_t batch, const int32_t *lang_tokens,
                               int32_t *tokens_out, int32_t *counts_out, int32_t *detected_lang_out, wq4_stream stream)
{
    return wq4_whisper_transcribe_ex(m, mel_dev, batch, lang_tokens, nullptr, tokens_out, counts_out, detected_lang_out, stream);
}

// transcribe() of src/transcribe.rs:34-107 for a batch of utterances, from samples to text-token ids:
// pad / truncate every clip to 480 000 samples (:46-52), log-mel [n_mels, 3000] (:55-76), WhisperModel::transcribe (:86),
// drop ids >= 50257 (:96-100).  Resampling (:38-43) and the tokenizer's decode (:102-105) stay on the host side of the
// reference (CPU code off the hot path): pcm must be 16 kHz mono f32.
//   pcm_dev   : device, clip b at pcm_dev + b*pcm_stride, n_samples[b] samples (host array; longer than 30 s is truncated)
//   language  : [batch] language token per clip (>= 0) or < 0 for auto-detect; one mode per call
//   text_tokens_out [batch][224] / text_counts_out [batch] (host): ids < 50257 only;  raw_* (may be NULL): everything emitted
//   elapsed_ms_out (may be NULL): wall time of the model part (encoder + decode loop), as transcribe.rs:84-87 measures it
int32_t wq4_transcribe(wq4_whisper *m, const wq4_logmel *mel, const float *pcm_dev, int64_t pcm_stride, const int64_t *n_samples,
                       int64_t batch, const int32_t *language, int32_t *text_tokens_out, int32_t *text_counts_out,
                       int32_t *raw_tokens_out, int32_t *raw_counts_out, double *elapsed_ms_out, wq4_stream stream)
{
    if (!m || !mel || !pcm_dev || !n_samples || !language || !text_tokens_out || !text_counts_out || batch <= 0 || pcm_stride <= 0) {
        wq4_set_error("wq4_transcribe: invalid argument");
        return WQ4_ERR_INVALID;
    }
    for (int64_t b = 0; b < batch; b++)
        if (n_samples[b] < 0 || n_samples[b] > pcm_stride) { wq4_set_error("wq4_transcribe: n_samples[%lld] = %lld outside 0..pcm_stride", (long long)b, (long long)n_samples[b]); return WQ4_ERR_INVALID; }
    // (an empty clip is padded to 30 s of silence like any short one, transcribe.rs:46-48: the "too short" bail of :61-63 cannot trigger)
    DeviceGuard g(m->device);
    cudaStream_t st = (cudaStream_t)stream;
    const int64_t n_mels = m->cfg.n_mels, frames = m->T_mel;
    float *mel_dev = nullptr;
    int64_t *lens_dev = nullptr;
    WQ4_CUDA(cudaMalloc(&mel_dev, (size_t)batch * n_mels * frames * 4));
    if (cudaMalloc(&lens_dev, (size_t)batch * 8) != cudaSuccess) { cudaFree(mel_dev); return wq4_cuda_fail(cudaGetLastError(), "wq4_transcribe alloc"); }
    int32_t rc = WQ4_OK;
    std::vector<int32_t> raw((size_t)batch * 224), cnt((size_t)batch);
    do {
        if (cudaMemcpyAsync(lens_dev, n_samples, (size_t)batch * 8, cudaMemcpyHostToDevice, st) != cudaSuccess) { rc = wq4_cuda_fail(cudaGetLastError(), "wq4_transcribe copy"); break; }
        if ((rc = wq4_logmel_forward_ragged(mel, pcm_dev, batch, pcm_stride, lens_dev, mel_dev, st)) != WQ4_OK) break;
        if (cudaStreamSynchronize(st) != cudaSuccess) { rc = wq4_cuda_fail(cudaGetLastError(), "wq4_transcribe sync"); break; }
        const auto t0 = std::chrono::steady_clock::now();
        if ((rc = wq4_whisper_transcribe_ex(m, mel_dev, batch, language, nullptr, raw.data(), cnt.data(), nullptr, st)) != WQ4_OK) break;
        if (elapsed_ms_out) *elapsed_ms_out = std::chrono::duration<double, std::milli>(std::chrono::steady_clock::now() - t0).count();
        for (int64_t b = 0; b < batch; b++) {
            int n = 0;
            for (int i = 0; i < cnt[b]; i++) {
                const int32_t t = raw[(size_t)b * 224 + i];
                if (t >= 0 && t < 50257) text_tokens_out[(size_t)b * 224 + n++] = t;     // transcribe.rs:96-100
            }
            text_counts_out[b] = n;
            if (raw_counts_out) raw_counts_out[b] = cnt[b];
            if (raw_tokens_out) memcpy(raw_tokens_out + (size_t)b * 224, raw.data() + (size_t)b * 224, (size_t)cnt[b] * 4);
        }
    } while (0);
    cudaFree(mel_dev);
    cudaFree(lens_dev);
    return rc;
}

// GGUF reader probes (host only, no device needed): GgufReader::{version, tensor_count, tensor_info,
// tensor_data} (reader.rs:190-223).
int32_t wq4_gguf_probe(const uint8_t *data, size_t nbytes, uint32_t *version, uint64_t *tensor_count, uint64_t *data_offset)
{
    if (!data) { wq4_set_error("wq4_gguf_probe: null data"); return WQ4_ERR_INVALID; }
    Gguf g;
    if (!g.parse(data, nbytes)) { wq4_set_error("Failed to parse GGUF: %s", g.err.c_str()); return WQ4_ERR_INVALID; }
    if (version) *version = g.version;
    if (tensor_count) *tensor_count = g.tensor_count;
    if (data_offset) *data_offset = g.data_offset;
    return WQ4_OK;
}

int32_t wq4_gguf_tensor_info(const uint8_t *data, size_t nbytes, const char *name, uint64_t *dims4, uint32_t *ndims,
                             uint32_t *dtype, uint64_t *abs_offset, uint64_t *byte_size)
{
    if (!data || !name) { wq4_set_error("wq4_gguf_tensor_info: null argument"); return WQ4_ERR_INVALID; }
    Gguf g;
    if (!g.parse(data, nbytes)) { wq4_set_error("Failed to parse GGUF: %s", g.err.c_str()); return WQ4_ERR_INVALID; }
    const GgufTensor *t = g.find(name);
    if (!t) { wq4_set_error("Tensor '%s' not found in GGUF", name); return WQ4_ERR_INVALID; }
    if (ndims) *ndims = (uint32_t)t->dims.size();
    if (dims4) for (size_t i = 0; i < 4; i++) dims4[i] = i < t->dims.size() ? t->dims[i] : 1;
    if (dtype) *dtype = t->dtype;
    if (abs_offset) *abs_offset = g.data_offset + t->offset;
    if (byte_size) *byte_size = t->byte_size();
    if (!g.tensor_data(*t)) { wq4_set_error("Tensor '%s' data out of range", name); return WQ4_ERR_BYTES; }
    return WQ4_OK;
}

// Test / bench hooks ------------------------------------------------------------------------------
// Runs the cross-K/V projection and ONE decoder pass with caller-given tokens; copies the logits out.
int32_t wq4_whisper_decoder_pass(wq4_whisper *m, const int32_t *tokens_host, int64_t batch, int64_t q_len, int64_t emb_pos0,
                                 int64_t kv_pos0, int32_t recompute_cross, float *logits_dev, wq4_stream stream)
{
    if (!m || !tokens_host || batch <= 0 || q_len <= 0) { wq4_set_error("wq4_whisper_decoder_pass: invalid argument"); return WQ4_ERR_INVALID; }
    DeviceGuard g(m->device);
    if (batch > m->max_batch) { wq4_set_error("wq4_whisper_decoder_pass: call wq4_whisper_encode first"); return WQ4_ERR_INVALID; }
    // validated BEFORE the token upload below: d_tokens holds max_batch * max_q entries
    if (q_len > m->max_q || emb_pos0 < 0 || kv_pos0 < 0 || emb_pos0 + q_len > m->cfg.n_text_ctx || kv_pos0 + q_len > m->cfg.n_text_ctx) {
        wq4_set_error("wq4_whisper_decoder_pass: q_len %lld (max %lld) / positions %lld, %lld outside the %d-token context", (long long)q_len,
                      (long long)m->max_q, (long long)emb_pos0, (long long)kv_pos0, m->cfg.n_text_ctx);
        return WQ4_ERR_INVALID;
    }
    cudaStream_t st = (cudaStream_t)stream;
    if (recompute_cross) CHK(cross_kv(m, batch, st));
    WQ4_CUDA(cudaMemcpyAsync(m->d_tokens, tokens_host, (size_t)batch * q_len * 4, cudaMemcpyHostToDevice, st));
    CHK(decoder_pass(m, batch, q_len, emb_pos0, kv_pos0, nullptr, st));
    if (logits_dev) WQ4_CUDA(cudaMemcpyAsync(logits_dev, m->logits, (size_t)batch * m->V * 4, cudaMemcpyDeviceToDevice, st));
    return WQ4_OK;
}

// Same for one single-query step over a SUBSET of the clips whose caches the model holds: row r of the pass is clip
// slots_host[r] (the active-set compaction of the greedy loop, exercised directly at the logits level).
int32_t wq4_whisper_decoder_pass_slots(wq4_whisper *m, const int32_t *tokens_host, const int32_t *slots_host, int64_t rows,
                                       int64_t emb_pos, int64_t kv_pos, float *logits_dev, wq4_stream stream)
{
    if (!m || !tokens_host || !slots_host || rows <= 0) { wq4_set_error("wq4_whisper_decoder_pass_slots: invalid argument"); return WQ4_ERR_INVALID; }
    DeviceGuard g(m->device);
    if (rows > m->kv_batch) { wq4_set_error("wq4_whisper_decoder_pass_slots: %lld rows but the caches hold %lld clips", (long long)rows, (long long)m->kv_batch); return WQ4_ERR_INVALID; }
    for (int64_t r = 0; r < rows; r++)
        if (slots_host[r] < 0 || slots_host[r] >= m->kv_batch) { wq4_set_error("wq4_whisper_decoder_pass_slots: slot %d outside 0..%lld", slots_host[r], (long long)m->kv_batch - 1); return WQ4_ERR_INVALID; }
    if (emb_pos < 0 || kv_pos < 0 || emb_pos >= m->cfg.n_text_ctx || kv_pos >= m->cfg.n_text_ctx) { wq4_set_error("wq4_whisper_decoder_pass_slots: position outside the context"); return WQ4_ERR_INVALID; }
    cudaStream_t st = (cudaStream_t)stream;
    WQ4_CUDA(cudaMemcpyAsync(m->d_tokens, tokens_host, (size_t)rows * 4, cudaMemcpyHostToDevice, st));
    WQ4_CUDA(cudaMemcpyAsync(m->d_slot, slots_host, (size_t)rows * 4, cudaMemcpyHostToDevice, st));
    CHK(decoder_pass(m, rows, 1, emb_pos, kv_pos, nullptr, st, m->d_slot));
    if (logits_dev) WQ4_CUDA(cudaMemcpyAsync(logits_dev, m->logits, (size_t)rows * m->V * 4, cudaMemcpyDeviceToDevice, st));
    return WQ4_OK;
}

// Stream-ordered single decode step on device-resident tokens (CUDA-graph capturable; used by bench.py):
// consumes d_next of the previous step, leaves the greedy pick in d_next again.
int32_t wq4_whisper_decode_step_device(wq4_whisper *m, int64_t batch, int64_t position, int64_t kv_len, wq4_stream stream)
{
    if (!m || batch <= 0 || batch > m->max_batch) { wq4_set_error("wq4_whisper_decode_step_device: invalid argument"); return WQ4_ERR_INVALID; }
    DeviceGuard g(m->device);
    cudaStream_t st = (cudaStream_t)stream;
    WQ4_CUDA(cudaMemcpyAsync(m->d_tokens, m->d_next, (size_t)batch * 4, cudaMemcpyDeviceToDevice, st));
    ArgReq req = {0, (int)m->V, m->d_sup_none, m->d_next};
    return decoder_pass(m, batch, 1, position, kv_len, &req, st);
}

}  // extern "C"
