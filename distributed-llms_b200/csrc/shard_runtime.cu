// The shard runtime behind the C ABI (include/dllm_b200.h): one handle = one pipeline stage
// (reference: one `ModelShard`, src/worker/node.py:13-32).  It owns the stage's paged KV cache, block
// table, activation and scratch buffers; it binds (does not own) the caller's weights; it turns a
// forward call into a CUDA-graph replay of the fused decode kernels.
#include <stdio.h>
#include <string.h>

#include <stdlib.h>

#include <algorithm>
#include <map>
#include <string>
#include <vector>

#include "../../include/dllm_b200.h"
#include "kernels.h"
#include "megakernel.h"

using namespace dllm;

// ---- error plumbing ------------------------------------------------------------------------------

static thread_local std::string g_err;

static int fail(const std::string& msg) {
    g_err = msg;
    return 1;
}

int dllm::set_last_error(const char* msg) { return fail(msg ? msg : ""); }

#define CUDA_TRY(expr)                                                                             \
    do {                                                                                           \
        cudaError_t _e = (expr);                                                                   \
        if (_e != cudaSuccess)                                                                     \
            return fail(std::string(#expr) + ": " + cudaGetErrorName(_e) + " (" + cudaGetErrorString(_e) + ")"); \
    } while (0)

extern "C" const char* dllm_last_error(void) { return g_err.c_str(); }
extern "C" int dllm_abi_version(void) { return DLLM_ABI_VERSION; }
extern "C" int dllm_device_count(void) {
    int n = 0;
    if (cudaGetDeviceCount(&n) != cudaSuccess) {
        cudaGetLastError();
        return 0;
    }
    return n;
}
extern "C" unsigned long long dllm_kernel_launch_count(void) { return launch_count(); }

static int ensure_device_ready() {
    // per-device one-time kernel attribute set-up
    static bool ready[64] = {false};
    int dev = 0;
    CUDA_TRY(cudaGetDevice(&dev));
    if (dev < 64 && ready[dev]) return 0;
    CUDA_TRY(gemv_init_device());
    if (dev < 64) ready[dev] = true;
    return 0;
}

// ---- the handle -----------------------------------------------------------------------------------

struct GraphEntry {
    cudaGraphExec_t exec = nullptr;
    unsigned long long kernels = 0;
};

struct dllm_shard {
    dllm_shard_config cfg;
    std::vector<dllm_layer_weights> layers;
    dllm_io_weights io;
    int device = 0, num_sms = 148;
    size_t esz = 2;
    // device memory
    char* arena = nullptr;
    size_t arena_bytes = 0;
    void* h = nullptr;         // residual stream [max_rows, d]
    void* qbuf = nullptr;      // [max_rows, H*hd]
    void* attn_out = nullptr;  // [max_rows, H*hd]
    void* act = nullptr;       // [max_rows, ffn]
    void* xn = nullptr;        // [max_rows, d]   normalised rows (prefill GEMM input)
    void* qkv_tmp = nullptr;   // [max_rows, (H+2KV)*hd] projected rows before rotary (prefill)
    float* gemm_ws = nullptr;          // fp32 partial tiles of the prefill GEMMs' split tails (one slice per SM at most)
    unsigned* gemm_tickets = nullptr;  // one per split tile, zero between launches
    size_t gemm_ws_bytes = 0;
    int32_t *pipe_pos = nullptr, *pipe_seq = nullptr;   // positions / KV slots of every micro-batch of a pipelined per-op loop
    int* pipe_status = nullptr;   // non-zero: a pipelined loop of this stage gave up waiting for a ticket
    int32_t *ids = nullptr, *pos = nullptr, *seq = nullptr, *logit_rows = nullptr, *arange = nullptr;
    int32_t *next_tok = nullptr, *block_table = nullptr, *tok_ring = nullptr, *step_ctr = nullptr;
    float* logits = nullptr;   // [max_logit_rows, V]
    float *part_o = nullptr, *part_ml = nullptr, *arg_val = nullptr;
    int32_t* arg_idx = nullptr;
    uint32_t *attn_counters = nullptr, *arg_counter = nullptr;
    // persistent decode kernel (bf16, <= 4 rows)
    MkPhase* mk_phases = nullptr;
    MkPhase* mk_phases_xg = nullptr;  // same step with the down projection in xg mode (rows streamed from global)
    int mk_part_units_xg = 1, mk_k_max_xg = 0;
    bool mk_xg_ok = false;
    MkLayer* mk_layers = nullptr;
    unsigned* mk_sync = nullptr;  // [0] barrier, [1] done, [2] argmax ticket
    void* mk_packed = nullptr;    // streaming-layout copy of the weights (megakernel.h)
    int mk_n_phases = 0, mk_k_max = 0, mk_part_units = 1, mk_out_phase = -1;
    bool mk_enabled = false, mk_packed_ok = false;
    std::vector<void*> kpool, vpool;
    // geometry
    int pages_per_seq = 0, page_shift = 0, num_pages = 0, max_logit_rows = 0, arg_grid_cap = 0;
    long max_partials = 0;
    size_t tok_ring_cap = 0;
    // host-side KV bookkeeping
    std::vector<int32_t> free_pages;
    std::vector<std::vector<int32_t>> seq_pages;
    int32_t* h_bt_row = nullptr;  // pinned staging, one block-table row per slot
    // graphs
    std::map<unsigned long long, GraphEntry> graphs;
    cudaStream_t cap_stream = nullptr;
    // true after a decode loop: `pos` already holds the position of the pending token (the loops advance after
    // every step, the last one included); false after a forward: `pos` is the position of the rows just processed
    bool pos_at_pending = false;
    int arange_rows = 0;  // logit_rows[0 .. arange_rows) currently holds 0, 1, 2, ... (skips the identity copy of single steps)
};

static size_t align_up(size_t v, size_t a) { return (v + a - 1) / a * a; }

// which blocks of local layer l this stage holds (a boundary may fall between a layer's attention and MLP blocks)
static bool has_attn(const dllm_shard_config& c, int l) { return !(l == 0 && c.first_layer_mlp_only); }
static bool has_mlp(const dllm_shard_config& c, int l) { return !(l == c.n_layers - 1 && c.last_layer_attn_only); }

static int pick_splits(const dllm_shard* s, int n_rows) {
    const int H = s->cfg.n_heads;
    const long target = 8L * s->num_sms;
    long S = (target + (long)n_rows * H - 1) / ((long)n_rows * H);
    const long cap_ctx = (s->cfg.max_context + 15) / 16;
    if (S > cap_ctx) S = cap_ctx;
    if (S > 64) S = 64;
    while (S > 1 && (long)n_rows * H * S > s->max_partials) --S;
    if (S < 1) S = 1;
    return (int)S;
}

extern "C" int dllm_shard_create(const dllm_shard_config* cfg, const dllm_layer_weights* layers,
                                 const dllm_io_weights* io, dllm_shard** out) {
    if (!cfg || !out || (cfg->n_layers > 0 && !layers) || !io) return fail("dllm_shard_create: null argument");
    *out = nullptr;
    const dllm_shard_config& c = *cfg;
    if (c.dtype != DLLM_DTYPE_BF16 && c.dtype != DLLM_DTYPE_F32) return fail("dllm_shard_create: bad dtype");
    if (c.arch < DLLM_ARCH_GPT2 || c.arch > DLLM_ARCH_QWEN2) return fail("dllm_shard_create: bad arch");
    if (c.head_dim != 32 && c.head_dim != 64 && c.head_dim != 128)
        return fail("dllm_shard_create: head_dim must be 32, 64 or 128");
    if (c.n_heads <= 0 || c.n_kv_heads <= 0 || c.n_heads % c.n_kv_heads != 0)
        return fail("dllm_shard_create: n_heads must be a positive multiple of n_kv_heads");
    const int vec = c.dtype == DLLM_DTYPE_BF16 ? 8 : 4;
    if (c.hidden % vec || c.ffn % vec || (c.n_heads * c.head_dim) % vec)
        return fail("dllm_shard_create: hidden, ffn and H*hd must be multiples of the 16-byte vector width");
    if (c.page_size < 8 || (c.page_size & (c.page_size - 1))) return fail("dllm_shard_create: page_size must be a power of two >= 8");
    if (c.max_rows <= 0 || c.max_seqs <= 0 || c.max_context <= 0) return fail("dllm_shard_create: max_rows/max_seqs/max_context must be positive");
    if ((c.first_layer_mlp_only || c.last_layer_attn_only) && c.n_layers < 1) return fail("dllm_shard_create: a split layer needs n_layers >= 1");
    if (c.first_layer_mlp_only && c.last_layer_attn_only && c.n_layers < 2) return fail("dllm_shard_create: one layer cannot be both MLP-only and attention-only");
    if (c.first_layer_mlp_only && c.has_embed) return fail("dllm_shard_create: the stage with the embedding starts with a whole layer");
    if (c.last_layer_attn_only && c.has_head) return fail("dllm_shard_create: the stage with the head ends with a whole layer");
    for (int l = 0; l < c.n_layers; ++l) {
        const dllm_layer_weights& w = layers[l];
        if (has_attn(c, l) && (!w.norm1_w || !w.wq || !w.wk || !w.wv || !w.wo)) return fail("dllm_shard_create: attention weights missing");
        if (has_mlp(c, l) && (!w.norm2_w || !w.w_up || !w.w_down)) return fail("dllm_shard_create: MLP weights missing");
    }
    if (c.has_embed && !io->wte) return fail("dllm_shard_create: has_embed but wte is null");
    if (c.has_head && (!io->lm_head || !io->final_norm_w)) return fail("dllm_shard_create: has_head but lm_head/final_norm_w is null");
    if (c.arch != DLLM_ARCH_GPT2 && c.n_layers > 0 && (!io->rope_cos || !io->rope_sin))
        return fail("dllm_shard_create: rope tables missing");
    int ndev = 0;
    if (cudaGetDeviceCount(&ndev) != cudaSuccess || ndev == 0) {
        cudaGetLastError();
        return fail("dllm_shard_create: no CUDA device (this library has no CPU path)");
    }
    if (ensure_device_ready()) return 1;

    dllm_shard* s = new dllm_shard();
    s->cfg = c;
    s->layers.assign(layers, layers + c.n_layers);
    s->io = *io;
    s->esz = c.dtype == DLLM_DTYPE_BF16 ? 2 : 4;
    cudaGetDevice(&s->device);
    cudaDeviceGetAttribute(&s->num_sms, cudaDevAttrMultiProcessorCount, s->device);
    s->pages_per_seq = (c.max_context + c.page_size - 1) / c.page_size;
    while ((1 << s->page_shift) < c.page_size) ++s->page_shift;
    s->num_pages = c.num_pages > 0 ? c.num_pages : c.max_seqs * s->pages_per_seq;
    s->max_logit_rows = c.has_head ? c.max_rows : 0;
    s->max_partials = (long)c.max_rows * c.n_heads + 16L * s->num_sms;
    s->arg_grid_cap = s->num_sms * 4 + 8;
    s->tok_ring_cap = 1u << 20;

    // carve one arena
    const size_t Hd = (size_t)c.n_heads * c.head_dim;
    struct Item { void** p; size_t bytes; };
    std::vector<Item> items;
    auto add = [&](void* pp, size_t bytes) { items.push_back({(void**)pp, align_up(bytes ? bytes : 16, 256)}); };
    add(&s->h, (size_t)c.max_rows * c.hidden * s->esz);
    add(&s->qbuf, (size_t)c.max_rows * Hd * s->esz);
    add(&s->attn_out, (size_t)c.max_rows * Hd * s->esz);
    add(&s->act, (size_t)std::max(c.max_rows, 8) * c.ffn * s->esz);
    const bool prefill_bufs = c.dtype == DLLM_DTYPE_BF16 && c.max_rows > 4;
    add(&s->xn, prefill_bufs ? (size_t)c.max_rows * c.hidden * s->esz : 0);
    add(&s->qkv_tmp, prefill_bufs ? (size_t)c.max_rows * (c.n_heads + 2 * c.n_kv_heads) * c.head_dim * s->esz : 0);
    s->gemm_ws_bytes = prefill_bufs && c.max_rows > 64 ? (size_t)s->num_sms * 128 * 256 * 4 : 0;
    add(&s->gemm_ws, s->gemm_ws_bytes);
    add(&s->gemm_tickets, (size_t)kGemmSplitTickets * 4);
    add(&s->ids, (size_t)c.max_rows * 4);
    add(&s->pipe_pos, (size_t)c.max_rows * 4);
    add(&s->pipe_seq, (size_t)c.max_rows * 4);
    add(&s->pipe_status, 4);
    add(&s->pos, (size_t)c.max_rows * 4);
    add(&s->seq, (size_t)c.max_rows * 4);
    add(&s->logit_rows, (size_t)c.max_rows * 4);
    add(&s->arange, (size_t)c.max_rows * 4);
    add(&s->next_tok, (size_t)c.max_rows * 4);
    add(&s->block_table, (size_t)c.max_seqs * s->pages_per_seq * 4);
    add(&s->tok_ring, c.has_embed ? s->tok_ring_cap * 4 : 0);
    add(&s->step_ctr, 4);
    add(&s->logits, (size_t)s->max_logit_rows * c.vocab * 4);
    add(&s->part_o, (size_t)s->max_partials * c.head_dim * 4);
    add(&s->part_ml, (size_t)s->max_partials * 2 * 4);
    add(&s->arg_val, (size_t)s->arg_grid_cap * c.max_rows * 4);
    add(&s->arg_idx, (size_t)s->arg_grid_cap * c.max_rows * 4);
    add(&s->attn_counters, (size_t)c.max_rows * c.n_heads * 4);
    add(&s->arg_counter, 4);
    const int mk_phase_cap = 5 * c.n_layers + 2;
    add(&s->mk_phases, (size_t)mk_phase_cap * sizeof(MkPhase));
    add(&s->mk_phases_xg, (size_t)mk_phase_cap * sizeof(MkPhase));
    add(&s->mk_layers, (size_t)(c.n_layers > 0 ? c.n_layers : 1) * sizeof(MkLayer));
    add(&s->mk_sync, 16);
    {   // the persistent decode kernel streams a packed copy of the weights (bf16, every K a multiple of 32)
        const char* env = getenv("DLLM_MEGAKERNEL");
        const int Hdd = c.n_heads * c.head_dim;
        s->mk_packed_ok = c.dtype == DLLM_DTYPE_BF16 && c.n_layers > 0 && !(env && env[0] == '0') &&
                          c.hidden % 32 == 0 && Hdd % 32 == 0 && c.ffn % 32 == 0 && c.n_heads / c.n_kv_heads <= 8;
        size_t pb = 0;
        if (s->mk_packed_ok) {
            const bool g2 = c.arch == DLLM_ARCH_GPT2;
            const size_t per_layer = align_up(mk_packed_bytes((c.n_heads + 2 * c.n_kv_heads) * (c.head_dim / 8), c.hidden), 256) +
                                     align_up(mk_packed_bytes(c.hidden / 8, Hdd), 256) +
                                     align_up(mk_packed_bytes(g2 ? c.ffn / 8 : c.ffn / 4, c.hidden), 256) +
                                     align_up(mk_packed_bytes(c.hidden / 8, c.ffn), 256);
            pb = per_layer * c.n_layers + (c.has_head ? align_up(mk_packed_bytes((c.vocab + 7) / 8, c.hidden), 256) : 0);
        }
        add(&s->mk_packed, pb);
    }
    const size_t kv_layer = (size_t)s->num_pages * c.n_kv_heads * c.page_size * c.head_dim * s->esz;
    s->kpool.resize(c.n_layers);
    s->vpool.resize(c.n_layers);
    for (int l = 0; l < c.n_layers; ++l) {
        add(&s->kpool[l], has_attn(c, l) ? kv_layer : 0);
        add(&s->vpool[l], has_attn(c, l) ? kv_layer : 0);
    }
    size_t total = 0;
    for (auto& it : items) total += it.bytes;
    cudaError_t e = cudaMalloc((void**)&s->arena, total);
    if (e != cudaSuccess) {
        delete s;
        char buf[160];
        snprintf(buf, sizeof buf, "dllm_shard_create: cudaMalloc of %zu bytes failed: %s", total, cudaGetErrorString(e));
        cudaGetLastError();
        return fail(buf);
    }
    s->arena_bytes = total;
    size_t off = 0;
    for (auto& it : items) {
        *it.p = s->arena + off;
        off += it.bytes;
    }
    // zero everything: counters must start at zero, and the tcgen05 attention multiplies whole 16-token blocks of V --
    // rows past the end of a context meet a zero probability, which only helps if they never held NaN bit patterns
    e = cudaMemset(s->arena, 0, total);
    if (e == cudaSuccess) e = cudaMemset(s->block_table, 0xff, (size_t)c.max_seqs * s->pages_per_seq * 4);  // -1
    std::vector<int32_t> ar(c.max_rows);
    for (int i = 0; i < c.max_rows; ++i) ar[i] = i;
    if (e == cudaSuccess) e = cudaMemcpy(s->arange, ar.data(), ar.size() * 4, cudaMemcpyHostToDevice);
    if (e == cudaSuccess) e = cudaMallocHost((void**)&s->h_bt_row, (size_t)c.max_seqs * s->pages_per_seq * 4);
    if (e == cudaSuccess) e = cudaStreamCreateWithFlags(&s->cap_stream, cudaStreamNonBlocking);
    if (e != cudaSuccess) {
        std::string m = std::string("dllm_shard_create: ") + cudaGetErrorString(e);
        cudaGetLastError();
        dllm_shard_destroy(s);
        return fail(m);
    }
    memset(s->h_bt_row, 0xff, (size_t)c.max_seqs * s->pages_per_seq * 4);
    if (s->mk_packed_ok) {
        // packed streaming copy of the weights + phase list of one decode step for the persistent kernel
        const bool gpt2 = c.arch == DLLM_ARCH_GPT2;
        const int Hd = c.n_heads * c.head_dim;
        const int G = s->num_sms;
        std::vector<MkPhase> ph;
        std::vector<MkLayer> ly(c.n_layers > 0 ? c.n_layers : 1);
        char* pk = (char*)s->mk_packed;
        cudaError_t me = cudaSuccess;
        int part_units = 1;
        auto weight_phase = [&](int kind, int layer, int n_panels, int K, int n0, const void* w0, const void* w1, const void* w2) {
            PackJob job{};
            job.src[0] = w0; job.src[1] = w1; job.src[2] = w2;
            job.dst = pk; job.kind = kind; job.n_panels = n_panels; job.K = K; job.n0 = n0;
            job.n_heads = c.n_heads; job.n_kv_heads = c.n_kv_heads; job.head_dim = c.head_dim;
            if (me == cudaSuccess) me = launch_pack(job, 0);
            MkPhase q{};
            q.wp = pk; q.n_panels = n_panels; q.K = K; q.n0 = n0; q.kind = (int16_t)kind; q.layer = (int16_t)layer;
            mk_set_stages(q, (K + MK_STAGE_K - 1) / MK_STAGE_K);
            ph.push_back(q);
            pk += align_up(mk_packed_bytes(n_panels, K), 256);
        };
        auto plain_phase = [&](int kind, int layer) {
            MkPhase q{};
            q.kind = (int16_t)kind; q.layer = (int16_t)layer; mk_set_stages(q, 1);
            ph.push_back(q);
        };
        if (c.has_embed) plain_phase(MK_EMBED, 0);
        for (int l = 0; l < c.n_layers; ++l) {
            const dllm_layer_weights& w = s->layers[l];
            ly[l] = MkLayer{w.norm1_w, w.norm1_b, w.bq, w.bk, w.bv, w.bo, w.norm2_w, w.norm2_b, w.b_up, w.b_down,
                            s->kpool[l], s->vpool[l]};
            if (has_attn(c, l)) {
                weight_phase(MK_QKV, l, (c.n_heads + 2 * c.n_kv_heads) * (c.head_dim / 8), c.hidden, Hd, w.wq, w.wk, w.wv);
                plain_phase(MK_ATTN, l);
                weight_phase(MK_OPROJ, l, c.hidden / 8, Hd, c.hidden, w.wo, nullptr, nullptr);
            }
            if (has_mlp(c, l)) {
                if (gpt2) weight_phase(MK_FC_GELU, l, c.ffn / 8, c.hidden, c.ffn, w.w_up, nullptr, nullptr);
                else weight_phase(MK_GATEUP, l, c.ffn / 4, c.hidden, c.ffn, w.w_gate, w.w_up, nullptr);
                weight_phase(MK_DOWN, l, c.hidden / 8, c.ffn, c.hidden, w.w_down, nullptr, nullptr);
            }
        }
        if (c.has_head) weight_phase(MK_HEAD, 0, (c.vocab + 7) / 8, c.hidden, c.vocab, s->io.lm_head, nullptr, nullptr);
        s->mk_n_phases = (int)ph.size();
        for (int i = 0; i < (int)ph.size(); ++i)
            if (ph[i].kind == MK_OPROJ || ph[i].kind == MK_DOWN) s->mk_out_phase = i;  // the stage's rows are final here
        s->mk_part_units = MK_PART_UNITS;
        s->mk_k_max = std::max(std::max(c.hidden, Hd), c.ffn);
        if (me == cudaSuccess) me = cudaMemcpy(s->mk_phases, ph.data(), ph.size() * sizeof(MkPhase), cudaMemcpyHostToDevice);
        if (me == cudaSuccess) me = cudaMemcpy(s->mk_layers, ly.data(), ly.size() * sizeof(MkLayer), cudaMemcpyHostToDevice);
        // variant: down projection with its input rows streamed from global memory (half-size stages)
        s->mk_xg_ok = c.ffn % 128 == 0;
        if (s->mk_xg_ok) {
            std::vector<MkPhase> px = ph;
            int pu = 1;
            for (auto& q : px) {
                if (q.kind == MK_DOWN) {
                    q.xg = 1;
                    mk_set_stages(q, q.K / (MK_STAGE_K / 2));
                }
            }
            s->mk_part_units_xg = MK_PART_UNITS;
            s->mk_k_max_xg = std::max(c.hidden, Hd);
            if (me == cudaSuccess) me = cudaMemcpy(s->mk_phases_xg, px.data(), px.size() * sizeof(MkPhase), cudaMemcpyHostToDevice);
        }
        if (me == cudaSuccess) me = cudaDeviceSynchronize();
        s->mk_enabled = me == cudaSuccess;
        cudaGetLastError();
    }
    s->free_pages.resize(s->num_pages);
    for (int i = 0; i < s->num_pages; ++i) s->free_pages[i] = s->num_pages - 1 - i;  // pop_back hands out 0,1,2,...
    s->seq_pages.resize(c.max_seqs);
    *out = s;
    return 0;
}

extern "C" int dllm_shard_destroy(dllm_shard* s) {
    if (!s) return 0;
    cudaSetDevice(s->device);
    cudaDeviceSynchronize();
    for (auto& kv : s->graphs)
        if (kv.second.exec) cudaGraphExecDestroy(kv.second.exec);
    if (s->cap_stream) cudaStreamDestroy(s->cap_stream);
    if (s->h_bt_row) cudaFreeHost(s->h_bt_row);
    if (s->arena) cudaFree(s->arena);
    cudaGetLastError();
    delete s;
    return 0;
}

extern "C" size_t dllm_shard_device_bytes(const dllm_shard* s) { return s ? s->arena_bytes : 0; }

// ---- KV pages ---------------------------------------------------------------------------------------

extern "C" int dllm_shard_kv_reserve(dllm_shard* s, int32_t seq_id, int32_t n_tokens, void* stream) {
    if (!s) return fail("dllm_shard_kv_reserve: null shard");
    if (seq_id < 0 || seq_id >= s->cfg.max_seqs) return fail("dllm_shard_kv_reserve: seq_id out of range");
    if (n_tokens < 0 || n_tokens > s->cfg.max_context) return fail("dllm_shard_kv_reserve: n_tokens exceeds max_context");
    auto& pages = s->seq_pages[seq_id];
    const int need = (n_tokens + s->cfg.page_size - 1) / s->cfg.page_size;
    if ((int)pages.size() >= need) return 0;
    if ((int)(need - pages.size()) > (int)s->free_pages.size()) return fail("dllm_shard_kv_reserve: KV page pool exhausted");
    int32_t* row = s->h_bt_row + (size_t)seq_id * s->pages_per_seq;
    const size_t first = pages.size();
    while ((int)pages.size() < need) {
        const int32_t pg = s->free_pages.back();
        s->free_pages.pop_back();
        row[pages.size()] = pg;
        pages.push_back(pg);
    }
    CUDA_TRY(cudaMemcpyAsync(s->block_table + (size_t)seq_id * s->pages_per_seq + first, row + first,
                             (pages.size() - first) * 4, cudaMemcpyHostToDevice, (cudaStream_t)stream));
    return 0;
}

extern "C" int dllm_shard_kv_release(dllm_shard* s, int32_t seq_id) {
    if (!s) return fail("dllm_shard_kv_release: null shard");
    if (seq_id >= s->cfg.max_seqs) return fail("dllm_shard_kv_release: seq_id out of range");
    const int lo = seq_id < 0 ? 0 : seq_id, hi = seq_id < 0 ? s->cfg.max_seqs : seq_id + 1;
    for (int q = lo; q < hi; ++q) {
        auto& pages = s->seq_pages[q];
        while (!pages.empty()) {
            s->free_pages.push_back(pages.back());
            pages.pop_back();
        }
    }
    // block-table rows keep their stale ids: a slot is always re-reserved before it is used again
    return 0;
}

extern "C" int dllm_shard_kv_free_pages(const dllm_shard* s) { return s ? (int)s->free_pages.size() : 0; }

// ---- kernel sequence of one stage step -------------------------------------------------------------

// Enqueue this stage's kernels for n_rows rows already staged in the handle's buffers.
static int enqueue_step(dllm_shard* s, int n_rows, int n_logit_rows, bool advance, cudaStream_t st) {
    const dllm_shard_config& c = s->cfg;
    const int dt = c.dtype;
    const bool gpt2 = c.arch == DLLM_ARCH_GPT2;
    const int norm = gpt2 ? NORM_LAYER : NORM_RMS;
    const int Hd = c.n_heads * c.head_dim;
    if (c.has_embed) {
        EmbedParams e{};
        e.ids = s->ids; e.positions = s->pos; e.wte = s->io.wte; e.wpe = gpt2 ? s->io.wpe : nullptr;
        e.out = s->h; e.n_rows = n_rows; e.hidden = c.hidden; e.vocab = c.vocab;
        CUDA_TRY(launch_embed(dt, e, st));
    }
    const int splits = pick_splits(s, n_rows);
    for (int l = 0; l < c.n_layers; ++l) {
        const dllm_layer_weights& w = s->layers[l];
        PagedKV kv{s->kpool[l], s->vpool[l], s->block_table, s->pages_per_seq, s->page_shift};
        if (has_attn(c, l)) {   // K1: norm -> QKV -> bias -> rope -> paged append
            GemvParams p{};
            p.x = s->h; p.n_rows = n_rows; p.K = c.hidden;
            p.norm = norm; p.norm_w = w.norm1_w; p.norm_b = w.norm1_b; p.eps = c.norm_eps;
            p.w[0] = w.wq; p.w[1] = w.wk; p.w[2] = w.wv; p.b[0] = w.bq; p.b[1] = w.bk; p.b[2] = w.bv;
            p.n[0] = Hd; p.n[1] = p.n[2] = c.n_kv_heads * c.head_dim;
            p.epi = EPI_QKV; p.y = s->qbuf;
            p.n_heads = c.n_heads; p.n_kv_heads = c.n_kv_heads; p.head_dim = c.head_dim; p.rope = !gpt2;
            p.rope_cos = s->io.rope_cos; p.rope_sin = s->io.rope_sin;
            p.positions = s->pos; p.seq_ids = s->seq; p.kv = kv;
            CUDA_TRY(launch_gemv(dt, p, s->num_sms, st));
        }
        if (has_attn(c, l)) {   // K2: split-K flash decoding
            AttnParams a{};
            a.q = s->qbuf; a.kv = kv; a.positions = s->pos; a.seq_ids = s->seq;
            a.n_rows = n_rows; a.n_heads = c.n_heads; a.n_kv_heads = c.n_kv_heads; a.head_dim = c.head_dim;
            a.n_splits = splits; a.part_o = s->part_o; a.part_ml = s->part_ml; a.counters = s->attn_counters;
            a.out = s->attn_out;
            CUDA_TRY(launch_attn_decode(dt, a, st));
        }
        if (has_attn(c, l)) {   // K3: o-proj + residual
            GemvParams p{};
            p.x = s->attn_out; p.n_rows = n_rows; p.K = Hd; p.norm = NORM_NONE;
            p.w[0] = w.wo; p.b[0] = w.bo; p.n[0] = c.hidden;
            p.epi = EPI_RESIDUAL; p.y = s->h; p.accumulate = 1;
            CUDA_TRY(launch_gemv(dt, p, s->num_sms, st));
        }
        if (has_mlp(c, l)) {   // K4: norm -> gate/up -> SiLU*up   |   norm -> fc -> gelu_new
            GemvParams p{};
            p.x = s->h; p.n_rows = n_rows; p.K = c.hidden;
            p.norm = norm; p.norm_w = w.norm2_w; p.norm_b = w.norm2_b; p.eps = c.norm_eps;
            if (gpt2) { p.w[0] = w.w_up; p.b[0] = w.b_up; p.n[0] = c.ffn; p.epi = EPI_FC_GELU; }
            else { p.w[0] = w.w_gate; p.w[1] = w.w_up; p.n[0] = p.n[1] = c.ffn; p.epi = EPI_GATEUP; }
            p.y = s->act;
            CUDA_TRY(launch_gemv(dt, p, s->num_sms, st));
        }
        if (has_mlp(c, l)) {   // K5: down-proj + residual
            GemvParams p{};
            p.x = s->act; p.n_rows = n_rows; p.K = c.ffn; p.norm = NORM_NONE;
            p.w[0] = w.w_down; p.b[0] = w.b_down; p.n[0] = c.hidden;
            p.epi = EPI_RESIDUAL; p.y = s->h; p.accumulate = 1;
            CUDA_TRY(launch_gemv(dt, p, s->num_sms, st));
        }
    }
    if (c.has_head && n_logit_rows > 0) {   // K6: final norm -> lm_head -> argmax
        GemvParams p{};
        p.x = s->h; p.row_map = s->logit_rows; p.n_rows = n_logit_rows; p.K = c.hidden;
        p.norm = norm; p.norm_w = s->io.final_norm_w; p.norm_b = s->io.final_norm_b; p.eps = c.norm_eps;
        p.w[0] = s->io.lm_head; p.n[0] = c.vocab; p.epi = EPI_LOGITS;
        p.logits = s->logits; p.arg_val = s->arg_val; p.arg_idx = s->arg_idx; p.counter = s->arg_counter;
        p.next_token = s->next_tok;
        if (gemv_grid_size(dt, p, s->num_sms) > s->arg_grid_cap) return fail("internal: argmax scratch too small");
        CUDA_TRY(launch_gemv(dt, p, s->num_sms, st));
    }
    if (advance) CUDA_TRY(launch_advance(s->ids, s->pos, s->next_tok, s->tok_ring, s->step_ctr, n_rows, st));
    return 0;
}

// One launch of the persistent decode kernel for `n_steps` steps (bf16, <= 16 rows, logits for every row).
// Returns 0 = not eligible, 1 = plain variant, 2 = variant with the down projection's rows streamed from global.
static int mk_variant(const dllm_shard* s, int n_rows, int n_logit_rows, bool identity_rows) {
    if (!s->mk_enabled || !mk_supported(n_rows, s->cfg.n_heads, s->cfg.n_kv_heads, s->cfg.head_dim)) return 0;
    if (mk_max_splits(n_rows, s->cfg.n_heads, s->cfg.head_dim) < 1) return 0;
    if (s->cfg.has_head && !(identity_rows && n_logit_rows == n_rows)) return 0;
    MkParams q{};
    q.n_phases = s->mk_n_phases; q.n_rows = n_rows; q.head_dim = s->cfg.head_dim; q.pages_per_seq = s->pages_per_seq;
    q.k_max = s->mk_k_max; q.part_units = s->mk_part_units;
    const char* force_xg = getenv("DLLM_MK_XG");  // tests: take the streamed-rows variant even when the plain one fits
    if (!(force_xg && atoi(force_xg) == 1 && s->mk_xg_ok && n_rows <= 8) &&
        mk_smem_bytes(q, 2) <= (size_t)MK_SMEM_BUDGET) return 1;
    if (s->mk_xg_ok && n_rows <= 8) {
        q.k_max = s->mk_k_max_xg; q.part_units = s->mk_part_units_xg;
        if (mk_smem_bytes(q, 2) <= (size_t)MK_SMEM_BUDGET) return 2;
    }
    return 0;
}

static bool mk_eligible(const dllm_shard* s, int n_rows, int n_logit_rows, bool identity_rows) {
    return mk_variant(s, n_rows, n_logit_rows, identity_rows) != 0;
}

extern "C" int dllm_handoff_peer_copy(void* dst_ptr, int dst_device, const void* src_ptr, int src_device, size_t bytes, void* stream);

// mailboxes of the fused hand-off (see include/dllm_b200.h)
struct dllm_mailbox {
    int device = 0;           // device the pointers are valid on
    int n_slots = 0, rows = 0, hidden = 0;
    char* base = nullptr;
    bool owner = false, ipc = false;
    unsigned* flags() const { return reinterpret_cast<unsigned*>(base); }
    int32_t* tokens() const { return reinterpret_cast<int32_t*>(base + align_up((size_t)n_slots * 128, 256)); }
    void* hidden_rows() const { return base + align_up((size_t)n_slots * 128, 256) + align_up((size_t)n_slots * rows * 4, 256); }
    size_t bytes() const {
        return align_up((size_t)n_slots * 128, 256) + align_up((size_t)n_slots * rows * 4, 256) + (size_t)n_slots * rows * hidden * 2;
    }
};
constexpr int kFlagStride = 32;  // one ticket per 128-byte line
constexpr int kPipeOpsMaxRows = 64;  // rows per micro-batch of the per-op pipelined loop (one CUDA graph per step)

// how long a stage waits for a hand-off ticket before it abandons the loop (DLLM_TICKET_TIMEOUT_MS, default 20 s:
// the stages' processes start seconds apart)
static unsigned long long ticket_timeout_ns() {
    const char* e = getenv("DLLM_TICKET_TIMEOUT_MS");
    const long ms = e ? atol(e) : 0;
    return (unsigned long long)(ms > 0 ? ms : 20000) * 1000000ull;
}

struct MkPipe {
    const dllm_mailbox *in, *out;
    int n_micro;
    unsigned step_base;
};

// a single step whose int32 inputs are read in place (pinned host or device memory) and whose sampled ids go straight out
struct MkHostIo {
    const int32_t *ids, *pos, *seq;   // device-visible addresses
    int32_t* next_token;              // device-visible address or null
};

// the address the device can use for `ptr` when it is pinned / registered host memory or device memory; null for pageable memory
static const void* device_visible(const void* ptr) {
    cudaPointerAttributes a{};
    if (cudaPointerGetAttributes(&a, ptr) != cudaSuccess) {
        cudaGetLastError();
        return nullptr;
    }
    if (a.type == cudaMemoryTypeHost || a.type == cudaMemoryTypeDevice || a.type == cudaMemoryTypeManaged) return a.devicePointer;
    return nullptr;
}

static int run_megakernel(dllm_shard* s, int n_rows, int n_steps, bool advance, cudaStream_t st, const MkPipe* pipe = nullptr,
                          const MkHostIo* io = nullptr) {
    const dllm_shard_config& c = s->cfg;
    const int variant = mk_variant(s, n_rows, n_rows, true);
    if (variant == 0) return fail("internal: persistent kernel not eligible");
    MkParams p{};
    p.phases = variant == 2 ? s->mk_phases_xg : s->mk_phases; p.layers = s->mk_layers; p.n_phases = s->mk_n_phases;
    p.act_xlayout = variant == 2 ? 1 : 0;
    p.n_steps = n_steps; p.n_rows = n_rows;
    p.hidden = c.hidden; p.n_heads = c.n_heads; p.n_kv_heads = c.n_kv_heads; p.head_dim = c.head_dim;
    p.vocab = c.vocab; p.k_max = variant == 2 ? s->mk_k_max_xg : s->mk_k_max;
    p.norm = c.arch == DLLM_ARCH_GPT2 ? NORM_LAYER : NORM_RMS; p.rope = c.arch != DLLM_ARCH_GPT2;
    long S = s->num_sms / ((long)n_rows * c.n_kv_heads);  // rows x KV heads x S units: at most one per CTA
    if (S < 1) S = 1;
    const long s_cap = mk_max_splits(n_rows, c.n_heads, c.head_dim);
    if (S > s_cap) S = s_cap;
    while (S > 1 && (long)n_rows * c.n_heads * S > s->max_partials) --S;
    if (const char* e = getenv("DLLM_MK_SPLITS")) { long v = atol(e); if (v >= 1 && v <= s_cap && (long)n_rows * c.n_heads * v <= s->max_partials) S = v; }
    p.part_units = variant == 2 ? s->mk_part_units_xg : s->mk_part_units;
    p.attn_out = s->attn_out; p.attn_counters = s->attn_counters;
    p.n_splits = (int)S; p.advance = advance ? 1 : 0; p.tok_base = 0; p.eps = c.norm_eps;
    p.h = s->h; p.q = s->qbuf; p.act = s->act; p.part_o = s->part_o; p.part_ml = s->part_ml;
    p.logits = c.has_head ? s->logits : nullptr; p.arg_val = s->arg_val; p.arg_idx = s->arg_idx;
    p.next_token = s->next_tok; p.ids = s->ids; p.pos = s->pos; p.seq = s->seq; p.tok_ring = s->tok_ring;
    p.block_table = s->block_table; p.pages_per_seq = s->pages_per_seq; p.page_shift = s->page_shift;
    p.wte = s->io.wte; p.wpe = c.arch == DLLM_ARCH_GPT2 ? s->io.wpe : nullptr;
    p.final_norm_w = s->io.final_norm_w; p.final_norm_b = s->io.final_norm_b;
    p.rope_cos = s->io.rope_cos; p.rope_sin = s->io.rope_sin;
    p.barrier = s->mk_sync; p.done = s->mk_sync + 1; p.arg_counter = s->mk_sync + 2;
    if (io) {  // read in place, mirror into the handle's device copies
        p.ids = const_cast<int32_t*>(io->ids); p.pos = const_cast<int32_t*>(io->pos); p.seq = io->seq;
        p.mirror_ids = s->ids; p.mirror_pos = s->pos; p.mirror_seq = s->seq;
        p.next_token_host = io->next_token;
    }
    p.n_micro = 1;
    p.blocked = 1;
    if (const char* e = getenv("DLLM_MK_BLOCKED")) p.blocked = atoi(e) != 0;  // 0: panel p -> CTA p % grid (A/B measurements)
    if (pipe) {
        p.n_micro = pipe->n_micro; p.ticket_base = pipe->step_base; p.flag_stride = kFlagStride;
        p.status = s->pipe_status; p.ticket_timeout_ns = ticket_timeout_ns();
        p.out_phase = s->mk_out_phase;
        p.in_flag = pipe->in->flags(); p.in_tokens = pipe->in->tokens();
        p.in_hidden = c.has_embed ? nullptr : pipe->in->hidden_rows();
        p.out_flag = pipe->out->flags(); p.out_tokens = pipe->out->tokens();
        p.out_hidden = c.has_head ? nullptr : pipe->out->hidden_rows();
        p.logits = nullptr;
    }
    if (s->num_sms > s->arg_grid_cap) return fail("internal: argmax scratch too small for the persistent kernel");
    const char* trace_path = getenv("DLLM_MK_TRACE");
    long long* d_trace = nullptr;
    const size_t trace_n = (size_t)n_steps * s->mk_n_phases * MK_TRACE_SLOTS;
    if (trace_path && n_steps > 1) {  // debug: per-phase SM-clock stamps of CTA 0, dumped after a device sync
        CUDA_TRY(cudaMalloc((void**)&d_trace, trace_n * 8));
        CUDA_TRY(cudaMemsetAsync(d_trace, 0, trace_n * 8, st));
        p.trace = d_trace;
    }
    CUDA_TRY(launch_decode_megakernel(p, s->num_sms, st));
    if (d_trace) {
        std::vector<long long> h(trace_n);
        CUDA_TRY(cudaStreamSynchronize(st));
        CUDA_TRY(cudaMemcpy(h.data(), d_trace, trace_n * 8, cudaMemcpyDeviceToHost));
        cudaFree(d_trace);
        std::vector<MkPhase> ph(s->mk_n_phases);
        CUDA_TRY(cudaMemcpy(ph.data(), s->mk_phases, ph.size() * sizeof(MkPhase), cudaMemcpyDeviceToHost));
        if (FILE* f = fopen(trace_path, "w")) {
            fprintf(f, "step,phase,kind,layer,t_start,t_prologue,t_main,t_epilogue,c_wait,c_mma,c_issue,n_stages,b_enter,b_arrived,b_done,"
                       "p_loaded,p_reduced,p_stored,e_summed,e_done\n");
            for (int i = 0; i < n_steps; ++i)
                for (int j = 0; j < s->mk_n_phases; ++j) {
                    const long long* t = &h[((size_t)i * s->mk_n_phases + j) * MK_TRACE_SLOTS];
                    fprintf(f, "%d,%d,%d,%d,%lld,%lld,%lld,%lld,%lld,%lld,%lld,%lld,%lld,%lld,%lld,%lld,%lld,%lld,%lld,%lld\n", i, j,
                            (int)ph[j].kind, (int)ph[j].layer, t[0], t[1], t[2], t[3], t[4], t[5], t[6], t[7], t[8], t[9], t[10], t[11],
                            t[12], t[13], t[14], t[15]);
                }
            fclose(f);
        }
    }
    return 0;
}

// Prefill (many rows, bf16): the dense contractions run as tcgen05 GEMMs, the rest as row-wise kernels.
// If the caller's positions / seq ids are host-readable and the rows are `n_seqs` runs of `len` consecutive positions of
// one sequence each (what a [B, T] prompt chunk is), return true and the run shape: the flash-style prefill attention
// kernel applies.  Otherwise the per-row split-K kernel is used.
static bool detect_runs(const int32_t* positions, const int32_t* seq_ids, int n_rows, int* n_seqs, int* len) {
    cudaPointerAttributes ap{}, as{};
    if (cudaPointerGetAttributes(&ap, positions) != cudaSuccess || cudaPointerGetAttributes(&as, seq_ids) != cudaSuccess) {
        cudaGetLastError();
        return false;
    }
    if ((ap.type != cudaMemoryTypeHost && ap.type != cudaMemoryTypeUnregistered) ||
        (as.type != cudaMemoryTypeHost && as.type != cudaMemoryTypeUnregistered))
        return false;
    int run = 1;
    while (run < n_rows && seq_ids[run] == seq_ids[0] && positions[run] == positions[0] + run) ++run;
    if (n_rows % run) return false;
    for (int b = 0; b < n_rows / run; ++b)
        for (int i = 0; i < run; ++i)
            if (seq_ids[b * run + i] != seq_ids[b * run] || positions[b * run + i] != positions[b * run] + i) return false;
    *n_seqs = n_rows / run;
    *len = run;
    return true;
}

static int enqueue_prefill(dllm_shard* s, int n_rows, int n_logit_rows, int run_seqs, int run_len, cudaStream_t st) {
    const dllm_shard_config& c = s->cfg;
    const bool gpt2 = c.arch == DLLM_ARCH_GPT2;
    const int norm = gpt2 ? NORM_LAYER : NORM_RMS;
    const int Hd = c.n_heads * c.head_dim, KVd = c.n_kv_heads * c.head_dim, QKVd = Hd + 2 * KVd;
    if (c.has_embed) {
        EmbedParams e{};
        e.ids = s->ids; e.positions = s->pos; e.wte = s->io.wte; e.wpe = gpt2 ? s->io.wpe : nullptr;
        e.out = s->h; e.n_rows = n_rows; e.hidden = c.hidden; e.vocab = c.vocab;
        CUDA_TRY(launch_embed(0, e, st));
    }
    const int splits = pick_splits(s, n_rows);
    char* qkv = static_cast<char*>(s->qkv_tmp);
    for (int l = 0; l < c.n_layers; ++l) {
        const dllm_layer_weights& w = s->layers[l];
        PagedKV kv{s->kpool[l], s->vpool[l], s->block_table, s->pages_per_seq, s->page_shift};
        if (has_attn(c, l)) {
        CUDA_TRY(launch_norm_rows(s->h, w.norm1_w, w.norm1_b, s->xn, n_rows, c.hidden, norm, c.norm_eps, st));
        const bool qkv_fused = s->esz == 2 && gemm_qkv_fusable(n_rows, c.head_dim, w.bq, w.bk, w.bv);
        {   // q, k, v projections: row segments of one launch
            GemmTcParams g{};
            g.split_ws = s->gemm_ws; g.split_tickets = s->gemm_tickets; g.split_ws_bytes = s->gemm_ws_bytes;
            g.a = s->xn; g.M = n_rows; g.K = c.hidden; g.ldc = QKVd; g.epi = qkv_fused ? GEMM_EPI_QKV : GEMM_EPI_BIAS;
            g.w = w.wq; g.bias = w.bq; g.c = qkv; g.N = Hd;
            g.w_more[0] = w.wk; g.bias_more[0] = w.bk; g.n_more[0] = KVd;
            g.w_more[1] = w.wv; g.bias_more[1] = w.bv; g.n_more[1] = KVd;
            if (qkv_fused) {   // rotary and the KV append happen in the projection's epilogue; no [R, QKVd] round trip
                g.q_out = s->qbuf; g.kv = kv; g.positions = s->pos; g.seq_ids = s->seq;
                g.rope_cos = s->io.rope_cos; g.rope_sin = s->io.rope_sin;
                g.n_heads = c.n_heads; g.n_kv_heads = c.n_kv_heads; g.head_dim = c.head_dim; g.rope = !gpt2;
            }
            CUDA_TRY(launch_gemm_tc(g, st));
        }
        if (!qkv_fused) {   // <= 64 rows (swap-AB projection) or an odd head size: separate rotary + append pass
            RopeAppendParams ra{};
            ra.qkv = qkv; ra.q = s->qbuf; ra.kv = kv; ra.positions = s->pos; ra.seq_ids = s->seq;
            ra.rope_cos = s->io.rope_cos; ra.rope_sin = s->io.rope_sin;
            ra.n_rows = n_rows; ra.n_heads = c.n_heads; ra.n_kv_heads = c.n_kv_heads; ra.head_dim = c.head_dim; ra.rope = !gpt2;
            CUDA_TRY(launch_rope_append(ra, st));
        }
        if (run_len > 0) {   // causal flash-style attention over each run of consecutive positions
            PrefillAttnParams a{};
            a.q = s->qbuf; a.kv = kv; a.positions = s->pos; a.seq_ids = s->seq; a.out = s->attn_out;
            a.n_seqs = run_seqs; a.rows_per_seq = run_len; a.n_heads = c.n_heads; a.n_kv_heads = c.n_kv_heads;
            a.head_dim = c.head_dim;
            a.num_pages = s->num_pages;
            CUDA_TRY(launch_prefill_attn(a, st));
        } else {
            AttnParams a{};
            a.q = s->qbuf; a.kv = kv; a.positions = s->pos; a.seq_ids = s->seq;
            a.n_rows = n_rows; a.n_heads = c.n_heads; a.n_kv_heads = c.n_kv_heads; a.head_dim = c.head_dim;
            a.n_splits = splits; a.part_o = s->part_o; a.part_ml = s->part_ml; a.counters = s->attn_counters;
            a.out = s->attn_out;
            CUDA_TRY(launch_attn_decode(0, a, st));
        }
        {   // o-proj + residual (in place on the residual stream)
            GemmTcParams g{};
            g.split_ws = s->gemm_ws; g.split_tickets = s->gemm_tickets; g.split_ws_bytes = s->gemm_ws_bytes;
            g.a = s->attn_out; g.w = w.wo; g.bias = w.bo; g.c = s->h; g.M = n_rows; g.N = c.hidden; g.K = Hd;
            g.ldc = c.hidden; g.epi = GEMM_EPI_RESIDUAL;
            CUDA_TRY(launch_gemm_tc(g, st));
        }
        }  // attention block
        if (!has_mlp(c, l)) continue;
        CUDA_TRY(launch_norm_rows(s->h, w.norm2_w, w.norm2_b, s->xn, n_rows, c.hidden, norm, c.norm_eps, st));
        {
            GemmTcParams g{};
            g.split_ws = s->gemm_ws; g.split_tickets = s->gemm_tickets; g.split_ws_bytes = s->gemm_ws_bytes;
            g.a = s->xn; g.c = s->act; g.M = n_rows; g.N = c.ffn; g.K = c.hidden; g.ldc = c.ffn;
            if (gpt2) { g.w = w.w_up; g.bias = w.b_up; g.epi = GEMM_EPI_GELU; }
            else { g.w = w.w_gate; g.w2 = w.w_up; g.epi = GEMM_EPI_SWIGLU; }
            CUDA_TRY(launch_gemm_tc(g, st));
        }
        {
            GemmTcParams g{};
            g.split_ws = s->gemm_ws; g.split_tickets = s->gemm_tickets; g.split_ws_bytes = s->gemm_ws_bytes;
            g.a = s->act; g.w = w.w_down; g.bias = w.b_down; g.c = s->h; g.M = n_rows; g.N = c.hidden; g.K = c.ffn;
            g.ldc = c.hidden; g.epi = GEMM_EPI_RESIDUAL;
            CUDA_TRY(launch_gemm_tc(g, st));
        }
    }
    if (c.has_head && n_logit_rows > 4 && n_logit_rows <= 64 && c.hidden <= 16384 && gemm_swapab_enabled()) {
        // lm_head for 5..64 rows: final norm of the wanted rows -> swap-AB GEMM (one pass over the head) -> argmax
        CUDA_TRY(launch_norm_rows(s->h, s->io.final_norm_w, s->io.final_norm_b, s->xn, n_logit_rows, c.hidden, norm,
                                  c.norm_eps, st, s->logit_rows));
        GemmTcParams g{};
            g.split_ws = s->gemm_ws; g.split_tickets = s->gemm_tickets; g.split_ws_bytes = s->gemm_ws_bytes;
        g.a = s->xn; g.w = s->io.lm_head; g.c = s->logits; g.M = n_logit_rows; g.N = c.vocab; g.K = c.hidden;
        g.ldc = c.vocab; g.epi = GEMM_EPI_LOGITS;
        CUDA_TRY(launch_gemm_tc(g, st));
        CUDA_TRY(launch_argmax_rows(s->logits, n_logit_rows, c.vocab, s->next_tok, st));
    } else if (c.has_head && n_logit_rows > 0) {
        GemvParams p{};
        p.x = s->h; p.row_map = s->logit_rows; p.n_rows = n_logit_rows; p.K = c.hidden;
        p.norm = norm; p.norm_w = s->io.final_norm_w; p.norm_b = s->io.final_norm_b; p.eps = c.norm_eps;
        p.w[0] = s->io.lm_head; p.n[0] = c.vocab; p.epi = EPI_LOGITS;
        p.logits = s->logits; p.arg_val = s->arg_val; p.arg_idx = s->arg_idx; p.counter = s->arg_counter;
        p.next_token = s->next_tok;
        if (gemv_grid_size(0, p, s->num_sms) > s->arg_grid_cap) return fail("internal: argmax scratch too small");
        CUDA_TRY(launch_gemv(0, p, s->num_sms, st));
    }
    return 0;
}

static bool prefill_eligible(const dllm_shard* s, int n_rows) {
    static const bool off = [] { const char* e = getenv("DLLM_PREFILL_GEMM"); return e && e[0] == '0'; }();
    const dllm_shard_config& c = s->cfg;
    // more than 4 rows: the GEMM streams every weight once (the GEMV kernels re-read them per 4 rows)
    return !off && c.dtype == DLLM_DTYPE_BF16 && n_rows > 4 && s->xn && c.n_layers > 0 &&
           c.hidden % 8 == 0 && c.ffn % 8 == 0 && (c.n_heads * c.head_dim) % 8 == 0;
}

// Replay (capturing on first use) the CUDA graph of one step.  mode 0: per-op decode kernels (large row counts launch
// directly); mode 1: the tcgen05 GEMM path (prefill chunks and decode batches the persistent kernel cannot take).
static int run_step(dllm_shard* s, int n_rows, int n_logit_rows, bool advance, cudaStream_t st, int mode = 0,
                    int run_seqs = 0, int run_len = 0) {
    constexpr int kGraphMaxRows = 64;
    auto enqueue = [&](cudaStream_t q) {
        return mode == 1 ? enqueue_prefill(s, n_rows, n_logit_rows, run_seqs, run_len, q)
                         : enqueue_step(s, n_rows, n_logit_rows, advance, q);
    };
    if (mode == 0 && n_rows > kGraphMaxRows) return enqueue(st);
    const unsigned long long key = ((unsigned long long)n_rows << 40) | ((unsigned long long)run_len << 20) |
                                   ((unsigned long long)n_logit_rows << 2) | ((unsigned long long)mode << 1) | (advance ? 1 : 0);
    auto it = s->graphs.find(key);
    if (it == s->graphs.end()) {
        cudaGraph_t graph = nullptr;
        const unsigned long long before = launch_count();
        CUDA_TRY(cudaStreamBeginCapture(s->cap_stream, cudaStreamCaptureModeThreadLocal));
        const int rc = enqueue(s->cap_stream);
        const cudaError_t ce = cudaStreamEndCapture(s->cap_stream, &graph);
        if (rc) {
            if (graph) cudaGraphDestroy(graph);
            cudaGetLastError();
            return rc;
        }
        CUDA_TRY(ce);
        GraphEntry ge;
        ge.kernels = launch_count() - before;
        count_launch(0ULL - ge.kernels);  // capture enqueued nothing; replays are what count
        const cudaError_t ie = cudaGraphInstantiate(&ge.exec, graph, 0);
        cudaGraphDestroy(graph);
        CUDA_TRY(ie);
        it = s->graphs.emplace(key, ge).first;
    }
    CUDA_TRY(cudaGraphLaunch(it->second.exec, st));
    count_launch(it->second.kernels);
    return 0;
}

// ---- the hot path ------------------------------------------------------------------------------------

extern "C" int dllm_shard_forward(dllm_shard* s, const void* input, const int32_t* positions,
                                  const int32_t* seq_ids, int32_t n_rows, void* hidden_out,
                                  const int32_t* logit_rows, int32_t n_logit_rows, float* logits_out,
                                  int32_t* next_token_out, void* stream) {
    if (!s) return fail("dllm_shard_forward: null shard");
    if (!input || !positions || !seq_ids) return fail("dllm_shard_forward: input, positions and seq_ids are required");
    const dllm_shard_config& c = s->cfg;
    if (n_rows <= 0 || n_rows > c.max_rows) return fail("dllm_shard_forward: n_rows out of range (max_rows)");
    cudaStream_t st = (cudaStream_t)stream;
    int dev = -1;
    cudaGetDevice(&dev);
    if (dev != s->device) CUDA_TRY(cudaSetDevice(s->device));
    int nl = 0;
    if (c.has_head) {
        if (logit_rows) {
            if (n_logit_rows <= 0 || n_logit_rows > n_rows) return fail("dllm_shard_forward: n_logit_rows out of range");
            nl = n_logit_rows;
        } else if (n_logit_rows >= 0) {
            nl = n_rows;
        }  // n_logit_rows < 0 with no row list: no logits this call (an inner chunk of a chunked prefill)
    }
    const bool use_mk = mk_eligible(s, n_rows, nl, logit_rows == nullptr);
    // A token-by-token host loop (the master's relay, bench.py's e2e figure) pays for every copy node in front of the launch.
    // When the stage owns the embedding and the caller's int32 buffers are pinned host (or device) memory, the persistent
    // kernel reads them in place and writes the sampled ids straight back: the call is one launch, no copies.
    MkHostIo io{};
    bool in_place = false;
    if (use_mk && c.has_embed && !hidden_out && !logits_out) {
        io.ids = static_cast<const int32_t*>(device_visible(input));
        io.pos = static_cast<const int32_t*>(device_visible(positions));
        io.seq = static_cast<const int32_t*>(device_visible(seq_ids));
        io.next_token = next_token_out ? static_cast<int32_t*>(const_cast<void*>(device_visible(next_token_out))) : nullptr;
        in_place = io.ids && io.pos && io.seq && (!next_token_out || io.next_token);
    }
    if (!in_place) {
        if (c.has_embed) CUDA_TRY(cudaMemcpyAsync(s->ids, input, (size_t)n_rows * 4, cudaMemcpyDefault, st));
        else CUDA_TRY(cudaMemcpyAsync(s->h, input, (size_t)n_rows * c.hidden * s->esz, cudaMemcpyDefault, st));
        CUDA_TRY(cudaMemcpyAsync(s->pos, positions, (size_t)n_rows * 4, cudaMemcpyDefault, st));
        CUDA_TRY(cudaMemcpyAsync(s->seq, seq_ids, (size_t)n_rows * 4, cudaMemcpyDefault, st));
    }
    if (c.has_head && logit_rows) {
        CUDA_TRY(cudaMemcpyAsync(s->logit_rows, logit_rows, (size_t)nl * 4, cudaMemcpyDefault, st));
        s->arange_rows = 0;
    }
    // every path leaves logit_rows describing this call's logit rows: a decode loop that follows gathers its
    // positions / slots through them (the persistent kernel itself takes every row's logits in place)
    if (c.has_head && !logit_rows && nl > s->arange_rows) {
        CUDA_TRY(cudaMemcpyAsync(s->logit_rows, s->arange, (size_t)nl * 4, cudaMemcpyDeviceToDevice, st));
        s->arange_rows = nl;  // (a shorter identity prefix is already in place)
    }
    s->pos_at_pending = false;
    if (use_mk) {
        if (run_megakernel(s, n_rows, 1, false, st, nullptr, in_place ? &io : nullptr)) return 1;
        if (in_place) return 0;  // the sampled ids are already where the caller wants them
    } else if (prefill_eligible(s, n_rows)) {
        int run_seqs = 0, run_len = 0;
        static const bool fa_off = [] { const char* e = getenv("DLLM_PREFILL_ATTN"); return e && e[0] == '0'; }();
        if (fa_off || !detect_runs(positions, seq_ids, n_rows, &run_seqs, &run_len)) run_seqs = run_len = 0;
        if (run_len < 8) run_seqs = run_len = 0;  // decode-like rows: the split-K kernel spreads one row's keys over CTAs
        if (run_step(s, n_rows, nl, false, st, 1, run_seqs, run_len)) return 1;
    } else if (run_step(s, n_rows, nl, false, st)) {
        return 1;
    }
    if (hidden_out) CUDA_TRY(cudaMemcpyAsync(hidden_out, s->h, (size_t)n_rows * c.hidden * s->esz, cudaMemcpyDefault, st));
    if (c.has_head && logits_out)
        CUDA_TRY(cudaMemcpyAsync(logits_out, s->logits, (size_t)nl * c.vocab * 4, cudaMemcpyDefault, st));
    if (c.has_head && next_token_out)
        CUDA_TRY(cudaMemcpyAsync(next_token_out, s->next_tok, (size_t)nl * 4, cudaMemcpyDefault, st));
    return 0;
}

extern "C" int dllm_shard_decode_loop(dllm_shard* s, int32_t n_rows, int32_t n_steps, int32_t* tokens_out,
                                      void* stream) {
    if (!s) return fail("dllm_shard_decode_loop: null shard");
    const dllm_shard_config& c = s->cfg;
    if (!c.has_embed || !c.has_head) return fail("dllm_shard_decode_loop: the shard must own both the embedding and the head");
    if (n_rows <= 0 || n_rows > c.max_rows || n_rows > 1024) return fail("dllm_shard_decode_loop: n_rows out of range");
    if (n_steps <= 0 || (size_t)n_steps * n_rows > s->tok_ring_cap) return fail("dllm_shard_decode_loop: n_steps out of range");
    cudaStream_t st = (cudaStream_t)stream;
    int dev = -1;
    cudaGetDevice(&dev);
    if (dev != s->device) CUDA_TRY(cudaSetDevice(s->device));
    // after a forward the pending token sits one past the processed rows; after a loop `pos` is already there
    CUDA_TRY(launch_decode_begin(s->ids, s->pos, s->seq, s->next_tok, s->logit_rows, n_rows, s->step_ctr,
                                 s->pos_at_pending ? 0 : 1, st));
    s->pos_at_pending = true;
    CUDA_TRY(cudaMemcpyAsync(s->logit_rows, s->arange, (size_t)n_rows * 4, cudaMemcpyDeviceToDevice, st));
    if (mk_eligible(s, n_rows, n_rows, true)) {
        if (run_megakernel(s, n_rows, n_steps, true, st)) return 1;
    } else {
        const bool gemm = prefill_eligible(s, n_rows);
        for (int i = 0; i < n_steps; ++i) {
            if (gemm) {  // GEMM-path graph for the layers, then the device-side token feedback
                if (run_step(s, n_rows, n_rows, false, st, 1, 0, 0)) return 1;
                CUDA_TRY(launch_advance(s->ids, s->pos, s->next_tok, s->tok_ring, s->step_ctr, n_rows, st));
            } else if (run_step(s, n_rows, n_rows, true, st)) {
                return 1;
            }
        }
    }
    if (tokens_out)
        CUDA_TRY(cudaMemcpyAsync(tokens_out, s->tok_ring, (size_t)n_steps * n_rows * 4, cudaMemcpyDefault, st));
    return 0;
}

// ---- hand-off ---------------------------------------------------------------------------------------

// rows: up to 16 for the persistent decode loop (checked again there); a prefill tile slot may hold a whole prompt chunk
constexpr int kMailboxMaxRows = 1 << 20;
static int mailbox_shape_ok(int n_slots, int rows, int hidden) {
    return n_slots >= 1 && n_slots <= 64 && rows >= 1 && rows <= kMailboxMaxRows && hidden >= 8 && hidden % 8 == 0 &&
           (size_t)n_slots * rows * hidden * 2 <= ((size_t)8 << 30);
}

extern "C" int dllm_mailbox_create(int32_t device, int32_t n_slots, int32_t rows, int32_t hidden, dllm_mailbox** out) {
    if (!out) return fail("dllm_mailbox_create: null out");
    if (!mailbox_shape_ok(n_slots, rows, hidden)) return fail("dllm_mailbox_create: bad shape (1..64 slots, rows >= 1, hidden % 8 == 0, at most 8 GiB)");
    if (dllm_device_count() == 0) return fail("dllm_mailbox_create: no CUDA device (this library has no CPU path)");
    CUDA_TRY(cudaSetDevice(device));
    dllm_mailbox* b = new dllm_mailbox();
    b->device = device; b->n_slots = n_slots; b->rows = rows; b->hidden = hidden; b->owner = true;
    cudaError_t e = cudaMalloc((void**)&b->base, b->bytes());
    if (e == cudaSuccess) e = cudaMemset(b->base, 0, b->bytes());
    if (e != cudaSuccess) {
        if (b->base) cudaFree(b->base);
        delete b;
        CUDA_TRY(e);
    }
    *out = b;
    return 0;
}

extern "C" int dllm_mailbox_export(dllm_mailbox* b, void* ipc_handle_out) {
    if (!b || !b->owner || !ipc_handle_out) return fail("dllm_mailbox_export: needs the owner's mailbox and a 64-byte buffer");
    static_assert(sizeof(cudaIpcMemHandle_t) == DLLM_MAILBOX_IPC_BYTES, "IPC handle size");
    cudaIpcMemHandle_t h;
    CUDA_TRY(cudaSetDevice(b->device));
    CUDA_TRY(cudaIpcGetMemHandle(&h, b->base));
    memcpy(ipc_handle_out, &h, sizeof(h));
    return 0;
}

extern "C" int dllm_mailbox_import(int32_t device, const void* ipc_handle, int32_t n_slots, int32_t rows, int32_t hidden,
                                   dllm_mailbox** out) {
    if (!out || !ipc_handle) return fail("dllm_mailbox_import: null argument");
    if (!mailbox_shape_ok(n_slots, rows, hidden)) return fail("dllm_mailbox_import: bad shape");
    CUDA_TRY(cudaSetDevice(device));
    cudaIpcMemHandle_t h;
    memcpy(&h, ipc_handle, sizeof(h));
    void* ptr = nullptr;
    CUDA_TRY(cudaIpcOpenMemHandle(&ptr, h, cudaIpcMemLazyEnablePeerAccess));
    dllm_mailbox* b = new dllm_mailbox();
    b->device = device; b->n_slots = n_slots; b->rows = rows; b->hidden = hidden; b->base = (char*)ptr; b->ipc = true;
    *out = b;
    return 0;
}

extern "C" int dllm_mailbox_view(dllm_mailbox* box, int32_t device, dllm_mailbox** out) {
    if (!box || !out) return fail("dllm_mailbox_view: null argument");
    if (device != box->device) {
        int can = 0;
        CUDA_TRY(cudaDeviceCanAccessPeer(&can, device, box->device));
        if (!can) return fail("dllm_mailbox_view: no peer access between the two devices");
        CUDA_TRY(cudaSetDevice(device));
        cudaError_t e = cudaDeviceEnablePeerAccess(box->device, 0);
        if (e != cudaSuccess && e != cudaErrorPeerAccessAlreadyEnabled) CUDA_TRY(e);
        cudaGetLastError();
    }
    dllm_mailbox* b = new dllm_mailbox(*box);
    b->device = device; b->owner = false; b->ipc = false;
    *out = b;
    return 0;
}

extern "C" int dllm_mailbox_post_tokens(dllm_mailbox* b, int32_t slot, const int32_t* tokens, uint32_t ticket, void* stream) {
    if (!b || !b->owner || !tokens || slot < 0 || slot >= b->n_slots) return fail("dllm_mailbox_post_tokens: bad argument");
    cudaStream_t st = (cudaStream_t)stream;
    CUDA_TRY(cudaSetDevice(b->device));
    CUDA_TRY(cudaMemcpyAsync(b->tokens() + (size_t)slot * b->rows, tokens, (size_t)b->rows * 4, cudaMemcpyDefault, st));
    // after the ids, stream-ordered (a copy from a pageable host word is staged before the call returns)
    CUDA_TRY(cudaMemcpyAsync(b->flags() + (size_t)slot * kFlagStride, &ticket, 4, cudaMemcpyHostToDevice, st));
    CUDA_TRY(cudaStreamSynchronize(st));
    return 0;
}

extern "C" int dllm_mailbox_read_tokens(dllm_mailbox* b, int32_t slot, int32_t* tokens_out, uint32_t* ticket_out, void* stream) {
    if (!b || !b->owner || !tokens_out || slot < 0 || slot >= b->n_slots) return fail("dllm_mailbox_read_tokens: bad argument");
    cudaStream_t st = (cudaStream_t)stream;
    CUDA_TRY(cudaSetDevice(b->device));
    CUDA_TRY(cudaMemcpyAsync(tokens_out, b->tokens() + (size_t)slot * b->rows, (size_t)b->rows * 4, cudaMemcpyDefault, st));
    if (ticket_out) CUDA_TRY(cudaMemcpyAsync(ticket_out, b->flags() + (size_t)slot * kFlagStride, 4, cudaMemcpyDeviceToHost, st));
    CUDA_TRY(cudaStreamSynchronize(st));
    return 0;
}

// Prefill hand-off through a mailbox slot: the producer stage copies its [n_rows, hidden] tile straight into the next
// stage's slot (peer memory, NVLink DMA), the consumer copies it out of its own slot.  Ordering between the two is the
// caller's (a control message after the producer's stream has drained); no ticket is involved.
extern "C" int dllm_mailbox_put_rows(dllm_mailbox* b, int32_t slot, const void* d_src, int32_t n_rows, void* stream) {
    if (!b || !d_src || slot < 0 || slot >= b->n_slots || n_rows < 1 || n_rows > b->rows) return fail("dllm_mailbox_put_rows: bad argument");
    char* dst = static_cast<char*>(b->hidden_rows()) + (size_t)slot * b->rows * b->hidden * 2;
    // the mailbox (own, peer view or IPC mapping) is addressable from b->device: the hand-off is one device-to-device copy
    return dllm_handoff_peer_copy(dst, b->device, d_src, b->device, (size_t)n_rows * b->hidden * 2, stream);
}

extern "C" int dllm_mailbox_get_rows(dllm_mailbox* b, int32_t slot, void* d_dst, int32_t n_rows, void* stream) {
    if (!b || !b->owner || !d_dst || slot < 0 || slot >= b->n_slots || n_rows < 1 || n_rows > b->rows) return fail("dllm_mailbox_get_rows: bad argument");
    const char* src = static_cast<const char*>(b->hidden_rows()) + (size_t)slot * b->rows * b->hidden * 2;
    return dllm_handoff_peer_copy(d_dst, b->device, src, b->device, (size_t)n_rows * b->hidden * 2, stream);
}

extern "C" int dllm_mailbox_destroy(dllm_mailbox* b) {
    if (!b) return 0;
    cudaSetDevice(b->device);
    if (b->owner && b->base) {
        cudaDeviceSynchronize();
        cudaFree(b->base);
    } else if (b->ipc && b->base) {
        cudaIpcCloseMemHandle(b->base);
    }
    delete b;
    return 0;
}

// Everything dllm_shard_decode_pipelined checks before it launches, without launching (also exported on its own: a master
// asks every stage first and starts the loops only when all of them can run -- a stage that refuses after its neighbours
// have started would leave them waiting for tickets that never come).
static int pipelined_validate(const char* who, dllm_shard* s, dllm_mailbox* inbox, dllm_mailbox* outbox, int32_t n_micro,
                              int32_t rows, int32_t n_steps) {
    const std::string w(who);
    if (!s || !inbox || !outbox) return fail(w + ": null argument");
    const dllm_shard_config& c = s->cfg;
    if (n_micro < 1 || rows < 1 || n_steps < 1) return fail(w + ": n_micro, rows and n_steps must be positive");
    if ((long)n_micro * rows > c.max_rows) return fail(w + ": n_micro * rows exceeds max_rows");
    if (!inbox->owner || inbox->device != s->device || outbox->device != s->device)
        return fail(w + ": the inbox must be this stage's own mailbox and the outbox a view mapped on this stage's device");
    for (const dllm_mailbox* b : {inbox, outbox})
        if (b->n_slots < n_micro || b->rows != rows || b->hidden != c.hidden) return fail(w + ": mailbox shape does not match");
    if (c.has_embed && (size_t)n_steps * n_micro * rows > s->tok_ring_cap) return fail(w + ": n_steps out of range");
    if (c.n_layers < 1) return fail(w + ": the stage has no layers");
    if (!mk_eligible(s, rows, rows, true) && !(c.dtype == DLLM_DTYPE_BF16 && rows <= kPipeOpsMaxRows))
        return fail(w + ": shape taken neither by the persistent kernel nor by the per-op loop (bf16, <= 64 rows)");
    return 0;
}

// The pipelined loop for shapes the persistent kernel does not take: every micro-batch step is [wait kernel] -> the
// step's CUDA graph (swap-AB tcgen05 GEMMs, grouped-head attention) -> [send kernel], all enqueued up front.
static int run_pipelined_ops(dllm_shard* s, const MkPipe& pipe, int rows, int n_steps, cudaStream_t st) {
    const dllm_shard_config& c = s->cfg;
    const int mode = prefill_eligible(s, rows) ? 1 : 0;
    const int nl = c.has_head ? rows : 0;
    if (c.has_head) {
        CUDA_TRY(cudaMemcpyAsync(s->logit_rows, s->arange, (size_t)rows * 4, cudaMemcpyDeviceToDevice, st));
        s->arange_rows = std::max(s->arange_rows, rows);
    }
    for (int step = 0; step < n_steps * pipe.n_micro; ++step) {
        const int m = step % pipe.n_micro, g = step / pipe.n_micro;
        PipeBeginParams b{};
        b.in_flag = pipe.in->flags() + (size_t)m * kFlagStride; b.want = pipe.step_base + g + 1;
        b.in_tokens = c.has_embed ? pipe.in->tokens() + (size_t)m * rows : nullptr;
        b.in_hidden = c.has_embed ? nullptr : static_cast<char*>(pipe.in->hidden_rows()) + (size_t)m * rows * c.hidden * 2;
        b.pos_all = s->pipe_pos + (size_t)m * rows; b.seq_all = s->pipe_seq + (size_t)m * rows; b.pos_add = g;
        b.ids = s->ids; b.pos = s->pos; b.seq = s->seq; b.h = s->h;
        b.tok_ring = c.has_embed ? s->tok_ring + (size_t)step * rows : nullptr;
        b.n_rows = rows; b.hidden = c.hidden; b.status = s->pipe_status; b.timeout_ns = ticket_timeout_ns();
        CUDA_TRY(launch_pipe_begin(b, st));
        if (run_step(s, rows, nl, false, st, mode, 0, 0)) return 1;
        PipeEndParams e{};
        e.out_flag = pipe.out->flags() + (size_t)m * kFlagStride;
        e.ticket = pipe.step_base + g + (c.has_head ? 2 : 1);   // the first stage waits for step g + 1's ids
        e.out_tokens = c.has_head ? pipe.out->tokens() + (size_t)m * rows : nullptr;
        e.next_tok = s->next_tok;
        e.out_hidden = c.has_head ? nullptr : static_cast<char*>(pipe.out->hidden_rows()) + (size_t)m * rows * c.hidden * 2;
        e.h = s->h; e.n_rows = rows; e.hidden = c.hidden; e.status = s->pipe_status;
        CUDA_TRY(launch_pipe_end(e, st));
    }
    s->pos_at_pending = false;
    return 0;
}

extern "C" int dllm_shard_decode_pipelined_check(dllm_shard* s, dllm_mailbox* inbox, dllm_mailbox* outbox, int32_t n_micro,
                                                 int32_t rows, int32_t n_steps) {
    return pipelined_validate("dllm_shard_decode_pipelined_check", s, inbox, outbox, n_micro, rows, n_steps);
}

extern "C" int dllm_shard_decode_pipelined(dllm_shard* s, dllm_mailbox* inbox, dllm_mailbox* outbox, int32_t n_micro,
                                           int32_t rows, int32_t n_steps, uint32_t step_base, const int32_t* positions,
                                           const int32_t* seq_ids, int32_t* fed_tokens_out, void* stream) {
    if (!positions || !seq_ids) return fail("dllm_shard_decode_pipelined: null argument");
    if (pipelined_validate("dllm_shard_decode_pipelined", s, inbox, outbox, n_micro, rows, n_steps)) return 1;
    const dllm_shard_config& c = s->cfg;
    cudaStream_t st = (cudaStream_t)stream;
    int dev = -1;
    cudaGetDevice(&dev);
    if (dev != s->device) CUDA_TRY(cudaSetDevice(s->device));
    MkPipe pipe{inbox, outbox, n_micro, step_base};
    if (mk_eligible(s, rows, rows, true)) {
        CUDA_TRY(cudaMemcpyAsync(s->pos, positions, (size_t)n_micro * rows * 4, cudaMemcpyDefault, st));
        CUDA_TRY(cudaMemcpyAsync(s->seq, seq_ids, (size_t)n_micro * rows * 4, cudaMemcpyDefault, st));
        if (run_megakernel(s, rows, n_steps * n_micro, false, st, &pipe)) return 1;
    } else {
        CUDA_TRY(cudaMemcpyAsync(s->pipe_pos, positions, (size_t)n_micro * rows * 4, cudaMemcpyDefault, st));
        CUDA_TRY(cudaMemcpyAsync(s->pipe_seq, seq_ids, (size_t)n_micro * rows * 4, cudaMemcpyDefault, st));
        if (run_pipelined_ops(s, pipe, rows, n_steps, st)) return 1;
    }
    if (c.has_embed && fed_tokens_out)
        CUDA_TRY(cudaMemcpyAsync(fed_tokens_out, s->tok_ring, (size_t)n_steps * n_micro * rows * 4, cudaMemcpyDefault, st));
    return 0;
}

extern "C" int dllm_shard_pipeline_status(dllm_shard* s, void* stream) {
    if (!s) return fail("dllm_shard_pipeline_status: null shard");
    cudaStream_t st = (cudaStream_t)stream;
    CUDA_TRY(cudaSetDevice(s->device));
    CUDA_TRY(cudaStreamSynchronize(st));
    int status = 0;
    CUDA_TRY(cudaMemcpy(&status, s->pipe_status, 4, cudaMemcpyDeviceToHost));
    if (status == 0) return 0;
    CUDA_TRY(cudaMemset(s->pipe_status, 0, 4));
    if (s->mk_sync) CUDA_TRY(cudaMemset(s->mk_sync, 0, 16));   // barrier / exit counters of the abandoned launch
    return fail("dllm_shard_pipeline_status: the pipelined decode loop was abandoned -- a hand-off ticket did not arrive within " +
                std::to_string(ticket_timeout_ns() / 1000000ull) + " ms (the previous stage is not running, or is itself waiting); "
                "tokens and KV rows of that loop are incomplete: reset its sequences before reusing them");
}

extern "C" int dllm_handoff_peer_copy(void* dst_ptr, int dst_device, const void* src_ptr, int src_device,
                                      size_t bytes, void* stream) {
    if (!dst_ptr || !src_ptr) return fail("dllm_handoff_peer_copy: null pointer");
    if (bytes == 0) return 0;
    if (dst_device == src_device) {
        CUDA_TRY(cudaMemcpyAsync(dst_ptr, src_ptr, bytes, cudaMemcpyDeviceToDevice, (cudaStream_t)stream));
        return 0;
    }
    int can = 0;
    CUDA_TRY(cudaDeviceCanAccessPeer(&can, src_device, dst_device));
    if (can) {
        int cur = 0;
        cudaGetDevice(&cur);
        cudaSetDevice(src_device);
        cudaError_t e = cudaDeviceEnablePeerAccess(dst_device, 0);
        if (e != cudaSuccess && e != cudaErrorPeerAccessAlreadyEnabled) {
            cudaSetDevice(cur);
            CUDA_TRY(e);
        }
        cudaGetLastError();
        cudaSetDevice(cur);
    }
    CUDA_TRY(cudaMemcpyPeerAsync(dst_ptr, dst_device, src_ptr, src_device, bytes, (cudaStream_t)stream));
    return 0;
}

// ---- single operators ----------------------------------------------------------------------------------

extern "C" int dllm_op_gemv(int32_t dtype, const void* d_x, const void* d_w, const void* d_bias, void* d_y,
                            int32_t n_rows, int32_t n_out, int32_t k, int32_t accumulate, void* stream) {
    if (dtype != DLLM_DTYPE_BF16 && dtype != DLLM_DTYPE_F32) return fail("dllm_op_gemv: bad dtype");
    if (!d_x || !d_w || !d_y || n_rows <= 0 || n_out <= 0 || k <= 0) return fail("dllm_op_gemv: bad argument");
    if (dllm_device_count() == 0) return fail("dllm_op_gemv: no CUDA device (this library has no CPU path)");
    if (k % (dtype == DLLM_DTYPE_BF16 ? 8 : 4)) return fail("dllm_op_gemv: k must be a multiple of the 16-byte vector width");
    if (ensure_device_ready()) return 1;
    int dev = 0, sms = 148;
    cudaGetDevice(&dev);
    cudaDeviceGetAttribute(&sms, cudaDevAttrMultiProcessorCount, dev);
    GemvParams p{};
    p.x = d_x; p.n_rows = n_rows; p.K = k; p.norm = NORM_NONE;
    p.w[0] = d_w; p.b[0] = d_bias; p.n[0] = n_out; p.epi = EPI_RESIDUAL; p.y = d_y; p.accumulate = accumulate;
    CUDA_TRY(launch_gemv(dtype, p, sms, (cudaStream_t)stream));
    return 0;
}

extern "C" int dllm_op_gemm(const void* d_x, const void* d_w, const void* d_bias, void* d_y, int32_t m, int32_t n_out,
                            int32_t k, int32_t accumulate, void* stream) {
    if (!d_x || !d_w || !d_y || m <= 0 || n_out <= 0 || k <= 0) return fail("dllm_op_gemm: bad argument");
    if (k % 8 || n_out % 2) return fail("dllm_op_gemm: k must be a multiple of 8 and n_out even");
    if (dllm_device_count() == 0) return fail("dllm_op_gemm: no CUDA device (this library has no CPU path)");
    GemmTcParams g{};
    g.a = d_x; g.w = d_w; g.bias = d_bias; g.c = d_y; g.M = m; g.N = n_out; g.K = k; g.ldc = n_out;
    g.epi = accumulate ? GEMM_EPI_RESIDUAL : GEMM_EPI_BIAS;
    CUDA_TRY(launch_gemm_tc(g, (cudaStream_t)stream));
    return 0;
}

extern "C" size_t dllm_op_decode_attention_scratch_bytes(int32_t n_rows, int32_t n_heads, int32_t head_dim,
                                                         int32_t n_splits) {
    const size_t rh = (size_t)n_rows * n_heads;
    return align_up(rh * 4, 256) + align_up(rh * n_splits * 2 * 4, 256) + align_up(rh * n_splits * head_dim * 4, 256);
}

extern "C" int dllm_op_decode_attention(int32_t dtype, const void* d_q, const void* d_kpool, const void* d_vpool,
                                        const int32_t* d_block_table, int32_t pages_per_seq, int32_t page_size,
                                        const int32_t* d_positions, const int32_t* d_seq_ids, int32_t n_rows,
                                        int32_t n_heads, int32_t n_kv_heads, int32_t head_dim, int32_t n_splits,
                                        void* d_scratch, void* d_out, void* stream) {
    if (dtype != DLLM_DTYPE_BF16 && dtype != DLLM_DTYPE_F32) return fail("dllm_op_decode_attention: bad dtype");
    if (!d_q || !d_kpool || !d_vpool || !d_block_table || !d_positions || !d_seq_ids || !d_scratch || !d_out)
        return fail("dllm_op_decode_attention: null pointer");
    if (n_rows <= 0 || n_splits <= 0 || n_heads <= 0 || n_kv_heads <= 0 || n_heads % n_kv_heads)
        return fail("dllm_op_decode_attention: bad shape");
    if (head_dim != 32 && head_dim != 64 && head_dim != 128) return fail("dllm_op_decode_attention: head_dim must be 32, 64 or 128");
    if (page_size < 8 || (page_size & (page_size - 1))) return fail("dllm_op_decode_attention: page_size must be a power of two >= 8");
    if (dllm_device_count() == 0) return fail("dllm_op_decode_attention: no CUDA device (this library has no CPU path)");
    const size_t rh = (size_t)n_rows * n_heads;
    char* base = static_cast<char*>(d_scratch);
    AttnParams a{};
    a.counters = reinterpret_cast<uint32_t*>(base);
    a.part_ml = reinterpret_cast<float*>(base + align_up(rh * 4, 256));
    a.part_o = reinterpret_cast<float*>(base + align_up(rh * 4, 256) + align_up(rh * n_splits * 2 * 4, 256));
    a.q = d_q;
    a.kv.kpool = const_cast<void*>(d_kpool);
    a.kv.vpool = const_cast<void*>(d_vpool);
    a.kv.block_table = d_block_table;
    a.kv.pages_per_seq = pages_per_seq;
    a.kv.page_shift = 0;
    while ((1 << a.kv.page_shift) < page_size) ++a.kv.page_shift;
    a.positions = d_positions; a.seq_ids = d_seq_ids;
    a.n_rows = n_rows; a.n_heads = n_heads; a.n_kv_heads = n_kv_heads; a.head_dim = head_dim; a.n_splits = n_splits;
    a.out = d_out;
    CUDA_TRY(launch_attn_decode(dtype, a, (cudaStream_t)stream));
    return 0;
}

extern "C" int dllm_op_legacy_matmul(const float* d_in, const float* d_param, float* d_out, int32_t m, int32_t k,
                                     int32_t n, void* stream) {
    if (!d_in || !d_param || !d_out) return fail("dllm_op_legacy_matmul: null pointer");
    if (dllm_device_count() == 0) return fail("dllm_op_legacy_matmul: no CUDA device (this library has no CPU path)");
    CUDA_TRY(launch_legacy_matmul(d_in, d_param, d_out, m, k, n, (cudaStream_t)stream));
    return 0;
}
