// Parameters of the persistent decode kernel (decode_megakernel.cu) and of the weight packer.
#pragma once

#include <cuda_runtime.h>
#include <stdint.h>

namespace dllm {

constexpr int MK_THREADS = 384;
constexpr int MK_STAGE_K = 256;                    // k-elements of one stage
constexpr int MK_STAGE_BYTES = 8 * MK_STAGE_K * 2; // one stage = 8 weight rows x 256 k, bf16 = 4 KB
constexpr int MK_MAX_ROWS = 16;
constexpr int MK_MAX_SPLITS = 16;  // the merge loads the partials in batches of eight 16-byte vectors per thread
constexpr int MK_TRACE_SLOTS = 16;   // clock stamps per phase in the debug trace (DLLM_MK_TRACE)
constexpr int MK_MAX_PHASES = 512;  // 5 per layer + 2: up to 102 layers per stage
constexpr int MK_SMEM_BUDGET = 227 * 1024 - 6 * 1024;  // dynamic shared memory: the opt-in maximum less the kernel's static arrays
constexpr int MK_ROUND_PANELS = 13; // panels a CTA works on between two epilogues
constexpr int MK_PART_UNITS = MK_ROUND_PANELS + MK_THREADS / 32 - 1;  // partial sums of a round: one per (panel, warp) piece

enum MkKind { MK_EMBED = 0, MK_QKV, MK_ATTN, MK_OPROJ, MK_GATEUP, MK_FC_GELU, MK_DOWN, MK_HEAD };

// Weights are streamed from a packed copy owned by the shard handle:
//   panel  = 8 weight rows; for the paired kinds (QKV rotary pairs, gate/up) rows (2i, 2i+1) of a panel are
//            the two rows of pair i, so that one thread of the epilogue sees both
//   layout = [panel][k / 32][row 0..7][32 elements]: every 256-k slice of a panel is 4 KB contiguous (one TMA
//            bulk copy = one ring stage) and, inside a stage, the 16-byte chunks a warp reads per mma are
//            512 contiguous bytes (bank-conflict free).  It is also the no-swizzle UMMA core-matrix order.
struct MkPhase {
    const void* wp;      // packed weights of this phase (null for EMBED / ATTN)
    int32_t n_panels;
    int32_t K;
    int32_t n0;          // logical output rows (un-padded)
    int16_t kind;        // MkKind
    int16_t layer;
    int16_t n_stages;    // ceil(K / 256); K / 128 when xg
    int16_t xg;          // 1: the input rows are streamed from global memory with the weights (half-size stages:
                         //    2 KB of weights + 2 KB of rows in fragment layout) instead of being staged in shared
                         //    memory by the prologue -- used for the down projection when R x ffn does not fit
    float inv_ns;        // 1 / n_stages: the kernel divides stage indices by multiplying (mk_set_stages)
};
static_assert(sizeof(MkPhase) == 32, "MkPhase is copied to shared memory word by word");

inline void mk_set_stages(MkPhase& q, int n_stages) {
    q.n_stages = (int16_t)n_stages;
    q.inv_ns = 1.0f / (float)(n_stages > 0 ? n_stages : 1);
}

struct MkLayer {
    const void *norm1_w, *norm1_b, *bq, *bk, *bv, *bo, *norm2_w, *norm2_b, *b_up, *b_down;
    void *kpool, *vpool;
};

struct MkParams {
    const MkPhase* phases;   // device array, n_phases entries = one decode step
    const MkLayer* layers;   // device array
    int32_t n_phases, n_steps, n_rows;
    int32_t hidden, n_heads, n_kv_heads, head_dim, vocab, k_max;
    int32_t norm, rope, stages, n_splits, advance, tok_base, part_units;
    int32_t blocked;         // 1: a CTA owns a contiguous run of every phase's panels, 0: panel p belongs to CTA p % grid
    float eps;
    void* h;                 // [R, d] residual stream
    void* q;                 // [R, H*hd]
    void* act;               // [R, ffn]; with act_xlayout: [ffn / 32][8 rows][32] (the fragment order the down phase streams)
    int32_t act_xlayout;
    void* attn_out;          // [R, H*hd] merged attention output
    unsigned* attn_counters; // [R, KV] tickets, zero between phases
    float* part_o;           // [R, H, S, hd]
    float* part_ml;          // [R, H, S, 2]
    float* logits;           // [R, V] or null
    float* arg_val;          // [grid, R]
    int32_t* arg_idx;        // [grid, R]
    int32_t* next_token;     // [R]
    int32_t* ids;            // [R]
    int32_t* pos;            // [R]
    const int32_t* seq;      // [R]
    int32_t* tok_ring;
    const int32_t* block_table;
    int32_t pages_per_seq, page_shift;
    const void *wte, *wpe, *final_norm_w, *final_norm_b;
    const float *rope_cos, *rope_sin;
    unsigned *barrier, *done, *arg_counter;  // zero between launches
    // Single steps driven by a host loop: ids / positions / slots may be read in place from pinned host memory (no copy
    // nodes in front of the launch) and the sampled ids written there; the device copies below are then refreshed by the
    // kernel itself, for a decode loop that may follow.
    int32_t *mirror_ids, *mirror_pos, *mirror_seq;   // null = p.ids / p.pos / p.seq ARE the device copies
    int32_t* next_token_host;                        // null = none
    long long* trace;        // optional [n_steps][n_phases][4] SM-clock stamps of CTA 0 (debug, DLLM_MK_TRACE)
    // ---- fused hand-off (null in_flag = off).  The kernel runs n_steps = decode steps x n_micro micro-batch steps;
    // micro-batch m = step % n_micro uses rows [m*R, (m+1)*R) of pos / seq and slot m of the mailboxes.  A slot's
    // ticket counts the inputs delivered to it: the consumer of decode step G waits for ticket >= G + 1.
    int32_t n_micro;
    uint32_t ticket_base;    // decode step number of this launch's first step
    int* status;             // non-zero: a stage of this loop gave up waiting for a ticket (see wait_ticket)
    unsigned long long ticket_timeout_ns;
    int32_t flag_stride;     // uint32 elements between two slots' tickets
    int32_t out_phase;       // index of the phase whose epilogue produces the stage's output rows (its last o- or down projection)
    unsigned* in_flag;       // local inbox: tickets,
    const int32_t* in_tokens;//   token ids [n_micro][R] (first stage),
    void* in_hidden;         //   hidden rows [n_micro][R][d] (other stages; used in place as the residual stream)
    unsigned* out_flag;      // the next stage's inbox (peer memory; the first stage's inbox for the last stage)
    int32_t* out_tokens;
    void* out_hidden;
};

// Packed layout, panel-major: a panel's stages are consecutive, so any k-range of a panel (one warp unit) is
// one contiguous run -- a single L2 prefetch instruction covers it.  `stage_bytes` = 4096 (256 k) or 2048 (128 k).
__host__ __device__ inline size_t mk_stage_offset(int K, int panel, int st, uint32_t stage_bytes, uint32_t* bytes) {
    const size_t panel_bytes = (size_t)K * 16;
    const size_t off = (size_t)st * stage_bytes;
    const size_t left = panel_bytes - off;
    *bytes = (uint32_t)(left < (size_t)stage_bytes ? left : (size_t)stage_bytes);
    return (size_t)panel * panel_bytes + off;
}

// packer: kind selects the row order (QKV pairs, gate/up pairs, plain)
struct PackJob {
    const void* src[3];   // QKV: wq, wk, wv;  GATEUP: gate, up;  else: w
    void* dst;
    int32_t kind, n_panels, K, n0;
    int32_t n_heads, n_kv_heads, head_dim;
};
size_t mk_packed_bytes(int n_panels, int K);
cudaError_t launch_pack(const PackJob& job, cudaStream_t st);

bool mk_supported(int n_rows, int n_heads, int n_kv_heads, int head_dim);
size_t mk_smem_bytes(const MkParams& p, int stages);
int mk_max_splits(int n_rows, int n_heads, int head_dim);  // largest split count the combine scratch allows
cudaError_t launch_decode_megakernel(const MkParams& p, int num_sms, cudaStream_t st);

}  // namespace dllm
