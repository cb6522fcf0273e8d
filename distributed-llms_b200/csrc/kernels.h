// Internal launcher interface between the shard runtime / C ABI and the sm_100a kernels.
#pragma once

#include <cuda_runtime.h>
#include <stdint.h>

namespace dllm {

enum NormKind { NORM_NONE = 0, NORM_RMS = 1, NORM_LAYER = 2 };
enum EpiKind { EPI_RESIDUAL = 0, EPI_QKV = 1, EPI_GATEUP = 2, EPI_FC_GELU = 3, EPI_LOGITS = 4 };

// Where a (sequence slot, position) lives in the paged KV pools of one layer.
//   pool layout: [num_pages][KV][page_size][hd]; block_table: [max_seqs][pages_per_seq]
struct PagedKV {
    void* kpool;
    void* vpool;
    const int32_t* block_table;
    int32_t pages_per_seq;
    int32_t page_shift;  // log2(page_size)
};

// One fused decode GEMV launch: prologue (stage + normalise the rows), y = W x, epilogue.
struct GemvParams {
    // rows
    const void* x;           // [*, K]   input rows (residual stream / attention output / activation)
    const int32_t* row_map;  // optional gather: logical row r reads x[row_map[r]]
    int32_t n_rows;          // R
    int32_t K;
    // prologue
    int32_t norm;            // NormKind
    const void* norm_w;
    const void* norm_b;
    float eps;
    // weights: up to three [n_i, K] matrices handled by one launch
    const void* w[3];
    const void* b[3];
    int32_t n[3];            // rows of each matrix
    // epilogue
    int32_t epi;             // EpiKind
    void* y;                 // RESIDUAL: [R, n0] (in/out); QKV: q [R, H*hd]; GATEUP/FC: act [R, ffn]
    int32_t accumulate;      // RESIDUAL: y += ...
    // QKV
    int32_t n_heads, n_kv_heads, head_dim, rope;
    const float* rope_cos;
    const float* rope_sin;
    const int32_t* positions;
    const int32_t* seq_ids;
    PagedKV kv;
    // LOGITS
    float* logits;           // [R, n0] fp32 or null
    float* arg_val;          // [grid, R] scratch
    int32_t* arg_idx;        // [grid, R] scratch
    uint32_t* counter;       // 1 counter, zero before the launch, left zero
    int32_t* next_token;     // [R] or null
};

struct AttnParams {
    const void* q;           // [R, H, hd]
    PagedKV kv;
    const int32_t* positions;
    const int32_t* seq_ids;
    int32_t n_rows, n_heads, n_kv_heads, head_dim, n_splits;
    float* part_o;           // [R, H, S, hd]
    float* part_ml;          // [R, H, S, 2]
    uint32_t* counters;      // [R, H], zero before the launch, left zero
    void* out;               // [R, H*hd]
};

struct EmbedParams {
    const int32_t* ids;
    const int32_t* positions;
    const void* wte;
    const void* wpe;         // null unless GPT-2
    void* out;               // [R, d]
    int32_t n_rows, hidden, vocab;
};

// Prefill GEMM on tcgen05 (prefill_gemm.cu): C[M, N] = A[M, K] . W[N, K]^T, bf16 in, fp32 accumulate in TMEM.
enum GemmEpi { GEMM_EPI_BIAS = 0, GEMM_EPI_RESIDUAL = 1, GEMM_EPI_SWIGLU = 2, GEMM_EPI_GELU = 3,
               GEMM_EPI_LOGITS = 4,   // LOGITS: c is fp32 [M, ldc], values rounded to bf16 (M <= 64 only)
               GEMM_EPI_QKV = 5 };    // q | k | v segments: bias, rotary embedding, q to its buffer, k / v into the KV pages (M > 64 only)
struct GemmTcParams {
    const void* a;      // [M, K] activations, row-major (lda = K)
    const void* w;      // [N, K] weights (nn.Linear layout)
    const void* w2;     // SWIGLU: the up projection (w = gate)
    const void* bias;   // [N] or null
    void* c;            // [M, ldc] output; RESIDUAL adds into it in place
    int32_t M, N, K, ldc, epi;
    // BIAS epilogue with M <= 64 only: up to two more weight matrices whose rows continue the logical [N, K] matrix
    // (q | k | v in one launch); their outputs continue in the columns of c after N
    const void* w_more[2];
    const void* bias_more[2];
    int32_t n_more[2];
    // GEMM_EPI_QKV: the projection's epilogue is the whole K1 tail -- rotary on q and k, q rows to `q_out`, k / v rows
    // appended to the paged cache at (seq_ids[row], positions[row]); `c` is not written
    void* q_out;             // [M, H*hd]
    PagedKV kv;
    const int32_t* positions;
    const int32_t* seq_ids;
    const float* rope_cos;   // [max_context, hd/2] fp32, null-safe when rope == 0
    const float* rope_sin;
    int32_t n_heads, n_kv_heads, head_dim, rope;
    // Optional scratch for splitting the tiles of the last, partly filled wave along K (M > 64 path): fp32 partial tiles
    // + one ticket per split tile (zero between launches).  Null: every tile is computed whole.  One launch at a time
    // may use a given scratch area (the shard handle owns one and runs its launches on one stream).
    float* split_ws;
    unsigned* split_tickets;
    size_t split_ws_bytes;
};
constexpr int kGemmSplitTickets = 512;
struct RopeAppendParams {
    const void* qkv;    // [R, (H+2KV)*hd] projected rows (bias added)
    void* q;            // [R, H*hd]
    PagedKV kv;
    const int32_t* positions;
    const int32_t* seq_ids;
    const float* rope_cos;
    const float* rope_sin;
    int32_t n_rows, n_heads, n_kv_heads, head_dim, rope;
};
cudaError_t launch_gemm_tc(const GemmTcParams& p, cudaStream_t st);
// whether the q|k|v projection of `n_rows` rows can take GEMM_EPI_QKV (rotary + KV append in the epilogue)
bool gemm_qkv_fusable(int n_rows, int head_dim, const void* bq, const void* bk, const void* bv);
bool gemm_swapab_enabled();  // false under DLLM_GEMM_SWAPAB=0: M <= 64 then takes the 128-row tiles, LOGITS is unavailable
cudaError_t launch_norm_rows(const void* x, const void* w, const void* b, void* y, int rows, int K, int norm, float eps,
                             cudaStream_t st, const int32_t* row_map = nullptr);  // y[r] = norm(x[row_map ? row_map[r] : r])
// next_token[r] = lowest index of the maximum of logits[r, 0..n) (fp32 rows of stride n)
cudaError_t launch_argmax_rows(const float* logits, int rows, int n, int32_t* next_token, cudaStream_t st);
cudaError_t launch_rope_append(const RopeAppendParams& p, cudaStream_t st);

// Causal prefill attention over the paged cache (prefill_attention.cu): rows are `n_seqs` runs of `rows_per_seq`
// consecutive positions of one sequence each (row = b * rows_per_seq + i), K/V of the run already appended.
struct PrefillAttnParams {
    const void* q;           // [R, H, hd] bf16
    PagedKV kv;
    const int32_t* positions;
    const int32_t* seq_ids;
    void* out;               // [R, H*hd] bf16
    int32_t n_seqs, rows_per_seq, n_heads, n_kv_heads, head_dim;
    int32_t num_pages;       // pages in each pool (the TMA tensor maps of the tcgen05 kernel cover the whole pool); 0 = unknown
};
cudaError_t launch_prefill_attn(const PrefillAttnParams& p, cudaStream_t st);
// tcgen05 / TMA form (prefill_attention_tc.cu): 64-token pages, head_dim 64 or 128; launch_prefill_attn dispatches to it
bool prefill_attn_tc_supported(const PrefillAttnParams& p);
cudaError_t launch_prefill_attn_tc(const PrefillAttnParams& p, cudaStream_t st);

// Each returns cudaGetLastError() of its launch.  dtype: 0 = bf16, 1 = fp32.
cudaError_t gemv_init_device();  // once per device, outside stream capture
cudaError_t launch_gemv(int dtype, const GemvParams& p, int num_sms, cudaStream_t s);
int gemv_grid_size(int dtype, const GemvParams& p, int num_sms);  // grid launch_gemv will use (sizes arg scratch)
cudaError_t launch_attn_decode(int dtype, const AttnParams& p, cudaStream_t s);
// bf16, grouped heads on the tensor cores (attention_gqa.cu); launch_attn_decode dispatches to it
bool attn_decode_gqa_supported(const AttnParams& p);
cudaError_t launch_attn_decode_gqa(const AttnParams& p, cudaStream_t s);
cudaError_t launch_embed(int dtype, const EmbedParams& p, cudaStream_t s);
cudaError_t launch_decode_begin(int32_t* ids, int32_t* positions, int32_t* seq_ids, const int32_t* next_token,
                                const int32_t* logit_rows, int32_t n, int32_t* step_ctr, int32_t pos_add,
                                cudaStream_t s);
cudaError_t launch_advance(int32_t* ids, int32_t* positions, const int32_t* next_token, int32_t* tok_ring,
                           int32_t* step_ctr, int32_t n_rows, cudaStream_t s);
cudaError_t launch_legacy_matmul(const float* in, const float* param, float* out, int m, int k, int n,
                                 cudaStream_t s);

// Function attributes (opt-in shared memory) are per device: remember, per kernel, on which devices they are set.
struct PerDeviceOnce {
    bool done[64] = {false};
    bool need() {  // true the first time on the current device
        int dev = 0;
        cudaGetDevice(&dev);
        if (dev < 0 || dev >= 64) return true;
        if (done[dev]) return false;
        done[dev] = true;
        return true;
    }
};

int set_last_error(const char* msg);  // stores the message dllm_last_error() returns; always returns 1
// Fused hand-off around a per-op decode step (pipe_ops.cu): wait for the inbox slot's ticket and set the step's inputs
// up; afterwards store the step's rows (or sampled ids) into the next stage's mailbox and release its ticket
struct PipeBeginParams {
    const unsigned* in_flag;    // this slot's ticket (own memory)
    unsigned want;
    const int32_t* in_tokens;   // slot's ids (first stage) or null
    const void* in_hidden;      // slot's rows (other stages) or null
    const int32_t *pos_all, *seq_all;   // the micro-batch's positions at step 0 / KV slots
    int32_t pos_add;
    int32_t *ids, *pos, *seq;   // the handle's working arrays
    void* h;
    int32_t* tok_ring;          // where to record the fed ids, or null
    int32_t n_rows, hidden;
    int* status;                // the stage's abort word (see mk_device.cuh: wait_ticket)
    unsigned long long timeout_ns;
};
struct PipeEndParams {
    unsigned* out_flag;         // the slot's ticket in the next stage's mailbox
    unsigned ticket;
    int32_t* out_tokens;        // last stage: the first stage's inbox ids
    const int32_t* next_tok;
    void* out_hidden;           // other stages: the next stage's inbox rows (peer memory)
    const void* h;
    int32_t n_rows, hidden;
    const int* status;
};
cudaError_t launch_pipe_begin(const PipeBeginParams& p, cudaStream_t st);
cudaError_t launch_pipe_end(const PipeEndParams& p, cudaStream_t st);

void count_launch(unsigned long long n = 1);
unsigned long long launch_count();

}  // namespace dllm
