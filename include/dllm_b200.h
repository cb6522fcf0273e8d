/*
 * dllm_b200.h -- C ABI of the B200-native shard executor (libdllm_b200.so).
 *
 * This is the drop-in boundary for the hot path of MihirPanpatil/Distributed-LLMs: the per-worker
 * forward pass over a contiguous shard of transformer layers, and the stage-to-stage hand-off of
 * the hidden-state tile.  The reference has no FFI of its own (it is pure Python);
 * the seam these entry points sit behind is the Python class `ModelShard`
 * (reference src/worker/node.py:13-32) as used by `WorkerNode` (src/worker/node.py:195-196, 230,
 * 253, 282-283, 301).  Each function below cites the reference interface it replaces.
 *
 * Conventions
 *   - plain C types only: pointers, sizes, POD structs.  No torch / C++ types cross this boundary.
 *   - every function returns 0 on success, non-zero on failure; `dllm_last_error()` then holds a
 *     human-readable message (thread-local).  The reference prints and returns False / {} in the
 *     same places (src/worker/node.py:285-287); the Python mirror turns non-zero into exceptions.
 *   - `stream` is a `cudaStream_t` passed as `void*` (NULL = the legacy default stream).  All work is
 *     enqueued on it; nothing synchronises the device unless stated.
 *   - pointers named `d_*` must be device memory.  Pointers documented "host or device" go through
 *     `cudaMemcpyAsync(..., cudaMemcpyDefault)`; host buffers should be pinned to stay asynchronous.
 *   - a handle is not re-entrant: serialise calls on one `dllm_shard`.  Different handles are
 *     independent (reference: unlocked `self.shards`, src/worker/node.py:195,209,229).
 *   - there is no CPU fallback: without a CUDA device every compute entry point fails.
 */
#ifndef DLLM_B200_H
#define DLLM_B200_H

#include <stddef.h>
#include <stdint.h>

#ifdef __cplusplus
extern "C" {
#endif

#define DLLM_ABI_VERSION 2

#if defined(__GNUC__)
#define DLLM_API __attribute__((visibility("default")))
#else
#define DLLM_API
#endif

enum { DLLM_ARCH_GPT2 = 0, DLLM_ARCH_LLAMA = 1, DLLM_ARCH_QWEN2 = 2 };
enum { DLLM_DTYPE_BF16 = 0, DLLM_DTYPE_F32 = 1 };

/* Geometry of one stage (one `ModelShard`, reference src/worker/node.py:14-17). */
typedef struct dllm_shard_config {
    int32_t arch;         /* DLLM_ARCH_*                                                          */
    int32_t dtype;        /* DLLM_DTYPE_*: type of weights, activations, KV and the hand-off tile  */
    int32_t hidden;       /* d                                                                     */
    int32_t n_heads;      /* H                                                                     */
    int32_t n_kv_heads;   /* KV (== H for GPT-2)                                                   */
    int32_t head_dim;     /* 32, 64 or 128                                                         */
    int32_t ffn;          /* intermediate size                                                     */
    int32_t vocab;        /* V (used when has_embed / has_head)                                    */
    int32_t n_layers;     /* layers held by THIS shard (contiguous range)                          */
    int32_t has_embed;    /* shard 0: token (+ learned position) embedding                         */
    int32_t has_head;     /* last shard: final norm + lm_head (+ greedy argmax)                    */
    float   norm_eps;
    int32_t max_rows;     /* max rows (batch x tokens) per forward call                            */
    int32_t max_seqs;     /* sequence slots with resident KV                                       */
    int32_t max_context;  /* tokens per sequence the KV cache and the rope tables cover            */
    int32_t page_size;    /* tokens per KV page (power of two, >= 8)                               */
    int32_t num_pages;    /* KV pages per layer; 0 = max_seqs * ceil(max_context / page_size)      */
    /* A stage boundary may fall INSIDE a decoder layer, between the attention block (norm, q/k/v, attention, o-proj,
     * residual) and the MLP block (norm, gate/up, down, residual): the residual row is the hand-off either way, and
     * stages balance better -- the blocks are the units, and the last stage also streams the lm_head.  (The
     * reference's sharder assigns whole layers, src/model/shard_manager.py:33-61.)                                  */
    int32_t first_layer_mlp_only;   /* layers[0] holds only its MLP block (its attention block ran on the stage before) */
    int32_t last_layer_attn_only;   /* layers[n_layers-1] holds only its attention block (its MLP runs on the next stage) */
} dllm_shard_config;

/* Device pointers to one layer's weights (for a layer split between two stages: of the block this stage holds, the
 * other block's pointers NULL), all in `dtype`, row-major [out, in] (nn.Linear layout;
 * GPT-2 Conv1D [in,out] tensors are transposed by the host before they get here).  The memory is
 * owned by the caller (reference: `ModelShard.parameters`, src/worker/node.py:16,19-22) and must
 * outlive the handle.  NULL = absent (no bias / not this architecture). */
typedef struct dllm_layer_weights {
    const void* norm1_w;  const void* norm1_b;   /* [d]   input_layernorm / ln_1                   */
    const void* wq;       const void* bq;        /* [H*hd, d]                                      */
    const void* wk;       const void* bk;        /* [KV*hd, d]                                     */
    const void* wv;       const void* bv;        /* [KV*hd, d]                                     */
    const void* wo;       const void* bo;        /* [d, H*hd]                                      */
    const void* norm2_w;  const void* norm2_b;   /* [d]   post_attention_layernorm / ln_2          */
    const void* w_gate;                          /* [ffn, d]  (NULL for GPT-2)                     */
    const void* w_up;     const void* b_up;      /* [ffn, d]  up_proj / c_fc                       */
    const void* w_down;   const void* b_down;    /* [d, ffn]  down_proj / mlp.c_proj               */
} dllm_layer_weights;

/* Device pointers to the tensors at the two ends of the model. */
typedef struct dllm_io_weights {
    const void*  wte;          /* [V, d]        token embedding        (has_embed)                 */
    const void*  wpe;          /* [max_pos, d]  learned positions      (GPT-2, has_embed)          */
    const void*  final_norm_w; /* [d]                                  (has_head)                  */
    const void*  final_norm_b; /* [d]           (GPT-2)                                            */
    const void*  lm_head;      /* [V, d]                               (has_head)                  */
    const float* rope_cos;     /* [max_context, hd/2] fp32             (Llama / Qwen2)             */
    const float* rope_sin;     /* [max_context, hd/2] fp32                                         */
} dllm_io_weights;

typedef struct dllm_shard dllm_shard;

/* ---- library -------------------------------------------------------------------------------- */

/* Last error message of the calling thread ("" if none). */
DLLM_API const char* dllm_last_error(void);
/* DLLM_ABI_VERSION the library was built with. */
DLLM_API int dllm_abi_version(void);
/* Number of CUDA devices visible (0 = none; compute entry points will fail). */
DLLM_API int dllm_device_count(void);
/* Kernels this library launched since load (all handles, all threads); bench.py reports it. */
DLLM_API unsigned long long dllm_kernel_launch_count(void);

/* ---- shard life cycle ----------------------------------------------------------------------- */

/* Replaces `ModelShard.__init__` + `to_device` (reference src/worker/node.py:14-22) and the
 * LOAD_SHARD branch of `WorkerNode._handle_command` (src/worker/node.py:191-203): binds the
 * caller's device-resident weights and allocates what the reference never had -- the stage's paged
 * KV cache, its block table and the activation/scratch buffers -- on the current CUDA device. */
DLLM_API int dllm_shard_create(const dllm_shard_config* cfg, const dllm_layer_weights* layers,
                      const dllm_io_weights* io, dllm_shard** out);

/* Replaces `del self.shards[shard_id]` (reference src/worker/node.py:209,292). Frees KV + scratch. */
DLLM_API int dllm_shard_destroy(dllm_shard* shard);

/* Bytes of device memory the handle owns (KV pages + scratch). */
DLLM_API size_t dllm_shard_device_bytes(const dllm_shard* shard);

/* ---- KV-cache state (no counterpart in the reference: it has no KV cache) -------- */

/* Make sure sequence slot `seq_id` owns pages for positions [0, n_tokens).  Pages come from the
 * handle's free list; the block-table row is updated on `stream`.  Fails when the pool is empty. */
DLLM_API int dllm_shard_kv_reserve(dllm_shard* shard, int32_t seq_id, int32_t n_tokens, void* stream);
/* Return all pages of `seq_id` (or of every slot when seq_id < 0) to the free list. */
DLLM_API int dllm_shard_kv_release(dllm_shard* shard, int32_t seq_id);
/* Free pages left in the pool. */
DLLM_API int dllm_shard_kv_free_pages(const dllm_shard* shard);

/* ---- the hot path ---------------------------------------------------------------------------- */

/* Replaces `ModelShard.compute` (reference src/worker/node.py:24-32) as called from RUN_INFERENCE /
 * SCHEDULE_COMPUTATION / `schedule_computation` (src/worker/node.py:230, 253, 301).
 *
 * Runs this stage's layers over `n_rows` rows.  Row r is one token of sequence slot `seq_ids[r]` at
 * absolute position `positions[r]`; its K/V are appended to the slot's pages (reserved beforehand)
 * and it attends to positions [0, positions[r]] of that slot.  Decode = one row per sequence;
 * chunked prefill = consecutive positions of one sequence in one call.
 *
 *   input        has_embed: int32 token ids [n_rows]; else hidden states [n_rows, d] in `dtype`
 *   positions    int32 [n_rows]          seq_ids   int32 [n_rows]
 *   hidden_out   [n_rows, d] in `dtype`: the hand-off tile for the next stage (NULL if has_head)
 *   logit_rows   int32 [n_logit_rows]: rows whose logits are wanted (has_head; NULL = all rows, or -- with
 *                n_logit_rows < 0 -- none: the inner chunks of a chunked prefill skip the head)
 *   logits_out   fp32 [n_logit_rows, V] or NULL;  next_token_out int32 [n_logit_rows] or NULL
 *                (greedy: lowest index of the maximum, like torch.argmax)
 * All seven data pointers may be host or device memory. */
DLLM_API int dllm_shard_forward(dllm_shard* shard, const void* input, const int32_t* positions,
                       const int32_t* seq_ids, int32_t n_rows, void* hidden_out,
                       const int32_t* logit_rows, int32_t n_logit_rows, float* logits_out,
                       int32_t* next_token_out, void* stream);

/* Replaces the master's (absent) decode loop for a stage that owns both ends of the model
 * (`MasterNode._process_tasks`, reference src/master/node.py:227-277, runs one broadcast and no
 * loop -- SURVEY.md F2): `n_steps` greedy decode steps for the `n_rows` sequences whose state the
 * previous `dllm_shard_forward` call left behind, entirely on the device: each step feeds the
 * argmax of the last one back into the embedding and advances the positions by one, without a host
 * round trip.  `tokens_out` (host or device) receives int32 [n_steps, n_rows].  KV pages for the new
 * positions must have been reserved. */
DLLM_API int dllm_shard_decode_loop(dllm_shard* shard, int32_t n_rows, int32_t n_steps, int32_t* tokens_out,
                           void* stream);

/* ---- stage-to-stage hand-off ------------------------------------------------------------------ */

/* Fused hand-off for the steady-state decode loop.  Replaces, for the data plane inside one NVLink box,
 * the request/response round trip `MessageProtocol.send_message` / `receive_message` (reference
 * src/network/protocol.py:38-129) that `MasterNode._process_tasks` (reference src/master/node.py:227-277)
 * would drive once per token: every stage runs ONE persistent kernel for the whole loop; a stage's last
 * down-projection epilogue stores its output rows straight into the next stage's *mailbox* (peer memory over
 * NVLink) and publishes a ticket; the last stage stores the sampled token ids into the first stage's mailbox.
 * No launch, no collective library call and no host involvement per token.
 *
 * A mailbox is device memory of its owner: per slot (micro-batch) a ticket, `rows` int32 token ids and
 * `rows x hidden` bf16 hidden rows (`rows` <= 64 for the decode loop; any number for a prefill tile slot).  Slot
 * tickets count delivered inputs: the consumer of decode step G waits for ticket >= G + 1.  One process per GPU: the owner exports a 64-byte IPC handle, the producer imports it. */
typedef struct dllm_mailbox dllm_mailbox;
#define DLLM_MAILBOX_IPC_BYTES 64
DLLM_API int dllm_mailbox_create(int32_t device, int32_t n_slots, int32_t rows, int32_t hidden, dllm_mailbox** out);
/* `ipc_handle_out`: DLLM_MAILBOX_IPC_BYTES bytes (host) to hand to the producer's process. */
DLLM_API int dllm_mailbox_export(dllm_mailbox* box, void* ipc_handle_out);
/* The producer's view of another process's mailbox, mapped on `device` (the producer's GPU). */
DLLM_API int dllm_mailbox_import(int32_t device, const void* ipc_handle, int32_t n_slots, int32_t rows, int32_t hidden,
                        dllm_mailbox** out);
/* Same process, different GPU (tests, single-process multi-GPU hosts): a view of `box` usable from `device`. */
DLLM_API int dllm_mailbox_view(dllm_mailbox* box, int32_t device, dllm_mailbox** out);
/* Owner side: put `rows` token ids (host or device) into `slot` and set its ticket (seeds decode step
 * `ticket - 1`), stream-ordered.  */
DLLM_API int dllm_mailbox_post_tokens(dllm_mailbox* box, int32_t slot, const int32_t* tokens, uint32_t ticket, void* stream);
/* Owner side: copy out the token ids currently in `slot` (host or device destination) and, if `ticket_out`
 * (host) is given, the slot's ticket; synchronises `stream`. */
DLLM_API int dllm_mailbox_read_tokens(dllm_mailbox* box, int32_t slot, int32_t* tokens_out, uint32_t* ticket_out, void* stream);
/* Prefill hand-off through a mailbox slot (a slot may be created with as many rows as a prompt chunk has): replaces the
 * pickled `{name: tensor}` RESULT / RUN_INFERENCE payloads that carry the hidden states from worker to worker through
 * the master (reference src/worker/node.py:224-238, src/master/node.py:256-269).  The producer copies its
 * [n_rows, hidden] bf16 tile (device memory) into the NEXT stage's slot -- `box` is the producer's view of that mailbox
 * (dllm_mailbox_import / _view), so the copy is one NVLink DMA -- and the consumer copies it out of its own slot.
 * Stream-ordered on `stream`; the two sides are ordered by the caller (a control message sent after the producer's
 * stream has drained), not by tickets. */
DLLM_API int dllm_mailbox_put_rows(dllm_mailbox* box, int32_t slot, const void* d_src, int32_t n_rows, void* stream);
DLLM_API int dllm_mailbox_get_rows(dllm_mailbox* box, int32_t slot, void* d_dst, int32_t n_rows, void* stream);
DLLM_API int dllm_mailbox_destroy(dllm_mailbox* box);

/* `n_steps` decode steps of `n_micro` micro-batches of `rows` rows each, bf16.  Shapes the persistent kernel takes
 * (rows <= 16 and the staged rows fit shared memory: <= 8 rows at hidden 4096 and wider) run as ONE persistent launch
 * whose epilogue does the hand-off.  Other shapes up to 64 rows run every micro-batch step as [wait kernel] -> the
 * step's CUDA graph (swap-AB tcgen05 GEMMs, grouped-head attention) -> [send kernel], all enqueued by this call: the
 * same mailboxes and tickets, still no host round trip and no collective library between steps -- at 8 rows and above
 * the separate kernels reach the same fraction of the memory roofline as the persistent kernel.
 * `inbox` is this stage's own mailbox, `outbox` the next stage's
 * (the first stage's for the last stage; for a single stage pass its own inbox twice).  `positions`, `seq_ids`
 * (host or device, int32 [n_micro * rows]): position of the token fed at this call's first step and the KV slot of
 * every row; KV pages must be reserved up to position + n_steps.  `step_base`: decode steps done by earlier calls
 * (tickets keep counting).  `fed_tokens_out` (first stage only, host or device, int32 [n_steps][n_micro][rows],
 * may be null): the token ids this stage fed at every step -- the seed at step 0, afterwards what the last stage
 * generated one step earlier; the ids generated by the final step are in the inbox (dllm_mailbox_read_tokens)
 * once every stage has finished. */
DLLM_API int dllm_shard_decode_pipelined(dllm_shard* shard, dllm_mailbox* inbox, dllm_mailbox* outbox, int32_t n_micro,
                                int32_t rows, int32_t n_steps, uint32_t step_base, const int32_t* positions,
                                const int32_t* seq_ids, int32_t* fed_tokens_out, void* stream);

/* Pre-flight of dllm_shard_decode_pipelined: the same checks, no launch, nothing enqueued.  Belongs to the master's decode
 * loop that the reference lacks (src/master/node.py:227-277 broadcasts once): the master asks every stage first and starts
 * the stages' loops only when all of them accept -- a stage that refused after its neighbours had started would leave them
 * waiting for tickets that never arrive (until they give up: dllm_shard_pipeline_status). */
DLLM_API int dllm_shard_decode_pipelined_check(dllm_shard* shard, dllm_mailbox* inbox, dllm_mailbox* outbox, int32_t n_micro,
                                      int32_t rows, int32_t n_steps);

/* Outcome of the pipelined loops enqueued on `stream` so far: waits for the stream, returns 0 when every hand-off ticket
 * arrived.  A stage whose input ticket does not arrive within DLLM_TICKET_TIMEOUT_MS (environment, default 20000 -- the
 * stages' processes start seconds apart) abandons its loop in an orderly way: its kernels return, nothing is sent on (so
 * the stages after it give up in turn), the CUDA context stays usable, and this call reports the error once and clears
 * it.  The reference has no counterpart (its master waits on a socket with a timeout, src/master/node.py:188-203); this
 * is that timeout for the device-side hand-off. */
DLLM_API int dllm_shard_pipeline_status(dllm_shard* shard, void* stream);

/* Replaces `MessageProtocol.send_message` / `receive_message` (reference
 * src/network/protocol.py:38-129) for the data plane inside one NVLink box: copy the hand-off tile
 * (`bytes` of hidden states) from this device to `dst_ptr` on device `dst_device` over NVLink
 * (peer-to-peer DMA; peer access is enabled on first use).  Stream-ordered on `stream`. */
DLLM_API int dllm_handoff_peer_copy(void* dst_ptr, int dst_device, const void* src_ptr, int src_device,
                           size_t bytes, void* stream);

/* ---- single operators (exported so that each kernel can be checked against the oracle) --------- */

/* y[r, n] (+)= sum_k x[r, k] * w[n, k] (+ bias[n]); `accumulate` adds into y (the residual add).
 * x [n_rows, k], w [n_out, k], y [n_rows, n_out], all device memory in `dtype`.
 * This is the decode GEMV: 128-bit coalesced weight loads, warp-shuffle reductions. */
DLLM_API int dllm_op_gemv(int32_t dtype, const void* d_x, const void* d_w, const void* d_bias, void* d_y,
                 int32_t n_rows, int32_t n_out, int32_t k, int32_t accumulate, void* stream);

/* y[m, n] (+)= sum_k x[m, k] * w[n, k] (+ bias[n]) for many rows: the prefill GEMM (TMA-fed tcgen05 tiles, fp32
 * accumulation in tensor memory).  bf16 only; x [m, k], w [n, k], y [m, n] device memory; k a multiple of 8,
 * n even.  Same contract as dllm_op_gemv, which is the decode (few rows) form of the same product. */
DLLM_API int dllm_op_gemm(const void* d_x, const void* d_w, const void* d_bias, void* d_y, int32_t m, int32_t n_out,
                 int32_t k, int32_t accumulate, void* stream);

/* Split-K flash-decoding attention over a paged KV cache (GQA).
 *   d_q [n_rows, H, hd] `dtype`; d_kpool/d_vpool [num_pages, KV, page_size, hd] `dtype`;
 *   d_block_table int32 [max_seqs, pages_per_seq]; d_positions / d_seq_ids int32 [n_rows]
 *   (row r attends to [0, positions[r]] of slot seq_ids[r]); d_out [n_rows, H*hd] `dtype`.
 *   d_scratch: `dllm_op_decode_attention_scratch_bytes(...)` bytes of device memory, zeroed once
 *   before first use (the kernel leaves its counters zeroed again). */
DLLM_API size_t dllm_op_decode_attention_scratch_bytes(int32_t n_rows, int32_t n_heads, int32_t head_dim,
                                              int32_t n_splits);
DLLM_API int dllm_op_decode_attention(int32_t dtype, const void* d_q, const void* d_kpool, const void* d_vpool,
                             const int32_t* d_block_table, int32_t pages_per_seq, int32_t page_size,
                             const int32_t* d_positions, const int32_t* d_seq_ids, int32_t n_rows,
                             int32_t n_heads, int32_t n_kv_heads, int32_t head_dim, int32_t n_splits,
                             void* d_scratch, void* d_out, void* stream);

/* No counterpart in the reference.  The same split-K attention as dllm_op_decode_attention, for callers that
 * cannot hand over a scratch buffer because they cannot allocate device memory at this point: a process whose
 * own memory pool already owns the whole GPU (a cudaMalloc from outside the pool fails), or a stage thread that
 * may not stall the other stages on the device (cudaMalloc synchronises it).  The scratch is a fixed area of
 * DLLM_STATIC_ATTN_SCRATCH_BYTES inside the library image; every GPU that loads the library holds its own copy
 * until the process exits.  Nothing is allocated or freed by this call.
 *   - Shapes that need more than the area (tickets + (m,l) pairs + partial outputs, i.e. roughly
 *     n_rows * n_heads * n_splits * (head_dim + 2) * 4 bytes) are refused with an error that names both sizes.
 *   - The first call on a device looks up the device address of the area; make that first call outside any
 *     stream capture.
 *   - All calls share the one area and it keeps the last call's contents; the call clears the part that must
 *     start at zero on `stream`.  Two calls that overlap on the GPU (different streams or threads) race on it:
 *     serialise them.
 * Arguments are those of dllm_op_decode_attention without d_scratch; inputs and output stay the caller's. */
#define DLLM_STATIC_ATTN_SCRATCH_BYTES (4 * 1024 * 1024)
DLLM_API int dllm_op_decode_attention_static(int32_t dtype, const void* d_q, const void* d_kpool, const void* d_vpool,
                                    const int32_t* d_block_table, int32_t pages_per_seq, int32_t page_size,
                                    const int32_t* d_positions, const int32_t* d_seq_ids, int32_t n_rows,
                                    int32_t n_heads, int32_t n_kv_heads, int32_t head_dim, int32_t n_splits,
                                    void* d_out, void* stream);

/* The reference's shipped placeholder, kept so that its own tests run on the GPU: fp32
 * out[m, n] = sum_k in[m, k] * param[k, n] (`torch.matmul(inputs[key], parameters[key])`,
 * reference src/worker/node.py:31). Device pointers. */
DLLM_API int dllm_op_legacy_matmul(const float* d_in, const float* d_param, float* d_out, int32_t m, int32_t k,
                          int32_t n, void* stream);

#ifdef __cplusplus
}
#endif
#endif /* DLLM_B200_H */
