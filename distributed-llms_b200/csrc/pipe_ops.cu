// Fused hand-off for decode steps that run as separate kernels (more rows than the persistent kernel takes at this
// width): the same mailboxes and tickets as decode_megakernel's, with the wait and the send as two small kernels in the
// stage's stream around the step's CUDA graph.  Nothing returns to the host between steps and no collective library is
// involved: the producer's send kernel stores the rows into the consumer's mailbox (peer memory over NVLink, plain
// 16-byte stores) and releases the slot's ticket; the consumer's wait kernel spins on the ticket in its own memory.
// Replaces, for these shapes, the per-step pickled hidden-state hop of the reference (src/worker/node.py:218-238,
// src/master/node.py:256-269).
#include "common.cuh"
#include "kernels.h"
#include "mk_device.cuh"

namespace dllm {
namespace {

using namespace mkd;

__global__ void __launch_bounds__(1024) pipe_begin_kernel(const PipeBeginParams p) {
    pdl_launch_dependents();   // the next kernel of the step may be placed as soon as every CTA of this one runs
    pdl_wait();                // ... and this one touches global memory only once its predecessor has completed
    bool ok = true;
    if (threadIdx.x == 0) ok = wait_ticket(p.in_flag, p.want, p.status, p.timeout_ns);
    if (__syncthreads_or(!ok)) return;   // gave up (or an earlier step of this stage did): the step runs on stale rows, nothing is sent on
    if ((int)threadIdx.x < p.n_rows) {
        const int r = threadIdx.x;
        p.pos[r] = __ldcg(p.pos_all + r) + p.pos_add;
        p.seq[r] = __ldcg(p.seq_all + r);
        if (p.in_tokens) {
            const int id = __ldcg(p.in_tokens + r);
            p.ids[r] = id;
            if (p.tok_ring) p.tok_ring[r] = id;   // the ids this stage fed, step-major
        }
    }
    if (p.in_hidden) {   // the previous stage's rows out of this stage's inbox
        const int vecs = p.n_rows * (p.hidden >> 3);
        const uint4* src = static_cast<const uint4*>(p.in_hidden);
        uint4* dst = static_cast<uint4*>(p.h);
        for (int i = threadIdx.x; i < vecs; i += blockDim.x) dst[i] = __ldcg(src + i);
    }
}

__global__ void __launch_bounds__(1024) pipe_end_kernel(const PipeEndParams p) {
    pdl_launch_dependents();   // the next kernel of the step may be placed as soon as every CTA of this one runs
    pdl_wait();                // ... and this one touches global memory only once its predecessor has completed
    if (*reinterpret_cast<const volatile int*>(p.status) != 0) return;   // the loop has been aborted: the next stage gives up in turn
    if (p.out_hidden) {   // rows into the next stage's inbox: peer memory
        const int vecs = p.n_rows * (p.hidden >> 3);
        const uint4* src = static_cast<const uint4*>(p.h);
        uint4* dst = static_cast<uint4*>(p.out_hidden);
        for (int i = threadIdx.x; i < vecs; i += blockDim.x) dst[i] = src[i];
    } else if ((int)threadIdx.x < p.n_rows) {   // the last stage: sampled ids into the first stage's inbox
        p.out_tokens[threadIdx.x] = p.next_tok[threadIdx.x];
    }
    __threadfence_system();
    __syncthreads();
    if (threadIdx.x == 0) st_release_sys(p.out_flag, p.ticket);
}

}  // namespace

cudaError_t launch_pipe_begin(const PipeBeginParams& p, cudaStream_t st) {
    if (p.n_rows < 1 || p.n_rows > 1024 || p.hidden % 8) return cudaErrorInvalidValue;
    const cudaError_t e = launch_pdl(pipe_begin_kernel, dim3(1), dim3(1024), 0, st, p);
    count_launch();
    return e;
}

cudaError_t launch_pipe_end(const PipeEndParams& p, cudaStream_t st) {
    if (p.n_rows < 1 || p.n_rows > 1024 || p.hidden % 8) return cudaErrorInvalidValue;
    const cudaError_t e = launch_pdl(pipe_end_kernel, dim3(1), dim3(1024), 0, st, p);
    count_launch();
    return e;
}

}  // namespace dllm
