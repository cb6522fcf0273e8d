// Split-K flash-decoding attention over the paged KV cache.
//
// grid = (n_splits, H, R): one CTA owns one query head of one row and one slice of that row's
// context.  Its four warps walk the slice token by token (lanes across head_dim, online softmax in
// fp32), merge through shared memory and write an un-normalised partial (o, m, l) to scratch.  The
// last CTA to finish a (row, head) -- found with an atomic ticket -- combines the splits and writes the
// normalised output, so the step needs one launch and the scratch counters end up zero again.
//
// Rounding follows HF eager attention (HF:llama/modeling_llama.py:212-219): the score is rounded to
// the activation dtype after the dot product and again after scaling; probabilities stay fp32 here
// (HF rounds them to bf16 before P@V -- documented difference, below bf16 output resolution).
#include <float.h>
#include <stdio.h>
#include <stdlib.h>

#include "../../include/dllm_b200.h"
#include "common.cuh"
#include "kernels.h"

namespace dllm {

constexpr int kAttnThreads = 128;
constexpr int kAttnWarps = kAttnThreads / 32;

template <typename T, int HD>
__global__ void __launch_bounds__(kAttnThreads) attn_decode_splitk_kernel(const AttnParams p) {
    pdl_launch_dependents();   // the next kernel of the step may be placed as soon as every CTA of this one runs
    pdl_wait();                // ... and this one touches global memory only once its predecessor has completed
    using D = DT<T>;
    constexpr int EPL = HD / 32;  // elements of head_dim per lane
    const int split = blockIdx.x, h = blockIdx.y, r = blockIdx.z;
    const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
    const int H = p.n_heads, KV = p.n_kv_heads, S = p.n_splits;
    const int kvh = h / (H / KV);
    const int len = p.positions[r] + 1;
    const int seq = p.seq_ids[r];
    const int chunk = (len + S - 1) / S;
    const int t0 = split * chunk;
    const int t1 = min(len, t0 + chunk);
    const float scale = rsqrtf((float)HD);
    const int page_mask = (1 << p.kv.page_shift) - 1;
    const int32_t* bt = p.kv.block_table + (size_t)seq * p.kv.pages_per_seq;
    const T* kpool = static_cast<const T*>(p.kv.kpool);
    const T* vpool = static_cast<const T*>(p.kv.vpool);

    float q[EPL];
    {
        const T* qp = static_cast<const T*>(p.q) + ((size_t)r * H + h) * HD + lane * EPL;
#pragma unroll
        for (int e = 0; e < EPL; ++e) q[e] = D::to_f(qp[e]);
    }

    float m = -FLT_MAX, l = 0.f, o[EPL];
#pragma unroll
    for (int e = 0; e < EPL; ++e) o[e] = 0.f;

    constexpr int UN = 4;  // tokens in flight per warp
    for (int tb = t0 + warp * UN; tb < t1; tb += kAttnWarps * UN) {
        float kf[UN][EPL], vf[UN][EPL];
#pragma unroll
        for (int u = 0; u < UN; ++u) {
            const int t = tb + u;
            if (t < t1) {
                const int page = bt[t >> p.kv.page_shift];
                const size_t base = ((((size_t)page * KV + kvh) << p.kv.page_shift) + (t & page_mask)) * HD + lane * EPL;
#pragma unroll
                for (int e = 0; e < EPL; ++e) {
                    kf[u][e] = D::to_f(kpool[base + e]);
                    vf[u][e] = D::to_f(vpool[base + e]);
                }
            } else {
#pragma unroll
                for (int e = 0; e < EPL; ++e) kf[u][e] = vf[u][e] = 0.f;
            }
        }
#pragma unroll
        for (int u = 0; u < UN; ++u) {
            if (tb + u >= t1) break;  // warp-uniform
            float d = 0.f;
#pragma unroll
            for (int e = 0; e < EPL; ++e) d = fmaf(q[e], kf[u][e], d);
            d = warp_sum(d);
            const float s = D::rnd(D::rnd(d) * scale);
            const float mn = fmaxf(m, s);
            const float corr = expf(m - mn);
            const float pe = expf(s - mn);
            l = l * corr + pe;
#pragma unroll
            for (int e = 0; e < EPL; ++e) o[e] = fmaf(pe, vf[u][e], o[e] * corr);
            m = mn;
        }
    }

    // merge the four warps
    __shared__ float sm_m[kAttnWarps], sm_l[kAttnWarps];
    __shared__ float sm_o[kAttnWarps][HD];
    __shared__ int s_last;
    if (lane == 0) {
        sm_m[warp] = m;
        sm_l[warp] = l;
    }
#pragma unroll
    for (int e = 0; e < EPL; ++e) sm_o[warp][lane * EPL + e] = o[e];
    __syncthreads();
    float M = -FLT_MAX;
#pragma unroll
    for (int w = 0; w < kAttnWarps; ++w) M = fmaxf(M, sm_m[w]);
    const size_t pidx = ((size_t)r * H + h) * S + split;
    if (threadIdx.x < HD) {
        float acc = 0.f;
#pragma unroll
        for (int w = 0; w < kAttnWarps; ++w) acc += sm_o[w][threadIdx.x] * expf(sm_m[w] - M);
        p.part_o[pidx * HD + threadIdx.x] = acc;
    }
    if (threadIdx.x == 0) {
        float L = 0.f;
#pragma unroll
        for (int w = 0; w < kAttnWarps; ++w) L += sm_l[w] * expf(sm_m[w] - M);
        p.part_ml[pidx * 2 + 0] = M;  // -FLT_MAX marks an empty split (L == 0)
        p.part_ml[pidx * 2 + 1] = L;
    }

    // last CTA of this (row, head) combines the splits
    __threadfence();
    __syncthreads();
    if (threadIdx.x == 0) {
        const unsigned ticket = atomicAdd(p.counters + (size_t)r * H + h, 1u);
        s_last = (ticket == (unsigned)S - 1);
    }
    __syncthreads();
    if (!s_last) return;
    __threadfence();
    const size_t pb = ((size_t)r * H + h) * S;
    float Mg = -FLT_MAX;
    for (int s = 0; s < S; ++s) Mg = fmaxf(Mg, __ldcg(p.part_ml + (pb + s) * 2));
    float Lg = 0.f;
    for (int s = 0; s < S; ++s)
        Lg += __ldcg(p.part_ml + (pb + s) * 2 + 1) * expf(__ldcg(p.part_ml + (pb + s) * 2) - Mg);
    if (threadIdx.x < HD) {
        float acc = 0.f;
        for (int s = 0; s < S; ++s)
            acc += __ldcg(p.part_o + (pb + s) * HD + threadIdx.x) * expf(__ldcg(p.part_ml + (pb + s) * 2) - Mg);
        static_cast<T*>(p.out)[((size_t)r * H + h) * HD + threadIdx.x] = D::from_f(acc / Lg);
    }
    if (threadIdx.x == 0) p.counters[(size_t)r * H + h] = 0u;
}

template <typename T>
static cudaError_t launch_t(const AttnParams& p, cudaStream_t s) {
    dim3 grid(p.n_splits, p.n_heads, p.n_rows);
    cudaError_t le = cudaSuccess;
    switch (p.head_dim) {
        case 32: le = launch_pdl(attn_decode_splitk_kernel<T, 32>, grid, dim3(kAttnThreads), 0, s, p); break;
        case 64: le = launch_pdl(attn_decode_splitk_kernel<T, 64>, grid, dim3(kAttnThreads), 0, s, p); break;
        case 128: le = launch_pdl(attn_decode_splitk_kernel<T, 128>, grid, dim3(kAttnThreads), 0, s, p); break;
        default: return cudaErrorInvalidValue;
    }
    count_launch();
    return le;
}

cudaError_t launch_attn_decode(int dtype, const AttnParams& p, cudaStream_t s) {
    if (p.n_rows <= 0 || p.n_splits <= 0 || p.n_heads % p.n_kv_heads != 0) return cudaErrorInvalidValue;
    static const bool gqa_off = [] { const char* e = getenv("DLLM_ATTN_GQA"); return e && e[0] == '0'; }();
    if (dtype == 0 && !gqa_off && attn_decode_gqa_supported(p)) return launch_attn_decode_gqa(p, s);
    return dtype == 0 ? launch_t<bf16>(p, s) : launch_t<float>(p, s);
}

}  // namespace dllm

// ---- split-K attention without any allocation -------------------------------------------------------------
//
// dllm_op_decode_attention() needs a scratch buffer from its caller.  Some callers have no way to get one:
// a host process that handed the whole GPU to its own memory pool (a foreign cudaMalloc fails there), or a
// stage thread that must never stall the other stages sharing the device (cudaMalloc synchronises it).
// For them the same kernel runs on a fixed scratch area that is part of the library image.
//
// The area is one file-scope __device__ array.  Each GPU that loads the library keeps its own copy of it until
// the process exits.  In host code the name of the array is only a stand-in -- it is not a device address --
// so the address is looked up once per device with cudaGetSymbolAddress(), while that device is current, and
// that pointer is what the kernel receives.
static __device__ __align__(256) unsigned char g_static_attn_scratch[DLLM_STATIC_ATTN_SCRATCH_BYTES];

namespace {
constexpr int kMaxDevices = 64;
// Not locked: two threads making the first call on a device at the same moment both look the address up and both
// store the same value (the symbol's address on a device never changes), so whichever store lands last is right.
void* g_static_scratch_ptr[kMaxDevices] = {nullptr};

size_t round_up_256(size_t v) { return (v + 255) / 256 * 256; }
}  // namespace

extern "C" int dllm_op_decode_attention_static(int32_t dtype, const void* d_q, const void* d_kpool, const void* d_vpool,
                                               const int32_t* d_block_table, int32_t pages_per_seq, int32_t page_size,
                                               const int32_t* d_positions, const int32_t* d_seq_ids, int32_t n_rows,
                                               int32_t n_heads, int32_t n_kv_heads, int32_t head_dim, int32_t n_splits,
                                               void* d_out, void* stream) {
    using dllm::set_last_error;
    if (dtype != DLLM_DTYPE_BF16 && dtype != DLLM_DTYPE_F32) return set_last_error("dllm_op_decode_attention_static: bad dtype");
    if (!d_q || !d_kpool || !d_vpool || !d_block_table || !d_positions || !d_seq_ids || !d_out)
        return set_last_error("dllm_op_decode_attention_static: null pointer");
    if (n_rows <= 0 || n_splits <= 0 || n_heads <= 0 || n_kv_heads <= 0 || n_heads % n_kv_heads)
        return set_last_error("dllm_op_decode_attention_static: bad shape");
    if (head_dim != 32 && head_dim != 64 && head_dim != 128)
        return set_last_error("dllm_op_decode_attention_static: head_dim must be 32, 64 or 128");
    if (page_size < 8 || (page_size & (page_size - 1)))
        return set_last_error("dllm_op_decode_attention_static: page_size must be a power of two >= 8");
    // same carve-up as dllm_op_decode_attention_scratch_bytes(): tickets | (m, l) pairs | partial outputs
    const size_t rh = (size_t)n_rows * n_heads;
    const size_t ticket_bytes = round_up_256(rh * 4);
    const size_t ml_bytes = round_up_256(rh * n_splits * 2 * 4);
    const size_t need = ticket_bytes + ml_bytes + round_up_256(rh * n_splits * head_dim * 4);
    if (need > (size_t)DLLM_STATIC_ATTN_SCRATCH_BYTES) {
        char msg[256];
        snprintf(msg, sizeof msg,
                 "dllm_op_decode_attention_static: this shape needs %zu bytes of scratch but the built-in area holds %d; "
                 "use fewer splits or rows, or call dllm_op_decode_attention with your own buffer",
                 need, (int)DLLM_STATIC_ATTN_SCRATCH_BYTES);
        return set_last_error(msg);
    }
    int dev = -1;
    cudaError_t e = cudaGetDevice(&dev);
    if (e != cudaSuccess || dev < 0 || dev >= kMaxDevices) {
        cudaGetLastError();
        return set_last_error("dllm_op_decode_attention_static: no usable CUDA device (this library has no CPU path)");
    }
    if (!g_static_scratch_ptr[dev]) {
        // first call on this device: resolve the device address of the array (see the note above)
        void* ptr = nullptr;
        e = cudaGetSymbolAddress((void**)&ptr, g_static_attn_scratch);
        if (e != cudaSuccess || !ptr) {
            cudaGetLastError();
            char msg[200];
            snprintf(msg, sizeof msg, "dllm_op_decode_attention_static: cudaGetSymbolAddress failed: %s", cudaGetErrorString(e));
            return set_last_error(msg);
        }
        g_static_scratch_ptr[dev] = ptr;
    }
    char* base = static_cast<char*>(g_static_scratch_ptr[dev]);
    cudaStream_t st = (cudaStream_t)stream;
    // The area still holds whatever the previous call left.  Only the tickets have to start at zero; the
    // partials are fully rewritten before they are read.
    e = cudaMemsetAsync(base, 0, ticket_bytes, st);
    if (e != cudaSuccess) {
        cudaGetLastError();
        return set_last_error("dllm_op_decode_attention_static: clearing the ticket counters failed");
    }
    dllm::AttnParams a{};
    a.counters = reinterpret_cast<uint32_t*>(base);
    a.part_ml = reinterpret_cast<float*>(base + ticket_bytes);
    a.part_o = reinterpret_cast<float*>(base + ticket_bytes + ml_bytes);
    a.q = d_q;
    a.kv.kpool = const_cast<void*>(d_kpool);
    a.kv.vpool = const_cast<void*>(d_vpool);
    a.kv.block_table = d_block_table;
    a.kv.pages_per_seq = pages_per_seq;
    a.kv.page_shift = 0;
    while ((1 << a.kv.page_shift) < page_size) ++a.kv.page_shift;
    a.positions = d_positions;
    a.seq_ids = d_seq_ids;
    a.n_rows = n_rows;
    a.n_heads = n_heads;
    a.n_kv_heads = n_kv_heads;
    a.head_dim = head_dim;
    a.n_splits = n_splits;
    a.out = d_out;
    e = dllm::launch_attn_decode(dtype, a, st);
    if (e != cudaSuccess) {
        char msg[200];
        snprintf(msg, sizeof msg, "dllm_op_decode_attention_static: launch failed: %s", cudaGetErrorString(e));
        return set_last_error(msg);
    }
    return 0;
}

