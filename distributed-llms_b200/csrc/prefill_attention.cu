// Causal prefill attention over the paged KV cache, flash-attention style.
//
// One CTA = 64 consecutive query rows of one sequence x one query head; its 4 warps own 16 rows each.  The
// sequence's K/V (already appended, rotary applied) are walked in 64-token tiles gathered from the pages into
// shared memory with cp.async; S = Q.K^T and O += P.V run on the tensor cores (mma.sync m16n8k16, bf16 in, fp32
// accumulate), the softmax is the usual online one in fp32 registers.  Rounding follows HF eager attention: the
// score is rounded to bf16 after the dot product and after scaling, P is bf16 for the second product.
// (The dense projections around it are the tcgen05 GEMMs of prefill_gemm.cu; attention is ~2 % of the prefill
// flops at these lengths, so it uses the simpler warp-level tensor-core path for now.)
#include <float.h>
#include <stdlib.h>

#include "common.cuh"
#include "kernels.h"
#include "ptx.cuh"

namespace dllm {

namespace {

using namespace ptx;

constexpr int kKT = 64;  // keys per tile; the query tile QT is 64 rows (4 warps) or, for long prompts, 128 rows (8 warps)


template <int HD, int kQT>
__global__ void __launch_bounds__(kQT * 2) prefill_attn_kernel(const PrefillAttnParams p) {
    pdl_launch_dependents();   // the next kernel of the step may be placed as soon as every CTA of this one runs
    pdl_wait();                // ... and this one touches global memory only once its predecessor has completed
    constexpr int kThreads = kQT * 2;  // one warp per 16 query rows
    constexpr int LD = HD + 8;  // padded rows: ldmatrix reads 8 rows x 16 bytes without bank conflicts
    extern __shared__ __align__(16) unsigned char smem_dyn[];
    bf16* sQ = reinterpret_cast<bf16*>(smem_dyn);
    bf16* sKV = sQ + kQT * LD;  // two buffers of {K tile, V tile}: tile i+1 is gathered while tile i is multiplied
    const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31, g = lane >> 2, t = lane & 3;
    const int b = blockIdx.z, h = blockIdx.y, qt = blockIdx.x;
    const int H = p.n_heads, KV = p.n_kv_heads, kvh = h / (H / KV);
    const int n = p.rows_per_seq;
    const int row0 = b * n + qt * kQT;                 // first row of this tile in the call's row array
    const int rows_here = min(kQT, n - qt * kQT);
    const int seq = p.seq_ids[row0];
    const int pos0 = p.positions[row0];                // positions in the run are consecutive
    const int32_t* bt = p.kv.block_table + (size_t)seq * p.kv.pages_per_seq;
    const int pmask = (1 << p.kv.page_shift) - 1;
    const bf16* kpool = static_cast<const bf16*>(p.kv.kpool);
    const bf16* vpool = static_cast<const bf16*>(p.kv.vpool);
    const float scale = rsqrtf((float)HD);

    // Q tile -> shared memory (rows past the end of the run are zero)
    for (int i = threadIdx.x; i < kQT * (HD / 8); i += kThreads) {
        const int r = i / (HD / 8), c = i % (HD / 8);
        uint4 v = make_uint4(0u, 0u, 0u, 0u);
        if (r < rows_here) v = ld_u4(static_cast<const bf16*>(p.q) + ((size_t)(row0 + r) * H + h) * HD + c * 8);
        *reinterpret_cast<uint4*>(sQ + r * LD + c * 8) = v;
    }

    float m_lo = -FLT_MAX, m_hi = -FLT_MAX, l_lo = 0.f, l_hi = 0.f;  // rows g and g+8 of this warp's 16
    float o[HD / 8][4];
#pragma unroll
    for (int i = 0; i < HD / 8; ++i) o[i][0] = o[i][1] = o[i][2] = o[i][3] = 0.f;
    const int q_lo = pos0 + warp * 16 + g, q_hi = q_lo + 8;   // absolute positions of the two rows
    const int last_key = pos0 + rows_here - 1;                // keys beyond the tile's last row are never needed

    auto gather_tile = [&](int kt0, int buf) {  // one commit group per tile
        bf16* dK = sKV + (size_t)buf * 2 * kKT * LD;
        bf16* dV = dK + kKT * LD;
        for (int i = threadIdx.x; i < kKT * (HD / 8); i += kThreads) {
            const int r = i / (HD / 8), c = i % (HD / 8);
            const int tok = kt0 + r;
            if (tok <= last_key) {
                const int page = bt[tok >> p.kv.page_shift];
                const size_t base = ((((size_t)page * KV + kvh) << p.kv.page_shift) + (tok & pmask)) * HD + c * 8;
                cp_async16(dK + r * LD + c * 8, kpool + base);
                cp_async16(dV + r * LD + c * 8, vpool + base);
            } else {
                *reinterpret_cast<uint4*>(dK + r * LD + c * 8) = make_uint4(0u, 0u, 0u, 0u);
                *reinterpret_cast<uint4*>(dV + r * LD + c * 8) = make_uint4(0u, 0u, 0u, 0u);
            }
        }
        cp_async_commit();
    };
    gather_tile(0, 0);
    int it = 0;
    for (int kt0 = 0; kt0 <= last_key; kt0 += kKT, ++it) {
        const bf16* sK = sKV + (size_t)(it & 1) * 2 * kKT * LD;
        const bf16* sV = sK + kKT * LD;
        if (kt0 + kKT <= last_key) {
            gather_tile(kt0 + kKT, (it + 1) & 1);  // that buffer was released by the barrier that ended tile it - 1
            cp_async_wait_group<1>();              // tile it has landed, tile it + 1 may still be in flight
        } else {
            cp_async_wait_group<0>();
        }
        __syncthreads();  // tile it visible to every warp (also orders the Q stores before the first reads)
        if (kt0 > pos0 + warp * 16 + 15) {  // every key of this tile lies after this warp's rows (warp-uniform)
            __syncthreads();
            continue;
        }

        // ---- S = Q K^T : 16 rows x 64 keys per warp ----
        float s[kKT / 8][4];
#pragma unroll
        for (int i = 0; i < kKT / 8; ++i) s[i][0] = s[i][1] = s[i][2] = s[i][3] = 0.f;
#pragma unroll
        for (int ks = 0; ks < HD / 16; ++ks) {
            uint32_t a[4];
            ldsm_x4(a, smem_u32(sQ + (warp * 16 + (lane & 7) + ((lane >> 3) & 1) * 8) * LD + ks * 16 + (lane >> 4) * 8));
#pragma unroll
            for (int nt = 0; nt < kKT / 8; nt += 2) {
                uint32_t bfr[4];  // two key tiles of 8: {b0, b1} each
                ldsm_x4(bfr, smem_u32(sK + (nt * 8 + (lane & 7) + (lane >> 4) * 8) * LD + ks * 16 + ((lane >> 3) & 1) * 8));
                mma_bf16(s[nt], a, bfr[0], bfr[1]);
                mma_bf16(s[nt + 1], a, bfr[2], bfr[3]);
            }
        }
        // ---- scale, causal mask, online softmax ----
        float mx_lo = -FLT_MAX, mx_hi = -FLT_MAX;
#pragma unroll
        for (int nt = 0; nt < kKT / 8; ++nt) {
#pragma unroll
            for (int e = 0; e < 4; ++e) {
                const int key = kt0 + nt * 8 + 2 * t + (e & 1);
                const int qpos = (e < 2) ? q_lo : q_hi;
                float v = bf16_round(bf16_round(s[nt][e]) * scale);
                if (key > qpos) v = -FLT_MAX;
                s[nt][e] = v;
            }
            mx_lo = fmaxf(mx_lo, fmaxf(s[nt][0], s[nt][1]));
            mx_hi = fmaxf(mx_hi, fmaxf(s[nt][2], s[nt][3]));
        }
        mx_lo = fmaxf(mx_lo, __shfl_xor_sync(0xffffffffu, mx_lo, 1));
        mx_lo = fmaxf(mx_lo, __shfl_xor_sync(0xffffffffu, mx_lo, 2));
        mx_hi = fmaxf(mx_hi, __shfl_xor_sync(0xffffffffu, mx_hi, 1));
        mx_hi = fmaxf(mx_hi, __shfl_xor_sync(0xffffffffu, mx_hi, 2));
        const float mn_lo = fmaxf(m_lo, mx_lo), mn_hi = fmaxf(m_hi, mx_hi);
        const float c_lo = expf(m_lo - mn_lo), c_hi = expf(m_hi - mn_hi);
        float sum_lo = 0.f, sum_hi = 0.f;
        // P as A fragments (16 rows x 16 keys each), rounded to bf16 -- where HF eager attention rounds it
        uint32_t pfr[kKT / 16][4];
#pragma unroll
        for (int nt = 0; nt < kKT / 8; ++nt) {
            const float p0 = s[nt][0] == -FLT_MAX ? 0.f : expf(s[nt][0] - mn_lo);
            const float p1 = s[nt][1] == -FLT_MAX ? 0.f : expf(s[nt][1] - mn_lo);
            const float p2 = s[nt][2] == -FLT_MAX ? 0.f : expf(s[nt][2] - mn_hi);
            const float p3 = s[nt][3] == -FLT_MAX ? 0.f : expf(s[nt][3] - mn_hi);
            sum_lo += p0 + p1;
            sum_hi += p2 + p3;
            // C fragment of key tile nt -> A fragment: tiles 2j and 2j+1 are the two k-halves of A step j
            pfr[nt >> 1][(nt & 1) * 2 + 0] = pack_bf16(p0, p1);
            pfr[nt >> 1][(nt & 1) * 2 + 1] = pack_bf16(p2, p3);
        }
        sum_lo += __shfl_xor_sync(0xffffffffu, sum_lo, 1);
        sum_lo += __shfl_xor_sync(0xffffffffu, sum_lo, 2);
        sum_hi += __shfl_xor_sync(0xffffffffu, sum_hi, 1);
        sum_hi += __shfl_xor_sync(0xffffffffu, sum_hi, 2);
        l_lo = l_lo * c_lo + sum_lo;
        l_hi = l_hi * c_hi + sum_hi;
        m_lo = mn_lo;
        m_hi = mn_hi;
#pragma unroll
        for (int i = 0; i < HD / 8; ++i) {
            o[i][0] *= c_lo;
            o[i][1] *= c_lo;
            o[i][2] *= c_hi;
            o[i][3] *= c_hi;
        }
        // ---- O += P V : V[token][col] is read transposed (ldmatrix.trans) ----
#pragma unroll
        for (int kk = 0; kk < kKT / 16; ++kk) {
            // A fragment order is (row g, k lo), (row g+8, k lo), (row g, k hi), (row g+8, k hi)
            const uint32_t a[4] = {pfr[kk][0], pfr[kk][1], pfr[kk][2], pfr[kk][3]};
#pragma unroll
            for (int nt = 0; nt < HD / 8; nt += 2) {
                uint32_t bfr[4];
                ldsm_x4_trans(bfr, smem_u32(sV + (kk * 16 + (lane & 7) + ((lane >> 3) & 1) * 8) * LD + nt * 8 + (lane >> 4) * 8));
                mma_bf16(o[nt], a, bfr[0], bfr[1]);
                mma_bf16(o[nt + 1], a, bfr[2], bfr[3]);
            }
        }
        __syncthreads();  // tile it fully consumed: its buffer is the target of the gather two tiles on
    }

    // ---- normalise and store ----
    const float inv_lo = l_lo > 0.f ? 1.0f / l_lo : 0.f, inv_hi = l_hi > 0.f ? 1.0f / l_hi : 0.f;
    const int r_lo = warp * 16 + g, r_hi = r_lo + 8;
    bf16* out = static_cast<bf16*>(p.out);
#pragma unroll
    for (int i = 0; i < HD / 8; ++i) {
        const int col = h * HD + i * 8 + 2 * t;
        if (r_lo < rows_here)
            *reinterpret_cast<uint32_t*>(out + (size_t)(row0 + r_lo) * H * HD + col) = pack_bf16(o[i][0] * inv_lo, o[i][1] * inv_lo);
        if (r_hi < rows_here)
            *reinterpret_cast<uint32_t*>(out + (size_t)(row0 + r_hi) * H * HD + col) = pack_bf16(o[i][2] * inv_hi, o[i][3] * inv_hi);
    }
}

}  // namespace

template <int HD, int QT>
static cudaError_t launch_attn_t(const PrefillAttnParams& p, cudaStream_t st) {
    const size_t smem = (size_t)(QT + 4 * kKT) * (HD + 8) * 2;
    static PerDeviceOnce attr_once;
    if (attr_once.need()) {
        cudaError_t e = cudaFuncSetAttribute(prefill_attn_kernel<HD, QT>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem);
        if (e != cudaSuccess) return e;
    }
    dim3 grid((p.rows_per_seq + QT - 1) / QT, p.n_heads, p.n_seqs);
    const cudaError_t le = launch_pdl(prefill_attn_kernel<HD, QT>, grid, dim3(QT * 2), smem, st, p);
    count_launch();
    return le;
}

cudaError_t launch_prefill_attn(const PrefillAttnParams& p, cudaStream_t st) {
    if (p.n_seqs <= 0 || p.rows_per_seq <= 0 || p.n_heads % p.n_kv_heads) return cudaErrorInvalidValue;
    if (prefill_attn_tc_supported(p)) return launch_prefill_attn_tc(p, st);  // TMA + tcgen05 (64-token pages, hd 64 / 128)
    // 128-row query tiles (8 warps share each K/V tile) once there are enough of them to fill the GPU
    const char* env = getenv("DLLM_PREFILL_QT128");  // "0": never, "1": always (tests), unset: by grid size
    const bool big = env ? env[0] == '1' : (long)((p.rows_per_seq + 127) / 128) * p.n_heads * p.n_seqs >= 296;
    switch (p.head_dim) {
        case 32: return launch_attn_t<32, 64>(p, st);
        case 64: return big ? launch_attn_t<64, 128>(p, st) : launch_attn_t<64, 64>(p, st);
        case 128: return big ? launch_attn_t<128, 128>(p, st) : launch_attn_t<128, 64>(p, st);
    }
    return cudaErrorInvalidValue;
}

}  // namespace dllm
