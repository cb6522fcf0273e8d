// Batched decode attention for grouped-query models on the tensor cores (bf16).
//
// grid = (n_splits, KV heads, rows): one CTA owns the Gq query heads that share one KV head, for one row and one
// slice of that row's context -- the K/V rows are read once per group, not once per query head.  The slice is walked
// in 64-token tiles gathered from the pages with cp.async (double-buffered); each of the four warps takes 16 tokens
// of a tile: S = Q.K^T (mma.sync m16n8k16, the 16 MMA rows are the group's heads, zero-padded) -> online softmax in
// fp32 -> O += P.V with P rounded to bf16, which is where HF eager attention rounds it.  The warps' (m, l, O) are merged
// through shared memory; with several splits the CTA writes an un-normalised partial and the last CTA of the
// (row, KV head) -- atomic ticket -- combines them, so the step needs one launch and the tickets end up zero again.
// Scratch layout is that of attention.cu: tickets [R, H] | (m, l) [R, H, S] | o [R, H, S, hd].
#include <float.h>

#include "common.cuh"
#include "kernels.h"
#include "ptx.cuh"

namespace dllm {

namespace {

using namespace ptx;

constexpr int kKT = 64, kThreads = 128, kWarps = 4, kRows = 16;
constexpr int kMergeUnroll = 8;   // splits whose partials the merging CTA requests in one go


template <int HD>
__global__ void __launch_bounds__(kThreads) attn_decode_gqa_kernel(const AttnParams p) {
    pdl_launch_dependents();   // the next kernel of the step may be placed as soon as every CTA of this one runs
    pdl_wait();                // ... and this one touches global memory only once its predecessor has completed
    constexpr int LD = HD + 8;            // padded rows: ldmatrix reads 8 rows x 16 bytes without bank conflicts
    constexpr int kTile = kKT * LD;       // elements of one K (or V) tile
    extern __shared__ __align__(16) unsigned char smem_dyn[];
    bf16* sQ = reinterpret_cast<bf16*>(smem_dyn);   // [16][LD]
    bf16* sKV = sQ + kRows * LD;                    // 2 buffers x {K tile, V tile}
    __shared__ float sm_m[kWarps][kRows], sm_l[kWarps][kRows];
    __shared__ int s_last;

    const int split = blockIdx.x, kvh = blockIdx.y, r = blockIdx.z;
    const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31, g = lane >> 2, t = lane & 3;
    const int H = p.n_heads, KV = p.n_kv_heads, S = p.n_splits, Gq = H / KV;
    const int len = p.positions[r] + 1;
    const int seq = p.seq_ids[r];
    int chunk = (len + S - 1) / S;
    chunk = (chunk + kKT - 1) / kKT * kKT;   // whole tiles per split
    const int t0 = min(len, split * chunk), t1 = min(len, t0 + chunk);
    const float scale = rsqrtf((float)HD);
    const int pmask = (1 << p.kv.page_shift) - 1;
    const int32_t* bt = p.kv.block_table + (size_t)seq * p.kv.pages_per_seq;
    const bf16* kpool = static_cast<const bf16*>(p.kv.kpool);
    const bf16* vpool = static_cast<const bf16*>(p.kv.vpool);

    // the group's query rows -> shared memory (MMA rows past Gq are zero)
    for (int i = threadIdx.x; i < kRows * (HD / 8); i += kThreads) {
        const int row = i / (HD / 8), c = i % (HD / 8);
        uint4 v = make_uint4(0u, 0u, 0u, 0u);
        if (row < Gq) v = ld_u4(static_cast<const bf16*>(p.q) + ((size_t)r * H + kvh * Gq + row) * HD + c * 8);
        *reinterpret_cast<uint4*>(sQ + row * LD + c * 8) = v;
    }

    auto gather_tile = [&](int kt0, int buf) {  // one commit group per tile; tokens past t1 are zero
        bf16* dK = sKV + (size_t)buf * 2 * kTile;
        bf16* dV = dK + kTile;
        for (int i = threadIdx.x; i < kKT * (HD / 8); i += kThreads) {
            const int row = i / (HD / 8), c = i % (HD / 8);
            const int tok = kt0 + row;
            if (tok < t1) {
                const int page = bt[tok >> p.kv.page_shift];
                const size_t base = ((((size_t)page * KV + kvh) << p.kv.page_shift) + (tok & pmask)) * HD + c * 8;
                cp_async16(dK + row * LD + c * 8, kpool + base);
                cp_async16(dV + row * LD + c * 8, vpool + base);
            } else {
                *reinterpret_cast<uint4*>(dK + row * LD + c * 8) = make_uint4(0u, 0u, 0u, 0u);
                *reinterpret_cast<uint4*>(dV + row * LD + c * 8) = make_uint4(0u, 0u, 0u, 0u);
            }
        }
        cp_async_commit();
    };

    float m_lo = -FLT_MAX, m_hi = -FLT_MAX, l_lo = 0.f, l_hi = 0.f;  // heads g and g + 8, this warp's tokens only
    float o[HD / 8][4];
#pragma unroll
    for (int i = 0; i < HD / 8; ++i) o[i][0] = o[i][1] = o[i][2] = o[i][3] = 0.f;

    if (t0 < t1) gather_tile(t0, 0);
    int it = 0;
    for (int kt0 = t0; kt0 < t1; kt0 += kKT, ++it) {
        const bf16* sK = sKV + (size_t)(it & 1) * 2 * kTile;
        const bf16* sV = sK + kTile;
        if (kt0 + kKT < t1) {
            gather_tile(kt0 + kKT, (it + 1) & 1);
            cp_async_wait_group<1>();
        } else {
            cp_async_wait_group<0>();
        }
        __syncthreads();  // tile `it` (and, the first time, Q) visible to every warp
        const int tok0 = kt0 + warp * 16;  // this warp's 16 tokens of the tile
        if (tok0 < t1) {
            // ---- S = Q K^T : 16 heads x 16 tokens ----
            float s[2][4] = {{0.f, 0.f, 0.f, 0.f}, {0.f, 0.f, 0.f, 0.f}};
#pragma unroll
            for (int ks = 0; ks < HD / 16; ++ks) {
                uint32_t a[4], bfr[4];
                ldsm_x4(a, smem_u32(sQ + ((lane & 7) + ((lane >> 3) & 1) * 8) * LD + ks * 16 + (lane >> 4) * 8));
                ldsm_x4(bfr, smem_u32(sK + (warp * 16 + (lane & 7) + (lane >> 4) * 8) * LD + ks * 16 + ((lane >> 3) & 1) * 8));
                mma_bf16(s[0], a, bfr[0], bfr[1]);
                mma_bf16(s[1], a, bfr[2], bfr[3]);
            }
            // ---- scale, mask the tokens past the end, online softmax ----
            float mx_lo = -FLT_MAX, mx_hi = -FLT_MAX;
#pragma unroll
            for (int nt = 0; nt < 2; ++nt) {
#pragma unroll
                for (int e = 0; e < 4; ++e) {
                    const int key = tok0 + nt * 8 + 2 * t + (e & 1);
                    float v = bf16_round(bf16_round(s[nt][e]) * scale);
                    if (key >= t1) v = -FLT_MAX;
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
            uint32_t pf[4];  // P as the A fragment of one 16-token k-step
#pragma unroll
            for (int nt = 0; nt < 2; ++nt) {
                const float p0 = s[nt][0] == -FLT_MAX ? 0.f : expf(s[nt][0] - mn_lo);
                const float p1 = s[nt][1] == -FLT_MAX ? 0.f : expf(s[nt][1] - mn_lo);
                const float p2 = s[nt][2] == -FLT_MAX ? 0.f : expf(s[nt][2] - mn_hi);
                const float p3 = s[nt][3] == -FLT_MAX ? 0.f : expf(s[nt][3] - mn_hi);
                sum_lo += p0 + p1;
                sum_hi += p2 + p3;
                pf[nt * 2 + 0] = pack_bf16(p0, p1);
                pf[nt * 2 + 1] = pack_bf16(p2, p3);
            }
            sum_lo += __shfl_xor_sync(0xffffffffu, sum_lo, 1);
            sum_lo += __shfl_xor_sync(0xffffffffu, sum_lo, 2);
            sum_hi += __shfl_xor_sync(0xffffffffu, sum_hi, 1);
            sum_hi += __shfl_xor_sync(0xffffffffu, sum_hi, 2);
            l_lo = l_lo * c_lo + sum_lo;
            l_hi = l_hi * c_hi + sum_hi;
            m_lo = mn_lo;
            m_hi = mn_hi;
            // ---- O = O * corr + P V : V[token][col] read transposed ----
#pragma unroll
            for (int nt = 0; nt < HD / 8; nt += 2) {
                uint32_t bfr[4];
                ldsm_x4_trans(bfr, smem_u32(sV + (warp * 16 + (lane & 7) + ((lane >> 3) & 1) * 8) * LD + nt * 8 + (lane >> 4) * 8));
                o[nt][0] *= c_lo; o[nt][1] *= c_lo; o[nt][2] *= c_hi; o[nt][3] *= c_hi;
                o[nt + 1][0] *= c_lo; o[nt + 1][1] *= c_lo; o[nt + 1][2] *= c_hi; o[nt + 1][3] *= c_hi;
                mma_bf16(o[nt], pf, bfr[0], bfr[1]);
                mma_bf16(o[nt + 1], pf, bfr[2], bfr[3]);
            }
        }
        __syncthreads();  // tile `it` consumed: its buffer is the target of the gather two tiles on
    }

    // ---- merge the four warps (each saw a quarter of every tile) through shared memory ----
    float* sm_o = reinterpret_cast<float*>(sKV);  // [warp][16][HD] fp32, the tile buffers are free now
    if (t == 0) {
        sm_m[warp][g] = m_lo; sm_m[warp][g + 8] = m_hi;
        sm_l[warp][g] = l_lo; sm_l[warp][g + 8] = l_hi;
    }
#pragma unroll
    for (int i = 0; i < HD / 8; ++i) {
        float* d = sm_o + ((size_t)warp * kRows + g) * HD + i * 8 + 2 * t;
        d[0] = o[i][0]; d[1] = o[i][1];
        d[8 * HD] = o[i][2]; d[8 * HD + 1] = o[i][3];
    }
    __syncthreads();
    bf16* out = static_cast<bf16*>(p.out);
    for (int i = threadIdx.x; i < Gq * HD; i += kThreads) {
        const int hq = i / HD, c = i - hq * HD;
        float M = -FLT_MAX;
#pragma unroll
        for (int w = 0; w < kWarps; ++w) M = fmaxf(M, sm_m[w][hq]);
        float L = 0.f, acc = 0.f;
#pragma unroll
        for (int w = 0; w < kWarps; ++w) {
            const float e = sm_m[w][hq] == -FLT_MAX ? 0.f : expf(sm_m[w][hq] - M);
            L = fmaf(sm_l[w][hq], e, L);
            acc = fmaf(sm_o[((size_t)w * kRows + hq) * HD + c], e, acc);
        }
        const int head = kvh * Gq + hq;
        if (S == 1) {
            out[((size_t)r * H + head) * HD + c] = __float2bfloat16_rn(L > 0.f ? acc / L : 0.f);
        } else {
            const size_t pidx = ((size_t)r * H + head) * S + split;
            p.part_o[pidx * HD + c] = acc;
            if (c == 0) {
                p.part_ml[pidx * 2 + 0] = M;  // -FLT_MAX marks an empty split (L == 0)
                p.part_ml[pidx * 2 + 1] = L;
            }
        }
    }
    if (S == 1) return;

    // ---- last CTA of this (row, KV head) combines the splits ----
    __threadfence();
    __syncthreads();
    if (threadIdx.x == 0) {
        const unsigned ticket = atomicAdd(p.counters + (size_t)r * H + kvh * Gq, 1u);
        s_last = (ticket == (unsigned)S - 1);
    }
    __syncthreads();
    if (!s_last) return;
    __threadfence();
    for (int i = threadIdx.x; i < Gq * HD; i += kThreads) {
        const int hq = i / HD, c = i - hq * HD;
        const size_t pb = ((size_t)r * H + kvh * Gq + hq) * S;
        float Mg = -FLT_MAX, Lg = 0.f, acc = 0.f;
        if (S <= kMergeUnroll) {
            // every (m, l, o) of this output element requested before the first is used: one L2 round trip, not 3 S
            float ms[kMergeUnroll], ls[kMergeUnroll], os[kMergeUnroll];
#pragma unroll
            for (int s2 = 0; s2 < kMergeUnroll; ++s2) {
                const bool on = s2 < S;
                const float2 ml = on ? __ldcg(reinterpret_cast<const float2*>(p.part_ml) + pb + s2) : make_float2(-FLT_MAX, 0.f);
                ms[s2] = ml.x;
                ls[s2] = ml.y;
                os[s2] = on ? __ldcg(p.part_o + (pb + s2) * HD + c) : 0.f;
            }
#pragma unroll
            for (int s2 = 0; s2 < kMergeUnroll; ++s2) Mg = fmaxf(Mg, ms[s2]);
#pragma unroll
            for (int s2 = 0; s2 < kMergeUnroll; ++s2) {   // same order and arithmetic as the loop below
                if (s2 < S) {
                    const float e = ms[s2] == -FLT_MAX ? 0.f : expf(ms[s2] - Mg);
                    Lg = fmaf(ls[s2], e, Lg);
                    acc = fmaf(os[s2], e, acc);
                }
            }
        } else {
            for (int s2 = 0; s2 < S; ++s2) Mg = fmaxf(Mg, __ldcg(p.part_ml + (pb + s2) * 2));
            for (int s2 = 0; s2 < S; ++s2) {
                const float ms = __ldcg(p.part_ml + (pb + s2) * 2);
                const float e = ms == -FLT_MAX ? 0.f : expf(ms - Mg);
                Lg = fmaf(__ldcg(p.part_ml + (pb + s2) * 2 + 1), e, Lg);
                acc = fmaf(__ldcg(p.part_o + (pb + s2) * HD + c), e, acc);
            }
        }
        out[((size_t)r * H + kvh * Gq + hq) * HD + c] = __float2bfloat16_rn(Lg > 0.f ? acc / Lg : 0.f);
    }
    if (threadIdx.x == 0) p.counters[(size_t)r * H + kvh * Gq] = 0u;
}

template <int HD>
cudaError_t launch_hd(const AttnParams& p, cudaStream_t st) {
    constexpr int LD = HD + 8;
    const size_t tiles = (size_t)(kRows + 4 * kKT) * LD * 2;
    const size_t merge = (size_t)kRows * LD * 2 + (size_t)kWarps * kRows * HD * 4;
    const size_t smem = tiles > merge ? tiles : merge;
    static PerDeviceOnce attr_once;
    if (attr_once.need()) {
        cudaError_t e = cudaFuncSetAttribute(attn_decode_gqa_kernel<HD>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem);
        if (e != cudaSuccess) return e;
    }
    dim3 grid(p.n_splits, p.n_kv_heads, p.n_rows);
    const cudaError_t le = launch_pdl(attn_decode_gqa_kernel<HD>, grid, dim3(kThreads), smem, st, p);
    count_launch();
    return le;
}

}  // namespace

bool attn_decode_gqa_supported(const AttnParams& p) {
    const int gq = p.n_kv_heads > 0 ? p.n_heads / p.n_kv_heads : 0;
    return gq >= 1 && gq <= kRows && (p.head_dim == 32 || p.head_dim == 64 || p.head_dim == 128);
}

cudaError_t launch_attn_decode_gqa(const AttnParams& p, cudaStream_t st) {
    switch (p.head_dim) {
        case 32: return launch_hd<32>(p, st);
        case 64: return launch_hd<64>(p, st);
        case 128: return launch_hd<128>(p, st);
    }
    return cudaErrorInvalidValue;
}

}  // namespace dllm
