// Persistent decode kernel for sm_100a: one cooperative launch runs `n_steps` whole decode steps of a
// stage (embedding, every layer's phases, head, greedy argmax, token feedback) for 1-16 rows.
//
// Why: at batch 1 a layer of TinyLlama is 84 MB = 13 us of HBM time; with one kernel per fused op the
// step was launch/latency-bound at 30 % of the HBM roofline (profiles/r1a_summary.md).  Here
//   * one CTA per SM stays resident; phases are separated by a grid barrier (monotonic counter in L2);
//   * weights are read from a packed copy (megakernel.h): every warp owns a ring of 4 KB stages filled by
//     TMA bulk copies (cp.async.bulk + mbarrier complete_tx), one copy per stage.  A stage is refilled the
//     moment the warp has consumed it and a warp's sequence of stages is static, so the ring runs ahead
//     *across* phase barriers, layers and decode steps: HBM keeps streaming while the CTA waits;
//   * the dot products run on the tensor cores as a swap-AB skinny GEMM (mma.sync m16n8k16: M = batch rows,
//     N = the 8 weight rows of a panel), which removes the bf16->fp32 unpack / FMA instruction stream that
//     made the first version issue-bound; accumulation is fp32;
//   * a CTA's panels are processed in rounds; the stages of a round are cut into equal per-warp chunks, the pieces
//     of a panel are summed in a fixed order in shared memory (deterministic), then one thread per (row pair,
//     batch row) runs the fused epilogue;
//   * attention: one unit per (row, KV head, split); the group's query heads are the 16 rows of mma.sync, K/V
//     tiles arrive through cp.async one tile ahead (the first one before the barrier that precedes the phase),
//     the last unit of a (row, KV head) merges the splits;
//   * activations (a few KB) move between phases through L2 (ld.global.cg);
//   * optional fused hand-off: the last down projection also stores its rows into the next stage's mailbox
//     (peer memory) and tickets replace kernel boundaries between pipeline stages (megakernel.h).
// Rounding points are those of the per-op kernels in decode_gemv.cu / attention_gqa.cu.
#include <float.h>
#include <stdlib.h>
#include <limits.h>

#include "mk_device.cuh"

namespace dllm {

namespace {

using namespace mkd;

// ---- prologues: build the phase's input rows in shared memory (bf16, row stride ldx) ------------------------

// Rows of the residual stream / an activation -> shared memory, optionally normalised.  One row is shared by
// 256 / 128 / 64 / 32 threads (1 / 2 / <=4 / more rows) so that a single row costs one batched L2 round trip.

// (not inlined: it has five call sites, and the kernel's code size is what the instruction caches feel -- 15.7 k -> 12.2 k
// SASS instructions, +1 % tokens/s; TR: per-phase time stamps compiled in)
template <bool TR, int RK, bool LL>   // RK = 1: the launch has exactly one row (the row loops and row masks fold away);
                                       // LL: the model normalises with RMSNorm (no LayerNorm code)
__device__ __noinline__ void stage_rows(const MkParams& p, const void* src, int K, int norm, const void* nw, const void* nb,
                                        bf16* xs, int ldx, float* red2, unsigned& red_flip, long long* tr_in = nullptr) {
    long long* const tr = TR ? tr_in : nullptr;
    // `red2`: two buffers of kWarps floats used alternately by successive block reductions, so that one __syncthreads per
    // reduction is enough (a buffer is rewritten two reductions later, with the barrier of the one in between in between)
    const int R = RK == 1 ? 1 : p.n_rows;
    const int tshift = R == 1 ? 8 : (R == 2 ? 7 : (R <= 4 ? 6 : 5)), tpr = 1 << tshift;   // threads per row: 256 / 128 / 64 / 32
    const int rows_par = kThreads >> tshift, wpr = tpr >> 5;
    const int sub = threadIdx.x >> tshift, lt = threadIdx.x & (tpr - 1);
    const int warp = threadIdx.x >> 5;
    const int stride = tpr * 8;
    const uint4 z = make_uint4(0u, 0u, 0u, 0u);
    for (int r0 = 0; r0 < R; r0 += rows_par) {
        const int r = r0 + sub;
        const bool active = r < R && sub < rows_par;
        const bf16* xr = static_cast<const bf16*>(src) + (size_t)(active ? r : 0) * K;
        bf16* dst = xs + (size_t)(active ? r : 0) * ldx;
        float s1 = 0.f, s2 = 0.f;
        if (norm != NORM_NONE && K <= 4 * stride) {
            // fast path (every decode shape of the BASELINE models): the whole row is held in registers by its
            // threads; the activation and the norm weights are fetched together -- one L2 round trip in all
            const int kb = lt * 8, k1 = kb + stride, k2 = kb + 2 * stride, k3 = kb + 3 * stride;
            const bf16* wq = static_cast<const bf16*>(nw);
            const bf16* bq = static_cast<const bf16*>(nb);
            const bool a0 = active && kb < K, a1 = active && k1 < K, a2 = active && k2 < K, a3 = active && k3 < K;
            uint4 v0 = z, v1 = z, v2 = z, v3 = z, w0 = z, w1 = z, w2 = z, w3 = z, b0 = z, b1 = z, b2 = z, b3 = z;
            if (a0) { v0 = ldcg_u4(xr + kb); w0 = ld_u4(wq + kb); if ((!LL && norm == NORM_LAYER)) b0 = ld_u4(bq + kb); }
            if (a1) { v1 = ldcg_u4(xr + k1); w1 = ld_u4(wq + k1); if ((!LL && norm == NORM_LAYER)) b1 = ld_u4(bq + k1); }
            if (a2) { v2 = ldcg_u4(xr + k2); w2 = ld_u4(wq + k2); if ((!LL && norm == NORM_LAYER)) b2 = ld_u4(bq + k2); }
            if (a3) { v3 = ldcg_u4(xr + k3); w3 = ld_u4(wq + k3); if ((!LL && norm == NORM_LAYER)) b3 = ld_u4(bq + k3); }
            if (a0) accum_stats(v0, s1, s2);
            if (a1) accum_stats(v1, s1, s2);
            if (a2) accum_stats(v2, s1, s2);
            if (a3) accum_stats(v3, s1, s2);
            if (tr) tr[11] = clock64();  // row and norm weights have arrived
            auto group_sum = [&](float a) {
                a = warp_sum(a);
                if (wpr > 1) {
                    float* red = red2 + (red_flip & 1u) * kWarps;
                    ++red_flip;
                    if ((threadIdx.x & 31) == 0) red[warp] = a;
                    __syncthreads();
                    a = 0.f;
                    if (sub < rows_par)
                        for (int w = 0; w < wpr; ++w) a += red[sub * wpr + w];
                }
                return a;
            };
            float mean = 0.f, rstd;
            if (LL || norm == NORM_RMS) {
                rstd = 1.0f / sqrtf(group_sum(s2) / (float)K + p.eps);
            } else {
                mean = group_sum(s1) / (float)K;
                float sq = 0.f, f[8];
                if (a0) { unpack<bf16>(v0, f);
#pragma unroll
                    for (int j = 0; j < 8; ++j) sq = fmaf(f[j] - mean, f[j] - mean, sq); }
                if (a1) { unpack<bf16>(v1, f);
#pragma unroll
                    for (int j = 0; j < 8; ++j) sq = fmaf(f[j] - mean, f[j] - mean, sq); }
                if (a2) { unpack<bf16>(v2, f);
#pragma unroll
                    for (int j = 0; j < 8; ++j) sq = fmaf(f[j] - mean, f[j] - mean, sq); }
                if (a3) { unpack<bf16>(v3, f);
#pragma unroll
                    for (int j = 0; j < 8; ++j) sq = fmaf(f[j] - mean, f[j] - mean, sq); }
                rstd = 1.0f / sqrtf(group_sum(sq) / (float)K + p.eps);
            }
            if (tr) tr[12] = clock64();  // statistics reduced
            if (a0) *reinterpret_cast<uint4*>(dst + kb) = norm_vec(v0, w0, b0, norm, mean, rstd);
            if (a1) *reinterpret_cast<uint4*>(dst + k1) = norm_vec(v1, w1, b1, norm, mean, rstd);
            if (a2) *reinterpret_cast<uint4*>(dst + k2) = norm_vec(v2, w2, b2, norm, mean, rstd);
            if (a3) *reinterpret_cast<uint4*>(dst + k3) = norm_vec(v3, w3, b3, norm, mean, rstd);
            if (tr) tr[13] = clock64();  // normalised row stored
            continue;
        }
        for (int kb = lt * 8; kb < K; kb += 4 * stride) {
            // four loads in flight per thread, then the stores (explicit scalars: no local arrays)
            const int k1 = kb + stride, k2 = kb + 2 * stride, k3 = kb + 3 * stride;
            uint4 v0 = z, v1 = z, v2 = z, v3 = z;
            if (active) {
                v0 = ldcg_u4(xr + kb);
                if (k1 < K) v1 = ldcg_u4(xr + k1);
                if (k2 < K) v2 = ldcg_u4(xr + k2);
                if (k3 < K) v3 = ldcg_u4(xr + k3);
                *reinterpret_cast<uint4*>(dst + kb) = v0;
                accum_stats(v0, s1, s2);
                if (k1 < K) { *reinterpret_cast<uint4*>(dst + k1) = v1; accum_stats(v1, s1, s2); }
                if (k2 < K) { *reinterpret_cast<uint4*>(dst + k2) = v2; accum_stats(v2, s1, s2); }
                if (k3 < K) { *reinterpret_cast<uint4*>(dst + k3) = v3; accum_stats(v3, s1, s2); }
            }
        }
        if (norm == NORM_NONE) continue;
        // row statistics: warp shuffle, then across the row's warps through shared memory
        float* red = red2;  // (rows wider than four vectors per thread: the plain two-barrier protocol on the first buffer)
        float mean = 0.f, rstd;
        {
            float a = warp_sum(norm == NORM_RMS ? s2 : s1);
            if (wpr > 1) {
                __syncthreads();
                if ((threadIdx.x & 31) == 0) red[warp] = a;
                __syncthreads();
                a = 0.f;
                if (sub < rows_par)
                    for (int w = 0; w < wpr; ++w) a += red[sub * wpr + w];
            }
            if (norm == NORM_RMS) rstd = 1.0f / sqrtf(a / (float)K + p.eps);  // HF LlamaRMSNorm: fp32 variance
            else mean = a / (float)K;
        }
        if ((!LL && norm == NORM_LAYER)) {
            float sq = 0.f;
            if (active)
                for (int k = lt * 8; k < K; k += stride) {
                    float f[8];
                    unpack<bf16>(*reinterpret_cast<const uint4*>(dst + k), f);
#pragma unroll
                    for (int j = 0; j < 8; ++j) sq = fmaf(f[j] - mean, f[j] - mean, sq);
                }
            float a = warp_sum(sq);
            if (wpr > 1) {
                __syncthreads();
                if ((threadIdx.x & 31) == 0) red[warp] = a;
                __syncthreads();
                a = 0.f;
                if (sub < rows_par)
                    for (int w = 0; w < wpr; ++w) a += red[sub * wpr + w];
            }
            rstd = 1.0f / sqrtf(a / (float)K + p.eps);
        }
        const bf16* wp = static_cast<const bf16*>(nw);
        const bf16* bp = static_cast<const bf16*>(nb);
        for (int kb = lt * 8; kb < K; kb += 2 * stride) {
            const int k1 = kb + stride;
            if (active) {
                uint4 w0 = ld_u4(wp + kb), w1 = z, b0 = z, b1 = z;
                if (k1 < K) w1 = ld_u4(wp + k1);
                if ((!LL && norm == NORM_LAYER)) {
                    b0 = ld_u4(bp + kb);
                    if (k1 < K) b1 = ld_u4(bp + k1);
                }
                *reinterpret_cast<uint4*>(dst + kb) = norm_vec(*reinterpret_cast<const uint4*>(dst + kb), w0, b0, norm, mean, rstd);
                if (k1 < K)
                    *reinterpret_cast<uint4*>(dst + k1) = norm_vec(*reinterpret_cast<const uint4*>(dst + k1), w1, b1, norm, mean, rstd);
            }
        }
    }
}


__device__ __forceinline__ void attention_phase(const MkParams& p, const StepCache& sc, const int* sbt, const MkLayer& L, const AttnSmem& sm, bool prefetched,
                                                const UnitId& my_unit, const int hd,
                                                long long* tr = nullptr) {
    __shared__ int s_attn_last;
    const int H = p.n_heads, KV = p.n_kv_heads, S = p.n_splits, Gq = H / KV, ld = hd + 8;
    const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
    const float scale = rsqrtf((float)hd);
    const int n_units = p.n_rows * KV * S;
    bool first = true;
    for (int u = blockIdx.x; u < n_units; u += gridDim.x, first = false) {
        const AttnUnit a = attn_unit(p, sc, first ? my_unit : unit_id(p, u));
        if (a.t0 >= a.t1) continue;  // split past the end of the context: the merge never reads it
        // Tensor-core unit: the Gq query heads of the KV head are the (zero-padded) 16 rows of mma.sync.m16n8k16; of a
        // 64-token tile, warp w < 4 takes tokens 16 w .. 16 w + 15 and keeps its own online-softmax state (m, l) and
        // output fragment; the four warps are merged once, after the last tile.  P is rounded to bf16 where HF rounds it.
        const int g = lane >> 2, t4 = lane & 3;
        bf16* sQ = reinterpret_cast<bf16*>(sm.q);  // [16][ld]
        // (a group has at most kMaxGq = 8 query heads: MMA rows 8..15 are padding, their half of every fragment is never used)
        static_assert(kMaxGq <= 8, "the upper half of the MMA rows is assumed to be padding");
        float m_lo = -FLT_MAX, l_lo = 0.f;  // head g, this warp's tokens only
        float o[16][4];  // [hd / 8 column tiles][C fragment]; head_dim 32 / 64 use the first 4 / 8
#pragma unroll
        for (int i = 0; i < 16; ++i) o[i][0] = o[i][1] = o[i][2] = o[i][3] = 0.f;
        for (int tile0 = a.t0; tile0 < a.t1; tile0 += kTT) {
            const int tend = min(a.t1, tile0 + kTT), nt = tend - tile0;
            const bool has_next = tile0 + kTT < a.t1;
            if (tile0 == a.t0) {
                // first tile of the unit: its K rows, then its V rows (later tiles were requested while the tile
                // before them was being multiplied)
                int from = tile0;
                if (first && prefetched) from = max(tile0, min(tend, attn_safe_end(sc, a)));
                attn_load_rows(p, L, sm, sbt + a.r * p.pages_per_seq, a.kvh, tile0, from, tend, true, false, hd);
                attn_load_rows(p, L, sm, sbt + a.r * p.pages_per_seq, a.kvh, tile0, from, tend, false, true, hd);
                // q of the Gq heads -> 16 bf16 rows in shared memory (rows past Gq zero): one 16-byte L2 load per
                // thread, in flight together with the K/V copies above
                const int vpr = hd >> 3;  // 16-byte vectors per row
                for (int i = threadIdx.x; i < 16 * vpr; i += kThreads) {
                    const int row = i / vpr, c = i - row * vpr;
                    uint4 v = make_uint4(0u, 0u, 0u, 0u);
                    if (row < Gq) v = __ldcg(reinterpret_cast<const uint4*>(static_cast<const bf16*>(p.q) + ((size_t)a.r * H + a.kvh * Gq + row) * hd) + c);
                    *reinterpret_cast<uint4*>(sQ + row * ld + c * 8) = v;
                }
                if (tr) {
                    __threadfence_block();
                    tr[4] = clock64();  // q staged
                }
            }
            cp_async_wait_but_one();  // everything but the newest group (this tile's V rows): the K rows are in
            __syncthreads();
            if (tr) tr[5] = clock64();  // K tile in shared memory
            const bool my_tokens = warp < 4 && warp * 16 < nt;
            float s[2][4] = {{0.f, 0.f, 0.f, 0.f}, {0.f, 0.f, 0.f, 0.f}};
            if (my_tokens) {  // S = Q K^T : 16 heads x 16 tokens
                for (int ks = 0; ks < (hd >> 4); ++ks) {
                    uint32_t qa[4], kb[4];
                    ldsm_x4(qa, smem_u32(sQ + ((lane & 7) + ((lane >> 3) & 1) * 8) * ld + ks * 16 + (lane >> 4) * 8));
                    ldsm_x4(kb, smem_u32(sm.k + (warp * 16 + (lane & 7) + (lane >> 4) * 8) * ld + ks * 16 + ((lane >> 3) & 1) * 8));
                    mma_bf16(s[0], qa[0], qa[1], qa[2], qa[3], kb[0], kb[1]);
                    mma_bf16(s[1], qa[0], qa[1], qa[2], qa[3], kb[2], kb[3]);
                }
            }
            __syncthreads();  // every warp is done with the K tile
            if (tr) tr[6] = clock64();  // scores done
            if (has_next) {  // the K buffer is free: request the next tile's K rows, then wait for this tile's V rows
                attn_load_rows(p, L, sm, sbt + a.r * p.pages_per_seq, a.kvh, tile0 + kTT, tile0 + kTT, min(a.t1, tile0 + 2 * kTT), true, false, hd);
                cp_async_wait_but_one();
            } else {
                cp_async_wait_all();
            }
            __syncthreads();
            if (my_tokens) {
                // scale, mask the tokens past the end of the tile, online softmax over this warp's 16 tokens
                float mx_lo = -FLT_MAX;
#pragma unroll
                for (int n2 = 0; n2 < 2; ++n2) {
#pragma unroll
                    for (int e = 0; e < 2; ++e) {
                        const int key = warp * 16 + n2 * 8 + 2 * t4 + e;
                        float v = bf16_round(bf16_round(s[n2][e]) * scale);
                        if (key >= nt) v = -FLT_MAX;
                        s[n2][e] = v;
                    }
                    mx_lo = fmaxf(mx_lo, fmaxf(s[n2][0], s[n2][1]));
                }
                mx_lo = fmaxf(mx_lo, __shfl_xor_sync(0xffffffffu, mx_lo, 1));
                mx_lo = fmaxf(mx_lo, __shfl_xor_sync(0xffffffffu, mx_lo, 2));
                const float mn_lo = fmaxf(m_lo, mx_lo);
                const float c_lo = expf(m_lo - mn_lo);
                float sum_lo = 0.f;
                uint32_t pf[4];  // P as the A fragment of one 16-token k-step
#pragma unroll
                for (int n2 = 0; n2 < 2; ++n2) {
                    const float p0 = s[n2][0] == -FLT_MAX ? 0.f : expf(s[n2][0] - mn_lo);
                    const float p1 = s[n2][1] == -FLT_MAX ? 0.f : expf(s[n2][1] - mn_lo);
                    sum_lo += p0 + p1;
                    pf[n2 * 2 + 0] = pack_bf16(p0, p1);
                    pf[n2 * 2 + 1] = 0u;   // rows 8..15 of P
                }
                sum_lo += __shfl_xor_sync(0xffffffffu, sum_lo, 1);
                sum_lo += __shfl_xor_sync(0xffffffffu, sum_lo, 2);
                l_lo = l_lo * c_lo + sum_lo;
                m_lo = mn_lo;
                // O = O * corr + P V : V[token][col] is read transposed
#pragma unroll
                for (int ct = 0; ct < 16; ct += 2) {
                    if (ct * 8 < hd) {
                        uint32_t vb[4];
                        ldsm_x4_trans(vb, smem_u32(sm.v + (warp * 16 + (lane & 7) + ((lane >> 3) & 1) * 8) * ld + ct * 8 + (lane >> 4) * 8));
                        o[ct][0] *= c_lo; o[ct][1] *= c_lo;
                        o[ct + 1][0] *= c_lo; o[ct + 1][1] *= c_lo;
                        mma_bf16(o[ct], pf[0], pf[1], pf[2], pf[3], vb[0], vb[1]);
                        mma_bf16(o[ct + 1], pf[0], pf[1], pf[2], pf[3], vb[2], vb[3]);
                    }
                }
            }
            __syncthreads();  // the V buffer is free
            if (has_next)
                attn_load_rows(p, L, sm, sbt + a.r * p.pages_per_seq, a.kvh, tile0 + kTT, tile0 + kTT, min(a.t1, tile0 + 2 * kTT), false, true, hd);
            if (tr) tr[7] = clock64();  // softmax + P.V done
        }
        // merge the four token slices (fixed order) through shared memory: the K/V tile buffers are free now
        float* mo = reinterpret_cast<float*>(sm.k);  // [4 warps][16 heads][hd] fp32  (<= the two tile buffers)
        float* mm = sm.s;                            // [4][16] maxima, then [4][16] sums
        if (warp < 4) {
            if (t4 == 0) {
                mm[warp * 16 + g] = m_lo;
                mm[64 + warp * 16 + g] = l_lo;
            }
#pragma unroll
            for (int ct = 0; ct < 16; ++ct) {
                if (ct * 8 < hd) {
                    float* d = mo + ((size_t)warp * 16 + g) * hd + ct * 8 + 2 * t4;  // rows past Gq are padding
                    if (g < Gq) *reinterpret_cast<float2*>(d) = make_float2(o[ct][0], o[ct][1]);
                }
            }
        }
        __syncthreads();
        for (int i = threadIdx.x; i < (Gq * hd) >> 2; i += kThreads) {  // four columns of one head per thread
            const int hq = (i * 4) / hd, c = i * 4 - hq * hd;
            float M = -FLT_MAX;
#pragma unroll
            for (int w = 0; w < 4; ++w) M = fmaxf(M, mm[w * 16 + hq]);
            float Lsum = 0.f;
            float4 acc = make_float4(0.f, 0.f, 0.f, 0.f);
#pragma unroll
            for (int w = 0; w < 4; ++w) {
                const float mw = mm[w * 16 + hq];
                const float e = mw == -FLT_MAX ? 0.f : expf(mw - M);
                Lsum = fmaf(mm[64 + w * 16 + hq], e, Lsum);
                const float4 v = *reinterpret_cast<const float4*>(mo + ((size_t)w * 16 + hq) * hd + c);
                acc.x = fmaf(v.x, e, acc.x); acc.y = fmaf(v.y, e, acc.y); acc.z = fmaf(v.z, e, acc.z); acc.w = fmaf(v.w, e, acc.w);
            }
            const size_t pidx = ((size_t)a.r * H + a.kvh * Gq + hq) * S + a.split;
            *reinterpret_cast<float4*>(p.part_o + pidx * hd + c) = acc;
            if (c == 0) {
                p.part_ml[pidx * 2 + 0] = M;
                p.part_ml[pidx * 2 + 1] = Lsum;
            }
        }
        // The last unit of this (row, KV head) to finish -- atomic ticket -- merges the splits and writes the bf16
        // attention rows the o-projection stages: one CTA loads Gq x S partial rows once, instead of every CTA
        // loading all of them in the next prologue.
        const int n_act = attn_active_splits(p, sc, a.r);
        __threadfence();
        __syncthreads();
        if (threadIdx.x == 0) {
            unsigned* ctr = p.attn_counters + (size_t)a.r * KV + a.kvh;
            const bool last = atomicAdd(ctr, 1u) == (unsigned)n_act - 1;
            if (last) *ctr = 0u;  // every unit of the group has arrived: leave the ticket at zero for the next phase
            s_attn_last = last;
        }
        __syncthreads();
        if (s_attn_last) {
            __threadfence();
            float* wgt = sm.s;  // [Gq][S] merge weights (the score buffer is free)
            const size_t pb = ((size_t)a.r * H + a.kvh * Gq) * S;
            // one L2 round trip: every thread asks for its (m, l) pair and for the first eight partial rows of its
            // four output columns before anything is used
            const int mi = threadIdx.x;  // (head, split) pair of this thread, Gq * S <= 128 < kThreads
            float2 ml = make_float2(-FLT_MAX, 0.f);
            if (mi < Gq * S && mi % S < n_act) ml = __ldcg(reinterpret_cast<const float2*>(p.part_ml) + pb + mi);
            const int oi = threadIdx.x;  // four output columns of this thread, Gq * hd / 4 <= 256 < kThreads
            const bool has_o = oi < ((Gq * hd) >> 2);
            const int ohq = has_o ? (oi * 4) / hd : 0, oc = oi * 4 - ohq * hd;
            const float* po = p.part_o + (pb + (size_t)ohq * S) * hd + oc;
            float4 v[8];
#pragma unroll
            for (int j = 0; j < 8; ++j)
                v[j] = has_o && j < n_act ? ldcg_f4(po + (size_t)j * hd) : make_float4(0.f, 0.f, 0.f, 0.f);
            if (mi < Gq * S) {
                wgt[mi] = ml.x;
                sm.q[mi] = ml.y;  // [Gq][S] split sums (q is free as well)
            }
            __syncthreads();
            for (int hq = threadIdx.x; hq < Gq; hq += kThreads) {
                float Mg = -FLT_MAX, Lg = 0.f;
                for (int s2 = 0; s2 < n_act; ++s2) Mg = fmaxf(Mg, wgt[hq * S + s2]);
                for (int s2 = 0; s2 < n_act; ++s2) {
                    const float w = expf(wgt[hq * S + s2] - Mg);
                    Lg = fmaf(sm.q[hq * S + s2], w, Lg);
                    wgt[hq * S + s2] = w;
                }
                const float inv = 1.0f / Lg;
                for (int s2 = 0; s2 < n_act; ++s2) wgt[hq * S + s2] *= inv;
            }
            __syncthreads();
            if (has_o) {
                float4 acc = make_float4(0.f, 0.f, 0.f, 0.f);
                for (int s0 = 0; s0 < n_act; s0 += 8) {
                    if (s0 > 0) {  // more than eight splits: the next batch of rows
#pragma unroll
                        for (int j = 0; j < 8; ++j)
                            v[j] = s0 + j < n_act ? ldcg_f4(po + (size_t)(s0 + j) * hd) : make_float4(0.f, 0.f, 0.f, 0.f);
                    }
#pragma unroll
                    for (int j = 0; j < 8; ++j)
                        if (s0 + j < n_act) {
                            const float w = wgt[ohq * S + s0 + j];
                            acc.x = fmaf(v[j].x, w, acc.x);
                            acc.y = fmaf(v[j].y, w, acc.y);
                            acc.z = fmaf(v[j].z, w, acc.z);
                            acc.w = fmaf(v[j].w, w, acc.w);
                        }
                }
                __nv_bfloat162 lo = __floats2bfloat162_rn(acc.x, acc.y), hi = __floats2bfloat162_rn(acc.z, acc.w);
                uint2 out;
                out.x = *reinterpret_cast<uint32_t*>(&lo);
                out.y = *reinterpret_cast<uint32_t*>(&hi);
                *reinterpret_cast<uint2*>(static_cast<bf16*>(p.attn_out) + ((size_t)a.r * H + a.kvh * Gq + ohq) * hd + oc) = out;
            }
        }
        __syncthreads();  // shared buffers are reused by the next unit
    }
}

// ---- the kernel ----------------------------------------------------------------------------------------


template <bool HI, bool TR, int RK, bool LL, int HD, int PP, bool XG>  // XG: the phase table may hold a streamed-rows phase; PP: fused hand-off off (0) / on (1) / decided at run time (2); HD: head_dim as a compile-time constant (0: run-time); LL: Llama-style model (RMSNorm, rotary, gated MLP): the GPT-2 paths are compiled out; HI: more than 8 batch rows (second half of the mma M dimension in use); TR: time stamps
                                      // compiled in; RK = 1: one row (batch 1 -- R and RP are the constant 1), 0: any
__global__ void __launch_bounds__(kThreads, 1) decode_megakernel(const __grid_constant__ MkParams p) {
    extern __shared__ __align__(128) unsigned char smem[];
    const int ldx = p.k_max + 8;  // padded row stride of xs: rows land in different banks
    // carve-up (mirrors mk_smem_bytes)
    unsigned char* sp = smem;
    MkPhase* phases = reinterpret_cast<MkPhase*>(sp);
    sp += (size_t)((p.n_phases * sizeof(MkPhase) + 127) / 128) * 128;
    const uint32_t full_all = smem_u32(sp);
    sp += (size_t)((kWarps * p.stages * 8 + 127) / 128) * 128;
    // the staged input rows of the weight phases and the K/V tiles of the attention phase are never live together:
    // one region serves both
    bf16* xs = reinterpret_cast<bf16*>(sp);
    AttnSmem at;
    {
        unsigned char* ap = sp;
        at.k = reinterpret_cast<bf16*>(ap);
        ap += (size_t)kTT * (p.head_dim + 8) * 2;
        at.v = reinterpret_cast<bf16*>(ap);
        ap += (size_t)kTT * (p.head_dim + 8) * 2;
        at.s = reinterpret_cast<float*>(ap);
        ap += kMaxGq * kTT * 4;
        at.q = reinterpret_cast<float*>(ap);
    }
    sp += mk_rows_region_bytes(p);
    float* spart = reinterpret_cast<float*>(sp);
    sp += mk_part_bytes(p);
    float* srope = reinterpret_cast<float*>(sp);  // [R][head_dim]: cos row then sin row of each batch row's position
    sp += (size_t)p.n_rows * p.head_dim * 4;
    int* sbt = reinterpret_cast<int*>(sp);        // [R][pages_per_seq]: block-table row of each batch row's sequence
    sp += (size_t)((p.n_rows * p.pages_per_seq * 4 + 15) / 16) * 16;
    sp = reinterpret_cast<unsigned char*>(((uintptr_t)sp + 127) & ~(uintptr_t)127);
    const uint32_t ring_all = smem_u32(sp);

    __shared__ StepCache sc;
    __shared__ int s_units[MK_MAX_PHASES], s_first[MK_MAX_PHASES];
    __shared__ float red[2 * kWarps];
    unsigned red_flip = 0;
    __shared__ float s_val[kWarps][MK_MAX_ROWS];
    __shared__ int s_idx[kWarps][MK_MAX_ROWS];
    __shared__ int s_last;
    __shared__ int s_xkey[kWarps][8], s_xst[kWarps][8];
    __shared__ int s_wfirst[MK_ROUND_PANELS], s_wlast[MK_ROUND_PANELS];  // warps holding partial sums of a round's panels

    const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
    const int g = lane >> 2, t4 = lane & 3;  // mma fragment coordinates
    const int HDv = HD ? HD : p.head_dim;
    const int R = RK == 1 ? 1 : p.n_rows;
    int RP = 1;
    if (RK != 1)
        while (RP < R) RP <<= 1;
    for (int i = threadIdx.x; i < p.n_phases * (int)(sizeof(MkPhase) / 4); i += kThreads)
        reinterpret_cast<uint32_t*>(phases)[i] = reinterpret_cast<const uint32_t*>(p.phases)[i];
    for (int i = threadIdx.x; i < p.n_phases; i += kThreads) {
        const MkPhase& q = p.phases[i];
        s_units[i] = is_weight_phase(q.kind) ? local_panels(q) : 0;
        s_first[i] = panel_first(q, p.blocked);
    }
    if (threadIdx.x == 0) {
        for (int i = 0; i < kWarps * p.stages; ++i) mbar_init(full_all + i * 8, 1);
        asm volatile("fence.mbarrier_init.release.cluster;" ::: "memory");
    }
    __syncthreads();

    Ctx c;
    c.n_steps = p.n_steps;
    c.n_phases = p.n_phases;
    c.phases = phases;
    c.act_x = static_cast<const unsigned char*>(p.act);
    c.cur_key = -1;
    c.xkey = s_xkey[warp];
    c.xst = s_xst[warp];
    if (lane < 8) s_xkey[warp][lane] = -1;
    __syncwarp();
    c.units = s_units;
    c.first = s_first;
    c.warp = warp;
    c.lane = lane;
    c.ring = ring_all + warp * p.stages * kStageBytes;
    c.full = full_all + warp * p.stages * 8;

    // prime this warp's ring: the first `stages` entries of its static sequence
    const int pstride = p.blocked ? 1 : (int)gridDim.x;
    WIter prod = {0, 0, 0, 0, 0, 0, 0, nullptr, 0, 0, 0, 0, pstride};
    witer_seek(c, prod);
    for (int i = 0; i < p.stages && prod.step < p.n_steps; ++i) {
        issue_stage<XG>(c, prod, i);
        witer_next(c, prod);
    }
    unsigned ring_slot = 0, ring_parity = 0;  // the ring slot this warp consumes next and its mbarrier parity (kept incrementally: no divisions in the loop)
    unsigned bar_target = 0;
    bool need_barrier = false;
    int kv_prefetch_layer = -1;  // >= 0: prefetch this layer's old K/V rows for the attention phase at the next barrier
    // rows this lane feeds into the A fragment (rows >= R produce output rows nobody reads)
    const uint32_t xrow_lo = smem_u32(xs) + (uint32_t)(min(g, R - 1) * ldx * 2);
    const uint32_t xrow_hi = smem_u32(xs) + (uint32_t)(min(g + 8, R - 1) * ldx * 2);

    const bool piped = PP == 2 ? p.in_flag != nullptr : PP == 1;  // fused hand-off: inputs arrive in / outputs leave through mailboxes
    const UnitId my_unit = unit_id(p, blockIdx.x);   // this CTA's first attention unit
    bool ticket_ok = true;
    for (int step = 0; step < p.n_steps; ++step) {
        // micro-batch of this step and its decode step inside the launch (classic mode: m = 0, gstep = step)
        const int m = piped ? step % p.n_micro : 0;
        const int gstep = piped ? step / p.n_micro : step;
        const int32_t* pos_m = p.pos + m * R;
        const int32_t* seq_m = p.seq + m * R;
        const int pos_add = piped ? gstep : 0;  // piped: the position arrays stay as the host wrote them
        const int32_t* ids_m = piped ? p.in_tokens + m * R : p.ids;
        bf16* h_m = static_cast<bf16*>(piped && p.in_hidden ? p.in_hidden : p.h) + (piped ? (size_t)m * R * p.hidden : 0);
        for (int phi = 0; phi < p.n_phases; ++phi) {
            const MkPhase& ph = phases[phi];
            const MkLayer* L = p.layers + ph.layer;
            long long* const tr = (TR && p.trace && blockIdx.x == 0 && threadIdx.x == 0) ? p.trace + ((size_t)step * p.n_phases + phi) * MK_TRACE_SLOTS : nullptr;
            if (need_barrier) {
                grid_barrier_arrive(p.barrier, bar_target, tr);
                if (kv_prefetch_layer >= 0) {
                    // between this CTA's arrival and everybody's: ask for the old K/V rows of its first attention tile
                    // (rows no CTA writes in this step -- independent of the barrier)
                    const MkLayer& PL = p.layers[kv_prefetch_layer];
                    kv_prefetch_layer = -1;
                    const AttnUnit a = attn_unit(p, sc, my_unit);
                    const int tend = min(a.t1, a.t0 + kTT);
                    const int upto = max(a.t0, min(tend, attn_safe_end(sc, a)));
                    if (upto > a.t0) {
                        attn_load_rows(p, PL, at, sbt + a.r * p.pages_per_seq, a.kvh, a.t0, a.t0, upto, true, false, HDv);
                        attn_load_rows(p, PL, at, sbt + a.r * p.pages_per_seq, a.kvh, a.t0, a.t0, upto, false, true, HDv);
                    }
                }
                if (grid_barrier_wait(p.barrier, bar_target, tr, piped ? p.status : nullptr)) goto mk_aborted;
            }
            need_barrier = true;
            c.cur_key = step * p.n_phases + phi;
            if (XG && ph.xg) issue_deferred_rows(c, p.stages);
            if (phi == 0) {
                if (piped && threadIdx.x == 0) {
                    // every CTA is past the barrier: the previous step's rows are complete in the next stage's inbox
                    if (blockIdx.x == 0 && step > 0 && p.out_hidden) {
                        const int pm = (step - 1) % p.n_micro, pg = (step - 1) / p.n_micro;
                        st_release_sys(p.out_flag + (size_t)pm * p.flag_stride, p.ticket_base + pg + 1);
                    }
                    ticket_ok = wait_ticket(p.in_flag + (size_t)m * p.flag_stride, p.ticket_base + gstep + 1, p.status, p.ticket_timeout_ns);
                }
                if (piped && __syncthreads_or(!ticket_ok)) goto mk_aborted;   // this stage's input did not arrive: leave
                // positions / slots / KV page of every row, and the rotary rows of those positions
                if ((int)threadIdx.x < R) {
                    const int ps = ldcg_i32(pos_m + threadIdx.x) + pos_add, sq = ldcg_i32(seq_m + threadIdx.x);
                    sc.pos[threadIdx.x] = ps;
                    sc.seq[threadIdx.x] = sq;
                    sc.page[threadIdx.x] = p.block_table[(size_t)sq * p.pages_per_seq + (ps >> p.page_shift)];
                    sc.off[threadIdx.x] = ps & ((1 << p.page_shift) - 1);
                    step_cache_splits(sc, threadIdx.x, p.n_splits);
                    if (p.mirror_pos && blockIdx.x == 0) {  // inputs were read in place from host memory: keep the device copies current
                        p.mirror_pos[threadIdx.x] = ps;
                        p.mirror_seq[threadIdx.x] = sq;
                        if (p.mirror_ids) p.mirror_ids[threadIdx.x] = ldcg_i32(ids_m + threadIdx.x);
                    }
                }
                __syncthreads();  // positions and slots are in shared memory (they may live in host memory: read them once)
                for (int i = threadIdx.x; i < R * p.pages_per_seq; i += kThreads) {
                    const int r = i / p.pages_per_seq;
                    sbt[i] = p.block_table[(size_t)sc.seq[r] * p.pages_per_seq + (i - r * p.pages_per_seq)];
                }
                if (LL || p.rope) {
                    const int half = HDv >> 1;
                    for (int i = threadIdx.x; i < R * HDv; i += kThreads) {
                        const int r = i / HDv, e = i - r * HDv;
                        const int ps = sc.pos[r];
                        srope[i] = e < half ? p.rope_cos[(size_t)ps * half + e] : p.rope_sin[(size_t)ps * half + e - half];
                    }
                }
                __syncthreads();
            }
            if (tr) tr[0] = clock64();

            if (ph.kind == MK_EMBED) {
                // h[r] = wte[ids[r]] (+ wpe[pos[r]]), spread over the grid in 16-byte vectors
                const int vecs = p.hidden >> 3;
                for (int i = blockIdx.x * kThreads + threadIdx.x; i < R * vecs; i += gridDim.x * kThreads) {
                    const int r = i / vecs, v = i - r * vecs;
                    int id = ldcg_i32(ids_m + r);
                    if (piped && v == 0) p.tok_ring[(size_t)step * R + r] = id;  // the ids this stage fed, step-major
                    id = min(max(id, 0), p.vocab - 1);
                    uint4 val = ld_u4(static_cast<const bf16*>(p.wte) + (size_t)id * p.hidden + v * 8);
                    if (!LL && p.wpe) {
                        float a[8], b[8];
                        unpack<bf16>(val, a);
                        unpack<bf16>(ld_u4(static_cast<const bf16*>(p.wpe) + (size_t)sc.pos[r] * p.hidden + v * 8), b);
#pragma unroll
                        for (int j = 0; j < 8; ++j) a[j] += b[j];
                        val = pack8(a);
                    }
                    *reinterpret_cast<uint4*>(h_m + (size_t)r * p.hidden + v * 8) = val;
                }
                if (tr) tr[1] = tr[2] = tr[3] = clock64();
                continue;
            }
            if (ph.kind == MK_ATTN) {
                attention_phase(p, sc, sbt, *L, at, /*prefetched=*/true, my_unit, HDv, tr);
                if (tr) tr[1] = tr[2] = tr[3] = clock64();
                continue;
            }

            // ---- weight-streaming phase: prologue ----
            {   // one call site: which rows, which norm
                const void* rows = h_m;
                const void *nw = nullptr, *nb = nullptr;
                int nk = LL ? (int)NORM_RMS : p.norm;
                bool stage = true;
                if (ph.kind == MK_QKV) { nw = L->norm1_w; nb = L->norm1_b; }
                else if (ph.kind == MK_GATEUP || ph.kind == MK_FC_GELU) { nw = L->norm2_w; nb = L->norm2_b; }
                else if (ph.kind == MK_HEAD) { nw = p.final_norm_w; nb = p.final_norm_b; }
                else if (ph.kind == MK_DOWN) { rows = p.act; nk = NORM_NONE; stage = !(XG && ph.xg); }
                else { rows = p.attn_out; nk = NORM_NONE; }   // o-proj: merged attention rows
                if (stage) stage_rows<TR, RK, LL>(p, rows, ph.K, nk, nw, nb, xs, ldx, red, red_flip, tr);
            }
            __syncthreads();
            if (tr) tr[1] = clock64();

            // ---- main: this warp's (panel slot, k-split) units on the tensor cores ----
            const int K = ph.K, ns = ph.n_stages;
            const int n_local = s_units[phi];
            // the epilogue of the residual phases needs the old residual: fetch this thread's first pair now,
            // so that the L2 round trip overlaps the main loop
            uint32_t old_pref = 0u;
            if ((ph.kind == MK_OPROJ || ph.kind == MK_DOWN) && (int)threadIdx.x < n_local * 4 * RP) {
                const int r = threadIdx.x & (RP - 1), pi = threadIdx.x / RP;
                const int n = (s_first[phi] + pstride * (pi >> 2)) * 8 + 2 * (pi & 3);
                if (r < R && n < ph.n0)
                    old_pref = __ldcg(reinterpret_cast<const uint32_t*>(h_m + (size_t)r * ph.n0 + n));
            }
            const bool send_rows = piped && p.out_hidden && phi == p.out_phase;
            ArgBest best = {-FLT_MAX, INT_MAX};
            // panels are processed in rounds of MK_ROUND_PANELS so that the partial buffer stays small however many
            // panels a CTA owns (the lm_head: ~100 per CTA); inside a round every warp takes an equal chunk of stages
            for (int base = 0; base < n_local; base += MK_ROUND_PANELS) {
            const int n_round = min(MK_ROUND_PANELS, n_local - base);
            const int T = n_round * ns;
            if ((int)threadIdx.x < n_round) {  // which warps hold a piece of panel jr: chunk_begin(w) <= x < chunk_begin(w + 1)
                s_wfirst[threadIdx.x] = (kWarps * ((int)threadIdx.x * ns + 1) - 1) / T;
                s_wlast[threadIdx.x] = (kWarps * ((int)threadIdx.x * ns + ns) - 1) / T;
            }
            int x = chunk_begin(warp, T);
            const int x_end = chunk_begin(warp + 1, T);
            int jr = x < x_end ? div_stages(x, ph) : 0;
            int st = x - jr * ns;
            while (x < x_end) {
                const int seg_end = min(x_end, (jr + 1) * ns);
                // four independent accumulator sets: the mma dependency chain of a stage is 4 long, not 16
                float acc[4] = {0.f, 0.f, 0.f, 0.f}, acc1[4] = {0.f, 0.f, 0.f, 0.f};
                float acc2[4] = {0.f, 0.f, 0.f, 0.f}, acc3[4] = {0.f, 0.f, 0.f, 0.f};
                for (; x < seg_end; ++x, ++st) {
                    const unsigned slot = ring_slot;
                    long long tq0 = 0, tq1 = 0, tq2 = 0;
                    if (tr) tq0 = clock64();
                    mbar_wait(c.full + slot * 8, ring_parity);
                    if (tr) tq1 = clock64();
                    const uint32_t sbase = c.ring + slot * kStageBytes + (uint32_t)(lane * 16);  // [32-k block][row g][chunk t4]
                    const int stage_k = (XG && ph.xg) ? kStageK / 2 : kStageK;
                    const int k0 = st * stage_k;
                    const int nkb = min(stage_k, K - k0) >> 5;  // 32-k blocks in this stage
                    // input rows: from the shared-memory copy (row-major), or from the second half of the stage
                    // (fragment layout: [32-k block][row g][chunk t4], 512 bytes per block)
                    const uint32_t xstep = (XG && ph.xg) ? 512u : 64u;
                    const uint32_t xlo = (XG && ph.xg) ? c.ring + slot * kStageBytes + kStageBytes / 2 + (uint32_t)(lane * 16)
                                               : xrow_lo + (uint32_t)((k0 + t4 * 8) * 2);
                    const uint32_t xhi = xrow_hi + (uint32_t)((k0 + t4 * 8) * 2);
                    // lane (g, t4) holds 8 consecutive k of weight row g and of batch row g: two mma k-steps per block
#define DLLM_MK_BLOCK2(KB)                                                                          \
    {                                                                                               \
        const uint4 wa = lds_u4(sbase + (KB) * 512), wb = lds_u4(sbase + (KB) * 512 + 512);         \
        const uint4 xa = lds_u4(xlo + (KB) * xstep), xb = lds_u4(xlo + (KB) * xstep + xstep);       \
        uint4 ha = make_uint4(0u, 0u, 0u, 0u), hb = ha;                                             \
        if (HI) {                                                                                   \
            ha = lds_u4(xhi + (KB) * 64);                                                           \
            hb = lds_u4(xhi + (KB) * 64 + 64);                                                      \
        }                                                                                           \
        mma_bf16(acc, xa.x, ha.x, xa.y, ha.y, wa.x, wa.y);                                          \
        mma_bf16(acc1, xa.z, ha.z, xa.w, ha.w, wa.z, wa.w);                                         \
        mma_bf16(acc2, xb.x, hb.x, xb.y, hb.y, wb.x, wb.y);                                         \
        mma_bf16(acc3, xb.z, hb.z, xb.w, hb.w, wb.z, wb.w);                                         \
    }
                    if (nkb == 8) {
                        DLLM_MK_BLOCK2(0) DLLM_MK_BLOCK2(2) DLLM_MK_BLOCK2(4) DLLM_MK_BLOCK2(6)
                    } else {
                        int kb = 0;
                        for (; kb + 1 < nkb; kb += 2) DLLM_MK_BLOCK2(kb)
                        if (kb < nkb) {
                            const uint4 wa = lds_u4(sbase + kb * 512);
                            const uint4 xa = lds_u4(xlo + kb * xstep);
                            uint4 ha = make_uint4(0u, 0u, 0u, 0u);
                            if (HI) ha = lds_u4(xhi + kb * 64);
                            mma_bf16(acc, xa.x, ha.x, xa.y, ha.y, wa.x, wa.y);
                            mma_bf16(acc1, xa.z, ha.z, xa.w, ha.w, wa.z, wa.w);
                        }
                    }
                    __syncwarp();
                    if (tr) tq2 = clock64();
                    if (++ring_slot == (unsigned)p.stages) {
                        ring_slot = 0;
                        ring_parity ^= 1u;
                    }
                    if (prod.step < p.n_steps) {  // refill the slot with the next stage of this warp's sequence
                        issue_stage<XG>(c, prod, (int)slot);
                        witer_next(c, prod);
                    }
                    if (tr) {
                        const long long tq3 = clock64();
                        tr[4] += tq1 - tq0;  // waiting for the stage
                        tr[5] += tq2 - tq1;  // LDS + mma
                        tr[6] += tq3 - tq2;  // refill issue + cursor advance
                        tr[7] += 1;
                    }
                }
#pragma unroll
                for (int q = 0; q < 4; ++q) acc[q] = (acc[q] + acc1[q]) + (acc2[q] + acc3[q]);
                // C[batch g][panel rows 2*t4, 2*t4+1] (and batch g+8) -> partial buffer [jr + warp][panel row][batch]
                // (jr and warp both grow along the round's stage sequence, so jr + warp is unique per piece)
                float* dstp = spart + (size_t)(jr + warp) * (8 * RP);
                if (g < R) {
                    dstp[(2 * t4) * RP + g] = acc[0];
                    dstp[(2 * t4 + 1) * RP + g] = acc[1];
                }
                if (HI && g + 8 < R) {
                    dstp[(2 * t4) * RP + g + 8] = acc[2];
                    dstp[(2 * t4 + 1) * RP + g + 8] = acc[3];
                }
                ++jr;
                st = 0;
            }
            __syncthreads();
            if (tr) tr[2] = clock64();

            // ---- epilogue: one thread per (local panel, row pair, batch row); pieces summed in fixed order ----
            for (int o = threadIdx.x; o < n_round * 4 * RP; o += kThreads) {
                const int r = o & (RP - 1);
                const int pi = o / RP, pj = pi >> 2, i = pi & 3;
                if (r >= R) continue;
                float v0 = 0.f, v1 = 0.f;
                const int w0 = s_wfirst[pj], w1 = s_wlast[pj];
                if (T >= kWarps) {
                    // every warp holds stages: pieces w0..w1 (at most kWarps) all exist.  All loads first, then the sum in
                    // warp order -- the order is what makes the result reproducible, the loads need not wait for each other.
                    float pa[kWarps], pb[kWarps];
#pragma unroll
                    for (int j = 0; j < kWarps; ++j) {
                        const float* src = spart + (size_t)(pj + min(w0 + j, w1)) * (8 * RP);
                        pa[j] = src[(2 * i) * RP + r];
                        pb[j] = src[(2 * i + 1) * RP + r];
                    }
#pragma unroll
                    for (int j = 0; j < kWarps; ++j)
                        if (w0 + j <= w1) {
                            v0 += pa[j];
                            v1 += pb[j];
                        }
                } else {
                    for (int w = w0; w <= w1; ++w) {
                        if (chunk_begin(w, T) == chunk_begin(w + 1, T)) continue;  // a warp without stages
                        const float* src = spart + (size_t)(pj + w) * (8 * RP);
                        v0 += src[(2 * i) * RP + r];
                        v1 += src[(2 * i + 1) * RP + r];
                    }
                }
                const int panel = s_first[phi] + pstride * (base + pj);
                if (tr && o == 0 && base == 0) tr[14] = clock64();  // thread 0's pieces summed
                if (ph.kind == MK_QKV) {
                    const int hd = HDv, half = hd >> 1, H = p.n_heads, KV = p.n_kv_heads;
                    const int ppb = half >> 2;  // panels per head
                    const int head = panel / ppb, jj = (panel - head * ppb) * 4 + i;
                    const int m = head < H ? 0 : (head < H + KV ? 1 : 2);
                    const int hl = head - (m == 0 ? 0 : (m == 1 ? H : H + KV));
                    const int row0 = hl * hd + jj;
                    const void* bias = m == 0 ? L->bq : (m == 1 ? L->bk : L->bv);
                    v0 = bf16_round(v0 + (bias ? ldf(bias, row0) : 0.f));
                    v1 = bf16_round(v1 + (bias ? ldf(bias, row0 + half) : 0.f));
                    if ((LL || p.rope) && m != 2) {
                        const float cs = bf16_round(srope[r * hd + jj]);
                        const float sn = bf16_round(srope[r * hd + half + jj]);
                        const float o0 = bf16_round(bf16_round(v0 * cs) + bf16_round(-v1 * sn));
                        const float o1 = bf16_round(bf16_round(v1 * cs) + bf16_round(v0 * sn));
                        v0 = o0;
                        v1 = o1;
                    }
                    if (m == 0) {
                        bf16* q = static_cast<bf16*>(p.q) + (size_t)r * H * hd + row0;
                        q[0] = __float2bfloat16_rn(v0);
                        q[half] = __float2bfloat16_rn(v1);
                    } else {
                        const int page = sc.page[r], off = sc.off[r];
                        bf16* pool = static_cast<bf16*>(m == 1 ? L->kpool : L->vpool);
                        bf16* dst = pool + ((((size_t)page * KV + hl) << p.page_shift) + off) * hd + jj;
                        dst[0] = __float2bfloat16_rn(v0);
                        dst[half] = __float2bfloat16_rn(v1);
                    }
                } else if (ph.kind == MK_GATEUP) {
                    const int n = panel * 4 + i;
                    const float gg = bf16_round(v0), uu = bf16_round(v1);
                    const size_t ai = p.act_xlayout ? ((size_t)(n >> 5) * 8 + r) * 32 + (n & 31) : (size_t)r * ph.n0 + n;
                    static_cast<bf16*>(p.act)[ai] = __float2bfloat16_rn(bf16_round(silu_f(gg)) * uu);
                    if (p.act_xlayout) asm volatile("fence.proxy.async.global;" ::: "memory");  // read back by bulk copies
                } else {
                    const int n = panel * 8 + 2 * i;  // rows n, n+1
                    if (n >= ph.n0) continue;
                    const bool valid1 = n + 1 < ph.n0;
                    if (!LL && ph.kind == MK_FC_GELU) {
                        bf16* y = static_cast<bf16*>(p.act) +
                                  (p.act_xlayout ? ((size_t)(n >> 5) * 8 + r) * 32 + (n & 31) : (size_t)r * ph.n0 + n);
                        y[0] = __float2bfloat16_rn(gelu_new_f(bf16_round(v0 + (L->b_up ? ldf(L->b_up, n) : 0.f))));
                        if (valid1) y[1] = __float2bfloat16_rn(gelu_new_f(bf16_round(v1 + (L->b_up ? ldf(L->b_up, n + 1) : 0.f))));
                        if (p.act_xlayout) asm volatile("fence.proxy.async.global;" ::: "memory");
                    } else if (ph.kind == MK_OPROJ || ph.kind == MK_DOWN) {
                        const void* bias = ph.kind == MK_OPROJ ? L->bo : L->b_down;
                        bf16* y = h_m + (size_t)r * ph.n0 + n;
                        const uint32_t old = (o == (int)threadIdx.x && base == 0) ? old_pref : __ldcg(reinterpret_cast<const uint32_t*>(y));  // two bf16 of the residual
                        v0 = bf16_round(v0 + (bias ? ldf(bias, n) : 0.f)) + __uint_as_float(old << 16);
                        v1 = bf16_round(v1 + (bias ? ldf(bias, n + 1) : 0.f)) + __uint_as_float(old & 0xffff0000u);
                        *reinterpret_cast<__nv_bfloat162*>(y) = __floats2bfloat162_rn(v0, v1);
                        if (send_rows)  // the stage's output row goes straight into the next stage's inbox (peer store)
                            *reinterpret_cast<__nv_bfloat162*>(static_cast<bf16*>(p.out_hidden) + ((size_t)m * R + r) * ph.n0 + n) =
                                __floats2bfloat162_rn(v0, v1);
                    } else {  // MK_HEAD: logits in the model dtype, argmax = first index of the maximum
                        v0 = bf16_round(v0);
                        v1 = bf16_round(v1);
                        if (p.logits) {
                            float* lg = p.logits + (size_t)r * ph.n0 + n;
                            lg[0] = v0;
                            if (valid1) lg[1] = v1;
                        }
                        arg_update(best, v0, n);
                        if (valid1) arg_update(best, v1, n + 1);
                    }
                }
            }

            if (tr && base == 0) tr[15] = clock64();  // thread 0 through its outputs of the first round
            __syncthreads();  // the partial buffer is reused by the next round
            }  // rounds
            if (send_rows) __threadfence_system();  // peer stores performed before this CTA arrives at the next barrier

            if (ph.kind == MK_QKV && phi + 1 < p.n_phases && phases[phi + 1].kind == MK_ATTN &&
                (int)blockIdx.x < R * p.n_kv_heads * p.n_splits)
                kv_prefetch_layer = ph.layer;  // issued right after this CTA has arrived at the barrier (see the loop head)

            if (tr) tr[3] = clock64();
            if (ph.kind == MK_HEAD) {
                // thread -> batch row (o & (RP-1)) is fixed: reduce over the lanes that share it, then warps, then CTAs
                for (int off = RP; off < 32; off <<= 1) {
                    const float ov = __shfl_xor_sync(0xffffffffu, best.val, off);
                    const int oi = __shfl_xor_sync(0xffffffffu, best.idx, off);
                    arg_update(best, ov, oi);
                }
                if (lane < RP && lane < MK_MAX_ROWS) {
                    s_val[warp][lane] = best.val;
                    s_idx[warp][lane] = best.idx;
                }
                __syncthreads();
                if ((int)threadIdx.x < R) {
                    ArgBest b = {-FLT_MAX, INT_MAX};
                    for (int w = 0; w < kWarps; ++w) arg_update(b, s_val[w][threadIdx.x], s_idx[w][threadIdx.x]);
                    p.arg_val[(size_t)blockIdx.x * R + threadIdx.x] = b.val;
                    p.arg_idx[(size_t)blockIdx.x * R + threadIdx.x] = b.idx;
                }
                __threadfence();
                __syncthreads();
                if (threadIdx.x == 0) s_last = (atomicAdd(p.arg_counter, 1u) == gridDim.x - 1);
                __syncthreads();
                if (s_last) {
                    __threadfence();
                    for (int r = warp; r < R; r += kWarps) {
                        ArgBest b = {-FLT_MAX, INT_MAX};
                        for (int gi = lane; gi < (int)gridDim.x; gi += 32)
                            arg_update(b, __ldcg(p.arg_val + (size_t)gi * R + r), __ldcg(p.arg_idx + (size_t)gi * R + r));
#pragma unroll
                        for (int off = 16; off > 0; off >>= 1) {
                            const float ov = __shfl_xor_sync(0xffffffffu, b.val, off);
                            const int oi = __shfl_xor_sync(0xffffffffu, b.idx, off);
                            arg_update(b, ov, oi);
                        }
                        if (lane == 0) {
                            p.next_token[r] = b.idx;
                            if (p.next_token_host) p.next_token_host[r] = b.idx;  // pinned host memory: the caller reads it after the stream
                            if (piped) p.out_tokens[(size_t)m * R + r] = b.idx;  // into the first stage's inbox
                            if (p.advance) {
                                p.ids[r] = b.idx;
                                p.pos[r] = ldcg_i32(p.pos + r) + 1;
                                p.tok_ring[(size_t)(p.tok_base + step) * R + r] = b.idx;
                            }
                        }
                    }
                    if (threadIdx.x == 0) *p.arg_counter = 0u;
                    __threadfence();
                    if (piped) {  // s_last is uniform in the CTA: the ids are out, publish the slot's next ticket
                        __threadfence_system();
                        __syncthreads();
                        if (threadIdx.x == 0) st_release_sys(p.out_flag + (size_t)m * p.flag_stride, p.ticket_base + gstep + 2);
                    }
                }
            }
        }
    }

    if (piped && p.out_hidden && p.n_steps > 0) {  // the last step's rows: complete once every CTA has arrived
        grid_barrier_arrive(p.barrier, bar_target);
        if (grid_barrier_wait(p.barrier, bar_target, nullptr, p.status)) goto mk_aborted;
        if (blockIdx.x == 0 && threadIdx.x == 0) {
            const int pm = (p.n_steps - 1) % p.n_micro, pg = (p.n_steps - 1) / p.n_micro;
            st_release_sys(p.out_flag + (size_t)pm * p.flag_stride, p.ticket_base + pg + 1);
        }
    }
    if (false) {
mk_aborted:
        abort_drain();   // nothing is sent on; the next stage gives up in turn and the host reads the status word
    }
    // leave the barrier words at zero for the next launch: the last CTA out resets them
    __syncthreads();
    if (threadIdx.x == 0) {
        __threadfence();
        if (atomicAdd(p.done, 1u) == gridDim.x - 1) {
            *p.barrier = 0u;
            *p.done = 0u;
            __threadfence();
        }
    }
}

// ---- weight packer -----------------------------------------------------------------------------------------
// dst[panel][k/8][row r][8]  <-  source row chosen by `kind` (see megakernel.h); rows past n0 are zero.
__global__ void pack_panels_kernel(const PackJob j) {
    const int kc_n = j.K >> 3;
    const size_t total = (size_t)j.n_panels * kc_n * 8;
    for (size_t i = (size_t)blockIdx.x * blockDim.x + threadIdx.x; i < total; i += (size_t)gridDim.x * blockDim.x) {
        const int r = (int)(i & 7);
        const int kc = (int)((i >> 3) % kc_n);
        const int panel = (int)(i / ((size_t)kc_n * 8));
        const bf16* src = nullptr;
        if (j.kind == MK_QKV) {
            const int hd = j.head_dim, half = hd >> 1, ppb = half >> 2;
            const int head = panel / ppb, jj = (panel - head * ppb) * 4 + (r >> 1) + (r & 1) * half;
            const int m = head < j.n_heads ? 0 : (head < j.n_heads + j.n_kv_heads ? 1 : 2);
            const int hl = head - (m == 0 ? 0 : (m == 1 ? j.n_heads : j.n_heads + j.n_kv_heads));
            src = static_cast<const bf16*>(j.src[m]) + (size_t)(hl * hd + jj) * j.K;
        } else if (j.kind == MK_GATEUP) {
            src = static_cast<const bf16*>(j.src[r & 1]) + (size_t)(panel * 4 + (r >> 1)) * j.K;
        } else {
            const int row = panel * 8 + r;
            if (row < j.n0) src = static_cast<const bf16*>(j.src[0]) + (size_t)row * j.K;
        }
        uint4 v = make_uint4(0u, 0u, 0u, 0u);
        if (src) v = *reinterpret_cast<const uint4*>(src + (size_t)kc * 8);
        // [panel][32-k block][row][4 chunks of 8 k]: the 32 lanes of a warp read 512 contiguous bytes
        const size_t off = (size_t)panel * j.K * 16 + (size_t)(kc >> 2) * 512 + r * 64 + (kc & 3) * 16;
        *reinterpret_cast<uint4*>(static_cast<unsigned char*>(j.dst) + off) = v;
    }
}

}  // namespace

// ---- host side -------------------------------------------------------------------------------------------

size_t mk_packed_bytes(int n_panels, int K) { return (size_t)n_panels * K * 16; }

cudaError_t launch_pack(const PackJob& job, cudaStream_t st) {
    if (job.K % 8 || job.n_panels <= 0) return cudaErrorInvalidValue;
    pack_panels_kernel<<<1184, 256, 0, st>>>(job);
    count_launch();
    return cudaGetLastError();
}

size_t mk_smem_bytes(const MkParams& p, int stages) {
    auto up = [](size_t v) { return (v + 127) / 128 * 128; };
    size_t b = up((size_t)p.n_phases * sizeof(MkPhase));
    b += up((size_t)kWarps * stages * 8);
    b += mk_rows_region_bytes(p);
    b += mk_part_bytes(p);
    b += (size_t)p.n_rows * p.head_dim * 4;
    b += (size_t)((p.n_rows * p.pages_per_seq * 4 + 15) / 16) * 16;
    b = up(b) + 128;
    b += (size_t)kWarps * stages * kStageBytes;
    return b;
}

static int mk_pick_stages(const MkParams& p) {
    int cap = 6;
    if (const char* e = getenv("DLLM_MK_STAGES")) { const int v = atoi(e); if (v >= 2 && v <= 6) cap = v; }  // experiments
    for (int s = cap; s >= 2; --s)
        if (mk_smem_bytes(p, s) <= (size_t)MK_SMEM_BUDGET) return s;
    return 0;
}

template <bool HI, bool TR, int RK, bool LL, int HD, int PP, bool XG = true>
static cudaError_t mk_launch_t(const MkParams& p, int grid, size_t smem, cudaStream_t st) {
    cudaError_t e = cudaFuncSetAttribute(decode_megakernel<HI, TR, RK, LL, HD, PP, XG>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem);
    if (e != cudaSuccess) return e;
    cudaLaunchConfig_t cfg{};
    cfg.gridDim = dim3(grid);
    cfg.blockDim = dim3(kThreads);
    cfg.dynamicSmemBytes = smem;
    cfg.stream = st;
    cudaLaunchAttribute attr[1];
    attr[0].id = cudaLaunchAttributeCooperative;  // all CTAs co-resident: the grid barrier cannot deadlock
    attr[0].val.cooperative = 1;
    cfg.attrs = attr;
    cfg.numAttrs = 1;
    e = cudaLaunchKernelEx(&cfg, decode_megakernel<HI, TR, RK, LL, HD, PP, XG>, p);
    count_launch();
    return e;
}

cudaError_t launch_decode_megakernel(const MkParams& p_in, int num_sms, cudaStream_t st) {
    MkParams p = p_in;
    if (!mk_supported(p.n_rows, p.n_heads, p.n_kv_heads, p.head_dim)) return cudaErrorInvalidValue;
    if (p.n_splits < 1 || p.n_splits > kMaxSplits || p.n_phases > MK_MAX_PHASES) return cudaErrorInvalidValue;
    if ((size_t)p.n_rows * p.n_heads * p.n_splits * 8 > (size_t)MK_PART_UNITS * kPartStride * 4) return cudaErrorInvalidValue;  // merge scratch
    p.stages = mk_pick_stages(p);
    if (p.stages == 0) return cudaErrorInvalidValue;
    const size_t smem = mk_smem_bytes(p, p.stages);
    const bool ll = p.norm == NORM_RMS && p.rope && p.wpe == nullptr;   // Llama / Qwen2 (GPT-2: LayerNorm, learned positions, gelu)
    // Variants: one generic kernel per (more than 8 rows | time stamps | one row | several rows), and for the shapes that
    // matter most -- Llama-style models, no time stamps -- specialised ones: a single row with head_dim 64 / 128 as
    // compile-time constants and the fused hand-off compiled in or out (the headline decode and the default pipeline
    // bench), several rows with a run-time head_dim.
    if (p.n_rows > 8) return mk_launch_t<true, false, 0, false, 0, 2>(p, num_sms, smem, st);   // (no time stamps in this variant)
    if (p.trace) return p.n_rows == 1 ? mk_launch_t<false, true, 1, false, 0, 2>(p, num_sms, smem, st) : mk_launch_t<false, true, 0, false, 0, 2>(p, num_sms, smem, st);
    if (p.n_rows == 1) {
        const bool piped = p.in_flag != nullptr;
        const bool plain = ll && !p.act_xlayout;   // (a single row never needs the streamed-rows variant; DLLM_MK_XG=1 can force it)
        if (plain && p.head_dim == 64)
            return piped ? mk_launch_t<false, false, 1, true, 64, 1, false>(p, num_sms, smem, st) : mk_launch_t<false, false, 1, true, 64, 0, false>(p, num_sms, smem, st);
        if (plain && p.head_dim == 128)
            return piped ? mk_launch_t<false, false, 1, true, 128, 1, false>(p, num_sms, smem, st) : mk_launch_t<false, false, 1, true, 128, 0, false>(p, num_sms, smem, st);
        return mk_launch_t<false, false, 1, false, 0, 2>(p, num_sms, smem, st);
    }
    return ll ? mk_launch_t<false, false, 0, true, 0, 2>(p, num_sms, smem, st) : mk_launch_t<false, false, 0, false, 0, 2>(p, num_sms, smem, st);
}

int mk_max_splits(int n_rows, int n_heads, int head_dim) {
    (void)head_dim;
    const size_t scratch = (size_t)MK_PART_UNITS * kPartStride * 4;  // the partial buffer doubles as merge scratch
    const long s = (long)(scratch / ((size_t)n_rows * n_heads * 8));
    return (int)(s > kMaxSplits ? kMaxSplits : s);
}

bool mk_supported(int n_rows, int n_heads, int n_kv_heads, int head_dim) {
    return n_rows >= 1 && n_rows <= MK_MAX_ROWS && n_heads % n_kv_heads == 0 && n_heads / n_kv_heads <= kMaxGq &&
           (head_dim == 32 || head_dim == 64 || head_dim == 128);
}

}  // namespace dllm
