// Causal prefill attention over the paged KV cache on the 5th-generation tensor cores.
//
// One CTA = 128 consecutive query rows of one sequence x one query head.  Both products of every 128-key tile run as
// tcgen05.mma with fp32 accumulators in tensor memory; all operands arrive by TMA:
//   * Q tile (128 x hd) once, K and V tiles page by page (a KV page of 64 tokens is one 64-row box of a 2-D tensor
//     map over the layer's pool; the block table gives the row coordinate), 128-byte swizzle, two-stage ring;
//   * S = Q K^T: A = Q (K-major), B = K (K-major), N = 128 keys, into one of two S buffers in tensor memory, so that
//     the scores of tile j+1 are computed while the softmax warps work on tile j;
//   * the softmax warps read the scores with tcgen05.ld (thread = TMEM lane = query row, half of the tile's keys
//     each), keep them as rounded bf16 pairs, exchange the row maximum, and store P = 2^((s - m) log2 e), rounded to
//     bf16, into shared memory in the swizzled K-major layout the next MMA reads;
//   * O_j = P_j V_j: A = P (K-major), B = V as it lies in the cache -- [token][hd], i.e. MN-major -- into a fresh
//     accumulator (two O buffers); the softmax warps add it into their fp32 registers with the online-softmax
//     correction, so no accumulator in tensor memory is ever rescaled.
// Warp roles: warp 0 = TMA producer, warp 1 = TMEM allocator + MMA issuer, warps 2-9 = softmax / accumulate / store
// (two threads per query row: the scalar work per tile, not the tensor cores, is what has to be spread out).
// Rounding follows HF eager attention exactly as prefill_attention.cu does: scores rounded to bf16 after the dot
// product and after scaling, P (unnormalised) rounded to bf16 for the second product, fp32 sums.
// Needs 64-token pages and head_dim 64 or 128; other shapes take the mma.sync kernel of prefill_attention.cu.
#include <cuda.h>
#include <float.h>
#include <stdlib.h>

#include "common.cuh"
#include "kernels.h"
#include "ptx.cuh"

namespace dllm {

namespace {

using namespace ptx;

constexpr int QT = 128, KT = 128, PG = 64;   // query rows per CTA, keys per tile, tokens per KV page
constexpr int kThreads = 320;             // TMA warp, MMA warp, eight softmax warps
constexpr int kSliceBytes = 128 * 128;       // one 64-column slice of a 128-row tile: 128 rows x 128 bytes

__device__ __forceinline__ float ex2f(float x) {   // 2^x, one MUFU instruction; 2^-inf = 0
    float y;
    asm("ex2.approx.ftz.f32 %0, %1;" : "=f"(y) : "f"(x));
    return y;
}

// K-major operand tile, rows of 128 bytes, 128-byte swizzle (what TMA SWIZZLE_128B writes; P is stored the same way):
// stride between 8-row groups 1024 B, descriptor version 1 (sm_100), layout type 2.
__device__ __forceinline__ uint64_t desc_kmajor(uint32_t smem_addr) {
    return (uint64_t)((smem_addr >> 4) & 0x3FFF) | (1ull << 16) | ((uint64_t)(1024 >> 4) << 32) | (1ull << 46) | (2ull << 61);
}
// MN-major operand (V: rows = tokens = the contraction, 64 head-dim columns = 128 bytes per row and slice): 8 token rows
// per swizzle atom (stride 1024 B between atoms along K), the next 64 columns `slice_bytes` further on.
__device__ __forceinline__ uint64_t desc_mnmajor(uint32_t smem_addr, uint32_t slice_bytes) {
    return (uint64_t)((smem_addr >> 4) & 0x3FFF) | ((uint64_t)((slice_bytes >> 4) & 0x3FFF) << 16) |
           ((uint64_t)(1024 >> 4) << 32) | (1ull << 46) | (2ull << 61);
}
// D fp32, A and B bf16, M = 128, N = n; b_mn: B is MN-major
constexpr uint32_t make_idesc(int n, bool b_mn) {
    return (1u << 4) | (1u << 7) | (1u << 10) | (b_mn ? (1u << 16) : 0u) | ((uint32_t)(n >> 3) << 17) | ((uint32_t)(128 >> 4) << 24);
}

struct TcAttnParams {
    const int32_t* positions;
    const int32_t* seq_ids;
    const int32_t* block_table;
    void* out;                 // [R, H*hd] bf16
    int32_t pages_per_seq, n_heads, n_kv_heads, rows_per_seq;
};

template <int HD>
__global__ void __launch_bounds__(kThreads, 1)
prefill_attn_tc_kernel(const __grid_constant__ CUtensorMap tm_q, const __grid_constant__ CUtensorMap tm_k,
                       const __grid_constant__ CUtensorMap tm_v, const TcAttnParams p) {
    constexpr int NS = HD / 64;                       // 64-column slices of a head
    constexpr int kTileBytes = NS * kSliceBytes;      // a 128-row x hd tile (Q, or one K / V stage)
    constexpr uint32_t kIdescQK = make_idesc(KT, false), kIdescPV = make_idesc(HD, true);
    extern __shared__ __align__(1024) unsigned char smem_raw[];
    unsigned char* smem = reinterpret_cast<unsigned char*>(((uintptr_t)smem_raw + 1023) & ~(uintptr_t)1023);
    unsigned char* sQ = smem;
    unsigned char* sK = sQ + kTileBytes;              // 2 stages
    unsigned char* sV = sK + 2 * kTileBytes;          // 2 stages
    unsigned char* sP = sV + 2 * kTileBytes;          // 2 slices of 64 keys: [128 rows][128 B] each
    __shared__ __align__(8) uint64_t bars[12];
    __shared__ uint32_t s_tmem;
    __shared__ float s_max[2][2][QT];   // [tile parity][column group][row]: the two halves of a row meet here
    const uint32_t b0 = smem_u32(bars);
    const uint32_t q_full = b0, kv_full = b0 + 8, kv_empty = b0 + 24, s_full = b0 + 40, s_free = b0 + 56, p_full = b0 + 72,
                   o_full = b0 + 80;

    const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
    const int H = p.n_heads, KV = p.n_kv_heads;
    const int n_qt = (p.rows_per_seq + QT - 1) / QT;
    const int qt = n_qt - 1 - (int)blockIdx.x;        // the longest tiles (last query rows) start first
    const int h = blockIdx.y, b = blockIdx.z, kvh = h / (H / KV);
    const int row0 = b * p.rows_per_seq + qt * QT;
    const int rows_here = min(QT, p.rows_per_seq - qt * QT);
    const int pos0 = p.positions[row0];               // positions in the run are consecutive
    const int last_key = pos0 + rows_here - 1;
    const int n_tiles = last_key / KT + 1;

    if (threadIdx.x == 0) {
        mbar_init(q_full, 1);
        for (int i = 0; i < 2; ++i) {
            mbar_init(kv_full + i * 8, 1);
            mbar_init(kv_empty + i * 8, 1);
            mbar_init(s_full + i * 8, 1);
            mbar_init(s_free + i * 8, 256);
            mbar_init(o_full + i * 8, 1);
        }
        mbar_init(p_full, 256);
        asm volatile("fence.mbarrier_init.release.cluster;" ::: "memory");
    }
    if (warp == 1) {
        asm volatile("tcgen05.alloc.cta_group::1.sync.aligned.shared::cta.b32 [%0], %1;" ::"r"(smem_u32(&s_tmem)), "r"(512));
        asm volatile("tcgen05.relinquish_alloc_permit.cta_group::1.sync.aligned;");
    }
    tc_fence_before();
    __syncthreads();
    tc_fence_after();
    const uint32_t tmem = s_tmem;
    pdl_launch_dependents();   // barriers and tensor memory are set up: from here on q / K / V of this step are read
    pdl_wait();
    const uint32_t tS = tmem, tO = tmem + 256;        // S buffers at columns 0 / 128, O buffers at 256 / 256 + HD

    if (warp == 0) {
        // ===== TMA producer =====
        if (lane == 0) {
            const int32_t* bt = p.block_table + (size_t)p.seq_ids[row0] * p.pages_per_seq;
            mbar_expect_tx(q_full, kTileBytes);
            for (int s = 0; s < NS; ++s) tma_load_2d(smem_u32(sQ + s * kSliceBytes), &tm_q, h * HD + s * 64, row0, q_full);
            for (int j = 0; j < n_tiles; ++j) {
                const int st = j & 1;
                if (j >= 2) mbar_wait(kv_empty + st * 8, ((j >> 1) & 1) ^ 1);
                const int n_pages = (j * KT + PG <= last_key) ? 2 : 1;   // the tile's second page holds a needed key?
                mbar_expect_tx(kv_full + st * 8, (uint32_t)(2 * n_pages * NS * PG * 128));
                for (int pg = 0; pg < n_pages; ++pg) {
                    const int page = bt[j * 2 + pg];
                    const int row = (page * KV + kvh) * PG;
                    for (int s = 0; s < NS; ++s) {
                        const uint32_t off = (uint32_t)(st * kTileBytes + s * kSliceBytes + pg * PG * 128);
                        tma_load_2d(smem_u32(sK) + off, &tm_k, s * 64, row, kv_full + st * 8);
                        tma_load_2d(smem_u32(sV) + off, &tm_v, s * 64, row, kv_full + st * 8);
                    }
                }
            }
        }
    } else if (warp == 1) {
        // ===== MMA issuer (one thread) =====
        if (lane == 0) {
            auto issue_qk = [&](int j) {
                const int st = j & 1;
                mbar_wait(kv_full + st * 8, (j >> 1) & 1);
                if (j >= 2) mbar_wait(s_free + st * 8, ((j >> 1) & 1) ^ 1);
                tc_fence_after();
                const uint32_t qb = smem_u32(sQ), kb = smem_u32(sK) + st * kTileBytes;
#pragma unroll
                for (int s = 0; s < NS; ++s)
#pragma unroll
                    for (int k = 0; k < 4; ++k)  // 16 of the head's columns per instruction: 32 bytes along the swizzled row
                        tc_mma(tS + st * KT, desc_kmajor(qb + s * kSliceBytes) + 2 * k, desc_kmajor(kb + s * kSliceBytes) + 2 * k,
                               kIdescQK, (s | k) ? 1u : 0u);
                tc_commit(s_full + st * 8);
            };
            mbar_wait(q_full, 0);
            issue_qk(0);
            for (int j = 0; j < n_tiles; ++j) {
                const int st = j & 1;
                if (j + 1 < n_tiles) issue_qk(j + 1);
                mbar_wait(p_full, j & 1);
                tc_fence_after();
                // keys of this tile that exist (rounded up to the 16 of one instruction; the rest of P is zero and is skipped)
                const int nk16 = (min(KT, last_key + 1 - j * KT) + 15) >> 4;
                const uint32_t pb = smem_u32(sP), vb = smem_u32(sV) + st * kTileBytes;
                for (int k = 0; k < nk16; ++k)  // 16 keys: 32 bytes along P's row, 16 token rows (2048 B) down V
                    tc_mma(tO + st * HD, desc_kmajor(pb + (k >> 2) * kSliceBytes) + 2 * (k & 3),
                           desc_mnmajor(vb + k * 2048, kSliceBytes), kIdescPV, k ? 1u : 0u);
                tc_commit(o_full + st * 8);   // O_j complete (and P free again)
                tc_commit(kv_empty + st * 8); // both products have read this stage
            }
        }
    } else {
        // ===== softmax / accumulate / store: eight warps, two threads per query row =====
        // A warp reaches the TMEM lanes 32 * (warp % 4) .. + 31, so warps 2-5 and 6-9 both cover all 128 rows: the first
        // group takes keys 0-63 of every tile and the first half of the head's columns, the second group the rest.
        // The two threads of a row exchange their maxima (and, at the end, their sums) through shared memory.
        const int grp = (warp - 2) >> 2;
        const int quarter = warp & 3;
        const int r = quarter * 32 + lane;
        const int qpos = pos0 + min(r, rows_here - 1);   // rows past the run repeat the last one (never stored)
        const uint32_t lane_base = (uint32_t)(quarter * 32) << 16;
        constexpr int OC = HD / 2;                        // output columns per thread
        const float scale = rsqrtf((float)HD);
        constexpr float kLog2e = 1.4426950408889634f;
        float m = -FLT_MAX, l = 0.f;
        float o[OC];
#pragma unroll
        for (int i = 0; i < OC; ++i) o[i] = 0.f;
        unsigned char* prow = sP + (size_t)grp * kSliceBytes + (size_t)r * 128;   // this thread's 64 keys: one swizzled row
        for (int j = 0; j < n_tiles; ++j) {
            const int st = j & 1, kt0 = j * KT + grp * 64;
            mbar_wait(s_full + st * 8, (j >> 1) & 1);
            tc_fence_after();
            const uint32_t ts = tS + st * KT + grp * 64 + lane_base;
            // pass 1: scores rounded where HF rounds them (after the product, after scaling), masked, kept as bf16 pairs
            uint32_t sv[32];
            float mx = -FLT_MAX;
#pragma unroll
            for (int c0 = 0; c0 < 64; c0 += 32) {
                uint32_t v[32];
                tc_ld32(ts + c0, v);
#pragma unroll
                for (int i = 0; i < 32; i += 2) {
                    float s0 = bf16_round(bf16_round(__uint_as_float(v[i])) * scale);
                    float s1 = bf16_round(bf16_round(__uint_as_float(v[i + 1])) * scale);
                    s0 = kt0 + c0 + i <= qpos ? s0 : -INFINITY;
                    s1 = kt0 + c0 + i + 1 <= qpos ? s1 : -INFINITY;
                    mx = fmaxf(mx, fmaxf(s0, s1));
                    sv[(c0 + i) >> 1] = pack_bf16(s0, s1);   // exact: both are bf16 values (or -inf)
                }
            }
            tc_fence_before();
            mbar_arrive(s_free + st * 8);            // this S buffer may be overwritten by the tile after next
            s_max[st][grp][r] = mx;
            asm volatile("bar.sync 1, 256;" ::: "memory");   // the eight softmax warps
            const float m_new = fmaxf(m, fmaxf(mx, s_max[st][grp ^ 1][r]));
            const float corr = ex2f((m - m_new) * kLog2e);
            // P of the previous tile has been consumed (and its product is ready) before P is overwritten
            if (j > 0) mbar_wait(o_full + (st ^ 1) * 8, ((j - 1) >> 1) & 1);
            // pass 2: p = exp(s - m) from the kept scores, rounded to bf16, into this thread's row of the swizzled P tile
            float sum = 0.f;
            const float mb = m_new * kLog2e;
#pragma unroll
            for (int c = 0; c < 8; ++c) {            // eight 16-byte chunks of 8 keys; chunk index xor (row % 8): 128-byte swizzle
                uint32_t pk[4];
#pragma unroll
                for (int q2 = 0; q2 < 4; ++q2) {
                    const uint32_t w = sv[c * 4 + q2];
                    const float p0 = ex2f(fmaf(__uint_as_float(w << 16), kLog2e, -mb));
                    const float p1 = ex2f(fmaf(__uint_as_float(w & 0xffff0000u), kLog2e, -mb));
                    sum += p0 + p1;
                    pk[q2] = pack_bf16(p0, p1);
                }
                *reinterpret_cast<uint4*>(prow + ((c ^ (r & 7)) << 4)) = make_uint4(pk[0], pk[1], pk[2], pk[3]);
            }
            asm volatile("fence.proxy.async.shared::cta;" ::: "memory");  // P: generic-proxy stores -> read by the MMA (async proxy)
            mbar_arrive(p_full);
            // o = (o + P_{j-1} V_{j-1}) * corr : both scaled to the previous maximum, then moved to the new one
            if (j > 0) {
                tc_fence_after();
                const uint32_t to = tO + (st ^ 1) * HD + grp * OC + lane_base;
#pragma unroll
                for (int c0 = 0; c0 < OC; c0 += 32) {
                    uint32_t v[32];
                    tc_ld32(to + c0, v);
#pragma unroll
                    for (int i = 0; i < 32; ++i) o[c0 + i] = (o[c0 + i] + __uint_as_float(v[i])) * corr;
                }
            }
            l = l * corr + sum;
            m = m_new;
        }
        {   // the last tile's product, the row sum of both halves, normalise, store
            const int st = (n_tiles - 1) & 1;
            s_max[st ^ 1][grp][r] = l;               // (the buffer of the other parity: nobody reads its maxima any more)
            asm volatile("bar.sync 1, 256;" ::: "memory");
            l += s_max[st ^ 1][grp ^ 1][r];
            mbar_wait(o_full + st * 8, ((n_tiles - 1) >> 1) & 1);
            tc_fence_after();
            const uint32_t to = tO + st * HD + grp * OC + lane_base;
            const float inv = l > 0.f ? 1.0f / l : 0.f;
            bf16* out = static_cast<bf16*>(p.out) + ((size_t)(row0 + r) * H + h) * HD + grp * OC;
#pragma unroll
            for (int c0 = 0; c0 < OC; c0 += 32) {
                uint32_t v[32];
                tc_ld32(to + c0, v);
                if (r < rows_here) {
#pragma unroll
                    for (int i = 0; i < 32; i += 8) {
                        uint4 w;
                        w.x = pack_bf16((o[c0 + i] + __uint_as_float(v[i])) * inv, (o[c0 + i + 1] + __uint_as_float(v[i + 1])) * inv);
                        w.y = pack_bf16((o[c0 + i + 2] + __uint_as_float(v[i + 2])) * inv, (o[c0 + i + 3] + __uint_as_float(v[i + 3])) * inv);
                        w.z = pack_bf16((o[c0 + i + 4] + __uint_as_float(v[i + 4])) * inv, (o[c0 + i + 5] + __uint_as_float(v[i + 5])) * inv);
                        w.w = pack_bf16((o[c0 + i + 6] + __uint_as_float(v[i + 6])) * inv, (o[c0 + i + 7] + __uint_as_float(v[i + 7])) * inv);
                        *reinterpret_cast<uint4*>(out + c0 + i) = w;
                    }
                }
            }
        }
        tc_fence_before();
    }
    __syncthreads();
    if (warp == 1) {
        tc_fence_after();
        asm volatile("tcgen05.dealloc.cta_group::1.sync.aligned.b32 %0, %1;" ::"r"(tmem), "r"(512));
    }
}

typedef CUresult (*EncodeTiledFn)(CUtensorMap*, CUtensorMapDataType, cuuint32_t, void*, const cuuint64_t*, const cuuint64_t*,
                                  const cuuint32_t*, const cuuint32_t*, CUtensorMapInterleave, CUtensorMapSwizzle,
                                  CUtensorMapL2promotion, CUtensorMapFloatOOBfill);

EncodeTiledFn get_encode() {
    static EncodeTiledFn fn = nullptr;
    if (!fn) {
        void* ptr = nullptr;
        cudaDriverEntryPointQueryResult q;
        if (cudaGetDriverEntryPoint("cuTensorMapEncodeTiled", &ptr, cudaEnableDefault, &q) == cudaSuccess &&
            q == cudaDriverEntryPointSuccess)
            fn = reinterpret_cast<EncodeTiledFn>(ptr);
    }
    return fn;
}

// [rows, cols] row-major bf16, box = box_rows x 64 columns (128 bytes), 128-byte swizzle, out-of-bounds reads return zero
bool make_map(CUtensorMap* map, const void* base, size_t rows, size_t cols, int box_rows) {
    EncodeTiledFn enc = get_encode();
    if (!enc) return false;
    cuuint64_t dims[2] = {(cuuint64_t)cols, (cuuint64_t)rows};
    cuuint64_t strides[1] = {(cuuint64_t)cols * 2};
    cuuint32_t box[2] = {64, (cuuint32_t)box_rows};
    cuuint32_t estr[2] = {1, 1};
    return enc(map, CU_TENSOR_MAP_DATA_TYPE_BFLOAT16, 2, const_cast<void*>(base), dims, strides, box, estr,
               CU_TENSOR_MAP_INTERLEAVE_NONE, CU_TENSOR_MAP_SWIZZLE_128B, CU_TENSOR_MAP_L2_PROMOTION_L2_256B,
               CU_TENSOR_MAP_FLOAT_OOB_FILL_NONE) == CUDA_SUCCESS;
}

template <int HD>
cudaError_t launch_t(const PrefillAttnParams& p, cudaStream_t st) {
    CUtensorMap q, k, v;
    const size_t R = (size_t)p.n_seqs * p.rows_per_seq;
    const size_t kv_rows = (size_t)p.num_pages * p.n_kv_heads * PG;
    if (!make_map(&q, p.q, R, (size_t)p.n_heads * HD, QT) || !make_map(&k, p.kv.kpool, kv_rows, HD, PG) ||
        !make_map(&v, p.kv.vpool, kv_rows, HD, PG))
        return cudaErrorNotSupported;
    // Q + two K stages + two V stages + P; at least 117 KB so that a second CTA (and its 512 TMEM columns) never shares the SM
    size_t smem = (size_t)(HD / 64) * kSliceBytes * 5 + 2 * kSliceBytes + 1024;
    if (smem < 117 * 1024) smem = 117 * 1024;
    static PerDeviceOnce attr_once;
    if (attr_once.need()) {
        cudaError_t e = cudaFuncSetAttribute(prefill_attn_tc_kernel<HD>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem);
        if (e != cudaSuccess) return e;
    }
    TcAttnParams a{p.positions, p.seq_ids, p.kv.block_table, p.out, p.kv.pages_per_seq, p.n_heads, p.n_kv_heads, p.rows_per_seq};
    dim3 grid((p.rows_per_seq + QT - 1) / QT, p.n_heads, p.n_seqs);
    const cudaError_t le = launch_pdl(prefill_attn_tc_kernel<HD>, grid, dim3(kThreads), smem, st, q, k, v, a);
    count_launch();
    return le;
}

}  // namespace

bool prefill_attn_tc_supported(const PrefillAttnParams& p) {
    const char* env = getenv("DLLM_PREFILL_ATTN_TC");  // "0": the mma.sync kernel (A/B measurements, tests)
    if (env && env[0] == '0') return false;
    return (p.head_dim == 64 || p.head_dim == 128) && p.kv.page_shift == 6 && p.num_pages > 0 && p.n_seqs > 0 &&
           p.rows_per_seq > 0 && p.n_heads % p.n_kv_heads == 0;
}

cudaError_t launch_prefill_attn_tc(const PrefillAttnParams& p, cudaStream_t st) {
    if (!prefill_attn_tc_supported(p)) return cudaErrorInvalidValue;
    return p.head_dim == 64 ? launch_t<64>(p, st) : launch_t<128>(p, st);
}

}  // namespace dllm
