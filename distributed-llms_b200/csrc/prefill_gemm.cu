// Prefill GEMM for sm_100a: C[M, N] = A[M, K] . W[N, K]^T on the 5th-generation tensor cores.
//
// Only prefill is a dense contraction: M = tokens of the prompt chunk, N = output features,
// both operands K-major in HBM exactly as the activations / nn.Linear weights already lie there.
//   * TMA (cp.async.bulk.tensor.2d, 128-byte swizzle) stages 128x64 tiles of A and W into a 4-deep ring;
//   * one elected thread issues tcgen05.mma.cta_group::1.kind::f16 (M=128, N=128, K=16) from shared-memory
//     descriptors; the fp32 accumulator lives in tensor memory (128 lanes x 128 or 256 columns);
//   * tcgen05.commit releases ring slots and finally signals the epilogue warps, which read the accumulator with
//     tcgen05.ld and apply the fused epilogue: bias | bias + residual (in place) | SwiGLU of a gate/up tile pair;
//   * warp roles: warp 0 = TMA producer, warp 1 = TMEM allocator + MMA issuer, warps 2-5 = epilogue (one TMEM lane
//     quarter each).  Rounding points are those of the decode kernels (bf16 after bias, after the residual add,
//     after SiLU and after the product).
#include <cuda.h>
#include <cuda_bf16.h>
#include <cuda_runtime.h>
#include <float.h>
#include <stdint.h>
#include <stdlib.h>

#include <stdio.h>

#include "common.cuh"
#include "kernels.h"
#include "ptx.cuh"

namespace dllm {

namespace {

using namespace ptx;

constexpr int BM = 128, BK = 64;                // tile rows / k-slice; BK * 2 bytes = one 128-byte swizzle row
// ring depth: as many stages as 200 KB of shared memory hold (4 for the 48 KB stages, 6 for 32 KB, 8 for 24 KB)
constexpr int gemm_stages(int stage_bytes) { return 200 * 1024 / stage_bytes > 8 ? 8 : 200 * 1024 / stage_bytes; }
constexpr int kTileBytes = BM * BK * 2;         // 16 KB per 128-row operand tile
// Output tile width BN: 128, or 256 when the grid still fills the machine -- the A tile is then read once per
// 256 columns (87 instead of 64 flop per shared-memory byte staged).
constexpr int kGemmThreads = 192;


// ---- CTA pairs (cta_group::2): two CTAs of a cluster on one TPC share one UMMA of M = 256 --------------------------
// Every CTA loads its own 128 A rows and HALF of the B tile; the leader (cluster rank 0) issues the MMAs, which read the
// operands from both CTAs' shared memory and write each CTA's 128 accumulator rows into its own tensor memory.  All TMA
// loads are plain (no multicast) 2-D tiled loads; those of the second CTA signal the LEADER's mbarrier.
__device__ __forceinline__ uint32_t cluster_ctarank() {
    uint32_t r;
    asm volatile("mov.u32 %0, %%cluster_ctarank;" : "=r"(r));
    return r;
}
__device__ __forceinline__ uint32_t mapa_u32(uint32_t local_addr, uint32_t rank) {   // the same offset in CTA `rank`'s shared memory
    uint32_t remote;
    asm volatile("mapa.shared::cluster.u32 %0, %1, %2;" : "=r"(remote) : "r"(local_addr), "r"(rank));
    return remote;
}
__device__ __forceinline__ void mbar_expect_tx_cluster(uint32_t cluster_bar, uint32_t bytes) {
    asm volatile("mbarrier.arrive.expect_tx.release.cluster.shared::cluster.b64 _, [%0], %1;" ::"r"(cluster_bar), "r"(bytes) : "memory");
}
__device__ __forceinline__ void mbar_arrive_cluster(uint32_t cluster_bar) {
    asm volatile("mbarrier.arrive.release.cluster.shared::cluster.b64 _, [%0];" ::"r"(cluster_bar) : "memory");
}
__device__ __forceinline__ void mbar_wait_cluster(uint32_t bar, uint32_t parity) {   // own barrier, arrivals from the pair
    uint32_t ok = 0;
    for (unsigned spins = 0; !ok; ++spins) {
        asm volatile(
            "{\n\t.reg .pred p;\n\t"
            "mbarrier.try_wait.parity.acquire.cluster.shared::cta.b64 p, [%1], %2;\n\t"
            "selp.u32 %0, 1, 0, p;\n\t}" : "=r"(ok) : "r"(bar), "r"(parity) : "memory");
        if (!ok && spins > (1u << 26)) __trap();
    }
}
__device__ __forceinline__ void tma_load_2d_pair(uint32_t dst, const CUtensorMap* map, int c0, int c1, uint32_t cluster_bar) {
    asm volatile("cp.async.bulk.tensor.2d.cta_group::2.shared::cluster.global.mbarrier::complete_tx::bytes [%0], [%1, {%3, %4}], [%2];"
                 ::"r"(dst), "l"(map), "r"(cluster_bar), "r"(c0), "r"(c1) : "memory");
}
__device__ __forceinline__ void tc_commit_pair(uint32_t bar) {   // arrives on the barrier at this offset in BOTH CTAs
    asm volatile("tcgen05.commit.cta_group::2.mbarrier::arrive::one.shared::cluster.multicast::cluster.b64 [%0], %1;"
                 ::"r"(bar), "h"((uint16_t)3) : "memory");
}
__device__ __forceinline__ void tc_mma_pair(uint32_t d_tmem, uint64_t a_desc, uint64_t b_desc, uint32_t idesc, uint32_t accumulate) {
    asm volatile(
        "{\n\t.reg .pred p;\n\t"
        "setp.ne.b32 p, %4, 0;\n\t"
        "tcgen05.mma.cta_group::2.kind::f16 [%0], %1, %2, %3, p;\n\t}"
        ::"r"(d_tmem), "l"(a_desc), "l"(b_desc), "r"(idesc), "r"(accumulate) : "memory");
}
__device__ __forceinline__ void cluster_sync_pair() {
    asm volatile("barrier.cluster.arrive.release.aligned;\n\tbarrier.cluster.wait.acquire.aligned;" ::: "memory");
}



// Instruction descriptor: D fp32, A and B bf16, both K-major, N = bn, M = 128.
constexpr uint32_t make_idesc(int bn, int m = BM) {
    return (1u << 4) | (1u << 7) | (1u << 10) | ((uint32_t)(bn >> 3) << 17) | ((uint32_t)(m >> 4) << 24);
}

__device__ __forceinline__ uint32_t pack2(float lo, float hi) {
    __nv_bfloat162 v = __floats2bfloat162_rn(lo, hi);
    return *reinterpret_cast<uint32_t*>(&v);
}
__device__ __forceinline__ float silu_f(float g) { return g / (1.0f + expf(-g)); }
__device__ __forceinline__ float gelu_new_f(float x) {
    return 0.5f * x * (1.0f + tanhf(0.7978845608028654f * (x + 0.044715f * x * x * x)));
}

// One tile's coordinates: which row segment of the logical weight matrix (q | k | v in one launch; BIAS epilogue only),
// its tensor map, its first row / column.  Tiles are numbered m-fastest, so that the CTAs working side by side share a
// weight tile (and L2 serves it once).
struct TileCoord {
    int seg, m0, n0, col0, seg_n;
};
template <int EPI, int BN, int CG = 1>
__device__ __forceinline__ TileCoord tile_coord(const GemmTcParams& p, int t, int tiles_m) {
    TileCoord c;
    c.seg = 0; c.col0 = 0; c.seg_n = p.N;
    int tn = t / tiles_m;
    c.m0 = (t - tn * tiles_m) * (BM * CG);   // a CTA pair covers 256 rows
    if (EPI != GEMM_EPI_SWIGLU) {
        const int t0 = (p.N + BN - 1) / BN, t1 = (p.n_more[0] + BN - 1) / BN;
        if (tn >= t0 + t1) { c.seg = 2; tn -= t0 + t1; c.col0 = p.N + p.n_more[0]; c.seg_n = p.n_more[1]; }
        else if (tn >= t0) { c.seg = 1; tn -= t0; c.col0 = p.N; c.seg_n = p.n_more[0]; }
    }
    c.n0 = tn * BN;
    return c;
}

// Work units of the persistent kernel.  Tiles of the full waves are computed whole; the tiles of the last, partly filled
// wave (`n_tiles - whole` of them, fewer than there are SMs) are cut along K into `splits` slices, one slice per CTA, so
// that the tail keeps every SM loading instead of a few of them for a whole tile's time -- at M = 512 a GEMM IS its
// tail, and what bounds it is the rate at which the SMs that hold a tile can fill their shared memory.  Every slice
// leaves its fp32 accumulator in the scratch area, waits until the tile's other slices have done so (they are running
// at the same time: the launch is cooperative and a slice is its CTA's last unit), then adds the slices in slice order
// -- the same sum on every run -- for ITS share of the tile's columns and runs the epilogue on them: reduction and
// epilogue are spread over the tile's CTAs as well.
struct UnitSched {
    int n_units, whole, splits, kb_per;
};
struct Unit {
    int tile, kb0, kb1, tail, slice;   // tail: index of the (tile, slice) pair among the split units, -1 for a whole tile
};
__device__ __forceinline__ Unit unit_of(const UnitSched& us, int u, int num_kb) {
    Unit r;
    if (u < us.whole) {
        r.tile = u; r.kb0 = 0; r.kb1 = num_kb; r.tail = -1; r.slice = 0;
    } else {
        const int v = u - us.whole, tt = v / us.splits, sl = v - tt * us.splits;
        r.tile = us.whole + tt; r.kb0 = sl * us.kb_per; r.kb1 = min(num_kb, r.kb0 + us.kb_per); r.tail = v; r.slice = sl;
    }
    return r;
}

// Where an epilogue thread's 32 accumulator columns come from: tensor memory, or the sum of a split tile's slices.
// A slice in the scratch area is [32-column chunk][4-column group j4][row] float4: a warp's 32 rows are 512 contiguous
// bytes in every access, and one chunk of one slice is one contiguous 16 KB block -- the size of a bulk copy.  The
// CTA's share of the tile (a few chunks of every slice, <= kRedBlocks blocks) is brought into the idle operand ring
// by bulk copies that are all in flight together; the sum then reads shared memory.
constexpr int kRedBlockBytes = BM * 32 * 4;                  // 16 KB
constexpr int kRedBlocks = 192 * 1024 / kRedBlockBytes;      // the ring holds 12 of them, whatever the tile width
struct AccSrc {
    uint32_t tbase;             // TMEM address of (lane quarter, column 0) of the accumulator
    const unsigned char* red;   // staged blocks [chunk of the share][slice], or null: read tensor memory
    int row;                    // this thread's row inside the tile
    int splits;
    int first_chunk;            // first 32-column chunk of this CTA's share of the output columns
};
__device__ __forceinline__ size_t ws_offset(int col32, int j4, int row) { return ((size_t)(col32 * 8 + j4) * BM + row) * 4; }
// index of accumulator chunk `cc` among the staged chunks: out chunks in order; the dual tile keeps (gate, up) together
template <bool kDual, int BN>
__device__ __forceinline__ int staged_chunk(int cc, int first_chunk) {
    constexpr int kOut = BN / 32;
    return kDual ? ((cc % kOut) - first_chunk) * 2 + cc / kOut : cc - first_chunk;
}
template <bool kDual, int BN>
__device__ __forceinline__ void acc_load32(const AccSrc& a, int col, uint32_t (&acc)[32]) {
    if (a.red == nullptr) {   // uniform over the epilogue warps
        tc_ld32(a.tbase + col, acc);
        return;
    }
    float f[32];
#pragma unroll
    for (int j = 0; j < 32; ++j) f[j] = 0.f;
    const unsigned char* blk = a.red + (size_t)staged_chunk<kDual, BN>(col >> 5, a.first_chunk) * a.splits * kRedBlockBytes + a.row * 16;
    for (int sl = 0; sl < a.splits; ++sl, blk += kRedBlockBytes) {   // slice order: the same sum on every run
#pragma unroll
        for (int j4 = 0; j4 < 8; ++j4) {
            const float4 v = *reinterpret_cast<const float4*>(blk + j4 * (BM * 16));
            f[4 * j4] += v.x; f[4 * j4 + 1] += v.y; f[4 * j4 + 2] += v.z; f[4 * j4 + 3] += v.w;
        }
    }
#pragma unroll
    for (int j = 0; j < 32; ++j) acc[j] = __float_as_uint(f[j]);
}
__device__ __forceinline__ void bulk_load(uint32_t dst, const void* src, uint32_t bytes, uint32_t bar) {
    asm volatile("cp.async.bulk.shared::cluster.global.mbarrier::complete_tx::bytes [%0], [%1], %2, [%3];"
                 ::"r"(dst), "l"(src), "r"(bytes), "r"(bar) : "memory");
}
__device__ __forceinline__ void fence_proxy_async_global() { asm volatile("fence.proxy.async.global;" ::: "memory"); }
__device__ __forceinline__ unsigned ld_acquire_gpu(const unsigned* p) {
    unsigned v;
    asm volatile("ld.acquire.gpu.global.u32 %0, [%1];" : "=r"(v) : "l"(p) : "memory");
    return v;
}

// GEMM_EPI_QKV: the tail of K1 in the projection's epilogue.  A thread owns one token (row) and walks the heads of the
// tile; it holds all HD columns of a head in registers, so both elements of every rotary pair (j, j + HD/2) are at hand.
// Bias, rounding and rotary (q and k segments) in rope_append_kernel's order of roundings; then the head leaves as one
// contiguous 128- or 256-byte row: q to its buffer, k / v into the KV page of (slot, position).  tcgen05.ld is
// warp-wide, so rows past M take the loads too and skip only the arithmetic and the stores.
template <int HD, int BN>
__device__ __forceinline__ void qkv_tail(const GemmTcParams& p, const AccSrc& src, int row, int seg, int n0, int seg_n,
                                         const bf16* bias, int c_begin, int c_end) {
    constexpr int kHalf = HD / 2;
    const bool live = row < p.M;
    const int pos = live ? p.positions[row] : 0;
    const float* cs = p.rope_cos + (size_t)pos * kHalf;
    const float* sn = p.rope_sin + (size_t)pos * kHalf;
    bf16* kv_row = nullptr;
    if (live && seg != 0) {
        const int page = p.kv.block_table[(size_t)p.seq_ids[row] * p.kv.pages_per_seq + (pos >> p.kv.page_shift)];
        const int off = pos & ((1 << p.kv.page_shift) - 1);
        bf16* pool = static_cast<bf16*>(seg == 1 ? p.kv.kpool : p.kv.vpool);
        kv_row = pool + ((((size_t)page * p.n_kv_heads) << p.kv.page_shift) + off) * HD;   // head 0 of that token
    }
    const bool rotate = p.rope && seg != 2;
#pragma unroll 1
    for (int h0 = c_begin; h0 + HD <= c_end; h0 += HD) {
        const int n = n0 + h0;            // first column of this head inside its segment (warp-uniform)
        if (n >= seg_n) break;
        float v[HD];
#pragma unroll
        for (int c0 = 0; c0 < HD; c0 += 32) {
            uint32_t acc[32];
            acc_load32<false, BN>(src, h0 + c0, acc);
#pragma unroll
            for (int j = 0; j < 32; ++j) v[c0 + j] = __uint_as_float(acc[j]);
        }
        if (!live) continue;
        if (bias) {
#pragma unroll
            for (int j8 = 0; j8 < HD; j8 += 8) {
                float bf[8];
                unpack<bf16>(ld_u4(bias + n + j8), bf);
#pragma unroll
                for (int j = 0; j < 8; ++j) v[j8 + j] += bf[j];
            }
        }
#pragma unroll
        for (int j = 0; j < HD; ++j) v[j] = bf16_round(v[j]);
        if (rotate) {
#pragma unroll
            for (int j4 = 0; j4 < kHalf; j4 += 4) {
                const float4 c4 = *reinterpret_cast<const float4*>(cs + j4), s4 = *reinterpret_cast<const float4*>(sn + j4);
                const float cc[4] = {c4.x, c4.y, c4.z, c4.w}, ss[4] = {s4.x, s4.y, s4.z, s4.w};
#pragma unroll
                for (int u = 0; u < 4; ++u) {
                    const int j = j4 + u;
                    const float c = bf16_round(cc[u]), s2 = bf16_round(ss[u]);
                    const float a = v[j], b = v[j + kHalf];
                    v[j] = bf16_round(bf16_round(a * c) + bf16_round(-b * s2));
                    v[j + kHalf] = bf16_round(bf16_round(b * c) + bf16_round(a * s2));
                }
            }
        }
        const int head = n / HD;
        bf16* dst = seg == 0 ? static_cast<bf16*>(p.q_out) + ((size_t)row * p.n_heads + head) * HD
                             : kv_row + ((size_t)head << p.kv.page_shift) * HD;
#pragma unroll
        for (int j = 0; j < HD; j += 8) {
            uint4 w;
            w.x = pack2(v[j], v[j + 1]); w.y = pack2(v[j + 2], v[j + 3]);
            w.z = pack2(v[j + 4], v[j + 5]); w.w = pack2(v[j + 6], v[j + 7]);
            *reinterpret_cast<uint4*>(dst + j) = w;
        }
    }
}

// Persistent: one CTA per SM walks tiles t = blockIdx.x, + gridDim.x, ...  The accumulator is double-buffered in tensor
// memory, so the MMA warp starts tile i+1 while the epilogue warps are still reading tile i out -- the epilogue
// (tcgen05.ld, SiLU / residual arithmetic, stores) no longer sits between two main loops -- and the TMA ring never drains
// at a tile boundary.
// CG = 2: the CTAs come in pairs (cluster of 2, cta_group::2).  A pair owns a 256-row tile; each CTA loads its 128 A rows
// and half of the B tile (BN / 2 weight rows), so a k-block costs a CTA 32 KB of shared-memory fill instead of 48 KB for
// the same 128 x BN x 64 MMA work -- what bounds these kernels is that fill rate (profiles/r2_gemm_m512_ncu.md).  The
// leader issues UMMAs of M = 256 that read both CTAs' shared memory; each CTA's epilogue warps read their own half of
// the accumulator out of their own tensor memory.
template <int EPI, int BN, int CG>
__global__ void __launch_bounds__(kGemmThreads, 1)
gemm_tc_kernel(const __grid_constant__ CUtensorMap tm_a, const __grid_constant__ CUtensorMap tm_b,
               const __grid_constant__ CUtensorMap tm_b2, const __grid_constant__ CUtensorMap tm_b3, const GemmTcParams p,
               const UnitSched us, const int tiles_m) {
    constexpr bool kDual = EPI == GEMM_EPI_SWIGLU;
    static_assert(!(kDual && BN != 128), "the dual-B tile is 128 wide");
    static_assert(CG == 1 || CG == 2, "one CTA or a pair");
    constexpr int kBRows = BN / CG;                        // rows of a B tile this CTA loads
    constexpr int kBTileBytes = kBRows * BK * 2;
    constexpr int kStageBytes = kTileBytes + kBTileBytes * (kDual ? 2 : 1);
    constexpr uint32_t kAccCols = kDual ? 256 : BN;       // columns of one accumulator
    constexpr uint32_t kTmemCols = 2 * kAccCols < 32 ? 32 : 2 * kAccCols;  // two of them
    constexpr uint32_t kIdesc = make_idesc(BN, BM * CG);
    constexpr int kStages = gemm_stages(kStageBytes);
    extern __shared__ __align__(1024) unsigned char smem_raw[];
    unsigned char* smem = reinterpret_cast<unsigned char*>(((uintptr_t)smem_raw + 1023) & ~(uintptr_t)1023);
    __shared__ __align__(8) uint64_t bars[2 * kStages + 5];
    __shared__ uint32_t s_tmem;

    const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
    const int rank = CG == 2 ? (int)cluster_ctarank() : 0;            // 0 = the leader of the pair
    const int unit0 = blockIdx.x / CG, unit_step = gridDim.x / CG;    // both CTAs of a pair walk the same units
    const int num_kb = (p.K + BK - 1) / BK;
    const uint32_t full0 = smem_u32(bars), empty0 = smem_u32(bars + kStages);
    const uint32_t acc_full = smem_u32(bars + 2 * kStages), acc_empty = smem_u32(bars + 2 * kStages + 2);
    const uint32_t red_bar = smem_u32(bars + 2 * kStages + 4);   // the staged slices of a split tile have landed

    if (threadIdx.x == 0) {
        for (int i = 0; i < kStages; ++i) {
            mbar_init(full0 + i * 8, 1);         // (pair: only the leader's is used; it expects both CTAs' bytes)
            mbar_init(empty0 + i * 8, 1);
        }
        for (int i = 0; i < 2; ++i) {
            mbar_init(acc_full + i * 8, 1);
            mbar_init(acc_empty + i * 8, 128 * CG);   // the four epilogue warps (of both CTAs: the leader's MMA thread waits)
        }
        mbar_init(red_bar, 1);
        asm volatile("fence.mbarrier_init.release.cluster;" ::: "memory");
    }
    if (warp == 1) {  // one warp (per CTA) allocates tensor memory and later frees it
        if (CG == 2) {
            asm volatile("tcgen05.alloc.cta_group::2.sync.aligned.shared::cta.b32 [%0], %1;" ::"r"(smem_u32(&s_tmem)), "r"(kTmemCols));
            asm volatile("tcgen05.relinquish_alloc_permit.cta_group::2.sync.aligned;");
        } else {
            asm volatile("tcgen05.alloc.cta_group::1.sync.aligned.shared::cta.b32 [%0], %1;" ::"r"(smem_u32(&s_tmem)), "r"(kTmemCols));
            asm volatile("tcgen05.relinquish_alloc_permit.cta_group::1.sync.aligned;");
        }
    }
    tc_fence_before();
    if (CG == 2) cluster_sync_pair();   // the partner's barriers are initialised before anything signals them
    else __syncthreads();
    tc_fence_after();
    const uint32_t tmem = s_tmem;
    pdl_launch_dependents();
    if (threadIdx.x != 0) pdl_wait();   // (the producer thread waits after it has requested the first weight tiles)

    if (warp == 0) {
        // ===== TMA producer: the ring runs on across tile boundaries =====
        if (lane == 0) {
            // the barrier a stage's loads signal: this CTA's, or (pair) the leader's
            auto full_of = [&](unsigned st) { return CG == 2 ? mapa_u32(full0 + st * 8, 0) : full0 + st * 8; };
            // the leader expects the bytes of both CTAs' loads on its barrier (the partner's complete_tx may come first:
            // the phase cannot complete before the leader's own arrival); no cluster-scope release per stage -- a
            // remote arrive.expect_tx.release.cluster here cost a fence per k-block and halved the kernel's rate
            auto arm = [&](unsigned st) {
                if (rank == 0) mbar_expect_tx(full0 + st * 8, CG * kStageBytes);
            };
            auto load = [&](uint32_t dst, const CUtensorMap* map, int c0, int c1, unsigned st) {
                if (CG == 2) tma_load_2d_pair(dst, map, c0, c1, full_of(st));
                else tma_load_2d(dst, map, c0, c1, full0 + st * 8);
            };
            unsigned it = 0;  // k-blocks issued so far
            bool waited = false;
            for (int u = unit0; u < us.n_units; u += unit_step) {
                const Unit un = unit_of(us, u, num_kb);
                const TileCoord tc = tile_coord<EPI, BN, CG>(p, un.tile, tiles_m);
                const CUtensorMap* map_b = tc.seg == 0 ? &tm_b : (tc.seg == 1 ? &tm_b2 : &tm_b3);
                const int a_row = tc.m0 + rank * BM, b_row = tc.n0 + rank * kBRows;
                int kb = un.kb0;
                if (!waited) {
                    // The first ring-full of WEIGHT tiles is requested before this kernel waits for its predecessor: no
                    // kernel of the step writes weights.  The activation tiles of those stages follow the wait.
                    const int pre = min(kStages, un.kb1 - un.kb0);
                    for (int i = 0; i < pre; ++i) {
                        const uint32_t dst = smem_u32(smem + (size_t)i * kStageBytes);
                        arm(i);
                        load(dst + kTileBytes, map_b, (kb + i) * BK, b_row, i);
                        if (kDual) load(dst + kTileBytes + kBTileBytes, &tm_b2, (kb + i) * BK, b_row, i);
                    }
                    pdl_wait();
                    waited = true;
                    for (int i = 0; i < pre; ++i)
                        load(smem_u32(smem + (size_t)i * kStageBytes), &tm_a, (kb + i) * BK, a_row, i);
                    kb += pre;
                    it += pre;
                }
                for (; kb < un.kb1; ++kb, ++it) {
                    const unsigned s = it % kStages;
                    mbar_wait(empty0 + s * 8, ((it / kStages) & 1) ^ 1);
                    const uint32_t dst = smem_u32(smem + (size_t)s * kStageBytes);
                    arm(s);
                    load(dst, &tm_a, kb * BK, a_row, s);
                    load(dst + kTileBytes, map_b, kb * BK, b_row, s);
                    if (kDual) load(dst + kTileBytes + kBTileBytes, &tm_b2, kb * BK, b_row, s);
                }
            }
        }
    } else if (warp == 1) {
        // ===== MMA issuer (one thread; of a pair, the leader's): tile i+1 goes into the other accumulator while tile i
        // is being read out =====
        if (lane == 0 && rank == 0) {
            unsigned it = 0, lt = 0;
            for (int u = unit0; u < us.n_units; u += unit_step, ++lt) {
                const Unit un = unit_of(us, u, num_kb);
                const uint32_t acc_t = tmem + (lt & 1) * kAccCols;
                if (lt >= 2) {   // the epilogue (of both CTAs) has drained this buffer
                    if (CG == 2) mbar_wait_cluster(acc_empty + (lt & 1) * 8, ((lt >> 1) & 1) ^ 1);
                    else mbar_wait(acc_empty + (lt & 1) * 8, ((lt >> 1) & 1) ^ 1);
                }
                tc_fence_after();
                for (int kb = un.kb0; kb < un.kb1; ++kb, ++it) {
                    const unsigned s = it % kStages;
                    mbar_wait(full0 + s * 8, (it / kStages) & 1);   // TMA-written operands, read by the tensor core only
                    tc_fence_after();
                    const uint32_t base = smem_u32(smem + (size_t)s * kStageBytes);
                    const uint64_t da = make_sw128_desc(base), db = make_sw128_desc(base + kTileBytes);
                    const uint64_t db2 = make_sw128_desc(base + kTileBytes + kBTileBytes);
#pragma unroll
                    for (int k = 0; k < BK / 16; ++k) {
                        // 16 k-elements = 32 bytes further along the swizzled row: +2 in the (address >> 4) field
                        const uint32_t acc = (kb > un.kb0 || k) ? 1u : 0u;
                        if (CG == 2) {
                            tc_mma_pair(acc_t, da + 2 * k, db + 2 * k, kIdesc, acc);
                            if (kDual) tc_mma_pair(acc_t + BN, da + 2 * k, db2 + 2 * k, kIdesc, acc);
                        } else {
                            tc_mma(acc_t, da + 2 * k, db + 2 * k, kIdesc, acc);
                            if (kDual) tc_mma(acc_t + BN, da + 2 * k, db2 + 2 * k, kIdesc, acc);
                        }
                    }
                    // the slot is free once these MMAs have read it (in both CTAs)
                    if (CG == 2) tc_commit_pair(empty0 + s * 8);
                    else tc_commit(empty0 + s * 8);
                }
                if (CG == 2) tc_commit_pair(acc_full + (lt & 1) * 8);  // accumulator complete
                else tc_commit(acc_full + (lt & 1) * 8);
            }
        }
    } else {
        // ===== epilogue: warps 2..5, TMEM lane quarter = warp % 4 =====
        const int quarter = warp & 3;
        // where this CTA's epilogue threads report an accumulator drained: the (leader's) MMA thread's barrier
        auto acc_drained = [&](unsigned buf) {
            if (CG == 2) mbar_arrive_cluster(mapa_u32(acc_empty + buf * 8, 0));
            else mbar_arrive(acc_empty + buf * 8);
        };
        unsigned lt = 0;
        for (int u = unit0; u < us.n_units; u += unit_step, ++lt) {
        const Unit un = unit_of(us, u, num_kb);
        const TileCoord tc = tile_coord<EPI, BN, CG>(p, un.tile, tiles_m);
        const int seg = tc.seg, col0 = tc.col0, seg_n = tc.seg_n, n0 = tc.n0;
        mbar_wait(acc_full + (lt & 1) * 8, (lt >> 1) & 1);
        tc_fence_after();
        const int row = tc.m0 + rank * BM + quarter * 32 + lane;
        const uint32_t tbase = tmem + (lt & 1) * kAccCols + ((uint32_t)(quarter * 32) << 16);
        AccSrc src{tbase, nullptr, quarter * 32 + lane, 1, 0};
        int c_begin = 0, c_end = BN;   // the output columns of the tile this CTA's epilogue covers
        if (un.tail >= 0) {
            // a K slice of a split tile (this CTA's last unit): the accumulator goes to the scratch area as it is; then
            // wait for the tile's other slices and take this slice's share of the columns
            constexpr size_t kSlice = (size_t)BM * kAccCols;
            const int tt = (un.tail / us.splits) * CG + rank;   // this CTA's half of the split tile is a tile of its own here
            float* mine = p.split_ws + ((size_t)tt * us.splits + un.slice) * kSlice;
#pragma unroll 1
            for (int c0 = 0; c0 < (int)kAccCols; c0 += 32) {
                uint32_t acc[32];
                tc_ld32(tbase + c0, acc);
#pragma unroll
                for (int j4 = 0; j4 < 8; ++j4)
                    *reinterpret_cast<uint4*>(mine + ws_offset(c0 >> 5, j4, src.row)) =
                        make_uint4(acc[4 * j4], acc[4 * j4 + 1], acc[4 * j4 + 2], acc[4 * j4 + 3]);
            }
            tc_fence_before();
            acc_drained(lt & 1);
            __threadfence();
            fence_proxy_async_global();   // the slices are read back by bulk copies (async proxy)
            asm volatile("bar.sync 1, 128;" ::: "memory");   // the four epilogue warps: every row of the slice is out
            // shares are whole heads for the q|k|v epilogue, 32-column chunks otherwise
            const int gcols = EPI == GEMM_EPI_QKV ? p.head_dim : 32, n_g = BN / gcols;
            c_begin = (un.slice * n_g / us.splits) * gcols;
            c_end = ((un.slice + 1) * n_g / us.splits) * gcols;
            if (threadIdx.x == 64) {
                unsigned* arrived = p.split_tickets + 2 * tt;
                atomicAdd(arrived, 1u);
                while (ld_acquire_gpu(arrived) < (unsigned)us.splits) __nanosleep(32);
                fence_proxy_async_global();
                // every slice of the tile is in the scratch area: this CTA's chunks of all of them into the operand ring
                // (idle: this unit's MMAs have completed and no further loads were issued)
                const float* tile_ws = p.split_ws + (size_t)tt * us.splits * kSlice;
                const int n_out = (c_end - c_begin) >> 5, per_out = kDual ? 2 : 1;
                if (n_out > 0) mbar_expect_tx(red_bar, (uint32_t)(n_out * per_out * us.splits * kRedBlockBytes));
                for (int oc = 0; oc < n_out; ++oc)
                    for (int h = 0; h < per_out; ++h) {
                        const int cc = (c_begin >> 5) + oc + h * (BN / 32);   // accumulator chunk (the up half follows the gate half)
                        for (int sl = 0; sl < us.splits; ++sl)
                            bulk_load(smem_u32(smem) + (uint32_t)(((oc * per_out + h) * us.splits + sl) * kRedBlockBytes),
                                      tile_ws + (size_t)sl * kSlice + ws_offset(cc, 0, 0), kRedBlockBytes, red_bar);
                    }
            }
            if (c_end > c_begin) mbar_wait(red_bar, 0);
            src.red = smem;
            src.splits = us.splits;
            src.first_chunk = c_begin >> 5;
        }
        bf16* out = static_cast<bf16*>(p.c) + col0;
        const bf16* bias = static_cast<const bf16*>(seg == 0 ? p.bias : p.bias_more[seg - 1]);
        if (EPI == GEMM_EPI_QKV) {
            if (p.head_dim == 64) qkv_tail<64, BN>(p, src, row, seg, n0, seg_n, bias, c_begin, c_end);
            else qkv_tail<128, BN>(p, src, row, seg, n0, seg_n, bias, c_begin, c_end);
        } else {
        // A thread owns one output row and, per tcgen05.ld, 32 consecutive columns of it: 64 bytes of bf16.  They leave
        // as four 16-byte stores (and the residual / bias arrive as 16-byte loads): every store fills half a sector,
        // where one bf16 pair per instruction touched 32 rows with 4 bytes each.
        const bool vec_ok = (p.ldc % 8 == 0) && (col0 % 8 == 0) && ((reinterpret_cast<uintptr_t>(p.c) & 15) == 0) &&
                            (!bias || (reinterpret_cast<uintptr_t>(bias) & 15) == 0);
#pragma unroll 1
        for (int c0 = c_begin; c0 < c_end; c0 += 32) {
            uint32_t acc[32], acc2[32];
            acc_load32<kDual, BN>(src, c0, acc);
            if (kDual) acc_load32<kDual, BN>(src, BN + c0, acc2);
            if (row < p.M) {
#pragma unroll
                for (int j8 = 0; j8 < 32; j8 += 8) {
                    const int n8 = n0 + c0 + j8;
                    if (n8 >= seg_n) break;
                    bf16* dst = out + (size_t)row * p.ldc + n8;
                    if (vec_ok && n8 + 8 <= seg_n) {
                        float v[8];
#pragma unroll
                        for (int j = 0; j < 8; ++j) v[j] = __uint_as_float(acc[j8 + j]);
                        if (EPI == GEMM_EPI_SWIGLU) {
#pragma unroll
                            for (int j = 0; j < 8; ++j)
                                v[j] = bf16_round(silu_f(bf16_round(v[j]))) * bf16_round(__uint_as_float(acc2[j8 + j]));
                        } else {
                            if (bias) {
                                float bf[8];
                                unpack<bf16>(ld_u4(bias + n8), bf);
#pragma unroll
                                for (int j = 0; j < 8; ++j) v[j] += bf[j];
                            }
#pragma unroll
                            for (int j = 0; j < 8; ++j) v[j] = bf16_round(v[j]);
                            if (EPI == GEMM_EPI_RESIDUAL) {
                                float of[8];
                                unpack<bf16>(*reinterpret_cast<const uint4*>(dst), of);
#pragma unroll
                                for (int j = 0; j < 8; ++j) v[j] += of[j];
                            } else if (EPI == GEMM_EPI_GELU) {
#pragma unroll
                                for (int j = 0; j < 8; ++j) v[j] = gelu_new_f(v[j]);
                            }
                        }
                        uint4 w;
                        w.x = pack2(v[0], v[1]); w.y = pack2(v[2], v[3]); w.z = pack2(v[4], v[5]); w.w = pack2(v[6], v[7]);
                        *reinterpret_cast<uint4*>(dst) = w;
                        continue;
                    }
#pragma unroll
                    for (int j = j8; j < j8 + 8; j += 2) {   // ragged edge / unaligned output: pairs
                        const int n = n0 + c0 + j;
                        if (n >= seg_n) break;
                        float v0 = __uint_as_float(acc[j]), v1 = __uint_as_float(acc[j + 1]);
                        if (EPI == GEMM_EPI_SWIGLU) {
                            const float g0 = bf16_round(v0), g1 = bf16_round(v1);
                            const float u0 = bf16_round(__uint_as_float(acc2[j])), u1 = bf16_round(__uint_as_float(acc2[j + 1]));
                            v0 = bf16_round(silu_f(g0)) * u0;
                            v1 = bf16_round(silu_f(g1)) * u1;
                        } else {
                            v0 = bf16_round(v0 + (bias ? __bfloat162float(bias[n]) : 0.f));
                            v1 = bf16_round(v1 + (bias && n + 1 < seg_n ? __bfloat162float(bias[n + 1]) : 0.f));
                            if (EPI == GEMM_EPI_RESIDUAL) {
                                v0 += __bfloat162float(out[(size_t)row * p.ldc + n]);
                                if (n + 1 < seg_n) v1 += __bfloat162float(out[(size_t)row * p.ldc + n + 1]);
                            } else if (EPI == GEMM_EPI_GELU) {
                                v0 = gelu_new_f(v0);
                                v1 = gelu_new_f(v1);
                            }
                        }
                        out[(size_t)row * p.ldc + n] = __float2bfloat16_rn(v0);
                        if (n + 1 < seg_n) out[(size_t)row * p.ldc + n + 1] = __float2bfloat16_rn(v1);
                    }
                }
            }
        }
        }  // generic epilogues
        if (un.tail < 0) {
            tc_fence_before();
            acc_drained(lt & 1);   // this accumulator may be overwritten (by the tile after next)
        } else {
            asm volatile("bar.sync 1, 128;" ::: "memory");   // this CTA has read what it needs of the tile's slices
            if (threadIdx.x == 64) {
                unsigned* t = p.split_tickets + 2 * ((un.tail / us.splits) * CG + rank);
                if (atomicAdd(t + 1, 1u) == (unsigned)us.splits - 1) {   // the last CTA to leave the tile: tickets back to zero
                    t[0] = 0u;
                    t[1] = 0u;
                }
            }
        }
        }  // tiles
    }
    tc_fence_before();
    if (CG == 2) cluster_sync_pair();   // the leader's MMAs read the partner's shared memory until the very end
    else __syncthreads();
    if (warp == 1) {
        tc_fence_after();
        if (CG == 2) asm volatile("tcgen05.dealloc.cta_group::2.sync.aligned.b32 %0, %1;" ::"r"(tmem), "r"(kTmemCols));
        else asm volatile("tcgen05.dealloc.cta_group::1.sync.aligned.b32 %0, %1;" ::"r"(tmem), "r"(kTmemCols));
    }
}

// ---- decode batches of 5..64 rows: swap-AB, split-K across a cluster ----------------------------------------------
// With a handful of rows the contraction is weight-streaming work again: what matters is that every SM keeps many
// 16 KB weight tiles in flight.  The weights take the M side of the UMMA (128 weight rows per CTA), the batch rows
// the N side (16 / 32 / 64), so a stage is one weight tile (two for gate/up) plus a 2-8 KB row tile and the ring is
// 5-10 stages deep.  Narrow matrices are split along K over the CTAs of a cluster (<= 8): every CTA reads its
// accumulator from tensor memory and stores it into the leader's shared memory (DSMEM), the leader adds the
// pieces in rank order -- the same sum on every run -- and applies the epilogue.
__device__ __forceinline__ uint32_t cluster_rank() {
    uint32_t r;
    asm volatile("mov.u32 %0, %%cluster_ctarank;" : "=r"(r));
    return r;
}
__device__ __forceinline__ void cluster_sync_all() {
    asm volatile("barrier.cluster.arrive.release.aligned;\n\tbarrier.cluster.wait.acquire.aligned;" ::: "memory");
}
__device__ __forceinline__ void st_dsmem_f32(uint32_t local_addr, uint32_t rank, float v) {
    uint32_t remote;
    asm volatile("mapa.shared::cluster.u32 %0, %1, %2;" : "=r"(remote) : "r"(local_addr), "r"(rank));
    asm volatile("st.shared::cluster.f32 [%0], %1;" ::"r"(remote), "f"(v) : "memory");
}

template <int EPI, int MP>
struct SwapAbCfg {
    static constexpr bool kDual = EPI == GEMM_EPI_SWIGLU;
    static constexpr int kCols = (kDual ? 2 : 1) * MP;                 // accumulator columns (fp32)
    static constexpr int kXBytes = MP * BK * 2;
    static constexpr int kStageBytes = kTileBytes * (kDual ? 2 : 1) + kXBytes;
    static constexpr int kStagesS = (200 * 1024 / kStageBytes) > 10 ? 10 : (200 * 1024 / kStageBytes);
    static constexpr uint32_t kTmemCols = kCols <= 32 ? 32 : (kCols <= 64 ? 64 : 128);
    static constexpr int kMaxSplit = 256 / kCols > 8 ? 8 : 256 / kCols;  // partial buffer (split x cols x 512 B) fits the ring
};

template <int EPI, int MP>
__global__ void __launch_bounds__(kGemmThreads, 1)
gemm_swapab_kernel(const __grid_constant__ CUtensorMap tm_w, const __grid_constant__ CUtensorMap tm_w2,
                   const __grid_constant__ CUtensorMap tm_w3, const __grid_constant__ CUtensorMap tm_x,
                   const GemmTcParams p, const int kb_per_cta) {
    using Cfg = SwapAbCfg<EPI, MP>;
    constexpr bool kDual = Cfg::kDual;
    constexpr int kCols = Cfg::kCols, kStageBytes = Cfg::kStageBytes, kSt = Cfg::kStagesS;
    constexpr uint32_t kIdesc = make_idesc(MP);  // M = 128 weight rows, N = MP batch rows
    extern __shared__ __align__(1024) unsigned char smem_raw[];
    unsigned char* smem = reinterpret_cast<unsigned char*>(((uintptr_t)smem_raw + 1023) & ~(uintptr_t)1023);
    __shared__ __align__(8) uint64_t bars[2 * kSt + 1];
    __shared__ uint32_t s_tmem;

    const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
    // row segment of the logical weight matrix this tile belongs to (q | k | v): its tensor map, bias, columns
    int seg = 0, tile = blockIdx.x, col0 = 0, seg_n = p.N;
    if (!kDual) {
        const int t0 = (p.N + BM - 1) / BM, t1 = (p.n_more[0] + BM - 1) / BM;
        if (tile >= t0 + t1) { seg = 2; tile -= t0 + t1; col0 = p.N + p.n_more[0]; seg_n = p.n_more[1]; }
        else if (tile >= t0) { seg = 1; tile -= t0; col0 = p.N; seg_n = p.n_more[0]; }
    }
    const CUtensorMap* map_w = seg == 0 ? &tm_w : (seg == 1 ? &tm_w2 : &tm_w3);
    const int n0 = tile * BM;
    const int ksplit = gridDim.y, krank = (int)cluster_rank();   // the cluster spans the whole y dimension
    const int total_kb = (p.K + BK - 1) / BK;
    const int kb0 = krank * kb_per_cta;
    const int num_kb = max(0, min(kb_per_cta, total_kb - kb0));
    const uint32_t full0 = smem_u32(bars), empty0 = smem_u32(bars + kSt), tfull = smem_u32(bars + 2 * kSt);

    if (threadIdx.x == 0) {
        for (int i = 0; i < kSt; ++i) {
            mbar_init(full0 + i * 8, 1);
            mbar_init(empty0 + i * 8, 1);
        }
        mbar_init(tfull, 1);
        asm volatile("fence.mbarrier_init.release.cluster;" ::: "memory");
    }
    if (warp == 1) {
        asm volatile("tcgen05.alloc.cta_group::1.sync.aligned.shared::cta.b32 [%0], %1;" ::"r"(smem_u32(&s_tmem)), "r"(Cfg::kTmemCols));
        asm volatile("tcgen05.relinquish_alloc_permit.cta_group::1.sync.aligned;");
    }
    tc_fence_before();
    __syncthreads();
    tc_fence_after();
    const uint32_t tmem = s_tmem;
    pdl_launch_dependents();
    if (threadIdx.x != 0) pdl_wait();   // (the producer thread waits after it has requested the first weight tiles)
    float resid[EPI == GEMM_EPI_RESIDUAL ? MP : 1];

    if (warp == 0) {
        if (lane == 0) {
            // The first ring-full of WEIGHT tiles is requested before this kernel waits for its predecessor (no kernel of
            // the step writes weights): at a handful of rows the projection is weight streaming, and this way the stream
            // starts while the previous kernel is still draining.  The row tiles of those stages follow the wait.
            const int pre = min(kSt, num_kb);
            for (int kb = 0; kb < pre; ++kb) {
                const uint32_t dst = smem_u32(smem + (size_t)kb * kStageBytes);
                mbar_expect_tx(full0 + kb * 8, kStageBytes);
                tma_load_2d(dst, map_w, (kb0 + kb) * BK, n0, full0 + kb * 8);
                if (kDual) tma_load_2d(dst + kTileBytes, &tm_w2, (kb0 + kb) * BK, n0, full0 + kb * 8);
            }
            pdl_wait();
            for (int kb = 0; kb < pre; ++kb)
                tma_load_2d(smem_u32(smem + (size_t)kb * kStageBytes) + kTileBytes * (kDual ? 2 : 1), &tm_x, (kb0 + kb) * BK, 0, full0 + kb * 8);
            for (int kb = pre; kb < num_kb; ++kb) {
                const int s = kb % kSt;
                mbar_wait(empty0 + s * 8, ((kb / kSt) & 1) ^ 1);
                const uint32_t dst = smem_u32(smem + (size_t)s * kStageBytes);
                mbar_expect_tx(full0 + s * 8, kStageBytes);
                tma_load_2d(dst, map_w, (kb0 + kb) * BK, n0, full0 + s * 8);
                if (kDual) tma_load_2d(dst + kTileBytes, &tm_w2, (kb0 + kb) * BK, n0, full0 + s * 8);
                tma_load_2d(dst + kTileBytes * (kDual ? 2 : 1), &tm_x, (kb0 + kb) * BK, 0, full0 + s * 8);
            }
        }
    } else if (warp == 1) {
        if (lane == 0 && num_kb > 0) {
            for (int kb = 0; kb < num_kb; ++kb) {
                const int s = kb % kSt;
                mbar_wait(full0 + s * 8, (kb / kSt) & 1);
                tc_fence_after();
                const uint32_t base = smem_u32(smem + (size_t)s * kStageBytes);
                const uint64_t dw = make_sw128_desc(base), dw2 = make_sw128_desc(base + kTileBytes);
                const uint64_t dx = make_sw128_desc(base + kTileBytes * (kDual ? 2 : 1));
#pragma unroll
                for (int k = 0; k < BK / 16; ++k) {
                    const uint32_t acc = (kb | k) ? 1u : 0u;
                    tc_mma(tmem, dw + 2 * k, dx + 2 * k, kIdesc, acc);
                    if (kDual) tc_mma(tmem + MP, dw2 + 2 * k, dx + 2 * k, kIdesc, acc);
                }
                tc_commit(empty0 + s * 8);
            }
            tc_commit(tfull);
        }
    } else {
        if (EPI == GEMM_EPI_RESIDUAL && krank == 0) {
            // the residual values this thread will add to -- one per batch row, the same output column -- requested
            // now, all at once: they arrive while the weights stream, instead of one dependent round trip per row in
            // the epilogue (the epilogue's own stores are the only writes to these addresses)
            const int n = n0 + (warp & 3) * 32 + lane;
            const bf16* y = static_cast<const bf16*>(p.c) + col0 + n;
#pragma unroll
            for (int j = 0; j < MP; ++j) resid[j] = (n < seg_n && j < p.M) ? __bfloat162float(y[(size_t)j * p.ldc]) : 0.f;
        }
        if (num_kb > 0) {
            mbar_wait(tfull, 0);  // every MMA of this CTA has read its operands: the ring is free
            tc_fence_after();
        }
    }
    // 1st cluster barrier: every CTA's ring is idle, so the leader's can take the partial sums
    cluster_sync_all();
    float* part = reinterpret_cast<float*>(smem);  // leader: [krank][column][128 weight rows]
    if (warp >= 2) {
        const int quarter = warp & 3, nl = quarter * 32 + lane;
        const uint32_t tbase = tmem + ((uint32_t)(quarter * 32) << 16);
#pragma unroll 1
        for (int c0 = 0; c0 < kCols; c0 += 32) {
            uint32_t acc[32];
            if (num_kb > 0) {
                tc_ld32(tbase + c0, acc);
            } else {
#pragma unroll
                for (int j = 0; j < 32; ++j) acc[j] = 0u;
            }
#pragma unroll
            for (int j = 0; j < 32; ++j)
                if (c0 + j < kCols)
                    st_dsmem_f32(smem_u32(part + ((size_t)krank * kCols + c0 + j) * BM + nl), 0u, __uint_as_float(acc[j]));
        }
        tc_fence_before();
    }
    // 2nd cluster barrier: all pieces are in the leader's shared memory
    cluster_sync_all();
    if (krank == 0 && warp >= 2) {
        const int quarter = warp & 3, nl = quarter * 32 + lane, n = n0 + nl;
        bf16* out = static_cast<bf16*>(p.c) + col0;
        const bf16* bias = static_cast<const bf16*>(seg == 0 ? p.bias : p.bias_more[seg - 1]);
        if (n < seg_n) {
            const float bn = (bias && EPI != GEMM_EPI_SWIGLU) ? __bfloat162float(bias[n]) : 0.f;
#pragma unroll
            for (int j = 0; j < MP; ++j) {
                if (j >= p.M) break;
                float v = 0.f, u = 0.f;
                for (int k = 0; k < ksplit; ++k) {
                    v += part[((size_t)k * kCols + j) * BM + nl];
                    if (kDual) u += part[((size_t)k * kCols + MP + j) * BM + nl];
                }
                if (EPI == GEMM_EPI_LOGITS) {  // fp32 row of logits, in the model dtype's precision
                    static_cast<float*>(p.c)[(size_t)j * p.ldc + col0 + n] = bf16_round(v);
                    continue;
                }
                bf16* y = out + (size_t)j * p.ldc + n;
                if (EPI == GEMM_EPI_SWIGLU) {
                    v = bf16_round(silu_f(bf16_round(v))) * bf16_round(u);
                } else {
                    v = bf16_round(v + bn);
                    if (EPI == GEMM_EPI_RESIDUAL) v += resid[j];
                    else if (EPI == GEMM_EPI_GELU) v = gelu_new_f(v);
                }
                *y = __float2bfloat16_rn(v);
            }
        }
    }
    __syncthreads();
    if (warp == 1) {
        tc_fence_after();
        asm volatile("tcgen05.dealloc.cta_group::1.sync.aligned.b32 %0, %1;" ::"r"(tmem), "r"(Cfg::kTmemCols));
    }
}

// ---- small row-wise kernels around the GEMMs -------------------------------------------------------------------

// xn[r] = norm(x[r]) for T rows (RMSNorm / LayerNorm with the rounding of the decode prologue); one warp per row
__global__ void __launch_bounds__(256) norm_rows_kernel(const bf16* __restrict__ x, const bf16* __restrict__ w,
                                                        const bf16* __restrict__ b, bf16* __restrict__ y, int rows, int K,
                                                        int norm, float eps) {
    pdl_launch_dependents();   // the next kernel of the step may be placed as soon as every CTA of this one runs
    pdl_wait();                // ... and this one touches global memory only once its predecessor has completed
    const int row = blockIdx.x * 8 + (threadIdx.x >> 5), lane = threadIdx.x & 31;
    if (row >= rows) return;
    const bf16* xr = x + (size_t)row * K;
    bf16* yr = y + (size_t)row * K;
    float s1 = 0.f, s2 = 0.f;
    for (int k = lane * 8; k < K; k += 256) {
        float f[8];
        unpack<bf16>(ld_u4(xr + k), f);
#pragma unroll
        for (int j = 0; j < 8; ++j) {
            s1 += f[j];
            s2 = fmaf(f[j], f[j], s2);
        }
    }
    float mean = 0.f, rstd;
    if (norm == NORM_RMS) {
        rstd = 1.0f / sqrtf(warp_sum(s2) / (float)K + eps);
    } else {
        mean = warp_sum(s1) / (float)K;
        float sq = 0.f;
        for (int k = lane * 8; k < K; k += 256) {
            float f[8];
            unpack<bf16>(ld_u4(xr + k), f);
#pragma unroll
            for (int j = 0; j < 8; ++j) sq = fmaf(f[j] - mean, f[j] - mean, sq);
        }
        rstd = 1.0f / sqrtf(warp_sum(sq) / (float)K + eps);
    }
    for (int k = lane * 8; k < K; k += 256) {
        float f[8], wf[8], bf[8];
        unpack<bf16>(ld_u4(xr + k), f);
        unpack<bf16>(ld_u4(w + k), wf);
        if (norm == NORM_RMS) {
#pragma unroll
            for (int j = 0; j < 8; ++j) f[j] = wf[j] * bf16_round(f[j] * rstd);
        } else {
            unpack<bf16>(ld_u4(b + k), bf);
#pragma unroll
            for (int j = 0; j < 8; ++j) f[j] = fmaf((f[j] - mean) * rstd, wf[j], bf[j]);
        }
        uint4 o;
        __nv_bfloat162 a0 = __floats2bfloat162_rn(f[0], f[1]), a1 = __floats2bfloat162_rn(f[2], f[3]);
        __nv_bfloat162 a2 = __floats2bfloat162_rn(f[4], f[5]), a3 = __floats2bfloat162_rn(f[6], f[7]);
        o.x = *reinterpret_cast<uint32_t*>(&a0);
        o.y = *reinterpret_cast<uint32_t*>(&a1);
        o.z = *reinterpret_cast<uint32_t*>(&a2);
        o.w = *reinterpret_cast<uint32_t*>(&a3);
        *reinterpret_cast<uint4*>(yr + k) = o;
    }
}

// The same for a handful of rows (decode batches): one CTA per row, the row stays in registers (K <= 16384), and the
// rows may be gathered (row_map) -- the final norm in front of the lm_head GEMM reads the rows whose logits are wanted.
__global__ void __launch_bounds__(256) norm_rows_cta_kernel(const bf16* __restrict__ x, const bf16* __restrict__ w,
                                                            const bf16* __restrict__ b, bf16* __restrict__ y,
                                                            const int32_t* __restrict__ row_map, int K, int norm, float eps) {
    pdl_launch_dependents();   // the next kernel of the step may be placed as soon as every CTA of this one runs
    pdl_wait();                // ... and this one touches global memory only once its predecessor has completed
    __shared__ float red[32];
    const int row = blockIdx.x, src = row_map ? row_map[row] : row;
    const bf16* xr = x + (size_t)src * K;
    bf16* yr = y + (size_t)row * K;
    constexpr int kMaxV = 8;  // 8 vectors of 8 per thread: K <= 16384
    float f[kMaxV][8];
    float s1 = 0.f, s2 = 0.f;
#pragma unroll
    for (int i = 0; i < kMaxV; ++i) {
        const int k = (i * 256 + threadIdx.x) * 8;
        if (k < K) {
            unpack<bf16>(ld_u4(xr + k), f[i]);
#pragma unroll
            for (int j = 0; j < 8; ++j) {
                s1 += f[i][j];
                s2 = fmaf(f[i][j], f[i][j], s2);
            }
        }
    }
    float mean = 0.f, rstd;
    if (norm == NORM_RMS) {
        rstd = 1.0f / sqrtf(block_sum(s2, red) / (float)K + eps);
    } else {
        mean = block_sum(s1, red) / (float)K;
        float sq = 0.f;
#pragma unroll
        for (int i = 0; i < kMaxV; ++i)
            if ((i * 256 + (int)threadIdx.x) * 8 < K) {
#pragma unroll
                for (int j = 0; j < 8; ++j) sq = fmaf(f[i][j] - mean, f[i][j] - mean, sq);
            }
        rstd = 1.0f / sqrtf(block_sum(sq, red) / (float)K + eps);
    }
#pragma unroll
    for (int i = 0; i < kMaxV; ++i) {
        const int k = (i * 256 + threadIdx.x) * 8;
        if (k < K) {
            float wf[8], bf[8], o8[8];
            unpack<bf16>(ld_u4(w + k), wf);
            if (norm == NORM_RMS) {
#pragma unroll
                for (int j = 0; j < 8; ++j) o8[j] = wf[j] * bf16_round(f[i][j] * rstd);
            } else {
                unpack<bf16>(ld_u4(b + k), bf);
#pragma unroll
                for (int j = 0; j < 8; ++j) o8[j] = fmaf((f[i][j] - mean) * rstd, wf[j], bf[j]);
            }
            uint4 o;
            __nv_bfloat162 a0 = __floats2bfloat162_rn(o8[0], o8[1]), a1 = __floats2bfloat162_rn(o8[2], o8[3]);
            __nv_bfloat162 a2 = __floats2bfloat162_rn(o8[4], o8[5]), a3 = __floats2bfloat162_rn(o8[6], o8[7]);
            o.x = *reinterpret_cast<uint32_t*>(&a0);
            o.y = *reinterpret_cast<uint32_t*>(&a1);
            o.z = *reinterpret_cast<uint32_t*>(&a2);
            o.w = *reinterpret_cast<uint32_t*>(&a3);
            *reinterpret_cast<uint4*>(yr + k) = o;
        }
    }
}

// next_token[r] = lowest index of the maximum of a row of fp32 logits (torch.argmax semantics); one CTA per row
__global__ void __launch_bounds__(256) argmax_rows_kernel(const float* __restrict__ logits, int n, int32_t* __restrict__ next_token) {
    pdl_launch_dependents();   // the next kernel of the step may be placed as soon as every CTA of this one runs
    pdl_wait();                // ... and this one touches global memory only once its predecessor has completed
    __shared__ float s_val[8];
    __shared__ int s_idx[8];
    const float* row = logits + (size_t)blockIdx.x * n;
    float best = -FLT_MAX;
    int bi = 0x7fffffff;
    for (int i = threadIdx.x; i < n; i += 256) {
        const float v = row[i];
        if (v > best || (v == best && i < bi)) {
            best = v;
            bi = i;
        }
    }
#pragma unroll
    for (int o = 16; o > 0; o >>= 1) {
        const float ov = __shfl_xor_sync(0xffffffffu, best, o);
        const int oi = __shfl_xor_sync(0xffffffffu, bi, o);
        if (ov > best || (ov == best && oi < bi)) {
            best = ov;
            bi = oi;
        }
    }
    if ((threadIdx.x & 31) == 0) {
        s_val[threadIdx.x >> 5] = best;
        s_idx[threadIdx.x >> 5] = bi;
    }
    __syncthreads();
    if (threadIdx.x == 0) {
        for (int wv = 1; wv < 8; ++wv)
            if (s_val[wv] > best || (s_val[wv] == best && s_idx[wv] < bi)) {
                best = s_val[wv];
                bi = s_idx[wv];
            }
        next_token[blockIdx.x] = bi;
    }
}

// qkv [T, (H+2KV)*hd] (already bias-added and rounded) -> rotary on q and k, q to its buffer, k/v into the pages
__global__ void __launch_bounds__(256) rope_append_kernel(const RopeAppendParams p) {
    pdl_launch_dependents();   // the next kernel of the step may be placed as soon as every CTA of this one runs
    pdl_wait();                // ... and this one touches global memory only once its predecessor has completed
    const int hd = p.head_dim, half = hd >> 1, H = p.n_heads, KV = p.n_kv_heads;
    const int heads = H + 2 * KV;
    const int per_row = heads * half;
    for (int i = blockIdx.x * blockDim.x + threadIdx.x; i < p.n_rows * per_row; i += gridDim.x * blockDim.x) {
        const int r = i / per_row, u = i - r * per_row;
        const int head = u / half, j = u - head * half;
        const bf16* src = static_cast<const bf16*>(p.qkv) + (size_t)r * heads * hd + (size_t)head * hd + j;
        float v0 = __bfloat162float(src[0]), v1 = __bfloat162float(src[half]);
        const int m = head < H ? 0 : (head < H + KV ? 1 : 2);
        const int hl = head - (m == 0 ? 0 : (m == 1 ? H : H + KV));
        const int pos = p.positions[r];
        if (p.rope && m != 2) {
            const float c = bf16_round(p.rope_cos[(size_t)pos * half + j]), s = bf16_round(p.rope_sin[(size_t)pos * half + j]);
            const float o0 = bf16_round(bf16_round(v0 * c) + bf16_round(-v1 * s));
            const float o1 = bf16_round(bf16_round(v1 * c) + bf16_round(v0 * s));
            v0 = o0;
            v1 = o1;
        }
        if (m == 0) {
            bf16* q = static_cast<bf16*>(p.q) + (size_t)r * H * hd + hl * hd + j;
            q[0] = __float2bfloat16_rn(v0);
            q[half] = __float2bfloat16_rn(v1);
        } else {
            const int seq = p.seq_ids[r];
            const int page = p.kv.block_table[(size_t)seq * p.kv.pages_per_seq + (pos >> p.kv.page_shift)];
            const int off = pos & ((1 << p.kv.page_shift) - 1);
            bf16* pool = static_cast<bf16*>(m == 1 ? p.kv.kpool : p.kv.vpool);
            bf16* dst = pool + ((((size_t)page * KV + hl) << p.kv.page_shift) + off) * hd + j;
            dst[0] = __float2bfloat16_rn(v0);
            dst[half] = __float2bfloat16_rn(v1);
        }
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

// [rows, K] row-major bf16, box = box_rows (128 or 256) rows x 64 k, 128-byte swizzle, out-of-bounds reads return zero
bool make_map(CUtensorMap* map, const void* base, int rows, int K, int box_rows = BM) {
    EncodeTiledFn enc = get_encode();
    if (!enc) return false;
    cuuint64_t dims[2] = {(cuuint64_t)K, (cuuint64_t)rows};
    cuuint64_t strides[1] = {(cuuint64_t)K * 2};
    cuuint32_t box[2] = {BK, (cuuint32_t)box_rows};
    cuuint32_t estr[2] = {1, 1};
    return enc(map, CU_TENSOR_MAP_DATA_TYPE_BFLOAT16, 2, const_cast<void*>(base), dims, strides, box, estr,
               CU_TENSOR_MAP_INTERLEAVE_NONE, CU_TENSOR_MAP_SWIZZLE_128B, CU_TENSOR_MAP_L2_PROMOTION_L2_256B,
               CU_TENSOR_MAP_FLOAT_OOB_FILL_NONE) == CUDA_SUCCESS;
}

int sm_count() {   // of the current device
    static int sms[64] = {0};
    int dev = 0;
    cudaGetDevice(&dev);
    if (dev >= 0 && dev < 64 && !sms[dev]) cudaDeviceGetAttribute(&sms[dev], cudaDevAttrMultiProcessorCount, dev);
    return (dev >= 0 && dev < 64 && sms[dev] > 0) ? sms[dev] : 148;
}

bool gemm_split_enabled() {
    const char* env = getenv("DLLM_GEMM_SPLIT");   // read per call: the tests compare split and whole tiles in one process
    return !(env && env[0] == '0');
}

bool gemm_pair_enabled() {
    const char* env = getenv("DLLM_GEMM_CG2");   // "0": one CTA per tile everywhere (A/B measurements, tests)
    return !(env && env[0] == '0');
}

// CG = 2: CTA pairs (clusters of 2) on 256-row tiles; the tensor maps of the weights then have boxes of BN / 2 rows.
template <int EPI, int BN, int CG>
cudaError_t launch_epi(const CUtensorMap& a, const CUtensorMap& b, const CUtensorMap& b2, const CUtensorMap& b3, const GemmTcParams& p,
                       cudaStream_t st) {
    constexpr int stage = kTileBytes + (BN / CG) * BK * 2 * (EPI == GEMM_EPI_SWIGLU ? 2 : 1);
    const size_t smem = (size_t)gemm_stages(stage) * stage + 1024;
    static PerDeviceOnce attr_once;
    if (attr_once.need()) {
        cudaError_t e = cudaFuncSetAttribute(gemm_tc_kernel<EPI, BN, CG>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem);
        if (e != cudaSuccess) return e;
    }
    int tiles = (p.N + BN - 1) / BN;
    if (EPI != GEMM_EPI_SWIGLU) tiles += (p.n_more[0] + BN - 1) / BN + (p.n_more[1] + BN - 1) / BN;
    const int tiles_m = (p.M + BM * CG - 1) / (BM * CG), n_tiles = tiles * tiles_m;
    // how many CTAs (CG = 2: pairs, each on one TPC) run at once
    int n_slots = sm_count() / CG;
    if (CG == 2) {
        static int pairs[64] = {0};
        int dev = 0;
        cudaGetDevice(&dev);
        if (dev >= 0 && dev < 64 && !pairs[dev]) {
            cudaLaunchConfig_t q{};
            q.gridDim = dim3(2 * n_slots);
            q.blockDim = dim3(kGemmThreads);
            q.dynamicSmemBytes = smem;
            cudaLaunchAttribute qa[1];
            qa[0].id = cudaLaunchAttributeClusterDimension;
            qa[0].val.clusterDim.x = 2; qa[0].val.clusterDim.y = 1; qa[0].val.clusterDim.z = 1;
            q.attrs = qa;
            q.numAttrs = 1;
            int n = 0;
            if (cudaOccupancyMaxActiveClusters(&n, gemm_tc_kernel<EPI, BN, CG>, &q) != cudaSuccess || n <= 0) {
                cudaGetLastError();
                return cudaErrorNotSupported;
            }
            pairs[dev] = n;
        }
        if (dev >= 0 && dev < 64 && pairs[dev] > 0 && pairs[dev] < n_slots) n_slots = pairs[dev];
    }
    // the last, partly filled wave is cut along K (see UnitSched)
    constexpr size_t kSliceBytes = (size_t)BM * (EPI == GEMM_EPI_SWIGLU ? 256 : BN) * 4;
    const int num_kb = (p.K + BK - 1) / BK;
    UnitSched us{n_tiles, n_tiles, 1, num_kb};
    const int rem = n_tiles % n_slots;
    if (p.split_ws && p.split_tickets && rem > 0 && 2 * rem * CG <= kGemmSplitTickets && gemm_split_enabled()) {
        int s = n_slots / rem;
        if (s > num_kb / 4) s = num_kb / 4;   // a slice keeps at least four k-blocks
        if (s > 8) s = 8;
        // a CTA's share of the tile (its column groups x every slice) must fit the operand ring as 16 KB blocks
        const int gcols = EPI == GEMM_EPI_QKV ? p.head_dim : 32, n_g = BN / gcols;
        const int blocks_per_group = (gcols / 32) * (EPI == GEMM_EPI_SWIGLU ? 2 : 1);
        for (; s >= 2; --s) {
            const int kb_per = (num_kb + s - 1) / s;
            if ((num_kb + kb_per - 1) / kb_per != s) continue;   // would leave an empty slice
            if (((n_g + s - 1) / s) * blocks_per_group * s > kRedBlocks) continue;
            if ((size_t)rem * s * CG * kSliceBytes > p.split_ws_bytes) continue;
            us.whole = n_tiles - rem; us.splits = s; us.kb_per = kb_per;
            us.n_units = us.whole + rem * s;
            break;
        }
    }
    const int tm = tiles_m;
    for (int attempt = 0; attempt < 2; ++attempt) {
        cudaLaunchConfig_t cfg{};
        cfg.gridDim = dim3(CG * (us.n_units < n_slots ? us.n_units : n_slots));
        cfg.blockDim = dim3(kGemmThreads);
        cfg.dynamicSmemBytes = smem;
        cfg.stream = st;
        cudaLaunchAttribute attr[3];
        int na = 0;
        if (CG == 2) {
            attr[na].id = cudaLaunchAttributeClusterDimension;
            attr[na].val.clusterDim.x = 2; attr[na].val.clusterDim.y = 1; attr[na].val.clusterDim.z = 1;
            ++na;
        }
        if (us.splits > 1) {   // the slices of a tile wait for each other: all CTAs must be resident together
            attr[na].id = cudaLaunchAttributeCooperative;
            attr[na].val.cooperative = 1;
            ++na;
        } else if (pdl_enabled()) {
            attr[na].id = cudaLaunchAttributeProgrammaticStreamSerialization;
            attr[na].val.programmaticStreamSerializationAllowed = 1;
            ++na;
        }
        cfg.attrs = attr;
        cfg.numAttrs = na;
        const cudaError_t e = cudaLaunchKernelEx(&cfg, gemm_tc_kernel<EPI, BN, CG>, a, b, b2, b3, p, us, tm);
        if (e == cudaSuccess) {
            count_launch();
            return e;
        }
        if (us.splits == 1) return e;
        cudaGetLastError();   // cannot be co-resident here: whole tiles
        us = UnitSched{n_tiles, n_tiles, 1, num_kb};
    }
    return cudaErrorUnknown;
}

// 256-wide tiles once they still give every SM a tile (and the shape has full 256-column tiles to give); with the scratch
// area for split tails at hand always -- the widest tile moves the fewest bytes into shared memory per flop, and the
// split spreads however few tiles there are over the SMs
bool wide_tiles(const GemmTcParams& p) {
    const char* env = getenv("DLLM_GEMM_BN256");  // read per call: the tests compare both tile widths in one process
    const bool off = env && env[0] == '0';
    if (off || p.epi == GEMM_EPI_SWIGLU || p.N % 256 || p.n_more[0] % 256 || p.n_more[1] % 256) return false;
    if (p.split_ws && gemm_split_enabled()) return true;
    return (long)((p.M + BM - 1) / BM) * ((p.N + p.n_more[0] + p.n_more[1]) / 256) >= 148;
}

// 64-wide tiles when 128-wide ones would leave SMs without a tile (mid-size prompts: M = 256 .. 1024) and there is no
// scratch area to split tiles along K instead
bool narrow_tiles(const GemmTcParams& p) {
    const char* env = getenv("DLLM_GEMM_BN64");
    const bool off = env && env[0] == '0';
    if (off || p.epi == GEMM_EPI_SWIGLU || (p.split_ws && gemm_split_enabled())) return false;
    const long n_all = (long)p.N + p.n_more[0] + p.n_more[1];
    return (long)((p.M + BM - 1) / BM) * ((n_all + 127) / 128) <= 74;  // at most half the SMs would get a 128-wide tile
}

template <int EPI, int MP>
cudaError_t launch_swapab(const GemmTcParams& p, cudaStream_t st) {
    using Cfg = SwapAbCfg<EPI, MP>;
    CUtensorMap w, w2, w3, x;
    if (!make_map(&w, p.w, p.N, p.K, BM) || !make_map(&x, p.a, p.M, p.K, MP)) return cudaErrorNotSupported;
    w2 = w3 = w;
    int tiles = (p.N + BM - 1) / BM;
    if (Cfg::kDual) {
        if (!make_map(&w2, p.w2, p.N, p.K, BM)) return cudaErrorNotSupported;
    } else {
        if (p.n_more[0] > 0 && !make_map(&w2, p.w_more[0], p.n_more[0], p.K, BM)) return cudaErrorNotSupported;
        if (p.n_more[1] > 0 && !make_map(&w3, p.w_more[1], p.n_more[1], p.K, BM)) return cudaErrorNotSupported;
        tiles += (p.n_more[0] + BM - 1) / BM + (p.n_more[1] + BM - 1) / BM;
    }
    const int total_kb = (p.K + BK - 1) / BK;
    int ksplit = 1;
    const int n_sm = sm_count();
    // one wave of CTAs: every CTA pays a fixed set-up (tensor memory, barriers, two cluster barriers, the DSMEM pass),
    // so 128 CTAs streaming twice as much each beat 256 CTAs in two waves
    while (ksplit < Cfg::kMaxSplit && tiles * ksplit * 2 <= n_sm + 12 && ksplit * 2 <= total_kb) ksplit *= 2;   // (clusters of 3: measured slower)
    const int kb_per_cta = (total_kb + ksplit - 1) / ksplit;
    const size_t smem = (size_t)Cfg::kStagesS * Cfg::kStageBytes + 1024;
    static PerDeviceOnce attr_once;
    if (attr_once.need()) {
        cudaError_t e = cudaFuncSetAttribute(gemm_swapab_kernel<EPI, MP>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem);
        if (e != cudaSuccess) return e;
    }
    cudaLaunchConfig_t cfg{};
    cfg.gridDim = dim3(tiles, ksplit);
    cfg.blockDim = dim3(kGemmThreads);
    cfg.dynamicSmemBytes = smem;
    cfg.stream = st;
    cudaLaunchAttribute attr[2];
    attr[0].id = cudaLaunchAttributeClusterDimension;
    attr[0].val.clusterDim.x = 1;
    attr[0].val.clusterDim.y = ksplit;
    attr[0].val.clusterDim.z = 1;
    attr[1].id = cudaLaunchAttributeProgrammaticStreamSerialization;   // see common.cuh: the kernel calls pdl_wait()
    attr[1].val.programmaticStreamSerializationAllowed = 1;
    cfg.attrs = attr;
    cfg.numAttrs = pdl_enabled() ? 2 : 1;
    cudaError_t e = cudaLaunchKernelEx(&cfg, gemm_swapab_kernel<EPI, MP>, w, w2, w3, x, p, kb_per_cta);
    count_launch();
    return e;
}

template <int EPI>
cudaError_t launch_swapab_rows(const GemmTcParams& p, cudaStream_t st) {
    if (p.M <= 16) return launch_swapab<EPI, 16>(p, st);
    if (p.M <= 32) return launch_swapab<EPI, 32>(p, st);
    return launch_swapab<EPI, 64>(p, st);
}

bool small_m(const GemmTcParams& p) { return gemm_swapab_enabled() && p.M <= 64; }

}  // namespace

bool gemm_swapab_enabled() {
    static const bool off = [] { const char* e = getenv("DLLM_GEMM_SWAPAB"); return e && e[0] == '0'; }();
    return !off;
}

bool gemm_qkv_fusable(int n_rows, int head_dim, const void* bq, const void* bk, const void* bv) {
    const char* env = getenv("DLLM_QKV_FUSED");
    if (env && env[0] == '0') return false;
    const auto al = [](const void* b) { return (reinterpret_cast<uintptr_t>(b) & 15) == 0; };
    return n_rows > 64 && (head_dim == 64 || head_dim == 128) && al(bq) && al(bk) && al(bv);
}

cudaError_t launch_gemm_tc(const GemmTcParams& p, cudaStream_t st) {
    if (p.M <= 0 || p.N <= 0 || p.K <= 0 || p.K % 8 || p.ldc < p.N || (p.ldc % 2 && p.epi != GEMM_EPI_LOGITS)) return cudaErrorInvalidValue;
    const bool more = p.n_more[0] > 0 || p.n_more[1] > 0;
    if (more && ((p.epi != GEMM_EPI_BIAS && p.epi != GEMM_EPI_QKV) || (p.n_more[1] > 0 && p.n_more[0] <= 0) || p.ldc < p.N + p.n_more[0] + p.n_more[1]))
        return cudaErrorInvalidValue;
    if (small_m(p)) {
        switch (p.epi) {
            case GEMM_EPI_BIAS: return launch_swapab_rows<GEMM_EPI_BIAS>(p, st);
            case GEMM_EPI_RESIDUAL: return launch_swapab_rows<GEMM_EPI_RESIDUAL>(p, st);
            case GEMM_EPI_SWIGLU: return launch_swapab_rows<GEMM_EPI_SWIGLU>(p, st);
            case GEMM_EPI_GELU: return launch_swapab_rows<GEMM_EPI_GELU>(p, st);
            case GEMM_EPI_LOGITS: return launch_swapab_rows<GEMM_EPI_LOGITS>(p, st);
        }
        return cudaErrorInvalidValue;
    }
    if (p.epi == GEMM_EPI_LOGITS) return cudaErrorInvalidValue;
    if (p.epi == GEMM_EPI_QKV && ((p.head_dim != 64 && p.head_dim != 128) || p.N % p.head_dim || p.n_more[0] % p.head_dim ||
                                  p.n_more[1] % p.head_dim || !p.q_out || !p.positions || !p.seq_ids || !gemm_qkv_fusable(p.M, p.head_dim, p.bias, p.bias_more[0], p.bias_more[1])))
        return cudaErrorInvalidValue;
    CUtensorMap a, b, b2, b3;
    int bn = wide_tiles(p) ? 256 : (narrow_tiles(p) ? 64 : 128);
    if (p.epi == GEMM_EPI_QKV && bn < p.head_dim) bn = 128;   // a tile holds whole heads
    // CTA pairs on 256-row tiles: full-width tiles (256, or the dual 128 + 128), more than one 128-row tile of rows
    const bool dual = p.epi == GEMM_EPI_SWIGLU;
    // (measured: +5 % at 4096 rows -- 1.34 against 1.27 PFLOP/s on 4096 x 4096 x 14336 -- nothing at 512 rows, where the
    // kernels are all prologue and tail; DLLM_GEMM_CG2=1 forces pairs from 129 rows on, for the tests)
    const char* pair_env = getenv("DLLM_GEMM_CG2");
    const int pair_min_rows = pair_env && pair_env[0] == '1' ? BM + 1 : 1024;
    const bool pair = gemm_pair_enabled() && p.M >= pair_min_rows && (dual ? p.N % 128 == 0 : bn == 256);
    const int box = pair ? (dual ? 64 : 128) : (dual ? BM : bn);   // weight rows one CTA loads per tile
    if (!make_map(&a, p.a, p.M, p.K) || !make_map(&b, p.w, p.N, p.K, box)) return cudaErrorNotSupported;
    b2 = b3 = b;
    if (dual) {
        if (!make_map(&b2, p.w2, p.N, p.K, box)) return cudaErrorNotSupported;
    } else {
        if (p.n_more[0] > 0 && !make_map(&b2, p.w_more[0], p.n_more[0], p.K, box)) return cudaErrorNotSupported;
        if (p.n_more[1] > 0 && !make_map(&b3, p.w_more[1], p.n_more[1], p.K, box)) return cudaErrorNotSupported;
    }
    if (pair) {
        cudaError_t e = cudaErrorNotSupported;
        switch (p.epi) {
            case GEMM_EPI_BIAS: e = launch_epi<GEMM_EPI_BIAS, 256, 2>(a, b, b2, b3, p, st); break;
            case GEMM_EPI_RESIDUAL: e = launch_epi<GEMM_EPI_RESIDUAL, 256, 2>(a, b, b2, b3, p, st); break;
            case GEMM_EPI_SWIGLU: e = launch_epi<GEMM_EPI_SWIGLU, 128, 2>(a, b, b2, b3, p, st); break;
            case GEMM_EPI_GELU: e = launch_epi<GEMM_EPI_GELU, 256, 2>(a, b, b2, b3, p, st); break;
            case GEMM_EPI_QKV: e = launch_epi<GEMM_EPI_QKV, 256, 2>(a, b, b2, b3, p, st); break;
        }
        return e;
    }
#define DLLM_GEMM_BY_BN(E)                                                                        \
    (bn == 256 ? launch_epi<E, 256, 1>(a, b, b2, b3, p, st)                                       \
               : (bn == 64 ? launch_epi<E, 64, 1>(a, b, b2, b3, p, st) : launch_epi<E, 128, 1>(a, b, b2, b3, p, st)))
    switch (p.epi) {
        case GEMM_EPI_BIAS: return DLLM_GEMM_BY_BN(GEMM_EPI_BIAS);
        case GEMM_EPI_RESIDUAL: return DLLM_GEMM_BY_BN(GEMM_EPI_RESIDUAL);
        case GEMM_EPI_SWIGLU: return launch_epi<GEMM_EPI_SWIGLU, 128, 1>(a, b, b2, b3, p, st);
        case GEMM_EPI_GELU: return DLLM_GEMM_BY_BN(GEMM_EPI_GELU);
        case GEMM_EPI_QKV: return DLLM_GEMM_BY_BN(GEMM_EPI_QKV);
    }
#undef DLLM_GEMM_BY_BN
    return cudaErrorInvalidValue;
}

cudaError_t launch_norm_rows(const void* x, const void* w, const void* b, void* y, int rows, int K, int norm, float eps,
                             cudaStream_t st, const int32_t* row_map) {
    if (rows <= 0 || K % 8) return cudaErrorInvalidValue;
    cudaError_t le = cudaSuccess;
    if ((rows <= 64 || row_map) && K <= 16384) {
        le = launch_pdl(norm_rows_cta_kernel, dim3(rows), dim3(256), 0, st, static_cast<const bf16*>(x), static_cast<const bf16*>(w),
                        static_cast<const bf16*>(b), static_cast<bf16*>(y), row_map, K, norm, eps);
    } else {
        if (row_map) return cudaErrorInvalidValue;
        le = launch_pdl(norm_rows_kernel, dim3((rows + 7) / 8), dim3(256), 0, st, static_cast<const bf16*>(x), static_cast<const bf16*>(w),
                        static_cast<const bf16*>(b), static_cast<bf16*>(y), rows, K, norm, eps);
    }
    count_launch();
    return le;
}

cudaError_t launch_argmax_rows(const float* logits, int rows, int n, int32_t* next_token, cudaStream_t st) {
    if (rows <= 0 || n <= 0) return cudaErrorInvalidValue;
    const cudaError_t le = launch_pdl(argmax_rows_kernel, dim3(rows), dim3(256), 0, st, logits, n, next_token);
    count_launch();
    return le;
}

cudaError_t launch_rope_append(const RopeAppendParams& p, cudaStream_t st) {
    const int total = p.n_rows * (p.n_heads + 2 * p.n_kv_heads) * (p.head_dim / 2);
    if (total <= 0) return cudaErrorInvalidValue;
    int blocks = (total + 255) / 256;
    if (blocks > 148 * 8) blocks = 148 * 8;
    const cudaError_t le = launch_pdl(rope_append_kernel, dim3(blocks), dim3(256), 0, st, p);
    count_launch();
    return le;
}

}  // namespace dllm
