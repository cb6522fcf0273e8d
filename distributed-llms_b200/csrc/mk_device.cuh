// Device code of the persistent decode kernel (decode_megakernel.cu) kept apart from its phase logic, and shared with the
// hand-off kernels of the per-op path (pipe_ops.cu): PTX wrappers, grid barrier and tickets, the per-warp ring of TMA
// bulk stages and its static stage sequence, the row-normalisation vector math, and the attention unit geometry.
#pragma once

#include <float.h>
#include <limits.h>

#include "common.cuh"
#include "kernels.h"
#include "megakernel.h"
#include "ptx.cuh"

namespace dllm {
namespace mkd {

using namespace ptx;

constexpr int kThreads = MK_THREADS;
constexpr int kWarps = MK_THREADS / 32;
constexpr int kStageBytes = MK_STAGE_BYTES;
constexpr int kStageK = MK_STAGE_K;
constexpr int kTT = 64;     // tokens per attention tile
constexpr int kMaxGq = 8;   // query heads per KV head handled by one attention unit
constexpr int kMaxSplits = MK_MAX_SPLITS;
constexpr int kPartStride = 8 * MK_MAX_ROWS;  // floats per unit in the partial buffer: [8 panel rows][16 batch]

// ---- PTX wrappers beyond ptx.cuh ------------------------------------------------------------------------
__device__ __forceinline__ void cp_async_wait_but_one() { asm volatile("cp.async.wait_group 1;" ::: "memory"); }
__device__ __forceinline__ uint4 lds_u4(uint32_t addr) {
    uint4 r;
    asm volatile("ld.shared.v4.u32 {%0, %1, %2, %3}, [%4];" : "=r"(r.x), "=r"(r.y), "=r"(r.z), "=r"(r.w) : "r"(addr));
    return r;
}

__device__ __forceinline__ unsigned ld_acquire_u32(const unsigned* p) {
    unsigned v;
    asm volatile("ld.acquire.gpu.global.u32 %0, [%1];" : "=r"(v) : "l"(p) : "memory");
    return v;
}
__device__ __forceinline__ void red_release_add(unsigned* p, unsigned v) {
    asm volatile("red.release.gpu.global.add.u32 [%0], %1;" ::"l"(p), "r"(v) : "memory");
}
__device__ __forceinline__ uint4 ldcg_u4(const void* p) { return __ldcg(reinterpret_cast<const uint4*>(p)); }
// positions / ids change between steps inside one launch: always read them from L2
__device__ __forceinline__ int ldcg_i32(const int32_t* p) { return __ldcg(p); }

// All CTAs of the (co-resident, cooperative) grid meet here.  `target` advances by gridDim.x per barrier.
// The barrier in two halves, so that work which does not depend on the other CTAs (cp.async prefetches of old K/V rows)
// can be issued between this CTA's arrival and the moment everybody has arrived.
__device__ __forceinline__ void grid_barrier_arrive(unsigned* bar, unsigned& target, long long* tr = nullptr) {
    __syncthreads();
    target += gridDim.x;
    if (threadIdx.x == 0) {
        // release: the CTA's earlier writes (ordered before by bar.sync) become visible before the arrival
        if (tr) tr[8] = clock64();
        red_release_add(bar, 1u);
        if (tr) tr[9] = clock64();
    }
}
// Returns true (to every thread of the CTA) when the launch has been aborted -- a stage gave up waiting for a ticket,
// `status` is non-zero -- and this CTA must leave instead of waiting for arrivals that will not come.
__device__ __forceinline__ bool grid_barrier_wait(const unsigned* bar, unsigned target, long long* tr = nullptr,
                                                  const int* status = nullptr) {
    int bad = 0;
    if (threadIdx.x == 0) {
        // acquire: nothing after the barrier is satisfied from stale data
        for (unsigned spins = 0; ld_acquire_u32(bar) < target; ++spins) {
            if (status && (spins & 0x3ffu) == 0x3ffu && *reinterpret_cast<const volatile int*>(status) != 0) {
                bad = 1;
                break;
            }
            if (spins > (1u << 26)) __trap();  // a lost arrival is a bug; do not hang the GPU
        }
        if (tr) tr[10] = clock64();
    }
    if (status == nullptr) {   // classic launches: nothing can abort them
        __syncthreads();
        return false;
    }
    return __syncthreads_or(bad) != 0;
}
__device__ __forceinline__ void grid_barrier(unsigned* bar, unsigned& target, long long* tr = nullptr) {
    grid_barrier_arrive(bar, target, tr);
    grid_barrier_wait(bar, target, tr);
}

// ---- tickets in (peer) mailboxes: system scope, the writer may be another GPU ------------------------------
__device__ __forceinline__ void st_release_sys(unsigned* p, unsigned v) {
    asm volatile("st.release.sys.global.u32 [%0], %1;" ::"l"(p), "r"(v) : "memory");
}
__device__ __forceinline__ unsigned ld_acquire_sys(const unsigned* p) {
    unsigned v;
    asm volatile("ld.acquire.sys.global.u32 %0, [%1];" : "=r"(v) : "l"(p) : "memory");
    return v;
}
__device__ __forceinline__ unsigned long long global_ns() {
    unsigned long long t;
    asm volatile("mov.u64 %0, %globaltimer;" : "=l"(t));
    return t;
}
// One thread waits until the slot's ticket reaches `want`.  The producer is another kernel (on another GPU): processes
// start seconds apart, so the wait is long (`timeout_ns`, 20 s by default) -- but it ends: on a timeout the thread marks
// `status` (the first one to do so names the reason) and returns false, as it does when another CTA or an earlier kernel
// of this stage has already given up.  The caller then leaves its kernel in an orderly way; nothing traps, the context
// stays usable and the host reads the status word (dllm_shard_pipeline_status).
__device__ __forceinline__ bool wait_ticket(const unsigned* flag, unsigned want, int* status, unsigned long long timeout_ns) {
    if ((int)(ld_acquire_sys(flag) - want) >= 0) return true;
    const unsigned long long t0 = global_ns();
    while ((int)(ld_acquire_sys(flag) - want) < 0) {
        __nanosleep(64);
        if (*reinterpret_cast<volatile int*>(status) != 0) return false;
        if (global_ns() - t0 > timeout_ns) {
            atomicCAS(status, 0, 1);
            return false;
        }
    }
    return true;
}
// after an abort: bulk copies into this CTA's shared memory may still be in flight -- let them land before the CTA exits
__device__ __forceinline__ void abort_drain() {
    const unsigned long long t0 = global_ns();
    while (global_ns() - t0 < 200000ull) __nanosleep(1000);
}

__device__ __forceinline__ uint4 pack8(const float (&in)[8]) {
    uint4 r;
    __nv_bfloat162 a = __floats2bfloat162_rn(in[0], in[1]);
    __nv_bfloat162 b = __floats2bfloat162_rn(in[2], in[3]);
    __nv_bfloat162 c = __floats2bfloat162_rn(in[4], in[5]);
    __nv_bfloat162 d = __floats2bfloat162_rn(in[6], in[7]);
    r.x = *reinterpret_cast<uint32_t*>(&a);
    r.y = *reinterpret_cast<uint32_t*>(&b);
    r.z = *reinterpret_cast<uint32_t*>(&c);
    r.w = *reinterpret_cast<uint32_t*>(&d);
    return r;
}

__device__ __forceinline__ float ldf(const void* p, int i) { return __bfloat162float(static_cast<const bf16*>(p)[i]); }
__device__ __forceinline__ float silu_f(float g) { return g / (1.0f + expf(-g)); }
__device__ __forceinline__ float gelu_new_f(float x) {
    return 0.5f * x * (1.0f + tanhf(0.7978845608028654f * (x + 0.044715f * x * x * x)));
}

struct ArgBest {
    float val;
    int idx;
};
__device__ __forceinline__ void arg_update(ArgBest& b, float v, int i) {
    if (v > b.val || (v == b.val && i < b.idx)) {
        b.val = v;
        b.idx = i;
    }
}

__device__ __forceinline__ bool is_weight_phase(int kind) {
    return kind == MK_QKV || kind == MK_OPROJ || kind == MK_GATEUP || kind == MK_FC_GELU || kind == MK_DOWN ||
           kind == MK_HEAD;
}

// Per-step constants every phase needs (positions do not change inside a step): cached in shared memory at
// the start of the step so that epilogues and attention set-up have no dependent global loads.
struct StepCache {
    int pos[MK_MAX_ROWS];
    int seq[MK_MAX_ROWS];
    int page[MK_MAX_ROWS];   // KV page that holds position pos[r] of sequence seq[r]
    int off[MK_MAX_ROWS];    // offset of pos[r] inside that page
    // the split geometry of each row's context, the same in every layer of the step: tokens per split (whole 64-token
    // tiles) and how many splits own a token -- worked out once per step, not per attention phase
    int chunk[MK_MAX_ROWS];
    int n_act[MK_MAX_ROWS];
};
__device__ __forceinline__ void step_cache_splits(StepCache& sc, int r, int n_splits) {
    const int len = sc.pos[r] + 1;
    int chunk = (len + n_splits - 1) / n_splits;
    chunk = (chunk + 63) / 64 * 64;   // kTT
    sc.chunk[r] = chunk;
    sc.n_act[r] = (len + chunk - 1) / chunk;
}

// ---- a warp's static sequence of stages -------------------------------------------------------------------
// Panel p belongs to CTA p % grid (local index p / grid).  A CTA works through its panels in rounds of up to
// MK_ROUND_PANELS; the stages of a round (panel-major) are cut into kWarps equal contiguous chunks, one per
// warp, so the warps finish together whatever the panel count (a chunk may end one panel and begin the next:
// the warp then writes two partial sums).  The issue side and the consume side walk the same sequence; the
// issue side is `stages` entries ahead.
struct WIter {
    int step, ph, j0;       // decode step, phase, first local panel of the round
    int x, x_end;           // position / end of this warp's chunk in the round's stage sequence
    int jr, st;             // local panel inside the round and stage inside the panel at x
    // cached at seek time so that advancing by one stage needs no division and no phase-table lookup
    const unsigned char* wp;
    int K, ns, xg;
    int p0, pstride;        // this CTA's panel with local index i in the phase at hand: p0 + pstride * i
};

__device__ __forceinline__ int local_panels(const MkPhase& ph) {
    const int c = blockIdx.x, G = gridDim.x;
    return c < ph.n_panels ? (ph.n_panels - c + G - 1) / G : 0;
}
// Panels -> CTAs.  Interleaved: panel p belongs to CTA p % grid.  Blocked: CTA c owns a contiguous run of panels (the first
// n % grid CTAs one more than the others -- the same counts as interleaved), i.e. one contiguous piece of every weight
// matrix: its TMA stream walks 2 MB pages one after the other instead of touching every page of the matrix.
__device__ __forceinline__ int panel_first(const MkPhase& ph, int blocked) {
    const int c = blockIdx.x, G = gridDim.x;
    if (!blocked) return c;
    const int q = ph.n_panels / G, rem = ph.n_panels - q * G;
    return c * q + min(c, rem);
}

struct Ctx {
    int n_steps, n_phases;
    const unsigned char* act_x;  // activation rows in fragment layout (xg phases)
    int cur_key;            // step * n_phases + phase the CTA is consuming now
    int* xkey;              // shared, per ring slot of this warp: key of the xg stage whose row half is still to be issued
    int* xst;               //         and its stage index (lane 0 only)
    const MkPhase* phases;  // shared-memory copy
    const int* units;       // shared: panels this CTA owns in each phase (0 = not a weight phase)
    const int* first;       // shared: the first of them (blocked assignment) -- see panel_first()
    int warp, lane;
    uint32_t ring;          // shared address of this warp's stages
    uint32_t full;          // shared address of this warp's mbarriers
};

// chunk of warp w in a round of T stages: [w * T / kWarps, (w + 1) * T / kWarps)
__device__ __forceinline__ int chunk_begin(int w, int T) { return (w * T) / kWarps; }
// x / ph.n_stages without an integer division: x is a stage index of a round (< 2^16), so (x + 0.5) * (1 / ns) is at
// least 0.5 / ns away from every integer -- far more than the rounding error of the product
__device__ __forceinline__ int div_stages(int x, const MkPhase& ph) { return (int)(((float)x + 0.5f) * ph.inv_ns); }

__device__ __forceinline__ void witer_seek(const Ctx& c, WIter& it) {
    while (it.step < c.n_steps) {
        while (it.ph < c.n_phases) {
            const int n_local = c.units[it.ph];
            while (it.j0 < n_local) {
                const MkPhase& ph = c.phases[it.ph];
                const int T = min(MK_ROUND_PANELS, n_local - it.j0) * ph.n_stages;
                it.x = chunk_begin(c.warp, T);
                it.x_end = chunk_begin(c.warp + 1, T);
                if (it.x < it.x_end) {
                    it.ns = ph.n_stages;
                    it.jr = div_stages(it.x, ph);
                    it.st = it.x - it.jr * it.ns;
                    it.wp = static_cast<const unsigned char*>(ph.wp);
                    it.K = ph.K;
                    it.xg = ph.xg;
                    it.p0 = c.first[it.ph];
                    return;
                }
                it.j0 += MK_ROUND_PANELS;
            }
            ++it.ph;
            it.j0 = 0;
        }
        it.ph = 0;
        ++it.step;
    }
}

// XG = false: the launch has no streamed-rows phase (the variant's code is compiled out)
template <bool XG = true>
__device__ __forceinline__ void issue_stage(const Ctx& c, const WIter& it, int slot) {
    if (c.lane == 0) {
        uint32_t bytes;
        const uint32_t dst = c.ring + slot * kStageBytes, bar = c.full + slot * 8;
        const int panel = it.p0 + it.pstride * (it.j0 + it.jr);
        if (!XG || !it.xg) {
            const unsigned char* src = it.wp + mk_stage_offset(it.K, panel, it.st, kStageBytes, &bytes);
            mbar_expect_tx(bar, bytes);
            bulk_g2s(dst, src, bytes, bar);
        } else {
            // half-size stage: 128 k of the panel, then the same 128 k of the (up to 8) input rows.
            // The weights may run ahead of the grid barrier, the rows may not: they are written by the phase before.
            // A stage issued early arms the barrier for both halves and leaves the row half to issue_deferred_rows().
            const unsigned char* src = it.wp + mk_stage_offset(it.K, panel, it.st, kStageBytes / 2, &bytes);
            mbar_expect_tx(bar, bytes + kStageBytes / 2);
            bulk_g2s(dst, src, bytes, bar);
            const int key = it.step * c.n_phases + it.ph;
            if (key == c.cur_key) {
                bulk_g2s(dst + kStageBytes / 2, c.act_x + (size_t)it.st * (kStageBytes / 2), kStageBytes / 2, bar);
            } else {
                c.xkey[slot] = key;
                c.xst[slot] = it.st;
            }
        }
    }
}

// Start of an xg phase (after the grid barrier): issue the row halves of the stages that were armed early.
__device__ __forceinline__ void issue_deferred_rows(const Ctx& c, int stages) {
    if (c.lane == 0) {
        asm volatile("fence.proxy.async.global;" ::: "memory");
        for (int slot = 0; slot < stages; ++slot)
            if (c.xkey[slot] == c.cur_key) {
                c.xkey[slot] = -1;
                bulk_g2s(c.ring + slot * kStageBytes + kStageBytes / 2, c.act_x + (size_t)c.xst[slot] * (kStageBytes / 2),
                         kStageBytes / 2, c.full + slot * 8);
            }
    }
}

__device__ __forceinline__ void witer_next(const Ctx& c, WIter& it) {
    if (++it.x == it.x_end) {
        it.j0 += MK_ROUND_PANELS;
        witer_seek(c, it);
    } else if (++it.st == it.ns) {
        it.st = 0;
        ++it.jr;
    }
}


__device__ __forceinline__ void accum_stats(const uint4& v, float& s1, float& s2) {
    float f[8];
    unpack<bf16>(v, f);
#pragma unroll
    for (int j = 0; j < 8; ++j) {
        s1 += f[j];
        s2 = fmaf(f[j], f[j], s2);
    }
}

__device__ __forceinline__ uint4 norm_vec(const uint4& x, const uint4& w, const uint4& b, int norm, float mean, float rstd) {
    float f[8], wf[8], bf[8];
    unpack<bf16>(x, f);
    unpack<bf16>(w, wf);
    if (norm == NORM_RMS) {
#pragma unroll
        for (int j = 0; j < 8; ++j) f[j] = wf[j] * bf16_round(f[j] * rstd);
    } else {
        unpack<bf16>(b, bf);
#pragma unroll
        for (int j = 0; j < 8; ++j) f[j] = fmaf((f[j] - mean) * rstd, wf[j], bf[j]);
    }
    return pack8(f);
}

// number of splits of row r that own at least one token (the others never run and are never read)
__device__ __forceinline__ int attn_active_splits(const MkParams&, const StepCache& sc, int r) { return sc.n_act[r]; }

__device__ __forceinline__ float4 ldcg_f4(const float* p) { return __ldcg(reinterpret_cast<const float4*>(p)); }

// ---- attention phase ------------------------------------------------------------------------------------

struct AttnSmem {
    bf16* k;      // [kTT][hd + 8]
    bf16* v;      // [kTT][hd + 8]
    float* s;     // [kMaxGq][kTT]
    float* q;     // 16 bf16 rows [16][hd + 8] while a unit walks its tiles, then fp32 scratch of the split merge
};

// cp.async the K and V rows of tokens [a, b) of (seq, kvh) into tile rows (token - tile0)
// (inlined at its six call sites: as a real call it spills the attention accumulators around every call -- measured 3 % slower)
__device__ __forceinline__ void attn_load_rows(const MkParams& p, const MkLayer& L, const AttnSmem& sm, const int* bt,
                                               int kvh, int tile0, int a, int b, bool load_k, bool load_v, int hd) {
    const int vecs = hd >> 3, ld = hd + 8;   // hd: p.head_dim, a compile-time constant in the specialised kernels
    const int vshift = 31 - __clz(vecs);   // head_dim is 32 / 64 / 128: vecs is a power of two
    const int mask = (1 << p.page_shift) - 1;
    for (int i = threadIdx.x; i < (b - a) * vecs; i += kThreads) {
        const int t = a + (i >> vshift), v = i & (vecs - 1);
        const int page = bt[t >> p.page_shift];
        const size_t base = ((((size_t)page * p.n_kv_heads + kvh) << p.page_shift) + (t & mask)) * hd + v * 8;
        if (load_k) cp_async16(sm.k + (size_t)(t - tile0) * ld + v * 8, static_cast<const bf16*>(L.kpool) + base);
        if (load_v) cp_async16(sm.v + (size_t)(t - tile0) * ld + v * 8, static_cast<const bf16*>(L.vpool) + base);
    }
    if (load_v) {
        // V rows between the last token and the end of its 16-token MMA step: their probabilities are zero, but
        // 0 x (stale NaN / Inf bits) is not -- clear them
        const int r0 = b - tile0, r1 = min(kTT, (r0 + 15) & ~15);
        for (int i = threadIdx.x; i < (r1 - r0) * vecs; i += kThreads)
            *reinterpret_cast<uint4*>(sm.v + (size_t)(r0 + (i >> vshift)) * ld + (i & (vecs - 1)) * 8) = make_uint4(0u, 0u, 0u, 0u);
    }
    cp_async_commit();  // one group per call: the K rows and the V rows of a tile are waited for separately
}

struct AttnUnit {
    int r, kvh, split, seq, t0, t1;
};

// (row, KV head, split) of unit u: fixed for the whole launch -- a CTA works its first unit's out once (UnitId) and only
// the token range, which moves with the positions, is recomputed per step
struct UnitId {
    int r, kvh, split;
};
__device__ __forceinline__ UnitId unit_id(const MkParams& p, int u) {
    const int S = p.n_splits, KV = p.n_kv_heads;
    UnitId id;
    id.r = u / (KV * S);
    id.kvh = (u / S) % KV;
    id.split = u % S;
    return id;
}
__device__ __forceinline__ AttnUnit attn_unit(const MkParams& p, const StepCache& sc, const UnitId& id) {
    AttnUnit a;
    a.r = id.r;
    a.kvh = id.kvh;
    a.split = id.split;
    const int len = sc.pos[a.r] + 1;
    a.seq = sc.seq[a.r];
    const int chunk = sc.chunk[a.r];  // whole tiles per split: no one-token tail tiles
    a.t0 = min(len, a.split * chunk);
    a.t1 = min(len, a.t0 + chunk);
    return a;
}

// Tokens that may be written during this very step (any row of the same sequence) start here; everything
// below is old and safe to prefetch before the barrier.
__device__ __forceinline__ int attn_safe_end(const StepCache& sc, const AttnUnit& a) {
    return sc.pos[a.r] - (MK_MAX_ROWS - 1);
}

// the query buffer of the attention phase: 16 bf16 rows (the MMA's M side) during the tiles, fp32 scratch afterwards
__host__ __device__ inline size_t mk_attn_q_bytes(int head_dim) {
    const size_t rows16 = (size_t)16 * (head_dim + 8) * 2, f32 = (size_t)kMaxGq * head_dim * 4;
    return rows16 > f32 ? rows16 : f32;
}

// partial sums of a round: one unit per (panel, warp) piece, [8 panel rows][rows rounded up to a power of two] floats each
__host__ __device__ inline size_t mk_part_bytes(const MkParams& p) {
    int rp = 1;
    while (rp < p.n_rows) rp <<= 1;
    return ((size_t)p.part_units * 8 * rp * 4 + 127) / 128 * 128;
}

// shared region that holds the staged input rows (weight phases) or the attention tiles (attention phase)
__host__ __device__ inline size_t mk_rows_region_bytes(const MkParams& p) {
    const size_t rows = (size_t)p.n_rows * (p.k_max + 8) * 2;
    const size_t attn = (size_t)2 * kTT * (p.head_dim + 8) * 2 + (size_t)kMaxGq * kTT * 4 + mk_attn_q_bytes(p.head_dim);
    return ((rows > attn ? rows : attn) + 127) / 128 * 128;
}

}  // namespace mkd
}  // namespace dllm
