// Fused decode GEMV family for sm_100a.
//
// One kernel template covers every weight-streaming step of a decode layer (SURVEY.md 7.2 K1,K3-K6):
//   prologue : stage the (few) input rows in shared memory, optionally RMSNorm / LayerNorm them
//   main     : each warp streams two weight rows with 128-bit coalesced, L1-bypassing loads
//              (8 vectors in flight per lane) and reduces with warp shuffles -- HBM-bound by design
//   epilogue : QKV -> bias + rotary + paged-KV append | residual add | SwiGLU | gelu_new |
//              logits + greedy argmax (cross-block, last-block-done)
// The reference's only arithmetic, `torch.matmul(inputs[key], parameters[key])`
// (reference src/worker/node.py:31), is what all of this replaces.
#include <float.h>
#include <limits.h>

#include <atomic>

#include "common.cuh"
#include "kernels.h"

namespace dllm {

static std::atomic<unsigned long long> g_launches{0};
void count_launch(unsigned long long n) { g_launches.fetch_add(n, std::memory_order_relaxed); }
bool pdl_enabled() {
    // Off unless DLLM_PDL=1 (read per launch).  Measured on B200, Llama-3-8B decode steps of 16 / 32 rows as CUDA graphs of
    // these kernels, same box, alternating runs: 3.41-3.59 k / 5.63-5.71 k tokens/s without the attribute, 2.97-3.17 k /
    // 4.46-4.54 k with it (profiles/r2b_decode_latency_experiments.md) -- every programmatic edge costs about 2 us more than
    // a plain one here, more than the overlapped prologues give back.
    const char* e = getenv("DLLM_PDL");
    return e && e[0] == '1';
}
unsigned long long launch_count() { return g_launches.load(std::memory_order_relaxed); }

constexpr int kGemvThreads = 256;
constexpr int kGemvWarps = kGemvThreads / 32;
constexpr int kGemvOcc = 4;  // CTAs per SM the grid is sized for

template <typename T>
__device__ __forceinline__ uint4 pack(const float (&in)[DT<T>::VEC]);

template <>
__device__ __forceinline__ uint4 pack<float>(const float (&in)[4]) {
    return make_uint4(__float_as_uint(in[0]), __float_as_uint(in[1]), __float_as_uint(in[2]),
                      __float_as_uint(in[3]));
}

template <>
__device__ __forceinline__ uint4 pack<bf16>(const float (&in)[8]) {
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

// ---- prologue: stage rows [r0, r0+rt) of x into shared memory, normalised ------------------------
template <typename T>
__device__ void stage_rows(const GemvParams& p, int r0, int rt, T* xs, float* red) {
    constexpr int VEC = DT<T>::VEC;
    const int K = p.K;
    const T* x = static_cast<const T*>(p.x);
    const T* nw = static_cast<const T*>(p.norm_w);
    const T* nb = static_cast<const T*>(p.norm_b);
    for (int i = 0; i < rt; ++i) {
        const int r = r0 + i;
        const int src = p.row_map ? p.row_map[r] : r;
        const T* xr = x + (size_t)src * K;
        T* dst = xs + (size_t)i * K;
        float s1 = 0.f, s2 = 0.f;
        for (int k = threadIdx.x * VEC; k < K; k += kGemvThreads * VEC) {
            uint4 v = ld_u4(xr + k);
            *reinterpret_cast<uint4*>(dst + k) = v;
            float f[VEC];
            unpack<T>(v, f);
#pragma unroll
            for (int j = 0; j < VEC; ++j) {
                s1 += f[j];
                s2 = fmaf(f[j], f[j], s2);
            }
        }
        if (p.norm == NORM_NONE) continue;
        if (p.norm == NORM_RMS) {
            // HF LlamaRMSNorm: fp32 variance; x * rsqrt(var + eps) rounded to T; then weight * that.
            const float var = block_sum(s2, red) / (float)K;
            const float rstd = 1.0f / sqrtf(var + p.eps);
            for (int k = threadIdx.x * VEC; k < K; k += kGemvThreads * VEC) {
                float f[VEC], w[VEC];
                unpack<T>(*reinterpret_cast<const uint4*>(dst + k), f);
                unpack<T>(ld_u4(nw + k), w);
#pragma unroll
                for (int j = 0; j < VEC; ++j) f[j] = w[j] * DT<T>::rnd(f[j] * rstd);
                *reinterpret_cast<uint4*>(dst + k) = pack<T>(f);
            }
        } else {
            // LayerNorm, two-pass variance in fp32, affine, one rounding.
            const float mean = block_sum(s1, red) / (float)K;
            float sq = 0.f;
            for (int k = threadIdx.x * VEC; k < K; k += kGemvThreads * VEC) {
                float f[VEC];
                unpack<T>(*reinterpret_cast<const uint4*>(dst + k), f);
#pragma unroll
                for (int j = 0; j < VEC; ++j) sq = fmaf(f[j] - mean, f[j] - mean, sq);
            }
            const float var = block_sum(sq, red) / (float)K;
            const float rstd = 1.0f / sqrtf(var + p.eps);
            for (int k = threadIdx.x * VEC; k < K; k += kGemvThreads * VEC) {
                float f[VEC], w[VEC], b[VEC];
                unpack<T>(*reinterpret_cast<const uint4*>(dst + k), f);
                unpack<T>(ld_u4(nw + k), w);
                unpack<T>(ld_u4(nb + k), b);
#pragma unroll
                for (int j = 0; j < VEC; ++j) f[j] = fmaf((f[j] - mean) * rstd, w[j], b[j]);
                *reinterpret_cast<uint4*>(dst + k) = pack<T>(f);
            }
        }
    }
}

// ---- epilogues (one lane per batch row) -----------------------------------------------------------

template <typename T>
__device__ __forceinline__ float load_bias(const void* b, int idx) {
    return b ? DT<T>::to_f(static_cast<const T*>(b)[idx]) : 0.f;
}

template <typename T>
__device__ __forceinline__ void epi_qkv(const GemvParams& p, int unit, int r, float v0, float v1) {
    using D = DT<T>;
    const int hd = p.head_dim, half = hd >> 1, H = p.n_heads, KV = p.n_kv_heads;
    const int head = unit / half, j = unit - head * half;
    int m, hl;
    if (head < H) { m = 0; hl = head; }
    else if (head < H + KV) { m = 1; hl = head - H; }
    else { m = 2; hl = head - H - KV; }
    const int row0 = hl * hd + j;
    v0 = D::rnd(v0 + load_bias<T>(p.b[m], row0));
    v1 = D::rnd(v1 + load_bias<T>(p.b[m], row0 + half));
    const int pos = p.positions[r];
    if (p.rope && m != 2) {
        // HF apply_rotary_pos_emb: (x*cos) + (rotate_half(x)*sin), cos/sin cast to T first.
        const float c = D::rnd(p.rope_cos[(size_t)pos * half + j]);
        const float s = D::rnd(p.rope_sin[(size_t)pos * half + j]);
        const float o0 = D::rnd(D::rnd(v0 * c) + D::rnd(-v1 * s));
        const float o1 = D::rnd(D::rnd(v1 * c) + D::rnd(v0 * s));
        v0 = o0;
        v1 = o1;
    }
    if (m == 0) {
        T* q = static_cast<T*>(p.y) + (size_t)r * H * hd + row0;
        q[0] = D::from_f(v0);
        q[half] = D::from_f(v1);
    } else {
        const int seq = p.seq_ids[r];
        const int page = p.kv.block_table[(size_t)seq * p.kv.pages_per_seq + (pos >> p.kv.page_shift)];
        const int off = pos & ((1 << p.kv.page_shift) - 1);
        T* pool = static_cast<T*>(m == 1 ? p.kv.kpool : p.kv.vpool);
        T* dst = pool + ((((size_t)page * KV + hl) << p.kv.page_shift) + off) * hd + j;
        dst[0] = D::from_f(v0);
        dst[half] = D::from_f(v1);
    }
}

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

// ---- the kernel -----------------------------------------------------------------------------------

template <typename T, int RT, int EPI>
__global__ void __launch_bounds__(kGemvThreads, (RT >= 4 ? 2 : kGemvOcc))
gemv_fused_kernel(const GemvParams p, const int n_units) {
    using D = DT<T>;
    constexpr int VEC = D::VEC;
    constexpr int STEP = 32 * VEC;
    extern __shared__ __align__(16) unsigned char smem_raw[];
    T* xs = reinterpret_cast<T*>(smem_raw);
    __shared__ float red[32];
    __shared__ float s_val[kGemvWarps][RT];
    __shared__ int s_idx[kGemvWarps][RT];
    __shared__ int s_last;

    const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
    const int K = p.K;

    for (int r0 = 0; r0 < p.n_rows; r0 += RT) {
        const int rt = min(RT, p.n_rows - r0);
        __syncthreads();  // previous pass is done with xs
        stage_rows<T>(p, r0, rt, xs, red);
        __syncthreads();

        ArgBest best[RT];
#pragma unroll
        for (int i = 0; i < RT; ++i) best[i] = {-FLT_MAX, INT_MAX};

        for (int unit = blockIdx.x * kGemvWarps + warp; unit < n_units; unit += gridDim.x * kGemvWarps) {
            // which two weight rows this warp streams
            const T* wr0;
            const T* wr1;
            bool valid1 = true;
            if (EPI == EPI_QKV) {
                const int hd = p.head_dim, half = hd >> 1;
                const int head = unit / half, j = unit - head * half;
                const int m = head < p.n_heads ? 0 : (head < p.n_heads + p.n_kv_heads ? 1 : 2);
                const int hl = head - (m == 0 ? 0 : (m == 1 ? p.n_heads : p.n_heads + p.n_kv_heads));
                const T* w = static_cast<const T*>(p.w[m]);
                wr0 = w + (size_t)(hl * hd + j) * K;
                wr1 = wr0 + (size_t)half * K;
            } else if (EPI == EPI_GATEUP) {
                wr0 = static_cast<const T*>(p.w[0]) + (size_t)unit * K;
                wr1 = static_cast<const T*>(p.w[1]) + (size_t)unit * K;
            } else {
                wr0 = static_cast<const T*>(p.w[0]) + (size_t)(2 * unit) * K;
                valid1 = (2 * unit + 1) < p.n[0];
                wr1 = valid1 ? wr0 + K : wr0;
            }

            float acc0[RT], acc1[RT];
#pragma unroll
            for (int i = 0; i < RT; ++i) acc0[i] = acc1[i] = 0.f;

            int k = lane * VEC;
            for (; k + 3 * STEP < K; k += 4 * STEP) {
                uint4 a[4], b[4];
#pragma unroll
                for (int u = 0; u < 4; ++u) {
                    a[u] = ldg_stream_u4(wr0 + k + u * STEP);
                    b[u] = ldg_stream_u4(wr1 + k + u * STEP);
                }
#pragma unroll
                for (int u = 0; u < 4; ++u) {
#pragma unroll
                    for (int i = 0; i < RT; ++i) {
                        const uint4 xv = *reinterpret_cast<const uint4*>(xs + (size_t)i * K + k + u * STEP);
                        acc0[i] = dot_vec<T>(a[u], xv, acc0[i]);
                        acc1[i] = dot_vec<T>(b[u], xv, acc1[i]);
                    }
                }
            }
            for (; k < K; k += STEP) {
                const uint4 a = ldg_stream_u4(wr0 + k), b = ldg_stream_u4(wr1 + k);
#pragma unroll
                for (int i = 0; i < RT; ++i) {
                    const uint4 xv = *reinterpret_cast<const uint4*>(xs + (size_t)i * K + k);
                    acc0[i] = dot_vec<T>(a, xv, acc0[i]);
                    acc1[i] = dot_vec<T>(b, xv, acc1[i]);
                }
            }
#pragma unroll
            for (int i = 0; i < RT; ++i) {
                acc0[i] = warp_sum(acc0[i]);
                acc1[i] = warp_sum(acc1[i]);
            }

            // epilogue: lane i finishes batch row r0+i
#pragma unroll
            for (int i = 0; i < RT; ++i) {
                if (lane != i || i >= rt) continue;
                const int r = r0 + i;
                float v0 = acc0[i], v1 = acc1[i];
                if (EPI == EPI_QKV) {
                    epi_qkv<T>(p, unit, r, v0, v1);
                } else if (EPI == EPI_GATEUP) {
                    const float g = D::rnd(v0), u = D::rnd(v1);
                    static_cast<T*>(p.y)[(size_t)r * p.n[0] + unit] = D::from_f(D::rnd(silu_f(g)) * u);
                } else if (EPI == EPI_FC_GELU) {
                    T* y = static_cast<T*>(p.y) + (size_t)r * p.n[0] + 2 * unit;
                    y[0] = D::from_f(gelu_new_f(D::rnd(v0 + load_bias<T>(p.b[0], 2 * unit))));
                    if (valid1) y[1] = D::from_f(gelu_new_f(D::rnd(v1 + load_bias<T>(p.b[0], 2 * unit + 1))));
                } else if (EPI == EPI_RESIDUAL) {
                    T* y = static_cast<T*>(p.y) + (size_t)r * p.n[0] + 2 * unit;
                    v0 = D::rnd(v0 + load_bias<T>(p.b[0], 2 * unit));
                    if (p.accumulate) v0 = D::rnd(v0 + D::to_f(y[0]));
                    y[0] = D::from_f(v0);
                    if (valid1) {
                        v1 = D::rnd(v1 + load_bias<T>(p.b[0], 2 * unit + 1));
                        if (p.accumulate) v1 = D::rnd(v1 + D::to_f(y[1]));
                        y[1] = D::from_f(v1);
                    }
                } else {  // EPI_LOGITS: HF returns logits in the model dtype; argmax = first index of the max
                    v0 = D::rnd(v0 + load_bias<T>(p.b[0], 2 * unit));
                    if (valid1) v1 = D::rnd(v1 + load_bias<T>(p.b[0], 2 * unit + 1));
                    if (p.logits) {
                        float* lg = p.logits + (size_t)r * p.n[0] + 2 * unit;
                        lg[0] = v0;
                        if (valid1) lg[1] = v1;
                    }
                    arg_update(best[i], v0, 2 * unit);
                    if (valid1) arg_update(best[i], v1, 2 * unit + 1);
                }
            }
        }

        if (EPI == EPI_LOGITS && p.next_token) {
            // block-level argmax for rows r0..r0+rt, then publish this block's candidates
#pragma unroll
            for (int i = 0; i < RT; ++i) {
                if (lane == i) {
                    s_val[warp][i] = best[i].val;
                    s_idx[warp][i] = best[i].idx;
                }
            }
            __syncthreads();
            if (threadIdx.x < rt) {
                ArgBest b = {-FLT_MAX, INT_MAX};
                for (int w = 0; w < kGemvWarps; ++w) arg_update(b, s_val[w][threadIdx.x], s_idx[w][threadIdx.x]);
                p.arg_val[(size_t)blockIdx.x * p.n_rows + r0 + threadIdx.x] = b.val;
                p.arg_idx[(size_t)blockIdx.x * p.n_rows + r0 + threadIdx.x] = b.idx;
            }
        }
    }

    if (EPI == EPI_LOGITS && p.next_token) {
        // last block done: reduce every block's candidates (ties -> lowest index)
        __threadfence();
        __syncthreads();
        if (threadIdx.x == 0) {
            const unsigned ticket = atomicAdd(p.counter, 1u);
            s_last = (ticket == gridDim.x - 1);
        }
        __syncthreads();
        if (s_last) {
            __threadfence();
            for (int r = warp; r < p.n_rows; r += kGemvWarps) {
                ArgBest b = {-FLT_MAX, INT_MAX};
                for (int g = lane; g < (int)gridDim.x; g += 32)
                    arg_update(b, __ldcg(p.arg_val + (size_t)g * p.n_rows + r),
                               __ldcg(p.arg_idx + (size_t)g * p.n_rows + r));
#pragma unroll
                for (int o = 16; o > 0; o >>= 1) {
                    const float ov = __shfl_xor_sync(0xffffffffu, b.val, o);
                    const int oi = __shfl_xor_sync(0xffffffffu, b.idx, o);
                    arg_update(b, ov, oi);
                }
                if (lane == 0) p.next_token[r] = b.idx;
            }
            if (threadIdx.x == 0) *p.counter = 0u;
        }
    }
}

// ---- host side --------------------------------------------------------------------------------------

static int units_of(const GemvParams& p) {
    switch (p.epi) {
        case EPI_QKV: return (p.n_heads + 2 * p.n_kv_heads) * (p.head_dim / 2);
        case EPI_GATEUP: return p.n[0];
        default: return (p.n[0] + 1) / 2;
    }
}

static int pick_rt(int dtype, const GemvParams& p) {
    const size_t esz = dtype == 0 ? 2 : 4;
    const size_t row = (size_t)p.K * esz;
    if (p.n_rows >= 4 && 4 * row <= 96 * 1024) return 4;
    if (p.n_rows >= 2 && 2 * row <= 96 * 1024) return 2;
    return 1;
}

static int grid_for(int n_units, int num_sms, int rt) {
    const int needed = (n_units + kGemvWarps - 1) / kGemvWarps;
    const int cap = num_sms * (rt >= 4 ? 2 : kGemvOcc);
    const int iters = (needed + cap - 1) / cap;
    return (needed + iters - 1) / iters;
}

int gemv_grid_size(int dtype, const GemvParams& p, int num_sms) {
    return grid_for(units_of(p), num_sms, pick_rt(dtype, p));
}

constexpr int kGemvMaxSmem = 227 * 1024 - 2048;  // dynamic part; static shared memory takes the rest

template <typename T, int RT, int EPI>
static cudaError_t launch_inst(const GemvParams& p, int n_units, int grid, size_t smem, cudaStream_t s) {
    gemv_fused_kernel<T, RT, EPI><<<grid, kGemvThreads, smem, s>>>(p, n_units);
    count_launch();
    return cudaGetLastError();
}

template <typename T, int RT, int EPI>
static cudaError_t set_attr() {
    return cudaFuncSetAttribute(gemv_fused_kernel<T, RT, EPI>, cudaFuncAttributeMaxDynamicSharedMemorySize,
                                kGemvMaxSmem);
}

template <typename T, int RT>
static cudaError_t set_attr_epi() {
    cudaError_t e;
    if ((e = set_attr<T, RT, EPI_RESIDUAL>()) != cudaSuccess) return e;
    if ((e = set_attr<T, RT, EPI_QKV>()) != cudaSuccess) return e;
    if ((e = set_attr<T, RT, EPI_GATEUP>()) != cudaSuccess) return e;
    if ((e = set_attr<T, RT, EPI_FC_GELU>()) != cudaSuccess) return e;
    return set_attr<T, RT, EPI_LOGITS>();
}

// Opt every instantiation into large dynamic shared memory.  Called once per device from
// dllm_shard_create / the op entry points, never during stream capture.
cudaError_t gemv_init_device() {
    cudaError_t e;
    if ((e = set_attr_epi<bf16, 1>()) != cudaSuccess) return e;
    if ((e = set_attr_epi<bf16, 2>()) != cudaSuccess) return e;
    if ((e = set_attr_epi<bf16, 4>()) != cudaSuccess) return e;
    if ((e = set_attr_epi<float, 1>()) != cudaSuccess) return e;
    if ((e = set_attr_epi<float, 2>()) != cudaSuccess) return e;
    return set_attr_epi<float, 4>();
}

template <typename T, int RT>
static cudaError_t launch_epi(const GemvParams& p, int n_units, int grid, size_t smem, cudaStream_t s) {
    switch (p.epi) {
        case EPI_RESIDUAL: return launch_inst<T, RT, EPI_RESIDUAL>(p, n_units, grid, smem, s);
        case EPI_QKV: return launch_inst<T, RT, EPI_QKV>(p, n_units, grid, smem, s);
        case EPI_GATEUP: return launch_inst<T, RT, EPI_GATEUP>(p, n_units, grid, smem, s);
        case EPI_FC_GELU: return launch_inst<T, RT, EPI_FC_GELU>(p, n_units, grid, smem, s);
        case EPI_LOGITS: return launch_inst<T, RT, EPI_LOGITS>(p, n_units, grid, smem, s);
    }
    return cudaErrorInvalidValue;
}

template <typename T>
static cudaError_t launch_rt(const GemvParams& p, int rt, int n_units, int grid, size_t smem, cudaStream_t s) {
    switch (rt) {
        case 1: return launch_epi<T, 1>(p, n_units, grid, smem, s);
        case 2: return launch_epi<T, 2>(p, n_units, grid, smem, s);
        default: return launch_epi<T, 4>(p, n_units, grid, smem, s);
    }
}

cudaError_t launch_gemv(int dtype, const GemvParams& p, int num_sms, cudaStream_t s) {
    const size_t esz = dtype == 0 ? 2 : 4;
    const int vec = dtype == 0 ? 8 : 4;
    if (p.K % vec != 0 || p.n_rows <= 0) return cudaErrorInvalidValue;
    const int rt = pick_rt(dtype, p);
    const size_t smem = (size_t)rt * p.K * esz;
    if (smem > (size_t)kGemvMaxSmem) return cudaErrorInvalidValue;
    const int n_units = units_of(p);
    const int grid = grid_for(n_units, num_sms, rt);
    return dtype == 0 ? launch_rt<bf16>(p, rt, n_units, grid, smem, s)
                      : launch_rt<float>(p, rt, n_units, grid, smem, s);
}

// ---- small kernels: embedding, decode-loop advance, the reference's placeholder matmul ---------------

template <typename T>
__global__ void embed_kernel(const EmbedParams p) {
    pdl_launch_dependents();   // the next kernel of the step may be placed as soon as every CTA of this one runs
    pdl_wait();                // ... and this one touches global memory only once its predecessor has completed
    using D = DT<T>;
    constexpr int VEC = D::VEC;
    const int r = blockIdx.x;
    int id = p.ids[r];
    id = min(max(id, 0), p.vocab - 1);
    const T* te = static_cast<const T*>(p.wte) + (size_t)id * p.hidden;
    T* out = static_cast<T*>(p.out) + (size_t)r * p.hidden;
    const T* pe = p.wpe ? static_cast<const T*>(p.wpe) + (size_t)p.positions[r] * p.hidden : nullptr;
    for (int k = threadIdx.x * VEC; k < p.hidden; k += blockDim.x * VEC) {
        uint4 v = ld_u4(te + k);
        if (pe) {  // GPT-2: wte[id] + wpe[pos], rounded once
            float a[VEC], b[VEC];
            unpack<T>(v, a);
            unpack<T>(ld_u4(pe + k), b);
#pragma unroll
            for (int j = 0; j < VEC; ++j) a[j] += b[j];
            v = pack<T>(a);
        }
        *reinterpret_cast<uint4*>(out + k) = v;
    }
}

cudaError_t launch_embed(int dtype, const EmbedParams& p, cudaStream_t s) {
    if (p.hidden % (dtype == 0 ? 8 : 4) != 0) return cudaErrorInvalidValue;
    const cudaError_t e = dtype == 0 ? launch_pdl(embed_kernel<bf16>, dim3(p.n_rows), dim3(128), 0, s, p)
                                     : launch_pdl(embed_kernel<float>, dim3(p.n_rows), dim3(128), 0, s, p);
    count_launch();
    return e;
}

// Decode-loop bookkeeping, on the device so that no step needs the host.
//   begin  : gather the state the previous call left at its logit rows into a dense batch; pos_add = 1 after a
//            forward (the pending token sits one past the processed row), 0 after a loop (already advanced)
//   advance: feed the sampled token back, move every position on by one, log the token
__global__ void decode_begin_kernel(int32_t* ids, int32_t* positions, int32_t* seq_ids, const int32_t* next_token,
                                    const int32_t* logit_rows, int n, int32_t* step_ctr, int pos_add) {
    pdl_launch_dependents();   // the next kernel of the step may be placed as soon as every CTA of this one runs
    pdl_wait();                // ... and this one touches global memory only once its predecessor has completed
    const int i = threadIdx.x;
    int t = 0, ps = 0, sq = 0;
    if (i < n) {
        const int src = logit_rows[i];
        t = next_token[i];
        ps = positions[src] + pos_add;
        sq = seq_ids[src];
    }
    __syncthreads();
    if (i < n) {
        ids[i] = t;
        positions[i] = ps;
        seq_ids[i] = sq;
    }
    if (i == 0) *step_ctr = 0;
}

cudaError_t launch_decode_begin(int32_t* ids, int32_t* positions, int32_t* seq_ids, const int32_t* next_token,
                                const int32_t* logit_rows, int32_t n, int32_t* step_ctr, int32_t pos_add,
                                cudaStream_t s) {
    if (n <= 0 || n > 1024) return cudaErrorInvalidValue;
    const cudaError_t e = launch_pdl(decode_begin_kernel, dim3(1), dim3(1024), 0, s, ids, positions, seq_ids, next_token, logit_rows, n, step_ctr, pos_add);
    count_launch();
    return e;
}

__global__ void advance_kernel(int32_t* ids, int32_t* positions, const int32_t* next_token, int32_t* tok_ring,
                               int32_t* step_ctr, int n_rows) {
    pdl_launch_dependents();   // the next kernel of the step may be placed as soon as every CTA of this one runs
    pdl_wait();                // ... and this one touches global memory only once its predecessor has completed
    const int r = threadIdx.x;
    const int step = *step_ctr;
    __syncthreads();
    if (r < n_rows) {
        const int t = next_token[r];
        ids[r] = t;
        positions[r] += 1;
        tok_ring[(size_t)step * n_rows + r] = t;
    }
    if (r == 0) *step_ctr = step + 1;
}

cudaError_t launch_advance(int32_t* ids, int32_t* positions, const int32_t* next_token, int32_t* tok_ring,
                           int32_t* step_ctr, int32_t n_rows, cudaStream_t s) {
    if (n_rows <= 0 || n_rows > 1024) return cudaErrorInvalidValue;
    const cudaError_t e = launch_pdl(advance_kernel, dim3(1), dim3(1024), 0, s, ids, positions, next_token, tok_ring, step_ctr, n_rows);
    count_launch();
    return e;
}

// out[m, n] = sum_k in[m, k] * param[k, n]   (reference src/worker/node.py:31, fp32)
__global__ void legacy_matmul_kernel(const float* __restrict__ in, const float* __restrict__ param,
                                     float* __restrict__ out, int m, int k, int n) {
    const int col = blockIdx.x * blockDim.x + threadIdx.x;
    const int row = blockIdx.y;
    if (col >= n || row >= m) return;
    float acc = 0.f;
    for (int i = 0; i < k; ++i) acc = fmaf(in[(size_t)row * k + i], param[(size_t)i * n + col], acc);
    out[(size_t)row * n + col] = acc;
}

cudaError_t launch_legacy_matmul(const float* in, const float* param, float* out, int m, int k, int n,
                                 cudaStream_t s) {
    if (m <= 0 || n <= 0 || k < 0) return cudaErrorInvalidValue;
    dim3 grid((n + 127) / 128, m);
    legacy_matmul_kernel<<<grid, 128, 0, s>>>(in, param, out, m, k, n);
    count_launch();
    return cudaGetLastError();
}

}  // namespace dllm
