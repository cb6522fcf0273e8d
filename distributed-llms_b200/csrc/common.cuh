// Shared device helpers for the sm_100a shard kernels: dtype traits, 128-bit loads, reductions.
#pragma once

#include <cuda_bf16.h>
#include <cuda_runtime.h>
#include <stdint.h>

namespace dllm {

using bf16 = __nv_bfloat16;

// ---- programmatic dependent launch ---------------------------------------------------------------------------------
// The kernels of a step follow each other in one stream (or one CUDA graph).  Launched through launch_pdl, a kernel may
// become resident while its predecessor is still draining: it sets its barriers and tensor memory up -- the GEMMs also
// start streaming WEIGHTS, which no kernel of the step writes -- and only then waits (pdl_wait) for the predecessor to
// have completed and its stores to be visible.  Everything a kernel reads that an earlier kernel of the step wrote, and
// every global store, comes after pdl_wait.  pdl_launch_dependents lets the runtime place the NEXT kernel as soon as every
// CTA of this one has started.  Both instructions are no-ops in a kernel launched the ordinary way.
__device__ __forceinline__ void pdl_wait() { asm volatile("griddepcontrol.wait;" ::: "memory"); }
__device__ __forceinline__ void pdl_launch_dependents() { asm volatile("griddepcontrol.launch_dependents;" ::: "memory"); }

bool pdl_enabled();   // DLLM_PDL=1 switches the launch attribute on; off by default: measured slower on B200 (see its definition)

// cudaLaunchKernelEx with the programmatic-serialization attribute (plus any the caller passes).  Only for kernels that
// call pdl_wait() before they touch what their predecessors wrote.
template <typename... KArgs, typename... Args>
cudaError_t launch_pdl(void (*kernel)(KArgs...), dim3 grid, dim3 block, size_t smem, cudaStream_t st, Args&&... args) {
    cudaLaunchConfig_t cfg{};
    cfg.gridDim = grid;
    cfg.blockDim = block;
    cfg.dynamicSmemBytes = smem;
    cfg.stream = st;
    cudaLaunchAttribute attr[1];
    attr[0].id = cudaLaunchAttributeProgrammaticStreamSerialization;
    attr[0].val.programmaticStreamSerializationAllowed = 1;
    cfg.attrs = attr;
    cfg.numAttrs = pdl_enabled() ? 1 : 0;
    return cudaLaunchKernelEx(&cfg, kernel, static_cast<KArgs>(args)...);
}

__device__ __forceinline__ float bf16_round(float v) { return __bfloat162float(__float2bfloat16_rn(v)); }

// Per-dtype traits.  VEC = elements in one 16-byte vector.  rnd() rounds an fp32 value to the storage
// type and back: the kernels round at the points where the HF bf16 model rounds (oracle/shard_oracle.py).
template <typename T>
struct DT;

template <>
struct DT<float> {
    static constexpr int VEC = 4;
    static constexpr bool kIsBf16 = false;
    static __device__ __forceinline__ float to_f(float v) { return v; }
    static __device__ __forceinline__ float from_f(float v) { return v; }
    static __device__ __forceinline__ float rnd(float v) { return v; }
};

template <>
struct DT<bf16> {
    static constexpr int VEC = 8;
    static constexpr bool kIsBf16 = true;
    static __device__ __forceinline__ float to_f(bf16 v) { return __bfloat162float(v); }
    static __device__ __forceinline__ bf16 from_f(float v) { return __float2bfloat16_rn(v); }
    static __device__ __forceinline__ float rnd(float v) { return bf16_round(v); }
};

// ---- 16-byte vector loads ------------------------------------------------------------------------

// Streaming load for weights: read-only path, do not allocate in L1 (each byte is used once).
__device__ __forceinline__ uint4 ldg_stream_u4(const void* p) {
    uint4 r;
    asm volatile("ld.global.nc.L1::no_allocate.v4.u32 {%0, %1, %2, %3}, [%4];"
                 : "=r"(r.x), "=r"(r.y), "=r"(r.z), "=r"(r.w)
                 : "l"(p));
    return r;
}

__device__ __forceinline__ uint4 ld_u4(const void* p) { return *reinterpret_cast<const uint4*>(p); }

// Unpack one raw 16-byte vector into VEC floats.
template <typename T>
__device__ __forceinline__ void unpack(const uint4& v, float (&out)[DT<T>::VEC]);

template <>
__device__ __forceinline__ void unpack<float>(const uint4& v, float (&out)[4]) {
    out[0] = __uint_as_float(v.x);
    out[1] = __uint_as_float(v.y);
    out[2] = __uint_as_float(v.z);
    out[3] = __uint_as_float(v.w);
}

template <>
__device__ __forceinline__ void unpack<bf16>(const uint4& v, float (&out)[8]) {
    out[0] = __uint_as_float(v.x << 16);
    out[1] = __uint_as_float(v.x & 0xffff0000u);
    out[2] = __uint_as_float(v.y << 16);
    out[3] = __uint_as_float(v.y & 0xffff0000u);
    out[4] = __uint_as_float(v.z << 16);
    out[5] = __uint_as_float(v.z & 0xffff0000u);
    out[6] = __uint_as_float(v.w << 16);
    out[7] = __uint_as_float(v.w & 0xffff0000u);
}

template <typename T>
__device__ __forceinline__ float dot_vec(const uint4& a, const uint4& b, float acc) {
    float fa[DT<T>::VEC], fb[DT<T>::VEC];
    unpack<T>(a, fa);
    unpack<T>(b, fb);
#pragma unroll
    for (int i = 0; i < DT<T>::VEC; ++i) acc = fmaf(fa[i], fb[i], acc);
    return acc;
}

// ---- reductions ----------------------------------------------------------------------------------

__device__ __forceinline__ float warp_sum(float v) {
#pragma unroll
    for (int o = 16; o > 0; o >>= 1) v += __shfl_xor_sync(0xffffffffu, v, o);
    return v;
}

__device__ __forceinline__ float warp_max(float v) {
#pragma unroll
    for (int o = 16; o > 0; o >>= 1) v = fmaxf(v, __shfl_xor_sync(0xffffffffu, v, o));
    return v;
}

// Sum over the whole block; every thread gets the result.  `red` = at least 32 floats of shared memory.
__device__ __forceinline__ float block_sum(float v, float* red) {
    const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5, nwarps = (blockDim.x + 31) >> 5;
    v = warp_sum(v);
    __syncthreads();  // protect `red` against the previous use
    if (lane == 0) red[warp] = v;
    __syncthreads();
    float t = (lane < nwarps) ? red[lane] : 0.f;
    return warp_sum(t);
}

}  // namespace dllm
