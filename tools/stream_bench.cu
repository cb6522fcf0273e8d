// Microbenchmark: how fast can 148 persistent CTAs stream HBM through per-warp TMA-bulk rings?
// Variants: bulk-copy size, ring depth, warps per CTA, access order; plus an LDG.128 baseline.
//   nvcc -gencode arch=compute_100a,code=sm_100a -O3 -o stream_bench tools/stream_bench.cu
#include <cuda_runtime.h>
#include <stdint.h>
#include <stdio.h>
#include <stdlib.h>

__device__ __forceinline__ uint32_t smem_u32(const void* p) { return (uint32_t)__cvta_generic_to_shared(p); }
__device__ __forceinline__ void mbar_init(uint32_t bar, uint32_t c) { asm volatile("mbarrier.init.shared::cta.b64 [%0], %1;" ::"r"(bar), "r"(c)); }
__device__ __forceinline__ void mbar_expect_tx(uint32_t bar, uint32_t b) { asm volatile("mbarrier.arrive.expect_tx.shared::cta.b64 _, [%0], %1;" ::"r"(bar), "r"(b) : "memory"); }
__device__ __forceinline__ bool mbar_try(uint32_t bar, uint32_t par) {
    uint32_t ok;
    asm volatile("{\n\t.reg .pred p;\n\tmbarrier.try_wait.parity.shared::cta.b64 p, [%1], %2;\n\tselp.u32 %0, 1, 0, p;\n\t}" : "=r"(ok) : "r"(bar), "r"(par) : "memory");
    return ok;
}
__device__ __forceinline__ void bulk(uint32_t dst, const void* src, uint32_t bytes, uint32_t bar) {
    asm volatile("cp.async.bulk.shared::cluster.global.mbarrier::complete_tx::bytes [%0], [%1], %2, [%3];" ::"r"(dst), "l"(src), "r"(bytes), "r"(bar) : "memory");
}

// mode 0: warp w of CTA c streams chunks c*W+w, +total_warps, ...   (chip reads a moving contiguous window)
// mode 1: each warp streams its own contiguous region
__global__ void __launch_bounds__(512, 1) tma_stream(const unsigned char* src, size_t n_chunks, int chunk, int depth, int mode, unsigned* sink) {
    extern __shared__ __align__(128) unsigned char smem[];
    const int W = blockDim.x / 32, warp = threadIdx.x / 32, lane = threadIdx.x % 32;
    uint64_t* bars = reinterpret_cast<uint64_t*>(smem);
    unsigned char* ring = smem + 1024 + (size_t)warp * depth * chunk;
    const uint32_t bar0 = smem_u32(bars) + warp * depth * 8;
    if (lane == 0) for (int i = 0; i < depth; ++i) mbar_init(bar0 + i * 8, 1);
    asm volatile("fence.mbarrier_init.release.cluster;" ::: "memory");
    __syncthreads();
    const size_t gw = (size_t)blockIdx.x * W + warp, nw = (size_t)gridDim.x * W;
    const size_t per = n_chunks / nw;
    auto addr = [&](size_t i) { return src + (mode == 0 ? (i * nw + gw) : (gw * per + i)) * (size_t)chunk; };
    size_t issued = 0;
    for (; issued < (size_t)depth && issued < per; ++issued)
        if (lane == 0) { mbar_expect_tx(bar0 + issued * 8, chunk); bulk(smem_u32(ring) + issued * chunk, addr(issued), chunk, bar0 + issued * 8); }
    unsigned acc = 0;
    for (size_t i = 0; i < per; ++i) {
        const int slot = i % depth;
        while (!mbar_try(bar0 + slot * 8, (i / depth) & 1)) {}
        acc += *reinterpret_cast<const unsigned*>(ring + (size_t)slot * chunk + lane * 16);
        __syncwarp();
        if (issued < per) {
            if (lane == 0) { mbar_expect_tx(bar0 + slot * 8, chunk); bulk(smem_u32(ring) + slot * chunk, addr(issued), chunk, bar0 + slot * 8); }
            ++issued;
        }
    }
    if (acc == 0x12345678u) *sink = acc;
}

__global__ void __launch_bounds__(512, 1) ldg_stream(const uint4* src, size_t n_vec, int unroll8, unsigned* sink) {
    const size_t tid = (size_t)blockIdx.x * blockDim.x + threadIdx.x, nt = (size_t)gridDim.x * blockDim.x;
    unsigned acc = 0;
    for (size_t i = tid; i + 7 * nt < n_vec; i += 8 * nt) {
        uint4 v[8];
#pragma unroll
        for (int u = 0; u < 8; ++u) asm volatile("ld.global.nc.L1::no_allocate.v4.u32 {%0,%1,%2,%3}, [%4];" : "=r"(v[u].x), "=r"(v[u].y), "=r"(v[u].z), "=r"(v[u].w) : "l"(src + i + u * nt));
#pragma unroll
        for (int u = 0; u < 8; ++u) acc += v[u].x ^ v[u].w;
    }
    if (acc == 0x12345678u) *sink = acc;
}

int main() {
    const size_t bytes = 4ull << 30;
    unsigned char* d; unsigned* sink;
    cudaMalloc(&d, bytes); cudaMalloc(&sink, 4); cudaMemset(d, 1, bytes);
    cudaEvent_t e0, e1; cudaEventCreate(&e0); cudaEventCreate(&e1);
    int sms = 148; cudaDeviceGetAttribute(&sms, cudaDevAttrMultiProcessorCount, 0);
    cudaFuncSetAttribute(tma_stream, cudaFuncAttributeMaxDynamicSharedMemorySize, 220 * 1024);
    auto run_tma = [&](int warps, int chunk, int depth, int mode, int grid_mult) {
        const size_t smem = 1024 + (size_t)warps * depth * chunk;
        if (smem > 220 * 1024) return;
        const size_t n_chunks = bytes / chunk;
        float best = 1e9;
        for (int it = 0; it < 3; ++it) {
            cudaEventRecord(e0);
            tma_stream<<<sms * grid_mult, warps * 32, smem>>>(d, n_chunks, chunk, depth, mode, sink);
            cudaEventRecord(e1); cudaEventSynchronize(e1);
            float ms; cudaEventElapsedTime(&ms, e0, e1); if (ms < best) best = ms;
        }
        cudaError_t e = cudaGetLastError();
        printf("tma  warps=%2d chunk=%5d depth=%d mode=%d inflight/SM=%4zu KB : %7.1f GB/s %s\n", warps, chunk, depth, mode,
               (size_t)warps * depth * chunk / 1024, bytes / best / 1e6, e == cudaSuccess ? "" : cudaGetErrorString(e));
    };
    for (int mode = 0; mode < 2; ++mode)
        for (int chunk : {2048, 4096, 8192, 16384})
            for (int warps : {4, 8, 16})
                for (int depth : {2, 4, 6, 8}) run_tma(warps, chunk, depth, mode, 1);
    for (int threads : {256, 512})
        for (int mult : {1, 2, 4}) {
            float best = 1e9;
            for (int it = 0; it < 3; ++it) {
                cudaEventRecord(e0);
                ldg_stream<<<sms * mult, threads>>>((const uint4*)d, bytes / 16, 1, sink);
                cudaEventRecord(e1); cudaEventSynchronize(e1);
                float ms; cudaEventElapsedTime(&ms, e0, e1); if (ms < best) best = ms;
            }
            printf("ldg  threads=%d ctas/SM=%d : %7.1f GB/s\n", threads, mult, bytes / best / 1e6);
        }
    return 0;
}
