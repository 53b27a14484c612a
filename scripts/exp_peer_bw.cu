// Experiment (not part of the product): where is the NVLink ceiling between two B200s and
// which access shape reaches it?  One process, two GPUs, peer access enabled.
//   push: kernel on GPU0 loads local, stores to GPU1         pull: kernel on GPU1 loads GPU0, stores local
//   ce:   cudaMemcpyPeerAsync / cudaMemcpyAsync
// build: nvcc -gencode arch=compute_100a,code=sm_100a -O3 -o build/exp_peer_bw scripts/exp_peer_bw.cu
#include <cstdint>
#include <cstdio>
#include <cstdlib>
#include <cuda_runtime.h>

#define CK(x) do { cudaError_t e_ = (x); if (e_ != cudaSuccess) { printf("CUDA error %s at %s:%d\n", cudaGetErrorString(e_), __FILE__, __LINE__); exit(1); } } while (0)

struct alignas(32) v32 { uint32_t r[8]; };

template <int ST> __device__ __forceinline__ void st16(void* p, uint4 v) {
    if (ST == 0) asm volatile("st.global.v4.b32 [%0], {%1,%2,%3,%4};" ::"l"(p), "r"(v.x), "r"(v.y), "r"(v.z), "r"(v.w) : "memory");
    if (ST == 1) asm volatile("st.global.L1::no_allocate.v4.b32 [%0], {%1,%2,%3,%4};" ::"l"(p), "r"(v.x), "r"(v.y), "r"(v.z), "r"(v.w) : "memory");
    if (ST == 2) asm volatile("st.global.cs.v4.b32 [%0], {%1,%2,%3,%4};" ::"l"(p), "r"(v.x), "r"(v.y), "r"(v.z), "r"(v.w) : "memory");
    if (ST == 3) asm volatile("st.global.wt.v4.b32 [%0], {%1,%2,%3,%4};" ::"l"(p), "r"(v.x), "r"(v.y), "r"(v.z), "r"(v.w) : "memory");
}
__device__ __forceinline__ uint4 ld16(const void* p) {
    uint4 v;
    asm volatile("ld.global.L1::no_allocate.v4.b32 {%0,%1,%2,%3}, [%4];" : "=r"(v.x), "=r"(v.y), "=r"(v.z), "=r"(v.w) : "l"(p) : "memory");
    return v;
}
__device__ __forceinline__ v32 ld32(const void* p) {
    v32 v;
    asm volatile("ld.global.L1::no_allocate.v8.b32 {%0,%1,%2,%3,%4,%5,%6,%7}, [%8];"
                 : "=r"(v.r[0]), "=r"(v.r[1]), "=r"(v.r[2]), "=r"(v.r[3]), "=r"(v.r[4]), "=r"(v.r[5]), "=r"(v.r[6]), "=r"(v.r[7]) : "l"(p) : "memory");
    return v;
}
__device__ __forceinline__ void st32(void* p, const v32& v) {
    asm volatile("st.global.L1::no_allocate.v8.b32 [%0], {%1,%2,%3,%4,%5,%6,%7,%8};"
                 ::"l"(p), "r"(v.r[0]), "r"(v.r[1]), "r"(v.r[2]), "r"(v.r[3]), "r"(v.r[4]), "r"(v.r[5]), "r"(v.r[6]), "r"(v.r[7]) : "memory");
}

// LAYOUT 0: tiles dealt round-robin to CTAs; 1: each CTA owns one contiguous range
template <int UNROLL, int ST, int LAYOUT>
__global__ void __launch_bounds__(512) copy16(uint4* dst, const uint4* src, size_t nvec) {
    const size_t tile = (size_t)blockDim.x * UNROLL;
    const size_t ntiles = nvec / tile;
    size_t t0, t1, step;
    if (LAYOUT == 0) { t0 = blockIdx.x; t1 = ntiles; step = gridDim.x; }
    else { const size_t per = (ntiles + gridDim.x - 1) / gridDim.x; t0 = per * blockIdx.x; t1 = t0 + per < ntiles ? t0 + per : ntiles; step = 1; }
    for (size_t t = t0; t < t1; t += step) {
        const uint4* s = src + t * tile + threadIdx.x;
        uint4* d = dst + t * tile + threadIdx.x;
        uint4 v[UNROLL];
#pragma unroll
        for (int u = 0; u < UNROLL; ++u) v[u] = ld16(s + (size_t)u * blockDim.x);
#pragma unroll
        for (int u = 0; u < UNROLL; ++u) st16<ST>(d + (size_t)u * blockDim.x, v[u]);
    }
}
template <int UNROLL>
__global__ void __launch_bounds__(512) copy32(v32* dst, const v32* src, size_t nvec) {
    const size_t tile = (size_t)blockDim.x * UNROLL;
    const size_t ntiles = nvec / tile;
    for (size_t t = blockIdx.x; t < ntiles; t += gridDim.x) {
        const v32* s = src + t * tile + threadIdx.x;
        v32* d = dst + t * tile + threadIdx.x;
        v32 v[UNROLL];
#pragma unroll
        for (int u = 0; u < UNROLL; ++u) v[u] = ld32(s + (size_t)u * blockDim.x);
#pragma unroll
        for (int u = 0; u < UNROLL; ++u) st32(d + (size_t)u * blockDim.x, v[u]);
    }
}

static double time_it(cudaStream_t s, int iters, const char* what, size_t bytes, auto fn) {
    cudaEvent_t e0, e1;
    CK(cudaEventCreate(&e0));
    CK(cudaEventCreate(&e1));
    for (int i = 0; i < 3; ++i) fn();
    CK(cudaStreamSynchronize(s));
    CK(cudaEventRecord(e0, s));
    for (int i = 0; i < iters; ++i) fn();
    CK(cudaEventRecord(e1, s));
    CK(cudaStreamSynchronize(s));
    float ms;
    CK(cudaEventElapsedTime(&ms, e0, e1));
    double gbs = (double)bytes * iters / (ms * 1e-3) / 1e9;
    printf("%-58s %8.1f GB/s  %8.1f us\n", what, gbs, ms * 1e3 / iters);
    fflush(stdout);
    return gbs;
}

int main(int argc, char** argv) {
    size_t bytes = argc > 1 ? strtoull(argv[1], nullptr, 0) : (256ull << 20);
    int n = 0;
    CK(cudaGetDeviceCount(&n));
    if (n < 2) { printf("need 2 GPUs\n"); return 1; }
    uint8_t *b0, *b1, *c0, *c1;
    cudaStream_t s0, s1;
    CK(cudaSetDevice(0)); CK(cudaDeviceEnablePeerAccess(1, 0)); CK(cudaMalloc(&b0, bytes)); CK(cudaMalloc(&c0, bytes)); CK(cudaMemset(b0, 1, bytes)); CK(cudaStreamCreateWithFlags(&s0, cudaStreamNonBlocking));
    CK(cudaSetDevice(1)); CK(cudaDeviceEnablePeerAccess(0, 0)); CK(cudaMalloc(&b1, bytes)); CK(cudaMalloc(&c1, bytes)); CK(cudaMemset(b1, 2, bytes)); CK(cudaStreamCreateWithFlags(&s1, cudaStreamNonBlocking));
    CK(cudaDeviceSynchronize());
    const size_t nv16 = bytes / 16, nv32 = bytes / 32;
    const int iters = 20;
    char name[128];
    printf("message %zu bytes; link nominal 900 GB/s per direction\n", bytes);

    CK(cudaSetDevice(0));
    time_it(s0, iters, "ce  cudaMemcpyPeerAsync 0->1 (issued on GPU0 stream)", bytes, [&] { CK(cudaMemcpyPeerAsync(b1, 1, b0, 0, bytes, s0)); });
    time_it(s0, iters, "ce  cudaMemcpyAsync default 0->1 (GPU0 stream)", bytes, [&] { CK(cudaMemcpyAsync(b1, b0, bytes, cudaMemcpyDefault, s0)); });
    CK(cudaSetDevice(1));
    time_it(s1, iters, "ce  cudaMemcpyAsync default 0->1 (GPU1 stream, pull)", bytes, [&] { CK(cudaMemcpyAsync(b1, b0, bytes, cudaMemcpyDefault, s1)); });

    const int grids[] = {8, 16, 24, 32, 48, 64, 96, 128, 148, 296};
    for (int g : grids) {
        CK(cudaSetDevice(0));
        snprintf(name, sizeof name, "push v16 u8 st.na rr      ctas=%d", g);
        time_it(s0, iters, name, bytes, [&] { copy16<8, 1, 0><<<g, 512, 0, s0>>>((uint4*)b1, (const uint4*)b0, nv16); });
        snprintf(name, sizeof name, "push v16 u8 st.na contig  ctas=%d", g);
        time_it(s0, iters, name, bytes, [&] { copy16<8, 1, 1><<<g, 512, 0, s0>>>((uint4*)b1, (const uint4*)b0, nv16); });
        snprintf(name, sizeof name, "push v16 u4 st.na rr      ctas=%d", g);
        time_it(s0, iters, name, bytes, [&] { copy16<4, 1, 0><<<g, 512, 0, s0>>>((uint4*)b1, (const uint4*)b0, nv16); });
        snprintf(name, sizeof name, "push v32 u4 st.na rr      ctas=%d", g);
        time_it(s0, iters, name, bytes, [&] { copy32<4><<<g, 512, 0, s0>>>((v32*)b1, (const v32*)b0, nv32); });
        CK(cudaSetDevice(1));
        snprintf(name, sizeof name, "pull v16 u8 rr            ctas=%d", g);
        time_it(s1, iters, name, bytes, [&] { copy16<8, 1, 0><<<g, 512, 0, s1>>>((uint4*)b1, (const uint4*)b0, nv16); });
        snprintf(name, sizeof name, "pull v32 u4 rr            ctas=%d", g);
        time_it(s1, iters, name, bytes, [&] { copy32<4><<<g, 512, 0, s1>>>((v32*)b1, (const v32*)b0, nv32); });
    }
    CK(cudaSetDevice(0));
    for (int g : {16, 32, 48, 96}) {
        snprintf(name, sizeof name, "push v16 u8 st default    ctas=%d", g);
        time_it(s0, iters, name, bytes, [&] { copy16<8, 0, 0><<<g, 512, 0, s0>>>((uint4*)b1, (const uint4*)b0, nv16); });
        snprintf(name, sizeof name, "push v16 u8 st.cs         ctas=%d", g);
        time_it(s0, iters, name, bytes, [&] { copy16<8, 2, 0><<<g, 512, 0, s0>>>((uint4*)b1, (const uint4*)b0, nv16); });
        snprintf(name, sizeof name, "push v16 u8 st.wt         ctas=%d", g);
        time_it(s0, iters, name, bytes, [&] { copy16<8, 3, 0><<<g, 512, 0, s0>>>((uint4*)b1, (const uint4*)b0, nv16); });
    }
    // both directions at once: push from both sides / pull from both sides / CE both ways
    auto both = [&](const char* what, auto f0, auto f1) {
        cudaEvent_t a0, a1, z0, z1;
        CK(cudaSetDevice(0)); CK(cudaEventCreate(&a0)); CK(cudaEventCreate(&z0));
        CK(cudaSetDevice(1)); CK(cudaEventCreate(&a1)); CK(cudaEventCreate(&z1));
        for (int i = 0; i < 3; ++i) { CK(cudaSetDevice(0)); f0(); CK(cudaSetDevice(1)); f1(); }
        CK(cudaSetDevice(0)); CK(cudaStreamSynchronize(s0)); CK(cudaSetDevice(1)); CK(cudaStreamSynchronize(s1));
        CK(cudaSetDevice(0)); CK(cudaEventRecord(a0, s0)); CK(cudaSetDevice(1)); CK(cudaEventRecord(a1, s1));
        for (int i = 0; i < iters; ++i) { CK(cudaSetDevice(0)); f0(); CK(cudaSetDevice(1)); f1(); }
        CK(cudaSetDevice(0)); CK(cudaEventRecord(z0, s0)); CK(cudaSetDevice(1)); CK(cudaEventRecord(z1, s1));
        CK(cudaSetDevice(0)); CK(cudaStreamSynchronize(s0)); CK(cudaSetDevice(1)); CK(cudaStreamSynchronize(s1));
        float m0, m1;
        CK(cudaEventElapsedTime(&m0, a0, z0)); CK(cudaEventElapsedTime(&m1, a1, z1));
        float ms = m0 > m1 ? m0 : m1;
        printf("%-58s %8.1f GB/s total (%.1f per direction)\n", what, 2.0 * bytes * iters / (ms * 1e-3) / 1e9, (double)bytes * iters / (ms * 1e-3) / 1e9);
    };
    for (int g : {16, 32, 48}) {
        snprintf(name, sizeof name, "bidir push+push v16 u8 ctas=%d", g);
        both(name, [&] { copy16<8, 1, 0><<<g, 512, 0, s0>>>((uint4*)b1, (const uint4*)b0, nv16); },
             [&] { copy16<8, 1, 0><<<g, 512, 0, s1>>>((uint4*)c0, (const uint4*)c1, nv16); });
        snprintf(name, sizeof name, "bidir pull+pull v16 u8 ctas=%d", g);
        both(name, [&] { copy16<8, 1, 0><<<g, 512, 0, s0>>>((uint4*)c0, (const uint4*)c1, nv16); },
             [&] { copy16<8, 1, 0><<<g, 512, 0, s1>>>((uint4*)b1, (const uint4*)b0, nv16); });
    }
    both("bidir ce+ce", [&] { CK(cudaMemcpyAsync(b1, b0, bytes, cudaMemcpyDefault, s0)); }, [&] { CK(cudaMemcpyAsync(c0, c1, bytes, cudaMemcpyDefault, s1)); });
    return 0;
}
