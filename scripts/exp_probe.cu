// Experiment (not part of the product): which of the mechanisms round 2 wants exist on this box?
// One process, two GPUs.  Every stage prints one line and never aborts the rest.
//   1. device attributes: VMM, POSIX-fd / fabric handles, multicast, 64-bit stream memory operations
//   2. stream memory operations: cuStreamWriteValue64 into PEER memory, cuStreamWaitValue64 on local memory,
//      and an 8-byte DMA into peer memory as the alternative flag write
//   3. NVSwitch multicast: cuMulticastCreate over both GPUs, bind, map, multimem.st from GPU 0, read both copies;
//      bandwidth of a multimem.st broadcast against two unicast pushes
//   4. NVLink shapes round 1 did not measure: TMA bulk-copy PULL (peer global -> shared -> local global),
//      and mixed directions at once (GPU 0 pushes out while it pulls in; push+pull on one flow pair)
// build: nvcc -gencode arch=compute_100a,code=sm_100a -std=c++17 -O3 -o build/exp_probe scripts/exp_probe.cu -L/usr/local/cuda/lib64/stubs -lcuda
#include <cstdint>
#include <cstdio>
#include <cstdlib>
#include <cstring>
#include <vector>
#include <time.h>
#include <cuda.h>
#include <cuda_runtime.h>

#define CK(x) do { cudaError_t e_ = (x); if (e_ != cudaSuccess) { printf("CUDA error %s at %s:%d\n", cudaGetErrorString(e_), __FILE__, __LINE__); exit(1); } } while (0)
#define CU(x) ([&] { CUresult r_ = (x); if (r_ != CUDA_SUCCESS) { const char* s_ = nullptr; cuGetErrorName(r_, &s_); printf("    %s -> %s (%d)\n", #x, s_ ? s_ : "?", (int)r_); } return r_; }())

__device__ __forceinline__ uint4 ld16(const void* p) {
    uint4 v;
    asm volatile("ld.global.L1::no_allocate.v4.b32 {%0,%1,%2,%3}, [%4];" : "=r"(v.x), "=r"(v.y), "=r"(v.z), "=r"(v.w) : "l"(p) : "memory");
    return v;
}
__device__ __forceinline__ void st16(void* p, uint4 v) {
    asm volatile("st.global.L1::no_allocate.v4.b32 [%0], {%1,%2,%3,%4};" ::"l"(p), "r"(v.x), "r"(v.y), "r"(v.z), "r"(v.w) : "memory");
}
__device__ __forceinline__ void mst16(void* p, uint4 v) {
    asm volatile("multimem.st.relaxed.sys.global.v4.f32 [%0], {%1,%2,%3,%4};" ::"l"(p), "r"(v.x), "r"(v.y), "r"(v.z), "r"(v.w) : "memory");
}

template <int MC>
__global__ void __launch_bounds__(512) copy16(uint4* dst, const uint4* src, size_t nvec) {
    constexpr int U = 8;
    const size_t tile = (size_t)blockDim.x * U;
    const size_t ntiles = nvec / tile;
    const size_t in_tile = (size_t)(threadIdx.x >> 5) * (32 * U) + (threadIdx.x & 31);
    for (size_t t = blockIdx.x; t < ntiles; t += gridDim.x) {
        const uint4* s = src + t * tile + in_tile;
        uint4* d = dst + t * tile + in_tile;
        uint4 v[U];
#pragma unroll
        for (int u = 0; u < U; ++u) v[u] = ld16(s + u * 32);
#pragma unroll
        for (int u = 0; u < U; ++u) { if (MC) mst16(d + u * 32, v[u]); else st16(d + u * 32, v[u]); }
    }
}

// ---- TMA ring copy (global -> shared -> global), one producer and one consumer lane per CTA
__device__ __forceinline__ uint32_t smem_u32(const void* p) { return (uint32_t)__cvta_generic_to_shared(p); }
__device__ __forceinline__ void mbar_init(uint64_t* bar, uint32_t count) { asm volatile("mbarrier.init.shared::cta.b64 [%0], %1;" ::"r"(smem_u32(bar)), "r"(count) : "memory"); }
__device__ __forceinline__ void mbar_expect(uint64_t* bar, uint32_t bytes) { asm volatile("mbarrier.arrive.expect_tx.shared::cta.b64 _, [%0], %1;" ::"r"(smem_u32(bar)), "r"(bytes) : "memory"); }
__device__ __forceinline__ void mbar_arrive(uint64_t* bar) { asm volatile("mbarrier.arrive.shared::cta.b64 _, [%0];" ::"r"(smem_u32(bar)) : "memory"); }
__device__ __forceinline__ void mbar_wait(uint64_t* bar, uint32_t parity) {
    asm volatile("{\n\t.reg .pred p;\n\tW_%=:\n\tmbarrier.try_wait.parity.shared::cta.b64 p, [%0], %1;\n\t@p bra D_%=;\n\tbra W_%=;\n\tD_%=:\n\t}" ::"r"(smem_u32(bar)), "r"(parity) : "memory");
}
__global__ void __launch_bounds__(128, 1) tma_copy(uint8_t* dst, const uint8_t* src, size_t n, uint32_t chunk, uint32_t stages) {
    extern __shared__ __align__(128) uint8_t ring[];
    __shared__ __align__(8) uint64_t full[16], empty[16];
    if (threadIdx.x == 0) {
        for (uint32_t i = 0; i < stages; ++i) { mbar_init(&full[i], 1); mbar_init(&empty[i], 1); }
        asm volatile("fence.mbarrier_init.release.cluster;" ::: "memory");
        asm volatile("fence.proxy.async;" ::: "memory");
    }
    __syncthreads();
    const size_t nchunks = (n + chunk - 1) / chunk;
    const uint32_t part = blockIdx.x, nparts = gridDim.x;
    const size_t mine = nchunks > part ? (nchunks - part + nparts - 1) / nparts : 0;
    if (threadIdx.x == 0) {
        for (size_t k = 0; k < mine; ++k) {
            const uint32_t st = k % stages, round = k / stages;
            if (round > 0) mbar_wait(&empty[st], (round - 1) & 1);
            const size_t off = (part + k * nparts) * (size_t)chunk;
            const uint32_t bytes = (uint32_t)(n - off < chunk ? n - off : chunk);
            mbar_expect(&full[st], bytes);
            asm volatile("cp.async.bulk.shared::cluster.global.mbarrier::complete_tx::bytes [%0], [%1], %2, [%3];" ::"r"(smem_u32(ring + (size_t)st * chunk)), "l"(src + off), "r"(bytes), "r"(smem_u32(&full[st])) : "memory");
        }
    } else if (threadIdx.x == 32) {
        for (size_t k = 0; k < mine; ++k) {
            const uint32_t st = k % stages, round = k / stages;
            mbar_wait(&full[st], round & 1);
            asm volatile("fence.proxy.async;" ::: "memory");
            const size_t off = (part + k * nparts) * (size_t)chunk;
            const uint32_t bytes = (uint32_t)(n - off < chunk ? n - off : chunk);
            asm volatile("cp.async.bulk.global.shared::cta.bulk_group [%0], [%1], %2;" ::"l"(dst + off), "r"(smem_u32(ring + (size_t)st * chunk)), "r"(bytes) : "memory");
            asm volatile("cp.async.bulk.commit_group;" ::: "memory");
            if (k > 0) { asm volatile("cp.async.bulk.wait_group.read 1;" ::: "memory"); mbar_arrive(&empty[(k - 1) % stages]); }
        }
        asm volatile("cp.async.bulk.wait_group 0;" ::: "memory");
    }
}

static cudaStream_t s0, s1;

template <class F>
static double time_one(int dev, cudaStream_t s, int iters, const char* what, size_t bytes, F fn) {
    CK(cudaSetDevice(dev));
    cudaEvent_t e0, e1;
    CK(cudaEventCreate(&e0)); CK(cudaEventCreate(&e1));
    for (int i = 0; i < 3; ++i) fn();
    CK(cudaStreamSynchronize(s));
    CK(cudaEventRecord(e0, s));
    for (int i = 0; i < iters; ++i) fn();
    CK(cudaEventRecord(e1, s));
    CK(cudaStreamSynchronize(s));
    float ms;
    CK(cudaEventElapsedTime(&ms, e0, e1));
    const double gbs = (double)bytes * iters / (ms * 1e-3) / 1e9;
    printf("  %-62s %8.1f GB/s  %8.1f us\n", what, gbs, ms * 1e3 / iters);
    fflush(stdout);
    return gbs;
}

// f0 runs on GPU 0 / stream s0, f1 on GPU 1 / stream s1, concurrently; GB/s = flows * bytes / max(time)
template <class F0, class F1>
static void both(const char* what, size_t bytes, double flows, int iters, F0 f0, F1 f1) {
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
    const float ms = m0 > m1 ? m0 : m1;
    printf("  %-62s %8.1f GB/s total (%.1f per flow; GPU0 %.1f us, GPU1 %.1f us)\n", what, flows * bytes * iters / (ms * 1e-3) / 1e9,
           (double)bytes * iters / (ms * 1e-3) / 1e9, m0 * 1e3 / iters, m1 * 1e3 / iters);
    fflush(stdout);
}

static int attr(CUdevice d, CUdevice_attribute a) {
    int v = -1;
    if (cuDeviceGetAttribute(&v, a, d) != CUDA_SUCCESS) v = -1;
    return v;
}

int main(int argc, char** argv) {
    const size_t bytes = argc > 1 ? strtoull(argv[1], nullptr, 0) : (256ull << 20);
    int n = 0;
    CK(cudaGetDeviceCount(&n));
    if (n < 2) { printf("need 2 GPUs\n"); return 1; }
    CK(cudaSetDevice(0)); CK(cudaFree(0)); CK(cudaDeviceEnablePeerAccess(1, 0)); CK(cudaStreamCreateWithFlags(&s0, cudaStreamNonBlocking));
    CK(cudaSetDevice(1)); CK(cudaFree(0)); CK(cudaDeviceEnablePeerAccess(0, 0)); CK(cudaStreamCreateWithFlags(&s1, cudaStreamNonBlocking));

    // ---- 1. attributes
    printf("== 1. device attributes\n");
    CUdevice d0, d1;
    CU(cuDeviceGet(&d0, 0)); CU(cuDeviceGet(&d1, 1));
    printf("  VMM %d  posix-fd handle %d  fabric handle %d  multicast %d  64-bit stream memops %d  wait-value NOR %d  GPUDirect RDMA w/ VMM %d\n",
           attr(d0, CU_DEVICE_ATTRIBUTE_VIRTUAL_MEMORY_MANAGEMENT_SUPPORTED), attr(d0, CU_DEVICE_ATTRIBUTE_HANDLE_TYPE_POSIX_FILE_DESCRIPTOR_SUPPORTED),
           attr(d0, CU_DEVICE_ATTRIBUTE_HANDLE_TYPE_FABRIC_SUPPORTED), attr(d0, CU_DEVICE_ATTRIBUTE_MULTICAST_SUPPORTED),
           attr(d0, CU_DEVICE_ATTRIBUTE_CAN_USE_64_BIT_STREAM_MEM_OPS), attr(d0, CU_DEVICE_ATTRIBUTE_CAN_USE_STREAM_WAIT_VALUE_NOR),
           attr(d0, CU_DEVICE_ATTRIBUTE_GPU_DIRECT_RDMA_WITH_CUDA_VMM_SUPPORTED));

    uint8_t *b0, *b1, *c0, *c1;
    CK(cudaSetDevice(0)); CK(cudaMalloc(&b0, bytes)); CK(cudaMalloc(&c0, bytes)); CK(cudaMemset(b0, 1, bytes));
    CK(cudaSetDevice(1)); CK(cudaMalloc(&b1, bytes)); CK(cudaMalloc(&c1, bytes)); CK(cudaMemset(b1, 2, bytes));
    CK(cudaDeviceSynchronize());

    // ---- 2. stream memory operations
    printf("== 2. stream memory operations\n");
    {
        uint64_t *f0, *f1, h = 0;
        CK(cudaSetDevice(0)); CK(cudaMalloc(&f0, 64)); CK(cudaMemset(f0, 0, 64));
        CK(cudaSetDevice(1)); CK(cudaMalloc(&f1, 64)); CK(cudaMemset(f1, 0, 64));
        CK(cudaDeviceSynchronize());
        CK(cudaSetDevice(0));
        CUresult r = CU(cuStreamWriteValue64((CUstream)s0, (CUdeviceptr)f1, 0x1122334455667788ull, 0));
        cudaError_t e = cudaStreamSynchronize(s0);
        CK(cudaMemcpy(&h, f1, 8, cudaMemcpyDefault));
        printf("  cuStreamWriteValue64 from GPU0's stream into GPU1 memory: enqueue %d, sync %s, value %s\n", (int)r, cudaGetErrorString(e),
               h == 0x1122334455667788ull ? "arrived" : "NOT there");
        // wait on a local flag that a peer's memop write releases
        r = CU(cuStreamWaitValue64((CUstream)s0, (CUdeviceptr)f0, 7, CU_STREAM_WAIT_VALUE_EQ));
        CK(cudaSetDevice(1));
        CUresult r2 = CU(cuStreamWriteValue64((CUstream)s1, (CUdeviceptr)f0, 7, 0));
        CK(cudaStreamSynchronize(s1));
        CK(cudaSetDevice(0));
        bool released = false;
        for (int i = 0; i < 2000 && !released; ++i) {  // never hang the box on a wait nobody releases
            released = cudaStreamQuery(s0) == cudaSuccess;
            if (!released) { struct timespec ts = {0, 1000000}; nanosleep(&ts, nullptr); }
        }
        if (!released) {
            uint64_t* pin7;
            CK(cudaMallocHost(&pin7, 8));
            *pin7 = 7;
            cudaStream_t rescue;
            CK(cudaStreamCreateWithFlags(&rescue, cudaStreamNonBlocking));
            CK(cudaMemcpyAsync(f0, pin7, 8, cudaMemcpyDefault, rescue));
            CK(cudaStreamSynchronize(rescue));
        }
        e = cudaStreamSynchronize(s0);
        printf("  cuStreamWaitValue64 EQ on local memory released by a peer's write: enqueue %d/%d, %s, sync %s\n", (int)r, (int)r2,
               released ? "released by the memop" : "NOT released (rescued by a DMA write)", cudaGetErrorString(e));
        // 8-byte DMA into peer memory from pinned host
        uint64_t* pin;
        CK(cudaMallocHost(&pin, 64));
        pin[0] = 0xabcdefull;
        CK(cudaMemcpyAsync(f1 + 1, pin, 8, cudaMemcpyDefault, s0));
        CK(cudaStreamSynchronize(s0));
        CK(cudaMemcpy(&h, f1 + 1, 8, cudaMemcpyDefault));
        printf("  8-byte DMA pinned host -> GPU1 memory on GPU0's stream: %s\n", h == 0xabcdefull ? "arrived" : "NOT there");
        // cost of a chain: write + wait(already true) + write, per iteration
        CK(cudaMemset(f0, 0, 64));
        CK(cudaDeviceSynchronize());
        cudaEvent_t e0, e1;
        CK(cudaEventCreate(&e0)); CK(cudaEventCreate(&e1));
        CK(cudaEventRecord(e0, s0));
        for (int i = 0; i < 200; ++i) {
            cuStreamWriteValue64((CUstream)s0, (CUdeviceptr)f1, i, 0);
            cuStreamWaitValue64((CUstream)s0, (CUdeviceptr)f0, 0, CU_STREAM_WAIT_VALUE_EQ);
            cuStreamWriteValue64((CUstream)s0, (CUdeviceptr)(f1 + 2), i, 0);
        }
        CK(cudaEventRecord(e1, s0));
        CK(cudaStreamSynchronize(s0));
        float ms;
        CK(cudaEventElapsedTime(&ms, e0, e1));
        printf("  write(peer) + wait(local, already true) + write(peer): %.2f us per chain\n", ms * 1e3 / 200);
    }

    // ---- 3. multicast
    printf("== 3. NVSwitch multicast\n");
    uint8_t *mc_ptr = nullptr, *uc_ptr[2] = {nullptr, nullptr};
    size_t mc_size = 0;
    if (attr(d0, CU_DEVICE_ATTRIBUTE_MULTICAST_SUPPORTED) == 1) {
        CUmulticastObjectProp mp = {};
        mp.numDevices = 2;
        mp.handleTypes = CU_MEM_HANDLE_TYPE_POSIX_FILE_DESCRIPTOR;
        mp.flags = 0;
        size_t gran = 0, min_gran = 0;
        mp.size = bytes;
        CU(cuMulticastGetGranularity(&min_gran, &mp, CU_MULTICAST_GRANULARITY_MINIMUM));
        CU(cuMulticastGetGranularity(&gran, &mp, CU_MULTICAST_GRANULARITY_RECOMMENDED));
        printf("  multicast granularity: minimum %zu, recommended %zu\n", min_gran, gran);
        if (gran) {
            mc_size = (bytes + gran - 1) / gran * gran;
            mp.size = mc_size;
            CUmemGenericAllocationHandle mc = 0, mem[2] = {0, 0};
            bool ok = CU(cuMulticastCreate(&mc, &mp)) == CUDA_SUCCESS;
            ok = ok && CU(cuMulticastAddDevice(mc, d0)) == CUDA_SUCCESS && CU(cuMulticastAddDevice(mc, d1)) == CUDA_SUCCESS;
            int fd = -1;
            if (ok) printf("  export of the multicast object as POSIX fd: %s\n", CU(cuMemExportToShareableHandle(&fd, mc, CU_MEM_HANDLE_TYPE_POSIX_FILE_DESCRIPTOR, 0)) == CUDA_SUCCESS ? "ok" : "failed");
            for (int d = 0; d < 2 && ok; ++d) {
                CUmemAllocationProp ap = {};
                ap.type = CU_MEM_ALLOCATION_TYPE_PINNED;
                ap.location.type = CU_MEM_LOCATION_TYPE_DEVICE;
                ap.location.id = d;
                ap.requestedHandleTypes = CU_MEM_HANDLE_TYPE_POSIX_FILE_DESCRIPTOR;
                size_t ag = 0;
                CU(cuMemGetAllocationGranularity(&ag, &ap, CU_MEM_ALLOC_GRANULARITY_RECOMMENDED));
                if (d == 0) printf("  allocation granularity (recommended): %zu\n", ag);
                ok = ok && CU(cuMemCreate(&mem[d], mc_size, &ap, 0)) == CUDA_SUCCESS;
                ok = ok && CU(cuMulticastBindMem(mc, 0, mem[d], 0, mc_size, 0)) == CUDA_SUCCESS;
                CUdeviceptr va = 0;
                ok = ok && CU(cuMemAddressReserve(&va, mc_size, gran, 0, 0)) == CUDA_SUCCESS;
                ok = ok && CU(cuMemMap(va, mc_size, 0, mem[d], 0)) == CUDA_SUCCESS;
                CUmemAccessDesc ad[2] = {};
                for (int k = 0; k < 2; ++k) { ad[k].location.type = CU_MEM_LOCATION_TYPE_DEVICE; ad[k].location.id = k; ad[k].flags = CU_MEM_ACCESS_FLAGS_PROT_READWRITE; }
                ok = ok && CU(cuMemSetAccess(va, mc_size, ad, 2)) == CUDA_SUCCESS;
                uc_ptr[d] = (uint8_t*)va;
            }
            if (ok) {
                CUdeviceptr va = 0;
                ok = CU(cuMemAddressReserve(&va, mc_size, gran, 0, 0)) == CUDA_SUCCESS;
                ok = ok && CU(cuMemMap(va, mc_size, 0, mc, 0)) == CUDA_SUCCESS;
                CUmemAccessDesc ad[2] = {};
                for (int k = 0; k < 2; ++k) { ad[k].location.type = CU_MEM_LOCATION_TYPE_DEVICE; ad[k].location.id = k; ad[k].flags = CU_MEM_ACCESS_FLAGS_PROT_READWRITE; }
                ok = ok && CU(cuMemSetAccess(va, mc_size, ad, 2)) == CUDA_SUCCESS;
                mc_ptr = ok ? (uint8_t*)va : nullptr;
            }
            printf("  multicast object over 2 GPUs, %zu bytes bound and mapped: %s\n", mc_size, mc_ptr ? "ok" : "FAILED");
            if (mc_ptr) {
                CK(cudaSetDevice(0));
                CK(cudaMemset(uc_ptr[0], 0, mc_size));
                CK(cudaMemset(uc_ptr[1], 0, mc_size));
                CK(cudaDeviceSynchronize());
                copy16<1><<<32, 512, 0, s0>>>((uint4*)mc_ptr, (const uint4*)b0, bytes / 16);
                cudaError_t e = cudaStreamSynchronize(s0);
                std::vector<uint8_t> h0(4096), h1(4096);
                CK(cudaMemcpy(h0.data(), uc_ptr[0] + bytes / 2, 4096, cudaMemcpyDefault));
                CK(cudaMemcpy(h1.data(), uc_ptr[1] + bytes / 2, 4096, cudaMemcpyDefault));
                bool g0 = true, g1 = true;
                for (int i = 0; i < 4096; ++i) { g0 = g0 && h0[i] == 1; g1 = g1 && h1[i] == 1; }
                printf("  multimem.st from GPU0: sync %s; copy on GPU0 %s, copy on GPU1 %s\n", cudaGetErrorString(e), g0 ? "written" : "NOT written", g1 ? "written" : "NOT written");
                for (int g : {16, 32, 64, 128})  {
                    char name[128];
                    snprintf(name, sizeof name, "multimem.st broadcast 0 -> {0,1} (bytes per destination) ctas=%d", g);
                    time_one(0, s0, 20, name, bytes, [&] { copy16<1><<<g, 512, 0, s0>>>((uint4*)mc_ptr, (const uint4*)b0, bytes / 16); });
                }
            }
        }
    } else {
        printf("  CU_DEVICE_ATTRIBUTE_MULTICAST_SUPPORTED is not 1: no multicast on this box\n");
    }

    // ---- 4. NVLink shapes
    printf("== 4. NVLink shapes (%zu bytes)\n", bytes);
    const size_t nv = bytes / 16;
    const int it = 20;
    time_one(0, s0, it, "push ld/st 0->1, 32 CTAs", bytes, [&] { copy16<0><<<32, 512, 0, s0>>>((uint4*)b1, (const uint4*)b0, nv); });
    time_one(1, s1, it, "pull ld/st 0->1, 64 CTAs", bytes, [&] { copy16<0><<<64, 512, 0, s1>>>((uint4*)b1, (const uint4*)b0, nv); });
    CK(cudaSetDevice(0)); CK(cudaFuncSetAttribute(tma_copy, cudaFuncAttributeMaxDynamicSharedMemorySize, 200 * 1024));
    CK(cudaSetDevice(1)); CK(cudaFuncSetAttribute(tma_copy, cudaFuncAttributeMaxDynamicSharedMemorySize, 200 * 1024));
    struct Geo { uint32_t chunk, stages; };
    for (Geo ge : {Geo{32768, 6}, Geo{65536, 3}, Geo{16384, 12}}) {
        for (int g : {16, 32, 64, 96, 148}) {
            char name[128];
            snprintf(name, sizeof name, "TMA pull 0->1 (kernel on GPU1) chunk=%u stages=%u ctas=%d", ge.chunk, ge.stages, g);
            time_one(1, s1, it, name, bytes, [&] { tma_copy<<<g, 128, ge.chunk * ge.stages, s1>>>(b1, b0, bytes, ge.chunk, ge.stages); });
        }
        for (int g : {32, 96, 148}) {
            char name[128];
            snprintf(name, sizeof name, "TMA push 0->1 (kernel on GPU0) chunk=%u stages=%u ctas=%d", ge.chunk, ge.stages, g);
            time_one(0, s0, it, name, bytes, [&] { tma_copy<<<g, 128, ge.chunk * ge.stages, s0>>>(b1, b0, bytes, ge.chunk, ge.stages); });
        }
    }
    // both directions at once, every pairing of mechanisms (flow A: 0 -> 1 into b1; flow B: 1 -> 0 into c0)
    both("bidir push(32)+push(32)", bytes, 2, it, [&] { copy16<0><<<32, 512, 0, s0>>>((uint4*)b1, (const uint4*)b0, nv); }, [&] { copy16<0><<<32, 512, 0, s1>>>((uint4*)c0, (const uint4*)c1, nv); });
    for (int g : {64, 96, 128})
    {
        char name[128];
        snprintf(name, sizeof name, "bidir pull(%d)+pull(%d)", g, g);
        both(name, bytes, 2, it, [&] { copy16<0><<<g, 512, 0, s0>>>((uint4*)c0, (const uint4*)c1, nv); }, [&] { copy16<0><<<g, 512, 0, s1>>>((uint4*)b1, (const uint4*)b0, nv); });
    }
    // GPU 0 moves both flows (pushes A, pulls B), GPU 1 idle — "one side pulls while it pushes"
    {
        cudaStream_t s0b;
        CK(cudaSetDevice(0)); CK(cudaStreamCreateWithFlags(&s0b, cudaStreamNonBlocking));
        for (int g : {32, 64}) {
            char name[128];
            snprintf(name, sizeof name, "GPU0 pushes A (32 CTAs) and pulls B (%d CTAs), GPU1 idle", g);
            cudaEvent_t fork, join;
            CK(cudaEventCreateWithFlags(&fork, cudaEventDisableTiming)); CK(cudaEventCreateWithFlags(&join, cudaEventDisableTiming));
            time_one(0, s0, it, name, 2 * bytes, [&] {
                CK(cudaEventRecord(fork, s0)); CK(cudaStreamWaitEvent(s0b, fork, 0));
                copy16<0><<<32, 512, 0, s0>>>((uint4*)b1, (const uint4*)b0, nv);
                copy16<0><<<g, 512, 0, s0b>>>((uint4*)c0, (const uint4*)c1, nv);
                CK(cudaEventRecord(join, s0b)); CK(cudaStreamWaitEvent(s0, join, 0));
            });
        }
    }
    both("bidir TMA pull(64, 32K x6)+TMA pull", bytes, 2, it,
         [&] { tma_copy<<<64, 128, 32768 * 6, s0>>>(c0, c1, bytes, 32768, 6); }, [&] { tma_copy<<<64, 128, 32768 * 6, s1>>>(b1, b0, bytes, 32768, 6); });
    both("bidir TMA pull(148, 32K x6)+TMA pull", bytes, 2, it,
         [&] { tma_copy<<<148, 128, 32768 * 6, s0>>>(c0, c1, bytes, 32768, 6); }, [&] { tma_copy<<<148, 128, 32768 * 6, s1>>>(b1, b0, bytes, 32768, 6); });
    both("bidir ce+ce", bytes, 2, it, [&] { CK(cudaMemcpyAsync(b1, b0, bytes, cudaMemcpyDefault, s0)); }, [&] { CK(cudaMemcpyAsync(c0, c1, bytes, cudaMemcpyDefault, s1)); });
    // the copy engine one way while SMs push the other way
    both("bidir ce (A, GPU0) + push ld/st 32 (B, GPU1)", bytes, 2, it, [&] { CK(cudaMemcpyAsync(b1, b0, bytes, cudaMemcpyDefault, s0)); },
         [&] { copy16<0><<<32, 512, 0, s1>>>((uint4*)c0, (const uint4*)c1, nv); });
    printf("done\n");
    return 0;
}
