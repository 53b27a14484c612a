// Profiling harness (not part of the product): the engine's REAL group kernels, launched with peer-mapped
// pointers from ONE process that drives two GPUs, so that `ncu` (which must not wrap a multi-rank command)
// can read NVLink, DRAM, L2 and LSU counters for each data path the selector chooses between.
//
//   push   p2p_group_kernel_ldst, mover role on GPU 0: local LDG.E.128 -> STG.E.128 into GPU 1's memory
//   pull   p2p_group_kernel_ldst, mover role on GPU 1: LDG.E.128 from GPU 0's memory -> local STG
//   tma    p2p_group_kernel_tma on GPU 0: cp.async.bulk local -> shared -> GPU 1's memory
//   tmapull p2p_group_kernel_tma on GPU 1: cp.async.bulk GPU 0's memory -> shared -> local
//   ll     eager path: OP_SEND_LL on GPU 0 writes flagged 16-byte lines into a ring in GPU 1's memory,
//          OP_RECV_LL on GPU 1 copies them out (the 8-byte latency path)
//
// The mover roles are launched as OP_COPY (the copy loop of a push / pull without the rendezvous flags, which
// need the peer's process): same kernel, same loop, same addresses as in production.
//
// build: nvcc -gencode arch=compute_100a,code=sm_100a -std=c++17 -O3 -lineinfo -Iinclude -Itest-nccl-p2p_b200/csrc \
//             -o build/prof_peer scripts/prof_peer.cu build/libb200p2p.a
// run:   build/prof_peer <push|pull|tma|tmapull|ll> [bytes] [ctas] [iters]
#include <cstdio>
#include <cstdlib>
#include <cstring>
#include <vector>
#include "p2p.h"
#include "p2p_internal.h"

using namespace p2p;

#define CK(x) do { cudaError_t e_ = (x); if (e_ != cudaSuccess) { printf("CUDA error %s at %s:%d\n", cudaGetErrorString(e_), __FILE__, __LINE__); exit(1); } } while (0)

int main(int argc, char** argv) {
    const char* mode = argc > 1 ? argv[1] : "pull";
    size_t bytes = argc > 2 ? strtoull(argv[2], nullptr, 0) : (64ull << 20);
    int ctas = argc > 3 ? atoi(argv[3]) : 0;
    const int iters = argc > 4 ? atoi(argv[4]) : 10;
    int n = 0;
    CK(cudaGetDeviceCount(&n));
    if (n < 2) { printf("need 2 GPUs\n"); return 1; }
    uint8_t *b0, *b1, *ring1;
    uint32_t* arrive1;
    uint64_t *ack0;
    cudaStream_t s0, s1;
    CK(cudaSetDevice(0)); CK(cudaDeviceEnablePeerAccess(1, 0)); CK(cudaMalloc(&b0, bytes + 256)); CK(cudaMemset(b0, 0x5a, bytes + 256));
    CK(cudaMalloc(&ack0, 256)); CK(cudaMemset(ack0, 0, 256));
    CK(cudaStreamCreateWithFlags(&s0, cudaStreamNonBlocking)); CK(configure_kernels(200 * 1024));
    CK(cudaSetDevice(1)); CK(cudaDeviceEnablePeerAccess(0, 0)); CK(cudaMalloc(&b1, bytes + 256)); CK(cudaMemset(b1, 0, bytes + 256));
    CK(cudaMalloc(&ring1, 2 << 20)); CK(cudaMemset(ring1, 0, 2 << 20)); CK(cudaMalloc(&arrive1, 256)); CK(cudaMemset(arrive1, 0, 256));
    CK(cudaStreamCreateWithFlags(&s1, cudaStreamNonBlocking)); CK(configure_kernels(200 * 1024));
    CK(cudaDeviceSynchronize());

    GroupArgs a;
    memset(&a, 0, sizeof a);
    a.h.nops = 1;
    a.h.tma_chunk = 64 * 1024;   // the engine's geometry for launches that cross NVLink
    a.h.tma_stages = 3;
    a.h.timeout_ns = 2000000000ull;
    OpDesc& op = a.ops[0];
    op.nbytes = bytes;
    op.seq = 1;

    int dev = 0, flavour = FLAVOUR_LDST;
    cudaStream_t s = s0;
    if (!strcmp(mode, "push") || !strcmp(mode, "tma")) {
        if (!ctas) ctas = 32;
        op.kind = OP_COPY; op.local = (uint64_t)b0; op.aux = (uint64_t)b1; op.nblocks = ctas;
        if (!strcmp(mode, "tma")) flavour = FLAVOUR_TMA;
    } else if (!strcmp(mode, "pull") || !strcmp(mode, "tmapull")) {
        if (!ctas) ctas = 64;
        dev = 1; s = s1;
        op.kind = OP_COPY; op.local = (uint64_t)b0; op.aux = (uint64_t)b1; op.nblocks = ctas;
        if (!strcmp(mode, "tmapull")) flavour = FLAVOUR_TMA;
    } else if (!strcmp(mode, "ll")) {
        if (bytes > (512 << 10)) bytes = 8;
        op.nbytes = bytes;
        ctas = 1;
    } else {
        printf("unknown mode %s\n", mode);
        return 2;
    }

    cudaEvent_t e0, e1;
    CK(cudaSetDevice(!strcmp(mode, "ll") ? 1 : dev));  // events live on the device whose stream records them
    CK(cudaEventCreate(&e0)); CK(cudaEventCreate(&e1));
    float ms = 0;
    if (strcmp(mode, "ll")) {
        const int dyn = flavour == FLAVOUR_TMA ? (int)(a.h.tma_chunk * a.h.tma_stages) : 0;
        for (int i = 0; i < 3; ++i) CK(launch_group(flavour, a, ctas, kLdstThreads, dyn, s));
        CK(cudaStreamSynchronize(s));
        CK(cudaEventRecord(e0, s));
        for (int i = 0; i < iters; ++i) CK(launch_group(flavour, a, ctas, kLdstThreads, dyn, s));
        CK(cudaEventRecord(e1, s));
        CK(cudaStreamSynchronize(s));
        CK(cudaEventElapsedTime(&ms, e0, e1));
        printf("%s: %zu bytes, %d CTAs, %s: %.2f us per launch, %.1f GB/s over NVLink\n", mode, bytes, ctas,
               flavour == FLAVOUR_TMA ? "p2p_group_kernel_tma" : "p2p_group_kernel_ldst", ms * 1e3 / iters, bytes * (double)iters / (ms * 1e-3) / 1e9);
        std::vector<uint8_t> h(4096);
        CK(cudaMemcpy(h.data(), b1 + bytes - 4096, 4096, cudaMemcpyDefault));
        bool ok = true;
        for (uint8_t v : h) ok = ok && v == 0x5a;
        printf("payload %s\n", ok ? "arrived" : "WRONG");
        return ok ? 0 : 1;
    }
    // eager: message k = one send launch on GPU 0, then one receive launch on GPU 1 (the data is already there,
    // so neither kernel waits for the other: safe to serialise under ncu)
    for (int k = 1; k <= iters && k <= kLLSlots; ++k) {
        GroupArgs snd = a, rcv = a;
        OpDesc& so = snd.ops[0];
        so.kind = OP_SEND_LL; so.local = (uint64_t)b0; so.aux = (uint64_t)(ring1 + (size_t)(k % kLLSlots) * (1 << 19)); so.f_wait = (uint64_t)ack0;
        so.seq = (uint64_t)k; so.nblocks = 1; so.nbytes = bytes;
        OpDesc& ro = rcv.ops[0];
        ro.kind = OP_RECV_LL; ro.local = (uint64_t)b1; ro.aux = so.aux; ro.f_signal = (uint64_t)ack0; ro.arrive = (uint64_t)arrive1;
        ro.seq = (uint64_t)k; ro.nblocks = 1; ro.nbytes = bytes;
        CK(cudaSetDevice(0));
        CK(launch_group(FLAVOUR_LDST, snd, 1, 128, 0, s0));
        CK(cudaStreamSynchronize(s0));
        CK(cudaSetDevice(1));
        CK(cudaEventRecord(e0, s1));
        CK(launch_group(FLAVOUR_LDST, rcv, 1, 128, 0, s1));
        CK(cudaEventRecord(e1, s1));
        CK(cudaStreamSynchronize(s1));
        CK(cudaEventElapsedTime(&ms, e0, e1));
    }
    std::vector<uint8_t> h(bytes);
    CK(cudaMemcpy(h.data(), b1, bytes, cudaMemcpyDefault));
    bool ok = true;
    for (uint8_t v : h) ok = ok && v == 0x5a;
    printf("ll: %zu bytes through the eager ring in GPU 1's memory, receive kernel %.2f us; payload %s\n", bytes, ms * 1e3, ok ? "arrived" : "WRONG");
    return ok ? 0 : 1;
}
