// Experiment (not part of the product): why does the engine's copy-engine chain move 256 MiB both ways at
// 1040 GB/s when two bare DMAs between the same two processes reach 1521 (scripts/exp_ce_windows.py)?
// Two ranks (MPI), registered windows, each rank copies `bytes` into the other's window, both at once:
//   A  bare cudaMemcpyAsync on each side                                              (the ceiling)
//   B  A, while a one-thread kernel polls a local flag on each GPU (ld.acquire.sys + nanosleep)
//   C  A, while a 65-CTA kernel of which one thread polls runs on each GPU             (the engine's receive kernel)
//   D  the chain with stream memory operations on BOTH sides, no kernel anywhere:
//        write "ready" into the peer, wait for the peer's "ready", DMA, write "done" into the peer, wait for its "done"
//   E  D, but the two waits are done by a polling kernel on another stream             (the engine's shape)
// build: nvcc -gencode arch=compute_100a,code=sm_100a -std=c++17 -O3 -Iinclude -Ithird_party/minimpi -o build/exp_ce_chain \
//             scripts/exp_ce_chain.cu build/libb200p2p.a -Lthird_party/minimpi -lmpi -lpthread -L/usr/local/cuda/lib64/stubs -lcuda
// run:   build/mpirun -n 2 build/exp_ce_chain [bytes]
#include <cstdint>
#include <cstdio>
#include <cstdlib>
#include <cuda.h>
#include <cuda_runtime.h>
#include "mpi.h"
#include "p2p.h"
#include "p2p_mpi.h"

#define CK(x) do { cudaError_t e_ = (x); if (e_ != cudaSuccess) { printf("CUDA error %s at %s:%d\n", cudaGetErrorString(e_), __FILE__, __LINE__); exit(1); } } while (0)
#define CU(x) do { CUresult r_ = (x); if (r_ != CUDA_SUCCESS) { printf("driver error %d at %s:%d\n", (int)r_, __FILE__, __LINE__); exit(1); } } while (0)

__global__ void poll_kernel(const uint64_t* flag, uint64_t want, uint64_t timeout_ns) {
    if (blockIdx.x != 0 || threadIdx.x != 0) return;
    uint64_t t0, t;
    asm volatile("mov.u64 %0, %%globaltimer;" : "=l"(t0));
    for (;;) {
        uint64_t v;
        asm volatile("ld.acquire.sys.global.u64 %0, [%1];" : "=l"(v) : "l"(flag) : "memory");
        if (v >= want) return;
        __nanosleep(40);
        asm volatile("mov.u64 %0, %%globaltimer;" : "=l"(t));
        if (t - t0 > timeout_ns) return;
    }
}

int main(int argc, char** argv) {
    int provided, rank, size;
    MPI_Init_thread(&argc, &argv, MPI_THREAD_MULTIPLE, &provided);
    MPI_Comm_rank(MPI_COMM_WORLD, &rank);
    MPI_Comm_size(MPI_COMM_WORLD, &size);
    if (size != 2) { if (!rank) printf("run with 2 ranks\n"); return 1; }
    const size_t bytes = argc > 1 ? strtoull(argv[1], nullptr, 0) : (256ull << 20);
    CK(cudaSetDevice(rank));
    P2PCHECK(p2p_init_mpi(MPI_COMM_WORLD, rank));
    const int peer = 1 - rank;
    uint8_t *src, *dst;
    uint64_t* flags;  // [0] ready, [1] done, [2] release of the polling kernels
    const size_t slack = getenv("EXP_DST_SLACK") ? 256 : 0;   // the matrix driver's receive buffer is bytes + 256
    CK(cudaMalloc(&src, bytes)); CK(cudaMalloc(&dst, bytes + slack)); CK(cudaMalloc(&flags, 4096));
    CK(cudaMemset(src, rank + 1, bytes)); CK(cudaMemset(flags, 0, 4096));
    CK(cudaDeviceSynchronize());
    int wdst = -1, wflags = -1, wsrc = -1;
    P2PCHECK(p2p_register_mpi(MPI_COMM_WORLD, dst, bytes + slack, &wdst));
    if (getenv("EXP_REGISTER_SRC")) {   // is a DMA slower when its SOURCE is a cudaIpc-exported buffer a peer has mapped?
        P2PCHECK(p2p_register_mpi(MPI_COMM_WORLD, src, bytes, &wsrc));
        if (!rank) printf("(source buffer registered as a window too)\n");
    }
    P2PCHECK(p2p_register_mpi(MPI_COMM_WORLD, flags, 4096, &wflags));
    void *pdst = nullptr, *pflags_v = nullptr;
    P2PCHECK(p2p_window_peer_ptr(wdst, peer, &pdst));
    P2PCHECK(p2p_window_peer_ptr(wflags, peer, &pflags_v));
    uint64_t* pflags = (uint64_t*)pflags_v;
    cudaStream_t s, s2;
    CK(cudaStreamCreateWithFlags(&s, cudaStreamNonBlocking));
    CK(cudaStreamCreateWithFlags(&s2, cudaStreamNonBlocking));
    cudaEvent_t e0, e1;
    CK(cudaEventCreate(&e0)); CK(cudaEventCreate(&e1));
    uint64_t* pin;
    CK(cudaMallocHost(&pin, 64));
    const int iters = 20;
    uint64_t seq = 0, release = 0;

    auto report = [&](const char* what, float ms) {
        double t = ms * 1e-3 / iters, tmax = 0;
        MPI_Allreduce(&t, &tmax, 1, MPI_DOUBLE, MPI_MAX, MPI_COMM_WORLD);
        if (!rank) printf("%-78s %8.1f us per step  %8.1f GB/s both ways\n", what, tmax * 1e6, 2.0 * bytes / tmax / 1e9);
        fflush(stdout);
    };
    auto run = [&](const char* what, auto step) {
        for (int i = 0; i < 3; ++i) step();
        CK(cudaStreamSynchronize(s)); CK(cudaStreamSynchronize(s2));
        MPI_Barrier(MPI_COMM_WORLD);
        CK(cudaEventRecord(e0, s));
        for (int i = 0; i < iters; ++i) step();
        CK(cudaEventRecord(e1, s));
        CK(cudaStreamSynchronize(s)); CK(cudaStreamSynchronize(s2));
        float ms;
        CK(cudaEventElapsedTime(&ms, e0, e1));
        report(what, ms);
        MPI_Barrier(MPI_COMM_WORLD);
    };
    // polling kernels are released by a host write once the step's DMA has finished
    auto with_poller = [&](int grid, auto body) {
        ++release;
        poll_kernel<<<grid, grid > 1 ? 512 : 32, 0, s2>>>(flags + 2, release, 2000000000ull);
        body();
        CK(cudaStreamSynchronize(s));
        *pin = release;
        CK(cudaMemcpyAsync(flags + 2, pin, 8, cudaMemcpyDefault, s));
        CK(cudaStreamSynchronize(s)); CK(cudaStreamSynchronize(s2));
    };

    run("A  bare DMA each way", [&] { CK(cudaMemcpyAsync(pdst, src, bytes, cudaMemcpyDefault, s)); });
    run("B  DMA each way + a one-thread polling kernel on each GPU", [&] { with_poller(1, [&] { CK(cudaMemcpyAsync(pdst, src, bytes, cudaMemcpyDefault, s)); }); });
    run("C  DMA each way + a 65-CTA kernel (one thread polls) on each GPU", [&] { with_poller(65, [&] { CK(cudaMemcpyAsync(pdst, src, bytes, cudaMemcpyDefault, s)); }); });
    run("A' bare DMA each way, host sync per step (what B and C pay besides the kernel)", [&] { CK(cudaMemcpyAsync(pdst, src, bytes, cudaMemcpyDefault, s)); CK(cudaStreamSynchronize(s)); });
    run("D  memop chain both sides: ready -> wait ready -> DMA -> done -> wait done", [&] {
        ++seq;
        CU(cuStreamWriteValue64((CUstream)s, (CUdeviceptr)(pflags + 0), seq, 0));
        CU(cuStreamWaitValue64((CUstream)s, (CUdeviceptr)(flags + 0), seq, CU_STREAM_WAIT_VALUE_GEQ));
        CK(cudaMemcpyAsync(pdst, src, bytes, cudaMemcpyDefault, s));
        CU(cuStreamWriteValue64((CUstream)s, (CUdeviceptr)(pflags + 1), seq, 0));
        CU(cuStreamWaitValue64((CUstream)s, (CUdeviceptr)(flags + 1), seq, CU_STREAM_WAIT_VALUE_GEQ));
    });
    run("E  chain on one stream, the wait for 'done' by a polling kernel on another stream", [&] {
        ++seq;
        poll_kernel<<<65, 512, 0, s2>>>(flags + 1, seq, 2000000000ull);
        CU(cuStreamWriteValue64((CUstream)s, (CUdeviceptr)(pflags + 0), seq, 0));
        CU(cuStreamWaitValue64((CUstream)s, (CUdeviceptr)(flags + 0), seq, CU_STREAM_WAIT_VALUE_GEQ));
        CK(cudaMemcpyAsync(pdst, src, bytes, cudaMemcpyDefault, s));
        CU(cuStreamWriteValue64((CUstream)s, (CUdeviceptr)(pflags + 1), seq, 0));
    });
    run("H  D with a local stream write between the wait and the DMA (the engine zeroes the consumed post slot there)", [&] {
        ++seq;
        CU(cuStreamWriteValue64((CUstream)s, (CUdeviceptr)(pflags + 0), seq, 0));
        CU(cuStreamWaitValue64((CUstream)s, (CUdeviceptr)(flags + 0), seq, CU_STREAM_WAIT_VALUE_GEQ));
        CU(cuStreamWriteValue64((CUstream)s, (CUdeviceptr)(flags + 8), seq, 0));
        CK(cudaMemcpyAsync(pdst, src, bytes, cudaMemcpyDefault, s));
        CU(cuStreamWriteValue64((CUstream)s, (CUdeviceptr)(pflags + 1), seq, 0));
        CU(cuStreamWaitValue64((CUstream)s, (CUdeviceptr)(flags + 1), seq, CU_STREAM_WAIT_VALUE_GEQ));
    });
    // F: the engine's shape - the chain on a SIDE stream, forked from and joined to the main stream by events every step
    // (the host enqueues many steps ahead, so each event wait sits in its queue for a whole step before it is satisfied)
    {
        cudaEvent_t fork, join;
        CK(cudaEventCreateWithFlags(&fork, cudaEventDisableTiming));
        CK(cudaEventCreateWithFlags(&join, cudaEventDisableTiming));
        run("F  D's chain on a side stream, forked/joined by events each step", [&] {
            ++seq;
            CK(cudaEventRecord(fork, s));
            CK(cudaStreamWaitEvent(s2, fork, 0));
            CU(cuStreamWriteValue64((CUstream)s2, (CUdeviceptr)(pflags + 0), seq, 0));
            CU(cuStreamWaitValue64((CUstream)s2, (CUdeviceptr)(flags + 0), seq, CU_STREAM_WAIT_VALUE_GEQ));
            CK(cudaMemcpyAsync(pdst, src, bytes, cudaMemcpyDefault, s2));
            CU(cuStreamWriteValue64((CUstream)s2, (CUdeviceptr)(pflags + 1), seq, 0));
            CU(cuStreamWaitValue64((CUstream)s2, (CUdeviceptr)(flags + 1), seq, CU_STREAM_WAIT_VALUE_GEQ));
            CK(cudaEventRecord(join, s2));
            CK(cudaStreamWaitEvent(s, join, 0));
        });
        run("F' as F, plus a 65-CTA polling kernel for 'done' on the main stream (engine: receive kernel beside the chain)", [&] {
            ++seq;
            CK(cudaEventRecord(fork, s));
            CK(cudaStreamWaitEvent(s2, fork, 0));
            poll_kernel<<<65, 512, 0, s>>>(flags + 1, seq, 2000000000ull);
            CU(cuStreamWriteValue64((CUstream)s2, (CUdeviceptr)(pflags + 0), seq, 0));
            CU(cuStreamWaitValue64((CUstream)s2, (CUdeviceptr)(flags + 0), seq, CU_STREAM_WAIT_VALUE_GEQ));
            CK(cudaMemcpyAsync(pdst, src, bytes, cudaMemcpyDefault, s2));
            CU(cuStreamWriteValue64((CUstream)s2, (CUdeviceptr)(pflags + 1), seq, 0));
            CK(cudaEventRecord(join, s2));
            CK(cudaStreamWaitEvent(s, join, 0));
        });
    }
    // G: one way only.  Rank 1's stream memory wait for 'done' lasts a whole DMA: does a long wait react late?
    {
        const uint64_t base = seq;
        for (int i = 0; i < 3 + iters; ++i) {
            if (i == 3) { CK(cudaStreamSynchronize(s)); MPI_Barrier(MPI_COMM_WORLD); CK(cudaEventRecord(e0, s)); }
            ++seq;
            if (rank == 0) {
                CU(cuStreamWaitValue64((CUstream)s, (CUdeviceptr)(flags + 0), seq, CU_STREAM_WAIT_VALUE_GEQ));
                CK(cudaMemcpyAsync(pdst, src, bytes, cudaMemcpyDefault, s));
                CU(cuStreamWriteValue64((CUstream)s, (CUdeviceptr)(pflags + 1), seq, 0));
            } else {
                CU(cuStreamWriteValue64((CUstream)s, (CUdeviceptr)(pflags + 0), seq, 0));
                CU(cuStreamWaitValue64((CUstream)s, (CUdeviceptr)(flags + 1), seq, CU_STREAM_WAIT_VALUE_GEQ));
            }
        }
        (void)base;
        CK(cudaEventRecord(e1, s));
        CK(cudaStreamSynchronize(s));
        float ms;
        CK(cudaEventElapsedTime(&ms, e0, e1));
        double t = ms * 1e-3 / iters, tmax = 0;
        MPI_Allreduce(&t, &tmax, 1, MPI_DOUBLE, MPI_MAX, MPI_COMM_WORLD);
        if (!rank) printf("%-78s %8.1f us per step  %8.1f GB/s one way\n", "G  one way: receiver = memop ready + memop wait for done (a DMA long)", tmax * 1e6, bytes / tmax / 1e9);
        MPI_Barrier(MPI_COMM_WORLD);
    }
    // payload check
    uint8_t probe[2];
    CK(cudaMemcpy(probe, dst, 1, cudaMemcpyDefault)); CK(cudaMemcpy(probe + 1, dst + bytes - 1, 1, cudaMemcpyDefault));
    if (probe[0] != peer + 1 || probe[1] != peer + 1) printf("rank %d: payload WRONG\n", rank);
    CK(cudaDeviceSynchronize());
    P2PCHECK(p2p_finalize_mpi(MPI_COMM_WORLD));
    MPI_Barrier(MPI_COMM_WORLD);
    return 0;
}
