// TEST INFRASTRUCTURE / BASELINE — not part of the product.
//
// The reference's own transfer path (libnccl's ncclSend/ncclRecv, called exactly as
// p2p_matrix.cc:157-169 and :212-227 call them) behind the same matrix driver the engine is
// measured with, so that sizes other than 32 MiB, device timing, all-pairs and the sweep can
// be run against NCCL with identical loops.  With no arguments it repeats the reference's
// measurement.  Payload fill and check are done ON THE CPU with oracle/pattern.c — nothing
// of the engine touches this data — so the digests it reports are the golden ones.
#include "matrix_driver.hpp"
#include "nccl.h"
#include "pattern.h"

#define NCCLCHECK(cmd) do {                         \
    ncclResult_t r = cmd;                             \
    if (r!= ncclSuccess) {                            \
        printf("Failed, NCCL error %s:%d '%s'\n",             \
                __FILE__,__LINE__,ncclGetErrorString(r));   \
        exit(EXIT_FAILURE);                             \
    }                                                 \
} while(0)

struct NcclBackend {
    ncclComm_t comm = nullptr;
    int rank = 0;
    std::vector<uint8_t> host;

    const char* name() const { return "nccl"; }

    void init(int rank_, int world_size, int) {
        rank = rank_;
        ncclUniqueId id;  // p2p_matrix.cc:111-120
        if (rank == 0) NCCLCHECK(ncclGetUniqueId(&id));
        MPICHECK(MPI_Bcast((void*)&id, sizeof(id), MPI_BYTE, 0, MPI_COMM_WORLD));
        NCCLCHECK(ncclCommInitRank(&comm, world_size, id, rank));
    }
    void buffers(void*, size_t, void*, size_t) {}  // plain cudaMalloc, unregistered: NCCL defaults, as in the reference

    void send(const void* buf, size_t bytes, int peer, cudaStream_t s) {
        NCCLCHECK(ncclSend(buf, bytes, ncclInt8, peer, comm, s));  // p2p_matrix.cc:157-161
    }
    void recv(void* buf, size_t bytes, int peer, cudaStream_t s) {
        NCCLCHECK(ncclRecv(buf, bytes, ncclInt8, peer, comm, s));  // p2p_matrix.cc:165-169
    }
    void group_start() { NCCLCHECK(ncclGroupStart()); }  // p2p_matrix.cc:212
    void group_end() { NCCLCHECK(ncclGroupEnd()); }      // p2p_matrix.cc:227
    int nranks = 0;
    // the grouped send/recv loop NCCL's documentation prescribes for all-to-all (nccl.h:438-440)
    void alltoall(const void* send, void* recv, size_t bytes, cudaStream_t s) {
        if (!nranks) NCCLCHECK(ncclCommCount(comm, &nranks));
        NCCLCHECK(ncclGroupStart());
        for (int p = 0; p < nranks; ++p) {
            NCCLCHECK(ncclRecv((uint8_t*)recv + (size_t)p * bytes, bytes, ncclInt8, p, comm, s));
            NCCLCHECK(ncclSend((const uint8_t*)send + (size_t)p * bytes, bytes, ncclInt8, p, comm, s));
        }
        NCCLCHECK(ncclGroupEnd());
    }

    // one-to-many the only way the reference's call shape allows it: N-1 sends from the root (p2p_matrix.cc:141-172),
    // grouped so that they overlap; NCCL_HARNESS_BCAST=collective times ncclBroadcast instead, for context
    void* symm_alloc(size_t bytes) {
        void* p = nullptr;
        CUDACHECK(cudaMalloc(&p, bytes));
        return p;
    }
    void bcast(const void* send, void* recv, size_t bytes, int root, cudaStream_t s) {
        if (!nranks) NCCLCHECK(ncclCommCount(comm, &nranks));
        const char* how = getenv("NCCL_HARNESS_BCAST");
        if (how && !strcmp(how, "collective")) {
            NCCLCHECK(ncclBroadcast(send, recv, bytes, ncclInt8, root, comm, s));
            return;
        }
        NCCLCHECK(ncclGroupStart());
        if (rank == root) {
            for (int p = 0; p < nranks; ++p)
                if (p != root) NCCLCHECK(ncclSend(send, bytes, ncclInt8, p, comm, s));
                else CUDACHECK(cudaMemcpyAsync(recv, send, bytes, cudaMemcpyDeviceToDevice, s));
        } else {
            NCCLCHECK(ncclRecv(recv, bytes, ncclInt8, root, comm, s));
        }
        NCCLCHECK(ncclGroupEnd());
    }

    void fill(void* buf, size_t bytes, uint64_t seed, int r, uint64_t iter, cudaStream_t s) {
        host.resize(bytes);
        oracle_pattern_fill(host.data(), bytes, seed, r, iter);
        CUDACHECK(cudaMemcpyAsync(buf, host.data(), bytes, cudaMemcpyHostToDevice, s));
        CUDACHECK(cudaStreamSynchronize(s));
    }
    uint64_t check(const void* buf, size_t bytes, uint64_t seed, int r, uint64_t iter, cudaStream_t s,
                   uint64_t* digest) {
        host.resize(bytes);
        CUDACHECK(cudaMemcpyAsync(host.data(), buf, bytes, cudaMemcpyDeviceToHost, s));
        CUDACHECK(cudaStreamSynchronize(s));
        uint64_t first = 0;
        const uint64_t bad = oracle_pattern_mismatches(host.data(), bytes, seed, r, iter, &first);
        if (digest) *digest = oracle_digest(host.data(), bytes);
        return bad;
    }
    void finalize() { ncclCommDestroy(comm); }  // p2p_matrix.cc:270
};

int main(int argc, char** argv) {
    NcclBackend be;
    return bench::run_main(be, argc, argv);
}
