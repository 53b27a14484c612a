// p2p_matrix — the matrix benchmark of AmadeusChan/test-nccl-p2p, with the B200 transfer
// engine where the reference calls NCCL.  Same launch (`mpirun -n N p2p_matrix`, reference
// README.md:5), same report; with no arguments it measures what the reference measures the
// way the reference measures it.  The driver is in matrix_driver.hpp; this file is only the
// engine's backend, i.e. the lines of p2p_matrix.cc that named ncclXxx.
#include "matrix_driver.hpp"
#include "p2p.h"
#include "p2p_mpi.h"

struct EngineBackend {
    const char* name() const { return "engine"; }

    // replaces ncclGetUniqueId + MPI_Bcast + ncclCommInitRank (p2p_matrix.cc:115-120)
    void init(int, int, int device) { P2PCHECK(p2p_init_mpi(MPI_COMM_WORLD, device)); }

    // the reference's cudaMalloc'd buffers stay as they are (p2p_matrix.cc:127-128); the one
    // added step is making them windows: peers store into the receive buffer (push) and, for
    // large messages, load from the send buffer (pull)
    void buffers(void* send, size_t sbytes, void* recv, size_t rbytes) {
        int win = -1;
        // P2P_NO_WINDOWS: exactly the reference's lines, nothing registered (everything above the eager limit then
        // goes through the staging ring); P2P_NO_SEND_WINDOW: only the receive buffer is a window
        if (getenv("P2P_NO_WINDOWS")) return;
        if (!getenv("P2P_NO_RECV_WINDOW")) P2PCHECK(p2p_register_mpi(MPI_COMM_WORLD, recv, rbytes, &win));
        if (!getenv("P2P_NO_SEND_WINDOW")) P2PCHECK(p2p_register_mpi(MPI_COMM_WORLD, send, sbytes, &win));
    }

    // ncclSend / ncclRecv with ncclInt8 (p2p_matrix.cc:157-161, 165-169)
    void send(const void* buf, size_t bytes, int peer, cudaStream_t s) { P2PCHECK(p2p_send(buf, bytes, P2P_INT8, peer, s)); }
    void recv(void* buf, size_t bytes, int peer, cudaStream_t s) { P2PCHECK(p2p_recv(buf, bytes, P2P_INT8, peer, s)); }
    void group_start() { P2PCHECK(p2p_group_start()); }  // p2p_matrix.cc:212
    void group_end() { P2PCHECK(p2p_group_end()); }      // p2p_matrix.cc:227
    void alltoall(const void* send, void* recv, size_t bytes, cudaStream_t s) { P2PCHECK(p2p_alltoall(send, recv, bytes, P2P_INT8, s)); }

    // one-to-many: a symmetric window (driver VMM memory, bound to the NVSwitch multicast object) and p2p_bcast;
    // P2P_BCAST_SENDS=1 does it the reference's way instead, as N-1 grouped sends from the root, for comparison
    void* symm_alloc(size_t bytes) {
        void* p = nullptr;
        int win = -1;
        P2PCHECK(p2p_symm_alloc_mpi(MPI_COMM_WORLD, bytes, &p, &win));
        return p;
    }
    void bcast(const void* send, void* recv, size_t bytes, int root, cudaStream_t s) {
        if (!getenv("P2P_BCAST_SENDS")) {
            P2PCHECK(p2p_bcast(send, recv, bytes, P2P_INT8, root, s));
            return;
        }
        const int me = p2p_rank(), n = p2p_nranks();
        P2PCHECK(p2p_group_start());
        if (me == root) {
            for (int p = 0; p < n; ++p)
                if (p != root) P2PCHECK(p2p_send(send, bytes, P2P_INT8, p, s));
                else CUDACHECK(cudaMemcpyAsync(recv, send, bytes, cudaMemcpyDeviceToDevice, s));
        } else {
            P2PCHECK(p2p_recv(recv, bytes, P2P_INT8, root, s));
        }
        P2PCHECK(p2p_group_end());
    }

    void fill(void* buf, size_t bytes, uint64_t seed, int rank, uint64_t iter, cudaStream_t s) {
        P2PCHECK(p2p_pattern_fill(buf, bytes, seed, rank, iter, s));
    }
    uint64_t check(const void* buf, size_t bytes, uint64_t seed, int rank, uint64_t iter, cudaStream_t s,
                   uint64_t* digest) {
        uint64_t bad = 0, first = 0;
        P2PCHECK(p2p_pattern_check(buf, bytes, seed, rank, iter, s, &bad, &first, digest));
        int aerr = 0;
        P2PCHECK(p2p_get_async_error(&aerr));
        if (aerr != P2P_SUCCESS) {
            printf("Failed, P2P async error '%s' (%s)\n", p2p_get_error_string(aerr), p2p_get_last_error());
            exit(EXIT_FAILURE);
        }
        return bad;
    }

    void finalize() {
        if (getenv("P2P_PRINT_STATS")) {  // experiments: how many kernels were launched, how many payloads went by DMA
            long long launches = 0, ce = 0;
            p2p_get_option("stat_launches", &launches);
            p2p_get_option("stat_ce_copies", &ce);
            printf("[rank %d] kernels launched %lld, payloads handed to the copy engine %lld\n", p2p_rank(), launches, ce);
        }
        int aerr = 0;
        if (p2p_get_async_error(&aerr) == P2P_SUCCESS && aerr != P2P_SUCCESS) {
            printf("Failed, P2P async error '%s' (%s)\n", p2p_get_error_string(aerr), p2p_get_last_error());
            exit(EXIT_FAILURE);
        }
        P2PCHECK(p2p_finalize_mpi(MPI_COMM_WORLD));  // ncclCommDestroy, p2p_matrix.cc:270
    }
};

int main(int argc, char** argv) {
    EngineBackend be;
    return bench::run_main(be, argc, argv);
}
