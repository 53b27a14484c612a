// TEST INFRASTRUCTURE / BASELINE — not part of the product.
//
// The CPU statement of the path: the same message passed between the same two ranks through
// HOST memory with MPI_Send/MPI_Recv, in the reference's loop shape (p2p_matrix.cc:141-186
// uni-directional, :196-267 bi-directional: barrier, wall clock on rank 0, `iters` messages,
// barrier, Gbps = bytes*8/t/1e9).  No GPU is touched, so this also runs on a CPU-only box.
// It is the "MPI host-memory" baseline BASELINE.json asks for next to the NCCL one; under
// the in-repo MPI shim the transport is a shared-memory ring (two memcpy per byte), and the
// program says which MPI carried it.
//
//   mpirun -n N mpi_host_p2p [--size BYTES] [--iters N] [--warmup N] [--mode uni|bidir|ref]
//                            [--first-pair] [--verify] [--json FILE]
#include <chrono>
#include <cstdint>
#include <cstdio>
#include <cstdlib>
#include <cstring>
#include <string>
#include <vector>
#include <unistd.h>

#include "mpi.h"
#include "pattern.h"

#define MPICHECK(cmd) do {                          \
    int e = cmd;                                      \
    if( e != MPI_SUCCESS ) {                          \
        printf("Failed: MPI error %s:%d '%d'\n",        \
                __FILE__,__LINE__, e);   \
        exit(EXIT_FAILURE);                             \
    }                                                 \
} while(0)

static const char* transport_name() {
#ifdef MINIMPI
    return MINIMPI_transport_name();
#else
    return "system MPI";
#endif
}

int main(int argc, char** argv) {
    int rank, size, provided;
    MPICHECK(MPI_Init_thread(&argc, &argv, MPI_THREAD_MULTIPLE, &provided));
    MPICHECK(MPI_Comm_rank(MPI_COMM_WORLD, &rank));
    MPICHECK(MPI_Comm_size(MPI_COMM_WORLD, &size));
    size_t bytes = 32u * 1024u * 1024u;  // p2p_matrix.cc:124
    int iters = 128, warmup = 0;          // p2p_matrix.cc:132 (and its absence of warm-up)
    std::string mode = "ref", json;
    bool first_pair = false, verify = false;
    for (int i = 1; i < argc; ++i) {
        std::string a = argv[i];
        if (a == "--size" && i + 1 < argc) bytes = (size_t)strtoull(argv[++i], nullptr, 0);
        else if (a == "--iters" && i + 1 < argc) iters = atoi(argv[++i]);
        else if (a == "--warmup" && i + 1 < argc) warmup = atoi(argv[++i]);
        else if (a == "--mode" && i + 1 < argc) mode = argv[++i];
        else if (a == "--json" && i + 1 < argc) json = argv[++i];
        else if (a == "--first-pair") first_pair = true;
        else if (a == "--verify") verify = true;
        else { if (rank == 0) printf("bad argument %s\n", a.c_str()); return 2; }
    }
    if (bytes > 0x7fffffffu) { if (rank == 0) printf("--size must fit an MPI int count\n"); return 2; }
    std::vector<uint8_t> send_buff(bytes ? bytes : 1, 0), recv_buff(bytes ? bytes : 1, 0);  // zeros, :129-130
    FILE* jf = nullptr;
    if (rank == 0 && !json.empty()) jf = fopen(json.c_str(), "w");
    long ncores = sysconf(_SC_NPROCESSORS_ONLN);
    uint64_t verify_bad = 0;

    for (int pass = 0; pass < 2; ++pass) {
        const bool bidir = pass == 1;
        if ((mode == "uni" && bidir) || (mode == "bidir" && !bidir)) continue;
        if (rank == 0) {
            printf("%sEvaluating the %s-Directional MPI host-memory P2P Bandwidth (Gbps) [%s, %ld host cores]\n",
                   bidir ? "\n" : "", bidir ? "Bi" : "Uni", transport_name(), ncores);
            printf("   D\\D");
            for (int i = 0; i < size; ++i) printf("%6d ", i);
            printf("\n");
        }
        for (int src = 0; src < size; ++src) {
            if (rank == 0) printf("%6d ", src);
            for (int dst = 0; dst < size; ++dst) {
                MPI_Barrier(MPI_COMM_WORLD);
                if (src == dst || (first_pair && !(src == 0 && dst == 1))) {
                    if (rank == 0) printf("%6.02f ", 0.);
                    continue;
                }
                auto one = [&]() {
                    if (!bidir) {
                        if (rank == src) MPICHECK(MPI_Send(send_buff.data(), (int)bytes, MPI_BYTE, dst, 0, MPI_COMM_WORLD));
                        else if (rank == dst) MPICHECK(MPI_Recv(recv_buff.data(), (int)bytes, MPI_BYTE, src, 0, MPI_COMM_WORLD, MPI_STATUS_IGNORE));
                    } else if (rank == src || rank == dst) {
                        const int peer = rank == src ? dst : src;
                        MPICHECK(MPI_Sendrecv(send_buff.data(), (int)bytes, MPI_BYTE, peer, 0, recv_buff.data(), (int)bytes,
                                              MPI_BYTE, peer, 0, MPI_COMM_WORLD, MPI_STATUS_IGNORE));
                    }
                };
                if (verify) {
                    const uint64_t iter = (uint64_t)(pass * size * size + src * size + dst);
                    oracle_pattern_fill(send_buff.data(), bytes, 0x5EED0000ull, rank, iter);
                    memset(recv_buff.data(), 0xA5, bytes);
                    MPI_Barrier(MPI_COMM_WORLD);
                    one();
                    uint64_t first;
                    if (rank == dst) verify_bad += oracle_pattern_mismatches(recv_buff.data(), bytes, 0x5EED0000ull, src, iter, &first);
                    if (bidir && rank == src) verify_bad += oracle_pattern_mismatches(recv_buff.data(), bytes, 0x5EED0000ull, dst, iter, &first);
                }
                for (int i = 0; i < warmup; ++i) one();
                MPI_Barrier(MPI_COMM_WORLD);
                auto start = std::chrono::system_clock::now();  // :153
                for (int i = 0; i < iters; ++i) one();
                MPI_Barrier(MPI_COMM_WORLD);                     // :173
                std::chrono::duration<double> el = std::chrono::system_clock::now() - start;
                const double t = el.count() / iters;
                const double gbps = oracle_gbps(bytes, t, bidir);
                if (rank == 0) {
                    printf("%6.02f ", gbps);
                    fflush(stdout);
                    if (jf) {
                        fprintf(jf,
                                "{\"backend\": \"mpi-host\", \"transport\": \"%s\", \"cores\": %d, \"host_cores\": %ld, "
                                "\"mode\": \"%s\", \"timing\": \"ref\", \"ranks\": %d, \"src\": %d, \"dst\": %d, \"bytes\": %zu, "
                                "\"seconds\": %.9g, \"us\": %.4f, \"GBps\": %.4f, \"iters\": %d, \"warmup\": %d}\n",
                                transport_name(), 2, ncores, bidir ? "bidir" : "uni", size, src, dst, bytes, t, t * 1e6,
                                (bidir ? 2.0 : 1.0) * bytes / t / 1e9, iters, warmup);
                        fflush(jf);
                    }
                }
            }
            if (rank == 0) printf("\n");
        }
    }
    unsigned long long b = verify_bad, total = 0;
    MPICHECK(MPI_Allreduce(&b, &total, 1, MPI_UINT64_T, MPI_SUM, MPI_COMM_WORLD));
    if (verify && rank == 0) printf(total ? "verify: FAILED (%llu bytes differ)\n" : "verify: OK\n", total);
    if (jf) fclose(jf);
    MPICHECK(MPI_Barrier(MPI_COMM_WORLD));
    return total ? 1 : 0;
}
