// matrix_driver.hpp — the matrix benchmark, written once over a "backend" (who moves the
// bytes).  The engine's backend lives in p2p_matrix.cc; the NCCL and MPI-host baselines in
// oracle/ plug into the same driver so all three are measured by the same loops.
//
// What is mirrored from the reference (p2p_matrix.cc), and where:
//   placement / device choice        :44-100   placement_device()
//   src x dst sweep, barriers, order :141-186, :196-267   run_ref_tables()
//   per-iteration host sync          :162, :170, :229-230, :250-251
//   wall-clock timing on rank 0      :153, :173-176
//   Gbps formula                     :177, :258
//   report text                      :133-140, :178-185, :188-195
// What is added (the reference has no flags, sizes, warm-up, verification or device timing):
//   --size/--iters/--warmup/--mode/--timing/--verify/--json and the GB/s, us, %-of-link tables.
#pragma once

#include <algorithm>
#include <cassert>
#include <chrono>
#include <cmath>
#include <cstdint>
#include <cstdio>
#include <cstdlib>
#include <cstring>
#include <set>
#include <string>
#include <vector>
#include <unistd.h>

#include "cuda_runtime.h"
#include "mpi.h"

#define MPICHECK(cmd) do {                          \
    int e = cmd;                                      \
    if( e != MPI_SUCCESS ) {                          \
        printf("Failed: MPI error %s:%d '%d'\n",        \
                __FILE__,__LINE__, e);   \
        exit(EXIT_FAILURE);                             \
    }                                                 \
} while(0)

#define CUDACHECK(cmd) do {                         \
    cudaError_t e = cmd;                              \
    if( e != cudaSuccess ) {                          \
        printf("Failed: Cuda error %s:%d '%s'\n",             \
                __FILE__,__LINE__,cudaGetErrorString(e));   \
        exit(EXIT_FAILURE);                             \
    }                                                 \
} while(0)

namespace bench {

constexpr double kLinkGBps = 900.0;  // NVLink 5, per direction per GPU (BASELINE.json)

// ------------------------------------------------------------------ placement (pure part testable)

inline uint64_t host_hash(const char* s) {
    uint64_t r = 5381;  // DJB2a, stops at the first '.', as getHostName+getHostHash do together
    for (int c = 0; s[c] != '\0' && s[c] != '.'; c++) r = ((r << 5) + r) ^ (uint64_t)s[c];
    return r;
}

// Returns the local device index for `rank`, or -1 (uneven hosts) / -2 (ranks of one host
// not contiguous).  Same rules as check_process_placement_policy, p2p_matrix.cc:78-99.
inline int placement_from_hashes(const std::vector<uint64_t>& h, int rank) {
    const int size = (int)h.size();
    std::set<uint64_t> hosts(h.begin(), h.end());
    const int num_hosts = (int)hosts.size();
    if (size % num_hosts != 0) return -1;
    const int per_host = size / num_hosts;
    for (int host = 0; host < num_hosts; ++host)
        for (int g = 1; g < per_host; ++g)
            if (h[(size_t)host * per_host + g] != h[(size_t)host * per_host + g - 1]) return -2;
    return rank % per_host;
}

inline int placement_device(int rank, int size) {
    char name[1024];
    if (gethostname(name, sizeof name) != 0) name[0] = '\0';
    name[sizeof name - 1] = '\0';
    uint64_t mine = host_hash(name);
    std::vector<uint64_t> all((size_t)size);
    MPICHECK(MPI_Allgather(&mine, 1, MPI_INT64_T, all.data(), 1, MPI_INT64_T, MPI_COMM_WORLD));
    const int dev = placement_from_hashes(all, rank);
    if (dev == -1) {
        fprintf(stderr, "Please make sure that each node has the same number of processes (i.e., used GPUs)\n");
        exit(-1);
    }
    if (dev == -2) {
        fprintf(stderr, "Please make sure that the MPI uses the round-robin process placement policy. For example, if there are 8 processes and 2 nodes, the fist node should handle processes 0-3 while the second node handling processes 4-7.\n");
        exit(-1);
    }
    return dev;
}

// ------------------------------------------------------------------ options

struct Options {
    size_t size = 32u * 1024u * 1024u;  // p2p_matrix.cc:124
    int iters = 128;                    // p2p_matrix.cc:132
    int warmup = -1;                    // -1: 0 for the reference method, 20 for device timing
    std::string mode = "ref";           // ref | uni | bidir | allpairs | sweep | latency | local
    std::string timing = "ref";         // ref (host clock, sync per message) | dev (cudaEvents)
    bool verify = false;
    bool extra = false;                 // print GB/s / us / % tables next to the Gbps ones
    bool first_pair_only = false;
    bool e2e = false;                   // ring mode: host buffers, H2D + transfer + D2H inside the timed region
    size_t sweep_min = 8, sweep_max = 1ull << 30;
    std::string json;
    uint64_t seed = 0x5EED0000ull;
    bool help = false;
};

inline bool parse_size(const char* s, size_t* out) {
    char* end = nullptr;
    double v = strtod(s, &end);
    if (end == s || v < 0) return false;
    size_t mul = 1;
    if (*end == 'k' || *end == 'K') mul = 1ull << 10, ++end;
    else if (*end == 'm' || *end == 'M') mul = 1ull << 20, ++end;
    else if (*end == 'g' || *end == 'G') mul = 1ull << 30, ++end;
    if (*end == 'i') ++end;
    if (*end == 'B' || *end == 'b') ++end;
    if (*end != '\0') return false;
    *out = (size_t)(v * (double)mul + 0.5);
    return true;
}

inline void usage(const char* prog, const char* backend) {
    printf("usage: mpirun -n N %s [options]      (backend: %s)\n"
           "  no options: the reference's behaviour - 32 MiB messages, 128 iterations, host-clock\n"
           "  timing with a sync per message, uni- then bi-directional Gbps matrices.\n"
           "  --size BYTES[K|M|G]   message size (default 32M)\n"
           "  --iters N             timed iterations per cell (default 128)\n"
           "  --warmup N            untimed iterations per cell (default 0 for --timing ref, 20 for dev)\n"
           "  --mode M              ref | uni | bidir | allpairs | alltoall | bcast | ring | sweep | latency | local\n"
           "  --timing T            ref (rank-0 wall clock, sync per message) | dev (cudaEvents, back to back)\n"
           "  --verify              patterned payload, byte-exact check of every cell's receive buffer\n"
           "  --extra               also print GB/s, us per message and %% of 900 GB/s tables\n"
           "  --first-pair          only the pair 0 -> 1 instead of all ordered pairs\n"
           "  --e2e                 (ring) payload starts and ends in pinned HOST memory: H2D + transfer + D2H per step\n"
           "  --sweep-min/--sweep-max BYTES   range of --mode sweep (default 8 .. 1G)\n"
           "  --json FILE           write every measurement as JSON lines\n"
           "  --seed N              pattern seed for --verify\n",
           prog, backend);
}

inline Options parse_args(int argc, char** argv, int rank, const char* backend) {
    Options o;
    for (int i = 1; i < argc; ++i) {
        std::string a = argv[i];
        auto need = [&](const char* what) -> const char* {
            if (i + 1 >= argc) {
                if (rank == 0) printf("%s needs a value\n", what);
                exit(2);
            }
            return argv[++i];
        };
        bool ok = true;
        if (a == "--size") ok = parse_size(need("--size"), &o.size);
        else if (a == "--iters") o.iters = atoi(need("--iters"));
        else if (a == "--warmup") o.warmup = atoi(need("--warmup"));
        else if (a == "--mode") o.mode = need("--mode");
        else if (a == "--timing") o.timing = need("--timing");
        else if (a == "--verify") o.verify = true;
        else if (a == "--extra") o.extra = true;
        else if (a == "--first-pair") o.first_pair_only = true;
        else if (a == "--e2e") o.e2e = true;
        else if (a == "--sweep-min") ok = parse_size(need("--sweep-min"), &o.sweep_min);
        else if (a == "--sweep-max") ok = parse_size(need("--sweep-max"), &o.sweep_max);
        else if (a == "--json") o.json = need("--json");
        else if (a == "--seed") o.seed = strtoull(need("--seed"), nullptr, 0);
        else if (a == "--help" || a == "-h") o.help = true;
        else ok = false;
        if (!ok) {
            if (rank == 0) { printf("bad argument '%s'\n", a.c_str()); usage(argv[0], backend); }
            exit(2);
        }
    }
    if (o.iters < 1) o.iters = 1;
    if (o.timing != "ref" && o.timing != "dev") {
        if (rank == 0) printf("--timing must be ref or dev\n");
        exit(2);
    }
    if (o.warmup < 0) o.warmup = (o.timing == "dev") ? 20 : 0;
    return o;
}

// ------------------------------------------------------------------ the driver

// Backend concept:
//   const char* name() const;
//   void init(int rank, int size, int device);                   // collective
//   void buffers(void* send, size_t sbytes, void* recv, size_t rbytes);  // collective, once
//   void send(const void* buf, size_t bytes, int peer, cudaStream_t s);
//   void recv(void* buf, size_t bytes, int peer, cudaStream_t s);
//   void group_start(); void group_end();
//   void alltoall(const void* send, void* recv, size_t bytes_per_rank, cudaStream_t s);
//   void* symm_alloc(size_t bytes);                               // collective: memory a broadcast may land in
//   void bcast(const void* send, void* recv, size_t bytes, int root, cudaStream_t s);
//   void fill(void* buf, size_t bytes, uint64_t seed, int rank, uint64_t iter, cudaStream_t s);
//   uint64_t check(const void* buf, size_t bytes, uint64_t seed, int rank, uint64_t iter, cudaStream_t s,
//                  uint64_t* digest);                            // mismatching bytes
//   void finalize();
template <class Backend>
class Driver {
public:
    Driver(Backend& be, const Options& opt, int rank, int size) : be_(be), o_(opt), rank_(rank), size_(size) {}

    void setup(size_t max_bytes, int recv_slots, int send_slots = 1) {
        CUDACHECK(cudaStreamCreateWithFlags(&s0_, cudaStreamNonBlocking));  // p2p_matrix.cc:121-122
        CUDACHECK(cudaStreamCreateWithFlags(&s1_, cudaStreamNonBlocking));
        cap_ = max_bytes ? max_bytes : 1;
        slots_ = recv_slots < 1 ? 1 : recv_slots;
        // Sizes are rounded up to whole 2 MiB pages (P2P_DRIVER_ODD_SIZES=1: not).  Measured on two B200s: a 256 MiB + 256 B
        // allocation as the destination of peer DMA copies in both directions moves 1084 GB/s where a 256 MiB one
        // moves 1521 - an allocation that is not a multiple of the large page size gets a slower peer mapping.
        const bool odd = getenv("P2P_DRIVER_ODD_SIZES") != nullptr;
        auto whole_pages = [odd](size_t n) { return odd ? n : (n + (2u << 20) - 1) & ~(size_t)((2u << 20) - 1); };
        send_cap_ = whole_pages(cap_ * (size_t)(send_slots < 1 ? 1 : send_slots));
        CUDACHECK(cudaMalloc(&send_, send_cap_));                           // p2p_matrix.cc:127-130
        recv_cap_ = whole_pages(cap_ * (size_t)slots_ + 256);  // slack: --verify keeps a poisoned guard zone after the payload
        CUDACHECK(cudaMalloc(&recv_, recv_cap_));
        CUDACHECK(cudaMemset(send_, 0, send_cap_));
        CUDACHECK(cudaMemset(recv_, 0, recv_cap_));
        CUDACHECK(cudaDeviceSynchronize());
        be_.buffers(send_, send_cap_, recv_, recv_cap_);
        for (auto& evt : ev_) CUDACHECK(cudaEventCreate(&evt));
        if (!o_.json.empty() && rank_ == 0) {
            json_ = fopen(o_.json.c_str(), "w");
            if (!json_) { printf("cannot open %s\n", o_.json.c_str()); exit(EXIT_FAILURE); }
        }
    }

    void teardown() {
        if (json_) fclose(json_);
        json_ = nullptr;
        // like the reference (p2p_matrix.cc:270-274) nothing else is freed before exit
    }

    // ---- one cell, reference method (p2p_matrix.cc:146-177 / :201-258): returns seconds per message
    double cell_ref(int src, int dst, size_t bytes, bool bidir) {
        for (int i = 0; i < o_.warmup; ++i) one_message(src, dst, bytes, bidir, true);
        MPI_Barrier(MPI_COMM_WORLD);
        auto start = std::chrono::system_clock::now();
        for (int i = 0; i < o_.iters; ++i) one_message(src, dst, bytes, bidir, true);
        MPI_Barrier(MPI_COMM_WORLD);
        auto end = std::chrono::system_clock::now();
        std::chrono::duration<double> elapsed = end - start;
        double t = elapsed.count() / o_.iters;
        MPICHECK(MPI_Bcast(&t, 1, MPI_DOUBLE, 0, MPI_COMM_WORLD));  // rank 0's clock is the one reported
        return t;
    }

    // ---- one cell, device timing: cudaEvents around `iters` back-to-back messages, max over ranks
    double cell_dev(int src, int dst, size_t bytes, bool bidir) {
        for (int i = 0; i < o_.warmup; ++i) one_message(src, dst, bytes, bidir, false);
        sync_streams();
        MPI_Barrier(MPI_COMM_WORLD);
        const bool active = (rank_ == src || rank_ == dst);
        float ms = 0.f;
        if (active) {
            CUDACHECK(cudaEventRecord(ev_[0], s0_));
            if (bidir) CUDACHECK(cudaEventRecord(ev_[2], s1_));
            for (int i = 0; i < o_.iters; ++i) one_message(src, dst, bytes, bidir, false);
            CUDACHECK(cudaEventRecord(ev_[1], s0_));
            if (bidir) CUDACHECK(cudaEventRecord(ev_[3], s1_));
            sync_streams();
            CUDACHECK(cudaEventElapsedTime(&ms, ev_[0], ev_[1]));
            if (bidir) {
                float ms1 = 0.f;
                CUDACHECK(cudaEventElapsedTime(&ms1, ev_[2], ev_[3]));
                ms = std::max(ms, ms1);
            }
        }
        double t = (double)ms * 1e-3 / o_.iters, tmax = 0;
        MPICHECK(MPI_Allreduce(&t, &tmax, 1, MPI_DOUBLE, MPI_MAX, MPI_COMM_WORLD));
        return tmax;
    }

    double cell(int src, int dst, size_t bytes, bool bidir) {
        return o_.timing == "dev" ? cell_dev(src, dst, bytes, bidir) : cell_ref(src, dst, bytes, bidir);
    }

    // ---- byte-exact verification of one cell (patterned send, poisoned receive)
    // returns mismatching bytes summed over the receivers of this cell
    uint64_t verify_cell(int src, int dst, size_t bytes, bool bidir, uint64_t iter) {
        const bool active = (rank_ == src || rank_ == dst);
        const size_t poison = std::min(bytes + 64, recv_cap_);  // payload + a guard zone
        if (active) {
            be_.fill(send_, bytes, o_.seed, rank_, iter, s0_);
            CUDACHECK(cudaMemsetAsync(recv_, 0xA5, poison, s0_));
            CUDACHECK(cudaStreamSynchronize(s0_));
        }
        MPI_Barrier(MPI_COMM_WORLD);
        one_message(src, dst, bytes, bidir, true);
        uint64_t bad = 0, digest = 0, digest_sum = 0;
        if (rank_ == dst) { bad += be_.check(recv_, bytes, o_.seed, src, iter, s0_, &digest); digest_sum += digest; }
        if (bidir && rank_ == src) { bad += be_.check(recv_, bytes, o_.seed, dst, iter, s0_, &digest); digest_sum += digest; }
        // the byte after the payload must still be poison (no overrun) when there is room
        if ((rank_ == dst || (bidir && rank_ == src)) && bytes < poison) {
            uint8_t guard[64] = {};
            CUDACHECK(cudaMemcpy(guard, (uint8_t*)recv_ + bytes, poison - bytes, cudaMemcpyDeviceToHost));
            for (size_t i = 0; i < poison - bytes; ++i)
                if (guard[i] != 0xA5) ++bad;
        }
        unsigned long long b = bad, total = 0, dg = digest_sum, dg_total = 0;
        MPICHECK(MPI_Allreduce(&b, &total, 1, MPI_UINT64_T, MPI_SUM, MPI_COMM_WORLD));
        MPICHECK(MPI_Allreduce(&dg, &dg_total, 1, MPI_UINT64_T, MPI_SUM, MPI_COMM_WORLD));
        if (json_) {
            // digest of what the receiver(s) hold after the transfer (sum over both ends when bidir)
            fprintf(json_,
                    "{\"backend\": \"%s\", \"mode\": \"verify\", \"dir\": \"%s\", \"ranks\": %d, \"src\": %d, \"dst\": %d, "
                    "\"bytes\": %zu, \"seed\": %llu, \"iter\": %llu, \"mismatches\": %llu, \"digest\": \"0x%016llx\"}\n",
                    be_.name(), bidir ? "bidir" : "uni", size_, src, dst, bytes, (unsigned long long)o_.seed,
                    (unsigned long long)iter, total, dg_total);
            fflush(json_);
        }
        verified_cells_++;
        verified_bytes_ += bytes * (bidir ? 2 : 1);
        verify_failures_ += total;
        if (total && rank_ == 0)
            printf("\nVERIFY FAILED: %s %d->%d, %zu bytes: %llu bytes differ\n", bidir ? "bidir" : "uni", src, dst,
                   bytes, total);
        return total;
    }

    // ---- the reference's two tables (p2p_matrix.cc:133-267), same text, same order
    void run_ref_tables(bool do_uni, bool do_bidir) {
        for (int pass = 0; pass < 2; ++pass) {
            const bool bidir = (pass == 1);
            if (bidir ? !do_bidir : !do_uni) continue;
            if (rank_ == 0) {
                printf(bidir ? "\nEvaluating the Bi-Directional NCCL P2P Bandwidth (Gbps)\n"
                             : "Evaluating the Uni-Directional NCCL P2P Bandwidth (Gbps)\n");
                printf("   D\\D");
                for (int i = 0; i < size_; ++i) printf("%6d ", i);
                printf("\n");
            }
            std::vector<double> secs((size_t)size_ * size_, 0.0);
            for (int src = 0; src < size_; ++src) {
                if (rank_ == 0) printf("%6d ", src);
                for (int dst = 0; dst < size_; ++dst) {
                    MPI_Barrier(MPI_COMM_WORLD);
                    if (src == dst) {
                        if (rank_ == 0) printf("%6.02f ", 0.);
                        continue;
                    }
                    if (o_.first_pair_only && !(src == 0 && dst == 1)) {
                        if (rank_ == 0) printf("%6.02f ", 0.);
                        continue;
                    }
                    if (o_.verify) verify_cell(src, dst, o_.size, bidir, (uint64_t)(pass * size_ * size_ + src * size_ + dst));
                    const double t = cell(src, dst, o_.size, bidir);
                    secs[(size_t)src * size_ + dst] = t;
                    double throughput = o_.size * 8. / t / 1e9;  // Gbps
                    if (bidir) throughput *= 2;
                    if (rank_ == 0) {
                        printf("%6.02f ", throughput);
                        fflush(stdout);
                    }
                    record(bidir ? "bidir" : "uni", src, dst, o_.size, t, bidir ? 2.0 : 1.0);
                }
                if (rank_ == 0) printf("\n");
            }
            if (o_.extra && rank_ == 0) print_extra(secs, bidir);
        }
    }

    // ---- all-pairs concurrent: every rank sends `bytes` to each peer at once (BASELINE cfg 4)
    void run_allpairs() {
        const size_t bytes = o_.size;
        const int np = size_ - 1;
        if (np < 1) { if (rank_ == 0) printf("all-pairs needs at least 2 ranks\n"); return; }
        auto round = [&](bool sync) {
            be_.group_start();
            int slot = 0;
            for (int p = 0; p < size_; ++p) {
                if (p == rank_) continue;
                be_.recv((uint8_t*)recv_ + (size_t)slot * bytes, bytes, p, s1_);
                be_.send(send_, bytes, p, s0_);
                ++slot;
            }
            be_.group_end();
            if (sync) sync_streams();
        };
        uint64_t bad_total = 0;
        if (o_.verify) {
            be_.fill(send_, bytes, o_.seed, rank_, 777, s0_);
            CUDACHECK(cudaMemsetAsync(recv_, 0xA5, bytes * (size_t)np, s0_));
            CUDACHECK(cudaStreamSynchronize(s0_));
            MPI_Barrier(MPI_COMM_WORLD);
            round(true);
            uint64_t bad = 0, digest;
            int slot = 0;
            for (int p = 0; p < size_; ++p) {
                if (p == rank_) continue;
                bad += be_.check((uint8_t*)recv_ + (size_t)slot * bytes, bytes, o_.seed, p, 777, s0_, &digest);
                ++slot;
            }
            unsigned long long b = bad, total = 0;
            MPICHECK(MPI_Allreduce(&b, &total, 1, MPI_UINT64_T, MPI_SUM, MPI_COMM_WORLD));
            bad_total = total;
            verified_cells_ += (uint64_t)size_ * np;
            verified_bytes_ += (uint64_t)size_ * np * bytes;
            verify_failures_ += total;
        }
        for (int i = 0; i < std::max(o_.warmup, 1); ++i) round(true);
        double t;
        if (o_.timing == "dev") {
            MPI_Barrier(MPI_COMM_WORLD);
            CUDACHECK(cudaEventRecord(ev_[0], s0_));
            CUDACHECK(cudaEventRecord(ev_[2], s1_));
            for (int i = 0; i < o_.iters; ++i) round(false);
            CUDACHECK(cudaEventRecord(ev_[1], s0_));
            CUDACHECK(cudaEventRecord(ev_[3], s1_));
            sync_streams();
            float ms0, ms1;
            CUDACHECK(cudaEventElapsedTime(&ms0, ev_[0], ev_[1]));
            CUDACHECK(cudaEventElapsedTime(&ms1, ev_[2], ev_[3]));
            double mine = std::max(ms0, ms1) * 1e-3 / o_.iters;
            MPICHECK(MPI_Allreduce(&mine, &t, 1, MPI_DOUBLE, MPI_MAX, MPI_COMM_WORLD));
        } else {
            MPI_Barrier(MPI_COMM_WORLD);
            auto start = std::chrono::system_clock::now();
            for (int i = 0; i < o_.iters; ++i) round(true);
            MPI_Barrier(MPI_COMM_WORLD);
            std::chrono::duration<double> el = std::chrono::system_clock::now() - start;
            t = el.count() / o_.iters;
            MPICHECK(MPI_Bcast(&t, 1, MPI_DOUBLE, 0, MPI_COMM_WORLD));
        }
        const double flows = (double)size_ * np;
        const double agg = flows * bytes / t / 1e9;
        if (rank_ == 0) {
            printf("All-pairs concurrent (%s, timing %s): %d ranks x %d peers x %zu bytes per round\n", be_.name(),
                   o_.timing.c_str(), size_, np, bytes);
            printf("  round %.2f us   aggregate %.2f GB/s   per-GPU egress %.2f GB/s (%.1f%% of %.0f)%s\n", t * 1e6, agg,
                   agg / size_, 100.0 * agg / size_ / kLinkGBps, kLinkGBps,
                   o_.verify ? (bad_total ? "   VERIFY FAILED" : "   verify OK") : "");
            fflush(stdout);
        }
        record("allpairs", -1, -1, bytes, t, flows);
    }

    // ---- all-to-all: block p of the send buffer to rank p, block p of the receive buffer from rank p
    // (the all-pairs mode with distinct payloads and the local block included)
    void run_alltoall() {
        const size_t bytes = o_.size;
        auto round = [&](bool sync) {
            be_.alltoall(send_, recv_, bytes, s0_);
            if (sync) CUDACHECK(cudaStreamSynchronize(s0_));
        };
        unsigned long long bad_total = 0;
        if (o_.verify) {
            for (int p = 0; p < size_; ++p) be_.fill((uint8_t*)send_ + (size_t)p * bytes, bytes, o_.seed, rank_, (uint64_t)p, s0_);
            CUDACHECK(cudaMemsetAsync(recv_, 0xA5, recv_cap_, s0_));
            CUDACHECK(cudaStreamSynchronize(s0_));
            MPI_Barrier(MPI_COMM_WORLD);
            round(true);
            uint64_t bad = 0, digest;
            for (int p = 0; p < size_; ++p)
                bad += be_.check((uint8_t*)recv_ + (size_t)p * bytes, bytes, o_.seed, p, (uint64_t)rank_, s0_, &digest);
            uint8_t guard[64];
            CUDACHECK(cudaMemcpy(guard, (uint8_t*)recv_ + (size_t)size_ * bytes, sizeof guard, cudaMemcpyDeviceToHost));
            for (uint8_t gb : guard) if (gb != 0xA5) ++bad;
            unsigned long long b = bad;
            MPICHECK(MPI_Allreduce(&b, &bad_total, 1, MPI_UINT64_T, MPI_SUM, MPI_COMM_WORLD));
            verified_cells_ += (uint64_t)size_ * size_;
            verified_bytes_ += (uint64_t)size_ * size_ * bytes;
            verify_failures_ += bad_total;
        }
        for (int i = 0; i < std::max(o_.warmup, 1); ++i) round(true);
        double t;
        MPI_Barrier(MPI_COMM_WORLD);
        if (o_.timing == "dev") {
            CUDACHECK(cudaEventRecord(ev_[0], s0_));
            for (int i = 0; i < o_.iters; ++i) round(false);
            CUDACHECK(cudaEventRecord(ev_[1], s0_));
            CUDACHECK(cudaStreamSynchronize(s0_));
            float ms;
            CUDACHECK(cudaEventElapsedTime(&ms, ev_[0], ev_[1]));
            double mine = ms * 1e-3 / o_.iters;
            MPICHECK(MPI_Allreduce(&mine, &t, 1, MPI_DOUBLE, MPI_MAX, MPI_COMM_WORLD));
        } else {
            auto start = std::chrono::system_clock::now();
            for (int i = 0; i < o_.iters; ++i) round(true);
            MPI_Barrier(MPI_COMM_WORLD);
            std::chrono::duration<double> el = std::chrono::system_clock::now() - start;
            t = el.count() / o_.iters;
            MPICHECK(MPI_Bcast(&t, 1, MPI_DOUBLE, 0, MPI_COMM_WORLD));
        }
        const double flows = (double)size_ * (size_ - 1);  // blocks that cross NVLink
        if (rank_ == 0) {
            printf("All-to-all (%s, timing %s): %d ranks, %zu bytes per block\n", be_.name(), o_.timing.c_str(), size_, bytes);
            printf("  round %.2f us   aggregate over NVLink %.2f GB/s   per-GPU egress %.2f GB/s%s\n", t * 1e6,
                   flows * bytes / t / 1e9, flows * bytes / t / 1e9 / size_,
                   o_.verify ? (bad_total ? "   VERIFY FAILED" : "   verify OK") : "");
            fflush(stdout);
        }
        record("alltoall", -1, -1, bytes, t, flows > 0 ? flows : 1.0);
    }

    // ---- one-to-many: rank 0's `bytes` into every rank's receive buffer.  The reference can
    // only write this as N-1 sends from the root (p2p_matrix.cc:141-172 walks the peers one ncclSend at a time); the
    // NCCL baseline does exactly that, grouped; the engine stores once into the NVSwitch multicast mapping.
    void run_bcast() {
        const size_t bytes = o_.size;
        const int root = 0;
        uint8_t* dst = (uint8_t*)be_.symm_alloc(bytes + 256);
        auto round = [&](bool sync) {
            be_.bcast(send_, dst, bytes, root, s0_);
            if (sync) CUDACHECK(cudaStreamSynchronize(s0_));
        };
        unsigned long long bad_total = 0;
        if (o_.verify) {
            for (int rep = 0; rep < 2; ++rep) {   // twice: flags and counters carry over from call to call
                if (rank_ == root) be_.fill(send_, bytes, o_.seed, root, 99 + (uint64_t)rep, s0_);
                CUDACHECK(cudaMemsetAsync(dst, 0xA5, bytes + 256, s0_));
                CUDACHECK(cudaStreamSynchronize(s0_));
                MPI_Barrier(MPI_COMM_WORLD);
                round(true);
                uint64_t digest = 0;
                unsigned long long b = be_.check(dst, bytes, o_.seed, root, 99 + (uint64_t)rep, s0_, &digest);
                uint8_t guard[64];
                CUDACHECK(cudaMemcpy(guard, dst + bytes, sizeof guard, cudaMemcpyDeviceToHost));
                for (uint8_t gb : guard) if (gb != 0xA5) ++b;
                unsigned long long total = 0;
                MPICHECK(MPI_Allreduce(&b, &total, 1, MPI_UINT64_T, MPI_SUM, MPI_COMM_WORLD));
                bad_total += total;
                std::vector<unsigned long long> digs((size_t)size_);
                unsigned long long dg = digest;
                MPICHECK(MPI_Allgather(&dg, 1, MPI_UINT64_T, digs.data(), 1, MPI_UINT64_T, MPI_COMM_WORLD));
                if (json_)
                    for (int r = 0; r < size_; ++r)
                        fprintf(json_,
                                "{\"backend\": \"%s\", \"mode\": \"verify\", \"dir\": \"bcast\", \"ranks\": %d, \"src\": %d, \"dst\": %d, "
                                "\"bytes\": %zu, \"seed\": %llu, \"iter\": %d, \"mismatches\": %llu, \"digest\": \"0x%016llx\"}\n",
                                be_.name(), size_, root, r, bytes, (unsigned long long)o_.seed, 99 + rep, total, digs[(size_t)r]);
                MPI_Barrier(MPI_COMM_WORLD);
            }
            verified_cells_ += 2 * (uint64_t)size_;
            verified_bytes_ += 2 * (uint64_t)size_ * bytes;
            verify_failures_ += bad_total;
        }
        for (int i = 0; i < std::max(o_.warmup, 3); ++i) round(true);
        double t;
        MPI_Barrier(MPI_COMM_WORLD);
        if (o_.timing == "dev") {
            CUDACHECK(cudaEventRecord(ev_[0], s0_));
            for (int i = 0; i < o_.iters; ++i) round(false);
            CUDACHECK(cudaEventRecord(ev_[1], s0_));
            CUDACHECK(cudaStreamSynchronize(s0_));
            float ms;
            CUDACHECK(cudaEventElapsedTime(&ms, ev_[0], ev_[1]));
            double mine = ms * 1e-3 / o_.iters;
            MPICHECK(MPI_Allreduce(&mine, &t, 1, MPI_DOUBLE, MPI_MAX, MPI_COMM_WORLD));
        } else {
            auto start = std::chrono::system_clock::now();
            for (int i = 0; i < o_.iters; ++i) round(true);
            MPI_Barrier(MPI_COMM_WORLD);
            std::chrono::duration<double> el = std::chrono::system_clock::now() - start;
            t = el.count() / o_.iters;
            MPICHECK(MPI_Bcast(&t, 1, MPI_DOUBLE, 0, MPI_COMM_WORLD));
        }
        const double receivers = (double)std::max(size_ - 1, 1);
        if (rank_ == 0) {
            printf("Broadcast from rank %d (%s, timing %s): %d ranks, %zu bytes\n", root, be_.name(), o_.timing.c_str(), size_, bytes);
            printf("  round %.2f us   %.2f GB/s per receiver   %.2f GB/s delivered in total%s\n", t * 1e6, bytes / t / 1e9,
                   receivers * bytes / t / 1e9, o_.verify ? (bad_total ? "   VERIFY FAILED" : "   verify OK") : "");
            fflush(stdout);
        }
        record("bcast", root, -1, bytes, t, receivers);
    }

    // ---- ring: every rank sends `bytes` to (rank+1)%N and receives from (rank-1)%N at once.
    // N = 1 is a local copy through the same send/recv path.  This is bench.py's step.
    void run_ring() {
        const size_t bytes = o_.size;
        const int right = (rank_ + 1) % size_, left = (rank_ + size_ - 1) % size_;
        uint8_t *hsrc = nullptr, *hdst = nullptr;
        if (o_.e2e) {
            CUDACHECK(cudaMallocHost(&hsrc, bytes ? bytes : 1));
            CUDACHECK(cudaMallocHost(&hdst, bytes ? bytes : 1));
            memset(hsrc, 0x3c, bytes);
        }
        auto step = [&](bool sync) {
            if (o_.e2e) CUDACHECK(cudaMemcpyAsync(send_, hsrc, bytes, cudaMemcpyHostToDevice, s0_));
            be_.group_start();
            be_.recv(recv_, bytes, left, s0_);
            be_.send(send_, bytes, right, s0_);
            be_.group_end();
            if (o_.e2e) CUDACHECK(cudaMemcpyAsync(hdst, recv_, bytes, cudaMemcpyDeviceToHost, s0_));
            if (sync) CUDACHECK(cudaStreamSynchronize(s0_));
        };
        if (o_.verify) {
            be_.fill(send_, bytes, o_.seed, rank_, 4242, s0_);
            CUDACHECK(cudaMemsetAsync(recv_, 0xA5, bytes, s0_));
            CUDACHECK(cudaStreamSynchronize(s0_));
            MPI_Barrier(MPI_COMM_WORLD);
            const bool keep = o_.e2e;
            o_.e2e = false;
            step(true);
            o_.e2e = keep;
            uint64_t digest = 0;
            unsigned long long b = be_.check(recv_, bytes, o_.seed, left, 4242, s0_, &digest), total = 0;
            MPICHECK(MPI_Allreduce(&b, &total, 1, MPI_UINT64_T, MPI_SUM, MPI_COMM_WORLD));
            // one record per receiver: what it holds after the step (compared across backends by bench.py)
            std::vector<unsigned long long> digs((size_t)size_), bads((size_t)size_);
            unsigned long long dg = digest;
            MPICHECK(MPI_Allgather(&dg, 1, MPI_UINT64_T, digs.data(), 1, MPI_UINT64_T, MPI_COMM_WORLD));
            MPICHECK(MPI_Allgather(&b, 1, MPI_UINT64_T, bads.data(), 1, MPI_UINT64_T, MPI_COMM_WORLD));
            if (json_)
                for (int r = 0; r < size_; ++r)
                    fprintf(json_,
                            "{\"backend\": \"%s\", \"mode\": \"verify\", \"dir\": \"ring\", \"ranks\": %d, \"src\": %d, \"dst\": %d, "
                            "\"bytes\": %zu, \"seed\": %llu, \"iter\": 4242, \"mismatches\": %llu, \"digest\": \"0x%016llx\"}\n",
                            be_.name(), size_, (r + size_ - 1) % size_, r, bytes, (unsigned long long)o_.seed, bads[(size_t)r], digs[(size_t)r]);
            verified_cells_ += (uint64_t)size_;
            verified_bytes_ += (uint64_t)size_ * bytes;
            verify_failures_ += total;
        }
        for (int i = 0; i < std::max(o_.warmup, 3); ++i) step(true);
        double t;
        MPI_Barrier(MPI_COMM_WORLD);
        if (o_.timing == "dev") {
            CUDACHECK(cudaEventRecord(ev_[0], s0_));
            for (int i = 0; i < o_.iters; ++i) step(false);
            CUDACHECK(cudaEventRecord(ev_[1], s0_));
            CUDACHECK(cudaStreamSynchronize(s0_));
            float ms;
            CUDACHECK(cudaEventElapsedTime(&ms, ev_[0], ev_[1]));
            double mine = ms * 1e-3 / o_.iters;
            MPICHECK(MPI_Allreduce(&mine, &t, 1, MPI_DOUBLE, MPI_MAX, MPI_COMM_WORLD));
        } else {
            auto start = std::chrono::system_clock::now();
            for (int i = 0; i < o_.iters; ++i) step(true);
            MPI_Barrier(MPI_COMM_WORLD);
            std::chrono::duration<double> el = std::chrono::system_clock::now() - start;
            t = el.count() / o_.iters;
            MPICHECK(MPI_Bcast(&t, 1, MPI_DOUBLE, 0, MPI_COMM_WORLD));
        }
        const double agg = (double)size_ * bytes / t / 1e9;
        if (rank_ == 0) {
            printf("Ring%s (%s, timing %s): %d ranks, %zu bytes each to the next rank\n", o_.e2e ? " host-to-host" : "",
                   be_.name(), o_.timing.c_str(), size_, bytes);
            printf("  step %.2f us   aggregate %.2f GB/s   per GPU %.2f GB/s\n", t * 1e6, agg, agg / size_);
            fflush(stdout);
        }
        record(o_.e2e ? "ring_e2e" : "ring", -1, -1, bytes, t, (double)size_);
    }

    // ---- size sweep on the pair 0 -> 1: streaming bandwidth and ping-pong latency per size
    void run_sweep(bool latency_only) {
        if (size_ < 2) { run_local(); return; }
        if (rank_ == 0) {
            printf("Message-size sweep, pair 0 <-> 1 (%s, device timing)\n", be_.name());
            printf("%12s %12s %12s %12s %10s\n", "bytes", "oneway_us", "stream_us", "GB/s", "%of900");
        }
        const int keep_iters = o_.iters, keep_warm = o_.warmup;
        const std::string keep_timing = o_.timing;
        o_.timing = "dev";
        for (size_t bytes = o_.sweep_min; bytes <= o_.sweep_max && bytes <= cap_; bytes *= 2) {
            // enough iterations for a stable figure, bounded for the large sizes
            int iters = (int)std::min<double>(keep_iters, std::max(8.0, 2e9 / (double)std::max<size_t>(bytes, 1)));
            if (bytes <= 65536) iters = std::max(iters, 200);
            o_.iters = iters;
            o_.warmup = std::max(5, std::min(keep_warm > 0 ? keep_warm : 20, iters));
            if (o_.verify) verify_cell(0, 1, bytes, false, bytes);
            const double pp = pingpong(0, 1, bytes);           // seconds per one-way hop
            double st = 0;
            if (!latency_only) st = cell_dev(0, 1, bytes, false);
            if (rank_ == 0) {
                if (latency_only) printf("%12zu %12.2f %12s %12s %10s\n", bytes, pp * 1e6, "-", "-", "-");
                else printf("%12zu %12.2f %12.2f %12.2f %9.1f%%\n", bytes, pp * 1e6, st * 1e6, bytes / st / 1e9,
                            100.0 * bytes / st / 1e9 / kLinkGBps);
                fflush(stdout);
            }
            record("latency", 0, 1, bytes, pp, 1.0);
            if (!latency_only) record("stream", 0, 1, bytes, st, 1.0);
        }
        o_.iters = keep_iters;
        o_.warmup = keep_warm;
        o_.timing = keep_timing;
    }

    // ---- N = 1: local device-to-device copy through the same send/recv path (self-send)
    void run_local() {
        if (rank_ == 0) {
            printf("Local copy on one GPU via self send/recv (%s, device timing); GB/s counts read+write\n", be_.name());
            printf("%12s %12s %12s\n", "bytes", "us", "GB/s(2B/t)");
        }
        for (size_t bytes = o_.sweep_min; bytes <= o_.sweep_max && bytes <= cap_; bytes *= 2) {
            int iters = (int)std::min<double>(o_.iters, std::max(8.0, 4e9 / (double)std::max<size_t>(bytes, 1)));
            auto one = [&]() {
                be_.group_start();
                be_.recv(recv_, bytes, rank_, s0_);
                be_.send(send_, bytes, rank_, s0_);
                be_.group_end();
            };
            if (o_.verify) {
                be_.fill(send_, bytes, o_.seed, rank_, bytes, s0_);
                CUDACHECK(cudaMemsetAsync(recv_, 0xA5, bytes, s0_));
                one();
                uint64_t digest;
                uint64_t bad = be_.check(recv_, bytes, o_.seed, rank_, bytes, s0_, &digest);
                verified_cells_++;
                verified_bytes_ += bytes;
                verify_failures_ += bad;
            }
            for (int i = 0; i < 5; ++i) one();
            CUDACHECK(cudaStreamSynchronize(s0_));
            CUDACHECK(cudaEventRecord(ev_[0], s0_));
            for (int i = 0; i < iters; ++i) one();
            CUDACHECK(cudaEventRecord(ev_[1], s0_));
            CUDACHECK(cudaStreamSynchronize(s0_));
            float ms;
            CUDACHECK(cudaEventElapsedTime(&ms, ev_[0], ev_[1]));
            const double t = ms * 1e-3 / iters;
            if (rank_ == 0) {
                printf("%12zu %12.2f %12.2f\n", bytes, t * 1e6, 2.0 * bytes / t / 1e9);
                fflush(stdout);
            }
            record("local", rank_, rank_, bytes, t, 2.0);
        }
    }

    void finish_report() {
        if (o_.verify && rank_ == 0) {
            if (verify_failures_ == 0)
                printf("verify: OK (%llu transfers, %llu payload bytes, every byte compared)\n",
                       (unsigned long long)verified_cells_, (unsigned long long)verified_bytes_);
            else
                printf("verify: FAILED (%llu bytes differ)\n", (unsigned long long)verify_failures_);
            fflush(stdout);
        }
    }
    uint64_t verify_failures() const { return verify_failures_; }

private:
    // One message exactly as the reference issues it (p2p_matrix.cc:155-171 / :210-252).
    void one_message(int src, int dst, size_t bytes, bool bidir, bool sync) {
        if (!bidir) {
            if (rank_ == src) {
                be_.send(send_, bytes, dst, s0_);
                if (sync) CUDACHECK(cudaStreamSynchronize(s0_));
            } else if (rank_ == dst) {
                be_.recv(recv_, bytes, src, s0_);
                if (sync) CUDACHECK(cudaStreamSynchronize(s0_));
            }
        } else if (rank_ == src || rank_ == dst) {
            const int peer = rank_ == src ? dst : src;
            be_.group_start();
            be_.send(send_, bytes, peer, s0_);
            be_.recv(recv_, bytes, peer, s1_);
            be_.group_end();
            if (sync) sync_streams();
        }
    }

    void sync_streams() {
        CUDACHECK(cudaStreamSynchronize(s0_));
        CUDACHECK(cudaStreamSynchronize(s1_));
    }

    // Ping-pong a <-> b with device timing on `a`; returns seconds per one-way message.
    double pingpong(int a, int b, size_t bytes) {
        auto hop = [&](int n) {
            for (int i = 0; i < n; ++i) {
                if (rank_ == a) { be_.send(send_, bytes, b, s0_); be_.recv(recv_, bytes, b, s0_); }
                else if (rank_ == b) { be_.recv(recv_, bytes, a, s0_); be_.send(send_, bytes, a, s0_); }
            }
        };
        hop(o_.warmup);
        CUDACHECK(cudaStreamSynchronize(s0_));
        MPI_Barrier(MPI_COMM_WORLD);
        float ms = 0.f;
        if (rank_ == a || rank_ == b) {
            CUDACHECK(cudaEventRecord(ev_[0], s0_));
            hop(o_.iters);
            CUDACHECK(cudaEventRecord(ev_[1], s0_));
            CUDACHECK(cudaStreamSynchronize(s0_));
            CUDACHECK(cudaEventElapsedTime(&ms, ev_[0], ev_[1]));
        }
        double t = (double)ms * 1e-3 / o_.iters / 2.0, tmax = 0;
        MPICHECK(MPI_Allreduce(&t, &tmax, 1, MPI_DOUBLE, MPI_MAX, MPI_COMM_WORLD));
        return tmax;
    }

    void print_extra(const std::vector<double>& secs, bool bidir) {
        const double dirs = bidir ? 2.0 : 1.0;
        const char* names[3] = {"GB/s", "us per message", "% of link rate"};
        for (int k = 0; k < 3; ++k) {
            printf("\n%s, %s (%s, timing %s, %zu bytes%s; link = %.0f GB/s per direction)\n",
                   bidir ? "Bi-directional" : "Uni-directional", names[k], be_.name(), o_.timing.c_str(), o_.size,
                   bidir ? " each way" : "", kLinkGBps);
            printf("   D\\D");
            for (int i = 0; i < size_; ++i) printf("%9d ", i);
            printf("\n");
            for (int s = 0; s < size_; ++s) {
                printf("%6d ", s);
                for (int d = 0; d < size_; ++d) {
                    const double t = secs[(size_t)s * size_ + d];
                    double v = 0;
                    if (t > 0) {
                        const double gbs = dirs * o_.size / t / 1e9;
                        v = k == 0 ? gbs : k == 1 ? t * 1e6 : 100.0 * gbs / (dirs * kLinkGBps);
                    }
                    printf("%9.2f ", v);
                }
                printf("\n");
            }
        }
        fflush(stdout);
    }

    void record(const char* mode, int src, int dst, size_t bytes, double secs, double flows) {
        if (!json_) return;
        fprintf(json_,
                "{\"backend\": \"%s\", \"mode\": \"%s\", \"timing\": \"%s\", \"ranks\": %d, \"src\": %d, \"dst\": %d, "
                "\"bytes\": %zu, \"seconds\": %.9g, \"us\": %.4f, \"GBps\": %.4f, \"iters\": %d, \"warmup\": %d}\n",
                be_.name(), mode, (!strcmp(mode, "latency") || !strcmp(mode, "stream") || !strcmp(mode, "local")) ? "dev" : o_.timing.c_str(),
                size_, src, dst, bytes, secs, secs * 1e6, secs > 0 ? flows * bytes / secs / 1e9 : 0.0, o_.iters, o_.warmup);
        fflush(json_);
    }

    Backend& be_;
    Options o_;
    int rank_, size_;
    cudaStream_t s0_ = nullptr, s1_ = nullptr;
    void* send_ = nullptr;
    void* recv_ = nullptr;
    size_t cap_ = 0, recv_cap_ = 0, send_cap_ = 0;
    int slots_ = 1;
    cudaEvent_t ev_[4] = {};
    FILE* json_ = nullptr;
    uint64_t verified_cells_ = 0, verified_bytes_ = 0, verify_failures_ = 0;
};

// Everything main() does after MPI is up and the backend exists.
template <class Backend>
int run_main(Backend& be, int argc, char** argv) {
    int rank, world_size, provided;
    MPICHECK(MPI_Init_thread(&argc, &argv, MPI_THREAD_MULTIPLE, &provided));  // p2p_matrix.cc:105-108
    assert(provided == MPI_THREAD_MULTIPLE);
    MPICHECK(MPI_Comm_rank(MPI_COMM_WORLD, &rank));
    MPICHECK(MPI_Comm_size(MPI_COMM_WORLD, &world_size));
    Options opt = parse_args(argc, argv, rank, be.name());
    if (opt.help) {
        if (rank == 0) usage(argv[0], be.name());
        return 0;
    }
    int gpu_id = placement_device(rank, world_size);  // p2p_matrix.cc:109
    int ndev = 0;
    CUDACHECK(cudaGetDeviceCount(&ndev));
    if (gpu_id >= ndev) {
        // The reference would fail in cudaSetDevice here.  Sharing a GPU between ranks is refused outright: two ranks
        // whose kernels wait for each other's flags on one GPU can starve each other for ever.
        if (rank == 0)
            printf("Failed: %d ranks on this host but only %d GPU%s; launch at most one rank per GPU\n", world_size, ndev,
                   ndev == 1 ? "" : "s");
        exit(EXIT_FAILURE);
    }
    CUDACHECK(cudaSetDevice(gpu_id));                 // p2p_matrix.cc:119
    be.init(rank, world_size, gpu_id);

    Driver<Backend> d(be, opt, rank, world_size);
    const bool sweepish = opt.mode == "sweep" || opt.mode == "latency" || opt.mode == "local";
    const size_t max_bytes = sweepish ? std::max(opt.size, opt.sweep_max) : opt.size;
    if (opt.mode == "alltoall") d.setup(max_bytes, world_size, world_size);
    else d.setup(max_bytes, opt.mode == "allpairs" ? world_size - 1 : 1);

    if (opt.mode == "ref") d.run_ref_tables(true, true);
    else if (opt.mode == "uni") d.run_ref_tables(true, false);
    else if (opt.mode == "bidir") d.run_ref_tables(false, true);
    else if (opt.mode == "allpairs") d.run_allpairs();
    else if (opt.mode == "ring") d.run_ring();
    else if (opt.mode == "alltoall") d.run_alltoall();
    else if (opt.mode == "bcast") d.run_bcast();
    else if (opt.mode == "sweep") d.run_sweep(false);
    else if (opt.mode == "latency") d.run_sweep(true);
    else if (opt.mode == "local") d.run_local();
    else {
        if (rank == 0) printf("unknown --mode %s\n", opt.mode.c_str());
        return 2;
    }
    d.finish_report();
    d.teardown();
    const int rc = d.verify_failures() ? 1 : 0;
    be.finalize();                                    // p2p_matrix.cc:270
    MPICHECK(MPI_Barrier(MPI_COMM_WORLD));            // p2p_matrix.cc:271
    // MPI_Finalize is not called, as in the reference (p2p_matrix.cc:272)
    return rc;
}

}  // namespace bench
