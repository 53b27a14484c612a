// Prints the results of the benchmark driver's pure host functions as JSON so pytest can
// compare them with the oracle.  No CUDA or MPI call is made.
//   driver_unit placement <rank> <h0> <h1> ...      -> {"device": d}
//   driver_unit hash <hostname>                       -> {"hash": h}
//   driver_unit size <text>                           -> {"ok": 0|1, "bytes": n}
//   driver_unit gbps <bytes> <seconds> <bidir>        -> formatted like the report cell
#include "matrix_driver.hpp"

int main(int argc, char** argv) {
    if (argc < 2) return 2;
    std::string cmd = argv[1];
    if (cmd == "placement" && argc >= 4) {
        int rank = atoi(argv[2]);
        std::vector<uint64_t> h;
        for (int i = 3; i < argc; ++i) h.push_back(strtoull(argv[i], nullptr, 0));
        printf("{\"device\": %d}\n", bench::placement_from_hashes(h, rank));
        return 0;
    }
    if (cmd == "hash" && argc == 3) {
        printf("{\"hash\": %llu}\n", (unsigned long long)bench::host_hash(argv[2]));
        return 0;
    }
    if (cmd == "size" && argc == 3) {
        size_t n = 0;
        bool ok = bench::parse_size(argv[2], &n);
        printf("{\"ok\": %d, \"bytes\": %zu}\n", ok ? 1 : 0, n);
        return 0;
    }
    if (cmd == "gbps" && argc == 5) {
        double bytes = strtod(argv[2], nullptr), t = strtod(argv[3], nullptr);
        double throughput = bytes * 8. / t / 1e9;
        if (atoi(argv[4])) throughput *= 2;
        printf("%6.02f \n", throughput);
        return 0;
    }
    return 2;
}
