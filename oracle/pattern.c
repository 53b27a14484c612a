/* TEST INFRASTRUCTURE — see pattern.h.  Plain C, one byte at a time where that is the
 * clearest statement; nothing here is tuned. */
#include "pattern.h"

#define ORACLE_GAMMA 0x9E3779B97F4A7C15ull

uint64_t oracle_mix64(uint64_t z) {
    z = (z ^ (z >> 30)) * 0xBF58476D1CE4E5B9ull;
    z = (z ^ (z >> 27)) * 0x94D049BB133111EBull;
    return z ^ (z >> 31);
}

uint64_t oracle_pattern_key(uint64_t seed, int rank, uint64_t iter) {
    return oracle_mix64(seed + ORACLE_GAMMA * (uint64_t)(rank + 1)) ^ oracle_mix64(iter + 0xD1B54A32D192ED03ull);
}

static uint64_t pattern_word(uint64_t key, uint64_t w) { return oracle_mix64(key + ORACLE_GAMMA * (w + 1)); }

uint8_t oracle_pattern_byte(uint64_t seed, int rank, uint64_t iter, uint64_t i) {
    const uint64_t word = pattern_word(oracle_pattern_key(seed, rank, iter), i >> 3);
    return (uint8_t)(word >> (8 * (i & 7)));
}

void oracle_pattern_fill(uint8_t* buf, size_t n, uint64_t seed, int rank, uint64_t iter) {
    const uint64_t key = oracle_pattern_key(seed, rank, iter);
    for (size_t i = 0; i < n; ++i) buf[i] = (uint8_t)(pattern_word(key, i >> 3) >> (8 * (i & 7)));
}

/* ncclSend/ncclRecv (p2p_matrix.cc:157-169): identity copy, so the expectation at dst is
 * simply what src put in its send buffer. */
void oracle_expected_recv(uint8_t* expect, size_t n, uint64_t seed, int src_rank, uint64_t iter) {
    oracle_pattern_fill(expect, n, seed, src_rank, iter);
}

uint64_t oracle_pattern_mismatches(const uint8_t* buf, size_t n, uint64_t seed, int rank, uint64_t iter,
                                   uint64_t* first_bad) {
    const uint64_t key = oracle_pattern_key(seed, rank, iter);
    uint64_t bad = 0, first = n;
    for (size_t i = 0; i < n; ++i) {
        const uint8_t want = (uint8_t)(pattern_word(key, i >> 3) >> (8 * (i & 7)));
        if (buf[i] != want) {
            if (!bad) first = i;
            ++bad;
        }
    }
    if (first_bad) *first_bad = first;
    return bad;
}

uint64_t oracle_digest(const uint8_t* buf, size_t n) {
    uint64_t sum = 0;
    const size_t nwords = (n + 7) / 8;
    for (size_t w = 0; w < nwords; ++w) {
        uint64_t word = 0;
        for (size_t b = 0; b < 8 && w * 8 + b < n; ++b) word |= (uint64_t)buf[w * 8 + b] << (8 * b);
        sum += oracle_mix64(word ^ (ORACLE_GAMMA * (w + 1)));
    }
    return oracle_mix64(sum ^ oracle_mix64((uint64_t)n));
}

uint64_t oracle_pattern_digest(size_t n, uint64_t seed, int rank, uint64_t iter) {
    const uint64_t key = oracle_pattern_key(seed, rank, iter);
    uint64_t sum = 0;
    const size_t nwords = (n + 7) / 8;
    for (size_t w = 0; w < nwords; ++w) {
        uint64_t word = pattern_word(key, w);
        const size_t left = n - w * 8;
        if (left < 8) word &= (1ull << (8 * left)) - 1;
        sum += oracle_mix64(word ^ (ORACLE_GAMMA * (w + 1)));
    }
    return oracle_mix64(sum ^ oracle_mix64((uint64_t)n));
}

uint64_t oracle_host_hash(const char* hostname) {
    /* p2p_matrix.cc:53-61 cuts the name at the first '.', :44-51 hashes what is left */
    uint64_t result = 5381;
    for (int c = 0; hostname[c] != '\0' && hostname[c] != '.'; c++)
        result = ((result << 5) + result) ^ (uint64_t)hostname[c];
    return result;
}

int oracle_placement(const uint64_t* hashes, int size, int rank) {
    /* count distinct hosts (the reference uses a std::set, p2p_matrix.cc:78-82) */
    int num_hosts = 0;
    for (int i = 0; i < size; ++i) {
        int seen = 0;
        for (int j = 0; j < i; ++j)
            if (hashes[j] == hashes[i]) seen = 1;
        if (!seen) ++num_hosts;
    }
    if (size % num_hosts != 0) return -1; /* :83-86 */
    const int per_host = size / num_hosts;
    for (int host = 0; host < num_hosts; ++host) /* :88-98 */
        for (int g = 1; g < per_host; ++g)
            if (hashes[host * per_host + g] != hashes[host * per_host + g - 1]) return -2;
    return rank % per_host; /* :99 */
}

double oracle_gbps(uint64_t msg_bytes, double seconds_per_message, int bidirectional) {
    double throughput = (double)msg_bytes * 8. / seconds_per_message / 1e9; /* :177 */
    if (bidirectional) throughput *= 2;                                       /* :258 */
    return throughput;
}
