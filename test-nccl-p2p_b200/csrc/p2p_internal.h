// Shared between the engine's host code and its kernels.  Not installed; the public
// surface is include/p2p.h.
#pragma once
#include <cstddef>
#include <cstdint>
#include <cuda_runtime.h>
#include "p2p_device.cuh"

namespace p2p {

enum Flavour { FLAVOUR_LDST = 1, FLAVOUR_TMA = 2 };

// ---- verification pattern (restated independently on the CPU in oracle/pattern.h) ----
//   mix      = splitmix64 finaliser
//   key      = mix(seed + GAMMA*(rank+1)) ^ mix(iter + 0xD1B54A32D192ED03)
//   word(w)  = mix(key + GAMMA*(w+1))          (w = byte offset / 8)
//   byte(i)  = little-endian byte (i & 7) of word(i >> 3)
//   digest   = mix( (sum_w mix(word_w ^ GAMMA*(w+1))) ^ mix(nbytes) ), tail word zero-padded
constexpr uint64_t kGamma = 0x9E3779B97F4A7C15ull;

__host__ __device__ __forceinline__ uint64_t mix64(uint64_t z) {
    z = (z ^ (z >> 30)) * 0xBF58476D1CE4E5B9ull;
    z = (z ^ (z >> 27)) * 0x94D049BB133111EBull;
    return z ^ (z >> 31);
}
__host__ __device__ __forceinline__ uint64_t pattern_key(uint64_t seed, int rank, uint64_t iter) {
    return mix64(seed + kGamma * (uint64_t)(rank + 1)) ^ mix64(iter + 0xD1B54A32D192ED03ull);
}
__host__ __device__ __forceinline__ uint64_t pattern_word(uint64_t key, uint64_t w) {
    return mix64(key + kGamma * (w + 1));
}
__host__ __device__ __forceinline__ uint64_t digest_term(uint64_t word, uint64_t w) {
    return mix64(word ^ (kGamma * (w + 1)));
}
__host__ __device__ __forceinline__ uint64_t digest_finish(uint64_t sum, uint64_t nbytes) {
    return mix64(sum ^ mix64(nbytes));
}

struct CheckPartial {
    uint64_t mismatches;
    uint64_t first_bad;
    uint64_t digest;
};

// records the text p2p_get_last_error() returns (this thread) and hands `code` back
int set_last_error(int code, const char* fmt, ...);

cudaError_t configure_kernels(int max_tma_smem);
cudaError_t occupancy(int flavour, int dyn_smem, int* ctas_per_sm);
cudaError_t launch_group(int flavour, const GroupArgs& args, uint32_t grid, uint32_t block, int dyn_smem, cudaStream_t stream);
cudaError_t launch_ce_wait(uint64_t* notice_flag, uint64_t notice_word, const uint64_t* post_slot, uint64_t seq,
                           uint64_t expect_word, const GroupHdr& h, cudaStream_t stream);
cudaError_t launch_ce_signal(uint64_t* done_flag, uint64_t seq, cudaStream_t stream);
uint32_t pattern_grid(size_t nbytes, int sm_count);
cudaError_t launch_pattern_fill(void* buf, size_t nbytes, uint64_t key, int sm_count, cudaStream_t stream);
cudaError_t launch_pattern_check(const void* buf, size_t nbytes, uint64_t key, int check, int sm_count,
                                 CheckPartial* partials, unsigned int* ticket, CheckPartial* out,
                                 cudaStream_t stream);

}  // namespace p2p
