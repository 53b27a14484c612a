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

// ---- what p2p_symm.cu (VMM windows, multicast broadcast) needs from the engine in p2p_engine.cu ----
struct EngineInfo {
    int rank, nranks, device, sm_count;
    uint32_t* async_err;   // host-mapped word the kernels report into
    uint64_t timeout_ns;
};
int engine_info(EngineInfo* out);  // P2P_INVALID_USAGE unless p2p_init + p2p_connect have run
// A window whose peer mappings the caller made itself (VMM handles instead of cudaIpc); usable by p2p_send/p2p_recv
// like any other.  p2p_deregister / p2p_finalize forget it without unmapping anything.
int engine_add_window(uint8_t* base, size_t bytes, uint8_t* const* peer_bases, int* window_id);
void symm_finalize(int nranks);    // unmaps and releases every symmetric window (called by p2p_finalize)

// ---- passing POSIX file descriptors between the ranks' processes (p2p_fdpass.cu) ----
// A CUDA VMM handle exported as a file descriptor means nothing in another process until the descriptor itself
// has been sent there over a Unix-domain socket (SCM_RIGHTS).  The exporting rank runs a short-lived server thread:
//   fdpass_serve   starts it for `nfds` descriptors and `requests` expected requests; `name_out` (at least 40 bytes)
//                  receives the abstract socket name to publish in the rank's blob
//   fdpass_fetch   connects to a peer's server and returns a duplicate of its descriptor `which` (or -1)
//   fdpass_finish  waits for the server thread (all expected requests served, or the timeout) and closes the socket
struct FdServer;
FdServer* fdpass_serve(const int* fds, int nfds, int requests, int timeout_ms, char* name_out);
int fdpass_fetch(const char* name, int which, int timeout_ms);
void fdpass_finish(FdServer* server);

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
