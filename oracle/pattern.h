/*
 * TEST INFRASTRUCTURE — not part of the product.  Only tests/, __graft_entry__.smoke() and
 * bench.py's baseline legs may use anything under oracle/.
 *
 * CPU restatement of what the transfer path must compute, used as the checker.
 *
 * The path itself is ncclSend/ncclRecv as called at p2p_matrix.cc:157-169 (and grouped at
 * :212-227): a byte-identity copy of `count * sizeof(dtype)` contiguous bytes from the
 * sender's buffer into the receiver's.  Its CPU statement is therefore one line
 * (oracle_expected_recv: what dst must hold == what src sent), and the work of an oracle is
 * in making wrong answers visible: the reference sends zeros (p2p_matrix.cc:129-130) and
 * never reads recv_buff, so a payload pattern is defined here — unique per (seed, rank,
 * iteration, byte offset) so that stale data, data from the wrong peer, a shifted copy and
 * a dropped tail all mismatch — together with an order-independent 64-bit digest.
 *
 * PARITY PINNING.  The reference has no tests, fixtures or golden vectors, and its
 * algorithm lives in libnccl (not under /root/reference; 2.27.3 in this image).  So:
 *   - the pattern/digest functions are pinned by tests/golden/pattern_vectors.json
 *     (written by oracle/make_golden.py from this C code; numpy and CUDA must agree);
 *   - the transfer itself is pinned against the reference's library by running
 *     oracle/_ref/p2p_matrix_nccl (the reference's loop + NCCL) on the GPU box and
 *     comparing its receive-buffer digests with the engine's (tests/test_gpu_multi.py).
 *   Without a GPU box run, transfer parity is "unpinned" in the sense of the build rules.
 */
#ifndef ORACLE_PATTERN_H
#define ORACLE_PATTERN_H

#include <stddef.h>
#include <stdint.h>

#ifdef __cplusplus
extern "C" {
#endif

/* splitmix64 finaliser */
uint64_t oracle_mix64(uint64_t z);
/* key(seed, rank, iter) = mix(seed + G*(rank+1)) ^ mix(iter + 0xD1B54A32D192ED03) */
uint64_t oracle_pattern_key(uint64_t seed, int rank, uint64_t iter);
/* byte i of the payload: little-endian byte (i & 7) of mix(key + G*((i>>3)+1)) */
uint8_t oracle_pattern_byte(uint64_t seed, int rank, uint64_t iter, uint64_t i);
/* fill buf[0..n) with the payload of (seed, rank, iter) */
void oracle_pattern_fill(uint8_t* buf, size_t n, uint64_t seed, int rank, uint64_t iter);
/* The transfer law: after send(src->dst) of n bytes, dst's receive buffer must equal src's
 * send buffer.  Writes that expected content (here: src's pattern) into `expect`. */
void oracle_expected_recv(uint8_t* expect, size_t n, uint64_t seed, int src_rank, uint64_t iter);
/* bytes of buf that differ from the payload; *first_bad = offset of the first (n if none) */
uint64_t oracle_pattern_mismatches(const uint8_t* buf, size_t n, uint64_t seed, int rank, uint64_t iter,
                                   uint64_t* first_bad);
/* digest(buf) = mix( sum_w mix(word_w ^ G*(w+1)) ^ mix(n) ), 8-byte LE words, tail zero-padded */
uint64_t oracle_digest(const uint8_t* buf, size_t n);
/* digest of the payload without materialising it */
uint64_t oracle_pattern_digest(size_t n, uint64_t seed, int rank, uint64_t iter);

/* The reference's host hash (getHostName + getHostHash, p2p_matrix.cc:44-61): DJB2a over the
 * host name up to the first '.'. */
uint64_t oracle_host_hash(const char* hostname);
/* check_process_placement_policy (p2p_matrix.cc:63-100) on given hashes: device index for
 * `rank`, -1 if hosts are uneven, -2 if a host's ranks are not consecutive. */
int oracle_placement(const uint64_t* hashes, int size, int rank);
/* Throughput figure the reference prints (p2p_matrix.cc:177, :258), in Gbps. */
double oracle_gbps(uint64_t msg_bytes, double seconds_per_message, int bidirectional);

#ifdef __cplusplus
}
#endif
#endif
