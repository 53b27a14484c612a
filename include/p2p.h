/*
 * p2p.h — C ABI of the B200 point-to-point transfer engine (libb200p2p.so).
 *
 * This is the drop-in boundary for the one path of AmadeusChan/test-nccl-p2p that the
 * engine replaces: the ncclSend/ncclRecv calls of the matrix benchmark.  Every entry
 * point below names the reference interface it stands in for (file:line into the
 * reference's p2p_matrix.cc, or into the NCCL header that file includes).  Plain
 * pointers, sizes and ints only; a stream is the CUDA driver's opaque `CUstream_st*`
 * (what `cudaStream_t` already is), so the library can be bound from C, C++, ctypes, cgo…
 *
 * Model.  One process per GPU, one process-global engine per process (the reference has
 * one `ncclComm_t` per process, p2p_matrix.cc:112,120).  All ranks live on one host and
 * reach each other over NVLink/NVSwitch.  Bootstrap is two-phase so any transport can
 * carry it (MPI_Allgather in include/p2p_mpi.h, torch.distributed in the Python binding):
 *
 *     p2p_init(rank, nranks, device);
 *     p2p_export(my_blob);      // p2p_blob_bytes() bytes
 *     <allgather the blobs in rank order>
 *     p2p_connect(all_blobs);
 *
 * Buffers.  The reference passes plain cudaMalloc memory and registers nothing (p2p_matrix.cc:127-128), and so may
 * a caller of this library: messages up to the eager limit travel through a ring in the engine's own memory, larger
 * ones through a staging ring there when the receive buffer is in no window (chunk by chunk, copied out by the
 * receiver: NCCL-like rates).  The fast path is one collective call per buffer, p2p_register_export /
 * p2p_register_import (same two-phase shape as bootstrap): payload then crosses NVLink exactly once, straight between
 * the two user buffers, through a peer mapping of the window.  Above the eager limit the sender's kernel stores into
 * the receiver's registered buffer (push); from the pull threshold up the receiver's kernel loads from the sender's
 * registered buffer instead when it is registered (pull, the faster direction on B200), else the sender pushes.
 * Registering both buffers gives the engine the choice.
 *
 * Matching rules are NCCL's (nccl.h:438-462): the k-th p2p_send(count, dtype, peer=B) on
 * rank A pairs with the k-th p2p_recv(count, dtype, peer=A) on rank B, byte counts must
 * agree, calls return after enqueue and complete in stream order, and a send and a recv
 * that must make progress together (bidirectional exchange, self-send) belong in one
 * p2p_group_start()/p2p_group_end() pair.
 *
 * Limits.  At most P2P_MAX_RANKS ranks, one rank per GPU, P2P_MAX_WINDOWS windows per rank,
 * P2P_MAX_GROUP_OPS operations per group, messages below 512 GiB.  Flag rings hold 64 rendezvous-sized
 * (4 eager-sized) messages per ordered pair of ranks: message k+64 (k+4) between two ranks must not be
 * issued before message k has completed — more than any one stream can have outstanding, since each
 * receive holds its stream until its payload has landed — and one message at a time per ordered pair may
 * go through the staging ring (a second one waits for the first).  One host thread at a time.  Calls are
 * not capturable into CUDA graphs (sequence numbers are assigned when a call is made).
 *
 * Launch overlap.  Kernels of large messages are launched with programmatic stream serialisation, and a
 * send whose payload the engine's own kernel ahead of it on the stream cannot be writing is announced to
 * the receiver before that kernel has retired ("early post").  The check covers engine launches only: a
 * kernel of the caller's that itself triggers programmatic launch completion early and then goes on to
 * write the buffer of the next p2p_send on that stream is outside it — set P2P_EARLY_POST=0 (or
 * P2P_PDL=1) in a program that does that.
 */
#ifndef B200_P2P_H
#define B200_P2P_H

#include <stddef.h>
#include <stdint.h>

#ifdef __cplusplus
extern "C" {
#endif

#if defined(__GNUC__)
#define P2P_API __attribute__((visibility("default")))
#else
#define P2P_API
#endif

/* Same opaque type cuda_runtime.h calls cudaStream_t. */
struct CUstream_st;
typedef struct CUstream_st* p2p_stream_t;

/* Result codes: numeric values follow ncclResult_t (nccl.h:41-49) so that code which
 * tests `!= ncclSuccess` keeps working.  The reference wraps calls in NCCLCHECK
 * (p2p_matrix.cc:35-42); P2PCHECK in include/p2p_mpi.h is the same macro. */
typedef enum {
    P2P_SUCCESS = 0,
    P2P_UNHANDLED_CUDA_ERROR = 1,
    P2P_SYSTEM_ERROR = 2,
    P2P_INTERNAL_ERROR = 3,
    P2P_INVALID_ARGUMENT = 4,
    P2P_INVALID_USAGE = 5,
    P2P_REMOTE_ERROR = 6, /* a peer did not answer within the timeout */
    P2P_IN_PROGRESS = 7,
    P2P_NUM_RESULTS = 8
} p2p_result_t;

/* Element types: numeric values are ncclDataType_t's (nccl.h:278-291), so the reference's
 * `ncclInt8` argument (p2p_matrix.cc:158) can be passed through unchanged.  The type only
 * scales `count` to bytes; payloads are opaque. */
typedef enum {
    P2P_INT8 = 0, P2P_CHAR = 0,
    P2P_UINT8 = 1,
    P2P_INT32 = 2, P2P_INT = 2,
    P2P_UINT32 = 3,
    P2P_INT64 = 4,
    P2P_UINT64 = 5,
    P2P_FLOAT16 = 6, P2P_HALF = 6,
    P2P_FLOAT32 = 7, P2P_FLOAT = 7,
    P2P_FLOAT64 = 8, P2P_DOUBLE = 8,
    P2P_BFLOAT16 = 9,
    P2P_FLOAT8E4M3 = 10,
    P2P_FLOAT8E5M2 = 11,
    P2P_NUM_TYPES = 12
} p2p_datatype_t;

#define P2P_MAX_RANKS 16      /* one NVSwitch domain on one host */
#define P2P_MAX_WINDOWS 254   /* registered windows per rank (ids 254 and 255 are protocol words) */
#define P2P_MAX_GROUP_OPS 32  /* sends + recvs in one group */

/* ---- lifecycle: replaces ncclGetUniqueId + MPI_Bcast + ncclCommInitRank
 *      (p2p_matrix.cc:115-120) and ncclCommDestroy (p2p_matrix.cc:270) ---- */

/* Create this process's engine on CUDA device `device` (the reference's
 * `rank % gpus_per_host`, p2p_matrix.cc:99,119).  Makes that device current. */
P2P_API int p2p_init(int rank, int nranks, int device);
/* Size of one rank's bootstrap blob (constant for a given library build). */
P2P_API size_t p2p_blob_bytes(void);
/* Fill `blob` with this rank's control-region handle and identity. */
P2P_API int p2p_export(void* blob);
/* `all_blobs` holds nranks blobs in rank order.  Maps every peer's control region and
 * enables peer access.  Fails with P2P_INVALID_USAGE if ranks span several hosts (NVLink
 * peer mappings do not cross hosts; the reference tolerates it, p2p_matrix.cc:78-99). */
P2P_API int p2p_connect(const void* all_blobs);
/* Teardown is two-phase like bootstrap, because CUDA IPC requires every importer to close its
 * mapping before the exporter frees the memory:
 *
 *     p2p_disconnect();         // drain this rank's device, close every mapping of peer memory
 *     <barrier over all ranks>
 *     p2p_finalize();           // free this rank's own regions
 *
 * p2p_finalize alone still works (it disconnects first) but then a peer may free memory this
 * rank has mapped, or the reverse; p2p_finalize_mpi in p2p_mpi.h and finalize() of the Python
 * binding run the three steps.  User buffers are never freed by the library. */
P2P_API int p2p_disconnect(void);
P2P_API int p2p_finalize(void);
P2P_API int p2p_rank(void);
P2P_API int p2p_nranks(void);

/* ---- windows: stand in for ncclCommWindowRegister / ncclCommRegister
 *      (nccl.h:243,251); the reference's buffers are plain cudaMalloc (p2p_matrix.cc:127-128),
 *      which is exactly what this accepts ---- */

/* Collective, two-phase like bootstrap.  `buf` may be NULL with bytes == 0 on ranks that
 * contribute nothing.  `buf` must come from cudaMalloc (any offset inside an allocation). */
P2P_API size_t p2p_register_blob_bytes(void);   /* 128 */
P2P_API int p2p_register_export(void* buf, size_t bytes, void* blob);
P2P_API int p2p_register_import(const void* all_blobs, int* window_id);
/* Collective by convention: every rank drops the same window. */
P2P_API int p2p_deregister(int window_id);
/* Where `peer`'s part of a window is mapped in THIS process (NULL if that rank contributed nothing): for
 * callers who run kernels or DMA copies of their own on peer memory.  No counterpart in the reference, which
 * never sees a peer address (NCCL keeps its mappings to itself). */
P2P_API int p2p_window_peer_ptr(int window_id, int peer, void** ptr);
/* Convenience for bindings that cannot call cudaMalloc themselves (ctypes, cgo). */
P2P_API int p2p_mem_alloc(void** ptr, size_t bytes);   /* nccl.h:130 ncclMemAlloc */
P2P_API int p2p_mem_free(void* ptr);                   /* nccl.h:140 ncclMemFree */

/* ---- the hot path: replaces ncclSend / ncclRecv (nccl.h:442,461; call sites
 *      p2p_matrix.cc:157-161, 165-169, 215-225, 236-246) ---- */

P2P_API int p2p_send(const void* buf, size_t count, int dtype, int peer, p2p_stream_t stream);
P2P_API int p2p_recv(void* buf, size_t count, int dtype, int peer, p2p_stream_t stream);
/* replace ncclGroupStart / ncclGroupEnd (nccl.h:493,503; p2p_matrix.cc:212,227,233,248).
 * All operations of a group run as ONE kernel launched on the stream of the group's first
 * operation; the other streams named in the group wait for it and it waits for them. */
P2P_API int p2p_group_start(void);
P2P_API int p2p_group_end(void);

/* ---- all-to-all on top of the all-pairs mode.  Block p of `sendbuf`
 *      (count elements) goes to rank p, block p of `recvbuf` arrives from rank p; the block for
 *      this rank itself is a local copy.  It is what NCCL users write as a grouped loop of
 *      ncclSend/ncclRecv over all peers (nccl.h:438-440, 455-457) — the reference's own
 *      bidirectional call shape (p2p_matrix.cc:212-228) widened from one peer to all — issued as
 *      ONE group, hence one fused kernel launch.  Same window rules as p2p_send/p2p_recv:
 *      register both buffers. ---- */
P2P_API int p2p_alltoall(const void* sendbuf, void* recvbuf, size_t count, int dtype, p2p_stream_t stream);

/* ---- symmetric windows on the driver's virtual memory API, and one-to-many over NVSwitch multicast
 *      (SURVEY section 8, rows f2 and f3).  p2p_register_* shares cudaMalloc memory through cudaIpc handles;
 *      this allocates instead: `bytes` of device memory on EVERY rank (cuMemCreate), exported as a POSIX
 *      file descriptor — handed to the peers' processes over a Unix-domain socket inside the library — or,
 *      with P2P_SYMM_HANDLE=fabric, as a fabric handle (CU_MEM_HANDLE_TYPE_FABRIC: plain bytes, the form that
 *      also crosses hosts within one NVLink domain).  Collective and two-phase like registration; the result is
 *      a window p2p_send/p2p_recv use like any other (stands in for ncclMemAlloc + ncclCommWindowRegister,
 *      nccl.h:130,251).  Where the fabric supports multicast the ranks' parts are also bound to one multicast
 *      object (cuMulticastCreate), which p2p_bcast stores into. ---- */
P2P_API size_t p2p_symm_blob_bytes(void);   /* 256 */
P2P_API int p2p_symm_alloc_export(size_t bytes, void* blob);
P2P_API int p2p_symm_alloc_import(const void* all_blobs, void** ptr, int* window_id);
P2P_API int p2p_symm_free(int window_id);   /* collective by convention, like p2p_deregister */
/* The window's multicast mapping (NULL when there is none): a store through it lands at the same offset of
 * every rank's part.  For callers who issue multimem instructions from kernels of their own. */
P2P_API int p2p_symm_multicast_ptr(int window_id, void** mc_ptr);
/* Broadcast, argument order of ncclBroadcast (nccl.h: sendbuff, recvbuff, count, datatype, root, comm, stream).
 * The reference has no one-to-many call: its loop shape reaches N-1 receivers with N-1 ncclSend calls
 * (p2p_matrix.cc:141-172), i.e. the root's link carries the payload N-1 times.  Here the root stores it ONCE,
 * with multimem.st into the multicast mapping, and the switch replicates it into every rank's `recvbuf`
 * (the root's own included); completion is published the same way.  Collective: every rank calls it with the
 * same count, dtype and root; `recvbuf` lies in a symmetric window at the same offset on every rank; `sendbuf`
 * is read on the root only (any device memory there; may equal recvbuf).  Stream-ordered like p2p_recv: a
 * receiver's stream holds until the payload has landed, the root's until it has been sent.  Without multicast
 * (P2P_NO_MULTICAST=1, or a fabric that has none) the root stores to each receiver in turn. ---- */
P2P_API int p2p_bcast(const void* sendbuf, void* recvbuf, size_t count, int dtype, int root, p2p_stream_t stream);

/* ---- errors: ncclGetErrorString / ncclGetLastError / ncclCommGetAsyncError
 *      (nccl.h:215,219,227) ---- */
P2P_API const char* p2p_get_error_string(int result);
P2P_API const char* p2p_get_last_error(void);
/* Non-blocking: P2P_SUCCESS, or the first error a device kernel reported (peer timeout). */
P2P_API int p2p_get_async_error(int* result);

/* ---- tuning knobs (the reference has none; NCCL reads NCCL_* environment variables).  Every option has an
 *      environment variable P2P_<NAME> read (and range-checked) in p2p_init; p2p_set_option changes it afterwards
 *      unless noted.  Protocol limits and ring shapes must be equal on all ranks (checked in p2p_connect) and
 *      are environment-only: eager_bytes (may shrink later), pull_min_bytes, stage_chunk, stage_lanes.
 *        path               0 auto, 1 push with ld/st, 2 push with TMA (P2P_PATH=auto|ldst|tma)
 *        nblocks, pull_nblocks, self_nblocks, eager_nblocks     CTAs per push / pull / local copy / eager message, 0 auto
 *        pull_total, push_total, fan_min     CTAs the flows of one group share, and the least each gets
 *        bidir_push         a lone large send pushes when its group also receives a large message: 0 never,
 *                           1 from that same peer (default), 2 from any peer, 3 also with several large sends
 *        pull_tma_min_bytes a rank that pulls while being pulled from uses the TMA ring from this size (-1 never)
 *        tma_chunk, tma_stages, tma_peer_chunk, tma_peer_stages, tma_min_bytes, self_tma_min_bytes     ring geometry, thresholds
 *        pdl                programmatic dependent launch: 0 never, 1 eager-only and local-copy groups, 2 every
 *                           group (trigger at kernel entry), 3 (default) as 1 plus groups of large messages
 *        early_post         see "Launch overlap" above (default 1)
 *        timeout_ms         device spin timeout (default 20000)
 *        loopback           send-to-self goes through the wire protocol (tests)
 *        ce_min_bytes, ce_wait_us, ce_flags     the experimental copy-engine path (off: measured slower)
 *      p2p_get_option also answers "sm_count", "ctas_per_sm_ldst", "ctas_per_sm_tma", "stat_launches" (kernels
 *      launched so far), "stat_ce_copies", and what the last launched group looked like: "stat_last_flavour"
 *      (1 ld/st, 2 TMA), "stat_last_grid", "stat_last_block", "stat_last_push_ops", "stat_last_push_ctas",
 *      "stat_last_offer_ops" (sends offered for pulling), "stat_last_recv_ops", "stat_last_recv_ctas",
 *      "stat_last_eager_ops", "stat_last_copy_ctas", "stat_last_ce_ops", "stat_last_pdl". ---- */
P2P_API int p2p_set_option(const char* name, long long value);
P2P_API int p2p_get_option(const char* name, long long* value);
/* What the selector would do for a message of `bytes` to `peer`: writes the path id
 * (1 push ld/st, 2 push TMA, 4 eager, 5 pull when the send buffer is registered) and CTA count. */
P2P_API int p2p_query_plan(size_t bytes, int peer, int* path, int* nblocks);

/* ---- verification support: the reference sends zeros and never reads recv_buff
 *      (p2p_matrix.cc:129-130); parity needs a pattern.  With mix = the splitmix64 finaliser and
 *      G = 0x9E3779B97F4A7C15:  key = mix(seed + G*(rank+1)) ^ mix(iter + 0xD1B54A32D192ED03),
 *      word(w) = mix(key + G*(w+1)), and byte i of the buffer is little-endian byte (i & 7) of
 *      word(i >> 3).  oracle/pattern.h is the CPU restatement these kernels are tested against. ---- */
P2P_API int p2p_pattern_fill(void* buf, size_t bytes, uint64_t seed, int rank, uint64_t iter,
                             p2p_stream_t stream);
/* Counts bytes of `buf` that differ from the pattern; synchronises `stream`.  Also returns
 * the offset of the first mismatch (or bytes if none) and a 64-bit digest of the buffer. */
P2P_API int p2p_pattern_check(const void* buf, size_t bytes, uint64_t seed, int rank, uint64_t iter,
                              p2p_stream_t stream, uint64_t* mismatches, uint64_t* first_bad,
                              uint64_t* digest);
/* 64-bit order-dependent digest of any device buffer (same function as above). */
P2P_API int p2p_digest(const void* buf, size_t bytes, p2p_stream_t stream, uint64_t* digest);

/* ---- no counterpart in the reference: p2p_digest for callers that may not allocate.
 *      Who that is: a process whose framework already owns all of HBM (a cudaMalloc from a
 *      library would fail or fragment its pool), a loop that cannot afford the device-wide
 *      synchronisation cudaMalloc/cudaFree imply, or a checker that runs without p2p_init.
 *      The scratch the digest kernel needs (per-CTA partial sums, an arrival ticket, the
 *      result) lives in a fixed array inside the library's device image, so this call — unlike
 *      every other entry point — needs no p2p_init and performs no allocation of any kind.
 *      Limits and rules:
 *        - bytes <= P2P_NOALLOC_MAX_BYTES (the array holds 4096 partial sums, one per 256 KiB
 *          of input); larger buffers are refused with P2P_INVALID_ARGUMENT;
 *        - `digest_dev` is DEVICE memory (8 bytes) the caller owns; the call is asynchronous
 *          on `stream`, the value is there once the stream has caught up;
 *        - the first call on each device looks up that device's copy of the array; make that
 *          first call with the device current and outside any stream capture;
 *        - all calls on a device share the one array: two calls whose kernels could overlap
 *          (different streams, no dependency) race and give garbage — serialise them. ---- */
#define P2P_NOALLOC_MAX_BYTES (1ull << 30)
P2P_API int p2p_digest_noalloc(const void* buf, size_t bytes, p2p_stream_t stream, uint64_t* digest_dev);

#ifdef __cplusplus
}
#endif
#endif /* B200_P2P_H */
