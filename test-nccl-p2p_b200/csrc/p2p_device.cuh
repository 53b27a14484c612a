// Device side of the transfer engine: hand-written sm_100a data movement + signalling.
//
// One launch (`p2p_group_kernel_*`) executes every operation of a group, the way one
// ncclDevKernel_SendRecv launch serves a ncclGroupStart/End pair in the reference's
// bidirectional loop (p2p_matrix.cc:212-228).  CTAs are assigned to operations by block
// range; the first CTA is the control CTA that posts and completes rendezvous receives.
//
// Protocols
//   push        receiver posts {seq, window, offset} into the SENDER's control region;
//               sender CTAs store payload straight into the receiver's buffer through the
//               peer mapping, then the last CTA to finish publishes `done[seq]` in the
//               RECEIVER's control region with a system-scope release store.
//   pull        (large messages) the mirror image: sender posts where its payload lies,
//               receiver CTAs LOAD it through the peer mapping (NVLink reads run ~8% faster
//               than SM-issued writes on B200) and publish `pull_done[seq]` at the sender.
//               The sender decides; a receiver of a large message adapts to what it is told.
//   eager (LL)  8 payload bytes + two 4-byte flags per 16-byte store into a ring in the
//               receiver's control region; the receiver copies out as flags turn valid.
//   copy        send and receive of one rank to itself, matched on the host: a plain copy.
#pragma once
#include <cstdint>
#include <cuda_runtime.h>

namespace p2p {

constexpr int kRing = 64;            // outstanding rendezvous messages per ordered pair
constexpr int kLLSlots = 4;          // eager ring depth per ordered pair
constexpr int kLdstThreads = 512;
constexpr int kTmaThreads = 128;
constexpr int kMaxGroupOps = 32;
constexpr int kTmaMaxStages = 16;

enum OpKind : uint32_t {
    OP_NONE = 0,
    OP_POST = 1,        // one thread of the control CTA: publish `word`, wait for `f_wait == seq`
    OP_PUSH = 2,        // mover CTAs: wait for the receiver's post, store local -> peer, signal done
    OP_SEND_LL = 3,
    OP_RECV_LL = 4,
    OP_COPY = 5,        // mover CTAs: local -> local, no flags
    OP_RECV_ADAPT = 6,  // mover CTAs: wait for the sender's word; pull peer -> local, or fall back to posting
    OP_RECV_STAGED = 7, // mover CTAs: this rank's buffer is in no window: ask the sender to go through the staging ring
};

enum AsyncError : uint32_t { AERR_NONE = 0, AERR_TIMEOUT_POST = 1, AERR_TIMEOUT_DONE = 2,
                             AERR_TIMEOUT_LL_DATA = 3, AERR_TIMEOUT_LL_ACK = 4, AERR_BAD_WINDOW = 5,
                             AERR_TIMEOUT_STAGE = 6, AERR_STAGE_BUSY = 7 };

struct alignas(8) OpDesc {
    uint64_t local;      // this GPU's buffer: source of a send, destination of a receive
    uint64_t nbytes;
    uint64_t seq;        // sequence number of this message on its (src, dst, protocol) lane
    uint64_t f_wait;     // local flag polled      (post: done slot; push: post slot; LL send: ack slot)
    uint64_t f_signal;   // peer-mapped flag written (post: peer's post slot; push: done slot; LL recv: ack slot)
    uint64_t word;       // post: the word to publish; adaptive receive: its own push post word (0: buffer not registered)
    uint64_t f_wait2;    // adaptive receive: local slot the sender's word arrives in
    uint64_t f_signal2;  // adaptive receive: sender's pull_done slot; push with notice: receiver's pull_post slot
    uint64_t word2;      // push with notice: the "I will push" word
    uint64_t aux;        // window table of the peer / LL ring slot / destination of a local copy
    uint64_t arrive;     // local u32 arrival counter of the mover CTAs
    uint32_t kind;
    uint32_t first_block;
    uint32_t nblocks;
    uint32_t peer;
    uint32_t flags;      // OPF_*
    uint32_t pad;
    uint64_t stage;      // StageInfo of `peer` in local memory (roles that may go through the staging ring)
};

// Staging ring between one ordered pair of ranks, for receive buffers that are in no window (plain cudaMalloc
// memory nobody registered, as in the reference, p2p_matrix.cc:127-128).  The message is cut into chunks; chunk c
// belongs to lane c % lanes, and each lane is a private ring of `slots` chunk-sized slots in the RECEIVER's control
// region with its own pair of counters: the sender stores the chunk into the lane's next slot and publishes
// `filled[lane]`, the receiver copies it out into its buffer and publishes `drained[lane]` at the sender.  A lane
// is served by one CTA on each side, so no chunk needs a barrier between CTAs: the lanes stream independently.
// Counters are (seq << 24 | chunks of this lane), so they never need resetting.  One such message per ordered pair
// at a time (`lock`, taken by the receiver before it asks).
constexpr int kStageLanes = 32;
constexpr int kStageSlots = 4;
struct StageHeader {                  // first 4 KiB of the ring area [region of rank r][from rank q]
    uint64_t filled[kStageLanes];     // q -> r traffic: chunks q has stored into each lane   (written by q)
    uint64_t drained[kStageLanes];    // r -> q traffic: chunks q has copied out of each lane (written by q)
};
constexpr size_t kStageHeaderBytes = 4096;
struct StageInfo {
    uint64_t send_ring;     // peer-mapped: lane 0, slot 0 of the ring in the peer's region that my sends to it go through
    uint64_t send_filled;   // peer-mapped u64[lanes] I publish into
    uint64_t send_drained;  // local u64[lanes] the peer publishes into
    uint64_t send_arrive;   // local u32: CTAs of the send that are through with all their chunks
    uint64_t recv_ring;     // local
    uint64_t recv_filled;   // local u64[lanes]
    uint64_t recv_drained;  // peer-mapped u64[lanes]
    uint64_t recv_arrive;   // local u32
    uint64_t lock;          // local u32
    uint32_t chunk, slots, lanes, pad;
};
__host__ __device__ __forceinline__ uint64_t stage_count(uint64_t seq, uint64_t chunks) { return (seq << 24) | chunks; }

enum OpFlags : uint32_t {
    OPF_EARLY_POST = 1,  // OP_POST of a send: the word may be published before griddepcontrol.wait when the kernel
                         // that is still running ahead of this one is the engine's own (token check, see pdl_*)
};

constexpr uint32_t kPdlLaunched = 1;     // launched with programmatic stream serialisation: griddepcontrol.wait applies
constexpr uint32_t kPdlLateTrigger = 2;  // mover CTAs let the next kernel in only once their copy loop is done
constexpr int kTokenSlots = 64;

struct GroupHdr {
    uint32_t nops;
    uint32_t n_ctrl;         // OP_POST ops served by the control CTA (block 0), or 0: no control CTA
    uint32_t my_token;       // identifies this launch in tokens[] (0: none)
    uint32_t tma_chunk;      // bytes per cp.async.bulk (TMA flavour)
    uint32_t tma_stages;
    uint32_t pdl;            // kPdl* bits
    uint64_t timeout_ns;
    uint32_t* async_err;     // host-mapped word; first error wins
    uint32_t* tokens;        // local u32[kTokenSlots]: which engine launches have passed their griddepcontrol.wait
    uint32_t prev_token;     // the engine launch enqueued right before this one on the same stream (0: unknown)
    uint32_t pad;
};

template <int MAXOPS>
struct GroupArgsT {
    GroupHdr h;
    OpDesc ops[MAXOPS];
};
using GroupArgs = GroupArgsT<kMaxGroupOps>;
constexpr int kSmallGroupOps = 4;
using GroupArgsSmall = GroupArgsT<kSmallGroupOps>;

// ---------------------------------------------------------------- PTX helpers

__device__ __forceinline__ void st_release_sys(uint64_t* p, uint64_t v) {
    asm volatile("st.release.sys.global.u64 [%0], %1;" ::"l"(p), "l"(v) : "memory");
}
__device__ __forceinline__ uint64_t ld_acquire_sys(const uint64_t* p) {
    uint64_t v;
    asm volatile("ld.acquire.sys.global.u64 %0, [%1];" : "=l"(v) : "l"(p) : "memory");
    return v;
}
__device__ __forceinline__ void st_relaxed_sys(uint64_t* p, uint64_t v) {
    asm volatile("st.relaxed.sys.global.u64 [%0], %1;" ::"l"(p), "l"(v) : "memory");
}
__device__ __forceinline__ void fence_sys() { asm volatile("fence.acq_rel.sys;" ::: "memory"); }
__device__ __forceinline__ uint64_t globaltimer_ns() {
    uint64_t t;
    asm volatile("mov.u64 %0, %%globaltimer;" : "=l"(t));
    return t;
}

// streaming 16-byte accesses: no L1 allocation, the payload is touched once.  (32-byte
// LDG/STG.256 were measured too: same NVLink and HBM rates, twice the registers.)
__device__ __forceinline__ uint4 ld_stream16(const void* p) {
    uint4 v;
    asm volatile("ld.global.L1::no_allocate.v4.b32 {%0,%1,%2,%3}, [%4];"
                 : "=r"(v.x), "=r"(v.y), "=r"(v.z), "=r"(v.w) : "l"(p) : "memory");
    return v;
}
__device__ __forceinline__ void st_stream16(void* p, const uint4& v) {
    asm volatile("st.global.L1::no_allocate.v4.b32 [%0], {%1,%2,%3,%4};"
                 ::"l"(p), "r"(v.x), "r"(v.y), "r"(v.z), "r"(v.w) : "memory");
}
// LL line: {data_lo, flag, data_hi, flag}; each 8-byte half is written atomically
__device__ __forceinline__ void st_ll_line(void* p, uint64_t data, uint32_t flag) {
    asm volatile("st.volatile.global.v4.u32 [%0], {%1,%2,%3,%4};"
                 ::"l"(p), "r"((uint32_t)data), "r"(flag), "r"((uint32_t)(data >> 32)), "r"(flag)
                 : "memory");
}
__device__ __forceinline__ uint4 ld_ll_line(const void* p) {
    uint4 v;
    asm volatile("ld.volatile.global.v4.u32 {%0,%1,%2,%3}, [%4];"
                 : "=r"(v.x), "=r"(v.y), "=r"(v.z), "=r"(v.w) : "l"(p) : "memory");
    return v;
}

__device__ __forceinline__ void report_error(const GroupHdr& h, uint32_t code) {
    if (h.async_err) atomicCAS_system(h.async_err, 0u, code);
}

// Programmatic dependent launch.  pdl_wait holds a kernel until everything before it in the stream has
// completed; nothing that depends on earlier stream work (user buffers) is touched before it — what a role
// may do earlier is poll its own flags and, for a send whose source the kernel ahead cannot be writing, publish
// where the payload lies (OPF_EARLY_POST).  pdl_trigger lets the next kernel of the stream become resident:
// the control CTA and the flag-only roles do it at once, mover CTAs after their copy loop (kPdlLateTrigger),
// so that the next kernel's CTAs arrive as this one's retire instead of crowding the SMs while it runs.
__device__ __forceinline__ void publish_token(const GroupHdr& h);
__device__ __forceinline__ void pdl_wait(const GroupHdr& h) {
    if (h.pdl & kPdlLaunched) asm volatile("griddepcontrol.wait;" ::: "memory");
    publish_token(h);
}
__device__ __forceinline__ void pdl_trigger(const GroupHdr& h) {
    if (h.pdl & kPdlLaunched) asm volatile("griddepcontrol.launch_dependents;" ::: "memory");
}
__device__ __forceinline__ void pdl_trigger_early(const GroupHdr& h) {
    if ((h.pdl & (kPdlLaunched | kPdlLateTrigger)) == kPdlLaunched) asm volatile("griddepcontrol.launch_dependents;" ::: "memory");
}
__device__ __forceinline__ void pdl_trigger_late(const GroupHdr& h) {
    if ((h.pdl & (kPdlLaunched | kPdlLateTrigger)) == (kPdlLaunched | kPdlLateTrigger)) asm volatile("griddepcontrol.launch_dependents;" ::: "memory");
}
__device__ __forceinline__ uint32_t ld_volatile_u32(const uint32_t* p) {
    uint32_t v;
    asm volatile("ld.volatile.global.u32 %0, [%1];" : "=r"(v) : "l"(p) : "memory");
    return v;
}
// Block 0 announces, once past its wait, that launch `my_token` is the one running on its stream
// (called from pdl_wait: the store never precedes the wait).
__device__ __forceinline__ void publish_token(const GroupHdr& h) {
    if (h.my_token && h.tokens && blockIdx.x == 0 && threadIdx.x == 0) {
        asm volatile("st.volatile.global.u32 [%0], %1;" ::"l"(h.tokens + (h.my_token % kTokenSlots)), "r"(h.my_token) : "memory");
        __threadfence();
    }
}
// Is the engine launch enqueued right before this one on this stream past its wait (and this kernel therefore
// running ahead of its completion, with everything older in the stream complete)?
__device__ __forceinline__ bool predecessor_is_ours(const GroupHdr& h) {
    return (h.pdl & kPdlLaunched) && h.prev_token && h.tokens &&
           ld_volatile_u32(h.tokens + (h.prev_token % kTokenSlots)) == h.prev_token;
}

// Poll helper.  `pred(value)` decides; returns false on timeout.
template <typename Pred>
__device__ __forceinline__ bool spin_flag(const uint64_t* flag, uint64_t timeout_ns, uint64_t& out,
                                          Pred pred) {
    uint64_t v = ld_acquire_sys(flag);
    if (pred(v)) { out = v; return true; }
    uint64_t t0 = globaltimer_ns();
    uint32_t polls = 0;
    for (;;) {
        v = ld_acquire_sys(flag);
        if (pred(v)) { out = v; return true; }
        if (++polls > 64) {
            __nanosleep(40);
            if ((polls & 255u) == 0 && globaltimer_ns() - t0 > timeout_ns) return false;
        }
    }
}

// ---------------------------------------------------------------- copy loops

// Vectors [0, nvec) of width sizeof(V); this CTA is `part` of `nparts`.  Tiles of
// blockDim.x*UNROLL vectors are dealt round-robin so that concurrently running CTAs
// stream neighbouring addresses (DRAM pages and NVLink packets stay dense).  Inside a tile
// each warp owns 32*UNROLL consecutive vectors and walks them in UNROLL coalesced steps of
// 32 vectors, so the per-step offsets are immediates and one address register pair serves
// all loads and one all stores.
// `wait_hdr` (may be null): a pull may LOAD its first tile from the sender's buffer — which the sender has declared
// ready — while the kernel ahead of this one on the stream is still retiring; only its STORES, into this rank's
// buffer, have to wait for that kernel (griddepcontrol.wait).  The first tile's loads are therefore issued before
// the wait: NVLink's ramp-up and the launch gap disappear behind the predecessor's tail.
template <typename V, int UNROLL, typename LD, typename ST>
__device__ __forceinline__ void copy_vectors(V* __restrict__ dst, const V* __restrict__ src,
                                             size_t nvec, uint32_t part, uint32_t nparts, LD ld, ST st,
                                             const GroupHdr* wait_hdr = nullptr) {
    const size_t tile = (size_t)blockDim.x * UNROLL;
    const size_t ntiles = nvec / tile;
    const size_t in_tile = (size_t)(threadIdx.x >> 5) * (32 * UNROLL) + (threadIdx.x & 31);
    V v[UNROLL];
    size_t t = part;
    if (t < ntiles) {
        const V* s = src + t * tile + in_tile;
#pragma unroll
        for (int u = 0; u < UNROLL; ++u) v[u] = ld(s + u * 32);
    }
    if (wait_hdr) pdl_wait(*wait_hdr);
    while (t < ntiles) {
        V* d = dst + t * tile + in_tile;
#pragma unroll
        for (int u = 0; u < UNROLL; ++u) st(d + u * 32, v[u]);
        t += nparts;
        if (t < ntiles) {
            const V* s = src + t * tile + in_tile;
#pragma unroll
            for (int u = 0; u < UNROLL; ++u) v[u] = ld(s + u * 32);
        }
    }
    if (part == (uint32_t)(ntiles % nparts)) {
        for (size_t i = ntiles * tile + threadIdx.x; i < nvec; i += blockDim.x) st(dst + i, ld(src + i));
    }
}

template <typename T>
__device__ __forceinline__ void copy_elems(T* __restrict__ dst, const T* __restrict__ src, size_t n,
                                           uint32_t part, uint32_t nparts) {
    const size_t stride = (size_t)nparts * blockDim.x;
    size_t i = (size_t)part * blockDim.x + threadIdx.x;
    // ld.cv: a staging-ring slot is read again and again within one kernel, so nothing of it may linger in L1
#pragma unroll 4
    for (; i < n; i += stride) dst[i] = __ldcv(src + i);
}

struct Ld16 { __device__ __forceinline__ uint4 operator()(const uint4* p) const { return ld_stream16(p); } };
struct St16 { __device__ __forceinline__ void operator()(uint4* p, const uint4& v) const { st_stream16(p, v); } };

// Byte-exact copy of n bytes for arbitrary alignment of either side.  Each kernel
// inlines it exactly once (see resolve_role), which keeps the hot loop free of spills.
// With `wait_hdr` nothing is stored before pdl_wait (see copy_vectors).
__device__ __forceinline__ void copy_bytes(uint8_t* dst, const uint8_t* src, size_t n, uint32_t part,
                                           uint32_t nparts, const GroupHdr* wait_hdr = nullptr) {
    if (n == 0) {
        if (wait_hdr) pdl_wait(*wait_hdr);
        return;
    }
    const uint64_t rel = ((uint64_t)dst ^ (uint64_t)src);
    if ((rel & 15) == 0) {
        // same phase: 16-byte body first (its first loads may run ahead of the wait), then the byte head up to a
        // 16-byte boundary and the byte tail
        size_t head = (16 - ((uint64_t)dst & 15)) & 15;
        if (head > n) head = n;
        uint8_t* d = dst + head;
        const uint8_t* s = src + head;
        const size_t left = n - head;
        const size_t body = left & ~(size_t)15;
        copy_vectors<uint4, 8>((uint4*)d, (const uint4*)s, body / 16, part, nparts, Ld16{}, St16{}, wait_hdr);
        if (part == 0)
            for (size_t i = threadIdx.x; i < head; i += blockDim.x) dst[i] = __ldcv(src + i);
        if (part == nparts - 1)
            for (size_t i = body + threadIdx.x; i < left; i += blockDim.x) d[i] = __ldcv(s + i);
        return;
    }
    if (wait_hdr) pdl_wait(*wait_hdr);
    // different phase: widest element both sides can be aligned to at once
    const uint32_t w = (rel & 7) == 0 ? 8 : (rel & 3) == 0 ? 4 : (rel & 1) == 0 ? 2 : 1;
    size_t head = (w - ((uint64_t)dst & (w - 1))) & (w - 1);
    if (head > n) head = n;
    if (part == 0)
        for (size_t i = threadIdx.x; i < head; i += blockDim.x) dst[i] = __ldcv(src + i);
    uint8_t* d = dst + head;
    const uint8_t* s = src + head;
    const size_t left = n - head;
    const size_t ne = left / w;
    switch (w) {
        case 8: copy_elems((uint64_t*)d, (const uint64_t*)s, ne, part, nparts); break;
        case 4: copy_elems((uint32_t*)d, (const uint32_t*)s, ne, part, nparts); break;
        case 2: copy_elems((uint16_t*)d, (const uint16_t*)s, ne, part, nparts); break;
        default: copy_elems(d, s, ne, part, nparts); break;
    }
    if (part == nparts - 1)
        for (size_t i = ne * w + threadIdx.x; i < left; i += blockDim.x) d[i] = __ldcv(s + i);
}

// ---------------------------------------------------------------- rendezvous roles

// Post word layout: [63:48] low 16 bits of seq, [47:40] window id, [39] valid (always 1, so
// a real post is never the all-zero word of a fresh slot), [38:0] byte offset (512 GiB).
// Window id 0xFF in a SENDER's word means "I will push, post me your buffer".
constexpr uint64_t kPostValid = 1ull << 39;
constexpr uint64_t kPostOffMask = kPostValid - 1;
constexpr uint32_t kWinPushPlease = 0xFFu;
constexpr uint32_t kWinStaged = 0xFEu;  // in a RECEIVER's post: "my buffer is in no window, send through the staging ring"
__host__ __device__ __forceinline__ uint64_t make_post_word(uint64_t seq, uint32_t win, uint64_t off) {
    return ((seq & 0xffffull) << 48) | ((uint64_t)(win & 0xffu) << 40) | kPostValid | (off & kPostOffMask);
}

// One control thread: publish a post word, then wait for the matching completion flag.
__device__ __forceinline__ void post_and_wait(const GroupHdr& h, const OpDesc& op, bool already_posted = false) {
    if (!already_posted) st_release_sys((uint64_t*)op.f_signal, op.word);
    uint64_t v;
    const uint64_t want = op.seq;
    if (!spin_flag((const uint64_t*)op.f_wait, h.timeout_ns, v, [want](uint64_t x) { return x == want; }))
        report_error(h, AERR_TIMEOUT_DONE);
}

// Thread 0 waits for the word tagged with this message's sequence number in `slot`.
__device__ __forceinline__ bool wait_word(const GroupHdr& h, const uint64_t* slot, uint64_t seq, uint64_t& word) {
    const uint64_t tag = seq & 0xffffull;
    return spin_flag(slot, h.timeout_ns, word, [tag](uint64_t x) { return (x >> 48) == tag && (x & kPostValid); });
}

constexpr uint64_t kStagedMarker = 1;  // "destination" of a push whose receiver asked for the staging ring
// All CTAs of a push learn the destination; returns nullptr on timeout / bad window.
__device__ __forceinline__ uint8_t* push_wait_post(const GroupHdr& h, const OpDesc& op, uint32_t part, uint64_t* smem_slot) {
    if (threadIdx.x == 0) {
        if (part == 0 && op.f_signal2) st_release_sys((uint64_t*)op.f_signal2, op.word2);  // "I will push"
        uint64_t v = 0, dst = 0;
        if (wait_word(h, (const uint64_t*)op.f_wait, op.seq, v)) {
            const uint32_t win = (uint32_t)(v >> 40) & 0xffu;
            const uint64_t base = win == kWinStaged ? 0 : ((const uint64_t*)op.aux)[win];
            if (win == kWinStaged) dst = kStagedMarker;
            else if (base) dst = base + (v & kPostOffMask);
            else report_error(h, AERR_BAD_WINDOW);
        } else {
            report_error(h, AERR_TIMEOUT_POST);
        }
        *smem_slot = dst;
    }
    __syncthreads();
    pdl_wait(h);  // the handshake above touched flags only; the payload is read from here on
    return (uint8_t*)(*smem_slot);
}

// After this CTA's accesses: the last CTA of the op to get here publishes completion.  Before that it
// zeroes the local slot the op's tagged word arrived in (`consumed`): every CTA of the op has read it by then,
// and a word left behind could be taken for that of the message 65536 sequence numbers later, which carries
// the same 16-bit tag (the slot's other users may all have been of the other protocol and never rewritten it).
__device__ __forceinline__ void movers_signal(const GroupHdr& h, const OpDesc& op, uint64_t flag, uint64_t consumed) {
    __syncthreads();
    pdl_trigger_late(h);
    if (threadIdx.x == 0) {
        fence_sys();  // this CTA's peer accesses are performed before the counter moves
        const uint32_t old = atomicInc((uint32_t*)op.arrive, op.nblocks - 1);
        if (old == op.nblocks - 1) {
            if (consumed) *(volatile uint64_t*)consumed = 0;
            st_release_sys((uint64_t*)flag, op.seq);  // release = fence + store
        }
    }
}

// Receive of a large message.  The sender's word says either where its payload lies (the
// caller then pulls it: returns the peer-mapped source) or that it will push (then CTA 0 posts
// this rank's buffer and waits like a control thread; returns nullptr).
__device__ __forceinline__ const uint8_t* recv_adaptive(const GroupHdr& h, const OpDesc& op, uint32_t part, uint64_t* smem_slot) {
    if (threadIdx.x == 0) {
        uint64_t v = 0;
        if (!wait_word(h, (const uint64_t*)op.f_wait2, op.seq, v)) {
            report_error(h, AERR_TIMEOUT_POST);
            v = 0;
        }
        *smem_slot = v;
    }
    __syncthreads();
    const uint64_t v = *smem_slot;
    const uint32_t win = (uint32_t)(v >> 40) & 0xffu;
    // Only flags so far.  A pull goes on without waiting for the kernel ahead on this stream: its first loads (from
    // the SENDER's buffer, declared ready by the word just read) may run ahead, its stores wait inside the copy
    // loop.  Every other outcome writes or offers this rank's buffer from here on: wait now.
    if (v == 0 || win == kWinPushPlease) pdl_wait(h);
    if (v == 0) return nullptr;
    if (win == kWinPushPlease) {
        if (!op.word) return (const uint8_t*)kStagedMarker;  // neither buffer is in a window: the staging ring
        if (threadIdx.x == 0) {
            if (part == 0) post_and_wait(h, op);
            // every CTA of the op has read the sender's word: the last one through clears the slot (see movers_signal)
            const uint32_t old = atomicInc((uint32_t*)op.arrive, op.nblocks - 1);
            if (old == op.nblocks - 1) *(volatile uint64_t*)op.f_wait2 = 0;
        }
        return nullptr;
    }
    const uint64_t base = ((const uint64_t*)op.aux)[win];
    if (!base) {
        pdl_wait(h);
        if (threadIdx.x == 0) report_error(h, AERR_BAD_WINDOW);
        return nullptr;
    }
    return (const uint8_t*)(base + (v & kPostOffMask));
}


// ---------------------------------------------------------------- staging ring (unregistered receive buffers)

struct CopyJob;

// Receiver, before anything else: take the pair's ring (one staged message per ordered pair at a time) and
// tell the sender to use it.  Done by one thread; the other CTAs go straight to waiting for chunks.
__device__ __forceinline__ void stage_ask(const GroupHdr& h, const OpDesc& op, uint32_t part) {
    if (part != 0 || threadIdx.x != 0) return;
    const StageInfo* st = (const StageInfo*)op.stage;
    const uint64_t t0 = globaltimer_ns();
    uint32_t polls = 0;
    while (atomicCAS((uint32_t*)st->lock, 0u, 1u) != 0u) {
        __nanosleep(100);
        if ((++polls & 255u) == 0 && globaltimer_ns() - t0 > h.timeout_ns) { report_error(h, AERR_STAGE_BUSY); return; }
    }
    st_release_sys((uint64_t*)op.f_signal, make_post_word(op.seq, kWinStaged, 0));
}

// ---------------------------------------------------------------- eager (LL) roles

__device__ __forceinline__ uint64_t load_8_bytes_partial(const uint8_t* p, size_t avail) {
    uint64_t v = 0;
    if (avail >= 8 && ((uint64_t)p & 7) == 0) return *(const uint64_t*)p;
    const int n = avail < 8 ? (int)avail : 8;
    for (int i = 0; i < n; ++i) v |= (uint64_t)p[i] << (8 * i);
    return v;
}
__device__ __forceinline__ void store_8_bytes_partial(uint8_t* p, uint64_t v, size_t avail) {
    if (avail >= 8 && ((uint64_t)p & 7) == 0) { *(uint64_t*)p = v; return; }
    const int n = avail < 8 ? (int)avail : 8;
    for (int i = 0; i < n; ++i) p[i] = (uint8_t)(v >> (8 * i));
}

// CTA `part` of `op.nblocks` handles lines part*blockDim + tid, + nblocks*blockDim, ...
__device__ __forceinline__ void ll_send(const GroupHdr& h, const OpDesc& op, uint32_t part, uint64_t* smem_slot) {
    // slot reuse: the receiver acknowledged message seq - kLLSlots in this slot
    if (threadIdx.x == 0) {
        uint64_t ok = 1, v;
        if (op.seq > (uint64_t)kLLSlots) {
            const uint64_t want = op.seq - kLLSlots;
            if (!spin_flag((const uint64_t*)op.f_wait, h.timeout_ns, v, [want](uint64_t x) { return x == want; })) {
                report_error(h, AERR_TIMEOUT_LL_ACK);
                ok = 0;
            }
        }
        *smem_slot = ok;
    }
    __syncthreads();
    if (!*smem_slot) return;
    const uint8_t* src = (const uint8_t*)op.local;
    uint8_t* ring = (uint8_t*)op.aux;
    const size_t nlines = op.nbytes ? (op.nbytes + 7) / 8 : 1;  // a 0-byte message still needs a flag
    const uint32_t flag = (uint32_t)op.seq;
    const size_t stride = (size_t)op.nblocks * blockDim.x;
    const bool fast = ((uint64_t)src & 7) == 0;
    size_t e = (size_t)part * blockDim.x + threadIdx.x;
    const size_t full = op.nbytes / 8;  // lines that carry 8 whole payload bytes
    if (fast) {
#pragma unroll 4
        for (; e < full; e += stride) st_ll_line(ring + e * 16, *(const uint64_t*)(src + e * 8), flag);
    }
    for (; e < nlines; e += stride) {
        const size_t off = e * 8;
        const uint64_t data = off < op.nbytes ? load_8_bytes_partial(src + off, op.nbytes - off) : 0;
        st_ll_line(ring + e * 16, data, flag);
    }
}

__device__ __forceinline__ void ll_recv(const GroupHdr& h, const OpDesc& op, uint32_t part) {
    uint8_t* dst = (uint8_t*)op.local;
    const uint8_t* ring = (const uint8_t*)op.aux;
    const size_t nlines = op.nbytes ? (op.nbytes + 7) / 8 : 1;
    const uint32_t flag = (uint32_t)op.seq;
    const size_t stride = (size_t)op.nblocks * blockDim.x;
    bool failed = false;
    for (size_t e = (size_t)part * blockDim.x + threadIdx.x; e < nlines && !failed; e += stride) {
        uint4 line = ld_ll_line(ring + e * 16);
        if (line.y != flag || line.w != flag) {
            const uint64_t t0 = globaltimer_ns();
            uint32_t polls = 0;
            for (;;) {
                line = ld_ll_line(ring + e * 16);
                if (line.y == flag && line.w == flag) break;
                if ((++polls & 1023u) == 0 && globaltimer_ns() - t0 > h.timeout_ns) { failed = true; break; }
            }
        }
        if (failed) { report_error(h, AERR_TIMEOUT_LL_DATA); break; }
        const size_t off = e * 8;
        if (off < op.nbytes)
            store_8_bytes_partial(dst + off, (uint64_t)line.x | ((uint64_t)line.z << 32), op.nbytes - off);
    }
    __syncthreads();
    // The slot may be overwritten once the sender sees the ack; the last CTA to finish sends it.
    // Every line of the slot has been loaded and consumed by then (the barrier above), so the
    // ack needs no fence of its own: a relaxed system-scope store keeps MEMBAR.SYS off the
    // latency path.
    if (threadIdx.x == 0) {
        if (op.nblocks == 1) {
            st_relaxed_sys((uint64_t*)op.f_signal, op.seq);
        } else {
            __threadfence();
            const uint32_t old = atomicInc((uint32_t*)op.arrive, op.nblocks - 1);
            if (old == op.nblocks - 1) st_relaxed_sys((uint64_t*)op.f_signal, op.seq);
        }
    }
}

// ---------------------------------------------------------------- dispatch

__device__ __forceinline__ const OpDesc* find_op(const GroupHdr& h, const OpDesc* ops, uint32_t bid) {
    for (uint32_t i = 0; i < h.nops; ++i) {
        const OpDesc& op = ops[i];
        if (op.kind != OP_POST && bid >= op.first_block && bid < op.first_block + op.nblocks) return &op;
    }
    return nullptr;
}

// Control CTA: thread t serves the t-th OP_POST of the group.
__device__ __forceinline__ void run_ctrl(const GroupHdr& h, const OpDesc* ops) {
    const OpDesc* mine = nullptr;
    uint32_t seen = 0;
    for (uint32_t i = 0; i < h.nops; ++i) {
        if (ops[i].kind != OP_POST) continue;
        if (seen++ == threadIdx.x) { mine = &ops[i]; break; }
    }
    bool posted = false;
    if (mine && (mine->flags & OPF_EARLY_POST) && predecessor_is_ours(h)) {
        st_release_sys((uint64_t*)mine->f_signal, mine->word);
        posted = true;
    }
    pdl_wait(h);
    pdl_trigger(h);  // the control CTA only waits on flags: the next kernel may come in at once
    if (mine) post_and_wait(h, *mine, posted);
}

// What a mover CTA has to copy once its role has been resolved.
struct CopyJob {
    uint8_t* dst;
    const uint8_t* src;
    size_t nbytes;       // bytes of this pass (the message, or one chunk of a staged one)
    uint64_t signal;     // flag the last CTA publishes afterwards (0: none)
    uint64_t consumed;   // local slot the op's tagged word arrived in, zeroed by that last CTA
    const OpDesc* op;
    uint32_t part;
    bool is_push;        // destination is peer memory (the TMA flavour stages these)
    uint32_t staged;     // 0 direct, 1 sending through the staging ring, 2 receiving through it
    uint32_t chunk;      // staged: index of the chunk this pass moves (== nchunks: this CTA has none left)
    uint32_t nchunks;
    const GroupHdr* wait_hdr;  // set: griddepcontrol.wait is still owed and happens inside the copy loop (pulls)
};

// Which chunks a CTA of a staged message serves: lane l belongs to CTA l % nblocks (on either side, whatever the
// two sides' CTA counts are), a CTA walks its chunks in increasing order.  Returns nchunks when there is none left.
__device__ __forceinline__ uint32_t stage_next_chunk(uint32_t from, uint32_t nchunks, uint32_t lanes, uint32_t part, uint32_t nblocks) {
    for (uint32_t c = from; c < nchunks; ++c)
        if ((c % lanes) % nblocks == part) return c;
    return nchunks;
}

// Staged message, first of all: how many chunks there are and which one this CTA starts with.
__device__ __forceinline__ void stage_start(CopyJob& job) {
    const StageInfo* st = (const StageInfo*)job.op->stage;
    job.nchunks = (uint32_t)((job.op->nbytes + st->chunk - 1) / st->chunk);
    job.chunk = stage_next_chunk(0, job.nchunks, st->lanes, job.part, job.op->nblocks);
}

// Staged message, before a chunk is copied: wait until its slot may be written (send: the receiver has drained the
// chunk that used it before) or read (receive: the sender has filled it); point the job at it.
__device__ __forceinline__ bool stage_begin(const GroupHdr& h, CopyJob& job, uint64_t* smem_slot) {
    const OpDesc& op = *job.op;
    const StageInfo* st = (const StageInfo*)op.stage;
    const uint32_t c = job.chunk, lane = c % st->lanes, k = c / st->lanes;
    const bool send = job.staged == 1;
    __syncthreads();  // smem_slot is reused from pass to pass
    if (threadIdx.x == 0) {
        uint64_t ok = 1, v;
        if (!send || k >= st->slots) {
            // counts of this very message only: an older message's count carries a smaller sequence number
            const uint64_t want = send ? stage_count(op.seq, k + 1 - st->slots) : stage_count(op.seq, k + 1);
            const uint64_t* flag = (const uint64_t*)(send ? st->send_drained : st->recv_filled) + lane;
            if (!spin_flag(flag, h.timeout_ns, v, [want](uint64_t x) { return x >= want; })) {
                report_error(h, AERR_TIMEOUT_STAGE);
                ok = 0;
            }
        }
        *smem_slot = ok;
    }
    __syncthreads();
    if (!*smem_slot) return false;
    const size_t off = (size_t)c * st->chunk;
    job.nbytes = op.nbytes - off < st->chunk ? op.nbytes - off : st->chunk;
    uint8_t* user = (uint8_t*)op.local + off;
    uint8_t* slot = (uint8_t*)(send ? st->send_ring : st->recv_ring) + ((size_t)lane * st->slots + k % st->slots) * st->chunk;
    job.dst = send ? slot : user;
    job.src = send ? user : slot;
    return true;
}

// After the chunk: publish the lane's count.  After a CTA's last chunk: the last CTA of the op to get there zeroes
// the consumed post slot and (receiver) frees the pair's ring.  Returns true while this CTA has chunks left.
__device__ __forceinline__ bool stage_end(const GroupHdr& h, CopyJob& job) {
    const OpDesc& op = *job.op;
    const StageInfo* st = (const StageInfo*)op.stage;
    const bool send = job.staged == 1;
    const uint32_t nchunks = job.nchunks;
    const uint32_t c = job.chunk;
    const uint32_t next = c < nchunks ? stage_next_chunk(c + 1, nchunks, st->lanes, job.part, op.nblocks) : nchunks;
    __syncthreads();  // every thread's accesses to the slot are issued (receive: its loads have returned)
    if (next >= nchunks) pdl_trigger_late(h);
    if (threadIdx.x == 0) {
        if (c < nchunks) {
            const uint32_t lane = c % st->lanes, k = c / st->lanes;
            if (send) st_release_sys((uint64_t*)st->send_filled + lane, stage_count(op.seq, k + 1));
            else st_relaxed_sys((uint64_t*)st->recv_drained + lane, stage_count(op.seq, k + 1));  // loads are done: no fence needed
        }
        if (next >= nchunks) {
            const uint32_t old = atomicInc((uint32_t*)(send ? st->send_arrive : st->recv_arrive), op.nblocks - 1);
            if (old == op.nblocks - 1) {
                if (job.consumed) *(volatile uint64_t*)job.consumed = 0;
                if (!send) {
                    __threadfence();
                    atomicExch((uint32_t*)st->lock, 0u);  // the pair's ring is free again
                }
            }
        }
    }
    job.chunk = next;
    return next < nchunks;
}

// Runs the control / eager roles to completion and resolves the mover roles to a CopyJob, so
// that each kernel has exactly one inlined copy loop.  Returns false when this CTA is done.
__device__ __forceinline__ bool resolve_role(const GroupHdr& h, const OpDesc* ops, uint64_t* slot, CopyJob& job) {
    const uint32_t bid = blockIdx.x;
    if (h.n_ctrl && bid == 0) { run_ctrl(h, ops); return false; }
    const OpDesc* op = find_op(h, ops, bid);
    if (!op) { pdl_wait(h); return false; }
    if (op->kind != OP_PUSH && op->kind != OP_RECV_ADAPT) pdl_wait(h);  // those two shake hands first
    pdl_trigger_early(h);
    const uint32_t part = bid - op->first_block;
    job.op = op;
    job.part = part;
    job.is_push = false;
    job.nbytes = op->nbytes;
    job.staged = 0;
    job.chunk = 0;
    job.nchunks = 0;
    job.wait_hdr = nullptr;
    switch (op->kind) {
        case OP_PUSH:
            job.dst = push_wait_post(h, *op, part, slot);
            job.src = (const uint8_t*)op->local;
            job.signal = op->f_signal;
            job.consumed = op->f_wait;
            job.is_push = true;
            if ((uint64_t)job.dst == kStagedMarker) { job.staged = 1; job.signal = 0; }
            return job.dst != nullptr;
        case OP_COPY:
            job.dst = (uint8_t*)op->aux;
            job.src = (const uint8_t*)op->local;
            job.signal = 0;
            job.consumed = 0;
            return true;
        case OP_RECV_ADAPT:
            job.src = recv_adaptive(h, *op, part, slot);
            job.dst = (uint8_t*)op->local;
            job.signal = op->f_signal2;
            job.consumed = op->f_wait2;
            if ((uint64_t)job.src == kStagedMarker) { stage_ask(h, *op, part); job.staged = 2; job.signal = 0; }
            else if (job.src) job.wait_hdr = &h;   // a pull: see recv_adaptive
            return job.src != nullptr;
        case OP_RECV_STAGED:
            stage_ask(h, *op, part);
            job.staged = 2;
            job.signal = 0;
            job.consumed = 0;
            job.dst = (uint8_t*)op->local;
            job.src = nullptr;
            return true;
        case OP_SEND_LL: ll_send(h, *op, part, slot); return false;
        case OP_RECV_LL: ll_recv(h, *op, part); return false;
        default: return false;
    }
}

}  // namespace p2p
