// Device side of the transfer engine: hand-written sm_100a data movement + signalling.
//
// One launch (`p2p_group_kernel_*`) executes every operation of a group, the way one
// ncclDevKernel_SendRecv launch serves a ncclGroupStart/End pair in the reference's
// bidirectional loop (p2p_matrix.cc:212-228).  CTAs are assigned to operations by block
// range; the first CTA is the control CTA that posts and completes rendezvous receives.
//
// Protocols
//   rendezvous  receiver posts {seq, window, offset} into the SENDER's control region;
//               sender CTAs push payload straight into the receiver's buffer through the
//               peer mapping, then the last CTA to finish publishes `done[seq]` in the
//               RECEIVER's control region with a system-scope release store.
//   eager (LL)  8 payload bytes + two 4-byte flags per 16-byte store into a ring in the
//               receiver's control region; the receiver copies out as flags turn valid.
#pragma once
#include <cstdint>
#include <cuda_runtime.h>

namespace p2p {

constexpr int kRing = 64;            // outstanding rendezvous messages per ordered pair
constexpr int kLLSlots = 8;          // eager ring depth per ordered pair
constexpr int kLdstThreads = 512;
constexpr int kTmaThreads = 128;
constexpr int kMaxGroupOps = 32;
constexpr int kTmaMaxStages = 16;

enum OpKind : uint32_t {
    OP_NONE = 0,
    OP_RECV_RDV = 1,   // handled by one thread of the control CTA
    OP_SEND_RDV = 2,   // push payload (ld/st or TMA, by kernel flavour)
    OP_SEND_LL = 3,
    OP_RECV_LL = 4,
};

enum AsyncError : uint32_t { AERR_NONE = 0, AERR_TIMEOUT_POST = 1, AERR_TIMEOUT_DONE = 2,
                             AERR_TIMEOUT_LL_DATA = 3, AERR_TIMEOUT_LL_ACK = 4, AERR_BAD_WINDOW = 5 };

struct alignas(16) OpDesc {
    uint64_t local;        // send: source pointer; recv: destination pointer (this GPU)
    uint64_t nbytes;
    uint64_t wait_flag;    // local flag this op polls (send: post slot, recv: done slot, LL send: ack slot)
    uint64_t signal_flag;  // peer-mapped flag this op writes (send: done, recv: post, LL recv: ack)
    uint64_t word;         // recv: the post word; others unused
    uint64_t seq;          // sequence number of this message on its (src,dst,protocol) lane
    uint64_t aux;          // send: table of the peer's window bases; LL: ring slot (peer-mapped for send)
    uint64_t arrive;       // send: local u32 arrival counter
    uint32_t kind;
    uint32_t first_block;
    uint32_t nblocks;
    uint32_t peer;
};

struct GroupArgs {
    uint32_t nops;
    uint32_t n_ctrl;         // rendezvous receives served by the control CTA (block 0) or 0
    uint32_t vec_bytes;      // 16 or 32: ld/st width of the push loop
    uint32_t tma_chunk;      // bytes per cp.async.bulk (TMA flavour)
    uint32_t tma_stages;
    uint32_t pad;
    uint64_t timeout_ns;
    uint32_t* async_err;     // host-mapped word; first error wins
    OpDesc ops[kMaxGroupOps];
};

// ---------------------------------------------------------------- PTX helpers

__device__ __forceinline__ void st_release_sys(uint64_t* p, uint64_t v) {
    asm volatile("st.release.sys.global.u64 [%0], %1;" ::"l"(p), "l"(v) : "memory");
}
__device__ __forceinline__ uint64_t ld_acquire_sys(const uint64_t* p) {
    uint64_t v;
    asm volatile("ld.acquire.sys.global.u64 %0, [%1];" : "=l"(v) : "l"(p) : "memory");
    return v;
}
__device__ __forceinline__ uint64_t ld_relaxed_sys(const uint64_t* p) {
    uint64_t v;
    asm volatile("ld.relaxed.sys.global.u64 %0, [%1];" : "=l"(v) : "l"(p) : "memory");
    return v;
}
__device__ __forceinline__ void fence_sys() { asm volatile("fence.acq_rel.sys;" ::: "memory"); }
__device__ __forceinline__ uint64_t globaltimer_ns() {
    uint64_t t;
    asm volatile("mov.u64 %0, %%globaltimer;" : "=l"(t));
    return t;
}

// streaming 16- and 32-byte accesses: no L1 allocation, the payload is touched once
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
struct alignas(32) vec32 { uint32_t r[8]; };
__device__ __forceinline__ vec32 ld_stream32(const void* p) {
    vec32 v;
    asm volatile("ld.global.L1::no_allocate.v8.b32 {%0,%1,%2,%3,%4,%5,%6,%7}, [%8];"
                 : "=r"(v.r[0]), "=r"(v.r[1]), "=r"(v.r[2]), "=r"(v.r[3]),
                   "=r"(v.r[4]), "=r"(v.r[5]), "=r"(v.r[6]), "=r"(v.r[7]) : "l"(p) : "memory");
    return v;
}
__device__ __forceinline__ void st_stream32(void* p, const vec32& v) {
    asm volatile("st.global.L1::no_allocate.v8.b32 [%0], {%1,%2,%3,%4,%5,%6,%7,%8};"
                 ::"l"(p), "r"(v.r[0]), "r"(v.r[1]), "r"(v.r[2]), "r"(v.r[3]),
                   "r"(v.r[4]), "r"(v.r[5]), "r"(v.r[6]), "r"(v.r[7]) : "memory");
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

__device__ __forceinline__ void report_error(const GroupArgs& a, uint32_t code) {
    if (a.async_err) atomicCAS_system(a.async_err, 0u, code);
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
// stream neighbouring addresses (DRAM pages and NVLink packets stay dense).
template <typename V, int UNROLL, typename LD, typename ST>
__device__ __forceinline__ void copy_vectors(V* __restrict__ dst, const V* __restrict__ src,
                                             size_t nvec, uint32_t part, uint32_t nparts, LD ld, ST st) {
    const size_t tile = (size_t)blockDim.x * UNROLL;
    const size_t ntiles = nvec / tile;
    for (size_t t = part; t < ntiles; t += nparts) {
        const V* s = src + t * tile + threadIdx.x;
        V* d = dst + t * tile + threadIdx.x;
        V v[UNROLL];
#pragma unroll
        for (int u = 0; u < UNROLL; ++u) v[u] = ld(s + (size_t)u * blockDim.x);
#pragma unroll
        for (int u = 0; u < UNROLL; ++u) st(d + (size_t)u * blockDim.x, v[u]);
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
#pragma unroll 4
    for (; i < n; i += stride) dst[i] = src[i];
}

// Byte-exact copy of n bytes for arbitrary alignment of either side.
__device__ __forceinline__ void copy_bytes(uint8_t* dst, const uint8_t* src, size_t n, uint32_t part,
                                           uint32_t nparts, uint32_t vec_bytes) {
    if (n == 0) return;
    const uint64_t rel = ((uint64_t)dst ^ (uint64_t)src);
    if ((rel & 15) == 0) {
        // same phase: byte head up to a 32-byte boundary, wide body, byte tail
        size_t head = (32 - ((uint64_t)dst & 31)) & 31;
        if ((rel & 31) != 0 || vec_bytes != 32) head = (16 - ((uint64_t)dst & 15)) & 15;
        if (head > n) head = n;
        if (part == 0)
            for (size_t i = threadIdx.x; i < head; i += blockDim.x) dst[i] = src[i];
        uint8_t* d = dst + head;
        const uint8_t* s = src + head;
        size_t left = n - head;
        size_t body;
        if ((rel & 31) == 0 && vec_bytes == 32) {
            body = left & ~(size_t)31;
            copy_vectors<vec32, 4>((vec32*)d, (const vec32*)s, body / 32, part, nparts,
                                   [](const vec32* p) { return ld_stream32(p); },
                                   [](vec32* p, const vec32& v) { st_stream32(p, v); });
        } else {
            body = left & ~(size_t)15;
            copy_vectors<uint4, 8>((uint4*)d, (const uint4*)s, body / 16, part, nparts,
                                   [](const uint4* p) { return ld_stream16(p); },
                                   [](uint4* p, const uint4& v) { st_stream16(p, v); });
        }
        if (part == nparts - 1)
            for (size_t i = body + threadIdx.x; i < left; i += blockDim.x) d[i] = s[i];
        return;
    }
    // different phase: widest element both sides can be aligned to at once
    const uint32_t w = (rel & 7) == 0 ? 8 : (rel & 3) == 0 ? 4 : (rel & 1) == 0 ? 2 : 1;
    size_t head = (w - ((uint64_t)dst & (w - 1))) & (w - 1);
    if (head > n) head = n;
    if (part == 0)
        for (size_t i = threadIdx.x; i < head; i += blockDim.x) dst[i] = src[i];
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
        for (size_t i = ne * w + threadIdx.x; i < left; i += blockDim.x) d[i] = s[i];
}

// ---------------------------------------------------------------- rendezvous roles

// Post word layout: [63:48] low 16 bits of seq, [47:40] window id, [39] valid (always 1, so
// a real post is never the all-zero word of a fresh slot), [38:0] byte offset (512 GiB).
constexpr uint64_t kPostValid = 1ull << 39;
constexpr uint64_t kPostOffMask = kPostValid - 1;
__host__ __device__ __forceinline__ uint64_t make_post_word(uint64_t seq, uint32_t win, uint64_t off) {
    return ((seq & 0xffffull) << 48) | ((uint64_t)(win & 0xffu) << 40) | kPostValid | (off & kPostOffMask);
}

// One thread per rendezvous receive: publish where the payload goes, then wait for it.
__device__ __forceinline__ void ctrl_recv(const GroupArgs& a, const OpDesc& op) {
    st_release_sys((uint64_t*)op.signal_flag, op.word);
    uint64_t v;
    const uint64_t want = op.seq;
    if (!spin_flag((const uint64_t*)op.wait_flag, a.timeout_ns, v, [want](uint64_t x) { return x == want; }))
        report_error(a, AERR_TIMEOUT_DONE);
}

// All CTAs of a send learn the destination; returns nullptr on timeout / bad window.
__device__ __forceinline__ uint8_t* send_wait_post(const GroupArgs& a, const OpDesc& op, uint64_t* smem_slot) {
    if (threadIdx.x == 0) {
        uint64_t v = 0;
        const uint64_t tag = op.seq & 0xffffull;
        uint64_t dst = 0;
        if (spin_flag((const uint64_t*)op.wait_flag, a.timeout_ns, v,
                      [tag](uint64_t x) { return (x >> 48) == tag && (x & kPostValid); })) {
            const uint32_t win = (uint32_t)(v >> 40) & 0xffu;
            const uint64_t base = ((const uint64_t*)op.aux)[win];
            if (base) dst = base + (v & kPostOffMask);
            else report_error(a, AERR_BAD_WINDOW);
        } else {
            report_error(a, AERR_TIMEOUT_POST);
        }
        *smem_slot = dst;
    }
    __syncthreads();
    return (uint8_t*)(*smem_slot);
}

// After this CTA's stores: the last CTA of the op to get here publishes completion.
__device__ __forceinline__ void send_signal_done(const OpDesc& op) {
    __syncthreads();
    if (threadIdx.x == 0) {
        fence_sys();  // this CTA's peer stores are performed before the counter moves
        const uint32_t old = atomicInc((uint32_t*)op.arrive, op.nblocks - 1);
        if (old == op.nblocks - 1) {
            fence_sys();
            st_release_sys((uint64_t*)op.signal_flag, op.seq);
        }
    }
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

__device__ __forceinline__ void ll_send(const GroupArgs& a, const OpDesc& op, uint64_t* smem_slot) {
    // slot reuse: the receiver acknowledged message seq - kLLSlots in this slot
    if (threadIdx.x == 0) {
        uint64_t ok = 1, v;
        if (op.seq > (uint64_t)kLLSlots) {
            const uint64_t want = op.seq - kLLSlots;
            if (!spin_flag((const uint64_t*)op.wait_flag, a.timeout_ns, v,
                           [want](uint64_t x) { return x == want; })) {
                report_error(a, AERR_TIMEOUT_LL_ACK);
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
    for (size_t e = threadIdx.x; e < nlines; e += blockDim.x) {
        const size_t off = e * 8;
        const uint64_t data = off < op.nbytes ? load_8_bytes_partial(src + off, op.nbytes - off) : 0;
        st_ll_line(ring + e * 16, data, flag);
    }
}

__device__ __forceinline__ void ll_recv(const GroupArgs& a, const OpDesc& op) {
    uint8_t* dst = (uint8_t*)op.local;
    const uint8_t* ring = (const uint8_t*)op.aux;
    const size_t nlines = op.nbytes ? (op.nbytes + 7) / 8 : 1;
    const uint32_t flag = (uint32_t)op.seq;
    bool failed = false;
    for (size_t e = threadIdx.x; e < nlines && !failed; e += blockDim.x) {
        uint4 line = ld_ll_line(ring + e * 16);
        if (line.y != flag || line.w != flag) {
            const uint64_t t0 = globaltimer_ns();
            uint32_t polls = 0;
            for (;;) {
                line = ld_ll_line(ring + e * 16);
                if (line.y == flag && line.w == flag) break;
                if ((++polls & 1023u) == 0 && globaltimer_ns() - t0 > a.timeout_ns) { failed = true; break; }
            }
        }
        if (failed) { report_error(a, AERR_TIMEOUT_LL_DATA); break; }
        const size_t off = e * 8;
        if (off < op.nbytes)
            store_8_bytes_partial(dst + off, (uint64_t)line.x | ((uint64_t)line.z << 32), op.nbytes - off);
    }
    __syncthreads();
    // slot may be overwritten once the sender sees this
    if (threadIdx.x == 0) st_release_sys((uint64_t*)op.signal_flag, op.seq);
}

// ---------------------------------------------------------------- dispatch

__device__ __forceinline__ const OpDesc* find_op(const GroupArgs& a, uint32_t bid) {
    for (uint32_t i = 0; i < a.nops; ++i) {
        const OpDesc& op = a.ops[i];
        if (op.kind != OP_RECV_RDV && bid >= op.first_block && bid < op.first_block + op.nblocks) return &op;
    }
    return nullptr;
}

// Control CTA: thread t serves the t-th rendezvous receive of the group.
__device__ __forceinline__ void run_ctrl(const GroupArgs& a) {
    uint32_t seen = 0;
    for (uint32_t i = 0; i < a.nops; ++i) {
        if (a.ops[i].kind != OP_RECV_RDV) continue;
        if (seen++ == threadIdx.x) { ctrl_recv(a, a.ops[i]); return; }
    }
}

}  // namespace p2p
