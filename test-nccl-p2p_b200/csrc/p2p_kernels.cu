// Kernels of the transfer engine and their launch wrappers (compiled for sm_100a only).
#include "p2p.h"
#include "p2p_internal.h"
#include "p2p_tma.cuh"

namespace p2p {

// ------------------------------------------------------------------ group kernels

// ld/st flavour: up to 512-thread CTAs, payload moves through registers as 16-byte
// coalesced vectors (LDG.E.128 + STG.E.128 on the peer-mapped address).
template <int MAXOPS>
__global__ void __launch_bounds__(kLdstThreads, 2)
p2p_group_kernel_ldst(const __grid_constant__ GroupArgsT<MAXOPS> a) {
    __shared__ uint64_t slot;
    CopyJob job;
    if (!resolve_role(a.h, a.ops, &slot, job)) return;
    if (job.staged) stage_start(job);
    for (;;) {  // one pass - or, for a message that goes through the staging ring, one per chunk of this CTA's lanes
        const bool idle = job.staged && job.chunk >= job.nchunks;   // more CTAs than lanes, or a short message
        if (job.staged && !idle && !stage_begin(a.h, job, &slot)) return;
        if (!idle) copy_bytes(job.dst, job.src, job.nbytes, job.staged ? 0 : job.part, job.staged ? 1 : job.op->nblocks, job.wait_hdr);
        if (!job.staged || !stage_end(a.h, job)) break;
    }
    if (job.signal) movers_signal(a.h, *job.op, job.signal, job.consumed);
    else if (!job.staged) pdl_trigger_late(a.h);
}

// TMA flavour: 128-thread CTAs, a push is staged global -> shared -> peer global by
// cp.async.bulk (UBLKCP); two elected threads drive the ring.
__global__ void __launch_bounds__(kTmaThreads, 1)
p2p_group_kernel_tma(const __grid_constant__ GroupArgs a) {
    extern __shared__ __align__(128) uint8_t smem_ring[];
    __shared__ uint64_t slot;
    __shared__ __align__(8) uint64_t full[kTmaMaxStages];
    __shared__ __align__(8) uint64_t empty[kTmaMaxStages];
    if (threadIdx.x == 0) {
        for (uint32_t i = 0; i < a.h.tma_stages; ++i) { mbar_init(&full[i], 1); mbar_init(&empty[i], 1); }
        asm volatile("fence.mbarrier_init.release.cluster;" ::: "memory");
        fence_proxy_async();
    }
    __syncthreads();
    CopyJob job;
    if (!resolve_role(a.h, a.ops, &slot, job)) return;
    // the ring moves pushes, pulls and local copies whose two sides share their 16-byte phase; anything else, and
    // the chunks of a message that goes through the staging ring, keep the register loop
    if (job.staged) stage_start(job);
    for (;;) {
        const bool idle = job.staged && job.chunk >= job.nchunks;
        if (job.staged && !idle && !stage_begin(a.h, job, &slot)) return;
        if (!job.staged && (((uint64_t)job.dst ^ (uint64_t)job.src) & 15) == 0)
            tma_push(job.dst, job.src, job.nbytes, job.part, job.op->nblocks, a.h.tma_chunk, a.h.tma_stages, smem_ring, full, empty, job.wait_hdr);
        else if (!idle)
            copy_bytes(job.dst, job.src, job.nbytes, job.staged ? 0 : job.part, job.staged ? 1 : job.op->nblocks, job.wait_hdr);
        if (!job.staged || !stage_end(a.h, job)) break;
    }
    if (job.signal) movers_signal(a.h, *job.op, job.signal, job.consumed);
    else if (!job.staged) pdl_trigger_late(a.h);
}

cudaError_t configure_kernels(int max_tma_smem) {
    return cudaFuncSetAttribute(p2p_group_kernel_tma, cudaFuncAttributeMaxDynamicSharedMemorySize, max_tma_smem);
}

cudaError_t occupancy(int flavour, int dyn_smem, int* ctas_per_sm) {
    if (flavour == FLAVOUR_TMA)
        return cudaOccupancyMaxActiveBlocksPerMultiprocessor(ctas_per_sm, p2p_group_kernel_tma, kTmaThreads, dyn_smem);
    return cudaOccupancyMaxActiveBlocksPerMultiprocessor(ctas_per_sm, p2p_group_kernel_ldst<kMaxGroupOps>, kLdstThreads, 0);
}

template <typename Args>
static cudaError_t launch_ex(void (*kernel)(Args), const Args& args, uint32_t grid, uint32_t block, int dyn_smem,
                             cudaStream_t stream, bool pdl) {
    cudaLaunchConfig_t cfg = {};
    cfg.gridDim = dim3(grid);
    cfg.blockDim = dim3(block);
    cfg.dynamicSmemBytes = (size_t)dyn_smem;
    cfg.stream = stream;
    cudaLaunchAttribute attr[1];
    attr[0].id = cudaLaunchAttributeProgrammaticStreamSerialization;
    attr[0].val.programmaticStreamSerializationAllowed = 1;
    cfg.attrs = attr;
    cfg.numAttrs = pdl ? 1 : 0;
    return cudaLaunchKernelEx(&cfg, kernel, args);
}

cudaError_t launch_group(int flavour, const GroupArgs& args, uint32_t grid, uint32_t block, int dyn_smem, cudaStream_t stream) {
    const bool pdl = args.h.pdl != 0;
    if (flavour == FLAVOUR_TMA) return launch_ex(p2p_group_kernel_tma, args, grid, kTmaThreads, dyn_smem, stream, pdl);
    if (args.h.nops <= (uint32_t)kSmallGroupOps) {
        // same kernel with a 0.5 KB parameter block instead of 3.4 KB: less to copy per launch
        GroupArgsSmall small;
        small.h = args.h;
        for (uint32_t i = 0; i < args.h.nops; ++i) small.ops[i] = args.ops[i];
        for (uint32_t i = args.h.nops; i < (uint32_t)kSmallGroupOps; ++i) small.ops[i] = OpDesc{};
        return launch_ex(p2p_group_kernel_ldst<kSmallGroupOps>, small, grid, block, 0, stream, pdl);
    }
    return launch_ex(p2p_group_kernel_ldst<kMaxGroupOps>, args, grid, block, 0, stream, pdl);
}

// ------------------------------------------------------------------ copy-engine path: the two flag kernels
//
// For the largest one-to-one messages the payload is moved by a DMA engine (cudaMemcpyAsync on the
// peer-mapped destination: 766 GB/s, against 700 for SM-issued stores); the protocol around it stays on
// the device.  ce_wait tells the receiver "I will push", then holds the stream until the receiver's
// post (buffer ready) has arrived; ce_signal publishes completion after the copy has finished.

__global__ void ce_wait_kernel(uint64_t* notice_flag, uint64_t notice_word, const uint64_t* post_slot, uint64_t seq,
                               uint64_t expect_word, GroupHdr h) {
    if (threadIdx.x != 0) return;
    st_release_sys(notice_flag, notice_word);
    uint64_t v = 0;
    if (!wait_word(h, post_slot, seq, v)) report_error(h, AERR_TIMEOUT_POST);
    else if (v != expect_word) report_error(h, AERR_BAD_WINDOW);  // the host was told another address
}

__global__ void ce_signal_kernel(uint64_t* done_flag, uint64_t seq) {
    if (threadIdx.x != 0) return;
    st_release_sys(done_flag, seq);  // fence.sys + store: ordered after the copy the stream completed before this launch
}

cudaError_t launch_ce_wait(uint64_t* notice_flag, uint64_t notice_word, const uint64_t* post_slot, uint64_t seq,
                           uint64_t expect_word, const GroupHdr& h, cudaStream_t stream) {
    ce_wait_kernel<<<1, 32, 0, stream>>>(notice_flag, notice_word, post_slot, seq, expect_word, h);
    return cudaGetLastError();
}

cudaError_t launch_ce_signal(uint64_t* done_flag, uint64_t seq, cudaStream_t stream) {
    ce_signal_kernel<<<1, 32, 0, stream>>>(done_flag, seq);
    return cudaGetLastError();
}

// ------------------------------------------------------------------ pattern / digest kernels

constexpr int kPatThreads = 256;

__global__ void __launch_bounds__(kPatThreads)
pattern_fill_kernel(uint8_t* buf, size_t nbytes, uint64_t key) {
    const size_t nwords = (nbytes + 7) / 8;
    const bool aligned = ((uint64_t)buf & 7) == 0;
    for (size_t w = (size_t)blockIdx.x * blockDim.x + threadIdx.x; w < nwords; w += (size_t)gridDim.x * blockDim.x) {
        const uint64_t v = pattern_word(key, w);
        const size_t off = w * 8;
        if (aligned && off + 8 <= nbytes) {
            *(uint64_t*)(buf + off) = v;
        } else {
            const int n = nbytes - off < 8 ? (int)(nbytes - off) : 8;
            for (int i = 0; i < n; ++i) buf[off + i] = (uint8_t)(v >> (8 * i));
        }
    }
}

__device__ __forceinline__ uint64_t load_word_le(const uint8_t* buf, size_t off, size_t nbytes, bool aligned) {
    if (aligned && off + 8 <= nbytes) return *(const uint64_t*)(buf + off);
    uint64_t v = 0;
    const int n = nbytes - off < 8 ? (int)(nbytes - off) : 8;
    for (int i = 0; i < n; ++i) v |= (uint64_t)buf[off + i] << (8 * i);
    return v;
}

// One pass: mismatching bytes vs the pattern (if check), first bad offset, digest.
// Per-CTA partials go to `partials`; the last CTA to finish folds them into `out`.
__global__ void __launch_bounds__(kPatThreads)
pattern_check_kernel(const uint8_t* buf, size_t nbytes, uint64_t key, int check, CheckPartial* partials,
                     unsigned int* ticket, CheckPartial* out) {
    const size_t nwords = (nbytes + 7) / 8;
    const bool aligned = ((uint64_t)buf & 7) == 0;
    uint64_t bad = 0, first = UINT64_MAX, dig = 0;
    for (size_t w = (size_t)blockIdx.x * blockDim.x + threadIdx.x; w < nwords; w += (size_t)gridDim.x * blockDim.x) {
        const size_t off = w * 8;
        const uint64_t v = load_word_le(buf, off, nbytes, aligned);
        dig += digest_term(v, w);
        if (check) {
            uint64_t want = pattern_word(key, w);
            if (nbytes - off < 8) want &= (1ull << (8 * (nbytes - off))) - 1;
            uint64_t x = v ^ want;
            if (x) {
                // fold every differing byte to its low bit, count, and locate the first
                uint64_t m = x | (x >> 4);
                m |= m >> 2;
                m |= m >> 1;
                m &= 0x0101010101010101ull;
                bad += __popcll(m);
                const uint64_t pos = off + ((__ffsll((long long)m) - 1) >> 3);
                first = pos < first ? pos : first;
            }
        }
    }
    // CTA reduction
    __shared__ uint64_t s_bad[kPatThreads / 32], s_first[kPatThreads / 32], s_dig[kPatThreads / 32];
    for (int o = 16; o > 0; o >>= 1) {
        bad += __shfl_xor_sync(0xffffffffu, bad, o);
        dig += __shfl_xor_sync(0xffffffffu, dig, o);
        const uint64_t f = __shfl_xor_sync(0xffffffffu, first, o);
        first = f < first ? f : first;
    }
    const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
    if (lane == 0) { s_bad[warp] = bad; s_first[warp] = first; s_dig[warp] = dig; }
    __syncthreads();
    __shared__ bool is_last;
    if (threadIdx.x == 0) {
        for (int i = 1; i < kPatThreads / 32; ++i) {
            bad += s_bad[i];
            dig += s_dig[i];
            first = s_first[i] < first ? s_first[i] : first;
        }
        partials[blockIdx.x].mismatches = bad;
        partials[blockIdx.x].first_bad = first;
        partials[blockIdx.x].digest = dig;
        __threadfence();
        const unsigned int t = atomicInc(ticket, gridDim.x - 1);
        is_last = (t == gridDim.x - 1);
    }
    __syncthreads();
    if (!is_last) return;
    __threadfence();
    bad = 0; first = UINT64_MAX; dig = 0;
    for (unsigned int i = threadIdx.x; i < gridDim.x; i += blockDim.x) {
        const volatile CheckPartial* p = &partials[i];
        bad += p->mismatches;
        dig += p->digest;
        const uint64_t f = p->first_bad;
        first = f < first ? f : first;
    }
    for (int o = 16; o > 0; o >>= 1) {
        bad += __shfl_xor_sync(0xffffffffu, bad, o);
        dig += __shfl_xor_sync(0xffffffffu, dig, o);
        const uint64_t f = __shfl_xor_sync(0xffffffffu, first, o);
        first = f < first ? f : first;
    }
    __syncthreads();
    if (lane == 0) { s_bad[warp] = bad; s_first[warp] = first; s_dig[warp] = dig; }
    __syncthreads();
    if (threadIdx.x == 0) {
        for (int i = 1; i < kPatThreads / 32; ++i) {
            bad += s_bad[i];
            dig += s_dig[i];
            first = s_first[i] < first ? s_first[i] : first;
        }
        out->mismatches = bad;
        out->first_bad = first < nbytes ? first : nbytes;
        out->digest = digest_finish(dig, nbytes);
    }
}

uint32_t pattern_grid(size_t nbytes, int sm_count) {
    const size_t nwords = (nbytes + 7) / 8;
    size_t want = (nwords + (size_t)kPatThreads * 4 - 1) / ((size_t)kPatThreads * 4);
    const size_t cap = (size_t)sm_count * 8;
    if (want > cap) want = cap;
    if (want < 1) want = 1;
    return (uint32_t)want;
}

cudaError_t launch_pattern_fill(void* buf, size_t nbytes, uint64_t key, int sm_count, cudaStream_t stream) {
    if (nbytes == 0) return cudaSuccess;
    pattern_fill_kernel<<<pattern_grid(nbytes, sm_count), kPatThreads, 0, stream>>>((uint8_t*)buf, nbytes, key);
    return cudaGetLastError();
}

cudaError_t launch_pattern_check(const void* buf, size_t nbytes, uint64_t key, int check, int sm_count,
                                 CheckPartial* partials, unsigned int* ticket, CheckPartial* out,
                                 cudaStream_t stream) {
    pattern_check_kernel<<<pattern_grid(nbytes, sm_count), kPatThreads, 0, stream>>>(
        (const uint8_t*)buf, nbytes, key, check, partials, ticket, out);
    return cudaGetLastError();
}

// ------------------------------------------------------------------ digest without allocation
//
// p2p_digest() keeps its per-CTA partial sums, its arrival ticket and its result in device
// memory that p2p_init() obtained with cudaMalloc.  Some callers of this library cannot have
// that: a serving process whose framework has already claimed every byte of HBM for its own
// pool (a late cudaMalloc simply fails there), a latency-critical loop that must not hit the
// implicit device-wide synchronisation of cudaMalloc/cudaFree, or a tool that wants to check a
// buffer without bringing the engine up at all.  For them the same kernel runs here against a
// scratch area that is part of the loaded module itself: it exists on every GPU that loads the
// library, from load until the process exits, and nobody has to ask for it.

constexpr size_t kNoAllocSpan = 256 * 1024;       // input bytes covered per CTA (one partial each)
constexpr unsigned kNoAllocMaxCtas = 4096;        // => P2P_NOALLOC_MAX_BYTES = 4096 * 256 KiB = 1 GiB
static_assert((unsigned long long)kNoAllocMaxCtas * kNoAllocSpan == P2P_NOALLOC_MAX_BYTES, "header and kernel disagree");

struct NoAllocScratch {
    CheckPartial partials[kNoAllocMaxCtas];
    CheckPartial out;
    unsigned int ticket;
};

// 96 KiB and change.  In host code this name is only a stand-in for the symbol; the address a
// kernel can use comes from cudaGetSymbolAddress, once per device (noalloc_scratch_for_device).
static __device__ __align__(256) unsigned char g_noalloc_scratch[(sizeof(NoAllocScratch) + 255) / 256 * 256];

}  // namespace p2p

namespace {
constexpr int kMaxDevices = 64;
p2p::NoAllocScratch* g_noalloc_ptr[kMaxDevices];  // per device, filled on first use
}

extern "C" int p2p_digest_noalloc(const void* buf, size_t bytes, p2p_stream_t stream, uint64_t* digest_dev) {
    using namespace p2p;
    if (!digest_dev) return set_last_error(P2P_INVALID_ARGUMENT, "p2p_digest_noalloc: null result pointer");
    if (bytes && !buf) return set_last_error(P2P_INVALID_ARGUMENT, "p2p_digest_noalloc: null buffer");
    if (bytes > P2P_NOALLOC_MAX_BYTES)
        return set_last_error(P2P_INVALID_ARGUMENT,
                              "p2p_digest_noalloc: %zu bytes exceed the %llu-byte limit of the built-in scratch area "
                              "(%u partial sums, one per %zu input bytes); use p2p_digest for larger buffers",
                              bytes, (unsigned long long)P2P_NOALLOC_MAX_BYTES, kNoAllocMaxCtas, kNoAllocSpan);
    int dev = -1;
    cudaError_t e = cudaGetDevice(&dev);
    if (e != cudaSuccess || dev < 0 || dev >= kMaxDevices)
        return set_last_error(P2P_UNHANDLED_CUDA_ERROR, "p2p_digest_noalloc: no usable current device (%s)", cudaGetErrorString(e));
    NoAllocScratch* scratch = g_noalloc_ptr[dev];
    if (!scratch) {
        // first call on this device: resolve the symbol to this device's copy of the array
        void* p = nullptr;
        e = cudaGetSymbolAddress(&p, g_noalloc_scratch);
        if (e != cudaSuccess || !p)
            return set_last_error(P2P_UNHANDLED_CUDA_ERROR, "p2p_digest_noalloc: cudaGetSymbolAddress failed on device %d: %s",
                                  dev, cudaGetErrorString(e));
        scratch = (NoAllocScratch*)p;
        g_noalloc_ptr[dev] = scratch;
    }
    cudaStream_t s = (cudaStream_t)stream;
    // the array still holds whatever the previous call left: the ticket must start from zero
    e = cudaMemsetAsync(&scratch->ticket, 0, sizeof(unsigned int), s);
    if (e != cudaSuccess) return set_last_error(P2P_UNHANDLED_CUDA_ERROR, "p2p_digest_noalloc: cudaMemsetAsync: %s", cudaGetErrorString(e));
    unsigned ctas = (unsigned)((bytes + kNoAllocSpan - 1) / kNoAllocSpan);
    if (ctas < 1) ctas = 1;
    pattern_check_kernel<<<ctas, kPatThreads, 0, s>>>((const uint8_t*)buf, bytes, 0, 0, scratch->partials, &scratch->ticket, &scratch->out);
    e = cudaGetLastError();
    if (e != cudaSuccess) return set_last_error(P2P_UNHANDLED_CUDA_ERROR, "p2p_digest_noalloc: launch failed: %s", cudaGetErrorString(e));
    e = cudaMemcpyAsync(digest_dev, &scratch->out.digest, sizeof(uint64_t), cudaMemcpyDeviceToDevice, s);
    if (e != cudaSuccess) return set_last_error(P2P_UNHANDLED_CUDA_ERROR, "p2p_digest_noalloc: result copy: %s", cudaGetErrorString(e));
    return P2P_SUCCESS;
}
