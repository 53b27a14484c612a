// TMA (bulk async copy) flavour of the movers, for large messages: pushes (local -> peer), pulls (peer -> local)
// and local copies all go global -> shared -> global through this ring.
//
// Per CTA: a ring of `stages` shared-memory buffers.  Warp 0's elected lane is the
// PRODUCER: cp.async.bulk global -> shared, completion counted on a `full` mbarrier
// (expect_tx).  Warp 1's elected lane is the CONSUMER: cp.async.bulk shared -> peer-mapped
// global (bulk_group), and once a group's shared-memory reads have drained
// (wait_group.read) it hands the stage back through an `empty` mbarrier.  No register
// staging, no LSU traffic for the payload; the SM only issues descriptors.
#pragma once
#include "p2p_device.cuh"

namespace p2p {

__device__ __forceinline__ uint32_t smem_u32(const void* p) {
    return (uint32_t)__cvta_generic_to_shared(p);
}
__device__ __forceinline__ void mbar_init(uint64_t* bar, uint32_t count) {
    asm volatile("mbarrier.init.shared::cta.b64 [%0], %1;" ::"r"(smem_u32(bar)), "r"(count) : "memory");
}
__device__ __forceinline__ void mbar_arrive_expect_tx(uint64_t* bar, uint32_t bytes) {
    asm volatile("mbarrier.arrive.expect_tx.shared::cta.b64 _, [%0], %1;" ::"r"(smem_u32(bar)), "r"(bytes)
                 : "memory");
}
__device__ __forceinline__ void mbar_arrive(uint64_t* bar) {
    asm volatile("mbarrier.arrive.shared::cta.b64 _, [%0];" ::"r"(smem_u32(bar)) : "memory");
}
__device__ __forceinline__ void mbar_wait(uint64_t* bar, uint32_t parity) {
    asm volatile(
        "{\n\t"
        ".reg .pred p;\n\t"
        "WAIT_%=:\n\t"
        "mbarrier.try_wait.parity.shared::cta.b64 p, [%0], %1;\n\t"
        "@p bra DONE_%=;\n\t"
        "bra WAIT_%=;\n\t"
        "DONE_%=:\n\t"
        "}" ::"r"(smem_u32(bar)), "r"(parity) : "memory");
}
__device__ __forceinline__ void bulk_g2s(void* smem_dst, const void* gsrc, uint32_t bytes, uint64_t* bar) {
    asm volatile(
        "cp.async.bulk.shared::cluster.global.mbarrier::complete_tx::bytes [%0], [%1], %2, [%3];"
        ::"r"(smem_u32(smem_dst)), "l"(gsrc), "r"(bytes), "r"(smem_u32(bar)) : "memory");
}
__device__ __forceinline__ void bulk_s2g(void* gdst, const void* smem_src, uint32_t bytes) {
    asm volatile("cp.async.bulk.global.shared::cta.bulk_group [%0], [%1], %2;"
                 ::"l"(gdst), "r"(smem_u32(smem_src)), "r"(bytes) : "memory");
}
__device__ __forceinline__ void bulk_commit() { asm volatile("cp.async.bulk.commit_group;" ::: "memory"); }
template <int N>
__device__ __forceinline__ void bulk_wait_read() {
    asm volatile("cp.async.bulk.wait_group.read %0;" ::"n"(N) : "memory");
}
template <int N>
__device__ __forceinline__ void bulk_wait_all() {
    asm volatile("cp.async.bulk.wait_group %0;" ::"n"(N) : "memory");
}
__device__ __forceinline__ void fence_proxy_async() { asm volatile("fence.proxy.async;" ::: "memory"); }

// Push `n` bytes src -> dst with this CTA as `part` of `nparts`.  dst/src share their
// 16-byte phase (checked by the caller); head and tail bytes go through the LSU.
// `wait_hdr` (may be null; pulls): griddepcontrol.wait is still owed.  The producer's bulk LOADS from the sender's
// buffer start at once and fill the ring while the kernel ahead on this stream retires; whatever STORES into this
// rank's buffer - the consumer, the byte head and tail - waits first.
__device__ __forceinline__ void tma_push(uint8_t* dst, const uint8_t* src, size_t n, uint32_t part,
                                         uint32_t nparts, uint32_t chunk, uint32_t stages,
                                         uint8_t* smem_ring, uint64_t* full, uint64_t* empty,
                                         const GroupHdr* wait_hdr = nullptr) {
    size_t head = (16 - ((uint64_t)dst & 15)) & 15;
    if (head > n) head = n;
    const size_t body = (n - head) & ~(size_t)15;
    const size_t tail = n - head - body;
    if (wait_hdr && threadIdx.x != 0) pdl_wait(*wait_hdr);   // the producer (thread 0) waits after its loop
    // byte head and tail: the upper half of the CTA, which drives nothing (the producer must not store before the wait)
    if (threadIdx.x >= 64) {
        const size_t lane = threadIdx.x - 64, step = blockDim.x - 64;
        if (part == 0)
            for (size_t i = lane; i < head; i += step) dst[i] = __ldcv(src + i);
        if (part == nparts - 1)
            for (size_t i = lane; i < tail; i += step) dst[head + body + i] = __ldcv(src + head + body + i);
    }

    const uint8_t* s = src + head;
    uint8_t* d = dst + head;
    const size_t nchunks = (body + chunk - 1) / chunk;
    // chunks part, part+nparts, ... belong to this CTA
    const size_t mine = nchunks > part ? (nchunks - part + nparts - 1) / nparts : 0;

    if (threadIdx.x == 0) {  // producer
        // the payload was written through the generic proxy (by this or a peer GPU) and made visible by the flag this
        // CTA acquired; the bulk loads below read it through the async proxy
        fence_proxy_async();
        for (size_t k = 0; k < mine; ++k) {
            const uint32_t st = (uint32_t)(k % stages);
            const uint32_t round = (uint32_t)(k / stages);
            if (round > 0) mbar_wait(&empty[st], (round - 1) & 1);
            const size_t off = (part + k * nparts) * (size_t)chunk;
            const uint32_t bytes = (uint32_t)(body - off < chunk ? body - off : chunk);
            mbar_arrive_expect_tx(&full[st], bytes);
            bulk_g2s(smem_ring + (size_t)st * chunk, s + off, bytes, &full[st]);
        }
        if (wait_hdr) pdl_wait(*wait_hdr);
    } else if (threadIdx.x == 32) {  // consumer
        for (size_t k = 0; k < mine; ++k) {
            const uint32_t st = (uint32_t)(k % stages);
            const uint32_t round = (uint32_t)(k / stages);
            mbar_wait(&full[st], round & 1);
            fence_proxy_async();  // stage was filled by the async proxy; keep its reader behind that
            const size_t off = (part + k * nparts) * (size_t)chunk;
            const uint32_t bytes = (uint32_t)(body - off < chunk ? body - off : chunk);
            bulk_s2g(d + off, smem_ring + (size_t)st * chunk, bytes);
            bulk_commit();
            if (k > 0) {
                bulk_wait_read<1>();  // group k-1 no longer reads its stage
                mbar_arrive(&empty[(uint32_t)((k - 1) % stages)]);
            }
        }
        bulk_wait_all<0>();   // every bulk store of this CTA has been performed
        fence_proxy_async();  // order them before the generic-proxy release that follows
    }
}

}  // namespace p2p
