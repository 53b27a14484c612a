// Symmetric windows on CUDA's virtual memory management API, and the NVSwitch multicast broadcast built on them
// (SURVEY section 8, rows f2 and f3).  C ABI in include/p2p.h.
//
// A symmetric window is `bytes` of device memory on EVERY rank, allocated with cuMemCreate, exported as a POSIX
// file descriptor (passed between processes over a Unix-domain socket, p2p_fdpass.cu) or as a fabric handle
// (CU_MEM_HANDLE_TYPE_FABRIC, a plain 64-byte blob that also crosses hosts inside one NVLink domain), imported and
// mapped by every peer.  It is registered with the engine like a cudaIpc window, so p2p_send/p2p_recv use it as
// usual.  Where the fabric supports it (CU_DEVICE_ATTRIBUTE_MULTICAST_SUPPORTED) rank 0 also creates a multicast
// object over all ranks' devices and every rank binds its part: one `multimem.st` to the multicast mapping then
// lands in every rank's copy, replicated inside the switch — the sender's link carries the payload once instead of
// once per receiver.  The reference can only do one-to-many as N-1 serial ncclSend calls (p2p_matrix.cc:141-172).
#include "p2p.h"
#include "p2p_internal.h"

#include <cstdio>
#include <cstdlib>
#include <cstring>
#include <mutex>
#include <vector>
#include <cuda.h>
#include <unistd.h>

namespace p2p {
namespace {

// ------------------------------------------------------------------ driver entry points (no link against libcuda)

struct Driver {
    decltype(&cuDeviceGet) DeviceGet = nullptr;
    decltype(&cuDeviceGetAttribute) DeviceGetAttribute = nullptr;
    decltype(&cuMemCreate) MemCreate = nullptr;
    decltype(&cuMemRelease) MemRelease = nullptr;
    decltype(&cuMemExportToShareableHandle) MemExport = nullptr;
    decltype(&cuMemImportFromShareableHandle) MemImport = nullptr;
    decltype(&cuMemAddressReserve) AddressReserve = nullptr;
    decltype(&cuMemAddressFree) AddressFree = nullptr;
    decltype(&cuMemMap) MemMap = nullptr;
    decltype(&cuMemUnmap) MemUnmap = nullptr;
    decltype(&cuMemSetAccess) SetAccess = nullptr;
    decltype(&cuMemGetAllocationGranularity) AllocGranularity = nullptr;
    decltype(&cuMulticastCreate) McCreate = nullptr;
    decltype(&cuMulticastAddDevice) McAddDevice = nullptr;
    decltype(&cuMulticastBindMem) McBindMem = nullptr;
    decltype(&cuMulticastUnbind) McUnbind = nullptr;
    decltype(&cuMulticastGetGranularity) McGranularity = nullptr;
    int state = 0;  // 0 not tried, 1 ok, -1 missing
};
Driver drv;

template <class F>
bool resolve(const char* name, F* fn) {
    void* p = nullptr;
    cudaDriverEntryPointQueryResult q;
    if (cudaGetDriverEntryPoint(name, &p, cudaEnableDefault, &q) != cudaSuccess || q != cudaDriverEntryPointSuccess || !p) return false;
    *fn = (F)p;
    return true;
}

bool driver_ready() {
    if (drv.state) return drv.state > 0;
    drv.state = -1;
    bool ok = resolve("cuDeviceGet", &drv.DeviceGet) && resolve("cuDeviceGetAttribute", &drv.DeviceGetAttribute) &&
              resolve("cuMemCreate", &drv.MemCreate) && resolve("cuMemRelease", &drv.MemRelease) &&
              resolve("cuMemExportToShareableHandle", &drv.MemExport) && resolve("cuMemImportFromShareableHandle", &drv.MemImport) &&
              resolve("cuMemAddressReserve", &drv.AddressReserve) && resolve("cuMemAddressFree", &drv.AddressFree) &&
              resolve("cuMemMap", &drv.MemMap) && resolve("cuMemUnmap", &drv.MemUnmap) && resolve("cuMemSetAccess", &drv.SetAccess) &&
              resolve("cuMemGetAllocationGranularity", &drv.AllocGranularity);
    if (!ok) return false;
    // multicast entry points may be missing on an old driver: the windows still work, the broadcast then uses unicast stores
    if (!(resolve("cuMulticastCreate", &drv.McCreate) && resolve("cuMulticastAddDevice", &drv.McAddDevice) &&
          resolve("cuMulticastBindMem", &drv.McBindMem) && resolve("cuMulticastUnbind", &drv.McUnbind) &&
          resolve("cuMulticastGetGranularity", &drv.McGranularity)))
        drv.McCreate = nullptr;
    drv.state = 1;
    return true;
}

#define CU_TRY(expr)                                                                                        \
    do {                                                                                                    \
        CUresult r_ = (expr);                                                                               \
        if (r_ != CUDA_SUCCESS) return set_last_error(P2P_UNHANDLED_CUDA_ERROR, "%s failed with CUresult %d (%s:%d)", #expr, (int)r_, __FILE__, __LINE__); \
    } while (0)
#define RT_TRY(expr)                                                                                        \
    do {                                                                                                    \
        cudaError_t e_ = (expr);                                                                            \
        if (e_ != cudaSuccess) return set_last_error(P2P_UNHANDLED_CUDA_ERROR, "%s failed: %s (%s:%d)", #expr, cudaGetErrorString(e_), __FILE__, __LINE__); \
    } while (0)

// ------------------------------------------------------------------ state

constexpr uint32_t kSymmMagic = 0x53593250u;  // "SY2P"
constexpr size_t kSymmBlobBytes = 256;
constexpr size_t kTrailerBytes = 4096;
constexpr int kMaxSymm = 16;

struct Trailer {                       // last 4 KiB of every rank's part of a symmetric window
    uint64_t done[P2P_MAX_RANKS];      // [root]: written by that root (multicast or unicast) when its broadcast has landed
    uint64_t ready[P2P_MAX_RANKS];     // [peer]: on the ROOT's copy, written by that peer: "my buffer may be overwritten"
};

struct SymmBlob {
    uint32_t magic;
    int32_t rank, pid, handle_type;    // 1 POSIX fd, 2 fabric
    uint64_t bytes, alloc_bytes;
    int32_t has_mc, pad;
    char socket_name[40];              // POSIX fd: where to ask for the descriptors
    unsigned char mem_fabric[64];
    unsigned char mc_fabric[64];
};
static_assert(sizeof(SymmBlob) <= kSymmBlobBytes, "symmetric-window blob too large");

struct SymmWindow {
    bool used = false, importing = false;
    int engine_window = -1;
    size_t bytes = 0, alloc_bytes = 0, gran = 0;
    int handle_type = 0;
    CUmemGenericAllocationHandle mem[P2P_MAX_RANKS] = {};   // [me] created here, others imported
    CUdeviceptr va[P2P_MAX_RANKS] = {};                     // unicast mappings of every rank's part
    CUmemGenericAllocationHandle mc = 0;
    CUdeviceptr mc_va = 0;
    bool mc_bound = false;
    int mem_fd = -1, mc_fd = -1;
    FdServer* server = nullptr;
    uint32_t* arrive = nullptr;                             // local counter of the broadcast kernel's CTAs
    uint64_t seq[P2P_MAX_RANKS] = {};                       // broadcasts so far, per root
};

SymmWindow symm[kMaxSymm];
SymmWindow* pending = nullptr;  // between export and import
std::mutex symm_mu;

int device_attr(int device, CUdevice_attribute a) {
    CUdevice d;
    int v = 0;
    if (drv.DeviceGet(&d, device) != CUDA_SUCCESS || drv.DeviceGetAttribute(&v, a, d) != CUDA_SUCCESS) return 0;
    return v;
}

CUmemAccessDesc rw_access(int device) {
    CUmemAccessDesc d = {};
    d.location.type = CU_MEM_LOCATION_TYPE_DEVICE;
    d.location.id = device;
    d.flags = CU_MEM_ACCESS_FLAGS_PROT_READWRITE;
    return d;
}

void release_window(SymmWindow& W, int nranks) {
    if (W.server) fdpass_finish(W.server);
    W.server = nullptr;
    if (W.mc_va) { drv.MemUnmap(W.mc_va, W.alloc_bytes); drv.AddressFree(W.mc_va, W.alloc_bytes); }
    for (int r = 0; r < nranks; ++r)
        if (W.va[r]) { drv.MemUnmap(W.va[r], W.alloc_bytes); drv.AddressFree(W.va[r], W.alloc_bytes); }
    for (int r = 0; r < nranks; ++r)
        if (W.mem[r]) drv.MemRelease(W.mem[r]);
    if (W.mc) drv.MemRelease(W.mc);
    if (W.mem_fd >= 0) close(W.mem_fd);
    if (W.mc_fd >= 0) close(W.mc_fd);
    if (W.arrive) cudaFree(W.arrive);
    W = SymmWindow();
}

// ------------------------------------------------------------------ the broadcast kernels

struct BcastArgs {
    const uint8_t* src;                  // root's payload (any local device memory)
    uint8_t* mc_dst;                     // multicast address of the destination, or nullptr: unicast stores
    uint8_t* uc_dst[P2P_MAX_RANKS];      // every rank's destination through its unicast mapping
    uint64_t* mc_done;                   // multicast address of trailer.done[root], or nullptr
    uint64_t* uc_done[P2P_MAX_RANKS];
    const uint64_t* ready;               // root's own trailer.ready[]
    uint32_t* arrive;
    uint32_t* async_err;
    uint64_t nbytes, seq, timeout_ns;
    int32_t nranks, me, in_place, pad;
};

__device__ __forceinline__ void multimem_st16(void* p, const uint4& v) {
    asm volatile("multimem.st.relaxed.sys.global.v4.f32 [%0], {%1,%2,%3,%4};" ::"l"(p), "r"(v.x), "r"(v.y), "r"(v.z), "r"(v.w) : "memory");
}
__device__ __forceinline__ void multimem_st_release_u64(uint64_t* p, uint64_t v) {
    asm volatile("multimem.st.release.sys.global.u64 [%0], %1;" ::"l"(p), "l"(v) : "memory");
}
struct StMc16 { __device__ __forceinline__ void operator()(uint4* p, const uint4& v) const { multimem_st16(p, v); } };

// Root: wait until every receiver has said its buffer may be overwritten, store the payload once into the
// multicast mapping (the switch replicates it into every rank's copy, the root's own included), publish `done`
// the same way.  Without multicast: one pass of unicast stores per receiver.
__global__ void __launch_bounds__(kLdstThreads, 1) p2p_bcast_root_kernel(const __grid_constant__ BcastArgs a) {
    __shared__ uint32_t failed;
    if (threadIdx.x == 0) failed = 0;
    __syncthreads();
    if (threadIdx.x < (uint32_t)a.nranks && (int)threadIdx.x != a.me) {
        uint64_t v;
        const uint64_t want = a.seq;
        if (!spin_flag(a.ready + threadIdx.x, a.timeout_ns, v, [want](uint64_t x) { return x >= want; })) {
            if (a.async_err) atomicCAS_system(a.async_err, 0u, (uint32_t)AERR_TIMEOUT_POST);
            failed = 1;
        }
    }
    __syncthreads();
    if (failed) return;
    const uint32_t part = blockIdx.x, nparts = gridDim.x;
    if (a.mc_dst && ((((uint64_t)a.mc_dst) | ((uint64_t)a.src)) & 15) == 0) {
        const size_t nvec = a.nbytes / 16;
        copy_vectors<uint4, 8>((uint4*)a.mc_dst, (const uint4*)a.src, nvec, part, nparts, Ld16{}, StMc16{});
        // multimem.st has no byte form: the last nbytes % 16 bytes go to every rank through its unicast mapping
        const size_t tail0 = nvec * 16;
        if (part == nparts - 1 && threadIdx.x < a.nbytes - tail0)
            for (int r = 0; r < a.nranks; ++r)
                if (!(a.in_place && r == a.me)) a.uc_dst[r][tail0 + threadIdx.x] = a.src[tail0 + threadIdx.x];
    } else {
        for (int r = 0; r < a.nranks; ++r)
            if (!(a.in_place && r == a.me)) copy_bytes(a.uc_dst[r], a.src, a.nbytes, part, nparts);
    }
    __syncthreads();
    if (threadIdx.x == 0) {
        fence_sys();
        const uint32_t old = atomicInc(a.arrive, nparts - 1);
        if (old == nparts - 1) {
            if (a.mc_done) {
                multimem_st_release_u64(a.mc_done, a.seq);
            } else {
                for (int r = 0; r < a.nranks; ++r) st_release_sys(a.uc_done[r], a.seq);
            }
        }
    }
}

// Everyone else: tell the root this rank's buffer may be overwritten, then hold the stream until the payload is there.
__global__ void p2p_bcast_recv_kernel(uint64_t* ready_at_root, const uint64_t* done, uint64_t seq, uint64_t timeout_ns, uint32_t* async_err) {
    if (threadIdx.x != 0) return;
    st_release_sys(ready_at_root, seq);
    uint64_t v;
    if (!spin_flag(done, timeout_ns, v, [seq](uint64_t x) { return x >= seq; }) && async_err)
        atomicCAS_system(async_err, 0u, (uint32_t)AERR_TIMEOUT_DONE);
}

SymmWindow* window_of(const void* p, size_t nbytes, int me, size_t* off) {
    for (SymmWindow& W : symm) {
        if (!W.used || W.importing) continue;
        const uint8_t* base = (const uint8_t*)W.va[me];
        if ((const uint8_t*)p >= base && (const uint8_t*)p + nbytes <= base + W.bytes) {
            *off = (size_t)((const uint8_t*)p - base);
            return &W;
        }
    }
    return nullptr;
}

}  // namespace
}  // namespace p2p

using namespace p2p;

extern "C" {

size_t p2p_symm_blob_bytes(void) { return kSymmBlobBytes; }

int p2p_symm_alloc_export(size_t bytes, void* blob) {
    EngineInfo e;
    int rc = engine_info(&e);
    if (rc) return rc;
    if (!blob || bytes == 0) return set_last_error(P2P_INVALID_ARGUMENT, "p2p_symm_alloc_export: null blob or zero bytes");
    std::lock_guard<std::mutex> lk(symm_mu);
    if (!driver_ready()) return set_last_error(P2P_SYSTEM_ERROR, "the CUDA driver lacks the virtual memory management entry points");
    if (pending) return set_last_error(P2P_INVALID_USAGE, "p2p_symm_alloc_export called twice without p2p_symm_alloc_import");
    RT_TRY(cudaSetDevice(e.device));
    if (!device_attr(e.device, CU_DEVICE_ATTRIBUTE_VIRTUAL_MEMORY_MANAGEMENT_SUPPORTED))
        return set_last_error(P2P_SYSTEM_ERROR, "device %d does not support virtual memory management", e.device);
    SymmWindow* W = nullptr;
    for (SymmWindow& c : symm)
        if (!c.used) { W = &c; break; }
    if (!W) return set_last_error(P2P_INVALID_USAGE, "all %d symmetric windows are in use", kMaxSymm);

    // handle type: fabric handles cross process (and host) boundaries as plain bytes; POSIX descriptors need the
    // socket hand-over but work everywhere on one host.  P2P_SYMM_HANDLE=fabric|fd chooses; default fd.
    const char* want = getenv("P2P_SYMM_HANDLE");
    int type = (want && !strcmp(want, "fabric")) ? 2 : 1;
    if (type == 2 && !device_attr(e.device, CU_DEVICE_ATTRIBUTE_HANDLE_TYPE_FABRIC_SUPPORTED))
        return set_last_error(P2P_SYSTEM_ERROR, "P2P_SYMM_HANDLE=fabric, but device %d reports no fabric-handle support", e.device);
    if (type == 1 && !device_attr(e.device, CU_DEVICE_ATTRIBUTE_HANDLE_TYPE_POSIX_FILE_DESCRIPTOR_SUPPORTED))
        return set_last_error(P2P_SYSTEM_ERROR, "device %d cannot export memory as POSIX file descriptors", e.device);
    const CUmemAllocationHandleType htype = type == 2 ? CU_MEM_HANDLE_TYPE_FABRIC : CU_MEM_HANDLE_TYPE_POSIX_FILE_DESCRIPTOR;

    CUmemAllocationProp prop = {};
    prop.type = CU_MEM_ALLOCATION_TYPE_PINNED;
    prop.location.type = CU_MEM_LOCATION_TYPE_DEVICE;
    prop.location.id = e.device;
    prop.requestedHandleTypes = htype;
    size_t gran = 0;
    CU_TRY(drv.AllocGranularity(&gran, &prop, CU_MEM_ALLOC_GRANULARITY_RECOMMENDED));
    const bool want_mc = drv.McCreate && device_attr(e.device, CU_DEVICE_ATTRIBUTE_MULTICAST_SUPPORTED) && !getenv("P2P_NO_MULTICAST");
    CUmulticastObjectProp mp = {};
    if (want_mc) {
        mp.numDevices = (unsigned)e.nranks;
        mp.handleTypes = htype;
        mp.size = bytes + kTrailerBytes;
        size_t mg = 0;
        if (drv.McGranularity(&mg, &mp, CU_MULTICAST_GRANULARITY_MINIMUM) == CUDA_SUCCESS && mg > gran) gran = mg;
    }
    if (gran == 0) gran = 2u << 20;
    *W = SymmWindow();
    W->bytes = bytes;
    W->gran = gran;
    W->alloc_bytes = (bytes + kTrailerBytes + gran - 1) / gran * gran;
    W->handle_type = type;

    SymmBlob b;
    memset(&b, 0, sizeof b);
    b.magic = kSymmMagic;
    b.rank = e.rank;
    b.pid = (int32_t)getpid();
    b.handle_type = type;
    b.bytes = bytes;
    b.alloc_bytes = W->alloc_bytes;
    CUresult cr = drv.MemCreate(&W->mem[e.rank], W->alloc_bytes, &prop, 0);
    if (cr != CUDA_SUCCESS) {
        *W = SymmWindow();
        return set_last_error(P2P_UNHANDLED_CUDA_ERROR, "cuMemCreate of %zu bytes (handle type %s) failed with CUresult %d", b.alloc_bytes,
                              type == 2 ? "fabric" : "POSIX fd", (int)cr);
    }
    if (type == 2) cr = drv.MemExport(b.mem_fabric, W->mem[e.rank], htype, 0);
    else cr = drv.MemExport(&W->mem_fd, W->mem[e.rank], htype, 0);
    if (cr != CUDA_SUCCESS) {
        release_window(*W, e.nranks);
        return set_last_error(P2P_UNHANDLED_CUDA_ERROR, "cuMemExportToShareableHandle (%s) failed with CUresult %d", type == 2 ? "fabric" : "POSIX fd", (int)cr);
    }
    // this rank's own part is mapped and cleared NOW: once the blob is out, a peer may write flags into it
    {
        const CUmemAccessDesc access = rw_access(e.device);
        cr = drv.AddressReserve(&W->va[e.rank], W->alloc_bytes, gran, 0, 0);
        if (cr == CUDA_SUCCESS) cr = drv.MemMap(W->va[e.rank], W->alloc_bytes, 0, W->mem[e.rank], 0);
        if (cr == CUDA_SUCCESS) cr = drv.SetAccess(W->va[e.rank], W->alloc_bytes, &access, 1);
        cudaError_t ce = cr == CUDA_SUCCESS ? cudaMemset((void*)W->va[e.rank], 0, W->alloc_bytes) : cudaErrorUnknown;
        if (ce == cudaSuccess) ce = cudaDeviceSynchronize();
        if (cr != CUDA_SUCCESS || ce != cudaSuccess) {
            release_window(*W, e.nranks);
            return set_last_error(P2P_UNHANDLED_CUDA_ERROR, "mapping and clearing this rank's part of the window failed (CUresult %d, %s)", (int)cr,
                                  cudaGetErrorString(ce));
        }
    }
    if (want_mc && e.rank == 0) {
        mp.size = W->alloc_bytes;
        if (drv.McCreate(&W->mc, &mp) == CUDA_SUCCESS) {
            cr = type == 2 ? drv.MemExport(b.mc_fabric, W->mc, htype, 0) : drv.MemExport(&W->mc_fd, W->mc, htype, 0);
            if (cr == CUDA_SUCCESS) b.has_mc = 1;
            else { drv.MemRelease(W->mc); W->mc = 0; }
        } else {
            W->mc = 0;  // no multicast object to be had (fabric manager, resources): the broadcast will use unicast stores
        }
    }
    if (type == 1 && e.nranks > 1) {
        const int fds[2] = {W->mem_fd, W->mc_fd};
        const int requests = (e.nranks - 1) * (b.has_mc ? 2 : 1);
        W->server = fdpass_serve(fds, b.has_mc ? 2 : 1, requests, 60000, b.socket_name);
        if (!W->server) {
            release_window(*W, e.nranks);
            return set_last_error(P2P_SYSTEM_ERROR, "cannot open the Unix-domain socket that hands the memory's file descriptor to the peers");
        }
    }
    W->used = W->importing = true;
    pending = W;
    memset(blob, 0, kSymmBlobBytes);
    memcpy(blob, &b, sizeof b);
    return P2P_SUCCESS;
}

int p2p_symm_alloc_import(const void* all_blobs, void** ptr, int* window_id) {
    EngineInfo e;
    int rc = engine_info(&e);
    if (rc) return rc;
    if (!all_blobs || !ptr || !window_id) return set_last_error(P2P_INVALID_ARGUMENT, "p2p_symm_alloc_import: null argument");
    std::lock_guard<std::mutex> lk(symm_mu);
    SymmWindow* W = pending;
    if (!W) return set_last_error(P2P_INVALID_USAGE, "p2p_symm_alloc_import without p2p_symm_alloc_export");
    pending = nullptr;
    auto fail_out = [&](int code) { release_window(*W, e.nranks); return code; };
    RT_TRY(cudaSetDevice(e.device));
    const CUmemAllocationHandleType htype = W->handle_type == 2 ? CU_MEM_HANDLE_TYPE_FABRIC : CU_MEM_HANDLE_TYPE_POSIX_FILE_DESCRIPTOR;
    std::vector<SymmBlob> blobs((size_t)e.nranks);
    for (int r = 0; r < e.nranks; ++r) {
        memcpy(&blobs[(size_t)r], (const uint8_t*)all_blobs + (size_t)r * kSymmBlobBytes, sizeof(SymmBlob));
        const SymmBlob& b = blobs[(size_t)r];
        if (b.magic != kSymmMagic || b.rank != r)
            return fail_out(set_last_error(P2P_INVALID_ARGUMENT, "symmetric-window blob %d is malformed or out of rank order", r));
        if (b.bytes != W->bytes || b.alloc_bytes != W->alloc_bytes || b.handle_type != W->handle_type)
            return fail_out(set_last_error(P2P_INVALID_USAGE, "rank %d allocates %llu bytes (handle type %d), this rank %zu (type %d): a symmetric "
                                           "window has one size and one handle type on all ranks", r, (unsigned long long)b.bytes, b.handle_type, W->bytes, W->handle_type));
    }
    // every peer's part: import, map, allow this device to read and write it
    const CUmemAccessDesc access = rw_access(e.device);
    for (int r = 0; r < e.nranks; ++r) {
        if (r != e.rank) {
            CUresult cr;
            if (W->handle_type == 2) {
                cr = drv.MemImport(&W->mem[r], (void*)blobs[(size_t)r].mem_fabric, htype);
            } else {
                const int fd = fdpass_fetch(blobs[(size_t)r].socket_name, 0, 30000);
                if (fd < 0) return fail_out(set_last_error(P2P_SYSTEM_ERROR, "rank %d (pid %d) did not hand over its memory's file descriptor", r, blobs[(size_t)r].pid));
                cr = drv.MemImport(&W->mem[r], (void*)(uintptr_t)fd, htype);
                close(fd);
            }
            if (cr != CUDA_SUCCESS) return fail_out(set_last_error(P2P_UNHANDLED_CUDA_ERROR, "cuMemImportFromShareableHandle for rank %d failed with CUresult %d", r, (int)cr));
        }
        if (r == e.rank) continue;  // mapped and cleared at export time
        if (drv.AddressReserve(&W->va[r], W->alloc_bytes, W->gran, 0, 0) != CUDA_SUCCESS)
            return fail_out(set_last_error(P2P_UNHANDLED_CUDA_ERROR, "cuMemAddressReserve of %zu bytes failed", W->alloc_bytes));
        CUresult cr = drv.MemMap(W->va[r], W->alloc_bytes, 0, W->mem[r], 0);
        if (cr == CUDA_SUCCESS) cr = drv.SetAccess(W->va[r], W->alloc_bytes, &access, 1);
        if (cr != CUDA_SUCCESS) return fail_out(set_last_error(P2P_UNHANDLED_CUDA_ERROR, "mapping rank %d's part of the window failed with CUresult %d", r, (int)cr));
    }
    // the multicast object: import, join with this device, bind this rank's part (the bind waits until every rank
    // has joined), map.  Whether there is a multicast path was decided by rank 0 at export time for all ranks alike; a
    // rank that cannot join it fails its import, because a root that stores through the multicast mapping would
    // never reach this rank's copy.
    if (blobs[0].has_mc) {
        CUresult cr = CUDA_SUCCESS;
        if (e.rank != 0) {
            if (W->handle_type == 2) {
                cr = drv.MemImport(&W->mc, (void*)blobs[0].mc_fabric, htype);
            } else {
                const int fd = fdpass_fetch(blobs[0].socket_name, 1, 30000);
                cr = fd >= 0 ? drv.MemImport(&W->mc, (void*)(uintptr_t)fd, htype) : CUDA_ERROR_UNKNOWN;
                if (fd >= 0) close(fd);
            }
        }
        CUdevice dev;
        if (cr == CUDA_SUCCESS) cr = drv.DeviceGet(&dev, e.device);
        if (cr == CUDA_SUCCESS) cr = drv.McAddDevice(W->mc, dev);
        if (cr == CUDA_SUCCESS) cr = drv.McBindMem(W->mc, 0, W->mem[e.rank], 0, W->alloc_bytes, 0);
        if (cr == CUDA_SUCCESS) {
            W->mc_bound = true;
            cr = drv.AddressReserve(&W->mc_va, W->alloc_bytes, W->gran, 0, 0);
            if (cr == CUDA_SUCCESS) cr = drv.MemMap(W->mc_va, W->alloc_bytes, 0, W->mc, 0);
            if (cr == CUDA_SUCCESS) cr = drv.SetAccess(W->mc_va, W->alloc_bytes, &access, 1);
        }
        if (cr != CUDA_SUCCESS)
            return fail_out(set_last_error(P2P_UNHANDLED_CUDA_ERROR, "joining the multicast object failed with CUresult %d (set P2P_NO_MULTICAST=1 on all "
                                           "ranks to do without it)", (int)cr));
    }
    if (W->server) { fdpass_finish(W->server); W->server = nullptr; }
    if (W->mem_fd >= 0) { close(W->mem_fd); W->mem_fd = -1; }
    if (W->mc_fd >= 0) { close(W->mc_fd); W->mc_fd = -1; }
    cudaError_t ce = cudaMalloc(&W->arrive, sizeof(uint32_t));
    if (ce == cudaSuccess) ce = cudaMemset(W->arrive, 0, sizeof(uint32_t));
    if (ce == cudaSuccess) ce = cudaDeviceSynchronize();
    if (ce != cudaSuccess) return fail_out(set_last_error(P2P_UNHANDLED_CUDA_ERROR, "clearing the new window failed: %s", cudaGetErrorString(ce)));
    uint8_t* bases[P2P_MAX_RANKS] = {};
    for (int r = 0; r < e.nranks; ++r) bases[r] = (uint8_t*)W->va[r];
    rc = engine_add_window((uint8_t*)W->va[e.rank], W->bytes, bases, &W->engine_window);
    if (rc) return fail_out(rc);
    W->importing = false;
    *ptr = (void*)W->va[e.rank];
    *window_id = W->engine_window;
    return P2P_SUCCESS;
}

int p2p_symm_free(int window_id) {
    EngineInfo e;
    int rc = engine_info(&e);
    if (rc) return rc;
    std::lock_guard<std::mutex> lk(symm_mu);
    for (SymmWindow& W : symm) {
        if (!W.used || W.importing || W.engine_window != window_id) continue;
        RT_TRY(cudaSetDevice(e.device));
        RT_TRY(cudaDeviceSynchronize());
        rc = p2p_deregister(window_id);
        if (W.mc_bound) {
            CUdevice d = 0;
            if (drv.DeviceGet(&d, e.device) == CUDA_SUCCESS) drv.McUnbind(W.mc, d, 0, W.alloc_bytes);
        }
        release_window(W, e.nranks);
        return rc;
    }
    return set_last_error(P2P_INVALID_ARGUMENT, "window %d is not a symmetric window", window_id);
}

}  // extern "C"

// p2p_finalize: whatever symmetric windows are left go with the engine
void p2p::symm_finalize(int nranks) {
    std::lock_guard<std::mutex> lk(symm_mu);
    pending = nullptr;
    for (SymmWindow& W : symm)
        if (W.used) release_window(W, nranks);
}

extern "C" {

int p2p_symm_multicast_ptr(int window_id, void** mc_ptr) {
    if (!mc_ptr) return set_last_error(P2P_INVALID_ARGUMENT, "null argument");
    std::lock_guard<std::mutex> lk(symm_mu);
    for (SymmWindow& W : symm)
        if (W.used && !W.importing && W.engine_window == window_id) {
            *mc_ptr = (void*)W.mc_va;
            return P2P_SUCCESS;
        }
    return set_last_error(P2P_INVALID_ARGUMENT, "window %d is not a symmetric window", window_id);
}

int p2p_bcast(const void* sendbuf, void* recvbuf, size_t count, int dtype, int root, p2p_stream_t stream) {
    EngineInfo e;
    int rc = engine_info(&e);
    if (rc) return rc;
    static const size_t dtype_size[P2P_NUM_TYPES] = {1, 1, 4, 4, 8, 8, 2, 4, 8, 2, 1, 1};
    if (dtype < 0 || dtype >= P2P_NUM_TYPES) return set_last_error(P2P_INVALID_ARGUMENT, "unknown datatype %d", dtype);
    if (root < 0 || root >= e.nranks) return set_last_error(P2P_INVALID_ARGUMENT, "root %d out of range [0,%d)", root, e.nranks);
    if (count > SIZE_MAX / dtype_size[dtype]) return set_last_error(P2P_INVALID_ARGUMENT, "count %zu overflows size_t", count);
    const size_t nbytes = count * dtype_size[dtype];
    if (nbytes == 0) return P2P_SUCCESS;
    if (!recvbuf || (e.rank == root && !sendbuf)) return set_last_error(P2P_INVALID_ARGUMENT, "null buffer with count %zu", count);
    std::lock_guard<std::mutex> lk(symm_mu);
    size_t off = 0;
    SymmWindow* W = window_of(recvbuf, nbytes, e.rank, &off);
    if (!W)
        return set_last_error(P2P_INVALID_USAGE, "p2p_bcast: receive buffer %p (+%zu bytes) is not inside a symmetric window "
                              "(p2p_symm_alloc_export/import); every rank passes the same offset of the same window", recvbuf, nbytes);
    const uint64_t seq = ++W->seq[root];
    const size_t trailer = W->alloc_bytes - kTrailerBytes;
    cudaStream_t s = (cudaStream_t)stream;
    if (e.rank != root) {
        Trailer* at_root = (Trailer*)((uint8_t*)W->va[root] + trailer);
        const Trailer* mine = (const Trailer*)((uint8_t*)W->va[e.rank] + trailer);
        p2p_bcast_recv_kernel<<<1, 32, 0, s>>>(&at_root->ready[e.rank], &mine->done[root], seq, e.timeout_ns, e.async_err);
        RT_TRY(cudaGetLastError());
        return P2P_SUCCESS;
    }
    BcastArgs a;
    memset(&a, 0, sizeof a);
    a.src = (const uint8_t*)sendbuf;
    a.nbytes = nbytes;
    a.seq = seq;
    a.timeout_ns = e.timeout_ns;
    a.async_err = e.async_err;
    a.nranks = e.nranks;
    a.me = e.rank;
    a.in_place = sendbuf == recvbuf;
    a.arrive = W->arrive;
    a.ready = ((const Trailer*)((uint8_t*)W->va[e.rank] + trailer))->ready;
    for (int r = 0; r < e.nranks; ++r) {
        a.uc_dst[r] = (uint8_t*)W->va[r] + off;
        a.uc_done[r] = &((Trailer*)((uint8_t*)W->va[r] + trailer))->done[root];
    }
    if (W->mc_va && !getenv("P2P_BCAST_UNICAST")) {
        a.mc_dst = (uint8_t*)W->mc_va + off;
        a.mc_done = &((Trailer*)((uint8_t*)W->mc_va + trailer))->done[root];
    }
    // 16 CTAs already saturate SM-issued NVLink stores; more help only the local read
    const size_t tile = (size_t)kLdstThreads * 128;
    long long ctas = (long long)((nbytes + tile - 1) / tile);
    const char* env = getenv("P2P_BCAST_NBLOCKS");
    const long long cap = env && atoll(env) > 0 ? atoll(env) : 32;
    if (ctas > cap) ctas = cap;
    if (ctas < 1) ctas = 1;
    p2p_bcast_root_kernel<<<(unsigned)ctas, kLdstThreads, 0, s>>>(a);
    RT_TRY(cudaGetLastError());
    return P2P_SUCCESS;
}

}  // extern "C"
