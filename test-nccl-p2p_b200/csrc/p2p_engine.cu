// Host side of the transfer engine: bootstrap, window registration, the group builder that
// turns p2p_send/p2p_recv calls into one fused kernel launch, plan selection, error reporting.
// C ABI in include/p2p.h; reference call sites it stands in for are cited there.
#include "p2p.h"
#include "p2p_debug.h"
#include "p2p_internal.h"

#include <algorithm>
#include <array>
#include <cstdarg>
#include <cstdio>
#include <cstdlib>
#include <cerrno>
#include <cstring>
#include <map>
#include <mutex>
#include <string>
#include <vector>
#include <atomic>
#include <chrono>
#include <fcntl.h>
#include <sys/mman.h>
#include <sys/stat.h>
#include <unistd.h>

namespace p2p {
namespace {

constexpr uint32_t kBlobMagic = 0x42325032u;  // "B2P2"
constexpr uint32_t kBlobVersion = 1;
constexpr size_t kBlobBytes = 256;
constexpr size_t kRegBlobBytes = 128;

struct BootBlob {
    uint32_t magic, version;
    int32_t rank, nranks;
    uint64_t host_hash;
    int32_t pid, device;
    uint64_t eager_bytes;
    int64_t pull_min_bytes;
    uint64_t ctl_bytes;
    char pci_bus_id[32];
    char hostbox_name[48];  // rank 0 only: POSIX shared-memory name of the host mailbox
    cudaIpcMemHandle_t ctl_handle;
    uint32_t stage_chunk, stage_lanes;  // shape of the staging ring, which both ends of a pair walk
};
static_assert(sizeof(BootBlob) <= kBlobBytes, "bootstrap blob too large");

struct RegBlob {
    uint32_t magic;
    int32_t has;
    uint64_t bytes;
    uint64_t offset;  // of the window inside its cudaMalloc allocation
    cudaIpcMemHandle_t handle;
};
static_assert(sizeof(RegBlob) <= kRegBlobBytes, "register blob too large");

// Section of a rank's control region that belongs to one peer.  Who writes what:
//   post[]   the peer, as receiver of my sends   (I poll it as sender)
//   done[]   the peer, as sender to me           (I poll it as receiver)
//   ll_ack[] the peer, as eager receiver         (I poll it as eager sender)
//   pull_post[] the peer, as sender of a large message to me: where its payload lies / "I will push"
//   pull_done[] the peer, as receiver that pulled from me   (I poll it as sender)
struct PeerCtl {
    uint64_t post[kRing];
    uint64_t done[kRing];
    uint64_t pull_post[kRing];
    uint64_t pull_done[kRing];
    uint64_t ll_ack[kLLSlots];
    uint64_t pad[8];
};

// Host mailbox (POSIX shared memory, one per job): the receiver's HOST publishes where a large receive goes at the
// moment p2p_recv is called, so that the sender's host can address a DMA copy.  It carries no ordering: readiness of
// the buffer is still signalled on the device.  Entry [dst][src][seq % kRing], seqlock-protected.
struct HostPost {
    std::atomic<uint64_t> seq;
    std::atomic<uint64_t> word;
};

struct Window {
    bool used = false;
    uint8_t* base = nullptr;  // local buffer (may be null if this rank contributed nothing)
    size_t bytes = 0;
    uint8_t* peer_base[P2P_MAX_RANKS] = {};  // mapped address of every rank's part of this window
    std::string peer_handle[P2P_MAX_RANKS];  // key into ipc_cache for unmapping
};

struct IpcMapping {
    void* ptr = nullptr;
    int refs = 0;
};

struct PendingOp {
    bool is_send;
    const void* buf;
    size_t nbytes;
    int peer;
    cudaStream_t stream;
};

struct Options {
    long long path = 0;            // 0 auto, 1 ldst, 2 tma, 3 ce
    long long nblocks = 0;         // CTAs per push to a peer, 0 auto
    long long self_nblocks = 0;    // CTAs of a local copy, 0 auto
    long long pull_nblocks = 0;    // CTAs that pull a large message, 0 auto
    long long eager_nblocks = 0;   // most CTAs an eager message may use, 0 auto
    long long eager_bytes = 512 << 10; // pairs: the eager ring beats the push up to 1 MiB (8.4 vs 14.5 us); all-pairs on 8 GPUs
                                       // wants 1 MiB pushed (27.9 vs 41.9 us, NCCL 32.8): 512 KiB serves both
    long long pull_min_bytes = 4 << 20;  // messages at least this large let the receiver pull (-1: never); same on all ranks
    long long timeout_ms = 20000;
    long long tma_chunk = 32 * 1024;
    long long tma_stages = 6;
    long long tma_peer_chunk = 64 * 1024;  // ring geometry of launches that move bytes over NVLink: 3 x 64 KiB reaches the
    long long tma_peer_stages = 3;         // link rate with 32 CTAs (6 x 32 KiB, the best LOCAL copy, needs 96 for a push)
    long long pull_tma_min_bytes = 32 << 20;  // auto: a large receive at or above this, in a group that is also being pulled
                                              // from, goes through the TMA ring (-1 never); see resolve_protocols
    long long tma_min_bytes = -1;  // auto: pushes at or above this use the TMA flavour (-1 never)
    long long self_tma_min_bytes = 32 << 20;  // auto: local copies at or above this use the TMA flavour (-1 never)
    long long pdl = 3;             // programmatic dependent launch: 0 never, 1 eager-only and local-copy groups (where it
                                   // measured faster), 2 every group with the trigger at kernel entry (push/pull kernels
                                   // measured slower with it), 3 as 1 plus groups of large messages, whose mover CTAs
                                   // trigger after their copy loop
    long long early_post = 1;      // with pdl 3: a send offered for pulling is published before the kernel ahead of it on the
                                   // stream has retired, when that kernel is the engine's own and does not write the payload
    long long loopback = 0;        // 1: a send to oneself goes through the wire protocol instead of a plain copy (tests)
    long long bidir_push = 1;      // a lone large send pushes instead of offering a pull when its group also holds a large
                                   // receive: 0 never, 1 from that same peer (pairwise exchange), 2 from any peer (ring),
                                   // 3 as 2 and also with several large sends in the group (all-pairs)
    long long pull_total = 64;     // CTAs all pulls of one group share (each flow gets pull_total / flows, at least fan_min)
    long long push_total = 64;     // the same for pushes (each flow at most 32)
    long long fan_min = 8;
    long long ce_min_bytes = -1;   // experimental, off: a send that is alone in its group and at least this large goes through
                                   // the copy engine when the receiver's host has already said where to.  Measured on 2xB200
                                   // it loses to the pull kernel (256 MiB: 708 vs 726 GB/s streamed, 496 vs 368 us per
                                   // ping-pong hop): kernel -> DMA -> kernel on one stream costs more than the DMA gains.
    long long ce_wait_us = 200;    // how long p2p_send's host side waits for that word before using the SM path
    long long stage_chunk = 0;     // bytes per slot of the staging ring's lanes (receive buffers that are in no window): 32 lanes x
                                   // 4 slots x this per ordered pair of ranks; same on all ranks.  0: 256 KiB up to 8 ranks, less
                                   // beyond, so that a rank's rings stay within 256 MiB (one lane moves ~18 GB/s whatever the
                                   // chunk: 16 lanes 292 GB/s, 8 lanes 157; 128 KiB chunks 188 GB/s, 512 KiB 411)
    long long stage_lanes = kStageLanes;
    long long ce_flags = 1;        // who handles the flags around the DMA: 0 two one-thread kernels (round 1), 1 stream memory
                                   // operations (cuStreamWaitValue64 / cuStreamWriteValue64: nothing to launch, so the chain
                                   // may run beside a group's kernel), 2 as 1 but flags in PEER memory are written by 8-byte DMA
};

struct Engine {
    bool inited = false, connected = false;
    int rank = -1, nranks = 0, device = -1, sm_count = 0;
    int ctas_ldst = 2, ctas_tma = 1;
    uint64_t host_hash = 0;
    Options opt;

    size_t ctl_bytes = 0, ll_slot_bytes = 0, ll_off = 0, stage_off = 0;
    StageInfo* stage_table = nullptr;           // device [nranks], filled in p2p_connect
    uint8_t* ctl = nullptr;                     // my control region (exported)
    uint8_t* peer_ctl[P2P_MAX_RANKS] = {};      // everyone's, peer-mapped (mine == ctl)
    uint32_t* arrive = nullptr;                 // local arrival counters: push [nranks][kRing], pull [nranks][kRing], eager [nranks][kLLSlots]
    uint64_t* win_table = nullptr;              // [nranks][P2P_MAX_WINDOWS] device table of mapped bases
    CheckPartial* chk_partials = nullptr;       // scratch for pattern_check
    unsigned int* chk_ticket = nullptr;
    CheckPartial* chk_out = nullptr;
    CheckPartial* chk_out_host = nullptr;       // pinned
    uint32_t* async_err_host = nullptr;         // pinned + mapped
    uint32_t* async_err_dev = nullptr;

    uint64_t rdv_send_seq[P2P_MAX_RANKS] = {}, rdv_recv_seq[P2P_MAX_RANKS] = {};
    uint64_t ll_send_seq[P2P_MAX_RANKS] = {}, ll_recv_seq[P2P_MAX_RANKS] = {};

    Window windows[P2P_MAX_WINDOWS];
    std::map<std::string, IpcMapping> ipc_cache[P2P_MAX_RANKS];
    void* pending_reg_buf = nullptr;
    size_t pending_reg_bytes = 0;
    bool has_pending_reg = false;

    HostPost* hostbox = nullptr;  // [nranks][nranks][kRing]
    size_t hostbox_bytes = 0;
    char hostbox_name[48] = "";
    cudaStream_t ce_stream[P2P_MAX_RANKS] = {};
    cudaEvent_t ce_join[P2P_MAX_RANKS] = {};
    cudaEvent_t ce_fork = nullptr;
    long long stat_launches = 0, stat_ce_copies = 0;
    std::vector<std::array<cudaEvent_t, 3>> ce_timing;  // P2P_CE_TIMING: chain start, DMA start, DMA end
    uint64_t* ce_vals = nullptr;                // pinned host ring of flag values for ce_flags 2 (source of the 8-byte DMAs)
    uint32_t ce_vals_next = 0;
    // what the last launched group looked like (bench.py reports the kernel that actually ran, not a forecast)
    long long last_stat[12] = {};
    uint32_t* tokens = nullptr;                 // device u32[kTokenSlots], see publish_token()
    uint32_t launch_counter = 0;
    struct StreamLast {
        uint32_t token = 0;
        std::vector<std::pair<const uint8_t*, size_t>> writes;  // user memory that launch writes (its receives)
    };
    std::map<cudaStream_t, StreamLast> stream_last;

    // plan-only mode (p2p_debug.h): groups are resolved exactly as usual but captured instead of launched
    bool plan_only = false;
    GroupArgs last_args;
    uint32_t last_grid = 0, last_block = 0;
    int last_flavour = 0;

    int group_depth = 0;
    std::vector<PendingOp> group_ops;
    cudaEvent_t ev_done = nullptr;
    std::vector<cudaEvent_t> ev_pool;

    std::mutex mu;
};

Engine g;
thread_local char t_last_error[512] = "";

int fail(int code, const char* fmt, ...) {
    va_list ap;
    va_start(ap, fmt);
    vsnprintf(t_last_error, sizeof t_last_error, fmt, ap);
    va_end(ap);
    return code;
}

}  // namespace

int set_last_error(int code, const char* fmt, ...) {
    va_list ap;
    va_start(ap, fmt);
    vsnprintf(t_last_error, sizeof t_last_error, fmt, ap);
    va_end(ap);
    return code;
}

namespace {

#define CUDA_TRY(expr)                                                                          \
    do {                                                                                        \
        cudaError_t e_ = (expr);                                                                \
        if (e_ != cudaSuccess)                                                                  \
            return fail(P2P_UNHANDLED_CUDA_ERROR, "%s failed: %s (%s:%d)", #expr,               \
                        cudaGetErrorString(e_), __FILE__, __LINE__);                            \
    } while (0)

// DJB2a over the short host name, as the reference does (p2p_matrix.cc:44-61).
uint64_t host_hash_now() {
    char name[1024];
    if (gethostname(name, sizeof name) != 0) name[0] = '\0';
    name[sizeof name - 1] = '\0';
    uint64_t h = 5381;
    for (int i = 0; name[i] != '\0' && name[i] != '.'; ++i) h = ((h << 5) + h) ^ (uint64_t)name[i];
    return h;
}

const size_t kDtypeSize[P2P_NUM_TYPES] = {1, 1, 4, 4, 8, 8, 2, 4, 8, 2, 1, 1};

inline PeerCtl* peer_section(uint8_t* region, int p) { return (PeerCtl*)region + p; }
inline uint8_t* ll_slot(uint8_t* region, int from, uint64_t seq) {
    return region + g.ll_off + ((size_t)from * kLLSlots + (size_t)(seq % kLLSlots)) * g.ll_slot_bytes;
}

long long env_ll(const char* name, long long dflt) {
    const char* s = getenv(name);
    if (!s || !*s) return dflt;
    if (!strcmp(name, "P2P_PATH")) {
        if (!strcmp(s, "auto")) return 0;
        if (!strcmp(s, "ldst")) return 1;
        if (!strcmp(s, "tma")) return 2;
        if (!strcmp(s, "ce")) return 3;
    }
    char* end = nullptr;
    long long v = strtoll(s, &end, 0);
    return end == s ? dflt : v;
}

int check_ready() {
    if (!g.inited) return fail(P2P_INVALID_USAGE, "p2p_init has not been called");
    if (!g.connected) return fail(P2P_INVALID_USAGE, "p2p_connect has not been called");
    return P2P_SUCCESS;
}


// ---- options: one table, one set of range checks for the environment and for p2p_set_option ----

using OptionSlot = long long Options::*;
struct OptionSpec {
    const char* name;
    const char* env;       // nullptr: not settable from the environment
    OptionSlot slot;
    long long lo, hi;
};

const OptionSpec kOptionSpecs[] = {
    {"path", "P2P_PATH", &Options::path, 0, 3},
    {"nblocks", "P2P_NBLOCKS", &Options::nblocks, 0, 4096},
    {"self_nblocks", "P2P_SELF_NBLOCKS", &Options::self_nblocks, 0, 4096},
    {"pull_nblocks", "P2P_PULL_NBLOCKS", &Options::pull_nblocks, 0, 4096},
    {"eager_nblocks", "P2P_EAGER_NBLOCKS", &Options::eager_nblocks, 0, 4096},
    {"eager_bytes", "P2P_EAGER_BYTES", &Options::eager_bytes, 0, 1 << 20},
    {"pull_min_bytes", "P2P_PULL_MIN_BYTES", &Options::pull_min_bytes, -1, 1ll << 40},
    {"timeout_ms", "P2P_TIMEOUT_MS", &Options::timeout_ms, 1, 3600ll * 1000},
    {"tma_chunk", "P2P_TMA_CHUNK", &Options::tma_chunk, 1024, 128 * 1024},
    {"tma_stages", "P2P_TMA_STAGES", &Options::tma_stages, 2, kTmaMaxStages},
    {"tma_min_bytes", "P2P_TMA_MIN_BYTES", &Options::tma_min_bytes, -1, 1ll << 40},
    {"tma_peer_chunk", "P2P_TMA_PEER_CHUNK", &Options::tma_peer_chunk, 1024, 128 * 1024},
    {"tma_peer_stages", "P2P_TMA_PEER_STAGES", &Options::tma_peer_stages, 2, kTmaMaxStages},
    {"pull_tma_min_bytes", "P2P_PULL_TMA_MIN_BYTES", &Options::pull_tma_min_bytes, -1, 1ll << 40},
    {"self_tma_min_bytes", "P2P_SELF_TMA_MIN_BYTES", &Options::self_tma_min_bytes, -1, 1ll << 40},
    {"pdl", "P2P_PDL", &Options::pdl, 0, 3},
    {"early_post", "P2P_EARLY_POST", &Options::early_post, 0, 1},
    {"loopback", "P2P_LOOPBACK", &Options::loopback, 0, 1},
    {"bidir_push", "P2P_BIDIR_PUSH", &Options::bidir_push, 0, 3},
    {"pull_total", "P2P_PULL_TOTAL", &Options::pull_total, 1, 1024},
    {"push_total", "P2P_PUSH_TOTAL", &Options::push_total, 1, 1024},
    {"fan_min", "P2P_FAN_MIN", &Options::fan_min, 1, 64},
    {"ce_min_bytes", "P2P_CE_MIN_BYTES", &Options::ce_min_bytes, -1, 1ll << 40},
    {"ce_wait_us", "P2P_CE_WAIT_US", &Options::ce_wait_us, 0, 10000000},
    {"ce_flags", "P2P_CE_FLAGS", &Options::ce_flags, 0, 2},
    {"stage_chunk", "P2P_STAGE_CHUNK", &Options::stage_chunk, 0, 16 << 20},
    {"stage_lanes", "P2P_STAGE_LANES", &Options::stage_lanes, 1, kStageLanes},
};

const OptionSpec* find_option(const char* name) {
    for (const OptionSpec& sp : kOptionSpecs)
        if (!strcmp(sp.name, name)) return &sp;
    return nullptr;
}

// Range of one value, and the constraints that tie options together, checked on the set `o` would become.
int validate_option(const OptionSpec& sp, long long value, const Options& o, const char* source) {
    if (value < sp.lo || value > sp.hi)
        return fail(P2P_INVALID_ARGUMENT, "%s%s = %lld is outside [%lld, %lld]", source, sp.name, value, sp.lo, sp.hi);
    Options t = o;
    t.*(sp.slot) = value;
    if (t.tma_chunk % 16)
        return fail(P2P_INVALID_ARGUMENT, "%stma_chunk = %lld is not a multiple of 16 (cp.async.bulk alignment)", source, t.tma_chunk);
    if (t.tma_chunk * t.tma_stages > 200 * 1024)
        return fail(P2P_INVALID_ARGUMENT, "%stma ring %lld x %lld bytes exceeds 200 KiB of shared memory", source, t.tma_stages, t.tma_chunk);
    if (t.tma_peer_chunk % 16)
        return fail(P2P_INVALID_ARGUMENT, "%stma_peer_chunk = %lld is not a multiple of 16 (cp.async.bulk alignment)", source, t.tma_peer_chunk);
    if (t.tma_peer_chunk * t.tma_peer_stages > 200 * 1024)
        return fail(P2P_INVALID_ARGUMENT, "%stma ring %lld x %lld bytes exceeds 200 KiB of shared memory", source, t.tma_peer_stages, t.tma_peer_chunk);
    return P2P_SUCCESS;
}

// The staging ring's chunk when left to the engine: a function of the rank count alone, so that all ranks agree.
void resolve_stage_chunk(Options& o, int nranks) {
    if (o.stage_chunk > 0) {
        o.stage_chunk = std::max<long long>(4096, o.stage_chunk & ~255ll);
        return;
    }
    long long chunk = 256 << 10;
    while (chunk > (64 << 10) && chunk * kStageLanes * kStageSlots * nranks > (256ll << 20)) chunk >>= 1;
    o.stage_chunk = chunk;
}

// Options from the environment, through the same checks as p2p_set_option.
int read_env_options(Options& o) {
    for (const OptionSpec& sp : kOptionSpecs) {
        if (!sp.env) continue;
        const long long v = env_ll(sp.env, o.*(sp.slot));
        if (v == o.*(sp.slot)) continue;
        char source[64];
        snprintf(source, sizeof source, "%s: ", sp.env);
        int rc = validate_option(sp, v, o, source);
        if (rc) return rc;
        o.*(sp.slot) = v;
    }
    return P2P_SUCCESS;
}

// ---- host mailbox ---------------------------------------------------------------------

inline HostPost* hostbox_entry(int dst, int src, uint64_t seq) {
    return g.hostbox + ((size_t)dst * g.nranks + src) * kRing + seq % kRing;
}

void hostbox_publish(int src, uint64_t seq, uint64_t word) {
    if (!g.hostbox) return;
    HostPost* e = hostbox_entry(g.rank, src, seq);
    e->seq.store(0, std::memory_order_relaxed);  // seqlock: invalidate, write, validate
    std::atomic_thread_fence(std::memory_order_release);
    e->word.store(word, std::memory_order_relaxed);
    e->seq.store(seq, std::memory_order_release);
}

// Has rank `dst` already said where message `seq` from me goes?  Waits at most `wait_us`.
bool hostbox_lookup(int dst, uint64_t seq, long long wait_us, uint64_t* word) {
    if (!g.hostbox) return false;
    HostPost* e = hostbox_entry(dst, g.rank, seq);
    const auto t0 = std::chrono::steady_clock::now();
    for (;;) {
        const uint64_t have = e->seq.load(std::memory_order_acquire);
        if (have > seq) return false;  // the receiver's host is more than a ring ahead: the entry is gone
        if (have == seq) {
            const uint64_t w = e->word.load(std::memory_order_relaxed);
            std::atomic_thread_fence(std::memory_order_acquire);
            if (e->seq.load(std::memory_order_relaxed) == seq) {
                *word = w;
                return true;
            }
        }
        if (std::chrono::duration_cast<std::chrono::microseconds>(std::chrono::steady_clock::now() - t0).count() >= wait_us)
            return false;
    }
}

int hostbox_create() {
    snprintf(g.hostbox_name, sizeof g.hostbox_name, "/b200p2p.%d.%llx", (int)getpid(),
             (unsigned long long)std::chrono::steady_clock::now().time_since_epoch().count());
    int fd = shm_open(g.hostbox_name, O_CREAT | O_EXCL | O_RDWR, 0600);
    if (fd < 0) return fail(P2P_SYSTEM_ERROR, "shm_open(%s) failed: %s", g.hostbox_name, strerror(errno));
    if (ftruncate(fd, (off_t)g.hostbox_bytes) != 0) {
        close(fd);
        shm_unlink(g.hostbox_name);
        return fail(P2P_SYSTEM_ERROR, "ftruncate(%s, %zu) failed: %s", g.hostbox_name, g.hostbox_bytes, strerror(errno));
    }
    void* m = mmap(nullptr, g.hostbox_bytes, PROT_READ | PROT_WRITE, MAP_SHARED, fd, 0);
    close(fd);
    if (m == MAP_FAILED) {
        shm_unlink(g.hostbox_name);
        return fail(P2P_SYSTEM_ERROR, "mmap of the host mailbox failed: %s", strerror(errno));
    }
    g.hostbox = (HostPost*)m;  // fresh shared memory is zero-filled: no entry is valid
    static bool registered = false;
    if (!registered) {  // the reference never tears anything down (p2p_matrix.cc:270-274): do not leak the name
        atexit([] { if (g.hostbox_name[0]) shm_unlink(g.hostbox_name); });
        registered = true;
    }
    return P2P_SUCCESS;
}

int hostbox_open(const char* name) {
    int fd = shm_open(name, O_RDWR, 0600);
    if (fd < 0) return fail(P2P_SYSTEM_ERROR, "shm_open(%s) failed: %s (ranks must share /dev/shm)", name, strerror(errno));
    void* m = mmap(nullptr, g.hostbox_bytes, PROT_READ | PROT_WRITE, MAP_SHARED, fd, 0);
    close(fd);
    if (m == MAP_FAILED) return fail(P2P_SYSTEM_ERROR, "mmap of the host mailbox failed: %s", strerror(errno));
    g.hostbox = (HostPost*)m;
    return P2P_SUCCESS;
}

// ---- plan selection -----------------------------------------------------------------

enum Proto { PROTO_EAGER = 4, PROTO_PUSH = 1, PROTO_LARGE = 5, PROTO_COPY = 6 };

inline bool is_large(size_t nbytes) { return g.opt.pull_min_bytes >= 0 && (long long)nbytes >= g.opt.pull_min_bytes; }

inline Proto proto_for(size_t nbytes) {
    if ((long long)nbytes <= g.opt.eager_bytes) return PROTO_EAGER;
    return is_large(nbytes) ? PROTO_LARGE : PROTO_PUSH;
}

inline int eager_ctas(size_t nbytes) {
    // 4096 lines (32 KiB of payload) per CTA
    const size_t n = (nbytes + 32767) / 32768;
    const long long cap = g.opt.eager_nblocks ? g.opt.eager_nblocks : 32;  // 1 MiB: 32 CTAs 8.6 us vs 16 CTAs 9.9 us per message
    return (int)std::max<size_t>(1, std::min<size_t>(n, (size_t)cap));
}

// flavour of the kernel a push of this size would like (1 ldst, 2 tma)
inline int push_flavour(size_t nbytes) {
    int path = (int)g.opt.path;
    if (path == 0 || path == 3) path = (g.opt.tma_min_bytes >= 0 && (long long)nbytes >= g.opt.tma_min_bytes) ? 2 : 1;
    return path;
}

// flavour a matched send-to-self (local copy) would like: measured on B200, the TMA ring with one CTA
// per SM is the fastest large HBM copy (96% of the measured copy peak at 256 MiB), ld/st the fastest small one
inline int self_copy_flavour(size_t nbytes) {
    if (g.opt.path == 1 || g.opt.path == 2) return (int)g.opt.path;
    return (g.opt.self_tma_min_bytes >= 0 && (long long)nbytes >= g.opt.self_tma_min_bytes) ? 2 : 1;
}

// CTAs that move `nbytes`: `flavour` of the launch, destination class decides the cap
inline int mover_ctas(size_t nbytes, int flavour, bool self, bool pull, int fanout = 1, long long tma_chunk = 0) {
    if (flavour == FLAVOUR_TMA && !pull) {
        if (tma_chunk <= 0) tma_chunk = g.opt.tma_chunk;
        const size_t chunks = (nbytes + (size_t)tma_chunk - 1) / (size_t)tma_chunk;
        const long long cap = self ? (g.opt.self_nblocks ? g.opt.self_nblocks : (long long)g.sm_count * g.ctas_tma)
                                   : (g.opt.nblocks ? g.opt.nblocks : 32);
        return (int)std::max<size_t>(1, std::min<size_t>((chunks + 1) / 2, (size_t)cap));  // >= 2 chunks per ring
    }
    const size_t tile = (size_t)kLdstThreads * 128;  // bytes one CTA moves per loop trip
    const size_t tiles = (nbytes + tile - 1) / tile;
    long long cap;
    if (self) cap = g.opt.self_nblocks ? g.opt.self_nblocks : (long long)g.sm_count;  // 1 CTA per SM beat 2 (6.17 vs 6.02 TB/s)
    // one flow wants many CTAs in flight (pull: 64 -> 725 GB/s, 32 -> 610); several concurrent flows per GPU
    // want few each (all-pairs on 8 GPUs: 8 CTAs per flow 4.83 TB/s, 32 per flow 4.58) but not fewer in total
    // than one flow would get (3 flows of 8 CTAs on 4 GPUs: 1.73 TB/s, of 32: 2.43): the flows share a total
    else if (pull) cap = g.opt.pull_nblocks ? g.opt.pull_nblocks : std::max(g.opt.fan_min, g.opt.pull_total / std::max(fanout, 1));
    else cap = g.opt.nblocks ? g.opt.nblocks : std::min<long long>(32, std::max(g.opt.fan_min, g.opt.push_total / std::max(fanout, 1)));
    return (int)std::max<size_t>(1, std::min<size_t>(tiles, (size_t)cap));
}

bool lookup_window(const void* buf, size_t nbytes, uint32_t* win, uint64_t* off) {
    const uint8_t* b = (const uint8_t*)buf;
    for (int w = 0; w < P2P_MAX_WINDOWS; ++w) {
        const Window& W = g.windows[w];
        if (!W.used || !W.base) continue;
        if (b >= W.base && b + nbytes <= W.base + W.bytes) {
            *win = (uint32_t)w;
            *off = (uint64_t)(b - W.base);
            return true;
        }
    }
    return false;
}

inline uint32_t* arrive_push(int p, uint64_t seq) { return g.arrive + (size_t)p * kRing + seq % kRing; }
inline uint32_t* arrive_pull(int p, uint64_t seq) { return g.arrive + (size_t)(g.nranks + p) * kRing + seq % kRing; }
inline uint32_t* arrive_ll(int p, uint64_t seq) { return g.arrive + (size_t)2 * g.nranks * kRing + (size_t)p * kLLSlots + seq % kLLSlots; }
// staging ring: per peer one send counter, one receive counter, one lock
inline size_t arrive_words(int nranks) { return (size_t)nranks * (2 * kRing + kLLSlots + 3); }
inline uint32_t* arrive_stage(int p) { return g.arrive + (size_t)g.nranks * (2 * kRing + kLLSlots) + (size_t)p * 3; }
inline size_t stage_area_bytes() { return kStageHeaderBytes + (size_t)kStageLanes * kStageSlots * (size_t)g.opt.stage_chunk; }
inline uint8_t* stage_area(uint8_t* region, int from) { return region + g.stage_off + (size_t)from * stage_area_bytes(); }

// Where everything of the staging ring between me and peer p lies, as the kernels need it.
StageInfo stage_info_for(int p) {
    StageInfo si;
    memset(&si, 0, sizeof si);
    StageHeader* mine = (StageHeader*)stage_area(g.ctl, p);               // in my region: what p writes
    StageHeader* theirs = (StageHeader*)stage_area(g.peer_ctl[p], g.rank);  // in p's region: what I write
    si.send_ring = (uint64_t)((uint8_t*)theirs + kStageHeaderBytes);
    si.send_filled = (uint64_t)theirs->filled;
    si.send_drained = (uint64_t)mine->drained;
    si.send_arrive = (uint64_t)arrive_stage(p);
    si.recv_ring = (uint64_t)((uint8_t*)mine + kStageHeaderBytes);
    si.recv_filled = (uint64_t)mine->filled;
    si.recv_drained = (uint64_t)theirs->drained;
    si.recv_arrive = (uint64_t)(arrive_stage(p) + 1);
    si.lock = (uint64_t)(arrive_stage(p) + 2);
    si.chunk = (uint32_t)g.opt.stage_chunk;
    si.slots = kStageSlots;
    si.lanes = (uint32_t)g.opt.stage_lanes;
    return si;
}

// ---- group launch --------------------------------------------------------------------
//
// launch_pending() turns the operations collected since p2p_group_start (or a single ungrouped call) into one
// kernel launch.  Steps, each a function below:
//   match_self_pairs   a send to oneself and a receive from oneself become one local copy
//   resolve_protocols  protocol per message, window lookup, who moves the bytes; everything that can be refused
//   assign_sequence    sequence numbers in program order; large receives tell the sender's host where they go
//   choose_copy_engine the opt-in DMA path for a lone large send
//   size_roles         CTAs per role, scaled so that the whole launch is co-resident
//   build_args         the kernel's argument block: roles, flags, block ranges
//   submit             stream dependencies, the launch, copy-engine chains

struct Resolved {
    Proto proto;
    uint32_t win;
    uint64_t off;
    bool registered;
    bool pull_send;   // large send that lets the receiver pull
    bool staged;      // receive into a buffer that is in no window, below the pull threshold: through the staging ring
    int nblocks;
    uint64_t seq;
    bool ce;          // large send handed to the copy engine
    uint64_t ce_word;
};

struct GroupShape {
    int flavour = FLAVOUR_LDST;
    long long tma_chunk = 32 * 1024, tma_stages = 6;  // ring geometry of this launch (TMA flavour)
    int n_large_sends = 0, n_large_recvs = 0;
    uint32_t n_ctrl = 0;
    size_t eager_total = 0;
    bool only_small = true;
    bool any_ce = false;
    bool peers = false;   // the group holds operations with other ranks (not only a local copy)
};

int match_self_pairs(const std::vector<PendingOp>& ops, std::vector<int>& partner) {
    partner.assign(ops.size(), -1);
    if (g.opt.loopback) return P2P_SUCCESS;
    std::vector<size_t> self_sends, self_recvs;
    for (size_t i = 0; i < ops.size(); ++i)
        if (ops[i].peer == g.rank) (ops[i].is_send ? self_sends : self_recvs).push_back(i);
    for (size_t k = 0; k < std::min(self_sends.size(), self_recvs.size()); ++k) {
        const size_t si = self_sends[k], ri = self_recvs[k];
        if (ops[si].nbytes != ops[ri].nbytes)
            return fail(P2P_INVALID_ARGUMENT, "send to self of %zu bytes is matched by a receive of %zu bytes", ops[si].nbytes,
                        ops[ri].nbytes);
        partner[si] = (int)ri;
        partner[ri] = (int)si;
    }
    return P2P_SUCCESS;
}

// Nothing here changes engine state: a group that is refused leaves every sequence number where it was.
int resolve_protocols(const std::vector<PendingOp>& ops, const std::vector<int>& partner, std::vector<Resolved>& res,
                      GroupShape& shape) {
    for (size_t i = 0; i < ops.size(); ++i)
        if (partner[i] < 0 && is_large(ops[i].nbytes) && (long long)ops[i].nbytes > g.opt.eager_bytes)
            (ops[i].is_send ? shape.n_large_sends : shape.n_large_recvs)++;
    bool peers_in_group = false;
    for (size_t j = 0; j < ops.size(); ++j)
        if (partner[j] < 0) peers_in_group = true;

    res.assign(ops.size(), Resolved{});
    size_t nops_out = 0;
    for (size_t i = 0; i < ops.size(); ++i) {
        const PendingOp& op = ops[i];
        Resolved& r = res[i];
        if (partner[i] >= 0) {
            r.proto = PROTO_COPY;
            if (op.is_send) {
                ++nops_out;
                // the TMA ring is the faster LOCAL copy; with peers in the group the whole launch stays ld/st
                // (128-thread CTAs and one CTA per SM would starve the peer roles)
                if (!peers_in_group && self_copy_flavour(op.nbytes) == 2) shape.flavour = FLAVOUR_TMA;
            }
            continue;
        }
        r.proto = proto_for(op.nbytes);
        ++nops_out;
        if (r.proto == PROTO_EAGER) continue;
        r.registered = lookup_window(op.buf, op.nbytes, &r.win, &r.off);
        // a receive buffer nobody registered (the reference's plain cudaMalloc, p2p_matrix.cc:127-128) cannot be
        // stored into by the sender: the payload then goes through the staging ring in this rank's control region
        if (!op.is_send && r.proto == PROTO_PUSH && !r.registered) r.staged = true;
        if (op.is_send && r.proto == PROTO_LARGE) {
            // The sender decides.  Let the receiver pull when it can map my buffer — except in a plain pairwise
            // exchange (one large send, and a large receive from that same peer in the group), where two pushes
            // interfere less than two pulls (1.32 vs 1.24 TB/s between two B200s).  With several large sends in
            // the group (fan-out, all-pairs) pulls with few CTAs win again (4.83 vs 4.28 TB/s on 8 GPUs).
            // A push needs the RECEIVER's buffer in a window.  What the peer registered is not known here, but in an
            // exchange both ends usually treat their buffers alike: if this rank's own large receives of the group
            // are in no window, a push would most likely end in the staging ring, where a pull goes straight.
            bool from_same_peer = false, my_recvs_registered = true;
            for (size_t j = 0; j < ops.size(); ++j) {
                if (ops[j].is_send || partner[j] >= 0 || !is_large(ops[j].nbytes)) continue;
                if (ops[j].peer == op.peer) from_same_peer = true;
                uint32_t w;
                uint64_t o;
                if (!lookup_window(ops[j].buf, ops[j].nbytes, &w, &o)) my_recvs_registered = false;
            }
            bool push_instead = false;
            switch (g.opt.bidir_push) {
                case 1: push_instead = from_same_peer && shape.n_large_sends == 1; break;
                case 2: push_instead = shape.n_large_recvs > 0 && shape.n_large_sends == 1; break;
                case 3: push_instead = shape.n_large_recvs > 0; break;
                default: break;
            }
            if (r.registered && !my_recvs_registered) push_instead = false;
            r.pull_send = r.registered && !push_instead && g.opt.path != 2 && g.opt.path != 1;
        }
        if (op.is_send && !(r.proto == PROTO_LARGE && r.pull_send) && push_flavour(op.nbytes) == 2) shape.flavour = FLAVOUR_TMA;
    }
    if (nops_out > (size_t)kMaxGroupOps)
        return fail(P2P_INVALID_USAGE, "a group holds at most %d operations (got %zu)", kMaxGroupOps, nops_out);
    // Traffic in both directions at once - this rank pulls a large message while a peer pulls one from it (ring step,
    // all-pairs) - goes through the bulk-copy ring: cp.async.bulk reads of peer memory put larger requests on the
    // link than the ld/st loop, which leaves more of the reverse direction to the other flow's payload (two B200s,
    // 256 MiB both ways: 1269-1285 GB/s against 1216; one way alone the two tie at 741-751 and the ld/st loop starts
    // faster).  Not when this rank pushes: the push then decides the flavour.
    if (g.opt.path == 0 && g.opt.pull_tma_min_bytes >= 0 && shape.flavour == FLAVOUR_LDST) {
        bool pulls = false, offers = false, pushes = false;
        for (size_t i = 0; i < ops.size(); ++i) {
            if (partner[i] >= 0 || res[i].proto != PROTO_LARGE) continue;
            if (!ops[i].is_send) pulls = pulls || (long long)ops[i].nbytes >= g.opt.pull_tma_min_bytes;
            else if (res[i].pull_send) offers = true;
            else pushes = true;
        }
        if (pulls && offers && !pushes) shape.flavour = FLAVOUR_TMA;
    }
    // ring geometry: launches that cross NVLink want few large stages, pure local copies more and smaller ones
    shape.peers = peers_in_group;
    shape.tma_chunk = peers_in_group ? g.opt.tma_peer_chunk : g.opt.tma_chunk;
    shape.tma_stages = peers_in_group ? g.opt.tma_peer_stages : g.opt.tma_stages;
    return P2P_SUCCESS;
}

// Sequence numbers in PROGRAM order (the k-th send to a peer pairs with its k-th receive, whatever protocol each
// message ends up using).  From here on nothing may fail for a reason the peer would not share.
void assign_sequence(const std::vector<PendingOp>& ops, std::vector<Resolved>& res) {
    for (size_t i = 0; i < ops.size(); ++i) {
        Resolved& r = res[i];
        if (r.proto == PROTO_COPY) continue;
        const int p = ops[i].peer;
        if (r.proto == PROTO_EAGER) r.seq = ops[i].is_send ? ++g.ll_send_seq[p] : ++g.ll_recv_seq[p];
        else r.seq = ops[i].is_send ? ++g.rdv_send_seq[p] : ++g.rdv_recv_seq[p];
        // a large receive also tells the sender's HOST where it goes (copy-engine path)
        if (!ops[i].is_send && r.proto == PROTO_LARGE && p != g.rank)
            hostbox_publish(p, r.seq, r.registered ? make_post_word(r.seq, r.win, r.off) : 0);
    }
}

bool memops_ready();

// Copy engine (opt-in) for a send that is alone in its group and whose destination the host already knows.
// Alone, because the chain runs beside the group's kernel on another stream and a receive in that kernel may
// depend on it: two launches that must overlap are not something a GPU promises.  Never waits longer than
// ce_wait_us for the receiver's host; otherwise the SM path is used as if the option were off.
void choose_copy_engine(const std::vector<PendingOp>& ops, std::vector<Resolved>& res, GroupShape& shape) {
    for (size_t i = 0; i < ops.size(); ++i) {
        Resolved& r = res[i];
        const PendingOp& op = ops[i];
        if (!(op.is_send && r.proto == PROTO_LARGE && op.peer != g.rank)) continue;
        if (g.opt.ce_min_bytes < 0 || (long long)op.nbytes < g.opt.ce_min_bytes) continue;
        if (g.opt.ce_flags == 0 && ops.size() != 1) continue;  // a chain with kernels in it must not have to overlap the group's
        if (g.opt.ce_flags != 0 && !memops_ready()) continue;
        if (g.opt.path != 0 && g.opt.path != 3) continue;
        uint64_t word = 0;
        if (!hostbox_lookup(op.peer, r.seq, g.opt.ce_wait_us, &word) || word == 0) continue;
        const uint32_t win = (uint32_t)(word >> 40) & 0xffu;
        if (win >= P2P_MAX_WINDOWS || !g.windows[win].used || !g.windows[win].peer_base[op.peer]) continue;
        r.ce = true;
        r.ce_word = word;
        r.pull_send = false;
        shape.any_ce = true;
    }
}

// CTAs per role.  Everything must be co-resident: roles wait for peers, and a CTA that cannot get an SM could
// keep a peer waiting for ever.  Eager roles may take at most half of the machine, movers share what is left.
void size_roles(const std::vector<PendingOp>& ops, std::vector<Resolved>& res, GroupShape& shape) {
    const int per_sm = shape.flavour == FLAVOUR_TMA ? g.ctas_tma : g.ctas_ldst;
    long long budget = (long long)g.sm_count * per_sm;
    long long fixed = 0, want = 0;
    for (size_t i = 0; i < ops.size(); ++i) {
        const PendingOp& op = ops[i];
        Resolved& r = res[i];
        const bool self = op.peer == g.rank;
        switch (r.proto) {
            case PROTO_COPY:
                r.nblocks = op.is_send ? mover_ctas(op.nbytes, shape.flavour, true, false, 1, shape.tma_chunk) : 0;
                // beside transfers over NVLink (all-to-all's own block) a local copy is the short job: it must not
                // take the SMs the link-bound roles need
                if (shape.peers && !g.opt.self_nblocks) r.nblocks = std::min(r.nblocks, 32);
                want += r.nblocks;
                if (op.nbytes > 4096) shape.only_small = false;
                break;
            case PROTO_EAGER:
                r.nblocks = eager_ctas(op.nbytes);
                fixed += r.nblocks;
                shape.eager_total += op.nbytes;
                break;
            case PROTO_PUSH:
                shape.only_small = false;
                if (op.is_send) { r.nblocks = mover_ctas(op.nbytes, shape.flavour, self, false, 1, shape.tma_chunk); want += r.nblocks; }
                else if (r.staged) { r.nblocks = std::min<int>(kStageLanes, mover_ctas(op.nbytes, FLAVOUR_LDST, self, true)); want += r.nblocks; }
                else ++shape.n_ctrl;
                break;
            case PROTO_LARGE:
                shape.only_small = false;
                if (op.is_send) {
                    if (r.ce) break;  // moved by the copy engine, outside the kernel
                    if (r.pull_send) ++shape.n_ctrl;
                    else { r.nblocks = mover_ctas(op.nbytes, shape.flavour, self, false, shape.n_large_sends, shape.tma_chunk); want += r.nblocks; }
                } else {
                    r.nblocks = mover_ctas(op.nbytes, FLAVOUR_LDST, self, true, shape.n_large_recvs);
                    // a lone pull through the bulk-copy ring (ring step) likes a few more, lighter CTAs: 96 of 128 threads
                    // 646-648 GB/s per GPU on 4 and 8 B200s, 64: 639-640, 147: 641
                    if (shape.flavour == FLAVOUR_TMA && shape.n_large_recvs == 1 && !g.opt.pull_nblocks && r.nblocks == 64) r.nblocks = 96;
                    want += r.nblocks;
                }
                break;
        }
    }
    budget -= (shape.n_ctrl ? 1 : 0);
    if (fixed > budget / 2) {
        const long long room = std::max<long long>(budget / 2, 1);
        long long now = 0;
        for (size_t i = 0; i < ops.size(); ++i)
            if (res[i].proto == PROTO_EAGER) {
                res[i].nblocks = (int)std::max<long long>(1, res[i].nblocks * room / fixed);
                now += res[i].nblocks;
            }
        fixed = now;
    }
    budget -= fixed;
    if (budget < 1) budget = 1;
    if (want > budget)
        for (size_t i = 0; i < ops.size(); ++i)
            if (res[i].nblocks > 0 && res[i].proto != PROTO_EAGER)
                res[i].nblocks = (int)std::max<long long>(1, res[i].nblocks * budget / want);
}

// Fills one role's descriptor: which flags it polls and writes, in whose control region.
void describe_role(const PendingOp& op, const Resolved& r, const PendingOp* copy_partner, bool may_post_early, OpDesc& d) {
    const int me = g.rank, p = op.peer;
    uint8_t* my = g.ctl;
    uint8_t* theirs = g.peer_ctl[p];
    const uint64_t seq = r.seq;
    d.local = (uint64_t)op.buf;
    d.nbytes = op.nbytes;
    d.peer = (uint32_t)p;
    d.nblocks = (uint32_t)r.nblocks;
    d.seq = seq;
    d.stage = (uint64_t)(g.stage_table + p);
    if (r.proto == PROTO_COPY) {
        d.kind = OP_COPY;
        d.aux = (uint64_t)copy_partner->buf;
    } else if (r.proto == PROTO_EAGER && !op.is_send) {
        d.kind = OP_RECV_LL;
        d.aux = (uint64_t)ll_slot(my, p, seq);
        d.f_signal = (uint64_t)&peer_section(theirs, me)->ll_ack[seq % kLLSlots];
        d.arrive = (uint64_t)arrive_ll(p, seq);
    } else if (r.proto == PROTO_EAGER) {
        d.kind = OP_SEND_LL;
        d.aux = (uint64_t)ll_slot(theirs, me, seq);
        d.f_wait = (uint64_t)&peer_section(my, p)->ll_ack[seq % kLLSlots];
    } else if (!op.is_send) {
        // what a pushing sender needs from me
        d.word = r.registered ? make_post_word(seq, r.win, r.off) : 0;
        d.f_signal = (uint64_t)&peer_section(theirs, me)->post[seq % kRing];
        d.f_wait = (uint64_t)&peer_section(my, p)->done[seq % kRing];
        if (r.proto == PROTO_PUSH) {
            d.kind = r.staged ? OP_RECV_STAGED : OP_POST;
        } else {
            d.kind = OP_RECV_ADAPT;
            d.f_wait2 = (uint64_t)&peer_section(my, p)->pull_post[seq % kRing];
            d.f_signal2 = (uint64_t)&peer_section(theirs, me)->pull_done[seq % kRing];
            d.aux = (uint64_t)(g.win_table + (size_t)p * P2P_MAX_WINDOWS);
            d.arrive = (uint64_t)arrive_pull(p, seq);
        }
    } else if (r.proto == PROTO_LARGE && r.pull_send) {
        d.kind = OP_POST;  // publish where my payload lies, wait until it has been read
        d.flags = may_post_early ? OPF_EARLY_POST : 0;
        d.word = make_post_word(seq, r.win, r.off);
        d.f_signal = (uint64_t)&peer_section(theirs, me)->pull_post[seq % kRing];
        d.f_wait = (uint64_t)&peer_section(my, p)->pull_done[seq % kRing];
    } else {
        d.kind = OP_PUSH;
        d.f_wait = (uint64_t)&peer_section(my, p)->post[seq % kRing];
        d.f_signal = (uint64_t)&peer_section(theirs, me)->done[seq % kRing];
        d.aux = (uint64_t)(g.win_table + (size_t)p * P2P_MAX_WINDOWS);
        d.arrive = (uint64_t)arrive_push(p, seq);
        if (r.proto == PROTO_LARGE) {  // the receiver is waiting to be told
            d.f_signal2 = (uint64_t)&peer_section(theirs, me)->pull_post[seq % kRing];
            d.word2 = make_post_word(seq, kWinPushPlease, 0);
        }
    }
}

// Order of CTAs: control CTA, eager sends (they wait for nothing that runs here), eager receives, large
// receives (they start by waiting), then everything that pushes or copies.  Returns the grid size.
uint32_t build_args(const std::vector<PendingOp>& ops, const std::vector<Resolved>& res, const std::vector<int>& partner,
                    const GroupShape& shape, GroupArgs& args) {
    memset(&args, 0, sizeof args);
    // Which engine launch precedes this one on its stream, and may this one's sends be offered before that
    // launch has retired?  Only when the whole group sits on one stream (no event waits in front of the kernel)
    // and the payload does not overlap anything the launch ahead receives into.
    cudaStream_t launch_stream = ops[0].stream;
    bool one_stream = true;
    for (const PendingOp& op : ops) one_stream = one_stream && op.stream == launch_stream;
    const Engine::StreamLast* prev = nullptr;
    if (one_stream) {
        auto it = g.stream_last.find(launch_stream);
        if (it != g.stream_last.end()) prev = &it->second;
    }
    auto early_ok = [&](const PendingOp& op) {
        if (!prev || !g.opt.early_post || g.opt.pdl != 3) return false;
        const uint8_t* b = (const uint8_t*)op.buf;
        for (const auto& w : prev->writes)
            if (b < w.first + w.second && w.first < b + op.nbytes) return false;
        return true;
    };
    args.h.tokens = g.tokens;
    if (++g.launch_counter == 0) ++g.launch_counter;
    args.h.my_token = g.launch_counter;
    args.h.prev_token = prev ? prev->token : 0;
    args.h.tma_chunk = (uint32_t)shape.tma_chunk;
    args.h.tma_stages = (uint32_t)shape.tma_stages;
    args.h.timeout_ns = (uint64_t)g.opt.timeout_ms * 1000000ull;
    args.h.async_err = g.async_err_dev;
    args.h.n_ctrl = shape.n_ctrl;
    uint32_t next_block = shape.n_ctrl ? 1 : 0;
    for (int pass = 0; pass < 5; ++pass) {
        for (size_t i = 0; i < ops.size(); ++i) {
            const PendingOp& op = ops[i];
            const Resolved& r = res[i];
            int cls;
            if (r.proto == PROTO_COPY) cls = op.is_send ? 4 : -1;
            else if (r.proto == PROTO_EAGER) cls = op.is_send ? 1 : 2;
            else if (r.proto == PROTO_PUSH) cls = op.is_send ? 4 : r.staged ? 3 : 0;
            else cls = op.is_send ? (r.ce ? -1 : r.pull_send ? 0 : 4) : 3;
            if (cls != pass) continue;
            OpDesc& d = args.ops[args.h.nops++];
            describe_role(op, r, r.proto == PROTO_COPY ? &ops[(size_t)partner[i]] : nullptr, op.is_send && early_ok(op), d);
            if (d.kind != OP_POST) {
                d.first_block = next_block;
                next_block += d.nblocks;
            }
        }
    }
    bool light = true;  // no rendezvous role in the group
    for (uint32_t i = 0; i < args.h.nops; ++i) {
        const uint32_t k = args.ops[i].kind;
        if (k == OP_POST || k == OP_PUSH || k == OP_RECV_ADAPT || k == OP_RECV_STAGED) light = false;
    }
    bool mid = false;  // a rendezvous message below the pull threshold: those measured slower with PDL (1-2 MiB: +1 us)
    for (size_t i = 0; i < ops.size(); ++i)
        if (res[i].proto == PROTO_PUSH) mid = true;
    if (shape.any_ce || g.opt.pdl == 0) args.h.pdl = 0;
    else if (light) args.h.pdl = kPdlLaunched;                      // trigger at kernel entry
    else if (g.opt.pdl == 2) args.h.pdl = kPdlLaunched;
    else if (g.opt.pdl == 3 && !mid) args.h.pdl = kPdlLaunched | kPdlLateTrigger;  // movers trigger after their copy loop
    else args.h.pdl = 0;
    // remember this launch as the latest on its stream; the other streams of the group now end in an event wait
    Engine::StreamLast mine;
    mine.token = args.h.my_token;
    for (const PendingOp& op : ops) {
        if (!op.is_send && op.nbytes) mine.writes.emplace_back((const uint8_t*)op.buf, op.nbytes);
        if (op.stream != launch_stream) g.stream_last.erase(op.stream);
    }
    if (g.stream_last.size() > 256) g.stream_last.clear();  // streams come and go; forgetting only costs an early post
    g.stream_last[launch_stream] = std::move(mine);
    return next_block;
}

// Stream memory operations of the driver API (the runtime API has no equivalent), resolved once.
typedef int (*cuStreamMemOp64_t)(cudaStream_t stream, unsigned long long addr, unsigned long long value, unsigned int flags);
struct MemOps {
    cuStreamMemOp64_t wait = nullptr, write = nullptr;
    int state = 0;  // 0 not tried, 1 usable, -1 unavailable
};
MemOps g_memops;
constexpr unsigned kWaitValueEq = 1;  // CU_STREAM_WAIT_VALUE_EQ

bool memops_ready() {
    if (g_memops.state) return g_memops.state > 0;
    g_memops.state = -1;
    void *w = nullptr, *x = nullptr;
    cudaDriverEntryPointQueryResult q1, q2;
    if (cudaGetDriverEntryPoint("cuStreamWaitValue64", &w, cudaEnableDefault, &q1) != cudaSuccess || q1 != cudaDriverEntryPointSuccess || !w)
        return false;
    if (cudaGetDriverEntryPoint("cuStreamWriteValue64", &x, cudaEnableDefault, &q2) != cudaSuccess || q2 != cudaDriverEntryPointSuccess || !x)
        return false;
    g_memops.wait = (cuStreamMemOp64_t)w;
    g_memops.write = (cuStreamMemOp64_t)x;
    g_memops.state = 1;
    return true;
}

// 8 bytes into a flag, in stream order, without a kernel: a stream memory operation, or (ce_flags 2, for flags in
// peer memory) a DMA from a pinned ring of values.
int stream_write_flag(cudaStream_t cs, uint64_t* flag, uint64_t value, bool in_peer_memory) {
    if (g.opt.ce_flags == 2 && in_peer_memory) {
        uint64_t* src = g.ce_vals + (g.ce_vals_next++ % 4096);
        *src = value;
        CUDA_TRY(cudaMemcpyAsync(flag, src, sizeof(uint64_t), cudaMemcpyDefault, cs));
        return P2P_SUCCESS;
    }
    const int r = g_memops.write(cs, (unsigned long long)flag, value, 0);
    if (r != 0) return fail(P2P_UNHANDLED_CUDA_ERROR, "cuStreamWriteValue64 on %s memory failed with CUresult %d", in_peer_memory ? "peer" : "local", r);
    return P2P_SUCCESS;
}

// One copy-engine chain on a side stream: tell the receiver "I will push", wait for its post (buffer ready),
// DMA the payload into the peer mapping, publish completion.  Readiness and completion stay on the device.
int submit_copy_engine_chain(const PendingOp& op, const Resolved& r, const GroupHdr& hdr, cudaStream_t launch_stream) {
    const int me = g.rank, p = op.peer;
    const uint64_t seq = r.seq, word = r.ce_word;
    if (!g.ce_stream[p]) {
        CUDA_TRY(cudaStreamCreateWithFlags(&g.ce_stream[p], cudaStreamNonBlocking));
        CUDA_TRY(cudaEventCreateWithFlags(&g.ce_join[p], cudaEventDisableTiming));
    }
    uint8_t* theirs = g.peer_ctl[p];
    uint8_t* dst = g.windows[(word >> 40) & 0xffu].peer_base[p] + (word & kPostOffMask);
    uint64_t* notice = &peer_section(theirs, me)->pull_post[seq % kRing];
    uint64_t* post = &peer_section(g.ctl, p)->post[seq % kRing];
    uint64_t* done = &peer_section(theirs, me)->done[seq % kRing];
    cudaStream_t cs = g.ce_stream[p];
    CUDA_TRY(cudaStreamWaitEvent(cs, g.ce_fork, 0));
    // P2P_CE_TIMING=1 (experiments): where a chain's time goes - events before the handshake, before and after the DMA
    static const bool timing = getenv("P2P_CE_TIMING") != nullptr;
    cudaEvent_t tev[3] = {};
    if (timing && g.ce_timing.size() < 512) {
        for (cudaEvent_t& e : tev) CUDA_TRY(cudaEventCreate(&e));
        CUDA_TRY(cudaEventRecord(tev[0], cs));
    }
    if (g.opt.ce_flags == 0) {
        CUDA_TRY(launch_ce_wait(notice, make_post_word(seq, kWinPushPlease, 0), post, seq, word, hdr, cs));
        CUDA_TRY(cudaMemcpyAsync(dst, op.buf, op.nbytes, cudaMemcpyDefault, cs));
        CUDA_TRY(launch_ce_signal(done, seq, cs));
        g.stat_launches += 2;
    } else {
        int rc = stream_write_flag(cs, notice, make_post_word(seq, kWinPushPlease, 0), true);
        if (rc) return rc;
        // the host was told the very word the device will post (choose_copy_engine): wait for exactly that
        const int wr = g_memops.wait(cs, (unsigned long long)post, word, kWaitValueEq);
        if (wr != 0) return fail(P2P_UNHANDLED_CUDA_ERROR, "cuStreamWaitValue64 failed with CUresult %d", wr);
        rc = stream_write_flag(cs, post, 0, false);  // consumed (see movers_signal)
        if (rc) return rc;
        if (tev[0]) CUDA_TRY(cudaEventRecord(tev[1], cs));
        CUDA_TRY(cudaMemcpyAsync(dst, op.buf, op.nbytes, cudaMemcpyDefault, cs));
        if (tev[0]) CUDA_TRY(cudaEventRecord(tev[2], cs));
        rc = stream_write_flag(cs, done, seq, true);
        if (rc) return rc;
        if (tev[0]) g.ce_timing.push_back({tev[0], tev[1], tev[2]});
    }
    CUDA_TRY(cudaEventRecord(g.ce_join[p], cs));
    CUDA_TRY(cudaStreamWaitEvent(launch_stream, g.ce_join[p], 0));
    g.stat_ce_copies++;
    return P2P_SUCCESS;
}

// Streams: launch on the first op's stream; the others wait for the kernel and it for them.
int submit(const std::vector<PendingOp>& ops, const std::vector<Resolved>& res, const GroupShape& shape, const GroupArgs& args,
           uint32_t grid, uint32_t block) {
    cudaStream_t launch_stream = ops[0].stream;
    std::vector<cudaStream_t> others;
    for (const PendingOp& op : ops)
        if (op.stream != launch_stream && std::find(others.begin(), others.end(), op.stream) == others.end())
            others.push_back(op.stream);
    while (g.ev_pool.size() < others.size()) {
        cudaEvent_t e;
        CUDA_TRY(cudaEventCreateWithFlags(&e, cudaEventDisableTiming));
        g.ev_pool.push_back(e);
    }
    for (size_t i = 0; i < others.size(); ++i) {
        CUDA_TRY(cudaEventRecord(g.ev_pool[i], others[i]));
        CUDA_TRY(cudaStreamWaitEvent(launch_stream, g.ev_pool[i], 0));
    }
    // copy-engine chains fork off BEFORE the kernel: the kernel may hold a receive that only completes once
    // the peer's chain has run, and the peer's chain may need ours
    if (shape.any_ce) CUDA_TRY(cudaEventRecord(g.ce_fork, launch_stream));
    if (grid > 0) {
        const int dyn_smem = shape.flavour == FLAVOUR_TMA ? (int)(shape.tma_chunk * shape.tma_stages) : 0;
        CUDA_TRY(launch_group(shape.flavour, args, grid, block, dyn_smem, launch_stream));
        g.stat_launches++;
    }
    if (shape.any_ce)
        for (size_t i = 0; i < ops.size(); ++i)
            if (res[i].ce) {
                int rc = submit_copy_engine_chain(ops[i], res[i], args.h, launch_stream);
                if (rc) return rc;
            }
    if (!others.empty()) {
        CUDA_TRY(cudaEventRecord(g.ev_done, launch_stream));
        for (cudaStream_t s : others) CUDA_TRY(cudaStreamWaitEvent(s, g.ev_done, 0));
    }
    return P2P_SUCCESS;
}

int launch_pending() {
    std::vector<PendingOp> ops;
    ops.swap(g.group_ops);
    if (ops.empty()) return P2P_SUCCESS;

    std::vector<int> partner;
    std::vector<Resolved> res;
    GroupShape shape;
    int rc = match_self_pairs(ops, partner);
    if (rc == P2P_SUCCESS) rc = resolve_protocols(ops, partner, res, shape);
    if (rc != P2P_SUCCESS) return rc;
    assign_sequence(ops, res);
    choose_copy_engine(ops, res, shape);
    size_roles(ops, res, shape);

    GroupArgs args;
    const uint32_t grid = build_args(ops, res, partner, shape, args);
    if (grid == 0 && !shape.any_ce) return P2P_SUCCESS;
    // small eager-only groups run with 128-thread CTAs: less to launch on the latency path
    const uint32_t block = (shape.flavour == FLAVOUR_LDST && shape.only_small && shape.eager_total <= 8192) ? 128 : kLdstThreads;

    {
        long long* st = g.last_stat;
        for (int i = 0; i < 12; ++i) st[i] = 0;
        st[0] = shape.flavour; st[1] = grid; st[2] = shape.flavour == FLAVOUR_TMA ? kTmaThreads : block; st[11] = args.h.pdl;
        for (uint32_t i = 0; i < args.h.nops; ++i) {
            const OpDesc& d = args.ops[i];
            switch (d.kind) {
                case OP_PUSH: st[3]++; st[4] += d.nblocks; break;
                case OP_RECV_ADAPT: st[6]++; st[7] += d.nblocks; break;
                case OP_SEND_LL: case OP_RECV_LL: st[8]++; break;
                case OP_COPY: st[9] += d.nblocks; break;
                default: break;
            }
        }
        for (size_t i = 0; i < ops.size(); ++i) {
            if (ops[i].is_send && res[i].proto == PROTO_LARGE && res[i].pull_send && !res[i].ce) st[5]++;
            if (res[i].ce) st[10]++;
        }
    }
    if (g.plan_only) {  // CPU tests of the host logic: no device, nothing to launch
        g.last_args = args;
        g.last_grid = grid;
        g.last_block = block;
        g.last_flavour = shape.flavour;
        return P2P_SUCCESS;
    }
    return submit(ops, res, shape, args, grid, block);
}

int enqueue(bool is_send, const void* buf, size_t count, int dtype, int peer, cudaStream_t stream) {
    int rc = check_ready();
    if (rc) return rc;
    if (dtype < 0 || dtype >= P2P_NUM_TYPES) return fail(P2P_INVALID_ARGUMENT, "unknown datatype %d", dtype);
    if (peer < 0 || peer >= g.nranks) return fail(P2P_INVALID_ARGUMENT, "peer %d out of range [0,%d)", peer, g.nranks);
    const size_t es = kDtypeSize[dtype];
    if (count > (SIZE_MAX / es)) return fail(P2P_INVALID_ARGUMENT, "count %zu overflows size_t", count);
    const size_t nbytes = count * es;
    if (nbytes > kPostOffMask) return fail(P2P_INVALID_ARGUMENT, "message of %zu bytes exceeds the 512 GiB limit", nbytes);
    if (nbytes && !buf) return fail(P2P_INVALID_ARGUMENT, "null buffer with count %zu", count);
    std::lock_guard<std::mutex> lk(g.mu);
    // after a device-side timeout counters and flag rings are no longer consistent: nothing more is issued
    if (g.async_err_host && *(volatile uint32_t*)g.async_err_host != AERR_NONE)
        return fail(P2P_REMOTE_ERROR, "the engine is in an error state (device code %u, see p2p_get_async_error); "
                    "no further transfers are issued until p2p_finalize", *(volatile uint32_t*)g.async_err_host);
    g.group_ops.push_back(PendingOp{is_send, buf, nbytes, peer, stream});
    if (g.group_depth == 0) return launch_pending();
    return P2P_SUCCESS;
}

void* ipc_open(int peer, const cudaIpcMemHandle_t& h, std::string* key_out, cudaError_t* err) {
    std::string key((const char*)&h, sizeof h);
    auto& cache = g.ipc_cache[peer];
    auto it = cache.find(key);
    if (it != cache.end()) {
        it->second.refs++;
        *key_out = key;
        *err = cudaSuccess;
        return it->second.ptr;
    }
    void* ptr = nullptr;
    *err = cudaIpcOpenMemHandle(&ptr, h, cudaIpcMemLazyEnablePeerAccess);
    if (*err != cudaSuccess) return nullptr;
    cache[key] = IpcMapping{ptr, 1};
    *key_out = key;
    return ptr;
}

void ipc_close(int peer, const std::string& key) {
    auto& cache = g.ipc_cache[peer];
    auto it = cache.find(key);
    if (it == cache.end()) return;
    if (--it->second.refs == 0) {
        cudaIpcCloseMemHandle(it->second.ptr);
        cache.erase(it);
    }
}

typedef int (*cuMemGetAddressRange_t)(unsigned long long* pbase, size_t* psize, unsigned long long dptr);

int allocation_range(void* ptr, uint8_t** base, size_t* size) {
    static cuMemGetAddressRange_t fn = nullptr;
    if (!fn) {
        void* sym = nullptr;
        cudaDriverEntryPointQueryResult q;
        cudaError_t e = cudaGetDriverEntryPoint("cuMemGetAddressRange", &sym, cudaEnableDefault, &q);
        if (e != cudaSuccess || q != cudaDriverEntryPointSuccess || !sym)
            return fail(P2P_SYSTEM_ERROR, "cannot resolve cuMemGetAddressRange from the driver");
        fn = (cuMemGetAddressRange_t)sym;
    }
    unsigned long long b = 0;
    size_t s = 0;
    int r = fn(&b, &s, (unsigned long long)ptr);
    if (r != 0) return fail(P2P_INVALID_ARGUMENT, "%p is not device memory known to the driver (cuMemGetAddressRange=%d)", ptr, r);
    *base = (uint8_t*)b;
    *size = s;
    return P2P_SUCCESS;
}

}  // namespace
}  // namespace p2p

namespace p2p {

int engine_info(EngineInfo* out) {
    std::lock_guard<std::mutex> lk(g.mu);
    int rc = check_ready();
    if (rc) return rc;
    if (g.plan_only) return fail(P2P_INVALID_USAGE, "not available in plan-only mode");
    out->rank = g.rank;
    out->nranks = g.nranks;
    out->device = g.device;
    out->sm_count = g.sm_count;
    out->async_err = g.async_err_dev;
    out->timeout_ns = (uint64_t)g.opt.timeout_ms * 1000000ull;
    return P2P_SUCCESS;
}

int engine_add_window(uint8_t* base, size_t bytes, uint8_t* const* peer_bases, int* window_id) {
    std::lock_guard<std::mutex> lk(g.mu);
    int rc = check_ready();
    if (rc) return rc;
    int w = -1;
    for (int i = 0; i < P2P_MAX_WINDOWS; ++i)
        if (!g.windows[i].used) { w = i; break; }
    if (w < 0) return fail(P2P_INVALID_USAGE, "all %d windows are in use", P2P_MAX_WINDOWS);
    Window W;
    W.used = true;
    W.base = base;
    W.bytes = bytes;
    for (int r = 0; r < g.nranks; ++r) {
        W.peer_base[r] = peer_bases[r];
        const uint64_t b = (uint64_t)peer_bases[r];
        CUDA_TRY(cudaMemcpy(g.win_table + (size_t)r * P2P_MAX_WINDOWS + w, &b, sizeof b, cudaMemcpyHostToDevice));
    }
    g.windows[w] = W;
    *window_id = w;
    return P2P_SUCCESS;
}

}  // namespace p2p

using namespace p2p;

// ======================================================================== C ABI

extern "C" {

int p2p_init(int rank, int nranks, int device) {
    std::lock_guard<std::mutex> lk(g.mu);
    if (g.inited) return fail(P2P_INVALID_USAGE, "p2p_init called twice");
    if (nranks < 1 || nranks > P2P_MAX_RANKS)
        return fail(P2P_INVALID_ARGUMENT, "nranks %d outside [1,%d]", nranks, P2P_MAX_RANKS);
    if (rank < 0 || rank >= nranks) return fail(P2P_INVALID_ARGUMENT, "rank %d outside [0,%d)", rank, nranks);
    int ndev = 0;
    cudaError_t e = cudaGetDeviceCount(&ndev);
    if (e != cudaSuccess || ndev == 0)
        return fail(P2P_UNHANDLED_CUDA_ERROR, "no CUDA device available (%s); this engine has no CPU path",
                    cudaGetErrorString(e));
    if (device < 0 || device >= ndev) return fail(P2P_INVALID_ARGUMENT, "device %d outside [0,%d)", device, ndev);
    CUDA_TRY(cudaSetDevice(device));
    cudaDeviceProp prop;
    CUDA_TRY(cudaGetDeviceProperties(&prop, device));
    if (prop.major != 10)
        return fail(P2P_INVALID_USAGE, "device %d is sm_%d%d; this library contains sm_100a code only", device,
                    prop.major, prop.minor);
    g.rank = rank;
    g.nranks = nranks;
    g.device = device;
    g.sm_count = prop.multiProcessorCount;
    g.host_hash = host_hash_now();

    Options& o = g.opt;
    o = Options();
    {
        int rc = read_env_options(o);
        if (rc) return rc;
    }

    resolve_stage_chunk(o, nranks);
    // control region: per-peer flag sections, then the eager rings
    g.ll_slot_bytes = (size_t)std::max<long long>(o.eager_bytes, 8) * 2;
    g.ll_slot_bytes = (g.ll_slot_bytes + 255) & ~(size_t)255;
    g.ll_off = (sizeof(PeerCtl) * (size_t)nranks + 4095) & ~(size_t)4095;
    g.stage_off = (g.ll_off + (size_t)nranks * kLLSlots * g.ll_slot_bytes + 4095) & ~(size_t)4095;
    g.ctl_bytes = g.stage_off + (size_t)nranks * stage_area_bytes();
    CUDA_TRY(cudaMalloc(&g.ctl, g.ctl_bytes));
    CUDA_TRY(cudaMemset(g.ctl, 0, g.ctl_bytes));
    CUDA_TRY(cudaMalloc(&g.arrive, sizeof(uint32_t) * arrive_words(nranks)));
    CUDA_TRY(cudaMemset(g.arrive, 0, sizeof(uint32_t) * arrive_words(nranks)));
    CUDA_TRY(cudaMalloc(&g.stage_table, sizeof(StageInfo) * (size_t)nranks));
    CUDA_TRY(cudaMemset(g.stage_table, 0, sizeof(StageInfo) * (size_t)nranks));
    CUDA_TRY(cudaMalloc(&g.win_table, sizeof(uint64_t) * (size_t)nranks * P2P_MAX_WINDOWS));
    CUDA_TRY(cudaMemset(g.win_table, 0, sizeof(uint64_t) * (size_t)nranks * P2P_MAX_WINDOWS));
    CUDA_TRY(cudaMalloc(&g.tokens, sizeof(uint32_t) * kTokenSlots));
    CUDA_TRY(cudaMemset(g.tokens, 0, sizeof(uint32_t) * kTokenSlots));
    g.launch_counter = 0;
    g.stream_last.clear();
    CUDA_TRY(cudaMalloc(&g.chk_partials, sizeof(CheckPartial) * (size_t)g.sm_count * 8));
    CUDA_TRY(cudaMalloc(&g.chk_ticket, sizeof(unsigned int)));
    CUDA_TRY(cudaMemset(g.chk_ticket, 0, sizeof(unsigned int)));
    CUDA_TRY(cudaMalloc(&g.chk_out, sizeof(CheckPartial)));
    CUDA_TRY(cudaHostAlloc(&g.chk_out_host, sizeof(CheckPartial), cudaHostAllocDefault));
    CUDA_TRY(cudaHostAlloc(&g.async_err_host, sizeof(uint32_t), cudaHostAllocMapped));
    *g.async_err_host = 0;
    CUDA_TRY(cudaHostGetDevicePointer((void**)&g.async_err_dev, g.async_err_host, 0));
    CUDA_TRY(cudaEventCreateWithFlags(&g.ev_done, cudaEventDisableTiming));
    CUDA_TRY(configure_kernels(200 * 1024));
    CUDA_TRY(occupancy(FLAVOUR_LDST, 0, &g.ctas_ldst));
    CUDA_TRY(occupancy(FLAVOUR_TMA, (int)std::max(o.tma_chunk * o.tma_stages, o.tma_peer_chunk * o.tma_peer_stages), &g.ctas_tma));
    if (g.ctas_ldst < 1) g.ctas_ldst = 1;
    if (g.ctas_tma < 1) g.ctas_tma = 1;
    CUDA_TRY(cudaDeviceSynchronize());
    g.peer_ctl[rank] = g.ctl;
    g.hostbox = nullptr;
    g.hostbox_name[0] = '\0';
    g.hostbox_bytes = sizeof(HostPost) * (size_t)nranks * nranks * kRing;
    if (nranks > 1 && rank == 0 && hostbox_create() != P2P_SUCCESS) {
        // only the opt-in copy-engine path needs the mailbox; without /dev/shm everything else still works
        g.hostbox = nullptr;
        g.hostbox_name[0] = '\0';
    }
    CUDA_TRY(cudaEventCreateWithFlags(&g.ce_fork, cudaEventDisableTiming));
    CUDA_TRY(cudaHostAlloc(&g.ce_vals, sizeof(uint64_t) * 4096, cudaHostAllocDefault));
    g.ce_vals_next = 0;
    g.stat_launches = g.stat_ce_copies = 0;
    g.inited = true;
    g.connected = false;
    return P2P_SUCCESS;
}

size_t p2p_blob_bytes(void) { return kBlobBytes; }
size_t p2p_register_blob_bytes(void) { return kRegBlobBytes; }

int p2p_export(void* blob) {
    std::lock_guard<std::mutex> lk(g.mu);
    if (!g.inited) return fail(P2P_INVALID_USAGE, "p2p_init has not been called");
    if (!blob) return fail(P2P_INVALID_ARGUMENT, "null blob");
    memset(blob, 0, kBlobBytes);
    BootBlob b;
    memset(&b, 0, sizeof b);
    b.magic = kBlobMagic;
    b.version = kBlobVersion;
    b.rank = g.rank;
    b.nranks = g.nranks;
    b.host_hash = g.host_hash;
    b.pid = (int32_t)getpid();
    b.device = g.device;
    b.eager_bytes = (uint64_t)g.opt.eager_bytes;
    b.pull_min_bytes = g.opt.pull_min_bytes;
    b.ctl_bytes = g.ctl_bytes;
    b.stage_chunk = (uint32_t)g.opt.stage_chunk;
    b.stage_lanes = (uint32_t)g.opt.stage_lanes;
    CUDA_TRY(cudaDeviceGetPCIBusId(b.pci_bus_id, sizeof b.pci_bus_id, g.device));
    memcpy(b.hostbox_name, g.hostbox_name, sizeof b.hostbox_name);
    if (g.nranks > 1) CUDA_TRY(cudaIpcGetMemHandle(&b.ctl_handle, g.ctl));
    memcpy(blob, &b, sizeof b);
    return P2P_SUCCESS;
}

int p2p_connect(const void* all_blobs) {
    std::lock_guard<std::mutex> lk(g.mu);
    if (!g.inited) return fail(P2P_INVALID_USAGE, "p2p_init has not been called");
    if (g.connected) return fail(P2P_INVALID_USAGE, "p2p_connect called twice");
    if (!all_blobs) return fail(P2P_INVALID_ARGUMENT, "null blobs");
    for (int r = 0; r < g.nranks; ++r) {
        BootBlob b;
        memcpy(&b, (const uint8_t*)all_blobs + (size_t)r * kBlobBytes, sizeof b);
        if (b.magic != kBlobMagic || b.version != kBlobVersion)
            return fail(P2P_INVALID_ARGUMENT, "blob %d is not a p2p bootstrap blob of this library version", r);
        if (b.rank != r || b.nranks != g.nranks)
            return fail(P2P_INVALID_ARGUMENT, "blob %d says rank %d of %d; blobs must be in rank order", r, b.rank, b.nranks);
        if (b.host_hash != g.host_hash)
            return fail(P2P_INVALID_USAGE,
                        "rank %d runs on another host; NVLink peer mappings do not cross hosts (single-host engine)", r);
        if (b.eager_bytes != (uint64_t)g.opt.eager_bytes)
            return fail(P2P_INVALID_USAGE, "rank %d uses eager_bytes=%llu, this rank %lld; they must match", r,
                        (unsigned long long)b.eager_bytes, g.opt.eager_bytes);
        if (b.stage_chunk != (uint32_t)g.opt.stage_chunk || b.stage_lanes != (uint32_t)g.opt.stage_lanes)
            return fail(P2P_INVALID_USAGE, "rank %d uses a staging ring of %u lanes x %u bytes, this rank %lld x %lld; they must match", r,
                        b.stage_lanes, b.stage_chunk, g.opt.stage_lanes, g.opt.stage_chunk);
        if (b.pull_min_bytes != g.opt.pull_min_bytes)
            return fail(P2P_INVALID_USAGE, "rank %d uses pull_min_bytes=%lld, this rank %lld; they must match", r,
                        (long long)b.pull_min_bytes, g.opt.pull_min_bytes);
        b.pci_bus_id[sizeof b.pci_bus_id - 1] = '\0';
        for (int q = 0; q < r; ++q) {
            BootBlob bq;
            memcpy(&bq, (const uint8_t*)all_blobs + (size_t)q * kBlobBytes, sizeof bq);
            bq.pci_bus_id[sizeof bq.pci_bus_id - 1] = '\0';
            if (b.pci_bus_id[0] && !strcmp(b.pci_bus_id, bq.pci_bus_id))
                return fail(P2P_INVALID_USAGE,
                            "ranks %d and %d both drive GPU %s: the engine is one process per GPU (kernels of two ranks that "
                            "wait for each other's flags cannot share a GPU); launch at most as many ranks as there are GPUs",
                            q, r, b.pci_bus_id);
        }
        if (r == g.rank) continue;
        if (b.ctl_bytes != g.ctl_bytes)
            return fail(P2P_INTERNAL_ERROR, "rank %d control region is %llu bytes, mine %zu", r,
                        (unsigned long long)b.ctl_bytes, g.ctl_bytes);
        if (b.pid == (int32_t)getpid())
            return fail(P2P_INVALID_USAGE, "rank %d lives in this process; the engine is one process per GPU", r);
        std::string key;
        cudaError_t e;
        void* p = ipc_open(r, b.ctl_handle, &key, &e);
        if (!p)
            return fail(P2P_UNHANDLED_CUDA_ERROR,
                        "cudaIpcOpenMemHandle for rank %d (pid %d, GPU %s) failed: %s — peers must share a PID/IPC "
                        "namespace and be P2P-capable",
                        r, b.pid, b.pci_bus_id, cudaGetErrorString(e));
        g.peer_ctl[r] = (uint8_t*)p;
    }
    if (g.nranks > 1 && !g.hostbox) {
        BootBlob b0;
        memcpy(&b0, all_blobs, sizeof b0);
        b0.hostbox_name[sizeof b0.hostbox_name - 1] = '\0';
        // without it only the copy-engine path is lost; every other path works
        if (b0.hostbox_name[0] && hostbox_open(b0.hostbox_name) != P2P_SUCCESS) g.hostbox = nullptr;
    }
    if (!g.plan_only) {
        std::vector<StageInfo> table((size_t)g.nranks);
        for (int r = 0; r < g.nranks; ++r) table[(size_t)r] = stage_info_for(r);
        CUDA_TRY(cudaMemcpy(g.stage_table, table.data(), sizeof(StageInfo) * table.size(), cudaMemcpyHostToDevice));
    }
    g.connected = true;
    return P2P_SUCCESS;
}

// Teardown in two steps, because CUDA IPC wants every importer to close its mapping before the exporter
// frees the memory: p2p_disconnect (drain, close every mapping of peer memory), a barrier supplied by the
// caller, then p2p_finalize (free this rank's own regions).
static int disconnect_locked() {
    if (!g.inited || g.plan_only) return P2P_SUCCESS;
    cudaSetDevice(g.device);
    cudaDeviceSynchronize();
    for (int w = 0; w < P2P_MAX_WINDOWS; ++w) {
        Window& W = g.windows[w];
        if (!W.used) continue;
        for (int r = 0; r < g.nranks; ++r)
            if (r != g.rank && !W.peer_handle[r].empty()) ipc_close(r, W.peer_handle[r]);
        W = Window();
    }
    for (int r = 0; r < g.nranks; ++r) {
        for (auto& kv : g.ipc_cache[r]) cudaIpcCloseMemHandle(kv.second.ptr);
        g.ipc_cache[r].clear();
        if (r != g.rank) g.peer_ctl[r] = nullptr;
    }
    g.connected = false;
    return P2P_SUCCESS;
}

int p2p_disconnect(void) {
    std::lock_guard<std::mutex> lk(g.mu);
    return disconnect_locked();
}

int p2p_finalize(void) {
    std::lock_guard<std::mutex> lk(g.mu);
    if (!g.inited) return P2P_SUCCESS;
    if (g.plan_only) {  // nothing was allocated
        for (int w = 0; w < P2P_MAX_WINDOWS; ++w) g.windows[w] = Window();
        for (int r = 0; r < P2P_MAX_RANKS; ++r) {
            g.peer_ctl[r] = nullptr;
            g.rdv_send_seq[r] = g.rdv_recv_seq[r] = g.ll_send_seq[r] = g.ll_recv_seq[r] = 0;
        }
        g.ctl = nullptr;
        g.arrive = nullptr;
        g.win_table = nullptr;
        g.stage_table = nullptr;
        g.tokens = nullptr;
        g.stream_last.clear();
        g.launch_counter = 0;
        g.group_ops.clear();
        g.group_depth = 0;
        g.plan_only = g.inited = g.connected = false;
        return P2P_SUCCESS;
    }
    disconnect_locked();  // a no-op when p2p_disconnect has run
    if (!g.ce_timing.empty()) {
        cudaDeviceSynchronize();
        double hs = 0, dma = 0, gap = 0;
        size_t n = 0, ngap = 0;
        for (size_t i = g.ce_timing.size() / 4; i < g.ce_timing.size(); ++i) {  // skip the warm-up quarter
            float a = 0, b = 0;
            if (cudaEventElapsedTime(&a, g.ce_timing[i][0], g.ce_timing[i][1]) != cudaSuccess) continue;
            if (cudaEventElapsedTime(&b, g.ce_timing[i][1], g.ce_timing[i][2]) != cudaSuccess) continue;
            hs += a; dma += b; ++n;
            float c = 0;
            if (i + 1 < g.ce_timing.size() && cudaEventElapsedTime(&c, g.ce_timing[i][2], g.ce_timing[i + 1][0]) == cudaSuccess) { gap += c; ++ngap; }
        }
        if (n) printf("[rank %d] copy-engine chains: handshake %.1f us, DMA %.1f us, from DMA end to the next chain's start %.1f us (mean of %zu)\n",
                      g.rank, hs / n * 1e3, dma / n * 1e3, ngap ? gap / ngap * 1e3 : 0.0, n);
        for (auto& t : g.ce_timing) for (cudaEvent_t e : t) cudaEventDestroy(e);
        g.ce_timing.clear();
    }
    symm_finalize(g.nranks);
    for (int r = 0; r < P2P_MAX_RANKS; ++r) {
        g.peer_ctl[r] = nullptr;
        g.rdv_send_seq[r] = g.rdv_recv_seq[r] = g.ll_send_seq[r] = g.ll_recv_seq[r] = 0;
    }
    if (g.hostbox) munmap(g.hostbox, g.hostbox_bytes);
    g.hostbox = nullptr;
    if (g.rank == 0 && g.hostbox_name[0]) shm_unlink(g.hostbox_name);
    g.hostbox_name[0] = '\0';
    for (int r = 0; r < P2P_MAX_RANKS; ++r) {
        if (g.ce_stream[r]) cudaStreamDestroy(g.ce_stream[r]);
        if (g.ce_join[r]) cudaEventDestroy(g.ce_join[r]);
        g.ce_stream[r] = nullptr;
        g.ce_join[r] = nullptr;
    }
    if (g.ce_fork) cudaEventDestroy(g.ce_fork);
    g.ce_fork = nullptr;
    if (g.ce_vals) cudaFreeHost(g.ce_vals);
    g.ce_vals = nullptr;
    cudaFree(g.ctl);
    cudaFree(g.arrive);
    cudaFree(g.win_table);
    cudaFree(g.stage_table);
    g.stage_table = nullptr;
    cudaFree(g.tokens);
    g.tokens = nullptr;
    g.stream_last.clear();
    cudaFree(g.chk_partials);
    cudaFree(g.chk_ticket);
    cudaFree(g.chk_out);
    cudaFreeHost(g.chk_out_host);
    cudaFreeHost(g.async_err_host);
    g.async_err_host = nullptr;
    g.async_err_dev = nullptr;
    if (g.ev_done) cudaEventDestroy(g.ev_done);
    for (cudaEvent_t e : g.ev_pool) cudaEventDestroy(e);
    g.ev_pool.clear();
    g.ev_done = nullptr;
    g.ctl = nullptr;
    g.group_ops.clear();
    g.group_depth = 0;
    g.has_pending_reg = false;
    g.inited = g.connected = false;
    return P2P_SUCCESS;
}

int p2p_rank(void) { return g.inited ? g.rank : -1; }
int p2p_nranks(void) { return g.inited ? g.nranks : 0; }

int p2p_register_export(void* buf, size_t bytes, void* blob) {
    std::lock_guard<std::mutex> lk(g.mu);
    int rc = check_ready();
    if (rc) return rc;
    if (!blob) return fail(P2P_INVALID_ARGUMENT, "null blob");
    // an export that was never imported (a peer failed in between) is simply superseded
    memset(blob, 0, kRegBlobBytes);
    RegBlob b;
    memset(&b, 0, sizeof b);
    b.magic = kBlobMagic;
    if (buf && bytes) {
        uint8_t* base = nullptr;
        size_t asize = 0;
        rc = allocation_range(buf, &base, &asize);
        if (rc) return rc;
        if ((uint8_t*)buf + bytes > base + asize)
            return fail(P2P_INVALID_ARGUMENT, "window %p+%zu runs past its allocation (%p+%zu)", buf, bytes, base, asize);
        b.has = 1;
        b.bytes = bytes;
        b.offset = (uint64_t)((uint8_t*)buf - base);
        if (g.nranks > 1) CUDA_TRY(cudaIpcGetMemHandle(&b.handle, base));
    }
    memcpy(blob, &b, sizeof b);
    g.pending_reg_buf = b.has ? buf : nullptr;
    g.pending_reg_bytes = b.has ? bytes : 0;
    g.has_pending_reg = true;
    return P2P_SUCCESS;
}

int p2p_register_import(const void* all_blobs, int* window_id) {
    std::lock_guard<std::mutex> lk(g.mu);
    int rc = check_ready();
    if (rc) return rc;
    if (!g.has_pending_reg) return fail(P2P_INVALID_USAGE, "p2p_register_import without p2p_register_export");
    if (!all_blobs || !window_id) return fail(P2P_INVALID_ARGUMENT, "null argument");
    int w = -1;
    for (int i = 0; i < P2P_MAX_WINDOWS; ++i)
        if (!g.windows[i].used) { w = i; break; }
    if (w < 0) return fail(P2P_INVALID_USAGE, "all %d windows are in use", P2P_MAX_WINDOWS);
    Window W;
    W.used = true;
    W.base = (uint8_t*)g.pending_reg_buf;
    W.bytes = g.pending_reg_bytes;
    g.has_pending_reg = false;
    std::vector<uint64_t> bases((size_t)g.nranks, 0);
    for (int r = 0; r < g.nranks; ++r) {
        RegBlob b;
        memcpy(&b, (const uint8_t*)all_blobs + (size_t)r * kRegBlobBytes, sizeof b);
        if (b.magic != kBlobMagic) return fail(P2P_INVALID_ARGUMENT, "register blob %d is malformed", r);
        if (!b.has) continue;
        if (r == g.rank) {
            W.peer_base[r] = W.base;
        } else {
            cudaError_t e;
            void* p = ipc_open(r, b.handle, &W.peer_handle[r], &e);
            if (!p) {
                for (int q = 0; q < r; ++q)
                    if (q != g.rank && !W.peer_handle[q].empty()) ipc_close(q, W.peer_handle[q]);
                return fail(P2P_UNHANDLED_CUDA_ERROR, "cudaIpcOpenMemHandle for rank %d's window failed: %s", r,
                            cudaGetErrorString(e));
            }
            W.peer_base[r] = (uint8_t*)p + b.offset;
        }
        bases[(size_t)r] = (uint64_t)W.peer_base[r];
    }
    // device table: row r, column w = where rank r's part of window w is mapped here
    for (int r = 0; r < g.nranks; ++r)
        CUDA_TRY(cudaMemcpy(g.win_table + (size_t)r * P2P_MAX_WINDOWS + w, &bases[(size_t)r], sizeof(uint64_t),
                            cudaMemcpyHostToDevice));
    g.windows[w] = W;
    *window_id = w;
    return P2P_SUCCESS;
}

int p2p_deregister(int window_id) {
    std::lock_guard<std::mutex> lk(g.mu);
    int rc = check_ready();
    if (rc) return rc;
    if (window_id < 0 || window_id >= P2P_MAX_WINDOWS || !g.windows[window_id].used)
        return fail(P2P_INVALID_ARGUMENT, "window %d is not registered", window_id);
    CUDA_TRY(cudaDeviceSynchronize());
    Window& W = g.windows[window_id];
    const uint64_t zero = 0;
    for (int r = 0; r < g.nranks; ++r) {
        if (r != g.rank && !W.peer_handle[r].empty()) ipc_close(r, W.peer_handle[r]);
        CUDA_TRY(cudaMemcpy(g.win_table + (size_t)r * P2P_MAX_WINDOWS + window_id, &zero, sizeof zero,
                            cudaMemcpyHostToDevice));
    }
    W = Window();
    return P2P_SUCCESS;
}

int p2p_window_peer_ptr(int window_id, int peer, void** ptr) {
    std::lock_guard<std::mutex> lk(g.mu);
    int rc = check_ready();
    if (rc) return rc;
    if (!ptr) return fail(P2P_INVALID_ARGUMENT, "null argument");
    if (window_id < 0 || window_id >= P2P_MAX_WINDOWS || !g.windows[window_id].used)
        return fail(P2P_INVALID_ARGUMENT, "window %d is not registered", window_id);
    if (peer < 0 || peer >= g.nranks) return fail(P2P_INVALID_ARGUMENT, "peer %d out of range [0,%d)", peer, g.nranks);
    *ptr = g.windows[window_id].peer_base[peer];
    return P2P_SUCCESS;
}

int p2p_mem_alloc(void** ptr, size_t bytes) {
    if (!g.inited) return fail(P2P_INVALID_USAGE, "p2p_init has not been called");
    if (!ptr) return fail(P2P_INVALID_ARGUMENT, "null ptr");
    CUDA_TRY(cudaSetDevice(g.device));
    // whole 2 MiB pages from 2 MiB up: an allocation of any other size gets a slower peer mapping (copy engine both
    // ways into a 256 MiB + 256 B buffer: 1084 GB/s, into a 256 MiB one: 1521)
    if (bytes >= (2u << 20)) bytes = (bytes + (2u << 20) - 1) & ~(size_t)((2u << 20) - 1);
    CUDA_TRY(cudaMalloc(ptr, bytes ? bytes : 1));
    return P2P_SUCCESS;
}

int p2p_mem_free(void* ptr) {
    if (!g.inited) return fail(P2P_INVALID_USAGE, "p2p_init has not been called");
    CUDA_TRY(cudaFree(ptr));
    return P2P_SUCCESS;
}

int p2p_send(const void* buf, size_t count, int dtype, int peer, p2p_stream_t stream) {
    return enqueue(true, buf, count, dtype, peer, (cudaStream_t)stream);
}

int p2p_recv(void* buf, size_t count, int dtype, int peer, p2p_stream_t stream) {
    return enqueue(false, buf, count, dtype, peer, (cudaStream_t)stream);
}

int p2p_alltoall(const void* sendbuf, void* recvbuf, size_t count, int dtype, p2p_stream_t stream) {
    int rc = check_ready();
    if (rc) return rc;
    if (dtype < 0 || dtype >= P2P_NUM_TYPES) return fail(P2P_INVALID_ARGUMENT, "unknown datatype %d", dtype);
    const size_t es = kDtypeSize[dtype];
    if (count > SIZE_MAX / es / (size_t)g.nranks) return fail(P2P_INVALID_ARGUMENT, "count %zu overflows size_t", count);
    const size_t block = count * es;
    if (block && (!sendbuf || !recvbuf)) return fail(P2P_INVALID_ARGUMENT, "null buffer with count %zu", count);
    if (2 * g.nranks - 1 > kMaxGroupOps) return fail(P2P_INVALID_USAGE, "all-to-all over %d ranks needs more than %d operations in a group", g.nranks, kMaxGroupOps);
    rc = p2p_group_start();
    if (rc) return rc;
    for (int p = 0; p < g.nranks && rc == P2P_SUCCESS; ++p) {
        rc = enqueue(false, (uint8_t*)recvbuf + (size_t)p * block, count, dtype, p, (cudaStream_t)stream);
        if (rc == P2P_SUCCESS) rc = enqueue(true, (const uint8_t*)sendbuf + (size_t)p * block, count, dtype, p, (cudaStream_t)stream);
    }
    if (rc != P2P_SUCCESS) {
        // drop what was queued and close the group without launching anything
        std::lock_guard<std::mutex> lk(g.mu);
        g.group_ops.clear();
        g.group_depth = 0;
        return rc;
    }
    return p2p_group_end();
}

int p2p_group_start(void) {
    std::lock_guard<std::mutex> lk(g.mu);
    int rc = check_ready();
    if (rc) return rc;
    g.group_depth++;
    return P2P_SUCCESS;
}

int p2p_group_end(void) {
    std::lock_guard<std::mutex> lk(g.mu);
    int rc = check_ready();
    if (rc) return rc;
    if (g.group_depth == 0) return fail(P2P_INVALID_USAGE, "p2p_group_end without p2p_group_start");
    if (--g.group_depth > 0) return P2P_SUCCESS;
    return launch_pending();
}

const char* p2p_get_error_string(int result) {
    switch (result) {
        case P2P_SUCCESS: return "no error";
        case P2P_UNHANDLED_CUDA_ERROR: return "unhandled cuda error";
        case P2P_SYSTEM_ERROR: return "unhandled system error";
        case P2P_INTERNAL_ERROR: return "internal error";
        case P2P_INVALID_ARGUMENT: return "invalid argument";
        case P2P_INVALID_USAGE: return "invalid usage";
        case P2P_REMOTE_ERROR: return "remote peer did not respond";
        case P2P_IN_PROGRESS: return "operation in progress";
        default: return "unknown result code";
    }
}

const char* p2p_get_last_error(void) { return t_last_error; }

int p2p_get_async_error(int* result) {
    if (!g.inited) return fail(P2P_INVALID_USAGE, "p2p_init has not been called");
    if (!result) return fail(P2P_INVALID_ARGUMENT, "null result");
    const uint32_t code = *(volatile uint32_t*)g.async_err_host;
    if (code == AERR_NONE) {
        *result = P2P_SUCCESS;
    } else {
        static const char* const what[] = {"", "send: receiver never posted", "recv: payload never completed",
                                           "eager recv: data never arrived", "eager send: ring slot never freed",
                                           "send: receiver posted an unknown window",
                                           "staging ring: the other side never filled / drained a chunk",
                                           "staging ring of this pair still in use by an earlier message (one unregistered "
                                           "transfer per ordered pair at a time)"};
        *result = code == AERR_BAD_WINDOW ? P2P_INVALID_USAGE : P2P_REMOTE_ERROR;
        fail(*result, "device reported: %s (code %u, timeout %lld ms)", code < 8 ? what[code] : "?", code,
             g.opt.timeout_ms);
    }
    return P2P_SUCCESS;
}

int p2p_set_option(const char* name, long long value) {
    std::lock_guard<std::mutex> lk(g.mu);
    if (!g.inited) return fail(P2P_INVALID_USAGE, "p2p_init has not been called");
    if (!name) return fail(P2P_INVALID_ARGUMENT, "null option name");
    const OptionSpec* sp = find_option(name);
    if (!sp) return fail(P2P_INVALID_ARGUMENT, "unknown option '%s'", name);
    if (!strcmp(name, "pull_min_bytes"))
        return fail(P2P_INVALID_USAGE, "pull_min_bytes is agreed between ranks at p2p_connect; set P2P_PULL_MIN_BYTES before p2p_init");
    if (!strcmp(name, "stage_chunk") || !strcmp(name, "stage_lanes"))
        return fail(P2P_INVALID_USAGE, "%s shapes the staging ring both ends share; set P2P_STAGE_CHUNK / P2P_STAGE_LANES before p2p_init (same on all ranks)", name);
    int rc = validate_option(*sp, value, g.opt, "");
    if (rc) return rc;
    // the rings were sized in p2p_init; the limit can only shrink afterwards, and on every rank alike
    if (!strcmp(name, "eager_bytes") && (size_t)value * 2 > g.ll_slot_bytes)
        return fail(P2P_INVALID_ARGUMENT, "eager_bytes %lld exceeds the ring slot (%zu bytes); set P2P_EAGER_BYTES before p2p_init",
                    value, g.ll_slot_bytes / 2);
    g.opt.*(sp->slot) = value;
    if (!strncmp(name, "tma_", 4) && !g.plan_only) {
        int n = 1;
        const long long smem = std::max(g.opt.tma_chunk * g.opt.tma_stages, g.opt.tma_peer_chunk * g.opt.tma_peer_stages);
        if (occupancy(FLAVOUR_TMA, (int)smem, &n) == cudaSuccess && n >= 1) g.ctas_tma = n;
    }
    return P2P_SUCCESS;
}

int p2p_get_option(const char* name, long long* value) {
    std::lock_guard<std::mutex> lk(g.mu);
    if (!g.inited) return fail(P2P_INVALID_USAGE, "p2p_init has not been called");
    if (!name || !value) return fail(P2P_INVALID_ARGUMENT, "null argument");
    if (!strcmp(name, "sm_count")) { *value = g.sm_count; return P2P_SUCCESS; }
    if (!strcmp(name, "stat_launches")) { *value = g.stat_launches; return P2P_SUCCESS; }
    if (!strcmp(name, "stat_ce_copies")) { *value = g.stat_ce_copies; return P2P_SUCCESS; }
    {
        static const char* const names[12] = {"stat_last_flavour", "stat_last_grid", "stat_last_block", "stat_last_push_ops",
                                              "stat_last_push_ctas", "stat_last_offer_ops", "stat_last_recv_ops", "stat_last_recv_ctas",
                                              "stat_last_eager_ops", "stat_last_copy_ctas", "stat_last_ce_ops", "stat_last_pdl"};
        for (int i = 0; i < 12; ++i)
            if (!strcmp(name, names[i])) { *value = g.last_stat[i]; return P2P_SUCCESS; }
    }
    if (!strcmp(name, "ctas_per_sm_ldst")) { *value = g.ctas_ldst; return P2P_SUCCESS; }
    if (!strcmp(name, "ctas_per_sm_tma")) { *value = g.ctas_tma; return P2P_SUCCESS; }
    const OptionSpec* sp = find_option(name);
    if (!sp) return fail(P2P_INVALID_ARGUMENT, "unknown option '%s'", name);
    *value = g.opt.*(sp->slot);
    return P2P_SUCCESS;
}

int p2p_query_plan(size_t bytes, int peer, int* path, int* nblocks) {
    std::lock_guard<std::mutex> lk(g.mu);
    int rc = check_ready();
    if (rc) return rc;
    if (peer < 0 || peer >= g.nranks) return fail(P2P_INVALID_ARGUMENT, "peer %d out of range", peer);
    const Proto proto = proto_for(bytes);
    const bool self = peer == g.rank;
    int pth, nb;
    if (proto == PROTO_EAGER) { pth = 4; nb = eager_ctas(bytes); }
    else if (proto == PROTO_LARGE && g.opt.path == 0 && (!self || g.opt.loopback)) { pth = 5; nb = mover_ctas(bytes, FLAVOUR_LDST, self, true); }
    else {
        pth = (self && !g.opt.loopback) ? self_copy_flavour(bytes) : push_flavour(bytes);
        nb = mover_ctas(bytes, pth == 2 ? FLAVOUR_TMA : FLAVOUR_LDST, self, false);
    }
    if (path) *path = pth;
    if (nblocks) *nblocks = nb;
    return P2P_SUCCESS;
}

int p2p_pattern_fill(void* buf, size_t bytes, uint64_t seed, int rank, uint64_t iter, p2p_stream_t stream) {
    if (!g.inited) return fail(P2P_INVALID_USAGE, "p2p_init has not been called");
    if (bytes && !buf) return fail(P2P_INVALID_ARGUMENT, "null buffer");
    CUDA_TRY(launch_pattern_fill(buf, bytes, pattern_key(seed, rank, iter), g.sm_count, (cudaStream_t)stream));
    return P2P_SUCCESS;
}

static int run_check(const void* buf, size_t bytes, uint64_t key, int check, cudaStream_t stream,
                     uint64_t* mismatches, uint64_t* first_bad, uint64_t* digest) {
    if (!g.inited) return fail(P2P_INVALID_USAGE, "p2p_init has not been called");
    if (bytes && !buf) return fail(P2P_INVALID_ARGUMENT, "null buffer");
    std::lock_guard<std::mutex> lk(g.mu);  // one scratch area
    CUDA_TRY(launch_pattern_check(buf, bytes, key, check, g.sm_count, g.chk_partials, g.chk_ticket, g.chk_out, stream));
    CUDA_TRY(cudaMemcpyAsync(g.chk_out_host, g.chk_out, sizeof(CheckPartial), cudaMemcpyDeviceToHost, stream));
    CUDA_TRY(cudaStreamSynchronize(stream));
    if (mismatches) *mismatches = g.chk_out_host->mismatches;
    if (first_bad) *first_bad = g.chk_out_host->first_bad;
    if (digest) *digest = g.chk_out_host->digest;
    return P2P_SUCCESS;
}

int p2p_pattern_check(const void* buf, size_t bytes, uint64_t seed, int rank, uint64_t iter, p2p_stream_t stream,
                      uint64_t* mismatches, uint64_t* first_bad, uint64_t* digest) {
    return run_check(buf, bytes, pattern_key(seed, rank, iter), 1, (cudaStream_t)stream, mismatches, first_bad, digest);
}

int p2p_digest(const void* buf, size_t bytes, p2p_stream_t stream, uint64_t* digest) {
    return run_check(buf, bytes, 0, 0, (cudaStream_t)stream, nullptr, nullptr, digest);
}

// ---- plan-only mode, for CPU tests of the host logic (declared in csrc/p2p_debug.h, not in p2p.h) ----

int p2p_debug_plan_init(int rank, int nranks, int sm_count, int ctas_ldst, int ctas_tma) {
    std::lock_guard<std::mutex> lk(g.mu);
    if (g.inited) return fail(P2P_INVALID_USAGE, "engine already initialised");
    if (nranks < 1 || nranks > P2P_MAX_RANKS || rank < 0 || rank >= nranks) return fail(P2P_INVALID_ARGUMENT, "bad rank/nranks");
    g.rank = rank;
    g.nranks = nranks;
    g.device = 0;
    g.sm_count = sm_count;
    g.ctas_ldst = ctas_ldst;
    g.ctas_tma = ctas_tma;
    Options& o = g.opt;
    o = Options();
    {
        int rc = read_env_options(o);
        if (rc) return rc;
    }
    resolve_stage_chunk(o, nranks);
    // the same layout arithmetic as p2p_init, over made-up base addresses: rank r's control region "lives" at
    // P2P_DEBUG_CTL_BASE + r * P2P_DEBUG_CTL_STRIDE, so tests can tell whose memory a flag address points into
    g.ll_slot_bytes = ((size_t)std::max<long long>(o.eager_bytes, 8) * 2 + 255) & ~(size_t)255;
    g.ll_off = (sizeof(PeerCtl) * (size_t)nranks + 4095) & ~(size_t)4095;
    g.stage_off = (g.ll_off + (size_t)nranks * kLLSlots * g.ll_slot_bytes + 4095) & ~(size_t)4095;
    g.ctl_bytes = g.stage_off + (size_t)nranks * stage_area_bytes();
    g.stage_table = (StageInfo*)0x500000000000ull;
    for (int r = 0; r < nranks; ++r) g.peer_ctl[r] = (uint8_t*)(0x100000000000ull + (uint64_t)r * 0x1000000000ull);
    g.ctl = g.peer_ctl[rank];
    g.arrive = (uint32_t*)0x200000000000ull;
    g.win_table = (uint64_t*)0x300000000000ull;
    g.tokens = (uint32_t*)0x400000000000ull;
    g.launch_counter = 0;
    g.stream_last.clear();
    g.hostbox = nullptr;
    g.async_err_dev = nullptr;
    g.async_err_host = nullptr;
    g.plan_only = g.inited = g.connected = true;
    return P2P_SUCCESS;
}

// Back to the state between p2p_init and p2p_connect, so that p2p_connect's checks of the peers' blobs
// (same host, same protocol limits, rank order) can be exercised without a device.
int p2p_debug_plan_disconnect(uint64_t* host_hash, size_t* ctl_bytes) {
    std::lock_guard<std::mutex> lk(g.mu);
    if (!g.plan_only) return fail(P2P_INVALID_USAGE, "not in plan-only mode");
    g.connected = false;
    g.host_hash = host_hash_now();
    if (host_hash) *host_hash = g.host_hash;
    if (ctl_bytes) *ctl_bytes = g.ctl_bytes;
    return P2P_SUCCESS;
}

// Pretend `base .. base+bytes` is a registered window of this rank (every rank contributing).
int p2p_debug_plan_window(uint64_t base, size_t bytes) {
    std::lock_guard<std::mutex> lk(g.mu);
    if (!g.plan_only) return fail(P2P_INVALID_USAGE, "not in plan-only mode");
    for (int w = 0; w < P2P_MAX_WINDOWS; ++w)
        if (!g.windows[w].used) {
            Window W;
            W.used = true;
            W.base = (uint8_t*)base;
            W.bytes = bytes;
            for (int r = 0; r < g.nranks; ++r) W.peer_base[r] = (uint8_t*)(base + ((uint64_t)(r + 1) << 44));
            g.windows[w] = W;
            return w;
        }
    return -1;
}

// The last group as it would have been launched: raw GroupArgs (layout in p2p_device.cuh) + launch shape.
int p2p_debug_plan_last(void* args_out, size_t args_bytes, uint32_t* grid, uint32_t* block, int* flavour) {
    std::lock_guard<std::mutex> lk(g.mu);
    if (!g.plan_only) return fail(P2P_INVALID_USAGE, "not in plan-only mode");
    if (args_out) memcpy(args_out, &g.last_args, std::min(args_bytes, sizeof g.last_args));
    if (grid) *grid = g.last_grid;
    if (block) *block = g.last_block;
    if (flavour) *flavour = g.last_flavour;
    return (int)sizeof g.last_args;
}

}  // extern "C"
