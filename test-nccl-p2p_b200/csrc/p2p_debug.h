/* p2p_debug.h — plan-only mode of the engine, for tests of its host logic on a machine without a GPU.
 * Not part of the public C ABI (include/p2p.h) and of no use to applications: after
 * p2p_debug_plan_init every p2p_send/p2p_recv/p2p_group_end/p2p_alltoall resolves its group exactly
 * as in production (protocol per message, roles, CTA counts, sequence numbers, flag addresses) but
 * records the kernel arguments instead of launching anything. */
#ifndef B200_P2P_DEBUG_H
#define B200_P2P_DEBUG_H
#include "p2p.h"
#ifdef __cplusplus
extern "C" {
#endif
/* made-up address map of plan-only mode */
#define P2P_DEBUG_CTL_BASE 0x100000000000ull   /* rank r's control region starts at BASE + r * STRIDE */
#define P2P_DEBUG_CTL_STRIDE 0x1000000000ull
#define P2P_DEBUG_PEER_WINDOW_SHIFT 44         /* rank r's part of a window is "mapped" at base + ((r+1) << 44) */
P2P_API int p2p_debug_plan_init(int rank, int nranks, int sm_count, int ctas_ldst, int ctas_tma);
P2P_API int p2p_debug_plan_disconnect(uint64_t* host_hash, size_t* ctl_bytes);
P2P_API int p2p_debug_plan_window(uint64_t base, size_t bytes);
P2P_API int p2p_debug_plan_last(void* args_out, size_t args_bytes, uint32_t* grid, uint32_t* block, int* flavour);
#ifdef __cplusplus
}
#endif
#endif
