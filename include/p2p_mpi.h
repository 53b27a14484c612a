/*
 * p2p_mpi.h — header-only MPI glue for p2p.h: the exchanges the engine's two-phase
 * bootstrap and window registration need, done with MPI_Allgather.  Kept out of the
 * library so that libb200p2p.so links against no particular MPI.
 *
 * It replaces the reference's communicator set-up — rank 0's ncclGetUniqueId, the
 * MPI_Bcast of the 128-byte id and ncclCommInitRank (p2p_matrix.cc:115-120) — with an
 * all-gather of per-rank handles, which is what mapping peer memory needs.
 */
#ifndef B200_P2P_MPI_H
#define B200_P2P_MPI_H

#include <stdio.h>
#include <stdlib.h>
#include "mpi.h"
#include "p2p.h"

/* Same shape as the reference's NCCLCHECK (p2p_matrix.cc:35-42): message on stdout,
 * then exit(EXIT_FAILURE). */
#define P2PCHECK(cmd) do {                                   \
    int r_ = (cmd);                                          \
    if (r_ != P2P_SUCCESS) {                                 \
        printf("Failed, P2P error %s:%d '%s' (%s)\n",        \
               __FILE__, __LINE__, p2p_get_error_string(r_), \
               p2p_get_last_error());                        \
        exit(EXIT_FAILURE);                                  \
    }                                                        \
} while (0)

/* Collective over `comm`.  `device` is the CUDA device this rank drives — the value the
 * reference computes in check_process_placement_policy (p2p_matrix.cc:63-100). */
static inline int p2p_init_mpi(MPI_Comm comm, int device) {
    int rank = 0, size = 0, rc;
    if (MPI_Comm_rank(comm, &rank) != MPI_SUCCESS || MPI_Comm_size(comm, &size) != MPI_SUCCESS)
        return P2P_SYSTEM_ERROR;
    rc = p2p_init(rank, size, device);
    /* every rank must reach the all-gather even if its own init failed */
    size_t nb = p2p_blob_bytes();
    char* mine = (char*)calloc(1, nb);
    char* all = (char*)calloc((size_t)size, nb);
    if (!mine || !all) { free(mine); free(all); return P2P_SYSTEM_ERROR; }
    if (rc == P2P_SUCCESS) rc = p2p_export(mine);
    int ok = (rc == P2P_SUCCESS), all_ok = 0;
    if (MPI_Allreduce(&ok, &all_ok, 1, MPI_INT, MPI_MIN, comm) != MPI_SUCCESS) all_ok = 0;
    if (all_ok) {
        if (MPI_Allgather(mine, (int)nb, MPI_BYTE, all, (int)nb, MPI_BYTE, comm) != MPI_SUCCESS)
            rc = P2P_SYSTEM_ERROR;
        else
            rc = p2p_connect(all);
    } else if (rc == P2P_SUCCESS) {
        rc = P2P_REMOTE_ERROR;
    }
    free(mine);
    free(all);
    return rc;
}

/* Collective: every rank registers one buffer (or NULL/0).  Stands in for
 * ncclCommWindowRegister (nccl.h:251). */
static inline int p2p_register_mpi(MPI_Comm comm, void* buf, size_t bytes, int* window_id) {
    int size = 0, rc;
    if (MPI_Comm_size(comm, &size) != MPI_SUCCESS) return P2P_SYSTEM_ERROR;
    char mine[128];
    char* all = (char*)calloc((size_t)size, sizeof mine);
    if (!all) return P2P_SYSTEM_ERROR;
    rc = p2p_register_export(buf, bytes, mine);
    int ok = (rc == P2P_SUCCESS), all_ok = 0;
    if (MPI_Allreduce(&ok, &all_ok, 1, MPI_INT, MPI_MIN, comm) != MPI_SUCCESS) all_ok = 0;
    if (all_ok) {
        if (MPI_Allgather(mine, (int)sizeof mine, MPI_BYTE, all, (int)sizeof mine, MPI_BYTE, comm) != MPI_SUCCESS)
            rc = P2P_SYSTEM_ERROR;
        else
            rc = p2p_register_import(all, window_id);
    } else if (rc == P2P_SUCCESS) {
        rc = P2P_REMOTE_ERROR;
    }
    free(all);
    return rc;
}

/* Collective: a symmetric window of `bytes` on every rank (ncclMemAlloc + ncclCommWindowRegister in one). */
static inline int p2p_symm_alloc_mpi(MPI_Comm comm, size_t bytes, void** ptr, int* window_id) {
    int size = 0, rc;
    if (MPI_Comm_size(comm, &size) != MPI_SUCCESS) return P2P_SYSTEM_ERROR;
    char mine[256];
    char* all = (char*)calloc((size_t)size, sizeof mine);
    if (!all) return P2P_SYSTEM_ERROR;
    rc = p2p_symm_alloc_export(bytes, mine);
    int ok = (rc == P2P_SUCCESS), all_ok = 0;
    if (MPI_Allreduce(&ok, &all_ok, 1, MPI_INT, MPI_MIN, comm) != MPI_SUCCESS) all_ok = 0;
    if (all_ok) {
        if (MPI_Allgather(mine, (int)sizeof mine, MPI_BYTE, all, (int)sizeof mine, MPI_BYTE, comm) != MPI_SUCCESS)
            rc = P2P_SYSTEM_ERROR;
        else
            rc = p2p_symm_alloc_import(all, ptr, window_id);
    } else if (rc == P2P_SUCCESS) {
        rc = P2P_REMOTE_ERROR;
    }
    free(all);
    return rc;
}

/* Collective teardown: close every mapping of peer memory, wait for all ranks to have done so,
 * then free this rank's own regions (stands in for ncclCommDestroy, p2p_matrix.cc:270). */
static inline int p2p_finalize_mpi(MPI_Comm comm) {
    int rc = p2p_disconnect();
    int brc = MPI_Barrier(comm);  /* reached even if the disconnect failed: peers are waiting in it */
    int frc = p2p_finalize();
    if (rc != P2P_SUCCESS) return rc;
    if (brc != MPI_SUCCESS) return P2P_SYSTEM_ERROR;
    return frc;
}

#endif /* B200_P2P_MPI_H */
