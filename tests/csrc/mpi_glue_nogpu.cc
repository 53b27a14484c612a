// The MPI glue (include/p2p_mpi.h) when the engine cannot start: every rank must get an error
// and none may be left waiting in a collective.  Run under mpirun on a box WITHOUT a GPU.
#include <cstdio>
#include "p2p_mpi.h"

int main(int argc, char** argv) {
    int provided = 0, rank = 0;
    MPI_Init_thread(&argc, &argv, MPI_THREAD_MULTIPLE, &provided);
    MPI_Comm_rank(MPI_COMM_WORLD, &rank);
    int rc = p2p_init_mpi(MPI_COMM_WORLD, 0);
    int win = -1;
    int rc2 = p2p_register_mpi(MPI_COMM_WORLD, nullptr, 0, &win);
    MPI_Barrier(MPI_COMM_WORLD);
    printf("rank %d init=%d (%s) register=%d\n", rank, rc, p2p_get_error_string(rc), rc2);
    return (rc != P2P_SUCCESS && rc2 != P2P_SUCCESS) ? 0 : 1;
}
