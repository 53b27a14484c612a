/* Exercises every minimpi call the build relies on; run as `mpirun -n N minimpi_selftest`.
 * Prints "OK <n>" on rank 0 and exits 0 only if every rank agrees. */
#include <stdio.h>
#include <stdlib.h>
#include <string.h>
#include "mpi.h"

#define REQUIRE(cond)                                                              \
    do {                                                                           \
        if (!(cond)) {                                                             \
            fprintf(stderr, "rank %d: FAILED %s (%s:%d)\n", rank, #cond, __FILE__, __LINE__); \
            MPI_Abort(MPI_COMM_WORLD, 3);                                          \
        }                                                                          \
    } while (0)

static unsigned char pat(int src, int dst, size_t i) {
    return (unsigned char)(src * 131 + dst * 17 + i * 7 + (i >> 8) * 13);
}

int main(int argc, char** argv) {
    int rank = -1, size = 0, provided = -1;
    if (MPI_Init_thread(&argc, &argv, MPI_THREAD_MULTIPLE, &provided) != MPI_SUCCESS) return 2;
    MPI_Comm_rank(MPI_COMM_WORLD, &rank);
    MPI_Comm_size(MPI_COMM_WORLD, &size);
    REQUIRE(provided == MPI_THREAD_MULTIPLE); /* the reference asserts this, p2p_matrix.cc:106 */
    if (argc > 1 && !strcmp(argv[1], "--fail-rank")) {
        /* failure propagation: one rank dies, the launcher must stop the rest */
        int who = atoi(argv[2]);
        if (rank == who) exit(7);
        MPI_Barrier(MPI_COMM_WORLD);
        MPI_Barrier(MPI_COMM_WORLD);
        return 0;
    }

    /* Allgather, the reference's use: one 64-bit word per rank (p2p_matrix.cc:71-75) */
    uint64_t mine = 0x1000u + (uint64_t)rank, *all = malloc(sizeof(uint64_t) * (size_t)size);
    REQUIRE(MPI_Allgather(&mine, 1, MPI_INT64_T, all, 1, MPI_INT64_T, MPI_COMM_WORLD) == MPI_SUCCESS);
    for (int i = 0; i < size; ++i) REQUIRE(all[i] == 0x1000u + (uint64_t)i);

    /* Allgather bigger than one collective slot */
    size_t big = 200 * 1000 + 3;
    unsigned char* sb = malloc(big);
    unsigned char* rb = malloc(big * (size_t)size);
    for (size_t i = 0; i < big; ++i) sb[i] = pat(rank, 0, i);
    REQUIRE(MPI_Allgather(sb, (int)big, MPI_BYTE, rb, (int)big, MPI_BYTE, MPI_COMM_WORLD) == MPI_SUCCESS);
    for (int r = 0; r < size; ++r)
        for (size_t i = 0; i < big; i += 997) REQUIRE(rb[(size_t)r * big + i] == pat(r, 0, i));

    /* Bcast: 128 bytes from rank 0 (p2p_matrix.cc:118) and a multi-slot one from the last rank */
    unsigned char id[128];
    if (rank == 0) for (int i = 0; i < 128; ++i) id[i] = (unsigned char)(i ^ 0x5a);
    else memset(id, 0, sizeof id);
    REQUIRE(MPI_Bcast(id, sizeof id, MPI_BYTE, 0, MPI_COMM_WORLD) == MPI_SUCCESS);
    for (int i = 0; i < 128; ++i) REQUIRE(id[i] == (unsigned char)(i ^ 0x5a));
    if (rank == size - 1) for (size_t i = 0; i < big; ++i) sb[i] = pat(99, 1, i);
    REQUIRE(MPI_Bcast(sb, (int)big, MPI_BYTE, size - 1, MPI_COMM_WORLD) == MPI_SUCCESS);
    for (size_t i = 0; i < big; i += 499) REQUIRE(sb[i] == pat(99, 1, i));

    /* Allreduce */
    double t = 1.5 * rank, tmax = -1, tsum = -1;
    REQUIRE(MPI_Allreduce(&t, &tmax, 1, MPI_DOUBLE, MPI_MAX, MPI_COMM_WORLD) == MPI_SUCCESS);
    REQUIRE(MPI_Allreduce(&t, &tsum, 1, MPI_DOUBLE, MPI_SUM, MPI_COMM_WORLD) == MPI_SUCCESS);
    REQUIRE(tmax == 1.5 * (size - 1));
    REQUIRE(tsum == 1.5 * (size - 1) * size / 2.0);
    long long cnt = rank + 1, csum = 0;
    REQUIRE(MPI_Allreduce(&cnt, &csum, 1, MPI_LONG_LONG, MPI_SUM, MPI_COMM_WORLD) == MPI_SUCCESS);
    REQUIRE(csum == (long long)size * (size + 1) / 2);

    /* Gather */
    int gi = rank * 3, *gall = malloc(sizeof(int) * (size_t)size);
    REQUIRE(MPI_Gather(&gi, 1, MPI_INT, gall, 1, MPI_INT, 0, MPI_COMM_WORLD) == MPI_SUCCESS);
    if (rank == 0) for (int i = 0; i < size; ++i) REQUIRE(gall[i] == i * 3);

    /* Send/Recv: every ordered pair in turn, message larger than the ring (forces wrap) */
    size_t msg = 9 * 1024 * 1024 + 5;
    unsigned char* m = malloc(msg);
    for (int src = 0; src < size; ++src)
        for (int dst = 0; dst < size; ++dst) {
            if (src == dst) continue;
            if (rank == src) {
                for (size_t i = 0; i < msg; ++i) m[i] = pat(src, dst, i);
                REQUIRE(MPI_Send(m, (int)msg, MPI_BYTE, dst, 5, MPI_COMM_WORLD) == MPI_SUCCESS);
            } else if (rank == dst) {
                MPI_Status st;
                memset(m, 0, msg);
                REQUIRE(MPI_Recv(m, (int)msg, MPI_BYTE, src, 5, MPI_COMM_WORLD, &st) == MPI_SUCCESS);
                int got = 0;
                MPI_Get_count(&st, MPI_BYTE, &got);
                REQUIRE((size_t)got == msg && st.MPI_SOURCE == src && st.MPI_TAG == 5);
                for (size_t i = 0; i < msg; ++i) REQUIRE(m[i] == pat(src, dst, i));
            }
        }

    /* Sendrecv around a ring, both directions busy at once */
    if (size > 1) {
        int right = (rank + 1) % size, left = (rank + size - 1) % size;
        unsigned char* r2 = malloc(msg);
        for (size_t i = 0; i < msg; ++i) m[i] = pat(rank, right, i);
        REQUIRE(MPI_Sendrecv(m, (int)msg, MPI_BYTE, right, 1, r2, (int)msg, MPI_BYTE, left, 1,
                             MPI_COMM_WORLD, MPI_STATUS_IGNORE) == MPI_SUCCESS);
        for (size_t i = 0; i < msg; i += 101) REQUIRE(r2[i] == pat(left, rank, i));
        free(r2);
    }

    /* tags: MPI_ANY_TAG accepts, a wrong tag is reported; a short receive buffer truncates */
    if (size > 1) {
        int v = 41, got = 0;
        MPI_Status st;
        if (rank == 0) {
            REQUIRE(MPI_Send(&v, 1, MPI_INT, 1, 9, MPI_COMM_WORLD) == MPI_SUCCESS);
            REQUIRE(MPI_Send(&v, 1, MPI_INT, 1, 10, MPI_COMM_WORLD) == MPI_SUCCESS);
            int two[2] = {1, 2};
            REQUIRE(MPI_Send(two, 2, MPI_INT, 1, 11, MPI_COMM_WORLD) == MPI_SUCCESS);
            REQUIRE(MPI_Send(&v, 1, MPI_INT, 1, 12, MPI_COMM_WORLD) == MPI_SUCCESS);
        } else if (rank == 1) {
            REQUIRE(MPI_Recv(&got, 1, MPI_INT, 0, MPI_ANY_TAG, MPI_COMM_WORLD, &st) == MPI_SUCCESS);
            REQUIRE(got == 41 && st.MPI_TAG == 9);
            REQUIRE(MPI_Recv(&got, 1, MPI_INT, 0, 77, MPI_COMM_WORLD, &st) == MPI_ERR_TAG);
            REQUIRE(MPI_Recv(&got, 1, MPI_INT, 0, 11, MPI_COMM_WORLD, &st) == MPI_ERR_TRUNCATE);
            REQUIRE(got == 1);
            REQUIRE(MPI_Recv(&got, 1, MPI_INT, 0, 12, MPI_COMM_WORLD, &st) == MPI_SUCCESS);  /* the stream is still in step */
            REQUIRE(got == 41);
        }
        /* argument errors */
        REQUIRE(MPI_Send(&v, 1, MPI_INT, size, 0, MPI_COMM_WORLD) == MPI_ERR_RANK);
        REQUIRE(MPI_Send(&v, 1, 12345, 0, 0, MPI_COMM_WORLD) == MPI_ERR_TYPE);
        REQUIRE(MPI_Send(&v, -1, MPI_INT, 0, 0, MPI_COMM_WORLD) == MPI_ERR_ARG);
        REQUIRE(MPI_Barrier(MPI_COMM_NULL) == MPI_ERR_COMM);
    }

    /* zero-byte message and many barriers */
    if (size > 1) {
        if (rank == 0) REQUIRE(MPI_Send(m, 0, MPI_BYTE, 1, 0, MPI_COMM_WORLD) == MPI_SUCCESS);
        if (rank == 1) REQUIRE(MPI_Recv(m, 0, MPI_BYTE, 0, 0, MPI_COMM_WORLD, MPI_STATUS_IGNORE) == MPI_SUCCESS);
    }
    for (int i = 0; i < 1000; ++i) MPI_Barrier(MPI_COMM_WORLD);

    double w0 = MPI_Wtime(), w1 = MPI_Wtime();
    REQUIRE(w1 >= w0);
    if (rank == 0) printf("OK %d\n", size);
    MPI_Finalize();
    return 0;
}
