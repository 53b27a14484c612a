#!/bin/bash
OUT=gpurun_out/r2_n2i
mkdir -p $OUT
MPI=build/mpirun
export P2P_PRINT_STATS=1 P2P_CE_MIN_BYTES=8388608 P2P_CE_WAIT_US=2000 P2P_CE_TIMING=1
: > $OUT/ce.log
t() { echo "== $*" >> $OUT/ce.log; env "$@" >> $OUT/ce.log 2>&1; }
t timeout 60 $MPI -n 2 ./p2p_matrix --mode ring --size 268435456 --iters 50 --warmup 5 --timing dev
t timeout 60 $MPI -n 2 ./p2p_matrix --mode ring --size 67108864 --iters 50 --warmup 5 --timing dev
t timeout 60 $MPI -n 2 ./p2p_matrix --mode uni --first-pair --size 268435456 --iters 50 --warmup 5 --timing dev
t timeout 60 $MPI -n 2 ./p2p_matrix --mode latency --sweep-min 256M --sweep-max 256M --iters 30
grep -v "D.D\|^$\|Evaluating\|^ *[01] " $OUT/ce.log | cut -c1-200
