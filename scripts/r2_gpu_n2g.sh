#!/bin/bash
OUT=gpurun_out/r2_n2g
mkdir -p $OUT
MPI=build/mpirun
export P2P_PRINT_STATS=1 P2P_CE_MIN_BYTES=8388608
t() { echo "== $*" >> $OUT/ce.log; env "$@" >> $OUT/ce.log 2>&1; }
: > $OUT/ce.log
t P2P_CE_WAIT_US=2000 timeout 60 $MPI -n 2 ./p2p_matrix --mode ring --size 268435456 --iters 50 --warmup 5 --timing dev
t P2P_CE_WAIT_US=2000 timeout 60 $MPI -n 2 ./p2p_matrix --mode ring --size 268435456 --iters 50 --warmup 5 --timing ref
t P2P_CE_WAIT_US=200 timeout 60 $MPI -n 2 ./p2p_matrix --mode ring --size 268435456 --iters 50 --warmup 5 --timing dev
t P2P_CE_WAIT_US=20000 timeout 60 $MPI -n 2 ./p2p_matrix --mode ring --size 268435456 --iters 10 --warmup 5 --timing dev
t P2P_CE_WAIT_US=2000 timeout 60 $MPI -n 2 ./p2p_matrix --mode ring --size 134217728 --iters 50 --warmup 5 --timing dev
t P2P_CE_WAIT_US=2000 timeout 60 $MPI -n 2 ./p2p_matrix --mode ring --size 67108864 --iters 50 --warmup 5 --timing dev
t P2P_CE_WAIT_US=2000 timeout 60 $MPI -n 2 ./p2p_matrix --mode uni --first-pair --size 268435456 --iters 50 --warmup 5 --timing dev
t P2P_CE_WAIT_US=2000 timeout 60 $MPI -n 2 ./p2p_matrix --mode uni --first-pair --size 268435456 --iters 50 --warmup 5 --timing ref
t P2P_CE_WAIT_US=2000 P2P_PULL_NBLOCKS=1 timeout 60 $MPI -n 2 ./p2p_matrix --mode ring --size 268435456 --iters 50 --warmup 5 --timing dev
grep -v "^ *[01] \|D.D\|^$" $OUT/ce.log
