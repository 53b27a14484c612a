#!/bin/bash
OUT=gpurun_out/r2_n2j
mkdir -p $OUT
MPI=build/mpirun
EXP_DST_SLACK=1 EXP_REGISTER_SRC=1 timeout 200 $MPI -n 2 build/exp_ce_chain 268435456 2>&1 | grep -E "^\(|^A |^D |^H |^F'" > $OUT/exp.log
export P2P_PRINT_STATS=1 P2P_CE_MIN_BYTES=8388608 P2P_CE_WAIT_US=2000 P2P_CE_TIMING=1
: > $OUT/ce.log
t() { echo "== $*" >> $OUT/ce.log; env "$@" 2>&1 | grep -E "step|chains" >> $OUT/ce.log; }
t P2P_CE_DEBUG_NOZERO=1 timeout 60 $MPI -n 2 ./p2p_matrix --mode ring --size 268435456 --iters 30 --warmup 5 --timing dev
t P2P_PDL=0 timeout 60 $MPI -n 2 ./p2p_matrix --mode ring --size 268435456 --iters 30 --warmup 5 --timing dev
t P2P_EAGER_BYTES=4096 P2P_STAGE_CHUNK=4096 timeout 60 $MPI -n 2 ./p2p_matrix --mode ring --size 268435456 --iters 30 --warmup 5 --timing dev
t P2P_CE_FLAGS=0 timeout 60 $MPI -n 2 ./p2p_matrix --mode ring --size 268435456 --iters 30 --warmup 5 --timing dev
t P2P_CE_FLAGS=2 timeout 60 $MPI -n 2 ./p2p_matrix --mode ring --size 268435456 --iters 30 --warmup 5 --timing dev
cat $OUT/exp.log $OUT/ce.log | cut -c1-200
