#!/bin/bash
OUT=gpurun_out/r2_n2h
mkdir -p $OUT
MPI=build/mpirun
EXP_REGISTER_SRC=1 timeout 200 $MPI -n 2 build/exp_ce_chain 268435456 2>&1 | grep -E "^\(|^A |^D |^F'" > $OUT/ce_chain_regsrc.log
export P2P_PRINT_STATS=1 P2P_CE_MIN_BYTES=8388608 P2P_CE_WAIT_US=2000
: > $OUT/ce.log
t() { echo "== $*" >> $OUT/ce.log; env "$@" >> $OUT/ce.log 2>&1; }
t P2P_NO_SEND_WINDOW=1 timeout 60 $MPI -n 2 ./p2p_matrix --mode ring --size 268435456 --iters 50 --warmup 5 --timing dev --verify
t P2P_NO_SEND_WINDOW=1 timeout 60 $MPI -n 2 ./p2p_matrix --mode ring --size 67108864 --iters 50 --warmup 5 --timing dev --verify
t P2P_NO_SEND_WINDOW=1 timeout 60 $MPI -n 2 ./p2p_matrix --mode uni --first-pair --size 268435456 --iters 50 --warmup 5 --timing dev --extra
cat $OUT/ce_chain_regsrc.log; grep -v "D.D\|^$\|Evaluating" $OUT/ce.log | cut -c1-160
