#!/bin/bash
OUT=gpurun_out/r2_n2k
mkdir -p $OUT
MPI=build/mpirun
export P2P_PRINT_STATS=1
: > $OUT/log.txt
t() { echo "== $*" >> $OUT/log.txt; env "$@" 2>&1 | grep -E "step|chains|verify|^ +[0-9]+ +[0-9]" >> $OUT/log.txt; }
for sz in 67108864 268435456; do
t P2P_X=sm_default timeout 60 $MPI -n 2 ./p2p_matrix --mode ring --size $sz --iters 40 --warmup 5 --timing dev --verify
t P2P_DRIVER_ODD_SIZES=1 timeout 60 $MPI -n 2 ./p2p_matrix --mode ring --size $sz --iters 40 --warmup 5 --timing dev --verify
t P2P_BIDIR_PUSH=0 timeout 60 $MPI -n 2 ./p2p_matrix --mode ring --size $sz --iters 40 --warmup 5 --timing dev --verify
t P2P_BIDIR_PUSH=0 P2P_DRIVER_ODD_SIZES=1 timeout 60 $MPI -n 2 ./p2p_matrix --mode ring --size $sz --iters 40 --warmup 5 --timing dev --verify
t P2P_CE_MIN_BYTES=8388608 P2P_CE_WAIT_US=2000 P2P_CE_TIMING=1 timeout 60 $MPI -n 2 ./p2p_matrix --mode ring --size $sz --iters 40 --warmup 5 --timing dev --verify
done
t P2P_CE_MIN_BYTES=8388608 P2P_CE_WAIT_US=2000 timeout 100 $MPI -n 2 ./p2p_matrix --mode sweep --sweep-min 4M --sweep-max 256M --iters 50 --verify
t P2P_X=sm_default timeout 100 $MPI -n 2 ./p2p_matrix --mode sweep --sweep-min 4M --sweep-max 256M --iters 50 --verify
t P2P_DRIVER_ODD_SIZES=1 timeout 100 $MPI -n 2 ./p2p_matrix --mode sweep --sweep-min 4M --sweep-max 256M --iters 50 --verify
cat $OUT/log.txt | cut -c1-200
