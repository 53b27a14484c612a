#!/bin/bash
# Short 8-GPU check of the all-pairs defaults (every second here costs eight).
N=8; OUT=gpurun_out/n8b; mkdir -p $OUT; MPI=build/mpirun; REF=oracle/_ref
{
echo "== engine default, 64 MiB, verify"; timeout 120 $MPI -n $N ./p2p_matrix --mode allpairs --size 67108864 --iters 30 --warmup 3 --timing dev --verify --json $OUT/cfg4_engine_dev.json | tail -3
echo "== nccl, 64 MiB"; timeout 120 $MPI -n $N $REF/p2p_matrix_nccl --mode allpairs --size 67108864 --iters 30 --warmup 3 --timing dev --json $OUT/cfg4_nccl_dev.json | tail -1
echo "== engine default, 64 MiB, reference timing"; timeout 120 $MPI -n $N ./p2p_matrix --mode allpairs --size 67108864 --iters 30 --warmup 3 --timing ref | tail -1
for sz in 1048576 8388608; do echo "== engine $sz"; timeout 120 $MPI -n $N ./p2p_matrix --mode allpairs --size $sz --iters 100 --warmup 5 --timing dev --verify | tail -2; done
echo "== engine 1 MiB with eager limit 512 KiB"; P2P_EAGER_BYTES=524288 timeout 120 $MPI -n $N ./p2p_matrix --mode allpairs --size 1048576 --iters 100 --warmup 5 --timing dev | tail -1
echo "== ring 256 MiB pull 32 / 16 CTAs"; for nb in 32 16; do P2P_PULL_NBLOCKS=$nb timeout 120 $MPI -n $N ./p2p_matrix --mode ring --size 268435456 --iters 30 --warmup 3 --timing dev | tail -1; done
} > $OUT/allpairs.log 2>&1
cat $OUT/allpairs.log
