#!/bin/bash
# Round 2, quick two-GPU check of the symmetric windows / broadcast and of the defaults after the second pass.
OUT=gpurun_out/r2_n2c
mkdir -p $OUT
MPI=build/mpirun
REF=oracle/_ref
run() { local name=$1 t=$2; shift 2; timeout $t "$@" > $OUT/$name.log 2>&1; echo "exit $?" >> $OUT/$name.log; }
run pytest_symm 600 python -m pytest tests/test_symm.py -m gpu -x -q -p timeout --timeout 200
run pytest_gpu 900 python -m pytest tests -m gpu -x -q -p timeout --timeout 300 --deselect tests/test_symm.py
: > $OUT/bcast.log
for sz in 65536 1048576 67108864 268435456; do
  echo "== bcast size=$sz: multicast / unicast stores / grouped p2p_send / NCCL grouped sends / ncclBroadcast" >> $OUT/bcast.log
  timeout 100 $MPI -n 2 ./p2p_matrix --mode bcast --size $sz --iters 50 --warmup 5 --timing dev --verify 2>&1 | tail -2 | head -1 >> $OUT/bcast.log
  P2P_NO_MULTICAST=1 timeout 100 $MPI -n 2 ./p2p_matrix --mode bcast --size $sz --iters 50 --warmup 5 --timing dev --verify 2>&1 | tail -2 | head -1 >> $OUT/bcast.log
  P2P_BCAST_SENDS=1 timeout 100 $MPI -n 2 ./p2p_matrix --mode bcast --size $sz --iters 50 --warmup 5 --timing dev --verify 2>&1 | tail -2 | head -1 >> $OUT/bcast.log
  timeout 100 $MPI -n 2 $REF/p2p_matrix_nccl --mode bcast --size $sz --iters 50 --warmup 5 --timing dev 2>&1 | tail -1 >> $OUT/bcast.log
  NCCL_HARNESS_BCAST=collective timeout 100 $MPI -n 2 $REF/p2p_matrix_nccl --mode bcast --size $sz --iters 50 --warmup 5 --timing dev 2>&1 | tail -1 >> $OUT/bcast.log
done
P2P_SYMM_HANDLE=fabric timeout 100 $MPI -n 2 ./p2p_matrix --mode bcast --size 1048576 --iters 5 --timing dev --verify > $OUT/fabric.log 2>&1; echo "exit $?" >> $OUT/fabric.log
: > $OUT/ring.log
for e in P2P_DEFAULTS=1 P2P_BIDIR_PUSH=0 "P2P_BIDIR_PUSH=0 P2P_PULL_NBLOCKS=96"; do echo "== $e" >> $OUT/ring.log; for sz in 67108864 268435456; do env $e timeout 100 $MPI -n 2 ./p2p_matrix --mode ring --size $sz --iters 50 --warmup 5 --timing dev --verify 2>&1 | tail -2 >> $OUT/ring.log; done; done
run sweep_default 200 $MPI -n 2 ./p2p_matrix --mode sweep --sweep-min 8 --sweep-max 256M --iters 200 --verify --json $OUT/sweep_default.json
tail -n 5 $OUT/pytest_symm.log $OUT/pytest_gpu.log
cat $OUT/bcast.log $OUT/fabric.log $OUT/ring.log
grep -E "^ +[0-9]+ " $OUT/sweep_default.log | awk '{print $1, $2, $3, $4}' | tr '\n' ';'; echo
