#!/bin/bash
# 2-GPU regression: whole GPU suite plus the all-to-all mode against NCCL.
OUT=gpurun_out/n2g
mkdir -p $OUT
MPI=build/mpirun
run() { local name=$1 t=$2; shift 2; timeout $t "$@" > $OUT/$name.log 2>&1; echo "exit $?" >> $OUT/$name.log; }
run pytest_all 900 python -m pytest tests -m gpu -x -q -p timeout --timeout 300
for sz in 65536 1048576 67108864; do
  run a2a_engine_$sz 120 $MPI -n 2 ./p2p_matrix --mode alltoall --size $sz --iters 50 --warmup 3 --timing dev --verify
  run a2a_nccl_$sz 120 $MPI -n 2 oracle/_ref/p2p_matrix_nccl --mode alltoall --size $sz --iters 50 --warmup 3 --timing dev --verify
done
tail -n 4 $OUT/pytest_all.log
for f in $OUT/a2a_*.log; do echo "== $f"; grep -A1 "All-to-all" $f | tail -1; grep -h "verify:\|Failed" $f; done
