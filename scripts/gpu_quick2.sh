#!/bin/bash
# 2-GPU regression: whole GPU suite, verified sweep, bench.py both arms.
OUT=gpurun_out/n2d
mkdir -p $OUT
MPI=build/mpirun
run() { local name=$1 t=$2; shift 2; timeout $t "$@" > $OUT/$name.log 2>&1; echo "exit $?" >> $OUT/$name.log; }
run pytest_all 1200 python -m pytest tests -m gpu -x -q -p timeout --timeout 600
run sweep_auto 600 $MPI -n 2 ./p2p_matrix --mode sweep --iters 200 --verify --json $OUT/sweep_auto.json
run allpairs2 200 $MPI -n 2 ./p2p_matrix --mode allpairs --size 1048576 --iters 50 --timing dev --verify
run bench_engine 900 python -m torch.distributed.run --nnodes=1 --nproc-per-node 2 --master-addr 127.0.0.1 --master-port 29541 bench.py --gpus 2 --steps 200 --warmup 5
tail -n 5 $OUT/pytest_all.log
echo "== sweep_auto"; grep -E "^ +[0-9]+ " $OUT/sweep_auto.log | awk '{printf "%s:%s/%s/%s ", $1, $2, $3, $4} END {print ""}'; tail -2 $OUT/sweep_auto.log
tail -3 $OUT/allpairs2.log
grep -h '"metric"' $OUT/bench_engine.log | cut -c1-1900
