#!/bin/bash
# 2-GPU regression: cross-GPU suite, BASELINE pair configs, bench.py.
OUT=gpurun_out/n2f
mkdir -p $OUT
MPI=build/mpirun
run() { local name=$1 t=$2; shift 2; timeout $t "$@" > $OUT/$name.log 2>&1; echo "exit $?" >> $OUT/$name.log; }
run pytest_multi 600 python -m pytest tests/test_gpu_multi.py tests/test_report.py tests/test_zz_no_alloc.py -m gpu -x -q -p timeout --timeout 300
run cfg3_bidir256_dev 200 $MPI -n 2 ./p2p_matrix --mode bidir --size 268435456 --iters 30 --timing dev --verify
run cfg2_uni256_dev 200 $MPI -n 2 ./p2p_matrix --mode uni --size 268435456 --iters 30 --timing dev --verify
run bench_engine 400 python -m torch.distributed.run --nnodes=1 --nproc-per-node 2 --master-addr 127.0.0.1 --master-port 29541 bench.py --gpus 2 --steps 100 --warmup 5
tail -n 4 $OUT/pytest_multi.log
for f in cfg2_uni256_dev cfg3_bidir256_dev; do echo "== $f"; grep -A3 "Gbps" $OUT/$f.log | grep -v "^--" | head -4; grep -h "verify\|Failed" $OUT/$f.log; done
grep -h '"metric"' $OUT/bench_engine.log | cut -c1-1200
