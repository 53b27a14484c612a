#!/bin/bash
OUT=gpurun_out/n2c
mkdir -p $OUT
MPI=build/mpirun
REF=oracle/_ref
run() { local name=$1 t=$2; shift 2; timeout $t "$@" > $OUT/$name.log 2>&1; echo "exit $?" >> $OUT/$name.log; }
run pytest_all 1200 python -m pytest tests -m gpu -x -q -p timeout --timeout 600
run sweep_auto 600 $MPI -n 2 ./p2p_matrix --mode sweep --iters 200 --verify --json $OUT/sweep_auto.json
run sweep_pdl2 600 env P2P_PDL=2 $MPI -n 2 ./p2p_matrix --mode sweep --sweep-min 1M --iters 200
for nb in 8 16 32; do run sweep_eager_nb$nb 300 env P2P_EAGER_NBLOCKS=$nb $MPI -n 2 ./p2p_matrix --mode sweep --sweep-min 32K --sweep-max 1M --iters 200; done
run cfg3_bidir256_dev 600 $MPI -n 2 ./p2p_matrix --mode bidir --size 268435456 --iters 30 --timing dev --extra --verify
run bench_engine 900 python -m torch.distributed.run --nnodes=1 --nproc-per-node 2 --master-addr 127.0.0.1 --master-port 29541 bench.py --gpus 2 --steps 200 --warmup 5
tail -n 5 $OUT/pytest_all.log
echo "== sweep_auto"; grep -E "^ +[0-9]+ " $OUT/sweep_auto.log; tail -2 $OUT/sweep_auto.log
echo "== sweep_pdl2"; grep -E "^ +[0-9]+ " $OUT/sweep_pdl2.log
for nb in 8 16 32; do echo "== eager nb $nb"; grep -E "^ +[0-9]+ " $OUT/sweep_eager_nb$nb.log | awk '{printf "%s:%s/%s ", $1, $2, $3} END {print ""}'; done
grep -A3 "Gbps" $OUT/cfg3_bidir256_dev.log | head -5
grep -h '"metric"' $OUT/bench_engine.log | cut -c1-1800
