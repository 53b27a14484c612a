#!/bin/bash
# Round 2, fourth two-GPU pass: pull prefetch ahead of griddepcontrol.wait, 32-lane staging ring, copy-engine
# mapping experiment, ncu at the bench's message size.
OUT=gpurun_out/r2_n2d
mkdir -p $OUT
MPI=build/mpirun
REF=oracle/_ref
run() { local name=$1 t=$2; shift 2; timeout $t "$@" > $OUT/$name.log 2>&1; echo "exit $?" >> $OUT/$name.log; }
run pytest_gpu 900 python -m pytest tests -m gpu -x -q -p timeout --timeout 300
run ce_windows 300 python -m torch.distributed.run --nnodes=1 --nproc-per-node 2 --master-addr 127.0.0.1 --master-port 29551 scripts/exp_ce_windows.py
sweep() { local name=$1; shift; env "$@" timeout 200 $MPI -n 2 ./p2p_matrix --mode sweep --sweep-min 1M --sweep-max 256M --iters 100 --verify --json $OUT/sweep_$name.json > $OUT/sweep_$name.log 2>&1; echo "exit $?" >> $OUT/sweep_$name.log; }
sweep default P2P_DEFAULTS=1
sweep pdl1 P2P_PDL=1
sweep noearly P2P_EARLY_POST=0
sweep nowindows P2P_NO_WINDOWS=1
sweep nowindows_512k P2P_NO_WINDOWS=1 P2P_STAGE_CHUNK=524288
sweep nowindows_lanes16 P2P_NO_WINDOWS=1 P2P_STAGE_LANES=16
: > $OUT/ring.log
for e in P2P_DEFAULTS=1 P2P_BIDIR_PUSH=0 "P2P_BIDIR_PUSH=0 P2P_PULL_NBLOCKS=96" "P2P_NO_WINDOWS=1"; do echo "== $e" >> $OUT/ring.log; for sz in 67108864 268435456; do env $e timeout 100 $MPI -n 2 ./p2p_matrix --mode ring --size $sz --iters 50 --warmup 5 --timing dev --verify 2>&1 | tail -2 >> $OUT/ring.log; done; done
P=$OUT/ncu
mkdir -p $P
NVL="nvlrx__bytes.sum,nvltx__bytes.sum,dram__bytes_read.sum,dram__bytes_write.sum,lts__t_bytes.sum,gpu__time_duration.sum"
for spec in "push 32" "pull 64" "tmapull 64"; do set -- $spec
  timeout 60 build/prof_peer $1 268435456 $2 10 > $P/plain_$1_256M.log 2>&1 && \
  timeout 300 ncu --metrics $NVL --clock-control none -k regex:p2p_group_kernel -s 4 -c 2 --csv --log-file $P/traffic_$1_256M.csv build/prof_peer $1 268435456 $2 6 > $P/ncu_$1_256M.log 2>&1
done
run bench_engine_n2 900 python -m torch.distributed.run --nnodes=1 --nproc-per-node 2 --master-addr 127.0.0.1 --master-port 29541 bench.py --gpus 2 --steps 100 --warmup 5
tail -n 4 $OUT/pytest_gpu.log
cat $OUT/ce_windows.log | grep -v "^\[" | tail -8
for f in $OUT/sweep_*.log; do echo "== $f"; grep -E "^ +[0-9]+ " $f | awk '{print $1, $2, $3, $4}' | tr '\n' ';'; echo; grep -h "verify\|Failed" $f | tail -1; done
cat $OUT/ring.log
tail -n 2 $P/plain_*.log; grep -h "p2p_group_kernel" $P/traffic_*.csv | cut -c1-400
grep -h '"metric"' $OUT/bench_*.log | cut -c1-3000
