#!/bin/bash
# 8-GPU measurement pass (also works for N=4):  scripts/gpu_multi8.sh N
# Engine, NCCL harness and the unmodified reference back to back in the same call.
N=${1:-8}
OUT=gpurun_out/n$N
mkdir -p $OUT
MPI=build/mpirun
REF=oracle/_ref
nvidia-smi -L > $OUT/gpus.txt 2>&1
nvidia-smi topo -m >> $OUT/gpus.txt 2>&1
echo "host cores: $(nproc)" >> $OUT/gpus.txt
grep -m1 "model name" /proc/cpuinfo >> $OUT/gpus.txt

run() { local name=$1 t=$2; shift 2; timeout $t "$@" > $OUT/$name.log 2>&1; echo "exit $?" >> $OUT/$name.log; }

run pytest_multi 1200 python -m pytest tests/test_gpu_multi.py tests/test_report.py -m gpu -x -q -p timeout --timeout 900

# the reference's own measurement (32 MiB, 128 iterations, host clock, sync per message)
run ref_unmodified 600 $MPI -n $N $REF/p2p_matrix_ref
run ref_nccl_harness 600 $MPI -n $N $REF/p2p_matrix_nccl --json $OUT/ref_nccl_harness.json
run ref_nccl_harness_warm 600 $MPI -n $N $REF/p2p_matrix_nccl --warmup 5 --json $OUT/ref_nccl_harness_warm.json
run ref_engine 600 $MPI -n $N ./p2p_matrix --extra --verify --json $OUT/ref_engine.json

for be in engine nccl; do
  if [ $be = engine ]; then BIN=./p2p_matrix; V=--verify; else BIN=$REF/p2p_matrix_nccl; V=; fi
  run cfg2_uni256_${be}_dev 900 $MPI -n $N $BIN --mode uni --size 268435456 --iters 20 --warmup 5 --timing dev --extra $V --json $OUT/cfg2_${be}_dev.json
  run cfg3_bidir256_${be}_dev 900 $MPI -n $N $BIN --mode bidir --size 268435456 --iters 20 --warmup 5 --timing dev --extra $V --json $OUT/cfg3_${be}_dev.json
  run cfg4_allpairs64_${be}_dev 600 $MPI -n $N $BIN --mode allpairs --size 67108864 --iters 30 --warmup 3 --timing dev $V --json $OUT/cfg4_${be}_dev.json
  run cfg4_allpairs64_${be}_ref 600 $MPI -n $N $BIN --mode allpairs --size 67108864 --iters 30 --warmup 3 --timing ref --json $OUT/cfg4_${be}_ref.json
  run ring256_${be}_dev 600 $MPI -n $N $BIN --mode ring --size 268435456 --iters 50 --warmup 5 --timing dev $V --json $OUT/ring_${be}_dev.json
  run sweep_${be} 600 $MPI -n $N $BIN --mode sweep --iters 200 --json $OUT/sweep_${be}.json
done

# all-pairs: how many CTAs per flow, push vs pull
: > $OUT/allpairs_scan.log
for nb in 4 8 16 32; do
  echo "push nblocks=$nb" >> $OUT/allpairs_scan.log
  P2P_NBLOCKS=$nb P2P_PULL_NBLOCKS=8 timeout 200 $MPI -n $N ./p2p_matrix --mode allpairs --size 67108864 --iters 30 --warmup 3 --timing dev | tail -1 >> $OUT/allpairs_scan.log
done
for nb in 8 16 32 64; do
  echo "pull nblocks=$nb" >> $OUT/allpairs_scan.log
  P2P_BIDIR_PUSH=0 P2P_PULL_NBLOCKS=$nb timeout 200 $MPI -n $N ./p2p_matrix --mode allpairs --size 67108864 --iters 30 --warmup 3 --timing dev | tail -1 >> $OUT/allpairs_scan.log
done
for sz in 4096 65536 1048576 8388608; do
  echo "allpairs size=$sz engine / nccl" >> $OUT/allpairs_scan.log
  timeout 200 $MPI -n $N ./p2p_matrix --mode allpairs --size $sz --iters 100 --warmup 5 --timing dev | tail -1 >> $OUT/allpairs_scan.log
  timeout 200 $MPI -n $N $REF/p2p_matrix_nccl --mode allpairs --size $sz --iters 100 --warmup 5 --timing dev | tail -1 >> $OUT/allpairs_scan.log
done

# bench.py both arms, at N and at N/2
for n in $N $((N/2)); do
  run bench_engine_n$n 900 python -m torch.distributed.run --nnodes=1 --nproc-per-node $n --master-addr 127.0.0.1 --master-port 29541 bench.py --gpus $n --steps 200 --warmup 5
  run bench_reference_n$n 900 python -m torch.distributed.run --nnodes=1 --nproc-per-node $n --master-addr 127.0.0.1 --master-port 29542 bench.py --gpus $n --steps 200 --warmup 5 --impl reference
done

tail -n 4 $OUT/pytest_multi.log
for f in ref_unmodified ref_nccl_harness_warm ref_engine; do echo "== $f"; grep -B1 -A$((N+1)) "Directional NCCL P2P Bandwidth (Gbps)" $OUT/$f.log | head -$((2*N+8)); grep -h verify $OUT/$f.log; done
for be in engine nccl; do for f in cfg4_allpairs64_${be}_dev cfg4_allpairs64_${be}_ref ring256_${be}_dev; do echo "== $f"; tail -4 $OUT/$f.log; done; done
cat $OUT/allpairs_scan.log
grep -h '"metric"' $OUT/bench_*.log | cut -c1-1200
