#!/bin/bash
# Multi-GPU measurement pass, run on the GPU box:  scripts/gpu_multi.sh N [quick]
# Writes everything under gpurun_out/n$N/.  Engine, NCCL harness and the unmodified reference
# are run back to back in the same call so that their numbers are comparable.
N=${1:-2}
QUICK=${2:-}
OUT=gpurun_out/n$N
mkdir -p $OUT
MPI=build/mpirun
REF=oracle/_ref
nvidia-smi -L > $OUT/gpus.txt 2>&1
nvidia-smi topo -m >> $OUT/gpus.txt 2>&1
echo "host cores: $(nproc)" >> $OUT/gpus.txt

run() { # name timeout cmd...
    local name=$1 t=$2; shift 2
    timeout $t "$@" > $OUT/$name.log 2>&1
    echo "exit $?" >> $OUT/$name.log
}

# 1. parity
run pytest_multi 900 python -m pytest tests/test_gpu_multi.py tests/test_report.py -m gpu -x -q -p timeout --timeout 600

# 2. the reference's own measurement, three ways: unmodified binary, NCCL harness, engine
run ref_unmodified 600 $MPI -n $N $REF/p2p_matrix_ref
run ref_nccl_harness 600 $MPI -n $N $REF/p2p_matrix_nccl --extra --json $OUT/ref_nccl_harness.json
run ref_engine 600 $MPI -n $N ./p2p_matrix --extra --verify --json $OUT/ref_engine.json

# 3. size sweep on the pair 0<->1 (device timing): engine per path, NCCL
run sweep_engine_ldst 600 env P2P_PATH=ldst $MPI -n 2 ./p2p_matrix --mode sweep --iters 200 --json $OUT/sweep_engine_ldst.json
run sweep_engine_ldst32 600 env P2P_PATH=ldst P2P_VEC_BYTES=32 $MPI -n 2 ./p2p_matrix --mode sweep --iters 200 --json $OUT/sweep_engine_ldst32.json
run sweep_engine_tma 600 env P2P_PATH=tma $MPI -n 2 ./p2p_matrix --mode sweep --iters 200 --json $OUT/sweep_engine_tma.json
run sweep_nccl 600 $MPI -n 2 $REF/p2p_matrix_nccl --mode sweep --iters 200 --json $OUT/sweep_nccl.json

# 4. CTA-count scan at 64 MiB and 256 MiB, uni-directional, device timing
if [ -z "$QUICK" ]; then
  : > $OUT/scan.jsonl
  for size in 67108864 268435456; do
    for nb in 4 8 16 24 32 48 64 96 128; do
      for cfg in "ldst 16" "ldst 32" "tma 16"; do
        set -- $cfg
        P2P_PATH=$1 P2P_VEC_BYTES=$2 P2P_NBLOCKS=$nb timeout 120 $MPI -n 2 ./p2p_matrix --mode uni --first-pair --timing dev \
            --size $size --iters 50 --warmup 5 --json $OUT/scan_tmp.json > /dev/null 2>&1
        sed "s/^{/{\"path\": \"$1\", \"vec\": $2, \"nblocks\": $nb, /" $OUT/scan_tmp.json >> $OUT/scan.jsonl
      done
    done
  done
  # TMA ring geometry at 32 CTAs
  for geo in "8192 16" "16384 8" "32768 4" "32768 6" "65536 3"; do
    set -- $geo
    P2P_PATH=tma P2P_TMA_CHUNK=$1 P2P_TMA_STAGES=$2 P2P_NBLOCKS=32 timeout 120 $MPI -n 2 ./p2p_matrix --mode uni --first-pair --timing dev \
        --size 268435456 --iters 50 --warmup 5 --json $OUT/scan_tmp.json > /dev/null 2>&1
    sed "s/^{/{\"path\": \"tma\", \"chunk\": $1, \"stages\": $2, \"nblocks\": 32, /" $OUT/scan_tmp.json >> $OUT/scan.jsonl
  done
  rm -f $OUT/scan_tmp.json
fi

# 5. configs of BASELINE.json at this N (device timing and reference method), engine vs NCCL
for be in engine nccl; do
  if [ $be = engine ]; then BIN=./p2p_matrix; else BIN=$REF/p2p_matrix_nccl; fi
  run cfg2_uni256_${be}_dev 900 $MPI -n $N $BIN --mode uni --size 268435456 --iters 30 --timing dev --extra --json $OUT/cfg2_${be}_dev.json
  run cfg3_bidir256_${be}_dev 900 $MPI -n $N $BIN --mode bidir --size 268435456 --iters 30 --timing dev --extra --json $OUT/cfg3_${be}_dev.json
  run cfg4_allpairs64_${be}_dev 600 $MPI -n $N $BIN --mode allpairs --size 67108864 --iters 30 --warmup 3 --timing dev --json $OUT/cfg4_${be}_dev.json
  run cfg4_allpairs64_${be}_ref 600 $MPI -n $N $BIN --mode allpairs --size 67108864 --iters 30 --warmup 3 --timing ref --json $OUT/cfg4_${be}_ref.json
  run cfg1_uni64_${be}_ref 600 $MPI -n 2 $BIN --mode uni --first-pair --size 67108864 --iters 128 --timing ref --json $OUT/cfg1_${be}_ref.json
  run cfg1_uni64_${be}_dev 600 $MPI -n 2 $BIN --mode uni --first-pair --size 67108864 --iters 128 --timing dev --json $OUT/cfg1_${be}_dev.json
done
run cfg1_uni64_mpihost 600 $MPI -n 2 $REF/mpi_host_p2p --mode uni --first-pair --size 67108864 --iters 40 --json $OUT/cfg1_mpihost.json

# 6. bench.py both arms
run bench_engine 900 python -m torch.distributed.run --nnodes=1 --nproc-per-node $N --master-addr 127.0.0.1 --master-port 29541 bench.py --gpus $N --steps 200 --warmup 5
run bench_reference 900 python -m torch.distributed.run --nnodes=1 --nproc-per-node $N --master-addr 127.0.0.1 --master-port 29542 bench.py --gpus $N --steps 200 --warmup 5 --impl reference

tail -n 12 $OUT/pytest_multi.log $OUT/ref_unmodified.log $OUT/ref_engine.log
grep -h '"metric"' $OUT/bench_engine.log $OUT/bench_reference.log | cut -c1-1500
