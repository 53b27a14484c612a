#!/bin/bash
# Round 2 on four B200s: the send+receive policy (ring), the all-pairs CTA budget, broadcast, cfg 5 at N=4.
N=${1:-4}
OUT=gpurun_out/r2_n$N
mkdir -p $OUT
MPI=build/mpirun
REF=oracle/_ref
nvidia-smi -L > $OUT/gpus.txt 2>&1
run() { local name=$1 t=$2; shift 2; timeout $t "$@" > $OUT/$name.log 2>&1; echo "exit $?" >> $OUT/$name.log; }

run pytest_multi 900 python -m pytest tests/test_symm.py tests/test_gpu_multi.py -m gpu -x -q -p timeout --timeout 300

# the lane-based staging ring (receive buffers in no window): parity on one GPU, rate on one pair
run pytest_staging 600 python -m pytest tests/test_gpu_self.py -m gpu -x -q -k "staging or pdl_modes or pull_path" -p timeout --timeout 300
for e in "P2P_NO_WINDOWS=1" "P2P_NO_WINDOWS=1 P2P_STAGE_CHUNK=131072" "P2P_NO_WINDOWS=1 P2P_STAGE_CHUNK=524288" "P2P_NO_WINDOWS=1 P2P_STAGE_LANES=8" "P2P_NO_RECV_WINDOW=1"; do
  name=$(echo "$e" | tr ' =' '__')
  env $e timeout 200 $MPI -n 2 ./p2p_matrix --mode sweep --sweep-min 1M --sweep-max 256M --iters 50 --verify --json $OUT/sweep_$name.json > $OUT/sweep_$name.log 2>&1; echo "exit $?" >> $OUT/sweep_$name.log
done
timeout 200 $MPI -n 2 $REF/p2p_matrix_nccl --mode sweep --sweep-min 1M --sweep-max 256M --iters 50 --json $OUT/sweep_nccl_n2.json > $OUT/sweep_nccl_n2.log 2>&1
# ---- ring step (bench.py's step): who moves the bytes when a GPU sends and receives at once
: > $OUT/ring.log
rng() { echo "== $*" >> $OUT/ring.log; for sz in 67108864 268435456; do env "$@" timeout 100 $MPI -n $N ./p2p_matrix --mode ring --size $sz --iters 40 --warmup 5 --timing dev --verify 2>&1 | tail -2 >> $OUT/ring.log; done; }
rng P2P_DEFAULTS=1
rng P2P_PULL_TMA_MIN_BYTES=-1
rng P2P_BIDIR_PUSH=2
rng P2P_BIDIR_PUSH=2 P2P_NBLOCKS=48
rng P2P_PULL_NBLOCKS=96
rng P2P_PULL_NBLOCKS=147
rng P2P_PULL_NBLOCKS=32
rng P2P_PDL=1
rng P2P_TMA_PEER_CHUNK=32768 P2P_TMA_PEER_STAGES=6
timeout 100 $MPI -n $N $REF/p2p_matrix_nccl --mode ring --size 268435456 --iters 40 --warmup 5 --timing dev 2>&1 | tail -2 >> $OUT/ring.log

# ---- all-pairs, 64 MiB per flow (BASELINE cfg 4 shape)
: > $OUT/allpairs.log
ap() { echo "== $*" >> $OUT/allpairs.log; env "$@" timeout 100 $MPI -n $N ./p2p_matrix --mode allpairs --size 67108864 --iters 30 --warmup 3 --timing dev --verify 2>&1 | tail -2 >> $OUT/allpairs.log; }
ap P2P_DEFAULTS=1
ap P2P_PULL_TMA_MIN_BYTES=-1
ap P2P_PULL_TOTAL=96
ap P2P_PULL_TOTAL=128
ap P2P_PULL_TOTAL=144
ap P2P_PULL_TMA_MIN_BYTES=-1 P2P_PULL_TOTAL=96
ap P2P_PULL_TMA_MIN_BYTES=-1 P2P_PULL_TOTAL=128
ap P2P_BIDIR_PUSH=3
ap P2P_BIDIR_PUSH=3 P2P_PUSH_TOTAL=96
echo "== nccl" >> $OUT/allpairs.log
timeout 100 $MPI -n $N $REF/p2p_matrix_nccl --mode allpairs --size 67108864 --iters 30 --warmup 3 --timing dev 2>&1 | tail -2 >> $OUT/allpairs.log
for sz in 65536 1048576 8388608; do
  echo "== allpairs size=$sz engine / nccl" >> $OUT/allpairs.log
  timeout 100 $MPI -n $N ./p2p_matrix --mode allpairs --size $sz --iters 100 --warmup 5 --timing dev --verify 2>&1 | tail -2 | head -1 >> $OUT/allpairs.log
  timeout 100 $MPI -n $N $REF/p2p_matrix_nccl --mode allpairs --size $sz --iters 100 --warmup 5 --timing dev 2>&1 | tail -1 >> $OUT/allpairs.log
done

# ---- all-to-all
: > $OUT/alltoall.log
for sz in 65536 1048576 67108864; do
  echo "== alltoall block=$sz engine / nccl" >> $OUT/alltoall.log
  timeout 100 $MPI -n $N ./p2p_matrix --mode alltoall --size $sz --iters 50 --warmup 5 --timing dev --verify 2>&1 | tail -2 | head -1 >> $OUT/alltoall.log
  timeout 100 $MPI -n $N $REF/p2p_matrix_nccl --mode alltoall --size $sz --iters 50 --warmup 5 --timing dev 2>&1 | tail -1 >> $OUT/alltoall.log
done

# ---- one-to-many
: > $OUT/bcast.log
for sz in 65536 1048576 16777216 67108864 268435456; do
  echo "== bcast size=$sz: multicast / unicast stores / grouped p2p_send / NCCL grouped sends / ncclBroadcast" >> $OUT/bcast.log
  timeout 100 $MPI -n $N ./p2p_matrix --mode bcast --size $sz --iters 50 --warmup 5 --timing dev --verify 2>&1 | tail -2 | head -1 >> $OUT/bcast.log
  P2P_NO_MULTICAST=1 timeout 100 $MPI -n $N ./p2p_matrix --mode bcast --size $sz --iters 50 --warmup 5 --timing dev --verify 2>&1 | tail -2 | head -1 >> $OUT/bcast.log
  P2P_BCAST_SENDS=1 timeout 100 $MPI -n $N ./p2p_matrix --mode bcast --size $sz --iters 50 --warmup 5 --timing dev --verify 2>&1 | tail -2 | head -1 >> $OUT/bcast.log
  timeout 100 $MPI -n $N $REF/p2p_matrix_nccl --mode bcast --size $sz --iters 50 --warmup 5 --timing dev 2>&1 | tail -1 >> $OUT/bcast.log
  NCCL_HARNESS_BCAST=collective timeout 100 $MPI -n $N $REF/p2p_matrix_nccl --mode bcast --size $sz --iters 50 --warmup 5 --timing dev 2>&1 | tail -1 >> $OUT/bcast.log
done
for nb in 16 64 128; do echo "== bcast 256 MiB multicast, ctas=$nb" >> $OUT/bcast.log; P2P_BCAST_NBLOCKS=$nb timeout 100 $MPI -n $N ./p2p_matrix --mode bcast --size 268435456 --iters 30 --warmup 5 --timing dev 2>&1 | tail -1 >> $OUT/bcast.log; done

# ---- BASELINE cfg 5 at this rank count: 8 B .. 1 GiB on the pair 0 <-> 1
run sweep_engine 300 $MPI -n $N ./p2p_matrix --mode sweep --iters 200 --verify --json $OUT/sweep_engine.json
run sweep_nccl 300 $MPI -n $N $REF/p2p_matrix_nccl --mode sweep --iters 200 --json $OUT/sweep_nccl.json

run bench_engine 900 python -m torch.distributed.run --nnodes=1 --nproc-per-node $N --master-addr 127.0.0.1 --master-port 29541 bench.py --gpus $N --steps 100 --warmup 5

tail -n 4 $OUT/pytest_multi.log
cat $OUT/ring.log $OUT/allpairs.log $OUT/alltoall.log $OUT/bcast.log
tail -n 3 $OUT/pytest_staging.log
for f in $OUT/sweep_*.log; do echo "== $f"; grep -E "^ +[0-9]+ " $f | awk '{print $1, $2, $3, $4}' | tr '\n' ';'; echo; grep -h "verify" $f | tail -1; done
grep -h '"metric"' $OUT/bench_*.log | cut -c1-3000
