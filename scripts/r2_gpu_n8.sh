#!/bin/bash
# Round 2 on eight B200s (kept short: every minute here costs eight): parity across GPUs, BASELINE cfg 2-5 with the
# round's defaults, NCCL beside the engine where round 1 has no figure, broadcast, bench.py.   scripts/r2_gpu_n8.sh
N=8
OUT=gpurun_out/r2_n8
mkdir -p $OUT
MPI=build/mpirun
REF=oracle/_ref
nvidia-smi -L > $OUT/gpus.txt 2>&1
run() { local name=$1 t=$2; shift 2; timeout $t "$@" > $OUT/$name.log 2>&1; echo "exit $?" >> $OUT/$name.log; }
run pytest_multi 600 python -m pytest tests/test_symm.py tests/test_gpu_multi.py -m gpu -x -q -p timeout --timeout 300 -k "digests or allpairs_and_ring or across_gpus_verified or alltoall"
run cfg2_uni256 300 $MPI -n $N ./p2p_matrix --mode uni --size 268435456 --iters 10 --warmup 3 --timing dev --extra --verify --json $OUT/cfg2_engine.json
run cfg3_bidir256 300 $MPI -n $N ./p2p_matrix --mode bidir --size 268435456 --iters 10 --warmup 3 --timing dev --extra --verify --json $OUT/cfg3_engine.json
: > $OUT/modes.log
m() { echo "== $*" >> $OUT/modes.log; env "$@" 2>&1 | grep -E "round|step|verify:" >> $OUT/modes.log; }
m P2P_X=default timeout 100 $MPI -n $N ./p2p_matrix --mode ring --size 268435456 --iters 40 --warmup 5 --timing dev --verify --json $OUT/ring_engine.json
m P2P_PULL_NBLOCKS=96 timeout 100 $MPI -n $N ./p2p_matrix --mode ring --size 268435456 --iters 40 --warmup 5 --timing dev
m P2P_BIDIR_PUSH=2 timeout 100 $MPI -n $N ./p2p_matrix --mode ring --size 268435456 --iters 40 --warmup 5 --timing dev
m P2P_CE_MIN_BYTES=8388608 P2P_CE_WAIT_US=2000 timeout 100 $MPI -n $N ./p2p_matrix --mode ring --size 268435456 --iters 40 --warmup 5 --timing dev --verify
m P2P_X=nccl timeout 100 $MPI -n $N $REF/p2p_matrix_nccl --mode ring --size 268435456 --iters 40 --warmup 5 --timing dev
m P2P_X=default timeout 100 $MPI -n $N ./p2p_matrix --mode allpairs --size 67108864 --iters 30 --warmup 3 --timing dev --verify --json $OUT/cfg4_engine.json
m P2P_PULL_TMA_MIN_BYTES=-1 timeout 100 $MPI -n $N ./p2p_matrix --mode allpairs --size 67108864 --iters 30 --warmup 3 --timing dev
m P2P_PULL_TOTAL=96 timeout 100 $MPI -n $N ./p2p_matrix --mode allpairs --size 67108864 --iters 30 --warmup 3 --timing dev
m P2P_X=nccl timeout 100 $MPI -n $N $REF/p2p_matrix_nccl --mode allpairs --size 67108864 --iters 30 --warmup 3 --timing dev --json $OUT/cfg4_nccl.json
for sz in 65536 1048576 67108864; do
m P2P_X=default timeout 100 $MPI -n $N ./p2p_matrix --mode alltoall --size $sz --iters 40 --warmup 5 --timing dev --verify
m P2P_X=nccl timeout 100 $MPI -n $N $REF/p2p_matrix_nccl --mode alltoall --size $sz --iters 40 --warmup 5 --timing dev
done
for sz in 1048576 67108864 268435456; do
m P2P_X=multicast timeout 100 $MPI -n $N ./p2p_matrix --mode bcast --size $sz --iters 30 --warmup 5 --timing dev --verify
m P2P_BCAST_SENDS=1 timeout 100 $MPI -n $N ./p2p_matrix --mode bcast --size $sz --iters 30 --warmup 5 --timing dev
m P2P_X=nccl_sends timeout 100 $MPI -n $N $REF/p2p_matrix_nccl --mode bcast --size $sz --iters 30 --warmup 5 --timing dev
m NCCL_HARNESS_BCAST=collective timeout 100 $MPI -n $N $REF/p2p_matrix_nccl --mode bcast --size $sz --iters 30 --warmup 5 --timing dev
done
run sweep_engine 300 $MPI -n $N ./p2p_matrix --mode sweep --iters 100 --verify --json $OUT/sweep_engine.json
run sweep_nccl 300 $MPI -n $N $REF/p2p_matrix_nccl --mode sweep --iters 100 --json $OUT/sweep_nccl.json
run ref_engine 300 $MPI -n $N ./p2p_matrix --verify --json $OUT/ref_engine.json
run bench_engine 900 python -m torch.distributed.run --nnodes=1 --nproc-per-node $N --master-addr 127.0.0.1 --master-port 29541 bench.py --gpus $N --steps 100 --warmup 5
tail -n 3 $OUT/pytest_multi.log
grep -A 9 "Uni-directional, GB/s" $OUT/cfg2_uni256.log | head -11; grep -h "verify" $OUT/cfg2_uni256.log
grep -A 9 "Bi-directional, GB/s" $OUT/cfg3_bidir256.log | head -11; grep -h "verify" $OUT/cfg3_bidir256.log
cat $OUT/modes.log
for f in $OUT/sweep_engine.log $OUT/sweep_nccl.log; do echo "== $f"; grep -E "^ +[0-9]+ " $f | awk '{print $1, $2, $3, $4}' | tr '\n' ';'; echo; grep -h "verify" $f | tail -1; done
grep -A 9 "Uni-Directional" $OUT/ref_engine.log | head -11
grep -h '"metric"' $OUT/bench_*.log | cut -c1-3500
