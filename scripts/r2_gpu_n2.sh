#!/bin/bash
# Round 2, first pass on two B200s: parity suite, mechanism probe, scans that decide defaults, ncu on the
# real kernels with peer pointers (one process, two GPUs), bench.py both arms.   scripts/r2_gpu_n2.sh
OUT=gpurun_out/r2_n2
mkdir -p $OUT
MPI=build/mpirun
REF=oracle/_ref
nvidia-smi -L > $OUT/gpus.txt 2>&1
echo "host cores: $(nproc)" >> $OUT/gpus.txt
run() { local name=$1 t=$2; shift 2; timeout $t "$@" > $OUT/$name.log 2>&1; echo "exit $?" >> $OUT/$name.log; }

run pytest_gpu 600 python -m pytest tests -m gpu -x -q -p timeout --timeout 300
run probe 240 build/exp_probe

# ---- per-message fixed cost: PDL modes, early post, pull threshold (pair 0 -> 1, ping-pong + streaming)
sweep() { local name=$1; shift; env "$@" timeout 200 $MPI -n 2 ./p2p_matrix --mode sweep --sweep-min 1M --sweep-max 256M --iters 100 --verify --json $OUT/sweep_$name.json > $OUT/sweep_$name.log 2>&1; echo "exit $?" >> $OUT/sweep_$name.log; }
sweep default P2P_PDL=1
sweep pdl2 P2P_PDL=2
sweep pdl3 P2P_PDL=3
sweep pdl3_early P2P_PDL=3 P2P_EARLY_POST=1
sweep pull2m P2P_PULL_MIN_BYTES=2097152
sweep pdl3_early_pull2m P2P_PDL=3 P2P_EARLY_POST=1 P2P_PULL_MIN_BYTES=2097152
sweep ce_memops P2P_CE_MIN_BYTES=8388608 P2P_CE_WAIT_US=2000 P2P_CE_FLAGS=1
sweep ce_dma P2P_CE_MIN_BYTES=8388608 P2P_CE_WAIT_US=2000 P2P_CE_FLAGS=2
sweep ce_kernels P2P_CE_MIN_BYTES=8388608 P2P_CE_WAIT_US=2000 P2P_CE_FLAGS=0
timeout 200 $MPI -n 2 $REF/p2p_matrix_nccl --mode sweep --sweep-min 1M --sweep-max 256M --iters 100 --json $OUT/sweep_nccl.json > $OUT/sweep_nccl.log 2>&1
# small-message end of the sweep: latency with and without PDL everywhere
env P2P_PDL=1 timeout 100 $MPI -n 2 ./p2p_matrix --mode sweep --sweep-min 8 --sweep-max 512K --iters 300 --json $OUT/small_default.json > $OUT/small_default.log 2>&1
env P2P_PDL=3 timeout 100 $MPI -n 2 ./p2p_matrix --mode sweep --sweep-min 8 --sweep-max 512K --iters 300 --json $OUT/small_pdl3.json > $OUT/small_pdl3.log 2>&1

# ---- both directions at once (reference's bidirectional cell) and the ring step at two sizes
: > $OUT/bidir.log
bid() { echo "== $*" >> $OUT/bidir.log; for sz in 67108864 268435456; do env "$@" timeout 100 $MPI -n 2 ./p2p_matrix --mode bidir --first-pair --size $sz --iters 30 --warmup 5 --timing dev --verify --extra 2>&1 | grep -A3 "GB/s (" | tail -2 >> $OUT/bidir.log; done; }
bid P2P_BIDIR_PUSH=1
bid P2P_BIDIR_PUSH=0
bid P2P_BIDIR_PUSH=0 P2P_PULL_NBLOCKS=96
bid P2P_BIDIR_PUSH=0 P2P_PULL_NBLOCKS=128
bid P2P_BIDIR_PUSH=1 P2P_NBLOCKS=48
bid P2P_BIDIR_PUSH=1 P2P_PDL=3
bid P2P_CE_MIN_BYTES=8388608 P2P_CE_WAIT_US=2000 P2P_CE_FLAGS=1
bid P2P_CE_MIN_BYTES=8388608 P2P_CE_WAIT_US=2000 P2P_CE_FLAGS=2
: > $OUT/ring.log
rng() { echo "== $*" >> $OUT/ring.log; for sz in 67108864 268435456; do env "$@" timeout 100 $MPI -n 2 ./p2p_matrix --mode ring --size $sz --iters 50 --warmup 5 --timing dev --verify 2>&1 | tail -2 >> $OUT/ring.log; done; }
rng P2P_BIDIR_PUSH=1
rng P2P_BIDIR_PUSH=1 P2P_PDL=3 P2P_EARLY_POST=1
rng P2P_BIDIR_PUSH=0 P2P_PDL=3 P2P_EARLY_POST=1
rng P2P_CE_MIN_BYTES=8388608 P2P_CE_WAIT_US=2000 P2P_CE_FLAGS=1
rng P2P_CE_MIN_BYTES=8388608 P2P_CE_WAIT_US=2000 P2P_CE_FLAGS=2

# ---- ncu: the real group kernels with peer pointers, one process driving both GPUs
P=$OUT/ncu
mkdir -p $P
NVL="nvlrx__bytes.sum,nvltx__bytes.sum,nvlrx__bytes_data_user.sum,nvltx__bytes_data_user.sum,nvlrx__bytes_data_protocol.sum,nvltx__bytes_data_protocol.sum,nvlrx__bytes_packet_request.sum,nvlrx__bytes_packet_response.sum,nvltx__bytes_packet_request.sum,nvltx__bytes_packet_response.sum,dram__bytes_read.sum,dram__bytes_write.sum,lts__t_bytes.sum,l1tex__t_bytes.sum,gpu__time_duration.sum"
prof() {  # mode bytes ctas skip count
  local m=$1 b=$2 c=$3 skip=$4 cnt=$5
  timeout 60 build/prof_peer $m $b $c 10 > $P/plain_$m.log 2>&1 || return
  timeout 300 ncu --set full --section Nvlink --section Nvlink_Tables --clock-control none --import-source on -k regex:p2p_group_kernel -s $skip -c $cnt \
      -o $P/prof_peer_$m build/prof_peer $m $b $c 6 > $P/ncu_full_$m.log 2>&1
  timeout 300 ncu --metrics $NVL --clock-control none -k regex:p2p_group_kernel -s $skip -c $cnt --csv --log-file $P/nvl_$m.csv \
      build/prof_peer $m $b $c 6 > $P/ncu_nvl_$m.log 2>&1
}
prof push 67108864 32 4 2
prof pull 67108864 64 4 2
prof tma 67108864 32 4 2
prof ll 8 1 0 4
ncu --query-metrics 2>/dev/null | grep -i "^nvl" | head -60 > $P/nvl_metrics_available.txt

# ---- bench.py, both arms
run bench_reference_n2 600 python -m torch.distributed.run --nnodes=1 --nproc-per-node 2 --master-addr 127.0.0.1 --master-port 29542 bench.py --gpus 2 --steps 100 --warmup 5 --impl reference
run bench_engine_n2 900 python -m torch.distributed.run --nnodes=1 --nproc-per-node 2 --master-addr 127.0.0.1 --master-port 29541 bench.py --gpus 2 --steps 100 --warmup 5

tail -n 3 $OUT/pytest_gpu.log
cat $OUT/probe.log
for f in $OUT/sweep_*.log; do echo "== $f"; grep -E "^ +[0-9]+ " $f | awk '{print $1, $2, $3, $4}' | tr '\n' ';'; echo; grep -h "verify" $f | tail -1; done
cat $OUT/bidir.log $OUT/ring.log
tail -n 3 $P/plain_*.log
grep -h '"metric"' $OUT/bench_*.log | cut -c1-3000
