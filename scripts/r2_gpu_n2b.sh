#!/bin/bash
# Round 2, second pass on two B200s: the bulk-copy-ring pull, the staging ring, their thresholds.   scripts/r2_gpu_n2b.sh
OUT=gpurun_out/r2_n2b
mkdir -p $OUT
MPI=build/mpirun
REF=oracle/_ref
run() { local name=$1 t=$2; shift 2; timeout $t "$@" > $OUT/$name.log 2>&1; echo "exit $?" >> $OUT/$name.log; }

run pytest_gpu 900 python -m pytest tests -m gpu -x -q -p timeout --timeout 300

sweep() { local name=$1; shift; env "$@" timeout 200 $MPI -n 2 ./p2p_matrix --mode sweep --sweep-min 1M --sweep-max 256M --iters 100 --verify --json $OUT/sweep_$name.json > $OUT/sweep_$name.log 2>&1; echo "exit $?" >> $OUT/sweep_$name.log; }
sweep default P2P_PDL=1
sweep ldst_pull P2P_PULL_TMA_MIN_BYTES=-1
sweep tma4m P2P_PULL_TMA_MIN_BYTES=4194304
sweep pdl3_early P2P_PDL=3 P2P_EARLY_POST=1
sweep pdl3_early_tma4m P2P_PDL=3 P2P_EARLY_POST=1 P2P_PULL_TMA_MIN_BYTES=4194304
sweep pdl3_early_ctas32 P2P_PDL=3 P2P_EARLY_POST=1 P2P_PULL_NBLOCKS=32
sweep pdl3_early_ctas96 P2P_PDL=3 P2P_EARLY_POST=1 P2P_PULL_NBLOCKS=96
sweep pdl3_early_ctas147 P2P_PDL=3 P2P_EARLY_POST=1 P2P_PULL_NBLOCKS=147
sweep pdl3_early_32kx6 P2P_PDL=3 P2P_EARLY_POST=1 P2P_TMA_PEER_CHUNK=32768 P2P_TMA_PEER_STAGES=6
sweep nowindows P2P_NO_WINDOWS=1
sweep nowindows_chunk512k P2P_NO_WINDOWS=1 P2P_STAGE_CHUNK=524288
sweep nowindows_chunk2m P2P_NO_WINDOWS=1 P2P_STAGE_CHUNK=2097152
sweep norecvwindow P2P_NO_RECV_WINDOW=1

: > $OUT/bidir.log
bid() { echo "== $*" >> $OUT/bidir.log; for sz in 16777216 67108864 268435456; do env "$@" timeout 100 $MPI -n 2 ./p2p_matrix --mode bidir --first-pair --size $sz --iters 30 --warmup 5 --timing dev --verify --json $OUT/tmp.json > $OUT/tmp.log 2>&1; grep -h '"mode": "bidir"' $OUT/tmp.json | sed 's/.*"bytes": \([0-9]*\).*"us": \([0-9.]*\).*"GBps": \([0-9.]*\).*/  \1 bytes  \2 us  \3 GB\/s/' >> $OUT/bidir.log; grep -h "verify:" $OUT/tmp.log >> $OUT/bidir.log; done; }
bid P2P_BIDIR_PUSH=1
bid P2P_BIDIR_PUSH=0
bid P2P_BIDIR_PUSH=0 P2P_PULL_NBLOCKS=96
bid P2P_BIDIR_PUSH=0 P2P_PULL_NBLOCKS=147
bid P2P_BIDIR_PUSH=0 P2P_PULL_TMA_MIN_BYTES=-1
bid P2P_BIDIR_PUSH=1 P2P_PATH=tma
bid P2P_BIDIR_PUSH=0 P2P_PDL=3 P2P_EARLY_POST=1
bid P2P_BIDIR_PUSH=1 P2P_PDL=3 P2P_EARLY_POST=1
bid P2P_NO_WINDOWS=1

# ---- ncu: the bulk-copy pull and the eager path with peer pointers
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
prof tmapull 67108864 64 4 2
prof tma 67108864 32 4 2
prof ll 8 1 0 4

run bench_engine_n2 900 python -m torch.distributed.run --nnodes=1 --nproc-per-node 2 --master-addr 127.0.0.1 --master-port 29541 bench.py --gpus 2 --steps 100 --warmup 5
run bench_engine_n1 600 python bench.py --steps 100 --warmup 5
run smoke 300 python -c "import __graft_entry__ as g; g.smoke()"

tail -n 3 $OUT/pytest_gpu.log
for f in $OUT/sweep_*.log; do echo "== $f"; grep -E "^ +[0-9]+ " $f | awk '{print $1, $2, $3, $4}' | tr '\n' ';'; echo; grep -h "verify\|Failed\|exit [1-9]" $f | tail -2; done
cat $OUT/bidir.log
tail -n 3 $P/plain_*.log
tail -n 2 $OUT/smoke.log
grep -h '"metric"' $OUT/bench_*.log | cut -c1-2500
