#!/bin/bash
# Second-pass scans on 2 GPUs: eager limit, pull CTA count / threshold, PDL, per-path sweeps.
OUT=gpurun_out/n2b
mkdir -p $OUT
MPI=build/mpirun
REF=oracle/_ref
run() { local name=$1 t=$2; shift 2; timeout $t "$@" > $OUT/$name.log 2>&1; echo "exit $?" >> $OUT/$name.log; }

run pytest_multi 900 python -m pytest tests/test_gpu_multi.py tests/test_report.py -m gpu -x -q -p timeout --timeout 600

# eager limit: sweep 8 B .. 4 MiB with different limits
for eb in 16384 65536 262144 1048576; do
  run sweep_eager_$eb 300 env P2P_EAGER_BYTES=$eb $MPI -n 2 ./p2p_matrix --mode sweep --sweep-max 4M --iters 200 --json $OUT/sweep_eager_$eb.json
done
run sweep_nopdl 300 env P2P_PDL=0 $MPI -n 2 ./p2p_matrix --mode sweep --sweep-max 4M --iters 200 --json $OUT/sweep_nopdl.json
run sweep_nccl_small 300 $MPI -n 2 $REF/p2p_matrix_nccl --mode sweep --sweep-max 4M --iters 200 --json $OUT/sweep_nccl_small.json

# full sweeps: auto (pull for large), forced push
run sweep_auto 600 $MPI -n 2 ./p2p_matrix --mode sweep --iters 200 --verify --json $OUT/sweep_auto.json
run sweep_push 600 env P2P_PATH=ldst $MPI -n 2 ./p2p_matrix --mode sweep --sweep-min 1M --iters 200 --json $OUT/sweep_push.json

# pull CTA scan
: > $OUT/scan.jsonl
for size in 8388608 67108864 268435456; do
  for nb in 16 32 48 64 96 128; do
    P2P_PULL_NBLOCKS=$nb timeout 120 $MPI -n 2 ./p2p_matrix --mode uni --first-pair --timing dev --size $size --iters 50 --warmup 5 --json $OUT/tmp.json > /dev/null 2>&1
    sed "s/^{/{\"path\": \"pull\", \"nblocks\": $nb, /" $OUT/tmp.json >> $OUT/scan.jsonl
  done
done
rm -f $OUT/tmp.json

# BASELINE configs at N=2 again with the new defaults, engine only (NCCL numbers are in n2/)
run cfg2_uni256_dev 600 $MPI -n 2 ./p2p_matrix --mode uni --size 268435456 --iters 30 --timing dev --extra --verify
run cfg3_bidir256_dev 600 $MPI -n 2 ./p2p_matrix --mode bidir --size 268435456 --iters 30 --timing dev --extra --verify
run cfg3_bidir256_dev_pullboth 600 env P2P_BIDIR_PUSH=0 $MPI -n 2 ./p2p_matrix --mode bidir --size 268435456 --iters 30 --timing dev --extra
run cfg1_uni64_ref 600 $MPI -n 2 ./p2p_matrix --mode uni --first-pair --size 67108864 --iters 128 --timing ref
run cfg1_uni64_dev 600 $MPI -n 2 ./p2p_matrix --mode uni --first-pair --size 67108864 --iters 128 --timing dev
run ref_default 600 $MPI -n 2 ./p2p_matrix --extra --verify
run ref_default_nccl_warm 600 $MPI -n 2 $REF/p2p_matrix_nccl --warmup 5
run bench_engine 900 python -m torch.distributed.run --nnodes=1 --nproc-per-node 2 --master-addr 127.0.0.1 --master-port 29541 bench.py --gpus 2 --steps 200 --warmup 5

tail -n 5 $OUT/pytest_multi.log
for f in sweep_eager_16384 sweep_eager_65536 sweep_eager_262144 sweep_eager_1048576 sweep_nopdl sweep_nccl_small; do echo "== $f"; grep -E "^ +[0-9]+ " $OUT/$f.log | awk '{printf "%s:%s/%s ", $1, $2, $3} END {print ""}'; done
echo "== sweep_auto"; grep -E "^ +[0-9]+ " $OUT/sweep_auto.log | tail -12; tail -2 $OUT/sweep_auto.log
echo "== sweep_push"; grep -E "^ +[0-9]+ " $OUT/sweep_push.log
python - <<'PY'
import json
rows=[json.loads(l) for l in open('gpurun_out/n2b/scan.jsonl')]
for size in (8388608, 67108864, 268435456):
    print("pull", size, " ".join(f"{r['nblocks']}:{r['GBps']:.0f}" for r in rows if r['bytes']==size))
PY
for f in cfg2_uni256_dev cfg3_bidir256_dev cfg3_bidir256_dev_pullboth cfg1_uni64_ref cfg1_uni64_dev ref_default ref_default_nccl_warm; do echo "== $f"; grep -A3 "Gbps" $OUT/$f.log | head -10; grep verify $OUT/$f.log; done
grep -h '"metric"' $OUT/bench_engine.log | cut -c1-1800
