#!/bin/bash
# ncu evidence for the kernel bench.py times at N=1 (one GPU only; never wrap a multi-rank command).
O=gpurun_out/prof
mkdir -p $O
CMD="python bench.py --steps 20 --warmup 3 --no-extras"
$CMD > $O/plain_tma.log 2>&1 && \
ncu --metrics gpu__time_duration.sum --clock-control none -c 400 --csv --log-file $O/launches_tma.csv $CMD > $O/ncu_launches_tma.log 2>&1
$CMD > $O/plain_tma2.log 2>&1 && \
ncu --set full --clock-control none --import-source on -k regex:p2p_group_kernel -s 6 -c 2 -o $O/prof_local_tma $CMD > $O/ncu_full_tma.log 2>&1
export P2P_SELF_TMA_MIN_BYTES=-1
$CMD > $O/plain_ldst.log 2>&1 && \
ncu --set full --clock-control none --import-source on -k regex:p2p_group_kernel -s 6 -c 2 -o $O/prof_local_ldst $CMD > $O/ncu_full_ldst.log 2>&1
unset P2P_SELF_TMA_MIN_BYTES
ls -la $O
tail -3 $O/ncu_full_tma.log $O/ncu_full_ldst.log
