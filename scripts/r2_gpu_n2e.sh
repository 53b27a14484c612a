#!/bin/bash
# Round 2, fifth two-GPU pass (short): where the copy-engine chain loses its bandwidth; early PDL trigger for pulls.
OUT=gpurun_out/r2_n2e
mkdir -p $OUT
MPI=build/mpirun
timeout 200 $MPI -n 2 build/exp_ce_chain 268435456 > $OUT/ce_chain_256M.log 2>&1; echo "exit $?" >> $OUT/ce_chain_256M.log
timeout 200 $MPI -n 2 build/exp_ce_chain 67108864 > $OUT/ce_chain_64M.log 2>&1; echo "exit $?" >> $OUT/ce_chain_64M.log
sweep() { local name=$1; shift; env "$@" timeout 200 $MPI -n 2 ./p2p_matrix --mode sweep --sweep-min 1M --sweep-max 256M --iters 100 --verify --json $OUT/sweep_$name.json > $OUT/sweep_$name.log 2>&1; echo "exit $?" >> $OUT/sweep_$name.log; }
sweep default P2P_DEFAULTS=1
sweep pdl2 P2P_PDL=2
sweep pdl2_ctas96 P2P_PDL=2 P2P_PULL_NBLOCKS=96
cat $OUT/ce_chain_256M.log $OUT/ce_chain_64M.log
for f in $OUT/sweep_*.log; do echo "== $f"; grep -E "^ +[0-9]+ " $f | awk '{print $1, $2, $3, $4}' | tr '\n' ';'; echo; grep -h "verify\|Failed" $f | tail -1; done
