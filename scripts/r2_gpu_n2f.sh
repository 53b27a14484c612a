#!/bin/bash
OUT=gpurun_out/r2_n2f
mkdir -p $OUT
for sz in 268435456 67108864; do timeout 200 build/mpirun -n 2 build/exp_ce_chain $sz > $OUT/ce_chain_$sz.log 2>&1; echo "exit $?" >> $OUT/ce_chain_$sz.log; cat $OUT/ce_chain_$sz.log; done
