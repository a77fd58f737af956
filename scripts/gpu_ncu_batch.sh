#!/bin/bash
# ncu launch list + one full capture of the tcgen05 GEMM for the batched path
OUT=gpurun_out/${1:-ncu_batch}; mkdir -p $OUT
PROF="python bench.py --batch 64 --frames 3 --steps 1 --warmup 1 --no-cpu"
$PROF > $OUT/plain.log 2>&1 &&
ncu --metrics gpu__time_duration.sum --clock-control none -s 3000 -c 3000 --csv --log-file $OUT/launches.csv $PROF > $OUT/ncu_list.log 2>&1
echo "list exit $?"
$PROF > $OUT/plain2.log 2>&1 &&
ncu --set full --clock-control none --import-source on -k regex:k_gemm_tc -s 2000 -c 8 -f -o $OUT/prof_gemm $PROF > $OUT/ncu_full.log 2>&1
echo "full exit $?"
