#!/bin/bash
# ncu full capture (with source) of the mode-2 kernel: one sub-talker launch and one talker launch
OUT=gpurun_out/${1:-ncu_ring}; mkdir -p $OUT
PROF="python bench.py --mode 2 --frames 4 --steps 1 --warmup 3 --no-cpu"
$PROF > $OUT/plain.log 2>&1 &&
ncu --set full --clock-control none --import-source on -k regex:k_lm_ring -s 40 -c 2 -f -o $OUT/prof_ring $PROF > $OUT/ncu_full.log 2>&1
echo "ncu exit $?"; tail -3 $OUT/ncu_full.log
