#!/bin/bash
# Evidence for the cluster kernel (mode 2): phase clocks, ncu launch list of an 8-frame utterance, one ncu --set full capture.
# usage: gpurun --timeout 1500 -- 'bash scripts/gpu_r2_cluster.sh <tag>'
set -u
TAG=${1:-cluster}
OUT=gpurun_out/$TAG
mkdir -p $OUT
timeout 200 python scripts/phase_profile_cluster.py 1.7b 190 3 > $OUT/phase_talker.txt 2>&1; tail -19 $OUT/phase_talker.txt
timeout 200 python scripts/phase_profile_cluster.py 1.7b -1 2 > $OUT/phase_sub.txt 2>&1; tail -19 $OUT/phase_sub.txt
PROF="python bench.py --frames 8 --steps 1 --warmup 3 --no-cpu"
$PROF > $OUT/plain.log 2>&1 &&
timeout 600 ncu --metrics gpu__time_duration.sum --clock-control none -c 4000 --csv --log-file $OUT/launches.csv $PROF > $OUT/ncu_list.log 2>&1
echo "ncu list exit $?"
$PROF > $OUT/plain2.log 2>&1 &&
timeout 900 ncu --set full --clock-control none --import-source on -k regex:k_lm_cluster -s 60 -c 4 -f -o $OUT/prof_cluster $PROF > $OUT/ncu_full.log 2>&1
echo "ncu full exit $?"
ls -la $OUT
