#!/bin/bash
# ncu evidence for the cluster kernel: launch list of an 8-frame utterance + one --set full capture (talker and sub-talker launches).
# ncu could not replay the cooperative + cluster launch (LaunchFailed), so the profiled runs drop the cooperative attribute (QTC_CL_COOP=0).
set -u
TAG=${1:-cluster_ncu}
OUT=gpurun_out/$TAG
mkdir -p $OUT
export QTC_CL_COOP=0
PROF="python bench.py --frames 8 --steps 1 --warmup 3 --no-cpu"
$PROF > $OUT/plain.log 2>&1; echo "plain exit $?"; tail -c 600 $OUT/plain.log
timeout 600 ncu --metrics gpu__time_duration.sum --clock-control none -c 4000 --csv --log-file $OUT/launches.csv $PROF > $OUT/ncu_list.log 2>&1
echo "ncu list exit $?"
timeout 900 ncu --set full --clock-control none --import-source on -k regex:k_lm_cluster -s 60 -c 4 -f -o $OUT/prof_cluster $PROF > $OUT/ncu_full.log 2>&1
echo "ncu full exit $?"; tail -5 $OUT/ncu_full.log
ls -la $OUT
