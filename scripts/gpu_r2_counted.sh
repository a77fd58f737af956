#!/bin/bash
# NOTE: the switch / kernel variant this script measures is not in the library any more: apply profiles/r02_counted_accumulator_experiment.patch to re-run it
# cluster kernel with counted accumulator words (the data is the barrier: no grid barrier between phases), and the prefill's split sums loaded in parallel
TAG=${1:-counted}; OUT=gpurun_out/$TAG; mkdir -p $OUT
timeout 150 python -m pytest tests/test_gpu_cluster.py -x -q > $OUT/pytest_cluster.log 2>&1; rc=$?; echo "pytest cluster exit $rc"; tail -5 $OUT/pytest_cluster.log
if [ $rc -ne 0 ]; then grep -n "Error\|error\|assert" $OUT/pytest_cluster.log | head -20; exit 1; fi
timeout 900 python -m pytest tests/test_gpu_fused.py tests/test_gpu_fullshape_parity.py tests/test_gpu_parity.py tests/test_gpu_paged_kv.py tests/test_gpu_stream.py -x -q > $OUT/pytest_more.log 2>&1; echo "pytest more exit $?"; tail -5 $OUT/pytest_more.log
timeout 600 python bench.py --no-batched --steps 3 --warmup 3 --no-cpu > $OUT/bench.json 2> $OUT/bench.err; echo "bench exit $?"
python - $OUT/bench.json <<'P'
import json,sys
d=json.loads(open(sys.argv[1]).read().strip().splitlines()[-1]); r=d['roofline']
print(round(d['value'],2), 'frame', round(d['ms_per_frame'],4), 'talker', round(r['ms_per_talker_step'],4), round(r['frac'],4), 'sub', round(r['subtalker']['ms_per_frame'],4), 'prefill', d['latency']['prefill_ms'], 'ttff', d['latency']['ttff_ms'])
P
timeout 300 python scripts/phase_profile_cluster.py 1.7b 190 3 > $OUT/phase_talker.txt 2>&1; head -18 $OUT/phase_talker.txt; grep "layer total\|talker step" $OUT/phase_talker.txt
timeout 300 python scripts/phase_profile_cluster.py 1.7b -1 2 > $OUT/phase_sub.txt 2>&1; head -18 $OUT/phase_sub.txt; grep "layer total\|sub-talker frame" $OUT/phase_sub.txt
