#!/bin/bash
# NOTE: the switch / kernel variant this script measures is not in the library any more: apply profiles/r02_l2_policy_experiment.patch to re-run it
# L2 eviction policy on the cluster kernel's weight loads (QTC_L2_PIN = n: sub-talker layers < n evict_last, everything else - the talker's
# stream included - evict_first; -1: no hint), same box; first the cluster kernel's parity tests on the tree as it stands
TAG=${1:-l2pin}; OUT=gpurun_out/$TAG; mkdir -p $OUT
timeout 900 python -m pytest tests/test_gpu_cluster.py tests/test_gpu_fullshape_parity.py tests/test_gpu_fused.py -x -q > $OUT/pytest.log 2>&1; echo "pytest exit $?"; tail -3 $OUT/pytest.log
for v in -1 0 2 3 4; do
  QTC_L2_PIN=$v timeout 600 python bench.py --no-batched --steps 2 --warmup 3 --no-cpu > $OUT/bench_pin$v.json 2> $OUT/bench_pin$v.err; echo "bench pin=$v exit $?"
  python - $OUT/bench_pin$v.json <<'P'
import json,sys
d=json.loads(open(sys.argv[1]).read().strip().splitlines()[-1]); r=d['roofline']
print(round(d['value'],2), 'frame', round(d['ms_per_frame'],4), 'talker', round(r['ms_per_talker_step'],4), round(r['frac'],4), 'sub', round(r['subtalker']['ms_per_frame'],4))
P
done
