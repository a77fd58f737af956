#!/bin/bash
# NOTE: the change this script measured (split sums loaded together) was not kept
# prefill: K-split partial sums loaded together in the consumers (pf_get); parity of both prefill paths, then the latency object
TAG=${1:-pfget}; OUT=gpurun_out/$TAG; mkdir -p $OUT
timeout 600 python -m pytest tests/test_gpu_fused.py tests/test_gpu_parity.py tests/test_gpu_batch.py -x -q > $OUT/pytest.log 2>&1; echo "pytest exit $?"; tail -3 $OUT/pytest.log
timeout 600 python bench.py --no-batched --steps 3 --warmup 3 --no-cpu > $OUT/bench.json 2> $OUT/bench.err; echo "bench exit $?"
python - $OUT/bench.json <<'P'
import json,sys
d=json.loads(open(sys.argv[1]).read().strip().splitlines()[-1]); r=d['roofline']
print(round(d['value'],2), 'frame', round(d['ms_per_frame'],4), 'talker', round(r['ms_per_talker_step'],4), round(r['frac'],4), 'sub', round(r['subtalker']['ms_per_frame'],4), 'prefill', d['latency']['prefill_ms'], 'ttff', d['latency']['ttff_ms'], 'ttfa', d['latency']['ttfa_ms'])
P
