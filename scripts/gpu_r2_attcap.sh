#!/bin/bash
# the talker's shared-memory K/V stage (QTC_ATT_CAP positions per CTA: 2 x 512 bytes each, taken from L1), same box
TAG=${1:-attcap}; OUT=gpurun_out/$TAG; mkdir -p $OUT
for v in 0 32 48 64; do
  QTC_ATT_CAP=$v timeout 600 python bench.py --no-batched --steps 2 --warmup 3 --no-cpu > $OUT/bench_$v.json 2> $OUT/bench_$v.err; echo "att_cap=$v exit $?"
  python - $OUT/bench_$v.json <<'P'
import json,sys
d=json.loads(open(sys.argv[1]).read().strip().splitlines()[-1]); r=d['roofline']
print(round(d['value'],2), 'frame', round(d['ms_per_frame'],4), 'talker', round(r['ms_per_talker_step'],4), round(r['frac'],4), 'sub', round(r['subtalker']['ms_per_frame'],4))
P
done
