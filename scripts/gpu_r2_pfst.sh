#!/bin/bash
# idle-cluster L2 prefetch of the sub-talker's MLP weights: fraction x pacing (QTC_PF_ST_FRAC / QTC_PF_ST_GAP), same box
TAG=${1:-pfst}; OUT=gpurun_out/$TAG; mkdir -p $OUT
IFS=',' read -ra SW <<< "${PF_SWEEP:-100 0,100 500,75 0,50 0,0 0}"
for v in "${SW[@]}"; do set -- $v
  QTC_PF_ST_FRAC=$1 QTC_PF_ST_GAP=$2 timeout 600 python bench.py --no-batched --steps 2 --warmup 3 --no-cpu > $OUT/bench_$1_$2.json 2> $OUT/bench_$1_$2.err; echo "frac=$1 gap=$2 exit $?"
  python - $OUT/bench_$1_$2.json <<'P'
import json,sys
d=json.loads(open(sys.argv[1]).read().strip().splitlines()[-1]); r=d['roofline']
print(round(d['value'],2), 'frame', round(d['ms_per_frame'],4), 'talker', round(r['ms_per_talker_step'],4), round(r['frac'],4), 'sub', round(r['subtalker']['ms_per_frame'],4))
P
done
