#!/bin/bash
# same box, two builds of the library (QTC_LIB): the committed tree vs the working tree, alternating
TAG=${1:-ablibs}; OUT=gpurun_out/$TAG; mkdir -p $OUT
i=0
for rep in 1 2; do
for v in "scripts/probe/libs/libA_head.so 0" "qwen3-tts-c_b200/csrc/libqwen_tts_cuda.so ${2:-0}"; do set -- $v; i=$((i+1))
  QTC_LIB=$PWD/$1 QTC_FUSED_FLAGS=$2 timeout 600 python bench.py --no-batched --steps 2 --warmup 3 --no-cpu > $OUT/bench_$i.json 2> $OUT/bench_$i.err; echo "bench lib=$1 flags=$2 exit $?"
  python - $OUT/bench_$i.json <<'P'
import json,sys
d=json.loads(open(sys.argv[1]).read().strip().splitlines()[-1]); r=d['roofline']
print(round(d['value'],2), 'frame', round(d['ms_per_frame'],4), 'talker', round(r['ms_per_talker_step'],4), round(r['frac'],4), 'sub', round(r['subtalker']['ms_per_frame'],4), 'p99', d['latency']['frame_ms_p99'])
P
done
done
