#!/bin/bash
# generate_queue: retired requests decoded on the second stream while the rows go on generating (QTC_QUEUE_OVERLAP=0: all decodes at the end)
TAG=${1:-qoverlap}; OUT=gpurun_out/$TAG; mkdir -p $OUT
timeout 200 python -m pytest tests/test_gpu_batch.py -x -q -k "queue" > $OUT/pytest.log 2>&1; echo "pytest exit $?"; tail -2 $OUT/pytest.log
for v in 1 0; do
  QTC_QUEUE_OVERLAP=$v timeout 300 python bench.py --queue 256 --frames 250 --steps 1 --warmup 1 --no-cpu > $OUT/bench_queue_$v.json 2> $OUT/bench_queue_$v.err; echo "overlap=$v exit $?"
  python - $OUT/bench_queue_$v.json <<'P'
import json,sys
d=json.loads(open(sys.argv[1]).read().strip().splitlines()[-1])
print(round(d['value'],1), 'ms', round(d['ms_per_step'],1), 'lockstep', round(d['lockstep_batches']['value'],1))
P
done
