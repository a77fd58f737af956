#!/bin/bash
# A/B of the batched path's gate/up variants
for e in 1 0; do
  QTC_BATCH_SWIGLU_EPI=$e python bench.py --batch 64 --frames 40 --steps 2 --warmup 1 --no-cpu 2>&1 | tail -1 > gpurun_out/batch_epi$e.json
  python -c "import json; d=json.load(open('gpurun_out/batch_epi$e.json')); print('epi', $e, round(d['value'],1), round(d['ms_per_frame'],3), d['gpu_launches'])"
done
QTC_BATCH_SWIGLU_EPI=0 timeout 300 python -m pytest tests/test_gpu_batch.py -x -q 2>&1 | tail -2
