#!/bin/bash
# batched path: parity tests, then the B = 64 bench line; default single-utterance bench after it
mkdir -p gpurun_out/$1
timeout 600 python -m pytest tests/test_gpu_batch.py -x -q 2>&1 | tail -3
python bench.py --batch 64 --frames 40 --steps 2 --warmup 1 --no-cpu 2>&1 | tail -1 > gpurun_out/$1/batch64.json
python -c "import json; d=json.load(open('gpurun_out/$1/batch64.json')); print('batch64', round(d['value'],1), round(d['ms_per_frame'],3), d['gpu_launches'])"
python bench.py --steps 2 --warmup 1 --no-cpu 2>&1 | tail -1 > gpurun_out/$1/single.json
python -c "import json; d=json.load(open('gpurun_out/$1/single.json')); print('single', round(d['value'],2), d.get('ms_per_step'))"
