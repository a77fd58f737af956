#!/bin/bash
# batched path after a change: parity tests, then B = 64 bench with and without programmatic dependent launch.
# usage: gpurun --timeout 1200 -- 'bash scripts/gpu_batch_pdl.sh <tag>'
OUT=gpurun_out/${1:-bp}; mkdir -p $OUT
timeout 600 python -m pytest tests/test_gpu_batch.py -x -q > $OUT/pytest.log 2>&1; echo "pytest exit $?"; tail -4 $OUT/pytest.log
for pdl in 1 0; do
  QTC_PDL=$pdl timeout 600 python bench.py --batch 64 --frames 100 --steps 2 --warmup 1 --no-cpu > $OUT/batch64_pdl$pdl.json 2> $OUT/batch64_pdl$pdl.err; echo "bench pdl=$pdl exit $?"
  python -c "import json; d=json.load(open('$OUT/batch64_pdl$pdl.json')); print('pdl', $pdl, round(d['value'],1), 'audio_s/s', round(d['ms_per_frame'],3), 'ms/frame', d['gpu_launches'])"; tail -2 $OUT/batch64_pdl$pdl.err
done
PB="python bench.py --batch 64 --frames 3 --steps 1 --warmup 1 --no-cpu"
timeout 900 ncu --metrics gpu__time_duration.sum --clock-control none -k regex:'k_b_|k_gemm_tc|k_sample|k_next_embed|k_advance' -s 6000 -c 2500 --csv --log-file $OUT/launches_batch.csv $PB > $OUT/ncu_list_batch.log 2>&1
echo "batch list exit $?"
