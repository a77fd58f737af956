#!/bin/bash
# batched talker attention with one CTA per (KV head, utterance) serving both query heads (k_b_attn_gqa2): parity, config 4 A/B, launch list of the batched frame loop
TAG=${1:-battn}; OUT=gpurun_out/$TAG; mkdir -p $OUT
timeout 900 python -m pytest tests/test_gpu_batch.py -x -q > $OUT/pytest.log 2>&1; echo "pytest exit $?"; tail -5 $OUT/pytest.log
timeout 900 python -m pytest tests/test_gpu_fullshape_parity.py -x -q -k batch > $OUT/pytest_full.log 2>&1; echo "pytest exit $?"; tail -5 $OUT/pytest_full.log
timeout 600 python bench.py --batch 64 --frames 250 --steps 2 --warmup 1 --no-cpu > $OUT/bench_batch64.json 2> $OUT/bench_batch64.err; echo "batch exit $?"; cat $OUT/bench_batch64.json; tail -3 $OUT/bench_batch64.err
QTC_BATCH_ATTN_GQA=0 timeout 600 python bench.py --batch 64 --frames 250 --steps 2 --warmup 1 --no-cpu > $OUT/bench_batch64_old.json 2> $OUT/bench_batch64_old.err; echo "batch exit $?"; cat $OUT/bench_batch64_old.json; tail -3 $OUT/bench_batch64_old.err
PROF="python bench.py --batch 64 --frames 140 --steps 1 --warmup 0 --no-cpu --codes-only"
timeout 600 ncu --metrics gpu__time_duration.sum --clock-control none -s 85000 -c 2200 --csv --log-file $OUT/launches_batch64.csv $PROF > $OUT/ncu_list.log 2>&1; echo "ncu exit $?"; tail -2 $OUT/ncu_list.log
