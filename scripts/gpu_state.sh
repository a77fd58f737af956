#!/bin/bash
# One GPU-box call that re-establishes the state of the tree: all GPU parity tests, the default bench line,
# the reference arm, the B = 64 batched line, phase clocks of the fused kernel.
# usage: gpurun --timeout 1500 -- 'bash scripts/gpu_state.sh <tag>'
set -u
TAG=${1:-state}
OUT=gpurun_out/$TAG
mkdir -p $OUT
nvidia-smi --query-gpu=name,clocks.max.sm,memory.total --format=csv > $OUT/gpu.txt
timeout 1200 python -m pytest tests -m gpu -x -q > $OUT/pytest.log 2>&1; echo "pytest exit $?" | tee -a $OUT/pytest.log
tail -3 $OUT/pytest.log
timeout 600 python bench.py --steps 3 --warmup 3 > $OUT/bench.json 2> $OUT/bench.err; echo "bench exit $?"; cat $OUT/bench.json
timeout 600 python bench.py --impl reference --steps 1 --warmup 1 > $OUT/bench_ref.json 2> $OUT/bench_ref.err; echo "ref exit $?"; cat $OUT/bench_ref.json
timeout 600 python bench.py --batch 64 --frames 250 --steps 2 --warmup 1 --no-cpu > $OUT/batch64.json 2> $OUT/batch64.err; echo "batch exit $?"; cat $OUT/batch64.json
QTC_MODE=1 timeout 300 python scripts/phase_profile.py 1.7b 190 5 > $OUT/phase_talker.txt 2>&1; cat $OUT/phase_talker.txt
QTC_MODE=1 timeout 300 python scripts/phase_profile.py 1.7b -1 2 > $OUT/phase_sub.txt 2>&1; cat $OUT/phase_sub.txt
ls -la $OUT
