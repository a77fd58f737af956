#!/bin/bash
# mode 3 (barrier-free fused step): parity tests under a timeout, phase clocks, short bench next to mode 1.
# usage: gpurun --timeout 900 -- 'bash scripts/gpu_ll.sh <tag>'
TAG=${1:-ll}; OUT=gpurun_out/$TAG; mkdir -p $OUT
timeout 300 python -m pytest tests/test_gpu_fused.py -x -q -k "ll" > $OUT/pytest.log 2>&1; echo "pytest exit $?"; tail -15 $OUT/pytest.log
QTC_MODE=3 timeout 120 python scripts/phase_profile.py 1.7b 190 5 > $OUT/phase_talker.txt 2>&1; cat $OUT/phase_talker.txt
QTC_MODE=3 timeout 120 python scripts/phase_profile.py 1.7b -1 2 > $OUT/phase_sub.txt 2>&1; cat $OUT/phase_sub.txt
timeout 300 python bench.py --mode 3 --steps 2 --warmup 3 --no-cpu > $OUT/bench3.json 2> $OUT/bench3.err; echo "bench exit $?"; cat $OUT/bench3.json; tail -3 $OUT/bench3.err
