#!/bin/bash
# mode 1 after a kernel change: fused + generate parity tests, phase clocks, short bench.  usage: gpurun --timeout 900 -- 'bash scripts/gpu_m1.sh <tag>'
TAG=${1:-m1}; OUT=gpurun_out/$TAG; mkdir -p $OUT
timeout 600 python -m pytest tests/test_gpu_fused.py tests/test_gpu_parity.py -x -q > $OUT/pytest.log 2>&1; echo "pytest exit $?"; tail -5 $OUT/pytest.log
QTC_MODE=1 timeout 120 python scripts/phase_profile.py 1.7b 190 5 > $OUT/phase_talker.txt 2>&1; cat $OUT/phase_talker.txt
QTC_MODE=1 timeout 120 python scripts/phase_profile.py 1.7b -1 2 > $OUT/phase_sub.txt 2>&1; cat $OUT/phase_sub.txt
timeout 300 python bench.py --steps 3 --warmup 3 --no-cpu > $OUT/bench.json 2> $OUT/bench.err; echo "bench exit $?"; cat $OUT/bench.json; tail -3 $OUT/bench.err
