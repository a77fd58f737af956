#!/bin/bash
# quick GPU check of one mode: fused-kernel parity tests, phase clocks, short bench.   usage: bash scripts/gpu_quick.sh <tag> <mode>
TAG=${1:-q}; MODE=${2:-2}
OUT=gpurun_out/$TAG; mkdir -p $OUT
timeout 600 python -m pytest tests/test_gpu_fused.py -x -q > $OUT/pytest.log 2>&1; echo "pytest exit $?"; tail -5 $OUT/pytest.log
QTC_MODE=$MODE timeout 300 python scripts/phase_profile.py 1.7b 190 5 > $OUT/phase_talker.txt 2>&1; cat $OUT/phase_talker.txt
QTC_MODE=$MODE timeout 300 python scripts/phase_profile.py 1.7b -1 2 > $OUT/phase_sub.txt 2>&1; cat $OUT/phase_sub.txt
timeout 600 python bench.py --mode $MODE --steps 2 --warmup 3 --no-cpu > $OUT/bench.json 2> $OUT/bench.err; echo "bench exit $?"; cat $OUT/bench.json; tail -3 $OUT/bench.err
