#!/bin/bash
# landing-zone kernel: parity tests, phase clocks, bench with the zone on / off.   usage: bash scripts/gpu_r2_lz.sh <tag> [quick]
TAG=${1:-lz}; OUT=gpurun_out/$TAG; mkdir -p $OUT
timeout 900 python -m pytest tests/test_gpu_fused.py tests/test_gpu_fullshape_parity.py -x -q > $OUT/pytest.log 2>&1; echo "pytest exit $?"; tail -6 $OUT/pytest.log
timeout 300 python scripts/phase_profile.py 1.7b 190 5 > $OUT/phase_talker.txt 2>&1; cat $OUT/phase_talker.txt
timeout 300 python scripts/phase_profile.py 1.7b -1 2 > $OUT/phase_sub.txt 2>&1; cat $OUT/phase_sub.txt
QTC_FUSED_FLAGS=128 timeout 300 python scripts/phase_profile.py 1.7b 190 5 > $OUT/phase_talker_nolz.txt 2>&1; cat $OUT/phase_talker_nolz.txt
QTC_FUSED_FLAGS=128 timeout 300 python scripts/phase_profile.py 1.7b -1 2 > $OUT/phase_sub_nolz.txt 2>&1; cat $OUT/phase_sub_nolz.txt
if [ "$2" != "quick" ]; then
timeout 600 python bench.py --steps 2 --warmup 3 --no-cpu > $OUT/bench.json 2> $OUT/bench.err; echo "bench exit $?"; cat $OUT/bench.json; tail -3 $OUT/bench.err
QTC_FUSED_FLAGS=128 timeout 600 python bench.py --steps 2 --warmup 3 --no-cpu > $OUT/bench_nolz.json 2> $OUT/bench_nolz.err; echo "bench(no zone) exit $?"; cat $OUT/bench_nolz.json
fi
