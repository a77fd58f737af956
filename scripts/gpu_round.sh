#!/bin/bash
# One GPU-box call: parity tests, the bench line, then the ncu launch list and one full capture of the fused kernel.
# usage (from the repo root): gpurun --timeout 1500 -- 'bash scripts/gpu_round.sh <tag>'
set -u
TAG=${1:-run}
OUT=gpurun_out/$TAG
mkdir -p $OUT
python -m pytest tests -m gpu -x -q > $OUT/pytest.log 2>&1; echo "pytest exit $?" | tee -a $OUT/pytest.log
tail -3 $OUT/pytest.log
python bench.py --steps 3 --warmup 3 > $OUT/bench.json 2> $OUT/bench.err; echo "bench exit $?"; cat $OUT/bench.json
python bench.py --impl reference --steps 1 --warmup 1 > $OUT/bench_ref.json 2> $OUT/bench_ref.err; echo "ref exit $?"; cat $OUT/bench_ref.json
PROF="python bench.py --frames 8 --steps 1 --warmup 3 --no-cpu"
$PROF > $OUT/plain.log 2>&1 &&
ncu --metrics gpu__time_duration.sum --clock-control none -c 4000 --csv --log-file $OUT/launches.csv $PROF > $OUT/ncu_list.log 2>&1
echo "ncu list exit $?"
$PROF > $OUT/plain2.log 2>&1 &&
ncu --set full --clock-control none --import-source on -k regex:k_lm_fused -s 60 -c 4 -f -o $OUT/prof_fused $PROF > $OUT/ncu_full.log 2>&1
echo "ncu full exit $?"
ls -la $OUT
