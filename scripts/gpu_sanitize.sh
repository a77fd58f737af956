#!/bin/bash
# compute-sanitizer over the fused kernels (shared-memory races, out-of-bounds) on the small presets.
# usage: gpurun --timeout 1500 -- 'bash scripts/gpu_sanitize.sh <tag>'
OUT=gpurun_out/${1:-san}; mkdir -p $OUT
timeout 600 compute-sanitizer --tool memcheck --error-exitcode 9 python __graft_entry__.py --smoke > $OUT/memcheck_smoke.log 2>&1; echo "memcheck smoke exit $?"; tail -4 $OUT/memcheck_smoke.log
timeout 900 compute-sanitizer --tool racecheck --error-exitcode 9 python -m pytest tests/test_gpu_fused.py -x -q -k "long_context and fused" > $OUT/racecheck_fused.log 2>&1; echo "racecheck fused exit $?"; tail -6 $OUT/racecheck_fused.log
timeout 600 compute-sanitizer --tool memcheck --error-exitcode 9 python -m pytest tests/test_gpu_batch.py -x -q -k "generate_batch" > $OUT/memcheck_batch.log 2>&1; echo "memcheck batch exit $?"; tail -4 $OUT/memcheck_batch.log
