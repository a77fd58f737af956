#!/bin/bash
# round 2, first call: full-shape parity tests + smoke on the tree as it stands.   usage: bash scripts/gpu_r2_parity.sh <tag>
TAG=${1:-r2a}; OUT=gpurun_out/$TAG; mkdir -p $OUT
nproc > $OUT/nproc.txt
timeout 1500 python -m pytest tests/test_gpu_fullshape_parity.py -x -q --durations=12 > $OUT/pytest_fullshape.log 2>&1; echo "fullshape exit $?"; tail -25 $OUT/pytest_fullshape.log
timeout 300 python -c 'import __graft_entry__ as g; g.smoke()' > $OUT/smoke.log 2>&1; echo "smoke exit $?"; tail -3 $OUT/smoke.log
