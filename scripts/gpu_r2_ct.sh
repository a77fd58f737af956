#!/bin/bash
TAG=${1:-ct}; OUT=gpurun_out/$TAG; mkdir -p $OUT
timeout 600 python -m pytest tests/test_gpu_conv_tc.py -q --durations=5 > $OUT/pytest.log 2>&1; echo "pytest exit $?"; tail -40 $OUT/pytest.log
