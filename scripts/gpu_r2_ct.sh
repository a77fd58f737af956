#!/bin/bash
TAG=${1:-ct}; OUT=gpurun_out/$TAG; mkdir -p $OUT
rm -f gpurun_out/snr_*.log; timeout 900 python -m pytest tests/test_gpu_conv_tc.py tests/test_gpu_fullshape_parity.py tests/test_gpu_fullsize.py tests/test_gpu_parity.py -q -x --durations=5 > $OUT/pytest.log 2>&1; echo "pytest exit $?"; tail -30 $OUT/pytest.log
cat gpurun_out/snr_conv_tc.log gpurun_out/snr_fullshape.log
