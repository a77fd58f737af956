#!/bin/bash
# A/B of builds / flags on the same box: phase clocks of the talker step and the sub-talker frame.  usage: gpu_ab.sh <tag> "<lib:flags> ..."
TAG=${1:-ab}; OUT=gpurun_out/$TAG; mkdir -p $OUT
for cfg in ${2:-libqwen_tts_cuda_v9.so:0 libqwen_tts_cuda.so:128 libqwen_tts_cuda.so:0}; do
  lib=${cfg%%:*}; fl=${cfg##*:}
  echo "== $lib flags=$fl"
  QTC_LIB=$PWD/qwen3-tts-c_b200/csrc/$lib QTC_FUSED_FLAGS=$fl timeout 300 python scripts/phase_profile.py 1.7b 190 5 2>&1 | grep -v "P3\|consume" | tee $OUT/phase_talker_${lib}_$fl.txt
  QTC_LIB=$PWD/qwen3-tts-c_b200/csrc/$lib QTC_FUSED_FLAGS=$fl timeout 300 python scripts/phase_profile.py 1.7b -1 2 2>&1 | grep -v "P3\|consume" | tee $OUT/phase_sub_${lib}_$fl.txt
done
