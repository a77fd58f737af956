#!/bin/bash
# A/B of QTC_FUSED_FLAGS values on the fused step: phase clocks + talker-step / sub-talker timing per flag value.
# usage: gpurun --timeout 900 -- 'bash scripts/gpu_flags.sh <tag> "0 4"'
TAG=${1:-fl}; OUT=gpurun_out/$TAG; mkdir -p $OUT
for f in ${2:-0 4}; do
  echo "=== QTC_FUSED_FLAGS=$f"
  QTC_FUSED_FLAGS=$f QTC_MODE=1 timeout 120 python scripts/phase_profile.py 1.7b 190 5 > $OUT/phase_talker_f$f.txt 2>&1; grep -v "qkv consume" $OUT/phase_talker_f$f.txt
  QTC_FUSED_FLAGS=$f QTC_MODE=1 timeout 120 python scripts/phase_profile.py 1.7b -1 2 > $OUT/phase_sub_f$f.txt 2>&1; grep -E "layer total|sub-talker frame" $OUT/phase_sub_f$f.txt
done
