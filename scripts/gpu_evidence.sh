#!/bin/bash
# ncu evidence for the kernels north_star names: batched tcgen05 GEMM (tensor-pipe utilisation), codec convolutions, plus the
# config-3 codec-only bench line.  usage: gpurun --timeout 1800 -- 'bash scripts/gpu_evidence.sh <tag>'
OUT=gpurun_out/${1:-evid}; mkdir -p $OUT
cd qwen3-tts-c_b200/csrc && make -j8 > /dev/null 2>&1; cd ../..
timeout 900 python bench.py --codec-frames 3600 --steps 3 --warmup 3 > $OUT/codec3600.json 2> $OUT/codec3600.err; echo "codec bench exit $?"; cat $OUT/codec3600.json; tail -3 $OUT/codec3600.err
# batched path: launch list of 3 frames at B = 64, then a full capture of the GEMM
PB="python bench.py --batch 64 --frames 3 --steps 1 --warmup 1 --no-cpu"
$PB > $OUT/batch_plain.log 2>&1 &&
timeout 900 ncu --metrics gpu__time_duration.sum --clock-control none -s 3000 -c 3000 --csv --log-file $OUT/launches_batch.csv $PB > $OUT/ncu_list_batch.log 2>&1
echo "batch list exit $?"
timeout 900 ncu --set full --clock-control none --import-source on -k regex:k_gemm_tc -s 2000 -c 8 -f -o $OUT/prof_gemm $PB > $OUT/ncu_full_batch.log 2>&1
echo "batch full exit $?"
# codec: launch list of one 360-frame decode, then a full capture of the conv kernel
PC="python bench.py --codec-frames 360 --steps 1 --warmup 1 --no-cpu"
$PC > $OUT/codec_plain.log 2>&1 &&
timeout 900 ncu --metrics gpu__time_duration.sum --clock-control none -c 1000 --csv --log-file $OUT/launches_codec.csv $PC > $OUT/ncu_list_codec.log 2>&1
echo "codec list exit $?"
timeout 900 ncu --set full --clock-control none --import-source on -k regex:k_conv_tile -s 40 -c 6 -f -o $OUT/prof_conv $PC > $OUT/ncu_full_codec.log 2>&1
echo "codec full exit $?"
ls -la $OUT
