# same box, two builds: QTC_LIB=old build vs the tree's build
for lib in scripts/probe/libs/libA_head.so qwen3-tts-c_b200/csrc/libqwen_tts_cuda.so; do
  echo "== $lib"; QTC_LIB=$PWD/$lib STEPS=4 timeout 200 python scripts/cluster_check.py small 1.7b 2>&1 | grep "mode 2:"
  QTC_LIB=$PWD/$lib timeout 200 python scripts/phase_profile_cluster.py 1.7b 190 3 2>&1 | grep "B: \|layer total\|talker step"
done
