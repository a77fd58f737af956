for lib in libA_head lib_stage; do
  echo "== $lib $(QTC_LIB=$PWD/scripts/probe/libs/$lib.so STEPS=40 timeout 200 python scripts/cluster_check.py small 1.7b 2>&1 | grep 'mode 2:\|worst\|codes equal\|identical')"
done
timeout 200 python scripts/phase_profile_cluster.py 1.7b 190 3 2>&1 | tail -19; timeout 200 python scripts/phase_profile_cluster.py 1.7b -1 2 2>&1 | tail -19
