for f in 0 8; do
  echo "== flags $f"; QTC_FUSED_FLAGS=$f STEPS=4 timeout 200 python scripts/cluster_check.py small 1.7b 2>&1 | grep "mode 2:"
  QTC_FUSED_FLAGS=$f timeout 200 python scripts/phase_profile_cluster.py 1.7b 190 3 2>&1 | grep "B: \|layer total"
done
