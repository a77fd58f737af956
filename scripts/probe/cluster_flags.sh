# A/B of the cluster kernel's experiment flags (QTC_FUSED_FLAGS): 1 = no idle-cluster L2 prefetch, 2 = no next-layer prefetch
for f in 3 2 1 0; do echo "== flags $f"; QTC_FUSED_FLAGS=$f STEPS=4 timeout 200 python scripts/cluster_check.py small 1.7b 2>&1 | grep "mode 2:\|worst\|codes equal"; done
