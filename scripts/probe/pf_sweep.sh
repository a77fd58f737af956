# idle-cluster L2 prefetch: fraction of every consumer's MLP slice x pacing; flag 4 = also for the talker, flag 2 = no next-layer attention prefetch
for cfg in "0 0 0" "4 100 0" "4 100 1000" "4 100 3000" "4 50 0" "4 50 2000" "4 25 0" "6 100 1000" "6 50 1000"; do
  set -- $cfg
  echo "flags=$1 frac=$2 gap=$3: $(QTC_FUSED_FLAGS=$1 QTC_PF_FRAC=$2 QTC_PF_GAP=$3 STEPS=4 timeout 200 python scripts/cluster_check.py small 1.7b 2>&1 | grep 'mode 2:')"
done
