for cfg in "0 -1 -1" "0 50 2000" "0 50 3000" "0 60 2000" "0 40 2000" "0 50 1500" "2 50 2000"; do
  set -- $cfg
  echo "flags=$1 frac=$2 gap=$3: $(QTC_FUSED_FLAGS=$1 QTC_PF_FRAC=$2 QTC_PF_GAP=$3 STEPS=4 timeout 200 python scripts/cluster_check.py small 1.7b 2>&1 | grep 'mode 2:')"
done
