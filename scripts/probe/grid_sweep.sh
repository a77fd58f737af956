for g in 148 128 112 96 64; do echo "== grid $g"; QTC_FUSED_CTAS=$g STEPS=4 timeout 200 python scripts/cluster_check.py small 1.7b 2>&1 | grep "mode 1:"; done
