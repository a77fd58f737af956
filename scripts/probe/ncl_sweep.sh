# sub-talker frame time against the number of 8-CTA clusters in the grid (the talker needs >= 12 and falls back to mode 1 below that)
for n in 15 14 12 10 8; do echo "clusters=$n: $(QTC_CLUSTERS=$n STEPS=4 timeout 200 python scripts/cluster_check.py small 1.7b 2>&1 | grep 'mode 2:')"; done
