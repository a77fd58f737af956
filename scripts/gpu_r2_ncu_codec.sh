#!/bin/bash
# evidence for the tensor-core vocoder: launch list of one 3600-frame decode, and ncu --set full of a few k_conv_tc launches
TAG=${1:-ncuc}; OUT=gpurun_out/$TAG; mkdir -p $OUT
timeout 600 python bench.py --codec-frames 3600 --steps 1 --warmup 1 --no-cpu > $OUT/plain.json 2> $OUT/plain.err; echo "plain exit $?"
timeout 900 ncu --metrics gpu__time_duration.sum --clock-control none -s 444 -c 444 --csv --log-file $OUT/launches_codec_3600f.csv python bench.py --codec-frames 3600 --steps 1 --warmup 1 --no-cpu > $OUT/ncu_list.log 2>&1; echo "ncu list exit $?"
timeout 1200 ncu --set full --clock-control none --import-source on -k regex:k_conv_tc -s 27 -c 8 -o $OUT/conv_tc_full python bench.py --codec-frames 360 --steps 1 --warmup 0 --no-cpu > $OUT/ncu_full.log 2>&1; echo "ncu full exit $?"
ncu -i $OUT/conv_tc_full.ncu-rep --page raw --csv > $OUT/ncu_full_k_conv_tc.csv 2>/dev/null; ls -la $OUT
python - $OUT/launches_codec_3600f.csv <<'PY'
import csv, collections, sys
rows = list(csv.reader(open(sys.argv[1] if len(sys.argv) > 1 else "gpurun_out/$TAG/launches_codec_3600f.csv", errors="ignore")))
hdr = None; agg = collections.defaultdict(lambda: [0, 0.0])
for r in rows:
    if "Kernel Name" in r: hdr = r; continue
    if hdr and len(r) == len(hdr):
        d = dict(zip(hdr, r))
        if d.get("Metric Name") == "gpu__time_duration.sum":
            v = float(d["Metric Value"].replace(",", "")); u = d.get("Metric Unit", "ns")
            v = v / 1e3 if u in ("ns", "nsecond") else (v * 1e3 if u in ("ms", "msecond") else v)
            k = d["Kernel Name"].split("(")[0]; agg[k][0] += 1; agg[k][1] += v
tot = sum(v[1] for v in agg.values())
for k, (n, t) in sorted(agg.items(), key=lambda x: -x[1][1]): print(f"{k:40s} {n:5d} launches {t/1e3:9.2f} ms {100*t/tot:5.1f} %")
print("total", tot / 1e3, "ms")
PY
