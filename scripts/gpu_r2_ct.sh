#!/bin/bash
TAG=${1:-ct}; OUT=gpurun_out/$TAG; mkdir -p $OUT
rm -f gpurun_out/snr_*.log; timeout 900 python -m pytest tests/test_gpu_conv_tc.py tests/test_gpu_stream.py tests/test_gpu_fullshape_parity.py -q -x --durations=5 > $OUT/pytest.log 2>&1; echo "pytest exit $?"; tail -12 $OUT/pytest.log
timeout 600 python bench.py --codec-frames 3600 --steps 3 --warmup 2 --no-cpu > $OUT/bench_codec.json 2> $OUT/bench_codec.err; echo "codec bench exit $?"; cat $OUT/bench_codec.json; tail -3 $OUT/bench_codec.err
timeout 900 ncu --metrics gpu__time_duration.sum --clock-control none -s 296 -c 148 --csv --log-file $OUT/launches_codec_3600f.csv python bench.py --codec-frames 3600 --steps 1 --warmup 2 --no-cpu > $OUT/ncu_list.log 2>&1; echo "ncu list exit $?"
python - $OUT/launches_codec_3600f.csv <<'PY'
import csv, collections, sys
rows = list(csv.reader(open(sys.argv[1], errors="ignore")))
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
cat gpurun_out/snr_fullshape.log | tail -3
