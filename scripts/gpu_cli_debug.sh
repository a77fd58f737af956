#!/bin/bash
# run the reference's CLI on the cuda backend in persistent mode and show its stderr
python - <<'PY'
import os, sys, subprocess, tempfile
sys.path.insert(0, "qwen3-tts-c_b200")
from harness import modelgen as mg
spec = mg.PRESETS["tiny"]; main, cod = mg.build_tensors(spec, seed=3)
d = tempfile.mkdtemp(); mg.write_model_dir(d, spec, main, cod)
ids = "151644,77091,198,2354,2244,2355,4041,2244,11752,13,151645,198,151644,77091,198"
for extra in (["-v", "-v"], ["--benchmark-runs", "2", "--benchmark-warmup", "1"]):
    r = subprocess.run(["oracle/_ref/qwen-tts-cuda", "--backend", "cuda", "-d", d, "-t", ids, "-s", "aiden", "-l", "english", "-o", d + "/o.wav", "--fixed-codec-tokens", "6", *extra], capture_output=True, text=True)
    print("rc", r.returncode); print(r.stderr[-1500:])
PY
