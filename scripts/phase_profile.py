"""Per-phase timing of the fused talker step (mode 1) from in-kernel clock64 stamps.
Usage (GPU box): python scripts/phase_profile.py [preset] [pos] [layer]"""
import ctypes as C
import os
import sys

import numpy as np

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, os.path.join(ROOT, "qwen3-tts-c_b200"))
from harness import binding as B, modelgen as mg  # noqa: E402

NAMES = ["rmsnorm1", "qkv gemv", "barrier", "attention", "barrier", "o gemv(+stage)", "barrier", "rmsnorm2", "gate/up", "barrier",
         "stage g", "down gemv", "barrier"]


def main():
    preset = sys.argv[1] if len(sys.argv) > 1 else "1.7b"
    pos = int(sys.argv[2]) if len(sys.argv) > 2 else 190
    layer = int(sys.argv[3]) if len(sys.argv) > 3 else 5
    cuda = B.load_cuda()
    assert cuda.qwen_tts_cuda_init(0) == 0
    spec = mg.PRESETS[preset]
    main_t, cod = mg.build_tensors(spec, seed=0, with_codec=False)
    hm = B.HostModel(spec, main_t, {})
    ctx = cuda.qwen_tts_cuda_upload(C.byref(hm.desc), 1)
    assert ctx
    cuda.qwen_tts_cuda_set_mode(ctx, int(os.environ.get("QTC_MODE", "1")))
    cuda.qwen_tts_cuda_debug_phase_clocks.argtypes = [C.c_void_p, C.c_int, C.c_int, C.POINTER(C.c_longlong)]
    out = np.zeros(148 * 32 * 2, dtype=np.int64)
    n = cuda.qwen_tts_cuda_debug_phase_clocks(ctx, pos, layer, out.ctypes.data_as(C.POINTER(C.c_longlong)))
    assert n > 0
    raw = out[: n * 32].reshape(n, 32).astype(np.float64)
    for nm, a, b in (("qkv consume only", 1, 16), ("stage (P3)", 5, 15)):
        dd = (raw[:, b] - raw[:, a]) / 1.965e3
        print(f"  [{nm}] {dd.min():7.2f} {np.median(dd):7.2f} {dd.max():7.2f}")
    if raw[:, 20].any():
        for nm, a_, b_ in (("P3 issue cp.async", 5, 20), ("P3 wait_all", 20, 21), ("P3 sync", 21, 22), ("P3 fac", 22, 23), ("P3 sync", 23, 24), ("P3 combine+sync", 24, 15)):
            dd = (raw[:, b_] - raw[:, a_]) / 1.965e3
            print(f"  [{nm:18s}] {dd.min():7.2f} {np.median(dd):7.2f} {dd.max():7.2f}")
    if raw[:, 17].any():
        names = {26: "before geom", 17: "consume entry", 18: "x slice loaded", 19: "quad0 loaded", 20: "quad0 mma done", 21: "quad0 released", 22: "chunk1 done",
                 23: "chunk2 full", 24: "chunk2 done", 25: "loop end", 16: "after consume", 27: "PRODUCER at qkv", 28: "PRODUCER got slot"}
        for k in (27, 28, 26, 17, 19, 20, 21, 25, 16):
            dd = (raw[:, k] - raw[:, 1]) / 1.965e3
            print(f"  [qkv +{names[k]:16s}] {dd.min():7.2f} {np.median(dd):7.2f} {dd.max():7.2f}")
    st = raw[:, :14]
    d = np.diff(st, axis=1) / 1.965e3          # us at the 1965 MHz the bench runs at
    print(f"{preset} pos={pos} layer={layer}: per-phase microseconds over {n} CTAs (min / median / max)")
    for i, name in enumerate(NAMES):
        print(f"  {name:16s} {d[:, i].min():7.2f} {np.median(d[:, i]):7.2f} {d[:, i].max():7.2f}")
    tot = (st[:, 13] - st[:, 0]) / 1.965e3
    print(f"  layer total      {tot.min():7.2f} {np.median(tot):7.2f} {tot.max():7.2f}")
    if pos >= 0:
        step = cuda.qwen_tts_cuda_bench_talker_step(ctx, pos, 20, 1)
        print(f"  talker step {step * 1e3:.1f} us, bytes {cuda.qwen_tts_cuda_talker_step_bytes(ctx, pos) / 1e6:.1f} MB")
    else:
        fr = cuda.qwen_tts_cuda_bench_subtalker(ctx, 20, 1)
        print(f"  sub-talker frame {fr * 1e3:.1f} us (16 steps), streamed bytes {cuda.qwen_tts_cuda_subtalker_frame_bytes(ctx) / 1e6:.1f} MB")
    cuda.qwen_tts_cuda_release(ctx)


if __name__ == "__main__":
    main()
