"""Per-phase timing of the cluster kernel (mode 2) from in-kernel clock64 stamps (PROF slots of qtc_cluster.cuh).
Usage (GPU box): python scripts/phase_profile_cluster.py [preset] [pos (<0: sub-talker)] [layer]"""
import ctypes as C
import os
import sys

import numpy as np

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, os.path.join(ROOT, "qwen3-tts-c_b200"))
from harness import binding as B, modelgen as mg  # noqa: E402

NAMES = ["A: zero + kv stage + rmsnorm", "A: qkv rows + push", "A: cluster sync 1", "A: attention + O rows requested", "A: cluster sync 2",
         "A: combine + cluster sync 3", "A: O partial + atomics", "grid barrier 1", "B: zero + rmsnorm", "B: gate/up + swiglu", "B: down partial",
         "B: cluster sync 4", "B: DSMEM fold + atomics", "grid barrier 2"]


def main():
    preset = sys.argv[1] if len(sys.argv) > 1 else "1.7b"
    pos = int(sys.argv[2]) if len(sys.argv) > 2 else 190
    layer = int(sys.argv[3]) if len(sys.argv) > 3 else 3
    cuda = B.load_cuda()
    assert cuda.qwen_tts_cuda_init(0) == 0
    spec = mg.PRESETS[preset]
    main_t, cod = mg.build_tensors(spec, seed=0, with_codec=False)
    hm = B.HostModel(spec, main_t, {})
    ctx = cuda.qwen_tts_cuda_upload(C.byref(hm.desc), 1)
    assert ctx
    cuda.qwen_tts_cuda_set_mode(ctx, 2)
    cuda.qwen_tts_cuda_debug_phase_clocks.argtypes = [C.c_void_p, C.c_int, C.c_int, C.POINTER(C.c_longlong)]
    out = np.zeros(148 * 32 * 2, dtype=np.int64)
    n = cuda.qwen_tts_cuda_debug_phase_clocks(ctx, pos, layer, out.ctypes.data_as(C.POINTER(C.c_longlong)))
    assert n > 0
    raw = out[: n * 32].reshape(n, 32).astype(np.float64)
    used = raw[:, 14] > 0
    raw = raw[used]
    att = raw[:, 3] > 0                      # clusters that own a KV head stamp slots 1..6
    print(f"{preset} pos={pos} layer={layer}: {used.sum()} CTAs stamped, {att.sum()} in KV-head clusters; microseconds (min / median / max)")
    d = np.diff(raw[:, :15], axis=1) / 1.965e3
    for i, name in enumerate(NAMES):
        rows = d[att, i] if 0 <= i <= 5 else d[:, i]
        if i == 6:
            rows = d[att, i]
        if i in (0, 6) or (1 <= i <= 5):
            pass
        print(f"  {name:34s} {rows.min():7.2f} {np.median(rows):7.2f} {rows.max():7.2f}")
    for nm, a_, b_ in (("attn: q/k norm + rope", 3, 16), ("attn: scores + softmax + V", 16, 17), ("attn: records + block barrier", 17, 18), ("attn: fold warps", 18, 22), ("attn: push records to peers", 22, 19),
                       ("attn: request O rows", 19, 4), ("O: request batch 2 + att load", 6, 20), ("O: batch 1 dots + butterflies", 20, 23), ("O: batch 1 convert + atomics", 23, 21), ("O: batch 2", 21, 7)):
        if raw[att, b_].any() and raw[att, a_].any():
            dd = (raw[att, b_] - raw[att, a_]) / 1.965e3
            print(f"    [{nm:30s}] {dd.min():7.2f} {np.median(dd):7.2f} {dd.max():7.2f}")
    if pos < 0 and raw[:, 16].any():
        for nm, a_, b_ in (("sampler: scale + moments", 16, 17), ("sampler: threshold + gather", 17, 18), ("sampler: rank candidates", 18, 19), ("sampler: expf", 19, 20),
                           ("sampler: sequential sum + walk", 20, 21), ("sampler: last barrier", 21, 22)):
            dd = (raw[:, b_] - raw[:, a_]) / 1.965e3
            print(f"    [{nm:36s}] {dd.min():7.2f} {np.median(dd):7.2f} {dd.max():7.2f}")
        print(f"    [sampler: candidates] {raw[:, 23].min():.0f} .. {raw[:, 23].max():.0f}")
    sm = raw[:, 24] > 0
    if sm.any():
        for nm, a_, b_ in (("small attn: entry after sync", 3, 24), ("small attn: norms + rope + kv write", 24, 25), ("small attn: scores", 25, 26), ("small attn: softmax", 26, 27),
                           ("small attn: V", 27, 28), ("small attn: push to peers", 28, 29)):
            dd = (raw[sm, b_] - raw[sm, a_]) / 1.965e3
            print(f"    [{nm:36s}] {dd.min():7.2f} {np.median(dd):7.2f} {dd.max():7.2f}")
    idle = ~att
    if idle.any():
        w = (raw[idle, 8] - raw[idle, 0]) / 1.965e3
        print(f"  (clusters without a KV head wait through phase A: {np.median(w):.2f} us)")
    tot = (raw[:, 14] - raw[:, 0]) / 1.965e3
    print(f"  layer total                        {tot.min():7.2f} {np.median(tot):7.2f} {tot.max():7.2f}")
    if pos >= 0:
        step = cuda.qwen_tts_cuda_bench_talker_step(ctx, pos, 20, 1)
        print(f"  talker step {step * 1e3:.1f} us, bytes {cuda.qwen_tts_cuda_talker_step_bytes(ctx, pos) / 1e6:.1f} MB")
    else:
        fr = cuda.qwen_tts_cuda_bench_subtalker(ctx, 20, 1)
        print(f"  sub-talker frame {fr * 1e3:.1f} us (16 steps), streamed bytes {cuda.qwen_tts_cuda_subtalker_frame_bytes(ctx) / 1e6:.1f} MB")
    cuda.qwen_tts_cuda_release(ctx)


if __name__ == "__main__":
    main()
