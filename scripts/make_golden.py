"""Generate tests/golden/*.npz by running the UNMODIFIED reference (oracle/_ref/libqtts_ref.so,
built by oracle/Makefile from /root/reference/c) on seeded random-init models written by
harness/modelgen.py.  Run in the build container (needs /root/reference):

    make -C oracle ref && python scripts/make_golden.py

The vectors pin the oracle (tests/test_oracle_golden.py) and, through it, the CUDA path on the
GPU box where /root/reference does not exist.  The reference ships no golden vectors of its own
(SURVEY.md section 4); its only fixture is the 15-id prompt test/tokens_great_power.txt, used here.
"""
import ctypes as C
import os
import sys
import tempfile

import numpy as np

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, os.path.join(ROOT, "qwen3-tts-c_b200"))
sys.path.insert(0, ROOT)
from harness import binding as B, modelgen as mg  # noqa: E402
from oracle import orc_binding as O  # noqa: E402

TOKENS = [151644, 77091, 198, 2354, 2244, 2355, 4041, 2244, 11752, 13, 151645, 198, 151644, 77091, 198]
CSV = ",".join(str(t) for t in TOKENS).encode()


def run_generate(ref, ctx, params, n_frames, speaker=b"aiden", language=b"english"):
    ref.ref_set_params(ctx, C.byref(params))
    ns = C.c_int(0)
    ptr = ref.ref_generate(ctx, CSV, speaker, language, C.byref(ns))
    pcm = B.take_floats(ref.ref_free_buf, ptr, ns.value)
    T, G = ref.ref_last_frames(), ref.ref_last_groups()
    codes = np.zeros((T, G), dtype=np.int32)
    ref.ref_last_codes(B.iptr(codes))
    assert T == n_frames, (T, n_frames)
    return codes, pcm


def main():
    out = {}
    ref = O.load_ref()
    spec = mg.PRESETS["tiny"]
    main_t, cod_t = mg.build_tensors(spec, seed=0)
    with tempfile.TemporaryDirectory() as d:
        mg.write_model_dir(d, spec, main_t, cod_t)
        ctx = ref.ref_load(d.encode(), 0)
        assert ctx
        # whole utterances through qwen_tts_generate (c/qwen_tts.c:1059)
        for name, p in (("greedy", B.greedy_params(fixed_codec_tokens=24)),
                        ("sampled", B.default_params(fixed_codec_tokens=24)),
                        ("sampled_seed7", B.default_params(fixed_codec_tokens=16, sample_seed=7, temperature=1.3,
                                                           top_k=8, subtalker_top_k=3, repetition_penalty=1.3)),
                        ("nospeaker", B.greedy_params(fixed_codec_tokens=8))):
            spk = None if name == "nospeaker" else b"aiden"
            lang = None if name == "nospeaker" else b"english"
            codes, pcm = run_generate(ref, ctx, p, p.fixed_codec_tokens, spk, lang)
            out[f"gen_{name}_codes"], out[f"gen_{name}_pcm"] = codes, pcm
        # EOS path: max_new_tokens stop (no fixed budget)
        p = B.greedy_params(max_new_tokens=5)
        codes, pcm = run_generate(ref, ctx, p, 5)
        out["gen_maxtok_codes"], out["gen_maxtok_pcm"] = codes, pcm

        # stage API (c/qwen_tts.h:483-502) on seeded inputs
        H, V, G = spec["talker"]["hidden"], spec["talker"]["vocab"], spec["groups"]
        rng = np.random.default_rng(123)
        embeds = (rng.standard_normal((7, H)) * 0.5).astype(np.float32)
        ref.ref_set_params(ctx, C.byref(B.greedy_params()))
        ref.ref_reset(ctx)
        ref.ref_prefill(ctx, B.fptr(embeds), 7)
        hid = np.ctypeslib.as_array(ref.ref_hidden(ctx), shape=(H,)).copy()
        logits = np.zeros((5, V), dtype=np.float32)
        hids = np.zeros((5, H), dtype=np.float32)
        steps = (rng.standard_normal((5, H)) * 0.5).astype(np.float32)
        for i in range(5):
            ref.ref_forward(ctx, B.fptr(steps[i]), B.fptr(logits[i]))
            hids[i] = np.ctypeslib.as_array(ref.ref_hidden(ctx), shape=(H,))
        out.update(stage_embeds=embeds, stage_prefill_hidden=hid, stage_steps=steps, stage_logits=logits,
                   stage_hiddens=hids)
        sub_codes = np.zeros((3, 32), dtype=np.int32)
        for i, (pp, fc) in enumerate(((B.greedy_params(), 17), (B.default_params(), 901),
                                      (B.default_params(sample_seed=5, subtalker_top_k=4), 2047))):
            ref.ref_set_params(ctx, C.byref(pp))
            ref.ref_subtalker(ctx, B.fptr(hids[i]), fc, B.iptr(sub_codes[i]))
        out["stage_sub_codes"] = sub_codes[:, :G].copy()
        # codec only, T > sliding window, includes an out-of-range code (-> index 0, c/qwen_tts_codec.c:169)
        cc = rng.integers(0, spec["codec"]["cb_size"], size=(21, 16), dtype=np.int32)
        cc[3, 5] = 99999
        ns = C.c_int(0)
        ptr = ref.ref_codec_decode(ctx, B.iptr(cc), 21, C.byref(ns))
        out["codec_codes"], out["codec_pcm"] = cc, B.take_floats(ref.ref_free_buf, ptr, ns.value)
        ref.ref_free(ctx)

    # kernel-level known answers (c/qwen_tts_kernels.c)
    rng = np.random.default_rng(7)
    lg = (rng.standard_normal(3072) * 4).astype(np.float32)
    lg[100] = lg[2000] = lg.max() + 0.5          # exact tie at the top: lower index must win
    picks = []
    for k, tp, temp, seed in ((1, 1.0, 1.0, 42), (50, 1.0, 0.9, 42), (50, 1.0, 0.9, 43), (5, 1.0, 2.0, 1),
                              (50, 0.8, 0.9, 42), (0, 0.5, 1.0, 9), (3072, 1.0, 1.0, 3), (50, 1.0, 0.0, 42)):
        st = np.array([seed], dtype=np.float32)
        seq = []
        for _ in range(6):
            seq.append(ref.kernel_sample_top_k(B.fptr(lg), 3072, k, C.c_float(tp), C.c_float(temp), B.fptr(st)))
        picks.append(seq + [int(st.view(np.uint32)[0])])
    out["sample_logits"], out["sample_picks"] = lg, np.array(picks, dtype=np.int64)
    A = mg.f32_to_bf16(rng.standard_normal((96, 256)).astype(np.float32))
    x = rng.standard_normal(256).astype(np.float32)
    y = np.zeros(96, dtype=np.float32)
    ref.kernel_matvec_bf16(B.fptr(y), B.uptr(A), B.fptr(x), 96, 256)
    out.update(mv_A=A, mv_x=x, mv_y=y)
    for name, (ci, co, k, L, dil, grp) in dict(conv_k7d3=(12, 10, 7, 50, 3, 1), conv_dw=(8, 8, 7, 33, 1, 8),
                                               conv_k1=(6, 9, 1, 20, 1, 1), conv_short=(4, 4, 7, 5, 9, 1)).items():
        xi = rng.standard_normal((ci, L)).astype(np.float32)
        w = rng.standard_normal((co, ci // grp, k)).astype(np.float32)
        b = rng.standard_normal(co).astype(np.float32)
        yo = np.zeros((co, L), dtype=np.float32)
        ref.kernel_causal_conv1d(B.fptr(yo), B.fptr(xi), B.fptr(w), B.fptr(b), ci, co, k, L, dil, grp)
        out.update({f"{name}_x": xi, f"{name}_w": w, f"{name}_b": b, f"{name}_y": yo,
                    f"{name}_p": np.array([ci, co, k, L, dil, grp])})
    for name, (ci, co, k, s, L) in dict(tconv_16_8=(6, 4, 16, 8, 11), tconv_2_2=(5, 5, 2, 2, 9)).items():
        xi = rng.standard_normal((ci, L)).astype(np.float32)
        w = rng.standard_normal((ci, co, k)).astype(np.float32)
        b = rng.standard_normal(co).astype(np.float32)
        yo = np.zeros((co, L * s + k), dtype=np.float32)
        nl = C.c_int(0)
        ref.kernel_transposed_conv1d(B.fptr(yo), B.fptr(xi), B.fptr(w), B.fptr(b), ci, co, k, s, L, C.byref(nl))
        out.update({f"{name}_x": xi, f"{name}_w": w, f"{name}_b": b,
                    f"{name}_y": yo.reshape(-1)[:co * nl.value].reshape(co, nl.value).copy(),
                    f"{name}_p": np.array([ci, co, k, s, L])})
    path = os.path.join(ROOT, "tests", "golden", "ref_tiny_seed0.npz")
    np.savez_compressed(path, **out)
    print("wrote", path, os.path.getsize(path), "bytes;", len(out), "arrays")


if __name__ == "__main__":
    main()
