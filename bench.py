#!/usr/bin/env python
"""bench.py - headline benchmark of the B200 backend for Qwen3-TTS-C's per-frame generation path.

Workload (BASELINE.json configs[1]): Qwen3-TTS-12Hz-1.7B-CustomVoice shapes, batch 1, one utterance of
360 frames (28.8 s of 24 kHz audio) from the reference's fixture prompt, default sampling, fixed frame
budget instead of EOS (random-init weights: no checkpoints offline).  One "step" = one whole utterance:
prompt build + prefill + 360 x {talker step, sample, 15-codebook sub-talker, next embedding} + codec decode.

  value : audio-seconds per second with the prompt already in HBM (CUDA events, prompt build .. last codec kernel)
  e2e   : the same through the C-ABI call a user makes (qwen_tts_cuda_generate: host token string in,
          host PCM out), wall clock around the call, H2D/D2H inside
  roofline : talker decode step, algorithmic bytes / event-timed step, against MEASURED_PEAKS.json hbm_gbs
  cpu_baseline / --impl reference : the reference's own C path (oracle/_ref, unmodified sources) on the
          host cores, bounded sample (see REF_FRAMES)

N > 1 (torchrun): utterances are independent, so every rank decodes its own utterance on its own GPU with
replicated weights and no collective on the data path ("weak" scaling); torch.distributed carries only the
barrier and the max-over-ranks of the device time.
"""
import argparse
import ctypes as C
import json
import os
import subprocess
import sys
import threading
import time

import numpy as np

ROOT = os.path.dirname(os.path.abspath(__file__))
sys.path.insert(0, os.path.join(ROOT, "qwen3-tts-c_b200"))
from harness import binding as B, modelgen as mg, shard  # noqa: E402

TOKENS = [151644, 77091, 198, 2354, 2244, 2355, 4041, 2244, 11752, 13, 151645, 198, 151644, 77091, 198]
CSV = ",".join(str(t) for t in TOKENS).encode()
FRAMES = 360
FRAME_S = 1920 / 24000.0
REF_FRAMES = 2          # frames per reference step: ~2.7 s of single-thread GEMV work per frame on the 1.7B shapes
METRIC = "audio-sec generated/sec (RTF), 1.7B-CustomVoice batch 1, 360 frames"


def ncu_traffic(mode):
    """dram__bytes_read.sum + dram__bytes_write.sum of one talker-step launch, from the committed ncu --set full capture"""
    p = os.path.join(ROOT, "profiles", "traffic.json")
    try:
        return json.load(open(p)).get(f"talker_step_mode{mode}")
    except Exception:
        return None


def peaks():
    p = os.path.join(ROOT, "MEASURED_PEAKS.json")
    if os.path.exists(p):
        return json.load(open(p)).get("hbm_gbs", 6650.0), "measured"
    return 6650.0, "fallback"


class ClockSampler(threading.Thread):
    """nvidia-smi clocks and throttle reasons DURING the timed region"""
    Q = "clocks.sm,clocks.max.sm,clocks_event_reasons.hw_slowdown,clocks_event_reasons.hw_thermal_slowdown," \
        "clocks_event_reasons.sw_thermal_slowdown,clocks_event_reasons.sw_power_cap"

    def __init__(self, index):
        super().__init__(daemon=True)
        self.index, self.rows, self.stop_flag = index, [], False

    def run(self):
        while not self.stop_flag:
            try:
                out = subprocess.check_output(["nvidia-smi", "-i", str(self.index), f"--query-gpu={self.Q}",
                                               "--format=csv,noheader,nounits"], timeout=5).decode().strip()
                self.rows.append([x.strip() for x in out.split(",")])
            except Exception:
                pass
            time.sleep(0.2)

    def summary(self):
        if not self.rows:
            return {"sm_mhz": None, "sm_max_mhz": None, "reasons": ["unavailable"]}
        sm = sorted(float(r[0]) for r in self.rows if r[0].replace(".", "").isdigit())
        names = ["hw_slowdown", "hw_thermal_slowdown", "sw_thermal_slowdown", "sw_power_cap"]
        reasons = [n for i, n in enumerate(names) if any(len(r) > 2 + i and r[2 + i].lower().startswith("active") for r in self.rows)]
        return {"sm_mhz": sm[len(sm) // 2] if sm else None, "sm_max_mhz": float(self.rows[0][1]) if self.rows[0][1].replace(".", "").isdigit() else None,
                "reasons": reasons, "samples": len(self.rows)}


def build_model(preset):
    spec = mg.PRESETS[preset]
    main, cod = mg.build_tensors(spec, seed=0)
    return spec, main, cod, B.HostModel(spec, main, cod)


def load_checker():
    """oracle/ is test infrastructure: only the CPU legs of this file (cpu_baseline, --impl reference) load it"""
    sys.path.insert(0, ROOT)
    from oracle import orc_binding as O
    return O


def run_reference(args, rank, world):
    """The reference's own C path on the host cores (oracle/_ref/libqtts_ref.so = unmodified c/*.c + shim).
    Setup (untimed): load the model directory, prefill the fixture prompt.  One step = REF_FRAMES frames of
    the frame loop (talker forward, logits policy, reference sampler, sub-talker) + codec decode of those
    frames.  Rank 0 only."""
    if rank != 0:
        return
    O = load_checker()
    if not O.have_ref():
        print(json.dumps({"impl": "reference", "unavailable": "oracle/_ref/libqtts_ref.so not built (no /root/reference at build time)"}))
        return
    import tempfile
    ref = O.load_ref()
    orc = O.load_oracle()
    spec, main, cod, hm = build_model(args.model)
    d = tempfile.mkdtemp(prefix="qtts_model_", dir="/dev/shm" if os.path.isdir("/dev/shm") else None)
    try:
        mg.write_model_dir(d, spec, main, cod)
        ctx = ref.ref_load(d.encode(), 0)
        assert ctx, "reference failed to load the model"
        H, V, G = spec["talker"]["hidden"], spec["talker"]["vocab"], spec["groups"]
        p = B.default_params(fixed_codec_tokens=REF_FRAMES)
        ref.ref_set_params(ctx, C.byref(p))
        # prompt through the oracle's restatement of the prompt builder (untimed setup; bit-identical to the reference's)
        s = orc.orc_create(hm.desc)
        pre, tr, pad = C.c_void_p(), C.c_void_p(), C.c_void_p()
        plen, ntr = C.c_int(0), C.c_int(0)
        tok = np.array(TOKENS, dtype=np.int32)
        orc.orc_build_prompt(s, B.iptr(tok), len(tok), mg.SPEAKERS["aiden"], mg.LANGUAGES["english"], C.byref(pre), C.byref(plen),
                             C.byref(tr), C.byref(ntr), C.byref(pad))
        trailing = np.ctypeslib.as_array(C.cast(tr, B.f32p), shape=(ntr.value, H)).copy()
        padv = np.ctypeslib.as_array(C.cast(pad, B.f32p), shape=(H,)).copy()
        ref.ref_reset(ctx)
        ref.ref_prefill(ctx, C.cast(pre, B.f32p), plen.value)
        logits = np.zeros(V, dtype=np.float32)
        hidden0 = np.ctypeslib.as_array(ref.ref_hidden(ctx), shape=(H,)).copy()
        ref.kernel_matvec_bf16(B.fptr(logits), hm.desc.codec_head, B.fptr(hidden0), V, H)
        emb_t = mg.bf16_to_f32(main["talker.model.codec_embedding.weight"])
        sub_emb = [main[f"talker.code_predictor.model.codec_embedding.{g}.weight"] for g in range(G - 1)]
        rng = np.array([p.sample_seed], dtype=np.float32)
        ref.kernel_apply_repetition_penalty.argtypes = [B.f32p, B.i32p, C.c_int, C.c_int, C.c_float]
        hist, step_idx = [], 0

        def one_step():
            nonlocal logits, step_idx
            codes = np.zeros((REF_FRAMES, G), dtype=np.int32)
            for f in range(REF_FRAMES):
                lg = logits.copy()
                lg[V - 1024:] = np.where(np.arange(V - 1024, V) == hm.desc.codec_eos_id, lg[V - 1024:], -1e9)
                h = np.array(hist, dtype=np.int32) if hist else np.zeros(1, dtype=np.int32)
                ref.kernel_apply_repetition_penalty(B.fptr(lg), B.iptr(h), len(hist), V, C.c_float(p.repetition_penalty))
                t = ref.kernel_sample_top_k(B.fptr(lg), V, p.top_k, C.c_float(p.top_p), C.c_float(p.temperature), B.fptr(rng))
                if t == hm.desc.codec_eos_id:
                    lg[t] = -1e9
                    t = ref.kernel_sample_top_k(B.fptr(lg), V, p.top_k, C.c_float(p.top_p), C.c_float(p.temperature), B.fptr(rng))
                hist.append(t)
                out = np.zeros(32, dtype=np.int32)
                hid = np.ctypeslib.as_array(ref.ref_hidden(ctx), shape=(H,)).copy()
                ref.ref_subtalker(ctx, B.fptr(hid), t, B.iptr(out))
                codes[f] = out[:G]
                nxt = emb_t[t].copy()
                for g in range(1, G):
                    nxt += mg.bf16_to_f32(sub_emb[g - 1][out[g]])
                nxt += trailing[step_idx] if step_idx < len(trailing) else padv
                step_idx += 1
                ref.ref_forward(ctx, B.fptr(np.ascontiguousarray(nxt, dtype=np.float32)), B.fptr(logits))
            ns = C.c_int(0)
            ptr = ref.ref_codec_decode(ctx, B.iptr(codes), REF_FRAMES, C.byref(ns))
            ref.ref_free_buf(ptr)

        for _ in range(args.warmup):
            one_step()
        t0 = time.time()
        for _ in range(args.steps):
            one_step()
        dt = time.time() - t0
        rtf = args.steps * REF_FRAMES * FRAME_S / dt
        cores = os.cpu_count()
        # the same path through the oracle port, whose GEMVs are OpenMP loops over rows (the reference's are single-threaded by
        # construction): both CPU figures travel in this line, so the GPU / CPU ratio is never quoted against the handicapped one alone
        port_rtf, port_dt = cpu_baseline_sample(hm, spec, main, frames=1)
        line = {"impl": "reference", "metric": METRIC, "value": rtf, "unit": "audio_s/s", "n_gpus": args.gpus, "steps": args.steps,
                "warmup": args.warmup, "ms_per_step": dt / args.steps * 1e3, "higher_is_better": True, "scaling": "weak",
                "vs_baseline": None, "dtype": "f32 (bf16 weights)", "data": "synthetic",
                "config": {"workload": "1.7B-CustomVoice shapes, batch 1, default sampling, random-init weights",
                           "sample": f"{REF_FRAMES} frames of the frame loop + codec decode of them per step, after an untimed prefill"},
                "cpu_baseline": {"value": rtf, "unit": "audio_s/s", "cores": cores, "kind": "reference",
                                 "sample": f"{args.steps} x {REF_FRAMES} frames; GEMVs single-threaded (no BLAS on the image), codec convs OpenMP x{cores}",
                                 "port_all_cores": {"value": port_rtf, "unit": "audio_s/s", "cores": cores, "kind": "port",
                                                    "sample": f"1 frame + its codec decode after an untimed prefill, {port_dt:.1f} s; oracle port, OpenMP over output rows"}},
                "ms_per_frame": dt / (args.steps * REF_FRAMES) * 1e3,
                "e2e": {"value": rtf, "unit": "audio_s/s", "h2d_bytes_per_step": 0, "d2h_bytes_per_step": 0}}
        print(json.dumps(line))
        ref.ref_free(ctx)
    finally:
        import shutil
        shutil.rmtree(d, ignore_errors=True)


def cpu_baseline_sample(hm, spec, main, frames=1):
    """bounded CPU sample for the default line: the oracle port (OpenMP over rows) on `frames` frames after prefill"""
    orc = load_checker().load_oracle()
    s = orc.orc_create(hm.desc)
    p = B.default_params(fixed_codec_tokens=frames)
    orc.orc_set_params(s, C.byref(p))
    tok = np.array(TOKENS, dtype=np.int32)
    pre, tr, pad = C.c_void_p(), C.c_void_p(), C.c_void_p()
    plen, ntr = C.c_int(0), C.c_int(0)
    orc.orc_build_prompt(s, B.iptr(tok), len(tok), mg.SPEAKERS["aiden"], mg.LANGUAGES["english"], C.byref(pre), C.byref(plen),
                         C.byref(tr), C.byref(ntr), C.byref(pad))
    H, V, G = spec["talker"]["hidden"], spec["talker"]["vocab"], spec["groups"]
    orc.orc_reset(s)
    orc.orc_talker_prefill(s, C.cast(pre, B.f32p), plen.value)
    t0 = time.time()
    logits = np.zeros(V, dtype=np.float32)
    codes = np.zeros((frames, G), dtype=np.int32)
    emb = np.zeros(H, dtype=np.float32)
    for f in range(frames):
        out = np.zeros(32, dtype=np.int32)
        orc.orc_subtalker_generate(s, orc.orc_hidden(s), 100 + f, B.iptr(out))
        codes[f] = out[:G]
        orc.orc_talker_forward(s, B.fptr(emb), B.fptr(logits))
    ns = C.c_int(0)
    ptr = orc.orc_codec_decode(s, B.iptr(codes), frames, C.byref(ns))
    orc.orc_free(ptr)
    dt = time.time() - t0
    orc.orc_destroy(s)
    return frames * FRAME_S / dt, dt


def measure_batch(args, cuda, ctx, spec, rank, world, local, dist, torch, n_utt, frames, steps, warmup, with_pcm=True):
    """configs 4-5: this rank's n_utt utterances decoded as lock-step batches of <= 64 (tcgen05 GEMM path), fixed frame budget,
    codes AND PCM (every utterance's codes go through the codec decoder; PCM lands in a host buffer).  Returns the numbers
    of the slowest rank (device time: prompt build .. last codec kernel of every batch)."""
    G = spec["groups"]
    cuda.qwen_tts_cuda_generate_batch.argtypes = [C.c_void_p, C.c_int, B.i32p, B.i32p, B.i32p, B.i32p, B.i32p, B.i32p, B.f32p, C.c_int]
    cuda.qwen_tts_cuda_set_batch_pieces.argtypes = [C.c_void_p, C.c_int]
    assert cuda.qwen_tts_cuda_set_batch_pieces(ctx, args.pieces) == 0
    p = B.default_params(fixed_codec_tokens=frames, sample_seed=1000 * rank + 42)
    cuda.qwen_tts_cuda_set_params(ctx, C.byref(p))
    rng = np.random.default_rng(rank)
    batches = []
    for lo, hi in shard.micro_batches(0, n_utt, 64):
        nb = hi - lo
        prompts = []
        for b in range(nb):                   # fixture prompt with a random-length run of text ids in the middle
            extra = [int(x) for x in rng.integers(1000, 5000, size=int(rng.integers(0, 24)))]
            prompts.append(TOKENS[:4] + extra + TOKENS[4:])
        batches.append(dict(nb=nb, toks=np.array(sum(prompts, []), dtype=np.int32),
                            offs=np.array([0] + list(np.cumsum([len(q) for q in prompts])), dtype=np.int32),
                            spk=np.full(nb, mg.SPEAKERS["aiden"], dtype=np.int32), lang=np.full(nb, mg.LANGUAGES["english"], dtype=np.int32),
                            codes=np.zeros((nb, frames, G), dtype=np.int32), nfr=np.zeros(nb, dtype=np.int32),
                            pcm=np.zeros((nb, frames * 1920), dtype=np.float32) if with_pcm else None))
    perf = B.Perf()

    def one():
        t0 = time.perf_counter()
        dev = loop = 0.0
        for bt in batches:
            rc = cuda.qwen_tts_cuda_generate_batch(ctx, bt["nb"], B.iptr(bt["toks"]), B.iptr(bt["offs"]), B.iptr(bt["spk"]), B.iptr(bt["lang"]), B.iptr(bt["codes"]),
                                                   B.iptr(bt["nfr"]), B.fptr(bt["pcm"]) if with_pcm else None, frames)
            if rc != 0 or int(bt["nfr"].min()) != frames:
                raise RuntimeError("generate_batch failed: " + cuda.qwen_tts_cuda_last_error().decode())
            cuda.qwen_tts_cuda_get_perf(ctx, C.byref(perf))
            dev += perf.device_ms / 1e3
            loop += perf.talker_ms / 1e3
        return time.perf_counter() - t0, dev, loop

    for _ in range(warmup):
        one()
    cuda.qwen_tts_cuda_get_perf(ctx, C.byref(perf))
    l0 = perf.kernel_launches
    sampler = ClockSampler(local)
    sampler.start()
    if dist is not None:
        dist.barrier()
    torch.cuda.synchronize()
    t0 = time.perf_counter()
    devs, loops = [], []
    for _ in range(steps):
        w, dv, lp = one()
        devs.append(dv); loops.append(lp)
    if dist is not None:
        dist.barrier()
    torch.cuda.synchronize()
    wall_total, dev_total = time.perf_counter() - t0, sum(devs)
    sampler.stop_flag = True
    cuda.qwen_tts_cuda_get_perf(ctx, C.byref(perf))
    dev_total, wall_total = shard.reduce_times(dist, [dev_total, wall_total], device=None)
    total_utt = sum(shard.gather_counts(dist, n_utt, device=None))
    audio_s = steps * total_utt * frames * FRAME_S
    hbm, which = peaks()
    nb0 = batches[0]["nb"]
    ms_frame = sum(loops) / (steps * frames * len(batches)) * 1e3
    pos_mid = 10 + frames // 2
    kv_one = cuda.qwen_tts_cuda_talker_step_bytes(ctx, pos_mid) - cuda.qwen_tts_cuda_talker_step_bytes(ctx, 0)   # fp32 KV read of ONE utterance
    step_bytes = cuda.qwen_tts_cuda_talker_step_bytes(ctx, pos_mid) + (nb0 - 1) * kv_one + cuda.qwen_tts_cuda_subtalker_frame_bytes(ctx)
    ach = step_bytes / (ms_frame * 1e-3) / 1e9
    pcm_bytes = sum(int(bt["pcm"].nbytes) for bt in batches) if with_pcm else 0
    return {"value": audio_s / dev_total, "e2e": audio_s / wall_total, "unit": "audio_s/s", "utterances": total_utt, "utterances_per_gpu": n_utt,
            "batches_per_gpu": len(batches), "frames": frames, "includes_pcm": bool(with_pcm), "ms_per_batch_frame": ms_frame,
            "ms_per_step": wall_total / steps * 1e3, "gpu_launches": int(perf.kernel_launches - l0),
            "h2d_bytes_per_step": sum(int(bt["toks"].nbytes + bt["offs"].nbytes) for bt in batches),
            "d2h_bytes_per_step": sum(int(bt["codes"].nbytes) for bt in batches) + pcm_bytes,
            "roofline": {"bound": "hbm", "kernel": "one batched frame = talker step + 16 sub-talker steps: weights streamed once per step for the whole batch + every utterance's fp32 KV at the mid position",
                         "achieved": ach, "peak": hbm, "unit": "GB/s", "frac": ach / hbm, "traffic": None, "peak_source": which, "bytes_per_frame": step_bytes},
            "clocks": sampler.summary()}


def run_batch(args, cuda, ctx, spec, rank, world, local, dist, torch):
    """`--batch N`: BASELINE config 4 as its own line (N utterances per GPU in lock-step, PCM included)"""
    r = measure_batch(args, cuda, ctx, spec, rank, world, local, dist, torch, args.batch, args.frames, args.steps, args.warmup, with_pcm=not args.codes_only)
    if rank == 0:
        print(json.dumps({
            "metric": "audio-sec generated/sec, 1.7B-CustomVoice lock-step batch", "value": r["value"], "unit": "audio_s/s", "n_gpus": world,
            "steps": args.steps, "warmup": args.warmup, "ms_per_step": r["ms_per_step"], "higher_is_better": True, "scaling": "weak",
            "vs_baseline": None, "dtype": "bf16 x bf16 -> f32 (tcgen05), activations as %d bf16 piece(s)" % args.pieces, "data": "synthetic",
            "config": {"workload": f"{args.model} CustomVoice shapes, {args.batch} utterances per GPU in lock-step batches of <= 64, {args.frames} frames each, "
                                   "default sampling, ragged prompts, " + ("codes only" if args.codes_only else "codes + codec decode to PCM"),
                       "batch": args.batch, "pieces": args.pieces},
            "ms_per_frame": r["ms_per_batch_frame"], "e2e": {"value": r["e2e"], "unit": "audio_s/s", "h2d_bytes_per_step": r["h2d_bytes_per_step"],
                                                              "d2h_bytes_per_step": r["d2h_bytes_per_step"]},
            "gpu_launches": r["gpu_launches"], "roofline": r["roofline"], "clocks": r["clocks"]}))


def run_queue(args, cuda, ctx, spec, rank, world, local, dist, torch):
    """`--queue N`: N requests per GPU with RAGGED lengths (uniform in [frames / 5, frames], fixed per request) through 64 lock-step rows.
    Two arms over the same requests: continuous batching (qwen_tts_cuda_generate_queue: a row whose request ended takes the next one)
    and lock-step batches of 64 (qwen_tts_cuda_generate_batch runs every row of a batch to the call's frame limit; only the
    requested frames count as audio).  Codes + codec decode to PCM in both."""
    G, frames, n_utt = spec["groups"], args.frames, args.queue
    cuda.qwen_tts_cuda_generate_batch.argtypes = [C.c_void_p, C.c_int, B.i32p, B.i32p, B.i32p, B.i32p, B.i32p, B.i32p, B.f32p, C.c_int]
    cuda.qwen_tts_cuda_generate_queue.argtypes = [C.c_void_p, C.c_int, C.c_int, B.i32p, B.i32p, B.i32p, B.i32p, B.i32p, B.i32p, B.i32p, B.f32p, C.c_int]
    cuda.qwen_tts_cuda_set_batch_pieces.argtypes = [C.c_void_p, C.c_int]
    assert cuda.qwen_tts_cuda_set_batch_pieces(ctx, args.pieces) == 0
    cuda.qwen_tts_cuda_set_params(ctx, C.byref(B.default_params(fixed_codec_tokens=frames, sample_seed=1000 * rank + 42)))
    rng = np.random.default_rng(rank)
    prompts = []
    for _ in range(n_utt):
        extra = [int(x) for x in rng.integers(1000, 5000, size=int(rng.integers(0, 24)))]
        prompts.append(TOKENS[:4] + extra + TOKENS[4:])
    budgets = rng.integers(max(1, frames // 5), frames + 1, size=n_utt).astype(np.int32)
    toks = np.array(sum(prompts, []), dtype=np.int32)
    offs = np.array([0] + list(np.cumsum([len(q) for q in prompts])), dtype=np.int32)
    spk, lang = np.full(n_utt, mg.SPEAKERS["aiden"], dtype=np.int32), np.full(n_utt, mg.LANGUAGES["english"], dtype=np.int32)
    codes, nfr = np.zeros((n_utt, frames, G), dtype=np.int32), np.zeros(n_utt, dtype=np.int32)
    pcm = np.zeros((n_utt, frames * 1920), dtype=np.float32)
    perf = B.Perf()

    def queue_arm():
        rc = cuda.qwen_tts_cuda_generate_queue(ctx, n_utt, 64, B.iptr(toks), B.iptr(offs), B.iptr(spk), B.iptr(lang), B.iptr(budgets), B.iptr(codes), B.iptr(nfr), B.fptr(pcm), frames)
        if rc != 0 or not np.array_equal(nfr, budgets):
            raise RuntimeError("generate_queue failed: " + cuda.qwen_tts_cuda_last_error().decode())

    def lockstep_arm():
        for lo, hi in shard.micro_batches(0, n_utt, 64):
            sub = np.ascontiguousarray(toks[offs[lo]:offs[hi]]); so = np.ascontiguousarray(offs[lo:hi + 1] - offs[lo])
            rc = cuda.qwen_tts_cuda_generate_batch(ctx, hi - lo, B.iptr(sub), B.iptr(so), B.iptr(spk[lo:hi].copy()), B.iptr(lang[lo:hi].copy()), B.iptr(codes[lo:hi]),
                                                   B.iptr(nfr[lo:hi]), B.fptr(pcm[lo:hi]), frames)
            if rc != 0:
                raise RuntimeError("generate_batch failed: " + cuda.qwen_tts_cuda_last_error().decode())

    def timed(fn):
        for _ in range(args.warmup):
            fn()
        if dist is not None:
            dist.barrier()
        torch.cuda.synchronize()
        t0 = time.perf_counter()
        for _ in range(args.steps):
            fn()
        if dist is not None:
            dist.barrier()
        torch.cuda.synchronize()
        return shard.reduce_times(dist, [time.perf_counter() - t0], device=None)[0]

    sampler = ClockSampler(local)
    sampler.start()
    cuda.qwen_tts_cuda_get_perf(ctx, C.byref(perf))
    l0 = perf.kernel_launches
    t_queue = timed(queue_arm)
    cuda.qwen_tts_cuda_get_perf(ctx, C.byref(perf))
    launches, admitted = int(perf.kernel_launches - l0), int(perf.queue_admitted)
    t_lock = timed(lockstep_arm)
    sampler.stop_flag = True
    useful = float(sum(shard.gather_counts(dist, int(budgets.sum()), device=None))) * FRAME_S * args.steps
    if rank == 0:
        print(json.dumps({
            "metric": "audio-sec generated/sec, 1.7B-CustomVoice, queue of ragged-length requests through 64 lock-step rows", "value": useful / t_queue, "unit": "audio_s/s",
            "n_gpus": world, "steps": args.steps, "warmup": args.warmup, "ms_per_step": t_queue / args.steps * 1e3, "higher_is_better": True, "scaling": "weak",
            "vs_baseline": None, "dtype": "bf16 x bf16 -> f32 (tcgen05), activations as %d bf16 piece(s)" % args.pieces, "data": "synthetic",
            "config": {"workload": f"{args.model} CustomVoice shapes, {n_utt} requests per GPU, lengths uniform in [{max(1, frames // 5)}, {frames}] frames (mean {float(budgets.mean()):.1f}), "
                                   "continuous batching over 64 rows, default sampling, ragged prompts, codes + codec decode to PCM; host wall clock",
                       "requests": n_utt, "rows": 64, "pieces": args.pieces, "admitted": admitted},
            "e2e": {"value": useful / t_queue, "unit": "audio_s/s", "h2d_bytes_per_step": int(toks.nbytes + offs.nbytes + budgets.nbytes), "d2h_bytes_per_step": int(codes.nbytes + pcm.nbytes)},
            "lockstep_batches": {"value": useful / t_lock, "unit": "audio_s/s", "ms_per_step": t_lock / args.steps * 1e3,
                                 "note": "the same requests as lock-step batches of 64, every row run to the frame limit; only the requested frames count"},
            "gpu_launches": launches, "clocks": sampler.summary()}))


def run_codec(args, cuda, ctx, spec, hm, rank, world, local, dist, torch):
    """BASELINE config 3: codec decoder only, 16 codebooks x T frames -> 24 kHz PCM (qwen_tts_codec_decode, c/qwen_tts_codec.c:581-749).
    value: audio-seconds per second, codes resident in HBM .. last codec kernel (CUDA events); e2e: host codes in, host PCM out."""
    T, G = args.codec_frames, spec["groups"]
    rng = np.random.default_rng(rank)
    codes = rng.integers(0, 2048, size=(T, G), dtype=np.int32)
    perf = B.Perf()

    def one():
        ns = C.c_int(0)
        t0 = time.perf_counter()
        ptr = cuda.qwen_tts_cuda_codec_decode(ctx, B.iptr(codes), T, C.byref(ns))
        wall = time.perf_counter() - t0
        if not ptr or ns.value != T * 1920:
            raise RuntimeError("codec_decode failed: " + cuda.qwen_tts_cuda_last_error().decode())
        cuda.qwen_tts_cuda_free_host(ptr)
        cuda.qwen_tts_cuda_get_perf(ctx, C.byref(perf))
        return wall, perf.device_ms / 1e3

    for _ in range(args.warmup):
        one()
    cuda.qwen_tts_cuda_get_perf(ctx, C.byref(perf))
    l0 = perf.kernel_launches
    sampler = ClockSampler(local)
    sampler.start()
    if dist is not None:
        dist.barrier()
    torch.cuda.synchronize()
    t0 = time.perf_counter()
    devs = [one()[1] for _ in range(args.steps)]
    if dist is not None:
        dist.barrier()
    torch.cuda.synchronize()
    wall_total, dev_total = time.perf_counter() - t0, sum(devs)
    sampler.stop_flag = True
    cuda.qwen_tts_cuda_get_perf(ctx, C.byref(perf))
    dev_total, wall_total = shard.reduce_times(dist, [dev_total, wall_total], device=None)
    if rank == 0:
        audio_s = world * args.steps * T * FRAME_S
        flops = 5.12e9 * T                        # 18.44 TFLOP at T = 3600, 93 % of it in the vocoder convolutions
        ach = flops / (dev_total / args.steps) / 1e12
        tc = os.environ.get("QTC_CODEC_TC", "1") != "0"
        pk = json.load(open(os.path.join(ROOT, "MEASURED_PEAKS.json"))) if os.path.exists(os.path.join(ROOT, "MEASURED_PEAKS.json")) else {}
        bf16_peak = pk.get("bf16_tflops_sustained", 1377.0)   # a decode is hundreds of back-to-back launches: the sustained figure
        fp32_peak = 148 * 128 * 2 * 1.965e9 / 1e12   # FFMA lanes x 2 flop x boost clock: the pipe k_conv_tile / k_sgemm_nt run on
        line = {"metric": "audio-sec decoded/sec, codec decoder only (16 codebooks x T frames -> 24 kHz PCM)", "value": audio_s / dev_total, "unit": "audio_s/s",
                "n_gpus": world, "steps": args.steps, "warmup": args.warmup, "ms_per_step": wall_total / args.steps * 1e3, "higher_is_better": True,
                "scaling": "weak", "vs_baseline": None, "dtype": "bf16 x2 pieces -> f32 (tcgen05)" if tc else "f32", "data": "synthetic",
                "config": {"workload": f"Qwen3-TTS-Tokenizer-12Hz decoder shapes, {T} frames ({T * FRAME_S:.0f} s audio), codes ~ U[0,2048), random-init f32 weights",
                           "l2": "activations (up to 2.65 GB per layer at T = 3600) exceed the 126 MB L2"},
                "e2e": {"value": audio_s / wall_total, "unit": "audio_s/s", "h2d_bytes_per_step": int(codes.nbytes), "d2h_bytes_per_step": T * 1920 * 4},
                "gpu_launches": int(perf.kernel_launches - l0),
                "roofline": ({"bound": "tensor", "kernel": "k_conv_tc (92 % of the decode: vocoder convolutions, codec transformer and ConvNeXt projections as tcgen05 implicit GEMMs)",
                              "achieved": 3 * ach, "peak": bf16_peak, "unit": "TFLOP/s", "frac": 3 * ach / bf16_peak, "traffic": None,
                              "peak_source": "MEASURED_PEAKS.json bf16_tflops_sustained" if pk else "fallback 1377",
                              "flop_per_step": flops, "algorithmic_tflops": ach,
                              "note": "every fp32 product is three bf16 MMAs (hi.hi + hi.lo + lo.hi): achieved = 3 x algorithmic FLOP/s of the whole decode "
                                      "(5.12 GFLOP per frame, SURVEY 8d) over the event-timed decode"} if tc else
                             {"bound": "fp32-fma (QTC_CODEC_TC=0: direct convolutions on the FFMA pipe)", "achieved": ach, "peak": fp32_peak, "unit": "TFLOP/s",
                              "frac": ach / fp32_peak, "traffic": None, "peak_source": "nominal 148 SM x 128 FFMA x 1.965 GHz", "flop_per_step": flops}),
                "clocks": sampler.summary()}
        if not args.no_cpu and world == 1:
            Tc = 4
            orc = load_checker().load_oracle()
            s = orc.orc_create(hm.desc)
            ns = C.c_int(0)
            t1 = time.perf_counter()
            ptr = orc.orc_codec_decode(s, B.iptr(codes[:Tc].copy()), Tc, C.byref(ns))
            dt = time.perf_counter() - t1
            orc.orc_free(ptr)
            orc.orc_destroy(s)
            line["cpu_baseline"] = {"value": Tc * FRAME_S / dt, "unit": "audio_s/s", "cores": os.cpu_count(), "kind": "port",
                                    "sample": f"codec decode of {Tc} frames, {dt:.1f} s, oracle port with OpenMP over output channels"}
        print(json.dumps(line))


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--gpus", type=int, default=1)
    ap.add_argument("--steps", type=int, default=5)
    ap.add_argument("--warmup", type=int, default=3)
    ap.add_argument("--impl", default="b200", choices=["b200", "reference"])
    ap.add_argument("--model", default="1.7b", choices=list(mg.PRESETS))
    ap.add_argument("--frames", type=int, default=FRAMES)
    ap.add_argument("--mode", type=int, default=2, help="2: cluster kernel, two grid-wide exchanges per layer (default); 1: 148-CTA fused kernel, five; 0: one kernel per op")
    ap.add_argument("--strict", type=int, default=0)
    ap.add_argument("--no-cpu", action="store_true", help="skip the cpu_baseline sample")
    ap.add_argument("--batch", type=int, default=0, help="> 0: lock-step batched decode of this many utterances per GPU (configs 4-5), codes only")
    ap.add_argument("--queue", type=int, default=0, help="> 0: this many ragged-length requests per GPU through 64 rows: continuous batching vs lock-step batches")
    ap.add_argument("--codes-only", action="store_true", help="with --batch: skip the codec decode (round-1 behaviour)")
    ap.add_argument("--no-batched", action="store_true", help="default line: skip the config 4 / 5 sub-measurement")
    ap.add_argument("--greedy", action="store_true", help="greedy decoding (BASELINE config 1: --model 0.6b --frames 74 --greedy)")
    ap.add_argument("--codec-frames", type=int, default=0, help="> 0: codec decoder only on this many frames (BASELINE config 3: 3600)")
    ap.add_argument("--stream-first", type=int, default=8, help="frames in the first PCM chunk of the streamed-latency utterances")
    ap.add_argument("--stream-chunk", type=int, default=160, help="frames in later PCM chunks")
    ap.add_argument("--pieces", type=int, default=3, help="batched GEMM activations: 3 = exact bf16 split, 1 = bf16")
    args = ap.parse_args()
    rank, world = int(os.environ.get("RANK", 0)), int(os.environ.get("WORLD_SIZE", 1))
    local = int(os.environ.get("LOCAL_RANK", 0))
    if args.impl == "reference":
        # torchrun exports OMP_NUM_THREADS=1 to every rank; the reference arm is rank 0 alone and is entitled to every host
        # thread (its convolutions are OpenMP loops), so undo that before libgomp is loaded
        os.environ["OMP_NUM_THREADS"] = str(os.cpu_count() or 1)
        run_reference(args, rank, world)
        return

    import torch
    dist = None
    if world > 1:
        import torch.distributed as dist_
        dist = dist_
        os.environ.setdefault("MASTER_ADDR", "127.0.0.1")
        # utterances are independent and the models fit one GPU: nothing of the data path crosses GPUs (north_star: "no NCCL").
        # The process group carries the barrier and the max-over-ranks of the device time only: gloo on the host.
        dist.init_process_group("gloo")
    torch.cuda.set_device(local)
    cuda = B.load_cuda()                       # raises if the extension was not built: there is no fallback
    if cuda.qwen_tts_cuda_init(local) != 0:
        raise RuntimeError(cuda.qwen_tts_cuda_last_error().decode())
    spec, main_t, cod, hm = build_model(args.model)
    default_line = args.batch == 0 and args.codec_frames == 0 and args.queue == 0
    want_batched = default_line and not args.no_batched and args.model == "1.7b" and not args.greedy and args.frames == FRAMES
    ctx = cuda.qwen_tts_cuda_upload(C.byref(hm.desc), 64 if (want_batched or args.batch > 0 or args.queue > 0) else 1)     # max_batch sizes the paged KV pool
    if not ctx:
        raise RuntimeError(cuda.qwen_tts_cuda_last_error().decode())
    cuda.qwen_tts_cuda_set_mode(ctx, args.mode)
    cuda.qwen_tts_cuda_set_arith(ctx, args.strict)
    frames = args.frames
    if args.codec_frames > 0:
        run_codec(args, cuda, ctx, spec, hm, rank, world, local, dist, torch)
        cuda.qwen_tts_cuda_release(ctx)
        if dist is not None:
            dist.destroy_process_group()
        return
    if args.queue > 0 or args.batch > 0:
        (run_queue if args.queue > 0 else run_batch)(args, cuda, ctx, spec, rank, world, local, dist, torch)
        cuda.qwen_tts_cuda_release(ctx)
        if dist is not None:
            dist.destroy_process_group()
        return
    p = (B.greedy_params if args.greedy else B.default_params)(fixed_codec_tokens=frames, sample_seed=42 + rank)     # every rank: its own utterance
    cuda.qwen_tts_cuda_set_params(ctx, C.byref(p))
    G = spec["groups"]
    codes = np.zeros((frames, G), dtype=np.int32)
    perf = B.Perf()

    lat = []

    def one_utterance(record=True):
        ns, nf = C.c_int(0), C.c_int(0)
        t0 = time.perf_counter()
        ptr = cuda.qwen_tts_cuda_generate(ctx, CSV, b"aiden", b"english", C.byref(ns), B.iptr(codes), frames, C.byref(nf))
        wall = time.perf_counter() - t0
        if not ptr or nf.value != frames or ns.value != frames * 1920:
            raise RuntimeError("generate failed: " + cuda.qwen_tts_cuda_last_error().decode())
        cuda.qwen_tts_cuda_free_host(ptr)
        cuda.qwen_tts_cuda_get_perf(ctx, C.byref(perf))
        if record:
            lat.append((perf.ttff_ms, perf.ttfa_ms, perf.frame_ms_p50, perf.frame_ms_p99, wall, perf.prefill_ms))
        return wall, perf.device_ms / 1e3, perf.talker_ms / 1e3, ns.value

    def barrier():
        if dist is not None:
            dist.barrier()
        torch.cuda.synchronize()

    # streaming latency (BASELINE config 2 is "streaming latency"): a few utterances with the chunked decode on (first PCM chunk
    # after 8 frames, then 160-frame chunks), measured before the throughput region, which runs in the reference's
    # decode-at-the-end order
    cuda.qwen_tts_cuda_set_streaming(ctx, args.stream_first, args.stream_chunk)
    for _ in range(3):
        one_utterance()
    cuda.qwen_tts_cuda_set_streaming(ctx, 0, 0)
    for _ in range(args.warmup):
        one_utterance(record=False)
    cuda.qwen_tts_cuda_get_perf(ctx, C.byref(perf))
    launches0 = perf.kernel_launches
    sampler = ClockSampler(local)
    sampler.start()
    barrier()
    t_all0 = time.perf_counter()
    walls, devs, loops = [], [], []
    pcm_bytes = 0
    for _ in range(args.steps):
        w, dv, lp, ns = one_utterance(record=False)
        walls.append(w); devs.append(dv); loops.append(lp)
        pcm_bytes = ns * 4
    barrier()
    t_all = time.perf_counter() - t_all0
    sampler.stop_flag = True
    cuda.qwen_tts_cuda_get_perf(ctx, C.byref(perf))
    launches = perf.kernel_launches - launches0

    dev_total, wall_total = sum(devs), t_all
    dev_total, wall_total = shard.reduce_times(dist, [dev_total, wall_total], device=None)   # slowest rank
    audio_s = world * args.steps * frames * FRAME_S
    value, e2e = audio_s / dev_total, audio_s / wall_total

    # BASELINE configs 4 / 5 inside the driver-run line: 1 GPU -> config 4 (64 utterances x 250 frames in lock-step on this GPU);
    # N > 1 -> config 5 (256 utterances split contiguously over the N GPUs, harness/shard.py).  Codes + PCM, device-timed like `value`.
    batched = None
    if want_batched:
        per_gpu = 64 if world == 1 else (shard.partition(256, world)[rank][1] - shard.partition(256, world)[rank][0])
        r = measure_batch(args, cuda, ctx, spec, rank, world, local, dist, torch, per_gpu, 250, 1, 1, with_pcm=True)
        if rank == 0:
            batched = {"config": ("BASELINE config 4: 64 utterances per GPU, 20 s each (250 frames), lock-step, codes + PCM" if world == 1 else
                                  f"BASELINE config 5: 256 utterances over {world} GPUs ({per_gpu} per GPU in lock-step batches of <= 64), 250 frames each, codes + PCM"),
                       **{k: r[k] for k in ("value", "e2e", "unit", "utterances", "utterances_per_gpu", "batches_per_gpu", "frames", "includes_pcm",
                                            "ms_per_batch_frame", "gpu_launches", "roofline")}}
        p = (B.greedy_params if args.greedy else B.default_params)(fixed_codec_tokens=frames, sample_seed=42 + rank)
        cuda.qwen_tts_cuda_set_params(ctx, C.byref(p))
    line = None
    if rank == 0:
        hbm, which = peaks()
        pos = 10 + frames // 2
        step_ms = cuda.qwen_tts_cuda_bench_talker_step(ctx, pos, 20, 1)
        sub_ms = cuda.qwen_tts_cuda_bench_subtalker(ctx, 10, 1)
        step_bytes = cuda.qwen_tts_cuda_talker_step_bytes(ctx, pos)
        sub_bytes = cuda.qwen_tts_cuda_subtalker_frame_bytes(ctx)
        ach = step_bytes / (step_ms * 1e-3) / 1e9
        metric = METRIC if (args.model == "1.7b" and frames == FRAMES and not args.greedy) else \
            f"audio-sec generated/sec (RTF), {args.model}-CustomVoice batch 1, {frames} frames{', greedy' if args.greedy else ''}"
        line = {"metric": metric, "value": value, "unit": "audio_s/s", "n_gpus": world, "steps": args.steps, "warmup": args.warmup,
                "ms_per_step": wall_total / args.steps * 1e3, "higher_is_better": True, "scaling": "weak", "vs_baseline": None,
                "dtype": "f32 (bf16 weights)", "data": "synthetic",
                "config": {"workload": f"{args.model} CustomVoice shapes, batch 1 per GPU, {frames} frames ({frames * FRAME_S:.1f} s audio), " +
                                       ("greedy, " if args.greedy else "default sampling (temp 0.9, top-k 50, rep-pen 1.05), ") + "random-init weights, fixed frame budget",
                           "mode": args.mode, "arith": "strict" if args.strict else "fast",
                           "l2": "weights per step (2.8 GB talker, 157 MB x16 sub-talker) exceed the 126 MB L2; the roofline probe flushes L2 between iterations"},
                "ms_per_frame": sum(loops) / (args.steps * frames) * 1e3,
                # streaming latency (config 2): device clock from the start of the prompt build, median over the timed utterances;
                # see "note"
                "latency": {"ttff_ms": float(np.median([x[0] for x in lat[1:]])), "ttfa_ms": float(np.median([x[1] for x in lat[1:]])),
                            "frame_ms_p50": float(np.median([x[2] for x in lat[1:]])), "frame_ms_p99": float(np.median([x[3] for x in lat[1:]])),
                            "prefill_ms": float(np.median([x[5] for x in lat[1:]])), "first_chunk_frames": args.stream_first, "chunk_frames": args.stream_chunk,
                            "e2e_streamed": frames * FRAME_S / float(np.median([x[4] for x in lat[1:]])),
                            "note": "qwen_tts_cuda_set_streaming on: PCM chunks decoded between frames and copied to the host while the loop goes on; "
                                    "value / e2e of this line are measured with it off (decode at the end, the reference's order)"},
                "e2e": {"value": e2e, "unit": "audio_s/s", "h2d_bytes_per_step": len(CSV), "d2h_bytes_per_step": pcm_bytes + frames * G * 4},
                "gpu_launches": int(launches),
                "roofline": {"bound": "hbm", "kernel": "talker decode step (all kernels of one step)", "achieved": ach, "peak": hbm, "unit": "GB/s",
                             "frac": ach / hbm, "traffic": ncu_traffic(args.mode) if args.model == "1.7b" else None, "peak_source": which, "bytes_per_step": step_bytes, "ms_per_talker_step": step_ms,
                             "subtalker": {"ms_per_frame": sub_ms, "streamed_bytes": sub_bytes, "achieved": sub_bytes / (sub_ms * 1e-3) / 1e9,
                                           "frac": sub_bytes / (sub_ms * 1e-3) / 1e9 / hbm}},
                "clocks": sampler.summary()}
        if batched is not None:
            line["batched"] = batched
        cuda.qwen_tts_cuda_get_perf(ctx, C.byref(perf))
        line["upload"] = {"ms": perf.upload_ms, "gb": perf.upload_bytes / 1e9, "gb_per_s": perf.upload_bytes / 1e6 / max(perf.upload_ms, 1e-9),
                          "note": "qwen_tts_cuda_upload: model from pageable host memory through pinned staging, incl. device-side re-layouts (untimed setup)"}
        if not args.no_cpu and world == 1:
            rtf_cpu, dt = cpu_baseline_sample(hm, spec, main_t, frames=1)
            line["cpu_baseline"] = {"value": rtf_cpu, "unit": "audio_s/s", "cores": os.cpu_count(), "kind": "port",
                                    "sample": f"1 frame (sub-talker + talker step) + codec decode of it after an untimed prefill, {dt:.1f} s; "
                                              "oracle port, OpenMP over output rows (the reference's own GEMVs are single-threaded: see --impl reference)"}
        print(json.dumps(line))
    cuda.qwen_tts_cuda_release(ctx)
    if dist is not None:
        dist.destroy_process_group()


if __name__ == "__main__":
    main()
