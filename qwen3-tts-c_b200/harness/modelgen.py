"""Random-init model writer for the named Qwen3-TTS-12Hz architectures.

No checkpoints exist offline, so tests and bench.py use seeded random weights of
the named shapes (BASELINE.json north_star).  Tensor names are
exactly those the reference loader binds (c/qwen_tts.c:440-766, SURVEY.md
Appendix B) so the same tensors can be (a) written as safetensors + config.json
for the unmodified reference to load and (b) handed to the oracle / the CUDA
library as a qtc_model_desc_t.

bf16 tensors are drawn through a 65536-entry inverse-CDF table (exactly Gaussian
at 16-bit resolution, ~10x faster than standard_normal for the 2.3 G parameters
of the 1.7B model).
"""
from __future__ import annotations

import json
import os
import struct

import numpy as np

TTS_PAD, TTS_BOS, TTS_EOS = 151671, 151672, 151673   # c/qwen_tts.h:67-69

PRESETS = {
    # tiny: every code path (input_proj because sub.hidden != talker.hidden, GQA,
    # windowed codec attention with T > window, odd vocoder channel counts)
    "tiny": dict(
        talker=dict(vocab=3072, hidden=256, inter=512, layers=2, heads=4, kv=2, hd=128),
        text_hidden=64, text_vocab=151936, groups=16, mrope=[24, 20, 20], theta=1000000.0,
        sub=dict(vocab=2048, hidden=128, inter=256, layers=2, heads=2, kv=1, hd=128),
        codec=dict(nq=16, cb_size=2048, cb_dim=64, hidden=128, latent=128, layers=2, heads=2,
                   kv=2, inter=256, window=8, dec=96, rates=[8, 5, 4, 3], ratios=[2, 2]),
    ),
    # codec at full size but a small LM: for codec-only tests
    "0.6b": dict(
        talker=dict(vocab=3072, hidden=1024, inter=3072, layers=28, heads=16, kv=8, hd=128),
        text_hidden=2048, text_vocab=151936, groups=16, mrope=[24, 20, 20], theta=1000000.0,
        sub=dict(vocab=2048, hidden=1024, inter=3072, layers=5, heads=16, kv=8, hd=128),
        codec=dict(nq=16, cb_size=2048, cb_dim=512, hidden=1024, latent=1024, layers=8, heads=16,
                   kv=16, inter=3072, window=72, dec=1536, rates=[8, 5, 4, 3], ratios=[2, 2]),
    ),
    "1.7b": dict(
        talker=dict(vocab=3072, hidden=2048, inter=6144, layers=28, heads=16, kv=8, hd=128),
        text_hidden=2048, text_vocab=151936, groups=16, mrope=[24, 20, 20], theta=1000000.0,
        sub=dict(vocab=2048, hidden=1024, inter=3072, layers=5, heads=16, kv=8, hd=128),
        codec=dict(nq=16, cb_size=2048, cb_dim=512, hidden=1024, latent=1024, layers=8, heads=16,
                   kv=16, inter=3072, window=72, dec=1536, rates=[8, 5, 4, 3], ratios=[2, 2]),
    ),
}
# a mid-size LM (real head/GQA geometry, few layers) for GPU parity runs the oracle finishes in seconds
PRESETS["small"] = dict(
    talker=dict(vocab=3072, hidden=1024, inter=3072, layers=4, heads=16, kv=8, hd=128),
    text_hidden=512, text_vocab=151936, groups=16, mrope=[24, 20, 20], theta=1000000.0,
    sub=dict(vocab=2048, hidden=1024, inter=3072, layers=2, heads=16, kv=8, hd=128),
    codec=PRESETS["tiny"]["codec"],
)
# one 1.7B-shaped layer stack cut to 4 + 2 layers: the row split over 148 CTAs at H = 2048 / I = 6144 and long KV ranges
# at a cost the oracle pays in seconds per hundred steps
PRESETS["1.7b-4L"] = dict(
    talker=dict(vocab=3072, hidden=2048, inter=6144, layers=4, heads=16, kv=8, hd=128),
    text_hidden=2048, text_vocab=151936, groups=16, mrope=[24, 20, 20], theta=1000000.0,
    sub=dict(vocab=2048, hidden=1024, inter=3072, layers=2, heads=16, kv=8, hd=128),
    codec=PRESETS["tiny"]["codec"],
)
SPEAKERS = {"aiden": 3010, "vivian": 3011}
LANGUAGES = {"english": 2050, "chinese": 2055}
CODEC_IDS = dict(codec_pad_id=2148, codec_bos_id=2149, codec_eos_token_id=2150, codec_think_id=2154,
                 codec_nothink_id=2155, codec_think_bos_id=2156, codec_think_eos_id=2157)

_LUT = None


def _normal_lut():
    global _LUT
    if _LUT is None:
        from scipy.special import ndtri
        q = (np.arange(65536, dtype=np.float64) + 0.5) / 65536.0
        _LUT = ndtri(q).astype(np.float32)
    return _LUT


def f32_to_bf16(x: np.ndarray) -> np.ndarray:
    """round-to-nearest-even to the upper 16 bits"""
    u = np.ascontiguousarray(x, dtype=np.float32).view(np.uint32)
    r = ((u >> 16) & 1) + 0x7FFF
    return ((u + r) >> 16).astype(np.uint16)


def bf16_to_f32(h: np.ndarray) -> np.ndarray:
    return (h.astype(np.uint32) << 16).view(np.float32)


class _Gen:
    def __init__(self, seed):
        self.rng = np.random.Generator(np.random.PCG64(seed))

    def bf16(self, shape, std):
        n = int(np.prod(shape))
        idx = self.rng.integers(0, 65536, size=n, dtype=np.uint16)
        lut = f32_to_bf16(_normal_lut() * np.float32(std))
        return lut[idx].reshape(shape)

    def f32(self, shape, std, mean=0.0):
        return (self.rng.standard_normal(size=shape, dtype=np.float32) * np.float32(std) + np.float32(mean))

    def uniform(self, shape, lo, hi):
        return self.rng.uniform(lo, hi, size=shape).astype(np.float32)


def build_tensors(spec: dict, seed: int = 0, with_codec: bool = True, with_lm: bool = True):
    """returns (main_tensors, codec_tensors): name -> ndarray (uint16 = bf16, float32 = f32)"""
    g = _Gen(seed)
    T, S, C = spec["talker"], spec["sub"], spec["codec"]
    main, cod = {}, {}
    H = T["hidden"]
    if with_lm:
        main["talker.model.codec_embedding.weight"] = g.bf16((T["vocab"], H), 0.3)
        main["talker.model.text_embedding.weight"] = g.bf16((spec["text_vocab"], spec["text_hidden"]), 0.02)
        th = spec["text_hidden"]
        main["talker.text_projection.linear_fc1.weight"] = g.bf16((th, th), 1.0 / np.sqrt(th))
        main["talker.text_projection.linear_fc1.bias"] = g.f32((th,), 0.02)
        main["talker.text_projection.linear_fc2.weight"] = g.bf16((H, th), 0.5 / np.sqrt(th))
        main["talker.text_projection.linear_fc2.bias"] = g.f32((H,), 0.02)

        def lm(prefix, L):
            h, qd, kvd, I = L["hidden"], L["heads"] * L["hd"], L["kv"] * L["hd"], L["inter"]
            for i in range(L["layers"]):
                p = f"{prefix}.layers.{i}."
                main[p + "self_attn.q_proj.weight"] = g.bf16((qd, h), 1.0 / np.sqrt(h))
                main[p + "self_attn.k_proj.weight"] = g.bf16((kvd, h), 1.0 / np.sqrt(h))
                main[p + "self_attn.v_proj.weight"] = g.bf16((kvd, h), 1.0 / np.sqrt(h))
                main[p + "self_attn.o_proj.weight"] = g.bf16((h, qd), 0.5 / np.sqrt(qd))
                main[p + "self_attn.q_norm.weight"] = g.f32((L["hd"],), 0.1, 1.0)
                main[p + "self_attn.k_norm.weight"] = g.f32((L["hd"],), 0.1, 1.0)
                main[p + "input_layernorm.weight"] = g.f32((h,), 0.1, 1.0)
                main[p + "post_attention_layernorm.weight"] = g.f32((h,), 0.1, 1.0)
                main[p + "mlp.gate_proj.weight"] = g.bf16((I, h), 1.0 / np.sqrt(h))
                main[p + "mlp.up_proj.weight"] = g.bf16((I, h), 1.0 / np.sqrt(h))
                main[p + "mlp.down_proj.weight"] = g.bf16((h, I), 0.5 / np.sqrt(I))
            main[f"{prefix}.norm.weight"] = g.f32((h,), 0.1, 1.0)

        lm("talker.model", T)
        # heads scaled so logits have std ~4: argmax decisions are not near-ties
        main["talker.codec_head.weight"] = g.bf16((T["vocab"], H), 4.0 / np.sqrt(H))
        lm("talker.code_predictor.model", S)
        for k in range(spec["groups"] - 1):
            main[f"talker.code_predictor.model.codec_embedding.{k}.weight"] = g.bf16((S["vocab"], H), 0.3)
            main[f"talker.code_predictor.lm_head.{k}.weight"] = g.bf16((S["vocab"], S["hidden"]),
                                                                        4.0 / np.sqrt(S["hidden"]))
        if S["hidden"] != H:
            main["talker.code_predictor.small_to_mtp_projection.weight"] = g.bf16((S["hidden"], H), 1.0 / np.sqrt(H))
            main["talker.code_predictor.small_to_mtp_projection.bias"] = g.f32((S["hidden"],), 0.02)
    if with_codec:
        vq, half, lat, hc = C["cb_dim"] // 2, C["latent"] // 2, C["latent"], C["hidden"]
        for q in range(C["nq"]):
            p = ("decoder.quantizer.rvq_first.vq.layers.0._codebook." if q == 0
                 else f"decoder.quantizer.rvq_rest.vq.layers.{q - 1}._codebook.")
            cod[p + "cluster_usage"] = g.uniform((C["cb_size"],), 0.5, 1.5)
            cod[p + "embedding_sum"] = g.f32((C["cb_size"], vq), 1.0)
        cod["decoder.quantizer.rvq_first.output_proj.weight"] = g.f32((half, vq, 1), 1.0 / np.sqrt(vq))
        cod["decoder.quantizer.rvq_rest.output_proj.weight"] = g.f32((half, vq, 1), 0.25 / np.sqrt(vq))
        cod["decoder.pre_conv.conv.weight"] = g.f32((lat, C["cb_dim"], 3), 1.0 / np.sqrt(C["cb_dim"] * 3))
        cod["decoder.pre_conv.conv.bias"] = g.f32((lat,), 0.02)
        pt = "decoder.pre_transformer."
        cod[pt + "input_proj.weight"] = g.f32((hc, lat), 1.0 / np.sqrt(lat))
        cod[pt + "input_proj.bias"] = g.f32((hc,), 0.02)
        cod[pt + "output_proj.weight"] = g.f32((lat, hc), 1.0 / np.sqrt(hc))
        cod[pt + "output_proj.bias"] = g.f32((lat,), 0.02)
        cod[pt + "norm.weight"] = g.f32((hc,), 0.1, 1.0)
        hd = hc // C["heads"]
        for i in range(C["layers"]):
            p = f"{pt}layers.{i}."
            cod[p + "input_layernorm.weight"] = g.f32((hc,), 0.1, 1.0)
            cod[p + "post_attention_layernorm.weight"] = g.f32((hc,), 0.1, 1.0)
            cod[p + "self_attn_layer_scale.scale"] = g.f32((hc,), 0.002, 0.01)
            cod[p + "mlp_layer_scale.scale"] = g.f32((hc,), 0.002, 0.01)
            cod[p + "self_attn.q_proj.weight"] = g.f32((C["heads"] * hd, hc), 1.0 / np.sqrt(hc))
            cod[p + "self_attn.k_proj.weight"] = g.f32((C["kv"] * hd, hc), 1.0 / np.sqrt(hc))
            cod[p + "self_attn.v_proj.weight"] = g.f32((C["kv"] * hd, hc), 1.0 / np.sqrt(hc))
            cod[p + "self_attn.o_proj.weight"] = g.f32((hc, C["heads"] * hd), 1.0 / np.sqrt(hc))
            cod[p + "mlp.gate_proj.weight"] = g.f32((C["inter"], hc), 1.0 / np.sqrt(hc))
            cod[p + "mlp.up_proj.weight"] = g.f32((C["inter"], hc), 1.0 / np.sqrt(hc))
            cod[p + "mlp.down_proj.weight"] = g.f32((hc, C["inter"]), 1.0 / np.sqrt(C["inter"]))
        for s in range(2):
            r = C["ratios"][s]
            cod[f"decoder.upsample.{s}.0.conv.weight"] = g.f32((lat, lat, r), 1.0 / np.sqrt(lat))
            cod[f"decoder.upsample.{s}.0.conv.bias"] = g.f32((lat,), 0.02)
            p = f"decoder.upsample.{s}.1."
            cod[p + "dwconv.conv.weight"] = g.f32((lat, 1, 7), 1.0 / np.sqrt(7))
            cod[p + "dwconv.conv.bias"] = g.f32((lat,), 0.02)
            cod[p + "norm.weight"] = g.f32((lat,), 0.1, 1.0)
            cod[p + "norm.bias"] = g.f32((lat,), 0.02)
            cod[p + "pwconv1.weight"] = g.f32((4 * lat, lat), 1.0 / np.sqrt(lat))
            cod[p + "pwconv1.bias"] = g.f32((4 * lat,), 0.02)
            cod[p + "pwconv2.weight"] = g.f32((lat, 4 * lat), 1.0 / np.sqrt(4 * lat))
            cod[p + "pwconv2.bias"] = g.f32((lat,), 0.02)
            cod[p + "gamma"] = g.f32((lat,), 0.002, 0.01)
        dec = C["dec"]
        cod["decoder.decoder.0.conv.weight"] = g.f32((dec, lat, 7), 1.0 / np.sqrt(lat * 7))
        cod["decoder.decoder.0.conv.bias"] = g.f32((dec,), 0.02)
        dim = dec
        for b in range(4):
            r, od = C["rates"][b], dim // 2
            p = f"decoder.decoder.{b + 1}.block."
            cod[p + "0.alpha"] = g.f32((dim,), 0.2)
            cod[p + "0.beta"] = g.f32((dim,), 0.2)
            cod[p + "1.conv.weight"] = g.f32((dim, od, 2 * r), 1.0 / np.sqrt(dim * 2))
            cod[p + "1.conv.bias"] = g.f32((od,), 0.02)
            for u in range(3):
                q = f"{p}{u + 2}."
                cod[q + "act1.alpha"] = g.f32((od,), 0.2)
                cod[q + "act1.beta"] = g.f32((od,), 0.2)
                cod[q + "conv1.conv.weight"] = g.f32((od, od, 7), 0.7 / np.sqrt(od * 7))
                cod[q + "conv1.conv.bias"] = g.f32((od,), 0.02)
                cod[q + "act2.alpha"] = g.f32((od,), 0.2)
                cod[q + "act2.beta"] = g.f32((od,), 0.2)
                cod[q + "conv2.conv.weight"] = g.f32((od, od, 1), 0.5 / np.sqrt(od))
                cod[q + "conv2.conv.bias"] = g.f32((od,), 0.02)
            dim = od
        cod["decoder.decoder.5.alpha"] = g.f32((dim,), 0.2)
        cod["decoder.decoder.5.beta"] = g.f32((dim,), 0.2)
        cod["decoder.decoder.6.conv.weight"] = g.f32((1, dim, 7), 0.04 / np.sqrt(dim * 7))
        cod["decoder.decoder.6.conv.bias"] = g.f32((1,), 0.01)
    return main, cod


def _write_safetensors(path, tensors):
    header, off = {}, 0
    for name, a in tensors.items():
        n = a.nbytes
        header[name] = {"dtype": "BF16" if a.dtype == np.uint16 else "F32",
                        "shape": list(a.shape), "data_offsets": [off, off + n]}
        off += n
    hj = json.dumps(header, separators=(",", ":")).encode()
    hj += b" " * ((8 - len(hj) % 8) % 8)
    with open(path, "wb") as f:
        f.write(struct.pack("<Q", len(hj)))
        f.write(hj)
        for a in tensors.values():
            f.write(np.ascontiguousarray(a).tobytes())


def configs(spec):
    T, S, C = spec["talker"], spec["sub"], spec["codec"]
    main = {"talker_config": dict(
        vocab_size=T["vocab"], hidden_size=T["hidden"], intermediate_size=T["inter"],
        num_hidden_layers=T["layers"], num_attention_heads=T["heads"], num_key_value_heads=T["kv"],
        head_dim=T["hd"], text_hidden_size=spec["text_hidden"], text_vocab_size=spec["text_vocab"],
        num_code_groups=spec["groups"], rms_norm_eps=1e-6, rope_theta=spec["theta"],
        rope_scaling={"mrope_section": spec["mrope"]},
        code_predictor_config=dict(vocab_size=S["vocab"], hidden_size=S["hidden"], intermediate_size=S["inter"],
                                   num_hidden_layers=S["layers"], num_attention_heads=S["heads"],
                                   num_key_value_heads=S["kv"], head_dim=S["hd"]),
        spk_id=SPEAKERS, codec_language_id=LANGUAGES, **CODEC_IDS)}
    codec = {"decoder_config": dict(
        num_quantizers=C["nq"], codebook_size=C["cb_size"], codebook_dim=C["cb_dim"], hidden_size=C["hidden"],
        latent_dim=C["latent"], num_hidden_layers=C["layers"], num_attention_heads=C["heads"],
        num_key_value_heads=C["kv"], intermediate_size=C["inter"], sliding_window=C["window"],
        decoder_dim=C["dec"], rms_norm_eps=1e-5, layer_scale_initial_scale=0.01,
        upsample_rates=C["rates"], upsampling_ratios=C["ratios"])}
    return main, codec


def write_model_dir(path, spec, main, cod):
    """safetensors + the two config.json files, as qwen_tts_load expects (c/qwen_tts.c:235-355, 862-921)"""
    os.makedirs(os.path.join(path, "speech_tokenizer"), exist_ok=True)
    cm, cc = configs(spec)
    with open(os.path.join(path, "config.json"), "w") as f:
        json.dump(cm, f, indent=1)
    with open(os.path.join(path, "speech_tokenizer", "config.json"), "w") as f:
        json.dump(cc, f, indent=1)
    _write_safetensors(os.path.join(path, "model.safetensors"), main)
    _write_safetensors(os.path.join(path, "speech_tokenizer", "model.safetensors"), cod)
