#!/usr/bin/env python3
"""Generate the committed golden fixtures (run here, in the build container; results committed).

The reference has no tests and no golden vectors for the generation path, and
its runtime (LM Studio / llama.cpp) cannot be installed offline.  The fixtures are therefore made
from the only independent implementation on this image: llama.cpp's own gguf-py package
(GGUFReader + gguf.quants.dequantize) driving a float64 numpy forward pass written here.  The C
oracle (oracle/oracle.c) is pinned against these files; the CUDA engine is then checked against
the oracle and against these files.

Outputs (tests/golden/):
  tiny-llama.gguf, tiny-granite.gguf   seeded synthetic models (tools/make_gguf.py)
  tiny-*.npz                           prompt, teacher-forced logits [n, V] (float64 -> float32),
                                       greedy continuation
  dequant_blocks.npz                   random block bytes per type + gguf-py dequantised values,
                                       including edge blocks (d=0, inf/nan fp16, all-0xFF)
"""
import os
import sys

import numpy as np
from gguf import GGUFReader, GGMLQuantizationType as QT
from gguf.quants import dequantize

HERE = os.path.dirname(os.path.abspath(__file__))
sys.path.insert(0, os.path.join(HERE, "..", "..", "tools"))
import make_gguf  # noqa: E402


def load(path):
    r = GGUFReader(path)
    kv = {}
    for f in r.fields.values():
        try:
            kv[f.name] = f.contents()
        except Exception:
            pass
    tens = {}
    for t in r.tensors:
        data = np.array(t.data)
        shape = [int(x) for x in t.shape]          # ne[0] (inner) first
        arr = dequantize(data, QT(t.tensor_type)).astype(np.float64)
        tens[t.name] = arr.reshape(shape[::-1])
    return kv, tens


def forward(kv, tens, toks):
    arch = kv["general.architecture"]
    g = lambda k, d=None: kv.get(f"{arch}.{k}", d)
    L, D, H, HK = g("block_count"), g("embedding_length"), g("attention.head_count"), g("attention.head_count_kv")
    HD = g("attention.key_length", D // H)
    eps, base, nrot = g("attention.layer_norm_rms_epsilon"), g("rope.freq_base", 10000.0), g("rope.dimension_count", HD)
    emb_s, res_s = g("embedding_scale", 0.0), g("residual_scale", 0.0)
    att_s, log_s = g("attention.scale", 0.0), g("logit_scale", 0.0)
    kq = att_s if att_s else 1.0 / np.sqrt(HD)
    res = res_s if res_s else 1.0
    n = len(toks)
    h = tens["token_embd.weight"][toks].copy()
    if emb_s:
        h *= emb_s
    pos = np.arange(n)[:, None]
    inv = base ** (-2.0 * np.arange(nrot // 2) / nrot)
    th = pos * inv[None, :]
    cos, sin = np.cos(th), np.sin(th)

    def rms(x, w):
        return x / np.sqrt((x * x).mean(axis=-1, keepdims=True) + eps) * w

    def rope(x):                                   # x: [n, heads, HD], adjacent pairs
        a, b = x[..., 0:nrot:2], x[..., 1:nrot:2]
        out = x.copy()
        out[..., 0:nrot:2] = a * cos[:, None, :] - b * sin[:, None, :]
        out[..., 1:nrot:2] = a * sin[:, None, :] + b * cos[:, None, :]
        return out

    mask = np.tril(np.ones((n, n), dtype=bool))
    for l in range(L):
        p = f"blk.{l}."
        x = rms(h, tens[p + "attn_norm.weight"])
        q = rope((x @ tens[p + "attn_q.weight"].T).reshape(n, H, HD))
        k = rope((x @ tens[p + "attn_k.weight"].T).reshape(n, HK, HD))
        v = (x @ tens[p + "attn_v.weight"].T).reshape(n, HK, HD)
        rep = H // HK
        out = np.zeros((n, H, HD))
        for hh in range(H):
            s = (q[:, hh] @ k[:, hh // rep].T) * kq
            s = np.where(mask, s, -np.inf)
            s = np.exp(s - s.max(axis=1, keepdims=True))
            s /= s.sum(axis=1, keepdims=True)
            out[:, hh] = s @ v[:, hh // rep]
        h = h + (out.reshape(n, H * HD) @ tens[p + "attn_output.weight"].T) * res
        x = rms(h, tens[p + "ffn_norm.weight"])
        gt = x @ tens[p + "ffn_gate.weight"].T
        up = x @ tens[p + "ffn_up.weight"].T
        h = h + ((gt / (1.0 + np.exp(-gt)) * up) @ tens[p + "ffn_down.weight"].T) * res
    x = rms(h, tens["output_norm.weight"])
    wout = tens.get("output.weight", tens["token_embd.weight"])
    logits = x @ wout.T
    if log_s:
        logits = logits / log_s
    return logits


def golden_model(preset, n_prompt=12, n_gen=12, **over):
    path = os.path.join(HERE, f"{preset}.gguf")
    spec = make_gguf.spec_from_preset(preset, **over)
    make_gguf.write_model(path, spec)
    kv, tens = load(path)
    rng = np.random.default_rng(1234)
    V = tens["token_embd.weight"].shape[0]
    toks = [int(t) for t in rng.integers(0, V, n_prompt)]
    for _ in range(n_gen):                          # greedy continuation (full recompute: tiny model)
        lg = forward(kv, tens, np.array(toks))
        toks.append(int(lg[-1].argmax()))
    logits = forward(kv, tens, np.array(toks))
    srt = np.sort(logits, axis=1)
    np.savez_compressed(os.path.join(HERE, f"{preset}.npz"), tokens=np.array(toks, dtype=np.int32),
                        n_prompt=np.int32(n_prompt), logits=logits.astype(np.float32),
                        margin=(srt[:, -1] - srt[:, -2]).astype(np.float32))
    print(preset, "tokens", toks, "min greedy margin", float((srt[:, -1] - srt[:, -2])[n_prompt - 1:-1].min()))


def golden_blocks():
    rng = np.random.default_rng(7)
    out = {}
    for qt in (QT.Q8_0, QT.Q4_K, QT.Q5_K, QT.Q6_K):
        be, bb = make_gguf.BLOCK[qt]
        raw = rng.integers(0, 256, size=(64, bb), dtype=np.uint8)
        dpos = {QT.Q8_0: [0], QT.Q4_K: [0, 2], QT.Q5_K: [0, 2], QT.Q6_K: [208]}[qt]
        # first 48 blocks: finite, modest fp16 scales; rest keep arbitrary bit patterns (inf/nan/subnormal)
        for p in dpos:
            raw[:48, p:p + 2] = (rng.standard_normal(48).astype(np.float16) * np.float16(0.01)).view(np.uint8).reshape(48, 2)
        raw[48, :] = 0xFF                                    # all-ones block (nan scale, max codes)
        raw[49, :] = 0x00                                    # all-zero block (d = 0)
        for p in dpos:
            raw[50, p:p + 2] = np.array([0x00, 0x7C], dtype=np.uint8)   # +inf
            raw[51, p:p + 2] = np.array([0x01, 0x00], dtype=np.uint8)   # smallest subnormal
            raw[52, p:p + 2] = np.array([0xFF, 0xFB], dtype=np.uint8)   # -65504
        with np.errstate(all="ignore"):
            out[f"{qt.name}_raw"] = raw
            out[f"{qt.name}_f32"] = dequantize(raw, qt).astype(np.float32)
    np.savez_compressed(os.path.join(HERE, "dequant_blocks.npz"), **out)
    print("dequant blocks written")


if __name__ == "__main__":
    golden_blocks()
    golden_model("tiny-llama")
    golden_model("tiny-granite", gain=3.0)
