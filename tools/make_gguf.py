#!/usr/bin/env python3
"""Synthetic random-init GGUF writer (fixtures for tests, smoke and bench).

There is no network on the build or GPU boxes, so no real checkpoint exists; BASELINE.json asks
for "GGUF files of the named architectures with random-init weights".  This tool writes them with
llama.cpp's own gguf-py `GGUFWriter`, so the container bytes are produced by an implementation that
is independent of both the engine's C++ parser and the oracle's C parser.

Two weight modes:
  * ``bytes``    – block bytes sampled directly (uniform nibbles / scales, fp16 d chosen so the
                   dequantised weights have std ~ gain/sqrt(K)).  Fast enough for 8B/70B shapes.
  * ``quantize`` – float weights N(0, gain/sqrt(K)) quantised by the small numpy quantisers below
                   (valid encodings, not ggml's exact search) – used for small test models so the
                   network behaves like a real quantised one.

The vocabulary is synthetic and byte-level: ids 0..255 are `<0xNN>` byte tokens, then the special
tokens, then filler tokens `▁t<id>`; no tokenizer files exist on the box.
"""
from __future__ import annotations

import argparse
import sys
from dataclasses import dataclass, field

import numpy as np
from gguf import GGUFWriter, GGMLQuantizationType as QT

QK_K = 256
BLOCK = {QT.Q8_0: (32, 34), QT.Q4_K: (256, 144), QT.Q5_K: (256, 176), QT.Q6_K: (256, 210)}

SPECIALS = ["<|bos|>", "<|eos|>", "<|system|>", "<|user|>", "<|assistant|>", "<|end|>"]


@dataclass
class ModelSpec:
    arch: str = "llama"            # "llama" | "granite"
    n_layer: int = 2
    n_embd: int = 256
    n_head: int = 4
    n_head_kv: int = 2
    head_dim: int = 0              # 0 -> n_embd // n_head
    n_ff: int = 512
    n_vocab: int = 512
    n_ctx: int = 2048
    rope_base: float = 500000.0
    rms_eps: float = 1e-5
    mix: str = "q4_k_m"            # "q4_k_m" | "q8_0" | "q6_k" | "q5_k" | "q4_k" | "q5_k_m"
    mode: str = "quantize"         # "quantize" | "bytes"
    seed: int = 1
    tied: bool = False             # no output.weight (granite-3.0-2b ties embeddings)
    reuse_layers: bool = False     # every layer repeats the bytes of the first layer with the same (tensor, type): fast to write, same size
    gain: float = 1.0
    out_gain: float = 1.0
    name: str = "synthetic"
    granite_scalars: dict = field(default_factory=lambda: dict(
        embedding_scale=12.0, residual_scale=0.22, attention_scale=0.015625, logit_scale=8.0))

    @property
    def hd(self) -> int:
        return self.head_dim or self.n_embd // self.n_head


PRESETS = {
    # shapes from the public HF configs
    "llama3-8b": dict(arch="llama", n_layer=32, n_embd=4096, n_head=32, n_head_kv=8, n_ff=14336, n_vocab=128256,
                      n_ctx=8192, mode="bytes"),
    "llama3-70b": dict(arch="llama", n_layer=80, n_embd=8192, n_head=64, n_head_kv=8, n_ff=28672, n_vocab=128256,
                       n_ctx=8192, mode="bytes"),
    "granite3-2b": dict(arch="granite", n_layer=40, n_embd=2048, n_head=32, n_head_kv=8, n_ff=8192, n_vocab=49155,
                        n_ctx=4096, rope_base=10000.0, mix="q8_0", tied=True, mode="bytes"),
    # head_dim 128, GQA 4:1, q4_k_m mix incl. a Q6_K attn_v next to Q4_K q/k: the 8B code paths at test size
    "small-llama": dict(arch="llama", n_layer=2, n_embd=1024, n_head=8, n_head_kv=2, n_ff=2816, n_vocab=2048),
    "tiny-llama": dict(arch="llama", n_layer=2, n_embd=256, n_head=4, n_head_kv=2, n_ff=512, n_vocab=512),
    "tiny-granite": dict(arch="granite", n_layer=2, n_embd=256, n_head=4, n_head_kv=2, n_ff=512, n_vocab=515,
                         mix="q8_0", tied=True, rope_base=10000.0),
}


# ----------------------------------------------------------------------------- numpy quantisers
def _f16(a):
    return np.asarray(a, dtype=np.float32).astype(np.float16)


def _pack_scale_min(sc: np.ndarray, mn: np.ndarray) -> np.ndarray:
    """Inverse of gguf/quants.py Q4_K.get_scale_min: 8+8 six-bit values -> 12 bytes."""
    n = sc.shape[0]
    out = np.zeros((n, 12), dtype=np.uint8)
    sc = sc.astype(np.uint8); mn = mn.astype(np.uint8)
    out[:, 0:4] = (sc[:, 0:4] & 63) | ((sc[:, 4:8] >> 4) << 6)
    out[:, 4:8] = (mn[:, 0:4] & 63) | ((mn[:, 4:8] >> 4) << 6)
    out[:, 8:12] = (sc[:, 4:8] & 0x0F) | ((mn[:, 4:8] & 0x0F) << 4)
    return out


def _quant_k45(x: np.ndarray, bits: int) -> np.ndarray:
    """Q4_K (bits=4) / Q5_K (bits=5) encoder; x is (rows, K) float32."""
    levels = (1 << bits) - 1
    xb = x.reshape(-1, 8, 32).astype(np.float32)
    mn = np.minimum(xb.min(axis=2), 0.0)
    mx = np.maximum(xb.max(axis=2), 0.0)
    scale = (mx - mn) / levels                      # (n, 8)
    mpos = -mn
    d = scale.max(axis=1, keepdims=True) / 63.0
    dmin = mpos.max(axis=1, keepdims=True) / 63.0
    d16 = _f16(d); dmin16 = _f16(dmin)
    df = d16.astype(np.float32); dminf = dmin16.astype(np.float32)
    with np.errstate(divide="ignore", invalid="ignore"):
        sc = np.where(df > 0, np.rint(scale / df), 0).clip(0, 63)
        mq = np.where(dminf > 0, np.rint(mpos / dminf), 0).clip(0, 63)
        eff = (df * sc)[:, :, None]
        q = np.where(eff > 0, np.rint((xb + (dminf * mq)[:, :, None]) / eff), 0).clip(0, levels).astype(np.uint8)
    n = q.shape[0]
    ql = q & 0x0F
    qs = (ql[:, 0::2, :] | (ql[:, 1::2, :] << 4)).reshape(n, 128)
    parts = [d16.view(np.uint8).reshape(n, 2), dmin16.view(np.uint8).reshape(n, 2), _pack_scale_min(sc, mq)]
    if bits == 5:
        hb = (q >> 4) & 1                            # (n, 8, 32)
        qh = np.zeros((n, 32), dtype=np.uint8)
        for j in range(8):
            qh |= (hb[:, j, :] << j).astype(np.uint8)
        parts.append(qh)
    parts.append(qs)
    return np.concatenate(parts, axis=1).reshape(x.shape[0], -1)


def _quant_q6k(x: np.ndarray) -> np.ndarray:
    xb = x.reshape(-1, 16, 16).astype(np.float32)
    n = xb.shape[0]
    amax_idx = np.abs(xb).argmax(axis=2)
    smax = np.take_along_axis(xb, amax_idx[:, :, None], axis=2)[:, :, 0]
    scale = smax / -32.0
    dd = np.abs(scale).max(axis=1, keepdims=True) / 127.0
    d16 = _f16(dd); df = d16.astype(np.float32)
    with np.errstate(divide="ignore", invalid="ignore"):
        sc = np.where(df > 0, np.rint(scale / df), 0).clip(-128, 127).astype(np.int8)
        eff = (df * sc.astype(np.float32))[:, :, None]
        q = (np.where(eff != 0, np.rint(xb / eff), 0).clip(-32, 31) + 32).astype(np.uint8)
    q = q.reshape(n, 2, 4, 32)                       # [half][grp][l]
    lo = q & 0x0F
    hi = (q >> 4) & 3
    ql = np.zeros((n, 2, 64), dtype=np.uint8)
    ql[:, :, 0:32] = lo[:, :, 0, :] | (lo[:, :, 2, :] << 4)
    ql[:, :, 32:64] = lo[:, :, 1, :] | (lo[:, :, 3, :] << 4)
    qh = hi[:, :, 0, :] | (hi[:, :, 1, :] << 2) | (hi[:, :, 2, :] << 4) | (hi[:, :, 3, :] << 6)
    out = np.concatenate([ql.reshape(n, 128), qh.reshape(n, 64).astype(np.uint8), sc.view(np.uint8).reshape(n, 16),
                          d16.view(np.uint8).reshape(n, 2)], axis=1)
    return out.reshape(x.shape[0], -1)


def _quant_q8_0(x: np.ndarray) -> np.ndarray:
    from gguf.quants import quantize
    return quantize(x.astype(np.float32), QT.Q8_0)


def quantize_rows(x: np.ndarray, qt: QT) -> np.ndarray:
    if qt == QT.Q8_0:
        return _quant_q8_0(x)
    if qt == QT.Q4_K:
        return _quant_k45(x, 4)
    if qt == QT.Q5_K:
        return _quant_k45(x, 5)
    if qt == QT.Q6_K:
        return _quant_q6k(x)
    raise ValueError(qt)


def random_block_bytes(rng: np.random.Generator, rows: int, K: int, qt: QT, sigma: float) -> np.ndarray:
    """Sample block bytes directly; fp16 scales set so dequantised std ~ sigma."""
    be, bb = BLOCK[qt]
    nb = rows * (K // be)
    raw = rng.integers(0, 256, size=(nb, bb), dtype=np.uint8)
    jitter = (0.5 + rng.random(nb, dtype=np.float32))
    if qt == QT.Q8_0:
        raw[:, 0:2] = _f16(sigma / 73.9 * jitter).view(np.uint8).reshape(nb, 2)
    elif qt in (QT.Q4_K, QT.Q5_K):
        qmean, std_u = (7.5, 258.0) if qt == QT.Q4_K else (15.5, 520.0)
        d = sigma / std_u * jitter
        raw[:, 0:2] = _f16(d).view(np.uint8).reshape(nb, 2)
        raw[:, 2:4] = _f16(d * qmean).view(np.uint8).reshape(nb, 2)
    elif qt == QT.Q6_K:
        raw[:, 208:210] = _f16(sigma / 1369.0 * jitter).view(np.uint8).reshape(nb, 2)
    return raw.reshape(rows, -1)


# ----------------------------------------------------------------------------- type mix
def tensor_type(spec: ModelSpec, name: str, layer: int) -> QT:
    """Per-tensor type.  q4_k_m follows llama.cpp's published mix rule ([external])."""
    mix = spec.mix
    simple = {"q8_0": QT.Q8_0, "q6_k": QT.Q6_K, "q5_k": QT.Q5_K, "q4_k": QT.Q4_K}
    if mix in simple:
        return simple[mix]
    base, more = (QT.Q4_K, QT.Q6_K) if mix == "q4_k_m" else (QT.Q5_K, QT.Q6_K)
    if name == "output":
        return QT.Q6_K
    if name in ("attn_v", "ffn_down"):
        L = spec.n_layer
        use_more = layer < L // 8 or layer >= 7 * L // 8 or (layer - L // 8) % 3 == 2
        if use_more:
            return more
        if name == "attn_v" and spec.n_layer >= 80 and mix == "q4_k_m":
            return QT.Q5_K
    return base


def vocab_tokens(n_vocab: int):
    toks, types = [], []
    for i in range(min(256, n_vocab)):
        toks.append(f"<0x{i:02X}>"); types.append(6)          # BYTE
    for s in SPECIALS:
        if len(toks) < n_vocab:
            toks.append(s); types.append(3)                   # CONTROL
    while len(toks) < n_vocab:
        toks.append(f"▁t{len(toks)}"); types.append(1)   # NORMAL
    return toks, types


def write_model(path: str, spec: ModelSpec, verbose: bool = False) -> dict:
    rng = np.random.Generator(np.random.PCG64(spec.seed))
    w = GGUFWriter(path, spec.arch)
    w.add_name(spec.name)
    w.add_block_count(spec.n_layer)
    w.add_context_length(spec.n_ctx)
    w.add_embedding_length(spec.n_embd)
    w.add_feed_forward_length(spec.n_ff)
    w.add_head_count(spec.n_head)
    w.add_head_count_kv(spec.n_head_kv)
    w.add_layer_norm_rms_eps(spec.rms_eps)
    w.add_rope_dimension_count(spec.hd)
    w.add_rope_freq_base(spec.rope_base)
    w.add_vocab_size(spec.n_vocab)
    if spec.head_dim:
        w.add_key_length(spec.hd); w.add_value_length(spec.hd)
    ftype = {"q8_0": 7, "q4_k_m": 15, "q5_k_m": 17, "q6_k": 18, "q4_k": 14, "q5_k": 16}.get(spec.mix, 0)
    w.add_file_type(ftype)
    if spec.arch == "granite":
        g = spec.granite_scalars
        w.add_embedding_scale(g["embedding_scale"]); w.add_residual_scale(g["residual_scale"])
        w.add_attention_scale(g["attention_scale"]); w.add_logit_scale(g["logit_scale"])
    toks, types = vocab_tokens(spec.n_vocab)
    w.add_tokenizer_model("llama")
    w.add_token_list(toks)
    w.add_token_types(types)
    w.add_bos_token_id(256 if spec.n_vocab > 256 else 0)
    w.add_eos_token_id(257 if spec.n_vocab > 257 else 0)
    w.add_chat_template("{% for m in messages %}<|{{m.role}}|>{{m.content}}<|end|>{% endfor %}<|assistant|>")

    D, F, HD = spec.n_embd, spec.n_ff, spec.hd
    total = 0

    cache = {}

    def add_matrix(name: str, rows: int, K: int, qt: QT, gain: float):
        nonlocal total
        sigma = gain / np.sqrt(K)
        key = (name.split(".")[-2], qt, rows, K)
        if spec.reuse_layers and key in cache:
            w.add_tensor(name, cache[key], raw_dtype=qt); total += cache[key].nbytes
            return
        if spec.mode == "bytes":
            data = random_block_bytes(rng, rows, K, qt, sigma)
        else:
            data = quantize_rows(rng.standard_normal((rows, K), dtype=np.float32) * np.float32(sigma), qt)
        w.add_tensor(name, data, raw_dtype=qt)
        total += data.nbytes
        if spec.reuse_layers:
            cache[key] = data
        if verbose:
            print(f"  {name:32s} {qt.name:5s} {rows}x{K} {data.nbytes/1e6:.1f} MB", file=sys.stderr)

    def add_norm(name: str, n: int):
        nonlocal total
        v = (1.0 + 0.1 * rng.standard_normal(n, dtype=np.float32)).astype(np.float32)
        w.add_tensor(name, v); total += v.nbytes

    emb_t = tensor_type(spec, "token_embd", 0) if not spec.tied else tensor_type(spec, "output", 0)
    if spec.mix in ("q4_k_m", "q5_k_m") and not spec.tied:
        emb_t = QT.Q4_K if spec.mix == "q4_k_m" else QT.Q5_K
    add_matrix("token_embd.weight", spec.n_vocab, D, emb_t, spec.out_gain if spec.tied else 1.0 * np.sqrt(D))
    for l in range(spec.n_layer):
        add_norm(f"blk.{l}.attn_norm.weight", D)
        add_matrix(f"blk.{l}.attn_q.weight", spec.n_head * HD, D, tensor_type(spec, "attn_q", l), spec.gain)
        add_matrix(f"blk.{l}.attn_k.weight", spec.n_head_kv * HD, D, tensor_type(spec, "attn_k", l), spec.gain)
        add_matrix(f"blk.{l}.attn_v.weight", spec.n_head_kv * HD, D, tensor_type(spec, "attn_v", l), spec.gain)
        add_matrix(f"blk.{l}.attn_output.weight", D, spec.n_head * HD, tensor_type(spec, "attn_output", l), spec.gain)
        add_norm(f"blk.{l}.ffn_norm.weight", D)
        add_matrix(f"blk.{l}.ffn_gate.weight", F, D, tensor_type(spec, "ffn_gate", l), spec.gain)
        add_matrix(f"blk.{l}.ffn_up.weight", F, D, tensor_type(spec, "ffn_up", l), spec.gain)
        add_matrix(f"blk.{l}.ffn_down.weight", D, F, tensor_type(spec, "ffn_down", l), spec.gain)
    add_norm("output_norm.weight", D)
    if not spec.tied:
        add_matrix("output.weight", spec.n_vocab, D, tensor_type(spec, "output", 0), spec.out_gain)

    w.write_header_to_file()
    w.write_kv_data_to_file()
    w.write_tensors_to_file()
    w.close()
    return {"path": path, "tensor_bytes": total}


def spec_from_preset(preset: str, **over) -> ModelSpec:
    kw = dict(PRESETS[preset]); kw.update(over); kw.setdefault("name", preset)
    return ModelSpec(**kw)


def main(argv=None):
    ap = argparse.ArgumentParser(description=__doc__, formatter_class=argparse.RawDescriptionHelpFormatter)
    ap.add_argument("out")
    ap.add_argument("--preset", default="tiny-llama", choices=sorted(PRESETS))
    ap.add_argument("--mix"); ap.add_argument("--mode"); ap.add_argument("--seed", type=int)
    ap.add_argument("--n-layer", type=int); ap.add_argument("-v", "--verbose", action="store_true")
    a = ap.parse_args(argv)
    over = {k: v for k, v in dict(mix=a.mix, mode=a.mode, seed=a.seed, n_layer=a.n_layer).items() if v is not None}
    info = write_model(a.out, spec_from_preset(a.preset, **over), verbose=a.verbose)
    print(f"wrote {info['path']}: {info['tensor_bytes']/1e9:.3f} GB of tensor data")


if __name__ == "__main__":
    main()
