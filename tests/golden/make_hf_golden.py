#!/usr/bin/env python3
"""Pin the forward pass to an IMPORTED implementation: Hugging Face transformers (run here, in the
build container; the outputs are committed, this script is how they were made).

The reference's runtime for this path (LM Studio's llama.cpp, called at
/root/reference/nats_llm_studio.go:158-179) is not installable offline, and the reference holds no
golden vectors.  `transformers` (5.5.0 on this image) ships independent
implementations of both architectures the path serves — `LlamaForCausalLM` and
`GraniteForCausalLM` — so the committed tiny GGUFs are loaded with llama.cpp's own gguf-py reader,
dequantised with `gguf.quants.dequantize`, copied into those classes (CPU, float64) and run.
`tests/test_oracle.py::test_forward_matches_transformers` holds oracle/oracle.c to these logits.

RoPE convention.  GGUF llama/granite files store attn_q / attn_k rows in the adjacent-pair order
(rotation pairs are rows (2i, 2i+1) of a head; llama.cpp's converter permutes HF checkpoints into
it).  transformers rotates pairs (i, i + head_dim/2).  The GGUF rows are therefore un-permuted on
the way in (`unpermute`, the inverse of convert_hf_to_gguf.py's `permute`); q·k is invariant under
a permutation applied to both, so the logits are those of the GGUF model.

Outputs (tests/golden/): tiny-llama.hf.npz, tiny-granite.hf.npz, small-llama.hf.npz
  tokens   the token ids of the numpy golden (same prompt + continuation), or a seeded prompt
  logits   teacher-forced logits [n, V] from transformers (float64 -> float32)
  versions transformers / torch / gguf versions used
"""
import os
import sys

import numpy as np
import torch
import transformers
import gguf
from gguf import GGUFReader, GGMLQuantizationType as QT
from gguf.quants import dequantize

HERE = os.path.dirname(os.path.abspath(__file__))
sys.path.insert(0, os.path.join(HERE, "..", "..", "tools"))


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
        shape = [int(x) for x in t.shape]
        tens[t.name] = dequantize(np.array(t.data), QT(t.tensor_type)).astype(np.float64).reshape(shape[::-1])
    return kv, tens


def unpermute(w, n_head):
    """GGUF (adjacent-pair) row order -> HF (split-half) row order, per head."""
    hd = w.shape[0] // n_head
    return w.reshape(n_head, hd // 2, 2, *w.shape[1:]).swapaxes(1, 2).reshape(w.shape)


def hf_model(kv, tens):
    arch = kv["general.architecture"]
    g = lambda k, d=None: kv.get(f"{arch}.{k}", d)
    L, D, H, HK = g("block_count"), g("embedding_length"), g("attention.head_count"), g("attention.head_count_kv")
    HD = g("attention.key_length", D // H)
    V = tens["token_embd.weight"].shape[0]
    tied = "output.weight" not in tens
    common = dict(vocab_size=V, hidden_size=D, intermediate_size=g("feed_forward_length"), num_hidden_layers=L,
                  num_attention_heads=H, num_key_value_heads=HK, head_dim=HD, rms_norm_eps=g("attention.layer_norm_rms_epsilon"),
                  rope_theta=g("rope.freq_base", 10000.0), max_position_embeddings=4096, attention_bias=False, mlp_bias=False,
                  tie_word_embeddings=tied, hidden_act="silu", attn_implementation="eager")
    if arch == "granite":
        cfg = transformers.GraniteConfig(**common, embedding_multiplier=g("embedding_scale"), residual_multiplier=g("residual_scale"),
                                         attention_multiplier=g("attention.scale"), logits_scaling=g("logit_scale"))
        m = transformers.GraniteForCausalLM(cfg)
    else:
        cfg = transformers.LlamaConfig(**common)
        m = transformers.LlamaForCausalLM(cfg)
    m = m.double().eval()
    sd = {"model.embed_tokens.weight": tens["token_embd.weight"], "model.norm.weight": tens["output_norm.weight"]}
    if not tied:
        sd["lm_head.weight"] = tens["output.weight"]
    for l in range(L):
        p, q = f"blk.{l}.", f"model.layers.{l}."
        sd[q + "input_layernorm.weight"] = tens[p + "attn_norm.weight"]
        sd[q + "post_attention_layernorm.weight"] = tens[p + "ffn_norm.weight"]
        sd[q + "self_attn.q_proj.weight"] = unpermute(tens[p + "attn_q.weight"], H)
        sd[q + "self_attn.k_proj.weight"] = unpermute(tens[p + "attn_k.weight"], HK)
        sd[q + "self_attn.v_proj.weight"] = tens[p + "attn_v.weight"]
        sd[q + "self_attn.o_proj.weight"] = tens[p + "attn_output.weight"]
        sd[q + "mlp.gate_proj.weight"] = tens[p + "ffn_gate.weight"]
        sd[q + "mlp.up_proj.weight"] = tens[p + "ffn_up.weight"]
        sd[q + "mlp.down_proj.weight"] = tens[p + "ffn_down.weight"]
    sd = {k: torch.from_numpy(np.ascontiguousarray(v)) for k, v in sd.items()}
    if tied:
        sd["lm_head.weight"] = sd["model.embed_tokens.weight"]
    missing, unexpected = m.load_state_dict(sd, strict=False)
    assert not unexpected and all("rotary" in k or "inv_freq" in k for k in missing), (missing, unexpected)
    return m


def hf_logits(path, toks):
    kv, tens = load(path)
    m = hf_model(kv, tens)
    with torch.no_grad():
        out = m(input_ids=torch.tensor(np.asarray(toks, dtype=np.int64))[None, :]).logits[0]
    return out.numpy()


def main():
    import make_gguf
    ver = f"transformers {transformers.__version__}, torch {torch.__version__}, gguf {getattr(gguf, '__version__', '0.19.0')}"
    for name in ("tiny-llama", "tiny-granite"):
        toks = np.load(os.path.join(HERE, f"{name}.npz"))["tokens"]
        lg = hf_logits(os.path.join(HERE, f"{name}.gguf"), toks)
        np.savez_compressed(os.path.join(HERE, f"{name}.hf.npz"), tokens=toks.astype(np.int32), logits=lg.astype(np.float32), versions=ver)
        print(name, lg.shape, "max|logit|", float(np.abs(lg).max()))
    # head_dim 128, GQA 4:1, Q4_K_M type mix (the 8B's code paths at test size); the GGUF itself is regenerated
    # deterministically by tools/make_gguf.py wherever it is needed (3.6 MB, not committed), only its logits are.
    path = "/tmp/small-llama.golden.gguf"
    make_gguf.write_model(path, make_gguf.spec_from_preset("small-llama"))
    toks = np.random.default_rng(4321).integers(0, 2048, 20).astype(np.int32)
    lg = hf_logits(path, toks)
    np.savez_compressed(os.path.join(HERE, "small-llama.hf.npz"), tokens=toks, logits=lg.astype(np.float32), versions=ver)
    print("small-llama", lg.shape, "max|logit|", float(np.abs(lg).max()))


if __name__ == "__main__":
    main()
