"""Pin the CPU oracle: dequantisers bit-exact vs gguf-py; forward vs logits produced by an IMPORTED implementation
(transformers LlamaForCausalLM / GraniteForCausalLM, tests/golden/make_hf_golden.py) and vs the numpy goldens."""
import os

import numpy as np
import pytest

import orc

TYPES = ["Q8_0", "Q4_K", "Q5_K", "Q6_K"]
BLOCK = {"Q8_0": (32, 34), "Q4_K": (256, 144), "Q5_K": (256, 176), "Q6_K": (256, 210)}


def _bits(a):
    return np.ascontiguousarray(a, dtype=np.float32).view(np.uint32)


def _same_bits(a, b):
    """Bit-exact, except that any NaN matches any NaN (payload propagation is not specified by numpy)."""
    a, b = np.asarray(a, np.float32), np.asarray(b, np.float32)
    nan = np.isnan(a) & np.isnan(b)
    return bool(np.all((_bits(a) == _bits(b)) | nan))


@pytest.mark.parametrize("t", TYPES)
def test_dequant_golden_blocks(golden_dir, t):
    g = np.load(os.path.join(golden_dir, "dequant_blocks.npz"))
    raw, want = g[f"{t}_raw"], g[f"{t}_f32"]
    got = orc.dequantize(t, raw, want.size).reshape(want.shape)
    assert _same_bits(got, want)


@pytest.mark.parametrize("t", TYPES)
def test_dequant_live_gguf_py(t):
    """Same check against gguf-py run live on fresh random blocks (gguf-py is on both boxes' image)."""
    gguf = pytest.importorskip("gguf")
    from gguf.quants import dequantize
    rng = np.random.default_rng(99)
    be, bb = BLOCK[t]
    raw = rng.integers(0, 256, size=(257, bb), dtype=np.uint8)
    with np.errstate(all="ignore"):
        want = dequantize(raw, getattr(gguf.GGMLQuantizationType, t)).astype(np.float32)
    got = orc.dequantize(t, raw, want.size).reshape(want.shape)
    assert _same_bits(got, want)


def test_dequant_rejects_ragged():
    with pytest.raises(ValueError):
        orc.dequantize("Q4_K", np.zeros(144, np.uint8), 100)


def test_dequant_empty():
    assert orc.dequantize("Q8_0", np.zeros(0, np.uint8), 0).size == 0


@pytest.mark.parametrize("name", ["tiny-llama", "tiny-granite"])
def test_forward_matches_numpy_golden(golden_dir, name):
    g = np.load(os.path.join(golden_dir, f"{name}.npz"))
    m = orc.Model(os.path.join(golden_dir, f"{name}.gguf"))
    logits = m.forward(g["tokens"])
    ref = g["logits"]
    scale = np.abs(ref).max()
    assert np.abs(logits - ref).max() <= 2e-5 * scale
    # greedy stream: oracle regenerates the golden continuation from the prompt
    n_prompt = int(g["n_prompt"])
    toks = g["tokens"]
    out = m.generate(toks[:n_prompt], len(toks) - n_prompt)
    safe = g["margin"][n_prompt - 1:-1] > 1e-3 * scale
    first_unsafe = int(np.argmin(safe)) if not safe.all() else len(out)
    assert np.array_equal(out[:first_unsafe], toks[n_prompt:n_prompt + first_unsafe])


@pytest.mark.parametrize("name", ["tiny-llama", "tiny-granite", "small-llama"])
def test_forward_matches_transformers(golden_dir, name, tmp_path):
    """The pin: oracle.c's forward against Hugging Face transformers run on the same GGUF tensors (fp64, CPU).
    Tolerance 1e-5 of the largest logit (measured 2e-7: the oracle accumulates in fp64 and rounds to fp32)."""
    g = np.load(os.path.join(golden_dir, f"{name}.hf.npz"))
    if name == "small-llama":                       # regenerated deterministically (not committed: 3.6 MB)
        import make_gguf
        path = str(tmp_path / "small-llama.gguf")
        make_gguf.write_model(path, make_gguf.spec_from_preset("small-llama"))
    else:
        path = os.path.join(golden_dir, f"{name}.gguf")
    logits = orc.Model(path).forward(g["tokens"])
    ref = g["logits"]
    assert logits.shape == ref.shape
    assert np.abs(logits - ref).max() <= 1e-5 * np.abs(ref).max()
    assert np.array_equal(logits.argmax(axis=1), ref.argmax(axis=1))


def test_incremental_equals_teacher_forced(golden_dir):
    g = np.load(os.path.join(golden_dir, "tiny-llama.npz"))
    m = orc.Model(os.path.join(golden_dir, "tiny-llama.gguf"))
    toks = g["tokens"][:10]
    full = m.forward(toks)
    m.reset()
    for i, t in enumerate(toks):
        lg = m.eval_token(t)
        assert np.array_equal(lg, full[i])


def test_engine_numerics_mode_is_close(golden_dir):
    """act_quant=1 / kv_f16=1 (the engine's GEMV + KV numerics) stays within the stated tolerance."""
    g = np.load(os.path.join(golden_dir, "tiny-llama.npz"))
    path = os.path.join(golden_dir, "tiny-llama.gguf")
    a = orc.Model(path).forward(g["tokens"])
    b = orc.Model(path, act_quant=1, kv_f16=1).forward(g["tokens"])
    # int8 activations (ggml quantises activations to Q8 before every quantised mat-vec) move the
    # logits by ~1-2 % of the largest logit on this 256-wide model; fp16 KV alone by ~2e-4.
    assert np.abs(a - b).max() <= 3e-2 * np.abs(a).max()
    assert np.abs(a - b).max() > 0
    c = orc.Model(path, act_quant=0, kv_f16=1).forward(g["tokens"])
    assert np.abs(a - c).max() <= 1e-3 * np.abs(a).max()


def test_load_rejects_garbage(tmp_path):
    p = tmp_path / "bad.gguf"
    p.write_bytes(b"NOTGGUF" + bytes(100))
    with pytest.raises(RuntimeError):
        orc.Model(str(p))
    q = tmp_path / "trunc.gguf"
    q.write_bytes(b"GGUF" + (3).to_bytes(4, "little") + (5).to_bytes(8, "little"))
    with pytest.raises(RuntimeError):
        orc.Model(str(q))
