"""GPU parity tests, model level: logits and greedy streams vs the CPU oracle and the committed
numpy goldens, through the C ABI; the NATS subject handlers end to end."""
import json
import os

import numpy as np
import pytest

import orc
import nats_llm_studio_b200 as b200

pytestmark = pytest.mark.gpu

# Tolerances (stated, see DESIGN.md "Numerics"), relative to the largest |logit|:
#  * TOL_EXACT: vs the oracle in engine-numerics mode (int8 activations per 256 block, fp16 KV) the
#    integer dot products are identical, so positions agree to fp32 ordering noise ...
TOL_EXACT = 1e-4
#  * ... until the first int8 rounding decision flips (a 1e-7 difference in an activation flips its
#    code with probability ~1e-5); from that position on the cached K/V differ and the two streams
#    agree only at quantisation-noise level, like any two implementations of ggml-style inference.
#  * TOL_QUANT: that level, and the distance to the fp32 oracle / numpy golden (int8 activation
#    rounding: ~1-2 % on these 256-wide toy models).
TOL_QUANT = 4e-2


@pytest.fixture(scope="module")
def engine():
    return b200.Engine(max_ctx=2048)


@pytest.fixture(scope="module", params=["tiny-llama", "tiny-granite"])
def loaded(request, engine, golden_dir):
    name = request.param
    path = os.path.join(golden_dir, f"{name}.gguf")
    m = engine.load_file(f"golden/{name}", path)
    yield name, path, m
    m.close()


def test_logits_match_oracle_and_golden(loaded, golden_dir):
    name, path, m = loaded
    g = np.load(os.path.join(golden_dir, f"{name}.npz"))
    toks = g["tokens"]
    got = m.eval(toks)
    ref_same = orc.Model(path, act_quant=1, kv_f16=1).forward(toks)
    ref_fp32 = g["logits"]
    scale = np.abs(ref_fp32).max()
    assert np.isfinite(got).all()
    per_pos = np.abs(got - ref_same).max(axis=1) / scale
    assert per_pos[0] <= TOL_EXACT                       # nothing upstream can flip at position 0
    assert (per_pos <= TOL_EXACT).mean() >= 0.5          # most positions precede the first flip
    assert per_pos.max() <= TOL_QUANT
    assert np.abs(got - ref_fp32).max() <= TOL_QUANT * scale


def test_greedy_stream_matches_oracle(loaded):
    """Greedy tokens equal the oracle's; a divergence is tolerated only where the oracle's own
    top-1/top-2 margin is below the logit tolerance (then both continuations are teacher-forced)."""
    name, path, m = loaded
    rng = np.random.default_rng(11)
    prompt = rng.integers(0, m.n_vocab, 9).astype(np.int32)
    n_gen = 24
    got, ttft, dec = m.generate(prompt, n_gen)
    assert len(got) == n_gen and ttft > 0
    o = orc.Model(path, act_quant=1, kv_f16=1)
    want, lg = o.generate(prompt, n_gen, want_logits=True)
    for i in range(n_gen):
        if got[i] != want[i]:
            srt = np.sort(lg[i]); margin = srt[-1] - srt[-2]
            assert margin <= TOL_QUANT * np.abs(lg[i]).max(), f"token {i}: {got[i]} != {want[i]} with margin {margin}"
            break
    # logits of the last step are also exposed
    full = np.concatenate([prompt, got])
    assert np.abs(m.eval(full[:-1], all_logits=False) - o.forward(full[:-1])[-1]).max() <= TOL_QUANT * np.abs(lg).max()


def test_small_llama_head_dim_128(engine, tmp_path):
    """8B code paths at test size (head_dim 128, 4:1 GQA, mixed Q4_K/Q6_K QKV phase): logits vs the oracle, and a
    forced-token prompt longer than one KV page followed by greedy decoding (regression: token-boundary race)."""
    import make_gguf
    path = str(tmp_path / "small.gguf")
    make_gguf.write_model(path, make_gguf.spec_from_preset("small-llama"))
    m = engine.load_file("t/small", path)
    rng = np.random.default_rng(3)
    toks = rng.integers(0, m.n_vocab, 40).astype(np.int32)
    got = m.eval(toks)
    ref = orc.Model(path, act_quant=1, kv_f16=1).forward(toks)
    per_pos = np.abs(got - ref).max(axis=1) / np.abs(ref).max()
    # wider model => more int8 codes per position => the first rounding flip comes earlier; the exact prefix must
    # still cover the first positions (attention over several cached positions, both KV pages of layer 0/1)
    assert (per_pos[:4] <= TOL_EXACT).all() and per_pos.max() <= TOL_QUANT
    out, _, _ = m.generate(toks, 20)
    want = orc.Model(path, act_quant=1, kv_f16=1).generate(toks, 20)
    assert out[0] == want[0] and len(out) == 20
    out2, _, _ = m.generate(toks, 20)
    assert np.array_equal(out, out2)
    m.close()


def test_eval_is_deterministic_and_incremental(loaded):
    name, path, m = loaded
    toks = np.arange(1, 40, dtype=np.int32) % m.n_vocab
    a = m.eval(toks)
    b = m.eval(toks)
    assert np.array_equal(a, b)
    assert np.array_equal(m.eval(toks[:17])[-1], a[16])      # prefix gives the same logits: KV cache path is consistent


def test_chunked_upload_equals_file_load(engine, golden_dir):
    """b200_model_open/upload/commit with ragged chunk sizes == load_file (the Object Store path)."""
    path = os.path.join(golden_dir, "tiny-llama.gguf")
    data = open(path, "rb").read()
    a = engine.load_file("t/file", path)
    b = engine.load_bytes("t/bytes", data, chunk=100_003)
    toks = np.array([3, 1, 4, 1, 5, 9, 2, 6], dtype=np.int32)
    assert np.array_equal(a.eval(toks), b.eval(toks))
    ids = sorted(e["id"] for e in engine.list_models()["data"])
    assert "t/bytes" in ids and "t/file" in ids
    a.close(); b.close()
    assert "t/file" not in [e["id"] for e in engine.list_models()["data"]]


def test_bad_models_are_rejected(engine, tmp_path):
    p = tmp_path / "junk.gguf"; p.write_bytes(b"GGUF" + bytes(64))
    with pytest.raises(b200.EngineError):
        engine.load_file("bad", str(p))
    with pytest.raises(b200.EngineError):
        engine.load_file("missing", str(tmp_path / "nope.gguf"))


def test_sampling_is_seeded_and_in_vocab(loaded):
    name, path, m = loaded
    prompt = np.array([1, 2, 3, 4], dtype=np.int32)
    a, _, _ = m.generate(prompt, 12, temperature=0.9, top_k=20, top_p=0.9, seed=42)
    b, _, _ = m.generate(prompt, 12, temperature=0.9, top_k=20, top_p=0.9, seed=42)
    c, _, _ = m.generate(prompt, 12, temperature=0.9, top_k=20, top_p=0.9, seed=43)
    assert np.array_equal(a, b) and ((a >= 0) & (a < m.n_vocab)).all()
    assert not np.array_equal(a, c)
    # top_k = 1 is greedy
    gk, _, _ = m.generate(prompt, 12, temperature=0.7, top_k=1, seed=1)
    gg, _, _ = m.generate(prompt, 12)
    assert np.array_equal(gk, gg)


def test_subject_handlers_end_to_end(engine, golden_dir):
    engine.set_models_dir(golden_dir)
    pulled = json.loads(engine.handle("lmstudio.pull_model", json.dumps({"identifier": "tiny-llama"})))
    assert pulled["ok"] and pulled["data"]["model"] == "tiny-llama" and "tensors" in pulled["data"]["output"]
    listed = json.loads(engine.handle("lmstudio.list_models"))
    assert listed["ok"] and listed["data"]["http_status"] == 200
    assert any(e["id"] == "tiny-llama" and e["arch"] == "llama" for e in listed["data"]["models"]["data"])
    req = {"model": "tiny-llama", "messages": [{"role": "system", "content": "be brief"}, {"role": "user", "content": "hi <there>"}],
           "temperature": 0, "max_tokens": 16}
    raw = engine.handle("lmstudio.chat_model", json.dumps(req))
    assert raw.startswith(b'{"ok":true,"data":{"http_status":200,"response":{"id":"chatcmpl-')
    rep = json.loads(raw)
    r = rep["data"]["response"]
    assert r["object"] == "chat.completion" and r["model"] == "tiny-llama"
    assert r["choices"][0]["message"]["role"] == "assistant" and isinstance(r["choices"][0]["message"]["content"], str)
    u = r["usage"]; assert u["total_tokens"] == u["prompt_tokens"] + u["completion_tokens"] and 1 <= u["completion_tokens"] <= 16
    assert r["stats"]["time_to_first_token"] > 0
    # greedy chat is reproducible and equals generate() on the rendered prompt
    rep2 = json.loads(engine.handle("lmstudio.chat_model", json.dumps(req)))
    assert rep2["data"]["response"]["choices"][0]["message"]["content"] == r["choices"][0]["message"]["content"]
    # OpenAI `stop`: the text ends before the first occurrence of a stop string (which may span tokens); usage counts the
    # tokens consumed; finish_reason becomes "stop".  The byte-level test vocabulary makes any 2-byte substring a valid stop.
    full = json.loads(engine.handle("lmstudio.chat_model", json.dumps({**req, "ignore_eos": True})))["data"]["response"]
    text = full["choices"][0]["message"]["content"]
    if len(text) >= 6:
        stop = text[3:5]; cut = text.find(stop)
        for st in (stop, ["\x00never\x00", stop]):
            r2 = json.loads(engine.handle("lmstudio.chat_model", json.dumps({**req, "ignore_eos": True, "stop": st})))["data"]["response"]
            assert r2["choices"][0]["message"]["content"] == text[:cut] and r2["choices"][0]["finish_reason"] == "stop"
            assert r2["usage"]["completion_tokens"] < full["usage"]["completion_tokens"] and r2["stats"]["stop_reason"] == "stopStringFound"
        same = json.loads(engine.handle("lmstudio.chat_model", json.dumps({**req, "ignore_eos": True, "stop": "\x00never\x00"})))["data"]["response"]
        assert same["choices"][0]["message"]["content"] == text and same["choices"][0]["finish_reason"] == full["choices"][0]["finish_reason"]
    unknown = json.loads(engine.handle("lmstudio.chat_model", json.dumps({"model": "ghost", "messages": []})))
    assert unknown["ok"] and unknown["data"]["http_status"] == 404          # status is forwarded, never inspected (:173-178)
    deleted = json.loads(engine.handle("lmstudio.delete_model", json.dumps({"model_id": "tiny-llama"})))
    assert deleted == {"ok": True, "data": {"deleted_dir": "hbm://tiny-llama", "model_id": "tiny-llama"}}
    again = json.loads(engine.handle("lmstudio.delete_model", json.dumps({"model_id": "tiny-llama"})))
    assert again["ok"] is False and again["error"] == "model directory not found: hbm://tiny-llama"


def test_submit_poll(engine, golden_dir):
    import ctypes as C
    from nats_llm_studio_b200 import engine as eng
    m = engine.load_file("q/tiny", os.path.join(golden_dir, "tiny-llama.gguf"))
    L = b200.lib()
    prompt = np.array([5, 6, 7], dtype=np.int32)
    s = eng.Sampling(0.0, 0, 1.0, 0, 8, -1, 1)
    for rid in (101, 102):
        assert L.b200_submit(m.h, prompt.ctypes.data, 3, C.byref(s), rid) == 0
    evs = []
    buf = (eng.Event * 64)()
    while sum(1 for e in evs if e[1] != 1) < 2:
        n = L.b200_poll(buf, 64, 5000)
        assert n > 0
        evs += [(buf[i].req_id, buf[i].kind, buf[i].token, buf[i].n_generated) for i in range(n)]
    toks = {rid: [e[2] for e in evs if e[0] == rid and e[1] == 1] for rid in (101, 102)}
    want, _, _ = m.generate(prompt, 8)
    assert toks[101] == list(want) and toks[102] == list(want)
    m.close()
