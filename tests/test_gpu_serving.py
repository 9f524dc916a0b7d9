"""Serving behaviour on the GPU, through the C ABI: the sampler (K8) against its numpy restatement, sampling inside
batched steps, lmstudio.chat_model requests sharing the continuous-batching scheduler
(/root/reference/nats_llm_studio.go:327-364 served concurrently), the handler deadline (:328), and a model deleted
while requests still hold it (the reference answers a delete during chat with a clean error)."""
import json
import os
import threading

import numpy as np
import pytest

import sampler_ref
import nats_llm_studio_b200 as b200
from nats_llm_studio_b200.engine import lib

pytestmark = pytest.mark.gpu
MAX_SEQS = 16


@pytest.fixture(scope="module")
def engine():
    lib().b200_shutdown()
    e = b200.Engine(max_ctx=1024, max_seqs=MAX_SEQS)
    yield e
    lib().b200_shutdown()


def _logit_sets():
    rng = np.random.default_rng(5)
    sets = {"normal515": rng.standard_normal(515).astype(np.float32) * 2,
            "peaked2048": np.concatenate([rng.standard_normal(2040), rng.standard_normal(8) + 9]).astype(np.float32),
            "vocab128k": (rng.standard_normal(128256) * 3).astype(np.float32)}
    t = rng.standard_normal(700).astype(np.float32); t[100:140] = t[7]            # ties: order falls back to the index
    sets["ties700"] = t
    return sets


@pytest.mark.parametrize("name", ["normal515", "peaked2048", "vocab128k", "ties700"])
def test_sampler_matches_numpy_restatement(engine, name):
    """Every stage has an observable effect on the pick: top-k cut, nucleus cut at T=1, temperature order, Philox draw.
    Draws whose uniform lands within 1e-5 of a cumulative boundary are skipped (fp32 scan vs float64 cumsum)."""
    lg = _logit_sets()[name]
    checked = 0
    for top_k, top_p, temp in [(0, 1.0, 1.0), (40, 0.95, 0.8), (5, 1.0, 1.5), (0, 0.5, 0.7), (1, 1.0, 1.0), (2000, 0.9, 1.2), (40, 0.1, 0.5), (3, 0.6, 2.0)]:
        for seed in range(6):
            for step in (0, 1, 77):
                want, kept, probs, u = sampler_ref.sample(lg, temp, top_k, top_p, seed, step, return_dist=True)
                cum = np.cumsum(probs)
                if np.abs(cum - u).min() < 1e-5:
                    continue
                got = engine.sample(lg, temp, top_k, top_p, seed, step)
                assert got == want, (name, top_k, top_p, temp, seed, step, got, want)
                checked += 1
    assert checked > 100


def test_sampler_distribution(engine):
    """4000 draws (steps 0..3999 of one seed) from a 12-way distribution: empirical frequencies follow
    softmax(logits / T) restricted to the nucleus — a biased draw or a wrong temperature order would not."""
    lg = np.array([2.0, 1.5, 1.0, 0.5, 0.0, -0.5, -1.0, -1.5, -2.0, -2.5, -3.0, -8.0], np.float32)
    temp, top_p = 0.7, 0.9
    _, kept, probs, _ = sampler_ref.sample(lg, temp, 0, top_p, 0, 0, return_dist=True)
    n = 4000
    counts = np.zeros(len(lg))
    for step in range(n):
        counts[engine.sample(lg, temp, 0, top_p, 1234, step)] += 1
    assert counts[[i for i in range(len(lg)) if i not in set(kept.tolist())]].sum() == 0      # nothing outside the nucleus
    exp = probs * n
    chi2 = ((counts[kept] - exp) ** 2 / exp).sum()
    assert chi2 < 40, (chi2, counts, exp)                   # dof <= 11: p(chi2 > 40) < 4e-5


def _collect(engine, ids):
    got = {i: [] for i in ids}; fin = {}
    while len(fin) < len(ids):
        evs = engine.poll(64, 5000)
        assert evs, "scheduler made no progress"
        for e in evs:
            assert e.kind != b200.EV_ERROR, e.finish_reason
            if e.kind == b200.EV_TOKEN:
                got[e.req_id].append(e.token)
            else:
                fin[e.req_id] = e
    return got, fin


def test_sampled_request_is_the_same_alone_and_in_a_batch(engine, golden_dir):
    """temperature > 0 through the batched path: the tokens are a function of (prompt, sampling fields, seed) only."""
    m = engine.load_file("sv/sampled", os.path.join(golden_dir, "tiny-llama.gguf"))
    rng = np.random.default_rng(8)
    p = rng.integers(0, m.n_vocab, 9).astype(np.int32)
    kw = dict(max_tokens=12, temperature=0.9, top_k=40, top_p=0.95, seed=77)
    m.submit(p, 1, **kw)
    alone, _ = _collect(engine, [1])
    others = [rng.integers(0, m.n_vocab, n).astype(np.int32) for n in (3, 14, 6, 21)]
    m.submit(others[0], 10, max_tokens=20, temperature=1.3, seed=5)
    m.submit(others[1], 11, max_tokens=7)
    m.submit(p, 2, **kw)
    m.submit(others[2], 12, max_tokens=15, temperature=0.5, top_k=3, seed=6)
    m.submit(others[3], 13, max_tokens=9)
    got, _ = _collect(engine, [2, 10, 11, 12, 13])
    assert got[2] == alone[1]
    m.submit(p, 3, **{**kw, "seed": 78})                   # another seed gives another stream (flat random-init logits)
    other_seed, _ = _collect(engine, [3])
    assert other_seed[3] != alone[1]
    m.submit(p, 4, max_tokens=12)                          # greedy rows are untouched by the sampling kernel
    greedy, _ = _collect(engine, [4])
    assert greedy[4] != alone[1]
    m.close()


def _chat(engine, model, content, **kw):
    req = {"model": model, "messages": [{"role": "user", "content": content}], "temperature": 0, "max_tokens": 8, "ignore_eos": True, **kw}
    return json.loads(engine.handle("lmstudio.chat_model", json.dumps(req)))


def test_concurrent_chat_requests_share_the_scheduler(engine, golden_dir):
    """64 concurrent lmstudio.chat_model calls (config 3's load shape) against 16 sequence slots: every envelope equals
    the one the same request gets when it is alone; a seeded sampled chat reproduces under load."""
    m = engine.load_file("sv/chat", os.path.join(golden_dir, "tiny-llama.gguf"))
    texts = [("q%d " % i) * (1 + i % 7) for i in range(64)]
    extra = [dict(temperature=0.8, seed=1000 + i, top_k=40) if i % 4 == 3 else {} for i in range(64)]
    _chat(engine, "sv/chat", "warm")                       # one-time weight-layout builds happen here
    l0 = engine.launch_count()
    solo = [_chat(engine, "sv/chat", t, **x) for t, x in zip(texts, extra)]
    launches_one_by_one = engine.launch_count() - l0
    out = [None] * 64
    def work(i):
        out[i] = _chat(engine, "sv/chat", texts[i], **extra[i])
    th = [threading.Thread(target=work, args=(i,)) for i in range(64)]
    l0 = engine.launch_count()
    [t.start() for t in th]; [t.join() for t in th]
    launches = engine.launch_count() - l0
    for i in range(64):
        r = out[i]
        assert r["ok"] and r["data"]["http_status"] == 200, r
        resp = r["data"]["response"]
        assert resp["usage"]["completion_tokens"] == 8 and resp["choices"][0]["finish_reason"] == "length"
        assert resp["choices"][0]["message"]["content"] == solo[i]["data"]["response"]["choices"][0]["message"]["content"], i
        assert resp["usage"] == solo[i]["data"]["response"]["usage"]
    # the concurrent requests shared batched steps: far fewer kernel launches than the same 64 requests one by one
    assert launches < 0.7 * launches_one_by_one, (launches, launches_one_by_one)
    m.close()


def test_chat_deadline_answers_ok_false(engine, golden_dir):
    m = engine.load_file("sv/deadline", os.path.join(golden_dir, "tiny-llama.gguf"))
    os.environ["B200_CHAT_TIMEOUT_MS"] = "1"
    try:
        r = _chat(engine, "sv/deadline", "hello", max_tokens=900)
    finally:
        del os.environ["B200_CHAT_TIMEOUT_MS"]
    assert r["ok"] is False and r["error"] == "error calling LM Studio: context deadline exceeded" and r["data"] == {"http_status": 0}
    assert _chat(engine, "sv/deadline", "hello")["ok"]    # the engine keeps serving
    m.submit(np.arange(5, dtype=np.int32), 71, max_tokens=900, timeout_ms=30)
    got, fin = _collect(engine, [71])
    assert fin[71].finish_reason == b200.FINISH_TIMEOUT and fin[71].n_generated < 900
    m.close()


def test_model_deleted_while_requests_hold_it(engine, golden_dir):
    """delete_model / b200_model_close while the scheduler is decoding that model: the in-flight requests end with an
    ERROR (or a FINISH if they were already done), nothing touches freed memory, and the engine keeps serving."""
    path = os.path.join(golden_dir, "tiny-llama.gguf")
    m = engine.load_file("sv/victim", path)
    for i in range(6):
        m.submit(np.arange(3 + i, dtype=np.int32), 200 + i, max_tokens=900)
    seen = 0
    while seen < 12:                                       # decoding is under way
        seen += sum(1 for e in engine.poll(64, 5000) if e.kind == b200.EV_TOKEN)
    rep = json.loads(engine.handle("lmstudio.delete_model", json.dumps({"model_id": "sv/victim"})))
    assert rep["ok"], rep
    ended = {}
    while len(ended) < 6:
        evs = engine.poll(64, 5000)
        assert evs, "requests of the deleted model never ended"
        for e in evs:
            if e.kind != b200.EV_TOKEN:
                ended[e.req_id] = e
    assert any(e.kind == b200.EV_ERROR and e.finish_reason == -7 for e in ended.values())      # B200_ERR_STATE
    m.h = None
    r = _chat(engine, "sv/victim", "hi")
    assert r["ok"] and r["data"]["http_status"] == 404     # gone from the registry; the reply is LM Studio's 404 shape
    m2 = engine.load_file("sv/after", path)                # no sticky CUDA fault: a fresh model loads and generates
    assert _chat(engine, "sv/after", "hi")["ok"]
    toks, _, _ = m2.generate(np.arange(4, dtype=np.int32), 4)
    assert len(toks) == 4
    m2.close()


def test_sync_path_and_scheduler_do_not_share_kv_pages(engine, golden_dir):
    """b200_model_generate (batch-1 kernel, its own sequence slot) while the scheduler decodes other requests of the same
    model, including one in slot 0: both streams equal their solo runs."""
    m = engine.load_file("sv/mixed", os.path.join(golden_dir, "tiny-llama.gguf"))
    rng = np.random.default_rng(31)
    pa, pb = rng.integers(0, m.n_vocab, 11).astype(np.int32), rng.integers(0, m.n_vocab, 7).astype(np.int32)
    solo_sync, _, _ = m.generate(pa, 24)
    m.submit(pb, 300, max_tokens=40)
    solo_sched, _ = _collect(engine, [300])
    m.submit(pb, 301, max_tokens=40)
    res = {}
    def work():
        res["sync"], _, _ = m.generate(pa, 24)
    t = threading.Thread(target=work); t.start()
    got, _ = _collect(engine, [301]); t.join()
    assert np.array_equal(res["sync"], solo_sync)
    assert got[301] == solo_sched[300]
    m.close()
