"""Batched decode (continuous batching, BASELINE.json configs[2]): one dense step advances every active sequence.
Checks: a row's result is independent of its batch (bit-exact B=1 vs B=k), the batched logits agree with the oracle
within the quantisation tolerance, and the b200_submit scheduler in continuous-batching mode returns, per request,
exactly what the same request returns when it runs alone."""
import os

import numpy as np
import pytest

import orc
import nats_llm_studio_b200 as b200
from nats_llm_studio_b200.engine import lib

pytestmark = pytest.mark.gpu
MAX_SEQS = 8
TOL_QUANT = 4e-2          # same bound as tests/test_gpu_model.py: bf16 weights/activations vs the fp32 oracle


@pytest.fixture(scope="module")
def engine():
    lib().b200_shutdown()                                  # this module needs max_seqs > 1: re-initialise the engine
    e = b200.Engine(max_ctx=1024, max_seqs=MAX_SEQS)
    yield e
    lib().b200_shutdown()                                  # later modules initialise with their own config


def _prompts(n_vocab, lens, seed=11):
    rng = np.random.default_rng(seed)
    return [rng.integers(0, n_vocab, n).astype(np.int32) for n in lens]


def _run(m, prompts, slots, n_gen, want_logits=False):
    """prefill each prompt into its slot, then step all of them together; returns tokens [B][n_gen] (and first logits)."""
    for p, s in zip(prompts, slots):
        m.prefill_slot(s, p[:-1])
    cur = np.array([p[-1] for p in prompts], np.int32); pos = np.array([len(p) - 1 for p in prompts], np.int32)
    out = np.zeros((len(prompts), n_gen), np.int32); first = None
    for t in range(n_gen):
        if want_logits and t == 0:
            cur, first = m.batch_step(cur, slots, pos, want_logits=True)
        else:
            cur = m.batch_step(cur, slots, pos)
        out[:, t] = cur; pos += 1
    return (out, first) if want_logits else out


@pytest.mark.parametrize("name", ["tiny-llama", "tiny-granite"])
def test_batch_rows_are_independent_and_match_oracle(engine, golden_dir, name):
    path = os.path.join(golden_dir, f"{name}.gguf")
    m = engine.load_file(f"bt/{name}", path)
    prompts = _prompts(m.n_vocab, [5, 17, 40, 1, 33])
    together, lg = _run(m, prompts, np.array([3, 0, 7, 1, 4], np.int32), 6, want_logits=True)
    for i, p in enumerate(prompts):                        # each sequence alone, in a different slot: identical tokens
        alone, lg1 = _run(m, [p], np.array([(i * 3 + 2) % MAX_SEQS], np.int32), 6, want_logits=True)
        assert np.array_equal(alone[0], together[i])
        assert np.array_equal(lg1[0], lg[i])               # bit-exact: batch composition does not change a row
    o = orc.Model(path)
    for i, p in enumerate(prompts):
        want, ref = o.generate(p, 1, want_logits=True)
        scale = np.abs(ref[0]).max()
        assert np.abs(lg[i] - ref[0]).max() <= TOL_QUANT * scale
        srt = np.sort(ref[0])
        if srt[-1] - srt[-2] > 2 * TOL_QUANT * scale:
            assert together[i, 0] == want[0]
    m.close()


def test_batch_step_argument_checks(engine, golden_dir):
    m = engine.load_file("bt/args", os.path.join(golden_dir, "tiny-llama.gguf"))
    one = np.zeros(1, np.int32)
    with pytest.raises(b200.EngineError):
        m.batch_step(np.array([m.n_vocab], np.int32), one, one)
    with pytest.raises(b200.EngineError):
        m.batch_step(one, np.array([MAX_SEQS], np.int32), one)
    with pytest.raises(b200.EngineError):
        m.batch_step(np.zeros(MAX_SEQS + 1, np.int32), np.zeros(MAX_SEQS + 1, np.int32), np.zeros(MAX_SEQS + 1, np.int32))
    with pytest.raises(b200.EngineError):
        m.prefill_slot(MAX_SEQS, one)
    m.close()


def test_scheduler_continuous_batching(engine, golden_dir):
    """More requests than slots, different prompt lengths and max_tokens: every request gets exactly the tokens it gets
    when it is the only sequence in the batch, in order, followed by one FINISH event."""
    m = engine.load_file("bt/sched", os.path.join(golden_dir, "tiny-llama.gguf"))
    lens = [3, 9, 30, 1, 12, 7, 21, 5, 16, 2, 11, 4, 19]   # 13 requests > 8 slots
    prompts = _prompts(m.n_vocab, lens, seed=3)
    n_gen = [4 + (i * 5) % 9 for i in range(len(prompts))]
    want = [_run(m, [p], np.array([0], np.int32), g)[0] for p, g in zip(prompts, n_gen)]
    for i, (p, g) in enumerate(zip(prompts, n_gen)):
        m.submit(p, 100 + i, max_tokens=g)
    got = {100 + i: [] for i in range(len(prompts))}; fin = {}
    while len(fin) < len(prompts):
        evs = engine.poll(64, 2000)
        assert evs, "scheduler made no progress"
        for e in evs:
            assert e.kind != b200.EV_ERROR, e.finish_reason
            if e.kind == b200.EV_TOKEN:
                assert e.req_id not in fin
                got[e.req_id].append(e.token); assert e.n_generated == len(got[e.req_id])
            else:
                fin[e.req_id] = e
    for i in range(len(prompts)):
        assert np.array_equal(np.array(got[100 + i], np.int32), want[i]), i
        f = fin[100 + i]
        assert f.finish_reason == b200.FINISH_LENGTH and f.n_generated == n_gen[i] and f.n_prompt == lens[i] and f.ttft_ms > 0
    m.close()


def test_scheduler_prompts_straddle_prefill_chunks(engine, golden_dir):
    """Prompts longer than one 512-row chunk, packed back to back: a request's prompt is split across chunks shared with
    other requests; its tokens must still equal the solo run (prefill numerics do not depend on the packing)."""
    m = engine.load_file("bt/straddle", os.path.join(golden_dir, "tiny-llama.gguf"))
    lens = [600, 300, 450, 20, 700]
    prompts = _prompts(m.n_vocab, lens, seed=17)
    want = [_run(m, [p], np.array([0], np.int32), 5)[0] for p in prompts]
    for i, p in enumerate(prompts):
        m.submit(p, 500 + i, max_tokens=5)
    got = {500 + i: [] for i in range(len(prompts))}; fin = set()
    while len(fin) < len(prompts):
        evs = engine.poll(64, 5000)
        assert evs, "scheduler made no progress"
        for e in evs:
            assert e.kind != b200.EV_ERROR, e.finish_reason
            if e.kind == b200.EV_TOKEN:
                got[e.req_id].append(e.token)
            else:
                fin.add(e.req_id)
    for i in range(len(prompts)):
        assert np.array_equal(np.array(got[500 + i], np.int32), want[i]), i
    m.close()


def test_cancel_frees_the_slot_and_the_others_finish(engine, golden_dir):
    """b200_cancel on a running request: it ends with FINISH/CANCELLED before its token budget, the other requests are
    unaffected (same tokens as alone), and its sequence slot is reused by a request submitted afterwards."""
    m = engine.load_file("bt/cancel", os.path.join(golden_dir, "tiny-llama.gguf"))
    prompts = _prompts(m.n_vocab, [12, 9, 20], seed=23)
    want1 = _run(m, [prompts[1]], np.array([0], np.int32), 6)[0]
    m.submit(prompts[0], 900, max_tokens=900)              # would run for a long time
    m.submit(prompts[1], 901, max_tokens=6)
    got, fin = {900: [], 901: [], 902: []}, {}
    cancelled = False
    while len(fin) < 2:
        evs = engine.poll(64, 5000)
        assert evs, "scheduler made no progress"
        for e in evs:
            assert e.kind != b200.EV_ERROR, e.finish_reason
            if e.kind == b200.EV_TOKEN:
                got[e.req_id].append(e.token)
                if e.req_id == 900 and len(got[900]) >= 3 and not cancelled:
                    assert engine.cancel(900) == 0; cancelled = True
            else:
                fin[e.req_id] = e
    assert fin[900].finish_reason == b200.FINISH_CANCELLED and 3 <= fin[900].n_generated < 900
    assert fin[901].finish_reason == b200.FINISH_LENGTH and np.array_equal(np.array(got[901], np.int32), want1)
    assert engine.cancel(12345) != 0                       # unknown request id
    m.submit(prompts[2], 902, max_tokens=3)                # the engine keeps serving
    while 902 not in fin:
        for e in engine.poll(64, 5000):
            if e.kind == b200.EV_FINISH:
                fin[e.req_id] = e
    assert fin[902].n_generated == 3
    m.close()


def test_tensor_core_decode_attention_matches_row_kernel_and_oracle(engine, tmp_path):
    """attn_decode_tc_kernel (csrc/attn_tc.cuh: head size 128, operands swapped): decode steps of sequences whose contexts
    end in different 128-position KV tiles (first tile only, tile boundaries 127/128/129, three tiles) agree with the
    CUDA-core row kernel (B200_NO_ATTN_TC_DEC=1: same GEMMs around it) and with the fp32 oracle; rows stay independent of
    the batch they are in."""
    import make_gguf
    path = str(tmp_path / "small.gguf"); make_gguf.write_model(path, make_gguf.spec_from_preset("small-llama"))
    m = engine.load_file("bt/tcdec", path)
    lens = [3, 128, 129, 130, 300, 17, 257]
    prompts = _prompts(m.n_vocab, lens, seed=23); slots = np.array([5, 0, 2, 7, 1, 3, 6], np.int32)
    tok_tc, lg_tc = _run(m, prompts, slots, 3, want_logits=True)
    alone, lg1 = _run(m, [prompts[4]], np.array([4], np.int32), 3, want_logits=True)
    assert np.array_equal(alone[0], tok_tc[4]) and np.array_equal(lg1[0], lg_tc[4])     # a row does not depend on its batch
    os.environ["B200_NO_ATTN_TC_DEC"] = "1"               # read when a step is launched or captured: a second instance of the
    try:                                                   # model, whose step graphs are captured with the row kernel
        m2 = engine.load_file("bt/tcdec-cc", path)
        tok_cc, lg_cc = _run(m2, prompts, slots, 3, want_logits=True)
        m2.close()
    finally:
        del os.environ["B200_NO_ATTN_TC_DEC"]
    assert np.abs(lg_tc - lg_cc).max() > 0                                # the two kernels really are different code
    o = orc.Model(path)
    for i, p in enumerate(prompts):
        ref = o.forward(p, all_logits=False); scale = np.abs(ref).max()
        d_tc, d_cc, d_x = np.abs(lg_tc[i] - ref).max() / scale, np.abs(lg_cc[i] - ref).max() / scale, np.abs(lg_tc[i] - lg_cc[i]).max() / scale
        assert np.isfinite(lg_tc[i]).all()
        assert d_tc <= TOL_QUANT and d_tc <= 1.15 * d_cc + 2e-3, (lens[i], d_tc, d_cc, d_x)     # no further from the oracle than the fp32-probability kernel
        assert d_x <= 2.5e-2, (lens[i], d_x)
    m.close()
