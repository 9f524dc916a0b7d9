"""KV page allocator (K5, paged KV): pages of 16 positions are handed to sequence slots on demand from a pool that may be
smaller than slots x max_ctx, returned when a request ends, and every attention kernel reads them through the block table.
(The default pool backs every slot fully; its physical page order is the reverse of the logical one, so every other GPU
test already runs through a non-identity table.)"""
import os

import numpy as np
import pytest

import nats_llm_studio_b200 as b200
from nats_llm_studio_b200.engine import lib

pytestmark = pytest.mark.gpu
MAX_SEQS, MAX_CTX = 8, 256                      # 16 pages per sequence, 9 slots (8 + the synchronous path's): 144 table entries
POOL = 40                                       # pages in the pool: 640 positions for all sequences together


@pytest.fixture(scope="module")
def engine():
    lib().b200_shutdown()
    e = b200.Engine(max_ctx=MAX_CTX, max_seqs=MAX_SEQS, kv_pages=POOL)
    yield e
    lib().b200_shutdown()


def _collect(engine, ids, allow_error=False):
    got = {i: [] for i in ids}; fin = {}
    while len(fin) < len(ids):
        evs = engine.poll(64, 5000)
        assert evs, "scheduler made no progress"
        for e in evs:
            if e.kind == b200.EV_TOKEN:
                got[e.req_id].append(e.token)
            else:
                assert allow_error or e.kind != b200.EV_ERROR, e.finish_reason
                fin[e.req_id] = e
    return got, fin


def test_pages_are_allocated_on_demand_and_returned(engine, golden_dir):
    m = engine.load_file("kv/a", os.path.join(golden_dir, "tiny-llama.gguf"))
    assert m.kv_free_pages() == POOL
    rng = np.random.default_rng(1)
    p = rng.integers(0, m.n_vocab, 40).astype(np.int32)
    m.prefill_slot(2, p[:-1])                                            # positions 0..38: 3 pages
    assert m.kv_free_pages() == POOL - 3
    cur = np.array([p[-1]], np.int32); pos = np.array([39], np.int32); toks_a = []
    for _ in range(12):                                                 # positions 39..50: crosses into a 4th page at 48
        cur = m.batch_step(cur, np.array([2], np.int32), pos); toks_a.append(int(cur[0])); pos += 1
    assert m.kv_free_pages() == POOL - 4
    # the same sequence in another slot, while slot 2 still holds its pages: other physical pages, identical tokens
    m.prefill_slot(5, p[:-1])
    cur = np.array([p[-1]], np.int32); pos = np.array([39], np.int32); toks_b = []
    for _ in range(12):
        cur = m.batch_step(cur, np.array([5], np.int32), pos); toks_b.append(int(cur[0])); pos += 1
    assert toks_a == toks_b and m.kv_free_pages() == POOL - 8
    m.release_slot(2); m.release_slot(5)
    assert m.kv_free_pages() == POOL
    # pool exhaustion is an error, not a write through a stale table entry: three sequences of 240 positions want 45 pages
    long = rng.integers(0, m.n_vocab, 240).astype(np.int32)
    m.prefill_slot(0, long); m.prefill_slot(1, long)
    with pytest.raises(b200.EngineError) as ei:
        m.prefill_slot(3, long)
    assert ei.value.code == -3                                           # B200_ERR_NOMEM
    for s in (0, 1, 3):
        m.release_slot(s)
    assert m.kv_free_pages() == POOL
    m.close()


def test_scheduler_admits_by_free_pages(engine, golden_dir):
    """12 requests of 3 + 100 positions (7 pages each) against a pool of 40 pages: at most 5 run at once, the others wait for
    pages; every request finishes with the tokens it produces alone, and the pool is whole again afterwards.  A request that
    can never fit fails with B200_ERR_NOMEM."""
    m = engine.load_file("kv/b", os.path.join(golden_dir, "tiny-llama.gguf"))
    prompts = [np.arange(3, dtype=np.int32) + 7 * i for i in range(12)]
    solo = []
    for i, p in enumerate(prompts):
        m.submit(p, 900 + i, max_tokens=100, ignore_eos=True)
        got, fin = _collect(engine, [900 + i]); solo.append(got[900 + i])
        assert fin[900 + i].n_generated == 100
    assert m.kv_free_pages() == POOL
    for i, p in enumerate(prompts):
        m.submit(p, 1000 + i, max_tokens=100, ignore_eos=True)
    got, fin = _collect(engine, [1000 + i for i in range(12)])
    for i in range(12):
        assert got[1000 + i] == solo[i], i
    assert m.kv_free_pages() == POOL
    # the synchronous path shares the pool (its own slot): pages taken by generate() stay with that slot for its next call
    toks, _, _ = m.generate(np.arange(5, dtype=np.int32), 8)
    assert len(toks) == 8 and m.kv_free_pages() == POOL - 1
    # a request larger than the whole pool: 250 positions = 16 pages fits; make the pool too small for it by holding pages
    for s in range(5):
        m.prefill_slot(s, np.zeros(96, np.int32))                       # 5 x 6 = 30 pages held outside the scheduler
    m.submit(np.arange(3, dtype=np.int32), 1100, max_tokens=240, ignore_eos=True)       # wants 16 pages, 9 are free, nothing is running
    got, fin = _collect(engine, [1100], allow_error=True)
    assert fin[1100].kind == b200.EV_ERROR and fin[1100].finish_reason == -3
    for s in range(5):
        m.release_slot(s)
    m.close()
