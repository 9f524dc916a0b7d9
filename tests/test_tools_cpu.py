"""Host-side tooling the measurements depend on (no GPU): the synthetic GGUF writer's fast mode and the bench's CPU arm."""
import os
import sys

import numpy as np

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT); sys.path.insert(0, os.path.join(ROOT, "tools"))
import make_gguf  # noqa: E402


def test_reuse_layers_writes_the_same_file_shape(tmp_path):
    """reuse_layers repeats the first layer's bytes (per tensor and quant type): same size, same tensor table, same header
    values as the fully random file — only the random bytes of layers > 0 differ."""
    from gguf import GGUFReader
    a, b = str(tmp_path / "a.gguf"), str(tmp_path / "b.gguf")
    make_gguf.write_model(a, make_gguf.spec_from_preset("tiny-llama", n_layer=4, seed=5))
    make_gguf.write_model(b, make_gguf.spec_from_preset("tiny-llama", n_layer=4, seed=5, reuse_layers=True))
    assert os.path.getsize(a) == os.path.getsize(b)
    ra, rb = GGUFReader(a), GGUFReader(b)
    ta = [(t.name, t.tensor_type, tuple(t.shape), t.n_bytes) for t in ra.tensors]
    tb = [(t.name, t.tensor_type, tuple(t.shape), t.n_bytes) for t in rb.tensors]
    assert ta == tb
    byname = {t.name: t for t in rb.tensors}
    assert np.array_equal(byname["blk.0.attn_q.weight"].data, byname["blk.3.attn_q.weight"].data)      # repeated bytes
    an = {t.name: t for t in ra.tensors}
    assert np.array_equal(an["blk.0.attn_q.weight"].data, byname["blk.0.attn_q.weight"].data)          # layer 0 is the same draw


def test_cpu_arm_ignores_omp_num_threads(monkeypatch, golden_dir):
    """torchrun exports OMP_NUM_THREADS=1; the reference arm must still use every core it may run on."""
    import bench
    monkeypatch.setenv("OMP_NUM_THREADS", "1")
    r = bench.cpu_reference_arm(os.path.join(golden_dir, "tiny-llama.gguf"), 4, 4)
    want = len(os.sched_getaffinity(0)) if hasattr(os, "sched_getaffinity") else os.cpu_count()
    assert r["cores"] == want and r["value"] > 0 and r["kind"] == "port"


def test_sampler_restatement_known_answers():
    """tests/sampler_ref.py (the sampler's checker): Philox4x32-10 known-answer vector (Random123 kat_vectors:
    counter 0, key 0), top_k=1 is the argmax, and the nucleus never keeps less than one candidate."""
    import numpy as np
    import sampler_ref
    assert sampler_ref.philox4x32_10(0, 0) == [0x6627E8D5, 0xE169C58D, 0xBC57AC4C, 0x9B00DBD8]
    lg = np.random.default_rng(0).standard_normal(300).astype(np.float32)
    for seed in range(5):
        assert sampler_ref.sample(lg, 0.8, top_k=1, seed=seed) == int(lg.argmax())
        assert sampler_ref.sample(lg, 0.8, top_p=1e-9, seed=seed) == int(lg.argmax())
        t, kept, probs, u = sampler_ref.sample(lg, 1.3, 40, 0.9, seed, 3, return_dist=True)
        assert t in kept and len(kept) <= 40 and abs(probs.sum() - 1) < 1e-9
