"""Parity at the BASELINE.json configurations themselves (not toy sizes), through the C ABI, against the CPU oracle:
  * configs[1]  Llama-3-8B Q4_K_M — the file bench.py times: teacher-forced logits of the batch-1 kernel and of the
                batched row path at 8 positions;
  * configs[2]  Llama-3-8B Q8_0 — the same two paths;
  * configs[0]  granite-3.0-2b Q8_0, 16-token prompt, 64 greedy tokens (the reference's own CPU-runnable case).
The GGUFs are the synthetic random-init files of tools/make_gguf.py (there is no network for real checkpoints); they are
written once per box under /dev/shm and shared with bench.py.  The oracle costs ~1 s per 8B token on 16 cores.

Tolerances (relative to the largest |logit| of the position):
  TOL_QUANT 4e-2  granite-2b (tied Q8_0) vs the plain fp32 oracle; TOL_INT8 6e-2 the batch-1 kernel on the 8B files (measured
                  3.4e-2 Q4_K_M, 4.2e-2 Q8_0: the oracle's own int8 restatement measures 3.3e-2 / 4.0e-2, see below);
  TOL_FP16  1e-2  the batched row path (fp16 weights and activations, fp32 accumulate) vs the fp32 oracle — BASELINE's figure;
  the batch-1 kernel quantises activations to int8 per 256-block the way ggml does before a K-quant mat-vec: at 32 layers
  x 4096 columns that alone moves the logits by 2-3e-2 of the largest logit (measured below on the oracle's own int8
  restatement, orc act_quant=1), so the kernel is additionally held to "no further from fp32 than 1.5x the oracle's int8
  restatement + 5e-3".  Bit-level agreement with that restatement (TOL_EXACT in test_gpu_model.py) exists only on small
  models: at this depth the first flipped int8 code (a 1e-7 difference flips one with probability ~1e-5 per element)
  cascades, and two correct implementations end up one quantisation-noise apart.  Measured numbers: BASELINE.md."""
import os
import sys

import numpy as np
import pytest

import orc
import nats_llm_studio_b200 as b200
from nats_llm_studio_b200.engine import lib

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
pytestmark = pytest.mark.gpu
TOL_QUANT, TOL_INT8, TOL_FP16 = 4e-2, 6e-2, 1e-2
N_POS = 8


def _cores():
    n = len(os.sched_getaffinity(0)) if hasattr(os, "sched_getaffinity") else (os.cpu_count() or 1)
    orc.lib().orc_set_threads(n)
    return n


@pytest.fixture(scope="module")
def engine():
    lib().b200_shutdown()
    e = b200.Engine(max_ctx=256, max_seqs=4)
    yield e
    lib().b200_shutdown()


@pytest.mark.parametrize("mix", ["q4_k_m", "q8_0"])
def test_llama3_8b_logits_match_oracle(engine, mix, capsys):
    import bench
    _cores()
    path = bench.model_path("llama3-8b", mix, 0, 0)
    m = engine.load_file(f"full/8b-{mix}", path)
    toks = np.random.default_rng(21).integers(0, m.n_vocab, N_POS).astype(np.int32)
    ref32 = orc.Model(path).forward(toks)                                   # fp32 oracle, [N_POS][V]
    scale = np.abs(ref32).max(axis=1)
    # ---- batch-1 path: the persistent decode kernel, teacher-forced
    got = m.eval(toks)
    assert np.isfinite(got).all()
    d32 = np.abs(got - ref32).max(axis=1) / scale
    ref_eng = orc.Model(path, act_quant=1, kv_f16=1).forward(toks)          # the oracle's restatement of the int8-activation numerics
    d_eng = np.abs(got - ref_eng).max(axis=1) / scale                      # engine vs that restatement
    d_int8 = np.abs(ref_eng - ref32).max(axis=1) / scale                   # what int8 activations cost by themselves
    # ---- batched row path: prompt ingestion by the tensor-core prefill, last position by one batched decode step
    m.prefill_slot(1, toks[:-1])
    nxt, lgb = m.batch_step(toks[-1:], np.array([1], np.int32), np.array([N_POS - 1], np.int32), want_logits=True)
    db = float(np.abs(lgb[0] - ref32[-1]).max() / scale[-1])
    with capsys.disabled():
        print(f"\n[fullsize] llama3-8b {mix}, max |dlogit| / max |logit| over {N_POS} positions: bs1 kernel vs fp32 oracle {d32.max():.4f} (mean {d32.mean():.4f}); "
              f"oracle int8 restatement vs fp32 oracle {d_int8.max():.4f} (mean {d_int8.mean():.4f}); bs1 kernel vs int8 restatement {d_eng.max():.4f}; batched fp16 path vs fp32 oracle {db:.4f}")
    assert d32.max() <= TOL_INT8
    assert d32.max() <= 1.5 * d_int8.max() + 5e-3                           # all of the distance is the stated activation rounding
    assert db <= TOL_FP16
    srt = np.sort(ref32[-1])
    if srt[-1] - srt[-2] > 2 * TOL_QUANT * scale[-1]:                       # a clear winner must be found by both paths
        assert int(got[-1].argmax()) == int(ref32[-1].argmax()) == int(nxt[0])
    m.close()


def test_granite3_2b_greedy_64_tokens(engine, capsys):
    """configs[0]: 16-token prompt, 64 greedy tokens; the engine's stream equals the CPU oracle's, token for token; a
    difference is tolerated only at a position where the oracle's own top-2 margin is inside the logit tolerance."""
    import make_gguf
    _cores()
    d = os.environ.get("B200_BENCH_DIR") or "/dev/shm"
    path = os.path.join(d, "b200-config1-granite3-2b-q8_0.gguf")
    if not os.path.exists(path):
        make_gguf.write_model(path + ".tmp", make_gguf.spec_from_preset("granite3-2b", seed=1)); os.replace(path + ".tmp", path)
    prompt = np.random.default_rng(1).integers(0, 49155, 16).astype(np.int32)
    want, lg = orc.Model(path, act_quant=1, kv_f16=1).generate(prompt, 64, want_logits=True)
    m = engine.load_file("full/granite", path)
    got, _, _ = m.generate(prompt, 64)
    same = int(np.argmax(np.concatenate([(got != want).astype(int), [1]])))
    with capsys.disabled():
        print(f"\n[fullsize] granite3-2b q8_0: {same} of 64 greedy tokens equal to the oracle's")
    if same < 64:
        srt = np.sort(lg[same]); assert srt[-1] - srt[-2] <= TOL_QUANT * np.abs(lg[same]).max(), f"token {same}: {got[same]} != {want[same]}"
    lg0 = m.eval(prompt, all_logits=False)
    ref0 = orc.Model(path).forward(prompt)[-1]
    assert np.abs(lg0 - ref0).max() <= TOL_QUANT * np.abs(ref0).max()
    m.close()
