"""Prompt ingestion path (tcgen05 GEMM + row kernels): GEMM vs numpy, prefill vs the token-by-token path and the oracle."""
import os

import numpy as np
import pytest

import orc
import nats_llm_studio_b200 as b200

pytestmark = pytest.mark.gpu


@pytest.fixture(scope="module")
def engine():
    return b200.Engine(max_ctx=2048)


def _bf16(a):
    u = np.ascontiguousarray(a, np.float32).view(np.uint32)
    u = ((u + 0x7FFF + ((u >> 16) & 1)) >> 16) << 16          # round to nearest even
    return u.view(np.float32)


# the last four shapes have enough tiles for the persistent kernels: 128 x 256 tiles (ragged M and N edges, several tiles per
# CTA so both accumulators and the ring wrap), 128 x 128 tiles, and the 8B gate/up shape
@pytest.mark.parametrize("shape", [(128, 128, 64), (128, 128, 256), (256, 384, 512), (100, 256, 4096), (511, 640, 1024), (1, 128, 64),
                                   (511, 19000, 192), (300, 12800, 128), (250, 4104, 320), (512, 28672, 4096)])
def test_gemm_tcgen05_matches_numpy(engine, shape):
    M, N, K = shape
    rng = np.random.default_rng(M * 7 + N * 3 + K)
    A = rng.standard_normal((M, K)).astype(np.float32); B = rng.standard_normal((N, K)).astype(np.float32)
    want = _bf16(A).astype(np.float64) @ _bf16(B).astype(np.float64).T
    got = engine.gemm_bf16(A, B)
    # identical bf16 operands, fp32 accumulation: only summation order differs
    assert np.abs(got - want).max() <= 2e-5 * np.sqrt(K) * np.abs(want).max() + 1e-4


def test_gemm_identity_and_zero(engine):
    K = 128
    A = np.eye(K, dtype=np.float32)[:128]; B = np.arange(256 * K, dtype=np.float32).reshape(256, K) % 13 - 6
    assert np.array_equal(engine.gemm_bf16(A, B), B[:, :128].T)           # small integers are exact in bf16
    assert not engine.gemm_bf16(np.zeros((64, K), np.float32), B).any()
    with pytest.raises(b200.EngineError):
        engine.gemm_bf16(np.zeros((8, 100), np.float32), np.zeros((8, 100), np.float32))


@pytest.mark.parametrize("name", ["tiny-llama", "tiny-granite"])
def test_prefill_matches_sequential_and_oracle(engine, golden_dir, name):
    """A 40-token prompt ingested by the batched path: the logits of the next position agree with the token-by-token
    path and with the fp32 oracle within the quantisation tolerance (bf16 GEMMs vs int8-activation GEMVs)."""
    path = os.path.join(golden_dir, f"{name}.gguf")
    m = engine.load_file(f"pf/{name}", path)
    rng = np.random.default_rng(5)
    prompt = rng.integers(0, m.n_vocab, 41).astype(np.int32)
    m.set_prefill_min(0)
    lg_seq = m.eval(prompt, all_logits=False)
    seq_tok, _, _ = m.generate(prompt, 6)
    m.set_prefill_min(8)
    lg_pf = m.eval(prompt, all_logits=False)
    pf_tok, _, _ = m.generate(prompt, 6)
    o = orc.Model(path)                                                   # fp32 oracle
    want, lg = o.generate(prompt, 6, want_logits=True)
    ref = np.abs(lg[0]).max()
    assert np.isfinite(lg_pf).all()
    assert np.abs(lg_pf - lg[0]).max() <= 4e-2 * ref                      # vs fp32 oracle (TOL_QUANT of test_gpu_model.py)
    assert np.abs(lg_pf - lg_seq).max() <= 4e-2 * ref                     # vs the int8-activation token-by-token path
    assert np.abs(lg_pf - lg_seq).max() > 0                               # the two paths really are different code
    srt = np.sort(lg[0]); margin = srt[-1] - srt[-2]
    scale = np.abs(lg[0]).max()
    if margin > 8e-2 * scale:                                             # unambiguous argmax: every path must agree
        assert pf_tok[0] == want[0] == seq_tok[0]
    assert len(pf_tok) == 6 and ((pf_tok >= 0) & (pf_tok < m.n_vocab)).all()
    m.close()
