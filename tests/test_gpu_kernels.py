"""GPU parity tests, kernel level: every call goes through the C ABI (libb200engine.so)."""
import os
import zlib

import numpy as np
import pytest

import orc
import nats_llm_studio_b200 as b200

pytestmark = pytest.mark.gpu

TYPES = ["Q8_0", "Q4_K", "Q5_K", "Q6_K"]
BLOCK = {"Q8_0": (32, 34), "Q4_K": (256, 144), "Q5_K": (256, 176), "Q6_K": (256, 210)}
DPOS = {"Q8_0": [0], "Q4_K": [0, 2], "Q5_K": [0, 2], "Q6_K": [208]}


@pytest.fixture(scope="module")
def engine():
    return b200.Engine(max_ctx=2048)


def _same_bits(a, b):
    a, b = np.asarray(a, np.float32), np.asarray(b, np.float32)
    nan = np.isnan(a) & np.isnan(b)
    return bool(np.all((a.view(np.uint32) == b.view(np.uint32)) | nan))


def random_rows(rng, t, n_rows, K, sane=True):
    be, bb = BLOCK[t]
    raw = rng.integers(0, 256, size=(n_rows * (K // be), bb), dtype=np.uint8)
    if sane:
        for p in DPOS[t]:
            d = (0.5 + rng.random(raw.shape[0])).astype(np.float32) * 0.01
            raw[:, p:p + 2] = d.astype(np.float16).view(np.uint8).reshape(-1, 2)
    return raw.reshape(n_rows, -1)


@pytest.mark.parametrize("t", TYPES)
def test_dequant_bit_exact_golden(engine, golden_dir, t):
    """K9 repack + dequant on the GPU is bit-identical to gguf-py on the committed block fixtures
    (includes d=0, inf/nan/subnormal fp16 scales, all-0xFF blocks)."""
    g = np.load(os.path.join(golden_dir, "dequant_blocks.npz"))
    raw, want = g[f"{t}_raw"], g[f"{t}_f32"]
    be, bb = BLOCK[t]
    K = 2048 if t == "Q8_0" else 256 * 8        # 64 blocks -> rows of K elements
    n_rows = raw.shape[0] * be // K
    got = engine.dequantize(t, raw.reshape(n_rows, -1), n_rows, K)
    assert _same_bits(got.reshape(-1), want.reshape(-1))


@pytest.mark.parametrize("t", TYPES)
@pytest.mark.parametrize("shape", [(1, 256), (3, 512), (16, 2048), (7, 4096), (4, 14336)])
def test_dequant_bit_exact_vs_oracle(engine, t, shape):
    """Ragged shapes: odd row counts, K not a multiple of the 2048 group (zero-padded groups)."""
    n_rows, K = shape
    rng = np.random.default_rng(zlib.crc32(repr((t, shape)).encode()))
    raw = random_rows(rng, t, n_rows, K, sane=False)
    want = orc.dequantize(t, raw, n_rows * K).reshape(n_rows, K)
    got = engine.dequantize(t, raw, n_rows, K)
    assert _same_bits(got, want)


def test_dequant_empty_and_bad_args(engine):
    assert engine.dequantize("Q4_K", np.zeros(0, np.uint8), 0, 256).size == 0
    with pytest.raises(b200.EngineError):
        engine.dequantize("Q4_K", np.zeros(144, np.uint8), 1, 100)


@pytest.mark.parametrize("t", TYPES)
@pytest.mark.parametrize("shape", [(2, 256), (64, 512), (1000, 2048), (4096, 4096), (512, 14336), (6000, 2304)])
def test_gemv_matches_oracle(engine, t, shape):
    """Streaming GEMV vs the oracle: same int8 activation rule => tight bound; vs fp32 => stated bound."""
    N, K = shape
    rng = np.random.default_rng(zlib.crc32(repr((t, shape, 1)).encode()))
    raw = random_rows(rng, t, N, K)
    x = rng.standard_normal(K).astype(np.float32)
    y, _ = engine.gemv(t, raw, N, K, x)
    ref_q = orc.matvec(t, raw, N, K, x, act_quant=1)
    ref_f = orc.matvec(t, raw, N, K, x, act_quant=0)
    scale = np.abs(ref_f).max()
    mag = (np.abs(orc.dequantize(t, raw, N * K).reshape(N, K)) @ np.abs(x)).max()      # sum |w||x|: what fp32 ordering noise scales with
    assert np.abs(y - ref_q).max() <= 2e-5 * scale + 1e-6 * mag         # same quantised activations: ordering noise only
    assert np.abs(y - ref_f).max() <= 3e-2 * scale          # int8 activations vs fp32 (ggml does the same)


def test_gemv_linearity_and_zero(engine):
    """Size-independent properties at a full-size shape: W.0 = 0 and W.(2x) = 2 W.x exactly
    (the int8 codes of x and 2x coincide; only the block scale doubles)."""
    N, K = 28672, 4096
    rng = np.random.default_rng(5)
    raw = random_rows(rng, "Q4_K", N, K)
    x = rng.standard_normal(K).astype(np.float32)
    y0, _ = engine.gemv("Q4_K", raw, N, K, np.zeros(K, np.float32))
    assert not y0.any()
    y1, _ = engine.gemv("Q4_K", raw, N, K, x)
    y2, _ = engine.gemv("Q4_K", raw, N, K, 2 * x)
    assert np.array_equal(2 * y1, y2)
    ya, _ = engine.gemv("Q4_K", raw, N, K, x)
    assert np.array_equal(y1, ya)                            # deterministic run to run


def test_gemv_bad_args(engine):
    with pytest.raises(b200.EngineError):
        engine.gemv("Q4_K", np.zeros(144 * 3, np.uint8), 3, 256, np.zeros(256, np.float32))   # odd N
