"""Dequant-in-GEMM kernel of batched decode (qgemm.cuh) through the C ABI: weights stay quantised, are expanded to
fp16 in shared memory and contracted with fp16 activations on the tensor cores (fp32 accumulate)."""
import zlib

import numpy as np
import pytest

import orc
import nats_llm_studio_b200 as b200
from test_gpu_kernels import random_rows, TYPES

pytestmark = pytest.mark.gpu


@pytest.fixture(scope="module")
def engine():
    return b200.Engine(max_ctx=2048)


def _bf16(a):
    u = np.ascontiguousarray(a, np.float32).view(np.uint32)
    u = ((u + 0x7FFF + ((u >> 16) & 1)) >> 16) << 16
    return u.view(np.float32)


@pytest.mark.parametrize("t", TYPES)
def test_one_hot_rows_read_back_every_weight(engine, t):
    """x = unit vectors: y[b][n] is the kernel's fp16 expansion of W[n][k_b]; every K position of every row is
    checked against the bit-exact dequantiser (layout maps, nibble/bit placement, scales), tolerance = fp16 rounding."""
    N, K = 130, 512                                    # two row tiles (second one ragged), two K-steps
    rng = np.random.default_rng(zlib.crc32(t.encode()))
    raw = random_rows(rng, t, N, K)
    W = orc.dequantize(t, raw, N * K).reshape(N, K)
    for k0 in range(0, K, 64):
        x = np.zeros((64, K), np.float32); x[np.arange(64), k0 + np.arange(64)] = 1.0
        y = engine.qgemm(t, raw, N, K, x)
        want = W[:, k0:k0 + 64].T
        err = np.abs(y - want)
        assert np.all(err <= 2.0 ** -8 * np.abs(W).max(axis=1)[None, :]), (t, k0)      # worst case incl. cancellation d*sc*q - dmin*m
        assert np.sqrt((err ** 2).mean()) <= 1e-3 * np.sqrt((W ** 2).mean())


@pytest.mark.parametrize("t", TYPES)
@pytest.mark.parametrize("shape", [(2, 256, 1), (1000, 2048, 5), (4096, 4096, 64), (512, 14336, 64), (6144, 4096, 33), (130, 256, 64)])
def test_qgemm_matches_numpy(engine, t, shape):
    N, K, B = shape
    rng = np.random.default_rng(zlib.crc32(repr((t, shape)).encode()))
    raw = random_rows(rng, t, N, K)
    x = rng.standard_normal((B, K)).astype(np.float32)
    y = engine.qgemm(t, raw, N, K, x)
    W = orc.dequantize(t, raw, N * K).reshape(N, K).astype(np.float64)
    xb = x.astype(np.float16).astype(np.float64)           # the kernel's activation format
    want = xb @ W.T
    mag = np.abs(xb) @ np.abs(W).T                     # what rounding noise scales with
    assert np.isfinite(y).all()
    assert np.all(np.abs(y - want) <= 1e-3 * mag + 1e-6)   # fp16 weight expansion (<= 3 roundings of 2^-11), fp32 accumulate
    assert np.array_equal(y, engine.qgemm(t, raw, N, K, x))  # stream-K partials are summed in a fixed order


def test_qgemm_many_tiles_small_batch(engine):
    """Regression: a matrix with many row tiles (every CTA's K range crosses a tile boundary) and only 1-2 activation rows,
    Q8_0 (largest blobs, shortest weight ring).  The activation box used to hang over the end of a B-row tensor."""
    N, K = 14336, 4096
    rng = np.random.default_rng(21)
    raw = random_rows(rng, "Q8_0", N, K)
    W = orc.dequantize("Q8_0", raw, N * K).reshape(N, K)
    for B in (1, 2):
        x = rng.standard_normal((B, K)).astype(np.float32)
        y = engine.qgemm("Q8_0", raw, N, K, x)
        xh = x.astype(np.float16).astype(np.float32)
        want = xh @ W.T
        mag = np.abs(xh) @ np.abs(W).T
        assert np.all(np.abs(y - want) <= 2e-3 * mag + 1e-6)


def test_qgemm_rows_do_not_depend_on_batch(engine):
    N, K = 1024, 2048
    rng = np.random.default_rng(9)
    raw = random_rows(rng, "Q4_K", N, K)
    x = rng.standard_normal((64, K)).astype(np.float32)
    full = engine.qgemm("Q4_K", raw, N, K, x)
    for B in (1, 7, 32):
        assert np.array_equal(engine.qgemm("Q4_K", raw, N, K, x[:B]), full[:B])


def test_qgemm_argument_checks(engine):
    raw = np.zeros(144 * 2, np.uint8)
    with pytest.raises(b200.EngineError):
        engine.qgemm("Q4_K", raw, 2, 256, np.zeros((65, 256), np.float32))
    with pytest.raises(b200.EngineError):
        engine.qgemm("Q4_K", raw, 2, 128, np.zeros((1, 128), np.float32))
