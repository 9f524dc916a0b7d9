"""b200_dequantize_noalloc: the dequantise entry point for callers that cannot allocate device memory (a worker whose
HBM is fully committed to models and KV pages, or code on the request path where cudaMalloc/cudaFree would stall the
device).  Its scratch is a fixed array inside the library, so this file checks, through the C ABI, that it returns
exactly what b200_dequantize returns — twice in a row with different shapes, so bytes left in the array by the first
call cannot leak into the second — and that oversize requests are refused.  Collected last on purpose."""
import ctypes as C

import numpy as np
import pytest

import nats_llm_studio_b200 as b200
from nats_llm_studio_b200.engine import GGML_TYPES

pytestmark = pytest.mark.gpu

BLOCK = {"Q8_0": (32, 34), "Q4_K": (256, 144), "Q5_K": (256, 176), "Q6_K": (256, 210)}


def _noalloc(type_name, raw, n_rows, K):
    import torch
    d_raw = torch.from_numpy(np.ascontiguousarray(raw, dtype=np.uint8).reshape(-1)).cuda()
    d_out = torch.full((n_rows * K,), float("nan"), dtype=torch.float32, device="cuda")
    torch.cuda.synchronize()                       # the engine uses its own stream: inputs must be complete
    rc = b200.lib().b200_dequantize_noalloc(GGML_TYPES[type_name], C.c_void_p(d_raw.data_ptr()), n_rows, K, C.c_void_p(d_out.data_ptr()))
    if rc < 0:
        raise b200.EngineError(rc, b200.lib().b200_last_error().decode())
    return d_out.cpu().numpy().reshape(n_rows, K)


def _same_bits(a, b):
    a, b = np.asarray(a, np.float32), np.asarray(b, np.float32)
    return bool(np.all((a.view(np.uint32) == b.view(np.uint32)) | (np.isnan(a) & np.isnan(b))))


def test_noalloc_matches_allocating_entry_point_twice():
    eng = b200.Engine(max_ctx=2048)                # b200_init (idempotent); the no-alloc call itself creates nothing
    rng = np.random.default_rng(2024)
    # two calls in a row, second smaller and of another type: stale bytes of the first call stay in the array
    for t, n_rows, K in (("Q6_K", 37, 4096), ("Q4_K", 5, 512), ("Q8_0", 64, 2304), ("Q5_K", 2, 256)):
        be, bb = BLOCK[t]
        raw = rng.integers(0, 256, size=(n_rows, K // be * bb), dtype=np.uint8)
        want = eng.dequantize(t, raw, n_rows, K)
        got = _noalloc(t, raw, n_rows, K)
        assert _same_bits(got, want), (t, n_rows, K)      # bit-exact, like the other dequant tests
        assert _same_bits(_noalloc(t, raw, n_rows, K), want)


def test_noalloc_rejects_oversize_and_bad_args():
    import torch
    b200.Engine(max_ctx=2048)
    L = b200.lib()
    dummy = torch.zeros(16, dtype=torch.uint8, device="cuda")
    # Q8_0 4096 x 4096 needs 17.8 MB of engine-layout scratch: beyond the 4 MiB built-in buffer
    rc = L.b200_dequantize_noalloc(GGML_TYPES["Q8_0"], C.c_void_p(dummy.data_ptr()), 4096, 4096, C.c_void_p(dummy.data_ptr()))
    assert rc == -1 and b"built-in buffer holds" in L.b200_last_error()
    assert L.b200_dequantize_noalloc(GGML_TYPES["Q4_K"], C.c_void_p(dummy.data_ptr()), 2, 100, C.c_void_p(dummy.data_ptr())) == -1
    assert L.b200_dequantize_noalloc(GGML_TYPES["Q4_K"], None, 2, 256, C.c_void_p(dummy.data_ptr())) == -1
    assert L.b200_dequantize_noalloc(GGML_TYPES["Q4_K"], None, 0, 256, None) == 0      # empty input: nothing to do
