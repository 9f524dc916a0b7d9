"""Kernel-level GEMV microbench through the C ABI (used under ncu): python tools/gemv_bench.py TYPE N K ITERS"""
import sys
import numpy as np
sys.path.insert(0, '.')
import nats_llm_studio_b200 as b200
t, N, K, iters = sys.argv[1], int(sys.argv[2]), int(sys.argv[3]), int(sys.argv[4])
bb = {"Q4_K": 144, "Q5_K": 176, "Q6_K": 210, "Q8_0": 34}[t]; be = 32 if t == "Q8_0" else 256
e = b200.Engine(max_ctx=1024)
rng = np.random.default_rng(0)
raw = rng.integers(0, 256, size=(N * (K // be), bb), dtype=np.uint8)
x = rng.standard_normal(K).astype(np.float32)
y, ms = e.gemv(t, raw.reshape(N, -1), N, K, x, iters=iters)
print(f"{t} {N}x{K}: {ms*1e3:.2f} us/iter  {raw.nbytes/ms/1e6:.0f} GB/s")
