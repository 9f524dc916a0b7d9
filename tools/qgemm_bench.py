#!/usr/bin/env python
"""Time the dequant-in-GEMM kernel on the 8B matrix shapes (B = 64 rows): GB/s of quantised weight bytes."""
import os, sys
import numpy as np
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
import nats_llm_studio_b200 as b200
BB = {"Q8_0": (32, 34), "Q4_K": (256, 144), "Q5_K": (256, 176), "Q6_K": (256, 210)}
eng = b200.Engine()
rng = np.random.default_rng(0)
shapes = [("gate_up", 28672, 4096), ("down", 4096, 14336), ("qkv_q", 4096, 4096), ("lm_head", 128256, 4096)]
types = sys.argv[1:] or ["Q4_K", "Q6_K", "Q8_0"]
for t in types:
    be, bb = BB[t]
    for name, N, K in shapes:
        raw = rng.integers(0, 256, size=(N, K // be * bb), dtype=np.uint8)
        x = rng.standard_normal((64, K)).astype(np.float32)
        _, ms = eng.qgemm(t, raw, N, K, x, iters=20)
        print(f"{t:5s} {name:8s} N={N:6d} K={K:5d}  {ms*1e3:8.1f} us  {raw.size/ms/1e6:8.1f} GB/s", flush=True)
