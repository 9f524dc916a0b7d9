#!/usr/bin/env python
"""Time prompt ingestion (prefill_slot of one 512-token prompt) — for the ncu launch list of the prefill path."""
import argparse, json, os, sys, time
import numpy as np
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT); sys.path.insert(0, os.path.join(ROOT, "tools"))
import bench  # noqa: E402
import nats_llm_studio_b200 as b200  # noqa: E402
ap = argparse.ArgumentParser()
ap.add_argument("--preset", default="llama3-8b"); ap.add_argument("--mix", default="q4_k_m"); ap.add_argument("--n-layer", type=int, default=0)
ap.add_argument("--tokens", type=int, default=512); ap.add_argument("--iters", type=int, default=3); ap.add_argument("--seqs", type=int, default=1)
a = ap.parse_args()
eng = b200.Engine(max_ctx=a.tokens + 64, max_seqs=a.seqs)
m = eng.load_file("probe", bench.model_path(a.preset, a.mix, a.n_layer, 0))
p = np.random.default_rng(1).integers(0, m.n_vocab, a.tokens).astype(np.int32)
m.prefill_slot(0, p)
t0 = time.perf_counter()
for _ in range(a.iters):
    m.prefill_slot(0, p)
ms = (time.perf_counter() - t0) * 1e3 / a.iters
print(json.dumps({"tokens": a.tokens, "n_layer": a.n_layer, "prefill_ms": round(ms, 2), "tok_s": round(a.tokens / ms * 1e3, 1)}))
