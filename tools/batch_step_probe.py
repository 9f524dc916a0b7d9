#!/usr/bin/env python
"""Time b200_model_batch_step alone (B sequences at a given context) — the device-side cost of one batched decode step."""
import argparse, json, os, sys, time
import numpy as np
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT); sys.path.insert(0, os.path.join(ROOT, "tools"))
import bench  # noqa: E402
import nats_llm_studio_b200 as b200  # noqa: E402

ap = argparse.ArgumentParser()
ap.add_argument("--preset", default="llama3-8b"); ap.add_argument("--mix", default="q8_0"); ap.add_argument("--n-layer", type=int, default=0)
ap.add_argument("--batch", type=int, default=64); ap.add_argument("--ctx", type=int, default=300); ap.add_argument("--iters", type=int, default=20)
a = ap.parse_args()
eng = b200.Engine(max_ctx=a.ctx + a.iters + 16, max_seqs=a.batch)
m = eng.load_file("probe", bench.model_path(a.preset, a.mix, a.n_layer, 0))
rng = np.random.default_rng(1)
for s in range(a.batch):
    m.prefill_slot(s, rng.integers(0, m.n_vocab, a.ctx).astype(np.int32))
slots = np.arange(a.batch, dtype=np.int32); pos = np.full(a.batch, a.ctx, np.int32); cur = rng.integers(0, m.n_vocab, a.batch).astype(np.int32)
for _ in range(3):
    cur = m.batch_step(cur, slots, pos); pos += 1
l0 = b200.lib().b200_launch_count(); t0 = time.perf_counter()
for _ in range(a.iters):
    cur = m.batch_step(cur, slots, pos); pos += 1
ms = (time.perf_counter() - t0) * 1e3 / a.iters
print(json.dumps({"mix": a.mix, "n_layer": a.n_layer, "batch": a.batch, "ctx": a.ctx, "step_ms": round(ms, 3), "tok_s": round(a.batch / ms * 1e3, 1),
                  "launches_per_step": (b200.lib().b200_launch_count() - l0) // a.iters}))
