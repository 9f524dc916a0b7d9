#!/usr/bin/env python
"""Config 4: Llama-3-70B Q4_K_M, tensor-parallel decode at bs1 and bs32, 512-token context.

    python -m torch.distributed.run --nnodes=1 --nproc-per-node N --master-addr 127.0.0.1 tools/bench_tp.py [--preset llama3-70b]

One rank per GPU; every rank keeps its shard (heads / FFN columns / vocab rows), all ranks call prefill_slot /
batch_step in lockstep; NCCL all-reduce after the two row-parallel products of a layer.  Rank 0 prints one JSON line:
tok/s = batch / (max over ranks of the mean step time), roofline fraction from the algorithmic bytes one GPU reads per step.
The GGUF is synthetic (random-init); with --reuse-layers (default for 70B) every layer repeats the first layer's bytes so
that the 42.5 GB file is written in seconds — same size, same types, same HBM traffic."""
import argparse, json, os, sys, time
import numpy as np
import torch
import torch.distributed as dist
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT); sys.path.insert(0, os.path.join(ROOT, "tools"))
import make_gguf  # noqa: E402
import nats_llm_studio_b200 as b200  # noqa: E402

ap = argparse.ArgumentParser()
ap.add_argument("--preset", default="llama3-70b"); ap.add_argument("--mix", default="q4_k_m"); ap.add_argument("--n-layer", type=int, default=0)
ap.add_argument("--ctx", type=int, default=512); ap.add_argument("--iters", type=int, default=20); ap.add_argument("--batches", default="1,32")
a = ap.parse_args()
rank, world, local = int(os.environ.get("RANK", 0)), int(os.environ.get("WORLD_SIZE", 1)), int(os.environ.get("LOCAL_RANK", 0))
dist.init_process_group("gloo")
d = os.environ.get("B200_BENCH_DIR") or "/dev/shm"
path = os.path.join(d, f"b200tp-{a.preset}-{a.mix}" + (f"-L{a.n_layer}" if a.n_layer else "") + ".gguf")
t0 = time.perf_counter()
if rank == 0 and not os.path.exists(path):
    over = dict(mix=a.mix, seed=2, reuse_layers=True)
    if a.n_layer:
        over["n_layer"] = a.n_layer
    make_gguf.write_model(path + ".tmp", make_gguf.spec_from_preset(a.preset, **over)); os.replace(path + ".tmp", path)
dist.barrier()
t_file = time.perf_counter() - t0
batches = [int(x) for x in a.batches.split(",")]
eng = b200.Engine(device=local, max_ctx=a.ctx + 3 * a.iters + 16, max_seqs=max(batches))
ids = [eng.tp_unique_id() if rank == 0 and world > 1 else None]
dist.broadcast_object_list(ids, src=0)
eng.tp_init(rank, world, ids[0])
t0 = time.perf_counter()
m = eng.load_file("tp/bench", path)
t_load = time.perf_counter() - t0
info = m.info()
rng = np.random.default_rng(5)
out = {"workload": f"{a.preset} {a.mix}" + (f" L{a.n_layer}" if a.n_layer else "") + f", tensor-parallel x{world}, ctx {a.ctx}", "n_gpus": world,
       "file_bytes": os.path.getsize(path), "file_s": round(t_file, 1), "load_s": round(t_load, 1), "shard_weight_bytes": info["weight_bytes"], "results": {}}
peak = 6548.5
try:
    peak = float(json.load(open(os.path.join(ROOT, "MEASURED_PEAKS.json"))).get("hbm_gbs", peak))
except Exception:
    pass
kv_per_pos_shard = info.get("kv_bytes_per_pos", 0)
for B in batches:
    prompt = rng.integers(0, m.n_vocab, a.ctx).astype(np.int32)
    for s in range(B):
        m.prefill_slot(s, prompt)
    slots = np.arange(B, dtype=np.int32); pos = np.full(B, a.ctx, np.int32); cur = rng.integers(0, m.n_vocab, B).astype(np.int32)
    for _ in range(3):
        cur = m.batch_step(cur, slots, pos); pos += 1
    dist.barrier(); torch.cuda.synchronize(local)
    t0 = time.perf_counter()
    for _ in range(a.iters):
        cur = m.batch_step(cur, slots, pos); pos += 1
    ms = (time.perf_counter() - t0) * 1e3 / a.iters
    t = torch.tensor([ms]); dist.all_reduce(t, op=dist.ReduceOp.MAX); ms = float(t[0])
    # algorithmic bytes one GPU reads per step: its weight shard (minus the replicated embedding table, plus B embedding rows) + its KV shard
    wb = info["weight_bytes"] - info.get("embd_bytes", 0)
    kvb = B * (a.ctx + a.iters // 2) * kv_per_pos_shard
    out["results"][f"bs{B}"] = {"step_ms": round(ms, 3), "tok_s": round(B / ms * 1e3, 1), "gbs_per_gpu": round((wb + kvb) / ms / 1e6, 1),
                                "roofline_frac_per_gpu": round((wb + kvb) / ms / 1e6 / peak, 3)}
if rank == 0:
    print(json.dumps(out), flush=True)
m.close(); eng.shutdown(); dist.destroy_process_group()
