#!/usr/bin/env python
"""Config 3: N concurrent requests through b200_submit / b200_poll with continuous batching.

Prompt lengths U[64,512], 128 generated each; prints one JSON line with aggregate tok/s (generated tokens of all
requests / wall time from the first submit to the last FINISH), p50/p90 TTFT, and the device-side step time of a
full batch (b200_model_batch_step at B = n_req).  A reporting tool next to bench.py, not the headline bench."""
import argparse
import json
import os
import sys
import time

import numpy as np

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT); sys.path.insert(0, os.path.join(ROOT, "tools"))
import bench  # noqa: E402  (model_path)
import nats_llm_studio_b200 as b200  # noqa: E402


def measure(eng, m, n_req=64, gen=128, prompt_min=64, prompt_max=512, step_iters=20, seed=7, peak=None, step_ctx=512):
    """n_req concurrent requests through b200_submit / b200_poll (continuous batching) + the bare batched-step time."""
    rng = np.random.default_rng(seed)
    lens = rng.integers(prompt_min, prompt_max + 1, n_req)
    prompts = [rng.integers(0, m.n_vocab, n).astype(np.int32) for n in lens]
    # warm-up: one request end to end (builds the tile-layout / bf16 weight copies), then a few full-batch steps
    m.submit(prompts[0], 1, max_tokens=4)
    done = False
    while not done:
        for e in eng.poll(64, 60000):
            if e.kind == b200.EV_ERROR:
                raise SystemExit(f"warm-up request failed: {e.finish_reason}")
            done |= e.kind == b200.EV_FINISH
    # the bare step, device-timed: CUDA events on the engine stream around step_iters graph replays, nothing crossing PCIe
    l0 = int(b200.lib().b200_launch_count())
    step_ms, step_bytes = m.bench_batch_step(n_req, step_ctx, step_iters, 3)
    launches_per_step = (int(b200.lib().b200_launch_count()) - l0) // (step_iters + 3)
    roofline = None
    if peak:
        ach = step_bytes / (step_ms * 1e-3) / 1e9
        roofline = {"bound": "hbm", "achieved": round(ach, 1), "peak": peak, "unit": "GB/s", "frac": round(ach / peak, 4), "traffic": None,
                    "algorithmic_bytes_per_step": step_bytes, "kernel": "batched decode step (qgemm_kernel + row kernels, one CUDA graph)",
                    "timing": f"CUDA events around {step_iters} steps on the engine stream"}

    t0 = time.perf_counter()
    for i, p in enumerate(prompts):
        m.submit(p, 1000 + i, max_tokens=gen)
    first, fin, n_tok = {}, {}, 0
    while len(fin) < n_req:
        evs = eng.poll(256, 20000)
        now = time.perf_counter()
        if not evs:
            raise SystemExit("scheduler stalled")
        for e in evs:
            if e.kind == b200.EV_TOKEN:
                n_tok += 1; first.setdefault(e.req_id, now - t0)
            elif e.kind == b200.EV_FINISH:
                fin[e.req_id] = (now - t0, e.ttft_ms, e.n_generated)
            else:
                raise SystemExit(f"request {e.req_id} failed: {e.finish_reason}")
    wall = max(v[0] for v in fin.values())
    ttft = np.array(sorted(first.values())) * 1e3
    return {"requests": n_req, "prompt_len": f"U[{prompt_min},{prompt_max}]", "generated_each": gen,
            "aggregate_tok_s": round(n_tok / wall, 1), "wall_s": round(wall, 3), "generated": n_tok, "prompt_tokens": int(lens.sum()),
            "ttft_ms_p50": round(float(np.percentile(ttft, 50)), 1), "ttft_ms_p90": round(float(np.percentile(ttft, 90)), 1), "ttft_ms_max": round(float(ttft.max()), 1),
            "batch_step_ms": round(step_ms, 3), "batch_step_tok_s": round(n_req / step_ms * 1e3, 1), "batch_step_ctx": step_ctx, "launches_per_step": launches_per_step,
            "roofline": roofline}


def measure_nats(eng, model_id, n_vocab, n_req=64, gen=128, prompt_min=64, prompt_max=512, seed=7):
    """Config 3 over the wire: n_req concurrent lmstudio.chat_model requests published to a NATS queue group (stub broker,
    tools/nats_stub.py; the worker is csrc/worker.cpp in this process), continuous batching behind it.  Prompt lengths
    U[prompt_min, prompt_max] tokens (byte-level vocabulary: one token per character), `gen` tokens each, greedy."""
    import threading
    import nats_stub
    rng = np.random.default_rng(seed)
    lens = rng.integers(prompt_min, prompt_max + 1, n_req)
    reqs = [json.dumps({"model": model_id, "messages": [{"role": "user", "content": "".join(chr(97 + int(c)) for c in rng.integers(0, 26, max(1, int(n) - 4)))}],
                        "temperature": 0, "max_tokens": gen, "ignore_eos": True}).encode() for n in lens]
    broker = nats_stub.Broker().start()
    t = threading.Thread(target=lambda: b200.Engine.worker_run(broker.url, None, 0), daemon=True); t.start()
    for _ in range(500):
        with broker.lock:
            if broker.subs.get("lmstudio.chat_model"):
                break
        time.sleep(0.01)
    clients = [nats_stub.Client(broker.url) for _ in range(8)]
    clients[0].request("lmstudio.chat_model", reqs[0][:0] + json.dumps({"model": model_id, "messages": [{"role": "user", "content": "warm"}], "temperature": 0, "max_tokens": 4, "ignore_eos": True}).encode(), timeout=300)
    out = [None] * n_req; lat = [0.0] * n_req
    def work(i):
        t0 = time.perf_counter(); out[i] = clients[i % 8].request("lmstudio.chat_model", reqs[i], timeout=600); lat[i] = time.perf_counter() - t0
    th = [threading.Thread(target=work, args=(i,)) for i in range(n_req)]
    t0 = time.perf_counter()
    [x.start() for x in th]; [x.join() for x in th]
    wall = time.perf_counter() - t0
    n_tok = 0; ttft = []; h2d = 0; d2h = 0
    for i, o in enumerate(out):
        r = json.loads(o)
        if not r["ok"] or r["data"]["http_status"] != 200:
            raise SystemExit(f"chat over NATS failed: {o[:200]}")
        n_tok += r["data"]["response"]["usage"]["completion_tokens"]; ttft.append(r["data"]["response"]["stats"]["time_to_first_token"] * 1e3)
        h2d += len(reqs[i]) + 4 * r["data"]["response"]["usage"]["prompt_tokens"]; d2h += len(o)
    for c in clients:
        c.close()
    b200.Engine.worker_stop(); t.join(timeout=20); broker.stop()
    return {"requests": n_req, "generated": n_tok, "wall_s": round(wall, 3), "aggregate_tok_s": round(n_tok / wall, 1), "ttft_ms_p50": round(float(np.percentile(ttft, 50)), 1),
            "ttft_ms_p90": round(float(np.percentile(ttft, 90)), 1), "latency_s_p50": round(float(np.percentile(lat, 50)), 3), "h2d_bytes": h2d, "d2h_bytes": d2h,
            "path": "client -> NATS (stub broker) -> queue group -> C++ worker -> lmstudio.chat_model handler -> continuous-batching scheduler"}


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--preset", default="llama3-8b"); ap.add_argument("--mix", default="q8_0"); ap.add_argument("--n-layer", type=int, default=0)
    ap.add_argument("--n-req", type=int, default=64); ap.add_argument("--gen", type=int, default=128)
    ap.add_argument("--prompt-min", type=int, default=64); ap.add_argument("--prompt-max", type=int, default=512)
    ap.add_argument("--step-iters", type=int, default=20)
    a = ap.parse_args()
    path = bench.model_path(a.preset, a.mix, a.n_layer, 0)
    eng = b200.Engine(max_ctx=a.prompt_max + a.gen + 8, max_seqs=a.n_req)
    m = eng.load_file("bench/batch", path)
    r = measure(eng, m, a.n_req, a.gen, a.prompt_min, a.prompt_max, a.step_iters)
    r["workload"] = f"{a.preset} {a.mix}" + (f" L{a.n_layer}" if a.n_layer else "") + ", continuous batching"
    print(json.dumps(r))
    m.close()


if __name__ == "__main__":
    main()
