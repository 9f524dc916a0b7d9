#!/usr/bin/env python3
"""bench.py — headline metric of BASELINE.json: batch-1 decode tokens/s, Llama-3-8B Q4_K_M,
512-token context, on B200 (configs[1]).

  python bench.py [--gpus N] [--steps K] [--warmup W] [--impl b200|reference]

A "step" is one decoded token (one pass of the hot path: all 32 layers + lm_head + argmax) at batch 1.
* value        device-timed (CUDA events on the engine's stream) tokens/s with the context already
               resident in HBM; N>1 = N independent replicas (request-level data parallel, what the
               reference's NATS queue group does), aggregate tokens/s, weak scaling.
* e2e          the same work through the reference-facing call b200_handle("lmstudio.chat_model")
               with HOST request/reply buffers: JSON in, tokenise, prompt ingestion, decode,
               detokenise, envelope out; completion tokens / wall seconds.
* roofline     algorithmic HBM bytes per token (all matrices but token_embd + one
               embedding row + norms + fp16 KV at the current context) x tokens per launch / launch time,
               against MEASURED_PEAKS.json hbm_gbs.
* cpu_baseline the CPU oracle (oracle/oracle.c, OpenMP, all host cores) on a bounded sample of the
               same model.  --impl reference prints that arm alone (the reference's own runtime,
               LM Studio/llama.cpp, is unavailable offline; the oracle port stands in, kind "port").
"""
import argparse
import json
import os
import subprocess
import sys
import threading
import time

ROOT = os.path.dirname(os.path.abspath(__file__))
sys.path.insert(0, ROOT)
sys.path.insert(0, os.path.join(ROOT, "tools"))
sys.path.insert(0, os.path.join(ROOT, "tests"))

METRIC = "decode_tokens_per_s_bs1_llama3_8b_q4_k_m"
UNIT = "tok/s"


def model_path(preset, mix, n_layer, rank):
    import make_gguf
    d = os.environ.get("B200_BENCH_DIR") or ("/dev/shm" if os.path.isdir("/dev/shm") and os.access("/dev/shm", os.W_OK) else "/tmp")
    tag = f"{preset}-{mix}" + (f"-L{n_layer}" if n_layer else "")
    path = os.path.join(d, f"b200bench-{tag}.gguf")
    if not os.path.exists(path):
        if rank == 0:
            over = dict(mix=mix, seed=2)
            if n_layer:
                over["n_layer"] = n_layer
            tmp = path + f".tmp{os.getpid()}"
            make_gguf.write_model(tmp, make_gguf.spec_from_preset(preset, **over))
            os.replace(tmp, path)
        else:
            while not os.path.exists(path):
                time.sleep(0.5)
    return path


class ClockSampler:
    """nvidia-smi clocks + throttle reasons during the timed region (B200_PROFILING.md recipe)."""
    Q = "clocks.sm,clocks.max.sm,power.draw,clocks_event_reasons.hw_slowdown,clocks_event_reasons.hw_thermal_slowdown,clocks_event_reasons.sw_thermal_slowdown,clocks_event_reasons.sw_power_cap"

    def __init__(self, index):
        self.rows = []; self.proc = None; self.index = index

    def start(self):
        try:
            self.proc = subprocess.Popen(["nvidia-smi", f"--id={self.index}", f"--query-gpu={self.Q}", "--format=csv,noheader,nounits", "-lms", "100"],
                                         stdout=subprocess.PIPE, stderr=subprocess.DEVNULL, text=True)
            threading.Thread(target=self._read, daemon=True).start()
        except OSError:
            self.proc = None

    def _read(self):
        for line in self.proc.stdout:
            self.rows.append([c.strip() for c in line.split(",")])

    def stop(self):
        if self.proc:
            self.proc.terminate()
            try:
                self.proc.wait(timeout=2)
            except Exception:
                self.proc.kill()
        sm = sorted(int(float(r[0])) for r in self.rows if r and r[0].replace(".", "").isdigit())
        mx = [int(float(r[1])) for r in self.rows if len(r) > 1 and r[1].replace(".", "").isdigit()]
        names = ["hw_slowdown", "hw_thermal_slowdown", "sw_thermal_slowdown", "sw_power_cap"]
        reasons = sorted({names[i] for r in self.rows if len(r) >= 7 for i in range(4) if r[3 + i].lower().startswith("active")})
        return {"sm_mhz": sm[len(sm) // 2] if sm else None, "sm_max_mhz": max(mx) if mx else None, "reasons": reasons, "samples": len(sm)}


def reduce_max(value, dist=None, device="cpu"):
    """Max over ranks of a per-rank time (the multi-GPU contract: never the rank-0 clock alone)."""
    if dist is None:
        return float(value)
    import torch
    t = torch.tensor([float(value)], dtype=torch.float64, device=device)
    dist.all_reduce(t, op=dist.ReduceOp.MAX)
    return float(t.item())


def aggregate_value(world, steps, ms_total_max):
    """N independent replicas, each decoding `steps` tokens: whole-job tokens/s over the slowest rank's time."""
    return world * steps / (ms_total_max * 1e-3)


REF_MAX_STEPS = 48          # cap of the CPU arm: one step is one decoded token at ~1 s on 16 cores


def cpu_reference_arm(path, warmup, steps):
    """Times the CPU oracle (the stand-in for the reference's runtime, which is not installable offline) decoding `steps`
    tokens one by one after `warmup` untimed ones: a step is one pass of the hot path at batch 1, like the GPU arm's,
    but at a context of a few tokens instead of 512 (the CPU pass is weight-bound: 4.6 GB of quantised weights
    dequantised row by row with fp64 dot products; 512 cached positions would add <1 % to it)."""
    import numpy as np
    import orc
    m = orc.Model(path)
    cores = len(os.sched_getaffinity(0)) if hasattr(os, "sched_getaffinity") else (os.cpu_count() or 1)
    orc.lib().orc_set_threads(cores)            # torchrun exports OMP_NUM_THREADS=1: use every host core the process may run on
    rng = np.random.default_rng(3)
    tok = int(rng.integers(0, m.n_vocab))
    for _ in range(warmup):
        tok = int(np.argmax(m.eval_token(tok)))
    t0 = time.perf_counter()
    for _ in range(steps):
        tok = int(np.argmax(m.eval_token(tok)))
    dt = time.perf_counter() - t0
    m.close()
    return {"value": steps / dt, "unit": UNIT, "cores": cores, "kind": "port",
            "sample": f"oracle/oracle.c (a plain fp32/fp64-accumulate restatement, OpenMP over rows, not a tuned CPU runtime): {steps} greedy tokens of the same GGUF decoded one by one after {warmup} warm-up tokens, context {warmup}..{warmup + steps}, {dt:.1f} s"}


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--gpus", type=int, default=1)
    ap.add_argument("--steps", type=int, default=256)
    ap.add_argument("--warmup", type=int, default=8)
    ap.add_argument("--impl", default="b200", choices=["b200", "reference"])
    ap.add_argument("--preset", default="llama3-8b")
    ap.add_argument("--mix", default="q4_k_m")
    ap.add_argument("--n-layer", type=int, default=0, help="debug: fewer layers (number is then NOT the headline config)")
    ap.add_argument("--ctx", type=int, default=512)
    ap.add_argument("--no-cpu", action="store_true")
    ap.add_argument("--no-e2e", action="store_true")
    ap.add_argument("--no-batched", action="store_true", help="skip the bs64 / continuous-batching measurements (N=1 only)")
    ap.add_argument("--no-tp", action="store_true", help="skip the Llama-3-70B tensor-parallel measurements (config 4)")
    ap.add_argument("--tp-preset", default="llama3-70b")
    ap.add_argument("--tp-layers", type=int, default=0, help="debug: fewer layers in the tensor-parallel model")
    a = ap.parse_args()
    rank = int(os.environ.get("RANK", "0")); world = int(os.environ.get("WORLD_SIZE", "1")); local = int(os.environ.get("LOCAL_RANK", "0"))
    W = max(a.warmup, 3); K = a.steps
    workload = f"{a.preset} {a.mix} GGUF (random-init), bs1 decode, ctx {a.ctx}+{K}" + (f", {a.n_layer} layers (debug)" if a.n_layer else "")
    config = {"workload": workload, "prompt_tokens": a.ctx, "generated_tokens": K, "parallelism": f"replica x{world}",
              "l2": "weights (4.6 GB/token) >> 126 MB L2: no flush needed"}

    if a.impl == "reference":
        if rank != 0:
            return
        path = model_path(a.preset, a.mix, a.n_layer, 0)
        Kr, Wr = min(K, REF_MAX_STEPS), min(W, 8)           # honoured up to the cap (a CPU token costs ~1 s); the line says what ran
        cb = cpu_reference_arm(path, Wr, Kr)
        config = dict(config, reference_sample=f"{Kr} timed + {Wr} warm-up tokens" + (f" (capped from --steps {K})" if Kr != K else ""))
        K, W = Kr, Wr
        line = {"metric": METRIC, "value": cb["value"], "unit": UNIT, "n_gpus": a.gpus, "steps": K, "warmup": W, "ms_per_step": 1e3 / cb["value"],
                "higher_is_better": True, "scaling": "weak", "vs_baseline": None, "dtype": "f32", "data": "synthetic", "impl": "reference",
                "config": config, "cpu_baseline": cb, "e2e": {"value": cb["value"], "unit": UNIT, "h2d_bytes_per_step": 0, "d2h_bytes_per_step": 0},
                "gpu_launches": 0, "note": "LM Studio / llama.cpp (the reference's runtime) is unavailable on this box; CPU oracle port timed instead"}
        print(json.dumps(line)); return

    import numpy as np
    import torch
    import nats_llm_studio_b200 as b200

    dist = None
    if world > 1:
        import torch.distributed as dist
        torch.cuda.set_device(local)
        dist.init_process_group("nccl", device_id=torch.device("cuda", local))
    path = model_path(a.preset, a.mix, a.n_layer, rank)
    if dist:
        dist.barrier()
    eng = b200.Engine(device=local, max_ctx=max(2048, a.ctx + K + W + 64))
    t0 = time.perf_counter()
    m = eng.load_file(f"bench/{a.preset}-{a.mix}", path)
    load_s = time.perf_counter() - t0
    info = m.info()

    def barrier():
        if dist:
            dist.barrier()
        torch.cuda.synchronize()

    # ---- device-timed decode (value, roofline)
    clocks = ClockSampler(local); clocks.start()
    barrier()
    l0 = eng.launch_count()
    ms_tok, bytes_tok = m.bench_decode(a.ctx, K, W)      # W warm-up steps, then exactly K timed steps (CUDA events on the engine stream)
    l1 = eng.launch_count()
    barrier()
    clk = clocks.stop()
    ms_total = ms_tok * K
    ms_total = reduce_max(ms_total, dist, "cuda")
    value = aggregate_value(world, K, ms_total)
    peaks = {}
    try:
        peaks = json.load(open(os.path.join(ROOT, "MEASURED_PEAKS.json")))
    except Exception:
        pass
    peak = float(peaks.get("hbm_gbs", 6650.0)); peak_src = "measured (MEASURED_PEAKS.json hbm_gbs)" if "hbm_gbs" in peaks else "fallback (B200_PROFILING.md 6.65 TB/s)"
    achieved = bytes_tok / (ms_tok * 1e-3) / 1e9
    traffic = None; traffic_src = None
    try:                                            # DRAM bytes/token measured by ncu --set full on this kernel (profiles/), scaled to this launch
        tj = json.load(open(os.path.join(ROOT, "profiles", "traffic.json")))
        if tj.get("workload_key") == f"{a.preset}-{a.mix}" and not a.n_layer:
            traffic = tj["dram_bytes_per_token"] * K; traffic_src = tj["source"]
    except Exception:
        pass
    roofline = {"bound": "hbm", "achieved": achieved, "peak": peak, "unit": "GB/s", "frac": achieved / peak, "traffic": traffic, "traffic_source": traffic_src,
                "algorithmic_bytes_per_launch": bytes_tok * K,
                "kernel": "decode_kernel (persistent, 1 launch = K tokens)", "algorithmic_bytes_per_token": bytes_tok, "peak_source": peak_src}

    # ---- end to end through the reference-facing handler, host buffers
    e2e = None
    if not a.no_e2e:
        content = "".join(chr(97 + (i * 7) % 26) for i in range(max(1, a.ctx - 4)))
        req = json.dumps({"model": f"bench/{a.preset}-{a.mix}", "messages": [{"role": "user", "content": content}], "temperature": 0,
                          "max_tokens": K, "ignore_eos": True}).encode()
        # warm-up request of the same shape (a serving worker is warm: prompt-ingestion buffers and weight copies exist)
        eng.handle("lmstudio.chat_model", json.dumps({"model": f"bench/{a.preset}-{a.mix}", "messages": [{"role": "user", "content": content[::-1]}], "temperature": 0, "max_tokens": 4, "ignore_eos": True}))
        barrier(); t0 = time.perf_counter()
        rep_raw = eng.handle("lmstudio.chat_model", req)
        wall = time.perf_counter() - t0
        rep = json.loads(rep_raw)["data"]["response"]
        ntok = rep["usage"]["completion_tokens"]; npr = rep["usage"]["prompt_tokens"]
        wall = reduce_max(wall, dist, "cuda")
        e2e = {"value": world * ntok / wall, "unit": UNIT, "h2d_bytes_per_step": (len(req) + 4 * npr) / ntok, "d2h_bytes_per_step": 4 + len(rep_raw) / ntok,
               "wall_s": wall, "prompt_tokens": npr, "completion_tokens": ntok, "ttft_s": rep["stats"]["time_to_first_token"],
               "decode_tok_s_inside_call": rep["stats"]["tokens_per_second"]}

    cb = None
    if rank == 0 and world == 1 and not a.no_cpu:
        cb = cpu_reference_arm(path, 3, 16)
    # ---- batched decode (BASELINE.json metric "bs64") and config 3 (64 concurrent requests, continuous batching): the engine is
    # re-initialised with 64 sequence slots; same files, same C ABI (b200_submit / b200_poll)
    batched = None
    if rank == 0 and world == 1 and not a.no_batched:
        try:
            sys.path.insert(0, os.path.join(ROOT, "tools"))
            import bench_batch
            m.close(); eng.shutdown()
            batched = {}
            for mix in dict.fromkeys([a.mix, "q8_0"]):             # the headline file, and config 3's all-Q8_0 file
                eng2 = b200.Engine(device=local, max_ctx=648, max_seqs=64)
                m2 = eng2.load_file(f"bench/batch-{mix}", model_path(a.preset, mix, a.n_layer, 0))
                batched[mix] = bench_batch.measure(eng2, m2, peak=peak)
                if mix == "q8_0":                                   # config 3 as the reference deploys it: requests over NATS
                    batched[mix]["nats"] = bench_batch.measure_nats(eng2, f"bench/batch-{mix}", m2.n_vocab)
                m2.close(); eng2.shutdown()
            m = None
        except (Exception, SystemExit) as ex:                      # reported, never fatal for the headline line
            batched = {"error": str(ex)}; m = None
    # ---- config 4: Llama-3-70B Q4_K_M.  N = 1: the single-GPU base (the file fits in 180 GB) through the persistent kernel.
    # N > 1: the model tensor-parallel over the N ranks — parity on a small model first (tools/tp_check.py: shards vs the
    # CPU oracle, results identical on all ranks), then bs1 through the persistent kernel with its in-kernel NVLink
    # exchange and bs32 through the batched row path (NCCL).  The replica `value` above stays the headline metric.
    tp = None
    if not a.no_tp and not a.n_layer:
        try:
            import bench_tp
            if m is not None:
                m.close(); m = None
            eng.shutdown()
            gloo = dist.new_group(backend="gloo") if dist else None
            if dist:
                import tp_check
                parity = tp_check.check(rank, world, local, gloo, verbose=False)
            else:
                parity = None
            tp = bench_tp.measure(rank, world, local, gloo, preset=a.tp_preset, n_layer=a.tp_layers, ctx=a.ctx, steps=min(K, 32), warmup=4,
                                  batches=(1, 32) if dist else (1,), peak=peak)
            tp["parity"] = parity
        except (Exception, SystemExit) as ex:
            tp = {"error": f"{type(ex).__name__}: {ex}"}
    if rank == 0:
        line = {"metric": METRIC, "value": value, "unit": UNIT, "n_gpus": world, "steps": K, "warmup": W, "ms_per_step": ms_total / K,
                "higher_is_better": True, "scaling": "weak", "vs_baseline": None, "dtype": "int8 x int4/6/8 dp4a, fp32 accumulate", "data": "synthetic",
                "config": config, "roofline": roofline, "cpu_baseline": cb, "e2e": e2e, "gpu_launches": l1 - l0, "clocks": clk,
                "load_s": load_s, "engine": {k: info[k] for k in ("grid", "ring_bytes", "n_phases", "n_splits", "weight_bytes", "hbm_bytes")},
                "batched": batched, "tp": tp}
        print(json.dumps(line))
    if m is not None:
        m.close()
    if dist:
        dist.destroy_process_group()


if __name__ == "__main__":
    main()
