#!/usr/bin/env python
"""Config 1 (BASELINE.json configs[0]): granite-3.0-2b-instruct architecture, random-init Q8_0 GGUF (seed 1),
16-token prompt, greedy 64 tokens, bs1.  The reference runs this on LM Studio's llama.cpp CPU runtime, which is unavailable
on this box; the CPU arm here is the oracle (oracle/oracle.c, fp32, OpenMP on all host cores), timed, and the same request
then goes through the engine on the GPU: tokens and first-position logits are compared.  One JSON line."""
import json, os, sys, time
import numpy as np
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT); sys.path.insert(0, os.path.join(ROOT, "tools")); sys.path.insert(0, os.path.join(ROOT, "tests"))
import make_gguf  # noqa: E402
import orc  # noqa: E402
import nats_llm_studio_b200 as b200  # noqa: E402

d = os.environ.get("B200_BENCH_DIR") or "/dev/shm"
path = os.path.join(d, "b200-config1-granite3-2b-q8_0.gguf")
if not os.path.exists(path):
    make_gguf.write_model(path + ".tmp", make_gguf.spec_from_preset("granite3-2b", seed=1)); os.replace(path + ".tmp", path)
prompt = np.random.default_rng(1).integers(0, 49155, 16).astype(np.int32)
n_gen = int(os.environ.get("CONFIG1_TOKENS", "64"))
o = orc.Model(path)
t0 = time.perf_counter(); want, lg = o.generate(prompt, n_gen, want_logits=True); cpu_s = time.perf_counter() - t0
eng = b200.Engine(max_ctx=256)
m = eng.load_file("config1/granite", path)
m.generate(prompt, 4)                                   # warm-up
t0 = time.perf_counter(); got, ttft, dec = m.generate(prompt, n_gen); gpu_s = time.perf_counter() - t0
lg0 = m.eval(prompt, all_logits=False)
agree = int(np.argmax(np.concatenate([(got != want).astype(int), [1]])))        # length of the common prefix
err = float(np.abs(lg0 - lg[0]).max() / np.abs(lg[0]).max())
print(json.dumps({"workload": "granite3-2b Q8_0 (random-init, seed 1), 16-token prompt, greedy, bs1", "generated": n_gen,
                  "cpu_oracle_tok_s": round(n_gen / cpu_s, 2), "cpu_cores": os.cpu_count(), "cpu_s": round(cpu_s, 2),
                  "gpu_tok_s_e2e": round(n_gen / gpu_s, 1), "gpu_decode_ms_per_token": round(dec / max(1, n_gen - 1), 3),
                  "first_logits_rel_err": err, "greedy_tokens_equal_prefix": agree, "of": n_gen}))
m.close()
