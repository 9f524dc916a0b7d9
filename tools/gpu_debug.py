import sys, os, json, time, numpy as np, faulthandler
faulthandler.dump_traceback_later(40, exit=True)
sys.path.insert(0,'.'); sys.path.insert(0,'tests'); sys.path.insert(0,'tools')
import nats_llm_studio_b200 as b200, make_gguf
e = b200.Engine(max_ctx=2048)
path = "/tmp/probe-2.gguf"
make_gguf.write_model(path, make_gguf.spec_from_preset("llama3-8b", n_layer=2, mix="q4_k_m", seed=2))
print("written", flush=True)
m = e.load_file("probe", path)
print("loaded", flush=True)
ms, by = m.bench_decode(512, 8, 2); print("bench_decode ok", ms, flush=True)
for n_prompt in (4, 40, 300):
    t0=time.time(); out, ttft, dec = m.generate(np.arange(1, n_prompt+1, dtype=np.int32), 8); print("generate", n_prompt, out[:4], round(time.time()-t0,3), flush=True)
req = {"model": "probe", "messages": [{"role": "user", "content": "warm"}], "temperature": 0, "max_tokens": 4, "ignore_eos": True}
t0=time.time(); r = e.handle("lmstudio.chat_model", json.dumps(req)); print("chat short", len(r), round(time.time()-t0,3), flush=True)
req["messages"][0]["content"] = "x"*500; req["max_tokens"] = 16
t0=time.time(); r = e.handle("lmstudio.chat_model", json.dumps(req)); print("chat long", len(r), round(time.time()-t0,3), flush=True)
