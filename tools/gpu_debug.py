import sys, os, numpy as np
sys.path.insert(0,'.'); sys.path.insert(0,'tests'); sys.path.insert(0,'tools')
import nats_llm_studio_b200 as b200, make_gguf
e = b200.Engine(max_ctx=4096)
path = "/tmp/probe-8.gguf"
make_gguf.write_model(path, make_gguf.spec_from_preset("tiny-llama", n_layer=8))
m = e.load_file("probe8", path)
os.environ["B200_TRACE"] = "gpurun_out/trace_tiny.bin"
ms, by = m.bench_decode(512, 16, 4)
print(f"tiny L=8 ctx=512: {ms*1e3:.1f} us/token")
