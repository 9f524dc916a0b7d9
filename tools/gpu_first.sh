#!/bin/bash
# first-contact script: progressively riskier steps, each under its own timeout
cd "$(dirname "$0")/.."
mkdir -p gpurun_out
nvidia-smi --query-gpu=name,memory.total,clocks.max.sm --format=csv,noheader
nproc; free -g | head -2
timeout 300 python -m pytest tests/test_gpu_kernels.py -x -q -m gpu -k "dequant" 2>&1 | tail -5
timeout 120 python - <<'PY' 2>&1 | tail -20
import sys, numpy as np
sys.path.insert(0,'.'); sys.path.insert(0,'tests')
import nats_llm_studio_b200 as b200, orc
e = b200.Engine(max_ctx=1024)
rng = np.random.default_rng(0)
for t,bb in (("Q4_K",144),("Q6_K",210),("Q8_0",34),("Q5_K",176)):
    N,K = 64,2048
    be = 32 if t=="Q8_0" else 256
    raw = rng.integers(0,256,size=(N*(K//be),bb),dtype=np.uint8)
    for p in {"Q8_0":[0],"Q4_K":[0,2],"Q5_K":[0,2],"Q6_K":[208]}[t]:
        raw[:,p:p+2] = np.full(raw.shape[0],0.01,np.float16).view(np.uint8).reshape(-1,2)
    x = rng.standard_normal(K).astype(np.float32)
    y,ms = e.gemv(t, raw.reshape(N,-1), N, K, x)
    r = orc.matvec(t, raw.reshape(N,-1), N, K, x, act_quant=1)
    print(t, "max err", np.abs(y-r).max(), "scale", np.abs(r).max(), flush=True)
PY
echo "--- kernels tests"; timeout 600 python -m pytest tests/test_gpu_kernels.py -x -q -m gpu 2>&1 | tail -15
echo "--- model tests"; timeout 600 python -m pytest tests/test_gpu_model.py -x -q -m gpu 2>&1 | tail -30
echo "--- smoke"; timeout 300 python -c "import __graft_entry__ as g; g.smoke()" 2>&1 | tail -5
echo "--- gemv perf"; timeout 300 python - <<'PY' 2>&1 | tail -20
import sys, numpy as np
sys.path.insert(0,'.')
import nats_llm_studio_b200 as b200
e = b200.Engine(max_ctx=1024)
rng = np.random.default_rng(0)
for t,bb,N,K in (("Q4_K",144,28672,4096),("Q6_K",210,128256,4096),("Q8_0",34,28672,4096),("Q4_K",144,4096,14336),("Q6_K",210,4096,14336),("Q4_K",144,6144,4096)):
    be = 32 if t=="Q8_0" else 256
    raw = rng.integers(0,256,size=(N*(K//be),bb),dtype=np.uint8)
    x = rng.standard_normal(K).astype(np.float32)
    y,ms = e.gemv(t, raw.reshape(N,-1), N, K, x, iters=20)
    print(f"{t} {N}x{K}: {ms*1e3:.1f} us  {raw.nbytes/ms/1e6:.0f} GB/s", flush=True)
PY
echo "--- bench 8B"; timeout 900 python bench.py --steps 64 --warmup 4 2>&1 | tail -5
