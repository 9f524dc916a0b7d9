#!/bin/bash
# round-2 probe on the GPU box: serving tests, GEMV microbench (HBM-sized and L2-resident), traced 8B decode
cd "$(dirname "$0")/.."
mkdir -p gpurun_out
if [ -z "$SKIP_TESTS" ]; then timeout 600 python -m pytest ${TESTS:-tests/test_gpu_serving.py} -x -q -m gpu 2>&1 | tail -${TAIL:-8}; fi
if [ -z "$SKIP_GEMV" ]; then
for t in "Q4_K 28672 4096" "Q4_K 4096 4096" "Q4_K 6144 4096" "Q6_K 4096 14336" "Q6_K 128256 4096" "Q8_0 28672 4096" "Q8_0 4096 4096"; do timeout 100 python tools/gemv_bench.py $t 100; done
fi
if [ -z "$SKIP_TRACE" ]; then
B200_TRACE=gpurun_out/trace.bin timeout 300 python bench.py --steps ${STEPS:-16} --warmup 3 --no-cpu --no-e2e --no-batched ${NLAYER:+--n-layer $NLAYER} 2>&1 | tail -1 | python -c "import sys,json; d=json.loads(sys.stdin.read()); print('tok/s', d['value'], 'ms', d['ms_per_step'], 'frac', d['roofline']['frac'])"
python tools/trace_report.py gpurun_out/trace.bin --attn
rm -f gpurun_out/trace.bin
fi
