#!/bin/bash
# run the traced 8B decode under several env settings: tools/gpu_sweep.sh "A=1 B=2" "A=2" ...   (first: quick parity tests)
cd "$(dirname "$0")/.."
mkdir -p gpurun_out
if [ -z "$SKIP_TESTS" ]; then timeout 600 python -m pytest ${TESTS:-tests/test_gpu_model.py tests/test_gpu_kernels.py} -x -q -m gpu 2>&1 | tail -${TAIL:-4}; fi
for cfg in "$@"; do
  echo "== $cfg"
  env $cfg B200_TRACE=gpurun_out/trace.bin timeout 300 python bench.py --steps ${STEPS:-16} --warmup 3 --no-cpu --no-e2e --no-batched --no-tp ${NLAYER:+--n-layer $NLAYER} 2>&1 | tail -1 | python -c "import sys,json; d=json.loads(sys.stdin.read()); print('tok/s', round(d['value'],1), 'ms', round(d['ms_per_step'],4), 'frac', round(d['roofline']['frac'],4))"
  python tools/trace_report.py gpurun_out/trace.bin | grep -v "^  attn"
done
if [ -n "$KEEP_TRACE" ]; then env $KEEP_TRACE B200_TRACE=gpurun_out/trace_keep.bin timeout 300 python bench.py --steps 4 --warmup 3 --no-cpu --no-e2e --no-batched --no-tp 2>&1 | tail -1 | cut -c1-120; fi
rm -f gpurun_out/trace.bin
