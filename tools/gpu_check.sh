#!/bin/bash
# quick regression on the GPU box: parity tests, microbench, then a traced 8-layer bench
cd "$(dirname "$0")/.."
timeout 300 python -m pytest tests -x -q -m gpu 2>&1 | tail -6
for t in "Q4_K 28672 4096" "Q6_K 28672 4096" "Q8_0 28672 4096" "Q4_K 4096 14336" "Q6_K 128256 4096"; do timeout 100 python tools/gemv_bench.py $t 50; done
for inf in ${INFLIGHTS:-2}; do
  echo "== inflight $inf"
  B200_INFLIGHT=$inf B200_TRACE=gpurun_out/trace.bin timeout 200 python bench.py --steps 16 --warmup 3 --no-cpu --no-e2e --n-layer ${NLAYER:-8} 2>&1 | tail -1 | python -c "import sys,json; d=json.loads(sys.stdin.read()); print('tok/s', d['value'], 'ms', d['ms_per_step'], 'frac', d['roofline']['frac'])"
  python tools/trace_report.py gpurun_out/trace.bin
done
