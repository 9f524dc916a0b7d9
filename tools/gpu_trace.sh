#!/bin/bash
cd "$(dirname "$0")/.."
timeout 100 python tools/gpu_debug.py 2>&1 | tail -2; python tools/trace_report.py gpurun_out/trace_tiny.bin
B200_TRACE=gpurun_out/trace.bin timeout 200 python bench.py --steps 16 --warmup 3 --no-cpu --no-e2e --n-layer ${NLAYER:-8} 2>&1 | tail -1 | cut -c1-160; python tools/trace_report.py gpurun_out/trace.bin
