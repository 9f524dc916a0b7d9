#!/bin/bash
# qgemm what-bounds-it matrix: the full kernel, then with one stage removed at a time via B200_QG_DEBUG
# (1 = no expansion math, 2 = no MMAs, 4 = no activation loads, 8 = no weight loads, 64 = no tensor-memory stores; results are wrong
# in those modes, only the time is meaningful).
cd "$(dirname "$0")/.."
for d in 0 1 2 3 4 8 14 15 64; do
    echo "B200_QG_DEBUG=$d"
    B200_QG_DEBUG=$d timeout 100 python tools/qgemm_bench.py ${1:-Q4_K} 2>&1 | tail -4 | grep -E "lm_head|gate_up"
done
