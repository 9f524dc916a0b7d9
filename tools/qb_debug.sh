timeout 100 python tools/qgemm_bench.py 2>&1 | tail -12
