timeout 100 python tools/qgemm_bench.py 2>&1 | tail -12
for d in 15 14; do echo DEBUG$d; B200_QG_DEBUG=$d timeout 100 python tools/qgemm_bench.py Q4_K 2>&1 | tail -4 | grep -E "lm_head"; done
