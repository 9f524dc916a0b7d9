for d in 3 67 66 65 64; do echo DEBUG$d; B200_QG_DEBUG=$d timeout 100 python tools/qgemm_bench.py Q4_K 2>&1 | tail -4 | grep -E "lm_head|gate_up"; done
