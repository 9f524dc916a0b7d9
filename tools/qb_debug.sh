timeout 100 python tools/qgemm_bench.py Q4_K Q8_0 2>&1 | tail -8
echo WHOLE; B200_QG_WHOLE=1 timeout 100 python tools/qgemm_bench.py Q4_K 2>&1 | tail -4
