#!/bin/bash
# per-round profile capture of the headline bench command (B200_PROFILING.md recipe): launch list + one full capture
cd "$(dirname "$0")/.."
R=${ROUND:-r2}
CMD="python bench.py --steps 8 --warmup 3 --no-cpu --no-e2e --no-batched --no-tp"
timeout 300 $CMD > gpurun_out/plain.log 2>&1 && timeout 400 ncu --metrics gpu__time_duration.sum --clock-control none -c 500 --csv --log-file gpurun_out/${R}_launches.csv $CMD > gpurun_out/ncu1.log 2>&1
timeout 300 $CMD > gpurun_out/plain2.log 2>&1 && timeout 600 ncu --set full --clock-control none --import-source on -k regex:decode_kernel -s 1 -c 1 -o gpurun_out/${R}_decode_8b $CMD > gpurun_out/ncu2.log 2>&1
# the prompt-ingestion side of the e2e number: launch list of one chat request with a 512-token prompt
CMD2="python bench.py --steps 8 --warmup 3 --no-cpu --no-batched --no-tp"
timeout 300 $CMD2 > gpurun_out/plain3.log 2>&1 && timeout 400 ncu --metrics gpu__time_duration.sum --clock-control none -c 3000 --csv --log-file gpurun_out/${R}_e2e_launches.csv $CMD2 > gpurun_out/ncu3.log 2>&1
tail -2 gpurun_out/ncu2.log; tail -1 gpurun_out/plain.log | cut -c1-200
