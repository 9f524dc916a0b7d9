#!/bin/bash
for i in $(seq 1 20); do
  out=$(/usr/local/graft/bin/gpurun --timeout 600 -- "$1" 2>&1)
  if echo "$out" | grep -q "status=transient"; then sleep 90; continue; fi
  echo "$out" | tail -${2:-12}; exit 0
done
echo "gave up"
