cd /root/repo
SKIP_GEMV=1 SKIP_TRACE=1 TESTS=tests TAIL=6 bash tools/gpu_probe.sh
for inf in 1 2 3; do echo "== inflight $inf"; B200_INFLIGHT=$inf SKIP_TESTS=1 SKIP_GEMV=1 bash tools/gpu_probe.sh; done
echo "== ring 120KB inflight 2"; B200_RING_KB=120 B200_INFLIGHT=2 SKIP_TESTS=1 SKIP_GEMV=1 bash tools/gpu_probe.sh
