"""N>1 host logic of bench.py on CPU (gloo, world_size 2): the replicas share one model file written by rank 0, the
reported time is the max over ranks and the value is the whole-job aggregate.  (The data path itself has no
collective: batch-1 decode of a model that fits one GPU runs as independent replicas — DESIGN.md §5.)"""
import os
import socket
import sys

import pytest

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))


def _free_port():
    s = socket.socket(); s.bind(("127.0.0.1", 0)); p = s.getsockname()[1]; s.close(); return p


def _worker(rank, world, port, tmpdir, q):
    sys.path.insert(0, ROOT); sys.path.insert(0, os.path.join(ROOT, "tools"))
    os.environ.update(MASTER_ADDR="127.0.0.1", MASTER_PORT=str(port), B200_BENCH_DIR=tmpdir)
    import torch.distributed as dist
    import bench
    dist.init_process_group("gloo", rank=rank, world_size=world)
    path = bench.model_path("tiny-llama", "q4_k_m", 0, rank)        # rank 0 writes, the others wait for the file
    dist.barrier()
    size = os.path.getsize(path)
    ms = bench.reduce_max(10.0 * (rank + 1), dist)                  # rank r "took" 10(r+1) ms
    val = bench.aggregate_value(world, 100, ms)
    q.put((rank, size, ms, val))
    dist.destroy_process_group()


def test_replica_aggregation_world2(tmp_path):
    torch = pytest.importorskip("torch")
    import torch.multiprocessing as mp
    ctx = mp.get_context("spawn")
    q = ctx.Queue(); port = _free_port()
    procs = [ctx.Process(target=_worker, args=(r, 2, port, str(tmp_path), q)) for r in range(2)]
    for p in procs:
        p.start()
    res = sorted(q.get(timeout=120) for _ in procs)
    for p in procs:
        p.join(timeout=60); assert p.exitcode == 0
    assert res[0][1] == res[1][1] > 0                               # both ranks see the same complete file
    assert res[0][2] == res[1][2] == 20.0                           # max over ranks, identical everywhere
    assert abs(res[0][3] - 2 * 100 / 0.020) < 1e-6                  # aggregate tokens/s over the slowest rank


def test_single_process_aggregation():
    sys.path.insert(0, ROOT)
    import bench
    assert bench.reduce_max(3.5) == 3.5 and bench.aggregate_value(1, 256, 512.0) == 500.0
