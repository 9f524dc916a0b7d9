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


# ---- tensor-parallel shard arithmetic on CPU: b200_tp_shard_window is the host function Model::commit
# uses to cut every matrix; here two gloo ranks cut float matrices with it and run one decoder block's worth of
# products: column-parallel results concatenate to the full product, row-parallel partial sums all-reduce to it.
def _shard_worker(rank, world, port, q):
    sys.path.insert(0, ROOT)
    os.environ.update(MASTER_ADDR="127.0.0.1", MASTER_PORT=str(port))
    import numpy as np
    import torch
    import torch.distributed as dist
    import nats_llm_studio_b200 as b200
    dist.init_process_group("gloo", rank=rank, world_size=world)
    rng = np.random.default_rng(42)                                   # same matrices on every rank
    D, QD, KVD, F, V = 512, 512, 256, 1536, 1000
    W = {"attn_q": rng.standard_normal((QD, D)), "attn_k": rng.standard_normal((KVD, D)), "attn_v": rng.standard_normal((KVD, D)),
         "attn_output": rng.standard_normal((D, QD)), "ffn_gate": rng.standard_normal((F, D)), "ffn_up": rng.standard_normal((F, D)),
         "ffn_down": rng.standard_normal((D, F)), "output": rng.standard_normal((V, D))}
    x = rng.standard_normal(D)

    def cut(kind):
        N, K = W[kind].shape
        r0, nr, c0, nc = b200.Engine.tp_shard_window(kind, N, K, rank, world)
        return W[kind][r0:r0 + nr, c0:c0 + nc], (r0, nr, c0, nc)

    def gather_rows(part):                                            # column-parallel: concatenate the ranks' row ranges
        parts = [None] * world; dist.all_gather_object(parts, part); return np.concatenate(parts)

    def reduce_sum(part):                                             # row-parallel: all-reduce of the partial sums
        t = torch.from_numpy(part.copy()); dist.all_reduce(t); return t.numpy()

    errs = {}
    q_, k_, v_ = [gather_rows(cut(n)[0] @ x) for n in ("attn_q", "attn_k", "attn_v")]
    errs["qkv"] = max(np.abs(q_ - W["attn_q"] @ x).max(), np.abs(k_ - W["attn_k"] @ x).max(), np.abs(v_ - W["attn_v"] @ x).max())
    att = rng.standard_normal(QD)                                     # stands for the attention output, sharded by heads like q
    wo, (r0, nr, c0, nc) = cut("attn_output")
    errs["wo"] = np.abs(reduce_sum(wo @ att[c0:c0 + nc]) - W["attn_output"] @ att).max()
    g, (gr0, gnr, _, _) = cut("ffn_gate"); u, _ = cut("ffn_up")
    act_l = (g @ x) / (1 + np.exp(-(g @ x))) * (u @ x)                # this rank's FFN columns
    act = (W["ffn_gate"] @ x) / (1 + np.exp(-(W["ffn_gate"] @ x))) * (W["ffn_up"] @ x)
    wd, (_, _, dc0, dnc) = cut("ffn_down")
    assert (dc0, dnc) == (gr0, gnr)                                   # the down window is exactly the gate/up row range
    errs["down"] = np.abs(reduce_sum(wd @ act_l) - W["ffn_down"] @ act).max()
    lg = cut("output")[0] @ x
    best = (float(lg.max()), int(lg.argmax()) + cut("output")[1][0])  # (value, global index) per rank: what greedy decode exchanges
    cands = [None] * world; dist.all_gather_object(cands, best)
    errs["argmax"] = 0.0 if max(cands)[1] == int((W["output"] @ x).argmax()) else 1.0
    q.put((rank, {k: float(v) for k, v in errs.items()}))
    dist.destroy_process_group()


def test_tensor_parallel_shard_windows_world2():
    pytest.importorskip("torch")
    import torch.multiprocessing as mp
    ctx = mp.get_context("spawn")
    q = ctx.Queue(); port = _free_port()
    procs = [ctx.Process(target=_shard_worker, args=(r, 2, port, q)) for r in range(2)]
    for p in procs:
        p.start()
    res = [q.get(timeout=180) for _ in procs]
    for p in procs:
        p.join(timeout=60); assert p.exitcode == 0
    for rank, errs in res:
        assert all(v < 1e-9 for v in errs.values()), (rank, errs)


def test_shard_window_rejects_bad_splits():
    sys.path.insert(0, ROOT)
    import nats_llm_studio_b200 as b200
    W = b200.Engine.tp_shard_window
    assert W("ffn_down", 8192, 28672, 3, 8) == (0, 8192, 3 * 3584, 3584)            # 70B at TP8: 14 super-blocks per rank
    assert W("attn_q", 8192, 8192, 7, 8) == (7 * 1024, 1024, 0, 8192)
    for args in (("ffn_down", 4096, 14336, 0, 8),      # 14336 / 8 = 1792 = 7 x 256: fine ...
                 ):
        W(*args)
    with pytest.raises(b200.EngineError):
        W("ffn_down", 4096, 14336, 0, 16)               # ... but 896 is not a multiple of 256
    with pytest.raises(b200.EngineError):
        W("output", 128257, 4096, 0, 2)
    with pytest.raises(b200.EngineError):
        W("token_embd", 128256, 4096, 0, 2)             # replicated, not sharded
    with pytest.raises(b200.EngineError):
        W("attn_q", 4096, 4096, 2, 2)
