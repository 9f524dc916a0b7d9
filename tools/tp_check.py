#!/usr/bin/env python
"""Tensor-parallel parity check, run under torchrun (one rank per GPU): every rank loads its shard of the same GGUF.
  * batched row path: all ranks ingest the same prompts and take batched decode steps in lockstep;
  * persistent batch-1 kernel: teacher-forced logits (each rank holds a vocabulary slice) and a greedy stream, with the
    partial sums and argmax candidates exchanged inside the kernel over NVLink.
Logits must agree with the CPU oracle within the stated tolerances and results must be bit-identical on all ranks.
Exit code 0 = pass.  bench.py --gpus N runs check() on its N ranks and reports the outcome in its `tp` object."""
import os, sys
import numpy as np
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT); sys.path.insert(0, os.path.join(ROOT, "tools")); sys.path.insert(0, os.path.join(ROOT, "tests"))


def check(rank, world, local, group=None, verbose=True):
    import torch.distributed as dist
    import make_gguf
    import nats_llm_studio_b200 as b200
    import orc
    say = (lambda *a: print(*a, flush=True)) if verbose and rank == 0 else (lambda *a: None)
    path = f"/tmp/b200-tp-check-{world}.gguf"
    if rank == 0:
        make_gguf.write_model(path, make_gguf.spec_from_preset("small-llama", n_ff=max(3072, 1024 * world), n_head_kv=max(2, world), n_head=max(8, 2 * world), head_dim=128))
    dist.barrier(group=group)
    eng = b200.Engine(device=local, max_ctx=256, max_seqs=4)
    ids = [eng.tp_unique_id() if rank == 0 else None]
    dist.broadcast_object_list(ids, src=0, group=group)
    eng.tp_init(rank, world, ids[0])
    m = eng.load_file("tp/check", path)
    rng = np.random.default_rng(3)
    prompts = [rng.integers(0, m.n_vocab, n).astype(np.int32) for n in (40, 7, 90)]
    slots = np.array([2, 0, 3], np.int32)
    for p, s in zip(prompts, slots):
        m.prefill_slot(int(s), p[:-1])
    cur = np.array([p[-1] for p in prompts], np.int32); pos = np.array([len(p) - 1 for p in prompts], np.int32)
    toks = []
    nxt, lg = m.batch_step(cur, slots, pos, want_logits=True)
    toks.append(nxt.copy()); cur = nxt; pos += 1
    for _ in range(4):
        cur = m.batch_step(cur, slots, pos); toks.append(cur.copy()); pos += 1
    toks = np.stack(toks, 1)
    ok = True; res = {"world": world}
    gl = [None] * world; dist.all_gather_object(gl, (lg.tobytes(), toks.tobytes()), group=group)
    ok &= all(g == gl[0] for g in gl)                                    # identical on all ranks
    if rank == 0:
        o = orc.Model(path); worst = 0.0
        for i, p in enumerate(prompts):
            want, ref = o.generate(p, 1, want_logits=True)
            scale = np.abs(ref[0]).max(); err = np.abs(lg[i] - ref[0]).max(); worst = max(worst, float(err / scale))
            say(f"seq {i}: max|logit err| = {err:.4g} ({err / scale:.3%} of max |logit|), token {toks[i, 0]} vs oracle {want[0]}")
            ok &= bool(err <= 4e-2 * scale)
            srt = np.sort(ref[0])
            if srt[-1] - srt[-2] > 8e-2 * scale:
                ok &= bool(toks[i, 0] == want[0])
        res["batched_logit_err_vs_fp32_oracle"] = worst
    # ---- the persistent batch-1 kernel on the same shards: partial sums and argmax candidates cross NVLink inside the kernel
    dist.barrier(group=group)                                            # rank 0 has just spent seconds in the CPU oracle: start the lockstep part together
    prompt = rng.integers(0, m.n_vocab, 12).astype(np.int32)
    sl = m.eval(prompt)                                                  # this rank's vocabulary columns are filled
    V_l = m.n_vocab // world
    parts = [None] * world; dist.all_gather_object(parts, sl[:, rank * V_l:(rank + 1) * V_l].copy(), group=group)
    lg1 = np.concatenate(parts, axis=1)
    toks1, _, _ = m.generate(prompt, 10)
    gl = [None] * world; dist.all_gather_object(gl, toks1.tobytes(), group=group)
    ok &= all(g == gl[0] for g in gl)                                    # every rank decoded the same stream
    if rank == 0:
        ref = orc.Model(path, act_quant=1, kv_f16=1).forward(prompt)
        scale = np.abs(ref).max(); per_pos = np.abs(lg1 - ref).max(axis=1) / scale
        say(f"persistent kernel, TP x{world}: logits vs oracle (engine numerics) per position: first {per_pos[0]:.2e}, max {per_pos.max():.2e}")
        ok &= bool(per_pos[0] <= 1e-4 and per_pos.max() <= 4e-2)
        want, wl = orc.Model(path, act_quant=1, kv_f16=1).generate(prompt, 10, want_logits=True)
        for i in range(10):
            if toks1[i] != want[i]:
                srt = np.sort(wl[i]); ok &= bool(srt[-1] - srt[-2] <= 4e-2 * np.abs(wl[i]).max())   # only a near-tie may differ
                break
        say(f"persistent kernel, TP x{world}: greedy {toks1.tolist()} oracle {want.tolist()}")
        res.update(bs1_logit_err_first_pos=float(per_pos[0]), bs1_logit_err_max=float(per_pos.max()), bs1_greedy_equal=bool(np.array_equal(toks1, want)))
    flag = [ok, res]; dist.broadcast_object_list(flag, src=0, group=group)
    m.close(); eng.shutdown()
    res = flag[1]; res["pass"] = bool(ok and flag[0]); res["identical_on_all_ranks"] = bool(ok or not flag[0])
    return res


if __name__ == "__main__":
    import torch.distributed as dist
    rank, world, local = int(os.environ["RANK"]), int(os.environ["WORLD_SIZE"]), int(os.environ.get("LOCAL_RANK", "0"))
    dist.init_process_group("gloo")
    r = check(rank, world, local)
    if rank == 0:
        print("TP CHECK", "PASS" if r["pass"] else "FAIL", f"world={world}", r, flush=True)
    dist.destroy_process_group()
    sys.exit(0 if r["pass"] else 1)
