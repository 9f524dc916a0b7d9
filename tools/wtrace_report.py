"""Summarise a per-warp cycle trace of ONE CTA of decode_kernel (build variant -DB200_WTRACE, env B200_WTRACE=<file>,
B200_WTRACE_CTA=<cta>).  Stamps (clock64) per (step, phase, warp): 0 phase start, 1 own first poll done, 2 past the first
prologue barrier, 3 own quantisation done, 4 prologue done, 5 first stage available, 6 stream loop done, 7 phase end.
Prints, per phase kind, each stamp relative to the EARLIEST warp's phase start: median over (steps, layers) of the
min / median / max over the consumer warps, in microseconds at the given clock."""
import sys
import numpy as np

raw = open(sys.argv[1], "rb").read()
mhz = float(sys.argv[2]) if len(sys.argv) > 2 else 1965.0
n_steps, n_phases, W = np.frombuffer(raw[:12], dtype=np.int32)
t = np.frombuffer(raw[12:], dtype=np.int64).reshape(n_steps, n_phases, W, 8).astype(np.float64)[1:, :, :14]   # consumer warps
names = ["embed"] + ["qkv", "attn", "wo", "gate_up", "down"] * ((n_phases - 2) // 5) + ["lm_head"]
lab = ["start", "poll", "bar1", "quant", "pro", "1st", "strm", "end"]
print(f"{'phase':8s} {'stat':4s} " + " ".join(f"{l:>6s}" for l in lab) + "   us from the first warp's phase start; next = gap to the next phase's first start")
for kind in ["qkv", "attn", "wo", "gate_up", "down", "lm_head"]:
    idx = [i for i, n in enumerate(names) if n == kind]
    x = t[:, idx]                                            # [steps, layers, warps, 8]
    t0 = x[..., 0].min(axis=2, keepdims=True)[..., None]     # earliest warp start
    rel = (x - t0) / mhz
    rel[x == 0] = np.nan
    for nm, fn in (("min", np.nanmin), ("med", np.nanmedian), ("max", np.nanmax)):
        v = fn(rel, axis=2)                                  # over warps
        print(f"{kind:8s} {nm:4s} " + " ".join(f"{np.nanmedian(v[..., k]):6.2f}" for k in range(8)))
