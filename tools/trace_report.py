"""Summarise a B200_TRACE dump: per-phase-kind mean cycles for prologue / stream / barrier (4 sampled CTAs)."""
import sys
import numpy as np
raw = open(sys.argv[1], "rb").read()
n_steps, n_phases = np.frombuffer(raw[:8], dtype=np.int32)
t = np.frombuffer(raw[8:], dtype=np.int64).reshape(n_steps, n_phases, 4, 4)
mhz = float(sys.argv[2]) if len(sys.argv) > 2 else 1965.0
t = t[1:]                                            # skip first step
pro = (t[..., 1] - t[..., 0]) / mhz; stream = (t[..., 2] - t[..., 1]) / mhz; bar = (t[..., 3] - t[..., 2]) / mhz
# phase index pattern: 0 embed, then per layer 5 phases (qkv, attn, wo, gu, down), last lm
names = ["embed"] + ["qkv", "attn", "wo", "gate_up", "down"] * ((n_phases - 2) // 5) + ["lm_head"]
print(f"{'phase':8s} {'prologue':>9s} {'stream':>9s} {'barrier':>9s} {'total':>9s}  us (mean over steps, layers, 4 CTAs); per-CTA stream max-min")
tot = 0
for kind in ["embed", "qkv", "attn", "wo", "gate_up", "down", "lm_head"]:
    idx = [i for i, n in enumerate(names) if n == kind]
    p, s, b = pro[:, idx].mean(), stream[:, idx].mean(), bar[:, idx].mean()
    spread = (stream[:, idx].max(axis=2) - stream[:, idx].min(axis=2)).mean()
    tot += (p + s + b) * len(idx)
    print(f"{kind:8s} {p:9.2f} {s:9.2f} {b:9.2f} {p+s+b:9.2f}   x{len(idx)}  spread {spread:.2f}")
print(f"sum over phases: {tot:.1f} us/token")
