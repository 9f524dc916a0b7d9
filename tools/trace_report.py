"""Summarise a B200_TRACE dump (globaltimer ns).  Stamps per (step, phase, CTA): 0 start, 1 inputs polled, 2 rms known,
3 prologue done, 4 first stage landed (warp 0), 5 stream loop done (warp 0), 6 phase done, 7 smid."""
import sys
import numpy as np
raw = open(sys.argv[1], "rb").read()
n_steps, n_phases, G = np.frombuffer(raw[:12], dtype=np.int32)
t = np.frombuffer(raw[12:], dtype=np.int64).reshape(n_steps, n_phases, G, 8).astype(np.float64)[1:] / 1e3
names = ["embed"] + ["qkv", "attn", "wo", "gate_up", "down"] * ((n_phases - 2) // 5) + ["lm_head"]
end = t[..., 6]
print(f"{'phase':8s} {'poll':>6s} {'rms':>6s} {'quant':>6s} {'1stld':>6s} {'strm':>6s} {'epi':>6s} | {'total':>6s} | {'wall':>7s} {'skew':>6s}   us, medians over CTAs with work; wall = last end - previous phase's last end")
tot = 0
for kind in ["embed", "qkv", "attn", "wo", "gate_up", "down", "lm_head"]:
    idx = [i for i, n in enumerate(names) if n == kind]
    x = t[:, idx]
    d = [np.median(x[..., k + 1] - x[..., k]) for k in range(6)]
    prev = [i - 1 for i in idx]
    wall = (end[:, idx].max(axis=2) - (end[:, prev].max(axis=2) if idx[0] > 0 else t[:, idx, :, 0].min(axis=2)))
    if idx[0] == 0 and n_steps > 2:   # embed follows the previous step's lm_head
        wall = end[1:, idx].max(axis=2) - end[:-1, [n_phases - 1]].max(axis=2)
    skew = end[:, idx].max(axis=2) - np.median(end[:, idx], axis=2)
    tot += wall.mean() * len(idx)
    print(f"{kind:8s} " + " ".join(f"{v:6.2f}" for v in d) + f" | {np.median(x[..., 6] - x[..., 0]):6.2f} | {wall.mean():7.2f} {np.median(skew):6.2f}  x{len(idx)}")
print(f"sum of phase wall times: {tot:.1f} us/token")
for kind in ["qkv", "wo", "gate_up", "down", "lm_head"]:      # warp 0 of each CTA: time blocked on the ring's full-barriers during the stream loop
    idx = [i for i, n in enumerate(names) if n == kind]
    w = t[:, idx, :, 7] * 1e3 / 1965.0                    # cycles (t was divided by 1e3) -> us at 1965 MHz
    print(f"  {kind:8s} warp 0 waited for data {np.median(w):6.2f} us (p90 {np.percentile(w, 90):6.2f}) of a {np.median(t[:, idx, :, 5] - t[:, idx, :, 3]):6.2f} us stream window")
if "--attn" in sys.argv:
    idx = [i for i, n in enumerate(names) if n == "attn"]
    x = t[:, idx]
    comb = (x[..., 4] - x[..., 3]) > 0
    lab = ["q polled", "batches done", "smem combine+publish", "combiner polled", "final published", "phase end"]
    for nm, sel in (("all items", x[..., 6] > x[..., 0]), ("combiner CTAs", comb)):
        d = [np.median((x[..., k + 1] - x[..., k])[sel]) for k in range(6)]
        print(f"  attn {nm:14s}: " + ", ".join(f"{l} {v:.2f}" for l, v in zip(lab, d)) + f" | total {np.median((x[..., 6] - x[..., 0])[sel]):.2f} us")
if "--wait" in sys.argv:
    for kind in ["qkv", "wo", "gate_up", "down", "lm_head"]:
        idx = [i for i, n in enumerate(names) if n == kind]
        w = t[:, idx, :, 7] * 1e3 / 1950.0; s_ = t[:, idx, :, 5] - t[:, idx, :, 3]
        ghz = np.median((t[:, idx, :, 4] - t[:, idx, :, 0]) * 1e3 / np.maximum((t[:, idx, :, 6] - t[:, idx, :, 0]) * 1e3, 1))
        print(f"  {kind:8s} SM clock during phase: {ghz:.3f} GHz")
        cyc = t[:, idx, :, 1] * 1e3; nst = np.maximum(t[:, idx, :, 2] * 1e3, 1)
        print(f"  {kind:8s} warp0: stream window {np.median(s_):6.2f} us, waiting for data {np.median(w):6.2f} us, stages {np.median(nst):.0f}, compute {np.median(cyc/nst):.0f} cycles/stage, total compute {np.median(cyc):.0f} cyc => {np.median(cyc)/max(np.median(s_-w),1e-9)/1e3:.2f} GHz-equivalent")
