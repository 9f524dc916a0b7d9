"""Timeline of qgemm_kernel's roles in CTA 0 (build variant -DB200_QG_TRACE, env B200_QG_TRACE=<file>; stamps are clock64).
Roles: 0 weight producer (0: ring slot free), 1 activation producer (0: slot free), 2 MMA thread (0: starts waiting, 1: A ready,
2: X ready, 3: MMAs issued + commit), 3 / 4 first warp of expansion group 0 / 1 (0: starts waiting for the blob, 1: blob landed,
2: math done, 3: A slot free, 4: tensor-memory stores issued, 5: stores complete + arrived)."""
import sys
import numpy as np
t = np.fromfile(sys.argv[1], dtype=np.int64).reshape(6, 64, 8).astype(np.float64)
mhz = 1965.0
t0 = t[t > 0].min()
r = lambda x: (x - t0) / mhz if x > 0 else float("nan")
n = int(sys.argv[2]) if len(sys.argv) > 2 else 24
print("unit | W free | X free | M wait  A-rdy  X-rdy  issued | grp: wait  blob  math  Afree  st  done   (us from the first stamp)")
for it in range(n):
    g = 3 + (it & 1)
    print(f"{it:4d} | {r(t[0, it, 0]):6.2f} | {r(t[1, it, 0]):6.2f} | {r(t[2, it, 0]):6.2f} {r(t[2, it, 1]):6.2f} {r(t[2, it, 2]):6.2f} {r(t[2, it, 3]):6.2f} | g{it & 1}: "
          + " ".join(f"{r(t[g, it, k]):6.2f}" for k in range(6)))
