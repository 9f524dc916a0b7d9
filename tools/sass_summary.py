#!/usr/bin/env python3
"""Per-kernel SASS evidence (B200_PROFILING.md "What proves a Blackwell-native kernel"): counts of the Blackwell-specific
mnemonics in every kernel of libb200engine.so, from `cuobjdump -sass`.  Writes profiles/<round>_sass_summary.md."""
import collections
import re
import subprocess
import sys

so = "nats_llm_studio_b200/libb200engine.so"
rnd = sys.argv[1] if len(sys.argv) > 1 else "r2"
sass = subprocess.run(["cuobjdump", "-sass", so], capture_output=True, text=True).stdout
WANT = [("UTCHMMA", "tcgen05.mma (kind::f16)"), ("UTCBAR", "tcgen05.commit"), ("LDTM", "tcgen05.ld"), ("STTM", "tcgen05.st"), ("UTMALDG", "cp.async.bulk.tensor (TMA load)"),
        ("UBLKCP", "cp.async.bulk (1-D bulk copy)"), ("UBLKPF", "cp.async.bulk.prefetch.L2"), ("SYNCS", "mbarrier"), ("IDP.4A", "dp4a"), ("HMMA", "mma.sync (legacy)"), ("SHFL", "warp shuffle"),
        ("LDL", "local-memory load (spill)"), ("STL", "local-memory store (spill)")]
cur = None; cnt = collections.OrderedDict(); n_ins = collections.Counter()
for line in sass.splitlines():
    m = re.match(r"\s*Function : (\S+)", line)
    if m:
        cur = subprocess.run(["c++filt", m.group(1)], capture_output=True, text=True).stdout.strip().split("(")[0]
        cnt.setdefault(cur, collections.Counter()); continue
    if cur and re.match(r"\s+/\*[0-9a-f]{4,}\*/", line):
        n_ins[cur] += 1
        for key, _ in WANT:
            if re.search(r"\b" + re.escape(key), line):
                cnt[cur][key] += 1
out = [f"# SASS summary of `{so}` ({rnd}; `cuobjdump -sass`, sm_100a)\n", "Mnemonic -> PTX it comes from: " + "; ".join(f"`{k}` = {v}" for k, v in WANT) + "\n",
       "| kernel | instructions | " + " | ".join(k for k, _ in WANT) + " |", "|---|---:|" + "---:|" * len(WANT)]
for k, c in sorted(cnt.items(), key=lambda kv: -n_ins[kv[0]]):
    out.append(f"| `{k}` | {n_ins[k]} | " + " | ".join(str(c[key]) if c[key] else "" for key, _ in WANT) + " |")
open(f"profiles/{rnd}_sass_summary.md", "w").write("\n".join(out) + "\n")
print("\n".join(out))
