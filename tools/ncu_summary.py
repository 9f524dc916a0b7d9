"""Turn gpurun_out ncu artefacts into the committed per-round summary under profiles/."""
import collections
import csv
import subprocess
import sys

rnd = sys.argv[1] if len(sys.argv) > 1 else "r1"
launches = f"gpurun_out/{rnd}_launches.csv"
rep = f"gpurun_out/{rnd}_decode_8b.ncu-rep"
out = []
# ---- launch list
rows = [r for r in csv.reader(l for l in open(launches) if not l.startswith("==")) if len(r) > 5]
hdr = rows[0]; col = {h: i for i, h in enumerate(hdr)}
agg = collections.OrderedDict()
for r in rows[1:]:
    if r[col["Metric Name"]] != "gpu__time_duration.sum":
        continue
    name = r[col["Kernel Name"]].split("(")[0]
    v = float(r[col["Metric Value"]].replace(",", "")); unit = r[col["Metric Unit"]]
    v_us = v / 1e3 if unit in ("ns", "nsecond") else v * 1e3 if unit in ("ms", "msecond") else v
    a = agg.setdefault(name, [0, 0.0]); a[0] += 1; a[1] += v_us
tot = sum(a[1] for a in agg.values())
out.append(f"# ncu launch list — `python bench.py --steps 8 --warmup 3 --no-cpu --no-e2e --no-batched --no-tp` ({rnd})\n")
out.append("Per-launch times under ncu are cold-cache and serialised: compare shares, not absolutes.\n")
out.append("| kernel | launches | total us | share |\n|---|---:|---:|---:|")
for k, (n, t) in sorted(agg.items(), key=lambda kv: -kv[1][1]):
    out.append(f"| `{k}` | {n} | {t:.1f} | {100*t/tot:.1f}% |")
dec = [float(r[col["Metric Value"]].replace(",", "")) for r in rows[1:] if "decode_kernel" in r[col["Kernel Name"]]]
out.append(f"\nThe timed region of the bench contains exactly one `decode_kernel` launch (8 tokens); the load-time kernels (`repack_kernel`) run once per model before it.\n")
# ---- full capture of the decode kernel
raw = subprocess.run(["ncu", "-i", rep, "--page", "raw", "--csv"], capture_output=True, text=True).stdout
rr = list(csv.reader(raw.splitlines()))
h = rr[0]; u = rr[1]; v = rr[2]
want = ["gpu__time_duration.sum", "dram__bytes_read.sum", "dram__bytes_write.sum", "gpu__dram_throughput.avg.pct_of_peak_sustained_elapsed",
        "dram__throughput.avg.pct_of_peak_sustained_elapsed", "sm__throughput.avg.pct_of_peak_sustained_elapsed", "smsp__issue_active.avg.pct_of_peak_sustained_active",
        "sm__warps_active.avg.pct_of_peak_sustained_active", "launch__registers_per_thread", "launch__grid_size", "launch__block_size",
        "launch__shared_mem_per_block_dynamic", "lts__t_sector_hit_rate.pct", "sm__pipe_alu_cycles_active.avg.pct_of_peak_sustained_active",
        "sm__pipe_fma_cycles_active.avg.pct_of_peak_sustained_active", "l1tex__data_bank_conflicts_pipe_lsu_mem_shared.sum", "smsp__inst_executed.sum"]
out.append(f"\n# ncu --set full: `decode_kernel`, the timed launch (8 tokens, Llama-3-8B Q4_K_M, ctx 512) ({rnd})\n")
out.append("| metric | value | unit |\n|---|---:|---|")
vals = {}
for i, name in enumerate(h):
    if name in want:
        vals[name] = (v[i], u[i]); out.append(f"| {name} | {v[i]} | {u[i]} |")
stall = []
for i, name in enumerate(h):
    if "pcsamp_warps_issue_stalled" in name and "not_issued" not in name:
        try:
            stall.append((float(v[i]), name.split("stalled_")[1]))
        except ValueError:
            pass
ts = sum(s for s, _ in stall)
out.append("\nWarp stall sampling (all warps, whole launch): " + ", ".join(f"{n} {100*s/ts:.0f}%" for s, n in sorted(stall, reverse=True)[:8]))
try:
    rd = float(vals["dram__bytes_read.sum"][0].replace(",", "")); wr = float(vals["dram__bytes_write.sum"][0].replace(",", ""))
    scale = {"Gbyte": 1e9, "Mbyte": 1e6, "Kbyte": 1e3, "byte": 1}
    rd *= scale.get(vals["dram__bytes_read.sum"][1], 1); wr *= scale.get(vals["dram__bytes_write.sum"][1], 1)
    out.append(f"\nDRAM traffic per launch: {rd/1e9:.3f} GB read + {wr/1e9:.3f} GB written = {(rd+wr)/1e9:.3f} GB for 8 tokens = {(rd+wr)/8/1e9:.3f} GB/token "
               f"(algorithmic: 4.685 GB/token at ctx 512 => traffic/algorithmic = {(rd+wr)/8/4.685e9:.3f}).")
except Exception as e:
    out.append(f"\n(traffic parse failed: {e})")
open(f"profiles/{rnd}_ncu_summary.md", "w").write("\n".join(out) + "\n")
print("\n".join(out))
