"""Digest of one ncu report: headline metrics, stall reasons per issue, and the SASS instructions with the most samples
grouped by the block barriers between them.  usage: python tools/ncu_digest.py report.ncu-rep"""
import csv, subprocess, sys, io
rep = sys.argv[1]
raw = subprocess.run(["ncu", "-i", rep, "--page", "raw", "--csv"], capture_output=True, text=True).stdout
rows = list(csv.reader(io.StringIO(raw))); H = rows[0]; V = rows[2] if len(rows) > 2 else rows[1]; d = dict(zip(H, V))
print(d.get("Kernel Name", "")[:120])
for k in ("gpu__time_duration.sum", "dram__bytes_read.sum", "dram__bytes_write.sum", "smsp__issue_active.avg.pct_of_peak_sustained_active", "sm__warps_active.avg.pct_of_peak_sustained_active",
          "launch__registers_per_thread", "launch__occupancy_limit_registers", "launch__occupancy_limit_shared_mem", "launch__occupancy_limit_warps", "lts__throughput.avg.pct_of_peak_sustained_elapsed",
          "dram__throughput.avg.pct_of_peak_sustained_elapsed", "l1tex__data_bank_conflicts_pipe_lsu_mem_shared.sum", "l1tex__data_pipe_lsu_wavefronts_mem_shared.sum", "sm__inst_executed.sum"):
    if k in d: print(f"  {k:70s} {d[k]} {rows[1][H.index(k)]}")
for h in H:
    if "average_warps_issue_stalled" in h and h.endswith("per_issue_active.ratio"):
        try:
            v = float(d[h].replace(",", ""))
            if v > 0.3: print(f"  stall {h.split('issue_stalled_')[1].split('_per_issue')[0]:28s} {v:.2f} per issue")
        except ValueError: pass
src = subprocess.run(["ncu", "-i", rep, "--page", "source", "--csv"], capture_output=True, text=True).stdout
rows = list(csv.reader(io.StringIO(src))); H = rows[1]; si = H.index("Warp Stall Sampling (All Samples)"); so = H.index("Source")
names = [h for h in H if h.startswith("stall_") and "Not Issued" not in h]
data = []
for n, r in enumerate(rows[2:]):
    try: v = float(r[si])
    except (ValueError, IndexError): continue
    top = max(names, key=lambda nm: float(r[H.index(nm)] or 0))
    data.append((v, n, r[so].strip()[:80], top))
tot = sum(v for v, *_ in data) or 1
print(f"  {len(data)} SASS instructions, {tot:.0f} samples")
bars = [n for v, n, s, t in data if "BAR.SYNC" in s]
cuts = [0] + bars + [len(data)]
for i in range(len(cuts) - 1):
    seg = [x for x in data if cuts[i] <= x[1] < cuts[i + 1]]
    print(f"  region {i}: instructions {cuts[i]}..{cuts[i + 1]}: {sum(x[0] for x in seg) / tot * 100:.1f} % of samples")
for v, n, s, t in sorted(sorted(data, reverse=True)[:int(sys.argv[2]) if len(sys.argv) > 2 else 16], key=lambda x: x[1]):
    print(f"  {v / tot * 100:5.1f} %  #{n:4d} {s:80s} {t}")
