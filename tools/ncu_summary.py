#!/usr/bin/env python
"""Digest an .ncu-rep (read here, no GPU needed) into the short text/JSON summaries kept under profiles/.
usage: python tools/ncu_summary.py gpurun_out/prof.ncu-rep [--json out.json] [--top 25]"""
import csv, io, json, subprocess, sys

def ncu_csv(rep, page, extra=()):
    out = subprocess.run(["ncu", "-i", rep, "--page", page, "--csv", *extra], capture_output=True, text=True).stdout
    return list(csv.reader(io.StringIO(out)))

KEYS = ["gpu__time_duration.sum", "dram__bytes_read.sum", "dram__bytes_write.sum", "gpu__dram_throughput.avg.pct_of_peak_sustained_elapsed",
        "lts__throughput.avg.pct_of_peak_sustained_elapsed", "sm__throughput.avg.pct_of_peak_sustained_elapsed",
        "sm__warps_active.avg.pct_of_peak_sustained_active", "launch__registers_per_thread", "launch__grid_size", "launch__block_size",
        "launch__occupancy_limit_registers", "smsp__inst_executed.sum", "smsp__issue_active.avg.pct_of_peak_sustained_active",
        "lts__t_sector_hit_rate.pct", "lts__t_sectors_srcunit_tex_op_red.sum", "lts__t_requests_srcunit_tex_op_red.sum",
        "lts__t_sectors_srcunit_tex_op_read.sum", "lts__t_sectors_srcunit_tex_op_write.sum", "lts__t_sectors_srcunit_tex_op_atom.sum",
        "sm__pipe_tensor_cycles_active.avg.pct_of_peak_sustained_active"]

def to_bytes(v, unit):
    v = float(v.replace(",", ""))
    return v * {"byte": 1, "Kbyte": 1e3, "Mbyte": 1e6, "Gbyte": 1e9, "Tbyte": 1e12}.get(unit, 1)

def main():
    rep = sys.argv[1]; top = 25; jout = None
    if "--top" in sys.argv: top = int(sys.argv[sys.argv.index("--top") + 1])
    if "--json" in sys.argv: jout = sys.argv[sys.argv.index("--json") + 1]
    rows = ncu_csv(rep, "raw")
    hdr, units = rows[0], rows[1]
    summ = []
    for r in rows[2:]:
        name = r[hdr.index("Kernel Name")]
        d = {"kernel": name}
        for k in KEYS:
            if k in hdr: d[k] = r[hdr.index(k)] + " " + units[hdr.index(k)]
        d["dram_bytes"] = to_bytes(r[hdr.index("dram__bytes_read.sum")], units[hdr.index("dram__bytes_read.sum")]) + \
            to_bytes(r[hdr.index("dram__bytes_write.sum")], units[hdr.index("dram__bytes_write.sum")])
        st = [(float(r[i].replace(",", "")), h.replace("smsp__average_warps_issue_stalled_", "").replace("_per_issue_active.ratio", ""))
              for i, h in enumerate(hdr) if h.startswith("smsp__average_warps_issue_stalled_") and r[i] not in ("", "n/a")]
        d["stalls_per_issue"] = {h: v for v, h in sorted(st, reverse=True)[:8]}
        summ.append(d)
    for d in summ:
        print("==", d["kernel"][:90])
        for k, v in d.items():
            if k not in ("kernel",): print(f"   {k:70s} {v}")
    # SASS hot spots of the first launch
    rows = ncu_csv(rep, "source", ["--print-source=sass"])
    hi = [i for i, r in enumerate(rows) if r and r[0] == "Address"]
    if hi:
        hdr = rows[hi[0]]; first = []
        for r in rows[hi[0] + 1:]:
            if r and r[0] == "Kernel Name": break
            if len(r) == len(hdr): first.append(r)
        ix, isamp, isrc = hdr.index("Instructions Executed"), hdr.index("# Samples"), hdr.index("Source")
        tots = sum(int(r[isamp]) for r in first) or 1
        print(f"== SASS: {len(first)} instructions, {sum(int(r[ix]) for r in first)} warp-instructions executed, {tots} samples")
        for r in sorted(first, key=lambda r: -int(r[isamp]))[:top]:
            print(f"   {100 * int(r[isamp]) / tots:5.1f}%  exec={int(r[ix]):10d}  {r[isrc][:100]}")
    if jout:
        json.dump(summ, open(jout, "w"), indent=1)

if __name__ == "__main__":
    main()
