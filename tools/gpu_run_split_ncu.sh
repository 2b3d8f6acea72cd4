#!/bin/bash
# launch list of the owner split by itself (count, scan, scatter) for world 2 / 8 / 64
cd "$(dirname "$0")/.."; mkdir -p gpurun_out
MEASURE_ONLY=A4 timeout -s KILL 300 python tools/measure_configs.py > gpurun_out/split_plain.log 2>&1 || { echo plain failed; tail -5 gpurun_out/split_plain.log; exit 1; }
grep "A4" gpurun_out/split_plain.log
MEASURE_ONLY=A4 timeout -s KILL 600 ncu --metrics gpu__time_duration.sum --clock-control none -k regex:"xs_|scan_" -c 200 --csv --log-file gpurun_out/split_launches.csv python tools/measure_configs.py > gpurun_out/split_ncu.log 2>&1
python - <<'PY'
import csv, collections
rows = [r for r in csv.reader(open("gpurun_out/split_launches.csv", errors="replace")) if len(r) > 5]
hdr = next(i for i, r in enumerate(rows) if "Kernel Name" in r); H = rows[hdr]; kn, mv = H.index("Kernel Name"), H.index("Metric Value")
agg = collections.OrderedDict()
for r in rows[hdr + 1:]:
    try: v = float(r[mv].replace(",", ""))
    except ValueError: continue
    a = agg.setdefault(r[kn][:90], []); a.append(v)
for k, v in agg.items(): print(f"{k:92s} x{len(v):3d}  avg {sum(v) / len(v) / 1e3:8.1f} us  min {min(v) / 1e3:8.1f}")
PY
