#!/usr/bin/env python
"""Records the measured DRAM traffic of the dominant kernel for bench.py's `roofline.traffic`.
usage: ncu_traffic.py <raw.csv from `ncu -i X.ncu-rep --page raw --csv`> <config> <math> [kernel-prefix]
Writes / updates profiles/traffic_ncu.json: {"<config>:<math>": {"kernel", "bytes" (read + write per launch, mean over the
captured launches), "read", "write", "duration_us", "source"}}."""
import csv, json, os, sys

raw, config, math = sys.argv[1:4]
prefix = sys.argv[4] if len(sys.argv) > 4 else "step_"
rows = list(csv.reader(open(raw)))
hdr, units = rows[0], rows[1]
col = {h: i for i, h in enumerate(hdr)}
scale = {"byte": 1.0, "Kbyte": 1e3, "Mbyte": 1e6, "Gbyte": 1e9}
rd = wr = dur = 0.0
cnt = 0
name = None
for r in rows[2:]:
    k = r[col["Kernel Name"]]
    if prefix not in k:
        continue
    name = k[k.index(prefix):].split("(")[0].split("<")[0]
    rd += float(r[col["dram__bytes_read.sum"]]) * scale[units[col["dram__bytes_read.sum"]]]
    wr += float(r[col["dram__bytes_write.sum"]]) * scale[units[col["dram__bytes_write.sum"]]]
    dur += float(r[col["gpu__time_duration.sum"]]) * {"us": 1.0, "ms": 1e3, "ns": 1e-3}[units[col["gpu__time_duration.sum"]]]
    cnt += 1
if not cnt:
    sys.exit("no launch of a kernel containing " + prefix)
root = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
path = os.path.join(root, "profiles", "traffic_ncu.json")
data = json.load(open(path)) if os.path.exists(path) else {}
data[f"{config}:{math}"] = {"kernel": name, "bytes": round((rd + wr) / cnt), "read": round(rd / cnt), "write": round(wr / cnt),
                           "duration_us": round(dur / cnt, 2), "launches": cnt, "source": os.path.relpath(os.path.abspath(raw), root)}
json.dump(data, open(path, "w"), indent=1, sort_keys=True)
print(data[f"{config}:{math}"])
