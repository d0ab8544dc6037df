#!/usr/bin/env python3
"""profiles/ncu_traffic.py <report.ncu-rep> <out.json> [commit] — dram__bytes_read.sum + dram__bytes_write.sum of every captured launch of the
match kernel (one `ncu --set full` capture of the five launches of one host-driven registration) -> the JSON bench.py reads for
`roofline.traffic` (mean per launch), with the per-launch list and the commit the capture was taken at."""
import csv
import io
import json
import subprocess
import sys

rep, out = sys.argv[1], sys.argv[2]
commit = sys.argv[3] if len(sys.argv) > 3 else "unknown"
txt = subprocess.run(["ncu", "-i", rep, "--page", "raw", "--csv"], capture_output=True, text=True).stdout
rows = list(csv.reader(io.StringIO(txt)))
hdr, units = rows[0], rows[1]
col = {k: hdr.index(k) for k in ("Kernel Name", "dram__bytes_read.sum", "dram__bytes_write.sum", "gpu__time_duration.sum", "lts__t_sector_hit_rate.pct",
                                  "sm__warps_active.avg.pct_of_peak_sustained_active", "smsp__issue_active.avg.pct_of_peak_sustained_active") if k in hdr}


def to_bytes(v, unit):
    v = float(v.replace(",", ""))
    return v * {"byte": 1, "Kbyte": 1e3, "Mbyte": 1e6, "Gbyte": 1e9}.get(unit, 1)


launches = []
for r in rows[2:]:
    if "knn1" not in r[col["Kernel Name"]]:
        continue
    rd = to_bytes(r[col["dram__bytes_read.sum"]], units[col["dram__bytes_read.sum"]])
    wr = to_bytes(r[col["dram__bytes_write.sum"]], units[col["dram__bytes_write.sum"]])
    launches.append({"kernel": r[col["Kernel Name"]][:48], "dram_read_bytes": rd, "dram_write_bytes": wr,
                     "duration": r[col["gpu__time_duration.sum"]] + " " + units[col["gpu__time_duration.sum"]],
                     "l2_hit_pct": r[col["lts__t_sector_hit_rate.pct"]] if "lts__t_sector_hit_rate.pct" in col else None,
                     "warps_active_pct": r[col["sm__warps_active.avg.pct_of_peak_sustained_active"]] if "sm__warps_active.avg.pct_of_peak_sustained_active" in col else None,
                     "issue_active_pct": r[col["smsp__issue_active.avg.pct_of_peak_sustained_active"]] if "smsp__issue_active.avg.pct_of_peak_sustained_active" in col else None})
mean = sum(l["dram_read_bytes"] + l["dram_write_bytes"] for l in launches) / max(len(launches), 1)
json.dump({"dram_bytes_per_launch": mean, "launches": launches, "commit": commit,
           "how": "ncu --set full --clock-control none -k regex:knn1 -c 5, B200REG_HOST_LOOP=1 python profiles/prof_knn.py (the five match launches of one registration of BASELINE config #2)"},
          open(out, "w"), indent=1)
print(json.dumps({"dram_bytes_per_launch": mean, "n": len(launches)}))
