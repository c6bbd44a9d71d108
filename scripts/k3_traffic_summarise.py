"""ncu --csv log of scripts/k3_traffic_target.py -> profiles/k3_traffic_r2.json (mean DRAM bytes per K3 launch)."""
import csv
import io
import json
import sys
from pathlib import Path

src, dst = Path(sys.argv[1]), Path(sys.argv[2])
text = src.read_text()
start = text.index('"ID"')
rows = list(csv.DictReader(io.StringIO(text[start:])))
per = {}
for r in rows:
    if "gram_kernel" not in r["Kernel Name"]:
        continue
    v = float(r["Metric Value"].replace(",", ""))
    unit = r["Metric Unit"].lower()
    scale = {"byte": 1.0, "kbyte": 1e3, "mbyte": 1e6, "gbyte": 1e9, "tbyte": 1e12, "ns": 1.0, "us": 1e3, "ms": 1e6, "s": 1e9,
             "nsecond": 1.0, "usecond": 1e3, "msecond": 1e6, "second": 1e9}.get(unit, 1.0)
    per.setdefault(r["ID"], {})[r["Metric Name"]] = v * scale
tot = [d.get("dram__bytes_read.sum", 0) + d.get("dram__bytes_write.sum", 0) for d in per.values()]
rd = [d.get("dram__bytes_read.sum", 0) for d in per.values()]
wr = [d.get("dram__bytes_write.sum", 0) for d in per.values()]
out = {"launches": len(tot), "dram_bytes_per_launch_mean": sum(tot) / max(len(tot), 1),
       "dram_read_bytes_per_launch_mean": sum(rd) / max(len(rd), 1), "dram_write_bytes_per_launch_mean": sum(wr) / max(len(wr), 1),
       "dram_bytes_per_launch_min": min(tot) if tot else None, "dram_bytes_per_launch_max": max(tot) if tot else None,
       "source": f"ncu --metrics dram__bytes_read.sum,dram__bytes_write.sum -k regex:gram_kernel over the {len(tot)} K3 launches of "
                 "one all-fragment build of the bench workload (scripts/k3_traffic_target.py); mean per launch"}
dst.write_text(json.dumps(out, indent=1))
print(json.dumps(out))
