#!/usr/bin/env python
"""profiles/traffic.json <- `ncu -i X.ncu-rep --page raw --csv` of ONE launch of a dominant kernel:
  python scripts/ncu_traffic.py <raw.csv> "<kernel key as bench.py names it>" <capture name>
Records dram__bytes_read.sum + dram__bytes_write.sum, the launch's element count (grid size / CTAs per element) and the
md5 of the kernel's source text, so bench.py reports roofline.traffic only for the source revision that was profiled."""
import csv
import json
import os
import sys

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
import bench  # noqa: E402

raw, key, capture = sys.argv[1], sys.argv[2], sys.argv[3]
rows = list(csv.reader(open(raw)))
h, units, v = rows[0], rows[1], rows[2]


def val(name):
    i = h.index(name)
    x = float(v[i].replace(",", ""))
    return x * {"Mbyte": 1e6, "Kbyte": 1e3, "Gbyte": 1e9, "byte": 1.0, "": 1.0}.get(units[i], 1.0)


grid = int(float(v[h.index("launch__grid_size")]))
ctas_per_element = 4 if "split" in key else 1
path = os.path.join(ROOT, "profiles", "traffic.json")
T = json.load(open(path)) if os.path.exists(path) else {}
T[key] = {"launch_elements": grid // ctas_per_element, "dram_read_bytes": val("dram__bytes_read.sum"),
          "dram_write_bytes": val("dram__bytes_write.sum"), "kernel_ms": float(v[h.index("gpu__time_duration.sum")]) * {"us": 1e-3, "ns": 1e-6, "ms": 1.0, "s": 1e3}.get(units[h.index("gpu__time_duration.sum")], 1.0),
          "source_md5": bench.source_md5(key), "capture": capture}
json.dump(T, open(path, "w"), indent=1, sort_keys=True)
print(json.dumps(T[key]))
