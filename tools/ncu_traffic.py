"""profiles/traffic.json from `ncu --set full` reports: DRAM bytes (dram__bytes_read.sum + dram__bytes_write.sum) of the
one render_frame launch each report holds, stamped with the hash of the kernel sources in this tree (bench.py refuses a
stamp that does not match the sources it runs: roofline.traffic = null instead of a stale figure).
    python tools/ncu_traffic.py render_map_4k=gpurun_out/r02/prof.ncu-rep spheres1024_4k=gpurun_out/r02/prof_wide.ncu-rep"""
import csv
import io
import json
import os
import subprocess
import sys

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
import bench  # noqa: E402  (kernel_source_hash)

UNIT = {"byte": 1.0, "Kbyte": 1e3, "Mbyte": 1e6, "Gbyte": 1e9, "Tbyte": 1e12}
path = os.path.join(ROOT, "profiles", "traffic.json")
try:
    out = json.load(open(path))
except Exception:
    out = {}
for arg in sys.argv[1:]:
    workload, rep = arg.split("=")
    txt = subprocess.run(["ncu", "-i", rep, "--page", "raw", "--csv"], capture_output=True, text=True, check=True).stdout
    rows = list(csv.reader(io.StringIO(txt)))
    head, units, vals = rows[0], rows[1], rows[2]
    tot = 0.0
    for name in ("dram__bytes_read.sum", "dram__bytes_write.sum"):
        i = head.index(name)
        tot += float(vals[i].replace(",", "")) * UNIT[units[i]]
    out[workload] = {"dram_bytes": tot, "source_hash": bench.kernel_source_hash(), "capture": "profiles/r02_" + ("render_frame_wide" if "wide" in rep else "render_frame") + "_details.txt (ncu --set full)",
                     "kernel": vals[head.index("Kernel Name")]}
    print(workload, out[workload])
json.dump(out, open(path, "w"), indent=1)
