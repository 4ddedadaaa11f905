"""Record the DRAM traffic of a workload's march kernel from an ncu capture into profiles/traffic.json, stamped with the
hash of the kernels' source (bench.py reports roofline.traffic only while that hash still matches the tree).
usage: python tools/update_traffic.py <workload> <capture.ncu-rep> [kernel-name-substring]"""
import csv
import io
import json
import os
import subprocess
import sys

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
import bench  # noqa: E402

workload, rep = sys.argv[1], sys.argv[2]
want = sys.argv[3] if len(sys.argv) > 3 else "march"
out = subprocess.run(["ncu", "-i", rep, "--page", "raw", "--csv"], capture_output=True, text=True).stdout
rows = list(csv.reader(io.StringIO(out)))
hdr, units = rows[0], rows[1]
total, n = 0.0, 0
scale = {"byte": 1.0, "Kbyte": 1e3, "Mbyte": 1e6, "Gbyte": 1e9}
for r in rows[2:]:
    d, u = dict(zip(hdr, r)), dict(zip(hdr, units))
    if want not in d.get("Kernel Name", ""):
        continue
    total += sum(float(d[k].replace(",", "")) * scale[u[k]] for k in ("dram__bytes_read.sum", "dram__bytes_write.sum"))
    n += 1
if n == 0:
    raise SystemExit("no kernel matching %r in %s" % (want, rep))
path = os.path.join(ROOT, "profiles", "traffic.json")
try:
    tj = json.load(open(path))
except Exception:
    tj = {}
sha = bench.kernel_source_sha16()
if tj.get("source_sha16") != sha:                 # captures of another source do not mix
    tj = {"source_sha16": sha, "captures": {}}
tj[workload] = int(round(total / n))
tj.setdefault("captures", {})[workload] = os.path.basename(rep)
json.dump(tj, open(path, "w"), indent=1, sort_keys=True)
print(workload, tj[workload], "bytes per launch (%d launches) ->" % n, path)
