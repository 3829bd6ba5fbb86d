"""Write profiles/ransac_traffic.json (what bench.py reports as roofline.traffic) from an ncu --set full report of
lslam_ransac_lines.  usage: ncu_traffic.py <report.ncu-rep> <scans in the captured launch> <name to cite>"""
import csv
import io
import json
import os
import subprocess
import sys

rep, scans, cite = sys.argv[1], int(sys.argv[2]), sys.argv[3]
raw = subprocess.run(["ncu", "-i", rep, "--page", "raw", "--csv"], capture_output=True, text=True).stdout
rows = list(csv.reader(io.StringIO(raw)))
hdr, units, vals = rows[0], rows[1], rows[-1]
scale = {"byte": 1.0, "Kbyte": 1e3, "Mbyte": 1e6, "Gbyte": 1e9}
out = {"file": cite, "scans": scans}
for h, u, v in zip(hdr, units, vals):
    if h == "dram__bytes_read.sum":
        out["dram_bytes_read"] = float(v.replace(",", "")) * scale[u]
    if h == "dram__bytes_write.sum":
        out["dram_bytes_write"] = float(v.replace(",", "")) * scale[u]
    if h == "Kernel Name":
        out["kernel"] = v
    if h == "gpu__time_duration.sum":
        out["ncu_duration"] = v + " " + u
    if h == "smsp__issue_active.avg.pct_of_peak_sustained_active":
        out["issue_active_pct"] = float(v.replace(",", ""))
    if h == "sm__warps_active.avg.pct_of_peak_sustained_active":
        out["warps_active_pct"] = float(v.replace(",", ""))
root = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
json.dump(out, open(os.path.join(root, "profiles", "ransac_traffic.json"), "w"), indent=1)
print(out)
