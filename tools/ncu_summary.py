"""Write the metric/value/unit summary of one ncu report: python tools/ncu_summary.py rep.ncu-rep metrics.txt > out.csv
(metrics.txt: one metric name per line; default = tools/ncu_metrics.txt)."""
import csv, subprocess, sys
rep = sys.argv[1]
import os
names = [l.strip() for l in open(sys.argv[2] if len(sys.argv) > 2 else os.path.join(os.path.dirname(os.path.abspath(__file__)), "ncu_metrics.txt"))]
raw = subprocess.run(["ncu", "-i", rep, "--page", "raw", "--csv"], capture_output=True, text=True).stdout
rows = list(csv.reader(raw.splitlines()))
h, u, v = rows[0], rows[1], rows[2]
m = {k: (vv, uu) for k, uu, vv in zip(h, u, v)}
w = csv.writer(sys.stdout)
w.writerow(["metric", "value", "unit"])
for n in names:
    if n in m: w.writerow([n, m[n][0], m[n][1]])
