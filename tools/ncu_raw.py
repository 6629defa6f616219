"""Prints selected metrics of an .ncu-rep raw-page CSV: python tools/ncu_raw.py file.csv [substring ...]"""
import csv
import sys

rows = list(csv.reader(open(sys.argv[1])))
keys = sys.argv[2:]
hdr, units = rows[0], rows[1]
for r in rows[2:]:
    print("==", r[4][:60], "grid", r[8], "block", r[7])
    for h, u, v in zip(hdr, units, r):
        if any(k in h for k in keys):
            print(f"  {h} [{u}] = {v}")
