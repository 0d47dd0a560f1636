#!/usr/bin/env python
"""Condense an `ncu --page source --print-source sass --csv` dump: per-instruction executed
counts and stall samples, top-N by samples and a region histogram.  Dev aid.
   ncu -i rep --page source --csv --print-source sass --kernel-name regex:X > x.csv
   python tools/sasshot.py x.csv [N]"""
import csv
import sys

rows = list(csv.reader(open(sys.argv[1])))
n = int(sys.argv[2]) if len(sys.argv) > 2 else 40
# several kernels may be concatenated: take blocks starting with "Kernel Name"
i = 0
while i < len(rows):
    if rows[i] and rows[i][0] == "Kernel Name":
        name = rows[i][1][:90]
        hdr = rows[i + 1]
        j = i + 2
        body = []
        while j < len(rows) and not (rows[j] and rows[j][0] == "Kernel Name"):
            if len(rows[j]) == len(hdr):
                body.append(dict(zip(hdr, rows[j])))
            j += 1
        i = j
        tot_s = sum(int(r["# Samples"] or 0) for r in body)
        tot_i = sum(int(r["Instructions Executed"] or 0) for r in body)
        print("==", name, "instr lines", len(body), "samples", tot_s, "warp-instr", tot_i)
        top = sorted(range(len(body)), key=lambda k: -int(body[k]["# Samples"] or 0))[:n]
        for k in sorted(top):
            r = body[k]
            st = {h[6:]: int(r[h]) for h in hdr if h.startswith("stall_") and "Not Issued" not in h and r[h] not in ("", "0")}
            st = sorted(st.items(), key=lambda kv: -kv[1])[:3]
            print("%5d %6s smp %5.1f%% exe %8s  %-70s %s" % (k, r["Address"][-5:], 100.0 * int(r["# Samples"] or 0) / max(tot_s, 1),
                                                            r["Instructions Executed"], r["Source"][:70], st))
    else:
        i += 1
