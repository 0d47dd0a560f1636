#!/usr/bin/env python
"""Group an ncu SASS source-page CSV into regions of equal execution count (loop bodies vs
straight-line code) with their share of executed warp instructions and stall samples."""
import csv
import sys

rows = list(csv.reader(open(sys.argv[1])))
hdr = rows[1]
body = [dict(zip(hdr, r)) for r in rows[2:] if len(r) == len(hdr) and r[0] != "Address"]
# the file may hold several launches of the same kernel back to back: keep the first
n = len(body)
for k in range(1, n):
    if body[k]["Address"] == body[0]["Address"]:
        body = body[:k]
        break
i, out = 0, []
while i < len(body):
    e = int(body[i]["Instructions Executed"] or 0)
    j = i
    while j < len(body) and int(body[j]["Instructions Executed"] or 0) == e:
        j += 1
    out.append((i, j - 1, e, (j - i) * e, sum(int(b["# Samples"] or 0) for b in body[i:j])))
    i = j
tot = sum(o[3] for o in out)
ts = sum(o[4] for o in out)
print("total warp-instr %d, samples %d" % (tot, ts))
for o in out:
    if o[3] > tot * 0.004 or o[4] > ts * 0.01:
        print("lines %5d-%5d (%4d) exec/line %7d  total %9d (%4.1f%%) samples %4.1f%%" % (
            o[0], o[1], o[1] - o[0] + 1, o[2], o[3], 100 * o[3] / tot, 100 * o[4] / ts))
