#!/usr/bin/env python
"""Condense an `ncu --csv --metrics ...` log into one line per launch (used for profiles/*.csv)."""
import csv
import sys


def main(path):
    rows = list(csv.reader(open(path)))
    hdr, data, order = None, {}, []
    for r in rows:
        if len(r) > 10 and r[0] == "ID":
            hdr = r
            continue
        if hdr and len(r) == len(hdr):
            d = dict(zip(hdr, r))
            key = (int(d["ID"]), d["Kernel Name"].split("(")[0][-60:])
            if key not in data:
                order.append(key)
            data.setdefault(key, {})[d["Metric Name"]] = d["Metric Value"]
    names = sorted({m for v in data.values() for m in v})
    print(",".join(["id", "kernel"] + names))
    for k in order:
        print(",".join([str(k[0]), k[1].replace(",", ";")] + [data[k].get(m, "") for m in names]))


if __name__ == "__main__":
    main(sys.argv[1])
