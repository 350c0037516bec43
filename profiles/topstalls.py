#!/usr/bin/env python
"""profiles/topstalls.py <file.ncu-rep> [N] — the N SASS instructions with the most warp-stall samples (needs -lineinfo
and --import-source on), with the dominant stall reason of each."""
import csv
import subprocess
import sys


def main(path, n=25):
    out = subprocess.run(["ncu", "-i", path, "--page", "source", "--csv"], capture_output=True, text=True).stdout
    rows = [r for r in csv.reader(out.splitlines()) if r]
    i = next(i for i, r in enumerate(rows) if r[0] == "Address")
    hdr = rows[i]
    body = rows[i + 1:]
    si, sa = hdr.index("Source"), hdr.index("# Samples")
    stall = [(j, h) for j, h in enumerate(hdr) if h.startswith("stall_") and "Not Issued" not in h]
    tot = sum(int(r[sa] or 0) for r in body)
    print(f"total samples {tot}, instructions {len(body)}")
    agg = {}
    for r in body:
        for j, h in stall:
            agg[h] = agg.get(h, 0) + int(r[j] or 0)
    print("by reason:", ", ".join(f"{h[6:]} {100 * v / max(tot, 1):.1f}%" for h, v in sorted(agg.items(), key=lambda kv: -kv[1]) if v))
    order = sorted(range(len(body)), key=lambda k: -int(body[k][sa] or 0))[:n]
    for k in sorted(order):
        r = body[k]
        top = sorted(((int(r[j] or 0), h[6:]) for j, h in stall), reverse=True)[:2]
        print(f"{k:5d} {100 * int(r[sa] or 0) / max(tot, 1):5.1f}%  {r[si][:90]:90s} {top}")


if __name__ == "__main__":
    main(sys.argv[1], int(sys.argv[2]) if len(sys.argv) > 2 else 25)
