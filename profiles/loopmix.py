#!/usr/bin/env python
"""profiles/loopmix.py <file.ncu-rep> — opcode mix of the hot loop (instructions executed >= 50 % of the median non-zero count)."""
import collections
import csv
import statistics
import subprocess
import sys


def main(path):
    out = subprocess.run(["ncu", "-i", path, "--page", "source", "--csv"], capture_output=True, text=True).stdout
    rows = [r for r in csv.reader(out.splitlines()) if r]
    i = next(i for i, r in enumerate(rows) if r[0] == "Address")
    hdr, body = rows[i], rows[i + 1:]
    si, ei = hdr.index("Source"), hdr.index("Instructions Executed")
    ex = [int(r[ei] or 0) for r in body]
    med = statistics.median(sorted(e for e in ex if e > 0)[len(ex) // 4:])
    cnt, tot = collections.Counter(), 0
    for r, e in zip(body, ex):
        if e >= 0.5 * med:
            tok = r[si].split()
            op = tok[1] if tok[0].startswith("@") else tok[0]
            cnt[op.split(".")[0]] += e / med
            tot += e / med
    print(f"hot-loop instructions per iteration ~ {tot:.0f} (median exec count {med:.0f})")
    for k, v in cnt.most_common(40):
        print(f"  {k:12s} {v:6.1f}")


if __name__ == "__main__":
    main(sys.argv[1])
