#!/usr/bin/env python
"""profiles/timeline_report.py FILE [--last N] — read the record file a -DB200_TIMELINE library wrote at exit
(B200_TIMELINE_OUT) and print the steady-state picture of one layer-op of the decode chain.  The k-th launch of every
kernel belongs to the k-th layer-op; all stamps are relative to the moment the main kernel's first CTA got past its
wait for the predecessor (stores: relative to the same moment).  Medians over the last N layer-ops (default 200),
nanoseconds (%globaltimer ticks in 32 ns steps on B200).  "period" is the distance to the next layer-op's same moment:
what bench.py's ms_per_step / layers measures."""
import statistics
import sys

NAMES = {1: "kv_write", 2: "plan", 3: "main", 4: "combine"}
FIELDS = {1: ["entry", "past-wait", "", "", "", "end"],
          2: ["entry", "", "", "", "", "end"],
          3: ["entry", "first-TMA-issue", "past-wait", "first-tile-done", "first-row-loop-end", "end"],
          4: ["entry", "past-wait", "", "", "", "end"]}


def main():
    path = sys.argv[1]
    last = int(sys.argv[sys.argv.index("--last") + 1]) if "--last" in sys.argv else 200
    by = {}
    for line in open(path):
        if line.startswith("#"):
            continue
        v = [int(x) for x in line.split()]
        by.setdefault((v[0], "first CTA" if v[1] == 0 else "last CTA"), []).append(v)
    for k in by:
        by[k].sort(key=lambda v: v[2])
    mains = by[(3, "first CTA")]
    n = len(mains)
    # kernels launched once per layer-op line up by index from the END (the plan kernel and cache-fill stores do not)
    cols = {}
    periods = []
    for back in range(1, min(last, n - 1) + 1):
        t0 = mains[n - back][4] or mains[n - back][2]
        if back > 1:
            nxt = mains[n - back + 1]
            periods.append((nxt[4] or nxt[2]) - t0)
        for key, rs in by.items():
            if key[0] == 2 or len(rs) < n // 2 or back > len(rs):
                continue
            v = rs[len(rs) - back]
            for j, name in enumerate(FIELDS[key[0]]):
                if name and v[2 + j]:
                    cols.setdefault(key + (name,), []).append(v[2 + j] - t0)
    print(f"{sum(len(r) for r in by.values())} records, {n} layer-ops; medians over the last {min(last, n - 1)}")
    print(f"period: {statistics.median(periods):.0f} ns")
    for key in sorted(cols, key=lambda k: statistics.median(cols[k])):
        print(f"  {statistics.median(cols[key]):8.0f} ns  {NAMES[key[0]]:9s} {key[1]:9s} {key[2]}")


if __name__ == "__main__":
    main()
