"""Summarises ncu launch lists (`--metrics gpu__time_duration.sum --csv --log-file X.csv`): per kernel the launch
count, mean and total duration and the SHARE of the listed time (per-launch times under ncu are cold-cache and
serialised, so shares are what carries over to a real run).  Usage: python tools/launch_summary.py a.csv b.csv ..."""
import collections
import csv
import sys

for path in sys.argv[1:]:
    rows = [r for r in csv.reader(open(path)) if len(r) > 10]
    hdr = rows[0]
    ki, vi, ui = hdr.index("Kernel Name"), hdr.index("Metric Value"), hdr.index("Metric Unit")
    d = collections.defaultdict(list)
    for r in rows[1:]:
        try:
            v = float(r[vi].replace(",", ""))
        except ValueError:
            continue
        scale = {"ns": 1e-3, "us": 1.0, "ms": 1e3, "s": 1e6}.get(r[ui], 1e-3)
        d[r[ki]].append(v * scale)
    total = sum(sum(v) for v in d.values())
    print("%s  (ncu gpu__time_duration.sum, --clock-control none; per-launch times are cold-cache and serialised: "
          "use the SHARES)" % path)
    for k, v in sorted(d.items(), key=lambda kv: -sum(kv[1])):
        print("  %-50s n=%4d  mean %9.2f us  total %10.1f us  share %5.1f %%" % (k[:50], len(v), sum(v) / len(v), sum(v),
                                                                               100.0 * sum(v) / total))
