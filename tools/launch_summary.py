"""Per-kernel summary of an ncu launch list (--metrics gpu__time_duration.sum --csv).
usage: python tools/launch_summary.py <launches.csv> [top]"""
import collections, csv, sys
rows = list(csv.reader(l for l in open(sys.argv[1]) if l.startswith('"')))
top = int(sys.argv[2]) if len(sys.argv) > 2 else 20
hdr = rows[0]
ki, vi = hdr.index("Kernel Name"), hdr.index("Metric Value")
agg = collections.defaultdict(list)
for r in rows[1:]:
    try:
        agg[r[ki][:70]].append(float(r[vi].replace(",", "")))
    except ValueError:
        pass
tot = sum(sum(v) for v in agg.values())
for k, v in sorted(agg.items(), key=lambda kv: -sum(kv[1]))[:top]:
    print(f"{k:70s} n={len(v):4d} avg={sum(v) / len(v) / 1e3:9.1f}us share={sum(v) / tot:.3f}")
