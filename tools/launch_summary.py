"""Per-kernel totals of the LAST pass in an `ncu --metrics gpu__time_duration.sum --csv` launch list.
Usage: python tools/launch_summary.py launches.csv [first-kernel-substring]   (default: embed_rows = start of a tensor-core pass)"""
import collections, csv, sys
rows = [r for r in csv.reader(open(sys.argv[1])) if len(r) > 10 and r[0].isdigit()]
key = sys.argv[2] if len(sys.argv) > 2 else "embed_rows"
idx = [i for i, r in enumerate(rows) if key in r[4]]
rows = rows[idx[-1]:] if idx else rows
tot = collections.defaultdict(float); cnt = collections.Counter()
for r in rows:
    k = r[4].split("(")[0][-46:]; tot[k] += float(r[-1]); cnt[k] += 1
s = sum(tot.values())
for k, v in sorted(tot.items(), key=lambda kv: -kv[1]):
    print(f"{k:48s} {cnt[k]:5d} {v/1e6:8.3f} ms {100*v/s:5.1f}%  avg {v/cnt[k]/1e3:7.1f} us")
print(f"total {s/1e6:.3f} ms over {sum(cnt.values())} launches")
