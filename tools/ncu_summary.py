"""Aggregates an ncu --csv launch list (gpu__time_duration.sum) by kernel name: total ms, share, launches."""
import collections, csv, sys
rows = list(csv.reader(open(sys.argv[1])))
hdr = [i for i, r in enumerate(rows) if r and r[0] == 'ID'][0]
H = rows[hdr]
ki, vi = H.index('Kernel Name'), H.index('Metric Value')
agg = collections.defaultdict(lambda: [0, 0.0])
for r in rows[hdr + 1:]:
    if len(r) <= vi:
        continue
    n = r[ki].split('(')[0][:70]
    agg[n][0] += 1
    agg[n][1] += float(r[vi].replace(',', ''))
tot = sum(v[1] for v in agg.values())
top = int(sys.argv[2]) if len(sys.argv) > 2 else 30
for n, (c, t) in sorted(agg.items(), key=lambda kv: -kv[1][1])[:top]:
    print(f'{t / 1e6:9.3f} ms {100 * t / tot:5.1f}% {c:5d}  {n}')
print(f'total {tot / 1e6:.3f} ms, {sum(v[0] for v in agg.values())} launches')
