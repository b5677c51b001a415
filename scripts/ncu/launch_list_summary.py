"""Summarise an ncu launch list (--metrics gpu__time_duration.sum --csv --log-file launches.csv) by kernel: share of the
listed time, launches, mean duration.  python scripts/ncu/launch_list_summary.py gpurun_out/launches.csv "<title>" """
import collections, csv, re, sys
rows = [r for r in csv.reader(l for l in open(sys.argv[1]) if not l.startswith('=='))]
hdr = rows[0]; ki = hdr.index('Kernel Name'); vi = hdr.index('Metric Value'); ui = hdr.index('Metric Unit')
agg = collections.defaultdict(lambda: [0, 0.0])
for r in rows[1:]:
    if len(r) <= vi:
        continue
    v = float(r[vi].replace(',', '')); u = r[ui]
    us = v / 1000 if u in ('nsecond', 'ns') else (v if u in ('usecond', 'us') else v * 1000)
    a = agg[re.sub(r'\(.*', '', r[ki]).strip()]; a[0] += 1; a[1] += us
tot = sum(a[1] for a in agg.values()); n = sum(a[0] for a in agg.values())
print(f"{sys.argv[2] if len(sys.argv) > 2 else sys.argv[1]}; total {tot / 1000:.2f} ms over {n} launches (cold-cache, serialised, boost clocks: shares, not absolutes)")
for k, a in sorted(agg.items(), key=lambda kv: -kv[1][1])[:30]:
    print(f"{100 * a[1] / tot:6.2f} %  {a[1] / 1000:8.3f} ms  {a[0]:5d} x  {a[1] / a[0]:8.1f} us  {k[:90]}")
