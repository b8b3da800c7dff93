"""Per-kernel totals of an ncu `--metrics gpu__time_duration.sum --csv` launch list."""
import csv, collections, sys
rows = [r for r in csv.reader(open(sys.argv[1])) if len(r) > 5]
hdr = None
agg = collections.OrderedDict()
for r in rows:
    if r[0] == 'ID':
        hdr = r
        continue
    if hdr is None:
        continue
    d = dict(zip(hdr, r))
    try:
        v = float(d['Metric Value'].replace(',', ''))
    except ValueError:
        continue
    unit = d['Metric Unit']
    v = v / 1e3 if unit == 'ns' else (v * 1e3 if unit == 'ms' else v)
    a = agg.setdefault(d['Kernel Name'][:100], [0, 0.0])
    a[0] += 1
    a[1] += v
for k, a in agg.items():
    print(f"{a[0]:4d} {a[1]:10.1f} us  {a[1] / a[0]:9.1f} us/launch  {k}")
print(f"total {sum(a[1] for a in agg.values()):.1f} us")
