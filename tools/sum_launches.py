import csv, sys, collections
with open(sys.argv[1]) as f:
    lines = [l for l in f if not l.startswith('==')]
agg = collections.OrderedDict()
for row in csv.DictReader(lines):
    if row.get('Metric Name') != 'gpu__time_duration.sum': continue
    v = float(row['Metric Value'].replace(',', '')); u = row['Metric Unit']
    v = v / 1e3 if u in ('ns', 'nsecond') else v
    k = row['Kernel Name'][:48] + " g" + row.get('Grid Size', '')
    a = agg.setdefault(k, [0, 0.0]); a[0] += 1; a[1] += v
for k, a in sorted(agg.items(), key=lambda x: -x[1][1])[:int(sys.argv[2]) if len(sys.argv) > 2 else 12]:
    print("%-75s n=%4d avg %8.1f us" % (k, a[0], a[1] / a[0]))
