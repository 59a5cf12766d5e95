import csv, collections, sys
rows = list(csv.reader(open(sys.argv[1])))
h = [i for i, r in enumerate(rows) if r and r[0] == 'ID'][0]
t = collections.defaultdict(lambda: collections.defaultdict(list))
for r in rows[h + 1:]:
    t[r[4].split('(')[0][:34]][r[-3]].append(float(r[-1].replace(',', '')))
for k, v in t.items():
    g = lambda m: sum(v[m]) / len(v[m]) if v[m] else 0
    print(f"{k:36} n={len(v['gpu__time_duration.sum']):3} {g('gpu__time_duration.sum')/1e3:8.1f} us  {g('smsp__inst_executed.sum')/1e6:7.2f} M inst  issue {g('smsp__issue_active.avg.pct_of_peak_sustained_active'):5.1f}%  warps {g('sm__warps_active.avg.pct_of_peak_sustained_active'):5.1f}%")
