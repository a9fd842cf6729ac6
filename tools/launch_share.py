"""summarise an ncu launch list (--metrics gpu__time_duration.sum --csv) of tools/verify_once.py 0: per-kernel share of one verify step and one draft level"""
import csv, re, collections, sys
rows = [r for r in csv.reader(open(sys.argv[1])) if len(r) > 10]
hdr = rows[0]; ki = hdr.index('Kernel Name'); vi = hdr.index('Metric Value'); ui = hdr.index('Metric Unit')
L = []
for r in rows[1:]:
    v = float(r[vi].replace(',', '')); u = r[ui]
    v = v / 1e3 if u == 'ns' else v * 1e3 if u == 'ms' else v
    n = re.sub(r'\(.*', '', r[ki]).replace('void ', '').replace('<unnamed>::', '')
    L.append((n, v))
idx = [i for i, (n, _) in enumerate(L) if n.startswith('get_rows')]
def agg(lo, hi, title):
    d = collections.defaultdict(lambda: [0, 0.0])
    for n, v in L[lo:hi]:
        d[n][0] += 1; d[n][1] += v
    tot = sum(x[1] for x in d.values())
    print(f"## {title}: {hi - lo} launches, {tot:.0f} us (serialised, cold cache)")
    print("| kernel | launches | avg us | us | share |\n|---|---|---|---|---|")
    for n, (c, t) in sorted(d.items(), key=lambda x: -x[1][1]):
        print(f"| {n[:70]} | {c} | {t / c:.1f} | {t:.0f} | {100 * t / tot:.1f} % |")
agg(idx[1], idx[2], "verify step (T=65, n_kv=4096, 7B Q4_K_M)")
agg(idx[3], idx[4], "one draft level (head forward T=8 + top-k + expansion)")
