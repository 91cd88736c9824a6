"""Bucket the ncu source page (SASS) of the sweep kernel by execution frequency.
usage: python tools/sass_buckets.py src.csv <chunk_warps>"""
import csv, sys
rows = list(csv.reader(open(sys.argv[1])))
# several captured kernels are concatenated: keep the first
ends = [i for i, r in enumerate(rows) if r and r[0] == 'Kernel Name']
if len(ends) > 1:
    rows = rows[:ends[1]]
cw = float(sys.argv[2])
hdr = rows[1]
ia = hdr.index('Source'); ie = hdr.index('Instructions Executed'); it = hdr.index('Avg. Threads Executed'); isamp = hdr.index('# Samples')
data = [(r[ia], int(r[ie]), float(r[it] or 0), int(r[isamp])) for r in rows[2:] if len(r) > ie]
tot = sum(d[1] for d in data); ts = sum(d[3] for d in data)
print('total inst', tot, 'inst per chunk-warp %.1f' % (tot / cw), 'samples', ts)
out = []; cur = None
for k, (s, e, t, sm) in enumerate(data):
    f = e / cw
    toks = s.strip().split()
    op = toks[1] if toks[0].startswith('@') else toks[0]
    if cur and abs(f - cur['f']) <= 0.02 * max(f, cur['f'], 0.01):
        cur['n'] += 1; cur['inst'] += f; cur['samp'] += sm; cur['ops'][op] = cur['ops'].get(op, 0) + 1; cur['thr'] += t; cur['end'] = k
    else:
        if cur: out.append(cur)
        cur = {'start': k, 'end': k, 'f': f, 'n': 1, 'inst': f, 'samp': sm, 'ops': {op: 1}, 'thr': t}
out.append(cur)
for c in out:
    if c['inst'] > 0.3 or c['samp'] > 0.005 * ts:
        ops = sorted(c['ops'].items(), key=lambda x: -x[1])[:5]
        print(f"{c['start']:5d}-{c['end']:5d} n={c['n']:3d} exec/cw={c['f']:.3f} inst/cw={c['inst']:6.2f} thr={c['thr']/c['n']:4.1f} samp={100*c['samp']/ts:5.1f}% {ops}")
