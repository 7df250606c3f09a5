"""Summarise an `ncu --page source --csv` dump: stall mix + top sampled SASS instructions."""
import csv, collections, sys
rows = list(csv.reader(open(sys.argv[1])))
starts = [i for i, r in enumerate(rows) if r and r[0] == 'Address']
sec = int(sys.argv[2]) if len(sys.argv) > 2 else 0
hdr = rows[starts[sec]]
end = starts[sec + 1] - 1 if len(starts) > sec + 1 else len(rows)
ie = hdr.index('Instructions Executed'); sa = hdr.index('# Samples')
stall_cols = [i for i, h in enumerate(hdr) if h.startswith('stall_') and 'Not Issued' not in h]
data = []
for r in rows[starts[sec] + 1:end]:
    try: data.append((int(r[ie]), int(r[sa]), r[1].strip(), [int(r[i] or 0) for i in stall_cols]))
    except Exception: pass
tot = sum(d[0] for d in data); ts = sum(d[1] for d in data)
print('sections', len(starts), 'instr', tot, 'samples', ts, 'nsass', len(data))
agg = collections.Counter()
for n, s, sx, st in data:
    for i, v in zip(stall_cols, st): agg[hdr[i]] += v
print(' '.join(f"{k[6:]}={100*v/ts:.1f}%" for k, v in agg.most_common(8)))
for k, (n, s, sx, st) in sorted(enumerate(data), key=lambda x: -x[1][1])[:int(sys.argv[3]) if len(sys.argv) > 3 else 20]:
    top = max(zip(st, [hdr[i] for i in stall_cols]))
    print(f"{k:4d} smp{100*s/ts:5.1f}% ex{n:8d} {top[1][6:]:16s} {sx[:80]}")
