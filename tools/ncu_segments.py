#!/usr/bin/env python3
"""Per-100-instruction segments of an ncu source page (csv): share of samples, executed instructions, top stall reasons."""
import csv, sys
rows = list(csv.reader(open(sys.argv[1])))
step = int(sys.argv[2]) if len(sys.argv) > 2 else 100
hdr = rows[1]; data = rows[2:]
ix = {h: i for i, h in enumerate(hdr)}
S = ix['# Samples']; E = ix['Instructions Executed']
stalls = [h for h in hdr if h.startswith('stall_') and 'Not Issued' not in h]
tot = sum(int(r[S] or 0) for r in data)
print('total samples', tot, 'instructions', len(data))
for a in range(0, len(data), step):
    seg = data[a:a + step]
    s = sum(int(r[S] or 0) for r in seg); e = sum(int(r[E] or 0) for r in seg)
    st = {h: sum(int(r[ix[h]] or 0) for r in seg) for h in stalls}
    top = sorted(st.items(), key=lambda x: -x[1])[:3]
    if s > tot * 0.002:
        print("%5d %5.1f%% %6.0fM  %s" % (a, 100 * s / tot, e / 1e6, " ".join("%s=%.1f" % (k[6:], 100 * v / tot) for k, v in top)))
