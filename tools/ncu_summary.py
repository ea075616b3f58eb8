#!/usr/bin/env python3
"""Summarise an ncu capture (raw.csv + src.csv exported with `ncu -i X.ncu-rep --page raw|source --csv`)."""
import csv, sys
raw, src = sys.argv[1], sys.argv[2]
rows = list(csv.reader(open(raw)))
d = dict(zip(rows[0], rows[2]))
for k in ['gpu__time_duration.sum', 'launch__grid_size', 'launch__registers_per_thread', 'sm__pipe_fmaheavy_cycles_active.avg.pct_of_peak_sustained_elapsed',
          'sm__inst_executed_pipe_alu.avg.pct_of_peak_sustained_active', 'smsp__issue_active.avg.pct_of_peak_sustained_active', 'dram__bytes_read.sum', 'dram__bytes_write.sum',
          'lts__t_sector_hit_rate.pct', 'smsp__inst_executed.sum', 'sm__warps_active.avg.pct_of_peak_sustained_active']:
    print(k, d.get(k), rows[1][rows[0].index(k)] if k in rows[0] else '')
tot, items = 0, []
for k in d:
    if 'pcsamp_warps_issue_stalled' in k and 'not_issued' not in k:
        try: v = float(d[k])
        except ValueError: continue
        items.append((v, k)); tot += v
for v, k in sorted(items, reverse=True)[:8]:
    print("%6.1f%% %s" % (100 * v / tot, k.replace('smsp__pcsamp_warps_issue_stalled_', '')))
rows = list(csv.reader(open(src)))
hdr = rows[1]; data = rows[2:]
ismp, iex = hdr.index("# Samples"), hdr.index("Instructions Executed")
tot = sum(int(r[ismp] or 0) for r in data)
for a in range(0, len(data), 1500):
    s = sum(int(r[ismp] or 0) for r in data[a:a + 1500]); e = sum(int(r[iex] or 0) for r in data[a:a + 1500])
    if e: print("instr %5d..: samples %.1f%%  executed %.2fG" % (a, 100 * s / tot, e / 1e9))
