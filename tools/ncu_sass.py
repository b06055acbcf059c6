#!/usr/bin/env python
"""Per-instruction stall-sample shares of one kernel in an .ncu-rep:
python tools/ncu_sass.py rep regex [min_pct [skip]]   (regex on the base name; skip = matching launches to pass over)"""
import csv, subprocess, sys
rep, rx = sys.argv[1], sys.argv[2]
minp = float(sys.argv[3]) if len(sys.argv) > 3 else 1.0
skip = sys.argv[4] if len(sys.argv) > 4 else '0'
raw = subprocess.run(['ncu', '-i', rep, '--page', 'source', '--csv', '--kernel-name', 'regex:' + rx, '--launch-skip', skip,
                      '--launch-count', '1'], capture_output=True, text=True).stdout
rows = list(csv.reader(raw.splitlines()))
hidx = [i for i, r in enumerate(rows) if r and r[0] == 'Address']
h = rows[hidx[0]]
si, j, ie = h.index('Source'), h.index('# Samples'), h.index('Instructions Executed')
end = hidx[1] if len(hidx) > 1 else len(rows)
body = [r for r in rows[hidx[0] + 1:end] if len(r) > j and r[j].isdigit()]
tot = sum(int(r[j]) for r in body)
print('total samples', tot)
for r in body:
    v = int(r[j])
    if 100.0 * v / tot >= minp or 'LDG' in r[si] or 'UBLKCP' in r[si] or 'SYNCS' in r[si] or 'BAR' in r[si]:
        print('%5.1f%% %8s  %s' % (100.0 * v / tot, r[ie], r[si][:110]))
