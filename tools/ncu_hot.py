"""List the hottest SASS instructions (warp-stall samples) of one kernel launch in an ncu report.
usage: python tools/ncu_hot.py report.ncu-rep <launch index> [min share]"""
import csv
import subprocess
import sys

rep, idx = sys.argv[1], int(sys.argv[2])
share = float(sys.argv[3]) if len(sys.argv) > 3 else 0.012
out = subprocess.run(['ncu', '-i', rep, '--page', 'source', '--csv', '--launch-skip', str(idx), '--launch-count', '1'],
                     capture_output=True, text=True).stdout
rows = list(csv.reader(out.splitlines()))
hdr = next(r for r in rows if 'Source' in r and '# Samples' in r)
si, so, ie = hdr.index('# Samples'), hdr.index('Source'), hdr.index('Instructions Executed')
stall_cols = [i for i, h in enumerate(hdr) if h.startswith('stall_') and 'Not Issued' not in h]
data = [r for r in rows if len(r) == len(hdr) and r[si].isdigit()]
tot = sum(int(r[si]) for r in data)
print('total samples', tot, 'instructions', len(data))
for i, r in enumerate(data):
    s = int(r[si])
    if s > tot * share:
        st = sorted([(int(r[c] or 0), hdr[c]) for c in stall_cols], reverse=True)[:2]
        print('%5d %6d %9s  %-64s %s' % (i, s, r[ie], r[so].strip()[:64], st))
