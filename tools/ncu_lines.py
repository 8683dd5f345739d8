"""Per-CUDA-source-line totals (instructions executed, stall samples) of one kernel launch in an ncu report
captured with --import-source on and compiled with -lineinfo.
usage: python tools/ncu_lines.py report.ncu-rep [launch index] [top N]"""
import collections
import csv
import subprocess
import sys

rep = sys.argv[1]
idx = int(sys.argv[2]) if len(sys.argv) > 2 else 0
top = int(sys.argv[3]) if len(sys.argv) > 3 else 40
out = subprocess.run(['ncu', '-i', rep, '--page', 'source', '--print-source', 'cuda,sass', '--csv', '--launch-skip', str(idx),
                      '--launch-count', '1'], capture_output=True, text=True).stdout
rows = list(csv.reader(out.splitlines()))
hdr = next(r for r in rows if r and r[0] == 'Line No')
li, si, ie = 0, hdr.index('# Samples'), hdr.index('Instructions Executed')
src_i = 1
sb = hdr.index('stall_barrier')
sl = hdr.index('stall_long_sb')
agg = collections.OrderedDict()
cur_file = ''
for r in rows:
    if r and r[0] == 'File Path':
        cur_file = r[1].split('/')[-1]
    if len(r) != len(hdr) or not r[li].isdigit():
        continue
    key = (cur_file, int(r[li]))
    a = agg.setdefault(key, [0, 0, 0, 0, r[src_i].strip()])
    a[0] += int(r[ie] or 0)
    a[1] += int(r[si] or 0)
    a[2] += int(r[sb] or 0)
    a[3] += int(r[sl] or 0)
ti = sum(a[0] for a in agg.values())
ts = sum(a[1] for a in agg.values())
print('total warp instructions', ti, 'samples', ts)
print('%-18s %7s %7s %7s %7s  %s' % ('file:line', 'inst%', 'smpl%', 'barr%', 'lsb%', 'source'))
for key, a in sorted(agg.items(), key=lambda kv: -kv[1][0])[:top]:
    print('%-18s %6.2f%% %6.2f%% %6.2f%% %6.2f%%  %s' % ('%s:%d' % key, 100 * a[0] / max(ti, 1), 100 * a[1] / max(ts, 1), 100 * a[2] / max(ts, 1),
                                                     100 * a[3] / max(ts, 1), a[4][:90]))
