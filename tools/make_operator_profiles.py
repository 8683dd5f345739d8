"""profiles/r01_per_operator_sf<N>.md from the outputs of tools/profile_queries.py --serial --summary (development aid).
usage: python tools/make_operator_profiles.py <prof.txt> <ops.md> <sf>"""
import os
import re
import sys

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
prof, ops, sf = open(sys.argv[1]).read(), open(sys.argv[2]).read(), int(sys.argv[3])
heads = [l for l in prof.splitlines() if l.startswith('==') or l.startswith('warehouse') or l.startswith('registered')]
walls = [float(re.search(r'wall ([0-9.]+) ms', l).group(1)) for l in heads if l.startswith('==')]
out = ['# Round 1 - per-operator achieved bandwidth, SF%d (B200, one GPU, encoded pages resident in HBM)' % sf, '',
       'Command: `python tools/profile_queries.py --sf %d --subset_tables --no_keep_host --reps 3 --serial --summary ...` (CUDA events per '
       'operator, `profile` option; the last of 3 repetitions is listed; `--serial` keeps every operator on one stream so that an '
       'operator\'s events bracket it alone - the production schedule overlaps build sides and the base scan on separate streams and '
       'is faster end to end).' % sf,
       'Seeded synthetic warehouse, only the 8 tables the six queries read. Algorithmic bytes per operator as defined in DESIGN.md section 3 '
       '(rows read back from the device after the query x bytes of the column slots the operator reads, plus what it writes).',
       'Profile mode records two events per operator, which adds a few microseconds of gap per operator; wall = host wall-clock of the whole '
       '`nds_engine_execute` call.', '', '```'] + heads + ['sum of the six queries (one power run, single stream): %.3f ms' % sum(walls), '```', '', ops]
if sf >= 100:
    q7 = [l for l in prof.split('== query7')[1].split('== query19')[0].splitlines() if l.strip() and 'phases' not in l and 'routes' not in l]
    out += ['', 'query7 at SF%d (55.4 M fact rows scanned after partition pruning, 8 columns):' % sf, '```'] + q7 + ['```']
open(os.path.join(ROOT, 'profiles', 'r01_per_operator_sf%d.md' % sf), 'w').write('\n'.join(out) + '\n')
print('wrote profiles/r01_per_operator_sf%d.md' % sf)
