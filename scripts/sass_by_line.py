#!/usr/bin/env python
"""Attribute an `ncu --page source --csv` SASS export to source lines.

usage: sass_by_line.py <sass.csv> <nvdisasm -g -c output> <mangled kernel name> [top]
The k-th instruction of the export is matched with the k-th instruction of the kernel's section in the
disassembly (same binary), whose `//## File "...", line N` comments carry the -lineinfo mapping.
"""
import csv, sys, re, collections
sass, dis, kern = sys.argv[1:4]
top = int(sys.argv[4]) if len(sys.argv) > 4 else 40
lines = open(dis).read().split('\n')
start = next(i for i, l in enumerate(lines) if l.startswith('.text.' + kern + ':'))
loc, cur = [], ('?', 0)
for l in lines[start + 1:]:
    if l.startswith('.text.') or l.lstrip().startswith('.section'): break
    m = re.search(r'//## File "([^"]+)", line (\d+)', l)
    if m: cur = (m.group(1).split('/')[-1], int(m.group(2))); continue
    if re.match(r'\s+/\*[0-9a-f]{4,}\*/', l): loc.append((cur, l.split('*/', 1)[1].strip()))
rows = list(csv.reader(open(sass)))
h = next(i for i, r in enumerate(rows) if r and r[0] == 'Address')
hdr = rows[h]; body = [r for r in rows[h + 1:] if len(r) == len(hdr)]
ix = {n: i for i, n in enumerate(hdr)}
print('instructions: export %d, disassembly %d' % (len(body), len(loc)))
agg = collections.defaultdict(lambda: collections.Counter())
tot = collections.Counter()
keys = ['# Samples', 'Instructions Executed', 'stall_barrier', 'stall_long_sb', 'stall_short_sb', 'stall_mio', 'stall_math', 'stall_wait',
        'stall_lg', 'stall_no_inst', 'stall_not_selected', 'stall_selected', 'L1 Wavefronts Shared', 'L1 Wavefronts Shared Ideal', 'L1 Tag Requests Global']
for (lc, ins), r in zip(loc, body):
    for k in keys:
        try: v = float(r[ix[k]])
        except ValueError: v = 0.0
        agg[lc][k] += v; tot[k] += v
print('TOTAL', {k: int(v) for k, v in tot.items()})
print('%-28s %8s %6s %10s | bar long short mio math wait lg noinst notsel sel | shW shIdeal gReq' % ('file:line', 'samples', '%', 'inst'))
for lc, c in sorted(agg.items(), key=lambda kv: -kv[1]['# Samples'])[:top]:
    print('%-28s %8d %5.1f%% %10d | %s' % ('%s:%d' % lc, c['# Samples'], 100 * c['# Samples'] / tot['# Samples'], c['Instructions Executed'],
          ' '.join('%d' % c[k] for k in keys[2:])))
