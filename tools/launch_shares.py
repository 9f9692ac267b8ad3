"""Per-kernel shares of an ncu launch list: python tools/launch_shares.py launches.csv out.csv "header comment"
(the --metrics gpu__time_duration.sum --clock-control none pass of /opt/skills/guides/B200_PROFILING.md)."""
import collections
import csv
import sys

src, out, note = sys.argv[1], sys.argv[2], sys.argv[3] if len(sys.argv) > 3 else ''
rows = [r for r in csv.reader(l for l in open(src) if not l.startswith('==')) if r]
hdr = rows[0]
ki, vi, ui = hdr.index('Kernel Name'), hdr.index('Metric Value'), hdr.index('Metric Unit')
tot, cnt = collections.Counter(), collections.Counter()
for r in rows[1:]:
    if len(r) <= vi:
        continue
    v = float(r[vi].replace(',', ''))
    us = v / 1e3 if r[ui] in ('ns', 'nsecond') else v * 1e3 if r[ui] in ('ms', 'msecond') else v
    name = r[ki].split('(')[0]
    tot[name] += us
    cnt[name] += 1
allus = sum(tot.values())
with open(out, 'w') as f:
    for l in note.split('\\n'):
        if l:
            f.write('# ' + l + '\n')
    f.write('kernel,launches,total_us,avg_us,share\n')
    for k, v in tot.most_common():
        f.write('%s,%d,%.1f,%.1f,%.4f\n' % (k, cnt[k], v, v / cnt[k], v / allus))
print(open(out).read()[:1800])
