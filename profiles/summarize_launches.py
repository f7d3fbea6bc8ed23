"""Aggregate an ncu launch list (gpu__time_duration.sum per launch) by kernel/grid and write a compact copy.
usage: summarize_launches.py launches.csv out_prefix "<command line that was profiled>" """
import csv, re, sys

def short(n):
    n = re.sub(r'void |\(.*', '', n)
    return n.replace('retr::', '').replace('(anonymous namespace)::', '').replace('unnamed>::', '').lstrip('<')

src, prefix, cmd = sys.argv[1], sys.argv[2], sys.argv[3]
rows = list(csv.DictReader(l for l in open(src) if not l.startswith('==')))
with open(prefix + '_launches.csv', 'w') as f:
    f.write('# ' + cmd + '\n# cold-cache serialised per-launch times: compare SHARES, not absolutes\nid,kernel,grid,block,ns\n')
    for r in rows:
        f.write("%s,%s,%s,%s,%s\n" % (r['ID'], short(r['Kernel Name']), r['Grid Size'].replace(',', ' '),
                                      r['Block Size'].replace(',', ' '), r['Metric Value']))
agg, tot = {}, 0.0
for r in rows:
    key = (short(r['Kernel Name']), r['Grid Size'])
    us = float(r['Metric Value']) / 1000
    a = agg.setdefault(key, [0, 0.0]); a[0] += 1; a[1] += us; tot += us
with open(prefix + '_summary.txt', 'w') as f:
    f.write(cmd + '\nper kernel/grid; serialised cold-cache durations -> shares\n')
    f.write('total %.1f ms over %d launches\n' % (tot / 1000, len(rows)))
    for k, (c, us) in sorted(agg.items(), key=lambda kv: -kv[1][1])[:28]:
        f.write("%-58s grid=%-16s n=%5d tot=%8.2f ms avg=%8.1f us share=%.3f\n" % (k[0][:58], k[1], c, us / 1000, us / c, us / tot))
print(open(prefix + '_summary.txt').read())
