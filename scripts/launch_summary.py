"""Summarise an `ncu --metrics gpu__time_duration.sum --csv` launch list per kernel name."""
import collections
import csv
import re
import sys


def main(path):
    rows = [r for r in csv.reader(open(path)) if len(r) > 10]
    hdr = rows[0]
    ki, vi = hdr.index('Kernel Name'), hdr.index('Metric Value')
    agg = collections.OrderedDict()
    for r in rows[1:]:
        name = re.sub(r'\(.*', '', r[ki])
        agg.setdefault(name, []).append(float(r[vi].replace(',', '')))
    tot = sum(sum(v) for v in agg.values())
    print('%-52s %5s %12s %10s %7s' % ('kernel', 'n', 'total_us', 'avg_us', 'share'))
    for k, v in agg.items():
        print('%-52s %5d %12.1f %10.1f %6.1f%%' % (k[:52], len(v), sum(v) / 1e3, sum(v) / len(v) / 1e3, 100 * sum(v) / tot))
    print('%-52s %5s %12.1f' % ('TOTAL', '', tot / 1e3))


if __name__ == '__main__':
    main(sys.argv[1])
