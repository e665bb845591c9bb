#!/usr/bin/env python
"""Summarises an `ncu --metrics gpu__time_duration.sum --csv` launch list: finds the last complete bench step
(from one kv_extra_rows launch pair to the next) and prints per-kernel totals and shares.
usage: python tools/summarize_launches.py gpurun_out/launches.csv [marker-substring]"""
import collections
import csv
import re
import sys


def short(n):
    n = re.sub(r'^void ', '', n)
    n = re.sub(r'\(.*$', '', n)
    n = n.replace('tnt::<unnamed>::', '').replace('tnt::', '').replace('__nv_bfloat16', 'bf16')
    return n[:80]


def main():
    path = sys.argv[1]
    marker = sys.argv[2] if len(sys.argv) > 2 else 'kv_extra_rows'
    lines = [l for l in open(path) if not l.startswith('==')]
    rows = list(csv.DictReader(lines))
    names = [short(r['Kernel Name']) for r in rows]
    ns = [float(r['Metric Value']) for r in rows]
    # a step starts at a marker launch that has no other marker launch among the 4 launches before it (one per source)
    starts = [i for i, n in enumerate(names) if marker in n and not any(marker in m for m in names[max(0, i - 4):i])]
    if len(starts) < 2:
        print('fewer than two steps found')
        return
    a, b = starts[-2], starts[-1]
    tot = sum(ns[a:b])
    agg = collections.OrderedDict()
    for n, t in zip(names[a:b], ns[a:b]):
        c = agg.setdefault(n, [0, 0.0])
        c[0] += 1
        c[1] += t
    print('step = launches [%d, %d): %d launches, %.1f us total (serialised, cold-cache ncu times)' % (a, b, b - a, tot / 1e3))
    print('%-82s %6s %10s %7s' % ('kernel', 'count', 'us', 'share'))
    for n, (c, t) in sorted(agg.items(), key=lambda kv: -kv[1][1]):
        print('%-82s %6d %10.1f %6.1f%%' % (n, c, t / 1e3, 100 * t / tot))
    big = max(range(a, b), key=lambda i: ns[i])
    print('largest single launch: #%d %s grid %s: %.1f us = %.1f%% of the step' %
          (big, names[big], rows[big]['Grid Size'], ns[big] / 1e3, 100 * ns[big] / tot))


if __name__ == '__main__':
    main()
