#!/usr/bin/env python
"""Summarises an `ncu -i X.ncu-rep --page raw --csv` export: one line per launch with the metrics the roofline
discussion uses (duration, DRAM bytes, DRAM / tensor-pipe / L2 / SM utilisation, L2->SM fill bytes).
usage: python tools/summarize_ncu_raw.py raw.csv [name-substring]"""
import csv
import re
import sys

WANT = [('gpu__time_duration.sum', 'us'), ('dram__bytes_read.sum', 'dram_rd'), ('dram__bytes_write.sum', 'dram_wr'),
        ('gpu__dram_throughput.avg.pct_of_peak_sustained_elapsed', 'dram%'),
        ('sm__pipe_tensor_cycles_active.avg.pct_of_peak_sustained_active', 'tensor%'),
        ('lts__throughput.avg.pct_of_peak_sustained_elapsed', 'l2%'),
        ('sm__throughput.avg.pct_of_peak_sustained_elapsed', 'sm%'),
        ('l1tex__m_xbar2l1tex_read_bytes.sum', 'l2->sm'), ('launch__grid_size', 'grid'),
        ('launch__registers_per_thread', 'regs'), ('sm__warps_active.avg.pct_of_peak_sustained_active', 'occ%')]


def short(n):
    n = re.sub(r'^void ', '', n)
    n = re.sub(r'\(.*$', '', n)
    return n.replace('tnt::', '').replace('<unnamed>::', '').replace('__nv_bfloat16', 'bf16')[:48]


def main():
    rows = list(csv.reader(l for l in open(sys.argv[1]) if not l.startswith('==')))
    hdr, units, data = rows[0], rows[1], rows[2:]
    idx = {h: i for i, h in enumerate(hdr)}
    sub = sys.argv[2] if len(sys.argv) > 2 else ''
    print('%-4s %-48s ' % ('id', 'kernel') + ' '.join('%10s' % lab for _, lab in WANT))
    for r in data:
        name = short(r[idx['Kernel Name']])
        if sub not in name:
            continue
        vals = []
        for k, _ in WANT:
            if k in idx:
                u = units[idx[k]]
                v = float(r[idx[k]].replace(',', ''))
                vals.append('%8.2f%s' % (v, {'Mbyte': 'MB', 'Gbyte': 'GB', 'Kbyte': 'KB', 'byte': 'B'}.get(u, '')[:2].ljust(2)))
            else:
                vals.append('%10s' % '-')
        print('%-4s %-48s ' % (r[idx['ID']], name) + ' '.join(vals))


if __name__ == '__main__':
    main()
