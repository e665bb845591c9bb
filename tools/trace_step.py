"""Ordered device timeline of ONE steady-state step of a bench workload (torch.profiler / CUPTI, one GPU): every kernel /
memset / copy with its stream, start offset, duration and the idle gap before it, then per-kernel totals.  In-situ (warm
cache, PDL overlap visible) -- what the serialised ncu launch list cannot show.  A profiler run is a diagnosis, never a
benchmark number.
  python tools/trace_step.py --workload fwd|fwd4|gen|train [--out file] [--decode-steps 1]"""
import argparse
import collections
import json
import os
import re
import sys
import tempfile

import torch

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import bench  # noqa: E402
from news_image_caption_b200 import DynamicConvDecoder  # noqa: E402
from news_image_caption_b200.decoder import random_state_dict  # noqa: E402


def short(n):
    n = re.sub(r'^void ', '', n).replace('(anonymous namespace)::', '').replace('tnt::', '').replace('__nv_bfloat16', 'bf16')
    return re.sub(r'\(.*$', '', n)[:64]


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument('--workload', default='fwd', choices=sorted(bench.WORKLOADS))
    ap.add_argument('--out', default=None)
    ap.add_argument('--max-lines', type=int, default=500)
    ap.add_argument('--simt-attention', action='store_true', help='force the SIMT attention kernels (diagnosis)')
    a = ap.parse_args()
    w = bench.WORKLOADS[a.workload]
    if a.simt_attention:
        from news_image_caption_b200 import _lib
        _lib.lib().tnt_debug_force_simt_attention(1)
    sd = random_state_dict(seed=0, sources=w['sources'])
    dec = DynamicConvDecoder(sources=w['sources'], dtype=torch.bfloat16).load_reference_state_dict(sd)
    prev, tgt, ctx = bench.make_inputs(a.workload, 1234)
    p, t = prev.cuda(), tgt.cuda()
    c = {k: (v.cuda().bfloat16() if v.is_floating_point() else v.cuda()) for k, v in ctx.items()}
    if a.workload == 'train':
        from news_image_caption_b200.trainer import Trainer
        tr = Trainer(dec, t_total=100000)
        dec.set_dropout(dropout=0.1, input_dropout=0.1, weight_dropout=0.1, attention_dropout=0.1, seed=1234)
        step = lambda: tr.step(p, t, c)
    elif a.workload == 'gen':
        seed = p[:, :1].contiguous()
        step = lambda: dec.generate(seed, c, gen_len=bench.T, trim=False)
    else:
        step = lambda: dec.nll(dec(p, c)[0], t)
    for _ in range(5):
        step()
    torch.cuda.synchronize()
    from torch.profiler import ProfilerActivity, profile
    with profile(activities=[ProfilerActivity.CUDA]) as prof:
        for _ in range(3):
            step()
        torch.cuda.synchronize()
    f = tempfile.mktemp(suffix='.json')
    prof.export_chrome_trace(f)
    ev = [e for e in json.load(open(f))['traceEvents'] if e.get('cat') in ('kernel', 'gpu_memcpy', 'gpu_memset')]
    ev.sort(key=lambda e: e['ts'])
    # a step starts at the first kv_extra_rows launch of its K/V projection (one per source, back to back); a decode step
    # of the generation ends with its gen_update_kernel
    if a.workload == 'gen':
        marks = [i + 1 for i, e in enumerate(ev) if 'gen_update_kernel' in e['name']]
        per = len(marks) // 3
        i0, i1 = marks[per + 10], marks[per + 11]
    else:
        marks = [i for i, e in enumerate(ev) if 'kv_extra_rows' in e['name'] and
                 not any('kv_extra_rows' in m['name'] for m in ev[max(0, i - 4):i])]
        i0, i1 = marks[1], marks[2]
    step_ev = ev[i0:i1]
    t0 = step_ev[0]['ts']
    wall = ev[i1]['ts'] - t0
    out = ['%s: one step = %d device ops, %.1f us wall (first op of the step -> first op of the next)' % (a.workload, len(step_ev), wall)]
    # under PDL a kernel's recorded duration includes its wait for the predecessor: `excl_us` = the time by which this op
    # moved the stream's completion front (end - max(previous end, start)) is what the op really costs the step
    out.append('%9s %8s %8s %7s %4s  %s' % ('start_us', 'dur_us', 'excl_us', 'gap_us', 'strm', 'op'))
    end = t0
    agg = collections.OrderedDict()
    for i, e in enumerate(step_ev):
        s = e['args'].get('stream')
        gap = e['ts'] - end
        fin = e['ts'] + e['dur']
        excl = max(0.0, fin - max(end, e['ts']))
        if i < a.max_lines:
            out.append('%9.1f %8.1f %8.1f %7.1f %4s  %s' % (e['ts'] - t0, e['dur'], excl, gap, s, short(e['name'])))
        end = max(end, fin)
        x = agg.setdefault(short(e['name']), [0, 0.0, 0.0])
        x[0] += 1
        x[1] += e['dur']
        x[2] += excl
    out.append('--- per kernel: count, summed duration us, summed exclusive us, share of the wall (%.1f us)' % wall)
    for nm, (n, us, ex) in sorted(agg.items(), key=lambda kv: -kv[1][2]):
        out.append('  %-66s %4d %9.1f %9.1f %6.1f%%' % (nm, n, us, ex, 100 * ex / wall))
    txt = '\n'.join(out)
    print(txt)
    if a.out:
        open(a.out, 'w').write(txt + '\n')


if __name__ == '__main__':
    main()
