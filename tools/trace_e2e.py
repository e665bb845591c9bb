"""Device timeline of the pipelined host-buffer scoring loop (bench.py's e2e leg of the forward workload): per stream, when
the H2D copies of batch i+1 and the kernels of batch i run.  Consecutive copies are collapsed.  Diagnosis, not a benchmark.
  python tools/trace_e2e.py [--out file]"""
import argparse
import json
import os
import sys
import tempfile

import torch

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import bench  # noqa: E402
from news_image_caption_b200 import DynamicConvDecoder  # noqa: E402
from news_image_caption_b200.decoder import random_state_dict  # noqa: E402


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument('--out', default=None)
    a = ap.parse_args()
    w = bench.WORKLOADS['fwd']
    dec = DynamicConvDecoder(sources=w['sources'], dtype=torch.bfloat16).load_reference_state_dict(random_state_dict(seed=0, sources=w['sources']))
    prev, tgt, ctx = bench.make_inputs('fwd', 1234)
    pin = lambda t: t.contiguous().pin_memory()
    prev_h, tgt_h = pin(prev), pin(tgt)
    ctx_h = {k: pin(v.bfloat16() if v.is_floating_point() else v) for k, v in ctx.items()}
    for _ in range(3):
        dec.score_host(prev_h, tgt_h, ctx_h)
    torch.cuda.synchronize()
    from torch.profiler import ProfilerActivity, profile
    n = 8
    with profile(activities=[ProfilerActivity.CUDA]) as prof:
        dec.score_host_submit(0, prev_h, tgt_h, ctx_h)
        for i in range(1, n):
            dec.score_host_submit(i & 1, prev_h, tgt_h, ctx_h)
            dec.score_host_wait((i - 1) & 1)
        dec.score_host_wait((n - 1) & 1)
        torch.cuda.synchronize()
    f = tempfile.mktemp(suffix='.json')
    prof.export_chrome_trace(f)
    ev = [e for e in json.load(open(f))['traceEvents'] if e.get('cat') in ('kernel', 'gpu_memcpy', 'gpu_memset')]
    ev.sort(key=lambda e: e['ts'])
    t0 = ev[0]['ts']
    out = []
    # collapse runs of copies on one stream; report kernels only at step boundaries (first / last kernel of a forward)
    run = None
    firsts = [e for e in ev if 'kv_extra_rows' in e['name']]
    for e in ev:
        s = e['args'].get('stream')
        if e['cat'] != 'kernel':
            if run and run['s'] == s and e['ts'] - run['end'] < 50:
                run['end'] = e['ts'] + e['dur']
                run['n'] += 1
                run['bytes'] += e['args'].get('bytes', 0)
            else:
                if run:
                    out.append('%9.1f %9.1f  stream %-3s %4d copies / memsets, %.1f MB' % (run['t'] - t0, run['end'] - run['t'], run['s'], run['n'], run['bytes'] / 1e6))
                run = {'s': s, 't': e['ts'], 'end': e['ts'] + e['dur'], 'n': 1, 'bytes': e['args'].get('bytes', 0)}
        elif 'kv_extra_rows' in e['name'] or 'nll_reduce' in e['name']:
            out.append('%9.1f %9.1f  stream %-3s %s' % (e['ts'] - t0, e['dur'], s, e['name'][:60]))
    if run:
        out.append('%9.1f %9.1f  stream %-3s %4d copies / memsets, %.1f MB' % (run['t'] - t0, run['end'] - run['t'], run['s'], run['n'], run['bytes'] / 1e6))
    out.sort(key=lambda l: float(l.split()[0]))
    txt = 'start_us    dur_us\n' + '\n'.join(out)
    print(txt)
    if a.out:
        open(a.out, 'w').write(txt + '\n')


if __name__ == '__main__':
    main()
