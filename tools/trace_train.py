"""Kernel timeline of the data-parallel training step (torch.profiler / CUPTI; run under torchrun): per stream, which kernels
ran when, for ONE steady-state step -- used to see how the gradient exchange overlaps the backward pass.
  python -m torch.distributed.run --nproc-per-node 2 ... tools/trace_train.py [out.txt]
NB a profiler run is a diagnosis, never a benchmark number."""
import json
import os
import sys
import tempfile

import torch
import torch.distributed as dist

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
from news_image_caption_b200 import DynamicConvDecoder  # noqa: E402
from news_image_caption_b200.decoder import random_state_dict  # noqa: E402
from news_image_caption_b200.trainer import Trainer  # noqa: E402
from oracle import synth  # noqa: E402


def main():
    rank, world, local = int(os.environ.get('RANK', 0)), int(os.environ.get('WORLD_SIZE', 1)), int(os.environ.get('LOCAL_RANK', 0))
    torch.cuda.set_device(local)
    if world > 1:
        dist.init_process_group('nccl', device_id=torch.device('cuda', local))
    sd = random_state_dict(seed=0)
    dec = DynamicConvDecoder(dtype=torch.bfloat16).load_reference_state_dict(sd)
    tr = Trainer(dec, t_total=100000, wire=os.environ.get('TNT_WIRE', 'bf16'), sparse_embed=os.environ.get('TNT_SPARSE', '1') == '1')
    dec.set_dropout(dropout=0.1, input_dropout=0.1, weight_dropout=0.1, attention_dropout=0.1, seed=1)
    prev, tgt, ctx = synth.make_batch(64, 32, 50265, ('image', 'article'), seed=1234 + rank)
    p, t = prev.cuda(), tgt.cuda()
    c = {k: (v.cuda().bfloat16() if v.is_floating_point() else v.cuda()) for k, v in ctx.items()}
    for _ in range(5):
        tr.step(p, t, c)
    torch.cuda.synchronize()
    from torch.profiler import ProfilerActivity, profile
    with profile(activities=[ProfilerActivity.CUDA, ProfilerActivity.CPU]) as prof:
        for _ in range(3):
            tr.step(p, t, c)
        torch.cuda.synchronize()
    if rank == 0:
        f = tempfile.mktemp(suffix='.json')
        prof.export_chrome_trace(f)
        ev = [e for e in json.load(open(f))['traceEvents'] if e.get('cat') in ('kernel', 'gpu_memcpy', 'gpu_memset')]
        ev.sort(key=lambda e: e['ts'])
        # the middle step: between the 1st and 2nd bert_adam kernels
        adam = [e for e in ev if 'bert_adam' in e['name']]
        t0, t1 = adam[0]['ts'] + adam[0]['dur'], adam[1]['ts'] + adam[1]['dur']
        step = [e for e in ev if t0 <= e['ts'] < t1]
        streams = sorted({e['args'].get('stream') for e in step})
        out = ['step wall (adam end -> adam end): %.3f ms; %d device ops on streams %s' % ((t1 - t0) / 1e3, len(step), streams)]
        for s in streams:
            es = [e for e in step if e['args'].get('stream') == s]
            busy = sum(e['dur'] for e in es)
            out.append('stream %s: %d ops, busy %.3f ms, first at +%.3f ms, last ends +%.3f ms' % (
                s, len(es), busy / 1e3, (es[0]['ts'] - t0) / 1e3, (es[-1]['ts'] + es[-1]['dur'] - t0) / 1e3))
        main_s = max(streams, key=lambda s: sum(1 for e in step if e['args'].get('stream') == s))
        out.append('--- ops NOT on the main stream %s, and main-stream gaps > 30 us' % main_s)
        prev_end = t0
        for e in step:
            s = e['args'].get('stream')
            if s != main_s:
                out.append('  +%8.3f ms  %7.1f us  stream %-3s %s' % ((e['ts'] - t0) / 1e3, e['dur'], s, e['name'][:90]))
            else:
                if e['ts'] - prev_end > 30:
                    out.append('  +%8.3f ms  GAP %.1f us on the main stream before %s' % ((prev_end - t0) / 1e3, e['ts'] - prev_end, e['name'][:70]))
                prev_end = max(prev_end, e['ts'] + e['dur'])
        # slowest main-stream kernels with overlap info: which were running concurrently with an NCCL kernel
        nccl = [e for e in step if 'nccl' in e['name'].lower()]
        def overl(e):
            return any(n['ts'] < e['ts'] + e['dur'] and e['ts'] < n['ts'] + n['dur'] for n in nccl)
        out.append('--- main-stream time overlapped by NCCL kernels: %.3f ms of %.3f ms' % (
            sum(e['dur'] for e in step if e['args'].get('stream') == main_s and overl(e)) / 1e3,
            sum(e['dur'] for e in step if e['args'].get('stream') == main_s) / 1e3))
        import collections
        import re
        agg = collections.OrderedDict()
        for e in step:
            if e['args'].get('stream') != main_s:
                continue
            nm = re.sub(r'\(.*$', '', e['name'].replace('void ', '')).replace('tnt::', '').replace('(anonymous namespace)::', '')[:70]
            a = agg.setdefault(nm, [0, 0.0, 0.0])
            a[0] += 1
            a[1] += e['dur']
            if overl(e):
                a[2] += e['dur']
        # idle accounting of the main stream: wall - union of its op intervals, and the largest gaps with what follows them
        ms = sorted((e for e in step if e['args'].get('stream') == main_s), key=lambda e: e['ts'])
        gaps, end = [], t0
        for e in ms:
            if e['ts'] > end:
                gaps.append((e['ts'] - end, (end - t0) / 1e3, e['name'][:60]))
            end = max(end, e['ts'] + e['dur'])
        out.append('--- main stream idle: %.3f ms of %.3f ms wall in %d gaps; gaps > 8 us:' % (sum(g[0] for g in gaps) / 1e3, (t1 - t0) / 1e3, len(gaps)))
        for g in sorted(gaps, reverse=True)[:30]:
            if g[0] > 8:
                out.append('  %7.1f us at +%.3f ms before %s' % g)
        out.append('--- main-stream kernels: count, total us, us while an NCCL kernel was running')
        for nm, (c, t, o) in sorted(agg.items(), key=lambda kv: -kv[1][1])[:28]:
            out.append('  %-72s %4d %9.1f %9.1f' % (nm, c, t, o))
        txt = '\n'.join(out)
        print(txt)
        if len(sys.argv) > 1:
            open(sys.argv[1], 'w').write(txt + '\n')
    if world > 1:
        dist.barrier()
        dist.destroy_process_group()


if __name__ == '__main__':
    main()
