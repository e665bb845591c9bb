"""Sweep of the SM share of the K/V projections when the teacher-forced forward runs them beside the layer chain
(engine_fwd.cu project_kv_beside, tnt_debug_kv_beside): C2 (2 sources, B=64) and C4 (4 sources, B=128) ms per step for
0 (= the plain order: projections first) and a range of shares, in one process; optionally the pipelined host-buffer loop."""
import os
import sys
import time

import torch

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import bench  # noqa: E402
from news_image_caption_b200 import DynamicConvDecoder, _lib  # noqa: E402
from news_image_caption_b200.decoder import random_state_dict  # noqa: E402


def timeit(step, n=30, w=5):
    for _ in range(w):
        step()
    torch.cuda.synchronize()
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    e0.record()
    for _ in range(n):
        step()
    e1.record()
    torch.cuda.synchronize()
    return e0.elapsed_time(e1) / n


def e2e(dec, prev_h, tgt_h, ctx_h, n=20):
    reps = []
    for _ in range(3):
        torch.cuda.synchronize()
        t0 = time.perf_counter()
        dec.score_host_submit(0, prev_h, tgt_h, ctx_h)
        for i in range(1, n):
            dec.score_host_submit(i & 1, prev_h, tgt_h, ctx_h)
            dec.score_host_wait((i - 1) & 1)
        dec.score_host_wait((n - 1) & 1)
        reps.append((time.perf_counter() - t0) / n * 1e3)
    return sorted(reps)[1]


def main():
    shares = [int(a) for a in sys.argv[1:]] or [0, 56, 64, 72, 80, 88, 96, 104, 112, 120, 132, 0, 88]
    L = _lib.lib()
    for wl, sources in (('fwd', ('image', 'article')), ('fwd4', ('image', 'article', 'faces', 'obj'))):
        sd = random_state_dict(seed=0, sources=sources)
        dec = DynamicConvDecoder(sources=list(sources), dtype=torch.bfloat16).load_reference_state_dict(sd)
        prev, tgt, ctx = bench.make_inputs(wl, 1234)
        p, t = prev.cuda(), tgt.cuda()
        c = {k: (v.cuda().bfloat16() if v.is_floating_point() else v.cuda()) for k, v in ctx.items()}
        prev_h, tgt_h = bench.pin(prev), bench.pin(tgt)
        ctx_h = {k: bench.pin(v.bfloat16() if v.is_floating_point() else v) for k, v in ctx.items()}
        step = lambda: dec.nll(dec(p, c)[0], t)
        ref = None
        for s in shares:
            L.tnt_debug_kv_beside(s)
            out = step()
            torch.cuda.synchronize()
            if ref is None:
                ref = out.clone()
            ms = timeit(step)
            print('%s  kv_beside_sms=%3d  %.4f ms/step  e2e %.4f ms/step  loss %.4f (first: %.4f)' %
                  (wl, s, ms, e2e(dec, prev_h, tgt_h, ctx_h), out[0].item(), ref[0].item()), flush=True)
        L.tnt_debug_kv_beside(-1)
        del dec
        torch.cuda.empty_cache()


if __name__ == '__main__':
    main()
