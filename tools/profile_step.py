"""One workload's step, a few times, nothing else -- the command ncu wraps (launch lists / --set full captures of the bench
workloads without bench.py's parity / sustained / CPU legs).
  python tools/profile_step.py --workload fwd|fwd4|gen|train [--steps 2] [--warmup 3]"""
import argparse
import os
import sys

import torch

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import bench  # noqa: E402
from news_image_caption_b200 import DynamicConvDecoder  # noqa: E402
from news_image_caption_b200.decoder import random_state_dict  # noqa: E402


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument('--workload', default='fwd', choices=sorted(bench.WORKLOADS))
    ap.add_argument('--steps', type=int, default=2)
    ap.add_argument('--warmup', type=int, default=3)
    a = ap.parse_args()
    w = bench.WORKLOADS[a.workload]
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
    for _ in range(a.warmup):
        step()
    torch.cuda.synchronize()
    from news_image_caption_b200 import _lib
    n0 = _lib.lib().tnt_launch_count()
    for _ in range(a.steps):
        step()
    torch.cuda.synchronize()
    print('launches per step: %d' % ((_lib.lib().tnt_launch_count() - n0) // a.steps))


if __name__ == '__main__':
    main()
