"""A/B of the dynamic tile scheduler of the single-CTA tcgen05 GEMM on the bench steps (1 GPU): fwd and train ms/step with
TNT_GEMM_DYNAMIC on / off in the same process (tnt_debug_gemm_dynamic)."""
import os
import sys

import torch

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import bench  # noqa: E402
from news_image_caption_b200 import DynamicConvDecoder, _lib  # noqa: E402
from news_image_caption_b200.decoder import random_state_dict  # noqa: E402
from news_image_caption_b200.trainer import Trainer  # noqa: E402


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


def main():
    L = _lib.lib()
    sd = random_state_dict(seed=0)
    dec = DynamicConvDecoder(dtype=torch.bfloat16).load_reference_state_dict(sd)
    prev, tgt, ctx = bench.make_inputs('fwd', 1234)
    p, t = prev.cuda(), tgt.cuda()
    c = {k: (v.cuda().bfloat16() if v.is_floating_point() else v.cuda()) for k, v in ctx.items()}
    fwd = lambda: dec.nll(dec(p, c)[0], t)
    p2, t2, c2 = bench.make_inputs('gen', 1234)
    c2 = {k: (v.cuda().bfloat16() if v.is_floating_point() else v.cuda()) for k, v in c2.items()}
    seed = p2[:, :1].contiguous().cuda()
    gen = lambda: dec.generate(seed, c2, gen_len=32, trim=False)
    for rnd in range(2):
        for on in (1, 0):
            L.tnt_debug_gemm_dynamic(on)
            print('dynamic=%d fwd %.4f ms  gen %.3f ms' % (on, timeit(fwd), timeit(gen, 5, 2)), flush=True)
    dect = DynamicConvDecoder(dtype=torch.bfloat16).load_reference_state_dict(sd)
    tr = Trainer(dect, t_total=100000)
    dect.set_dropout(dropout=0.1, input_dropout=0.1, weight_dropout=0.1, attention_dropout=0.1, seed=1)
    trn = lambda: tr.step(p, t, c)
    for rnd in range(2):
        for on in (1, 0):
            L.tnt_debug_gemm_dynamic(on)
            print('dynamic=%d train %.4f ms' % (on, timeit(trn, 20, 5)), flush=True)


if __name__ == '__main__':
    main()
