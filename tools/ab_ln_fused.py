"""A/B of the fused linear + residual + LayerNorm launches (gemm_ln.cu, tnt_debug_ln_fused) on the bench steps, one process:
C2 / C4 forward + loss, C5 greedy generation."""
import os
import sys

import torch

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import bench  # noqa: E402
from news_image_caption_b200 import DynamicConvDecoder, _lib  # noqa: E402
from news_image_caption_b200.decoder import random_state_dict  # noqa: E402
from tools.sweep_kv_beside import timeit  # noqa: E402


def main():
    L = _lib.lib()
    for wl, sources in (('fwd', ('image', 'article')), ('fwd4', ('image', 'article', 'faces', 'obj'))):
        sd = random_state_dict(seed=0, sources=sources)
        dec = DynamicConvDecoder(sources=list(sources), dtype=torch.bfloat16).load_reference_state_dict(sd)
        prev, tgt, ctx = bench.make_inputs(wl, 1234)
        p, t = prev.cuda(), tgt.cuda()
        c = {k: (v.cuda().bfloat16() if v.is_floating_point() else v.cuda()) for k, v in ctx.items()}
        step = lambda: dec.nll(dec(p, c)[0], t)
        for on in (0, 1, 0, 1):
            L.tnt_debug_ln_fused(on)
            out = step()
            print('%s  ln_fused=%d  %.4f ms/step  loss %.4f' % (wl, on, timeit(step), out[0].item()), flush=True)
        if wl == 'fwd':
            p2, t2, c2 = bench.make_inputs('gen', 1234)
            c2 = {k: (v.cuda().bfloat16() if v.is_floating_point() else v.cuda()) for k, v in c2.items()}
            seed = p2[:, :1].contiguous().cuda()
            gen = lambda: dec.generate(seed, c2, gen_len=32, trim=False)
            for on in (0, 1, 0, 1):
                L.tnt_debug_ln_fused(on)
                lp, ids = gen()
                print('gen  ln_fused=%d  %.3f ms per 32-step generation  sum(ids) %d' % (on, timeit(gen, 5, 2), int(ids.sum())), flush=True)
        L.tnt_debug_ln_fused(-1)
        del dec
        torch.cuda.empty_cache()


if __name__ == '__main__':
    main()
