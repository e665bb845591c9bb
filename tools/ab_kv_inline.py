"""A/B of the inline per-layer K/V projection of the large sources (engine_fwd.cu kv_layer, tnt_debug_kv_inline) against the
one-GEMM-per-source form on the C2 / C4 forward + loss steps, in one process."""
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
            L.tnt_debug_kv_inline(on)
            out = step()
            print('%s  kv_inline=%d  %.4f ms/step  loss %.4f' % (wl, on, timeit(step), out[0].item()), flush=True)
        L.tnt_debug_kv_inline(-1)
        del dec
        torch.cuda.empty_cache()


if __name__ == '__main__':
    main()
