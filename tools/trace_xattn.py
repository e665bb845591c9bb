"""clock64 trace of the tcgen05 attention forward (debug hook tnt_debug_xattn_trace): where the softmax warps, the MMA warp
and the TMA producer of CTAs 0..3 spend an item.  The stamped launch is the LAST attention launch of a forward step
(layer 3, article source; need_attn=False so that it is the plain kernel).
  python tools/trace_xattn.py [--source image]"""
import argparse
import ctypes as C
import os
import sys

import torch

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import bench  # noqa: E402
from news_image_caption_b200 import DynamicConvDecoder, _lib  # noqa: E402
from news_image_caption_b200.decoder import random_state_dict  # noqa: E402

ap = argparse.ArgumentParser()
ap.add_argument('--train', action='store_true')
a = ap.parse_args()
sd = random_state_dict(seed=0)
dec = DynamicConvDecoder(dtype=torch.bfloat16).load_reference_state_dict(sd)
prev, tgt, ctx = bench.make_inputs('fwd', 1234)
p = prev.cuda()
c = {k: (v.cuda().bfloat16() if v.is_floating_point() else v.cuda()) for k, v in ctx.items()}
for _ in range(3):
    dec(p, c, need_attn=False)
torch.cuda.synchronize()
buf = torch.zeros(4 * 3 * 8 * 8, dtype=torch.int64, device='cuda')
L = _lib.lib()
L.tnt_debug_xattn_trace.argtypes = [C.c_void_p]
L.tnt_debug_xattn_trace.restype = None
L.tnt_debug_xattn_trace(C.c_void_p(buf.data_ptr()))
dec(p, c, need_attn=False)
torch.cuda.synchronize()
L.tnt_debug_xattn_trace(None)
t = buf.cpu().view(4, 3, 8, 8)
mhz = 1.9
names = {0: ['loop top', 'pre s_ready', 's_ready', 'p1 done', 'p2 done', 'bar3(sums)', 'o_ready', 'epi done'],
         1: ['top', 'q_full', 's_empty', 'QK issued', 'o_empty', 'V0 full', 'P0 ready', 'PV issued'],
         2: ['q_empty', 'ld0', 'ld1', 'ld2', 'ld3', 'ld4', 'ld5', 'ld6+']}
for cta in (0, 1):
    t0 = int(t[cta][t[cta] > 0].min())
    for role in (0, 1, 2):
        print('CTA %d role %d (%s)' % (cta, role, ['softmax warp 2', 'MMA warp', 'TMA producer'][role]))
        for it in range(8):
            row = t[cta, role, it]
            if int(row.max()) == 0:
                continue
            print('  item %d: ' % it + '  '.join('%s %.2f' % (n, (int(v) - t0) / mhz / 1000) if int(v) else '%s -' % n
                                                  for n, v in zip(names[role], row)) + '  (us)')
