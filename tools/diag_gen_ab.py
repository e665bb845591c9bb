"""GPU diagnostic: greedy generation (bench shape, B=256 x 32 steps) under the GEMM tuning hooks."""
import os, sys
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import torch
from news_image_caption_b200 import DynamicConvDecoder, _lib, build
from news_image_caption_b200.decoder import random_state_dict
from oracle import synth
build.build(); _lib.require_device()
Lb = _lib.lib()
B, L, V = 256, 32, 50265
flush = torch.empty(256 << 20, dtype=torch.uint8, device='cuda')
dec = DynamicConvDecoder(dtype=torch.bfloat16)
dec.load_reference_state_dict(random_state_dict(seed=0))
dec.prepare()
prev, tgt, ctx = synth.make_batch(B, L, V, seed=1234)
ctx = {k: (v.cuda().bfloat16() if v.is_floating_point() else v.cuda()) for k, v in ctx.items()}
seed = prev[:, :1].cuda()


def timed(n=5):
    fn = lambda: dec.generate(seed, ctx, gen_len=L, trim=False)
    for _ in range(3): fn()
    ts = []
    for _ in range(n):
        flush.fill_(0)
        e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        e0.record(); fn(); e1.record(); torch.cuda.synchronize()
        ts.append(e0.elapsed_time(e1))
    return sorted(ts)[n // 2]


for tag, sk in (('default', -1), ('split-K off', 0), ('split-K 2', 2), ('split-K 4', 4), ('default again', -1)):
    Lb.tnt_debug_gemm_split_k(sk)
    print('%-24s %7.2f ms' % (tag, timed()), flush=True)
Lb.tnt_debug_gemm_split_k(-1)
