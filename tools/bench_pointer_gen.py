"""GPU diagnostic: greedy generation through the copy heads (pointer / entity-pointer decoders) at the bench shape
(B=256 captions x 32 steps, d=1024, V=50265, bf16) next to the plain decoder; CUDA-event timed, L2 flushed between runs."""
import os, sys
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import torch
from news_image_caption_b200 import (DynamicConvDecoder, DynamicConvDecoderEntityPointer, DynamicConvDecoderPointer, _lib,
                                      build)
from news_image_caption_b200.decoder import random_state_dict
from oracle import synth          # input generator only

build.build(); _lib.require_device()
B, L, V = 256, 32, 50265
flush = torch.empty(256 << 20, dtype=torch.uint8, device='cuda')


def timed(fn, n=5):
    for _ in range(3):
        fn()
    ts = []
    for _ in range(n):
        flush.fill_(0)
        e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        e0.record(); fn(); e1.record(); torch.cuda.synchronize()
        ts.append(e0.elapsed_time(e1))
    return sorted(ts)[n // 2]


def head_weights(dec, names, g):
    for n in names:
        m = getattr(dec, n)
        m.weight_v.data.normal_(0, 0.02, generator=g)
        m.weight_g.data.fill_(1.0)


for kind in ('plain', 'pointer', 'entity_pointer'):
    sources = ('image', 'article', 'entity') if kind == 'entity_pointer' else ('image', 'article')
    cls = {'plain': DynamicConvDecoder, 'pointer': DynamicConvDecoderPointer, 'entity_pointer': DynamicConvDecoderEntityPointer}[kind]
    dec = cls(sources=sources, dtype=torch.bfloat16)
    sd = random_state_dict(seed=0, sources=sources)
    own = dict(dec.named_parameters())
    with torch.no_grad():
        for k, v in sd.items():
            if k in own:
                own[k].copy_(v.to(own[k].device).reshape(own[k].shape))
    g = torch.Generator(device='cuda').manual_seed(1)
    if kind != 'plain':
        head_weights(dec, ('project_p_gens',) + (('project_first_p_gens',) if kind == 'entity_pointer' else ()), g)
    dec._dirty = True
    dec.prepare()
    prev, tgt, ctx = synth.make_batch(B, L, V, sources, seed=1234)
    ctx = {k: (v.cuda().bfloat16() if v.is_floating_point() else v.cuda()) for k, v in ctx.items()}
    seed = prev[:, :1].cuda()
    src = 'entity' if kind == 'entity_pointer' else 'article'
    copy_ids = torch.randint(3, V, (B, ctx[src].shape[0]), device='cuda')
    n0 = _lib.lib().tnt_launch_count()
    if kind == 'plain':
        fn = lambda: dec.generate(seed, ctx, gen_len=L, trim=False)
    else:
        fn = lambda: dec.generate(seed, ctx, copy_ids, gen_len=L, trim=False)
    fn(); torch.cuda.synchronize()
    per_call = _lib.lib().tnt_launch_count() - n0
    ms = timed(fn)
    print('%-15s B=%d x %d steps: %7.2f ms  %8.0f captions/s  (%d kernel launches per call)' % (kind, B, L, ms, B / ms * 1e3, per_call),
          flush=True)
