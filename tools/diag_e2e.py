"""GPU diagnostic: where the end-to-end (host buffers) step loses time against the device-resident step:
H2D alone, compute alone, both concurrently, and the pipelined score_host loop."""
import os, sys, time
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import torch
from news_image_caption_b200 import DynamicConvDecoder, build, _lib
from news_image_caption_b200.decoder import random_state_dict
from oracle import synth
build.build(); _lib.require_device()
dec = DynamicConvDecoder(dtype=torch.bfloat16); dec.load_reference_state_dict(random_state_dict(seed=0)); dec.prepare()
prev, tgt, ctx = synth.make_batch(64, 32, 50265, ('image', 'article'), seed=1)
pin = lambda t: t.contiguous().pin_memory()
prev_h, tgt_h = pin(prev), pin(tgt)
ctx_h = {k: pin(v.bfloat16() if v.is_floating_point() else v) for k, v in ctx.items()}
ctx_d = {k: v.cuda() for k, v in ctx_h.items()}
prev_d, tgt_d = prev.cuda(), tgt.cuda()
N = 30
def compute():
    X, _ = dec(prev_d, ctx_d, need_attn=False); return dec.nll(X, tgt_d)
side = torch.cuda.Stream()
dst = {k: torch.empty_like(v, device='cuda') for k, v in ctx_h.items()}
def copies():
    with torch.cuda.stream(side):
        for k, v in ctx_h.items(): dst[k].copy_(v, non_blocking=True)
def timeit(fn, n=N):
    for _ in range(3): fn()
    torch.cuda.synchronize(); t0 = time.perf_counter()
    for _ in range(n): fn()
    torch.cuda.synchronize(); return (time.perf_counter() - t0) / n * 1e3
print('compute only        %.3f ms/step' % timeit(compute))
print('H2D only (80 MB)    %.3f ms/step' % timeit(copies))
print('both concurrently   %.3f ms/step' % timeit(lambda: (copies(), compute())))
t0 = time.perf_counter()
for _ in range(N): compute()
t_enq = (time.perf_counter() - t0) / N * 1e3; torch.cuda.synchronize()
print('host enqueue time of one step (python + C launches) %.3f ms' % t_enq)
for _ in range(2): dec.score_host(prev_h, tgt_h, ctx_h)
torch.cuda.synchronize(); t0 = time.perf_counter()
dec.score_host_submit(0, prev_h, tgt_h, ctx_h)
for i in range(1, N):
    dec.score_host_submit(i & 1, prev_h, tgt_h, ctx_h); dec.score_host_wait((i - 1) & 1)
dec.score_host_wait((N - 1) & 1)
print('pipelined score_host %.3f ms/step' % ((time.perf_counter() - t0) / N * 1e3))
t0 = time.perf_counter()
for _ in range(N): dec.score_host(prev_h, tgt_h, ctx_h)
print('synchronous score_host %.3f ms/step' % ((time.perf_counter() - t0) / N * 1e3))
