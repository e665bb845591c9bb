"""GPU diagnostic: the article K/V projection GEMM (M=32768 N=8192 K=1024) under the conditions of the bench step:
with / without bias, single-CTA vs CTA-pair, back-to-back vs cold L2 (a 512 MB write between launches), sustained."""
import sys, os
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import torch
from news_image_caption_b200 import build, ops, _lib
build.build(); _lib.require_device()
L = _lib.lib()
M, N, K = 32768, 8192, 1024
g = torch.Generator().manual_seed(1)
a = torch.randn(M, K, generator=g).bfloat16().cuda()
w = (torch.randn(N, K, generator=g) / K ** 0.5).bfloat16().cuda()
b = torch.randn(N, generator=g).cuda()
flush = torch.empty(512 << 20, dtype=torch.uint8, device='cuda')

def timed(fn, n, cold=False):
    ev = [(torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)) for _ in range(n)]
    for _ in range(3): fn()
    torch.cuda.synchronize()
    for e0, e1 in ev:
        if cold: flush.fill_(1)
        e0.record(); fn(); e1.record()
    torch.cuda.synchronize()
    ts = sorted(e0.elapsed_time(e1) for e0, e1 in ev)
    return ts[len(ts) // 2], ts[0], ts[-1]

for cg in (1, 2):
    L.tnt_debug_gemm_cta_group(cg)
    for name, fn in (('no bias', lambda: ops.linear(a, w)), ('bias', lambda: ops.linear(a, w, b))):
        for cold in (False, True):
            for n in (10, 300):
                med, lo, hi = timed(fn, n, cold)
                print('cta_group %d %-8s %s n=%3d: median %.3f ms (%.0f TF/s)  min %.3f max %.3f' %
                      (cg, name, 'cold-L2' if cold else 'warm   ', n, med, 2e-9 * M * N * K / med, lo, hi), flush=True)
L.tnt_debug_gemm_cta_group(0)
