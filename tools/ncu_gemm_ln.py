"""The three linear + residual + LayerNorm sites of a decoder layer at the bench's B*T = 2048 rows, fused (gemm_ln.cu) and as
the two launches they replace -- the command `ncu --set full -k regex:"gemm_ln_kernel|gemm_tc_kernel|layernorm_fwd"` wraps;
without ncu it prints back-to-back timings of both forms."""
import os
import sys

import torch

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
from news_image_caption_b200 import _lib, ops  # noqa: E402

_lib.require_device()
g = torch.Generator().manual_seed(1)


def site(M, N, K, groups):
    a = torch.randn(M, groups * K, generator=g).bfloat16().cuda()
    w = (torch.randn(groups, N, K, generator=g) / K ** 0.5).bfloat16().cuda()
    b = torch.randn(groups, N, generator=g).cuda()
    r = torch.randn(M, N, generator=g).bfloat16().cuda()
    gm, bt = torch.rand(groups, N, generator=g).cuda() + 0.5, torch.randn(groups, N, generator=g).cuda()
    if groups == 1:
        fused = lambda: ops.linear_ln(a, w[0], b[0], r, gm[0], bt[0])
        two = lambda: ops.layernorm(ops.linear(a, w[0], b[0], epilogue=3, residual=r), gm[0], bt[0])
    else:
        fused = lambda: ops.linear_ln(a, w, b, r, gm, bt, groups=groups)
        two = lambda: [ops.layernorm(ops.linear(a[:, s * K:(s + 1) * K], w[s], b[s], epilogue=3, residual=r), gm[s], bt[s]) for s in range(groups)]
    return fused, two


def timeit(f, n=50):
    for _ in range(5):
        f()
    torch.cuda.synchronize()
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    e0.record()
    for _ in range(n):
        f()
    e1.record()
    torch.cuda.synchronize()
    return e0.elapsed_time(e1) / n * 1e3


for M, N, K, G in ((2048, 1024, 1024, 1), (2048, 1024, 1024, 2), (2048, 1024, 4096, 1), (256, 1024, 1024, 1)):
    fused, two = site(M, N, K, G)
    print('M=%d N=%d K=%d groups=%d: fused %.1f us, two launches %.1f us (back to back, host overhead included)' %
          (M, N, K, G, timeit(fused), timeit(two)), flush=True)
