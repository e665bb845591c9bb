"""The log-sum-exp GEMMs of the adaptive softmax at the bench's shapes (rows = 64 * 32; head 5002 x 1024, tail clusters
15000 x 256 and 30265 x 64), a few times each -- the command `ncu --set full --import-source on -k regex:gemm_tc_kernel` wraps;
without ncu it prints the launch times (CUDA events)."""
import os
import sys

import torch

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
from news_image_caption_b200 import _lib, ops  # noqa: E402

_lib.require_device()
g = torch.Generator().manual_seed(1)
M = 2048
shapes = [(5002, 1024), (15000, 256), (30265, 64)]
which = int(sys.argv[1]) if len(sys.argv) > 1 else -1
for i, (N, K) in enumerate(shapes):
    if which >= 0 and i != which:
        continue
    x = torch.randn(M, K, generator=g).bfloat16().cuda()
    w = (torch.randn(N, K, generator=g) / K ** 0.5).bfloat16().cuda()
    tgt = torch.randint(0, N, (M,), generator=g).cuda()
    for _ in range(3):
        s = ops.lse(x, w, None, tgt)
    torch.cuda.synchronize()
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    e0.record()
    for _ in range(20):
        s = ops.lse(x, w, None, tgt)
    e1.record()
    torch.cuda.synchronize()
    ref = torch.logsumexp(x.float() @ w.float().t(), -1)
    st = s if isinstance(s, torch.Tensor) else s[0]
    print('N=%d K=%d: %.1f us per call (GEMM + finish)' % (N, K, e0.elapsed_time(e1) * 50), 'stats', tuple(st.shape))
