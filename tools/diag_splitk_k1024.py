"""GPU diagnostic: does the cluster split-K kernel pay for K = 1024 problems with few tiles (decode-step projections at
M = 256, the tap-logit GEMM at M = 2048)?  CUDA-graph timed, S = 0 (persistent kernel) vs forced split factors."""
import sys, os
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import torch
from news_image_caption_b200 import build, ops, _lib
build.build(); _lib.require_device()
L = _lib.lib()
flush = torch.empty(256 << 20, dtype=torch.uint8, device='cuda')
def t(a, w, b):
    s = torch.cuda.Stream()
    with torch.cuda.stream(s):
        for _ in range(2): ops.linear(a, w, b)
        s.synchronize()
        g = torch.cuda.CUDAGraph()
        with torch.cuda.graph(g, stream=s):
            for _ in range(20): ops.linear(a, w, b)
        ts = []
        for _ in range(5):
            flush.fill_(0)
            e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
            e0.record(s); g.replay(); e1.record(s); s.synchronize()
            ts.append(e0.elapsed_time(e1) / 20 * 1e3)
    return sorted(ts)[2]
for (M, N, K) in ((256, 1024, 1024), (256, 2048, 1024), (256, 496, 1024), (256, 48, 1024), (2048, 48, 1024), (2048, 112, 1024),
                  (2048, 240, 1024), (2048, 496, 1024), (2048, 1024, 1024), (2048, 1024, 2048), (2048, 1024, 4096)):
    a = torch.randn(M, K, device='cuda').bfloat16(); w = (torch.randn(N, K, device='cuda') / K ** 0.5).bfloat16(); b = torch.randn(N, device='cuda')
    row = []
    for S in (0, 2, 4, 8):
        L.tnt_debug_gemm_split_k(S)
        row.append('S=%d %5.1f' % (S, t(a, w, b)))
    L.tnt_debug_gemm_split_k(-1)
    print('M=%d N=%d K=%d us: %s' % (M, N, K, ' | '.join(row)), flush=True)
