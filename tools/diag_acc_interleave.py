"""GPU diagnostic: k-loop cost of narrow tcgen05 tiles with one TMEM accumulator vs. interleaved accumulators
(gemm.cu acc_count): CUDA-graph timed (20 launches per graph, L2 flushed between graphs), split-K off."""
import os, sys
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


L.tnt_debug_gemm_split_k(0)
for (M, N, K) in ((256, 1024, 1024), (256, 1024, 4096), (256, 4096, 1024), (2048, 1024, 1024), (2048, 4096, 1024), (2048, 1024, 4096)):
    a = torch.randn(M, K, device='cuda').bfloat16(); w = (torch.randn(N, K, device='cuda') / K ** 0.5).bfloat16(); b = torch.randn(N, device='cuda')
    ref = a.float() @ w.float().t() + b
    row = []
    for single in (1, 0):
        L.tnt_debug_gemm_single_acc(single)
        c = ops.linear(a, w, b)
        err = ((c.float() - ref).abs().max() / ref.abs().max()).item()
        row.append('%s %6.1f us (err %.1e)' % ('1 acc' if single else 'interleaved', t(a, w, b), err))
    print('M=%d N=%d K=%d: %s' % (M, N, K, ' | '.join(row)), flush=True)
L.tnt_debug_gemm_single_acc(0)
L.tnt_debug_gemm_split_k(-1)
