"""GPU diagnostic: the decode-step (M=256) GEMM k-loop: time vs K and vs smem ring depth (CUDA-graph timed)."""
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
M, N = 256, 1024
for K in (512, 1024, 2048, 4096, 8192):
    a = torch.randn(M, K, device='cuda').bfloat16(); w = (torch.randn(N, K, device='cuda') / K ** 0.5).bfloat16(); b = torch.randn(N, device='cuda')
    row = []
    for bn in (32, 128):
        for st in (2, 4, 8):
            L.tnt_debug_gemm_tile_n(bn); L.tnt_debug_gemm_cta_group(1); L.tnt_debug_gemm_stages(st)
            row.append('BN%d/st%d %5.1f' % (bn, st, t(a, w, b)))
    L.tnt_debug_gemm_tile_n(0); L.tnt_debug_gemm_cta_group(0); L.tnt_debug_gemm_stages(0)
    print('M=256 N=1024 K=%5d us: %s' % (K, ' | '.join(row)), flush=True)
# cluster split-K (gemm_splitk.cu) against the persistent kernel at its automatic tile
for (M, N, K) in ((256, 1024, 2048), (256, 1024, 4096), (256, 4096, 1024), (256, 1024, 8192), (64, 1024, 4096), (512, 1024, 4096)):
    a = torch.randn(M, K, device='cuda').bfloat16(); w = (torch.randn(N, K, device='cuda') / K ** 0.5).bfloat16(); b = torch.randn(N, device='cuda')
    row = []
    for S in (0, -1, 2, 3, 4, 6, 8):
        L.tnt_debug_gemm_split_k(S)
        row.append('S=%d %5.1f' % (S, t(a, w, b)))
    L.tnt_debug_gemm_split_k(-1)
    print('split-K M=%d N=%d K=%d us: %s' % (M, N, K, ' | '.join(row)), flush=True)
