"""GPU diagnostic: tile-shape sweep of the tcgen05 GEMM on the decoder's layer shapes, timed by replaying a CUDA graph
of 20 launches (no host launch cost in the number).  Prints us per launch for each (cta_group, BN)."""
import sys, os
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import torch
from news_image_caption_b200 import build, ops, _lib
build.build(); _lib.require_device()
L = _lib.lib()
shapes = [('linear1+GLU', 2048, 1024, _lib.EPI_GLU), ('taps K=31', 496, 1024, 0), ('taps K=3', 48, 1024, 0), ('linear2', 1024, 1024, 0),
          ('Wq 2src', 2048, 1024, 0), ('ctx_fc', 1024, 2048, 0), ('fc1', 4096, 1024, _lib.EPI_RELU), ('fc2', 1024, 4096, 0)]
Ms = [int(a) for a in sys.argv[1:]] or [256, 2048, 4096]
cfgs = [(1, 32), (1, 64), (1, 128), (1, 256), (2, 64), (2, 128), (2, 256)]
flush = torch.empty(256 << 20, dtype=torch.uint8, device='cuda')
for M in Ms:
    for name, N, K, epi in shapes:
        g = torch.Generator().manual_seed(1)
        a = torch.randn(M, K, generator=g).bfloat16().cuda()
        w = (torch.randn(N, K, generator=g) / K ** 0.5).bfloat16().cuda()
        b = torch.randn(N, generator=g).cuda()
        res = []
        for cg, bn in [(0, 0)] + cfgs:
            L.tnt_debug_gemm_cta_group(cg); L.tnt_debug_gemm_tile_n(bn)
            try:
                s = torch.cuda.Stream()
                with torch.cuda.stream(s):
                    for _ in range(2): ops.linear(a, w, b, epi)
                    s.synchronize()
                    gr = torch.cuda.CUDAGraph()
                    with torch.cuda.graph(gr, stream=s):
                        for _ in range(20): ops.linear(a, w, b, epi)
                    gr.replay(); s.synchronize()
                    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
                    ts = []
                    for _ in range(5):
                        flush.fill_(0)
                        e0.record(s); gr.replay(); e1.record(s); s.synchronize()
                        ts.append(e0.elapsed_time(e1) / 20 * 1e3)
                res.append('%s %5.1f' % ('auto' if cg == 0 else '%d/%d' % (cg, bn), sorted(ts)[2]))
            except Exception as e:
                res.append('%d/%d ERR' % (cg, bn))
        L.tnt_debug_gemm_cta_group(0); L.tnt_debug_gemm_tile_n(0)
        print('M=%5d %-12s N=%5d K=%5d us: %s' % (M, name, N, K, ' | '.join(res)), flush=True)
