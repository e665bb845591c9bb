"""GPU diagnostic: tcgen05 GEMM vs torch.matmul per shape (error + time).  Not a test; prints a table."""
import sys, os, time
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import torch
from news_image_caption_b200 import build, ops, _lib
build.build(); _lib.require_device()
shapes = [(128, 64, 64), (128, 256, 64), (256, 256, 128), (2048, 2048, 1024), (2048, 1024, 4096), (2048, 496, 1024),
          (32768, 8192, 1024), (131072, 8192, 1024), (8192, 8192, 8192), (3136, 8192, 2048), (256, 1024, 1024), (2048, 5002, 1024), (2048, 30265, 1024)]
for M, N, K in shapes:
    g = torch.Generator().manual_seed(1)
    a = torch.randn(M, K, generator=g).bfloat16().cuda()
    w = (torch.randn(N, K, generator=g) / K ** 0.5).bfloat16().cuda()
    try:
        c = ops.linear(a, w); torch.cuda.synchronize()
    except Exception as e:
        print(M, N, K, 'FAILED', e); continue
    ref = (a.float() @ w.float().t())
    err = ((c.float() - ref).abs().max() / ref.abs().max()).item()
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    for _ in range(3): ops.linear(a, w)
    e0.record()
    for _ in range(10): ops.linear(a, w)
    e1.record(); torch.cuda.synchronize()
    ms = e0.elapsed_time(e1) / 10
    ms_cg = {}
    for cg in (1, 2):
        _lib.lib().tnt_debug_gemm_cta_group(cg)
        for _ in range(3): ops.linear(a, w)
        e0.record()
        for _ in range(10): ops.linear(a, w)
        e1.record(); torch.cuda.synchronize()
        ms_cg[cg] = e0.elapsed_time(e1) / 10
    _lib.lib().tnt_debug_gemm_cta_group(0)
    e0.record()
    for _ in range(10): torch.matmul(a, w.t())
    e1.record(); torch.cuda.synchronize()
    ms_t = e0.elapsed_time(e1) / 10
    print('M=%6d N=%6d K=%5d  rel_err %.2e  ours %.3f ms %.0f TF/s (1-CTA %.3f ms %.0f TF/s, pair %.3f ms %.0f TF/s) | cublas %.3f ms %.0f TF/s' %
          (M, N, K, err, ms, 2e-9 * M * N * K / ms, ms_cg[1], 2e-9 * M * N * K / ms_cg[1], ms_cg[2], 2e-9 * M * N * K / ms_cg[2],
           ms_t, 2e-9 * M * N * K / ms_t), flush=True)
