"""GPU measurement: achieved HBM bandwidth of the memory-bound kernels against the measured copy peak
(MEASURED_PEAKS.json hbm_gbs), at the bench size (B=64, T=32: tiny, latency bound) and at a size that fills the GPU.
Each kernel is timed by replaying a CUDA graph of `reps` launches (no host cost), after a 512 MB L2 flush.
ALGORITHMIC bytes per launch are the figures of DESIGN.md section 4.  Prints one line per (kernel, size)."""
import json
import os
import sys

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import torch  # noqa: E402
from news_image_caption_b200 import build, ops, _lib  # noqa: E402

build.build()
_lib.require_device()
peak = 6550.0
pk = os.path.join(os.path.dirname(os.path.dirname(os.path.abspath(__file__))), 'MEASURED_PEAKS.json')
if os.path.exists(pk):
    peak = json.load(open(pk)).get('hbm_gbs', peak)
flush = torch.empty(512 << 20, dtype=torch.uint8, device='cuda')
bf = torch.bfloat16


def timed(fn, reps=10):
    s = torch.cuda.Stream()
    with torch.cuda.stream(s):
        fn()
        s.synchronize()
        g = torch.cuda.CUDAGraph()
        with torch.cuda.graph(g, stream=s):
            for _ in range(reps):
                fn()
        ts = []
        for _ in range(5):
            flush.fill_(0)
            e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
            e0.record(s)
            g.replay()
            e1.record(s)
            s.synchronize()
            ts.append(e0.elapsed_time(e1) / reps)
    return sorted(ts)[2]


def report(name, size, nbytes, ms):
    gbs = nbytes / (ms * 1e-3) / 1e9
    print('%-34s %-26s %9.2f MB %8.1f us %8.0f GB/s  %5.1f %% of measured peak (%.0f)' %
          (name, size, nbytes / 1e6, ms * 1e3, gbs, 100 * gbs / peak, peak), flush=True)


H, C = 16, 1024
for T, B in ((32, 64), (32, 1024)):
    for K in (3, 31):
        x = torch.randn(T, B, C, device='cuda').to(bf)
        lg = torch.randn(T, B, H * K, device='cuda').to(bf)
        dy = torch.randn(T, B, C, device='cuda').to(bf)
        n = T * B
        report('dynconv_fwd K=%d' % K, 'T=%d B=%d' % (T, B), n * (2 * C + H * K) * 2, timed(lambda: ops.dynconv(x, lg, H, K)))
        report('dynconv_bwd K=%d' % K, 'T=%d B=%d' % (T, B), n * (4 * C + 2 * H * K) * 2,
               timed(lambda: ops.dynconv_bwd(x, lg, dy, H, K)))
    x = torch.randn(T * B, C, device='cuda').to(bf)
    g_, b_ = torch.ones(C, device='cuda'), torch.zeros(C, device='cuda')
    report('layernorm_fwd', 'rows=%d' % (T * B), T * B * 2 * C * 2, timed(lambda: ops.layernorm(x, g_, b_)))
    report('layernorm_bwd', 'rows=%d' % (T * B), T * B * 3 * C * 2, timed(lambda: ops.layernorm_bwd(x, x, g_)))

# decode-step attention over the cached article K/V (all keys live): K and V rows once
for B in (256,):
    S, E = 514, 1024
    q = torch.randn(1, B, E, device='cuda').to(bf)
    k = torch.randn(S, B, E, device='cuda').to(bf)
    v = torch.randn(S, B, E, device='cuda').to(bf)
    report('xattn_step (decode, S=514)', 'B=%d H=16' % B, (2 * S * B * E + 2 * B * E) * 2,
           timed(lambda: ops.xattn(q, k, v, 16, 0.125)))
    ring = torch.zeros(30, B, C, device='cuda').to(bf)
    pos = torch.full((1,), 40, dtype=torch.int32, device='cuda')
    xs = torch.randn(B, C, device='cuda').to(bf)
    lg = torch.randn(B, H * 31, device='cuda').to(bf)
    report('dynconv_step K=31', 'B=%d' % B, (30 * B * C + 3 * B * C + B * H * 31) * 2, timed(lambda: ops.dynconv_step(xs, lg, ring, pos, H, 31)))

# RoBERTa hidden-layer mix (context assembly): 25 layers read once, one written
for B in (8, 64):
    hs = [torch.randn(B, 512, 1024, device='cuda').to(bf) for _ in range(25)]
    w = torch.randn(25).cuda()
    report('context_mix 25 layers', 'B=%d S=512 E=1024' % B, 26 * B * 512 * 1024 * 2, timed(lambda: ops.context_mix(hs, w), reps=3))
