"""GPU diagnostic: host -> device copy of the article tensor [S=512, B=64, E=1024] bf16 (67 MB padded, ~42 MB live):
one contiguous copy vs the live-row 2-D copies (one per caption) on one stream vs the captions split over two streams."""
import os, sys
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import torch
from news_image_caption_b200 import ops, _lib
from oracle import synth
_lib.require_device()
prev, tgt, ctx = synth.make_batch(64, 32, 50265, ('image', 'article'), seed=1)
art = ctx['article'].bfloat16().contiguous().pin_memory()          # [S, B, E]
mask = ctx['article_mask'].contiguous().pin_memory()                # [B, S]
S, B, E = art.shape
dst = torch.zeros_like(art, device='cuda')
live = int((~mask.bool()).sum()) * E * 2


def timeit(fn, n=20):
    for _ in range(3):
        fn()
    torch.cuda.synchronize()
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    e0.record()
    for _ in range(n):
        fn()
    e1.record()
    torch.cuda.synchronize()
    return e0.elapsed_time(e1) / n


t = timeit(lambda: dst.copy_(art, non_blocking=True))
print('contiguous  %5.1f MB  %.3f ms  %.1f GB/s' % (art.numel() * 2 / 1e6, t, art.numel() * 2 / t / 1e6))
t = timeit(lambda: ops.h2d_live_rows(dst, art, mask))
print('live rows   %5.1f MB  %.3f ms  %.1f GB/s (one 2-D copy per caption, one stream)' % (live / 1e6, t, live / t / 1e6))
s2 = torch.cuda.Stream()
halves = [(0, B // 2), (B // 2, B)]


def two():
    cur = torch.cuda.current_stream()
    s2.wait_stream(cur)
    for i, (a, b) in enumerate(halves):
        with torch.cuda.stream(s2 if i else cur):
            ops.h2d_live_rows(dst[:, a:b], art[:, a:b], mask[a:b])
    cur.wait_stream(s2)


try:
    t = timeit(two)
    print('live rows, captions split over two streams: %.3f ms  %.1f GB/s' % (t, live / t / 1e6))
except Exception as e:
    print('two-stream form failed:', e)
