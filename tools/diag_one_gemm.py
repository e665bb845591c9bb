"""GPU diagnostic: one decode-shaped GEMM (M=256, N=1024, K=4096, no split-K) launched a few times -- the target of an
`ncu --set full --import-source on` capture that shows where the producer / MMA threads of the k-loop wait."""
import os, sys
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import torch
from news_image_caption_b200 import build, ops, _lib
build.build(); _lib.require_device()
L = _lib.lib()
M, N, K = 256, 1024, int(os.environ.get('K', 4096))
a = torch.randn(M, K, device='cuda').bfloat16(); w = (torch.randn(N, K, device='cuda') / K ** 0.5).bfloat16(); b = torch.randn(N, device='cuda')
L.tnt_debug_gemm_split_k(0)
for _ in range(6):
    c = ops.linear(a, w, b)
torch.cuda.synchronize()
print('ok', c.float().abs().mean().item())
