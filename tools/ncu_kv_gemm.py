"""The article K/V projection GEMM of the bench step (M = 512*64 = 32768 rows, N = 4 layers * 2 * 1024 = 8192, K = 1024, bias,
bf16) a few times -- the command `ncu --set full --import-source on -k regex:gemm_tc_kernel` wraps."""
import os
import sys

import torch

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
from news_image_caption_b200 import _lib, ops  # noqa: E402

_lib.require_device()
M, N, K = 32768, 8192, 1024
g = torch.Generator().manual_seed(1)
a = torch.randn(M, K, generator=g).bfloat16().cuda()
w = (torch.randn(N, K, generator=g) / K ** 0.5).bfloat16().cuda()
b = torch.randn(N, generator=g).cuda()
for _ in range(4):
    c = ops.linear(a, w, b)
torch.cuda.synchronize()
print('ok', float(c[0, 0]))
