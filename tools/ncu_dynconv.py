"""ncu target: one launch each of the DynamicConv forward / backward (bf16) at the GPU-filling size of
tools/bench_hbm_kernels.py (T=32, B=1024, C=1024, H=16) for K = 31 and K = 3.
  ncu --set full --clock-control none --import-source on -k regex:dynconv -o gpurun_out/dynconv python tools/ncu_dynconv.py"""
import os
import sys

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import torch  # noqa: E402
from news_image_caption_b200 import build, ops, _lib  # noqa: E402

build.build()
_lib.require_device()
T, B, C, H = 32, 1024, 1024, 16
x = torch.randn(T, B, C, device='cuda', dtype=torch.bfloat16)
dy = torch.randn(T, B, C, device='cuda', dtype=torch.bfloat16)
for K in (31, 3):
    lg = torch.randn(T, B, H * K, device='cuda', dtype=torch.bfloat16)
    ops.dynconv(x, lg, H, K)
    ops.dynconv_bwd(x, lg, dy, H, K)
torch.cuda.synchronize()
print('done')
