"""final* model (custom_v4.py:484: 6 layers, d=1024, 8 heads of 128, ffn 4096, 196 image tokens, 512 article tokens) forward +
loss at the script's batch 12, bf16: tcgen05 D=128 attention vs the SIMT kernel (TNT_XATTN_TC_D128=0), and one training step."""
import os
import sys

import torch

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
from news_image_caption_b200 import transformer_v2 as tv  # noqa: E402
from tools.ab_dynamic import timeit  # noqa: E402

g = torch.Generator().manual_seed(1)
V, B, T = 50266, 12, 32
m = tv.make_model_news(V, 6, 1024, 4096, 8, img_dim=1024, sent_dim=1024, dtype=torch.bfloat16)
with torch.no_grad():
    for k, p in m.named_parameters():
        if p.dim() > 1:
            p.copy_((torch.randn(p.shape, generator=g) * p.shape[-1] ** -0.5).cuda())
        elif k.endswith('a_2'):
            p.fill_(1.0)
ids = torch.randint(0, V, (B, T), generator=g).cuda()
y = torch.randint(2, V, (B, T), generator=g).cuda()
mem = {'image': torch.randn(B, 196, 1024, generator=g).cuda(), 'article': torch.randn(B, 512, 1024, generator=g).cuda(),
       'article_mask': (torch.rand(B, 1, 512, generator=g) > 0.3).cuda()}
tm = tv.subsequent_mask(T).cuda()
lc = tv.SimpleLossCompute(m.generator, 1)
with torch.no_grad():
    fwd = timeit(lambda: lc(m.forward(mem, ids, None, tm), y, float(B * T)), 10, 3)
m.enable_training()
opt = tv.FlatAdam(m.parameters())


def train():
    out = m.forward(mem, ids, None, tm)
    loss, node = lc(out, y, float(B * T))
    opt.zero_grad()
    node.backward()
    opt.step()


trn = timeit(train, 5, 2)
print('final* B=12 T=32 bf16 [TNT_XATTN_TC_D128=%s]: forward+loss %.3f ms (%.0f tokens/s), training step %.2f ms' % (
    os.environ.get('TNT_XATTN_TC_D128', '1'), fwd, B * T / fwd * 1e3, trn))
