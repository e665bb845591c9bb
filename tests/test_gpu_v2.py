"""GPU parity of the `final*/transformer_v2.py` mirror against the golden fixture
frozen from the unmodified reference module and against the oracle at the script's real width."""
import os

import pytest
import torch

from oracle import restate as O

pytestmark = pytest.mark.gpu
G = os.path.join(os.path.dirname(__file__), 'golden')


def rel_err(a, b):
    a, b = a.float().cpu(), b.float().cpu()
    return ((a - b).abs().max() / b.abs().max().clamp_min(1e-6)).item()


def build(cfg, sd, dtype, img_dim, sent_dim):
    from news_image_caption_b200 import build as bld
    from news_image_caption_b200 import transformer_v2 as tv
    bld.build()
    m = tv.make_model_news(cfg['vocab'], cfg['layers'], cfg['d'], cfg['ffn'], cfg['h'], img_dim=img_dim,
                           sent_dim=sent_dim, dtype=dtype)
    missing, unexpected = m.load_state_dict(sd, strict=False)
    assert not missing, missing                    # every parameter of the mirror exists under the reference's key
    assert all('pe' in k for k in unexpected), unexpected
    return m, tv


@pytest.mark.parametrize('dtype', [torch.float32, torch.bfloat16])
def test_v2_golden(dtype):
    f = torch.load(os.path.join(G, 'v2_small.pt'), weights_only=False)
    cfg, sd, mem = f['config'], f['state_dict'], f['memory']
    m, tv = build(cfg, sd, dtype, mem['image'].shape[-1], mem['article'].shape[-1])
    memc = {k: v.cuda() for k, v in mem.items()}
    tol = 1e-4 if dtype == torch.float32 else 2e-2
    out = m.forward(memc, f['ids'].cuda(), None, f['tgt_mask'].cuda())
    assert out.shape == f['out'].shape and rel_err(out, f['out']) < tol
    pad = 1
    total, per = tv.SimpleLossCompute(m.generator, pad)(out, f['y'].cuda(), float(f['ntokens']))
    lo = O.v2_generator_loss(f['out'], f['y'], sd, pad, float(f['ntokens']))
    assert abs(per.item() - lo.item()) < tol * abs(lo.item())
    lp = m.generator(out[:, -1])
    z = torch.log_softmax(torch.nn.functional.linear(f['out'][:, -1], sd['generator.proj.weight'], sd['generator.proj.bias']), -1)
    assert rel_err(lp, z) < tol
    if dtype == torch.float32:
        ys = tv.translate_sentence(m, memc, f['greedy'].shape[1], int(f['greedy'][0, 0]))
        assert torch.equal(ys.cpu(), f['greedy'])


def test_v2_full_width_bf16():
    """custom_v4.py model width (d=1024, 8 heads of 128, ffn 4096; 2 of the 6 layers, 196 image tokens of
    1024, 60 article tokens of 1024) against the oracle on the CPU."""
    g = torch.Generator().manual_seed(9)
    cfg = dict(vocab=1000, d=1024, h=8, layers=2, ffn=4096)
    from news_image_caption_b200 import transformer_v2 as tv0
    ref = tv0.make_model_news(cfg['vocab'], cfg['layers'], cfg['d'], cfg['ffn'], cfg['h'], img_dim=1024, sent_dim=1024,
                              dtype=torch.float32)
    sd = {}
    for k, v in ref.state_dict().items():
        if v.dim() > 1:
            sd[k] = torch.randn(v.shape, generator=g) * v.shape[-1] ** -0.5
        elif k.endswith('a_2'):
            sd[k] = 1 + 0.1 * torch.randn(v.shape, generator=g)
        else:
            sd[k] = 0.1 * torch.randn(v.shape, generator=g)
    B, T = 3, 9
    ids = torch.randint(0, cfg['vocab'], (B, T), generator=g)
    mem = {'image': torch.randn(B, 196, 1024, generator=g), 'article': torch.randn(B, 60, 1024, generator=g),
           'article_mask': torch.rand(B, 1, 60, generator=g) > 0.3}
    want = O.v2_decode(ids, mem, O.v2_subsequent_mask(T), sd, cfg['h'])
    m, tv = build(cfg, sd, torch.bfloat16, 1024, 1024)
    out = m.forward({k: v.cuda() for k, v in mem.items()}, ids.cuda(), None, tv.subsequent_mask(T).cuda())
    assert rel_err(out, want) < 2e-2
