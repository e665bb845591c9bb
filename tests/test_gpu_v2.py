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


def l2_err(a, b):
    a, b = a.float().cpu(), b.float().cpu()
    return ((a - b).norm() / b.norm().clamp_min(1e-12)).item()


@pytest.mark.parametrize('dtype', [torch.float32, torch.bfloat16])
def test_v2_training_gradients_and_adam_noam(dtype):
    """Rows a12-a14, training: `run_epoch` of the fork's script (custom_v4.py:287-304) -- out = model.forward(...);
    loss, loss_node = loss_compute(out, y, ntokens); loss_node.backward(); optimizer.step(); scheduler.step() -- on the
    mirror after enable_training().  Every gradient against autograd through the oracle (golden fixture model: pinned on
    the unmodified reference module), then Adam(betas=(0.9, 0.98), eps=1e-9) under the Noam schedule: the fused FlatAdam
    kernel and a stock torch.optim.Adam + LambdaLR(rate) on the oracle's leaves must move the loss identically (fp32)."""
    f = torch.load(os.path.join(G, 'v2_small.pt'), weights_only=False)
    cfg, sd0, mem = f['config'], f['state_dict'], f['memory']
    m, tv = build(cfg, sd0, dtype, mem['image'].shape[-1], mem['article'].shape[-1])
    m.enable_training()
    memc = {k: v.cuda() for k, v in mem.items()}
    ids, y, norm, pad = f['ids'], f['y'], float(f['ntokens']), 1
    leaves = {k: v.clone().requires_grad_(True) for k, v in sd0.items() if v.is_floating_point() and 'pe' not in k}

    def oracle_loss():
        out = O.v2_decode(ids, mem, f['tgt_mask'], leaves, cfg['h'])
        return O.v2_generator_loss(out, y, leaves, pad, norm)

    loss_compute = tv.SimpleLossCompute(m.generator, pad)
    own = dict(m.named_parameters())
    opt = tv.FlatAdam(m.parameters(), lr=1.0, betas=(0.9, 0.98), eps=1e-9)
    sched = tv.NoamScheduler(opt, cfg['d'], factor=1.0, warmup=4)
    ref_opt = torch.optim.Adam(list(leaves.values()), lr=1.0, betas=(0.9, 0.98), eps=1e-9)
    ref_sched = torch.optim.lr_scheduler.LambdaLR(ref_opt, lambda s: tv.rate(s, cfg['d'], 1.0, 4))
    tol = 1e-4 if dtype == torch.float32 else 2e-2
    losses = []
    for step in range(3):
        out = m.forward(memc, ids.cuda(), None, f['tgt_mask'].cuda())
        loss, loss_node = loss_compute(out, y.cuda(), norm)
        opt.zero_grad(set_to_none=True)
        loss_node.backward()
        ref_opt.zero_grad(set_to_none=True)
        lo = oracle_loss()
        lo.backward()
        if dtype == torch.float32 or step == 0:
            assert abs(loss_node.item() - lo.item()) < tol * abs(lo.item()), (step, loss_node.item(), lo.item())
        if step == 0:                                  # every gradient, first step (identical parameters on both sides)
            err = rel_err if dtype == torch.float32 else l2_err
            gmax = max(v.grad.abs().max().item() for v in leaves.values())
            # the key-projection biases have a TRUE gradient of zero (softmax is shift invariant): only rounding noise there
            dead = [k for k in leaves if k.endswith('linear_kv.0.bias')]
            assert len(dead) == 3 * cfg['layers']
            for k in dead:
                assert leaves[k].grad.abs().max().item() < 1e-4 * gmax
                assert own[k].grad.abs().max().item() < (1e-4 if dtype == torch.float32 else 2e-2) * gmax
            worst = sorted(((err(own[k].grad, v.grad), k) for k, v in leaves.items() if k not in dead), reverse=True)
            # bf16: relative L2 per tensor through 2 layers of bf16 activation gradients (d = 32 toy model)
            assert worst[0][0] < (5e-4 if dtype == torch.float32 else 1.5e-1), '\n'.join('%.3e %s' % w for w in worst[:10])
        assert abs(opt.param_groups[0]['lr'] - ref_opt.param_groups[0]['lr']) < 1e-12
        opt.step()
        sched.step()
        ref_opt.step()
        ref_sched.step()
        losses.append(loss.item())
    if dtype == torch.float32:
        # parameters after 3 Adam steps; the key biases (true gradient 0) take Adam's +-lr steps on rounding noise on both
        # sides -- the loss does not depend on them -- so they are left out
        worst = max((rel_err(own[k], v.detach()), k) for k, v in leaves.items() if not k.endswith('linear_kv.0.bias'))
        assert worst[0] < 2e-3, worst
    # the fast no-grad path sees the updated parameters (cast / table caches key on the version counters)
    with torch.no_grad():
        out2 = m.forward(memc, ids.cuda(), None, f['tgt_mask'].cuda())
        _, per = loss_compute(out2, y.cuda(), norm)
    with torch.no_grad():
        lo2 = oracle_loss()
    if dtype == torch.float32:
        assert abs(per.item() - lo2.item()) < 1e-3 * abs(lo2.item())
