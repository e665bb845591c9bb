"""GPU parity of the backward kernels and the training step (north_star kernel group 4) against autograd
through the oracle (oracle/restate.py is plain differentiable torch; its gradients are pinned against the
reference modules' own autograd in tests/test_oracle_vs_reference.py)."""
import os

import pytest
import torch

from oracle import restate as O
from oracle import synth

pytestmark = pytest.mark.gpu
G = os.path.join(os.path.dirname(__file__), 'golden')


def rel_err(a, b):
    a, b = a.float().cpu(), b.float().cpu()
    return ((a - b).abs().max() / b.abs().max().clamp_min(1e-6)).item()


def l2_err(a, b):
    a, b = a.double().cpu(), b.double().cpu()
    return ((a - b).norm() / b.norm().clamp_min(1e-12)).item()


@pytest.fixture(scope='module')
def ops():
    from news_image_caption_b200 import build, _lib, ops as o
    build.build()
    _lib.require_device()
    return o


@pytest.mark.parametrize('T,S,H,D,mode', [(7, 20, 4, 16, 'pad'), (32, 130, 2, 64, 'pad'), (40, 33, 2, 128, 'causal'),
                                          (5, 9, 2, 64, 'none')])
@pytest.mark.parametrize('dtype', [torch.float32, torch.bfloat16])
def test_xattn_backward(ops, T, S, H, D, mode, dtype):
    g = torch.Generator().manual_seed(T * 100 + S)
    B, E = 3, H * D
    q = torch.randn(T, B, E, generator=g).to(dtype)
    k = torch.randn(S, B, E, generator=g).to(dtype)
    v = torch.randn(S, B, E, generator=g).to(dtype)
    do = torch.randn(T, B, E, generator=g).to(dtype)
    scale = D ** -0.5
    pad = v2 = None
    qd, kd, vd = [t.double().requires_grad_(True) for t in (q, k, v)]
    s = torch.einsum('tbhd,sbhd->bhts', qd.view(T, B, H, D), kd.view(S, B, H, D)) * scale
    if mode == 'pad':
        pad = torch.rand(B, S, generator=g) > 0.6
        pad[:, -2:] = False                       # the two extra keys are never masked
        pad[0, :-2] = True                        # a fully masked source
        s = s.masked_fill(pad[:, None, None, :], float('-inf'))
    elif mode == 'causal':
        v2 = torch.tril(torch.ones(T, S)).bool()[None]
        s = s.masked_fill(~v2[:, None], -1e9)
    o = torch.einsum('bhts,sbhd->tbhd', torch.softmax(s, -1), vd.view(S, B, H, D)).reshape(T, B, E)
    gq, gk, gv = torch.autograd.grad(o, [qd, kd, vd], do.double())
    out, _ = ops.xattn(q.cuda(), k.cuda(), v.cuda(), H, scale, pad_mask=None if pad is None else pad.cuda(),
                       v2_mask=None if v2 is None else v2.cuda(), mask_S=S - 2 if pad is not None else 0)
    dq, dk, dv = ops.xattn_bwd(q.cuda(), k.cuda(), v.cuda(), out, do.cuda(), H, scale,
                               pad_mask=None if pad is None else pad.cuda(),
                               v2_mask=None if v2 is None else v2.cuda(), mask_S=S - 2 if pad is not None else 0)
    tol = 2e-5 if dtype == torch.float32 else 3e-2
    assert rel_err(out, o.detach()) < tol
    assert rel_err(dq, gq) < tol and rel_err(dk, gk) < tol and rel_err(dv, gv) < tol


@pytest.mark.parametrize('T,S,B,H,D', [(7, 20, 3, 4, 16), (32, 130, 3, 2, 64), (40, 514, 2, 4, 64), (64, 51, 2, 2, 64)])
@pytest.mark.parametrize('dtype', [torch.float32, torch.bfloat16])
def test_xattn_probability_dropout(ops, T, S, B, H, D, dtype):
    """attention_dropout (multi_head.py:463): forward and backward of the SIMT (fp32, D != 64) and tcgen05 (bf16, D = 64)
    kernels against fp64 autograd with the same counter-based mask; the soft-max normaliser stays un-dropped."""
    g = torch.Generator().manual_seed(T * 100 + S)
    E = H * D
    q = torch.randn(T, B, E, generator=g).to(dtype)
    k = torch.randn(S, B, E, generator=g).to(dtype)
    v = torch.randn(S, B, E, generator=g).to(dtype)
    do = torch.randn(T, B, E, generator=g).to(dtype)
    scale = D ** -0.5
    pad = torch.rand(B, S, generator=g) > 0.6
    pad[:, -2:] = False
    qd, kd, vd = [t.double().requires_grad_(True) for t in (q, k, v)]
    s = torch.einsum('tbhd,sbhd->bhts', qd.view(T, B, H, D), kd.view(S, B, H, D)) * scale
    s = s.masked_fill(pad[:, None, None, :], float('-inf'))
    drop = (0.2, 4242, 6)
    m = O.dropout_mask(*drop, torch.arange(B * H * T * S).view(B, H, T, S)).double()
    o = torch.einsum('bhts,sbhd->tbhd', torch.softmax(s, -1) * m, vd.view(S, B, H, D)).reshape(T, B, E)
    gq, gk, gv = torch.autograd.grad(o, [qd, kd, vd], do.double())
    lse = torch.empty(B, H, T, dtype=torch.float32, device='cuda')
    out, _ = ops.xattn(q.cuda(), k.cuda(), v.cuda(), H, scale, pad_mask=pad.cuda(), mask_S=S - 2, lse=lse, drop=drop)
    dq, dk, dv = ops.xattn_bwd(q.cuda(), k.cuda(), v.cuda(), out, do.cuda(), H, scale, pad_mask=pad.cuda(), mask_S=S - 2,
                               lse=lse, drop=drop)
    tol = 2e-5 if dtype == torch.float32 else 3e-2
    assert rel_err(lse, torch.logsumexp(s.detach(), -1)) < (1e-4 if dtype == torch.float32 else 2e-2)
    assert rel_err(out, o.detach()) < tol
    assert rel_err(dq, gq) < tol and rel_err(dk, gk) < tol and rel_err(dv, gv) < tol


def test_xattn_skips_all_padding_key_tiles(ops):
    """The tcgen05 forward / backward visit only the 128-key tiles that hold a live key (tnt_xattn_tile_mask): same
    numbers as visiting every tile, and as fp64, for suffix padding, a dead tile in the MIDDLE and a fully masked row."""
    g = torch.Generator().manual_seed(31)
    T, S, B, H, D = 32, 514, 6, 4, 64
    E = H * D
    q, k, v, do = [torch.randn(n, B, E, generator=g).bfloat16() for n in (T, S, S, T)]
    pad = torch.zeros(B, S - 2, dtype=torch.bool)
    pad[0, 130:] = True                    # tiles 2, 3 dead (tile 4 holds the 2 extra keys)
    pad[1, 300:] = True
    pad[2, 128:256] = True                 # a dead tile in the middle
    pad[3, :] = True                       # only the extra keys remain
    pad[4, 5:] = True
    tl = ops.xattn_tile_mask(pad.cuda(), S, S - 2)
    assert [v & 0xFFFF for v in tl.cpu().tolist()] == [0b10011, 0b10111, 0b11101, 0b10000, 0b10001, 0b11111]   # live tiles
    assert [(v >> 16) & 0xFFFF for v in tl.cpu().tolist()] == [132, 302, 386, 2, 7, 514]                        # live keys
    scale = D ** -0.5
    lse0 = torch.empty(B, H, T, device='cuda')
    lse1 = torch.empty(B, H, T, device='cuda')
    full, _ = ops.xattn(q.cuda(), k.cuda(), v.cuda(), H, scale, pad_mask=pad.cuda(), mask_S=S - 2, lse=lse0)
    skip, _ = ops.xattn(q.cuda(), k.cuda(), v.cuda(), H, scale, pad_mask=pad.cuda(), mask_S=S - 2, lse=lse1, tile_live=tl)
    assert torch.equal(full, skip) and torch.equal(lse0, lse1)
    qd, kd, vd = [t.double().requires_grad_(True) for t in (q, k, v)]
    s = torch.einsum('tbhd,sbhd->bhts', qd.view(T, B, H, D), kd.view(S, B, H, D)) * scale
    fm = torch.cat([pad, torch.zeros(B, 2, dtype=torch.bool)], 1)
    s = s.masked_fill(fm[:, None, None, :], float('-inf'))
    o = torch.einsum('bhts,sbhd->tbhd', torch.softmax(s, -1), vd.view(S, B, H, D)).reshape(T, B, E)
    gq, gk, gv = torch.autograd.grad(o, [qd, kd, vd], do.double())
    assert rel_err(skip, o.detach()) < 3e-2
    a = ops.xattn_bwd(q.cuda(), k.cuda(), v.cuda(), full, do.cuda(), H, scale, pad_mask=pad.cuda(), mask_S=S - 2, lse=lse0)
    b = ops.xattn_bwd(q.cuda(), k.cuda(), v.cuda(), full, do.cuda(), H, scale, pad_mask=pad.cuda(), mask_S=S - 2, lse=lse0,
                      tile_live=tl)
    for x, y, z in zip(a, b, (gq, gk, gv)):
        assert rel_err(y, x) < 1e-6 and rel_err(y, z) < 3e-2
    assert float(b[1][:, 3][:S - 2].abs().max()) == 0.0          # masked keys receive exactly zero gradient


@pytest.mark.parametrize('T,S,B,H', [(32, 514, 3, 4), (7, 130, 2, 2), (40, 51, 5, 16), (64, 1000, 2, 2)])
def test_xattn_backward_tcgen05(ops, T, S, B, H):
    """The tcgen05 / TMEM backward kernel (bf16, D = 64, T <= 64, lse saved by the forward) against fp64
    autograd and against the library's own SIMT backward on identical inputs."""
    from news_image_caption_b200 import _lib
    g = torch.Generator().manual_seed(T * 7 + S)
    D, E = 64, H * 64
    q = torch.randn(T, B, E, generator=g).bfloat16()
    k = torch.randn(S, B, E, generator=g).bfloat16()
    v = torch.randn(S, B, E, generator=g).bfloat16()
    do = torch.randn(T, B, E, generator=g).bfloat16()
    scale = D ** -0.5
    pad = torch.rand(B, S, generator=g) > 0.7
    pad[:, -2:] = False
    pad[0, :-2] = True
    qd, kd, vd = [t.double().requires_grad_(True) for t in (q, k, v)]
    s = torch.einsum('tbhd,sbhd->bhts', qd.view(T, B, H, D), kd.view(S, B, H, D)) * scale
    s = s.masked_fill(pad[:, None, None, :], float('-inf'))
    o = torch.einsum('bhts,sbhd->tbhd', torch.softmax(s, -1), vd.view(S, B, H, D)).reshape(T, B, E)
    gq, gk, gv = torch.autograd.grad(o, [qd, kd, vd], do.double())
    lse = torch.empty(B, H, T, dtype=torch.float32, device='cuda')
    args = dict(pad_mask=pad.cuda(), mask_S=S - 2)
    out, _ = ops.xattn(q.cuda(), k.cuda(), v.cuda(), H, scale, lse=lse, **args)
    assert rel_err(lse, torch.logsumexp(s, -1).detach()) < 1e-2
    dq, dk, dv = ops.xattn_bwd(q.cuda(), k.cuda(), v.cuda(), out, do.cuda(), H, scale, lse=lse, **args)
    torch.cuda.synchronize()
    for a_, b_ in ((dq, gq), (dk, gk), (dv, gv)):
        assert rel_err(a_, b_) < 3e-2
    _lib.lib().tnt_debug_force_simt_attention_bwd(1)
    try:
        sq, sk, sv = ops.xattn_bwd(q.cuda(), k.cuda(), v.cuda(), out, do.cuda(), H, scale, **args)
    finally:
        _lib.lib().tnt_debug_force_simt_attention_bwd(0)
    for a_, b_ in ((dq, sq), (dk, sk), (dv, sv)):
        assert rel_err(a_, b_) < 2e-2


def make_decoder(cfg, sd, dtype):
    from news_image_caption_b200 import DynamicConvDecoder, build
    build.build()
    dec = DynamicConvDecoder(decoder_output_dim=cfg['d'], decoder_attention_heads=cfg['heads'],
                             decoder_ffn_embed_dim=cfg['ffn'], decoder_kernel_size_list=cfg['kernels'],
                             vocab_size=cfg['vocab'], adaptive_softmax_cutoff=cfg['cutoffs'][:-1],
                             sources=cfg['sources'], dtype=dtype)
    dec.load_reference_state_dict(sd)
    return dec


def oracle_grads(f, scale, drop=None):
    cfg = f['config']
    sd = O.leaf_state_dict(f['state_dict'], O.tied_output_keys(len(cfg['cutoffs'])))
    X, _ = O.decoder_forward(f['prev'], f['contexts'], sd, cfg['heads'], cfg['cutoffs'], cfg['sources'], drop=drop)
    loss, n = O.adaptive_nll(X, f['target'], sd, cfg['cutoffs'])
    (loss * scale).backward()
    return loss.detach(), n, sd


@pytest.mark.parametrize('name', ['tell2_small.pt', 'tell4_small.pt'])
@pytest.mark.parametrize('dtype', [torch.float32, torch.bfloat16])
def test_train_step_gradients(name, dtype):
    f = torch.load(os.path.join(G, name), weights_only=False)
    scale = 0.37
    loss_o, n_o, sd = oracle_grads(f, scale)
    dec = make_decoder(f['config'], f['state_dict'], dtype).enable_training()
    ctx = {k: v.cuda() for k, v in f['contexts'].items()}
    out = dec.train_step(f['prev'].cuda(), f['target'].cuda(), ctx, loss_scale=scale)
    torch.cuda.synchronize()
    tol = 1e-4 if dtype == torch.float32 else 2e-2
    assert int(out[1].item()) == n_o
    assert abs(out[0].item() - loss_o.item()) < tol * abs(loss_o.item())
    # fp32: max-norm relative error per tensor; bf16 (activation gradients are stored in bf16 through
    # 2 layers x ~15 ops): relative L2 error per tensor -- north_star states tolerances for logits / loss only
    gtol = 2e-4 if dtype == torch.float32 else 1e-1
    err = rel_err if dtype == torch.float32 else l2_err
    worst = []
    for k, p in dec.named_parameters():
        if 'token_embedder_position' in k:
            continue
        go = sd[k].grad
        assert go is not None, k
        worst.append((err(p.grad, go), k))
    worst.sort(reverse=True)
    assert worst[0][0] < gtol, '\n'.join('%.3e %s' % w for w in worst[:12])


@pytest.mark.parametrize('art_dim', [1024, 128])
def test_train_step_long_context_weight_gradients(art_dim):
    """An article of S*B = 8192 rows (bf16): the K/V projection weight gradient splits its reduction into groups of one
    grouped GEMM launch whose fp32 partials are summed (engine_train.cu kv_param_grads); both destinations -- separate
    k/v_proj_weight (kdim != d) and the in_proj_weight rows [d, 3d) (kdim == d) -- against autograd through the oracle."""
    from news_image_caption_b200 import DynamicConvDecoder, build
    from news_image_caption_b200.decoder import random_state_dict
    from oracle import synth
    build.build()
    d, H, V, cut = 128, 2, 300, [100, 200, 300]
    sdim = {'article': art_dim}
    sd0 = random_state_dict(d=d, heads=H, ffn=128, kernels=(3,), vocab=V, cutoffs=cut, source_dim=sdim, seed=5)
    dec = DynamicConvDecoder(decoder_output_dim=d, decoder_attention_heads=H, decoder_ffn_embed_dim=128,
                             decoder_kernel_size_list=[3], vocab_size=V, adaptive_softmax_cutoff=cut[:-1], source_dim=sdim,
                             dtype=torch.bfloat16)
    dec.load_reference_state_dict(sd0)
    dec.enable_training()
    B, T = 16, 4
    prev, tgt, ctx = synth.make_batch(B, T, V, ('image', 'article'), seed=9, S_img=3, S_art=512, d_art=art_dim)
    f = dict(config=dict(heads=H, cutoffs=cut, sources=['image', 'article']), state_dict=sd0, prev=prev, target=tgt, contexts=ctx)
    scale = 0.5
    loss_o, n_o, sd = oracle_grads(f, scale)
    out = dec.train_step(prev.cuda(), tgt.cuda(), {k: v.cuda() for k, v in ctx.items()}, loss_scale=scale)
    torch.cuda.synchronize()
    assert int(out[1].item()) == n_o
    assert abs(out[0].item() - loss_o.item()) < 2e-2 * abs(loss_o.item())
    keys = [k for k, _ in dec.named_parameters() if 'context_attns.article' in k]
    assert any('k_proj_weight' in k for k in keys) == (art_dim != d)
    worst = sorted(((l2_err(p.grad, sd[k].grad), k) for k, p in dec.named_parameters() if k in keys), reverse=True)
    assert worst[0][0] < 1e-1, '\n'.join('%.3e %s' % w for w in worst[:8])


DROP_CFG = dict(dropout=0.1, input_dropout=0.1, weight_dropout=0.1, attention_dropout=0.1, relu_dropout=0.05)


@pytest.mark.parametrize('name', ['tell2_small.pt', 'tell4_small.pt'])
@pytest.mark.parametrize('dtype', [torch.float32, torch.bfloat16])
def test_train_step_gradients_with_dropout(name, dtype):
    """Training mode (config.yaml:51-69 dropout rates; relu_dropout raised from 0 to exercise its site): loss and every
    gradient against autograd through the oracle fed the SAME masks (the counter-based generator restated in numpy),
    for two consecutive steps (step k draws its masks with seed + k)."""
    f = torch.load(os.path.join(G, name), weights_only=False)
    scale, seed = 0.37, 0x1234567890ABCDEF
    dec = make_decoder(f['config'], f['state_dict'], dtype).enable_training()
    dec.set_dropout(seed=seed, **DROP_CFG)
    ctx = {k: v.cuda() for k, v in f['contexts'].items()}
    loss_plain = oracle_grads(f, scale)[0]
    for step in range(2):
        loss_o, n_o, sd = oracle_grads(f, scale, O.DropoutMasks(seed=seed + step, **DROP_CFG))
        assert abs(loss_o.item() - loss_plain.item()) > 1e-3 * abs(loss_plain.item())      # the masks do something
        out = dec.train_step(f['prev'].cuda(), f['target'].cuda(), ctx, loss_scale=scale)
        torch.cuda.synchronize()
        tol = 1e-4 if dtype == torch.float32 else 2e-2
        assert int(out[1].item()) == n_o
        assert abs(out[0].item() - loss_o.item()) < tol * abs(loss_o.item()), (step, out[0].item(), loss_o.item())
        # bf16: relative L2 per tensor as in test_train_step_gradients; the 1/(1-p) factors of 6 stacked dropout sites
        # amplify the bf16 rounding of the stored activation gradients, hence 2e-1 (fp32 pins the math at 2e-4)
        gtol = 2e-4 if dtype == torch.float32 else 2e-1
        err = rel_err if dtype == torch.float32 else l2_err
        worst = []
        for k, p in dec.named_parameters():
            if 'token_embedder_position' in k:
                continue
            worst.append((err(p.grad, sd[k].grad), k))
        worst.sort(reverse=True)
        assert worst[0][0] < gtol, 'step %d\n' % step + '\n'.join('%.3e %s' % w for w in worst[:12])
    # switching dropout off again restores the parity configuration
    dec.set_dropout()
    out = dec.train_step(f['prev'].cuda(), f['target'].cuda(), ctx, loss_scale=scale)
    assert abs(out[0].item() - loss_plain.item()) < (1e-4 if dtype == torch.float32 else 2e-2) * abs(loss_plain.item())


def test_gradients_are_additive_over_batch_shards():
    """Data-parallel invariant: summed gradients of two half-batches == gradients of the whole
    batch (each rank back-propagates the un-normalised sum loss), and the losses / token counts add up."""
    f = torch.load(os.path.join(G, 'tell4_small.pt'), weights_only=False)
    dec = make_decoder(f['config'], f['state_dict'], torch.float32).enable_training()
    prev, tgt = f['prev'].cuda(), f['target'].cuda()
    ctx = {k: v.cuda() for k, v in f['contexts'].items()}
    B = prev.shape[0]
    full = dec.train_step(prev, tgt, ctx).clone()
    g_full = dec.flat_grads.clone()
    acc, outs = torch.zeros_like(g_full), torch.zeros_like(full)
    for sl in (slice(0, B // 2), slice(B // 2, B)):
        c = {k: (v[sl] if k.endswith('_mask') else v[:, sl]).contiguous() for k, v in ctx.items()}
        outs += dec.train_step(prev[sl].contiguous(), tgt[sl].contiguous(), c)
        acc += dec.flat_grads
    assert outs[1].item() == full[1].item()
    assert abs(outs[0].item() - full[0].item()) < 1e-5 * abs(full[0].item())
    assert rel_err(acc, g_full) < 1e-4


def test_trainer_two_steps_match_oracle_bert_adam():
    """Trainer.step (engine backward + fused BertAdam with per-tensor clipping and the 1/(n_tokens ln 2) loss
    normalisation) against the oracle: autograd through the restated decoder + the restated BertAdam."""
    import math
    from news_image_caption_b200.trainer import Trainer
    f = torch.load(os.path.join(G, 'tell2_small.pt'), weights_only=False)
    cfg = f['config']
    dec = make_decoder(cfg, f['state_dict'], torch.float32)
    tr = Trainer(dec, lr=1e-2, max_grad_norm=0.1, weight_decay=1e-3, t_total=10, warmup=0.2)
    ctx = {k: v.cuda() for k, v in f['contexts'].items()}
    ties = O.tied_output_keys(len(cfg['cutoffs']))
    sd = O.leaf_state_dict(f['state_dict'], ties)
    own = dict(dec.named_parameters())
    trainable = {k: v for k, v in sd.items() if k in own and k not in ties and 'token_embedder_position' not in k}
    state = {}
    for step in range(2):
        out = tr.step(f['prev'].cuda(), f['target'].cuda(), ctx)
        X, _ = O.decoder_forward(f['prev'], f['contexts'], sd, cfg['heads'], cfg['cutoffs'], cfg['sources'])
        loss, n = O.adaptive_nll(X, f['target'], sd, cfg['cutoffs'])
        assert abs(out[0].item() - loss.item()) < 1e-4 * abs(loss.item()) and int(out[1].item()) == n
        grads = dict(zip(trainable, torch.autograd.grad(loss / n / math.log(2.0), list(trainable.values()))))
        with torch.no_grad():
            O.bert_adam_step(trainable, grads, state, lr=1e-2 * O.warmup_linear(step / 10, 0.2), weight_decay=1e-3)
    torch.cuda.synchronize()
    worst = max((rel_err(own[k], v.detach()), k) for k, v in trainable.items())
    # parameters whose true gradient is ~0 (key-projection biases: softmax is shift invariant) get Adam-normalised
    # rounding noise, hence the looser bound
    assert worst[0] < 1e-3, worst


@pytest.mark.parametrize('dtype,tol', [(torch.float32, 1e-5), (torch.bfloat16, 2e-3)])
def test_parameter_gradient_streams_do_not_change_the_gradients(dtype, tol):
    """The training backward issues every parameter gradient on side streams beside the data-gradient chain
    (engine_train.cu); with all of it on the caller's stream (0 side streams) the step must give the same loss and gradients."""
    from news_image_caption_b200 import _lib
    f = torch.load(os.path.join(G, 'tell2_small.pt'), weights_only=False)
    cfg = f['config']
    ctx = {k: v.cuda() for k, v in f['contexts'].items()}
    L = _lib.lib()
    res = {}
    try:
        for n in (0, 3, 1):
            L.tnt_debug_train_side_streams(n)
            dec = make_decoder(cfg, f['state_dict'], dtype).enable_training()
            for _ in range(2):                      # the second step reuses the per-parity buffers of the first
                out = dec.train_step(f['prev'].cuda(), f['target'].cuda(), ctx).clone()
            torch.cuda.synchronize()
            res[n] = (out, dec.flat_grads.clone())
    finally:
        L.tnt_debug_train_side_streams(-1)
    for n in (3, 1):
        assert rel_err(res[n][0], res[0][0]) < tol
        assert rel_err(res[n][1], res[0][1]) < tol


def test_stage_wise_bert_adam_equals_the_single_pass(monkeypatch):
    """Trainer with TNT_STAGE_ADAM=1 (per-tensor clipping makes a stage's slice of the arena steppable as soon as its
    gradients are final: norms + BertAdam slice by slice behind the stage events, the band tables last) must end with the
    same parameters and moments as the single pass over the whole arena after the backward."""
    from news_image_caption_b200.trainer import Trainer
    f = torch.load(os.path.join(G, 'tell2_small.pt'), weights_only=False)
    cfg = f['config']
    ctx = {k: v.cuda() for k, v in f['contexts'].items()}
    res = {}
    for mode in ('0', '1'):
        monkeypatch.setenv('TNT_STAGE_ADAM', mode)
        dec = make_decoder(cfg, f['state_dict'], torch.float32)
        tr = Trainer(dec, lr=1e-2, max_grad_norm=0.1, weight_decay=1e-3, t_total=10, warmup=0.2)
        assert tr.stage_adam == (mode == '1')
        for _ in range(3):
            out = tr.step(f['prev'].cuda(), f['target'].cuda(), ctx)
        torch.cuda.synchronize()
        res[mode] = (dec.flat_params.clone(), tr.m.clone(), tr.v.clone(), out.clone())
    for a, b in zip(res['0'], res['1']):
        assert rel_err(a, b) < 1e-6


def test_train_step_full_width_bf16():
    """BASELINE.json model width (d=1024, 16 heads of 64, ffn 4096, V=50265, cutoffs 5000/20000), 2 layers:
    the tcgen05 paths (NT / NN / TN GEMMs, attention forward + backward, softmax-gradient epilogue) against
    fp32 autograd through the oracle on the CPU."""
    from news_image_caption_b200 import DynamicConvDecoder, build
    from news_image_caption_b200.decoder import random_state_dict
    build.build()
    kernels, cut = (3, 31), (5000, 20000, 50265)
    sd0 = random_state_dict(kernels=kernels, seed=3)
    prev, tgt, ctx = synth.make_batch(3, 12, 50265, ('image', 'article'), seed=21, S_img=49, S_art=140)
    sd = O.leaf_state_dict(sd0, O.tied_output_keys(3))
    X, _ = O.decoder_forward(prev, ctx, sd, 16, list(cut), ('image', 'article'))
    loss, n = O.adaptive_nll(X, tgt, sd, list(cut))
    loss.backward()
    dec = DynamicConvDecoder(decoder_kernel_size_list=kernels, dtype=torch.bfloat16)
    dec.load_reference_state_dict(sd0)
    dec.enable_training()
    out = dec.train_step(prev.cuda(), tgt.cuda(), {k: v.cuda() for k, v in ctx.items()})
    torch.cuda.synchronize()
    assert int(out[1].item()) == n and abs(out[0].item() - loss.item()) < 2e-2 * abs(loss.item())
    worst = sorted(((l2_err(p.grad, sd[k].grad), k) for k, p in dec.named_parameters()
                    if 'token_embedder_position' not in k), reverse=True)
    assert worst[0][0] < 1e-1, '\n'.join('%.3e %s' % w for w in worst[:10])


@pytest.mark.parametrize('dtype', [torch.float32, torch.bfloat16])
def test_sparse_embedding_gradient_equals_dense_scatter(dtype):
    """Data-parallel form of the tied band-table gradient (tnt_decoder_set_sparse_embed_grad + tnt_embed_grad_scatter,
    trainer.SparseRowExchange): the step packs the <= B*T embedding rows into a list instead of scattering them; the dense
    tables then hold the softmax part only.  Scattering the list afterwards must reproduce the dense-mode gradients
    (every other gradient is untouched), incl. the skipped band-local padding row."""
    import ctypes as C
    from news_image_caption_b200 import _lib
    from news_image_caption_b200.trainer import SparseRowExchange
    f = torch.load(os.path.join(G, 'tell2_small.pt'), weights_only=False)
    cfg = f['config']
    dec = make_decoder(cfg, f['state_dict'], dtype).enable_training()
    prev, tgt = f['prev'].clone(), f['target'].cuda()
    prev[0, 1] = 0                          # band-local row 0 of band 0 (the padding row) appears as an input id
    prev[1, 2] = cfg['cutoffs'][0]          # ... and of band 1
    ctx = {k: v.cuda() for k, v in f['contexts'].items()}
    dec.train_step(prev.cuda(), tgt, ctx)
    dense = dec.flat_grads.clone()
    n = prev.numel()
    sx = SparseRowExchange(n, cfg['d'], dec.torch_dtype, 'cuda')
    rows, meta, count = sx.views(sx.mine)
    L = _lib.lib()
    _lib.check(L.tnt_decoder_set_sparse_embed_grad(dec._h(), C.c_void_p(rows.data_ptr()), C.c_void_p(meta.data_ptr()),
                                                   C.c_void_p(count.data_ptr()), n), 'set_sparse')
    dec.train_step(prev.cuda(), tgt, ctx)
    assert int(count[0].item()) == n                       # every token lives in exactly one band
    sparse_only = dec.flat_grads.clone()
    tabs = dec.band_table_grads()
    assert not torch.equal(sparse_only, dense)             # the rows really were withheld
    sx.gather()
    sx.scatter_into(tabs, dec.embed_padding_idx)
    torch.cuda.synchronize()
    assert rel_err(dec.flat_grads, dense) < (1e-5 if dtype == torch.float32 else 1e-2)
    # back to the dense scatter
    _lib.check(L.tnt_decoder_set_sparse_embed_grad(dec._h(), None, None, None, 0), 'set_sparse off')
    dec.train_step(prev.cuda(), tgt, ctx)
    assert rel_err(dec.flat_grads, dense) < 1e-5


@pytest.mark.parametrize('name', ['tell2_small.pt', 'tell4_small.pt'])
def test_autograd_connected_forward_and_loss(name):
    """the reference trainer's three calls -- decoder(...), criterion(...), loss.backward()
    (transformer_faces_objects.py:79-88, callback_apex_trainer.py:229-233) -- work on the mirror: gradients in `.grad`
    equal autograd through the oracle for the reference's normalised loss (sum-NLL / ln 2 / n_tokens), they ACCUMULATE over
    two backward calls, survive optimizer.zero_grad(set_to_none=True), and a stock torch optimizer stepping
    decoder.parameters() moves the loss exactly as it moves the oracle's."""
    import math
    from news_image_caption_b200 import AdaptiveLoss
    f = torch.load(os.path.join(G, name), weights_only=False)
    cfg = f['config']
    dec = make_decoder(cfg, f['state_dict'], torch.float32).enable_training()
    dec.train()
    crit = AdaptiveLoss(1)
    ctx = {k: v.cuda() for k, v in f['contexts'].items()}
    prev, tgt = f['prev'].cuda(), f['target'].cuda()
    ties = O.tied_output_keys(len(cfg['cutoffs']))
    sd = O.leaf_state_dict(f['state_dict'], ties)
    own = dict(dec.named_parameters())
    trainable = {k: v for k, v in sd.items() if k in own and k not in ties and 'token_embedder_position' not in k}

    def oracle_loss():
        X, _ = O.decoder_forward(f['prev'], f['contexts'], sd, cfg['heads'], cfg['cutoffs'], cfg['sources'])
        loss, n = O.adaptive_nll(X, f['target'], sd, cfg['cutoffs'])
        return loss / math.log(2.0) / n, n

    def mirror_loss():
        X, _ = dec({'roberta': prev}, ctx)
        assert X.requires_grad
        loss, n = crit(dec, (X, None), tgt)
        return loss / math.log(2) / n, n

    lo, n_o = oracle_loss()
    go = dict(zip(trainable, torch.autograd.grad(lo, list(trainable.values()))))
    opt = torch.optim.SGD([p for k, p in dec.named_parameters() if k in trainable], lr=0.05)
    opt.zero_grad(set_to_none=True)
    lm, n = mirror_loss()
    assert n == n_o and abs(lm.item() - lo.item()) < 1e-4 * abs(lo.item())
    lm.backward()
    worst = max((rel_err(own[k].grad, g), k) for k, g in go.items())
    assert worst[0] < 2e-4, worst
    lm2, _ = mirror_loss()                       # a second backward accumulates: .grad == 3 x the single-step gradient
    (2 * lm2).backward()
    worst = max((rel_err(own[k].grad, 3 * g), k) for k, g in go.items())
    assert worst[0] < 2e-4, worst
    # one SGD step through a stock torch optimizer on both sides, then the losses still agree
    opt.zero_grad(set_to_none=True)
    lm, _ = mirror_loss()
    lm.backward()
    opt.step()
    with torch.no_grad():
        for k, v in trainable.items():
            v -= 0.05 * go[k]
    lo2, _ = oracle_loss()
    with torch.no_grad():
        dec.eval()
        X, _ = dec({'roberta': prev}, ctx)
        l2, n2 = crit(dec, (X, None), tgt)
        dec.train()
    assert abs(l2.item() / math.log(2) / n2 - lo2.item()) < 1e-4 * abs(lo2.item())
    assert lo2.item() < lo.item()                # and the step went downhill
    # a criterion call on anything but the tensor forward() returned is refused loudly
    X, _ = dec({'roberta': prev}, ctx)
    with pytest.raises(RuntimeError):
        crit(dec, (X * 1.0, None), tgt)


# ----------------------------------------------------------------------------- training through the copy heads (row f4)
def _pointer_case(kind):
    from news_image_caption_b200 import DynamicConvDecoderEntityPointer, DynamicConvDecoderPointer, build
    build.build()
    ent = kind == 'entity_pointer'
    f = torch.load(os.path.join(G, 'entity_pointer_small.pt' if ent else 'pointer_small.pt'), weights_only=False)
    cfg = f['config']
    kw = dict(decoder_output_dim=cfg['d'], decoder_attention_heads=cfg['heads'], decoder_ffn_embed_dim=cfg['ffn'],
              decoder_kernel_size_list=cfg['kernels'], vocab_size=cfg['vocab'], adaptive_softmax_cutoff=cfg['cutoffs'][:-1],
              sources=cfg['sources'])
    make = (lambda dt: DynamicConvDecoderEntityPointer(hidden=cfg['hidden'], dtype=dt, **kw)) if ent else \
           (lambda dt: DynamicConvDecoderPointer(dtype=dt, **kw))
    copy_ids = f['entity_tokens'] if ent else f['article_ids']
    head = O.entity_pointer_forward if ent else O.pointer_decoder_forward
    return f, cfg, make, copy_ids, head


@pytest.mark.parametrize('kind', ['pointer', 'entity_pointer'])
@pytest.mark.parametrize('dtype', [torch.float32, torch.bfloat16])
def test_pointer_train_step_gradients(kind, dtype):
    """ttl/run.sh trains expt/goodnews/h2_entity_pointer: loss = F.nll_loss(lprobs of the copy mix, target, ignore_index=pad,
    sum) (transformer_entity_pointer.py:123-131; decoder_pointer.py for the plain pointer).  train_step's loss and EVERY
    gradient -- decoder stack, embedding bands, full (not adaptive) softmax gradient, the returned attention weights'
    path into the last layer's attention, p_gen head with its weight-norm -- against autograd through the oracle (pinned on
    the unmodified reference classes: test_pointer_decoder_matches_reference / test_entity_pointer_decoder_matches_reference)."""
    import torch.nn.functional as F
    f, cfg, make, copy_ids, head = _pointer_case(kind)
    tgt = f['target'].clone()
    tgt[0, 2] = copy_ids[0, 1]                   # targets that ARE copy candidates: the attention path carries gradient
    tgt[1, 4] = copy_ids[1, 0]
    tgt[2, 1] = cfg['cutoffs'][0] + 1            # tail-local id 1: counts in the full log-softmax (no in-cluster ignore_index)
    scale = 0.5
    sd = O.leaf_state_dict(f['state_dict'], O.tied_output_keys(len(cfg['cutoffs'])))
    lp_o, _, _ = head(f['prev'], f['contexts'], copy_ids, sd, cfg['heads'], cfg['cutoffs'])
    loss_o = F.nll_loss(lp_o.view(-1, lp_o.shape[-1]), tgt.view(-1), ignore_index=1, reduction='sum')
    (loss_o * scale).backward()
    dec = make(dtype)
    dec.load_reference_state_dict(f['state_dict'])
    dec.enable_training()
    ctx = {k: v.cuda() for k, v in f['contexts'].items()}
    for _ in range(2):                           # twice: gradients are WRITTEN, not accumulated
        out = dec.train_step(f['prev'].cuda(), tgt.cuda(), ctx, copy_ids.cuda(), loss_scale=scale)
    torch.cuda.synchronize()
    tol = 1e-4 if dtype == torch.float32 else 2e-2
    assert int(out[1].item()) == int((tgt != 1).sum())
    assert abs(out[0].item() - loss_o.item()) < tol * abs(loss_o.item())
    gtol = 5e-4 if dtype == torch.float32 else 1.5e-1
    err = rel_err if dtype == torch.float32 else l2_err
    worst = sorted(((err(p.grad, sd[k].grad), k) for k, p in dec.named_parameters()
                    if 'token_embedder_position' not in k and sd[k].grad is not None), reverse=True)
    assert any('p_gens' in k for _, k in worst)
    assert worst[0][0] < gtol, '\\n'.join('%.3e %s' % w for w in worst[:12])
