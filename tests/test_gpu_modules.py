"""GPU parity (forward AND autograd backward) of the module-level drop-ins in news_image_caption_b200/modules.py
against the oracle's restatement of the same reference modules (oracle/restate.py)."""
import pytest
import torch
import torch.nn.functional as F

from oracle import restate as O

pytestmark = pytest.mark.gpu


def rel_err(a, b):
    a, b = a.float().cpu(), b.float().cpu()
    return ((a - b).abs().max() / b.abs().max().clamp_min(1e-6)).item()


@pytest.fixture(scope='module')
def M():
    from news_image_caption_b200 import build, _lib, modules
    build.build()
    _lib.require_device()
    return modules


def test_dynamic_conv_module(M):
    g = torch.Generator().manual_seed(1)
    T, B, C, H, K = 11, 3, 64, 4, 7
    m = M.DynamicConv1dTBC(C, K, padding_l=K - 1, num_heads=H, weight_softmax=True).cuda().eval()
    assert list(m.state_dict().keys()) == ['weight_linear.weight']
    x = torch.randn(T, B, C, generator=g)
    w = m.weight_linear.weight.detach().cpu().clone().requires_grad_(True)
    xr = x.clone().requires_grad_(True)
    yo = O.dynamic_conv(xr, w, H, K)
    gy = torch.randn(T, B, C, generator=g)
    yo.backward(gy)
    xc = x.cuda().requires_grad_(True)
    y = m(xc)
    y.backward(gy.cuda())
    assert rel_err(y, yo.detach()) < 1e-5
    assert rel_err(xc.grad, xr.grad) < 1e-4 and rel_err(m.weight_linear.weight.grad, w.grad) < 1e-4
    # incremental buffer (dynamic.py:95-99): step-by-step == full sequence; reorder = index_select(1, order)
    state = {}
    with torch.no_grad():
        steps = torch.cat([m(x[t:t + 1].cuda(), incremental_state=state) for t in range(T)], 0)
    assert rel_err(steps, yo.detach()) < 1e-5
    key = [k for k in state if k.endswith('input_buffer')][0]
    assert key.startswith('DynamicConv1dTBC.') and state[key].shape == (K - 1, B, C)
    before = state[key].clone()
    m.reorder_incremental_state(state, torch.tensor([2, 0, 1]).cuda())
    assert torch.equal(state[key], before[:, [2, 0, 1]])


@pytest.mark.parametrize('kdim', [64, 40])
def test_multi_head_attention_module(M, kdim):
    g = torch.Generator().manual_seed(2)
    T, B, E, H, S = 6, 3, 64, 4, 13
    m = M.MultiHeadAttention(E, H, kdim=kdim, vdim=kdim).cuda().eval()
    sd = {'a.' + k: v.detach().cpu().clone().requires_grad_(True) for k, v in m.state_dict().items()}
    expect = {'in_proj_bias', 'bias_k', 'bias_v', 'out_proj.weight', 'out_proj.bias'} | \
        ({'in_proj_weight'} if kdim == E else {'q_proj_weight', 'k_proj_weight', 'v_proj_weight'})
    assert set(m.state_dict().keys()) == expect
    q = torch.randn(T, B, E, generator=g)
    mem = torch.randn(S, B, kdim, generator=g)
    pad = torch.rand(B, S, generator=g) > 0.6
    pad[1] = True                                             # fully masked memory: the 2 extra keys remain
    qo = q.clone().requires_grad_(True)
    out_o, w_o = O.mha(qo, mem, pad, sd, 'a.', H)
    gy = torch.randn(T, B, E, generator=g)
    out_o.backward(gy)
    qc = q.cuda().requires_grad_(True)
    out, w = m(qc, mem.cuda(), mem.cuda(), key_padding_mask=pad.cuda(), static_kv=True)
    out.backward(gy.cuda())
    assert rel_err(out, out_o.detach()) < 1e-4 and rel_err(w, w_o.detach()) < 1e-4
    assert rel_err(qc.grad, qo.grad) < 1e-4
    for k, p in m.named_parameters():
        assert rel_err(p.grad, sd['a.' + k].grad) < 2e-4, k


def _next_seed(s):
    """The seed the module's next dropout call will draw after torch.manual_seed(s)."""
    torch.manual_seed(s)
    seed = int(torch.empty((), dtype=torch.int64).random_().item())
    torch.manual_seed(s)
    return seed


def test_modules_in_training_mode_with_dropout(M):
    """train(): DynamicConv1dTBC weight_dropout (dynamic.py:305) and MultiHeadAttention dropout (multi_head.py:463),
    forward, returned (dropped) weights and autograd backward against the oracle fed the same counter-based masks."""
    g = torch.Generator().manual_seed(4)
    T, B, C, H, K = 11, 3, 64, 4, 7
    m = M.DynamicConv1dTBC(C, K, padding_l=K - 1, num_heads=H, weight_softmax=True, weight_dropout=0.25).cuda().train()
    x = torch.randn(T, B, C, generator=g)
    gy = torch.randn(T, B, C, generator=g)
    w = m.weight_linear.weight.detach().cpu().clone().requires_grad_(True)
    xr = x.clone().requires_grad_(True)
    seed = _next_seed(77)
    yo = O.dynamic_conv(xr, w, H, K, lambda t: t * O.dropout_mask(0.25, seed, m._id, torch.arange(t.numel()).view(t.shape)))
    yo.backward(gy)
    xc = x.cuda().requires_grad_(True)
    y = m(xc)
    y.backward(gy.cuda())
    assert rel_err(y, yo.detach()) < 1e-5
    assert rel_err(xc.grad, xr.grad) < 1e-4 and rel_err(m.weight_linear.weight.grad, w.grad) < 1e-4
    assert rel_err(m.eval()(x.cuda()), O.dynamic_conv(x, w.detach(), H, K)) < 1e-5        # eval(): identity

    E, S = 64, 13
    a = M.MultiHeadAttention(E, H, kdim=40, vdim=40, dropout=0.3).cuda().train()
    sd = {'a.' + k: v.detach().cpu().clone().requires_grad_(True) for k, v in a.state_dict().items()}
    q = torch.randn(T, B, E, generator=g)
    mem = torch.randn(S, B, 40, generator=g)
    pad = torch.rand(B, S, generator=g) > 0.6
    qo = q.clone().requires_grad_(True)
    seed = _next_seed(78)
    kk, vv = O.mha_project_kv(mem, sd, 'a.', E)
    out_o, w_o = O.mha_attend(qo, kk, vv, pad, sd, 'a.', H,
                              lambda pr: pr * O.dropout_mask(0.3, seed, 0, torch.arange(pr.numel()).view(pr.shape)))
    gy = torch.randn(T, B, E, generator=g)
    out_o.backward(gy)
    qc = q.cuda().requires_grad_(True)
    out, wts = a(qc, mem.cuda(), mem.cuda(), key_padding_mask=pad.cuda(), static_kv=True)
    out.backward(gy.cuda())
    assert rel_err(out, out_o.detach()) < 1e-4 and rel_err(wts, w_o.detach()) < 1e-4
    assert rel_err(qc.grad, qo.grad) < 1e-4
    for k, p in a.named_parameters():
        assert rel_err(p.grad, sd['a.' + k].grad) < 2e-4, k


def test_gehring_linear_and_adaptive_softmax_modules(M):
    g = torch.Generator().manual_seed(3)
    lin = M.GehringLinear(48, 72).cuda()
    assert set(lin.state_dict().keys()) == {'weight_g', 'weight_v', 'bias'}
    sd = {'l.' + k: v.detach().cpu().clone().requires_grad_(True) for k, v in lin.state_dict().items()}
    x = torch.randn(5, 7, 48, generator=g)
    yo = O.gehring_linear(x, sd, 'l.')
    yo.sum().backward()
    y = lin(x.cuda())
    y.sum().backward()
    assert rel_err(y, yo.detach()) < 1e-5
    assert rel_err(lin.weight_v.grad, sd['l.weight_v'].grad) < 1e-4 and rel_err(lin.weight_g.grad, sd['l.weight_g'].grad) < 1e-4
    V, d, cut = 300, 64, [100, 200]
    ada = M.AdaptiveSoftmax(V, d, cut, dropout=0, factor=1).cuda().eval()
    assert ada.cutoff == [100, 200, 300]
    sdo = {'adaptive_softmax.' + k: v.detach().cpu() for k, v in ada.state_dict().items()}
    X = torch.randn(4, 6, d, generator=g)
    tgt = torch.randint(0, V, (4, 6), generator=g)
    tgt[0, :2] = 1
    logits, targets = ada(X.cuda(), tgt.cuda())
    loss = sum(F.cross_entropy(z.float(), t, ignore_index=1, reduction='sum') for z, t in zip(logits, targets) if z is not None)
    lo, _ = O.adaptive_nll(X, tgt, sdo, [100, 200, 300])
    assert abs(loss.item() - lo.item()) < 1e-4 * abs(lo.item())
    assert rel_err(ada.get_log_prob(X.cuda()), O.adaptive_log_prob(X, sdo, [100, 200, 300])) < 1e-4


@pytest.mark.parametrize('separate,bias_row,zero_row', [(False, True, True), (True, True, False), (False, False, False)])
@pytest.mark.parametrize('dtype,tol', [(torch.float32, 1e-4), (torch.bfloat16, 2e-2)])
def test_attention_score_forward(M, separate, bias_row, zero_row, dtype, tol):
    """multi_head_attention_score_forward (multi_head.py:14-204): head-averaged copy-attention weights only."""
    g = torch.Generator().manual_seed(31)
    T, B, E, H, S = 7, 3, 128, 4, 45
    q = torch.randn(T, B, E, generator=g)
    mem = torch.randn(S, B, E, generator=g)
    w = torch.randn(2 * E, E, generator=g) * 0.1
    b = torch.randn(2 * E, generator=g) * 0.1
    bias_k = torch.randn(1, 1, E, generator=g) if bias_row else None
    pad = torch.rand(B, S, generator=g) > 0.6
    pad[:, 0] = False
    if bias_row:
        pad[1] = True                                             # fully masked memory: only the extra keys remain
    ref = O.mha_score(q, mem, w[:E], w[E:], b[:E], b[E:], bias_k, zero_row, pad, H)
    wc = w.cuda()
    kw = dict(use_separate_proj_weight=True, q_proj_weight=wc[:E], k_proj_weight=wc[E:]) if separate else {}
    got = M.multi_head_attention_score_forward(q.cuda().to(dtype), mem.cuda().to(dtype), E, H, None if separate else wc, b.cuda(),
                                               None if bias_k is None else bias_k.cuda(), zero_row, 0.0, None, None,
                                               training=False, key_padding_mask=pad.cuda(), **kw)
    assert got.shape == ref.shape
    assert rel_err(got, ref) < tol
    assert torch.allclose(got.float().sum(-1).cpu(), torch.ones(B, T), atol=1e-3)
    with pytest.raises(NotImplementedError):
        M.multi_head_attention_score_forward(q.cuda(), mem.cuda(), E, H, wc, b.cuda(), None, False, 0.0,
                                             attn_mask=torch.zeros(T, S).cuda())


@pytest.mark.parametrize('dtype,tol', [(torch.float32, 1e-5), (torch.bfloat16, 2e-2)])
def test_lightweight_conv_module(M, dtype, tol):
    """LightweightConv1dTBC drop-in (lightweight.py:86-253; `decoder_conv_type: lightweight`): forward, gradients w.r.t. the
    input and the static [H, 1, K] weight, bias, K > T, and the incremental buffer, against the oracle (pinned on the
    unmodified module: test_lightweight_conv_matches_reference)."""
    g = torch.Generator().manual_seed(3)
    for T, K in ((11, 7), (3, 15)):
        B, C, H = 3, 64, 4
        m = M.LightweightConv1dTBC(C, K, padding_l=K - 1, num_heads=H, weight_softmax=True, bias=True).cuda().eval()
        assert sorted(m.state_dict().keys()) == ['bias', 'weight'] and tuple(m.weight.shape) == (H, 1, K)
        with torch.no_grad():
            m.bias.normal_(generator=None)
        x = torch.randn(T, B, C, generator=g)
        w = m.weight.detach().cpu().clone().requires_grad_(True)
        xr = x.clone().requires_grad_(True)
        yo = O.lightweight_conv(xr, w, H, K, m.bias.detach().cpu())
        gy = torch.randn(T, B, C, generator=g)
        yo.backward(gy)
        xc = x.cuda().to(dtype).requires_grad_(True)
        y = m(xc)
        y.backward(gy.cuda().to(dtype))
        assert rel_err(y, yo.detach()) < tol
        assert rel_err(xc.grad, xr.grad) < max(tol, 1e-4) and rel_err(m.weight.grad, w.grad) < max(5 * tol, 1e-4)
        if dtype == torch.float32:
            state = {}
            with torch.no_grad():
                steps = torch.cat([m(x[t:t + 1].cuda(), incremental_state=state) for t in range(T)], 0)
            assert rel_err(steps, yo.detach()) < 1e-5
            key = [k for k in state if k.endswith('input_buffer')][0]
            assert key.startswith('LightweightConv1dTBC.') and state[key].shape[0] == min(K - 1, T)
