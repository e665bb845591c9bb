"""CPU, build container only: the oracle against the UNMODIFIED reference modules imported from
/root/reference at fresh seeds / shapes the fixtures do not cover (skipped on the GPU box)."""
import warnings

import pytest
import torch

from oracle import ref_import as ri
from oracle import restate as O
from oracle import synth

pytestmark = pytest.mark.reference
warnings.filterwarnings('ignore')


def close(a, b, tol=5e-5):
    err = (a - b).abs().max().item()
    assert err <= tol * max(1.0, b.abs().max().item()), err


@pytest.mark.parametrize('nsrc,srcs', [(2, ('image', 'article')), (4, ('image', 'article', 'faces', 'obj'))])
def test_decoder_loss_logprob(nsrc, srcs):
    torch.set_grad_enabled(False)
    R = ri.load_tell()
    dec = ri.build_ref_decoder(nsrc, d=64, heads=8, ffn=96, kernels=(3, 31), vocab=500, cutoff=(120, 300),
                               seed=3)
    sd = dec.state_dict()
    prev, tgt, ctx = synth.make_batch(5, 12, 500, srcs, seed=99, S_img=4, S_art=20, S_obj=5)
    X, ex = dec({'roberta': prev}, ctx)
    Xo, attn = O.decoder_forward(prev, ctx, sd, 8, [120, 300, 500], srcs)
    close(Xo, X)
    close(attn, ex['attn'])
    loss, n = R.AdaptiveLoss(1)(dec.adaptive_softmax, (X, None), tgt)
    lo, no = O.adaptive_nll(Xo, tgt, sd, [120, 300, 500])
    assert n == no and abs(lo.item() - loss.item()) < 1e-4 * abs(loss.item())
    close(O.adaptive_log_prob(Xo, sd, [120, 300, 500]), dec.get_normalized_probs((X, None), log_probs=True))
    torch.set_grad_enabled(True)


def test_gradients_match_reference():
    """The oracle is plain differentiable torch code; its autograd gradients are the checker for
    kernel group 4 (backward).  Pin them against the reference module's own autograd."""
    R = ri.load_tell()
    dec = ri.build_ref_decoder(2, d=32, heads=4, ffn=48, kernels=(7,), vocab=300, cutoff=(100, 200), seed=5)
    prev, tgt, ctx = synth.make_batch(3, 8, 300, ('image', 'article'), seed=4, S_img=3, S_art=10)
    X, _ = dec({'roberta': prev}, ctx)
    loss, _ = R.AdaptiveLoss(1)(dec.adaptive_softmax, (X, None), tgt)
    names = ['layers.0.conv.weight_linear.weight', 'layers.0.linear1.weight_v', 'layers.0.fc2.weight_g',
             'layers.0.context_attns.article.bias_k', 'adaptive_softmax.tail.1.0.weight']
    params = dict(dec.named_parameters())
    g_ref = torch.autograd.grad(loss, [params[n] for n in names])
    sd = {k: v.detach().clone().requires_grad_(v.is_floating_point()) for k, v in dec.state_dict().items()}
    Xo, _ = O.decoder_forward(prev, ctx, sd, 4, [100, 200, 300], ('image', 'article'))
    lo, _ = O.adaptive_nll(Xo, tgt, sd, [100, 200, 300])
    g_or = torch.autograd.grad(lo, [sd[n] for n in names])
    for a, b in zip(g_or, g_ref):
        close(a, b, 2e-4)


@pytest.mark.parametrize('nsrc,srcs', [(2, ('image', 'article')), (4, ('image', 'article', 'faces', 'obj'))])
def test_all_gradients_match_reference(nsrc, srcs):
    """Every trainable parameter, with the tied embedding / output rows sharing one leaf and the
    band-local padding rows (ids 0, cutoff[0], cutoff[1]) present in the batch."""
    R = ri.load_tell()
    dec = ri.build_ref_decoder(nsrc, d=32, heads=4, ffn=48, kernels=(3, 7), vocab=300, cutoff=(100, 200), seed=7)
    prev, tgt, ctx = synth.make_batch(4, 9, 300, srcs, seed=11, S_img=3, S_art=10, S_obj=4)
    prev[1, 3], prev[2, 4] = 100, 200                      # local row 0 of bands 1 and 2
    X, _ = dec({'roberta': prev}, ctx)
    loss, _ = R.AdaptiveLoss(1)(dec.adaptive_softmax, (X, None), tgt)
    params = dict(dec.named_parameters())
    g_ref = dict(zip(params, torch.autograd.grad(loss, list(params.values()), allow_unused=True)))
    sd = O.leaf_state_dict(dec.state_dict())
    Xo, _ = O.decoder_forward(prev, ctx, sd, 4, [100, 200, 300], srcs)
    lo, _ = O.adaptive_nll(Xo, tgt, sd, [100, 200, 300])
    for name, gr in g_ref.items():
        if gr is None:
            continue
        go, = torch.autograd.grad(lo, sd[name], retain_graph=True, allow_unused=True)
        assert go is not None, name
        close(go, gr, 2e-4)


def test_v2_other_variant():
    torch.set_grad_enabled(False)
    tv = ri.load_final_v2('final_roberta2')
    torch.manual_seed(2)
    m = tv.make_model_news(40, 1, 32, 64, attn_heads=4, img_dim=16, sent_dim=24).eval()
    g = torch.Generator().manual_seed(1)
    ids = torch.randint(0, 40, (2, 5), generator=g)
    mem = {'image': torch.randn(2, 3, 16, generator=g), 'article': torch.randn(2, 6, 24, generator=g),
           'article_mask': torch.rand(2, 1, 6, generator=g) > 0.5}
    out = m.forward(mem, ids, None, tv.subsequent_mask(5))
    sd = {k: v for k, v in m.state_dict().items()}
    close(O.v2_decode(ids, mem, tv.subsequent_mask(5), sd, 4), out)
    torch.set_grad_enabled(True)
