"""GPU: edge cases of the decoder engine against the oracle -- single row / single step, odd batch and ragged source
lengths, a caption whose article is padding only, an empty source, the longest caption the position table allows, and
the error behaviour of the boundary (bad shapes, missing contexts, too-small workspaces fail loudly, nothing falls back)."""
import ctypes as C

import pytest
import torch

from oracle import restate as O
from oracle import synth

pytestmark = pytest.mark.gpu

D, H, V, CUT = 128, 2, 300, [100, 200, 300]


def rel_err(a, b):
    a, b = a.float().cpu(), b.float().cpu()
    return ((a - b).abs().max() / b.abs().max().clamp_min(1e-6)).item()


def make(dtype, sources=('image', 'article'), kernels=(3, 7), seed=2):
    from news_image_caption_b200 import DynamicConvDecoder, build
    from news_image_caption_b200.decoder import random_state_dict
    build.build()
    sd = random_state_dict(d=D, heads=H, ffn=128, kernels=kernels, vocab=V, cutoffs=CUT, sources=sources, seed=seed)
    dec = DynamicConvDecoder(decoder_output_dim=D, decoder_attention_heads=H, decoder_ffn_embed_dim=128,
                             decoder_kernel_size_list=list(kernels), vocab_size=V, adaptive_softmax_cutoff=CUT[:-1],
                             sources=sources, dtype=dtype)
    dec.load_reference_state_dict(sd)
    return dec, sd


def check(dec, sd, prev, tgt, ctx, dtype, sources=('image', 'article')):
    X, ex = dec(prev.cuda(), {k: v.cuda() for k, v in ctx.items()})
    out = dec.nll(X, tgt.cuda())
    Xo, attn = O.decoder_forward(prev, ctx, sd, H, CUT, sources)
    loss, n = O.adaptive_nll(Xo, tgt, sd, CUT)
    tol = 1e-4 if dtype == torch.float32 else 3e-2
    assert rel_err(X, Xo) < tol
    assert rel_err(ex['attn'], attn) < tol
    assert int(out[1].item()) == n
    if n:
        assert abs(out[0].item() - loss.item()) < tol * abs(loss.item())
    else:
        assert out[0].item() == 0.0
    return X


@pytest.mark.parametrize('dtype', [torch.float32, torch.bfloat16])
@pytest.mark.parametrize('B,T,S_img,S_art', [(1, 1, 1, 1), (3, 5, 7, 77), (5, 33, 49, 130)])
def test_odd_shapes(dtype, B, T, S_img, S_art):
    dec, sd = make(dtype)
    prev, tgt, ctx = synth.make_batch(B, T, V, seed=B * 100 + T, S_img=S_img, S_art=S_art)
    check(dec, sd, prev, tgt, ctx, dtype)


@pytest.mark.parametrize('dtype', [torch.float32, torch.bfloat16])
def test_article_of_padding_only_and_all_pad_targets(dtype):
    """Row 1's article is entirely padding: its attention sees only the two extra keys (bias_k and the zero key,
    multi_head.py:212-214, 393-411).  Row 2's targets are all <pad>: it contributes nothing to the loss."""
    dec, sd = make(dtype)
    prev, tgt, ctx = synth.make_batch(4, 6, V, seed=41, S_img=5, S_art=140)
    ctx['article_mask'][1] = True
    ctx['article'][:, 1] = 0
    tgt[2] = 1
    X = check(dec, sd, prev, tgt, ctx, dtype)
    assert torch.isfinite(X.float()).all()
    # every target padded: loss 0, count 0
    out = dec.nll(X, torch.ones_like(tgt).cuda())
    assert out[0].item() == 0.0 and int(out[1].item()) == 0


@pytest.mark.parametrize('dtype', [torch.float32, torch.bfloat16])
def test_empty_source(dtype):
    """A batch without any detected face: the faces context is [0, B, 512] and the attention reduces to the two extra keys
    (decoder_faces_objects.py:341-359 with an empty key tensor)."""
    sources = ('image', 'article', 'faces', 'obj')
    dec, sd = make(dtype, sources, kernels=(3,))
    prev, tgt, ctx = synth.make_batch(3, 4, V, sources, seed=43, S_img=5, S_art=20, S_faces=0, S_obj=6)
    assert ctx['faces'].shape[0] == 0
    check(dec, sd, prev, tgt, ctx, dtype, sources)


def test_longest_caption_and_position_limit():
    """T = 511 is the longest input the 513-row position table holds (positions start at padding_idx + 1); T = 512 is
    refused by the engine instead of reading past the table."""
    from news_image_caption_b200 import TntError
    dec, sd = make(torch.bfloat16, kernels=(31,))
    prev, tgt, ctx = synth.make_batch(2, 511, V, seed=47, S_img=3, S_art=9)
    check(dec, sd, prev, tgt, ctx, torch.bfloat16)
    prev2 = torch.cat([prev, prev[:, :1]], 1)
    with pytest.raises(TntError):
        dec(prev2.cuda(), {k: v.cuda() for k, v in ctx.items()})
    with pytest.raises(TntError):
        dec.generate(prev[:, :1].cuda(), {k: v.cuda() for k, v in ctx.items()}, gen_len=512)


def test_boundary_errors_are_loud():
    from news_image_caption_b200 import TntError, _lib
    dec, sd = make(torch.bfloat16)
    prev, tgt, ctx = synth.make_batch(2, 4, V, seed=53, S_img=3, S_art=9)
    cc = {k: v.cuda() for k, v in ctx.items()}
    with pytest.raises(KeyError):                                   # a context the decoder was built for is missing
        dec(prev.cuda(), {k: v for k, v in cc.items() if not k.startswith('article')})
    with pytest.raises(ValueError):                                 # wrong feature width
        dec(prev.cuda(), dict(cc, image=cc['image'][:, :, :100]))
    with pytest.raises(ValueError):                                 # mask of the wrong shape
        dec(prev.cuda(), dict(cc, article_mask=cc['article_mask'][:, :5]))
    # the C ABI itself: a workspace that is too small is refused with an error code and a message
    L, h = _lib.lib(), dec._h()
    c, keep = dec._contexts(cc, 2, 4)
    X = torch.empty(2, 4, D, dtype=torch.bfloat16, device='cuda')
    ws = torch.empty(1024, dtype=torch.uint8, device='cuda')
    ids = prev.cuda().contiguous()
    rc = L.tnt_decoder_forward(h, ids.data_ptr(), C.byref(c), X.data_ptr(), None, ws.data_ptr(), ws.numel(), None)
    assert rc != 0 and b'workspace' in L.tnt_last_error()
    with pytest.raises(TntError):
        _lib.check(rc, 'tnt_decoder_forward')
    rc = L.tnt_decoder_forward(h, None, C.byref(c), X.data_ptr(), None, ws.data_ptr(), ws.numel(), None)
    assert rc != 0


def test_out_of_range_token_id_is_loud():
    """nn.Embedding raises IndexError for an id outside [0, vocab) (adaptive.py:61-76).  The engine never reads outside
    its table: the offending row is written as NaN, a device error flag is raised and the mirror turns it into an
    IndexError wherever it synchronises (loss .item(), trimmed generation); the flag clears on read."""
    from news_image_caption_b200 import AdaptiveLoss
    dec, sd = make(torch.bfloat16)
    prev, tgt, ctx = synth.make_batch(3, 4, V, seed=59, S_img=3, S_art=9)
    cc = {k: v.cuda() for k, v in ctx.items()}
    bad = prev.clone()
    bad[1, 2] = V + 7
    X, _ = dec(bad.cuda(), cc)
    assert torch.isnan(X[1].float()).any() and torch.isfinite(X[0].float()).all()
    with pytest.raises(IndexError):
        AdaptiveLoss(1)(dec, (X, None), tgt.cuda())
    bad[1, 2] = -3
    dec(bad.cuda(), cc)
    with pytest.raises(IndexError):
        dec.check_input_errors()
    X, _ = dec(prev.cuda(), cc)                                       # a clean batch afterwards is clean
    loss, n = AdaptiveLoss(1)(dec, (X, None), tgt.cuda())
    assert torch.isfinite(loss).all() and n == int((tgt != 1).sum())


@pytest.mark.parametrize('dtype', [torch.float32, torch.bfloat16])
def test_host_scoring_copies_live_rows_only(dtype):
    """tnt_decoder_score_host / _submit: ids, targets, contexts and masks in PINNED HOST memory.  Only the live rows of a
    masked source cross PCIe (one strided 2-D copy per caption up to its last un-masked position) when skipping the padding
    pays for the copies' fixed cost (S_art = 300); a source too small for that (S_art = 37) crosses as one contiguous copy
    and its padded tails are cleared on the device.  Either way the result must equal the device-resident forward + loss --
    incl. a caption whose article is all padding, one without padding, masks with holes, NaN in the padded tail of the host
    tensor (never seen by a kernel), and a second batch reusing the staging buffer."""
    from news_image_caption_b200 import _lib
    dec, sd = make(dtype)
    pin = lambda t: t.contiguous().pin_memory()
    for seed, poison, S_art in ((61, True, 37), (67, False, 37), (71, True, 300), (73, False, 300)):
        prev, tgt, ctx = synth.make_batch(5, 6, V, seed=seed, S_img=4, S_art=S_art)
        ctx['article_mask'][0] = True                     # nothing live
        ctx['article_mask'][1] = False                    # everything live
        ctx['article_mask'][2, 3:9] = True                # a hole inside the live prefix (those rows are copied, masked later)
        cc = {k: v.cuda() for k, v in ctx.items()}
        X, _ = dec(prev.cuda(), cc)
        want = dec.nll(X, tgt.cuda()).cpu()
        host = {k: pin(v.to(dec.torch_dtype) if v.is_floating_point() else v) for k, v in ctx.items()}
        if poison:                                        # the padded TAIL of every caption never reaches a kernel
            m = ctx['article_mask']
            for b in range(5):
                live = (~m[b]).nonzero()
                n = int(live.max()) + 1 if len(live) else 0
                host['article'][n:, b] = float('nan')
        l, n_tok = dec.score_host(pin(prev), pin(tgt), host)
        assert n_tok == want[1].item() and abs(l - want[0].item()) < 1e-5 * abs(want[0].item())
        full = sum(v.numel() * v.element_size() for v in host.values()) + 2 * prev.numel() * 8
        sent = _lib.lib().tnt_decoder_last_h2d_bytes(dec._h())
        assert sent < full if S_art == 300 else sent <= full     # fewer bytes than the padded tensors hold
        dec.score_host_submit(0, pin(prev), pin(tgt), host)
        l2, n2 = dec.score_host_wait(0)
        assert n2 == n_tok and abs(l2 - l) < 1e-6 * abs(l)
