"""CPU, build container only: the oracle against the UNMODIFIED reference modules imported from
/root/reference at fresh seeds / shapes the fixtures do not cover (skipped on the GPU box)."""
import os
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


@pytest.mark.parametrize('nsrc,srcs', [(2, ('image', 'article')), (4, ('image', 'article', 'faces', 'obj'))])
def test_training_mode_dropout_sites_match_reference(monkeypatch, nsrc, srcs):
    """Pins WHERE the oracle (and so the CUDA training step) applies dropout: the unmodified reference decoder in
    train() mode with F.dropout patched to hand out the counter-based masks in the order the reference calls it
    (decoder_flattened.py:109,283,289,308,326,340,342, dynamic.py:305, multi_head.py:463) -- forward, loss and every
    gradient must equal the oracle's with DropoutMasks of the same seed."""
    import torch.nn.functional as F
    R = ri.load_tell()
    d, H, kernels, V, cut = 32, 4, (3, 7), 300, (100, 200)
    dec = ri.build_ref_decoder(nsrc, d=d, heads=H, ffn=48, kernels=kernels, vocab=V, cutoff=cut, seed=7, relu_dropout=0.05)
    dec.train()
    B, T = 3, 8
    prev, tgt, ctx = synth.make_batch(B, T, V, srcs, seed=11, S_img=3, S_art=10, S_obj=4)
    seed = 0xC0FFEE1234
    cfg = dict(dropout=0.1, input_dropout=0.1, weight_dropout=0.1, attention_dropout=0.1, relu_dropout=0.05)
    ns = len(srcs)
    nat = lambda shape: torch.arange(int(torch.tensor(shape).prod())).view(*shape)
    queue = [(cfg['dropout'], O.DROP_SITE_EMBED, lambda x: nat((T, B, d)).transpose(0, 1))]
    for l in range(len(kernels)):
        st = lambda kind, l=l: l * 16 + kind
        queue.append((cfg['input_dropout'], st(O.DROP_INPUT), lambda x: nat(x.shape)))
        queue.append((cfg['weight_dropout'], st(O.DROP_TAPS), lambda x: nat(x.shape)))
        queue.append((cfg['dropout'], st(O.DROP_CONV_OUT), lambda x: nat(x.shape)))
        for si in range(ns):
            queue.append((cfg['attention_dropout'], st(O.DROP_ATTN_PROB + si), lambda x: nat(x.shape)))
            queue.append((cfg['dropout'], st(O.DROP_ATTN_OUT),
                          lambda x, si=si: (torch.arange(T * B)[:, None] * (ns * d) + si * d + torch.arange(d)[None, :]).view(T, B, d)))
        queue.append((cfg['relu_dropout'], st(O.DROP_RELU), lambda x: nat(x.shape)))
        queue.append((cfg['dropout'], st(O.DROP_FFN_OUT), lambda x: nat(x.shape)))
    calls = []

    def fake_dropout(x, p=0.5, training=True, inplace=False):
        if not training or p == 0:
            return x
        want_p, site, index = queue[len(calls)]
        assert abs(p - want_p) < 1e-12, (len(calls), p, want_p)
        calls.append(site)
        return x * O.dropout_mask(p, seed, site, index(x)).view(x.shape)

    monkeypatch.setattr(F, 'dropout', fake_dropout)
    X, _ = dec({'roberta': prev}, ctx)
    loss, n = R.AdaptiveLoss(1)(dec.adaptive_softmax, (X, None), tgt)
    assert len(calls) == len(queue)
    params = dict(dec.named_parameters())
    g_ref = dict(zip(params, torch.autograd.grad(loss, list(params.values()), allow_unused=True)))
    monkeypatch.undo()
    sd = O.leaf_state_dict(dec.state_dict(), O.tied_output_keys(len(cut) + 1))
    Xo, _ = O.decoder_forward(prev, ctx, sd, H, list(cut) + [V], srcs, drop=O.DropoutMasks(seed=seed, **cfg))
    close(Xo.detach(), X.detach())
    lo, no = O.adaptive_nll(Xo, tgt, sd, list(cut) + [V])
    assert n == no and abs(lo.item() - loss.item()) < 1e-4 * abs(loss.item())
    lo.backward()
    checked = 0
    for k, gr in g_ref.items():
        if gr is None or k not in sd or sd[k].grad is None:
            continue
        close(sd[k].grad, gr, 2e-4)
        checked += 1
    assert checked > 20 * len(kernels)
    # and the masks matter: the eval-mode forward differs
    dec.eval()
    with torch.no_grad():
        Xe, _ = dec({'roberta': prev}, ctx)
    assert (Xe - X.detach()).abs().max() > 1e-2


def test_pointer_decoder_matches_reference():
    """Row f4: the copy head of the unmodified DynamicConvDecoderPointer (decoder_pointer.py) against the oracle's
    pointer_decoder_forward, incl. repeated article token ids (scatter_add) and the nll_loss the model applies."""
    import torch.nn.functional as F
    torch.set_grad_enabled(False)
    ri.load_tell()
    d, H, V, cut = 64, 8, 500, (120, 300)
    dec = ri.build_ref_decoder('pointer', d=d, heads=H, ffn=96, kernels=(3, 15), vocab=V, cutoff=cut, seed=13)
    torch.manual_seed(1)
    dec.project_p_gens.weight_v.normal_(0, 0.3)
    dec.project_p_gens.bias.fill_(0.2)
    sd = dec.state_dict()
    prev, tgt, ctx = synth.make_batch(4, 9, V, ('image', 'article'), seed=21, S_img=4, S_art=17)
    art_ids = torch.randint(0, V, (4, 17), generator=torch.Generator().manual_seed(3))
    art_ids[:, 5] = art_ids[:, 2]                              # a token that occurs twice in the article
    lp, ex = dec({'roberta': prev}, ctx, art_ids)
    lo, attn, pg = O.pointer_decoder_forward(prev, ctx, art_ids, sd, H, list(cut) + [V])
    close(lo, lp)
    close(attn, ex['attn'])
    assert (pg > 0.05).all() and (pg < 0.95).all()            # the mix is exercised on both sides
    loss = F.nll_loss(lp.view(-1, V), tgt.view(-1), ignore_index=1, reduction='sum')
    mask = tgt.view(-1) != 1
    assert abs((-lo.view(-1, V)[mask, tgt.view(-1)[mask]].sum() - loss).item()) < 1e-4 * abs(loss.item())
    torch.set_grad_enabled(True)


@pytest.mark.parametrize('weigh_bert', [True, False])
def test_context_assembly_matches_reference(weigh_bert):
    """Row f2: the UNMODIFIED `TransformerFacesObjectModel._forward` (transformer_faces_objects.py:312-397) -- RoBERTa
    hidden-layer mix with softmax(bert_weight), NaN face / object rows -> masks, the transposes and the target shift --
    run on an instance whose RoBERTa / ResNet members are fakes (the hubs are unreachable), against the oracle's
    context_mix / nan_padding_mask."""
    torch.set_grad_enabled(False)
    Cls = ri.load_faces_objects_model_class()
    g = torch.Generator().manual_seed(5)
    B, S, E, n_hid = 3, 11, 16, 25
    hid = [torch.randn(B, S, E, generator=g) for _ in range(n_hid)]
    feat = torch.randn(B, 2048, 7, 7, generator=g).abs()

    class FakeRoberta:
        def extract_features(self, ids, return_all_hiddens=False):
            assert return_all_hiddens
            return hid

    m = object.__new__(Cls)
    torch.nn.Module.__init__(m)
    m.index, m.padding_idx, m.weigh_bert = 'roberta', 1, weigh_bert
    m.bert_weight = torch.nn.Parameter(torch.randn(n_hid, generator=g))
    object.__setattr__(m, 'roberta', FakeRoberta())
    object.__setattr__(m, 'resnet', lambda image: feat)
    art_ids = torch.randint(3, 100, (B, S), generator=g)
    art_ids[0, 7:] = 1
    cap = torch.randint(3, 100, (B, 9), generator=g)
    faces = torch.randn(B, 4, 512, generator=g)
    faces[0, 2:] = float('nan')
    faces[1, :] = float('nan')
    objs = torch.randn(B, 6, 2048, generator=g)
    objs[2, 1:] = float('nan')
    cap_ids, tgt_ids, ctx = m._forward({'roberta': art_ids}, torch.zeros(B, 3, 224, 224), {'roberta': cap.clone()},
                                       faces.clone(), objs.clone())
    assert torch.equal(cap_ids, cap[:, :-1]) and torch.equal(tgt_ids, cap[:, 1:])
    want_article = O.context_mix(hid, m.bert_weight) if weigh_bert else hid[-1].transpose(0, 1)
    close(want_article, ctx['article'], 1e-6)
    assert torch.equal(ctx['article_mask'], art_ids == 1)
    assert torch.equal(ctx['image'], feat.permute(0, 2, 3, 1).reshape(B, 49, 2048).transpose(0, 1))
    assert ctx['image_mask'].shape == (B, 49) and not ctx['image_mask'].any()
    for key, src in (('faces', faces), ('obj', objs)):
        e, mk = O.nan_padding_mask(src)
        assert torch.equal(e, ctx[key]) and torch.equal(mk, ctx[key + '_mask'])
    torch.set_grad_enabled(True)


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


# ----------------------------------------------------------------------------- beam search (row f1)
def _reference_caption_model():
    """The UNMODIFIED models/CaptionModel.py of the reference (its only non-torch import, misc.utils, is unused by
    beam_search and stubbed)."""
    import importlib.util
    import sys
    import types
    for name in ('misc', 'misc.utils'):
        sys.modules.setdefault(name, types.ModuleType(name))
    path = os.path.join(ri.REF_ROOT, 'models', 'CaptionModel.py')
    spec = importlib.util.spec_from_file_location('ref_caption_model', path)
    mod = importlib.util.module_from_spec(spec)
    spec.loader.exec_module(mod)
    return mod.CaptionModel


@pytest.mark.parametrize('seed,eos_bias', [(0, -30.0), (1, 0.0), (2, 1.5), (3, 3.0)])
def test_beam_search_matches_reference(monkeypatch, seed, eos_bias):
    """oracle.beam_search vs the reference's CaptionModel.beam_search on a small recurrent toy model (the search logic
    is model-agnostic).  End token 0 as in the reference.  The reference's final sort is blind under torch >= 0.4
    (its stored score is a view that the next line overwrites, see the oracle docstring), so it returns the first
    `beam` finished beams in ARRIVAL order: that list must equal the head of the oracle's arrival-order list; with
    no early end token (eos_bias = -30) arrival order == score order and the sorted outputs must agree too."""
    CaptionModel = _reference_caption_model()
    monkeypatch.setattr(torch.Tensor, 'cuda', lambda self, *a, **k: self)
    g = torch.Generator().manual_seed(seed)
    V, Hd, W, L = 23, 16, 4, 7
    emb = torch.randn(V, Hd, generator=g)
    Wh = torch.randn(Hd, Hd, generator=g) * 0.5
    Wo = torch.randn(V, Hd, generator=g)
    bias = torch.zeros(V)
    bias[0] = eos_bias

    def cell(it, h):                                   # h [1, W, Hd]: "dimension one is the beam" as in the reference
        h2 = torch.tanh(emb[it][None] + h @ Wh)
        return torch.log_softmax(h2[0] @ Wo.t() + bias, -1), h2

    class Toy(CaptionModel):
        seq_length = L

        def get_logprobs_state(self, it, state):
            lp, h2 = cell(it, state[0])
            return lp, [h2]

    h0 = torch.zeros(1, W, Hd)
    lp0, h1 = cell(torch.full((W,), 5), h0)
    ref = Toy().beam_search([h1.clone()], lp0.clone(), opt={'beam_size': W})

    st = {'h': h0.clone()}

    def step(tokens):
        it = torch.full((W,), 5) if tokens is None else tokens
        lp, st['h'] = cell(it, st['h'])
        return lp

    def reorder(order):
        st['h'] = st['h'].index_select(1, order)

    mine_all = O.beam_search(step, reorder, W, L, eos=0, return_all=True)
    assert len(ref) == W and len(mine_all) >= W
    for r, m in zip(ref, mine_all[:W]):
        assert torch.equal(r['seq'], m['seq'])
        assert torch.allclose(r['logps'], m['logps'], atol=1e-6)
    st['h'] = h0.clone()
    mine = O.beam_search(step, reorder, W, L, eos=0)
    ps = [float(m['p']) for m in mine]
    assert ps == sorted(ps, reverse=True)
    for m in mine:                                      # p = sum of the per-word log-probs up to the end token
        n = int((m['seq'] != 0).sum()) + (1 if (m['seq'] == 0).any() else 0)
        assert abs(float(m['logps'][:min(n, L)].sum()) - float(m['p'])) < 1e-4
    if eos_bias < -10:
        for r, m in zip(ref, mine):
            assert torch.equal(r['seq'], m['seq'])


def _entity_pointer_reference(d=64, H=8, V=500, cut=(120, 300), hidden=48, seed=17):
    """The unmodified entity-pointer decoder + an instance of the unmodified TransformerEntityPointerModel class that
    holds it (the class's own __init__ needs the RoBERTa / ResNet hubs and hard-codes 2048 -> 1024 head sizes)."""
    R = ri.load_tell()
    Cls = ri.load_entity_pointer_model_class()
    dec = ri.build_ref_decoder('entity_pointer', d=d, heads=H, ffn=96, kernels=(3, 15), vocab=V, cutoff=cut, seed=seed)
    m = object.__new__(Cls)
    torch.nn.Module.__init__(m)
    m.decoder = dec
    torch.manual_seed(seed + 1)
    m.project_first_p_gens = R.GehringLinear(2 * d, hidden, dropout=0.1)
    m.project_p_gens = R.GehringLinear(hidden, 1)
    m.project_first_p_gens.weight_v.normal_(0, 0.2)
    m.project_p_gens.weight_v.normal_(0, 0.5)
    m.project_p_gens.bias.fill_(0.1)
    m.relu_dropout, m.vocab_size, m.index, m.padding_idx = 0.1, V, 'roberta', 1
    m.sampling_topk, m.sampling_temp = 1, 1.0
    m.eval()
    sd = dict(dec.state_dict())
    for k, v in m.state_dict().items():
        if k.startswith('project_'):
            sd[k] = v
    return m, sd


def test_entity_pointer_decoder_matches_reference():
    """Row f4: the 3-source entity-pointer decoder (decoder_entity_pointer.py) and the model's copy head
    (transformer_entity_pointer.py:185-261, unmodified) against the oracle's entity_pointer_forward."""
    torch.set_grad_enabled(False)
    d, H, V, cut = 64, 8, 500, (120, 300)
    m, sd = _entity_pointer_reference(d, H, V, cut)
    prev, tgt, ctx = synth.make_batch(4, 9, V, O.ENTITY_SOURCES, seed=23, S_img=4, S_art=17, S_ent=11, d_ent=1024)
    ctx['sections'], ctx['sections_mask'] = None, None
    ent = torch.randint(0, V, (4, 11), generator=torch.Generator().manual_seed(4))
    ent[:, 6] = ent[:, 1]                                       # an entity token that occurs twice
    out = m.decoder({'roberta': prev}, ctx)
    lp = m.output_layer(out, ent)
    lo, attn, pg = O.entity_pointer_forward(prev, ctx, ent, sd, H, list(cut) + [V])
    close(attn, out[1]['attn'])
    close(lo, lp)
    assert (pg > 0.05).all() and (pg < 0.95).all()
    torch.set_grad_enabled(True)


def test_pointer_greedy_generation_matches_reference():
    """The unmodified `_generate` loop of the entity-pointer model (transformer_entity_pointer.py:375-470; 100 steps,
    sampling_topk = 1 so the multinomial draw is deterministic) against the oracle's pointer_greedy_generate."""
    torch.set_grad_enabled(False)
    d, H, V, cut = 64, 8, 500, (120, 300)
    m, sd = _entity_pointer_reference(d, H, V, cut, seed=29)
    sd['adaptive_softmax.head.class_proj.weight'] *= 0.2       # keep most of the mass in the head: EOS (id 2) can win
    m.decoder.load_state_dict(sd, strict=False)
    prev, _, ctx = synth.make_batch(5, 3, V, O.ENTITY_SOURCES, seed=31, S_img=4, S_art=17, S_ent=11, d_ent=1024)
    ent = torch.randint(0, V, (5, 11), generator=torch.Generator().manual_seed(6))
    ent[:, :3] = 2                                              # EOS among the copy candidates: some rows finish early
    lp_r, ids_r = m._generate(prev[:, :1].clone(), ctx, ent)
    lp_o, ids_o = O.pointer_greedy_generate(prev[:, :1], ctx, ent, sd, H, list(cut) + [V], entity=True)
    assert ids_r.shape == ids_o.shape and torch.equal(ids_r, ids_o)
    close(lp_o, lp_r)
    torch.set_grad_enabled(True)


@pytest.mark.parametrize('separate,bias_row,zero_row', [(False, True, True), (True, True, False), (False, False, False)])
def test_attention_score_forward_matches_reference(separate, bias_row, zero_row):
    """oracle `mha_score` against the unmodified `multi_head_attention_score_forward` (multi_head.py:14-204)."""
    import importlib
    ri.load_tell()
    mh = importlib.import_module('tell.modules.attention.multi_head')
    g = torch.Generator().manual_seed(31)
    T, B, E, H, S = 5, 3, 64, 4, 11
    q = torch.randn(T, B, E, generator=g)
    mem = torch.randn(S, B, E, generator=g)
    w = torch.randn(2 * E, E, generator=g) * 0.1
    b = torch.randn(2 * E, generator=g) * 0.1
    bias_k = torch.randn(1, 1, E, generator=g) if bias_row else None
    pad = torch.rand(B, S, generator=g) > 0.6
    pad[:, 0] = False
    kw = dict(use_separate_proj_weight=True, q_proj_weight=w[:E], k_proj_weight=w[E:]) if separate else {}
    ref = mh.multi_head_attention_score_forward(q, mem, E, H, None if separate else w, b, bias_k, zero_row, 0.0, None, None,
                                                training=False, key_padding_mask=pad, **kw)
    got = O.mha_score(q, mem, w[:E], w[E:], b[:E], b[E:], bias_k, zero_row, pad, H)
    assert ref.shape == (B, T, S + int(bias_row) + int(zero_row))
    close(got, ref, 1e-6)
    ref_self = mh.multi_head_attention_score_forward(q, q, E, H, w, b, bias_k, zero_row, 0.0, None, None, training=False)
    close(O.mha_score(q, q, w[:E], w[E:], b[:E], b[E:], bias_k, zero_row, None, H), ref_self, 1e-6)   # qk_same branch (:92-95)


@pytest.mark.parametrize('nsrc,srcs', [(2, ('image', 'article')), (4, ('image', 'article', 'faces', 'obj'))])
def test_module_generate_loop_around_reference_decoder(nsrc, srcs):
    """`O.module_generate_loop` (the reference `_generate` body, transformer_flattened.py:240-325, restated around the module
    protocol: filter_incremental_state -> decoder(prev, contexts_i, incremental_state) -> get_normalized_probs) driven
    around the UNMODIFIED reference decoder must reproduce the functional oracle's greedy loop -- ids exactly, log-probs
    5e-5 -- including rows that finish early and leave the batch."""
    torch.set_grad_enabled(False)
    d, H, V, cut = 64, 4, 400, [100, 250]
    dec = ri.build_ref_decoder(nsrc, d=d, heads=H, ffn=96, kernels=[3, 7], vocab=V, cutoff=list(cut)).eval()
    sd = {k: v for k, v in dec.state_dict().items()}
    sd['adaptive_softmax.head.class_proj.weight'] *= 0.2         # mass stays in the head: EOS (id 2) can win for some rows
    sd['embedder.token_embedder_adaptive.embeddings.0.0.weight'][2] *= 3.0
    dec.load_state_dict(sd, strict=False)
    sd = dec.state_dict()
    prev, _, ctx = synth.make_batch(6, 3, V, srcs, seed=13, S_img=5, S_art=19, S_obj=6, S_faces=3)
    lp_m, ids_m = O.module_generate_loop(dec, prev[:, :1].clone(), ctx, srcs, gen_len=14)
    lp_o, ids_o = O.greedy_generate(prev[:, :1], ctx, sd, H, list(cut) + [V], srcs, gen_len=14)
    assert ids_m.shape == ids_o.shape and torch.equal(ids_m, ids_o)
    close(lp_o, lp_m)
    torch.set_grad_enabled(True)


@pytest.mark.parametrize('T,K,unfold', [(9, 5, False), (9, 5, True), (3, 7, False)])
def test_lightweight_conv_matches_reference(T, K, unfold):
    """O.lightweight_conv against the unmodified LightweightConv1dTBC (lightweight.py:86-253), both of its paths
    (band-matrix and unfolded) incl. K > T, with bias, and gradients."""
    install = ri.load_tell()          # stubs + imports
    del install
    import tell.modules as M
    torch.manual_seed(5)
    B, C, H = 3, 24, 4
    m = M.LightweightConv1dTBC(C, K, padding_l=K - 1, num_heads=H, weight_softmax=True, bias=True).eval()
    with torch.no_grad():
        m.bias.normal_()
    x = torch.randn(T, B, C, requires_grad=True)
    y = m(x, unfold=unfold)
    gy = torch.randn_like(y)
    gx, gw = torch.autograd.grad(y, [x, m.weight], gy)
    x2 = x.detach().clone().requires_grad_(True)
    w2 = m.weight.detach().clone().requires_grad_(True)
    yo = O.lightweight_conv(x2, w2, H, K, m.bias.detach())
    gx2, gw2 = torch.autograd.grad(yo, [x2, w2], gy)
    close(yo, y.detach())
    close(gx2, gx)
    close(gw2, gw)
