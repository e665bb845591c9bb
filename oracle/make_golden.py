"""TEST INFRASTRUCTURE -- freezes outputs of the UNMODIFIED reference (imported from
/root/reference via oracle/ref_import.py) into small fixtures under tests/golden/.

Run in the build container only:   python -m oracle.make_golden
The fixtures travel to the GPU box; the reference does not.
"""
import os
import warnings

import torch

from . import ref_import as ri
from . import synth

OUT = os.path.join(os.path.dirname(os.path.dirname(os.path.abspath(__file__))), 'tests', 'golden')


def ref_greedy(dec, seed, ctx, srcs, gen_len):
    """The reference `_generate` loop body (transformer_flattened.py:238-325) driven around the REAL reference decoder
    object; top-1 sampling == argmax.  (The loop lives in oracle/restate.py so the GPU tests can drive the mirror with it.)"""
    from .restate import module_generate_loop
    return module_generate_loop(dec, seed, ctx, srcs, gen_len)


def tell_case(name, nsrc, srcs, d, H, ffn, kernels, V, cut, B, T, S_img, S_art, gen_len):
    R = ri.load_tell()
    dec = ri.build_ref_decoder(nsrc, d=d, heads=H, ffn=ffn, kernels=kernels, vocab=V, cutoff=cut)
    prev, tgt, ctx = synth.make_batch(B, T, V, srcs, S_img=S_img, S_art=S_art, S_obj=6, S_faces=4)
    X, ex = dec({'roberta': prev}, ctx)
    loss, ntok = R.AdaptiveLoss(1)(dec.adaptive_softmax, (X, None), tgt)
    lp = dec.get_normalized_probs((X[:, -1:], None), log_probs=True)
    g_lp, g_ids = ref_greedy(dec, prev[:, :1], ctx, srcs, gen_len)
    # per-module probes on layer 0
    L0 = dec.layers[0]
    xin = torch.randn(T, B, d, generator=torch.Generator().manual_seed(7))
    conv_y = L0.conv(xin)
    fix = dict(
        config=dict(n_sources=nsrc, sources=list(srcs), d=d, heads=H, ffn=ffn,
                    kernels=list(kernels), vocab=V, cutoffs=list(cut) + [V]),
        state_dict={k: v.clone() for k, v in dec.state_dict().items()},
        prev=prev, target=tgt, contexts=ctx, X=X, attn=ex['attn'], loss=loss, ntokens=ntok,
        logprob_last=lp, greedy_logprobs=g_lp, greedy_ids=g_ids, conv_in=xin, conv_out=conv_y)
    torch.save(fix, os.path.join(OUT, name))
    print(name, 'loss', float(loss), 'greedy', g_ids.shape,
          os.path.getsize(os.path.join(OUT, name)) // 1024, 'KiB')


def dynconv_cases():
    """Module-level DynamicConv1dTBC vectors incl. K > T (tap narrowing) and both code paths."""
    R = ri.load_tell()
    cases = []
    g = torch.Generator().manual_seed(11)
    for (T, B, C, H, K) in [(5, 2, 32, 4, 3), (4, 3, 32, 4, 7), (9, 2, 64, 8, 15), (40, 1, 32, 2, 31),
                            (1, 2, 32, 4, 7)]:
        torch.manual_seed(T * 100 + K)
        m = R.DynamicConv1dTBC(C, K, padding_l=K - 1, num_heads=H, weight_softmax=True).eval()
        x = torch.randn(T, B, C, generator=g)
        y_exp = m(x, unfold=False)
        y_unf = m(x, unfold=True)
        # NB the two reference paths AGREE only for K <= T: for K > T the unfold path narrows
        # the tap logits BEFORE the softmax (dynamic.py:149-151,165-166) while the expanded
        # path -- the one every decoder call takes (T <= 512, :102) -- narrows AFTER (:304,324-326).
        if K <= T:
            assert (y_exp - y_unf).abs().max() < 1e-5
        # incremental: feed one step at a time
        st, ys = {}, []
        for t in range(T):
            ys.append(m(x[t:t + 1], incremental_state=st))
        cases.append(dict(T=T, B=B, C=C, H=H, K=K, w=m.weight_linear.weight.detach().clone(),
                          x=x, y=y_exp, y_unfold_path=y_unf, y_incremental=torch.cat(ys, 0)))
    torch.save(cases, os.path.join(OUT, 'dynconv_cases.pt'))
    print('dynconv_cases.pt', len(cases))


def mha_cases():
    """MultiHeadAttention vectors: packed and separate projections, a fully masked source."""
    R = ri.load_tell()
    cases = []
    g = torch.Generator().manual_seed(12)
    for (T, B, E, H, S, kdim, full_mask_row) in [(6, 3, 32, 4, 10, 32, False), (5, 2, 32, 4, 4, 48, True),
                                                 (1, 4, 64, 8, 33, 24, False)]:
        torch.manual_seed(S)
        m = R.MultiHeadAttention(E, H, kdim=kdim, vdim=kdim, dropout=0.1).eval()
        with torch.no_grad():
            m.in_proj_bias.normal_(0, 0.1, generator=g)
            m.out_proj.bias.normal_(0, 0.1, generator=g)
        q = torch.randn(T, B, E, generator=g)
        mem = torch.randn(S, B, kdim, generator=g)
        n = torch.randint(1, S + 1, (B,), generator=g)
        if full_mask_row:
            n[0] = 0
        mask = torch.arange(S)[None, :] >= n[:, None]
        out, w = m(q, mem, mem, key_padding_mask=mask, incremental_state=None, static_kv=True,
                   need_weights=True)
        cases.append(dict(T=T, B=B, E=E, H=H, S=S, kdim=kdim,
                          state_dict={k: v.clone() for k, v in m.state_dict().items()},
                          query=q, memory=mem, mask=mask, out=out, weights=w))
    torch.save(cases, os.path.join(OUT, 'mha_cases.pt'))
    print('mha_cases.pt', len(cases))


def v2_case():
    tv = ri.load_final_v2('final_roberta2')
    torch.manual_seed(0)
    V, d, h = 60, 64, 4
    m = tv.make_model_news(V, 2, d, 128, attn_heads=h, img_dim=32, sent_dim=48).eval()
    g = torch.Generator().manual_seed(5)
    B, T = 3, 7
    ids = torch.randint(0, V, (B, T), generator=g)
    y = torch.randint(0, V, (B, T), generator=g)
    y[0, 5:] = 1
    mem = {'image': torch.randn(B, 5, 32, generator=g), 'article': torch.randn(B, 9, 48, generator=g),
           'article_mask': torch.rand(B, 1, 9, generator=g) > 0.3}
    tm = tv.subsequent_mask(T)
    out = m.forward(mem, ids, None, tm)
    logits = m.generator.proj(out)
    ntok = int((y != 1).sum())
    loss = torch.nn.CrossEntropyLoss(ignore_index=1, reduction='sum')(
        logits.reshape(-1, V), y.reshape(-1)) / ntok
    # translate_sentence (custom_v4.py:647-662) around the real model
    ys = torch.zeros(B, 1, dtype=torch.long)
    for _ in range(5):
        o = m.decode(mem, None, ys, tv.subsequent_mask(ys.size(1)))
        prob = m.generator(o[:, -1])
        ys = torch.cat([ys, prob.max(dim=1)[1].unsqueeze(1)], 1)
    torch.save(dict(config=dict(vocab=V, d=d, h=h, layers=2, ffn=128),
                    state_dict={k: v.clone() for k, v in m.state_dict().items() if not k.endswith('.pe')},
                    ids=ids, y=y, memory=mem, tgt_mask=tm, out=out, loss=loss, ntokens=ntok,
                    greedy=ys), os.path.join(OUT, 'v2_small.pt'))
    print('v2_small.pt loss', float(loss))


def pointer_case():
    """Row f4: the unmodified DynamicConvDecoderPointer (decoder_pointer.py): lprobs of the copy head + the model's loss."""
    import torch.nn.functional as F
    ri.load_tell()
    d, H, V, cut = 32, 4, 300, (100, 200)
    dec = ri.build_ref_decoder('pointer', d=d, heads=H, ffn=64, kernels=(3, 7), vocab=V, cutoff=cut, seed=17)
    torch.manual_seed(2)
    dec.project_p_gens.weight_v.normal_(0, 0.4)
    dec.project_p_gens.bias.fill_(-0.3)
    prev, tgt, ctx = synth.make_batch(4, 10, V, ('image', 'article'), S_img=5, S_art=12)
    art_ids = torch.randint(0, V, (4, 12), generator=torch.Generator().manual_seed(5))
    art_ids[:, 7] = art_ids[:, 1]
    lp, ex = dec({'roberta': prev}, ctx, art_ids)
    loss = F.nll_loss(lp.view(-1, V), tgt.view(-1), ignore_index=1, reduction='sum')
    # greedy generation through the unmodified model `_generate` (transformer_context_pointer.py:467-560, sampling_topk = 1)
    m = object.__new__(ri.load_context_pointer_model_class())
    torch.nn.Module.__init__(m)
    m.decoder, m.index, m.padding_idx, m.sampling_topk, m.sampling_temp = dec, 'roberta', 1, 1, 1.0
    m.eval()
    gen_ids_in = art_ids.clone()
    gen_ids_in[:, :2] = 2                                      # EOS among the copy candidates
    gen_lp, gen_ids = m._generate(prev[:, :1].clone(), ctx, gen_ids_in)
    fix = dict(config=dict(n_sources=2, sources=['image', 'article'], d=d, heads=H, ffn=64, kernels=[3, 7], vocab=V,
                           cutoffs=list(cut) + [V]),
               state_dict={k: v.clone() for k, v in dec.state_dict().items()}, prev=prev, target=tgt, contexts=ctx,
               article_ids=art_ids, lprobs=lp, attn=ex['attn'], loss=loss, ntokens=int((tgt != 1).sum()),
               gen_article_ids=gen_ids_in, gen_logp=gen_lp, gen_ids=gen_ids)
    torch.save(fix, os.path.join(OUT, 'pointer_small.pt'))
    print('pointer_small.pt loss', float(loss), os.path.getsize(os.path.join(OUT, 'pointer_small.pt')) // 1024, 'KiB')


def entity_pointer_case():
    """Row f4: the unmodified DynamicConvDecoderEntityPointer (decoder_entity_pointer.py) under the unmodified
    TransformerEntityPointerModel.output_layer / ._generate (transformer_entity_pointer.py:185-261, 375-470)."""
    import torch.nn.functional as F
    R = ri.load_tell()
    d, H, V, cut, hidden = 32, 4, 300, (100, 200), 24
    dec = ri.build_ref_decoder('entity_pointer', d=d, heads=H, ffn=64, kernels=(3, 7), vocab=V, cutoff=cut, seed=19)
    m = object.__new__(ri.load_entity_pointer_model_class())
    torch.nn.Module.__init__(m)
    m.decoder = dec
    torch.manual_seed(3)
    m.project_first_p_gens = R.GehringLinear(2 * d, hidden, dropout=0.1)
    m.project_p_gens = R.GehringLinear(hidden, 1)
    m.project_first_p_gens.weight_v.normal_(0, 0.3)
    m.project_p_gens.weight_v.normal_(0, 0.6)
    m.project_p_gens.bias.fill_(-0.2)
    m.relu_dropout, m.vocab_size, m.index, m.padding_idx, m.sampling_topk, m.sampling_temp = 0.1, V, 'roberta', 1, 1, 1.0
    m.eval()
    sources = ('image', 'article', 'entity')
    prev, tgt, ctx = synth.make_batch(4, 10, V, sources, S_img=5, S_art=12, S_ent=9, d_ent=1024)
    ent = torch.randint(0, V, (4, 9), generator=torch.Generator().manual_seed(7))
    ent[:, 5] = ent[:, 0]
    out = dec({'roberta': prev}, ctx)
    lp = m.output_layer(out, ent)
    loss = F.nll_loss(lp.view(-1, V), tgt.view(-1), ignore_index=1, reduction='sum')
    gen_ent = ent.clone()
    gen_ent[:, :2] = 2
    gen_lp, gen_ids = m._generate(prev[:, :1].clone(), ctx, gen_ent)
    sd = {k: v.clone() for k, v in dec.state_dict().items()}
    sd.update({k: v.clone() for k, v in m.state_dict().items() if k.startswith('project_')})
    fix = dict(config=dict(sources=list(sources), d=d, heads=H, ffn=64, kernels=[3, 7], vocab=V, cutoffs=list(cut) + [V],
                           hidden=hidden),
               state_dict=sd, prev=prev, target=tgt, contexts=ctx, entity_tokens=ent, lprobs=lp, attn=out[1]['attn'],
               predictors=out[1]['predictors'], loss=loss, ntokens=int((tgt != 1).sum()),
               gen_entity_tokens=gen_ent, gen_logp=gen_lp, gen_ids=gen_ids)
    torch.save(fix, os.path.join(OUT, 'entity_pointer_small.pt'))
    print('entity_pointer_small.pt loss', float(loss), 'gen', tuple(gen_ids.shape),
          os.path.getsize(os.path.join(OUT, 'entity_pointer_small.pt')) // 1024, 'KiB')


def main():
    warnings.filterwarnings('ignore')
    os.makedirs(OUT, exist_ok=True)
    torch.set_grad_enabled(False)
    tell_case('tell2_small.pt', 2, ('image', 'article'), d=32, H=4, ffn=64, kernels=(3, 7),
              V=300, cut=(100, 200), B=4, T=10, S_img=5, S_art=12, gen_len=12)
    tell_case('tell4_small.pt', 4, ('image', 'article', 'faces', 'obj'), d=32, H=4, ffn=64,
              kernels=(15,), V=300, cut=(100, 200), B=8, T=6, S_img=3, S_art=9, gen_len=8)
    dynconv_cases()
    mha_cases()
    v2_case()
    pointer_case()
    entity_pointer_case()


if __name__ == '__main__':
    main()
