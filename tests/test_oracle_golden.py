"""CPU: the oracle (oracle/restate.py) against the golden fixtures frozen from the UNMODIFIED
reference (oracle/make_golden.py).  Runs anywhere (no /root/reference, no GPU)."""
import os

import pytest
import torch

from oracle import restate as O

G = os.path.join(os.path.dirname(__file__), 'golden')


def load(name):
    return torch.load(os.path.join(G, name), weights_only=False)


def close(a, b, tol=2e-5):
    assert a.shape == b.shape, (a.shape, b.shape)
    err = (a - b).abs().max().item()
    scale = max(1.0, b.abs().max().item())
    assert err <= tol * scale, 'max abs err %g (scale %g)' % (err, scale)


def test_dynconv_cases():
    for c in load('dynconv_cases.pt'):
        y = O.dynamic_conv(c['x'], c['w'], c['H'], c['K'])
        close(y, c['y'])
        # incremental: one step at a time == the expanded path of the full sequence
        hist, ys = None, []
        for t in range(c['T']):
            yt, hist = O.dynamic_conv_step(c['x'][t:t + 1], hist, c['w'], c['H'], c['K'])
            ys.append(yt)
        close(torch.cat(ys, 0), c['y_incremental'])


def test_mha_cases():
    for c in load('mha_cases.pt'):
        sd = c['state_dict']
        out, w = O.mha(c['query'], c['memory'], c['mask'], sd, '', c['H'])
        close(out, c['out'])
        close(w, c['weights'])


@pytest.mark.parametrize('name', ['tell2_small.pt', 'tell4_small.pt'])
def test_tell_decoder(name):
    f = load(name)
    cfg, sd = f['config'], f['state_dict']
    H, cut, srcs = cfg['heads'], cfg['cutoffs'], cfg['sources']
    K0 = cfg['kernels'][0]
    close(O.dynamic_conv(f['conv_in'], sd['layers.0.conv.weight_linear.weight'], H, K0), f['conv_out'])
    X, attn = O.decoder_forward(f['prev'], f['contexts'], sd, H, cut, srcs)
    close(X, f['X'], 5e-5)
    close(attn, f['attn'], 5e-5)
    loss, ntok = O.adaptive_nll(X, f['target'], sd, cut)
    assert ntok == f['ntokens']
    assert abs(loss.item() - f['loss'].item()) <= 1e-4 * abs(f['loss'].item())
    close(O.adaptive_log_prob(X[:, -1:], sd, cut), f['logprob_last'], 5e-5)
    L = f['greedy_logprobs'].shape[1]
    lp, ids = O.greedy_generate(f['prev'][:, :1], f['contexts'], sd, H, cut, srcs, gen_len=L)
    assert torch.equal(ids, f['greedy_ids'])          # bit-exact ids
    close(lp, f['greedy_logprobs'], 5e-5)


def test_incremental_equals_full():
    f = load('tell2_small.pt')
    cfg, sd = f['config'], f['state_dict']
    full, _ = O.decoder_forward(f['prev'], f['contexts'], sd, cfg['heads'], cfg['cutoffs'], cfg['sources'])
    st, outs = {}, []
    for t in range(f['prev'].shape[1]):
        x, _ = O.decoder_forward(f['prev'][:, t:t + 1], f['contexts'], sd, cfg['heads'], cfg['cutoffs'],
                                 cfg['sources'], st)
        outs.append(x)
    inc = torch.cat(outs, 1)
    # pad inputs take position row 1 in both modes, so only the non-pad prefix offsets agree when
    # pads are absent before t; compare rows without pads
    rows = (f['prev'] != 1).all(1)
    close(inc[rows], full[rows], 5e-5)


def test_v2():
    f = load('v2_small.pt')
    sd, h = f['state_dict'], f['config']['h']
    out = O.v2_decode(f['ids'], f['memory'], f['tgt_mask'], sd, h)
    close(out, f['out'], 5e-5)
    loss = O.v2_generator_loss(out, f['y'], sd, 1, f['ntokens'])
    assert abs(loss.item() - f['loss'].item()) <= 1e-4 * abs(f['loss'].item())
    ys = O.v2_translate(f['memory'], sd, h, f['greedy'].shape[1])
    assert torch.equal(ys, f['greedy'])


def test_pointer_decoder():
    """Row f4: the oracle's copy head against the fixture frozen from the unmodified DynamicConvDecoderPointer."""
    f = load('pointer_small.pt')
    cfg = f['config']
    lp, attn, pg = O.pointer_decoder_forward(f['prev'], f['contexts'], f['article_ids'], f['state_dict'], cfg['heads'],
                                             cfg['cutoffs'])
    close(lp, f['lprobs'])
    close(attn, f['attn'])
    tgt = f['target'].view(-1)
    m = tgt != 1
    loss = -lp.view(-1, cfg['vocab'])[m, tgt[m]].sum()
    assert int(m.sum()) == f['ntokens'] and abs(loss.item() - f['loss'].item()) < 1e-4 * abs(f['loss'].item())
    glp, gids = O.pointer_greedy_generate(f['prev'][:, :1], f['contexts'], f['gen_article_ids'], f['state_dict'],
                                          cfg['heads'], cfg['cutoffs'])
    assert torch.equal(gids, f['gen_ids'])
    close(glp, f['gen_logp'])


def test_entity_pointer_decoder():
    """Row f4: the oracle's 3-source entity-pointer decoder + copy head and its greedy generation against the fixture
    frozen from the unmodified DynamicConvDecoderEntityPointer / TransformerEntityPointerModel."""
    f = load('entity_pointer_small.pt')
    cfg = f['config']
    lp, attn, pg = O.entity_pointer_forward(f['prev'], f['contexts'], f['entity_tokens'], f['state_dict'], cfg['heads'],
                                            cfg['cutoffs'])
    close(lp, f['lprobs'])
    close(attn, f['attn'])
    tgt = f['target'].view(-1)
    m = tgt != 1
    loss = -lp.view(-1, cfg['vocab'])[m, tgt[m]].sum()
    assert int(m.sum()) == f['ntokens'] and abs(loss.item() - f['loss'].item()) < 1e-4 * abs(f['loss'].item())
    glp, gids = O.pointer_greedy_generate(f['prev'][:, :1], f['contexts'], f['gen_entity_tokens'], f['state_dict'],
                                          cfg['heads'], cfg['cutoffs'], entity=True)
    assert torch.equal(gids, f['gen_ids'])
    close(glp, f['gen_logp'])
