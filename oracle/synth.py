"""TEST INFRASTRUCTURE -- synthetic GoodNews / NYTimes800k-shaped inputs.

Pure torch-CPU generator code shared by tests, golden-vector generation and bench.py, so that the
oracle, the reference and the CUDA path all see byte-identical inputs.  No reference code here.
"""
import torch

PAD, EOS, BOS = 1, 2, 0


def caption_ids(B, T, vocab, g, ragged=True):
    """[B, T+1] ids: col 0 = <s>(0); body uniform in [3, vocab); 25% of the rows are cut short at a
    random length L in [min(8,T//2), T] with </s> then pad(1)."""
    ids = torch.randint(3, vocab, (B, T + 1), generator=g)
    ids[:, 0] = BOS
    if ragged:
        short = torch.rand(B, generator=g) < 0.25
        lo = max(1, min(8, T // 2))
        for b in range(B):
            if short[b]:
                L = int(torch.randint(lo, T + 1, (1,), generator=g))
                if L < T + 1:
                    ids[b, L] = EOS
                    ids[b, L + 1:] = PAD
    return ids


def contexts(B, g, sources=('image', 'article'), S_img=49, S_art=512, S_faces=4, S_obj=64,
             d_img=2048, d_art=1024, d_faces=512, d_obj=2048, S_ent=32, d_ent=1024):
    """contexts dict in the reference layout: src [S,B,kdim] float, src_mask [B,S] bool True=pad."""
    ctx = {}
    if 'image' in sources:
        ctx['image'] = torch.randn(S_img, B, d_img, generator=g).abs()
        ctx['image_mask'] = torch.zeros(B, S_img, dtype=torch.bool)
    if 'article' in sources:
        a = torch.randn(S_art, B, d_art, generator=g)
        lo = max(1, S_art // 4)
        n = torch.randint(lo, S_art + 1, (B,), generator=g)
        m = torch.arange(S_art)[None, :] >= n[:, None]
        a = a.masked_fill(m.t()[:, :, None], 0.0)
        ctx['article'], ctx['article_mask'] = a, m
    if 'entity' in sources:      # entity-token embeddings; a row may have no entity at all (transformer_entity_pointer.py:354-355)
        e = torch.randn(S_ent, B, d_ent, generator=g)
        n = torch.randint(0, S_ent + 1, (B,), generator=g)
        m = torch.arange(S_ent)[None, :] >= n[:, None]
        ctx['entity'], ctx['entity_mask'] = e.masked_fill(m.t()[:, :, None], 0.0), m
    if 'faces' in sources:
        f = torch.randn(S_faces, B, d_faces, generator=g)
        f = f / f.norm(dim=-1, keepdim=True)
        n = torch.randint(0, S_faces + 1, (B,), generator=g)
        n[:: 8] = 0                                   # rows with no face at all: only the 2 extra keys
        m = torch.arange(S_faces)[None, :] >= n[:, None]
        ctx['faces'], ctx['faces_mask'] = f.masked_fill(m.t()[:, :, None], 0.0), m
    if 'obj' in sources:
        o = torch.randn(S_obj, B, d_obj, generator=g).abs()
        n = torch.randint(0, S_obj + 1, (B,), generator=g)
        m = torch.arange(S_obj)[None, :] >= n[:, None]
        ctx['obj'], ctx['obj_mask'] = o.masked_fill(m.t()[:, :, None], 0.0), m
    return ctx


def make_batch(B, T, vocab, sources=('image', 'article'), seed=1234, **kw):
    g = torch.Generator().manual_seed(seed)
    ids = caption_ids(B, T, vocab, g)
    ctx = contexts(B, g, sources, **kw)
    prev, target = ids[:, :-1].contiguous(), ids[:, 1:].contiguous()
    return prev, target, ctx
