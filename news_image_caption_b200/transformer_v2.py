"""Host-side mirror of the fork's own decoder, `final*/transformer_v2.py`, on the
B200 kernels: same class names, constructor meaning, `forward` signatures and `state_dict` keys
(final_roberta2/transformer_v2.py:40-62, 82-140, 186-266, 345-397), so a `v5/latest.pt` checkpoint of
`custom_v3/v4.py` loads unchanged and `run_epoch` / `translate_sentence` call sites keep working
(final_roberta2/custom_v4.py:287-289, 404-419, 647-662).

Every computation is a C-ABI call (include/tnt_b200.h): tcgen05 GEMMs with fused bias / ReLU / residual
epilogues, the attention kernel in its `mask == 0 -> -1e9` mode (transformer_v2.py:217-218), the LayerNorm
kernel in its `variant = 1` mode (unbiased std, eps outside the sqrt, :91-94), the embedding gather and the
log-sum-exp GEMM epilogue for Generator + CrossEntropy (logits [N, V] are never written).  torch supplies memory and
the stream only.  No CPU fallback.

Training (`run_epoch`, custom_v4.py:287-304: `out = model.forward(...)`, `loss, loss_node = loss_compute(out, y, ntokens)`,
`loss_node.backward()`, `optimizer.step()`, `scheduler.step()`): after `model.enable_training()` the same calls build an
autograd graph whose nodes are the CUDA kernels -- tcgen05 GEMMs in NT / NN / TN layouts (_LinFn), the attention kernel and its
backward with the `mask == 0 -> -1e9` rule (_AttnFn), the variant-1 LayerNorm forward / backward (_LNFn), the embedding
gather / scatter (_EmbedFn) and Generator + CrossEntropy through tnt_lse_fwd / tnt_ce_bwd (_CEFn) -- so a stock
`torch.optim.Adam(model.parameters(), lr=1, betas=(0.9, 0.98), eps=1e-9)` under `LambdaLR(rate)` (custom_v4.py:491-498)
works unchanged; `FlatAdam` is the same optimizer as ONE fused kernel over a flat arena (tnt_adam_step).  Dropout is the
identity (p = 0 parity configuration)."""
import math

import torch
import torch.nn as nn

from . import _lib, ops
from ._lib import EPI_NONE, EPI_RELU, EPI_RESIDUAL


def subsequent_mask(size):
    """transformer_v2.py:391-397: True = visible (lower triangle incl. the diagonal)."""
    return torch.triu(torch.ones(1, size, size), diagonal=1).type(torch.uint8) == 0


def positional_table(n, d):
    """PositionalEncoding buffer: interleaved sin / cos (transformer_v2.py PositionalEncoding)."""
    pe = torch.zeros(n, d)
    pos = torch.arange(0, n, dtype=torch.float).unsqueeze(1)
    div = torch.exp(torch.arange(0, d, 2).float() * -(math.log(10000.0) / d))
    pe[:, 0::2] = torch.sin(pos * div)
    pe[:, 1::2] = torch.cos(pos * div)
    return pe


def _param(*shape):
    return nn.Parameter(torch.zeros(*shape, dtype=torch.float32, device='cuda'), requires_grad=False)


def rate(step, model_size, factor, warmup):
    """custom_v4.py:421-430 (the Noam schedule handed to LambdaLR)."""
    if step == 0:
        step = 1
    return factor * (model_size ** (-0.5) * min(step ** (-0.5), step * warmup ** (-1.5)))


# ------------------------------------------------------------------------------------------ autograd nodes (training)
class _LinFn(torch.autograd.Function):
    """y = epilogue(x W^T + b) [+ residual]: tcgen05 GEMM forward (NT), data gradient NN, weight gradient TN (fp32)."""

    @staticmethod
    def forward(ctx, x2d, weight, bias, wc, epilogue, residual):
        y = ops.linear(x2d, wc, bias, epilogue, residual)
        ctx.save_for_backward(x2d, wc, y if epilogue == EPI_RELU else None)
        ctx.meta = (epilogue, residual is not None)
        return y

    @staticmethod
    def backward(ctx, dy):
        from ._lib import LAYOUT_NN, LAYOUT_TN
        x2d, wc, y = ctx.saved_tensors
        epilogue, has_res = ctx.meta
        dy = dy.contiguous()
        dz = dy * (y > 0) if epilogue == EPI_RELU else dy                 # relu'(z) mask (elementwise glue)
        dx = ops.linear(dz, wc, layout=LAYOUT_NN) if ctx.needs_input_grad[0] else None
        dw = ops.linear(dz, x2d, layout=LAYOUT_TN, out_dtype=torch.float32)
        db = dz.float().sum(0)
        return dx, dw, db, None, None, (dy if has_res else None)


class _LNFn(torch.autograd.Function):
    @staticmethod
    def forward(ctx, x, a_2, b_2, eps):
        ctx.save_for_backward(x, a_2)
        ctx.eps = eps
        return ops.layernorm(x, a_2, b_2, eps, variant=1)

    @staticmethod
    def backward(ctx, dy):
        x, a_2 = ctx.saved_tensors
        dx, dg, db = ops.layernorm_bwd(x, dy.contiguous(), a_2, ctx.eps, variant=1)
        return dx, dg, db, None


class _AttnFn(torch.autograd.Function):
    """attention() of transformer_v2.py:213-222 on projected q / k / v addressed by (time, batch) strides."""

    @staticmethod
    def forward(ctx, q, k, v, geom, mask):
        q_tb, k_sb, v_sb, B, T, S, H, D, scale = geom
        out = ops.xattn_strided(q, q_tb, k, k_sb, v, v_sb, B, T, S, H, D, scale, v2_mask=mask)
        ctx.save_for_backward(q, k, v, out)
        ctx.geom, ctx.mask = geom, mask
        return out

    @staticmethod
    def backward(ctx, dout):
        q, k, v, out = ctx.saved_tensors
        q_tb, k_sb, v_sb, B, T, S, H, D, scale = ctx.geom
        dq, dk, dv = ops.xattn_bwd_strided(q, q_tb, k, k_sb, v, v_sb, out, dout, B, T, S, H, D, scale, v2_mask=ctx.mask)
        return dq, dk, dv, None, None


class _EmbedFn(torch.autograd.Function):
    """Embeddings.lut(x) * sqrt(d) + PositionalEncoding: gather forward, scatter-add backward."""

    @staticmethod
    def forward(ctx, lut, ids, table, pe, dtype):
        ctx.save_for_backward(ids)
        ctx.meta = (lut.shape[0], math.sqrt(lut.shape[1]))
        return ops.embed(ids, table, pe, -1, dtype)

    @staticmethod
    def backward(ctx, dy):
        ids, = ctx.saved_tensors
        vocab, scale = ctx.meta
        T, B, d = dy.shape                                           # rows t*B + b  <->  ids[b, t]
        return ops.embedding_bwd(ids.t().contiguous(), dy.contiguous().view(T * B, d), scale, vocab), None, None, None, None


class _CEFn(torch.autograd.Function):
    """Generator.proj + CrossEntropyLoss(ignore_index=pad, reduction='sum') (custom_v4.py:404-419, 489): the forward is the
    log-sum-exp GEMM epilogue, the backward rebuilds dz = softmax - onehot in the same GEMM's epilogue (tnt_ce_bwd)."""

    @staticmethod
    def forward(ctx, x2d, weight, bias, wc, y, pad):
        stats = ops.lse_stats(x2d, wc, bias, y.clamp_min(0))
        ctx.save_for_backward(x2d, wc, bias, y, stats)
        ctx.pad = pad
        nll = (stats[:, 0] + stats[:, 1].log() - stats[:, 3]) * (y != pad)
        return nll.sum()

    @staticmethod
    def backward(ctx, g):
        x2d, wc, bias, y, stats = ctx.saved_tensors
        dx, dW, db = ops.ce_bwd(x2d, wc, bias, y, stats, ctx.pad, float(g))      # one host read of the upstream scalar
        return dx, dW, db, None, None, None


class _Linear(nn.Module):
    """weight [out, in] + bias [out] under the nn.Linear / GehringLinear(weight_norm=False) keys."""

    def __init__(self, d_in, d_out):
        super().__init__()
        self.weight, self.bias = _param(d_out, d_in), _param(d_out)

        self._cast = None

    def w(self, dtype):
        """weight in the compute dtype, cast once per parameter version."""
        if dtype == torch.float32:
            return self.weight
        key = (dtype, self.weight._version, self.weight.data_ptr())
        if self._cast is None or self._cast[0] != key:
            self._cast = (key, self.weight.detach().to(dtype))
        return self._cast[1]

    def run(self, x2d, dtype, epilogue=EPI_NONE, residual=None):
        if torch.is_grad_enabled() and self.weight.requires_grad:
            return _LinFn.apply(x2d, self.weight, self.bias, self.w(dtype), epilogue, residual)
        return ops.linear(x2d, self.w(dtype), self.bias, epilogue, residual)


class LayerNorm(nn.Module):
    """transformer_v2.py:82-94: a_2 * (x - mean) / (std_unbiased + eps) + b_2."""

    def __init__(self, features, eps=1e-6):
        super().__init__()
        self.a_2, self.b_2, self.eps = _param(features), _param(features), eps

    def forward(self, x, transpose_tb=None):
        if torch.is_grad_enabled() and self.a_2.requires_grad:
            y = _LNFn.apply(x, self.a_2, self.b_2, self.eps)
            if transpose_tb is not None:                                   # rows t*B+b -> [B, T, d] (a view, as the reference's)
                T, B = transpose_tb
                y = y.view(T, B, -1).transpose(0, 1)
            return y
        return ops.layernorm(x, self.a_2, self.b_2, self.eps, variant=1, transpose_tb=transpose_tb)


def _attend(q, q_tb, k, k_sb, v, v_sb, B, T, S, H, D, mask):
    scale = 1.0 / math.sqrt(D)
    if torch.is_grad_enabled() and (q.requires_grad or k.requires_grad or v.requires_grad):
        return _AttnFn.apply(q, k, v, (q_tb, k_sb, v_sb, B, T, S, H, D, scale), mask)
    return ops.xattn_strided(q, q_tb, k, k_sb, v, v_sb, B, T, S, H, D, scale, v2_mask=mask)


class MultiHeadedAttention(nn.Module):
    """transformer_v2.py:224-266: linears[0] = query, linears[1] = output, linear_kv[0/1] = key / value
    projections of a memory of width d_key; attention() = softmax(QK^T / sqrt(d_k) + (mask == 0 ? -1e9)) V."""

    def __init__(self, h, d_model, d_key, dropout=0.1):
        super().__init__()
        assert d_model % h == 0
        self.h, self.d_k, self.d_model = h, d_model // h, d_model
        self.linears = nn.ModuleList([_Linear(d_model, d_model), _Linear(d_model, d_model)])
        self.linear_kv = nn.ModuleList([_Linear(d_key, d_model), _Linear(d_key, d_model)])

    def project_kv(self, mem_bf, dtype):
        """mem [B, S, d_key] (batch first, as the reference passes it) -> (k, v) [B*S, d_model]."""
        m2 = mem_bf.to(dtype).contiguous().view(-1, mem_bf.shape[-1])
        return self.linear_kv[0].run(m2, dtype), self.linear_kv[1].run(m2, dtype)

    def attend(self, xn, k, v, B, T, S, mask, residual, dtype):
        """xn [T*B, d] (t*B+b rows, already normed); k, v [B*S, d] (b*S+s rows).  Returns residual + out-proj."""
        E = self.d_model
        q = self.linears[0].run(xn, dtype)
        o = _attend(q, (B * E, E), k, (E, S * E), v, (E, S * E), B, T, S, self.h, self.d_k, mask)
        return self.linears[1].run(o.view(T * B, E), dtype, EPI_RESIDUAL, residual)


class PositionwiseFeedForward(nn.Module):
    def __init__(self, d_model, d_ff, dropout=0.1):
        super().__init__()
        self.w_1, self.w_2 = _Linear(d_model, d_ff), _Linear(d_ff, d_model)


class SublayerConnection(nn.Module):
    """transformer_v2.py:96-109: x + dropout(sublayer(norm(x))) -- only the norm carries parameters."""

    def __init__(self, size, dropout=0.1):
        super().__init__()
        self.norm = LayerNorm(size)


class DecoderLayer(nn.Module):
    """transformer_v2.py:186-211: pre-norm blocks -- causal self-attention; image attention and article attention
    both on the post-self-attention x; cat -> context_fc (GehringLinear(2d, d, weight_norm=False)); FFN."""

    def __init__(self, size, self_attn, img_attn, article_attn, feed_forward, gh_linear, dropout=0.1):
        super().__init__()
        self.size = size
        self.self_attn, self.img_attn, self.article_attn = self_attn, img_attn, article_attn
        self.feed_forward, self.context_fc = feed_forward, gh_linear
        self.sublayer = nn.ModuleList([SublayerConnection(size, dropout) for _ in range(4)])

    def run(self, x, kv, B, T, S_img, S_art, article_mask, tgt_mask, dtype):
        """x [T*B, d] in t*B+b row order; kv = ((k_img, v_img), (k_art, v_art)) projected memories."""
        d = self.size
        n = self.sublayer[0].norm(x)
        k_self, v_self = self.self_attn.linear_kv[0].run(n, dtype), self.self_attn.linear_kv[1].run(n, dtype)
        # self-attention keys are the normed x itself: rows t*B+b -> strides (B*d, d)
        q = self.self_attn.linears[0].run(n, dtype)
        o = _attend(q, (B * d, d), k_self, (B * d, d), v_self, (B * d, d), B, T, T, self.self_attn.h, self.self_attn.d_k, tgt_mask)
        x = self.self_attn.linears[1].run(o.view(T * B, d), dtype, EPI_RESIDUAL, x)
        xi = self.img_attn.attend(self.sublayer[1].norm(x), kv[0][0], kv[0][1], B, T, S_img, None, x, dtype)
        xa = self.article_attn.attend(self.sublayer[2].norm(x), kv[1][0], kv[1][1], B, T, S_art, article_mask, x, dtype)
        x = self.context_fc.run(torch.cat([xi, xa], -1), dtype)
        f = self.feed_forward.w_1.run(self.sublayer[3].norm(x), dtype, EPI_RELU)
        return self.feed_forward.w_2.run(f, dtype, EPI_RESIDUAL, x)


class Decoder(nn.Module):
    """transformer_v2.py:111-140: N layers + final LayerNorm."""

    def __init__(self, layers, size):
        super().__init__()
        self.layers = nn.ModuleList(layers)
        self.norm = LayerNorm(size)


class Embeddings(nn.Module):
    def __init__(self, d_model, vocab):
        super().__init__()
        self.lut = nn.Module()
        self.lut.weight = _param(vocab, d_model)
        self.d_model = d_model


class Generator(nn.Module):
    """transformer_v2.py:54-62: proj = Linear(d_model, vocab); forward = log_softmax(proj(x))."""

    def __init__(self, d_model, vocab):
        super().__init__()
        self.proj = _Linear(d_model, vocab)
        self.dtype = torch.float32

    def stats(self, x2d, target=None):
        """(max, sumexp, argmax, target logit) per row of proj(x) through the LSE GEMM epilogue."""
        return ops.lse(x2d.to(self.dtype).contiguous(), self.proj.w(self.dtype), self.proj.bias, target)

    def forward(self, x):
        shp = x.shape[:-1]
        z = self.proj.run(x.reshape(-1, x.shape[-1]).to(self.dtype).contiguous(), self.dtype).float()
        m, s, _, _ = self.stats(x.reshape(-1, x.shape[-1]))
        return (z - (m + s.log())[:, None]).view(*shp, -1)


class OnlyDecoder(nn.Module):
    """transformer_v2.py:40-51.  `memory` = {'image': [B, S_img, img_dim], 'article': [B, S_art, sent_dim],
    'article_mask': [B, 1, S_art] (nonzero = visible)} as encoder.py:189-204 builds it."""

    def __init__(self, decoder, tgt_embed, generator, h, dtype=torch.bfloat16, max_len=5000):
        super().__init__()
        self.decoder, self.tgt_embed, self.generator = decoder, tgt_embed, generator
        self.h, self.dtype = h, dtype
        self.generator.dtype = dtype
        self._pe = positional_table(max_len, tgt_embed[0].d_model).cuda()
        self._table = None

    def _emb_table(self):
        w = self.tgt_embed[0].lut.weight
        key = (w._version, w.data_ptr())
        if self._table is None or self._table[0] != key:
            self._table = (key, (w.detach() * math.sqrt(w.shape[1])).to(self.dtype))   # Embeddings: lut(x) * sqrt(d)
        return self._table[1]

    def project_memory(self, memory):
        """K / V of both memories for every layer (the reference recomputes them in every call and, in
        translate_sentence, every step; same numbers)."""
        return [((l.img_attn.project_kv(memory['image'], self.dtype)), (l.article_attn.project_kv(memory['article'], self.dtype)))
                for l in self.decoder.layers]

    def forward(self, memory, tgt, src_mask, tgt_mask):
        return self.decode(memory, src_mask, tgt, tgt_mask)

    def enable_training(self, on=True):
        """Parameters become autograd leaves: forward() / SimpleLossCompute then build a graph of CUDA-kernel nodes."""
        for p in self.parameters():
            p.requires_grad_(on)
        return self

    def decode(self, memory, src_mask, tgt, tgt_mask, kv=None):
        _lib.require_device()
        B, T = tgt.shape
        kv = kv if kv is not None else self.project_memory(memory)
        lut = self.tgt_embed[0].lut.weight
        if torch.is_grad_enabled() and lut.requires_grad:
            x = _EmbedFn.apply(lut, tgt.cuda(), self._emb_table(), self._pe, self.dtype).view(T * B, -1)
        else:
            x = ops.embed(tgt.cuda(), self._emb_table(), self._pe, -1, self.dtype).view(T * B, -1)
        am = memory.get('article_mask')
        am = am.cuda() if am is not None else None
        tm = tgt_mask.cuda() if tgt_mask is not None else None
        for layer, lkv in zip(self.decoder.layers, kv):
            x = layer.run(x, lkv, B, T, memory['image'].shape[1], memory['article'].shape[1], am, tm, self.dtype)
        return self.decoder.norm(x, transpose_tb=(T, B))


def make_model_news(tgt_vocab, num_enc_dec=6, dim_model=512, dim_feedfwd=2048, attn_heads=8, dropout=0.1,
                    img_dim=2048, sent_dim=300, dtype=torch.bfloat16):
    """transformer_v2.py:345-382 (zero-initialised: load a reference state_dict)."""
    layers = [DecoderLayer(dim_model, MultiHeadedAttention(attn_heads, dim_model, dim_model),
                           MultiHeadedAttention(attn_heads, dim_model, img_dim),
                           MultiHeadedAttention(attn_heads, dim_model, sent_dim),
                           PositionwiseFeedForward(dim_model, dim_feedfwd), _Linear(2 * dim_model, dim_model), dropout)
              for _ in range(num_enc_dec)]
    tgt_embed = nn.Sequential(Embeddings(dim_model, tgt_vocab))
    return OnlyDecoder(Decoder(layers, dim_model), tgt_embed, Generator(dim_model, tgt_vocab), attn_heads, dtype)


class SimpleLossCompute:
    """custom_v4.py:404-419 with criterion = CrossEntropyLoss(ignore_index=pad, reduction='sum') (:489):
    returns (loss_sum, loss_sum / norm); the logits are reduced inside the GEMM epilogue."""

    def __init__(self, generator, pad):
        self.generator, self.pad = generator, pad

    def __call__(self, x, y, norm):
        y = y.cuda().reshape(-1)
        g = self.generator
        if torch.is_grad_enabled() and g.proj.weight.requires_grad:
            x2d = x.reshape(-1, x.shape[-1]).to(g.dtype).contiguous()
            total = _CEFn.apply(x2d, g.proj.weight, g.proj.bias, g.proj.w(g.dtype), y, self.pad)
            sloss = total / norm
            return sloss.detach() * norm, sloss                              # custom_v4.py:419: (sloss.data * norm, sloss)
        m, s, _, tz = self.generator.stats(x.reshape(-1, x.shape[-1]), y.clamp_min(0))
        nll = (m + s.log() - tz) * (y != self.pad)
        total = nll.sum()
        return total, total / norm


def translate_sentence(model, memory, max_len, bos):
    """custom_v4.py:647-662: greedy, fixed max_len - 1 steps, no EOS stop, the whole prefix is re-decoded each
    step (the K / V projections of the memories are computed once).  Returns ys [B, max_len]."""
    B = memory['image'].shape[0]
    ys = torch.full((B, 1), bos, dtype=torch.long, device='cuda')
    kv = model.project_memory(memory)
    for _ in range(max_len - 1):
        out = model.decode(memory, None, ys, subsequent_mask(ys.shape[1]), kv=kv)
        _, _, arg, _ = model.generator.stats(out[:, -1])
        ys = torch.cat([ys, arg.long().unsqueeze(1)], 1)
    return ys


class FlatAdam:
    """torch.optim.Adam(model.parameters(), lr, betas=(0.9, 0.98), eps=1e-9) (custom_v4.py:491) as ONE fused kernel over a
    flat fp32 arena (tnt_adam_step): parameters and their `.grad` become views of two flat tensors.  `param_groups[0]['lr']`
    is the base lr a scheduler scales (`NoamScheduler` = LambdaLR(rate), custom_v4.py:493-498)."""

    def __init__(self, params, lr=1.0, betas=(0.9, 0.98), eps=1e-9, weight_decay=0.0):
        self.params = [p for p in params if p.requires_grad]
        n = sum((p.numel() + 3) // 4 * 4 for p in self.params)
        dev = self.params[0].device
        self.flat_p = torch.zeros(n, dtype=torch.float32, device=dev)
        self.flat_g = torch.zeros(n, dtype=torch.float32, device=dev)
        self.m, self.v = torch.zeros_like(self.flat_p), torch.zeros_like(self.flat_p)
        self.slices, off = [], 0
        with torch.no_grad():
            for p in self.params:
                k = p.numel()
                view = self.flat_p[off:off + k].view(p.shape)
                view.copy_(p.data)
                p.data = view
                self.slices.append((off, k))
                off += (k + 3) // 4 * 4
        self.param_groups = [dict(lr=lr, initial_lr=lr, betas=betas, eps=eps, weight_decay=weight_decay)]
        self.t = 0
        self._attach()

    def _attach(self):
        for p, (off, k) in zip(self.params, self.slices):
            p.grad = self.flat_g[off:off + k].view(p.shape)

    def zero_grad(self, set_to_none=True):
        self.flat_g.zero_()
        self._attach()                      # the arena views stay attached whatever `set_to_none` says

    def step(self):
        for p, (off, k) in zip(self.params, self.slices):       # a gradient autograd attached as a fresh tensor
            if p.grad is not None and p.grad.data_ptr() != self.flat_g.data_ptr() + 4 * off:
                self.flat_g[off:off + k].copy_(p.grad.reshape(-1))
        g = self.param_groups[0]
        self.t += 1
        ops.adam_step(self.flat_p, self.flat_g, self.m, self.v, g['lr'], self.t, g['betas'], g['eps'], g['weight_decay'])
        _bump_versions(self.params)


def _bump_versions(params):
    """The fused kernel wrote the parameters through raw pointers: bump their autograd version counters so per-parameter
    caches (`_Linear.w`, the scaled embedding table) notice."""
    with torch.no_grad():
        torch._foreach_add_(list(params), 0)


class NoamScheduler:
    """LambdaLR(optimizer, lambda step: rate(step, d_model, factor, warmup)) (custom_v4.py:493-498)."""

    def __init__(self, optimizer, model_size, factor=1.0, warmup=30000):
        self.opt, self.args, self.n = optimizer, (model_size, factor, warmup), 0
        self._apply()

    def _apply(self):
        for g in self.opt.param_groups:
            g['lr'] = g.get('initial_lr', 1.0) * rate(self.n, *self.args)

    def step(self):
        self.n += 1
        self._apply()
