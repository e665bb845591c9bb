"""Module-level drop-ins with the reference's constructor / forward signatures and state_dict keys
(SURVEY 8b "Python surface"), each a thin `torch.autograd.Function` over the C-ABI kernels:

  GehringLinear       ttl/tell/modules/linear.py:7-33
  DynamicConv1dTBC    ttl/tell/modules/convolutions/dynamic.py:25-361 (+ incremental buffer, utils/state.py:10-34)
  MultiHeadAttention  ttl/tell/modules/attention/multi_head.py:207-552
  AdaptiveSoftmax     ttl/tell/modules/softmax.py:43-222

They exist for callers that build their own layer stacks out of these modules (the reference's pointer /
entity decoders do); the fused `DynamicConvDecoder` engine is the fast path for the stock decoders.  Forward and
backward both run on the CUDA kernels; torch autograd only chains them.  Dropout arguments are accepted and must
be inactive (eval mode or p = 0): the kernels implement the deterministic path."""
import math

import torch
import torch.nn as nn
import torch.nn.functional as F

from . import ops
from ._lib import LAYOUT_NN, LAYOUT_TN


def _c(t):
    return t.contiguous()


class LinearFn(torch.autograd.Function):
    """y = x W^T (+ b): tcgen05 GEMM forward, NN / TN layouts backward (no transposed weight copies)."""

    @staticmethod
    def forward(ctx, x, w, b):
        x2 = _c(x.reshape(-1, x.shape[-1]))
        wc = _c(w.to(x2.dtype))
        ctx.save_for_backward(x2, wc)
        ctx.meta = (x.shape, w.dtype, b is not None)
        y = ops.linear(x2, wc, b.float() if b is not None else None)
        return y.view(*x.shape[:-1], w.shape[0])

    @staticmethod
    def backward(ctx, dy):
        x2, wc = ctx.saved_tensors
        shape, wdtype, has_b = ctx.meta
        dy2 = _c(dy.reshape(-1, dy.shape[-1]).to(x2.dtype))
        dx = ops.linear(dy2, wc, layout=LAYOUT_NN).view(shape) if ctx.needs_input_grad[0] else None
        dw = ops.linear(dy2, x2, layout=LAYOUT_TN, out_dtype=torch.float32).to(wdtype) if ctx.needs_input_grad[1] else None
        db = dy2.float().sum(0) if has_b and ctx.needs_input_grad[2] else None
        return dx, dw, db


def linear(x, w, b=None):
    return LinearFn.apply(x, w, b)


class GehringLinear(nn.Module):
    """linear.py:7-33: N(0, sqrt((1-p)/in)) init, weight_norm reparametrisation (weight_g [out,1], weight_v)."""

    def __init__(self, in_features, out_features, dropout=0, bias=True, weight_norm=True):
        super().__init__()
        self.in_features, self.out_features, self.weight_norm = in_features, out_features, weight_norm
        v = torch.empty(out_features, in_features).normal_(0, math.sqrt((1 - dropout) / in_features))
        if weight_norm:
            self.weight_g = nn.Parameter(v.norm(dim=1, keepdim=True))
            self.weight_v = nn.Parameter(v)
        else:
            self.weight = nn.Parameter(v)
        self.bias = nn.Parameter(torch.zeros(out_features)) if bias else None

    def folded_weight(self):
        if not self.weight_norm:
            return self.weight
        return self.weight_v * (self.weight_g / self.weight_v.norm(dim=1, keepdim=True))

    def forward(self, x):
        return linear(x, self.folded_weight(), self.bias)


def _draw_seed():
    """A 63-bit seed for one dropout call, drawn from torch's CPU generator (so torch.manual_seed controls it)."""
    return int(torch.empty((), dtype=torch.int64).random_().item())


class DynConvFn(torch.autograd.Function):
    @staticmethod
    def forward(ctx, x, logits, H, K, hist, drop=None):
        x, logits = _c(x), _c(logits)
        ctx.save_for_backward(x, logits)
        ctx.meta = (H, K, hist is not None, drop)
        return ops.dynconv(x, logits, H, K, hist=hist, drop=drop)

    @staticmethod
    def backward(ctx, dy):
        x, logits = ctx.saved_tensors
        H, K, had_hist, drop = ctx.meta
        if had_hist:
            raise RuntimeError('DynamicConv1dTBC: backward through the incremental path is not supported')
        dx, dl = ops.dynconv_bwd(x, logits, _c(dy), H, K, drop=drop)      # the mask is recomputed from (p, seed, site)
        return dx, dl, None, None, None, None


_INSTANCE_ID = [0]


class DynamicConv1dTBC(nn.Module):
    """dynamic.py:25-361.  Supported configuration = what every decoder uses (decoder_flattened.py:226-236):
    weight_softmax=True, padding_l=kernel_size-1 (causal), renorm_padding=False, conv_bias=False, in_proj=False."""

    def __init__(self, input_size, kernel_size=1, padding_l=None, num_heads=1, weight_dropout=0., weight_softmax=False,
                 renorm_padding=False, bias=False, conv_bias=False, query_size=None, in_proj=False):
        super().__init__()
        if not weight_softmax or renorm_padding or conv_bias or in_proj:
            raise NotImplementedError('DynamicConv1dTBC: only weight_softmax=True, renorm_padding=False, '
                                      'conv_bias=False, in_proj=False is implemented')
        if padding_l is not None and padding_l != kernel_size - 1:
            raise NotImplementedError('DynamicConv1dTBC: causal padding_l = kernel_size - 1 only')
        self.input_size, self.kernel_size, self.num_heads = input_size, kernel_size, num_heads
        self.query_size = input_size if query_size is None else query_size
        self.weight_dropout = weight_dropout
        self.weight_linear = nn.Linear(self.query_size, num_heads * kernel_size, bias)
        _INSTANCE_ID[0] += 1
        self._id = _INSTANCE_ID[0]

    def _key(self):       # utils/state.py:17-26 key scheme
        return '%s.%d.%s' % (type(self).__name__, self._id, 'input_buffer')

    def forward(self, X, incremental_state=None, query=None, unfold=None):
        """X [T, B, C] -> [T, B, C].  `unfold` is accepted and ignored: the kernel is one algorithm for every T
        (the reference switches to the unfolded path only for T > 512, dynamic.py:102; same numbers)."""
        drop = None
        if self.training and self.weight_dropout > 0:       # dynamic.py:305: DropConnect on the soft-maxed taps
            if incremental_state is not None:
                raise NotImplementedError('DynamicConv1dTBC: weight_dropout on the incremental path')
            drop = (self.weight_dropout, _draw_seed(), self._id)
        q = X if query is None else query
        logits = linear(q, self.weight_linear.weight, self.weight_linear.bias)
        hist = None
        if incremental_state is not None:
            hist = incremental_state.get(self._key())
            window = X if hist is None else torch.cat([hist, X], 0)
            keep = self.kernel_size - 1
            incremental_state[self._key()] = window[-keep:].detach() if keep > 0 else window[:0]
        return DynConvFn.apply(X, logits, self.num_heads, self.kernel_size, hist, drop)

    def reorder_incremental_state(self, incremental_state, new_order):
        buf = incremental_state.get(self._key())
        if buf is not None:
            incremental_state[self._key()] = buf.index_select(1, new_order)


class LightweightConv1dTBC(nn.Module):
    """lightweight.py:86-253 (`decoder_conv_type: lightweight`, decoder_flattened.py:226-230): the same causal depthwise
    convolution as DynamicConv1dTBC with STATIC taps -- one learnable [num_heads, 1, K] weight, soft-maxed over K, shared by
    every timestep and caption.  It is the DynamicConv kernel fed with the weight broadcast as its tap logits (forward and
    backward: autograd's expand sums the per-position tap gradients into the weight).  Supported configuration = what the
    decoder builds: weight_softmax=True, padding_l = kernel_size - 1."""

    def __init__(self, input_size, kernel_size=1, padding_l=None, num_heads=1, weight_dropout=0., weight_softmax=False,
                 bias=False):
        super().__init__()
        if not weight_softmax:
            raise NotImplementedError('LightweightConv1dTBC: only weight_softmax=True is implemented')
        if padding_l is not None and padding_l != kernel_size - 1:
            raise NotImplementedError('LightweightConv1dTBC: causal padding_l = kernel_size - 1 only')
        self.input_size, self.kernel_size, self.num_heads = input_size, kernel_size, num_heads
        self.weight_dropout = weight_dropout
        self.weight = nn.Parameter(torch.empty(num_heads, 1, kernel_size))
        nn.init.xavier_uniform_(self.weight)
        self.bias = nn.Parameter(torch.zeros(input_size)) if bias else None
        _INSTANCE_ID[0] += 1
        self._id = _INSTANCE_ID[0]

    def _key(self):
        return '%s.%d.%s' % (type(self).__name__, self._id, 'input_buffer')

    def forward(self, x, incremental_state=None, unfold=False):
        """x [T, B, C] -> [T, B, C]; `unfold` accepted and ignored (one algorithm for both reference paths)."""
        T, B, _ = x.shape
        drop = None
        if self.training and self.weight_dropout > 0:
            if incremental_state is not None:
                raise NotImplementedError('LightweightConv1dTBC: weight_dropout on the incremental path')
            drop = (self.weight_dropout, _draw_seed(), self._id)
        logits = self.weight.view(1, 1, -1).expand(T, B, self.num_heads * self.kernel_size).to(x.dtype)
        hist = None
        if incremental_state is not None:
            hist = incremental_state.get(self._key())
            window = x if hist is None else torch.cat([hist, x], 0)
            keep = self.kernel_size - 1
            incremental_state[self._key()] = window[-keep:].detach() if keep > 0 else window[:0]
        y = DynConvFn.apply(x, logits, self.num_heads, self.kernel_size, hist, drop)
        return y if self.bias is None else y + self.bias.to(y.dtype).view(1, 1, -1)

    def reorder_incremental_state(self, incremental_state, new_order):
        buf = incremental_state.get(self._key())
        if buf is not None:
            incremental_state[self._key()] = buf.index_select(1, new_order)


class XattnFn(torch.autograd.Function):
    @staticmethod
    def forward(ctx, q, k, v, H, scale, pad_mask, mask_S, need_weights, drop=None):
        q, k, v = _c(q), _c(k), _c(v)
        T, B, E = q.shape
        lse = torch.empty(B, H, T, dtype=torch.float32, device=q.device)
        out, w = ops.xattn(q, k, v, H, scale, pad_mask=pad_mask, mask_S=mask_S, need_weights=need_weights, lse=lse, drop=drop)
        ctx.save_for_backward(q, k, v, out, lse)
        ctx.meta = (H, scale, pad_mask, mask_S, drop)
        ctx.mark_non_differentiable(w) if w is not None else None
        return out, w

    @staticmethod
    def backward(ctx, dout, _dw):
        q, k, v, out, lse = ctx.saved_tensors
        H, scale, pad_mask, mask_S, drop = ctx.meta
        dq, dk, dv = ops.xattn_bwd(q, k, v, out, _c(dout), H, scale, pad_mask=pad_mask, mask_S=mask_S, lse=lse, drop=drop)
        return dq, dk, dv, None, None, None, None, None, None


class MultiHeadAttention(nn.Module):
    """multi_head.py:207-552 (the manual path the decoders take: static_kv=True, key = value = the memory).
    Defaults add_bias_kv=True, add_zero_attn=True as in the reference (:212-214): two extra keys after the memory
    rows -- the learned bias_k / bias_v row and an all-zero row -- which the key-padding mask never covers."""

    def __init__(self, embed_dim, num_heads, kdim=None, vdim=None, dropout=0., bias=True, add_bias_kv=True,
                 add_zero_attn=True, self_attention=False, encoder_decoder_attention=False):
        super().__init__()
        self.embed_dim, self.num_heads, self.dropout = embed_dim, num_heads, dropout
        self.kdim = embed_dim if kdim is None else kdim
        self.vdim = embed_dim if vdim is None else vdim
        self.qkv_same_dim = self.kdim == embed_dim and self.vdim == embed_dim
        self.head_dim = embed_dim // num_heads
        assert self.head_dim * num_heads == embed_dim and add_bias_kv and add_zero_attn and bias
        self.scaling = self.head_dim ** -0.5
        if self.qkv_same_dim:
            self.in_proj_weight = nn.Parameter(torch.empty(3 * embed_dim, embed_dim))
            nn.init.xavier_uniform_(self.in_proj_weight)
        else:
            self.k_proj_weight = nn.Parameter(torch.empty(embed_dim, self.kdim))
            self.v_proj_weight = nn.Parameter(torch.empty(embed_dim, self.vdim))
            self.q_proj_weight = nn.Parameter(torch.empty(embed_dim, embed_dim))
            for w in (self.k_proj_weight, self.v_proj_weight, self.q_proj_weight):
                nn.init.xavier_uniform_(w)
        self.in_proj_bias = nn.Parameter(torch.zeros(3 * embed_dim))
        self.out_proj = nn.Linear(embed_dim, embed_dim, bias=bias)
        nn.init.xavier_uniform_(self.out_proj.weight)
        nn.init.constant_(self.out_proj.bias, 0.)
        self.bias_k = nn.Parameter(torch.empty(1, 1, embed_dim))
        self.bias_v = nn.Parameter(torch.empty(1, 1, embed_dim))
        nn.init.xavier_normal_(self.bias_k)
        nn.init.xavier_normal_(self.bias_v)

    def _weights(self):
        E = self.embed_dim
        if self.qkv_same_dim:
            w = self.in_proj_weight
            return w[:E], w[E:2 * E], w[2 * E:]
        return self.q_proj_weight, self.k_proj_weight, self.v_proj_weight

    def forward(self, query, key, value, key_padding_mask=None, incremental_state=None, need_weights=True,
                static_kv=False, attn_mask=None):
        """query [T,B,E], key/value [S,B,kdim] -> (attn [T,B,E], head-averaged weights [B,T,S+2] | None)."""
        if attn_mask is not None:
            raise NotImplementedError('MultiHeadAttention: attn_mask is not used by the decoders and not implemented')
        drop = None
        if self.training and self.dropout > 0:              # multi_head.py:463: dropout on the probabilities
            drop = (self.dropout, _draw_seed(), 0)      # the returned weights are the dropped ones, as in the reference
        T, B, E = query.shape
        S = key.shape[0]
        wq, wk, wv = self._weights()
        b = self.in_proj_bias
        q = linear(query, wq, b[:E])
        k = linear(key, wk, b[E:2 * E])
        v = linear(value, wv, b[2 * E:])
        zeros = k.new_zeros(1, B, E)
        k = torch.cat([k, self.bias_k.to(k.dtype).expand(1, B, E), zeros], 0)      # :355-374, 416-421
        v = torch.cat([v, self.bias_v.to(v.dtype).expand(1, B, E), zeros], 0)
        out, w = XattnFn.apply(q, k, v, self.num_heads, self.scaling, key_padding_mask, S if key_padding_mask is not None else 0,
                               need_weights, drop)
        return linear(out, self.out_proj.weight, self.out_proj.bias), w

    def reorder_incremental_state(self, incremental_state, new_order):
        return          # static K/V: nothing is cached per beam


def multi_head_attention_score_forward(query, key, embed_dim_to_check, num_heads, in_proj_weight, in_proj_bias, bias_k,
                                       add_zero_attn, dropout_p, out_proj_weight=None, out_proj_bias=None, training=True,
                                       key_padding_mask=None, attn_mask=None, use_separate_proj_weight=False,
                                       q_proj_weight=None, k_proj_weight=None, static_k=None):
    """multi_head.py:14-204 -- the copy-attention score head the pointer decoders import: head-averaged attention
    probabilities [B,T,S(+1 bias_k row)(+1 zero row)] of query [T,B,E] over key [S,B,E]; no values, no output
    projection (out_proj_* are accepted and unused, as in the reference).  Same argument order and meaning.
    Forward only: the weights come from the attention kernel's head-averaging epilogue (its value operand is the
    key tile, the attended output is discarded).  attn_mask / static_k are not used by any caller and raise."""
    if attn_mask is not None or static_k is not None:
        raise NotImplementedError('multi_head_attention_score_forward: attn_mask / static_k are not implemented')
    T, B, E = query.shape
    assert E == embed_dim_to_check                                              # :81
    assert (E // num_heads) * num_heads == E, 'embed_dim must be divisible by num_heads'      # :85
    if use_separate_proj_weight:
        wq, wk = q_proj_weight, k_proj_weight
        assert tuple(wq.shape) == (E, query.size(-1)) and tuple(wk.shape) == (E, key.size(-1))   # :113-118
    else:
        wq, wk = in_proj_weight[:E], in_proj_weight[E:2 * E]                      # :96-110 (q | k halves)
    bq = bk = None
    if in_proj_bias is not None:
        bq, bk = in_proj_bias[:E], in_proj_bias[E:2 * E]
    S = key.shape[0]
    if key_padding_mask is not None:
        assert key_padding_mask.size(0) == B and key_padding_mask.size(1) == S    # :161-163
    with torch.no_grad():
        q = linear(query, wq, bq)
        k = linear(key, wk, bk)
        rows = [k]
        if bias_k is not None:
            rows.append(bias_k.to(k.dtype).expand(1, B, E))                     # :130-135
        if add_zero_attn:
            rows.append(k.new_zeros(1, B, E))                                   # :165-168
        if len(rows) > 1:
            k = torch.cat(rows, 0)
        drop = (dropout_p, _draw_seed(), 0) if training and dropout_p > 0 else None      # :199
        _, w = ops.xattn(q, k, k, num_heads, (E // num_heads) ** -0.5, pad_mask=key_padding_mask,
                         mask_S=S if key_padding_mask is not None else 0, need_weights=True, drop=drop)
    return w


class AdaptiveSoftmax(nn.Module):
    """softmax.py:43-222 (Grave et al.): head over cutoff[0] + n_tails classes, tail i = Linear(d, d / factor^(i+1),
    no bias) -> dropout -> Linear(., band size, no bias).  forward(X, target) -> (logit list, target list) exactly as
    the reference (rows of tail i = the rows whose target falls in band i, ascending); get_log_prob -> [B,T,V]."""

    def __init__(self, vocab_size, input_dim, cutoff, dropout=0., factor=4., adaptive_inputs=None, tie_proj=False):
        super().__init__()
        cutoff = list(cutoff)
        if vocab_size > cutoff[-1]:
            cutoff = cutoff + [vocab_size]
        assert vocab_size == cutoff[-1]
        self.vocab_size, self.cutoff, self.dropout, self.input_dim, self.factor = vocab_size, cutoff, dropout, input_dim, factor
        nt = len(cutoff) - 1
        self.head = nn.Module()
        if adaptive_inputs is not None:          # tied: softmax.py:83-86, 111-132
            self.head.word_proj = nn.Module()
            self.head.word_proj.weight = adaptive_inputs.embeddings[0][0].weight
        else:
            self.head.word_proj = nn.Linear(input_dim, cutoff[0], bias=False)
        self.head.class_proj = nn.Linear(input_dim, nt, bias=False)
        self.tail = nn.ModuleList()
        for i in range(nt):
            dim = int(input_dim // factor ** (i + 1))
            proj = nn.Linear(input_dim, dim, bias=False)
            if adaptive_inputs is not None:
                out = nn.Module()
                out.weight = adaptive_inputs.embeddings[i + 1][0].weight
            else:
                out = nn.Linear(dim, cutoff[i + 1] - cutoff[i], bias=False)
            self.tail.append(nn.ModuleList([proj, nn.Dropout(dropout), out]))

    def adapt_target(self, target):
        """softmax.py:144-167."""
        target = target.view(-1)
        new_target = [target.clone()]
        target_idxs = []
        for i in range(len(self.cutoff) - 1):
            mask = target.ge(self.cutoff[i]) & target.lt(self.cutoff[i + 1])
            new_target[0][mask] = self.cutoff[0] + i
            if mask.any():
                target_idxs.append(mask.nonzero(as_tuple=False).squeeze(1))
                new_target.append(target[mask] - self.cutoff[i])
            else:
                target_idxs.append(None)
                new_target.append(None)
        return new_target, target_idxs

    def _head_weight(self):
        return torch.cat([self.head.word_proj.weight, self.head.class_proj.weight], 0)

    def forward(self, X, target):
        X = X.reshape(-1, X.shape[-1])
        new_target, target_idxs = self.adapt_target(target)
        output = [linear(X, self._head_weight())]
        for i, idxs in enumerate(target_idxs):
            if idxs is None:
                output.append(None)
            else:
                h = linear(X.index_select(0, idxs), self.tail[i][0].weight)
                output.append(linear(h, self.tail[i][2].weight))
        return output, new_target

    def get_log_prob(self, X, target=None):
        """softmax.py:193-222."""
        B, T, d = X.shape
        X2 = X.reshape(-1, d)
        head = F.log_softmax(linear(X2, self._head_weight()).float(), -1)
        parts = [head[:, :self.cutoff[0]]]
        for i in range(len(self.cutoff) - 1):
            z = linear(linear(X2, self.tail[i][0].weight), self.tail[i][2].weight).float()
            parts.append(F.log_softmax(z, -1) + head[:, self.cutoff[0] + i, None])
        return torch.cat(parts, 1).view(B, T, self.vocab_size)
