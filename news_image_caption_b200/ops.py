"""Thin Python callers of the fine-grained C-ABI ops (include/tnt_b200.h).  torch tensors only supply the
device memory and the stream; every computation happens in libtnt_b200.so."""
import ctypes as C

import torch

from . import _lib
from ._lib import (EPI_GLU, EPI_NONE, EPI_RELU, EPI_RESIDUAL, MASK_KEY_PADDING, MASK_NONE, MASK_V2, TNT_BF16,
                   TNT_F32)


def _code(t):
    if t.dtype == torch.bfloat16:
        return TNT_BF16
    if t.dtype == torch.float32:
        return TNT_F32
    raise TypeError('tensor dtype must be float32 or bfloat16, got %s' % t.dtype)


def _p(t):
    return C.c_void_p(t.data_ptr()) if t is not None else None


def _st():
    return C.c_void_p(torch.cuda.current_stream().cuda_stream)


def _cuda(t):
    if not t.is_cuda:
        raise TypeError('tnt ops need CUDA tensors (there is no CPU path)')
    return t


def _drop(spec):
    """(p, seed, site[, ld, group_stride]) -> tnt_dropout (None: off)."""
    d = _lib.Dropout()
    if spec is not None:
        d.p, d.seed, d.site = float(spec[0]), int(spec[1]), int(spec[2])
        if len(spec) > 3:
            d.ld = int(spec[3])
        if len(spec) > 4:
            d.group_stride = int(spec[4])
    return d


def dropout_bits(seed, site, index):
    """Host function: the 32 random bits of element `index` of dropout site `site` (kept iff bits >= p * 2^32)."""
    return int(_lib.lib().tnt_dropout_bits(int(seed), int(site), int(index)))


def dropout(x, p, seed, site, residual=None):
    """dropout(x) (+ residual) with the counter-based mask of (seed, site); x [rows, cols], index = row * cols + col."""
    x = _cuda(x).contiguous()
    x2 = x.view(-1, x.shape[-1])
    out = torch.empty_like(x2)
    d = _drop((p, seed, site))
    r = residual.contiguous().view(-1, x.shape[-1]) if residual is not None else None
    _lib.check(_lib.lib().tnt_dropout_apply(x2.data_ptr(), x2.shape[1], r.data_ptr() if r is not None else None, x2.shape[1],
                                            out.data_ptr(), x2.shape[1], x2.shape[0], x2.shape[1], _code(x), C.byref(d), _st()),
               'tnt_dropout_apply')
    return out.view(x.shape)


def linear(a, w, bias=None, epilogue=EPI_NONE, residual=None, out_dtype=None, layout=0, alpha=0.0, m_dyn=None, drop=None):
    """epilogue(a[M,K] @ w[N,K]^T + bias).  For EPI_GLU `w`/`bias` rows must be pair-interleaved.
    layout 1 (NN): a[M,K] @ w[K,N];  layout 2 (TN): a[K,M]^T @ w[K,N] (weight gradient)."""
    a, w = _cuda(a).contiguous(), _cuda(w).contiguous()
    if layout == 0:
        (M, K), N = a.shape, w.shape[0]
    elif layout == 1:
        (M, K), N = a.shape, w.shape[1]
    else:
        (K, M), N = a.shape, w.shape[1]
    out_dtype = out_dtype or a.dtype
    n_out = N // 2 if epilogue == EPI_GLU else N
    c = torch.empty(M, n_out, dtype=out_dtype, device=a.device)
    p = _lib.LinearParams()
    p.a, p.lda, p.w, p.ldw = a.data_ptr(), a.shape[1], w.data_ptr(), w.shape[1]
    p.bias = bias.data_ptr() if bias is not None else None
    p.layout, p.alpha = layout, alpha
    if m_dyn is not None:
        p.m_dyn = m_dyn.data_ptr()
    if residual is not None:
        residual = residual.contiguous()
        p.residual, p.ldr = residual.data_ptr(), residual.shape[1]
    p.c, p.ldc, p.M, p.N, p.K, p.groups = c.data_ptr(), n_out, M, N, K, 1
    p.dtype_in, p.dtype_out, p.epilogue = _code(a), _code(c), epilogue
    p.drop = _drop(drop)
    _lib.check(_lib.lib().tnt_linear_fwd(C.byref(p), _st()), 'tnt_linear_fwd')
    return c


def linear_ln(a, w, bias, residual, gamma, beta, eps=1e-5, groups=1, transpose_tb=None):
    """LayerNorm(a @ w^T + bias + residual) * gamma + beta in one launch (tnt_linear_ln_fwd; bf16).
    groups == 1: a [M, K], w [N, K], residual [M, N] or None, gamma / beta [N] -> [M, N].
    groups > 1 (the per-source out_proj + context_attn_lns): a [M, groups * K], w [groups, N, K], bias / gamma / beta
    [groups, N], residual [M, N] shared by the groups -> [M, groups * N], every N-wide block normalised on its own.
    transpose_tb = (T, B): rows are in t*B + b order, the result is written as [B, T, N]."""
    a, w = _cuda(a).contiguous(), _cuda(w).contiguous()
    M = a.shape[0]
    N, K = w.shape[-2], w.shape[-1]
    out = torch.empty(M, groups * N, dtype=a.dtype, device=a.device) if transpose_tb is None else \
        torch.empty(transpose_tb[1], transpose_tb[0], N, dtype=a.dtype, device=a.device)
    p = _lib.LinearParams()
    p.a, p.lda, p.w, p.ldw = a.data_ptr(), a.shape[1], w.data_ptr(), K
    bias = bias.float().contiguous() if bias is not None else None
    p.bias = bias.data_ptr() if bias is not None else None
    if residual is not None:
        residual = residual.contiguous()
        p.residual, p.ldr = residual.data_ptr(), residual.shape[1]
    p.c, p.ldc, p.M, p.N, p.K, p.groups = out.data_ptr(), groups * N, M, N, K, groups
    p.stride_a, p.stride_w, p.stride_bias, p.stride_res, p.stride_c = (K, N * K, N, 0, N) if groups > 1 else (0, 0, 0, 0, 0)
    p.dtype_in, p.dtype_out = _code(a), _code(out)
    p.epilogue = EPI_RESIDUAL if residual is not None else EPI_NONE
    gamma, beta = gamma.float().contiguous(), beta.float().contiguous()
    L = _lib.lib()
    if not L.tnt_linear_ln_supported(C.byref(p), gamma.data_ptr(), beta.data_ptr()):
        raise RuntimeError('tnt_linear_ln_fwd: unsupported problem M=%d N=%d K=%d groups=%d %s' % (M, N, K, groups, a.dtype))
    tT, tB = transpose_tb if transpose_tb is not None else (0, 0)
    _lib.check(L.tnt_linear_ln_fwd(C.byref(p), gamma.data_ptr(), beta.data_ptr(), N, eps, tT, tB, _st()), 'tnt_linear_ln_fwd')
    return out


def wn_fold(v, g=None, glu_interleave=False, dtype=torch.float32):
    v = _cuda(v).float().contiguous()
    out = torch.empty(v.shape, dtype=dtype, device=v.device)
    gg = g.float().contiguous() if g is not None else None
    _lib.check(_lib.lib().tnt_wn_fold(_p(v), _p(gg), _p(out), v.shape[0], v.shape[1], int(glu_interleave),
                                      _code(out), _st()), 'tnt_wn_fold')
    return out


def glu_interleave_bias(b):
    b = _cuda(b).float().contiguous()
    out = torch.empty_like(b)
    _lib.check(_lib.lib().tnt_glu_interleave_bias(_p(b), _p(out), b.numel(), _st()), 'tnt_glu_interleave_bias')
    return out


def dynconv(x, logits, H, K, hist=None, drop=None):
    """x [T,B,C], logits [T,B,H*K] -> y [T,B,C]; hist [h,B,C] = previous conv inputs (oldest first)."""
    x, logits = _cuda(x).contiguous(), logits.contiguous()
    T, B, Cc = x.shape
    y = torch.empty_like(x)
    p = _lib.DynconvParams()
    p.x, p.ldx, p.logits, p.ldl, p.y, p.ldy = x.data_ptr(), Cc, logits.data_ptr(), logits.shape[-1], y.data_ptr(), Cc
    if hist is not None and hist.shape[0] > 0:
        hist = hist.contiguous()
        p.hist, p.hist_len = hist.data_ptr(), hist.shape[0]
    p.T, p.B, p.C, p.H, p.K, p.dtype = T, B, Cc, H, K, _code(x)
    p.drop = _drop(drop)
    _lib.check(_lib.lib().tnt_dynconv_fwd(C.byref(p), _st()), 'tnt_dynconv_fwd')
    return y


def dynconv_bwd(x, logits, dy, H, K, drop=None):
    x, logits, dy = _cuda(x).contiguous(), logits.contiguous(), dy.contiguous()
    T, B, Cc = x.shape
    dx, dl = torch.empty_like(x), torch.empty_like(logits)
    p = _lib.DynconvBwdParams()
    p.x, p.ldx, p.logits, p.ldl = x.data_ptr(), Cc, logits.data_ptr(), logits.shape[-1]
    p.dy, p.lddy, p.dx, p.lddx, p.dlogits, p.lddl = dy.data_ptr(), Cc, dx.data_ptr(), Cc, dl.data_ptr(), logits.shape[-1]
    p.T, p.B, p.C, p.H, p.K, p.dtype = T, B, Cc, H, K, _code(x)
    p.drop = _drop(drop)
    _lib.check(_lib.lib().tnt_dynconv_bwd(C.byref(p), _st()), 'tnt_dynconv_bwd')
    return dx, dl


def dynconv_step(x, logits, ring, pos, H, K):
    """x [B,C], logits [B,H*K], ring [(K-1),B,C] (updated in place), pos int32[1] device counter."""
    x, logits = _cuda(x).contiguous(), logits.contiguous()
    B, Cc = x.shape
    y = torch.empty_like(x)
    p = _lib.DynconvStepParams()
    p.x, p.ldx, p.logits, p.ldl, p.y, p.ldy = x.data_ptr(), Cc, logits.data_ptr(), logits.shape[-1], y.data_ptr(), Cc
    p.ring, p.pos = ring.data_ptr(), pos.data_ptr()
    p.B, p.C, p.H, p.K, p.dtype = B, Cc, H, K, _code(x)
    _lib.check(_lib.lib().tnt_dynconv_step(C.byref(p), _st()), 'tnt_dynconv_step')
    return y


def layernorm(x, gamma, beta, eps=1e-5, variant=0, transpose_tb=None):
    """LayerNorm over the last dim.  variant 1 = final*/transformer_v2.py LayerNorm (unbiased std, eps outside).
    transpose_tb=(T, B): x rows are in t*B+b order and the output is written as [B, T, C]."""
    x = _cuda(x).contiguous()
    Cc = x.shape[-1]
    y = torch.empty_like(x) if transpose_tb is None else torch.empty(transpose_tb[1], transpose_tb[0], Cc, dtype=x.dtype, device=x.device)
    p = _lib.LayernormParams()
    p.x, p.ldx, p.y, p.ldy = x.data_ptr(), Cc, y.data_ptr(), Cc
    gamma, beta = gamma.float().contiguous(), beta.float().contiguous()
    p.gamma, p.beta = gamma.data_ptr(), beta.data_ptr()
    p.rows, p.C, p.groups, p.variant, p.eps, p.dtype = x.numel() // Cc, Cc, 1, variant, eps, _code(x)
    if transpose_tb is not None:
        p.transpose_T, p.transpose_B = transpose_tb
    _lib.check(_lib.lib().tnt_layernorm_fwd(C.byref(p), _st()), 'tnt_layernorm_fwd')
    return y


def layernorm_bwd(x, dy, gamma, eps=1e-5, variant=0):
    """Backward of `layernorm` (variant 1 = the final* LayerNorm, transformer_v2.py:82-94) -> (dx, dgamma, dbeta)."""
    x, dy = _cuda(x).contiguous(), dy.contiguous()
    Cc = x.shape[-1]
    dx = torch.empty_like(x)
    gamma = gamma.float().contiguous()
    dg = torch.full((Cc,), 7.0, dtype=torch.float32, device=x.device)      # stale contents: the call zeroes them itself
    db = torch.full((Cc,), 7.0, dtype=torch.float32, device=x.device)
    p = _lib.LayernormBwdParams()
    p.zero_param_grads = 1
    p.x, p.ldx, p.dy, p.lddy, p.dx, p.lddx = x.data_ptr(), Cc, dy.data_ptr(), Cc, dx.data_ptr(), Cc
    p.gamma, p.dgamma, p.dbeta = gamma.data_ptr(), dg.data_ptr(), db.data_ptr()
    p.rows, p.C, p.eps, p.dtype, p.variant = x.numel() // Cc, Cc, eps, _code(x), variant
    _lib.check(_lib.lib().tnt_layernorm_bwd(C.byref(p), _st()), 'tnt_layernorm_bwd')
    return dx, dg, db


def xattn(q, k, v, H, scale, pad_mask=None, bias_k=None, bias_v=None, v2_mask=None, need_weights=False, mask_S=0,
          lse=None, drop=None, tile_live=None):
    """q [T,B,E], k/v [S,B,E] (projected) -> out [T,B,E] (+ head-averaged weights [B,T,S+n_extra]).
    pad_mask [B,S] bool True=pad (tell);  v2_mask [B,1|T,S] 'visible' mask (final*), 0 -> -1e9."""
    q, k, v = _cuda(q).contiguous(), k.contiguous(), v.contiguous()
    T, B, E = q.shape
    S = k.shape[0]
    out = torch.empty_like(q)
    p = _lib.XattnParams()
    p.q, p.q_stride_t, p.q_stride_b = q.data_ptr(), B * E, E
    p.k, p.k_stride_s, p.k_stride_b = k.data_ptr(), B * E, E
    p.v, p.v_stride_s, p.v_stride_b = v.data_ptr(), B * E, E
    p.out, p.o_stride_t, p.o_stride_b = out.data_ptr(), B * E, E
    keep = []
    n_extra = 0
    if bias_k is not None:
        bk, bv = bias_k.float().contiguous().view(-1), bias_v.float().contiguous().view(-1)
        keep += [bk, bv]
        p.bias_k, p.bias_v, n_extra = bk.data_ptr(), bv.data_ptr(), 2
    p.mask_mode = MASK_NONE
    if pad_mask is not None:
        m = pad_mask.to(torch.bool).contiguous()
        keep.append(m)
        p.mask, p.m_stride_b, p.m_stride_t, p.mask_mode = m.data_ptr(), m.shape[1], 0, MASK_KEY_PADDING
        p.mask_S = mask_S
    elif v2_mask is not None:
        m = v2_mask.to(torch.bool).expand(B, v2_mask.shape[1], S).contiguous()
        keep.append(m)
        p.mask, p.m_stride_b, p.mask_mode = m.data_ptr(), m.shape[1] * S, MASK_V2
        p.m_stride_t = S if m.shape[1] > 1 else 0
    w = None
    if need_weights:
        w = torch.empty(B, T, S + n_extra, dtype=torch.float32, device=q.device)
        p.weights = w.data_ptr()
    if lse is not None:
        p.lse = lse.data_ptr()
    p.B, p.T, p.S, p.H, p.D, p.n_extra, p.scale, p.dtype = B, T, S, H, E // H, n_extra, scale, _code(q)
    p.drop = _drop(drop)
    if tile_live is not None:
        p.tile_live = tile_live.data_ptr()
    _lib.check(_lib.lib().tnt_xattn_fwd(C.byref(p), _st()), 'tnt_xattn_fwd')
    return out, w


def xattn_tile_mask(pad_mask, S, mask_S=0):
    """pad_mask [B, >= mask_S] bool -> int32 [B]: bit i = keys [128 i, 128 i + 128) of row b hold an un-masked key."""
    m = _cuda(pad_mask).to(torch.bool).contiguous()
    out = torch.empty(m.shape[0], dtype=torch.int32, device=m.device)
    _lib.check(_lib.lib().tnt_xattn_tile_mask(m.data_ptr(), m.shape[1], m.shape[0], S, mask_S, out.data_ptr(), _st()),
               'tnt_xattn_tile_mask')
    return out


def xattn_bwd(q, k, v, out, dout, H, scale, pad_mask=None, v2_mask=None, mask_S=0, lse=None, drop=None, tile_live=None,
              dweights=None):
    """Backward of `xattn` (n_extra = 0: extra rows are ordinary k / v rows).  Returns (dq, dk, dv).  `dweights` [B, T, S]
    fp32 = gradient w.r.t. the head-averaged weights the forward returned (optional)."""
    q, k, v, out, dout = [_cuda(t).contiguous() for t in (q, k, v, out, dout)]
    T, B, E = q.shape
    S = k.shape[0]
    dq, dk, dv = torch.empty_like(q), torch.empty_like(k), torch.empty_like(v)
    bp = _lib.XattnBwdParams()
    p = bp.fwd
    p.q, p.q_stride_t, p.q_stride_b = q.data_ptr(), B * E, E
    p.k, p.k_stride_s, p.k_stride_b = k.data_ptr(), B * E, E
    p.v, p.v_stride_s, p.v_stride_b = v.data_ptr(), B * E, E
    p.out, p.o_stride_t, p.o_stride_b = out.data_ptr(), B * E, E
    keep = []
    p.mask_mode = MASK_NONE
    if pad_mask is not None:
        m = pad_mask.to(torch.bool).contiguous()
        keep.append(m)
        p.mask, p.m_stride_b, p.m_stride_t, p.mask_mode, p.mask_S = m.data_ptr(), m.shape[1], 0, MASK_KEY_PADDING, mask_S
    elif v2_mask is not None:
        m = v2_mask.to(torch.bool).expand(B, v2_mask.shape[1], S).contiguous()
        keep.append(m)
        p.mask, p.m_stride_b, p.mask_mode = m.data_ptr(), m.shape[1] * S, MASK_V2
        p.m_stride_t = S if m.shape[1] > 1 else 0
    p.B, p.T, p.S, p.H, p.D, p.n_extra, p.scale, p.dtype = B, T, S, H, E // H, 0, scale, _code(q)
    p.drop = _drop(drop)
    if tile_live is not None:
        p.tile_live = tile_live.data_ptr()
    if lse is not None:
        p.lse = lse.data_ptr()
    L = _lib.lib()
    scr = torch.empty(L.tnt_xattn_bwd_scratch_bytes(B, T, H, E // H) + 256, dtype=torch.uint8, device=q.device)
    bp.dout, bp.dq, bp.dk, bp.dv = dout.data_ptr(), dq.data_ptr(), dk.data_ptr(), dv.data_ptr()
    bp.scratch, bp.scratch_bytes = scr.data_ptr(), scr.numel()
    if dweights is not None:
        dweights = _cuda(dweights).float().contiguous()
        bp.dweights = dweights.data_ptr()
    _lib.check(L.tnt_xattn_bwd(C.byref(bp), _st()), 'tnt_xattn_bwd')
    return dq, dk, dv


def embed(ids, table, pos_table, pad_idx, dtype):
    """ids [B,T] -> [T,B,d] = table[ids] + pos_table[id == pad_idx ? pad_idx : t + pad_idx + 1]  (tnt_embed_fwd)."""
    ids = _cuda(ids).to(torch.long).contiguous()
    B, T = ids.shape
    d = table.shape[1]
    out = torch.empty(T, B, d, dtype=dtype, device=ids.device)
    p = _lib.EmbedParams()
    p.ids, p.table, p.pos_table, p.out, p.ldo = ids.data_ptr(), table.data_ptr(), pos_table.data_ptr(), out.data_ptr(), d
    p.B, p.T, p.d, p.n_pos, p.pad_idx, p.dtype = B, T, d, pos_table.shape[0], pad_idx, _code(out)
    _lib.check(_lib.lib().tnt_embed_fwd(C.byref(p), _st()), 'tnt_embed_fwd')
    return out


def xattn_strided(q, q_tb, k, k_sb, v, v_sb, B, T, S, H, D, scale, v2_mask=None):
    """Attention over tensors addressed by explicit (time, batch) element strides: q_tb = (stride_t, stride_b) etc.
    Output [T, B, H*D] (t*B+b rows).  v2_mask: bool, broadcastable to [B, T|1, S], 0 -> -1e9 (transformer_v2.py:217)."""
    E = H * D
    out = torch.empty(T, B, E, dtype=q.dtype, device=q.device)
    p = _lib.XattnParams()
    p.q, p.q_stride_t, p.q_stride_b = q.data_ptr(), q_tb[0], q_tb[1]
    p.k, p.k_stride_s, p.k_stride_b = k.data_ptr(), k_sb[0], k_sb[1]
    p.v, p.v_stride_s, p.v_stride_b = v.data_ptr(), v_sb[0], v_sb[1]
    p.out, p.o_stride_t, p.o_stride_b = out.data_ptr(), B * E, E
    p.mask_mode = MASK_NONE
    m = None
    if v2_mask is not None:
        m = v2_mask.to(torch.bool).expand(B, v2_mask.shape[-2], S).contiguous()
        p.mask, p.m_stride_b, p.mask_mode = m.data_ptr(), m.shape[1] * S, MASK_V2
        p.m_stride_t = S if m.shape[1] > 1 else 0
    p.B, p.T, p.S, p.H, p.D, p.n_extra, p.scale, p.dtype = B, T, S, H, D, 0, scale, _code(q)
    _lib.check(_lib.lib().tnt_xattn_fwd(C.byref(p), _st()), 'tnt_xattn_fwd')
    return out


def lse(x, w, bias=None, target=None):
    """Row statistics of z = x @ w^T (+bias) without materialising z:
    returns (max, sumexp, argmax, target_logit) as tensors [M]."""
    x, w = _cuda(x).contiguous(), w.contiguous()
    M, K = x.shape
    N = w.shape[0]
    L = _lib.lib()
    ws = torch.empty(L.tnt_lse_workspace_bytes(M, N, _code(x)) + 256, dtype=torch.uint8, device=x.device)
    stats = torch.empty(M, 4, dtype=torch.float32, device=x.device)
    p = _lib.LseParams()
    p.x, p.ldx, p.w, p.ldw = x.data_ptr(), K, w.data_ptr(), K
    p.bias = bias.data_ptr() if bias is not None else None
    tg = target.to(torch.int32).contiguous() if target is not None else None
    p.target = tg.data_ptr() if tg is not None else None
    p.stats, p.workspace, p.workspace_bytes = stats.data_ptr(), ws.data_ptr(), ws.numel()
    p.M, p.n_cols, p.K, p.dtype = M, N, K, _code(x)
    _lib.check(L.tnt_lse_fwd(C.byref(p), _st()), 'tnt_lse_fwd')
    return stats[:, 0], stats[:, 1], stats[:, 2].contiguous().view(torch.int32), stats[:, 3]


def context_mix(hiddens, weight_logits, out_dtype=None):
    """transformer_faces_objects.py:355-366: sum_l softmax(bert_weight)[l] * hiddens[l], transposed.
    hiddens: list of [B,S,E] CUDA tensors (same dtype) -> [S,B,E]."""
    hs = [_cuda(h).contiguous() for h in hiddens]
    B, S, E = hs[0].shape
    out_dtype = out_dtype or hs[0].dtype
    out = torch.empty(S, B, E, dtype=out_dtype, device=hs[0].device)
    wl = weight_logits.to(hs[0].device, torch.float32).contiguous()
    p = _lib.ContextMixParams()
    for i, h in enumerate(hs):
        if h.shape != hs[0].shape or h.dtype != hs[0].dtype:
            raise ValueError('context_mix: hidden %d differs in shape / dtype' % i)
        p.layers[i] = h.data_ptr()
    p.weight_logits, p.out = wl.data_ptr(), out.data_ptr()
    p.n_layers, p.B, p.S, p.E = len(hs), B, S, E
    p.dtype_in, p.dtype_out = _code(hs[0]), _code(out)
    _lib.check(_lib.lib().tnt_context_mix(C.byref(p), _st()), 'tnt_context_mix')
    return out


def nan_mask(x, out_dtype=None):
    """transformer_faces_objects.py:371-377: ([n,B,d] with NaN rows zeroed, mask [B,n] bool True = padding)."""
    x = _cuda(x).contiguous()
    B, n, d = x.shape
    out = torch.empty(n, B, d, dtype=out_dtype or x.dtype, device=x.device)
    mask = torch.empty(B, n, dtype=torch.bool, device=x.device)
    p = _lib.NanMaskParams()
    p.x, p.out, p.mask, p.B, p.n, p.d = x.data_ptr(), out.data_ptr(), mask.data_ptr(), B, n, d
    p.dtype_in, p.dtype_out = _code(x), _code(out)
    _lib.check(_lib.lib().tnt_nan_mask(C.byref(p), _st()), 'tnt_nan_mask')
    return out, mask


def _v2_mask(p, v2_mask, B, S):
    m = None
    p.mask_mode = MASK_NONE
    if v2_mask is not None:
        m = v2_mask.to(torch.bool).expand(B, v2_mask.shape[-2], S).contiguous()
        p.mask, p.m_stride_b, p.mask_mode = m.data_ptr(), m.shape[1] * S, MASK_V2
        p.m_stride_t = S if m.shape[1] > 1 else 0
    return m


def xattn_bwd_strided(q, q_tb, k, k_sb, v, v_sb, out, dout, B, T, S, H, D, scale, v2_mask=None):
    """Backward of `xattn_strided`: (dq, dk, dv) in the layouts (strides) of q, k, v; out / dout are [T, B, H*D]."""
    E = H * D
    dq, dk, dv = torch.empty_like(q), torch.empty_like(k), torch.empty_like(v)
    out, dout = out.contiguous(), dout.contiguous()
    bp = _lib.XattnBwdParams()
    p = bp.fwd
    p.q, p.q_stride_t, p.q_stride_b = q.data_ptr(), q_tb[0], q_tb[1]
    p.k, p.k_stride_s, p.k_stride_b = k.data_ptr(), k_sb[0], k_sb[1]
    p.v, p.v_stride_s, p.v_stride_b = v.data_ptr(), v_sb[0], v_sb[1]
    p.out, p.o_stride_t, p.o_stride_b = out.data_ptr(), B * E, E
    keep = _v2_mask(p, v2_mask, B, S)
    p.B, p.T, p.S, p.H, p.D, p.n_extra, p.scale, p.dtype = B, T, S, H, D, 0, scale, _code(q)
    L = _lib.lib()
    scr = torch.empty(L.tnt_xattn_bwd_scratch_bytes(B, T, H, D) + 256, dtype=torch.uint8, device=q.device)
    bp.dout, bp.dq, bp.dk, bp.dv = dout.data_ptr(), dq.data_ptr(), dk.data_ptr(), dv.data_ptr()
    bp.scratch, bp.scratch_bytes = scr.data_ptr(), scr.numel()
    _lib.check(L.tnt_xattn_bwd(C.byref(bp), _st()), 'tnt_xattn_bwd')
    del keep
    return dq, dk, dv


def lse_stats(x, w, bias, target):
    """`lse` returning the raw [M, 4] statistics tensor (kept for `ce_bwd`)."""
    x, w = _cuda(x).contiguous(), w.contiguous()
    M, K = x.shape
    N = w.shape[0]
    L = _lib.lib()
    ws = torch.empty(L.tnt_lse_workspace_bytes(M, N, _code(x)) + 256, dtype=torch.uint8, device=x.device)
    stats = torch.empty(M, 4, dtype=torch.float32, device=x.device)
    p = _lib.LseParams()
    p.x, p.ldx, p.w, p.ldw = x.data_ptr(), K, w.data_ptr(), K
    p.bias = bias.data_ptr() if bias is not None else None
    tg = target.to(torch.int32).contiguous()
    p.target = tg.data_ptr()
    p.stats, p.workspace, p.workspace_bytes = stats.data_ptr(), ws.data_ptr(), ws.numel()
    p.M, p.n_cols, p.K, p.dtype = M, N, K, _code(x)
    _lib.check(L.tnt_lse_fwd(C.byref(p), _st()), 'tnt_lse_fwd')
    return stats


def ce_bwd(x, w, bias, target, stats, pad, gscale):
    """Backward of sum-CrossEntropy(x w^T + bias, target, ignore_index=pad) * gscale -> (dx [M,K] dtype, dW [N,K] fp32, db [N] fp32)."""
    x, w = x.contiguous(), w.contiguous()
    M, K = x.shape
    N = w.shape[0]
    L = _lib.lib()
    dx = torch.empty_like(x)
    dW = torch.empty(N, K, dtype=torch.float32, device=x.device)
    db = torch.empty(N, dtype=torch.float32, device=x.device) if bias is not None else None
    tg = target.to(torch.int32).contiguous()
    ws = torch.empty(L.tnt_ce_bwd_workspace_bytes(M, N, _code(x)) + 256, dtype=torch.uint8, device=x.device)
    _lib.check(L.tnt_ce_bwd(_p(x), K, _p(w), K, _p(bias) if bias is not None else None, _p(tg), _p(stats), M, N, K, pad, float(gscale),
                            _p(dx), K, _p(dW), _p(db) if db is not None else None, _code(x), _p(ws), ws.numel(), _st()), 'tnt_ce_bwd')
    return dx, dW, db


def embedding_bwd(ids, dy, scale, vocab):
    """grad [vocab, d] fp32 with grad[ids[r]] += scale * dy[r] (nn.Embedding backward)."""
    ids = _cuda(ids).to(torch.long).contiguous().view(-1)
    dy = dy.contiguous().view(ids.numel(), -1)
    g = torch.zeros(vocab, dy.shape[1], dtype=torch.float32, device=dy.device)
    _lib.check(_lib.lib().tnt_embedding_bwd(_p(ids), ids.numel(), _p(dy), dy.shape[1], float(scale), _p(g), dy.shape[1], dy.shape[1],
                                            vocab, _code(dy), _st()), 'tnt_embedding_bwd')
    return g


def adam_step(params, grads, m, v, lr, step, betas=(0.9, 0.98), eps=1e-9, weight_decay=0.0, grad_scale=1.0):
    """torch.optim.Adam over flat fp32 arenas (tnt_adam_step): the final* scripts' optimizer (custom_v4.py:491)."""
    _lib.check(_lib.lib().tnt_adam_step(_p(params), _p(grads), _p(m), _p(v), params.numel(), None, float(grad_scale), float(lr),
                                        betas[0], betas[1], eps, weight_decay, int(step), _st()), 'tnt_adam_step')


def h2d_live_rows(dst, src_host, mask_host=None):
    """dst [S, B, E] (device) <- src_host [S, B, E] (pinned), only the rows s < len_b of every caption b when a host
    key-padding mask [B, S] (True = padding) is given (tnt_h2d_live_rows); on the current stream.  Returns the bytes copied."""
    S, B, E = src_host.shape
    if not (src_host.is_pinned() and src_host.is_contiguous() and dst.is_contiguous() and dst.shape == src_host.shape):
        raise ValueError('h2d_live_rows: contiguous pinned source and a device tensor of the same shape')
    m = None
    if mask_host is not None:
        m = mask_host.contiguous()
        if m.dtype != torch.bool or tuple(m.shape) != (B, S) or m.is_cuda:
            raise ValueError('h2d_live_rows: mask must be a host bool tensor [B, S]')
    n = _lib.lib().tnt_h2d_live_rows(_p(dst), _p(src_host), _p(m) if m is not None else None, S, B, E, _code(dst), _st())
    if n < 0:
        _lib.check(int(n), 'tnt_h2d_live_rows')
    return int(n)
