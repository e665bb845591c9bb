"""TEST INFRASTRUCTURE ONLY -- CPU restatement (the "oracle") of the caption-decoder hot path.

Plain functional fp32 torch-on-CPU code over a flat ``state_dict`` (no nn.Module, no autograd
tricks), one function per reference symbol, each citing the reference file:line it follows
(paths relative to /root/reference).  `torch` is the arithmetic provider exactly as it is for the
reference itself (every op on this path is in-tree Python over ATen).

PINNING: the reference ships no tests / golden vectors for this path, so the
oracle is pinned against the reference ITSELF: `tests/test_oracle_vs_reference.py` imports the
unmodified reference modules in the build container (oracle/ref_import.py) and compares; and
`oracle/make_golden.py` froze reference outputs into `tests/golden/*.pt`, which
`tests/test_oracle_golden.py` checks on any machine.

Only `tests/`, `__graft_entry__.smoke()` and `bench.py`'s cpu_baseline / `--impl reference` leg may
import this module.  The product package never does.
"""
import math
from typing import Dict, List, Optional, Sequence

import numpy as np
import torch
import torch.nn.functional as F

SD = Dict[str, torch.Tensor]


# ----------------------------------------------------------------------------- training dropout
# The reference draws its dropout masks from torch's RNG (F.dropout).  The CUDA path uses a counter-based generator
# (include/tnt_b200.h, tnt_dropout_bits) so that the backward can recompute the masks; the functions below restate
# that generator in numpy so a test can feed the SAME masks to this oracle (and, in the build container, to the
# unmodified reference through a patched F.dropout: tests/test_oracle_vs_reference.py).
DROP_INPUT, DROP_TAPS, DROP_CONV_OUT, DROP_ATTN_OUT, DROP_FFN_OUT, DROP_RELU, DROP_ATTN_PROB = range(7)
DROP_SITE_EMBED = 0xFFFF


def dropout_bits(seed: int, site: int, index) -> np.ndarray:
    """32 random bits per element index (uint32 arithmetic, wrap-around)."""
    M = np.uint64(0xFFFFFFFF)
    idx = np.asarray(index, dtype=np.uint64)
    k0 = ((seed & 0xFFFFFFFF) * 0x9E3779B1 + site * 0x7FEB352D + 0x165667B1) & 0xFFFFFFFF
    k1 = ((seed >> 32) & 0xFFFFFFFF) ^ ((site * 0x846CA68B + 0x27D4EB2F) & 0xFFFFFFFF)
    h = np.uint64(k0) ^ (idx & M)
    h = (h * np.uint64(0x9E3779B1)) & M
    h ^= h >> np.uint64(16)
    h = (h + np.uint64(k1) + (((idx >> np.uint64(32)) * np.uint64(0x85EBCA6B)) & M)) & M
    h ^= h >> np.uint64(15)
    h = (h * np.uint64(0x2C1B3C6D)) & M
    h ^= h >> np.uint64(12)
    h = (h * np.uint64(0x297A2D39)) & M
    h ^= h >> np.uint64(15)
    return h.astype(np.uint32)


def dropout_mask(p: float, seed: int, site: int, index: torch.Tensor) -> torch.Tensor:
    """F.dropout's multiplier (0 or 1/(1-p), fp32) for the elements whose counter index is `index`."""
    p32 = np.float32(p)
    t = float(p32) * 4294967296.0
    thr = 0xFFFFFFFF if t >= 4294967295.0 else max(int(t), 1)
    keep = dropout_bits(seed, site, index.numpy()) >= np.uint32(thr)
    scale = np.float32(1.0) / (np.float32(1.0) - p32)
    return torch.from_numpy(keep.astype(np.float32) * scale)


class DropoutMasks:
    """The decoder's dropout configuration (decoder_flattened.py:24-31) with counter-based masks."""

    def __init__(self, dropout=0.0, input_dropout=0.0, weight_dropout=0.0, attention_dropout=0.0, relu_dropout=0.0, seed=0):
        self.dropout, self.input_dropout, self.weight_dropout = dropout, input_dropout, weight_dropout
        self.attention_dropout, self.relu_dropout, self.seed = attention_dropout, relu_dropout, seed

    def __call__(self, x: torch.Tensor, p: float, site: int, index: Optional[torch.Tensor] = None) -> torch.Tensor:
        if p <= 0.0:
            return x
        if index is None:
            index = torch.arange(x.numel()).view(x.shape)
        return x * dropout_mask(p, self.seed, site, index).to(x.dtype)


# ----------------------------------------------------------------------------- linear pieces
def fold_weight_norm(g: torch.Tensor, v: torch.Tensor) -> torch.Tensor:
    """w = g * v / ||v||_row  -- ttl/tell/modules/linear.py:30-33 (nn.utils.weight_norm, dim=0)."""
    return v * (g / v.norm(dim=1, keepdim=True))


def gehring_linear(x: torch.Tensor, sd: SD, p: str) -> torch.Tensor:
    """GehringLinear forward -- ttl/tell/modules/linear.py:7-33."""
    if p + 'weight_g' in sd:
        w = fold_weight_norm(sd[p + 'weight_g'], sd[p + 'weight_v'])
    else:
        w = sd[p + 'weight']
    return F.linear(x, w, sd.get(p + 'bias'))


def layer_norm(x: torch.Tensor, sd: SD, p: str, eps: float = 1e-5) -> torch.Tensor:
    """nn.LayerNorm(embed_dim) -- ttl/tell/models/decoder_flattened.py:245,254,259,268."""
    mu = x.mean(-1, keepdim=True)
    var = ((x - mu) ** 2).mean(-1, keepdim=True)
    return (x - mu) / torch.sqrt(var + eps) * sd[p + 'weight'] + sd[p + 'bias']


# ----------------------------------------------------------------------------- DynamicConv
def dynconv_taps(x: torch.Tensor, w_tap: torch.Tensor, H: int, K: int) -> torch.Tensor:
    """softmax over ALL K tap logits per (t, b, head) -- dynamic.py:300-304.  [T,B,H,K]"""
    T, B, C = x.shape
    return F.softmax(F.linear(x, w_tap).view(T, B, H, K), dim=-1)


def dynamic_conv(x: torch.Tensor, w_tap: torch.Tensor, H: int, K: int, tap_drop=None) -> torch.Tensor:
    """Causal DynamicConv1dTBC (padding_l = K-1, weight_softmax, no renorm, no bias).

    y[t,b,h*R+r] = sum_j w[t,b,h,j] * x[t-(K-1)+j, b, h*R+r], x[<0] = 0; the taps that reach
    before t=0 are DROPPED AFTER the softmax (not renormalised) -- ttl/tell/modules/convolutions/
    dynamic.py:285-336 (band matrix + narrow(2, P, T)); same result as the unfold path :122-190.
    """
    T, B, C = x.shape
    R = C // H
    w = dynconv_taps(x, w_tap, H, K)                      # [T,B,H,K]
    if tap_drop is not None:                              # weight_dropout on the soft-maxed taps -- dynamic.py:305
        w = tap_drop(w)
    xp = F.pad(x, (0, 0, 0, 0, K - 1, 0))                 # left-pad time with K-1 zeros
    win = xp.unfold(0, K, 1)                              # [T,B,C,K]; win[t,...,j] = x[t-(K-1)+j]
    y = torch.einsum('tbhrk,tbhk->tbhr', win.reshape(T, B, H, R, K), w)
    return y.reshape(T, B, C)


def lightweight_conv(x: torch.Tensor, weight: torch.Tensor, H: int, K: int, bias: Optional[torch.Tensor] = None) -> torch.Tensor:
    """LightweightConv1dTBC (padding_l = K-1, weight_softmax) -- ttl/tell/modules/convolutions/lightweight.py:149-220:
    softmax over the K static taps of every head (in fp32, :174-175 / :193-194), the same taps for every timestep and
    caption, zero padding on the left (taps reaching before t = 0 dropped after the softmax, :201-203), optional bias (:145-146).
    weight [H, 1, K]."""
    T, B, C = x.shape
    R = C // H
    w = F.softmax(weight.view(H, K).float(), dim=1).type_as(weight)
    xp = F.pad(x, (0, 0, 0, 0, K - 1, 0))
    win = xp.unfold(0, K, 1)                              # [T,B,C,K]
    y = torch.einsum('tbhrk,hk->tbhr', win.reshape(T, B, H, R, K), w).reshape(T, B, C)
    return y if bias is None else y + bias.view(1, 1, -1)


def dynamic_conv_step(x_new: torch.Tensor, hist: Optional[torch.Tensor], w_tap: torch.Tensor,
                      H: int, K: int):
    """Incremental DynamicConv -- dynamic.py:95-99,114-116: keep the last K-1 conv inputs,
    re-run the conv over [hist ; new] and keep only the new rows.  Returns (y_new, new_hist).
    When the window is shorter than K the reference narrows the tap vector to its LAST `len`
    entries (:324-326), which is what left zero padding computes."""
    xs = x_new if hist is None else torch.cat([hist, x_new], 0)
    y = dynamic_conv(xs, w_tap, H, K)
    y = y if hist is None else y[hist.shape[0]:]
    new_hist = xs[-(K - 1):] if K > 1 else xs[:0]
    return y, new_hist


# ----------------------------------------------------------------------------- attention
def mha_params(sd: SD, p: str, E: int):
    """q/k/v weights either packed `in_proj_weight` (kdim == embed_dim) or separate --
    ttl/tell/modules/attention/multi_head.py:234-240, 488-526."""
    b = sd[p + 'in_proj_bias']
    if p + 'in_proj_weight' in sd:
        w = sd[p + 'in_proj_weight']
        wq, wk, wv = w[:E], w[E:2 * E], w[2 * E:]
    else:
        wq, wk, wv = sd[p + 'q_proj_weight'], sd[p + 'k_proj_weight'], sd[p + 'v_proj_weight']
    return wq, wk, wv, b[:E], b[E:2 * E], b[2 * E:]


def mha_project_kv(mem: torch.Tensor, sd: SD, p: str, E: int):
    """K/V rows for one source incl. the learned bias row and the all-zero row:
    k = [Wk m + bk ; bias_k ; 0]  -- multi_head.py:347-374, 416-421.  mem [S,B,kdim] -> [S+2,B,E]"""
    _, wk, wv, _, bk, bv = mha_params(sd, p, E)
    S, B, _ = mem.shape
    k = F.linear(mem, wk, bk)
    v = F.linear(mem, wv, bv)
    z = mem.new_zeros(1, B, E)
    k = torch.cat([k, sd[p + 'bias_k'].expand(1, B, E), z], 0)
    v = torch.cat([v, sd[p + 'bias_v'].expand(1, B, E), z], 0)
    return k, v


def mha_attend(query: torch.Tensor, k: torch.Tensor, v: torch.Tensor, pad_mask: torch.Tensor,
               sd: SD, p: str, H: int, prob_drop=None):
    """softmax_fp32((Wq x + bq) * d^-1/2 . k^T + mask(-inf)) . v -> out_proj; also the
    head-averaged probabilities [B,T,S+2] -- multi_head.py:348-353, 376-486."""
    T, B, E = query.shape
    D = E // H
    wq, _, _, bq, _, _ = mha_params(sd, p, E)
    q = F.linear(query, wq, bq) * (D ** -0.5)
    S2 = k.shape[0]
    qh = q.reshape(T, B, H, D).permute(1, 2, 0, 3)          # [B,H,T,D]
    kh = k.reshape(S2, B, H, D).permute(1, 2, 0, 3)
    vh = v.reshape(S2, B, H, D).permute(1, 2, 0, 3)
    s = qh @ kh.transpose(-1, -2)                           # [B,H,T,S+2]
    full_mask = torch.cat([pad_mask, pad_mask.new_zeros(B, 2)], 1)   # extra 2 keys never masked
    s = s.masked_fill(full_mask[:, None, None, :], float('-inf'))
    pr = torch.softmax(s.float(), -1).to(s.dtype)
    if prob_drop is not None:                               # dropout on the probabilities -- multi_head.py:463-464
        pr = prob_drop(pr)
    o = (pr @ vh).permute(2, 0, 1, 3).reshape(T, B, E)
    out = F.linear(o, sd[p + 'out_proj.weight'], sd[p + 'out_proj.bias'])
    return out, pr.sum(1) / H


def mha(query, mem, pad_mask, sd: SD, p: str, H: int):
    k, v = mha_project_kv(mem, sd, p, query.shape[-1])
    return mha_attend(query, k, v, pad_mask, sd, p, H)


def mha_score(query, key, wq, wk, bq, bk, bias_k, add_zero_attn: bool, pad_mask, H: int, prob_drop=None):
    """`multi_head_attention_score_forward` (multi_head.py:14-204): the head-averaged attention
    probabilities only -- no values, no output projection.  q = (Wq x + bq) d^-1/2 (:128),
    k = [Wk m + bk ; bias_k ; 0] (:130-148, 165-168), key padding -> -inf on the memory rows only (:187-194),
    softmax (:197), dropout (:199), sum over heads / H (:203-204).  query [T,B,E], key [S,B,E] -> [B,T,S(+1)(+1)]"""
    T, B, E = query.shape
    D = E // H
    q = F.linear(query, wq, bq) * (D ** -0.5)
    k = F.linear(key, wk, bk)
    if bias_k is not None:
        k = torch.cat([k, bias_k.expand(1, B, E)], 0)
    if add_zero_attn:
        k = torch.cat([k, k.new_zeros(1, B, E)], 0)
    S2 = k.shape[0]
    qh = q.reshape(T, B, H, D).permute(1, 2, 0, 3)
    kh = k.reshape(S2, B, H, D).permute(1, 2, 0, 3)
    s = qh @ kh.transpose(-1, -2)                            # [B,H,T,S2]
    if pad_mask is not None:
        full = torch.cat([pad_mask, pad_mask.new_zeros(B, S2 - pad_mask.shape[1])], 1)   # extra keys never masked
        s = s.masked_fill(full[:, None, None, :], float('-inf'))
    pr = torch.softmax(s, -1)
    if prob_drop is not None:
        pr = prob_drop(pr)
    return pr.sum(1) / H


# ----------------------------------------------------------------------------- decoder layer
def decoder_layer(X: torch.Tensor, contexts: dict, sd: SD, p: str, H: int, K: int,
                  sources: Sequence[str], conv_hist: Optional[list] = None, kv_cache=None,
                  drop: Optional[DropoutMasks] = None, layer: int = 0):
    """DynamicConvDecoderLayer.forward, post-LN (decoder_normalize_before = false) --
    ttl/tell/models/decoder_flattened.py:271-345 (2 sources),
    ttl/tell/models/decoder_faces_objects.py:283-393 (4 sources; concat order image, article,
    faces, obj).  `conv_hist` is a 1-element list holding the incremental conv buffer.
    `drop` (training mode): the F.dropout calls at :283 (input), dynamic.py:305 (taps), :289, :308/:326 (attention
    outputs), :340 (ReLU), :342, and multi_head.py:463 (probabilities), with the masks of DropoutMasks."""
    site = lambda kind: layer * 16 + kind
    res = X
    if drop is not None:
        X = drop(X, drop.input_dropout, site(DROP_INPUT))
    h = F.glu(gehring_linear(X, sd, p + 'linear1.'), dim=-1)
    w_tap = sd[p + 'conv.weight_linear.weight']
    if conv_hist is None:
        tap_drop = None if drop is None else (lambda w: drop(w, drop.weight_dropout, site(DROP_TAPS)))
        h = dynamic_conv(h, w_tap, H, K, tap_drop)
    else:
        h, conv_hist[0] = dynamic_conv_step(h, conv_hist[0], w_tap, H, K)
    h = gehring_linear(h, sd, p + 'linear2.')
    if drop is not None:
        h = drop(h, drop.dropout, site(DROP_CONV_OUT))
    X = layer_norm(res + h, sd, p + 'conv_layer_norm.')

    outs, attn = [], None
    Tn, Bn, dm = X.shape
    ns = len(sources)
    for si, s in enumerate(sources):
        ap = p + 'context_attns.%s.' % s
        if kv_cache is not None and (p, s) in kv_cache:
            k, v = kv_cache[(p, s)]
        else:
            k, v = mha_project_kv(contexts[s], sd, ap, X.shape[-1])
            if kv_cache is not None:
                kv_cache[(p, s)] = (k, v)
        prob_drop = None if drop is None else (lambda pr, si=si: drop(pr, drop.attention_dropout, site(DROP_ATTN_PROB + si)))
        a, w = mha_attend(X, k, v, contexts[s + '_mask'], sd, ap, H, prob_drop)
        if drop is not None:      # element (t, b, c) of source si has counter index (t*B + b)*(ns*d) + si*d + c
            idx = (torch.arange(Tn * Bn)[:, None] * (ns * dm) + si * dm + torch.arange(dm)[None, :]).view(Tn, Bn, dm)
            a = drop(a, drop.dropout, site(DROP_ATTN_OUT), idx)
        if s == 'entity' or (s == 'article' and 'entity' not in sources):   # decoder_entity_pointer.py:326-334: the entity attention is returned
            attn = w
        outs.append(layer_norm(X + a, sd, p + 'context_attn_lns.%s.' % s))
    X = gehring_linear(torch.cat(outs, -1), sd, p + 'context_fc.')
    f = F.relu(gehring_linear(X, sd, p + 'fc1.'))
    if drop is not None:
        f = drop(f, drop.relu_dropout, site(DROP_RELU))
    f = gehring_linear(f, sd, p + 'fc2.')
    if drop is not None:
        f = drop(f, drop.dropout, site(DROP_FFN_OUT))
    X = layer_norm(X + f, sd, p + 'final_layer_norm.')
    return X, attn


def sinusoid_table(n: int, d: int, padding_idx: int = 1) -> torch.Tensor:
    """tensor2tensor sinusoids [sin | cos], row `padding_idx` zeroed --
    ttl/tell/modules/token_embedders/positional.py:125-165."""
    half = d // 2
    inc = math.log(10000.0) / (half - 1)
    inv = torch.exp(torch.arange(half, dtype=torch.float) * -inc)
    ang = torch.arange(n, dtype=torch.float)[:, None] * inv[None, :]
    tab = torch.cat([ang.sin(), ang.cos()], 1)
    tab[padding_idx] = 0
    return tab


def embed(ids: torch.Tensor, sd: SD, cutoffs: Sequence[int], start_pos: int = 0,
          p: str = 'embedder.') -> torch.Tensor:
    """SumTextFieldEmbedder(adaptive + sinusoidal position).
    adaptive: band i row (id - cutoff[i-1]) of Embedding_i, times Linear_i, times sqrt(d); the
    band-local row 0 is the nn.Embedding padding row (zeros) -- adaptive.py:36-76.
    position: pad ids (==1) -> table row 1 (zeros); else column index + 2 + start_pos --
    positional.py:167-211, 231-268; sum -- sum_text_field_embedder.py:117."""
    B, T = ids.shape
    d = sd[p + 'token_embedder_adaptive.embeddings.0.1.weight'].shape[0]
    out = torch.zeros(B, T, d, device=ids.device)       # device-agnostic: bench.py also runs this on CUDA tensors (eager comparator)
    lo = 0
    for i, hi in enumerate(cutoffs):
        m = (ids >= lo) & (ids < hi)
        if m.any():
            e = sd[p + 'token_embedder_adaptive.embeddings.%d.0.weight' % i]
            pr = sd[p + 'token_embedder_adaptive.embeddings.%d.1.weight' % i]
            # nn.Embedding(padding_idx=0): band-local row 0 receives no gradient from this path (adaptive.py:42,51)
            out[m] = F.linear(F.embedding(ids[m] - lo, e, padding_idx=0), pr).to(out.dtype)   # (.to: a no-op except under autocast)
        lo = hi
    out = out * math.sqrt(d)
    tab = sd[p + 'token_embedder_position.weights']
    pos = torch.arange(T, device=ids.device)[None, :].expand(B, T) + 2 + start_pos
    pos = torch.where(ids == 1, torch.ones_like(pos), pos)
    return out + tab[pos]


def tied_output_keys(n_clusters: int) -> Dict[str, str]:
    """alias -> owner of the tied output rows (tie_adaptive_weights, softmax.py:83-86,111-132)."""
    e = 'embedder.token_embedder_adaptive.embeddings.%d.0.weight'
    ties = {'adaptive_softmax.head.word_proj.weight': e % 0}
    for i in range(n_clusters - 1):
        ties['adaptive_softmax.tail.%d.2.weight' % i] = e % (i + 1)
    return ties


def leaf_state_dict(sd: SD, ties: Optional[Dict[str, str]] = None) -> SD:
    """Differentiable copy of a state_dict that keeps parameter TYING: entries sharing storage in `sd`
    (the tied embedding / output rows, softmax.py:83-86,111-132) share one leaf, so autograd sums their
    gradients exactly as the reference module does."""
    leaves, out = {}, {}
    for k, v in sd.items():
        if not torch.is_tensor(v) or not v.is_floating_point():
            out[k] = v
            continue
        key = (v.data_ptr(), tuple(v.shape))
        if key not in leaves:
            leaves[key] = v.detach().clone().requires_grad_(True)
        out[k] = leaves[key]
    for alias, owner in (ties or {}).items():       # fixtures saved from a state_dict lose the storage sharing
        out[alias] = out[owner]
    return out


# ----------------------------------------------------------------------------- adaptive softmax
def head_weight(sd: SD, p: str = 'adaptive_softmax.') -> torch.Tensor:
    """TiedHeadModule: [tied band-0 embedding ; class_proj] -- softmax.py:11-40."""
    return torch.cat([sd[p + 'head.word_proj.weight'], sd[p + 'head.class_proj.weight']], 0)


def adaptive_targets(target: torch.Tensor, cutoffs: Sequence[int]):
    """adapt_target -- softmax.py:144-167. Returns head target [N] and per tail (rows, local ids)."""
    t = target.reshape(-1)
    head_t = t.clone()
    tails = []
    for i in range(len(cutoffs) - 1):
        m = (t >= cutoffs[i]) & (t < cutoffs[i + 1])
        head_t[m] = cutoffs[0] + i
        tails.append((m.nonzero().squeeze(1), t[m] - cutoffs[i]))
    return head_t, tails


def adaptive_nll(x: torch.Tensor, target: torch.Tensor, sd: SD, cutoffs: Sequence[int],
                 pad: int = 1, p: str = 'adaptive_softmax.'):
    """AdaptiveSoftmax.forward + AdaptiveLoss.forward(reduction='sum') --
    softmax.py:169-191, criteria/adaptive_loss.py:27-73.  NB ignore_index=pad is applied INSIDE
    every cluster, so tail-local id 1 (tokens cutoff[i]+1) contributes no tail term.
    Returns (loss_sum, n_nonpad_tokens)."""
    X = x.reshape(-1, x.shape[-1])
    head_t, tails = adaptive_targets(target, cutoffs)
    loss = F.cross_entropy(F.linear(X, head_weight(sd, p)), head_t, ignore_index=pad,
                           reduction='sum')
    for i, (rows, loc) in enumerate(tails):
        if rows.numel():
            z = F.linear(F.linear(X[rows], sd[p + 'tail.%d.0.weight' % i]),
                         sd[p + 'tail.%d.2.weight' % i])
            loss = loss + F.cross_entropy(z, loc, ignore_index=pad, reduction='sum')
    return loss, int((target != pad).sum())


def adaptive_log_prob(x: torch.Tensor, sd: SD, cutoffs: Sequence[int],
                      p: str = 'adaptive_softmax.') -> torch.Tensor:
    """get_log_prob: full-vocabulary log-probs -- softmax.py:193-222.  x [B,T,d] -> [B,T,V]"""
    B, T, d = x.shape
    X = x.reshape(-1, d)
    hl = F.log_softmax(F.linear(X, head_weight(sd, p)), 1)
    parts = [hl[:, :cutoffs[0]]]
    for i in range(len(cutoffs) - 1):
        z = F.linear(F.linear(X, sd[p + 'tail.%d.0.weight' % i]), sd[p + 'tail.%d.2.weight' % i])
        parts.append(F.log_softmax(z, 1) + hl[:, cutoffs[0] + i, None])
    return torch.cat(parts, 1).view(B, T, -1)


# ----------------------------------------------------------------------------- whole decoder
def n_layers(sd: SD) -> int:
    return 1 + max(int(k.split('.')[1]) for k in sd if k.startswith('layers.'))


def layer_kernel(sd: SD, i: int, H: int) -> int:
    return sd['layers.%d.conv.weight_linear.weight' % i].shape[0] // H


def decoder_forward(ids: torch.Tensor, contexts: dict, sd: SD, H: int, cutoffs: Sequence[int],
                    sources: Sequence[str] = ('image', 'article'), state: Optional[dict] = None,
                    drop: Optional[DropoutMasks] = None):
    """DynamicConvDecoder.forward -- decoder_flattened.py:96-146 / decoder_faces_objects.py:95-144
    (eval mode: every dropout is the identity; `drop` = training mode with given masks).  ids [B,T] -> X [B,T,d], article attn [B,T,S+2].
    `state` (dict) switches on the incremental path: position offset + per-layer conv buffers
    (+ a K/V cache, which the reference recomputes each step -- same numbers)."""
    start = 0
    if state is not None:
        start = state.get('pos', 0)
        state['pos'] = start + ids.shape[1]
    X = embed(ids, sd, cutoffs, start).transpose(0, 1)
    if drop is not None:                                   # decoder_flattened.py:109 (counter index in T x B order)
        X = drop(X.contiguous(), drop.dropout, DROP_SITE_EMBED)
    attn = None
    for i in range(n_layers(sd)):
        hist = None
        if state is not None:
            hist = state.setdefault(('conv', i), [None])
        X, attn = decoder_layer(X, contexts, sd, 'layers.%d.' % i, H, layer_kernel(sd, i, H),
                                sources, hist, None if state is None else state.setdefault('kv', {}), drop, i)
    return X.transpose(0, 1), attn


def pointer_output_layer(X: torch.Tensor, inp_embed: torch.Tensor, attn: torch.Tensor, article_ids: torch.Tensor,
                         sd: SD, cutoffs: Sequence[int]):
    """Copy head of DynamicConvDecoderPointer -- decoder_pointer.py:138-141 (p_gens = project_p_gens(cat(inp_embed, X)))
    and :151-216 (output_layer): probs = sigmoid(p_gens) * softmax_vocab(X) + scatter_add over the article token ids of
    (1 - sigmoid(p_gens)) * attn[:, :, :-2]; lprobs = log(clamp(probs, 1e-10, 1)).  X, inp_embed [B,T,d]; attn [B,T,S+2]
    (head-averaged article attention of the last layer); article_ids [B,S].  Returns (lprobs [B,T,V], p_gen [B,T,1])."""
    p_gens = torch.sigmoid(gehring_linear(torch.cat((inp_embed, X), 2), sd, 'project_p_gens.'))
    return copy_mix(adaptive_log_prob(X, sd, cutoffs).exp(), p_gens, attn, article_ids), p_gens


def pointer_decoder_forward(ids: torch.Tensor, contexts: dict, article_ids: torch.Tensor, sd: SD, H: int,
                            cutoffs: Sequence[int]):
    """DynamicConvDecoderPointer.forward (decoder_pointer.py:98-149), eval mode: lprobs [B,T,V], attn, p_gen."""
    X, attn = decoder_forward(ids, contexts, sd, H, cutoffs, ('image', 'article'))
    lp, pg = pointer_output_layer(X, embed(ids, sd, cutoffs), attn, article_ids, sd, cutoffs)
    return lp, attn, pg


ENTITY_SOURCES = ('image', 'article', 'entity')


def copy_mix(gen_probs: torch.Tensor, p_gens: torch.Tensor, attn: torch.Tensor, copy_ids: torch.Tensor) -> torch.Tensor:
    """The mix shared by both pointer heads (decoder_pointer.py:176-216 / transformer_entity_pointer.py:212-261):
    log(clamp(p_gen * gen + scatter_add(copy_ids, (1 - p_gen) * attn[:, :, :-2]), 1e-10, 1))."""
    B, T, V = gen_probs.shape
    copy = attn[:, :, :-2].float() * (1 - p_gens)
    index = copy_ids[:, None, :].expand(B, T, copy_ids.shape[1])
    dist = torch.zeros(B, T, V).scatter_add_(2, index.long(), copy)
    return (gen_probs * p_gens + dist).clamp(1e-10, 1.0).log()


def entity_pointer_output_layer(X: torch.Tensor, inp_embed: torch.Tensor, attn: torch.Tensor, entity_tokens: torch.Tensor,
                                sd: SD, cutoffs: Sequence[int]):
    """Copy head of TransformerEntityPointerModel.output_layer (transformer_entity_pointer.py:185-261), eval mode:
    p_gens = sigmoid(project_p_gens(relu(project_first_p_gens(predictors)))), predictors = cat(inp_embed, X)
    (decoder_entity_pointer.py:144); attn = the ENTITY attention of the last layer (decoder_entity_pointer.py:326-334).
    `sd` holds the decoder's keys plus the model's `project_first_p_gens.*` / `project_p_gens.*`."""
    hid = torch.relu(gehring_linear(torch.cat((inp_embed, X), 2), sd, 'project_first_p_gens.'))
    p_gens = torch.sigmoid(gehring_linear(hid, sd, 'project_p_gens.'))
    return copy_mix(adaptive_log_prob(X, sd, cutoffs).exp(), p_gens, attn, entity_tokens), p_gens


def entity_pointer_forward(ids: torch.Tensor, contexts: dict, entity_tokens: torch.Tensor, sd: SD, H: int,
                           cutoffs: Sequence[int]):
    """DynamicConvDecoderEntityPointer.forward (decoder_entity_pointer.py:97-145) + the model's output_layer."""
    X, attn = decoder_forward(ids, contexts, sd, H, cutoffs, ENTITY_SOURCES)
    lp, pg = entity_pointer_output_layer(X, embed(ids, sd, cutoffs), attn, entity_tokens, sd, cutoffs)
    return lp, attn, pg


def pointer_greedy_generate(seed_ids: torch.Tensor, contexts: dict, copy_ids: torch.Tensor, sd: SD, H: int,
                            cutoffs: Sequence[int], entity: bool = False, gen_len: int = 100, eos: int = 2, pad: int = 1):
    """`_generate` of the pointer models with sampling_topk = 1 (transformer_entity_pointer.py:375-470; the context
    pointer model runs the same loop over the article ids): per step the incremental decoder, the copy head on the
    step's (X, inp_embed, attn), arg-max of the mixed log-probabilities.  Finished rows are masked instead of dropped
    (rows are independent).  Returns (log_probs [B, L], token_ids [B, L+1])."""
    sources = ENTITY_SOURCES if entity else ('image', 'article')
    head = entity_pointer_output_layer if entity else pointer_output_layer
    state: dict = {}
    active = seed_ids[:, -1] != eos
    toks, lps = [seed_ids[:, :1]], []
    cur = seed_ids[:, :1]
    for _ in range(gen_len):
        start = state.get('pos', 0)
        X, attn = decoder_forward(cur, contexts, sd, H, cutoffs, sources, state)
        lp, _ = head(X[:, -1:], embed(cur, sd, cutoffs, start)[:, -1:], attn[:, -1:], copy_ids, sd, cutoffs)
        best_lp, best = lp[:, 0].max(-1)
        tok = torch.where(active, best, torch.full_like(best, pad))
        lps.append(torch.where(active, best_lp, torch.zeros_like(best_lp))[:, None])
        toks.append(tok[:, None])
        active = active & (best != eos)
        cur = best[:, None]
        if not bool(active.any()):
            break
    return torch.cat(lps, 1), torch.cat(toks, 1)


def greedy_generate(seed_ids: torch.Tensor, contexts: dict, sd: SD, H: int,
                    cutoffs: Sequence[int], sources=('image', 'article'), gen_len: int = 100,
                    eos: int = 2, pad: int = 1):
    """`_generate` with sampling_topk=1 (= arg-max) -- ttl/tell/models/transformer_flattened.py:
    238-325 / transformer_faces_objects.py:399-494.  Finished rows are dropped from the batch
    (here: masked, which is numerically identical since rows are independent); outputs are padded
    with `pad` ids / 0 log-probs.  Returns (log_probs [B,L], token_ids [B,L+1])."""
    B = seed_ids.shape[0]
    state: dict = {}
    active = seed_ids[:, -1] != eos
    toks, lps = [seed_ids[:, :1]], []
    cur = seed_ids[:, :1]
    for _ in range(gen_len):
        X, _ = decoder_forward(cur, contexts, sd, H, cutoffs, sources, state)
        lp = adaptive_log_prob(X[:, -1:], sd, cutoffs)[:, 0]
        best_lp, best = lp.max(-1)
        tok = torch.where(active, best, torch.full_like(best, pad))
        lps.append(torch.where(active, best_lp, torch.zeros_like(best_lp))[:, None])
        toks.append(tok[:, None])
        active = active & (best != eos)
        cur = best[:, None]
        if not bool(active.any()):
            break
    return torch.cat(lps, 1), torch.cat(toks, 1)


def module_generate_loop(decoder, seed_ids: torch.Tensor, contexts: dict, sources=('image', 'article'),
                         gen_len: int = 100, eos: int = 2, pad: int = 1, index: str = 'roberta'):
    """The reference's `_generate` loop body (transformer_flattened.py:240-325 / transformer_faces_objects.py:399-494 with
    sampling_topk = 1, where the multinomial draw over one candidate is deterministic) driven around ANY decoder object
    that speaks the reference's module protocol: `decoder.filter_incremental_state(state, active_idx)` (:256-257),
    `decoder(prev_target, contexts_i, incremental_state=state)` on the batch compacted to the active rows (:259-271),
    `decoder.get_normalized_probs(...)` (:276-277), scatter of the step's ids / log-probs into [B, 1] (:303-307), EOS
    bookkeeping (:314-323).  oracle/make_golden.py ran it around the UNMODIFIED reference decoder to freeze the
    `greedy_ids` fixtures; tests run it around the B200 mirror.  Returns (log_probs [B, L], token_ids [B, L+1])."""
    state: dict = {}
    B = seed_ids.shape[0]
    active = seed_ids[:, -1] != eos
    full_active = active.clone()
    seed_input = seed_ids
    lps, toks = [], [seed_ids]
    for i in range(gen_len):
        prev = {index: seed_input if i == 0 else seed_input[:, -1:]}
        decoder.filter_incremental_state(state, active)
        c = {}
        for s in sources:
            c[s] = contexts[s][:, full_active]
            c[s + '_mask'] = contexts[s + '_mask'][full_active]
        out = decoder(prev, c, incremental_state=state)
        lp = decoder.get_normalized_probs((out[0][:, -1:], None), log_probs=True).squeeze(1)
        top_lp, top_i = lp.topk(1)
        log_prob = top_lp.new_zeros(B, 1)
        log_prob[full_active] = top_lp
        path = top_i.new_full((B, 1), pad)
        path[full_active] = top_i
        lps.append(log_prob)
        toks.append(path)
        seed_input = torch.cat([seed_input, top_i], -1)
        active = top_i.squeeze(-1) != eos
        full_active[full_active.nonzero()[~active]] = 0
        seed_input = seed_input[active]
        if active.sum().item() == 0:
            break
    return torch.cat(lps, -1), torch.cat(toks, -1)


# ----------------------------------------------------------------------------- beam search
def beam_search(step_fn, reorder_fn, beam_size: int, seq_length: int, eos: int, return_all: bool = False):
    """`CaptionModel.beam_search` -- models/CaptionModel.py:23-122, restated for ONE caption.
    step_fn(tokens [W] or None for the seed step) -> log-probs [W, V] of the next word for every beam;
    reorder_fn(order [W]) forks the recurrent state (new beam vix continues old beam order[vix]; the reference
    copies `state[:, q]`, for the tell decoder this is `reorder_incremental_state`).  `eos` is the end token (0 there).
    Returns the list of finished beams sorted by score, at most beam_size: dicts seq [L], logps [L], p
    (`return_all`: every finished beam in arrival order, unsorted).
    Quirk of the reference under torch >= 0.4: `'p': beam_logprobs_sum[vix]` stores a 0-dim VIEW, which the next line
    (`beam_logprobs_sum[vix] = -1000`) overwrites, so its final sort sees equal keys and keeps arrival order; the score
    the code means (the value at finishing time, what old torch returned as a Python float) is what is restated here."""
    W, L = beam_size, seq_length
    beam_seq = torch.zeros(L, W, dtype=torch.long)
    beam_seq_logprobs = torch.zeros(L, W)
    beam_logprobs_sum = torch.zeros(W)
    done_beams = []
    logprobs = step_fn(None)
    for t in range(L):
        logprobsf = logprobs.float()
        ys, ix = torch.sort(logprobsf, 1, True)                      # :41
        candidates = []
        cols = min(W, ys.size(1))
        rows = 1 if t == 0 else W                                    # :44-46
        for c in range(cols):
            for q in range(rows):
                local = ys[q, c]
                candidates.append({'c': ix[q, c], 'q': q, 'p': beam_logprobs_sum[q] + local, 'r': local})
        candidates = sorted(candidates, key=lambda x: -x['p'])       # stable, :53
        if t >= 1:
            seq_prev = beam_seq[:t].clone()
            lp_prev = beam_seq_logprobs[:t].clone()
        order = []
        for vix in range(W):
            v = candidates[vix]
            if t >= 1:
                beam_seq[:t, vix] = seq_prev[:, v['q']]
                beam_seq_logprobs[:t, vix] = lp_prev[:, v['q']]
            order.append(v['q'])
            beam_seq[t, vix] = v['c']
            beam_seq_logprobs[t, vix] = v['r']
            beam_logprobs_sum[vix] = v['p']
        reorder_fn(torch.tensor(order))
        for vix in range(W):                                         # :102-112
            if beam_seq[t, vix] == eos or t == L - 1:
                done_beams.append({'seq': beam_seq[:, vix].clone(), 'logps': beam_seq_logprobs[:, vix].clone(),
                                   'p': beam_logprobs_sum[vix].clone()})
                beam_logprobs_sum[vix] = -1000
        logprobs = step_fn(beam_seq[t])
    if return_all:
        return done_beams
    return sorted(done_beams, key=lambda x: -x['p'])[:W]


def decoder_beam_search(seed_id: int, contexts: dict, sd: SD, H: int, cutoffs: Sequence[int], sources, beam_size: int,
                        gen_len: int, eos: int = 2):
    """Beam search of ONE caption over the incremental tell decoder: contexts hold that caption only ([S,1,dim]);
    the W beams form the batch.  State fork = reorder_incremental_state (dynamic.py:338-348): index_select(1, order)
    on every conv input buffer."""
    W = beam_size
    ctx = {k: (v.expand(v.shape[0], W, v.shape[2]) if v.dim() == 3 else v.expand(W, v.shape[1])) for k, v in contexts.items()
           if v is not None}
    state: dict = {}

    def step(tokens):
        cur = torch.full((W, 1), seed_id, dtype=torch.long) if tokens is None else tokens.view(W, 1)
        X, _ = decoder_forward(cur, ctx, sd, H, cutoffs, sources, state)
        return adaptive_log_prob(X[:, -1:], sd, cutoffs)[:, 0]

    def reorder(order):
        for k, v in state.items():
            if isinstance(k, tuple) and k[0] == 'conv' and v[0] is not None:
                v[0] = v[0].index_select(1, order)

    return beam_search(step, reorder, W, gen_len, eos)


# ----------------------------------------------------------------------------- context assembly
def context_mix(hiddens: Sequence[torch.Tensor], bert_weight: torch.Tensor) -> torch.Tensor:
    """RoBERTa hidden-layer mix -- ttl/tell/models/transformer_faces_objects.py:355-366 (weigh_bert) and the
    `.transpose(0, 1)` of :383.  hiddens: n tensors [B,S,E]; bert_weight [n] -> [S,B,E]."""
    X = torch.stack(list(hiddens), dim=2)
    w = F.softmax(bert_weight, dim=0).unsqueeze(0).unsqueeze(1).unsqueeze(3)
    return (X * w).sum(dim=2).transpose(0, 1)


def nan_padding_mask(embeds: torch.Tensor):
    """faces / objects: NaN rows are padding -- ttl/tell/models/transformer_faces_objects.py:371-377 (+ :388,390).
    embeds [B,n,d] -> ([n,B,d] with padding rows zeroed, mask [B,n] True = padding)."""
    m = torch.isnan(embeds).any(dim=-1)
    e = embeds.clone()
    e[m] = 0
    return e.transpose(0, 1), m


# ----------------------------------------------------------------------------- final*/transformer_v2
def v2_layer_norm(x, a, b, eps=1e-6):
    """final*/transformer_v2.py:82-94: UNBIASED std, eps added to the std (outside the sqrt)."""
    mu = x.mean(-1, keepdim=True)
    sdv = x.std(-1, keepdim=True)
    return a * (x - mu) / (sdv + eps) + b


def v2_mha(query, mem, mask, sd: SD, p: str, h: int):
    """MultiHeadedAttention.forward + attention() -- final_roberta2/transformer_v2.py:213-266.
    scores/sqrt(d_k), masked_fill(mask == 0, -1e9), softmax, PV, output linear.
    query [B,T,d]; mem [B,S,d_key]; mask broadcastable to [B,1,T,S] or None."""
    B, T, d = query.shape
    dk = d // h
    q = F.linear(query, sd[p + 'linears.0.weight'], sd[p + 'linears.0.bias'])
    k = F.linear(mem, sd[p + 'linear_kv.0.weight'], sd[p + 'linear_kv.0.bias'])
    v = F.linear(mem, sd[p + 'linear_kv.1.weight'], sd[p + 'linear_kv.1.bias'])
    q = q.view(B, T, h, dk).transpose(1, 2)
    k = k.view(B, -1, h, dk).transpose(1, 2)
    v = v.view(B, -1, h, dk).transpose(1, 2)
    s = q @ k.transpose(-1, -2) / math.sqrt(dk)
    if mask is not None:
        s = s.masked_fill(mask.unsqueeze(1) == 0, -1e9)
    o = (F.softmax(s, -1) @ v).transpose(1, 2).reshape(B, T, d)
    return F.linear(o, sd[p + 'linears.1.weight'], sd[p + 'linears.1.bias'])


def v2_decoder_layer(x, memory, tgt_mask, sd: SD, p: str, h: int):
    """DecoderLayer.forward, pre-norm SublayerConnection (x + sublayer(norm(x))) --
    final_roberta2/transformer_v2.py:96-109, 186-211.  img-attn and article-attn both read the
    post-self-attention x; their residual outputs are concatenated into context_fc."""
    def nrm(t, i):
        return v2_layer_norm(t, sd[p + 'sublayer.%d.norm.a_2' % i], sd[p + 'sublayer.%d.norm.b_2' % i])
    n = nrm(x, 0)
    x = x + v2_mha(n, n, tgt_mask, sd, p + 'self_attn.', h)
    xi = x + v2_mha(nrm(x, 1), memory['image'], None, sd, p + 'img_attn.', h)
    xa = x + v2_mha(nrm(x, 2), memory['article'], memory.get('article_mask'), sd,
                    p + 'article_attn.', h)
    x = F.linear(torch.cat([xi, xa], -1), sd[p + 'context_fc.weight'], sd[p + 'context_fc.bias'])
    n = nrm(x, 3)
    f = F.linear(F.relu(F.linear(n, sd[p + 'feed_forward.w_1.weight'], sd[p + 'feed_forward.w_1.bias'])),
                 sd[p + 'feed_forward.w_2.weight'], sd[p + 'feed_forward.w_2.bias'])
    return x + f


def v2_positional(T: int, d: int) -> torch.Tensor:
    """PositionalEncoding: interleaved sin/cos -- final_roberta2/transformer_v2.py (PositionalEncoding)."""
    pe = torch.zeros(T, d)
    pos = torch.arange(0, T, dtype=torch.float).unsqueeze(1)
    div = torch.exp(torch.arange(0, d, 2).float() * -(math.log(10000.0) / d))
    pe[:, 0::2] = torch.sin(pos * div)
    pe[:, 1::2] = torch.cos(pos * div)
    return pe


def v2_decode(ids, memory, tgt_mask, sd: SD, h: int):
    """OnlyDecoder.decode: Embeddings(x sqrt d) + PositionalEncoding -> N layers -> LayerNorm --
    final_roberta2/transformer_v2.py:45-51, 111-140."""
    lut = sd['tgt_embed.0.lut.weight']
    d = lut.shape[1]
    x = lut[ids] * math.sqrt(d) + v2_positional(ids.shape[1], d)[None]
    L = 1 + max(int(k.split('.')[2]) for k in sd if k.startswith('decoder.layers.'))
    for i in range(L):
        x = v2_decoder_layer(x, memory, tgt_mask, sd, 'decoder.layers.%d.' % i, h)
    return v2_layer_norm(x, sd['decoder.norm.a_2'], sd['decoder.norm.b_2'])


def v2_generator_loss(out, y, sd: SD, pad: int, norm: float):
    """Generator.proj + CrossEntropyLoss(ignore_index=pad, reduction='sum') / norm --
    final_roberta2/transformer_v2.py:54-62, final_roberta2/custom_v4.py:404-419."""
    z = F.linear(out, sd['generator.proj.weight'], sd['generator.proj.bias'])
    return F.cross_entropy(z.reshape(-1, z.shape[-1]), y.reshape(-1), ignore_index=pad,
                           reduction='sum') / norm


def v2_subsequent_mask(T: int) -> torch.Tensor:
    """final_roberta2/transformer_v2.py:391-397: lower-triangular (incl. diagonal) = visible."""
    return torch.tril(torch.ones(1, T, T, dtype=torch.bool))


def v2_translate(memory, sd: SD, h: int, max_len: int, bos: int = 0):
    """translate_sentence: fixed max_len-1 greedy steps, whole prefix re-decoded each step,
    no EOS stop -- final_roberta2/custom_v4.py:647-662."""
    B = memory['image'].shape[0]
    ys = torch.full((B, 1), bos, dtype=torch.long)
    for _ in range(max_len - 1):
        out = v2_decode(ys, memory, v2_subsequent_mask(ys.shape[1]), sd, h)
        z = F.linear(out[:, -1], sd['generator.proj.weight'], sd['generator.proj.bias'])
        ys = torch.cat([ys, F.log_softmax(z, -1).argmax(-1, keepdim=True)], 1)
    return ys


# ----------------------------------------------------------------------------- optimizer
def warmup_linear(x: float, warmup: float) -> float:
    return x / warmup if x < warmup else max((x - 1.0) / (warmup - 1.0), 0.0)


def bert_adam_step(params: SD, grads: SD, state: dict, lr: float, b1: float = 0.9, b2: float = 0.98,
                   e: float = 1e-6, weight_decay: float = 1e-5, max_grad_norm: float = 0.1):
    """One BertAdam update, the optimizer the reference configures (config.yaml:126-136, AllenNLP `bert_adam`
    = pytorch_pretrained_bert.optimization.BertAdam 0.6.x -- un-vendored third party; published algorithm):
    per-parameter clip_grad_norm_, no bias correction, L2 weight decay added to the update.  In place."""
    for k, p in params.items():
        g = grads[k].clone()
        if max_grad_norm > 0:
            coef = max_grad_norm / (g.norm() + 1e-6)
            if coef < 1:
                g = g * coef
        m = state.setdefault(('m', k), torch.zeros_like(p))
        v = state.setdefault(('v', k), torch.zeros_like(p))
        m.mul_(b1).add_(g, alpha=1 - b1)
        v.mul_(b2).addcmul_(g, g, value=1 - b2)
        upd = m / (v.sqrt() + e)
        if weight_decay > 0:
            upd = upd + weight_decay * p
        p.sub_(lr * upd)
