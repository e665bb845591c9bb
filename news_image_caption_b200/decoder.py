"""Host-side mirror of the reference decoder interface, backed by the C-ABI engine.

`DynamicConvDecoder` keeps the reference's constructor meaning, `forward(prev_target, contexts,
incremental_state)` signature, `get_normalized_probs`, and -- through an attribute tree with the same
names -- the reference's `state_dict` keys, so a reference checkpoint loads unchanged
(ttl/tell/models/decoder_flattened.py:22-207, decoder_faces_objects.py:21-208).  `AdaptiveLoss` mirrors
ttl/tell/modules/criteria/adaptive_loss.py:10-91.  torch is used for device memory and the stream only.
"""
import ctypes as C
import math

import torch
import torch.nn as nn

from . import _lib
from . import state as S_
from ._lib import TNT_BF16, TNT_F32

# concat order: decoder_faces_objects.py:296-383 (image, article, faces, obj), decoder_entity_pointer.py:287-340 (image, article, entity)
SOURCE_ORDER = ('image', 'article', 'entity', 'faces', 'obj')
DEFAULT_SOURCE_DIM = {'image': 2048, 'article': 1024, 'entity': 1024, 'faces': 512, 'obj': 2048}


def _ptr(t):
    return C.c_void_p(t.data_ptr()) if t is not None else None


def _stream():
    return C.c_void_p(torch.cuda.current_stream().cuda_stream)


def _dtype_code(dtype):
    if dtype in (torch.bfloat16, 'bf16', TNT_BF16):
        return TNT_BF16, torch.bfloat16
    if dtype in (torch.float32, 'fp32', 'f32', TNT_F32):
        return TNT_F32, torch.float32
    raise ValueError('dtype must be torch.bfloat16 or torch.float32')


def _register(root, key, tensor):
    """Registers `tensor` as parameter `key` (dotted) under nested container modules."""
    mod = root
    parts = key.split('.')
    for p in parts[:-1]:
        if not hasattr(mod, p):
            mod.add_module(p, nn.Module())
        mod = getattr(mod, p)
    mod.register_parameter(parts[-1], nn.Parameter(tensor, requires_grad=False))


def reference_shapes(d, heads, ffn, kernels, vocab, cutoffs, sources, source_dim, n_pos=513):
    """state_dict layout of the reference decoder."""
    sh = {}
    lo = 0
    for i, hi in enumerate(cutoffs):
        sh['embedder.token_embedder_adaptive.embeddings.%d.0.weight' % i] = (hi - lo, d)
        sh['embedder.token_embedder_adaptive.embeddings.%d.1.weight' % i] = (d, d)
        lo = hi
    sh['embedder.token_embedder_position.weights'] = (n_pos, d)
    for l, K in enumerate(kernels):
        p = 'layers.%d.' % l
        for name, (o, i) in (('linear1', (2 * d, d)), ('linear2', (d, d)), ('context_fc', (d, len(sources) * d)),
                             ('fc1', (ffn, d)), ('fc2', (d, ffn))):
            sh[p + name + '.bias'] = (o,)
            sh[p + name + '.weight_g'] = (o, 1)
            sh[p + name + '.weight_v'] = (o, i)
        sh[p + 'conv.weight_linear.weight'] = (heads * K, d)
        for ln in ('conv_layer_norm', 'final_layer_norm'):
            sh[p + ln + '.weight'] = (d,)
            sh[p + ln + '.bias'] = (d,)
        for s in sources:
            a = p + 'context_attns.%s.' % s
            kd = source_dim[s]
            if kd == d:
                sh[a + 'in_proj_weight'] = (3 * d, d)
            else:
                sh[a + 'k_proj_weight'] = (d, kd)
                sh[a + 'v_proj_weight'] = (d, kd)
                sh[a + 'q_proj_weight'] = (d, d)
            sh[a + 'in_proj_bias'] = (3 * d,)
            sh[a + 'bias_k'] = (1, 1, d)
            sh[a + 'bias_v'] = (1, 1, d)
            sh[a + 'out_proj.weight'] = (d, d)
            sh[a + 'out_proj.bias'] = (d,)
            sh[p + 'context_attn_lns.%s.weight' % s] = (d,)
            sh[p + 'context_attn_lns.%s.bias' % s] = (d,)
    sh['adaptive_softmax.head.word_proj.weight'] = (cutoffs[0], d)
    if len(cutoffs) > 1:
        sh['adaptive_softmax.head.class_proj.weight'] = (len(cutoffs) - 1, d)
    for i in range(len(cutoffs) - 1):
        sh['adaptive_softmax.tail.%d.0.weight' % i] = (d, d)
        sh['adaptive_softmax.tail.%d.2.weight' % i] = (cutoffs[i + 1] - cutoffs[i], d)
    return sh


IGNORED_KEYS = ('version', 'adaptive_softmax.version', 'adaptive_softmax.head._float_tensor')


class DynamicConvDecoder(nn.Module):
    """B200 engine behind the reference `DynamicConvDecoder` / `DynamicConvFacesObjectsDecoder` API
    (post-LN, GLU, dynamic conv, weight_softmax, adaptive softmax with tied weights: the configuration
    every shipped experiment uses, ttl/expt/*/config.yaml)."""
    HOST_SIDE_KEYS = ('\0',)          # prefixes of parameters that are not fed to the engine handle (none here)

    def __init__(self, decoder_output_dim=1024, decoder_attention_heads=16, decoder_ffn_embed_dim=4096,
                 decoder_kernel_size_list=(3, 7, 15, 31), vocab_size=50265, adaptive_softmax_cutoff=(5000, 20000),
                 sources=('image', 'article'), source_dim=None, padding_idx=1, max_positions=513,
                 dtype=torch.bfloat16, device='cuda', init_std=None, embed_padding_idx=0):
        super().__init__()
        self.d, self.heads, self.ffn = decoder_output_dim, decoder_attention_heads, decoder_ffn_embed_dim
        self.kernels = list(decoder_kernel_size_list)
        self.cutoffs = list(adaptive_softmax_cutoff) + [vocab_size]
        self.vocab = vocab_size
        self.sources = [s for s in SOURCE_ORDER if s in sources]
        self.source_dim = dict(DEFAULT_SOURCE_DIM)
        self.source_dim.update(source_dim or {})
        self.pad = padding_idx
        self.embed_padding_idx = -1 if embed_padding_idx is None else int(embed_padding_idx)   # config.yaml:39 (adaptive.py:42)
        self.max_positions = max_positions
        self.dtype_code, self.torch_dtype = _dtype_code(dtype)
        shapes = reference_shapes(self.d, self.heads, self.ffn, self.kernels, self.vocab, self.cutoffs,
                                  self.sources, self.source_dim, max_positions)
        for k, shp in shapes.items():
            _register(self, k, torch.zeros(shp, dtype=torch.float32, device=device))
        self._handle = None
        self._weights = None
        self._dirty = True
        self._ws = {}
        self._pending = None
        # incremental_state keys, reference scheme (utils/state.py:17-26): one DynamicConv1dTBC input buffer per layer, the
        # positional embedder's step counter, and (this mirror's own) the projected K/V of the contexts
        self._conv_keys = [S_.full_key('DynamicConv1dTBC', S_.new_instance_id('DynamicConv1dTBC'), 'input_buffer')
                           for _ in self.kernels]
        self._pos_key = S_.full_key('SinusoidalPositionalEmbedding', S_.new_instance_id('SinusoidalPositionalEmbedding'),
                                    'position')
        self._kv_key = S_.full_key(type(self).__name__, S_.new_instance_id(type(self).__name__), 'projected_kv')
        self.register_load_state_dict_post_hook(lambda m, ik: setattr(m, '_dirty', True))

    # ------------------------------------------------------------------ state_dict compatibility
    def load_reference_state_dict(self, sd):
        """Loads a reference checkpoint: tied output weights appear under both names there; buffers
        that carry no numbers (`version`, `_float_tensor`) are ignored."""
        own = dict(self.named_parameters())
        missing = [k for k in own if k not in sd]
        if missing:
            raise KeyError('reference state_dict lacks %s' % missing[:4])
        unexpected = [k for k in sd if k not in own and k not in IGNORED_KEYS]
        if unexpected:
            raise KeyError('unexpected keys %s' % unexpected[:4])
        with torch.no_grad():
            for k, p in own.items():
                p.copy_(sd[k].to(p.device, torch.float32).reshape(p.shape))
        self._dirty = True
        return self

    # ------------------------------------------------------------------ engine plumbing
    def _config(self):
        c = _lib.DecoderConfig()
        c.n_layers, c.d, c.heads, c.ffn, c.vocab = len(self.kernels), self.d, self.heads, self.ffn, self.vocab
        for i, k in enumerate(self.kernels):
            c.kernels[i] = k
        c.n_sources = len(self.sources)
        for i, s in enumerate(self.sources):
            c.source_names[i].value = s.encode()
            c.source_dim[i] = self.source_dim[s]
        c.n_clusters = len(self.cutoffs)
        for i, v in enumerate(self.cutoffs):
            c.cutoffs[i] = v
        c.pad_idx, c.max_positions, c.dtype = self.pad, self.max_positions, self.dtype_code
        c.embed_padding_idx = self.embed_padding_idx
        return c

    def prepare(self):
        """(Re)builds the prepared weights from the current parameters (weight-norm fold, fusion)."""
        L = _lib.lib()
        _lib.require_device()
        if self._handle is None:
            cfg = self._config()
            self._handle = L.tnt_decoder_create(C.byref(cfg))
            if not self._handle:
                _lib.check(-1, 'tnt_decoder_create')
            nbytes = L.tnt_decoder_weight_bytes(self._handle)
            self._weights = torch.empty(nbytes, dtype=torch.uint8, device='cuda')
            _lib.check(L.tnt_decoder_bind_weights(self._handle, _ptr(self._weights), nbytes), 'bind_weights')
        keep, temps, ptrs = [], False, {}
        for k, p in self.named_parameters(remove_duplicate=False):
            if k.startswith(self.HOST_SIDE_KEYS):       # parameters of heads that live outside the engine handle
                continue
            t = p.detach().contiguous()
            temps = temps or t.data_ptr() != p.data_ptr()
            keep.append((k, t))
            ptrs[k] = t.data_ptr()
        # the engine keeps the source pointers: re-register them only when one moved (a training step refreshes the prepared
        # weights every step -- ~150 registration calls and a stream synchronise per step were 0.25 ms of GPU idle time)
        if temps or ptrs != getattr(self, '_loaded_ptrs', None):
            for k, t in keep:
                shp = (C.c_int64 * t.dim())(*t.shape)
                _lib.check(L.tnt_decoder_load(self._handle, k.encode(), _ptr(t), shp, t.dim(), _stream()), 'load ' + k)
            self._loaded_ptrs = None if temps else ptrs
        _lib.check(L.tnt_decoder_finalize(self._handle, _stream()), 'tnt_decoder_finalize')
        if temps:
            torch.cuda.current_stream().synchronize()     # the contiguous temporaries may be released after this
        self._dirty = False
        self._seen_version = self._param_version()

    def _param_version(self):
        """Sum of the parameters' in-place version counters: any torch optimizer / in-place edit of a parameter bumps it, so
        the prepared (folded, cast, fused) weights are rebuilt before the next engine call without the caller saying so."""
        return sum(p._version for p in self.parameters())

    def _h(self):
        if self._dirty or self._handle is None or self._param_version() != getattr(self, '_seen_version', -1):
            self.prepare()
        return self._handle

    def __del__(self):
        try:
            if self._handle:
                _lib.lib().tnt_decoder_destroy(self._handle)
        except Exception:
            pass

    def check_input_errors(self):
        """Raises if a kernel of this decoder saw a token id outside [0, vocab) (nn.Embedding's IndexError in the
        reference, adaptive.py:61-76) or a position beyond the sinusoid table.  Synchronises the current stream;
        called wherever the mirror already synchronises (loss `.item()`, trimmed generation, host scoring)."""
        flags = _lib.lib().tnt_decoder_error_flags(self._h(), _stream())
        if flags < 0:
            _lib.check(flags, 'tnt_decoder_error_flags')
        if flags & 1:
            raise IndexError('token id outside [0, %d) in the decoder input (the affected rows were written as NaN)' % self.vocab)
        if flags & 2:
            raise IndexError('position beyond the sinusoid table (max_positions=%d)' % self.max_positions)

    def _workspace(self, key, nbytes):
        t = self._ws.get(key)
        if t is None or t.numel() < nbytes:
            t = torch.empty(nbytes + 256, dtype=torch.uint8, device='cuda')
            self._ws[key] = t
        return t

    def _contexts(self, contexts, B, T):
        c = _lib.Contexts()
        c.B, c.T = B, T
        keep = []
        for i, s in enumerate(self.sources):
            mem = contexts[s]
            if mem.dim() != 3 or mem.shape[1] != B or mem.shape[2] != self.source_dim[s]:
                raise ValueError('context %s must be [S, %d, %d], got %s' % (s, B, self.source_dim[s],
                                                                             tuple(mem.shape)))
            mem = mem.to(self.torch_dtype).contiguous()
            mask = contexts.get(s + '_mask')
            c.S[i] = mem.shape[0]
            c.memory[i] = mem.data_ptr() if mem.numel() else None
            if mask is not None and mem.shape[0] > 0:
                mask = mask.to(torch.bool).contiguous()
                if tuple(mask.shape) != (B, mem.shape[0]):
                    raise ValueError('%s_mask must be [B, S]' % s)
                c.pad_mask[i] = mask.data_ptr()
            keep += [mem, mask]
        return c, keep

    # ------------------------------------------------------------------ reference API
    def forward(self, prev_target, contexts, incremental_state=None, need_attn=True):
        """ids [B,T] (or {'roberta': ids}) -> (X [B,T,d], {'attn': [B,T,S_article+2] | None}; the entity attention when
        the decoder has an 'entity' source, decoder_entity_pointer.py:326-334) --
        decoder_flattened.py:96-146.  With an `incremental_state` dict the call is one step of the reference's step-wise
        protocol (`_generate`, transformer_flattened.py:250-275): ids are the NEW tokens only and the state in the dict is
        advanced (see `_forward_incremental`).  `generate()` is the fast device-resident loop; this is the drop-in one."""
        ids = prev_target['roberta'] if isinstance(prev_target, dict) else prev_target
        ids = ids.to('cuda', torch.long).contiguous()
        if incremental_state is not None:
            return self._forward_incremental(ids, contexts, incremental_state, need_attn)
        if self.training and torch.is_grad_enabled() and getattr(self, 'flat_params', None) is not None:
            return self._forward_autograd(ids, contexts), {'attn': None, 'inner_states': None}
        B, T = ids.shape
        L, h = _lib.lib(), self._h()
        ctx, keep = self._contexts(contexts, B, T)
        S = (C.c_int * 4)(*[ctx.S[i] for i in range(4)])
        ws = self._workspace('fwd', L.tnt_decoder_workspace_bytes(h, B, T, S))
        X = torch.empty(B, T, self.d, dtype=self.torch_dtype, device='cuda')
        attn = None
        src = 'entity' if 'entity' in self.sources else 'article'    # the source whose attention the reference's layer returns
        if need_attn and src in self.sources:
            attn = torch.empty(B, T, ctx.S[self.sources.index(src)] + 2, dtype=torch.float32, device='cuda')
        _lib.check(L.tnt_decoder_forward(h, _ptr(ids), C.byref(ctx), _ptr(X), _ptr(attn), _ptr(ws), ws.numel(),
                                         _stream()), 'tnt_decoder_forward')
        return X, {'attn': attn, 'inner_states': None}

    # ------------------------------------------------------------------ step-wise protocol (incremental_state)
    def _forward_incremental(self, ids, contexts, state, need_attn=True):
        """One call of `decoder(prev_target, contexts_i, incremental_state)`.  State kept in the caller's dict, under the
        reference's key scheme so its filter / reorder callers keep working:
          'DynamicConv1dTBC.<id>.input_buffer'  [<= K-1, B, d] per layer -- the conv inputs of the previous steps, oldest
                                                first, exactly the reference's buffer (dynamic.py:95-99, 338-348);
          'SinusoidalPositionalEmbedding.<id>.position'  tokens consumed so far (positional.py:170-176);
          '<DecoderClass>.<id>.projected_kv'    this mirror's own: K/V of the contexts for all layers, projected ONCE (the
                                                reference re-projects the memories in every step); re-made whenever the
                                                batch size or the source lengths of `contexts` change."""
        B, T = ids.shape
        L, h = _lib.lib(), self._h()
        ctx, keep = self._contexts(contexts, B, T)
        S = (C.c_int * 4)(*[ctx.S[i] for i in range(4)])
        ns, Ls = len(self.sources), len(self.kernels)
        sig = (B, tuple(ctx.S[i] for i in range(ns)))
        kv = state.get(self._kv_key)
        if kv is None or kv['sig'] != sig:
            bufs = [torch.empty(ctx.S[i] + 2, B, Ls * 2 * self.d, dtype=self.torch_dtype, device='cuda') for i in range(ns)]
            ptrs = (C.c_void_p * 4)(*([b.data_ptr() for b in bufs] + [None] * (4 - ns)))
            _lib.check(L.tnt_decoder_project_kv(h, C.byref(ctx), ptrs, _stream()), 'tnt_decoder_project_kv')
            kv = {'sig': sig, 'bufs': bufs}
            state[self._kv_key] = kv
        start = int(state.get(self._pos_key, 0))
        hist_in, hist_len, hist_out, new_bufs = [], [], [], []
        for k, K in zip(self._conv_keys, self.kernels):
            buf = state.get(k)
            if buf is not None and buf.shape[1] != B:
                raise ValueError('incremental_state holds conv buffers for %d rows, the batch has %d: call '
                                 'filter_incremental_state / reorder_incremental_state first' % (buf.shape[1], B))
            n = 0 if buf is None else buf.shape[0]
            if buf is not None:
                buf = buf.to(self.torch_dtype).contiguous()
            out = torch.empty(min(K - 1, n + T), B, self.d, dtype=self.torch_dtype, device='cuda')
            hist_in.append(buf.data_ptr() if n else None)
            hist_len.append(n)
            hist_out.append(out.data_ptr() if out.numel() else None)
            new_bufs.append(out)
            keep.append(buf)
        ws = self._workspace('inc', L.tnt_decoder_incremental_workspace_bytes(h, B, T, S))
        X = torch.empty(B, T, self.d, dtype=self.torch_dtype, device='cuda')
        attn = None
        src = 'entity' if 'entity' in self.sources else 'article'
        if need_attn and src in self.sources:
            attn = torch.empty(B, T, ctx.S[self.sources.index(src)] + 2, dtype=torch.float32, device='cuda')
        kvp = (C.c_void_p * 4)(*([b.data_ptr() for b in kv['bufs']] + [None] * (4 - ns)))
        _lib.check(L.tnt_decoder_forward_incremental(
            h, _ptr(ids), C.byref(ctx), kvp, start, (C.c_void_p * Ls)(*hist_in), (C.c_int * Ls)(*hist_len),
            (C.c_void_p * Ls)(*hist_out), _ptr(X), _ptr(attn), _ptr(ws), ws.numel(), _stream()), 'tnt_decoder_forward_incremental')
        for k, out in zip(self._conv_keys, new_bufs):
            state[k] = out
        state[self._pos_key] = start + T
        return X, {'attn': attn, 'inner_states': None}

    def filter_incremental_state(self, incremental_state, active_idx):
        """decoder_flattened.py:202-207: keep the rows of `active_idx` (bool [B]) in every per-row state tensor -- the conv
        input buffers, as the reference does, and this mirror's projected K/V (skipped when every row stays)."""
        if incremental_state is None:
            return
        idx = None
        for key in list(incremental_state):
            if 'DynamicConv1dTBC' in key:
                incremental_state[key] = incremental_state[key][:, active_idx]
            elif key.endswith('.projected_kv'):
                kv = incremental_state[key]
                if idx is None:
                    idx = active_idx.nonzero().view(-1) if active_idx.dtype == torch.bool else active_idx
                if idx.numel() != kv['sig'][0]:
                    incremental_state[key] = {'sig': (idx.numel(), kv['sig'][1]),
                                              'bufs': [b.index_select(1, idx.to(b.device)) for b in kv['bufs']]}

    def reorder_incremental_state(self, incremental_state, new_order):
        """Beam search (dynamic.py:338-348 on every layer's buffer): state rows re-gathered in `new_order`."""
        if incremental_state is None:
            return
        for key in list(incremental_state):
            if 'DynamicConv1dTBC' in key:
                incremental_state[key] = incremental_state[key].index_select(1, new_order)
            elif key.endswith('.projected_kv'):
                kv = incremental_state[key]
                incremental_state[key] = {'sig': (new_order.numel(), kv['sig'][1]),
                                          'bufs': [b.index_select(1, new_order.to(b.device)) for b in kv['bufs']]}

    def nll(self, X, target):
        """AdaptiveSoftmax.forward + AdaptiveLoss(reduction='sum'): (loss_sum [1], n_non_pad) as device tensors."""
        L, h = _lib.lib(), self._h()
        X = X.to(self.torch_dtype).contiguous().view(-1, self.d)
        target = target.to('cuda', torch.long).contiguous().view(-1)
        n = X.shape[0]
        ws = self._workspace('loss', L.tnt_decoder_loss_workspace_bytes(h, n))
        out = torch.empty(2, dtype=torch.float32, device='cuda')
        _lib.check(L.tnt_decoder_nll(h, _ptr(X), _ptr(target), n, _ptr(out), _ptr(ws), ws.numel(), _stream()),
                   'tnt_decoder_nll')
        return out

    def get_normalized_probs(self, net_output, log_probs=True, sample=None):
        """decoder_flattened.py:187-200 -> AdaptiveSoftmax.get_log_prob (softmax.py:193-222)."""
        X = net_output[0]
        B, T, _ = X.shape
        L, h = _lib.lib(), self._h()
        Xc = X.to(self.torch_dtype).contiguous().view(-1, self.d)
        ws = self._workspace('loss', L.tnt_decoder_loss_workspace_bytes(h, B * T))
        out = torch.empty(B, T, self.vocab, dtype=torch.float32, device='cuda')
        _lib.check(L.tnt_decoder_log_prob(h, _ptr(Xc), B * T, _ptr(out), _ptr(ws), ws.numel(), _stream()),
                   'tnt_decoder_log_prob')
        return out if log_probs else out.exp_()

    def generate(self, seed, contexts, gen_len=100, eos=2, trim=True):
        """Greedy `_generate` (transformer_flattened.py:238-325): returns (log_probs [B,L'], ids [B,L'+1]);
        with trim=True the outputs stop at the step after which every row had produced EOS, as the
        reference's early `break` does."""
        seed = seed.to('cuda', torch.long)
        if seed.dim() == 2:
            seed = seed[:, -1]
        seed = seed.contiguous()
        B = seed.shape[0]
        L, h = _lib.lib(), self._h()
        ctx, keep = self._contexts(contexts, B, 1)
        S = (C.c_int * 4)(*[ctx.S[i] for i in range(4)])
        ws = self._workspace('gen', L.tnt_decoder_generate_workspace_bytes(h, B, gen_len, S))
        ids = torch.empty(B, gen_len + 1, dtype=torch.long, device='cuda')
        lp = torch.empty(B, gen_len, dtype=torch.float32, device='cuda')
        steps = C.c_int(0)
        _lib.check(L.tnt_decoder_generate(h, _ptr(seed), C.byref(ctx), gen_len, eos, _ptr(ids), _ptr(lp),
                                          C.byref(steps) if trim else None, _ptr(ws), ws.numel(), _stream()),
                   'tnt_decoder_generate')
        if trim:
            n = steps.value
            self.check_input_errors()
            return lp[:, :n], ids[:, :n + 1]
        return lp, ids

    def generate_beam(self, seed, contexts, gen_len=20, beam_size=5, eos=2):
        """Beam search with the scoring / bookkeeping of `CaptionModel.beam_search` (models/CaptionModel.py:23-122)
        over the incremental decoder; `reorder_incremental_state` happens on the device.  Returns, best first:
        (seq [B, beam, gen_len] -- 0 after a hypothesis ended, logps [B, beam, gen_len], p [B, beam])."""
        seed = seed.to('cuda', torch.long)
        if seed.dim() == 2:
            seed = seed[:, -1]
        seed = seed.contiguous()
        B = seed.shape[0]
        L, h = _lib.lib(), self._h()
        ctx, keep = self._contexts(contexts, B, 1)
        S = (C.c_int * 4)(*[ctx.S[i] for i in range(4)])
        nbytes = L.tnt_decoder_beam_workspace_bytes(h, B, beam_size, gen_len, S)
        if nbytes == 0:
            raise ValueError('generate_beam: beam_size must be in [1, 16] and gen_len in [1, 128]')
        ws = self._workspace('beam', nbytes)
        seq = torch.empty(B, beam_size, gen_len, dtype=torch.long, device='cuda')
        lps = torch.empty(B, beam_size, gen_len, dtype=torch.float32, device='cuda')
        p = torch.empty(B, beam_size, dtype=torch.float32, device='cuda')
        _lib.check(L.tnt_decoder_generate_beam(h, _ptr(seed), C.byref(ctx), gen_len, eos, beam_size, _ptr(seq), _ptr(lps),
                                               _ptr(p), _ptr(ws), ws.numel(), _stream()), 'tnt_decoder_generate_beam')
        return seq, lps, p

    # ------------------------------------------------------------------ training (kernel group 4)
    def tied_keys(self):
        """Output-softmax rows tied to the embedding bands (softmax.py:83-86,111-132): alias -> owner."""
        e = 'embedder.token_embedder_adaptive.embeddings.%d.0.weight'
        ties = {'adaptive_softmax.head.word_proj.weight': e % 0}
        for i in range(len(self.cutoffs) - 1):
            ties['adaptive_softmax.tail.%d.2.weight' % i] = e % (i + 1)
        return ties

    def enable_training(self):
        """Ties the output rows to the embedding bands (one Parameter, as in the reference), moves every
        trainable parameter into ONE flat fp32 arena with a matching flat gradient arena (`flat_params`,
        `flat_grads`; `.grad` of each parameter is a view), and registers the gradient buffers with the engine.
        The flat gradient arena is what data-parallel training all-reduces (one NCCL call per bucket)."""
        if getattr(self, 'flat_params', None) is not None:
            return self
        own = dict(self.named_parameters(remove_duplicate=False))
        for alias, owner in self.tied_keys().items():
            mod = self
            parts = alias.split('.')
            for q in parts[:-1]:
                mod = getattr(mod, q)
            mod._parameters[parts[-1]] = own[owner]
        named = [(k, p) for k, p in self.named_parameters() if 'token_embedder_position' not in k]
        # arena order = the order in which the backward pass finishes gradients (tnt_decoder_train_stages):
        # stage 0: softmax-only parameters AND the band tables (tied output rows) -- their dense softmax-side gradient is final
        # there; the <= B*T embedding rows added when the backward ends are exchanged as a packed row list in data-parallel
        # training (trainer.py; with the dense scatter the tables are final at the last stage) -- then layers last -> first,
        # then the band projections
        n_layers = len(self.kernels)
        self.band_tables_in_stage0 = True

        def stage_of(k):
            if k.startswith('adaptive_softmax.') or (k.startswith('embedder.token_embedder_adaptive.embeddings.') and
                                                     k.endswith('.0.weight')):
                return 0
            if k.startswith('layers.'):
                return n_layers - int(k.split('.')[1])
            return n_layers + 1
        named.sort(key=lambda kp: stage_of(kp[0]))
        total = sum((p.numel() + 63) // 64 * 64 for _, p in named)
        self.flat_params = torch.zeros(total, dtype=torch.float32, device='cuda')
        self.flat_grads = torch.zeros(total, dtype=torch.float32, device='cuda')
        self.param_slices = {}
        self.stage_slices = [[None, None] for _ in range(n_layers + 2)]
        off = 0
        L = _lib.lib()
        with torch.no_grad():
            for k, p in named:
                n = p.numel()
                view = self.flat_params[off:off + n].view(p.shape)
                view.copy_(p.data)
                p.data = view
                p.requires_grad_(True)
                p.grad = self.flat_grads[off:off + n].view(p.shape)
                self.param_slices[k] = (off, n)
                st = self.stage_slices[stage_of(k)]
                st[0] = off if st[0] is None else st[0]
                off += (n + 63) // 64 * 64
                st[1] = off
        self._dirty = True
        h = self._h()
        for k, p in named:
            _lib.check(L.tnt_decoder_set_grad(h, k.encode(), _ptr(p.grad)), 'tnt_decoder_set_grad ' + k)
        return self

    def band_table_grads(self):
        """fp32 gradient tensors of the vocabulary band tables, band order (views of `flat_grads`)."""
        e = 'embedder.token_embedder_adaptive.embeddings.%d.0.weight'
        own = dict(self.named_parameters())
        return [own[e % i].grad for i in range(len(self.cutoffs))]

    def mark_updated(self):
        """Call after the optimizer changed the parameters: the prepared (folded, cast) weights are rebuilt."""
        self._dirty = True

    def set_dropout(self, dropout=0.0, input_dropout=0.0, weight_dropout=0.0, attention_dropout=0.0, relu_dropout=0.0,
                    seed=0, rank=None):
        """Dropout of `train_step` (the constructor arguments of DynamicConvDecoder, decoder_flattened.py:24-31;
        config.yaml:51-69).  Masks are counter based: step k after this call uses seed + k, and every mask bit can
        be reproduced with ops.dropout_bits(seed + k, site, index).  All zero switches dropout off (parity mode)."""
        # data-parallel replicas must not drop the same positions on their different shards (the reference draws from a
        # per-process RNG): the rank (default: torch.distributed's) is mixed into the seed; rank 0 keeps `seed` itself
        if rank is None:
            import torch.distributed as dist
            rank = dist.get_rank() if dist.is_available() and dist.is_initialized() else 0
        seed = (int(seed) + int(rank) * 0x9E3779B97F4A7C15) & 0xFFFFFFFFFFFFFFFF
        cfg = _lib.DropoutConfig()
        cfg.dropout, cfg.input_dropout, cfg.weight_dropout = dropout, input_dropout, weight_dropout
        cfg.attention_dropout, cfg.relu_dropout, cfg.seed = attention_dropout, relu_dropout, seed
        _lib.check(_lib.lib().tnt_decoder_set_dropout(self._h(), C.byref(cfg)), 'tnt_decoder_set_dropout')

    def train_step(self, prev_target, target, contexts, loss_scale=1.0):
        """Forward + AdaptiveLoss + backward.  Gradients of `loss_scale * sum-NLL` are written into `.grad`
        of every parameter (views of `flat_grads`).  Returns a device tensor [loss_sum, n_non_pad]."""
        if getattr(self, 'flat_params', None) is None:
            self.enable_training()
        ids = prev_target['roberta'] if isinstance(prev_target, dict) else prev_target
        ids = ids.to('cuda', torch.long).contiguous()
        target = target.to('cuda', torch.long).contiguous()
        B, T = ids.shape
        L, h = _lib.lib(), self._h()
        self._target_grads('grad')
        ctx, keep = self._contexts(contexts, B, T)
        S = (C.c_int * 4)(*[ctx.S[i] for i in range(4)])
        ws = self._workspace('train', L.tnt_decoder_train_workspace_bytes(h, B, T, S))
        out = torch.empty(2, dtype=torch.float32, device='cuda')
        _lib.check(L.tnt_decoder_train_step(h, _ptr(ids), _ptr(target), C.byref(ctx), float(loss_scale), _ptr(out),
                                            _ptr(ws), ws.numel(), _stream()), 'tnt_decoder_train_step')
        return out

    # ------------------------------------------------------------------ autograd-connected training
    # The reference trainer does `loss.backward()` on what `model.forward` returns (callback_apex_trainer.py:229-233) after
    # `decoder_out = self.decoder(...)`, `loss, n = self.criterion(self.decoder.adaptive_softmax, decoder_out, target)`
    # (transformer_faces_objects.py:79-88).  After `enable_training()`, in train() mode with grad enabled, the same three
    # calls work here: forward() runs tnt_decoder_train_forward (activations kept in the engine workspace) and returns an X
    # that requires grad; AdaptiveLoss.forward() runs tnt_decoder_train_backward -- loss AND every parameter gradient for an
    # upstream gradient of 1, into a staging arena -- behind a torch.autograd.Function whose backward adds
    # upstream * staged into `.grad` (tnt_grad_accumulate), so loss scaling, gradient accumulation, zero_grad(set_to_none) and
    # any torch optimizer over `decoder.parameters()` behave as with the reference modules.
    def _target_grads(self, which):
        """Point the engine's gradient outputs at `.grad` (train_step / Trainer) or at the staging arena (autograd)."""
        if getattr(self, '_grad_target', 'grad') == which:
            return
        L, h = _lib.lib(), self._h()
        if which == 'stage' and getattr(self, 'flat_stage', None) is None:
            self.flat_stage = torch.zeros_like(self.flat_grads)
        base = self.flat_grads if which == 'grad' else self.flat_stage
        for k, (off, n) in self.param_slices.items():
            _lib.check(L.tnt_decoder_set_grad(h, k.encode(), C.c_void_p(base.data_ptr() + 4 * off)), 'tnt_decoder_set_grad ' + k)
        self._grad_target = which

    def _forward_autograd(self, ids, contexts):
        self._dirty = True                      # an optimizer may have stepped the parameters since the last call
        L, h = _lib.lib(), self._h()
        self._target_grads('stage')
        B, T = ids.shape
        ctx, keep = self._contexts(contexts, B, T)
        S = (C.c_int * 4)(*[ctx.S[i] for i in range(4)])
        ws = self._workspace('train', L.tnt_decoder_train_workspace_bytes(h, B, T, S))
        X = torch.empty(B, T, self.d, dtype=self.torch_dtype, device='cuda')
        _lib.check(L.tnt_decoder_train_forward(h, _ptr(ids), C.byref(ctx), _ptr(X), _ptr(ws), ws.numel(), _stream()),
                   'tnt_decoder_train_forward')
        self._pending = dict(ids=ids, ctx=ctx, keep=keep, ws=ws, X=X)
        X.requires_grad_(True)                  # so that the loss built on it is part of the autograd graph
        return X

    def _loss_autograd(self, X, target):
        """-> out2 [loss_sum, n_tokens]; the staging arena then holds d(loss_sum)/d(parameter)."""
        pend = self._pending
        if pend is None or X is not pend['X']:
            raise RuntimeError('AdaptiveLoss: in training mode the criterion must be given the tensor decoder.forward returned '
                               '(the fused engine backward runs from the activations of THAT forward call)')
        L, h = _lib.lib(), self._h()
        target = target.to('cuda', torch.long).contiguous()
        out = torch.empty(2, dtype=torch.float32, device='cuda')
        ws = pend['ws']
        _lib.check(L.tnt_decoder_train_backward(h, _ptr(pend['ids']), _ptr(target), C.byref(pend['ctx']), 1.0, _ptr(out), _ptr(ws),
                                                ws.numel(), _stream()), 'tnt_decoder_train_backward')
        self._pending = None
        return out

    def _accumulate_staged(self, upstream):
        """.grad += upstream * staged for every parameter (one kernel over the flat arenas)."""
        lost = [p for _, p in self._trainable() if p.grad is None]
        if lost:                                # optimizer.zero_grad(set_to_none=True): re-attach the arena views, zeroed
            self.flat_grads.zero_()
            for k, p in self._trainable():
                off, n = self.param_slices[k]
                p.grad = self.flat_grads[off:off + n].view(p.shape)
        g = upstream.reshape(-1)[:1].to(torch.float32).contiguous()
        _lib.check(_lib.lib().tnt_grad_accumulate(_ptr(self.flat_grads), _ptr(self.flat_stage), self.flat_grads.numel(), 1.0, _ptr(g),
                                                  _stream()), 'tnt_grad_accumulate')

    def _trainable(self):
        return [(k, p) for k, p in self.named_parameters() if k in self.param_slices]

    def score_host(self, ids_host, target_host, contexts_host):
        """End-to-end teacher-forced scoring from PINNED HOST buffers (ids/targets int64 [B,T], contexts in
        the engine dtype, masks bool): H2D copies + forward + AdaptiveLoss + D2H of (loss_sum, n_tokens),
        synchronous.  Returns (loss_sum, n_tokens) as Python floats."""
        L, h = _lib.lib(), self._h()
        B, T = ids_host.shape
        c = _lib.Contexts()
        c.B, c.T = B, T
        for i, s in enumerate(self.sources):
            mem = contexts_host[s]
            if mem.dtype != self.torch_dtype or not mem.is_contiguous() or mem.is_cuda:
                raise ValueError('host context %s must be a contiguous CPU tensor of dtype %s' % (s, self.torch_dtype))
            c.S[i] = mem.shape[0]
            c.memory[i] = mem.data_ptr()
            m = contexts_host.get(s + '_mask')
            c.pad_mask[i] = m.data_ptr() if m is not None else None
        S = (C.c_int * 4)(*[c.S[i] for i in range(4)])
        stg = self._workspace('stage', L.tnt_decoder_score_host_staging_bytes(h, B, T, S))
        ws = self._workspace('score', L.tnt_decoder_workspace_bytes(h, B, T, S) +
                             L.tnt_decoder_loss_workspace_bytes(h, B * T))
        if getattr(self, '_out2', None) is None:
            self._out2 = torch.zeros(2, dtype=torch.float32).pin_memory()
        _lib.check(L.tnt_decoder_score_host(h, _ptr(ids_host), _ptr(target_host), C.byref(c), _ptr(self._out2),
                                            _ptr(stg), stg.numel(), _ptr(ws), ws.numel(), _stream()),
                   'tnt_decoder_score_host')
        self.check_input_errors()
        return float(self._out2[0]), float(self._out2[1])

    def _host_contexts(self, ids_host, contexts_host):
        B, T = ids_host.shape
        c = _lib.Contexts()
        c.B, c.T = B, T
        for i, s in enumerate(self.sources):
            mem = contexts_host[s]
            if mem.dtype != self.torch_dtype or not mem.is_contiguous() or mem.is_cuda:
                raise ValueError('host context %s must be a contiguous CPU tensor of dtype %s' % (s, self.torch_dtype))
            c.S[i] = mem.shape[0]
            c.memory[i] = mem.data_ptr()
            m = contexts_host.get(s + '_mask')
            c.pad_mask[i] = m.data_ptr() if m is not None else None
        return c

    def score_host_submit(self, slot, ids_host, target_host, contexts_host):
        """Pipelined `score_host`: enqueues batch `slot` (0 or 1) -- its H2D copies overlap the compute of
        the other slot -- and returns immediately; `score_host_wait(slot)` returns (loss_sum, n_tokens)."""
        L, h = _lib.lib(), self._h()
        B, T = ids_host.shape
        c = self._host_contexts(ids_host, contexts_host)
        S = (C.c_int * 4)(*[c.S[i] for i in range(4)])
        stg = self._workspace('stage%d' % slot, L.tnt_decoder_score_host_staging_bytes(h, B, T, S))
        ws = self._workspace('score', L.tnt_decoder_workspace_bytes(h, B, T, S) +
                             L.tnt_decoder_loss_workspace_bytes(h, B * T))
        if getattr(self, '_out2s', None) is None:
            self._out2s = [torch.zeros(2, dtype=torch.float32).pin_memory() for _ in range(2)]
        _lib.check(L.tnt_decoder_score_host_submit(h, slot, _ptr(ids_host), _ptr(target_host), C.byref(c),
                                                   _ptr(self._out2s[slot]), _ptr(stg), stg.numel(), _ptr(ws),
                                                   ws.numel(), _stream()), 'tnt_decoder_score_host_submit')

    def score_host_wait(self, slot):
        _lib.check(_lib.lib().tnt_decoder_score_host_wait(self._h(), slot), 'tnt_decoder_score_host_wait')
        return float(self._out2s[slot][0]), float(self._out2s[slot][1])


class DynamicConvDecoderPointer(DynamicConvDecoder):
    """`DynamicConvDecoderPointer` (ttl/tell/models/decoder_pointer.py:23-297; expt/goodnews/h3_context_pointer): the
    2-source decoder plus a copy head -- p_gen = sigmoid(project_p_gens(cat(inp_embed, X))) mixes the vocabulary
    distribution with the last layer's article attention scattered onto the article's token ids.  Inference / scoring
    path; forward(prev_target, contexts, article_ids) -> (lprobs [B, T, V], {'attn', 'p_gens'})."""
    HOST_SIDE_KEYS = ('project_p_gens.',)

    def __init__(self, *args, **kwargs):
        super().__init__(*args, **kwargs)
        dev = next(self.parameters()).device
        _register(self, 'project_p_gens.bias', torch.zeros(1, device=dev))
        _register(self, 'project_p_gens.weight_g', torch.ones(1, 1, device=dev))
        _register(self, 'project_p_gens.weight_v', torch.zeros(1, 2 * self.d, device=dev))

    def _embed(self, ids):
        B, T = ids.shape
        emb = torch.empty(T * B, self.d, dtype=self.torch_dtype, device='cuda')
        _lib.check(_lib.lib().tnt_decoder_embed(self._h(), _ptr(ids), B, T, _ptr(emb), _stream()), 'tnt_decoder_embed')
        return emb

    def _p_gen_inputs(self, X, emb, B, T):
        """-> (x, ldx, inp_embed | None, lde, w_pgen fp32, bias, d, keep-alive list) of tnt_pointer_params."""
        from . import ops
        pg = self.project_p_gens
        w = ops.wn_fold(pg.weight_v.detach(), pg.weight_g.detach()).view(-1).contiguous()     # g * v / ||v||, fp32 [2d]
        return X, self.d, emb, self.d, w, float(pg.bias.item()), self.d, [w]

    def _mix(self, X, emb, attn, copy_ids, B, T):
        """decoder_pointer.py:151-216 output_layer, in place on the full-vocabulary log-probabilities."""
        L = _lib.lib()
        lp = self.get_normalized_probs((X, None), log_probs=True)                 # [B, T, V] fp32
        ids = copy_ids.to('cuda', torch.long).contiguous()
        if tuple(ids.shape) != (B, attn.shape[2] - 2):
            raise ValueError('the copied token ids must be [B, S] of the attended source (S = %d)' % (attn.shape[2] - 2))
        x, ldx, e, lde, w, bias, d, keep = self._p_gen_inputs(X, emb, B, T)
        p = _lib.PointerParams()
        p.x, p.ldx, p.inp_embed, p.lde = x.data_ptr(), ldx, (e.data_ptr() if e is not None else None), lde
        p.w_pgen, p.b_pgen = w.data_ptr(), bias
        p.attn, p.attn_ld, p.article_ids = attn.data_ptr(), attn.shape[2], ids.data_ptr()
        p_gen = torch.empty(B, T, 1, dtype=torch.float32, device='cuda')
        p.logprob, p.ldp, p.p_gen = lp.data_ptr(), self.vocab, p_gen.data_ptr()
        p.B, p.T, p.S, p.d, p.V, p.dtype = B, T, ids.shape[1], d, self.vocab, self.dtype_code
        _lib.check(L.tnt_pointer_mix(C.byref(p), _stream()), 'tnt_pointer_mix')
        return lp, p_gen

    def forward(self, prev_target, contexts, article_ids, incremental_state=None, need_attn=True):
        X, extra = super().forward(prev_target, contexts, incremental_state, need_attn=True)
        ids = prev_target['roberta'] if isinstance(prev_target, dict) else prev_target
        ids = ids.to('cuda', torch.long).contiguous()
        B, T = ids.shape
        attn = extra['attn']
        lp, p_gen = self._mix(X, self._embed(ids), attn, article_ids, B, T)
        return lp, {'attn': attn, 'p_gens': p_gen, 'inner_states': None}

    def nll(self, lprobs, target):
        """F.nll_loss(lprobs, target, ignore_index=pad, reduction='sum') (transformer_context_pointer.py:128):
        device tensor [loss_sum, n_non_pad]."""
        lp = lprobs.contiguous().view(-1, lprobs.shape[-1])
        tg = target.to('cuda', torch.long).contiguous().view(-1)
        out = torch.empty(2, dtype=torch.float32, device='cuda')
        _lib.check(_lib.lib().tnt_nll_rows(_ptr(lp), lp.shape[1], _ptr(tg), lp.shape[0], self.pad, _ptr(out), _stream()),
                   'tnt_nll_rows')
        return out

    # ------------------------------------------------------------------ training through the copy head (row f4)
    def _p_gen_autograd(self, emb, X):
        """p_gen [B, T] from autograd leaves emb / X [B, T, d] (fp32): sigmoid(project_p_gens(cat(inp_embed, X))),
        decoder_pointer.py:138-141, with the weight-norm fold of GehringLinear (linear.py:30-33) in the graph."""
        pg = self.project_p_gens
        w = pg.weight_v * (pg.weight_g / pg.weight_v.norm(dim=1, keepdim=True))           # [1, 2d]
        return torch.sigmoid(torch.cat((emb, X), 2) @ w.view(-1) + pg.bias)

    def train_step(self, prev_target, target, contexts, copy_ids, loss_scale=1.0):
        """Forward + the pointer models' loss F.nll_loss(log(mixed probs), target, ignore_index=pad, reduction='sum')
        (transformer_entity_pointer.py:123-131, transformer_context_pointer.py:128) + backward.  Gradients of
        loss_scale * loss are WRITTEN into `.grad` of every parameter (decoder, embedding bands, softmax, copy head).
        The decoder stack, the adaptive-softmax term (log-prob of the target without the [rows, V] matrix, weighted softmax
        gradient) and the attention backward -- incl. the gradient through the returned head-averaged weights -- run in the
        engine (tnt_decoder_train_forward_ext / _target_logprob / _train_backward_ext); the mix arithmetic on [rows] and
        [rows, S] tensors and the small p_gen head are chained by torch autograd.  Returns [loss_sum, n_non_pad]."""
        if getattr(self, 'flat_params', None) is None:
            self.enable_training()
        ids = prev_target['roberta'] if isinstance(prev_target, dict) else prev_target
        ids = ids.to('cuda', torch.long).contiguous()
        target = target.to('cuda', torch.long).contiguous()
        B, T = ids.shape
        L, h = _lib.lib(), self._h()
        self._target_grads('grad')
        ctx, keep = self._contexts(contexts, B, T)
        S = (C.c_int * 4)(*[ctx.S[i] for i in range(4)])
        src = 'entity' if 'entity' in self.sources else 'article'
        S_w = ctx.S[self.sources.index(src)]
        cids = copy_ids.to('cuda', torch.long).contiguous()
        if tuple(cids.shape) != (B, S_w):
            raise ValueError('the copied token ids must be [B, S] of the attended source (S = %d)' % S_w)
        ws = self._workspace('train', L.tnt_decoder_train_workspace_bytes(h, B, T, S))
        X = torch.empty(B, T, self.d, dtype=self.torch_dtype, device='cuda')
        emb = torch.empty(B, T, self.d, dtype=self.torch_dtype, device='cuda')
        attn = torch.empty(B, T, S_w + 2, dtype=torch.float32, device='cuda')
        lp = torch.empty(B, T, dtype=torch.float32, device='cuda')
        _lib.check(L.tnt_decoder_train_forward_ext(h, _ptr(ids), C.byref(ctx), _ptr(X), _ptr(attn), _ptr(emb), _ptr(ws), ws.numel(),
                                                   _stream()), 'tnt_decoder_train_forward_ext')
        _lib.check(L.tnt_decoder_target_logprob(h, _ptr(target), C.byref(ctx), _ptr(lp), _ptr(ws), ws.numel(), _stream()),
                   'tnt_decoder_target_logprob')
        head = [p for k, p in self.named_parameters() if k.startswith(self.HOST_SIDE_KEYS)]
        for p_ in head:
            p_.grad.zero_()                     # autograd accumulates into the arena views; the step WRITES gradients
        with torch.enable_grad():
            Xl, el = X.float().requires_grad_(), emb.float().requires_grad_()
            al, ll = attn.requires_grad_(), lp.requires_grad_()
            p_gen = self._p_gen_autograd(el, Xl)                                            # [B, T]
            # probability mass the copy distribution puts on the target: attention over the positions holding that token
            # (the 2 extra keys are dropped, decoder_pointer.py:191-196)
            hit = (cids[:, None, :] == target[:, :, None]).to(torch.float32)
            copy_mass = (al[:, :, :S_w] * hit).sum(-1)
            probs = p_gen * ll.exp() + (1.0 - p_gen) * copy_mass
            logp = probs.clamp(1e-10, 1.0).log()                                            # decoder_pointer.py:213
            nonpad = target != self.pad
            loss = -(logp * nonpad).sum()
            (loss * float(loss_scale)).backward()
        dX, de = Xl.grad.to(self.torch_dtype).contiguous(), el.grad.to(self.torch_dtype).contiguous()
        row_w = ll.grad.contiguous().view(-1)
        dattn = al.grad.contiguous()
        _lib.check(L.tnt_decoder_train_backward_ext(h, _ptr(ids), _ptr(target), C.byref(ctx), _ptr(row_w), _ptr(dX), _ptr(dattn),
                                                    _ptr(de), _ptr(ws), ws.numel(), _stream()), 'tnt_decoder_train_backward_ext')
        return torch.stack((loss.detach().float(), nonpad.sum().float()))

    def _head(self, copy_ids):
        """-> (tnt_pointer_head, keep-alive list) for the generation loop."""
        from . import ops
        pg = self.project_p_gens
        w = ops.wn_fold(pg.weight_v.detach(), pg.weight_g.detach()).view(-1).contiguous()
        hd = _lib.PointerHead()
        hd.w_pgen, hd.b_pgen = w.data_ptr(), float(pg.bias.item())
        hd.w_hidden, hd.b_hidden, hd.d_hidden = None, None, 0
        hd.copy_ids, hd.S_copy = copy_ids.data_ptr(), copy_ids.shape[1]
        return hd, [w]

    def generate(self, seed, contexts, copy_ids, gen_len=100, eos=2, trim=True):
        """Greedy `_generate` of the pointer models (sampling_topk = 1; transformer_entity_pointer.py:375-470): every
        step runs the incremental decoder, the full-vocabulary log-probabilities, the copy mix and the arg-max of the
        mixed row on the device.  Returns (log_probs [B, L'], ids [B, L'+1]) as DynamicConvDecoder.generate."""
        seed = seed.to('cuda', torch.long)
        if seed.dim() == 2:
            seed = seed[:, -1]
        seed = seed.contiguous()
        B = seed.shape[0]
        L, h = _lib.lib(), self._h()
        ctx, keep = self._contexts(contexts, B, 1)
        S = (C.c_int * 4)(*[ctx.S[i] for i in range(4)])
        cids = copy_ids.to('cuda', torch.long).contiguous()
        hd, keep2 = self._head(cids)
        ws = self._workspace('gen_ptr', L.tnt_decoder_generate_pointer_workspace_bytes(h, B, gen_len, S, hd.d_hidden))
        ids = torch.empty(B, gen_len + 1, dtype=torch.long, device='cuda')
        lp = torch.empty(B, gen_len, dtype=torch.float32, device='cuda')
        steps = C.c_int(0)
        _lib.check(L.tnt_decoder_generate_pointer(h, _ptr(seed), C.byref(ctx), C.byref(hd), gen_len, eos, _ptr(ids),
                                                  _ptr(lp), C.byref(steps) if trim else None, _ptr(ws), ws.numel(),
                                                  _stream()), 'tnt_decoder_generate_pointer')
        if trim:
            n = steps.value
            return lp[:, :n], ids[:, :n + 1]
        return lp, ids

    def generate_beam(self, *a, **k):
        raise NotImplementedError('%s: beam search through the copy head is not implemented' % type(self).__name__)


class DynamicConvDecoderEntityPointer(DynamicConvDecoderPointer):
    """`DynamicConvDecoderEntityPointer` + the copy head of `TransformerEntityPointerModel` (ttl/tell/models/
    decoder_entity_pointer.py:22-388, transformer_entity_pointer.py:78-79,185-261; the experiment ttl/run.sh trains):
    a 3-source decoder (image, article, entity) whose last layer returns the ENTITY attention, and
    p_gen = sigmoid(project_p_gens(relu(project_first_p_gens(cat(inp_embed, X))))) mixes the vocabulary distribution
    with that attention scattered onto the entity token ids.
    forward(prev_target, contexts, entity_tokens) -> (lprobs [B, T, V], {'attn', 'p_gens', 'predictors'})."""
    HOST_SIDE_KEYS = ('project_p_gens.', 'project_first_p_gens.')

    def __init__(self, *args, sources=('image', 'article', 'entity'), hidden=1024, **kwargs):
        DynamicConvDecoder.__init__(self, *args, sources=sources, **kwargs)
        if 'entity' not in self.sources:
            raise ValueError('DynamicConvDecoderEntityPointer needs an "entity" source')
        dev = next(self.parameters()).device
        self.hidden = hidden
        _register(self, 'project_first_p_gens.bias', torch.zeros(hidden, device=dev))
        _register(self, 'project_first_p_gens.weight_g', torch.ones(hidden, 1, device=dev))
        _register(self, 'project_first_p_gens.weight_v', torch.zeros(hidden, 2 * self.d, device=dev))
        _register(self, 'project_p_gens.bias', torch.zeros(1, device=dev))
        _register(self, 'project_p_gens.weight_g', torch.ones(1, 1, device=dev))
        _register(self, 'project_p_gens.weight_v', torch.zeros(1, hidden, device=dev))

    def _first(self):
        from . import ops
        f = self.project_first_p_gens
        w1 = ops.wn_fold(f.weight_v.detach(), f.weight_g.detach(), dtype=self.torch_dtype)       # [hidden, 2d]
        return w1, f.bias.detach().float().contiguous()

    def _p_gen_inputs(self, X, emb, B, T):
        from . import ops
        # predictors = cat(inp_embed, X) (decoder_entity_pointer.py:144); hidden = relu(project_first_p_gens(.))
        pred = torch.cat((emb.view(T, B, self.d).transpose(0, 1), X), 2).contiguous().view(B * T, 2 * self.d)
        w1, b1 = self._first()
        hid = ops.linear(pred, w1, b1, epilogue=ops.EPI_RELU)
        pg = self.project_p_gens
        w2 = ops.wn_fold(pg.weight_v.detach(), pg.weight_g.detach()).view(-1).contiguous()        # fp32 [hidden]
        self._predictors = pred.view(B, T, 2 * self.d)
        return hid, self.hidden, None, 0, w2, float(pg.bias.item()), self.hidden, [hid, w2]

    def forward(self, prev_target, contexts, entity_tokens, incremental_state=None, need_attn=True):
        lp, extra = super().forward(prev_target, contexts, entity_tokens, incremental_state, need_attn)
        extra['predictors'] = self._predictors
        return lp, extra

    def _p_gen_autograd(self, emb, X):
        """sigmoid(project_p_gens(relu(project_first_p_gens(cat(inp_embed, X))))) (transformer_entity_pointer.py:198-200): the
        hidden layer is a tcgen05 GEMM forward and backward (modules.LinearFn), weight-norm folds stay in the graph."""
        from .modules import linear
        f, pg = self.project_first_p_gens, self.project_p_gens
        w1 = f.weight_v * (f.weight_g / f.weight_v.norm(dim=1, keepdim=True))            # [hidden, 2d]
        pred = torch.cat((emb, X), 2).to(self.torch_dtype)
        hid = torch.relu(linear(pred, w1, f.bias).float())
        w2 = pg.weight_v * (pg.weight_g / pg.weight_v.norm(dim=1, keepdim=True))          # [1, hidden]
        return torch.sigmoid(hid @ w2.view(-1) + pg.bias)

    def _head(self, copy_ids):
        hd, keep = super()._head(copy_ids)
        w1, b1 = self._first()
        hd.w_hidden, hd.b_hidden, hd.d_hidden = w1.data_ptr(), b1.data_ptr(), self.hidden
        return hd, keep + [w1, b1]


class AdaptiveLoss(nn.Module):
    """criteria/adaptive_loss.py:10-91: forward(adaptive_softmax_owner, net_output, target, reduction='sum')
    -> (loss [1], sample_size).  `adaptive_softmax_owner` is the engine decoder (it owns the tied head)."""

    def __init__(self, padding_idx=1):
        super().__init__()
        self.padding_idx = padding_idx

    def forward(self, decoder, net_output, decoder_target, reduction='sum'):
        if reduction != 'sum':
            raise NotImplementedError('reduction=%r' % reduction)
        if getattr(decoder, '_pending', None) is not None and net_output[0].requires_grad:
            loss, cnt = _EngineLossFn.apply(net_output[0], decoder, decoder_target)     # training: loss.backward() works
            n = int(cnt.item())
            decoder.check_input_errors()
            return loss, n
        out = decoder.nll(net_output[0], decoder_target)
        n = int(out[1].item())
        decoder.check_input_errors()
        return out[:1], n


class _EngineLossFn(torch.autograd.Function):
    """AdaptiveSoftmax + AdaptiveLoss + the WHOLE decoder backward as one autograd node (see
    DynamicConvDecoder._forward_autograd): forward = tnt_decoder_train_backward with an upstream gradient of 1 (loss and
    parameter gradients in one fused pass), backward = `.grad += upstream * staged`."""

    @staticmethod
    def forward(ctx, X, decoder, target):
        out = decoder._loss_autograd(X, target)
        ctx.decoder = decoder
        cnt = out[1:2].clone()
        ctx.mark_non_differentiable(cnt)
        return out[:1].clone(), cnt

    @staticmethod
    def backward(ctx, g_loss, _g_cnt):
        ctx.decoder._accumulate_staged(g_loss)
        return None, None, None


def sinusoid_table(n, d, padding_idx=1):
    """tensor2tensor sinusoids [sin | cos] with the padding row zeroed --
    ttl/tell/modules/token_embedders/positional.py:125-165 (the reference builds this buffer at init)."""
    half = d // 2
    inc = math.log(10000.0) / (half - 1)
    inv = torch.exp(torch.arange(half, dtype=torch.float) * -inc)
    ang = torch.arange(n, dtype=torch.float)[:, None] * inv[None, :]
    tab = torch.cat([ang.sin(), ang.cos()], 1)
    if d % 2 == 1:
        tab = torch.cat([tab, torch.zeros(n, 1)], 1)
    tab[padding_idx] = 0
    return tab


def random_state_dict(d=1024, heads=16, ffn=4096, kernels=(3, 7, 15, 31), vocab=50265, cutoffs=(5000, 20000, 50265),
                      sources=('image', 'article'), source_dim=None, seed=0, n_pos=513):
    """Random-init weights with the reference's state_dict layout (synthetic benchmark / test weights;
    no checkpoint exists offline).  Scales keep activations O(1); tied output rows alias the embedding
    bands as in the reference (softmax.py:83-86,111-132)."""
    sdim = dict(DEFAULT_SOURCE_DIM)
    sdim.update(source_dim or {})
    g = torch.Generator().manual_seed(seed)
    sd = {}
    for k, shp in reference_shapes(d, heads, ffn, list(kernels), vocab, list(cutoffs), list(sources), sdim,
                                   n_pos).items():
        if k.endswith('weight_g'):
            sd[k] = torch.rand(shp, generator=g) + 0.5
        elif (k.endswith('layer_norm.weight') or ('_lns.' in k and k.endswith('weight'))):
            sd[k] = 1 + 0.1 * torch.randn(shp, generator=g)
        elif len(shp) == 1 or 'bias' in k:
            sd[k] = 0.1 * torch.randn(shp, generator=g)
        elif 'token_embedder_position' in k:
            sd[k] = sinusoid_table(shp[0], shp[1])
        else:
            sd[k] = torch.randn(shp, generator=g) * (shp[-1] ** -0.5)
    sd['adaptive_softmax.head.word_proj.weight'] = sd['embedder.token_embedder_adaptive.embeddings.0.0.weight']
    for i in range(len(cutoffs) - 1):
        sd['adaptive_softmax.tail.%d.2.weight' % i] = sd['embedder.token_embedder_adaptive.embeddings.%d.0.weight' % (i + 1)]
    return sd
