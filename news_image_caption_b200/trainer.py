"""Data-parallel training driver for the B200 decoder engine (BASELINE.json config 3).

One process per GPU; every rank holds a full replica and a shard of the batch.  A step is
  engine forward + AdaptiveLoss + backward (tnt_decoder_train_step, all CUDA kernels of this repo)
  -> the gradient exchange (torch.distributed / NCCL over NVLink: the ONLY collective of the path),
     stage by stage on a communication stream that overlaps the rest of the backward pass:
       * every stage's slice of the flat fp32 gradient arena is cast to bf16 (tnt_cast), sum-all-reduced in bf16 -- half the
         wire bytes of the fp32 arena -- and cast back into the fp32 arena BertAdam reads;
       * the embedding bands are tied to the softmax output rows, so their dense (softmax-side) gradient is final at stage 0
         and is exchanged FIRST, under the whole backward; the <= B*T embedding rows that only exist when the backward ends
         travel as a packed row list (tnt_decoder_set_sparse_embed_grad: 4 MB per rank in bf16, one all-gather) that every
         rank scatters into its table gradient (tnt_embed_grad_scatter) -- instead of a second, un-overlappable pass over
         the 206 MB tables;
       * a 2-scalar all-reduce of (loss_sum, n_tokens): the reference normalises the loss by the GLOBAL non-pad token count
         (criteria/adaptive_loss.py:62-65, transformer_faces_objects.py:85-88)
  -> fused BertAdam over the flat arenas (tnt_bert_adam_step; config.yaml:126-136).
torch supplies memory, streams and the process group only."""
import ctypes as C
import math

import torch
import torch.distributed as dist

from . import _lib
from ._lib import TNT_BF16, TNT_F32


def warmup_linear(x, warmup):
    """pytorch_pretrained_bert.optimization.warmup_linear (the schedule named in config.yaml:131)."""
    if x < warmup:
        return x / warmup
    return max((x - 1.0) / (warmup - 1.0), 0.0)


# measured (2 GPUs, B=64): casting on a FEW CTAs removes the 20-50 us stalls a full-width cast causes at every stage, but the
# slow casts stretch the exchange chain and compete with the backward for longer: 9.17 ms (full width) vs 9.72 / 10.34 /
# 10.75 ms at 48 / 24 / 8 CTAs.  Full width stays the default (0); the compute stream's PRIORITY is what keeps the
# backward ahead of the exchange (Trainer.__init__).
CAST_CTAS = int(__import__('os').environ.get('TNT_EXCHANGE_CAST_CTAS', '0'))


def _cast(src, dst):
    """dst <- src across fp32 / bf16 on the current stream: the library's own kernel on the GPU (its background form -- a few
    CTAs -- because the exchange runs beside the backward pass), torch on the CPU (gloo tests)."""
    if src.is_cuda:
        code = {torch.float32: TNT_F32, torch.bfloat16: TNT_BF16}
        st = C.c_void_p(torch.cuda.current_stream().cuda_stream)
        ctas = CAST_CTAS if (CAST_CTAS > 0 and src.numel() > (4 << 20)) else 1 << 20
        _lib.check(_lib.lib().tnt_cast_background(C.c_void_p(src.data_ptr()), C.c_void_p(dst.data_ptr()), src.numel(),
                                                  code[src.dtype], code[dst.dtype], ctas, st), 'tnt_cast_background')
    else:
        dst.copy_(src)


class GradientExchange:
    """Sum all-reduce of slices of one flat fp32 gradient buffer, on the wire in `wire` precision ('bf16' halves the bytes;
    'fp32' is exact).  Pure torch.distributed plumbing: works with CPU tensors over gloo (the world_size-2 CPU tests) and
    with CUDA tensors over NCCL."""

    def __init__(self, flat_grads, slices, group=None, wire='bf16'):
        self.flat, self.slices, self.group = flat_grads, [tuple(s) for s in slices], group
        self.world = dist.get_world_size(group) if dist.is_initialized() else 1
        if wire not in ('bf16', 'fp32'):
            raise ValueError('wire must be "bf16" or "fp32"')
        self.wire = wire
        self.wire_buf = torch.empty_like(flat_grads, dtype=torch.bfloat16) if (wire == 'bf16' and self.world > 1) else None
        self.pending = []

    def bytes_on_wire(self):
        n = sum(b - a for a, b in self.slices if a is not None and b > a)
        return n * (2 if self.wire == 'bf16' else 4)

    def reduce_slice(self, i, async_op=True):
        a, b = self.slices[i]
        if self.world == 1 or a is None or b <= a:
            return
        if self.wire_buf is not None:
            # cast -> all-reduce -> cast back, all in order on the calling (communication) stream: the next stage only
            # becomes ready ~1 ms later, so nothing queues up, and only the LAST stage's 3 small ops trail the backward
            # (deferring the cast-backs to finish() left 0.19 ms of them serialised after the last collective)
            _cast(self.flat[a:b], self.wire_buf[a:b])
            w = dist.all_reduce(self.wire_buf[a:b], op=dist.ReduceOp.SUM, group=self.group, async_op=async_op)
            if async_op:
                w.wait()                # stream-side wait (NCCL) / host wait (gloo)
            _cast(self.wire_buf[a:b], self.flat[a:b])
            return
        w = dist.all_reduce(self.flat[a:b], op=dist.ReduceOp.SUM, group=self.group, async_op=async_op)
        if async_op:
            w.wait()                    # stream-side: what follows on the calling stream (the slice's norm pass) sees the sum

    def reduce_scalars(self, t):
        if self.world > 1:
            dist.all_reduce(t, op=dist.ReduceOp.SUM, group=self.group)
        return t

    def finish(self):
        for w, a, b in self.pending:
            w.wait()
        self.pending = []


class SparseRowExchange:
    """All-gather of every rank's packed embedding-gradient rows (rows [N, d] + meta [N] int64 + count, one byte buffer per
    rank, one collective) and their scatter into the dense fp32 band-table gradients."""

    def __init__(self, n_rows, d, dtype, device, group=None):
        self.n, self.d, self.dtype, self.group = n_rows, d, dtype, group
        self.world = dist.get_world_size(group) if dist.is_initialized() else 1
        es = 2 if dtype == torch.bfloat16 else 4
        self.rows_bytes, self.meta_bytes = n_rows * d * es, n_rows * 8
        self.bytes = self.rows_bytes + self.meta_bytes + 64
        self.mine = torch.zeros(self.bytes, dtype=torch.uint8, device=device)
        self.all = torch.zeros(self.world * self.bytes, dtype=torch.uint8, device=device)

    def views(self, buf):
        rows = buf[:self.rows_bytes].view(self.dtype).view(self.n, self.d)
        meta = buf[self.rows_bytes:self.rows_bytes + self.meta_bytes].view(torch.int64)
        count = buf[self.rows_bytes + self.meta_bytes:self.rows_bytes + self.meta_bytes + 4].view(torch.int32)
        return rows, meta, count

    def gather(self, async_op=False):
        if self.world == 1:
            self.all.copy_(self.mine)
            return None
        return dist.all_gather_into_tensor(self.all, self.mine, group=self.group, async_op=async_op)

    def scatter_into(self, band_grads, skip_row):
        """band_grads: list of fp32 [rows_i, d] gradient tensors (views of the flat arena), one per vocabulary band."""
        for r in range(self.world):
            rows, meta, count = self.views(self.all[r * self.bytes:(r + 1) * self.bytes])
            if rows.is_cuda:
                ptrs = (C.c_void_p * len(band_grads))(*[g.data_ptr() for g in band_grads])
                dims = (C.c_int * len(band_grads))(*[g.shape[1] for g in band_grads])
                code = TNT_BF16 if self.dtype == torch.bfloat16 else TNT_F32
                st = C.c_void_p(torch.cuda.current_stream().cuda_stream)
                _lib.check(_lib.lib().tnt_embed_grad_scatter(C.c_void_p(rows.data_ptr()), self.d, C.c_void_p(meta.data_ptr()),
                                                             C.c_void_p(count.data_ptr()), self.n, ptrs, dims, len(band_grads),
                                                             skip_row, code, st), 'tnt_embed_grad_scatter')
            else:                                   # CPU restatement of the kernel for the gloo tests
                n = int(count[0])
                band, loc = meta[:n] >> 32, meta[:n] & 0xffffffff
                for i, g in enumerate(band_grads):
                    m = (band == i) & (loc != skip_row)
                    g.index_add_(0, loc[m], rows[:n][m].float())


class Trainer:
    def __init__(self, decoder, lr=1e-4, betas=(0.9, 0.98), eps=1e-6, weight_decay=1e-5, max_grad_norm=0.1,
                 warmup=0.05, t_total=-1, group=None, overlap=True, wire='bf16', sparse_embed=True):
        self.dec = decoder.enable_training()
        self.lr, self.betas, self.eps, self.wd, self.max_norm = lr, betas, eps, weight_decay, max_grad_norm
        self.warmup, self.t_total, self.step_no = warmup, t_total, 0
        self.overlap = overlap
        n = self.dec.flat_params.numel()
        self.m = torch.zeros(n, dtype=torch.float32, device='cuda')
        self.v = torch.zeros(n, dtype=torch.float32, device='cuda')
        seg = torch.zeros(n // 64, dtype=torch.int32)
        self.n_segments = len(self.dec.param_slices)
        for i, (k, (off, cnt)) in enumerate(self.dec.param_slices.items()):
            seg[off // 64:(off + cnt + 63) // 64] = i
        self.seg_of_chunk = seg.cuda()
        self.seg_sq = torch.zeros(self.n_segments, dtype=torch.float32, device='cuda')
        self.exchange = GradientExchange(self.dec.flat_grads, self.dec.stage_slices, group, wire=wire)
        self.group = group
        self.comm_stream = torch.cuda.Stream() if self.exchange.world > 1 else None
        # multi-GPU: the step itself runs on a HIGH-priority stream, the exchange (casts on comm_stream, NCCL on its own
        # stream) at the default, lower priority: when SMs free up the block scheduler hands them to the backward pass first,
        # so the exchange fills in behind it instead of delaying it
        self.compute_stream = None
        if self.exchange.world > 1 and __import__('os').environ.get('TNT_COMPUTE_PRIORITY', '1') != '0':
            self.compute_stream = torch.cuda.Stream(priority=-1)
        self.gscale = torch.ones(1, dtype=torch.float32, device='cuda')
        # the per-tensor gradient norms BertAdam clips by can be summed slice by slice as soon as a stage's gradients are
        # final (tnt_grad_sqnorm): data-parallel on the communication stream right after the slice's cast-back, so only the
        # slices that are final last stay in the step's tail.  On ONE GPU (side stream beside the backward pass) it measured
        # slightly slower than the single pass before BertAdam (7.61 vs 7.58 ms: the 0.1 ms HBM pass competes with the
        # backward and costs two stream hops per stage), so it is off there unless TNT_STAGE_NORMS=1
        env = __import__('os').environ.get('TNT_STAGE_NORMS')
        self.stage_norms = max_grad_norm > 0 and (env == '1' or (env is None and self.exchange.world > 1))
        self.norm_stream = torch.cuda.Stream() if (self.stage_norms and self.exchange.world == 1) else None
        # BertAdam clips PER TENSOR, so a stage's tensors can be stepped as soon as that stage's gradients (and norms) are
        # final: slice by slice behind the stage events -- on the communication stream right behind the slice's all-reduce
        # (or on a side stream on one GPU).  What stays in the step's tail: the slices that are only final when the backward
        # ends (the tied band tables: a third of the parameters).  Measured: 2 GPUs 7.05 -> 6.97 ms per step; ONE GPU 6.39 ->
        # 6.39..6.44 ms (the HBM-bound update slows the backward it overlaps by as much as it saves; with a small grid --
        # TNT_ADAM_BG_CTAS -- it becomes the tail itself: 6.69 ms at 148 CTAs), so it is on for world > 1 only
        # (TNT_STAGE_ADAM=0 / 1 overrides).
        env_sa = __import__('os').environ.get('TNT_STAGE_ADAM')
        self.stage_adam = overlap and (env_sa == '1' or (env_sa is None and self.exchange.world > 1))
        self.bg_ctas = int(__import__('os').environ.get('TNT_ADAM_BG_CTAS', '0'))
        if self.stage_adam and self.exchange.world == 1:
            self.stage_norms = max_grad_norm > 0
            self.norm_stream = torch.cuda.Stream()
        # multi-GPU: the embedding rows of the tied band tables travel as a packed list (the arena keeps the band tables in
        # stage 0, where their dense softmax part is final); a band whose dim differs from d keeps the dense path
        self.sparse = None
        self.sparse_embed = bool(sparse_embed and self.exchange.world > 1 and self.dec.band_tables_in_stage0)

    def exchange_description(self):
        ex = self.exchange
        return ('%s sum all-reduce of the %.0f MB flat gradient arena in %d stages on a comm stream (%s wire, %.0f MB per step)%s '
                '+ 2-scalar all-reduce (loss_sum, n_tokens); world %d' % (
                    'NCCL' if self.dec.flat_grads.is_cuda else 'gloo', self.dec.flat_grads.numel() * 4 / 1e6, len(ex.slices), ex.wire,
                    ex.bytes_on_wire() / 1e6,
                    ' + all-gather of the packed embedding-row gradients (%.1f MB per rank)' % (self.sparse.bytes / 1e6) if self.sparse else '',
                    ex.world))

    def lr_now(self):
        if self.t_total > 0:
            return self.lr * warmup_linear(self.step_no / self.t_total, self.warmup)
        return self.lr

    def _ensure_sparse(self, n_rows):
        if not self.sparse_embed:
            return
        if self.sparse is None or self.sparse.n != n_rows:
            dec = self.dec
            self.sparse = SparseRowExchange(n_rows, dec.d, dec.torch_dtype, 'cuda', self.group)
            rows, meta, count = self.sparse.views(self.sparse.mine)
            _lib.check(_lib.lib().tnt_decoder_set_sparse_embed_grad(dec._h(), C.c_void_p(rows.data_ptr()), C.c_void_p(meta.data_ptr()),
                                                                    C.c_void_p(count.data_ptr()), n_rows),
                       'tnt_decoder_set_sparse_embed_grad')

    def step(self, prev_target, target, contexts):
        """One optimisation step on this rank's shard.  Returns the GLOBAL [loss_sum (nats), n_tokens] tensor
        (device); the reported loss of the reference is loss_sum / n_tokens / ln 2 (bits per token)."""
        if self.compute_stream is not None and torch.cuda.current_stream() != self.compute_stream:
            caller = torch.cuda.current_stream()
            self.compute_stream.wait_stream(caller)
            with torch.cuda.stream(self.compute_stream):
                out = self.step(prev_target, target, contexts)
            caller.wait_stream(self.compute_stream)
            for t in list(contexts.values()) + [out]:
                if torch.is_tensor(t) and t.is_cuda:
                    t.record_stream(self.compute_stream)
            return out
        dec, L = self.dec, _lib.lib()
        ids = prev_target['roberta'] if isinstance(prev_target, dict) else prev_target
        self._ensure_sparse(ids.shape[0] * ids.shape[1])
        cur = torch.cuda.current_stream()
        stage_norms = self.stage_norms and (self.exchange.world == 1 or self.overlap)
        if stage_norms:
            self.seg_sq.zero_()                 # on the step's stream: after the previous step's BertAdam read it
        out = dec.train_step(prev_target, target, contexts, loss_scale=1.0)
        h = dec._h()
        n_stages = L.tnt_decoder_train_stages(h)
        if not self.sparse_embed and dec.band_tables_in_stage0:
            # dense embedding scatter: the band tables (arena stage 0) are only final when the backward ends
            order = list(range(1, n_stages)) + [0]
            wait_on = {0: n_stages - 1}
        else:
            order, wait_on = list(range(n_stages)), {}
        late = []                               # slices whose norms can only be summed on the step's stream, at the end

        def sqnorm(s, stream, max_ctas=0):
            a, b = self.exchange.slices[s]
            if a is not None and b > a:
                _lib.check(L.tnt_grad_sqnorm(C.c_void_p(dec.flat_grads.data_ptr()), C.c_void_p(self.seg_of_chunk.data_ptr()),
                                             C.c_void_p(self.seg_sq.data_ptr()), a, b - a, max_ctas, C.c_void_p(stream.cuda_stream)),
                           'tnt_grad_sqnorm')
        done = set()                            # slices BertAdam has already stepped (stage-wise)
        ln2 = math.log(2.0)

        def gscale_now():
            # gradient of the reference loss (sum-NLL / n_tokens / ln 2): scale by the global token count on the device
            # (an all-padding global batch has n_tokens == 0: every gradient is 0 then, keep the scale finite)
            torch.reciprocal(out[1:2].clamp_min(1.0) * ln2, out=self.gscale)

        bg = self.bg_ctas                       # grid cap of the passes that run beside the backward

        def adam(stream, s=None, max_ctas=0):
            a = _lib.AdamParams()
            a.max_ctas = max_ctas
            lo, n = 0, dec.flat_params.numel()
            if s is not None:
                lo, hi = self.exchange.slices[s]
                if lo is None or hi <= lo:
                    return
                n = hi - lo
            a.params, a.grads = dec.flat_params.data_ptr() + 4 * lo, dec.flat_grads.data_ptr() + 4 * lo
            a.m, a.v = self.m.data_ptr() + 4 * lo, self.v.data_ptr() + 4 * lo
            a.n = n
            a.seg_of_chunk, a.seg_sq, a.n_segments = self.seg_of_chunk.data_ptr() + 4 * (lo // 64), self.seg_sq.data_ptr(), self.n_segments
            a.grad_scale_dev, a.grad_scale = self.gscale.data_ptr(), 1.0
            a.max_grad_norm, a.lr = self.max_norm, self.lr_now()
            a.beta1, a.beta2, a.eps, a.weight_decay = self.betas[0], self.betas[1], self.eps, self.wd
            a.norms_ready = 1 if (stage_norms or self.max_norm <= 0) else 0
            _lib.check(L.tnt_bert_adam_step(C.byref(a), C.c_void_p(stream.cuda_stream)), 'tnt_bert_adam_step')

        if self.exchange.world == 1 and (stage_norms or self.stage_adam):
            if self.norm_stream is None:
                self.norm_stream = torch.cuda.Stream()
            ns = self.norm_stream
            out.record_stream(ns)
            for j, s in enumerate(order):       # the stage events order this after seg_sq.zero_() too
                _lib.check(L.tnt_decoder_stream_wait_stage(h, wait_on.get(s, s), C.c_void_p(ns.cuda_stream)),
                           'tnt_decoder_stream_wait_stage')
                last = j == len(order) - 1      # nothing left to overlap with: full width
                if stage_norms:
                    sqnorm(s, ns, 0 if last else bg)
                if self.stage_adam:
                    if j == 0:
                        with torch.cuda.stream(ns):
                            gscale_now()        # (loss_sum, n_tokens) exist before any stage event
                    adam(ns, s, 0 if last else bg)
                    done.add(s)
            cur.wait_stream(ns)
        if self.exchange.world > 1:
            if self.overlap:
                with torch.cuda.stream(self.comm_stream):
                    for j, s in enumerate(order):
                        _lib.check(L.tnt_decoder_stream_wait_stage(h, wait_on.get(s, s), C.c_void_p(self.comm_stream.cuda_stream)),
                                   'tnt_decoder_stream_wait_stage')
                        if j == 0:      # (loss_sum, n_tokens) exist once the forward has ended: exchanged here, off the step's tail
                            self.exchange.reduce_scalars(out)
                            if self.stage_adam:
                                out.record_stream(self.comm_stream)
                                gscale_now()
                        self.exchange.reduce_slice(s)
                        if s == 0 and self.sparse is not None:
                            late.append(s)              # the band tables still get the gathered embedding rows
                            continue
                        if stage_norms:
                            sqnorm(s, self.comm_stream, bg if self.stage_adam else 0)
                        if self.stage_adam and (stage_norms or self.max_norm <= 0):
                            adam(self.comm_stream, s, bg)
                            done.add(s)
                    if self.sparse is not None:
                        self.sparse.gather()
                    self.exchange.finish()
                cur.wait_stream(self.comm_stream)
            else:
                self.exchange.reduce_scalars(out)
                for s in order:
                    self.exchange.reduce_slice(s, async_op=False)
                if self.sparse is not None:
                    self.sparse.gather()
            if self.sparse is not None:
                self.sparse.scatter_into(dec.band_table_grads(), dec.embed_padding_idx)
            for s in late:
                if stage_norms:
                    sqnorm(s, cur)
        if not done:
            gscale_now()
            adam(cur)                           # the whole arena in one pass
        else:
            for s in range(len(self.exchange.slices)):
                if s not in done:
                    adam(cur, s)
        self.step_no += 1
        dec.mark_updated()
        return out
