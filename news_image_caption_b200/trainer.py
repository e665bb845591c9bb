"""Data-parallel training driver for the B200 decoder engine (BASELINE.json config 3).

One process per GPU; every rank holds a full replica and a shard of the batch.  A step is
  engine forward + AdaptiveLoss + backward (tnt_decoder_train_step, all CUDA kernels of this repo)
  -> sum all-reduce of the flat gradient arena, stage by stage, on a communication stream that overlaps the
     rest of the backward pass (torch.distributed / NCCL over NVLink: the ONLY collective of the path,
     SURVEY 8e) plus a 2-scalar all-reduce of (loss_sum, n_tokens): the reference normalises the loss by the
     GLOBAL non-pad token count (criteria/adaptive_loss.py:62-65, transformer_faces_objects.py:85-88)
  -> fused BertAdam over the flat arenas (tnt_bert_adam_step; config.yaml:126-136).
torch supplies memory, streams and the process group only."""
import ctypes as C
import math

import torch
import torch.distributed as dist

from . import _lib


def warmup_linear(x, warmup):
    """pytorch_pretrained_bert.optimization.warmup_linear (the schedule named in config.yaml:131)."""
    if x < warmup:
        return x / warmup
    return max((x - 1.0) / (warmup - 1.0), 0.0)


class GradientExchange:
    """Sum all-reduce of slices of one flat gradient buffer.  Pure torch.distributed plumbing: works with
    CPU tensors over gloo (the world_size-2 CPU tests) and with CUDA tensors over NCCL."""

    def __init__(self, flat_grads, slices, group=None):
        self.flat, self.slices, self.group = flat_grads, [tuple(s) for s in slices], group
        self.world = dist.get_world_size(group) if dist.is_initialized() else 1
        self.pending = []

    def reduce_slice(self, i, async_op=True):
        a, b = self.slices[i]
        if self.world == 1 or a is None or b <= a:
            return
        w = dist.all_reduce(self.flat[a:b], op=dist.ReduceOp.SUM, group=self.group, async_op=async_op)
        if async_op:
            self.pending.append(w)

    def reduce_scalars(self, t):
        if self.world > 1:
            dist.all_reduce(t, op=dist.ReduceOp.SUM, group=self.group)
        return t

    def finish(self):
        for w in self.pending:
            w.wait()
        self.pending = []


class Trainer:
    def __init__(self, decoder, lr=1e-4, betas=(0.9, 0.98), eps=1e-6, weight_decay=1e-5, max_grad_norm=0.1,
                 warmup=0.05, t_total=-1, group=None, overlap=True):
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
        self.exchange = GradientExchange(self.dec.flat_grads, self.dec.stage_slices, group)
        self.comm_stream = torch.cuda.Stream() if self.exchange.world > 1 else None
        self.gscale = torch.ones(1, dtype=torch.float32, device='cuda')

    def lr_now(self):
        if self.t_total > 0:
            return self.lr * warmup_linear(self.step_no / self.t_total, self.warmup)
        return self.lr

    def step(self, prev_target, target, contexts):
        """One optimisation step on this rank's shard.  Returns the GLOBAL [loss_sum (nats), n_tokens] tensor
        (device); the reported loss of the reference is loss_sum / n_tokens / ln 2 (bits per token)."""
        dec, L = self.dec, _lib.lib()
        out = dec.train_step(prev_target, target, contexts, loss_scale=1.0)
        cur = torch.cuda.current_stream()
        if self.exchange.world > 1:
            h = dec._h()
            n_stages = L.tnt_decoder_train_stages(h)
            if self.overlap:
                with torch.cuda.stream(self.comm_stream):
                    for s in range(n_stages):
                        _lib.check(L.tnt_decoder_stream_wait_stage(h, s, C.c_void_p(self.comm_stream.cuda_stream)),
                                   'tnt_decoder_stream_wait_stage')
                        self.exchange.reduce_slice(s)
                    self.exchange.finish()
                cur.wait_stream(self.comm_stream)
            else:
                for s in range(n_stages):
                    self.exchange.reduce_slice(s, async_op=False)
            self.exchange.reduce_scalars(out)
        # gradient of the reference loss (sum-NLL / n_tokens / ln 2): scale by the global token count on the device
        torch.reciprocal(out[1:2] * math.log(2.0), out=self.gscale)
        a = _lib.AdamParams()
        a.params, a.grads, a.m, a.v = dec.flat_params.data_ptr(), dec.flat_grads.data_ptr(), self.m.data_ptr(), self.v.data_ptr()
        a.n = dec.flat_params.numel()
        a.seg_of_chunk, a.seg_sq, a.n_segments = self.seg_of_chunk.data_ptr(), self.seg_sq.data_ptr(), self.n_segments
        a.grad_scale_dev, a.grad_scale = self.gscale.data_ptr(), 1.0
        a.max_grad_norm, a.lr = self.max_norm, self.lr_now()
        a.beta1, a.beta2, a.eps, a.weight_decay = self.betas[0], self.betas[1], self.eps, self.wd
        _lib.check(L.tnt_bert_adam_step(C.byref(a), C.c_void_p(cur.cuda_stream)), 'tnt_bert_adam_step')
        self.step_no += 1
        dec.mark_updated()
        return out
