"""World-size-2 gloo test (CPU) of the gradient exchange used by data-parallel training: the
stage-by-stage sum all-reduce of the flat gradient arena and the (loss_sum, n_tokens) scalar all-reduce."""
import os
import socket

import torch
import torch.distributed as dist
import torch.multiprocessing as mp


def _free_port():
    s = socket.socket()
    s.bind(('127.0.0.1', 0))
    p = s.getsockname()[1]
    s.close()
    return p


def _worker(rank, world, port, q):
    os.environ.update(MASTER_ADDR='127.0.0.1', MASTER_PORT=str(port))
    dist.init_process_group('gloo', rank=rank, world_size=world)
    from news_image_caption_b200.trainer import GradientExchange
    g = torch.Generator().manual_seed(100 + rank)
    flat = torch.randn(64 * 50, generator=g)
    mine = flat.clone()
    slices = [(0, 640), (640, 640), (640, 2560), (None, None), (2560, 3200)]      # incl. empty stages
    ex = GradientExchange(flat, slices)
    for i in range(len(slices)):
        ex.reduce_slice(i, async_op=(i % 2 == 0))
    ex.finish()
    scal = ex.reduce_scalars(torch.tensor([10.0 * (rank + 1), 7.0 + rank]))
    other = torch.randn(64 * 50, generator=torch.Generator().manual_seed(100 + (1 - rank)))
    ok = torch.allclose(flat, mine + other) and scal.tolist() == [30.0, 15.0]
    q.put((rank, bool(ok)))
    dist.destroy_process_group()


def test_gradient_exchange_world2():
    ctx = mp.get_context('spawn')
    q = ctx.Queue()
    port = _free_port()
    ps = [ctx.Process(target=_worker, args=(r, 2, port, q)) for r in range(2)]
    for p in ps:
        p.start()
    res = sorted(q.get(timeout=120) for _ in ps)
    for p in ps:
        p.join(timeout=60)
    assert res == [(0, True), (1, True)]
