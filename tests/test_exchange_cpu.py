"""World-size-2 gloo tests (CPU) of the gradient exchange used by data-parallel training: the stage-by-stage sum
all-reduce of the flat gradient arena (exact fp32 wire and the default bf16 wire), the (loss_sum, n_tokens) scalar
all-reduce, and the all-gather + scatter of the packed embedding-row gradients (trainer.SparseRowExchange)."""
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


def _flat(rank):
    return torch.randn(64 * 50, generator=torch.Generator().manual_seed(100 + rank))


def _sparse_case(rank, n=24, d=16):
    g = torch.Generator().manual_seed(7 + rank)
    rows = torch.randn(n, d, generator=g).bfloat16()
    band = torch.randint(0, 3, (n,), generator=g)
    loc = torch.randint(0, 10, (n,), generator=g)
    count = n - 3 * rank                                     # ranks hold different numbers of valid rows
    return rows, band, loc, count


def _worker(rank, world, port, q):
    os.environ.update(MASTER_ADDR='127.0.0.1', MASTER_PORT=str(port))
    dist.init_process_group('gloo', rank=rank, world_size=world)
    from news_image_caption_b200.trainer import GradientExchange, SparseRowExchange
    slices = [(0, 640), (640, 640), (640, 2560), (None, None), (2560, 3200)]      # incl. empty stages
    want = _flat(0) + _flat(1)
    ok = True
    for wire, tol in (('fp32', 0.0), ('bf16', 2e-2)):
        flat = _flat(rank)
        ex = GradientExchange(flat, slices, wire=wire)
        for i in range(len(slices)):
            ex.reduce_slice(i, async_op=(i % 2 == 0))
        ex.finish()
        ok = ok and (torch.equal(flat, want) if tol == 0.0 else bool(((flat - want).abs() <= tol * want.abs().clamp_min(1.0)).all()))
        if wire == 'bf16':          # the wire really carried bf16: every value is a bf16-representable sum of two bf16 numbers
            exp = (_flat(0).bfloat16() + _flat(1).bfloat16()).float()
            ok = ok and torch.equal(flat, exp)
    scal = ex.reduce_scalars(torch.tensor([10.0 * (rank + 1), 7.0 + rank]))
    ok = ok and scal.tolist() == [30.0, 15.0]
    # packed embedding-row gradients: every rank ends with the sum of ALL ranks' rows scattered into its band tables
    n, d, skip = 24, 16, 0
    sx = SparseRowExchange(n, d, torch.bfloat16, 'cpu')
    rows, meta, count = sx.views(sx.mine)
    r, band, loc, cnt = _sparse_case(rank)
    rows.copy_(r)
    meta.copy_((band << 32) | loc)
    count[0] = cnt
    sx.gather()
    tables = [torch.zeros(10, d) for _ in range(3)]
    sx.scatter_into(tables, skip)
    exp = [torch.zeros(10, d) for _ in range(3)]
    for rr in range(world):
        r2, b2, l2, c2 = _sparse_case(rr)
        for j in range(c2):
            if int(l2[j]) != skip:
                exp[int(b2[j])][int(l2[j])] += r2[j].float()
    ok = ok and all(torch.allclose(a, b, atol=1e-5) for a, b in zip(tables, exp))
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
