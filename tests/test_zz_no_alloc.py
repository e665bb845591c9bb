"""tnt_lse_fwd_static: the workspace-free variant of tnt_lse_fwd (scratch in a device array built into the library; added
for callers that cannot allocate).  Checked through the C ABI against tnt_lse_fwd itself, on two calls in a row that share
the built-in array.  Collected last on purpose."""
import ctypes as C

import pytest
import torch

pytestmark = pytest.mark.gpu
TOL = {torch.float32: 1e-4, torch.bfloat16: 2e-2}      # the tolerances of tests/test_gpu_ops.py


@pytest.fixture(scope='module')
def L():
    from news_image_caption_b200 import build, _lib
    build.build()
    _lib.require_device()
    return _lib


def _params(L, x, w, b, tg, stats, ws=None):
    code = {torch.float32: L.TNT_F32, torch.bfloat16: L.TNT_BF16}[x.dtype]
    p = L.LseParams()
    p.x, p.ldx, p.w, p.ldw = x.data_ptr(), x.shape[1], w.data_ptr(), w.shape[1]
    p.bias, p.target, p.stats = b.data_ptr(), tg.data_ptr(), stats.data_ptr()
    if ws is not None:
        p.workspace, p.workspace_bytes = ws.data_ptr(), ws.numel()
    p.M, p.n_cols, p.K, p.dtype = x.shape[0], w.shape[0], x.shape[1], code
    return p


def _case(M, N, K, dtype, seed):
    g = torch.Generator().manual_seed(seed)
    x = torch.randn(M, K, generator=g).to(dtype).cuda()
    w = (torch.randn(N, K, generator=g) / K ** 0.5).to(dtype).cuda()
    b = torch.randn(N, generator=g).cuda()
    tg = torch.randint(0, N, (M,), generator=g).to(torch.int32).cuda()
    return x, w, b, tg


def _check(a, b, tol):
    a, b = a.cpu(), b.cpu()
    for col in (0, 1, 3):                                   # max, sum exp(z - max), target logit
        err = (a[:, col] - b[:, col]).abs().max() / b[:, col].abs().max().clamp_min(1e-6)
        assert err.item() <= tol, (col, err.item())
    assert torch.equal(a[:, 2].view(torch.int32), b[:, 2].view(torch.int32))       # argmax: index work, exact


@pytest.mark.parametrize('dtype', [torch.bfloat16, torch.float32])
def test_static_scratch_lse_matches_workspace_lse_twice(L, dtype):
    lib = L.lib()
    st = C.c_void_p(torch.cuda.current_stream().cuda_stream)
    # two calls in a row on one stream, the second smaller than the first: the shared array still holds the first call's
    # partials and none of them may leak into the second result
    for M, N, K, seed in ((256, 50265, 1024, 1), (77, 5002, 256, 2)):
        x, w, b, tg = _case(M, N, K, dtype, seed)
        assert lib.tnt_lse_workspace_bytes(M, N, 0) <= lib.tnt_lse_fwd_static_max_bytes()
        ws = torch.empty(lib.tnt_lse_workspace_bytes(M, N, 0) + 256, dtype=torch.uint8, device='cuda')
        ref = torch.zeros(M, 4, device='cuda')
        out = torch.zeros(M, 4, device='cuda')
        p_ref, p_out = _params(L, x, w, b, tg, ref, ws), _params(L, x, w, b, tg, out)
        L.check(lib.tnt_lse_fwd(C.byref(p_ref), st), 'tnt_lse_fwd')
        torch.cuda.synchronize()
        free0 = torch.cuda.mem_get_info()[0]
        L.check(lib.tnt_lse_fwd_static(C.byref(p_out), st), 'tnt_lse_fwd_static')
        torch.cuda.synchronize()
        if seed == 2:                                       # (the first call of a process may page the kernels in)
            assert torch.cuda.mem_get_info()[0] == free0    # the call allocated nothing on the device
        _check(out, ref, TOL[dtype])
        # and against plain torch
        z = x.float() @ w.float().t() + b
        assert torch.equal(out[:, 2].view(torch.int32).long().cpu(), z.argmax(1).cpu()) or dtype == torch.bfloat16
        lse = out[:, 0] + out[:, 1].log()
        assert ((lse - torch.logsumexp(z, 1)).abs().max() / torch.logsumexp(z, 1).abs().max()).item() < TOL[dtype]


def test_static_scratch_lse_rejects_sizes_beyond_the_array(L):
    lib = L.lib()
    M, N, K = 2048, 50265, 64                               # 2048 x 1571 records = 51 MB > the 8 MiB array
    assert lib.tnt_lse_workspace_bytes(M, N, 0) > lib.tnt_lse_fwd_static_max_bytes()
    x, w, b, tg = _case(8, 16, K, torch.float32, 3)         # small real buffers: the call must fail before touching them
    stats = torch.zeros(M, 4, device='cuda')
    p = _params(L, x, w, b, tg, stats)
    p.M, p.n_cols = M, N
    rc = lib.tnt_lse_fwd_static(C.byref(p), C.c_void_p(torch.cuda.current_stream().cuda_stream))
    assert rc != 0
    msg = lib.tnt_last_error().decode()
    assert 'lse_static' in msg and 'tnt_lse_fwd' in msg
    torch.cuda.synchronize()
