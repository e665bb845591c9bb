"""GPU parity tests of the fine-grained C-ABI ops against the oracle (oracle/restate.py) and the golden
fixtures frozen from the unmodified reference.  Tolerances: fp32 1e-4 relative, bf16 2e-2 (north_star)."""
import os

import pytest
import torch
import torch.nn.functional as F

from oracle import restate as O

pytestmark = pytest.mark.gpu
G = os.path.join(os.path.dirname(__file__), 'golden')
TOL = {torch.float32: 1e-4, torch.bfloat16: 2e-2}


def rel_err(a, b):
    a, b = a.float().cpu(), b.float().cpu()
    return ((a - b).abs().max() / b.abs().max().clamp_min(1e-6)).item()


@pytest.fixture(scope='module')
def ops():
    from news_image_caption_b200 import build, _lib, ops as o
    build.build()
    _lib.require_device()
    return o


# ----------------------------------------------------------------------------- GEMM
@pytest.mark.parametrize('M,N,K', [(128, 256, 64), (200, 1000, 1024), (2048, 2048, 1024), (77, 48, 32),
                                   (256, 5002, 1024), (3136, 8192, 2048), (1, 496, 1024)])
@pytest.mark.parametrize('dtype', [torch.bfloat16, torch.float32])
def test_linear(ops, M, N, K, dtype):
    g = torch.Generator().manual_seed(M + N + K)
    a = torch.randn(M, K, generator=g).to(dtype)
    w = (torch.randn(N, K, generator=g) / K ** 0.5).to(dtype)
    b = torch.randn(N, generator=g)
    ref = a.double() @ w.double().t() + b.double()          # plain torch reference of the same op
    out = ops.linear(a.cuda(), w.cuda(), b.cuda())
    torch.cuda.synchronize()
    assert rel_err(out, ref) < (1e-5 if dtype == torch.float32 else 8e-3)


@pytest.mark.parametrize('dtype', [torch.bfloat16, torch.float32])
def test_linear_epilogues(ops, dtype):
    from news_image_caption_b200 import _lib
    g = torch.Generator().manual_seed(3)
    M, N, K = 300, 512, 256
    a = torch.randn(M, K, generator=g).to(dtype)
    w = (torch.randn(N, K, generator=g) / K ** 0.5).to(dtype)
    b = torch.randn(N, generator=g)
    r = torch.randn(M, N, generator=g).to(dtype)
    z = a.double() @ w.double().t() + b.double()
    tol = 1e-5 if dtype == torch.float32 else 1e-2
    assert rel_err(ops.linear(a.cuda(), w.cuda(), b.cuda(), _lib.EPI_RELU), z.relu()) < tol
    assert rel_err(ops.linear(a.cuda(), w.cuda(), b.cuda(), _lib.EPI_RESIDUAL, residual=r.cuda()), z + r.double()) < tol
    # GLU: interleave rows with the library's own fold so the pairing convention is what is tested
    wi = ops.wn_fold(w.float().cuda(), None, glu_interleave=True, dtype=dtype)
    bi = ops.glu_interleave_bias(b.cuda())
    glu = F.glu(z, dim=-1)
    assert rel_err(ops.linear(a.cuda(), wi, bi, _lib.EPI_GLU), glu) < tol


@pytest.mark.parametrize('M,N,K', [(128, 64, 64), (2048, 1024, 1024), (300, 1000, 520), (2048, 5008, 1024), (1000, 72, 4096)])
@pytest.mark.parametrize('dtype', [torch.bfloat16, torch.float32])
def test_linear_backward_layouts(ops, M, N, K, dtype):
    """NN (dX = dY . W) and TN (dW = dY^T . X) operand layouts, used by the backward pass, incl. alpha,
    the relu-gradient epilogue, in-place accumulation and a device-side reduction length."""
    from news_image_caption_b200 import _lib
    g = torch.Generator().manual_seed(M * 3 + N + K)
    tol = 1e-5 if dtype == torch.float32 else 8e-3
    dy = torch.randn(M, K, generator=g).to(dtype)                 # [tokens, out]
    w = (torch.randn(K, N, generator=g) / K ** 0.5).to(dtype)     # [out, in]
    ref = dy.double() @ w.double()
    out = ops.linear(dy.cuda(), w.cuda(), layout=_lib.LAYOUT_NN, alpha=0.5)
    assert rel_err(out, 0.5 * ref) < tol
    act = torch.randn(M, N, generator=g).to(dtype)
    out = ops.linear(dy.cuda(), w.cuda(), epilogue=_lib.EPI_RELU_BWD, residual=act.cuda(), layout=_lib.LAYOUT_NN)
    assert rel_err(out, ref * (act.double() > 0)) < tol
    # weight gradient: tokens = M is the reduction
    x = torch.randn(M, N, generator=g).to(dtype)                  # [tokens, in]
    refw = dy.double().t() @ x.double()                           # [out=K, in=N]
    dw = ops.linear(dy.cuda(), x.cuda(), layout=_lib.LAYOUT_TN, out_dtype=torch.float32)
    assert rel_err(dw, refw) < tol
    acc0 = torch.randn(K, N, generator=g)
    dw2 = ops.linear(dy.cuda(), x.cuda(), epilogue=_lib.EPI_RESIDUAL, residual=acc0.cuda(), layout=_lib.LAYOUT_TN,
                     out_dtype=torch.float32)
    assert rel_err(dw2, refw + acc0.double()) < tol
    for cnt in (0, 1, 64 * (M // 128), M - 3):                    # device-side token count, rows beyond are zero
        cnt = max(0, min(cnt, M))
        dyz = dy.clone(); dyz[cnt:] = 0
        c = torch.tensor([cnt], dtype=torch.int32).cuda()
        dwc = ops.linear(dyz.cuda(), x.cuda(), layout=_lib.LAYOUT_TN, out_dtype=torch.float32, m_dyn=c)
        refc = dy[:cnt].double().t() @ x[:cnt].double()
        assert (dwc.cpu().double() - refc).abs().max().item() <= tol * max(1.0, refc.abs().max().item())


def test_tc_matches_simt(ops):
    """tcgen05 GEMM against the library's own fp32-accumulating SIMT kernel on identical bf16 inputs."""
    from news_image_caption_b200 import _lib
    g = torch.Generator().manual_seed(5)
    a = torch.randn(513, 1024, generator=g).bfloat16().cuda()
    w = (torch.randn(777, 1024, generator=g) / 32).bfloat16().cuda()
    tc = ops.linear(a, w, out_dtype=torch.float32)
    _lib.lib().tnt_debug_force_simt_gemm(1)
    try:
        sm = ops.linear(a, w, out_dtype=torch.float32)
    finally:
        _lib.lib().tnt_debug_force_simt_gemm(0)
    assert rel_err(tc, sm) < 1e-5


@pytest.mark.parametrize('M,N,K', [(256, 256, 64), (300, 520, 200), (1000, 777, 1024), (2048, 5002, 1024), (4096, 2048, 512)])
def test_linear_cta_pair(ops, M, N, K):
    """The CTA-pair (cta_group::2, 256 x 256 tiles) kernel, forced, against the single-CTA kernel and fp64:
    NT with every epilogue, NN, TN (ragged M / N / K: OOB rows are TMA zero fill)."""
    from news_image_caption_b200 import _lib
    g = torch.Generator().manual_seed(M + 7 * N + K)
    a = torch.randn(M, K, generator=g).bfloat16()
    K8 = K
    w = (torch.randn(N, K8, generator=g) / K ** 0.5).bfloat16()
    b = torch.randn(N, generator=g)
    r = torch.randn(M, N, generator=g).bfloat16()
    z = a.double() @ w.double().t() + b.double()
    L = _lib.lib()
    one = ops.linear(a.cuda(), w.cuda(), b.cuda(), out_dtype=torch.float32)
    L.tnt_debug_gemm_cta_group(2)
    try:
        two = ops.linear(a.cuda(), w.cuda(), b.cuda(), out_dtype=torch.float32)
        assert rel_err(two, z) < 8e-3
        assert rel_err(two, one) < 1e-5
        assert rel_err(ops.linear(a.cuda(), w.cuda(), b.cuda(), _lib.EPI_RELU), z.relu()) < 1e-2
        assert rel_err(ops.linear(a.cuda(), w.cuda(), b.cuda(), _lib.EPI_RESIDUAL, residual=r.cuda()), z + r.double()) < 1e-2
        if N % 2 == 0:
            wi = ops.wn_fold(w.float().cuda(), None, glu_interleave=True, dtype=torch.bfloat16)
            bi = ops.glu_interleave_bias(b.cuda())
            assert rel_err(ops.linear(a.cuda(), wi, bi, _lib.EPI_GLU), F.glu(z, dim=-1)) < 1e-2
        if N % 8 == 0 and K % 8 == 0:
            # backward layouts: dX = dY . W (NN), dW = dY^T . X (TN)
            dy, wkn = a, (torch.randn(K, N, generator=g) / K ** 0.5).bfloat16()
            assert rel_err(ops.linear(dy.cuda(), wkn.cuda(), layout=_lib.LAYOUT_NN), dy.double() @ wkn.double()) < 8e-3
            x = torch.randn(M, N, generator=g).bfloat16()
            dw = ops.linear(dy.cuda(), x.cuda(), layout=_lib.LAYOUT_TN, out_dtype=torch.float32)
            assert rel_err(dw, dy.double().t() @ x.double()) < 8e-3
        # the log-sum-exp epilogue (adaptive-softmax clusters) on the pair kernel
        tgt = torch.randint(0, N, (M,), generator=g)
        mx, se, arg, picked = ops.lse(a.cuda(), w.cuda(), b.cuda(), tgt.cuda())
        assert rel_err(mx + se.log(), torch.logsumexp(z, -1)) < 1e-3
        assert rel_err(picked, z.gather(1, tgt[:, None])[:, 0]) < 1e-2
    finally:
        L.tnt_debug_gemm_cta_group(0)


@pytest.mark.parametrize('M,N,K,S', [(256, 1024, 4096, -1), (256, 1024, 2048, -1), (1, 64, 2048, -1), (200, 1000, 4096 + 64, 4),
                                     (130, 96, 1024, 2), (64, 320, 8192, 8), (300, 40, 3000 + 8, 3), (256, 1024, 1024, -1), (2048, 112, 1024, -1)])
def test_linear_split_k(ops, M, N, K, S):
    """The cluster split-K kernel (decode-step GEMMs: few rows, long K) against fp64 and the persistent kernel:
    automatic plan (S = -1) and forced split factors, ragged M / N / K, every epilogue it serves."""
    from news_image_caption_b200 import _lib
    g = torch.Generator().manual_seed(M + 7 * N + K)
    a = torch.randn(M, K, generator=g).bfloat16()
    w = (torch.randn(N, K, generator=g) / K ** 0.5).bfloat16()
    b = torch.randn(N, generator=g)
    r = torch.randn(M, N, generator=g).bfloat16()
    z = a.double() @ w.double().t() + b.double()
    L = _lib.lib()
    L.tnt_debug_gemm_split_k(0)
    try:
        one = ops.linear(a.cuda(), w.cuda(), b.cuda(), out_dtype=torch.float32)
        L.tnt_debug_gemm_split_k(S)
        two = ops.linear(a.cuda(), w.cuda(), b.cuda(), out_dtype=torch.float32)
        assert rel_err(two, z) < 8e-3
        assert rel_err(two, one) < 2e-5          # same products, different fp32 summation order
        assert rel_err(ops.linear(a.cuda(), w.cuda(), None), z - b.double()) < 1e-2
        assert rel_err(ops.linear(a.cuda(), w.cuda(), b.cuda(), _lib.EPI_RELU), z.relu()) < 1e-2
        assert rel_err(ops.linear(a.cuda(), w.cuda(), b.cuda(), _lib.EPI_RESIDUAL, residual=r.cuda()), z + r.double()) < 1e-2
        if N % 2 == 0:
            wi = ops.wn_fold(w.float().cuda(), None, glu_interleave=True, dtype=torch.bfloat16)
            bi = ops.glu_interleave_bias(b.cuda())
            assert rel_err(ops.linear(a.cuda(), wi, bi, _lib.EPI_GLU), F.glu(z, dim=-1)) < 1e-2
    finally:
        L.tnt_debug_gemm_split_k(-1)


def test_dropout_masks_match_host_generator(ops):
    """tnt_dropout_apply against the numpy restatement of the counter-based generator: bit-exact mask, keep rate,
    1/(1-p) scaling, residual add, distinct sites / seeds give distinct masks."""
    g = torch.Generator().manual_seed(5)
    x = torch.randn(300, 1024, generator=g) + 3.0
    r = torch.randn(300, 1024, generator=g)
    for p, seed, site in ((0.1, 7, 3), (0.5, 2 ** 63 + 11, 0xFFFF), (0.25, 0, 17)):
        m = O.dropout_mask(p, seed, site, torch.arange(x.numel()).view(x.shape))
        y = ops.dropout(x.cuda(), p, seed, site).cpu()
        assert torch.equal(y, x * m)
        assert abs((m == 0).float().mean().item() - p) < 5e-3
        assert torch.equal(ops.dropout(x.cuda(), p, seed, site, residual=r.cuda()).cpu(), x * m + r)
        yb = ops.dropout(x.bfloat16().cuda(), p, seed, site).cpu()
        assert torch.equal(yb == 0, m == 0)
    m1 = O.dropout_mask(0.1, 7, 3, torch.arange(100000))
    assert (m1 != O.dropout_mask(0.1, 7, 4, torch.arange(100000))).float().mean() > 0.1
    assert (m1 != O.dropout_mask(0.1, 8, 3, torch.arange(100000))).float().mean() > 0.1
    assert torch.equal(ops.dropout(x.cuda(), 0.0, 1, 1).cpu(), x)


@pytest.mark.parametrize('dtype', [torch.float32, torch.bfloat16])
def test_linear_epilogue_dropout(ops, dtype):
    """dropout(a w^T + b) + residual and dropout(relu(.)) in the GEMM epilogues (tcgen05 persistent, split-K and SIMT)."""
    from news_image_caption_b200 import _lib
    g = torch.Generator().manual_seed(11)
    tol = 1e-5 if dtype == torch.float32 else 1e-2
    for (M, N, K) in ((200, 1000, 256), (256, 1024, 4096)):
        a = torch.randn(M, K, generator=g).to(dtype)
        w = (torch.randn(N, K, generator=g) / K ** 0.5).to(dtype)
        b = torch.randn(N, generator=g)
        r = torch.randn(M, N, generator=g).to(dtype)
        z = a.double() @ w.double().t() + b.double()
        m = O.dropout_mask(0.2, 99, 5, torch.arange(M * N).view(M, N)).double()
        y = ops.linear(a.cuda(), w.cuda(), b.cuda(), _lib.EPI_RESIDUAL, residual=r.cuda(), drop=(0.2, 99, 5))
        assert rel_err(y, z * m + r.double()) < tol
        y = ops.linear(a.cuda(), w.cuda(), b.cuda(), _lib.EPI_RELU, drop=(0.2, 99, 5))
        assert rel_err(y, z.relu() * m) < tol
        # explicit index layout: ld / group stride
        y = ops.linear(a.cuda(), w.cuda(), b.cuda(), _lib.EPI_RESIDUAL, residual=r.cuda(), drop=(0.2, 99, 5, 3 * N, N))
        m2g = O.dropout_mask(0.2, 99, 5, (torch.arange(M)[:, None] * (3 * N) + torch.arange(N)[None, :])).double()
        assert rel_err(y, z * m2g + r.double()) < tol      # group 0 of a grouped call: offset 0


@pytest.mark.parametrize('dtype', [torch.float32, torch.bfloat16])
@pytest.mark.parametrize('T,B,H,K', [(32, 3, 4, 7), (45, 2, 16, 31), (5, 2, 2, 3)])
def test_dynconv_tap_dropout(ops, dtype, T, B, H, K):
    """weight_dropout (dynamic.py:305): forward and backward against autograd through the oracle with the same mask."""
    g = torch.Generator().manual_seed(T + K)
    Cc = H * 64
    x = torch.randn(T, B, Cc, generator=g).to(dtype)
    lg = torch.randn(T, B, H * K, generator=g).to(dtype)
    dy = torch.randn(T, B, Cc, generator=g).to(dtype)
    xr, lr = x.double().requires_grad_(True), lg.double().requires_grad_(True)
    m = O.dropout_mask(0.3, 21, 1, torch.arange(T * B * H * K).view(T, B, H, K)).double()
    wts = torch.softmax(lr.view(T, B, H, K), -1) * m
    win = F.pad(xr, (0, 0, 0, 0, K - 1, 0)).unfold(0, K, 1)
    y = torch.einsum('tbhrk,tbhk->tbhr', win.reshape(T, B, H, 64, K), wts).reshape(T, B, Cc)
    gx, gl = torch.autograd.grad(y, [xr, lr], dy.double())
    tol = 1e-5 if dtype == torch.float32 else 2e-2
    assert rel_err(ops.dynconv(x.cuda(), lg.cuda(), H, K, drop=(0.3, 21, 1)), y.detach()) < tol
    dx, dl = ops.dynconv_bwd(x.cuda(), lg.cuda(), dy.cuda(), H, K, drop=(0.3, 21, 1))
    assert rel_err(dx, gx) < tol
    assert rel_err(dl, gl) < (1e-4 if dtype == torch.float32 else 3e-2)


def test_wn_fold(ops):
    g = torch.Generator().manual_seed(2)
    v, gg = torch.randn(64, 40, generator=g), torch.randn(64, 1, generator=g)
    assert rel_err(ops.wn_fold(v.cuda(), gg.cuda()), O.fold_weight_norm(gg, v)) < 1e-6


# ----------------------------------------------------------------------------- DynamicConv
@pytest.mark.parametrize('dtype', [torch.float32, torch.bfloat16])
def test_dynconv_golden(ops, dtype):
    for c in torch.load(os.path.join(G, 'dynconv_cases.pt'), weights_only=False):
        x = c['x'].to(dtype)
        logits = F.linear(x.float(), c['w']).to(dtype)           # the tap GEMM is tested separately
        y = ops.dynconv(x.cuda(), logits.cuda(), c['H'], c['K'])
        if dtype == torch.float32:
            assert rel_err(y, c['y']) < 1e-5, (c['T'], c['K'])
        else:
            w = F.softmax(logits.float().view(c['T'], c['B'], c['H'], c['K']), -1)
            xp = F.pad(x.float(), (0, 0, 0, 0, c['K'] - 1, 0)).unfold(0, c['K'], 1)
            ref = torch.einsum('tbhrk,tbhk->tbhr', xp.reshape(c['T'], c['B'], c['H'], -1, c['K']), w).reshape(x.shape)
            assert rel_err(y, ref) < 1e-2


def test_dynconv_full_size_and_step(ops):
    """T=32, B=8, C=1024, H=16, K=31 vs the oracle; incremental ring-buffer steps == full sequence."""
    g = torch.Generator().manual_seed(9)
    T, B, Cc, H, K = 32, 8, 1024, 16, 31
    x = torch.randn(T, B, Cc, generator=g)
    wt = torch.randn(H * K, Cc, generator=g) / 32
    ref = O.dynamic_conv(x, wt, H, K)
    logits = F.linear(x, wt)
    y = ops.dynconv(x.cuda(), logits.cuda(), H, K)
    assert rel_err(y, ref) < 1e-5
    ring = torch.zeros(K - 1, B, Cc, device='cuda')
    pos = torch.zeros(1, dtype=torch.int32, device='cuda')
    ys = []
    for t in range(T):
        ys.append(ops.dynconv_step(x[t].cuda(), logits[t].cuda(), ring, pos, H, K))
        pos += 1
    assert rel_err(torch.stack(ys), ref) < 1e-5
    # history form (dynamic.py:95-99): last K-1 inputs + 3 new steps
    y2 = ops.dynconv(x[20:23].cuda(), logits[20:23].cuda(), H, K, hist=x[:20][-(K - 1):].cuda())
    assert rel_err(y2, ref[20:23]) < 1e-5


def test_dynconv_backward(ops):
    g = torch.Generator().manual_seed(10)
    T, B, Cc, H, K = 12, 3, 64, 4, 7
    x = torch.randn(T, B, Cc, generator=g, requires_grad=True)
    logits = torch.randn(T, B, H * K, generator=g, requires_grad=True)
    w = F.softmax(logits.view(T, B, H, K), -1)
    xp = F.pad(x, (0, 0, 0, 0, K - 1, 0)).unfold(0, K, 1)
    y = torch.einsum('tbhrk,tbhk->tbhr', xp.reshape(T, B, H, Cc // H, K), w).reshape(T, B, Cc)
    dy = torch.randn(T, B, Cc, generator=g)
    gx, gl = torch.autograd.grad(y, [x, logits], dy)
    dx, dl = ops.dynconv_bwd(x.detach().cuda(), logits.detach().cuda(), dy.cuda(), H, K)
    assert rel_err(dx, gx) < 1e-5
    assert rel_err(dl, gl) < 1e-5


# ----------------------------------------------------------------------------- LayerNorm
def test_layernorm(ops):
    g = torch.Generator().manual_seed(4)
    x = torch.randn(70, 1024, generator=g) * 3 + 1
    gm, bt = torch.randn(1024, generator=g), torch.randn(1024, generator=g)
    ref = F.layer_norm(x, (1024,), gm, bt)
    assert rel_err(ops.layernorm(x.cuda(), gm.cuda(), bt.cuda()), ref) < 1e-5
    assert rel_err(ops.layernorm(x.bfloat16().cuda(), gm.cuda(), bt.cuda()), ref) < 2e-2
    v2 = O.v2_layer_norm(x, gm, bt)
    assert rel_err(ops.layernorm(x.cuda(), gm.cuda(), bt.cuda(), eps=1e-6, variant=1), v2) < 1e-5
    xr = x.clone().requires_grad_(True)
    dy = torch.randn(70, 1024, generator=g)
    gmr, btr = gm.clone().requires_grad_(True), bt.clone().requires_grad_(True)
    gx, gg, gb = torch.autograd.grad(F.layer_norm(xr, (1024,), gmr, btr), [xr, gmr, btr], dy)
    dx, dg, db = ops.layernorm_bwd(x.cuda(), dy.cuda(), gm.cuda())
    assert rel_err(dx, gx) < 1e-4 and rel_err(dg, gg) < 1e-4 and rel_err(db, gb) < 1e-4


# ----------------------------------------------------------------------------- Linear + residual + LayerNorm, one launch
@pytest.mark.parametrize('M,N,K,groups', [(2048, 1024, 1024, 1),      # a layer site at the bench's B*T: 16 clusters of 8 CTAs
                                          (300, 1024, 4096, 1),       # ragged last block, long K (fc2)
                                          (2048, 1024, 1024, 2),      # per-source out_proj + context_attn_lns: 4 CTAs x 256 columns
                                          (256, 1024, 1024, 4),       # decode-step rows, four sources
                                          (100, 512, 1024, 1),        # 8 CTAs x 64 columns
                                          (130, 2048, 512, 1)])       # 8 CTAs x 256 columns
def test_linear_residual_layernorm_fused(ops, M, N, K, groups):
    """tnt_linear_ln_fwd (gemm_ln.cu: a cluster of CTAs per 128-row block, row statistics through distributed shared memory)
    against fp32 torch on the bf16-rounded operands, and against the two-launch form it replaces (tnt_linear_fwd with the
    residual epilogue, then tnt_layernorm_fwd).  bf16 outputs: 2e-2 of the largest value; against the two-launch form the
    only difference is that the fused kernel does not round the pre-LayerNorm row to bf16."""
    g = torch.Generator().manual_seed(11)
    a = torch.randn(M, groups * K, generator=g).bfloat16()
    w = (torch.randn(groups, N, K, generator=g) / K ** 0.5).bfloat16()
    bias = torch.randn(groups, N, generator=g)
    res = (torch.randn(M, N, generator=g) * 2 + 0.5).bfloat16()
    gm, bt = torch.rand(groups, N, generator=g) + 0.5, torch.randn(groups, N, generator=g)
    ref = torch.cat([F.layer_norm(a[:, s * K:(s + 1) * K].float() @ w[s].float().t() + bias[s] + res.float(), (N,), gm[s], bt[s])
                     for s in range(groups)], 1)
    if groups == 1:
        y = ops.linear_ln(a.cuda(), w[0].cuda(), bias[0].cuda(), res.cuda(), gm[0].cuda(), bt[0].cuda())
    else:
        y = ops.linear_ln(a.cuda(), w.cuda(), bias.cuda(), res.cuda(), gm.cuda(), bt.cuda(), groups=groups)
    assert y.shape == (M, groups * N)
    assert rel_err(y, ref) < 2e-2
    two = torch.cat([ops.layernorm(ops.linear(a[:, s * K:(s + 1) * K].cuda(), w[s].cuda(), bias[s].cuda(), epilogue=3, residual=res.cuda()),
                                   gm[s].cuda(), bt[s].cuda()) for s in range(groups)], 1)
    assert rel_err(y, two) < 2e-2
    # tighter, on the error mass: both forms agree with the fp32 result to bf16 rounding, the fused one at least as well
    e_f = (y.float().cpu() - ref).norm() / ref.norm()
    e_t = (two.float().cpu() - ref).norm() / ref.norm()
    assert e_f < 4e-3 and e_f < 1.2 * e_t + 1e-4, (e_f, e_t)


def test_linear_residual_layernorm_fused_transposed_store_and_no_residual(ops):
    """the last layer's site writes batch-major rows: row t*B + b of the problem lands in row b*T + t of the output."""
    g = torch.Generator().manual_seed(12)
    T, B, N, K = 32, 5, 1024, 1024
    a = torch.randn(T * B, K, generator=g).bfloat16()
    w = (torch.randn(N, K, generator=g) / K ** 0.5).bfloat16()
    bias, gm, bt = torch.randn(N, generator=g), torch.rand(N, generator=g) + 0.5, torch.randn(N, generator=g)
    res = torch.randn(T * B, N, generator=g).bfloat16()
    ref = F.layer_norm(a.float() @ w.float().t() + bias + res.float(), (N,), gm, bt)
    y = ops.linear_ln(a.cuda(), w.cuda(), bias.cuda(), res.cuda(), gm.cuda(), bt.cuda(), transpose_tb=(T, B))
    assert y.shape == (B, T, N)
    assert rel_err(y, ref.view(T, B, N).transpose(0, 1)) < 2e-2
    ref0 = F.layer_norm(a.float() @ w.float().t(), (N,), gm, bt)
    assert rel_err(ops.linear_ln(a.cuda(), w.cuda(), None, None, gm.cuda(), bt.cuda()), ref0) < 2e-2
    with pytest.raises(RuntimeError):                       # fp32 operands: not a tcgen05 problem, the caller keeps two launches
        ops.linear_ln(a.float().cuda(), w.float().cuda(), bias.cuda(), res.float().cuda(), gm.cuda(), bt.cuda())


# ----------------------------------------------------------------------------- attention
@pytest.mark.parametrize('dtype', [torch.float32, torch.bfloat16])
def test_xattn_golden(ops, dtype):
    for c in torch.load(os.path.join(G, 'mha_cases.pt'), weights_only=False):
        sd, E, H = c['state_dict'], c['E'], c['H']
        wq, wk, wv, bq, bk, bv = O.mha_params(sd, '', E)
        q = F.linear(c['query'], wq, bq).to(dtype)
        k = F.linear(c['memory'], wk, bk).to(dtype)
        v = F.linear(c['memory'], wv, bv).to(dtype)
        o, w = ops.xattn(q.cuda(), k.cuda(), v.cuda(), H, (E // H) ** -0.5, pad_mask=c['mask'].cuda(),
                         bias_k=sd['bias_k'].cuda(), bias_v=sd['bias_v'].cuda(), need_weights=True)
        out = F.linear(o.float().cpu(), sd['out_proj.weight'], sd['out_proj.bias'])
        assert rel_err(out, c['out']) < TOL[dtype]
        assert rel_err(w, c['weights']) < TOL[dtype]


@pytest.mark.parametrize('B,S', [(5, 514), (3, 51), (2, 6), (9, 1000)])
def test_xattn_decode_step(ops, B, S):
    """T == 1 (one query per caption and head: the `_generate` step): the HBM-streaming decode kernel vs the
    oracle's attention core; ragged padding incl. captions whose only live keys are the 2 trailing extra rows."""
    g = torch.Generator().manual_seed(B * 1000 + S)
    E, H = 1024, 16
    q, k, v = (torch.randn(n, B, E, generator=g).bfloat16() for n in (1, S, S))
    n = torch.randint(0, S - 1, (B,), generator=g)
    n[0], n[-1] = S - 2, 0                                          # nothing masked / everything but the extras masked
    mask = torch.arange(S - 2)[None] >= n[:, None]                  # covers keys [0, S-2): the last two rows are extras
    sd = {'in_proj_bias': torch.zeros(3 * E), 'in_proj_weight': torch.eye(E).repeat(3, 1),
          'out_proj.weight': torch.eye(E), 'out_proj.bias': torch.zeros(E)}
    full_mask = torch.cat([mask, torch.zeros(B, 2, dtype=torch.bool)], 1)
    ref, _ = O.mha_attend(q.float(), k.float(), v.float(), full_mask[:, :S - 2], sd, '', H)
    lse = torch.empty(B, H, 1, device='cuda')
    o, _ = ops.xattn(q.cuda(), k.cuda(), v.cuda(), H, 64 ** -0.5, pad_mask=mask.cuda(), mask_S=S - 2, lse=lse)
    assert rel_err(o, ref) < 1e-2
    # log-sum-exp of the scaled masked scores (what the backward consumes), vs plain torch
    sc = torch.einsum('bhd,sbhd->bhs', q[0].float().view(B, H, 64), k.float().view(S, B, H, 64)) * 64 ** -0.5
    sc = sc.masked_fill(full_mask[:, None, :], float('-inf'))
    assert rel_err(lse[:, :, 0], torch.logsumexp(sc, -1)) < 1e-3
    from news_image_caption_b200 import _lib
    _lib.lib().tnt_debug_force_simt_attention(1)
    try:
        o2, _ = ops.xattn(q.cuda(), k.cuda(), v.cuda(), H, 64 ** -0.5, pad_mask=mask.cuda(), mask_S=S - 2)
    finally:
        _lib.lib().tnt_debug_force_simt_attention(0)
    assert rel_err(o, o2) < 1e-2


def test_xattn_full_size(ops):
    """T=32, B=3, H=16, D=64, S=512 with ragged padding, vs the oracle's attention core."""
    g = torch.Generator().manual_seed(6)
    T, B, E, H, S = 32, 3, 1024, 16, 512
    q, k, v = (torch.randn(n, B, E, generator=g) for n in (T, S, S))
    bk, bv = torch.randn(1, 1, E, generator=g), torch.randn(1, 1, E, generator=g)
    n = torch.tensor([512, 130, 0])
    mask = torch.arange(S)[None] >= n[:, None]
    sd = {'in_proj_bias': torch.zeros(3 * E), 'in_proj_weight': torch.eye(E).repeat(3, 1), 'bias_k': bk, 'bias_v': bv,
          'out_proj.weight': torch.eye(E), 'out_proj.bias': torch.zeros(E)}
    kk = torch.cat([k, bk.expand(1, B, E), torch.zeros(1, B, E)])
    vv = torch.cat([v, bv.expand(1, B, E), torch.zeros(1, B, E)])
    ref, wref = O.mha_attend(q, kk, vv, mask, sd, '', H)
    o, w = ops.xattn(q.cuda(), k.cuda(), v.cuda(), H, 64 ** -0.5, pad_mask=mask.cuda(), bias_k=bk.cuda(),
                     bias_v=bv.cuda(), need_weights=True)
    assert rel_err(o, ref) < 1e-4 and rel_err(w, wref) < 1e-4
    ob, _ = ops.xattn(q.bfloat16().cuda(), k.bfloat16().cuda(), v.bfloat16().cuda(), H, 64 ** -0.5,
                      pad_mask=mask.cuda(), bias_k=bk.cuda(), bias_v=bv.cuda())
    assert rel_err(ob, ref) < 2e-2


@pytest.mark.parametrize('T,B,S', [(32, 3, 514), (7, 2, 130), (1, 5, 51), (40, 2, 300)])
def test_xattn_tcgen05(ops, T, B, S):
    """tcgen05/TMEM attention kernel (bf16, D=64, extra rows materialised as ordinary keys, mask covering
    only the first S-2 keys) vs a plain fp32 torch reference of the same op and vs the SIMT kernel."""
    from news_image_caption_b200 import _lib
    g = torch.Generator().manual_seed(T * 1000 + S)
    E, H = 1024, 16
    q = torch.randn(T, B, E, generator=g).bfloat16()
    k = torch.randn(S, B, E, generator=g).bfloat16()
    v = torch.randn(S, B, E, generator=g).bfloat16()
    n = torch.randint(0, S - 1, (B,), generator=g)
    n[0] = S - 2
    mask = torch.arange(S - 2)[None] >= n[:, None]                      # [B, S-2]; last two keys never masked
    full = torch.cat([mask, torch.zeros(B, 2, dtype=torch.bool)], 1)
    qh = q.float().view(T, B, H, 64).permute(1, 2, 0, 3) * 0.125
    kh = k.float().view(S, B, H, 64).permute(1, 2, 0, 3)
    vh = v.float().view(S, B, H, 64).permute(1, 2, 0, 3)
    s = (qh @ kh.transpose(-1, -2)).masked_fill(full[:, None, None, :], float('-inf'))
    ref = (torch.softmax(s, -1) @ vh).permute(2, 0, 1, 3).reshape(T, B, E)
    o, _ = ops.xattn(q.cuda(), k.cuda(), v.cuda(), H, 0.125, pad_mask=mask.cuda(), mask_S=S - 2)
    torch.cuda.synchronize()
    assert rel_err(o, ref) < 2e-2
    _lib.lib().tnt_debug_force_simt_attention(1)
    try:
        o2, _ = ops.xattn(q.cuda(), k.cuda(), v.cuda(), H, 0.125, pad_mask=mask.cuda(), mask_S=S - 2)
    finally:
        _lib.lib().tnt_debug_force_simt_attention(0)
    assert rel_err(o, o2) < 1.5e-2


@pytest.mark.parametrize('T,B,S', [(32, 3, 51), (32, 5, 6), (40, 2, 66), (9, 2, 128), (2, 3, 17), (33, 2, 100)])
def test_xattn_short_sources(ops, T, B, S):
    """Warp-level mma kernel for memory sources of <= 128 keys (attention_short.cu; image regions / faces / objects): output,
    LSE and the dropped-out probabilities vs a plain fp32 torch reference and vs the SIMT kernel (same dropout stream)."""
    from news_image_caption_b200 import _lib
    L = _lib.lib()
    g = torch.Generator().manual_seed(T * 1000 + S)
    E, H = 1024, 16
    q = torch.randn(T, B, E, generator=g).bfloat16()
    k = torch.randn(S, B, E, generator=g).bfloat16()
    v = torch.randn(S, B, E, generator=g).bfloat16()
    n = torch.randint(0, max(S - 2, 1), (B,), generator=g)
    n[0] = S - 2
    mask = torch.arange(S - 2)[None] >= n[:, None]                      # [B, S-2]; last two keys never masked
    full = torch.cat([mask, torch.zeros(B, 2, dtype=torch.bool)], 1)
    qh = q.float().view(T, B, H, 64).permute(1, 2, 0, 3) * 0.125
    kh = k.float().view(S, B, H, 64).permute(1, 2, 0, 3)
    vh = v.float().view(S, B, H, 64).permute(1, 2, 0, 3)
    s = (qh @ kh.transpose(-1, -2)).masked_fill(full[:, None, None, :], float('-inf'))
    ref = (torch.softmax(s, -1) @ vh).permute(2, 0, 1, 3).reshape(T, B, E)
    lse_ref = torch.logsumexp(s, -1)                                    # [B, H, T]
    n0 = L.tnt_variant_count(b'xattn.short')
    lse = torch.zeros(B, H, T, device='cuda')
    o, _ = ops.xattn(q.cuda(), k.cuda(), v.cuda(), H, 0.125, pad_mask=mask.cuda(), mask_S=S - 2, lse=lse)
    torch.cuda.synchronize()
    assert L.tnt_variant_count(b'xattn.short') == n0 + 1
    assert rel_err(o, ref) < 2e-2
    assert (lse.cpu() - lse_ref).abs().max().item() < 2e-2
    # dropout on the probabilities: same counter-based mask as the SIMT kernel
    drop = (0.3, 1234, 77)
    od, _ = ops.xattn(q.cuda(), k.cuda(), v.cuda(), H, 0.125, pad_mask=mask.cuda(), mask_S=S - 2, drop=drop)
    L.tnt_debug_force_simt_attention(1)
    try:
        o2, _ = ops.xattn(q.cuda(), k.cuda(), v.cuda(), H, 0.125, pad_mask=mask.cuda(), mask_S=S - 2)
        od2, _ = ops.xattn(q.cuda(), k.cuda(), v.cuda(), H, 0.125, pad_mask=mask.cuda(), mask_S=S - 2, drop=drop)
    finally:
        L.tnt_debug_force_simt_attention(0)
    assert rel_err(o, o2) < 1.5e-2
    assert rel_err(od, od2) < 2e-2
    assert rel_err(od, o) > 5e-2                                        # the mask did something


def test_xattn_v2_causal(ops):
    """final*/transformer_v2 attention(): mask == 0 -> -1e9, causal self-attention, d_k = 128."""
    g = torch.Generator().manual_seed(8)
    B, T, d, h = 2, 9, 256, 2
    x = torch.randn(B, T, d, generator=g)
    sd = {'linears.0.weight': torch.eye(d), 'linears.0.bias': torch.zeros(d), 'linears.1.weight': torch.eye(d),
          'linears.1.bias': torch.zeros(d), 'linear_kv.0.weight': torch.eye(d), 'linear_kv.0.bias': torch.zeros(d),
          'linear_kv.1.weight': torch.eye(d), 'linear_kv.1.bias': torch.zeros(d)}
    tm = O.v2_subsequent_mask(T)
    ref = O.v2_mha(x, x, tm, sd, '', h)
    xt = x.transpose(0, 1).contiguous().cuda()
    o, _ = ops.xattn(xt, xt, xt, h, (d // h) ** -0.5, v2_mask=tm.cuda())
    assert rel_err(o.transpose(0, 1), ref) < 1e-4


# ----------------------------------------------------------------------------- LSE head
@pytest.mark.parametrize('dtype', [torch.float32, torch.bfloat16])
def test_lse(ops, dtype):
    g = torch.Generator().manual_seed(12)
    M, N, K = 300, 5002, 256
    x = torch.randn(M, K, generator=g).to(dtype)
    w = (torch.randn(N, K, generator=g) / 8).to(dtype)
    tgt = torch.randint(0, N, (M,), generator=g)
    z = x.double() @ w.double().t()
    m, s, arg, tz = ops.lse(x.cuda(), w.cuda(), target=tgt.cuda())
    lse = (m.double() + s.double().log()).cpu()
    tol = 1e-5 if dtype == torch.float32 else 1e-3
    assert ((lse - torch.logsumexp(z, 1)).abs().max() / lse.abs().max()).item() < tol
    assert rel_err(tz, z.gather(1, tgt[:, None])[:, 0]) < (1e-5 if dtype == torch.float32 else 1e-2)
    if dtype == torch.float32:
        assert torch.equal(arg.cpu().long(), z.argmax(1))


# ----------------------------------------------------------------------------- context assembly (row f2)
@pytest.mark.parametrize('dtype', [torch.float32, torch.bfloat16])
@pytest.mark.parametrize('B,S,E,n', [(3, 17, 1024, 25), (2, 512, 1024, 25), (1, 5, 64, 13), (4, 33, 256, 1)])
def test_context_mix(ops, dtype, B, S, E, n):
    g = torch.Generator().manual_seed(B + S + n)
    hs = [torch.randn(B, S, E, generator=g).to(dtype) for _ in range(n)]
    w = torch.randn(n, generator=g)
    ref = O.context_mix([h.float() for h in hs], w)
    out = ops.context_mix([h.cuda() for h in hs], w)
    assert out.shape == (S, B, E) and out.dtype == dtype
    assert rel_err(out, ref) < (1e-5 if dtype == torch.float32 else 8e-3)
    out32 = ops.context_mix([h.cuda() for h in hs], w, out_dtype=torch.float32)
    assert rel_err(out32, ref) < 1e-5


@pytest.mark.parametrize('dtype', [torch.float32, torch.bfloat16])
def test_nan_mask(ops, dtype):
    g = torch.Generator().manual_seed(4)
    for B, n, d in ((5, 4, 512), (3, 64, 2048), (2, 1, 40)):
        x = torch.randn(B, n, d, generator=g).to(dtype)
        pad = torch.rand(B, n, generator=g) < 0.4
        x[pad] = float('nan')
        x[0, 0, d // 2] = float('nan')                       # a single NaN element also marks the row
        ref, mref = O.nan_padding_mask(x.float())
        out, m = ops.nan_mask(x.cuda())
        assert torch.equal(m.cpu(), mref)
        assert torch.equal(out.float().cpu(), ref)           # pure copy / zero fill: bit exact


@pytest.mark.parametrize('case', ['key_mask', 'causal', 'no_mask'])
def test_xattn_tc_head_dim_128_v2_mask(case):
    """The `final*` attention (8 heads of 128, attention() of final_roberta2/transformer_v2.py:213-222: mask == 0 -> -1e9) on
    the tcgen05 kernel's D = 128 instantiation against fp32 torch on the CPU and against the SIMT kernel: per-key article
    masks incl. a FULLY masked caption (uniform attention, not zero), the causal [B, T, T] self-attention mask over two query
    tiles, three key tiles with a ragged tail, and the unmasked image attention."""
    import math
    from news_image_caption_b200 import _lib, build, ops
    build.build()
    L = _lib.lib()
    g = torch.Generator().manual_seed(11)
    B, H, D = 3, 8, 128
    E = H * D
    T, S = (40, 40) if case == 'causal' else (40, 300)
    q = (torch.randn(T, B, E, generator=g) * 0.5).bfloat16()
    k = (torch.randn(B, S, E, generator=g) * 0.5).bfloat16()
    v = torch.randn(B, S, E, generator=g).bfloat16()
    mask = None
    if case == 'key_mask':
        mask = torch.rand(B, 1, S, generator=g) > 0.3
        mask[1] = False                                      # nothing visible: softmax over -1e9 everywhere = uniform
    elif case == 'causal':
        mask = torch.tril(torch.ones(1, T, T, dtype=torch.bool)).expand(B, T, T).clone()
    qf = q.float().view(T, B, H, D).permute(1, 2, 0, 3)
    kf = k.float().view(B, S, H, D).permute(0, 2, 1, 3)
    vf = v.float().view(B, S, H, D).permute(0, 2, 1, 3)
    s = qf @ kf.transpose(-1, -2) / math.sqrt(D)
    if mask is not None:
        s = s.masked_fill(mask[:, None] == 0, -1e9)
    want = (torch.softmax(s, -1) @ vf).permute(2, 0, 1, 3).reshape(T, B, E)
    n0 = L.tnt_variant_count(b'xattn.tc')
    args = (q.cuda(), (B * E, E), k.cuda(), (E, S * E), v.cuda(), (E, S * E), B, T, S, H, D, 1.0 / math.sqrt(D))
    out = ops.xattn_strided(*args, v2_mask=mask.cuda() if mask is not None else None)
    torch.cuda.synchronize()
    assert L.tnt_variant_count(b'xattn.tc') == n0 + 1          # the tcgen05 kernel ran
    assert rel_err(out, want) < 2e-2
    L.tnt_debug_force_simt_attention(1)
    try:
        ref = ops.xattn_strided(*args, v2_mask=mask.cuda() if mask is not None else None)
    finally:
        L.tnt_debug_force_simt_attention(0)
    assert rel_err(out, ref) < 2e-2
