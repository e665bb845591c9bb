"""CPU tests of the host-side logic that needs no GPU: the incremental-state key scheme, the Noam schedule and its
scheduler, BertAdam's warm-up schedule, bench.py's row sampling / peak selection, and the C-ABI structs the round-2 entry
points take (sizes must match the header's layout on this platform)."""
import ctypes as C
import math
import os
import sys

import torch

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)


def test_incremental_state_keys_follow_the_reference_scheme():
    """utils/state.py:17-26: '{ClassName}.{instance id}.{key}', one counter per class name, ids handed out once."""
    from news_image_caption_b200 import state as S
    a = S.new_instance_id('DynamicConv1dTBC')
    b = S.new_instance_id('DynamicConv1dTBC')
    c = S.new_instance_id('SinusoidalPositionalEmbedding')
    assert b == a + 1 and c >= 1
    k = S.full_key('DynamicConv1dTBC', a, 'input_buffer')
    assert k == 'DynamicConv1dTBC.%d.input_buffer' % a
    assert 'DynamicConv1dTBC' in k                 # what decoder_flattened.py:205 tests for


def test_noam_rate_and_scheduler():
    """custom_v4.py:421-430: rate(step) = factor * d^-0.5 * min(step^-0.5, step * warmup^-1.5), step 0 treated as 1;
    NoamScheduler == torch's LambdaLR(rate) on the same base lr."""
    from news_image_caption_b200.transformer_v2 import NoamScheduler, rate
    d, f, w = 1024, 1.0, 30000
    assert rate(0, d, f, w) == rate(1, d, f, w)
    for s in (1, 10, 29999, 30000, 30001, 10 ** 6):
        want = f * d ** -0.5 * min(s ** -0.5, s * w ** -1.5)
        assert abs(rate(s, d, f, w) - want) < 1e-15
    assert rate(30000, d, f, w) > rate(29000, d, f, w) and rate(30000, d, f, w) > rate(31000, d, f, w)   # peak at warm-up

    class Opt:
        param_groups = [dict(lr=1.0, initial_lr=1.0)]
    p = torch.nn.Parameter(torch.zeros(1))
    ref = torch.optim.Adam([p], lr=1.0)
    lam = torch.optim.lr_scheduler.LambdaLR(ref, lambda s: rate(s, d, f, 7))
    o = Opt()
    sch = NoamScheduler(o, d, f, 7)
    for _ in range(12):
        assert abs(o.param_groups[0]['lr'] - ref.param_groups[0]['lr']) < 1e-15
        ref.step()
        lam.step()
        sch.step()


def test_bert_adam_warmup_linear():
    """pytorch_pretrained_bert's warmup_linear (config.yaml:131): linear ramp to 1 over the warm-up fraction, then linear decay."""
    from news_image_caption_b200.trainer import warmup_linear
    assert warmup_linear(0.0, 0.05) == 0.0 and abs(warmup_linear(0.025, 0.05) - 0.5) < 1e-12
    assert warmup_linear(0.05, 0.05) == (0.05 - 1.0) / (0.05 - 1.0) and warmup_linear(1.0, 0.05) == 0.0
    assert warmup_linear(0.5, 0.05) > warmup_linear(0.9, 0.05) > 0.0


def test_bench_row_sampling_and_peak_choice():
    import bench
    from oracle import synth
    prev, tgt, ctx = synth.make_batch(64, 32, 50265, ('image', 'article'), seed=1234)
    rows = bench.sample_rows(ctx, tgt, 4)
    live = (~ctx['article_mask']).sum(1)
    assert len(rows) == len(set(rows)) and rows == sorted(rows) and len(rows) >= 3
    assert int(live.argmax()) in rows and int(live.argmin()) in rows
    sub = bench.sub_batch(ctx, rows)
    assert sub['article'].shape[1] == len(rows) and sub['article_mask'].shape[0] == len(rows)
    assert torch.equal(sub['article'][:, 0], ctx['article'][:, rows[0]])
    # burst peak when the SM clock sat at its maximum, the sustained one when the part had settled lower
    pk, src = bench.peak_for({'sm_mhz': 1965, 'sm_max_mhz': 1965}, 1683.1, 1396.3, 'measured')
    assert pk == 1683.1 and 'burst' in src
    pk, src = bench.peak_for({'sm_mhz': 1327, 'sm_max_mhz': 1965}, 1683.1, 1396.3, 'measured')
    assert pk == 1396.3 and 'sustained' in src


def test_exchange_accounting_and_sparse_row_layout():
    """GradientExchange.bytes_on_wire (what bench.py prints as the collective's size) and the byte layout of the packed
    embedding-row buffer (rows | meta | count) the engine writes into."""
    from news_image_caption_b200.trainer import GradientExchange, SparseRowExchange
    flat = torch.zeros(64 * 100)
    ex = GradientExchange(flat, [(0, 640), (None, None), (640, 6400)], wire='bf16')
    assert ex.bytes_on_wire() == 6400 * 2
    assert GradientExchange(flat, [(0, 6400)], wire='fp32').bytes_on_wire() == 6400 * 4
    sx = SparseRowExchange(10, 16, torch.bfloat16, 'cpu')
    rows, meta, count = sx.views(sx.mine)
    assert rows.shape == (10, 16) and rows.dtype == torch.bfloat16 and meta.shape == (10,) and meta.dtype == torch.int64
    assert count.dtype == torch.int32 and sx.bytes == 10 * 16 * 2 + 10 * 8 + 64
    rows.fill_(1.0)
    meta.fill_((2 << 32) | 5)
    count[0] = 7
    assert sx.mine[:320].view(torch.bfloat16).float().sum().item() == 160.0 and int(sx.mine[320:328].view(torch.int64)[0]) == (2 << 32) | 5


def test_round2_struct_layouts():
    """ctypes mirrors of the structs that grew this round: field offsets the C side relies on."""
    from news_image_caption_b200 import _lib
    assert _lib.EmbedParams.vocab.offset == _lib.EmbedParams.dtype.offset + 4
    assert _lib.EmbedParams.error_flags.offset % 8 == 0
    assert _lib.DecoderConfig.embed_padding_idx.offset == _lib.DecoderConfig.dtype.offset + 4
    assert _lib.LayernormBwdParams.variant.offset == _lib.LayernormBwdParams.dtype.offset + 4
    assert _lib.XattnBwdParams.dweights.offset == _lib.XattnBwdParams.scratch_bytes.offset + C.sizeof(C.c_size_t)
    L = _lib.lib()                                   # loads and binds every symbol (no GPU needed)
    for name in ('tnt_decoder_forward_incremental', 'tnt_decoder_project_kv', 'tnt_decoder_train_forward', 'tnt_decoder_train_backward',
                 'tnt_decoder_train_forward_ext', 'tnt_decoder_target_logprob', 'tnt_decoder_train_backward_ext',
                 'tnt_decoder_set_sparse_embed_grad', 'tnt_embed_grad_scatter', 'tnt_ce_bwd', 'tnt_embedding_bwd', 'tnt_adam_step',
                 'tnt_grad_accumulate', 'tnt_variant_count', 'tnt_decoder_error_flags', 'tnt_cast_background'):
        assert hasattr(L, name)
    assert L.tnt_variant_count(b'gemm.tc.pair') == 0 and L.tnt_variant_count(b'no.such.variant') == -1


def test_linear_ln_support_rule_is_host_logic():
    """tnt_linear_ln_supported (gemm_ln.cu): which GEMM -> +residual -> LayerNorm sites get the one-launch cluster kernel is
    decided on the host from the parameter block alone (no launch, no GPU needed: the cluster-occupancy query falls back to an
    estimate): bf16 NT problems whose row is C <= 8 tiles of 64 / 128 / 256 columns, no dropout, no device row count."""
    import ctypes as C
    from news_image_caption_b200 import _lib
    L = _lib.lib()
    for name in ('tnt_linear_ln_fwd', 'tnt_linear_ln_supported', 'tnt_debug_ln_fused', 'tnt_debug_ln_tile_n', 'tnt_debug_kv_beside',
                 'tnt_debug_kv_beside_get', 'tnt_debug_kv_inline'):
        assert hasattr(L, name)
    assert L.tnt_variant_count(b'gemm.ln_fused') == 0 and L.tnt_variant_count(b'kv.beside') == 0
    buf = (C.c_char * 4096)()
    base = (C.addressof(buf) + 255) & ~255             # a 256-byte aligned fake address: only alignment is looked at

    def params(M=2048, N=1024, K=1024, dtype=_lib.TNT_BF16, epilogue=_lib.EPI_RESIDUAL, groups=1, **kw):
        p = _lib.LinearParams()
        p.a = p.w = p.c = p.residual = base
        p.lda, p.ldw, p.ldc, p.ldr = groups * K, K, groups * N, N
        p.M, p.N, p.K, p.groups = M, N, K, groups
        if groups > 1:
            p.stride_a, p.stride_w, p.stride_bias, p.stride_c = K, N * K, N, N
        p.dtype_in = p.dtype_out = dtype
        p.epilogue = epilogue
        for k, v in kw.items():
            setattr(p, k, v)
        return p
    ok = lambda p: bool(L.tnt_linear_ln_supported(C.byref(p), base, base))
    assert ok(params())                                                  # a layer site at B*T = 2048
    assert ok(params(M=256, groups=2)) and ok(params(M=100, N=512)) and ok(params(M=130, N=2048, K=512))
    assert ok(params(epilogue=_lib.EPI_NONE))
    assert not ok(params(dtype=_lib.TNT_F32))                            # not a tcgen05 problem
    assert not ok(params(N=1000)) and not ok(params(N=4096))             # not C <= 8 tiles of 64 / 128 / 256 columns
    assert not ok(params(epilogue=_lib.EPI_RELU))
    assert not ok(params(layout=1))                                      # NN / TN operands
    assert not ok(params(m_dyn=base))                                    # device-side row count
    assert not ok(params(ldc=1028))                                      # rows without 16-byte alignment
    pd = params()
    pd.drop.p = 0.1                                                      # training dropout: two launches
    assert not ok(pd)
    pa = params()
    pa.residual = base + 2
    assert not ok(pa)
    assert not L.tnt_linear_ln_supported(C.byref(params()), None, base)  # gamma missing


def test_bench_roofline_arithmetic():
    """bench.py kv_roofline: achieved = algorithmic FLOPs of ONE launch / that launch's mean duration.  The forward issues the
    article K/V projection once for all layers (N = L * 2d); with the per-layer schedules (TNT_KV_BESIDE_SMS / TNT_KV_INLINE)
    it is four launches of a quarter of the columns -- the FLOPs per launch must follow, share_of_step must count all four."""
    import bench
    hbm, tf_burst, tf_sus, which = bench.peaks()
    clocks = {'sm_mhz': 1965, 'sm_max_mhz': 1965}
    one = bench.kv_roofline(64, {'prof': {'kv_proj.article': (20 * 0.4, 20)}, 'clocks': clocks, 'steps': 20}, 1.5)
    assert one['shape'] == [512 * 64, 8192, 1024] and one['launches_timed'] == 20
    assert abs(one['achieved'] - 2.0 * 512 * 64 * 8192 * 1024 / 0.4e-3 / 1e12) < 1e-6
    assert abs(one['frac'] - one['achieved'] / tf_burst) < 1e-9 and abs(one['share_of_step'] - 0.4 / 1.5) < 1e-9
    assert 'sm_share' not in one
    four = bench.kv_roofline(64, {'prof': {'kv_proj.article': (80 * 0.15, 80)}, 'clocks': clocks, 'steps': 20}, 1.5)
    assert four['shape'] == [512 * 64, 2048, 1024]
    assert abs(four['achieved'] - 2.0 * 512 * 64 * 2048 * 1024 / 0.15e-3 / 1e12) < 1e-6
    assert abs(four['share_of_step'] - 4 * 0.15 / 1.5) < 1e-9
    # a part that settled well below its maximum clock is compared with the sustained peak
    slow = bench.kv_roofline(64, {'prof': {'kv_proj.article': (20 * 0.5, 20)}, 'clocks': {'sm_mhz': 1300, 'sm_max_mhz': 1965}, 'steps': 20}, 1.5)
    assert abs(slow['peak'] - tf_sus) < 1e-9
