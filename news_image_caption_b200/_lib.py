"""ctypes binding of libtnt_b200.so (include/tnt_b200.h).  No torch types cross this boundary: only raw
device pointers, sizes and the CUDA stream handle.  Loading fails loudly -- there is no CPU fallback."""
import ctypes as C
import os

HERE = os.path.dirname(os.path.abspath(__file__))
LIB_PATH = os.environ.get('TNT_B200_LIB') or os.path.join(HERE, 'libtnt_b200.so')     # (override: A/B of two builds)

TNT_F32, TNT_BF16 = 0, 1
EPI_NONE, EPI_RELU, EPI_GLU, EPI_RESIDUAL, EPI_RELU_BWD = 0, 1, 2, 3, 5
LAYOUT_NT, LAYOUT_NN, LAYOUT_TN = 0, 1, 2
MASK_NONE, MASK_KEY_PADDING, MASK_V2 = 0, 1, 2
MAX_SOURCES, MAX_LAYERS, MAX_CLUSTERS = 4, 16, 4

vp, i64, i32, f32 = C.c_void_p, C.c_int64, C.c_int, C.c_float


class TntError(RuntimeError):
    pass


class Dropout(C.Structure):
    """tnt_dropout: counter-based dropout site (p == 0: off)."""
    _fields_ = [('p', f32), ('site', C.c_uint32), ('seed', C.c_uint64), ('ld', i64), ('group_stride', i64)]


class DropoutConfig(C.Structure):
    """tnt_dropout_config: the decoder's training dropout (config.yaml:51-69)."""
    _fields_ = [('dropout', f32), ('input_dropout', f32), ('weight_dropout', f32), ('attention_dropout', f32),
                ('relu_dropout', f32), ('seed', C.c_uint64)]


DROP_INPUT, DROP_TAPS, DROP_CONV_OUT, DROP_ATTN_OUT, DROP_FFN_OUT, DROP_RELU, DROP_ATTN_PROB = range(7)
DROP_SITE_EMBED = 0xFFFF


def drop_site(layer, kind):
    return layer * 16 + kind


class PointerParams(C.Structure):
    _fields_ = [('x', vp), ('ldx', i64), ('inp_embed', vp), ('lde', i64), ('w_pgen', vp), ('b_pgen', f32), ('attn', vp),
                ('attn_ld', i64), ('article_ids', vp), ('logprob', vp), ('ldp', i64), ('p_gen', vp), ('B', i32), ('T', i32),
                ('S', i32), ('d', i32), ('V', i32), ('dtype', i32), ('best_tok', vp), ('best_logp', vp)]


class PointerHead(C.Structure):
    _fields_ = [('w_pgen', vp), ('b_pgen', f32), ('w_hidden', vp), ('b_hidden', vp), ('d_hidden', i32), ('copy_ids', vp),
                ('S_copy', i32)]


class LinearParams(C.Structure):
    _fields_ = [('a', vp), ('lda', i64), ('w', vp), ('ldw', i64), ('bias', vp), ('residual', vp), ('ldr', i64),
                ('c', vp), ('ldc', i64), ('M', i32), ('N', i32), ('K', i32), ('m_dyn', vp), ('groups', i32),
                ('stride_a', i64), ('stride_w', i64), ('stride_bias', i64), ('stride_res', i64), ('stride_c', i64),
                ('dtype_in', i32), ('dtype_out', i32), ('epilogue', i32), ('layout', i32), ('alpha', f32),
                ('prefetch', vp), ('prefetch_bytes', C.c_uint64), ('drop', Dropout)]


class DynconvParams(C.Structure):
    _fields_ = [('x', vp), ('ldx', i64), ('logits', vp), ('ldl', i64), ('y', vp), ('ldy', i64), ('hist', vp),
                ('hist_len', i32), ('T', i32), ('B', i32), ('C', i32), ('H', i32), ('K', i32), ('dtype', i32),
                ('drop', Dropout)]


class DynconvBwdParams(C.Structure):
    _fields_ = [('x', vp), ('ldx', i64), ('logits', vp), ('ldl', i64), ('dy', vp), ('lddy', i64), ('dx', vp),
                ('lddx', i64), ('dlogits', vp), ('lddl', i64), ('T', i32), ('B', i32), ('C', i32), ('H', i32),
                ('K', i32), ('dtype', i32), ('drop', Dropout)]


class DynconvStepParams(C.Structure):
    _fields_ = [('x', vp), ('ldx', i64), ('logits', vp), ('ldl', i64), ('y', vp), ('ldy', i64), ('ring', vp),
                ('pos', vp), ('B', i32), ('C', i32), ('H', i32), ('K', i32), ('dtype', i32)]


class LayernormParams(C.Structure):
    _fields_ = [('x', vp), ('ldx', i64), ('y', vp), ('ldy', i64), ('gamma', vp), ('beta', vp), ('rows', i32),
                ('C', i32), ('groups', i32), ('variant', i32), ('eps', f32), ('dtype', i32), ('transpose_T', i32),
                ('transpose_B', i32)]


class LayernormBwdParams(C.Structure):
    _fields_ = [('x', vp), ('ldx', i64), ('dy', vp), ('lddy', i64), ('dx', vp), ('lddx', i64), ('gamma', vp),
                ('dgamma', vp), ('dbeta', vp), ('rows', i32), ('C', i32), ('eps', f32), ('dtype', i32), ('variant', i32),
                ('zero_param_grads', i32)]


class EmbedParams(C.Structure):
    _fields_ = [('ids', vp), ('table', vp), ('pos_table', vp), ('out', vp), ('ldo', i64), ('start_pos', vp),
                ('B', i32), ('T', i32), ('d', i32), ('n_pos', i32), ('pad_idx', i32), ('dtype', i32), ('vocab', i32),
                ('error_flags', vp)]


class ContextMixParams(C.Structure):
    _fields_ = [('layers', vp * 32), ('weight_logits', vp), ('out', vp), ('n_layers', i32), ('B', i32), ('S', i32),
                ('E', i32), ('dtype_in', i32), ('dtype_out', i32)]


class NanMaskParams(C.Structure):
    _fields_ = [('x', vp), ('out', vp), ('mask', vp), ('B', i32), ('n', i32), ('d', i32), ('dtype_in', i32),
                ('dtype_out', i32)]


class XattnParams(C.Structure):
    _fields_ = [('q', vp), ('q_stride_t', i64), ('q_stride_b', i64), ('k', vp), ('k_stride_s', i64),
                ('k_stride_b', i64), ('v', vp), ('v_stride_s', i64), ('v_stride_b', i64), ('out', vp),
                ('o_stride_t', i64), ('o_stride_b', i64), ('mask', vp), ('m_stride_b', i64), ('m_stride_t', i64),
                ('bias_k', vp), ('bias_v', vp), ('weights', vp), ('B', i32), ('T', i32), ('S', i32), ('H', i32),
                ('D', i32), ('n_extra', i32), ('mask_mode', i32), ('scale', f32), ('dtype', i32), ('mask_S', i32), ('lse', vp), ('kv_group', i32),
                ('drop', Dropout), ('tile_live', vp)]


class XattnBwdParams(C.Structure):
    _fields_ = [('fwd', XattnParams), ('dout', vp), ('dq', vp), ('dk', vp), ('dv', vp), ('scratch', vp),
                ('scratch_bytes', C.c_size_t), ('dweights', vp)]


class AdamParams(C.Structure):
    _fields_ = [('params', vp), ('grads', vp), ('m', vp), ('v', vp), ('n', i64), ('seg_of_chunk', vp), ('seg_sq', vp),
                ('n_segments', i32), ('grad_scale_dev', vp), ('grad_scale', f32), ('max_grad_norm', f32), ('lr', f32),
                ('beta1', f32), ('beta2', f32), ('eps', f32), ('weight_decay', f32), ('norms_ready', i32), ('max_ctas', i32)]


class LseParams(C.Structure):
    _fields_ = [('x', vp), ('ldx', i64), ('w', vp), ('ldw', i64), ('bias', vp), ('target', vp), ('row_map', vp),
                ('m_dyn', vp), ('stats', vp), ('workspace', vp), ('workspace_bytes', C.c_size_t), ('M', i32),
                ('n_cols', i32), ('K', i32), ('dtype', i32)]


class DecoderConfig(C.Structure):
    _fields_ = [('n_layers', i32), ('d', i32), ('heads', i32), ('ffn', i32), ('vocab', i32),
                ('kernels', i32 * MAX_LAYERS), ('n_sources', i32), ('source_names', (C.c_char * 16) * MAX_SOURCES),
                ('source_dim', i32 * MAX_SOURCES), ('n_clusters', i32), ('cutoffs', i32 * MAX_CLUSTERS),
                ('pad_idx', i32), ('max_positions', i32), ('dtype', i32), ('embed_padding_idx', i32)]


class Contexts(C.Structure):
    _fields_ = [('B', i32), ('T', i32), ('S', i32 * MAX_SOURCES), ('memory', vp * MAX_SOURCES),
                ('pad_mask', vp * MAX_SOURCES)]


_SIGS = {
    'tnt_version': (i32, []),
    'tnt_last_error': (C.c_char_p, []),
    'tnt_device_check': (i32, []),
    'tnt_launch_count': (C.c_uint64, []),
    'tnt_variant_count': (C.c_int64, [C.c_char_p]),
    'tnt_prof_enable': (None, [i32]),
    'tnt_prof_reset': (None, []),
    'tnt_prof_read': (i32, [C.c_char_p, C.POINTER(C.c_double), C.POINTER(i32)]),
    'tnt_linear_fwd': (i32, [C.POINTER(LinearParams), vp]),
    'tnt_wn_fold': (i32, [vp, vp, vp, i32, i32, i32, i32, vp]),
    'tnt_glu_interleave_bias': (i32, [vp, vp, i32, vp]),
    'tnt_cast': (i32, [vp, vp, i64, i32, i32, vp]),
    'tnt_cast_background': (i32, [vp, vp, i64, i32, i32, i32, vp]),
    'tnt_dynconv_fwd': (i32, [C.POINTER(DynconvParams), vp]),
    'tnt_dynconv_bwd': (i32, [C.POINTER(DynconvBwdParams), vp]),
    'tnt_dynconv_step': (i32, [C.POINTER(DynconvStepParams), vp]),
    'tnt_layernorm_fwd': (i32, [C.POINTER(LayernormParams), vp]),
    'tnt_linear_ln_supported': (i32, [C.POINTER(LinearParams), vp, vp]),
    'tnt_linear_ln_fwd': (i32, [C.POINTER(LinearParams), vp, vp, i64, f32, i32, i32, vp]),
    'tnt_layernorm_bwd': (i32, [C.POINTER(LayernormBwdParams), vp]),
    'tnt_lse_fwd_static': (i32, [C.POINTER(LseParams), vp]),
    'tnt_lse_fwd_static_max_bytes': (C.c_size_t, []),
    'tnt_grad_sqnorm': (i32, [vp, vp, vp, i64, i64, i32, vp]),
    'tnt_embed_fwd': (i32, [C.POINTER(EmbedParams), vp]),
    'tnt_context_mix': (i32, [C.POINTER(ContextMixParams), vp]),
    'tnt_nan_mask': (i32, [C.POINTER(NanMaskParams), vp]),
    'tnt_xattn_fwd': (i32, [C.POINTER(XattnParams), vp]),
    'tnt_xattn_tile_mask': (i32, [vp, i64, i32, i32, i32, vp, vp]),
    'tnt_xattn_bwd_scratch_bytes': (C.c_size_t, [i32, i32, i32, i32]),
    'tnt_xattn_bwd': (i32, [C.POINTER(XattnBwdParams), vp]),
    'tnt_lse_workspace_bytes': (C.c_size_t, [i32, i32, i32]),
    'tnt_lse_fwd': (i32, [C.POINTER(LseParams), vp]),
    'tnt_ce_bwd_workspace_bytes': (C.c_size_t, [i32, i32, i32]),
    'tnt_ce_bwd': (i32, [vp, i64, vp, i64, vp, vp, vp, i32, i32, i32, i32, f32, vp, i64, vp, vp, i32, vp, C.c_size_t, vp]),
    'tnt_embedding_bwd': (i32, [vp, i32, vp, i64, f32, vp, i64, i32, i32, i32, vp]),
    'tnt_decoder_create': (vp, [C.POINTER(DecoderConfig)]),
    'tnt_decoder_destroy': (None, [vp]),
    'tnt_decoder_weight_bytes': (C.c_size_t, [vp]),
    'tnt_decoder_bind_weights': (i32, [vp, vp, C.c_size_t]),
    'tnt_decoder_load': (i32, [vp, C.c_char_p, vp, C.POINTER(i64), i32, vp]),
    'tnt_decoder_finalize': (i32, [vp, vp]),
    'tnt_decoder_workspace_bytes': (C.c_size_t, [vp, i32, i32, C.POINTER(i32)]),
    'tnt_decoder_loss_workspace_bytes': (C.c_size_t, [vp, i32]),
    'tnt_decoder_forward': (i32, [vp, vp, C.POINTER(Contexts), vp, vp, vp, C.c_size_t, vp]),
    'tnt_decoder_kv_cache_bytes': (C.c_size_t, [vp, i32, i32]),
    'tnt_decoder_project_kv': (i32, [vp, C.POINTER(Contexts), C.POINTER(vp), vp]),
    'tnt_decoder_incremental_workspace_bytes': (C.c_size_t, [vp, i32, i32, C.POINTER(i32)]),
    'tnt_decoder_forward_incremental': (i32, [vp, vp, C.POINTER(Contexts), C.POINTER(vp), i32, C.POINTER(vp), C.POINTER(i32),
                                              C.POINTER(vp), vp, vp, vp, C.c_size_t, vp]),
    'tnt_decoder_nll': (i32, [vp, vp, vp, i32, vp, vp, C.c_size_t, vp]),
    'tnt_decoder_log_prob': (i32, [vp, vp, i32, vp, vp, C.c_size_t, vp]),
    'tnt_decoder_embed': (i32, [vp, vp, i32, i32, vp, vp]),
    'tnt_decoder_error_flags': (i32, [vp, vp]),
    'tnt_pointer_mix': (i32, [C.POINTER(PointerParams), vp]),
    'tnt_nll_rows': (i32, [vp, i64, vp, i32, i32, vp, vp]),
    'tnt_decoder_generate_workspace_bytes': (C.c_size_t, [vp, i32, i32, C.POINTER(i32)]),
    'tnt_decoder_generate_pointer_workspace_bytes': (C.c_size_t, [vp, i32, i32, C.POINTER(i32), i32]),
    'tnt_decoder_generate_pointer': (i32, [vp, vp, C.POINTER(Contexts), C.POINTER(PointerHead), i32, i32, vp, vp,
                                           C.POINTER(i32), vp, C.c_size_t, vp]),
    'tnt_decoder_generate': (i32, [vp, vp, C.POINTER(Contexts), i32, i32, vp, vp, C.POINTER(i32), vp, C.c_size_t, vp]),
    'tnt_decoder_beam_workspace_bytes': (C.c_size_t, [vp, i32, i32, i32, C.POINTER(i32)]),
    'tnt_decoder_generate_beam': (i32, [vp, vp, C.POINTER(Contexts), i32, i32, i32, vp, vp, vp, vp, C.c_size_t, vp]),
    'tnt_decoder_score_host_staging_bytes': (C.c_size_t, [vp, i32, i32, C.POINTER(i32)]),
    'tnt_decoder_score_host': (i32, [vp, vp, vp, C.POINTER(Contexts), vp, vp, C.c_size_t, vp, C.c_size_t, vp]),
    'tnt_decoder_score_host_submit': (i32, [vp, i32, vp, vp, C.POINTER(Contexts), vp, vp, C.c_size_t, vp, C.c_size_t, vp]),
    'tnt_decoder_score_host_wait': (i32, [vp, i32]),
    'tnt_decoder_last_h2d_bytes': (C.c_uint64, [vp]),
    'tnt_h2d_live_rows': (C.c_int64, [vp, vp, vp, i32, i32, i32, i32, vp]),
    'tnt_dropout_bits': (C.c_uint32, [C.c_uint64, C.c_uint32, C.c_uint64]),
    'tnt_dropout_apply': (i32, [vp, i64, vp, i64, vp, i64, i32, i32, i32, C.POINTER(Dropout), vp]),
    'tnt_decoder_set_dropout': (i32, [vp, C.POINTER(DropoutConfig)]),
    'tnt_decoder_train_workspace_bytes': (C.c_size_t, [vp, i32, i32, C.POINTER(i32)]),
    'tnt_decoder_set_grad': (i32, [vp, C.c_char_p, vp]),
    'tnt_decoder_set_sparse_embed_grad': (i32, [vp, vp, vp, vp, i32]),
    'tnt_embed_grad_scatter': (i32, [vp, i64, vp, vp, i32, C.POINTER(vp), C.POINTER(i32), i32, i32, i32, vp]),
    'tnt_decoder_train_step': (i32, [vp, vp, vp, C.POINTER(Contexts), f32, vp, vp, C.c_size_t, vp]),
    'tnt_decoder_train_forward': (i32, [vp, vp, C.POINTER(Contexts), vp, vp, C.c_size_t, vp]),
    'tnt_decoder_train_backward': (i32, [vp, vp, vp, C.POINTER(Contexts), f32, vp, vp, C.c_size_t, vp]),
    'tnt_decoder_train_forward_ext': (i32, [vp, vp, C.POINTER(Contexts), vp, vp, vp, vp, C.c_size_t, vp]),
    'tnt_decoder_target_logprob': (i32, [vp, vp, C.POINTER(Contexts), vp, vp, C.c_size_t, vp]),
    'tnt_decoder_train_backward_ext': (i32, [vp, vp, vp, C.POINTER(Contexts), vp, vp, vp, vp, vp, C.c_size_t, vp]),
    'tnt_decoder_train_stages': (i32, [vp]),
    'tnt_decoder_stream_wait_stage': (i32, [vp, i32, vp]),
    'tnt_bert_adam_step': (i32, [C.POINTER(AdamParams), vp]),
    'tnt_adam_step': (i32, [vp, vp, vp, vp, i64, vp, f32, f32, f32, f32, f32, f32, i32, vp]),
    'tnt_grad_accumulate': (i32, [vp, vp, i64, f32, vp, vp]),
    'tnt_debug_force_simt_gemm': (None, [i32]),
    'tnt_debug_gemm_cta_group': (None, [i32]),
    'tnt_debug_pdl': (None, [i32]),
    'tnt_debug_gemm_tile_n': (None, [i32]),
    'tnt_debug_gemm_stages': (None, [i32]),
    'tnt_debug_gemm_dynamic': (None, [i32]),
    'tnt_debug_gemm_single_acc': (None, [i32]),
    'tnt_debug_gemm_split_k': (None, [i32]),
    'tnt_debug_sub_batches': (None, [i32]),
    'tnt_debug_ln_fused': (None, [i32]),
    'tnt_debug_kv_beside': (None, [i32]),
    'tnt_debug_kv_inline': (None, [i32]),
    'tnt_debug_kv_beside_get': (i32, []),
    'tnt_debug_train_side_streams': (None, [i32]),
    'tnt_debug_force_simt_attention': (None, [i32]),
    'tnt_debug_force_simt_attention_bwd': (None, [i32]),
}

_lib = None


def exported_symbols():
    """Every entry point include/tnt_b200.h declares (plus the one debug hook)."""
    return sorted(_SIGS)


def lib():
    """Loads the CUDA library; raises if it has not been built (python -m news_image_caption_b200.build)."""
    global _lib
    if _lib is None:
        if not os.path.exists(LIB_PATH):
            raise TntError('libtnt_b200.so is missing: run `python news_image_caption_b200/build.py` '
                           '(there is no CPU fallback)')
        L = C.CDLL(LIB_PATH)
        for name, (res, args) in _SIGS.items():
            fn = getattr(L, name)           # AttributeError if the symbol is not exported
            fn.restype, fn.argtypes = res, args
        _lib = L
    return _lib


def check(status, what=''):
    if status != 0:
        msg = lib().tnt_last_error().decode('utf-8', 'replace')
        raise TntError('%s failed (%d): %s' % (what or 'tnt call', status, msg))


def require_device():
    check(lib().tnt_device_check(), 'tnt_device_check')
