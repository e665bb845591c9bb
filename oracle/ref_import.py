"""TEST INFRASTRUCTURE ONLY -- imports the UNMODIFIED reference modules from /root/reference.

This file exists only inside the build container (the GPU box has no /root/reference).
It is used by `oracle/make_golden.py` (fixture generation) and by the `not gpu` tests that
pin `oracle/restate.py` against the real reference.  Nothing in the product package
(`news_image_caption_b200/`) may import it.

Recipe: the hot-path files of `ttl/tell` need only *base-class names*
from AllenNLP (`Registrable`, `TokenEmbedder`, `TextFieldEmbedder`) and `overrides`; we register
tiny stub modules for those, plus empty namespace packages for `tell` / `tell.models` so that
their heavy `__init__`s (MongoDB readers, fairseq hub, every model) never execute.
"""
import importlib
import os
import sys
import types

import torch.nn as nn

REF_ROOT = os.environ.get('TNT_REFERENCE_ROOT', '/root/reference')


def reference_available() -> bool:
    return os.path.isdir(os.path.join(REF_ROOT, 'ttl', 'tell', 'modules'))


def _mod(name, **attrs):
    m = types.ModuleType(name)
    m.__dict__.update(attrs)
    sys.modules[name] = m
    return m


def install_stubs():
    if 'tell' in sys.modules and getattr(sys.modules['tell'], '_tnt_stub', False):
        return

    class Registrable:
        @classmethod
        def register(cls, name, *a, **k):
            return lambda c: c

    class Params(dict):
        pass

    class ConfigurationError(Exception):
        pass

    class Vocabulary:
        pass

    class TokenEmbedder(nn.Module, Registrable):
        pass

    class TextFieldEmbedder(nn.Module, Registrable):
        pass

    class TimeDistributed(nn.Module):
        def __init__(self, m):
            super().__init__()
            self._module = m

    _mod('overrides', overrides=lambda f: f)
    _mod('allennlp')
    _mod('allennlp.common', Registrable=Registrable, Params=Params)
    _mod('allennlp.common.registrable', Registrable=Registrable)
    _mod('allennlp.common.checks', ConfigurationError=ConfigurationError)
    _mod('allennlp.data', Vocabulary=Vocabulary)
    _mod('allennlp.modules')
    _mod('allennlp.modules.token_embedders', TokenEmbedder=TokenEmbedder)
    _mod('allennlp.modules.token_embedders.token_embedder', TokenEmbedder=TokenEmbedder)
    _mod('allennlp.modules.text_field_embedders', TextFieldEmbedder=TextFieldEmbedder)
    _mod('allennlp.modules.text_field_embedders.text_field_embedder',
         TextFieldEmbedder=TextFieldEmbedder)
    _mod('allennlp.modules.time_distributed', TimeDistributed=TimeDistributed)

    tell = _mod('tell', _tnt_stub=True)
    tell.__path__ = [os.path.join(REF_ROOT, 'ttl', 'tell')]
    models = _mod('tell.models')
    models.__path__ = [os.path.join(REF_ROOT, 'ttl', 'tell', 'models')]


def load_tell():
    """Returns a namespace with the reference hot-path classes."""
    install_stubs()
    import tell.utils  # noqa: F401
    import tell.modules as M
    from tell.modules.criteria import AdaptiveLoss
    from tell.modules.token_embedders import (AdaptiveEmbedding, SinusoidalPositionalEmbedding,
                                              SumTextFieldEmbedder)
    dec2 = importlib.import_module('tell.models.decoder_flattened')
    dec4 = importlib.import_module('tell.models.decoder_faces_objects')
    decp = importlib.import_module('tell.models.decoder_pointer')
    dece = importlib.import_module('tell.models.decoder_entity_pointer')
    ns = types.SimpleNamespace(
        DynamicConvDecoderPointer=decp.DynamicConvDecoderPointer,
        DynamicConvDecoderEntityPointer=dece.DynamicConvDecoderEntityPointer,
        DynamicConv1dTBC=M.DynamicConv1dTBC, MultiHeadAttention=M.MultiHeadAttention,
        AdaptiveSoftmax=M.AdaptiveSoftmax, GehringLinear=M.GehringLinear,
        AdaptiveLoss=AdaptiveLoss, AdaptiveEmbedding=AdaptiveEmbedding,
        SinusoidalPositionalEmbedding=SinusoidalPositionalEmbedding,
        SumTextFieldEmbedder=SumTextFieldEmbedder,
        DynamicConvDecoder=dec2.DynamicConvDecoder,
        DynamicConvDecoderLayer=dec2.DynamicConvDecoderLayer,
        DynamicConvFacesObjectsDecoder=dec4.DynamicConvFacesObjectsDecoder,
    )
    return ns


def load_faces_objects_model_class():
    """The unmodified `TransformerFacesObjectModel` class (ttl/tell/models/transformer_faces_objects.py) with the
    AllenNLP / pycocoevalcap base classes stubbed: only its `_forward` (context assembly, :312-397) is used, on an
    instance whose RoBERTa / ResNet members are replaced by fakes (the hubs are not reachable here)."""
    install_stubs()
    import torch.nn as nn

    class Model(nn.Module):
        @classmethod
        def register(cls, name, *a, **k):
            return lambda c: c

    if 'allennlp.models.model' not in sys.modules:
        _mod('allennlp.data.vocabulary', Vocabulary=sys.modules['allennlp.data'].Vocabulary)
        _mod('allennlp.models')
        _mod('allennlp.models.model', Model=Model)
        _mod('allennlp.nn')
        _mod('allennlp.nn.initializers', InitializerApplicator=object)
        _mod('pycocoevalcap')
        _mod('pycocoevalcap.bleu')
        _mod('pycocoevalcap.bleu.bleu_scorer', BleuScorer=object)
    mod = importlib.import_module('tell.models.transformer_faces_objects')
    return mod.TransformerFacesObjectModel


def load_entity_pointer_model_class():
    """The unmodified `TransformerEntityPointerModel` class (ttl/tell/models/transformer_entity_pointer.py) with the
    AllenNLP / pycocoevalcap bases stubbed: its `output_layer` (:185-261) and `_generate` (:375-470) are run on an
    instance whose decoder is the real reference decoder and whose RoBERTa / ResNet members are absent."""
    load_faces_objects_model_class()      # registers the allennlp.models / pycocoevalcap stubs
    mod = importlib.import_module('tell.models.transformer_entity_pointer')
    return mod.TransformerEntityPointerModel


def load_context_pointer_model_class():
    """The unmodified model class of ttl/tell/models/transformer_context_pointer.py (the model that owns
    DynamicConvDecoderPointer); only its `_generate` (:467-560) is run, on an instance holding the reference decoder."""
    load_faces_objects_model_class()
    mod = importlib.import_module('tell.models.transformer_context_pointer')
    return mod.TransformerContextPointerModel


def load_final_v2(variant='final_roberta2'):
    path = os.path.join(REF_ROOT, variant)
    name = 'ref_transformer_v2_' + variant
    if name in sys.modules:
        return sys.modules[name]
    install_stubs()
    spec = importlib.util.spec_from_file_location(name, os.path.join(path, 'transformer_v2.py'))
    mod = importlib.util.module_from_spec(spec)
    sys.modules[name] = mod
    sys.path.insert(0, path)
    try:
        spec.loader.exec_module(mod)
    finally:
        sys.path.remove(path)
    return mod


def build_ref_decoder(n_sources=2, d=1024, heads=16, ffn=4096, kernels=(3, 7, 15, 31),
                      vocab=50265, cutoff=(5000, 20000), seed=0, article_dim=1024, relu_dropout=0.0):
    """Reference decoder at the YAML hyper-parameters
    (ttl/expt/nytimes/9_transformer_objects/config.yaml:28-76), dropout irrelevant in eval()."""
    import torch
    R = load_tell()
    torch.manual_seed(seed)
    ada = R.AdaptiveEmbedding(None, 'bpe', padding_idx=0, initial_dim=d, factor=1, output_dim=d,
                              cutoff=list(cutoff), vocab_size=vocab, scale_embeds=True)
    pos = R.SinusoidalPositionalEmbedding(None, embedding_dim=d, padding_idx=1, left_pad=False,
                                          init_size=512)
    emb = R.SumTextFieldEmbedder({'adaptive': ada, 'position': pos},
                                 {'adaptive': ['roberta'], 'position': ['roberta']}, True)
    common = dict(vocab=None, embedder=emb, max_target_positions=512, dropout=0.1,
                  share_decoder_input_output_embed=False, decoder_output_dim=d,
                  decoder_conv_dim=d, decoder_glu=True, decoder_conv_type='dynamic',
                  weight_softmax=True, decoder_attention_heads=heads, weight_dropout=0.1,
                  relu_dropout=relu_dropout, input_dropout=0.1, decoder_normalize_before=False,
                  attention_dropout=0.1, decoder_ffn_embed_dim=ffn,
                  decoder_kernel_size_list=list(kernels), adaptive_softmax_cutoff=list(cutoff),
                  tie_adaptive_weights=True, adaptive_softmax_dropout=0, tie_adaptive_proj=False,
                  adaptive_softmax_factor=1, decoder_layers=len(kernels), final_norm=False,
                  padding_idx=0, vocab_size=vocab)
    if n_sources == 2:
        dec = R.DynamicConvDecoder(article_embed_size=article_dim, **common)
    elif n_sources == 4:
        dec = R.DynamicConvFacesObjectsDecoder(**common)
    elif n_sources == 'pointer':      # 2 sources + copy head (decoder_pointer.py)
        dec = R.DynamicConvDecoderPointer(article_embed_size=article_dim, **common)
    elif n_sources == 'entity_pointer':   # image + article + entity, returns the entity attention (decoder_entity_pointer.py)
        dec = R.DynamicConvDecoderEntityPointer(article_embed_size=article_dim, **common)
    else:
        raise ValueError(n_sources)
    dec.eval()
    return dec
