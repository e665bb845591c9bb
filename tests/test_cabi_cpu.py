"""CPU: the C-ABI library builds for sm_100a, loads, and exports every symbol include/tnt_b200.h
declares (no compute calls without a GPU)."""
import os
import re

import pytest

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))


@pytest.fixture(scope='module')
def built():
    from news_image_caption_b200 import build
    return build.build()


def test_header_symbols_exported(built):
    from news_image_caption_b200 import _lib
    hdr = open(os.path.join(ROOT, 'include', 'tnt_b200.h')).read()
    hdr = re.sub(r'/\*.*?\*/', '', hdr, flags=re.S)
    declared = set(re.findall(r'\b(tnt_[a-z0-9_]+)\s*\(', hdr))
    assert len(declared) > 25
    L = _lib.lib()                       # binds every symbol of _SIGS or raises
    for name in declared:
        assert hasattr(L, name), name
    assert declared <= set(_lib.exported_symbols()), declared - set(_lib.exported_symbols())
    assert L.tnt_version() >= 100


def test_no_device_is_loud(built):
    import torch
    from news_image_caption_b200 import _lib
    if torch.cuda.is_available():
        pytest.skip('GPU present')
    with pytest.raises(_lib.TntError):
        _lib.require_device()


def test_product_never_imports_oracle():
    pkg = os.path.join(ROOT, 'news_image_caption_b200')
    for dp, _, fs in os.walk(pkg):
        for f in fs:
            if f.endswith(('.py', '.cu', '.cuh', '.h')):
                src = open(os.path.join(dp, f)).read()
                assert 'oracle' not in src.replace('the oracle', ''), os.path.join(dp, f)
