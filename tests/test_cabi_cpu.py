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


def test_dropout_generator_restatement(built):
    """tnt_dropout_bits is a HOST function of the C ABI: the oracle's numpy restatement (oracle/restate.py dropout_bits)
    must agree bit for bit -- it is what lets the training-step tests feed the kernels' masks to the oracle."""
    import numpy as np
    from news_image_caption_b200 import _lib
    from oracle import restate as O
    L = _lib.lib()
    rng = np.random.default_rng(0)
    idx = np.concatenate([np.arange(64, dtype=np.uint64), rng.integers(0, 2 ** 63, 200, dtype=np.uint64),
                          np.array([2 ** 32 - 1, 2 ** 32, 2 ** 64 - 1], dtype=np.uint64)])
    for seed, site in ((0, 0), (0x1234567890ABCDEF, 35), (2 ** 64 - 1, 0xFFFF), (42, 6 + 16 * 3)):
        want = [L.tnt_dropout_bits(seed, site, int(i)) for i in idx]
        assert O.dropout_bits(seed, site, idx).tolist() == want
    bits = O.dropout_bits(7, 3, np.arange(1 << 20))
    assert abs((bits < np.uint32(int(0.1 * 2 ** 32))).mean() - 0.1) < 2e-3          # keep rate
    assert abs(np.corrcoef(bits[:-1].astype(np.float64), bits[1:].astype(np.float64))[0, 1]) < 5e-3   # neighbours uncorrelated
    m = O.dropout_mask(0.1, 7, 3, __import__('torch').arange(1000))
    assert set(m.unique().tolist()) <= {0.0, float(np.float32(1.0) / (np.float32(1.0) - np.float32(0.1)))}
