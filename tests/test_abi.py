"""The C-ABI library loads here (no GPU), exports every symbol include/drynx_gpu.h declares, and refuses to run
without a device instead of falling back to the CPU."""
import ctypes
import os
import re

import pytest

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))


def _declared():
    text = open(os.path.join(ROOT, 'include', 'drynx_gpu.h')).read()
    text = re.sub(r'/\*.*?\*/', '', text, flags=re.S)
    return sorted(set(re.findall(r'\b(drx_[a-z0-9_]+)\s*\(', text)))


def test_header_symbols_are_exported_and_bound():
    from drynx_b200 import _lib, build
    build.build_gpu()
    lib = _lib.load()
    names = _declared()
    assert len(names) >= 25
    for n in names:
        assert hasattr(lib, n), n
    assert set(names) == set(_lib.SIGNATURES), set(names) ^ set(_lib.SIGNATURES)


def test_product_never_imports_oracle():
    for dirpath, _, files in os.walk(os.path.join(ROOT, 'drynx_b200')):
        for f in files:
            if f.endswith(('.py', '.cu', '.cuh', '.h', '.inc')) and f != 'build.py':
                src = open(os.path.join(dirpath, f)).read()
                assert 'bn256_ref' not in src and 'liboracle' not in src and 'hostemu' not in src.replace('DRX_HOST_EMU', ''), f


def test_init_fails_loudly_without_gpu():
    import torch
    if torch.cuda.is_available():
        pytest.skip('a GPU is present')
    from drynx_b200 import _lib
    lib = _lib.load()
    h = ctypes.c_void_p()
    rc = lib.drx_init(ctypes.byref(h), 0, None, None)
    assert rc != 0 and not h.value
    assert b'no CUDA device' in lib.drx_last_error() or b'CPU' in lib.drx_last_error()
    with pytest.raises(_lib.DrynxGpuError):
        from drynx_b200 import Backend
        Backend(0)


def test_size_helpers_need_no_gpu():
    from drynx_b200 import _lib
    lib = _lib.load()
    assert lib.drx_rp_proof_size(3, 4) == 256 + 32 * 4 + 544 * 3 * 4 == 6912
    assert lib.drx_rp_proof_size(-1, 0) == 128
    assert lib.drx_rp_ndigits(0, 16, 4) == 4 and lib.drx_rp_ndigits(-5, 16, 4) == 4
    assert lib.drx_rp_ndigits(65535, 16, 4) == 4 and lib.drx_rp_ndigits(65536, 16, 4) == 5
