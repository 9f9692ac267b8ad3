"""-m gpu, collected last: the allocation-free aggregation entry point against the ordinary one.

drx_aggregate_noalloc_dev is for hosts that cannot allocate GPU memory once they are running (the whole device
was handed to one pool at start-up, or allocation after initialisation is forbidden because a failing or
device-synchronising cudaMalloc in the middle of a survey is not acceptable). It keeps its partial sums in a fixed
array inside the library, which all calls share and which keeps the previous call's contents, so this test calls it
several times in a row with different shapes and checks every result byte-for-byte against drx_aggregate_dev."""
import ctypes

import numpy as np
import pytest

pytestmark = pytest.mark.gpu


def _vp(t):
    return ctypes.c_void_p(t.data_ptr())


def _encrypt_vectors(backend, seed, ndp, v):
    import torch
    g = np.random.default_rng(seed)
    m = g.integers(-1000, 1000, size=ndp * v, dtype=np.int64)
    r = np.frombuffer(g.bytes(32 * ndp * v), np.uint8).copy()
    r[::32] &= 0x7f
    return torch.from_numpy(backend.encrypt_int_vector(m, r)).cuda()


def test_no_alloc_aggregate_equals_existing_entry_point_on_consecutive_calls(backend):
    import torch
    from drynx_b200 import _lib
    lib = backend.lib
    sk = (123456789).to_bytes(32, 'big')
    g1 = bytes(31) + b'\x01' + (65000549695646603732796438742359905742825358107623003571877145026864184071783 - 2).to_bytes(32, 'big')
    backend.set_pubkey(backend.g1_mul(sk, g1))
    # the documented warm-up call (no input, one column): it is the one that places the library's device code and its
    # fixed array on this GPU; the result of an empty sum is the point at infinity (zeros)
    warm = torch.full((128,), 0xAB, dtype=torch.uint8, device='cuda')
    _lib.check(lib.drx_aggregate_noalloc_dev(0, 1, None, _vp(warm), None))
    torch.cuda.synchronize()
    assert int(warm.max()) == 0
    # shapes chosen so that consecutive calls use different slice counts and leave stale partial sums behind
    for call, (ndp, v) in enumerate([(37, 50), (5, 3), (130, 700), (37, 50), (1, 1)]):
        cts = _encrypt_vectors(backend, 100 + call, ndp, v)
        want = torch.empty(v * 128, dtype=torch.uint8, device='cuda')
        got = torch.full((v * 128,), 0xAB, dtype=torch.uint8, device='cuda')
        _lib.check(lib.drx_aggregate_dev(backend.h, ndp, v, _vp(cts), _vp(want)))
        backend.synchronize()
        torch.cuda.synchronize()
        free_before = torch.cuda.mem_get_info()[0]
        _lib.check(lib.drx_aggregate_noalloc_dev(ndp, v, _vp(cts), _vp(got), None))
        torch.cuda.synchronize()
        # after the warm-up no call may take device memory: the driver's free-byte count does not move
        assert torch.cuda.mem_get_info()[0] == free_before, (call, ndp, v)
        assert torch.equal(got, want), (call, ndp, v)   # bit-exact, the tolerance every test here uses


def test_no_alloc_aggregate_rejects_oversize_and_malformed(backend):
    import torch
    lib = backend.lib
    dummy = torch.zeros(256, dtype=torch.uint8, device='cuda')
    rc = lib.drx_aggregate_noalloc_dev(1, 16385, _vp(dummy), _vp(dummy), None)
    assert rc == -5 and b'16384' in lib.drx_last_error()
    bad = torch.zeros(128, dtype=torch.uint8, device='cuda')
    bad[31] = 5
    bad[63] = 7          # (5, 7) is not on the curve
    out = torch.zeros(128, dtype=torch.uint8, device='cuda')
    assert lib.drx_aggregate_noalloc_dev(1, 1, _vp(bad), _vp(out), None) == -2
    # and the next call is clean again: the counter is cleared on every call
    ok = _encrypt_vectors(backend, 7, 2, 1)
    assert lib.drx_aggregate_noalloc_dev(2, 1, _vp(ok), _vp(out), None) == 0
