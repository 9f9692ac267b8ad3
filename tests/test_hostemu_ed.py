"""The edwards25519 DEVICE headers (csrc/fe25519.cuh, ed25519.cuh) compiled for the host, where the PTX carry-flag
wrappers are modelled in C: the exact instruction sequences of the sm_100a kernels, checked on the CPU against the
big-integer definition oracle/ed25519_ref.py (itself pinned to RFC 8032)."""
import ctypes

import numpy as np
import pytest

import ed25519_ref as E


@pytest.fixture(scope='module')
def emu():
    from drynx_b200 import build
    lib = ctypes.CDLL(build.build_hostemu())
    lib.emu_ed_scalar_i64.argtypes = [ctypes.c_int64, ctypes.c_void_p]
    return lib


def _limbs(x):
    return np.array([(x >> (32 * k)) & 0xffffffff for k in range(8)], np.uint32)


def _val(a):
    return sum(int(v) << (32 * k) for k, v in enumerate(a))


def _fe(emu, op, a, b=0):
    A, B, Rr = _limbs(a), _limbs(b), np.zeros(8, np.uint32)
    emu.emu_fe25519(A.ctypes.data, B.ctypes.data, Rr.ctypes.data, op)
    return _val(Rr)


def test_field_ops_on_weakly_reduced_inputs(emu, rng):
    p = E.p
    edge = [0, 1, 2, 18, 19, 37, 38, p - 1, p, p + 1, p + 18, 2 * p - 1, 2 * p, 2 * p + 37, 2 ** 255 - 1, 2 ** 255, 2 ** 256 - 1,
            2 ** 256 - 38, 2 ** 256 - 39, 2 ** 32 - 1, 2 ** 224]
    vals = edge + [rng.scalar() | (rng.u64() << 192) for _ in range(60)]
    for a in vals:
        assert _fe(emu, 3, a) == a % p                       # canonical form
        assert _fe(emu, 6, a) % p == a * a % p               # dedicated squaring
        for b in edge[:19:3] + vals[-6:]:
            assert _fe(emu, 0, a, b) % p == a * b % p
            assert _fe(emu, 1, a, b) % p == (a + b) % p
            assert _fe(emu, 2, a, b) % p == (a - b) % p
    for a in vals[:30]:
        assert _fe(emu, 4, a) == (pow(a % p, -1, p) if a % p else 0)
        assert _fe(emu, 5, a) % p == pow(a % p, (p - 5) // 8, p)


def test_scalars(emu):
    out = ctypes.create_string_buffer(32)
    for m in [0, 1, -1, 255, -256, 2 ** 63 - 1, -(2 ** 63), 123456789, -987654321]:
        emu.emu_ed_scalar_i64(m, out)
        assert int.from_bytes(out.raw, 'little') == m % E.l


def test_point_ops_match_definition(emu, rng):
    B = E.to_bytes(E.BASE)
    out = ctypes.create_string_buffer(32)
    ks = [0, 1, 2, 255, 256, E.l - 1, E.l, E.l + 7, 2 ** 252, 2 ** 253 - 1, 2 ** 256 - 1, rng.scalar(), rng.scalar() >> 3]
    for k in ks:
        want = E.to_bytes(E.mul(k, E.BASE))
        assert emu.emu_ed_mul(k.to_bytes(32, 'little'), B, out, 0) == 0 and out.raw == want
    for k in ks[3:6] + ks[-2:]:
        assert emu.emu_ed_mul(k.to_bytes(32, 'little'), B, out, 1) == 0 and out.raw == E.to_bytes(E.mul(k, E.BASE))
    P = E.mul(rng.scalar(), E.BASE)
    Q = E.mul(rng.scalar(), E.BASE)
    for A, C in [(P, Q), (P, P), (P, E.neg(P)), (E.IDENT, P), (E.IDENT, E.IDENT), (Q, E.IDENT)]:
        assert emu.emu_ed_add(E.to_bytes(A), E.to_bytes(C), out) == 0 and out.raw == E.to_bytes(E.add(A, C))
    k = rng.scalar()
    assert emu.emu_ed_mul(k.to_bytes(32, 'little'), E.to_bytes(P), out, 0) == 0 and out.raw == E.to_bytes(E.mul(k, P))


def test_decoding_rules(emu):
    out = ctypes.create_string_buffer(32)
    one = (1).to_bytes(32, 'little')
    assert emu.emu_ed_mul(one, (2).to_bytes(32, 'little'), out, 0) == -1          # y = 2: no x
    assert emu.emu_ed_mul(one, (E.p + 1).to_bytes(32, 'little'), out, 0) == 0 and out.raw == E.to_bytes(E.IDENT)
    assert emu.emu_ed_mul(one, (1 | 1 << 255).to_bytes(32, 'little'), out, 0) == 0 and out.raw == E.to_bytes(E.IDENT)
    # every decodable y below 64, both signs: round trip through decode -> 1*P -> encode
    for y in range(64):
        for sign in (0, 1):
            enc = (y | sign << 255).to_bytes(32, 'little')
            P = E.from_bytes(enc)
            rc = emu.emu_ed_mul(one, enc, out, 0)
            if P is None:
                assert rc == -1
            else:
                # a small-order or non-prime-order point times (1 mod l) is itself
                assert rc == 0 and out.raw == E.to_bytes(P)
