"""Runs the DEVICE headers (drynx_b200/csrc/*.cuh) compiled for the host, where the PTX carry-chain wrappers are
modelled in C, against the oracle. This checks the GPU algorithms (Montgomery rows, tower, curve, Miller loop,
final exponentiation, comb tables) on the CPU; the real kernels are checked by the -m gpu tests."""
import ctypes
import os

import numpy as np
import pytest

import bn256_ref as R


@pytest.fixture(scope='module')
def emu():
    from drynx_b200 import build
    return ctypes.CDLL(build.build_hostemu())


def _limbs(x):
    return [(x >> (32 * k)) & 0xffffffff for k in range(8)]


def test_montgomery_rows(emu, rng):
    for which, mod in ((0, R.p), (1, R.n)):
        rinv = pow(1 << 256, -1, mod)
        cases = [(mod - 1, mod - 1), (0, 5), (1, 1), (mod - 1, 2 ** 255), (2 ** 256 - 1 - (2 ** 256 - mod), mod - 2)]
        cases += [(rng.scalar() % mod, rng.scalar() % mod) for _ in range(300)]
        for a, b in cases:
            A = np.array(_limbs(a), np.uint32); B = np.array(_limbs(b), np.uint32); Rr = np.zeros(8, np.uint32)
            emu.emu_fp_mul_raw(A.ctypes.data, B.ctypes.data, Rr.ctypes.data, which)
            got = sum(int(v) << (32 * k) for k, v in enumerate(Rr))
            assert got == a * b * rinv % mod


def test_dual_product_montgomery_pass(emu, rng):
    """fp.cuh mont_mul2: (a*b + c*d) / R mod p in one pass, a, c < p and b, d any 256-bit value (the Fp2 product passes
    p - b1, which is p itself for b1 = 0). The extreme operands drive the accumulator to its 3p bound."""
    mod = R.p
    rinv = pow(1 << 256, -1, mod)
    top = 2 ** 256 - 1
    cases = [(mod - 1, top, mod - 1, top), (mod - 1, mod - 1, mod - 1, mod - 1), (0, 0, 0, 0), (1, 1, 0, 5), (mod - 1, top, 0, 0),
             (0, 7, mod - 1, top), (mod - 1, mod, mod - 1, mod), (5, mod, 7, 0), (mod - 1, 2 ** 255, mod - 2, 2 ** 255 + 12345)]
    cases += [(rng.scalar() % mod, rng.scalar() % (1 << 256), rng.scalar() % mod, (rng.scalar() * 3) % (1 << 256)) for _ in range(600)]
    for a, b, c, d in cases:
        arrs = [np.array(_limbs(v), np.uint32) for v in (a, b, c, d)]
        out = np.zeros(8, np.uint32)
        emu.emu_fp_mul2_raw(arrs[0].ctypes.data, arrs[1].ctypes.data, arrs[2].ctypes.data, arrs[3].ctypes.data, out.ctypes.data)
        got = sum(int(v) << (32 * k) for k, v in enumerate(out))
        assert got == (a * b + c * d) * rinv % mod, (a, b, c, d)


def test_fp2_product_dual_equals_karatsuba(emu, rng):
    """tower.cuh fp2_mul_dual against fp2_mul (Karatsuba) and the definition, incl. zero and p - 1 components."""
    mod = R.p
    rinv = pow(1 << 256, -1, mod)
    edge = [0, 1, mod - 1, mod - 2, 2 ** 255]
    cases = [(a0, a1, b0, b1) for a0 in edge for a1 in (0, mod - 1) for b0 in (0, mod - 1, 5) for b1 in edge]
    cases += [tuple(rng.scalar() % mod for _ in range(4)) for _ in range(300)]
    for a0, a1, b0, b1 in cases:
        A = np.array(_limbs(a0) + _limbs(a1), np.uint32); B = np.array(_limbs(b0) + _limbs(b1), np.uint32)
        d, k = np.zeros(16, np.uint32), np.zeros(16, np.uint32)
        emu.emu_fp2_mul(A.ctypes.data, B.ctypes.data, d.ctypes.data, k.ctypes.data)
        assert (d == k).all()
        c0 = sum(int(v) << (32 * i) for i, v in enumerate(d[:8])); c1 = sum(int(v) << (32 * i) for i, v in enumerate(d[8:]))
        assert c0 == (a0 * b0 - a1 * b1) * rinv % mod and c1 == (a0 * b1 + a1 * b0) * rinv % mod


def test_division_step_inverse(emu, rng):
    """fp.cuh mod_inv_plain (batched division steps, 20 x 30) against Python's pow(a, -1, m) for both moduli, on the
    boundary values and 2000 random residues each; the Montgomery wrapper against a^(p-2) on the same inputs."""
    def call(a, which):
        A = np.array(_limbs(a), np.uint32); Rr = np.zeros(8, np.uint32)
        emu.emu_mod_inv(A.ctypes.data, Rr.ctypes.data, which)
        return sum(int(v) << (32 * k) for k, v in enumerate(Rr))
    for which, mod in ((0, R.p), (1, R.n)):
        edge = [0, 1, 2, 3, mod - 1, mod - 2, (mod + 1) // 2, (mod - 1) // 2, 2 ** 255, 2 ** 254 - 1, 2 ** 30, 2 ** 30 - 1,
                2 ** 60, 2 ** 240, (1 << 255) - 19, mod - 2 ** 30, mod - 2 ** 128]
        edge += [2 ** k for k in range(0, 256, 7)] + [mod - 2 ** k for k in range(1, 255, 11)]
        for a in edge + [rng.scalar() % mod for _ in range(2000)]:
            a %= mod
            got = call(a, which)
            assert got == (pow(a, -1, mod) if a else 0), hex(a)
    for a in [0, 1, R.p - 1] + [rng.scalar() % R.p for _ in range(200)]:
        assert call(a, 2) == call(a, 3)


def test_g1_g2_scalar_mult(emu, oracle, rng):
    for method in (0, 1):
        for k in (1, 255, 256, R.n - 1, rng.scalar()):
            P = oracle.g1_mul(rng.scalar().to_bytes(32, 'big'), oracle.generators()[0])
            out = ctypes.create_string_buffer(64)
            assert emu.emu_g1_mul(k.to_bytes(32, 'big'), P, out, method) == 0
            assert out.raw == oracle.g1_mul(k.to_bytes(32, 'big'), P)
    Q = oracle.g2_mul(rng.scalar().to_bytes(32, 'big'), oracle.generators()[1])
    k = rng.scalar()
    out = ctypes.create_string_buffer(128)
    assert emu.emu_g2_mul(k.to_bytes(32, 'big'), Q, out) == 0
    assert out.raw == oracle.g2_mul(k.to_bytes(32, 'big'), Q)


def test_g1_add_exceptional_cases(emu, rng):
    P = R.g1_mul(5, R.G1_GEN)
    for A, B in [(P, P), (P, R.g1_neg(P)), (None, P), (P, None), (None, None), (P, R.g1_mul(7, R.G1_GEN))]:
        out = ctypes.create_string_buffer(64)
        assert emu.emu_g1_add(R.g1_to_bytes(A), R.g1_to_bytes(B), out) == 0
        assert out.raw == R.g1_to_bytes(R.g1_add(A, B))


def test_pairing_tower_and_gt_pow(emu, oracle, rng):
    g1, g2 = oracle.generators()
    P = oracle.g1_mul(rng.scalar().to_bytes(32, 'big'), g1)
    Q = oracle.g2_mul(rng.scalar().to_bytes(32, 'big'), g2)
    out = ctypes.create_string_buffer(384)
    assert emu.emu_pair(P, Q, out) == 0
    assert out.raw == oracle.pair(P, Q)
    assert emu.emu_tower_selfcheck(out.raw) == 0
    k = rng.scalar()
    o2 = ctypes.create_string_buffer(384)
    assert emu.emu_gt_pow(out.raw, k.to_bytes(32, 'big'), o2) == 0
    assert o2.raw == R.f12_to_bytes(R.f12pow(R.f12_from_bytes(out.raw), k))
    bad = bytes(63) + b'\x05' + bytes(63) + b'\x07'
    assert emu.emu_pair(bad[:64], Q, out) == -1


def test_g2_membership_verdicts(emu, rng):
    """Device headers on the CPU: both forms of the G2 membership test accept points of G2 and reject twist points
    outside it (composite order, cofactor order, order 13), like kyber's [n]Q == O in twistPoint.IsOnCurve."""
    import bn256_ref as R
    P = R.g1_to_bytes(R.g1_mul(rng.scalar(), R.G1_GEN))
    for k in (1, 2, rng.scalar()):
        assert emu.emu_g2_membership(P, R.g2_to_bytes(R.g2_mul(k, R.G2_GEN))) == 3
    T = R.g2_twist_point(7)
    for Q in (T, R.g2_mul_raw(R.n, T), R.g2_mul_raw(R.n * R.TWIST_COFACTOR // 13, T), R.g2_mul_raw(13, T)):
        assert Q is not None and R.g2_on_curve(Q) and not R.g2_in_subgroup(Q)
        assert emu.emu_g2_membership(P, R.g2_to_bytes(Q)) == 0


def test_interpreter_runs_the_emitted_pairing_program(emu, oracle, rng):
    """csrc/pairing_ucode_g4.inc executed on the CPU by the interpreter's own slot access and phase body
    (csrc/ucode_interp.cuh ucs_phase: class-bit dispatch, two threads per program lane, dual-product passes): the
    result is the oracle's pairing, byte for byte, and the verdict slot is zero exactly for points of G2 (for the
    others the output is the all-zero element, which no comparison with a GT value can pass)."""
    g1, g2 = oracle.generators()
    for _ in range(2):
        P = oracle.g1_mul(rng.scalar().to_bytes(32, 'big'), g1)
        Q = oracle.g2_mul(rng.scalar().to_bytes(32, 'big'), g2)
        out = ctypes.create_string_buffer(384)
        assert emu.emu_ucode_pair(P, Q, out) == 0
        assert out.raw == oracle.pair(P, Q)
    T = R.g2_twist_point(11)
    for Qo in (T, R.g2_mul_raw(R.n * R.TWIST_COFACTOR // 13, T)):
        assert R.g2_on_curve(Qo) and not R.g2_in_subgroup(Qo)
        assert emu.emu_ucode_pair(P, R.g2_to_bytes(Qo), out) == 1
        assert out.raw == bytes(384)


def test_glv_split_and_product(emu, oracle, rng):
    """curve.cuh glv_split / glv_half_mul on the CPU: the constants follow from u, every scalar splits as
    k = +-k1 +- k2 lambda (mod n) with both parts below 2^128 (the kernel walks 33 nibbles), and the two half products
    add up to the oracle's k * P, infinity and the edge scalars included."""
    u = 6518589491078791937
    lam = (36 * u ** 4 - 1) % R.n
    assert (lam * lam + lam + 1) % R.n == 0
    a1, a2 = 6 * u * u + 2 * u, 2 * u + 1
    b2 = a1 + a2
    assert (a1 - a2 * lam) % R.n == 0 and (a2 + b2 * lam) % R.n == 0
    src = open(os.path.join(os.path.dirname(__file__), '..', 'drynx_b200', 'csrc', 'curve.cuh')).read()
    def const(name):
        body = src.split('c_%s(int i) { const uint32_t v[' % name)[1].split('{')[1].split('}')[0]
        return sum(int(x.strip().rstrip('u'), 16) << (32 * i) for i, x in enumerate(body.split(',')))
    assert (const('a1'), const('a2'), const('b2')) == (a1, a2, b2)
    assert const('g1') == (b2 << 256) // R.n and const('g2') == (a2 << 256) // R.n
    g1, _ = oracle.generators()
    scalars = [0, 1, 2, R.n - 1, R.n - 2, lam, lam + 1, lam - 1, R.n // 2, 1 << 255, a1, b2, (1 << 256) - 1]
    scalars += [rng.scalar() for _ in range(300)]
    k1 = (ctypes.c_uint32 * 8)()
    k2 = (ctypes.c_uint32 * 8)()
    P = oracle.g1_mul(rng.scalar().to_bytes(32, 'big'), g1)
    out = ctypes.create_string_buffer(64)
    for n_, k in enumerate(scalars):
        kb = k.to_bytes(32, 'big')
        signs = emu.emu_glv_split(kb, k1, k2)
        v1 = sum(int(v) << (32 * i) for i, v in enumerate(k1))
        v2 = sum(int(v) << (32 * i) for i, v in enumerate(k2))
        assert v1 < 1 << 128 and v2 < 1 << 128, k
        s1 = -v1 if signs & 1 else v1
        s2 = -v2 if signs & 2 else v2
        assert (s1 + s2 * lam - k) % R.n == 0, k
        if n_ < 40:
            assert emu.emu_g1_mul_glv(kb, P, out) == 0
            assert out.raw == oracle.g1_mul((k % R.n).to_bytes(32, 'big'), P), k
    assert emu.emu_g1_mul_glv((12345).to_bytes(32, 'big'), bytes(64), out) == 0 and out.raw == bytes(64)
