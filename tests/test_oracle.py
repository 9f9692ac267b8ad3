"""Pins the C++ oracle (oracle/oracle.cpp: kyber / unlynx / lib/range algorithms restated) against the
mathematical definition in oracle/bn256_ref.py and the committed golden vectors. CPU only."""
import json
import os

import numpy as np
import pytest

import bn256_ref as R
from conftest import SplitMix

GOLD = os.path.join(os.path.dirname(__file__), 'golden')


def test_generators_and_order(oracle):
    g1, g2 = oracle.generators()
    assert g1 == R.g1_to_bytes(R.G1_GEN) and g2 == R.g2_to_bytes(R.G2_GEN)
    assert R.g1_on_curve(R.G1_GEN) and R.g2_on_curve(R.G2_GEN)
    assert oracle.g1_mul(R.n.to_bytes(32, 'big'), g1) == bytes(64)  # n reduces to 0 -> infinity
    assert oracle.g2_mul((R.n - 1).to_bytes(32, 'big'), g2) == R.g2_to_bytes(R.g2_neg(R.G2_GEN))


def test_scalar_mults_match_definition(oracle, rng):
    for k in [0, 1, 2, 3, 255, 256, R.n - 1, rng.scalar(), rng.scalar()]:
        P = R.g1_mul(rng.scalar(), R.G1_GEN)
        Q = R.g2_mul(rng.scalar(), R.G2_GEN)
        kb = k.to_bytes(32, 'big')
        assert oracle.g1_mul(kb, R.g1_to_bytes(P)) == R.g1_to_bytes(R.g1_mul(k, P))
        assert oracle.g2_mul(kb, R.g2_to_bytes(Q)) == R.g2_to_bytes(R.g2_mul(k, Q))


def test_pairing_matches_definition_and_is_bilinear(oracle, rng):
    a, b = rng.scalar(), rng.scalar()
    P, Q = R.g1_mul(a, R.G1_GEN), R.g2_mul(b, R.G2_GEN)
    e = oracle.pair(R.g1_to_bytes(P), R.g2_to_bytes(Q))
    assert e == R.f12_to_bytes(R.pair(P, Q))
    base = R.f12_from_bytes(oracle.pair(*oracle.generators()))
    assert R.f12_to_bytes(R.f12pow(base, a * b % R.n)) == e
    assert R.f12pow(base, R.n) == R.F12_ONE
    assert oracle.pair(bytes(64), R.g2_to_bytes(Q)) == R.f12_to_bytes(R.F12_ONE)  # infinity -> 1


def test_malformed_points_are_rejected(oracle):
    bad = (5).to_bytes(32, 'big') + (7).to_bytes(32, 'big')
    out = np.zeros(64, np.uint8)
    assert oracle.lib.orc_g1_mul((1).to_bytes(32, 'big'), bad, out.ctypes.data) == -1
    big = (R.p + 1).to_bytes(32, 'big') + (R.p - 2).to_bytes(32, 'big')  # x = p + 1 == 1 mod p, non-canonical
    assert oracle.lib.orc_g1_mul((1).to_bytes(32, 'big'), big, out.ctypes.data) == -1


def test_encrypt_matches_definition(oracle, rng):
    sk = rng.scalar()
    P = R.g1_mul(sk, R.G1_GEN)
    ms = [0, 1, -1, 9, -12000, 2 ** 40, -(2 ** 62)]
    rs = [rng.scalar() for _ in ms]
    ct = oracle.encrypt(R.g1_to_bytes(P), ms, b''.join(r.to_bytes(32, 'big') for r in rs))
    for i, (m, r) in enumerate(zip(ms, rs)):
        assert ct[128 * i:128 * i + 128] == R.encrypt(P, m, r)


def test_encrypt_scalar_matches_definition(oracle, rng):
    """libunlynx.EncryptScalar (lib/encoding/OR_AND.go:52,105): scalar messages incl. zero and an unreduced value."""
    sk = rng.scalar()
    Pp = R.g1_mul(sk, R.G1_GEN)
    ms = [0, 1, R.n - 1, R.n + 3, rng.scalar()]
    rs = [rng.scalar() for _ in ms]
    got = oracle.encrypt_scalar(R.g1_to_bytes(Pp), b''.join(m.to_bytes(32, 'big') for m in ms),
                                b''.join(r.to_bytes(32, 'big') for r in rs), threads=2)
    assert got == b''.join(R.encrypt_scalar(Pp, m, r) for m, r in zip(ms, rs))
    assert list(oracle.decrypt_check_zero(sk.to_bytes(32, 'big'), got)) == [0, 1, 1, 1, 1]


def test_sum_round_trip_like_reference_test(oracle, rng):
    """lib/encoding/sum_test.go:15-31: inputs {0..9, -12000} -> -11955 after encrypt/decrypt."""
    sk = rng.scalar()
    P = R.g1_to_bytes(R.g1_mul(sk, R.G1_GEN))
    inputs = [0, 1, 2, 3, 4, 5, 6, 7, 8, 9, -12000]
    ct = oracle.encrypt(P, [sum(inputs)], rng.scalars_bytes(1))
    oracle.dlog_reset()
    out, found = oracle.decrypt(sk.to_bytes(32, 'big'), ct, with_neg=True)
    assert out[0] == -11955 and found[0] == 1


def test_aggregate_is_homomorphic(oracle, rng):
    sk = rng.scalar()
    P = R.g1_to_bytes(R.g1_mul(sk, R.G1_GEN))
    vals = [[rng.randint(-50, 50) for _ in range(4)] for _ in range(3)]
    vecs = [oracle.encrypt(P, v, rng.scalars_bytes(4)) for v in vals]
    rc, agg = oracle.aggregate(vecs)
    assert rc == 0
    oracle.dlog_reset()
    out, _ = oracle.decrypt(sk.to_bytes(32, 'big'), agg)
    assert list(out) == [sum(v[j] for v in vals) for j in range(4)]
    # definition: component-wise point addition
    K = None
    for v in vecs:
        K = R.g1_add(K, R.g1_from_bytes(v[:64]))
    assert agg[:64] == R.g1_to_bytes(K)


def test_discrete_log_table_semantics(oracle, rng):
    """unlynx discreteLog: lazily grown table, negatives only while checkNeg extended it, bound 100000."""
    sk = rng.scalar()
    P = R.g1_to_bytes(R.g1_mul(sk, R.G1_GEN))
    skb = sk.to_bytes(32, 'big')
    ct = oracle.encrypt(P, [5, -3, 99999, 100000, -7], rng.scalars_bytes(5))
    oracle.dlog_reset()
    out, found = oracle.decrypt(skb, ct[:128], with_neg=False)          # grows 0..5 without negatives
    assert (out[0], found[0]) == (5, 1)
    out, found = oracle.decrypt(skb, ct[128:256], with_neg=False)       # -3 not stored: walks to the bound
    assert (out[0], found[0]) == (0, 0)
    out, found = oracle.decrypt(skb, ct[256:384], with_neg=True)        # 99999 is already in the table
    assert (out[0], found[0]) == (99999, 1)
    out, found = oracle.decrypt(skb, ct[384:512], with_neg=True)        # out of bound
    assert (out[0], found[0]) == (0, 0)
    out, found = oracle.decrypt(skb, ct[512:640], with_neg=True)        # -7 never stored (table already full)
    assert (out[0], found[0]) == (0, 0)


def test_signatures_match_definition(oracle):
    Y, A = oracle.make_sigs([12], 4)  # InitRangeProofSignatureDeterministic uses x = 12
    Yd, Ad = R.init_range_sig(12, 4)
    assert Y == R.g1_to_bytes(Yd)
    assert A == b''.join(R.g2_to_bytes(a) for a in Ad)


def _small_case(oracle, seed, ns, u, l, secret):
    rng = SplitMix(seed)
    sk = rng.scalar()
    Pp = R.g1_mul(sk, R.G1_GEN)
    P = R.g1_to_bytes(Pp)
    xs = [rng.scalar() for _ in range(ns)]
    Y, A = oracle.make_sigs(xs, u)
    r = rng.scalar()
    ct = oracle.encrypt(P, [secret], r.to_bytes(32, 'big'))
    nd = len(R.to_base(secret, u, l))
    rnd = [rng.scalar() for _ in range(nd * (3 + ns))]
    rndb = b''.join(x.to_bytes(32, 'big') for x in rnd)
    return dict(sk=sk, Pp=Pp, P=P, xs=xs, Y=Y, A=A, r=r, ct=ct, nd=nd, rnd=rnd, rndb=rndb)


@pytest.mark.parametrize('ns,u,l,secret', [(1, 2, 1, 1), (2, 4, 2, 11)])
def test_range_proof_matches_definition(oracle, ns, u, l, secret):
    c = _small_case(oracle, 99 + ns, ns, u, l, secret)
    proofs, off = oracle.rp_create(ns, u, l, c['Y'], c['A'], c['P'], [secret], c['r'].to_bytes(32, 'big'), c['ct'], [c['rndb']])
    sigs = [R.init_range_sig(x, u) for x in c['xs']]
    want = R.rp_create(sigs, u, l, secret, c['r'], c['Pp'], c['ct'], c['rnd'])
    assert proofs == want
    assert len(proofs) == oracle.proof_size(ns, c['nd'])
    assert list(oracle.rp_verify(ns, u, l, c['Y'], c['P'], proofs, off)) == [1]
    assert R.rp_verify(proofs, u, l, [s[0] for s in sigs], c['Pp']) is True


def test_range_proof_like_reference_test_and_tamper(oracle):
    """lib/range/range_proof_test.go:43-77 (u=2, l=1, 5 servers, create -> bytes -> verify == true), plus the
    negative cases the reference never tests."""
    ns, u, l, secret = 5, 2, 1, 1
    c = _small_case(oracle, 7, ns, u, l, secret)
    proofs, off = oracle.rp_create(ns, u, l, c['Y'], c['A'], c['P'], [secret], c['r'].to_bytes(32, 'big'), c['ct'], [c['rndb']])
    assert list(oracle.rp_verify(ns, u, l, c['Y'], c['P'], proofs, off)) == [1]
    for pos in (130, 170, 200, 260, 256 + 32 + 5, 256 + 32 + 128 + 9, len(proofs) - 3):
        bad = bytearray(proofs)
        bad[pos] ^= 1
        assert list(oracle.rp_verify(ns, u, l, c['Y'], c['P'], bytes(bad), off)) == [0], pos
    assert list(oracle.rp_verify(ns, u, 2, c['Y'], c['P'], proofs, off)) == [0]      # size check :512
    assert list(oracle.rp_verify(ns, 0, 0, c['Y'], c['P'], proofs, off)) == [1]      # u = l = 0 bypass :508
    # a value outside [0, u^l) needs more digits than l, so its proof fails the size check
    p2, off2 = oracle.rp_create(ns, u, l, c['Y'], c['A'], c['P'], [2], c['r'].to_bytes(32, 'big'),
                                oracle.encrypt(c['P'], [2], c['r'].to_bytes(32, 'big')), [c['rndb'] * 2])
    assert list(oracle.rp_verify(ns, u, l, c['Y'], c['P'], p2, off2)) == [0]


def test_golden_vectors(oracle):
    """tests/golden/*.json were produced by tools/make_golden.py from the Python definition."""
    with open(os.path.join(GOLD, 'bn256_golden.json')) as f:
        g = json.load(f)
    for v in g['g1_mul']:
        assert oracle.g1_mul(bytes.fromhex(v['k']), bytes.fromhex(v['p'])).hex() == v['out']
    for v in g['g2_mul']:
        assert oracle.g2_mul(bytes.fromhex(v['k']), bytes.fromhex(v['p'])).hex() == v['out']
    for v in g['pair']:
        assert oracle.pair(bytes.fromhex(v['g1']), bytes.fromhex(v['g2'])).hex() == v['out']
    for v in g['encrypt']:
        ct = oracle.encrypt(bytes.fromhex(v['P']), [v['m']], bytes.fromhex(v['r']))
        assert ct.hex() == v['ct']
    for v in g['range_proof']:
        proofs, off = oracle.rp_create(v['ns'], v['u'], v['l'], bytes.fromhex(v['Y']), bytes.fromhex(v['A']),
                                       bytes.fromhex(v['P']), [v['secret']], bytes.fromhex(v['r']),
                                       bytes.fromhex(v['ct']), [bytes.fromhex(v['rnd'])])
        assert proofs.hex() == v['proof']
        ok = oracle.rp_verify(v['ns'], v['u'], v['l'], bytes.fromhex(v['Y']), bytes.fromhex(v['P']), proofs, off)
        assert list(ok) == [1 if v['valid'] else 0]


def test_key_switch_matches_definition_and_decrypts_under_target_key(oracle, rng):
    """libunlynxkeyswitch.KeySwitchSequence (call sites data/data.go:80, services/service.go:565-616): the C++
    restatement against the big-integer definition, then the protocol's point: ciphertexts under the collective key
    of three servers, switched by each server's contribution and summed at the root, decrypt under the querier's key."""
    xs = [rng.scalar() for _ in range(3)]
    coll = None
    for x in xs:
        coll = R.g1_add(coll, R.g1_mul(x, R.G1_GEN))
    q = rng.scalar()
    Q = R.g1_mul(q, R.G1_GEN)
    ms = [0, 5, -7, 1234]
    rs = [rng.scalar() for _ in ms]
    cts = [R.encrypt(coll, m, r) for m, r in zip(ms, rs)]
    rbs = b''.join(ct[:64] for ct in cts)
    contribs = [b''.join(b'\0' * 64 + ct[64:] for ct in cts)]     # the root starts from (0, C)
    for x in xs:
        vs = [rng.scalar() for _ in ms]
        vb = b''.join(v.to_bytes(32, 'big') for v in vs)
        rc, cv, neg = oracle.key_switch(R.g1_to_bytes(Q), x.to_bytes(32, 'big'), rbs, vb, threads=2)
        assert rc == 0
        for i, v in enumerate(vs):
            want_cv, want_neg = R.key_switch(Q, R.g1_from_bytes(rbs[64 * i:64 * i + 64]), x, v)
            assert cv[128 * i:128 * i + 128] == want_cv and neg[64 * i:64 * i + 64] == want_neg
        contribs.append(cv)
    rc, switched = oracle.aggregate(contribs, threads=2)
    assert rc == 0
    oracle.dlog_reset()
    vals, found = oracle.decrypt(q.to_bytes(32, 'big'), switched)
    assert list(vals) == ms and all(found)
    # a malformed K component is reported, not switched
    bad = bytearray(rbs)
    bad[5] ^= 1
    rc, _, _ = oracle.key_switch(R.g1_to_bytes(Q), xs[0].to_bytes(32, 'big'), bytes(bad), b'\0' * 31 + b'\1' + vb[32:])
    assert rc == -2


def test_obfuscation_scalar_mul_matches_definition(oracle, rng):
    """CipherText.MulCipherTextbyScalar (protocols/obfuscation_protocol.go:240-243): a zero stays an encryption of
    zero, a non-zero value stays non-zero (the OR / AND / min / max decodings rely on exactly that)."""
    sk = rng.scalar()
    P = R.g1_mul(sk, R.G1_GEN)
    ms = [0, 1, 7, 0]
    cts = b''.join(R.encrypt(P, m, rng.scalar()) for m in ms)
    ss = [rng.scalar() for _ in ms]
    rc, out = oracle.ct_scalar_mul(b''.join(s.to_bytes(32, 'big') for s in ss), cts, threads=2)
    assert rc == 0
    for i, s in enumerate(ss):
        assert out[128 * i:128 * i + 128] == R.ct_scalar_mul(cts[128 * i:128 * i + 128], s)
    assert list(oracle.decrypt_check_zero(sk.to_bytes(32, 'big'), out)) == [0, 1, 1, 0]
    bad = bytearray(cts)
    bad[70] ^= 1
    assert oracle.ct_scalar_mul(b''.join(s.to_bytes(32, 'big') for s in ss), bytes(bad))[0] == -2


def test_schnorr_matches_definition(oracle, rng):
    """kyber sign/schnorr over G1 (VerifyProofSignature, lib/proof/structs_proofs.go:498-505): signatures made by
    the C++ restatement equal the definition's and verify; a flipped message / R / s / key, s >= n and R off the curve
    are rejected by both."""
    x, k = rng.scalar(), rng.scalar()
    X = R.g1_mul(x, R.G1_GEN)
    Xb = R.g1_to_bytes(X)
    msg = b'range proofs of survey 7' * 40
    sig = oracle.schnorr_sign(x.to_bytes(32, 'big'), k.to_bytes(32, 'big'), Xb, msg)
    assert sig == R.schnorr_sign(x, k, msg) and R.schnorr_verify(X, msg, sig)
    other = R.g1_to_bytes(R.g1_mul(rng.scalar(), R.G1_GEN))
    flip = lambda b, i: bytes(b[:i]) + bytes([b[i] ^ 1]) + bytes(b[i + 1:])
    big_s = sig[:64] + (int.from_bytes(sig[64:], 'big') + R.n).to_bytes(32, 'big') if int.from_bytes(sig[64:], 'big') + R.n < 2 ** 256 else sig
    cases = [(Xb, sig, msg, 1), (Xb, sig, msg + b'!', 0), (Xb, flip(sig, 95), msg, 0), (Xb, flip(sig, 10), msg, 0),
             (other, sig, msg, 0), (Xb, big_s, msg, 1 if big_s == sig else 0), (Xb, bytes(64) + sig[64:], msg, 0)]
    got = oracle.schnorr_verify([c[0] for c in cases], [c[1] for c in cases], [c[2] for c in cases], threads=2)
    assert list(got) == [c[3] for c in cases]
    for pub, sg, m, want in cases:
        assert R.schnorr_verify(R.g1_from_bytes(pub), m, sg) == bool(want)


def test_shuffle_rerandomize_matches_definition(oracle, rng):
    """libunlynxshuffle.ShuffleSequence (data/data.go:68): lines permuted as a whole, every ciphertext re-randomised;
    the plaintexts survive in permuted order."""
    sk = rng.scalar()
    P = R.g1_mul(sk, R.G1_GEN)
    vals = [[1, 2, 3], [4, 5, 6], [7, 8, 9], [0, -1, 99]]
    lines = [b''.join(R.encrypt(P, m, rng.scalar()) for m in row) for row in vals]
    betas = [[rng.scalar() for _ in row] for row in vals]
    pi = [2, 0, 3, 1]
    rc, out = oracle.shuffle_rerandomize(R.g1_to_bytes(P), lines, [b''.join(b.to_bytes(32, 'big') for b in row) for row in betas], pi, threads=2)
    assert rc == 0
    for i in range(4):
        for j in range(3):
            assert out[i][128 * j:128 * j + 128] == R.rerandomize(lines[pi[i]][128 * j:128 * j + 128], P, betas[pi[i]][j])
        assert out[i] != lines[pi[i]]
    oracle.dlog_reset()
    for i in range(4):
        got, found = oracle.decrypt(sk.to_bytes(32, 'big'), out[i])
        assert list(got) == vals[pi[i]] and all(found)


# ---------------------------------------------------------------------------------------------- G2 membership
def _twist_points_outside_g2(seed):
    T = R.g2_twist_point(seed)
    pts = [T, R.g2_mul_raw(R.n, T), R.g2_mul_raw(R.n * R.TWIST_COFACTOR // 13, T), R.g2_mul_raw(13, T)]
    for Q in pts:
        assert Q is not None and R.g2_on_curve(Q) and not R.g2_in_subgroup(Q)
    return pts


def test_g2_membership_relation_is_sound_on_this_curve():
    """The GPU decides membership in G2 by [6u+2]Q + pi(Q) - pi^2(Q) + pi^3(Q) == O instead of kyber's [n]Q == O.
    Any Q on the twist is killed by the characteristic polynomial chi(X) = X^2 - tX + p of pi; a Q passing the test
    is also killed by h(X) = X^3 - X^2 + X + 6u+2, hence by the integer Res(h, chi). The order of Q then divides
    gcd(Res, #E'(Fp2)), and #E'(Fp2) = n (2p - n): if that gcd is n the two tests accept the same points."""
    from fractions import Fraction
    from math import gcd
    u, p, n = R.u, R.p, R.n
    t = 6 * u * u + 1
    assert n == p + 1 - t and (p ** 3 - p ** 2 + p + 6 * u + 2) % n == 0
    h, chi = [1, -1, 1, 6 * u + 2], [1, -t, p]
    M = [[Fraction(x) for x in row] for row in (h + [0], [0] + h, chi + [0, 0], [0] + chi + [0], [0, 0] + chi)]
    det = Fraction(1)
    for i in range(5):
        piv = next(r for r in range(i, 5) if M[r][i] != 0)
        if piv != i:
            M[i], M[piv] = M[piv], M[i]
            det = -det
        det *= M[i][i]
        for r in range(i + 1, 5):
            f = M[r][i] / M[i][i]
            M[r] = [x - f * y for x, y in zip(M[r], M[i])]
    res = abs(int(det))
    assert det.denominator == 1 and gcd(res, n * (2 * p - n)) == n
    # and the twist's group order really is n (2p - n): a random twist point is killed by it
    assert R.g2_mul_raw(n * (2 * p - n), R.g2_twist_point(3)) is None


def test_twist_points_outside_g2_are_malformed(oracle):
    """kyber's twistPoint.IsOnCurve multiplies by Order (recalled), so UnmarshalBinary rejects these: the C++ oracle and
    the Python definition must both refuse them wherever a G2 point is parsed."""
    one = (1).to_bytes(32, 'big')
    out128, out384 = np.zeros(128, np.uint8), np.zeros(384, np.uint8)
    g1 = R.g1_to_bytes(R.G1_GEN)
    for Q in _twist_points_outside_g2(11):
        b = R.g2_to_bytes(Q)
        with pytest.raises(ValueError):
            R.g2_from_bytes(b)
        assert oracle.lib.orc_g2_mul(one, b, out128.ctypes.data) == -1
        assert oracle.lib.orc_pair(g1, b, out384.ctypes.data) == -1
    good = R.g2_to_bytes(R.g2_mul(5, R.G2_GEN))
    assert oracle.lib.orc_g2_mul(one, good, out128.ctypes.data) == 0


def test_range_proof_with_v_outside_g2_is_rejected(oracle):
    """lib/range/range_proof.go:225-236: FromBytes aborts on a V_ij that kyber cannot unmarshal; here the proof is
    reported invalid. Every other field of the proof is left as the prover made it."""
    ns, u, l, secret = 2, 4, 2, 9
    c = _small_case(oracle, 21, ns, u, l, secret)
    proofs, off = oracle.rp_create(ns, u, l, c['Y'], c['A'], c['P'], [secret], c['r'].to_bytes(32, 'big'), c['ct'], [c['rndb']])
    assert list(oracle.rp_verify(ns, u, l, c['Y'], c['P'], proofs, off)) == [1]
    v00 = 256 + 32 * l
    for Q in _twist_points_outside_g2(5):
        bad = bytearray(proofs)
        bad[v00:v00 + 128] = R.g2_to_bytes(Q)
        assert list(oracle.rp_verify(ns, u, l, c['Y'], c['P'], bytes(bad), off)) == [0]
        assert R.rp_verify(bytes(bad), u, l, [R.g1_from_bytes(c['Y'][64 * i:64 * i + 64]) for i in range(ns)], c['Pp']) is False
