"""The proof.Rep list proofs of SURVEY 8(f)-2: obfuscation (lib/obfuscation/obfuscation_proof.go:37-91) and key
switching (libunlynxkeyswitch, verified from lib/proof/structs_proofs.go:469-482). The CPU tests pin the oracle's three
steps to the Python definition; the GPU tests run whole list proofs through the CUDA entry points and the oracle and
compare every byte, then tamper with each part of a proof in turn. The reference's own tests for these proofs
(lib/obfuscation/obfuscation_proof_test.go) check create -> verify == true and a wrong ciphertext / wrong secret ->
false; both are here."""
import numpy as np
import pytest

import bn256_ref as R
import rep_proofs as RP
from conftest import SplitMix


def _pt(k):
    return R.g1_to_bytes(R.g1_mul(k, R.G1_GEN))


def _obf_inputs(rng, n, engine_mul):
    """n ciphertexts (arbitrary points), their obfuscations by s_i, blinding values; the edge cases the protocol can
    meet: s = 0 (co = identity), s = 1, a ciphertext holding the identity, an unreduced secret."""
    cts = bytearray(b''.join(_pt(rng.scalar()) + _pt(rng.scalar()) for _ in range(n)))
    if n > 3:
        cts[128 * 3:128 * 3 + 64] = bytes(64)
    ss = [rng.scalar() for _ in range(n)]
    if n > 2:
        ss[0], ss[1], ss[2] = 0, 1, R.n - 1
    sb = b''.join(s.to_bytes(32, 'big') for s in ss)
    vs = rng.scalars_bytes(n)
    cos = engine_mul(sb, bytes(cts))
    return bytes(cts), cos, sb, vs


def test_oracle_rep_steps_match_the_definition(oracle, rng):
    """rep_commit / rep_respond / rep_verify of the oracle against bn256_ref's group law and Python integers."""
    g = SplitMix(4242)
    n, nt = 6, 2
    sc = [g.scalar() for _ in range(n * nt)]
    ks = [g.scalar() for _ in range(n * nt)]
    pts = b''.join(_pt(k) for k in ks)
    rc, V = oracle.rep_commit(nt, b''.join(s.to_bytes(32, 'big') for s in sc), pts)
    assert rc == 0
    for i in range(n):
        want = sum(sc[nt * i + j] * ks[nt * i + j] for j in range(nt)) % R.n
        assert V[64 * i:64 * i + 64] == _pt(want)
    v, c, x = [g.scalar() for _ in range(n)], [g.scalar() for _ in range(3)], [g.scalar() for _ in range(n)]
    r = oracle.rep_respond(2, b''.join(a.to_bytes(32, 'big') for a in v), b''.join(a.to_bytes(32, 'big') for a in c),
                           b''.join(a.to_bytes(32, 'big') for a in x))
    assert [int.from_bytes(r[32 * i:32 * i + 32], 'big') for i in range(n)] == [(v[i] - c[i // 2] * x[i]) % R.n for i in range(n)]
    # a relation that holds, and the same one with each input disturbed
    xs, vs, bs = [g.scalar(), g.scalar()], [g.scalar(), g.scalar()], [g.scalar(), g.scalar()]
    ch = g.scalar()
    bases = b''.join(_pt(b) for b in bs)
    P = _pt(sum(a * b for a, b in zip(xs, bs)) % R.n)
    Vv = _pt(sum(a * b for a, b in zip(vs, bs)) % R.n)
    rs = b''.join(((vs[j] - ch * xs[j]) % R.n).to_bytes(32, 'big') for j in range(2))
    chb = ch.to_bytes(32, 'big')
    assert list(oracle.rep_verify(2, rs, chb, bases, P, Vv)) == [1]
    assert list(oracle.rep_verify(2, rs, (ch + 1).to_bytes(32, 'big'), bases, P, Vv)) == [0]
    assert list(oracle.rep_verify(2, rs[32:] + rs[:32], chb, bases, P, Vv)) == [0]
    assert list(oracle.rep_verify(2, rs, chb, bases, Vv, P)) == [0]
    big = (int.from_bytes(rs[:32], 'big') + R.n)
    if big < 2 ** 256:  # the same residue, unreduced: kyber's UnmarshalBinary refuses it
        assert list(oracle.rep_verify(2, big.to_bytes(32, 'big') + rs[32:], chb, bases, P, Vv)) == [0]
    off = bytearray(P)
    off[63] ^= 1
    assert list(oracle.rep_verify(2, rs, chb, bases, bytes(off), Vv)) == [0]


def test_oracle_obfuscation_and_keyswitch_proofs_round_trip(oracle, rng):
    eng = RP.OracleEngine(oracle)
    cts, cos, sb, vs = _obf_inputs(rng, 9, lambda s, c: oracle.ct_scalar_mul(s, c, threads=8)[1])
    proofs = RP.obfuscation_create(eng, cts, cos, sb, vs)
    assert all(len(p) == 160 for p in proofs)
    assert RP.obfuscation_verify(eng, cts, cos, proofs).all()
    wrong = bytearray(cos)
    wrong[128 * 4:128 * 5] = cos[128 * 5:128 * 6]          # somebody else's obfuscated ciphertext
    bits = RP.obfuscation_verify(eng, cts, bytes(wrong), proofs)
    assert list(bits) == [i != 4 for i in range(9)]
    assert RP.list_verdict(bits, 0.4) and not RP.list_verdict(bits, 0.6) and RP.list_verdict(bits, 0.0)
    # key switching: n ciphertexts' K components, one server key k, the querier's key Q
    n = 7
    B = R.g1_to_bytes(R.G1_GEN)
    k, q = rng.scalar(), rng.scalar()
    Q = _pt(q)
    rbs = [rng.scalar() for _ in range(n)]
    rbnegs = b''.join(_pt(R.n - r) for r in rbs)
    vis = [rng.scalar() for _ in range(n)]
    publics, proofs = RP.keyswitch_create(eng, B, Q, k.to_bytes(32, 'big'), b''.join(v.to_bytes(32, 'big') for v in vis),
                                          rbnegs, rng.scalars_bytes(2 * n))
    for i in range(n):                                       # the publics are what KeySwitchSequence publishes
        assert publics[i] == _pt(vis[i]) + _pt(k) + _pt((vis[i] * q - k * rbs[i]) % R.n)
    assert RP.keyswitch_verify(eng, B, Q, publics, rbnegs, proofs).all()
    other = _pt(rng.scalar())
    assert not RP.keyswitch_verify(eng, B, other, publics, rbnegs, proofs).any()
    for part in range(5):                                    # V1, V2, V3, r_vi, r_k of proof 2
        t = [bytearray(p) for p in proofs]
        t[2][(64 * part if part < 3 else 192 + 32 * (part - 3)) + 31] ^= 1
        bits = RP.keyswitch_verify(eng, B, Q, publics, rbnegs, [bytes(p) for p in t])
        assert list(bits) == [i != 2 for i in range(n)], part


@pytest.mark.gpu
def test_gpu_rep_steps_match_oracle(backend, oracle, rng):
    """Bytes of the three entry points against the oracle: 1, 2 and 3 terms, with the identity, a zero scalar, an
    unreduced scalar and a repeated point among the inputs; a malformed point is an error of rep_commit and a zero
    accept byte of rep_verify."""
    for nt in (1, 2, 3):
        n = 130
        sc = bytearray(rng.scalars_bytes(n * nt))
        sc[0:32] = bytes(32)
        sc[32:64] = (R.n + 5).to_bytes(32, 'big')
        pts = bytearray(b''.join(_pt(rng.scalar()) for _ in range(n * nt)))
        pts[64 * 2:64 * 3] = bytes(64)
        pts[64 * 4:64 * 5] = pts[64 * 3:64 * 4]
        got = backend.rep_commit(nt, bytes(sc), bytes(pts)).tobytes()
        rc, want = oracle.rep_commit(nt, bytes(sc), bytes(pts), threads=8)
        assert rc == 0 and got == want
        per = nt
        chal = rng.scalars_bytes((n * nt + per - 1) // per)
        x = rng.scalars_bytes(n * nt)
        r = backend.rep_respond(per, bytes(sc), chal, x).tobytes()
        assert r == oracle.rep_respond(per, bytes(sc), chal, x)
        # relations P = sum x_j B_j with the commitment above: they hold; then disturb a third of them
        rcP, P = oracle.rep_commit(nt, x, bytes(pts), threads=8)
        assert rcP == 0
        resp, V, Pm = bytearray(r), bytearray(got), bytearray(P)
        for i in range(0, n, 3):
            kind = (i // 3) % 5
            if kind == 0:
                resp[32 * nt * i + 31] ^= 1
            elif kind == 1:
                V[64 * i + 40] ^= 1                       # leaves the curve
            elif kind == 2:
                Pm[64 * i:64 * i + 64] = Pm[64 * (i + 1):64 * (i + 2)]
            elif kind == 3:
                resp[32 * nt * i:32 * nt * i + 32] = b'\xff' * 32   # >= n
            else:
                V[64 * i:64 * i + 64] = bytes(64)         # identity
        chal_rel = b''.join(chal[32 * ((i * nt) // per):32 * ((i * nt) // per) + 32] for i in range(n))
        ok = backend.rep_verify(nt, bytes(resp), chal_rel, bytes(pts), bytes(Pm), bytes(V))
        want_ok = oracle.rep_verify(nt, bytes(resp), chal_rel, bytes(pts), bytes(Pm), bytes(V), threads=8)
        assert list(ok) == list(want_ok)
        assert all(ok[i] == 1 for i in range(n) if i % 3) and sum(ok) >= n - (n + 2) // 3
        bad = bytearray(pts)
        bad[64 * 7 + 1] ^= 1
        with pytest.raises(RuntimeError):
            backend.rep_commit(nt, bytes(sc), bytes(bad))
        assert backend.rep_verify(nt, bytes(r), chal_rel, bytes(bad), P, got)[7 // nt] == 0


@pytest.mark.gpu
def test_gpu_line_compression_of_the_shuffle_proof(backend, oracle, rng):
    """The list compression that opens unlynx's shuffle proof (sum_j e_j * ct_ij per line, K and C apart) is
    rep_commit with one term per ciphertext of the line: 12 lines of 40 ciphertexts, bytes against the oracle, and
    the plaintext of the compressed ciphertext is the same combination of the plaintexts."""
    lines, nq = 12, 40
    sk = rng.scalar()
    P = _pt(sk)
    backend.set_pubkey(P)
    g = np.random.default_rng(5)
    ms = [[int(v) for v in g.integers(0, 50, size=nq)] for _ in range(lines)]
    cts = [backend.encrypt_int_vector(m, rng.scalars_bytes(nq)).tobytes() for m in ms]
    e = [int(v) for v in g.integers(0, 40, size=nq)]                      # small, so that the result decrypts
    eb = b''.join(v.to_bytes(32, 'big') for v in e)
    Ks = b''.join(ct[128 * j:128 * j + 64] for ct in cts for j in range(nq))
    Cs = b''.join(ct[128 * j + 64:128 * j + 128] for ct in cts for j in range(nq))
    K = backend.rep_commit(nq, eb * lines, Ks).tobytes()
    C = backend.rep_commit(nq, eb * lines, Cs).tobytes()
    assert (0, K) == oracle.rep_commit(nq, eb * lines, Ks, threads=8) and (0, C) == oracle.rep_commit(nq, eb * lines, Cs, threads=8)
    comp = b''.join(K[64 * i:64 * i + 64] + C[64 * i:64 * i + 64] for i in range(lines))
    out = backend.decrypt_int_vector(sk.to_bytes(32, 'big'), comp)
    assert list(out) == [sum(a * b for a, b in zip(e, m)) for m in ms]
    with pytest.raises(RuntimeError):
        backend.rep_commit(5000, bytes(32 * 5000), bytes(64 * 5000))


@pytest.mark.gpu
def test_gpu_obfuscation_list_proof(backend, oracle, rng):
    """ObfuscationListProofCreation / Verification on 300 ciphertexts: proofs byte-identical to the oracle's, accepted by
    both, and each kind of tampering rejected by both alike."""
    gpu, cpu = RP.BackendEngine(backend), RP.OracleEngine(oracle)
    n = 300
    cts, cos, sb, vs = _obf_inputs(rng, n, lambda s, c: backend.ct_scalar_mul(s, c).tobytes())
    assert cos == oracle.ct_scalar_mul(sb, cts, threads=8)[1]
    proofs = RP.obfuscation_create(gpu, cts, cos, sb, vs)
    assert proofs == RP.obfuscation_create(cpu, cts, cos, sb, vs)
    assert RP.obfuscation_verify(gpu, cts, cos, proofs).all()
    t = [bytearray(p) for p in proofs]
    for i in range(0, n, 4):
        t[i][(31, 95, 159, 130)[(i // 4) % 4]] ^= 1        # V1, V2, r (low byte), r (high byte)
    t = [bytes(p) for p in t]
    t[6] = t[6][:100]                                        # a truncated proof
    bits = RP.obfuscation_verify(gpu, cts, cos, t)
    assert list(bits) == list(RP.obfuscation_verify(cpu, cts, cos, t))
    assert list(bits) == [i % 4 != 0 and i != 6 for i in range(n)]
    # a wrong secret: the prover claims s + 1
    s2 = b''.join(((int.from_bytes(sb[32 * i:32 * i + 32], 'big') + 1) % R.n).to_bytes(32, 'big') for i in range(n))
    assert not RP.obfuscation_verify(gpu, cts, cos, RP.obfuscation_create(gpu, cts, cos, s2, vs))[4:].any()


@pytest.mark.gpu
def test_gpu_keyswitch_list_proof(backend, oracle, rng):
    """KeySwitchListProofCreation / Verification over the output of drx_keyswitch: the publics of the proof are the
    bytes key switching produced (K' = viB, C' = ks2), proofs are byte-identical to the oracle's."""
    gpu, cpu = RP.BackendEngine(backend), RP.OracleEngine(oracle)
    n = 200
    B = R.g1_to_bytes(R.G1_GEN)
    k, q = rng.scalar(), rng.scalar()
    Q = _pt(q)
    backend.set_pubkey(_pt(rng.scalar()))
    backend.set_target_pubkey(Q)
    rbs = b''.join(_pt(rng.scalar()) for _ in range(n))
    vis = rng.scalars_bytes(n)
    cv, neg = backend.key_switch(k.to_bytes(32, 'big'), rbs, vis, want_rb_neg=True)
    cv, neg = cv.tobytes(), neg.tobytes()
    blind = rng.scalars_bytes(2 * n)
    publics, proofs = RP.keyswitch_create(gpu, B, Q, k.to_bytes(32, 'big'), vis, neg, blind)
    publics2, proofs2 = RP.keyswitch_create(cpu, B, Q, k.to_bytes(32, 'big'), vis, neg, blind)
    assert publics == publics2 and proofs == proofs2
    for i in range(n):
        assert publics[i][:64] == cv[128 * i:128 * i + 64] and publics[i][128:] == cv[128 * i + 64:128 * i + 128]
        assert publics[i][64:128] == _pt(k)
    assert RP.keyswitch_verify(gpu, B, Q, publics, neg, proofs).all()
    t = [bytearray(p) for p in proofs]
    for i in range(0, n, 5):
        part = (i // 5) % 5
        t[i][(64 * part if part < 3 else 192 + 32 * (part - 3)) + 17] ^= 1
    t = [bytes(p) for p in t]
    bits = RP.keyswitch_verify(gpu, B, Q, publics, neg, t)
    assert list(bits) == list(RP.keyswitch_verify(cpu, B, Q, publics, neg, t))
    assert list(bits) == [i % 5 != 0 for i in range(n)]


@pytest.mark.gpu
@pytest.mark.parametrize('nt,n', [(32, 3), (33, 3), (64, 2), (65, 5), (300, 2), (4096, 1)])
def test_gpu_wide_relations_match_oracle(backend, oracle, rng, nt, n):
    """drx_rep_commit sums relations of 33 terms and more in two levels (chunks, then the partial sums): the bytes are
    those of the oracle's single loop on both sides of the threshold, at the chunk boundaries and at the widest
    relation the ABI takes; zero scalars, the identity among the points and a whole chunk of zero terms included; a
    malformed point anywhere is still an error of the call."""
    sc = bytearray(rng.scalars_bytes(n * nt))
    base = [_pt(rng.scalar()) for _ in range(min(nt, 24))]
    pts = bytearray(b''.join(base[(i * 7 + i // 5) % len(base)] for i in range(n * nt)))
    sc[0:32] = bytes(32)
    pts[64:128] = bytes(64)
    if nt >= 65:
        sc[32 * 16:32 * 32] = bytes(32 * 16)                       # the second 16 terms of relation 0: all zero
    got = backend.rep_commit(nt, bytes(sc), bytes(pts)).tobytes()
    rc, want = oracle.rep_commit(nt, bytes(sc), bytes(pts), threads=8)
    assert rc == 0 and got == want
    bad = bytearray(pts)
    bad[64 * (n * nt - 1) + 3] ^= 1                                 # the last point of the last relation
    with pytest.raises(RuntimeError):
        backend.rep_commit(nt, bytes(sc), bytes(bad))
    with pytest.raises(RuntimeError):
        backend.rep_commit(4097, bytes(32 * 4097), bytes(64 * 4097))
