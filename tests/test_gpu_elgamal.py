"""-m gpu: the CUDA ElGamal path through the C ABI against the oracle (bit-exact), following the reference's
round-trip tests (lib/encoding/sum_test.go, mean_test.go, frequency_count_test.go) and adding the byte-level and
edge cases the reference never pins."""
import json
import os

import numpy as np
import pytest

import bn256_ref as R
from conftest import SplitMix

pytestmark = pytest.mark.gpu
GOLD = os.path.join(os.path.dirname(__file__), 'golden')


def _limbs(x):
    return [(x >> (32 * k)) & 0xffffffff for k in range(8)]


def _keys(rng):
    sk = rng.scalar()
    return sk, R.g1_to_bytes(R.g1_mul(sk, R.G1_GEN))


def test_field_ops_bit_exact(backend, rng):
    N = 2048
    for which, mod in ((0, R.p), (1, R.n)):
        xs = [rng.scalar() % mod for _ in range(N)]
        ys = [rng.scalar() % mod for _ in range(N)]
        xs[:4] = [mod - 1, 0, 1, mod - 1]
        ys[:4] = [mod - 1, mod - 1, 1, 2 ** 255]
        a = np.array([_limbs(x) for x in xs], np.uint32).ravel()
        b = np.array([_limbs(y) for y in ys], np.uint32).ravel()
        un = lambda r: [sum(int(v) << (32 * k) for k, v in enumerate(row)) for row in r.reshape(N, 8)]
        rinv = pow(1 << 256, -1, mod)
        assert un(backend.mont_mul(which, a, b)) == [x * y * rinv % mod for x, y in zip(xs, ys)]
        assert un(backend.mont_mul(2 + 2 * which, a, b)) == [(x + y) % mod for x, y in zip(xs, ys)]
        assert un(backend.mont_mul(3 + 2 * which, a, b)) == [(x - y) % mod for x, y in zip(xs, ys)]
    # the dual-product pass (fp.cuh mont_mul2) with multipliers up to 2^256 - 1: the second conditional subtraction
    # is a rarely taken branch, (p - 1)(2^256 - 1) twice over takes it
    xs = [R.p - 1, R.p - 1, 0, 1, R.p - 1] + [rng.scalar() % R.p for _ in range(N - 5)]
    ys = [2 ** 256 - 1, R.p, 2 ** 256 - 1, 2 ** 256 - 1, 2 ** 255] + [(rng.scalar() * 3) % 2 ** 256 for _ in range(N - 5)]
    a = np.array([_limbs(x) for x in xs], np.uint32).ravel()
    b = np.array([_limbs(y) for y in ys], np.uint32).ravel()
    rinv = pow(1 << 256, -1, R.p)
    assert un(backend.mont_mul(10, a, b)) == [2 * x * y * rinv % R.p for x, y in zip(xs, ys)]


def test_generators_match_oracle(backend, oracle):
    g1, g2 = oracle.generators()
    one = (1).to_bytes(32, 'big')
    assert backend.g1_mul(one, g1).tobytes() == g1 and backend.g2_mul(one, g2).tobytes() == g2


def test_group_ops_and_pairing_match_oracle(backend, oracle, rng):
    g1, g2 = oracle.generators()
    ks = [0, 1, 2, R.n - 1] + [rng.scalar() for _ in range(12)]
    pts1 = [oracle.g1_mul(rng.scalar().to_bytes(32, 'big'), g1) for _ in ks]
    pts2 = [oracle.g2_mul(rng.scalar().to_bytes(32, 'big'), g2) for _ in ks]
    kb = b''.join(k.to_bytes(32, 'big') for k in ks)
    o1 = backend.g1_mul(kb, b''.join(pts1)).tobytes()
    o2 = backend.g2_mul(kb, b''.join(pts2)).tobytes()
    e = backend.pair(b''.join(pts1), b''.join(pts2)).tobytes()
    for i, k in enumerate(ks):
        assert o1[64 * i:64 * i + 64] == oracle.g1_mul(k.to_bytes(32, 'big'), pts1[i])
        assert o2[128 * i:128 * i + 128] == oracle.g2_mul(k.to_bytes(32, 'big'), pts2[i])
        assert e[384 * i:384 * i + 384] == oracle.pair(pts1[i], pts2[i])
    # infinity in either slot -> 1
    one = R.f12_to_bytes(R.F12_ONE)
    assert backend.pair(bytes(64) + pts1[4], pts2[4] + bytes(128)).tobytes() == one + one


def test_malformed_points_rejected(backend):
    from drynx_b200 import DrynxGpuError
    bad = (5).to_bytes(32, 'big') + (7).to_bytes(32, 'big')
    with pytest.raises(DrynxGpuError):
        backend.g1_mul((1).to_bytes(32, 'big'), bad)
    noncanon = (R.p + 1).to_bytes(32, 'big') + (R.p - 2).to_bytes(32, 'big')
    with pytest.raises(DrynxGpuError):
        backend.g1_mul((1).to_bytes(32, 'big'), noncanon)
    with pytest.raises(DrynxGpuError):
        backend.set_pubkey(bad)
    with pytest.raises(DrynxGpuError):
        backend.aggregate([bad + bad])


def test_encrypt_bytes_match_oracle(backend, oracle, rng):
    sk, P = _keys(rng)
    backend.set_pubkey(P)
    ms = [0, 1, -1, 9, -12000, 99999, 100000, 2 ** 40 + 3, -(2 ** 62), 2 ** 63 - 1, -(2 ** 63)]
    ms += [rng.randint(-10 ** 6, 10 ** 6) for _ in range(500)]
    rs = [rng.scalar() for _ in ms]
    rs[3] = 0
    rs[4] = R.n - 1
    rb = b''.join(r.to_bytes(32, 'big') for r in rs)
    got = backend.encrypt_int_vector(ms, rb).tobytes()
    assert got == oracle.encrypt(P, ms, rb, threads=8)
    assert backend.encrypt_int_vector([], b'').size == 0
    # a scalar given unreduced (r + n < 2^256) encrypts like r: kyber SetBytes reduces mod n
    r_small = 12345
    a = backend.encrypt_int_vector([7], (r_small + R.n).to_bytes(32, 'big')).tobytes()
    assert a == oracle.encrypt(P, [7], r_small.to_bytes(32, 'big'))


def test_encrypt_lane_split_and_one_thread_kernels_agree(backend, oracle, rng):
    """Up to 16384 encryptions per call run eight lanes each (k_encrypt_lanes), more run one thread each
    (k_encrypt). The same 17000 inputs through one call and through two calls of 8500 give the same bytes, and
    a sample of them matches the oracle."""
    sk, P = _keys(rng)
    backend.set_pubkey(P)
    n = 17000
    ms = np.array([rng.randint(-10 ** 9, 10 ** 9) for _ in range(n)], np.int64)
    ms[:6] = [0, -1, 2 ** 63 - 1, -(2 ** 63), 255, 256]
    rb = rng.scalars_bytes(n)
    one = backend.encrypt_int_vector(ms, rb).tobytes()
    h = n // 2
    two = backend.encrypt_int_vector(ms[:h], rb[:32 * h]).tobytes() + backend.encrypt_int_vector(ms[h:], rb[32 * h:]).tobytes()
    assert one == two
    assert one[:128 * 64] == oracle.encrypt(P, [int(x) for x in ms[:64]], rb[:32 * 64], threads=8)


def test_encrypt_scalar_matches_oracle_and_definition(backend, oracle, rng):
    """libunlynx.EncryptScalar (OR_AND.go:52,105): full-width scalar messages, incl. 0, n-1 and an unreduced value."""
    sk, P = _keys(rng)
    backend.set_pubkey(P)
    ms = [0, 1, R.n - 1, R.n + 5, 2 ** 255] + [rng.scalar() for _ in range(40)]
    rs = [rng.scalar() for _ in ms]
    mb = b''.join(m.to_bytes(32, 'big') for m in ms)
    rb = b''.join(r.to_bytes(32, 'big') for r in rs)
    got = backend.encrypt_scalar_vector(mb, rb).tobytes()
    assert got == oracle.encrypt_scalar(P, mb, rb, threads=8)
    Pp = R.g1_from_bytes(P)
    for i in (0, 2, 3, 7):
        assert got[128 * i:128 * i + 128] == R.encrypt_scalar(Pp, ms[i], rs[i])
    # DecryptCheckZero semantics of the OR / AND encodings: zero scalar <-> identity after decryption
    z = backend.decrypt_check_zero(sk.to_bytes(32, 'big'), got[:128 * 6])
    assert list(z) == [0, 1, 1, 1, 1, 1]


def test_key_switch_matches_oracle_and_decrypts_under_target_key(backend, oracle, rng):
    """SURVEY 8(f)-3, libunlynxkeyswitch.KeySwitchSequence: bytes identical to the oracle for 300 ciphertexts (with
    the identity, a repeated point and an unreduced v among them), -rB as well; then the whole protocol on the device
    entry points: three servers' contributions summed onto (0, C) by drx_aggregate decrypt under the querier's key."""
    xs = [rng.scalar() for _ in range(3)]
    coll = None
    for x in xs:
        coll = R.g1_add(coll, R.g1_mul(x, R.G1_GEN))
    P = R.g1_to_bytes(coll)
    q = rng.scalar()
    Q = R.g1_to_bytes(R.g1_mul(q, R.G1_GEN))
    backend.set_pubkey(P)
    backend.set_target_pubkey(Q)
    n = 300
    g = np.random.default_rng(33)
    ms = [int(v) for v in g.integers(-5000, 5000, size=n)]
    rs = rng.scalars_bytes(n)
    cts = backend.encrypt_int_vector(ms, rs).tobytes()
    rbs = bytearray(b''.join(cts[128 * i:128 * i + 64] for i in range(n)))
    rbs[64:128] = bytes(64)                 # K = identity
    rbs[128:192] = rbs[0:64]                # a repeated point
    rbs = bytes(rbs)
    contribs = [b''.join(bytes(64) + cts[128 * i + 64:128 * i + 128] for i in range(n))]
    for k, x in enumerate(xs):
        vs = bytearray(rng.scalars_bytes(n))
        vs[0:32] = (R.n + 9).to_bytes(32, 'big')   # reduced mod n like Scalar.SetBytes
        vs[32:64] = bytes(32)                       # v = 0
        cv, neg = backend.key_switch(x.to_bytes(32, 'big'), rbs, bytes(vs), want_rb_neg=True)
        rc, want_cv, want_neg = oracle.key_switch(Q, x.to_bytes(32, 'big'), rbs, bytes(vs), threads=8)
        assert rc == 0
        assert cv.tobytes() == want_cv and neg.tobytes() == want_neg
        if k == 0:
            one, _ = R.key_switch(R.g1_from_bytes(Q), R.g1_from_bytes(rbs[192:256]), x, int.from_bytes(vs[96:128], 'big'))
            assert cv.tobytes()[384:512] == one
        # the protocol uses the true K components
        true_rbs = b''.join(cts[128 * i:128 * i + 64] for i in range(n))
        contribs.append(backend.key_switch(x.to_bytes(32, 'big'), true_rbs, bytes(vs)).tobytes())
    switched = backend.aggregate(contribs)
    backend.create_decryption_table(10000)
    assert list(backend.decrypt_int_vector(q.to_bytes(32, 'big'), switched)) == ms
    # malformed K component -> -2, like a failed UnmarshalBinary
    bad = bytearray(rbs)
    bad[200] ^= 1
    with pytest.raises(RuntimeError):
        backend.key_switch(xs[0].to_bytes(32, 'big'), bytes(bad), rng.scalars_bytes(n))


def test_obfuscation_scalar_mul_matches_oracle(backend, oracle, rng):
    """CipherText.MulCipherTextbyScalar (protocols/obfuscation_protocol.go:240-243) on 200 ciphertexts, with the
    scalars 0, 1, n-1 and an unreduced one and an identity component among them; DecryptCheckZero keeps its meaning."""
    sk, P = _keys(rng)
    backend.set_pubkey(P)
    n = 200
    ms = [0 if i % 3 == 0 else i for i in range(n)]
    cts = bytearray(backend.encrypt_int_vector(ms, rng.scalars_bytes(n)).tobytes())
    cts[128 * 5:128 * 5 + 64] = bytes(64)
    ss = [0, 1, R.n - 1, R.n + 2] + [rng.scalar() for _ in range(n - 4)]
    sb = b''.join(s.to_bytes(32, 'big') for s in ss)
    got = backend.ct_scalar_mul(sb, bytes(cts)).tobytes()
    rc, want = oracle.ct_scalar_mul(sb, bytes(cts), threads=8)
    assert rc == 0 and got == want
    assert got[128 * 7:128 * 8] == R.ct_scalar_mul(bytes(cts[128 * 7:128 * 8]), ss[7])
    z = backend.decrypt_check_zero(sk.to_bytes(32, 'big'), got[128 * 6:128 * 30])
    assert list(z) == [0 if i % 3 == 0 else 1 for i in range(6, 30)]
    bad = bytearray(cts)
    bad[128 * 9 + 3] ^= 1
    with pytest.raises(RuntimeError):
        backend.ct_scalar_mul(sb, bytes(bad))


def test_shuffle_rerandomize_matches_oracle(backend, oracle, rng):
    """libunlynxshuffle.ShuffleSequence with beta and pi injected (data/data.go:68, services/service.go:618-660): 40
    lines of 25 ciphertexts; bytes equal the oracle's, plaintexts come back in permuted order, beta = 0 leaves a
    ciphertext unchanged, a non-permutation and a malformed point are refused."""
    sk, P = _keys(rng)
    backend.set_pubkey(P)
    k, nq = 40, 25
    g = np.random.default_rng(41)
    vals = g.integers(-100, 100, size=(k, nq))
    lines = [backend.encrypt_int_vector([int(v) for v in vals[i]], rng.scalars_bytes(nq)).tobytes() for i in range(k)]
    betas = [bytearray(rng.scalars_bytes(nq)) for _ in range(k)]
    betas[3][0:32] = bytes(32)
    betas[4][0:32] = (R.n + 1).to_bytes(32, 'big')
    pi = [int(v) for v in g.permutation(k)]
    out = backend.shuffle_rerandomize(lines, betas, pi)
    rc, want = oracle.shuffle_rerandomize(P, lines, betas, pi, threads=8)
    assert rc == 0 and [o.tobytes() for o in out] == want
    at3 = pi.index(3)
    assert out[at3].tobytes()[:128] == lines[3][:128] and out[at3].tobytes()[128:256] != lines[3][128:256]
    assert out[0].tobytes()[:128] == R.rerandomize(lines[pi[0]][:128], R.g1_from_bytes(P), int.from_bytes(betas[pi[0]][:32], 'big'))
    backend.create_decryption_table(10000)
    for i in (0, 7, 39):
        assert list(backend.decrypt_int_vector(sk.to_bytes(32, 'big'), out[i])) == [int(v) for v in vals[pi[i]]]
    with pytest.raises(RuntimeError):
        backend.shuffle_rerandomize(lines, betas, [0] * k)
    bad = [bytearray(x) for x in lines]
    bad[5][130] ^= 1
    with pytest.raises(RuntimeError):
        backend.shuffle_rerandomize(bad, betas, pi)


def test_schnorr_batch_matches_oracle(backend, oracle, rng):
    """kyber sign/schnorr Verify as VerifyProofSignature uses it (lib/proof/structs_proofs.go:498-505): 64 signatures
    by different keys over messages of different lengths, then one tampering of each kind; accept bits equal the
    oracle's and the definition's."""
    pubs, sigs, msgs, want = [], [], [], []
    g = np.random.default_rng(98)
    for i in range(64):
        x, k = rng.scalar(), rng.scalar()
        Xb = R.g1_to_bytes(R.g1_mul(x, R.G1_GEN)) if i < 4 else oracle.g1_mul(x.to_bytes(32, 'big'), oracle.generators()[0])
        msg = g.bytes(int(g.integers(0, 3000)))
        sig = oracle.schnorr_sign(x.to_bytes(32, 'big'), k.to_bytes(32, 'big'), Xb, msg)
        if i < 2:
            assert sig == R.schnorr_sign(x, k, msg)
        kind = i % 8 if i >= 32 else 0
        if kind == 1:
            msg = msg + b'x'
        elif kind == 2:
            sig = sig[:95] + bytes([sig[95] ^ 1])
        elif kind == 3:
            sig = sig[:20] + bytes([sig[20] ^ 1]) + sig[21:]          # R leaves the curve (or changes)
        elif kind == 4:
            Xb = oracle.g1_mul(rng.scalar().to_bytes(32, 'big'), oracle.generators()[0])
        elif kind == 5:
            sig = bytes(64) + sig[64:]                                  # R = identity
        elif kind == 6:
            sig = sig[:64] + b'\xff' * 32                               # s >= n
        pubs.append(Xb); sigs.append(sig); msgs.append(msg); want.append(1 if kind in (0, 7) else 0)
    got = backend.schnorr_verify(pubs, sigs, msgs)
    assert list(got) == list(oracle.schnorr_verify(pubs, sigs, msgs, threads=8)) == want
    for i in (0, 33, 34, 38):
        assert R.schnorr_verify(R.g1_from_bytes(pubs[i]), msgs[i], sigs[i]) == bool(want[i])


def test_reduce_rows_host_buffers(backend):
    """drx_reduce_rows against the Go loops' arithmetic (sum.go:19-23, mean.go:19-24, variance.go:19-27), wrap-around
    of int64 included, ragged sizes and an empty provider list."""
    g = np.random.default_rng(3)
    for ndp, nrows in [(1, 1), (3, 7), (100, 10000), (5, 0), (2, 257)]:
        rows = g.integers(-2 ** 40, 2 ** 40, size=(ndp, nrows), dtype=np.int64)
        s = rows.sum(axis=1)
        q = (rows * rows).sum(axis=1)  # wraps like Go's int64
        cnt = np.full(ndp, nrows, np.int64)
        assert (backend.reduce_rows(rows, 0) == s[:, None]).all()
        assert (backend.reduce_rows(rows, 1) == np.stack([s, cnt], 1)).all()
        assert (backend.reduce_rows(rows, 2) == np.stack([s, cnt, q], 1)).all()
        assert (backend.reduce_rows(rows, 3) == np.stack([s, cnt, s, cnt, q], 1)).all()
    assert backend.reduce_rows(np.empty((0, 4), np.int64), 3).shape == (0, 5)


def test_golden_encrypt_vectors(backend):
    with open(os.path.join(GOLD, 'bn256_golden.json')) as f:
        g = json.load(f)
    backend.set_pubkey(bytes.fromhex(g['encrypt'][0]['P']))
    for v in g['encrypt']:
        assert backend.encrypt_int_vector([v['m']], bytes.fromhex(v['r'])).tobytes().hex() == v['ct']
    for v in g['g1_mul']:
        assert backend.g1_mul(bytes.fromhex(v['k']), bytes.fromhex(v['p'])).tobytes().hex() == v['out']
    for v in g['g2_mul']:
        assert backend.g2_mul(bytes.fromhex(v['k']), bytes.fromhex(v['p'])).tobytes().hex() == v['out']
    for v in g['pair']:
        assert backend.pair(bytes.fromhex(v['g1']), bytes.fromhex(v['g2'])).tobytes().hex() == v['out']


@pytest.mark.parametrize('ndp,v', [(1, 1), (2, 3), (10, 1), (7, 33), (65, 5), (130, 2)])
def test_aggregate_bytes_match_oracle(backend, oracle, rng, ndp, v):
    sk, P = _keys(rng)
    backend.set_pubkey(P)
    vecs = []
    for d in range(ndp):
        m = [rng.randint(-1000, 1000) for _ in range(v)]
        vecs.append(backend.encrypt_int_vector(m, rng.scalars_bytes(v)).tobytes())
    rc, want = oracle.aggregate(vecs, threads=8)
    assert rc == 0
    assert backend.aggregate(vecs).tobytes() == want


def test_aggregate_exceptional_points(backend, oracle, rng):
    """P + P (doubling inside an add), P + (-P) (identity), identity inputs."""
    sk, P = _keys(rng)
    backend.set_pubkey(P)
    ct = backend.encrypt_int_vector([5], rng.scalars_bytes(1)).tobytes()
    K, C = R.g1_from_bytes(ct[:64]), R.g1_from_bytes(ct[64:])
    neg = R.g1_to_bytes(R.g1_neg(K)) + R.g1_to_bytes(R.g1_neg(C))
    zero = bytes(128)
    for vecs in ([ct, ct], [ct, neg], [zero, ct], [ct, zero, neg], [zero, zero], [ct, ct, ct, neg]):
        rc, want = oracle.aggregate(vecs)
        assert rc == 0
        assert backend.aggregate(vecs).tobytes() == want


def test_sum_mean_frequency_round_trips(backend, oracle, rng):
    """lib/encoding/sum_test.go:15-31 (-> -11955), mean_test.go:15-32 (sum and count),
    frequency_count_test.go:15-40 (histogram), with CreateDecryptionTable(10000) as services/api.go:50."""
    sk, P = _keys(rng)
    skb = sk.to_bytes(32, 'big')
    backend.set_pubkey(P)
    backend.create_decryption_table(10000)
    inputs = [0, 1, 2, 3, 4, 5, 6, 7, 8, 9, -12000]
    ct = backend.encrypt_int_vector([sum(inputs)], rng.scalars_bytes(1))
    assert list(backend.decrypt_int_vector(skb, ct)) == [-11955]
    ct = backend.encrypt_int_vector([sum(inputs), len(inputs)], rng.scalars_bytes(2))
    s, cnt = backend.decrypt_int_vector(skb, ct)
    assert s / cnt == sum(inputs) / len(inputs)
    data = [1, 2, 2, 3, 5, 5, 5, 9]
    hist = [data.count(b) for b in range(0, 10)]
    dps = [backend.encrypt_int_vector(hist, rng.scalars_bytes(10)).tobytes() for _ in range(3)]
    agg = backend.aggregate(dps)
    assert list(backend.decrypt_int_vector(skb, agg)) == [3 * h for h in hist]
    assert list(backend.decrypt_check_zero(skb, agg)) == [0 if h == 0 else 1 for h in hist]


def test_decrypt_matches_oracle_including_table_state(oracle, rng):
    """Fresh contexts on both sides; the same call sequence must give the same values and found flags,
    including the reference's lazily-grown-table quirks (negatives only stored while checkNeg grew it)."""
    from drynx_b200 import Backend
    be = Backend(0)
    try:
        sk, P = _keys(rng)
        skb = sk.to_bytes(32, 'big')
        be.set_pubkey(P)
        seq = [([5, 7, 2], False), ([-3, 4], False), ([-3, -20000, 20000], True), ([99999, -99999, 100000, -100000], True),
               ([-7, 0, 12], True), ([-1], False)]
        oracle.dlog_reset()
        for ms, with_neg in seq:
            ct = oracle.encrypt(P, ms, rng.scalars_bytes(len(ms)))
            want, wf = oracle.decrypt(skb, ct, with_neg=with_neg)
            got, gf = be.decrypt_int_vector(skb, ct, with_neg=with_neg, return_found=True)
            assert list(got) == list(want) and list(gf) == list(wf), (ms, with_neg)
    finally:
        be.close()


def test_full_size_frequency_count_properties(backend, rng):
    """BASELINE config 4 at full size: 1,000 DPs x 1,000 bins = 10^6 encryptions (128 MB of ciphertexts), root
    aggregation 1000 x 1000 -> 1000, decryption with the 10,000-entry table request. Property checks:
    decrypt(aggregate(encrypt(x_d))) == sum_d x_d, and aggregation is order-independent."""
    sk, P = _keys(rng)
    skb = sk.to_bytes(32, 'big')
    backend.set_pubkey(P)
    ndp, v = 1000, 1000
    backend.create_decryption_table(10000)
    g = np.random.default_rng(4)
    counts = g.integers(0, 10, size=(ndp, v), dtype=np.int64)
    r = np.frombuffer(g.bytes(32 * ndp * v), np.uint8).copy()
    r[::32] &= 0x7f
    cts = backend.encrypt_int_vector(counts.ravel(), r).reshape(ndp, v * 128)
    agg = backend.aggregate([cts[d] for d in range(ndp)])
    assert list(backend.decrypt_int_vector(skb, agg)) == list(counts.sum(axis=0))
    perm = g.permutation(ndp)
    assert backend.aggregate([cts[d] for d in perm]).tobytes() == agg.tobytes()
