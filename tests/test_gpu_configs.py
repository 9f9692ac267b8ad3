"""-m gpu: the remaining BASELINE.json configs as parity cases, at sizes the oracle can
follow in seconds plus size-independent properties at larger sizes."""
import numpy as np
import pytest

import bn256_ref as R
from conftest import SplitMix

pytestmark = pytest.mark.gpu


def _keys(rng):
    sk = rng.scalar()
    return sk, R.g1_to_bytes(R.g1_mul(sk, R.G1_GEN))


def test_config1_sum_10_dps_100_rows(backend, oracle, rng):
    """Config 1: sum query, 10 DPs x 100 rows U{1,2}, proofs off: encrypt(10) -> aggregate(10x1) -> decrypt(1);
    ciphertext bytes equal the oracle's at every stage."""
    sk, P = _keys(rng)
    backend.set_pubkey(P)
    backend.create_decryption_table(10000)
    g = np.random.default_rng(1)
    rows = g.integers(1, 3, size=(10, 100), dtype=np.int64)
    sums = rows.sum(axis=1)
    rs = rng.scalars_bytes(10)
    cts = backend.encrypt_int_vector(sums, rs).tobytes()
    assert cts == oracle.encrypt(P, sums, rs)
    vecs = [cts[128 * d:128 * d + 128] for d in range(10)]
    agg = backend.aggregate(vecs).tobytes()
    assert agg == oracle.aggregate(vecs)[1]
    oracle.dlog_reset()
    want, _ = oracle.decrypt(sk.to_bytes(32, 'big'), agg)
    got = backend.decrypt_int_vector(sk.to_bytes(32, 'big'), agg)
    assert list(got) == list(want) == [int(rows.sum())]


def test_config1_over_edwards25519(backend, oracle):
    """Config 1 as BASELINE.json words it (Ed25519 ElGamal, proofs off): 10 DPs x 100 rows U{1,2}, per-provider sum on
    the device, encrypt(10) -> aggregate(10 x 1) -> decrypt(1); bytes equal the edwards25519 oracle's at every stage."""
    import ed25519_ref as E
    g = np.random.default_rng(1)
    sk = int.from_bytes(g.bytes(31), 'little') % E.l
    skb = sk.to_bytes(32, 'little')
    P = oracle.ed_mul(skb)
    backend.ed_set_pubkey(P)
    backend.ed_create_decryption_table(10000)
    rows = g.integers(1, 3, size=(10, 100), dtype=np.int64)
    sums = backend.reduce_rows(rows, 0).ravel()
    assert list(sums) == list(rows.sum(axis=1))
    rs = g.bytes(32 * 10)
    cts = backend.ed_encrypt_int_vector(sums, rs).tobytes()
    assert cts == oracle.ed_encrypt(P, sums, rs)
    vecs = [cts[64 * d:64 * d + 64] for d in range(10)]
    agg = backend.ed_aggregate(vecs).tobytes()
    assert agg == oracle.ed_aggregate(vecs)[1]
    oracle.ed_dlog_reset()
    rc, want, _ = oracle.ed_decrypt(skb, agg)
    assert rc == 0
    assert list(backend.ed_decrypt_int_vector(skb, agg)) == list(want) == [int(rows.sum())]
    oracle.ed_dlog_reset()


def test_config4_from_rows(backend, rng):
    """Config 4 from raw rows: 200 providers x 5,000 rows over 1,000 bins -> per-provider histogram on the device
    (frequency_count.go:18-36) -> encrypt -> aggregate -> decrypt == the global histogram; out-of-range rows are an
    error like the Go index panic."""
    from drynx_b200 import DrynxGpuError
    sk, P = _keys(rng)
    backend.set_pubkey(P)
    backend.create_decryption_table(10000)
    g = np.random.default_rng(44)
    ndp, nrows, lo, hi = 200, 5000, -3, 996
    rows = g.integers(lo, hi + 1, size=(ndp, nrows), dtype=np.int64)
    counts = backend.freq_count(rows, lo, hi)
    want = np.stack([np.bincount(rows[d] - lo, minlength=hi - lo + 1) for d in range(ndp)])
    assert (counts == want).all()
    rb = np.frombuffer(g.bytes(32 * counts.size), np.uint8).copy()
    rb[::32] &= 0x7f
    cts = backend.encrypt_int_vector(counts.ravel(), rb).reshape(ndp, -1)
    agg = backend.aggregate([cts[d] for d in range(ndp)])
    assert list(backend.decrypt_int_vector(sk.to_bytes(32, 'big'), agg)) == list(want.sum(axis=0))
    assert backend.freq_count(np.empty((0, 7), np.int64), 0, 3).shape == (0, 4)
    assert (backend.freq_count(np.empty((2, 0), np.int64), 0, 3) == 0).all()
    rows[5, 17] = hi + 1
    with pytest.raises(DrynxGpuError):
        backend.freq_count(rows, lo, hi)


def _all_coeffs(x, y, k=2):
    """logistic_regression.go:367-403 for k = 2 on an augmented record: level 1 = x(2y-1), level 2 = -(x_a x_b)."""
    lvl1 = x * (2.0 * y - 1.0)
    ypart = float(y - y * 1 - 1)  # (-1)^2 = 1
    lvl2 = ypart * np.outer(1.0 * x, x).ravel()
    return lvl1, lvl2


def test_config3_logreg_shape(backend, oracle, rng):
    """Config 3 shape (d = 8, k = 2 -> 90 coefficients per DP; 10 DPs), 2,000 synthetic records: per-DP float64
    coefficient sums in record order, rounded with precision 1, encrypted, aggregated, decrypted == plaintext sum."""
    sk, P = _keys(rng)
    backend.set_pubkey(P)
    backend.create_decryption_table(10000)
    g = np.random.default_rng(3)
    ndp, per_dp, d = 10, 200, 8
    X = g.integers(0, 4, size=(ndp, per_dp, d)).astype(np.float64)
    Y = g.integers(0, 2, size=(ndp, per_dp)).astype(np.int64)
    means, sds = np.zeros(d), np.full(d, 100.0)
    packed = np.zeros((ndp, 90), dtype=np.int64)
    for dp in range(ndp):
        a1, a2 = np.zeros(9), np.zeros(81)
        for r in range(per_dp):  # serial, index order: the float64 order is part of the result
            x = np.concatenate([[1.0], (X[dp, r] - means) / sds])
            l1, l2 = _all_coeffs(x, int(Y[dp, r]))
            a1 += l1
            a2 += l2
        packed[dp, :9] = np.round(a1 * 1.0).astype(np.int64)
        packed[dp, 9:] = np.round(a2 * 1.0).astype(np.int64)
    rs = rng.scalars_bytes(ndp * 90)
    cts = backend.encrypt_int_vector(packed.ravel(), rs).tobytes()
    assert cts[:128 * 90] == oracle.encrypt(P, packed[0], rs[:32 * 90])
    vecs = [cts[128 * 90 * dp:128 * 90 * (dp + 1)] for dp in range(ndp)]
    agg = backend.aggregate(vecs)
    assert list(backend.decrypt_int_vector(sk.to_bytes(32, 'big'), agg)) == list(packed.sum(axis=0))


@pytest.mark.parametrize('ns,l', [(1, 2), (3, 2), (1, 8), (3, 8)])
def test_config5_verify_sweep_point(backend, oracle, rng, ns, l):
    """Config 5 points (u = 16, l in {2, 8}, ns in {1, 3}): 96 proofs, 1 % (here 3) tampered in Zr / A / V;
    accept / reject vector identical to the oracle's on a subsample and exactly the tampered set overall."""
    u = 16
    sk, P = _keys(rng)
    backend.set_pubkey(P)
    xs = [rng.scalar() for _ in range(ns)]
    Y, A = backend.init_range_proof_signature(xs, u)
    backend.set_signatures(40, ns, u, Y, A)
    n = 96
    g = np.random.default_rng(5)
    secrets = g.integers(0, u ** l, size=n, dtype=np.int64)
    rs = rng.scalars_bytes(n)
    cts = backend.encrypt_int_vector(secrets, rs).tobytes()
    rnd = [rng.scalars_bytes(l * (3 + ns)) for _ in range(n)]
    proofs, off = backend.create_range_proofs([40] * n, {40: ns}, {40: u}, [l] * n, secrets, rs, cts, rnd)
    size = int(off[1])
    bad = bytearray(proofs.tobytes())
    blk = 256 + 32 * l
    bad[7 * size + 165] ^= 2                      # Zr
    bad[40 * size + blk + 128 * l + 77] ^= 1      # A_00
    bad[95 * size + blk + 9] ^= 1                 # V_00 (almost surely leaves the curve)
    ok = backend.verify_range_proofs([40] * n, [l] * n, bytes(bad), off)
    assert [i for i in range(n) if not ok[i]] == [7, 40, 95]
    sub = [0, 7, 40, 41, 95]
    for i in sub:
        want = oracle.rp_verify(ns, u, l, Y.tobytes(), P, bytes(bad[i * size:(i + 1) * size]), np.array([0, size], np.uint64))
        assert want[0] == ok[i]
