"""Row a-7, the float64 pre-pass of logistic regression: the numpy oracle (oracle/logreg_ref.py) against the reference's
own known answers and against an operation-by-operation Python loop (CPU), and the CUDA kernel k_logreg_coeffs through
the C ABI against the oracle, bit for bit (-m gpu)."""
import numpy as np
import pytest

import logreg_ref as L


def test_oracle_reproduces_reference_known_answers():
    """logistic_regression_test.go:20-72 (ComputeDistinctApproxCoefficients on X = [0,1,2,3,4]): the All variant holds
    the same products at the positions of the distinct ones; with one record, means 0, sd 1 and the leading 1.0 the
    packed vector is exactly [X*(2y-1) | -(x_a x_b)]."""
    x = [[1.0, 2.0, 3.0, 4.0]]  # augmented record = [1, 1, 2, 3, 4]; the test's X is [0, 1, 2, 3, 4] without Augment
    for yv, s in ((1, 1), (0, -1)):
        out = L.coefficients(x, [yv], [0, 0, 0, 0], [1, 1, 1, 1], 2, 1.0)
        X = [1, 1, 2, 3, 4]
        assert list(out[:5]) == [s * v for v in X]
        assert list(out[5:]) == [-(a * b) for a in X for b in X]
    # the packed vector of two records used by host/test_host.cpp (derived by hand from :367-455)
    assert list(L.coefficients([[1, 2], [3, 4]], [1, 0], [0, 0], [1, 1], 2, 1.0)) == [0, -2, -2, -2, -4, -6, -4, -10, -14, -6, -14, -20]


def test_numpy_form_equals_operation_by_operation_loop():
    g = np.random.default_rng(1)
    for n, d in [(1, 2), (7, 3), (300, 8), (0, 2), (33, 1)]:
        x = g.integers(0, 4, size=(n, d)).astype(np.float64) + g.random((n, d))
        y = g.integers(0, 2, size=n)
        means, sds = list(g.random(d)), list(1 + g.random(d) * 50)
        for k in (1, 2):
            for prec in (1.0, 100.0, 1234.5):
                assert (L.coefficients(x, y, means, sds, k, prec) == L.coefficients_loop(x, y, means, sds, k, prec)).all()
    # rounding: halves go away from zero (math.Round), sums 3.0 and 4.5, then -4.5
    assert list(L.coefficients([[0.5], [1.5], [2.5]], [1, 1, 1], [0], [1], 1, 1.0)) == [3, 5]
    assert list(L.coefficients([[0.5], [1.5], [2.5]], [0, 0, 0], [0], [1], 1, 1.0)) == [-3, -5]


@pytest.mark.gpu
def test_device_prepass_matches_oracle_bit_for_bit(backend):
    g = np.random.default_rng(7)
    for d, k, sizes in [(2, 2, [2, 0, 5]), (8, 2, [1, 127, 128, 129, 1000]), (8, 1, [40, 3]), (31, 2, [70]), (1, 2, [500])]:
        means, sds = list(g.random(d) * 3), list(0.5 + g.random(d) * 100)
        xs = [g.integers(0, 4, size=(n, d)).astype(np.float64) + g.random((n, d)) * (g.random() < 0.7) for n in sizes]
        ys = [g.integers(0, 2, size=n).astype(np.int64) for n in sizes]
        for prec in (1.0, 1e3, 12345.678):
            got = backend.logreg_coefficients(xs, ys, means, sds, k, prec)
            for p in range(len(sizes)):
                want = L.coefficients(xs[p], ys[p], means, sds, k, prec)
                assert (got[p] == want).all(), (d, k, sizes[p], prec)
    # values whose sums are not exactly representable: the summation order is visible in the last bit
    n, d = 5000, 3
    x = g.random((n, d)) * 1e-3 + np.array([1e6, 1.0, 1e-6])
    y = g.integers(0, 2, size=n).astype(np.int64)
    got = backend.logreg_coefficients([x], [y], [0.1, 0.2, 0.3], [3.0, 7.0, 11.0], 2, 1e3)
    assert (got[0] == L.coefficients(x, y, [0.1, 0.2, 0.3], [3.0, 7.0, 11.0], 2, 1e3)).all()
    with pytest.raises(Exception):
        backend.logreg_coefficients([x], [y], [0.1, 0.2, 0.3], [3.0, 7.0, 11.0], 3, 1.0)


@pytest.mark.gpu
def test_config3_full_size(backend):
    """BASELINE config 3: 10^6 records x 8 features U{0..3} as float64, labels U{0,1}, 10 providers x 10^5 records,
    means 0, sd 100, k = 2, precision 1: 10 x 90 coefficients identical to the oracle; then encrypted, aggregated and
    decrypted they give the sums of the providers' coefficients."""
    g = np.random.default_rng(3)
    ndp, n, d = 10, 100000, 8
    xs = [g.integers(0, 4, size=(n, d)).astype(np.float64) for _ in range(ndp)]
    ys = [g.integers(0, 2, size=n).astype(np.int64) for _ in range(ndp)]
    means, sds = [0.0] * d, [100.0] * d
    got = backend.logreg_coefficients(xs, ys, means, sds, 2, 1.0)
    for p in range(ndp):
        assert (got[p] == L.coefficients(xs[p], ys[p], means, sds, 2, 1.0)).all()
    # a slice small enough for the decryption table (|sum| < 100000: 10 providers x 5,000 records): the homomorphic
    # sum of the encrypted coefficient vectors decrypts to the sum of the plaintext ones
    got = backend.logreg_coefficients([a[:5000] for a in xs], [a[:5000] for a in ys], means, sds, 2, 1.0)
    import bn256_ref as R
    sk = 0x1234567
    backend.set_pubkey(R.g1_to_bytes(R.g1_mul(sk, R.G1_GEN)))
    backend.create_decryption_table(10000)
    r = g.bytes(32 * got.size)
    rb = np.frombuffer(r, np.uint8).copy()
    rb[::32] &= 0x7f
    cts = backend.encrypt_int_vector(got.ravel(), rb).reshape(ndp, -1)
    agg = backend.aggregate([cts[p] for p in range(ndp)])
    dec = backend.decrypt_int_vector(sk.to_bytes(32, 'big'), agg)
    assert list(dec) == list(got.sum(axis=0))
