"""-m gpu: the CUDA edwards25519 ElGamal path (kernels_ed.cu) through the C ABI against the C++ oracle (bit-exact),
the committed golden vectors and RFC 8032's known answers; full-size runs are checked through round-trip properties."""
import json
import os

import numpy as np
import pytest

import ed25519_ref as E
from drynx_b200 import DrynxGpuError

pytestmark = pytest.mark.gpu
GOLD = os.path.join(os.path.dirname(__file__), 'golden', 'ed25519_golden.json')


def _sc(v):
    return int(v).to_bytes(32, 'little')


def _keys(rng):
    sk = rng.scalar() % E.l
    return sk, E.to_bytes(E.mul(sk, E.BASE))


def test_rfc8032_and_golden_scalar_mults(backend):
    with open(GOLD) as f:
        g = json.load(f)
    B = E.to_bytes(E.BASE)
    for v in g['rfc8032']:
        assert backend.ed_point_mul(bytes.fromhex(v['scalar_mod_l']), B).tobytes().hex() == v['public']
    ks = b''.join(bytes.fromhex(v['k']) for v in g['mul'])
    ps = b''.join(bytes.fromhex(v['p']) for v in g['mul'])
    assert backend.ed_point_mul(ks, ps).tobytes().hex() == ''.join(v['out'] for v in g['mul'])


def test_golden_encrypt_aggregate_decrypt(backend):
    with open(GOLD) as f:
        g = json.load(f)
    backend.ed_set_pubkey(bytes.fromhex(g['encrypt'][0]['P']))
    ms = [v['m'] for v in g['encrypt']]
    rb = b''.join(bytes.fromhex(v['r']) for v in g['encrypt'])
    assert backend.ed_encrypt_int_vector(ms, rb).tobytes().hex() == ''.join(v['ct'] for v in g['encrypt'])
    a = g['aggregate']
    agg = backend.ed_aggregate([bytes.fromhex(x) for x in a['inputs']])
    assert agg.tobytes().hex() == a['out']
    sk = bytes.fromhex(g['encrypt'][0]['sk'])
    assert list(backend.ed_decrypt_int_vector(sk, agg)) == [a['plain_sum']]


def test_point_decoding_rules(backend):
    one = _sc(1)
    with pytest.raises(DrynxGpuError):
        backend.ed_point_mul(one, _sc(2))                       # y = 2: no x on the curve
    with pytest.raises(DrynxGpuError):
        backend.ed_set_pubkey(_sc(2))
    ident = E.to_bytes(E.IDENT)
    assert backend.ed_point_mul(one, _sc(E.p + 1)).tobytes() == ident     # non-canonical y accepted (ref10)
    assert backend.ed_point_mul(one, _sc(1 | 1 << 255)).tobytes() == ident  # "negative zero"
    # all decodable y < 48, both signs, including the small-order points: decode -> 1 * P -> encode is the identity map
    encs, want = [], []
    for y in range(48):
        for sign in (0, 1):
            P = E.from_bytes(_sc(y | sign << 255))
            if P is not None:
                encs.append(_sc(y | sign << 255))
                want.append(E.to_bytes(P))
    got = backend.ed_point_mul(one * len(encs), b''.join(encs)).tobytes()
    assert got == b''.join(want)


def test_encrypt_bytes_match_oracle(backend, oracle, rng):
    sk, P = _keys(rng)
    backend.ed_set_pubkey(P)
    ms = [0, 1, -1, 9, -12000, 99999, 100000, 2 ** 40 + 3, -(2 ** 62), 2 ** 63 - 1, -(2 ** 63)]
    ms += [rng.randint(-10 ** 6, 10 ** 6) for _ in range(300)]
    rs = [rng.scalar() | (rng.u64() << 192) for _ in ms]      # full 256-bit inputs: reduced mod l like SetBytes
    rs[3] = 0
    rs[4] = E.l - 1
    rs[5] = E.l
    rs[6] = 2 ** 256 - 1
    rb = b''.join(_sc(r) for r in rs)
    got = backend.ed_encrypt_int_vector(ms, rb).tobytes()
    assert got == oracle.ed_encrypt(P, ms, rb, threads=8)
    assert backend.ed_encrypt_int_vector([], b'').size == 0


def test_encrypt_lane_split_and_one_thread_kernels_agree(backend, oracle, rng):
    sk, P = _keys(rng)
    backend.ed_set_pubkey(P)
    n = 17000
    g = np.random.default_rng(8)
    ms = g.integers(-10 ** 9, 10 ** 9, size=n, dtype=np.int64)
    ms[:6] = [0, -1, 2 ** 63 - 1, -(2 ** 63), 255, 256]
    rb = g.bytes(32 * n)
    one = backend.ed_encrypt_int_vector(ms, rb).tobytes()
    h = n // 2
    two = backend.ed_encrypt_int_vector(ms[:h], rb[:32 * h]).tobytes() + backend.ed_encrypt_int_vector(ms[h:], rb[32 * h:]).tobytes()
    assert one == two
    assert one[:64 * 64] == oracle.ed_encrypt(P, [int(x) for x in ms[:64]], rb[:32 * 64], threads=8)


@pytest.mark.parametrize('ndp,v', [(1, 1), (2, 3), (10, 1), (37, 5), (100, 64), (1500, 2)])
def test_aggregate_bytes_match_oracle(backend, oracle, rng, ndp, v):
    sk, P = _keys(rng)
    backend.ed_set_pubkey(P)
    g = np.random.default_rng(ndp * 131 + v)
    ms = g.integers(-1000, 1000, size=ndp * v, dtype=np.int64)
    cts = backend.ed_encrypt_int_vector(ms, g.bytes(32 * ndp * v)).reshape(ndp, v * 64)
    vecs = [cts[d].tobytes() for d in range(ndp)]
    rc, want = oracle.ed_aggregate(vecs, threads=8)
    assert rc == 0
    assert backend.ed_aggregate(vecs).tobytes() == want


def test_aggregate_exceptional_and_malformed(backend, oracle, rng):
    P1 = E.mul(5, E.BASE)
    ident = E.to_bytes(E.IDENT)
    ct = lambda a, b: E.to_bytes(a) + E.to_bytes(b)
    # P + P, P + (-P), identity + P, small-order point + P
    T = next(E.from_bytes(_sc(y)) for y in range(3, 60) if E.from_bytes(_sc(y)) is not None)
    vecs = [ct(P1, P1) + ct(P1, E.IDENT) + ct(T, P1), ct(P1, E.neg(P1)) + ct(E.IDENT, E.IDENT) + ct(P1, T)]
    rc, want = oracle.ed_aggregate(vecs)
    assert rc == 0
    assert backend.ed_aggregate(vecs).tobytes() == want
    assert want[32:64] == ident
    bad = bytearray(vecs[0])
    bad[0:32] = _sc(2)
    with pytest.raises(DrynxGpuError):
        backend.ed_aggregate([bytes(bad), vecs[1]])


def test_decrypt_matches_oracle_including_table_state(oracle, rng):
    """Same lazily-grown-table semantics as the BN256 path (unlynx discreteLog): a fresh context per scenario."""
    from drynx_b200 import Backend
    sk, P = _keys(rng)
    skb = _sc(sk)
    scenarios = [[5, -3, 250, 0, -251, 7], [99999, 3, -99999, 100000, -100000, 12], [-40, 40, 2 ** 40, -7, 60]]
    for ms in scenarios:
        be = Backend(0)
        try:
            be.ed_set_pubkey(P)
            oracle.ed_dlog_reset()
            g = np.random.default_rng(len(ms))
            ct = oracle.ed_encrypt(P, ms, g.bytes(32 * len(ms)))
            for i in range(len(ms)):  # one element at a time: the table state between calls matters
                rc, want, wfound = oracle.ed_decrypt(skb, ct[64 * i:64 * i + 64])
                assert rc == 0
                got, found = be.ed_decrypt_int_vector(skb, ct[64 * i:64 * i + 64], return_found=True)
                assert (int(got[0]), int(found[0])) == (int(want[0]), int(wfound[0])), (ms, i)
        finally:
            be.close()
    oracle.ed_dlog_reset()


def test_full_size_frequency_count_properties(backend, rng):
    """BASELINE config 4 shape over edwards25519: 1,000 DPs x 1,000 bins = 10^6 encryptions (64 MB of ciphertexts),
    aggregation 1000 x 1000 -> 1000, decryption with the 10,000-entry table: decrypt(aggregate(encrypt(x_d))) is
    sum_d x_d and aggregation does not depend on the order of the providers."""
    sk, P = _keys(rng)
    backend.ed_set_pubkey(P)
    ndp, v = 1000, 1000
    backend.ed_create_decryption_table(10000)
    g = np.random.default_rng(4)
    counts = g.integers(0, 10, size=(ndp, v), dtype=np.int64)
    cts = backend.ed_encrypt_int_vector(counts.ravel(), g.bytes(32 * ndp * v)).reshape(ndp, v * 64)
    agg = backend.ed_aggregate([cts[d] for d in range(ndp)])
    assert list(backend.ed_decrypt_int_vector(_sc(sk), agg)) == list(counts.sum(axis=0))
    perm = g.permutation(ndp)
    assert backend.ed_aggregate([cts[d] for d in perm]).tobytes() == agg.tobytes()
