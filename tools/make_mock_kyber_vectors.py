"""Writes a file with the SCHEMA of tests/golden/kyber_vectors.json from this repository's own Python definition
(oracle/bn256_ref.py). It is a MOCK: it exercises tests/test_kyber_vectors.py end to end (the replay of the recorded
random stream, every comparison) so that the first real vector file, produced by go/vectors/dump_test.go on a machine
with Go, meets a loader that is known to run. It pins nothing and must never be committed as the golden file.

usage: python tools/make_mock_kyber_vectors.py /tmp/mock_vectors.json
       DRYNX_KYBER_VECTORS=/tmp/mock_vectors.json python -m pytest tests/test_kyber_vectors.py
"""
import hashlib
import json
import os
import sys

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, os.path.join(ROOT, 'oracle'))
import bn256_ref as R  # noqa: E402

hx = lambda b: bytes(b).hex()
sc = lambda v: (v % R.n).to_bytes(32, 'big')


def main(path):
    v = {'kyber': 'MOCK (oracle/bn256_ref.py, not kyber)', 'suite': 'bn256.G1'}
    v['g1_base'], v['g2_base'], v['g1_null'] = hx(R.g1_to_bytes(R.G1_GEN)), hx(R.g2_to_bytes(R.G2_GEN)), hx(bytes(64))
    big = int.from_bytes(b'a fixed scalar well above 2^64 for the vector file, reduced mod n by SetBytes', 'big') % R.n
    ks = [1, 2, 3, 12, 65537, 1234567891011, big]
    v['g1_mul_base'] = [{'k': hx(sc(k)), 'out': hx(R.g1_to_bytes(R.g1_mul(k, R.G1_GEN)))} for k in ks + [R.n - 5]]
    v['g2_mul_base'] = [{'k': hx(sc(k)), 'out': hx(R.g2_to_bytes(R.g2_mul(k, R.G2_GEN)))} for k in ks]
    v['scalar_setint64_minus5'] = hx(sc(-5))
    s64 = b'0123456789abcdef0123456789abcdef0123456789abcdef0123456789abcdef'
    v['scalar_setbytes_64'] = {'in': hx(s64), 'out': hx(sc(int.from_bytes(s64, 'big')))}
    v['pair'] = []
    for a, b in ((1, 1), (2, 3), (65537, 12)):
        P, Q = R.g1_mul(a, R.G1_GEN), R.g2_mul(b, R.G2_GEN)
        v['pair'].append({'g1': hx(R.g1_to_bytes(P)), 'g2': hx(R.g2_to_bytes(Q)), 'out': hx(R.f12_to_bytes(R.pair(P, Q)))})
    e = R.pair(R.G1_GEN, R.G2_GEN)
    v['gt_mul'] = {'k': hx(sc(big)), 'out': hx(R.f12_to_bytes(R.f12pow(e, big))), 'add_e_e': hx(R.f12_to_bytes(R.f12mul(e, e)))}
    sk = int.from_bytes(b'collective secret key of the vector file', 'big') % R.n
    P = R.g1_mul(sk, R.G1_GEN)
    encs = []
    for i, m in enumerate((0, 1, 25, -11955, 99999)):
        r = int.from_bytes(hashlib.sha256(b'r%d' % i).digest(), 'big') % R.n
        encs.append({'m': m, 'r': hx(sc(r)), 'ct': hx(R.encrypt(P, m, r)), 'dec': m})
    v['elgamal'] = {'sk': hx(sc(sk)), 'pub': hx(R.g1_to_bytes(P)), 'enc': encs}
    a, b = R.encrypt(P, 40, 77), R.encrypt(P, 2, 99)
    K = R.g1_add(R.g1_from_bytes(a[:64]), R.g1_from_bytes(b[:64]))
    C = R.g1_add(R.g1_from_bytes(a[64:]), R.g1_from_bytes(b[64:]))
    v['aggregate'] = {'a': hx(a), 'b': hx(b), 'sum': hx(R.g1_to_bytes(K) + R.g1_to_bytes(C))}
    Y, A = R.init_range_sig(12, 16)
    v['sig_deterministic_u16'] = {'Y': hx(R.g1_to_bytes(Y)), 'A': hx(b''.join(R.g2_to_bytes(x) for x in A))}
    ns, u, l, secret = 3, 16, 4, 40000
    xs = [int.from_bytes(hashlib.sha256(b'x%d' % i).digest(), 'big') % R.n for i in range(ns)]
    sigs = [R.init_range_sig(x, u) for x in xs]
    r = int.from_bytes(hashlib.sha256(b'rp r').digest(), 'big') % R.n
    ct = R.encrypt(P, secret, r)
    # a random stream with a rejected attempt in it (a value >= n), as the real recording will have now and then
    stream, picks = b'', []
    k = 0
    while len(picks) < l * (3 + ns):
        chunk = hashlib.sha256(b'stream%d' % k).digest()
        k += 1
        if k == 3:
            chunk = b'\xff' * 32
        stream += chunk
        val = int.from_bytes(chunk, 'big')
        if 0 < val < R.n:
            picks.append(val)
    proof = R.rp_create(sigs, u, l, secret, r, P, ct, picks)
    v['range_proof'] = {'ns': ns, 'u': u, 'l': l, 'secret': secret, 'pub': hx(R.g1_to_bytes(P)), 'r': hx(sc(r)), 'ct': hx(ct),
                        'Y': [hx(R.g1_to_bytes(s[0])) for s in sigs], 'A': [hx(b''.join(R.g2_to_bytes(x) for x in s[1])) for s in sigs],
                        'stream': hx(stream), 'proof': hx(proof), 'verifies': True}
    priv = int.from_bytes(b'schnorr private key of the vector file', 'big') % R.n
    msg = b'a serialised proof request'
    v['schnorr'] = {'priv': hx(sc(priv)), 'pub': hx(R.g1_to_bytes(R.g1_mul(priv, R.G1_GEN))), 'msg': hx(msg),
                    'sig': hx(R.schnorr_sign(priv, 424242, msg))}
    T = R.g2_twist_point(1)
    v['g2_outside_subgroup'] = [{'point': hx(R.g2_to_bytes(Q)), 'accepted': False, 'error': 'mock'}
                                for Q in (T, R.g2_mul_raw(R.n * R.TWIST_COFACTOR // 13, T))]
    over = bytearray(R.g1_to_bytes(R.G1_GEN))
    over[0] |= 0x80
    v['g1_coordinate_above_p'] = {'point': hx(over), 'accepted': False}
    with open(path, 'w') as f:
        json.dump(v, f, indent=1)
    print('wrote MOCK vectors to', path)


if __name__ == '__main__':
    main(sys.argv[1])
