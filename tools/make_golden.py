#!/usr/bin/env python3
"""Writes tests/golden/bn256_golden.json from the Python big-integer definition (oracle/bn256_ref.py).

The reference holds no golden vectors for this path and cannot run here, so these vectors pin
the *definition* every implementation in this repo (C++ oracle, host-emulated device code, CUDA kernels) must
reproduce; they are NOT outputs of the Go reference ("parity unpinned")."""
import json
import os
import sys

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, os.path.join(ROOT, 'oracle'))
sys.path.insert(0, os.path.join(ROOT, 'tests'))
import bn256_ref as R  # noqa: E402
from conftest import SplitMix  # noqa: E402

rng = SplitMix(2026)
g = {'g1_mul': [], 'g2_mul': [], 'pair': [], 'encrypt': [], 'range_proof': []}
for k in [1, 2, R.n - 1, rng.scalar(), rng.scalar()]:
    P = R.g1_mul(rng.scalar(), R.G1_GEN)
    g['g1_mul'].append({'k': k.to_bytes(32, 'big').hex(), 'p': R.g1_to_bytes(P).hex(), 'out': R.g1_to_bytes(R.g1_mul(k, P)).hex()})
    Q = R.g2_mul(rng.scalar(), R.G2_GEN)
    g['g2_mul'].append({'k': k.to_bytes(32, 'big').hex(), 'p': R.g2_to_bytes(Q).hex(), 'out': R.g2_to_bytes(R.g2_mul(k, Q)).hex()})
for _ in range(3):
    P = R.g1_mul(rng.scalar(), R.G1_GEN)
    Q = R.g2_mul(rng.scalar(), R.G2_GEN)
    g['pair'].append({'g1': R.g1_to_bytes(P).hex(), 'g2': R.g2_to_bytes(Q).hex(), 'out': R.f12_to_bytes(R.pair(P, Q)).hex()})
g['pair'].append({'g1': R.g1_to_bytes(R.G1_GEN).hex(), 'g2': R.g2_to_bytes(R.G2_GEN).hex(),
                  'out': R.f12_to_bytes(R.pair(R.G1_GEN, R.G2_GEN)).hex()})
sk = rng.scalar()
Pp = R.g1_mul(sk, R.G1_GEN)
for m in [0, 1, -1, 42, -11955, 99999, 2 ** 40 + 3]:
    r = rng.scalar()
    g['encrypt'].append({'P': R.g1_to_bytes(Pp).hex(), 'sk': sk.to_bytes(32, 'big').hex(), 'm': m,
                         'r': r.to_bytes(32, 'big').hex(), 'ct': R.encrypt(Pp, m, r).hex()})
for (ns, u, l, secret) in [(1, 2, 1, 1), (2, 4, 2, 9), (3, 16, 2, 200), (1, 16, 1, 300)]:
    xs = [rng.scalar() for _ in range(ns)]
    sigs = [R.init_range_sig(x, u) for x in xs]
    r = rng.scalar()
    ct = R.encrypt(Pp, secret, r)
    nd = len(R.to_base(secret, u, l))
    rnd = [rng.scalar() for _ in range(nd * (3 + ns))]
    proof = R.rp_create(sigs, u, l, secret, r, Pp, ct, rnd)
    ok = R.rp_verify(proof, u, l, [s[0] for s in sigs], Pp)
    assert ok == (nd == l), (ns, u, l, secret, ok)
    g['range_proof'].append({
        'ns': ns, 'u': u, 'l': l, 'secret': secret, 'P': R.g1_to_bytes(Pp).hex(),
        'x': b''.join(x.to_bytes(32, 'big') for x in xs).hex(),
        'Y': b''.join(R.g1_to_bytes(s[0]) for s in sigs).hex(),
        'A': b''.join(R.g2_to_bytes(a) for s in sigs for a in s[1]).hex(),
        'r': r.to_bytes(32, 'big').hex(), 'ct': ct.hex(),
        'rnd': b''.join(x.to_bytes(32, 'big') for x in rnd).hex(), 'proof': proof.hex(), 'valid': ok})
os.makedirs(os.path.join(ROOT, 'tests', 'golden'), exist_ok=True)
with open(os.path.join(ROOT, 'tests', 'golden', 'bn256_golden.json'), 'w') as f:
    json.dump(g, f, indent=0)
print('wrote', sum(len(v) for v in g.values()), 'vectors')
