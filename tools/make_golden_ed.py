#!/usr/bin/env python3
"""Writes tests/golden/ed25519_golden.json from the Python big-integer definition (oracle/ed25519_ref.py), which
checks itself against RFC 8032 section 7.1 on import. The RFC's own vectors are stored alongside (published known
answers); the ElGamal vectors pin unlynx's encryptPoint over this suite as restated in this repo."""
import hashlib
import json
import os
import sys

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, os.path.join(ROOT, 'oracle'))
sys.path.insert(0, os.path.join(ROOT, 'tests'))
import ed25519_ref as E  # noqa: E402
from conftest import SplitMix  # noqa: E402

rng = SplitMix(25519)
sc = lambda v: int(v).to_bytes(32, 'little').hex()
g = {'rfc8032': [], 'mul': [], 'encrypt': [], 'aggregate': {}}
for sk, pk in [('9d61b19deffd5a60ba844af492ec2cc44449c5697b326919703bac031cae7f60',
                'd75a980182b10ab7d54bfed3c964073a0ee172f3daa62325af021a68f707511a'),
               ('4ccd089b28ff96da9db6c346ec114e0f5b8a319f35aba624da8cf6ed4fb8a6fb',
                '3d4017c3e843895a92b70aa74d1b7ebc9c982ccf2ec4968cc0cd55f12af4660c'),
               ('c5aa8df43f9f837bedb7442f31dcb7b166d38535076f094b85ce3a2e0b4458f7',
                'fc51cd8e6218a1a38da47ed00230f0580816ed13ba3303ac5deb911548908025')]:
    h = hashlib.sha512(bytes.fromhex(sk)).digest()
    a = int.from_bytes(h[:32], 'little')
    a &= (1 << 254) - 8
    a |= 1 << 254
    assert E.to_bytes(E.mul_raw(a, E.BASE)).hex() == pk
    g['rfc8032'].append({'secret': sk, 'scalar_mod_l': sc(a % E.l), 'public': pk})
for k in [1, 2, E.l - 1, rng.scalar(), rng.scalar() >> 4]:
    P = E.mul(rng.scalar(), E.BASE)
    g['mul'].append({'k': sc(k), 'p': E.to_bytes(P).hex(), 'out': E.to_bytes(E.mul(k, P)).hex()})
sk = rng.scalar() % E.l
Pp = E.mul(sk, E.BASE)
cts = []
for m in [0, 1, -1, 42, -11955, 99999, 2 ** 40 + 3, -(2 ** 63)]:
    r = rng.scalar()
    ct = E.encrypt(Pp, m, r % E.l)
    cts.append(ct)
    g['encrypt'].append({'P': E.to_bytes(Pp).hex(), 'sk': sc(sk), 'm': m, 'r': sc(r), 'ct': E.ct_to_bytes(ct).hex()})
s = cts[0]
for ct in cts[1:6]:
    s = E.ct_add(s, ct)
g['aggregate'] = {'inputs': [E.ct_to_bytes(c).hex() for c in cts[:6]], 'out': E.ct_to_bytes(s).hex(),
                  'plain_sum': sum([0, 1, -1, 42, -11955, 99999])}
path = os.path.join(ROOT, 'tests', 'golden', 'ed25519_golden.json')
with open(path, 'w') as f:
    json.dump(g, f, indent=1)
print('wrote', path)
