"""TEST INFRASTRUCTURE ONLY -- textbook big-integer model of ElGamal over edwards25519.

UnLynx's default suite (`libunlynx.SuiTe = suites.MustFind("Ed25519")`, github.com/ldsec/unlynx v1.4.0
lib/crypto.go, not under /root/reference: go.mod:26) is kyber v3.0.8 `group/edwards25519`; BASELINE.json's north_star
names this curve. The same ElGamal entry points as the BN256 path (row a-1 encrypt, a-3 aggregate, a-4/a-5 decrypt)
are restated here over it.

*** parity PINNED for the group arithmetic *** : edwards25519 is a published standard with known answers. This file
checks itself on import against RFC 8032 section 7.1 (base point encoding, TEST 1 and TEST 2 public keys) and the
group-order identity l*B = O; the C++ restatement (oracle/oracle_ed.cpp) and the CUDA kernels are checked against it.
What stays recalled rather than pinned is kyber's *encoding of scalars* (32 bytes little-endian, reduced mod l) and
unlynx's ciphertext layout K || C (32 + 32 bytes).

Only tests/, __graft_entry__.smoke() and bench.py's cpu_baseline leg may import anything under oracle/.

Conventions (kyber group/edwards25519, after the ref10 code it was derived from):
  field p = 2^255 - 19; curve -x^2 + y^2 = 1 + d x^2 y^2, d = -121665/121666
  point wire form: 32 bytes, y little-endian, bit 255 = x mod 2 (RFC 8032 5.1.2). Decoding ignores bit 255 of y's
    bytes beyond the sign, accepts y >= p (reduced) and "negative zero" x, as ref10 ge_frombytes does.
  scalar wire form: 32 bytes little-endian, reduced mod l on input
  identity = (0, 1) -> bytes 01 00 ... 00
"""
import hashlib

p = 2 ** 255 - 19
l = 2 ** 252 + 27742317777372353535851937790883648493
d = (-121665 * pow(121666, -1, p)) % p
SQRT_M1 = pow(2, (p - 1) // 4, p)
By = 4 * pow(5, -1, p) % p


def _recover_x(y, sign):
    """x with x^2 = (y^2 - 1) / (d y^2 + 1) and x mod 2 == sign, or None (ref10 ge_frombytes semantics)."""
    y %= p
    u, v = (y * y - 1) % p, (d * y * y + 1) % p
    x = (u * pow(v, 3, p) * pow(u * pow(v, 7, p) % p, (p - 5) // 8, p)) % p
    vxx = v * x * x % p
    if vxx != u:
        if vxx != (-u) % p:
            return None
        x = x * SQRT_M1 % p
    if (x & 1) != sign:
        x = (-x) % p   # x = 0 stays 0 whatever the sign bit says
    return x


Bx = _recover_x(By, 0)
BASE = (Bx, By)
IDENT = (0, 1)


def on_curve(P):
    x, y = P
    return (-x * x + y * y - 1 - d * x * x * y * y) % p == 0


def add(P, Q):
    """Complete affine addition law on the twisted Edwards curve (a = -1)."""
    x1, y1 = P
    x2, y2 = Q
    t = d * x1 * x2 * y1 * y2 % p
    x3 = (x1 * y2 + x2 * y1) * pow(1 + t, -1, p) % p
    y3 = (y1 * y2 + x1 * x2) * pow(1 - t, -1, p) % p
    return (x3, y3)


def neg(P):
    return ((-P[0]) % p, P[1])


def mul(k, P):
    k %= l
    R, Q = IDENT, P
    while k:
        if k & 1:
            R = add(R, Q)
        Q = add(Q, Q)
        k >>= 1
    return R


def mul_raw(k, P):
    """Scalar multiplication by an integer that is NOT reduced mod l (RFC 8032 clamped secrets, cofactor tests)."""
    R, Q = IDENT, P
    while k:
        if k & 1:
            R = add(R, Q)
        Q = add(Q, Q)
        k >>= 1
    return R


def to_bytes(P):
    x, y = P
    return (y | ((x & 1) << 255)).to_bytes(32, 'little')


def from_bytes(b):
    """Point or None; mirrors kyber Point.UnmarshalBinary -> extendedGroupElement.FromBytes."""
    assert len(b) == 32
    v = int.from_bytes(b, 'little')
    sign = v >> 255
    y = (v & ((1 << 255) - 1)) % p
    x = _recover_x(y, sign)
    if x is None:
        return None
    return (x, y)


# ------------------------------------------------------------------ scalars
def scalar_from_int64(v):
    """kyber Scalar.SetInt64: negative values wrap mod l."""
    return v % l


def scalar_from_bytes(b):
    return int.from_bytes(b, 'little') % l


def scalar_to_bytes(s):
    return (s % l).to_bytes(32, 'little')


# ------------------------------------------------------------------ ElGamal (unlynx lib/crypto.go encryptPoint etc.)
def encrypt(pub, m, r):
    """(K, C) = (r*B, m*B + r*P); m an int64, r a scalar."""
    K = mul(r, BASE)
    C = add(mul(scalar_from_int64(m), BASE), mul(r, pub))
    return K, C


def ct_to_bytes(ct):
    return to_bytes(ct[0]) + to_bytes(ct[1])


def ct_from_bytes(b):
    K, C = from_bytes(b[:32]), from_bytes(b[32:64])
    if K is None or C is None:
        return None
    return K, C


def ct_add(a, b):
    return add(a[0], b[0]), add(a[1], b[1])


def decrypt_point(sk, ct):
    return add(ct[1], neg(mul(sk, ct[0])))


# ------------------------------------------------------------------ known answers (RFC 8032 section 7.1)
def _rfc8032_public(secret_hex):
    h = hashlib.sha512(bytes.fromhex(secret_hex)).digest()
    a = int.from_bytes(h[:32], 'little')
    a &= (1 << 254) - 8
    a |= 1 << 254
    return to_bytes(mul_raw(a, BASE)).hex()


def self_check():
    assert on_curve(BASE) and Bx == 15112221349535400772501151409588531511454012693041857206046113283949847762202
    assert to_bytes(BASE).hex() == '58' + '66' * 31
    assert _rfc8032_public('9d61b19deffd5a60ba844af492ec2cc44449c5697b326919703bac031cae7f60') == \
        'd75a980182b10ab7d54bfed3c964073a0ee172f3daa62325af021a68f707511a'
    assert _rfc8032_public('4ccd089b28ff96da9db6c346ec114e0f5b8a319f35aba624da8cf6ed4fb8a6fb') == \
        '3d4017c3e843895a92b70aa74d1b7ebc9c982ccf2ec4968cc0cd55f12af4660c'
    assert mul_raw(l, BASE) == IDENT and mul_raw(l - 1, BASE) == neg(BASE)
    assert from_bytes(to_bytes(BASE)) == BASE and to_bytes(IDENT) == b'\x01' + bytes(31)


self_check()
