"""TEST INFRASTRUCTURE ONLY — textbook big-integer model of the BN256 arithmetic under Drynx.

*** parity unpinned *** : the reference (`/root/reference`, Go) cannot be built here (no Go toolchain, the
arithmetic lives in the un-vendored modules go.dedis.ch/kyber/v3 v3.0.8 `pairing/bn256` and
github.com/ldsec/unlynx v1.4.0 `lib/crypto.go`, go.mod:14,26) and holds no golden vectors.
This file states the *mathematical definition* of every value on the path with Python integers, written
independently of both the C++ oracle (oracle/oracle.cpp, which restates kyber's algorithms) and the CUDA
kernels, so that an arithmetic slip in either shows up as a mismatch. Only tests/, __graft_entry__.smoke()
and bench.py's cpu_baseline leg may import anything under oracle/.

Conventions (kyber pairing/bn256, recalled):
  Fp2  = Fp[i]/(i^2+1)           kyber gfP2{x,y}  = x*i + y          here: (c0, c1) = c0 + c1*i
  Fp6  = Fp2[tau]/(tau^3 - xi)   xi = i + 3;  gfP6{x,y,z} = x*tau^2 + y*tau + z
  Fp12 = Fp6[w]/(w^2 - tau)      gfP12{x,y} = x*w + y                here: 6 Fp2 coefficients of w^0..w^5
  G1: y^2 = x^3 + 3, generator (1, -2) [kyber curveGen]; wire = x||y, 32-byte big-endian each, infinity = zeros
  G2: y^2 = x^3 + 3/xi over Fp2, wire = x.x||x.y||y.x||y.y (i-coefficient first)
  GT: wire = 12 x 32 B in order x.x.x x.x.y x.y.x x.y.y x.z.x x.z.y y.x.x ... y.z.y
  e(P, Q) = (f_{6u+2,Q}(P) * l_{[6u+2]Q, pi(Q)}(P) * l_{[6u+2]Q+pi(Q), -pi^2(Q)}(P)) ^ ((p^12-1)/n)
"""
import hashlib

u = 6518589491078791937
p = 36 * u**4 + 36 * u**3 + 24 * u**2 + 6 * u + 1
n = 36 * u**4 + 36 * u**3 + 18 * u**2 + 6 * u + 1
assert p == 65000549695646603732796438742359905742825358107623003571877145026864184071783
assert n == 65000549695646603732796438742359905742570406053903786389881062969044166799969
ATE_LOOP = 6 * u + 2

# ------------------------------------------------------------------ Fp2
XI = (3, 1)  # 3 + i


def f2add(a, b): return ((a[0] + b[0]) % p, (a[1] + b[1]) % p)
def f2sub(a, b): return ((a[0] - b[0]) % p, (a[1] - b[1]) % p)
def f2neg(a): return ((-a[0]) % p, (-a[1]) % p)
def f2mul(a, b): return ((a[0] * b[0] - a[1] * b[1]) % p, (a[0] * b[1] + a[1] * b[0]) % p)
def f2scal(a, k): return ((a[0] * k) % p, (a[1] * k) % p)
def f2conj(a): return (a[0], (-a[1]) % p)


def f2inv(a):
    d = pow(a[0] * a[0] + a[1] * a[1], -1, p)
    return ((a[0] * d) % p, (-a[1] * d) % p)


def f2pow(a, e):
    r = (1, 0)
    while e:
        if e & 1:
            r = f2mul(r, a)
        a = f2mul(a, a)
        e >>= 1
    return r


F2_ZERO, F2_ONE = (0, 0), (1, 0)

# ------------------------------------------------------------------ Fp12 as Fp2[w]/(w^6 - xi)
F12_ONE = [F2_ONE] + [F2_ZERO] * 5


def f12mul(a, b):
    t = [F2_ZERO] * 11
    for i in range(6):
        if a[i] == F2_ZERO:
            continue
        for j in range(6):
            if b[j] == F2_ZERO:
                continue
            t[i + j] = f2add(t[i + j], f2mul(a[i], b[j]))
    r = list(t[:6])
    for k in range(6, 11):
        r[k - 6] = f2add(r[k - 6], f2mul(t[k], XI))
    return r


def f12pow(a, e):
    r = list(F12_ONE)
    for bit in bin(e)[2:]:
        r = f12mul(r, r)
        if bit == '1':
            r = f12mul(r, a)
    return r


def f12_to_bytes(a):
    """kyber GT MarshalBinary order: x.x x.y x.z y.x y.y y.z, each Fp2 as (i-coefficient, constant)."""
    # w-power -> tower slot: w^5=x.x w^3=x.y w^1=x.z w^4=y.x w^2=y.y w^0=y.z
    out = b''
    for k in (5, 3, 1, 4, 2, 0):
        out += a[k][1].to_bytes(32, 'big') + a[k][0].to_bytes(32, 'big')
    return out


def f12_from_bytes(b):
    a = [None] * 6
    for idx, k in enumerate((5, 3, 1, 4, 2, 0)):
        c1 = int.from_bytes(b[64 * idx:64 * idx + 32], 'big')
        c0 = int.from_bytes(b[64 * idx + 32:64 * idx + 64], 'big')
        a[k] = (c0, c1)
    return a


# ------------------------------------------------------------------ G1 (affine, None = infinity)
G1_GEN = (1, p - 2)
B1 = 3


def g1_on_curve(P):
    return P is None or (P[1] * P[1] - P[0] ** 3 - B1) % p == 0


def g1_add(P, Q):
    if P is None: return Q
    if Q is None: return P
    if P[0] == Q[0]:
        if (P[1] + Q[1]) % p == 0:
            return None
        lam = 3 * P[0] * P[0] * pow(2 * P[1], -1, p) % p
    else:
        lam = (Q[1] - P[1]) * pow(Q[0] - P[0], -1, p) % p
    x = (lam * lam - P[0] - Q[0]) % p
    return (x, (lam * (P[0] - x) - P[1]) % p)


def g1_neg(P): return None if P is None else (P[0], (-P[1]) % p)


def g1_mul(k, P):
    k %= n
    R = None
    while k:
        if k & 1:
            R = g1_add(R, P)
        P = g1_add(P, P)
        k >>= 1
    return R


def g1_to_bytes(P):
    if P is None:
        return bytes(64)
    return P[0].to_bytes(32, 'big') + P[1].to_bytes(32, 'big')


def g1_from_bytes(b):
    x, y = int.from_bytes(b[:32], 'big'), int.from_bytes(b[32:64], 'big')
    if x == 0 and y == 0:
        return None
    P = (x, y)
    if x >= p or y >= p or not g1_on_curve(P):
        raise ValueError("bn256.G1: malformed point")
    return P


# ------------------------------------------------------------------ G2 (affine over Fp2)
B2 = f2mul((3, 0), f2inv(XI))  # 3/xi
# x/crypto bn256 twist generator (same point kyber's twistGen encodes in Montgomery form; on-curve and
# order-n checked by tests/test_oracle_ref.py). Fp2 here is (constant, i-coefficient).
G2_GEN = (
    (64746500191241794695844075326670126197795977525365406531717464316923369116492,
     21167961636542580255011770066570541300993051739349375019639421053990175267184),
    (17778617556404439934652658462602675281523610326338642107814333856843981424549,
     20666913350058776956210519119118544732556678129809273996262322366050359951122),
)


def g2_on_curve(Q):
    if Q is None: return True
    x, y = Q
    return f2sub(f2mul(y, y), f2add(f2mul(f2mul(x, x), x), B2)) == F2_ZERO


def g2_add(P, Q):
    if P is None: return Q
    if Q is None: return P
    if P[0] == Q[0]:
        if f2add(P[1], Q[1]) == F2_ZERO:
            return None
        lam = f2mul(f2scal(f2mul(P[0], P[0]), 3), f2inv(f2scal(P[1], 2)))
    else:
        lam = f2mul(f2sub(Q[1], P[1]), f2inv(f2sub(Q[0], P[0])))
    x = f2sub(f2sub(f2mul(lam, lam), P[0]), Q[0])
    return (x, f2sub(f2mul(lam, f2sub(P[0], x)), P[1]))


def g2_neg(P): return None if P is None else (P[0], f2neg(P[1]))


def g2_mul(k, P):
    k %= n
    R = None
    while k:
        if k & 1:
            R = g2_add(R, P)
        P = g2_add(P, P)
        k >>= 1
    return R


def g2_mul_raw(k, P):
    """[k]P without reducing k mod n (P may lie outside the order-n subgroup)."""
    R = None
    while k:
        if k & 1:
            R = g2_add(R, P)
        P = g2_add(P, P)
        k >>= 1
    return R


def g2_in_subgroup(Q):
    """Membership in G2 = E'(Fp2)[n], by definition: [n]Q = O. kyber's twistPoint.IsOnCurve (pairing/bn256 twist.go,
    recalled; not in /root/reference) multiplies by Order after the curve equation and rejects otherwise, so
    UnmarshalBinary fails for a twist point outside the subgroup and lib/range/range_proof.go:225-236 aborts."""
    return g2_mul_raw(n, Q) is None


def f2sqrt(a):
    """A square root in Fp2 (p = 3 mod 4), or None."""
    if a == F2_ZERO:
        return a
    s = pow((a[0] * a[0] + a[1] * a[1]) % p, (p + 1) // 4, p)
    if (s * s - a[0] * a[0] - a[1] * a[1]) % p:
        return None
    for t in (s, -s):
        h = (a[0] + t) * pow(2, -1, p) % p
        x0 = pow(h, (p + 1) // 4, p)
        if (x0 * x0 - h) % p == 0 and x0:
            r = (x0, a[1] * pow(2 * x0, -1, p) % p)
            if f2mul(r, r) == a:
                return r
    return None


def g2_twist_point(seed):
    """A deterministic point on the twist curve E'(Fp2) (almost surely OUTSIDE the order-n subgroup: #E' = n (2p - n))."""
    k = 0
    while True:
        h = hashlib.sha256(b'twist point %d %d' % (seed, k)).digest()
        x = (int.from_bytes(h, 'big') % p, int.from_bytes(hashlib.sha256(h).digest(), 'big') % p)
        y = f2sqrt(f2add(f2mul(f2mul(x, x), x), B2))
        if y is not None:
            return (x, y)
        k += 1


TWIST_COFACTOR = 2 * p - n   # = 13 * 7369 * ...: points of small order exist on the twist


def g2_to_bytes(Q):
    if Q is None:
        return bytes(128)
    (x0, x1), (y0, y1) = Q
    return b''.join(v.to_bytes(32, 'big') for v in (x1, x0, y1, y0))


def g2_from_bytes(b):
    v = [int.from_bytes(b[32 * k:32 * k + 32], 'big') for k in range(4)]
    if not any(v):
        return None
    Q = ((v[1], v[0]), (v[3], v[2]))
    if max(v) >= p or not g2_on_curve(Q) or not g2_in_subgroup(Q):
        raise ValueError("bn256.G2: malformed point")
    return Q


# ------------------------------------------------------------------ optimal ate pairing (definition)
XI_P_1_3 = f2pow(XI, (p - 1) // 3)
XI_P_1_2 = f2pow(XI, (p - 1) // 2)


def _line(T, lam, P):
    """Line through untwisted T with twist-slope lam, evaluated at P in G1: yP - lam*xP*w + (lam*xT - yT)*w^3."""
    r = [F2_ZERO] * 6
    r[0] = (P[1], 0)
    r[1] = f2neg(f2scal(lam, P[0]))
    r[3] = f2sub(f2mul(lam, T[0]), T[1])
    return r


def _dbl_step(T, P):
    lam = f2mul(f2scal(f2mul(T[0], T[0]), 3), f2inv(f2scal(T[1], 2)))
    return g2_add(T, T), _line(T, lam, P)


def _add_step(T, Q, P):
    lam = f2mul(f2sub(Q[1], T[1]), f2inv(f2sub(Q[0], T[0])))
    return g2_add(T, Q), _line(T, lam, P)


def miller(P, Q):
    f = list(F12_ONE)
    T = Q
    for bit in bin(ATE_LOOP)[3:]:
        T, l = _dbl_step(T, P)
        f = f12mul(f12mul(f, f), l)
        if bit == '1':
            T, l = _add_step(T, Q, P)
            f = f12mul(f, l)
    Q1 = (f2mul(f2conj(Q[0]), XI_P_1_3), f2mul(f2conj(Q[1]), XI_P_1_2))
    Q2 = (f2mul(f2conj(Q1[0]), XI_P_1_3), f2mul(f2conj(Q1[1]), XI_P_1_2))
    T, l = _add_step(T, Q1, P)
    f = f12mul(f, l)
    T, l = _add_step(T, g2_neg(Q2), P)
    f = f12mul(f, l)
    return f


FINAL_EXP = (p ** 12 - 1) // n


def pair(P, Q):
    """e(P in G1, Q in G2) as 6 Fp2 coefficients of w^0..w^5."""
    if P is None or Q is None:
        return list(F12_ONE)
    return f12pow(miller(P, Q), FINAL_EXP)


# ------------------------------------------------------------------ scalars / hashing
def scalar_from_int64(v):
    return v % n


def scalar_to_bytes(s):
    return (s % n).to_bytes(32, 'big')


def sha3_512(data):
    return hashlib.sha3_512(data).digest()


# ------------------------------------------------------------------ ElGamal (unlynx lib/crypto.go, recalled)
def encrypt(pub, m, r):
    """K = r*B, C = m*B + r*P (EncryptIntGetR with the ephemeral scalar r injected)."""
    K = g1_mul(r, G1_GEN)
    C = g1_add(g1_mul(scalar_from_int64(m), G1_GEN), g1_mul(r, pub))
    return g1_to_bytes(K) + g1_to_bytes(C)


def encrypt_scalar(pub, m, r):
    """libunlynx.EncryptScalar: the message is a full scalar, C = m*B + r*P (lib/encoding/OR_AND.go:52,105)."""
    K = g1_mul(r, G1_GEN)
    C = g1_add(g1_mul(m % n, G1_GEN), g1_mul(r, pub))
    return g1_to_bytes(K) + g1_to_bytes(C)


def decrypt_point(sk, ct):
    K, C = g1_from_bytes(ct[:64]), g1_from_bytes(ct[64:128])
    return g1_add(C, g1_neg(g1_mul(sk, K)))


def ct_scalar_mul(ct, s):
    """CipherText.MulCipherTextbyScalar (protocols/obfuscation_protocol.go:240-243): (s*K, s*C)."""
    K, C = g1_from_bytes(ct[:64]), g1_from_bytes(ct[64:128])
    return g1_to_bytes(g1_mul(s % n, K)) + g1_to_bytes(g1_mul(s % n, C))


def rerandomize(ct, pub, b):
    """CipherText.Rerandomize as ShuffleSequence uses it (data/data.go:68): (K + b*B, C + b*P)."""
    K, C = g1_from_bytes(ct[:64]), g1_from_bytes(ct[64:128])
    return g1_to_bytes(g1_add(K, g1_mul(b, G1_GEN))) + g1_to_bytes(g1_add(C, g1_mul(b, pub)))


def schnorr_sign(x, k, msg):
    """kyber sign/schnorr Sign over G1 with the nonce k injected (recalled): R = k*B, h = SHA-512(R || X || msg) mod n
    (big-endian), s = k + x*h, signature R || s."""
    import hashlib
    X = g1_mul(x, G1_GEN)
    Rb = g1_to_bytes(g1_mul(k, G1_GEN))
    h = int.from_bytes(hashlib.sha512(Rb + g1_to_bytes(X) + msg).digest(), 'big') % n
    return Rb + ((k + x * h) % n).to_bytes(32, 'big')


def schnorr_verify(X, msg, sig):
    """kyber sign/schnorr Verify (lib/proof/structs_proofs.go:498-505): s*B == R + h*X; s >= n or a bad R rejects."""
    import hashlib
    if len(sig) != 96:
        return False
    try:
        R = g1_from_bytes(sig[:64])
    except Exception:
        return False
    s = int.from_bytes(sig[64:], 'big')
    if s >= n:
        return False
    h = int.from_bytes(hashlib.sha512(sig[:64] + g1_to_bytes(X) + msg).digest(), 'big') % n
    return g1_mul(s, G1_GEN) == g1_add(R, g1_mul(h, X))


def key_switch(target_pub, rb, sk, v):
    """libunlynxkeyswitch.KeySwitchSequence for one element (unlynx v1.4.0 lib/key_switch/key_switch.go, recalled;
    call sites data/data.go:80, services/service.go:565-616): contribution (v*B, v*Q - sk*rB) and -rB."""
    rb_neg = g1_neg(rb)
    ks2 = g1_add(g1_mul(sk, rb_neg), g1_mul(v, target_pub))
    return g1_to_bytes(g1_mul(v, G1_GEN)) + g1_to_bytes(ks2), g1_to_bytes(rb_neg)


# ------------------------------------------------------------------ range proof (lib/range/range_proof.go)
def to_base(v, b, l):
    """lib/range/range_proof.go:584-595."""
    d = []
    while v > 0:
        d.append(v % b)
        v //= b
    while len(d) < l:
        d.append(0)
    return d


def init_range_sig(x, ubase):
    """lib/range/range_proof.go:249-288: Y = x*B, A_k = (x+k)^-1 * g2."""
    Y = g1_mul(x, G1_GEN)
    A = [g2_mul(pow((x + k) % n, -1, n), G2_GEN) for k in range(ubase)]
    return Y, A


def rp_create(sigs, ubase, l, secret, r, ca_pub, ct, rnd):
    """lib/range/range_proof.go:320-407. sigs = [(Y_i, [A_i0..]), ...]; rnd = iterator of scalars in the
    reference's draw order (s_j, t_j, m_j, then v_ij for each server). Returns serialised proof bytes in the
    field order Commit|Challenge|Zr|D|Zphi|per server V_i, A_i, Zv_i."""
    rnd = iter(rnd)
    digits = to_base(secret, ubase, l)
    C = g1_from_bytes(ct[64:128])
    sumY = None
    for (Y, _) in sigs:
        sumY = g1_add(sumY, Y)
    c = int.from_bytes(sha3_512(g1_to_bytes(G1_GEN) + g1_to_bytes(C) + g1_to_bytes(sumY)), 'big') % n
    ns = len(sigs)
    D = None
    msum = 0
    Zphi, Zv, V, A = [], [[] for _ in range(ns)], [[] for _ in range(ns)], [[] for _ in range(ns)]
    gtB = pair(G1_GEN, G2_GEN)
    for j, phi in enumerate(digits):
        sj, tj, mj = next(rnd), next(rnd), next(rnd)
        msum = (msum + mj) % n
        D = g1_add(D, g1_mul(sj * (ubase ** j % n) % n, G1_GEN))
        D = g1_add(D, g1_mul(mj, ca_pub))
        Zphi.append((sj - c * phi) % n)
        for i, (_, Ai) in enumerate(sigs):
            v = next(rnd)
            Vij = g2_mul(v, Ai[phi])
            a = f12mul(pair(g1_mul((-sj) % n, G1_GEN), Vij), f12pow(gtB, tj))
            V[i].append(Vij)
            A[i].append(a)
            Zv[i].append((tj - c * v) % n)
    Zr = (msum - c * r) % n
    out = ct + scalar_to_bytes(c) + scalar_to_bytes(Zr) + g1_to_bytes(D)
    out += b''.join(scalar_to_bytes(z) for z in Zphi)
    for i in range(ns):
        out += b''.join(g2_to_bytes(q) for q in V[i])
        out += b''.join(f12_to_bytes(a) for a in A[i])
        out += b''.join(scalar_to_bytes(z) for z in Zv[i])
    return out


def rp_verify(proof, ubase, l, Ys, P):
    """lib/range/range_proof.go:504-565 on serialised bytes."""
    if ubase == 0 and l == 0:
        return True
    ns = len(Ys)
    per = 32 + 544 * ns
    if (len(proof) - 256) % per != 0:
        return False
    nd = (len(proof) - 256) // per
    if nd != l:
        return False
    try:
        C = g1_from_bytes(proof[64:128])
        c = int.from_bytes(proof[128:160], 'big') % n
        Zr = int.from_bytes(proof[160:192], 'big') % n
        D = g1_from_bytes(proof[192:256])
        off = 256
        Zphi = [int.from_bytes(proof[off + 32 * j:off + 32 * j + 32], 'big') % n for j in range(nd)]
        off += 32 * nd
        V, A, Zv = [], [], []
        for i in range(ns):
            V.append([g2_from_bytes(proof[off + 128 * j:off + 128 * j + 128]) for j in range(nd)])
            off += 128 * nd
            A.append([f12_from_bytes(proof[off + 384 * j:off + 384 * j + 384]) for j in range(nd)])
            off += 384 * nd
            Zv.append([int.from_bytes(proof[off + 32 * j:off + 32 * j + 32], 'big') % n for j in range(nd)])
            off += 32 * nd
    except ValueError:
        return False
    Dp = g1_add(g1_mul(c, C), g1_mul(Zr, P))
    ok = True
    for j in range(nd):
        Dp = g1_add(Dp, g1_mul(Zphi[j], g1_mul(ubase ** j % n, G1_GEN)))
        for i in range(ns):
            ap = pair(g1_mul(c, Ys[i]), V[i][j])
            ap = f12mul(ap, pair(g1_mul((-Zphi[j]) % n, G1_GEN), V[i][j]))
            ap = f12mul(ap, pair(g1_mul(Zv[i][j], G1_GEN), G2_GEN))
            if ap != A[i][j]:
                ok = False
    return ok and Dp == D
