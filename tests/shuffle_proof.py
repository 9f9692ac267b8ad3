"""Neff's verifiable shuffle of ElGamal pairs as unlynx's shuffle proof runs it (libunlynxshuffle.ShuffleProofCreation /
ShuffleProofVerification, unlynx v1.4.0, called from lib/proof/structs_proofs.go:389-420 and the shuffling protocol):
the lines are compressed to one ciphertext each (tests/test_rep_proofs.py::test_gpu_line_compression_of_the_shuffle_proof),
then kyber's shuffle.PairShuffle proves that (Xbar_i, Ybar_i) = (X_pi(i) + beta_pi(i) g, Y_pi(i) + beta_pi(i) h).

kyber v3.0.8 shuffle/pair.go and shuffle/simple.go are not under /root/reference: the message flow below (ega1..ega6,
the embedded SimpleShuffle with its Theta / alpha vectors, checks (31)-(35) of the verifier) is RESTATED FROM MEMORY and
pinned by nothing but its own algebra (an honest proof verifies, damaged ones do not). As in rep_proofs.py the
Fiat-Shamir challenges are a stand-in for kyber's XOF transcript; scalar arithmetic mod n stays on the host (kyber
scalars in Go), and EVERY group operation goes through the engine's rep_commit / rep_verify, i.e. through
drx_rep_commit / drx_rep_verify on the GPU: what a Go shim over the C ABI would call, with no new entry point.

Group work per proof of k pairs: prover 7k + 3 fixed-base products and two sums of k + 1 terms; verifier 2k two-term
sums, 3k one-term relations and two sums of 2k + 1 terms. One thread walks a relation's terms, so the wide sums are
taken up a tree of 16-term relations (`_sum`)."""
import hashlib

import numpy as np

import bn256_ref as R

N = R.n


def sb(v):
    return (v % N).to_bytes(32, 'big')


class Transcript:
    """Stand-in for proof.HashProve / HashVerify: everything Put so far is hashed for each public random value."""

    def __init__(self, name):
        self.h = hashlib.sha3_512(name)
        self.count = 0

    def put(self, data):
        self.h.update(bytes(data))

    def pub_rand(self, k=1):
        out = []
        for _ in range(k):
            h = self.h.copy()
            h.update(self.count.to_bytes(4, 'big'))
            self.count += 1
            out.append(int.from_bytes(h.digest(), 'big') % N)
        return out


def _mul_base(engine, scalars, base):
    """[s * base for s in scalars] as one batch of one-term relations."""
    if not scalars:
        return []
    out = engine.rep_commit(1, b''.join(sb(s) for s in scalars), bytes(base) * len(scalars))
    return [out[64 * i:64 * i + 64] for i in range(len(scalars))]


CHUNK = 16


def _sum(engine, scalars, points, pad_point):
    """sum_j scalars[j] * points[j]. A relation's terms are walked by one thread, so a wide sum goes up a tree: CHUNK
    terms per relation (all relations of a level in one call), then the partial sums with scalar 1, until one is left.
    `pad_point` (any valid point) fills the last chunk with zero terms."""
    scalars, points = [sb(s) for s in scalars], list(points)
    one, zero = sb(1), sb(0)
    while len(scalars) > CHUNK:
        pad = (-len(scalars)) % CHUNK
        scalars += [zero] * pad
        points += [pad_point] * pad
        part = engine.rep_commit(CHUNK, b''.join(scalars), b''.join(points))
        points = [part[64 * i:64 * i + 64] for i in range(len(scalars) // CHUNK)]
        scalars = [one] * len(points)
    return engine.rep_commit(len(scalars), b''.join(scalars), b''.join(points))


def shuffle_pairs(engine, pi, g, h, beta, X, Y):
    """The shuffle itself: (Xbar_i, Ybar_i) = (X[pi[i]] + beta[pi[i]] g, Y[pi[i]] + beta[pi[i]] h)."""
    k = len(pi)
    one = sb(1)
    sc = b''.join(one + sb(beta[pi[i]]) for i in range(k))
    Xbar = engine.rep_commit(2, sc, b''.join(X[pi[i]] + g for i in range(k)))
    Ybar = engine.rep_commit(2, sc, b''.join(Y[pi[i]] + h for i in range(k)))
    return [Xbar[64 * i:64 * i + 64] for i in range(k)], [Ybar[64 * i:64 * i + 64] for i in range(k)]


# ---------------------------------------------------------------- SimpleShuffle (shuffle/simple.go)
def _simple_prove(engine, tr, G, gamma, x, y, theta):
    k = len(x)
    thlen = 2 * k - 1
    XY = _mul_base(engine, list(x) + list(y), G)                      # step 0: X_i = x_i G, Y_i = y_i G
    p0 = b''.join(XY)
    tr.put(p0)
    t = tr.pub_rand()[0]                                              # V step 1
    gt = gamma * t % N
    xh = [(v - t) % N for v in x]
    yh = [(v - gt) % N for v in y]
    th = []                                                           # P step 2: Theta = thenc(...) * G
    th.append(-theta[0] * yh[0])
    for i in range(1, k):
        th.append(theta[i - 1] * xh[i] - theta[i] * yh[i])
    for i in range(k, thlen):
        th.append(theta[i - 1] * gamma - theta[i])
    th.append(theta[thlen - 1] * gamma)
    p2 = b''.join(_mul_base(engine, th, G))
    tr.put(p2)
    c = tr.pub_rand()[0]                                              # V step 3
    alpha = [0] * thlen                                               # P step 4
    run = c
    for i in range(k):
        run = run * xh[i] % N * pow(yh[i], -1, N) % N
        alpha[i] = (theta[i] + run) % N
    ginv = pow(gamma, -1, N)
    run = c
    for i in range(1, k):
        run = run * ginv % N
        alpha[thlen - i] = (theta[thlen - i] + run) % N
    p4 = b''.join(sb(a) for a in alpha)
    tr.put(p4)
    return p0 + p2 + p4


def _simple_verify(engine, tr, G, Gamma, k, blob):
    """blob = X[k] Y[k] | Theta[2k] | alpha[2k-1]. Returns (ok, X list, Y list)."""
    thlen = 2 * k - 1
    p0, p2, p4 = blob[:128 * k], blob[128 * k:128 * k + 64 * 2 * k], blob[128 * k + 128 * k:]
    X = [p0[64 * i:64 * i + 64] for i in range(k)]
    Y = [p0[64 * (k + i):64 * (k + i) + 64] for i in range(k)]
    Theta = [p2[64 * i:64 * i + 64] for i in range(2 * k)]
    alpha = [int.from_bytes(p4[32 * i:32 * i + 32], 'big') for i in range(thlen)]
    tr.put(p0)
    t = tr.pub_rand()[0]
    tr.put(p2)
    c = tr.pub_rand()[0]
    tr.put(p4)
    # V step 5: Xhat_i = X_i - t G, Yhat_i = Y_i - t Gamma
    sc = (sb(1) + sb(-t)) * k
    Xh = engine.rep_commit(2, sc, b''.join(X[i] + G for i in range(k)))
    Yh = engine.rep_commit(2, sc, b''.join(Y[i] + Gamma for i in range(k)))
    # thver(A, B, T, a, b): a A - b B == T, 2k of them in one call (resp = a, bases = A, chal = -b, P = B, V = T)
    resp, chal, A, B, T = [], [], [], [], []

    def thver(Apt, Bpt, Tpt, a, b):
        resp.append(sb(a)); chal.append(sb(-b)); A.append(Apt); B.append(Bpt); T.append(Tpt)
    thver(Xh[:64], Yh[:64], Theta[0], c, alpha[0])
    for i in range(1, k):
        thver(Xh[64 * i:64 * i + 64], Yh[64 * i:64 * i + 64], Theta[i], alpha[i - 1], alpha[i])
    for i in range(k, thlen):
        thver(Gamma, G, Theta[i], alpha[i - 1], alpha[i])
    thver(Gamma, G, Theta[thlen], alpha[thlen - 1], c)
    ok = engine.rep_verify(1, b''.join(resp), b''.join(chal), b''.join(A), b''.join(B), b''.join(T))
    return bool(np.all(np.asarray(ok) == 1)), X, Y


# ---------------------------------------------------------------- PairShuffle (shuffle/pair.go)
def pair_shuffle_prove(engine, pi, g, h, beta, X, Y, secrets, name=b'PairShuffle'):
    """secrets = dict(u, w, a: k scalars each; tau0, nu, gamma; theta: 2k - 1 scalars) -- what ctx.PriRand draws.
    Returns the proof bytes in the order the messages are Put."""
    k = len(pi)
    assert k > 1 and len(X) == len(Y) == len(beta) == k
    piinv = [0] * k
    for i in range(k):
        piinv[pi[i]] = i
    u, w, a, tau0, gamma = secrets['u'], secrets['w'], secrets['a'], secrets['tau0'], secrets['gamma']
    tr = Transcript(name)
    # P step 1
    base = _mul_base(engine, [gamma] + list(a) + [gamma * a[pi[i]] for i in range(k)] + list(u) + [gamma * w[i] for i in range(k)], g)
    Gamma, A, C, U, W = base[0], base[1:1 + k], base[1 + k:1 + 2 * k], base[1 + 2 * k:1 + 3 * k], base[1 + 3 * k:1 + 4 * k]
    wbetasum = (tau0 + sum(w[i] * beta[pi[i]] for i in range(k))) % N
    coef = [(w[piinv[i]] - u[i]) % N for i in range(k)]
    Lambda1 = _sum(engine, coef + [wbetasum], list(X) + [g], g)
    Lambda2 = _sum(engine, coef + [wbetasum], list(Y) + [h], g)
    p1 = Gamma + b''.join(A) + b''.join(C) + b''.join(U) + b''.join(W) + Lambda1 + Lambda2
    tr.put(p1)
    rho = tr.pub_rand(k)                                              # V step 2
    b = [(rho[i] - u[i]) % N for i in range(k)]                       # P step 3
    p3 = b''.join(_mul_base(engine, [gamma * b[pi[i]] for i in range(k)], g))
    tr.put(p3)
    lam = tr.pub_rand()[0]                                            # V step 4
    r = [(a[i] + lam * b[i]) % N for i in range(k)]                   # P step 5
    s = [gamma * r[pi[i]] % N for i in range(k)]
    sigma = [(w[i] + b[pi[i]]) % N for i in range(k)]
    tau = (-tau0 + sum(b[i] * beta[i] for i in range(k))) % N
    p5 = b''.join(sb(v) for v in sigma) + sb(tau)
    tr.put(p5)
    p6 = _simple_prove(engine, tr, g, gamma, r, s, secrets['theta'])  # P, V step 6
    return p1 + p3 + p5 + p6


def proof_size(k):
    return 64 * (4 * k + 3) + 64 * k + 32 * (k + 1) + 128 * k + 128 * k + 32 * (2 * k - 1)


def pair_shuffle_verify(engine, g, h, X, Y, Xbar, Ybar, proof, name=b'PairShuffle', link_simple_shuffle=False):
    """shuffle.Verifier: True / False. link_simple_shuffle additionally checks that the embedded simple shuffle is
    about R_i = A_i + lambda B_i and S_i = C_i + lambda D_i (Neff's step; kyber's Verify, as recalled, leaves it out)."""
    k = len(X)
    if k < 2 or len(proof) != proof_size(k):
        return False
    pos = [0]

    def take(nbytes):
        out = proof[pos[0]:pos[0] + nbytes]
        pos[0] += nbytes
        return out
    p1 = take(64 * (4 * k + 3))
    Gamma = p1[:64]
    A, C, U, W = ([p1[64 * (1 + j * k + i):64 * (2 + j * k + i)] for i in range(k)] for j in range(4))
    Lambda1, Lambda2 = p1[64 * (4 * k + 1):64 * (4 * k + 2)], p1[64 * (4 * k + 2):]
    tr = Transcript(name)
    tr.put(p1)
    rho = tr.pub_rand(k)
    p3 = take(64 * k)
    D = [p3[64 * i:64 * i + 64] for i in range(k)]
    tr.put(p3)
    lam = tr.pub_rand()[0]
    p5 = take(32 * (k + 1))
    sigma = [int.from_bytes(p5[32 * i:32 * i + 32], 'big') for i in range(k)]
    tau = int.from_bytes(p5[32 * k:], 'big')
    tr.put(p5)
    try:
        ok6, Rs, Ss = _simple_verify(engine, tr, g, Gamma, k, take(len(proof) - pos[0]))
        if not ok6:
            return False
        # (33): sigma_i Gamma == W_i + D_i
        ok33 = engine.rep_verify(1, b''.join(sb(v) for v in sigma), sb(-1) * k, Gamma * k, b''.join(D), b''.join(W))
        # (31)-(35): Lambda == sum sigma_i bar_i - sum rho_i orig_i - tau base, the sums of 2k + 1 terms up a tree
        coef = list(sigma) + [-v for v in rho] + [-tau]
        Phi1 = _sum(engine, coef, list(Xbar) + list(X) + [g], g)
        Phi2 = _sum(engine, coef, list(Ybar) + list(Y) + [h], g)
        ok = bool(np.all(np.asarray(ok33) == 1)) and Phi1 == Lambda1 and Phi2 == Lambda2
        if ok and link_simple_shuffle:
            # R_i = A_i + lambda B_i with B_i = rho_i g - U_i; S_i = C_i + lambda D_i
            sc = b''.join(sb(1) + sb(lam * rho[i]) + sb(-lam) for i in range(k))
            Rv = engine.rep_commit(3, sc, b''.join(A[i] + g + U[i] for i in range(k)))
            Sv = engine.rep_commit(2, (sb(1) + sb(lam)) * k, b''.join(C[i] + D[i] for i in range(k)))
            ok = Rv == b''.join(Rs) and Sv == b''.join(Ss)
        return ok
    except RuntimeError:    # a point of the proof does not unmarshal
        return False
