"""The two proof.Rep list proofs on the path, written once over an `engine` that offers rep_commit / rep_respond /
rep_verify -- the CPU oracle or the CUDA backend -- so the tests can run the same proof through both and compare bytes.

Obfuscation proof (lib/obfuscation/obfuscation_proof.go:37-91): And(Rep(co1 = s*c1), Rep(co2 = s*c2)); proof bytes in
kyber's order = V1 || V2 || r_s (160 bytes). Key-switch proof (libunlynxkeyswitch, unlynx v1.4.0; verified from
lib/proof/structs_proofs.go:469-482): And(Rep(viB = vi*B), Rep(K = k*B), Rep(ks2 = vi*Q + k*rBNeg)); proof bytes =
V1 || V2 || V3 || r_vi || r_k (256 bytes).

The challenge comes from kyber's hashProver (the suite's XOF seeded with the protocol name, reseeded with the
commitments). That XOF is not reproducible here, so `challenge` below is a stand-in (SHA3-512 of name || commitments,
read big-endian mod n) -- the engines never see the difference: the challenge is an input of the C ABI."""
import hashlib

import numpy as np

import bn256_ref as R


def challenge(name, commitments):
    return (int.from_bytes(hashlib.sha3_512(name + bytes(commitments)).digest(), 'big') % R.n).to_bytes(32, 'big')


class OracleEngine:
    def __init__(self, oracle, threads=8):
        self.o, self.t = oracle, threads

    def rep_commit(self, nterms, scalars, points):
        rc, out = self.o.rep_commit(nterms, scalars, points, threads=self.t)
        if rc != 0:
            raise RuntimeError('malformed point')
        return out

    def rep_respond(self, per, v, chal, x):
        return self.o.rep_respond(per, v, chal, x)

    def rep_verify(self, nterms, resp, chal, bases, P, V):
        return np.asarray(self.o.rep_verify(nterms, resp, chal, bases, P, V, threads=self.t))


class BackendEngine:
    def __init__(self, backend):
        self.b = backend

    def rep_commit(self, nterms, scalars, points):
        return self.b.rep_commit(nterms, scalars, points).tobytes()

    def rep_respond(self, per, v, chal, x):
        return self.b.rep_respond(per, v, chal, x).tobytes()

    def rep_verify(self, nterms, resp, chal, bases, P, V):
        return np.asarray(self.b.rep_verify(nterms, resp, chal, bases, P, V))


def _chunks(b, size):
    b = bytes(b)
    return [b[i:i + size] for i in range(0, len(b), size)]


# ---------------------------------------------------------------- obfuscation
def obfuscation_create(engine, cts, cos, ss, vs, name=b'proofTest'):
    """ObfuscationListProofCreation: cts, cos n*128 bytes, ss (secrets) and vs (blinding values) n*32 bytes ->
    list of n 160-byte proofs."""
    n = len(cts) // 128
    v2 = b''.join(v + v for v in _chunks(vs, 32))          # both relations of a proof share the blinding value
    V = engine.rep_commit(1, v2, cts)                      # a ciphertext is already (c1 || c2)
    chal = b''.join(challenge(name, V[128 * i:128 * i + 128]) for i in range(n))
    r = engine.rep_respond(1, vs, chal, ss)
    return [V[128 * i:128 * i + 128] + r[32 * i:32 * i + 32] for i in range(n)]


def obfuscation_verify(engine, cts, cos, proofs, name=b'proofTest'):
    """ObfuscationProofVerification per proof -> n accept bits."""
    n = len(proofs)
    ok = np.ones(n, dtype=bool)
    V, resp, chal = [], [], []
    for i, pr in enumerate(proofs):
        if len(pr) != 160:
            ok[i] = False
            pr = bytes(160)
        V.append(pr[:128])
        resp.append(pr[128:160] * 2)
        chal.append(challenge(name, pr[:128]) * 2)
    acc = engine.rep_verify(1, b''.join(resp), b''.join(chal), cts, cos, b''.join(V)).reshape(n, 2)
    return ok & (acc[:, 0] == 1) & (acc[:, 1] == 1)


def list_verdict(bits, percent):
    """The *ListProofVerification rule: the first ceil(percent * n) proofs must all hold."""
    import math
    k = int(math.ceil(percent * len(bits)))
    return all(bool(b) for b in bits[:k])


# ---------------------------------------------------------------- key switching
def keyswitch_create(engine, B, Q, k, vis, rbnegs, blind, name=b'proofTest'):
    """KeySwitchListProofCreation(pubKey = k*B, targetPubKey = Q, secretKey = k, ks2s, rBNegs, vis): vis n*32, rbnegs
    n*64, blind n*64 (the blinding values of vi and of k for proof i). Returns (publics, proofs): publics[i] =
    viB || K || ks2 (192 bytes), proofs[i] 256 bytes."""
    n = len(vis) // 32
    kb = bytes(k)
    x = b''.join(vi + kb for vi in _chunks(vis, 32))                    # secrets in variable order (vi, k)
    pub1 = engine.rep_commit(1, x, (B + B) * n)                         # viB, K
    pts2 = b''.join(Q + rb for rb in _chunks(rbnegs, 64))
    pub2 = engine.rep_commit(2, x, pts2)                                # ks2
    V12 = engine.rep_commit(1, blind, (B + B) * n)
    V3 = engine.rep_commit(2, blind, pts2)
    Vs = [V12[128 * i:128 * i + 128] + V3[64 * i:64 * i + 64] for i in range(n)]
    chal = b''.join(challenge(name, v) for v in Vs)
    r = engine.rep_respond(2, blind, chal, x)
    publics = [pub1[128 * i:128 * i + 128] + pub2[64 * i:64 * i + 64] for i in range(n)]
    return publics, [Vs[i] + r[64 * i:64 * i + 64] for i in range(n)]


def keyswitch_verify(engine, B, Q, publics, rbnegs, proofs, name=b'proofTest'):
    n = len(proofs)
    ok = np.ones(n, dtype=bool)
    r1, c1, P1, V1, r2, c2, P2, V2 = [], [], [], [], [], [], [], []
    for i, pr in enumerate(proofs):
        if len(pr) != 256:
            ok[i] = False
            pr = bytes(256)
        c = challenge(name, pr[:192])
        r1.append(pr[192:256]); c1.append(c + c); P1.append(publics[i][:128]); V1.append(pr[:128])
        r2.append(pr[192:256]); c2.append(c); P2.append(publics[i][128:192]); V2.append(pr[128:192])
    a1 = engine.rep_verify(1, b''.join(r1), b''.join(c1), (B + B) * n, b''.join(P1), b''.join(V1)).reshape(n, 2)
    pts2 = b''.join(Q + rb for rb in _chunks(rbnegs, 64))
    a2 = engine.rep_verify(2, b''.join(r2), b''.join(c2), pts2, b''.join(P2), b''.join(V2))
    return ok & (a1[:, 0] == 1) & (a1[:, 1] == 1) & (a2 == 1)
