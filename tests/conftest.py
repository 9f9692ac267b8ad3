import ctypes
import os
import sys

import numpy as np
import pytest

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
sys.path.insert(0, os.path.join(ROOT, 'oracle'))


def pytest_configure(config):
    config.addinivalue_line('markers', 'gpu: needs a B200 (run with -m gpu on the GPU box)')


def _have_gpu():
    try:
        import torch
        return torch.cuda.is_available()
    except Exception:
        return False


def pytest_collection_modifyitems(config, items):
    if _have_gpu():
        return
    skip = pytest.mark.skip(reason='no GPU in this container')
    for it in items:
        if 'gpu' in it.keywords:
            it.add_marker(skip)


class Oracle:
    """ctypes view of oracle/_build/liboracle.so (the CPU restatement; test infrastructure)."""

    def __init__(self):
        from drynx_b200 import build
        path = build.ORACLE_SO
        if not os.path.exists(path):
            path = build.build_oracle()
        self.lib = ctypes.CDLL(path)
        self.lib.orc_rp_proof_size.restype = ctypes.c_size_t
        self.lib.orc_rp_proof_size.argtypes = [ctypes.c_int, ctypes.c_int64]

    @staticmethod
    def _p(a):
        return a.ctypes.data_as(ctypes.c_void_p)

    def generators(self):
        g1, g2 = np.zeros(64, np.uint8), np.zeros(128, np.uint8)
        self.lib.orc_generators(self._p(g1), self._p(g2))
        return g1.tobytes(), g2.tobytes()

    def g1_mul(self, k, pt):
        out = np.zeros(64, np.uint8)
        assert self.lib.orc_g1_mul(bytes(k), bytes(pt), self._p(out)) == 0
        return out.tobytes()

    def g2_mul(self, k, pt):
        out = np.zeros(128, np.uint8)
        assert self.lib.orc_g2_mul(bytes(k), bytes(pt), self._p(out)) == 0
        return out.tobytes()

    def pair(self, g1, g2):
        out = np.zeros(384, np.uint8)
        assert self.lib.orc_pair(bytes(g1), bytes(g2), self._p(out)) == 0
        return out.tobytes()

    def encrypt(self, P, m, r, threads=1):
        m = np.ascontiguousarray(m, dtype=np.int64)
        r = np.frombuffer(bytes(r), np.uint8).copy()
        ct = np.zeros(128 * m.size, np.uint8)
        assert self.lib.orc_encrypt(ctypes.c_size_t(m.size), self._p(m), self._p(r), bytes(P), self._p(ct), threads) == 0
        return ct.tobytes()

    def encrypt_scalar(self, P, m, r, threads=1):
        m = np.frombuffer(bytes(m), np.uint8).copy()
        r = np.frombuffer(bytes(r), np.uint8).copy()
        n = m.size // 32
        ct = np.zeros(128 * n, np.uint8)
        assert self.lib.orc_encrypt_scalar(ctypes.c_size_t(n), self._p(m), self._p(r), bytes(P), self._p(ct), threads) == 0
        return ct.tobytes()

    def ct_scalar_mul(self, scalars, cts, threads=1):
        sc = np.frombuffer(bytes(scalars), np.uint8).copy()
        ct = np.frombuffer(bytes(cts), np.uint8).copy()
        n = ct.size // 128
        out = np.zeros(128 * n, np.uint8)
        rc = self.lib.orc_ct_scalar_mul(ctypes.c_size_t(n), self._p(sc), self._p(ct), self._p(out), threads)
        return rc, out.tobytes()

    def shuffle_rerandomize(self, P, lines, betas, pi, threads=1):
        k = len(lines)
        nq = len(bytes(lines[0])) // 128
        data = np.frombuffer(b''.join(bytes(x) for x in lines), np.uint8).copy()
        beta = np.frombuffer(b''.join(bytes(x) for x in betas), np.uint8).copy()
        perm = np.ascontiguousarray(pi, dtype=np.int32)
        out = np.zeros(k * nq * 128, np.uint8)
        rc = self.lib.orc_shuffle_rerandomize(ctypes.c_size_t(k), ctypes.c_size_t(nq), self._p(perm), self._p(data),
                                              self._p(beta), bytes(P), self._p(out), threads)
        return rc, [out[i * nq * 128:(i + 1) * nq * 128].tobytes() for i in range(k)]

    def schnorr_sign(self, x, k, pub, msg):
        """Sign with the nonce injected: the group and scalar steps in the C++ oracle, SHA-512 here."""
        import hashlib
        R, s = np.zeros(64, np.uint8), np.zeros(32, np.uint8)
        assert self.lib.orc_schnorr_commit(bytes(k), self._p(R)) == 0
        dig = hashlib.sha512(R.tobytes() + bytes(pub) + bytes(msg)).digest()
        assert self.lib.orc_schnorr_finish(bytes(x), bytes(k), dig, self._p(s)) == 0
        return R.tobytes() + s.tobytes()

    def schnorr_verify(self, pubs, sigs, msgs, threads=1):
        import hashlib
        n = len(msgs)
        pub = np.frombuffer(b''.join(bytes(p) for p in pubs), np.uint8).copy()
        sig = np.frombuffer(b''.join(bytes(s) for s in sigs), np.uint8).copy()
        dig = np.frombuffer(b''.join(hashlib.sha512(bytes(sigs[i])[:64] + bytes(pubs[i]) + bytes(msgs[i])).digest()
                                     for i in range(n)), np.uint8).copy()
        ok = np.zeros(n, np.uint8)
        assert self.lib.orc_schnorr_verify(ctypes.c_size_t(n), self._p(pub), self._p(sig), self._p(dig), self._p(ok), threads) == 0
        return ok

    def rep_commit(self, nterms, scalars, points, threads=1):
        sc = np.frombuffer(bytes(scalars), np.uint8).copy()
        pt = np.frombuffer(bytes(points), np.uint8).copy()
        n = pt.size // (64 * nterms)
        out = np.zeros(64 * n, np.uint8)
        rc = self.lib.orc_rep_commit(ctypes.c_size_t(n), nterms, self._p(sc), self._p(pt), self._p(out), threads)
        return rc, out.tobytes()

    def rep_respond(self, per, v, chal, x):
        v = np.frombuffer(bytes(v), np.uint8).copy()
        c = np.frombuffer(bytes(chal), np.uint8).copy()
        x = np.frombuffer(bytes(x), np.uint8).copy()
        n = v.size // 32
        out = np.zeros(32 * n, np.uint8)
        assert self.lib.orc_rep_respond(ctypes.c_size_t(n), per, self._p(v), self._p(c), self._p(x), self._p(out)) == 0
        return out.tobytes()

    def rep_verify(self, nterms, resp, chal, bases, P, V, threads=1):
        arrs = [np.frombuffer(bytes(a), np.uint8).copy() for a in (resp, chal, bases, P, V)]
        n = arrs[3].size // 64
        ok = np.zeros(n, np.uint8)
        assert self.lib.orc_rep_verify(ctypes.c_size_t(n), nterms, *[self._p(a) for a in arrs], self._p(ok), threads) == 0
        return ok

    def key_switch(self, Q, sk, rbs, vs, threads=1):
        rb = np.frombuffer(bytes(rbs), np.uint8).copy()
        v = np.frombuffer(bytes(vs), np.uint8).copy()
        n = rb.size // 64
        cv, neg = np.zeros(128 * n, np.uint8), np.zeros(64 * n, np.uint8)
        rc = self.lib.orc_keyswitch(ctypes.c_size_t(n), bytes(Q), bytes(sk), self._p(rb), self._p(v), self._p(cv),
                                    self._p(neg), threads)
        return rc, cv.tobytes(), neg.tobytes()

    def aggregate(self, vectors, threads=1):
        v = len(vectors[0]) // 128
        cts = np.frombuffer(b''.join(bytes(x) for x in vectors), np.uint8).copy()
        out = np.zeros(128 * v, np.uint8)
        rc = self.lib.orc_aggregate(ctypes.c_size_t(len(vectors)), ctypes.c_size_t(v), self._p(cts), self._p(out), threads)
        return rc, out.tobytes()

    def dlog_reset(self):
        self.lib.orc_dlog_reset()

    def decrypt(self, sk, ct, with_neg=True, threads=1):
        ct = np.frombuffer(bytes(ct), np.uint8).copy()
        n = ct.size // 128
        out = np.zeros(n, np.int64)
        found = np.zeros(n, np.uint8)
        rc = self.lib.orc_decrypt(ctypes.c_size_t(n), bytes(sk), self._p(ct), 1 if with_neg else 0, self._p(out), self._p(found), threads)
        assert rc == 0
        return out, found

    def decrypt_check_zero(self, sk, ct):
        ct = np.frombuffer(bytes(ct), np.uint8).copy()
        n = ct.size // 128
        out = np.zeros(n, np.int64)
        assert self.lib.orc_decrypt_check_zero(ctypes.c_size_t(n), bytes(sk), self._p(ct), self._p(out)) == 0
        return out

    def make_sigs(self, xs, u):
        ns = len(xs)
        x = np.frombuffer(b''.join(int(v).to_bytes(32, 'big') for v in xs), np.uint8).copy()
        Y, A = np.zeros(64 * ns, np.uint8), np.zeros(128 * ns * u, np.uint8)
        assert self.lib.orc_make_sigs(ns, ctypes.c_int64(u), self._p(x), self._p(Y), self._p(A)) == 0
        return Y.tobytes(), A.tobytes()

    def proof_size(self, ns, nd):
        return int(self.lib.orc_rp_proof_size(ns, nd))

    def rp_create(self, ns, u, l, Y, A, P, secrets, rs, cts, rnd_list, threads=1):
        import bn256_ref as R
        n = len(secrets)
        secrets = np.ascontiguousarray(secrets, dtype=np.int64)
        off = np.zeros(n + 1, np.uint64)
        roff = np.zeros(n + 1, np.uint64)
        for p in range(n):
            nd = len(R.to_base(int(secrets[p]), u, l)) if u else 0
            off[p + 1] = off[p] + (self.proof_size(ns, nd) if u else 128)
            roff[p + 1] = roff[p] + len(rnd_list[p])
        rnd = np.frombuffer(b''.join(bytes(x) for x in rnd_list) + b'\0', np.uint8).copy()
        r = np.frombuffer(bytes(rs), np.uint8).copy()
        ct = np.frombuffer(bytes(cts), np.uint8).copy()
        proofs = np.zeros(int(off[n]), np.uint8)
        rc = self.lib.orc_rp_create(ctypes.c_size_t(n), ns, ctypes.c_int64(u), ctypes.c_int64(l), bytes(Y), bytes(A), bytes(P),
                                    self._p(secrets), self._p(r), self._p(ct), self._p(rnd), self._p(roff), self._p(proofs),
                                    self._p(off), threads)
        assert rc == 0
        return proofs.tobytes(), off

    def rp_verify(self, ns, u, l, Y, P, proofs, off, threads=1):
        proofs = np.frombuffer(bytes(proofs), np.uint8).copy()
        off = np.ascontiguousarray(off, dtype=np.uint64)
        n = off.size - 1
        ok = np.zeros(n, np.uint8)
        rc = self.lib.orc_rp_verify(ctypes.c_size_t(n), ns, ctypes.c_int64(u), ctypes.c_int64(l), bytes(Y), bytes(P),
                                    self._p(proofs), self._p(off), self._p(ok), threads)
        assert rc == 0
        return ok


    # ---- edwards25519 suite (oracle/oracle_ed.cpp)
    def ed_mul(self, k, pt=None):
        out = np.zeros(32, np.uint8)
        rc = self.lib.orc_ed_mul(bytes(k), None if pt is None else bytes(pt), self._p(out))
        return None if rc else out.tobytes()

    def ed_add(self, a, b):
        out = np.zeros(32, np.uint8)
        rc = self.lib.orc_ed_add(bytes(a), bytes(b), self._p(out))
        return None if rc else out.tobytes()

    def ed_encrypt(self, P, m, r, threads=1):
        m = np.ascontiguousarray(m, dtype=np.int64)
        r = np.frombuffer(bytes(r), np.uint8).copy()
        ct = np.zeros(64 * m.size, np.uint8)
        assert self.lib.orc_ed_encrypt(ctypes.c_size_t(m.size), self._p(m), self._p(r), bytes(P), self._p(ct), threads) == 0
        return ct.tobytes()

    def ed_aggregate(self, vectors, threads=1):
        v = len(vectors[0]) // 64
        cts = np.frombuffer(b''.join(bytes(x) for x in vectors), np.uint8).copy()
        out = np.zeros(64 * v, np.uint8)
        rc = self.lib.orc_ed_aggregate(ctypes.c_size_t(len(vectors)), ctypes.c_size_t(v), self._p(cts), self._p(out), threads)
        return rc, out.tobytes()

    def ed_dlog_reset(self):
        self.lib.orc_ed_dlog_reset()

    def ed_decrypt(self, sk, ct, with_neg=True, threads=1):
        ct = np.frombuffer(bytes(ct), np.uint8).copy()
        n = ct.size // 64
        out = np.zeros(n, np.int64)
        found = np.zeros(n, np.uint8)
        rc = self.lib.orc_ed_decrypt(ctypes.c_size_t(n), bytes(sk), self._p(ct), 1 if with_neg else 0, self._p(out),
                                     self._p(found), threads)
        return rc, out, found


@pytest.fixture(scope='session')
def oracle():
    return Oracle()


@pytest.fixture(scope='session')
def backend():
    from drynx_b200 import Backend
    be = Backend(0)
    yield be
    be.close()


class SplitMix:
    """SplitMix64 -> scalars by kyber-style rejection sampling (32 random bytes, accept if < n)."""

    def __init__(self, seed):
        self.s = seed & 0xffffffffffffffff

    def u64(self):
        self.s = (self.s + 0x9e3779b97f4a7c15) & 0xffffffffffffffff
        z = self.s
        z = ((z ^ (z >> 30)) * 0xbf58476d1ce4e5b9) & 0xffffffffffffffff
        z = ((z ^ (z >> 27)) * 0x94d049bb133111eb) & 0xffffffffffffffff
        return z ^ (z >> 31)

    def scalar(self):
        import bn256_ref as R
        while True:
            v = (self.u64() << 192) | (self.u64() << 128) | (self.u64() << 64) | self.u64()
            if v < R.n:
                return v

    def scalars_bytes(self, k):
        return b''.join(self.scalar().to_bytes(32, 'big') for _ in range(k))

    def randint(self, lo, hi):
        return lo + self.u64() % (hi - lo + 1)


@pytest.fixture()
def rng():
    return SplitMix(12345)
