"""CPU: the differential fuzzer (tests/fuzz_parity.py) run with the oracle on both sides -- checks the fuzzer's own
plumbing (shapes, offsets, mutations, that both accepted and rejected proofs occur), not the CUDA path."""
import numpy as np

import fuzz_parity


class OracleAsBackend:
    """The few Backend methods the fuzzer calls, answered by the oracle."""

    def __init__(self, oracle):
        self.o = oracle
        self.sets = {}
        self.P = None

    def set_pubkey(self, P):
        self.P = bytes(P)

    def set_option(self, name, value):
        pass

    def set_signatures(self, sid, ns, u, Y, A):
        self.sets[sid] = (ns, u, bytes(Y), bytes(A))

    def encrypt_int_vector(self, m, r):
        return np.frombuffer(self.o.encrypt(self.P, m, r), np.uint8)

    def create_range_proofs(self, sids, ns_of, u_of, ls, secrets, rs, cts, rnd):
        parts = []
        for i, sid in enumerate(sids):
            if sid < 0:
                parts.append(bytes(cts[128 * i:128 * i + 128]))
                continue
            ns, u, Y, A = self.sets[sid]
            one, _ = self.o.rp_create(ns, u, ls[i], Y, A, self.P, [secrets[i]], rs[32 * i:32 * i + 32], cts[128 * i:128 * i + 128], [rnd[i]])
            parts.append(one)
        off = np.zeros(len(sids) + 1, np.uint64)
        off[1:] = np.cumsum([len(x) for x in parts])
        return np.frombuffer(b''.join(parts), np.uint8), off

    def verify_range_proofs(self, sids, ls, proofs, off):
        ok = []
        for i, sid in enumerate(sids):
            if sid < 0:
                ok.append(1)
                continue
            ns, u, Y, A = self.sets[sid]
            one = bytes(proofs[int(off[i]):int(off[i + 1])])
            ok.append(int(self.o.rp_verify(ns, u, ls[i], Y, self.P, one, np.array([0, len(one)], np.uint64))[0]))
        return np.array(ok, np.uint8)

    def aggregate(self, vectors):
        rc, out = self.o.aggregate(vectors)
        if rc:
            raise ValueError('malformed point')
        return np.frombuffer(out, np.uint8)

    def decrypt_int_vector(self, sk, ct, with_neg=True, return_found=False):
        # the oracle has one table: ask once and hand the fuzzer's second (oracle) call the same state by not growing it twice
        return self.o.decrypt(sk, ct, with_neg=with_neg)

    def g1_mul(self, k, pt):
        k, pt = bytes(k), bytes(pt)
        return np.frombuffer(b''.join(self.o.g1_mul(k[32 * i:32 * i + 32], pt[64 * i:64 * i + 64]) for i in range(len(k) // 32)), np.uint8)

    def ct_scalar_mul(self, sc, cts):
        rc, out = self.o.ct_scalar_mul(sc, cts)
        if rc:
            raise ValueError('malformed point')
        return np.frombuffer(out, np.uint8)

    def set_target_pubkey(self, Q):
        self.Q = bytes(Q)

    def key_switch(self, sk, rbs, vs, want_rb_neg=False):
        rc, cv, neg = self.o.key_switch(self.Q, sk, rbs, vs)
        return np.frombuffer(cv, np.uint8), np.frombuffer(neg, np.uint8)

    def schnorr_verify(self, pubs, sigs, msgs):
        return self.o.schnorr_verify(pubs, sigs, msgs)

    def rep_commit(self, nterms, sc, pts):
        return np.frombuffer(self.o.rep_commit(nterms, sc, pts)[1], np.uint8)

    def rep_respond(self, per, v, chal, x):
        return np.frombuffer(self.o.rep_respond(per, v, chal, x), np.uint8)

    def rep_verify(self, nterms, resp, chal, bases, P, V):
        return self.o.rep_verify(nterms, resp, chal, bases, P, V)

    def encrypt_scalar_vector(self, m, r):
        return np.frombuffer(self.o.encrypt_scalar(self.P, m, r), np.uint8)

    def decrypt_check_zero(self, sk, ct):
        return self.o.decrypt_check_zero(sk, ct)

    def shuffle_rerandomize(self, lines, betas, pi):
        return [np.frombuffer(x, np.uint8) for x in self.o.shuffle_rerandomize(self.P, lines, betas, pi)[1]]

    def ed_set_pubkey(self, P):
        self.eP = bytes(P)

    def ed_encrypt_int_vector(self, m, r):
        return np.frombuffer(self.o.ed_encrypt(self.eP, m, r), np.uint8)

    def ed_aggregate(self, vectors):
        return np.frombuffer(self.o.ed_aggregate(vectors)[1], np.uint8)

    def ed_decrypt_int_vector(self, sk, ct, with_neg=True, return_found=False):
        rc, out, found = self.o.ed_decrypt(sk, ct, with_neg=with_neg)
        return out, found

    def reduce_rows(self, rows, mode):
        with np.errstate(over='ignore'):
            s_, q, cnt = rows.sum(axis=1), (rows * rows).sum(axis=1), np.full(rows.shape[0], rows.shape[1], np.int64)
        return np.stack([s_, cnt, q] if mode == 2 else [s_, cnt, s_, cnt, q], 1)

    def freq_count(self, rows, qmin, qmax):
        return np.stack([np.bincount(r - qmin, minlength=qmax - qmin + 1) for r in rows]).astype(np.int64)

    def close(self):
        pass


def test_fuzzer_plumbing_rangeproofs(oracle):
    stats = fuzz_parity.fuzz_rangeproofs(OracleAsBackend(oracle), oracle, 5, seconds=120, max_cases=3, threads=4)
    assert stats['cases'] == 3 and stats['proofs_verified'] == stats['accepted'] + stats['rejected']
    assert stats['rejected'] > 0 and stats['accepted'] > 0 and sum(stats['mutations'].values()) > 0


def test_fuzzer_plumbing_elgamal(oracle):
    stats = fuzz_parity.fuzz_elgamal(None, oracle, 5, seconds=60, max_cases=6, threads=4, make_backend=lambda: OracleAsBackend(oracle))
    assert stats['cases'] == 6 and stats['ciphertexts'] > 0


def test_fuzzer_plumbing_mixed_batches(oracle):
    stats = fuzz_parity.fuzz_mixed_batches(OracleAsBackend(oracle), oracle, 5, seconds=120, max_cases=2, threads=4)
    assert stats['cases'] == 2 and stats['proofs_verified'] == stats['accepted'] + stats['rejected']


def test_fuzzer_plumbing_group_operations(oracle):
    stats = fuzz_parity.fuzz_group_ops(OracleAsBackend(oracle), oracle, 5, seconds=120, max_cases=2, threads=4)
    assert stats['cases'] == 2 and stats['rep'] > 0


def test_fuzzer_plumbing_other_entry_points(oracle):
    stats = fuzz_parity.fuzz_misc(None, oracle, 5, seconds=120, max_cases=2, threads=4, make_backend=lambda: OracleAsBackend(oracle))
    assert stats['cases'] == 2 and stats['ed_ciphertexts'] > 0
