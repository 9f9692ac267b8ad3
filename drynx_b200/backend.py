"""Python host mirror of the reference's call surface over the C ABI (bytes in, bytes out).

Names follow the Go functions they stand in for (libunlynx.EncryptIntVector, CipherVector.Add,
DecryptIntVectorWithNeg, CreateDecryptionTable; libdrynxrange.InitRangeProofSignature,
CreatePredicateRangeProofListForAllServers, RangeProofListVerification). Buffers are bytes / numpy arrays;
no arithmetic happens here."""
import ctypes

import numpy as np

from . import _lib


def _ptr(a):
    if a is None:
        return None
    if isinstance(a, (bytes, bytearray)):
        return ctypes.cast(ctypes.c_char_p(bytes(a)), ctypes.c_void_p)
    return ctypes.c_void_p(a.ctypes.data)


def _u8(b):
    return np.frombuffer(bytes(b), dtype=np.uint8).copy() if not isinstance(b, np.ndarray) else np.ascontiguousarray(b, dtype=np.uint8)


class Backend:
    """One drx_ctx on one GPU."""

    def __init__(self, device=0, g1_base=None, g2_base=None):
        self.lib = _lib.load()
        h = ctypes.c_void_p()
        g1 = _u8(g1_base) if g1_base is not None else None
        g2 = _u8(g2_base) if g2_base is not None else None
        _lib.check(self.lib.drx_init(ctypes.byref(h), device, _ptr(g1), _ptr(g2)))
        self.h = h
        self.device = device

    def close(self):
        if getattr(self, 'h', None):
            self.lib.drx_free(self.h)
            self.h = None

    def __del__(self):
        try:
            self.close()
        except Exception:
            pass

    # ---- plumbing
    def set_stream(self, cuda_stream):
        _lib.check(self.lib.drx_set_stream(self.h, ctypes.c_void_p(cuda_stream)))

    def set_option(self, name, value):
        _lib.check(self.lib.drx_set_option(self.h, name.encode(), int(value)))

    def take_bad(self):
        """Malformed input points met by the device-buffer (`_dev`) calls since the last take_bad (waits for the stream)."""
        n = ctypes.c_int(0)
        _lib.check(self.lib.drx_take_bad(self.h, ctypes.byref(n)))
        return n.value

    def synchronize(self):
        _lib.check(self.lib.drx_synchronize(self.h))

    @property
    def launches(self):
        return int(self.lib.drx_launch_count(self.h))

    # ---- ElGamal (libunlynx lib/crypto.go call surface)
    def set_pubkey(self, P):
        p = _u8(P)
        _lib.check(self.lib.drx_set_pubkey(self.h, _ptr(p)))

    def reduce_rows(self, rows, mode):
        """Per-provider plaintext reduction (EncodeSum / EncodeMean / EncodeVariance before encryption):
        rows int64[ndp][nrows] -> int64[ndp][w]; mode 0 [sum], 1 [sum, N], 2 [sum, N, sum of squares],
        3 mean vector followed by variance vector."""
        rows = np.ascontiguousarray(rows, dtype=np.int64)
        assert rows.ndim == 2
        w = (1, 2, 3, 5)[mode]
        out = np.empty((rows.shape[0], w), dtype=np.int64)
        _lib.check(self.lib.drx_reduce_rows(self.h, rows.shape[0], rows.shape[1], _ptr(rows), mode, _ptr(out)))
        return out

    def freq_count(self, rows, qmin, qmax):
        """EncodeFreqCount's histogram per provider: rows int64[ndp][nrows] -> int64[ndp][qmax - qmin + 1]."""
        rows = np.ascontiguousarray(rows, dtype=np.int64)
        assert rows.ndim == 2
        out = np.empty((rows.shape[0], int(qmax - qmin + 1)), dtype=np.int64)
        _lib.check(self.lib.drx_freq_count(self.h, rows.shape[0], rows.shape[1], _ptr(rows), int(qmin), int(qmax), _ptr(out)))
        return out

    def logreg_coefficients(self, x_list, y_list, means, sds, k, precision):
        """EncodeLogisticRegression's plaintext step for several data providers: x_list[p] float64[n_p][d],
        y_list[p] int64[n_p] -> int64[ndp][(d+1) + (d+1)^2] (k = 2) packed coefficient vectors."""
        d = len(means)
        ndp = len(x_list)
        off = np.zeros(ndp + 1, dtype=np.uint64)
        for p in range(ndp):
            off[p + 1] = off[p] + len(y_list[p])
        x = np.ascontiguousarray(np.concatenate([np.asarray(a, np.float64).reshape(-1, d) for a in x_list]) if ndp else np.zeros((0, d)))
        y = np.ascontiguousarray(np.concatenate([np.asarray(a, np.int64) for a in y_list]) if ndp else np.zeros(0, np.int64))
        m = np.ascontiguousarray(means, dtype=np.float64)
        s = np.ascontiguousarray(sds, dtype=np.float64)
        ncoef = (d + 1) + ((d + 1) ** 2 if k >= 2 else 0)
        out = np.empty((ndp, ncoef), dtype=np.int64)
        _lib.check(self.lib.drx_logreg_coefficients(self.h, ndp, _ptr(off), d, k, _ptr(x), _ptr(y), _ptr(m), _ptr(s),
                                                    float(precision), _ptr(out)))
        return out

    def encrypt_int_vector(self, m, r):
        """EncryptIntVectorGetRs with injected scalars: m int64[n], r bytes n*32 -> n*128 bytes."""
        m = np.ascontiguousarray(m, dtype=np.int64)
        r = _u8(r)
        assert r.size == 32 * m.size
        ct = np.empty(128 * m.size, dtype=np.uint8)
        _lib.check(self.lib.drx_encrypt(self.h, m.size, _ptr(m), _ptr(r), _ptr(ct)))
        return ct

    def encrypt_scalar_vector(self, m, r):
        """libunlynx.EncryptScalar per element: m and r are n*32 bytes of big-endian scalars -> n*128 bytes."""
        m, r = _u8(m), _u8(r)
        assert m.size == r.size and m.size % 32 == 0
        ct = np.empty(4 * m.size, dtype=np.uint8)
        _lib.check(self.lib.drx_encrypt_scalar(self.h, m.size // 32, _ptr(m), _ptr(r), _ptr(ct)))
        return ct

    def ct_scalar_mul(self, scalars, cts):
        """CipherText.MulCipherTextbyScalar per element: scalars n*32 bytes, cts n*128 bytes -> n*128 bytes."""
        sc, ct = _u8(scalars), _u8(cts)
        n = ct.size // 128
        assert ct.size == 128 * n and sc.size == 32 * n
        out = np.empty(128 * n, dtype=np.uint8)
        _lib.check(self.lib.drx_ct_scalar_mul(self.h, n, _ptr(sc), _ptr(ct), _ptr(out)))
        return out

    def shuffle_rerandomize(self, lines, betas, pi):
        """libunlynxshuffle.ShuffleSequence with beta and pi injected: lines = list of serialised CipherVectors of one
        length, betas = list of nq*32-byte scalar blocks (indexed like lines), pi = permutation of the lines."""
        k = len(lines)
        nq = len(bytes(lines[0])) // 128 if k else 0
        data, beta = _u8(b''.join(bytes(x) for x in lines)), _u8(b''.join(bytes(x) for x in betas))
        assert data.size == k * nq * 128 and beta.size == k * nq * 32
        perm = np.ascontiguousarray(pi, dtype=np.int32)
        out = np.empty(k * nq * 128, dtype=np.uint8)
        _lib.check(self.lib.drx_shuffle_rerandomize(self.h, k, nq, _ptr(perm), _ptr(data), _ptr(beta), _ptr(out)))
        return [out[i * nq * 128:(i + 1) * nq * 128] for i in range(k)]

    def schnorr_verify(self, pubs, sigs, msgs):
        """kyber sign/schnorr Verify over G1 for n (key, signature, message) triples; the SHA-512 of R || P || msg is
        taken here on the host. Returns n accept bytes."""
        import hashlib
        n = len(msgs)
        pub, sig = _u8(b''.join(bytes(p) for p in pubs)), _u8(b''.join(bytes(s) for s in sigs))
        assert pub.size == 64 * n and sig.size == 96 * n
        dig = _u8(b''.join(hashlib.sha512(bytes(sigs[i])[:64] + bytes(pubs[i]) + bytes(msgs[i])).digest() for i in range(n)))
        ok = np.zeros(n, dtype=np.uint8)
        _lib.check(self.lib.drx_schnorr_verify(self.h, n, _ptr(pub), _ptr(sig), _ptr(dig), _ptr(ok)))
        return ok

    def rep_commit(self, nterms, scalars, points):
        """Commitments of n proof.Rep relations with nterms terms each: V_i = sum_j scalars[i][j] * points[i][j].
        scalars n*nterms*32 bytes, points n*nterms*64 bytes -> n*64 bytes."""
        sc, pt = _u8(scalars), _u8(points)
        n = pt.size // (64 * nterms)
        assert pt.size == 64 * nterms * n and sc.size == 32 * nterms * n
        out = np.empty(64 * n, dtype=np.uint8)
        _lib.check(self.lib.drx_rep_commit(self.h, n, nterms, _ptr(sc), _ptr(pt), _ptr(out)))
        return out

    def rep_respond(self, per, v, chal, x):
        """Responses r_i = v_i - chal[i // per] * x_i (all 32-byte big-endian scalars) -> n*32 bytes."""
        v, c, x = _u8(v), _u8(chal), _u8(x)
        n = v.size // 32
        assert v.size == 32 * n and x.size == 32 * n and c.size == 32 * ((n + per - 1) // per)
        out = np.empty(32 * n, dtype=np.uint8)
        _lib.check(self.lib.drx_rep_respond(self.h, n, per, _ptr(v), _ptr(c), _ptr(x), _ptr(out)))
        return out

    def rep_verify(self, nterms, resp, chal, bases, P, V):
        """Accept bytes of n proof.Rep relations: sum_j resp[i][j] * bases[i][j] + chal[i] * P[i] == V[i]."""
        r, c, b, p, v = _u8(resp), _u8(chal), _u8(bases), _u8(P), _u8(V)
        n = p.size // 64
        assert p.size == 64 * n and v.size == 64 * n and c.size == 32 * n
        assert r.size == 32 * nterms * n and b.size == 64 * nterms * n
        ok = np.zeros(n, dtype=np.uint8)
        _lib.check(self.lib.drx_rep_verify(self.h, n, nterms, _ptr(r), _ptr(c), _ptr(b), _ptr(p), _ptr(v), _ptr(ok)))
        return ok

    def set_target_pubkey(self, Q):
        """The querier's public key (64 bytes) for key switching."""
        q = _u8(Q)
        assert q.size == 64
        _lib.check(self.lib.drx_set_target_pubkey(self.h, _ptr(q)))

    def key_switch(self, sk, rbs, vs, want_rb_neg=False):
        """libunlynxkeyswitch.KeySwitchSequence with injected vi: rbs n*64 bytes (K components), vs n*32 bytes ->
        n*128 bytes of contributions (vi*B || vi*Q - sk*rB), plus n*64 bytes of -rB when asked."""
        rb, v, k = _u8(rbs), _u8(vs), _u8(sk)
        n = rb.size // 64
        assert rb.size == 64 * n and v.size == 32 * n and k.size == 32
        cv = np.empty(128 * n, dtype=np.uint8)
        neg = np.empty(64 * n, dtype=np.uint8) if want_rb_neg else None
        _lib.check(self.lib.drx_keyswitch(self.h, n, _ptr(k), _ptr(rb), _ptr(v), _ptr(cv),
                                          _ptr(neg) if want_rb_neg else None))
        return (cv, neg) if want_rb_neg else cv

    def aggregate(self, vectors):
        """CipherVector.Add over a list of serialised CipherVectors (each v*128 bytes)."""
        vs = [_u8(v) for v in vectors]
        v = vs[0].size // 128 if vs else 0
        assert all(x.size == v * 128 for x in vs)
        ptrs = (ctypes.c_void_p * len(vs))(*[x.ctypes.data for x in vs])
        out = np.empty(128 * v, dtype=np.uint8)
        _lib.check(self.lib.drx_aggregate(self.h, len(vs), v, ctypes.cast(ptrs, ctypes.c_void_p), _ptr(out)))
        return out

    def create_decryption_table(self, limit):
        _lib.check(self.lib.drx_dlog_table(self.h, int(limit)))

    def decrypt_int_vector(self, sk, ct, with_neg=True, return_found=False):
        ct = _u8(ct)
        n = ct.size // 128
        out = np.empty(n, dtype=np.int64)
        found = np.empty(n, dtype=np.uint8)
        skb = _u8(sk)
        _lib.check(self.lib.drx_decrypt(self.h, n, _ptr(skb), _ptr(ct), 1 if with_neg else 0, _ptr(out), _ptr(found)))
        return (out, found) if return_found else out

    def decrypt_check_zero(self, sk, ct):
        ct = _u8(ct)
        n = ct.size // 128
        out = np.empty(n, dtype=np.int64)
        skb = _u8(sk)
        _lib.check(self.lib.drx_decrypt_check_zero(self.h, n, _ptr(skb), _ptr(ct), _ptr(out)))
        return out

    # ---- the same ElGamal operations over edwards25519 (UnLynx's default suite): 32-byte points, 64-byte ciphertexts,
    # 32-byte little-endian scalars
    def ed_set_pubkey(self, P):
        p = _u8(P)
        assert p.size == 32
        _lib.check(self.lib.drx_ed_set_pubkey(self.h, _ptr(p)))

    def ed_encrypt_int_vector(self, m, r):
        m = np.ascontiguousarray(m, dtype=np.int64)
        r = _u8(r)
        assert r.size == 32 * m.size
        ct = np.empty(64 * m.size, dtype=np.uint8)
        _lib.check(self.lib.drx_ed_encrypt(self.h, m.size, _ptr(m), _ptr(r), _ptr(ct)))
        return ct

    def ed_aggregate(self, vectors):
        vs = [_u8(v) for v in vectors]
        v = vs[0].size // 64 if vs else 0
        assert all(x.size == v * 64 for x in vs)
        ptrs = (ctypes.c_void_p * len(vs))(*[x.ctypes.data for x in vs])
        out = np.empty(64 * v, dtype=np.uint8)
        _lib.check(self.lib.drx_ed_aggregate(self.h, len(vs), v, ctypes.cast(ptrs, ctypes.c_void_p), _ptr(out)))
        return out

    def ed_create_decryption_table(self, limit):
        _lib.check(self.lib.drx_ed_dlog_table(self.h, int(limit)))

    def ed_decrypt_int_vector(self, sk, ct, with_neg=True, return_found=False):
        ct = _u8(ct)
        n = ct.size // 64
        out = np.empty(n, dtype=np.int64)
        found = np.empty(n, dtype=np.uint8)
        skb = _u8(sk)
        _lib.check(self.lib.drx_ed_decrypt(self.h, n, _ptr(skb), _ptr(ct), 1 if with_neg else 0, _ptr(out), _ptr(found)))
        return (out, found) if return_found else out

    def ed_point_mul(self, k, pt):
        k, pt = _u8(k), _u8(pt)
        n = k.size // 32
        assert pt.size == 32 * n
        out = np.empty(32 * n, dtype=np.uint8)
        _lib.check(self.lib.drx_ed_point_mul(self.h, n, _ptr(k), _ptr(pt), _ptr(out)))
        return out

    # ---- range proofs (lib/range/range_proof.go call surface)
    def init_range_proof_signature(self, x_list, u):
        """InitRangeProofSignature for len(x_list) servers from caller-drawn private keys: (Y bytes, A bytes)."""
        ns = len(x_list)
        x = _u8(b''.join(int(v).to_bytes(32, 'big') for v in x_list))
        Y = np.empty(64 * ns, dtype=np.uint8)
        A = np.empty(128 * ns * u, dtype=np.uint8)
        _lib.check(self.lib.drx_rp_make_sigs(self.h, ns, u, _ptr(x), _ptr(Y), _ptr(A)))
        return Y, A

    def set_signatures(self, set_id, ns, u, Y, A):
        Y, A = _u8(Y), _u8(A)
        assert Y.size == 64 * ns and A.size == 128 * ns * u
        _lib.check(self.lib.drx_rp_set_sigs(self.h, set_id, ns, u, _ptr(Y), _ptr(A)))

    def ndigits(self, secret, u, l):
        return int(self.lib.drx_rp_ndigits(secret, u, l))

    def proof_size(self, ns, nd):
        return int(self.lib.drx_rp_proof_size(ns, nd))

    def create_range_proofs(self, set_ids, ns_of_set, u_of_set, ls, secrets, rs, cts, rnd_list):
        """CreatePredicateRangeProofListForAllServers. rnd_list[p] = bytes of nd*(3+ns) scalars in draw order.
        Returns (proof bytes, offsets[n+1])."""
        n = len(set_ids)
        set_ids = np.ascontiguousarray(set_ids, dtype=np.int32)
        ls = np.ascontiguousarray(ls, dtype=np.int64)
        secrets = np.ascontiguousarray(secrets, dtype=np.int64)
        rs, cts = _u8(rs), _u8(cts)
        off = np.zeros(n + 1, dtype=np.uint64)
        roff = np.zeros(n + 1, dtype=np.uint64)
        for p in range(n):
            if set_ids[p] < 0:
                sz = 128
            else:
                nd = self.ndigits(int(secrets[p]), int(u_of_set[int(set_ids[p])]), int(ls[p]))
                sz = self.proof_size(int(ns_of_set[int(set_ids[p])]), nd)
            off[p + 1] = off[p] + sz
            roff[p + 1] = roff[p] + len(rnd_list[p])
        rnd = _u8(b''.join(bytes(x) for x in rnd_list)) if n else np.empty(0, np.uint8)
        proofs = np.empty(int(off[n]), dtype=np.uint8)
        _lib.check(self.lib.drx_rp_create(self.h, n, _ptr(set_ids), _ptr(ls), _ptr(secrets), _ptr(rs), _ptr(cts),
                                          _ptr(rnd), _ptr(roff), _ptr(proofs), _ptr(off)))
        return proofs, off

    def verify_range_proofs(self, set_ids, ls, proofs, off):
        n = len(set_ids)
        set_ids = np.ascontiguousarray(set_ids, dtype=np.int32)
        ls = np.ascontiguousarray(ls, dtype=np.int64)
        proofs = _u8(proofs)
        off = np.ascontiguousarray(off, dtype=np.uint64)
        ok = np.empty(n, dtype=np.uint8)
        _lib.check(self.lib.drx_rp_verify(self.h, n, _ptr(set_ids), _ptr(ls), _ptr(proofs), _ptr(off), _ptr(ok)))
        return ok

    # ---- diagnostics
    def pair(self, g1, g2):
        g1, g2 = _u8(g1), _u8(g2)
        n = g1.size // 64
        out = np.empty(384 * n, dtype=np.uint8)
        _lib.check(self.lib.drx_pair(self.h, n, _ptr(g1), _ptr(g2), _ptr(out)))
        return out

    def g1_mul(self, k, pt):
        k, pt = _u8(k), _u8(pt)
        n = k.size // 32
        out = np.empty(64 * n, dtype=np.uint8)
        _lib.check(self.lib.drx_g1_mul(self.h, n, _ptr(k), _ptr(pt), _ptr(out)))
        return out

    def g2_mul(self, k, pt):
        k, pt = _u8(k), _u8(pt)
        n = k.size // 32
        out = np.empty(128 * n, dtype=np.uint8)
        _lib.check(self.lib.drx_g2_mul(self.h, n, _ptr(k), _ptr(pt), _ptr(out)))
        return out

    def mont_mul(self, which, a, b):
        a = np.ascontiguousarray(a, dtype=np.uint32)
        b = np.ascontiguousarray(b, dtype=np.uint32)
        r = np.empty_like(a)
        _lib.check(self.lib.drx_mont_mul(self.h, a.size // 8, which, _ptr(a), _ptr(b), _ptr(r)))
        return r

    def microbench(self, kind, iters):
        ms, work = ctypes.c_double(), ctypes.c_double()
        _lib.check(self.lib.drx_microbench(self.h, kind, iters, ctypes.byref(ms), ctypes.byref(work)))
        return ms.value, work.value
