"""Differential fuzzing of the CUDA path against the oracle through the C ABI (test infrastructure, like the oracle).

Every case draws a range-proof shape (ns, u, l), keys, secrets (in range, at the edges, out of range, negative), and the
prover's randomness; compares the created proof bytes with the oracle's; mutates about half of the proofs (bit flips
anywhere, 32-byte words zeroed / set to 0xff.. / replaced / swapped / copied from another proof) and compares the
accept bits of a randomly chosen verification path with the oracle's (lib/range/range_proof.go:504-565 as restated in
oracle/oracle.cpp). A second family does the same for ElGamal vectors: encryption, aggregation with damaged points,
decryption with negatives and out-of-table plaintexts.

    python tests/fuzz_parity.py --seconds 240 --seed 1 [--log gpurun_out/fuzz.log]

tests/test_gpu_fuzz.py runs a few seeded seconds of both families with `-m gpu`.
"""
import argparse
import os
import sys
import time

import numpy as np

HERE = os.path.dirname(os.path.abspath(__file__))
sys.path.insert(0, HERE)
sys.path.insert(0, os.path.join(HERE, '..', 'oracle'))
sys.path.insert(0, os.path.join(HERE, '..'))

import bn256_ref as R  # noqa: E402

SHAPES_U = [2, 3, 4, 5, 8, 10, 16]


def _scalar_bytes(g, k):
    """k scalars below n as big-endian bytes (top bit cleared, then reduced by rejection like kyber's Pick)."""
    out = []
    while len(out) < k:
        v = int.from_bytes(g.bytes(32), 'big') >> 2
        if v < R.n:
            out.append(v.to_bytes(32, 'big'))
    return b''.join(out)


def _mutate(g, buf, lo, hi, other):
    """One random mutation of proof bytes buf[lo:hi]; `other` = (lo, hi) of another proof of the same shape or None."""
    kind = int(g.integers(0, 7))
    size = hi - lo
    words = size // 32
    if kind == 0:
        pos = lo + int(g.integers(0, size))
        buf[pos] ^= 1 << int(g.integers(0, 8))
        return 'bit'
    w = lo + 32 * int(g.integers(0, words))
    if kind == 1:
        buf[w:w + 32] = bytes(32)
        return 'zero word'
    if kind == 2:
        buf[w:w + 32] = b'\xff' * 32
        return 'ff word'
    if kind == 3:
        buf[w:w + 32] = g.bytes(32)
        return 'random word'
    if kind == 4:
        w2 = lo + 32 * int(g.integers(0, words))
        a, b = bytes(buf[w:w + 32]), bytes(buf[w2:w2 + 32])
        buf[w:w + 32], buf[w2:w2 + 32] = b, a
        return 'swap words'
    if kind == 5 and other is not None and other[1] - other[0] == size:
        buf[w:w + 32] = buf[other[0] + (w - lo):other[0] + (w - lo) + 32]
        return 'word of another proof'
    # a whole 128-byte block (a G2 point or a third of a GT element) cleared
    blocks = size // 128
    b0 = lo + 128 * int(g.integers(0, blocks))
    buf[b0:b0 + 128] = bytes(128)
    return 'zero block'


def fuzz_rangeproofs(backend, oracle, seed, seconds, log=None, max_cases=None, threads=None):
    g = np.random.default_rng(seed)
    threads = threads or min(16, os.cpu_count() or 1)
    deadline = time.time() + seconds
    stats = {'cases': 0, 'proofs_created': 0, 'proofs_verified': 0, 'accepted': 0, 'rejected': 0, 'mutations': {}}
    while time.time() < deadline and (max_cases is None or stats['cases'] < max_cases):
        ns = int(g.integers(1, 6))
        u = int(g.choice(SHAPES_U))
        lmax = 1
        while lmax < 8 and u ** (lmax + 1) < (1 << 40):
            lmax += 1
        l = int(g.integers(1, lmax + 1))
        n = int(g.integers(1, 25))
        sid = int(g.integers(0, 64))
        sk = int.from_bytes(_scalar_bytes(g, 1), 'big') or 1
        P = R.g1_to_bytes(R.g1_mul(sk, R.G1_GEN))
        backend.set_pubkey(P)
        xs = [int.from_bytes(_scalar_bytes(g, 1), 'big') or 1 for _ in range(ns)]
        Y, A = oracle.make_sigs(xs, u)
        backend.set_signatures(sid, ns, u, Y, A)
        top = u ** l
        secrets = []
        for _ in range(n):
            c = int(g.integers(0, 10))
            if c == 0:
                secrets.append(int(g.choice([0, 1, top - 1, top // 2])))
            elif c == 1:
                secrets.append(int(top + g.integers(0, top)))        # needs more digits: the size check fails
            elif c == 2:
                secrets.append(-int(g.integers(1, 1000)))            # all-zero digits, inconsistent D
            else:
                secrets.append(int(g.integers(0, top)))
        rs = _scalar_bytes(g, n)
        cts = backend.encrypt_int_vector(secrets, rs).tobytes()
        assert cts == oracle.encrypt(P, secrets, rs, threads=threads), ('encrypt', seed, stats['cases'])
        rnd = [_scalar_bytes(g, len(R.to_base(s, u, l)) * (3 + ns)) for s in secrets]
        opts = {'rp_create_path': int(g.integers(0, 4)),
                'rp_verify_path': int(g.choice([0, 1, 2])),
                'rp_ucode_lanes': int(g.choice([0, 4, 8])),
                'rp_ucode_mode': int(g.choice([0, 1, 2])),
                'rp_verify_gt': int(g.choice([0, 2])),
                'rp_verify_hdr_early': int(g.integers(0, 2))}
        for k, v in opts.items():
            backend.set_option(k, v)
        try:
            proofs, off = backend.create_range_proofs([sid] * n, {sid: ns}, {sid: u}, [l] * n, secrets, rs, cts, rnd)
            want, woff = oracle.rp_create(ns, u, l, Y, A, P, secrets, rs, cts, rnd, threads=threads)
            assert list(off) == list(woff), ('offsets', seed, stats['cases'], ns, u, l, opts)
            assert proofs.tobytes() == want, ('created bytes', seed, stats['cases'], ns, u, l, secrets, opts)
            stats['proofs_created'] += n
            bad = bytearray(want)
            for p in range(n):
                if g.integers(0, 2):
                    continue
                q = int(g.integers(0, n))
                other = (int(off[q]), int(off[q + 1])) if q != p else None
                for _ in range(int(g.integers(1, 3))):
                    m = _mutate(g, bad, int(off[p]), int(off[p + 1]), other)
                    stats['mutations'][m] = stats['mutations'].get(m, 0) + 1
            ok = backend.verify_range_proofs([sid] * n, [l] * n, bytes(bad), off)
            ref = oracle.rp_verify(ns, u, l, Y, P, bytes(bad), off, threads=threads)
            assert list(ok) == list(ref), ('accept bits', seed, stats['cases'], ns, u, l, secrets, opts, list(ok), list(ref))
        finally:
            for k in opts:
                backend.set_option(k, 0)
        stats['cases'] += 1
        stats['proofs_verified'] += n
        stats['accepted'] += int(np.sum(ref))
        stats['rejected'] += n - int(np.sum(ref))
        if log:
            log.write('rp case %d: ns=%d u=%d l=%d n=%d %s accepted %d/%d\n' % (stats['cases'], ns, u, l, n, opts, int(np.sum(ref)), n))
    return stats


def fuzz_mixed_batches(backend, oracle, seed, seconds, log=None, max_cases=None, threads=None):
    """One call over several signature sets: per-proof (set, l), commit-only proofs (u = l = 0,
    lib/range/range_proof.go:322-324, :508-510) in between, some proofs verified with another l than they were made
    with. The oracle judges each (set, l) group on its own."""
    g = np.random.default_rng(seed ^ 0xba7c4)
    threads = threads or min(16, os.cpu_count() or 1)
    deadline = time.time() + seconds
    stats = {'cases': 0, 'proofs_verified': 0, 'accepted': 0, 'rejected': 0, 'commit_only': 0, 'wrong_l': 0}
    while time.time() < deadline and (max_cases is None or stats['cases'] < max_cases):
        sk = int.from_bytes(_scalar_bytes(g, 1), 'big') or 1
        P = R.g1_to_bytes(R.g1_mul(sk, R.G1_GEN))
        backend.set_pubkey(P)
        sets = {}
        for sid in g.choice(np.arange(64, 128), size=int(g.integers(1, 4)), replace=False):
            ns, u = int(g.integers(1, 5)), int(g.choice(SHAPES_U))
            xs = [int.from_bytes(_scalar_bytes(g, 1), 'big') or 1 for _ in range(ns)]
            Y, A = oracle.make_sigs(xs, u)
            backend.set_signatures(int(sid), ns, u, Y, A)
            sets[int(sid)] = (ns, u, Y, A)
        n = int(g.integers(1, 49))
        plan = []
        for _ in range(n):
            if g.integers(0, 6) == 0:
                plan.append((-1, 0, int(g.integers(0, 1000))))
                continue
            sid = int(g.choice(list(sets)))
            u = sets[sid][1]
            lmax = 1
            while lmax < 6 and u ** (lmax + 1) < (1 << 32):
                lmax += 1
            l = int(g.integers(1, lmax + 1))
            top = u ** l
            s_ = int(g.integers(0, top)) if g.integers(0, 8) else int(top + g.integers(0, top))
            plan.append((sid, l, s_))
        sids = [p[0] for p in plan]; ls = [p[1] for p in plan]; secrets = [p[2] for p in plan]
        rs = _scalar_bytes(g, n)
        cts = backend.encrypt_int_vector(secrets, rs).tobytes()
        rnd = [b'' if sid < 0 else _scalar_bytes(g, len(R.to_base(s_, sets[sid][1], l)) * (3 + sets[sid][0])) for sid, l, s_ in plan]
        opts = {'rp_create_path': int(g.integers(0, 4)), 'rp_verify_path': int(g.choice([0, 1, 2])),
                'rp_ucode_lanes': int(g.choice([0, 4, 8])), 'rp_ucode_mode': int(g.choice([0, 1, 2])),
                'rp_verify_gt': int(g.choice([0, 2]))}
        for k, v in opts.items():
            backend.set_option(k, v)
        try:
            proofs, off = backend.create_range_proofs(sids, {k: v[0] for k, v in sets.items()}, {k: v[1] for k, v in sets.items()},
                                                      ls, secrets, rs, cts, rnd)
            pb = bytearray(proofs.tobytes())
            for i, (sid, l, s_) in enumerate(plan):
                mine = bytes(pb[int(off[i]):int(off[i + 1])])
                if sid < 0:
                    assert mine == cts[128 * i:128 * i + 128], ('commit-only bytes', seed, stats['cases'], i)
                    continue
                ns, u, Y, A = sets[sid]
                want, _ = oracle.rp_create(ns, u, l, Y, A, P, [s_], rs[32 * i:32 * i + 32], cts[128 * i:128 * i + 128], [rnd[i]])
                assert mine == want, ('created bytes', seed, stats['cases'], i, sid, ns, u, l, s_, opts)
            ls_v = list(ls)
            for i, (sid, l, s_) in enumerate(plan):
                if sid < 0:
                    continue
                if g.integers(0, 20) == 0:
                    ls_v[i] = l + 1 if l == 1 or g.integers(0, 2) else l - 1
                    stats['wrong_l'] += 1
                elif g.integers(0, 3) == 0:
                    _mutate(g, pb, int(off[i]), int(off[i + 1]), None)
            ok = list(backend.verify_range_proofs(sids, ls_v, bytes(pb), off))
            ref = [None] * n
            groups = {}
            for i, (sid, l, s_) in enumerate(plan):
                if sid < 0:
                    ref[i] = 1
                    stats['commit_only'] += 1
                else:
                    groups.setdefault((sid, ls_v[i]), []).append(i)
            for (sid, l), idx in groups.items():
                ns, u, Y, A = sets[sid]
                blob = b''.join(bytes(pb[int(off[i]):int(off[i + 1])]) for i in idx)
                goff = np.zeros(len(idx) + 1, np.uint64)
                for k, i in enumerate(idx):
                    goff[k + 1] = goff[k] + off[i + 1] - off[i]
                r = oracle.rp_verify(ns, u, l, Y, P, blob, goff, threads=threads)
                for k, i in enumerate(idx):
                    ref[i] = int(r[k])
            assert ok == ref, ('accept bits', seed, stats['cases'], plan, ls_v, opts, ok, ref)
        finally:
            for k in opts:
                backend.set_option(k, 0)
        stats['cases'] += 1
        stats['proofs_verified'] += n
        stats['accepted'] += sum(ref)
        stats['rejected'] += n - sum(ref)
        if log:
            log.write('mixed case %d: %d sets, n=%d %s accepted %d/%d\n' % (stats['cases'], len(sets), n, opts, sum(ref), n))
    return stats


def _special_scalars():
    """Scalars around the places where the device code branches: zero, the group order's neighbourhood, the
    endomorphism's eigenvalue lambda = 36u^4 - 1 and the lattice basis of its split, window and limb boundaries."""
    u = R.u
    lam = 36 * u ** 4 - 1
    a1, a2 = 6 * u * u + 2 * u, 2 * u + 1
    out = [0, 1, 2, 3, R.n - 1, R.n - 2, lam, lam + 1, lam - 1, R.n - lam, (lam * lam) % R.n, a1, a2, a1 + a2,
           (a1 * lam) % R.n, (a2 * lam) % R.n, 1 << 127, (1 << 127) - 1, 1 << 128, (1 << 128) + 1, (1 << 128) - 1,
           (1 << 192) - 1, 1 << 253, R.n >> 1, (R.n >> 1) + 1]
    out += [(1 << (32 * k)) - 1 for k in range(1, 8)] + [1 << (32 * k) for k in range(1, 8)]
    out += [int('f' * k, 16) << (4 * j) for k in (1, 2, 8) for j in (0, 4, 60)]
    return [v % R.n for v in out]


def fuzz_group_ops(backend, oracle, seed, seconds, log=None, max_cases=None, threads=None):
    """Variable-base G1 products (endomorphism split), the ciphertext-by-scalar product, key switching, Schnorr
    verification and the proof.Rep relations against the oracle, with the special scalars above mixed in and some
    points damaged."""
    g = np.random.default_rng(seed ^ 0x9047)
    threads = threads or min(16, os.cpu_count() or 1)
    deadline = time.time() + seconds
    special = _special_scalars()
    stats = {'cases': 0, 'g1_mul': 0, 'ct_scalar_mul': 0, 'key_switch': 0, 'schnorr': 0, 'rep': 0}

    def scalars(k):
        vals = [int.from_bytes(_scalar_bytes(g, 1), 'big') if g.integers(0, 3) else int(g.choice(len(special))) for _ in range(k)]
        vals = [special[v] if v < len(special) else v for v in vals]
        return vals, b''.join(v.to_bytes(32, 'big') for v in vals)

    def points(k):
        ks = _scalar_bytes(g, k)
        gen = R.g1_to_bytes(R.G1_GEN)
        return backend.g1_mul(ks, gen * k).tobytes()

    while time.time() < deadline and (max_cases is None or stats['cases'] < max_cases):
        k = int(g.integers(1, 40))
        # a) k * P
        vals, sc = scalars(k)
        pts = points(k)
        got = backend.g1_mul(sc, pts).tobytes()
        for i in range(k):
            want = oracle.g1_mul(sc[32 * i:32 * i + 32], pts[64 * i:64 * i + 64])
            assert got[64 * i:64 * i + 64] == want, ('g1_mul', seed, stats['cases'], hex(vals[i]))
        stats['g1_mul'] += k
        # b) MulCipherTextbyScalar
        sk = int.from_bytes(_scalar_bytes(g, 1), 'big') or 1
        P = R.g1_to_bytes(R.g1_mul(sk, R.G1_GEN))
        backend.set_pubkey(P)
        cts = bytearray(backend.encrypt_int_vector([int(x) for x in g.integers(-1000, 1000, size=k)], _scalar_bytes(g, k)).tobytes())
        damaged = g.integers(0, 5) == 0
        if damaged:
            cts[int(g.integers(0, len(cts)))] ^= 1 << int(g.integers(0, 8))
        vals, sc = scalars(k)
        rc, want = oracle.ct_scalar_mul(sc, bytes(cts), threads=threads)
        try:
            got = backend.ct_scalar_mul(sc, bytes(cts)).tobytes()
            refused = False
        except Exception:
            refused = True
        assert refused == (rc != 0) == bool(damaged), ('ct_scalar_mul refusal', seed, stats['cases'], rc, damaged)
        if not refused:
            assert got == want, ('ct_scalar_mul', seed, stats['cases'])
        stats['ct_scalar_mul'] += k
        # c) key switching contributions
        Q = points(1)
        backend.set_target_pubkey(Q)
        skv, skb = scalars(1)
        rbs = points(k)
        vs_v, vs = scalars(k)
        rc, wcv, wneg = oracle.key_switch(Q, skb, rbs, vs, threads=threads)
        cv, neg = backend.key_switch(skb, rbs, vs, want_rb_neg=True)
        assert rc == 0 and cv.tobytes() == wcv and neg.tobytes() == wneg, ('key_switch', seed, stats['cases'], hex(skv[0]))
        stats['key_switch'] += k
        # d) Schnorr signatures, some of them damaged
        pubs, sigs, msgs = [], [], []
        for i in range(k):
            x = int.from_bytes(_scalar_bytes(g, 1), 'big') or 1
            xb = x.to_bytes(32, 'big')
            pub = oracle.g1_mul(xb, R.g1_to_bytes(R.G1_GEN))
            msg = g.bytes(int(g.integers(0, 80)))
            sig = bytearray(oracle.schnorr_sign(xb, _scalar_bytes(g, 1), pub, msg))
            c = int(g.integers(0, 6))
            if c == 0:
                sig[int(g.integers(0, 96))] ^= 1 << int(g.integers(0, 8))
            elif c == 1:
                sig[64:96] = b'\xff' * 32
            elif c == 2:
                msg = msg + b'x'
            pubs.append(pub); sigs.append(bytes(sig)); msgs.append(msg)
        assert list(backend.schnorr_verify(pubs, sigs, msgs)) == list(oracle.schnorr_verify(pubs, sigs, msgs, threads=threads)), \
            ('schnorr', seed, stats['cases'])
        stats['schnorr'] += k
        # e) proof.Rep relations: commit, respond, verify; a third of the responses damaged
        nterms = int(g.integers(1, 5))
        xv, xs = scalars(k * nterms)
        vv, vs = scalars(k * nterms)
        bases = points(k * nterms)
        rc, V = oracle.rep_commit(nterms, vs, bases, threads=threads)
        assert rc == 0 and backend.rep_commit(nterms, vs, bases).tobytes() == V, ('rep_commit', seed, stats['cases'])
        rc, Pp = oracle.rep_commit(nterms, xs, bases, threads=threads)
        cv_, chal = scalars(k)
        resp = oracle.rep_respond(nterms, vs, chal, xs)
        assert backend.rep_respond(nterms, vs, chal, xs).tobytes() == resp, ('rep_respond', seed, stats['cases'])
        resp = bytearray(resp)
        for i in range(k):
            if g.integers(0, 3) == 0:
                resp[32 * nterms * i + 31] ^= 1
        want = oracle.rep_verify(nterms, bytes(resp), chal, bases, Pp, V, threads=threads)
        assert list(backend.rep_verify(nterms, bytes(resp), chal, bases, Pp, V)) == list(want), ('rep_verify', seed, stats['cases'])
        stats['rep'] += k
        stats['cases'] += 1
        if log:
            log.write('group case %d: k=%d nterms=%d\n' % (stats['cases'], k, nterms))
    return stats


def fuzz_misc(backend, oracle, seed, seconds, log=None, max_cases=None, threads=None, make_backend=None):
    """What the other families leave out: EncryptScalar, DecryptCheckZero, the shuffle's re-randomisation, the
    edwards25519 suite (full 256-bit little-endian scalars, reduced mod l like SetBytes) and the plaintext reductions
    against numpy's wrapping int64 arithmetic (lib/encoding/sum.go:19-23, mean.go:19-24, variance.go:19-27)."""
    g = np.random.default_rng(seed ^ 0x315c)
    threads = threads or min(16, os.cpu_count() or 1)
    deadline = time.time() + seconds
    special = _special_scalars()
    stats = {'cases': 0, 'encrypt_scalar': 0, 'shuffle_lines': 0, 'ed_ciphertexts': 0, 'reduced_rows': 0}
    backend = make_backend() if make_backend else backend
    oracle.ed_dlog_reset()
    ed_base = None
    while time.time() < deadline and (max_cases is None or stats['cases'] < max_cases):
        k = int(g.integers(1, 30))
        sk = int.from_bytes(_scalar_bytes(g, 1), 'big') or 1
        skb = sk.to_bytes(32, 'big')
        P = R.g1_to_bytes(R.g1_mul(sk, R.G1_GEN))
        backend.set_pubkey(P)
        # a) EncryptScalar with special scalars as messages and as randomness
        ms = b''.join((special[int(g.integers(0, len(special)))] if g.integers(0, 3) == 0 else int.from_bytes(_scalar_bytes(g, 1), 'big')).to_bytes(32, 'big')
                      for _ in range(k))
        rs = b''.join((special[int(g.integers(0, len(special)))] if g.integers(0, 3) == 0 else int.from_bytes(_scalar_bytes(g, 1), 'big')).to_bytes(32, 'big')
                      for _ in range(k))
        assert backend.encrypt_scalar_vector(ms, rs).tobytes() == oracle.encrypt_scalar(P, ms, rs, threads=threads), ('encrypt_scalar', seed, stats['cases'])
        stats['encrypt_scalar'] += k
        # b) DecryptCheckZero on a vector with zeros in it
        vals = [0 if g.integers(0, 3) == 0 else int(g.integers(-50, 50)) for _ in range(k)]
        cts = backend.encrypt_int_vector(vals, _scalar_bytes(g, k)).tobytes()
        assert list(backend.decrypt_check_zero(skb, cts)) == list(oracle.decrypt_check_zero(skb, cts)), ('check_zero', seed, stats['cases'])
        # c) shuffle + re-randomisation of k lines of nq ciphertexts
        nq = int(g.integers(1, 6))
        lines = [backend.encrypt_int_vector([int(x) for x in g.integers(-9, 9, size=nq)], _scalar_bytes(g, nq)).tobytes() for _ in range(k)]
        betas = [_scalar_bytes(g, nq) for _ in range(k)]
        pi = [int(x) for x in g.permutation(k)]
        rc, want = oracle.shuffle_rerandomize(P, lines, betas, pi, threads=threads)
        got = backend.shuffle_rerandomize(lines, betas, pi)
        assert rc == 0 and [bytes(x.tobytes()) for x in got] == want, ('shuffle', seed, stats['cases'])
        stats['shuffle_lines'] += k
        # d) edwards25519: encryption with unreduced scalars, aggregation, decryption (table state follows the calls)
        esk = g.bytes(32)
        if ed_base is None:
            ed_base = oracle.ed_mul((1).to_bytes(32, 'little'))
        eP = oracle.ed_mul(esk)
        backend.ed_set_pubkey(eP)
        ndp, v = int(g.integers(1, 12)), int(g.integers(1, 8))
        msgs = [int(x) for x in g.integers(-2000, 2000, size=ndp * v)]
        ers = g.bytes(32 * ndp * v)
        ect = backend.ed_encrypt_int_vector(msgs, ers).tobytes()
        assert ect == oracle.ed_encrypt(eP, msgs, ers, threads=threads), ('ed encrypt', seed, stats['cases'])
        vecs = [ect[d * v * 64:(d + 1) * v * 64] for d in range(ndp)]
        rc, eagg = oracle.ed_aggregate(vecs, threads=threads)
        assert rc == 0 and backend.ed_aggregate(vecs).tobytes() == eagg, ('ed aggregate', seed, stats['cases'])
        with_neg = bool(g.integers(0, 4))
        dec, found = backend.ed_decrypt_int_vector(esk, eagg, with_neg=with_neg, return_found=True)
        rc, ref, rfound = oracle.ed_decrypt(esk, eagg, with_neg=with_neg)
        assert rc == 0 and list(found) == list(rfound), ('ed decrypt flags', seed, stats['cases'])
        assert [int(x) for x, f in zip(dec, found) if f] == [int(x) for x, f in zip(ref, rfound) if f], ('ed decrypt', seed, stats['cases'])
        stats['ed_ciphertexts'] += ndp * v
        # e) plaintext reductions, wrap-around of int64 included
        ndp, nrows = int(g.integers(1, 20)), int(g.integers(0, 600))
        span = int(g.choice([3, 1 << 20, 1 << 40, (1 << 62) - 1]))
        rows = g.integers(-span, span, size=(ndp, nrows), dtype=np.int64)
        with np.errstate(over='ignore'):
            ssum = rows.sum(axis=1)
            q = (rows * rows).sum(axis=1)
        cnt = np.full(ndp, nrows, np.int64)
        assert (backend.reduce_rows(rows, 2) == np.stack([ssum, cnt, q], 1)).all(), ('reduce_rows', seed, stats['cases'])
        assert (backend.reduce_rows(rows, 3) == np.stack([ssum, cnt, ssum, cnt, q], 1)).all(), ('reduce_rows 3', seed, stats['cases'])
        qmin = int(g.integers(-5, 5))
        qmax = qmin + int(g.integers(0, 40))
        frows = g.integers(qmin, qmax + 1, size=(ndp, nrows), dtype=np.int64)
        want = np.stack([np.bincount(frows[d] - qmin, minlength=qmax - qmin + 1) for d in range(ndp)]).astype(np.int64)
        assert (backend.freq_count(frows, qmin, qmax) == want).all(), ('freq_count', seed, stats['cases'])
        stats['reduced_rows'] += ndp * nrows
        stats['cases'] += 1
        if log:
            log.write('misc case %d: k=%d nq=%d\n' % (stats['cases'], k, nq))
    if make_backend:
        backend.close()
    return stats


def fuzz_elgamal(backend, oracle, seed, seconds, log=None, max_cases=None, threads=None, make_backend=None):
    g = np.random.default_rng(seed ^ 0x5eed)
    threads = threads or min(16, os.cpu_count() or 1)
    deadline = time.time() + seconds
    stats = {'cases': 0, 'ciphertexts': 0, 'damaged_points': 0}
    # the reference's discrete-log map grows lazily and remembers what earlier calls made it store
    # (tests/test_gpu_elgamal.py::test_decrypt_matches_oracle_including_table_state): fresh state on both sides,
    # then the same sequence of calls
    backend = make_backend() if make_backend else backend
    oracle.dlog_reset()
    while time.time() < deadline and (max_cases is None or stats['cases'] < max_cases):
        sk = int.from_bytes(_scalar_bytes(g, 1), 'big') or 1
        P = R.g1_to_bytes(R.g1_mul(sk, R.G1_GEN))
        backend.set_pubkey(P)
        ndp = int(g.integers(1, 40))
        v = int(g.integers(1, 20))
        msgs = g.integers(-3000, 3000, size=ndp * v)
        big = g.integers(0, ndp * v, size=max(1, ndp * v // 16))
        msgs[big] = g.integers(-(1 << 40), 1 << 40, size=big.size)        # outside every table: reported as not found
        msgs = [int(m) for m in msgs]
        rs = _scalar_bytes(g, ndp * v)
        cts = backend.encrypt_int_vector(msgs, rs).tobytes()
        assert cts == oracle.encrypt(P, msgs, rs, threads=threads), ('encrypt', seed, stats['cases'])
        vectors = [bytearray(cts[k * v * 128:(k + 1) * v * 128]) for k in range(ndp)]
        damaged = 0
        with_damage = g.integers(0, 3) == 0
        for k in range(ndp):
            if with_damage and g.integers(0, 8) == 0:                                     # a point that leaves the curve / a coordinate >= p
                pos = 64 * int(g.integers(0, 2 * v))
                if g.integers(0, 2):
                    vectors[k][pos + int(g.integers(0, 64))] ^= 1 << int(g.integers(0, 8))
                else:
                    vectors[k][pos:pos + 32] = b'\xff' * 32
                damaged += 1
        vectors = [bytes(x) for x in vectors]
        rc, want = oracle.aggregate(vectors, threads=threads)
        try:
            got = backend.aggregate(vectors).tobytes()
            refused = False
        except Exception:
            refused = True
        # a flipped bit can land on another curve point only with negligible probability: damaged input is refused
        # by both (a failed UnmarshalBinary in CipherVector.FromBytes), clean input by neither
        assert refused == (rc != 0), ('aggregate refusal', seed, stats['cases'], ndp, v, damaged, rc)
        assert refused == (damaged > 0), ('aggregate refusal vs damage', seed, stats['cases'], damaged)
        if not refused:
            assert got == want, ('aggregate', seed, stats['cases'], ndp, v)
            skb = sk.to_bytes(32, 'big')
            with_neg = bool(g.integers(0, 4))
            dec, found = backend.decrypt_int_vector(skb, got, with_neg=with_neg, return_found=True)
            ref, rfound = oracle.decrypt(skb, got, with_neg=with_neg, threads=1)
            assert list(found) == list(rfound), ('decrypt found flags', seed, stats['cases'])
            assert [int(x) for x, f in zip(dec, found) if f] == [int(x) for x, f in zip(ref, rfound) if f], ('decrypt', seed, stats['cases'])
        stats['cases'] += 1
        stats['ciphertexts'] += ndp * v
        stats['damaged_points'] += damaged
        if log:
            log.write('eg case %d: ndp=%d v=%d damaged=%d\n' % (stats['cases'], ndp, v, damaged))
    if make_backend:
        backend.close()
    return stats


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument('--seconds', type=float, default=120)
    ap.add_argument('--seed', type=int, default=1)
    ap.add_argument('--log', default=None)
    args = ap.parse_args()
    from conftest import Oracle
    from drynx_b200 import Backend
    backend, oracle = Backend(0), Oracle()
    log = open(args.log, 'w') if args.log else None
    t = time.time()
    eg = fuzz_elgamal(None, oracle, args.seed, args.seconds * 0.15, log, make_backend=lambda: Backend(0))
    rp = fuzz_rangeproofs(backend, oracle, args.seed, args.seconds * 0.45, log)
    mx = fuzz_mixed_batches(backend, oracle, args.seed, args.seconds * 0.25, log)
    gr = fuzz_group_ops(backend, oracle, args.seed, args.seconds * 0.15, log)
    mi = fuzz_misc(None, oracle, args.seed, args.seconds * 0.15, log, make_backend=lambda: Backend(0))
    line = 'fuzz seed %d, %.0f s: ElGamal %s; range proofs %s; mixed batches %s; group operations %s; other entry points %s; 0 mismatches' % (
        args.seed, time.time() - t, eg, rp, mx, gr, mi)
    print(line)
    if log:
        log.write(line + '\n')
        log.close()
    backend.close()


if __name__ == '__main__':
    main()
