"""-m gpu: range-proof creation and verification through the C ABI against the oracle.
Mirrors lib/range/range_proof_test.go:43-77 (create -> bytes -> verify == true) and the *WithProofs encoder tests
(lib/encoding/sum_test.go:34-57), and adds byte-level equality plus the rejection cases the reference lacks."""
import json
import os

import numpy as np
import pytest

import bn256_ref as R
from conftest import SplitMix

pytestmark = pytest.mark.gpu
GOLD = os.path.join(os.path.dirname(__file__), 'golden')


def _setup(backend, oracle, rng, ns, u, set_id=0):
    sk = rng.scalar()
    P = R.g1_to_bytes(R.g1_mul(sk, R.G1_GEN))
    backend.set_pubkey(P)
    xs = [rng.scalar() for _ in range(ns)]
    Y, A = backend.init_range_proof_signature(xs, u)
    Yo, Ao = oracle.make_sigs(xs, u)
    assert Y.tobytes() == Yo and A.tobytes() == Ao
    backend.set_signatures(set_id, ns, u, Y, A)
    return sk, P, Yo, Ao


def _make(backend, oracle, rng, P, ns, u, l, secrets):
    rs = b''.join(rng.scalar().to_bytes(32, 'big') for _ in secrets)
    cts = backend.encrypt_int_vector(secrets, rs).tobytes()
    rnd = [rng.scalars_bytes(len(R.to_base(s, u, l)) * (3 + ns)) for s in secrets]
    return rs, cts, rnd


def test_deterministic_signature(backend, oracle):
    """InitRangeProofSignatureDeterministic: x = 12 (lib/range/range_proof.go:249-267)."""
    Y, A = backend.init_range_proof_signature([12], 16)
    Yo, Ao = oracle.make_sigs([12], 16)
    assert Y.tobytes() == Yo and A.tobytes() == Ao


@pytest.mark.parametrize('ns,u,l', [(1, 2, 1), (5, 2, 1), (2, 4, 2), (3, 16, 4), (1, 16, 8)])
def test_create_bytes_match_oracle_and_verify(backend, oracle, rng, ns, u, l):
    sk, P, Y, A = _setup(backend, oracle, rng, ns, u)
    secrets = [0, 1, u ** l - 1, (u ** l) // 2] + [rng.randint(0, u ** l - 1) for _ in range(4)]
    rs, cts, rnd = _make(backend, oracle, rng, P, ns, u, l, secrets)
    n = len(secrets)
    proofs, off = backend.create_range_proofs([0] * n, {0: ns}, {0: u}, [l] * n, secrets, rs, cts, rnd)
    want, woff = oracle.rp_create(ns, u, l, Y, A, P, secrets, rs, cts, rnd, threads=8)
    assert list(off) == list(woff)
    assert proofs.tobytes() == want
    ok = backend.verify_range_proofs([0] * n, [l] * n, proofs, off)
    assert list(ok) == [1] * n
    assert list(oracle.rp_verify(ns, u, l, Y, P, proofs.tobytes(), off, threads=8)) == [1] * n


def test_out_of_range_secrets_and_negative(backend, oracle, rng):
    """A secret >= u^l needs more digits (ToBase grows the list) so verification fails the size check (:512);
    a negative secret gets all-zero digits (:587) and an inconsistent D."""
    ns, u, l = 2, 4, 2
    sk, P, Y, A = _setup(backend, oracle, rng, ns, u)
    secrets = [16, 255, -3, 15]
    rs, cts, rnd = _make(backend, oracle, rng, P, ns, u, l, secrets)
    n = len(secrets)
    proofs, off = backend.create_range_proofs([0] * n, {0: ns}, {0: u}, [l] * n, secrets, rs, cts, rnd)
    want, woff = oracle.rp_create(ns, u, l, Y, A, P, secrets, rs, cts, rnd)
    assert proofs.tobytes() == want and list(off) == list(woff)
    ok = backend.verify_range_proofs([0] * n, [l] * n, proofs, off)
    assert list(ok) == list(oracle.rp_verify(ns, u, l, Y, P, want, woff)) == [0, 0, 0, 1]


def test_tampered_and_malformed_proofs_rejected_like_oracle(backend, oracle, rng):
    ns, u, l = 3, 16, 2
    sk, P, Y, A = _setup(backend, oracle, rng, ns, u)
    rs, cts, rnd = _make(backend, oracle, rng, P, ns, u, l, [77])
    proof, off = backend.create_range_proofs([0], {0: ns}, {0: u}, [l], [77], rs, cts, rnd)
    proof = proof.tobytes()
    size = len(proof)
    nd = l
    blk = 256 + 32 * nd
    positions = [5, 70, 130, 165, 200, 250, 260, 300, blk + 3, blk + 128 + 7, blk + 128 * nd + 5, blk + 128 * nd + 384 + 100,
                 blk + 512 * nd + 9, blk + 544 * nd + 64, size - 1]
    batch = [proof]
    for pos in positions:
        b = bytearray(proof)
        b[pos] ^= 0x01
        batch.append(bytes(b))
    z = bytearray(proof); z[192:256] = bytes(64); batch.append(bytes(z))           # D = infinity
    z = bytearray(proof); z[blk:blk + 128] = bytes(128); batch.append(bytes(z))    # V_00 = infinity
    z = bytearray(proof); z[128:160] = bytes(32); batch.append(bytes(z))           # c = 0
    n = len(batch)
    offs = np.arange(n + 1, dtype=np.uint64) * size
    ok = backend.verify_range_proofs([0] * n, [l] * n, b''.join(batch), offs)
    want = oracle.rp_verify(ns, u, l, Y, P, b''.join(batch), offs, threads=8)
    assert list(ok) == list(want)
    assert ok[0] == 1 and sum(ok[1:]) == 0


def test_mixed_batch_sets_sizes_and_bypass(backend, oracle, rng):
    """Different signature sets / (u, l) per proof (QueryToProofsNbrs-style ranges), commit-only proofs
    (u = l = 0, lib/range/range_proof.go:322-324 and :508-510) and a wrong-l proof, in one call."""
    sk = rng.scalar()
    P = R.g1_to_bytes(R.g1_mul(sk, R.G1_GEN))
    backend.set_pubkey(P)
    sets = {}
    for sid, (ns, u) in {3: (2, 4), 5: (1, 16), 6: (2, 2)}.items():
        xs = [rng.scalar() for _ in range(ns)]
        Y, A = oracle.make_sigs(xs, u)
        backend.set_signatures(sid, ns, u, Y, A)
        sets[sid] = (ns, u, Y, A)
    plan = [(3, 2, 9), (-1, 0, 5), (5, 3, 4000), (6, 4, 13), (3, 3, 63), (-1, 0, 0), (5, 1, 15)]
    sids = [p[0] for p in plan]; ls = [p[1] for p in plan]; secrets = [p[2] for p in plan]
    rs = b''.join(rng.scalar().to_bytes(32, 'big') for _ in plan)
    cts = backend.encrypt_int_vector(secrets, rs).tobytes()
    rnd = []
    for sid, l, s in plan:
        rnd.append(b'' if sid < 0 else rng.scalars_bytes(len(R.to_base(s, sets[sid][1], l)) * (3 + sets[sid][0])))
    proofs, off = backend.create_range_proofs(sids, {k: v[0] for k, v in sets.items()}, {k: v[1] for k, v in sets.items()},
                                              ls, secrets, rs, cts, rnd)
    pb = proofs.tobytes()
    for i, (sid, l, s) in enumerate(plan):
        mine = pb[int(off[i]):int(off[i + 1])]
        if sid < 0:
            assert mine == cts[128 * i:128 * i + 128]
            continue
        ns, u, Y, A = sets[sid]
        want, _ = oracle.rp_create(ns, u, l, Y, A, P, [s], rs[32 * i:32 * i + 32], cts[128 * i:128 * i + 128], [rnd[i]])
        assert mine == want
    ok = backend.verify_range_proofs(sids, ls, proofs, off)
    assert list(ok) == [1] * len(plan)
    ls_bad = list(ls); ls_bad[0] = 3; ls_bad[3] = 1
    ok = backend.verify_range_proofs(sids, ls_bad, proofs, off)
    assert list(ok) == [0, 1, 1, 0, 1, 1, 1]
    from drynx_b200 import DrynxGpuError
    with pytest.raises(DrynxGpuError):
        backend.verify_range_proofs([9], [2], pb[:int(off[1])], off[:2])  # unknown signature set


def test_golden_range_proofs(backend):
    with open(os.path.join(GOLD, 'bn256_golden.json')) as f:
        g = json.load(f)
    for i, v in enumerate(g['range_proof']):
        backend.set_pubkey(bytes.fromhex(v['P']))
        Y, A = backend.init_range_proof_signature([int(v['x'][64 * k:64 * k + 64], 16) for k in range(v['ns'])], v['u'])
        assert Y.tobytes().hex() == v['Y'] and A.tobytes().hex() == v['A']
        backend.set_signatures(20 + i, v['ns'], v['u'], Y, A)
        proofs, off = backend.create_range_proofs([20 + i], {20 + i: v['ns']}, {20 + i: v['u']}, [v['l']], [v['secret']],
                                                  bytes.fromhex(v['r']), bytes.fromhex(v['ct']), [bytes.fromhex(v['rnd'])])
        assert proofs.tobytes().hex() == v['proof']
        assert list(backend.verify_range_proofs([20 + i], [v['l']], proofs, off)) == [1 if v['valid'] else 0]


def test_config2_shape_create_verify_properties(backend, oracle, rng):
    """BASELINE config 2 shape: u=16, l=4, ns=3, mean (2 outputs) + variance (3 outputs) per DP, here 40 DPs
    (200 proofs, 2400 pairing checks). Oracle on a subsample; verify(create) == 1 for all; one flipped byte in
    1% of the proofs flips exactly those bits."""
    ns, u, l = 3, 16, 4
    sk = rng.scalar()
    P = R.g1_to_bytes(R.g1_mul(sk, R.G1_GEN))
    backend.set_pubkey(P)
    nout, ndp = 5, 40
    sets = {}
    for o in range(nout):
        xs = [rng.scalar() for _ in range(ns)]
        Y, A = backend.init_range_proof_signature(xs, u)
        backend.set_signatures(o, ns, u, Y, A)
        sets[o] = (Y.tobytes(), A.tobytes())
    g = np.random.default_rng(2)
    secrets = g.integers(0, 40001, size=ndp * nout).astype(np.int64)
    sids = [i % nout for i in range(ndp * nout)]
    n = len(sids)
    rs = b''.join(rng.scalar().to_bytes(32, 'big') for _ in range(n))
    cts = backend.encrypt_int_vector(secrets, rs).tobytes()
    rnd = [rng.scalars_bytes(l * (3 + ns)) for _ in range(n)]
    proofs, off = backend.create_range_proofs(sids, {o: ns for o in sets}, {o: u for o in sets}, [l] * n, secrets, rs, cts, rnd)
    pb = proofs.tobytes()
    for i in (0, 7, 101, n - 1):
        Y, A = sets[sids[i]]
        want, _ = oracle.rp_create(ns, u, l, Y, A, P, [int(secrets[i])], rs[32 * i:32 * i + 32], cts[128 * i:128 * i + 128], [rnd[i]])
        assert pb[int(off[i]):int(off[i + 1])] == want
    assert list(backend.verify_range_proofs(sids, [l] * n, proofs, off)) == [1] * n
    bad = bytearray(pb)
    flipped = [3, 150]
    size = int(off[1])
    bad[3 * size + 170] ^= 4          # Zr
    bad[150 * size + size - 40] ^= 1  # last Zv
    ok = backend.verify_range_proofs(sids, [l] * n, bytes(bad), off)
    assert [i for i in range(n) if not ok[i]] == flipped


def _outside_g2(seed):
    """Twist points outside the order-n subgroup: generic (composite order), cofactor order, order 13, 13 * generic."""
    T = R.g2_twist_point(seed)
    pts = [T, R.g2_mul_raw(R.n, T), R.g2_mul_raw(R.n * R.TWIST_COFACTOR // 13, T), R.g2_mul_raw(13, T)]
    assert all(Q is not None and R.g2_on_curve(Q) and not R.g2_in_subgroup(Q) for Q in pts)
    return [R.g2_to_bytes(Q) for Q in pts]


@pytest.mark.parametrize('path,lanes,mode', [(1, 0, 0), (2, 4, 1), (2, 8, 1), (2, 4, 2), (2, 8, 2)])
def test_all_verification_paths_match_oracle(backend, oracle, rng, path, lanes, mode):
    """rp_verify_path 1 (one thread per pairing check) and 2 (lane groups running the generated micro-program, four or
    eight program lanes, static or task-queue launch) must give the oracle's accept / reject bits on valid, tampered
    and degenerate proofs -- including V_ij that lie on the twist but outside G2, which kyber's UnmarshalBinary refuses
    (lib/range/range_proof.go:225-236 aborts there; here the proof is rejected)."""
    ns, u, l = 3, 16, 2
    sk, P, Y, A = _setup(backend, oracle, rng, ns, u, set_id=60)
    secrets = [0, 1, 77, 255, 200, 13, 99, 254, 5, 6, 7, 8, 9, 10]
    rs, cts, rnd = _make(backend, oracle, rng, P, ns, u, l, secrets)
    n = len(secrets)
    proofs, off = backend.create_range_proofs([60] * n, {60: ns}, {60: u}, [l] * n, secrets, rs, cts, rnd)
    size = int(off[1])
    blk = 256 + 32 * l
    bad = bytearray(proofs.tobytes())
    bad[1 * size + 170] ^= 1                                  # Zr
    bad[2 * size + blk + 128 * l + 5] ^= 1                    # A_00
    bad[3 * size + blk + 544 * l + 128 * l + 384 + 100] ^= 1  # A_11
    bad[4 * size + blk + 3] ^= 1                              # V_00 leaves the curve
    bad[5 * size + blk:5 * size + blk + 128] = bytes(128)     # V_00 = infinity
    bad[6 * size + blk + 512 * l + 9] ^= 1                    # Zv_00
    outside = _outside_g2(60 + path)
    for k, Q in enumerate(outside):                           # V_00, V_11, V_20, V_01 outside G2
        i, j = [(0, 0), (1, 1), (2, 0), (0, 1)][k]
        pos = (8 + k) * size + blk + i * 544 * l + 128 * j
        bad[pos:pos + 128] = Q
    pos = 12 * size + blk + 544 * l + 128 * l + 64                # A_10: one coordinate >= p (malformed GT element)
    bad[pos:pos + 32] = b'\xff' * 32
    want = oracle.rp_verify(ns, u, l, Y, P, bytes(bad), off, threads=8)
    backend.set_option('rp_verify_path', path)
    backend.set_option('rp_ucode_lanes', lanes)
    backend.set_option('rp_ucode_mode', mode)
    backend.set_option('rp_verify_gt', 2 if mode == 2 else 0)   # the lane-group GT factor kernel on the task-queue runs
    backend.set_option('rp_verify_hdr_early', 1 if lanes == 8 else 0)   # the header check's other place in the stream order
    try:
        ok = backend.verify_range_proofs([60] * n, [l] * n, bytes(bad), off)
    finally:
        backend.set_option('rp_verify_path', 0)
        backend.set_option('rp_ucode_lanes', 0)
        backend.set_option('rp_ucode_mode', 0)
        backend.set_option('rp_verify_gt', 0)
        backend.set_option('rp_verify_hdr_early', 0)
    assert list(ok) == list(want) == [1, 0, 0, 0, 0, 0, 0, 1, 0, 0, 0, 0, 0, 1]


def test_g2_inputs_outside_the_subgroup_are_malformed(backend, oracle, rng):
    """Everywhere a G2 point crosses the ABI it must be a point kyber can unmarshal: on the twist AND of order n."""
    from drynx_b200._lib import DrynxGpuError
    one = (1).to_bytes(32, 'big')
    g1 = R.g1_to_bytes(R.G1_GEN)
    good = R.g2_to_bytes(R.g2_mul(rng.scalar(), R.G2_GEN))
    assert backend.g2_mul(one, good).tobytes() == good
    for Q in _outside_g2(3):
        with pytest.raises(DrynxGpuError):
            backend.g2_mul(one, Q)
        with pytest.raises(DrynxGpuError):
            backend.pair(g1, Q)
    ns, u = 2, 4
    Y, A = oracle.make_sigs([rng.scalar() for _ in range(ns)], u)
    Ab = bytearray(A)
    Ab[128 * 5:128 * 6] = _outside_g2(4)[0]
    with pytest.raises(DrynxGpuError):
        backend.set_signatures(63, ns, u, Y, bytes(Ab))
    backend.set_signatures(63, ns, u, Y, A)


def test_creation_kernels_agree(backend, oracle, rng):
    """The smallest batches are created by the lane-cooperative kernel (eight lanes per V_ij, sixteen per a_ij;
    rp_create_path 2 forces it), mid-size ones by the scratch-based twelve-way split (rp_create_path 1); both must
    give the oracle's bytes."""
    ns, u, l = 2, 8, 3
    sk, P, Y, A = _setup(backend, oracle, rng, ns, u, set_id=61)
    secrets = [0, 1, 7, 8, 63, 64, 100, 511, 300]
    rs, cts, rnd = _make(backend, oracle, rng, P, ns, u, l, secrets)
    n = len(secrets)
    backend.set_option('rp_create_path', 2)
    try:
        a, off = backend.create_range_proofs([61] * n, {61: ns}, {61: u}, [l] * n, secrets, rs, cts, rnd)
        backend.set_option('rp_create_path', 1)
        b, _ = backend.create_range_proofs([61] * n, {61: ns}, {61: u}, [l] * n, secrets, rs, cts, rnd)
        backend.set_option('rp_create_path', 3)   # GT factors on lane groups (the automatic choice up to 36,000 items)
        c3, _ = backend.create_range_proofs([61] * n, {61: ns}, {61: u}, [l] * n, secrets, rs, cts, rnd)
    finally:
        backend.set_option('rp_create_path', 0)
    assert a.tobytes() == b.tobytes() == c3.tobytes()
    want, _ = oracle.rp_create(ns, u, l, Y, A, P, secrets, rs, cts, rnd, threads=8)
    assert a.tobytes() == bytes(want)


@pytest.mark.parametrize('n', [600, 700])
def test_dynamic_segment_kernel(backend, oracle, rng, n):
    """600 and 700 proofs x 12 checks through the task-queue launch of the lane-group kernel (rp_ucode_mode 2: 32 program
    segments dealt from a queue to persistent blocks, slot images handed over through global memory). Every proof
    verifies, the tampered ones (one per kind of field) are rejected, and a sample is checked against the oracle."""
    ns, u, l = 3, 16, 4
    sk, P, Y, A = _setup(backend, oracle, rng, ns, u, set_id=62)
    g = np.random.default_rng(62)
    secrets = [int(x) for x in g.integers(0, u ** l, size=n)]
    rs, cts, rnd = _make(backend, oracle, rng, P, ns, u, l, secrets)
    proofs, off = backend.create_range_proofs([62] * n, {62: ns}, {62: u}, [l] * n, secrets, rs, cts, rnd)
    size = int(off[1])
    blk = 256 + 32 * l
    bad = bytearray(proofs.tobytes())
    flipped = {7: 170, 123: blk + 5, 300: blk + 128 * l + 77, 451: blk + 512 * l + 9, 599: size - 40}
    for p_, o_ in flipped.items():
        bad[p_ * size + o_] ^= 1
    backend.set_option('rp_verify_path', 2)
    backend.set_option('rp_ucode_lanes', 4)
    backend.set_option('rp_ucode_mode', 2)
    try:
        ok = backend.verify_range_proofs([62] * n, [l] * n, bytes(bad), off)
        ok_clean = backend.verify_range_proofs([62] * n, [l] * n, proofs, off)
    finally:
        backend.set_option('rp_verify_path', 0)
        backend.set_option('rp_ucode_lanes', 0)
        backend.set_option('rp_ucode_mode', 0)
    assert list(ok_clean) == [1] * n
    assert [i for i in range(n) if not ok[i]] == sorted(flipped)
    idx = [5, 6, 7, 8, 122, 123, 124, 450, 451, 598, 599]
    sub = b''.join(bytes(bad[i * size:(i + 1) * size]) for i in idx)
    suboff = np.arange(len(idx) + 1, dtype=np.uint64) * size
    want = oracle.rp_verify(ns, u, l, Y, P, sub, suboff, threads=8)
    assert [int(ok[i]) for i in idx] == list(want)


def test_ragged_batch_through_every_path(backend, oracle, rng):
    """A ragged batch large enough for the dynamic lane-group kernel (1,200 proofs over three signature sets with
    different server counts, bases and digit counts, commit-only proofs in between, 7,200 pairing checks): the
    one-thread path, the lane-group path and the oracle (on a sample) agree on every accept / reject bit."""
    sk = rng.scalar()
    P = R.g1_to_bytes(R.g1_mul(sk, R.G1_GEN))
    backend.set_pubkey(P)
    sets = {}
    for sid, (ns, u) in {80: (3, 16), 81: (1, 4), 82: (2, 8)}.items():
        xs = [rng.scalar() for _ in range(ns)]
        Y, A = oracle.make_sigs(xs, u)
        backend.set_signatures(sid, ns, u, Y, A)
        sets[sid] = (ns, u, Y, A)
    g = np.random.default_rng(80)
    n = 1200
    kinds = [(80, 4), (81, 6), (82, 3), (-1, 0), (80, 2)]
    plan = []
    for i in range(n):
        sid, l = kinds[i % len(kinds)]
        top = sets[sid][1] ** l if sid >= 0 else 10
        plan.append((sid, l, int(g.integers(0, top))))
    sids = [p[0] for p in plan]; ls = [p[1] for p in plan]; secrets = [p[2] for p in plan]
    rb = np.frombuffer(g.bytes(32 * n), np.uint8).copy()
    rb[::32] &= 0x7f
    rs = rb.tobytes()
    cts = backend.encrypt_int_vector(secrets, rs).tobytes()
    rnd = []
    for sid, l, s_ in plan:
        if sid < 0:
            rnd.append(b'')
        else:
            k = np.frombuffer(g.bytes(32 * l * (3 + sets[sid][0])), np.uint8).copy()
            k[::32] &= 0x7f
            rnd.append(k.tobytes())
    proofs, off = backend.create_range_proofs(sids, {k: v[0] for k, v in sets.items()}, {k: v[1] for k, v in sets.items()},
                                              ls, secrets, rs, cts, rnd)
    bad = bytearray(proofs.tobytes())
    flipped = [0, 11, 101, 452, 1197]      # proofs of sets 80, 81, 81, 82, 82 (no commit-only ones)
    for i in flipped:
        assert sids[i] >= 0
        bad[int(off[i + 1]) - 40] ^= 1      # inside the last Zv of the proof
    results = {}
    for path, lanes in ((1, 0), (2, 4)):
        backend.set_option('rp_verify_path', path)
        backend.set_option('rp_ucode_lanes', lanes)
        try:
            results[path] = list(backend.verify_range_proofs(sids, ls, bytes(bad), off))
        finally:
            backend.set_option('rp_verify_path', 0)
            backend.set_option('rp_ucode_lanes', 0)
    assert results[1] == results[2]
    assert [i for i in range(n) if not results[2][i]] == flipped
    for i in (0, 1, 2, 11, 452):
        sid, l, s_ = plan[i]
        ns, u, Y, A = sets[sid]
        one = bytes(bad[int(off[i]):int(off[i + 1])])
        want = oracle.rp_verify(ns, u, l, Y, P, one, np.array([0, len(one)], np.uint64))
        assert int(want[0]) == results[2][i]


def test_config5_ten_thousand_proofs(backend, oracle, rng):
    """Config 5 at 10^4 proofs (u=16, l=4, ns=3: 120,000 pairing checks, 69 MB of proofs): every created proof
    verifies, exactly the 1 % that get one byte flipped are rejected, on the automatic path and on both forced paths,
    and a sample of 128 proofs (every tampered one among the first 4,000 plus their neighbours) is judged by the oracle."""
    ns, u, l, n = 3, 16, 4, 10000
    sk = rng.scalar()
    P = R.g1_to_bytes(R.g1_mul(sk, R.G1_GEN))
    backend.set_pubkey(P)
    Y, A = backend.init_range_proof_signature([rng.scalar() for _ in range(ns)], u)
    backend.set_signatures(70, ns, u, Y, A)
    g = np.random.default_rng(55)
    secrets = g.integers(0, u ** l, size=n, dtype=np.int64)

    def scal(k):
        b = np.frombuffer(g.bytes(32 * k), np.uint8).copy()
        b[::32] &= 0x7f
        return b
    rs = scal(n)
    cts = backend.encrypt_int_vector(secrets, rs)
    per = l * (3 + ns) * 32
    rnd_all = scal(n * l * (3 + ns)).tobytes()
    rnd = [rnd_all[per * i:per * (i + 1)] for i in range(n)]
    proofs, off = backend.create_range_proofs([70] * n, {70: ns}, {70: u}, [l] * n, secrets, rs, cts, rnd)
    ok = backend.verify_range_proofs([70] * n, [l] * n, proofs, off)
    assert int(ok.sum()) == n
    size = int(off[1])
    bad = proofs.copy()
    flipped = sorted(set(int(x) for x in g.integers(0, n, size=n // 100)))
    for i in flipped:
        pos = [170, 300, size - 40, 256 + 32 * l + 200][i % 4]
        bad[i * size + pos] ^= 1
    results = {}
    for path in (0, 1, 2):
        backend.set_option('rp_verify_path', path)
        try:
            results[path] = backend.verify_range_proofs([70] * n, [l] * n, bad, off)
        finally:
            backend.set_option('rp_verify_path', 0)
        assert [i for i in range(n) if not results[path][i]] == flipped, path
    sample = []
    for i in flipped:
        if i < 4000:
            sample += [i, i + 1]
    sample = sorted(set(sample + list(range(0, 128))))[:128]
    sub = b''.join(bad[i * size:(i + 1) * size].tobytes() for i in sample)
    want = oracle.rp_verify(ns, u, l, Y.tobytes(), P, sub, np.arange(len(sample) + 1, dtype=np.uint64) * size, threads=16)
    assert [int(results[0][i]) for i in sample] == list(want)
    assert 0 < sum(want) < len(sample)
