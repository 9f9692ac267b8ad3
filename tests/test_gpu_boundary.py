"""-m gpu: behaviour of the C-ABI boundary itself -- the status path of the device-buffer (`_dev`) entry points, the
checks on device-side signature-set ids, and one process driving two GPUs (the shape of a Go process with one context
per device calling from arbitrary OS threads). The reference aborts on malformed input (kyber UnmarshalBinary errors
end in log.Fatal, e.g. lib/range/range_proof.go:225-236); here the caller gets a count or a rejected proof."""
import ctypes
import threading

import numpy as np
import pytest

import bn256_ref as R

pytestmark = pytest.mark.gpu


def _vp(t):
    return ctypes.c_void_p(t.data_ptr())


def _scalars(g, k):
    b = np.frombuffer(g.bytes(32 * k), np.uint8).copy()
    b[::32] &= 0x7f
    return b


def test_dev_calls_report_malformed_points_through_take_bad(backend, oracle, rng):
    import torch
    from drynx_b200 import _lib
    lib, ck = backend.lib, _lib.check
    sk = rng.scalar()
    P = R.g1_to_bytes(R.g1_mul(sk, R.G1_GEN))
    backend.set_pubkey(P)
    backend.set_target_pubkey(R.g1_to_bytes(R.g1_mul(rng.scalar(), R.G1_GEN)))
    g = np.random.default_rng(9)
    n = 64
    m = g.integers(0, 100, size=n, dtype=np.int64)
    cts = backend.encrypt_int_vector(m, _scalars(g, n)).copy()
    assert backend.take_bad() == 0
    bad_rows = [3, 40]
    for i in bad_rows:
        cts[128 * i:128 * i + 64] = np.frombuffer((5).to_bytes(32, 'big') + (7).to_bytes(32, 'big'), np.uint8)  # K off the curve
    d_ct = torch.from_numpy(cts).cuda()
    skb = np.frombuffer(sk.to_bytes(32, 'big'), np.uint8).copy()
    skp = ctypes.c_void_p(skb.ctypes.data)

    # aggregation: a malformed POINT is skipped (it enters its sum as the point at infinity) and counted; the result
    # equals the oracle's sum over the same rows with those points replaced by infinity
    d_out = torch.full((128,), 0xAB, dtype=torch.uint8, device='cuda')
    ck(lib.drx_aggregate_dev(backend.h, n, 1, _vp(d_ct), _vp(d_out)))
    assert backend.take_bad() == len(bad_rows)
    rows = [cts[128 * i:128 * i + 128].tobytes() for i in range(n)]
    rc, want = oracle.aggregate([bytes(64) + rows[i][64:] if i in bad_rows else rows[i] for i in range(n)])
    assert rc == 0 and d_out.cpu().numpy().tobytes() == want
    assert backend.take_bad() == 0                                  # cleared by the previous call

    # decryption: INT64_MIN for the rejected rows; two calls accumulate into one count
    d_dec = torch.zeros(n, dtype=torch.int64, device='cuda')
    ck(lib.drx_decrypt_dev(backend.h, n, skp, _vp(d_ct), _vp(d_dec)))
    ck(lib.drx_decrypt_dev(backend.h, n, skp, _vp(d_ct), _vp(d_dec)))
    assert backend.take_bad() == 2 * len(bad_rows)
    dec = d_dec.cpu().numpy()
    assert [i for i in range(n) if dec[i] == np.iinfo(np.int64).min] == bad_rows
    assert all(dec[i] == m[i] for i in range(n) if i not in bad_rows)

    # element-wise calls: rejected rows come back as zeros, never as whatever the buffer held before
    s = torch.from_numpy(_scalars(g, n)).cuda()
    d_mul = torch.full((n * 128,), 0xCD, dtype=torch.uint8, device='cuda')
    ck(lib.drx_ct_scalar_mul_dev(backend.h, n, _vp(s), _vp(d_ct), _vp(d_mul)))
    assert backend.take_bad() == len(bad_rows)
    mul = d_mul.cpu().numpy().reshape(n, 128)
    assert all(not mul[i].any() for i in bad_rows) and all(mul[i].any() for i in range(n) if i not in bad_rows)
    pi = torch.arange(n, dtype=torch.int32, device='cuda')
    d_sh = torch.full((n * 128,), 0xCD, dtype=torch.uint8, device='cuda')
    ck(lib.drx_shuffle_rerandomize_dev(backend.h, n, 1, _vp(pi), _vp(d_ct), _vp(s), _vp(d_sh)))
    assert backend.take_bad() == len(bad_rows)
    sh = d_sh.cpu().numpy().reshape(n, 128)
    assert all(not sh[i].any() for i in bad_rows)
    d_rb = torch.from_numpy(np.ascontiguousarray(cts.reshape(n, 128)[:, :64]).ravel()).cuda()
    d_cv = torch.full((n * 128,), 0xCD, dtype=torch.uint8, device='cuda')
    d_neg = torch.full((n * 64,), 0xCD, dtype=torch.uint8, device='cuda')
    ck(lib.drx_keyswitch_dev(backend.h, n, skp, _vp(d_rb), _vp(s), _vp(d_cv), _vp(d_neg)))
    assert backend.take_bad() == len(bad_rows)
    cv, neg = d_cv.cpu().numpy().reshape(n, 128), d_neg.cpu().numpy().reshape(n, 64)
    assert all(not cv[i].any() and not neg[i].any() for i in bad_rows)

    # proof.Rep commitments over device buffers: V_i = s_i*K_i + s_i*C_i per ciphertext (two terms); the rejected rows are
    # zeroed and counted, the others equal the oracle's; verification over device buffers accepts exactly the rows that hold
    s2 = torch.from_numpy(np.repeat(s.cpu().numpy().reshape(n, 32), 2, axis=0).ravel()).cuda()
    d_V = torch.full((n * 64,), 0xCD, dtype=torch.uint8, device='cuda')
    ck(lib.drx_rep_commit_dev(backend.h, n, 2, _vp(s2), _vp(d_ct), _vp(d_V)))
    assert backend.take_bad() == len(bad_rows)
    V = d_V.cpu().numpy().reshape(n, 64)
    good = [i for i in range(n) if i not in bad_rows]
    rcV, wantV = oracle.rep_commit(2, b''.join(s2.cpu().numpy().reshape(n, 64)[i].tobytes() for i in good),
                                   b''.join(rows[i] for i in good), threads=8)
    assert rcV == 0 and all(not V[i].any() for i in bad_rows)
    assert b''.join(V[i].tobytes() for i in good) == wantV
    # relations P = x*K + x*C with x = 1: P is V for s = 1; check r*K + r*C + c*P == V' for a made-up transcript
    ones = torch.from_numpy(np.tile(np.frombuffer((1).to_bytes(32, 'big'), np.uint8), 2 * n)).cuda()
    d_P = torch.empty(n * 64, dtype=torch.uint8, device='cuda')
    ck(lib.drx_rep_commit_dev(backend.h, n, 2, _vp(ones), _vp(d_ct), _vp(d_P)))
    assert backend.take_bad() == len(bad_rows)
    chal = torch.from_numpy(_scalars(g, n)).cuda()
    resp = backend.rep_respond(1, s.cpu().numpy(), chal.cpu().numpy(), np.tile(np.frombuffer((1).to_bytes(32, 'big'), np.uint8), n))
    d_resp = torch.from_numpy(np.repeat(resp.reshape(n, 32), 2, axis=0).ravel()).cuda()
    d_ok = torch.full((n,), 7, dtype=torch.uint8, device='cuda')
    ck(lib.drx_rep_verify_dev(backend.h, n, 2, _vp(d_resp), _vp(chal), _vp(d_ct), _vp(d_P), _vp(d_V), _vp(d_ok)))
    backend.synchronize()                                           # a _dev call returns once its work is queued
    ok = d_ok.cpu().numpy()
    assert [i for i in range(n) if ok[i] == 0] == bad_rows and all(ok[i] == 1 for i in good)


def test_rp_dev_calls_validate_shape_and_set_ids(backend, oracle, rng):
    import torch
    from drynx_b200 import _lib
    from drynx_b200._lib import DrynxGpuError
    lib, ck = backend.lib, _lib.check
    ns, u, l, n = 2, 4, 2, 6
    P = R.g1_to_bytes(R.g1_mul(rng.scalar(), R.G1_GEN))
    backend.set_pubkey(P)
    sid = torch.zeros(n, dtype=torch.int32, device='cuda')
    dummy = torch.zeros(n * backend.proof_size(ns, l), dtype=torch.uint8, device='cuda')
    ok = torch.full((n,), 7, dtype=torch.uint8, device='cuda')
    with pytest.raises(DrynxGpuError):          # nothing registered yet on this context
        ck(lib.drx_rp_verify_dev(backend.h, n, _vp(sid), ns, l, _vp(dummy), _vp(ok)))
    Y, A = oracle.make_sigs([rng.scalar() for _ in range(ns)], u)
    backend.set_signatures(5, ns, u, Y, A)
    with pytest.raises(DrynxGpuError):          # no registered set has three servers
        ck(lib.drx_rp_verify_dev(backend.h, n, _vp(sid), 3, l, _vp(dummy), _vp(ok)))
    with pytest.raises(DrynxGpuError):          # 4^(40-1) is not exact in a float64
        ck(lib.drx_rp_verify_dev(backend.h, n, _vp(sid), ns, 40, _vp(dummy), _vp(ok)))
    secrets = [1, 2, 3, 4, 5, 6]
    g = np.random.default_rng(3)
    rs = _scalars(g, n)
    cts = backend.encrypt_int_vector(secrets, rs)
    rnd = [_scalars(g, l * (3 + ns)).tobytes() for _ in secrets]
    proofs, off = backend.create_range_proofs([5] * n, {5: ns}, {5: u}, [l] * n, secrets, rs, cts, rnd)
    d_pr = torch.from_numpy(proofs).cuda()
    # ids: 5 = the registered set; 0 = unregistered slot below it; 6 and 10**6 = beyond the array
    sid = torch.tensor([5, 0, 5, 6, 1000000, 5], dtype=torch.int32, device='cuda')
    for path in (1, 2):
        backend.set_option('rp_verify_path', path)
        try:
            ck(lib.drx_rp_verify_dev(backend.h, n, _vp(sid), ns, l, _vp(d_pr), _vp(ok)))
            backend.synchronize()
        finally:
            backend.set_option('rp_verify_path', 0)
        assert ok.cpu().tolist() == [1, 0, 1, 0, 0, 1], path


def test_two_devices_in_one_process(oracle, rng):
    """One context per GPU in one process, verification launched from two threads at once: the lane-group kernels need
    their shared-memory opt-in on every device they run on."""
    import torch
    if torch.cuda.device_count() < 2:
        pytest.skip('needs two GPUs in one box')
    from drynx_b200 import Backend
    ns, u, l, n = 3, 16, 4, 700
    results, errors = {}, []

    def work(dev):
        try:
            be = Backend(dev)
            r2 = np.random.default_rng(100 + dev)
            sk = int.from_bytes(r2.bytes(31), 'big') + 1
            P = R.g1_to_bytes(R.g1_mul(sk, R.G1_GEN))
            be.set_pubkey(P)
            xs = [int.from_bytes(r2.bytes(31), 'big') + 1 for _ in range(ns)]
            Y, A = be.init_range_proof_signature(xs, u)
            be.set_signatures(0, ns, u, Y, A)
            secrets = [int(x) for x in r2.integers(0, u ** l, size=n)]
            rs = _scalars(r2, n)
            cts = be.encrypt_int_vector(secrets, rs)
            rnd = [_scalars(r2, l * (3 + ns)).tobytes() for _ in secrets]
            proofs, off = be.create_range_proofs([0] * n, {0: ns}, {0: u}, [l] * n, secrets, rs, cts, rnd)
            bad = proofs.copy()
            size = int(off[1])
            bad[17 * size + 170] ^= 1
            ok = None
            for mode in (1, 2):                     # static launch, then the task queue
                be.set_option('rp_ucode_mode', mode)
                got = be.verify_range_proofs([0] * n, [l] * n, bad, off)
                assert ok is None or list(ok) == list(got)
                ok = got
            sub = bad[16 * size:19 * size].tobytes()
            want = oracle.rp_verify(ns, u, l, Y.tobytes(), P, sub, np.arange(4, dtype=np.uint64) * size)
            results[dev] = (list(ok), list(want))
            be.close()
        except Exception as e:  # surfaced below
            errors.append((dev, e))
    threads = [threading.Thread(target=work, args=(d,)) for d in (0, 1)]
    for t in threads:
        t.start()
    for t in threads:
        t.join()
    assert not errors, errors
    for dev in (0, 1):
        ok, want = results[dev]
        assert [i for i in range(n) if not ok[i]] == [17] and want == [1, 0, 1]


def test_empty_inputs_on_every_entry_point(backend, oracle, rng):
    """n = 0 (an empty CipherVector, no proofs, no signatures ...) is a successful call that launches nothing and
    touches no buffer; the reference's loops over empty slices do the same (e.g. lib/range/range_proof.go:171-205
    ranges over the proof list, unlynx CipherVector.Add over the vector)."""
    P = R.g1_to_bytes(R.g1_mul(rng.scalar(), R.G1_GEN))
    backend.set_pubkey(P)
    backend.set_target_pubkey(P)
    launches = backend.launches
    skb = (5).to_bytes(32, 'big')
    assert backend.encrypt_int_vector([], b'').size == 0
    assert backend.encrypt_scalar_vector(b'', b'').size == 0
    assert backend.aggregate([b'', b'', b'']).size == 0                 # three providers, empty vectors
    zero_dps = backend.aggregate([])                                      # no provider at all
    assert zero_dps.size == 0
    assert backend.decrypt_int_vector(skb, b'').size == 0
    assert backend.decrypt_check_zero(skb, b'').size == 0
    assert backend.ct_scalar_mul(b'', b'').size == 0
    assert backend.key_switch(skb, b'', b'').size == 0
    assert backend.shuffle_rerandomize([], [], []) == []
    assert backend.schnorr_verify([], [], []).size == 0
    assert backend.rep_commit(2, b'', b'').size == 0
    assert backend.rep_respond(2, b'', b'', b'').size == 0
    assert backend.rep_verify(2, b'', b'', b'', b'', b'').size == 0
    assert backend.g1_mul(b'', b'').size == 0 and backend.g2_mul(b'', b'').size == 0
    proofs, off = backend.create_range_proofs([], {}, {}, [], [], b'', b'', [])
    assert proofs.size == 0 and list(off) == [0]
    assert backend.verify_range_proofs([], [], b'', off).size == 0
    assert backend.reduce_rows(np.empty((0, 4), np.int64), 3).shape == (0, 5)
    assert backend.launches == launches
    # a provider vector of zero ciphertexts next to the aggregate of one provider: the identity of the fold
    ct = backend.encrypt_int_vector([7, -7], rng.scalars_bytes(2)).tobytes()
    assert backend.aggregate([ct]).tobytes() == ct
