#!/usr/bin/env python3
"""bench.py -- BASELINE.json metric on BASELINE config 2 (mean/variance: 100 DPs x 10k rows, range proofs u=16, l=4).

One "step" = one pass of the hot path for one survey on every rank's shard of data providers:
  rows -> per-DP reduction (mean: [sum, N]; variance: [sum, N, sum of squares]) -> ElGamal encrypt (5 per DP)
       -> range-proof creation (5 per DP, ns = 3 servers) -> range-proof verification -> per-rank aggregation
       -> (N > 1) NCCL all-gather of the per-rank partial aggregates, fold on rank 0 -> decrypt.
`value` is proofs verified per second of WHOLE step time (every phase inside the timed region), device-resident
inputs; `e2e` is the same step through the host-buffer C ABI (H2D/D2H inside). ElGamal encryptions/s and the other
per-kernel rates are reported under `secondary`. `--impl reference` times the CPU restatement (oracle/, all host
threads) on a bounded sample of the same workload.
"""
import argparse
import ctypes
import json
import os
import subprocess
import sys
import threading
import time

import numpy as np

# stdout carries the one JSON line and nothing else: file descriptor 1 is pointed at stderr for the run (NCCL prints its
# "NCCL version ..." banner, and more with NCCL_DEBUG=INFO, straight to fd 1) and the line goes to the saved descriptor
_RESULT_OUT = os.fdopen(os.dup(1), 'w')
os.dup2(2, 1)


def emit(line):
    _RESULT_OUT.write(line + '\n')
    _RESULT_OUT.flush()


ROOT = os.path.dirname(os.path.abspath(__file__))
sys.path.insert(0, ROOT)

NS, U, L = 3, 16, 4
NOUT = 5  # mean (2 outputs) + variance (3 outputs)
METRIC = 'range-proof verifies/sec over the full config-2 step (ElGamal encs/sec in secondary)'
MAC32_PER_M = 128
# SURVEY.md section 8(d) algorithmic costs in Montgomery products (M)
M_ENCRYPT = 824
M_VERIFY_ITEM = 19700                      # per (server, digit): Miller 8,800 + final exp 9,100 + GT fixed 1,728 + 72
M_PAIRING = 17900                          # Miller loop + final exponentiation alone (the cooperative kernel's share)
M_VERIFY_PROOF = 3550 + 352 * (NS + L) + M_VERIFY_ITEM * NS * L
M_CREATE_PROOF = L * (736 + NS * 4458)
M_AGG_CT = 22
M_DECRYPT = 2840
G1_BYTES = bytes(31) + b'\x01' + (65000549695646603732796438742359905742825358107623003571877145026864184071783 - 2).to_bytes(32, 'big')
M_OBFUSCATE = 2 * 2816 + 16               # two 4-bit-window products, two affine outputs sharing one inversion
M_SHUFFLE = 352 + 352 + 11 + 11 + 16       # two comb walks, two mixed additions, two affine outputs
M_REP1 = 2 * 2816 + 16 + 11             # one proof.Rep relation with one term: r*B and c*P by 4-bit windows, one addition, one mixed addition
M_KEYSWITCH = 352 + 352 + 2816 + 16 + 16   # v*B and v*Q by comb, x*rB by 4-bit windows, one addition, two affine outputs


NCU_SUMMARY = 'profiles/r02_ncu_pair_ucode.txt'   # ncu --set full capture of the dominant kernel at the headline size


def ncu_traffic(rel):
    """dram__bytes_read.sum + dram__bytes_write.sum (bytes per launch) out of a committed `ncu --page raw` summary."""
    import re
    scale = {'byte': 1, 'Kbyte': 1e3, 'Mbyte': 1e6, 'Gbyte': 1e9}
    try:
        text = open(os.path.join(ROOT, rel)).read()
        tot = 0.0
        for name in ('dram__bytes_read.sum', 'dram__bytes_write.sum'):
            m = re.search(re.escape(name) + r' \[(\w+)\] = ([0-9.]+)', text)
            tot += float(m.group(2)) * scale[m.group(1)]
        return int(tot)
    except Exception:
        return None


def parse():
    ap = argparse.ArgumentParser()
    ap.add_argument('--gpus', type=int, default=1)
    ap.add_argument('--steps', type=int, default=20)
    ap.add_argument('--warmup', type=int, default=5)
    ap.add_argument('--impl', default='gpu', choices=['gpu', 'reference'])
    ap.add_argument('--ndp', type=int, default=100)
    ap.add_argument('--rows', type=int, default=10000)
    ap.add_argument('--cpu-sample-ndp', type=int, default=0, help='DPs in the CPU sample (0 = about 3 per host core, at most --ndp)')
    ap.add_argument('--no-secondary', action='store_true')
    return ap.parse_args()


class Clocks(threading.Thread):
    """Samples SM clocks and throttle reasons during the timed region: through NVML every 2 ms when nvidia-ml-py is
    there (the timed region of a default run is ~30 ms), else through nvidia-smi (one sample per ~0.25 s)."""
    Q = 'clocks.sm,clocks.max.sm,clocks_event_reasons.hw_slowdown,clocks_event_reasons.hw_thermal_slowdown,' \
        'clocks_event_reasons.sw_thermal_slowdown,clocks_event_reasons.sw_power_cap'
    NAMES = ['hw_slowdown', 'hw_thermal_slowdown', 'sw_thermal_slowdown', 'sw_power_cap']
    NVML_BITS = [0x8, 0x40, 0x20, 0x4]   # nvmlClocksThrottleReason{HwSlowdown, HwThermalSlowdown, SwThermalSlowdown, SwPowerCap}

    def __init__(self, index):
        super().__init__(daemon=True)
        self.index, self.stop_flag, self.rows, self.source = index, False, [], 'nvidia-smi'
        self.nvml = None
        try:
            import pynvml
            pynvml.nvmlInit()
            try:   # CUDA and NVML may number the devices differently: go by UUID when torch exposes it
                import torch
                uuid = 'GPU-' + str(torch.cuda.get_device_properties(index).uuid)
                self.handle = pynvml.nvmlDeviceGetHandleByUUID(uuid.encode())
            except Exception:
                self.handle = pynvml.nvmlDeviceGetHandleByIndex(index)
            self.max_mhz = pynvml.nvmlDeviceGetMaxClockInfo(self.handle, pynvml.NVML_CLOCK_SM)
            self.nvml, self.source = pynvml, 'nvml'
        except Exception:
            self.nvml = None

    def run(self):
        while not self.stop_flag:
            try:
                if self.nvml is not None:
                    mhz = self.nvml.nvmlDeviceGetClockInfo(self.handle, self.nvml.NVML_CLOCK_SM)
                    bits = self.nvml.nvmlDeviceGetCurrentClocksThrottleReasons(self.handle)
                    self.rows.append([str(mhz), str(self.max_mhz)] + ['Active' if bits & b else 'Not Active' for b in self.NVML_BITS])
                    time.sleep(0.002)
                    continue
                out = subprocess.run(['nvidia-smi', '-i', str(self.index), '--query-gpu=' + self.Q, '--format=csv,noheader,nounits'],
                                     capture_output=True, text=True, timeout=5).stdout.strip()
                if out:
                    self.rows.append([x.strip() for x in out.split(',')])
            except Exception:
                pass
            time.sleep(0.2)

    def summary(self):
        if not self.rows:
            return {'sm_mhz': None, 'sm_max_mhz': None, 'reasons': ['nvidia-smi unavailable']}
        sm = sorted(int(r[0]) for r in self.rows if r[0].isdigit())
        reasons = [n for k, n in enumerate(self.NAMES) if any(r[2 + k].lower().startswith('active') for r in self.rows if len(r) > 2 + k)]
        mx = [int(r[1]) for r in self.rows if r[1].isdigit()]
        return {'sm_mhz': sm[len(sm) // 2] if sm else None, 'sm_max_mhz': max(mx) if mx else None, 'reasons': reasons,
                'samples': len(self.rows), 'source': self.source}


def synth(seed, ndp, rows):
    """Config-2 synthetic input: rows U{0,1,2}; every scalar is 32 random bytes with the top bit cleared (< n)."""
    g = np.random.default_rng(seed)
    data = g.integers(0, 3, size=(ndp, rows), dtype=np.int64)
    n = ndp * NOUT

    def scalars(k):
        b = np.frombuffer(g.bytes(32 * k), dtype=np.uint8).copy()
        b[::32] &= 0x7f
        return b
    return dict(rows=data, r=scalars(n), rnd=scalars(n * L * (3 + NS)), set_id=(np.arange(n, dtype=np.int32) % NOUT),
                x=[int.from_bytes(g.bytes(31), 'big') + 1 for _ in range(NOUT * NS)], sk=int.from_bytes(g.bytes(31), 'big') + 1)


def host_reduce(rows):
    """lib/encoding/mean.go:19-24 + variance.go:19-27 on the host: [sum, N, sum, N, sum of squares] per DP."""
    s = rows.sum(axis=1)
    q = (rows * rows).sum(axis=1)
    n = np.full_like(s, rows.shape[1])
    return np.stack([s, n, s, n, q], axis=1).astype(np.int64).ravel()


# ---------------------------------------------------------------------------------------------------- reference arm
def run_reference(args):
    rank = int(os.environ.get('RANK', '0'))
    if rank != 0:
        return
    sys.path.insert(0, os.path.join(ROOT, 'tests'))
    sys.path.insert(0, os.path.join(ROOT, 'oracle'))
    from conftest import Oracle
    orc = Oracle()
    cores = len(os.sched_getaffinity(0)) if hasattr(os, 'sched_getaffinity') else (os.cpu_count() or 1)
    ndp = args.cpu_sample_ndp or min(args.ndp, max(2, 3 * cores))
    d = synth(2, ndp, args.rows)
    g1, _ = orc.generators()
    P = orc.g1_mul(d['sk'].to_bytes(32, 'big'), g1)
    sigs = [orc.make_sigs(d['x'][NS * o:NS * o + NS], U) for o in range(NOUT)]
    n = ndp * NOUT
    per_rnd = L * (3 + NS) * 32

    def step():
        m = host_reduce(d['rows'])
        ct = orc.encrypt(P, m, d['r'], threads=cores)
        nver = 0
        for o in range(NOUT):  # proofs of one output column share a signature set
            idx = [i for i in range(n) if i % NOUT == o]
            rs = b''.join(d['r'][32 * i:32 * i + 32].tobytes() for i in idx)
            cts = b''.join(ct[128 * i:128 * i + 128] for i in idx)
            rnd = [d['rnd'][per_rnd * i:per_rnd * (i + 1)].tobytes() for i in idx]
            proofs, off = orc.rp_create(NS, U, L, sigs[o][0], sigs[o][1], P, [int(m[i]) for i in idx], rs, cts, rnd, threads=cores)
            ok = orc.rp_verify(NS, U, L, sigs[o][0], P, proofs, off, threads=cores)
            assert ok.all()
            nver += len(idx)
        rc, agg = orc.aggregate([ct[128 * NOUT * k:128 * NOUT * (k + 1)] for k in range(ndp)], threads=cores)
        out, _ = orc.decrypt(d['sk'].to_bytes(32, 'big'), agg, threads=1)
        return nver, out

    # services/api.go:50 CreateDecryptionTable(10000): once per client, outside the step
    orc.dlog_reset()
    orc.decrypt(d['sk'].to_bytes(32, 'big'), orc.encrypt(P, [10000], d['r'][:32]), threads=1)

    for _ in range(max(args.warmup, 0)):
        step()
    t0 = time.perf_counter()
    nver = 0
    for _ in range(args.steps):
        k, out = step()
        nver += k
    dt = time.perf_counter() - t0
    val = nver / dt
    # the edwards25519 ElGamal suite, timed outside the step: 4,096 encryptions on all cores
    ed_n = 4096
    ed_P = orc.ed_mul((12345).to_bytes(32, 'little'))
    ed_m = np.arange(ed_n, dtype=np.int64) % 10
    ed_r = np.random.default_rng(7).bytes(32 * ed_n)
    t1 = time.perf_counter()
    orc.ed_encrypt(ed_P, ed_m, ed_r, threads=cores)
    ed_rate = ed_n / (time.perf_counter() - t1)
    # config 3 pre-pass on the host: the numpy restatement on one provider's 10^5 records
    import logreg_ref
    gl = np.random.default_rng(3)
    lx = gl.integers(0, 4, size=(100000, 8)).astype(np.float64)
    ly = gl.integers(0, 2, size=100000)
    t1 = time.perf_counter()
    logreg_ref.coefficients(lx, ly, [0.0] * 8, [100.0] * 8, 2, 1.0)
    lr_rate = 100000 / (time.perf_counter() - t1)
    # key switching: one server's contribution for 2,048 ciphertexts on all cores
    ks_n = 2048
    ks_ct = orc.encrypt(P, np.arange(ks_n, dtype=np.int64) % 10, np.random.default_rng(8).bytes(32 * ks_n), threads=cores)
    ks_rb = b''.join(ks_ct[128 * i:128 * i + 64] for i in range(ks_n))
    ks_v = np.random.default_rng(9).bytes(32 * ks_n)   # Scalar.SetBytes reduces mod n
    t1 = time.perf_counter()
    rc, _, _ = orc.key_switch(P, d['sk'].to_bytes(32, 'big'), ks_rb, ks_v, threads=cores)
    ks_rate = ks_n / (time.perf_counter() - t1)
    assert rc == 0
    t1 = time.perf_counter()
    rc, _ = orc.ct_scalar_mul(ks_v, ks_ct, threads=cores)
    obf_rate = ks_n / (time.perf_counter() - t1)
    assert rc == 0
    sg_x, sg_k = (0x1234567).to_bytes(32, 'big'), (0x777).to_bytes(32, 'big')
    sg_pub = orc.g1_mul(sg_x, g1)
    sg_msgs = [b'proof list %d' % i for i in range(1024)]
    sg_sigs = [orc.schnorr_sign(sg_x, sg_k, sg_pub, m_) for m_ in sg_msgs[:8]] * 128
    sg_msgs = sg_msgs[:8] * 128
    t1 = time.perf_counter()
    sg_ok = orc.schnorr_verify([sg_pub] * 1024, sg_sigs, sg_msgs, threads=cores)
    sg_rate = 1024 / (time.perf_counter() - t1)
    assert bool(sg_ok.all())
    sample = '%d of %d DPs x %d rows (%d proofs/step), oracle/oracle.cpp C++ restatement of the Go path' % (ndp, args.ndp, args.rows, n)
    emit(json.dumps({
        'impl': 'reference', 'metric': METRIC, 'value': val, 'unit': 'proofs/s', 'n_gpus': args.gpus, 'steps': args.steps,
        'warmup': args.warmup, 'ms_per_step': dt / args.steps * 1e3, 'higher_is_better': True, 'scaling': 'weak',
        'vs_baseline': None, 'dtype': 'u64 (4x64-bit Montgomery limbs)', 'data': 'synthetic',
        'config': {'workload': 'config 2: mean+variance, %d DPs x %d rows, u=16 l=4 ns=3 (bounded sample: %d DPs)' % (args.ndp, args.rows, ndp)},
        'cpu_baseline': {'value': val, 'unit': 'proofs/s', 'cores': cores, 'kind': 'port', 'sample': sample,
                         'ed25519_encs_per_s': ed_rate, 'logreg_prepass_records_per_s': lr_rate,
                         'keyswitch_cts_per_s': ks_rate, 'obfuscations_per_s': obf_rate,
                         'schnorr_verifies_per_s': sg_rate},
        'e2e': {'value': val, 'unit': 'proofs/s', 'h2d_bytes_per_step': 0, 'd2h_bytes_per_step': 0},
        'gpu_launches': 0,
    }))


# ---------------------------------------------------------------------------------------------------- GPU arm
def run_gpu(args):
    import torch
    import torch.distributed as dist
    from drynx_b200 import Backend, _lib
    from drynx_b200.sharding import gather_and_fold, gather_accept_bytes, shard_range

    rank = int(os.environ.get('RANK', '0'))
    world = int(os.environ.get('WORLD_SIZE', '1'))
    local = int(os.environ.get('LOCAL_RANK', '0'))
    if not torch.cuda.is_available():
        raise SystemExit('bench.py: no CUDA device; the GPU arm has no CPU fallback (use --impl reference)')
    torch.cuda.set_device(local)
    dev = torch.device('cuda', local)
    if world > 1:
        dist.init_process_group('nccl', device_id=dev)
    be = Backend(local)
    lib = be.lib
    stream = torch.cuda.current_stream()
    # the legacy default stream's handle is 0, which drx_set_stream reads as "own stream": pass cudaStreamLegacy (1)
    be.set_stream(stream.cuda_stream or 1)

    ndp, rows = args.ndp, args.rows
    n = ndp * NOUT
    d = synth(2 + rank, ndp, rows)
    g1 = bytes(31) + b'\x01' + (65000549695646603732796438742359905742825358107623003571877145026864184071783 - 2).to_bytes(32, 'big')
    # one survey key for all ranks
    d0 = synth(2, 1, 1)
    skb = d0['sk'].to_bytes(32, 'big')
    P = be.g1_mul(skb, g1).tobytes()
    be.set_pubkey(P)
    t_setup = time.perf_counter()
    for o in range(NOUT):
        Y, A = be.init_range_proof_signature(d0['x'][NS * o:NS * o + NS], U)
        be.set_signatures(o, NS, U, Y, A)
    be.create_decryption_table(10000)
    be.synchronize()
    t_setup = time.perf_counter() - t_setup

    proof_size = be.proof_size(NS, L)
    T = lambda a: torch.from_numpy(a).to(dev)
    d_rows, d_r, d_rnd, d_sid = T(d['rows']), T(d['r']), T(d['rnd']), T(d['set_id'])
    d_m = torch.empty(n, dtype=torch.int64, device=dev)
    d_ct = torch.empty(n * 128, dtype=torch.uint8, device=dev)
    d_proofs = torch.empty(n * proof_size, dtype=torch.uint8, device=dev)
    d_ok = torch.empty(n, dtype=torch.uint8, device=dev)
    d_agg = torch.empty(NOUT * 128, dtype=torch.uint8, device=dev)
    d_gather = torch.empty(world * NOUT * 128, dtype=torch.uint8, device=dev)
    d_fold = torch.empty(NOUT * 128, dtype=torch.uint8, device=dev)
    d_dec = torch.empty(NOUT, dtype=torch.int64, device=dev)
    d_flush = torch.empty(256 << 20, dtype=torch.uint8, device=dev)  # > 126 MB L2
    vp = lambda t: ctypes.c_void_p(t.data_ptr())
    skbuf = np.frombuffer(skb, np.uint8).copy()
    skp = ctypes.c_void_p(skbuf.ctypes.data)
    ck = _lib.check

    # a second context on a second stream: aggregation, the NCCL gather and decryption only depend on the
    # ciphertexts, so they run next to proof creation / verification (what two goroutines with one context each do)
    be2 = Backend(local)
    be2.set_pubkey(P)
    be2.create_decryption_table(10000)
    stream2 = torch.cuda.Stream(device=dev)
    be2.set_stream(stream2.cuda_stream)
    ev_ct = torch.cuda.Event()

    def fold_dev(gathered):
        ck(lib.drx_aggregate_dev(be2.h, world, NOUT, vp(gathered), vp(d_fold)))
        return d_fold

    def step_dev():
        d_flush.zero_()
        ck(lib.drx_reduce_rows_dev(be.h, ndp, rows, vp(d_rows), 3, vp(d_m)))
        ck(lib.drx_encrypt_dev(be.h, n, vp(d_m), vp(d_r), vp(d_ct)))
        ev_ct.record(stream)
        ck(lib.drx_rp_create_dev(be.h, n, vp(d_sid), NS, L, vp(d_m), vp(d_r), vp(d_ct), vp(d_rnd), vp(d_proofs)))
        ck(lib.drx_rp_verify_dev(be.h, n, vp(d_sid), NS, L, vp(d_proofs), vp(d_ok)))
        with torch.cuda.stream(stream2):
            stream2.wait_event(ev_ct)
            ck(lib.drx_aggregate_dev(be2.h, ndp, NOUT, vp(d_ct), vp(d_agg)))
            # the one exchange of the sharded survey (drynx_b200/sharding.py; tests/test_sharding.py runs it over gloo)
            final = gather_and_fold(d_agg, rank, world, dist, fold_dev, gathered=d_gather)
            if rank == 0:
                ck(lib.drx_decrypt_dev(be2.h, NOUT, skp, vp(final), vp(d_dec)))
        stream.wait_stream(stream2)

    def barrier():
        torch.cuda.synchronize()
        if world > 1:
            dist.barrier()
        torch.cuda.synchronize()

    for _ in range(max(args.warmup, 3)):
        step_dev()
    barrier()
    # correctness of the warm-up result before anything is timed
    assert bool((d_ok == 1).all()), 'a created proof failed verification'
    expect = host_reduce(d['rows']).reshape(ndp, NOUT).sum(axis=0)
    if world > 1:
        tot = torch.from_numpy(expect).to(dev)
        dist.all_reduce(tot)
        expect = tot.cpu().numpy()
    if rank == 0:
        got = d_dec.cpu().numpy()
        want = np.where(np.abs(expect) < 100000, expect, np.iinfo(np.int64).min)
        assert (got == want).all(), (got, expect)

    clocks = Clocks(local)
    clocks.start()
    # inside the timed region only the dominant kernel is bracketed with events (its duration is the roofline's
    # denominator); the other kernels are timed in three more steps after the region
    ck(lib.drx_profile_enable(be.h, 2))
    launches0 = be.launches + be2.launches
    barrier()
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    e0.record(stream)
    for _ in range(args.steps):
        step_dev()
    e1.record(stream)
    barrier()
    ms = e0.elapsed_time(e1)
    launches = be.launches + be2.launches - launches0
    clocks.stop_flag = True
    KERNELS = ('k_reduce_rows', 'k_encrypt', 'k_rp_create_hdr', 'k_rp_create_g2', 'k_rp_create_items', 'k_rp_verify_hdr', 'k_rp_verify_items',
               'k_rp_verify_prep', 'k_pair_ucode', 'k_rp_verify_gt', 'k_rp_verify_finish', 'k_rp_flags', 'k_agg_partial',
               'k_agg_final', 'k_decrypt')

    def read_prof(into, names):
        for k in names:
            tms, cnt = ctypes.c_double(), ctypes.c_uint64()
            ck(lib.drx_profile_get(be.h, k.encode(), ctypes.byref(tms), ctypes.byref(cnt)))
            if not cnt.value:
                ck(lib.drx_profile_get(be2.h, k.encode(), ctypes.byref(tms), ctypes.byref(cnt)))
            if cnt.value:
                into[k] = {'ms_per_launch': tms.value / cnt.value, 'launches': cnt.value}
    prof = {}
    read_prof(prof, ('k_pair_ucode', 'k_rp_verify_items'))     # from the timed region
    ck(lib.drx_profile_enable(be.h, 1))
    ck(lib.drx_profile_enable(be2.h, 1))
    for _ in range(3):
        step_dev()
    barrier()
    others = {}
    read_prof(others, KERNELS)
    for k, v in others.items():
        prof.setdefault(k, v)
    ck(lib.drx_profile_enable(be.h, 0))
    ck(lib.drx_profile_enable(be2.h, 0))
    if world > 1:
        t = torch.tensor([ms], dtype=torch.float64, device=dev)
        dist.all_reduce(t, op=dist.ReduceOp.MAX)
        ms = float(t.item())
    value = world * n * args.steps / (ms / 1e3)

    # ------------------------------------------------------------------ e2e: host buffers through the C ABI
    # Two sets of host buffers: pinned ones (`e2e`, what the bench contract prescribes) and PAGEABLE numpy arrays
    # (`e2e_pageable`: what a cgo caller's Go slices are; they travel through the driver's bounce buffers).
    keep = []

    def pinned(a):
        t = torch.from_numpy(a).pin_memory()
        keep.append(t)
        return t.numpy()

    def host_set(alloc):
        h = {k: alloc(d[k].copy()) for k in ('rows', 'r', 'rnd', 'set_id')}
        h.update(ct=alloc(np.empty(n * 128, np.uint8)), pr=alloc(np.empty(n * proof_size, np.uint8)), ok=alloc(np.empty(n, np.uint8)),
                 ag=alloc(np.empty(NOUT * 128, np.uint8)), m=alloc(np.empty(n, np.int64)))
        h['ptrs'] = (ctypes.c_void_p * ndp)(*[h['ct'].ctypes.data + k * NOUT * 128 for k in range(ndp)])
        return h
    h_dec = np.empty(NOUT, np.int64)
    h_found = np.empty(NOUT, np.uint8)
    h_l = np.full(n, L, np.int64)
    poff = (np.arange(n + 1, dtype=np.uint64) * proof_size)
    roff = (np.arange(n + 1, dtype=np.uint64) * (L * (3 + NS) * 32))
    np_p = lambda a: ctypes.c_void_p(a.ctypes.data)
    be.set_stream(0)
    be2.set_stream(0)
    cur = {}

    def agg_dec():
        h = cur['h']
        ck(lib.drx_aggregate(be2.h, ndp, NOUT, ctypes.cast(h['ptrs'], ctypes.c_void_p), np_p(h['ag'])))
        if rank == 0 and world == 1:
            ck(lib.drx_decrypt(be2.h, NOUT, skp, np_p(h['ag']), 1, np_p(h_dec), np_p(h_found)))

    # the second goroutine: a long-lived worker thread woken once per step (ctypes calls release the GIL)
    go, done_ev, quit_flag, side_err = threading.Event(), threading.Event(), [False], []

    def side_worker():
        while True:
            go.wait()
            go.clear()
            if quit_flag[0]:
                return
            try:
                agg_dec()
            except Exception as e:  # surfaced by the main thread after the join
                side_err.append(e)
            done_ev.set()
    side_thread = threading.Thread(target=side_worker, daemon=True)
    side_thread.start()

    breakdown = {} if os.environ.get('DRX_E2E_BREAKDOWN') else None   # per-call wall time of the e2e leg, to stderr

    def call(name, fn):
        if breakdown is None:
            return fn()
        t = time.perf_counter()
        fn()
        breakdown[name] = breakdown.get(name, 0.0) + time.perf_counter() - t

    def step_e2e():
        # what a Go data provider / root / verifying node would do with two goroutines and one context each:
        # aggregation + decryption only need the ciphertexts and run next to proof creation + verification
        h = cur['h']
        call('reduce_rows', lambda: ck(lib.drx_reduce_rows(be.h, ndp, rows, np_p(h['rows']), 3, np_p(h['m']))))
        call('encrypt', lambda: ck(lib.drx_encrypt(be.h, n, np_p(h['m']), np_p(h['r']), np_p(h['ct']))))
        go.set()
        call('rp_create', lambda: ck(lib.drx_rp_create(be.h, n, np_p(h['set_id']), np_p(h_l), np_p(h['m']), np_p(h['r']), np_p(h['ct']),
                                                       np_p(h['rnd']), np_p(roff), np_p(h['pr']), np_p(poff))))
        call('rp_verify', lambda: ck(lib.drx_rp_verify(be.h, n, np_p(h['set_id']), np_p(h_l), np_p(h['pr']), np_p(poff), np_p(h['ok']))))
        call('join', done_ev.wait)
        done_ev.clear()
        if side_err:
            raise side_err[0]
    pr_bytes = n * proof_size
    h2d = d['rows'].nbytes + (n * 8 + n * 32) + (n * 4 + n * 8 * 2 + n * 32 + n * 128 + d['rnd'].nbytes) + (n * 4 + n * 8 + pr_bytes + n * 4) \
        + n * 128 + (NOUT * 128 + 32)  # encrypt + create + verify + aggregate + decrypt inputs
    d2h = n * 8 + n * 128 + pr_bytes + n + NOUT * 128 + NOUT * 8
    e2e_s = {}
    for kind, alloc in (('pageable', lambda a: a), ('pinned', pinned)):
        cur['h'] = host_set(alloc)
        for _ in range(3):
            step_e2e()
        assert cur['h']['ok'].all() and cur['h']['pr'].tobytes() == d_proofs.cpu().numpy().tobytes(), 'e2e and device-resident legs disagree'
        barrier()
        t0 = time.perf_counter()
        for _ in range(args.steps):
            step_e2e()
        torch.cuda.synchronize()
        dt = time.perf_counter() - t0
        if world > 1:
            t = torch.tensor([dt], dtype=torch.float64, device=dev)
            dist.all_reduce(t, op=dist.ReduceOp.MAX)
            dt = float(t.item())
        e2e_s[kind] = dt
        if breakdown is not None and rank == 0:
            sys.stderr.write('e2e %s, ms per step: %s\n' % (kind, ', '.join('%s %.3f' % (k, v / (args.steps + 3) * 1e3) for k, v in breakdown.items())))
            breakdown.clear()
    e2e_val = world * n * args.steps / e2e_s['pinned']
    quit_flag[0] = True
    go.set()
    be.set_stream(stream.cuda_stream or 1)
    be2.set_stream(stream2.cuda_stream)

    # ------------------------------------------------------------------ roofline of the dominant kernel
    mac_ms, mac_work = be.microbench(0, 20000)
    peak_tmac = mac_work / mac_ms / 1e9
    dom = 'k_pair_ucode' if 'k_pair_ucode' in prof else 'k_rp_verify_items'
    vi = prof.get(dom, {'ms_per_launch': float('nan')})
    per_item = M_PAIRING if dom == 'k_pair_ucode' else M_VERIFY_ITEM
    alg_mac = n * NS * L * per_item * MAC32_PER_M
    achieved = alg_mac / (vi['ms_per_launch'] * 1e-3) / 1e12
    roofline = {'bound': 'int32 IMAD.WIDE pipe', 'kernel': dom, 'achieved': achieved, 'peak': peak_tmac,
                'unit': 'TMAC32/s', 'frac': achieved / peak_tmac,
                # dram__bytes_read.sum + dram__bytes_write.sum of one launch at this size, read from the committed ncu --set full
                # summary (algorithmic bytes: 6,000 x 208 B in, 384 B out = 3.6 MB)
                'traffic': ncu_traffic(NCU_SUMMARY) if dom == 'k_pair_ucode' else None, 'traffic_source': NCU_SUMMARY,
                'peak_source': 'measured live: register-resident IMAD.WIDE.U32.X chains (drx_microbench kind 0); MEASURED_PEAKS.json has no integer figure',
                'algorithmic': '%d proofs x %d (server,digit) pairs x %d M x 128 MAC32 per launch' % (n, NS * L, per_item)}

    secondary = {}
    if not args.no_secondary and rank == 0:
        secondary = secondary_rates(be, lib, torch, dev, stream, ck, vp, peak_tmac, skp)

    # ------------------------------------------------------------------ strong scaling of configs 5 and 4 (every rank)
    if not args.no_secondary:
        strong = strong_scaling(be, lib, torch, dist, dev, stream, ck, vp, rank, world, skp, peak_tmac)
        if rank == 0:
            secondary.update(strong)

    # ------------------------------------------------------------------ CPU baseline (rank 0, N = 1)
    cpu = None
    if rank == 0 and world == 1:
        cpu = cpu_baseline(args)

    if world > 1:
        dist.barrier()
        dist.destroy_process_group()
    if rank != 0:
        return
    out = {
        'metric': METRIC, 'value': value, 'unit': 'proofs/s', 'n_gpus': world, 'steps': args.steps, 'warmup': max(args.warmup, 3),
        'ms_per_step': ms / args.steps, 'higher_is_better': True, 'scaling': 'weak', 'vs_baseline': None,
        'dtype': 'u32 (8x32-bit Montgomery limbs, 256-bit modular integers)', 'data': 'synthetic',
        'config': {'workload': 'config 2: mean+variance, %d DPs x %d rows per GPU, u=16 l=4 ns=3, %d encryptions + %d proofs created + %d verified per GPU per step'
                   % (ndp, rows, n, n, n), 'l2': 'flushed between steps (256 MiB memset inside the timed region)', 'sharding': 'DPs per rank; all-gather of 5x128 B partial aggregates'},
        'clocks': clocks.summary(),
        'e2e': {'value': e2e_val, 'unit': 'proofs/s', 'h2d_bytes_per_step': int(h2d), 'd2h_bytes_per_step': int(d2h),
                'ms_per_step': e2e_s['pinned'] / args.steps * 1e3, 'host_buffers': 'pinned (torch pin_memory)'},
        'e2e_pageable': {'value': world * n * args.steps / e2e_s['pageable'], 'unit': 'proofs/s',
                         'ms_per_step': e2e_s['pageable'] / args.steps * 1e3,
                         'host_buffers': 'pageable numpy arrays (as a cgo caller passes Go slices); same bytes per step'},
        'gpu_launches': int(launches),
        'roofline': roofline,
        'kernels_ms_per_launch': {k: round(v['ms_per_launch'], 4) for k, v in prof.items()},
        'kernels_timing': 'the roofline kernel: CUDA events inside the timed region; the others: three instrumented steps after it',
        'setup_s': round(t_setup, 3),
        'secondary': secondary,
    }
    if cpu is not None:
        out['cpu_baseline'] = cpu
    emit(json.dumps(out))


def secondary_rates(be, lib, torch, dev, stream, ck, vp, peak_tmac, skp):
    """Per-kernel throughputs at sizes that fill the GPU: config 4 (10^6 encryptions, 1000x1000 aggregation,
    decryption) and config 5 (verification sweep point: 20,000 proofs, u=16 l=4 ns=3)."""
    out = {}
    g = np.random.default_rng(4)

    def timed(fn, reps=3):
        fn()
        torch.cuda.synchronize()
        e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        e0.record(stream)
        for _ in range(reps):
            fn()
        e1.record(stream)
        torch.cuda.synchronize()
        return e0.elapsed_time(e1) / reps

    ndp, v = 1000, 1000
    n = ndp * v
    m = torch.from_numpy(g.integers(0, 10, size=n, dtype=np.int64)).to(dev)
    rb = np.frombuffer(g.bytes(32 * n), np.uint8).copy()
    rb[::32] &= 0x7f
    r = torch.from_numpy(rb).to(dev)
    ct = torch.empty(n * 128, dtype=torch.uint8, device=dev)
    agg = torch.empty(v * 128, dtype=torch.uint8, device=dev)
    dec = torch.empty(v, dtype=torch.int64, device=dev)
    ms = timed(lambda: ck(lib.drx_encrypt_dev(be.h, n, vp(m), vp(r), vp(ct))))
    out['elgamal_encs_per_s'] = n / ms * 1e3
    out['encrypt_roofline_frac'] = n * M_ENCRYPT * MAC32_PER_M / (ms * 1e-3) / 1e12 / peak_tmac
    ms = timed(lambda: ck(lib.drx_aggregate_dev(be.h, ndp, v, vp(ct), vp(agg))))
    out['ciphertext_adds_per_s'] = n / ms * 1e3
    out['aggregate_hbm_gbs'] = n * 128 / (ms * 1e-3) / 1e9
    out['aggregate_roofline_frac'] = n * M_AGG_CT * MAC32_PER_M / (ms * 1e-3) / 1e12 / peak_tmac
    nd = 1 << 17
    dec2 = torch.empty(nd, dtype=torch.int64, device=dev)
    ms = timed(lambda: ck(lib.drx_decrypt_dev(be.h, nd, skp, vp(ct), vp(dec2))))
    out['decrypts_per_s'] = nd / ms * 1e3
    out['decrypt_roofline_frac'] = nd * M_DECRYPT * MAC32_PER_M / (ms * 1e-3) / 1e12 / peak_tmac
    # key switching: one server's contribution for 2^17 ciphertexts, K components taken from ct
    be.set_target_pubkey(be.g1_mul((0x7654321).to_bytes(32, 'big'), G1_BYTES))
    rbk = ct[:nd * 128].view(nd, 128)[:, :64].contiguous()
    cv = torch.empty(nd * 128, dtype=torch.uint8, device=dev)
    ms = timed(lambda: ck(lib.drx_keyswitch_dev(be.h, nd, skp, vp(rbk), vp(r), vp(cv), None)))
    out['keyswitch_cts_per_s'] = nd / ms * 1e3
    out['keyswitch_roofline_frac'] = nd * M_KEYSWITCH * MAC32_PER_M / (ms * 1e-3) / 1e12 / peak_tmac
    # obfuscation (MulCipherTextbyScalar) over the same 2^17 ciphertexts
    obf = torch.empty(nd * 128, dtype=torch.uint8, device=dev)
    ms = timed(lambda: ck(lib.drx_ct_scalar_mul_dev(be.h, nd, vp(r), vp(ct), vp(obf))))
    out['obfuscations_per_s'] = nd / ms * 1e3
    out['obfuscation_roofline_frac'] = nd * M_OBFUSCATE * MAC32_PER_M / (ms * 1e-3) / 1e12 / peak_tmac
    # obfuscation list proof over the same ciphertexts (lib/obfuscation/obfuscation_proof.go): commitments V = v*c, responses
    # r = v - ch*s, and the verifier's 2^18 relations r*c_k + ch*co_k == V_k; the transcript hash is the caller's, a stand-in here
    v_bl, ch = r.view(-1, 32).roll(1, 0).contiguous().view(-1), r.view(-1, 32).roll(2, 0).contiguous().view(-1)
    V = torch.empty(nd * 128, dtype=torch.uint8, device=dev)
    ck(lib.drx_ct_scalar_mul_dev(be.h, nd, vp(v_bl), vp(ct), vp(V)))
    resp_h = be.rep_respond(1, v_bl[:nd * 32].cpu().numpy(), ch[:nd * 32].cpu().numpy(), r[:nd * 32].cpu().numpy())
    resp2 = torch.from_numpy(resp_h).to(dev).view(nd, 32).repeat_interleave(2, dim=0).contiguous()
    ch2 = ch[:nd * 32].view(nd, 32).repeat_interleave(2, dim=0).contiguous()
    okr = torch.zeros(2 * nd, dtype=torch.uint8, device=dev)
    ms = timed(lambda: ck(lib.drx_rep_verify_dev(be.h, 2 * nd, 1, vp(resp2), vp(ch2), vp(ct), vp(obf), vp(V), vp(okr))))
    be.synchronize()
    assert bool((okr == 1).all()), 'an obfuscation proof relation failed'
    out['obfuscation_proofs_verified_per_s'] = nd / ms * 1e3
    out['obfuscation_proof_roofline_frac'] = 2 * nd * M_REP1 * MAC32_PER_M / (ms * 1e-3) / 1e12 / peak_tmac
    # shuffling (ShuffleSequence re-randomisation) of the same ciphertexts as 2^10 lines of 2^7
    perm = torch.from_numpy(np.random.default_rng(6).permutation(1 << 10).astype(np.int32)).to(dev)
    ms = timed(lambda: ck(lib.drx_shuffle_rerandomize_dev(be.h, 1 << 10, 1 << 7, vp(perm), vp(ct), vp(r), vp(obf))))
    out['shuffle_rerandomizations_per_s'] = nd / ms * 1e3
    out['shuffle_roofline_frac'] = nd * M_SHUFFLE * MAC32_PER_M / (ms * 1e-3) / 1e12 / peak_tmac
    # Schnorr proof-request signatures (host buffers, digests given): 2^14 signatures by one key
    import hashlib
    ns_ = 1 << 14
    sig_x = 0x1234567
    sig_pub = be.g1_mul(sig_x.to_bytes(32, 'big'), G1_BYTES)
    sig_R = be.g1_mul((0x777).to_bytes(32, 'big'), G1_BYTES)
    n_ord = 65000549695646603732796438742359905742570406053903786389881062969044166799969
    sigs, digs = bytearray(), bytearray()
    for i in range(64):       # 64 distinct messages, tiled: the kernel does not know
        dgst = hashlib.sha512(bytes(sig_R) + bytes(sig_pub) + b'proof list %d' % i).digest()
        sigs += bytes(sig_R) + ((0x777 + sig_x * (int.from_bytes(dgst, 'big') % n_ord)) % n_ord).to_bytes(32, 'big')
        digs += dgst
    pubs = np.frombuffer(bytes(sig_pub) * ns_, np.uint8).copy()
    sg = np.frombuffer(bytes(sigs) * (ns_ // 64), np.uint8).copy()
    dg = np.frombuffer(bytes(digs) * (ns_ // 64), np.uint8).copy()
    okb = np.zeros(ns_, np.uint8)
    cp = lambda a: ctypes.c_void_p(a.ctypes.data)
    t0 = time.perf_counter()
    ck(lib.drx_schnorr_verify(be.h, ns_, cp(pubs), cp(sg), cp(dg), cp(okb)))
    t0 = time.perf_counter()
    ck(lib.drx_schnorr_verify(be.h, ns_, cp(pubs), cp(sg), cp(dg), cp(okb)))
    dt = time.perf_counter() - t0
    assert bool(okb.all())
    out['schnorr_verifies_per_s_host_buffers'] = ns_ / dt
    # verification sweep point
    npf = 20000
    sid = torch.from_numpy((np.arange(npf, dtype=np.int32) % NOUT)).to(dev)
    sec = torch.from_numpy(g.integers(0, 16 ** L, size=npf, dtype=np.int64)).to(dev)
    rb2 = np.frombuffer(g.bytes(32 * npf * (1 + L * (3 + NS))), np.uint8).copy()
    rb2[::32] &= 0x7f
    r2 = torch.from_numpy(rb2[:32 * npf]).to(dev)
    rnd = torch.from_numpy(rb2[32 * npf:]).to(dev)
    ct2 = torch.empty(npf * 128, dtype=torch.uint8, device=dev)
    size = be.proof_size(NS, L)
    proofs = torch.empty(npf * size, dtype=torch.uint8, device=dev)
    ok = torch.empty(npf, dtype=torch.uint8, device=dev)
    ck(lib.drx_encrypt_dev(be.h, npf, vp(sec), vp(r2), vp(ct2)))
    ms = timed(lambda: ck(lib.drx_rp_create_dev(be.h, npf, vp(sid), NS, L, vp(sec), vp(r2), vp(ct2), vp(rnd), vp(proofs))), reps=2)
    out['proofs_created_per_s'] = npf / ms * 1e3
    out['create_roofline_frac'] = npf * M_CREATE_PROOF * MAC32_PER_M / (ms * 1e-3) / 1e12 / peak_tmac
    ms = timed(lambda: ck(lib.drx_rp_verify_dev(be.h, npf, vp(sid), NS, L, vp(proofs), vp(ok))), reps=2)
    assert bool((ok == 1).all())
    out['proofs_verified_per_s_20k_batch'] = npf / ms * 1e3
    out['verify_roofline_frac_20k_batch'] = npf * M_VERIFY_PROOF * MAC32_PER_M / (ms * 1e-3) / 1e12 / peak_tmac
    out.update(secondary_rates_ed(be, lib, torch, dev, timed, ck, vp, peak_tmac, g))
    return out


def strong_scaling(be, lib, torch, dist, dev, stream, ck, vp, rank, world, skp, peak_tmac):
    """FIXED total work split over the ranks (the headline step is weak scaling: every rank brings its own providers).
    config 5: 10^5 range proofs (ns=3, u=16, l=4, 1 % tampered) split by proof index, accept bytes gathered over NCCL and
    compared with the expected vector; config 4: 1,000 providers x 1,000 bins split by provider, local fold, all-gather,
    root fold, decryption. With N > 1 rank 0 afterwards runs the whole job alone, so each line carries its own
    efficiency t_1 / (N t_N) measured in the same process on the same GPU."""
    from drynx_b200.sharding import gather_and_fold, gather_accept_bytes, shard_range
    out = {}

    def timed(fn, reps, everyone=True):
        fn()
        torch.cuda.synchronize()
        if everyone and world > 1:
            dist.barrier()
        e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        e0.record(stream)
        for _ in range(reps):
            fn()
        e1.record(stream)
        torch.cuda.synchronize()
        ms = e0.elapsed_time(e1) / reps
        if everyone and world > 1:
            t = torch.tensor([ms], dtype=torch.float64, device=dev)
            dist.all_reduce(t, op=dist.ReduceOp.MAX)
            ms = float(t.item())
        return ms

    # ---- config 5
    NP = 100000
    size = be.proof_size(NS, L)
    tampered = lambda idx: idx % 100 == 7

    def make_proofs(lo, hi):
        k = hi - lo
        g = np.random.default_rng([5, lo])
        rb = np.frombuffer(g.bytes(32 * k * (1 + L * (3 + NS))), np.uint8).copy()
        rb[::32] &= 0x7f
        sid = torch.from_numpy(((np.arange(lo, hi)) % NOUT).astype(np.int32)).to(dev)
        sec = torch.from_numpy(g.integers(0, U ** L, size=k, dtype=np.int64)).to(dev)
        r = torch.from_numpy(rb[:32 * k]).to(dev)
        rnd = torch.from_numpy(rb[32 * k:]).to(dev)
        ct = torch.empty(k * 128, dtype=torch.uint8, device=dev)
        proofs = torch.empty(k * size, dtype=torch.uint8, device=dev)
        ck(lib.drx_encrypt_dev(be.h, k, vp(sec), vp(r), vp(ct)))
        ck(lib.drx_rp_create_dev(be.h, k, vp(sid), NS, L, vp(sec), vp(r), vp(ct), vp(rnd), vp(proofs)))
        bad = torch.from_numpy(np.nonzero(tampered(np.arange(lo, hi)))[0]).to(dev)
        proofs.view(k, size)[bad, 170] ^= 1          # a byte of Zr
        return sid, proofs

    def verify_job(lo, hi, everyone):
        sid, proofs = make_proofs(lo, hi)
        ok = torch.empty(hi - lo, dtype=torch.uint8, device=dev)
        res = {}

        def run():
            ck(lib.drx_rp_verify_dev(be.h, hi - lo, vp(sid), NS, L, vp(proofs), vp(ok)))
            res['all'] = gather_accept_bytes(ok, NP, rank, world, dist) if everyone else ok
        ms = timed(run, 2, everyone)
        want = torch.from_numpy((~tampered(np.arange(0, NP) if everyone else np.arange(lo, hi))).astype(np.uint8)).to(dev)
        assert torch.equal(res['all'], want), 'config 5: accept bytes differ from the expected vector'
        return ms
    lo, hi = shard_range(NP, rank, world)
    tN = verify_job(lo, hi, True)
    out['verify_1e5_strong_ms'] = tN
    out['verify_1e5_strong_proofs_per_s'] = NP / tN * 1e3
    out['verify_1e5_strong_roofline_frac'] = NP * M_VERIFY_PROOF * MAC32_PER_M / (tN * 1e-3) / 1e12 / (peak_tmac * world)
    if world > 1:
        t1 = verify_job(0, NP, False) if rank == 0 else 0.0
        out['verify_1e5_one_gpu_ms'] = t1
        out['verify_1e5_strong_eff'] = t1 / (world * tN) if rank == 0 else None
    torch.cuda.empty_cache()

    # ---- config 4
    NDP, V = 1000, 1000
    g = np.random.default_rng(4)                       # the same stream on every rank: ranks slice their block out of it
    counts_all = g.integers(0, 10, size=(NDP, V), dtype=np.int64)
    want = counts_all.sum(axis=0)

    def freq_job(lo, hi, everyone):
        mine = hi - lo
        counts = torch.from_numpy(counts_all[lo:hi].copy()).to(dev).reshape(-1)
        rb = np.frombuffer(np.random.default_rng([44, lo]).bytes(32 * mine * V), np.uint8).copy()
        rb[::32] &= 0x7f
        r = torch.from_numpy(rb).to(dev)
        ct = torch.empty(mine * V * 128, dtype=torch.uint8, device=dev)
        partial = torch.empty(V * 128, dtype=torch.uint8, device=dev)
        total = torch.empty(V * 128, dtype=torch.uint8, device=dev)
        dec = torch.empty(V, dtype=torch.int64, device=dev)
        w = world if everyone else 1

        def fold(gathered):
            ck(lib.drx_aggregate_dev(be.h, w, V, vp(gathered), vp(total)))
            return total

        res = {}

        def run():   # what BASELINE config 4 names: batched encryption + root aggregation (local fold, gather, root fold)
            ck(lib.drx_encrypt_dev(be.h, mine * V, vp(counts), vp(r), vp(ct)))
            ck(lib.drx_aggregate_dev(be.h, mine, V, vp(ct), vp(partial)))
            res['final'] = gather_and_fold(partial, rank, w, dist, fold)
        ms = timed(run, 3, everyone)
        if rank == 0:   # outside the timed region: the querier's decryption, as the check that the aggregate is right
            ck(lib.drx_decrypt_dev(be.h, V, skp, vp(res['final']), vp(dec)))
            torch.cuda.synchronize()
            assert (dec.cpu().numpy() == want).all(), 'config 4: the decrypted aggregate is not the global histogram'
        return ms
    lo, hi = shard_range(NDP, rank, world)
    tN = freq_job(lo, hi, True)
    out['config4_strong_ms'] = tN
    out['config4_strong_encs_per_s'] = NDP * V / tN * 1e3
    if world > 1:
        t1 = freq_job(0, NDP, False) if rank == 0 else 0.0
        out['config4_one_gpu_ms'] = t1
        out['config4_strong_eff'] = t1 / (world * tN) if rank == 0 else None
    torch.cuda.empty_cache()
    if world > 1:
        dist.barrier()
    return out


# edwards25519 suite: algorithmic field products (72 MAC32 each: 64 + 8 for the 38-fold) per unit
MAC32_PER_M_ED = 72
M_ED_ENCRYPT = 72 * 7 + 8 + 265      # 72 comb additions, two serialisations sharing one inversion (a^(p-2): 265)
M_ED_AGG_CT = 2 * (276 + 9)          # two point decodings (square root: 252 squarings + 24 products) + two additions
M_ED_DECRYPT = 2 * 276 + 252 * 8 + 78 * 9 + 9 + 265 + 2   # decodings, 4-bit window multiplication, one serialisation


def secondary_rates_ed(be, lib, torch, dev, timed, ck, vp, peak_tmac, g):
    out = {}
    sk = (987654321).to_bytes(32, 'little')
    base = bytes([0x58] + [0x66] * 31)
    be.ed_set_pubkey(be.ed_point_mul(sk, base))
    ndp, v = 1000, 1000
    n = ndp * v
    m = torch.from_numpy(g.integers(0, 10, size=n, dtype=np.int64)).to(dev)
    r = torch.from_numpy(np.frombuffer(g.bytes(32 * n), np.uint8).copy()).to(dev)
    ct = torch.empty(n * 64, dtype=torch.uint8, device=dev)
    agg = torch.empty(v * 64, dtype=torch.uint8, device=dev)
    ms = timed(lambda: ck(lib.drx_ed_encrypt_dev(be.h, n, vp(m), vp(r), vp(ct))))
    out['ed25519_encs_per_s'] = n / ms * 1e3
    out['ed25519_encrypt_roofline_frac'] = n * M_ED_ENCRYPT * MAC32_PER_M_ED / (ms * 1e-3) / 1e12 / peak_tmac
    ms = timed(lambda: ck(lib.drx_ed_aggregate_dev(be.h, ndp, v, vp(ct), vp(agg))))
    out['ed25519_ciphertext_adds_per_s'] = n / ms * 1e3
    out['ed25519_aggregate_roofline_frac'] = n * M_ED_AGG_CT * MAC32_PER_M_ED / (ms * 1e-3) / 1e12 / peak_tmac
    nd = 1 << 17
    dec = torch.empty(nd, dtype=torch.int64, device=dev)
    skp = ctypes.c_char_p(sk)
    ms = timed(lambda: ck(lib.drx_ed_decrypt_dev(be.h, nd, skp, vp(ct), vp(dec))))
    out['ed25519_decrypts_per_s'] = nd / ms * 1e3
    out['ed25519_decrypt_roofline_frac'] = nd * M_ED_DECRYPT * MAC32_PER_M_ED / (ms * 1e-3) / 1e12 / peak_tmac
    # config 3: the float64 pre-pass of logistic regression, 10 providers x 10^5 records x 8 features, k = 2
    nprov, nrec, dfeat = 10, 100000, 8
    xr = torch.from_numpy(g.integers(0, 4, size=(nprov * nrec, dfeat)).astype(np.float64)).to(dev)
    yr = torch.from_numpy(g.integers(0, 2, size=nprov * nrec).astype(np.int64)).to(dev)
    offs = torch.from_numpy((np.arange(nprov + 1, dtype=np.uint64) * nrec).astype(np.int64)).to(dev)
    mean_t = torch.zeros(dfeat, dtype=torch.float64, device=dev)
    sd_t = torch.full((dfeat,), 100.0, dtype=torch.float64, device=dev)
    coef = torch.empty(nprov * 90, dtype=torch.int64, device=dev)
    ms = timed(lambda: ck(lib.drx_logreg_coefficients_dev(be.h, nprov, nprov * nrec, vp(offs), dfeat, 2, vp(xr), vp(yr), vp(mean_t), vp(sd_t),
                                                          ctypes.c_double(1.0), vp(coef))))
    out['logreg_prepass_records_per_s'] = nprov * nrec / ms * 1e3
    out['logreg_prepass_hbm_gbs'] = nprov * nrec * (dfeat + 1) * 8 / (ms * 1e-3) / 1e9
    # the aggregate of 1,000 providers' 0..9 counts decrypts to the column sums
    be.synchronize()
    want = m.view(ndp, v).sum(dim=0).cpu().numpy()
    got = be.ed_decrypt_int_vector(sk, agg.cpu().numpy())
    assert (got == want).all(), 'edwards25519 aggregate does not decrypt to the column sums'
    return out


def cpu_baseline(args):
    """Times `bench.py --impl reference` logic in-process on a bounded sample (rank 0, N = 1 only)."""
    p = subprocess.run([sys.executable, os.path.abspath(__file__), '--impl', 'reference', '--steps', '2', '--warmup', '1',
                        '--ndp', str(args.ndp), '--rows', str(args.rows), '--cpu-sample-ndp', str(args.cpu_sample_ndp)],
                       capture_output=True, text=True)
    try:
        return json.loads(p.stdout.strip().splitlines()[-1])['cpu_baseline']
    except Exception:
        return {'value': None, 'unit': 'proofs/s', 'cores': os.cpu_count(), 'kind': 'port', 'sample': 'failed: ' + p.stderr[-200:]}


if __name__ == '__main__':
    a = parse()
    if a.impl == 'reference':
        run_reference(a)
    else:
        run_gpu(a)
