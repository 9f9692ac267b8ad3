"""Sweep of the range-proof verification paths over batch sizes (experiment harness, one GPU).

usage: python tools/ucode_sweep.py [n ...]           # proofs per batch (ns=3, u=16, l=4 -> 12 pairing checks each)
env:   SWEEP_CONFIGS="name:opt=val,opt=val;..."     # options passed to drx_set_option per configuration
Prints one line per (n, configuration): verify ms, pairing-kernel ms, proofs/s, fraction of the INT roof.
"""
import ctypes, os, sys
import numpy as np
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
import torch
from drynx_b200 import Backend, _lib

NS, U, L = 3, 16, 4
DEFAULT = ("lanes:rp_verify_path=2;"
           "thread:rp_verify_path=1")
be = Backend(0); lib, ck = be.lib, _lib.check
g = np.random.default_rng(5)
P = 65000549695646603732796438742359905742825358107623003571877145026864184071783
g1 = bytes(31) + b'\x01' + (P - 2).to_bytes(32, 'big')
be.set_pubkey(be.g1_mul(b'\0' + g.bytes(31), g1))
Y, A = be.init_range_proof_signature([int.from_bytes(g.bytes(31), 'big') + 1 for _ in range(NS)], U)
be.set_signatures(0, NS, U, Y, A)
dev = torch.device('cuda', 0)
st = torch.cuda.current_stream(); be.set_stream(st.cuda_stream or 1)
vp = lambda t: ctypes.c_void_p(t.data_ptr())
peak = float(os.environ.get('INT_PEAK', '8.83e12'))
MAC_PER_PROOF = (3550 + 352 * (NS + L) + NS * L * 19700) * 128


def scal(k):
    b = np.frombuffer(g.bytes(32 * k), np.uint8).copy(); b[::32] &= 0x7f
    return torch.from_numpy(b).to(dev)


def timed(fn, reps):
    fn(); torch.cuda.synchronize()
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    e0.record(st)
    for _ in range(reps): fn()
    e1.record(st); torch.cuda.synchronize()
    return e0.elapsed_time(e1) / reps


ALL_OPTS = ('rp_verify_path', 'rp_ucode_lanes', 'rp_ucode_mode', 'rp_ucode_resident', 'rp_ucode_nseg')
configs = []
for part in os.environ.get('SWEEP_CONFIGS', DEFAULT).split(';'):
    name, _, rest = part.partition(':')
    configs.append((name, dict((kv.split('=')[0], int(kv.split('=')[1])) for kv in rest.split(',') if kv)))

for n in [int(x) for x in (sys.argv[1:] or ['40', '200', '500', '1000', '2000', '4096'])]:
    sid = torch.zeros(n, dtype=torch.int32, device=dev)
    sec = torch.from_numpy(g.integers(0, U ** L, size=n, dtype=np.int64)).to(dev)
    r, rnd = scal(n), scal(n * L * (3 + NS))
    ct = torch.empty(n * 128, dtype=torch.uint8, device=dev)
    psz = be.proof_size(NS, L)
    proofs = torch.empty(n * psz, dtype=torch.uint8, device=dev)
    ok = torch.empty(n, dtype=torch.uint8, device=dev)
    for o in ALL_OPTS: be.set_option(o, 0)
    ck(lib.drx_encrypt_dev(be.h, n, vp(sec), vp(r), vp(ct)))
    ck(lib.drx_rp_create_dev(be.h, n, vp(sid), NS, L, vp(sec), vp(r), vp(ct), vp(rnd), vp(proofs)))
    be.synchronize()
    # tamper every 7th proof (a byte of A_00 / Zv / V), expected reject exactly there
    bad = torch.zeros(n, dtype=torch.bool, device=dev)
    pv = proofs.view(n, psz)
    for k, off in enumerate((256 + 32 * L + 128 * L + 5, 256 + 32 * L + 128 * L + 384 * L + 3, 70)):
        idx = torch.arange(3 + k, n, 21, device=dev)
        pv[idx, off] ^= 1
        bad[idx] = True
    expect = (~bad).to(torch.uint8)
    for name, opts in configs:
        for o in ALL_OPTS: be.set_option(o, opts.get(o, 0))
        reps = 5 if n <= 2000 else 2
        ok.zero_()
        tv = timed(lambda: ck(lib.drx_rp_verify_dev(be.h, n, vp(sid), NS, L, vp(proofs), vp(ok))), reps)
        good = bool((ok == expect).all())
        ck(lib.drx_profile_enable(be.h, 1))
        ck(lib.drx_rp_verify_dev(be.h, n, vp(sid), NS, L, vp(proofs), vp(ok))); be.synchronize()
        parts = []
        for k in ('k_rp_verify_hdr', 'k_rp_verify_items', 'k_rp_verify_prep', 'k_pair_ucode', 'k_rp_verify_gt', 'k_rp_verify_finish'):
            tms, cnt = ctypes.c_double(), ctypes.c_uint64()
            ck(lib.drx_profile_get(be.h, k.encode(), ctypes.byref(tms), ctypes.byref(cnt)))
            if cnt.value: parts.append('%s=%.2f' % (k[2:].replace('rp_verify_', ''), tms.value))
        ck(lib.drx_profile_enable(be.h, 0))
        print('n=%-7d %-14s verify %8.3f ms  %8.1f k proofs/s  frac %.3f  %s  [%s]' % (
            n, name, tv, n / tv, n / tv * 1e3 * MAC_PER_PROOF / peak, 'OK' if good else 'MISMATCH', ' '.join(parts)), flush=True)
