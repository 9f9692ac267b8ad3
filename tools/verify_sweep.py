"""BASELINE config 5: range-proof verification sweep. N proofs are created on the device,
1 % get one byte flipped, and verification is timed with CUDA events on device-resident buffers; the accept vector
must reject exactly the flipped proofs. Writes one JSON object per (ns, l, N) to stdout and, with --out, to a file.

Usage: python tools/verify_sweep.py [--out profiles/r01_verify_sweep.json] [--max 1000000]"""
import argparse
import ctypes
import json
import os
import sys

import numpy as np

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
import torch  # noqa: E402
from drynx_b200 import Backend, _lib  # noqa: E402

ap = argparse.ArgumentParser()
ap.add_argument('--out', default='')
ap.add_argument('--max', type=int, default=1000000)
args = ap.parse_args()

U = 16
M_ITEM, MAC = 19700, 128  # products per (server, digit) check, MAC32 per product
be = Backend(0)
lib, ck = be.lib, _lib.check
g = np.random.default_rng(5)
P_ORDER = 65000549695646603732796438742359905742825358107623003571877145026864184071783
g1 = bytes(31) + b'\x01' + (P_ORDER - 2).to_bytes(32, 'big')
be.set_pubkey(be.g1_mul(b'\0' + g.bytes(31), g1))
dev = torch.device('cuda', 0)
st = torch.cuda.current_stream()
be.set_stream(st.cuda_stream or 1)
vp = lambda t: ctypes.c_void_p(t.data_ptr())
mac_ms, mac_work = be.microbench(0, 20000)  # register-resident IMAD.WIDE chains: the integer-pipe peak, measured live
peak_tmac = mac_work / mac_ms / 1e9


def scal(k):
    out = torch.empty(32 * k, dtype=torch.uint8, device=dev)
    step = 1 << 24
    for o in range(0, k, step):
        c = min(step, k - o)
        b = np.frombuffer(g.bytes(32 * c), np.uint8).copy()
        b[::32] &= 0x7f
        out[32 * o:32 * (o + c)] = torch.from_numpy(b).to(dev)
    return out


results = []
set_id = 0
for ns in (3, 1):
    Y, A = be.init_range_proof_signature([int.from_bytes(g.bytes(31), 'big') + 1 for _ in range(ns)], U)
    be.set_signatures(set_id, ns, U, Y, A)
    for l in (4, 2, 8):
        sizes = [1000, 10000, 100000, 1000000] if (ns, l) == (3, 4) else [1000, 10000, 100000]
        for n in sizes:
            if n > args.max:
                continue
            size = be.proof_size(ns, l)
            sid = torch.full((n,), set_id, dtype=torch.int32, device=dev)
            sec = torch.from_numpy(g.integers(0, U ** l, size=n, dtype=np.int64)).to(dev)
            r, rnd = scal(n), scal(n * l * (3 + ns))
            ct = torch.empty(n * 128, dtype=torch.uint8, device=dev)
            proofs = torch.empty(n * size, dtype=torch.uint8, device=dev)
            ok = torch.empty(n, dtype=torch.uint8, device=dev)
            ck(lib.drx_encrypt_dev(be.h, n, vp(sec), vp(r), vp(ct)))
            e0, e1, e2 = (torch.cuda.Event(enable_timing=True) for _ in range(3))
            e0.record(st)
            ck(lib.drx_rp_create_dev(be.h, n, vp(sid), ns, l, vp(sec), vp(r), vp(ct), vp(rnd), vp(proofs)))
            e1.record(st)
            # 1 % tampered: one bit of Zr (offset 170) or of the last Zv (size - 40), alternating
            nbad = max(1, n // 100)
            idx = torch.from_numpy(np.sort(g.choice(n, size=nbad, replace=False))).to(dev)
            offs = torch.where(torch.arange(nbad, device=dev) % 2 == 0, 170, size - 40)
            pos = idx * size + offs
            proofs[pos] ^= 1
            ck(lib.drx_rp_verify_dev(be.h, n, vp(sid), ns, l, vp(proofs), vp(ok)))  # warm-up (tables, arena)
            torch.cuda.synchronize()
            reps = 3 if n <= 100000 else 1
            e1.record(st)
            for _ in range(reps):
                ck(lib.drx_rp_verify_dev(be.h, n, vp(sid), ns, l, vp(proofs), vp(ok)))
            e2.record(st)
            torch.cuda.synchronize()
            ms = e1.elapsed_time(e2) / reps
            rejected = torch.nonzero(ok == 0).flatten()
            assert rejected.numel() == nbad and bool((rejected == idx).all()), 'accept vector does not match the tampered set'
            rate = n / ms * 1e3
            frac = rate * ns * l * M_ITEM * MAC / 1e12 / peak_tmac
            row = {'ns': ns, 'u': U, 'l': l, 'proofs': n, 'pairing_checks': n * ns * l, 'verify_ms': round(ms, 3),
                   'proofs_per_s': round(rate, 1), 'int_roofline_frac': round(frac, 4), 'rejected': int(nbad),
                   'proof_bytes': size}
            results.append(row)
            print(json.dumps(row), flush=True)
            del sid, sec, r, rnd, ct, proofs, ok
            torch.cuda.empty_cache()
    set_id += 1
if args.out:
    with open(os.path.join(ROOT, args.out), 'w') as f:
        json.dump({'peak_tmac32_per_s': peak_tmac, 'note': 'BASELINE config 5 on one B200; device-resident proofs; 1 % tampered, '
                   'accept vector checked; roofline fraction uses SURVEY 8(d): 19,700 products x 128 MAC32 per pairing check',
                   'rows': results}, f, indent=1)
