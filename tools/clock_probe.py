"""Experiment: loops proof verification at a given batch size for ~2 s while nvidia-smi samples SM clock and power."""
import ctypes, os, subprocess, sys, time
import numpy as np
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
import torch
from drynx_b200 import Backend, _lib
NS, U, L = 3, 16, 4
be = Backend(0); lib, ck = be.lib, _lib.check
g = np.random.default_rng(5)
g1 = bytes(31) + b'\x01' + (65000549695646603732796438742359905742825358107623003571877145026864184071783 - 2).to_bytes(32, 'big')
be.set_pubkey(be.g1_mul(b'\0' + g.bytes(31), g1))
Y, A = be.init_range_proof_signature([int.from_bytes(g.bytes(31), 'big') + 1 for _ in range(NS)], U)
be.set_signatures(0, NS, U, Y, A)
be.set_option('rp_verify_path', 2); be.set_option('rp_ucode_lanes', 4)
dev = torch.device('cuda', 0)
st = torch.cuda.current_stream(); be.set_stream(st.cuda_stream or 1)
vp = lambda t: ctypes.c_void_p(t.data_ptr())
def scal(k):
    b = np.frombuffer(g.bytes(32 * k), np.uint8).copy(); b[::32] &= 0x7f
    return torch.from_numpy(b).to(dev)
for n in [int(x) for x in sys.argv[1:]]:
    sid = torch.zeros(n, dtype=torch.int32, device=dev)
    sec = torch.from_numpy(g.integers(0, U ** L, size=n, dtype=np.int64)).to(dev)
    r, rnd = scal(n), scal(n * L * (3 + NS))
    ct = torch.empty(n * 128, dtype=torch.uint8, device=dev)
    proofs = torch.empty(n * be.proof_size(NS, L), dtype=torch.uint8, device=dev)
    ok = torch.empty(n, dtype=torch.uint8, device=dev)
    ck(lib.drx_encrypt_dev(be.h, n, vp(sec), vp(r), vp(ct)))
    ck(lib.drx_rp_create_dev(be.h, n, vp(sid), NS, L, vp(sec), vp(r), vp(ct), vp(rnd), vp(proofs)))
    torch.cuda.synchronize()
    p = subprocess.Popen(['nvidia-smi', '--query-gpu=clocks.sm,power.draw,clocks_throttle_reasons.active', '--format=csv,noheader', '-lms', '50'],
                         stdout=subprocess.PIPE, text=True)
    time.sleep(0.3)
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    e0.record(st)
    reps = 300
    for _ in range(reps):
        ck(lib.drx_rp_verify_dev(be.h, n, vp(sid), NS, L, vp(proofs), vp(ok)))
    e1.record(st); torch.cuda.synchronize()
    time.sleep(0.1)
    p.terminate()
    rows = [l.strip() for l in p.stdout.read().splitlines() if l.strip()]
    print('n=%d verify %.2f ms/call; samples: %s' % (n, e0.elapsed_time(e1) / reps, ' | '.join(rows[4:40:4])))
