"""Small driver for ncu: one signature set (ns=3, u=16), N proofs created on the device, then verified `reps` times.
Usage: python tools/prof_verify.py [n_proofs] [reps]"""
import ctypes, os, sys
import numpy as np
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
import torch
from drynx_b200 import Backend, _lib

n = int(sys.argv[1]) if len(sys.argv) > 1 else 4096
reps = int(sys.argv[2]) if len(sys.argv) > 2 else 2
NS, U, L = 3, 16, 4
be = Backend(0)
lib, ck = be.lib, _lib.check
g = np.random.default_rng(5)
g1 = bytes(31) + b'\x01' + (65000549695646603732796438742359905742825358107623003571877145026864184071783 - 2).to_bytes(32, 'big')
sk = g.bytes(31)
be.set_pubkey(be.g1_mul(b'\0' + sk, g1))
Y, A = be.init_range_proof_signature([int.from_bytes(g.bytes(31), 'big') + 1 for _ in range(NS)], U)
be.set_signatures(0, NS, U, Y, A)
be.set_option('rp_verify_path', int(os.environ.get('DRX_RP_PATH', '0')))
for _o, _e in (('rp_ucode_lanes', 'DRX_UC_LANES'), ('rp_ucode_mode', 'DRX_UC_MODE'),
               ('rp_ucode_resident', 'DRX_UC_RESIDENT'), ('rp_ucode_nseg', 'DRX_UC_NSEG')):
    be.set_option(_o, int(os.environ.get(_e, '0')))
dev = torch.device('cuda', 0)
vp = lambda t: ctypes.c_void_p(t.data_ptr())
def scal(k):
    b = np.frombuffer(g.bytes(32 * k), np.uint8).copy(); b[::32] &= 0x7f
    return torch.from_numpy(b).to(dev)
sid = torch.zeros(n, dtype=torch.int32, device=dev)
sec = torch.from_numpy(g.integers(0, U ** L, size=n, dtype=np.int64)).to(dev)
r, rnd = scal(n), scal(n * L * (3 + NS))
ct = torch.empty(n * 128, dtype=torch.uint8, device=dev)
size = be.proof_size(NS, L)
proofs = torch.empty(n * size, dtype=torch.uint8, device=dev)
ok = torch.empty(n, dtype=torch.uint8, device=dev)
ck(lib.drx_encrypt_dev(be.h, n, vp(sec), vp(r), vp(ct)))
ck(lib.drx_rp_create_dev(be.h, n, vp(sid), NS, L, vp(sec), vp(r), vp(ct), vp(rnd), vp(proofs)))
for _ in range(reps):
    ck(lib.drx_rp_verify_dev(be.h, n, vp(sid), NS, L, vp(proofs), vp(ok)))
be.synchronize()
assert bool((ok == 1).all())
print('verified', n, 'x', reps)
