"""Times drx_rp_verify_dev / create / encrypt for the library named by DRYNX_LIB (experiment harness)."""
import ctypes, os, sys, time
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
be.set_option('rp_verify_path', int(os.environ.get('DRX_RP_PATH', '0')))
be.set_option('rp_ucode_lanes', int(os.environ.get('DRX_UC_LANES', '0')))
be.set_option('rp_create_path', int(os.environ.get('DRX_CREATE_PATH', '0')))
dev = torch.device('cuda', 0)
st = torch.cuda.current_stream(); be.set_stream(st.cuda_stream or 1)
vp = lambda t: ctypes.c_void_p(t.data_ptr())
def scal(k):
    b = np.frombuffer(g.bytes(32 * k), np.uint8).copy(); b[::32] &= 0x7f
    return torch.from_numpy(b).to(dev)
def timed(fn, reps=3):
    fn(); torch.cuda.synchronize()
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    e0.record(st)
    for _ in range(reps): fn()
    e1.record(st); torch.cuda.synchronize()
    return e0.elapsed_time(e1) / reps
out = [os.environ.get('DRYNX_LIB', 'x/default/lib.so').split('/')[-2]]
for n in [int(x) for x in (sys.argv[1:] or ['500', '8192'])]:
    sid = torch.zeros(n, dtype=torch.int32, device=dev)
    sec = torch.from_numpy(g.integers(0, U ** L, size=n, dtype=np.int64)).to(dev)
    r, rnd = scal(n), scal(n * L * (3 + NS))
    ct = torch.empty(n * 128, dtype=torch.uint8, device=dev)
    proofs = torch.empty(n * be.proof_size(NS, L), dtype=torch.uint8, device=dev)
    ok = torch.empty(n, dtype=torch.uint8, device=dev)
    te = timed(lambda: ck(lib.drx_encrypt_dev(be.h, n, vp(sec), vp(r), vp(ct))))
    tc = timed(lambda: ck(lib.drx_rp_create_dev(be.h, n, vp(sid), NS, L, vp(sec), vp(r), vp(ct), vp(rnd), vp(proofs))))
    tv = timed(lambda: ck(lib.drx_rp_verify_dev(be.h, n, vp(sid), NS, L, vp(proofs), vp(ok))))
    assert bool((ok == 1).all())
    out.append('n=%d enc %.3f ms create %.2f ms verify %.2f ms (%.1f k proofs/s)' % (n, te, tc, tv, n / tv))
    ck(lib.drx_profile_enable(be.h, 1))
    ck(lib.drx_rp_create_dev(be.h, n, vp(sid), NS, L, vp(sec), vp(r), vp(ct), vp(rnd), vp(proofs))); be.synchronize()
    ck(lib.drx_rp_verify_dev(be.h, n, vp(sid), NS, L, vp(proofs), vp(ok))); be.synchronize()
    parts = []
    for k in ('k_rp_create_hdr', 'k_rp_create_items', 'k_rp_create_g2', 'k_rp_verify_hdr', 'k_rp_verify_items', 'k_rp_verify_prep', 'k_pair_ucode', 'k_rp_verify_gt', 'k_rp_verify_finish'):
        tms, cnt = ctypes.c_double(), ctypes.c_uint64()
        ck(lib.drx_profile_get(be.h, k.encode(), ctypes.byref(tms), ctypes.byref(cnt)))
        if cnt.value: parts.append('%s=%.2f' % (k[2:], tms.value))
    ck(lib.drx_profile_enable(be.h, 0))
    out.append('[' + ' '.join(parts) + ']')
n = 1 << 20
m = torch.from_numpy(g.integers(0, 10, size=n, dtype=np.int64)).to(dev); r = scal(n)
ct = torch.empty(n * 128, dtype=torch.uint8, device=dev)
te = timed(lambda: ck(lib.drx_encrypt_dev(be.h, n, vp(m), vp(r), vp(ct))))
out.append('enc 1M: %.2f ms (%.1f M/s)' % (te, n / te / 1e3))
dec = torch.empty(1 << 17, dtype=torch.int64, device=dev)
skb = np.frombuffer(b'\0' + g.bytes(31), np.uint8).copy()
td = timed(lambda: ck(lib.drx_decrypt_dev(be.h, 1 << 17, ctypes.c_void_p(skb.ctypes.data), vp(ct), vp(dec))))
out.append('dec 128k: %.2f ms (%.1f M/s)' % (td, (1 << 17) / td / 1e3))
agg = torch.empty(1024 * 128, dtype=torch.uint8, device=dev)
ta = timed(lambda: ck(lib.drx_aggregate_dev(be.h, 1024, 1024, vp(ct), vp(agg))))
out.append('agg 1024x1024: %.2f ms (%.1f M adds/s)' % (ta, n / ta / 1e3))
print(' | '.join(out), flush=True)
