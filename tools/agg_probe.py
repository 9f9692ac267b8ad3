"""Times drx_aggregate_dev (stage 1 / stage 2 separately) for a few shapes and slice counts (experiment harness)."""
import ctypes, os, sys
import numpy as np
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
import torch
from drynx_b200 import Backend, _lib
be = Backend(0); lib, ck = be.lib, _lib.check
g = np.random.default_rng(4)
P = 65000549695646603732796438742359905742825358107623003571877145026864184071783
g1 = bytes(31) + b'\x01' + (P - 2).to_bytes(32, 'big')
be.set_pubkey(be.g1_mul(b'\0' + g.bytes(31), g1))
dev = torch.device('cuda', 0)
st = torch.cuda.current_stream(); be.set_stream(st.cuda_stream or 1)
vp = lambda t: ctypes.c_void_p(t.data_ptr())
peak = float(os.environ.get('INT_PEAK', '8.83e12'))
n = 1 << 20
m = torch.from_numpy(g.integers(0, 10, size=n, dtype=np.int64)).to(dev)
rb = np.frombuffer(g.bytes(32 * n), np.uint8).copy(); rb[::32] &= 0x7f
r = torch.from_numpy(rb).to(dev)
ct = torch.empty(n * 128, dtype=torch.uint8, device=dev)
ck(lib.drx_encrypt_dev(be.h, n, vp(m), vp(r), vp(ct)))
for ndp, v in ((1000, 1000), (1024, 1024), (125, 1000), (500, 1000), (100, 5), (10, 1), (8, 1000), (4096, 256)):
    out = torch.empty(v * 128, dtype=torch.uint8, device=dev)
    ref = None
    for slices in [int(x) for x in os.environ.get('SLICES', '0,16,32,48,64,96,128,192,256').split(',')]:
        be.set_option('agg_slices', slices)
        def run(): ck(lib.drx_aggregate_dev(be.h, ndp, v, vp(ct), vp(out)))
        run(); torch.cuda.synchronize()
        e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        e0.record(st)
        for _ in range(5): run()
        e1.record(st); torch.cuda.synchronize()
        ms = e0.elapsed_time(e1) / 5
        if ref is None: ref = out.clone()
        same = bool(torch.equal(ref, out))
        ck(lib.drx_profile_enable(be.h, 1)); run(); be.synchronize()
        parts = []
        for k in ('k_agg_partial', 'k_agg_final'):
            tms, cnt = ctypes.c_double(), ctypes.c_uint64()
            ck(lib.drx_profile_get(be.h, k.encode(), ctypes.byref(tms), ctypes.byref(cnt)))
            parts.append('%s=%.3f' % (k[6:], tms.value))
        ck(lib.drx_profile_enable(be.h, 0))
        print('%5d x %-5d slices %3d: %.3f ms  %.0f M adds/s  %.0f GB/s  frac %.3f  %s [%s]' % (
            ndp, v, slices, ms, ndp * v / ms / 1e3, ndp * v * 128 / ms / 1e6, ndp * v * 22 * 128 / (ms * 1e-3) / peak,
            'same' if same else 'DIFFERENT', ' '.join(parts)), flush=True)
    be.set_option('agg_slices', 0)
