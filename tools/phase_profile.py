"""Cycles per phase of the lane-group pairing kernel, by opcode, as the first warp of the grid sees them.

    python tools/phase_profile.py build [name [nvcc flags]]   # here: build/prof/<name>/libdrynx_gpu_prof.so (s4 kernel, -DUCS_PROFILE)
    DRYNX_LIB=build/prof/prof/libdrynx_gpu_prof.so python tools/phase_profile.py 40 500 8192     # on the GPU box

The instrumented kernel reads the clock once per phase (warp 0 of block 0 adds the difference to a per-opcode counter);
it is an experiment build and never the library the tests or the bench load."""
import ctypes
import os
import re
import subprocess
import sys

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
PROF_DIR = os.path.join(ROOT, 'build', 'prof')


def build(name='prof', extra=()):
    """build/prof/<name>/libdrynx_gpu_prof.so: the shipped objects with the s4 pairing kernel rebuilt with
    -DUCS_PROFILE (not for names that start with "plain": A/B builds for tools/ucode_sweep.py) and any extra flags
    (e.g. -DUCS_PROFILE=2 for the four-way cut of the addition phases)."""
    from drynx_b200 import build as B
    B.build_gpu()
    out_dir = os.path.join(PROF_DIR, name)
    os.makedirs(out_dir, exist_ok=True)
    obj = os.path.join(out_dir, 'kernels_rp_ucode_s4.o')
    src = os.path.join(B.CSRC, 'kernels_rp_ucode_s4.cu')
    subprocess.check_call([B.NVCC] + B.NVCC_FLAGS + (['-DUCS_PROFILE'] if not name.startswith('plain') and not any(f.startswith('-DUCS_PROFILE') for f in extra) else []) + list(extra) + ['-I', B.CSRC, '-c', src, '-o', obj],
                          stdout=subprocess.DEVNULL, stderr=subprocess.DEVNULL)
    objs = [obj if s == 'kernels_rp_ucode_s4.cu' else os.path.join(ROOT, 'build', s.replace('.cu', '.o')) for s in B.GPU_SOURCES]
    so = os.path.join(out_dir, 'libdrynx_gpu_prof.so')
    subprocess.check_call([B.NVCC, '-shared', '-o', so] + objs + ['-lcudart'], stdout=subprocess.DEVNULL, stderr=subprocess.DEVNULL)
    print(so)


def opcode_names():
    inc = open(os.path.join(ROOT, 'drynx_b200', 'csrc', 'pairing_ucode_g4.inc')).read()
    names = {}
    for k, v in re.findall(r'#define UC_([A-Z0-9]+) (\d+)\n', inc):
        if k not in ('G', 'NSLOTS', 'NPHASES', 'GEN_SLOTS', 'GEN_WINDOW') and not k.startswith('IN_') and not k.startswith('OUT_'):
            names[int(v)] = k
    return names


def main():
    if sys.argv[1:2] == ['build']:
        return build(*(sys.argv[2:3] or ['prof']), extra=sys.argv[3:])
    import numpy as np
    import torch
    from drynx_b200 import Backend, _lib
    NS, U, L = 3, 16, 4
    be = Backend(0)
    lib, ck = be.lib, _lib.check
    g = np.random.default_rng(5)
    g1 = bytes(31) + b'\x01' + (65000549695646603732796438742359905742825358107623003571877145026864184071783 - 2).to_bytes(32, 'big')
    be.set_pubkey(be.g1_mul(b'\0' + g.bytes(31), g1))
    Y, A = be.init_range_proof_signature([int.from_bytes(g.bytes(31), 'big') + 1 for _ in range(NS)], U)
    be.set_signatures(0, NS, U, Y, A)
    be.set_option('rp_ucode_lanes', 4)
    be.set_option('rp_ucode_mode', 1)
    dev = torch.device('cuda', 0)
    vp = lambda t: ctypes.c_void_p(t.data_ptr())

    def scal(k):
        b = np.frombuffer(g.bytes(32 * k), np.uint8).copy()
        b[::32] &= 0x7f
        return torch.from_numpy(b).to(dev)
    names = opcode_names()
    for n in [int(a) for a in sys.argv[1:]] or [40, 500]:
        sid = torch.zeros(n, dtype=torch.int32, device=dev)
        sec = torch.from_numpy(g.integers(0, U ** L, size=n, dtype=np.int64)).to(dev)
        r, rnd = scal(n), scal(n * L * (3 + NS))
        ct = torch.empty(n * 128, dtype=torch.uint8, device=dev)
        proofs = torch.empty(n * be.proof_size(NS, L), dtype=torch.uint8, device=dev)
        ok = torch.empty(n, dtype=torch.uint8, device=dev)
        ck(lib.drx_encrypt_dev(be.h, n, vp(sec), vp(r), vp(ct)))
        ck(lib.drx_rp_create_dev(be.h, n, vp(sid), NS, L, vp(sec), vp(r), vp(ct), vp(rnd), vp(proofs)))
        out = (ctypes.c_ulonglong * 64)()
        for rep in range(2):                                      # the first pass warms up; the read clears the counters
            ck(lib.drx_rp_verify_dev(be.h, n, vp(sid), NS, L, vp(proofs), vp(ok)))
            be.synchronize()
            assert lib.drx_debug_phase_profile(out) == 0
        assert bool((ok == 1).all())
        cyc, cnt = list(out[:32]), list(out[32:])
        print('== %d proofs: %d phases, %.3f M cycles in all for the first warp' % (n, sum(cnt), sum(cyc) / 1e6))
        if cnt[24]:   # -DUCS_PROFILE=2: the addition phases of the first thread, cut in four
            k = cnt[24]
            print('  ADD phases cut in four: dispatch %.0f, loads + sum %.0f, barrier %.0f, store + barrier + loop %.0f cycles'
                  % (cyc[24] / k, cyc[25] / k, cyc[26] / k, cyc[27] / k))
            cnt[24] = 0
        for op in sorted(range(24), key=lambda o: -cyc[o]):
            if cnt[op]:
                print('  %-9s %5d phases  %7.0f cycles each  %5.1f %% of the time' % (names.get(op, op), cnt[op], cyc[op] / cnt[op],
                                                                                   100.0 * cyc[op] / sum(cyc)))


if __name__ == '__main__':
    main()
