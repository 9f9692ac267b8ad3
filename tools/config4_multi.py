"""BASELINE config 4 on N GPUs: frequency count, 1,000 data providers x 1,000 bins.
Every rank owns a contiguous block of providers: it encrypts their counts (device-resident), folds them into one
partial ciphertext vector, the partials are all-gathered with NCCL (V x 128 B per rank) and rank 0 folds and decrypts.
Run for BN256 (Drynx's suite) and edwards25519 (UnLynx's default). Times are CUDA events, max over ranks.

  python -m torch.distributed.run --nnodes=1 --nproc-per-node N --master-addr 127.0.0.1 --master-port P \
      tools/config4_multi.py [--out profiles/r01_config4_Ngpu.json]
(also runs as a plain `python tools/config4_multi.py` on one GPU)"""
import argparse
import ctypes
import json
import os
import sys

import numpy as np

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
import torch  # noqa: E402
import torch.distributed as dist  # noqa: E402
from drynx_b200 import Backend, _lib  # noqa: E402
from drynx_b200.sharding import gather_and_fold, shard_range  # noqa: E402

ap = argparse.ArgumentParser()
ap.add_argument('--out', default='')
ap.add_argument('--ndp', type=int, default=1000)
ap.add_argument('--bins', type=int, default=1000)
ap.add_argument('--reps', type=int, default=5)
args = ap.parse_args()

rank = int(os.environ.get('RANK', '0'))
world = int(os.environ.get('WORLD_SIZE', '1'))
local = int(os.environ.get('LOCAL_RANK', '0'))
torch.cuda.set_device(local)
dev = torch.device('cuda', local)
if world > 1:
    dist.init_process_group('nccl', device_id=dev)
be = Backend(local)
lib, ck = be.lib, _lib.check
st = torch.cuda.current_stream()
be.set_stream(st.cuda_stream or 1)
vp = lambda t: ctypes.c_void_p(t.data_ptr())

NDP, V = args.ndp, args.bins
lo, hi = shard_range(NDP, rank, world)
mine = hi - lo
g = np.random.default_rng(4)                       # the same stream on every rank: ranks slice their block out of it
counts_all = g.integers(0, 10, size=(NDP, V), dtype=np.int64)
want = counts_all.sum(axis=0)
counts = torch.from_numpy(counts_all[lo:hi].copy()).to(dev).reshape(-1)
r = torch.from_numpy(np.frombuffer(np.random.default_rng(100 + rank).bytes(32 * mine * V), np.uint8).copy())
r[::32] &= 0x7f
r = r.to(dev)


def timed(fn, reps):
    fn()
    torch.cuda.synchronize()
    if world > 1:
        dist.barrier()
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    e0.record(st)
    for _ in range(reps):
        fn()
    e1.record(st)
    torch.cuda.synchronize()
    ms = e0.elapsed_time(e1) / reps
    if world > 1:
        t = torch.tensor([ms], dtype=torch.float64, device=dev)
        dist.all_reduce(t, op=dist.ReduceOp.MAX)
        ms = float(t.item())
    return ms


out = {'n_gpus': world, 'providers': NDP, 'bins': V, 'encryptions': NDP * V, 'scaling': 'strong (fixed 10^6 encryptions split over the ranks)'}
for suite in ('bn256', 'ed25519'):
    ctb = 128 if suite == 'bn256' else 64
    if suite == 'bn256':
        sk = (0x1234567).to_bytes(32, 'big')
        P_ORDER = 65000549695646603732796438742359905742825358107623003571877145026864184071783
        g1 = bytes(31) + b'\x01' + (P_ORDER - 2).to_bytes(32, 'big')
        be.set_pubkey(be.g1_mul(sk, g1))
        be.create_decryption_table(10000)
        enc, agg_dev = lib.drx_encrypt_dev, lib.drx_aggregate_dev
    else:
        sk = (0x1234567).to_bytes(32, 'little')
        be.ed_set_pubkey(be.ed_point_mul(sk, bytes([0x58] + [0x66] * 31)))
        be.ed_create_decryption_table(10000)
        enc, agg_dev = lib.drx_ed_encrypt_dev, lib.drx_ed_aggregate_dev
    ct = torch.empty(mine * V * ctb, dtype=torch.uint8, device=dev)
    partial = torch.empty(V * ctb, dtype=torch.uint8, device=dev)
    gathered = torch.empty(world * V * ctb, dtype=torch.uint8, device=dev)
    total = torch.empty(V * ctb, dtype=torch.uint8, device=dev)
    t_enc = timed(lambda: ck(enc(be.h, mine * V, vp(counts), vp(r), vp(ct))), args.reps)
    t_agg = timed(lambda: ck(agg_dev(be.h, mine, V, vp(ct), vp(partial))), args.reps)

    def fold(g_):
        ck(agg_dev(be.h, world, V, vp(g_), vp(total)))
        return total

    def gather_fold():
        res = gather_and_fold(partial, rank, world, dist, fold, gathered=gathered)
        if world == 1:
            total.copy_(res)
    t_gf = timed(gather_fold, args.reps)

    def whole():
        ck(enc(be.h, mine * V, vp(counts), vp(r), vp(ct)))
        ck(agg_dev(be.h, mine, V, vp(ct), vp(partial)))
        gather_fold()
    t_all = timed(whole, args.reps)
    if rank == 0:
        be.synchronize()
        host = total.cpu().numpy()
        dec = be.decrypt_int_vector(sk, host) if suite == 'bn256' else be.ed_decrypt_int_vector(sk, host)
        assert list(dec) == list(want), suite + ': the decrypted aggregate is not the global histogram'
        out[suite] = {'encrypt_ms': round(t_enc, 3), 'encryptions_per_s': round(NDP * V / t_enc * 1e3),
                      'local_aggregate_ms': round(t_agg, 3), 'ciphertext_adds_per_s': round(NDP * V / t_agg * 1e3),
                      'gather_and_fold_ms': round(t_gf, 3), 'gathered_bytes_per_rank': V * ctb,
                      'encrypt_aggregate_gather_ms': round(t_all, 3),
                      'encryptions_per_s_whole_pipeline': round(NDP * V / t_all * 1e3), 'decrypted_ok': True}
if rank == 0:
    print(json.dumps(out))
    if args.out:
        with open(os.path.join(ROOT, args.out), 'w') as f:
            json.dump(out, f, indent=1)
if world > 1:
    dist.barrier()
    dist.destroy_process_group()
