"""World-size-2 gloo test of the multi-GPU host logic (block partition of DPs, all-gather of partial aggregates,
fold on rank 0). The aggregation itself is done by the CPU oracle here; on GPUs the same code path calls
Backend.aggregate / drx_aggregate_dev (bench.py)."""
import os
import socket
import subprocess
import sys
import textwrap

import numpy as np

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))


def test_shard_range_partitions_exactly():
    from drynx_b200.sharding import shard_range
    for n in (0, 1, 7, 100, 1001):
        for world in (1, 2, 3, 8):
            spans = [shard_range(n, r, world) for r in range(world)]
            assert spans[0][0] == 0 and spans[-1][1] == n
            assert all(spans[r][1] == spans[r + 1][0] for r in range(world - 1))
            sizes = [hi - lo for lo, hi in spans]
            assert max(sizes) - min(sizes) <= 1


WORKER = textwrap.dedent('''
    import os, sys
    sys.path.insert(0, {root!r}); sys.path.insert(0, os.path.join({root!r}, 'tests')); sys.path.insert(0, os.path.join({root!r}, 'oracle'))
    import numpy as np, torch.distributed as dist
    import bn256_ref as R
    from conftest import Oracle, SplitMix
    from drynx_b200.sharding import shard_range, gather_partials, fold_on_root, gather_and_fold, gather_accept_bytes
    import torch
    dist.init_process_group('gloo')
    rank, world = dist.get_rank(), dist.get_world_size()
    orc = Oracle(); rng = SplitMix(77)
    sk = rng.scalar(); P = R.g1_to_bytes(R.g1_mul(sk, R.G1_GEN))
    ndp, v = 7, 3
    vals = [[rng.randint(-20, 20) for _ in range(v)] for _ in range(ndp)]
    rs = [rng.scalars_bytes(v) for _ in range(ndp)]
    lo, hi = shard_range(ndp, rank, world)
    mine = [orc.encrypt(P, vals[d], rs[d]) for d in range(lo, hi)]
    partial = np.frombuffer(orc.aggregate(mine)[1], np.uint8)
    gathered = gather_partials(partial, world, dist)
    total = fold_on_root(gathered, rank, lambda vecs: orc.aggregate([x.tobytes() for x in vecs])[1])
    # the call bench.py and tools/config4_multi.py make (there with CUDA tensors and drx_aggregate_dev as the fold)
    total2 = gather_and_fold(torch.from_numpy(partial.copy()), rank, world, dist,
                             lambda g: orc.aggregate([g[r].numpy().tobytes() for r in range(world)])[1])
    assert (total2 is None) == (rank != 0)
    # accept bytes of a verification split by proof index come back in proof order on every rank
    nproofs = 11
    plo, phi = shard_range(nproofs, rank, world)
    mine_ok = torch.tensor([(i * 7 + 3) % 5 != 0 for i in range(plo, phi)], dtype=torch.uint8)
    all_ok = gather_accept_bytes(mine_ok, nproofs, rank, world, dist)
    assert all_ok.tolist() == [int((i * 7 + 3) % 5 != 0) for i in range(nproofs)]
    if rank == 0:
        whole = orc.aggregate([orc.encrypt(P, vals[d], rs[d]) for d in range(ndp)])[1]
        assert total == whole == total2
        orc.dlog_reset()
        out, _ = orc.decrypt(sk.to_bytes(32, 'big'), total)
        assert list(out) == [sum(vals[d][j] for d in range(ndp)) for j in range(v)]
        print('SHARD_OK')
    # the same plumbing carries edwards25519 ciphertexts (64 bytes each): config 1 / 4 over UnLynx's default suite
    import ed25519_ref as E
    esk = (rng.scalar() % E.l).to_bytes(32, 'little'); eP = orc.ed_mul(esk)
    ers = [rng.scalars_bytes(v) for _ in range(ndp)]
    emine = [orc.ed_encrypt(eP, vals[d], ers[d]) for d in range(lo, hi)]
    epartial = np.frombuffer(orc.ed_aggregate(emine)[1], np.uint8)
    egathered = gather_partials(epartial, world, dist)
    etotal = fold_on_root(egathered, rank, lambda vecs: orc.ed_aggregate([x.tobytes() for x in vecs])[1])
    if rank == 0:
        assert etotal == orc.ed_aggregate([orc.ed_encrypt(eP, vals[d], ers[d]) for d in range(ndp)])[1]
        orc.ed_dlog_reset()
        rc, out, _ = orc.ed_decrypt(esk, etotal)
        assert rc == 0 and list(out) == [sum(vals[d][j] for d in range(ndp)) for j in range(v)]
        print('SHARD_ED_OK')
    dist.barrier(); dist.destroy_process_group()
''')


def test_two_rank_gather_and_fold(tmp_path):
    from drynx_b200 import build
    build.build_oracle()
    script = tmp_path / 'worker.py'
    script.write_text(WORKER.format(root=ROOT))
    s = socket.socket(); s.bind(('127.0.0.1', 0)); port = s.getsockname()[1]; s.close()
    p = subprocess.run([sys.executable, '-m', 'torch.distributed.run', '--nnodes=1', '--nproc-per-node', '2', '--master-addr',
                        '127.0.0.1', '--master-port', str(port), str(script)], capture_output=True, text=True, timeout=600)
    assert p.returncode == 0, p.stderr[-2000:]
    assert 'SHARD_OK' in p.stdout and 'SHARD_ED_OK' in p.stdout
