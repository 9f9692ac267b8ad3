"""Time of the pair-shuffle argument (tests/shuffle_proof.py) through the CUDA library for k pairs: host Python included.
    python tests/shuffle_proof_rate.py [k ...]"""
import os
import sys
import time

HERE = os.path.dirname(os.path.abspath(__file__))
sys.path.insert(0, HERE)
sys.path.insert(0, os.path.join(HERE, '..', 'oracle'))
sys.path.insert(0, os.path.join(HERE, '..'))

import rep_proofs as RP  # noqa: E402
import shuffle_proof as SP  # noqa: E402
from conftest import SplitMix  # noqa: E402
from test_shuffle_proof import _case  # noqa: E402
from drynx_b200 import Backend  # noqa: E402

be = Backend(0)
eng = RP.BackendEngine(be)
for k in [int(a) for a in sys.argv[1:]] or [64, 1024, 2047]:
    g, h, X, Y, beta, pi, secrets = _case(SplitMix(k), k)
    Xbar, Ybar = SP.shuffle_pairs(eng, pi, g, h, beta, X, Y)
    SP.pair_shuffle_prove(eng, pi, g, h, beta, X, Y, secrets)
    t0 = time.perf_counter()
    proof = SP.pair_shuffle_prove(eng, pi, g, h, beta, X, Y, secrets)
    t1 = time.perf_counter()
    ok = SP.pair_shuffle_verify(eng, g, h, X, Y, Xbar, Ybar, proof)
    t2 = time.perf_counter()
    launches = be.launches
    ok2 = SP.pair_shuffle_verify(eng, g, h, X, Y, Xbar, Ybar, proof)
    print('k = %5d: prove %.2f ms, verify %.2f ms (%d kernel launches), proof %d bytes, accepted %s' %
          (k, (t1 - t0) * 1e3, (t2 - t1) * 1e3, be.launches - launches, len(proof), ok and ok2))
    assert ok
be.close()
