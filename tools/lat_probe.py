"""Latency probes: cycles per Montgomery product for one warp per SM / per sub-partition / two per sub-partition."""
import os, sys
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
from drynx_b200 import Backend
be = Backend(0)
it = 20000
for kind, what in ((11, '1 chain, 1 warp/SM'), (12, '3 chains, 1 warp/SM'), (21, '1 chain, 1 warp/SMSP'), (22, '3 chains, 1 warp/SMSP'),
                   (31, '1 chain, 2 warps/SMSP'), (32, '3 chains, 2 warps/SMSP'), (1, '1 chain, full GPU'), (2, '3 chains, full GPU')):
    ms, work = be.microbench(kind, it)
    per_thread = work / (148 * (32 if kind < 20 else 128 if kind < 30 else 256)) if kind >= 10 else work / (148 * 8 * 256)
    print('%-24s %.3f ms, %.0f cycles per product per warp (at 1965 MHz)' % (what, ms, ms * 1.965e6 / per_thread))
for kind, what, it2 in ((13, 'fp_inv, 1 warp/SM', 200), (33, 'fp_inv, 2 warps/SMSP', 200), (3, 'fp_inv, full GPU', 200),
                        (14, 'jac_add, 1 warp/SM', 2000), (34, 'jac_add, 2 warps/SMSP', 2000)):
    ms, work = be.microbench(kind, it2)
    print('%-24s %.3f ms, %.0f cycles per operation per warp (at 1965 MHz)' % (what, ms, ms * 1.965e6 / it2))
ms0, w0 = be.microbench(0, 20000)
ms5, w5 = be.microbench(5, 20000)
print('IMAD.WIDE chains: %.2f T/s; DFMA (round toward zero) chains: %.2f T/s; ratio %.2f' % (w0 / ms0 / 1e9, w5 / ms5 / 1e9, (w5 / ms5) / (w0 / ms0)))
