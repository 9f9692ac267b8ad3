"""The lane-group micro-program of the pairing (tools/microcode/gen.py -> csrc/pairing_ucode.inc), checked on the CPU:
the generator's own simulator must reproduce the big-integer definition, and the committed tables must be what the
generator emits for the committed parameters."""
import importlib.util
import os
import re

import bn256_ref as R

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))


def _gen():
    spec = importlib.util.spec_from_file_location('ucode_gen', os.path.join(ROOT, 'tools', 'microcode', 'gen.py'))
    m = importlib.util.module_from_spec(spec)
    spec.loader.exec_module(m)
    return m


def test_scheduled_program_computes_the_pairing_and_matches_committed_tables(rng):
    gen = _gen()
    inc = open(os.path.join(ROOT, 'drynx_b200', 'csrc', 'pairing_ucode_g4.inc')).read()
    G = int(re.search(r'#define UC_G (\d+)', inc).group(1))
    nslots = int(re.search(r'#define UC_NSLOTS (\d+)', inc).group(1))
    nphases = int(re.search(r'#define UC_NPHASES (\d+)', inc).group(1))
    g = gen.Graph()
    gen.build_pairing(g)
    gen.fix_user_counts(g)
    gen.fuse(g)
    phases, slot_of, used = gen.schedule(g, G, nslots, 60)
    assert len(phases) == nphases and (used | 1) == nslots
    # no phase writes a slot that another lane of the same phase reads (the kernel only synchronises between phases)
    for op, lanes in phases:
        ops = [n for n in lanes if n is not None]
        for n in ops:
            for m in ops:
                if m is not n:
                    assert slot_of[n.id] not in [slot_of[x.id] for x in (m.a, m.b, m.c) if x is not None]
    a, b = rng.scalar(), rng.scalar()
    P, Q = R.g1_mul(a, R.G1_GEN), R.g2_mul(b, R.G2_GEN)
    out = gen.run_program(g, phases, slot_of, nslots, {'P': (P[0], P[1]), 'QX': Q[0], 'QY': Q[1], 'ONE': (1, 0), 'ZERO': (0, 0)})
    got = [out['E00'], out['E10'], out['E01'], out['E11'], out['E02'], out['E12']]
    assert got == R.pair(P, Q)
    # the committed opcode stream is the one just generated
    committed = [int(x) for x in re.findall(r'\d+', re.search(r'uc_ops\[UC_NPHASES\] = \{(.*?)\};', inc, re.S).group(1))]
    assert committed == [gen.OPC[o] for o, _ in phases]
