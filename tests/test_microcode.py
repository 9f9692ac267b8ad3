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
    gen_slots = int(re.search(r'#define UC_GEN_SLOTS (\d+)', inc).group(1))
    gen_window = int(re.search(r'#define UC_GEN_WINDOW (\d+)', inc).group(1))
    g = gen.Graph()
    gen.build_pairing(g)
    gen.fix_user_counts(g)
    gen.fuse(g)
    phases, slot_of, used = gen.schedule(g, G, gen_slots, gen_window)
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
    run = lambda P_, Q_: gen.run_program(g, phases, slot_of, nslots, {'P': ((-P_[0]) % R.p, P_[1]), 'QX': Q_[0], 'QY': Q_[1]})
    out = run(P, Q)
    got = [out['E00'], out['E10'], out['E01'], out['E11'], out['E02'], out['E12']]
    assert got == R.pair(P, Q)
    # the G2 membership verdict the program computes on the side: zero for G2, non-zero for twist points outside it
    assert out['SG'] == (0, 0)
    T = R.g2_twist_point(int(a % 1000))
    for Qo in (T, R.g2_mul_raw(R.n, T), R.g2_mul_raw(R.n * R.TWIST_COFACTOR // 13, T)):
        assert R.g2_on_curve(Qo) and not R.g2_in_subgroup(Qo)
        assert run(P, Qo)['SG'] != (0, 0)
    # the committed opcode stream is the one just generated
    committed = [int(x) for x in re.findall(r'\d+', re.search(r'uc_ops\[UC_NPHASES \+ UC_PAD_PHASES\] = \{(.*?)\};', inc, re.S).group(1))]
    pad = int(re.search(r'#define UC_PAD_PHASES (\d+)', inc).group(1))
    assert pad >= 8 and committed == [gen.op_byte(o) for o, _ in phases] + [gen.op_byte('NOP')] * pad   # idle phases close the table
    # ... and so is every lane word (the generator is deterministic)
    words = re.search(r'uc_prog\[[^\]]*\] = \{(.*?)\n\};', inc, re.S).group(1)
    committed_words = [(int(x, 16), int(y, 16)) for x, y in re.findall(r'\{0x([0-9a-f]{8})u, 0x([0-9a-f]{8})u\}', words)]
    assert committed_words == [w for ph in gen.prog_words(phases, slot_of, G) for w in ph]
    # opcode in the low five bits, dispatch class above
    assert all(gen.op_byte(o) & 31 == gen.OPC[o] and gen.op_byte(o) >> 5 == gen.OP_CLASS.get(o, 0) for o in gen.OPS)
