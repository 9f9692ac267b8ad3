#!/usr/bin/env python3
"""Generates the lane-group micro-program for the optimal ate pairing (drynx_b200/csrc/pairing_ucode.inc).

The cooperative pairing kernel (kernels_rp_ucode.cu) keeps ONE copy of an item's state as an array of Fp2 slots in
shared memory and lets a group of G lanes execute a static schedule over it: every phase all lanes run the same
Fp2 operation (MUL, ADD, ...) on their own (dst, a, b) slots. This script

  1. traces the pairing (the same formulas as csrc/tower.cuh / pairing.cuh) on symbolic Fp2 values into a DAG,
  2. list-schedules the DAG onto G lanes, one opcode per phase, and allocates slots by liveness,
  3. EXECUTES the emitted program with Python integers and checks it against oracle/bn256_ref.pair, and
  4. writes the tables as a C include.

Nothing here is copied from the reference; the formulas are the ones already used and tested in this repo.
"""
import os
import sys

ROOT = os.path.dirname(os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
sys.path.insert(0, os.path.join(ROOT, 'oracle'))
import bn256_ref as R  # noqa: E402

# ---------------------------------------------------------------------------------------------- symbolic Fp2 DAG
OPS = ['NOP', 'MUL', 'ADD', 'SUB', 'DBL', 'NEG', 'MULXI', 'CONJ', 'MULC', 'MULFP0', 'MULFP1', 'INV', 'COPY', 'CONJMULC', 'MULFPC',
       'SUB2', 'AXI', 'X3', 'X4', 'X8', 'CYCM', 'CYCP', 'SQR', 'CHK']
OPC = {n: i for i, n in enumerate(OPS)}


def uniq(*xs):
    """The distinct non-None operands in argument order (a set of nodes would iterate in address order and make the
    slot assignment differ from run to run)."""
    out = []
    for x in xs:
        if x is not None and all(x is not y for y in out):
            out.append(x)
    return out
# dispatch classes of the interpreter (three bits above the opcode in uc_ops), by how often a phase of the pairing has the
# opcode: ADD 1,888 of 6,815, SUB2 1,159, MUL 1,131, AXI 692, SQR 419, CYCP / CYCM / SUB / MULXI 1,106 together
# (class 3: the dual-product pass, class 2: one Montgomery product per thread -- each body exists once in the kernel)
OP_CLASS = {'ADD': 6, 'SUB2': 5, 'AXI': 4, 'MUL': 3, 'MULC': 3, 'CONJMULC': 3, 'SQR': 2, 'MULFP0': 2, 'MULFP1': 2, 'MULFPC': 2,
            'CYCP': 1, 'CYCM': 1, 'SUB': 1, 'MULXI': 1}


def op_byte(op):
    return OPC[op] | OP_CLASS.get(op, 0) << 5


PAD_PHASES = 16   # idle phases closing the emitted tables (read-ahead and prefetch of the interpreter)
COST = {'MUL': 10, 'SQR': 7, 'INV': 120, 'MULC': 10, 'CONJMULC': 10, 'MULFP0': 7, 'MULFP1': 7, 'MULFPC': 7}


class Node:
    __slots__ = ('op', 'a', 'b', 'c', 'imm', 'id', 'users', 'name', 'prio', 'slot', 'phase', '_dead')

    def __init__(self, g, op, a=None, b=None, imm=0, name=None):
        self.op, self.a, self.b, self.imm, self.name = op, a, b, imm, name
        self.c = None
        self._dead = False
        self.id = len(g.nodes)
        self.users = []
        self.slot = None
        self.phase = None
        g.nodes.append(self)
        for x in (a, b):
            if x is not None:
                x.users.append(self)


class Graph:
    def __init__(self):
        self.nodes = []
        self.inputs = {}
        self.outputs = {}
        self.consts2 = []   # Fp2 constants (c0, c1)
        self.constsp = []   # Fp constants

    def inp(self, name):
        n = Node(self, 'IN', name=name)
        self.inputs[name] = n
        return n

    def out(self, name, node):
        self.outputs[name] = node

    def c2(self, val):
        if val not in self.consts2:
            self.consts2.append(val)
        return self.consts2.index(val)

    def cp(self, val):
        if val not in self.constsp:
            self.constsp.append(val)
        return self.constsp.index(val)


class S:
    """Symbolic Fp2 value. The inputs named ZERO and ONE are folded away wherever an identity applies, so they only
    cost a slot if some operation really needs them as an operand."""

    def __init__(self, g, n):
        self.g, self.n = g, n

    def _is(self, name):
        return self.n is self.g.inputs.get(name)

    def _const(self, name):
        return S(self.g, self.g.inputs[name])

    def _bin(self, op, o):
        return S(self.g, Node(self.g, op, self.n, o.n))

    def _un(self, op, imm=0):
        return S(self.g, Node(self.g, op, self.n, None, imm))

    def __mul__(self, o):
        if self._is('ZERO') or o._is('ZERO'): return self._const('ZERO')
        if self._is('ONE'): return o
        if o._is('ONE'): return self
        return self._bin('MUL', o)

    def __add__(self, o):
        if self._is('ZERO'): return o
        if o._is('ZERO'): return self
        return self._bin('ADD', o)

    def __sub__(self, o):
        if o._is('ZERO'): return self
        if self._is('ZERO'): return o.neg()
        return self._bin('SUB', o)

    def sqr(self):
        if self._is('ZERO') or self._is('ONE'): return self
        return self._bin('MUL', self)

    def dbl(self):
        if self._is('ZERO'): return self
        return self._bin('ADD', self)            # same opcode as ADD: fuller phases

    def neg(self):
        if self._is('ZERO'): return self
        return self._un('NEG')

    def xi(self):
        if self._is('ZERO'): return self
        return self._un('MULXI')

    def conj(self):
        if self._is('ZERO') or self._is('ONE'): return self
        return self._un('CONJ')

    def inv(self): return self._un('INV')

    def mulc(self, val):
        if self._is('ZERO'): return self
        return self._un('MULC', self.g.c2(val))

    def conjmulc(self, val):
        if self._is('ZERO'): return self
        return self._un('CONJMULC', self.g.c2(val))

    def mulfpc(self, val):
        if self._is('ZERO'): return self
        return self._un('MULFPC', self.g.cp(val))

    def mulfp(self, pt, which):
        if self._is('ZERO'): return self
        return S(self.g, Node(self.g, 'MULFP%d' % which, self.n, pt.n))


def chk(dx, dy, z):
    """(1, 0) unless dx == 0, dy == 0 and z != 0, then (0, 0): the verdict of the subgroup test as one slot."""
    n = Node(dx.g, 'CHK', dx.n, dy.n)
    n.c = z.n
    z.n.users.append(n)
    return S(dx.g, n)


# ---------------------------------------------------------------------------------------------- tower formulas
def f6_add(a, b): return [a[i] + b[i] for i in range(3)]
def f6_sub(a, b): return [a[i] - b[i] for i in range(3)]
def f6_mul_tau(a): return [a[2].xi(), a[0], a[1]]


def f6_mul(a, b):
    v0, v1, v2 = a[0] * b[0], a[1] * b[1], a[2] * b[2]
    t0 = (a[1] + a[2]) * (b[1] + b[2]) - v1 - v2
    t1 = (a[0] + a[1]) * (b[0] + b[1]) - v0 - v1
    t2 = (a[0] + a[2]) * (b[0] + b[2]) - v0 - v2
    return [v0 + t0.xi(), t1 + v2.xi(), t2 + v1]


def f6_mul_fp2(a, k): return [a[0] * k, a[1] * k, a[2] * k]


def f6_mul_01(a, b0, b1):
    v0, v1 = a[0] * b0, a[1] * b1
    t = (a[0] + a[1]) * (b0 + b1) - v0 - v1
    return [v0 + (a[2] * b1).xi(), t, v1 + a[2] * b0]


def f12_mul(a, b):  # a = (c0[3], c1[3])
    v0, v1 = f6_mul(a[0], b[0]), f6_mul(a[1], b[1])
    t = f6_mul(f6_add(a[0], a[1]), f6_add(b[0], b[1]))
    return (f6_add(v0, f6_mul_tau(v1)), f6_sub(f6_sub(t, v0), v1))


def f12_sqr(a):
    v = f6_mul(a[0], a[1])
    t = f6_mul(f6_add(a[0], a[1]), f6_add(a[0], f6_mul_tau(a[1])))
    return (f6_sub(f6_sub(t, v), f6_mul_tau(v)), [x.dbl() for x in v])


def f12_mul_line(a, l0, l1, l3):
    t0 = f6_mul_fp2(a[0], l0)
    t1 = f6_mul_01(a[1], l1, l3)
    t2 = f6_mul_01(f6_add(a[0], a[1]), l0 + l1, l3)
    return (f6_add(t0, f6_mul_tau(t1)), f6_sub(f6_sub(t2, t0), t1))


def f4_sqr(a, b):
    a2, b2 = a.sqr(), b.sqr()
    return a2 + b2.xi(), (a + b).sqr() - a2 - b2


def f12_cyclo_sqr(a):
    g0, g1, g2, g3, g4, g5 = a[0][0], a[1][0], a[0][1], a[1][1], a[0][2], a[1][2]
    t00, t01 = f4_sqr(g0, g3)
    t10, t11 = f4_sqr(g1, g4)
    t20, t21 = f4_sqr(g2, g5)
    x = t21.xi()
    c00 = (t00 - g0).dbl() + t00
    c11 = (t01 + g3).dbl() + t01
    c01 = (t10 - g2).dbl() + t10
    c12 = (t11 + g5).dbl() + t11
    c02 = (t20 - g4).dbl() + t20
    c10 = (x + g1).dbl() + x
    return ([c00, c01, c02], [c10, c11, c12])


def f12_conj(a): return (a[0], [x.neg() for x in a[1]])


def f6_sqr(a):
    s0 = a[0].sqr()
    s1 = (a[0] * a[1]).dbl()
    s2 = (a[0] - a[1] + a[2]).sqr()
    s3 = (a[1] * a[2]).dbl()
    s4 = a[2].sqr()
    return [s0 + s3.xi(), s1 + s4.xi(), s1 + s2 + s3 - s0 - s4]


def f6_inv(a):
    A = a[0].sqr() - (a[1] * a[2]).xi()
    B = a[2].sqr().xi() - a[0] * a[1]
    C = a[1].sqr() - a[0] * a[2]
    F = ((a[2] * B + a[1] * C).xi() + a[0] * A).inv()
    return [A * F, B * F, C * F]


def f12_inv(a):
    t = f6_inv(f6_sub(f6_sqr(a[0]), f6_mul_tau(f6_sqr(a[1]))))
    return (f6_mul(a[0], t), [x.neg() for x in f6_mul(a[1], t)])


def frob_consts(e):
    xi = R.XI
    return [R.f2pow(xi, k * (R.p ** e - 1) // 6) for k in range(6)]


def f12_frob(a, e):
    g = frob_consts(e)
    w = [a[0][0], a[1][0], a[0][1], a[1][1], a[0][2], a[1][2]]
    out = []
    for k in range(6):
        if e % 2 == 1:
            out.append(w[k].conj() if k == 0 else w[k].conjmulc(g[k]))
        else:
            out.append(w[k] if k == 0 else w[k].mulfpc(g[k][0]))
    return ([out[0], out[2], out[4]], [out[1], out[3], out[5]])


def line_dbl(r, P):
    x, y, z, t = r
    A, B = x.sqr(), y.sqr()
    C = B.sqr()
    D = ((x + B).sqr() - A - C).dbl()
    E = A.dbl() + A
    G = E.sqr()
    ox = G - D - D
    oz = y.dbl() if z._is('ONE') else (y + z).sqr() - B - t   # (y + 1)^2 - y^2 - 1
    oy = (D - ox) * E - C.dbl().dbl().dbl()
    ot = oz.sqr()
    b = (E * t).dbl().mulfp(P, 0)            # P holds (-x_P, y_P)
    a = (x + E).sqr() - A - G - B.dbl().dbl()
    c = (oz * t).dbl().mulfp(P, 1)
    return (ox, oy, oz, ot), (a, b, c)


def line_add(r, q, r2, P):
    x, y, z, t = r
    B = q[0] * t
    D = ((q[1] + z).sqr() - r2 - t) * t
    H = B - x
    I = H.sqr()
    E = I.dbl().dbl()
    J = H * E
    L1 = D - y - y
    V = x * E
    ox = L1.sqr() - J - V - V
    oz = (z + H).sqr() - t - I
    oy = (V - ox) * L1 - (y * J).dbl()
    ot = oz.sqr()
    tt = (q[1] + oz).sqr() - r2 - ot
    a = (L1 * q[0]).dbl() - tt
    c = oz.mulfp(P, 1).dbl()
    b = L1.mulfp(P, 0).dbl()
    return (ox, oy, oz, ot), (a, b, c)


def naf(x):
    d = []
    while x:
        if x & 1:
            z = 2 - (x & 3)
            x -= z
        else:
            z = 0
        d.append(z)
        x >>= 1
    return d


def wnaf(x, w):
    d = []
    while x:
        if x & 1:
            z = x % (1 << w)
            if z >= 1 << (w - 1):
                z -= 1 << w
            x -= z
        else:
            z = 0
        d.append(z)
        x >>= 1
    return d


POW_U_WINDOW = 3   # width-3 NAF of u: digits +-1, +-3, 15 products + a^3 instead of the 22 of the plain NAF


def f12_pow_u(a):
    """a^u for a in the cyclotomic subgroup (the inverse is the conjugate)."""
    digits = wnaf(R.u, POW_U_WINDOW)
    odd = {1: a}
    if POW_U_WINDOW > 2:
        a2 = f12_cyclo_sqr(a)
        for k in range(3, 1 << (POW_U_WINDOW - 1), 2):
            odd[k] = f12_mul(odd[k - 2], a2)
    top = digits[-1]
    assert top > 0
    acc = odd[top]
    for i in range(len(digits) - 2, -1, -1):
        acc = f12_cyclo_sqr(acc)
        if digits[i] > 0:
            acc = f12_mul(acc, odd[digits[i]])
        elif digits[i] < 0:
            acc = f12_mul(acc, f12_conj(odd[-digits[i]]))
    return acc


def final_exp(f):
    a = f12_mul(f12_conj(f), f12_inv(f))
    a = f12_mul(a, f12_frob(a, 2))
    fu = f12_pow_u(a)
    fu2 = f12_pow_u(fu)
    fu3 = f12_pow_u(fu2)
    y6 = f12_conj(f12_mul(f12_frob(fu3, 1), fu3))
    y4 = f12_conj(f12_mul(f12_frob(fu2, 1), fu))
    y5 = f12_conj(fu2)
    t0 = f12_mul(f12_mul(f12_cyclo_sqr(y6), y4), y5)
    y3 = f12_conj(f12_frob(fu, 1))
    t1 = f12_mul(f12_mul(y3, y5), t0)
    t0 = f12_mul(t0, f12_frob(fu2, 2))
    t1 = f12_cyclo_sqr(f12_mul(f12_cyclo_sqr(t1), t0))
    t0 = f12_mul(t1, f12_conj(a))
    y0 = f12_mul(f12_mul(f12_frob(a, 1), f12_frob(a, 2)), f12_frob(a, 3))
    t1 = f12_mul(t1, y0)
    return f12_mul(f12_cyclo_sqr(t0), t1)


def build_pairing(g):
    """Inputs: P (one slot: c0 = -x_P, c1 = y_P; the line functions only ever need -x_P), QX, QY; ONE and ZERO are
    folded away. Outputs: 6 Fp2 of e(P, Q) in w-power order pairs, and SG = (0, 0) iff Q lies in G2."""
    P, qx, qy, one, zero = (S(g, g.inp(n)) for n in ('P', 'QX', 'QY', 'ONE', 'ZERO'))
    mqy = qy.neg()
    r = (qx, qy, one, one)
    r2 = qy.sqr()
    f = ([one, zero, zero], [zero, zero, zero])
    digits = naf(6 * R.u + 2)
    L = len(digits)
    first = True
    for i in range(L - 1, 0, -1):
        r, (a, b, c) = line_dbl(r, P)
        if first:
            # f = 1: the first line is f itself (c + b w + a w^3)
            f = ([c, zero, zero], [b, a, zero])
            first = False
        else:
            f = f12_mul_line(f12_sqr(f), c, b, a)
        d = digits[i - 1]
        if d:
            r, (a, b, c) = line_add(r, (qx, qy if d > 0 else mqy), r2, P)
            f = f12_mul_line(f, c, b, a)
    q1 = (qx.conjmulc(R.XI_P_1_3), qy.conjmulc(R.XI_P_1_2))
    mq2 = (qx.mulfpc(R.f2pow(R.XI, (R.p * R.p - 1) // 3)[0]), qy)
    r, (a, b, c) = line_add(r, q1, q1[1].sqr(), P)
    f = f12_mul_line(f, c, b, a)
    r, (a, b, c) = line_add(r, mq2, mq2[1].sqr(), P)
    f = f12_mul_line(f, c, b, a)
    # G2 membership for free: r is now [6u+2]Q + pi(Q) - pi^2(Q), which equals -pi^3(Q) exactly when Q has order n
    # (6u+2 + p - p^2 + p^3 = 0 mod n is the relation the optimal ate pairing is built on; the resultant of
    # X^3 - X^2 + X + 6u+2 and the characteristic polynomial of pi shares only the factor n with #E'(Fp2), checked in
    # tests/test_oracle.py, so the converse holds on this curve). An exceptional case of the incomplete addition
    # formulas, which only a point outside G2 can trigger, leaves z = 0 for good, and z != 0 is part of the test.
    g3 = frob_consts(3)
    q3 = (qx.conjmulc(g3[2]), qy.conjmulc(g3[3]))
    rx, ry, rz, rt = r
    sg = chk(rx - q3[0] * rt, ry + q3[1] * (rt * rz), rz)
    e = final_exp(f)
    names = ['E00', 'E01', 'E02', 'E10', 'E11', 'E12']
    vals = e[0] + e[1]
    for n, v in zip(names, vals):
        g.out(n, v.n)
    g.out('SG', sg.n)


def build_fp12_mul(g):
    """A <- A * B in Fp12 (tower order c0.c0 c0.c1 c0.c2 c1.c0 c1.c1 c1.c2, as csrc/tower.cuh fp12_mul). The A slots are
    pinned: main() appends copy phases that move the product back into them, so a kernel can multiply a running value
    by one table entry after another (fixed-base GT exponentiation) without touching slot numbers."""
    a = [S(g, g.inp('A%d' % k)) for k in range(6)]
    b = [S(g, g.inp('B%d' % k)) for k in range(6)]
    g.pinned = [x.n.id for x in a]
    c = f12_mul((a[0:3], a[3:6]), (b[0:3], b[3:6]))
    for k, v in enumerate(c[0] + c[1]):
        g.out('C%d' % k, v.n)


def append_copy_back(g, phases, slot_of, G):
    """C_k -> slot of A_k, after everything else (every A_k is dead by then)."""
    todo = []
    for k in range(6):
        n = Node(g, 'COPY', g.outputs['C%d' % k])
        slot_of[n.id] = slot_of[g.inputs['A%d' % k].id]
        todo.append(n)
    for i in range(0, 6, G):
        phases.append(('COPY', todo[i:i + G]))


# ---------------------------------------------------------------------------------------------- linear-op fusion
def fuse(g):
    """Peephole fusion of chains of cheap linear operations into single micro-ops (fewer phases, fewer slot
    round trips): a-b-c, a+xi*b, 3a, 4a, 8a, 3a-2b, 3a+2b. Only single-use intermediates are folded."""
    def users(n):
        return [u_ for u_ in n.users if not getattr(u_, '_dead', False)]

    def rewire(n, op, a, b=None, c=None):
        for x in (n.a, n.b, n.c):
            if x is not None and n in x.users:
                x.users.remove(n)
        n.op, n.a, n.b, n.c = op, a, b, c
        for x in uniq(a, b, c):
            if n not in x.users:
                x.users.append(n)

    changed = True
    while changed:
        changed = False
        for n in g.nodes:
            if getattr(n, '_dead', False):
                continue
            a, b = n.a, n.b
            if n.op == 'SUB' and a is not None and a.op == 'SUB' and users(a) == [n] and a is not b:
                a._dead = True
                rewire(n, 'SUB2', a.a, a.b, b)
                changed = True
            elif n.op == 'ADD' and a is not b and b.op == 'MULXI' and users(b) == [n]:
                b._dead = True
                rewire(n, 'AXI', a, b.a)
                changed = True
            elif n.op == 'ADD' and a is not b and a.op == 'MULXI' and users(a) == [n]:
                a._dead = True
                rewire(n, 'AXI', b, a.a)
                changed = True
            elif n.op == 'ADD' and a is b and a.op == 'ADD' and a.a is a.b and users(a) == [n]:
                a._dead = True
                rewire(n, 'X4', a.a)
                changed = True
            elif n.op == 'ADD' and a is b and a.op == 'X4' and users(a) == [n]:
                a._dead = True
                rewire(n, 'X8', a.a)
                changed = True
            elif n.op == 'ADD' and a is not b and a.op == 'ADD' and a.a is a.b and users(a) == [n]:
                d = a.a  # n = 2d + b
                if d is b:
                    a._dead = True
                    rewire(n, 'X3', b)
                    changed = True
                elif d.op == 'SUB' and d.a is b and users(d) == [a]:
                    a._dead = True
                    d._dead = True
                    rewire(n, 'CYCM', b, d.b)      # 2(b - g) + b = 3b - 2g
                    changed = True
                elif d.op == 'ADD' and d.a is not d.b and (d.a is b or d.b is b) and users(d) == [a]:
                    other = d.b if d.a is b else d.a
                    a._dead = True
                    d._dead = True
                    rewire(n, 'CYCP', b, other)    # 2(b + g) + b = 3b + 2g
                    changed = True
    for n in g.nodes:
        n.users = [u_ for u_ in n.users if not getattr(u_, '_dead', False)]


# ---------------------------------------------------------------------------------------------- evaluation
def eval_node(n, val, g, P):
    a = val[n.a.id] if n.a is not None else None
    b = val[n.b.id] if n.b is not None else None
    c = val[n.c.id] if n.c is not None else None
    op = n.op
    if op == 'SUB2': return R.f2sub(R.f2sub(a, b), c)
    if op == 'AXI': return R.f2add(a, R.f2mul(b, R.XI))
    if op == 'X3': return R.f2scal(a, 3)
    if op == 'X4': return R.f2scal(a, 4)
    if op == 'X8': return R.f2scal(a, 8)
    if op == 'CYCM': return R.f2sub(R.f2scal(a, 3), R.f2scal(b, 2))
    if op == 'CYCP': return R.f2add(R.f2scal(a, 3), R.f2scal(b, 2))
    if op == 'MUL': return R.f2mul(a, b)
    if op == 'ADD': return R.f2add(a, b)
    if op == 'SUB': return R.f2sub(a, b)
    if op == 'DBL': return R.f2add(a, a)
    if op == 'NEG': return R.f2neg(a)
    if op == 'MULXI': return R.f2mul(a, R.XI)
    if op == 'CONJ': return R.f2conj(a)
    if op == 'MULC': return R.f2mul(a, g.consts2[n.imm])
    if op == 'CONJMULC': return R.f2mul(R.f2conj(a), g.consts2[n.imm])
    if op == 'MULFPC': return R.f2scal(a, g.constsp[n.imm])
    if op == 'MULFP0': return R.f2scal(a, b[0])
    if op == 'MULFP1': return R.f2scal(a, b[1])
    if op == 'INV': return R.f2inv(a) if a != (0, 0) else (0, 0)   # the device inverse maps 0 to 0
    if op == 'COPY': return a
    if op == 'CHK': return (0, 0) if (a == (0, 0) and b == (0, 0) and c != (0, 0)) else (1, 0)
    raise ValueError(op)


# ---------------------------------------------------------------------------------------------- scheduling
def schedule(g, G, nslots, window=96):
    nodes = [n for n in g.nodes]
    # dead-code elimination from outputs
    live = set()
    stack = [n for n in g.outputs.values()]
    while stack:
        n = stack.pop()
        if n.id in live:
            continue
        live.add(n.id)
        for x in (n.a, n.b, n.c):
            if x is not None:
                stack.append(x)
    ops = [n for n in nodes if n.id in live and n.op != 'IN' and not getattr(n, '_dead', False)]
    # critical-path priority
    prio = {}
    for n in reversed(nodes):
        if n.id not in live:
            continue
        best = 0
        for u_ in n.users:
            if u_.id in live:
                best = max(best, prio[u_.id])
        prio[n.id] = best + COST.get(n.op, 1)
    remaining_users = {n.id: sum(1 for u_ in n.users if u_.id in live) for n in nodes if n.id in live}
    out_ids = set(n.id for n in g.outputs.values()) | set(getattr(g, 'pinned', ()))   # never freed, never lent in place
    # slots: inputs pinned first
    slot_of = {}
    free = list(range(nslots - 1, -1, -1))
    for name, n in g.inputs.items():   # inputs are ordinary values: a slot while live, none if nothing reads them
        if n.id in live:
            slot_of[n.id] = free.pop()
    done = set(n.id for n in g.inputs.values())
    pending = {n.id: n for n in ops}
    phases = []
    max_used = len(slot_of)
    while pending:
        lo = min(pending)  # ids are in trace order: only look `window` ops ahead so the live set stays small
        ready = [n for i, n in pending.items() if i < lo + window and all(x is None or x.id in done for x in (n.a, n.b, n.c))]
        by_op = {}
        for n in ready:
            by_op.setdefault(n.op, []).append(n)
        # choose the opcode whose best-G ready ops carry the most critical work
        mul_ops = [o for o in by_op if o in COST]
        lin_ops = [o for o in by_op if o not in COST]
        full_mul = [o for o in mul_ops if len(by_op[o]) >= G]
        if full_mul:          # a full multiplication phase is always worth issuing
            op = max(full_mul, key=lambda o: max(prio[n.id] for n in by_op[o]))
        elif lin_ops:         # otherwise drain the cheap linear ops first: they unlock further products
            op = max(lin_ops, key=lambda o: (len(by_op[o]) >= G, max(prio[n.id] for n in by_op[o]), len(by_op[o])))
        else:
            op = max(mul_ops, key=lambda o: (len(by_op[o]), max(prio[n.id] for n in by_op[o])))
        chosen = sorted(by_op[op], key=lambda n: -prio[n.id])[:G]
        phase_op = op
        if op == 'MUL':
            # squarings (a * a) need two Fp products instead of three, but only when the whole phase squares (the
            # lanes of a warp run one code path): fill a phase with squarings, or with none, whenever the ready list
            # allows it without leaving a lane idle that the mixed choice would have filled
            ready_mul = sorted(by_op[op], key=lambda n: -prio[n.id])
            sq = [n for n in ready_mul if n.a is n.b]
            nsq = [n for n in ready_mul if n.a is not n.b]
            want = min(G, len(ready_mul))
            if G <= 4 and len(sq) >= want:      # (with eight lanes the regrouping costs more phases than it saves)
                chosen = sq[:want]
            elif G <= 4 and len(nsq) >= want:
                chosen = nsq[:want]
            if all(n.a is n.b for n in chosen):
                phase_op = 'SQR'
        # shared-memory banks: a slot's 16-byte chunks are rotated by (slot & 3), so the four lanes of a group are
        # conflict-free when the slots they touch at the same time differ mod 4. Pick the lane order that spreads the
        # operand slots best, then give every lane a destination slot of its own class when one is free.
        import itertools

        CLS = 4 if G <= 4 else 8

        def conflicts(order):
            c = 0
            for sel in ('a', 'b', 'c'):
                seen = {}
                for n in order:
                    if n is None:
                        continue
                    x = getattr(n, sel)
                    if x is None or (sel == 'b' and n.op in ('MULC', 'CONJMULC', 'MULFPC')):
                        continue
                    k = slot_of[x.id] & (CLS - 1)
                    seen[k] = seen.get(k, 0) + 1
                c += sum(v - 1 for v in seen.values())
            return c
        CLS = 4 if G <= 4 else 8
        padded = chosen + [None] * (G - len(chosen))
        if G <= 4:
            padded = list(min(itertools.permutations(padded), key=conflicts))
        lanes = []
        to_free = []
        # operands read in this phase, by how many of the phase's ops
        readers = {}
        for n in padded:
            if n is None:
                continue
            for x in uniq(n.a, n.b, n.c):
                readers[x.id] = readers.get(x.id, 0) + 1
        for lane_id, n in enumerate(padded):
            if n is None:
                lanes.append(None)
                continue
            # in-place: an operand that dies here and that no other lane reads this phase lends its slot to the result
            s = None
            for x in (n.a, n.b, n.c):
                if x is not None and x.id not in out_ids and remaining_users[x.id] == 1 \
                        and readers[x.id] == 1 and slot_of[x.id] not in [slot_of.get(m.id) for m in lanes if m is not None]:
                    s = slot_of[x.id]
                    remaining_users[x.id] = -1  # consumed, slot handed over
                    break
            if s is None:
                if not free:
                    raise RuntimeError('out of slots (%d)' % nslots)
                pick = None
                for idx in range(len(free) - 1, -1, -1):
                    if free[idx] & (CLS - 1) == lane_id & (CLS - 1):
                        pick = idx
                        break
                s = free.pop(pick if pick is not None else -1)
            slot_of[n.id] = s
            max_used = max(max_used, max(slot_of.values()) + 1)
            lanes.append(n)
            for x in uniq(n.a, n.b, n.c):
                if remaining_users[x.id] == -1:
                    continue
                remaining_users[x.id] -= 1
                if remaining_users[x.id] == 0 and x.id not in out_ids:
                    to_free.append(slot_of[x.id])
        phases.append((phase_op, lanes))
        for n in chosen:
            done.add(n.id)
            del pending[n.id]
        for s in to_free:   # reusable from the next phase on (no write-after-read inside a phase)
            free.append(s)
    return phases, slot_of, max_used


def fix_user_counts(g):
    """A node used twice by the same consumer (a*a) must count once per consumer."""
    for n in g.nodes:
        seen = []
        for u_ in n.users:
            if u_ not in seen:
                seen.append(u_)
        n.users = seen


def run_program(g, phases, slot_of, nslots, inputs, want_slots=False):
    slots = [None] * nslots
    for name, n in g.inputs.items():
        if n.id in slot_of:
            slots[slot_of[n.id]] = inputs[name]
    P = inputs.get('P')
    for op, lanes in phases:
        lanes = [n for n in lanes if n is not None]
        reads = []
        for n in lanes:
            a = slots[slot_of[n.a.id]] if n.a is not None else None
            b = slots[slot_of[n.b.id]] if n.b is not None else None
            c = slots[slot_of[n.c.id]] if n.c is not None else None
            reads.append((a, b, c))
        writes = []
        for n, (a, b, c) in zip(lanes, reads):
            class V:  # minimal adapter for eval_node
                pass
            val = {}
            if n.a is not None:
                val[n.a.id] = a
            if n.b is not None:
                val[n.b.id] = b
            if n.c is not None:
                val[n.c.id] = c
            writes.append((slot_of[n.id], eval_node(n, val, g, P)))
        for s, v in writes:
            slots[s] = v
    if want_slots:
        return slots
    return {name: slots[slot_of[n.id]] for name, n in g.outputs.items()}


def main_gtmul(G, nslots, window):
    """python gen.py gtmul G nslots window [--emit]: the Fp12 multiplication program (csrc/gtmul_ucode_g<G>.inc)."""
    g = Graph()
    build_fp12_mul(g)
    fix_user_counts(g)
    fuse(g)
    phases, slot_of, used = schedule(g, G, nslots, window)
    append_copy_back(g, phases, slot_of, G)
    nops = sum(1 for _, l in phases for n in l if n is not None)
    print('gtmul G=%d: %d ops in %d phases, %d slots used' % (G, nops, len(phases), used))
    import random
    rng = random.Random(7)
    for _ in range(3):
        A = [(rng.randrange(R.p), rng.randrange(R.p)) for _ in range(6)]   # w^0 .. w^5
        B = [(rng.randrange(R.p), rng.randrange(R.p)) for _ in range(6)]
        tower = lambda w: [w[0], w[2], w[4], w[1], w[3], w[5]]
        inputs = {}
        for k, (x, y) in enumerate(zip(tower(A), tower(B))):
            inputs['A%d' % k], inputs['B%d' % k] = x, y
        slots = run_program(g, phases, slot_of, nslots, inputs, want_slots=True)
        got = [slots[slot_of[g.inputs['A%d' % k].id]] for k in range(6)]
        assert got == tower(R.f12mul(A, B)), 'gtmul program disagrees with the definition'
    print('gtmul program == bn256_ref.f12mul on random inputs (result read back from the A slots)')
    if '--emit' in sys.argv:
        emit(g, phases, slot_of, used, G, nslots, window, name='gtmul')
    return g, phases, slot_of, used


def main():
    if len(sys.argv) > 1 and sys.argv[1] == 'gtmul':
        return main_gtmul(int(sys.argv[2]), int(sys.argv[3]), int(sys.argv[4]))
    G = int(sys.argv[1]) if len(sys.argv) > 1 else 4
    nslots = int(sys.argv[2]) if len(sys.argv) > 2 else 96
    window = int(sys.argv[3]) if len(sys.argv) > 3 else 96
    g = Graph()
    build_pairing(g)
    fix_user_counts(g)
    fuse(g)
    phases, slot_of, used = schedule(g, G, nslots, window)
    cnt = lambda l: sum(1 for n in l if n is not None)
    nops = sum(cnt(l) for _, l in phases)
    nmul = sum(cnt(l) for o, l in phases if o in COST)
    mulph = sum(1 for o, _ in phases if o in COST)
    print('G=%d: %d ops (%d mul-class) in %d phases (%d mul-class phases), %d slots used' % (G, nops, nmul, len(phases), mulph, used))
    print('mul-class lane utilisation %.2f' % (nmul / (mulph * G)))
    # execute and compare with the definition
    import random
    rng = random.Random(5)
    for _ in range(1 if '--quick' in sys.argv else 2):
        a, b = rng.randrange(1, R.n), rng.randrange(1, R.n)
        Pp, Q = R.g1_mul(a, R.G1_GEN), R.g2_mul(b, R.G2_GEN)
        out = run_program(g, phases, slot_of, nslots, {'P': ((-Pp[0]) % R.p, Pp[1]), 'QX': Q[0], 'QY': Q[1], 'ONE': (1, 0), 'ZERO': (0, 0)})
        e = R.pair(Pp, Q)
        got = [out['E00'], out['E10'], out['E01'], out['E11'], out['E02'], out['E12']]
        assert got == e, 'micro-program disagrees with the definition'
        assert out['SG'] == (0, 0), 'subgroup verdict wrong for a point of G2'
    print('micro-program == bn256_ref.pair on random inputs')
    Pp = R.g1_mul(7, R.G1_GEN)
    T = R.g2_twist_point(1)
    outside = [T, R.g2_mul_raw(R.n, T), R.g2_mul_raw(R.n * R.TWIST_COFACTOR // 13, T), R.g2_mul_raw(R.n * R.TWIST_COFACTOR // (13 * 7369), T)]
    for Q in outside:
        assert Q is not None and R.g2_on_curve(Q) and not R.g2_in_subgroup(Q)
        out = run_program(g, phases, slot_of, nslots, {'P': ((-Pp[0]) % R.p, Pp[1]), 'QX': Q[0], 'QY': Q[1], 'ONE': (1, 0), 'ZERO': (0, 0)})
        assert out['SG'] == (1, 0), 'subgroup verdict wrong for a twist point outside G2'
    print('subgroup verdict: rejects twist points of composite, cofactor, 13 and 7369-multiple order')
    if '--emit' in sys.argv:
        emit(g, phases, slot_of, used, G, nslots, window)
    return g, phases, slot_of, used, G


def limbs(x):
    return ', '.join('0x%08xu' % ((x >> (32 * k)) & 0xffffffff) for k in range(8))


def mont(x):
    return (x % R.p) * (1 << 256) % R.p


def slot_offset(s_, G):
    """Byte offset of a slot inside an item with the swizzle folded in. With two threads per program lane the 8 (G = 4) or
    16 (G = 8) threads of an item read, in one instruction, chunk 2h + k of their lane's slot; bank group of that access =
    4 (slot & 1) + ((2h + k) ^ swz(slot)). For lanes whose slots differ mod 4 (mod 8) the rotation below makes the groups
    of one item all distinct (each twice for G = 8): no replays. (slot & 3, the round-1 rule for one thread per lane, put
    slots s and s + 2 of the two halves on the same groups: ncu counted 2.4x the ideal wavefronts.)"""
    rot = (s_ >> 1) & 1 if G <= 4 else (s_ >> 1) & 3
    return s_ * 64 + rot * 16


def prog_words(phases, slot_of, G):
    """The emitted lane words, one list of G (x, y) pairs per phase, closing idle phases included."""
    out = []
    for op, lanes in list(phases) + [('NOP', [])] * PAD_PHASES:
        words = []
        for k in range(G):
            if k < len(lanes) and lanes[k] is not None:
                n = lanes[k]
                a = slot_offset(slot_of[n.a.id], G) if n.a is not None else 0
                b = slot_offset(slot_of[n.b.id], G) if n.b is not None else n.imm
                c = slot_offset(slot_of[n.c.id], G) if n.c is not None else 0
                d = slot_offset(slot_of[n.id], G)
                assert max(a, b, c, d) < 65535
                words.append((d | a << 16, b | c << 16))
            else:
                words.append((0xffffffff, 0xffffffff))
        out.append(words)
    return out


def emit(g, phases, slot_of, used, G, nslots_arg, window, name='pairing'):
    path = os.path.join(ROOT, 'drynx_b200', 'csrc', '%s_ucode_g%d.inc' % (name, G))
    with open(path, 'w') as f:
        f.write('// GENERATED by tools/microcode/gen.py %s%d %d ... --emit -- do not edit.\n' % ('' if name == 'pairing' else name + ' ', G, used))
        what = 'the optimal ate pairing plus the G2 membership verdict' if name == 'pairing' else 'one Fp12 multiplication A <- A * B'
        f.write('// Lane-group micro-program of %s: %d Fp2 operations in\n'
                '// %d phases for %d program lanes,\n'
                % (what, sum(1 for _, l in phases for n in l if n is not None), len(phases), G))
        f.write('// %d Fp2 slots per item. Executed with Python integers against the definition of the pairing by the generator before emission.\n' % used)
        used |= 1  # odd slot count per item: adjacent items then sit in opposite bank halves
        f.write('#define UC_G %d\n#define UC_NSLOTS %d\n#define UC_NPHASES %d\n' % (G, used, len(phases)))
        f.write('// generator arguments (tests/test_microcode.py regenerates the program from them)\n')
        f.write('#define UC_GEN_SLOTS %d\n#define UC_GEN_WINDOW %d\n' % (nslots_arg, window))
        f.write('// A slot is 64 bytes = four 16-byte chunks; logical chunk c of slot s lives at physical chunk c ^ UC_SWZ(s)\n')
        f.write('#define UC_SWZ(s) %s\n' % ('(((s) >> 1) & 1u)' if G <= 4 else '(((s) >> 1) & 3u)'))
        for i, name in enumerate(OPS):
            f.write('#define UC_%s %d\n' % (name, i))
        f.write('// uc_ops byte = opcode | class << 5: the interpreter walks the class bits first (7, then 6, then 5), so the\n'
                '// frequent operations are two or three tests away instead of the five of a balanced search over all opcodes\n')
        for name, cls in sorted(OP_CLASS.items(), key=lambda kv: -kv[1]):
            f.write('#define UC_CLS_%s %d\n' % (name, cls))
        for name, n in g.inputs.items():
            if n.id in slot_of:
                f.write('#define UC_IN_%s %d\n' % (name, slot_of[n.id]))
        for name, n in g.outputs.items():
            f.write('#define UC_OUT_%s %d\n' % (name, slot_of[n.id]))
        f.write('__device__ const uint32_t uc_const2[%d][16] = {\n' % max(1, len(g.consts2)))
        for c in g.consts2:
            f.write('  {%s, %s},\n' % (limbs(mont(c[0])), limbs(mont(c[1]))))
        f.write('};\n__device__ const uint32_t uc_constp[%d][8] = {\n' % max(1, len(g.constsp)))
        for c in g.constsp:
            f.write('  {%s},\n' % limbs(mont(c)))
        f.write('};\n// one opcode per phase (warp-uniform: read through the constant cache); both tables end with UC_PAD_PHASES idle\n'
                '// phases so that the interpreter can read and prefetch a few phases ahead without a bound check\n')
        f.write('#define UC_PAD_PHASES %d\n' % PAD_PHASES)
        f.write('__constant__ uint8_t uc_ops[UC_NPHASES + UC_PAD_PHASES] = {\n')
        for k in range(0, len(phases), 32):
            f.write('  ' + ', '.join(str(op_byte(o)) for o, _ in phases[k:k + 32]) + ',\n')
        f.write('  ' + ', '.join([str(op_byte('NOP'))] * PAD_PHASES) + ',\n')
        f.write('};\n// per program lane: x = dst | a << 16, y = b | c << 16, each the BYTE offset of the slot inside the item with the\n'
                '// swizzle folded in (slot * 64 + UC_SWZ(slot) * 16); x = 0xffffffff: idle; b = constant index for MULC / CONJMULC / MULFPC\n')
        f.write('__device__ const uint2 uc_prog[(UC_NPHASES + UC_PAD_PHASES) * UC_G] = {\n')
        for words in prog_words(phases, slot_of, G):
            f.write('  ' + ', '.join('{0x%08xu, 0x%08xu}' % w for w in words) + ',\n')
        f.write('};\n')
    print('wrote', path, os.path.getsize(path), 'bytes')


if __name__ == '__main__':
    main()
