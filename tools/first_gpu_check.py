"""First on-GPU sanity run: field ops vs Python ints, ElGamal round trip vs the big-int model, pairing vs model,
and the integer-pipe probes. Prints timings; exits non-zero on any mismatch."""
import os, sys, time, random
import numpy as np
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT); sys.path.insert(0, os.path.join(ROOT, 'oracle'))
import bn256_ref as R
from drynx_b200 import Backend

t0 = time.time()
be = Backend(0)
print('init %.2fs' % (time.time() - t0), flush=True)
rng = random.Random(7)
fails = 0

def limbs(x): return [(x >> (32 * k)) & 0xffffffff for k in range(8)]
def unl(a): return sum(int(v) << (32 * k) for k, v in enumerate(a))
N = 4096
for which, mod in ((0, R.p), (1, R.n)):
    xs = [rng.randrange(mod) for _ in range(N)]; ys = [rng.randrange(mod) for _ in range(N)]
    xs[0] = mod - 1; ys[0] = mod - 1; xs[1] = 0; ys[2] = 1; xs[3] = mod - 1; ys[3] = 2**255
    a = np.array([limbs(x) for x in xs], dtype=np.uint32).ravel(); b = np.array([limbs(y) for y in ys], dtype=np.uint32).ravel()
    r = be.mont_mul(which, a, b).reshape(N, 8)
    rinv = pow(1 << 256, -1, mod)
    bad = sum(1 for i in range(N) if unl(r[i]) != xs[i] * ys[i] * rinv % mod)
    ra = be.mont_mul(2 + 2 * which, a, b).reshape(N, 8); rs = be.mont_mul(3 + 2 * which, a, b).reshape(N, 8)
    bad += sum(1 for i in range(N) if unl(ra[i]) != (xs[i] + ys[i]) % mod or unl(rs[i]) != (xs[i] - ys[i]) % mod)
    print('mont/add/sub mod', 'pn'[which], 'bad =', bad, flush=True); fails += bad

sk = rng.randrange(1, R.n); P = R.g1_mul(sk, R.G1_GEN)
be.set_pubkey(R.g1_to_bytes(P))
ms = [0, 1, -1, 5, -12000, 99999, 2**40 + 3, -(2**62)] + [rng.randrange(-50000, 50000) for _ in range(24)]
rs = [rng.randrange(1, R.n) for _ in ms]; rs[3] = 0
ct = be.encrypt_int_vector(ms, b''.join(x.to_bytes(32, 'big') for x in rs)).tobytes()
bad = sum(1 for i, (m, r) in enumerate(zip(ms, rs)) if ct[128 * i:128 * i + 128] != R.encrypt(P, m, r))
print('encrypt bad =', bad, flush=True); fails += bad
# aggregate 5 DPs x 32
vecs = []
allm = []
for d in range(5):
    mm = [rng.randrange(-1000, 1000) for _ in ms]; rr = [rng.randrange(1, R.n) for _ in ms]
    allm.append(mm)
    vecs.append(be.encrypt_int_vector(mm, b''.join(x.to_bytes(32, 'big') for x in rr)))
agg = be.aggregate(vecs)
be.create_decryption_table(10000)
dec = be.decrypt_int_vector(sk.to_bytes(32, 'big'), agg)
want = [sum(allm[d][j] for d in range(5)) for j in range(len(ms))]
print('aggregate+decrypt ok =', list(dec) == want, flush=True); fails += list(dec) != want
dec2, found = be.decrypt_int_vector(sk.to_bytes(32, 'big'), ct, return_found=True)
want2 = [m if abs(m) < 100000 else 0 for m in ms]
print('decrypt ok =', list(dec2) == want2, list(found), flush=True); fails += list(dec2) != want2
# pairing
a, b = rng.randrange(1, R.n), rng.randrange(1, R.n)
g1 = R.g1_to_bytes(R.g1_mul(a, R.G1_GEN)); g2 = R.g2_to_bytes(R.g2_mul(b, R.G2_GEN))
gt = be.pair(g1 + R.g1_to_bytes(R.G1_GEN), g2 + R.g2_to_bytes(R.G2_GEN)).tobytes()
e = R.pair(R.G1_GEN, R.G2_GEN)
ok = gt[384:] == R.f12_to_bytes(e) and gt[:384] == R.f12_to_bytes(R.f12pow(e, a * b % R.n))
print('pairing ok =', ok, flush=True); fails += not ok
k = rng.randrange(R.n)
ok = be.g2_mul(k.to_bytes(32, 'big'), g2).tobytes() == R.g2_to_bytes(R.g2_mul(k * b % R.n, R.G2_GEN))
ok &= be.g1_mul(k.to_bytes(32, 'big'), g1).tobytes() == R.g1_to_bytes(R.g1_mul(k * a % R.n, R.G1_GEN))
print('g1/g2 mul ok =', ok, flush=True); fails += not ok

# probes
for kind, name in ((0, 'IMAD.WIDE chains'), (1, 'montgomery products')):
    for iters in (2000, 20000):
        ms_, work = be.microbench(kind, iters)
        print('%s iters=%d: %.3f ms, %.3f T/s' % (name, iters, ms_, work / ms_ / 1e9), flush=True)
# throughput of kernels via host API (includes copies)
import torch
for n in (1 << 14, 1 << 18):
    m = np.arange(n, dtype=np.int64) % 1000; r = np.frombuffer(os.urandom(32 * n), dtype=np.uint8).copy(); r[::32] &= 0x7f
    t = time.time(); c = be.encrypt_int_vector(m, r); dt = time.time() - t
    print('encrypt n=%d host-API %.1f ms -> %.2f M enc/s' % (n, dt * 1e3, n / dt / 1e6), flush=True)
n = 4096
g1s = g1 * n; g2s = g2 * n
t = time.time(); be.pair(g1s, g2s); dt = time.time() - t
print('pair n=%d %.1f ms -> %.1f k pairings/s' % (n, dt * 1e3, n / dt / 1e3), flush=True)
n = 1 << 15
g1s = g1 * n; g2s = g2 * n
t = time.time(); be.pair(g1s, g2s); dt = time.time() - t
print('pair n=%d %.1f ms -> %.1f k pairings/s' % (n, dt * 1e3, n / dt / 1e3), flush=True)
n = 1 << 16
cts = (ct[:128] * n)
t = time.time(); be.decrypt_int_vector(sk.to_bytes(32, 'big'), cts); dt = time.time() - t
print('decrypt n=%d %.1f ms -> %.2f M dec/s' % (n, dt * 1e3, n / dt / 1e6), flush=True)
print('launches', be.launches)
print('FAILS', fails)
sys.exit(1 if fails else 0)
