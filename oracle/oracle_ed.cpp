// TEST INFRASTRUCTURE ONLY -- CPU restatement of the reference's algorithms for ElGamal over edwards25519
// (UnLynx's default suite, the curve BASELINE.json's north_star names).
//
// The arithmetic is in the un-vendored modules go.dedis.ch/kyber/v3 v3.0.8 `group/edwards25519` (fe.go, ge.go,
// point.go, scalar.go -- itself a port of the ref10 code of SUPERCOP) and github.com/ldsec/unlynx v1.4.0
// `lib/crypto.go` (go.mod:14,26); neither is under /root/reference and there is no Go toolchain. This file
// restates, as recalled:
//   * fe.go: field elements as ten signed limbs of alternating 26/25 bits, feMul with the 19-fold wrap and the
//     doubled odd-odd terms, feInvert / fePow22523 addition chains, feToBytes canonical reduction
//   * ge.go: extended / projective / completed / precomputed / cached group elements and the ref10 formulas
//     (geAdd, geSub, geMixedAdd, p2 doubling), geScalarMultBase (signed radix-16, table of (j+1)*256^i*B) and
//     geScalarMult (signed radix-16 over cached multiples 1..8 of the point), FromBytes / ToBytes
//   * scalar.go: 32-byte little-endian scalars reduced mod l; SetInt64 wraps negatives
//   * unlynx lib/crypto.go: encryptPoint, IntToPoint, CipherVector.Add, decryptPoint, discreteLog
// *** parity pinned (group arithmetic) ***: tests/test_oracle_ed.py checks this file against RFC 8032 known
// answers and against the big-integer definition oracle/ed25519_ref.py; the scalar byte order and the K || C
// ciphertext layout are recalled conventions.
// Only tests/, __graft_entry__.smoke() and bench.py's cpu_baseline / --impl reference legs may load this.
#include <atomic>
#include <cstdint>
#include <cstring>
#include <mutex>
#include <string>
#include <thread>
#include <unordered_map>
#include <vector>

namespace {

// ------------------------------------------------------------------ fe.go
struct fe {
  int32_t v[10];
};
static inline int shift_of(int i) { return (i & 1) ? 25 : 26; }

static void feZero(fe& h) { memset(h.v, 0, sizeof(h.v)); }
static void feOne(fe& h) { feZero(h); h.v[0] = 1; }
static void feAdd(fe& h, const fe& f, const fe& g) { for (int i = 0; i < 10; i++) h.v[i] = f.v[i] + g.v[i]; }
static void feSub(fe& h, const fe& f, const fe& g) { for (int i = 0; i < 10; i++) h.v[i] = f.v[i] - g.v[i]; }
static void feNeg(fe& h, const fe& f) { for (int i = 0; i < 10; i++) h.v[i] = -f.v[i]; }

// carry chain on 64-bit accumulators: leaves |h_i| <= 2^25 / 2^24 like ref10's interleaved carries
static void feCombine(fe& out, int64_t h[10]) {
  for (int pass = 0; pass < 2; pass++) {
    for (int i = 0; i < 10; i++) {
      const int s = shift_of(i);
      const int64_t c = (h[i] + ((int64_t)1 << (s - 1))) >> s;
      h[i] -= c << s;
      if (i == 9) h[0] += c * 19; else h[i + 1] += c;
    }
  }
  const int64_t c = (h[0] + ((int64_t)1 << 25)) >> 26;
  h[0] -= c << 26;
  h[1] += c;
  for (int i = 0; i < 10; i++) out.v[i] = (int32_t)h[i];
}
// h = f * g: limb i has weight 2^ceil(25.5 i); a product of two odd limbs is doubled, terms past limb 9 wrap
// with the factor 19 (2^255 = 19 mod p)
static void feMul(fe& out, const fe& f, const fe& g) {
  int64_t h[10] = {0};
  for (int i = 0; i < 10; i++)
    for (int j = 0; j < 10; j++) {
      int64_t t = (int64_t)f.v[i] * g.v[j];
      if ((i & 1) && (j & 1)) t *= 2;
      int k = i + j;
      if (k >= 10) { k -= 10; t *= 19; }
      h[k] += t;
    }
  feCombine(out, h);
}
static void feSquare(fe& out, const fe& f) { feMul(out, f, f); }
static void feSquareN(fe& out, const fe& f, int n) {
  feSquare(out, f);
  for (int i = 1; i < n; i++) feSquare(out, out);
}
// z^(2^250 - 1) and z^11, the shared prefix of both addition chains
static void fePow250(fe& t250, fe& z11, const fe& z) {
  fe t0, t1, t2, t3;
  feSquare(t0, z);             // 2
  feSquareN(t1, t0, 2);        // 8
  feMul(t1, z, t1);            // 9
  feMul(t0, t0, t1);           // 11
  z11 = t0;
  feSquare(t2, t0);            // 22
  feMul(t1, t1, t2);           // 2^5 - 1
  feSquareN(t2, t1, 5);
  feMul(t1, t2, t1);           // 2^10 - 1
  feSquareN(t2, t1, 10);
  feMul(t2, t2, t1);           // 2^20 - 1
  feSquareN(t3, t2, 20);
  feMul(t2, t3, t2);           // 2^40 - 1
  feSquareN(t2, t2, 10);
  feMul(t1, t2, t1);           // 2^50 - 1
  feSquareN(t2, t1, 50);
  feMul(t2, t2, t1);           // 2^100 - 1
  feSquareN(t3, t2, 100);
  feMul(t2, t3, t2);           // 2^200 - 1
  feSquareN(t2, t2, 50);
  feMul(t250, t2, t1);         // 2^250 - 1
}
static void feInvert(fe& out, const fe& z) {  // z^(p-2) = z^(2^255 - 21)
  fe t, z11;
  fePow250(t, z11, z);
  feSquareN(t, t, 5);
  feMul(out, t, z11);
}
static void fePow22523(fe& out, const fe& z) {  // z^((p-5)/8) = z^(2^252 - 3)
  fe t, z11;
  fePow250(t, z11, z);
  feSquareN(t, t, 2);
  feMul(out, t, z);
}
static void feFromBytes(fe& h, const uint8_t s[32]) {  // ignores bit 255, does not reject values >= p
  int64_t a[10];
  int bit = 0;
  for (int i = 0; i < 10; i++) {
    const int s_i = shift_of(i);
    int64_t x = 0;
    for (int b = 0; b < s_i; b++) {
      const int pos = bit + b;
      if (pos < 255) x |= (int64_t)((s[pos >> 3] >> (pos & 7)) & 1) << b;
    }
    a[i] = x;
    bit += s_i;
  }
  feCombine(h, a);
}
static void feToBytes(uint8_t s[32], const fe& f) {
  int32_t h[10];
  for (int i = 0; i < 10; i++) h[i] = f.v[i];
  int32_t q = (19 * h[9] + ((int32_t)1 << 24)) >> 25;
  for (int i = 0; i < 10; i++) q = (h[i] + q) >> shift_of(i);
  h[0] += 19 * q;
  for (int i = 0; i < 10; i++) {
    const int s_i = shift_of(i);
    const int32_t c = h[i] >> s_i;
    h[i] -= c << s_i;
    if (i < 9) h[i + 1] += c;  // the carry out of limb 9 is the multiple of 2^255 being dropped
  }
  memset(s, 0, 32);
  int bit = 0;
  for (int i = 0; i < 10; i++) {
    const int s_i = shift_of(i);
    for (int b = 0; b < s_i; b++)
      if ((h[i] >> b) & 1) s[(bit + b) >> 3] |= (uint8_t)(1u << ((bit + b) & 7));
    bit += s_i;
  }
}
static int feIsNegative(const fe& f) { uint8_t s[32]; feToBytes(s, f); return s[0] & 1; }
static int feIsNonZero(const fe& f) {
  uint8_t s[32];
  feToBytes(s, f);
  uint8_t x = 0;
  for (int i = 0; i < 32; i++) x |= s[i];
  return x != 0;
}

static fe fe_d, fe_d2, fe_sqrtM1;

// ------------------------------------------------------------------ ge.go
struct projectiveGE { fe X, Y, Z; };
struct extendedGE { fe X, Y, Z, T; };
struct completedGE { fe X, Y, Z, T; };
struct preComputedGE { fe yPlusX, yMinusX, xy2d; };
struct cachedGE { fe yPlusX, yMinusX, Z, T2d; };

static void extZero(extendedGE& p) { feZero(p.X); feOne(p.Y); feOne(p.Z); feZero(p.T); }
static void extToCached(cachedGE& r, const extendedGE& p) {
  feAdd(r.yPlusX, p.Y, p.X);
  feSub(r.yMinusX, p.Y, p.X);
  r.Z = p.Z;
  feMul(r.T2d, p.T, fe_d2);
}
static void extToProjective(projectiveGE& r, const extendedGE& p) { r.X = p.X; r.Y = p.Y; r.Z = p.Z; }
static void completedToProjective(projectiveGE& r, const completedGE& p) {
  feMul(r.X, p.X, p.T);
  feMul(r.Y, p.Y, p.Z);
  feMul(r.Z, p.Z, p.T);
}
static void completedToExtended(extendedGE& r, const completedGE& p) {
  feMul(r.X, p.X, p.T);
  feMul(r.Y, p.Y, p.Z);
  feMul(r.Z, p.Z, p.T);
  feMul(r.T, p.X, p.Y);
}
static void projDouble(completedGE& r, const projectiveGE& p) {
  fe t0;
  feSquare(r.X, p.X);
  feSquare(r.Z, p.Y);
  feSquare(r.T, p.Z);
  feAdd(r.T, r.T, r.T);
  feAdd(r.Y, p.X, p.Y);
  feSquare(t0, r.Y);
  feAdd(r.Y, r.Z, r.X);
  feSub(r.Z, r.Z, r.X);
  feSub(r.X, t0, r.Y);
  feSub(r.T, r.T, r.Z);
}
static void extDouble(completedGE& r, const extendedGE& p) {
  projectiveGE q;
  extToProjective(q, p);
  projDouble(r, q);
}
static void geAdd(completedGE& r, const extendedGE& p, const cachedGE& q) {
  fe t0;
  feAdd(r.X, p.Y, p.X);
  feSub(r.Y, p.Y, p.X);
  feMul(r.Z, r.X, q.yPlusX);
  feMul(r.Y, r.Y, q.yMinusX);
  feMul(r.T, q.T2d, p.T);
  feMul(r.X, p.Z, q.Z);
  feAdd(t0, r.X, r.X);
  feSub(r.X, r.Z, r.Y);
  feAdd(r.Y, r.Z, r.Y);
  feAdd(r.Z, t0, r.T);
  feSub(r.T, t0, r.T);
}
static void geMixedAdd(completedGE& r, const extendedGE& p, const preComputedGE& q) {
  fe t0;
  feAdd(r.X, p.Y, p.X);
  feSub(r.Y, p.Y, p.X);
  feMul(r.Z, r.X, q.yPlusX);
  feMul(r.Y, r.Y, q.yMinusX);
  feMul(r.T, q.xy2d, p.T);
  feAdd(t0, p.Z, p.Z);
  feSub(r.X, r.Z, r.Y);
  feAdd(r.Y, r.Z, r.Y);
  feAdd(r.Z, t0, r.T);
  feSub(r.T, t0, r.T);
}
static void extAdd(extendedGE& r, const extendedGE& a, const extendedGE& b) {  // point.Add
  cachedGE c;
  completedGE t;
  extToCached(c, b);
  geAdd(t, a, c);
  completedToExtended(r, t);
}
static void extNeg(extendedGE& r, const extendedGE& a) {
  feNeg(r.X, a.X);
  r.Y = a.Y;
  r.Z = a.Z;
  feNeg(r.T, a.T);
}
static void extToBytes(uint8_t s[32], const extendedGE& p) {
  fe recip, x, y;
  feInvert(recip, p.Z);
  feMul(x, p.X, recip);
  feMul(y, p.Y, recip);
  feToBytes(s, y);
  s[31] ^= (uint8_t)(feIsNegative(x) << 7);
}
static bool extFromBytes(extendedGE& p, const uint8_t s[32]) {
  fe u, v, v3, vxx, check;
  feFromBytes(p.Y, s);
  feOne(p.Z);
  feSquare(u, p.Y);
  feMul(v, u, fe_d);
  feSub(u, u, p.Z);  // u = y^2 - 1
  feAdd(v, v, p.Z);  // v = d y^2 + 1
  feSquare(v3, v);
  feMul(v3, v3, v);
  feSquare(p.X, v3);
  feMul(p.X, p.X, v);
  feMul(p.X, p.X, u);   // u v^7
  fePow22523(p.X, p.X);
  feMul(p.X, p.X, v3);
  feMul(p.X, p.X, u);   // u v^3 (u v^7)^((p-5)/8)
  feSquare(vxx, p.X);
  feMul(vxx, vxx, v);
  feSub(check, vxx, u);
  if (feIsNonZero(check)) {
    feAdd(check, vxx, u);
    if (feIsNonZero(check)) return false;
    feMul(p.X, p.X, fe_sqrtM1);
  }
  if (feIsNegative(p.X) != (s[31] >> 7)) feNeg(p.X, p.X);
  feMul(p.T, p.X, p.Y);
  return true;
}

// ------------------------------------------------------------------ scalars (scalar.go: 32 bytes little-endian mod l)
struct Sc { uint64_t v[4]; };  // little-endian words, always < l
static const uint64_t L_WORDS[4] = {0x5812631a5cf5d3edULL, 0x14def9dea2f79cd6ULL, 0, 0x1000000000000000ULL};
static bool geq(const uint64_t a[4], const uint64_t b[4]) {
  for (int i = 3; i >= 0; i--) {
    if (a[i] != b[i]) return a[i] > b[i];
  }
  return true;
}
static void sub_in_place(uint64_t a[4], const uint64_t b[4]) {
  unsigned __int128 borrow = 0;
  for (int i = 0; i < 4; i++) {
    unsigned __int128 t = (unsigned __int128)a[i] - b[i] - borrow;
    a[i] = (uint64_t)t;
    borrow = (t >> 64) & 1;
  }
}
static void sc_reduce(Sc& s) {  // value < 2^256 < 16 l
  for (int k = 3; k >= 0; k--) {
    uint64_t m[4];
    for (int i = 0; i < 4; i++) m[i] = (L_WORDS[i] << k) | (k && i ? L_WORDS[i - 1] >> (64 - k) : 0);
    if (geq(s.v, m)) sub_in_place(s.v, m);
  }
}
static void sc_from_bytes(Sc& s, const uint8_t b[32]) {
  for (int i = 0; i < 4; i++) {
    s.v[i] = 0;
    for (int k = 0; k < 8; k++) s.v[i] |= (uint64_t)b[8 * i + k] << (8 * k);
  }
  sc_reduce(s);
}
static Sc sc_int64(int64_t m) {  // SetInt64: negative values wrap mod l
  Sc s = {{0, 0, 0, 0}};
  if (m >= 0) { s.v[0] = (uint64_t)m; return s; }
  uint64_t a[4] = {(uint64_t)0 - (uint64_t)m, 0, 0, 0};
  memcpy(s.v, L_WORDS, 32);
  sub_in_place(s.v, a);
  return s;
}
static void sc_to_bytes(uint8_t b[32], const Sc& s) {
  for (int i = 0; i < 32; i++) b[i] = (uint8_t)(s.v[i >> 3] >> (8 * (i & 7)));
}

// signed radix-16 digits e[0..63] in [-8, 8] of a scalar < 2^255 (ge.go geScalarMultBase / geScalarMult)
static void radix16(int8_t e[64], const Sc& s) {
  uint8_t a[32];
  sc_to_bytes(a, s);
  for (int i = 0; i < 32; i++) {
    e[2 * i] = (int8_t)(a[i] & 15);
    e[2 * i + 1] = (int8_t)((a[i] >> 4) & 15);
  }
  int8_t carry = 0;
  for (int i = 0; i < 63; i++) {
    e[i] = (int8_t)(e[i] + carry);
    carry = (int8_t)((e[i] + 8) >> 4);
    e[i] = (int8_t)(e[i] - (carry << 4));
  }
  e[63] = (int8_t)(e[63] + carry);
}

// ------------------------------------------------------------------ scalar multiplication (ge.go)
static preComputedGE base_table[32][8];  // (j+1) * 256^i * B, the table ref10 ships as constants
static extendedGE ext_base;

static void to_precomputed(preComputedGE& r, const extendedGE& p) {
  fe recip, x, y, xy;
  feInvert(recip, p.Z);
  feMul(x, p.X, recip);
  feMul(y, p.Y, recip);
  feAdd(r.yPlusX, y, x);
  feSub(r.yMinusX, y, x);
  feMul(xy, x, y);
  feMul(r.xy2d, xy, fe_d2);
  // carry the sums so that table entries are reduced like the shipped constants
  fe one;
  feOne(one);
  feMul(r.yPlusX, r.yPlusX, one);
  feMul(r.yMinusX, r.yMinusX, one);
}
static void select_base(preComputedGE& t, int pos, int8_t b) {
  const int neg = b < 0;
  const int babs = neg ? -b : b;
  feOne(t.yPlusX);
  feOne(t.yMinusX);
  feZero(t.xy2d);
  if (babs) t = base_table[pos][babs - 1];
  if (neg) {
    const fe tmp = t.yPlusX;
    t.yPlusX = t.yMinusX;
    t.yMinusX = tmp;
    feNeg(t.xy2d, t.xy2d);
  }
}
static void geScalarMultBase(extendedGE& h, const Sc& a) {
  int8_t e[64];
  radix16(e, a);
  extZero(h);
  preComputedGE t;
  completedGE r;
  for (int i = 1; i < 64; i += 2) {
    select_base(t, i / 2, e[i]);
    geMixedAdd(r, h, t);
    completedToExtended(h, r);
  }
  projectiveGE s;
  extDouble(r, h);
  completedToProjective(s, r);
  projDouble(r, s);
  completedToProjective(s, r);
  projDouble(r, s);
  completedToProjective(s, r);
  projDouble(r, s);
  completedToExtended(h, r);
  for (int i = 0; i < 64; i += 2) {
    select_base(t, i / 2, e[i]);
    geMixedAdd(r, h, t);
    completedToExtended(h, r);
  }
}
static void geScalarMult(extendedGE& h, const Sc& a, const extendedGE& A) {
  cachedGE Ai[8];
  extendedGE u;
  completedGE r;
  extToCached(Ai[0], A);
  u = A;
  for (int j = 1; j < 8; j++) {
    geAdd(r, A, Ai[j - 1]);
    completedToExtended(u, r);
    extToCached(Ai[j], u);
  }
  int8_t e[64];
  radix16(e, a);
  extZero(h);
  for (int i = 63; i >= 0; i--) {
    projectiveGE s;
    extDouble(r, h);
    completedToProjective(s, r);
    projDouble(r, s);
    completedToProjective(s, r);
    projDouble(r, s);
    completedToProjective(s, r);
    projDouble(r, s);
    completedToExtended(h, r);
    const int neg = e[i] < 0, babs = neg ? -e[i] : e[i];
    if (babs) {
      cachedGE c = Ai[babs - 1];
      if (neg) {
        const fe tmp = c.yPlusX;
        c.yPlusX = c.yMinusX;
        c.yMinusX = tmp;
        feNeg(c.T2d, c.T2d);
      }
      geAdd(r, h, c);
      completedToExtended(h, r);
    }
  }
}

static std::once_flag g_once;
static void init_tables() {
  // d = -121665/121666, sqrt(-1) = 2^((p-1)/4); derived here instead of shipped as limb constants
  fe num, den, inv, two, e;
  feZero(num); num.v[0] = -121665;
  feZero(den); den.v[0] = 121666;
  feInvert(inv, den);
  feMul(fe_d, num, inv);
  feAdd(fe_d2, fe_d, fe_d);
  fe one;
  feOne(one);
  feMul(fe_d2, fe_d2, one);
  // 2^((p-1)/4) = 2^(2^253 - 5) = (2^(2^252 - 3))^2 * 2
  feZero(two); two.v[0] = 2;
  fePow22523(e, two);
  feSquare(e, e);
  feMul(fe_sqrtM1, e, two);
  // base point: y = 4/5, x even
  uint8_t b[32];
  memset(b, 0x66, 32);
  b[0] = 0x58;
  extFromBytes(ext_base, b);
  extendedGE row = ext_base;  // 256^i * B
  for (int i = 0; i < 32; i++) {
    extendedGE acc = row;
    for (int j = 0; j < 8; j++) {
      to_precomputed(base_table[i][j], acc);
      extAdd(acc, acc, row);
    }
    for (int k = 0; k < 8; k++) {
      completedGE r;
      extDouble(r, row);
      completedToExtended(row, r);
    }
  }
}
static void ensure_init() { std::call_once(g_once, init_tables); }

template <class Fn>
static void parallel_for(size_t n, int threads, Fn fn) {
  if (threads <= 1 || n <= 1) {
    for (size_t i = 0; i < n; i++) fn(i);
    return;
  }
  std::atomic<size_t> next(0);
  std::vector<std::thread> pool;
  for (int t = 0; t < threads; t++)
    pool.emplace_back([&] {
      for (;;) {
        size_t i = next.fetch_add(1);
        if (i >= n) return;
        fn(i);
      }
    });
  for (auto& th : pool) th.join();
}

// ------------------------------------------------------------------ ElGamal (unlynx lib/crypto.go)
static void encrypt_one(uint8_t ct[64], const extendedGE& pub, int64_t m, const uint8_t r32[32]) {
  Sc r;
  sc_from_bytes(r, r32);
  extendedGE K, S, M, C;
  geScalarMultBase(K, r);            // K = r*B
  geScalarMult(S, r, pub);           // S = r*P
  geScalarMultBase(M, sc_int64(m));  // IntToPoint
  extAdd(C, S, M);
  extToBytes(ct, K);
  extToBytes(ct + 32, C);
}

struct DlogTable {  // PointToInt map keyed by the marshalled point, grown lazily like discreteLog
  std::unordered_map<std::string, int64_t> map;
  extendedGE currentGreatestM;
  int64_t currentGreatestInt = 0;
  std::mutex mu;
};
static DlogTable g_table;
static const int64_t MaxHomomorphicInt = 100000;

static int64_t discreteLog(const extendedGE& P, bool checkNeg, bool* found) {
  uint8_t key[32];
  extToBytes(key, P);
  std::string ks((const char*)key, 32);
  std::lock_guard<std::mutex> lk(g_table.mu);
  auto it = g_table.map.find(ks);
  if (it != g_table.map.end()) { *found = true; return it->second; }
  if (g_table.currentGreatestInt == 0) extZero(g_table.currentGreatestM);
  bool foundPos = false, foundNeg = false;
  extendedGE guess = g_table.currentGreatestM, guessNeg;
  int64_t guessInt = g_table.currentGreatestInt, guessIntMinus = 0;
  bool start = true;
  for (int64_t i = guessInt; i < MaxHomomorphicInt && !foundPos && !foundNeg; i++) {
    if (!start) extAdd(guess, guess, ext_base);
    start = false;
    guessInt = i;
    uint8_t gk[32];
    extToBytes(gk, guess);
    g_table.map[std::string((const char*)gk, 32)] = guessInt;
    if (checkNeg) {
      guessIntMinus = -guessInt;
      extNeg(guessNeg, guess);
      uint8_t nk[32];
      extToBytes(nk, guessNeg);
      g_table.map[std::string((const char*)nk, 32)] = guessIntMinus;
      if (!memcmp(nk, key, 32)) foundNeg = true;
    }
    if (!foundPos && !memcmp(gk, key, 32)) foundPos = true;
  }
  g_table.currentGreatestM = guess;
  g_table.currentGreatestInt = guessInt;
  if (!foundPos && !foundNeg) { *found = false; return 0; }
  *found = true;
  return foundNeg ? guessIntMinus : guessInt;
}

}  // namespace

extern "C" {

// k * P (k: 32 bytes little-endian, reduced mod l); pt == NULL means the base point
int orc_ed_mul(const uint8_t* k32, const uint8_t* pt, uint8_t* out) {
  ensure_init();
  Sc k;
  sc_from_bytes(k, k32);
  extendedGE r;
  if (!pt) {
    geScalarMultBase(r, k);
  } else {
    extendedGE p;
    if (!extFromBytes(p, pt)) return -2;
    geScalarMult(r, k, p);
  }
  extToBytes(out, r);
  return 0;
}
int orc_ed_add(const uint8_t* a, const uint8_t* b, uint8_t* out) {
  ensure_init();
  extendedGE p, q, r;
  if (!extFromBytes(p, a) || !extFromBytes(q, b)) return -2;
  extAdd(r, p, q);
  extToBytes(out, r);
  return 0;
}
int orc_ed_encrypt(size_t n, const int64_t* m, const uint8_t* r, const uint8_t* P, uint8_t* ct, int threads) {
  ensure_init();
  extendedGE pub;
  if (!extFromBytes(pub, P)) return -2;
  parallel_for(n, threads, [&](size_t i) { encrypt_one(ct + 64 * i, pub, m[i], r + 32 * i); });
  return 0;
}
// FromBytes + CipherVector.Add folded over ndp providers; cts is [ndp][v][64] contiguous
int orc_ed_aggregate(size_t ndp, size_t v, const uint8_t* cts, uint8_t* out, int threads) {
  ensure_init();
  std::atomic<int> bad(0);
  parallel_for(v, threads, [&](size_t j) {
    extendedGE K, C, k, c;
    extZero(K);
    extZero(C);
    for (size_t d = 0; d < ndp; d++) {
      const uint8_t* w = cts + (d * v + j) * 64;
      if (!extFromBytes(k, w) || !extFromBytes(c, w + 32)) { bad++; continue; }
      extAdd(K, K, k);
      extAdd(C, C, c);
    }
    extToBytes(out + 64 * j, K);
    extToBytes(out + 64 * j + 32, C);
  });
  return bad ? -2 : 0;
}
void orc_ed_dlog_reset() {
  std::lock_guard<std::mutex> lk(g_table.mu);
  g_table.map.clear();
  g_table.currentGreatestInt = 0;
}
int orc_ed_decrypt(size_t n, const uint8_t* sk32, const uint8_t* ct, int with_neg, int64_t* out, uint8_t* found,
                   int threads) {
  ensure_init();
  Sc sk;
  sc_from_bytes(sk, sk32);
  std::atomic<int> bad(0);
  parallel_for(n, threads, [&](size_t i) {
    extendedGE K, C, S, M;
    if (!extFromBytes(K, ct + 64 * i) || !extFromBytes(C, ct + 64 * i + 32)) { bad++; out[i] = 0; return; }
    geScalarMult(S, sk, K);
    extNeg(S, S);
    extAdd(M, C, S);
    bool f;
    out[i] = discreteLog(M, with_neg != 0, &f);
    if (found) found[i] = f;
  });
  return bad ? -2 : 0;
}

}  // extern "C"
