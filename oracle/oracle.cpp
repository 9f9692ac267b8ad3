// TEST INFRASTRUCTURE ONLY -- CPU restatement of the reference's algorithms for the Drynx crypto hot path.
//
// *** parity unpinned ***: the reference (Go) cannot be compiled here (no Go toolchain; the arithmetic is in
// the un-vendored modules go.dedis.ch/kyber/v3 v3.0.8 `pairing/bn256` and github.com/ldsec/unlynx v1.4.0
// `lib/crypto.go`, go.mod:14,26) and its tests hold no golden vectors. This file restates
//   * kyber pairing/bn256: gfp.go (4 x u64 Montgomery), gfp2.go, gfp6.go, gfp12.go, curve.go, twist.go,
//     optate.go (miller / finalExponentiation), point.go (Marshal/Unmarshal)            [recalled, ⟂upstream]
//   * unlynx lib/crypto.go: encryptPoint, IntToPoint, CipherVector.Add, decryptPoint, discreteLog [recalled]
//   * /root/reference/lib/range/range_proof.go: ToBase :584-595, InitRangeProofSignature :249-288,
//     CreatePredicateRangeProofForAllServ :320-407, RangeProofVerification :504-565, ToBytes :88-155 [in tree]
// with the reference's own choice of algorithms (double-and-add scalar multiplication, general Jacobian
// addition, schoolbook gfP2 products, three pairings per (server, digit) on verification, one inversion per
// MarshalBinary), so it doubles as the timed CPU baseline. It is pinned against oracle/bn256_ref.py (the
// mathematical definition in Python integers) by tests/test_oracle.py. Only tests/, __graft_entry__.smoke()
// and bench.py's cpu_baseline / --impl reference legs may load the library built from this file.
#include <immintrin.h>

#include <algorithm>
#include <atomic>
#include <cstdint>
#include <cstring>
#include <mutex>
#include <string>
#include <thread>
#include <unordered_map>
#include <vector>

typedef unsigned __int128 u128;
typedef uint64_t u64;

// ------------------------------------------------------------------ gfP (kyber gfp.go, gfp_generic.go)
struct Mod {
  u64 m[4], np, r1[4], r2[4], r3[4];
};
static const Mod MP = {{0x185cac6c5e089667ULL, 0xee5b88d120b5b59eULL, 0xaa6fecb86184dc21ULL, 0x8fb501e34aa387f9ULL},
                       0x2387f9007f17daa9ULL,
                       {0xe7a35393a1f76999ULL, 0x11a4772edf4a4a61ULL, 0x559013479e7b23deULL, 0x704afe1cb55c7806ULL},
                       {0x9c21c3ff7e444f56ULL, 0x409ed151b2efb0c2ULL, 0x0c6dc37b80fb1651ULL, 0x7c36e0e62c2380b7ULL},
                       {0x2af2dfb9324a5bb8ULL, 0x388f899054f538a4ULL, 0xdf2ff66396b107a7ULL, 0x24ebbbb3a2529292ULL}};
static Mod MN;  // group order; filled by init_consts()

struct gfP {
  u64 v[4];
  bool operator==(const gfP& o) const { return !memcmp(v, o.v, 32); }
};

static inline bool geq(const u64 a[4], const u64 b[4]) {
  for (int i = 3; i >= 0; i--) {
    if (a[i] > b[i]) return true;
    if (a[i] < b[i]) return false;
  }
  return true;
}
static inline u64 sub4(u64 r[4], const u64 a[4], const u64 b[4]) {
  unsigned long long t0, t1, t2, t3;
  unsigned char bw = _subborrow_u64(0, a[0], b[0], &t0);
  bw = _subborrow_u64(bw, a[1], b[1], &t1);
  bw = _subborrow_u64(bw, a[2], b[2], &t2);
  bw = _subborrow_u64(bw, a[3], b[3], &t3);
  r[0] = t0; r[1] = t1; r[2] = t2; r[3] = t3;
  return bw;
}
static inline u64 add4(u64 r[4], const u64 a[4], const u64 b[4]) {
  unsigned long long t0, t1, t2, t3;
  unsigned char c = _addcarry_u64(0, a[0], b[0], &t0);
  c = _addcarry_u64(c, a[1], b[1], &t1);
  c = _addcarry_u64(c, a[2], b[2], &t2);
  c = _addcarry_u64(c, a[3], b[3], &t3);
  r[0] = t0; r[1] = t1; r[2] = t2; r[3] = t3;
  return c;
}
static inline void mod_add(const Mod& M, u64 r[4], const u64 a[4], const u64 b[4]) {
  u64 t[4], d[4];
  u64 c = add4(t, a, b);
  u64 bw = sub4(d, t, M.m);
  const u64 use_d = (u64)0 - (u64)(c | !bw);
  for (int i = 0; i < 4; i++) r[i] = (d[i] & use_d) | (t[i] & ~use_d);
}
static inline void mod_sub(const Mod& M, u64 r[4], const u64 a[4], const u64 b[4]) {
  u64 t[4], m[4];
  const u64 mask = (u64)0 - sub4(t, a, b);
  for (int i = 0; i < 4; i++) m[i] = M.m[i] & mask;
  add4(r, t, m);
}
static inline void mont_mul(const Mod& M, u64 r[4], const u64 a[4], const u64 b[4]) {
  u64 t[6] = {0, 0, 0, 0, 0, 0};
  for (int i = 0; i < 4; i++) {
    u64 c = 0;
    for (int j = 0; j < 4; j++) {
      u128 s = (u128)a[j] * b[i] + t[j] + c;
      t[j] = (u64)s;
      c = (u64)(s >> 64);
    }
    u128 s = (u128)t[4] + c;
    t[4] = (u64)s;
    t[5] = (u64)(s >> 64);
    u64 m = t[0] * M.np;
    s = (u128)m * M.m[0] + t[0];
    c = (u64)(s >> 64);
    for (int j = 1; j < 4; j++) {
      s = (u128)m * M.m[j] + t[j] + c;
      t[j - 1] = (u64)s;
      c = (u64)(s >> 64);
    }
    s = (u128)t[4] + c;
    t[3] = (u64)s;
    t[4] = t[5] + (u64)(s >> 64);
  }
  u64 d[4];
  u64 bw = sub4(d, t, M.m);
  const u64 use_d = (u64)0 - (u64)((t[4] != 0) | !bw);
  for (int i = 0; i < 4; i++) r[i] = (d[i] & use_d) | (t[i] & ~use_d);
}

static inline void gfpAdd(gfP& c, const gfP& a, const gfP& b) { mod_add(MP, c.v, a.v, b.v); }
static inline void gfpSub(gfP& c, const gfP& a, const gfP& b) { mod_sub(MP, c.v, a.v, b.v); }
static inline void gfpMul(gfP& c, const gfP& a, const gfP& b) { mont_mul(MP, c.v, a.v, b.v); }
static inline void gfpNeg(gfP& c, const gfP& a) {
  gfP z = {{0, 0, 0, 0}};
  gfpSub(c, z, a);
}
static inline bool gfpIsZero(const gfP& a) { return !(a.v[0] | a.v[1] | a.v[2] | a.v[3]); }
static gfP GFP_ZERO = {{0, 0, 0, 0}}, GFP_ONE;
static void montEncode(gfP& c, const gfP& a) {
  gfP r2;
  memcpy(r2.v, MP.r2, 32);
  gfpMul(c, a, r2);
}
static void montDecode(gfP& c, const gfP& a) {
  gfP one = {{1, 0, 0, 0}};
  gfpMul(c, a, one);
}
static gfP newGFp(int64_t x) {
  gfP out;
  if (x >= 0) {
    gfP t = {{(u64)x, 0, 0, 0}};
    montEncode(out, t);
  } else {
    gfP t = {{(u64)(-x), 0, 0, 0}};
    montEncode(out, t);
    gfpNeg(out, out);
  }
  return out;
}
// gfP.Invert: a^(p-2) by square-and-multiply over the bits of p-2 (kyber gfp.go Invert)
static void gfpInvert(gfP& e, const gfP& f) {
  u64 ex[4];
  memcpy(ex, MP.m, 32);
  ex[0] -= 2;
  gfP sum = GFP_ONE, power = f;
  for (int word = 0; word < 4; word++)
    for (int bit = 0; bit < 64; bit++) {
      if ((ex[word] >> bit) & 1) gfpMul(sum, sum, power);
      gfpMul(power, power, power);
    }
  e = sum;
}
static void gfpMarshal(uint8_t out[32], const gfP& a) {  // big-endian
  for (int w = 0; w < 4; w++)
    for (int b = 0; b < 8; b++) out[8 * w + b] = (uint8_t)(a.v[3 - w] >> (56 - 8 * b));
}
static void gfpUnmarshal(gfP& a, const uint8_t in[32]) {
  for (int w = 0; w < 4; w++) {
    a.v[3 - w] = 0;
    for (int b = 0; b < 8; b++) a.v[3 - w] |= (u64)in[8 * w + b] << (56 - 8 * b);
  }
}

// ------------------------------------------------------------------ gfP2 (kyber gfp2.go): x*i + y
struct gfP2 {
  gfP x, y;
  bool operator==(const gfP2& o) const { return x == o.x && y == o.y; }
  bool IsZero() const { return gfpIsZero(x) && gfpIsZero(y); }
  void SetZero() { x = GFP_ZERO; y = GFP_ZERO; }
  void SetOne() { x = GFP_ZERO; y = GFP_ONE; }
  void Conjugate(const gfP2& a) { y = a.y; gfpNeg(x, a.x); }
  void Neg(const gfP2& a) { gfpNeg(x, a.x); gfpNeg(y, a.y); }
  void Add(const gfP2& a, const gfP2& b) { gfpAdd(x, a.x, b.x); gfpAdd(y, a.y, b.y); }
  void Sub(const gfP2& a, const gfP2& b) { gfpSub(x, a.x, b.x); gfpSub(y, a.y, b.y); }
  void Mul(const gfP2& a, const gfP2& b) {  // schoolbook, four products
    gfP tx, t, ty;
    gfpMul(tx, a.x, b.y);
    gfpMul(t, b.x, a.y);
    gfpAdd(tx, tx, t);
    gfpMul(ty, a.y, b.y);
    gfpMul(t, a.x, b.x);
    gfpSub(ty, ty, t);
    x = tx;
    y = ty;
  }
  void MulScalar(const gfP2& a, const gfP& b) { gfpMul(x, a.x, b); gfpMul(y, a.y, b); }
  void MulXi(const gfP2& a) {  // (xi + y)(i + 3) = (3x + y)i + (3y - x)
    gfP tx, ty;
    gfpAdd(tx, a.x, a.x);
    gfpAdd(tx, tx, a.x);
    gfpAdd(tx, tx, a.y);
    gfpAdd(ty, a.y, a.y);
    gfpAdd(ty, ty, a.y);
    gfpSub(ty, ty, a.x);
    x = tx;
    y = ty;
  }
  void Square(const gfP2& a) {
    gfP tx, ty;
    gfpSub(tx, a.y, a.x);
    gfpAdd(ty, a.x, a.y);
    gfpMul(ty, tx, ty);
    gfpMul(tx, a.x, a.y);
    gfpAdd(tx, tx, tx);
    x = tx;
    y = ty;
  }
  void Invert(const gfP2& a) {
    gfP t1, t2, inv;
    gfpMul(t1, a.x, a.x);
    gfpMul(t2, a.y, a.y);
    gfpAdd(t1, t1, t2);
    gfpInvert(inv, t1);
    gfpNeg(t1, a.x);
    gfpMul(x, t1, inv);
    gfpMul(y, a.y, inv);
  }
};

// ------------------------------------------------------------------ gfP6 (kyber gfp6.go): x*tau^2 + y*tau + z
static gfP2 xiToPMinus1Over6, xiToPMinus1Over3, xiToPMinus1Over2, xiTo2PMinus2Over3;
static gfP xiToPSquaredMinus1Over3, xiTo2PSquaredMinus2Over3, xiToPSquaredMinus1Over6;

struct gfP6 {
  gfP2 x, y, z;
  bool operator==(const gfP6& o) const { return x == o.x && y == o.y && z == o.z; }
  void SetZero() { x.SetZero(); y.SetZero(); z.SetZero(); }
  void SetOne() { x.SetZero(); y.SetZero(); z.SetOne(); }
  bool IsZero() const { return x.IsZero() && y.IsZero() && z.IsZero(); }
  void Neg(const gfP6& a) { x.Neg(a.x); y.Neg(a.y); z.Neg(a.z); }
  void Add(const gfP6& a, const gfP6& b) { x.Add(a.x, b.x); y.Add(a.y, b.y); z.Add(a.z, b.z); }
  void Sub(const gfP6& a, const gfP6& b) { x.Sub(a.x, b.x); y.Sub(a.y, b.y); z.Sub(a.z, b.z); }
  void Frobenius(const gfP6& a) {
    x.Conjugate(a.x); y.Conjugate(a.y); z.Conjugate(a.z);
    x.Mul(x, xiTo2PMinus2Over3);
    y.Mul(y, xiToPMinus1Over3);
  }
  void FrobeniusP2(const gfP6& a) {
    x.MulScalar(a.x, xiTo2PSquaredMinus2Over3);
    y.MulScalar(a.y, xiToPSquaredMinus1Over3);
    z = a.z;
  }
  void Mul(const gfP6& a, const gfP6& b) {  // Karatsuba, six gfP2 products
    gfP2 v0, v1, v2, t0, t1, tz, ty, tx;
    v0.Mul(a.z, b.z);
    v1.Mul(a.y, b.y);
    v2.Mul(a.x, b.x);
    t0.Add(a.x, a.y);
    t1.Add(b.x, b.y);
    tz.Mul(t0, t1);
    tz.Sub(tz, v1);
    tz.Sub(tz, v2);
    tz.MulXi(tz);
    tz.Add(tz, v0);
    t0.Add(a.y, a.z);
    t1.Add(b.y, b.z);
    ty.Mul(t0, t1);
    t0.MulXi(v2);
    ty.Sub(ty, v0);
    ty.Sub(ty, v1);
    ty.Add(ty, t0);
    t0.Add(a.x, a.z);
    t1.Add(b.x, b.z);
    tx.Mul(t0, t1);
    tx.Sub(tx, v0);
    tx.Add(tx, v1);
    tx.Sub(tx, v2);
    x = tx;
    y = ty;
    z = tz;
  }
  void MulScalar(const gfP6& a, const gfP2& b) { x.Mul(a.x, b); y.Mul(a.y, b); z.Mul(a.z, b); }
  void MulGFP(const gfP6& a, const gfP& b) { x.MulScalar(a.x, b); y.MulScalar(a.y, b); z.MulScalar(a.z, b); }
  void MulTau(const gfP6& a) {
    gfP2 tz, ty;
    tz.MulXi(a.x);
    ty = a.y;
    y = a.z;
    x = ty;
    z = tz;
  }
  void Square(const gfP6& a) {
    gfP2 v0, v1, v2, c0, c1, c2, xiV2;
    v0.Square(a.z);
    v1.Square(a.y);
    v2.Square(a.x);
    c0.Add(a.x, a.y);
    c0.Square(c0);
    c0.Sub(c0, v1);
    c0.Sub(c0, v2);
    c0.MulXi(c0);
    c0.Add(c0, v0);
    c1.Add(a.y, a.z);
    c1.Square(c1);
    c1.Sub(c1, v0);
    c1.Sub(c1, v1);
    xiV2.MulXi(v2);
    c1.Add(c1, xiV2);
    c2.Add(a.x, a.z);
    c2.Square(c2);
    c2.Sub(c2, v0);
    c2.Add(c2, v1);
    c2.Sub(c2, v2);
    x = c2;
    y = c1;
    z = c0;
  }
  void Invert(const gfP6& a) {
    gfP2 t1, A, B, C, F;
    A.Square(a.z);
    t1.Mul(a.x, a.y);
    t1.MulXi(t1);
    A.Sub(A, t1);
    B.Square(a.x);
    B.MulXi(B);
    t1.Mul(a.y, a.z);
    B.Sub(B, t1);
    C.Square(a.y);
    t1.Mul(a.x, a.z);
    C.Sub(C, t1);
    F.Mul(C, a.y);
    F.MulXi(F);
    t1.Mul(A, a.z);
    F.Add(F, t1);
    t1.Mul(B, a.x);
    t1.MulXi(t1);
    F.Add(F, t1);
    F.Invert(F);
    x.Mul(C, F);
    y.Mul(B, F);
    z.Mul(A, F);
  }
};

// ------------------------------------------------------------------ gfP12 (kyber gfp12.go): x*w + y
struct gfP12 {
  gfP6 x, y;
  bool operator==(const gfP12& o) const { return x == o.x && y == o.y; }
  void SetOne() { x.SetZero(); y.SetOne(); }
  void Conjugate(const gfP12& a) { x.Neg(a.x); y = a.y; }
  void Frobenius(const gfP12& a) {
    x.Frobenius(a.x);
    y.Frobenius(a.y);
    x.MulScalar(x, xiToPMinus1Over6);
  }
  void FrobeniusP2(const gfP12& a) {
    x.FrobeniusP2(a.x);
    x.MulGFP(x, xiToPSquaredMinus1Over6);
    y.FrobeniusP2(a.y);
  }
  void Mul(const gfP12& a, const gfP12& b) {  // four gfP6 products
    gfP6 tx, t, ty;
    tx.Mul(a.x, b.y);
    t.Mul(b.x, a.y);
    tx.Add(tx, t);
    ty.Mul(a.y, b.y);
    t.Mul(a.x, b.x);
    t.MulTau(t);
    x = tx;
    y.Add(ty, t);
  }
  void Square(const gfP12& a) {  // complex squaring
    gfP6 v0, t, ty;
    v0.Mul(a.x, a.y);
    t.MulTau(a.x);
    t.Add(a.y, t);
    ty.Add(a.x, a.y);
    ty.Mul(ty, t);
    ty.Sub(ty, v0);
    t.MulTau(v0);
    ty.Sub(ty, t);
    x.Add(v0, v0);
    y = ty;
  }
  void Invert(const gfP12& a) {
    gfP6 t1, t2;
    t1.Square(a.x);
    t2.Square(a.y);
    t1.MulTau(t1);
    t1.Sub(t2, t1);
    t2.Invert(t1);
    gfP12 e;
    e.x.Neg(a.x);
    e.y = a.y;
    x.Mul(e.x, t2);
    y.Mul(e.y, t2);
  }
  // Exp by a little-endian multi-word exponent, square-and-multiply from the top bit (kyber gfP12.Exp)
  void Exp(const gfP12& a, const u64* e, int words) {
    gfP12 sum, t;
    sum.SetOne();
    int top = words * 64 - 1;
    while (top >= 0 && !((e[top >> 6] >> (top & 63)) & 1)) top--;
    for (int i = top; i >= 0; i--) {
      t.Square(sum);
      if ((e[i >> 6] >> (i & 63)) & 1) sum.Mul(t, a);
      else sum = t;
    }
    *this = sum;
  }
};

// ------------------------------------------------------------------ curvePoint (kyber curve.go)
static gfP curveB;
struct curvePoint {
  gfP x, y, z, t;
  bool IsInfinity() const { return gfpIsZero(z); }
  void SetInfinity() { x = GFP_ZERO; y = GFP_ONE; z = GFP_ZERO; t = GFP_ZERO; }
  bool IsOnCurve() {
    MakeAffine();
    if (IsInfinity()) return true;
    gfP y2, x3;
    gfpMul(y2, y, y);
    gfpMul(x3, x, x);
    gfpMul(x3, x3, x);
    gfpAdd(x3, x3, curveB);
    return y2 == x3;
  }
  void Double(const curvePoint& a) {  // dbl-2009-l
    gfP A, B, C, t, t2, d, e, f;
    gfpMul(A, a.x, a.x);
    gfpMul(B, a.y, a.y);
    gfpMul(C, B, B);
    gfpAdd(t, a.x, B);
    gfpMul(t2, t, t);
    gfpSub(t, t2, A);
    gfpSub(t2, t, C);
    gfpAdd(d, t2, t2);
    gfpAdd(t, A, A);
    gfpAdd(e, t, A);
    gfpMul(f, e, e);
    gfpAdd(t, d, d);
    gfP nx, ny, nz;
    gfpSub(nx, f, t);
    gfpMul(nz, a.y, a.z);
    gfpAdd(nz, nz, nz);
    gfpAdd(t, C, C);
    gfpAdd(t2, t, t);
    gfpAdd(t, t2, t2);
    gfpSub(ny, d, nx);
    gfpMul(t2, e, ny);
    gfpSub(ny, t2, t);
    x = nx;
    y = ny;
    z = nz;
  }
  void Add(const curvePoint& a, const curvePoint& b) {  // add-2007-bl
    if (a.IsInfinity()) { *this = b; return; }
    if (b.IsInfinity()) { *this = a; return; }
    gfP z12, z22, u1, u2, t, s1, s2, h, i, j, r, v, t4, t6;
    gfpMul(z12, a.z, a.z);
    gfpMul(z22, b.z, b.z);
    gfpMul(u1, a.x, z22);
    gfpMul(u2, b.x, z12);
    gfpMul(t, b.z, z22);
    gfpMul(s1, a.y, t);
    gfpMul(t, a.z, z12);
    gfpMul(s2, b.y, t);
    gfpSub(h, u2, u1);
    bool xEqual = gfpIsZero(h);
    gfpAdd(t, h, h);
    gfpMul(i, t, t);
    gfpMul(j, h, i);
    gfpSub(t, s2, s1);
    bool yEqual = gfpIsZero(t);
    if (xEqual && yEqual) { Double(a); return; }
    gfpAdd(r, t, t);
    gfpMul(v, u1, i);
    gfpMul(t4, r, r);
    gfpAdd(t, v, v);
    gfpSub(t6, t4, j);
    gfP nx, ny, nz;
    gfpSub(nx, t6, t);
    gfpSub(t, v, nx);
    gfpMul(t4, s1, j);
    gfpAdd(t6, t4, t4);
    gfpMul(t4, r, t);
    gfpSub(ny, t4, t6);
    gfpAdd(t, a.z, b.z);
    gfpMul(t4, t, t);
    gfpSub(t, t4, z12);
    gfpSub(t4, t, z22);
    gfpMul(nz, t4, h);
    x = nx;
    y = ny;
    z = nz;
  }
  // Mul: double-and-add over the bits of the (reduced) scalar, most significant first
  void Mul(const curvePoint& a, const u64 k[4]) {
    curvePoint sum, t;
    sum.SetInfinity();
    int top = 255;
    while (top >= 0 && !((k[top >> 6] >> (top & 63)) & 1)) top--;
    for (int i = top; i >= 0; i--) {
      t.Double(sum);
      if ((k[i >> 6] >> (i & 63)) & 1) sum.Add(t, a);
      else sum = t;
    }
    *this = sum;
  }
  void MakeAffine() {
    if (z == GFP_ONE) return;
    if (gfpIsZero(z)) { x = GFP_ZERO; y = GFP_ONE; t = GFP_ZERO; return; }
    gfP zInv, t1, zInv2;
    gfpInvert(zInv, z);
    gfpMul(t1, y, zInv);
    gfpMul(zInv2, zInv, zInv);
    gfpMul(x, x, zInv2);
    gfpMul(y, t1, zInv2);
    z = GFP_ONE;
    t = GFP_ONE;
  }
  void Neg(const curvePoint& a) { x = a.x; gfpNeg(y, a.y); z = a.z; t = GFP_ZERO; }
};

// ------------------------------------------------------------------ twistPoint (kyber twist.go)
static gfP2 twistB;
struct twistPoint {
  gfP2 x, y, z, t;
  bool IsInfinity() const { return z.IsZero(); }
  void SetInfinity() { x.SetZero(); y.SetOne(); z.SetZero(); t.SetZero(); }
  // kyber pairing/bn256 twist.go IsOnCurve (recalled; the module is not in /root/reference): the curve equation, then
  // `cneg.Mul(c, Order); return cneg.z.IsZero()` -- a twist point outside the order-n subgroup is NOT "on the curve"
  // for kyber, so UnmarshalBinary fails on it and lib/range/range_proof.go:225-236 aborts. Defined after Mul below.
  bool IsOnCurve();
  void Double(const twistPoint& a) {
    gfP2 A, B, C, t, t2, d, e, f, nx, ny, nz;
    A.Square(a.x);
    B.Square(a.y);
    C.Square(B);
    t.Add(a.x, B);
    t2.Square(t);
    t.Sub(t2, A);
    t2.Sub(t, C);
    d.Add(t2, t2);
    t.Add(A, A);
    e.Add(t, A);
    f.Square(e);
    t.Add(d, d);
    nx.Sub(f, t);
    nz.Mul(a.y, a.z);
    nz.Add(nz, nz);
    t.Add(C, C);
    t2.Add(t, t);
    t.Add(t2, t2);
    ny.Sub(d, nx);
    t2.Mul(e, ny);
    ny.Sub(t2, t);
    x = nx;
    y = ny;
    z = nz;
  }
  void Add(const twistPoint& a, const twistPoint& b) {
    if (a.IsInfinity()) { *this = b; return; }
    if (b.IsInfinity()) { *this = a; return; }
    gfP2 z12, z22, u1, u2, t, s1, s2, h, i, j, r, v, t4, t6, nx, ny, nz;
    z12.Square(a.z);
    z22.Square(b.z);
    u1.Mul(a.x, z22);
    u2.Mul(b.x, z12);
    t.Mul(b.z, z22);
    s1.Mul(a.y, t);
    t.Mul(a.z, z12);
    s2.Mul(b.y, t);
    h.Sub(u2, u1);
    bool xEqual = h.IsZero();
    t.Add(h, h);
    i.Square(t);
    j.Mul(h, i);
    t.Sub(s2, s1);
    bool yEqual = t.IsZero();
    if (xEqual && yEqual) { Double(a); return; }
    r.Add(t, t);
    v.Mul(u1, i);
    t4.Square(r);
    t.Add(v, v);
    t6.Sub(t4, j);
    nx.Sub(t6, t);
    t.Sub(v, nx);
    t4.Mul(s1, j);
    t6.Add(t4, t4);
    t4.Mul(r, t);
    ny.Sub(t4, t6);
    t.Add(a.z, b.z);
    t4.Square(t);
    t.Sub(t4, z12);
    t4.Sub(t, z22);
    nz.Mul(t4, h);
    x = nx;
    y = ny;
    z = nz;
  }
  void Mul(const twistPoint& a, const u64 k[4]) {
    twistPoint sum, t;
    sum.SetInfinity();
    int top = 255;
    while (top >= 0 && !((k[top >> 6] >> (top & 63)) & 1)) top--;
    for (int i = top; i >= 0; i--) {
      t.Double(sum);
      if ((k[i >> 6] >> (i & 63)) & 1) sum.Add(t, a);
      else sum = t;
    }
    *this = sum;
  }
  void MakeAffine() {
    gfP2 one;
    one.SetOne();
    if (z == one) return;
    if (z.IsZero()) { x.SetZero(); y.SetOne(); t.SetZero(); return; }
    gfP2 zInv, t1, zInv2;
    zInv.Invert(z);
    t1.Mul(y, zInv);
    zInv2.Square(zInv);
    y.Mul(t1, zInv2);
    t1.Mul(x, zInv2);
    x = t1;
    z.SetOne();
    t.SetOne();
  }
  void Neg(const twistPoint& a) { x = a.x; y.Neg(a.y); z = a.z; t.SetZero(); }
};

bool twistPoint::IsOnCurve() {
  MakeAffine();
  if (IsInfinity()) return true;
  gfP2 y2, x3;
  y2.Square(y);
  x3.Square(x);
  x3.Mul(x3, x);
  x3.Add(x3, twistB);
  if (!(y2 == x3)) return false;
  static const u64 order[4] = {0x1a2ef45b57ac7261ULL, 0x2e8d8e12f82b3924ULL, 0xaa6fecb86184dc21ULL, 0x8fb501e34aa387f9ULL};
  twistPoint cneg;
  cneg.Mul(*this, order);
  return cneg.z.IsZero();
}

// ------------------------------------------------------------------ optimal ate (kyber optate.go)
static void lineFunctionAdd(const twistPoint& r, const twistPoint& p, const curvePoint& q, const gfP2& r2, gfP2& a,
                            gfP2& b, gfP2& c, twistPoint& rOut) {
  gfP2 B, D, H, I, E, J, L1, V, t, t2;
  B.Mul(p.x, r.t);
  D.Add(p.y, r.z);
  D.Square(D);
  D.Sub(D, r2);
  D.Sub(D, r.t);
  D.Mul(D, r.t);
  H.Sub(B, r.x);
  I.Square(H);
  E.Add(I, I);
  E.Add(E, E);
  J.Mul(H, E);
  L1.Sub(D, r.y);
  L1.Sub(L1, r.y);
  V.Mul(r.x, E);
  rOut.x.Square(L1);
  rOut.x.Sub(rOut.x, J);
  rOut.x.Sub(rOut.x, V);
  rOut.x.Sub(rOut.x, V);
  rOut.z.Add(r.z, H);
  rOut.z.Square(rOut.z);
  rOut.z.Sub(rOut.z, r.t);
  rOut.z.Sub(rOut.z, I);
  t.Sub(V, rOut.x);
  t.Mul(t, L1);
  t2.Mul(r.y, J);
  t2.Add(t2, t2);
  rOut.y.Sub(t, t2);
  rOut.t.Square(rOut.z);
  t.Add(p.y, rOut.z);
  t.Square(t);
  t.Sub(t, r2);
  t.Sub(t, rOut.t);
  t2.Mul(L1, p.x);
  t2.Add(t2, t2);
  a.Sub(t2, t);
  c.MulScalar(rOut.z, q.y);
  c.Add(c, c);
  b.Neg(L1);
  b.MulScalar(b, q.x);
  b.Add(b, b);
}
static void lineFunctionDouble(const twistPoint& r, const curvePoint& q, gfP2& a, gfP2& b, gfP2& c, twistPoint& rOut) {
  gfP2 A, B, C, D, E, G, t;
  A.Square(r.x);
  B.Square(r.y);
  C.Square(B);
  D.Add(r.x, B);
  D.Square(D);
  D.Sub(D, A);
  D.Sub(D, C);
  D.Add(D, D);
  E.Add(A, A);
  E.Add(E, A);
  G.Square(E);
  rOut.x.Sub(G, D);
  rOut.x.Sub(rOut.x, D);
  rOut.z.Add(r.y, r.z);
  rOut.z.Square(rOut.z);
  rOut.z.Sub(rOut.z, B);
  rOut.z.Sub(rOut.z, r.t);
  rOut.y.Sub(D, rOut.x);
  rOut.y.Mul(rOut.y, E);
  t.Add(C, C);
  t.Add(t, t);
  t.Add(t, t);
  rOut.y.Sub(rOut.y, t);
  rOut.t.Square(rOut.z);
  t.Mul(E, r.t);
  t.Add(t, t);
  b.Neg(t);
  b.MulScalar(b, q.x);
  a.Add(r.x, E);
  a.Square(a);
  a.Sub(a, A);
  a.Sub(a, G);
  t.Add(B, B);
  t.Add(t, t);
  a.Sub(a, t);
  c.Mul(rOut.z, r.t);
  c.Add(c, c);
  c.MulScalar(c, q.y);
}
static void mulLine(gfP12& ret, const gfP2& a, const gfP2& b, const gfP2& c) {
  gfP6 a2, t3, t2;
  a2.x.SetZero();
  a2.y = a;
  a2.z = b;
  a2.Mul(a2, ret.x);
  t3.MulScalar(ret.y, c);
  gfP2 t;
  t.Add(b, c);
  t2.x.SetZero();
  t2.y = a;
  t2.z = t;
  ret.x.Add(ret.x, ret.y);
  ret.y = t3;
  ret.x.Mul(ret.x, t2);
  ret.x.Sub(ret.x, a2);
  ret.x.Sub(ret.x, ret.y);
  a2.MulTau(a2);
  ret.y.Add(ret.y, a2);
}

static std::vector<int8_t> sixuPlus2NAF;
static u64 U_PARAM = 6518589491078791937ULL;

static void miller(gfP12& ret, const twistPoint& q, const curvePoint& p) {
  ret.SetOne();
  twistPoint aAffine = q;
  aAffine.MakeAffine();
  curvePoint bAffine = p;
  bAffine.MakeAffine();
  twistPoint minusA;
  minusA.Neg(aAffine);
  minusA.t = aAffine.t;
  twistPoint r = aAffine;
  gfP2 r2;
  r2.Square(aAffine.y);
  gfP2 a, b, c;
  twistPoint newR;
  const int L = (int)sixuPlus2NAF.size();
  for (int i = L - 1; i > 0; i--) {
    lineFunctionDouble(r, bAffine, a, b, c, newR);
    if (i != L - 1) ret.Square(ret);
    mulLine(ret, a, b, c);
    r = newR;
    switch (sixuPlus2NAF[i - 1]) {
      case 1: lineFunctionAdd(r, aAffine, bAffine, r2, a, b, c, newR); break;
      case -1: lineFunctionAdd(r, minusA, bAffine, r2, a, b, c, newR); break;
      default: continue;
    }
    mulLine(ret, a, b, c);
    r = newR;
  }
  twistPoint q1, minusQ2;
  q1.x.Conjugate(aAffine.x);
  q1.x.Mul(q1.x, xiToPMinus1Over3);
  q1.y.Conjugate(aAffine.y);
  q1.y.Mul(q1.y, xiToPMinus1Over2);
  q1.z.SetOne();
  q1.t.SetOne();
  minusQ2.x.MulScalar(aAffine.x, xiToPSquaredMinus1Over3);
  minusQ2.y = aAffine.y;
  minusQ2.z.SetOne();
  minusQ2.t.SetOne();
  r2.Square(q1.y);
  lineFunctionAdd(r, q1, bAffine, r2, a, b, c, newR);
  mulLine(ret, a, b, c);
  r = newR;
  r2.Square(minusQ2.y);
  lineFunctionAdd(r, minusQ2, bAffine, r2, a, b, c, newR);
  mulLine(ret, a, b, c);
}

static void finalExponentiation(gfP12& out, const gfP12& in) {
  gfP12 t1, inv, t2, fp, fp2, fp3, fu, fu2, fu3, y0, y1, y2, y3, y4, y5, y6, fu2p, fu3p, t0;
  t1.x.Neg(in.x);
  t1.y = in.y;
  inv.Invert(in);
  t1.Mul(t1, inv);
  t2.FrobeniusP2(t1);
  t1.Mul(t1, t2);
  fp.Frobenius(t1);
  fp2.FrobeniusP2(t1);
  fp3.Frobenius(fp2);
  fu.Exp(t1, &U_PARAM, 1);
  fu2.Exp(fu, &U_PARAM, 1);
  fu3.Exp(fu2, &U_PARAM, 1);
  y3.Frobenius(fu);
  fu2p.Frobenius(fu2);
  fu3p.Frobenius(fu3);
  y2.FrobeniusP2(fu2);
  y0.Mul(fp, fp2);
  y0.Mul(y0, fp3);
  y1.Conjugate(t1);
  y5.Conjugate(fu2);
  y3.Conjugate(y3);
  y4.Mul(fu, fu2p);
  y4.Conjugate(y4);
  y6.Mul(fu3, fu3p);
  y6.Conjugate(y6);
  t0.Square(y6);
  t0.Mul(t0, y4);
  t0.Mul(t0, y5);
  t1.Mul(y3, y5);
  t1.Mul(t1, t0);
  t0.Mul(t0, y2);
  t1.Square(t1);
  t1.Mul(t1, t0);
  t1.Square(t1);
  t0.Mul(t1, y1);
  t1.Mul(t1, y0);
  t0.Square(t0);
  out.Mul(t0, t1);
}
static void optimalAte(gfP12& out, const twistPoint& a, const curvePoint& b) {
  if (a.IsInfinity() || b.IsInfinity()) { out.SetOne(); return; }
  gfP12 e;
  miller(e, a, b);
  finalExponentiation(out, e);
}

// ------------------------------------------------------------------ scalars (kyber group/mod Int, big-endian)
struct Sc {
  u64 v[4];
};
static void sc_from_bytes(Sc& s, const uint8_t in[32]) {  // SetBytes / UnmarshalBinary reduce mod n
  gfP t;
  gfpUnmarshal(t, in);
  memcpy(s.v, t.v, 32);
  if (geq(s.v, MN.m)) sub4(s.v, s.v, MN.m);
}
static void sc_to_bytes(uint8_t out[32], const Sc& s) {
  gfP t;
  memcpy(t.v, s.v, 32);
  gfpMarshal(out, t);
}
static Sc sc_add(const Sc& a, const Sc& b) { Sc r; mod_add(MN, r.v, a.v, b.v); return r; }
static Sc sc_sub(const Sc& a, const Sc& b) { Sc r; mod_sub(MN, r.v, a.v, b.v); return r; }
static Sc sc_mul(const Sc& a, const Sc& b) {
  Sc t, r;
  mont_mul(MN, t.v, a.v, b.v);
  mont_mul(MN, r.v, t.v, MN.r2);
  return r;
}
static Sc sc_neg(const Sc& a) { Sc z = {{0, 0, 0, 0}}; return sc_sub(z, a); }
static Sc sc_int64(int64_t x) {  // SetInt64
  Sc s = {{0, 0, 0, 0}};
  if (x >= 0) { s.v[0] = (u64)x; return s; }
  s.v[0] = (u64)0 - (u64)x;
  return sc_neg(s);
}
static Sc sc_inv(const Sc& a) {  // a^(n-2)
  u64 ex[4];
  memcpy(ex, MN.m, 32);
  ex[0] -= 2;
  Sc sum = {{1, 0, 0, 0}}, power = a;
  for (int i = 0; i < 256; i++) {
    if ((ex[i >> 6] >> (i & 63)) & 1) sum = sc_mul(sum, power);
    power = sc_mul(power, power);
  }
  return sum;
}
static Sc sc_from_wide(const uint8_t in[64]) {  // 64 big-endian bytes mod n
  Sc hi, lo, t;
  sc_from_bytes(hi, in);
  sc_from_bytes(lo, in + 32);
  mont_mul(MN, t.v, hi.v, MN.r2);  // hi * 2^256
  return sc_add(t, lo);
}

// ------------------------------------------------------------------ kyber point (de)serialisation (point.go)
static curvePoint curveGen;
static twistPoint twistGen;
static gfP12 gtBase;  // e(curveGen, twistGen)

static void g1_marshal(uint8_t out[64], const curvePoint& p) {
  curvePoint t = p;
  t.MakeAffine();
  memset(out, 0, 64);
  if (t.IsInfinity()) return;
  gfP tmp;
  montDecode(tmp, t.x);
  gfpMarshal(out, tmp);
  montDecode(tmp, t.y);
  gfpMarshal(out + 32, tmp);
}
static bool g1_unmarshal(curvePoint& p, const uint8_t in[64]) {
  gfP x, y;
  gfpUnmarshal(x, in);
  gfpUnmarshal(y, in + 32);
  if (gfpIsZero(x) && gfpIsZero(y)) { p.SetInfinity(); return true; }
  if (geq(x.v, MP.m) || geq(y.v, MP.m)) return false;
  montEncode(p.x, x);
  montEncode(p.y, y);
  p.z = GFP_ONE;
  p.t = GFP_ONE;
  return p.IsOnCurve();
}
static void fp2_marshal(uint8_t out[64], const gfP2& a) {
  gfP tmp;
  montDecode(tmp, a.x);
  gfpMarshal(out, tmp);
  montDecode(tmp, a.y);
  gfpMarshal(out + 32, tmp);
}
static bool fp2_unmarshal(gfP2& a, const uint8_t in[64]) {
  gfP x, y;
  gfpUnmarshal(x, in);
  gfpUnmarshal(y, in + 32);
  bool ok = !geq(x.v, MP.m) && !geq(y.v, MP.m);
  montEncode(a.x, x);
  montEncode(a.y, y);
  return ok;
}
static void g2_marshal(uint8_t out[128], const twistPoint& p) {
  twistPoint t = p;
  t.MakeAffine();
  memset(out, 0, 128);
  if (t.IsInfinity()) return;
  fp2_marshal(out, t.x);
  fp2_marshal(out + 64, t.y);
}
static bool g2_unmarshal(twistPoint& p, const uint8_t in[128]) {
  bool any = false;
  for (int i = 0; i < 128; i++) any |= in[i] != 0;
  if (!any) { p.SetInfinity(); return true; }
  if (!fp2_unmarshal(p.x, in) || !fp2_unmarshal(p.y, in + 64)) return false;
  p.z.SetOne();
  p.t.SetOne();
  return p.IsOnCurve();
}
static void gt_marshal(uint8_t out[384], const gfP12& a) {
  fp2_marshal(out, a.x.x);
  fp2_marshal(out + 64, a.x.y);
  fp2_marshal(out + 128, a.x.z);
  fp2_marshal(out + 192, a.y.x);
  fp2_marshal(out + 256, a.y.y);
  fp2_marshal(out + 320, a.y.z);
}
static bool gt_unmarshal(gfP12& a, const uint8_t in[384]) {
  bool ok = fp2_unmarshal(a.x.x, in);
  ok &= fp2_unmarshal(a.x.y, in + 64);
  ok &= fp2_unmarshal(a.x.z, in + 128);
  ok &= fp2_unmarshal(a.y.x, in + 192);
  ok &= fp2_unmarshal(a.y.y, in + 256);
  ok &= fp2_unmarshal(a.y.z, in + 320);
  return ok;
}

// ------------------------------------------------------------------ SHA3-512 (golang.org/x/crypto/sha3 New512)
static void keccakf(u64 st[25]) {
  static const u64 RC[24] = {0x0000000000000001ULL, 0x0000000000008082ULL, 0x800000000000808aULL, 0x8000000080008000ULL,
                             0x000000000000808bULL, 0x0000000080000001ULL, 0x8000000080008081ULL, 0x8000000000008009ULL,
                             0x000000000000008aULL, 0x0000000000000088ULL, 0x0000000080008009ULL, 0x000000008000000aULL,
                             0x000000008000808bULL, 0x800000000000008bULL, 0x8000000000008089ULL, 0x8000000000008003ULL,
                             0x8000000000008002ULL, 0x8000000000000080ULL, 0x000000000000800aULL, 0x800000008000000aULL,
                             0x8000000080008081ULL, 0x8000000000008080ULL, 0x0000000080000001ULL, 0x8000000080008008ULL};
  static const int rotc[24] = {1, 3, 6, 10, 15, 21, 28, 36, 45, 55, 2, 14, 27, 41, 56, 8, 25, 43, 62, 18, 39, 61, 20, 44};
  static const int piln[24] = {10, 7, 11, 17, 18, 3, 5, 16, 8, 21, 24, 4, 15, 23, 19, 13, 12, 2, 20, 14, 22, 9, 6, 1};
  for (int round = 0; round < 24; round++) {
    u64 bc[5], t;
    for (int i = 0; i < 5; i++) bc[i] = st[i] ^ st[i + 5] ^ st[i + 10] ^ st[i + 15] ^ st[i + 20];
    for (int i = 0; i < 5; i++) {
      t = bc[(i + 4) % 5] ^ ((bc[(i + 1) % 5] << 1) | (bc[(i + 1) % 5] >> 63));
      for (int j = 0; j < 25; j += 5) st[j + i] ^= t;
    }
    t = st[1];
    for (int i = 0; i < 24; i++) {
      int j = piln[i];
      u64 b = st[j];
      st[j] = (t << rotc[i]) | (t >> (64 - rotc[i]));
      t = b;
    }
    for (int j = 0; j < 25; j += 5) {
      for (int i = 0; i < 5; i++) bc[i] = st[j + i];
      for (int i = 0; i < 5; i++) st[j + i] ^= (~bc[(i + 1) % 5]) & bc[(i + 2) % 5];
    }
    st[0] ^= RC[round];
  }
}
static void sha3_512(uint8_t out[64], const uint8_t* in, size_t len) {
  u64 st[25];
  memset(st, 0, sizeof st);
  const size_t rate = 72;
  uint8_t block[72];
  while (len >= rate) {
    for (size_t i = 0; i < rate / 8; i++) {
      u64 w;
      memcpy(&w, in + 8 * i, 8);
      st[i] ^= w;
    }
    keccakf(st);
    in += rate;
    len -= rate;
  }
  memset(block, 0, rate);
  memcpy(block, in, len);
  block[len] ^= 0x06;
  block[rate - 1] ^= 0x80;
  for (size_t i = 0; i < rate / 8; i++) {
    u64 w;
    memcpy(&w, block + 8 * i, 8);
    st[i] ^= w;
  }
  keccakf(st);
  memcpy(out, st, 64);
}

// ------------------------------------------------------------------ constants
static gfP2 fp2_pow(const gfP2& a, const std::vector<u64>& e) {
  gfP2 r, b = a;
  r.SetOne();
  for (size_t i = 0; i < e.size() * 64; i++) {
    if ((e[i >> 6] >> (i & 63)) & 1) r.Mul(r, b);
    b.Square(b);
  }
  return r;
}
// little-endian multiword helpers for the exponents (p-1)/k, (p^2-1)/k
static std::vector<u64> mw_mul(const std::vector<u64>& a, const std::vector<u64>& b) {
  std::vector<u64> r(a.size() + b.size(), 0);
  for (size_t i = 0; i < a.size(); i++) {
    u64 c = 0;
    for (size_t j = 0; j < b.size(); j++) {
      u128 t = (u128)a[i] * b[j] + r[i + j] + c;
      r[i + j] = (u64)t;
      c = (u64)(t >> 64);
    }
    r[i + b.size()] += c;
  }
  return r;
}
static std::vector<u64> mw_sub1(std::vector<u64> a) {
  for (size_t i = 0; i < a.size(); i++)
    if (a[i]--) break;
  return a;
}
static std::vector<u64> mw_div_small(const std::vector<u64>& a, u64 d) {
  std::vector<u64> q(a.size());
  u128 rem = 0;
  for (size_t i = a.size(); i-- > 0;) {
    u128 cur = (rem << 64) | a[i];
    q[i] = (u64)(cur / d);
    rem = cur % d;
  }
  return q;
}
static std::vector<u64> mw_mul_small(std::vector<u64> a, u64 k) {
  u64 c = 0;
  for (auto& w : a) {
    u128 t = (u128)w * k + c;
    w = (u64)t;
    c = (u64)(t >> 64);
  }
  a.push_back(c);
  return a;
}

static std::once_flag g_init;
static void init_consts() {
  // n = 36u^4 + 36u^3 + 18u^2 + 6u + 1
  u64 nn[4] = {0x1a2ef45b57ac7261ULL, 0x2e8d8e12f82b3924ULL, 0xaa6fecb86184dc21ULL, 0x8fb501e34aa387f9ULL};
  memcpy(MN.m, nn, 32);
  u64 inv = 1;
  for (int i = 0; i < 6; i++) inv *= 2 - MN.m[0] * inv;
  MN.np = (u64)0 - inv;
  // 2^256 mod n, then R^2, R^3 by doubling / multiplication
  u64 r[4] = {0, 0, 0, 0};
  sub4(r, r, MN.m);  // 2^256 - n  (n > 2^255)
  memcpy(MN.r1, r, 32);
  u64 acc[4];
  memcpy(acc, r, 32);
  for (int i = 0; i < 256; i++) mod_add(MN, acc, acc, acc);
  memcpy(MN.r2, acc, 32);
  mont_mul(MN, MN.r3, MN.r2, MN.r2);
  memcpy(GFP_ONE.v, MP.r1, 32);
  curveB = newGFp(3);
  gfP2 xi;
  xi.x = newGFp(1);
  xi.y = newGFp(3);
  std::vector<u64> p(MP.m, MP.m + 4);
  std::vector<u64> pm1 = mw_sub1(p), p2m1 = mw_sub1(mw_mul(p, p));
  xiToPMinus1Over6 = fp2_pow(xi, mw_div_small(pm1, 6));
  xiToPMinus1Over3 = fp2_pow(xi, mw_div_small(pm1, 3));
  xiToPMinus1Over2 = fp2_pow(xi, mw_div_small(pm1, 2));
  xiTo2PMinus2Over3 = fp2_pow(xi, mw_div_small(mw_mul_small(pm1, 2), 3));
  xiToPSquaredMinus1Over3 = fp2_pow(xi, mw_div_small(p2m1, 3)).y;
  xiTo2PSquaredMinus2Over3 = fp2_pow(xi, mw_div_small(mw_mul_small(p2m1, 2), 3)).y;
  xiToPSquaredMinus1Over6 = fp2_pow(xi, mw_div_small(p2m1, 6)).y;
  gfP2 three;
  three.x = GFP_ZERO;
  three.y = newGFp(3);
  gfP2 xinv;
  xinv.Invert(xi);
  twistB.Mul(three, xinv);
  // NAF of 6u+2
  u128 s = (u128)6 * U_PARAM + 2;
  while (s) {
    int8_t d = 0;
    if (s & 1) {
      d = (int8_t)(2 - (int)(s & 3));
      if (d == 1) s -= 1; else s += 1;
    }
    sixuPlus2NAF.push_back(d);
    s >>= 1;
  }
  // generators: curveGen = (1, -2); twistGen = the x/crypto bn256 twist generator
  curveGen.x = newGFp(1);
  curveGen.y = newGFp(-2);
  curveGen.z = GFP_ONE;
  curveGen.t = GFP_ONE;
  static const uint8_t g2b[128] = {
#include "g2_gen_bytes.inc"
  };
  g2_unmarshal(twistGen, g2b);
  optimalAte(gtBase, twistGen, curveGen);
}
static inline void ensure_init() { std::call_once(g_init, init_consts); }

// ------------------------------------------------------------------ thread fan-out (libunlynx StartParallelize)
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
static void encrypt_one(uint8_t ct[128], const curvePoint& pub, int64_t m, const uint8_t r32[32]) {
  Sc r, ms = sc_int64(m);
  sc_from_bytes(r, r32);
  curvePoint K, S, M, C;
  K.Mul(curveGen, r.v);   // K = r*B
  S.Mul(pub, r.v);        // S = r*P
  M.Mul(curveGen, ms.v);  // IntToPoint
  C.Add(S, M);
  g1_marshal(ct, K);
  g1_marshal(ct + 64, C);
}

// libunlynx.EncryptScalar: IntToPoint replaced by scalar * B (lib/encoding/OR_AND.go:52,105)
static void encrypt_scalar_one(uint8_t ct[128], const curvePoint& pub, const uint8_t m32[32], const uint8_t r32[32]) {
  Sc r, ms;
  sc_from_bytes(r, r32);
  sc_from_bytes(ms, m32);
  curvePoint K, S, M, C;
  K.Mul(curveGen, r.v);
  S.Mul(pub, r.v);
  M.Mul(curveGen, ms.v);
  C.Add(S, M);
  g1_marshal(ct, K);
  g1_marshal(ct + 64, C);
}

// libunlynxkeyswitch.KeySwitchSequence (github.com/ldsec/unlynx v1.4.0 lib/key_switch/key_switch.go, not under
// /root/reference; call sites data/data.go:80 and, through protocolsunlynx.KeySwitchingProtocol,
// services/service.go:565-616), restated as recalled, one element: vi <- random (injected here);
// rbNeg = -rB; rbkNeg = secretKey * rbNeg; viB = vi * B; viNewK = vi * targetPubKey; ks2 = rbkNeg + viNewK;
// the contribution is the ciphertext (K, C) = (viB, ks2); rbNeg is returned for the proof.
static bool keyswitch_one(uint8_t cv[128], uint8_t* rbneg64, const curvePoint& target, const Sc& x, const uint8_t rb64[64],
                          const uint8_t v32[32]) {
  curvePoint rB;
  if (!g1_unmarshal(rB, rb64)) return false;
  Sc vi;
  sc_from_bytes(vi, v32);
  curvePoint rbNeg, rbkNeg, viB, viNewK, ks2;
  rbNeg.Neg(rB);
  rbkNeg.Mul(rbNeg, x.v);
  viB.Mul(curveGen, vi.v);
  viNewK.Mul(target, vi.v);
  ks2.Add(rbkNeg, viNewK);
  g1_marshal(cv, viB);
  g1_marshal(cv + 64, ks2);
  if (rbneg64) g1_marshal(rbneg64, rbNeg);
  return true;
}

// CipherText.MulCipherTextbyScalar (unlynx lib/crypto.go, recalled: K = s*K, C = s*C; in-tree call sites
// protocols/obfuscation_protocol.go:240-243, data/data.go:52-56)
static bool ct_scalar_mul_one(uint8_t out[128], const uint8_t ct[128], const uint8_t s32[32]) {
  curvePoint K, C, sK, sC;
  if (!g1_unmarshal(K, ct) || !g1_unmarshal(C, ct + 64)) return false;
  Sc s;
  sc_from_bytes(s, s32);
  sK.Mul(K, s.v);
  sC.Mul(C, s.v);
  g1_marshal(out, sK);
  g1_marshal(out + 64, sC);
  return true;
}

// kyber v3.0.8 sign/schnorr (not under /root/reference; called by VerifyProofSignature,
// lib/proof/structs_proofs.go:498-505, and by the New*ProofRequest constructors :117 etc.), restated as recalled:
//   Sign(s, x, msg):   k <- random; R = k*B; h = SHA-512(R || X || msg) as a scalar (SetBytes: big-endian mod n);
//                      s = k + x*h; signature = R || s
//   Verify(g, X, msg, sig): len(sig) == 64 + 32; R, s unmarshal (a point off the curve or s >= n is an error);
//                      h as above; accept iff s*B == R + h*X
// The digest is an input here: the caller hashes (tests use hashlib, the host mirror its own SHA-512).
static bool schnorr_verify_one(const uint8_t pub[64], const uint8_t sig[96], const uint8_t digest[64]) {
  curvePoint X, R, S, Ah, RAs;
  if (!g1_unmarshal(X, pub) || !g1_unmarshal(R, sig)) return false;
  gfP sraw;
  gfpUnmarshal(sraw, sig + 64);
  if (geq(sraw.v, MN.m)) return false;  // mod.Int.UnmarshalBinary: value out of range
  Sc sc;
  memcpy(sc.v, sraw.v, 32);
  const Sc h = sc_from_wide(digest);
  S.Mul(curveGen, sc.v);
  Ah.Mul(X, h.v);
  RAs.Add(R, Ah);
  uint8_t a[64], b[64];
  g1_marshal(a, S);
  g1_marshal(b, RAs);
  return memcmp(a, b, 64) == 0;
}
// kyber v3.0.8 proof/rep.go (dependency, recalled): a proof.Rep("P", "x1", "B1", ..., "xT", "BT") predicate.
//   commit:  V = v_1*B_1 + ... + v_T*B_T, v_j = the blinding scalar of secret x_j (shared by every Rep that names x_j)
//   respond: r_j = v_j - c*x_j ("ri.Mul(c, x); ri.Sub(v, ri)")
//   verify:  V' = c*P + r_1*B_1 + ... + r_T*B_T; V'.Equal(V)
// The challenge c is the caller's (kyber draws it from the hashProver's XOF after every V of the proof was Put). Used by
// lib/obfuscation/obfuscation_proof.go:37-91 and by unlynx's key-switch proof (lib/proof/structs_proofs.go:469-482).
static bool rep_commit_one(uint8_t out[64], int nt, const uint8_t* sc, const uint8_t* pts) {
  curvePoint V, t;
  V.SetInfinity();
  for (int j = 0; j < nt; j++) {
    curvePoint B;
    if (!g1_unmarshal(B, pts + 64 * j)) return false;
    Sc v;
    sc_from_bytes(v, sc + 32 * j);
    t.Mul(B, v.v);
    V.Add(V, t);
  }
  g1_marshal(out, V);
  return true;
}
static bool rep_verify_one(int nt, const uint8_t* resp, const uint8_t chal[32], const uint8_t* bases, const uint8_t P[64],
                           const uint8_t V[64]) {
  curvePoint Pp, Vv, acc, t;
  if (!g1_unmarshal(Pp, P) || !g1_unmarshal(Vv, V)) return false;
  Sc c;
  sc_from_bytes(c, chal);
  acc.Mul(Pp, c.v);
  for (int j = 0; j < nt; j++) {
    curvePoint B;
    if (!g1_unmarshal(B, bases + 64 * j)) return false;
    gfP raw;
    gfpUnmarshal(raw, resp + 32 * j);
    if (geq(raw.v, MN.m)) return false;  // mod.Int.UnmarshalBinary: value out of range
    Sc r;
    memcpy(r.v, raw.v, 32);
    t.Mul(B, r.v);
    acc.Add(acc, t);
  }
  uint8_t a[64], b[64];
  g1_marshal(a, acc);
  g1_marshal(b, Vv);
  return memcmp(a, b, 64) == 0;
}
struct DlogTable {  // PointToInt map keyed by the marshalled point, grown lazily like discreteLog
  std::unordered_map<std::string, int64_t> map;
  curvePoint currentGreatestM;
  int64_t currentGreatestInt = 0;
  std::mutex mu;
};
static DlogTable g_table;
static const int64_t MaxHomomorphicInt = 100000;

static int64_t discreteLog(const curvePoint& P, bool checkNeg, bool* found) {
  uint8_t key[64];
  g1_marshal(key, P);
  std::string ks((const char*)key, 64);
  std::lock_guard<std::mutex> lk(g_table.mu);
  auto it = g_table.map.find(ks);
  if (it != g_table.map.end()) { *found = true; return it->second; }
  if (g_table.currentGreatestInt == 0) g_table.currentGreatestM.SetInfinity();
  bool foundPos = false, foundNeg = false;
  curvePoint guess = g_table.currentGreatestM, guessNeg;
  int64_t guessInt = g_table.currentGreatestInt, guessIntMinus = 0;
  bool start = true;
  for (int64_t i = guessInt; i < MaxHomomorphicInt && !foundPos && !foundNeg; i++) {
    if (!start) guess.Add(guess, curveGen);
    start = false;
    guessInt = i;
    uint8_t gk[64];
    g1_marshal(gk, guess);
    g_table.map[std::string((const char*)gk, 64)] = guessInt;
    if (checkNeg) {
      guessIntMinus = -guessInt;
      guessNeg.Neg(guess);
      uint8_t nk[64];
      g1_marshal(nk, guessNeg);
      g_table.map[std::string((const char*)nk, 64)] = guessIntMinus;
      if (!memcmp(nk, key, 64)) foundNeg = true;
    }
    if (!foundPos && !memcmp(gk, key, 64)) foundPos = true;
  }
  g_table.currentGreatestM = guess;
  g_table.currentGreatestInt = guessInt;
  if (!foundPos && !foundNeg) { *found = false; return 0; }
  *found = true;
  return foundNeg ? guessIntMinus : guessInt;
}

// ------------------------------------------------------------------ range proofs (lib/range/range_proof.go)
static std::vector<int64_t> ToBase(int64_t n, int64_t b, int l) {  // :584-595
  std::vector<int64_t> digits;
  while (n > 0) {
    digits.push_back(n % b);
    n /= b;
  }
  while ((int)digits.size() < l) digits.push_back(0);
  return digits;
}
static int64_t ipow(int64_t u, int j) {
  int64_t r = 1;
  for (int k = 0; k < j; k++) r *= u;
  return r;
}

static size_t rp_create_one(uint8_t* out, int ns, int64_t u, int64_t l, const uint8_t* Y, const uint8_t* A,
                            const uint8_t* capub, int64_t secret, const uint8_t r32[32], const uint8_t ct[128],
                            const uint8_t* rnd) {
  if (u == 0 && l == 0) { memcpy(out, ct, 128); return 128; }  // :322-324
  std::vector<int64_t> base = ToBase(secret, u, (int)l);          // :329
  const int nd = (int)base.size();
  curvePoint commit, caPub;
  g1_unmarshal(commit, ct + 64);
  g1_unmarshal(caPub, capub);
  std::vector<curvePoint> Ys(ns);
  for (int i = 0; i < ns; i++) g1_unmarshal(Ys[i], Y + 64 * i);
  // c = Hash(B, Commitment, sum Y)  :349-374
  uint8_t hin[192], hout[64];
  g1_marshal(hin, curveGen);
  g1_marshal(hin + 64, commit);
  curvePoint sumPub;
  sumPub.SetInfinity();
  for (int i = 0; i < ns; i++) sumPub.Add(sumPub, Ys[i]);
  g1_marshal(hin + 128, sumPub);
  sha3_512(hout, hin, 192);
  Sc c = sc_from_wide(hout);
  curvePoint D;
  D.SetInfinity();
  Sc m = {{0, 0, 0, 0}};
  std::vector<Sc> Zphi(nd);
  std::vector<std::vector<Sc>> ZV(ns, std::vector<Sc>(nd));
  std::vector<std::vector<twistPoint>> V(ns, std::vector<twistPoint>(nd));
  std::vector<std::vector<gfP12>> a(ns, std::vector<gfP12>(nd));
  const uint8_t* rp = rnd;
  for (int j = 0; j < nd; j++) {
    Sc sj, tj, mj;
    sc_from_bytes(sj, rp); rp += 32;   // :377
    sc_from_bytes(tj, rp); rp += 32;   // :378
    sc_from_bytes(mj, rp); rp += 32;   // :379
    m = sc_add(m, mj);
    curvePoint firstT, secondT;
    Sc e = sc_mul(sj, sc_int64(ipow(u, j)));
    firstT.Mul(curveGen, e.v);         // :383
    D.Add(D, firstT);
    secondT.Mul(caPub, mj.v);          // :385
    D.Add(D, secondT);
    Zphi[j] = sc_sub(sj, sc_mul(c, sc_int64(base[j])));  // :388
    for (int i = 0; i < ns; i++) {
      Sc v;
      sc_from_bytes(v, rp); rp += 32;  // :390
      twistPoint sig;
      g2_unmarshal(sig, A + 128 * ((size_t)i * u + base[j]));
      V[i][j].Mul(sig, v.v);           // :392
      curvePoint p1, p2;
      Sc ns_ = sc_neg(sj);
      p1.Mul(curveGen, ns_.v);
      gfP12 e1, e2;
      optimalAte(e1, V[i][j], p1);     // :396
      p2.Mul(curveGen, tj.v);
      optimalAte(e2, twistGen, p2);    // :397
      a[i][j].Mul(e1, e2);
      ZV[i][j] = sc_sub(tj, sc_mul(c, v));  // :398
    }
  }
  Sc rr;
  sc_from_bytes(rr, r32);
  Sc Zr = sc_sub(m, sc_mul(c, rr));    // :401
  // ToBytes :88-155
  uint8_t* o = out;
  memcpy(o, ct, 128); o += 128;
  sc_to_bytes(o, c); o += 32;
  sc_to_bytes(o, Zr); o += 32;
  g1_marshal(o, D); o += 64;
  for (int j = 0; j < nd; j++) { sc_to_bytes(o, Zphi[j]); o += 32; }
  for (int i = 0; i < ns; i++) {
    for (int j = 0; j < nd; j++) { g2_marshal(o, V[i][j]); o += 128; }
    for (int j = 0; j < nd; j++) { gt_marshal(o, a[i][j]); o += 384; }
    for (int j = 0; j < nd; j++) { sc_to_bytes(o, ZV[i][j]); o += 32; }
  }
  return (size_t)(o - out);
}

static bool rp_verify_one(const uint8_t* proof, size_t len, int ns, int64_t u, int64_t l, const uint8_t* Y,
                          const uint8_t* Pb) {
  if (l == 0 && u == 0) return true;  // :508-510
  const size_t per = 32 + 544 * (size_t)ns;
  if (len < 256 + per || (len - 256) % per) return false;
  const int nd = (int)((len - 256) / per);
  if (nd != l) return false;          // :512
  curvePoint C, K, D, P;
  if (!g1_unmarshal(K, proof) || !g1_unmarshal(C, proof + 64) || !g1_unmarshal(D, proof + 192)) return false;
  g1_unmarshal(P, Pb);
  Sc c, Zr;
  sc_from_bytes(c, proof + 128);
  sc_from_bytes(Zr, proof + 160);
  std::vector<Sc> Zphi(nd);
  for (int j = 0; j < nd; j++) sc_from_bytes(Zphi[j], proof + 256 + 32 * j);
  std::vector<curvePoint> Ys(ns);
  for (int i = 0; i < ns; i++) g1_unmarshal(Ys[i], Y + 64 * i);
  // Dp = c*C + Zr*P + sum_j Zphi_j * (u^j * B)   :520-529
  curvePoint Dp, t1;
  Dp.Mul(C, c.v);
  t1.Mul(P, Zr.v);
  Dp.Add(Dp, t1);
  bool result = true;
  for (int j = 0; j < nd; j++) {
    curvePoint point;
    Sc uj = sc_int64(ipow(u, j));
    point.Mul(curveGen, uj.v);        // IntToPoint(u^j)
    point.Mul(point, Zphi[j].v);
    Dp.Add(Dp, point);
    for (int i = 0; i < ns; i++) {    // :539-550
      const uint8_t* blk = proof + 256 + 32 * nd + (size_t)i * 544 * nd;
      twistPoint Vij;
      gfP12 Aij;
      if (!g2_unmarshal(Vij, blk + 128 * j) || !gt_unmarshal(Aij, blk + 128 * nd + 384 * j)) return false;
      Sc Zv;
      sc_from_bytes(Zv, blk + 512 * nd + 32 * j);
      curvePoint p1, p2, p3;
      gfP12 ap, e2, e3;
      p1.Mul(Ys[i], c.v);
      optimalAte(ap, Vij, p1);        // :540
      Sc nz = sc_neg(Zphi[j]);
      p2.Mul(curveGen, nz.v);
      optimalAte(e2, Vij, p2);        // :542
      ap.Mul(ap, e2);
      p3.Mul(curveGen, Zv.v);
      optimalAte(e3, twistGen, p3);   // :544
      ap.Mul(ap, e3);
      if (!(ap == Aij)) result = false;  // :546-549
    }
  }
  uint8_t d1[64], d2[64];
  g1_marshal(d1, Dp);
  g1_marshal(d2, D);
  if (memcmp(d1, d2, 64)) return false;  // :561
  return result;
}

// ------------------------------------------------------------------ C entry points (ctypes)
extern "C" {

int orc_g1_mul(const uint8_t* k32, const uint8_t* pt, uint8_t* out) {
  ensure_init();
  curvePoint p, r;
  if (!g1_unmarshal(p, pt)) return -1;
  Sc k;
  sc_from_bytes(k, k32);
  r.Mul(p, k.v);
  g1_marshal(out, r);
  return 0;
}
int orc_g2_mul(const uint8_t* k32, const uint8_t* pt, uint8_t* out) {
  ensure_init();
  twistPoint p, r;
  if (!g2_unmarshal(p, pt)) return -1;
  Sc k;
  sc_from_bytes(k, k32);
  r.Mul(p, k.v);
  g2_marshal(out, r);
  return 0;
}
int orc_pair(const uint8_t* g1, const uint8_t* g2, uint8_t* gt) {
  ensure_init();
  curvePoint p;
  twistPoint q;
  if (!g1_unmarshal(p, g1) || !g2_unmarshal(q, g2)) return -1;
  gfP12 e;
  optimalAte(e, q, p);
  gt_marshal(gt, e);
  return 0;
}
int orc_sha3_512(const uint8_t* in, size_t len, uint8_t* out) {
  sha3_512(out, in, len);
  return 0;
}
void orc_generators(uint8_t* g1, uint8_t* g2) {
  ensure_init();
  g1_marshal(g1, curveGen);
  g2_marshal(g2, twistGen);
}

// EncryptIntVector with injected r (one task per value, like the goroutine fan-out of lib/encoding/*.go)
int orc_encrypt(size_t n, const int64_t* m, const uint8_t* r, const uint8_t* P, uint8_t* ct, int threads) {
  ensure_init();
  curvePoint pub;
  if (!g1_unmarshal(pub, P)) return -1;
  parallel_for(n, threads, [&](size_t i) { encrypt_one(ct + 128 * i, pub, m[i], r + 32 * i); });
  return 0;
}

int orc_encrypt_scalar(size_t n, const uint8_t* m, const uint8_t* r, const uint8_t* P, uint8_t* ct, int threads) {
  ensure_init();
  curvePoint pub;
  if (!g1_unmarshal(pub, P)) return -2;
  parallel_for(n, threads, [&](size_t i) { encrypt_scalar_one(ct + 128 * i, pub, m + 32 * i, r + 32 * i); });
  return 0;
}

int orc_ct_scalar_mul(size_t n, const uint8_t* s, const uint8_t* ct, uint8_t* out, int threads) {
  ensure_init();
  std::atomic<int> bad(0);
  parallel_for(n, threads, [&](size_t i) {
    if (!ct_scalar_mul_one(out + 128 * i, ct + 128 * i, s + 32 * i)) bad++;
  });
  return bad ? -2 : 0;
}

// libunlynxshuffle.ShuffleSequence(inputList, g, h, nil) with beta and pi injected (unlynx v1.4.0 lib/shuffle, recalled;
// call sites data/data.go:68, services/service.go:618-660): outputList[i][j].Rerandomize(inputList[pi[i]][j], b, b, _, g, h)
// with b = beta[pi[i]][j]: K = K1 + b*g, C = C1 + b*h
int orc_shuffle_rerandomize(size_t lines, size_t nq, const int32_t* pi, const uint8_t* in, const uint8_t* beta,
                            const uint8_t* P, uint8_t* out, int threads) {
  ensure_init();
  curvePoint h;
  if (!g1_unmarshal(h, P)) return -2;
  std::atomic<int> bad(0);
  parallel_for(lines * nq, threads, [&](size_t t) {
    const size_t i = t / nq, j = t % nq, src = (size_t)pi[i] * nq + j;
    curvePoint K1, C1, tmp1, tmp2, K, C;
    if (!g1_unmarshal(K1, in + 128 * src) || !g1_unmarshal(C1, in + 128 * src + 64)) { bad++; return; }
    Sc b;
    sc_from_bytes(b, beta + 32 * src);
    tmp1.Mul(curveGen, b.v);
    tmp2.Mul(h, b.v);
    K.Add(K1, tmp1);
    C.Add(C1, tmp2);
    g1_marshal(out + 128 * t, K);
    g1_marshal(out + 128 * t + 64, C);
  });
  return bad ? -2 : 0;
}

int orc_schnorr_verify(size_t n, const uint8_t* pub, const uint8_t* sig, const uint8_t* digest, uint8_t* ok, int threads) {
  ensure_init();
  parallel_for(n, threads, [&](size_t i) { ok[i] = schnorr_verify_one(pub + 64 * i, sig + 96 * i, digest + 64 * i) ? 1 : 0; });
  return 0;
}
int orc_rep_commit(size_t n, int nt, const uint8_t* sc, const uint8_t* pts, uint8_t* out, int threads) {
  ensure_init();
  std::atomic<int> bad(0);
  parallel_for(n, threads, [&](size_t i) {
    if (!rep_commit_one(out + 64 * i, nt, sc + 32 * nt * i, pts + 64 * nt * i)) bad++;
  });
  return bad ? -2 : 0;
}
int orc_rep_respond(size_t n, int per, const uint8_t* v, const uint8_t* chal, const uint8_t* x, uint8_t* r) {
  ensure_init();
  for (size_t i = 0; i < n; i++) {
    Sc vi, ci, xi;
    sc_from_bytes(vi, v + 32 * i);
    sc_from_bytes(ci, chal + 32 * (i / per));
    sc_from_bytes(xi, x + 32 * i);
    sc_to_bytes(r + 32 * i, sc_sub(vi, sc_mul(ci, xi)));
  }
  return 0;
}
int orc_rep_verify(size_t n, int nt, const uint8_t* resp, const uint8_t* chal, const uint8_t* bases, const uint8_t* P,
                   const uint8_t* V, uint8_t* ok, int threads) {
  ensure_init();
  parallel_for(n, threads, [&](size_t i) {
    ok[i] = rep_verify_one(nt, resp + 32 * nt * i, chal + 32 * i, bases + 64 * nt * i, P + 64 * i, V + 64 * i) ? 1 : 0;
  });
  return 0;
}
// first half of Sign: R = k*B (the caller hashes R || X || msg and calls orc_schnorr_finish)
int orc_schnorr_commit(const uint8_t* k32, uint8_t* R64) {
  ensure_init();
  Sc k;
  sc_from_bytes(k, k32);
  curvePoint R;
  R.Mul(curveGen, k.v);
  g1_marshal(R64, R);
  return 0;
}
// second half of Sign: s = k + x*h
int orc_schnorr_finish(const uint8_t* x32, const uint8_t* k32, const uint8_t* digest64, uint8_t* s32) {
  ensure_init();
  Sc x, k;
  sc_from_bytes(x, x32);
  sc_from_bytes(k, k32);
  sc_to_bytes(s32, sc_add(k, sc_mul(x, sc_from_wide(digest64))));
  return 0;
}

// KeySwitchSequence over n ciphertexts' K components (rb: n x 64 B) with injected randomness v (n x 32 B)
int orc_keyswitch(size_t n, const uint8_t* target, const uint8_t* sk32, const uint8_t* rb, const uint8_t* v, uint8_t* cv,
                  uint8_t* rb_neg, int threads) {
  ensure_init();
  curvePoint Q;
  if (!g1_unmarshal(Q, target)) return -2;
  Sc x;
  sc_from_bytes(x, sk32);
  std::atomic<int> bad(0);
  parallel_for(n, threads, [&](size_t i) {
    if (!keyswitch_one(cv + 128 * i, rb_neg ? rb_neg + 64 * i : nullptr, Q, x, rb + 64 * i, v + 32 * i)) bad++;
  });
  return bad ? -2 : 0;
}

// FromBytes + CipherVector.Add folded over ndp providers (protocols/data_collection_protocol.go:145-166);
// cts is [ndp][v][128] contiguous. The running sums stay Jacobian; ToBytes at the end.
int orc_aggregate(size_t ndp, size_t v, const uint8_t* cts, uint8_t* out, int threads) {
  ensure_init();
  std::atomic<int> bad(0);
  parallel_for(v, threads, [&](size_t j) {
    curvePoint K, C, k, c;
    K.SetInfinity();
    C.SetInfinity();
    for (size_t d = 0; d < ndp; d++) {
      const uint8_t* w = cts + (d * v + j) * 128;
      if (!g1_unmarshal(k, w) || !g1_unmarshal(c, w + 64)) { bad++; continue; }
      K.Add(K, k);
      C.Add(C, c);
    }
    g1_marshal(out + 128 * j, K);
    g1_marshal(out + 128 * j + 64, C);
  });
  return bad ? -2 : 0;
}

void orc_dlog_reset() {
  std::lock_guard<std::mutex> lk(g_table.mu);
  g_table.map.clear();
  g_table.currentGreatestInt = 0;
}
// DecryptInt(WithNeg) per element: decryptPoint then discreteLog
int orc_decrypt(size_t n, const uint8_t* sk32, const uint8_t* ct, int with_neg, int64_t* out, uint8_t* found,
                int threads) {
  ensure_init();
  Sc sk;
  sc_from_bytes(sk, sk32);
  std::atomic<int> bad(0);
  parallel_for(n, threads, [&](size_t i) {
    curvePoint K, C, S, M;
    if (!g1_unmarshal(K, ct + 128 * i) || !g1_unmarshal(C, ct + 128 * i + 64)) { bad++; out[i] = 0; return; }
    S.Mul(K, sk.v);
    S.Neg(S);
    M.Add(C, S);
    bool f;
    out[i] = discreteLog(M, with_neg != 0, &f);
    if (found) found[i] = f;
  });
  return bad ? -2 : 0;
}
int orc_decrypt_check_zero(size_t n, const uint8_t* sk32, const uint8_t* ct, int64_t* out) {
  ensure_init();
  Sc sk;
  sc_from_bytes(sk, sk32);
  for (size_t i = 0; i < n; i++) {
    curvePoint K, C, S, M;
    if (!g1_unmarshal(K, ct + 128 * i) || !g1_unmarshal(C, ct + 128 * i + 64)) return -2;
    S.Mul(K, sk.v);
    S.Neg(S);
    M.Add(C, S);
    M.MakeAffine();
    out[i] = M.IsInfinity() ? 0 : 1;
  }
  return 0;
}

// InitRangeProofSignature from caller-drawn x_i (:249-288)
int orc_make_sigs(int ns, int64_t u, const uint8_t* x, uint8_t* Y, uint8_t* A) {
  ensure_init();
  for (int i = 0; i < ns; i++) {
    Sc xi;
    sc_from_bytes(xi, x + 32 * i);
    curvePoint y;
    y.Mul(curveGen, xi.v);
    g1_marshal(Y + 64 * i, y);
    for (int64_t k = 0; k < u; k++) {
      Sc e = sc_inv(sc_add(xi, sc_int64(k)));
      twistPoint t;
      t.Mul(twistGen, e.v);
      g2_marshal(A + 128 * ((size_t)i * u + k), t);
    }
  }
  return 0;
}

size_t orc_rp_proof_size(int ns, int64_t nd) { return 256 + 32 * (size_t)nd + 544 * (size_t)ns * (size_t)nd; }

// CreatePredicateRangeProofListForAllServers: one task per proof (:305-317). All proofs share (ns, u, l, sigs).
// rnd holds, per proof, nd*(3+ns) scalars; rnd_off / proof_off are byte offsets (n+1 entries).
int orc_rp_create(size_t n, int ns, int64_t u, int64_t l, const uint8_t* Y, const uint8_t* A, const uint8_t* P,
                  const int64_t* secret, const uint8_t* r, const uint8_t* ct, const uint8_t* rnd, const size_t* rnd_off,
                  uint8_t* proofs, const size_t* proof_off, int threads) {
  ensure_init();
  parallel_for(n, threads, [&](size_t p) {
    rp_create_one(proofs + proof_off[p], ns, u, l, Y, A, P, secret[p], r + 32 * p, ct + 128 * p, rnd + rnd_off[p]);
  });
  return 0;
}
// RangeProofListVerification: one task per proof (:484-501)
int orc_rp_verify(size_t n, int ns, int64_t u, int64_t l, const uint8_t* Y, const uint8_t* P, const uint8_t* proofs,
                  const size_t* proof_off, uint8_t* ok, int threads) {
  ensure_init();
  parallel_for(n, threads, [&](size_t p) {
    ok[p] = rp_verify_one(proofs + proof_off[p], proof_off[p + 1] - proof_off[p], ns, u, l, Y, P) ? 1 : 0;
  });
  return 0;
}

}  // extern "C"
