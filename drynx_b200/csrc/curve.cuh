// Short-Weierstrass y^2 = x^3 + b (a = 0) point arithmetic shared by G1 (over Fp, b = 3) and the sextic twist
// G2 (over Fp2, b = 3/xi). Replaces kyber v3.0.8 pairing/bn256 curve.go / twist.go (curvePoint / twistPoint
// Add, Double, Mul, MakeAffine; not in /root/reference) and the kyber.Point operations that libunlynx and
// lib/range call.
//
// GPU-side choices: Jacobian coordinates without kyber's cached t = z^2; mixed additions against affine comb
// tables; 4-bit fixed windows for variable-base multiplication. Group elements are canonical once affine, so
// the wire bytes match any other correct algorithm.
#pragma once
#include "tower.cuh"

namespace drx {

// field-generic spellings
DRX_HD Fp f_add(const Fp& a, const Fp& b) { return fp_add(a, b); }
DRX_HD Fp f_sub(const Fp& a, const Fp& b) { return fp_sub(a, b); }
DRX_HD Fp f_mul(const Fp& a, const Fp& b) { return fp_mul(a, b); }
DRX_HD Fp f_sqr(const Fp& a) { return fp_sqr(a); }
DRX_HD Fp f_dbl(const Fp& a) { return fp_dbl(a); }
DRX_HD Fp f_neg(const Fp& a) { return fp_neg(a); }
DRX_HD bool f_is_zero(const Fp& a) { return fp_is_zero(a); }
DRX_HD bool f_eq(const Fp& a, const Fp& b) { return fp_eq(a, b); }
DRX_HD Fp f_inv(const Fp& a) { return fp_inv(a); }
DRX_HD void f_set_zero(Fp& a) { a = fp_zero(); }
DRX_HD void f_set_one(Fp& a) { a = fp_one(); }
DRX_HD Fp f_curve_b(const Fp&) { return K_CURVE_B(); }

DRX_HD Fp2 f_add(const Fp2& a, const Fp2& b) { return fp2_add(a, b); }
DRX_HD Fp2 f_sub(const Fp2& a, const Fp2& b) { return fp2_sub(a, b); }
DRX_HD Fp2 f_mul(const Fp2& a, const Fp2& b) { return fp2_mul(a, b); }
DRX_HD Fp2 f_sqr(const Fp2& a) { return fp2_sqr(a); }
DRX_HD Fp2 f_dbl(const Fp2& a) { return fp2_dbl(a); }
DRX_HD Fp2 f_neg(const Fp2& a) { return fp2_neg(a); }
DRX_HD bool f_is_zero(const Fp2& a) { return fp2_is_zero(a); }
DRX_HD bool f_eq(const Fp2& a, const Fp2& b) { return fp2_eq(a, b); }
DRX_HD Fp2 f_inv(const Fp2& a) { return fp2_inv(a); }
DRX_HD void f_set_zero(Fp2& a) { a = fp2_zero(); }
DRX_HD void f_set_one(Fp2& a) { a = fp2_one(); }
DRX_HD Fp2 f_curve_b(const Fp2&) { return K_TWIST_B(); }

template <class F>
struct Aff {  // affine; the point at infinity is (0, 0), as on the wire
  F x, y;
};
template <class F>
struct Jac {  // (X/Z^2, Y/Z^3); infinity has Z = 0
  F x, y, z;
};
typedef Aff<Fp> G1Aff;
typedef Jac<Fp> G1Jac;
typedef Aff<Fp2> G2Aff;
typedef Jac<Fp2> G2Jac;

template <class F>
DRX_HD bool aff_is_inf(const Aff<F>& p) { return f_is_zero(p.x) & f_is_zero(p.y); }
template <class F>
DRX_HD bool jac_is_inf(const Jac<F>& p) { return f_is_zero(p.z); }
template <class F>
DRX_HD Jac<F> jac_inf() {
  Jac<F> r;
  f_set_zero(r.x);
  f_set_one(r.y);
  f_set_zero(r.z);
  return r;
}
template <class F>
DRX_HD Jac<F> jac_from_aff(const Aff<F>& p) {
  Jac<F> r;
  r.x = p.x;
  r.y = p.y;
  if (aff_is_inf(p)) {
    f_set_one(r.y);
    f_set_zero(r.z);
  } else {
    f_set_one(r.z);
  }
  return r;
}
template <class F>
DRX_HD Aff<F> aff_neg(const Aff<F>& p) { return Aff<F>{p.x, f_neg(p.y)}; }
template <class F>
DRX_HD Jac<F> jac_neg(const Jac<F>& p) { return Jac<F>{p.x, f_neg(p.y), p.z}; }
template <class F>
DRX_HD bool aff_on_curve(const Aff<F>& p) {
  if (aff_is_inf(p)) return true;
  const F lhs = f_sqr(p.y);
  const F rhs = f_add(f_mul(f_sqr(p.x), p.x), f_curve_b(p.x));
  return f_eq(lhs, rhs);
}

// dbl-2009-l
template <class F>
DRX_NI Jac<F> jac_dbl(const Jac<F>& p) {
  const F A = f_sqr(p.x);
  const F B = f_sqr(p.y);
  const F C = f_sqr(B);
  F D = f_sub(f_sub(f_sqr(f_add(p.x, B)), A), C);
  D = f_dbl(D);
  const F E = f_add(f_dbl(A), A);
  const F Fq = f_sqr(E);
  Jac<F> r;
  r.x = f_sub(Fq, f_dbl(D));
  F C8 = f_dbl(C);
  C8 = f_dbl(C8);
  C8 = f_dbl(C8);
  r.z = f_dbl(f_mul(p.y, p.z));
  r.y = f_sub(f_mul(E, f_sub(D, r.x)), C8);
  return r;
}

// madd-2007-bl with the exceptional cases (p infinite, q infinite, p == q, p == -q) handled
template <class F>
DRX_NI Jac<F> jac_add_aff(const Jac<F>& p, const Aff<F>& q) {
  if (aff_is_inf(q)) return p;
  if (jac_is_inf(p)) return jac_from_aff(q);
  const F Z1Z1 = f_sqr(p.z);
  const F U2 = f_mul(q.x, Z1Z1);
  const F S2 = f_mul(f_mul(q.y, p.z), Z1Z1);
  const F H = f_sub(U2, p.x);
  const F rr = f_sub(S2, p.y);
  if (f_is_zero(H)) {
    if (f_is_zero(rr)) return jac_dbl(p);
    return jac_inf<F>();
  }
  const F HH = f_sqr(H);
  F I = f_dbl(HH);
  I = f_dbl(I);
  const F J = f_mul(H, I);
  const F r2 = f_dbl(rr);
  const F V = f_mul(p.x, I);
  Jac<F> r;
  r.x = f_sub(f_sub(f_sqr(r2), J), f_dbl(V));
  r.y = f_sub(f_mul(r2, f_sub(V, r.x)), f_dbl(f_mul(p.y, J)));
  r.z = f_sub(f_sub(f_sqr(f_add(p.z, H)), Z1Z1), HH);
  return r;
}

// add-2007-bl, complete via the same exceptional-case branches
template <class F>
DRX_NI Jac<F> jac_add(const Jac<F>& p, const Jac<F>& q) {
  if (jac_is_inf(q)) return p;
  if (jac_is_inf(p)) return q;
  const F Z1Z1 = f_sqr(p.z);
  const F Z2Z2 = f_sqr(q.z);
  const F U1 = f_mul(p.x, Z2Z2);
  const F U2 = f_mul(q.x, Z1Z1);
  const F S1 = f_mul(f_mul(p.y, q.z), Z2Z2);
  const F S2 = f_mul(f_mul(q.y, p.z), Z1Z1);
  const F H = f_sub(U2, U1);
  const F rr = f_sub(S2, S1);
  if (f_is_zero(H)) {
    if (f_is_zero(rr)) return jac_dbl(p);
    return jac_inf<F>();
  }
  const F I = f_sqr(f_dbl(H));
  const F J = f_mul(H, I);
  const F r2 = f_dbl(rr);
  const F V = f_mul(U1, I);
  Jac<F> r;
  r.x = f_sub(f_sub(f_sqr(r2), J), f_dbl(V));
  r.y = f_sub(f_mul(r2, f_sub(V, r.x)), f_dbl(f_mul(S1, J)));
  r.z = f_mul(f_sub(f_sub(f_sqr(f_add(p.z, q.z)), Z1Z1), Z2Z2), H);
  return r;
}

template <class F>
DRX_HD Aff<F> jac_to_aff_with_zinv(const Jac<F>& p, const F& zinv) {
  Aff<F> r;
  if (jac_is_inf(p)) {
    f_set_zero(r.x);
    f_set_zero(r.y);
    return r;
  }
  const F zi2 = f_sqr(zinv);
  r.x = f_mul(p.x, zi2);
  r.y = f_mul(f_mul(p.y, zi2), zinv);
  return r;
}
template <class F>
DRX_HD Aff<F> jac_to_aff(const Jac<F>& p) {
  if (jac_is_inf(p)) {
    Aff<F> r;
    f_set_zero(r.x);
    f_set_zero(r.y);
    return r;
  }
  return jac_to_aff_with_zinv(p, f_inv(p.z));
}
// projective equality (kyber Point.Equal compares after MakeAffine)
template <class F>
DRX_HD bool jac_eq_aff(const Jac<F>& p, const Aff<F>& q) {
  if (jac_is_inf(p) | aff_is_inf(q)) return jac_is_inf(p) & aff_is_inf(q);
  const F zz = f_sqr(p.z);
  return f_eq(p.x, f_mul(q.x, zz)) & f_eq(p.y, f_mul(f_mul(q.y, zz), p.z));
}

// ------------------------------------------------------------------ scalars: 256-bit plain (non-Montgomery) mod n
struct Scalar {
  uint32_t v[8];
};
DRX_HD Scalar sc_zero() {
  Scalar s;
  DRX_UNROLL
  for (int k = 0; k < 8; k++) s.v[k] = 0;
  return s;
}
DRX_HD Scalar sc_add(const Scalar& a, const Scalar& b) { Scalar r; mod_add<ModN>(r.v, a.v, b.v); return r; }
DRX_HD Scalar sc_sub(const Scalar& a, const Scalar& b) { Scalar r; mod_sub<ModN>(r.v, a.v, b.v); return r; }
DRX_HD Scalar sc_neg(const Scalar& a) { Scalar r; mod_neg<ModN>(r.v, a.v); return r; }
DRX_HD Scalar sc_mul(const Scalar& a, const Scalar& b) {  // plain * plain -> plain: (a*b/R) * R^2 / R
  Scalar t, r2, r;
  mont_mul<ModN>(t.v, a.v, b.v);
  DRX_UNROLL
  for (int k = 0; k < 8; k++) r2.v[k] = ModN::r2(k);
  mont_mul<ModN>(r.v, t.v, r2.v);
  return r;
}
DRX_HD Scalar sc_from_u64(uint64_t x) {
  Scalar s = sc_zero();
  s.v[0] = (uint32_t)x;
  s.v[1] = (uint32_t)(x >> 32);
  return s;
}
// kyber mod.Int.SetInt64: negative values map to n - |v|
DRX_HD Scalar sc_from_i64(int64_t x) {
  if (x >= 0) return sc_from_u64((uint64_t)x);
  return sc_neg(sc_from_u64((uint64_t)0 - (uint64_t)x));
}
// 32 big-endian bytes (as 8 words in memory order) -> scalar reduced mod n (kyber Scalar.SetBytes /
// UnmarshalBinary reduce silently). One conditional subtraction suffices: 2^256 < 2n.
DRX_HD Scalar sc_from_wire(const uint32_t w[8]) {
  Scalar t, d;
  limbs_from_be_words(t.v, w);
  d.v[0] = sub_cc(t.v[0], ModN::mod(0));
  DRX_UNROLL
  for (int k = 1; k < 8; k++) d.v[k] = subc_cc(t.v[k], ModN::mod(k));
  const bool lt = subc(0, 0) != 0;
  DRX_UNROLL
  for (int k = 0; k < 8; k++) d.v[k] = lt ? t.v[k] : d.v[k];
  return d;
}
// 64-byte digest (16 words in memory order) read as a big-endian integer, reduced mod n: kyber Scalar.SetBytes on
// hash.Sum (lib/range/range_proof.go:374 for SHA3-512; sign/schnorr hash() for SHA-512)
DRX_HD Scalar sc_from_digest(const uint32_t d[16]) {
  const Scalar hi = sc_from_wire(d);
  const Scalar lo = sc_from_wire(d + 8);
  Scalar r2, t;
  DRX_UNROLL
  for (int k = 0; k < 8; k++) r2.v[k] = ModN::r2(k);
  mont_mul<ModN>(t.v, hi.v, r2.v);  // hi * 2^256 mod n
  return sc_add(t, lo);
}
DRX_HD void sc_to_wire(uint32_t w[8], const Scalar& s) { be_words_from_limbs(w, s.v); }
DRX_HD uint32_t sc_window(const Scalar& s, int bit, int width) {  // bits [bit, bit+width), width <= 16
  const int limb = bit >> 5, off = bit & 31;
  uint64_t x = s.v[limb];
  if (limb < 7) x |= (uint64_t)s.v[limb + 1] << 32;
  return (uint32_t)(x >> off) & ((1u << width) - 1u);
}

// ------------------------------------------------------------------ scalar multiplication
// Fixed base: table[w * 255 + (d - 1)] = d * 2^(8w) * Base, affine, 32 windows of 8 bits.
#define DRX_COMB_WINDOWS 32
#define DRX_COMB_ENTRIES 255
template <class F>
DRX_HD Jac<F> comb_mul(const Aff<F>* __restrict__ table, const Scalar& k, int windows = DRX_COMB_WINDOWS) {
  Jac<F> acc = jac_inf<F>();
  for (int w = 0; w < windows; w++) {
    const uint32_t d = (k.v[w >> 2] >> ((w & 3) * 8)) & 0xffu;
    if (d) {
      const Aff<F> q = table[w * DRX_COMB_ENTRIES + (int)d - 1];
      acc = jac_add_aff(acc, q);
    }
  }
  return acc;
}

// Variable base, 4-bit fixed window, most significant window first; `nibbles` = how many windows the scalar can occupy.
template <class F>
DRX_HD Jac<F> window_mul(const Jac<F>& base, const Scalar& k, int nibbles = 64) {
  Jac<F> tab[16];
  tab[0] = jac_inf<F>();
  tab[1] = base;
  for (int i = 2; i < 16; i++) tab[i] = (i & 1) ? jac_add(tab[i - 1], base) : jac_dbl(tab[i >> 1]);
  Jac<F> acc = jac_inf<F>();
  for (int w = nibbles - 1; w >= 0; w--) {
    if (w != nibbles - 1) {
      acc = jac_dbl(acc);
      acc = jac_dbl(acc);
      acc = jac_dbl(acc);
      acc = jac_dbl(acc);
    }
    const uint32_t d = (k.v[w >> 3] >> ((w & 7) * 4)) & 15u;
    if (d) acc = jac_add(acc, tab[d]);
  }
  return acc;
}


// ------------------------------------------------------------------ GLV on G1 (verification header check)
// y^2 = x^3 + 3 over Fp with p = 1 mod 3 has the endomorphism phi(x, y) = (beta x, y), beta^3 = 1, and on the group of
// order n phi(P) = lambda P with lambda = 36u^4 - 1 (lambda^2 + lambda + 1 = 0 mod n). A scalar splits as
// k = k1 + k2 lambda (mod n) with |k1|, |k2| < 2^128: from the lattice basis (a1, -a2), (a2, b2) of {(x, y): x + y lambda = 0},
// a1 = 6u^2 + 2u, a2 = 2u + 1, b2 = a1 + a2, take c1 = round(b2 k / n), c2 = round(a2 k / n) and
// k1 = k - c1 a1 - c2 a2, k2 = c1 a2 - c2 b2. k P = k1 P + k2 phi(P): two products of half length that two threads can do
// side by side. The roundings use g1 = floor(2^256 b2 / n), g2 = floor(2^256 a2 / n); being off by one only makes k1, k2
// a little longer, the identity k = k1 + k2 lambda holds for any c1, c2 (the CPU tests recompute the constants from u
// and check the split and the product through the DRX_HOST_EMU build).
struct GlvSplit {
  Scalar k1, k2;     // magnitudes, below 2^132
  bool neg1, neg2;   // signs
};
#define DRX_GLV_NIBBLES 33
namespace glv_detail {
DRX_HD uint32_t c_a1(int i) { const uint32_t v[4] = {0xdf0d2a08u, 0xb31b1546u, 0x288a7c7au, 0xbfcdfabeu}; return v[i]; }
DRX_HD uint32_t c_a2(int i) { const uint32_t v[2] = {0xd8b10603u, 0xb4ed5d35u}; return v[i]; }
DRX_HD uint32_t c_b2(int i) { const uint32_t v[4] = {0xb7be300bu, 0x6808727cu, 0x288a7c7bu, 0xbfcdfabeu}; return v[i]; }
DRX_HD uint32_t c_g1(int i) { const uint32_t v[5] = {0xf0575aa6u, 0xf06129ebu, 0xffd56e3eu, 0x55ae596cu, 0x00000001u}; return v[i]; }
DRX_HD uint32_t c_g2(int i) { const uint32_t v[3] = {0x740d1660u, 0x424dd4afu, 0x00000001u}; return v[i]; }
// out[0..nout) = limbs 8.. of (k * g + 2^255), g given by the accessor G over ng limbs
template <uint32_t (*G)(int), int NG>
DRX_HD void mul_round_hi(uint32_t* out, const uint32_t k[8]) {
  uint32_t prod[8 + NG + 1];
  for (int i = 0; i < 8 + NG + 1; i++) prod[i] = 0;
  for (int j = 0; j < NG; j++) {
    uint64_t carry = 0;
    for (int i = 0; i < 8; i++) {
      const uint64_t t = (uint64_t)k[i] * G(j) + prod[i + j] + carry;
      prod[i + j] = (uint32_t)t;
      carry = t >> 32;
    }
    prod[8 + j] = (uint32_t)carry;
  }
  uint64_t c = (uint64_t)prod[7] + 0x80000000u;  // + 2^255
  c >>= 32;
  for (int i = 8; i < 8 + NG + 1; i++) {
    c += prod[i];
    prod[i] = (uint32_t)c;
    c >>= 32;
  }
  for (int i = 0; i < NG + 1; i++) out[i] = prod[8 + i];
}
// acc (8 limbs, mod 2^256) -= x * y, x over nx limbs, y by accessor over NY limbs
template <uint32_t (*Y)(int), int NY>
DRX_HD void sub_mul(uint32_t acc[8], const uint32_t* x, int nx) {
  uint32_t prod[8];
  for (int i = 0; i < 8; i++) prod[i] = 0;
  for (int j = 0; j < NY; j++) {
    uint64_t carry = 0;
    for (int i = 0; i < nx && i + j < 8; i++) {
      const uint64_t t = (uint64_t)x[i] * Y(j) + prod[i + j] + carry;
      prod[i + j] = (uint32_t)t;
      carry = t >> 32;
    }
    if (nx + j < 8) prod[nx + j] = (uint32_t)carry;
  }
  uint64_t borrow = 0;
  for (int i = 0; i < 8; i++) {
    const uint64_t t = (uint64_t)acc[i] - prod[i] - borrow;
    acc[i] = (uint32_t)t;
    borrow = (t >> 32) & 1u;
  }
}
DRX_HD bool take_sign(uint32_t v[8]) {  // two's complement -> magnitude; returns "was negative"
  if (!(v[7] >> 31)) return false;
  uint64_t c = 1;
  for (int i = 0; i < 8; i++) {
    c += (uint32_t)~v[i];
    v[i] = (uint32_t)c;
    c >>= 32;
  }
  return true;
}
}  // namespace glv_detail
DRX_HD GlvSplit glv_split(const Scalar& k) {
  using namespace glv_detail;
  uint32_t c1[6], c2[4];
  mul_round_hi<c_g1, 5>(c1, k.v);
  mul_round_hi<c_g2, 3>(c2, k.v);
  GlvSplit r;
  for (int i = 0; i < 8; i++) {
    r.k1.v[i] = k.v[i];
    r.k2.v[i] = 0;
  }
  sub_mul<c_a1, 4>(r.k1.v, c1, 6);   // k1 = k - c1 a1 - c2 a2
  sub_mul<c_a2, 2>(r.k1.v, c2, 4);
  sub_mul<c_b2, 4>(r.k2.v, c2, 4);   // k2 = -c2 b2 + c1 a2: subtract, then negate the part with the other sign
  uint32_t t[8];
  for (int i = 0; i < 8; i++) t[i] = 0;
  sub_mul<c_a2, 2>(t, c1, 6);        // t = -c1 a2
  uint64_t borrow = 0;
  for (int i = 0; i < 8; i++) {      // k2 = k2 - t
    const uint64_t d = (uint64_t)r.k2.v[i] - t[i] - borrow;
    r.k2.v[i] = (uint32_t)d;
    borrow = (d >> 32) & 1u;
  }
  r.neg1 = take_sign(r.k1.v);
  r.neg2 = take_sign(r.k2.v);
  return r;
}
// beta in Montgomery form: phi(x, y) = (beta x, y) = lambda (x, y)
DRX_HD Fp glv_beta() {
  const uint32_t v[8] = {0xe1ada57du, 0x12d3cef5u, 0x3753babbu, 0xe2eca146u, 0xddccf750u, 0x0ca41e40u, 0x0397e04cu, 0x55133706u};
  Fp r;
  for (int i = 0; i < 8; i++) r.v[i] = v[i];
  return r;
}
// k * P in one thread: the two half-length products walked together (Straus), 33 windows of four doublings with up to
// two additions each; the second table is the first under phi (one field product per entry). 2,220 M against 3,040 M
// for the plain 64-window walk.
DRX_HD Jac<Fp> window_mul_glv(const Aff<Fp>& P, const Scalar& k) {
  const GlvSplit s = glv_split(k);
  const Jac<Fp> base = jac_from_aff(s.neg1 ? aff_neg(P) : P);
  Jac<Fp> tab1[16], tab2[16];
  tab1[0] = jac_inf<Fp>();
  tab1[1] = base;
  for (int i = 2; i < 16; i++) tab1[i] = (i & 1) ? jac_add(tab1[i - 1], base) : jac_dbl(tab1[i >> 1]);
  const Fp beta = glv_beta();
  const bool flip = s.neg1 != s.neg2;
  for (int i = 0; i < 16; i++) {
    tab2[i].x = fp_mul(tab1[i].x, beta);
    tab2[i].y = flip ? fp_neg(tab1[i].y) : tab1[i].y;
    tab2[i].z = tab1[i].z;
  }
  Jac<Fp> acc = jac_inf<Fp>();
  for (int w = DRX_GLV_NIBBLES - 1; w >= 0; w--) {
    if (w != DRX_GLV_NIBBLES - 1) {
      acc = jac_dbl(acc);
      acc = jac_dbl(acc);
      acc = jac_dbl(acc);
      acc = jac_dbl(acc);
    }
    const uint32_t d1 = (s.k1.v[w >> 3] >> ((w & 7) * 4)) & 15u, d2 = (s.k2.v[w >> 3] >> ((w & 7) * 4)) & 15u;
    if (d1) acc = jac_add(acc, tab1[d1]);
    if (d2) acc = jac_add(acc, tab2[d2]);
  }
  return acc;
}
// half h (0 or 1) of k * P: k1 * P or k2 * phi(P), the sign folded into the point; the two halves add up to k * P
DRX_HD Jac<Fp> glv_half_mul(const Aff<Fp>& P, const GlvSplit& s, int h) {
  Aff<Fp> Q = P;
  if (h) Q.x = fp_mul(Q.x, glv_beta());
  if (h ? s.neg2 : s.neg1) Q = aff_neg(Q);
  return window_mul(jac_from_aff(Q), h ? s.k2 : s.k1, DRX_GLV_NIBBLES);
}

}  // namespace drx
