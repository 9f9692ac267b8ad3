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

// Variable base, 4-bit fixed window, most significant window first.
template <class F>
DRX_HD Jac<F> window_mul(const Jac<F>& base, const Scalar& k) {
  Jac<F> tab[16];
  tab[0] = jac_inf<F>();
  tab[1] = base;
  for (int i = 2; i < 16; i++) tab[i] = (i & 1) ? jac_add(tab[i - 1], base) : jac_dbl(tab[i >> 1]);
  Jac<F> acc = jac_inf<F>();
  for (int w = 63; w >= 0; w--) {
    if (w != 63) {
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

}  // namespace drx
