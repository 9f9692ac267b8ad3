// Fp2 / Fp6 / Fp12 tower over the BN256 base field, in the basis kyber v3.0.8 pairing/bn256 serialises
// (gfp2.go, gfp6.go, gfp12.go; not in /root/reference):
//   Fp2  = Fp[i]/(i^2+1)            kyber gfP2{x,y}   = x*i + y              here {c0, c1} = c0 + c1*i
//   Fp6  = Fp2[tau]/(tau^3 - xi)    kyber gfP6{x,y,z} = x*tau^2 + y*tau + z  here {c0, c1, c2} = c0 + c1*tau + c2*tau^2
//   Fp12 = Fp6[w]/(w^2 - tau)       kyber gfP12{x,y}  = x*w + y              here {c0, c1} = c0 + c1*w
// with xi = i + 3. The algorithms (Karatsuba at every level, complex squaring, Granger-Scott cyclotomic
// squaring, sparse line multiplication) are chosen for the GPU and are not kyber's; the values are canonical.
//
// Fp2 multiplication is the call granularity on the device: everything above it is built from out-of-line
// fp2_mul/fp2_sqr calls so that the pairing kernels stay a few tens of KB of SASS.
#pragma once
#include "fp.cuh"

namespace drx {

struct Fp2 {
  Fp c0, c1;
};
struct Fp6 {
  Fp2 c0, c1, c2;
};
struct Fp12 {
  Fp6 c0, c1;
};

// DRX_FP2_OUTLINE: the cheap Fp2 operations (add, sub, ...) are out-of-line too, which shrinks every tower
// function to a list of calls; operands then travel through (L1-resident) thread-local memory.
#ifndef DRX_FP2_OUTLINE
#define DRX_FP2_OUTLINE 1
#endif
#if DRX_FP2_OUTLINE
#define DRX_FP2_ATTR DRX_NI
#else
#define DRX_FP2_ATTR DRX_HD
#endif

// ------------------------------------------------------------------ Fp2
DRX_HD Fp2 fp2_zero() { return Fp2{fp_zero(), fp_zero()}; }
DRX_HD Fp2 fp2_one() { return Fp2{fp_one(), fp_zero()}; }
DRX_FP2_ATTR Fp2 fp2_add(const Fp2& a, const Fp2& b) { return Fp2{fp_add(a.c0, b.c0), fp_add(a.c1, b.c1)}; }
DRX_FP2_ATTR Fp2 fp2_sub(const Fp2& a, const Fp2& b) { return Fp2{fp_sub(a.c0, b.c0), fp_sub(a.c1, b.c1)}; }
DRX_FP2_ATTR Fp2 fp2_neg(const Fp2& a) { return Fp2{fp_neg(a.c0), fp_neg(a.c1)}; }
DRX_FP2_ATTR Fp2 fp2_dbl(const Fp2& a) { return Fp2{fp_dbl(a.c0), fp_dbl(a.c1)}; }
DRX_FP2_ATTR Fp2 fp2_conj(const Fp2& a) { return Fp2{a.c0, fp_neg(a.c1)}; }
DRX_HD bool fp2_is_zero(const Fp2& a) { return fp_is_zero(a.c0) & fp_is_zero(a.c1); }
DRX_HD bool fp2_eq(const Fp2& a, const Fp2& b) { return fp_eq(a.c0, b.c0) & fp_eq(a.c1, b.c1); }
DRX_HD Fp2 fp2_sel(bool c, const Fp2& a, const Fp2& b) { return Fp2{fp_sel(c, a.c0, b.c0), fp_sel(c, a.c1, b.c1)}; }
DRX_FP2_ATTR Fp2 fp2_mul_fp(const Fp2& a, const Fp& k) { return Fp2{fp_mul_inl(a.c0, k), fp_mul_inl(a.c1, k)}; }
// (c0 + c1 i)(3 + i) = (3c0 - c1) + (c0 + 3c1) i
DRX_FP2_ATTR Fp2 fp2_mul_xi(const Fp2& a) {
  const Fp t0 = fp_add(fp_dbl(a.c0), a.c0);
  const Fp t1 = fp_add(fp_dbl(a.c1), a.c1);
  return Fp2{fp_sub(t0, a.c1), fp_add(t1, a.c0)};
}
// (a0 + a1 i)(b0 + b1 i) as two dual-product Montgomery passes (fp.cuh mont_mul2): c0 = a0 b0 + a1 (p - b1),
// c1 = a0 b1 + a1 b0. p - b1 is left unreduced (p for b1 = 0): mont_mul2 only needs its multiplicands below p.
DRX_HD Fp2 fp2_mul_dual(const Fp2& a, const Fp2& b) {
  uint32_t nb1[8];
  nb1[0] = sub_cc(ModP::mod(0), b.c1.v[0]);
  DRX_UNROLL
  for (int k = 1; k < 7; k++) nb1[k] = subc_cc(ModP::mod(k), b.c1.v[k]);
  nb1[7] = subc(ModP::mod(7), b.c1.v[7]);
  Fp2 r;
  mont_mul2<ModP>(r.c0.v, a.c0.v, b.c0.v, a.c1.v, nb1);
  mont_mul2<ModP>(r.c1.v, a.c0.v, b.c1.v, a.c1.v, b.c0.v);
  return r;
}
// DRX_FP2_DUAL: 1 = the lane-group pairing kernels use the dual-product form (measured on a B200: the four-lane
// kernel 5.46 -> 5.39 ms at 6,000 checks, the eight-lane one unchanged), 2 = the one-thread kernels too (measured:
// much worse, 8,192 proofs verify in 86 ms instead of 58 and creation 13.4 instead of 11.1 ms: five live operand
// arrays per product spill where Karatsuba's three do not)
#ifndef DRX_FP2_DUAL
#define DRX_FP2_DUAL 1
#endif
DRX_NI Fp2 fp2_mul(const Fp2& a, const Fp2& b) {
#if DRX_FP2_DUAL >= 2
  return fp2_mul_dual(a, b);
#endif
  const Fp v0 = fp_mul_inl(a.c0, b.c0);
  const Fp v1 = fp_mul_inl(a.c1, b.c1);
  const Fp s = fp_mul_inl(fp_add(a.c0, a.c1), fp_add(b.c0, b.c1));
  return Fp2{fp_sub(v0, v1), fp_sub(fp_sub(s, v0), v1)};
}
DRX_NI Fp2 fp2_sqr(const Fp2& a) {
  const Fp t = fp_mul_inl(a.c0, a.c1);
  const Fp c0 = fp_mul_inl(fp_add(a.c0, a.c1), fp_sub(a.c0, a.c1));
  return Fp2{c0, fp_dbl(t)};
}
DRX_NI Fp2 fp2_inv(const Fp2& a) {
  const Fp d = fp_inv(fp_add(fp_sqr(a.c0), fp_sqr(a.c1)));
  return Fp2{fp_mul(a.c0, d), fp_neg(fp_mul(a.c1, d))};
}

#include "bn256_curve_consts.h"

// ------------------------------------------------------------------ Fp6
DRX_HD Fp6 fp6_zero() { return Fp6{fp2_zero(), fp2_zero(), fp2_zero()}; }
DRX_HD Fp6 fp6_one() { return Fp6{fp2_one(), fp2_zero(), fp2_zero()}; }
DRX_HD Fp6 fp6_add(const Fp6& a, const Fp6& b) { return Fp6{fp2_add(a.c0, b.c0), fp2_add(a.c1, b.c1), fp2_add(a.c2, b.c2)}; }
DRX_HD Fp6 fp6_sub(const Fp6& a, const Fp6& b) { return Fp6{fp2_sub(a.c0, b.c0), fp2_sub(a.c1, b.c1), fp2_sub(a.c2, b.c2)}; }
DRX_HD Fp6 fp6_neg(const Fp6& a) { return Fp6{fp2_neg(a.c0), fp2_neg(a.c1), fp2_neg(a.c2)}; }
DRX_HD Fp6 fp6_dbl(const Fp6& a) { return Fp6{fp2_dbl(a.c0), fp2_dbl(a.c1), fp2_dbl(a.c2)}; }
DRX_HD bool fp6_eq(const Fp6& a, const Fp6& b) { return fp2_eq(a.c0, b.c0) & fp2_eq(a.c1, b.c1) & fp2_eq(a.c2, b.c2); }
DRX_HD bool fp6_is_zero(const Fp6& a) { return fp2_is_zero(a.c0) & fp2_is_zero(a.c1) & fp2_is_zero(a.c2); }
// tau * (c0 + c1 tau + c2 tau^2) = xi c2 + c0 tau + c1 tau^2
DRX_HD Fp6 fp6_mul_tau(const Fp6& a) { return Fp6{fp2_mul_xi(a.c2), a.c0, a.c1}; }

DRX_NI void fp6_mul(Fp6& r, const Fp6& a, const Fp6& b) {
  const Fp2 v0 = fp2_mul(a.c0, b.c0);
  const Fp2 v1 = fp2_mul(a.c1, b.c1);
  const Fp2 v2 = fp2_mul(a.c2, b.c2);
  Fp2 t0 = fp2_mul(fp2_add(a.c1, a.c2), fp2_add(b.c1, b.c2));
  Fp2 t1 = fp2_mul(fp2_add(a.c0, a.c1), fp2_add(b.c0, b.c1));
  Fp2 t2 = fp2_mul(fp2_add(a.c0, a.c2), fp2_add(b.c0, b.c2));
  t0 = fp2_sub(fp2_sub(t0, v1), v2);
  t1 = fp2_sub(fp2_sub(t1, v0), v1);
  t2 = fp2_sub(fp2_sub(t2, v0), v2);
  r.c0 = fp2_add(v0, fp2_mul_xi(t0));
  r.c1 = fp2_add(t1, fp2_mul_xi(v2));
  r.c2 = fp2_add(t2, v1);
}
// Chung-Hasan SQR2: 2 multiplications + 3 squarings in Fp2
DRX_NI void fp6_sqr(Fp6& r, const Fp6& a) {
  const Fp2 s0 = fp2_sqr(a.c0);
  const Fp2 s1 = fp2_dbl(fp2_mul(a.c0, a.c1));
  const Fp2 s2 = fp2_sqr(fp2_add(fp2_sub(a.c0, a.c1), a.c2));
  const Fp2 s3 = fp2_dbl(fp2_mul(a.c1, a.c2));
  const Fp2 s4 = fp2_sqr(a.c2);
  r.c0 = fp2_add(s0, fp2_mul_xi(s3));
  r.c1 = fp2_add(s1, fp2_mul_xi(s4));
  r.c2 = fp2_sub(fp2_sub(fp2_add(fp2_add(s1, s2), s3), s0), s4);
}
DRX_NI void fp6_mul_fp2(Fp6& r, const Fp6& a, const Fp2& k) {
  r.c0 = fp2_mul(a.c0, k);
  r.c1 = fp2_mul(a.c1, k);
  r.c2 = fp2_mul(a.c2, k);
}
// a * (b0 + b1 tau): 5 multiplications
DRX_NI void fp6_mul_01(Fp6& r, const Fp6& a, const Fp2& b0, const Fp2& b1) {
  const Fp2 v0 = fp2_mul(a.c0, b0);
  const Fp2 v1 = fp2_mul(a.c1, b1);
  const Fp2 t = fp2_sub(fp2_sub(fp2_mul(fp2_add(a.c0, a.c1), fp2_add(b0, b1)), v0), v1);
  const Fp2 w0 = fp2_mul(a.c2, b1);
  const Fp2 w1 = fp2_mul(a.c2, b0);
  r.c0 = fp2_add(v0, fp2_mul_xi(w0));
  r.c1 = t;
  r.c2 = fp2_add(v1, w1);
}
DRX_NI void fp6_inv(Fp6& r, const Fp6& a) {
  const Fp2 A = fp2_sub(fp2_sqr(a.c0), fp2_mul_xi(fp2_mul(a.c1, a.c2)));
  const Fp2 B = fp2_sub(fp2_mul_xi(fp2_sqr(a.c2)), fp2_mul(a.c0, a.c1));
  const Fp2 C = fp2_sub(fp2_sqr(a.c1), fp2_mul(a.c0, a.c2));
  Fp2 F = fp2_mul_xi(fp2_add(fp2_mul(a.c2, B), fp2_mul(a.c1, C)));
  F = fp2_inv(fp2_add(F, fp2_mul(a.c0, A)));
  r.c0 = fp2_mul(A, F);
  r.c1 = fp2_mul(B, F);
  r.c2 = fp2_mul(C, F);
}

// ------------------------------------------------------------------ Fp12
DRX_HD Fp12 fp12_one() { return Fp12{fp6_one(), fp6_zero()}; }
DRX_HD bool fp12_eq(const Fp12& a, const Fp12& b) { return fp6_eq(a.c0, b.c0) & fp6_eq(a.c1, b.c1); }
DRX_HD bool fp12_is_one(const Fp12& a) { return fp6_eq(a.c0, fp6_one()) & fp6_is_zero(a.c1); }
DRX_HD void fp12_conj(Fp12& r, const Fp12& a) {
  r.c0 = a.c0;
  r.c1 = fp6_neg(a.c1);
}
DRX_NI void fp12_mul(Fp12& r, const Fp12& a, const Fp12& b) {
  Fp6 v0, v1, t;
  fp6_mul(v0, a.c0, b.c0);
  fp6_mul(v1, a.c1, b.c1);
  fp6_mul(t, fp6_add(a.c0, a.c1), fp6_add(b.c0, b.c1));
  r.c1 = fp6_sub(fp6_sub(t, v0), v1);
  r.c0 = fp6_add(v0, fp6_mul_tau(v1));
}
DRX_NI void fp12_sqr(Fp12& r, const Fp12& a) {
  Fp6 v, t;
  fp6_mul(v, a.c0, a.c1);
  fp6_mul(t, fp6_add(a.c0, a.c1), fp6_add(a.c0, fp6_mul_tau(a.c1)));
  r.c0 = fp6_sub(fp6_sub(t, v), fp6_mul_tau(v));
  r.c1 = fp6_dbl(v);
}
DRX_NI void fp12_inv(Fp12& r, const Fp12& a) {
  Fp6 t0, t1;
  fp6_sqr(t0, a.c0);
  fp6_sqr(t1, a.c1);
  t0 = fp6_sub(t0, fp6_mul_tau(t1));
  fp6_inv(t1, t0);
  fp6_mul(r.c0, a.c0, t1);
  fp6_mul(t0, a.c1, t1);
  r.c1 = fp6_neg(t0);
}
// r = a * (l0 + l1*w + l3*w^3): the shape of every Miller-loop line (w^3 = tau*w). 13 Fp2 multiplications.
DRX_NI void fp12_mul_line(Fp12& r, const Fp12& a, const Fp2& l0, const Fp2& l1, const Fp2& l3) {
  Fp6 t0, t1, t2;
  fp6_mul_fp2(t0, a.c0, l0);                          // a0 * l0
  fp6_mul_01(t1, a.c1, l1, l3);                       // a1 * (l1 + l3 tau)
  fp6_mul_01(t2, fp6_add(a.c0, a.c1), fp2_add(l0, l1), l3);
  r.c1 = fp6_sub(fp6_sub(t2, t0), t1);
  r.c0 = fp6_add(t0, fp6_mul_tau(t1));
}

// Frobenius maps: coefficient of w^k picks up xi^(k(p^e-1)/6); odd powers of p also conjugate.
DRX_NI void fp12_frob1(Fp12& r, const Fp12& a) {
  r.c0.c0 = fp2_conj(a.c0.c0);                                   // w^0
  r.c1.c0 = fp2_mul(fp2_conj(a.c1.c0), K_FROB1_1());             // w^1
  r.c0.c1 = fp2_mul(fp2_conj(a.c0.c1), K_FROB1_2());             // w^2
  r.c1.c1 = fp2_mul(fp2_conj(a.c1.c1), K_FROB1_3());             // w^3
  r.c0.c2 = fp2_mul(fp2_conj(a.c0.c2), K_FROB1_4());             // w^4
  r.c1.c2 = fp2_mul(fp2_conj(a.c1.c2), K_FROB1_5());             // w^5
}
DRX_NI void fp12_frob2(Fp12& r, const Fp12& a) {
  r.c0.c0 = a.c0.c0;
  r.c1.c0 = fp2_mul_fp(a.c1.c0, K_FROB2_1());
  r.c0.c1 = fp2_mul_fp(a.c0.c1, K_FROB2_2());
  r.c1.c1 = fp2_mul_fp(a.c1.c1, K_FROB2_3());
  r.c0.c2 = fp2_mul_fp(a.c0.c2, K_FROB2_4());
  r.c1.c2 = fp2_mul_fp(a.c1.c2, K_FROB2_5());
}
DRX_NI void fp12_frob3(Fp12& r, const Fp12& a) {
  r.c0.c0 = fp2_conj(a.c0.c0);
  r.c1.c0 = fp2_mul(fp2_conj(a.c1.c0), K_FROB3_1());
  r.c0.c1 = fp2_mul(fp2_conj(a.c0.c1), K_FROB3_2());
  r.c1.c1 = fp2_mul(fp2_conj(a.c1.c1), K_FROB3_3());
  r.c0.c2 = fp2_mul(fp2_conj(a.c0.c2), K_FROB3_4());
  r.c1.c2 = fp2_mul(fp2_conj(a.c1.c2), K_FROB3_5());
}

// Squaring in the cyclotomic subgroup (elements with a^(p^6+1) = 1 after the easy part of the final
// exponentiation), Granger-Scott 2010: three Fp4 squarings = 9 Fp2 squarings.
//   with a = g0 + g1 w + g2 w^2 + g3 w^3 + g4 w^4 + g5 w^5 the Fp4 = Fp2[s]/(s^2 - xi) pairs are
//   (g0,g3) [s = w^3], (g1,g4), (g2,g5).
DRX_HD void fp4_sqr(Fp2& r0, Fp2& r1, const Fp2& a, const Fp2& b) {  // (a + b s)^2 = (a^2 + xi b^2) + 2ab s
  const Fp2 a2 = fp2_sqr(a);
  const Fp2 b2 = fp2_sqr(b);
  r0 = fp2_add(a2, fp2_mul_xi(b2));
  r1 = fp2_sub(fp2_sub(fp2_sqr(fp2_add(a, b)), a2), b2);
}
DRX_NI void fp12_cyclo_sqr(Fp12& r, const Fp12& a) {
  const Fp2 g0 = a.c0.c0, g1 = a.c1.c0, g2 = a.c0.c1, g3 = a.c1.c1, g4 = a.c0.c2, g5 = a.c1.c2;
  Fp2 t00, t01, t10, t11, t20, t21;
  fp4_sqr(t00, t01, g0, g3);  // (g0 + g3 s)^2
  fp4_sqr(t10, t11, g1, g4);  // (g1 + g4 s)^2
  fp4_sqr(t20, t21, g2, g5);  // (g2 + g5 s)^2
  // h0 = 3 t00 - 2 g0 ; h3 = 3 t01 + 2 g3
  // h2 = 3 t10 - 2 g2 ; h5 = 3 t11 + 2 g5   (from the (g1,g4) square)
  // h4 = 3 t20 - 2 g4 ; h1 = 3 xi t21 + 2 g1 (from the (g2,g5) square)
  Fp2 h;
  h = fp2_sub(t00, g0); r.c0.c0 = fp2_add(fp2_dbl(h), t00);
  h = fp2_add(t01, g3); r.c1.c1 = fp2_add(fp2_dbl(h), t01);
  h = fp2_sub(t10, g2); r.c0.c1 = fp2_add(fp2_dbl(h), t10);
  h = fp2_add(t11, g5); r.c1.c2 = fp2_add(fp2_dbl(h), t11);
  h = fp2_sub(t20, g4); r.c0.c2 = fp2_add(fp2_dbl(h), t20);
  const Fp2 x = fp2_mul_xi(t21);
  h = fp2_add(x, g1); r.c1.c0 = fp2_add(fp2_dbl(h), x);
}

// ------------------------------------------------------------------ wire format (kyber GT/G2 MarshalBinary)
// Fp element <-> 8 big-endian words; the caller moves words to/from memory.
DRX_HD void fp_to_wire(uint32_t w[8], const Fp& a) {
  const Fp t = fp_from_mont(a);
  be_words_from_limbs(w, t.v);
}
// returns false when the encoded integer is >= p
DRX_HD bool fp_from_wire(Fp& r, const uint32_t w[8]) {
  Fp t;
  limbs_from_be_words(t.v, w);
  const bool bad = mod_geq<ModP>(t.v);
  r = fp_to_mont(t);
  return !bad;
}

}  // namespace drx
