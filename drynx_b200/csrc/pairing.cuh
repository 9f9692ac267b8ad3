// Optimal ate pairing e: G1 x G2 -> GT on BN256, plus GT exponentiation.
// Replaces kyber v3.0.8 pairing/bn256 optate.go (miller, finalExponentiation, optimalAte) and the GT
// "Add"/"Mul" used by lib/range/range_proof.go:396-397 and :540-546.
//   e(P, Q) = ( f_{6u+2,Q}(P) * l_{[6u+2]Q, pi(Q)}(P) * l_{[6u+2]Q + pi(Q), -pi^2(Q)}(P) ) ^ ((p^12 - 1)/n)
// The reduced value is canonical, so this file is free to differ from kyber's algorithm: lines are
// multiplied in as sparse elements (13 Fp2 products), the hard part of the final exponentiation runs on
// Granger-Scott cyclotomic squarings, and inverses in the cyclotomic subgroup are conjugations.
#pragma once
#include "curve.cuh"

namespace drx {

struct TwistJ {  // Jacobian point on the twist with cached t = z^2
  Fp2 x, y, z, t;
};

// Tangent at r evaluated at the G1 point q; r <- 2r. Line = c + b*w + a*w^3.
DRX_NI void line_dbl(Fp2& a, Fp2& b, Fp2& c, TwistJ& r, const G1Aff& q) {
  const Fp2 A = fp2_sqr(r.x);
  const Fp2 B = fp2_sqr(r.y);
  const Fp2 C = fp2_sqr(B);
  Fp2 D = fp2_sub(fp2_sub(fp2_sqr(fp2_add(r.x, B)), A), C);
  D = fp2_dbl(D);
  const Fp2 E = fp2_add(fp2_dbl(A), A);
  const Fp2 G = fp2_sqr(E);
  TwistJ o;
  o.x = fp2_sub(fp2_sub(G, D), D);
  o.z = fp2_sub(fp2_sub(fp2_sqr(fp2_add(r.y, r.z)), B), r.t);
  Fp2 t = fp2_dbl(C);
  t = fp2_dbl(t);
  t = fp2_dbl(t);
  o.y = fp2_sub(fp2_mul(fp2_sub(D, o.x), E), t);
  o.t = fp2_sqr(o.z);
  t = fp2_dbl(fp2_mul(E, r.t));
  b = fp2_mul_fp(fp2_neg(t), q.x);
  a = fp2_sub(fp2_sub(fp2_sqr(fp2_add(r.x, E)), A), G);
  t = fp2_dbl(B);
  t = fp2_dbl(t);
  a = fp2_sub(a, t);
  c = fp2_mul_fp(fp2_dbl(fp2_mul(o.z, r.t)), q.y);
  r = o;
}

// Chord through r and the affine twist point p (with p.y^2 = r2) evaluated at q; r <- r + p.
DRX_NI void line_add(Fp2& a, Fp2& b, Fp2& c, TwistJ& r, const G2Aff& p, const G1Aff& q, const Fp2& r2) {
  const Fp2 B = fp2_mul(p.x, r.t);
  Fp2 D = fp2_sub(fp2_sub(fp2_sqr(fp2_add(p.y, r.z)), r2), r.t);
  D = fp2_mul(D, r.t);
  const Fp2 H = fp2_sub(B, r.x);
  const Fp2 I = fp2_sqr(H);
  Fp2 E = fp2_dbl(I);
  E = fp2_dbl(E);
  const Fp2 J = fp2_mul(H, E);
  const Fp2 L1 = fp2_sub(fp2_sub(D, r.y), r.y);
  const Fp2 V = fp2_mul(r.x, E);
  TwistJ o;
  o.x = fp2_sub(fp2_sub(fp2_sub(fp2_sqr(L1), J), V), V);
  o.z = fp2_sub(fp2_sub(fp2_sqr(fp2_add(r.z, H)), r.t), I);
  Fp2 t = fp2_mul(fp2_sub(V, o.x), L1);
  Fp2 t2 = fp2_dbl(fp2_mul(r.y, J));
  o.y = fp2_sub(t, t2);
  o.t = fp2_sqr(o.z);
  t = fp2_sub(fp2_sub(fp2_sqr(fp2_add(p.y, o.z)), r2), o.t);
  t2 = fp2_dbl(fp2_mul(L1, p.x));
  a = fp2_sub(t2, t);
  c = fp2_dbl(fp2_mul_fp(o.z, q.y));
  b = fp2_dbl(fp2_mul_fp(fp2_neg(L1), q.x));
  r = o;
}

// Miller function value for affine, finite P in G1 and Q on the twist. Returns whether Q lies in G2 (the order-n
// subgroup), which the loop decides on the side: after the last two line steps r = [6u+2]Q + pi(Q) - pi^2(Q), and
// r = -pi^3(Q) exactly when Q has order n. (6u+2 + p - p^2 + p^3 = 0 mod n is the relation the optimal ate pairing
// rests on, so the equality holds on G2; conversely the resultant of X^3 - X^2 + X + 6u+2 with the characteristic
// polynomial of pi shares only the factor n with #E'(Fp2) = n (2p - n) on this curve -- checked with integers in
// tests/test_oracle.py -- so any twist point passing the test has order n.) The line formulas are incomplete, but
// only a point outside G2 can reach one of their exceptional cases, and those leave z = 0 for good, which fails
// the test as well. kyber's UnmarshalBinary rejects such points ([n]Q != O in twistPoint.IsOnCurve), so
// lib/range/range_proof.go:225-236 never hands one to the pairing.
DRX_NI bool miller_loop(Fp12& f, const G1Aff& P, const G2Aff& Q) {
  f = fp12_one();
  TwistJ r;
  r.x = Q.x;
  r.y = Q.y;
  r.z = fp2_one();
  r.t = fp2_one();
  const G2Aff mQ = aff_neg(Q);
  Fp2 r2 = fp2_sqr(Q.y);
  Fp2 a, b, c;
  for (int i = DRX_ATE_NAF_LEN - 1; i > 0; i--) {
    line_dbl(a, b, c, r, P);
    if (i != DRX_ATE_NAF_LEN - 1) fp12_sqr(f, f);
    fp12_mul_line(f, f, c, b, a);
    const int d = ate_naf(i - 1);
    if (d == 1) {
      line_add(a, b, c, r, Q, P, r2);
      fp12_mul_line(f, f, c, b, a);
    } else if (d == -1) {
      line_add(a, b, c, r, mQ, P, r2);
      fp12_mul_line(f, f, c, b, a);
    }
  }
  // pi(Q) and -pi^2(Q) on the twist
  G2Aff q1, mq2;
  q1.x = fp2_mul(fp2_conj(Q.x), K_XI_P_1_3());
  q1.y = fp2_mul(fp2_conj(Q.y), K_XI_P_1_2());
  mq2.x = fp2_mul_fp(Q.x, K_XI_P2_1_3());
  mq2.y = Q.y;
  r2 = fp2_sqr(q1.y);
  line_add(a, b, c, r, q1, P, r2);
  fp12_mul_line(f, f, c, b, a);
  r2 = fp2_sqr(mq2.y);
  line_add(a, b, c, r, mq2, P, r2);
  fp12_mul_line(f, f, c, b, a);
  // r == -pi^3(Q), compared projectively (r.t = r.z^2)
  const Fp2 x3 = fp2_mul(fp2_mul(fp2_conj(Q.x), K_FROB3_2()), r.t);
  const Fp2 y3 = fp2_mul(fp2_mul(fp2_conj(Q.y), K_FROB3_3()), fp2_mul(r.t, r.z));
  return fp2_eq(r.x, x3) & fp2_eq(r.y, fp2_neg(y3)) & !fp2_is_zero(r.z);
}

// Membership in G2 alone (set-up and diagnostic paths; verification gets it from its Miller loop): the same relation
// [6u+2]Q + pi(Q) - pi^2(Q) + pi^3(Q) = O, evaluated with the complete group law.
DRX_NI bool g2_in_subgroup(const G2Aff& Q) {
  if (aff_is_inf(Q)) return true;
  const G2Aff mQ = aff_neg(Q);
  G2Jac r = jac_from_aff(Q);
  for (int i = DRX_ATE_NAF_LEN - 1; i > 0; i--) {
    r = jac_dbl(r);
    const int d = ate_naf(i - 1);
    if (d == 1) r = jac_add_aff(r, Q);
    else if (d == -1) r = jac_add_aff(r, mQ);
  }
  G2Aff q1, mq2, q3;
  q1.x = fp2_mul(fp2_conj(Q.x), K_XI_P_1_3());
  q1.y = fp2_mul(fp2_conj(Q.y), K_XI_P_1_2());
  mq2.x = fp2_mul_fp(Q.x, K_XI_P2_1_3());
  mq2.y = Q.y;
  q3.x = fp2_mul(fp2_conj(Q.x), K_FROB3_2());
  q3.y = fp2_mul(fp2_conj(Q.y), K_FROB3_3());
  r = jac_add_aff(jac_add_aff(jac_add_aff(r, q1), mq2), q3);
  return jac_is_inf(r);
}

// a^u for a in the cyclotomic subgroup (u = 6518589491078791937): signed-digit (NAF, weight 23 instead of 30)
// square-and-multiply; a^-1 is the conjugate there, so negative digits cost the same as positive ones.
DRX_NI void fp12_pow_u(Fp12& r, const Fp12& a) {
  Fp12 acc = a, inv;
  fp12_conj(inv, a);
  for (int i = DRX_U_NAF_LEN - 2; i >= 0; i--) {  // the top digit is +1
    fp12_cyclo_sqr(acc, acc);
    const int d = u_naf(i);
    if (d > 0) fp12_mul(acc, acc, a);
    else if (d < 0) fp12_mul(acc, acc, inv);
  }
  r = acc;
}

// f^((p^12-1)/n): easy part (p^6-1)(p^2+1), then the Scott et al. addition chain for (p^4-p^2+1)/n
//   result = y0 * y1^2 * y2^6 * y3^12 * y4^18 * y5^30 * y6^36 with
//   y0 = a^p a^(p^2) a^(p^3), y1 = 1/a, y2 = (a^(u^2))^(p^2), y3 = 1/(a^u)^p, y4 = 1/(a^u (a^(u^2))^p),
//   y5 = 1/a^(u^2), y6 = 1/(a^(u^3) (a^(u^3))^p).  Ordered so that at most eight Fp12 values are live (the
//   stack of this function is the largest contributor to the kernels' thread-local footprint).
DRX_NI void final_exp(Fp12& out, const Fp12& in) {
  Fp12 a, fu, fu2, t0, t1, x;
  fp12_conj(a, in);
  fp12_inv(x, in);
  fp12_mul(a, a, x);      // in^(p^6-1)
  fp12_frob2(x, a);
  fp12_mul(a, a, x);      // ^(p^2+1): cyclotomic from here on
  fp12_pow_u(fu, a);
  fp12_pow_u(fu2, fu);
  fp12_pow_u(x, fu2);     // fu3
  fp12_frob1(t0, x);
  fp12_mul(t0, t0, x);
  fp12_conj(t0, t0);      // y6
  fp12_cyclo_sqr(t0, t0);
  fp12_frob1(x, fu2);
  fp12_mul(x, x, fu);
  fp12_conj(x, x);        // y4
  fp12_mul(t0, t0, x);
  fp12_conj(x, fu2);      // y5
  fp12_mul(t0, t0, x);    // t0 = y6^2 y4 y5
  fp12_frob1(t1, fu);
  fp12_conj(t1, t1);      // y3
  fp12_mul(t1, t1, x);
  fp12_mul(t1, t1, t0);   // t1 = y3 y5 t0
  fp12_frob2(x, fu2);     // y2
  fp12_mul(t0, t0, x);
  fp12_cyclo_sqr(t1, t1);
  fp12_mul(t1, t1, t0);
  fp12_cyclo_sqr(t1, t1);
  fp12_conj(x, a);        // y1
  fp12_mul(t0, t1, x);
  fp12_frob1(x, a);       // y0 = a^p a^(p^2) a^(p^3)
  fp12_frob2(fu, a);
  fp12_mul(x, x, fu);
  fp12_frob1(fu, fu);
  fp12_mul(x, x, fu);
  fp12_mul(t1, t1, x);
  fp12_cyclo_sqr(t0, t0);
  fp12_mul(out, t0, t1);
}

// e(P, Q); either argument at infinity gives 1 (kyber optimalAte). Returns false, and leaves out = 1, when Q is a
// twist point outside G2 (no such point can come out of kyber's UnmarshalBinary, see miller_loop).
DRX_NI bool pairing(Fp12& out, const G1Aff& P, const G2Aff& Q) {
  if (aff_is_inf(Q)) {
    out = fp12_one();
    return true;
  }
  if (aff_is_inf(P)) {
    out = fp12_one();
    return g2_in_subgroup(Q);
  }
  Fp12 f;
  if (!miller_loop(f, P, Q)) {
    out = fp12_one();
    return false;
  }
  final_exp(out, f);
  return true;
}

// Fixed-base GT exponentiation: table[w * 255 + d - 1] = g^(d * 2^(8w)).
DRX_NI void gt_comb_pow(Fp12& out, const Fp12* __restrict__ table, const Scalar& k) {
  Fp12 acc = fp12_one();
  bool any = false;
  for (int w = 0; w < DRX_COMB_WINDOWS; w++) {
    const uint32_t d = (k.v[w >> 2] >> ((w & 3) * 8)) & 0xffu;
    if (d) {
      const Fp12* e = &table[w * DRX_COMB_ENTRIES + (int)d - 1];
      if (any) {
        fp12_mul(acc, acc, *e);
      } else {
        acc = *e;
        any = true;
      }
    }
  }
  out = acc;
}

// The same over 16-bit windows: table16[w * 65535 + d - 1] = g^(d * 2^(16w)), w < 16 (the one base every proof shares,
// e(B, g2): kernels_rp.cu k_comb_table_gt16)
#define DRX_GT16_PER_WINDOW 65535
DRX_NI void gt_comb_pow16(Fp12& out, const Fp12* __restrict__ table16, const Scalar& k) {
  Fp12 acc = fp12_one();
  bool any = false;
  for (int w = 0; w < 16; w++) {
    const uint32_t d = (k.v[w >> 1] >> ((w & 1) * 16)) & 0xffffu;
    if (d) {
      const Fp12* e = &table16[(unsigned long long)w * DRX_GT16_PER_WINDOW + d - 1];
      if (any) {
        fp12_mul(acc, acc, *e);
      } else {
        acc = *e;
        any = true;
      }
    }
  }
  out = acc;
}

// ------------------------------------------------------------------ wire helpers for G2 / GT
// kyber GT MarshalBinary order x.x x.y x.z y.x y.y y.z = w^5 w^3 w^1 w^4 w^2 w^0, each Fp2 as (c1, c0).
DRX_HD const Fp2& fp12_wire_slot(const Fp12& a, int s) {
  switch (s) {
    case 0: return a.c1.c2;
    case 1: return a.c1.c1;
    case 2: return a.c1.c0;
    case 3: return a.c0.c2;
    case 4: return a.c0.c1;
    default: return a.c0.c0;
  }
}
DRX_HD Fp2& fp12_wire_slot(Fp12& a, int s) {
  switch (s) {
    case 0: return a.c1.c2;
    case 1: return a.c1.c1;
    case 2: return a.c1.c0;
    case 3: return a.c0.c2;
    case 4: return a.c0.c1;
    default: return a.c0.c0;
  }
}

}  // namespace drx
