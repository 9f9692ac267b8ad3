// Arithmetic mod p = 2^255 - 19 in 8 x 32-bit limbs for sm_100a (the edwards25519 suite).
//
// Replaces kyber v3.0.8 group/edwards25519 fe.go (ten signed 25.5-bit limbs, ref10) on the GPU side. Values are kept
// "weakly reduced": any 256-bit integer congruent to the element, not necessarily below p; fe_canon gives the unique
// representative in [0, p) for serialisation and comparisons. A product is a plain 8 x 8 schoolbook (64 multiply-
// accumulates, paired into IMAD.WIDE.X by the same even/odd accumulator layout as fp.cuh) followed by the fold
// 2^256 = 38 (mod p): 72 MAC32 against 128 for the Montgomery product of the BN field. Every PTX instruction goes
// through the carry-flag wrappers of fp.cuh, so the CPU test build (DRX_HOST_EMU) runs the same sequences.
#pragma once
#include "fp.cuh"

namespace drx {

struct Mod25519 {  // the parameter pack mod_inv_plain<M> needs
  static DRX_HD constexpr uint32_t mod(int i) { return i == 0 ? 0xffffffedu : i == 7 ? 0x7fffffffu : 0xffffffffu; }
  static DRX_HD constexpr uint32_t np0() { return 0x286bca1bu; }  // -p^-1 mod 2^32 = 19^-1
};

struct alignas(16) Fe {
  uint32_t v[8];
};

DRX_HD Fe fe_zero() {
  Fe r;
  DRX_UNROLL
  for (int k = 0; k < 8; k++) r.v[k] = 0;
  return r;
}
DRX_HD Fe fe_one() {
  Fe r = fe_zero();
  r.v[0] = 1;
  return r;
}
DRX_HD Fe fe_const(uint32_t a0, uint32_t a1, uint32_t a2, uint32_t a3, uint32_t a4, uint32_t a5, uint32_t a6, uint32_t a7) {
  Fe r;
  r.v[0] = a0; r.v[1] = a1; r.v[2] = a2; r.v[3] = a3; r.v[4] = a4; r.v[5] = a5; r.v[6] = a6; r.v[7] = a7;
  return r;
}
DRX_HD Fe fe_d() { return fe_const(0x135978a3u, 0x75eb4dcau, 0x4141d8abu, 0x00700a4du, 0x7779e898u, 0x8cc74079u, 0x2b6ffe73u, 0x52036ceeu); }
DRX_HD Fe fe_d2() { return fe_const(0x26b2f159u, 0xebd69b94u, 0x8283b156u, 0x00e0149au, 0xeef3d130u, 0x198e80f2u, 0x56dffce7u, 0x2406d9dcu); }
DRX_HD Fe fe_sqrtm1() { return fe_const(0x4a0ea0b0u, 0xc4ee1b27u, 0xad2fe478u, 0x2f431806u, 0x3dfbd7a7u, 0x2b4d0099u, 0x4fc1df0bu, 0x2b832480u); }

// r = a + b; the carry out of 2^256 comes back as 38
DRX_HD void fe_add_raw(uint32_t r[8], const uint32_t a[8], const uint32_t b[8]) {
  uint32_t t[8];
  t[0] = add_cc(a[0], b[0]);
  DRX_UNROLL
  for (int k = 1; k < 8; k++) t[k] = addc_cc(a[k], b[k]);
  const uint32_t c = addc(0, 0);
  t[0] = add_cc(t[0], c * 38u);
  DRX_UNROLL
  for (int k = 1; k < 8; k++) t[k] = addc_cc(t[k], 0);
  const uint32_t c2 = addc(0, 0);
  t[0] += c2 * 38u;  // after a second wrap the low limb is below 38: cannot overflow
  DRX_UNROLL
  for (int k = 0; k < 8; k++) r[k] = t[k];
}
// r = a - b; a borrow means 2^256 was added: take 38 away again
DRX_HD void fe_sub_raw(uint32_t r[8], const uint32_t a[8], const uint32_t b[8]) {
  uint32_t t[8];
  t[0] = sub_cc(a[0], b[0]);
  DRX_UNROLL
  for (int k = 1; k < 8; k++) t[k] = subc_cc(a[k], b[k]);
  const uint32_t bw = subc(0, 0) & 1u;
  t[0] = sub_cc(t[0], bw * 38u);
  DRX_UNROLL
  for (int k = 1; k < 8; k++) t[k] = subc_cc(t[k], 0);
  const uint32_t bw2 = subc(0, 0) & 1u;
  t[0] -= bw2 * 38u;  // after a second wrap the low limb is at least 2^32 - 38: cannot underflow
  DRX_UNROLL
  for (int k = 0; k < 8; k++) r[k] = t[k];
}

// lo + 38 * hi of a 16-limb value (shared tail of the product and the square)
DRX_HD void fe_fold16(uint32_t r[8], const uint32_t t[16]) {
  uint32_t s[9];
  s[0] = mad_lo_cc(t[8], 38u, t[0]);
  DRX_UNROLL
  for (int k = 1; k < 8; k++) s[k] = madc_lo_cc(t[8 + k], 38u, t[k]);
  s[8] = addc(0, 0);
  s[1] = add_cc(s[1], mul_hi(t[8], 38u));
  DRX_UNROLL
  for (int k = 2; k < 8; k++) s[k] = addc_cc(s[k], mul_hi(t[7 + k], 38u));
  s[8] = addc(s[8], mul_hi(t[15], 38u));
  s[0] = add_cc(s[0], s[8] * 38u);
  DRX_UNROLL
  for (int k = 1; k < 8; k++) s[k] = addc_cc(s[k], 0);
  const uint32_t c2 = addc(0, 0);
  s[0] += c2 * 38u;
  DRX_UNROLL
  for (int k = 0; k < 8; k++) r[k] = s[k];
}

// 16-limb product, even/odd accumulators (see fp.cuh mont_mul): E collects a[even] * b[i], O collects a[odd] * b[i],
// each row one carry chain over positions i .. i+8
DRX_HD void fe_mul_raw(uint32_t r[8], const uint32_t a[8], const uint32_t b[8]) {
  uint32_t e[16], o[16];
  DRX_UNROLL
  for (int k = 0; k < 16; k++) {
    e[k] = 0;
    o[k] = 0;
  }
  DRX_UNROLL
  for (int i = 0; i < 8; i++) {
    const uint32_t bi = b[i];
    // even limbs of a land on positions i, i+1, ..., i+7; the carry starts position i+8
    e[i] = mad_lo_cc(a[0], bi, e[i]);
    e[i + 1] = madc_hi_cc(a[0], bi, e[i + 1]);
    DRX_UNROLL
    for (int j = 2; j < 8; j += 2) {
      e[i + j] = madc_lo_cc(a[j], bi, e[i + j]);
      e[i + j + 1] = madc_hi_cc(a[j], bi, e[i + j + 1]);
    }
    e[i + 8] = addc(0, 0);
    // odd limbs of a land on positions i+1 .. i+8; a carry into position i+9 is kept while it exists (i < 7)
    o[i + 1] = mad_lo_cc(a[1], bi, o[i + 1]);
    o[i + 2] = madc_hi_cc(a[1], bi, o[i + 2]);
    DRX_UNROLL
    for (int j = 3; j < 8; j += 2) {
      o[i + j] = madc_lo_cc(a[j], bi, o[i + j]);
      o[i + j + 1] = madc_hi_cc(a[j], bi, o[i + j + 1]);
    }
    if (i < 7) o[i + 9] = addc(0, 0);
  }
  // t = E + O
  uint32_t t[16];
  t[0] = e[0];
  t[1] = add_cc(e[1], o[1]);
  DRX_UNROLL
  for (int k = 2; k < 16; k++) t[k] = addc_cc(e[k], o[k]);
  fe_fold16(r, t);
}

// Square: the 28 cross products a_i a_j (i < j) once, doubled by a one-bit shift, plus the 8 squares: 36 wide
// multiply-accumulates instead of 64. Cross products with i + j even collect in E, with i + j odd in O (so every
// lo/hi couple stays pair-aligned); a row's carry lands on a limb that so far holds at most earlier rows' carries.
DRX_HD void fe_sqr_raw(uint32_t r[8], const uint32_t a[8]) {
  uint32_t e[16], o[16];
  DRX_UNROLL
  for (int k = 0; k < 16; k++) {
    e[k] = 0;
    o[k] = 0;
  }
  DRX_UNROLL
  for (int i = 0; i < 7; i++) {
    // i + j odd: j = i+1, i+3, ...
    {
      const int j0 = i + 1;
      o[i + j0] = mad_lo_cc(a[i], a[j0], o[i + j0]);
      o[i + j0 + 1] = madc_hi_cc(a[i], a[j0], o[i + j0 + 1]);
      int top = i + j0 + 2;
      DRX_UNROLL
      for (int j = j0 + 2; j < 8; j += 2) {
        o[i + j] = madc_lo_cc(a[i], a[j], o[i + j]);
        o[i + j + 1] = madc_hi_cc(a[i], a[j], o[i + j + 1]);
        top = i + j + 2;
      }
      if (top < 16) o[top] = addc(o[top], 0);
    }
    // i + j even: j = i+2, i+4, ...
    if (i + 2 < 8) {
      const int j0 = i + 2;
      e[i + j0] = mad_lo_cc(a[i], a[j0], e[i + j0]);
      e[i + j0 + 1] = madc_hi_cc(a[i], a[j0], e[i + j0 + 1]);
      int top = i + j0 + 2;
      DRX_UNROLL
      for (int j = j0 + 2; j < 8; j += 2) {
        e[i + j] = madc_lo_cc(a[i], a[j], e[i + j]);
        e[i + j + 1] = madc_hi_cc(a[i], a[j], e[i + j + 1]);
        top = i + j + 2;
      }
      if (top < 16) e[top] = addc(e[top], 0);
    }
  }
  // x = E + O (limb 0 is empty: the lowest cross product sits at position 1)
  uint32_t x[16];
  x[0] = 0;
  x[1] = o[1];
  x[2] = add_cc(e[2], o[2]);
  DRX_UNROLL
  for (int k = 3; k < 16; k++) x[k] = addc_cc(e[k], o[k]);
  // t = 2x + sum a_i^2 2^(64 i)
  uint32_t d[16], t[16];
  d[0] = 0;
  DRX_UNROLL
  for (int k = 1; k < 16; k++) d[k] = (x[k] << 1) | (x[k - 1] >> 31);
  t[0] = mad_lo_cc(a[0], a[0], d[0]);
  t[1] = madc_hi_cc(a[0], a[0], d[1]);
  DRX_UNROLL
  for (int i = 1; i < 8; i++) {
    t[2 * i] = madc_lo_cc(a[i], a[i], d[2 * i]);
    t[2 * i + 1] = madc_hi_cc(a[i], a[i], d[2 * i + 1]);
  }
  fe_fold16(r, t);
}

DRX_FPMUL_ATTR Fe fe_mul(Fe a, Fe b) { Fe r; fe_mul_raw(r.v, a.v, b.v); return r; }
#ifndef DRX_FE_SQR_DEDICATED
#define DRX_FE_SQR_DEDICATED 1
#endif
#if DRX_FE_SQR_DEDICATED
DRX_FPMUL_ATTR Fe fe_sqr(Fe a) { Fe r; fe_sqr_raw(r.v, a.v); return r; }
#else
DRX_HD Fe fe_sqr(const Fe& a) { return fe_mul(a, a); }
#endif
DRX_HD Fe fe_add(const Fe& a, const Fe& b) { Fe r; fe_add_raw(r.v, a.v, b.v); return r; }
DRX_HD Fe fe_sub(const Fe& a, const Fe& b) { Fe r; fe_sub_raw(r.v, a.v, b.v); return r; }
DRX_HD Fe fe_neg(const Fe& a) { return fe_sub(fe_zero(), a); }
DRX_HD Fe fe_dbl(const Fe& a) { return fe_add(a, a); }

// the representative in [0, p)
DRX_HD Fe fe_canon(const Fe& a) {
  uint32_t t[8], s[8];
  const uint32_t top = a.v[7] >> 31;
  t[0] = add_cc(a.v[0], top * 19u);
  DRX_UNROLL
  for (int k = 1; k < 7; k++) t[k] = addc_cc(a.v[k], 0);
  t[7] = addc(a.v[7] & 0x7fffffffu, 0);  // t < 2^255 + 19
  // t >= p  <=>  t + 19 >= 2^255
  s[0] = add_cc(t[0], 19u);
  DRX_UNROLL
  for (int k = 1; k < 7; k++) s[k] = addc_cc(t[k], 0);
  s[7] = addc(t[7], 0);
  const bool ge = (s[7] >> 31) != 0;
  Fe r;
  DRX_UNROLL
  for (int k = 0; k < 7; k++) r.v[k] = ge ? s[k] : t[k];
  r.v[7] = ge ? (s[7] & 0x7fffffffu) : t[7];
  return r;
}
DRX_HD bool fe_is_zero(const Fe& a) {
  const Fe c = fe_canon(a);
  uint32_t x = 0;
  DRX_UNROLL
  for (int k = 0; k < 8; k++) x |= c.v[k];
  return x == 0;
}
DRX_HD bool fe_eq(const Fe& a, const Fe& b) { return fe_is_zero(fe_sub(a, b)); }
DRX_HD uint32_t fe_is_negative(const Fe& a) { return fe_canon(a).v[0] & 1u; }  // ref10: low bit of the canonical form
DRX_HD Fe fe_sel(bool c, const Fe& a, const Fe& b) {
  Fe r;
  DRX_UNROLL
  for (int k = 0; k < 8; k++) r.v[k] = c ? a.v[k] : b.v[k];
  return r;
}

DRX_HD Fe fe_sqr_n(Fe a, int n) {
  for (int i = 0; i < n; i++) a = fe_sqr(a);
  return a;
}
// z^(2^252 - 3) = z^((p-5)/8), the exponent of the square-root step in point decoding (fe.go fePow22523)
DRX_NI Fe fe_pow22523(Fe z) {
  const Fe z2 = fe_sqr(z);
  const Fe z9 = fe_mul(z, fe_sqr_n(z2, 2));
  const Fe z11 = fe_mul(z2, z9);
  const Fe a5 = fe_mul(z9, fe_sqr(z11));            // 2^5 - 1
  const Fe a10 = fe_mul(fe_sqr_n(a5, 5), a5);       // 2^10 - 1
  const Fe a20 = fe_mul(fe_sqr_n(a10, 10), a10);    // 2^20 - 1
  const Fe a40 = fe_mul(fe_sqr_n(a20, 20), a20);    // 2^40 - 1
  const Fe a50 = fe_mul(fe_sqr_n(a40, 10), a10);    // 2^50 - 1
  const Fe a100 = fe_mul(fe_sqr_n(a50, 50), a50);   // 2^100 - 1
  const Fe a200 = fe_mul(fe_sqr_n(a100, 100), a100);
  const Fe a250 = fe_mul(fe_sqr_n(a200, 50), a50);  // 2^250 - 1
  return fe_mul(fe_sqr_n(a250, 2), z);
}
// 1/a by division steps on the canonical form (fp.cuh mod_inv_plain); 0 -> 0
DRX_NI Fe fe_inv(Fe a) {
  const Fe c = fe_canon(a);
  Fe r;
  mod_inv_plain<Mod25519>(r.v, c.v);
  return r;
}

// wire form: 32 bytes little-endian = the 8 limbs in memory order; bit 255 is not part of the field element
DRX_HD Fe fe_from_words(const uint32_t w[8]) {
  Fe r;
  DRX_UNROLL
  for (int k = 0; k < 8; k++) r.v[k] = w[k];
  r.v[7] &= 0x7fffffffu;
  return r;
}

}  // namespace drx
