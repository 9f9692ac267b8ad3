// 256-bit Montgomery arithmetic in 8 x 32-bit limbs for sm_100a.
//
// Replaces kyber v3.0.8 pairing/bn256 `gfP` (4 x u64 Montgomery, gfp.go / gfp_amd64.s; not in /root/reference,
// see SURVEY.md §0.2) on the GPU side. One template serves both moduli on the path: the base field p
// (points, pairing tower) and the group order n (proof scalars: lib/range/range_proof.go:383-401).
//
// Layout: little-endian limbs, value = sum v[k] * 2^(32k), always fully reduced (< modulus) and, for field
// elements, in Montgomery form (x * 2^256 mod p).
//
// The multiplier is a one-limb-per-row interleaved Montgomery product whose partial sums live in two
// pair-aligned accumulators ("e": limb positions 0..7, "o": positions 1..8) so that every
// mad.lo.cc/madc.hi.cc couple is one 64-bit IMAD.WIDE.X in SASS; both moduli have their top bit set, so the
// 290-bit intermediate needs an explicit carry word (hi9). Every PTX instruction goes through a one-line
// wrapper; compiled for the host (tests only, DRX_HOST_EMU) the same wrappers model the carry flag in C, which
// lets the CPU test-suite execute the exact instruction sequence the GPU runs.
#pragma once
#include <stdint.h>

#if defined(__CUDACC__)
// nvcc builds are device-only: the shipped library has no host arithmetic at all
#define DRX_HD __device__ __forceinline__
#define DRX_D __device__ __forceinline__
#else
#define DRX_HD inline __attribute__((always_inline))
#define DRX_D inline __attribute__((always_inline))
#endif

#if defined(__CUDACC__)
#define DRX_NI static __device__ __noinline__
#else
#define DRX_NI inline
#endif

// Code-size knobs (the pairing kernels are instruction-fetch bound when everything is inlined):
//   DRX_FP_OUTLINE >= 1: fp_mul / fp_sqr are out-of-line functions taking their operands by value (registers)
//   DRX_FP_OUTLINE >= 2: fp_add / fp_sub / fp_neg too
#ifndef DRX_FP_OUTLINE
#define DRX_FP_OUTLINE 1
#endif
#if DRX_FP_OUTLINE >= 1
#define DRX_FPMUL_ATTR DRX_NI
#else
#define DRX_FPMUL_ATTR DRX_HD
#endif
#if DRX_FP_OUTLINE >= 2
#define DRX_FPADD_ATTR DRX_NI
#else
#define DRX_FPADD_ATTR DRX_HD
#endif

#if defined(__CUDA_ARCH__)
#define DRX_UNROLL _Pragma("unroll")
#else
#define DRX_UNROLL
#endif

namespace drx {

// ------------------------------------------------------------------ carry-flag primitives
#if defined(__CUDA_ARCH__)
DRX_D uint32_t add_cc(uint32_t a, uint32_t b) { uint32_t r; asm volatile("add.cc.u32 %0, %1, %2;" : "=r"(r) : "r"(a), "r"(b)); return r; }
DRX_D uint32_t addc_cc(uint32_t a, uint32_t b) { uint32_t r; asm volatile("addc.cc.u32 %0, %1, %2;" : "=r"(r) : "r"(a), "r"(b)); return r; }
DRX_D uint32_t addc(uint32_t a, uint32_t b) { uint32_t r; asm volatile("addc.u32 %0, %1, %2;" : "=r"(r) : "r"(a), "r"(b)); return r; }
DRX_D uint32_t sub_cc(uint32_t a, uint32_t b) { uint32_t r; asm volatile("sub.cc.u32 %0, %1, %2;" : "=r"(r) : "r"(a), "r"(b)); return r; }
DRX_D uint32_t subc_cc(uint32_t a, uint32_t b) { uint32_t r; asm volatile("subc.cc.u32 %0, %1, %2;" : "=r"(r) : "r"(a), "r"(b)); return r; }
DRX_D uint32_t subc(uint32_t a, uint32_t b) { uint32_t r; asm volatile("subc.u32 %0, %1, %2;" : "=r"(r) : "r"(a), "r"(b)); return r; }
DRX_D uint32_t mul_lo(uint32_t a, uint32_t b) { uint32_t r; asm volatile("mul.lo.u32 %0, %1, %2;" : "=r"(r) : "r"(a), "r"(b)); return r; }
DRX_D uint32_t mul_hi(uint32_t a, uint32_t b) { uint32_t r; asm volatile("mul.hi.u32 %0, %1, %2;" : "=r"(r) : "r"(a), "r"(b)); return r; }
DRX_D uint32_t mad_lo_cc(uint32_t a, uint32_t b, uint32_t c) { uint32_t r; asm volatile("mad.lo.cc.u32 %0, %1, %2, %3;" : "=r"(r) : "r"(a), "r"(b), "r"(c)); return r; }
DRX_D uint32_t madc_lo_cc(uint32_t a, uint32_t b, uint32_t c) { uint32_t r; asm volatile("madc.lo.cc.u32 %0, %1, %2, %3;" : "=r"(r) : "r"(a), "r"(b), "r"(c)); return r; }
DRX_D uint32_t madc_hi_cc(uint32_t a, uint32_t b, uint32_t c) { uint32_t r; asm volatile("madc.hi.cc.u32 %0, %1, %2, %3;" : "=r"(r) : "r"(a), "r"(b), "r"(c)); return r; }
DRX_D uint32_t madc_hi(uint32_t a, uint32_t b, uint32_t c) { uint32_t r; asm volatile("madc.hi.u32 %0, %1, %2, %3;" : "=r"(r) : "r"(a), "r"(b), "r"(c)); return r; }
#else
// Host model of the PTX condition-code register (tests only).
static thread_local uint32_t g_cf = 0;
inline uint32_t add_cc(uint32_t a, uint32_t b) { uint64_t t = (uint64_t)a + b; g_cf = (uint32_t)(t >> 32); return (uint32_t)t; }
inline uint32_t addc_cc(uint32_t a, uint32_t b) { uint64_t t = (uint64_t)a + b + g_cf; g_cf = (uint32_t)(t >> 32); return (uint32_t)t; }
inline uint32_t addc(uint32_t a, uint32_t b) { return a + b + g_cf; }
inline uint32_t sub_cc(uint32_t a, uint32_t b) { uint64_t t = (uint64_t)a - b; g_cf = (uint32_t)(t >> 32) & 1u; return (uint32_t)t; }
inline uint32_t subc_cc(uint32_t a, uint32_t b) { uint64_t t = (uint64_t)a - b - g_cf; g_cf = (uint32_t)(t >> 32) & 1u; return (uint32_t)t; }
inline uint32_t subc(uint32_t a, uint32_t b) { return a - b - g_cf; }
inline uint32_t mul_lo(uint32_t a, uint32_t b) { return a * b; }
inline uint32_t mul_hi(uint32_t a, uint32_t b) { return (uint32_t)(((uint64_t)a * b) >> 32); }
inline uint32_t mad_lo_cc(uint32_t a, uint32_t b, uint32_t c) { uint64_t t = (uint64_t)(uint32_t)(a * b) + c; g_cf = (uint32_t)(t >> 32); return (uint32_t)t; }
inline uint32_t madc_lo_cc(uint32_t a, uint32_t b, uint32_t c) { uint64_t t = (uint64_t)(uint32_t)(a * b) + c + g_cf; g_cf = (uint32_t)(t >> 32); return (uint32_t)t; }
inline uint32_t madc_hi_cc(uint32_t a, uint32_t b, uint32_t c) { uint64_t t = (((uint64_t)a * b) >> 32) + c + g_cf; g_cf = (uint32_t)(t >> 32); return (uint32_t)t; }
inline uint32_t madc_hi(uint32_t a, uint32_t b, uint32_t c) { return (uint32_t)(((uint64_t)a * b) >> 32) + c + g_cf; }
#endif

// ------------------------------------------------------------------ modulus parameter packs (tools/gen_consts.py)
#include "bn256_consts.h"

// A residue mod M::mod (8 limbs). For ModP it is in Montgomery form; scalars mod n are kept in plain form and
// multiplied through mont_mul(mont_mul(a,b), R2).
template <class M>
struct Res {
  uint32_t v[8];
};

template <class M>
DRX_HD void mont_mul(uint32_t r[8], const uint32_t a[8], const uint32_t b[8]) {
  uint32_t e[8], o[8], hi9, m;
  // row 0: plain products
  {
    const uint32_t bi = b[0];
    DRX_UNROLL
    for (int j = 0; j < 8; j += 2) {
      o[j] = mul_lo(a[j + 1], bi);
      o[j + 1] = mul_hi(a[j + 1], bi);
      e[j] = mul_lo(a[j], bi);
      e[j + 1] = mul_hi(a[j], bi);
    }
    m = mul_lo(e[0], M::np0());
    o[0] = mad_lo_cc(M::mod(1), m, o[0]);
    o[1] = madc_hi_cc(M::mod(1), m, o[1]);
    DRX_UNROLL
    for (int j = 2; j < 8; j += 2) {
      o[j] = madc_lo_cc(M::mod(j + 1), m, o[j]);
      o[j + 1] = madc_hi_cc(M::mod(j + 1), m, o[j + 1]);
    }
    hi9 = addc(0, 0);
    e[0] = mad_lo_cc(M::mod(0), m, e[0]);
    e[1] = madc_hi_cc(M::mod(0), m, e[1]);
    DRX_UNROLL
    for (int j = 2; j < 8; j += 2) {
      e[j] = madc_lo_cc(M::mod(j), m, e[j]);
      e[j + 1] = madc_hi_cc(M::mod(j), m, e[j + 1]);
    }
    o[7] = addc_cc(o[7], 0);
    hi9 = addc(hi9, 0);
  }
  DRX_UNROLL
  for (int i = 1; i < 8; i++) {
    const uint32_t bi = b[i];
    // shift one limb: positions 1..8 (o) become 0..7 (ne); e[2..7] become positions 1..6 (no)
    uint32_t ne[8], no[8];
    DRX_UNROLL
    for (int k = 0; k < 8; k++) ne[k] = o[k];
    DRX_UNROLL
    for (int k = 0; k < 6; k++) no[k] = e[k + 2];
    no[6] = 0;
    no[7] = hi9;
    ne[0] = add_cc(ne[0], e[1]);
    // odd limbs of a, carry-in from the position-0 add
    DRX_UNROLL
    for (int j = 0; j < 6; j += 2) {
      no[j] = madc_lo_cc(a[j + 1], bi, no[j]);
      no[j + 1] = madc_hi_cc(a[j + 1], bi, no[j + 1]);
    }
    no[6] = madc_lo_cc(a[7], bi, no[6]);
    no[7] = madc_hi(a[7], bi, no[7]);  // cannot carry out: a[7] <= mod[7] < 2^32 - 2
    // even limbs of a
    ne[0] = mad_lo_cc(a[0], bi, ne[0]);
    ne[1] = madc_hi_cc(a[0], bi, ne[1]);
    DRX_UNROLL
    for (int j = 2; j < 8; j += 2) {
      ne[j] = madc_lo_cc(a[j], bi, ne[j]);
      ne[j + 1] = madc_hi_cc(a[j], bi, ne[j + 1]);
    }
    no[7] = addc(no[7], 0);
    m = mul_lo(ne[0], M::np0());
    no[0] = mad_lo_cc(M::mod(1), m, no[0]);
    no[1] = madc_hi_cc(M::mod(1), m, no[1]);
    DRX_UNROLL
    for (int j = 2; j < 8; j += 2) {
      no[j] = madc_lo_cc(M::mod(j + 1), m, no[j]);
      no[j + 1] = madc_hi_cc(M::mod(j + 1), m, no[j + 1]);
    }
    hi9 = addc(0, 0);
    ne[0] = mad_lo_cc(M::mod(0), m, ne[0]);
    ne[1] = madc_hi_cc(M::mod(0), m, ne[1]);
    DRX_UNROLL
    for (int j = 2; j < 8; j += 2) {
      ne[j] = madc_lo_cc(M::mod(j), m, ne[j]);
      ne[j + 1] = madc_hi_cc(M::mod(j), m, ne[j + 1]);
    }
    no[7] = addc_cc(no[7], 0);
    hi9 = addc(hi9, 0);
    DRX_UNROLL
    for (int k = 0; k < 8; k++) {
      e[k] = ne[k];
      o[k] = no[k];
    }
  }
  // merge (drop the zero limb at position 0) and reduce once
  uint32_t t[8], d[8], top, bw;
  t[0] = add_cc(o[0], e[1]);
  DRX_UNROLL
  for (int k = 1; k < 7; k++) t[k] = addc_cc(o[k], e[k + 1]);
  t[7] = addc_cc(o[7], 0);
  top = addc(hi9, 0);
  d[0] = sub_cc(t[0], M::mod(0));
  DRX_UNROLL
  for (int k = 1; k < 8; k++) d[k] = subc_cc(t[k], M::mod(k));
  bw = subc(0, 0);  // 0xffffffff when t < mod
  const bool use_d = (top != 0) | (bw == 0);
  DRX_UNROLL
  for (int k = 0; k < 8; k++) r[k] = use_d ? d[k] : t[k];
}

template <class M>
DRX_HD void mod_add(uint32_t r[8], const uint32_t a[8], const uint32_t b[8]) {
  uint32_t t[8], d[8];
  t[0] = add_cc(a[0], b[0]);
  DRX_UNROLL
  for (int k = 1; k < 8; k++) t[k] = addc_cc(a[k], b[k]);
  const uint32_t top = addc(0, 0);
  d[0] = sub_cc(t[0], M::mod(0));
  DRX_UNROLL
  for (int k = 1; k < 8; k++) d[k] = subc_cc(t[k], M::mod(k));
  const uint32_t bw = subc(0, 0);
  const bool use_d = (top != 0) | (bw == 0);
  DRX_UNROLL
  for (int k = 0; k < 8; k++) r[k] = use_d ? d[k] : t[k];
}

template <class M>
DRX_HD void mod_sub(uint32_t r[8], const uint32_t a[8], const uint32_t b[8]) {
  uint32_t t[8];
  t[0] = sub_cc(a[0], b[0]);
  DRX_UNROLL
  for (int k = 1; k < 8; k++) t[k] = subc_cc(a[k], b[k]);
  const uint32_t mask = subc(0, 0);  // all ones on borrow
  r[0] = add_cc(t[0], M::mod(0) & mask);
  DRX_UNROLL
  for (int k = 1; k < 7; k++) r[k] = addc_cc(t[k], M::mod(k) & mask);
  r[7] = addc(t[7], M::mod(7) & mask);
}

template <class M>
DRX_HD bool mod_is_zero(const uint32_t a[8]) {
  uint32_t x = 0;
  DRX_UNROLL
  for (int k = 0; k < 8; k++) x |= a[k];
  return x == 0;
}

template <class M>
DRX_HD void mod_neg(uint32_t r[8], const uint32_t a[8]) {
  const uint32_t mask = mod_is_zero<M>(a) ? 0u : 0xffffffffu;
  uint32_t t[8];
  t[0] = sub_cc(M::mod(0), a[0]);
  DRX_UNROLL
  for (int k = 1; k < 7; k++) t[k] = subc_cc(M::mod(k), a[k]);
  t[7] = subc(M::mod(7), a[7]);
  DRX_UNROLL
  for (int k = 0; k < 8; k++) r[k] = t[k] & mask;
}

// true when a >= modulus (input validation of wire values)
template <class M>
DRX_HD bool mod_geq(const uint32_t a[8]) {
  uint32_t d = sub_cc(a[0], M::mod(0));
  DRX_UNROLL
  for (int k = 1; k < 8; k++) d = subc_cc(a[k], M::mod(k));
  (void)d;
  return subc(0, 0) == 0;
}

// ------------------------------------------------------------------ Fp: the base field, Montgomery form
struct alignas(16) Fp {
  uint32_t v[8];
};

DRX_HD Fp fp_zero() {
  Fp r;
  DRX_UNROLL
  for (int k = 0; k < 8; k++) r.v[k] = 0;
  return r;
}
DRX_HD Fp fp_one() {
  Fp r;
  DRX_UNROLL
  for (int k = 0; k < 8; k++) r.v[k] = ModP::r1(k);
  return r;
}
DRX_FPMUL_ATTR Fp fp_mul(Fp a, Fp b) { Fp r; mont_mul<ModP>(r.v, a.v, b.v); return r; }
// always-inlined product for the three hot Fp2 leaves, where interleaving independent products pays
#ifndef DRX_FP2_INLINE_MUL
#define DRX_FP2_INLINE_MUL 0
#endif
#if DRX_FP2_INLINE_MUL
DRX_HD Fp fp_mul_inl(const Fp& a, const Fp& b) { Fp r; mont_mul<ModP>(r.v, a.v, b.v); return r; }
#else
DRX_HD Fp fp_mul_inl(const Fp& a, const Fp& b) { return fp_mul(a, b); }
#endif
DRX_HD Fp fp_sqr(const Fp& a) { return fp_mul(a, a); }
DRX_FPADD_ATTR Fp fp_add(Fp a, Fp b) { Fp r; mod_add<ModP>(r.v, a.v, b.v); return r; }
DRX_FPADD_ATTR Fp fp_sub(Fp a, Fp b) { Fp r; mod_sub<ModP>(r.v, a.v, b.v); return r; }
DRX_FPADD_ATTR Fp fp_neg(Fp a) { Fp r; mod_neg<ModP>(r.v, a.v); return r; }
DRX_HD Fp fp_dbl(const Fp& a) { return fp_add(a, a); }
DRX_HD bool fp_is_zero(const Fp& a) { return mod_is_zero<ModP>(a.v); }
DRX_HD bool fp_eq(const Fp& a, const Fp& b) {
  uint32_t x = 0;
  DRX_UNROLL
  for (int k = 0; k < 8; k++) x |= a.v[k] ^ b.v[k];
  return x == 0;
}
DRX_HD Fp fp_sel(bool c, const Fp& a, const Fp& b) {  // c ? a : b
  Fp r;
  DRX_UNROLL
  for (int k = 0; k < 8; k++) r.v[k] = c ? a.v[k] : b.v[k];
  return r;
}
DRX_HD Fp fp_to_mont(const Fp& a) {
  Fp r2;
  DRX_UNROLL
  for (int k = 0; k < 8; k++) r2.v[k] = ModP::r2(k);
  return fp_mul(a, r2);
}
DRX_HD Fp fp_from_mont(const Fp& a) {
  Fp one = fp_zero();
  one.v[0] = 1;
  return fp_mul(a, one);
}

// a^(p-2); 4-bit fixed window over the compile-time exponent (252 squarings + 64 multiplications).
DRX_HD Fp fp_inv(const Fp& a) {
  Fp tab[16];
  tab[0] = fp_one();
  tab[1] = a;
  for (int k = 2; k < 16; k++) tab[k] = fp_mul(tab[k - 1], a);
  Fp r = fp_one();
  bool started = false;
  for (int w = 63; w >= 0; w--) {
    const uint32_t limb = ModP::pm2(w >> 3);
    const uint32_t nib = (limb >> ((w & 7) * 4)) & 15u;
    if (started) {
      r = fp_sqr(r); r = fp_sqr(r); r = fp_sqr(r); r = fp_sqr(r);
    }
    if (nib) {
      r = started ? fp_mul(r, tab[nib]) : tab[nib];
      started = true;
    }
  }
  return r;
}

// ------------------------------------------------------------------ wire format: 32-byte big-endian
DRX_HD uint32_t bswap32(uint32_t x) {
  return (x >> 24) | ((x >> 8) & 0xff00u) | ((x << 8) & 0xff0000u) | (x << 24);
}
// raw (non-Montgomery) limbs from 32 big-endian bytes held as 8 u32 words loaded in memory order
DRX_HD void limbs_from_be_words(uint32_t v[8], const uint32_t w[8]) {
  DRX_UNROLL
  for (int k = 0; k < 8; k++) v[k] = bswap32(w[7 - k]);
}
DRX_HD void be_words_from_limbs(uint32_t w[8], const uint32_t v[8]) {
  DRX_UNROLL
  for (int k = 0; k < 8; k++) w[k] = bswap32(v[7 - k]);
}

}  // namespace drx
