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

// r = (a*b + c*d) / 2^256 mod M: one Montgomery pass over the sum of two products (the two halves of an Fp2
// product, c0 = a0*b0 + a1*(p - b1) and c1 = a0*b1 + a1*b0, need no Karatsuba pre-additions, no subtractions of
// reduced products and one reduction each: the same 384 multiply-accumulates as three Montgomery products, about
// 120 fewer carry-chain instructions and a shorter dependent tail).
// Requires a, c < M (b and d may be any 256-bit values, e.g. p - 0 = p). Bounds: before each row t < 3M; a row adds
// a*b_i + c*d_i + m*M < 3M(2^32 - 1), so t stays below 3M*2^32 < 2^290 inside the row -- position 8 can carry into
// position 9 in every sweep, which is why each sweep ends by collecting its carry in hi9 -- and below 3M after the
// shift. Two conditional subtractions at the end. Same even / odd accumulator layout as mont_mul: E holds positions
// 0..7, O positions 1..8, hi9 position 9.
template <class M>
DRX_HD void mont_mul2(uint32_t r[8], const uint32_t a[8], const uint32_t b[8], const uint32_t c[8], const uint32_t d[8]) {
  uint32_t e[8], o[8], hi9 = 0;
  DRX_UNROLL
  for (int k = 0; k < 8; k++) {
    e[k] = 0;
    o[k] = 0;
  }
  DRX_UNROLL
  for (int i = 0; i < 8; i++) {
    const uint32_t bi = b[i], di = d[i];
    // shift one limb: positions 1..8 (o) become 0..7 (ne); e[2..7] become positions 1..6, hi9 position 8
    uint32_t ne[8], no[8], top;
    DRX_UNROLL
    for (int k = 0; k < 8; k++) ne[k] = o[k];
    DRX_UNROLL
    for (int k = 0; k < 6; k++) no[k] = e[k + 2];
    no[6] = 0;
    no[7] = hi9;
    ne[0] = add_cc(ne[0], e[1]);
    // odd limbs of a (carry-in from the position-0 add), positions 1..8
    DRX_UNROLL
    for (int j = 0; j < 8; j += 2) {
      no[j] = madc_lo_cc(a[j + 1], bi, no[j]);
      no[j + 1] = madc_hi_cc(a[j + 1], bi, no[j + 1]);
    }
    top = addc(0, 0);
    // even limbs of a, positions 0..7, carry into position 8
    ne[0] = mad_lo_cc(a[0], bi, ne[0]);
    ne[1] = madc_hi_cc(a[0], bi, ne[1]);
    DRX_UNROLL
    for (int j = 2; j < 8; j += 2) {
      ne[j] = madc_lo_cc(a[j], bi, ne[j]);
      ne[j + 1] = madc_hi_cc(a[j], bi, ne[j + 1]);
    }
    no[7] = addc_cc(no[7], 0);
    top = addc(top, 0);
    // second product: odd limbs of c
    no[0] = mad_lo_cc(c[1], di, no[0]);
    no[1] = madc_hi_cc(c[1], di, no[1]);
    DRX_UNROLL
    for (int j = 2; j < 8; j += 2) {
      no[j] = madc_lo_cc(c[j + 1], di, no[j]);
      no[j + 1] = madc_hi_cc(c[j + 1], di, no[j + 1]);
    }
    top = addc(top, 0);
    // even limbs of c
    ne[0] = mad_lo_cc(c[0], di, ne[0]);
    ne[1] = madc_hi_cc(c[0], di, ne[1]);
    DRX_UNROLL
    for (int j = 2; j < 8; j += 2) {
      ne[j] = madc_lo_cc(c[j], di, ne[j]);
      ne[j + 1] = madc_hi_cc(c[j], di, ne[j + 1]);
    }
    no[7] = addc_cc(no[7], 0);
    top = addc(top, 0);
    // reduction row
    const uint32_t m = mul_lo(ne[0], M::np0());
    no[0] = mad_lo_cc(M::mod(1), m, no[0]);
    no[1] = madc_hi_cc(M::mod(1), m, no[1]);
    DRX_UNROLL
    for (int j = 2; j < 8; j += 2) {
      no[j] = madc_lo_cc(M::mod(j + 1), m, no[j]);
      no[j + 1] = madc_hi_cc(M::mod(j + 1), m, no[j + 1]);
    }
    top = addc(top, 0);
    ne[0] = mad_lo_cc(M::mod(0), m, ne[0]);
    ne[1] = madc_hi_cc(M::mod(0), m, ne[1]);
    DRX_UNROLL
    for (int j = 2; j < 8; j += 2) {
      ne[j] = madc_lo_cc(M::mod(j), m, ne[j]);
      ne[j + 1] = madc_hi_cc(M::mod(j), m, ne[j + 1]);
    }
    no[7] = addc_cc(no[7], 0);
    hi9 = addc(top, 0);
    DRX_UNROLL
    for (int k = 0; k < 8; k++) {
      e[k] = ne[k];
      o[k] = no[k];
    }
  }
  // merge (drop the zero limb at position 0); the value is below 3M: subtract M up to twice
  uint32_t t[8], s[8], top, bw;
  t[0] = add_cc(o[0], e[1]);
  DRX_UNROLL
  for (int k = 1; k < 7; k++) t[k] = addc_cc(o[k], e[k + 1]);
  t[7] = addc_cc(o[7], 0);
  top = addc(hi9, 0);
  DRX_UNROLL
  for (int pass = 0; pass < 2; pass++) {
    // The second subtraction is needed only when the value was at least 2M, which takes both products and m near their
    // maxima: look at the top limbs first and skip the pass when they already say "below M".
    if (pass == 1 && top == 0 && t[7] < M::mod(7)) break;
    s[0] = sub_cc(t[0], M::mod(0));
    DRX_UNROLL
    for (int k = 1; k < 8; k++) s[k] = subc_cc(t[k], M::mod(k));
    bw = subc(0, 0);  // 0xffffffff when t (low 256 bits) < mod
    const bool use_s = (top != 0) | (bw == 0);
    DRX_UNROLL
    for (int k = 0; k < 8; k++) t[k] = use_s ? s[k] : t[k];
    top = use_s ? top + bw : top;  // the borrow comes out of the ninth limb
  }
  DRX_UNROLL
  for (int k = 0; k < 8; k++) r[k] = t[k];
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

// a^(p-2); 4-bit fixed window over the compile-time exponent (252 squarings + 64 multiplications). Kept as the
// cross-check of the division-step inversion below (tests, diagnostics kernel); the product path uses fp_inv.
DRX_HD Fp fp_inv_fermat(const Fp& a) {
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

// ------------------------------------------------------------------ modular inverse by batched division steps
// Bernstein-Yang "safegcd" with the half-delta rule: 20 rounds, each deriving a 2x2 transition matrix from 30
// division steps on the low 30 bits of (f, g) and applying it to the full-width (f, g) and to the Bezout pair
// (d, e) mod M. 600 steps cover every 256-bit odd modulus (bound 590). Values are 9 signed limbs of 30 bits.
// About 13 k instructions against 105 k for the exponentiation a^(M-2): one thread inverts in ~1/8 of the time,
// which is what the small-batch kernels (one inversion on every item's critical path) are bound by.
// Plain (non-Montgomery) in and out: r = a^-1 mod M for 0 < a < M, and 0 for a = 0 (as a^(M-2) gives).
template <class M>
DRX_HD void s30_from_limbs(int32_t r[9], const uint32_t a[8]) {
  DRX_UNROLL
  for (int k = 0; k < 9; k++) {
    const int bit = 30 * k, w = bit >> 5, sh = bit & 31;
    const uint64_t lo = a[w], hi = (w + 1 < 8) ? a[w + 1] : 0u;
    r[k] = (int32_t)((uint32_t)(((hi << 32) | lo) >> sh) & 0x3fffffffu);
  }
}
template <class M>
DRX_HD void mod_inv_plain(uint32_t out[8], const uint32_t a[8]) {
  const int32_t M30 = 0x3fffffff;
  int32_t mod[9];
  {
    uint32_t mm[8];
    DRX_UNROLL
    for (int k = 0; k < 8; k++) mm[k] = M::mod(k);
    s30_from_limbs<M>(mod, mm);
  }
  const uint32_t inv30 = (0u - M::np0()) & 0x3fffffffu;  // M^-1 mod 2^30
  int32_t d[9], e[9], f[9], g[9];
  DRX_UNROLL
  for (int k = 0; k < 9; k++) {
    d[k] = 0;
    e[k] = k == 0 ? 1 : 0;
    f[k] = mod[k];
  }
  s30_from_limbs<M>(g, a);
  int32_t zeta = -1;  // -(delta + 1/2)
#if defined(__CUDA_ARCH__)
#pragma unroll 1
#endif
  for (int round = 0; round < 20; round++) {
    // ---- 30 division steps on the low limbs; [u v; q r] * (f, g) = 2^30 * (f', g')
    uint32_t u = 1, v = 0, q = 0, r = 1;
    uint32_t fl = (uint32_t)f[0], gl = (uint32_t)g[0];
#if defined(__CUDA_ARCH__)
#pragma unroll 6
#endif
    for (int i = 0; i < 30; i++) {
      uint32_t c1 = (uint32_t)(zeta >> 31);  // zeta < 0
      const uint32_t c2 = 0u - (gl & 1u);    // g odd
      const uint32_t x = (fl ^ c1) - c1, y = (u ^ c1) - c1, z = (v ^ c1) - c1;
      gl += x & c2;
      q += y & c2;
      r += z & c2;
      c1 &= c2;
      zeta = (int32_t)((uint32_t)zeta ^ c1) - 1;
      fl += gl & c1;
      u += q & c1;
      v += r & c1;
      gl >>= 1;
      u <<= 1;
      v <<= 1;
    }
    const int64_t U = (int32_t)u, V = (int32_t)v, Q = (int32_t)q, R = (int32_t)r;
    // ---- (d, e) <- [u v; q r] * (d, e) / 2^30 mod M, staying inside (-2M, M)
    {
      const int32_t sd = d[8] >> 31, se = e[8] >> 31;
      int32_t md = ((int32_t)u & sd) + ((int32_t)v & se);
      int32_t me = ((int32_t)q & sd) + ((int32_t)r & se);
      int64_t cd = U * d[0] + V * e[0];
      int64_t ce = Q * d[0] + R * e[0];
      md -= (int32_t)((inv30 * (uint32_t)cd + (uint32_t)md) & 0x3fffffffu);
      me -= (int32_t)((inv30 * (uint32_t)ce + (uint32_t)me) & 0x3fffffffu);
      cd += (int64_t)mod[0] * md;
      ce += (int64_t)mod[0] * me;
      cd >>= 30;
      ce >>= 30;
      DRX_UNROLL
      for (int k = 1; k < 9; k++) {
        cd += U * d[k] + V * e[k] + (int64_t)mod[k] * md;
        ce += Q * d[k] + R * e[k] + (int64_t)mod[k] * me;
        d[k - 1] = (int32_t)cd & M30;
        e[k - 1] = (int32_t)ce & M30;
        cd >>= 30;
        ce >>= 30;
      }
      d[8] = (int32_t)cd;
      e[8] = (int32_t)ce;
    }
    // ---- (f, g) <- [u v; q r] * (f, g) / 2^30 (exact)
    {
      int64_t cf = U * f[0] + V * g[0];
      int64_t cg = Q * f[0] + R * g[0];
      cf >>= 30;
      cg >>= 30;
      DRX_UNROLL
      for (int k = 1; k < 9; k++) {
        cf += U * f[k] + V * g[k];
        cg += Q * f[k] + R * g[k];
        f[k - 1] = (int32_t)cf & M30;
        g[k - 1] = (int32_t)cg & M30;
        cf >>= 30;
        cg >>= 30;
      }
      f[8] = (int32_t)cf;
      g[8] = (int32_t)cg;
    }
  }
  // ---- g = 0 and f = +-1 now: the inverse is sign(f) * d, brought into [0, M)
  {
    int32_t add = d[8] >> 31;
    const int32_t neg = f[8] >> 31;
    DRX_UNROLL
    for (int k = 0; k < 9; k++) d[k] = ((d[k] + (mod[k] & add)) ^ neg) - neg;
    DRX_UNROLL
    for (int k = 0; k < 8; k++) {
      d[k + 1] += d[k] >> 30;
      d[k] &= M30;
    }
    add = d[8] >> 31;
    DRX_UNROLL
    for (int k = 0; k < 9; k++) d[k] += mod[k] & add;
    DRX_UNROLL
    for (int k = 0; k < 8; k++) {
      d[k + 1] += d[k] >> 30;
      d[k] &= M30;
    }
  }
  DRX_UNROLL
  for (int j = 0; j < 8; j++) {
    const int bit = 32 * j, k = bit / 30, sh = bit - 30 * k;
    uint64_t x = (uint64_t)(uint32_t)d[k] >> sh;
    x |= (uint64_t)(uint32_t)d[k + 1] << (30 - sh);
    if (k + 2 < 9) x |= (uint64_t)(uint32_t)d[k + 2] << (60 - sh);
    out[j] = (uint32_t)x;
  }
}

// Montgomery-form inverse: (aR)^-1 = a^-1 R^-1 plain, times R^3 through one Montgomery product gives a^-1 R.
DRX_NI Fp fp_inv(Fp a) {
  Fp r2, t;
  DRX_UNROLL
  for (int k = 0; k < 8; k++) r2.v[k] = ModP::r2(k);
  mod_inv_plain<ModP>(t.v, a.v);
  const Fp r3 = fp_mul(r2, r2);
  return fp_mul(t, r3);
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
