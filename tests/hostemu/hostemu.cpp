// TEST-ONLY host build of the device headers (drynx_b200/csrc/*.cuh compiled with g++ and DRX_HOST_EMU).
// The carry-flag wrappers in fp.cuh model PTX add.cc/madc.* in C, so these entry points execute the same
// instruction sequences the sm_100a kernels run. Nothing in the drynx_b200 package or the shipped
// libdrynx_gpu.so links or loads this file; it exists so `pytest -m "not gpu"` can check the GPU algorithms
// against the oracle without a GPU.
#define DRX_HOST_EMU 1
#include "wire.cuh"
#include "ed25519.cuh"
#include <cstring>
// The lane-group pairing program (csrc/pairing_ucode_g4.inc) through the interpreter's own slot access and phase body
// (csrc/ucode_interp.cuh); the tables' CUDA storage qualifiers mean nothing to g++.
struct uint2 { uint32_t x, y; };
struct uint4 { uint32_t x, y, z, w; };
#define __device__
#define __constant__
#define UC_PROGRAM_INC "pairing_ucode_g4.inc"
namespace drx {
#include "ucode_interp.cuh"
}
#undef __device__
#undef __constant__
using namespace drx;

static void load_words(uint32_t* w, const uint8_t* b, int nbytes) { memcpy(w, b, nbytes); }

extern "C" {

int emu_fp_mul_raw(const uint32_t* a, const uint32_t* b, uint32_t* r, int which) {
  if (which == 0) mont_mul<ModP>(r, a, b); else mont_mul<ModN>(r, a, b);
  return 0;
}

// r = (a*b + c*d) / 2^256 mod p (fp.cuh mont_mul2: the dual-product Montgomery pass behind the Fp2 product)
int emu_fp_mul2_raw(const uint32_t* a, const uint32_t* b, const uint32_t* c, const uint32_t* d, uint32_t* r) {
  mont_mul2<ModP>(r, a, b, c, d);
  return 0;
}

// Fp2 product by the dual-product passes (tower.cuh fp2_mul_dual) and by Karatsuba (fp2_mul): 16 words each
int emu_fp2_mul(const uint32_t* a, const uint32_t* b, uint32_t* r_dual, uint32_t* r_kara) {
  Fp2 x, y;
  memcpy(&x, a, 64); memcpy(&y, b, 64);
  const Fp2 u = fp2_mul_dual(x, y), v = fp2_mul(x, y);
  memcpy(r_dual, &u, 64); memcpy(r_kara, &v, 64);
  return 0;
}

// which: 0 = plain inverse mod p, 1 = plain inverse mod n (division steps); 2 = Montgomery fp_inv (division steps),
// 3 = Montgomery a^(p-2) (the cross-check)
int emu_mod_inv(const uint32_t* a, uint32_t* r, int which) {
  if (which == 0) mod_inv_plain<ModP>(r, a);
  else if (which == 1) mod_inv_plain<ModN>(r, a);
  else {
    Fp x; memcpy(x.v, a, 32);
    const Fp y = which == 2 ? fp_inv(x) : fp_inv_fermat(x);
    memcpy(r, y.v, 32);
  }
  return 0;
}

// ---- edwards25519 (fe25519.cuh / ed25519.cuh). op 0: a*b, 1: a+b, 2: a-b, 3: canonical form of a, 4: 1/a,
// 5: a^((p-5)/8), 6: a^2 (dedicated squaring); raw little-endian limbs in, weakly reduced limbs out (the test reduces mod p)
int emu_fe25519(const uint32_t* a, const uint32_t* b, uint32_t* r, int op) {
  Fe x, y, z;
  memcpy(x.v, a, 32); memcpy(y.v, b, 32);
  if (op == 0) z = fe_mul(x, y);
  else if (op == 1) z = fe_add(x, y);
  else if (op == 2) z = fe_sub(x, y);
  else if (op == 3) z = fe_canon(x);
  else if (op == 4) z = fe_inv(x);
  else if (op == 6) z = fe_sqr(x);
  else z = fe_pow22523(x);
  memcpy(r, z.v, 32);
  return 0;
}
// out = k * P (method 0: 4-bit window; 1: comb table built on the fly); -1 when P does not decode
int emu_ed_mul(const uint8_t* k32, const uint8_t* pt32, uint8_t* out32, int method) {
  uint32_t kw[8], pw[8], ow[8];
  load_words(kw, k32, 32); load_words(pw, pt32, 32);
  EdExt p;
  if (!ed_from_wire(p, pw)) return -1;
  const Scalar k = ed_sc_from_wire(kw);
  EdExt r;
  if (method == 0) {
    r = ed_window_mul(p, k);
  } else {
    static thread_local EdNiels table[DRX_COMB_WINDOWS * DRX_COMB_ENTRIES];
    EdExt base = p;
    for (int w = 0; w < DRX_COMB_WINDOWS; w++) {
      EdExt acc = base;
      for (int d = 1; d <= 255; d++) {
        table[w * 255 + d - 1] = ed_to_niels_zinv(acc, fe_inv(acc.Z));
        acc = ed_add(acc, base);
      }
      for (int s = 0; s < 8; s++) base = ed_dbl(base);
    }
    r = ed_comb_mul(table, k);
  }
  ed_to_wire(ow, r); memcpy(out32, ow, 32);
  return 0;
}
// ciphertext of m under P with the scalar r, the way k_ed_encrypt computes it from comb tables of B and P
int emu_ed_scalar_i64(int64_t m, uint8_t* out32) {
  const Scalar s = ed_sc_from_i64(m);
  memcpy(out32, s.v, 32);
  return 0;
}
int emu_ed_add(const uint8_t* a32, const uint8_t* b32, uint8_t* out32) {
  uint32_t aw[8], bw[8], ow[8];
  load_words(aw, a32, 32); load_words(bw, b32, 32);
  EdExt a, b;
  if (!ed_from_wire(a, aw) || !ed_from_wire(b, bw)) return -1;
  ed_to_wire(ow, ed_add(a, b)); memcpy(out32, ow, 32);
  return 0;
}

int emu_g1_mul(const uint8_t* k32, const uint8_t* pt64, uint8_t* out64, int method) {
  uint32_t kw[8], pw[16], ow[16];
  load_words(kw, k32, 32); load_words(pw, pt64, 64);
  G1Aff p; if (!g1_from_wire(p, pw)) return -1;
  Scalar k = sc_from_wire(kw);
  G1Jac r;
  if (method == 0) {
    r = window_mul(jac_from_aff(p), k);
  } else {  // comb: build the table on the fly
    static thread_local G1Aff table[DRX_COMB_WINDOWS * DRX_COMB_ENTRIES];
    G1Jac base = jac_from_aff(p);
    for (int w = 0; w < DRX_COMB_WINDOWS; w++) {
      G1Jac acc = base;
      for (int d = 1; d <= 255; d++) {
        table[w * 255 + d - 1] = jac_to_aff(acc);
        acc = jac_add(acc, base);
      }
      for (int s = 0; s < 8; s++) base = jac_dbl(base);
    }
    r = comb_mul(table, k);
  }
  g1_to_wire(ow, jac_to_aff(r)); memcpy(out64, ow, 64);
  return 0;
}

int emu_g1_add(const uint8_t* a64, const uint8_t* b64, uint8_t* out64) {
  uint32_t aw[16], bw[16], ow[16];
  load_words(aw, a64, 64); load_words(bw, b64, 64);
  G1Aff a, b; if (!g1_from_wire(a, aw) || !g1_from_wire(b, bw)) return -1;
  G1Jac r = jac_add_aff(jac_from_aff(a), b);
  g1_to_wire(ow, jac_to_aff(r)); memcpy(out64, ow, 64);
  return 0;
}

// GLV on G1 (curve.cuh glv_split / glv_half_mul): the split (magnitudes as 8 limbs each, signs in bits 0 and 1 of the
// return value) and k * P as the sum of its two half products
int emu_glv_split(const uint8_t* k32, uint32_t* k1, uint32_t* k2) {
  uint32_t kw[8]; load_words(kw, k32, 32);
  const GlvSplit s = glv_split(sc_from_wire(kw));
  memcpy(k1, s.k1.v, 32); memcpy(k2, s.k2.v, 32);
  return (s.neg1 ? 1 : 0) | (s.neg2 ? 2 : 0);
}
int emu_g1_mul_glv(const uint8_t* k32, const uint8_t* pt64, uint8_t* out64) {
  uint32_t kw[8], pw[16], ow[16];
  load_words(kw, k32, 32); load_words(pw, pt64, 64);
  G1Aff p; if (!g1_from_wire(p, pw)) return -1;
  const GlvSplit s = glv_split(sc_from_wire(kw));
  const G1Jac r = jac_add(glv_half_mul(p, s, 0), glv_half_mul(p, s, 1));
  g1_to_wire(ow, jac_to_aff(r)); memcpy(out64, ow, 64);
  uint32_t ow2[16];
  g1_to_wire(ow2, jac_to_aff(window_mul_glv(p, sc_from_wire(kw))));  // the one-thread form must agree
  return memcmp(ow, ow2, 64) ? -2 : 0;
}

int emu_g2_mul(const uint8_t* k32, const uint8_t* pt128, uint8_t* out128) {
  uint32_t kw[8], pw[32], ow[32];
  load_words(kw, k32, 32); load_words(pw, pt128, 128);
  G2Aff p; if (!g2_from_wire(p, pw)) return -1;
  G2Jac r = window_mul(jac_from_aff(p), sc_from_wire(kw));
  g2_to_wire(ow, jac_to_aff(r)); memcpy(out128, ow, 128);
  return 0;
}

int emu_pair(const uint8_t* g1, const uint8_t* g2, uint8_t* gt384) {
  uint32_t aw[16], bw[32], ow[96];
  load_words(aw, g1, 64); load_words(bw, g2, 128);
  G1Aff p; G2Aff q;
  if (!g1_from_wire(p, aw) || !g2_from_wire(q, bw)) return -1;
  Fp12 e; pairing(e, p, q);
  gt_to_wire(ow, e); memcpy(gt384, ow, 384);
  return 0;
}

// G2 membership two ways: bit 0 = the stand-alone test (complete group law), bit 1 = the verdict the Miller loop
// produces on the side (incomplete line formulas). -1: not even on the twist.
int emu_g2_membership(const uint8_t* g1, const uint8_t* g2) {
  uint32_t aw[16], bw[32];
  load_words(aw, g1, 64); load_words(bw, g2, 128);
  G1Aff p; G2Aff q;
  if (!g1_from_wire(p, aw) || !g2_from_wire(q, bw)) return -1;
  Fp12 f;
  return (g2_in_subgroup(q) ? 1 : 0) | (miller_loop(f, p, q) ? 2 : 0);
}

// The emitted micro-program executed the way a warp does it: per phase every thread (program lane x component) of the
// item computes its result from the slots, then all of them store. Returns 0 / 1 = the G2 verdict slot is zero / not.
int emu_ucode_pair(const uint8_t* g1, const uint8_t* g2, uint8_t* gt384) {
  uint32_t aw[16], bw[32], ow[96];
  load_words(aw, g1, 64); load_words(bw, g2, 128);
  G1Aff p; G2Aff q;
  if (!g1_from_wire(p, aw) || !g2_from_wire(q, bw)) return -1;
  static thread_local unsigned char smem[UC_NSLOTS * 64];
  memset(smem, 0xA5, sizeof(smem));
  ucs_host_smem = smem;
  ucs_put_inputs(0, p, q);
  for (int ph = 0; ph < UC_NPHASES; ph++) {
    Fp r[UC_G][2];
    bool act[UC_G];
    for (int lane = 0; lane < UC_G; lane++) {
      const uint2 w = uc_prog[ph * UC_G + lane];
      act[lane] = w.x != 0xffffffffu;
      for (uint32_t h = 0; h < 2 && act[lane]; h++) r[lane][h] = ucs_phase(0, uc_ops[ph], w, h);
    }
    for (int lane = 0; lane < UC_G; lane++)
      for (uint32_t h = 0; h < 2 && act[lane]; h++) ucs_store_c(0, uc_prog[ph * UC_G + lane].x & 0xffffu, h, r[lane][h]);
  }
  const bool in_g2 = fp2_is_zero(ucs_load(0, ucs_slot_off(UC_OUT_SG)));
  const Fp12 e = ucs_get_output(0);
  gt_to_wire(ow, e); memcpy(gt384, ow, 384);
  return in_g2 ? 0 : 1;
}

// checks internal identities that have no wire-level witness: cyclotomic squaring == generic squaring on a
// pairing value, Frobenius maps compose, inverse. Returns a bitmask of failures.
int emu_tower_selfcheck(const uint8_t* gt384) {
  uint32_t w[96]; load_words(w, gt384, 384);
  Fp12 a; if (!gt_from_wire(a, w)) return -1;
  int bad = 0;
  Fp12 s1, s2; fp12_sqr(s1, a); fp12_cyclo_sqr(s2, a); if (!fp12_eq(s1, s2)) bad |= 1;
  Fp12 m; fp12_mul(m, a, a); if (!fp12_eq(s1, m)) bad |= 2;
  Fp12 f1, f2, f11; fp12_frob1(f1, a); fp12_frob1(f11, f1); fp12_frob2(f2, a); if (!fp12_eq(f11, f2)) bad |= 4;
  Fp12 f3, f21; fp12_frob3(f3, a); fp12_frob1(f21, f2); if (!fp12_eq(f3, f21)) bad |= 8;
  Fp12 iv, one; fp12_inv(iv, a); fp12_mul(one, iv, a); if (!fp12_is_one(one)) bad |= 16;
  Fp12 cj; fp12_conj(cj, a); fp12_mul(one, cj, a); if (!fp12_is_one(one)) bad |= 32;  // unitary: conj = inverse
  Fp6 x = a.c0, xi, xo; fp6_inv(xi, x); fp6_mul(xo, xi, x); if (!fp6_eq(xo, fp6_one())) bad |= 64;
  Fp6 q1, q2; fp6_sqr(q1, x); fp6_mul(q2, x, x); if (!fp6_eq(q1, q2)) bad |= 128;
  return bad;
}

int emu_gt_pow(const uint8_t* gt384, const uint8_t* k32, uint8_t* out384) {
  uint32_t w[96], kw[8], ow[96]; load_words(w, gt384, 384); load_words(kw, k32, 32);
  Fp12 g; if (!gt_from_wire(g, w)) return -1;
  static thread_local Fp12 table[DRX_COMB_WINDOWS * DRX_COMB_ENTRIES];
  Fp12 base = g;
  for (int win = 0; win < DRX_COMB_WINDOWS; win++) {
    Fp12 acc = base;
    for (int d = 1; d <= 255; d++) { table[win * 255 + d - 1] = acc; fp12_mul(acc, acc, base); }
    base = acc;  // base^256
  }
  Fp12 r; gt_comb_pow(r, table, sc_from_wire(kw));
  gt_to_wire(ow, r); memcpy(out384, ow, 384);
  return 0;
}

}  // extern "C"
