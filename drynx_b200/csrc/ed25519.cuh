// The edwards25519 group for the ElGamal path: -x^2 + y^2 = 1 + d x^2 y^2 over 2^255 - 19.
//
// Replaces kyber v3.0.8 group/edwards25519 ge.go / point.go / scalar.go (not under /root/reference, go.mod:14) on
// the GPU side. Points live in extended coordinates (X : Y : Z : T), T = XY/Z; the addition law is complete, so
// there are no exceptional cases anywhere (doubling, inverses and the identity go through the same formulas).
// Fixed-base work walks a comb table of affine "Niels" triples (y+x, y-x, 2dxy), 32 windows of 8 bits like the
// BN256 path. Wire form: 32 bytes, y little-endian, bit 255 = parity of x (RFC 8032 5.1.2); decoding follows
// ref10's ge_frombytes (y >= p and "negative zero" accepted, off-curve rejected).
#pragma once
#include "curve.cuh"
#include "fe25519.cuh"

namespace drx {

struct EdExt { Fe X, Y, Z, T; };
struct alignas(16) EdNiels { Fe yPlusX, yMinusX, xy2d; };  // affine (Z = 1) table entry, 96 bytes

DRX_HD EdExt ed_identity() { return EdExt{fe_zero(), fe_one(), fe_one(), fe_zero()}; }
DRX_HD EdExt ed_neg(const EdExt& p) { return EdExt{fe_neg(p.X), p.Y, p.Z, fe_neg(p.T)}; }
DRX_HD EdNiels ed_niels_neg(const EdNiels& q) { return EdNiels{q.yMinusX, q.yPlusX, fe_neg(q.xy2d)}; }

// p + q, q an affine table entry: 7 products
DRX_NI EdExt ed_add_niels(const EdExt& p, const EdNiels& q) {
  const Fe a = fe_mul(fe_sub(p.Y, p.X), q.yMinusX);
  const Fe b = fe_mul(fe_add(p.Y, p.X), q.yPlusX);
  const Fe c = fe_mul(p.T, q.xy2d);
  const Fe d = fe_dbl(p.Z);
  const Fe e = fe_sub(b, a), f = fe_sub(d, c), g = fe_add(d, c), h = fe_add(b, a);
  return EdExt{fe_mul(e, f), fe_mul(g, h), fe_mul(f, g), fe_mul(e, h)};
}
// p + q, both extended: 9 products
DRX_NI EdExt ed_add(const EdExt& p, const EdExt& q) {
  const Fe a = fe_mul(fe_sub(p.Y, p.X), fe_sub(q.Y, q.X));
  const Fe b = fe_mul(fe_add(p.Y, p.X), fe_add(q.Y, q.X));
  const Fe c = fe_mul(fe_mul(p.T, q.T), fe_d2());
  const Fe d = fe_dbl(fe_mul(p.Z, q.Z));
  const Fe e = fe_sub(b, a), f = fe_sub(d, c), g = fe_add(d, c), h = fe_add(b, a);
  return EdExt{fe_mul(e, f), fe_mul(g, h), fe_mul(f, g), fe_mul(e, h)};
}
// 2p: 4 squarings + 4 products (T of the input is not used)
DRX_NI EdExt ed_dbl(const EdExt& p) {
  const Fe xx = fe_sqr(p.X), yy = fe_sqr(p.Y), zz2 = fe_dbl(fe_sqr(p.Z));
  const Fe aa = fe_sqr(fe_add(p.X, p.Y));
  const Fe h = fe_add(yy, xx), g = fe_sub(yy, xx);
  const Fe e = fe_sub(aa, h), f = fe_sub(zz2, g);
  return EdExt{fe_mul(e, f), fe_mul(h, g), fe_mul(g, f), fe_mul(e, h)};
}

// ------------------------------------------------------------------ scalars mod l = 2^252 + 277423...493
DRX_HD constexpr uint32_t ed_l(int i) {
  return i == 0 ? 0x5cf5d3edu : i == 1 ? 0x5812631au : i == 2 ? 0xa2f79cd6u : i == 3 ? 0x14def9deu : i == 7 ? 0x10000000u : 0u;
}
// (l << k) limb i, k in 0..3
DRX_HD constexpr uint32_t ed_l_shl(int i, int k) {
  return (ed_l(i) << k) | ((k && i) ? (ed_l(i - 1) >> (32 - k)) : 0u);
}
template <int K>
DRX_HD void ed_sc_cond_sub(uint32_t v[8]) {
  uint32_t t[8];
  t[0] = sub_cc(v[0], ed_l_shl(0, K));
  DRX_UNROLL
  for (int i = 1; i < 8; i++) t[i] = subc_cc(v[i], ed_l_shl(i, K));
  const bool keep = subc(0, 0) != 0;  // borrow: v < l << K
  DRX_UNROLL
  for (int i = 0; i < 8; i++) v[i] = keep ? v[i] : t[i];
}
// any 256-bit value -> [0, l) (2^256 < 16 l)
DRX_HD Scalar ed_sc_reduce(Scalar s) {
  ed_sc_cond_sub<3>(s.v);
  ed_sc_cond_sub<2>(s.v);
  ed_sc_cond_sub<1>(s.v);
  ed_sc_cond_sub<0>(s.v);
  return s;
}
// kyber scalar.SetBytes: 32 bytes little-endian, reduced
DRX_HD Scalar ed_sc_from_wire(const uint32_t w[8]) {
  Scalar s;
  DRX_UNROLL
  for (int k = 0; k < 8; k++) s.v[k] = w[k];
  return ed_sc_reduce(s);
}
// kyber scalar.SetInt64: negative values wrap mod l
DRX_HD Scalar ed_sc_from_i64(int64_t m) {
  const uint64_t am = m < 0 ? (uint64_t)0 - (uint64_t)m : (uint64_t)m;
  Scalar s = sc_from_u64(am);
  if (m < 0) {
    Scalar r;
    r.v[0] = sub_cc(ed_l(0), s.v[0]);
    DRX_UNROLL
    for (int i = 1; i < 7; i++) r.v[i] = subc_cc(ed_l(i), s.v[i]);
    r.v[7] = subc(ed_l(7), s.v[7]);
    return r;
  }
  return s;
}

// ------------------------------------------------------------------ scalar multiplication
// fixed base: table[w * 255 + d - 1] = d * 256^w * B
DRX_HD EdExt ed_comb_mul(const EdNiels* __restrict__ table, const Scalar& k, int windows = DRX_COMB_WINDOWS) {
  EdExt acc = ed_identity();
  for (int w = 0; w < windows; w++) {
    const uint32_t d = (k.v[w >> 2] >> ((w & 3) * 8)) & 0xffu;
    if (d) acc = ed_add_niels(acc, table[w * DRX_COMB_ENTRIES + (int)d - 1]);
  }
  return acc;
}
// variable base, 4-bit fixed window, most significant window first (scalars are below 2^253: 64 windows)
DRX_HD EdExt ed_window_mul(const EdExt& base, const Scalar& k) {
  EdExt tab[16];
  tab[0] = ed_identity();
  tab[1] = base;
  for (int i = 2; i < 16; i++) tab[i] = (i & 1) ? ed_add(tab[i - 1], base) : ed_dbl(tab[i >> 1]);
  EdExt acc = ed_identity();
  for (int w = 63; w >= 0; w--) {
    if (w != 63) {
      acc = ed_dbl(acc);
      acc = ed_dbl(acc);
      acc = ed_dbl(acc);
      acc = ed_dbl(acc);
    }
    const uint32_t d = (k.v[w >> 3] >> ((w & 7) * 4)) & 15u;
    if (d) acc = ed_add(acc, tab[d]);
  }
  return acc;
}

// ------------------------------------------------------------------ wire form
// point.MarshalBinary / extendedGroupElement.ToBytes with the inverse of Z supplied (callers batch inversions)
DRX_HD void ed_to_wire_zinv(uint32_t out[8], const EdExt& p, const Fe& zinv) {
  const Fe x = fe_mul(p.X, zinv);
  const Fe y = fe_canon(fe_mul(p.Y, zinv));
  DRX_UNROLL
  for (int k = 0; k < 8; k++) out[k] = y.v[k];
  out[7] |= fe_is_negative(x) << 31;
}
DRX_HD void ed_to_wire(uint32_t out[8], const EdExt& p) { ed_to_wire_zinv(out, p, fe_inv(p.Z)); }
// two points, one inversion
DRX_HD void ed_pair_to_wire(uint32_t out[16], const EdExt& a, const EdExt& b) {
  const Fe inv = fe_inv(fe_mul(a.Z, b.Z));
  ed_to_wire_zinv(out, a, fe_mul(inv, b.Z));
  ed_to_wire_zinv(out + 8, b, fe_mul(inv, a.Z));
}
// point.UnmarshalBinary / extendedGroupElement.FromBytes: false when no x exists for this y
DRX_NI bool ed_from_wire(EdExt& p, const uint32_t w[8]) {
  const uint32_t sign = w[7] >> 31;
  const Fe y = fe_from_words(w);
  const Fe yy = fe_sqr(y);
  const Fe u = fe_sub(yy, fe_one());
  const Fe v = fe_add(fe_mul(yy, fe_d()), fe_one());
  const Fe v3 = fe_mul(fe_sqr(v), v);
  const Fe uv7 = fe_mul(fe_mul(fe_sqr(v3), v), u);
  Fe x = fe_mul(fe_mul(fe_pow22523(uv7), v3), u);
  const Fe vxx = fe_mul(fe_sqr(x), v);
  if (!fe_eq(vxx, u)) {
    if (!fe_is_zero(fe_add(vxx, u))) return false;
    x = fe_mul(x, fe_sqrtm1());
  }
  if (fe_is_negative(x) != sign) x = fe_neg(x);
  p = EdExt{x, y, fe_one(), fe_mul(x, y)};
  return true;
}
// affine Niels entry of an extended point, the inverse of Z supplied
DRX_HD EdNiels ed_to_niels_zinv(const EdExt& p, const Fe& zinv) {
  const Fe x = fe_mul(p.X, zinv), y = fe_mul(p.Y, zinv);
  return EdNiels{fe_add(y, x), fe_sub(y, x), fe_mul(fe_mul(x, y), fe_d2())};
}

}  // namespace drx
