// Wire codecs: the fixed-width big-endian byte forms that cross the C ABI.
//   scalar 32 B | G1 64 B (x||y, infinity = zeros) | G2 128 B (x.i||x.1||y.i||y.1) | GT 384 B
// following kyber v3.0.8 pairing/bn256 point.go MarshalBinary/UnmarshalBinary (not in /root/reference) as
// used by libunlynx CipherText.ToBytes and lib/range/range_proof.go:72-246. All buffers are 4-byte aligned
// (device allocations, offsets are multiples of 32), so fields move as 32-bit words.
#pragma once
#include "pairing.cuh"

namespace drx {

// Returns false for a malformed encoding (coordinate >= p or point off the curve), like kyber's
// UnmarshalBinary error; out is then the point at infinity.
DRX_HD bool g1_from_wire(G1Aff& out, const uint32_t* w) {
  uint32_t any = 0;
  uint32_t a[8], b[8];
  DRX_UNROLL
  for (int k = 0; k < 8; k++) { a[k] = w[k]; b[k] = w[8 + k]; any |= a[k] | b[k]; }
  if (!any) { out.x = fp_zero(); out.y = fp_zero(); return true; }
  bool ok = fp_from_wire(out.x, a);
  ok &= fp_from_wire(out.y, b);
  ok = ok && aff_on_curve(out);
  if (!ok) { out.x = fp_zero(); out.y = fp_zero(); }
  return ok;
}
DRX_HD void g1_to_wire(uint32_t* w, const G1Aff& p) {
  uint32_t a[8], b[8];
  fp_to_wire(a, p.x);
  fp_to_wire(b, p.y);
  DRX_UNROLL
  for (int k = 0; k < 8; k++) { w[k] = a[k]; w[8 + k] = b[k]; }
}
DRX_HD bool fp2_from_wire(Fp2& out, const uint32_t* w) {  // i-coefficient first
  uint32_t a[8], b[8];
  DRX_UNROLL
  for (int k = 0; k < 8; k++) { a[k] = w[k]; b[k] = w[8 + k]; }
  bool ok = fp_from_wire(out.c1, a);
  ok &= fp_from_wire(out.c0, b);
  return ok;
}
DRX_HD void fp2_to_wire(uint32_t* w, const Fp2& v) {
  uint32_t a[8], b[8];
  fp_to_wire(a, v.c1);
  fp_to_wire(b, v.c0);
  DRX_UNROLL
  for (int k = 0; k < 8; k++) { w[k] = a[k]; w[8 + k] = b[k]; }
}
DRX_HD bool g2_from_wire(G2Aff& out, const uint32_t* w) {
  uint32_t any = 0;
  for (int k = 0; k < 32; k++) any |= w[k];
  if (!any) { out.x = fp2_zero(); out.y = fp2_zero(); return true; }
  bool ok = fp2_from_wire(out.x, w);
  ok &= fp2_from_wire(out.y, w + 16);
  ok = ok && aff_on_curve(out);
  if (!ok) { out.x = fp2_zero(); out.y = fp2_zero(); }
  return ok;
}
// g2_from_wire plus membership in the order-n subgroup: what kyber's UnmarshalBinary accepts. For the set-up and
// diagnostic paths; range-proof verification takes the verdict from its Miller loop instead (pairing.cuh).
DRX_HD bool g2_from_wire_checked(G2Aff& out, const uint32_t* w) {
  bool ok = g2_from_wire(out, w);
  ok = ok && g2_in_subgroup(out);
  if (!ok) { out.x = fp2_zero(); out.y = fp2_zero(); }
  return ok;
}
DRX_HD void g2_to_wire(uint32_t* w, const G2Aff& p) {
  fp2_to_wire(w, p.x);
  fp2_to_wire(w + 16, p.y);
}
DRX_HD bool gt_from_wire(Fp12& out, const uint32_t* w) {
  bool ok = true;
  for (int s = 0; s < 6; s++) ok &= fp2_from_wire(fp12_wire_slot(out, s), w + 16 * s);
  return ok;
}
DRX_HD void gt_to_wire(uint32_t* w, const Fp12& a) {
  for (int s = 0; s < 6; s++) fp2_to_wire(w + 16 * s, fp12_wire_slot(a, s));
}

}  // namespace drx
