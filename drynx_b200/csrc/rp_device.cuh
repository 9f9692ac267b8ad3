// Device-side range-proof logic shared by the one-thread-per-item kernels (kernels_rp.cu) and the cooperative
// lane-group pairing (kernels_rp_ucode.cu): proof layout, item location, and the bodies of the creation and
// verification items. lib/range/range_proof.go:320-407 and :504-565.
#pragma once
#include "launch.h"
#include "keccak.cuh"
#include "wire.cuh"

namespace drx {

// ------------------------------------------------------------------ scalar helpers (mod n)
__device__ __noinline__ static Scalar sc_inv(const Scalar& a) {  // a^-1 mod n (division steps, fp.cuh); 0 -> 0
  Scalar out;
  mod_inv_plain<ModN>(out.v, a.v);
  return out;
}

// c = SHA3-512(B || C || sum Y) mod n (lib/range/range_proof.go:349-374); C is the second half of the ciphertext.
__device__ __forceinline__ Scalar rp_challenge(const uint32_t* __restrict__ baseB_wire, const uint32_t* __restrict__ ct,
                                              const uint8_t* __restrict__ sumY) {
  uint32_t msg[48], dig[16];
  for (int k = 0; k < 16; k++) {
    msg[k] = baseB_wire[k];
    msg[16 + k] = ct[16 + k];
    msg[32 + k] = ((const uint32_t*)sumY)[k];
  }
  sha3_512_words(dig, msg, 48);
  return sc_from_digest(dig);
}

// ------------------------------------------------------------------ proof layout (lib/range/range_proof.go:72-155)
struct ProofView {
  int ns, nd;
  __device__ __forceinline__ static int c_off() { return 32; }
  __device__ __forceinline__ static int zr_off() { return 40; }
  __device__ __forceinline__ static int d_off() { return 48; }
  __device__ __forceinline__ int zphi_off(int j) const { return 64 + 8 * j; }
  __device__ __forceinline__ int blk(int i) const { return 64 + 8 * nd + i * 136 * nd; }
  __device__ __forceinline__ int v_off(int i, int j) const { return blk(i) + 32 * j; }
  __device__ __forceinline__ int a_off(int i, int j) const { return blk(i) + 32 * nd + 96 * j; }
  __device__ __forceinline__ int zv_off(int i, int j) const { return blk(i) + 128 * nd + 8 * j; }
};

__device__ __forceinline__ int64_t rp_ndigits_dev(int64_t secret, int64_t u, int64_t l) {
  int64_t d = 0;
  while (secret > 0 && u > 1) {
    d++;
    secret /= u;
  }
  return d > l ? d : l;
}
__device__ __forceinline__ uint32_t rp_digit(int64_t secret, int64_t u, int j) {
  if (secret <= 0) return 0;
  for (int k = 0; k < j; k++) secret /= u;
  return (uint32_t)(secret % u);
}

struct ItemRef {
  size_t p;
  int i, j, ns, nd;
  const uint32_t* proof;
};
// Set ids of the device-buffer entry points come from device memory the host never saw: an id outside the registered
// array, an unregistered slot, or (uniform batch) a set whose server count is not the batch's must not be followed.
__device__ __forceinline__ bool proof_set_valid(const RpBatch& b, const RpSetDev* sets, size_t p) {
  const int sid = b.d_set_id[p];
  if (sid < 0) return true;  // commit-only proof (u = l = 0)
  if (sid >= b.nsets || sets[sid].ns <= 0) return false;
  return b.d_proof_off != nullptr || sets[sid].ns == b.uni_ns;
}
// false: nothing to do for this proof (commit-only, or an invalid set id, which the verification header kernel rejects)
__device__ __forceinline__ bool locate_proof(const RpBatch& b, const RpSetDev* sets, size_t p, int& ns, int& nd,
                                             size_t& byte_off) {
  int sid = b.d_set_id[p];
  if (!proof_set_valid(b, sets, p)) sid = -1;
  if (b.d_proof_off) {
    byte_off = b.d_proof_off[p];
    if (sid < 0) { ns = 0; nd = 0; return false; }
    ns = sets[sid].ns;
    nd = (int)((b.d_proof_off[p + 1] - byte_off - 256) / (size_t)(32 + 544 * ns));
  } else {
    byte_off = p * b.uni_size;
    if (sid < 0) { ns = 0; nd = 0; return false; }
    ns = b.uni_ns;
    nd = b.uni_nd;
  }
  return true;
}
__device__ __forceinline__ bool locate_item(const RpBatch& b, const RpSetDev* sets, size_t t, size_t& p, int& i, int& j,
                                            int& ns, int& nd, size_t& byte_off) {
  size_t local;
  if (b.d_item_proof) {
    p = b.d_item_proof[t];
    local = t - b.d_item_start[p];
  } else {
    const size_t per = (size_t)b.uni_ns * b.uni_nd;
    p = t / per;
    local = t - p * per;
  }
  if (!locate_proof(b, sets, p, ns, nd, byte_off)) return false;
  i = (int)(local / nd);
  j = (int)(local - (size_t)i * nd);
  return true;
}




// Threads that compute a per-(proof, digit) value (the factor e(B, g2)^(t_j), shared by all servers): in a uniform
// batch thread k handles (proof k / nd, digit k % nd) and `t` becomes server 0's item of that pair, so the threads
// stay dense; in a ragged batch every item thread comes here and the callers drop those with i != 0.
__device__ __forceinline__ bool rp_shared_factor_item(const RpBatch& b, size_t& t) {
  if (b.d_item_proof) return t < b.items;
  const size_t nd = (size_t)b.uni_nd;
  if (t >= b.n * nd) return false;
  t = (t / nd) * (size_t)b.uni_ns * nd + (t % nd);
  return true;
}

// The same item split into independent roles that run in different thread blocks (three times the threads, a
// third of the latency): role 0 writes V_ij and Zv_ij, role 1 computes e(B, A_{i,phi})^(-s_j v), role 2 computes
// e(B, g2)^(t_j); the two GT factors go to scratch and rp_create_finish multiplies and serialises them.
struct GtPair {
  Fp12 a, b;
};
__device__ __forceinline__ void rp_create_role(const RpBatch& b, const RpSetDev* __restrict__ sets,
                                               const int64_t* __restrict__ secret, const uint32_t* __restrict__ rnd,
                                               const size_t* __restrict__ rnd_off, const Fp12* __restrict__ tabGtB,
                                               const uint32_t* __restrict__ baseB_wire, const uint32_t* __restrict__ ct,
                                               uint32_t* __restrict__ proofs, GtPair* __restrict__ scratch, size_t t,
                                               int role) {
  size_t p, off;
  int i, j, ns, nd;
  if (role == 2 && !rp_shared_factor_item(b, t)) return;
  if (!locate_item(b, sets, t, p, i, j, ns, nd, off)) return;
  if (role == 2 && i != 0) return;  // e(B, g2)^(t_j) does not depend on the server: computed once per (proof, digit)
  uint32_t* out = proofs + off / 4;
  const RpSetDev set = sets[b.d_set_id[p]];
  const int u = set.u;
  ProofView pv{ns, nd};
  const uint32_t* rp = rnd + (b.d_proof_off ? rnd_off[p] / 4 : p * (size_t)nd * (3 + ns) * 8);
  const uint32_t* row = rp + (size_t)j * (3 + ns) * 8;
  const Scalar v = sc_from_wire(row + 8 * (3 + i));
  const uint32_t phi = rp_digit(secret[p], u, j);
  const size_t tab = ((size_t)i * u + phi) * (DRX_COMB_WINDOWS * DRX_COMB_ENTRIES);
  if (role == 0) {
    const Scalar tj = sc_from_wire(row + 8);
    const Scalar c = rp_challenge(baseB_wire, ct + 32 * p, set.sumY);
    const G2Jac V = comb_mul((const G2Aff*)set.tabA + tab, v);
    g2_to_wire(out + pv.v_off(i, j), jac_to_aff(V));
    sc_to_wire(out + pv.zv_off(i, j), sc_sub(tj, sc_mul(c, v)));
  } else if (role == 1) {
    const Scalar sj = sc_from_wire(row);
    gt_comb_pow(scratch[t].a, (const Fp12*)set.tabGtA + tab, sc_neg(sc_mul(sj, v)));
  } else {
    const Scalar tj = sc_from_wire(row + 8);
    gt_comb_pow(scratch[t].b, tabGtB, tj);
  }
}
__device__ __forceinline__ void rp_create_finish(const RpBatch& b, const RpSetDev* __restrict__ sets,
                                                 const GtPair* __restrict__ scratch, uint32_t* __restrict__ proofs,
                                                 size_t t) {
  size_t p, off;
  int i, j, ns, nd;
  if (!locate_item(b, sets, t, p, i, j, ns, nd, off)) return;
  ProofView pv{ns, nd};
  Fp12 a;
  fp12_mul(a, scratch[t].a, scratch[t - (size_t)i * nd].b);  // the t_j factor lives with server 0's item
  gt_to_wire(proofs + off / 4 + pv.a_off(i, j), a);
}

// Finer split for small batches, where creation is pure latency: the 32 comb windows of each of the three
// fixed-base walks (V_ij in G2, the two GT factors) are dealt to four thread blocks each; partial results go to
// scratch and two finish roles fold them (3 G2 additions + one inversion; 7 GT products).
struct CreateFine {
  Fp12 g[8];     // g[0..3]: partial products of e(B, A_{i,phi})^(-s v); g[4..7]: of e(B, g2)^t
  G2Jac vp[4];   // partial sums of v * A_{i,phi}
};
__device__ __forceinline__ void gt_comb_partial(Fp12& out, const Fp12* __restrict__ table, const Scalar& k, int first,
                                                int stride = 4) {
  Fp12 acc = fp12_one();
  bool any = false;
  for (int w = first; w < DRX_COMB_WINDOWS; w += stride) {
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
__device__ __forceinline__ void rp_create_fine_role(const RpBatch& b, const RpSetDev* __restrict__ sets,
                                                    const int64_t* __restrict__ secret, const uint32_t* __restrict__ rnd,
                                                    const size_t* __restrict__ rnd_off, const Fp12* __restrict__ tabGtB,
                                                    const uint32_t* __restrict__ baseB_wire,
                                                    const uint32_t* __restrict__ ct, uint32_t* __restrict__ proofs,
                                                    CreateFine* __restrict__ scratch, size_t t, int role) {
  size_t p, off;
  int i, j, ns, nd;
  if (role >= 8 && !rp_shared_factor_item(b, t)) return;
  if (!locate_item(b, sets, t, p, i, j, ns, nd, off)) return;
  if (role >= 8 && i != 0) return;  // the e(B, g2)^(t_j) partials: once per (proof, digit)
  const RpSetDev set = sets[b.d_set_id[p]];
  const int u = set.u;
  ProofView pv{ns, nd};
  const uint32_t* rp = rnd + (b.d_proof_off ? rnd_off[p] / 4 : p * (size_t)nd * (3 + ns) * 8);
  const uint32_t* row = rp + (size_t)j * (3 + ns) * 8;
  const Scalar v = sc_from_wire(row + 8 * (3 + i));
  const uint32_t phi = rp_digit(secret[p], u, j);
  const size_t tab = ((size_t)i * u + phi) * (DRX_COMB_WINDOWS * DRX_COMB_ENTRIES);
  if (role < 4) {
    const G2Aff* tA = (const G2Aff*)set.tabA + tab;
    G2Jac acc = jac_inf<Fp2>();
    for (int w = role; w < DRX_COMB_WINDOWS; w += 4) {
      const uint32_t d = (v.v[w >> 2] >> ((w & 3) * 8)) & 0xffu;
      if (d) acc = jac_add_aff(acc, tA[w * DRX_COMB_ENTRIES + (int)d - 1]);
    }
    scratch[t].vp[role] = acc;
    if (role == 3) {  // the scalar Zv_ij = t_j - c v rides along
      const Scalar tj = sc_from_wire(row + 8);
      const Scalar c = rp_challenge(baseB_wire, ct + 32 * p, set.sumY);
      sc_to_wire(proofs + off / 4 + pv.zv_off(i, j), sc_sub(tj, sc_mul(c, v)));
    }
  } else if (role < 8) {
    const Scalar sj = sc_from_wire(row);
    gt_comb_partial(scratch[t].g[role - 4], (const Fp12*)set.tabGtA + tab, sc_neg(sc_mul(sj, v)), role - 4);
  } else {
    const Scalar tj = sc_from_wire(row + 8);
    gt_comb_partial(scratch[t].g[role - 4], tabGtB, tj, role - 8);
  }
}
__device__ __forceinline__ void rp_create_fine_finish(const RpBatch& b, const RpSetDev* __restrict__ sets,
                                                      const CreateFine* __restrict__ scratch,
                                                      uint32_t* __restrict__ proofs, size_t t, int role) {
  size_t p, off;
  int i, j, ns, nd;
  if (!locate_item(b, sets, t, p, i, j, ns, nd, off)) return;
  ProofView pv{ns, nd};
  uint32_t* out = proofs + off / 4;
  if (role == 0) {
    G2Jac V = jac_add(jac_add(scratch[t].vp[0], scratch[t].vp[1]), jac_add(scratch[t].vp[2], scratch[t].vp[3]));
    g2_to_wire(out + pv.v_off(i, j), jac_to_aff(V));
  } else {
    const size_t t0 = t - (size_t)i * nd;  // server 0's item of this digit holds the t_j partials
    Fp12 a, c2;
    fp12_mul(a, scratch[t].g[0], scratch[t].g[1]);
    fp12_mul(c2, scratch[t].g[2], scratch[t].g[3]);
    fp12_mul(a, a, c2);
    fp12_mul(c2, scratch[t0].g[4], scratch[t0].g[5]);
    fp12_mul(a, a, c2);
    fp12_mul(c2, scratch[t0].g[6], scratch[t0].g[7]);
    fp12_mul(a, a, c2);
    gt_to_wire(out + pv.a_off(i, j), a);
  }
}

// Lane-cooperative creation for small batches (one launch, no scratch). blockIdx.y selects the part of the item:
//   0: V_ij = v * A_{i,phi} in G2, eight lanes per item: four comb windows each, then a three-level butterfly of
//      full additions over warp shuffles; lane 0 converts to affine and serialises.
//   1: a_ij = e(B, A_{i,phi})^(-s v) * e(B, g2)^t, sixteen lanes per item: lanes 0-7 take four windows each of the
//      first power, lanes 8-15 of the second, then a four-level butterfly of Fp12 products (7 dependent products
//      per item instead of 15 through the scratch-based split, 62 on one thread).
//   2: Zv_ij = t_j - c v, one thread per item.
template <int NW, class T>
__device__ __forceinline__ T shfl_xor_words(const T& x, int m, uint32_t mask) {
  static_assert(sizeof(T) == NW * 4, "word count");
  T r;
  const uint32_t* src = reinterpret_cast<const uint32_t*>(&x);
  uint32_t* dst = reinterpret_cast<uint32_t*>(&r);
#pragma unroll 8
  for (int k = 0; k < NW; k++) dst[k] = __shfl_xor_sync(mask, src[k], m);
  return r;
}
__device__ __forceinline__ void rp_create_lanes(const RpBatch& b, const RpSetDev* __restrict__ sets,
                                                const int64_t* __restrict__ secret, const uint32_t* __restrict__ rnd,
                                                const size_t* __restrict__ rnd_off, const Fp12* __restrict__ tabGtB,
                                                const uint32_t* __restrict__ baseB_wire, const uint32_t* __restrict__ ct,
                                                uint32_t* __restrict__ proofs, int role) {
  const size_t gtid = (size_t)blockIdx.x * blockDim.x + threadIdx.x;
  const int shift = role == 0 ? 3 : role == 1 ? 4 : 0;
  const size_t t = gtid >> shift;
  if (t >= b.items) return;  // whole lane groups leave together
  size_t p, off;
  int i, j, ns, nd;
  if (!locate_item(b, sets, t, p, i, j, ns, nd, off)) return;
  const RpSetDev set = sets[b.d_set_id[p]];
  ProofView pv{ns, nd};
  uint32_t* out = proofs + off / 4;
  const uint32_t* rp = rnd + (b.d_proof_off ? rnd_off[p] / 4 : p * (size_t)nd * (3 + ns) * 8);
  const uint32_t* row = rp + (size_t)j * (3 + ns) * 8;
  const Scalar v = sc_from_wire(row + 8 * (3 + i));
  if (role == 2) {
    const Scalar tj = sc_from_wire(row + 8);
    const Scalar c = rp_challenge(baseB_wire, ct + 32 * p, set.sumY);
    sc_to_wire(out + pv.zv_off(i, j), sc_sub(tj, sc_mul(c, v)));
    return;
  }
  const uint32_t phi = rp_digit(secret[p], set.u, j);
  const size_t tab = ((size_t)i * set.u + phi) * (DRX_COMB_WINDOWS * DRX_COMB_ENTRIES);
  if (role == 0) {
    const int lane = threadIdx.x & 7;
    const uint32_t mask = 0xFFu << (threadIdx.x & 24u);
    const G2Aff* tA = (const G2Aff*)set.tabA + tab;
    G2Jac acc = jac_inf<Fp2>();
    for (int w = lane; w < DRX_COMB_WINDOWS; w += 8) {
      const uint32_t d = (v.v[w >> 2] >> ((w & 3) * 8)) & 0xffu;
      if (d) acc = jac_add_aff(acc, tA[w * DRX_COMB_ENTRIES + (int)d - 1]);
    }
    for (int s = 1; s <= 4; s <<= 1) {
      const G2Jac o = shfl_xor_words<48>(acc, s, mask);
      acc = (lane & s) ? jac_add(o, acc) : jac_add(acc, o);
    }
    if (lane == 0) g2_to_wire(out + pv.v_off(i, j), jac_to_aff(acc));
  } else {
    const int lane = threadIdx.x & 15;
    const uint32_t mask = 0xFFFFu << (threadIdx.x & 16u);
    Fp12 acc;
    if (lane < 8) {
      const Scalar sj = sc_from_wire(row);
      gt_comb_partial(acc, (const Fp12*)set.tabGtA + tab, sc_neg(sc_mul(sj, v)), lane, 8);
    } else {
      gt_comb_partial(acc, tabGtB, sc_from_wire(row + 8), lane - 8, 8);
    }
    for (int s = 1; s <= 8; s <<= 1) {
      const Fp12 o = shfl_xor_words<96>(acc, s, mask);
      fp12_mul(acc, acc, o);
    }
    if (lane == 0) gt_to_wire(out + pv.a_off(i, j), acc);
  }
}

// Verification item in two roles: role 0 computes e = e(c*Y_i - Zphi_j*B, V_ij), role 1 computes
// T = A_ij * e(B, g2)^(-Zv_ij); rp_verify_finish accepts iff e == T (the same equation as below, moved across).
__device__ __forceinline__ void rp_verify_role(const RpBatch& b, const RpSetDev* __restrict__ sets,
                                               const G1Aff* __restrict__ tabB, const Fp12* __restrict__ tabGtB,
                                               const uint32_t* __restrict__ proofs, uint32_t* __restrict__ flags,
                                               GtPair* __restrict__ scratch, size_t t, int role) {
  size_t p, off;
  int i, j, ns, nd;
  if (!locate_item(b, sets, t, p, i, j, ns, nd, off)) return;
  const uint32_t* in = proofs + off / 4;
  const RpSetDev set = sets[b.d_set_id[p]];
  ProofView pv{ns, nd};
  if (role == 0) {
    G2Aff V;
    if (!g2_from_wire(V, in + pv.v_off(i, j))) {
      atomicAnd(&flags[p], 0u);
      scratch[t].a = fp12_one();
      return;
    }
    const Scalar c = sc_from_wire(in + ProofView::c_off());
    const Scalar zphi = sc_from_wire(in + pv.zphi_off(j));
    G1Jac P1 = comb_mul((const G1Aff*)set.tabY + (size_t)i * (DRX_COMB_WINDOWS * DRX_COMB_ENTRIES), c);
    P1 = jac_add(P1, comb_mul(tabB, sc_neg(zphi)));
    if (!pairing(scratch[t].a, jac_to_aff(P1), V)) atomicAnd(&flags[p], 0u);  // V_ij outside G2: malformed for kyber
  } else {
    Fp12 A, g;
    if (!gt_from_wire(A, in + pv.a_off(i, j))) {
      atomicAnd(&flags[p], 0u);
      scratch[t].b = fp12_one();
      return;
    }
    const Scalar zv = sc_from_wire(in + pv.zv_off(i, j));
    gt_comb_pow(g, tabGtB, sc_neg(zv));
    fp12_mul(scratch[t].b, A, g);
  }
}
__device__ __forceinline__ void rp_verify_finish(const RpBatch& b, const RpSetDev* __restrict__ sets,
                                                 const GtPair* __restrict__ scratch, uint32_t* __restrict__ flags,
                                                 size_t t) {
  size_t p, off;
  int i, j, ns, nd;
  if (!locate_item(b, sets, t, p, i, j, ns, nd, off)) return;
  if (!fp12_eq(scratch[t].a, scratch[t].b)) atomicAnd(&flags[p], 0u);
}



}  // namespace drx
