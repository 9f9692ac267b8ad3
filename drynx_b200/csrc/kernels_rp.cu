// BN256 pairing-based range proofs on sm_100a: creation and verification, plus the G2 / GT set-up kernels.
// Replaces lib/range/range_proof.go:249-288 (signatures), :320-407 (CreatePredicateRangeProofForAllServ) and
// :504-565 (RangeProofVerification) of the reference, and the kyber pairing/bn256 calls they make.
#include <cstdio>
#include <cstdlib>
#include <mutex>
#include "rp_device.cuh"

namespace drx {

static inline int grid_for(size_t n, int threads) { return (int)((n + threads - 1) / threads); }

// Kernels that run beside the lane-group pairing ask for the shared-memory-heavy L1 split (they use no shared memory
// themselves): an SM they reach first then still has room for pairing blocks. Per device, once (a hint, errors ignored).
template <class K>
static void prefer_shared_carveout(K kernel, bool (&done)[64]) {
  static std::mutex mu;
  int dev = 0;
  if (cudaGetDevice(&dev) != cudaSuccess || dev < 0 || dev >= 64) return;
  std::lock_guard<std::mutex> lk(mu);
  if (done[dev]) return;
  if (cudaFuncSetAttribute(kernel, cudaFuncAttributePreferredSharedMemoryCarveout, (int)cudaSharedmemCarveoutMaxShared) != cudaSuccess)
    cudaGetLastError();
  done[dev] = true;
}


// ------------------------------------------------------------------ set-up
__global__ void k_g2_from_wire(size_t n, const uint32_t* __restrict__ wire, G2Aff* __restrict__ aff, int* bad) {
  const size_t i = (size_t)blockIdx.x * blockDim.x + threadIdx.x;
  if (i >= n) return;
  G2Aff p;
  if (!g2_from_wire_checked(p, wire + 32 * i)) atomicAdd(bad, 1);
  aff[i] = p;
}
int launch_g2_from_wire(cudaStream_t st, size_t n, const uint8_t* d_wire, void* d_aff, int* d_bad) {
  if (!n) return 0;
  k_g2_from_wire<<<grid_for(n, 64), 64, 0, st>>>(n, (const uint32_t*)d_wire, (G2Aff*)d_aff, d_bad);
  return 1;
}

__global__ void __launch_bounds__(64)
k_comb_table_g2(int nbases, const G2Aff* __restrict__ bases, G2Aff* __restrict__ tables) {
  const int idx = blockIdx.x * blockDim.x + threadIdx.x;
  if (idx >= nbases * DRX_COMB_WINDOWS * DRX_COMB_ENTRIES) return;
  const int b = idx / (DRX_COMB_WINDOWS * DRX_COMB_ENTRIES);
  const int rem = idx - b * (DRX_COMB_WINDOWS * DRX_COMB_ENTRIES);
  const int w = rem / DRX_COMB_ENTRIES;
  const int d = rem - w * DRX_COMB_ENTRIES + 1;
  G2Jac base = jac_from_aff(bases[b]);
  for (int s = 0; s < 8 * w; s++) base = jac_dbl(base);
  G2Jac acc = jac_inf<Fp2>();
  for (int bit = 7; bit >= 0; bit--) {
    acc = jac_dbl(acc);
    if ((d >> bit) & 1) acc = jac_add(acc, base);
  }
  tables[idx] = jac_to_aff(acc);
}
int launch_comb_table_g2(cudaStream_t st, int nbases, const void* d_bases_aff, void* d_tables) {
  const size_t n = (size_t)nbases * kCombPoints;
  if (!n) return 0;
  k_comb_table_g2<<<grid_for(n, 64), 64, 0, st>>>(nbases, (const G2Aff*)d_bases_aff, (G2Aff*)d_tables);
  return 1;
}

// gt[i] = e(g1[i * stride], g2[i]) on Montgomery affine inputs
__global__ void __launch_bounds__(64)
k_pair_aff(size_t n, const G1Aff* __restrict__ g1, size_t stride, const G2Aff* __restrict__ g2, Fp12* __restrict__ gt) {
  const size_t i = (size_t)blockIdx.x * blockDim.x + threadIdx.x;
  if (i >= n) return;
  const G1Aff P = g1[i * stride];
  const G2Aff Q = g2[i];
  Fp12 e;
  pairing(e, P, Q);
  gt[i] = e;
}
int launch_pair_aff(cudaStream_t st, size_t n, const void* d_g1_aff, size_t stride, const void* d_g2_aff, void* d_gt) {
  if (!n) return 0;
  k_pair_aff<<<grid_for(n, 64), 64, 0, st>>>(n, (const G1Aff*)d_g1_aff, stride, (const G2Aff*)d_g2_aff, (Fp12*)d_gt);
  return 1;
}

// tables[b][w][d-1] = base_b ^ (d * 2^(8w)); one thread per (b, w) walks its 255 entries
__global__ void __launch_bounds__(32)
k_comb_table_gt(int nbases, const Fp12* __restrict__ bases, Fp12* __restrict__ tables) {
  const int idx = blockIdx.x * blockDim.x + threadIdx.x;
  if (idx >= nbases * DRX_COMB_WINDOWS) return;
  const int b = idx / DRX_COMB_WINDOWS;
  const int w = idx - b * DRX_COMB_WINDOWS;
  Fp12 base = bases[b];
  for (int s = 0; s < 8 * w; s++) fp12_cyclo_sqr(base, base);
  Fp12 acc = base;
  Fp12* out = tables + ((size_t)b * DRX_COMB_WINDOWS + w) * DRX_COMB_ENTRIES;
  for (int d = 1; d <= DRX_COMB_ENTRIES; d++) {
    out[d - 1] = acc;
    fp12_mul(acc, acc, base);
  }
}
int launch_comb_table_gt(cudaStream_t st, int nbases, const void* d_bases_gt, void* d_tables) {
  const size_t n = (size_t)nbases * kCombWindows;
  if (!n) return 0;
  k_comb_table_gt<<<grid_for(n, 32), 32, 0, st>>>(nbases, (const Fp12*)d_bases_gt, (Fp12*)d_tables);
  return 1;
}

// 16-bit windows for the one base every proof shares, e(B, g2): table16[w][d-1] = base ^ (d * 2^(16w)), w < 16,
// d = 1 .. 65535 (403 MB), each entry the product of the two 8-bit entries it consists of. Halves the Fp12 products of
// the fixed-base exponentiations e(B, g2)^t (creation) and e(B, g2)^(-Zv) (verification) in the lane-group GT kernels.
__global__ void __launch_bounds__(64)
k_comb_table_gt16(const Fp12* __restrict__ tab8, Fp12* __restrict__ tab16) {
  const size_t idx = (size_t)blockIdx.x * blockDim.x + threadIdx.x;
  if (idx >= kGt16Entries) return;
  const int w = (int)(idx / kGt16PerWindow);
  const uint32_t d = (uint32_t)(idx - (size_t)w * kGt16PerWindow) + 1u, lo = d & 0xffu, hi = d >> 8;
  const Fp12* elo = tab8 + (size_t)(2 * w) * DRX_COMB_ENTRIES + lo - 1;
  const Fp12* ehi = tab8 + (size_t)(2 * w + 1) * DRX_COMB_ENTRIES + hi - 1;
  Fp12 r;
  if (lo && hi) {
    fp12_mul(r, *elo, *ehi);
  } else {
    r = lo ? *elo : *ehi;
  }
  tab16[idx] = r;
}
int launch_comb_table_gt16(cudaStream_t st, const void* d_tab8, void* d_tab16) {
  k_comb_table_gt16<<<grid_for(kGt16Entries, 64), 64, 0, st>>>((const Fp12*)d_tab8, (Fp12*)d_tab16);
  return 1;
}

// ------------------------------------------------------------------ diagnostics on wire bytes
__global__ void __launch_bounds__(64)
k_pair_wire(size_t n, const uint32_t* __restrict__ g1, const uint32_t* __restrict__ g2, uint32_t* __restrict__ gt,
            int* bad) {
  const size_t i = (size_t)blockIdx.x * blockDim.x + threadIdx.x;
  if (i >= n) return;
  G1Aff P;
  G2Aff Q;
  int nbad = !g1_from_wire(P, g1 + 16 * i);
  nbad += !g2_from_wire_checked(Q, g2 + 32 * i);
  if (nbad) atomicAdd(bad, nbad);
  Fp12 e;
  pairing(e, P, Q);
  gt_to_wire(gt + 96 * i, e);
}
int launch_pair_wire(cudaStream_t st, size_t n, const uint8_t* d_g1, const uint8_t* d_g2, uint8_t* d_gt, int* d_bad) {
  if (!n) return 0;
  k_pair_wire<<<grid_for(n, 64), 64, 0, st>>>(n, (const uint32_t*)d_g1, (const uint32_t*)d_g2, (uint32_t*)d_gt, d_bad);
  return 1;
}

__global__ void __launch_bounds__(64)
k_g2_mul(size_t n, const uint32_t* __restrict__ k, const uint32_t* __restrict__ pt, uint32_t* __restrict__ out, int* bad) {
  const size_t i = (size_t)blockIdx.x * blockDim.x + threadIdx.x;
  if (i >= n) return;
  G2Aff p;
  if (!g2_from_wire_checked(p, pt + 32 * i)) atomicAdd(bad, 1);
  const G2Jac r = window_mul(jac_from_aff(p), sc_from_wire(k + 8 * i));
  g2_to_wire(out + 32 * i, jac_to_aff(r));
}
int launch_g2_mul(cudaStream_t st, size_t n, const uint8_t* d_k, const uint8_t* d_pt, uint8_t* d_out, int* d_bad) {
  if (!n) return 0;
  k_g2_mul<<<grid_for(n, 64), 64, 0, st>>>(n, (const uint32_t*)d_k, (const uint32_t*)d_pt, (uint32_t*)d_out, d_bad);
  return 1;
}

}  // namespace drx

namespace drx {

// ------------------------------------------------------------------ signatures (a-10)
__global__ void __launch_bounds__(64)
k_make_sigs(int ns, int64_t u, const uint32_t* __restrict__ x, const G1Aff* __restrict__ tabB,
            const G2Aff* __restrict__ tabG2, uint32_t* __restrict__ Y, uint32_t* __restrict__ A) {
  const int t = blockIdx.x * blockDim.x + threadIdx.x;
  const int per = (int)u + 1;
  if (t >= ns * per) return;
  const int i = t / per, k = t - i * per;
  const Scalar xi = sc_from_wire(x + 8 * i);
  if (k == (int)u) {
    g1_to_wire(Y + 16 * i, jac_to_aff(comb_mul(tabB, xi)));
    return;
  }
  const Scalar e = sc_inv(sc_add(xi, sc_from_u64((uint64_t)k)));
  g2_to_wire(A + 32 * ((size_t)i * u + k), jac_to_aff(comb_mul(tabG2, e)));
}
int launch_make_sigs(cudaStream_t st, int ns, int64_t u, const uint8_t* d_x, const void* d_tabB, const void* d_tabG2,
                     uint8_t* d_Y, uint8_t* d_A) {
  const size_t n = (size_t)ns * (u + 1);
  if (!n) return 0;
  k_make_sigs<<<grid_for(n, 64), 64, 0, st>>>(ns, u, (const uint32_t*)d_x, (const G1Aff*)d_tabB, (const G2Aff*)d_tabG2,
                                             (uint32_t*)d_Y, (uint32_t*)d_A);
  return 1;
}

// ------------------------------------------------------------------ creation (a-6)
// Header per proof: Commit, Challenge, Zr, D, Zphi. The scalar part (challenge hash, Zphi_j, Zr, the two exponents
// of D) is shared by the one-thread kernel and the eight-lane kernel; `write` selects the lane that stores.
__device__ __forceinline__ bool rp_create_hdr_scalars(const RpBatch& b, const RpSetDev* __restrict__ sets,
                                                      const int64_t* __restrict__ secret, const uint32_t* __restrict__ r,
                                                      const uint32_t* __restrict__ ct, const uint32_t* __restrict__ rnd,
                                                      const size_t* __restrict__ rnd_off,
                                                      const uint32_t* __restrict__ baseB_wire, uint32_t* __restrict__ proofs,
                                                      size_t p, bool write, Scalar& ssum, Scalar& msum, uint32_t*& out) {
  int ns, nd;
  size_t off;
  const bool ranged = locate_proof(b, sets, p, ns, nd, off);
  out = proofs + off / 4;
  if (write)
    for (int k = 0; k < 32; k++) out[k] = ct[32 * p + k];
  if (!ranged) return false;
  const RpSetDev set = sets[b.d_set_id[p]];
  const int64_t u = set.u;
  const int64_t sec = secret[p];
  // c = SHA3-512(B || C || sum Y) mod n
  const Scalar c = rp_challenge(baseB_wire, ct + 32 * p, set.sumY);
  const uint32_t* rp = rnd + (b.d_proof_off ? rnd_off[p] / 4 : p * (size_t)nd * (3 + ns) * 8);
  ProofView pv{ns, nd};
  msum = sc_zero();
  ssum = sc_zero();
  uint64_t upow = 1;
  int64_t rest = sec > 0 ? sec : 0;
  for (int j = 0; j < nd; j++) {
    const uint32_t* row = rp + (size_t)j * (3 + ns) * 8;
    const Scalar sj = sc_from_wire(row);
    const Scalar mj = sc_from_wire(row + 16);
    msum = sc_add(msum, mj);
    ssum = sc_add(ssum, sc_mul(sj, sc_from_u64(upow)));
    const uint64_t phi = (uint64_t)(rest % u);
    rest /= u;
    const Scalar zphi = sc_sub(sj, sc_mul(c, sc_from_u64(phi)));
    if (write) sc_to_wire(out + pv.zphi_off(j), zphi);
    upow *= (uint64_t)u;
  }
  if (write) {
    const Scalar zr = sc_sub(msum, sc_mul(c, sc_from_wire(r + 8 * p)));
    sc_to_wire(out + ProofView::zr_off(), zr);
    sc_to_wire(out + ProofView::c_off(), c);
  }
  return true;
}
__global__ void __launch_bounds__(128)
k_rp_create_hdr(RpBatch b, const RpSetDev* __restrict__ sets, const int64_t* __restrict__ secret,
                const uint32_t* __restrict__ r, const uint32_t* __restrict__ ct, const uint32_t* __restrict__ rnd,
                const size_t* __restrict__ rnd_off, const G1Aff* __restrict__ tabB, const G1Aff* __restrict__ tabP,
                const uint32_t* __restrict__ baseB_wire, uint32_t* __restrict__ proofs) {
  const size_t p = (size_t)blockIdx.x * blockDim.x + threadIdx.x;
  if (p >= b.n) return;
  Scalar ssum, msum;
  uint32_t* out;
  if (!rp_create_hdr_scalars(b, sets, secret, r, ct, rnd, rnd_off, baseB_wire, proofs, p, true, ssum, msum, out)) return;
  const G1Jac D = jac_add(comb_mul(tabB, ssum), comb_mul(tabP, msum));
  g1_to_wire(out + ProofView::d_off(), jac_to_aff(D));
}
// Small batches: eight lanes per proof; lanes 0-3 walk every fourth window of ssum*B, lanes 4-7 of msum*P, a
// butterfly of full additions joins them (the scalars are computed by every lane, stored by lane 0).
__global__ void __launch_bounds__(64)
k_rp_create_hdr_lanes(RpBatch b, const RpSetDev* __restrict__ sets, const int64_t* __restrict__ secret,
                      const uint32_t* __restrict__ r, const uint32_t* __restrict__ ct, const uint32_t* __restrict__ rnd,
                      const size_t* __restrict__ rnd_off, const G1Aff* __restrict__ tabB,
                      const G1Aff* __restrict__ tabP, const uint32_t* __restrict__ baseB_wire,
                      uint32_t* __restrict__ proofs) {
  const size_t p = ((size_t)blockIdx.x * blockDim.x + threadIdx.x) >> 3;
  const int lane = threadIdx.x & 7;
  const uint32_t mask = 0xFFu << (threadIdx.x & 24u);
  if (p >= b.n) return;  // whole groups of eight leave together
  Scalar ssum, msum;
  uint32_t* out;
  if (!rp_create_hdr_scalars(b, sets, secret, r, ct, rnd, rnd_off, baseB_wire, proofs, p, lane == 0, ssum, msum, out))
    return;  // uniform across the eight lanes
  const Scalar k = lane < 4 ? ssum : msum;
  const G1Aff* tab = lane < 4 ? tabB : tabP;
  G1Jac acc = jac_inf<Fp>();
  for (int w = lane & 3; w < DRX_COMB_WINDOWS; w += 4) {
    const uint32_t d = (k.v[w >> 2] >> ((w & 3) * 8)) & 0xffu;
    if (d) acc = jac_add_aff(acc, tab[w * DRX_COMB_ENTRIES + (int)d - 1]);
  }
#pragma unroll
  for (int s = 1; s <= 4; s <<= 1) {
    G1Jac o;
#pragma unroll
    for (int q = 0; q < 8; q++) {
      o.x.v[q] = __shfl_xor_sync(mask, acc.x.v[q], s);
      o.y.v[q] = __shfl_xor_sync(mask, acc.y.v[q], s);
      o.z.v[q] = __shfl_xor_sync(mask, acc.z.v[q], s);
    }
    acc = (lane & s) ? jac_add(o, acc) : jac_add(acc, o);
  }
  if (lane == 0) g1_to_wire(out + ProofView::d_off(), jac_to_aff(acc));
}

// Role-split creation: blockIdx.y = role (see rp_device.cuh).
__global__ void __launch_bounds__(64)
k_rp_create_roles(RpBatch b, const RpSetDev* __restrict__ sets, const int64_t* __restrict__ secret,
                  const uint32_t* __restrict__ rnd, const size_t* __restrict__ rnd_off, const Fp12* __restrict__ tabGtB,
                  const uint32_t* __restrict__ baseB_wire, const uint32_t* __restrict__ ct, uint32_t* __restrict__ proofs,
                  GtPair* __restrict__ scratch) {
  const size_t t = (size_t)blockIdx.x * blockDim.x + threadIdx.x;
  if (t >= b.items) return;
  rp_create_role(b, sets, secret, rnd, rnd_off, tabGtB, baseB_wire, ct, proofs, scratch, t, (int)blockIdx.y);
}
__global__ void __launch_bounds__(64)
k_rp_create_finish(RpBatch b, const RpSetDev* __restrict__ sets, const GtPair* __restrict__ scratch,
                   uint32_t* __restrict__ proofs) {
  const size_t t = (size_t)blockIdx.x * blockDim.x + threadIdx.x;
  if (t >= b.items) return;
  rp_create_finish(b, sets, scratch, proofs, t);
}
__global__ void __launch_bounds__(64)
k_rp_create_fine(RpBatch b, const RpSetDev* __restrict__ sets, const int64_t* __restrict__ secret,
                 const uint32_t* __restrict__ rnd, const size_t* __restrict__ rnd_off, const Fp12* __restrict__ tabGtB,
                 const uint32_t* __restrict__ baseB_wire, const uint32_t* __restrict__ ct, uint32_t* __restrict__ proofs,
                 CreateFine* __restrict__ scratch) {
  const size_t t = (size_t)blockIdx.x * blockDim.x + threadIdx.x;
  if (t >= b.items) return;
  rp_create_fine_role(b, sets, secret, rnd, rnd_off, tabGtB, baseB_wire, ct, proofs, scratch, t, (int)blockIdx.y);
}
__global__ void __launch_bounds__(64)
k_rp_create_fine_finish(RpBatch b, const RpSetDev* __restrict__ sets, const CreateFine* __restrict__ scratch,
                        uint32_t* __restrict__ proofs) {
  const size_t t = (size_t)blockIdx.x * blockDim.x + threadIdx.x;
  if (t >= b.items) return;
  rp_create_fine_finish(b, sets, scratch, proofs, t, (int)blockIdx.y);
}
// scratch for either creation path (the fine one needs 3,840 B per item, the role-split one 768 B)
size_t rp_scratch_bytes(size_t items) {
  return items * (items <= kCreateFineMaxItems ? sizeof(CreateFine) : sizeof(GtPair));
}
int launch_rp_create_fine(cudaStream_t st, const RpBatch& b, const RpSetDev* d_sets, const int64_t* d_secret,
                          const uint8_t* d_rnd, const size_t* d_rnd_off, const void* d_tabGtB,
                          const uint8_t* d_baseB_wire, const uint8_t* d_ct, uint8_t* d_proofs, void* d_scratch,
                          bool g2_only) {
  if (!b.items) return 0;
  // roles 0-3 are the four window quarters of V_ij (role 3 also writes Zv_ij), finish role 0 folds them
  dim3 grid(grid_for(b.items, 64), g2_only ? 4 : 12), fin(grid_for(b.items, 64), g2_only ? 1 : 2);
  k_rp_create_fine<<<grid, 64, 0, st>>>(b, d_sets, d_secret, (const uint32_t*)d_rnd, d_rnd_off, (const Fp12*)d_tabGtB,
                                        (const uint32_t*)d_baseB_wire, (const uint32_t*)d_ct, (uint32_t*)d_proofs,
                                        (CreateFine*)d_scratch);
  k_rp_create_fine_finish<<<fin, 64, 0, st>>>(b, d_sets, (const CreateFine*)d_scratch, (uint32_t*)d_proofs);
  return 2;
}
__global__ void __launch_bounds__(64)
k_rp_create_lanes(RpBatch b, const RpSetDev* __restrict__ sets, const int64_t* __restrict__ secret,
                  const uint32_t* __restrict__ rnd, const size_t* __restrict__ rnd_off, const Fp12* __restrict__ tabGtB,
                  const uint32_t* __restrict__ baseB_wire, const uint32_t* __restrict__ ct, uint32_t* __restrict__ proofs) {
  rp_create_lanes(b, sets, secret, rnd, rnd_off, tabGtB, baseB_wire, ct, proofs, (int)blockIdx.y);
}
int launch_rp_create_lanes(cudaStream_t st, const RpBatch& b, const RpSetDev* d_sets, const int64_t* d_secret,
                           const uint8_t* d_rnd, const size_t* d_rnd_off, const void* d_tabGtB,
                           const uint8_t* d_baseB_wire, const uint8_t* d_ct, uint8_t* d_proofs) {
  if (!b.items) return 0;
  dim3 grid(grid_for(b.items * 16, 64), 3);  // sized for the sixteen-lane part; the other parts' spare blocks exit
  k_rp_create_lanes<<<grid, 64, 0, st>>>(b, d_sets, d_secret, (const uint32_t*)d_rnd, d_rnd_off, (const Fp12*)d_tabGtB,
                                         (const uint32_t*)d_baseB_wire, (const uint32_t*)d_ct, (uint32_t*)d_proofs);
  return 1;
}
int launch_rp_create_roles(cudaStream_t st, const RpBatch& b, const RpSetDev* d_sets, const int64_t* d_secret,
                           const uint8_t* d_rnd, const size_t* d_rnd_off, const void* d_tabGtB,
                           const uint8_t* d_baseB_wire, const uint8_t* d_ct, uint8_t* d_proofs, void* d_scratch) {
  if (!b.items) return 0;
  dim3 grid(grid_for(b.items, 64), 3);
  k_rp_create_roles<<<grid, 64, 0, st>>>(b, d_sets, d_secret, (const uint32_t*)d_rnd, d_rnd_off, (const Fp12*)d_tabGtB,
                                         (const uint32_t*)d_baseB_wire, (const uint32_t*)d_ct, (uint32_t*)d_proofs,
                                         (GtPair*)d_scratch);
  k_rp_create_finish<<<grid_for(b.items, 64), 64, 0, st>>>(b, d_sets, (const GtPair*)d_scratch, (uint32_t*)d_proofs);
  return 2;
}

int launch_rp_create_hdr(cudaStream_t st, const RpBatch& b, const RpSetDev* d_sets, const int64_t* d_secret,
                         const uint8_t* d_r, const uint8_t* d_ct, const uint8_t* d_rnd, const size_t* d_rnd_off,
                         const void* d_tabB, const void* d_tabP, const uint8_t* d_baseB_wire, uint8_t* d_proofs) {
  if (!b.n) return 0;
  if (b.n <= kEncryptLanesMaxItems) {
    k_rp_create_hdr_lanes<<<grid_for(b.n * 8, 64), 64, 0, st>>>(b, d_sets, d_secret, (const uint32_t*)d_r,
                                                               (const uint32_t*)d_ct, (const uint32_t*)d_rnd, d_rnd_off,
                                                               (const G1Aff*)d_tabB, (const G1Aff*)d_tabP,
                                                               (const uint32_t*)d_baseB_wire, (uint32_t*)d_proofs);
    return 1;
  }
  k_rp_create_hdr<<<grid_for(b.n, 128), 128, 0, st>>>(b, d_sets, d_secret, (const uint32_t*)d_r, (const uint32_t*)d_ct,
                                                     (const uint32_t*)d_rnd, d_rnd_off, (const G1Aff*)d_tabB,
                                                     (const G1Aff*)d_tabP, (const uint32_t*)d_baseB_wire,
                                                     (uint32_t*)d_proofs);
  return 1;
}

// ------------------------------------------------------------------ verification (a-8)
// Header check per proof: parse, D check (lib/range/range_proof.go:520-529, :561). flags[p] starts at the host's
// size-check verdict and is cleared on any failure. Two threads per proof: c*C is split by the curve's endomorphism into
// two products of half length (curve.cuh glv_split), one per thread, and each thread adds one of the comb walks Zr*P and
// (sum_j Zphi_j u^j)*B; a shuffle brings the halves together. The kernel runs beside the pairing kernel; as one thread
// per proof (3,550 M in a row) it outlasted that kernel at the headline size once it got below 4 ms; now its chain is
// 2,000 M. (Four threads per proof, 1,650 M, finish sooner but put twice the warps beside the pairing kernel, which
// then takes 4.19 instead of 3.94 ms: measured, not kept.)
__global__ void __launch_bounds__(128)
k_rp_verify_hdr(RpBatch b, const RpSetDev* __restrict__ sets, const G1Aff* __restrict__ tabB,
                const G1Aff* __restrict__ tabP, const uint32_t* __restrict__ proofs, uint32_t* __restrict__ flags) {
  const size_t p = ((size_t)blockIdx.x * blockDim.x + threadIdx.x) >> 1;
  const int half = threadIdx.x & 1;
  const uint32_t pair_mask = 3u << (threadIdx.x & 30u);
  if (p >= b.n) return;                                // both threads of a pair leave together, here and below
  int ns, nd;
  size_t off;
  if (!proof_set_valid(b, sets, p)) {                  // unknown signature set (device-buffer calls only): rejected
    if (half == 0) flags[p] = 0;
    return;
  }
  if (!locate_proof(b, sets, p, ns, nd, off)) return;  // u = l = 0: accepted
  // size check failed on the host? One thread reads the flag for both: the kernels running beside this one may clear
  // it at any moment, and the two threads must take the same way.
  const uint32_t seen = __shfl_sync(pair_mask, half == 0 ? flags[p] : 0u, (int)(threadIdx.x & 30u));
  if (!seen) return;
  const uint32_t* in = proofs + off / 4;
  const int64_t u = sets[b.d_set_id[p]].u;
  ProofView pv{ns, nd};
  G1Aff K, C, D;
  bool ok = g1_from_wire(K, in);
  ok &= g1_from_wire(C, in + 16);
  ok &= g1_from_wire(D, in + ProofView::d_off());
  if (!ok) {
    if (half == 0) flags[p] = 0;
    return;
  }
  G1Jac part = glv_half_mul(C, glv_split(sc_from_wire(in + ProofView::c_off())), half);
  if (half == 0) {
    part = jac_add(part, comb_mul(tabP, sc_from_wire(in + ProofView::zr_off())));
  } else {
    Scalar e = sc_zero();
    uint64_t upow = 1;
    for (int j = 0; j < nd; j++) {
      e = sc_add(e, sc_mul(sc_from_wire(in + pv.zphi_off(j)), sc_from_u64(upow)));
      upow *= (uint64_t)u;
    }
    part = jac_add(part, comb_mul(tabB, e));
  }
  G1Jac other;
#pragma unroll
  for (int k = 0; k < 8; k++) {
    other.x.v[k] = __shfl_xor_sync(pair_mask, part.x.v[k], 1);
    other.y.v[k] = __shfl_xor_sync(pair_mask, part.y.v[k], 1);
    other.z.v[k] = __shfl_xor_sync(pair_mask, part.z.v[k], 1);
  }
  if (half == 0 && !jac_eq_aff(jac_add(part, other), D)) flags[p] = 0;
}

#ifndef DRX_VERIFY_MINBLOCKS
#define DRX_VERIFY_MINBLOCKS 5
#endif
__global__ void k_flags_to_bytes(size_t n, const uint32_t* __restrict__ flags, uint8_t* __restrict__ ok) {
  const size_t p = (size_t)blockIdx.x * blockDim.x + threadIdx.x;
  if (p < n) ok[p] = flags[p] ? 1 : 0;
}

int launch_rp_verify_hdr(cudaStream_t st, const RpBatch& b, const RpSetDev* d_sets, const void* d_tabB,
                         const void* d_tabP, const uint8_t* d_proofs, uint32_t* d_flags) {
  if (!b.n) return 0;
  // one warp per block: these kernels run next to the lane-group pairing, whose pace is set by its most loaded SM;
  // small blocks spread the extra warps over many SMs instead of stacking four on a few
  static bool carveout_set[64] = {false};
  prefer_shared_carveout(k_rp_verify_hdr, carveout_set);
  k_rp_verify_hdr<<<grid_for(b.n * 2, 32), 32, 0, st>>>(b, d_sets, (const G1Aff*)d_tabB, (const G1Aff*)d_tabP,
                                                     (const uint32_t*)d_proofs, d_flags);
  return 1;
}
// Role-split verification: blockIdx.y = role.
__global__ void __launch_bounds__(64, DRX_VERIFY_MINBLOCKS)
k_rp_verify_roles(RpBatch b, const RpSetDev* __restrict__ sets, const G1Aff* __restrict__ tabB,
                  const Fp12* __restrict__ tabGtB, const uint32_t* __restrict__ proofs, uint32_t* __restrict__ flags,
                  GtPair* __restrict__ scratch) {
  const size_t t = (size_t)blockIdx.x * blockDim.x + threadIdx.x;
  if (t >= b.items) return;
  rp_verify_role(b, sets, tabB, tabGtB, proofs, flags, scratch, t, (int)blockIdx.y);
}
__global__ void __launch_bounds__(128)
k_rp_verify_finish(RpBatch b, const RpSetDev* __restrict__ sets, const GtPair* __restrict__ scratch,
                   uint32_t* __restrict__ flags) {
  const size_t t = (size_t)blockIdx.x * blockDim.x + threadIdx.x;
  if (t >= b.items) return;
  rp_verify_finish(b, sets, scratch, flags, t);
}
int launch_rp_verify_roles(cudaStream_t st, const RpBatch& b, const RpSetDev* d_sets, const void* d_tabB,
                           const void* d_tabGtB, const uint8_t* d_proofs, uint32_t* d_flags, void* d_scratch) {
  if (!b.items) return 0;
  dim3 grid(grid_for(b.items, 64), 2);
  k_rp_verify_roles<<<grid, 64, 0, st>>>(b, d_sets, (const G1Aff*)d_tabB, (const Fp12*)d_tabGtB,
                                         (const uint32_t*)d_proofs, d_flags, (GtPair*)d_scratch);
  k_rp_verify_finish<<<grid_for(b.items, 128), 128, 0, st>>>(b, d_sets, (const GtPair*)d_scratch, d_flags);
  return 2;
}

// The GT factor alone (role 1) and the final comparison, used when the pairing runs in the cooperative kernel.
__global__ void __launch_bounds__(64)
k_rp_verify_gt(RpBatch b, const RpSetDev* __restrict__ sets, const G1Aff* __restrict__ tabB,
               const Fp12* __restrict__ tabGtB, const Fp12* __restrict__ tabGtB16, const uint32_t* __restrict__ proofs,
               uint32_t* __restrict__ flags, GtPair* __restrict__ scratch) {
  const size_t t = (size_t)blockIdx.x * blockDim.x + threadIdx.x;
  if (t >= b.items) return;
  if (!tabGtB16) {
    rp_verify_role(b, sets, tabB, tabGtB, proofs, flags, scratch, t, 1);
    return;
  }
  // role 1 of rp_verify_role with the shared base walked in 16-bit windows: T = A_ij * e(B, g2)^(-Zv_ij)
  size_t p, off;
  int i, j, ns, nd;
  if (!locate_item(b, sets, t, p, i, j, ns, nd, off)) return;
  const uint32_t* in = proofs + off / 4;
  ProofView pv{ns, nd};
  Fp12 A, g;
  if (!gt_from_wire(A, in + pv.a_off(i, j))) {
    atomicAnd(&flags[p], 0u);
    scratch[t].b = fp12_one();
    return;
  }
  gt_comb_pow16(g, tabGtB16, sc_neg(sc_from_wire(in + pv.zv_off(i, j))));
  fp12_mul(scratch[t].b, A, g);
}
int launch_rp_verify_gt(cudaStream_t st, const RpBatch& b, const RpSetDev* d_sets, const void* d_tabB,
                        const void* d_tabGtB, const void* d_tabGtB16, const uint8_t* d_proofs, uint32_t* d_flags,
                        void* d_scratch) {
  if (!b.items) return 0;
  static bool carveout_set[64] = {false};
  prefer_shared_carveout(k_rp_verify_gt, carveout_set);
  k_rp_verify_gt<<<grid_for(b.items, 32), 32, 0, st>>>(b, d_sets, (const G1Aff*)d_tabB, (const Fp12*)d_tabGtB,
                                                      (const Fp12*)d_tabGtB16, (const uint32_t*)d_proofs, d_flags,
                                                      (GtPair*)d_scratch);
  return 1;
}
int launch_rp_verify_finish(cudaStream_t st, const RpBatch& b, const RpSetDev* d_sets, const void* d_scratch,
                            uint32_t* d_flags) {
  if (!b.items) return 0;
  k_rp_verify_finish<<<grid_for(b.items, 128), 128, 0, st>>>(b, d_sets, (const GtPair*)d_scratch, d_flags);
  return 1;
}

int launch_rp_flags(cudaStream_t st, size_t n, const uint32_t* d_flags, uint8_t* d_ok) {
  if (!n) return 0;
  k_flags_to_bytes<<<grid_for(n, 256), 256, 0, st>>>(n, d_flags, d_ok);
  return 1;
}

}  // namespace drx
