// Fixed-base GT exponentiations of the range proofs on lane groups: the GT factors of proof creation
// (a_ij = e(B, A_{i,phi})^(-s_j v_ij) * e(B, g2)^(t_j), lib/range/range_proof.go:396-397) and of verification
// (T_ij = A_ij * e(B, g2)^(-Zv_ij), the right-hand side of :540-546). One thread per factor walks 32 comb windows with
// 31 dependent Fp12 products of ~50,000 cycles each; here a group of eight threads (four program lanes of two threads,
// the interpreter of ucode_interp.cuh) runs each product as a 24-phase micro-program over slots in shared memory
// (tools/microcode/gen.py gtmul), about 12,000 cycles, and the running value never leaves shared memory.
#include <cstdio>
#include <cstdlib>

#define DRX_FP2_INLINE_MUL 1
#define UC_PROGRAM_INC "gtmul_ucode_g4.inc"
#include "ucode_common.cuh"

namespace drx {

#include "ucode_interp.cuh"

static_assert(UC_G == 4, "eight threads per group");
constexpr int kGtGroupsPerWarp = 4;
#define GTL_BLOCK 64
#define GTL_GROUPS_PER_BLOCK ((GTL_BLOCK / 32) * kGtGroupsPerWarp)

__device__ __forceinline__ uint32_t gtl_slot_a(int k) {
  return k == 0 ? UC_IN_A0 : k == 1 ? UC_IN_A1 : k == 2 ? UC_IN_A2 : k == 3 ? UC_IN_A3 : k == 4 ? UC_IN_A4 : UC_IN_A5;
}
__device__ __forceinline__ uint32_t gtl_slot_b(int k) {
  return k == 0 ? UC_IN_B0 : k == 1 ? UC_IN_B1 : k == 2 ? UC_IN_B2 : k == 3 ? UC_IN_B3 : k == 4 ? UC_IN_B4 : UC_IN_B5;
}
// An Fp12 in memory is six Fp2 of 64 bytes in tower order, i.e. 24 chunks of 16 bytes; chunk q belongs to Fp2 q / 4,
// logical chunk q % 4. The eight threads of a group move three chunks each (128 contiguous bytes per step).
__device__ __forceinline__ void gtl_load_entry(uint32_t S, int sub, const Fp12* __restrict__ e, bool into_b) {
  const uint4* src = reinterpret_cast<const uint4*>(e);
#pragma unroll
  for (int r = 0; r < 3; r++) {
    const int q = sub + 8 * r, k = q >> 2, c = q & 3;
    const uint4 v = __ldg(src + q);
    const uint32_t w[4] = {v.x, v.y, v.z, v.w};
    ucs_sts((S + ucs_slot_off(into_b ? gtl_slot_b(k) : gtl_slot_a(k))) ^ ((uint32_t)c << 4), w);
  }
}
// A slots of the group at S_from -> B slots of the group at S (shared to shared)
__device__ __forceinline__ void gtl_copy_a_to_b(uint32_t S, uint32_t S_from, int sub) {
#pragma unroll
  for (int r = 0; r < 3; r++) {
    const int q = sub + 8 * r, k = q >> 2, c = q & 3;
    const uint4 v = ucs_lds((S_from + ucs_slot_off(gtl_slot_a(k))) ^ ((uint32_t)c << 4));
    const uint32_t w[4] = {v.x, v.y, v.z, v.w};
    ucs_sts((S + ucs_slot_off(gtl_slot_b(k))) ^ ((uint32_t)c << 4), w);
  }
}
__device__ __forceinline__ void gtl_store_acc(uint32_t S, int sub, Fp12* __restrict__ out) {
  uint4* dst = reinterpret_cast<uint4*>(out);
#pragma unroll
  for (int r = 0; r < 3; r++) {
    const int q = sub + 8 * r, k = q >> 2, c = q & 3;
    dst[q] = ucs_lds((S + ucs_slot_off(gtl_slot_a(k))) ^ ((uint32_t)c << 4));
  }
}
__device__ __forceinline__ void gtl_set_one(uint32_t S, int sub) {
  if (sub < 6) ucs_store(S, ucs_slot_off(gtl_slot_a(sub)), sub == 0 ? fp2_one() : fp2_zero());
}
// One comb step of a group: acc (the group's A slots) <- acc * *entry, or acc <- *entry while the group holds no value
// yet (`have` false); a null entry (zero digit, or a group that is not live) skips the step. The whole warp calls this
// together: groups without work keep the others company at the barriers. Returns the new `have`.
__device__ __forceinline__ bool gtl_step(uint32_t S, int sub, const Fp12* __restrict__ entry, bool have) {
  const bool mul = entry != nullptr && have;
  if (entry) {
    gtl_load_entry(S, sub, entry, have);
    have = true;
  }
  __syncwarp();
  if (__any_sync(0xffffffffu, mul)) ucs_run(S, sub >> 1, (uint32_t)(sub & 1), mul, 0, UC_NPHASES);
  return have;
}
// entry of an 8-bit-window table (32 windows x 255 entries) / of the 16-bit-window table of e(B, g2) (16 x 65,535) for
// window w of scalar k; null for a zero digit
__device__ __forceinline__ const Fp12* gtl_entry8(const Fp12* __restrict__ table, const Scalar& k, int w) {
  const uint32_t d = (k.v[w >> 2] >> ((w & 3) * 8)) & 0xffu;
  return d ? table + (size_t)w * DRX_COMB_ENTRIES + d - 1 : nullptr;
}
__device__ __forceinline__ const Fp12* gtl_entry16(const Fp12* __restrict__ table16, const Scalar& k, int w) {
  const uint32_t d = (k.v[w >> 1] >> ((w & 1) * 16)) & 0xffffu;
  return d ? table16 + (size_t)w * kGt16PerWindow + d - 1 : nullptr;
}

// ------------------------------------------------------------------ creation: a_ij, sixteen threads per item
// Two groups of eight threads share the 32 + 16 products of an item evenly: the even group walks windows 0..23 of
// e(B, A_{i,phi})^(-s v) (8-bit windows, per-set table), the odd group the 16 windows of e(B, g2)^t (16-bit windows) and
// then windows 24..31 of the first factor; one more product joins the halves.
__global__ void __launch_bounds__(GTL_BLOCK, 8)
k_rp_create_gt_lanes(RpBatch b, const RpSetDev* __restrict__ sets, const int64_t* __restrict__ secret,
                     const uint32_t* __restrict__ rnd, const size_t* __restrict__ rnd_off, const Fp12* __restrict__ tabGtB16,
                     uint32_t* __restrict__ proofs) {
  extern __shared__ __align__(128) unsigned char uc_smem[];
  const int wl = threadIdx.x & 31, warp = threadIdx.x >> 5, grp = wl >> 3, sub = wl & 7;
  const uint32_t S = (uint32_t)__cvta_generic_to_shared(uc_smem) + (warp * kGtGroupsPerWarp + grp) * UCS_ITEM_BYTES;
  // groups 0, 1 of a warp serve one item, groups 2, 3 the next
  const size_t t = ((size_t)blockIdx.x * (GTL_BLOCK / 32) + warp) * 2 + (grp >> 1);
  const bool even = (grp & 1) == 0;
  size_t p = 0, off = 0;
  int i = 0, j = 0, ns = 0, nd = 0;
  const bool live = t < b.items && locate_item(b, sets, t, p, i, j, ns, nd, off);
  Scalar ka = sc_zero(), kt = sc_zero();
  const Fp12* tableA = nullptr;
  if (live) {
    const RpSetDev set = sets[b.d_set_id[p]];
    const uint32_t* rp = rnd + (b.d_proof_off ? rnd_off[p] / 4 : p * (size_t)nd * (3 + ns) * 8);
    const uint32_t* row = rp + (size_t)j * (3 + ns) * 8;
    const uint32_t phi = rp_digit(secret[p], set.u, j);
    tableA = (const Fp12*)set.tabGtA + ((size_t)i * set.u + phi) * (DRX_COMB_WINDOWS * DRX_COMB_ENTRIES);
    ka = sc_neg(sc_mul(sc_from_wire(row), sc_from_wire(row + 8 * (3 + i))));  // -s_j v_ij
    kt = sc_from_wire(row + 8);                                              // t_j
  }
  bool have = false;
  for (int s = 0; s < 24; s++) {
    const Fp12* e = nullptr;
    if (live) e = even ? gtl_entry8(tableA, ka, s) : s < 16 ? gtl_entry16(tabGtB16, kt, s) : gtl_entry8(tableA, ka, s + 8);
    have = gtl_step(S, sub, e, have);
  }
  if (live && !have) gtl_set_one(S, sub);  // exponent 0
  __syncwarp();
  // a_ij = the even group's partial product * the odd group's: the even group takes the odd group's value as its B operand
  if (live && even) gtl_copy_a_to_b(S, S + UCS_ITEM_BYTES, sub);
  __syncwarp();
  ucs_run(S, sub >> 1, (uint32_t)(sub & 1), live && even, 0, UC_NPHASES);
  // serialise: 12 field elements, one per thread of the item's sixteen (wire order x.x .. y.z, i-coefficient first)
  if (live) {
    const uint32_t S0 = even ? S : S - UCS_ITEM_BYTES;
    const int e = (even ? 0 : 8) + sub;  // 0 .. 15; elements 0 .. 11 exist
    if (e < 12) {
      const int s = e >> 1;               // wire slot: x.x x.y x.z y.x y.y y.z = tower c1.c2 c1.c1 c1.c0 c0.c2 c0.c1 c0.c0
      const int tower = s == 0 ? 5 : s == 1 ? 4 : s == 2 ? 3 : s == 3 ? 2 : s == 4 ? 1 : 0;
      const Fp v = ucs_load_c(S0, ucs_slot_off(gtl_slot_a(tower)), (e & 1) ? 0u : 1u);
      ProofView pv{ns, nd};
      fp_to_wire(proofs + off / 4 + pv.a_off(i, j) + 8 * e, v);
    }
  }
}

// ------------------------------------------------------------------ verification: T_ij, eight threads per item
__global__ void __launch_bounds__(GTL_BLOCK, 8)
k_rp_verify_gt_lanes(RpBatch b, const RpSetDev* __restrict__ sets, const Fp12* __restrict__ tabGtB16,
                     const uint32_t* __restrict__ proofs, uint32_t* __restrict__ flags, GtPair* __restrict__ scratch) {
  extern __shared__ __align__(128) unsigned char uc_smem[];
  const int wl = threadIdx.x & 31, warp = threadIdx.x >> 5, grp = wl >> 3, sub = wl & 7;
  const uint32_t S = (uint32_t)__cvta_generic_to_shared(uc_smem) + (warp * kGtGroupsPerWarp + grp) * UCS_ITEM_BYTES;
  const size_t t = ((size_t)blockIdx.x * (GTL_BLOCK / 32) + warp) * kGtGroupsPerWarp + grp;
  size_t p = 0, off = 0;
  int i = 0, j = 0, ns = 0, nd = 0;
  bool live = t < b.items && locate_item(b, sets, t, p, i, j, ns, nd, off);
  const bool present = live;  // the item exists (its scratch entry must be written whatever happens)
  Scalar k = sc_zero();
  bool parsed = true;
  if (live) {
    const uint32_t* in = proofs + off / 4;
    ProofView pv{ns, nd};
    // A_ij from the wire: threads 0 .. 5 parse one Fp2 each (wire slot s = tower slot 5 - s)
    if (sub < 6) {
      Fp2 v;
      parsed = fp2_from_wire(v, in + pv.a_off(i, j) + 16 * sub);
      ucs_store(S, ucs_slot_off(gtl_slot_a(5 - sub)), v);
    }
    k = sc_neg(sc_from_wire(in + pv.zv_off(i, j)));
  }
  // every thread votes (threads with nothing to parse say yes); a group's verdict is its eight bits
  const uint32_t vote = __ballot_sync(0xffffffffu, parsed);
  if (live && ((vote >> (wl & 24)) & 0xffu) != 0xffu) {  // a coordinate >= p: malformed for kyber, the proof is rejected
    if (sub == 0) atomicAnd(&flags[p], 0u);
    live = false;
  }
  __syncwarp();
  for (int w = 0; w < 16; w++) gtl_step(S, sub, live ? gtl_entry16(tabGtB16, k, w) : nullptr, true);  // A_ij * e(B, g2)^(-Zv)
  __syncwarp();
  if (present) {
    if (live) {
      gtl_store_acc(S, sub, &scratch[t].b);
    } else if (sub == 0) {
      scratch[t].b = fp12_one();
    }
  }
}

static bool gtl_configure(bool (&done)[kUcMaxDevices], const void* kernel, int smem) {
  int dev = 0;
  if (cudaGetDevice(&dev) != cudaSuccess || dev < 0 || dev >= kUcMaxDevices) return false;
  std::lock_guard<std::mutex> lk(uc_dev_mutex());
  if (done[dev]) return true;
  if (cudaFuncSetAttribute(kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, smem) != cudaSuccess) return false;
  done[dev] = true;
  return true;
}
constexpr int kGtlSmem = GTL_GROUPS_PER_BLOCK * UC_NSLOTS * 64;

int launch_rp_create_gt_lanes(cudaStream_t st, const RpBatch& b, const RpSetDev* d_sets, const int64_t* d_secret,
                              const uint8_t* d_rnd, const size_t* d_rnd_off, const void* d_tabGtB16, uint8_t* d_proofs) {
  if (!b.items) return 0;
  static bool done[kUcMaxDevices] = {false};
  if (!gtl_configure(done, (const void*)k_rp_create_gt_lanes, kGtlSmem)) return 0;
  const size_t items_per_block = (GTL_BLOCK / 32) * 2;
  k_rp_create_gt_lanes<<<uc_grid_for(b.items, (int)items_per_block), GTL_BLOCK, kGtlSmem, st>>>(
      b, d_sets, d_secret, (const uint32_t*)d_rnd, d_rnd_off, (const Fp12*)d_tabGtB16, (uint32_t*)d_proofs);
  return 1;
}
int launch_rp_verify_gt_lanes(cudaStream_t st, const RpBatch& b, const RpSetDev* d_sets, const void* d_tabGtB16,
                              const uint8_t* d_proofs, uint32_t* d_flags, void* d_scratch) {
  if (!b.items) return 0;
  static bool done[kUcMaxDevices] = {false};
  if (!gtl_configure(done, (const void*)k_rp_verify_gt_lanes, kGtlSmem)) return 0;
  k_rp_verify_gt_lanes<<<uc_grid_for(b.items, GTL_GROUPS_PER_BLOCK), GTL_BLOCK, kGtlSmem, st>>>(
      b, d_sets, (const Fp12*)d_tabGtB16, (const uint32_t*)d_proofs, d_flags, (GtPair*)d_scratch);
  return 1;
}

}  // namespace drx
