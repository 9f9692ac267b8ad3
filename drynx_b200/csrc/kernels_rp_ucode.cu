// Preparation of the lane-group pairing's arguments (range-proof verification, lib/range/range_proof.go:540-544):
// per (proof, server, digit) item, parse V_ij and build the G1 argument c*Y_i - Zphi_j*B. The pairing itself runs in
// kernels_rp_ucode_s4.cu / _s8.cu (ucode_split.cuh).
#include <cstdio>
#include <cstdlib>

#include "ucode_common.cuh"

namespace drx {

static inline int grid_for(size_t n, int threads) { return uc_grid_for(n, threads); }

// ---- four lanes per item: parse V_ij and build the G1 argument c*Y_i - Zphi_j*B. Each lane walks every fourth
// window of the two comb tables (16 mixed additions instead of 64), a two-level butterfly of full additions
// leaves the sum in all four lanes (same operand order on both partners, so the copies are bit-identical), and
// lane 0 writes the affine result.
__device__ __forceinline__ G1Jac prep_shfl_xor(const G1Jac& p, int m, uint32_t mask) {
  G1Jac r;
#pragma unroll
  for (int k = 0; k < 8; k++) {
    r.x.v[k] = __shfl_xor_sync(mask, p.x.v[k], m);
    r.y.v[k] = __shfl_xor_sync(mask, p.y.v[k], m);
    r.z.v[k] = __shfl_xor_sync(mask, p.z.v[k], m);
  }
  return r;
}
__global__ void __launch_bounds__(64)
k_rp_verify_prep(RpBatch b, const RpSetDev* __restrict__ sets, const G1Aff* __restrict__ tabB,
                 const uint32_t* __restrict__ proofs, uint32_t* __restrict__ flags, UcPrep* __restrict__ prep) {
  const size_t t = ((size_t)blockIdx.x * blockDim.x + threadIdx.x) >> 2;
  const int lane = threadIdx.x & 3;
  const uint32_t mask = 0xFu << (threadIdx.x & 28u);
  if (t >= b.items) return;  // whole quads leave together
  size_t p, off;
  int i, j, ns, nd;
  if (!locate_item(b, sets, t, p, i, j, ns, nd, off)) return;
  const uint32_t* in = proofs + off / 4;
  const RpSetDev set = sets[b.d_set_id[p]];
  ProofView pv{ns, nd};
  const Scalar c = sc_from_wire(in + ProofView::c_off());
  const Scalar nz = sc_neg(sc_from_wire(in + pv.zphi_off(j)));
  const G1Aff* tabY = (const G1Aff*)set.tabY + (size_t)i * (DRX_COMB_WINDOWS * DRX_COMB_ENTRIES);
  G1Jac acc = jac_inf<Fp>();
  for (int w = lane; w < DRX_COMB_WINDOWS; w += 4) {
    const uint32_t d0 = (c.v[w >> 2] >> ((w & 3) * 8)) & 0xffu;
    if (d0) acc = jac_add_aff(acc, tabY[w * DRX_COMB_ENTRIES + (int)d0 - 1]);
    const uint32_t d1 = (nz.v[w >> 2] >> ((w & 3) * 8)) & 0xffu;
    if (d1) acc = jac_add_aff(acc, tabB[w * DRX_COMB_ENTRIES + (int)d1 - 1]);
  }
#pragma unroll
  for (int m = 1; m <= 2; m <<= 1) {
    const G1Jac o = prep_shfl_xor(acc, m, mask);
    acc = (lane & m) ? jac_add(o, acc) : jac_add(acc, o);
  }
  if (lane != 0) return;
  G2Aff V;
  if (!g2_from_wire(V, in + pv.v_off(i, j))) {
    atomicAnd(&flags[p], 0u);
    prep[t].state = 2;
    return;
  }
  const G1Aff P = jac_to_aff(acc);
  prep[t].P = P;
  prep[t].V = V;
  uint32_t state = (aff_is_inf(P) | aff_is_inf(V)) ? 1u : 0u;
  // the pairing kernel decides G2 membership of V inside its Miller loop; an item that skips the pairing because its
  // G1 argument is the point at infinity (crafted input only) needs the verdict from here
  if (state == 1u && !aff_is_inf(V) && !g2_in_subgroup(V)) {
    atomicAnd(&flags[p], 0u);
    state = 2u;
  }
  prep[t].state = state;
}



// prepared arguments, then (dynamic launches only) the control words and the groups' slot images
size_t rp_ucode_prep_bytes(size_t items) {
  return uc_align256(items * sizeof(UcPrep)) + uc_dyn_ctrl_bytes() + uc_dyn_state_bytes(items);
}

int launch_rp_verify_prep(cudaStream_t st, const RpBatch& b, const RpSetDev* d_sets, const void* d_tabB,
                          const uint8_t* d_proofs, uint32_t* d_flags, void* d_prep) {
  if (!b.items) return 0;
  // The header check may still be running on some SMs when the pairing kernel arrives; an SM keeps the L1 / shared split
  // of the first block that reached it while it was empty, so this kernel too asks for the shared-heavy split (it uses no
  // shared memory itself). Per device, once; a hint, errors ignored.
  {
    static bool done[kUcMaxDevices] = {false};
    int dev = 0;
    if (cudaGetDevice(&dev) == cudaSuccess && dev >= 0 && dev < kUcMaxDevices) {
      std::lock_guard<std::mutex> lk(uc_dev_mutex());
      if (!done[dev]) {
        if (cudaFuncSetAttribute(k_rp_verify_prep, cudaFuncAttributePreferredSharedMemoryCarveout,
                                 (int)cudaSharedmemCarveoutMaxShared) != cudaSuccess)
          cudaGetLastError();
        done[dev] = true;
      }
    }
  }
  k_rp_verify_prep<<<grid_for(b.items * 4, 64), 64, 0, st>>>(b, d_sets, (const G1Aff*)d_tabB, (const uint32_t*)d_proofs,
                                                        d_flags, (UcPrep*)d_prep);
  return 1;
}

}  // namespace drx
