// BN256 pairing-based range proofs on sm_100a: creation and verification, plus the G2 / GT set-up kernels.
// Replaces lib/range/range_proof.go:249-288 (signatures), :320-407 (CreatePredicateRangeProofForAllServ) and
// :504-565 (RangeProofVerification) of the reference, and the kyber pairing/bn256 calls they make.
#include "launch.h"
#include "keccak.cuh"
#include "wire.cuh"

namespace drx {

static inline int grid_for(size_t n, int threads) { return (int)((n + threads - 1) / threads); }


// ------------------------------------------------------------------ set-up
__global__ void k_g2_from_wire(size_t n, const uint32_t* __restrict__ wire, G2Aff* __restrict__ aff, int* bad) {
  const size_t i = (size_t)blockIdx.x * blockDim.x + threadIdx.x;
  if (i >= n) return;
  G2Aff p;
  if (!g2_from_wire(p, wire + 32 * i)) atomicAdd(bad, 1);
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

// ------------------------------------------------------------------ diagnostics on wire bytes
__global__ void __launch_bounds__(64)
k_pair_wire(size_t n, const uint32_t* __restrict__ g1, const uint32_t* __restrict__ g2, uint32_t* __restrict__ gt,
            int* bad) {
  const size_t i = (size_t)blockIdx.x * blockDim.x + threadIdx.x;
  if (i >= n) return;
  G1Aff P;
  G2Aff Q;
  int nbad = !g1_from_wire(P, g1 + 16 * i);
  nbad += !g2_from_wire(Q, g2 + 32 * i);
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
  if (!g2_from_wire(p, pt + 32 * i)) atomicAdd(bad, 1);
  const G2Jac r = window_mul(jac_from_aff(p), sc_from_wire(k + 8 * i));
  g2_to_wire(out + 32 * i, jac_to_aff(r));
}
int launch_g2_mul(cudaStream_t st, size_t n, const uint8_t* d_k, const uint8_t* d_pt, uint8_t* d_out, int* d_bad) {
  if (!n) return 0;
  k_g2_mul<<<grid_for(n, 64), 64, 0, st>>>(n, (const uint32_t*)d_k, (const uint32_t*)d_pt, (uint32_t*)d_out, d_bad);
  return 1;
}

}  // namespace drx
