// Small-batch range-proof kernels: four lanes per (proof, server, digit) item.
// With a few thousand items (BASELINE config 2: 500 proofs -> 6,000 pairing checks) one thread per item leaves the
// 592 SM sub-partitions of a B200 with less than one warp each and the kernel is pure latency. Here an aligned
// quad of lanes shares one item (tower.cuh, DRX_COOP4): Fp2 products are split four ways, the rest is replicated.
// Same item bodies, same results as kernels_rp.cu; chosen by the C ABI when items <= kCoopMaxItems.
#define DRX_COOP4 1
#define DRX_FP2_INLINE_MUL 1
#include "rp_device.cuh"

namespace drx {

static inline int grid_for(size_t n, int threads) { return (int)((n + threads - 1) / threads); }

__global__ void __launch_bounds__(64)
k_rp_verify_items_coop4(RpBatch b, const RpSetDev* __restrict__ sets, const G1Aff* __restrict__ tabB,
                        const Fp12* __restrict__ tabGtB, const uint32_t* __restrict__ proofs,
                        uint32_t* __restrict__ flags) {
  const size_t t = ((size_t)blockIdx.x * blockDim.x + threadIdx.x) >> 2;
  if (t >= b.items) return;  // a whole quad leaves together
  rp_verify_item(b, sets, tabB, tabGtB, proofs, flags, t);
}

__global__ void __launch_bounds__(64)
k_rp_create_items_coop4(RpBatch b, const RpSetDev* __restrict__ sets, const int64_t* __restrict__ secret,
                        const uint32_t* __restrict__ rnd, const size_t* __restrict__ rnd_off,
                        const Fp12* __restrict__ tabGtB, const uint32_t* __restrict__ baseB_wire,
                        const uint32_t* __restrict__ ct, uint32_t* __restrict__ proofs) {
  const size_t t = ((size_t)blockIdx.x * blockDim.x + threadIdx.x) >> 2;
  if (t >= b.items) return;
  rp_create_item(b, sets, secret, rnd, rnd_off, tabGtB, baseB_wire, ct, proofs, t);  // the four lanes store identical bytes
}

int launch_rp_verify_items_coop4(cudaStream_t st, const RpBatch& b, const RpSetDev* d_sets, const void* d_tabB,
                                 const void* d_tabGtB, const uint8_t* d_proofs, uint32_t* d_flags) {
  if (!b.items) return 0;
  k_rp_verify_items_coop4<<<grid_for(b.items * 4, 64), 64, 0, st>>>(b, d_sets, (const G1Aff*)d_tabB,
                                                                   (const Fp12*)d_tabGtB, (const uint32_t*)d_proofs,
                                                                   d_flags);
  return 1;
}
int launch_rp_create_items_coop4(cudaStream_t st, const RpBatch& b, const RpSetDev* d_sets, const int64_t* d_secret,
                                 const uint8_t* d_rnd, const size_t* d_rnd_off, const void* d_tabGtB,
                                 const uint8_t* d_baseB_wire, const uint8_t* d_ct, uint8_t* d_proofs) {
  if (!b.items) return 0;
  k_rp_create_items_coop4<<<grid_for(b.items * 4, 64), 64, 0, st>>>(b, d_sets, d_secret, (const uint32_t*)d_rnd,
                                                                   d_rnd_off, (const Fp12*)d_tabGtB,
                                                                   (const uint32_t*)d_baseB_wire, (const uint32_t*)d_ct,
                                                                   (uint32_t*)d_proofs);
  return 1;
}

}  // namespace drx
