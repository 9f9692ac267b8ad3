// Lane-group pairing, two threads per program lane. One copy of a pairing check's state lives in shared memory as
// UC_NSLOTS Fp2 slots, and a generated micro-program (pairing_ucode_g*.inc, tools/microcode/gen.py) runs over it: every
// phase the program lanes execute one Fp2 operation each on their own slots. Each operation of a program lane is carried
// out by a PAIR of threads, thread h of the pair producing component h of the result: an Fp2 product is one
// dual-product Montgomery pass per thread (c0 = a0 b0 + a1 (p - b1) on h = 0, c1 = a0 b1 + a1 b0 on h = 1), a squaring
// one product per thread, a sum one Fp addition per thread. Both threads read whatever operand components they need
// straight from the item's slots, so there are no shuffles and no cross-thread carries. (Round 1 gave a program lane
// one thread; two halve the dependent instruction chain of every phase and double the resident warps per item for the
// same shared-memory footprint.) The interpreter itself is ucode_interp.cuh; this file holds the two kernels and their
// launcher. Replaces kyber pairing/bn256 optate.go for lib/range/range_proof.go:540-544.
#pragma once
#include <cstdio>
#include <cstdlib>

#define DRX_FP2_INLINE_MUL 1
#include "ucode_common.cuh"

namespace drx {

#include "ucode_interp.cuh"

__global__ void __launch_bounds__(UCS_BLOCK, UCS_MIN_BLOCKS)
UCS_KERNEL(size_t items, const UcPrep* __restrict__ prep, GtPair* __restrict__ scratch) {
  extern __shared__ __align__(128) unsigned char uc_smem[];
  const int wl = threadIdx.x & 31, warp = threadIdx.x >> 5;
  const int in_warp = wl / UCS_TPI, sub = wl % UCS_TPI;
  const bool seated = in_warp < UCS_WARP_ITEMS;            // false for the spare threads of a six-thread layout
  const int grp = warp * UCS_WARP_ITEMS + (seated ? in_warp : 0);
  const size_t t = (size_t)blockIdx.x * UCS_ITEMS_PER_BLOCK + grp;
  const bool alive = seated && t < items && prep[t].state == 0;
  const uint32_t S = (uint32_t)__cvta_generic_to_shared(uc_smem) + grp * UCS_ITEM_BYTES;
  if (alive && sub == 0) ucs_put_inputs(S, prep[t].P, prep[t].V);
  __syncwarp();
  ucs_run(S, sub >> 1, (uint32_t)(sub & 1), alive, 0, UC_NPHASES);
  if (seated && t < items && sub == 0) scratch[t].a = alive ? ucs_get_output(S) : fp12_one();
}

// Dynamic variant: the program is cut into nseg segments, a task is
// (warp-sized group of items, segment), persistent warps take tasks from an atomic counter segment-major, and a
// group's slot image travels through global memory between segments.
#define UCS_GROUP_U4 (UCS_WARP_ITEMS * UC_NSLOTS * 4)  // uint4 words of one group's slots
__global__ void __launch_bounds__(UCS_BLOCK, UCS_MIN_BLOCKS)
UCS_KERNEL_DYN(size_t items, const UcPrep* __restrict__ prep, GtPair* __restrict__ scratch, uint4* __restrict__ state,
               unsigned int* __restrict__ ctrl, unsigned int nseg, unsigned int* __restrict__ err) {
  extern __shared__ __align__(128) unsigned char uc_smem[];
  const int wl = threadIdx.x & 31, warp = threadIdx.x >> 5;
  const int in_warp = wl / UCS_TPI, sub = wl % UCS_TPI;
  const bool seated = in_warp < UCS_WARP_ITEMS;
  const int grp = warp * UCS_WARP_ITEMS + (seated ? in_warp : 0);
  const uint32_t S = (uint32_t)__cvta_generic_to_shared(uc_smem) + grp * UCS_ITEM_BYTES;
  uint4* W = reinterpret_cast<uint4*>(uc_smem + warp * UCS_WARP_ITEMS * UCS_ITEM_BYTES);
  const unsigned int ngroups = (unsigned int)((items + UCS_WARP_ITEMS - 1) / UCS_WARP_ITEMS);
  const unsigned int ntasks = ngroups * nseg;
  volatile unsigned int* done = ctrl + 4;
  for (;;) {
    unsigned int task = 0;
    if (wl == 0) task = atomicAdd(&ctrl[0], 1u);
    task = __shfl_sync(0xffffffffu, task, 0);
    if (task >= ntasks) break;
    const unsigned int g = task % ngroups, seg = task / ngroups;
    const size_t t = (size_t)g * UCS_WARP_ITEMS + (seated ? in_warp : 0);
    const bool alive = seated && t < items && prep[t].state == 0;
    if (seg == 0) {
      if (alive && sub == 0) ucs_put_inputs(S, prep[t].P, prep[t].V);
    } else {
      unsigned int polls = 0;
      if (wl == 0) {
        while (done[g] < seg && polls < kUcMaxPolls) {
          __nanosleep(200);
          polls++;
        }
        if (polls >= kUcMaxPolls) atomicOr(err, 1u);
      }
      polls = __shfl_sync(0xffffffffu, polls, 0);
      if (polls >= kUcMaxPolls) break;
      __threadfence();
      const uint4* src = state + (size_t)g * UCS_GROUP_U4;
      for (int i = wl; i < UCS_GROUP_U4; i += 32) W[i] = __ldcg(src + i);
    }
    __syncwarp();
    ucs_run(S, sub >> 1, (uint32_t)(sub & 1), alive, (int)(seg * UC_NPHASES / nseg), (int)((seg + 1) * UC_NPHASES / nseg));
    if (seg + 1 < nseg) {
      uint4* dst = state + (size_t)g * UCS_GROUP_U4;
      for (int i = wl; i < UCS_GROUP_U4; i += 32) __stcg(dst + i, W[i]);
      __threadfence();
      __syncwarp();
      if (wl == 0) done[g] = seg + 1;
    } else if (seated && t < items && sub == 0) {
      scratch[t].a = alive ? ucs_get_output(S) : fp12_one();
    }
    __syncwarp();
  }
}

// tune.mode: 0 = by batch size, 1 = static launch, 2 = dynamic launch; tune.resident: persistent blocks per SM of the
// dynamic launch (0 = as many as fit); tune.nseg: program segments of the dynamic launch (0 = 32)
int UCS_LAUNCH(cudaStream_t st, size_t items, const void* d_prep, void* d_scratch, unsigned int* d_err, const UcTune& tune) {
  if (!items) return 0;
  UcDevInfo di;
  int dev = 0;
  if (!uc_dev_info(di, dev)) return 0;
  static bool static_configured[kUcMaxDevices] = {false}, dyn_configured[kUcMaxDevices] = {false};
  const int natural = UCS_ITEMS_PER_BLOCK * UC_NSLOTS * (int)sizeof(Fp2);
  const int nblocks = uc_grid_for(items, UCS_ITEMS_PER_BLOCK);
  const int max_resident = di.smem_sm / (natural + 1024);
  // placement by shared-memory request: c = ceil(blocks / usable SMs) blocks per SM at most
  const int usable = di.nsm - (di.nsm + 9) / 10;
  const int cap = (nblocks + usable - 1) / usable;
  // Static launch while everything is resident with room to spare, and again for very large batches (thousands of
  // small blocks balance themselves); the task queue in between, where a static launch would run a ragged second
  // wave (measured, whole verification, 12 checks per proof: 500 proofs 4.57 static against 4.76 ms dynamic, 600
  // proofs 6.24 against 5.48, 1,000 proofs 8.90 against 8.21, 4,096 proofs 31.2 against 30.6, 8,192 proofs 58.8 against
  // 60.1, 20,000 proofs 141.5 against 146.5)
  const bool dynamic = tune.mode == 2 || (tune.mode == 0 && nblocks > (max_resident - 2) * usable && nblocks <= 8192);
  if (!dynamic) {
    if (!uc_optin_once(UCS_KERNEL, static_configured, dev, di.smem_optin)) return 0;
    int smem = natural;
    if (cap >= 1 && cap * (natural + 1024) <= di.smem_sm) {
      int want = (di.smem_sm / cap - 1024) & ~127;
      if (want > di.smem_optin) want = di.smem_optin;
      if (want > smem) smem = want;
    }
    UCS_KERNEL<<<nblocks, UCS_BLOCK, smem, st>>>(items, (const UcPrep*)d_prep, (GtPair*)d_scratch);
    return 1;
  }
  if (!uc_optin_once(UCS_KERNEL_DYN, dyn_configured, dev, di.smem_optin)) return 0;
  int resident = tune.resident > 0 ? tune.resident : max_resident;
  if (resident > max_resident) resident = max_resident;
  int smem = (di.smem_sm / resident - 1024) & ~127;
  if (smem > di.smem_optin) smem = di.smem_optin;
  if (smem < natural) smem = natural;
  const unsigned int nseg = tune.nseg > 0 ? (unsigned int)tune.nseg : 32u;
  // chunks bound the slot-image area (kUcDynChunkItems items at a time; consecutive launches on one stream reuse it)
  int launches = 0;
  unsigned char* base = (unsigned char*)d_prep + uc_align256(items * sizeof(UcPrep));
  for (size_t first = 0; first < items; first += kUcDynChunkItems) {
    const size_t n = items - first < kUcDynChunkItems ? items - first : kUcDynChunkItems;
    const size_t ngroups = (n + UCS_WARP_ITEMS - 1) / UCS_WARP_ITEMS;
    unsigned int* ctrl = (unsigned int*)base;
    uint4* state = (uint4*)(base + uc_dyn_ctrl_bytes());
    cudaMemsetAsync(ctrl, 0, 16 + 4 * ngroups, st);
    int grid = resident * di.nsm;
    const int nb = uc_grid_for(n, UCS_ITEMS_PER_BLOCK);
    if (grid > nb) grid = nb;
    UCS_KERNEL_DYN<<<grid, UCS_BLOCK, smem, st>>>(n, (const UcPrep*)d_prep + first, (GtPair*)d_scratch + first, state, ctrl,
                                                  nseg, d_err);
    launches++;
  }
  return launches;
}

#ifdef UCS_PROFILE
// Experiment builds only (tools/phase_profile.py): reads and clears the per-opcode cycle counters.
extern "C" int drx_debug_phase_profile(unsigned long long* out64) {
  static const unsigned long long zero[64] = {0};
  if (cudaDeviceSynchronize() != cudaSuccess) return -1;
  if (cudaMemcpyFromSymbol(out64, uc_prof, sizeof(zero)) != cudaSuccess) return -1;
  return cudaMemcpyToSymbol(uc_prof, zero, sizeof(zero)) == cudaSuccess ? 0 : -1;
}
#endif

}  // namespace drx
