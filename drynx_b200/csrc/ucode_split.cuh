// Lane-group pairing, two threads per program lane. The micro-program (pairing_ucode_g*.inc) and the slot image in
// shared memory are those of ucode_body.cuh; here each Fp2 operation of a program lane is carried out by a PAIR of
// threads, thread h of the pair producing component h of the result: an Fp2 product is one dual-product Montgomery
// pass per thread (c0 = a0 b0 + a1 (p - b1) on h = 0, c1 = a0 b1 + a1 b0 on h = 1), a squaring one product per thread,
// a sum one Fp addition per thread. Both threads read whatever operand components they need straight from the item's
// slots, so there are no shuffles and no cross-thread carries. Compared with one thread per lane this halves the
// dependent instruction chain of every phase (the lane-group kernels are latency-bound: a lone warp issues one
// instruction per ~4.5 cycles) and doubles the resident warps per item for the same shared-memory footprint.
// Replaces kyber pairing/bn256 optate.go for lib/range/range_proof.go:540-544, like ucode_body.cuh.
#pragma once
#include <cstdio>
#include <cstdlib>

#define DRX_FP2_INLINE_MUL 1
#include "ucode_common.cuh"

namespace drx {

#include UC_PROGRAM_INC

#define UCS_TPI (2 * UC_G)  // threads per item
#ifndef UCS_BLOCK
#define UCS_BLOCK 64
#endif
// An item never straddles a warp (its threads meet at warp barriers): a warp holds 32 / UCS_TPI whole items; a thread
// count that does not divide 32 (three program lanes = six threads per item: built and measured in round 2, 4.7-4.9 ms
// against 4.34 at 6,000 checks, 129 against 123 ms at 240,000, because the product phases per item do not shrink) leaves
// the last threads of the warp unseated.
#define UCS_WARP_ITEMS (32 / UCS_TPI)
#define UCS_ITEMS_PER_BLOCK ((UCS_BLOCK / 32) * UCS_WARP_ITEMS)
#ifndef UCS_MIN_BLOCKS
#define UCS_MIN_BLOCKS 8  // 16 warps per SM: at most 128 registers per thread
#endif
#define UCS_ITEM_BYTES (UC_NSLOTS * 64)
static_assert(UC_NSLOTS <= (int)kUcMaxSlots, "slot-image area is sized for kUcMaxSlots");

// Shared-memory slot access by 32-bit shared-window address. A slot is 64 bytes = four 16-byte chunks; logical chunk c
// lives at physical chunk c ^ UC_SWZ(slot). The program (uc_prog) carries each operand as the byte offset of its slot
// inside the item with the swizzle already folded in (slot * 64 + UC_SWZ(slot) * 16), items are 64-byte aligned, so the
// address of chunk c is (item + offset) ^ (c << 4): one XOR per chunk. The four program lanes of an item touch slots
// that differ mod 4 (the generator sees to it), so their 16-byte accesses fall into different bank groups.
__device__ __forceinline__ uint4 ucs_lds(uint32_t addr) {
  uint4 v;
  asm volatile("ld.shared.v4.u32 {%0, %1, %2, %3}, [%4];" : "=r"(v.x), "=r"(v.y), "=r"(v.z), "=r"(v.w) : "r"(addr) : "memory");
  return v;
}
__device__ __forceinline__ void ucs_sts(uint32_t addr, const uint32_t* w) {
  asm volatile("st.shared.v4.u32 [%0], {%1, %2, %3, %4};" ::"r"(addr), "r"(w[0]), "r"(w[1]), "r"(w[2]), "r"(w[3]) : "memory");
}
__device__ __forceinline__ uint32_t ucs_slot_off(uint32_t slot) { return slot * 64u + UC_SWZ(slot) * 16u; }
// component `comp` (chunks 2 comp, 2 comp + 1) of the slot at item address S + off
__device__ __forceinline__ Fp ucs_load_c(uint32_t S, uint32_t off, uint32_t comp) {
  const uint32_t a0 = (S + off) ^ (comp << 5);
  const uint4 lo = ucs_lds(a0), hi = ucs_lds(a0 ^ 16u);
  Fp r;
  r.v[0] = lo.x; r.v[1] = lo.y; r.v[2] = lo.z; r.v[3] = lo.w;
  r.v[4] = hi.x; r.v[5] = hi.y; r.v[6] = hi.z; r.v[7] = hi.w;
  return r;
}
__device__ __forceinline__ void ucs_store_c(uint32_t S, uint32_t off, uint32_t comp, const Fp& r) {
  const uint32_t a0 = (S + off) ^ (comp << 5);
  ucs_sts(a0, &r.v[0]);
  ucs_sts(a0 ^ 16u, &r.v[4]);
}
__device__ __forceinline__ Fp2 ucs_load(uint32_t S, uint32_t off) {
  return Fp2{ucs_load_c(S, off, 0u), ucs_load_c(S, off, 1u)};
}
__device__ __forceinline__ void ucs_store(uint32_t S, uint32_t off, const Fp2& r) {
  ucs_store_c(S, off, 0u, r.c0);
  ucs_store_c(S, off, 1u, r.c1);
}
__device__ __forceinline__ Fp uc_constp_(uint32_t k) {
  Fp r;
#pragma unroll
  for (int q = 0; q < 8; q++) r.v[q] = uc_constp[k][q];
  return r;
}
// p - b, left unreduced (p for b = 0): only ever a multiplier of mont_mul2, which accepts any 256-bit value there
__device__ __forceinline__ Fp ucs_neg_raw(const Fp& b) {
  Fp r;
  r.v[0] = sub_cc(ModP::mod(0), b.v[0]);
#pragma unroll
  for (int k = 1; k < 7; k++) r.v[k] = subc_cc(ModP::mod(k), b.v[k]);
  r.v[7] = subc(ModP::mod(7), b.v[7]);
  return r;
}
__device__ __forceinline__ Fp ucs_dual(const Fp& a0, const Fp& x, const Fp& a1, const Fp& y) {
  Fp r;
  mont_mul2<ModP>(r.v, a0.v, x.v, a1.v, y.v);
  return r;
}
__device__ __forceinline__ Fp ucs_const2_c(uint32_t k, uint32_t comp) {
  Fp r;
#pragma unroll
  for (int q = 0; q < 8; q++) r.v[q] = uc_const2[k][8 * comp + q];
  return r;
}
// component h of xi * a = (3 a0 - a1) + (3 a1 + a0) i, from a's own component ah and the other one ao
__device__ __forceinline__ Fp ucs_xi_c(const Fp& ah, const Fp& ao, uint32_t h) {
  const Fp t = fp_add(fp_add(ah, ah), ah);
  return h ? fp_add(t, ao) : fp_sub(t, ao);
}

// Phases [ph0, ph1) of the program on the item at shared address S; all lanes of the warp call it together. Slots are
// reused in place (a phase may overwrite one of its own operands), and the two threads of a pair read each other's
// components, so every load of a phase is separated from every store by a warp barrier.
__device__ __forceinline__ void ucs_run(uint32_t S, int lane, uint32_t h, bool alive, int ph0, int ph1) {
  uint32_t op = uc_ops[ph0];
  uint2 w = uc_prog[ph0 * UC_G + lane];
  for (int ph = ph0; ph < ph1; ph++) {
    const int nx = ph + 1 < ph1 ? ph + 1 : ph;
    const uint32_t op_n = uc_ops[nx];
    const uint2 w_n = uc_prog[nx * UC_G + lane];
    const bool act = alive && w.x != 0xffffffffu;
    const uint32_t d = w.x & 0xffffu, ia = w.x >> 16, ib = w.y & 0xffffu, ic = w.y >> 16;
    Fp r;
    if (act) {
      if (op == UC_ADD) {
        r = fp_add(ucs_load_c(S, ia, h), ucs_load_c(S, ib, h));
      } else if (op == UC_MUL) {
        const Fp2 a = ucs_load(S, ia);
        const Fp x = ucs_load_c(S, ib, h), yr = ucs_load_c(S, ib, h ^ 1u);
        r = ucs_dual(a.c0, x, a.c1, fp_sel(h != 0, yr, ucs_neg_raw(yr)));
      } else if (op == UC_SQR) {
        // (a0 + a1)(a0 - a1) on h = 0, (a0 + a0) a1 on h = 1: one product each
        const Fp2 a = ucs_load(S, ia);
        const Fp x = fp_add(a.c0, fp_sel(h != 0, a.c0, a.c1));
        const Fp y = fp_sel(h != 0, a.c1, fp_sub(a.c0, a.c1));
        r = fp_mul_inl(x, y);
      } else if (op == UC_SUB2) {
        r = fp_sub(fp_sub(ucs_load_c(S, ia, h), ucs_load_c(S, ib, h)), ucs_load_c(S, ic, h));
      } else if (op == UC_AXI) {
        r = fp_add(ucs_load_c(S, ia, h), ucs_xi_c(ucs_load_c(S, ib, h), ucs_load_c(S, ib, h ^ 1u), h));
      } else if (op == UC_SUB) {
        r = fp_sub(ucs_load_c(S, ia, h), ucs_load_c(S, ib, h));
      } else if (op == UC_CYCP || op == UC_CYCM) {
        const Fp a = ucs_load_c(S, ia, h), b = ucs_load_c(S, ib, h);
        const Fp s = op == UC_CYCP ? fp_add(a, b) : fp_sub(a, b);
        r = fp_add(fp_add(s, s), a);
      } else if (op == UC_MULXI) {
        r = ucs_xi_c(ucs_load_c(S, ia, h), ucs_load_c(S, ia, h ^ 1u), h);
      } else if (op == UC_MULFP0 || op == UC_MULFP1) {
        r = fp_mul(ucs_load_c(S, ia, h), ucs_load_c(S, ib, op == UC_MULFP1 ? 1u : 0u));
      } else if (op == UC_X4 || op == UC_X8 || op == UC_X3) {
        const Fp a = ucs_load_c(S, ia, h);
        const Fp d2 = fp_add(a, a);
        if (op == UC_X3) {
          r = fp_add(d2, a);
        } else {
          r = fp_add(d2, d2);
          if (op == UC_X8) r = fp_add(r, r);
        }
      } else if (op == UC_NEG) {
        r = fp_neg(ucs_load_c(S, ia, h));
      } else if (op == UC_CONJMULC || op == UC_MULC) {
        // (a0 +- a1 i)(k0 + k1 i): MULC c0 = a0 k0 + a1 (p - k1), c1 = a0 k1 + a1 k0;
        // CONJMULC c0 = a0 k0 + a1 k1, c1 = a0 k1 + a1 (p - k0)
        const Fp2 a = ucs_load(S, ia);
        const Fp x = ucs_const2_c(ib, h), yr = ucs_const2_c(ib, h ^ 1u);
        const bool negate = (op == UC_MULC) == (h == 0);
        r = ucs_dual(a.c0, x, a.c1, fp_sel(negate, ucs_neg_raw(yr), yr));
      } else if (op == UC_MULFPC) {
        r = fp_mul(ucs_load_c(S, ia, h), uc_constp_(ib));
      } else if (op == UC_CONJ) {
        const Fp a = ucs_load_c(S, ia, h);
        r = h ? fp_neg(a) : a;
      } else if (op == UC_INV) {
        Fp2 tmp = ucs_load(S, ia);  // both threads invert (same instructions either way) and keep their component
        const Fp2 v = fp2_inv(tmp);
        r = fp_sel(h != 0, v.c1, v.c0);
      } else if (op == UC_CHK) {
        // verdict of the G2 membership test: zero iff a == 0, b == 0 and c != 0 (see tools/microcode/gen.py)
        const bool good = fp2_is_zero(ucs_load(S, ia)) & fp2_is_zero(ucs_load(S, ib)) & !fp2_is_zero(ucs_load(S, ic));
        r = good ? fp_zero() : fp_one();
      } else {
        r = ucs_load_c(S, ia, h);  // UC_COPY
      }
    }
    __syncwarp();
    if (act) ucs_store_c(S, d, h, r);
    __syncwarp();
    op = op_n;
    w = w_n;
  }
}
// the line functions only ever use -x_P: the P slot holds (-x_P, y_P)
__device__ __forceinline__ void ucs_put_inputs(uint32_t S, const UcPrep& in) {
  ucs_store(S, ucs_slot_off(UC_IN_P), Fp2{fp_neg(in.P.x), in.P.y});
  ucs_store(S, ucs_slot_off(UC_IN_QX), in.V.x);
  ucs_store(S, ucs_slot_off(UC_IN_QY), in.V.y);
}
// e(P, V), or the all-zero element (no member of GT, so the comparison that follows fails) when V is outside G2
__device__ __forceinline__ Fp12 ucs_get_output(uint32_t S) {
  Fp12 e;
  e.c0.c0 = ucs_load(S, ucs_slot_off(UC_OUT_E00));
  e.c0.c1 = ucs_load(S, ucs_slot_off(UC_OUT_E01));
  e.c0.c2 = ucs_load(S, ucs_slot_off(UC_OUT_E02));
  e.c1.c0 = ucs_load(S, ucs_slot_off(UC_OUT_E10));
  e.c1.c1 = ucs_load(S, ucs_slot_off(UC_OUT_E11));
  e.c1.c2 = ucs_load(S, ucs_slot_off(UC_OUT_E12));
  if (!fp2_is_zero(ucs_load(S, ucs_slot_off(UC_OUT_SG)))) {
    const Fp6 z = fp6_zero();
    e.c0 = z;
    e.c1 = z;
  }
  return e;
}

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
  if (alive && sub == 0) ucs_put_inputs(S, prep[t]);
  __syncwarp();
  ucs_run(S, sub >> 1, (uint32_t)(sub & 1), alive, 0, UC_NPHASES);
  if (seated && t < items && sub == 0) scratch[t].a = alive ? ucs_get_output(S) : fp12_one();
}

// Dynamic variant (same scheme as UC_KERNEL_DYN in ucode_body.cuh): the program is cut into nseg segments, a task is
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
      if (alive && sub == 0) ucs_put_inputs(S, prep[t]);
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
  // placement by shared-memory request, as in ucode_body.cuh: c = ceil(blocks / usable SMs) blocks per SM at most
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

}  // namespace drx
