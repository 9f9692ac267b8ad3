// Cooperative pairing for range-proof verification: a group of UC_G lanes per (proof, server, digit) item, the
// item's state held ONCE as Fp2 slots in shared memory, the whole optimal ate pairing (Miller loop and final
// exponentiation) executed as a static schedule of Fp2 operations (pairing_ucode.inc, generated and checked against
// the big-integer definition by tools/microcode/gen.py). Every phase all lanes run the same operation on their own
// (dst, a, b) slots; no thread-local stack is involved, which is what bounded the one-thread-per-item kernel
// (DESIGN.md section 5). Replaces kyber pairing/bn256 optate.go for lib/range/range_proof.go:540-544.
#pragma once
#include <cstdio>
#include <cstdlib>

#define DRX_FP2_INLINE_MUL 1
#include "rp_device.cuh"

namespace drx {

#include UC_PROGRAM_INC

static inline int grid_for(size_t n, int threads) { return (int)((n + threads - 1) / threads); }

struct UcPrep {      // per item: the two pairing arguments in Montgomery affine form
  G1Aff P;           // c*Y_i - Zphi_j*B
  G2Aff V;
  uint32_t state;    // 0: run the pairing, 1: an argument is the point at infinity (value 1), 2: malformed V
  uint32_t pad[3];
};

#ifdef UC_WITH_PREP
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
  prep[t].state = (aff_is_inf(P) | aff_is_inf(V)) ? 1u : 0u;
}

#endif

// ---- Fp2 operations on register values (products go through the shared out-of-line fp_mul)
__device__ __forceinline__ Fp2 uc_mul(const Fp2& a, const Fp2& b) {
#if DRX_FP2_DUAL >= 1
  return fp2_mul_dual(a, b);
#endif
  // the three products are inlined here (and only here) so that ptxas interleaves their carry chains
  const Fp v0 = fp_mul_inl(a.c0, b.c0);
  const Fp v1 = fp_mul_inl(a.c1, b.c1);
  const Fp s = fp_mul_inl(fp_add(a.c0, a.c1), fp_add(b.c0, b.c1));
  return Fp2{fp_sub(v0, v1), fp_sub(fp_sub(s, v0), v1)};
}
// a^2 = (a0 + a1)(a0 - a1) + 2 a0 a1 i: two products instead of three; only reached when every lane of the phase
// squares (the generator labels such phases UC_SQR), so the warp never runs both forms
__device__ __forceinline__ Fp2 uc_sqr(const Fp2& a) {
  const Fp m = fp_mul_inl(a.c0, a.c1);
  const Fp d = fp_mul_inl(fp_add(a.c0, a.c1), fp_sub(a.c0, a.c1));
  return Fp2{d, fp_add(m, m)};
}
__device__ __forceinline__ Fp2 uc_add(const Fp2& a, const Fp2& b) { return Fp2{fp_add(a.c0, b.c0), fp_add(a.c1, b.c1)}; }
__device__ __forceinline__ Fp2 uc_sub(const Fp2& a, const Fp2& b) { return Fp2{fp_sub(a.c0, b.c0), fp_sub(a.c1, b.c1)}; }
__device__ __forceinline__ Fp2 uc_xi(const Fp2& a) {
  const Fp t0 = fp_add(fp_add(a.c0, a.c0), a.c0);
  const Fp t1 = fp_add(fp_add(a.c1, a.c1), a.c1);
  return Fp2{fp_sub(t0, a.c1), fp_add(t1, a.c0)};
}
__device__ __forceinline__ Fp2 uc_load_const2(uint32_t k) {
  Fp2 r;
#pragma unroll
  for (int q = 0; q < 8; q++) {
    r.c0.v[q] = uc_const2[k][q];
    r.c1.v[q] = uc_const2[k][8 + q];
  }
  return r;
}
__device__ __forceinline__ Fp uc_constp_(uint32_t k) {
  Fp r;
#pragma unroll
  for (int q = 0; q < 8; q++) r.v[q] = uc_constp[k][q];
  return r;
}

// Shared-memory slot access. A slot is 64 bytes = four 16-byte chunks; logical chunk c of slot s lives at physical
// chunk (c + s) & 3, and the per-item slot count is odd. A 128-bit shared load is served a quarter-warp (two lane
// groups) at a time: the two groups run the same micro-instruction, so lane l of both reads the same slot number,
// which lands in opposite bank halves (odd stride), and inside a group the generator keeps the slots touched
// together distinct mod 4, so the eight 16-byte accesses hit eight different bank groups.
__device__ __forceinline__ Fp2 uc_load(const Fp2* S, uint32_t s) {
  const uint4* base = reinterpret_cast<const uint4*>(S + s);
  Fp2 r;
#pragma unroll
  for (int c = 0; c < 4; c++) {
    const uint4 v = base[(c + UC_ROT(s)) & 3u];
    uint32_t* dst = c < 2 ? &r.c0.v[4 * c] : &r.c1.v[4 * (c - 2)];
    dst[0] = v.x; dst[1] = v.y; dst[2] = v.z; dst[3] = v.w;
  }
  return r;
}
__device__ __forceinline__ void uc_store(Fp2* S, uint32_t s, const Fp2& r) {
  uint4* base = reinterpret_cast<uint4*>(S + s);
#pragma unroll
  for (int c = 0; c < 4; c++) {
    const uint32_t* src = c < 2 ? &r.c0.v[4 * c] : &r.c1.v[4 * (c - 2)];
    base[(c + UC_ROT(s)) & 3u] = make_uint4(src[0], src[1], src[2], src[3]);
  }
}

#ifndef UC_BLOCK
#define UC_BLOCK 64
#endif
#define UC_ITEMS_PER_BLOCK (UC_BLOCK / UC_G)

#ifdef UC_CLOCKS
// Experiment build (-DUC_CLOCKS, never the shipped library): cycle accounting of the phase loop for thread 0 of
// block 0, printed when the kernel ends. The clock reads take a data dependence on the values they follow, so
// "operand" ends when the first operand has arrived and "compute" when the result exists.
__device__ __forceinline__ long long uc_clk(uint32_t dep) {
  long long t;
  asm volatile("mov.u64 %0, %%clock64; // %1" : "=l"(t) : "r"(dep) : "memory");
  return t;
}
#endif

// Phases [ph0, ph1) of the program on one item's slots (S); all lanes of the warp call it together.
__device__ __forceinline__ void uc_run(Fp2* S, int lane, bool alive, int ph0, int ph1) {
  uint32_t op = uc_ops[ph0], w = uc_lanes[ph0 * UC_G + lane];
#ifdef UC_CLOCKS
  long long acc[3][3] = {{0}}, whole = 0;
  int cnt[3] = {0, 0, 0};
#endif
  for (int ph = ph0; ph < ph1; ph++) {
#ifdef UC_CLOCKS
    const long long c0 = uc_clk(op);
#endif
    // fetch the next phase's words while this one computes
    const int nx = ph + 1 < ph1 ? ph + 1 : ph;
    const uint32_t op_n = uc_ops[nx], w_n = uc_lanes[nx * UC_G + lane];
    if (alive && w != 0xffffffffu) {
      const uint32_t d = w & 255u, ia = (w >> 8) & 255u, ib = (w >> 16) & 255u, ic = w >> 24;
      const Fp2 a = uc_load(S, ia);
#ifdef UC_CLOCKS
      const long long c1 = uc_clk(a.c0.v[0] ^ a.c1.v[7]);
#endif
      Fp2 r;
      // ordered by how often the generated program uses each operation (a compare-and-branch chain is cheaper
      // than the balanced search ptxas builds for a switch, and the hot cases come first)
      if (op == UC_ADD) {
        r = uc_add(a, uc_load(S, ib));
      } else if (op == UC_MUL) {
        r = uc_mul(a, uc_load(S, ib));
      } else if (op == UC_SQR) {
        r = uc_sqr(a);
      } else if (op == UC_SUB2) {
        r = uc_sub(uc_sub(a, uc_load(S, ib)), uc_load(S, ic));
      } else if (op == UC_AXI) {
        r = uc_add(a, uc_xi(uc_load(S, ib)));
      } else if (op == UC_SUB) {
        r = uc_sub(a, uc_load(S, ib));
      } else if (op == UC_CYCP) {
        r = uc_add(a, uc_load(S, ib));
        r = uc_add(uc_add(r, r), a);
      } else if (op == UC_CYCM) {
        r = uc_sub(a, uc_load(S, ib));
        r = uc_add(uc_add(r, r), a);
      } else if (op == UC_MULXI) {
        r = uc_xi(a);
      } else if (op == UC_MULFP0 || op == UC_MULFP1) {
        const Fp2 pt = uc_load(S, ib);
        const Fp k = op == UC_MULFP0 ? pt.c0 : pt.c1;
        r = Fp2{fp_mul(a.c0, k), fp_mul(a.c1, k)};
      } else if (op == UC_X4 || op == UC_X8 || op == UC_X3) {
        const Fp2 d2 = uc_add(a, a);
        if (op == UC_X3) {
          r = uc_add(d2, a);
        } else {
          r = uc_add(d2, d2);
          if (op == UC_X8) r = uc_add(r, r);
        }
      } else if (op == UC_CONJMULC || op == UC_MULC) {
        const Fp2 x = op == UC_CONJMULC ? Fp2{a.c0, fp_neg(a.c1)} : a;
        r = uc_mul(x, uc_load_const2(ib));
      } else if (op == UC_MULFPC) {
        const Fp k = uc_constp_(ib);
        r = Fp2{fp_mul(a.c0, k), fp_mul(a.c1, k)};
      } else if (op == UC_CONJ) {
        r = Fp2{a.c0, fp_neg(a.c1)};
      } else if (op == UC_INV) {
        Fp2 tmp = a;  // only this copy has its address taken (fp2_inv is out of line)
        r = fp2_inv(tmp);
      } else {
        r = a;  // UC_COPY
      }
#ifdef UC_CLOCKS
      const long long c2 = uc_clk(r.c0.v[0] ^ r.c1.v[7] ^ r.c1.v[0] ^ r.c0.v[7]);
#endif
      uc_store(S, d, r);
#ifdef UC_CLOCKS
      const long long c3 = uc_clk(d);
      const int cls = op == UC_ADD ? 0 : op == UC_MUL ? 1 : 2;
      acc[cls][0] += c1 - c0;
      acc[cls][1] += c2 - c1;
      acc[cls][2] += c3 - c2;
      cnt[cls]++;
#endif
    }
    __syncwarp();
    op = op_n;
    w = w_n;
#ifdef UC_CLOCKS
    whole += uc_clk(op) - c0;
#endif
  }
#ifdef UC_CLOCKS
  if (blockIdx.x == 0 && threadIdx.x == 0)
    printf("UC_CLOCKS G=%d cycles (operand / compute / store) | ADD n=%d %lld %lld %lld | MUL n=%d %lld %lld %lld | other n=%d %lld "
           "%lld %lld | all phases %lld\n", UC_G, cnt[0], acc[0][0], acc[0][1], acc[0][2], cnt[1], acc[1][0], acc[1][1], acc[1][2],
           cnt[2], acc[2][0], acc[2][1], acc[2][2], whole);
#endif
}
__device__ __forceinline__ void uc_put_inputs(Fp2* S, const UcPrep& in) {
  uc_store(S, UC_IN_P, Fp2{in.P.x, in.P.y});
  uc_store(S, UC_IN_QX, in.V.x);
  uc_store(S, UC_IN_QY, in.V.y);
  uc_store(S, UC_IN_ONE, fp2_one());
  uc_store(S, UC_IN_ZERO, fp2_zero());
}
__device__ __forceinline__ Fp12 uc_get_output(const Fp2* S) {
  Fp12 e;
  e.c0.c0 = uc_load(S, UC_OUT_E00);
  e.c0.c1 = uc_load(S, UC_OUT_E01);
  e.c0.c2 = uc_load(S, UC_OUT_E02);
  e.c1.c0 = uc_load(S, UC_OUT_E10);
  e.c1.c1 = uc_load(S, UC_OUT_E11);
  e.c1.c2 = uc_load(S, UC_OUT_E12);
  return e;
}

__global__ void __launch_bounds__(UC_BLOCK)
UC_KERNEL(size_t items, const UcPrep* __restrict__ prep, GtPair* __restrict__ scratch) {
  extern __shared__ __align__(16) unsigned char uc_smem[];
  Fp2* slots = (Fp2*)uc_smem;
  const int grp = threadIdx.x / UC_G, lane = threadIdx.x % UC_G;
  const size_t t = (size_t)blockIdx.x * UC_ITEMS_PER_BLOCK + grp;
  const bool alive = t < items && prep[t].state == 0;
  Fp2* S = slots + grp * UC_NSLOTS;
  if (alive && lane == 0) uc_put_inputs(S, prep[t]);
  __syncwarp();
  uc_run(S, lane, alive, 0, UC_NPHASES);
  if (t < items && lane == 0) scratch[t].a = alive ? uc_get_output(S) : fp12_one();
}

#ifdef UC_WITH_DYN
// Dynamic variant for batches that would need a fourth block per SM or a second wave. The program is cut into UC_NSEG
// segments of equal phase count; a task is (group of 32 / UC_G items, segment). Persistent warps (three blocks per
// SM) take tasks from one atomic counter, segment-major, so every SM stays busy until the queue is empty instead of
// the kernel running wave after wave at the pace of its fullest SM (measured, static -> dynamic: 789 proofs
// 8.8 -> 7.1 ms, 1,000 proofs 10.3 -> 9.0 ms; at 500 proofs the static launch is faster, 5.45 against 5.56 ms). Between segments the
// group's slots travel through global memory (the shared-memory image, copied as it is). A task waits for the
// previous segment of its group on a flag word; that segment was handed out earlier, so its warp is already
// running and the wait is bounded (it traps instead of spinning forever).
#define UC_NSEG 32
#define UC_GROUP_ITEMS (32 / UC_G)
#define UC_GROUP_U4 (UC_GROUP_ITEMS * UC_NSLOTS * 4)  // uint4 words of one group's slots
__global__ void __launch_bounds__(UC_BLOCK)
UC_KERNEL_DYN(size_t items, const UcPrep* __restrict__ prep, GtPair* __restrict__ scratch, uint4* __restrict__ state,
              unsigned int* __restrict__ ctrl, unsigned int nseg) {
  extern __shared__ __align__(16) unsigned char uc_smem[];
  Fp2* slots = (Fp2*)uc_smem;
  const int grp = threadIdx.x / UC_G, lane = threadIdx.x % UC_G, wl = threadIdx.x & 31, warp = threadIdx.x >> 5;
  Fp2* S = slots + grp * UC_NSLOTS;
  uint4* W = reinterpret_cast<uint4*>(slots + warp * UC_GROUP_ITEMS * UC_NSLOTS);
  const unsigned int ngroups = (unsigned int)((items + UC_GROUP_ITEMS - 1) / UC_GROUP_ITEMS);
  const unsigned int ntasks = ngroups * nseg;
  volatile unsigned int* done = ctrl + 4;
  for (;;) {
    unsigned int task = 0;
    if (wl == 0) task = atomicAdd(&ctrl[0], 1u);
    task = __shfl_sync(0xffffffffu, task, 0);
    if (task >= ntasks) break;
    const unsigned int g = task % ngroups, seg = task / ngroups;
    const size_t t = (size_t)g * UC_GROUP_ITEMS + wl / UC_G;
    const bool alive = t < items && prep[t].state == 0;
    if (seg == 0) {
      if (alive && lane == 0) uc_put_inputs(S, prep[t]);
    } else {
      if (wl == 0) {
        unsigned int polls = 0;
        while (done[g] < seg) {
          __nanosleep(200);
          if (++polls > (1u << 24)) __trap();
        }
      }
      __syncwarp();
      __threadfence();
      const uint4* src = state + (size_t)g * UC_GROUP_U4;
      for (int i = wl; i < UC_GROUP_U4; i += 32) W[i] = __ldcg(src + i);
    }
    __syncwarp();
    uc_run(S, lane, alive, (int)(seg * UC_NPHASES / nseg), (int)((seg + 1) * UC_NPHASES / nseg));
    if (seg + 1 < nseg) {
      uint4* dst = state + (size_t)g * UC_GROUP_U4;
      for (int i = wl; i < UC_GROUP_U4; i += 32) __stcg(dst + i, W[i]);
      __threadfence();
      __syncwarp();
      if (wl == 0) done[g] = seg + 1;
    } else if (t < items && lane == 0) {
      scratch[t].a = alive ? uc_get_output(S) : fp12_one();
    }
    __syncwarp();
  }
}
#endif

#ifdef UC_WITH_PREP
// prepared arguments, then (dynamic kernel only) the control words and the groups' slot images
static size_t uc_align256(size_t x) { return (x + 255) & ~(size_t)255; }
size_t rp_ucode_prep_bytes(size_t items) {
  const size_t ngroups = (items + UC_GROUP_ITEMS - 1) / UC_GROUP_ITEMS;
  return uc_align256(items * sizeof(UcPrep)) + uc_align256(16 + 4 * ngroups) + ngroups * UC_GROUP_U4 * 16;
}
#endif

#ifdef UC_WITH_PREP
int launch_rp_verify_prep(cudaStream_t st, const RpBatch& b, const RpSetDev* d_sets, const void* d_tabB,
                          const uint8_t* d_proofs, uint32_t* d_flags, void* d_prep) {
  if (!b.items) return 0;
  k_rp_verify_prep<<<grid_for(b.items * 4, 64), 64, 0, st>>>(b, d_sets, (const G1Aff*)d_tabB, (const uint32_t*)d_proofs,
                                                        d_flags, (UcPrep*)d_prep);
  return 1;
}
#endif
// cooperative pairing over prepared items; writes e into scratch[t].a (the slot role 0 of k_rp_verify_roles fills)
static int uc_smem_optin = 0;
int UC_LAUNCH(cudaStream_t st, size_t items, const void* d_prep, void* d_scratch) {
  if (!items) return 0;
  // Shared memory per block doubles as the placement control. The block scheduler fills SMs GPC by GPC, and the
  // GPCs of a B200 differ in size, so with the natural request (up to 4-7 blocks per SM) some SMs end up with more
  // blocks than ceil(blocks / SMs) while others idle -- and the kernel runs at the pace of its fullest SM. Asking
  // for 1/c of an SM's shared memory, c = ceil(blocks / usable SMs), caps every SM at c blocks and forces an even
  // spread (measured at 375 blocks: 5.70 -> 5.45 ms).
  static int nsm = 0, smem_sm = 0;
  const int natural = UC_ITEMS_PER_BLOCK * UC_NSLOTS * (int)sizeof(Fp2);
  if (!nsm) {
    int dev = 0;
    cudaGetDevice(&dev);
    cudaDeviceGetAttribute(&nsm, cudaDevAttrMultiProcessorCount, dev);
    cudaDeviceGetAttribute(&smem_sm, cudaDevAttrMaxSharedMemoryPerMultiprocessor, dev);
    int optin = 0;
    cudaDeviceGetAttribute(&optin, cudaDevAttrMaxSharedMemoryPerBlockOptin, dev);
    cudaFuncSetAttribute(UC_KERNEL, cudaFuncAttributeMaxDynamicSharedMemorySize, optin);
    uc_smem_optin = optin;
  }
  const int nblocks = grid_for(items, UC_ITEMS_PER_BLOCK);
  // 10 % of the SMs are left out of the count: a (nearly) exact fit (blocks = c * SMs) leaves no room for the SMs
  // the side-stream kernels configured for L1 before the pairing blocks arrived (an SM changes its L1 / shared split
  // only when it is empty), and a block that finds no SM waits for a whole wave (measured: 5.6 -> 7.2 ms)
  const int usable = nsm - (nsm + 9) / 10;
  const int cap = (nblocks + usable - 1) / usable;
  int smem = natural;
  if (cap >= 1 && cap * (natural + 1024) <= smem_sm) {
    int want = (smem_sm / cap - 1024) & ~127;  // 1 KB per resident block is reserved by the system
    if (want > uc_smem_optin) want = uc_smem_optin;
    if (want > smem) smem = want;
  }
#ifdef UC_WITH_DYN
  // more than three blocks per usable SM: deal the work dynamically over three or four persistent blocks per SM
  // (a few blocks early: within eight blocks of three per usable SM the static launch already loses a block to a
  // second wave -- 390 blocks 5.29 ms, 395 blocks 6.59, 399 blocks 7.15 against 5.3-5.4 dynamic)
  if (nblocks > 3 * usable - 8) {
    static bool dyn_configured = false;
    if (!dyn_configured) {
      cudaFuncSetAttribute(UC_KERNEL_DYN, cudaFuncAttributeMaxDynamicSharedMemorySize, uc_smem_optin);
      dyn_configured = true;
    }
    const size_t ngroups = (items + UC_GROUP_ITEMS - 1) / UC_GROUP_ITEMS;
    unsigned char* base = (unsigned char*)d_prep + uc_align256(items * sizeof(UcPrep));
    unsigned int* ctrl = (unsigned int*)base;
    uint4* state = (uint4*)(base + uc_align256(16 + 4 * ngroups));
    cudaMemsetAsync(ctrl, 0, 16 + 4 * ngroups, st);
    // persistent blocks per SM: three up to ~1,000 warp tasks, four above (measured after the dual-product Fp2
    // form: 900 tasks 5.51 against 6.29 ms, 1,050 tasks 6.16 against 5.89, 3,000 tasks 17.3 against 14.5)
    int resident = ngroups >= 1000 ? 4 : 3;
    if (resident * (natural + 1024) > smem_sm) resident = 3;
    const int smem3 = (smem_sm / resident - 1024) & ~127;
    int grid = resident * nsm;
    if (grid > nblocks) grid = nblocks;
    const unsigned int nseg = UC_NSEG;
    UC_KERNEL_DYN<<<grid, UC_BLOCK, smem3, st>>>(items, (const UcPrep*)d_prep, (GtPair*)d_scratch, state, ctrl, nseg);
    return 1;
  }
#endif
  UC_KERNEL<<<grid_for(items, UC_ITEMS_PER_BLOCK), UC_BLOCK, smem, st>>>(items, (const UcPrep*)d_prep, (GtPair*)d_scratch);
  return 1;
}

}  // namespace drx
