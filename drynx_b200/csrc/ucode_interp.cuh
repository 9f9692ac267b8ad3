// Interpreter of the lane-group micro-programs (tools/microcode/gen.py): slot access in shared memory and the phase loop
// with two threads per program lane. Included inside namespace drx by a translation unit that has defined
// UC_PROGRAM_INC (the generated tables it runs): ucode_split.cuh for the pairing, kernels_rp_gtlanes.cu for the Fp12
// multiplication behind fixed-base GT exponentiations.
// The slot access and ucs_phase also compile for the host (the DRX_HOST_EMU test build runs the emitted program through
// the very code the kernels execute, shared memory replaced by a byte array); the phase loop itself is device code.
#if defined(__CUDACC__)
#define UCS_FN __device__ __forceinline__
#else
#define UCS_FN inline
#endif
#include UC_PROGRAM_INC
#ifdef UCS_PROFILE
static __device__ unsigned long long uc_prof[64];  // [op]: cycles, [32 + op]: phases
#endif

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
static_assert(UC_PAD_PHASES >= 1, "the interpreter reads one phase ahead");
#if defined(__CUDACC__)
static_assert(UC_NSLOTS <= (int)kUcMaxSlots, "slot-image area is sized for kUcMaxSlots");
#endif

// Shared-memory slot access by 32-bit shared-window address. A slot is 64 bytes = four 16-byte chunks; logical chunk c
// lives at physical chunk c ^ UC_SWZ(slot). The program (uc_prog) carries each operand as the byte offset of its slot
// inside the item with the swizzle already folded in (slot * 64 + UC_SWZ(slot) * 16), items are 64-byte aligned, so the
// address of chunk c is (item + offset) ^ (c << 4): one XOR per chunk. The four program lanes of an item touch slots
// that differ mod 4 (the generator sees to it), so their 16-byte accesses fall into different bank groups.
#if defined(__CUDACC__)
__device__ __forceinline__ uint4 ucs_lds(uint32_t addr) {
  uint4 v;
  asm volatile("ld.shared.v4.u32 {%0, %1, %2, %3}, [%4];" : "=r"(v.x), "=r"(v.y), "=r"(v.z), "=r"(v.w) : "r"(addr) : "memory");
  return v;
}
__device__ __forceinline__ void ucs_sts(uint32_t addr, const uint32_t* w) {
  asm volatile("st.shared.v4.u32 [%0], {%1, %2, %3, %4};" ::"r"(addr), "r"(w[0]), "r"(w[1]), "r"(w[2]), "r"(w[3]) : "memory");
}
#else   // host emulation: `addr` is an offset into ucs_host_smem
static thread_local unsigned char* ucs_host_smem = nullptr;
inline uint4 ucs_lds(uint32_t addr) {
  uint4 v;
  memcpy(&v, ucs_host_smem + addr, 16);
  return v;
}
inline void ucs_sts(uint32_t addr, const uint32_t* w) { memcpy(ucs_host_smem + addr, w, 16); }
#endif
UCS_FN uint32_t ucs_slot_off(uint32_t slot) { return slot * 64u + UC_SWZ(slot) * 16u; }
// component `comp` (chunks 2 comp, 2 comp + 1) of the slot at item address S + off
UCS_FN Fp ucs_load_c(uint32_t S, uint32_t off, uint32_t comp) {
  const uint32_t a0 = (S + off) ^ (comp << 5);
  const uint4 lo = ucs_lds(a0), hi = ucs_lds(a0 ^ 16u);
  Fp r;
  r.v[0] = lo.x; r.v[1] = lo.y; r.v[2] = lo.z; r.v[3] = lo.w;
  r.v[4] = hi.x; r.v[5] = hi.y; r.v[6] = hi.z; r.v[7] = hi.w;
  return r;
}
UCS_FN void ucs_store_c(uint32_t S, uint32_t off, uint32_t comp, const Fp& r) {
  const uint32_t a0 = (S + off) ^ (comp << 5);
  ucs_sts(a0, &r.v[0]);
  ucs_sts(a0 ^ 16u, &r.v[4]);
}
UCS_FN Fp2 ucs_load(uint32_t S, uint32_t off) {
  return Fp2{ucs_load_c(S, off, 0u), ucs_load_c(S, off, 1u)};
}
UCS_FN void ucs_store(uint32_t S, uint32_t off, const Fp2& r) {
  ucs_store_c(S, off, 0u, r.c0);
  ucs_store_c(S, off, 1u, r.c1);
}
UCS_FN Fp uc_constp_(uint32_t k) {
  Fp r;
#pragma unroll
  for (int q = 0; q < 8; q++) r.v[q] = uc_constp[k][q];
  return r;
}
// p - b, left unreduced (p for b = 0): only ever a multiplier of mont_mul2, which accepts any 256-bit value there
UCS_FN Fp ucs_neg_raw(const Fp& b) {
  Fp r;
  r.v[0] = sub_cc(ModP::mod(0), b.v[0]);
#pragma unroll
  for (int k = 1; k < 7; k++) r.v[k] = subc_cc(ModP::mod(k), b.v[k]);
  r.v[7] = subc(ModP::mod(7), b.v[7]);
  return r;
}
UCS_FN Fp ucs_dual(const Fp& a0, const Fp& x, const Fp& a1, const Fp& y) {
  Fp r;
  mont_mul2<ModP>(r.v, a0.v, x.v, a1.v, y.v);
  return r;
}
UCS_FN Fp ucs_const2_c(uint32_t k, uint32_t comp) {
  Fp r;
#pragma unroll
  for (int q = 0; q < 8; q++) r.v[q] = uc_const2[k][8 * comp + q];
  return r;
}
// component h of xi * a = (3 a0 - a1) + (3 a1 + a0) i, from a's own component ah and the other one ao
UCS_FN Fp ucs_xi_c(const Fp& ah, const Fp& ao, uint32_t h) {
  const Fp t = fp_add(fp_add(ah, ah), ah);
  return h ? fp_add(t, ao) : fp_sub(t, ao);
}

// One phase of one thread: component h of the lane's result. All loads of the phase happen in here, before the caller's
// barrier. `op` is the uc_ops byte: the class bits are walked first (7, 6, 5 -- ADD is two tests away, SUB2 / AXI / MUL /
// SQR three), the opcode proper only for the rest; a balanced search over all opcodes, which is what a plain switch
// compiles to, costs five dependent compare-and-branch pairs per phase, ~175 cycles of a lone warp's ~400-cycle phase.
UCS_FN Fp ucs_phase(uint32_t S, uint32_t op, uint2 w, uint32_t h) {
  const uint32_t ia = w.x >> 16, ib = w.y & 0xffffu, ic = w.y >> 16;
  if (op & 0x80u) {
    if (op & 0x40u) return fp_add(ucs_load_c(S, ia, h), ucs_load_c(S, ib, h));                                      // ADD
    if (op & 0x20u) return fp_sub(fp_sub(ucs_load_c(S, ia, h), ucs_load_c(S, ib, h)), ucs_load_c(S, ic, h));      // SUB2
    return fp_add(ucs_load_c(S, ia, h), ucs_xi_c(ucs_load_c(S, ib, h), ucs_load_c(S, ib, h ^ 1u), h));             // AXI
  }
  const uint32_t code = op & 31u;
  if (op & 0x40u) {
    if (op & 0x20u) {
      // the dual-product pass. MUL: c0 = a0 b0 + a1 (p - b1) on h = 0, c1 = a0 b1 + a1 b0 on h = 1; MULC the same with
      // b a constant; CONJMULC (a0 - a1 i)(k0 + k1 i): c0 = a0 k0 + a1 k1, c1 = a0 k1 + a1 (p - k0)
      const Fp2 a = ucs_load(S, ia);
      Fp x, yr;
      if (code == UC_MUL) {
        x = ucs_load_c(S, ib, h);
        yr = ucs_load_c(S, ib, h ^ 1u);
      } else {
        x = ucs_const2_c(ib, h);
        yr = ucs_const2_c(ib, h ^ 1u);
      }
      const bool negate = (code == UC_CONJMULC) != (h == 0);
      return ucs_dual(a.c0, x, a.c1, fp_sel(negate, ucs_neg_raw(yr), yr));
    }
    // one product per thread. SQR: (a0 + a1)(a0 - a1) on h = 0, (a0 + a0) a1 on h = 1; MULFP0 / MULFP1 / MULFPC: the
    // thread's component of a times component 0 / 1 of b / times a constant of Fp
    Fp x, y;
    if (code == UC_SQR) {
      const Fp2 a = ucs_load(S, ia);
      x = fp_add(a.c0, fp_sel(h != 0, a.c0, a.c1));
      y = fp_sel(h != 0, a.c1, fp_sub(a.c0, a.c1));
    } else {
      x = ucs_load_c(S, ia, h);
      y = code == UC_MULFPC ? uc_constp_(ib) : ucs_load_c(S, ib, code == UC_MULFP1 ? 1u : 0u);
    }
    return fp_mul_inl(x, y);
  }
  if (op & 0x20u) {
    if (code == UC_MULXI) return ucs_xi_c(ucs_load_c(S, ia, h), ucs_load_c(S, ia, h ^ 1u), h);
    const Fp a = ucs_load_c(S, ia, h), b = ucs_load_c(S, ib, h);
    if (code == UC_SUB) return fp_sub(a, b);
    const Fp s = code == UC_CYCP ? fp_add(a, b) : fp_sub(a, b);  // CYCP / CYCM: 3a +- 2b
    return fp_add(fp_add(s, s), a);
  }
  switch (code) {
    case UC_X4:
    case UC_X8:
    case UC_X3: {
      const Fp a = ucs_load_c(S, ia, h);
      const Fp d2 = fp_add(a, a);
      if (code == UC_X3) return fp_add(d2, a);
      const Fp d4 = fp_add(d2, d2);
      return code == UC_X8 ? fp_add(d4, d4) : d4;
    }
    case UC_NEG:
      return fp_neg(ucs_load_c(S, ia, h));
    case UC_CONJ: {
      const Fp a = ucs_load_c(S, ia, h);
      return h ? fp_neg(a) : a;
    }
    case UC_INV: {
      Fp2 tmp = ucs_load(S, ia);  // both threads invert (same instructions either way) and keep their component
      const Fp2 v = fp2_inv(tmp);
      return fp_sel(h != 0, v.c1, v.c0);
    }
    case UC_CHK: {
      // verdict of the G2 membership test: zero iff a == 0, b == 0 and c != 0 (see tools/microcode/gen.py)
      const bool good = fp2_is_zero(ucs_load(S, ia)) & fp2_is_zero(ucs_load(S, ib)) & !fp2_is_zero(ucs_load(S, ic));
      return good ? fp_zero() : fp_one();
    }
    default:
      return ucs_load_c(S, ia, h);  // UC_COPY
  }
}

#ifdef UC_IN_QX   // the pairing programs
// the line functions only ever use -x_P: the P slot holds (-x_P, y_P)
UCS_FN void ucs_put_inputs(uint32_t S, const G1Aff& P, const G2Aff& V) {
  ucs_store(S, ucs_slot_off(UC_IN_P), Fp2{fp_neg(P.x), P.y});
  ucs_store(S, ucs_slot_off(UC_IN_QX), V.x);
  ucs_store(S, ucs_slot_off(UC_IN_QY), V.y);
}
// e(P, V), or the all-zero element (no member of GT, so the comparison that follows fails) when V is outside G2
UCS_FN Fp12 ucs_get_output(uint32_t S) {
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
#endif

#if defined(__CUDACC__)
// Phases [ph0, ph1) of the program on the item at shared address S; all lanes of the warp call it together. Slots are
// reused in place (a phase may overwrite one of its own operands), and the two threads of a pair read each other's
// components, so every load of a phase is separated from every store by a warp barrier. Lanes a phase leaves idle and
// threads with `alive` false skip it. (Letting them run the phase on a spare slot instead, to drop the divergent region
// around the dispatch, was measured: 480 checks 2.45 -> 3.14 ms, 6,000 checks 4.07 -> 4.78 ms. Dropped.)
__device__ __forceinline__ void ucs_run(uint32_t S, int lane, uint32_t h, bool alive, int ph0, int ph1) {
  // the next phase's opcode and lane word are fetched one phase ahead; both tables end with idle phases, so the read
  // ahead needs no bound check. (An L1 prefetch of the lane words four phases ahead saves a lone warp 25 cycles per
  // phase and costs a full machine 1-2 %: not kept.)
  const uint2* pw = uc_prog + ph0 * UC_G + lane;
  uint32_t op = uc_ops[ph0];
  uint2 w = *pw;
#ifdef UCS_PROFILE  // tools/phase_profile.py: cycles per opcode as the first warp of the grid sees them
  const bool prof = blockIdx.x == 0 && threadIdx.x == 0;
  long long prof_t0 = clock64();
#endif
  for (int ph = ph0; ph < ph1; ph++) {
    pw += UC_G;
    const uint32_t op_n = uc_ops[ph + 1];
    const uint2 w_n = *pw;
    const bool act = alive && w.x != 0xffffffffu;
    Fp r;
    if (act) r = ucs_phase(S, op, w, h);
    __syncwarp();
    if (act) ucs_store_c(S, w.x & 0xffffu, h, r);
    __syncwarp();
#ifdef UCS_PROFILE
    {
      const long long t1 = clock64();
      if (prof) {
        uc_prof[op & 31u] += (unsigned long long)(t1 - prof_t0);
        uc_prof[32u + (op & 31u)] += 1ull;
      }
      prof_t0 = t1;
    }
#endif
    op = op_n;
    w = w_n;
  }
}
#endif  // __CUDACC__
