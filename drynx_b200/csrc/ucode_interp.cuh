// Interpreter of the lane-group micro-programs (tools/microcode/gen.py): slot access in shared memory and the phase loop
// with two threads per program lane. Included inside namespace drx by a translation unit that has defined
// UC_PROGRAM_INC (the generated tables it runs): ucode_split.cuh for the pairing, kernels_rp_gtlanes.cu for the Fp12
// multiplication behind fixed-base GT exponentiations.
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
  // the next phase's opcode and lane word are fetched one phase ahead; both tables end with one idle phase, so the
  // read ahead needs no bound check
  const uint2* pw = uc_prog + ph0 * UC_G + lane;
  uint32_t op = uc_ops[ph0];
  uint2 w = *pw;
  for (int ph = ph0; ph < ph1; ph++) {
    pw += UC_G;
    const uint32_t op_n = uc_ops[ph + 1];
    const uint2 w_n = *pw;
    const bool act = alive && w.x != 0xffffffffu;
    const uint32_t d = w.x & 0xffffu, ia = w.x >> 16, ib = w.y & 0xffffu, ic = w.y >> 16;
    Fp r;
    if (act) {
      switch (op) {
        case UC_ADD:
          r = fp_add(ucs_load_c(S, ia, h), ucs_load_c(S, ib, h));
          break;
        case UC_MUL: {
          const Fp2 a = ucs_load(S, ia);
          const Fp x = ucs_load_c(S, ib, h), yr = ucs_load_c(S, ib, h ^ 1u);
          r = ucs_dual(a.c0, x, a.c1, fp_sel(h != 0, yr, ucs_neg_raw(yr)));
          break;
        }
        case UC_SQR: {
          // (a0 + a1)(a0 - a1) on h = 0, (a0 + a0) a1 on h = 1: one product each
          const Fp2 a = ucs_load(S, ia);
          const Fp x = fp_add(a.c0, fp_sel(h != 0, a.c0, a.c1));
          const Fp y = fp_sel(h != 0, a.c1, fp_sub(a.c0, a.c1));
          r = fp_mul_inl(x, y);
          break;
        }
        case UC_SUB2:
          r = fp_sub(fp_sub(ucs_load_c(S, ia, h), ucs_load_c(S, ib, h)), ucs_load_c(S, ic, h));
          break;
        case UC_AXI:
          r = fp_add(ucs_load_c(S, ia, h), ucs_xi_c(ucs_load_c(S, ib, h), ucs_load_c(S, ib, h ^ 1u), h));
          break;
        case UC_SUB:
          r = fp_sub(ucs_load_c(S, ia, h), ucs_load_c(S, ib, h));
          break;
        case UC_CYCP:
        case UC_CYCM: {
          const Fp a = ucs_load_c(S, ia, h), b = ucs_load_c(S, ib, h);
          const Fp s = op == UC_CYCP ? fp_add(a, b) : fp_sub(a, b);
          r = fp_add(fp_add(s, s), a);
          break;
        }
        case UC_MULXI:
          r = ucs_xi_c(ucs_load_c(S, ia, h), ucs_load_c(S, ia, h ^ 1u), h);
          break;
        case UC_MULFP0:
        case UC_MULFP1:
          r = fp_mul(ucs_load_c(S, ia, h), ucs_load_c(S, ib, op == UC_MULFP1 ? 1u : 0u));
          break;
        case UC_X4:
        case UC_X8:
        case UC_X3: {
          const Fp a = ucs_load_c(S, ia, h);
          const Fp d2 = fp_add(a, a);
          if (op == UC_X3) {
            r = fp_add(d2, a);
          } else {
            r = fp_add(d2, d2);
            if (op == UC_X8) r = fp_add(r, r);
          }
          break;
        }
        case UC_NEG:
          r = fp_neg(ucs_load_c(S, ia, h));
          break;
        case UC_CONJMULC:
        case UC_MULC: {
          // (a0 +- a1 i)(k0 + k1 i): MULC c0 = a0 k0 + a1 (p - k1), c1 = a0 k1 + a1 k0;
          // CONJMULC c0 = a0 k0 + a1 k1, c1 = a0 k1 + a1 (p - k0)
          const Fp2 a = ucs_load(S, ia);
          const Fp x = ucs_const2_c(ib, h), yr = ucs_const2_c(ib, h ^ 1u);
          const bool negate = (op == UC_MULC) == (h == 0);
          r = ucs_dual(a.c0, x, a.c1, fp_sel(negate, ucs_neg_raw(yr), yr));
          break;
        }
        case UC_MULFPC:
          r = fp_mul(ucs_load_c(S, ia, h), uc_constp_(ib));
          break;
        case UC_CONJ: {
          const Fp a = ucs_load_c(S, ia, h);
          r = h ? fp_neg(a) : a;
          break;
        }
        case UC_INV: {
          Fp2 tmp = ucs_load(S, ia);  // both threads invert (same instructions either way) and keep their component
          const Fp2 v = fp2_inv(tmp);
          r = fp_sel(h != 0, v.c1, v.c0);
          break;
        }
        case UC_CHK: {
          // verdict of the G2 membership test: zero iff a == 0, b == 0 and c != 0 (see tools/microcode/gen.py)
          const bool good = fp2_is_zero(ucs_load(S, ia)) & fp2_is_zero(ucs_load(S, ib)) & !fp2_is_zero(ucs_load(S, ic));
          r = good ? fp_zero() : fp_one();
          break;
        }
        default:
          r = ucs_load_c(S, ia, h);  // UC_COPY
          break;
      }
    }
    __syncwarp();
    if (act) ucs_store_c(S, d, h, r);
    __syncwarp();
    op = op_n;
    w = w_n;
  }
}
