// ElGamal over edwards25519 on sm_100a: the same four operations as kernels_g1.cu -- batched
// encryption, ciphertext-vector aggregation, discrete-log table, decryption -- for UnLynx's default suite
// (libunlynx v1.4.0 lib/crypto.go over kyber group/edwards25519; neither under /root/reference). Ciphertexts are
// K || C, 32 + 32 bytes of compressed points. Integer-pipe bound like the BN256 kernels; decoding a compressed
// point costs a square root (~265 products), which is what aggregation and decryption mostly pay for.
#include "launch.h"
#include "ed25519.cuh"

namespace drx {

static inline int grid_for(size_t n, int threads) { return (int)((n + threads - 1) / threads); }

// ------------------------------------------------------------------ set-up
// points[i] <- decoded wire[i] (extended, Z = 1); malformed encodings are counted
__global__ void k_ed_from_wire(size_t n, const uint32_t* __restrict__ wire, EdExt* __restrict__ pts, int* bad) {
  const size_t i = (size_t)blockIdx.x * blockDim.x + threadIdx.x;
  if (i >= n) return;
  EdExt p;
  if (!ed_from_wire(p, wire + 8 * i)) {
    atomicAdd(bad, 1);
    p = ed_identity();
  }
  pts[i] = p;
}
// table[w][d-1] = d * 256^w * base as affine Niels triples
__global__ void __launch_bounds__(64)
k_ed_comb_table(const EdExt* __restrict__ base_pt, EdNiels* __restrict__ table) {
  const int idx = blockIdx.x * blockDim.x + threadIdx.x;
  if (idx >= DRX_COMB_WINDOWS * DRX_COMB_ENTRIES) return;
  const int w = idx / DRX_COMB_ENTRIES;
  const int d = idx - w * DRX_COMB_ENTRIES + 1;
  EdExt base = *base_pt;
  for (int s = 0; s < 8 * w; s++) base = ed_dbl(base);
  EdExt acc = ed_identity();
  for (int bit = 7; bit >= 0; bit--) {
    acc = ed_dbl(acc);
    if ((d >> bit) & 1) acc = ed_add(acc, base);
  }
  table[idx] = ed_to_niels_zinv(acc, fe_inv(acc.Z));
}

// ------------------------------------------------------------------ encryption: K = r*B, C = m*B + r*P
__global__ void __launch_bounds__(128)
k_ed_encrypt(size_t n, const int64_t* __restrict__ m, const uint32_t* __restrict__ r, const EdNiels* __restrict__ tabB,
             const EdNiels* __restrict__ tabP, uint32_t* __restrict__ ct) {
  const size_t i = (size_t)blockIdx.x * blockDim.x + threadIdx.x;
  if (i >= n) return;
  const Scalar ri = ed_sc_from_wire(r + 8 * i);
  const EdExt K = ed_comb_mul(tabB, ri);
  EdExt C = ed_comb_mul(tabP, ri);
  const int64_t mi = m[i];
  const uint64_t am = mi < 0 ? (uint64_t)0 - (uint64_t)mi : (uint64_t)mi;
  for (int w = 0; w < 8; w++) {  // |m| * B over its eight windows, entries negated for m < 0
    const uint32_t d = (uint32_t)(am >> (8 * w)) & 0xffu;
    if (d) {
      const EdNiels q = tabB[w * DRX_COMB_ENTRIES + (int)d - 1];
      C = ed_add_niels(C, mi < 0 ? ed_niels_neg(q) : q);
    }
  }
  ed_pair_to_wire(ct + 16 * i, K, C);
}
// Small batches: eight lanes per ciphertext (same split as k_encrypt_lanes in kernels_g1.cu).
__device__ __forceinline__ EdExt ed_shfl_xor(const EdExt& p, int m, uint32_t mask) {
  EdExt r;
#pragma unroll
  for (int k = 0; k < 8; k++) {
    r.X.v[k] = __shfl_xor_sync(mask, p.X.v[k], m);
    r.Y.v[k] = __shfl_xor_sync(mask, p.Y.v[k], m);
    r.Z.v[k] = __shfl_xor_sync(mask, p.Z.v[k], m);
    r.T.v[k] = __shfl_xor_sync(mask, p.T.v[k], m);
  }
  return r;
}
__global__ void __launch_bounds__(64)
k_ed_encrypt_lanes(size_t n, const int64_t* __restrict__ m, const uint32_t* __restrict__ r,
                   const EdNiels* __restrict__ tabB, const EdNiels* __restrict__ tabP, uint32_t* __restrict__ ct) {
  const size_t i = ((size_t)blockIdx.x * blockDim.x + threadIdx.x) >> 3;
  const int lane = threadIdx.x & 7, sub = lane & 3;
  const uint32_t mask = 0xFFu << (threadIdx.x & 24u);
  if (i >= n) return;  // whole groups of eight leave together
  const Scalar ri = ed_sc_from_wire(r + 8 * i);
  const EdNiels* tab = lane < 4 ? tabB : tabP;
  EdExt acc = ed_identity();
  for (int w = sub; w < DRX_COMB_WINDOWS; w += 4) {
    const uint32_t d = (ri.v[w >> 2] >> ((w & 3) * 8)) & 0xffu;
    if (d) acc = ed_add_niels(acc, tab[w * DRX_COMB_ENTRIES + (int)d - 1]);
  }
  if (lane >= 4) {
    const int64_t mi = m[i];
    const uint64_t am = mi < 0 ? (uint64_t)0 - (uint64_t)mi : (uint64_t)mi;
    for (int w = sub; w < 8; w += 4) {
      const uint32_t d = (uint32_t)(am >> (8 * w)) & 0xffu;
      if (d) {
        const EdNiels q = tabB[w * DRX_COMB_ENTRIES + (int)d - 1];
        acc = ed_add_niels(acc, mi < 0 ? ed_niels_neg(q) : q);
      }
    }
  }
#pragma unroll
  for (int s = 1; s <= 2; s <<= 1) acc = ed_add(acc, ed_shfl_xor(acc, s, mask));
  const EdExt other = ed_shfl_xor(acc, 4, mask);
  if (lane == 0) ed_pair_to_wire(ct + 16 * i, acc, other);
}

// ------------------------------------------------------------------ aggregation: column sums over data providers
// Stage 1: thread (slice s, column j) decodes and folds providers s, s+S, ...; stage 2: one warp per column folds
// the slices by a butterfly and serialises.
struct EdAggPartial {
  EdExt K, C;
};
__global__ void __launch_bounds__(128)
k_ed_agg_partial(size_t ndp, size_t v, const uint32_t* __restrict__ cts, int slices, EdAggPartial* __restrict__ part,
                 int* bad) {
  const size_t t = (size_t)blockIdx.x * blockDim.x + threadIdx.x;
  if (t >= v * (size_t)slices) return;
  const size_t j = t % v;
  const int s = (int)(t / v);
  EdExt aK = ed_identity(), aC = ed_identity();
  int nbad = 0;
  for (size_t d = s; d < ndp; d += slices) {
    const uint32_t* w = cts + (d * v + j) * 16;
    EdExt K, C;
    const bool okK = ed_from_wire(K, w), okC = ed_from_wire(C, w + 8);
    nbad += !okK + !okC;
    if (okK) aK = ed_add(aK, K);
    if (okC) aC = ed_add(aC, C);
  }
  part[t].K = aK;
  part[t].C = aC;
  if (nbad) atomicAdd(bad, nbad);
}
__global__ void __launch_bounds__(128)
k_ed_agg_final(size_t v, int slices, const EdAggPartial* __restrict__ part, uint32_t* __restrict__ out) {
  const size_t j = ((size_t)blockIdx.x * blockDim.x + threadIdx.x) >> 5;
  const int lane = threadIdx.x & 31;
  if (j >= v) return;  // whole warps leave together
  EdExt aK = ed_identity(), aC = ed_identity();
  for (int s = lane; s < slices; s += 32) {
    aK = ed_add(aK, part[(size_t)s * v + j].K);
    aC = ed_add(aC, part[(size_t)s * v + j].C);
  }
  const int levels = slices < 32 ? slices : 32;
  for (int m = 1; m < levels; m <<= 1) {
    aK = ed_add(aK, ed_shfl_xor(aK, m, 0xffffffffu));
    aC = ed_add(aC, ed_shfl_xor(aC, m, 0xffffffffu));
  }
  if (lane == 0) ed_pair_to_wire(out + 16 * j, aK, aC);
}

// ------------------------------------------------------------------ discrete-log table and decryption
// Open addressing, linear probing. Key = canonical y of i*B (8 words); value word = (i << 1) | parity(x),
// 0xffffffff = empty. -i*B shares y and has the other x parity.
#define DRX_ED_DLOG_EMPTY 0xffffffffu
#define DRX_ED_DLOG_CHUNK 16

__device__ __forceinline__ uint32_t ed_dlog_hash(const Fe& y) {
  return (y.v[0] ^ (y.v[1] * 0x9e3779b1u) ^ (y.v[2] * 0x85ebca6bu)) & (kDlogCap - 1);
}
__global__ void __launch_bounds__(64)
k_ed_dlog_build(const EdNiels* __restrict__ tabB, Fe* __restrict__ keys, uint32_t* __restrict__ vals) {
  const int t = blockIdx.x * blockDim.x + threadIdx.x;
  const int64_t first = 1 + (int64_t)t * DRX_ED_DLOG_CHUNK;
  if (first >= kMaxHomomorphicInt) return;
  const EdNiels B = tabB[0];  // 1 * 256^0 * B
  EdExt pts[DRX_ED_DLOG_CHUNK];
  Fe pre[DRX_ED_DLOG_CHUNK];
  EdExt cur = ed_comb_mul(tabB, sc_from_u64((uint64_t)first), 3);
  int cnt = 0;
  for (; cnt < DRX_ED_DLOG_CHUNK && first + cnt < kMaxHomomorphicInt; cnt++) {
    if (cnt) cur = ed_add_niels(cur, B);
    pts[cnt] = cur;
    pre[cnt] = cnt ? fe_mul(pre[cnt - 1], cur.Z) : cur.Z;
  }
  Fe inv = fe_inv(pre[cnt - 1]);
  for (int k = cnt - 1; k >= 0; k--) {
    const Fe zinv = k ? fe_mul(inv, pre[k - 1]) : inv;
    if (k) inv = fe_mul(inv, pts[k].Z);
    const Fe x = fe_mul(pts[k].X, zinv);
    const Fe y = fe_canon(fe_mul(pts[k].Y, zinv));
    const uint32_t val = ((uint32_t)(first + k) << 1) | fe_is_negative(x);
    uint32_t h = ed_dlog_hash(y);
    while (atomicCAS(&vals[h], DRX_ED_DLOG_EMPTY, val) != DRX_ED_DLOG_EMPTY) h = (h + 1) & (kDlogCap - 1);
    keys[h] = y;
  }
}
// i with i*B == M for |i| < kMaxHomomorphicInt, else INT64_MIN
__device__ __forceinline__ int64_t ed_dlog_lookup(const Fe& x, const Fe& y, const Fe* __restrict__ keys,
                                                  const uint32_t* __restrict__ vals) {
  if (fe_is_zero(x)) {
    // x = 0: the identity (y = 1) is 0*B; (0, -1) has order two and is no multiple below the bound
    const Fe one = fe_one();
    bool is_one = true;
    for (int k = 0; k < 8; k++) is_one &= y.v[k] == one.v[k];
    return is_one ? 0 : INT64_MIN;
  }
  uint32_t h = ed_dlog_hash(y);
  for (;;) {
    const uint32_t val = vals[h];
    if (val == DRX_ED_DLOG_EMPTY) return INT64_MIN;
    const Fe k = keys[h];
    bool same = true;
    for (int q = 0; q < 8; q++) same &= k.v[q] == y.v[q];
    if (same) {
      const int64_t i = (int64_t)(val >> 1);
      return ((val & 1u) == fe_is_negative(x)) ? i : -i;
    }
    h = (h + 1) & (kDlogCap - 1);
  }
}
__global__ void __launch_bounds__(128)
k_ed_decrypt(size_t n, const uint32_t* __restrict__ sk, const uint32_t* __restrict__ ct, const Fe* __restrict__ keys,
             const uint32_t* __restrict__ vals, int64_t* __restrict__ out, int* bad) {
  const size_t i = (size_t)blockIdx.x * blockDim.x + threadIdx.x;
  if (i >= n) return;
  const Scalar k = ed_sc_from_wire(sk);
  EdExt K, C;
  int nbad = !ed_from_wire(K, ct + 16 * i);
  nbad += !ed_from_wire(C, ct + 16 * i + 8);
  if (nbad) {
    atomicAdd(bad, nbad);
    out[i] = INT64_MIN;
    return;
  }
  const EdExt S = ed_window_mul(K, k);
  const EdExt M = ed_add(C, ed_neg(S));
  const Fe zinv = fe_inv(M.Z);
  const Fe x = fe_mul(M.X, zinv);
  const Fe y = fe_canon(fe_mul(M.Y, zinv));
  out[i] = ed_dlog_lookup(x, y, keys, vals);
}

// ------------------------------------------------------------------ diagnostics: k * P per element
__global__ void __launch_bounds__(64)
k_ed_mul(size_t n, const uint32_t* __restrict__ k, const uint32_t* __restrict__ pt, uint32_t* __restrict__ out, int* bad) {
  const size_t i = (size_t)blockIdx.x * blockDim.x + threadIdx.x;
  if (i >= n) return;
  EdExt p;
  if (!ed_from_wire(p, pt + 8 * i)) {
    atomicAdd(bad, 1);
    p = ed_identity();
  }
  ed_to_wire(out + 8 * i, ed_window_mul(p, ed_sc_from_wire(k + 8 * i)));
}

// ------------------------------------------------------------------ launchers
int launch_ed_from_wire(cudaStream_t st, size_t n, const uint8_t* d_wire, void* d_pts, int* d_bad) {
  if (!n) return 0;
  k_ed_from_wire<<<grid_for(n, 64), 64, 0, st>>>(n, (const uint32_t*)d_wire, (EdExt*)d_pts, d_bad);
  return 1;
}
int launch_ed_comb_table(cudaStream_t st, const void* d_base_pt, void* d_table) {
  k_ed_comb_table<<<grid_for(kCombPoints, 64), 64, 0, st>>>((const EdExt*)d_base_pt, (EdNiels*)d_table);
  return 1;
}
int launch_ed_encrypt(cudaStream_t st, size_t n, const int64_t* d_m, const uint8_t* d_r, const void* d_tabB,
                      const void* d_tabP, uint8_t* d_ct) {
  if (!n) return 0;
  if (n <= kEncryptLanesMaxItems)
    k_ed_encrypt_lanes<<<grid_for(n * 8, 64), 64, 0, st>>>(n, d_m, (const uint32_t*)d_r, (const EdNiels*)d_tabB,
                                                            (const EdNiels*)d_tabP, (uint32_t*)d_ct);
  else
    k_ed_encrypt<<<grid_for(n, 128), 128, 0, st>>>(n, d_m, (const uint32_t*)d_r, (const EdNiels*)d_tabB,
                                                  (const EdNiels*)d_tabP, (uint32_t*)d_ct);
  return 1;
}
size_t ed_aggregate_scratch_bytes(size_t ndp, size_t v, int* slices_out) {
  // enough (slice, column) threads to fill the machine, at most one slice per provider
  int slices = 1;
  const size_t target = 148 * 1024;
  while ((size_t)slices * 2 <= ndp && v * (size_t)slices * 2 <= target && slices < 1024) slices *= 2;
  if (slices_out) *slices_out = slices;
  return v * (size_t)slices * sizeof(EdAggPartial);
}
int launch_ed_aggregate(cudaStream_t st, size_t ndp, size_t v, const uint8_t* d_cts, void* d_scratch, int slices,
                        uint8_t* d_out, int* d_bad) {
  if (!v) return 0;
  k_ed_agg_partial<<<grid_for(v * (size_t)slices, 128), 128, 0, st>>>(ndp, v, (const uint32_t*)d_cts, slices,
                                                                      (EdAggPartial*)d_scratch, d_bad);
  k_ed_agg_final<<<grid_for(v * 32, 128), 128, 0, st>>>(v, slices, (const EdAggPartial*)d_scratch, (uint32_t*)d_out);
  return 2;
}
int launch_ed_dlog_build(cudaStream_t st, const void* d_tabB, void* d_keys, uint32_t* d_vals) {
  cudaMemsetAsync(d_vals, 0xff, (size_t)kDlogCap * sizeof(uint32_t), st);
  const size_t threads = (kMaxHomomorphicInt + DRX_ED_DLOG_CHUNK - 1) / DRX_ED_DLOG_CHUNK;
  k_ed_dlog_build<<<grid_for(threads, 64), 64, 0, st>>>((const EdNiels*)d_tabB, (Fe*)d_keys, d_vals);
  return 1;
}
int launch_ed_decrypt(cudaStream_t st, size_t n, const uint8_t* d_sk, const uint8_t* d_ct, const void* d_keys,
                      const uint32_t* d_vals, int64_t* d_out, int* d_bad) {
  if (!n) return 0;
  k_ed_decrypt<<<grid_for(n, 128), 128, 0, st>>>(n, (const uint32_t*)d_sk, (const uint32_t*)d_ct, (const Fe*)d_keys,
                                                d_vals, d_out, d_bad);
  return 1;
}
int launch_ed_mul(cudaStream_t st, size_t n, const uint8_t* d_k, const uint8_t* d_pt, uint8_t* d_out, int* d_bad) {
  if (!n) return 0;
  k_ed_mul<<<grid_for(n, 64), 64, 0, st>>>(n, (const uint32_t*)d_k, (const uint32_t*)d_pt, (uint32_t*)d_out, d_bad);
  return 1;
}

}  // namespace drx
