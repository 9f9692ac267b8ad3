// ElGamal over BN256-G1 on sm_100a: batched encryption, ciphertext-vector aggregation, discrete-log table and
// decryption. Replaces libunlynx v1.4.0 lib/crypto.go (encryptPoint, IntToPoint, CipherVector.Add,
// decryptPoint, discreteLog, CreateDecryptionTable; not in /root/reference) at the call sites listed in
// SURVEY.md §8 rows a-1, a-3, a-4, a-5. One thread per ciphertext; fixed-base multiplications walk 8-bit comb
// tables that stay L2-resident (522 KB each); all kernels are integer-pipe bound.
#include <mutex>

#include "launch.h"
#include "keccak.cuh"
#include "wire.cuh"

namespace drx {

static inline int grid_for(size_t n, int threads) { return (int)((n + threads - 1) / threads); }

// ------------------------------------------------------------------ set-up kernels
__global__ void k_g1_from_wire(size_t n, const uint32_t* __restrict__ wire, G1Aff* __restrict__ aff, int* bad) {
  const size_t i = (size_t)blockIdx.x * blockDim.x + threadIdx.x;
  if (i >= n) return;
  G1Aff p;
  if (!g1_from_wire(p, wire + 16 * i)) atomicAdd(bad, 1);
  aff[i] = p;
}

// table[b][w][d-1] = d * 2^(8w) * base_b
template <class F>
__global__ void k_comb_table(int nbases, const Aff<F>* __restrict__ bases, Aff<F>* __restrict__ tables) {
  const int idx = blockIdx.x * blockDim.x + threadIdx.x;
  if (idx >= nbases * DRX_COMB_WINDOWS * DRX_COMB_ENTRIES) return;
  const int b = idx / (DRX_COMB_WINDOWS * DRX_COMB_ENTRIES);
  const int rem = idx - b * (DRX_COMB_WINDOWS * DRX_COMB_ENTRIES);
  const int w = rem / DRX_COMB_ENTRIES;
  const int d = rem - w * DRX_COMB_ENTRIES + 1;
  Jac<F> base = jac_from_aff(bases[b]);
  for (int s = 0; s < 8 * w; s++) base = jac_dbl(base);
  Jac<F> acc = jac_inf<F>();
  for (int bit = 7; bit >= 0; bit--) {
    acc = jac_dbl(acc);
    if ((d >> bit) & 1) acc = jac_add(acc, base);
  }
  tables[idx] = jac_to_aff(acc);
}
template __global__ void k_comb_table<Fp>(int, const Aff<Fp>*, Aff<Fp>*);

int launch_g1_from_wire(cudaStream_t st, size_t n, const uint8_t* d_wire, void* d_aff, int* d_bad) {
  if (!n) return 0;
  k_g1_from_wire<<<grid_for(n, 128), 128, 0, st>>>(n, (const uint32_t*)d_wire, (G1Aff*)d_aff, d_bad);
  return 1;
}
int launch_comb_table_g1(cudaStream_t st, int nbases, const void* d_bases_aff, void* d_tables) {
  const size_t n = (size_t)nbases * kCombPoints;
  if (!n) return 0;
  k_comb_table<Fp><<<grid_for(n, 64), 64, 0, st>>>(nbases, (const G1Aff*)d_bases_aff, (G1Aff*)d_tables);
  return 1;
}

// ------------------------------------------------------------------ encryption (a-1)
// Writes two affine points sharing one field inversion.
__device__ __forceinline__ void store_pair_wire(uint32_t* out, const G1Jac& A, const G1Jac& B) {
  const bool ia = jac_is_inf(A), ib = jac_is_inf(B);
  G1Aff a, b;
  if (ia | ib) {
    a = jac_to_aff(A);
    b = jac_to_aff(B);
  } else {
    const Fp inv = fp_inv(fp_mul(A.z, B.z));
    a = jac_to_aff_with_zinv(A, fp_mul(inv, B.z));
    b = jac_to_aff_with_zinv(B, fp_mul(inv, A.z));
  }
  g1_to_wire(out, a);
  g1_to_wire(out + 16, b);
}

__global__ void __launch_bounds__(128)
k_encrypt(size_t n, const int64_t* __restrict__ m, const uint32_t* __restrict__ r, const G1Aff* __restrict__ tabB,
          const G1Aff* __restrict__ tabP, uint32_t* __restrict__ ct) {
  const size_t i = (size_t)blockIdx.x * blockDim.x + threadIdx.x;
  if (i >= n) return;
  const Scalar ri = sc_from_wire(r + 8 * i);
  const G1Jac K = comb_mul(tabB, ri);
  const G1Jac S = comb_mul(tabP, ri);
  const int64_t mi = m[i];
  const uint64_t am = mi < 0 ? (uint64_t)0 - (uint64_t)mi : (uint64_t)mi;
  G1Jac M = comb_mul(tabB, sc_from_u64(am), 8);
  if (mi < 0) M = jac_neg(M);
  const G1Jac C = jac_add(S, M);
  store_pair_wire(ct + 32 * i, K, C);
}

// libunlynx.EncryptScalar: the message is a full scalar (32 comb windows instead of 8)
__global__ void __launch_bounds__(128)
k_encrypt_scalar(size_t n, const uint32_t* __restrict__ m, const uint32_t* __restrict__ r,
                 const G1Aff* __restrict__ tabB, const G1Aff* __restrict__ tabP, uint32_t* __restrict__ ct) {
  const size_t i = (size_t)blockIdx.x * blockDim.x + threadIdx.x;
  if (i >= n) return;
  const Scalar ri = sc_from_wire(r + 8 * i);
  const G1Jac K = comb_mul(tabB, ri);
  const G1Jac S = comb_mul(tabP, ri);
  const G1Jac M = comb_mul(tabB, sc_from_wire(m + 8 * i));
  store_pair_wire(ct + 32 * i, K, jac_add(S, M));
}
int launch_encrypt_scalar(cudaStream_t st, size_t n, const uint8_t* d_m, const uint8_t* d_r, const void* d_tabB,
                          const void* d_tabP, uint8_t* d_ct) {
  if (!n) return 0;
  k_encrypt_scalar<<<grid_for(n, 128), 128, 0, st>>>(n, (const uint32_t*)d_m, (const uint32_t*)d_r, (const G1Aff*)d_tabB,
                                                    (const G1Aff*)d_tabP, (uint32_t*)d_ct);
  return 1;
}

// Small batches: eight lanes per ciphertext. Lanes 0-3 walk every fourth window of r*B, lanes 4-7 every fourth
// window of r*K plus their share of the message windows of m*B; a two-level butterfly of full additions joins
// each half, lane 0 fetches C from lane 4 and serialises both points with one shared inversion. Each lane runs
// 8-10 dependent mixed additions instead of 66-72; the sums are the same group elements, so the bytes match the
// one-thread kernel's.
__device__ __forceinline__ G1Jac jac_shfl_xor_mask(const G1Jac& p, int m, uint32_t mask) {
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
k_encrypt_lanes(size_t n, const int64_t* __restrict__ m, const uint32_t* __restrict__ r,
                const G1Aff* __restrict__ tabB, const G1Aff* __restrict__ tabP, uint32_t* __restrict__ ct) {
  const size_t i = ((size_t)blockIdx.x * blockDim.x + threadIdx.x) >> 3;
  const int lane = threadIdx.x & 7, sub = lane & 3;
  const uint32_t mask = 0xFFu << (threadIdx.x & 24u);
  if (i >= n) return;  // whole groups of eight leave together
  const Scalar ri = sc_from_wire(r + 8 * i);
  const G1Aff* tab = lane < 4 ? tabB : tabP;
  G1Jac acc = jac_inf<Fp>();
  for (int w = sub; w < DRX_COMB_WINDOWS; w += 4) {
    const uint32_t d = (ri.v[w >> 2] >> ((w & 3) * 8)) & 0xffu;
    if (d) acc = jac_add_aff(acc, tab[w * DRX_COMB_ENTRIES + (int)d - 1]);
  }
  if (lane >= 4) {
    const int64_t mi = m[i];
    const uint64_t am = mi < 0 ? (uint64_t)0 - (uint64_t)mi : (uint64_t)mi;
    for (int w = sub; w < 8; w += 4) {
      const uint32_t d = (uint32_t)(am >> (8 * w)) & 0xffu;
      if (d) {
        const G1Aff q = tabB[w * DRX_COMB_ENTRIES + (int)d - 1];
        acc = jac_add_aff(acc, mi < 0 ? aff_neg(q) : q);
      }
    }
  }
#pragma unroll
  for (int s = 1; s <= 2; s <<= 1) {
    const G1Jac o = jac_shfl_xor_mask(acc, s, mask);
    acc = (lane & s) ? jac_add(o, acc) : jac_add(acc, o);
  }
  const G1Jac other = jac_shfl_xor_mask(acc, 4, mask);
  if (lane == 0) store_pair_wire(ct + 32 * i, acc, other);
}

int launch_encrypt(cudaStream_t st, size_t n, const int64_t* d_m, const uint8_t* d_r, const void* d_tabB,
                   const void* d_tabP, uint8_t* d_ct) {
  if (!n) return 0;
  if (n <= kEncryptLanesMaxItems) {
    k_encrypt_lanes<<<grid_for(n * 8, 64), 64, 0, st>>>(n, d_m, (const uint32_t*)d_r, (const G1Aff*)d_tabB,
                                                         (const G1Aff*)d_tabP, (uint32_t*)d_ct);
    return 1;
  }
  k_encrypt<<<grid_for(n, 128), 128, 0, st>>>(n, d_m, (const uint32_t*)d_r, (const G1Aff*)d_tabB,
                                             (const G1Aff*)d_tabP, (uint32_t*)d_ct);
  return 1;
}

// ------------------------------------------------------------------ aggregation (a-3)
// A ciphertext vector is 2v points of 64 bytes back to back (K_0 C_0 K_1 C_1 ...), and K and C sum independently, so
// the unit of work is a POINT COLUMN jp in [0, 2v). Stage 1: thread (slice s, point column jp) folds data providers
// d = s, s + S, ... of its column with mixed additions straight from wire bytes (one accumulator per thread keeps
// the kernel at five blocks per SM; the next point is fetched while the current one is added). Stage 2: one warp per
// point column folds the S partial sums and serialises.
#ifndef DRX_AGG_MIN_BLOCKS
#define DRX_AGG_MIN_BLOCKS 5
#endif
__global__ void __launch_bounds__(128, DRX_AGG_MIN_BLOCKS)
k_agg_partial(size_t ndp, size_t v2, const uint32_t* __restrict__ cts, int slices, G1Jac* __restrict__ part, int* bad) {
  const size_t t = (size_t)blockIdx.x * blockDim.x + threadIdx.x;
  if (t >= v2 * (size_t)slices) return;
  const size_t jp = t % v2;
  const size_t s = t / v2;
  const size_t stride = (size_t)slices * v2 * 4;  // uint4 words between consecutive data providers of this thread
  const uint4* w = reinterpret_cast<const uint4*>(cts) + (s * v2 + jp) * 4;
  G1Jac acc = jac_inf<Fp>();
  int nbad = 0;
  uint4 cur[4];
  if (s < ndp) {
#pragma unroll
    for (int k = 0; k < 4; k++) cur[k] = __ldg(w + k);
  }
  for (size_t d = s; d < ndp; d += slices) {
    uint4 nxt[4];
    w += stride;
    if (d + slices < ndp) {
#pragma unroll
      for (int k = 0; k < 4; k++) nxt[k] = __ldg(w + k);
    }
    uint32_t words[16];
#pragma unroll
    for (int k = 0; k < 4; k++) {
      words[4 * k] = cur[k].x; words[4 * k + 1] = cur[k].y; words[4 * k + 2] = cur[k].z; words[4 * k + 3] = cur[k].w;
    }
    G1Aff P;
    nbad += !g1_from_wire(P, words);  // a malformed point enters the sum as the point at infinity and is counted
    acc = jac_add_aff(acc, P);
#pragma unroll
    for (int k = 0; k < 4; k++) cur[k] = nxt[k];
  }
  part[t] = acc;
  if (nbad) atomicAdd(bad, nbad);
}

// Stage 2, L = 4 / 8 / 16 / 32 lanes per point column (the smallest that covers the slices, so that a warp serves up to
// eight columns and the multiply pipe is not spent on lanes that hold nothing): lane l of a column's group folds
// slices l, l + L, ...; a log2(L)-level butterfly of full additions leaves the column sum in every lane of the group;
// lane 0 converts to affine and serialises.
__device__ __forceinline__ G1Jac jac_shfl_xor(const G1Jac& p, int m) {
  G1Jac r;
#pragma unroll
  for (int k = 0; k < 8; k++) {
    r.x.v[k] = __shfl_xor_sync(0xffffffffu, p.x.v[k], m);
    r.y.v[k] = __shfl_xor_sync(0xffffffffu, p.y.v[k], m);
    r.z.v[k] = __shfl_xor_sync(0xffffffffu, p.z.v[k], m);
  }
  return r;
}
// (measured at 1,000 x 1,000: 32 lanes per column 0.336 ms, 16 lanes 0.142, 8 lanes 0.084 -- the multiply pipe is charged per
// warp instruction, so fewer, fuller groups win until the serial walk per lane outgrows the butterfly it saves)
static inline int agg_group_lanes(int slices) {
  return slices > 160 ? 32 : slices > 64 ? 16 : slices > 4 ? 8 : 4;
}
__global__ void __launch_bounds__(128)
k_agg_final(size_t v2, int slices, int L, const G1Jac* __restrict__ part, uint32_t* __restrict__ out) {
  const size_t gt = (size_t)blockIdx.x * blockDim.x + threadIdx.x;
  const size_t jp = gt / L;
  const int lane = (int)(gt % L);
  const bool live = jp < v2;  // every lane of a warp takes part in the shuffles
  G1Jac acc = jac_inf<Fp>();
  if (live)
    for (int s = lane; s < slices; s += L) acc = jac_add(acc, part[(size_t)s * v2 + jp]);
  for (int m = L >> 1; m >= 1; m >>= 1) {
    if (m < slices) {  // uniform across the warp
      const G1Jac o = jac_shfl_xor(acc, m);
      // same operand order on both partners keeps their copies bit-identical
      acc = (lane & m) == 0 ? jac_add(acc, o) : jac_add(o, acc);
    }
  }
  if (live && lane == 0) g1_to_wire(out + 16 * jp, jac_to_aff(acc));
}

// SM count of the current device (cached per ordinal; the slicing below sizes stage 1 for the GPU it runs on)
static int agg_sm_count() {
  static std::mutex mu;
  static int sms[64] = {0};
  int dev = 0;
  if (cudaGetDevice(&dev) != cudaSuccess || dev < 0 || dev >= 64) return 148;
  std::lock_guard<std::mutex> lk(mu);
  if (!sms[dev] && cudaDeviceGetAttribute(&sms[dev], cudaDevAttrMultiProcessorCount, dev) != cudaSuccess) sms[dev] = 148;
  return sms[dev];
}
// Slices for stage 1: enough (slice, point column) threads to fill every SM with DRX_AGG_MIN_BLOCKS blocks of 128, a
// whole number of such waves when there is more work than that, never more slices than data providers.
static int agg_slices_for(size_t ndp, size_t v, size_t cap) {
  if (!ndp || !v) return 1;
  const size_t resident = (size_t)agg_sm_count() * DRX_AGG_MIN_BLOCKS * 128;
  size_t want = resident / (2 * v);  // rounded down: one thread too many costs a whole second wave
  // ... but at least eight data providers per thread: below that the second stage (a few dependent full additions
  // and an inversion per column whatever the slice count) costs more than the extra parallelism returns
  if (want > (ndp + 7) / 8) want = (ndp + 7) / 8;
  if (want > cap) want = cap;
  if (want < 1) want = 1;
  return (int)want;
}
size_t aggregate_scratch_bytes(size_t ndp, size_t v, int* slices_out) {
  const int slices = agg_slices_for(ndp, v, kAggMaxSlices);
  if (slices_out) *slices_out = slices;
  return (size_t)slices * 2 * v * sizeof(G1Jac);
}

int launch_agg_partial(cudaStream_t st, size_t ndp, size_t v, const uint8_t* d_cts, void* d_scratch, int slices,
                       int* d_bad) {
  if (!v) return 0;
  k_agg_partial<<<grid_for(2 * v * (size_t)slices, 128), 128, 0, st>>>(ndp, 2 * v, (const uint32_t*)d_cts, slices,
                                                                      (G1Jac*)d_scratch, d_bad);
  return 1;
}
int launch_agg_final(cudaStream_t st, size_t v, int slices, const void* d_scratch, uint8_t* d_out) {
  if (!v) return 0;
  const int L = agg_group_lanes(slices);
  k_agg_final<<<grid_for(2 * v * L, 128), 128, 0, st>>>(2 * v, slices, L, (const G1Jac*)d_scratch, (uint32_t*)d_out);
  return 1;
}

// ------------------------------------------------------------------ aggregation without any allocation
// Same two kernels as above, but the partial sums live in a fixed array that is part of the loaded module, so a
// caller that must not (or cannot) allocate device memory can still aggregate. The array is shared by every
// call on a GPU: it holds whatever the previous call left, and two calls in flight at once would overwrite each
// other's partial sums.
#define DRX_NOALLOC_MAX_PARTIALS 16384  // (slice, column) pairs: 16384 x 2 points x 96 B = 3 MiB per GPU
static __device__ __align__(256) unsigned char g_noalloc_scratch[256 + DRX_NOALLOC_MAX_PARTIALS * 2 * sizeof(G1Jac)];

// In host code `g_noalloc_scratch` is only a handle for the symbol, not a device address; the address each GPU
// gave the array is looked up once per device (with that device current) and remembered.
static std::mutex g_noalloc_mu;
static void* g_noalloc_addr[64];

static cudaError_t noalloc_scratch(void** out) {
  int dev = -1;
  cudaError_t e = cudaGetDevice(&dev);
  if (e != cudaSuccess) return e;
  if (dev < 0 || dev >= 64) return cudaErrorInvalidDevice;
  std::lock_guard<std::mutex> lk(g_noalloc_mu);
  if (!g_noalloc_addr[dev]) {
    void* p = nullptr;
    e = cudaGetSymbolAddress((void**)&p, g_noalloc_scratch);
    if (e != cudaSuccess) return e;
    g_noalloc_addr[dev] = p;
  }
  *out = g_noalloc_addr[dev];
  return cudaSuccess;
}

size_t aggregate_static_max_columns() { return DRX_NOALLOC_MAX_PARTIALS; }

// Returns 0, -5 (v beyond the fixed array), -2 (malformed input points; *n_bad says how many) or a CUDA error code.
int launch_aggregate_static(cudaStream_t st, size_t ndp, size_t v, const uint8_t* d_cts, uint8_t* d_out, int* n_bad,
                            int* launches) {
  *n_bad = 0;
  *launches = 0;
  if (!v) return 0;
  if (v > (size_t)DRX_NOALLOC_MAX_PARTIALS) return -5;
  const size_t slices = (size_t)agg_slices_for(ndp, v, (size_t)DRX_NOALLOC_MAX_PARTIALS / v);
  void* base = nullptr;
  cudaError_t e = noalloc_scratch(&base);
  if (e != cudaSuccess) return (int)e;
  int* d_bad = (int*)base;                                   // first 256 bytes: the malformed-point counter
  G1Jac* part = (G1Jac*)((unsigned char*)base + 256);
  e = cudaMemsetAsync(d_bad, 0, 256, st);                    // the counter must start at zero; the partials are
  if (e != cudaSuccess) return (int)e;                       // all rewritten by stage 1 before stage 2 reads them
  k_agg_partial<<<grid_for(2 * v * slices, 128), 128, 0, st>>>(ndp, 2 * v, (const uint32_t*)d_cts, (int)slices, part, d_bad);
  const int L = agg_group_lanes((int)slices);
  k_agg_final<<<grid_for(2 * v * L, 128), 128, 0, st>>>(2 * v, (int)slices, L, part, (uint32_t*)d_out);
  *launches = 2;
  e = cudaGetLastError();
  if (e != cudaSuccess) return (int)e;
  e = cudaMemcpyAsync(n_bad, d_bad, sizeof(int), cudaMemcpyDeviceToHost, st);
  if (e != cudaSuccess) return (int)e;
  e = cudaStreamSynchronize(st);
  if (e != cudaSuccess) return (int)e;
  return *n_bad ? -2 : 0;
}

// ------------------------------------------------------------------ plaintext reduction (a-2)
// lib/encoding/sum.go:19-23, mean.go:19-24, variance.go:19-27: per data provider sum, count and sum of squares
// of its int64 rows. One block per provider; HBM-bound (8 bytes per row).
__global__ void __launch_bounds__(256)
k_reduce_rows(size_t nrows, const int64_t* __restrict__ rows, int mode, int64_t* __restrict__ out) {
  const size_t d = blockIdx.x;
  const int64_t* r = rows + d * nrows;
  int64_t s = 0, q = 0;
  for (size_t i = threadIdx.x; i < nrows; i += blockDim.x) {
    const int64_t x = r[i];
    s += x;
    q += x * x;
  }
  for (int o = 16; o > 0; o >>= 1) {
    s += __shfl_down_sync(0xffffffffu, s, o);
    q += __shfl_down_sync(0xffffffffu, q, o);
  }
  __shared__ int64_t ss[8], qq[8];
  if ((threadIdx.x & 31) == 0) {
    ss[threadIdx.x >> 5] = s;
    qq[threadIdx.x >> 5] = q;
  }
  __syncthreads();
  if (threadIdx.x == 0) {
    for (int w = 1; w < 8; w++) {
      s += ss[w];
      q += qq[w];
    }
    const int64_t n = (int64_t)nrows;
    if (mode == 0) {
      out[d] = s;
    } else if (mode == 1) {
      out[2 * d] = s;
      out[2 * d + 1] = n;
    } else if (mode == 2) {
      out[3 * d] = s;
      out[3 * d + 1] = n;
      out[3 * d + 2] = q;
    } else {
      out[5 * d] = s;
      out[5 * d + 1] = n;
      out[5 * d + 2] = s;
      out[5 * d + 3] = n;
      out[5 * d + 4] = q;
    }
  }
}
int launch_reduce_rows(cudaStream_t st, size_t ndp, size_t nrows, const int64_t* d_rows, int mode, int64_t* d_out) {
  if (!ndp) return 0;
  k_reduce_rows<<<(unsigned)ndp, 256, 0, st>>>(nrows, d_rows, mode, d_out);
  return 1;
}

// lib/encoding/frequency_count.go:18-36: per provider, one counter per value of [min, max]. One block per provider
// with a shared-memory histogram (32-bit counters: a provider has fewer than 2^31 rows); a value outside the
// range is an index panic in the Go loop and an error here.
#define DRX_FC_MAXBINS 12000
__global__ void __launch_bounds__(256)
k_freq_count(size_t nrows, const int64_t* __restrict__ rows, int64_t qmin, int nbins, int64_t* __restrict__ out, int* bad) {
  __shared__ unsigned int hist[DRX_FC_MAXBINS];
  const size_t dp = blockIdx.x;
  for (int b = threadIdx.x; b < nbins; b += blockDim.x) hist[b] = 0;
  __syncthreads();
  const int64_t* r = rows + dp * nrows;
  int nbad = 0;
  for (size_t i = threadIdx.x; i < nrows; i += blockDim.x) {
    const int64_t v = r[i] - qmin;
    if (v < 0 || v >= nbins) nbad++;
    else atomicAdd(&hist[(int)v], 1u);
  }
  if (nbad) atomicAdd(bad, nbad);
  __syncthreads();
  for (int b = threadIdx.x; b < nbins; b += blockDim.x) out[dp * (size_t)nbins + b] = (int64_t)hist[b];
}
int launch_freq_count(cudaStream_t st, size_t ndp, size_t nrows, const int64_t* d_rows, int64_t qmin, int nbins,
                      int64_t* d_out, int* d_bad) {
  if (!ndp) return 0;
  k_freq_count<<<(unsigned)ndp, 256, 0, st>>>(nrows, d_rows, qmin, nbins, d_out, d_bad);
  return 1;
}

// ------------------------------------------------------------------ logistic-regression pre-pass (a-7)
// lib/encoding/logistic_regression.go: StandardiseWith (:964-978), Augment (:1014-1023), ComputeAllApproxCoefficients
// (:367-403), AggregateApproxCoefficients (:433-455), Float64ToInt642DArrayWithPrecision (:1558-1574). The result
// depends on the order of the float64 operations. k_logreg_standardise does the divisions once, one thread per
// (record, feature); then every (provider, coefficient) pair is ONE thread that adds
// its provider's records in index order with explicitly rounded, never fused operations (__dmul_rn / __dadd_rn /
// __ddiv_rn). A block (one warp) serves one provider and 32 coefficients; all its threads stage tiles of standardised,
// augmented records in shared memory, then each walks the tile for its own coefficient: two float64 operations per
// record and thread, the addition dependent on the previous one (latency-bound; HBM only sees 8(d+1) bytes per
// record and block).
#define DRX_LR_TILE 128
#define DRX_LR_MAXD 31
#define DRX_LR_BLOCK 32
// StandardiseWith, one thread per (record, feature): the divisions are the expensive float64 operations of the
// pre-pass and every block of the accumulation kernel would otherwise repeat them
__global__ void __launch_bounds__(256)
k_logreg_standardise(size_t n, int d, const double* __restrict__ x, const double* __restrict__ means,
                     const double* __restrict__ sds, double* __restrict__ xs) {
  const size_t e = (size_t)blockIdx.x * blockDim.x + threadIdx.x;
  if (e >= n * (size_t)d) return;
  const int f = (int)(e % (size_t)d);
  xs[e] = __ddiv_rn(__dsub_rn(x[e], means[f]), sds[f]);
}
__global__ void __launch_bounds__(DRX_LR_BLOCK)
k_logreg_coeffs(const size_t* __restrict__ rec_off, int d, int k, const double* __restrict__ xs,
                const int64_t* __restrict__ y, double precision, int64_t* __restrict__ out) {
  __shared__ double X[DRX_LR_TILE * (DRX_LR_MAXD + 1)];
  __shared__ double ys[DRX_LR_TILE];
  const int dp = blockIdx.x, d1 = d + 1;
  const int ncoef = d1 + (k >= 2 ? d1 * d1 : 0);
  const int coef = blockIdx.y * DRX_LR_BLOCK + threadIdx.x;
  const size_t r0 = rec_off[dp], r1 = rec_off[dp + 1];
  // which product this thread accumulates: level 1 -> X[s] * (2y - 1); level 2 -> ypart * ((1.0 * X[i1]) * X[i2])
  const bool active = coef < ncoef, lvl2 = coef >= d1;
  const int i1 = lvl2 ? (coef - d1) / d1 : coef, i2 = lvl2 ? (coef - d1) % d1 : 0;
  double acc = 0.0;
  for (size_t base = r0; base < r1; base += DRX_LR_TILE) {
    const int cnt = (int)((r1 - base) < (size_t)DRX_LR_TILE ? (r1 - base) : (size_t)DRX_LR_TILE);
    __syncthreads();
    for (int e = threadIdx.x; e < cnt * d; e += DRX_LR_BLOCK) {  // coalesced copy of the tile, Augment's 1.0 in column 0
      const int r = e / d, f = e - r * d;
      X[r * d1 + 1 + f] = xs[base * (size_t)d + e];
    }
    for (int r = threadIdx.x; r < cnt; r += DRX_LR_BLOCK) {
      X[r * d1] = 1.0;
      ys[r] = __dsub_rn(__dmul_rn(2.0, (double)y[base + r]), 1.0);  // 2y - 1
    }
    __syncthreads();
    if (active) {
      // Exact shortcuts, each giving the same double as the reference's expression: 1.0 * X = X; for level 2
      // ypart = y - y*(-1)^2 - 1 = -1.0 whatever y is, and (-1.0) * p is p with its sign flipped.
      if (!lvl2) {
#pragma unroll 4
        for (int r = 0; r < cnt; r++) acc = __dadd_rn(acc, __dmul_rn(X[r * d1 + i1], ys[r]));
      } else {
#pragma unroll 4
        for (int r = 0; r < cnt; r++) acc = __dadd_rn(acc, -__dmul_rn(X[r * d1 + i1], X[r * d1 + i2]));
      }
    }
  }
  if (active) {
    // level j sits at offset j * (d+1)^j of the packed vector (logistic_regression.go:306-313)
    const int pos = lvl2 ? d1 + (coef - d1) : coef;
    out[(size_t)dp * ncoef + pos] = (int64_t)round(__dmul_rn(acc, precision));  // math.Round: halves away from zero
  }
}
int launch_logreg_coeffs(cudaStream_t st, size_t ndp, size_t nrec, const size_t* d_rec_off, int d, int k, const double* d_x,
                         const int64_t* d_y, const double* d_means, const double* d_sds, double precision, double* d_xs,
                         int64_t* d_out) {
  if (!ndp) return 0;
  const int d1 = d + 1, ncoef = d1 + (k >= 2 ? d1 * d1 : 0);
  int launches = 1;
  if (nrec) {
    k_logreg_standardise<<<grid_for(nrec * (size_t)d, 256), 256, 0, st>>>(nrec, d, d_x, d_means, d_sds, d_xs);
    launches++;
  }
  dim3 grid((unsigned)ndp, (unsigned)((ncoef + DRX_LR_BLOCK - 1) / DRX_LR_BLOCK));
  k_logreg_coeffs<<<grid, DRX_LR_BLOCK, 0, st>>>(d_rec_off, d, k, d_xs, d_y, precision, d_out);
  return launches;
}

// ------------------------------------------------------------------ discrete-log table (a-5) and decryption (a-4)
// Open addressing, linear probing. Key = Montgomery-form x coordinate of i*B (8 words); value word =
// (i << 1) | parity(y), 0xffffffff = empty. -i*B shares the x coordinate and has the other y parity.
#define DRX_DLOG_EMPTY 0xffffffffu
#define DRX_DLOG_CHUNK 16

__device__ __forceinline__ uint32_t dlog_hash(const Fp& x) {
  return (x.v[0] ^ (x.v[1] * 0x9e3779b1u) ^ (x.v[2] * 0x85ebca6bu)) & (kDlogCap - 1);
}

__global__ void __launch_bounds__(64)
k_dlog_build(const G1Aff* __restrict__ tabB, const G1Aff* __restrict__ baseB, Fp* __restrict__ keys,
             uint32_t* __restrict__ vals) {
  const int t = blockIdx.x * blockDim.x + threadIdx.x;
  const int64_t first = 1 + (int64_t)t * DRX_DLOG_CHUNK;
  if (first >= kMaxHomomorphicInt) return;
  const G1Aff B = *baseB;
  G1Jac pts[DRX_DLOG_CHUNK];
  Fp pre[DRX_DLOG_CHUNK];
  G1Jac cur = comb_mul(tabB, sc_from_u64((uint64_t)first), 3);
  int cnt = 0;
  for (; cnt < DRX_DLOG_CHUNK && first + cnt < kMaxHomomorphicInt; cnt++) {
    if (cnt) cur = jac_add_aff(cur, B);
    pts[cnt] = cur;
    pre[cnt] = cnt ? fp_mul(pre[cnt - 1], cur.z) : cur.z;
  }
  Fp inv = fp_inv(pre[cnt - 1]);
  for (int k = cnt - 1; k >= 0; k--) {
    const Fp zinv = k ? fp_mul(inv, pre[k - 1]) : inv;
    if (k) inv = fp_mul(inv, pts[k].z);
    const G1Aff a = jac_to_aff_with_zinv(pts[k], zinv);
    const uint32_t par = fp_from_mont(a.y).v[0] & 1u;
    const uint32_t val = ((uint32_t)(first + k) << 1) | par;
    uint32_t slot = dlog_hash(a.x);
    for (;;) {
      const uint32_t old = atomicCAS(&vals[slot], DRX_DLOG_EMPTY, val);
      if (old == DRX_DLOG_EMPTY) {
        keys[slot] = a.x;
        break;
      }
      slot = (slot + 1) & (kDlogCap - 1);
    }
  }
}

// returns the signed discrete log of M in (-100000, 100000), or INT64_MIN when it is not in the table
__device__ __forceinline__ int64_t dlog_lookup(const G1Aff& M, const Fp* __restrict__ keys,
                                               const uint32_t* __restrict__ vals) {
  if (aff_is_inf(M)) return 0;
  uint32_t slot = dlog_hash(M.x);
  for (;;) {
    const uint32_t val = vals[slot];
    if (val == DRX_DLOG_EMPTY) return INT64_MIN;
    if (fp_eq(keys[slot], M.x)) {
      const uint32_t par = fp_from_mont(M.y).v[0] & 1u;
      const int64_t i = (int64_t)(val >> 1);
      return par == (val & 1u) ? i : -i;
    }
    slot = (slot + 1) & (kDlogCap - 1);
  }
}

__global__ void __launch_bounds__(128)
k_decrypt(size_t n, const uint32_t* __restrict__ sk, const uint32_t* __restrict__ ct, const Fp* __restrict__ keys,
          const uint32_t* __restrict__ vals, int64_t* __restrict__ out, int* bad) {
  const size_t i = (size_t)blockIdx.x * blockDim.x + threadIdx.x;
  if (i >= n) return;
  const Scalar k = sc_from_wire(sk);
  G1Aff K, C;
  int nbad = !g1_from_wire(K, ct + 32 * i);
  nbad += !g1_from_wire(C, ct + 32 * i + 16);
  if (nbad) {
    atomicAdd(bad, nbad);
    out[i] = INT64_MIN;
    return;
  }
  const G1Jac S = window_mul_glv(K, k);
  const G1Jac M = jac_add_aff(jac_neg(S), C);
  out[i] = dlog_lookup(jac_to_aff(M), keys, vals);
}

int launch_dlog_build(cudaStream_t st, const void* d_tabB, const void* d_baseB_aff, void* d_keys, uint32_t* d_vals) {
  cudaMemsetAsync(d_vals, 0xff, (size_t)kDlogCap * sizeof(uint32_t), st);
  const size_t threads = (kMaxHomomorphicInt + DRX_DLOG_CHUNK - 1) / DRX_DLOG_CHUNK;
  k_dlog_build<<<grid_for(threads, 64), 64, 0, st>>>((const G1Aff*)d_tabB, (const G1Aff*)d_baseB_aff, (Fp*)d_keys,
                                                    d_vals);
  return 1;
}
int launch_decrypt(cudaStream_t st, size_t n, const uint8_t* d_sk, const uint8_t* d_ct, const void* d_keys,
                   const uint32_t* d_vals, int64_t* d_out, int* d_bad) {
  if (!n) return 0;
  k_decrypt<<<grid_for(n, 128), 128, 0, st>>>(n, (const uint32_t*)d_sk, (const uint32_t*)d_ct, (const Fp*)d_keys,
                                             d_vals, d_out, d_bad);
  return 1;
}

// ------------------------------------------------------------------ key switching
// One server's contribution to switching ciphertext i from the collective key to the target key Q
// (libunlynxkeyswitch.KeySwitchSequence, unlynx v1.4.0; call sites data/data.go:80, services/service.go:565-616):
// K' = v_i*B, C' = v_i*Q - x*rB_i with rB_i the K component of the ciphertext, x the server's secret and v_i the
// injected randomness. Two comb walks and one windowed variable-base product per thread, one shared inversion;
// -rB_i (needed by the proof) costs a field negation on the way out.
__global__ void __launch_bounds__(128)
k_keyswitch(size_t n, const uint32_t* __restrict__ sk, const uint32_t* __restrict__ rb, const uint32_t* __restrict__ v,
            const G1Aff* __restrict__ tabB, const G1Aff* __restrict__ tabQ, uint32_t* __restrict__ cv,
            uint32_t* __restrict__ rb_neg, int* bad) {
  const size_t i = (size_t)blockIdx.x * blockDim.x + threadIdx.x;
  if (i >= n) return;
  G1Aff R;
  if (!g1_from_wire(R, rb + 16 * i)) {  // rejected element: counted, and its output rows are zeroed (never stale)
    atomicAdd(bad, 1);
    for (int k = 0; k < 32; k++) cv[32 * i + k] = 0;
    if (rb_neg)
      for (int k = 0; k < 16; k++) rb_neg[16 * i + k] = 0;
    return;
  }
  const G1Aff Rn = aff_neg(R);
  if (rb_neg) g1_to_wire(rb_neg + 16 * i, Rn);
  const Scalar vi = sc_from_wire(v + 8 * i);
  const G1Jac K = comb_mul(tabB, vi);
  const G1Jac S = comb_mul(tabQ, vi);
  const G1Jac X = window_mul_glv(Rn, sc_from_wire(sk));
  store_pair_wire(cv + 32 * i, K, jac_add(X, S));
}
int launch_keyswitch(cudaStream_t st, size_t n, const uint8_t* d_sk, const uint8_t* d_rb, const uint8_t* d_v,
                     const void* d_tabB, const void* d_tabQ, uint8_t* d_cv, uint8_t* d_rb_neg, int* d_bad) {
  if (!n) return 0;
  k_keyswitch<<<grid_for(n, 128), 128, 0, st>>>(n, (const uint32_t*)d_sk, (const uint32_t*)d_rb, (const uint32_t*)d_v,
                                               (const G1Aff*)d_tabB, (const G1Aff*)d_tabQ, (uint32_t*)d_cv,
                                               (uint32_t*)d_rb_neg, d_bad);
  return 1;
}

// ------------------------------------------------------------------ obfuscation
// CipherText.MulCipherTextbyScalar per element (protocols/obfuscation_protocol.go:240-243; data/data.go:52-56):
// out_i = (s_i*K_i, s_i*C_i). Two windowed variable-base products per thread, one shared inversion.
__global__ void __launch_bounds__(128)
k_ct_scalar_mul(size_t n, const uint32_t* __restrict__ sc, const uint32_t* __restrict__ ct, uint32_t* __restrict__ out,
                int* bad) {
  const size_t i = (size_t)blockIdx.x * blockDim.x + threadIdx.x;
  if (i >= n) return;
  G1Aff K, C;
  int nbad = !g1_from_wire(K, ct + 32 * i);
  nbad += !g1_from_wire(C, ct + 32 * i + 16);
  if (nbad) {  // rejected element: counted, and its output row is zeroed (never stale)
    atomicAdd(bad, nbad);
    for (int k = 0; k < 32; k++) out[32 * i + k] = 0;
    return;
  }
  const Scalar k = sc_from_wire(sc + 8 * i);
  store_pair_wire(out + 32 * i, window_mul_glv(K, k), window_mul_glv(C, k));
}
int launch_ct_scalar_mul(cudaStream_t st, size_t n, const uint8_t* d_s, const uint8_t* d_ct, uint8_t* d_out, int* d_bad) {
  if (!n) return 0;
  k_ct_scalar_mul<<<grid_for(n, 128), 128, 0, st>>>(n, (const uint32_t*)d_s, (const uint32_t*)d_ct, (uint32_t*)d_out, d_bad);
  return 1;
}

// ------------------------------------------------------------------ shuffling
// libunlynxshuffle.ShuffleSequence (unlynx lib/shuffle; call sites data/data.go:68 and, through
// protocolsunlynx.ShufflingProtocol, services/service.go:618-660): line i of the output is line pi[i] of the input with
// every ciphertext re-randomised, CipherText.Rerandomize: K' = K + beta*g, C' = C + beta*h with g the base point and
// h the collective key. Two comb walks and two mixed additions per thread, one shared inversion.
__global__ void __launch_bounds__(128)
k_shuffle_rerandomize(size_t lines, size_t nq, const int32_t* __restrict__ pi, const uint32_t* __restrict__ in,
                      const uint32_t* __restrict__ beta, const G1Aff* __restrict__ tabB, const G1Aff* __restrict__ tabP,
                      uint32_t* __restrict__ out, int* bad) {
  const size_t t = (size_t)blockIdx.x * blockDim.x + threadIdx.x;
  if (t >= lines * nq) return;
  const size_t i = t / nq, j = t - i * nq;
  const size_t src = (size_t)pi[i] * nq + j;
  G1Aff K, C;
  int nbad = !g1_from_wire(K, in + 32 * src);
  nbad += !g1_from_wire(C, in + 32 * src + 16);
  if (nbad) {  // rejected element: counted, and its output row is zeroed (never stale)
    atomicAdd(bad, nbad);
    for (int k = 0; k < 32; k++) out[32 * t + k] = 0;
    return;
  }
  const Scalar b = sc_from_wire(beta + 8 * src);
  store_pair_wire(out + 32 * t, jac_add_aff(comb_mul(tabB, b), K), jac_add_aff(comb_mul(tabP, b), C));
}
int launch_shuffle_rerandomize(cudaStream_t st, size_t lines, size_t nq, const int32_t* d_pi, const uint8_t* d_in,
                               const uint8_t* d_beta, const void* d_tabB, const void* d_tabP, uint8_t* d_out, int* d_bad) {
  if (!lines || !nq) return 0;
  k_shuffle_rerandomize<<<grid_for(lines * nq, 128), 128, 0, st>>>(lines, nq, d_pi, (const uint32_t*)d_in,
                                                                  (const uint32_t*)d_beta, (const G1Aff*)d_tabB,
                                                                  (const G1Aff*)d_tabP, (uint32_t*)d_out, d_bad);
  return 1;
}

// ------------------------------------------------------------------ Schnorr signatures
// kyber sign/schnorr Verify over G1, as VerifyProofSignature calls it for every proof request
// (lib/proof/structs_proofs.go:498-505): sig = R (64 B) || s (32 B); h = SHA-512(R || P || msg) read big-endian and
// reduced mod n (the caller hashes: the message is the proof list it already holds in host memory); accept when
// s*B == R + h*P. A malformed R or P, or s >= n (mod.Int.UnmarshalBinary rejects it), fails the signature, it is
// not an error of the call.
__global__ void __launch_bounds__(128)
k_schnorr_verify(size_t n, const uint32_t* __restrict__ pub, const uint32_t* __restrict__ sig,
                 const uint32_t* __restrict__ digest, const G1Aff* __restrict__ tabB, uint8_t* __restrict__ ok) {
  const size_t i = (size_t)blockIdx.x * blockDim.x + threadIdx.x;
  if (i >= n) return;
  G1Aff P, R;
  Scalar s;
  limbs_from_be_words(s.v, sig + 24 * i + 16);
  if (!g1_from_wire(P, pub + 16 * i) || !g1_from_wire(R, sig + 24 * i) || mod_geq<ModN>(s.v)) {
    ok[i] = 0;
    return;
  }
  const Scalar h = sc_from_digest(digest + 16 * i);
  // s*B - h*P - R must be the identity
  G1Jac T = jac_add(comb_mul(tabB, s), jac_neg(window_mul_glv(P, h)));
  T = jac_add_aff(T, aff_neg(R));
  ok[i] = jac_is_inf(T) ? 1 : 0;
}
int launch_schnorr_verify(cudaStream_t st, size_t n, const uint8_t* d_pub, const uint8_t* d_sig, const uint8_t* d_digest,
                          const void* d_tabB, uint8_t* d_ok) {
  if (!n) return 0;
  k_schnorr_verify<<<grid_for(n, 128), 128, 0, st>>>(n, (const uint32_t*)d_pub, (const uint32_t*)d_sig,
                                                    (const uint32_t*)d_digest, (const G1Aff*)d_tabB, d_ok);
  return 1;
}

// ------------------------------------------------------------------ proof.Rep relations
// The group arithmetic of kyber's proof.Rep predicates as the obfuscation proof (lib/obfuscation/obfuscation_proof.go:
// 37-91: co1 = s*c1 and co2 = s*c2) and unlynx's key-switch proof (libunlynxkeyswitch, verified from
// lib/proof/structs_proofs.go:443-482: viB = vi*B, K = k*B, ks2 = vi*Q + k*rBNeg) use them. A relation is
// P = x_1*B_1 + ... + x_T*B_T. kyber (proof/rep.go: commit, respond, verify; dependency, v3.0.8) has the prover send
// V = v_1*B_1 + ... + v_T*B_T, draw the challenge c from the hash of every V of the proof, answer r_j = v_j - c*x_j,
// and the verifier accept when r_1*B_1 + ... + r_T*B_T + c*P == V. The transcript hash stays with the caller (kyber's
// hashProver over the suite's XOF); these kernels are the three group / scalar steps, one thread per relation.
__global__ void __launch_bounds__(128)
k_rep_commit(size_t n, int nt, const uint32_t* __restrict__ sc, const uint32_t* __restrict__ pts, uint32_t* __restrict__ out,
             int* bad) {
  const size_t i = (size_t)blockIdx.x * blockDim.x + threadIdx.x;
  if (i >= n) return;
  G1Jac acc = jac_inf<Fp>();
  for (int j = 0; j < nt; j++) {
    G1Aff B;
    if (!g1_from_wire(B, pts + 16 * (i * nt + j))) {  // rejected element: counted, and its output row is zeroed
      atomicAdd(bad, 1);
      for (int k = 0; k < 16; k++) out[16 * i + k] = 0;
      return;
    }
    acc = jac_add(acc, window_mul_glv(B, sc_from_wire(sc + 8 * (i * nt + j))));
  }
  g1_to_wire(out + 16 * i, jac_to_aff(acc));
}
// r_i = v_i - c_(i / per) * x_i: `per` secrets of one proof share its challenge
__global__ void __launch_bounds__(128)
k_rep_respond(size_t n, int per, const uint32_t* __restrict__ v, const uint32_t* __restrict__ c, const uint32_t* __restrict__ x,
              uint32_t* __restrict__ r) {
  const size_t i = (size_t)blockIdx.x * blockDim.x + threadIdx.x;
  if (i >= n) return;
  const Scalar ci = sc_from_wire(c + 8 * (i / per));
  sc_to_wire(r + 8 * i, sc_sub(sc_from_wire(v + 8 * i), sc_mul(ci, sc_from_wire(x + 8 * i))));
}
// A malformed point or a response >= n (mod.Int.UnmarshalBinary rejects it) fails the relation, not the call.
__global__ void __launch_bounds__(128)
k_rep_verify(size_t n, int nt, const uint32_t* __restrict__ resp, const uint32_t* __restrict__ chal,
             const uint32_t* __restrict__ bases, const uint32_t* __restrict__ P, const uint32_t* __restrict__ V,
             uint8_t* __restrict__ ok) {
  const size_t i = (size_t)blockIdx.x * blockDim.x + threadIdx.x;
  if (i >= n) return;
  G1Aff Pi, Vi;
  if (!g1_from_wire(Pi, P + 16 * i) || !g1_from_wire(Vi, V + 16 * i)) {
    ok[i] = 0;
    return;
  }
  G1Jac acc = window_mul_glv(Pi, sc_from_wire(chal + 8 * i));
  for (int j = 0; j < nt; j++) {
    G1Aff B;
    Scalar rj;
    limbs_from_be_words(rj.v, resp + 8 * (i * nt + j));
    if (!g1_from_wire(B, bases + 16 * (i * nt + j)) || mod_geq<ModN>(rj.v)) {
      ok[i] = 0;
      return;
    }
    acc = jac_add(acc, window_mul_glv(B, rj));
  }
  acc = jac_add_aff(acc, aff_neg(Vi));
  ok[i] = jac_is_inf(acc) ? 1 : 0;
}
int launch_rep_commit(cudaStream_t st, size_t n, int nt, const uint8_t* d_sc, const uint8_t* d_pts, uint8_t* d_out, int* d_bad) {
  if (!n) return 0;
  k_rep_commit<<<grid_for(n, 128), 128, 0, st>>>(n, nt, (const uint32_t*)d_sc, (const uint32_t*)d_pts, (uint32_t*)d_out, d_bad);
  return 1;
}
int launch_rep_respond(cudaStream_t st, size_t n, int per, const uint8_t* d_v, const uint8_t* d_c, const uint8_t* d_x,
                       uint8_t* d_r) {
  if (!n) return 0;
  k_rep_respond<<<grid_for(n, 128), 128, 0, st>>>(n, per, (const uint32_t*)d_v, (const uint32_t*)d_c, (const uint32_t*)d_x,
                                                 (uint32_t*)d_r);
  return 1;
}
int launch_rep_verify(cudaStream_t st, size_t n, int nt, const uint8_t* d_resp, const uint8_t* d_chal, const uint8_t* d_bases,
                      const uint8_t* d_P, const uint8_t* d_V, uint8_t* d_ok) {
  if (!n) return 0;
  k_rep_verify<<<grid_for(n, 128), 128, 0, st>>>(n, nt, (const uint32_t*)d_resp, (const uint32_t*)d_chal,
                                                (const uint32_t*)d_bases, (const uint32_t*)d_P, (const uint32_t*)d_V, d_ok);
  return 1;
}

// ------------------------------------------------------------------ diagnostics
__global__ void k_g1_mul(size_t n, const uint32_t* __restrict__ k, const uint32_t* __restrict__ pt,
                         uint32_t* __restrict__ out, int* bad) {
  const size_t i = (size_t)blockIdx.x * blockDim.x + threadIdx.x;
  if (i >= n) return;
  G1Aff p;
  if (!g1_from_wire(p, pt + 16 * i)) atomicAdd(bad, 1);
  const G1Jac r = window_mul_glv(p, sc_from_wire(k + 8 * i));
  g1_to_wire(out + 16 * i, jac_to_aff(r));
}
int launch_g1_mul(cudaStream_t st, size_t n, const uint8_t* d_k, const uint8_t* d_pt, uint8_t* d_out, int* d_bad) {
  if (!n) return 0;
  k_g1_mul<<<grid_for(n, 64), 64, 0, st>>>(n, (const uint32_t*)d_k, (const uint32_t*)d_pt, (uint32_t*)d_out, d_bad);
  return 1;
}

__global__ void k_mont_mul(size_t n, int which, const uint32_t* __restrict__ a, const uint32_t* __restrict__ b,
                           uint32_t* __restrict__ r) {
  const size_t i = (size_t)blockIdx.x * blockDim.x + threadIdx.x;
  if (i >= n) return;
  uint32_t x[8], y[8], z[8];
  for (int k = 0; k < 8; k++) {
    x[k] = a[8 * i + k];
    y[k] = b[8 * i + k];
  }
  if (which == 0) mont_mul<ModP>(z, x, y);
  else if (which == 1) mont_mul<ModN>(z, x, y);
  else if (which == 2) mod_add<ModP>(z, x, y);
  else if (which == 3) mod_sub<ModP>(z, x, y);
  else if (which == 4) mod_add<ModN>(z, x, y);
  else if (which == 5) mod_sub<ModN>(z, x, y);
  else if (which == 10) mont_mul2<ModP>(z, x, y, x, y);  // the dual-product pass on one pair: 2xy/R (x < p)
  else {  // which 6/7: halves of SHA3-512(x || y); 8/9: of SHA3-512((x || y) * 3)
    uint32_t msg[48], dig[16];
    for (int rep = 0; rep < 3; rep++)
      for (int k = 0; k < 8; k++) {
        msg[16 * rep + k] = x[k];
        msg[16 * rep + 8 + k] = y[k];
      }
    sha3_512_words(dig, msg, which >= 8 ? 48 : 16);
    for (int k = 0; k < 8; k++) z[k] = dig[((which & 1) ? 8 : 0) + k];
  }
  for (int k = 0; k < 8; k++) r[8 * i + k] = z[k];
}
int launch_mont_mul(cudaStream_t st, size_t n, int which, const uint32_t* d_a, const uint32_t* d_b, uint32_t* d_r) {
  if (!n) return 0;
  k_mont_mul<<<grid_for(n, 128), 128, 0, st>>>(n, which, d_a, d_b, d_r);
  return 1;
}

// Integer-pipe probes. kind 0: four independent 64-bit multiply-accumulate chains per thread with no memory
// traffic (16 IMAD.WIDE.U32.X per step); kind 1: dependent Montgomery products.
__global__ void __launch_bounds__(256) k_probe_imad(int iters, uint32_t* sink) {
  uint32_t a = threadIdx.x * 2654435761u + 1u, b = blockIdx.x * 40503u + 7u;
  uint32_t c0 = a, c1 = b, c2 = a ^ b, c3 = a + b, c4 = ~a, c5 = ~b, c6 = a * b, c7 = a - b;
  for (int it = 0; it < iters; it++) {
#pragma unroll
    for (int k = 0; k < 4; k++) {
      c0 = mad_lo_cc(a, b, c0);
      c1 = madc_hi_cc(a, b, c1);
      c2 = madc_lo_cc(b, c7, c2);
      c3 = madc_hi_cc(b, c7, c3);
      c4 = mad_lo_cc(a, c3, c4);
      c5 = madc_hi_cc(a, c3, c5);
      c6 = madc_lo_cc(b, c1, c6);
      c7 = madc_hi_cc(b, c1, c7);
    }
  }
  sink[blockIdx.x * blockDim.x + threadIdx.x] = c0 ^ c1 ^ c2 ^ c3 ^ c4 ^ c5 ^ c6 ^ c7;
}
__global__ void __launch_bounds__(256) k_probe_montmul(int iters, uint32_t* sink) {
  Fp x, y;
  for (int k = 0; k < 8; k++) {
    x.v[k] = (threadIdx.x + 1) * 2654435761u + k;
    y.v[k] = (blockIdx.x + 1) * 40503u + k * 77u;
  }
  x.v[7] &= 0x7fffffffu;
  y.v[7] &= 0x7fffffffu;
  for (int it = 0; it < iters; it++) {
    x = fp_mul(x, y);
    y = fp_mul(y, x);
  }
  uint32_t s = 0;
  for (int k = 0; k < 8; k++) s ^= x.v[k] ^ y.v[k];
  sink[blockIdx.x * blockDim.x + threadIdx.x] = s;
}
// three independent chains of dependent Montgomery products per thread (what an Fp2 product offers the scheduler)
__global__ void __launch_bounds__(256) k_probe_montmul3(int iters, uint32_t* sink) {
  Fp x[3], y;
  for (int k = 0; k < 8; k++) {
    for (int q = 0; q < 3; q++) x[q].v[k] = (threadIdx.x + 1 + q) * 2654435761u + k;
    y.v[k] = (blockIdx.x + 1) * 40503u + k * 77u;
  }
  for (int q = 0; q < 3; q++) x[q].v[7] &= 0x7fffffffu;
  y.v[7] &= 0x7fffffffu;
  for (int it = 0; it < iters; it++) {
#pragma unroll
    for (int q = 0; q < 3; q++) mont_mul<ModP>(x[q].v, x[q].v, y.v);
  }
  uint32_t s = 0;
  for (int k = 0; k < 8; k++) s ^= x[0].v[k] ^ x[1].v[k] ^ x[2].v[k];
  sink[blockIdx.x * blockDim.x + threadIdx.x] = s;
}
// dependent field inversions (kind 3) / dependent full point additions (kind 4): the latency of the building blocks
// on the critical path of the small-batch kernels and of the aggregation's second stage
__global__ void __launch_bounds__(256) k_probe_inv(int iters, uint32_t* sink) {
  Fp x;
  for (int k = 0; k < 8; k++) x.v[k] = (threadIdx.x + 1) * 2654435761u + k * 40503u + blockIdx.x;
  x.v[7] &= 0x7fffffffu;
  for (int it = 0; it < iters; it++) {
    x = fp_inv(x);
    x.v[0] ^= 1u;
  }
  uint32_t s = 0;
  for (int k = 0; k < 8; k++) s ^= x.v[k];
  sink[blockIdx.x * blockDim.x + threadIdx.x] = s;
}
__global__ void __launch_bounds__(256) k_probe_jac_add(int iters, uint32_t* sink) {
  G1Jac a, b;
  for (int k = 0; k < 8; k++) {
    a.x.v[k] = (threadIdx.x + 1) * 2654435761u + k; a.y.v[k] = a.x.v[k] * 3u + 1u; a.z.v[k] = a.x.v[k] * 5u + 2u;
    b.x.v[k] = (blockIdx.x + 1) * 40503u + k * 77u; b.y.v[k] = b.x.v[k] * 7u + 3u; b.z.v[k] = b.x.v[k] * 11u + 4u;
  }
  a.x.v[7] &= 0x7fffffffu; a.y.v[7] &= 0x7fffffffu; a.z.v[7] &= 0x7fffffffu;
  b.x.v[7] &= 0x7fffffffu; b.y.v[7] &= 0x7fffffffu; b.z.v[7] &= 0x7fffffffu;
  for (int it = 0; it < iters; it++) a = jac_add(a, b);   // the formulas run on any field elements; no curve needed
  uint32_t s = 0;
  for (int k = 0; k < 8; k++) s ^= a.x.v[k] ^ a.y.v[k] ^ a.z.v[k];
  sink[blockIdx.x * blockDim.x + threadIdx.x] = s;
}
// kind 5: eight independent chains of double-precision fused multiply-adds (round toward zero, the form the
// 52-bit-limb product technique uses: high part fma_rz(a, b, 2^104), low part fma_rz(a, b, 2^104 + 2^52 - high)) with
// no memory traffic; 16 DFMA per step. Measures what the FP64 pipe offers beside the IMAD.WIDE figure of kind 0.
__global__ void __launch_bounds__(256) k_probe_dfma(int iters, uint32_t* sink) {
  double a = 1.0 + threadIdx.x * 1e-9, b = 1.0 + blockIdx.x * 1e-9;
  double c0 = a, c1 = b, c2 = a + b, c3 = a - b, c4 = a * 3, c5 = b * 3, c6 = a * 5, c7 = b * 5;
  for (int it = 0; it < iters; it++) {
#pragma unroll
    for (int k = 0; k < 2; k++) {
      c0 = __fma_rz(a, b, c0); c1 = __fma_rz(a, c0, c1); c2 = __fma_rz(b, c1, c2); c3 = __fma_rz(a, c2, c3);
      c4 = __fma_rz(b, c3, c4); c5 = __fma_rz(a, c4, c5); c6 = __fma_rz(b, c5, c6); c7 = __fma_rz(a, c6, c7);
    }
  }
  const double s = c0 + c1 + c2 + c3 + c4 + c5 + c6 + c7;
  sink[blockIdx.x * blockDim.x + threadIdx.x] = (uint32_t)__double_as_longlong(s);
}
int launch_microbench(cudaStream_t st, int kind, int iters, int blocks, int threads, uint32_t* d_sink) {
  if (kind == 0) k_probe_imad<<<blocks, threads, 0, st>>>(iters, d_sink);
  else if (kind == 2) k_probe_montmul3<<<blocks, threads, 0, st>>>(iters, d_sink);
  else if (kind == 3) k_probe_inv<<<blocks, threads, 0, st>>>(iters, d_sink);
  else if (kind == 4) k_probe_jac_add<<<blocks, threads, 0, st>>>(iters, d_sink);
  else if (kind == 5) k_probe_dfma<<<blocks, threads, 0, st>>>(iters, d_sink);
  else k_probe_montmul<<<blocks, threads, 0, st>>>(iters, d_sink);
  return 1;
}

}  // namespace drx

namespace drx {
// wire form of the sum of n affine points (sum of the servers' public keys, lib/range/range_proof.go:363-366)
__global__ void k_g1_sum_wire(int n, const G1Aff* __restrict__ pts, uint32_t* __restrict__ out) {
  if (blockIdx.x || threadIdx.x) return;
  G1Jac acc = jac_inf<Fp>();
  for (int i = 0; i < n; i++) acc = jac_add_aff(acc, pts[i]);
  g1_to_wire(out, jac_to_aff(acc));
}
int launch_g1_sum_wire(cudaStream_t st, int n, const void* d_aff, uint8_t* d_out_wire) {
  k_g1_sum_wire<<<1, 32, 0, st>>>(n, (const G1Aff*)d_aff, (uint32_t*)d_out_wire);
  return 1;
}
}  // namespace drx
