// SHA3-512 (FIPS 202) for the Fiat-Shamir challenge c = H(B || C || sum Y) of lib/range/range_proof.go:349-374
// (golang.org/x/crypto/sha3 New512 in the reference). Messages on this path are exactly 192 bytes.
// Both functions are deliberately out of line with a rolled round loop: the fully unrolled, inlined form
// (constant message length propagated into 3 x 24 rounds) produced wrong digests on sm_100a with nvcc 12.9
// inside k_rp_create_hdr while the same source gave right ones in a smaller kernel.
#pragma once
#include <stdint.h>
#include "tower.cuh"

namespace drx {

DRX_HD uint64_t rotl64(uint64_t x, int s) { return s ? (x << s) | (x >> (64 - s)) : x; }

// -DDRX_KECCAK_INLINE=1 (experiment builds only) restores the inlined, fully unrolled form; see tools/keccak_repro.sh
#ifndef DRX_KECCAK_INLINE
#define DRX_KECCAK_INLINE 0
#endif
#if DRX_KECCAK_INLINE
#define DRX_KECCAK_ATTR DRX_HD
#else
#define DRX_KECCAK_ATTR DRX_NI
#endif
DRX_KECCAK_ATTR void keccak_f1600(uint64_t st[25]) {
  const uint64_t RC[24] = {
      0x0000000000000001ULL, 0x0000000000008082ULL, 0x800000000000808aULL, 0x8000000080008000ULL,
      0x000000000000808bULL, 0x0000000080000001ULL, 0x8000000080008081ULL, 0x8000000000008009ULL,
      0x000000000000008aULL, 0x0000000000000088ULL, 0x0000000080008009ULL, 0x000000008000000aULL,
      0x000000008000808bULL, 0x800000000000008bULL, 0x8000000000008089ULL, 0x8000000000008003ULL,
      0x8000000000008002ULL, 0x8000000000000080ULL, 0x000000000000800aULL, 0x800000008000000aULL,
      0x8000000080008081ULL, 0x8000000000008080ULL, 0x0000000080000001ULL, 0x8000000080008008ULL};
  const int ROT[25] = {0, 1, 62, 28, 27, 36, 44, 6, 55, 20, 3, 10, 43, 25, 39, 41, 45, 15, 21, 8, 18, 2, 61, 56, 14};
#if defined(__CUDA_ARCH__) && !DRX_KECCAK_INLINE
#pragma unroll 1
#elif defined(__CUDA_ARCH__)
#pragma unroll
#endif
  for (int round = 0; round < 24; round++) {
    uint64_t c[5], d[5], b[25];
    for (int x = 0; x < 5; x++) c[x] = st[x] ^ st[x + 5] ^ st[x + 10] ^ st[x + 15] ^ st[x + 20];
    for (int x = 0; x < 5; x++) d[x] = c[(x + 4) % 5] ^ rotl64(c[(x + 1) % 5], 1);
    for (int i = 0; i < 25; i++) st[i] ^= d[i % 5];
    for (int x = 0; x < 5; x++)
      for (int y = 0; y < 5; y++) b[y + 5 * ((2 * x + 3 * y) % 5)] = rotl64(st[x + 5 * y], ROT[x + 5 * y]);
    for (int y = 0; y < 5; y++)
      for (int x = 0; x < 5; x++) st[x + 5 * y] = b[x + 5 * y] ^ (~b[(x + 1) % 5 + 5 * y] & b[(x + 2) % 5 + 5 * y]);
    st[0] ^= RC[round];
  }
}

// msg: nwords 32-bit words in memory (little-endian byte order within each word, as loaded from a byte
// buffer); nwords*4 bytes total, any length. digest: 16 words = 64 bytes in memory order.
DRX_KECCAK_ATTR void sha3_512_words(uint32_t digest[16], const uint32_t* msg, int nwords) {
  uint64_t st[25];
  for (int i = 0; i < 25; i++) st[i] = 0;
  const int rate_words = 18;  // 72-byte rate
  int pos = 0;                // word position inside the current block
  for (int i = 0; i < nwords; i++) {
    st[pos >> 1] ^= (uint64_t)msg[i] << ((pos & 1) * 32);
    if (++pos == rate_words) {
      keccak_f1600(st);
      pos = 0;
    }
  }
  st[pos >> 1] ^= (uint64_t)0x06 << ((pos & 1) * 32);
  st[8] ^= 0x8000000000000000ULL;
  keccak_f1600(st);
  for (int i = 0; i < 8; i++) {
    digest[2 * i] = (uint32_t)st[i];
    digest[2 * i + 1] = (uint32_t)(st[i] >> 32);
  }
}

}  // namespace drx
