// drynx_host.hpp -- C++ host mirror of the reference's call surface over the C ABI (include/drynx_gpu.h).
//
// The reference is Go; Go is not installed here, so the host side above the ABI is C++ (header-only). Names,
// argument meaning and results follow the Go functions so that host/test_host.cpp reads like the reference's own
// tests:
//   libunlynx        (unlynx v1.4.0 lib/crypto.go)  CipherText, CipherVector, EncryptInt(GetR), EncryptIntVectorGetRs,
//                                                   CipherVector Add, DecryptInt(WithNeg), DecryptIntVectorWithNeg,
//                                                   DecryptCheckZero, CreateDecryptionTable
//   libdrynxrange    (lib/range/range_proof.go)     ToBase :584, InitRangeProofSignature :270, CreateProof :61,
//                                                   CreatePredicateRangeProofListForAllServers :305,
//                                                   RangeProofListVerification :484
//   libdrynxencoding (lib/encoding/*.go)            Encode*/Decode* for sum, mean, variance, frequency count,
//                                                   linear regression (dims) and logistic regression
// No arithmetic on group elements happens here: every such operation is one batched call into libdrynx_gpu.so.
// The randomness a Go node would draw from crypto/rand comes from a caller-supplied RandomStream.
#pragma once
#include <algorithm>
#include <cmath>
#include <cstdint>
#include <cstring>
#include <functional>
#include <map>
#include <stdexcept>
#include <string>
#include <utility>
#include <vector>

#include "../include/drynx_gpu.h"

namespace drynx {

using Bytes = std::vector<uint8_t>;
struct Scalar {
  uint8_t b[32];
};
struct Point {
  uint8_t b[64];
};
// Fills 32 bytes with a uniformly random scalar < n, big-endian (kyber Scalar.Pick).
using RandomStream = std::function<void(uint8_t out[32])>;

struct Error : std::runtime_error {
  int code;
  Error(int c, const std::string& m) : std::runtime_error(m), code(c) {}
};
inline void check(int rc) {
  if (rc != 0) throw Error(rc, drx_last_error());
}

// One context per process, like the package-level suite of lib/suite.go.
class Gpu {
 public:
  static Gpu& get(int device = 0) {
    static Gpu g(device);
    return g;
  }
  drx_ctx* ctx() const { return c_; }
  RandomStream random;  // set by the application; used wherever the reference calls RandomStream()

 private:
  explicit Gpu(int device) { check(drx_init(&c_, device, nullptr, nullptr)); }
  ~Gpu() { drx_free(c_); }
  drx_ctx* c_ = nullptr;
};
inline Scalar pick() {
  Scalar s;
  if (!Gpu::get().random) throw Error(-1, "drynx::Gpu::random is not set");
  Gpu::get().random(s.b);
  return s;
}
inline Scalar scalar_from_u64(uint64_t v) {
  Scalar s{};
  for (int i = 0; i < 8; i++) s.b[31 - i] = (uint8_t)(v >> (8 * i));
  return s;
}
inline Point base_point() {  // SuiTe.Point().Base() = (1, -2)
  static const uint8_t y[32] = {0x8f, 0xb5, 0x01, 0xe3, 0x4a, 0xa3, 0x87, 0xf9, 0xaa, 0x6f, 0xec, 0xb8, 0x61, 0x84, 0xdc, 0x21,
                                0xee, 0x5b, 0x88, 0xd1, 0x20, 0xb5, 0xb5, 0x9e, 0x18, 0x5c, 0xac, 0x6c, 0x5e, 0x08, 0x96, 0x65};
  Point p{};
  p.b[31] = 1;
  memcpy(p.b + 32, y, 32);
  return p;
}
inline Point point_mul(const Scalar& k, const Point& p) {
  Point r;
  check(drx_g1_mul(Gpu::get().ctx(), 1, k.b, p.b, r.b));
  return r;
}
struct KeyPair {  // key.NewKeyPair(libunlynx.SuiTe)
  Scalar Private;
  Point Public;
};
inline KeyPair NewKeyPair() {
  KeyPair k;
  k.Private = pick();
  k.Public = point_mul(k.Private, base_point());
  return k;
}

}  // namespace drynx

// =====================================================================================================================
namespace libunlynx {
using drynx::Gpu;
using drynx::Point;
using drynx::Scalar;

constexpr int64_t MaxHomomorphicInt = 100000;

struct CipherText {
  uint8_t b[128];  // K || C, the ToBytes form
};
using CipherVector = std::vector<CipherText>;

// The collective key the following calls encrypt under (the `pubkey` argument of the Go functions; the device
// keeps one fixed-base table, so switching keys costs a table build).
inline void use_pubkey(const Point& pub) {
  static Point cur{};
  static bool have = false;
  if (!have || memcmp(cur.b, pub.b, 64)) {
    drynx::check(drx_set_pubkey(Gpu::get().ctx(), pub.b));
    cur = pub;
    have = true;
  }
}

// EncryptIntVectorGetRs: one scalar per value drawn from the random stream, one batched encryption.
inline std::pair<CipherVector, std::vector<Scalar>> EncryptIntVectorGetRs(const Point& pubkey, const std::vector<int64_t>& ints) {
  use_pubkey(pubkey);
  std::vector<Scalar> rs(ints.size());
  for (auto& r : rs) r = drynx::pick();
  CipherVector cv(ints.size());
  if (!ints.empty())
    drynx::check(drx_encrypt(Gpu::get().ctx(), ints.size(), ints.data(), rs[0].b, cv[0].b));
  return {cv, rs};
}
inline CipherVector EncryptIntVector(const Point& pubkey, const std::vector<int64_t>& ints) {
  return EncryptIntVectorGetRs(pubkey, ints).first;
}
inline std::pair<CipherText, Scalar> EncryptIntGetR(const Point& pubkey, int64_t v) {
  auto r = EncryptIntVectorGetRs(pubkey, {v});
  return {r.first[0], r.second[0]};
}
inline CipherText EncryptInt(const Point& pubkey, int64_t v) { return EncryptIntGetR(pubkey, v).first; }

// EncryptScalar per element (lib/encoding/OR_AND.go:52,105): the message is a scalar, one batched call
inline CipherVector EncryptScalarVector(const Point& pubkey, const std::vector<Scalar>& ms) {
  use_pubkey(pubkey);
  std::vector<Scalar> rs(ms.size());
  for (auto& r : rs) r = drynx::pick();
  CipherVector cv(ms.size());
  if (!ms.empty()) drynx::check(drx_encrypt_scalar(Gpu::get().ctx(), ms.size(), ms[0].b, rs[0].b, cv[0].b));
  return cv;
}
inline CipherText EncryptScalar(const Point& pubkey, const Scalar& m) { return EncryptScalarVector(pubkey, {m})[0]; }

// CipherText.MulCipherTextbyScalar per element (protocols/obfuscation_protocol.go:240-243): (s_i*K_i, s_i*C_i)
inline CipherVector MulCipherVectorByScalars(const CipherVector& cv, const std::vector<Scalar>& ss) {
  if (cv.size() != ss.size()) throw drynx::Error(-1, "MulCipherTextbyScalar: length mismatch");
  CipherVector out(cv.size());
  if (!cv.empty()) drynx::check(drx_ct_scalar_mul(Gpu::get().ctx(), cv.size(), ss[0].b, cv[0].b, out[0].b));
  return out;
}
inline CipherText MulCipherTextbyScalar(const CipherText& c, const Scalar& s) { return MulCipherVectorByScalars({c}, {s})[0]; }

// CipherVector.Add over any number of vectors (the root node's fold, protocols/data_collection_protocol.go:145-166).
inline CipherVector AddAll(const std::vector<CipherVector>& vs) {
  if (vs.empty()) return {};
  const size_t v = vs[0].size();
  std::vector<const uint8_t*> ptrs;
  for (auto& x : vs) {
    if (x.size() != v) throw drynx::Error(-1, "CipherVector.Add: length mismatch");
    ptrs.push_back(v ? x[0].b : nullptr);
  }
  CipherVector out(v);
  if (v) drynx::check(drx_aggregate(Gpu::get().ctx(), vs.size(), v, ptrs.data(), out[0].b));
  return out;
}
inline CipherVector Add(const CipherVector& a, const CipherVector& b) { return AddAll({a, b}); }

inline std::vector<int64_t> DecryptIntVectorImpl(const Scalar& sk, const CipherVector& cv, bool neg) {
  std::vector<int64_t> out(cv.size());
  if (!cv.empty()) drynx::check(drx_decrypt(Gpu::get().ctx(), cv.size(), sk.b, cv[0].b, neg ? 1 : 0, out.data(), nullptr));
  return out;
}
inline std::vector<int64_t> DecryptIntVector(const Scalar& sk, const CipherVector& cv) { return DecryptIntVectorImpl(sk, cv, false); }
inline std::vector<int64_t> DecryptIntVectorWithNeg(const Scalar& sk, const CipherVector& cv) { return DecryptIntVectorImpl(sk, cv, true); }
inline int64_t DecryptInt(const Scalar& sk, const CipherText& c) { return DecryptIntVector(sk, {c})[0]; }
inline int64_t DecryptIntWithNeg(const Scalar& sk, const CipherText& c) { return DecryptIntVectorWithNeg(sk, {c})[0]; }
inline std::vector<int64_t> DecryptCheckZeroVector(const Scalar& sk, const CipherVector& cv) {
  std::vector<int64_t> out(cv.size());
  if (!cv.empty()) drynx::check(drx_decrypt_check_zero(Gpu::get().ctx(), cv.size(), sk.b, cv[0].b, out.data()));
  return out;
}
inline int64_t DecryptCheckZero(const Scalar& sk, const CipherText& c) {
  int64_t out;
  drynx::check(drx_decrypt_check_zero(Gpu::get().ctx(), 1, sk.b, c.b, &out));
  return out;
}
// CreateDecryptionTable(limit, pub, priv): the reference encrypts and decrypts `limit`; here the table is built
// directly, the keys are accepted for signature compatibility.
inline void CreateDecryptionTable(int64_t limit, const Point&, const Scalar&) {
  drynx::check(drx_dlog_table(Gpu::get().ctx(), limit));
}

}  // namespace libunlynx

// =====================================================================================================================
// The verifying nodes' remaining per-list work that is pure group arithmetic: the Schnorr signature
// on every proof request, and the aggregation proof, which is a recomputation.
namespace drynxproof {
using drynx::Gpu;
using drynx::Point;

// SHA-512 (FIPS 180-4) for the Schnorr challenge; the message is the serialised proof list, already in host memory.
inline void sha512(const uint8_t* msg0, size_t len0, const uint8_t* msg1, size_t len1, const uint8_t* msg2, size_t len2,
                   uint8_t out[64]) {
  static const uint64_t K[80] = {
      0x428a2f98d728ae22ull, 0x7137449123ef65cdull, 0xb5c0fbcfec4d3b2full, 0xe9b5dba58189dbbcull, 0x3956c25bf348b538ull,
      0x59f111f1b605d019ull, 0x923f82a4af194f9bull, 0xab1c5ed5da6d8118ull, 0xd807aa98a3030242ull, 0x12835b0145706fbeull,
      0x243185be4ee4b28cull, 0x550c7dc3d5ffb4e2ull, 0x72be5d74f27b896full, 0x80deb1fe3b1696b1ull, 0x9bdc06a725c71235ull,
      0xc19bf174cf692694ull, 0xe49b69c19ef14ad2ull, 0xefbe4786384f25e3ull, 0x0fc19dc68b8cd5b5ull, 0x240ca1cc77ac9c65ull,
      0x2de92c6f592b0275ull, 0x4a7484aa6ea6e483ull, 0x5cb0a9dcbd41fbd4ull, 0x76f988da831153b5ull, 0x983e5152ee66dfabull,
      0xa831c66d2db43210ull, 0xb00327c898fb213full, 0xbf597fc7beef0ee4ull, 0xc6e00bf33da88fc2ull, 0xd5a79147930aa725ull,
      0x06ca6351e003826full, 0x142929670a0e6e70ull, 0x27b70a8546d22ffcull, 0x2e1b21385c26c926ull, 0x4d2c6dfc5ac42aedull,
      0x53380d139d95b3dfull, 0x650a73548baf63deull, 0x766a0abb3c77b2a8ull, 0x81c2c92e47edaee6ull, 0x92722c851482353bull,
      0xa2bfe8a14cf10364ull, 0xa81a664bbc423001ull, 0xc24b8b70d0f89791ull, 0xc76c51a30654be30ull, 0xd192e819d6ef5218ull,
      0xd69906245565a910ull, 0xf40e35855771202aull, 0x106aa07032bbd1b8ull, 0x19a4c116b8d2d0c8ull, 0x1e376c085141ab53ull,
      0x2748774cdf8eeb99ull, 0x34b0bcb5e19b48a8ull, 0x391c0cb3c5c95a63ull, 0x4ed8aa4ae3418acbull, 0x5b9cca4f7763e373ull,
      0x682e6ff3d6b2b8a3ull, 0x748f82ee5defb2fcull, 0x78a5636f43172f60ull, 0x84c87814a1f0ab72ull, 0x8cc702081a6439ecull,
      0x90befffa23631e28ull, 0xa4506cebde82bde9ull, 0xbef9a3f7b2c67915ull, 0xc67178f2e372532bull, 0xca273eceea26619cull,
      0xd186b8c721c0c207ull, 0xeada7dd6cde0eb1eull, 0xf57d4f7fee6ed178ull, 0x06f067aa72176fbaull, 0x0a637dc5a2c898a6ull,
      0x113f9804bef90daeull, 0x1b710b35131c471bull, 0x28db77f523047d84ull, 0x32caab7b40c72493ull, 0x3c9ebe0a15c9bebcull,
      0x431d67c49c100d4cull, 0x4cc5d4becb3e42b6ull, 0x597f299cfc657e2aull, 0x5fcb6fab3ad6faecull, 0x6c44198c4a475817ull};
  uint64_t h[8] = {0x6a09e667f3bcc908ull, 0xbb67ae8584caa73bull, 0x3c6ef372fe94f82bull, 0xa54ff53a5f1d36f1ull,
                   0x510e527fade682d1ull, 0x9b05688c2b3e6c1full, 0x1f83d9abfb41bd6bull, 0x5be0cd19137e2179ull};
  auto rotr = [](uint64_t x, int n) { return (x >> n) | (x << (64 - n)); };
  auto block = [&](const uint8_t* b) {
    uint64_t w[80];
    for (int i = 0; i < 16; i++) {
      w[i] = 0;
      for (int k = 0; k < 8; k++) w[i] = (w[i] << 8) | b[8 * i + k];
    }
    for (int i = 16; i < 80; i++) {
      const uint64_t s0 = rotr(w[i - 15], 1) ^ rotr(w[i - 15], 8) ^ (w[i - 15] >> 7);
      const uint64_t s1 = rotr(w[i - 2], 19) ^ rotr(w[i - 2], 61) ^ (w[i - 2] >> 6);
      w[i] = w[i - 16] + s0 + w[i - 7] + s1;
    }
    uint64_t a = h[0], bb = h[1], c = h[2], d = h[3], e = h[4], f = h[5], g = h[6], hh = h[7];
    for (int i = 0; i < 80; i++) {
      const uint64_t S1 = rotr(e, 14) ^ rotr(e, 18) ^ rotr(e, 41), ch = (e & f) ^ (~e & g);
      const uint64_t t1 = hh + S1 + ch + K[i] + w[i];
      const uint64_t S0 = rotr(a, 28) ^ rotr(a, 34) ^ rotr(a, 39), mj = (a & bb) ^ (a & c) ^ (bb & c);
      hh = g; g = f; f = e; e = d + t1; d = c; c = bb; bb = a; a = t1 + S0 + mj;
    }
    h[0] += a; h[1] += bb; h[2] += c; h[3] += d; h[4] += e; h[5] += f; h[6] += g; h[7] += hh;
  };
  uint8_t buf[128];
  size_t fill = 0, total = 0;
  auto feed = [&](const uint8_t* p, size_t n) {
    total += n;
    while (n) {
      const size_t take = std::min(n, (size_t)128 - fill);
      memcpy(buf + fill, p, take);
      fill += take; p += take; n -= take;
      if (fill == 128) { block(buf); fill = 0; }
    }
  };
  feed(msg0, len0);
  feed(msg1, len1);
  feed(msg2, len2);
  const uint64_t bits = (uint64_t)total * 8;
  uint8_t pad[256] = {0x80};
  const size_t padlen = (fill < 112 ? 112 : 240) - fill;
  uint8_t lenb[16] = {0};
  for (int k = 0; k < 8; k++) lenb[15 - k] = (uint8_t)(bits >> (8 * k));
  feed(pad, padlen);
  feed(lenb, 16);
  for (int i = 0; i < 8; i++)
    for (int k = 0; k < 8; k++) out[8 * i + k] = (uint8_t)(h[i] >> (56 - 8 * k));
}

struct SignedData {  // one proof request's (sender key, serialised proofs, R || s)
  Point pubKey;
  const uint8_t* data;
  size_t len;
  uint8_t signature[96];
};
// VerifyProofSignature (lib/proof/structs_proofs.go:498-505) for a batch of proof requests: host SHA-512 per request,
// one device call for the group equations. true = signature correct.
inline std::vector<bool> VerifyProofSignatures(const std::vector<SignedData>& reqs) {
  const size_t n = reqs.size();
  std::vector<uint8_t> pub(64 * n), sig(96 * n), dig(64 * n), ok(n);
  for (size_t i = 0; i < n; i++) {
    memcpy(&pub[64 * i], reqs[i].pubKey.b, 64);
    memcpy(&sig[96 * i], reqs[i].signature, 96);
    sha512(reqs[i].signature, 64, reqs[i].pubKey.b, 64, reqs[i].data, reqs[i].len, &dig[64 * i]);
  }
  if (n) drynx::check(drx_schnorr_verify(Gpu::get().ctx(), n, pub.data(), sig.data(), dig.data(), ok.data()));
  std::vector<bool> out(n);
  for (size_t i = 0; i < n; i++) out[i] = ok[i] != 0;
  return out;
}
inline bool VerifyProofSignature(const Point& pubKey, const uint8_t* data, size_t len, const uint8_t signature[96]) {
  SignedData r;
  r.pubKey = pubKey;
  r.data = data;
  r.len = len;
  memcpy(r.signature, signature, 96);
  return VerifyProofSignatures({r})[0];
}

// kyber proof.HashProve / proof.HashVerify (kyber v3.0.8 proof/hash.go) draw the challenge of a Sigma protocol from the
// suite's XOF (blake2xb) seeded with the protocol name and reseeded with every commitment Put so far. That XOF is kyber's
// own construction and is not restated here: on the Go side the cgo shim keeps using kyber for this one step
// (go/gpubackend/proofs.go), and the device entry points take the challenge as an input. For a C++ caller the default
// below is a stand-in with the same shape (a hash of name || commitments): SHA-512, first 32 bytes, top bit cleared
// (< 2^255 < n). Replace it to interoperate with another transcript.
using ChallengeFn = std::function<drynx::Scalar(const std::string& protocol, const uint8_t* commitments, size_t len)>;
inline drynx::Scalar default_challenge(const std::string& protocol, const uint8_t* commitments, size_t len) {
  uint8_t d[64];
  sha512((const uint8_t*)protocol.data(), protocol.size(), commitments, len, nullptr, 0, d);
  drynx::Scalar c;
  memcpy(c.b, d, 32);
  c.b[0] &= 0x7f;
  return c;
}
inline ChallengeFn& challenge_fn() {
  static ChallengeFn f = default_challenge;
  return f;
}

}  // namespace drynxproof

namespace libunlynxaggr {
// PublishedAggregationProof (unlynx lib/aggregation, used at data/data.go:41-47 and services/service.go:533-541):
// the data and the claimed sum; verification recomputes Data.Acum() and compares.
struct PublishedAggregationProof {
  libunlynx::CipherVector Data;
  libunlynx::CipherText AggregationResult;
};
struct PublishedAggregationListProof {
  std::vector<PublishedAggregationProof> List;
};
inline PublishedAggregationListProof AggregationListProofCreation(const std::vector<libunlynx::CipherVector>& data,
                                                                  const libunlynx::CipherVector& results) {
  PublishedAggregationListProof l;
  for (size_t i = 0; i < data.size(); i++) l.List.push_back({data[i], results[i]});
  return l;
}
// AggregationListProofVerification(palp, percent): the first ceil(percent * len) proofs, all sums in one device call
// when the vectors have one length (the usual case), else one call per proof.
inline bool AggregationListProofVerification(const PublishedAggregationListProof& l, double percent) {
  const size_t nb = (size_t)std::ceil(percent * (double)l.List.size());
  bool ok = true;
  for (size_t i = 0; i < nb && i < l.List.size(); i++) {
    const auto& p = l.List[i];
    std::vector<libunlynx::CipherVector> cols;
    for (auto& c : p.Data) cols.push_back({c});
    libunlynx::CipherText sum;
    memset(sum.b, 0, 128);
    if (!cols.empty()) sum = libunlynx::AddAll(cols)[0];
    ok = ok && !memcmp(sum.b, p.AggregationResult.b, 128);
  }
  return ok;
}
}  // namespace libunlynxaggr

// =====================================================================================================================
namespace libunlynxshuffle {
using drynx::Gpu;
using drynx::Point;
using drynx::Scalar;

struct Shuffled {  // the three return values of ShuffleSequence
  std::vector<libunlynx::CipherVector> outputList;
  std::vector<int> pi;
  std::vector<std::vector<Scalar>> beta;
};
// ShuffleSequence(inputList, g = B, h, nil) (unlynx lib/shuffle; data/data.go:68, services/service.go:618-660): a fresh
// blinding factor per ciphertext from the random stream, a random permutation of the lines (Fisher-Yates over the
// same stream), one device call for the re-randomisation. `h` is the collective key.
inline Shuffled ShuffleSequence(const std::vector<libunlynx::CipherVector>& inputList, const Point& h) {
  libunlynx::use_pubkey(h);
  Shuffled s;
  const size_t k = inputList.size(), nq = k ? inputList[0].size() : 0;
  s.beta.assign(k, std::vector<Scalar>(nq));
  std::vector<uint8_t> in(k * nq * 128), beta(k * nq * 32), out(k * nq * 128);
  for (size_t i = 0; i < k; i++) {
    if (inputList[i].size() != nq) throw drynx::Error(-1, "ShuffleSequence: lines of different lengths");
    for (size_t j = 0; j < nq; j++) {
      s.beta[i][j] = drynx::pick();
      memcpy(&in[(i * nq + j) * 128], inputList[i][j].b, 128);
      memcpy(&beta[(i * nq + j) * 32], s.beta[i][j].b, 32);
    }
  }
  s.pi.resize(k);
  for (size_t i = 0; i < k; i++) s.pi[i] = (int)i;
  for (size_t i = k; i > 1; i--) {
    const Scalar r = drynx::pick();
    uint64_t v = 0;
    for (int b = 24; b < 32; b++) v = (v << 8) | r.b[b];
    std::swap(s.pi[i - 1], s.pi[v % i]);
  }
  std::vector<int32_t> pi32(s.pi.begin(), s.pi.end());
  if (k && nq)
    drynx::check(drx_shuffle_rerandomize(Gpu::get().ctx(), k, nq, pi32.data(), in.data(), beta.data(), out.data()));
  s.outputList.assign(k, libunlynx::CipherVector(nq));
  for (size_t i = 0; i < k; i++)
    for (size_t j = 0; j < nq; j++) memcpy(s.outputList[i][j].b, &out[(i * nq + j) * 128], 128);
  return s;
}
}  // namespace libunlynxshuffle

// =====================================================================================================================
namespace libunlynxkeyswitch {
using drynx::Gpu;
using drynx::Point;
using drynx::Scalar;

struct Sequence {  // the four return values of KeySwitchSequence
  libunlynx::CipherVector cv;   // (vi*B, ks2_i)
  std::vector<Point> ks2s;      // vi*targetPubKey - secretKey*rB_i
  std::vector<Point> rBNegs;    // -rB_i
  std::vector<Scalar> vis;
};
// KeySwitchSequence(targetPubKey, rBs, secretKey) (unlynx lib/key_switch/key_switch.go; data/data.go:80,
// services/service.go:565-616): vi drawn from the random stream per element, one batched device call.
inline Sequence KeySwitchSequence(const Point& targetPubKey, const std::vector<Point>& rBs, const Scalar& secretKey) {
  static Point cur{};
  static bool have = false;
  if (!have || memcmp(cur.b, targetPubKey.b, 64)) {
    drynx::check(drx_set_target_pubkey(Gpu::get().ctx(), targetPubKey.b));
    cur = targetPubKey;
    have = true;
  }
  Sequence s;
  const size_t n = rBs.size();
  s.vis.resize(n);
  for (auto& v : s.vis) v = drynx::pick();
  s.cv.resize(n);
  s.ks2s.resize(n);
  s.rBNegs.resize(n);
  if (n) drynx::check(drx_keyswitch(Gpu::get().ctx(), n, secretKey.b, rBs[0].b, s.vis[0].b, s.cv[0].b, s.rBNegs[0].b));
  for (size_t i = 0; i < n; i++) memcpy(s.ks2s[i].b, s.cv[i].b + 64, 64);
  return s;
}
// What the root of protocolsunlynx.KeySwitchingProtocol ends with: (0, C_i) plus every server's contribution.
inline libunlynx::CipherVector Combine(const libunlynx::CipherVector& original, const std::vector<libunlynx::CipherVector>& contributions) {
  libunlynx::CipherVector start = original;
  for (auto& c : start) memset(c.b, 0, 64);
  std::vector<libunlynx::CipherVector> all{start};
  all.insert(all.end(), contributions.begin(), contributions.end());
  return libunlynx::AddAll(all);
}


// PublishedKSProof / PublishedKSListProof (unlynx lib/key_switch proofs; verified by the VNs at
// lib/proof/structs_proofs.go:469-482): the predicate is And(Rep(viB = vi*B), Rep(K = k*B), Rep(ks2 = vi*Q + k*rBNeg)),
// the proof bytes in kyber's order V1 || V2 || V3 || r_vi || r_k (256 bytes).
struct PublishedKSProof {
  drynx::Bytes Proof;
  Point K, ViB, Ks2, RbNeg, Q;
};
struct PublishedKSListProof {
  std::vector<PublishedKSProof> List;
};
// KeySwitchListProofCreation(pubKey, targetPubKey, secretKey, ks2s, rBNegs, vis) (services/service.go:581,
// data/data.go:81): four batched device calls for the whole list (commitments with one and two terms, viB, responses).
inline PublishedKSListProof KeySwitchListProofCreation(const Point& pubKey, const Point& targetPubKey, const Scalar& secretKey,
                                                       const std::vector<Point>& ks2s, const std::vector<Point>& rBNegs,
                                                       const std::vector<Scalar>& vis) {
  const size_t n = vis.size();
  PublishedKSListProof l;
  l.List.resize(n);
  if (!n) return l;
  const Point B = drynx::base_point();
  std::vector<uint8_t> x(64 * n), blind(64 * n), bb(128 * n), qr(128 * n), v12(128 * n), v3(64 * n), vib(128 * n), r(64 * n),
      chal(32 * n);
  for (size_t i = 0; i < n; i++) {
    memcpy(&x[64 * i], vis[i].b, 32);
    memcpy(&x[64 * i + 32], secretKey.b, 32);
    const Scalar a = drynx::pick(), b = drynx::pick();
    memcpy(&blind[64 * i], a.b, 32);
    memcpy(&blind[64 * i + 32], b.b, 32);
    memcpy(&bb[128 * i], B.b, 64);
    memcpy(&bb[128 * i + 64], B.b, 64);
    memcpy(&qr[128 * i], targetPubKey.b, 64);
    memcpy(&qr[128 * i + 64], rBNegs[i].b, 64);
  }
  drx_ctx* g = Gpu::get().ctx();
  drynx::check(drx_rep_commit(g, 2 * n, 1, x.data(), bb.data(), vib.data()));       // viB (and K again)
  drynx::check(drx_rep_commit(g, 2 * n, 1, blind.data(), bb.data(), v12.data()));
  drynx::check(drx_rep_commit(g, n, 2, blind.data(), qr.data(), v3.data()));
  for (size_t i = 0; i < n; i++) {
    uint8_t V[192];
    memcpy(V, &v12[128 * i], 128);
    memcpy(V + 128, &v3[64 * i], 64);
    const Scalar c = drynxproof::challenge_fn()("proofTest", V, 192);
    memcpy(&chal[32 * i], c.b, 32);
  }
  drynx::check(drx_rep_respond(g, 2 * n, 2, blind.data(), chal.data(), x.data(), r.data()));
  for (size_t i = 0; i < n; i++) {
    auto& p = l.List[i];
    p.Proof.resize(256);
    memcpy(p.Proof.data(), &v12[128 * i], 128);
    memcpy(p.Proof.data() + 128, &v3[64 * i], 64);
    memcpy(p.Proof.data() + 192, &r[64 * i], 64);
    p.K = pubKey;
    memcpy(p.ViB.b, &vib[128 * i], 64);
    p.Ks2 = ks2s[i];
    p.RbNeg = rBNegs[i];
    p.Q = targetPubKey;
  }
  return l;
}
// Accept bit of each of the first `count` proofs: two batched device calls.
inline std::vector<uint8_t> KeySwitchVerifyEach(const PublishedKSListProof& l, size_t count) {
  const size_t n = std::min(count, l.List.size());
  std::vector<uint8_t> out(n, 1);
  if (!n) return out;
  const Point B = drynx::base_point();
  std::vector<uint8_t> r1(64 * n), c1(64 * n), b1(128 * n), p1(128 * n), v1(128 * n), c2(32 * n), b2(128 * n), p2(64 * n),
      v2(64 * n), ok1(2 * n), ok2(n);
  for (size_t i = 0; i < n; i++) {
    const auto& p = l.List[i];
    if (p.Proof.size() != 256) {
      out[i] = 0;
      continue;
    }
    const Scalar c = drynxproof::challenge_fn()("proofTest", p.Proof.data(), 192);
    memcpy(&r1[64 * i], p.Proof.data() + 192, 64);
    memcpy(&c1[64 * i], c.b, 32);
    memcpy(&c1[64 * i + 32], c.b, 32);
    memcpy(&c2[32 * i], c.b, 32);
    memcpy(&b1[128 * i], B.b, 64);
    memcpy(&b1[128 * i + 64], B.b, 64);
    memcpy(&p1[128 * i], p.ViB.b, 64);
    memcpy(&p1[128 * i + 64], p.K.b, 64);
    memcpy(&v1[128 * i], p.Proof.data(), 128);
    memcpy(&b2[128 * i], p.Q.b, 64);
    memcpy(&b2[128 * i + 64], p.RbNeg.b, 64);
    memcpy(&p2[64 * i], p.Ks2.b, 64);
    memcpy(&v2[64 * i], p.Proof.data() + 128, 64);
  }
  drx_ctx* g = Gpu::get().ctx();
  drynx::check(drx_rep_verify(g, 2 * n, 1, r1.data(), c1.data(), b1.data(), p1.data(), v1.data(), ok1.data()));
  drynx::check(drx_rep_verify(g, n, 2, r1.data(), c2.data(), b2.data(), p2.data(), v2.data(), ok2.data()));
  for (size_t i = 0; i < n; i++) out[i] = out[i] && ok1[2 * i] && ok1[2 * i + 1] && ok2[i];
  return out;
}
// KeySwitchListProofVerification(pkslp, percent): the first ceil(percent * len) proofs must all hold.
inline bool KeySwitchListProofVerification(const PublishedKSListProof& l, double percent) {
  const size_t nb = (size_t)std::ceil(percent * (double)l.List.size());
  for (uint8_t b : KeySwitchVerifyEach(l, nb))
    if (!b) return false;
  return true;
}

}  // namespace libunlynxkeyswitch

// =====================================================================================================================
// lib/obfuscation/obfuscation_proof.go: the proof that co = s*c componentwise for a secret s, the predicate
// And(Rep(co1 = s*c1), Rep(co2 = s*c2)); proof bytes in kyber's order V1 || V2 || r_s (160 bytes).
namespace libdrynxobfuscation {
using drynx::Gpu;
using drynx::Scalar;
using libunlynx::CipherText;

struct PublishedObfuscationProof {  // obfuscation_proof.go:12-16
  CipherText C, Co;
  drynx::Bytes Proof;
};
struct PublishedListObfuscationProof {  // obfuscation_proof.go:26-28
  std::vector<PublishedObfuscationProof> Prs;
};
// ObfuscationListProofCreation(c, co, s) (obfuscation_proof.go:62-77; protocols/obfuscation_protocol.go:259,
// data/data.go:57): where the Go code starts one goroutine per proof, one commitment call and one response call
// cover the list. A ciphertext is c1 || c2, so the commitments V1 || V2 = v*c1 || v*c2 are drx_ct_scalar_mul of c by v.
inline PublishedListObfuscationProof ObfuscationListProofCreation(const libunlynx::CipherVector& c, const libunlynx::CipherVector& co,
                                                                  const std::vector<Scalar>& s) {
  const size_t n = s.size();
  PublishedListObfuscationProof l;
  l.Prs.resize(n);
  if (!n) return l;
  std::vector<Scalar> v(n), chal(n), r(n);
  for (auto& x : v) x = drynx::pick();
  libunlynx::CipherVector V(n);
  drx_ctx* g = Gpu::get().ctx();
  drynx::check(drx_ct_scalar_mul(g, n, v[0].b, c[0].b, V[0].b));
  for (size_t i = 0; i < n; i++) chal[i] = drynxproof::challenge_fn()("proofTest", V[i].b, 128);
  drynx::check(drx_rep_respond(g, n, 1, v[0].b, chal[0].b, s[0].b, r[0].b));
  for (size_t i = 0; i < n; i++) {
    l.Prs[i].C = c[i];
    l.Prs[i].Co = co[i];
    l.Prs[i].Proof.resize(160);
    memcpy(l.Prs[i].Proof.data(), V[i].b, 128);
    memcpy(l.Prs[i].Proof.data() + 128, r[i].b, 32);
  }
  return l;
}
inline PublishedObfuscationProof ObfuscationProofCreation(const CipherText& c, const CipherText& co, const Scalar& s) {
  return ObfuscationListProofCreation({c}, {co}, {s}).Prs[0];
}
// Accept bit of each of the first `count` proofs: one device call over 2 * count relations.
inline std::vector<uint8_t> VerifyEach(const PublishedListObfuscationProof& l, size_t count) {
  const size_t n = std::min(count, l.Prs.size());
  std::vector<uint8_t> out(n, 1);
  if (!n) return out;
  std::vector<uint8_t> resp(64 * n), chal(64 * n), bases(128 * n), P(128 * n), V(128 * n), ok(2 * n);
  for (size_t i = 0; i < n; i++) {
    const auto& p = l.Prs[i];
    if (p.Proof.size() != 160) {
      out[i] = 0;
      continue;
    }
    const Scalar c = drynxproof::challenge_fn()("proofTest", p.Proof.data(), 128);
    for (int h = 0; h < 2; h++) {
      memcpy(&resp[64 * i + 32 * h], p.Proof.data() + 128, 32);
      memcpy(&chal[64 * i + 32 * h], c.b, 32);
    }
    memcpy(&bases[128 * i], p.C.b, 128);
    memcpy(&P[128 * i], p.Co.b, 128);
    memcpy(&V[128 * i], p.Proof.data(), 128);
  }
  drynx::check(drx_rep_verify(Gpu::get().ctx(), 2 * n, 1, resp.data(), chal.data(), bases.data(), P.data(), V.data(), ok.data()));
  for (size_t i = 0; i < n; i++) out[i] = out[i] && ok[2 * i] && ok[2 * i + 1];
  return out;
}
inline bool ObfuscationProofVerification(const PublishedObfuscationProof& pop) {  // obfuscation_proof.go:80-91
  PublishedListObfuscationProof l;
  l.Prs.push_back(pop);
  return VerifyEach(l, 1)[0] != 0;
}
// ObfuscationListProofVerification(plop, percent) (obfuscation_proof.go:94-111)
inline bool ObfuscationListProofVerification(const PublishedListObfuscationProof& plop, double percent) {
  const size_t nb = (size_t)std::ceil(percent * (double)plop.Prs.size());
  for (uint8_t b : VerifyEach(plop, nb))
    if (!b) return false;
  return true;
}
}  // namespace libdrynxobfuscation

// =====================================================================================================================
namespace libdrynx {
// PublishSignatureBytes (lib/structs.go:131-141): one server's public key and its u signatures A_k, in wire form.
struct PublishSignatureBytes {
  drynx::Point Public;
  drynx::Bytes Signature;  // u * 128 bytes
};
}  // namespace libdrynx

namespace libdrynxrange {
using drynx::Gpu;
using drynx::Point;
using drynx::Scalar;

// ToBase (lib/range/range_proof.go:584-595)
inline std::vector<int64_t> ToBase(int64_t n, int64_t b, int l) {
  std::vector<int64_t> digits;
  while (n > 0) {
    digits.push_back(n % b);
    n /= b;
  }
  while ((int)digits.size() < l) digits.push_back(0);
  return digits;
}

// InitRangeProofSignature(u) (:270-288): draws the server's private key from the random stream.
inline libdrynx::PublishSignatureBytes InitRangeProofSignatureFromKey(const Scalar& x, int64_t u) {
  libdrynx::PublishSignatureBytes s;
  s.Signature.resize((size_t)u * 128);
  drynx::check(drx_rp_make_sigs(Gpu::get().ctx(), 1, u, x.b, s.Public.b, s.Signature.data()));
  return s;
}
inline libdrynx::PublishSignatureBytes InitRangeProofSignature(int64_t u) { return InitRangeProofSignatureFromKey(drynx::pick(), u); }
// InitRangeProofSignatureDeterministic (:249-267): x = 12
inline libdrynx::PublishSignatureBytes InitRangeProofSignatureDeterministic(int64_t u) {
  return InitRangeProofSignatureFromKey(drynx::scalar_from_u64(12), u);
}

// The signatures of all servers for one output column, registered with the device (ReadColumn(sigs, i), :598-612).
struct SignatureColumn {
  int set_id = -1;
  int ns = 0;
  int64_t u = 0;
  std::vector<Point> Ys;
};
inline SignatureColumn RegisterColumn(const std::vector<libdrynx::PublishSignatureBytes>& servers, int64_t u) {
  static int next_id = 0;
  SignatureColumn c;
  c.set_id = next_id++;
  c.ns = (int)servers.size();
  c.u = u;
  drynx::Bytes Y, A;
  for (auto& s : servers) {
    Y.insert(Y.end(), s.Public.b, s.Public.b + 64);
    A.insert(A.end(), s.Signature.begin(), s.Signature.end());
    c.Ys.push_back(s.Public);
  }
  drynx::check(drx_rp_set_sigs(Gpu::get().ctx(), c.set_id, c.ns, u, Y.data(), A.data()));
  return c;
}

// CreateProof (:61-69). Sigs is the registered column; U = L = 0 means "no proof" (:322).
struct CreateProof {
  SignatureColumn Sigs;
  int64_t U = 0, L = 0, Secret = 0;
  Scalar R{};
  Point CaPub{};
  libunlynx::CipherText Cipher{};
};
// RangeProof in its serialised form (RangeProofBytes field order, :88-155).
struct RangeProof {
  drynx::Bytes bytes;
};
struct RangeProofList {
  std::vector<RangeProof> Data;
};

// CreatePredicateRangeProofListForAllServers (:305-317): one batched creation for the whole list.
inline std::vector<RangeProof> CreatePredicateRangeProofListForAllServers(const std::vector<CreateProof>& cps) {
  const size_t n = cps.size();
  std::vector<RangeProof> out(n);
  if (!n) return out;
  libunlynx::use_pubkey(cps[0].CaPub);
  std::vector<int32_t> sid(n);
  std::vector<int64_t> l(n), secret(n);
  std::vector<size_t> poff(n + 1, 0), roff(n + 1, 0);
  drynx::Bytes r(32 * n), ct(128 * n), rnd;
  for (size_t p = 0; p < n; p++) {
    const CreateProof& c = cps[p];
    const bool none = c.U == 0 && c.L == 0;
    sid[p] = none ? -1 : c.Sigs.set_id;
    l[p] = c.L;
    secret[p] = c.Secret;
    memcpy(&r[32 * p], c.R.b, 32);
    memcpy(&ct[128 * p], c.Cipher.b, 128);
    int64_t nd = none ? 0 : drx_rp_ndigits(c.Secret, c.U, c.L);
    poff[p + 1] = poff[p] + (none ? 128 : drx_rp_proof_size(c.Sigs.ns, nd));
    // draw order of :377-379 and :390: per digit s_j, t_j, m_j, then one v per server
    for (int64_t k = 0; k < nd * (3 + (none ? 0 : c.Sigs.ns)); k++) {
      Scalar s = drynx::pick();
      rnd.insert(rnd.end(), s.b, s.b + 32);
    }
    roff[p + 1] = rnd.size();
  }
  drynx::Bytes proofs(poff[n]);
  rnd.push_back(0);
  drynx::check(drx_rp_create(Gpu::get().ctx(), n, sid.data(), l.data(), secret.data(), r.data(), ct.data(), rnd.data(), roff.data(),
                             proofs.data(), poff.data()));
  for (size_t p = 0; p < n; p++) out[p].bytes.assign(proofs.begin() + poff[p], proofs.begin() + poff[p + 1]);
  return out;
}
inline RangeProof CreatePredicateRangeProofForAllServ(const CreateProof& cp) {
  return CreatePredicateRangeProofListForAllServers({cp})[0];
}

// RangeProofListVerification (:484-501): verifies the first ceil(threshold * N) proofs; ranges[i] = {u, l};
// columns[i] = the signature column of proof i (its Ys are what ReadColumnYs returns).
inline std::vector<uint8_t> VerifyEach(const RangeProofList& list, const std::vector<std::pair<int64_t, int64_t>>& ranges,
                                       const std::vector<SignatureColumn>& columns, const Point& P, size_t count) {
  std::vector<uint8_t> ok(count);
  if (!count) return ok;
  libunlynx::use_pubkey(P);
  std::vector<int32_t> sid(count);
  std::vector<int64_t> l(count);
  std::vector<size_t> off(count + 1, 0);
  drynx::Bytes all;
  for (size_t i = 0; i < count; i++) {
    const bool none = ranges[i].first == 0 && ranges[i].second == 0;
    sid[i] = none ? -1 : columns[i].set_id;
    l[i] = ranges[i].second;
    // a commit-only proof carries 128 bytes; anything else is passed through as received
    if (none && list.Data[i].bytes.size() != 128) {
      all.insert(all.end(), list.Data[i].bytes.begin(), list.Data[i].bytes.begin() + 128);
    } else {
      all.insert(all.end(), list.Data[i].bytes.begin(), list.Data[i].bytes.end());
    }
    off[i + 1] = all.size();
  }
  drynx::check(drx_rp_verify(Gpu::get().ctx(), count, sid.data(), l.data(), all.data(), off.data(), ok.data()));
  return ok;
}
inline bool RangeProofListVerification(const RangeProofList& list, const std::vector<std::pair<int64_t, int64_t>>& ranges,
                                       const std::vector<SignatureColumn>& columns, const Point& P, double verifThreshold) {
  const size_t nbr = (size_t)std::ceil(verifThreshold * (double)list.Data.size());
  bool result = true;
  for (uint8_t v : VerifyEach(list, ranges, columns, P, nbr)) result = result && v;
  return result;
}
inline bool RangeProofVerification(const RangeProof& rp, int64_t u, int64_t l, const SignatureColumn& col, const Point& P) {
  RangeProofList lst;
  lst.Data.push_back(rp);
  return RangeProofListVerification(lst, {{u, l}}, {col}, P, 1.0);
}

}  // namespace libdrynxrange

// =====================================================================================================================
namespace libdrynxencoding {
using drynx::Point;
using drynx::Scalar;
using libdrynxrange::CreateProof;
using libdrynxrange::SignatureColumn;
using libunlynx::CipherText;
using libunlynx::CipherVector;

struct Encoded {
  CipherVector enc;
  std::vector<int64_t> clear;
  std::vector<CreateProof> proofs;
};
using Ranges = std::vector<std::pair<int64_t, int64_t>>;  // lu[i] = {u, l}

// Shared tail of every Encode*WithProofs: encrypt the plaintext vector in one call, then describe the proofs.
inline Encoded encrypt_and_describe(const std::vector<int64_t>& clear, const Point& pubKey, const std::vector<SignatureColumn>* sigs,
                                    const Ranges* lu) {
  Encoded e;
  e.clear = clear;
  auto er = libunlynx::EncryptIntVectorGetRs(pubKey, clear);
  e.enc = er.first;
  if (!sigs) return e;
  e.proofs.resize(clear.size());
  for (size_t i = 0; i < clear.size(); i++) {
    CreateProof& cp = e.proofs[i];
    cp.Sigs = (*sigs)[i];
    cp.U = (*lu)[i].first;
    cp.L = (*lu)[i].second;
    cp.Secret = clear[i];
    cp.R = er.second[i];
    cp.CaPub = pubKey;
    cp.Cipher = e.enc[i];
  }
  return e;
}

// ---- sum (lib/encoding/sum.go:17-43)
inline Encoded EncodeSumWithProofs(const std::vector<int64_t>& input, const Point& pubKey, const std::vector<SignatureColumn>* sigs, const Ranges* lu) {
  int64_t sum = 0;
  for (int64_t el : input) sum += el;
  return encrypt_and_describe({sum}, pubKey, sigs, lu);
}
inline Encoded EncodeSum(const std::vector<int64_t>& input, const Point& pubKey) { return EncodeSumWithProofs(input, pubKey, nullptr, nullptr); }
inline int64_t DecodeSum(const CipherText& result, const Scalar& secKey) { return libunlynx::DecryptIntWithNeg(secKey, result); }

// ---- mean (mean.go:17-70): [sum, N]
inline Encoded EncodeMeanWithProofs(const std::vector<int64_t>& input, const Point& pubKey, const std::vector<SignatureColumn>* sigs, const Ranges* lu) {
  int64_t sum = 0;
  for (int64_t el : input) sum += el;
  return encrypt_and_describe({sum, (int64_t)input.size()}, pubKey, sigs, lu);
}
inline Encoded EncodeMean(const std::vector<int64_t>& input, const Point& pubKey) { return EncodeMeanWithProofs(input, pubKey, nullptr, nullptr); }
inline double DecodeMean(const CipherVector& result, const Scalar& secKey) {
  auto c = libunlynx::DecryptIntVectorWithNeg(secKey, result);
  return (double)c[0] / (double)c[1];
}

// ---- variance (variance.go:17-75): [sum, N, sum of squares]; Var = E[x^2] - E[x]^2
inline Encoded EncodeVarianceWithProofs(const std::vector<int64_t>& input, const Point& pubKey, const std::vector<SignatureColumn>* sigs, const Ranges* lu) {
  int64_t sum = 0, sq = 0;
  for (int64_t el : input) {
    sum += el;
    sq += el * el;
  }
  return encrypt_and_describe({sum, (int64_t)input.size(), sq}, pubKey, sigs, lu);
}
inline Encoded EncodeVariance(const std::vector<int64_t>& input, const Point& pubKey) { return EncodeVarianceWithProofs(input, pubKey, nullptr, nullptr); }
inline double DecodeVariance(const CipherVector& result, const Scalar& secKey) {
  auto c = libunlynx::DecryptIntVectorWithNeg(secKey, result);
  const double mean = (double)c[0] / (double)c[1];
  return (double)c[2] / (double)c[1] - mean * mean;
}

// ---- frequency count (frequency_count.go:18-78): one counter per value in [min, max]
inline Encoded EncodeFreqCountWithProofs(const std::vector<int64_t>& input, int64_t min, int64_t max, const Point& pubKey,
                                         const std::vector<SignatureColumn>* sigs, const Ranges* lu) {
  std::vector<int64_t> freq((size_t)(max - min + 1), 0);
  for (int64_t el : input) freq[(size_t)(el - min)]++;
  return encrypt_and_describe(freq, pubKey, sigs, lu);
}
inline Encoded EncodeFreqCount(const std::vector<int64_t>& input, int64_t min, int64_t max, const Point& pubKey) {
  return EncodeFreqCountWithProofs(input, min, max, pubKey, nullptr, nullptr);
}
inline std::vector<int64_t> DecodeFreqCount(const CipherVector& result, const Scalar& secKey) {
  return libunlynx::DecryptIntVectorWithNeg(secKey, result);
}

// ---- bits under OR / AND (OR_AND.go:17-144). Without signatures a set OR-bit (a cleared AND-bit) is the encryption
// of a random non-zero scalar and the other value the encryption of zero; with signatures it is the integer 1 / 0
// (AND: the complement) so that the range proof applies. A whole vector of bits is one batched encryption.
enum class BitOp { Or, And };
inline Scalar nonzero_scalar() {
  for (;;) {
    Scalar s = drynx::pick();
    for (int i = 0; i < 32; i++)
      if (s.b[i]) return s;
  }
}
inline Encoded EncodeBits(const std::vector<bool>& input, BitOp op, const Point& pubKey, const std::vector<SignatureColumn>* sigs,
                          const Ranges* lu) {
  std::vector<int64_t> clear(input.size());
  for (size_t i = 0; i < input.size(); i++)  // OR: toEncrypt = input ? 1 : 0 (:30-33); AND: input ? 0 : 1 (:83-86)
    clear[i] = sigs ? ((op == BitOp::Or) == input[i] ? 1 : 0) : (op == BitOp::Or ? 0 : 1);
  if (sigs) return encrypt_and_describe(clear, pubKey, sigs, lu);
  std::vector<Scalar> ms(input.size());
  for (size_t i = 0; i < input.size(); i++) {
    const bool random = op == BitOp::Or ? input[i] : !input[i];  // :42, :95
    if (random) ms[i] = nonzero_scalar(); else memset(ms[i].b, 0, 32);
  }
  Encoded e;
  e.clear = clear;  // the reference returns its initial toEncrypt here (0 for OR, 1 for AND), whatever the bit
  e.enc = libunlynx::EncryptScalarVector(pubKey, ms);
  return e;
}
inline std::pair<CipherText, int64_t> EncodeBitOr(bool input, const Point& pubKey) {
  auto e = EncodeBits({input}, BitOp::Or, pubKey, nullptr, nullptr);
  return {e.enc[0], e.clear[0]};
}
inline std::pair<CipherText, int64_t> EncodeBitAND(bool input, const Point& pubKey) {
  auto e = EncodeBits({input}, BitOp::And, pubKey, nullptr, nullptr);
  return {e.enc[0], e.clear[0]};
}
inline Encoded EncodeBitOrWithProof(bool input, const Point& pubKey, const SignatureColumn& sigs, int64_t l, int64_t u) {
  std::vector<SignatureColumn> s = {sigs};
  Ranges lu = {{u, l}};
  return EncodeBits({input}, BitOp::Or, pubKey, &s, &lu);
}
inline Encoded EncodeBitANDWithProof(bool input, const Point& pubKey, const SignatureColumn& sigs, int64_t l, int64_t u) {
  std::vector<SignatureColumn> s = {sigs};
  Ranges lu = {{u, l}};
  return EncodeBits({input}, BitOp::And, pubKey, &s, &lu);
}
inline bool DecodeBitOR(const CipherText& result, const Scalar& secKey) { return libunlynx::DecryptCheckZero(secKey, result) != 0; }
inline bool DecodeBitAND(const CipherText& result, const Scalar& secKey) { return libunlynx::DecryptCheckZero(secKey, result) == 0; }
inline bool LocalResultOR(const std::vector<bool>& input) {
  for (bool b : input)
    if (b) return true;
  return false;
}
inline bool LocalResultAND(const std::vector<bool>& input) {
  for (bool b : input)
    if (!b) return false;
  return true;
}

// ---- min / max (min_max.go:12-146): one OR-bit (min) or AND-bit (max) per value of [min, max]; bit i is
// "i >= local extreme". The reference takes (max, min) in that order.
inline Encoded EncodeMinWithProofs(const std::vector<int64_t>& input, int64_t max, int64_t min, const Point& pubKey,
                                   const std::vector<SignatureColumn>* sigs, const Ranges* lu) {
  int64_t localMin = input[0];
  for (int64_t v : input) localMin = v < localMin ? v : localMin;
  std::vector<bool> bits((size_t)(max - min + 1));
  for (int64_t i = min; i <= max; i++) bits[(size_t)(i - min)] = i >= localMin;
  return EncodeBits(bits, BitOp::Or, pubKey, sigs, lu);
}
inline Encoded EncodeMin(const std::vector<int64_t>& input, int64_t max, int64_t min, const Point& pubKey) {
  return EncodeMinWithProofs(input, max, min, pubKey, nullptr, nullptr);
}
inline int64_t DecodeMin(const CipherVector& result, int64_t globalMin, const Scalar& secKey) {
  auto z = libunlynx::DecryptCheckZeroVector(secKey, result);
  for (size_t i = 0; i < z.size(); i++)
    if (z[i] != 0) return (int64_t)i + globalMin;  // first set OR-bit
  return 0;
}
inline Encoded EncodeMaxWithProofs(const std::vector<int64_t>& input, int64_t max, int64_t min, const Point& pubKey,
                                   const std::vector<SignatureColumn>* sigs, const Ranges* lu) {
  int64_t localMax = input[0];
  for (int64_t v : input) localMax = v > localMax ? v : localMax;
  std::vector<bool> bits((size_t)(max - min + 1));
  for (int64_t i = min; i <= max; i++) bits[(size_t)(i - min)] = i >= localMax;
  return EncodeBits(bits, BitOp::And, pubKey, sigs, lu);
}
inline Encoded EncodeMax(const std::vector<int64_t>& input, int64_t max, int64_t min, const Point& pubKey) {
  return EncodeMaxWithProofs(input, max, min, pubKey, nullptr, nullptr);
}
inline int64_t DecodeMax(const CipherVector& result, int64_t globalMin, const Scalar& secKey) {
  auto z = libunlynx::DecryptCheckZeroVector(secKey, result);
  for (size_t i = 0; i < z.size(); i++)
    if (z[i] == 0) return (int64_t)i + globalMin;  // first set AND-bit
  return 0;
}

// ---- set union / intersection (set_union_intersection.go:12-169): membership bits over [min, max]
inline std::vector<int64_t> Unique(const std::vector<int64_t>& in) {
  std::vector<int64_t> out;
  for (int64_t v : in) {
    bool seen = false;
    for (int64_t w : out) seen |= w == v;
    if (!seen) out.push_back(v);
  }
  return out;
}
inline Encoded EncodeSetWithProofs(const std::vector<int64_t>& input, int64_t min, int64_t max, BitOp op, const Point& pubKey,
                                   const std::vector<SignatureColumn>* sigs, const Ranges* lu) {
  std::vector<bool> bits((size_t)(max - min + 1), false);
  for (int64_t v : Unique(input)) bits[(size_t)(v - min)] = true;
  return EncodeBits(bits, op, pubKey, sigs, lu);
}
inline Encoded EncodeUnion(const std::vector<int64_t>& input, int64_t min, int64_t max, const Point& pubKey) {
  return EncodeSetWithProofs(input, min, max, BitOp::Or, pubKey, nullptr, nullptr);
}
inline Encoded EncodeInter(const std::vector<int64_t>& input, int64_t min, int64_t max, const Point& pubKey) {
  return EncodeSetWithProofs(input, min, max, BitOp::And, pubKey, nullptr, nullptr);
}
inline std::vector<int64_t> DecodeUnion(const CipherVector& result, const Scalar& secKey) {
  auto z = libunlynx::DecryptCheckZeroVector(secKey, result);
  for (auto& v : z) v = v != 0 ? 1 : 0;
  return z;
}
inline std::vector<int64_t> DecodeInter(const CipherVector& result, const Scalar& secKey) {
  auto z = libunlynx::DecryptCheckZeroVector(secKey, result);
  for (auto& v : z) v = v == 0 ? 1 : 0;
  return z;
}

// ---- cosine similarity (cosim.go:11-86): [sum a, sum b, sum a^2, sum b^2, sum ab]
inline Encoded EncodeCosimWithProofs(const std::vector<int64_t>& rijs, const std::vector<int64_t>& riks, const Point& pubKey,
                                     const std::vector<SignatureColumn>* sigs, const Ranges* lu) {
  int64_t sa = 0, sb = 0, sa2 = 0, sb2 = 0, sab = 0;
  for (size_t i = 0; i < rijs.size(); i++) {
    const int64_t a = rijs[i], b = riks[i];
    sa += a; sb += b; sa2 += a * a; sb2 += b * b; sab += a * b;
  }
  return encrypt_and_describe({sa, sb, sa2, sb2, sab}, pubKey, sigs, lu);
}
inline Encoded EncodeCosim(const std::vector<int64_t>& rijs, const std::vector<int64_t>& riks, const Point& pubKey) {
  return EncodeCosimWithProofs(rijs, riks, pubKey, nullptr, nullptr);
}
inline double DecodeCosim(const CipherVector& result, const Scalar& secKey) {
  auto c = libunlynx::DecryptIntVectorWithNeg(secKey, result);
  return (double)c[4] / (std::sqrt((double)c[2]) * std::sqrt((double)c[3]));
}

// ---- model evaluation, R^2 (model_evaluation.go:11-96): [N, sum y, sum y^2, sum (pred - y)^2]
inline Encoded EncodeModelEvaluationWithProofs(const std::vector<int64_t>& y, const std::vector<int64_t>& preds, const Point& pubKey,
                                               const std::vector<SignatureColumn>* sigs, const Ranges* lu) {
  int64_t sy = 0, sy2 = 0, sd2 = 0;
  for (size_t i = 0; i < y.size(); i++) {
    sy += y[i];
    sy2 += y[i] * y[i];
    sd2 += (preds[i] - y[i]) * (preds[i] - y[i]);
  }
  Encoded e = encrypt_and_describe({(int64_t)y.size(), sy, sy2, sd2}, pubKey, sigs, lu);
  e.clear = {0};  // the reference returns []int64{0} as the clear vector (:63, :76)
  return e;
}
inline Encoded EncodeModelEvaluation(const std::vector<int64_t>& y, const std::vector<int64_t>& preds, const Point& pubKey) {
  return EncodeModelEvaluationWithProofs(y, preds, pubKey, nullptr, nullptr);
}
inline double DecodeModelEvaluation(const CipherVector& result, const Scalar& secKey) {
  auto c = libunlynx::DecryptIntVectorWithNeg(secKey, result);
  const double B = (double)c[2] - (double)(c[1] * c[1] / c[0]);  // integer division, as in the reference (:93)
  return 1.0 - (double)c[3] / B;
}

// ---- linear regression, d dimensions (linear_regression_dims.go:23-106)
// plaintext order: N | sum x_j (d) | sum x_j x_k for j <= k | sum y | sum x_j y (d)   -> (d^2 + 5d + 4) / 2 values
inline std::vector<int64_t> LinearRegressionDimsPlaintext(const std::vector<std::vector<int64_t>>& X, const std::vector<int64_t>& y) {
  const size_t d = X[0].size(), N = X.size();
  std::vector<int64_t> out, stored;
  out.push_back((int64_t)N);
  for (size_t j = 0; j < d; j++) {
    int64_t sx = 0, sxy = 0;
    for (size_t i = 0; i < N; i++) {
      sx += X[i][j];
      sxy += y[i] * X[i][j];
    }
    out.push_back(sx);
    stored.push_back(sxy);
  }
  for (size_t j = 0; j < d; j++)
    for (size_t k = j; k < d; k++) {
      int64_t s = 0;
      for (size_t i = 0; i < N; i++) s += X[i][j] * X[i][k];
      out.push_back(s);
    }
  int64_t sy = 0;
  for (int64_t el : y) sy += el;
  out.push_back(sy);
  out.insert(out.end(), stored.begin(), stored.end());
  return out;
}
inline Encoded EncodeLinearRegressionDimsWithProofs(const std::vector<std::vector<int64_t>>& X, const std::vector<int64_t>& y, const Point& pubKey,
                                                    const std::vector<SignatureColumn>* sigs, const Ranges* lu) {
  Encoded e = encrypt_and_describe(LinearRegressionDimsPlaintext(X, y), pubKey, sigs, lu);
  e.clear = {0};  // the reference returns []int64{0} as the clear result (:94, :105)
  return e;
}
inline Encoded EncodeLinearRegressionDims(const std::vector<std::vector<int64_t>>& X, const std::vector<int64_t>& y, const Point& pubKey) {
  return EncodeLinearRegressionDimsWithProofs(X, y, pubKey, nullptr, nullptr);
}
// DecodeLinearRegressionDims (:110-163): rebuild the symmetric normal-equation matrix and solve it. The reference
// solves over the rationals; this mirror uses Gaussian elimination with partial pivoting in long double.
inline std::vector<double> DecodeLinearRegressionDims(const CipherVector& result, const Scalar& secKey) {
  const auto v = libunlynx::DecryptIntVectorWithNeg(secKey, result);
  const int len = (int)v.size();
  const int d = (int)std::llround((-5.0 + std::sqrt(25.0 - 4.0 * (4.0 - 2.0 * len))) / 2.0);
  std::vector<std::vector<long double>> M(d + 1, std::vector<long double>(d + 2, 0));
  int s = 0, l = d + 1, k = d + 1, i = 0;
  for (int j = 0; j < len - d - 1; j++) {
    if (j == l) {
      k--;
      l += k;
      i++;
      s = 0;
    }
    M[i][i + s] = (long double)v[j];
    if (s != 0) M[i + s][i] = (long double)v[j];
    s++;
  }
  for (int j = len - d - 1; j < len; j++) M[j - len + d + 1][d + 1] = (long double)v[j];
  const int n = d + 1;
  for (int c = 0; c < n; c++) {
    int piv = c;
    for (int r = c + 1; r < n; r++)
      if (fabsl(M[r][c]) > fabsl(M[piv][c])) piv = r;
    std::swap(M[c], M[piv]);
    for (int r = 0; r < n; r++) {
      if (r == c || M[c][c] == 0) continue;
      const long double f = M[r][c] / M[c][c];
      for (int q = c; q <= n; q++) M[r][q] -= f * M[c][q];
    }
  }
  std::vector<double> coeffs(n);
  for (int r = 0; r < n; r++) coeffs[r] = (double)(M[r][n] / M[r][r]);
  return coeffs;
}

// ---- logistic regression pre-pass (logistic_regression.go; float64, evaluated in the reference's order)
struct LogisticRegressionParameters {  // the fields of libdrynx.LogisticRegressionParameters this path reads
  int64_t NbrFeatures = 0;
  int64_t NbrRecords = 0;
  int K = 2;
  double PrecisionApproxCoefficients = 1;
  std::vector<double> Means, StandardDeviations;
};
inline int64_t Factorial(int64_t n) { return n > 0 ? n * Factorial(n - 1) : 1; }
// CombinationsWithRepetition (:281-303) = C(n + k - 1, k)
inline int64_t CombinationsWithRepetition(int64_t n, int64_t k) {
  if (k > n) {  // the reference's reduction loop only runs for k <= n and otherwise returns 1
    return 1;
  }
  int64_t numerator = 1, denominator = Factorial(k), divisor = 2;
  for (int64_t i = 1; i <= k; i++) {
    numerator *= (n + k - i);
    while (divisor <= k && numerator % divisor == 0 && denominator % divisor == 0) {
      numerator /= divisor;
      denominator /= divisor;
      divisor++;
    }
  }
  return numerator / denominator;
}
inline int getNumberApproxCoefficients(int64_t d, int level) { return (int)std::pow((double)(d + 1), (double)(level + 1)); }
inline int getTotalNumberApproxCoefficients(int64_t d, int k) {
  int c = 0;
  for (int j = 0; j < k; j++) c += (int)std::pow((double)(d + 1), (double)(j + 1));
  return c;
}
// CartesianProduct(start, end, dim): all index tuples in lexicographic order (logistic_regression_test.go:544-558)
inline std::vector<std::vector<int64_t>> CartesianProduct(int64_t start, int64_t end, int dim) {
  std::vector<std::vector<int64_t>> out;
  std::vector<int64_t> cur((size_t)dim, start);
  if (end <= start) return out;
  for (;;) {
    out.push_back(cur);
    int pos = dim - 1;
    while (pos >= 0 && ++cur[(size_t)pos] == end) cur[(size_t)pos--] = start;
    if (pos < 0) break;
  }
  return out;
}
// ComputeDistinctApproxCoefficients (:322-364)
inline std::vector<std::vector<double>> ComputeDistinctApproxCoefficients(const std::vector<double>& X, int64_t y, int k) {
  const int d = (int)X.size() - 1;
  std::vector<std::vector<double>> a((size_t)k);
  for (int j = 0; j < k; j++) a[(size_t)j].assign((size_t)CombinationsWithRepetition(d + 1, j + 1), 0.0);
  for (int s = 0; s <= d; s++) a[0][(size_t)s] = X[(size_t)s] * (2 * (double)y - 1);
  for (int j = 1; j < k; j++) {
    const int total = (int)a[(size_t)j].size();
    const int64_t ypart = y - y * (int64_t)std::pow(-1.0, (double)(j + 1)) - 1;
    int idx1 = 0, idx2 = 0;
    int64_t c = CombinationsWithRepetition(d + 1 - idx2, j), cPrev = c;
    for (int ri = 0; ri < total; ri++) {
      a[(size_t)j][(size_t)ri] = a[(size_t)j - 1][(size_t)idx1] * X[(size_t)idx2] * (double)ypart;
      idx1++;
      c--;
      if (c == 0) {
        idx2++;
        c = CombinationsWithRepetition(d + 1 - idx2, j);
        idx1 = (int)(cPrev - c);
      }
    }
  }
  return a;
}
// ComputeAllApproxCoefficients (:367-403)
inline std::vector<std::vector<double>> ComputeAllApproxCoefficients(const std::vector<double>& X, int64_t y, int k) {
  const int d = (int)X.size() - 1;
  std::vector<std::vector<double>> a((size_t)k);
  for (int j = 0; j < k; j++) a[(size_t)j].assign((size_t)std::pow((double)(d + 1), (double)(j + 1)), 0.0);
  for (int s = 0; s <= d; s++) a[0][(size_t)s] = X[(size_t)s] * (2 * (double)y - 1);
  for (int j = 2; j <= k; j++) {
    const double ypart = (double)(y - y * (int64_t)std::pow(-1.0, (double)j) - 1);
    const auto comb = CartesianProduct(0, d + 1, j);
    for (size_t ri = 0; ri < comb.size(); ri++) {
      double xp = 1.0;
      for (int64_t idx : comb[ri]) xp *= X[(size_t)idx];
      a[(size_t)j - 1][ri] = ypart * xp;
    }
  }
  return a;
}
// AggregateApproxCoefficients (:433-455): record-by-record sums in index order (the float64 order is part of the result)
inline std::vector<std::vector<double>> AggregateApproxCoefficients(const std::vector<std::vector<std::vector<double>>>& all) {
  std::vector<std::vector<double>> agg(all[0].size());
  for (size_t j = 0; j < agg.size(); j++) agg[j].assign(all[0][j].size(), 0.0);
  for (auto& rec : all)
    for (size_t j = 0; j < rec.size(); j++)
      for (size_t ri = 0; ri < rec[j].size(); ri++) agg[j][ri] += rec[j][ri];
  return agg;
}
inline std::vector<std::vector<double>> StandardiseWith(const std::vector<std::vector<double>>& data, const std::vector<double>& means,
                                                        const std::vector<double>& sds) {
  auto out = data;
  for (auto& rec : out)
    for (size_t i = 0; i < rec.size(); i++) rec[i] = (rec[i] - means[i]) / sds[i];
  return out;
}
inline std::vector<std::vector<double>> Augment(const std::vector<std::vector<double>>& m) {
  auto out = m;
  for (auto& rec : out) rec.insert(rec.begin(), 1.0);
  return out;
}
inline std::vector<std::vector<int64_t>> Float64ToInt642DArrayWithPrecision(const std::vector<std::vector<double>>& a, double precision) {
  std::vector<std::vector<int64_t>> out(a.size());
  for (size_t i = 0; i < a.size(); i++)
    for (double x : a[i]) out[i].push_back((int64_t)std::round(x * precision));
  return out;
}
// The packed int64 coefficient vector of EncodeLogisticRegression (:46-118): level j at offset j * (d+1)^j.
inline std::vector<int64_t> LogisticRegressionPlaintext(const std::vector<std::vector<double>>& xData, const std::vector<int64_t>& yData,
                                                        const LogisticRegressionParameters& lr) {
  const int64_t d = lr.NbrFeatures;
  // the reference places level j at offset j * (d+1)^j (:306-313), which only tiles the vector up to k = 2
  if (lr.K < 1 || lr.K > 2) throw drynx::Error(-1, "logistic regression: the coefficient packing is only defined for k <= 2");
  std::vector<int64_t> packed((size_t)getTotalNumberApproxCoefficients(d, lr.K), 0);
  if (xData.empty()) return packed;
  auto Xs = Augment(StandardiseWith(xData, lr.Means, lr.StandardDeviations));
  std::vector<std::vector<std::vector<double>>> all;
  for (size_t i = 0; i < Xs.size(); i++) all.push_back(ComputeAllApproxCoefficients(Xs[i], yData[i], lr.K));
  const auto ints = Float64ToInt642DArrayWithPrecision(AggregateApproxCoefficients(all), lr.PrecisionApproxCoefficients);
  for (int j = 0; j < lr.K; j++) {
    const int nLevel = getNumberApproxCoefficients(d, j), nPrev = getNumberApproxCoefficients(d, j - 1);
    for (int i = 0; i < nLevel; i++) packed[(size_t)(j * nPrev + i)] = ints[(size_t)j][(size_t)i];
  }
  return packed;
}
inline Encoded EncodeLogisticRegressionWithProofs(const std::vector<std::vector<double>>& xData, const std::vector<int64_t>& yData,
                                                  const LogisticRegressionParameters& lr, const Point& pubKey,
                                                  const std::vector<SignatureColumn>* sigs, const Ranges* lu) {
  // the float64 pre-pass runs on the device (k_logreg_coeffs adds in record order with unfused, correctly rounded
  // operations: the integers equal LogisticRegressionPlaintext's, which host/test_host.cpp checks)
  if (lr.K < 1 || lr.K > 2) throw drynx::Error(-1, "logistic regression: the coefficient packing is only defined for k <= 2");
  const size_t d = (size_t)lr.NbrFeatures;
  std::vector<int64_t> packed((size_t)getTotalNumberApproxCoefficients(lr.NbrFeatures, lr.K), 0);
  if (!xData.empty()) {
    std::vector<double> flat;
    flat.reserve(xData.size() * d);
    for (auto& rec : xData) flat.insert(flat.end(), rec.begin(), rec.end());
    const size_t off[2] = {0, xData.size()};
    drynx::check(drx_logreg_coefficients(drynx::Gpu::get().ctx(), 1, off, (int)d, lr.K, flat.data(), yData.data(), lr.Means.data(),
                                         lr.StandardDeviations.data(), lr.PrecisionApproxCoefficients, packed.data()));
  }
  for (int64_t v : packed)
    if (v > libunlynx::MaxHomomorphicInt) throw drynx::Error(-1, "coefficient exceeds MaxHomomorphicInt");  // log.Fatalf at :415
  return encrypt_and_describe(packed, pubKey, sigs, lu);
}
inline Encoded EncodeLogisticRegression(const std::vector<std::vector<double>>& xData, const std::vector<int64_t>& yData,
                                        const LogisticRegressionParameters& lr, const Point& pubKey) {
  return EncodeLogisticRegressionWithProofs(xData, yData, lr, pubKey, nullptr, nullptr);
}

// ---- the dispatcher (encode_decode.go:14-160 Encode, :163-240 Decode): what a data provider / the querier calls
struct Operation {  // the fields of libdrynx.Operation this path reads
  std::string NameOp;
  int64_t QueryMin = 0, QueryMax = 0;
};
inline Encoded Encode(const std::vector<std::vector<int64_t>>& datas, const Point& pubKey, const std::vector<SignatureColumn>* signatures,
                      const Ranges* ranges, const Operation& operation) {
  const bool withProofs = signatures && ranges && !signatures->empty() && !ranges->empty();
  const std::vector<SignatureColumn>* sg = withProofs ? signatures : nullptr;
  const Ranges* lu = withProofs ? ranges : nullptr;
  const std::string& op = operation.NameOp;
  if (op == "sum") return EncodeSumWithProofs(datas[0], pubKey, sg, lu);
  if (op == "cosim") return EncodeCosimWithProofs(datas[0], datas[1], pubKey, sg, lu);
  if (op == "mean") return EncodeMeanWithProofs(datas[0], pubKey, sg, lu);
  if (op == "variance") return EncodeVarianceWithProofs(datas[0], pubKey, sg, lu);
  if (op == "lin_reg") {  // datas is column-major: d-1 feature columns, then y (:56-66)
    const size_t d = datas.size(), n = datas[0].size();
    std::vector<std::vector<int64_t>> X(n, std::vector<int64_t>(d - 1));
    std::vector<int64_t> y(n);
    for (size_t j = 0; j < n; j++) {
      for (size_t i = 0; i + 1 < d; i++) X[j][i] = datas[i][j];
      y[j] = datas[d - 1][j];
    }
    return EncodeLinearRegressionDimsWithProofs(X, y, pubKey, sg, lu);
  }
  if (op == "frequencyCount") return EncodeFreqCountWithProofs(datas[0], operation.QueryMin, operation.QueryMax, pubKey, sg, lu);
  if (op == "bool_AND") return EncodeBits({datas[0][0] == 1}, BitOp::And, pubKey, sg, lu);
  if (op == "bool_OR") return EncodeBits({datas[0][0] == 1}, BitOp::Or, pubKey, sg, lu);
  if (op == "min") return EncodeMinWithProofs(datas[0], operation.QueryMax, operation.QueryMin, pubKey, sg, lu);
  if (op == "max") return EncodeMaxWithProofs(datas[0], operation.QueryMax, operation.QueryMin, pubKey, sg, lu);
  if (op == "union") return EncodeSetWithProofs(datas[0], operation.QueryMin, operation.QueryMax, BitOp::Or, pubKey, sg, lu);
  if (op == "inter") return EncodeSetWithProofs(datas[0], operation.QueryMin, operation.QueryMax, BitOp::And, pubKey, sg, lu);
  if (op == "MLeval") return EncodeModelEvaluationWithProofs(datas[0], datas[1], pubKey, sg, lu);
  return Encoded{};  // unknown operation: empty response, as the Go switch falls through
}
inline std::vector<double> Decode(const CipherVector& ciphers, const Scalar& secKey, const Operation& operation) {
  const std::string& op = operation.NameOp;
  auto as_doubles = [](const std::vector<int64_t>& v) { return std::vector<double>(v.begin(), v.end()); };
  if (op == "sum") return {(double)DecodeSum(ciphers[0], secKey)};
  if (op == "cosim") return {DecodeCosim(ciphers, secKey)};
  if (op == "mean") return {DecodeMean(ciphers, secKey)};
  if (op == "variance") return {DecodeVariance(ciphers, secKey)};
  if (op == "lin_reg") return DecodeLinearRegressionDims(ciphers, secKey);
  if (op == "frequencyCount") return as_doubles(DecodeFreqCount(ciphers, secKey));
  if (op == "min") return {(double)DecodeMin(ciphers, operation.QueryMin, secKey)};
  if (op == "max") return {(double)DecodeMax(ciphers, operation.QueryMin, secKey)};
  if (op == "bool_AND") return {DecodeBitAND(ciphers[0], secKey) ? 1.0 : 0.0};
  if (op == "bool_OR") return {DecodeBitOR(ciphers[0], secKey) ? 1.0 : 0.0};
  if (op == "union") return as_doubles(DecodeUnion(ciphers, secKey));
  if (op == "inter") return as_doubles(DecodeInter(ciphers, secKey));
  if (op == "MLeval") return {DecodeModelEvaluation(ciphers, secKey)};
  return as_doubles(libunlynx::DecryptIntVectorWithNeg(secKey, ciphers));  // default branch (:228-237)
}

}  // namespace libdrynxencoding
