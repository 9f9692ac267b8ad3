// Reference-style tests of the C++ host mirror (host/drynx_host.hpp) against libdrynx_gpu.so.
// Each test cites the Go test it follows. `--cpu` runs only the plaintext tests (no GPU needed).
#include <cstdio>
#include <cstdlib>

#include "drynx_host.hpp"

using namespace libdrynxencoding;
static int failures = 0;
#define EXPECT(cond)                                                     \
  do {                                                                   \
    if (!(cond)) {                                                       \
      printf("FAIL %s:%d: %s\n", __FILE__, __LINE__, #cond);             \
      failures++;                                                        \
    }                                                                    \
  } while (0)

// SplitMix64 + kyber-style rejection sampling below the group order
static uint64_t sm_state = 2026;
static uint64_t sm_next() {
  uint64_t z = (sm_state += 0x9e3779b97f4a7c15ULL);
  z = (z ^ (z >> 30)) * 0xbf58476d1ce4e5b9ULL;
  z = (z ^ (z >> 27)) * 0x94d049bb133111ebULL;
  return z ^ (z >> 31);
}
static void random_scalar(uint8_t out[32]) {
  static const uint8_t n[32] = {0x8f, 0xb5, 0x01, 0xe3, 0x4a, 0xa3, 0x87, 0xf9, 0xaa, 0x6f, 0xec, 0xb8, 0x61, 0x84, 0xdc, 0x21,
                                0x2e, 0x8d, 0x8e, 0x12, 0xf8, 0x2b, 0x39, 0x24, 0x1a, 0x2e, 0xf4, 0x5b, 0x57, 0xac, 0x72, 0x61};
  for (;;) {
    for (int w = 0; w < 4; w++) {
      uint64_t v = sm_next();
      for (int b = 0; b < 8; b++) out[8 * w + b] = (uint8_t)(v >> (56 - 8 * b));
    }
    if (memcmp(out, n, 32) < 0) return;
  }
}

// ---------------------------------------------------------------- plaintext tests (no GPU)
static void TestComputeApproxCoefficients() {  // logistic_regression_test.go:20-72
  const std::vector<double> X = {0, 1, 2, 3, 4};
  using VV = std::vector<std::vector<double>>;
  EXPECT(ComputeDistinctApproxCoefficients(X, 1, 1) == VV({{0, 1, 2, 3, 4}}));
  EXPECT(ComputeDistinctApproxCoefficients(X, 0, 1) == VV({{0, -1, -2, -3, -4}}));
  EXPECT(ComputeDistinctApproxCoefficients(X, 1, 2) == VV({{0, 1, 2, 3, 4}, {0, 0, 0, 0, 0, -1, -2, -3, -4, -4, -6, -8, -9, -12, -16}}));
  EXPECT(ComputeDistinctApproxCoefficients(X, 0, 2) == VV({{0, -1, -2, -3, -4}, {0, 0, 0, 0, 0, 1, 2, 3, 4, 4, 6, 8, 9, 12, 16}}));
  const std::vector<double> k3 = {0, 0, 0, 0, 0, 0, 0, 0, 0, 0, 0, 0, 0, 0, 0, -1, -2, -3, -4, -4, -6, -8, -9, -12, -16, -8, -12, -16, -18, -24, -32, -27, -36, -48, -64};
  EXPECT(ComputeDistinctApproxCoefficients(X, 1, 3) == VV({{0, 1, 2, 3, 4}, {0, 0, 0, 0, 0, -1, -2, -3, -4, -4, -6, -8, -9, -12, -16}, k3}));
  EXPECT(ComputeDistinctApproxCoefficients(X, 0, 3) == VV({{0, -1, -2, -3, -4}, {0, 0, 0, 0, 0, 1, 2, 3, 4, 4, 6, 8, 9, 12, 16}, k3}));
}
static void TestCartesianProduct() {  // logistic_regression_test.go:544-558
  using VV = std::vector<std::vector<int64_t>>;
  EXPECT(CartesianProduct(0, 2, 3) == VV({{0, 0, 0}, {0, 0, 1}, {0, 1, 0}, {0, 1, 1}, {1, 0, 0}, {1, 0, 1}, {1, 1, 0}, {1, 1, 1}}));
  EXPECT(CartesianProduct(0, 3, 2) == VV({{0, 0}, {0, 1}, {0, 2}, {1, 0}, {1, 1}, {1, 2}, {2, 0}, {2, 1}, {2, 2}}));
}
static void TestShapesAndToBase() {
  EXPECT(libdrynxrange::ToBase(11, 4, 2) == std::vector<int64_t>({3, 2}));
  EXPECT(libdrynxrange::ToBase(0, 16, 4) == std::vector<int64_t>({0, 0, 0, 0}));
  EXPECT(libdrynxrange::ToBase(-5, 16, 2) == std::vector<int64_t>({0, 0}));
  EXPECT(libdrynxrange::ToBase(300, 16, 1) == std::vector<int64_t>({12, 2, 1}));
  // lib/structs.go:629: (d^2 + 5d + 4) / 2 values for lin_reg
  std::vector<std::vector<int64_t>> X = {{1, 2}, {0, 1}, {1, 0}, {2, 1}, {3, 5}};
  std::vector<int64_t> y = {11, 5, 3, 9, 27};
  auto pt = LinearRegressionDimsPlaintext(X, y);
  EXPECT(pt.size() == (2 * 2 + 5 * 2 + 4) / 2);
  EXPECT(pt == std::vector<int64_t>({5, 7, 9, 15, 19, 31, 55, 113, 171}));
  // logistic regression: (d+1) + (d+1)^2 = 90 coefficients for d = 8, k = 2 (logistic_regression.go:306-313)
  EXPECT(getTotalNumberApproxCoefficients(8, 2) == 90);
  LogisticRegressionParameters lr;
  lr.NbrFeatures = 2;
  lr.K = 2;
  lr.Means = {0, 0};
  lr.StandardDeviations = {1, 1};
  lr.PrecisionApproxCoefficients = 1;
  auto packed = LogisticRegressionPlaintext({{1, 2}, {3, 4}}, {1, 0}, lr);
  // records [1,1,2] (y=1) and [1,3,4] (y=0): level 1 = x(2y-1) summed; level 2 = -(x_a x_b) summed (ypart = -1 for both labels)
  EXPECT(packed == std::vector<int64_t>({0, -2, -2, -2, -4, -6, -4, -10, -14, -6, -14, -20}));
  // All vs Distinct agree on the coefficients they share (first level)
  EXPECT(ComputeAllApproxCoefficients({1, 2, 3}, 1, 2)[0] == ComputeDistinctApproxCoefficients({1, 2, 3}, 1, 2)[0]);
}

// ---------------------------------------------------------------- GPU tests
static void TestEncodeDecodeSum() {  // sum_test.go:15-31
  auto keys = drynx::NewKeyPair();
  auto e = EncodeSum({0, 1, 2, 3, 4, 5, 6, 7, 8, 9, -12000}, keys.Public);
  EXPECT(DecodeSum(e.enc[0], keys.Private) == -11955);
  EXPECT(e.clear == std::vector<int64_t>({-11955}));
}
static void TestEncodeDecodeSumWithProofs() {  // sum_test.go:34-57: u = 2, l = 6, one server
  auto keys = drynx::NewKeyPair();
  const int64_t u = 2, l = 6;
  auto col = libdrynxrange::RegisterColumn({libdrynxrange::InitRangeProofSignature(u)}, u);
  std::vector<libdrynxrange::SignatureColumn> sigs = {col};
  Ranges lu = {{u, l}};
  auto e = EncodeSumWithProofs({0, 10, 9}, keys.Public, &sigs, &lu);
  EXPECT(DecodeSum(e.enc[0], keys.Private) == 19);
  auto rp = libdrynxrange::CreatePredicateRangeProofForAllServ(e.proofs[0]);
  EXPECT(libdrynxrange::RangeProofVerification(rp, u, l, col, keys.Public));
  // and what the reference never checks: a sum outside [0, u^l) must not verify
  auto e2 = EncodeSumWithProofs({60, 10}, keys.Public, &sigs, &lu);
  EXPECT(!libdrynxrange::RangeProofVerification(libdrynxrange::CreatePredicateRangeProofForAllServ(e2.proofs[0]), u, l, col, keys.Public));
}
static void TestEncodeDecodeMeanVariance() {  // mean_test.go:15-32, variance_test.go:16-46
  auto keys = drynx::NewKeyPair();
  std::vector<int64_t> in = {0, 1, 2, 3, 4, 5, 6, 7, 8, 9, -12000};
  double sum = 0, sq = 0;
  for (auto v : in) { sum += v; sq += (double)v * v; }
  libunlynx::CreateDecryptionTable(10000, keys.Public, keys.Private);
  EXPECT(DecodeMean(EncodeMean(in, keys.Public).enc, keys.Private) == sum / in.size());
  std::vector<int64_t> small = {1, 2, 3, 4, 5, 6, 7, 8, 9, 10, -120};
  sum = sq = 0;
  for (auto v : small) { sum += v; sq += (double)v * v; }
  const double mean = sum / small.size();
  EXPECT(DecodeVariance(EncodeVariance(small, keys.Public).enc, keys.Private) == sq / small.size() - mean * mean);
}
static void TestEncodeDecodeFrequencyCountWithProofs() {  // frequency_count_test.go:43-100: 2 servers, u = 2, l = 10
  auto keys = drynx::NewKeyPair();
  libunlynx::CreateDecryptionTable(10000, keys.Public, keys.Private);
  const int64_t min = -1, max = 12, u = 2, l = 10;
  std::vector<int64_t> in = {-1, 0, 1, 2, 3, 4, 6, 7, 8, 9, 10, 12, 3, 2};
  std::vector<int64_t> expect((size_t)(max - min + 1), 0);
  for (auto v : in) expect[(size_t)(v - min)]++;
  std::vector<libdrynxrange::SignatureColumn> sigs;
  Ranges lu;
  for (size_t i = 0; i < expect.size(); i++) {
    sigs.push_back(libdrynxrange::RegisterColumn({libdrynxrange::InitRangeProofSignature(u), libdrynxrange::InitRangeProofSignature(u)}, u));
    lu.push_back({u, l});
  }
  auto e = EncodeFreqCountWithProofs(in, min, max, keys.Public, &sigs, &lu);
  EXPECT(DecodeFreqCount(e.enc, keys.Private) == expect);
  libdrynxrange::RangeProofList list;
  list.Data = libdrynxrange::CreatePredicateRangeProofListForAllServers(e.proofs);
  EXPECT(libdrynxrange::RangeProofListVerification(list, lu, sigs, keys.Public, 1.0));
  list.Data[3].bytes[170] ^= 1;  // tamper with Zr of one proof
  EXPECT(!libdrynxrange::RangeProofListVerification(list, lu, sigs, keys.Public, 1.0));
  EXPECT(libdrynxrange::RangeProofListVerification(list, lu, sigs, keys.Public, 0.2));  // threshold: only the first 3 are checked
}
static void TestEncodeDecodeLinearRegressionDims() {  // linear_regression_dims_test.go:15-90: c = (1, 2, 4)
  auto keys = drynx::NewKeyPair();
  std::vector<std::vector<int64_t>> X = {{1, 2}, {0, 1}, {1, 0}, {2, 1}, {3, 5}};
  std::vector<int64_t> y = {11, 5, 3, 9, 27};
  auto c = DecodeLinearRegressionDims(EncodeLinearRegressionDims(X, y, keys.Public).enc, keys.Private);
  EXPECT(c.size() == 3 && fabs(c[0] - 1) < 1e-9 && fabs(c[1] - 2) < 1e-9 && fabs(c[2] - 4) < 1e-9);
}
static void TestLogisticRegressionAggregation() {  // logistic_regression_test.go:74-108: encrypted sums of coefficients
  auto keys = drynx::NewKeyPair();
  LogisticRegressionParameters lr;
  lr.NbrFeatures = 2;
  lr.K = 2;
  lr.Means = {0, 0};
  lr.StandardDeviations = {1, 1};
  std::vector<std::vector<double>> x1 = {{1, 2}, {3, 4}}, x2 = {{2, 2}, {0, 1}, {5, 1}};
  std::vector<int64_t> y1 = {1, 0}, y2 = {1, 1, 0};
  auto e1 = EncodeLogisticRegression(x1, y1, lr, keys.Public), e2 = EncodeLogisticRegression(x2, y2, lr, keys.Public);
  auto agg = libunlynx::AddAll({e1.enc, e2.enc});
  auto got = libunlynx::DecryptIntVectorWithNeg(keys.Private, agg);
  std::vector<int64_t> want(e1.clear.size());
  for (size_t i = 0; i < want.size(); i++) want[i] = e1.clear[i] + e2.clear[i];
  EXPECT(got == want);
  EXPECT(libunlynx::DecryptCheckZero(keys.Private, libunlynx::EncryptInt(keys.Public, 0)) == 0);
  EXPECT(libunlynx::DecryptCheckZero(keys.Private, libunlynx::EncryptInt(keys.Public, 7)) == 1);
}

int main(int argc, char** argv) {
  const bool cpu_only = argc > 1 && !strcmp(argv[1], "--cpu");
  TestComputeApproxCoefficients();
  TestCartesianProduct();
  TestShapesAndToBase();
  if (!cpu_only) {
    try {
      drynx::Gpu::get().random = random_scalar;
      TestEncodeDecodeSum();
      TestEncodeDecodeSumWithProofs();
      TestEncodeDecodeMeanVariance();
      TestEncodeDecodeFrequencyCountWithProofs();
      TestEncodeDecodeLinearRegressionDims();
      TestLogisticRegressionAggregation();
    } catch (const drynx::Error& e) {
      printf("FAIL: exception %d: %s\n", e.code, e.what());
      failures++;
    }
  }
  printf(failures ? "HOST_TESTS_FAILED %d\n" : "HOST_TESTS_OK\n", failures);
  return failures ? 1 : 0;
}
