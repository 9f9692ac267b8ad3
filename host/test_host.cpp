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

static void TestLogisticRegressionDevicePrepass() {  // drx_logreg_coefficients == the host's float64 loop, bit for bit
  LogisticRegressionParameters lr;
  lr.NbrFeatures = 3;
  lr.K = 2;
  lr.Means = {0.25, 1.5, -0.75};
  lr.StandardDeviations = {3.0, 7.0, 0.3};
  lr.PrecisionApproxCoefficients = 1000.0;
  std::vector<std::vector<double>> x[2];
  std::vector<int64_t> y[2];
  std::vector<double> flat;
  std::vector<int64_t> yflat;
  for (int p = 0; p < 2; p++)
    for (int r = 0; r < 500 + 37 * p; r++) {
      std::vector<double> rec(3);
      for (auto& v : rec) v = (double)(sm_next() % 4000) / 997.0;
      x[p].push_back(rec);
      y[p].push_back((int64_t)(sm_next() & 1));
      flat.insert(flat.end(), rec.begin(), rec.end());
      yflat.push_back(y[p].back());
    }
  const size_t off[3] = {0, x[0].size(), x[0].size() + x[1].size()};
  std::vector<int64_t> out(2 * 20);
  drynx::check(drx_logreg_coefficients(drynx::Gpu::get().ctx(), 2, off, 3, 2, flat.data(), yflat.data(), lr.Means.data(),
                                       lr.StandardDeviations.data(), lr.PrecisionApproxCoefficients, out.data()));
  for (int p = 0; p < 2; p++) {
    const auto want = LogisticRegressionPlaintext(x[p], y[p], lr);
    EXPECT(want.size() == 20 && std::vector<int64_t>(out.begin() + 20 * p, out.begin() + 20 * (p + 1)) == want);
  }
}
static void TestEncodeDecodeBit() {  // OR_AND_test.go:14-38 and :40-69 (u = 2, l = 1, one server)
  auto keys = drynx::NewKeyPair();
  for (auto in : {std::vector<bool>{false, true, false}, std::vector<bool>{true, true}, std::vector<bool>{false, false}}) {
    const bool expectOr = LocalResultOR(in), expectAnd = LocalResultAND(in);
    EXPECT(DecodeBitOR(EncodeBitOr(expectOr, keys.Public).first, keys.Private) == expectOr);
    EXPECT(DecodeBitAND(EncodeBitAND(expectAnd, keys.Public).first, keys.Private) == expectAnd);
    const int64_t u = 2, l = 1;
    auto col = libdrynxrange::RegisterColumn({libdrynxrange::InitRangeProofSignature(u)}, u);
    auto eo = EncodeBitOrWithProof(expectOr, keys.Public, col, l, u), ea = EncodeBitANDWithProof(expectAnd, keys.Public, col, l, u);
    EXPECT(DecodeBitOR(eo.enc[0], keys.Private) == expectOr);
    EXPECT(DecodeBitAND(ea.enc[0], keys.Private) == expectAnd);
    EXPECT(libdrynxrange::RangeProofVerification(libdrynxrange::CreatePredicateRangeProofForAllServ(eo.proofs[0]), u, l, col, keys.Public));
    EXPECT(libdrynxrange::RangeProofVerification(libdrynxrange::CreatePredicateRangeProofForAllServ(ea.proofs[0]), u, l, col, keys.Public));
  }
}
static void TestEncodeDecodeMinMax() {  // min_max_test.go:14-47 (401 bits) and :49-116 (two servers, u = 2, l = 1)
  auto keys = drynx::NewKeyPair();
  std::vector<int64_t> in = {1, 2, 3, 4, 5, 6, 7, 8, 9, 10, 12, 3, 2, 15, 6, 17, 2, -5, 72, -20, 100, -190, 200};
  const int64_t gmax = 200, gmin = -200;
  EXPECT(DecodeMin(EncodeMin(in, gmax, gmin, keys.Public).enc, gmin, keys.Private) == -190);
  EXPECT(DecodeMax(EncodeMax(in, gmax, gmin, keys.Public).enc, gmin, keys.Private) == 200);
  // the aggregate of two providers' encodings decodes to the global extreme (what the protocol does with them)
  auto a = EncodeMin({5, 9}, 20, 0, keys.Public), b = EncodeMin({7, 3, 11}, 20, 0, keys.Public);
  EXPECT(DecodeMin(libunlynx::Add(a.enc, b.enc), 0, keys.Private) == 3);
  auto c = EncodeMax({5, 9}, 20, 0, keys.Public), d = EncodeMax({7, 3, 11}, 20, 0, keys.Public);
  EXPECT(DecodeMax(libunlynx::Add(c.enc, d.enc), 0, keys.Private) == 11);
  const int64_t u = 2, l = 1, lo = 0, hi = 10;
  std::vector<libdrynxrange::SignatureColumn> sigs;
  Ranges lu;
  for (int64_t i = lo; i <= hi; i++) {
    sigs.push_back(libdrynxrange::RegisterColumn({libdrynxrange::InitRangeProofSignature(u), libdrynxrange::InitRangeProofSignature(u)}, u));
    lu.push_back({u, l});
  }
  auto emin = EncodeMinWithProofs({1, 2, 10}, hi, lo, keys.Public, &sigs, &lu);
  auto emax = EncodeMaxWithProofs({1, 2, 10}, hi, lo, keys.Public, &sigs, &lu);
  EXPECT(DecodeMin(emin.enc, lo, keys.Private) == 1);
  EXPECT(DecodeMax(emax.enc, lo, keys.Private) == 10);
  libdrynxrange::RangeProofList lmin, lmax;
  lmin.Data = libdrynxrange::CreatePredicateRangeProofListForAllServers(emin.proofs);
  lmax.Data = libdrynxrange::CreatePredicateRangeProofListForAllServers(emax.proofs);
  EXPECT(libdrynxrange::RangeProofListVerification(lmin, lu, sigs, keys.Public, 1.0));
  EXPECT(libdrynxrange::RangeProofListVerification(lmax, lu, sigs, keys.Public, 1.0));
}
static void TestEncodeDecodeUnionInter() {  // set_union_intersection_test.go:15-50
  auto keys = drynx::NewKeyPair();
  std::vector<int64_t> in = {1, 2, 3, 1, 5, 1, 2, 3};
  const int64_t min = 0, max = 10;
  std::vector<int64_t> expect((size_t)(max - min + 1), 0);
  for (auto v : Unique(in)) expect[(size_t)(v - min)] = 1;
  EXPECT(Unique(in) == std::vector<int64_t>({1, 2, 3, 5}));
  EXPECT(DecodeUnion(EncodeUnion(in, min, max, keys.Public).enc, keys.Private) == expect);
  EXPECT(DecodeInter(EncodeInter(in, min, max, keys.Public).enc, keys.Private) == expect);
  // two providers: union = either has it, intersection = both have it
  auto u1 = EncodeUnion({1, 2}, min, max, keys.Public), u2 = EncodeUnion({2, 7}, min, max, keys.Public);
  EXPECT(DecodeUnion(libunlynx::Add(u1.enc, u2.enc), keys.Private) == std::vector<int64_t>({0, 1, 1, 0, 0, 0, 0, 1, 0, 0, 0}));
  auto i1 = EncodeInter({1, 2}, min, max, keys.Public), i2 = EncodeInter({2, 7}, min, max, keys.Public);
  EXPECT(DecodeInter(libunlynx::Add(i1.enc, i2.enc), keys.Private) == std::vector<int64_t>({0, 0, 1, 0, 0, 0, 0, 0, 0, 0, 0}));
}
static void TestEncodeDecodeCosimAndModelEvaluation() {  // cosim_test.go:17-55, model_evaluation_test.go:14-50
  auto keys = drynx::NewKeyPair();
  libunlynx::CreateDecryptionTable(10000, keys.Public, keys.Private);
  std::vector<int64_t> a = {0, 1, 2, 3, 4, 5, 6, 7, 8, 9, 12};
  auto e = EncodeCosim(a, a, keys.Public);
  EXPECT(e.clear == std::vector<int64_t>({57, 57, 429, 429, 429}));
  EXPECT(DecodeCosim(e.enc, keys.Private) == 429.0 / (std::sqrt(429.0) * std::sqrt(429.0)));
  std::vector<int64_t> y = {1, 2, 3, 4, 5, 6, 7, 8, 1, 67, -72}, preds = {32, 12, 23, 4, 13, -72, 12, 8, 23, 67, 2};
  int64_t sy = 0, sy2 = 0, sd2 = 0;
  for (size_t i = 0; i < y.size(); i++) { sy += y[i]; sy2 += y[i] * y[i]; sd2 += (preds[i] - y[i]) * (preds[i] - y[i]); }
  const double B = (double)sy2 - (double)(sy * sy / (int64_t)y.size());
  auto m = EncodeModelEvaluation(y, preds, keys.Public);
  EXPECT(m.clear == std::vector<int64_t>({0}));
  EXPECT(DecodeModelEvaluation(m.enc, keys.Private) == 1.0 - (double)sd2 / B);
}

static void TestEncodeDecodeDispatcher() {  // encode_decode.go Encode / Decode over every operation name
  auto keys = drynx::NewKeyPair();
  libunlynx::CreateDecryptionTable(10000, keys.Public, keys.Private);
  std::vector<std::vector<int64_t>> one = {{3, 1, 4, 1, 5, 9, 2, 6}};
  auto run = [&](const char* name, const std::vector<std::vector<int64_t>>& datas, int64_t qmin = 0, int64_t qmax = 0) {
    Operation op;
    op.NameOp = name;
    op.QueryMin = qmin;
    op.QueryMax = qmax;
    return Decode(Encode(datas, keys.Public, nullptr, nullptr, op).enc, keys.Private, op);
  };
  EXPECT(run("sum", one) == std::vector<double>({31}));
  EXPECT(run("mean", one) == std::vector<double>({31.0 / 8}));
  EXPECT(run("variance", one) == std::vector<double>({173.0 / 8 - (31.0 / 8) * (31.0 / 8)}));
  EXPECT(run("min", one, 0, 12) == std::vector<double>({1}));
  EXPECT(run("max", one, 0, 12) == std::vector<double>({9}));
  EXPECT(run("frequencyCount", one, 0, 9) == std::vector<double>({0, 2, 1, 1, 1, 1, 1, 0, 0, 1}));
  EXPECT(run("union", one, 0, 9) == std::vector<double>({0, 1, 1, 1, 1, 1, 1, 0, 0, 1}));
  EXPECT(run("inter", one, 0, 9) == std::vector<double>({0, 1, 1, 1, 1, 1, 1, 0, 0, 1}));
  EXPECT(run("bool_OR", {{1}}) == std::vector<double>({1}) && run("bool_OR", {{0}}) == std::vector<double>({0}));
  EXPECT(run("bool_AND", {{1}}) == std::vector<double>({1}) && run("bool_AND", {{0}}) == std::vector<double>({0}));
  EXPECT(run("cosim", {{1, 2, 3}, {1, 2, 3}}) == std::vector<double>({14.0 / (std::sqrt(14.0) * std::sqrt(14.0))}));
  EXPECT(run("MLeval", {{1, 2, 3, 4}, {1, 2, 3, 4}}) == std::vector<double>({1.0}));
  // lin_reg takes columns: two feature columns then y = 1 + 2 x1 + 4 x2
  auto c = run("lin_reg", {{1, 0, 1, 2, 3}, {2, 1, 0, 1, 5}, {11, 5, 3, 9, 27}});
  EXPECT(c.size() == 3 && fabs(c[0] - 1) < 1e-9 && fabs(c[1] - 2) < 1e-9 && fabs(c[2] - 4) < 1e-9);
  // with proofs through the dispatcher: mean = [sum, N], u = 4, l = 3 (values below 64)
  const int64_t u = 4, l = 3;
  std::vector<libdrynxrange::SignatureColumn> sigs;
  Ranges lu;
  for (int i = 0; i < 2; i++) {
    sigs.push_back(libdrynxrange::RegisterColumn({libdrynxrange::InitRangeProofSignature(u)}, u));
    lu.push_back({u, l});
  }
  Operation op;
  op.NameOp = "mean";
  auto e = Encode(one, keys.Public, &sigs, &lu, op);
  EXPECT(e.proofs.size() == 2 && e.clear == std::vector<int64_t>({31, 8}));
  libdrynxrange::RangeProofList list;
  list.Data = libdrynxrange::CreatePredicateRangeProofListForAllServers(e.proofs);
  EXPECT(libdrynxrange::RangeProofListVerification(list, lu, sigs, keys.Public, 1.0));
}

// The UnLynx key-switching protocol as Drynx runs it (services/service.go:565-616, 853-862; data/data.go:73-83):
// the aggregate is encrypted under the collective key of three servers; every server contributes
// KeySwitchSequence(querierKey, rBs, itsSecret); the root adds the contributions to (0, C); the querier decrypts.
static void TestKeySwitching() {
  std::vector<drynx::KeyPair> servers = {drynx::NewKeyPair(), drynx::NewKeyPair(), drynx::NewKeyPair()};
  // collective key = sum of the servers' public keys (summed as the two halves of ciphertext-shaped values)
  std::vector<libunlynx::CipherVector> shaped;
  for (auto& k : servers) {
    libunlynx::CipherText c;
    memcpy(c.b, k.Public.b, 64);
    memcpy(c.b + 64, k.Public.b, 64);
    shaped.push_back({c});
  }
  drynx::Point collective;
  memcpy(collective.b, libunlynx::AddAll(shaped)[0].b, 64);
  auto querier = drynx::NewKeyPair();
  libunlynx::CreateDecryptionTable(10000, querier.Public, querier.Private);
  std::vector<int64_t> values = {1, 2, 0, -3, 9999};
  auto cv = libunlynx::EncryptIntVector(collective, values);
  std::vector<drynx::Point> rBs(cv.size());
  for (size_t i = 0; i < cv.size(); i++) memcpy(rBs[i].b, cv[i].b, 64);
  std::vector<libunlynx::CipherVector> contributions;
  for (auto& k : servers) {
    auto seq = libunlynxkeyswitch::KeySwitchSequence(querier.Public, rBs, k.Private);
    EXPECT(seq.cv.size() == values.size() && seq.ks2s.size() == values.size() && seq.rBNegs.size() == values.size() &&
           seq.vis.size() == values.size());
    EXPECT(!memcmp(seq.ks2s[1].b, seq.cv[1].b + 64, 64));
    // rBNeg + rB = identity: adding them as ciphertext halves gives 128 zero bytes
    libunlynx::CipherText a, b;
    memcpy(a.b, rBs[0].b, 64); memcpy(a.b + 64, rBs[0].b, 64);
    memcpy(b.b, seq.rBNegs[0].b, 64); memcpy(b.b + 64, seq.rBNegs[0].b, 64);
    uint8_t zero[128] = {0};
    EXPECT(!memcmp(libunlynx::Add({a}, {b})[0].b, zero, 128));
    contributions.push_back(seq.cv);
  }
  auto switched = libunlynxkeyswitch::Combine(cv, contributions);
  EXPECT(libunlynx::DecryptIntVectorWithNeg(querier.Private, switched) == values);
  // not decryptable by a single server, nor by the querier before switching
  EXPECT(libunlynx::DecryptIntVectorWithNeg(querier.Private, cv) != values);
}

static void TestObfuscationProofVerification() {  // lib/obfuscation/obfuscation_proof_test.go:11-34
  auto keys = drynx::NewKeyPair();
  auto ev = libunlynx::EncryptInt(keys.Public, 2);
  const drynx::Scalar obfuscationFactor = keys.Private;
  auto evo = libunlynx::MulCipherTextbyScalar(ev, obfuscationFactor);
  auto op = libdrynxobfuscation::ObfuscationProofCreation(ev, evo, obfuscationFactor);
  EXPECT(op.Proof.size() == 160);
  EXPECT(libdrynxobfuscation::ObfuscationProofVerification(op));
  auto nop = op;  // the ToBytes / FromBytes round trip of the Go test: the C++ structs already hold the byte forms
  libdrynxobfuscation::PublishedListObfuscationProof lop;
  lop.Prs = {op, nop};
  EXPECT(libdrynxobfuscation::ObfuscationListProofVerification(lop, 1));
  auto wrong = libdrynxobfuscation::ObfuscationProofCreation(ev, evo, drynx::scalar_from_u64(1));
  EXPECT(!libdrynxobfuscation::ObfuscationProofVerification(wrong));
  // a longer list, as protocols/obfuscation_protocol.go:231-259 builds it; the threshold reads only a prefix
  std::vector<int64_t> vals = {0, 1, 0, 7, 0, 3};
  auto cv = libunlynx::EncryptIntVector(keys.Public, vals);
  std::vector<drynx::Scalar> ss(cv.size());
  for (auto& x : ss) x = drynx::pick();
  auto co = libunlynx::MulCipherVectorByScalars(cv, ss);
  auto list = libdrynxobfuscation::ObfuscationListProofCreation(cv, co, ss);
  EXPECT(libdrynxobfuscation::ObfuscationListProofVerification(list, 1.0));
  list.Prs[4].Proof[150] ^= 1;
  EXPECT(!libdrynxobfuscation::ObfuscationListProofVerification(list, 1.0));
  EXPECT(libdrynxobfuscation::ObfuscationListProofVerification(list, 0.5));
  auto bits = libdrynxobfuscation::VerifyEach(list, list.Prs.size());
  EXPECT(bits == std::vector<uint8_t>({1, 1, 1, 1, 0, 1}));
  std::swap(list.Prs[0].Co, list.Prs[1].Co);
  EXPECT(!libdrynxobfuscation::ObfuscationListProofVerification(list, 0.2));
}

static void TestKeySwitchListProof() {  // the proof services/service.go:581 and data/data.go:81 create
  auto server = drynx::NewKeyPair();
  auto querier = drynx::NewKeyPair();
  auto coll = drynx::NewKeyPair();
  auto cv = libunlynx::EncryptIntVector(coll.Public, {5, 0, -2, 77});
  std::vector<drynx::Point> rBs(cv.size());
  for (size_t i = 0; i < cv.size(); i++) memcpy(rBs[i].b, cv[i].b, 64);
  auto seq = libunlynxkeyswitch::KeySwitchSequence(querier.Public, rBs, server.Private);
  auto proof = libunlynxkeyswitch::KeySwitchListProofCreation(server.Public, querier.Public, server.Private, seq.ks2s, seq.rBNegs, seq.vis);
  EXPECT(proof.List.size() == cv.size() && proof.List[0].Proof.size() == 256);
  for (size_t i = 0; i < cv.size(); i++) EXPECT(!memcmp(proof.List[i].ViB.b, seq.cv[i].b, 64));  // viB is the K component sent on
  EXPECT(libunlynxkeyswitch::KeySwitchListProofVerification(proof, 1.0));
  auto t = proof;
  t.List[2].Ks2 = t.List[1].Ks2;
  EXPECT(!libunlynxkeyswitch::KeySwitchListProofVerification(t, 1.0));
  EXPECT(libunlynxkeyswitch::KeySwitchListProofVerification(t, 0.5));
  t = proof;
  t.List[0].K = querier.Public;  // somebody else's key
  EXPECT(!libunlynxkeyswitch::KeySwitchListProofVerification(t, 0.1));
  t = proof;
  t.List[3].Proof[200] ^= 1;
  EXPECT(libunlynxkeyswitch::KeySwitchVerifyEach(t, 4) == std::vector<uint8_t>({1, 1, 1, 0}));
  // a proof made with the wrong secret key does not verify
  auto other = drynx::NewKeyPair();
  auto forged = libunlynxkeyswitch::KeySwitchListProofCreation(server.Public, querier.Public, other.Private, seq.ks2s, seq.rBNegs, seq.vis);
  EXPECT(!libunlynxkeyswitch::KeySwitchListProofVerification(forged, 1.0));
}

static void TestShuffleSequence() {  // data/data.go:60-70: two lines shuffled under the collective key
  auto keys = drynx::NewKeyPair();
  libunlynx::CreateDecryptionTable(10000, keys.Public, keys.Private);
  std::vector<std::vector<int64_t>> tabs = {{1, 2, 3, 6}, {2, 4, 8, 6}, {0, -5, 7, 7}};
  std::vector<libunlynx::CipherVector> responses;
  for (auto& t : tabs) responses.push_back(libunlynx::EncryptIntVector(keys.Public, t));
  auto sh = libunlynxshuffle::ShuffleSequence(responses, keys.Public);
  EXPECT(sh.outputList.size() == 3 && sh.pi.size() == 3 && sh.beta.size() == 3 && sh.beta[0].size() == 4);
  std::vector<int> seen(3, 0);
  for (size_t i = 0; i < 3; i++) {
    EXPECT(sh.pi[i] >= 0 && sh.pi[i] < 3);
    seen[(size_t)sh.pi[i]]++;
    EXPECT(libunlynx::DecryptIntVectorWithNeg(keys.Private, sh.outputList[i]) == tabs[(size_t)sh.pi[i]]);
    EXPECT(memcmp(sh.outputList[i][0].b, responses[(size_t)sh.pi[i]][0].b, 128) != 0);
  }
  EXPECT(seen == std::vector<int>({1, 1, 1}));
}

static std::string hex(const uint8_t* b, size_t n) {
  static const char* d = "0123456789abcdef";
  std::string s;
  for (size_t i = 0; i < n; i++) { s += d[b[i] >> 4]; s += d[b[i] & 15]; }
  return s;
}
static void TestSha512() {  // FIPS 180-4 "abc" vector and a two-block message, split across the three inputs
  uint8_t out[64];
  drynxproof::sha512((const uint8_t*)"a", 1, (const uint8_t*)"b", 1, (const uint8_t*)"c", 1, out);
  EXPECT(hex(out, 64) == "ddaf35a193617abacc417349ae20413112e6fa4e89a97ea20a9eeee64b55d39a2192992a274fc1a836ba3c23a3feebbd454d4423643ce80e2a9ac94fa54ca49f");
  std::string a(200, 'a');
  drynxproof::sha512((const uint8_t*)a.data(), 64, (const uint8_t*)a.data() + 64, 64, (const uint8_t*)a.data() + 128, 72, out);
  EXPECT(hex(out, 64) == "4b11459c33f52a22ee8236782714c150a3b2c60994e9acee17fe68947a3e6789f31e7668394592da7bef827cddca88c4e6f86e4df7ed1ae6cba71f3e98faee9f");
}
// VerifyProofSignature (lib/proof/structs_proofs.go:498-505) on a signature made by the big-integer definition
// (oracle/bn256_ref.py schnorr_sign, x = 12, nonce 34), then on tampered copies; and the aggregation proof
// (data/data.go:41-47: Data = the vector, AggregationResult = its Acum()).
static void TestProofSignatureAndAggregationProof() {
  static const uint8_t pub[64] = {0x24, 0x0c, 0x76, 0x0d, 0x59, 0xa5, 0xe7, 0x9d, 0x0d, 0xd4, 0xde, 0xb2, 0x00, 0x02, 0xeb, 0x9d, 0xe3, 0x9d, 0xfe, 0x7f, 0xcc, 0xaa, 0xd7, 0xc7, 0x1c, 0x4b, 0xcc, 0x0b, 0x42, 0xa2, 0x5a, 0x6c, 0x86, 0xad, 0x33, 0x99, 0x8d, 0x8f, 0x86, 0x27, 0x3b, 0xbe, 0xd0, 0x70, 0x6f, 0xb1, 0xb5, 0xe3, 0x13, 0x55, 0x9d, 0x3f, 0xbc, 0x54, 0xf7, 0xf4, 0x41, 0xea, 0x54, 0xf6, 0x56, 0x89, 0xe1, 0x8b};
  static const uint8_t sig[96] = {0x75, 0xe3, 0xc7, 0x4b, 0x76, 0xa3, 0x4d, 0x8f, 0xa0, 0x1e, 0xc5, 0x25, 0x37, 0x09, 0x81, 0x8b, 0x5e, 0x08, 0xce, 0xc1, 0x14, 0x42, 0xdd, 0x08, 0x24, 0xb5, 0xf1, 0xb2, 0xfb, 0x40, 0x32, 0xbd, 0x6e, 0x3a, 0x60, 0x98, 0x4d, 0x16, 0xf9, 0xd0, 0xf2, 0x6c, 0x88, 0x09, 0xa5, 0x42, 0xc7, 0x7d, 0xcd, 0x44, 0x10, 0xde, 0x9a, 0xa0, 0xda, 0x90, 0x5b, 0x83, 0xba, 0x78, 0x5a, 0x94, 0x02, 0x53, 0x1a, 0xdc, 0xbd, 0xa0, 0x82, 0xdb, 0x73, 0x26, 0x7e, 0x62, 0x76, 0x9b, 0x2f, 0x5c, 0xbe, 0x33, 0x81, 0x58, 0xb5, 0xe9, 0x46, 0x84, 0x86, 0xf5, 0x3e, 0x9c, 0x91, 0xad, 0x15, 0x10, 0xb2, 0xb9};
  const char* msg = "drynx proof request";
  drynx::Point P;
  memcpy(P.b, pub, 64);
  EXPECT(memcmp(drynx::point_mul(drynx::scalar_from_u64(12), drynx::base_point()).b, pub, 64) == 0);
  EXPECT(drynxproof::VerifyProofSignature(P, (const uint8_t*)msg, strlen(msg), sig));
  EXPECT(!drynxproof::VerifyProofSignature(P, (const uint8_t*)msg, strlen(msg) - 1, sig));
  uint8_t bad[96];
  memcpy(bad, sig, 96);
  bad[95] ^= 1;
  EXPECT(!drynxproof::VerifyProofSignature(P, (const uint8_t*)msg, strlen(msg), bad));
  memcpy(bad, sig, 96);
  memset(bad + 64, 0xff, 32);  // s >= n
  EXPECT(!drynxproof::VerifyProofSignature(P, (const uint8_t*)msg, strlen(msg), bad));
  // aggregation proof
  auto keys = drynx::NewKeyPair();
  auto ev = libunlynx::EncryptIntVector(keys.Public, {1, 2, 3, 4, 5});
  std::vector<libunlynx::CipherVector> cols;
  for (auto& c : ev) cols.push_back({c});
  auto acum = libunlynx::AddAll(cols)[0];
  auto list = libunlynxaggr::AggregationListProofCreation({ev, ev}, {acum, acum});
  EXPECT(libunlynxaggr::AggregationListProofVerification(list, 1.0));
  list.List[1].AggregationResult.b[100] ^= 1;
  EXPECT(!libunlynxaggr::AggregationListProofVerification(list, 1.0));
  EXPECT(libunlynxaggr::AggregationListProofVerification(list, 0.5));  // only the first proof is looked at
  // obfuscation: an encrypted zero stays zero, a one stays non-zero (OR_AND.go decoding)
  libunlynx::CreateDecryptionTable(10000, keys.Public, keys.Private);
  auto bits = libunlynx::EncryptIntVector(keys.Public, {0, 1});
  auto obf = libunlynx::MulCipherVectorByScalars(bits, {drynx::pick(), drynx::pick()});
  EXPECT(libunlynx::DecryptCheckZeroVector(keys.Private, obf) == std::vector<int64_t>({0, 1}));
  EXPECT(memcmp(obf[1].b, bits[1].b, 128) != 0);
}

int main(int argc, char** argv) {
  const bool cpu_only = argc > 1 && !strcmp(argv[1], "--cpu");
  TestComputeApproxCoefficients();
  TestCartesianProduct();
  TestShapesAndToBase();
  TestSha512();
  if (!cpu_only) {
    try {
      drynx::Gpu::get().random = random_scalar;
      TestEncodeDecodeSum();
      TestEncodeDecodeSumWithProofs();
      TestEncodeDecodeMeanVariance();
      TestEncodeDecodeFrequencyCountWithProofs();
      TestEncodeDecodeLinearRegressionDims();
      TestLogisticRegressionAggregation();
      TestLogisticRegressionDevicePrepass();
      TestEncodeDecodeBit();
      TestEncodeDecodeMinMax();
      TestEncodeDecodeUnionInter();
      TestEncodeDecodeCosimAndModelEvaluation();
      TestEncodeDecodeDispatcher();
      TestKeySwitching();
      TestProofSignatureAndAggregationProof();
      TestShuffleSequence();
      TestObfuscationProofVerification();
      TestKeySwitchListProof();
    } catch (const drynx::Error& e) {
      printf("FAIL: exception %d: %s\n", e.code, e.what());
      failures++;
    }
  }
  printf(failures ? "HOST_TESTS_FAILED %d\n" : "HOST_TESTS_OK\n", failures);
  return failures ? 1 : 0;
}
