// Host-side launch wrappers implemented in kernels_g1.cu / kernels_rp.cu and called from capi.cu.
// Plain pointers only; every function enqueues on `st` and returns the number of kernels it launched.
#pragma once
#include <cuda_runtime.h>
#include <stddef.h>
#include <stdint.h>

namespace drx {

constexpr int kCombWindows = 32;
constexpr int kCombEntries = 255;
constexpr size_t kCombPoints = (size_t)kCombWindows * kCombEntries;  // 8160
constexpr size_t kG1AffBytes = 64, kG2AffBytes = 128, kGtBytes = 384, kG1JacBytes = 96;
constexpr int64_t kMaxHomomorphicInt = 100000;  // libunlynx.MaxHomomorphicInt
constexpr uint32_t kDlogCap = 1u << 18;         // open-addressing slots for 100001 keys
constexpr int kAggMaxSlices = 512;  // stage-1 slices of an aggregation (sized from the SM count, see kernels_g1.cu)
constexpr size_t kEncryptLanesMaxItems = 16384; // encryptions up to this many use eight lanes each (latency)
constexpr size_t kCreateLanesMaxItems = 800;   // creation items up to this many: lane-cooperative kernel (8 / 16 lanes per item)
constexpr size_t kGtLanesCreateMaxItems = 12000; // creation items up to this many: GT factors on lane groups (measured, 12 items per proof: 100 proofs 0.44 against 0.76 ms, 500 proofs 1.05 against 1.24, 2,000 proofs 3.49 against 3.05)
constexpr size_t kGtLanesVerifyMinItems = 12000; // verification items from this many on: GT factor on lane groups (below, its shared memory and warps get in the way of the pairing kernel: 500 proofs 4.91 against 4.57 ms; 2,000 proofs 15.1 against 15.7)
constexpr size_t kCreateFineMaxItems = 36000; // creation items up to this many are split twelve ways (latency)
constexpr size_t kUcodeG8MaxItems = 1000;      // pairing checks up to this many use eight program lanes per check instead of four (measured, pairing kernel: 480 checks 2.00 against 2.56 ms, 1,200 checks 2.61 against 2.60)

// ---- kernels_g1.cu
// wire bytes -> Montgomery affine (returns malformed count in *d_bad)
int launch_g1_from_wire(cudaStream_t st, size_t n, const uint8_t* d_wire, void* d_aff, int* d_bad);
int launch_comb_table_g1(cudaStream_t st, int nbases, const void* d_bases_aff, void* d_tables);
constexpr int kFreqCountMaxBins = 12000;
int launch_freq_count(cudaStream_t st, size_t ndp, size_t nrows, const int64_t* d_rows, int64_t qmin, int nbins,
                      int64_t* d_out, int* d_bad);
constexpr int kLogRegMaxFeatures = 31;
int launch_logreg_coeffs(cudaStream_t st, size_t ndp, size_t nrec, const size_t* d_rec_off, int d, int k, const double* d_x,
                         const int64_t* d_y, const double* d_means, const double* d_sds, double precision, double* d_xs,
                         int64_t* d_out);
int launch_encrypt_scalar(cudaStream_t st, size_t n, const uint8_t* d_m, const uint8_t* d_r, const void* d_tabB,
                          const void* d_tabP, uint8_t* d_ct);
int launch_encrypt(cudaStream_t st, size_t n, const int64_t* d_m, const uint8_t* d_r, const void* d_tabB,
                   const void* d_tabP, uint8_t* d_ct);
// scratch: slices * 2v Jacobian points
size_t aggregate_scratch_bytes(size_t ndp, size_t v, int* slices_out);
int launch_agg_partial(cudaStream_t st, size_t ndp, size_t v, const uint8_t* d_cts, void* d_scratch, int slices,
                       int* d_bad);
int launch_agg_final(cudaStream_t st, size_t v, int slices, const void* d_scratch, uint8_t* d_out);
// allocation-free aggregation over a fixed module-resident scratch array (kernels_g1.cu)
size_t aggregate_static_max_columns();
int launch_aggregate_static(cudaStream_t st, size_t ndp, size_t v, const uint8_t* d_cts, uint8_t* d_out, int* n_bad,
                            int* launches);
// per-provider statistics of int64 rows [ndp][nrows] for the sum / mean / variance encoders:
// mode 0: [sum]; 1: [sum, N]; 2: [sum, N, sum of squares]; 3: mean then variance = [sum, N, sum, N, sumsq]
int launch_reduce_rows(cudaStream_t st, size_t ndp, size_t nrows, const int64_t* d_rows, int mode, int64_t* d_out);
int launch_dlog_build(cudaStream_t st, const void* d_tabB, const void* d_baseB_aff, void* d_keys, uint32_t* d_vals);
int launch_decrypt(cudaStream_t st, size_t n, const uint8_t* d_sk, const uint8_t* d_ct, const void* d_keys,
                   const uint32_t* d_vals, int64_t* d_out, int* d_bad);
// key switching: cv[i] = (v_i*B) || (v_i*Q - sk*rb_i), rb_neg[i] = -rb_i (optional)
int launch_keyswitch(cudaStream_t st, size_t n, const uint8_t* d_sk, const uint8_t* d_rb, const uint8_t* d_v,
                     const void* d_tabB, const void* d_tabQ, uint8_t* d_cv, uint8_t* d_rb_neg, int* d_bad);
// ShuffleSequence: out line i = in line pi[i], every ciphertext re-randomised with beta[pi[i]][j]
int launch_shuffle_rerandomize(cudaStream_t st, size_t lines, size_t nq, const int32_t* d_pi, const uint8_t* d_in,
                               const uint8_t* d_beta, const void* d_tabB, const void* d_tabP, uint8_t* d_out, int* d_bad);
// MulCipherTextbyScalar per element: out_i = (s_i*K_i, s_i*C_i)
int launch_ct_scalar_mul(cudaStream_t st, size_t n, const uint8_t* d_s, const uint8_t* d_ct, uint8_t* d_out, int* d_bad);
// Schnorr over G1: ok_i = (s_i*B == R_i + h_i*P_i), sig = R || s (96 B), digest = SHA-512(R || P || msg) (64 B)
int launch_rep_commit(cudaStream_t st, size_t n, int nt, const uint8_t* d_sc, const uint8_t* d_pts, uint8_t* d_out, int* d_bad);
int launch_rep_respond(cudaStream_t st, size_t n, int per, const uint8_t* d_v, const uint8_t* d_c, const uint8_t* d_x,
                       uint8_t* d_r);
int launch_rep_verify(cudaStream_t st, size_t n, int nt, const uint8_t* d_resp, const uint8_t* d_chal, const uint8_t* d_bases,
                      const uint8_t* d_P, const uint8_t* d_V, uint8_t* d_ok);
int launch_schnorr_verify(cudaStream_t st, size_t n, const uint8_t* d_pub, const uint8_t* d_sig, const uint8_t* d_digest,
                          const void* d_tabB, uint8_t* d_ok);
int launch_g1_mul(cudaStream_t st, size_t n, const uint8_t* d_k, const uint8_t* d_pt, uint8_t* d_out, int* d_bad);
int launch_mont_mul(cudaStream_t st, size_t n, int which, const uint32_t* d_a, const uint32_t* d_b, uint32_t* d_r);
int launch_microbench(cudaStream_t st, int kind, int iters, int blocks, int threads, uint32_t* d_sink);
int launch_g1_sum_wire(cudaStream_t st, int n, const void* d_aff, uint8_t* d_out_wire);

// ---- kernels_rp.cu
int launch_g2_from_wire(cudaStream_t st, size_t n, const uint8_t* d_wire, void* d_aff, int* d_bad);
int launch_comb_table_g2(cudaStream_t st, int nbases, const void* d_bases_aff, void* d_tables);
int launch_pair_aff(cudaStream_t st, size_t n, const void* d_g1_aff, size_t g1_stride_items, const void* d_g2_aff,
                    void* d_gt);  // g1_stride_items = 0 broadcasts one G1 point
int launch_comb_table_gt(cudaStream_t st, int nbases, const void* d_bases_gt, void* d_tables);
constexpr size_t kGt16PerWindow = 65535 /* = DRX_GT16_PER_WINDOW */, kGt16Entries = 16 * kGt16PerWindow;  // 16-bit windows of e(B, g2): 403 MB of Fp12
int launch_comb_table_gt16(cudaStream_t st, const void* d_tab8, void* d_tab16);
int launch_pair_wire(cudaStream_t st, size_t n, const uint8_t* d_g1, const uint8_t* d_g2, uint8_t* d_gt, int* d_bad);
int launch_g2_mul(cudaStream_t st, size_t n, const uint8_t* d_k, const uint8_t* d_pt, uint8_t* d_out, int* d_bad);
int launch_make_sigs(cudaStream_t st, int ns, int64_t u, const uint8_t* d_x, const void* d_tabB, const void* d_tabG2,
                     uint8_t* d_Y, uint8_t* d_A);

struct RpSetDev {        // device-resident per-signature-set data (array indexed by set_id on the device)
  const void* tabY;      // ns comb tables of Y_i (G1)
  const void* tabA;      // ns*u comb tables of A_{i,k} (G2)
  const void* tabGtA;    // ns*u comb tables of e(B, A_{i,k}) (GT)
  const uint8_t* sumY;   // 64 B wire form of sum_i Y_i
  int32_t ns;
  int32_t u;
};

struct RpBatch {          // one create/verify call
  size_t n;               // proofs
  size_t items;           // sum over proofs of ns*nd
  const int32_t* d_set_id;
  const int64_t* d_l;
  const size_t* d_proof_off;   // n+1, or nullptr for a uniform batch
  const uint32_t* d_item_proof;  // items, or nullptr for a uniform batch
  const uint32_t* d_item_start;  // n (first item of each proof), or nullptr
  int nsets;              // length of the device array of signature sets (ids at or beyond it are invalid)
  int uni_ns;             // uniform batch: servers
  int uni_nd;             // uniform batch: digits
  size_t uni_size;        // uniform batch: bytes per proof
};

int launch_rp_create_hdr(cudaStream_t st, const RpBatch& b, const RpSetDev* d_sets, const int64_t* d_secret,
                         const uint8_t* d_r, const uint8_t* d_ct, const uint8_t* d_rnd, const size_t* d_rnd_off,
                         const void* d_tabB, const void* d_tabP, const uint8_t* d_baseB_wire, uint8_t* d_proofs);
int launch_rp_verify_hdr(cudaStream_t st, const RpBatch& b, const RpSetDev* d_sets, const void* d_tabB,
                         const void* d_tabP, const uint8_t* d_proofs, uint32_t* d_flags);
int launch_rp_flags(cudaStream_t st, size_t n, const uint32_t* d_flags, uint8_t* d_ok);
// role-split item kernels (independent parts of one item in different thread blocks) + their finish kernels;
// d_scratch holds rp_scratch_bytes(items) bytes (two GT elements per item)
size_t rp_scratch_bytes(size_t items);
int launch_rp_create_lanes(cudaStream_t st, const RpBatch& b, const RpSetDev* d_sets, const int64_t* d_secret,
                           const uint8_t* d_rnd, const size_t* d_rnd_off, const void* d_tabGtB,
                           const uint8_t* d_baseB_wire, const uint8_t* d_ct, uint8_t* d_proofs);
// g2_only: just the V_ij (G2) and Zv_ij parts; the GT factors then come from launch_rp_create_gt_lanes
int launch_rp_create_fine(cudaStream_t st, const RpBatch& b, const RpSetDev* d_sets, const int64_t* d_secret,
                          const uint8_t* d_rnd, const size_t* d_rnd_off, const void* d_tabGtB,
                          const uint8_t* d_baseB_wire, const uint8_t* d_ct, uint8_t* d_proofs, void* d_scratch,
                          bool g2_only = false);
// fixed-base GT exponentiations on lane groups (kernels_rp_gtlanes.cu): a_ij of creation, T_ij of verification
int launch_rp_create_gt_lanes(cudaStream_t st, const RpBatch& b, const RpSetDev* d_sets, const int64_t* d_secret,
                              const uint8_t* d_rnd, const size_t* d_rnd_off, const void* d_tabGtB16, uint8_t* d_proofs);
int launch_rp_verify_gt_lanes(cudaStream_t st, const RpBatch& b, const RpSetDev* d_sets, const void* d_tabGtB16,
                              const uint8_t* d_proofs, uint32_t* d_flags, void* d_scratch);
int launch_rp_create_roles(cudaStream_t st, const RpBatch& b, const RpSetDev* d_sets, const int64_t* d_secret,
                           const uint8_t* d_rnd, const size_t* d_rnd_off, const void* d_tabGtB,
                           const uint8_t* d_baseB_wire, const uint8_t* d_ct, uint8_t* d_proofs, void* d_scratch);
int launch_rp_verify_roles(cudaStream_t st, const RpBatch& b, const RpSetDev* d_sets, const void* d_tabB,
                           const void* d_tabGtB, const uint8_t* d_proofs, uint32_t* d_flags, void* d_scratch);
// cooperative lane-group pairing driven by a generated micro-program (kernels_rp_ucode.cu)
size_t rp_ucode_prep_bytes(size_t items);
int launch_rp_verify_prep(cudaStream_t st, const RpBatch& b, const RpSetDev* d_sets, const void* d_tabB,
                          const uint8_t* d_proofs, uint32_t* d_flags, void* d_prep);
// four / eight program lanes of two threads each per pairing check (kernels_rp_ucode_s4.cu / _s8.cu);
// d_err: one device word; bit 0 is raised if a dynamic launch gave up waiting (see ucode_common.cuh)
struct UcTune {
  int mode = 0;      // 0 = by batch size, 1 = static launch, 2 = dynamic (task queue) launch
  int resident = 0;  // dynamic launch: persistent blocks per SM (0 = as many as fit)
  int nseg = 0;      // dynamic launch: program segments (0 = 32)
};
int launch_pair_ucode_s4(cudaStream_t st, size_t items, const void* d_prep, void* d_scratch, unsigned int* d_err,
                         const UcTune& tune);
int launch_pair_ucode_s8(cudaStream_t st, size_t items, const void* d_prep, void* d_scratch, unsigned int* d_err,
                         const UcTune& tune);
int launch_rp_verify_gt(cudaStream_t st, const RpBatch& b, const RpSetDev* d_sets, const void* d_tabB,
                        const void* d_tabGtB, const void* d_tabGtB16 /* may be null */, const uint8_t* d_proofs,
                        uint32_t* d_flags, void* d_scratch);
int launch_rp_verify_finish(cudaStream_t st, const RpBatch& b, const RpSetDev* d_sets, const void* d_scratch,
                            uint32_t* d_flags);
// edwards25519 ElGamal (kernels_ed.cu); points travel as 32-byte compressed encodings, ciphertexts as K || C (64 B)
int launch_ed_from_wire(cudaStream_t st, size_t n, const uint8_t* d_wire, void* d_pts, int* d_bad);
int launch_ed_comb_table(cudaStream_t st, const void* d_base_pt, void* d_table);
int launch_ed_encrypt(cudaStream_t st, size_t n, const int64_t* d_m, const uint8_t* d_r, const void* d_tabB,
                      const void* d_tabP, uint8_t* d_ct);
size_t ed_aggregate_scratch_bytes(size_t ndp, size_t v, int* slices_out);
int launch_ed_aggregate(cudaStream_t st, size_t ndp, size_t v, const uint8_t* d_cts, void* d_scratch, int slices,
                        uint8_t* d_out, int* d_bad);
int launch_ed_dlog_build(cudaStream_t st, const void* d_tabB, void* d_keys, uint32_t* d_vals);
int launch_ed_decrypt(cudaStream_t st, size_t n, const uint8_t* d_sk, const uint8_t* d_ct, const void* d_keys,
                      const uint32_t* d_vals, int64_t* d_out, int* d_bad);
int launch_ed_mul(cudaStream_t st, size_t n, const uint8_t* d_k, const uint8_t* d_pt, uint8_t* d_out, int* d_bad);
constexpr size_t kEdNielsBytes = 96, kEdExtBytes = 128;
}  // namespace drx
