/* drynx_gpu.h -- C ABI of the B200 (sm_100a) backend for Drynx's data-provider / verifying-node crypto path.
 *
 * The reference (cgrigis/drynx, Go) has no FFI: the seam is ordinary Go functions over kyber.Point /
 * kyber.Scalar with the global suite bn256.G1 (lib/suite.go:10-15). Every type on the path already has a
 * fixed-width big-endian byte form, so this ABI is bytes in, bytes out:
 *
 *   scalar      32 B  big-endian, reduced mod n on input (kyber Scalar.SetBytes / UnmarshalBinary)
 *   G1 point    64 B  x || y, infinity = 64 zero bytes (kyber pointG1.MarshalBinary)
 *   G2 point   128 B  x.i || x.1 || y.i || y.1          (kyber pointG2.MarshalBinary)
 *   GT element 384 B  12 x 32 B, order x.x.x ... y.z.y  (kyber pointGT.MarshalBinary)
 *   ciphertext 128 B  K || C                            (libunlynx CipherText.ToBytes)
 *   range proof       Commit(128) | Challenge(32) | Zr(32) | D(64) | Zphi(32*l) |
 *                     per server i: V_i(128*l) | A_i(384*l) | Zv_i(32*l)
 *                     (lib/range/range_proof.go:72-155 field order; 256 + 32*l + 544*ns*l bytes)
 *
 * Rules: every pointer is caller-owned memory that is only read or written during the call; nothing is
 * retained (cgo pointer rules). Randomness is an INPUT: the host draws r, s_j, t_j, m_j, v_ij and passes them
 * in (the reference draws them from crypto/rand inside goroutines). Return value:
 * 0 = success, negative = invalid argument / malformed input, positive = CUDA error code; text through
 * drx_last_error(). Nothing here falls back to the CPU: without a usable sm_100a device drx_init fails.
 * A context is bound to one device and one stream and serialises its calls with an internal lock.
 *
 * Functions ending in _dev take DEVICE pointers on the context's device, enqueue on the context's stream and
 * return without synchronising; they exist so that a caller that keeps ciphertexts resident in HBM (and the
 * benchmark's device-resident leg) skips the host staging copies.
 */
#ifndef DRYNX_GPU_H
#define DRYNX_GPU_H
#include <stddef.h>
#include <stdint.h>
#ifdef __cplusplus
extern "C" {
#endif

typedef struct drx_ctx drx_ctx;

/* Creates a context on CUDA device `device`. g1_base / g2_base are the suite generators in wire form as the
 * Go side sees them (lib.Suite.Point().Base().MarshalBinary(), bn256.NewSuite().G2().Point().Base()); NULL
 * selects (1, -2) and the x/crypto twist generator that kyber's constants encode. Builds the fixed-base
 * tables for B and for e(B, g2). Replaces: package init of lib/suite.go:12-15. */
int drx_init(drx_ctx** out, int device, const uint8_t* g1_base /*64 or NULL*/, const uint8_t* g2_base /*128 or NULL*/);
void drx_free(drx_ctx* ctx);
/* Thread-local text of the last failure on this thread; never NULL. */
const char* drx_last_error(void);
/* Use an existing cudaStream_t (e.g. torch's current stream) for every later call; NULL restores the
 * context's own non-blocking stream. To name the legacy default stream pass cudaStreamLegacy ((void*)1). */
int drx_set_stream(drx_ctx* ctx, void* cuda_stream);
/* Waits for everything enqueued on the context's stream. Returns -5 if a range-proof verification enqueued since the
 * last synchronising call had to abandon its pairing kernel (a device-side wait that never completed: the results of
 * that call are invalid, repeat it); 0 otherwise. */
int drx_synchronize(drx_ctx* ctx);
/* Status path of the device-buffer (`_dev`) entry points, which only enqueue work and so cannot return a count:
 * malformed input points they meet (off the curve, coordinate >= p: what kyber's UnmarshalBinary refuses and the
 * reference aborts on) are added to a per-context counter; the affected output rows are written as ZEROS (the point at
 * infinity), an aggregate skips the malformed term, drx_decrypt_dev writes INT64_MIN. drx_take_bad waits for the stream,
 * stores the count accumulated since the previous drx_take_bad in *n_bad and clears it. A caller that needs the
 * reference's "malformed input is an error" behaviour calls it once after its batch of `_dev` calls and treats
 * *n_bad != 0 like the -2 of the host-buffer entry points. No counterpart in the reference (its calls are synchronous). */
int drx_take_bad(drx_ctx* ctx, int* n_bad);
/* Number of kernel launches issued by this context so far (for the benchmark's gpu_launches field). */
uint64_t drx_launch_count(const drx_ctx* ctx);

/* Tuning knobs (no counterpart in the reference). "rp_verify_path": 0 = choose by batch size (default),
 * 1 = one thread per pairing check, 2 = cooperative lane groups. Results are identical on every path. */
int drx_set_option(drx_ctx* ctx, const char* name, int value);

/* Per-kernel timing with CUDA events on the launching stream (bench.py's roofline leg): enable (which clears
 * the totals), run, then read the summed duration and launch count of a kernel by name (k_encrypt,
 * k_agg_partial, k_agg_final, k_decrypt, k_reduce_rows, k_rp_create_hdr, k_rp_create_items, k_rp_verify_hdr,
 * k_rp_verify_items or k_rp_verify_prep + k_pair_ucode + k_rp_verify_gt + k_rp_verify_finish, k_rp_flags). on = 1 brackets every kernel,
 * on = 2 only the pairing kernel of verification (what bench.py's timed region uses: one event pair per step), 0 none.
 * No counterpart in the reference (it times with onet monitor.TimeMeasure). */
int drx_profile_enable(drx_ctx* ctx, int on);
int drx_profile_get(drx_ctx* ctx, const char* kernel, double* total_ms, uint64_t* launches);

/* Per-provider plaintext reduction of int64 rows resident on the device, d_rows[ndp][nrows]
 * (lib/encoding/sum.go:19-23, mean.go:19-24, variance.go:19-27). mode 0: [sum]; 1: [sum, N];
 * 2: [sum, N, sum of squares]; 3: the mean vector followed by the variance vector (5 values). */
int drx_reduce_rows_dev(drx_ctx* ctx, size_t ndp, size_t nrows, const int64_t* d_rows, int mode, int64_t* d_out);
/* The same with host buffers: rows[ndp][nrows] in, out[ndp][w] back, w = 1, 2, 3 or 5 by mode (what a data
 * provider calls in place of the Go loops of EncodeSum / EncodeMean / EncodeVariance before it encrypts). */
int drx_reduce_rows(drx_ctx* ctx, size_t ndp, size_t nrows, const int64_t* rows, int mode, int64_t* out);

/* EncodeFreqCount's plaintext step (lib/encoding/frequency_count.go:18-36) for ndp providers of nrows rows each:
 * out[p][v - qmin] = number of rows of provider p equal to v, for v in [qmin, qmax] (at most 12,000 values). A row
 * outside the range (an index panic in the Go loop) is error -2. */
int drx_freq_count(drx_ctx* ctx, size_t ndp, size_t nrows, const int64_t* rows, int64_t qmin, int64_t qmax, int64_t* out);

/* The float64 pre-pass of logistic regression (lib/encoding/logistic_regression.go:46-118: StandardiseWith :964-978,
 * Augment :1014-1023, ComputeAllApproxCoefficients :367-403, AggregateApproxCoefficients :433-455,
 * Float64ToInt642DArrayWithPrecision :1558-1574) for ndp data providers at once. Provider p owns records
 * rec_off[p] .. rec_off[p+1]-1 of x[record][d] (float64) and y[record] (0/1); out[p] is its packed vector of
 * (d+1) + (d+1)^2 int64 coefficients for k = 2 (d+1 for k = 1), ready for drx_encrypt. Every sum runs in record
 * order with correctly rounded, unfused float64 operations: the integers equal the Go loop's bit for bit.
 * d <= 31, k <= 2 (the reference's packing offset j*(d+1)^j is only consistent up to k = 2). */
int drx_logreg_coefficients(drx_ctx* ctx, size_t ndp, const size_t* rec_off /*ndp+1*/, int d, int k, const double* x,
                            const int64_t* y, const double* means /*d*/, const double* sds /*d*/, double precision,
                            int64_t* out);
/* device-resident form; nrec = total number of records = d_rec_off[ndp] (d_rec_off[0] must be 0) */
int drx_logreg_coefficients_dev(drx_ctx* ctx, size_t ndp, size_t nrec, const size_t* d_rec_off, int d, int k,
                                const double* d_x, const int64_t* d_y, const double* d_means, const double* d_sds,
                                double precision, int64_t* d_out);

/* Collective public key P for EncryptInt* and range proofs (CreateProof.CaPub, lib/range/range_proof.go:67);
 * builds its fixed-base table. */
int drx_set_pubkey(drx_ctx* ctx, const uint8_t P[64]);

/* libunlynx.EncryptIntGetR / EncryptIntVectorGetRs with the ephemeral scalars injected (call sites
 * lib/encoding/sum.go:24, mean.go:33, variance.go:36, frequency_count.go:36, linear_regression_dims.go:40-85,
 * logistic_regression.go:420): ct[i] = (r_i*B) || (m_i*B + r_i*P). Negative m maps to n-|m| (SetInt64). */
int drx_encrypt(drx_ctx* ctx, size_t n, const int64_t* m, const uint8_t* r /*n*32*/, uint8_t* ct /*n*128*/);
int drx_encrypt_dev(drx_ctx* ctx, size_t n, const int64_t* d_m, const uint8_t* d_r, uint8_t* d_ct);
/* libunlynx.EncryptScalar (call sites lib/encoding/OR_AND.go:52,105: the bit encodings of OR / AND / min / max /
 * set union / intersection): the message is a scalar (32 bytes big-endian, reduced mod n), C = m*B + r*P. */
int drx_encrypt_scalar(drx_ctx* ctx, size_t n, const uint8_t* m /*n*32*/, const uint8_t* r /*n*32*/, uint8_t* ct /*n*128*/);

/* CipherVector.FromBytes + CipherVector.Add over all data providers
 * (protocols/data_collection_protocol.go:145-166, lib/structs.go:338-350): out[j] = sum_d cts[d][j].
 * cts holds ndp pointers to v*128 bytes each. Returns -2 if any point is malformed. */
int drx_aggregate(drx_ctx* ctx, size_t ndp, size_t v, const uint8_t* const* cts, uint8_t* out /*v*128*/);
/* Same with the ndp vectors contiguous in device memory (d_cts[d*v*128 ...]); malformed points: see drx_take_bad. */
int drx_aggregate_dev(drx_ctx* ctx, size_t ndp, size_t v, const uint8_t* d_cts, uint8_t* d_out);

/* Aggregation for callers that must not allocate GPU memory. It has no counterpart in the reference; it computes
 * exactly what drx_aggregate_dev computes (d_cts = ndp vectors of v ciphertexts back to back in device memory,
 * d_out = v ciphertexts), but needs no context and calls no cudaMalloc/cudaFree or any other allocating routine:
 * the partial sums go into a fixed 3 MiB array inside this library's device code. CUDA loads a library's device code
 * (and reserves such arrays) lazily, at the first use on each device -- for a caller without a context that is the
 * first drx_aggregate_noalloc_dev call itself. So make ONE warm-up call per device at start-up (ndp = 0, v = 1 is
 * enough), BEFORE the memory pool takes the rest of the GPU and before any stream capture: that call is where the
 * 3 MiB and the code are placed, and the only one that can fail for lack of memory; from then on the GPU holds them
 * until the process exits and no later call allocates anything. Meant for hosts that give the whole GPU to one memory pool at start-up and have nothing left
 * to allocate from later, for services whose policy is "no allocation after initialisation" because a failed
 * cudaMalloc in the middle of a survey cannot be handled, and for latency-sensitive loops where the device-wide
 * synchronisation hidden in cudaMalloc/cudaFree would stall the other streams.
 *   - runs on the CURRENT device, on `cuda_stream` (NULL = the default stream), and returns after the result is
 *     complete (it synchronises that stream to read the malformed-point count);
 *   - limit: v <= 16384 columns per call; larger v returns -5 with a message, split the vector instead;
 *   - the first call on each device looks up the array's device address (cudaGetSymbolAddress); make that first
 *     call before starting any stream capture on that device;
 *   - all calls on one GPU share the array, which still holds the previous call's partial sums: calls must not
 *     overlap on the GPU (from different streams or threads) or they corrupt each other's results. The one word
 *     that has to start at zero, the malformed-point counter, is cleared by every call.
 * Returns 0, -1 (null buffer), -2 (malformed points), -5 (v too large) or a CUDA error code. */
int drx_aggregate_noalloc_dev(size_t ndp, size_t v, const uint8_t* d_cts, uint8_t* d_out, void* cuda_stream);

/* libunlynx.CreateDecryptionTable (services/api.go:50): makes i*B, 0 <= i <= limit, available to drx_decrypt
 * with negatives. The device table always spans the reference's MaxHomomorphicInt (100000); `limit` only
 * feeds the host-side model of which entries the reference's lazily grown map would hold. */
int drx_dlog_table(drx_ctx* ctx, int64_t limit);
/* libunlynx.DecryptInt / DecryptIntWithNeg / DecryptIntVector(WithNeg) (lib/encoding/sum.go:38, mean.go:66,
 * variance.go:68, frequency_count.go:70, encode_decode.go:224): M = C - sk*K, then the discrete log by table.
 * out[i] = value, or 0 with found[i] = 0 when |value| >= 100000 (the reference logs "out of bound" and
 * returns 0). found may be NULL. */
int drx_decrypt(drx_ctx* ctx, size_t n, const uint8_t sk[32], const uint8_t* ct /*n*128*/, int with_neg,
                int64_t* out, uint8_t* found);
int drx_decrypt_dev(drx_ctx* ctx, size_t n, const uint8_t sk[32], const uint8_t* d_ct, int64_t* d_out /*raw signed dlog, INT64_MIN = miss*/);
/* libunlynx.DecryptCheckZero (lib/encoding/OR_AND.go:61,114): out[i] = 0 if the plaintext point is the
 * identity, else 1. */
int drx_decrypt_check_zero(drx_ctx* ctx, size_t n, const uint8_t sk[32], const uint8_t* ct, int64_t* out);

/* Key switching: one server's contribution to moving n ciphertexts from the collective
 * key to the querier's key Q. Replaces libunlynxkeyswitch.KeySwitchSequence(targetPubKey, rBs, secretKey)
 * (github.com/ldsec/unlynx v1.4.0 lib/key_switch/key_switch.go; call sites data/data.go:80 and, through
 * protocolsunlynx.KeySwitchingProtocol, services/service.go:565-616, 853-862) with the per-element randomness vi
 * injected: cv[i] = (v_i*B) || (v_i*Q - sk*rb_i), where rb_i is the K component of ciphertext i; rb_neg[i] = -rb_i
 * (the proof's rBNegs; may be NULL). The root adds the contributions of all servers to (0, C_i) with
 * drx_aggregate; the result decrypts under the querier's secret. Returns -2 on a malformed point. */
int drx_set_target_pubkey(drx_ctx* ctx, const uint8_t Q[64]);
int drx_keyswitch(drx_ctx* ctx, size_t n, const uint8_t sk[32], const uint8_t* rb /*n*64*/, const uint8_t* v /*n*32*/,
                  uint8_t* cv /*n*128*/, uint8_t* rb_neg /*n*64 or NULL*/);
int drx_keyswitch_dev(drx_ctx* ctx, size_t n, const uint8_t sk[32], const uint8_t* d_rb, const uint8_t* d_v, uint8_t* d_cv,
                      uint8_t* d_rb_neg);

/* Obfuscation data path: CipherText.MulCipherTextbyScalar per element
 * (protocols/obfuscation_protocol.go:240-243, data/data.go:52-56): out[i] = (s_i*K_i) || (s_i*C_i), s 32 bytes
 * big-endian reduced mod n. Returns -2 on a malformed point. */
int drx_ct_scalar_mul(drx_ctx* ctx, size_t n, const uint8_t* s /*n*32*/, const uint8_t* ct /*n*128*/, uint8_t* out /*n*128*/);
int drx_ct_scalar_mul_dev(drx_ctx* ctx, size_t n, const uint8_t* d_s, const uint8_t* d_ct, uint8_t* d_out);

/* Shuffling data path: libunlynxshuffle.ShuffleSequence(inputList, g, h, nil) with the
 * blinding factors and the permutation injected (unlynx v1.4.0 lib/shuffle; call sites data/data.go:68 and, through
 * protocolsunlynx.ShufflingProtocol, services/service.go:618-660). `in` holds `lines` CipherVectors of `nq` ciphertexts;
 * out line i = in line pi[i] with every ciphertext re-randomised (CipherText.Rerandomize):
 * K' = K + beta[pi[i]][j]*B, C' = C + beta[pi[i]][j]*P, P the key of drx_set_pubkey. beta is indexed like `in`.
 * Returns -1 when pi is not a permutation of 0..lines-1 (host-buffer form), -2 on a malformed point. */
int drx_shuffle_rerandomize(drx_ctx* ctx, size_t lines, size_t nq, const int32_t* pi, const uint8_t* in /*lines*nq*128*/,
                            const uint8_t* beta /*lines*nq*32*/, uint8_t* out /*lines*nq*128*/);
int drx_shuffle_rerandomize_dev(drx_ctx* ctx, size_t lines, size_t nq, const int32_t* d_pi, const uint8_t* d_in,
                                const uint8_t* d_beta, uint8_t* d_out);

/* Schnorr signature check over G1, batched: kyber sign/schnorr Verify as
 * VerifyProofSignature calls it for every proof request (lib/proof/structs_proofs.go:498-505).
 * pub: n*64 signer keys; sig: n*96, R (64) || s (32); digest: n*64, SHA-512(R || P || msg) computed by the caller
 * (the message is the proof list the caller already holds). ok[i] = 1 when s*B == R + h*P with h the digest read
 * big-endian mod n. A malformed R or P, or s >= n, gives ok[i] = 0 (the reference returns an error for that
 * signature), never an error of the call. */
int drx_schnorr_verify(drx_ctx* ctx, size_t n, const uint8_t* pub, const uint8_t* sig, const uint8_t* digest, uint8_t* ok /*n*/);

/* proof.Rep relations, batched: the group and scalar steps of the two kyber proof.Rep
 * proofs on the path -- the obfuscation proof (lib/obfuscation/obfuscation_proof.go:37-91, created at
 * protocols/obfuscation_protocol.go:259 and data/data.go:57, verified at lib/proof/structs_proofs.go:318-338:
 * co1 = s*c1, co2 = s*c2) and unlynx's key-switch proof (libunlynxkeyswitch.KeySwitchListProofCreation /
 * KeySwitchListProofVerification, unlynx v1.4.0; call sites services/service.go:581, data/data.go:81,
 * lib/proof/structs_proofs.go:469-482: viB = vi*B, K = k*B, ks2 = vi*Q + k*rBNeg). A relation is
 * P = x_1*B_1 + ... + x_T*B_T (T = nterms, 1..4096). As kyber v3.0.8 proof/rep.go does it: the prover sends
 * V = sum_j v_j*B_j (drx_rep_commit; with nterms = 1 on both halves of a ciphertext it equals drx_ct_scalar_mul),
 * the challenge c comes out of the transcript hash of every V of the proof, the answers are r_j = v_j - c*x_j
 * (drx_rep_respond: `per` consecutive secrets share challenge i / per), and the verifier accepts when
 * sum_j r_j*B_j + c*P == V (drx_rep_verify). The transcript (kyber proof.HashProve / HashVerify over the suite's XOF)
 * stays on the caller's side: the challenge is an input here, the proof bytes are V's then r's in kyber's order.
 * scalars, resp, chal: 32 bytes big-endian; points 64 bytes. drx_rep_commit returns -2 on a malformed point;
 * drx_rep_verify gives ok[i] = 0 for a malformed point or a response >= n (kyber fails that proof), never an error
 * of the call. drx_rep_commit with nterms = the length of a line also is the list compression that opens unlynx's shuffle
 * proof (libunlynxshuffle.ShuffleProofCreation / ShuffleProofVerification, call sites services/service.go:634,
 * data/data.go:69, lib/proof/structs_proofs.go:392-403: sum_j e_j*K_ij and sum_j e_j*C_ij per line i); the Neff
 * pair-shuffle argument that follows it (kyber shuffle.PairShuffle) consists of sums and checks of these two shapes
 * and runs over these calls as well (go/gpubackend/shuffle.go, tests/shuffle_proof.py). One device thread walks the terms
 * of a relation (~1.2 ms per term when n is small), so drx_rep_commit sums relations of 33 terms and more in two levels
 * (chunks of about sqrt(nterms) terms, then the partial sums; same bytes). drx_rep_commit_dev and drx_rep_verify do not:
 * callers with few, wide relations there split them themselves, as wideSum in go/gpubackend/shuffle.go does. The `_dev`
 * forms only enqueue work:
 * drx_rep_commit_dev zeroes the row of a relation with a malformed point and counts it for drx_take_bad. */
int drx_rep_commit(drx_ctx* ctx, size_t n, int nterms, const uint8_t* scalars /*n*nterms*32*/,
                   const uint8_t* points /*n*nterms*64*/, uint8_t* out /*n*64*/);
int drx_rep_commit_dev(drx_ctx* ctx, size_t n, int nterms, const uint8_t* d_scalars, const uint8_t* d_points, uint8_t* d_out);
int drx_rep_respond(drx_ctx* ctx, size_t n, int per, const uint8_t* v /*n*32*/, const uint8_t* chal /*ceil(n/per)*32*/,
                    const uint8_t* x /*n*32*/, uint8_t* r /*n*32*/);
int drx_rep_verify(drx_ctx* ctx, size_t n, int nterms, const uint8_t* resp /*n*nterms*32*/, const uint8_t* chal /*n*32*/,
                   const uint8_t* bases /*n*nterms*64*/, const uint8_t* P /*n*64*/, const uint8_t* V /*n*64*/,
                   uint8_t* ok /*n*/);
int drx_rep_verify_dev(drx_ctx* ctx, size_t n, int nterms, const uint8_t* d_resp, const uint8_t* d_chal, const uint8_t* d_bases,
                       const uint8_t* d_P, const uint8_t* d_V, uint8_t* d_ok);

/* Server signatures for one output column: Y_i and A_{i,k} = (x_i + k)^-1 * g2, k < u
 * (lib/range/range_proof.go:249-302 InitRangeProofSignature, PublishSignatureBytesToPublishSignatures;
 * libdrynx.PublishSignatureBytes lib/structs.go:131-141). Builds the per-set tables; set_id is the caller's
 * handle (0 <= set_id < 65536) used by create/verify. */
int drx_rp_set_sigs(drx_ctx* ctx, int set_id, int ns, int64_t u, const uint8_t* Y /*ns*64*/,
                    const uint8_t* A /*ns*u*128*/);
/* InitRangeProofSignature on the device from caller-drawn private keys x_i (:254-266, :275-287):
 * Y_out[i] = x_i*B, A_out[i][k] = (x_i + k)^-1 * g2. */
int drx_rp_make_sigs(drx_ctx* ctx, int ns, int64_t u, const uint8_t* x /*ns*32*/, uint8_t* Y_out /*ns*64*/,
                     uint8_t* A_out /*ns*u*128*/);
/* Number of base-u digits ToBase(secret, u, l) yields (lib/range/range_proof.go:584-595): max(l, digits). */
int64_t drx_rp_ndigits(int64_t secret, int64_t u, int64_t l);
/* Serialised size of one proof with nd digits for ns servers (128 when set_id < 0: commit only). */
size_t drx_rp_proof_size(int ns, int64_t nd);

/* CreatePredicateRangeProofListForAllServers (lib/range/range_proof.go:305-407) for n values.
 * Per proof p: set_id[p] (< 0: u = l = 0, emit the ciphertext only, :322-324), l[p], secret[p], the
 * encryption scalar r[p], the ciphertext ct[p], and nd*(3+ns) random scalars at rnd + rnd_off[p] in the
 * reference's draw order: for each digit j: s_j, t_j, m_j (:377-379) then v_0j .. v_{ns-1,j} (:390).
 * Proof p is written at proofs + proof_off[p]; proof_off[p+1]-proof_off[p] must equal drx_rp_proof_size. */
int drx_rp_create(drx_ctx* ctx, size_t n, const int32_t* set_id, const int64_t* l, const int64_t* secret,
                  const uint8_t* r /*n*32*/, const uint8_t* ct /*n*128*/, const uint8_t* rnd,
                  const size_t* rnd_off /*n+1, bytes*/, uint8_t* proofs, const size_t* proof_off /*n+1*/);
/* RangeProofVerification for n proofs (lib/range/range_proof.go:504-565; list driver :484-501, caller
 * lib/proof/structs_proofs.go:171): ok[p] = 1 iff the size check (:512), every pairing equation (:540-549)
 * and the D check (:561) hold. set_id[p] < 0 means u = l = 0 -> 1 (:508-510). Malformed group elements
 * yield 0 (the reference aborts in FromBytes, :225-236). */
int drx_rp_verify(drx_ctx* ctx, size_t n, const int32_t* set_id, const int64_t* l, const uint8_t* proofs,
                  const size_t* proof_off /*n+1*/, uint8_t* ok /*n*/);
/* Device-resident uniform batch: n proofs of identical size (ns servers, l digits) back to back. */
int drx_rp_verify_dev(drx_ctx* ctx, size_t n, const int32_t* d_set_id, int ns, int64_t l, const uint8_t* d_proofs,
                      uint8_t* d_ok);
int drx_rp_create_dev(drx_ctx* ctx, size_t n, const int32_t* d_set_id, int ns, int64_t l, const int64_t* d_secret,
                      const uint8_t* d_r, const uint8_t* d_ct, const uint8_t* d_rnd, uint8_t* d_proofs);

/* Diagnostics used by tests and bench.py: raw pairing / group operations on wire bytes, n items each. */
int drx_pair(drx_ctx* ctx, size_t n, const uint8_t* g1 /*n*64*/, const uint8_t* g2 /*n*128*/, uint8_t* gt /*n*384*/);
int drx_g1_mul(drx_ctx* ctx, size_t n, const uint8_t* k /*n*32*/, const uint8_t* pt /*n*64*/, uint8_t* out /*n*64*/);
int drx_g2_mul(drx_ctx* ctx, size_t n, const uint8_t* k /*n*32*/, const uint8_t* pt /*n*128*/, uint8_t* out /*n*128*/);
/* Field self-test: r[i] = a[i]*b[i]*2^-256 mod (which ? n : p), raw little-endian limbs, n items. */
int drx_mont_mul(drx_ctx* ctx, size_t n, int which, const uint32_t* a, const uint32_t* b, uint32_t* r);
/* Integer-pipe roofline probes: runs `iters` dependent steps per thread over a full grid and returns the
 * kernel time in ms. kind 0: register-resident IMAD.WIDE chains (64 MAC32 per step per thread);
 * kind 1: Montgomery products (one per step per thread). *work_per_ms = MAC32 (kind 0) or products (kind 1). */
int drx_microbench(drx_ctx* ctx, int kind, int iters, double* ms, double* work);

/* ------------------------------------------------------------------------------------------------------------
 * The same ElGamal operations over edwards25519, UnLynx's default suite (libunlynx.SuiTe = Ed25519 in
 * github.com/ldsec/unlynx v1.4.0 lib/crypto.go; kyber v3.0.8 group/edwards25519 -- go.mod:14,26, not under
 * /root/reference). Drynx itself fixes bn256 (lib/suite.go:10), so these replace the UnLynx functions as UnLynx's own
 * protocols call them; argument meaning is that of the entry points above. Wire forms: point = 32 bytes (y little-
 * endian, bit 255 = parity of x, RFC 8032 5.1.2); scalar = 32 bytes little-endian, reduced mod l on input;
 * ciphertext = K || C = 64 bytes. Decoding follows ref10 ge_frombytes: y >= p and "negative zero" are accepted,
 * an encoding with no x on the curve is an error (-2). */
int drx_ed_set_pubkey(drx_ctx* ctx, const uint8_t P[32]);
int drx_ed_encrypt(drx_ctx* ctx, size_t n, const int64_t* m, const uint8_t* r /*n*32*/, uint8_t* ct /*n*64*/);
int drx_ed_encrypt_dev(drx_ctx* ctx, size_t n, const int64_t* d_m, const uint8_t* d_r, uint8_t* d_ct);
int drx_ed_aggregate(drx_ctx* ctx, size_t ndp, size_t v, const uint8_t* const* cts /*ndp x (v*64)*/, uint8_t* out /*v*64*/);
int drx_ed_aggregate_dev(drx_ctx* ctx, size_t ndp, size_t v, const uint8_t* d_cts /*[ndp][v][64]*/, uint8_t* d_out);
int drx_ed_dlog_table(drx_ctx* ctx, int64_t limit);
int drx_ed_decrypt(drx_ctx* ctx, size_t n, const uint8_t sk[32], const uint8_t* ct /*n*64*/, int with_neg, int64_t* out,
                   uint8_t* found);
/* raw discrete logs on the device: d_out[i] in (-100000, 100000), or INT64_MIN when outside or malformed */
int drx_ed_decrypt_dev(drx_ctx* ctx, size_t n, const uint8_t sk[32], const uint8_t* d_ct, int64_t* d_out);
/* Diagnostics: out[i] = k[i] * pt[i]. */
int drx_ed_point_mul(drx_ctx* ctx, size_t n, const uint8_t* k /*n*32*/, const uint8_t* pt /*n*32*/, uint8_t* out /*n*32*/);

#ifdef __cplusplus
}
#endif
#endif
