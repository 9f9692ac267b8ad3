// C ABI of libdrynx_gpu.so (include/drynx_gpu.h): context, device memory, host<->device staging and the
// host-side bookkeeping that mirrors state the reference keeps in Go globals (the lazily grown PointToInt map
// of libunlynx v1.4.0 lib/crypto.go discreteLog). All arithmetic happens in the kernels; there is no CPU
// implementation behind any entry point.
#include <cuda_runtime.h>

#include <cstdarg>
#include <cstdio>
#include <cstdlib>
#include <cstring>
#include <map>
#include <mutex>
#include <string>
#include <utility>
#include <vector>

#include "../../include/drynx_gpu.h"
#include "launch.h"

using namespace drx;

namespace {

thread_local std::string g_err = "";

int fail(int code, const char* fmt, ...) {
  char buf[512];
  va_list ap;
  va_start(ap, fmt);
  vsnprintf(buf, sizeof buf, fmt, ap);
  va_end(ap);
  g_err = buf;
  return code;
}

#define CU(x)                                                                                     \
  do {                                                                                            \
    cudaError_t e_ = (x);                                                                         \
    if (e_ != cudaSuccess) return fail((int)e_, "%s: %s (%s:%d)", #x, cudaGetErrorString(e_), __FILE__, __LINE__); \
  } while (0)

// default generators in wire form: B = (1, -2); g2 = the x/crypto bn256 twist generator (x.i, x.1, y.i, y.1)
const uint8_t kDefaultG1[64] = {
    0, 0, 0, 0, 0, 0, 0, 0, 0, 0, 0, 0, 0, 0, 0, 0, 0, 0, 0, 0, 0, 0, 0, 0, 0, 0, 0, 0, 0, 0, 0, 1,
    0x8f, 0xb5, 0x01, 0xe3, 0x4a, 0xa3, 0x87, 0xf9, 0xaa, 0x6f, 0xec, 0xb8, 0x61, 0x84, 0xdc, 0x21,
    0xee, 0x5b, 0x88, 0xd1, 0x20, 0xb5, 0xb5, 0x9e, 0x18, 0x5c, 0xac, 0x6c, 0x5e, 0x08, 0x96, 0x65};
#include "g2_default.inc"

struct DevBuf {
  void* p = nullptr;
  size_t cap = 0;
  int ensure(size_t bytes) {
    if (bytes <= cap) return 0;
    if (p) cudaFree(p);
    p = nullptr;
    cap = 0;
    size_t want = bytes + bytes / 4 + 4096;
    cudaError_t e = cudaMalloc(&p, want);
    if (e != cudaSuccess) return fail((int)e, "cudaMalloc(%zu): %s", want, cudaGetErrorString(e));
    cap = want;
    return 0;
  }
  void release() {
    if (p) cudaFree(p);
    p = nullptr;
    cap = 0;
  }
};

// Carves 256-byte aligned regions out of one growable device allocation.
struct Arena {
  DevBuf buf;
  size_t used = 0;
  std::vector<size_t> sizes;
  void plan_reset() { used = 0; sizes.clear(); }
  size_t plan(size_t bytes) {
    size_t off = used;
    used += (bytes + 255) & ~(size_t)255;
    return off;
  }
  int commit() { return buf.ensure(used ? used : 256); }
  template <class T>
  T* at(size_t off) { return (T*)((char*)buf.p + off); }
};

// Host model of which integers the reference's lazily grown PointToInt map holds (libunlynx discreteLog).
struct DlogModel {
  int64_t pos_hi = -1;                                // map holds 0..pos_hi
  std::vector<std::pair<int64_t, int64_t>> neg;       // inclusive ranges of i with -i stored
  bool has_neg(int64_t i) const {
    for (auto& r : neg)
      if (i >= r.first && i <= r.second) return true;
    return false;
  }
  void add_neg(int64_t lo, int64_t hi) {
    if (!neg.empty() && neg.back().second + 1 >= lo && neg.back().first <= lo) {
      if (hi > neg.back().second) neg.back().second = hi;
    } else {
      neg.push_back({lo, hi});
    }
  }
  // raw: true discrete log in (-100000, 100000) or INT64_MIN when outside that range
  int64_t resolve(int64_t raw, bool check_neg, bool* found) {
    const bool known = raw != INT64_MIN;
    if (known && raw >= 0 && raw <= pos_hi) { *found = true; return raw; }
    if (known && raw < 0 && has_neg(-raw)) { *found = true; return raw; }
    const int64_t g0 = pos_hi < 0 ? 0 : pos_hi;
    int64_t stop = kMaxHomomorphicInt - 1;
    bool hit = false;
    if (known && raw >= g0) { stop = raw; hit = true; }
    else if (known && raw < 0 && check_neg && -raw >= g0) { stop = -raw; hit = true; }
    pos_hi = stop;
    if (check_neg) add_neg(g0, stop);
    *found = hit;
    return hit ? raw : 0;
  }
};

// Optional per-kernel timing: CUDA event pairs recorded on the launching stream around selected kernels.
struct ProfSpan {
  std::string name;
  cudaEvent_t a, b;
};
struct Profiler {
  int on = 0;  // 0 off, 1 every kernel, 2 only the pairing kernel of verification (k_pair_ucode / k_rp_verify_items)
  std::vector<ProfSpan> spans;
  std::map<std::string, std::pair<double, uint64_t>> totals;  // name -> (ms, launches)
};

struct RpSetHost {
  int ns = 0;
  int64_t u = 0;
  DevBuf tabY, tabA, tabGtA, sumY;
};

}  // namespace

struct drx_ctx {
  int device = 0;
  cudaStream_t own_stream = nullptr, stream = nullptr, side_stream = nullptr, side_stream2 = nullptr;
  // creation's two short kernels (header, G2 part) run on high-priority streams beside the GT kernel on `stream`, see run_create
  cudaStream_t hi_stream = nullptr, hi_stream2 = nullptr;
  cudaEvent_t ev_fork = nullptr, ev_fork0 = nullptr, ev_join = nullptr, ev_join2 = nullptr;  // fork/join of the side streams (independent small kernels)
  std::mutex mu;
  uint64_t launches = 0;
  bool have_pub = false, have_dlog = false, have_target = false;
  int opt_verify_path = 0;  // drx_set_option "rp_verify_path"
  int opt_create_path = 0;  // drx_set_option "rp_create_path": 0 = by batch size, 1 / 2 / 3 force one of the small-batch kernels
  int opt_ucode_lanes = 0;  // drx_set_option "rp_ucode_lanes": 0 = by batch size, 4 or 8
  int opt_verify_gt = 0;    // drx_set_option "rp_verify_gt": 0 = by batch size, 2 = always the lane-group GT kernel (tests)
  int opt_hdr_early = 0;    // drx_set_option "rp_verify_hdr_early" (experiments): 1 = the header check starts beside the preparation, 0 = with the pairing kernel
  int opt_agg_slices = 0;   // drx_set_option "agg_slices": 0 = sized from the SM count, else forced (experiments)
  UcTune uc_tune;           // drx_set_option "rp_ucode_mode" / "rp_ucode_resident" / "rp_ucode_nseg"
  DevBuf baseB_aff, baseB_wire, g2_aff, g2_wire, gtB, tabB, tabP, pub_aff, tabGtB, tabG2, dlog_keys, dlog_vals, bad,
      sets_dev, sink, tabQ, target_aff;  // tabQ: comb table of the key-switching target key
  DevBuf tabGtB16;  // 16-bit windows of e(B, g2) for the lane-group GT kernels, built at their first use (403 MB)
  Arena arena;
  Profiler prof;
  DlogModel dlog;
  std::map<int, RpSetHost> sets;
  int sets_dev_count = 0;
  // edwards25519 suite (capi_ed.inc): its own base table, key table and discrete-log table
  bool ed_have_base = false, ed_have_pub = false, ed_have_dlog = false;
  DevBuf ed_pt, ed_tabB, ed_tabP, ed_dlog_keys, ed_dlog_vals;
  DlogModel ed_dlog;
};

namespace {

struct Guard {
  drx_ctx* c;
  explicit Guard(drx_ctx* ctx) : c(ctx) { c->mu.lock(); cudaSetDevice(c->device); }
  ~Guard() { c->mu.unlock(); }
};

#define RET(x)            \
  do {                    \
    int r_ = (x);         \
    if (r_) return r_;    \
  } while (0)

// Transfers of the host-buffer entry points. Callers hand over pageable memory (Go slices, numpy arrays,
// std::vector), which cudaMemcpyAsync moves through the driver's own pinned bounce buffers, chunk by chunk with the
// copy and the DMA overlapped. A staging area inside the context was built and measured in round 2 and lost to that
// path (config-2 step from pageable numpy arrays: 7.21 ms with the driver's path, 7.66 ms staged with one host thread,
// 8.26 ms staged with four; from pinned memory 6.5-6.9 ms either way), so the wrappers below copy directly; they
// remain the one place where a different transfer policy would go.
int h2d(drx_ctx* c, void* dst, const void* src, size_t n) {
  if (!n) return 0;
  cudaError_t e = cudaMemcpyAsync(dst, src, n, cudaMemcpyHostToDevice, c->stream);
  if (e != cudaSuccess) return fail((int)e, "cudaMemcpyAsync (to device): %s", cudaGetErrorString(e));
  return 0;
}
int d2h(drx_ctx* c, void* dst, const void* src, size_t n) {
  if (!n) return 0;
  cudaError_t e = cudaMemcpyAsync(dst, src, n, cudaMemcpyDeviceToHost, c->stream);
  if (e != cudaSuccess) return fail((int)e, "cudaMemcpyAsync (to host): %s", cudaGetErrorString(e));
  return 0;
}
int sync_out(drx_ctx* c) {
  cudaError_t e = cudaStreamSynchronize(c->stream);
  if (e != cudaSuccess) return fail((int)e, "cudaStreamSynchronize: %s", cudaGetErrorString(e));
  return 0;
}

// Device-buffer (`_dev`) calls are asynchronous and cannot return a malformed-input count; they add to a second
// counter (word 1 of the buffer) that drx_take_bad reads and clears. Host-buffer calls use word 0, reset per call.
int* dev_bad(drx_ctx* c) { return (int*)c->bad.p + 1; }
int reset_bad(drx_ctx* c) {
  CU(cudaMemsetAsync(c->bad.p, 0, sizeof(int), c->stream));
  return 0;
}
int read_bad(drx_ctx* c, int* out) {
  RET(d2h(c, out, c->bad.p, sizeof(int)));
  RET(sync_out(c));
  return 0;
}
int check_launch(const char* what) {
  cudaError_t e = cudaGetLastError();
  if (e != cudaSuccess) return fail((int)e, "%s launch: %s", what, cudaGetErrorString(e));
  return 0;
}

// Brackets the kernels `fn` launches with an event pair when profiling is on; always counts launches.
template <class Fn>
void timed_on(drx_ctx* c, cudaStream_t st, const char* name, Fn fn) {
  if (!c->prof.on || (c->prof.on == 2 && strcmp(name, "k_pair_ucode") != 0 && strcmp(name, "k_rp_verify_items") != 0)) {
    c->launches += fn();
    return;
  }
  ProfSpan sp;
  sp.name = name;
  cudaEventCreate(&sp.a);
  cudaEventCreate(&sp.b);
  cudaEventRecord(sp.a, st);
  c->launches += fn();
  cudaEventRecord(sp.b, st);
  c->prof.spans.push_back(sp);
}
template <class Fn>
void timed(drx_ctx* c, const char* name, Fn fn) {
  timed_on(c, c->stream, name, fn);
}
void prof_collect(drx_ctx* c) {
  // DRX_PROFILE_TIMELINE=1: every span's start (relative to the first span collected with it) and duration to stderr
  static const bool timeline = getenv("DRX_PROFILE_TIMELINE") != nullptr;
  if (timeline)
    for (auto& sp : c->prof.spans) {
      cudaEventSynchronize(sp.b);
      float at = 0, ms = 0;
      cudaEventElapsedTime(&at, c->prof.spans.front().a, sp.a);
      cudaEventElapsedTime(&ms, sp.a, sp.b);
      fprintf(stderr, "[drx timeline %p] %-22s start %9.3f ms  duration %8.3f ms\n", (void*)c, sp.name.c_str(), at, ms);
    }
  for (auto& sp : c->prof.spans) {
    cudaEventSynchronize(sp.b);
    float ms = 0;
    cudaEventElapsedTime(&ms, sp.a, sp.b);
    auto& t = c->prof.totals[sp.name];
    t.first += ms;
    t.second += 1;
    cudaEventDestroy(sp.a);
    cudaEventDestroy(sp.b);
  }
  c->prof.spans.clear();
}

int ensure_dlog(drx_ctx* c) {
  if (c->have_dlog) return 0;
  RET(c->dlog_keys.ensure((size_t)kDlogCap * 32));
  RET(c->dlog_vals.ensure((size_t)kDlogCap * 4));
  c->launches += launch_dlog_build(c->stream, c->tabB.p, c->baseB_aff.p, c->dlog_keys.p, (uint32_t*)c->dlog_vals.p);
  RET(check_launch("dlog_build"));
  c->have_dlog = true;
  return 0;
}

}  // namespace

extern "C" {

const char* drx_last_error(void) { return g_err.c_str(); }

int drx_init(drx_ctx** out, int device, const uint8_t* g1_base, const uint8_t* g2_base) {
  if (!out) return fail(-1, "drx_init: out is NULL");
  *out = nullptr;
  int ndev = 0;
  cudaError_t e = cudaGetDeviceCount(&ndev);
  if (e != cudaSuccess || ndev == 0)
    return fail(e != cudaSuccess ? (int)e : -3, "drx_init: no CUDA device (%s); this library has no CPU path",
                cudaGetErrorString(e));
  if (device < 0 || device >= ndev) return fail(-1, "drx_init: device %d out of range (0..%d)", device, ndev - 1);
  cudaDeviceProp prop;
  CU(cudaGetDeviceProperties(&prop, device));
  if (prop.major != 10)
    return fail(-3, "drx_init: device %d is sm_%d%d; libdrynx_gpu is built for sm_100a only", device, prop.major,
                prop.minor);
  CU(cudaSetDevice(device));
  drx_ctx* c = new drx_ctx();
  c->device = device;
  e = cudaStreamCreateWithFlags(&c->own_stream, cudaStreamNonBlocking);
  if (e != cudaSuccess) { delete c; return fail((int)e, "cudaStreamCreate: %s", cudaGetErrorString(e)); }
  c->stream = c->own_stream;
  int prio_least = 0, prio_greatest = 0;
  cudaDeviceGetStreamPriorityRange(&prio_least, &prio_greatest);
  if (cudaStreamCreateWithFlags(&c->side_stream, cudaStreamNonBlocking) != cudaSuccess ||
      cudaStreamCreateWithFlags(&c->side_stream2, cudaStreamNonBlocking) != cudaSuccess ||
      cudaStreamCreateWithPriority(&c->hi_stream, cudaStreamNonBlocking, prio_greatest) != cudaSuccess ||
      cudaStreamCreateWithPriority(&c->hi_stream2, cudaStreamNonBlocking, prio_greatest) != cudaSuccess ||
      cudaEventCreateWithFlags(&c->ev_join2, cudaEventDisableTiming) != cudaSuccess ||
      cudaEventCreate(&c->ev_fork) != cudaSuccess ||  // a TIMING event on purpose, see run_verify (capi_rp.inc)
      cudaEventCreateWithFlags(&c->ev_fork0, cudaEventDisableTiming) != cudaSuccess ||
      cudaEventCreateWithFlags(&c->ev_join, cudaEventDisableTiming) != cudaSuccess) {
    drx_free(c);
    return fail(-3, "drx_init: cannot create side stream");
  }
  int rc = 0;
  auto cleanup = [&](int code) { drx_free(c); return code; };
  if ((rc = c->bad.ensure(256))) return cleanup(rc);
  if ((rc = c->baseB_wire.ensure(64)) || (rc = c->baseB_aff.ensure(64)) || (rc = c->g2_wire.ensure(128)) ||
      (rc = c->g2_aff.ensure(128)) || (rc = c->gtB.ensure(kGtBytes)) || (rc = c->tabB.ensure(kCombPoints * 64)) ||
      (rc = c->tabGtB.ensure(kCombPoints * kGtBytes)) || (rc = c->tabG2.ensure(kCombPoints * 128)))
    return cleanup(rc);
  cudaMemcpyAsync(c->baseB_wire.p, g1_base ? g1_base : kDefaultG1, 64, cudaMemcpyHostToDevice, c->stream);
  cudaMemcpyAsync(c->g2_wire.p, g2_base ? g2_base : kDefaultG2, 128, cudaMemcpyHostToDevice, c->stream);
  cudaMemsetAsync(c->bad.p, 0, 256, c->stream);  // word 0: malformed-input counter; bytes 64..95: key-switching secret; word 48: lane-group pairing error word
  c->launches += launch_g1_from_wire(c->stream, 1, (const uint8_t*)c->baseB_wire.p, c->baseB_aff.p, (int*)c->bad.p);
  c->launches += launch_g2_from_wire(c->stream, 1, (const uint8_t*)c->g2_wire.p, c->g2_aff.p, (int*)c->bad.p);
  c->launches += launch_comb_table_g1(c->stream, 1, c->baseB_aff.p, c->tabB.p);
  c->launches += launch_comb_table_g2(c->stream, 1, c->g2_aff.p, c->tabG2.p);
  c->launches += launch_pair_aff(c->stream, 1, c->baseB_aff.p, 0, c->g2_aff.p, c->gtB.p);
  c->launches += launch_comb_table_gt(c->stream, 1, c->gtB.p, c->tabGtB.p);
  if ((rc = check_launch("init"))) return cleanup(rc);
  int bad = 0;
  if ((rc = read_bad(c, &bad))) return cleanup(rc);
  if (bad) return cleanup(fail(-2, "drx_init: generator bytes are not valid curve points"));
  *out = c;
  return 0;
}

void drx_free(drx_ctx* c) {
  if (!c) return;
  cudaSetDevice(c->device);
  cudaDeviceSynchronize();
  for (DevBuf* b : {&c->baseB_aff, &c->baseB_wire, &c->g2_aff, &c->g2_wire, &c->gtB, &c->tabB, &c->tabP, &c->pub_aff,
                    &c->tabGtB, &c->tabG2, &c->dlog_keys, &c->dlog_vals, &c->bad, &c->sets_dev, &c->sink, &c->tabQ,
                    &c->target_aff, &c->tabGtB16, &c->ed_pt,
                    &c->ed_tabB, &c->ed_tabP, &c->ed_dlog_keys, &c->ed_dlog_vals})
    b->release();
  c->arena.buf.release();
  for (auto& kv : c->sets) {
    kv.second.tabY.release();
    kv.second.tabA.release();
    kv.second.tabGtA.release();
    kv.second.sumY.release();
  }
  if (c->own_stream) cudaStreamDestroy(c->own_stream);
  if (c->side_stream) cudaStreamDestroy(c->side_stream);
  if (c->side_stream2) cudaStreamDestroy(c->side_stream2);
  if (c->hi_stream) cudaStreamDestroy(c->hi_stream);
  if (c->hi_stream2) cudaStreamDestroy(c->hi_stream2);
  if (c->ev_join2) cudaEventDestroy(c->ev_join2);
  if (c->ev_fork) cudaEventDestroy(c->ev_fork);
  if (c->ev_fork0) cudaEventDestroy(c->ev_fork0);
  if (c->ev_join) cudaEventDestroy(c->ev_join);
  delete c;
}

int drx_set_stream(drx_ctx* c, void* cuda_stream) {
  if (!c) return fail(-1, "null context");
  Guard g(c);
  c->stream = cuda_stream ? (cudaStream_t)cuda_stream : c->own_stream;
  return 0;
}
// Reads and clears the word a task-queue launch of the lane-group pairing raises when one of its warps gave up
// waiting for a predecessor task (ucode_common.cuh): the verification results of that call are not valid then.
static int take_pairing_error(drx_ctx* c) {
  unsigned int err = 0;
  unsigned int* d_err = (unsigned int*)c->bad.p + 48;  // kUcErrWord (capi_rp.inc); bytes 64..95 of the buffer hold the key-switching secret
  RET(d2h(c, &err, d_err, sizeof(err)));
  RET(sync_out(c));
  if (!err) return 0;
  CU(cudaMemsetAsync(d_err, 0, sizeof(err), c->stream));
  return fail(-5, "range-proof verification: the pairing kernel abandoned a wait on the device (results invalid; retry the call)");
}
int drx_synchronize(drx_ctx* c) {
  if (!c) return fail(-1, "null context");
  Guard g(c);
  RET(sync_out(c));
  return take_pairing_error(c);
}
int drx_take_bad(drx_ctx* c, int* n_bad) {
  if (!c || !n_bad) return fail(-1, "null argument");
  Guard g(c);
  RET(d2h(c, n_bad, dev_bad(c), sizeof(int)));
  CU(cudaMemsetAsync(dev_bad(c), 0, sizeof(int), c->stream));
  RET(sync_out(c));
  return 0;
}
uint64_t drx_launch_count(const drx_ctx* c) { return c ? c->launches : 0; }

int drx_set_option(drx_ctx* c, const char* name, int value) {
  if (!c || !name) return fail(-1, "null argument");
  Guard g(c);
  if (!strcmp(name, "rp_verify_path")) {
    if (value < 0 || value > 2) return fail(-1, "rp_verify_path must be 0 (auto), 1 (thread per item) or 2 (lane groups)");
    c->opt_verify_path = value;
    return 0;
  }
  if (!strcmp(name, "rp_create_path")) {
    if (value < 0 || value > 3)
      return fail(-1, "rp_create_path must be 0 (auto), 1 (scratch-based split), 2 (shuffle-cooperative) or 3 (GT factors on lane groups)");
    c->opt_create_path = value;
    return 0;
  }
  if (!strcmp(name, "rp_ucode_lanes")) {
    if (value != 0 && value != 4 && value != 8) return fail(-1, "rp_ucode_lanes must be 0 (auto), 4 or 8");
    c->opt_ucode_lanes = value;
    return 0;
  }
  if (!strcmp(name, "rp_verify_gt")) {
    if (value != 0 && value != 2) return fail(-1, "rp_verify_gt must be 0 (auto) or 2 (lane groups)");
    c->opt_verify_gt = value;
    return 0;
  }
  if (!strcmp(name, "agg_slices")) {
    if (value < 0 || value > kAggMaxSlices) return fail(-1, "agg_slices must be 0 (auto) .. %d", kAggMaxSlices);
    c->opt_agg_slices = value;
    return 0;
  }
  if (!strcmp(name, "rp_ucode_mode")) {
    if (value < 0 || value > 2) return fail(-1, "rp_ucode_mode must be 0 (auto), 1 (static launch) or 2 (task queue)");
    c->uc_tune.mode = value;
    return 0;
  }
  if (!strcmp(name, "rp_ucode_resident")) {
    if (value < 0 || value > 32) return fail(-1, "rp_ucode_resident must be 0 (auto) .. 32 blocks per SM");
    c->uc_tune.resident = value;
    return 0;
  }
  if (!strcmp(name, "rp_verify_hdr_early")) {
    if (value != 0 && value != 1) return fail(-1, "rp_verify_hdr_early must be 0 or 1");
    c->opt_hdr_early = value;
    return 0;
  }
  if (!strcmp(name, "rp_ucode_nseg")) {
    if (value < 0 || value > 1024) return fail(-1, "rp_ucode_nseg must be 0 (auto) .. 1024");
    c->uc_tune.nseg = value;
    return 0;
  }
  return fail(-1, "unknown option %s", name);
}

int drx_profile_enable(drx_ctx* c, int on) {
  if (!c) return fail(-1, "null context");
  Guard g(c);
  prof_collect(c);
  c->prof.totals.clear();
  c->prof.on = on == 2 ? 2 : on != 0;
  return 0;
}
int drx_profile_get(drx_ctx* c, const char* kernel, double* total_ms, uint64_t* launches) {
  if (!c || !kernel || !total_ms || !launches) return fail(-1, "null argument");
  Guard g(c);
  prof_collect(c);
  auto it = c->prof.totals.find(kernel);
  *total_ms = it == c->prof.totals.end() ? 0.0 : it->second.first;
  *launches = it == c->prof.totals.end() ? 0 : it->second.second;
  return 0;
}

int drx_reduce_rows_dev(drx_ctx* c, size_t ndp, size_t nrows, const int64_t* d_rows, int mode, int64_t* d_out) {
  if (!c) return fail(-1, "null context");
  if (mode < 0 || mode > 3) return fail(-1, "drx_reduce_rows_dev: mode %d", mode);
  Guard g(c);
  timed(c, "k_reduce_rows", [&] { return launch_reduce_rows(c->stream, ndp, nrows, d_rows, mode, d_out); });
  return check_launch("reduce_rows");
}

int drx_reduce_rows(drx_ctx* c, size_t ndp, size_t nrows, const int64_t* rows, int mode, int64_t* out) {
  if (!c) return fail(-1, "null context");
  if (mode < 0 || mode > 3) return fail(-1, "drx_reduce_rows: mode %d", mode);
  if (ndp == 0) return 0;
  if (!rows || !out) return fail(-1, "drx_reduce_rows: null buffer");
  Guard g(c);
  static const size_t width[4] = {1, 2, 3, 5};
  const size_t w = width[mode];
  Arena& a = c->arena;
  a.plan_reset();
  size_t oi = a.plan(ndp * nrows * 8), oo = a.plan(ndp * w * 8);
  RET(a.commit());
  if (nrows) RET(h2d(c, a.at<void>(oi), rows, ndp * nrows * 8));
  timed(c, "k_reduce_rows", [&] { return launch_reduce_rows(c->stream, ndp, nrows, a.at<int64_t>(oi), mode, a.at<int64_t>(oo)); });
  RET(check_launch("reduce_rows"));
  RET(d2h(c, out, a.at<void>(oo), ndp * w * 8));
  RET(sync_out(c));
  return 0;
}

int drx_freq_count(drx_ctx* c, size_t ndp, size_t nrows, const int64_t* rows, int64_t qmin, int64_t qmax, int64_t* out) {
  if (!c) return fail(-1, "null context");
  if (qmax < qmin || qmax - qmin + 1 > kFreqCountMaxBins)
    return fail(-1, "drx_freq_count: range [%lld, %lld] must hold 1..%d values", (long long)qmin, (long long)qmax, kFreqCountMaxBins);
  if (ndp == 0) return 0;
  if (!out || (nrows && !rows)) return fail(-1, "drx_freq_count: null buffer");
  Guard g(c);
  const int nbins = (int)(qmax - qmin + 1);
  Arena& a = c->arena;
  a.plan_reset();
  size_t oi = a.plan(ndp * nrows * 8), oo = a.plan(ndp * (size_t)nbins * 8);
  RET(a.commit());
  RET(reset_bad(c));
  if (nrows) RET(h2d(c, a.at<void>(oi), rows, ndp * nrows * 8));
  timed(c, "k_freq_count", [&] {
    return launch_freq_count(c->stream, ndp, nrows, a.at<int64_t>(oi), qmin, nbins, a.at<int64_t>(oo), (int*)c->bad.p);
  });
  RET(check_launch("freq_count"));
  RET(d2h(c, out, a.at<void>(oo), ndp * (size_t)nbins * 8));
  int bad = 0;
  RET(read_bad(c, &bad));
  if (bad) return fail(-2, "drx_freq_count: %d values outside [%lld, %lld]", bad, (long long)qmin, (long long)qmax);
  return 0;
}

int drx_logreg_coefficients(drx_ctx* c, size_t ndp, const size_t* rec_off, int d, int k, const double* x, const int64_t* y,
                            const double* means, const double* sds, double precision, int64_t* out) {
  if (!c) return fail(-1, "null context");
  if (ndp == 0) return 0;
  if (!rec_off || !means || !sds || !out) return fail(-1, "drx_logreg_coefficients: null buffer");
  if (d < 1 || d > kLogRegMaxFeatures) return fail(-1, "drx_logreg_coefficients: d = %d outside 1..%d", d, kLogRegMaxFeatures);
  if (k < 1 || k > 2) return fail(-1, "drx_logreg_coefficients: k = %d (the reference's coefficient packing is only defined for k <= 2)", k);
  for (size_t i = 0; i < ndp; i++)
    if (rec_off[i + 1] < rec_off[i]) return fail(-1, "drx_logreg_coefficients: record offsets must not decrease");
  const size_t nrec = rec_off[ndp] - rec_off[0], d1 = (size_t)d + 1, ncoef = d1 + (k >= 2 ? d1 * d1 : 0);
  if (nrec && (!x || !y)) return fail(-1, "drx_logreg_coefficients: null records");
  Guard g(c);
  Arena& a = c->arena;
  a.plan_reset();
  size_t oo = a.plan((ndp + 1) * 8), ox = a.plan(nrec * d * 8), oy = a.plan(nrec * 8), om = a.plan(d * 8), os = a.plan(d * 8),
         oc = a.plan(ndp * ncoef * 8), oxs = a.plan(nrec * d * 8);
  RET(a.commit());
  std::vector<size_t> off(rec_off, rec_off + ndp + 1);
  for (auto& v : off) v -= rec_off[0];
  cudaStream_t st = c->stream;
  RET(h2d(c, a.at<void>(oo), off.data(), (ndp + 1) * 8));
  if (nrec) {
    RET(h2d(c, a.at<void>(ox), x + rec_off[0] * d, nrec * d * 8));
    RET(h2d(c, a.at<void>(oy), y + rec_off[0], nrec * 8));
  }
  RET(h2d(c, a.at<void>(om), means, d * 8));
  RET(h2d(c, a.at<void>(os), sds, d * 8));
  RET(sync_out(c));  // `off` is a stack-owned staging vector
  timed(c, "k_logreg_coeffs", [&] {
    return launch_logreg_coeffs(st, ndp, nrec, a.at<size_t>(oo), d, k, a.at<double>(ox), a.at<int64_t>(oy), a.at<double>(om),
                                a.at<double>(os), precision, a.at<double>(oxs), a.at<int64_t>(oc));
  });
  RET(check_launch("logreg_coefficients"));
  RET(d2h(c, out, a.at<void>(oc), ndp * ncoef * 8));
  RET(sync_out(c));
  return 0;
}

int drx_logreg_coefficients_dev(drx_ctx* c, size_t ndp, size_t nrec, const size_t* d_rec_off, int d, int k,
                                const double* d_x, const int64_t* d_y, const double* d_means, const double* d_sds,
                                double precision, int64_t* d_out) {
  if (!c) return fail(-1, "null context");
  if (d < 1 || d > kLogRegMaxFeatures) return fail(-1, "drx_logreg_coefficients_dev: d = %d outside 1..%d", d, kLogRegMaxFeatures);
  if (k < 1 || k > 2) return fail(-1, "drx_logreg_coefficients_dev: k = %d (only k <= 2)", k);
  Guard g(c);
  Arena& a = c->arena;
  a.plan_reset();
  size_t oxs = a.plan(nrec * (size_t)d * 8);  // the standardised records
  RET(a.commit());
  timed(c, "k_logreg_coeffs", [&] {
    return launch_logreg_coeffs(c->stream, ndp, nrec, d_rec_off, d, k, d_x, d_y, d_means, d_sds, precision, a.at<double>(oxs),
                                d_out);
  });
  return check_launch("logreg_coefficients");
}

int drx_set_pubkey(drx_ctx* c, const uint8_t P[64]) {
  if (!c || !P) return fail(-1, "drx_set_pubkey: null argument");
  Guard g(c);
  RET(c->pub_aff.ensure(64));
  RET(c->tabP.ensure(kCombPoints * 64));
  Arena& a = c->arena;
  a.plan_reset();
  size_t o = a.plan(64);
  RET(a.commit());
  RET(reset_bad(c));
  RET(h2d(c, a.at<uint8_t>(o), P, 64));
  c->launches += launch_g1_from_wire(c->stream, 1, a.at<uint8_t>(o), c->pub_aff.p, (int*)c->bad.p);
  c->launches += launch_comb_table_g1(c->stream, 1, c->pub_aff.p, c->tabP.p);
  RET(check_launch("set_pubkey"));
  int bad = 0;
  RET(read_bad(c, &bad));
  if (bad) { c->have_pub = false; return fail(-2, "drx_set_pubkey: malformed point"); }
  c->have_pub = true;
  return 0;
}

int drx_encrypt_dev(drx_ctx* c, size_t n, const int64_t* d_m, const uint8_t* d_r, uint8_t* d_ct) {
  if (!c) return fail(-1, "null context");
  Guard g(c);
  if (!c->have_pub) return fail(-1, "drx_encrypt: call drx_set_pubkey first");
  timed(c, "k_encrypt", [&] { return launch_encrypt(c->stream, n, d_m, d_r, c->tabB.p, c->tabP.p, d_ct); });
  return check_launch("encrypt");
}

int drx_encrypt(drx_ctx* c, size_t n, const int64_t* m, const uint8_t* r, uint8_t* ct) {
  if (!c) return fail(-1, "null context");
  if (n == 0) return 0;
  if (!m || !r || !ct) return fail(-1, "drx_encrypt: null buffer");
  Guard g(c);
  if (!c->have_pub) return fail(-1, "drx_encrypt: call drx_set_pubkey first");
  Arena& a = c->arena;
  a.plan_reset();
  size_t om = a.plan(n * 8), orr = a.plan(n * 32), oc = a.plan(n * 128);
  RET(a.commit());
  RET(h2d(c, a.at<void>(om), m, n * 8));
  RET(h2d(c, a.at<void>(orr), r, n * 32));
  timed(c, "k_encrypt", [&] {
    return launch_encrypt(c->stream, n, a.at<int64_t>(om), a.at<uint8_t>(orr), c->tabB.p, c->tabP.p, a.at<uint8_t>(oc));
  });
  RET(check_launch("encrypt"));
  RET(d2h(c, ct, a.at<void>(oc), n * 128));
  RET(sync_out(c));
  return 0;
}

int drx_encrypt_scalar(drx_ctx* c, size_t n, const uint8_t* m, const uint8_t* r, uint8_t* ct) {
  if (!c) return fail(-1, "null context");
  if (n == 0) return 0;
  if (!m || !r || !ct) return fail(-1, "drx_encrypt_scalar: null buffer");
  Guard g(c);
  if (!c->have_pub) return fail(-1, "drx_encrypt_scalar: call drx_set_pubkey first");
  Arena& a = c->arena;
  a.plan_reset();
  size_t om = a.plan(n * 32), orr = a.plan(n * 32), oc = a.plan(n * 128);
  RET(a.commit());
  RET(h2d(c, a.at<void>(om), m, n * 32));
  RET(h2d(c, a.at<void>(orr), r, n * 32));
  timed(c, "k_encrypt", [&] {
    return launch_encrypt_scalar(c->stream, n, a.at<uint8_t>(om), a.at<uint8_t>(orr), c->tabB.p, c->tabP.p, a.at<uint8_t>(oc));
  });
  RET(check_launch("encrypt_scalar"));
  RET(d2h(c, ct, a.at<void>(oc), n * 128));
  RET(sync_out(c));
  return 0;
}

static int aggregate_on_device(drx_ctx* c, size_t ndp, size_t v, const uint8_t* d_cts, void* d_scratch, int slices,
                               uint8_t* d_out, int* d_bad) {
  timed(c, "k_agg_partial", [&] { return launch_agg_partial(c->stream, ndp, v, d_cts, d_scratch, slices, d_bad); });
  timed(c, "k_agg_final", [&] { return launch_agg_final(c->stream, v, slices, d_scratch, d_out); });
  return check_launch("aggregate");
}

int drx_aggregate_dev(drx_ctx* c, size_t ndp, size_t v, const uint8_t* d_cts, uint8_t* d_out) {
  if (!c) return fail(-1, "null context");
  if (v == 0) return 0;
  Guard g(c);
  int slices = 1;
  size_t sb = aggregate_scratch_bytes(ndp, v, &slices);
  if (c->opt_agg_slices > 0 && ndp) {
    slices = c->opt_agg_slices < (int)ndp ? c->opt_agg_slices : (int)ndp;
    sb = (size_t)slices * 2 * v * kG1JacBytes;
  }
  Arena& a = c->arena;
  a.plan_reset();
  size_t os = a.plan(sb);
  RET(a.commit());
  return aggregate_on_device(c, ndp, v, d_cts, a.at<void>(os), slices, d_out, dev_bad(c));
}

int drx_aggregate(drx_ctx* c, size_t ndp, size_t v, const uint8_t* const* cts, uint8_t* out) {
  if (!c) return fail(-1, "null context");
  if (v == 0) return 0;
  if (!out || (ndp && !cts)) return fail(-1, "drx_aggregate: null buffer");
  Guard g(c);
  int slices = 1;
  size_t sb = aggregate_scratch_bytes(ndp, v, &slices);
  Arena& a = c->arena;
  a.plan_reset();
  size_t oi = a.plan(ndp * v * 128), os = a.plan(sb), oo = a.plan(v * 128);
  RET(a.commit());
  RET(reset_bad(c));
  for (size_t d = 0; d < ndp; d++) {
    if (!cts[d]) return fail(-1, "drx_aggregate: cts[%zu] is NULL", d);
    RET(h2d(c, a.at<uint8_t>(oi) + d * v * 128, cts[d], v * 128));
  }
  RET(aggregate_on_device(c, ndp, v, a.at<uint8_t>(oi), a.at<void>(os), slices, a.at<uint8_t>(oo), (int*)c->bad.p));
  RET(d2h(c, out, a.at<void>(oo), v * 128));
  int bad = 0;
  RET(read_bad(c, &bad));
  if (bad) return fail(-2, "drx_aggregate: %d malformed points", bad);
  return 0;
}

int drx_aggregate_noalloc_dev(size_t ndp, size_t v, const uint8_t* d_cts, uint8_t* d_out, void* cuda_stream) {
  if (v == 0) return 0;
  if (!d_out || (ndp && !d_cts)) return fail(-1, "drx_aggregate_noalloc_dev: null buffer");
  int n_bad = 0, launches = 0;
  const int rc = launch_aggregate_static((cudaStream_t)cuda_stream, ndp, v, d_cts, d_out, &n_bad, &launches);
  if (rc == -5)
    return fail(-5, "drx_aggregate_noalloc_dev: v = %zu exceeds the fixed scratch array (at most %zu columns per call)", v,
                aggregate_static_max_columns());
  if (rc == -2) return fail(-2, "drx_aggregate_noalloc_dev: %d malformed points", n_bad);
  if (rc > 0) return fail(rc, "drx_aggregate_noalloc_dev: %s", cudaGetErrorString((cudaError_t)rc));
  return rc;
}

int drx_dlog_table(drx_ctx* c, int64_t limit) {
  if (!c) return fail(-1, "null context");
  if (limit < 0) return fail(-1, "drx_dlog_table: negative limit");
  Guard g(c);
  RET(ensure_dlog(c));
  bool found;
  c->dlog.resolve(limit < kMaxHomomorphicInt ? limit : INT64_MIN, true, &found);
  RET(sync_out(c));
  return 0;
}

// ---- key switching: libunlynxkeyswitch.KeySwitchSequence(targetPubKey, rBs, secretKey)
int drx_set_target_pubkey(drx_ctx* c, const uint8_t Q[64]) {
  if (!c || !Q) return fail(-1, "drx_set_target_pubkey: null argument");
  Guard g(c);
  RET(c->target_aff.ensure(64));
  RET(c->tabQ.ensure(kCombPoints * 64));
  Arena& a = c->arena;
  a.plan_reset();
  size_t o = a.plan(64);
  RET(a.commit());
  RET(reset_bad(c));
  RET(h2d(c, a.at<uint8_t>(o), Q, 64));
  c->launches += launch_g1_from_wire(c->stream, 1, a.at<uint8_t>(o), c->target_aff.p, (int*)c->bad.p);
  c->launches += launch_comb_table_g1(c->stream, 1, c->target_aff.p, c->tabQ.p);
  RET(check_launch("set_target_pubkey"));
  int bad = 0;
  RET(read_bad(c, &bad));
  if (bad) { c->have_target = false; return fail(-2, "drx_set_target_pubkey: malformed point"); }
  c->have_target = true;
  return 0;
}

int drx_keyswitch_dev(drx_ctx* c, size_t n, const uint8_t sk[32], const uint8_t* d_rb, const uint8_t* d_v, uint8_t* d_cv,
                      uint8_t* d_rb_neg) {
  if (!c || !sk) return fail(-1, "null argument");
  Guard g(c);
  if (!c->have_target) return fail(-1, "drx_keyswitch: call drx_set_target_pubkey first");
  // the key travels through the tail of the 256-byte flag buffer
  RET(h2d(c, (uint8_t*)c->bad.p + 64, sk, 32));
  timed(c, "k_keyswitch", [&] {
    return launch_keyswitch(c->stream, n, (uint8_t*)c->bad.p + 64, d_rb, d_v, c->tabB.p, c->tabQ.p, d_cv, d_rb_neg,
                            dev_bad(c));
  });
  return check_launch("keyswitch");
}

int drx_keyswitch(drx_ctx* c, size_t n, const uint8_t sk[32], const uint8_t* rb, const uint8_t* v, uint8_t* cv,
                  uint8_t* rb_neg) {
  if (!c) return fail(-1, "null context");
  if (n == 0) return 0;
  if (!sk || !rb || !v || !cv) return fail(-1, "drx_keyswitch: null buffer");
  Guard g(c);
  if (!c->have_target) return fail(-1, "drx_keyswitch: call drx_set_target_pubkey first");
  Arena& a = c->arena;
  a.plan_reset();
  size_t orb = a.plan(n * 64), ov = a.plan(n * 32), oc = a.plan(n * 128), on = a.plan(rb_neg ? n * 64 : 0);
  RET(a.commit());
  RET(reset_bad(c));
  RET(h2d(c, a.at<void>(orb), rb, n * 64));
  RET(h2d(c, a.at<void>(ov), v, n * 32));
  RET(h2d(c, (uint8_t*)c->bad.p + 64, sk, 32));
  timed(c, "k_keyswitch", [&] {
    return launch_keyswitch(c->stream, n, (uint8_t*)c->bad.p + 64, a.at<uint8_t>(orb), a.at<uint8_t>(ov), c->tabB.p,
                            c->tabQ.p, a.at<uint8_t>(oc), rb_neg ? a.at<uint8_t>(on) : nullptr, (int*)c->bad.p);
  });
  RET(check_launch("keyswitch"));
  RET(d2h(c, cv, a.at<void>(oc), n * 128));
  if (rb_neg) RET(d2h(c, rb_neg, a.at<void>(on), n * 64));
  int bad = 0;
  RET(read_bad(c, &bad));
  if (bad) return fail(-2, "drx_keyswitch: %d malformed points", bad);
  return 0;
}

// ---- obfuscation data path: CipherText.MulCipherTextbyScalar per element (protocols/obfuscation_protocol.go:240-243)
int drx_ct_scalar_mul_dev(drx_ctx* c, size_t n, const uint8_t* d_s, const uint8_t* d_ct, uint8_t* d_out) {
  if (!c) return fail(-1, "null context");
  Guard g(c);
  timed(c, "k_ct_scalar_mul", [&] { return launch_ct_scalar_mul(c->stream, n, d_s, d_ct, d_out, dev_bad(c)); });
  return check_launch("ct_scalar_mul");
}

int drx_ct_scalar_mul(drx_ctx* c, size_t n, const uint8_t* s, const uint8_t* ct, uint8_t* out) {
  if (!c) return fail(-1, "null context");
  if (n == 0) return 0;
  if (!s || !ct || !out) return fail(-1, "drx_ct_scalar_mul: null buffer");
  Guard g(c);
  Arena& a = c->arena;
  a.plan_reset();
  size_t os = a.plan(n * 32), oc = a.plan(n * 128), oo = a.plan(n * 128);
  RET(a.commit());
  RET(reset_bad(c));
  RET(h2d(c, a.at<void>(os), s, n * 32));
  RET(h2d(c, a.at<void>(oc), ct, n * 128));
  timed(c, "k_ct_scalar_mul", [&] {
    return launch_ct_scalar_mul(c->stream, n, a.at<uint8_t>(os), a.at<uint8_t>(oc), a.at<uint8_t>(oo), (int*)c->bad.p);
  });
  RET(check_launch("ct_scalar_mul"));
  RET(d2h(c, out, a.at<void>(oo), n * 128));
  int bad = 0;
  RET(read_bad(c, &bad));
  if (bad) return fail(-2, "drx_ct_scalar_mul: %d malformed points", bad);
  return 0;
}

// ---- shuffling data path: libunlynxshuffle.ShuffleSequence with beta and pi injected
int drx_shuffle_rerandomize(drx_ctx* c, size_t lines, size_t nq, const int32_t* pi, const uint8_t* in, const uint8_t* beta,
                            uint8_t* out) {
  if (!c) return fail(-1, "null context");
  const size_t n = lines * nq;
  if (n == 0) return 0;
  if (!pi || !in || !beta || !out) return fail(-1, "drx_shuffle_rerandomize: null buffer");
  Guard g(c);
  if (!c->have_pub) return fail(-1, "drx_shuffle_rerandomize: call drx_set_pubkey first");
  std::vector<char> seen(lines, 0);
  for (size_t i = 0; i < lines; i++) {
    if (pi[i] < 0 || (size_t)pi[i] >= lines || seen[pi[i]]) return fail(-1, "drx_shuffle_rerandomize: pi is not a permutation");
    seen[pi[i]] = 1;
  }
  Arena& a = c->arena;
  a.plan_reset();
  size_t op = a.plan(lines * 4), oi = a.plan(n * 128), ob = a.plan(n * 32), oo = a.plan(n * 128);
  RET(a.commit());
  RET(reset_bad(c));
  RET(h2d(c, a.at<void>(op), pi, lines * 4));
  RET(h2d(c, a.at<void>(oi), in, n * 128));
  RET(h2d(c, a.at<void>(ob), beta, n * 32));
  timed(c, "k_shuffle_rerandomize", [&] {
    return launch_shuffle_rerandomize(c->stream, lines, nq, a.at<int32_t>(op), a.at<uint8_t>(oi), a.at<uint8_t>(ob),
                                      c->tabB.p, c->tabP.p, a.at<uint8_t>(oo), (int*)c->bad.p);
  });
  RET(check_launch("shuffle_rerandomize"));
  RET(d2h(c, out, a.at<void>(oo), n * 128));
  int bad = 0;
  RET(read_bad(c, &bad));
  if (bad) return fail(-2, "drx_shuffle_rerandomize: %d malformed points", bad);
  return 0;
}
int drx_shuffle_rerandomize_dev(drx_ctx* c, size_t lines, size_t nq, const int32_t* d_pi, const uint8_t* d_in,
                                const uint8_t* d_beta, uint8_t* d_out) {
  if (!c) return fail(-1, "null context");
  Guard g(c);
  if (!c->have_pub) return fail(-1, "drx_shuffle_rerandomize: call drx_set_pubkey first");
  timed(c, "k_shuffle_rerandomize", [&] {
    return launch_shuffle_rerandomize(c->stream, lines, nq, d_pi, d_in, d_beta, c->tabB.p, c->tabP.p, d_out, dev_bad(c));
  });
  return check_launch("shuffle_rerandomize");
}

// ---- Schnorr signatures over G1 (lib/proof/structs_proofs.go:498-505): the caller supplies SHA-512(R || P || msg)
int drx_schnorr_verify(drx_ctx* c, size_t n, const uint8_t* pub, const uint8_t* sig, const uint8_t* digest, uint8_t* ok) {
  if (!c) return fail(-1, "null context");
  if (n == 0) return 0;
  if (!pub || !sig || !digest || !ok) return fail(-1, "drx_schnorr_verify: null buffer");
  Guard g(c);
  Arena& a = c->arena;
  a.plan_reset();
  size_t op = a.plan(n * 64), os = a.plan(n * 96), od = a.plan(n * 64), oo = a.plan(n);
  RET(a.commit());
  RET(h2d(c, a.at<void>(op), pub, n * 64));
  RET(h2d(c, a.at<void>(os), sig, n * 96));
  RET(h2d(c, a.at<void>(od), digest, n * 64));
  timed(c, "k_schnorr_verify", [&] {
    return launch_schnorr_verify(c->stream, n, a.at<uint8_t>(op), a.at<uint8_t>(os), a.at<uint8_t>(od), c->tabB.p,
                                 a.at<uint8_t>(oo));
  });
  RET(check_launch("schnorr_verify"));
  RET(d2h(c, ok, a.at<void>(oo), n));
  RET(sync_out(c));
  return 0;
}

// ---- proof.Rep relations (obfuscation and key-switch proofs): commit, respond, verify
static const int kRepTreeMinTerms = 33;  // drx_rep_commit sums relations at least this wide in two levels
static const int kRepMaxTerms = 4096;  // the widest use is the list compression of the shuffle proof: one term per ciphertext of a line
int drx_rep_commit_dev(drx_ctx* c, size_t n, int nterms, const uint8_t* d_scalars, const uint8_t* d_points, uint8_t* d_out) {
  if (!c) return fail(-1, "null context");
  if (nterms < 1 || nterms > kRepMaxTerms) return fail(-1, "drx_rep_commit: nterms %d outside 1..%d", nterms, kRepMaxTerms);
  Guard g(c);
  timed(c, "k_rep_commit", [&] { return launch_rep_commit(c->stream, n, nterms, d_scalars, d_points, d_out, dev_bad(c)); });
  return check_launch("rep_commit");
}

int drx_rep_commit(drx_ctx* c, size_t n, int nterms, const uint8_t* scalars, const uint8_t* points, uint8_t* out) {
  if (!c) return fail(-1, "null context");
  if (nterms < 1 || nterms > kRepMaxTerms) return fail(-1, "drx_rep_commit: nterms %d outside 1..%d", nterms, kRepMaxTerms);
  if (n == 0) return 0;
  if (!scalars || !points || !out) return fail(-1, "drx_rep_commit: null buffer");
  Guard g(c);
  Arena& a = c->arena;
  a.plan_reset();
  const size_t nt = (size_t)nterms;
  if (nterms >= kRepTreeMinTerms) {
    // One thread walks the terms of a relation (~1.2 ms each), so a wide relation is summed in two levels: `chunk`
    // terms per thread (chunk = the power of two at or above sqrt(nterms)), then the m partial sums with scalar 1.
    // Rows are padded to m * chunk terms with zero scalars on the row's own first point.
    size_t chunk = 8;
    while (chunk * chunk < nt) chunk *= 2;
    const size_t m = (nt + chunk - 1) / chunk, T = m * chunk;
    std::vector<uint8_t> hs(n * T * 32, 0), hp(n * T * 64), ones(n * m * 32, 0);
    for (size_t i = 0; i < n; i++) {
      memcpy(&hs[i * T * 32], scalars + i * nt * 32, nt * 32);
      memcpy(&hp[i * T * 64], points + i * nt * 64, nt * 64);
      for (size_t j = nt; j < T; j++) memcpy(&hp[(i * T + j) * 64], points + i * nt * 64, 64);
    }
    for (size_t k = 0; k < n * m; k++) ones[k * 32 + 31] = 1;
    size_t os = a.plan(n * T * 32), op = a.plan(n * T * 64), o1 = a.plan(n * m * 32), opart = a.plan(n * m * 64), oo = a.plan(n * 64);
    RET(a.commit());
    RET(reset_bad(c));
    RET(h2d(c, a.at<void>(os), hs.data(), n * T * 32));
    RET(h2d(c, a.at<void>(op), hp.data(), n * T * 64));
    RET(h2d(c, a.at<void>(o1), ones.data(), n * m * 32));
    timed(c, "k_rep_commit", [&] {
      return launch_rep_commit(c->stream, n * m, (int)chunk, a.at<uint8_t>(os), a.at<uint8_t>(op), a.at<uint8_t>(opart), (int*)c->bad.p) +
             launch_rep_commit(c->stream, n, (int)m, a.at<uint8_t>(o1), a.at<uint8_t>(opart), a.at<uint8_t>(oo), (int*)c->bad.p);
    });
    RET(check_launch("rep_commit"));
    RET(d2h(c, out, a.at<void>(oo), n * 64));
    RET(sync_out(c));  // the staging vectors above are read by the copies until here
    int bad = 0;
    RET(read_bad(c, &bad));
    if (bad) return fail(-2, "drx_rep_commit: %d malformed points", bad);
    return 0;
  }
  size_t os = a.plan(n * nt * 32), op = a.plan(n * nt * 64), oo = a.plan(n * 64);
  RET(a.commit());
  RET(reset_bad(c));
  RET(h2d(c, a.at<void>(os), scalars, n * nt * 32));
  RET(h2d(c, a.at<void>(op), points, n * nt * 64));
  timed(c, "k_rep_commit", [&] {
    return launch_rep_commit(c->stream, n, nterms, a.at<uint8_t>(os), a.at<uint8_t>(op), a.at<uint8_t>(oo), (int*)c->bad.p);
  });
  RET(check_launch("rep_commit"));
  RET(d2h(c, out, a.at<void>(oo), n * 64));
  int bad = 0;
  RET(read_bad(c, &bad));
  if (bad) return fail(-2, "drx_rep_commit: %d malformed points", bad);
  return 0;
}

int drx_rep_respond(drx_ctx* c, size_t n, int per, const uint8_t* v, const uint8_t* chal, const uint8_t* x, uint8_t* r) {
  if (!c) return fail(-1, "null context");
  if (per < 1) return fail(-1, "drx_rep_respond: per must be at least 1");
  if (n == 0) return 0;
  if (!v || !chal || !x || !r) return fail(-1, "drx_rep_respond: null buffer");
  Guard g(c);
  Arena& a = c->arena;
  a.plan_reset();
  const size_t nc = (n + (size_t)per - 1) / (size_t)per;
  size_t ov = a.plan(n * 32), oc = a.plan(nc * 32), ox = a.plan(n * 32), orr = a.plan(n * 32);
  RET(a.commit());
  RET(h2d(c, a.at<void>(ov), v, n * 32));
  RET(h2d(c, a.at<void>(oc), chal, nc * 32));
  RET(h2d(c, a.at<void>(ox), x, n * 32));
  timed(c, "k_rep_respond", [&] {
    return launch_rep_respond(c->stream, n, per, a.at<uint8_t>(ov), a.at<uint8_t>(oc), a.at<uint8_t>(ox), a.at<uint8_t>(orr));
  });
  RET(check_launch("rep_respond"));
  RET(d2h(c, r, a.at<void>(orr), n * 32));
  RET(sync_out(c));
  // the secrets and their blinding values do not stay behind in the arena
  cudaMemsetAsync(a.at<void>(ov), 0, n * 32, c->stream);
  cudaMemsetAsync(a.at<void>(ox), 0, n * 32, c->stream);
  return 0;
}

int drx_rep_verify_dev(drx_ctx* c, size_t n, int nterms, const uint8_t* d_resp, const uint8_t* d_chal, const uint8_t* d_bases,
                       const uint8_t* d_P, const uint8_t* d_V, uint8_t* d_ok) {
  if (!c) return fail(-1, "null context");
  if (nterms < 1 || nterms > kRepMaxTerms) return fail(-1, "drx_rep_verify: nterms %d outside 1..%d", nterms, kRepMaxTerms);
  Guard g(c);
  timed(c, "k_rep_verify", [&] { return launch_rep_verify(c->stream, n, nterms, d_resp, d_chal, d_bases, d_P, d_V, d_ok); });
  return check_launch("rep_verify");
}

int drx_rep_verify(drx_ctx* c, size_t n, int nterms, const uint8_t* resp, const uint8_t* chal, const uint8_t* bases,
                   const uint8_t* P, const uint8_t* V, uint8_t* ok) {
  if (!c) return fail(-1, "null context");
  if (nterms < 1 || nterms > kRepMaxTerms) return fail(-1, "drx_rep_verify: nterms %d outside 1..%d", nterms, kRepMaxTerms);
  if (n == 0) return 0;
  if (!resp || !chal || !bases || !P || !V || !ok) return fail(-1, "drx_rep_verify: null buffer");
  Guard g(c);
  Arena& a = c->arena;
  a.plan_reset();
  const size_t nt = (size_t)nterms;
  size_t orr = a.plan(n * nt * 32), oc = a.plan(n * 32), ob = a.plan(n * nt * 64), op = a.plan(n * 64), ov = a.plan(n * 64),
         oo = a.plan(n);
  RET(a.commit());
  RET(h2d(c, a.at<void>(orr), resp, n * nt * 32));
  RET(h2d(c, a.at<void>(oc), chal, n * 32));
  RET(h2d(c, a.at<void>(ob), bases, n * nt * 64));
  RET(h2d(c, a.at<void>(op), P, n * 64));
  RET(h2d(c, a.at<void>(ov), V, n * 64));
  timed(c, "k_rep_verify", [&] {
    return launch_rep_verify(c->stream, n, nterms, a.at<uint8_t>(orr), a.at<uint8_t>(oc), a.at<uint8_t>(ob), a.at<uint8_t>(op),
                             a.at<uint8_t>(ov), a.at<uint8_t>(oo));
  });
  RET(check_launch("rep_verify"));
  RET(d2h(c, ok, a.at<void>(oo), n));
  RET(sync_out(c));
  return 0;
}

int drx_decrypt_dev(drx_ctx* c, size_t n, const uint8_t sk[32], const uint8_t* d_ct, int64_t* d_out) {
  if (!c || !sk) return fail(-1, "null argument");
  Guard g(c);
  RET(ensure_dlog(c));
  // the key travels through the tail of the 256-byte flag buffer
  RET(h2d(c, (uint8_t*)c->bad.p + 64, sk, 32));
  timed(c, "k_decrypt", [&] {
    return launch_decrypt(c->stream, n, (uint8_t*)c->bad.p + 64, d_ct, c->dlog_keys.p, (uint32_t*)c->dlog_vals.p, d_out,
                          dev_bad(c));
  });
  return check_launch("decrypt");
}

static int decrypt_raw(drx_ctx* c, size_t n, const uint8_t sk[32], const uint8_t* ct, std::vector<int64_t>& raw) {
  RET(ensure_dlog(c));
  Arena& a = c->arena;
  a.plan_reset();
  size_t oc = a.plan(n * 128), oo = a.plan(n * 8);
  RET(a.commit());
  RET(reset_bad(c));
  RET(h2d(c, a.at<void>(oc), ct, n * 128));
  RET(h2d(c, (uint8_t*)c->bad.p + 64, sk, 32));
  timed(c, "k_decrypt", [&] {
    return launch_decrypt(c->stream, n, (uint8_t*)c->bad.p + 64, a.at<uint8_t>(oc), c->dlog_keys.p,
                          (uint32_t*)c->dlog_vals.p, a.at<int64_t>(oo), (int*)c->bad.p);
  });
  RET(check_launch("decrypt"));
  raw.resize(n);
  RET(d2h(c, raw.data(), a.at<void>(oo), n * 8));
  int bad = 0;
  RET(read_bad(c, &bad));
  if (bad) return fail(-2, "drx_decrypt: %d malformed points", bad);
  return 0;
}

int drx_decrypt(drx_ctx* c, size_t n, const uint8_t sk[32], const uint8_t* ct, int with_neg, int64_t* out,
                uint8_t* found) {
  if (!c) return fail(-1, "null context");
  if (n == 0) return 0;
  if (!sk || !ct || !out) return fail(-1, "drx_decrypt: null buffer");
  Guard g(c);
  std::vector<int64_t> raw;
  RET(decrypt_raw(c, n, sk, ct, raw));
  for (size_t i = 0; i < n; i++) {
    bool f;
    out[i] = c->dlog.resolve(raw[i], with_neg != 0, &f);
    if (found) found[i] = f ? 1 : 0;
  }
  return 0;
}

int drx_decrypt_check_zero(drx_ctx* c, size_t n, const uint8_t sk[32], const uint8_t* ct, int64_t* out) {
  if (!c) return fail(-1, "null context");
  if (n == 0) return 0;
  if (!sk || !ct || !out) return fail(-1, "drx_decrypt_check_zero: null buffer");
  Guard g(c);
  std::vector<int64_t> raw;
  RET(decrypt_raw(c, n, sk, ct, raw));
  for (size_t i = 0; i < n; i++) out[i] = raw[i] == 0 ? 0 : 1;
  return 0;
}

// ------------------------------------------------------------------ diagnostics
static int run_wire_op(drx_ctx* c, size_t n, const void* in1, size_t s1, const void* in2, size_t s2, void* out,
                       size_t so, int kind) {
  Arena& a = c->arena;
  a.plan_reset();
  size_t o1 = a.plan(n * s1), o2 = a.plan(n * s2), oo = a.plan(n * so);
  RET(a.commit());
  RET(reset_bad(c));
  RET(h2d(c, a.at<void>(o1), in1, n * s1));
  RET(h2d(c, a.at<void>(o2), in2, n * s2));
  int* bad = (int*)c->bad.p;
  if (kind == 0) c->launches += launch_pair_wire(c->stream, n, a.at<uint8_t>(o1), a.at<uint8_t>(o2), a.at<uint8_t>(oo), bad);
  if (kind == 1) c->launches += launch_g1_mul(c->stream, n, a.at<uint8_t>(o1), a.at<uint8_t>(o2), a.at<uint8_t>(oo), bad);
  if (kind == 2) c->launches += launch_g2_mul(c->stream, n, a.at<uint8_t>(o1), a.at<uint8_t>(o2), a.at<uint8_t>(oo), bad);
  RET(check_launch("wire op"));
  RET(d2h(c, out, a.at<void>(oo), n * so));
  int nb = 0;
  RET(read_bad(c, &nb));
  if (nb) return fail(-2, "%d malformed points", nb);
  return 0;
}
int drx_pair(drx_ctx* c, size_t n, const uint8_t* g1, const uint8_t* g2, uint8_t* gt) {
  if (!c) return fail(-1, "null context");
  if (!n) return 0;
  Guard g(c);
  return run_wire_op(c, n, g1, 64, g2, 128, gt, 384, 0);
}
int drx_g1_mul(drx_ctx* c, size_t n, const uint8_t* k, const uint8_t* pt, uint8_t* out) {
  if (!c) return fail(-1, "null context");
  if (!n) return 0;
  Guard g(c);
  return run_wire_op(c, n, k, 32, pt, 64, out, 64, 1);
}
int drx_g2_mul(drx_ctx* c, size_t n, const uint8_t* k, const uint8_t* pt, uint8_t* out) {
  if (!c) return fail(-1, "null context");
  if (!n) return 0;
  Guard g(c);
  return run_wire_op(c, n, k, 32, pt, 128, out, 128, 2);
}
int drx_mont_mul(drx_ctx* c, size_t n, int which, const uint32_t* x, const uint32_t* y, uint32_t* r) {
  if (!c) return fail(-1, "null context");
  if (!n) return 0;
  Guard g(c);
  Arena& a = c->arena;
  a.plan_reset();
  size_t o1 = a.plan(n * 32), o2 = a.plan(n * 32), oo = a.plan(n * 32);
  RET(a.commit());
  RET(h2d(c, a.at<void>(o1), x, n * 32));
  RET(h2d(c, a.at<void>(o2), y, n * 32));
  c->launches += launch_mont_mul(c->stream, n, which, a.at<uint32_t>(o1), a.at<uint32_t>(o2), a.at<uint32_t>(oo));
  RET(check_launch("mont_mul"));
  RET(d2h(c, r, a.at<void>(oo), n * 32));
  RET(sync_out(c));
  return 0;
}
int drx_microbench(drx_ctx* c, int kind, int iters, double* ms, double* work) {
  if (!c || !ms || !work) return fail(-1, "null argument");
  Guard g(c);
  // kind 0: wide multiply-accumulate chains, 1: dependent Montgomery products (both fill the GPU: the roofline peaks);
  // kinds 10 + k / 20 + k / 30 + k: probe k (1: one chain of products, 2: three chains) with one warp per SM / one
  // warp per sub-partition / two warps per sub-partition (latency probes, profiles/r01_experiments.md); k = 3: dependent
  // field inversions, k = 4: dependent full point additions; kind 5 (full GPU only): chains of double-precision FMAs
  int blocks = 148 * 8, threads = 256;
  if (kind >= 10) {
    blocks = 148;
    threads = kind >= 30 ? 256 : kind >= 20 ? 128 : 32;
    kind %= 10;
  }
  RET(c->sink.ensure((size_t)148 * 8 * 256 * 4));
  cudaEvent_t e0, e1;
  CU(cudaEventCreate(&e0));
  CU(cudaEventCreate(&e1));
  launch_microbench(c->stream, kind, iters / 8 + 1, blocks, threads, (uint32_t*)c->sink.p);  // warm-up
  CU(cudaEventRecord(e0, c->stream));
  c->launches += 1 + launch_microbench(c->stream, kind, iters, blocks, threads, (uint32_t*)c->sink.p);
  CU(cudaEventRecord(e1, c->stream));
  CU(cudaEventSynchronize(e1));
  float t = 0;
  CU(cudaEventElapsedTime(&t, e0, e1));
  cudaEventDestroy(e0);
  cudaEventDestroy(e1);
  *ms = t;
  // wide MACs, Montgomery products, or (kinds 3 / 4) inversions / full point additions
  const double per_thread = kind == 0 || kind == 5 ? 16.0 * iters : kind == 2 ? 3.0 * iters : kind >= 3 ? 1.0 * iters : 2.0 * iters;
  *work = per_thread * blocks * threads;
  return check_launch("microbench");
}

}  // extern "C"

// ---- range-proof entry points: see capi_rp.inc
#include "capi_rp.inc"
// ---- edwards25519 ElGamal entry points: see capi_ed.inc
#include "capi_ed.inc"
