// Shared between the lane-group pairing kernels (ucode_split.cuh, instantiated for four and for eight program lanes)
// and the kernel that prepares their arguments: the prepared-argument record, per-device launch state, and the layout
// of the dynamic kernels' control words.
#pragma once
#include <mutex>

#include "rp_device.cuh"

namespace drx {

struct UcPrep {      // per item: the two pairing arguments in Montgomery affine form
  G1Aff P;           // c*Y_i - Zphi_j*B
  G2Aff V;
  uint32_t state;    // 0: run the pairing, 1: an argument is the point at infinity (value 1), 2: malformed V
  uint32_t pad[3];
};

static inline size_t uc_align256(size_t x) { return (x + 255) & ~(size_t)255; }
static inline int uc_grid_for(size_t n, int threads) { return (int)((n + threads - 1) / threads); }

// Device attributes and per-device kernel configuration. cudaFuncSetAttribute is a per-device setting and a process
// may hold one context per GPU (and call from several threads), so everything here is keyed by the device ordinal
// that is current at the launch and guarded by one mutex.
constexpr int kUcMaxDevices = 64;
struct UcDevInfo {
  int nsm = 0, smem_sm = 0, smem_optin = 0;
};
static inline std::mutex& uc_dev_mutex() {
  static std::mutex m;
  return m;
}
static inline bool uc_dev_info(UcDevInfo& out, int& dev) {
  static UcDevInfo info[kUcMaxDevices];
  if (cudaGetDevice(&dev) != cudaSuccess || dev < 0 || dev >= kUcMaxDevices) return false;
  std::lock_guard<std::mutex> lk(uc_dev_mutex());
  UcDevInfo& d = info[dev];
  if (!d.nsm) {
    cudaDeviceGetAttribute(&d.nsm, cudaDevAttrMultiProcessorCount, dev);
    cudaDeviceGetAttribute(&d.smem_sm, cudaDevAttrMaxSharedMemoryPerMultiprocessor, dev);
    cudaDeviceGetAttribute(&d.smem_optin, cudaDevAttrMaxSharedMemoryPerBlockOptin, dev);
  }
  out = d;
  return d.nsm > 0;
}
// opt one kernel into the device's full shared-memory carve-out, once per device
template <class K>
static inline bool uc_optin_once(K kernel, bool (&done)[kUcMaxDevices], int dev, int bytes) {
  std::lock_guard<std::mutex> lk(uc_dev_mutex());
  if (done[dev]) return true;
  if (cudaFuncSetAttribute(kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, bytes) != cudaSuccess) return false;
  done[dev] = true;
  return true;
}

// Control words of a dynamic (task-queue) launch: [0] next task, [1..3] unused, [4 + g] segments finished for item
// group g. A warp that waits for a predecessor segment longer than kUcMaxPolls polls (minutes; the predecessor was
// handed out earlier and is running, so this only happens when the device is wedged) raises bit 0 of *err, which the
// host reads at the next synchronising call, and leaves; it never traps, so a slow wait (time-slicing, a profiler
// replay) cannot poison the CUDA context.
constexpr unsigned int kUcMaxPolls = 1u << 30;
// A dynamic launch handles at most kUcDynChunkItems items (larger batches are a sequence of launches on one stream
// that reuse the control words and the slot-image area), which bounds that area at 256 MB whatever the batch size.
constexpr size_t kUcDynChunkItems = 65536;
constexpr size_t kUcMaxSlots = 64;  // upper bound of UC_NSLOTS over the generated programs (static_assert'ed per kernel)
static inline size_t uc_dyn_max_groups() { return kUcDynChunkItems / 2; }  // smallest group: two items per warp
// bytes behind the UcPrep array: control words, then slot images
static inline size_t uc_dyn_ctrl_bytes() { return uc_align256(16 + 4 * uc_dyn_max_groups()); }
static inline size_t uc_dyn_state_bytes(size_t items) {
  const size_t n = items < kUcDynChunkItems ? items : kUcDynChunkItems;
  return ((n + 7) & ~(size_t)7) * kUcMaxSlots * 64;
}

}  // namespace drx
