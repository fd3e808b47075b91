// LGBM_B200_RootHistogramNoAlloc — the histogram kernel (K4) for callers that are not allowed to allocate device memory.
//
// Who cannot allocate: code that runs inside somebody else's memory budget or allocator discipline — a Spark executor
// plugin whose GPU memory is handed out by an external pool (RMM-style arenas, a fixed per-task budget the plugin may
// not exceed), or a latency-critical loop that must not take the driver's allocation lock.  For them cudaMalloc /
// cudaFree (which also synchronise the whole device) are off limits, so the scratch this kernel needs — the int64
// accumulation area the CTAs flush into plus the small leaf / control descriptors — lives in a fixed array in the
// library's own device image instead.
//
// Consequences the caller must know:
//   * every GPU that loads the library carries the array (about 1 MB) until the process exits;
//   * the array is shared by all calls on a device and keeps what the last call left in it; this entry point clears the
//     part that has to start at zero, but two calls that overlap on the same GPU (different streams or threads) race on
//     it — serialise them;
//   * sizes beyond the fixed capacity are refused with an error, never silently truncated;
//   * the first call on each device resolves the array's device address (and sets the kernel's shared-memory attribute),
//     so make that first call before any stream capture begins.
#include <algorithm>
#include <cstdio>
#include <mutex>

#include "../../include/lightgbm_b200_c_api.h"
#include "hist_kernel.cuh"

namespace {

using namespace lgb;

constexpr int kNoAllocMaxFeatures = 256;                                          // stated capacity
constexpr size_t kNoAllocHistBytes = size_t(kNoAllocMaxFeatures) * kBins * 2 * sizeof(long long);   // 1 MiB
constexpr size_t kNoAllocDescBytes = 4096;                                        // LeafDev[1] + TreeCtrl
constexpr size_t kNoAllocBytes = kNoAllocHistBytes + kNoAllocDescBytes;

// Host code must never use this name as a pointer: on the host it is only a stand-in for the device object.
static __device__ __align__(256) unsigned char g_noalloc_scratch[kNoAllocBytes];

constexpr int kMaxDevices = 64;
std::mutex g_mu;
unsigned char* g_scratch_ptr[kMaxDevices] = {nullptr};

thread_local char g_msg[512];
}  // namespace
namespace lgb { void SetLastErrorMessage(const char* m); }
namespace {

__global__ void k_noalloc_init(LeafDev* leaf, TreeCtrl* ctrl, int n) {
  LeafDev z;
  z.begin = 0; z.count = n; z.buf = -1; z.global_count = n; z.int_g = z.int_h = 0; z.sum_g = z.sum_h = 0.0;
  z.depth = 0; z.hist_slot = 0;
  z.best.gain = 0.0; z.best.feature = -1; z.best.real_feature = -1; z.best.threshold = 0; z.best.default_left = 1;
  z.best.left_count = z.best.right_count = 0; z.best.left_output = z.best.right_output = 0.0;
  z.best.left_sum_g = z.best.left_sum_h = z.best.right_sum_g = z.best.right_sum_h = 0.0;
  z.best.left_int_g = z.best.left_int_h = z.best.right_int_g = z.best.right_int_h = 0;
  leaf[0] = z;
  ctrl->done = 0; ctrl->find = 1; ctrl->smaller = 0; ctrl->hist_rows = 0; ctrl->hist_launches = 0;
}

// Set-up, once per device, with that device current: resolve the device address of the scratch array.
cudaError_t ScratchForCurrentDevice(int dev, unsigned char** out) {
  std::lock_guard<std::mutex> lock(g_mu);
  if (g_scratch_ptr[dev] == nullptr) {
    unsigned char* p = nullptr;
    cudaError_t e = cudaGetSymbolAddress(reinterpret_cast<void**>(&p), g_noalloc_scratch);
    if (e != cudaSuccess) return e;
    e = cudaFuncSetAttribute(k_hist<4, 4, false, true>, cudaFuncAttributeMaxDynamicSharedMemorySize, static_cast<int>(kHistSmemBytes));
    if (e != cudaSuccess) return e;
    e = cudaFuncSetAttribute(k_hist<2, 4, false, true>, cudaFuncAttributeMaxDynamicSharedMemorySize, static_cast<int>(kHistSmemBytes));
    if (e != cudaSuccess) return e;
    e = cudaFuncSetAttribute(k_hist<1, 4, false, true>, cudaFuncAttributeMaxDynamicSharedMemorySize, static_cast<int>(kHistSmemBytes));
    if (e != cudaSuccess) return e;
    g_scratch_ptr[dev] = p;
  }
  *out = g_scratch_ptr[dev];
  return cudaSuccess;
}

}  // namespace

extern "C" {

// Histogram of ALL rows of a binned row-major matrix with quantized gradients (the root-leaf case of kernel K4).
//   d_rows      device, uint8 [num_rows][row_stride]: one record per row, feature f's bin at byte f (row_stride is 32 for
//               <= 32 features, 64 for <= 64, else a multiple of 128 — the layout LGBM_B200_DatasetCopyRowMajor reports)
//   d_gh        device, uint16 [num_rows]: (int8 gradient << 8) | uint8 hessian
//   d_out_hist  device, int64 [num_features][256][2]: (sum gradient, sum hessian) per (feature, bin); overwritten
//   stream      a cudaStream_t (NULL = default stream); the call only enqueues work, it does not synchronise
// Allocates nothing (no cudaMalloc / cudaFree, no allocating library call).  Returns 0, or -1 with a message available
// from LGBM_GetLastError().  Capacity: num_features <= 256.  See the header of this file for the sharing rules.
int LGBM_B200_RootHistogramNoAlloc(const uint8_t* d_rows, int32_t row_stride, const uint16_t* d_gh, int32_t num_rows,
                                   int32_t num_features, int32_t num_grad_quant_bins, int64_t* d_out_hist, void* stream) {
  auto fail = [](const char* what, cudaError_t e) {
    if (e != cudaSuccess) snprintf(g_msg, sizeof(g_msg), "%s: %s", what, cudaGetErrorString(e));
    else snprintf(g_msg, sizeof(g_msg), "%s", what);
    lgb::SetLastErrorMessage(g_msg);
    return -1;
  };
  if (d_rows == nullptr || d_gh == nullptr || d_out_hist == nullptr) return fail("RootHistogramNoAlloc: null device pointer", cudaSuccess);
  if (num_features < 1 || num_features > kNoAllocMaxFeatures) {
    snprintf(g_msg, sizeof(g_msg), "RootHistogramNoAlloc: num_features=%d exceeds the fixed scratch capacity of %d features "
             "(the scratch array is static; use the allocating path for wider data)", num_features, kNoAllocMaxFeatures);
    lgb::SetLastErrorMessage(g_msg);
    return -1;
  }
  if (num_rows < 1) return fail("RootHistogramNoAlloc: num_rows must be positive", cudaSuccess);
  if (num_grad_quant_bins < 2 || num_grad_quant_bins > 64) return fail("RootHistogramNoAlloc: num_grad_quant_bins outside [2, 64]", cudaSuccess);
  int R;
  if (row_stride == 32 && num_features <= 32) R = 4;
  else if (row_stride == 64 && num_features <= 64) R = 2;
  else if (row_stride % 128 == 0 && row_stride >= num_features) R = 1;
  else return fail("RootHistogramNoAlloc: row_stride does not match the row-record layout (32, 64 or a multiple of 128)", cudaSuccess);
  int dev = -1;
  cudaError_t e = cudaGetDevice(&dev);
  if (e != cudaSuccess) return fail("RootHistogramNoAlloc: cudaGetDevice", e);
  if (dev < 0 || dev >= kMaxDevices) return fail("RootHistogramNoAlloc: device ordinal out of range", cudaSuccess);
  unsigned char* scratch = nullptr;
  e = ScratchForCurrentDevice(dev, &scratch);
  if (e != cudaSuccess) return fail("RootHistogramNoAlloc: cudaGetSymbolAddress / set-up failed", e);
  cudaStream_t s = static_cast<cudaStream_t>(stream);
  long long* hist = reinterpret_cast<long long*>(scratch);
  LeafDev* leaf = reinterpret_cast<LeafDev*>(scratch + kNoAllocHistBytes);
  TreeCtrl* ctrl = reinterpret_cast<TreeCtrl*>(scratch + kNoAllocHistBytes + 1024);
  const size_t hist_bytes = static_cast<size_t>(num_features) * kBins * 2 * sizeof(long long);
  // the accumulation area must start at zero (it still holds the previous call's sums)
  e = cudaMemsetAsync(hist, 0, hist_bytes, s);
  if (e != cudaSuccess) return fail("RootHistogramNoAlloc: cudaMemsetAsync", e);
  k_noalloc_init<<<1, 1, 0, s>>>(leaf, ctrl, num_rows);
  const int ft = 32 / R * 4;
  const int tiles = (num_features + ft - 1) / ft;
  dim3 grid(std::max(1, 148 / tiles), tiles);
  const int per_copy = std::min(32767 / (num_grad_quant_bins / 2 + 1), 65535 / (num_grad_quant_bins + 1));
  const int rows_per_flush = std::max(32, (per_copy * R) / 32 * 32);
  if (R == 4) k_hist<4, 4, false, true><<<grid, kHistThreads, kHistSmemBytes, s>>>(d_rows, static_cast<unsigned>(row_stride), d_gh, nullptr, nullptr, nullptr, leaf, ctrl, hist, num_features, rows_per_flush);
  else if (R == 2) k_hist<2, 4, false, true><<<grid, kHistThreads, kHistSmemBytes, s>>>(d_rows, static_cast<unsigned>(row_stride), d_gh, nullptr, nullptr, nullptr, leaf, ctrl, hist, num_features, rows_per_flush);
  else k_hist<1, 4, false, true><<<grid, kHistThreads, kHistSmemBytes, s>>>(d_rows, static_cast<unsigned>(row_stride), d_gh, nullptr, nullptr, nullptr, leaf, ctrl, hist, num_features, rows_per_flush);
  e = cudaGetLastError();
  if (e != cudaSuccess) return fail("RootHistogramNoAlloc: kernel launch", e);
  e = cudaMemcpyAsync(d_out_hist, hist, hist_bytes, cudaMemcpyDeviceToDevice, s);
  if (e != cudaSuccess) return fail("RootHistogramNoAlloc: cudaMemcpyAsync", e);
  return 0;
}

}  // extern "C"
