// Host-side helpers shared by all translation units: error plumbing for the C ABI.
#pragma once
#include <cuda_runtime.h>
#include <nvtx3/nvToolsExt.h>
#include <stdarg.h>
#include <stdio.h>
#include <stdlib.h>

#include "../include/dmlp_b200.h"

namespace dmlp {

// Stores a formatted message in the calling thread's error slot and returns `code`.
int set_error(int code, const char* fmt, ...) __attribute__((format(printf, 2, 3)));

#define DMLP_CUDA_CHECK(expr)                                                              \
  do {                                                                                     \
    cudaError_t _e = (expr);                                                               \
    if (_e != cudaSuccess)                                                                 \
      return ::dmlp::set_error(DMLP_ERR_CUDA, "%s failed: %s (%s:%d)", #expr,              \
                               cudaGetErrorString(_e), __FILE__, __LINE__);                \
  } while (0)

#define DMLP_REQUIRE(cond, ...)                                          \
  do {                                                                   \
    if (!(cond)) return ::dmlp::set_error(DMLP_ERR_INVALID, __VA_ARGS__); \
  } while (0)

inline cudaStream_t as_stream(void* s) { return reinterpret_cast<cudaStream_t>(s); }

// NVTX ranges around the host-side phases (trainer runs, graph capture, collective launches), for timelines taken with
// Nsight Systems / ncu --nvtx. Off unless DMLP_NVTX=1: the reference has no tracing at all.
inline bool nvtx_enabled() {
  static const bool on = [] { const char* e = getenv("DMLP_NVTX"); return e && atoi(e) > 0; }();
  return on;
}
struct NvtxRange {
  bool on;
  explicit NvtxRange(const char* name) : on(nvtx_enabled()) { if (on) nvtxRangePushA(name); }
  ~NvtxRange() { if (on) nvtxRangePop(); }
  NvtxRange(const NvtxRange&) = delete;
  NvtxRange& operator=(const NvtxRange&) = delete;
};

// Selects a device for the current scope and restores the caller's device afterwards, so the
// library never changes the current device behind PyTorch's back.
struct DeviceGuard {
  int prev = -1;
  bool ok = true;
  explicit DeviceGuard(int dev) {
    if (cudaGetDevice(&prev) != cudaSuccess) prev = -1;
    if (prev != dev) ok = cudaSetDevice(dev) == cudaSuccess;
  }
  ~DeviceGuard() {
    if (prev >= 0) cudaSetDevice(prev);
  }
};

}  // namespace dmlp

// Words of a rank's peer-visible signal pad (64 KiB at the start of its communicator segment) that more than one
// translation unit addresses. collectives.cu owns the layout and static_asserts these against it.
namespace dmlp {
constexpr int kPadAbortWord = 1280;          // abort word (== kPadAbort in collectives.cu)
constexpr int kPadFusedFlags = 8192;         // fused step kernel: [source rank][cta] arrival epochs (64 words)
constexpr int kPadFusedEpoch = 8192 + 64;    // fused step kernel: this rank's step epoch (local use)
// global-batch mode of the fused step (mlp_fused.cu) and its input copy kernel (input_exchange.cu)
constexpr int kXPushCtas = 16;               // CTAs of the input copy kernel (one ready word each)
constexpr int kPadXReady = 8192 + 256;       // [source rank][copy CTA]: steps whose inputs that rank has delivered to this rank
constexpr int kPadXDone = 8192 + 512;        // [source rank]: steps that rank has finished (its input slots may be overwritten)
}  // namespace dmlp
