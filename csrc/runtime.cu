// Device / memory / stream / IPC layer of the runtime (host code; sm_100a build).
//
// Replaces the reference's simulated device memory (`map[uint64][]byte`,
// DSML/gpu_device_service/gpu_device_server.go:26-49) and the coordinator's connection set-up
// (`grpc.Dial` per device, DSML/gpu_coordinator_service/gpu_coordinator_server.go:131-164) with
// real CUDA devices: allocations, peer access over NVLink/NVSwitch, topology queries, pinned
// staging, and CUDA-IPC handles so one-process-per-GPU jobs can map each other's buffers.
#include "common.h"

#include <cuda.h>
#include <string.h>


using namespace dmlp;

extern "C" int dmlp_device_count(int* n) {
  DMLP_REQUIRE(n != nullptr, "dmlp_device_count: null out pointer");
  cudaError_t e = cudaGetDeviceCount(n);
  if (e != cudaSuccess) {
    *n = 0;
    cudaGetLastError();
    return set_error(DMLP_ERR_CUDA, "cudaGetDeviceCount: %s", cudaGetErrorString(e));
  }
  return DMLP_OK;
}

extern "C" int dmlp_device_info(int dev, dmlp_device_info_t* out) {
  DMLP_REQUIRE(out != nullptr, "dmlp_device_info: null out pointer");
  cudaDeviceProp p;
  DMLP_CUDA_CHECK(cudaGetDeviceProperties(&p, dev));
  memset(out, 0, sizeof(*out));
  snprintf(out->name, sizeof(out->name), "%s", p.name);  // (truncates a longer name, always terminated)
  const unsigned char* u = reinterpret_cast<const unsigned char*>(p.uuid.bytes);
  snprintf(out->uuid, sizeof(out->uuid),
           "GPU-%02x%02x%02x%02x-%02x%02x-%02x%02x-%02x%02x-%02x%02x%02x%02x%02x%02x", u[0], u[1],
           u[2], u[3], u[4], u[5], u[6], u[7], u[8], u[9], u[10], u[11], u[12], u[13], u[14], u[15]);
  out->sm_count = p.multiProcessorCount;
  out->cc_major = p.major;
  out->cc_minor = p.minor;
  out->total_mem = p.totalGlobalMem;
  out->l2_bytes = p.l2CacheSize;
  out->smem_per_block_optin = static_cast<int>(p.sharedMemPerBlockOptin);
  int v = 0;
  if (cudaDeviceGetAttribute(&v, cudaDevAttrClockRate, dev) == cudaSuccess) out->clock_khz = v;
  CUdevice cudev;
  // Multicast (NVLS) capability is a driver attribute.
  typedef CUresult (*GetAttrFn)(int*, CUdevice_attribute, CUdevice);
  typedef CUresult (*GetDevFn)(CUdevice*, int);
  void* fa = nullptr;
  void* fd = nullptr;
  cudaDriverEntryPointQueryResult q;
  if (cudaGetDriverEntryPoint("cuDeviceGetAttribute", &fa, cudaEnableDefault, &q) == cudaSuccess && fa &&
      cudaGetDriverEntryPoint("cuDeviceGet", &fd, cudaEnableDefault, &q) == cudaSuccess && fd) {
    if (reinterpret_cast<GetDevFn>(fd)(&cudev, dev) == CUDA_SUCCESS) {
      int mc = 0;
      if (reinterpret_cast<GetAttrFn>(fa)(&mc, CU_DEVICE_ATTRIBUTE_MULTICAST_SUPPORTED, cudev) == CUDA_SUCCESS)
        out->multicast_supported = mc;
    }
  }
  return DMLP_OK;
}

// access[i*n+j] = 1 if devs[i] can map devs[j]'s memory; atomics / perf_rank from the P2P attributes.
extern "C" int dmlp_peer_matrix(int n, const int* devs, int* access, int* atomics, int* perf_rank) {
  DMLP_REQUIRE(n > 0 && devs && access, "dmlp_peer_matrix: bad arguments");
  for (int i = 0; i < n; ++i)
    for (int j = 0; j < n; ++j) {
      int a = 1, at = 1, pr = 0;
      if (devs[i] != devs[j]) {
        DMLP_CUDA_CHECK(cudaDeviceCanAccessPeer(&a, devs[i], devs[j]));
        if (a) {
          cudaDeviceGetP2PAttribute(&at, cudaDevP2PAttrNativeAtomicSupported, devs[i], devs[j]);
          cudaDeviceGetP2PAttribute(&pr, cudaDevP2PAttrPerformanceRank, devs[i], devs[j]);
        } else {
          at = 0;
        }
      }
      access[i * n + j] = a;
      if (atomics) atomics[i * n + j] = at;
      if (perf_rank) perf_rank[i * n + j] = pr;
    }
  return DMLP_OK;
}

extern "C" int dmlp_enable_peer_access(int n, const int* devs) {
  DMLP_REQUIRE(n > 0 && devs, "dmlp_enable_peer_access: bad arguments");
  for (int i = 0; i < n; ++i) {
    DeviceGuard g(devs[i]);
    if (!g.ok) return set_error(DMLP_ERR_CUDA, "cannot select device %d", devs[i]);
    for (int j = 0; j < n; ++j) {
      if (devs[i] == devs[j]) continue;
      int can = 0;
      DMLP_CUDA_CHECK(cudaDeviceCanAccessPeer(&can, devs[i], devs[j]));
      if (!can)
        return set_error(DMLP_ERR_UNSUPPORTED, "device %d cannot access peer %d (no NVLink/PCIe P2P path)",
                         devs[i], devs[j]);
      cudaError_t e = cudaDeviceEnablePeerAccess(devs[j], 0);
      if (e == cudaErrorPeerAccessAlreadyEnabled) {
        cudaGetLastError();
      } else if (e != cudaSuccess) {
        return set_error(DMLP_ERR_CUDA, "cudaDeviceEnablePeerAccess(%d->%d): %s", devs[i], devs[j],
                         cudaGetErrorString(e));
      }
    }
  }
  return DMLP_OK;
}

extern "C" int dmlp_malloc(int dev, size_t bytes, void** ptr) {
  DMLP_REQUIRE(ptr != nullptr, "dmlp_malloc: null out pointer");
  DeviceGuard g(dev);
  if (!g.ok) return set_error(DMLP_ERR_CUDA, "cannot select device %d", dev);
  if (bytes == 0) bytes = 16;
  DMLP_CUDA_CHECK(cudaMalloc(ptr, bytes));
  // The zero-fill runs on the legacy default stream, which does not order against the library's non-blocking streams
  // (or other devices): it must have finished before anybody can touch the buffer. Allocation is not a hot path.
  DMLP_CUDA_CHECK(cudaMemsetAsync(*ptr, 0, bytes, nullptr));
  DMLP_CUDA_CHECK(cudaStreamSynchronize(nullptr));
  return DMLP_OK;
}

extern "C" int dmlp_free(int dev, void* ptr) {
  if (ptr == nullptr) return DMLP_OK;
  DeviceGuard g(dev);
  DMLP_CUDA_CHECK(cudaFree(ptr));
  return DMLP_OK;
}

extern "C" int dmlp_host_alloc(size_t bytes, void** ptr) {
  DMLP_REQUIRE(ptr != nullptr, "dmlp_host_alloc: null out pointer");
  DMLP_CUDA_CHECK(cudaHostAlloc(ptr, bytes ? bytes : 16, cudaHostAllocPortable | cudaHostAllocMapped));
  return DMLP_OK;
}

extern "C" int dmlp_host_free(void* ptr) {
  if (ptr) DMLP_CUDA_CHECK(cudaFreeHost(ptr));
  return DMLP_OK;
}

// Copies between host and device. stream == NULL: synchronous on the device's default stream.
extern "C" int dmlp_memcpy_h2d(int dev, void* dst_dev, const void* src_host, size_t bytes, void* stream) {
  if (bytes == 0) return DMLP_OK;
  DeviceGuard g(dev);
  if (stream) {
    DMLP_CUDA_CHECK(cudaMemcpyAsync(dst_dev, src_host, bytes, cudaMemcpyHostToDevice, as_stream(stream)));
  } else {
    DMLP_CUDA_CHECK(cudaMemcpy(dst_dev, src_host, bytes, cudaMemcpyHostToDevice));
  }
  return DMLP_OK;
}

extern "C" int dmlp_memcpy_d2h(int dev, void* dst_host, const void* src_dev, size_t bytes, void* stream) {
  if (bytes == 0) return DMLP_OK;
  DeviceGuard g(dev);
  if (stream) {
    DMLP_CUDA_CHECK(cudaMemcpyAsync(dst_host, src_dev, bytes, cudaMemcpyDeviceToHost, as_stream(stream)));
  } else {
    DMLP_CUDA_CHECK(cudaMemcpy(dst_host, src_dev, bytes, cudaMemcpyDeviceToHost));
  }
  return DMLP_OK;
}

extern "C" int dmlp_memcpy_d2d(int dev, void* dst, const void* src, size_t bytes, void* stream) {
  if (bytes == 0) return DMLP_OK;
  DeviceGuard g(dev);
  if (stream) {
    DMLP_CUDA_CHECK(cudaMemcpyAsync(dst, src, bytes, cudaMemcpyDeviceToDevice, as_stream(stream)));
  } else {
    DMLP_CUDA_CHECK(cudaMemcpy(dst, src, bytes, cudaMemcpyDeviceToDevice));
  }
  return DMLP_OK;
}

extern "C" int dmlp_memset(int dev, void* dst, int value, size_t bytes, void* stream) {
  if (bytes == 0) return DMLP_OK;
  DeviceGuard g(dev);
  if (stream) {
    DMLP_CUDA_CHECK(cudaMemsetAsync(dst, value, bytes, as_stream(stream)));
  } else {
    DMLP_CUDA_CHECK(cudaMemset(dst, value, bytes));
  }
  return DMLP_OK;
}

// ---------------------------------------------------------------- streams / events
extern "C" int dmlp_stream_create(int dev, void** stream) {
  DMLP_REQUIRE(stream != nullptr, "dmlp_stream_create: null out pointer");
  DeviceGuard g(dev);
  cudaStream_t s;
  DMLP_CUDA_CHECK(cudaStreamCreateWithFlags(&s, cudaStreamNonBlocking));
  *stream = s;
  return DMLP_OK;
}
extern "C" int dmlp_stream_destroy(int dev, void* stream) {
  DeviceGuard g(dev);
  if (stream) DMLP_CUDA_CHECK(cudaStreamDestroy(as_stream(stream)));
  return DMLP_OK;
}
extern "C" int dmlp_stream_sync(int dev, void* stream) {
  DeviceGuard g(dev);
  DMLP_CUDA_CHECK(cudaStreamSynchronize(as_stream(stream)));
  return DMLP_OK;
}
// 0 = complete, 1 = still running.
extern "C" int dmlp_stream_query(int dev, void* stream, int* busy) {
  DeviceGuard g(dev);
  cudaError_t e = cudaStreamQuery(as_stream(stream));
  if (e == cudaSuccess) { *busy = 0; return DMLP_OK; }
  if (e == cudaErrorNotReady) { cudaGetLastError(); *busy = 1; return DMLP_OK; }
  return set_error(DMLP_ERR_CUDA, "cudaStreamQuery: %s", cudaGetErrorString(e));
}
extern "C" int dmlp_device_sync(int dev) {
  DeviceGuard g(dev);
  DMLP_CUDA_CHECK(cudaDeviceSynchronize());
  return DMLP_OK;
}
extern "C" int dmlp_event_create(int dev, int timing, void** ev) {
  DMLP_REQUIRE(ev != nullptr, "dmlp_event_create: null out pointer");
  DeviceGuard g(dev);
  cudaEvent_t e;
  DMLP_CUDA_CHECK(cudaEventCreateWithFlags(&e, timing ? cudaEventDefault : cudaEventDisableTiming));
  *ev = e;
  return DMLP_OK;
}
extern "C" int dmlp_event_destroy(void* ev) {
  if (ev) DMLP_CUDA_CHECK(cudaEventDestroy(reinterpret_cast<cudaEvent_t>(ev)));
  return DMLP_OK;
}
extern "C" int dmlp_event_record(int dev, void* ev, void* stream) {
  DeviceGuard g(dev);
  DMLP_CUDA_CHECK(cudaEventRecord(reinterpret_cast<cudaEvent_t>(ev), as_stream(stream)));
  return DMLP_OK;
}
extern "C" int dmlp_event_query(void* ev, int* busy) {
  cudaError_t e = cudaEventQuery(reinterpret_cast<cudaEvent_t>(ev));
  if (e == cudaSuccess) { *busy = 0; return DMLP_OK; }
  if (e == cudaErrorNotReady) { cudaGetLastError(); *busy = 1; return DMLP_OK; }
  return set_error(DMLP_ERR_CUDA, "cudaEventQuery: %s", cudaGetErrorString(e));
}
extern "C" int dmlp_event_sync(void* ev) {
  DMLP_CUDA_CHECK(cudaEventSynchronize(reinterpret_cast<cudaEvent_t>(ev)));
  return DMLP_OK;
}
extern "C" int dmlp_event_elapsed_ms(void* start, void* stop, float* ms) {
  DMLP_CUDA_CHECK(cudaEventElapsedTime(ms, reinterpret_cast<cudaEvent_t>(start), reinterpret_cast<cudaEvent_t>(stop)));
  return DMLP_OK;
}
extern "C" int dmlp_stream_wait_event(int dev, void* stream, void* ev) {
  DeviceGuard g(dev);
  DMLP_CUDA_CHECK(cudaStreamWaitEvent(as_stream(stream), reinterpret_cast<cudaEvent_t>(ev), 0));
  return DMLP_OK;
}

// ---------------------------------------------------------------- CUDA IPC
extern "C" int dmlp_ipc_get_handle(int dev, void* ptr_dev, unsigned char out[64]) {
  DeviceGuard g(dev);
  static_assert(sizeof(cudaIpcMemHandle_t) == 64, "unexpected IPC handle size");
  cudaIpcMemHandle_t h;
  DMLP_CUDA_CHECK(cudaIpcGetMemHandle(&h, ptr_dev));
  memcpy(out, &h, 64);
  return DMLP_OK;
}
extern "C" int dmlp_ipc_open_handle(int dev, const unsigned char in[64], void** ptr_dev) {
  DeviceGuard g(dev);
  cudaIpcMemHandle_t h;
  memcpy(&h, in, 64);
  DMLP_CUDA_CHECK(cudaIpcOpenMemHandle(ptr_dev, h, cudaIpcMemLazyEnablePeerAccess));
  return DMLP_OK;
}
extern "C" int dmlp_ipc_close_handle(int dev, void* ptr_dev) {
  DeviceGuard g(dev);
  DMLP_CUDA_CHECK(cudaIpcCloseMemHandle(ptr_dev));
  return DMLP_OK;
}
