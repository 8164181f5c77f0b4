// NVLS: all-reduce through NVSwitch multicast (multimem.ld_reduce / multimem.st), sm_100a.
//
// The fourth variant of the reference's AllReduceRing
// (DSML/gpu_coordinator_service/gpu_coordinator_server.go:272-566), next to ring / one-shot /
// two-shot in collectives.cu. Every rank binds a buffer of the same size to ONE multicast object
// (cuMulticastCreate / AddDevice / BindMem); a load from the multicast address makes the switch
// fetch the word from every replica and return the SUM (in-switch reduction), a store to it writes
// every replica. Rank r reduces slice r of the gradient with multimem.ld_reduce, applies
// average + SGD, and broadcasts the updated weights with multimem.st — (1 + 1)/n of the bytes per
// GPU instead of two-shot's 2(n-1)/n, and no kernel ever touches a peer's unicast address.
//
// Set-up is driver-API VMM (cuMemCreate/Map + multicast bind). One process owning all GPUs does
// everything locally; with one process per GPU the creating rank exports the multicast object as
// a POSIX file descriptor that the Python layer passes to the peers over a Unix socket.
#include "common.h"
#include "ptx_sm100.cuh"

#include <cuda.h>
#include <string.h>
#include <unistd.h>

#include <new>
#include <vector>

namespace dmlp {

struct DriverApi {
  bool ok = false;
  CUresult (*MemCreate)(CUmemGenericAllocationHandle*, size_t, const CUmemAllocationProp*, unsigned long long);
  CUresult (*MemRelease)(CUmemGenericAllocationHandle);
  CUresult (*MemAddressReserve)(CUdeviceptr*, size_t, size_t, CUdeviceptr, unsigned long long);
  CUresult (*MemAddressFree)(CUdeviceptr, size_t);
  CUresult (*MemMap)(CUdeviceptr, size_t, size_t, CUmemGenericAllocationHandle, unsigned long long);
  CUresult (*MemUnmap)(CUdeviceptr, size_t);
  CUresult (*MemSetAccess)(CUdeviceptr, size_t, const CUmemAccessDesc*, size_t);
  CUresult (*MemGetAllocationGranularity)(size_t*, const CUmemAllocationProp*, CUmemAllocationGranularity_flags);
  CUresult (*MemExportToShareableHandle)(void*, CUmemGenericAllocationHandle, CUmemAllocationHandleType, unsigned long long);
  CUresult (*MemImportFromShareableHandle)(CUmemGenericAllocationHandle*, void*, CUmemAllocationHandleType);
  CUresult (*MulticastCreate)(CUmemGenericAllocationHandle*, const CUmulticastObjectProp*);
  CUresult (*MulticastAddDevice)(CUmemGenericAllocationHandle, CUdevice);
  CUresult (*MulticastBindMem)(CUmemGenericAllocationHandle, size_t, CUmemGenericAllocationHandle, size_t, size_t, unsigned long long);
  CUresult (*MulticastUnbind)(CUmemGenericAllocationHandle, CUdevice, size_t, size_t);
  CUresult (*MulticastGetGranularity)(size_t*, const CUmulticastObjectProp*, CUmulticastGranularity_flags);
  CUresult (*DeviceGet)(CUdevice*, int);
  CUresult (*DeviceGetAttribute)(int*, CUdevice_attribute, CUdevice);
  CUresult (*GetErrorString)(CUresult, const char**);
};

static DriverApi& drv() {
  static DriverApi d;
  static bool tried = false;
  if (tried) return d;
  tried = true;
  bool ok = true;
  auto get = [&](const char* name, void** fn) {
    cudaDriverEntryPointQueryResult q;
    if (cudaGetDriverEntryPoint(name, fn, cudaEnableDefault, &q) != cudaSuccess || q != cudaDriverEntryPointSuccess || !*fn)
      ok = false;
  };
  get("cuMemCreate", reinterpret_cast<void**>(&d.MemCreate));
  get("cuMemRelease", reinterpret_cast<void**>(&d.MemRelease));
  get("cuMemAddressReserve", reinterpret_cast<void**>(&d.MemAddressReserve));
  get("cuMemAddressFree", reinterpret_cast<void**>(&d.MemAddressFree));
  get("cuMemMap", reinterpret_cast<void**>(&d.MemMap));
  get("cuMemUnmap", reinterpret_cast<void**>(&d.MemUnmap));
  get("cuMemSetAccess", reinterpret_cast<void**>(&d.MemSetAccess));
  get("cuMemGetAllocationGranularity", reinterpret_cast<void**>(&d.MemGetAllocationGranularity));
  get("cuMemExportToShareableHandle", reinterpret_cast<void**>(&d.MemExportToShareableHandle));
  get("cuMemImportFromShareableHandle", reinterpret_cast<void**>(&d.MemImportFromShareableHandle));
  get("cuMulticastCreate", reinterpret_cast<void**>(&d.MulticastCreate));
  get("cuMulticastAddDevice", reinterpret_cast<void**>(&d.MulticastAddDevice));
  get("cuMulticastBindMem", reinterpret_cast<void**>(&d.MulticastBindMem));
  get("cuMulticastUnbind", reinterpret_cast<void**>(&d.MulticastUnbind));
  get("cuMulticastGetGranularity", reinterpret_cast<void**>(&d.MulticastGetGranularity));
  get("cuDeviceGet", reinterpret_cast<void**>(&d.DeviceGet));
  get("cuDeviceGetAttribute", reinterpret_cast<void**>(&d.DeviceGetAttribute));
  get("cuGetErrorString", reinterpret_cast<void**>(&d.GetErrorString));
  d.ok = ok;
  return d;
}

static const char* cu_err(CUresult r) {
  const char* s = nullptr;
  if (drv().GetErrorString && drv().GetErrorString(r, &s) == CUDA_SUCCESS && s) return s;
  return "unknown driver error";
}

#define DMLP_CU_CHECK(expr)                                                                      \
  do {                                                                                           \
    CUresult _r = (expr);                                                                        \
    if (_r != CUDA_SUCCESS)                                                                      \
      return set_error(_r == CUDA_ERROR_NOT_SUPPORTED || _r == CUDA_ERROR_NOT_PERMITTED ? DMLP_ERR_UNSUPPORTED \
                                                                                          : DMLP_ERR_CUDA,  \
                       "%s failed: %s (%d)", #expr, cu_err(_r), (int)_r);                        \
  } while (0)

struct NvlsBinding {
  int dev = -1;
  CUmemGenericAllocationHandle phys = 0;
  CUdeviceptr uc_va = 0;
};

struct Nvls {
  int world = 0;
  size_t bytes = 0;  // rounded to the multicast granularity
  CUmemGenericAllocationHandle mc = 0;
  CUdeviceptr mc_va = 0;
  std::vector<NvlsBinding> bound;
  std::vector<int> devices;
  int export_fd = -1;
};

static int mc_prop(int world, size_t bytes, CUmulticastObjectProp* p) {
  memset(p, 0, sizeof(*p));
  p->numDevices = world;
  p->size = bytes;
  p->handleTypes = CU_MEM_HANDLE_TYPE_POSIX_FILE_DESCRIPTOR;
  p->flags = 0;
  return DMLP_OK;
}

static int round_to_granularity(int world, size_t* bytes) {
  CUmulticastObjectProp p;
  mc_prop(world, *bytes, &p);
  size_t g = 0;
  DMLP_CU_CHECK(drv().MulticastGetGranularity(&g, &p, CU_MULTICAST_GRANULARITY_RECOMMENDED));
  if (g == 0) g = 2 << 20;
  *bytes = (*bytes + g - 1) / g * g;
  return DMLP_OK;
}

// ---------------------------------------------------------------- multimem kernel
__device__ __forceinline__ float4 multimem_ld_reduce_add_f32x4(const void* mc) {
  float4 v;
  asm volatile("multimem.ld_reduce.relaxed.sys.global.add.v4.f32 {%0, %1, %2, %3}, [%4];"
               : "=f"(v.x), "=f"(v.y), "=f"(v.z), "=f"(v.w) : "l"(mc) : "memory");
  return v;
}
__device__ __forceinline__ float multimem_ld_reduce_add_f32(const void* mc) {
  float v;
  asm volatile("multimem.ld_reduce.relaxed.sys.global.add.f32 %0, [%1];" : "=f"(v) : "l"(mc) : "memory");
  return v;
}
__device__ __forceinline__ void multimem_st_f32x4(void* mc, float4 v) {
  asm volatile("multimem.st.relaxed.sys.global.v4.f32 [%0], {%1, %2, %3, %4};" ::"l"(mc), "f"(v.x), "f"(v.y),
               "f"(v.z), "f"(v.w) : "memory");
}
__device__ __forceinline__ void multimem_st_f32(void* mc, float v) {
  asm volatile("multimem.st.relaxed.sys.global.f32 [%0], %1;" ::"l"(mc), "f"(v) : "memory");
}
__device__ __forceinline__ void multimem_st_b32x2(void* mc, uint32_t a, uint32_t b) {
  asm volatile("multimem.st.relaxed.sys.global.v2.f32 [%0], {%1, %2};" ::"l"(mc), "f"(__uint_as_float(a)),
               "f"(__uint_as_float(b)) : "memory");
}

struct NvlsArgs {
  int world, rank;
  unsigned long long count;
  const float* mc_in;    // multicast address of the inputs (gradients)
  float* mc_out;         // multicast address of the outputs (weights when fuse_sgd)
  void* mc_out_bf16;     // optional multicast address of the bf16 weight shadow
  const float* uc_out;   // this rank's own replica of the output (read for the SGD update)
  const float* lr;
  float scale;
  int fuse_sgd;
  // flag protocol (same pads as the other algorithms)
  uint32_t* pad[DMLP_MAX_RANKS];
  uint32_t* state;
  uint32_t* status;
  unsigned long long timeout_ns;
  uint32_t dead_mask;
  int oneshot;  // every rank reduces EVERYTHING through the switch and updates only its own replica (no broadcast)
};

constexpr int kNvlsThreads = 512;
constexpr int kNvlsMaxBlocks = 128;
constexpr int kNvlsPadBarrier = 0;  // same layout as collectives.cu: [block][src rank]
constexpr int kNvlsPadAbort = 128 * DMLP_MAX_RANKS + 2 * 128;  // == kPadAbort in collectives.cu

__device__ __forceinline__ bool nvls_wait(const uint32_t* flag, uint32_t target, const NvlsArgs& a, int peer, int phase,
                                          unsigned long long deadline) {
  uint32_t spins = 0;
  while (static_cast<int32_t>(ld_relaxed_sys(flag) - target) < 0) {
    if ((++spins & 0x3F) == 0) {
      const uint32_t ab = ld_relaxed_sys(a.pad[a.rank] + kNvlsPadAbort);
      if (ab != 0u) { record_abort(a.status, a.rank, ab, static_cast<uint32_t>(phase)); return false; }
      if (globaltimer_ns() > deadline) {
        a.status[1] = a.rank; a.status[2] = peer; a.status[3] = phase;
        __threadfence_system();
        a.status[0] = DMLP_ERR_TIMEOUT;
        for (int p = 0; p < a.world; ++p) st_relaxed_sys(a.pad[p] + kNvlsPadAbort, abort_word_for(a.rank, peer));
        __threadfence_system();
        return false;
      }
    }
  }
  __threadfence_system();  // multimem loads are not volatile peer loads: keep the acquire here
  return true;
}

__device__ __forceinline__ bool nvls_barrier(const NvlsArgs& a, uint32_t epoch, int phase, unsigned long long deadline,
                                             int* s_fail, bool release = true) {
  __syncthreads();
  if (threadIdx.x < a.world) {
    const int p = threadIdx.x;
    if (release) __threadfence_system();
    st_relaxed_sys(a.pad[p] + kNvlsPadBarrier + blockIdx.x * DMLP_MAX_RANKS + a.rank, epoch);
    if (!nvls_wait(a.pad[a.rank] + kNvlsPadBarrier + blockIdx.x * DMLP_MAX_RANKS + p, epoch, a, p, phase, deadline))
      *s_fail = 1;
  }
  __syncthreads();
  return *s_fail == 0;
}

__global__ void __launch_bounds__(kNvlsThreads) allreduce_nvls_kernel(const __grid_constant__ NvlsArgs a) {
  asm volatile("griddepcontrol.launch_dependents;" ::: "memory");  // a PDL successor may begin its prologue now
  asm volatile("griddepcontrol.wait;" ::: "memory");               // predecessors' gradients are complete and visible
  if ((a.dead_mask >> a.rank) & 1u) return;
  __shared__ int s_fail;
  __shared__ uint32_t s_epoch;
  if (threadIdx.x == 0) { s_fail = 0; s_epoch = a.state[blockIdx.x]; }
  __syncthreads();
  uint32_t epoch = s_epoch;
  const unsigned long long deadline = globaltimer_ns() + a.timeout_ns;
  if (!nvls_barrier(a, ++epoch, 1, deadline, &s_fail, false)) return;  // every replica holds its gradients

  const size_t nfull = a.count / 4;
  const int tail = static_cast<int>(a.count % 4);
  if (a.oneshot) {
    // Latency-bound sizes: one rendezvous, then every rank pulls the switch-reduced sum of ALL elements
    // (inbound traffic = 1 x bytes instead of the (n-1) x bytes of a peer-memory one-shot) and updates its own
    // replica. No broadcast, no trailing rendezvous: the caller double-buffers the inputs by step parity, and a
    // peer cannot reach the next-but-one step before this rank has passed the next rendezvous.
    const float step1 = a.fuse_sgd ? a.lr[0] * a.scale : 0.f;
    float* out = const_cast<float*>(a.uc_out);
    const size_t stride1 = static_cast<size_t>(gridDim.x) * blockDim.x;
    for (size_t i = blockIdx.x * static_cast<size_t>(blockDim.x) + threadIdx.x; i < nfull; i += stride1) {
      float4 v = multimem_ld_reduce_add_f32x4(a.mc_in + 4 * i);
      if (a.fuse_sgd) {
        const float4 w = *reinterpret_cast<const float4*>(out + 4 * i);
        v = make_float4(w.x - step1 * v.x, w.y - step1 * v.y, w.z - step1 * v.z, w.w - step1 * v.w);
      } else if (a.scale != 1.f) {
        v = make_float4(v.x * a.scale, v.y * a.scale, v.z * a.scale, v.w * a.scale);
      }
      *reinterpret_cast<float4*>(out + 4 * i) = v;
    }
    if (tail && blockIdx.x == 0 && threadIdx.x < tail) {
      const size_t e = nfull * 4 + threadIdx.x;
      const float v = multimem_ld_reduce_add_f32(a.mc_in + e);
      out[e] = a.fuse_sgd ? out[e] - step1 * v : v * a.scale;
    }
    if (threadIdx.x == 0) a.state[blockIdx.x] = epoch;
    return;
  }
  const size_t per = (nfull + a.world - 1) / a.world;
  const size_t lo = static_cast<size_t>(a.rank) * per;
  const size_t hi = (lo + per < nfull) ? lo + per : nfull;
  const float step = a.fuse_sgd ? a.lr[0] * a.scale : 0.f;
  const size_t stride = static_cast<size_t>(gridDim.x) * blockDim.x;
  for (size_t i = lo + blockIdx.x * static_cast<size_t>(blockDim.x) + threadIdx.x; i < hi; i += stride) {
    float4 v = multimem_ld_reduce_add_f32x4(a.mc_in + 4 * i);  // the switch returns the sum over all replicas
    if (a.fuse_sgd) {
      const float4 w = *reinterpret_cast<const float4*>(a.uc_out + 4 * i);
      v = make_float4(w.x - step * v.x, w.y - step * v.y, w.z - step * v.z, w.w - step * v.w);
      if (a.mc_out_bf16) {
        __nv_bfloat162 l2 = __floats2bfloat162_rn(v.x, v.y), h2 = __floats2bfloat162_rn(v.z, v.w);
        multimem_st_b32x2(static_cast<unsigned short*>(a.mc_out_bf16) + 4 * i, *reinterpret_cast<uint32_t*>(&l2),
                          *reinterpret_cast<uint32_t*>(&h2));
      }
    } else if (a.scale != 1.f) {
      v = make_float4(v.x * a.scale, v.y * a.scale, v.z * a.scale, v.w * a.scale);
    }
    multimem_st_f32x4(a.mc_out + 4 * i, v);  // one store updates every replica
  }
  if (tail && a.rank == 0 && blockIdx.x == 0 && threadIdx.x < tail) {
    const size_t e = nfull * 4 + threadIdx.x;
    float v = multimem_ld_reduce_add_f32(a.mc_in + e);
    if (a.fuse_sgd) {
      v = a.uc_out[e] - step * v;  // (a bf16 shadow requires count % 4 == 0: checked on the host)
    } else {
      v *= a.scale;
    }
    multimem_st_f32(a.mc_out + e, v);
  }
  if (!nvls_barrier(a, ++epoch, 2, deadline, &s_fail)) return;  // every replica has every slice
  if (threadIdx.x == 0) a.state[blockIdx.x] = epoch;
}

}  // namespace dmlp

using namespace dmlp;

extern "C" int dmlp_nvls_supported(int dev, int* supported) {
  DMLP_REQUIRE(supported != nullptr, "null out pointer");
  *supported = 0;
  if (!drv().ok) return DMLP_OK;
  CUdevice d;
  if (drv().DeviceGet(&d, dev) != CUDA_SUCCESS) return DMLP_OK;
  int v = 0;
  if (drv().DeviceGetAttribute(&v, CU_DEVICE_ATTRIBUTE_MULTICAST_SUPPORTED, d) == CUDA_SUCCESS) *supported = v;
  return DMLP_OK;
}

// Creates the multicast object (one rank / the owning process). *fd_out receives a POSIX file
// descriptor that other processes import (-1 is never returned on success).
extern "C" int dmlp_nvls_create(int world, size_t bytes, int* fd_out, dmlp_nvls_t* out) {
  DMLP_REQUIRE(out && world >= 1 && world <= DMLP_MAX_RANKS && bytes > 0, "dmlp_nvls_create: bad arguments");
  if (!drv().ok) return set_error(DMLP_ERR_UNSUPPORTED, "driver lacks the multicast / VMM entry points");
  cudaFree(nullptr);  // make sure a context exists
  int rc = round_to_granularity(world, &bytes);
  if (rc) return rc;
  Nvls* n = new (std::nothrow) Nvls();
  if (!n) return set_error(DMLP_ERR_CUDA, "out of host memory");
  n->world = world;
  n->bytes = bytes;
  CUmulticastObjectProp p;
  mc_prop(world, bytes, &p);
  CUresult r = drv().MulticastCreate(&n->mc, &p);
  if (r != CUDA_SUCCESS) {
    delete n;
    return set_error(r == CUDA_ERROR_NOT_SUPPORTED || r == CUDA_ERROR_NOT_PERMITTED ? DMLP_ERR_UNSUPPORTED : DMLP_ERR_CUDA,
                     "cuMulticastCreate(%d devices, %zu bytes) failed: %s (%d)", world, bytes, cu_err(r), (int)r);
  }
  if (fd_out) {
    int fd = -1;
    r = drv().MemExportToShareableHandle(&fd, n->mc, CU_MEM_HANDLE_TYPE_POSIX_FILE_DESCRIPTOR, 0);
    if (r != CUDA_SUCCESS) {
      drv().MemRelease(n->mc);
      delete n;
      return set_error(DMLP_ERR_CUDA, "cuMemExportToShareableHandle(multicast) failed: %s (%d)", cu_err(r), (int)r);
    }
    n->export_fd = fd;
    *fd_out = fd;
  }
  *out = reinterpret_cast<dmlp_nvls_t>(n);
  return DMLP_OK;
}

extern "C" int dmlp_nvls_import(int world, size_t bytes, int fd, dmlp_nvls_t* out) {
  DMLP_REQUIRE(out && world >= 1 && bytes > 0 && fd >= 0, "dmlp_nvls_import: bad arguments");
  if (!drv().ok) return set_error(DMLP_ERR_UNSUPPORTED, "driver lacks the multicast / VMM entry points");
  cudaFree(nullptr);
  int rc = round_to_granularity(world, &bytes);
  if (rc) return rc;
  Nvls* n = new (std::nothrow) Nvls();
  if (!n) return set_error(DMLP_ERR_CUDA, "out of host memory");
  n->world = world;
  n->bytes = bytes;
  CUresult r = drv().MemImportFromShareableHandle(&n->mc, reinterpret_cast<void*>(static_cast<intptr_t>(fd)),
                                                  CU_MEM_HANDLE_TYPE_POSIX_FILE_DESCRIPTOR);
  if (r != CUDA_SUCCESS) {
    delete n;
    return set_error(DMLP_ERR_CUDA, "cuMemImportFromShareableHandle(multicast fd %d) failed: %s (%d)", fd, cu_err(r), (int)r);
  }
  *out = reinterpret_cast<dmlp_nvls_t>(n);
  return DMLP_OK;
}

extern "C" int dmlp_nvls_size(dmlp_nvls_t h, size_t* bytes) {
  Nvls* n = reinterpret_cast<Nvls*>(h);
  DMLP_REQUIRE(n && bytes, "bad arguments");
  *bytes = n->bytes;
  return DMLP_OK;
}

// Every participating device must be added (by whichever process drives it) before any bind.
extern "C" int dmlp_nvls_add_device(dmlp_nvls_t h, int dev) {
  Nvls* n = reinterpret_cast<Nvls*>(h);
  DMLP_REQUIRE(n, "null nvls handle");
  CUdevice d;
  DMLP_CU_CHECK(drv().DeviceGet(&d, dev));
  DMLP_CU_CHECK(drv().MulticastAddDevice(n->mc, d));
  n->devices.push_back(dev);
  return DMLP_OK;
}

// Allocates this device's replica, binds it to the multicast object and maps both views:
// *uc_ptr is the device's own memory, *mc_ptr the multicast address (same for all local devices).
extern "C" int dmlp_nvls_bind(dmlp_nvls_t h, int dev, void** uc_ptr, void** mc_ptr) {
  Nvls* n = reinterpret_cast<Nvls*>(h);
  DMLP_REQUIRE(n && uc_ptr && mc_ptr, "dmlp_nvls_bind: bad arguments");
  DeviceGuard g(dev);
  cudaFree(nullptr);
  CUmemAllocationProp prop;
  memset(&prop, 0, sizeof(prop));
  prop.type = CU_MEM_ALLOCATION_TYPE_PINNED;
  prop.location.type = CU_MEM_LOCATION_TYPE_DEVICE;
  prop.location.id = dev;
  prop.requestedHandleTypes = CU_MEM_HANDLE_TYPE_POSIX_FILE_DESCRIPTOR;
  NvlsBinding b;
  b.dev = dev;
  DMLP_CU_CHECK(drv().MemCreate(&b.phys, n->bytes, &prop, 0));
  DMLP_CU_CHECK(drv().MemAddressReserve(&b.uc_va, n->bytes, n->bytes > (2u << 20) ? (2u << 20) : 0, 0, 0));
  DMLP_CU_CHECK(drv().MemMap(b.uc_va, n->bytes, 0, b.phys, 0));
  std::vector<CUmemAccessDesc> acc;
  const std::vector<int>& devs = n->devices.empty() ? std::vector<int>{dev} : n->devices;
  bool has_self = false;
  for (int d : devs) {
    CUmemAccessDesc a;
    memset(&a, 0, sizeof(a));
    a.location.type = CU_MEM_LOCATION_TYPE_DEVICE;
    a.location.id = d;
    a.flags = CU_MEM_ACCESS_FLAGS_PROT_READWRITE;
    acc.push_back(a);
    has_self |= d == dev;
  }
  if (!has_self) {
    CUmemAccessDesc a;
    memset(&a, 0, sizeof(a));
    a.location.type = CU_MEM_LOCATION_TYPE_DEVICE;
    a.location.id = dev;
    a.flags = CU_MEM_ACCESS_FLAGS_PROT_READWRITE;
    acc.push_back(a);
  }
  DMLP_CU_CHECK(drv().MemSetAccess(b.uc_va, n->bytes, acc.data(), acc.size()));
  DMLP_CUDA_CHECK(cudaMemset(reinterpret_cast<void*>(b.uc_va), 0, n->bytes));
  DMLP_CU_CHECK(drv().MulticastBindMem(n->mc, 0, b.phys, 0, n->bytes, 0));
  if (n->mc_va == 0) {
    DMLP_CU_CHECK(drv().MemAddressReserve(&n->mc_va, n->bytes, n->bytes > (2u << 20) ? (2u << 20) : 0, 0, 0));
    DMLP_CU_CHECK(drv().MemMap(n->mc_va, n->bytes, 0, n->mc, 0));
  }
  DMLP_CU_CHECK(drv().MemSetAccess(n->mc_va, n->bytes, acc.data(), acc.size()));
  n->bound.push_back(b);
  *uc_ptr = reinterpret_cast<void*>(b.uc_va);
  *mc_ptr = reinterpret_cast<void*>(n->mc_va);
  return DMLP_OK;
}

extern "C" int dmlp_nvls_destroy(dmlp_nvls_t h) {
  Nvls* n = reinterpret_cast<Nvls*>(h);
  if (!n) return DMLP_OK;
  for (auto& b : n->bound) {
    cudaSetDevice(b.dev);
    cudaDeviceSynchronize();
  }
  if (n->mc_va) {
    drv().MemUnmap(n->mc_va, n->bytes);
    drv().MemAddressFree(n->mc_va, n->bytes);
  }
  for (auto& b : n->bound) {
    CUdevice d;
    if (drv().DeviceGet(&d, b.dev) == CUDA_SUCCESS) drv().MulticastUnbind(n->mc, d, 0, n->bytes);
    drv().MemUnmap(b.uc_va, n->bytes);
    drv().MemAddressFree(b.uc_va, n->bytes);
    drv().MemRelease(b.phys);
  }
  if (n->mc) drv().MemRelease(n->mc);
  if (n->export_fd >= 0) close(n->export_fd);
  cudaGetLastError();
  delete n;
  return DMLP_OK;
}

// Implemented in collectives.cu: hands out the communicator's pads / state / watchdog words.
extern "C" int dmlp_comm_internal_view(dmlp_comm_t comm, int local_idx, uint32_t* pad[DMLP_MAX_RANKS], uint32_t** state,
                                       uint32_t** host_flags, unsigned long long* timeout_ns, uint32_t* dead_mask,
                                       int* rank, int* world, int* dev);

extern "C" int dmlp_allreduce_nvls(dmlp_comm_t comm, int local_idx, const dmlp_nvls_allreduce_desc_t* d, void* stream) {
  DMLP_REQUIRE(comm && d, "dmlp_allreduce_nvls: null argument");
  NvtxRange nvtx_ar("dmlp.allreduce_nvls");
  DMLP_REQUIRE(d->mc_in && d->mc_out, "dmlp_allreduce_nvls: multicast addresses are required");
  DMLP_REQUIRE(((reinterpret_cast<uintptr_t>(d->mc_in) | reinterpret_cast<uintptr_t>(d->mc_out) |
                 reinterpret_cast<uintptr_t>(d->uc_out) | reinterpret_cast<uintptr_t>(d->mc_out_bf16)) & 15) == 0,
               "dmlp_allreduce_nvls: buffers must be 16-byte aligned");
  if (d->fuse_sgd) DMLP_REQUIRE(d->uc_out && d->lr, "dmlp_allreduce_nvls: fused SGD needs the local replica and lr");
  if (d->mc_out_bf16) DMLP_REQUIRE(d->count % 4 == 0, "dmlp_allreduce_nvls: a bf16 shadow needs count to be a multiple of 4");
  if (d->count == 0) return DMLP_OK;
  NvlsArgs a;
  memset(&a, 0, sizeof(a));
  uint32_t* host_flags = nullptr;
  uint32_t* state = nullptr;
  int dev = 0;
  int rc = dmlp_comm_internal_view(comm, local_idx, a.pad, &state, &host_flags, &a.timeout_ns, &a.dead_mask, &a.rank,
                                   &a.world, &dev);
  if (rc) return rc;
  volatile uint32_t* hf = host_flags;
  if (hf[4] != 0 || hf[0] != 0) return set_error(DMLP_ERR_ABORTED, "communicator is FAILED");
  a.state = state;
  a.status = host_flags + 4;
  a.count = d->count;
  a.mc_in = static_cast<const float*>(d->mc_in);
  a.mc_out = static_cast<float*>(d->mc_out);
  a.mc_out_bf16 = d->mc_out_bf16;
  a.uc_out = static_cast<const float*>(d->uc_out);
  a.lr = d->lr;
  a.scale = d->scale != 0.f ? d->scale : 1.f;
  a.fuse_sgd = d->fuse_sgd;
  // one-shot flavour: outputs land in the caller's own replica only (uc_out), so it needs one and cannot feed a shadow
  a.oneshot = (d->oneshot && d->uc_out != nullptr && d->mc_out_bf16 == nullptr) ? 1 : 0;
  const size_t work = a.oneshot ? d->count / 4 : (d->count / 4 + a.world - 1) / a.world;
  // one-shot: one 16-byte pack per thread (a switch-reduced load is a ~3 us round trip: no serial loop over packs)
  int blocks = a.oneshot ? static_cast<int>((work + kNvlsThreads - 1) / kNvlsThreads)
                         : static_cast<int>((work + kNvlsThreads * 2 - 1) / (kNvlsThreads * 2));
  if (blocks < 1) blocks = 1;
  if (blocks > kNvlsMaxBlocks) blocks = kNvlsMaxBlocks;
  DeviceGuard guard(dev);
  {
    // Programmatic stream serialization: may be scheduled while its predecessor drains (griddepcontrol.wait inside)
    cudaLaunchConfig_t cfg = {};
    cfg.gridDim = dim3(blocks);
    cfg.blockDim = dim3(kNvlsThreads);
    cfg.stream = as_stream(stream);
    cudaLaunchAttribute attr[1];
    attr[0].id = cudaLaunchAttributeProgrammaticStreamSerialization;
    attr[0].val.programmaticStreamSerializationAllowed = 1;
    cfg.attrs = attr;
    cfg.numAttrs = 1;
    DMLP_CUDA_CHECK(cudaLaunchKernelEx(&cfg, allreduce_nvls_kernel, a));
  }
  DMLP_CUDA_CHECK(cudaGetLastError());
  return DMLP_OK;
}
