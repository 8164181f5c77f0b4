// Peer-memory collectives over NVLink 5 / NVSwitch, written for sm_100a. No NCCL.
//
// Replaces the reference's coordinator-driven ring all-reduce
// (DSML/gpu_coordinator_service/gpu_coordinator_server.go:272-566: goroutine per device, five
// gRPCs per ring step, byte-wise add on the host) and its naive gather/reduce/broadcast mode
// (:611-717) with GPU kernels that read and write peer buffers directly:
//   one-shot : every rank loads all peers' inputs and reduces in registers (latency-bound sizes)
//   two-shot : reduce-scatter by peer loads, then all-gather by peer stores
//   ring     : 2(n-1) chunked steps through per-rank staging slots with data/ack sequence flags
// Every variant can fuse the gradient average and the SGD update (client.go:254-267) so that no
// separate elementwise kernel runs:  w -= lr * scale * sum_r g_r   (lr read from device memory).
//
// Synchronisation is a flag protocol in peer-visible "signal pads": same-index CTAs of all ranks
// meet at barriers using monotonically increasing epochs (relaxed sys-scope flags, see wait_flag). The
// epochs live in device memory, so a captured CUDA graph can be replayed. Every spin-wait has a
// globaltimer deadline and also polls the rank's abort word (device memory): a dead peer makes the kernel
// record {rank, peer, phase} in a host-mapped status block and exit instead of hanging the GPU.
//
// "Emulation": when all ranks of a communicator live on ONE device the same kernels are launched
// once, cooperatively, with gridDim.y = world and rank = blockIdx.y. That is how multi-rank logic
// is tested on a single GPU and how N logical devices are served from one physical GPU.
#include "common.h"
#include "ptx_sm100.cuh"

#include <cooperative_groups.h>
#include <stdlib.h>
#include <string.h>

#include <stdio.h>

#include <mutex>
#include <new>
#include <type_traits>
#include <vector>

namespace dmlp {

constexpr int kMaxRanks = DMLP_MAX_RANKS;
constexpr int kMaxCollBlocks = 128;
constexpr int kCollThreads = 512;
constexpr int kInplacePerThread = 4;

// Signal pad layout in 32-bit words (one pad per rank, peer-visible).
constexpr int kPadBarrier = 0;                                 // [block][src rank]
constexpr int kPadRingData = kMaxCollBlocks * kMaxRanks;       // [block] data sequence (written by prev)
constexpr int kPadRingAck = kPadRingData + kMaxCollBlocks;     // [block] ack sequence (written by next)
constexpr int kPadAbort = kPadRingAck + kMaxCollBlocks;         // [1] abort word: set by a failing peer or by the host
constexpr int kPadWords = kPadAbort + 8;
constexpr size_t kPadBytes = 64 * 1024;                        // reserved at the start of a segment
static_assert(kPadWords * 4 <= kPadBytes, "signal pad overflow");
static_assert(kPadAbort == kPadAbortWord && kPadWords <= kPadFusedFlags && (kPadFusedEpoch + 1) * 4 <= kPadBytes,
              "common.h pad words out of sync with this layout");
// Local (never peer-accessed) per-rank state: barrier epochs then ring sequence numbers.
constexpr int kStateEpoch = 0;
constexpr int kStateRingSeq = kMaxCollBlocks;
constexpr int kStateWords = 2 * kMaxCollBlocks;

enum { PHASE_BARRIER_IN = 1, PHASE_BARRIER_OUT = 2, PHASE_RING_DATA = 3, PHASE_RING_ACK = 4 };

struct CollArgs {
  int world;
  int rank;  // < 0: emulation, rank = blockIdx.y
  int op;
  int skip_end_barrier;
  int inplace;  // one-shot with out == in: reduce into registers, rendezvous, then store
  unsigned long long count;  // elements
  const void* in[kMaxRanks];
  void* out[kMaxRanks];
  void* out_bf16[kMaxRanks];
  uint32_t* pad[kMaxRanks];
  uint8_t* staging[kMaxRanks];
  unsigned long long staging_bytes;
  uint32_t* state[kMaxRanks];
  const float* lr[kMaxRanks];
  float scale;
  uint32_t dead_mask;  // fault injection: these ranks never show up
  uint32_t* status;      // host-mapped, written only on failure: {code, rank, peer, phase}
  unsigned long long timeout_ns;
  // optional phase timestamps: [block][4] x u64 globaltimer ns {entry, after barrier-in, after data, exit}
  unsigned long long* prof;
  uint32_t strict;  // env DMLP_COLL_STRICT: 2 (default) flag polls are acquire loads; 1 adds system fences both sides; 0 relaxed polls
};

#define DMLP_PROF(slot)                                                               \
  do {                                                                                \
    if (a.prof != nullptr && threadIdx.x == 0 && a.rank >= 0) a.prof[blockIdx.x * 4 + (slot)] = globaltimer_ns(); \
  } while (0)

// ------------------------------------------------------------------ 16-byte packs
template <typename T> struct Pack { static constexpr int N = 16 / sizeof(T); };

__device__ __forceinline__ uint4 ld_pack(const void* base, size_t i) {
  // volatile: never served from a stale L1 line (peer / staging data changes during the kernel)
  uint4 v;
  const uint4* p = reinterpret_cast<const uint4*>(base) + i;
  asm volatile("ld.volatile.global.v4.u32 {%0, %1, %2, %3}, [%4];"
               : "=r"(v.x), "=r"(v.y), "=r"(v.z), "=r"(v.w) : "l"(p) : "memory");
  return v;
}
__device__ __forceinline__ void st_pack(void* base, size_t i, uint4 v) {
  reinterpret_cast<uint4*>(base)[i] = v;
}
template <typename T>
__device__ __forceinline__ uint4 ld_pack_partial(const void* base, size_t i, int valid) {
  union { uint4 v; T e[Pack<T>::N]; } u;
  u.v = make_uint4(0, 0, 0, 0);
  const volatile T* p = reinterpret_cast<const volatile T*>(base) + i * Pack<T>::N;
  for (int k = 0; k < valid; ++k) {
    if constexpr (sizeof(T) == 2) {
      unsigned short raw = reinterpret_cast<const volatile unsigned short*>(p)[k];
      reinterpret_cast<unsigned short*>(u.e)[k] = raw;
    } else {
      u.e[k] = p[k];
    }
  }
  return u.v;
}
template <typename T>
__device__ __forceinline__ void st_pack_partial(void* base, size_t i, uint4 v, int valid) {
  union { uint4 v; T e[Pack<T>::N]; } u;
  u.v = v;
  T* p = reinterpret_cast<T*>(base) + i * Pack<T>::N;
  for (int k = 0; k < valid; ++k) p[k] = u.e[k];
}

template <typename T> __device__ __forceinline__ T red1(T a, T b, int op);
template <> __device__ __forceinline__ float red1<float>(float a, float b, int op) {
  switch (op) { case DMLP_PROD: return a * b; case DMLP_MIN: return fminf(a, b); case DMLP_MAX: return fmaxf(a, b); default: return a + b; }
}
template <> __device__ __forceinline__ int red1<int>(int a, int b, int op) {
  switch (op) { case DMLP_PROD: return a * b; case DMLP_MIN: return min(a, b); case DMLP_MAX: return max(a, b); default: return a + b; }
}
template <> __device__ __forceinline__ unsigned char red1<unsigned char>(unsigned char a, unsigned char b, int op) {
  switch (op) {  // SUM wraps modulo 256 like the reference's byte adds (gpu_coordinator_server.go:541-543)
    case DMLP_PROD: return static_cast<unsigned char>(a * b);
    case DMLP_MIN: return a < b ? a : b;
    case DMLP_MAX: return a > b ? a : b;
    default: return static_cast<unsigned char>(a + b);
  }
}
template <> __device__ __forceinline__ __nv_bfloat16 red1<__nv_bfloat16>(__nv_bfloat16 a, __nv_bfloat16 b, int op) {
  return __float2bfloat16_rn(red1<float>(__bfloat162float(a), __bfloat162float(b), op));
}

template <typename T>
__device__ __forceinline__ uint4 red_pack(uint4 a, uint4 b, int op) {
  union U { uint4 v; T e[Pack<T>::N]; };
  U x, y;
  x.v = a; y.v = b;
#pragma unroll
  for (int k = 0; k < Pack<T>::N; ++k) x.e[k] = red1<T>(x.e[k], y.e[k], op);
  return x.v;
}
template <typename T>
__device__ __forceinline__ uint4 scale_pack(uint4 a, float s) {
  if constexpr (sizeof(T) == 4 && Pack<T>::N == 4) {
    union { uint4 v; T e[4]; } x;
    x.v = a;
#pragma unroll
    for (int k = 0; k < 4; ++k) x.e[k] = static_cast<T>(static_cast<float>(x.e[k]) * s);
    return x.v;
  } else if constexpr (sizeof(T) == 2) {
    union { uint4 v; __nv_bfloat16 e[8]; } x;
    x.v = a;
#pragma unroll
    for (int k = 0; k < 8; ++k) x.e[k] = __float2bfloat16_rn(__bfloat162float(x.e[k]) * s);
    return x.v;
  } else {
    return a;
  }
}
// w - step * g on a float4 pack; optionally produces the bf16 shadow (8 bytes).
__device__ __forceinline__ uint4 sgd_pack(uint4 w, uint4 g, float step, uint2* shadow) {
  union { uint4 v; float e[4]; } a, b;
  a.v = w; b.v = g;
#pragma unroll
  for (int k = 0; k < 4; ++k) a.e[k] = a.e[k] - step * b.e[k];
  if (shadow) {
    __nv_bfloat162 lo = __floats2bfloat162_rn(a.e[0], a.e[1]), hi = __floats2bfloat162_rn(a.e[2], a.e[3]);
    shadow->x = *reinterpret_cast<uint32_t*>(&lo);
    shadow->y = *reinterpret_cast<uint32_t*>(&hi);
  }
  return a.v;
}
__device__ __forceinline__ void st_shadow(void* base16, size_t i, uint2 s, int valid) {
  if (valid >= 4) {
    reinterpret_cast<uint2*>(base16)[i] = s;
  } else {
    unsigned short* p = reinterpret_cast<unsigned short*>(base16) + i * 4;
    const unsigned short e[4] = {static_cast<unsigned short>(s.x & 0xffff), static_cast<unsigned short>(s.x >> 16),
                                 static_cast<unsigned short>(s.y & 0xffff), static_cast<unsigned short>(s.y >> 16)};
    for (int k = 0; k < valid; ++k) p[k] = e[k];
  }
}

// ------------------------------------------------------------------ watchdog + flags
__device__ __forceinline__ void report_failure(const CollArgs& a, int rank, int peer, int phase) {
  a.status[1] = static_cast<uint32_t>(rank);
  a.status[2] = static_cast<uint32_t>(peer);
  a.status[3] = static_cast<uint32_t>(phase);
  __threadfence_system();
  a.status[0] = DMLP_ERR_TIMEOUT;
  // Release everybody quickly: the abort word sits in each rank's (peer-visible) signal pad.
  for (int p = 0; p < a.world; ++p) st_relaxed_sys(a.pad[p] + kPadAbort, abort_word_for(rank, peer));
  __threadfence_system();
}

// Spins until *flag >= target (wrap-safe). False on abort or deadline.
// Cost notes (measured with the in-kernel stamps, profiles/r1_allreduce_phase_experiments.md):
//  * polling a HOST-mapped abort flag from many CTAs serialises on PCIe and stalls barriers by
//    40-230 us -> the abort word lives in device memory (this rank's signal pad);
//  * a system-scope fence (MEMBAR.SYS) costs ~3 us. The poll is a relaxed sys-scope load and no
//    acquire fence follows: every read of peer-written data in these kernels is a volatile
//    (sys-scope, L1-bypassing) load issued after the flag was observed (control dependency +
//    bar.sync), and data a peer pushed into local memory is only read by later kernels.
//    Since round 2 the poll itself is an acquire load by default (+0.05 us per rendezvous); DMLP_COLL_STRICT=1
//    restores the fence, =0 the relaxed poll.
__device__ __forceinline__ bool wait_flag(const uint32_t* flag, uint32_t target, const CollArgs& a,
                                          int rank, int peer, int phase, unsigned long long deadline) {
  uint32_t spins = 0;
  while (static_cast<int32_t>((a.strict == 2 ? ld_acquire_sys(flag) : ld_relaxed_sys(flag)) - target) < 0) {
    if ((++spins & 0x3F) == 0) {
      const uint32_t ab = ld_relaxed_sys(a.pad[rank] + kPadAbort);
      if (ab != 0u) { record_abort(a.status, rank, ab, static_cast<uint32_t>(phase)); return false; }
      if (globaltimer_ns() > deadline) {
        report_failure(a, rank, peer, phase);
        return false;
      }
    }
  }
  if (a.strict == 1) __threadfence_system();
  return true;
}

// Barrier between the CTAs with this blockIdx.x on every rank. Returns false if it failed; the
// result is uniform across the CTA. `release` = this CTA wrote data (to peers) in this kernel that
// the peers may read once they pass: a system fence precedes the flag store. The entry barrier of a
// kernel needs no release: everything earlier kernels wrote is visible at the kernel boundary.
__device__ __forceinline__ bool coll_barrier(const CollArgs& a, int rank, uint32_t epoch, int phase,
                                             unsigned long long deadline, int* s_fail, bool release = true) {
  __syncthreads();
  if (threadIdx.x < a.world) {
    const int p = threadIdx.x;
    if (release || a.strict == 1) __threadfence_system();  // (strict mode: every barrier releases, entry barriers included)
    st_relaxed_sys(a.pad[p] + kPadBarrier + blockIdx.x * kMaxRanks + rank, epoch);
    if (!wait_flag(a.pad[rank] + kPadBarrier + blockIdx.x * kMaxRanks + p, epoch, a, rank, p, phase, deadline))
      *s_fail = 1;
  }
  __syncthreads();
  return *s_fail == 0;
}

struct Prologue {
  int rank;
  uint32_t epoch;
  unsigned long long deadline;
};

// ------------------------------------------------------------------ one-shot
template <typename T, bool SGD>
__global__ void __launch_bounds__(kCollThreads)
allreduce_oneshot_kernel(const __grid_constant__ CollArgs a) {
  asm volatile("griddepcontrol.launch_dependents;" ::: "memory");  // a PDL successor may begin its prologue now
  asm volatile("griddepcontrol.wait;" ::: "memory");               // predecessors' gradients are complete and visible
  constexpr int V = Pack<T>::N;
  const int rank = a.rank >= 0 ? a.rank : blockIdx.y;
  if ((a.dead_mask >> rank) & 1u) return;
  __shared__ int s_fail;
  __shared__ uint32_t s_epoch;
  if (threadIdx.x == 0) {
    s_fail = 0;
    s_epoch = a.state[rank][kStateEpoch + blockIdx.x];
  }
  __syncthreads();
  uint32_t epoch = s_epoch;
  const unsigned long long deadline = globaltimer_ns() + a.timeout_ns;
  DMLP_PROF(0);

  if (!coll_barrier(a, rank, ++epoch, PHASE_BARRIER_IN, deadline, &s_fail, /*release=*/false)) return;
  DMLP_PROF(1);

  const size_t nfull = a.count / V;
  const int tail = static_cast<int>(a.count % V);
  const size_t npack = nfull + (tail ? 1 : 0);
  float step = 0.f;
  if constexpr (SGD) step = a.lr[rank][0] * a.scale;
  const size_t stride = static_cast<size_t>(gridDim.x) * blockDim.x;
  const size_t tid = blockIdx.x * static_cast<size_t>(blockDim.x) + threadIdx.x;
  if (a.inplace) {
    // out aliases in: nobody may store before every rank has finished loading. The host sized the
    // grid so that each thread owns at most kInplacePerThread packs, which wait in registers.
    uint4 acc[kInplacePerThread];
#pragma unroll
    for (int k = 0; k < kInplacePerThread; ++k) {
      const size_t i = tid + k * stride;
      if (i < npack) {
        const bool partial = (i == nfull);
        acc[k] = partial ? ld_pack_partial<T>(a.in[0], i, tail) : ld_pack(a.in[0], i);
        for (int p = 1; p < a.world; ++p)
          acc[k] = red_pack<T>(acc[k], partial ? ld_pack_partial<T>(a.in[p], i, tail) : ld_pack(a.in[p], i), a.op);
        if (a.scale != 1.f) acc[k] = scale_pack<T>(acc[k], a.scale);
      }
    }
    if (!coll_barrier(a, rank, ++epoch, PHASE_BARRIER_OUT, deadline, &s_fail, /*release=*/false)) return;
#pragma unroll
    for (int k = 0; k < kInplacePerThread; ++k) {
      const size_t i = tid + k * stride;
      if (i < npack) {
        if (i == nfull) st_pack_partial<T>(a.out[rank], i, acc[k], tail);
        else st_pack(a.out[rank], i, acc[k]);
      }
    }
    if (threadIdx.x == 0) a.state[rank][kStateEpoch + blockIdx.x] = epoch;
    return;
  }
  for (size_t i = tid; i < npack; i += stride) {
    const bool partial = (i == nfull);
    const int valid = partial ? tail : V;
    uint4 v[kMaxRanks];
#pragma unroll
    for (int p = 0; p < kMaxRanks; ++p)
      if (p < a.world) v[p] = partial ? ld_pack_partial<T>(a.in[p], i, valid) : ld_pack(a.in[p], i);
    uint4 acc = v[0];
#pragma unroll
    for (int p = 1; p < kMaxRanks; ++p)
      if (p < a.world) acc = red_pack<T>(acc, v[p], a.op);  // fixed order: replicas stay bit-identical
    if constexpr (SGD) {
      uint4 w = partial ? ld_pack_partial<T>(a.out[rank], i, valid) : ld_pack(a.out[rank], i);
      uint2 sh;
      acc = sgd_pack(w, acc, step, a.out_bf16[rank] ? &sh : nullptr);
      if (a.out_bf16[rank]) st_shadow(a.out_bf16[rank], i, sh, valid);
    } else if (a.scale != 1.f) {
      acc = scale_pack<T>(acc, a.scale);
    }
    if (partial) st_pack_partial<T>(a.out[rank], i, acc, valid);
    else st_pack(a.out[rank], i, acc);
  }

  __syncthreads();
  DMLP_PROF(2);
  // Peers may not overwrite their inputs until everyone has read them; callers that
  // double-buffer their inputs can drop this second rendezvous.
  // (no release fence: this rank wrote nothing its peers will read; it only says "done reading")
  if (!a.skip_end_barrier)
    if (!coll_barrier(a, rank, ++epoch, PHASE_BARRIER_OUT, deadline, &s_fail, /*release=*/false)) return;
  if (threadIdx.x == 0) a.state[rank][kStateEpoch + blockIdx.x] = epoch;
  DMLP_PROF(3);
}

// ------------------------------------------------------------------ two-shot
template <typename T, bool SGD>
__global__ void __launch_bounds__(kCollThreads)
allreduce_twoshot_kernel(const __grid_constant__ CollArgs a) {
  asm volatile("griddepcontrol.launch_dependents;" ::: "memory");  // a PDL successor may begin its prologue now
  asm volatile("griddepcontrol.wait;" ::: "memory");               // predecessors' gradients are complete and visible
  constexpr int V = Pack<T>::N;
  const int rank = a.rank >= 0 ? a.rank : blockIdx.y;
  if ((a.dead_mask >> rank) & 1u) return;
  __shared__ int s_fail;
  __shared__ uint32_t s_epoch;
  if (threadIdx.x == 0) {
    s_fail = 0;
    s_epoch = a.state[rank][kStateEpoch + blockIdx.x];
  }
  __syncthreads();
  uint32_t epoch = s_epoch;
  const unsigned long long deadline = globaltimer_ns() + a.timeout_ns;

  if (!coll_barrier(a, rank, ++epoch, PHASE_BARRIER_IN, deadline, &s_fail, /*release=*/false)) return;

  const size_t nfull = a.count / V;
  const int tail = static_cast<int>(a.count % V);
  const size_t npack = nfull + (tail ? 1 : 0);
  const size_t per = (npack + a.world - 1) / a.world;
  const size_t lo = static_cast<size_t>(rank) * per;
  const size_t hi = (lo + per < npack) ? lo + per : npack;
  float step = 0.f;
  if constexpr (SGD) step = a.lr[rank][0] * a.scale;
  const size_t stride = static_cast<size_t>(gridDim.x) * blockDim.x;
  for (size_t i = lo + blockIdx.x * static_cast<size_t>(blockDim.x) + threadIdx.x; i < hi; i += stride) {
    const bool partial = (i == nfull);
    const int valid = partial ? tail : V;
    uint4 v[kMaxRanks];
#pragma unroll
    for (int p = 0; p < kMaxRanks; ++p)
      if (p < a.world) v[p] = partial ? ld_pack_partial<T>(a.in[p], i, valid) : ld_pack(a.in[p], i);
    uint4 acc = v[0];
#pragma unroll
    for (int p = 1; p < kMaxRanks; ++p)
      if (p < a.world) acc = red_pack<T>(acc, v[p], a.op);
    uint2 sh;
    bool has_shadow = false;
    if constexpr (SGD) {
      uint4 w = partial ? ld_pack_partial<T>(a.out[rank], i, valid) : ld_pack(a.out[rank], i);
      has_shadow = a.out_bf16[rank] != nullptr;
      acc = sgd_pack(w, acc, step, has_shadow ? &sh : nullptr);
    } else if (a.scale != 1.f) {
      acc = scale_pack<T>(acc, a.scale);
    }
#pragma unroll
    for (int p = 0; p < kMaxRanks; ++p) {
      if (p < a.world) {
        if (partial) st_pack_partial<T>(a.out[p], i, acc, valid);
        else st_pack(a.out[p], i, acc);
        if (has_shadow) st_shadow(a.out_bf16[p], i, sh, valid);
      }
    }
  }
  if (!coll_barrier(a, rank, ++epoch, PHASE_BARRIER_OUT, deadline, &s_fail)) return;
  if (threadIdx.x == 0) a.state[rank][kStateEpoch + blockIdx.x] = epoch;
}

// ------------------------------------------------------------------ ring
// One persistent kernel per rank runs all 2(n-1) steps (the reference runs each step as five
// gRPCs per device, gpu_coordinator_server.go:379-566). Same-index CTAs of neighbouring ranks form
// independent rings. Data is pushed into the next rank's staging slot (slot = seq & 1), then the
// data flag is released; the consumer acks so the producer may reuse the slot two sends later.
template <typename T, bool SGD>
__global__ void __launch_bounds__(kCollThreads)
allreduce_ring_kernel(const __grid_constant__ CollArgs a) {
  asm volatile("griddepcontrol.wait;" ::: "memory");
  constexpr int V = Pack<T>::N;
  const int rank = a.rank >= 0 ? a.rank : blockIdx.y;
  if ((a.dead_mask >> rank) & 1u) return;
  const int n = a.world;
  const int next = (rank + 1) % n, prev = (rank + n - 1) % n;
  __shared__ int s_fail;
  __shared__ uint32_t s_seq;
  if (threadIdx.x == 0) {
    s_fail = 0;
    s_seq = a.state[rank][kStateRingSeq + blockIdx.x];
  }
  __syncthreads();
  uint32_t seq_send = s_seq, seq_recv = s_seq;
  const unsigned long long deadline = globaltimer_ns() + a.timeout_ns;
  float step = 0.f;
  if constexpr (SGD) step = a.lr[rank][0] * a.scale;

  uint32_t* my_data_flag = a.pad[rank] + kPadRingData + blockIdx.x;
  uint32_t* my_ack_flag = a.pad[rank] + kPadRingAck + blockIdx.x;
  uint32_t* next_data_flag = a.pad[next] + kPadRingData + blockIdx.x;
  uint32_t* prev_ack_flag = a.pad[prev] + kPadRingAck + blockIdx.x;

  const size_t nfull = a.count / V;
  const int tail = static_cast<int>(a.count % V);
  const size_t npack = nfull + (tail ? 1 : 0);
  const size_t slot_packs = a.staging_bytes / 2 / 16;   // capacity of one staging slot
  const size_t slab_packs = slot_packs * n;             // packs all-reduced per outer iteration

  for (size_t slab0 = 0; slab0 < npack; slab0 += slab_packs) {
    const size_t slab_n = (npack - slab0 < slab_packs) ? npack - slab0 : slab_packs;
    const size_t seg = (slab_n + n - 1) / n;            // packs per segment (<= slot_packs)
    const size_t bw = (seg + gridDim.x - 1) / gridDim.x;  // packs of a segment owned by this CTA
    const size_t b_lo = blockIdx.x * bw;
    const size_t b_hi = (b_lo + bw < seg) ? b_lo + bw : seg;

    // step indices: 0 = initial send; 1..n-1 reduce-scatter receives; n..2n-2 all-gather receives
    for (int s = 0; s <= 2 * (n - 1); ++s) {
      const bool has_recv = s >= 1;
      const bool rs = s >= 1 && s <= n - 1;
      const bool last_rs = s == n - 1;
      const bool has_send = s < 2 * (n - 1);
      int segidx;
      if (s == 0) segidx = rank;
      else if (rs) segidx = ((rank - s) % n + n) % n;              // (rank-1-t), t = s-1
      else segidx = ((rank - (s - n)) % n + n) % n;                 // all-gather: (rank - t')
      if (threadIdx.x == 0) {
        if (has_recv && !wait_flag(my_data_flag, seq_recv + 1, a, rank, prev, PHASE_RING_DATA, deadline)) s_fail = 1;
        if (has_send && seq_send + 1 > 2 &&
            !wait_flag(my_ack_flag, seq_send + 1 - 2, a, rank, next, PHASE_RING_ACK, deadline)) s_fail = 1;
      }
      __syncthreads();
      if (s_fail) return;
      const uint4* rslot = reinterpret_cast<const uint4*>(a.staging[rank] + ((seq_recv + 1) & 1) * (a.staging_bytes / 2));
      uint4* sslot = reinterpret_cast<uint4*>(a.staging[next] + ((seq_send + 1) & 1) * (a.staging_bytes / 2));
      const size_t seg_base = slab0 + static_cast<size_t>(segidx) * seg;  // global pack index of segment start
      for (size_t j = b_lo + threadIdx.x; j < b_hi; j += blockDim.x) {
        const size_t i = seg_base + j;
        if (i >= slab0 + slab_n) break;  // ragged last segment
        const bool partial = (i == nfull);
        const int valid = partial ? tail : V;
        uint4 val;
        if (s == 0) {
          val = partial ? ld_pack_partial<T>(a.in[rank], i, valid) : ld_pack(a.in[rank], i);
        } else {
          val = ld_pack(rslot, j);
          if (rs) {
            uint4 mine = partial ? ld_pack_partial<T>(a.in[rank], i, valid) : ld_pack(a.in[rank], i);
            val = red_pack<T>(mine, val, a.op);
            if (last_rs) {
              if constexpr (SGD) {
                uint4 w = partial ? ld_pack_partial<T>(a.out[rank], i, valid) : ld_pack(a.out[rank], i);
                uint2 sh;
                val = sgd_pack(w, val, step, a.out_bf16[rank] ? &sh : nullptr);
                if (a.out_bf16[rank]) st_shadow(a.out_bf16[rank], i, sh, valid);
              } else if (a.scale != 1.f) {
                val = scale_pack<T>(val, a.scale);
              }
            }
          }
          if (!rs || last_rs) {
            if (partial) st_pack_partial<T>(a.out[rank], i, val, valid);
            else st_pack(a.out[rank], i, val);
            if constexpr (SGD) {
              if (!rs && a.out_bf16[rank]) {  // all-gather of updated weights: refresh the shadow too
                union { uint4 v; float e[4]; } u;
                u.v = val;
                __nv_bfloat162 lo = __floats2bfloat162_rn(u.e[0], u.e[1]), hi = __floats2bfloat162_rn(u.e[2], u.e[3]);
                uint2 sh;
                sh.x = *reinterpret_cast<uint32_t*>(&lo);
                sh.y = *reinterpret_cast<uint32_t*>(&hi);
                st_shadow(a.out_bf16[rank], i, sh, valid);
              }
            }
          }
        }
        if (has_send) st_pack(sslot, j, val);
      }
      __syncthreads();
      if (threadIdx.x == 0) {
        __threadfence_system();
        if (has_send) st_relaxed_sys(next_data_flag, seq_send + 1);
        if (has_recv) st_relaxed_sys(prev_ack_flag, seq_recv + 1);
      }
      if (has_send) ++seq_send;
      if (has_recv) ++seq_recv;
    }
  }
  if (threadIdx.x == 0) a.state[rank][kStateRingSeq + blockIdx.x] = seq_send;
}

// ------------------------------------------------------------------ helpers: p2p copy / local reduce
// Vectorised push copy into (possibly peer-mapped) memory: the device-side half of
// BeginSend/BeginReceive/StreamSend (gpu_device_server.go:64-181).
__global__ void __launch_bounds__(kCollThreads)
p2p_copy_kernel(uint8_t* __restrict__ dst, const uint8_t* __restrict__ src, size_t bytes) {
  const size_t stride = static_cast<size_t>(gridDim.x) * blockDim.x;
  const size_t tid = blockIdx.x * static_cast<size_t>(blockDim.x) + threadIdx.x;
  if (((reinterpret_cast<uintptr_t>(dst) | reinterpret_cast<uintptr_t>(src)) & 15) == 0) {
    const size_t nv = bytes / 16;
    const uint4* s4 = reinterpret_cast<const uint4*>(src);
    uint4* d4 = reinterpret_cast<uint4*>(dst);
    size_t i = tid;
    for (; i + 3 * stride < nv; i += 4 * stride) {  // 4 independent 16 B loads in flight per thread
      const uint4 v0 = s4[i], v1 = s4[i + stride], v2 = s4[i + 2 * stride], v3 = s4[i + 3 * stride];
      d4[i] = v0; d4[i + stride] = v1; d4[i + 2 * stride] = v2; d4[i + 3 * stride] = v3;
    }
    for (; i < nv; i += stride) d4[i] = s4[i];
    for (size_t b = nv * 16 + tid; b < bytes; b += stride) dst[b] = src[b];
  } else {
    for (size_t b = tid; b < bytes; b += stride) dst[b] = src[b];
  }
}

// dst = reduce(srcs[0..n)) elementwise on one device (root step of the naive all-reduce,
// gpu_coordinator_server.go:681-686).
struct ReduceArgs {
  const void* src[kMaxRanks];
  void* dst;
  unsigned long long count;
  int n, op;
};
template <typename T>
__global__ void __launch_bounds__(kCollThreads) reduce_n_kernel(const __grid_constant__ ReduceArgs a) {
  const size_t stride = static_cast<size_t>(gridDim.x) * blockDim.x;
  for (size_t i = blockIdx.x * static_cast<size_t>(blockDim.x) + threadIdx.x; i < a.count; i += stride) {
    T acc = reinterpret_cast<const T*>(a.src[0])[i];
    for (int p = 1; p < a.n; ++p) acc = red1<T>(acc, reinterpret_cast<const T*>(a.src[p])[i], a.op);
    reinterpret_cast<T*>(a.dst)[i] = acc;
  }
}

// ------------------------------------------------------------------ communicator (host)
struct Comm {
  int world = 0;
  int nlocal = 0;
  int local_rank[kMaxRanks];
  int local_dev[kMaxRanks];
  // segment[li][r]: base of rank r's peer-visible segment as addressable from local rank li.
  // Segment layout: [signal pad 64 KiB][staging][arena].
  uint8_t* segment[kMaxRanks][kMaxRanks];
  bool imported[kMaxRanks][kMaxRanks];
  size_t staging_bytes = 0, arena_bytes = 0, segment_bytes = 0;
  uint32_t* state[kMaxRanks];  // per local rank, local device memory
  uint32_t* host_flags = nullptr;  // host-mapped: [0] abort requested (host-side), [4..8) status written by kernels
  cudaStream_t abort_stream[kMaxRanks] = {};
  unsigned long long timeout_ns = 10ull * 1000 * 1000 * 1000;
  uint32_t dead_mask = 0;
  bool emulated = false;  // all ranks on one device
  std::mutex mu;
};

static int local_index_of(const Comm* c, int rank) {
  for (int i = 0; i < c->nlocal; ++i)
    if (c->local_rank[i] == rank) return i;
  return -1;
}

template <typename K>
static int launch_coll(K kern, const CollArgs& args, int blocks, int world_y, cudaStream_t st) {
  dim3 grid(blocks, world_y);
  if (world_y > 1) {
    void* params[] = {const_cast<CollArgs*>(&args)};
    DMLP_CUDA_CHECK(cudaLaunchCooperativeKernel(reinterpret_cast<void*>(kern), grid, dim3(kCollThreads), params, 0, st));
  } else {
    // Programmatic stream serialization: the kernel may be scheduled while its predecessor drains; it
    // executes griddepcontrol.wait before touching any data.
    cudaLaunchConfig_t cfg = {};
    cfg.gridDim = grid;
    cfg.blockDim = dim3(kCollThreads);
    cfg.stream = st;
    cudaLaunchAttribute attr[1];
    attr[0].id = cudaLaunchAttributeProgrammaticStreamSerialization;
    attr[0].val.programmaticStreamSerializationAllowed = 1;
    cfg.attrs = attr;
    cfg.numAttrs = 1;
    DMLP_CUDA_CHECK(cudaLaunchKernelEx(&cfg, kern, args));
    DMLP_CUDA_CHECK(cudaGetLastError());
  }
  return DMLP_OK;
}

template <typename T>
static int dispatch_algo(int algo, bool sgd, const CollArgs& args, int blocks, int world_y, cudaStream_t st) {
  if constexpr (std::is_same<T, float>::value) {
    if (sgd) {
      switch (algo) {
        case DMLP_ALGO_ONESHOT: return launch_coll(allreduce_oneshot_kernel<float, true>, args, blocks, world_y, st);
        case DMLP_ALGO_TWOSHOT: return launch_coll(allreduce_twoshot_kernel<float, true>, args, blocks, world_y, st);
        default: return launch_coll(allreduce_ring_kernel<float, true>, args, blocks, world_y, st);
      }
    }
  }
  switch (algo) {
    case DMLP_ALGO_ONESHOT: return launch_coll(allreduce_oneshot_kernel<T, false>, args, blocks, world_y, st);
    case DMLP_ALGO_TWOSHOT: return launch_coll(allreduce_twoshot_kernel<T, false>, args, blocks, world_y, st);
    default: return launch_coll(allreduce_ring_kernel<T, false>, args, blocks, world_y, st);
  }
}

}  // namespace dmlp

using namespace dmlp;

extern "C" int dmlp_comm_create(int world, int nlocal, const int* local_ranks, const int* local_devs,
                                size_t staging_bytes, size_t arena_bytes, dmlp_comm_t* out) {
  DMLP_REQUIRE(out != nullptr, "dmlp_comm_create: null out pointer");
  DMLP_REQUIRE(world >= 1 && world <= kMaxRanks, "dmlp_comm_create: world %d not in [1,%d]", world, kMaxRanks);
  DMLP_REQUIRE(nlocal >= 1 && nlocal <= world, "dmlp_comm_create: nlocal %d not in [1,%d]", nlocal, world);
  Comm* c = new (std::nothrow) Comm();
  if (!c) return set_error(DMLP_ERR_CUDA, "out of host memory");
  c->world = world;
  c->nlocal = nlocal;
  staging_bytes = (staging_bytes + 4095) & ~static_cast<size_t>(4095);
  if (staging_bytes < 64 * 1024) staging_bytes = 64 * 1024;
  arena_bytes = (arena_bytes + 4095) & ~static_cast<size_t>(4095);
  c->staging_bytes = staging_bytes;
  c->arena_bytes = arena_bytes;
  c->segment_bytes = kPadBytes + staging_bytes + arena_bytes;
  memset(c->segment, 0, sizeof(c->segment));
  memset(c->imported, 0, sizeof(c->imported));
  memset(c->state, 0, sizeof(c->state));
  bool same = true;
  for (int i = 0; i < nlocal; ++i) {
    c->local_rank[i] = local_ranks[i];
    c->local_dev[i] = local_devs[i];
    if (local_devs[i] != local_devs[0]) same = false;
    if (local_ranks[i] < 0 || local_ranks[i] >= world) {
      delete c;
      return set_error(DMLP_ERR_INVALID, "dmlp_comm_create: local rank %d out of range", local_ranks[i]);
    }
  }
  c->emulated = same && nlocal == world && world > 1;
  int rc = DMLP_OK;
  void* hf = nullptr;
  if (cudaHostAlloc(&hf, 64, cudaHostAllocPortable | cudaHostAllocMapped) != cudaSuccess) {
    delete c;
    return set_error(DMLP_ERR_CUDA, "cudaHostAlloc for communicator flags failed: %s", cudaGetErrorString(cudaGetLastError()));
  }
  memset(hf, 0, 64);
  c->host_flags = reinterpret_cast<uint32_t*>(hf);
  for (int i = 0; i < nlocal && rc == DMLP_OK; ++i) {
    void* seg = nullptr;
    rc = dmlp_malloc(local_devs[i], c->segment_bytes, &seg);
    if (rc) break;
    void* stt = nullptr;
    rc = dmlp_malloc(local_devs[i], kStateWords * 4, &stt);
    if (rc) break;
    c->state[i] = reinterpret_cast<uint32_t*>(stt);
    // In one process a device pointer is valid on every device (UVA + peer access).
    for (int j = 0; j < nlocal; ++j) c->segment[j][local_ranks[i]] = reinterpret_cast<uint8_t*>(seg);
  }
  if (rc == DMLP_OK && nlocal > 1 && !same) {
    // distinct devices in one process: peer access must be on
    int devs[kMaxRanks];
    int nd = 0;
    for (int i = 0; i < nlocal; ++i) {
      bool seen = false;
      for (int k = 0; k < nd; ++k) seen |= devs[k] == local_devs[i];
      if (!seen) devs[nd++] = local_devs[i];
    }
    if (nd > 1) rc = dmlp_enable_peer_access(nd, devs);
  }
  if (rc) {
    dmlp_comm_destroy(reinterpret_cast<dmlp_comm_t>(c));
    return rc;
  }
  *out = reinterpret_cast<dmlp_comm_t>(c);
  return DMLP_OK;
}

extern "C" int dmlp_comm_destroy(dmlp_comm_t h) {
  Comm* c = reinterpret_cast<Comm*>(h);
  if (!c) return DMLP_OK;
  for (int i = 0; i < c->nlocal; ++i) {
    cudaSetDevice(c->local_dev[i]);
    cudaDeviceSynchronize();
  }
  for (int li = 0; li < c->nlocal; ++li)
    for (int r = 0; r < c->world; ++r)
      if (c->imported[li][r] && c->segment[li][r]) dmlp_ipc_close_handle(c->local_dev[li], c->segment[li][r]);
  for (int i = 0; i < c->nlocal; ++i) {
    uint8_t* seg = c->segment[i][c->local_rank[i]];
    if (seg) dmlp_free(c->local_dev[i], seg);
    if (c->state[i]) dmlp_free(c->local_dev[i], c->state[i]);
  }
  for (int i = 0; i < c->nlocal; ++i)
    if (c->abort_stream[i]) { cudaSetDevice(c->local_dev[i]); cudaStreamDestroy(c->abort_stream[i]); }
  if (c->host_flags) cudaFreeHost(c->host_flags);
  cudaGetLastError();
  delete c;
  return DMLP_OK;
}

extern "C" int dmlp_comm_export(dmlp_comm_t h, int local_idx, unsigned char out[64]) {
  Comm* c = reinterpret_cast<Comm*>(h);
  DMLP_REQUIRE(c && local_idx >= 0 && local_idx < c->nlocal, "dmlp_comm_export: bad arguments");
  return dmlp_ipc_get_handle(c->local_dev[local_idx], c->segment[local_idx][c->local_rank[local_idx]], out);
}

extern "C" int dmlp_comm_import(dmlp_comm_t h, int rank, const unsigned char in[64]) {
  Comm* c = reinterpret_cast<Comm*>(h);
  DMLP_REQUIRE(c && rank >= 0 && rank < c->world, "dmlp_comm_import: bad arguments");
  if (local_index_of(c, rank) >= 0) return DMLP_OK;  // our own segment
  for (int li = 0; li < c->nlocal; ++li) {
    void* p = nullptr;
    int rc = dmlp_ipc_open_handle(c->local_dev[li], in, &p);
    if (rc) return rc;
    c->segment[li][rank] = reinterpret_cast<uint8_t*>(p);
    c->imported[li][rank] = true;
  }
  return DMLP_OK;
}

extern "C" int dmlp_comm_arena_ptr(dmlp_comm_t h, int local_idx, int rank, void** ptr, size_t* bytes) {
  Comm* c = reinterpret_cast<Comm*>(h);
  DMLP_REQUIRE(c && local_idx >= 0 && local_idx < c->nlocal && rank >= 0 && rank < c->world && ptr,
               "dmlp_comm_arena_ptr: bad arguments");
  uint8_t* seg = c->segment[local_idx][rank];
  if (!seg) return set_error(DMLP_ERR_NOT_FOUND, "rank %d's segment has not been imported", rank);
  *ptr = seg + kPadBytes + c->staging_bytes;
  if (bytes) *bytes = c->arena_bytes;
  return DMLP_OK;
}

extern "C" int dmlp_comm_set_timeout_ms(dmlp_comm_t h, uint64_t ms) {
  Comm* c = reinterpret_cast<Comm*>(h);
  DMLP_REQUIRE(c, "null communicator");
  c->timeout_ns = ms * 1000000ull;
  return DMLP_OK;
}

extern "C" int dmlp_comm_abort(dmlp_comm_t h) {
  Comm* c = reinterpret_cast<Comm*>(h);
  DMLP_REQUIRE(c, "null communicator");
  reinterpret_cast<volatile uint32_t*>(c->host_flags)[0] = 1u;  // sticky host-side state (and the copy source)
  // The kernels poll an abort word in device memory (their own signal pad): push it there with a
  // copy on a side stream, which runs while the spinning kernels occupy the compute streams.
  for (int li = 0; li < c->nlocal; ++li) {
    uint8_t* seg = c->segment[li][c->local_rank[li]];
    if (!seg) continue;
    DeviceGuard g(c->local_dev[li]);
    if (c->abort_stream[li] == nullptr) DMLP_CUDA_CHECK(cudaStreamCreateWithFlags(&c->abort_stream[li], cudaStreamNonBlocking));
    DMLP_CUDA_CHECK(cudaMemcpyAsync(reinterpret_cast<uint32_t*>(seg) + kPadAbort, c->host_flags, 4,
                                    cudaMemcpyHostToDevice, c->abort_stream[li]));
  }
  return DMLP_OK;
}

// out[0] = failure code (0 healthy), out[1] = reporting rank, out[2] = peer waited for, out[3] = phase,
// out[4] = abort flag.
extern "C" int dmlp_comm_status(dmlp_comm_t h, uint32_t out[5]) {
  Comm* c = reinterpret_cast<Comm*>(h);
  DMLP_REQUIRE(c && out, "dmlp_comm_status: bad arguments");
  volatile uint32_t* f = c->host_flags;
  out[0] = f[4]; out[1] = f[5]; out[2] = f[6]; out[3] = f[7]; out[4] = f[0];
  return DMLP_OK;
}

// Fault injection: ranks in the mask stop taking part in collectives, as if the GPU died.
extern "C" int dmlp_comm_inject_dead_ranks(dmlp_comm_t h, uint32_t mask) {
  Comm* c = reinterpret_cast<Comm*>(h);
  DMLP_REQUIRE(c, "null communicator");
  c->dead_mask = mask;
  return DMLP_OK;
}

// Used by nvls.cu: the NVLS kernel shares this communicator's signal pads, epochs and watchdog words.
extern "C" int dmlp_comm_internal_view(dmlp_comm_t h, int local_idx, uint32_t* pad[DMLP_MAX_RANKS], uint32_t** state,
                                       uint32_t** host_flags, unsigned long long* timeout_ns, uint32_t* dead_mask,
                                       int* rank, int* world, int* dev) {
  Comm* c = reinterpret_cast<Comm*>(h);
  DMLP_REQUIRE(c && local_idx >= 0 && local_idx < c->nlocal, "dmlp_comm_internal_view: bad arguments");
  for (int r = 0; r < c->world; ++r) {
    if (!c->segment[local_idx][r]) return set_error(DMLP_ERR_NOT_FOUND, "rank %d's segment is not mapped", r);
    pad[r] = reinterpret_cast<uint32_t*>(c->segment[local_idx][r]);
  }
  *state = c->state[local_idx];
  *host_flags = c->host_flags;
  *timeout_ns = c->timeout_ns;
  *dead_mask = c->dead_mask;
  *rank = c->local_rank[local_idx];
  *world = c->world;
  *dev = c->local_dev[local_idx];
  return DMLP_OK;
}

// 1 when all ranks of the communicator sit on one device (collectives then run as ONE cooperative launch; anything
// that would make separate per-rank launches wait on each other must not be used).
extern "C" int dmlp_comm_is_emulated(dmlp_comm_t h, int* emulated) {
  Comm* c = reinterpret_cast<Comm*>(h);
  DMLP_REQUIRE(c && emulated, "dmlp_comm_is_emulated: bad arguments");
  *emulated = c->emulated ? 1 : 0;
  return DMLP_OK;
}

// Algorithm choice, table-driven from the graph-timed sweeps under profiles/ (fp32 SUM, one 8 x B200 NVSwitch box).
// One row per world size: largest message (bytes) for which one-shot still wins over two-shot, and the smallest for which
// the in-switch reduction (NVLS: multimem.ld_reduce + multimem.st) wins - it needs both buffers in multicast-bound
// memory (runtime/nvls.py), which the caller states with `nvls_available`.
//   * one-shot: one rendezvous, (n-1) x bytes pulled per rank - latency-bound sizes;
//   * two-shot: reduce-scatter by peer loads + all-gather by peer stores, 2(n-1)/n x bytes per rank;
//   * NVLS    : every byte crosses each link once and the switch adds; extra latency of the multimem round trip;
//   * ring    : never picked. 2(n-1) dependent hops, each a flag handshake over the same all-to-all fabric on which
//               two-shot moves the same bytes in two hops: on NVSwitch a ring has the latency of a ring and the
//               bandwidth of nothing better. It stays as the RPC's namesake and the comparison mode (algo = ring).
struct AlgoRow { int world; size_t oneshot_max; size_t nvls_min; };
static const AlgoRow kAlgoTable[] = {
    {2, 4u << 20, ~static_cast<size_t>(0)},   // two ranks: a pull of the peer's buffer IS the reduce-scatter + gather
    {3, 1u << 20, 64u << 20},
    {4, 1u << 20, 64u << 20},                 // profiles/r1_allreduce_sweep_4gpu_graph.jsonl: NVLS 173 vs 180 us at 64 MiB
    {6, 512u << 10, 64u << 20},
    {8, 512u << 10, 64u << 20},               // profiles/r2_allreduce_sweep_8gpu_graph.jsonl: two-shot 61 vs NVLS 75 us at 16 MiB,
                                              // NVLS 178 vs 201 us at 64 MiB, 588 vs 761 us at 256 MiB
};

// The table in force: the built-in rows, overridden per world size by DMLP_ALLREDUCE_TABLE (read once) and by
// dmlp_allreduce_set_table. Spec: "world:oneshot_max:nvls_min[,world:...]" with byte counts that take a K / M / G
// suffix (binary) and "never" for "do not pick NVLS", e.g. "8:256K:32M,2:1M:never".
static std::mutex g_table_mu;
static std::vector<AlgoRow> g_table;
static bool g_table_ready = false;

static bool parse_bytes(const char* s, const char* end, size_t* out) {
  if (end - s == 5 && strncmp(s, "never", 5) == 0) { *out = ~static_cast<size_t>(0); return true; }
  if (s == end) return false;
  size_t v = 0;
  const char* p = s;
  for (; p < end && *p >= '0' && *p <= '9'; ++p) v = v * 10 + static_cast<size_t>(*p - '0');
  if (p == s) return false;
  if (p < end) {
    const char c = *p++;
    if (c == 'K' || c == 'k') v <<= 10; else if (c == 'M' || c == 'm') v <<= 20; else if (c == 'G' || c == 'g') v <<= 30; else return false;
  }
  if (p != end) return false;
  *out = v;
  return true;
}

// Applies a spec on top of `rows` (kept sorted by world). Returns false on a malformed spec (rows untouched).
static bool apply_table_spec(const char* spec, std::vector<AlgoRow>* rows) {
  std::vector<AlgoRow> out = *rows;
  const char* p = spec;
  while (*p) {
    const char* e = strchr(p, ',');
    if (!e) e = p + strlen(p);
    const char* c1 = static_cast<const char*>(memchr(p, ':', e - p));
    const char* c2 = c1 ? static_cast<const char*>(memchr(c1 + 1, ':', e - c1 - 1)) : nullptr;
    size_t w = 0, one = 0, nv = 0;
    if (!c1 || !c2 || !parse_bytes(p, c1, &w) || !parse_bytes(c1 + 1, c2, &one) || !parse_bytes(c2 + 1, e, &nv)) return false;
    if (w < 2 || w > static_cast<size_t>(kMaxRanks)) return false;
    AlgoRow r{static_cast<int>(w), one, nv};
    bool found = false;
    for (AlgoRow& x : out)
      if (x.world == r.world) { x = r; found = true; }
    if (!found) {
      size_t i = 0;
      while (i < out.size() && out[i].world < r.world) ++i;
      out.insert(out.begin() + i, r);
    }
    p = *e ? e + 1 : e;
  }
  *rows = out;
  return true;
}

static void table_init_locked() {
  if (g_table_ready) return;
  g_table.assign(kAlgoTable, kAlgoTable + sizeof(kAlgoTable) / sizeof(kAlgoTable[0]));
  const char* e = getenv("DMLP_ALLREDUCE_TABLE");
  if (e && *e && !apply_table_spec(e, &g_table))
    fprintf(stderr, "dmlp: ignoring malformed DMLP_ALLREDUCE_TABLE=%s (want world:oneshot_max:nvls_min[,...])\n", e);
  g_table_ready = true;
}

// Overrides rows of the dispatcher table for this process (spec as above); NULL or "" restores the built-in table
// (+ DMLP_ALLREDUCE_TABLE). The thresholds come from sweeps on one box: a different fabric can re-tune without a rebuild.
extern "C" int dmlp_allreduce_set_table(const char* spec) {
  std::lock_guard<std::mutex> lk(g_table_mu);
  if (!spec || !*spec) {
    g_table_ready = false;
    table_init_locked();
    return DMLP_OK;
  }
  table_init_locked();
  if (!apply_table_spec(spec, &g_table))
    return set_error(DMLP_ERR_INVALID, "dmlp_allreduce_set_table: malformed spec '%s' (want world:oneshot_max:nvls_min[,...])", spec);
  return DMLP_OK;
}

extern "C" int dmlp_allreduce_pick_algo_ex(int world, size_t bytes, int nvls_available) {
  if (world <= 1) return DMLP_ALGO_ONESHOT;
  std::lock_guard<std::mutex> lk(g_table_mu);
  table_init_locked();
  const AlgoRow* row = &g_table[0];
  for (const AlgoRow& r : g_table)
    if (r.world <= world) row = &r;  // the largest tabulated world not above this one
  if (nvls_available && bytes >= row->nvls_min) return DMLP_ALGO_NVLS;
  return bytes <= row->oneshot_max ? DMLP_ALGO_ONESHOT : DMLP_ALGO_TWOSHOT;
}

extern "C" int dmlp_allreduce_pick_algo(int world, size_t bytes) { return dmlp_allreduce_pick_algo_ex(world, bytes, 0); }

extern "C" int dmlp_allreduce(dmlp_comm_t h, int local_idx, const dmlp_allreduce_desc_t* d, void* stream) {
  Comm* c = reinterpret_cast<Comm*>(h);
  DMLP_REQUIRE(c && d, "dmlp_allreduce: null argument");
  NvtxRange nvtx_ar("dmlp.allreduce");
  DMLP_REQUIRE(d->dtype == DMLP_F32 || d->dtype == DMLP_BF16 || d->dtype == DMLP_U8 || d->dtype == DMLP_I32,
               "dmlp_allreduce: unsupported dtype %d", d->dtype);
  DMLP_REQUIRE(d->op >= DMLP_SUM && d->op <= DMLP_AVG, "dmlp_allreduce: unsupported op %d", d->op);
  volatile uint32_t* hf = c->host_flags;
  if (hf[4] != 0 || hf[0] != 0)
    return set_error(DMLP_ERR_ABORTED, "communicator is FAILED (code %u: rank %u waited for rank %u in phase %u)",
                     hf[4], hf[5], hf[6], hf[7]);
  const bool sgd = d->fuse_sgd != 0;
  if (sgd) DMLP_REQUIRE(d->dtype == DMLP_F32 && d->op == DMLP_SUM, "fused SGD needs f32 SUM");
  const bool all = local_idx < 0;
  if (all) DMLP_REQUIRE(c->nlocal == c->world, "local_idx -1 needs every rank to be local");
  else DMLP_REQUIRE(local_idx < c->nlocal, "dmlp_allreduce: local_idx %d out of range", local_idx);
  const int world = c->world;
  const size_t esz = d->dtype == DMLP_U8 ? 1 : d->dtype == DMLP_BF16 ? 2 : 4;
  const size_t bytes = d->count * esz;
  if (d->count == 0) return DMLP_OK;

  int algo = d->algo;
  if (algo == DMLP_ALGO_AUTO) algo = dmlp_allreduce_pick_algo(world, bytes);
  if (world == 1) algo = DMLP_ALGO_ONESHOT;  // degenerate: local copy / update, self-barrier
  DMLP_REQUIRE(algo == DMLP_ALGO_ONESHOT || algo == DMLP_ALGO_TWOSHOT || algo == DMLP_ALGO_RING,
               "dmlp_allreduce: algorithm %d is not available in this build", algo);

  CollArgs a;
  memset(&a, 0, sizeof(a));
  a.world = world;
  a.op = d->op == DMLP_AVG ? DMLP_SUM : d->op;
  a.scale = d->scale != 0.f ? d->scale : 1.f;
  if (d->op == DMLP_AVG && !sgd) a.scale *= 1.f / world;
  a.skip_end_barrier = d->skip_end_barrier;
  a.count = d->count;
  a.staging_bytes = c->staging_bytes;
  a.dead_mask = c->dead_mask;
  a.status = c->host_flags + 4;
  a.timeout_ns = c->timeout_ns;
  a.prof = reinterpret_cast<unsigned long long*>(d->prof_dev);
  {
    static int strict = -1;
    if (strict < 0) { const char* e = getenv("DMLP_COLL_STRICT"); strict = e ? atoi(e) : 2; }  // 2: acquire polls (default)
    a.strict = static_cast<uint32_t>(strict);
  }
  for (int r = 0; r < world; ++r) {
    a.in[r] = d->in[r];
    a.out[r] = d->out[r];
    a.out_bf16[r] = d->out_bf16[r];
    a.lr[r] = d->lr[r];
    if (((reinterpret_cast<uintptr_t>(d->in[r]) | reinterpret_cast<uintptr_t>(d->out[r]) |
          reinterpret_cast<uintptr_t>(d->out_bf16[r])) & 15) != 0)
      return set_error(DMLP_ERR_INVALID, "dmlp_allreduce: rank %d buffers must be 16-byte aligned", r);
  }

  // grid: identical on every rank for a given (count, dtype, algo, world)
  const size_t npack = (bytes + 15) / 16;
  bool inplace = false;
  for (int r = 0; r < world; ++r) inplace |= (d->in[r] != nullptr && d->in[r] == d->out[r]);
  const int hard_cap = c->emulated ? 16 : kMaxCollBlocks;
  if (algo == DMLP_ALGO_ONESHOT && inplace && world > 1) {
    DMLP_REQUIRE(!sgd, "fused SGD cannot run in place (weights and gradients must be distinct buffers)");
    // In-place one-shot keeps its results in registers across a rendezvous; larger messages go
    // through two-shot, which is alias-safe by construction (each rank owns a disjoint slice).
    if (npack > static_cast<size_t>(hard_cap) * kCollThreads * kInplacePerThread) algo = DMLP_ALGO_TWOSHOT;
  }
  size_t work = npack;
  if (algo != DMLP_ALGO_ONESHOT) work = (npack + world - 1) / world;
  // one-shot is latency bound (every load is a peer round trip): one pack per thread keeps all of a
  // message's loads in flight at once; the bandwidth-bound algorithms take two packs per thread.
  const int per_thread = algo == DMLP_ALGO_ONESHOT ? 1 : 2;
  int blocks = static_cast<int>((work + kCollThreads * per_thread - 1) / (kCollThreads * per_thread));
  if (blocks < 1) blocks = 1;
  const int cap = algo == DMLP_ALGO_RING ? (hard_cap < 32 ? hard_cap : 32) : hard_cap;
  if (blocks > cap) blocks = cap;
  a.inplace = (algo == DMLP_ALGO_ONESHOT && inplace && world > 1) ? 1 : 0;
  if (a.inplace) {
    const int need = static_cast<int>((npack + kCollThreads * kInplacePerThread - 1) / (kCollThreads * kInplacePerThread));
    if (blocks < need) blocks = need;
  }

  const int first = all ? 0 : local_idx;
  const int last = all ? c->nlocal : local_idx + 1;
  const bool emu = all && c->emulated;
  for (int li = first; li < last; ++li) {
    const int rank = c->local_rank[li];
    for (int r = 0; r < world; ++r) {
      uint8_t* seg = c->segment[li][r];
      if (!seg) return set_error(DMLP_ERR_NOT_FOUND, "rank %d's segment is not mapped (missing dmlp_comm_import)", r);
      a.pad[r] = reinterpret_cast<uint32_t*>(seg);
      a.staging[r] = seg + kPadBytes;
      const int lr = local_index_of(c, r);
      a.state[r] = lr >= 0 ? c->state[lr] : nullptr;
    }
    a.rank = emu ? -1 : rank;
    if (!emu) {
      DMLP_REQUIRE(a.in[rank] && a.out[rank], "dmlp_allreduce: rank %d has no buffers", rank);
      if (sgd) DMLP_REQUIRE(a.lr[rank] != nullptr, "dmlp_allreduce: fused SGD needs lr for rank %d", rank);
    }
    DeviceGuard guard(c->local_dev[li]);
    cudaStream_t st = as_stream(all && d->streams[li] ? d->streams[li] : stream);
    int rc;
    switch (d->dtype) {
      case DMLP_F32: rc = dispatch_algo<float>(algo, sgd, a, blocks, emu ? world : 1, st); break;
      case DMLP_BF16: rc = dispatch_algo<__nv_bfloat16>(algo, false, a, blocks, emu ? world : 1, st); break;
      case DMLP_U8: rc = dispatch_algo<unsigned char>(algo, false, a, blocks, emu ? world : 1, st); break;
      default: rc = dispatch_algo<int>(algo, false, a, blocks, emu ? world : 1, st); break;
    }
    if (rc) return rc;
    if (emu) break;
  }
  return DMLP_OK;
}

extern "C" int dmlp_p2p_copy(int src_dev, void* dst, const void* src, size_t bytes, void* stream) {
  if (bytes == 0) return DMLP_OK;
  DMLP_REQUIRE(dst && src, "dmlp_p2p_copy: null pointer");
  DeviceGuard guard(src_dev);
  size_t blocks = (bytes / 16 + kCollThreads * 4 - 1) / (kCollThreads * 4);
  if (blocks < 1) blocks = 1;
  if (blocks > 148 * 2) blocks = 148 * 2;
  p2p_copy_kernel<<<static_cast<int>(blocks), kCollThreads, 0, as_stream(stream)>>>(
      reinterpret_cast<uint8_t*>(dst), reinterpret_cast<const uint8_t*>(src), bytes);
  DMLP_CUDA_CHECK(cudaGetLastError());
  return DMLP_OK;
}

extern "C" int dmlp_reduce_n(int dev, void* dst, const void* const* srcs, int n, size_t count, int dtype,
                             int op, void* stream) {
  DMLP_REQUIRE(dst && srcs && n >= 1 && n <= kMaxRanks, "dmlp_reduce_n: bad arguments");
  if (count == 0) return DMLP_OK;
  ReduceArgs a;
  memset(&a, 0, sizeof(a));
  for (int i = 0; i < n; ++i) a.src[i] = srcs[i];
  a.dst = dst; a.count = count; a.n = n; a.op = op == DMLP_AVG ? DMLP_SUM : op;
  DeviceGuard guard(dev);
  size_t blocks = (count + kCollThreads * 4 - 1) / (kCollThreads * 4);
  if (blocks < 1) blocks = 1;
  if (blocks > 148 * 4) blocks = 148 * 4;
  cudaStream_t st = as_stream(stream);
  switch (dtype) {
    case DMLP_F32: reduce_n_kernel<float><<<static_cast<int>(blocks), kCollThreads, 0, st>>>(a); break;
    case DMLP_BF16: reduce_n_kernel<__nv_bfloat16><<<static_cast<int>(blocks), kCollThreads, 0, st>>>(a); break;
    case DMLP_U8: reduce_n_kernel<unsigned char><<<static_cast<int>(blocks), kCollThreads, 0, st>>>(a); break;
    case DMLP_I32: reduce_n_kernel<int><<<static_cast<int>(blocks), kCollThreads, 0, st>>>(a); break;
    default: return set_error(DMLP_ERR_INVALID, "dmlp_reduce_n: unsupported dtype %d", dtype);
  }
  DMLP_CUDA_CHECK(cudaGetLastError());
  return DMLP_OK;
}

// ------------------------------------------------------------------ allocation-free all-reduce
// Some users of this library must not allocate device memory when they call it. Two we know of:
// a serving process that gave all of the GPU's memory to its own tensor pool at start-up (a later
// cudaMalloc can simply fail there), and a latency-critical loop that cannot afford cudaMalloc /
// cudaFree because both synchronise the whole device and stall every stream. For them
// dmlp_allreduce_sum_f32_noalloc runs the ring all-reduce kernel with its signal pads, sequence
// state, watchdog words and staging slots carved out of the fixed array below instead of a
// communicator's cudaMalloc'd segment. The array costs every GPU that loads this library
// kNoAllocBytes (about 4.1 MiB) until the process exits, and its size bounds the message.
namespace dmlp {

constexpr size_t kNoAllocFlagBytes = 256;                // [0] abort word, [4..8) watchdog status
constexpr size_t kNoAllocPadBytes = 8 * 1024;            // one signal pad (uses kPadWords * 4)
constexpr size_t kNoAllocStateBytes = 1024;              // one rank's epochs + ring sequence numbers
constexpr size_t kNoAllocSlotBytes = 256 * 1024;         // one ring staging slot; two per rank
constexpr size_t kNoAllocCtrlBytes = kNoAllocFlagBytes + kMaxRanks * (kNoAllocPadBytes + kNoAllocStateBytes);
constexpr size_t kNoAllocBytes = kNoAllocCtrlBytes + kMaxRanks * 2 * kNoAllocSlotBytes;
constexpr size_t kNoAllocMaxFloatsPerRank = kNoAllocSlotBytes / sizeof(float);  // x world = largest count
static_assert(kPadWords * 4 <= kNoAllocPadBytes, "signal pad does not fit its static slice");
static_assert(kStateWords * 4 <= kNoAllocStateBytes, "rank state does not fit its static slice");
static_assert(kNoAllocCtrlBytes % 256 == 0, "staging slices must stay 256-byte aligned");

// Only device code may use this name as an address. On the host it is a stand-in symbol: the
// launcher below asks the runtime for the real device address, once per device.
static __device__ __align__(256) unsigned char g_noalloc_scratch[kNoAllocBytes];

constexpr int kNoAllocMaxDevices = 64;
static std::mutex g_noalloc_mu;
static unsigned char* g_noalloc_dev_ptr[kNoAllocMaxDevices];

// Device address of the scratch array on the CURRENT device (resolved on first use per device).
static int noalloc_scratch_ptr(unsigned char** out) {
  int dev = -1;
  DMLP_CUDA_CHECK(cudaGetDevice(&dev));
  if (dev < 0 || dev >= kNoAllocMaxDevices)
    return set_error(DMLP_ERR_UNSUPPORTED, "device ordinal %d is outside the %d supported by the static scratch table",
                     dev, kNoAllocMaxDevices);
  std::lock_guard<std::mutex> lock(g_noalloc_mu);
  if (g_noalloc_dev_ptr[dev] == nullptr) {
    void* p = nullptr;
    cudaError_t e = cudaGetSymbolAddress(&p, g_noalloc_scratch);
    if (e != cudaSuccess || p == nullptr)
      return set_error(DMLP_ERR_CUDA, "cudaGetSymbolAddress(static all-reduce scratch) failed on device %d: %s", dev,
                       cudaGetErrorString(e));
    g_noalloc_dev_ptr[dev] = static_cast<unsigned char*>(p);
  }
  *out = g_noalloc_dev_ptr[dev];
  return DMLP_OK;
}

}  // namespace dmlp

extern "C" int dmlp_allreduce_sum_f32_noalloc(int world, const float* const* in_dev, float* const* out_dev,
                                              size_t count, void* stream) {
  DMLP_REQUIRE(world >= 1 && world <= kMaxRanks, "dmlp_allreduce_sum_f32_noalloc: world %d not in [1,%d]", world, kMaxRanks);
  DMLP_REQUIRE(in_dev != nullptr && out_dev != nullptr, "dmlp_allreduce_sum_f32_noalloc: null pointer table");
  const size_t max_count = static_cast<size_t>(world) * kNoAllocMaxFloatsPerRank;
  if (count > max_count)
    return set_error(DMLP_ERR_INVALID,
                     "dmlp_allreduce_sum_f32_noalloc: count %zu exceeds the static scratch capacity of %zu floats "
                     "(%zu per rank x %d ranks); use dmlp_allreduce with a communicator for larger messages",
                     count, max_count, kNoAllocMaxFloatsPerRank, world);
  for (int r = 0; r < world; ++r) {
    DMLP_REQUIRE(in_dev[r] != nullptr && out_dev[r] != nullptr, "dmlp_allreduce_sum_f32_noalloc: rank %d has a null buffer", r);
    DMLP_REQUIRE(((reinterpret_cast<uintptr_t>(in_dev[r]) | reinterpret_cast<uintptr_t>(out_dev[r])) & 15) == 0,
                 "dmlp_allreduce_sum_f32_noalloc: rank %d buffers must be 16-byte aligned", r);
  }
  if (count == 0) return DMLP_OK;
  unsigned char* base = nullptr;
  int rc = noalloc_scratch_ptr(&base);
  if (rc) return rc;
  {
    // the kernel runs on the current device and dereferences every rank's buffers there
    int dev = -1;
    DMLP_CUDA_CHECK(cudaGetDevice(&dev));
    for (int r = 0; r < world; ++r) {
      cudaPointerAttributes at;
      if (cudaPointerGetAttributes(&at, in_dev[r]) == cudaSuccess && at.type == cudaMemoryTypeDevice && at.device != dev)
        return set_error(DMLP_ERR_INVALID, "dmlp_allreduce_sum_f32_noalloc: rank %d's input lives on device %d, the call runs on device %d",
                         r, at.device, dev);
      cudaGetLastError();
    }
  }
  cudaStream_t st = as_stream(stream);
  // The array keeps whatever the previous call left behind. Flags, epochs, sequence numbers and
  // the watchdog words must start from zero; staging slots are always written before they are read.
  DMLP_CUDA_CHECK(cudaMemsetAsync(base, 0, kNoAllocCtrlBytes, st));

  CollArgs a;
  memset(&a, 0, sizeof(a));
  a.world = world;
  a.rank = world > 1 ? -1 : 0;  // every rank lives on this device: one launch, rank = blockIdx.y
  a.op = DMLP_SUM;
  a.scale = 1.f;
  a.count = count;
  a.staging_bytes = 2 * kNoAllocSlotBytes;
  a.status = reinterpret_cast<uint32_t*>(base) + 4;
  a.timeout_ns = 10ull * 1000 * 1000 * 1000;
  unsigned char* pads = base + kNoAllocFlagBytes;
  unsigned char* states = pads + kMaxRanks * kNoAllocPadBytes;
  unsigned char* staging = base + kNoAllocCtrlBytes;
  for (int r = 0; r < world; ++r) {
    a.in[r] = in_dev[r];
    a.out[r] = out_dev[r];
    a.pad[r] = reinterpret_cast<uint32_t*>(pads + r * kNoAllocPadBytes);
    a.state[r] = reinterpret_cast<uint32_t*>(states + r * kNoAllocStateBytes);
    a.staging[r] = staging + static_cast<size_t>(r) * 2 * kNoAllocSlotBytes;
  }
  const size_t npack = (count * sizeof(float) + 15) / 16;
  const size_t work = (npack + world - 1) / world;
  int blocks = static_cast<int>((work + kCollThreads * 2 - 1) / (kCollThreads * 2));
  if (blocks < 1) blocks = 1;
  if (blocks > 16) blocks = 16;
  if (world == 1) return launch_coll(allreduce_oneshot_kernel<float, false>, a, blocks, 1, st);
  return launch_coll(allreduce_ring_kernel<float, false>, a, blocks, world, st);
}

// The allocation-free all-reduce keeps its watchdog record {code, rank, peer, phase} in the static array, where the
// host cannot see it: without this getter a kernel that gave up (10 s without a peer, or an abort) would still look
// like a success. Call it after synchronising the stream the all-reduce ran on; out[0] == 0 means the last call on the
// current device completed. Performs one small synchronous device-to-host copy, no allocation.
extern "C" int dmlp_allreduce_noalloc_status(uint32_t out[4]) {
  DMLP_REQUIRE(out != nullptr, "dmlp_allreduce_noalloc_status: null output");
  unsigned char* base = nullptr;
  int rc = noalloc_scratch_ptr(&base);
  if (rc) return rc;
  DMLP_CUDA_CHECK(cudaMemcpy(out, reinterpret_cast<uint32_t*>(base) + 4, 4 * sizeof(uint32_t), cudaMemcpyDeviceToHost));
  return DMLP_OK;
}
