// Input all-gather for the global-batch mode of the fused training step (mlp_fused.cu, sm_100a).
//
// In that mode every rank computes the layer-1 weight gradient of the GLOBAL batch, so it needs every rank's
// input batches. They are known long before they are used: after a rank's host->device copy of a chunk of
// batches has landed, this kernel (on the copy stream, beside the compute kernel) writes the chunk into the
// same ring slots of every peer's gathered input buffer over NVLink (plain 16-byte stores to peer-mapped
// memory: posted writes, no round trip) and then raises one "ready" word per CTA in each peer's signal pad.
// Flow control the other way: a ring slot is overwritten only after every peer's step kernel has reported
// (its "done" word in MY pad) that it finished the steps which last read that slot.
//
// The reference moves every batch's gradient through host RPCs (DSML/client/client.go:619-647); inputs
// never leave the client there. This exchange has no counterpart in the reference: it is what lets the step
// kernel trade the 401 KB dW1 all-reduce for a 32 KB dZ1 all-gather.
#include "common.h"
#include "ptx_sm100.cuh"

namespace dmlp {

struct XPushParams {
  const float4* src;   // this rank's chunk (already in its own slots of its own gathered buffer)
  float4* dst[8];      // the same slots in rank r's gathered buffer, as mapped here (own rank: unused)
  size_t n16;          // 16-byte packs
  uint32_t* pad[8];    // rank r's signal pad
  int rank, world;
  uint32_t need_done;  // every peer's done word in my pad must have reached this step count
  uint32_t ready;      // step count to publish in the peers' ready words
  unsigned long long timeout_ns;
  uint32_t* status;    // host-mapped {code, rank, peer, phase} (nullable)
};

__global__ void __launch_bounds__(256) input_exchange_kernel(const XPushParams a) {
  __shared__ int s_fail;
  if (threadIdx.x == 0) s_fail = 0;
  __syncthreads();
  if (threadIdx.x < a.world && static_cast<int>(threadIdx.x) != a.rank) {
    const uint32_t* f = a.pad[a.rank] + kPadXDone + threadIdx.x;
    const unsigned long long deadline = globaltimer_ns() + a.timeout_ns;
    uint32_t spins = 0;
    while (static_cast<int32_t>(ld_relaxed_sys(f) - a.need_done) < 0) {
      if ((++spins & 0x3F) == 0) {
        const uint32_t ab = ld_relaxed_sys(a.pad[a.rank] + kPadAbortWord);
        if (ab != 0u) { record_abort(a.status, a.rank, ab, 12u); s_fail = 1; break; }
        if (globaltimer_ns() > deadline) {
          if (a.status != nullptr) {
            a.status[1] = static_cast<uint32_t>(a.rank);
            a.status[2] = threadIdx.x;
            a.status[3] = 12u;  // phase: input exchange, waiting for a peer to release ring slots
            __threadfence_system();
            a.status[0] = DMLP_ERR_TIMEOUT;
          }
          for (int q = 0; q < a.world; ++q) st_relaxed_sys(a.pad[q] + kPadAbortWord, abort_word_for(a.rank, static_cast<int>(threadIdx.x)));
          s_fail = 1;
          break;
        }
      }
    }
  }
  __syncthreads();
  if (s_fail != 0) return;  // (no ready word: the peers' watchdogs report this rank)
  const size_t stride = static_cast<size_t>(gridDim.x) * blockDim.x;
  for (size_t i0 = static_cast<size_t>(blockIdx.x) * blockDim.x + threadIdx.x; i0 < a.n16; i0 += 4 * stride) {
    float4 v[4];
#pragma unroll
    for (int k = 0; k < 4; ++k)
      if (i0 + k * stride < a.n16) v[k] = __ldcs(a.src + i0 + k * stride);
#pragma unroll
    for (int q = 0; q < 8; ++q) {
      if (q >= a.world || q == a.rank) continue;
#pragma unroll
      for (int k = 0; k < 4; ++k)
        if (i0 + k * stride < a.n16) a.dst[q][i0 + k * stride] = v[k];
    }
  }
  __syncthreads();
  if (threadIdx.x == 0) {
    __threadfence_system();  // this CTA's stores (ordered before it by the barrier) are performed at the peers ...
    for (int q = 0; q < a.world; ++q)  // ... before its ready word becomes visible there
      if (q != a.rank) st_relaxed_sys(a.pad[q] + kPadXReady + a.rank * kXPushCtas + blockIdx.x, a.ready);
  }
}

}  // namespace dmlp

using namespace dmlp;

extern "C" int dmlp_input_exchange(const void* src_local, void* const* dst_peers, size_t bytes, uint32_t* const* pads,
                                   int rank, int world, uint32_t need_done, uint32_t ready_value,
                                   unsigned long long timeout_ns, uint32_t* status, void* stream) {
  DMLP_REQUIRE(src_local && dst_peers && pads && world >= 2 && world <= 8 && rank >= 0 && rank < world && bytes % 16 == 0,
               "dmlp_input_exchange: bad arguments");
  XPushParams a;
  a.src = static_cast<const float4*>(src_local);
  a.n16 = bytes / 16;
  for (int r = 0; r < 8; ++r) {
    a.dst[r] = r < world ? static_cast<float4*>(dst_peers[r]) : nullptr;
    a.pad[r] = r < world ? pads[r] : nullptr;
    DMLP_REQUIRE(r >= world || (a.pad[r] && (r == rank || a.dst[r])), "dmlp_input_exchange: rank %d is not mapped", r);
  }
  a.rank = rank;
  a.world = world;
  a.need_done = need_done;
  a.ready = ready_value;
  a.timeout_ns = timeout_ns ? timeout_ns : 10ull * 1000 * 1000 * 1000;
  a.status = status;
  input_exchange_kernel<<<kXPushCtas, 256, 0, as_stream(stream)>>>(a);
  DMLP_CUDA_CHECK(cudaGetLastError());
  return DMLP_OK;
}
