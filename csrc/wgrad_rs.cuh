// Consumer side of the GEMM's reduce-scatter epilogue: reduce + average + SGD + weight all-gather, tile by tile (sm_100a).
// Runs on four extra warps of every CTA of the wgrad GEMM (gemm_tcgen05.cu, E2M == 2 instantiations).
//
// Data-parallel weight gradient of a large layer as ONE kernel over peer memory instead of "GEMM, then all-reduce":
//
//   epilogue warps of every rank's GEMM                    consumer warps of every rank's GEMM
//   ------------------------------------                    -----------------------------------------------------------
//   tile t done -> its 4 KiB staging chunks go by bulk      for every tile this rank OWNS (t % world == rank), in the
//   copy to slot [my rank][t / world] of the owner          order the GEMMs produce them: wait until all `world` sources
//   (t % world), then the owner's arrival counter of        have delivered it, sum the slots in rank order, w -= lr/world
//   (source, tile) is bumped                                * sum, write the new fp32 weights (+ bf16 shadow) to EVERY rank
//
// so gradient traffic, the update and the weight broadcast run while the tensor cores are still producing later tiles.
// The slots are the layer's own region of the (peer-mapped, parity double-buffered) gradient buffer: a tile-ordered layout
// needs exactly M x N floats. The MMA / TMA / epilogue warps never wait for the consumer warps, and the consumer warps only
// wait for what epilogue warps (here and on the peers) produce: no cyclic wait. A separate consumer kernel launched beside
// a CTA-pair GEMM was measured first: its CTAs were only scheduled after the GEMM's had left (profiles/r2_wgrad_rs.md).
//
// Why a weight write is safe as soon as a tile has arrived from every rank: a rank launches its wgrad GEMM after the
// dgrad GEMM that reads W_l (stream order in trainer.cu), so a counter bumped by rank r proves r no longer reads W_l.
//
// The reference reduces every gradient through host RPCs after the whole backward pass (DSML/client/client.go:619-647,
// gpu_coordinator_server.go:272-376); the bucketed all-reduce in collectives.cu is the unfused form of this code.
#pragma once
#include "common.h"
#include "ptx_sm100.cuh"
#include "wgrad_rs.h"

#include <cuda_bf16.h>

namespace dmlp {

constexpr int kRsWarps = 4;
constexpr int kRsThreads = kRsWarps * 32;
// Debug timeline (DMLP_RS_PROF=1): per CTA {start, first tile arrived, last chunk written, end} of the latest launch
__device__ unsigned long long g_rs_prof[148 * 4];
#define RSPROF(slot)                                                                                         \
  do {                                                                                                       \
    if (a.prof && tid == 0 && blockIdx.x < 148) g_rs_prof[blockIdx.x * 4 + (slot)] = globaltimer_ns();       \
  } while (0)

__device__ __forceinline__ float4 ld_cg_f4(const float* p) {
  float4 v;
  asm volatile("ld.global.cg.v4.f32 {%0, %1, %2, %3}, [%4];" : "=f"(v.x), "=f"(v.y), "=f"(v.z), "=f"(v.w) : "l"(p));
  return v;
}
__device__ __forceinline__ void rs_bar() { asm volatile("bar.sync 3, %0;" ::"n"(kRsThreads) : "memory"); }

// Waits until *f has reached `target` (wrapping compare); false on watchdog expiry / abort (recorded in a.status).
__device__ __noinline__ bool rs_wait(const RsReduceParams& a, const uint32_t* f, uint32_t target, int peer, uint32_t phase) {
  const unsigned long long deadline = globaltimer_ns() + a.timeout_ns;
  uint32_t spins = 0;
  while (static_cast<int32_t>(ld_acquire_sys(f) - target) < 0) {
    if ((++spins & 0x3F) == 0) {
      const uint32_t ab = ld_relaxed_sys(a.pad[a.rank] + kPadAbortWord);
      if (ab != 0u) { record_abort(a.status, a.rank, ab, phase); return false; }
      if (globaltimer_ns() > deadline) {
        if (a.status != nullptr) {
          a.status[1] = static_cast<uint32_t>(a.rank);
          a.status[2] = static_cast<uint32_t>(peer);
          a.status[3] = phase;
          __threadfence_system();
          a.status[0] = DMLP_ERR_TIMEOUT;
        }
        for (int q = 0; q < a.world; ++q) st_relaxed_sys(a.pad[q] + kPadAbortWord, abort_word_for(a.rank, peer));
        return false;
      }
    }
  }
  return true;
}

// Body of the consumer warps (tid = 0 .. kRsThreads-1 among them). s_fail / s_epoch: two words of shared memory.
__device__ __forceinline__ void rs_consume(const RsReduceParams& a, const int tid, int* s_fail, uint32_t* s_epoch) {
  if (tid == 0) { *s_fail = 0; *s_epoch = __ldcg(a.epoch); }
  rs_bar();
  RSPROF(0);
  const uint32_t epoch = *s_epoch;                     // steps this pipeline has completed
  const uint32_t target = (epoch + 1u) * a.per_tile;   // counter value of a fully delivered tile
  const int cph = a.bn / 64;                            // 128-byte chunks per column half
  const int chunks_per_sub = 2 * cph * 4;
  const int chunks_per_tile = a.subs * chunks_per_sub;
  const long long total_items = static_cast<long long>(a.tiles_per_owner) * chunks_per_tile / 2;  // items of 2 chunks
  const size_t tile_elems = static_cast<size_t>(a.tm) * a.bn;
  const size_t src_stride = static_cast<size_t>(a.tiles_per_owner) * tile_elems;
  const float step = a.lr[0] * a.scale;
  int known = -1;                                       // owned tiles [0, known] have arrived from every source
  bool ok = true;
  for (long long it = blockIdx.x; it < total_items; it += gridDim.x) {
    const int j = static_cast<int>((2 * it) / chunks_per_tile);
    const int rem = static_cast<int>(2 * it - static_cast<long long>(j) * chunks_per_tile);
    if (j > known) {
      if (tid < a.world && !rs_wait(a, a.counts + static_cast<size_t>(tid) * a.tiles_per_owner + j, target, tid, 13u))
        *s_fail = 1;
      rs_bar();
      if (*s_fail != 0) { ok = false; break; }
      if (known < 0) RSPROF(1);
      known = j;
    }
    // place of the tile in the weight matrix (same tile order as the GEMM: bands of raster_g M-tiles)
    const int tile = j * a.world + a.rank;
    const int per_band = a.raster_g * a.tiles_n;
    const int band = tile / per_band;
    const int first_m = band * a.raster_g;
    const int rows = min(a.tiles_m - first_m, a.raster_g);
    const int rr = tile - band * per_band;
    const int tm_i = first_m + rr % rows, tn_i = rr / rows;
    const float* tsrc = a.slots + static_cast<size_t>(j) * tile_elems;
    // four 16-byte units per thread: units tid and tid + 128 of chunks rem and rem + 1
    size_t woff[4];
    const float* src[4];
    float4 w[4], s[4];
#pragma unroll
    for (int i = 0; i < 4; ++i) {
      const int c = rem + (i >> 1);
      const int sub = c / chunks_per_sub, cid = c - sub * chunks_per_sub;
      const int u = tid + (i & 1) * kRsThreads, r = u >> 3, k = (u & 7) ^ (r & 7);  // row r, logical 16-byte piece k
      const int half = cid / (cph * 4), ci = (cid >> 2) % cph, q = cid & 3;
      const long long row = static_cast<long long>(tm_i) * a.tm + sub * 128 + q * 32 + r;
      const int col = tn_i * a.bn + half * (a.bn / 2) + ci * 32 + 4 * k;
      woff[i] = static_cast<size_t>(row) * a.ld + col;
      src[i] = tsrc + static_cast<size_t>(sub) * 128 * a.bn + static_cast<size_t>(cid) * 1024 + u * 4;
      s[i] = make_float4(0.f, 0.f, 0.f, 0.f);
    }
#pragma unroll
    for (int i = 0; i < 4; ++i) w[i] = *reinterpret_cast<const float4*>(a.w[a.rank] + woff[i]);
#pragma unroll 1
    for (int g2 = 0; g2 < a.world; g2 += 2) {  // sources in rank order, two at a time (8 loads in flight per thread)
      float4 v[2][4];
#pragma unroll
      for (int i = 0; i < 4; ++i) {
        v[0][i] = ld_cg_f4(src[i] + g2 * src_stride);
        if (g2 + 1 < a.world) v[1][i] = ld_cg_f4(src[i] + (g2 + 1) * src_stride);
      }
#pragma unroll
      for (int i = 0; i < 4; ++i) {
        s[i].x += v[0][i].x; s[i].y += v[0][i].y; s[i].z += v[0][i].z; s[i].w += v[0][i].w;
        if (g2 + 1 < a.world) { s[i].x += v[1][i].x; s[i].y += v[1][i].y; s[i].z += v[1][i].z; s[i].w += v[1][i].w; }
      }
    }
#pragma unroll
    for (int i = 0; i < 4; ++i) {
      const float4 wn = make_float4(w[i].x - step * s[i].x, w[i].y - step * s[i].y, w[i].z - step * s[i].z, w[i].w - step * s[i].w);
      const __nv_bfloat162 lo = __floats2bfloat162_rn(wn.x, wn.y), hi = __floats2bfloat162_rn(wn.z, wn.w);
      const uint2 pk = make_uint2(*reinterpret_cast<const uint32_t*>(&lo), *reinterpret_cast<const uint32_t*>(&hi));
#pragma unroll 1
      for (int q2 = 0; q2 < a.world; ++q2) {
        *reinterpret_cast<float4*>(a.w[q2] + woff[i]) = wn;
        if (a.w16[q2] != nullptr) *reinterpret_cast<uint2*>(a.w16[q2] + woff[i]) = pk;
      }
    }
  }
  // every weight these warps wrote is performed at the peers before their done count becomes visible there
  rs_bar();
  RSPROF(2);
  if (ok && tid == 0) {
    __threadfence_system();
    for (int q2 = 0; q2 < a.world; ++q2)
      asm volatile("red.relaxed.sys.global.add.u32 [%0], 1;" ::"l"(a.done[q2] + a.rank) : "memory");
  }
  RSPROF(3);
  if (blockIdx.x != 0 || !ok) return;
  // CTA 0 holds the launch open until every rank's consumers have finished: when this kernel completes, the layer's
  // weights on this GPU are the new ones everywhere (the next forward pass follows in stream order)
  if (tid < a.world && !rs_wait(a, a.done[a.rank] + tid, (epoch + 1u) * gridDim.x, tid, 14u)) *s_fail = 1;
  rs_bar();
  if (tid == 0 && *s_fail == 0) *a.epoch = epoch + 1u;
  RSPROF(3);
}

}  // namespace dmlp
