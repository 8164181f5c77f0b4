// Parameters of the reduce + SGD + all-gather consumer of the GEMM's reduce-scatter epilogue (wgrad_rs.cuh).
#pragma once
#include <cuda_bf16.h>
#include <cuda_runtime.h>
#include <stdint.h>

namespace dmlp {

// Words of the signal pad used by the pipeline (one block per fused layer, see trainer.cu):
//   [0..7] done counters by source rank, [8] steps completed (local), [16 ..) arrival counters [source][tile / world]
constexpr int kPadRsBase = 9216;
constexpr int kPadRsWords = 16384 - kPadRsBase;
constexpr int kRsDone = 0, kRsEpoch = 8, kRsCounts = 16;

struct RsReduceParams {
  int world, rank;
  long long ld;               // row pitch of the weight matrix (elements)
  int tm, bn, subs;           // GEMM tile: rows, columns, CTAs per tile (128 rows each)
  int tiles_m, tiles_n, raster_g, tiles_per_owner;
  const float* slots;         // my slot area: [source][tile / world][tm * bn], chunks as the GEMM staged them
  const uint32_t* counts;     // my arrival counters [source][tile / world]
  uint32_t per_tile;          // increments one source adds per tile and step (epilogue warps x CTAs per tile)
  float* w[8];                // the layer's fp32 weights on rank r, as mapped here (own rank included)
  __nv_bfloat16* w16[8];      // bf16 shadow (nullable)
  const float* lr;
  float scale;                // 1 / world
  uint32_t* epoch;            // local word: steps completed
  uint32_t* done[8];          // rank r's done counters [source]
  uint32_t* pad[8];           // rank r's signal pad (abort word)
  uint32_t* status;           // host-mapped {code, rank, peer, phase} (nullable)
  unsigned long long timeout_ns;
  int prof;                   // debug timeline (DMLP_RS_PROF=1)
};

}  // namespace dmlp
