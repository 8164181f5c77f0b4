// A GEMM launch prepared once (TMA descriptors encoded, tile/stage configuration chosen) and
// replayed every step: the native trainer builds its plans at set-up so the per-step path is
// kernel launches only (and is then captured into a CUDA graph).
#pragma once
#include <cuda.h>
#include <cuda_bf16.h>
#include <cuda_runtime.h>
#include <stdint.h>

#include "../include/dmlp_b200.h"

#include "wgrad_rs.h"

namespace dmlp {

// Epilogue transpose patch: 32 rows x 32 fp32 columns per epilogue warp, rows padded to 36 floats so that
// both the row-wise float4 writes and the column-group float4 reads are bank-conflict free.
constexpr int kEpiLd = 36;
constexpr uint32_t kEpiStageBytes = 4 * 32 * kEpiLd * 4;

struct GemmParams {
  int M, N, K;
  int block_n;
  int num_k_blocks;
  int stages;
  uint32_t tmem_cols;      // allocated TMEM columns (two accumulator buffers)
  uint32_t tmem_buf_cols;  // columns of one buffer
  uint32_t a_stage_bytes, b_stage_bytes;
  uint32_t a_tx_bytes;  // bytes TMA actually delivers for A per stage (half a tile when M <= 64)
  uint32_t idesc;
  int raster_g;            // tile order: height (in tiles) of the M bands, see the kernel
  uint32_t epi_stage_off;  // byte offset (from the aligned smem base) of the epilogue's transpose patches
  float* out_f32;
  long long ld_out;
  __nv_bfloat16* out_bf16;
  long long ld_out_bf16;
  const float* bias;
  const void* mask;
  long long ld_mask;
  int mask_is_bf16;
  float* colsum;
  int relu;
  int accumulate;
  int prof;  // debug: CTA 0 stamps a per-tile timeline into g_gemm_prof
  int pdl;  // launched with programmatic stream serialization: the kernel executes griddepcontrol.wait
  int epi2;  // TMA-store epilogue with eight epilogue warps (E2 kernel instantiations); see gemm_plan_create
  // fused softmax cross-entropy epilogue (last layer): non-null labels switch it on
  const int* ce_labels;
  int ce_classes;
  float ce_scale;
  float* ce_logits;
  float* ce_loss_sum;
  int* ce_correct;
  // reduce-scatter epilogue (data-parallel weight gradients, E2 plans only): instead of storing the tile into out_f32,
  // every 4 KiB staging chunk is pushed by the bulk copy engine into the slot [my rank][tile / world] of the tile's
  // OWNER (rank tile % world; peer-mapped memory, own rank included) and the owner's arrival counter of that
  // (source, tile) is bumped once per epilogue warp. Four more warps of the same CTAs (wgrad_rs.cuh) consume the tiles
  // this rank owns: rsq.
  int rs_world;  // 0: off
  int rs_rank;
  int rs_tiles_per_owner;
  float* rs_slot[8];
  uint32_t* rs_count[8];
  RsReduceParams rsq;
};

// Geometry a consumer of the reduce-scatter epilogue needs to find a chunk's place in the output matrix.
struct GemmRsGeometry {
  int tm, bn, subs;  // tile rows (128 per CTA of a pair), tile columns, CTAs per tile
  int tiles_m, tiles_n, raster_g, total_tiles;
  int grid;          // CTAs of the GEMM launch
};

struct GemmPlan {
  CUtensorMap tmA, tmB;
  CUtensorMap tmC;  // output tensor (box: 32 rows x 128 bytes, 128B swizzle) of the TMA-store epilogue, when p.epi2
  GemmParams p;
  int grid_x, grid_y;
  int total_tiles;
  size_t smem;
  int key;  // kernel instantiation: (bf16 << 2) | (a_mn << 1) | b_mn
};

int gemm_plan_create(const dmlp_gemm_desc_t* desc, GemmPlan* plan);
int gemm_plan_run(const GemmPlan* plan, cudaStream_t stream);
// True when the plan's epilogue can push whole tiles to their owners (TMA-store epilogue, fp32 output, no partial tiles,
// tiles divisible among the ranks).
bool gemm_plan_rs_eligible(const GemmPlan* plan, int world);
// Switches the plan to the reduce-scatter epilogue. slots[r] / counts[r]: rank r's slot area and counters as mapped on
// this device. Shrinks the shared-memory ring so that a small consumer CTA fits on the same SM.
int gemm_plan_enable_rs(GemmPlan* plan, int world, int rank, float* const* slots, uint32_t* const* counts,
                        GemmRsGeometry* geo);

}  // namespace dmlp
