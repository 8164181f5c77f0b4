// tcgen05 / TMEM / TMA GEMM with fused MLP epilogues (sm_100a only).
//
// One CTA computes a 128 x block_n output tile:
//   warp 0      : TMA producer  (cp.async.bulk.tensor, 128B swizzle, mbarrier complete_tx)
//   warp 1      : TMEM allocator + single-thread tcgen05.mma issuer (accumulator in TMEM)
//   warps 2..5  : epilogue (tcgen05.ld -> bias / ReLU / ReLU-mask / column-sum -> global)
//   warps 2..9  : (E2 instantiations: plain / bias / ReLU outputs) eight epilogue warps, two per TMEM lane quadrant,
//                 each staging 32 rows x 128 B of finished output in shared memory and handing it to the TMA store
//                 engine (cp.async.bulk.tensor shared -> global): no transposing patch, no per-lane global stores
// The smem ring has `stages` slots, each holding one 128-byte-wide K slab of A and B
// (32 tf32 or 64 bf16 elements of K). block_n is a *runtime* value: the instruction
// descriptor, the TMEM column count and the smem carve-up are all computed at launch.
//
// Provides (reference DSML/client/client.go): forward GEMM+bias+ReLU (:112-141),
// dgrad + ReLU mask (:178-189), wgrad (:167-176, :191-199), bias-gradient column sums
// (:173-175, :197-199).
#include "common.h"
#include "gemm_plan.h"
#include "wgrad_rs.cuh"
#include "ptx_sm100.cuh"

#include <stdlib.h>
#include <string.h>

#include <atomic>

namespace dmlp {


constexpr int kMaxStages = 8;
constexpr int kGemmThreads = 192;    // producer + MMA + 4 epilogue warps
constexpr int kGemmThreadsRS = 448;  // ... + 4 consumer warps (reduce-scatter epilogue, wgrad_rs.cuh)
constexpr int kGemmThreadsE2 = 320;  // producer + MMA + 8 epilogue warps (TMA-store epilogue)
constexpr uint32_t kE2StageBytes = 8 * 4096;  // one 32-row x 128 B staging buffer per epilogue warp

// Per-tile timeline of CTA 0 (debug, DMLP_GEMM_PROF=1 at plan creation): 8 globaltimer stamps per tile.
//  0 producer: first smem slot of the tile acquired   1 producer: last TMA of the tile issued
//  2 MMA: accumulator buffer acquired                 3 MMA: first operands landed      4 MMA: last commit issued
//  5 epilogue: accumulator ready                      6 epilogue: accumulator drained   7 epilogue: tile done
constexpr int kProfTiles = 64;
__device__ unsigned long long g_gemm_prof[kProfTiles * 16];
#define GPROF(slot, tile_i)                                                                          \
  do {                                                                                               \
    if (p.prof && blockIdx.x == 0 && (tile_i) < kProfTiles && (threadIdx.x & 31) == 0)               \
      g_gemm_prof[(tile_i) * 16 + (slot)] = globaltimer_ns();                                         \
  } while (0)

// TWO: CTA pairs (cluster of 2) run tcgen05.mma.cta_group::2 on 256 x block_n tiles: each CTA stages its 128 rows of A
// and HALF of the B tile, which halves the shared-memory fill + read traffic per SM that bounds the 1-CTA mainloop.
// E2M: 0 = four epilogue warps, generic epilogue; 1 = eight epilogue warps, TMA-store epilogue; 2 = eight epilogue warps,
// reduce-scatter epilogue (fp32 tiles pushed to their owner ranks; a leaner instantiation: see gemm_plan_enable_rs).
template <int EB, bool A_MN, bool B_MN, bool TWO, int E2M>
__global__ void __launch_bounds__(E2M == 2 ? kGemmThreadsRS : E2M ? kGemmThreadsE2 : kGemmThreads, 1)
gemm_tcgen05_kernel(const __grid_constant__ CUtensorMap tmA, const __grid_constant__ CUtensorMap tmB,
                    const __grid_constant__ CUtensorMap tmC, const __grid_constant__ GemmParams p) {
  constexpr bool E2 = E2M != 0;
  constexpr bool RS = E2M == 2;
  constexpr int BM = 128;
  constexpr int KS = 128 / EB;            // K elements per pipeline stage (one 128 B swizzle row)
  constexpr int UK = 32 / EB;             // K elements per tcgen05.mma
  constexpr int ATOM_MN = 128 / EB;       // MN elements in one MN-major swizzle atom
  constexpr uint32_t MN_ATOM_COL_BYTES = KS * 128;  // one MN-atom column of a stage (KS k-rows)
  // MN-major swizzle flavour: 16-bit types use SWIZZLE_128B (8 K-rows per atom, 1024 B); 32-bit
  // types must use SWIZZLE_128B_BASE32B (32 B chunks, 4 K-rows per atom, 512 B).
  constexpr uint32_t MN_SBO = (EB == 4) ? 512 : 1024;
  constexpr uint32_t MN_LAYOUT = (EB == 4) ? 1 : 2;

  extern __shared__ uint8_t smem_raw[];
  // 1024-byte alignment by pointer arithmetic on the shared array itself (a round trip through an integer makes
  // the compiler fall back to generic LD/ST for everything addressed off this base)
  uint8_t* smem = smem_raw + ((1024u - (smem_u32(smem_raw) & 1023u)) & 1023u);
  uint8_t* smem_a = smem;
  uint8_t* smem_b = smem + static_cast<size_t>(p.stages) * p.a_stage_bytes;

  __shared__ __align__(8) uint64_t full_bar[kMaxStages];
  __shared__ __align__(8) uint64_t empty_bar[kMaxStages];
  __shared__ __align__(8) uint64_t tmem_full_bar[2];   // accumulator buffer b holds a finished tile
  __shared__ __align__(8) uint64_t tmem_empty_bar[2];  // the epilogue has drained buffer b
  __shared__ uint32_t tmem_base_slot;
  __shared__ float s_part_raw[E2 ? 1 : 4][256];  // per-epilogue-warp column partials (E2 plans have no column sums:
  auto s_part = [&](int w) -> float* { return s_part_raw[E2 ? 0 : w]; };  // their code is never reached there)
  __shared__ float s_loss[4];
  __shared__ int s_rs_fail;
  __shared__ uint32_t s_rs_epoch;
  __shared__ int s_hit[4];

  // Warp index through a shuffle: the compiler then knows the role branches below are warp-uniform and
  // keeps descriptors, barrier addresses and loop counters in uniform registers. With a thread-dependent
  // branch (`lane == 0`) every tcgen05.mma / TMA issue is wrapped in a per-lane "waterfall" loop of
  // R2UR moves and costs ~170 cycles instead of ~40 (scripts/probe_mma_issue.cu, profiles/r1_mma_issue_probe.md).
  const int warp = __shfl_sync(0xffffffffu, threadIdx.x >> 5, 0);
  const int lane = threadIdx.x & 31;
  // Persistent CTA: tiles blockIdx.x, blockIdx.x + gridDim.x, ... (M fastest, so CTAs that run together share
  // the same B tile through L2). Two TMEM accumulator buffers let the epilogue of one tile overlap the
  // mainloop of the next.
  constexpr int TM = TWO ? 2 * BM : BM;  // rows of a (pair's) tile
  const int tiles_m = (p.M + TM - 1) / TM;
  const int total_tiles = tiles_m * ((p.N + p.block_n - 1) / p.block_n);
  const int crank = TWO ? static_cast<int>(cl_rank()) : 0;  // 0: leader (issues the MMAs)
  const int tile_first = TWO ? static_cast<int>(blockIdx.x >> 1) : static_cast<int>(blockIdx.x);
  const int tile_step = TWO ? static_cast<int>(gridDim.x >> 1) : static_cast<int>(gridDim.x);
  const int b_cols = TWO ? p.block_n / 2 : p.block_n;       // N columns of the B tile this CTA stages
  // Tile order: bands of `raster_g` M-tiles, N fastest inside a band, M fastest inside a column of the band. The CTAs
  // that run together then cover a near-square block of tiles (shared A rows AND shared B columns stay in L2)
  // instead of one full column of M tiles: 3x less DRAM traffic at 8192^3, which is power, i.e. clock.
  const int tiles_n = (p.N + p.block_n - 1) / p.block_n;
  auto tile_mn = [&](int tile, int& tm_i, int& tn_i) {
    const int per_band = p.raster_g * tiles_n;
    const int band = tile / per_band;
    const int first_m = band * p.raster_g;
    const int rows = min(tiles_m - first_m, p.raster_g);
    const int r = tile - band * per_band;
    tm_i = first_m + r % rows;
    tn_i = r / rows;
  };

  if (warp == 0 && lane == 0) {
    tma_prefetch_desc(&tmA);
    tma_prefetch_desc(&tmB);
    if constexpr (E2) tma_prefetch_desc(&tmC);
    for (int s = 0; s < p.stages; ++s) {
      mbar_init(&full_bar[s], 1);
      mbar_init(&empty_bar[s], 1);
    }
    for (int b = 0; b < 2; ++b) {
      mbar_init(&tmem_full_bar[b], 1);
      mbar_init(&tmem_empty_bar[b], (TWO ? 2 : 1) * (E2 ? 256 : 128));  // pair: the leader's barrier collects both CTAs' epilogues
    }
    fence_mbar_init();
  }
  if (warp == 1) {
    if constexpr (TWO) {
      tmem_alloc_2sm(&tmem_base_slot, p.tmem_cols);
      tmem_relinquish_2sm();
    } else {
      tmem_alloc(&tmem_base_slot, p.tmem_cols);
      tmem_relinquish();
    }
  }
  tc_fence_before_sync();
  __syncthreads();
  if constexpr (TWO) cl_sync();  // the peer's barriers are initialised before anything signals them
  tc_fence_after_sync();
  const uint32_t tmem_base = __shfl_sync(0xffffffffu, tmem_base_slot, 0);
  if (p.pdl) {
    // Programmatic dependent launch: everything above (barrier init, TMEM allocation, descriptor
    // prefetch) overlapped the tail of the previous kernel in the stream. From here on this kernel
    // reads what its predecessors wrote, so wait for them to complete and flush; then let OUR
    // dependent start its own prologue right away (single-wave grids only, see launch_gemm).
    asm volatile("griddepcontrol.wait;" ::: "memory");
    asm volatile("griddepcontrol.launch_dependents;" ::: "memory");
  }

  if (warp == 0) {
    // ------------------------------------------------------------ TMA producer (whole warp loops, one lane issues)
    {
      const uint32_t stage_bytes = p.a_tx_bytes + p.b_stage_bytes;  // bytes TMA will deliver per stage
      const int n_atoms = (b_cols + ATOM_MN - 1) / ATOM_MN;
      // k-block counter across all tiles of this CTA: the smem ring never drains between tiles
      int s = 0;
      uint32_t ph = 0;
      int ti = 0;
      for (int tile = tile_first; tile < total_tiles; tile += tile_step, ++ti) {
        int tm_i, tn_i;
        tile_mn(tile, tm_i, tn_i);
        const int m0 = tm_i * TM + crank * BM;                 // my 128 rows of the (pair's) tile
        const int n0 = tn_i * p.block_n + crank * b_cols;      // my share of the tile's B columns
        for (int kb = 0; kb < p.num_k_blocks; ++kb) {
          mbar_wait(&empty_bar[s], ph ^ 1);
          if (kb == 0) GPROF(0, ti);
          if (elect_one()) {
            // pair: both CTAs' loads complete bytes on the LEADER's barrier, which only the leader arms
            if (crank == 0) mbar_arrive_expect_tx(&full_bar[s], TWO ? 2 * stage_bytes : stage_bytes);
            const uint32_t fb = TWO ? cl_mapa(smem_u32(&full_bar[s]), 0) : 0u;
            auto load = [&](void* dst, const CUtensorMap* tm, int c0, int c1) {
              if constexpr (TWO) tma_load_2d_2sm(dst, tm, fb, c0, c1);
              else tma_load_2d(dst, tm, &full_bar[s], c0, c1);
            };
            uint8_t* a_dst = smem_a + static_cast<size_t>(s) * p.a_stage_bytes;
            uint8_t* b_dst = smem_b + static_cast<size_t>(s) * p.b_stage_bytes;
            const int k0 = kb * KS;
            if constexpr (!A_MN) {
              load(a_dst, &tmA, k0, m0);
            } else {
#pragma unroll
              for (int j = 0; j < BM / ATOM_MN; ++j)
                load(a_dst + j * MN_ATOM_COL_BYTES, &tmA, m0 + j * ATOM_MN, k0);
            }
            if constexpr (!B_MN) {
              load(b_dst, &tmB, k0, n0);
            } else {
              for (int j = 0; j < n_atoms; ++j)
                load(b_dst + j * MN_ATOM_COL_BYTES, &tmB, n0 + j * ATOM_MN, k0);
            }
          }
          __syncwarp();
          if (kb == p.num_k_blocks - 1) GPROF(1, ti);
          if (++s == p.stages) { s = 0; ph ^= 1; }
        }
      }
    }
  } else if (warp == 1) {
    // ------------------------------------------------------------ MMA issuer (whole warp loops, one lane issues)
    if (crank == 0) {  // pair: the leader issues for both CTAs
      const uint32_t a_base = smem_u32(smem_a), b_base = smem_u32(smem_b);
      // descriptor bits that never change; the 14-bit start address (>> 4) is added per stage / k step
      const uint64_t adesc0 = A_MN ? make_smem_desc(0, MN_ATOM_COL_BYTES, MN_SBO, MN_LAYOUT) : make_smem_desc(0, 16, 1024);
      const uint64_t bdesc0 = B_MN ? make_smem_desc(0, MN_ATOM_COL_BYTES, MN_SBO, MN_LAYOUT) : make_smem_desc(0, 16, 1024);
      constexpr uint32_t A_KSTEP = (A_MN ? UK * 128 : 32) >> 4, B_KSTEP = (B_MN ? UK * 128 : 32) >> 4;
      int s = 0;
      uint32_t ph = 0, tcount = 0;
      for (int tile = tile_first; tile < total_tiles; tile += tile_step, ++tcount) {
        const uint32_t buf = tcount & 1;
        mbar_wait(&tmem_empty_bar[buf], ((tcount >> 1) & 1) ^ 1);  // epilogue drained this buffer (free at first use)
        tc_fence_after_sync();
        GPROF(2, tcount);
        const uint32_t tmem_acc = tmem_base + buf * p.tmem_buf_cols;
        for (int kb = 0; kb < p.num_k_blocks; ++kb) {
          mbar_wait(&full_bar[s], ph);
          tc_fence_after_sync();
          if (kb == 0) GPROF(3, tcount);
          // K-major: step 32 B inside the 128 B swizzle row. MN-major: step UK k-rows of 128 B.
          const uint64_t adesc = adesc0 + ((a_base + s * p.a_stage_bytes) >> 4);
          const uint64_t bdesc = bdesc0 + ((b_base + s * p.b_stage_bytes) >> 4);
          if (elect_one()) {
#pragma unroll
            for (int k = 0; k < KS / UK; ++k) {
              if constexpr (TWO) tc_mma_2sm<EB>(tmem_acc, adesc + k * A_KSTEP, bdesc + k * B_KSTEP, p.idesc, (kb | k) != 0 ? 1u : 0u);
              else tc_mma<EB>(tmem_acc, adesc + k * A_KSTEP, bdesc + k * B_KSTEP, p.idesc, (kb | k) != 0 ? 1u : 0u);
            }
            // frees the smem slot (in both CTAs of a pair) once these MMAs have read it
            if constexpr (TWO) tc_commit_2sm(&empty_bar[s], 3); else tc_commit(&empty_bar[s]);
            if (kb == p.num_k_blocks - 1) {
              if constexpr (TWO) tc_commit_2sm(&tmem_full_bar[buf], 3); else tc_commit(&tmem_full_bar[buf]);
            }
          }
          __syncwarp();
          if (kb == p.num_k_blocks - 1) GPROF(4, tcount);
          if (++s == p.stages) { s = 0; ph ^= 1; }
        }
      }
    }
  } else if (RS && warp >= 10) {
    // ------------------------------------------------------------ consumer warps of the reduce-scatter epilogue
    if constexpr (RS) rs_consume(p.rsq, static_cast<int>(threadIdx.x) - kGemmThreadsE2, &s_rs_fail, &s_rs_epoch);
  } else {
    // ------------------------------------------------------------ epilogue warps
    const int q = warp & 3;  // TMEM lane quadrant this warp may access
    const int row = q * 32 + lane;
    const int etid = (warp - 2) * 32 + lane;   // 0..127 among the epilogue threads
    // A single M tile sees every row of the problem: its reductions (bias gradient, loss, hits) are
    // final and can be stored; otherwise they are accumulated atomically into zeroed buffers.
    const bool direct = tiles_m == 1 && !TWO;  // (a pair's two CTAs each hold half the rows of the tile)
    uint32_t tcount = 0;
    for (int tile = tile_first; tile < total_tiles; tile += tile_step, ++tcount) {
    int tm_i, tn_i;
    tile_mn(tile, tm_i, tn_i);
    const int m0 = tm_i * TM + crank * BM;
    const int n0 = tn_i * p.block_n;
    const int m = m0 + row;
    const bool m_ok = m < p.M;
    const uint32_t buf = tcount & 1;
    bool handled = false;
    if constexpr (E2) {
      // ---- TMA-store epilogue (plans without mask / column sums / accumulate / softmax-CE, exactly one output).
      // Interior tiles: warp w drains TMEM lanes 32*(w%4).. of the column half (w-2)/4 in chunks of 128 output bytes
      // (64 bf16 / 32 fp32 columns): tcgen05.ld -> bias / ReLU in registers -> one 128 B row per lane into a private
      // 4 KiB staging buffer (128B-swizzled: conflict-free) -> one cp.async.bulk.tensor store per chunk by one lane.
      const bool fast2 = m0 + BM <= p.M && n0 + p.block_n <= p.N;  // (uniform over the CTA)
      const uint32_t e2buf = smem_u32(smem + p.epi_stage_off) + static_cast<uint32_t>(warp - 2) * 4096u;
      if (elect_one()) tma_store_wait_read<0>();  // my previous tile's stores have left the staging buffer
      __syncwarp();
      if (!fast2) {
        // edge tile: the bounds-checked path below, run by warps 2..5 over patches that overlay all staging buffers
        asm volatile("bar.sync 2, 256;" ::: "memory");
        handled = warp >= 6;
      } else {
        handled = true;
        constexpr int CH = 128 / 4;                       // fp32 accumulator columns per fp32 chunk ...
        const bool ob = !RS && p.out_bf16 != nullptr;     // ... twice that per bf16 chunk
        const int chunk = ob ? 2 * CH : CH;
        const int ncols = p.block_n >> 1;
        const int cbase = ((warp - 2) >> 2) * ncols;
        mbar_wait(&tmem_full_bar[buf], (tcount >> 1) & 1);
        tc_fence_after_sync();
        if (warp == 2) GPROF(5, tcount);
        const uint32_t t_row = tmem_base + buf * p.tmem_buf_cols + (static_cast<uint32_t>(q * 32) << 16);
        const uint32_t wrow = e2buf + lane * 128;
        const uint32_t sw = static_cast<uint32_t>(lane & 7);
        // reduce-scatter epilogue: this CTA's half of slot [my rank][tile / world] at the tile's owner, chunk by chunk
        // (chunk id = (column half * chunks per half + chunk) * 4 + lane quadrant; 4 KiB each, kept 128B-swizzled)
        float* rs_dst = nullptr;
        if constexpr (RS) {
          const int owner = tile % p.rs_world, jt = tile / p.rs_world;
          rs_dst = p.rs_slot[owner] +
                   (static_cast<size_t>(p.rs_rank) * p.rs_tiles_per_owner + jt) * (static_cast<size_t>(TM) * p.block_n) +
                   static_cast<size_t>(crank) * BM * p.block_n +
                   (static_cast<size_t>((warp - 2) >> 2) * (ncols / CH) * 4 + q) * 1024;
        }
        for (int c = 0; c < ncols; c += chunk) {
          const int col = cbase + c;
          float v[64];
#pragma unroll
          for (int g = 0; g < 4; ++g) {
            if (g < chunk / 16) {
              uint32_t r[16];
              tmem_ld_x16(t_row + col + 16 * g, r);
              tmem_ld_wait();
#pragma unroll
              for (int jj = 0; jj < 16; ++jj) v[16 * g + jj] = __uint_as_float(r[jj]);
            }
          }
          if (p.bias != nullptr) {
#pragma unroll
            for (int g = 0; g < 16; ++g) {
              if (4 * g < chunk) {
                const float4 b4 = __ldg(reinterpret_cast<const float4*>(p.bias + n0 + col) + g);
                v[4 * g] += b4.x; v[4 * g + 1] += b4.y; v[4 * g + 2] += b4.z; v[4 * g + 3] += b4.w;
              }
            }
          }
          if (p.relu) {
#pragma unroll
            for (int jj = 0; jj < 64; ++jj) v[jj] = fmaxf(v[jj], 0.f);
          }
          if (c != 0) {
            if (elect_one()) tma_store_wait_read<0>();  // the previous chunk's store has read the buffer
            __syncwarp();
          }
          if (ob) {
#pragma unroll
            for (int k = 0; k < 8; ++k) {  // 8 x 16 B = this lane's 128 B row: 64 bf16
              uint32_t pk[4];
#pragma unroll
              for (int e = 0; e < 4; ++e) {
                const __nv_bfloat162 h2 = __floats2bfloat162_rn(v[8 * k + 2 * e], v[8 * k + 2 * e + 1]);
                pk[e] = *reinterpret_cast<const uint32_t*>(&h2);
              }
              asm volatile("st.shared.v4.b32 [%0], {%1, %2, %3, %4};" ::"r"(wrow + ((static_cast<uint32_t>(k) ^ sw) << 4)), "r"(pk[0]),
                           "r"(pk[1]), "r"(pk[2]), "r"(pk[3]) : "memory");
            }
          } else {
#pragma unroll
            for (int k = 0; k < 8; ++k)  // 32 fp32
              sts128(wrow + ((static_cast<uint32_t>(k) ^ sw) << 4), v[4 * k], v[4 * k + 1], v[4 * k + 2], v[4 * k + 3]);
          }
          fence_proxy_async_smem();
          __syncwarp();
          if (elect_one()) {  // (converged warp: always the same lane, which also owns the bulk group)
            if constexpr (RS) {
              asm volatile("cp.async.bulk.global.shared::cta.bulk_group [%0], [%1], 4096;" ::"l"(rs_dst + static_cast<size_t>(c / CH) * 4096),
                           "r"(e2buf) : "memory");
            } else {
              tma_store_2d(&tmC, smem + p.epi_stage_off + static_cast<size_t>(warp - 2) * 4096u, n0 + col, m0 + q * 32);
            }
            tma_store_commit();
          }
        }
        if constexpr (RS) {
          if (elect_one()) {
            tma_store_wait<0>();  // my chunks of this tile have been written ...
            asm volatile("fence.proxy.async;" ::: "memory");
            __threadfence_system();  // ... and are ordered before the owner sees the counter move
            uint32_t* cnt = p.rs_count[tile % p.rs_world] + static_cast<size_t>(p.rs_rank) * p.rs_tiles_per_owner + tile / p.rs_world;
            asm volatile("red.relaxed.sys.global.add.u32 [%0], 1;" ::"l"(cnt) : "memory");
          }
          __syncwarp();
        }
        if (warp == 2) GPROF(6, tcount);
      }
    }
    if (!handled) {
    float* stg = reinterpret_cast<float*>(smem + p.epi_stage_off) + q * (32 * kEpiLd);
    const int rsub = lane >> 3, c4 = (lane & 7) * 4;
    const bool vec_f32 = p.out_f32 != nullptr && ((reinterpret_cast<uintptr_t>(p.out_f32) & 15) == 0) && (p.ld_out & 3) == 0;
    const bool vec_b16 = p.out_bf16 != nullptr && ((reinterpret_cast<uintptr_t>(p.out_bf16) & 7) == 0) && (p.ld_out_bf16 & 3) == 0;
    const bool vec_mask = p.mask != nullptr && (p.ld_mask & 3) == 0 &&
                          ((reinterpret_cast<uintptr_t>(p.mask) & (p.mask_is_bf16 ? 7 : 15)) == 0);
    // Interior tile with vectorisable pointers: straight-line code, no per-element bounds checks, loads
    // batched ahead of their uses. (One epilogue warp per scheduler: nothing hides latency but ILP.)
    const bool fast = !E2 && m0 + BM <= p.M && n0 + p.block_n <= p.N && (p.block_n & 31) == 0 &&
                    (p.out_f32 == nullptr || vec_f32) && (p.out_bf16 == nullptr || vec_b16) &&
                    (p.mask == nullptr || vec_mask) && ((reinterpret_cast<uintptr_t>(p.bias) & 15) == 0);
    const uint32_t stg_s = smem_u32(stg);
    const uint32_t wr_s = stg_s + lane * (kEpiLd * 4);
    const uint32_t rd_s = stg_s + (rsub * kEpiLd + c4) * 4;
    const long long rbase = m0 + q * 32 + rsub;
    float* of = p.out_f32 != nullptr ? p.out_f32 + rbase * p.ld_out + n0 + c4 : nullptr;
    __nv_bfloat16* oh = p.out_bf16 != nullptr ? p.out_bf16 + rbase * p.ld_out_bf16 + n0 + c4 : nullptr;
    const __nv_bfloat16* mh = (p.mask != nullptr && p.mask_is_bf16)
                                  ? reinterpret_cast<const __nv_bfloat16*>(p.mask) + rbase * p.ld_mask + n0 + c4 : nullptr;
    const float* mf = (p.mask != nullptr && !p.mask_is_bf16)
                          ? reinterpret_cast<const float*>(p.mask) + rbase * p.ld_mask + n0 + c4 : nullptr;
    const long long of_step = 4 * p.ld_out, oh_step = 4 * p.ld_out_bf16, mk_step = 4 * p.ld_mask;
    // first chunk's bias / mask values: requested before waiting for the accumulator
    float4 b_nx = make_float4(0.f, 0.f, 0.f, 0.f);
    uint2 mh_nx[8];
    float4 mf_nx[8];
    if (fast && p.ce_labels == nullptr) {
      if (p.bias != nullptr) b_nx = __ldg(reinterpret_cast<const float4*>(p.bias + n0 + c4));
      if (mh != nullptr) {
#pragma unroll
        for (int i = 0; i < 8; ++i) mh_nx[i] = __ldg(reinterpret_cast<const uint2*>(mh + i * mk_step));
      }
      if (mf != nullptr) {
#pragma unroll
        for (int i = 0; i < 8; ++i) mf_nx[i] = __ldg(reinterpret_cast<const float4*>(mf + i * mk_step));
      }
    }
    mbar_wait(&tmem_full_bar[buf], (tcount >> 1) & 1);
    tc_fence_after_sync();
    if (warp == 2) GPROF(5, tcount);
    const uint32_t t_row = tmem_base + buf * p.tmem_buf_cols + (static_cast<uint32_t>(q * 32) << 16);

    if (p.ce_labels != nullptr) {
      // ---- fused softmax cross-entropy (reference client.go:76-93,144-165,352-382): the whole logits
      // row (block_n == N <= 32) sits in this thread's registers; logits never travel through memory.
      float v[32];
#pragma unroll
      for (int c = 0; c < 32; c += 16) {
        if (c < p.block_n) {
          uint32_t r[16];
          tmem_ld_x16(t_row + c, r);
          tmem_ld_wait();
#pragma unroll
          for (int j = 0; j < 16; ++j) {
            float x = __uint_as_float(r[j]);
            if (p.bias != nullptr && c + j < p.N) x += __ldg(p.bias + c + j);
            v[c + j] = x;
          }
        } else {
#pragma unroll
          for (int j = 0; j < 16; ++j) v[c + j] = 0.f;
        }
      }
      const int C = p.ce_classes;
      if (p.ce_logits != nullptr && m_ok) {
        float* lo = p.ce_logits + static_cast<long long>(m) * p.ld_out;
#pragma unroll
        for (int j = 0; j < 32; ++j)
          if (j < p.N) lo[j] = v[j];
      }
      float mx = -INFINITY;
      int arg = 0;
#pragma unroll
      for (int j = 0; j < 32; ++j)
        if (j < C && v[j] > mx) { mx = v[j]; arg = j; }
      float sum = 0.f;
#pragma unroll
      for (int j = 0; j < 32; ++j) {
        v[j] = (j < C) ? __expf(v[j] - mx) : 0.f;
        sum += v[j];
      }
      const float inv = 1.f / sum;
      const int y = m_ok ? p.ce_labels[m] : -1;
      float loss = 0.f;
      int hit = 0;
#pragma unroll
      for (int j = 0; j < 32; ++j) {
        const float pr = v[j] * inv;
        if (j == y) loss = -__logf(pr + 1e-10f);
        v[j] = (m_ok && j < C) ? (pr - (j == y ? 1.f : 0.f)) * p.ce_scale : 0.f;
      }
      if (m_ok && arg == y) hit = 1;
      if (m_ok) {
        if (p.out_f32 != nullptr) {
          float* o = p.out_f32 + static_cast<long long>(m) * p.ld_out;
#pragma unroll
          for (int j = 0; j < 32; ++j)
            if (j < p.N) o[j] = v[j];
        }
        if (p.out_bf16 != nullptr) {
          __nv_bfloat16* o = p.out_bf16 + static_cast<long long>(m) * p.ld_out_bf16;
#pragma unroll
          for (int j = 0; j < 32; ++j)
            if (j < p.N) o[j] = __float2bfloat16_rn(v[j]);
        }
      }
      // reductions over the rows of this tile
#pragma unroll
      for (int j = 0; j < 32; ++j) {
        float sj = v[j];
        sj += __shfl_xor_sync(0xffffffffu, sj, 16);
        sj += __shfl_xor_sync(0xffffffffu, sj, 8);
        sj += __shfl_xor_sync(0xffffffffu, sj, 4);
        sj += __shfl_xor_sync(0xffffffffu, sj, 2);
        sj += __shfl_xor_sync(0xffffffffu, sj, 1);
        if (lane == j) s_part(q)[j] = sj;
      }
#pragma unroll
      for (int o = 16; o > 0; o >>= 1) {
        loss += __shfl_xor_sync(0xffffffffu, loss, o);
        hit += __shfl_xor_sync(0xffffffffu, hit, o);
      }
      if (lane == 0) { s_loss[q] = loss; s_hit[q] = hit; }
      asm volatile("bar.sync 1, 128;" ::: "memory");
      if (etid < p.N && p.colsum != nullptr) {
        const float t = s_part(0)[etid] + s_part(1)[etid] + s_part(2)[etid] + s_part(3)[etid];
        if (direct) p.colsum[etid] = t; else atomicAdd(p.colsum + etid, t);
      }
      if (etid == 0) {
        const float l = s_loss[0] + s_loss[1] + s_loss[2] + s_loss[3];
        const int h = s_hit[0] + s_hit[1] + s_hit[2] + s_hit[3];
        if (p.ce_loss_sum != nullptr) { if (direct) *p.ce_loss_sum = l; else atomicAdd(p.ce_loss_sum, l); }
        if (p.ce_correct != nullptr) { if (direct) *p.ce_correct = h; else if (h) atomicAdd(p.ce_correct, h); }
      }
    } else {
      // ---- generic epilogue. A thread owns one accumulator ROW in TMEM, but rows are ld_out apart in
      // memory: storing from that layout makes every warp-wide access touch 32 cache lines. So each warp
      // turns its 32 rows x 32 columns around through a private shared-memory patch: rows go in, then
      // lane l picks up columns 4*(l%8)..+3 of rows l/8 + 4*i. In that layout one instruction covers
      // 4 rows x 128 contiguous bytes: mask / accumulate loads and the stores are coalesced, and a column
      // sum is 8 register adds + 2 shuffles instead of 5 shuffles per element.
      int cols_done = 0;
      if (fast) {
        for (int c = 0; c < p.block_n; c += 32) {
          const float4 b = b_nx;
          uint2 mh_cur[8];
          float4 mf_cur[8];
          if (mh != nullptr) {
#pragma unroll
            for (int i = 0; i < 8; ++i) mh_cur[i] = mh_nx[i];
          }
          if (mf != nullptr) {
#pragma unroll
            for (int i = 0; i < 8; ++i) mf_cur[i] = mf_nx[i];
          }
          if (c + 32 < p.block_n) {  // next chunk's values travel while this chunk is processed
            if (p.bias != nullptr) b_nx = __ldg(reinterpret_cast<const float4*>(p.bias + n0 + c + 32 + c4));
            if (mh != nullptr) {
#pragma unroll
              for (int i = 0; i < 8; ++i) mh_nx[i] = __ldg(reinterpret_cast<const uint2*>(mh + c + 32 + i * mk_step));
            }
            if (mf != nullptr) {
#pragma unroll
              for (int i = 0; i < 8; ++i) mf_nx[i] = __ldg(reinterpret_cast<const float4*>(mf + c + 32 + i * mk_step));
            }
          }
          {
            uint32_t r0[16], r1[16];
            tmem_ld_x16(t_row + c, r0);
            tmem_ld_x16(t_row + c + 16, r1);
            tmem_ld_wait();
#pragma unroll
            for (int j = 0; j < 4; ++j) {
              sts128(wr_s + j * 16, __uint_as_float(r0[4 * j]), __uint_as_float(r0[4 * j + 1]), __uint_as_float(r0[4 * j + 2]),
                     __uint_as_float(r0[4 * j + 3]));
              sts128(wr_s + 64 + j * 16, __uint_as_float(r1[4 * j]), __uint_as_float(r1[4 * j + 1]),
                     __uint_as_float(r1[4 * j + 2]), __uint_as_float(r1[4 * j + 3]));
            }
          }
          __syncwarp();
          float4 t[8];
#pragma unroll
          for (int i = 0; i < 8; ++i) t[i] = lds128(rd_s + i * (4 * kEpiLd * 4));
          if (p.bias != nullptr) {
#pragma unroll
            for (int i = 0; i < 8; ++i) { t[i].x += b.x; t[i].y += b.y; t[i].z += b.z; t[i].w += b.w; }
          }
          if (p.relu) {
#pragma unroll
            for (int i = 0; i < 8; ++i) {
              t[i].x = fmaxf(t[i].x, 0.f); t[i].y = fmaxf(t[i].y, 0.f); t[i].z = fmaxf(t[i].z, 0.f); t[i].w = fmaxf(t[i].w, 0.f);
            }
          }
          if (mh != nullptr) {
            const uint2* mr = mh_cur;
#pragma unroll
            for (int i = 0; i < 8; ++i) {
              // bf16 > 0  <=>  sign clear and magnitude non-zero (NaN masks count as positive: never produced here)
              if (!(__uint_as_float(mr[i].x << 16) > 0.f)) t[i].x = 0.f;
              if (!(__uint_as_float(mr[i].x & 0xffff0000u) > 0.f)) t[i].y = 0.f;
              if (!(__uint_as_float(mr[i].y << 16) > 0.f)) t[i].z = 0.f;
              if (!(__uint_as_float(mr[i].y & 0xffff0000u) > 0.f)) t[i].w = 0.f;
            }
          } else if (mf != nullptr) {
            const float4* mr = mf_cur;
#pragma unroll
            for (int i = 0; i < 8; ++i) {
              if (!(mr[i].x > 0.f)) t[i].x = 0.f;
              if (!(mr[i].y > 0.f)) t[i].y = 0.f;
              if (!(mr[i].z > 0.f)) t[i].z = 0.f;
              if (!(mr[i].w > 0.f)) t[i].w = 0.f;
            }
          }
          if (of != nullptr) {
            if (p.accumulate) {
              float4 o[8];
#pragma unroll
              for (int i = 0; i < 8; ++i) o[i] = *reinterpret_cast<const float4*>(of + c + i * of_step);
#pragma unroll
              for (int i = 0; i < 8; ++i) { t[i].x += o[i].x; t[i].y += o[i].y; t[i].z += o[i].z; t[i].w += o[i].w; }
            }
#pragma unroll
            for (int i = 0; i < 8; ++i) *reinterpret_cast<float4*>(of + c + i * of_step) = t[i];
          }
          if (oh != nullptr) {
#pragma unroll
            for (int i = 0; i < 8; ++i) {
              const __nv_bfloat162 h0 = __floats2bfloat162_rn(t[i].x, t[i].y);
              const __nv_bfloat162 h1 = __floats2bfloat162_rn(t[i].z, t[i].w);
              uint2 pk;
              pk.x = *reinterpret_cast<const uint32_t*>(&h0);
              pk.y = *reinterpret_cast<const uint32_t*>(&h1);
              *reinterpret_cast<uint2*>(oh + c + i * oh_step) = pk;
            }
          }
          if (p.colsum != nullptr) {
            float cs[4] = {0.f, 0.f, 0.f, 0.f};
#pragma unroll
            for (int i = 0; i < 8; ++i) { cs[0] += t[i].x; cs[1] += t[i].y; cs[2] += t[i].z; cs[3] += t[i].w; }
#pragma unroll
            for (int k = 0; k < 4; ++k) {
              cs[k] += __shfl_xor_sync(0xffffffffu, cs[k], 8);
              cs[k] += __shfl_xor_sync(0xffffffffu, cs[k], 16);
            }
            if (lane < 8) {
#pragma unroll
              for (int k = 0; k < 4; ++k) s_part(q)[c + c4 + k] = cs[k];
            }
          }
          __syncwarp();  // the patch is rewritten by the next chunk
        }
        cols_done = p.block_n;
      } else
      for (int c = 0; c < p.block_n; c += 32) {
        const int nb = n0 + c;
        if (nb >= p.N) break;  // uniform across the CTA
        {
          uint32_t r0[16], r1[16];
          tmem_ld_x16(t_row + c, r0);
          const bool second = c + 16 < p.block_n;  // block_n is a multiple of 16, not always of 32
          if (second) tmem_ld_x16(t_row + c + 16, r1);
          tmem_ld_wait();
          float4* dst = reinterpret_cast<float4*>(stg + lane * kEpiLd);
#pragma unroll
          for (int j = 0; j < 4; ++j)
            dst[j] = make_float4(__uint_as_float(r0[4 * j]), __uint_as_float(r0[4 * j + 1]), __uint_as_float(r0[4 * j + 2]),
                                 __uint_as_float(r0[4 * j + 3]));
          if (second) {
#pragma unroll
            for (int j = 0; j < 4; ++j)
              dst[4 + j] = make_float4(__uint_as_float(r1[4 * j]), __uint_as_float(r1[4 * j + 1]),
                                       __uint_as_float(r1[4 * j + 2]), __uint_as_float(r1[4 * j + 3]));
          }
        }
        __syncwarp();
        const int cw = min(32, p.block_n - c);  // live columns of this patch
        cols_done = c + cw;
        const int col = nb + c4;
        const int nvalid = (c4 < cw) ? min(4, p.N - col) : 0;  // columns of mine that exist (<= 0: none)
        float bia[4] = {0.f, 0.f, 0.f, 0.f};
        if (p.bias != nullptr) {
#pragma unroll
          for (int k = 0; k < 4; ++k)
            if (k < nvalid) bia[k] = __ldg(p.bias + col + k);
        }
        float cs[4] = {0.f, 0.f, 0.f, 0.f};
#pragma unroll
        for (int i = 0; i < 8; ++i) {
          const int rr = i * 4 + rsub;
          const int mm = m0 + q * 32 + rr;
          const float4 t = *reinterpret_cast<const float4*>(stg + rr * kEpiLd + c4);
          float v[4] = {t.x + bia[0], t.y + bia[1], t.z + bia[2], t.w + bia[3]};
          if (p.relu) {
#pragma unroll
            for (int k = 0; k < 4; ++k) v[k] = fmaxf(v[k], 0.f);
          }
          if (mm < p.M && nvalid > 0) {
            const bool full4 = nvalid == 4;
            if (p.mask != nullptr) {
              if (p.mask_is_bf16) {
                const __nv_bfloat16* mk = reinterpret_cast<const __nv_bfloat16*>(p.mask) + static_cast<long long>(mm) * p.ld_mask + col;
                if (full4 && vec_mask) {
                  const uint2 raw = *reinterpret_cast<const uint2*>(mk);
                  const __nv_bfloat162 a = *reinterpret_cast<const __nv_bfloat162*>(&raw.x);
                  const __nv_bfloat162 b = *reinterpret_cast<const __nv_bfloat162*>(&raw.y);
                  if (!(__low2float(a) > 0.f)) v[0] = 0.f;
                  if (!(__high2float(a) > 0.f)) v[1] = 0.f;
                  if (!(__low2float(b) > 0.f)) v[2] = 0.f;
                  if (!(__high2float(b) > 0.f)) v[3] = 0.f;
                } else {
#pragma unroll
                  for (int k = 0; k < 4; ++k)
                    if (k < nvalid && !(__bfloat162float(mk[k]) > 0.f)) v[k] = 0.f;
                }
              } else {
                const float* mk = reinterpret_cast<const float*>(p.mask) + static_cast<long long>(mm) * p.ld_mask + col;
                if (full4 && vec_mask) {
                  const float4 mv = *reinterpret_cast<const float4*>(mk);
                  if (!(mv.x > 0.f)) v[0] = 0.f;
                  if (!(mv.y > 0.f)) v[1] = 0.f;
                  if (!(mv.z > 0.f)) v[2] = 0.f;
                  if (!(mv.w > 0.f)) v[3] = 0.f;
                } else {
#pragma unroll
                  for (int k = 0; k < 4; ++k)
                    if (k < nvalid && !(mk[k] > 0.f)) v[k] = 0.f;
                }
              }
            }
            if (p.out_f32 != nullptr) {
              float* o = p.out_f32 + static_cast<long long>(mm) * p.ld_out + col;
              if (full4 && vec_f32) {
                if (p.accumulate) {
                  const float4 old = *reinterpret_cast<const float4*>(o);
                  v[0] += old.x; v[1] += old.y; v[2] += old.z; v[3] += old.w;
                }
                *reinterpret_cast<float4*>(o) = make_float4(v[0], v[1], v[2], v[3]);
              } else {
#pragma unroll
                for (int k = 0; k < 4; ++k)
                  if (k < nvalid) {
                    if (p.accumulate) v[k] += o[k];
                    o[k] = v[k];
                  }
              }
            }
            if (p.out_bf16 != nullptr) {
              __nv_bfloat16* o = p.out_bf16 + static_cast<long long>(mm) * p.ld_out_bf16 + col;
              if (full4 && vec_b16) {
                const __nv_bfloat162 h0 = __floats2bfloat162_rn(v[0], v[1]);
                const __nv_bfloat162 h1 = __floats2bfloat162_rn(v[2], v[3]);
                uint2 pk;
                pk.x = *reinterpret_cast<const uint32_t*>(&h0);
                pk.y = *reinterpret_cast<const uint32_t*>(&h1);
                *reinterpret_cast<uint2*>(o) = pk;
              } else {
#pragma unroll
                for (int k = 0; k < 4; ++k)
                  if (k < nvalid) o[k] = __float2bfloat16_rn(v[k]);
              }
            }
#pragma unroll
            for (int k = 0; k < 4; ++k) cs[k] += v[k];  // rows past M and columns past N never get here
          }
        }
        if (p.colsum != nullptr) {
          // Bias gradient: column sums over this warp's 32 rows
#pragma unroll
          for (int k = 0; k < 4; ++k) {
            cs[k] += __shfl_xor_sync(0xffffffffu, cs[k], 8);
            cs[k] += __shfl_xor_sync(0xffffffffu, cs[k], 16);
          }
          if (lane < 8) {
#pragma unroll
            for (int k = 0; k < 4; ++k) s_part(q)[c + c4 + k] = cs[k];
          }
        }
        __syncwarp();  // the patch is rewritten by the next chunk
      }
      if (warp == 2) GPROF(6, tcount);
      if (p.colsum != nullptr) {
        // combine the four warps' partials: one store (or one atomic) per column per CTA
        asm volatile("bar.sync 1, 128;" ::: "memory");
        for (int cidx = etid; cidx < cols_done; cidx += 128) {
          const int n = n0 + cidx;
          if (n < p.N) {
            const float t = s_part(0)[cidx] + s_part(1)[cidx] + s_part(2)[cidx] + s_part(3)[cidx];
            if (direct) p.colsum[n] = t; else atomicAdd(p.colsum + n, t);
          }
        }
      }
    }
    if (p.colsum != nullptr || p.ce_labels != nullptr) asm volatile("bar.sync 1, 128;" ::: "memory");  // s_part reuse
    }  // !handled
    if constexpr (E2) {
      // an edge tile's patches overlay every warp's staging buffer: nobody stages again before they are all done
      if (!(m0 + BM <= p.M && n0 + p.block_n <= p.N)) {
        if (warp >= 6) {  // (sat the tile out: still in step with the accumulator hand-over)
          mbar_wait(&tmem_full_bar[buf], (tcount >> 1) & 1);
        }
        asm volatile("bar.sync 2, 256;" ::: "memory");
      }
    }
    // this tile's accumulator has been read out: hand the TMEM buffer back to the MMA warp
    tc_fence_before_sync();
    if (TWO && crank != 0) mbar_arrive_cluster(cl_mapa(smem_u32(&tmem_empty_bar[buf]), 0));  // the leader's barrier
    else mbar_arrive(&tmem_empty_bar[buf]);
    if (warp == 2) GPROF(7, tcount);
    }
  }

  if constexpr (E2) {
    if (warp >= 2 && elect_one()) tma_store_wait<0>();  // my bulk stores are complete before this CTA's shared memory goes away
  }
  tc_fence_before_sync();
  __syncthreads();
  if constexpr (TWO) cl_sync();  // the peer may still be signalling my barriers / the pair's MMAs use both TMEMs
  if (warp == 1) {
    tc_fence_after_sync();
    if constexpr (TWO) tmem_dealloc_2sm(tmem_base, p.tmem_cols);
    else tmem_dealloc(tmem_base, p.tmem_cols);
  }
}

// ---------------------------------------------------------------------------- host side
typedef CUresult (*EncodeTiledFn)(CUtensorMap*, CUtensorMapDataType, cuuint32_t, void*,
                                  const cuuint64_t*, const cuuint64_t*, const cuuint32_t*,
                                  const cuuint32_t*, CUtensorMapInterleave, CUtensorMapSwizzle,
                                  CUtensorMapL2promotion, CUtensorMapFloatOOBfill);

static EncodeTiledFn get_encode_fn() {
  static EncodeTiledFn fn = nullptr;
  if (fn == nullptr) {
    void* p = nullptr;
    cudaDriverEntryPointQueryResult qres;
    if (cudaGetDriverEntryPoint("cuTensorMapEncodeTiled", &p, cudaEnableDefault, &qres) ==
            cudaSuccess &&
        qres == cudaDriverEntryPointSuccess)
      fn = reinterpret_cast<EncodeTiledFn>(p);
  }
  return fn;
}

// 2-D tensor map over a row-major [rows][ld] matrix (inner dimension `cols` contiguous).
int make_tensor_map_2d(CUtensorMap* out, const void* base, int elem_bytes, bool is_bf16,
                       uint64_t cols, uint64_t rows, uint64_t ld_elems, uint32_t box_cols,
                       uint32_t box_rows, int swizzle /*0 none, 1 128B, 2 128B with 32B atoms*/) {
  EncodeTiledFn fn = get_encode_fn();
  if (fn == nullptr) return set_error(DMLP_ERR_CUDA, "cuTensorMapEncodeTiled is unavailable");
  if ((reinterpret_cast<uintptr_t>(base) & 15) != 0)
    return set_error(DMLP_ERR_INVALID, "TMA operand base %p is not 16-byte aligned", base);
  if ((ld_elems * elem_bytes) % 16 != 0)
    return set_error(DMLP_ERR_INVALID,
                     "TMA operand leading dimension %llu elements is not a multiple of 16 bytes",
                     (unsigned long long)ld_elems);
  cuuint64_t gdim[2] = {cols, rows};
  cuuint64_t gstride[1] = {ld_elems * static_cast<uint64_t>(elem_bytes)};
  cuuint32_t box[2] = {box_cols, box_rows};
  cuuint32_t estr[2] = {1, 1};
  CUtensorMapDataType dt = is_bf16 ? CU_TENSOR_MAP_DATA_TYPE_BFLOAT16 : CU_TENSOR_MAP_DATA_TYPE_FLOAT32;
  CUresult r = fn(out, dt, 2, const_cast<void*>(base), gdim, gstride, box, estr,
                  CU_TENSOR_MAP_INTERLEAVE_NONE,
                  swizzle == 1 ? CU_TENSOR_MAP_SWIZZLE_128B
                  : swizzle == 2 ? CU_TENSOR_MAP_SWIZZLE_128B_ATOM_32B
                                 : CU_TENSOR_MAP_SWIZZLE_NONE,
                  CU_TENSOR_MAP_L2_PROMOTION_L2_256B, CU_TENSOR_MAP_FLOAT_OOB_FILL_NONE);
  if (r != CUDA_SUCCESS)
    return set_error(DMLP_ERR_CUDA, "cuTensorMapEncodeTiled failed with CUresult %d", (int)r);
  return DMLP_OK;
}

template <int EB, bool A_MN, bool B_MN, bool TWO, int E2>
static int launch_gemm_t(const CUtensorMap& tmA, const CUtensorMap& tmB, const CUtensorMap& tmC, const GemmParams& p,
                         dim3 grid, size_t smem_bytes, cudaStream_t stream) {
  auto kern = gemm_tcgen05_kernel<EB, A_MN, B_MN, TWO, E2>;
  constexpr int kThreads = E2 == 2 ? kGemmThreadsRS : E2 ? kGemmThreadsE2 : kGemmThreads;
  // devices on which the smem opt-in has been applied (gRPC worker threads launch GEMMs on different devices at once)
  static std::atomic<unsigned long long> attr_mask{0};
  int dev = 0;
  DMLP_CUDA_CHECK(cudaGetDevice(&dev));
  if (!(attr_mask.load(std::memory_order_acquire) & (1ull << dev))) {
    cudaFuncAttributes fa;
    DMLP_CUDA_CHECK(cudaFuncGetAttributes(&fa, kern));  // 227 KiB per CTA = static + dynamic
    DMLP_CUDA_CHECK(cudaFuncSetAttribute(kern, cudaFuncAttributeMaxDynamicSharedMemorySize,
                                         227 * 1024 - static_cast<int>((fa.sharedSizeBytes + 1023) / 1024 * 1024)));
    attr_mask.fetch_or(1ull << dev, std::memory_order_release);
  }
  if (p.pdl || TWO) {
    cudaLaunchConfig_t cfg = {};
    cfg.gridDim = grid;
    cfg.blockDim = dim3(kThreads);
    cfg.dynamicSmemBytes = smem_bytes;
    cfg.stream = stream;
    cudaLaunchAttribute attr[2];
    int na = 0;
    if (p.pdl) {
      attr[na].id = cudaLaunchAttributeProgrammaticStreamSerialization;
      attr[na].val.programmaticStreamSerializationAllowed = 1;
      ++na;
    }
    if (TWO) {  // CTA pairs: clusters of 2 (ranks 2i, 2i+1 share a TPC)
      attr[na].id = cudaLaunchAttributeClusterDimension;
      attr[na].val.clusterDim.x = 2;
      attr[na].val.clusterDim.y = 1;
      attr[na].val.clusterDim.z = 1;
      ++na;
    }
    cfg.attrs = attr;
    cfg.numAttrs = na;
    DMLP_CUDA_CHECK(cudaLaunchKernelEx(&cfg, kern, tmA, tmB, tmC, p));
  } else {
    kern<<<grid, kThreads, smem_bytes, stream>>>(tmA, tmB, tmC, p);
  }
  DMLP_CUDA_CHECK(cudaGetLastError());
  return DMLP_OK;
}

template <int EB, bool A_MN, bool B_MN, bool TWO = false>
static int launch_gemm(const GemmPlan* pl, dim3 grid, cudaStream_t stream) {
  if constexpr (A_MN && B_MN) {  // (weight gradients: both operands MN-major)
    if (pl->p.rs_world > 0) return launch_gemm_t<EB, A_MN, B_MN, TWO, 2>(pl->tmA, pl->tmB, pl->tmC, pl->p, grid, pl->smem, stream);
  }
  if (pl->p.epi2) return launch_gemm_t<EB, A_MN, B_MN, TWO, 1>(pl->tmA, pl->tmB, pl->tmC, pl->p, grid, pl->smem, stream);
  return launch_gemm_t<EB, A_MN, B_MN, TWO, 0>(pl->tmA, pl->tmB, pl->tmA, pl->p, grid, pl->smem, stream);
}

static int ceil_div(int a, int b) { return (a + b - 1) / b; }

int choose_block_n(int M, int N) {
  const int tiles_m = ceil_div(M, 128);
  int bn = ((N + 15) / 16) * 16;
  if (bn > 256) bn = 256;
  // Small problems are latency bound: spread N over more CTAs (down to the N=16 MMA minimum).
  while (bn > 16 && (bn / 2) % 16 == 0 && tiles_m * ceil_div(N, bn) < 148) bn /= 2;
  return bn;
}

}  // namespace dmlp

using namespace dmlp;

int dmlp::gemm_plan_create(const dmlp_gemm_desc_t* d, GemmPlan* plan) {
  DMLP_REQUIRE(d != nullptr && plan != nullptr, "dmlp_gemm: null descriptor");
  DMLP_REQUIRE(d->dtype == DMLP_F32 || d->dtype == DMLP_BF16, "dmlp_gemm: dtype must be F32 (tf32) or BF16");
  DMLP_REQUIRE(d->M > 0 && d->N > 0 && d->K > 0, "dmlp_gemm: empty problem %dx%dx%d", d->M, d->N, d->K);
  DMLP_REQUIRE(d->A_dev && d->B_dev, "dmlp_gemm: null operand");
  DMLP_REQUIRE(d->out_f32_dev || d->out_bf16_dev || d->colsum_dev || d->ce_labels_dev, "dmlp_gemm: no output requested");
  const bool bf16 = d->dtype == DMLP_BF16;
  const int EB = bf16 ? 2 : 4;
  const int KS = 128 / EB;
  const int ATOM = 128 / EB;
  int bn = d->block_n > 0 ? d->block_n : choose_block_n(d->M, d->N);
  if (d->ce_labels_dev != nullptr) bn = ((d->N + 15) / 16) * 16;  // fused softmax-CE: one N tile holds the whole row
  DMLP_REQUIRE(bn % 16 == 0 && bn >= 16 && bn <= 256, "dmlp_gemm: block_n %d must be a multiple of 16 in [16,256]", bn);

  // CTA pairs (tcgen05.mma.cta_group::2, 256 x 256 tiles) for problems with at least one full wave of pair tiles
  bool two = false;
  {
    static int pair_env = -1;
    if (pair_env < 0) { const char* e = getenv("DMLP_GEMM_PAIR"); pair_env = e ? atoi(e) : 1; }
    two = pair_env != 0 && bn == 256 && d->ce_labels_dev == nullptr && d->M >= 256 &&
          ceil_div(d->M, 256) * ceil_div(d->N, 256) >= 74;
  }
  const int b_cols = two ? bn / 2 : bn;  // B columns one CTA stages
  int rc;
  const int mn_swz = bf16 ? 1 : 2;  // fp32 MN-major operands need the 32B-atom swizzle
  // A problem with M <= 64 (the reference's batch) only fills half of the 128-row MMA tile. Loading
  // just the live rows halves the TMA traffic of these latency-bound launches; the MMA still reads
  // 128 smem rows, and whatever the dead rows hold only reaches accumulator rows that are never stored.
  const int a_box_rows = (d->a_major == 0 && d->M <= 64) ? 64 : 128;
  if (d->a_major == 0)
    rc = make_tensor_map_2d(&plan->tmA, d->A_dev, EB, bf16, d->K, d->M, d->lda, KS, a_box_rows, 1);
  else
    rc = make_tensor_map_2d(&plan->tmA, d->A_dev, EB, bf16, d->M, d->K, d->lda, ATOM, KS, mn_swz);
  if (rc) return rc;
  if (d->b_major == 0)
    rc = make_tensor_map_2d(&plan->tmB, d->B_dev, EB, bf16, d->K, d->N, d->ldb, KS, b_cols, 1);
  else
    rc = make_tensor_map_2d(&plan->tmB, d->B_dev, EB, bf16, d->N, d->K, d->ldb, ATOM, KS, mn_swz);
  if (rc) return rc;

  GemmParams& p = plan->p;
  p = GemmParams{};
  p.M = d->M; p.N = d->N; p.K = d->K;
  p.block_n = bn;
  p.num_k_blocks = ceil_div(d->K, KS);
  p.a_stage_bytes = 128 * 128;
  p.a_tx_bytes = (d->a_major == 0) ? a_box_rows * 128 : 128 * 128;
  p.b_stage_bytes = d->b_major == 0 ? b_cols * 128 : ceil_div(b_cols, ATOM) * KS * 128;  // per CTA
  const size_t stage_bytes = p.a_stage_bytes + p.b_stage_bytes;
  // TMA-store epilogue with eight epilogue warps (E2 kernels): outputs that are a plain / bias / ReLU function of the
  // accumulator, one output tensor, tiles whose column halves split into whole 128-byte chunks.
  {
    static int e2_env = -1;
    if (e2_env < 0) { const char* e = getenv("DMLP_GEMM_EPI2"); e2_env = e ? atoi(e) : 1; }
    const bool one_out = (d->out_f32_dev != nullptr) != (d->out_bf16_dev != nullptr);
    const void* optr = d->out_f32_dev ? static_cast<const void*>(d->out_f32_dev) : d->out_bf16_dev;
    const long long old = d->out_f32_dev ? d->ld_out * 4 : d->ld_out_bf16 * 2;  // output row pitch in bytes
    const int half_bytes = (bn / 2) * (d->out_f32_dev ? 4 : 2);
    p.epi2 = (e2_env != 0 && one_out && d->mask_dev == nullptr && d->colsum_dev == nullptr && !d->accumulate &&
              d->ce_labels_dev == nullptr && d->M >= 128 && half_bytes % 128 == 0 && old % 16 == 0 &&
              (reinterpret_cast<uintptr_t>(optr) & 15) == 0 && (reinterpret_cast<uintptr_t>(d->bias_dev) & 15) == 0) ? 1 : 0;
    if (p.epi2) {
      if (d->out_f32_dev) rc = make_tensor_map_2d(&plan->tmC, d->out_f32_dev, 4, false, d->N, d->M, d->ld_out, 32, 32, 1);
      else rc = make_tensor_map_2d(&plan->tmC, d->out_bf16_dev, 2, true, d->N, d->M, d->ld_out_bf16, 64, 32, 1);
      if (rc) return rc;
    }
  }
  const size_t epi_bytes = p.epi2 ? kE2StageBytes : kEpiStageBytes;
  // stages + epilogue staging (E2: 32 KiB, 8 store buffers; else 18 KiB of transposing patches) + 1 KiB alignment slack
  const size_t budget = (p.epi2 ? 225 : 214) * 1024 - epi_bytes;  // static shared memory: ~1 KiB (E2) / ~5 KiB
  int stages = static_cast<int>(budget / stage_bytes);
  if (stages > kMaxStages) stages = kMaxStages;
  if (stages > p.num_k_blocks) stages = p.num_k_blocks;
  if (stages < 1) stages = 1;
  {
    const char* e = getenv("DMLP_GEMM_STAGES");  // experiment knob
    if (e && atoi(e) > 0 && atoi(e) < stages) stages = atoi(e);
  }
  p.stages = stages;
  {
    const int tm = two ? 256 : 128, tiles_m = ceil_div(d->M, tm);
    int g = 2048 / tm;  // band height: 2048 rows
    const char* e = getenv("DMLP_GEMM_RASTER");  // experiment knob: band height in tiles (0: plain M-fastest)
    if (e) g = atoi(e);
    p.raster_g = (g <= 0 || g > tiles_m) ? tiles_m : g;
  }
  {
    const char* e = getenv("DMLP_GEMM_PROF");
    p.prof = (e && atoi(e) > 0) ? 1 : 0;
  }
  uint32_t cols = 32;
  while (cols < static_cast<uint32_t>(bn)) cols <<= 1;
  p.tmem_buf_cols = cols;   // one accumulator buffer
  p.tmem_cols = 2 * cols;   // two of them (<= 512 columns)
  p.idesc = make_instr_desc(EB, d->a_major != 0, d->b_major != 0, bn);
  if (two) p.idesc = (p.idesc & ~(0x1Fu << 24)) | (static_cast<uint32_t>(256 >> 4) << 24);  // M = 256 across the pair
  p.out_f32 = d->out_f32_dev; p.ld_out = d->ld_out;
  p.out_bf16 = reinterpret_cast<__nv_bfloat16*>(d->out_bf16_dev); p.ld_out_bf16 = d->ld_out_bf16;
  p.bias = d->bias_dev;
  p.mask = d->mask_dev; p.ld_mask = d->ld_mask; p.mask_is_bf16 = d->mask_dtype == DMLP_BF16;
  p.colsum = d->colsum_dev;
  p.relu = d->relu; p.accumulate = d->accumulate;
  p.ce_labels = d->ce_labels_dev; p.ce_classes = d->ce_classes; p.ce_scale = d->ce_scale;
  p.ce_logits = d->ce_logits_dev; p.ce_loss_sum = d->ce_loss_sum_dev; p.ce_correct = d->ce_correct_dev;
  if (d->ce_labels_dev != nullptr) {
    DMLP_REQUIRE(d->N <= 32 && d->ce_classes >= 1 && d->ce_classes <= d->N, "fused softmax-CE needs classes <= N <= 32 (got N=%d classes=%d)", d->N, d->ce_classes);
    DMLP_REQUIRE(!d->relu && !d->mask_dev && !d->accumulate, "fused softmax-CE excludes relu/mask/accumulate");
  }

  {
    const int total = ceil_div(d->N, bn) * ceil_div(d->M, two ? 256 : 128);
    plan->total_tiles = total;
    plan->grid_x = total < 148 ? total : 148;  // persistent: at most one CTA per SM
    if (two) plan->grid_x = 2 * (total < 74 ? total : 74);  // one CTA pair (cluster of 2) per TPC
    plan->grid_y = 1;
  }
  p.epi_stage_off = static_cast<uint32_t>(stages * stage_bytes);
  plan->smem = stages * stage_bytes + epi_bytes + 1024;
  plan->key = (two ? 8 : 0) | (bf16 ? 4 : 0) | (d->a_major ? 2 : 0) | (d->b_major ? 1 : 0);
  // PDL only for single-wave grids: an early-launched dependent must never compete with CTAs of
  // this grid that are still waiting for an SM.
  {
    static int pdl_env = -1;
    if (pdl_env < 0) { const char* e = getenv("DMLP_PDL"); pdl_env = e ? atoi(e) : 1; }
    p.pdl = (pdl_env && !two && plan->total_tiles <= 148) ? 1 : 0;
  }
  return DMLP_OK;
}

bool dmlp::gemm_plan_rs_eligible(const GemmPlan* plan, int world) {
  const GemmParams& p = plan->p;
  const bool two = (plan->key & 8) != 0;
  const int tm = two ? 256 : 128;
  // Validated on hardware for CTA-pair plans with 256-column tiles (2 and 8 GPUs, bf16 and tf32); the single-CTA / narrower
  // tile instantiations follow the same index arithmetic but stay opt-in (DMLP_WGRAD_RS_ANY=1) until they have been run.
  static int any_env = -1;
  if (any_env < 0) { const char* e = getenv("DMLP_WGRAD_RS_ANY"); any_env = (e && atoi(e) > 0) ? 1 : 0; }
  if (!any_env && !(two && p.block_n == 256)) return false;
  return world >= 2 && world <= 8 && p.epi2 && (plan->key & 3) == 3 && p.out_f32 != nullptr && p.out_bf16 == nullptr &&
         p.M % tm == 0 && p.N % p.block_n == 0 && (p.block_n / 2) % 32 == 0 && plan->total_tiles % world == 0;
}

int dmlp::gemm_plan_enable_rs(GemmPlan* plan, int world, int rank, float* const* slots, uint32_t* const* counts,
                              GemmRsGeometry* geo) {
  DMLP_REQUIRE(plan && slots && counts && geo && rank >= 0 && rank < world, "gemm_plan_enable_rs: bad arguments");
  DMLP_REQUIRE(gemm_plan_rs_eligible(plan, world), "gemm_plan_enable_rs: the plan cannot push whole tiles to %d owners", world);
  GemmParams& p = plan->p;
  const bool two = (plan->key & 8) != 0;
  p.rs_world = world;
  p.rs_rank = rank;
  p.rs_tiles_per_owner = plan->total_tiles / world;
  for (int r = 0; r < 8; ++r) {
    p.rs_slot[r] = r < world ? slots[r] : nullptr;
    p.rs_count[r] = r < world ? counts[r] : nullptr;
    DMLP_REQUIRE(r >= world || (p.rs_slot[r] && p.rs_count[r]), "gemm_plan_enable_rs: rank %d is not mapped", r);
  }
  p.pdl = 0;
  { const char* pe = getenv("DMLP_RS_PROF"); if (pe && atoi(pe) > 0) p.prof = 1; }
  geo->tm = two ? 256 : 128;
  geo->bn = p.block_n;
  geo->subs = two ? 2 : 1;
  geo->tiles_m = p.M / geo->tm;
  geo->tiles_n = p.N / p.block_n;
  geo->raster_g = p.raster_g;
  geo->total_tiles = plan->total_tiles;
  geo->grid = plan->grid_x;
  return DMLP_OK;
}

int dmlp::gemm_plan_run(const GemmPlan* plan, cudaStream_t st) {
  const dim3 grid(plan->grid_x, plan->grid_y);
  switch (plan->key) {
    case 0: return launch_gemm<4, false, false>(plan, grid, st);
    case 1: return launch_gemm<4, false, true>(plan, grid, st);
    case 2: return launch_gemm<4, true, false>(plan, grid, st);
    case 3: return launch_gemm<4, true, true>(plan, grid, st);
    case 4: return launch_gemm<2, false, false>(plan, grid, st);
    case 5: return launch_gemm<2, false, true>(plan, grid, st);
    case 6: return launch_gemm<2, true, false>(plan, grid, st);
    case 7: return launch_gemm<2, true, true>(plan, grid, st);
    case 8: return launch_gemm<4, false, false, true>(plan, grid, st);
    case 9: return launch_gemm<4, false, true, true>(plan, grid, st);
    case 10: return launch_gemm<4, true, false, true>(plan, grid, st);
    case 11: return launch_gemm<4, true, true, true>(plan, grid, st);
    case 12: return launch_gemm<2, false, false, true>(plan, grid, st);
    case 13: return launch_gemm<2, false, true, true>(plan, grid, st);
    case 14: return launch_gemm<2, true, false, true>(plan, grid, st);
    default: return launch_gemm<2, true, true, true>(plan, grid, st);
  }
}

extern "C" int dmlp_gemm(const dmlp_gemm_desc_t* d, void* stream) {
  GemmPlan plan;
  int rc = gemm_plan_create(d, &plan);
  if (rc) return rc;
  return gemm_plan_run(&plan, as_stream(stream));
}

// Debug: the consumer warps' per-CTA timeline of the most recent reduce-scatter GEMM (DMLP_RS_PROF=1), 148 x 4 stamps.
extern "C" int dmlp_wgrad_rs_prof_read(unsigned long long* host, int n_words) {
  DMLP_REQUIRE(host != nullptr && n_words > 0 && n_words <= 148 * 4, "dmlp_wgrad_rs_prof_read: bad arguments");
  DMLP_CUDA_CHECK(cudaMemcpyFromSymbol(host, dmlp::g_rs_prof, sizeof(unsigned long long) * n_words));
  return DMLP_OK;
}

// Debug: copies the per-tile timeline of CTA 0 of the most recent profiled GEMM (DMLP_GEMM_PROF=1) to the host.
extern "C" int dmlp_gemm_prof_read(unsigned long long* host, int n_words) {
  DMLP_REQUIRE(host != nullptr && n_words > 0 && n_words <= dmlp::kProfTiles * 16, "dmlp_gemm_prof_read: bad arguments");
  DMLP_CUDA_CHECK(cudaMemcpyFromSymbol(host, dmlp::g_gemm_prof, sizeof(unsigned long long) * n_words));
  return DMLP_OK;
}
