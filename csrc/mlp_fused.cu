// Cluster-fused training step for the reference-sized MLP (sm_100a): the whole forward, loss and
// backward of one batch in ONE launch of one 8-CTA thread-block cluster.
//
// Why: at batch 64 every layer is latency bound: the layer-1 forward is 98 dependent tcgen05.mma in a
// row and each per-layer kernel pays ~3.3 us of fixed cost (profiles/r1_gemm_kloop_probe.md). Here the K
// dimension of every layer is split over the 8 CTAs of a cluster; partial sums travel through
// distributed shared memory (st.async: the bytes complete on an mbarrier of the DESTINATION CTA, no
// cluster barrier on the chain) and the output slice a CTA owns after the reduce-scatter is exactly the K
// slice it needs for the next layer:
//
//   P0  TMA: X[:, k-slice], W1[k-slice, :]            F1: P1 = X_j . W1_j            -> reduce-scatter
//   P1  H1_j = relu(sum P1 + b1_j)                     F2: P2 = H1_j . W2_j           -> reduce-scatter
//   P2  H2_j = relu(sum P2 + b2_j)                     F3: P3 = H2_j . W3_j           -> all-reduce
//   P3  logits -> softmax-CE -> dZ3 (every CTA)        dW3_j = H2_j^T dZ3 ; dH2_j = dZ3 W3_j^T
//       dZ2_j = dH2_j * (H2_j > 0)                                                    -> all-gather
//   P4  dW2_j = H1_j^T dZ2 ; dH1_j = dZ2 W2_j^T ; dZ1_j = dH1_j * (H1_j > 0)          -> all-gather
//   P5  dW1_j = X_j^T dZ1
//
// Every product is a tcgen05.mma (kind::tf32, M=128, accumulator in TMEM); weights and inputs come
// in by TMA, activations / gradients are written to shared memory by the epilogue threads in the
// canonical swizzled layouts the MMA descriptors expect.
//
// What happens to the gradients (FusedParams):
//   * wparams set, world 1      : w -= lr*g in place (red.global.add.v4.f32); a launch loops over several steps
//   * wparams null              : gradients are stored in the flat layout; a separate all-reduce+SGD kernel follows
//   * red_ctas > 0 (world > 1)  : persistent data-parallel launch - this cluster computes, the other CTAs of the grid
//                                 run the one-shot all-reduce + average + SGD over peer memory (fused_reducer_loop)
//   * world > 1, red_ctas == 0  : gradient exchange inside the cluster (bulk-copy push, rank-ordered reduce): opt-in
//   * gx (world > 1)            : global-batch weight gradients - the ranks exchange dZ1 (32 KB) and their inputs instead of
//                                 dW1 (401 KB); every rank computes the layer-1 gradient of the GLOBAL batch itself, in rank
//                                 order, and applies it in place. See the "global-batch" block below (default on real GPUs).
//
// Reference math: DSML/client/client.go:112-202 (forward, softmax-CE, backward).
// Supported shapes: batch <= 64 (multiple of 8), in0 % 8 == 0 and <= 832, H1 = 128, H2 in {64, none},
// classes <= 16. Anything else uses the per-layer kernels.
#include "common.h"
#include "ptx_sm100.cuh"

#include <stdlib.h>
#include <string.h>

#include <memory>
#include <new>

namespace dmlp {

int make_tensor_map_2d(CUtensorMap* out, const void* base, int elem_bytes, bool is_bf16, uint64_t cols,
                       uint64_t rows, uint64_t ld_elems, uint32_t box_cols, uint32_t box_rows, int swizzle);

constexpr int kFC = 8;          // CTAs per cluster (one cluster per step)
constexpr int kFThreads = 160;  // warp 0: TMA + MMA issue; warps 1..4: one thread per batch row
constexpr int kFBK = 64;        // batch rows a tile holds
constexpr int kKGroupsMax = 13; // 8-float K groups of layer 1 per CTA (784/8 = 98 groups over 8 CTAs)

struct FusedParams {
  int B, in0, C;
  const float* params;
  float* grads;
  unsigned long long w_off[3], b_off[3];
  const int* labels;
  float* metrics;  // [0] loss sum, [1] hit count (int bits); may be host-mapped memory
  float* wparams;  // non-null: single-GPU mode, apply w -= lr*g in place instead of writing gradients
  const float* lr;
  float* dbg;      // optional dump of intermediates (tests)
  unsigned long long* prof;  // optional [8 CTAs][64] globaltimer stamps at phase boundaries
  // ---- in-kernel data-parallel gradient exchange (world > 1), see the "gradient exchange" block in the kernel
  int world, rank;
  float* xrecv[8];          // rank r's receive area as mapped here: [2 parities][world sources][total floats]
  uint32_t* xpad[8];        // rank r's signal pad (flags at kPadFusedFlags + src*8 + cta; epoch word at kPadFusedEpoch)
  uint32_t* xstatus;        // host-mapped failure record {code, rank, peer, phase} (watchdog), may be null
  unsigned long long xtimeout_ns;
  unsigned long long total;  // floats in the flat parameter vector
  // ---- persistent data-parallel mode (red_ctas > 0): the grid is [1 compute cluster | red_ctas reducer CTAs]. The compute
  // cluster writes its gradients to xgrads[parity][rank] and raises local flags; the reducers (spinning since launch)
  // rendezvous with the same reducer CTA on every rank, pull all ranks' gradients over NVLink, apply w -= lr/world * sum
  // and raise a local counter the compute cluster waits on before the next step's weight loads. No kernel boundary
  // between step and all-reduce, several steps per launch.
  int red_ctas;
  float* xgrads[2][8];
  int bulk_update; // single GPU: apply the update with bulk reduce-adds from a staged image (else red.global per thread)
  int nsteps;      // steps this launch performs back to back (> 1 only with in-place SGD: wparams != null)
  int slot0;       // first input slot: step s reads rows [(slot0+s)*B, +B) of the X maps, labels + s*B, metrics + 2*s
  // ---- global-batch mode (gx != 0, world > 1, wparams != null). The mean gradient of data-parallel SGD is the gradient of
  // the global batch: dW1 = sum_r X_r^T dZ1_r. For this model the factors are far smaller than the product (64 x (784 + 128)
  // floats per rank against 784 x 128), so the ranks exchange the factors: inputs are all-gathered ahead of time by a copy
  // kernel (input_exchange.cu), dZ1 slices are pushed into every peer's receive area the moment they exist, and every rank
  // accumulates the K = world x batch product itself in rank order - bit-identical replicas by construction, no reducers,
  // no rendezvous. The small tensors (W2, W3, b2, b3: 36 KB) are all-reduced inside the kernel: each CTA pushes its slices
  // to every rank and sums the arrivals in rank order. Arrival is detected on the data itself (the receive areas hold a
  // NaN pattern no arithmetic produces until real values land), so a transfer costs one NVLink traversal: no flag, no fence.
  int gx;
  int xslots;                   // input ring slots per rank: rank r's batch of ring slot s starts at row (r*xslots + s)*B of the X maps
  unsigned long long g0;        // global index of this launch's first step (parity of the receive areas)
  long long fault_step;         // fault injection: from this global step on this rank sends nothing and waits for nothing (-1: never)
  unsigned int xneed;           // the peers' inputs for this launch are in place once their ready words reach this step count
  // DMLP_COLL_STRICT (same switch as collectives.cu). 2 (default): flag polls are ld.acquire.sys; flags are published
  // behind a gpu-scope fence, which relies on this GPU's L2 being the point of coherence for its own memory (peer reads
  // of it are served there). 1: the conservative protocol - system-scope fence before every flag published to a peer and
  // after every flag wait. 0: relaxed polls, no acquire (round 1's protocol). docs/FUSED_STEP.md, section 6a.
  int strict;
};

#define FPROF(slot)                                                                                   \
  do {                                                                                                \
    if (p.prof != nullptr && threadIdx.x == 32) p.prof[j * 64 + (slot)] = globaltimer_ns();             \
  } while (0)
// fine-grained stamps of the global-batch tail (slots 16..63): thread `tid` only
#define FPROFT(tid, slot)                                                                             \
  do {                                                                                                \
    if (GX && p.prof != nullptr && threadIdx.x == (tid)) p.prof[j * 64 + (slot)] = globaltimer_ns();    \
  } while (0)

// ---- shared-memory map (bytes). A-operand tiles are written for 64 rows (8 KiB per 128-byte-wide
// K block); the MMA reads 128 rows, i.e. 8 KiB past each tile: whatever follows it, harmlessly.
constexpr uint32_t kTileK = 8192;                      // K-major tile: 64 rows x 128 B
constexpr uint32_t kColM = kFBK * 128;                 // MN-major atom column: 64 K-rows x 128 B
constexpr uint32_t kW1Col = kKGroupsMax * 8 * 128;     // W1 slice atom column: 104 K-rows x 128 B
constexpr uint32_t oXA = 0;                            // 4 K blocks (P0)            | RS2, DZ2K later
constexpr uint32_t oW1B = oXA + 4 * kTileK;            // 4 atoms x 13 KiB (P0)      | AR3, DZ2M, DZ1M later
constexpr uint32_t oRS1 = oW1B + 4 * kW1Col;           // 8 src x 64 rows x 64 B     | H2K, DZ3K, H2M, DZ3M later
constexpr uint32_t oXT = oRS1 + 8 * 64 * 64;           // 4 atoms x 8 KiB (used in P5)
constexpr uint32_t oH1K = oXT + 4 * kColM;
constexpr uint32_t oH1M = oH1K + kTileK;
constexpr uint32_t oW2B = oH1M + kColM;                // 2 atoms x 16 rows x 128 B (+ slack)
constexpr uint32_t oW2K = oW2B + 4096;                 // 2 K blocks x 32 rows x 128 B
constexpr uint32_t oWLB = oW2K + 8192;
constexpr uint32_t oWLK = oWLB + 4096;
constexpr uint32_t oRED = oWLK + 4096;                 // reductions scratch
constexpr uint32_t oEND = oRED + 4096;
constexpr uint32_t kFusedSmem = oEND + 32768 + 1024;   // slack: MMA over-read of MN-major A atoms + alignment
// aliases (lifetimes are disjoint, see the phase table in the header comment)
// Gradient exchange (data parallel): staging image of this CTA's gradients and the receive stages of the reduction.
// All of it aliases regions that are dead by then (the W1 image is written after the last MMA has completed; the
// small images sit in the layer-2/3 weight tiles, which no MMA reads after dgrad2).
constexpr uint32_t kXsW1 = oXA;              // [nk1*8][128] fp32 <= 53,248 B
constexpr uint32_t kXsW2 = oW2B;             // [16][64]
constexpr uint32_t kXsWL = oW2B + 4096;      // [SL][16]
constexpr uint32_t kXsB1 = oW2B + 5120, kXsB2 = oW2B + 5184, kXsBL = oW2B + 5248;
constexpr uint32_t kXChunk = kKGroupsMax * 2 * 128 * 4;  // 13,312 B: a quarter of the W1 slice
constexpr uint32_t kXr = kKGroupsMax * 8 * 128 * 4;      // receive stages start right after the W1 image ...
static_assert(kXr + 8 * kXChunk <= oW2B, "receive stages run into the small staging images");  // ... 8 chunk buffers
constexpr uint32_t oRS2 = oXA;                         // 8 src x 64 rows x 32 B = 16 KiB
constexpr uint32_t oDZ2K = oXA + 16384;                // 2 K blocks x 8 KiB
constexpr uint32_t oAR3 = oW1B;                        // 8 src x 64 rows x 64 B = 32 KiB
constexpr uint32_t oDZ2M = oW1B + 32768;               // 2 atoms x 8 KiB
constexpr uint32_t oDZ1M = oW1B;                       // 4 atoms x 8 KiB (three-layer model)
constexpr uint32_t oDZ1M2 = oXA;                       // same, two-layer model: a peer may deliver dZ1 while this CTA
                                                       // still reads the logits exchange (AR3), so it must not alias it
constexpr uint32_t oH2K = oRS1;
constexpr uint32_t oDZ3K = oRS1 + 8192;
constexpr uint32_t oH2M = oRS1 + 16384;
constexpr uint32_t oDZ3M = oRS1 + 24576;
// Global-batch mode: two stages for the peers' operands of the layer-1 weight gradient (X_r^T tile + dZ1_r image, 32 KiB
// each), in regions whose last readers are the weight-gradient GEMMs that run beside the chain (complete before P5).
constexpr uint32_t oGX0 = oXA;               // stage 0 X^T (three-layer model; the two-layer model keeps its dZ1 in XA ...
constexpr uint32_t oGX0b = oW1B;             // ... and uses the W1 region instead)
constexpr uint32_t oGD0 = oRS1;              // stage 0 dZ1
constexpr uint32_t oGX1 = oH1K;              // stage 1 X^T: H1K, H1M, W2B, W2K, WLB
constexpr uint32_t oGD1 = oH1K + 32768;      // stage 1 dZ1: WLK, RED and 24 KiB of the over-read slack
static_assert(oGD1 == oWLK && oGD1 + 32768 <= kFusedSmem - 1024, "stage 1 runs past the shared-memory map");
// Staging for pushes over NVLink (the last 8 KiB of the over-read slack: MMAs only ever READ there, into accumulator rows
// nobody uses). An SM gets a remote write out every ~11 ns whatever its size, so what is pushed is first laid out
// contiguously here and then stored with full-line, lane-contiguous 16-byte stores: 32 transactions per 4 KiB, not 512.
constexpr uint32_t oSTG = oEND + 24576;     // [0, 4K): dZ1 slice, 2 KiB per warp holding rows < 64 (later: last-layer dW slice)
constexpr uint32_t oSTG2 = oSTG + 4096;     // [4K, 8K): dW2 slice
static_assert(oSTG2 + 4096 <= kFusedSmem - 1024, "push staging runs past the shared-memory map");
static_assert(kFusedSmem <= 227 * 1024, "fused step does not fit in shared memory");

// TMEM columns
constexpr uint32_t cF1 = 0, cF2 = 128, cF3 = 192, cWGL = 208, cDGL = 224, cWG2 = 256, cDG2 = 320, cWG1 = 384;

// ---- layouts written by threads -------------------------------------------------------------
// K-major, SWIZZLE_128B: row r (batch), element k (< 32 per 128-byte block).
__device__ __forceinline__ uint32_t koff(int r, int k) {
  return (k >> 5) * kTileK + (r >> 3) * 1024 + (r & 7) * 128 + ((((k & 31) >> 2) ^ (r & 7)) << 4) + (k & 3) * 4;
}
// MN-major, SWIZZLE_128B_BASE32B: K-row r (batch), MN element n (32 per atom column of kColM bytes).
__device__ __forceinline__ uint32_t moff(int r, int n) {
  return (n >> 5) * kColM + r * 128 + ((((n & 31) >> 3) ^ (r & 3)) << 5) + (n & 7) * 4;
}
__device__ __forceinline__ uint32_t mapa(uint32_t a, uint32_t cta) {
  uint32_t r;
  asm volatile("mapa.shared::cluster.u32 %0, %1, %2;" : "=r"(r) : "r"(a), "r"(cta));
  return r;
}
__device__ __forceinline__ void st_cluster_f4(uint32_t a, float x, float y, float z, float w) {
  asm volatile("st.shared::cluster.v4.f32 [%0], {%1, %2, %3, %4};" ::"r"(a), "f"(x), "f"(y), "f"(z), "f"(w) : "memory");
}
// Remote store whose completion is counted (in bytes) on an mbarrier in the destination CTA.
__device__ __forceinline__ void st_async_f4(uint32_t a, uint32_t bar, float x, float y, float z, float w) {
  asm volatile("st.async.weak.shared::cluster.mbarrier::complete_tx::bytes.v4.f32 [%0], {%1, %2, %3, %4}, [%5];" ::"r"(a),
               "f"(x), "f"(y), "f"(z), "f"(w), "r"(bar)
               : "memory");
}
__device__ __forceinline__ void cluster_sync_all() {
  asm volatile("barrier.cluster.arrive.release.aligned;\n\tbarrier.cluster.wait.acquire.aligned;" ::: "memory");
}
__device__ __forceinline__ uint32_t cluster_ctarank() {
  uint32_t r;
  asm volatile("mov.u32 %0, %%cluster_ctarank;" : "=r"(r));
  return r;
}
__device__ __forceinline__ void tmem_ld_x8(uint32_t taddr, uint32_t (&r)[8]) {
  asm volatile("tcgen05.ld.sync.aligned.32x32b.x8.b32 {%0, %1, %2, %3, %4, %5, %6, %7}, [%8];"
               : "=r"(r[0]), "=r"(r[1]), "=r"(r[2]), "=r"(r[3]), "=r"(r[4]), "=r"(r[5]), "=r"(r[6]), "=r"(r[7])
               : "r"(taddr) : "memory");
}
// Bulk (non-tensor) asynchronous copies: global -> shared with mbarrier completion, shared -> global in bulk groups.
__device__ __forceinline__ void bulk_g2s(uint32_t dst_s, const void* src, uint32_t bytes, uint64_t* bar) {
  asm volatile("cp.async.bulk.shared::cluster.global.mbarrier::complete_tx::bytes [%0], [%1], %2, [%3];" ::"r"(dst_s), "l"(src),
               "r"(bytes), "r"(smem_u32(bar))
               : "memory");
}
__device__ __forceinline__ void bulk_s2g(void* dst, uint32_t src_s, uint32_t bytes) {
  asm volatile("cp.async.bulk.global.shared::cta.bulk_group [%0], [%1], %2;" ::"l"(dst), "r"(src_s), "r"(bytes) : "memory");
}
// global[i] += shared[i] (fp32) for a contiguous range, performed by the bulk copy engine
__device__ __forceinline__ void bulk_red_add_f32(void* dst, uint32_t src_s, uint32_t bytes) {
  asm volatile("cp.reduce.async.bulk.global.shared::cta.bulk_group.add.f32 [%0], [%1], %2;" ::"l"(dst), "r"(src_s), "r"(bytes)
               : "memory");
}
__device__ __forceinline__ void bulk_commit() { asm volatile("cp.async.bulk.commit_group;" ::: "memory"); }
__device__ __forceinline__ void bulk_wait_all() { asm volatile("cp.async.bulk.wait_group 0;" ::: "memory"); }
__device__ __forceinline__ void bulk_wait_read() { asm volatile("cp.async.bulk.wait_group.read 0;" ::: "memory"); }
__device__ __forceinline__ void fence_proxy_async_all() { asm volatile("fence.proxy.async;" ::: "memory"); }
// L2 prefetch of one TMA box (no shared-memory destination, no barrier)
__device__ __forceinline__ void tma_prefetch_2d(const CUtensorMap* m, int c0, int c1) {
  asm volatile("cp.async.bulk.prefetch.tensor.2d.L2.global.tile [%0, {%1, %2}];" ::"l"(reinterpret_cast<uint64_t>(m)), "r"(c0),
               "r"(c1)
               : "memory");
}
__device__ __forceinline__ uint64_t kdesc(uint32_t tile, int kstep) {  // K-major A/B, 8 floats per step
  return make_smem_desc(tile + (kstep >> 2) * kTileK + (kstep & 3) * 32, 16, 1024, 2);
}
__device__ __forceinline__ uint64_t mdesc(uint32_t tile, int kstep, uint32_t col_bytes) {  // MN-major, 8 K-rows per step
  return make_smem_desc(tile + kstep * 1024, col_bytes, 512, 1);
}

// Store that is performed at system scope right away (peer-visible data whose arrival the receiver polls for)
__device__ __forceinline__ void st_volatile_f4(float* p, float a, float b, float c, float d) {
  asm volatile("st.volatile.global.v4.f32 [%0], {%1, %2, %3, %4};" ::"l"(p), "f"(a), "f"(b), "f"(c), "f"(d) : "memory");
}
// Emits 4 consecutive gradient values: either into the gradient buffer (multi-GPU: the all-reduce
// kernel consumes it) or straight into the weights as w -= lr*g (single GPU).
// Exchange mode (xs != 0): the gradient goes to this CTA's staging image in shared memory (byte offset soff); the bulk
// copy engine pushes the image to the peers afterwards.
__device__ __forceinline__ void emit_grad4(const FusedParams& p, float step, size_t off, uint32_t xs, uint32_t soff, float a,
                                           float b, float c, float d, float* gdst = nullptr, long long gpush = -1) {
  if (gpush >= 0) {
    // global-batch mode, small tensors: this rank's contribution goes to slot [parity][my rank] of EVERY rank's receive
    // area (own included: the reduction then reads all contributions the same way); gpush = float offset of that slot
#pragma unroll
    for (int q = 0; q < 8; ++q)
      if (q < p.world) st_volatile_f4(p.xrecv[q] + gpush + off, a, b, c, d);
  } else if (gdst != nullptr) {  // persistent data-parallel mode: this rank's peer-visible gradient buffer
    *reinterpret_cast<float4*>(gdst + off) = make_float4(a, b, c, d);
  } else if (xs != 0u) {
    // staged: raw gradients (exchange mode) or -lr*g (single GPU: one bulk reduce-add per range applies them)
    const float gs = p.world > 1 ? 1.f : -step;
    sts128(xs + soff, gs * a, gs * b, gs * c, gs * d);
  } else if (p.wparams != nullptr) {
    // w += (-lr*g) as a fire-and-forget vector reduction: no load latency in the epilogue. Every
    // element is touched by exactly one thread, so the result equals w - lr*g bit for bit.
    asm volatile("red.global.add.v4.f32 [%0], {%1, %2, %3, %4};" ::"l"(p.wparams + off), "f"(-step * a), "f"(-step * b),
                 "f"(-step * c), "f"(-step * d) : "memory");
  } else {
    *reinterpret_cast<float4*>(p.grads + off) = make_float4(a, b, c, d);
  }
}

// Column sums over the batch rows held one-per-thread by the 128 worker threads; result (n <= 16
// columns) is valid in the first n worker threads of worker-warp 0.
template <int N>
__device__ __forceinline__ void worker_colsum(const float (&v)[N], float* red, int wwarp, int lane, float (&out)[1],
                                              int& out_col) {
#pragma unroll
  for (int c = 0; c < N; ++c) {
    float s = v[c];
    s += __shfl_xor_sync(0xffffffffu, s, 16);
    s += __shfl_xor_sync(0xffffffffu, s, 8);
    s += __shfl_xor_sync(0xffffffffu, s, 4);
    s += __shfl_xor_sync(0xffffffffu, s, 2);
    s += __shfl_xor_sync(0xffffffffu, s, 1);
    if (lane == c) red[wwarp * 32 + c] = s;
  }
  asm volatile("bar.sync 1, 128;" ::: "memory");
  out_col = -1;
  if (wwarp == 0 && lane < N) {
    out[0] = red[lane] + red[32 + lane] + red[64 + lane] + red[96 + lane];
    out_col = lane;
  }
  asm volatile("bar.sync 1, 128;" ::: "memory");
}

// Pad words of the persistent data-parallel mode (relative to kPadFusedFlags; all but the last group are only used locally)
constexpr int kPfEpoch = 64;        // steps completed (shared with the in-kernel exchange mode)
constexpr int kPfWeights = 80;      // counter: reducer CTAs that finished applying an update (red_ctas per step)
constexpr int kPfRedFlags = 128;    // [source rank][compute CTA]: epoch whose gradient slices are complete (written by every rank)
constexpr uint32_t kGxEmpty = 0xFFFFFFFFu;  // "nothing has arrived": a NaN pattern that arithmetic never produces (canonical NaN is 0x7FFFFFFF)
constexpr int kMaxRedCtas = 136;  // 8 + 136 CTAs <= 148 SMs: every CTA of the launch must be resident (the roles wait on each other)

// 16-byte asynchronous copy global -> shared that bypasses L1 (LDGSTS), and the wait for all of this thread's copies
__device__ __forceinline__ void cp_async16(uint32_t dst_s, const float* src) {
  asm volatile("cp.async.cg.shared.global [%0], [%1], 16;" ::"r"(dst_s), "l"(src) : "memory");
}
__device__ __forceinline__ void cp_async_wait_all() { asm volatile("cp.async.wait_all;" ::: "memory"); }
// Polling load for data a peer writes into THIS GPU's memory: served by L2 (where incoming NVLink writes land), never by
// L1. A plain .cg load coalesces like any other load; system-scope strong loads (ld.volatile) were measured at ~13 GB/s
// per SM here. `asm volatile` keeps the compiler from hoisting it out of the polling loop.
__device__ __forceinline__ float4 ld_poll_f4(const float* p) {
  float4 v;
  asm volatile("ld.global.cg.v4.f32 {%0, %1, %2, %3}, [%4];" : "=f"(v.x), "=f"(v.y), "=f"(v.z), "=f"(v.w) : "l"(p) : "memory");
  return v;
}
__device__ __forceinline__ float4 ld_volatile_f4(const float* p) {
  float4 v;
  asm volatile("ld.volatile.global.v4.f32 {%0, %1, %2, %3}, [%4];" : "=f"(v.x), "=f"(v.y), "=f"(v.z), "=f"(v.w) : "l"(p) : "memory");
  return v;
}

// Spin until *flag >= target (wrap-safe) with the communicator's watchdog; false on abort / deadline (status recorded).
__device__ __forceinline__ bool pf_wait(const uint32_t* flag, uint32_t target, const FusedParams& p, int peer, uint32_t phase,
                                        unsigned long long deadline) {
  uint32_t spins = 0;
  // (strict == 2: the poll itself is an acquire load - the variant measured against the relaxed default)
  while (static_cast<int32_t>((p.strict == 2 ? ld_acquire_sys(flag) : ld_relaxed_sys(flag)) - target) < 0) {
    if ((++spins & 0x3F) == 0) {
      const uint32_t ab = ld_relaxed_sys(p.xpad[p.rank] + kPadAbortWord);
      if (ab != 0u) { record_abort(p.xstatus, p.rank, ab, phase); return false; }
      if (globaltimer_ns() > deadline) {
        if (p.xstatus != nullptr) {
          p.xstatus[1] = static_cast<uint32_t>(p.rank);
          p.xstatus[2] = static_cast<uint32_t>(peer);
          p.xstatus[3] = phase;
          __threadfence_system();
          p.xstatus[0] = DMLP_ERR_TIMEOUT;
        }
        for (int q = 0; q < p.world; ++q) st_relaxed_sys(p.xpad[q] + kPadAbortWord, abort_word_for(p.rank, peer));
        return false;
      }
    }
  }
  if (p.strict == 1) __threadfence_system();  // acquire side of the conservative protocol
  return true;
}

// Global-batch mode: waits until the 16 bytes at `src` (this GPU's memory, written by a peer) hold data, i.e. no word is the
// "empty" pattern any more. Watchdog as in pf_wait; `fail` (shared) short-cuts every later wait of a failed launch.
__device__ __forceinline__ bool gx_valid(const float4& v) {
  return __float_as_uint(v.x) != kGxEmpty && __float_as_uint(v.y) != kGxEmpty && __float_as_uint(v.z) != kGxEmpty &&
         __float_as_uint(v.w) != kGxEmpty;
}
__device__ __forceinline__ bool gx_timeout(const FusedParams& p, int peer, uint32_t phase, unsigned long long deadline) {
  const uint32_t ab = ld_relaxed_sys(p.xpad[p.rank] + kPadAbortWord);
  if (ab != 0u) { record_abort(p.xstatus, p.rank, ab, phase); return true; }
  if (globaltimer_ns() > deadline) {
    if (p.xstatus != nullptr) {
      p.xstatus[1] = static_cast<uint32_t>(p.rank);
      p.xstatus[2] = static_cast<uint32_t>(peer);
      p.xstatus[3] = phase;
      __threadfence_system();
      p.xstatus[0] = DMLP_ERR_TIMEOUT;
    }
    for (int q = 0; q < p.world; ++q) st_relaxed_sys(p.xpad[q] + kPadAbortWord, abort_word_for(p.rank, peer));
    return true;
  }
  return false;
}
__device__ __forceinline__ float4 gx_wait4(const float* src, const FusedParams& p, int peer, uint32_t phase,
                                           unsigned long long deadline, volatile int* fail) {
  float4 v = ld_volatile_f4(src);
  if (*fail != 0) return v;
  uint32_t spins = 0;
  while (!gx_valid(v)) {
    if ((++spins & 0x3F) == 0 && (*fail != 0 || gx_timeout(p, peer, phase, deadline))) { *fail = 1; break; }
    v = ld_volatile_f4(src);
  }
  return v;
}

// Reducer CTA of the persistent data-parallel mode: one-shot all-reduce + average + SGD per step, fed by local flags.
__device__ void fused_reducer_loop(const FusedParams& p) {
  __shared__ int s_fail;
  __shared__ uint32_t s_epoch0;
  const int b = static_cast<int>(blockIdx.x) - kFC, nb = p.red_ctas;
  uint32_t* mypad = p.xpad[p.rank] + kPadFusedFlags;
  asm volatile("griddepcontrol.launch_dependents;" ::: "memory");
  asm volatile("griddepcontrol.wait;" ::: "memory");
  if (threadIdx.x == 0) { s_fail = 0; s_epoch0 = __ldcg(mypad + kPfEpoch); }
  __syncthreads();
  const uint32_t epoch0 = s_epoch0;
  const int nsteps = p.nsteps > 0 ? p.nsteps : 1;
  const size_t npacks = p.total / 4;  // the flat parameter vector is padded to 64-float multiples
  for (int st = 0; st < nsteps; ++st) {
    const uint32_t epoch = epoch0 + static_cast<uint32_t>(st) + 1u;
    const uint32_t par = epoch & 1u;
    const unsigned long long deadline = globaltimer_ns() + p.xtimeout_ns;
    // fault injection: a "dead" rank neither reduces nor waits (its own compute cluster is still released below)
    if (p.fault_step >= 0 && static_cast<long long>(p.g0) + st >= p.fault_step) {
      __syncthreads();
      if (threadIdx.x == 0) { s_fail = 1; atomicAdd(mypad + kPfWeights, 1u); }
      continue;
    }
    // every rank's compute CTAs raise [rank][cta] in every rank's pad once their gradient slices are complete:
    // (world x 8) flags to poll locally, no relay through this rank's own flags
    if (threadIdx.x < p.world * kFC) {
      if (!pf_wait(mypad + kPfRedFlags + threadIdx.x, epoch, p, threadIdx.x / kFC, 7u, deadline)) s_fail = 1;
    }
    __syncthreads();
    if (s_fail == 0) {
      const float sc = __ldcg(p.lr) / static_cast<float>(p.world);
      // kRedPk packs per thread and pass, every load of a pass issued before the first add: one NVLink round trip per pass
      constexpr int kRedPk = 4;
      const size_t stride = static_cast<size_t>(nb) * blockDim.x;
      for (size_t i0 = static_cast<size_t>(b) * blockDim.x + threadIdx.x; i0 < npacks; i0 += stride * kRedPk) {
        float4 g[kRedPk][8];
        float4 w[kRedPk];
#pragma unroll
        for (int k = 0; k < kRedPk; ++k) {
          const size_t i = i0 + k * stride;
          if (i < npacks) {
#pragma unroll
            for (int r = 0; r < 8; ++r)
              if (r < p.world) g[k][r] = ld_volatile_f4(p.xgrads[par][r] + 4 * i);  // rank order: identical sums everywhere
            w[k] = __ldcg(reinterpret_cast<const float4*>(p.wparams + 4 * i));
          }
        }
#pragma unroll
        for (int k = 0; k < kRedPk; ++k) {
          const size_t i = i0 + k * stride;
          if (i < npacks) {
            float4 acc = g[k][0];
#pragma unroll
            for (int r = 1; r < 8; ++r)
              if (r < p.world) { acc.x += g[k][r].x; acc.y += g[k][r].y; acc.z += g[k][r].z; acc.w += g[k][r].w; }
            w[k].x -= sc * acc.x; w[k].y -= sc * acc.y; w[k].z -= sc * acc.z; w[k].w -= sc * acc.w;
            *reinterpret_cast<float4*>(p.wparams + 4 * i) = w[k];
          }
        }
      }
    }
    if (p.strict == 1) __threadfence_system(); else
    __threadfence();
    __syncthreads();
    if (threadIdx.x == 0) atomicAdd(mypad + kPfWeights, 1u);  // (also on failure: the compute cluster must not hang)
  }
}

// S1, S2: slice widths H1/8 and H2/8 (S2 == 0: two-layer model). GX: compiled with the global-batch data-parallel mode
// (a separate instantiation, so the single-GPU kernel carries none of its registers or code).
template <int S1, int S2, bool GX>
__global__ void __cluster_dims__(kFC, 1, 1) __launch_bounds__(kFThreads, 1)
mlp_fused_step_kernel(const __grid_constant__ CUtensorMap tmXA, const __grid_constant__ CUtensorMap tmXT,
                      const __grid_constant__ CUtensorMap tmW1, const __grid_constant__ CUtensorMap tmW2B,
                      const __grid_constant__ CUtensorMap tmW2K, const __grid_constant__ CUtensorMap tmWLB,
                      const __grid_constant__ CUtensorMap tmWLK, const FusedParams p) {
  constexpr int H1 = S1 * kFC;
  constexpr int H2 = S2 * kFC;
  constexpr bool kMid = S2 > 0;
  constexpr int SL = kMid ? S2 : S1;  // K-slice of the last layer
  constexpr int LAST = kMid ? 2 : 1;
  static_assert(S1 == 16 && (S2 == 8 || S2 == 0), "slice widths this kernel was laid out for");
  extern __shared__ uint8_t smem_raw[];
  // 1024-byte alignment by pointer arithmetic on the shared array itself (a round trip through an integer makes
  // the compiler fall back to generic LD/ST for everything addressed off this base)
  uint8_t* smem = smem_raw + ((1024u - (smem_u32(smem_raw) & 1023u)) & 1023u);
  const uint32_t sb = smem_u32(smem);
  // bar_mma: the critical chain F1 -> F2 -> F3 -> dgrad_last -> dgrad2 -> wgrad1 (phase toggles per use);
  // bar_wgl / bar_wg2: weight-gradient GEMMs that run beside the chain; bar_ex[k]: arrival of exchange k
  // (remote st.async stores complete transaction bytes on the DESTINATION's barrier: no cluster-wide barrier).
  __shared__ __align__(8) uint64_t bar_tma0, bar_tma1, bar_mma, bar_wgl, bar_wg2, bar_ex[5], bar_xr[4];
  // global-batch mode, two peer stages: X^T tile landed (TMA) / stage complete (128 workers) / stage consumed (MMA commit)
  __shared__ __align__(8) uint64_t bar_gx[2], bar_gf[2], bar_ge[2];
  __shared__ uint32_t tmem_slot;

  // Warp index through a shuffle: the role branches are then provably warp-uniform and the control warp's
  // descriptors live in uniform registers (a `threadIdx.x == 0` branch wraps every tcgen05.mma / TMA issue in a
  // per-lane R2UR loop: ~170 instead of ~40 cycles per instruction, profiles/r1_mma_issue_probe.md).
  // Control work is done by warp 0: all lanes wait, one elected lane issues.
  const int warp = __shfl_sync(0xffffffffu, threadIdx.x >> 5, 0), lane = threadIdx.x & 31;
  const int j = static_cast<int>(cluster_ctarank());
  const int B = p.B;
  const int ngroups = p.in0 / 8;
  const int g0 = (ngroups * j) / kFC, g1 = (ngroups * (j + 1)) / kFC;
  const int k0 = g0 * 8, nk1 = g1 - g0;  // layer-1 K slice: floats [k0, k0 + 8*nk1)
  if (!GX && p.red_ctas > 0 && blockIdx.x >= kFC) {  // persistent data-parallel mode: everything past the first cluster reduces
    fused_reducer_loop(p);
    return;
  }
  const bool pmode = !GX && p.red_ctas > 0;  // (the global-batch instantiation carries none of the other exchange modes)
  const bool worker = warp >= 1;
  const int wwarp = warp - 1;
  const int row = (warp & 3) * 32 + lane;  // TMEM lane == batch row handled by this worker thread
  const bool rok = worker && row < B;
  float* red = reinterpret_cast<float*>(smem + oRED);
  const uint32_t tq = (static_cast<uint32_t>((warp & 3) * 32) << 16);

  // arm the exchange barriers: bytes every CTA of the cluster (this one included) will deliver in one step
  auto arm_exchanges = [&]() {
    mbar_arrive_expect_tx(&bar_ex[0], kFC * B * S1 * 4);
    mbar_arrive_expect_tx(&bar_ex[2], kFC * B * 64);
    mbar_arrive_expect_tx(&bar_ex[4], kFC * B * S1 * 4);
    if (kMid) {
      mbar_arrive_expect_tx(&bar_ex[1], kFC * B * S2 * 4);
      mbar_arrive_expect_tx(&bar_ex[3], kFC * B * S2 * 4 * 2);
    }
  };
  if (threadIdx.x == 0) {
    mbar_init(&bar_tma0, 1);
    mbar_init(&bar_tma1, 1);
    mbar_init(&bar_mma, 1);
    mbar_init(&bar_wgl, 1);
    mbar_init(&bar_wg2, 1);
    for (int k = 0; k < 5; ++k) mbar_init(&bar_ex[k], 1);
    for (int k = 0; k < 4; ++k) mbar_init(&bar_xr[k], 1);
    for (int k = 0; k < 2; ++k) { mbar_init(&bar_gx[k], 1); mbar_init(&bar_gf[k], 96); mbar_init(&bar_ge[k], 1); }
    fence_mbar_init();
    arm_exchanges();
  }
  if (warp == 0) {
    tmem_alloc(&tmem_slot, 512);
    tmem_relinquish();
  }
  tc_fence_before_sync();
  __syncthreads();
  tc_fence_after_sync();
  const uint32_t tm = __shfl_sync(0xffffffffu, tmem_slot, 0);
  uint32_t mma_phase = 0;
  FPROF(0);
  cluster_sync_all();  // every CTA's shared memory and barriers are live before anyone writes into them
  FPROF(1);

  // Programmatic dependent launch: this grid may start while its predecessor in the stream (the previous
  // step's kernel, or its all-reduce) is still finishing. Everything so far — and the input loads
  // just below — touch nothing the predecessor writes; parameters and gradients are only accessed
  // after griddepcontrol.wait.
  asm volatile("griddepcontrol.launch_dependents;" ::: "memory");
  // ================================================================== step loop
  // One launch can run several optimisation steps back to back (single-GPU mode, SGD applied in place): the
  // launch gap, TMEM allocation and barrier set-up are paid once; steps are separated by one cluster barrier
  // that also publishes the weight updates (red.global + __threadfence) to the next step's loads.
  const int nsteps = p.nsteps > 0 ? p.nsteps : 1;
  constexpr bool gmode = GX;  // (launch_fused<.., true> is only reached through plans with gx set and world > 1)
  const bool xmode = !GX && p.world > 1 && !pmode;
  const int W = p.world > 1 ? p.world : 1;
  const int npeer = gmode ? W - 1 : 0;
  uint32_t guse0 = 0, guse1 = 0;  // global-batch mode: uses of peer stage 0 / 1 so far (barrier parity = use & 1)
  uint32_t gmma0 = 0, gmma1 = 0;  // ... as counted by the MMA issuer (warp 0 fills ahead of its own consumption)
  uint32_t xr_phase = 0;  // bit s: parity the next wait on receive stage s expects
  uint32_t epoch0 = 0;  // steps this rank completed before this launch (every rank holds the same count)
  __shared__ int s_xfail;
  if (threadIdx.x == 0) s_xfail = 0;
  volatile int* xfail = &s_xfail;
  for (int st = 0; st < nsteps; ++st) {
  const uint32_t sp = st & 1;  // parity of the barriers that complete once per step
  FPROF(13);
  if (warp == 0 && elect_one()) {
    const int xr = ((gmode ? p.rank * p.xslots : 0) + p.slot0 + st) * B;
    mbar_arrive_expect_tx(&bar_tma0, 4 * kTileK + (H1 / 32) * kW1Col);
    for (int b = 0; b < 4; ++b) tma_load_2d(smem + oXA + b * kTileK, &tmXA, &bar_tma0, k0 + 32 * b, xr);
    uint32_t bytes1 = 4 * kColM + 2 * SL * 128;
    if (kMid) bytes1 += (H2 / 32) * S1 * 128 + (H2 / 32) * S1 * 128;
    mbar_arrive_expect_tx(&bar_tma1, bytes1);
    for (int a = 0; a < 4; ++a) tma_load_2d(smem + oXT + a * kColM, &tmXT, &bar_tma1, k0 + 32 * a, xr);
    if (st + 1 < nsteps)  // pull the next step's inputs from HBM into L2 while this step computes
      for (int b = 0; b < 4; ++b) tma_prefetch_2d(&tmXA, k0 + 32 * b, xr + B);
  }
  if (st == 0) {
    asm volatile("griddepcontrol.wait;" ::: "memory");
    if (xmode || pmode) epoch0 = __ldcg(p.xpad[p.rank] + kPadFusedEpoch);
    if (gmode) epoch0 = static_cast<uint32_t>(p.g0);
  }
  const uint32_t epoch = epoch0 + static_cast<uint32_t>(st) + 1u;
  // fault injection: a "dead" rank sends nothing and waits for nothing from its fault step on (its peers' watchdogs fire)
  const bool gdead = gmode && p.fault_step >= 0 && static_cast<long long>(p.g0) + st >= p.fault_step;
  const uint32_t xpar = epoch & 1u;  // receive areas are double-buffered by step parity
  // gradients are staged in shared memory in exchange mode and (single GPU) when the update is applied by bulk reduce-adds
  const bool bmode = !GX && !xmode && !pmode && p.wparams != nullptr && p.bulk_update != 0;
  const uint32_t xs = (xmode || bmode) ? sb : 0u;
  float* gdst = pmode ? p.xgrads[xpar][p.rank] : nullptr;
  if (gdead && threadIdx.x == 0) s_xfail = 1;  // (published by the barriers that precede the first cross-GPU wait)
  if (pmode && st > 0) {
    // the reducers have applied the previous step's update (red_ctas increments per step) before any weight is loaded
    if (threadIdx.x == 0) {
      const unsigned long long deadline = globaltimer_ns() + p.xtimeout_ns;
      if (!pf_wait(p.xpad[p.rank] + kPadFusedFlags + kPfWeights, (epoch - 1u) * static_cast<uint32_t>(p.red_ctas), p, p.rank, 8u, deadline))
        s_xfail = 1;
      fence_proxy_async_all();  // the weights the reducers wrote are re-read by TMA (async proxy)
    }
    __syncthreads();
  }
  if (warp == 0) {
    if (elect_one()) {
    for (int a = 0; a < H1 / 32; ++a) tma_load_2d(smem + oW1B + a * kW1Col, &tmW1, &bar_tma0, 32 * a, k0);
    if (kMid) {
      for (int a = 0; a < H2 / 32; ++a) tma_load_2d(smem + oW2B + a * S1 * 128, &tmW2B, &bar_tma1, 32 * a, S1 * j);
      for (int kb = 0; kb < H2 / 32; ++kb) tma_load_2d(smem + oW2K + kb * 4096, &tmW2K, &bar_tma1, 32 * kb, S1 * j);
    }
    tma_load_2d(smem + oWLB, &tmWLB, &bar_tma1, 0, SL * j);
    tma_load_2d(smem + oWLK, &tmWLK, &bar_tma1, 0, SL * j);
    }
    __syncwarp();
    // ---------------------------------------------------------------- F1: P1 = X_j . W1_j
    mbar_wait(&bar_tma0, sp);
    tc_fence_after_sync();
    if (elect_one()) {
      const uint32_t idesc = make_instr_desc(4, false, true, H1);
#pragma unroll
      for (int t = 0; t < kKGroupsMax; ++t)
        if (t < nk1) tc_mma<4>(tm + cF1, kdesc(sb + oXA, t), mdesc(sb + oW1B, t, kW1Col), idesc, t != 0);
      tc_commit(&bar_mma);
    }
    __syncwarp();
  }
  // (after griddepcontrol.wait) global-batch mode: the in-place update applies the gradient of the global batch, lr / world
  const float step = p.wparams != nullptr ? __ldcg(p.lr) / static_cast<float>(gmode ? W : 1) : 0.f;
  // slot [parity][my rank] of a receive area, in floats (the small tensors keep their offsets of the flat parameter vector,
  // the dZ1 image sits where W1 would be)
  const long long gslot = gmode ? static_cast<long long>((static_cast<size_t>(xpar) * W + p.rank) * p.total) : -1;
  const long long gpush = (gmode && !gdead) ? gslot : -1;
  const unsigned long long gdeadline = gmode ? globaltimer_ns() + p.xtimeout_ns : 0ull;
  if (gmode && worker) {
    // The dZ1 images of the other parity were consumed by all 8 CTAs in the previous step (its end-of-step barrier is
    // behind us) and are written again by the peers at the end of the NEXT step: mark them empty. Each CTA clears an eighth.
    const float4 e4 = make_float4(__uint_as_float(kGxEmpty), __uint_as_float(kGxEmpty), __uint_as_float(kGxEmpty), __uint_as_float(kGxEmpty));
    for (int q = 0; q < W; ++q) {
      if (q == p.rank) continue;
      float* img = p.xrecv[p.rank] + (static_cast<size_t>(xpar ^ 1u) * W + q) * p.total + p.w_off[0];
      const int wt = threadIdx.x - 32;
      *reinterpret_cast<float4*>(img + 4 * (j * 256 + wt)) = e4;
      *reinterpret_cast<float4*>(img + 4 * (j * 256 + 128 + wt)) = e4;
    }
  }
  // Biases of this CTA's slices and this row's label: requested now, consumed phases later (a cold global load
  // on the critical chain costs ~0.6 us each time).
  float b1r[S1], b2r[kMid ? S2 : 1], blr[16];
  int ylab = 0;
  if (rok) {
    ylab = __ldg(p.labels + st * B + row);
#pragma unroll
    for (int c = 0; c < S1; c += 4) {
      const float4 v = __ldcg(reinterpret_cast<const float4*>(p.params + p.b_off[0] + S1 * j + c));
      b1r[c] = v.x; b1r[c + 1] = v.y; b1r[c + 2] = v.z; b1r[c + 3] = v.w;
    }
    if (kMid) {
#pragma unroll
      for (int c = 0; c < S2; c += 4) {
        const float4 v = __ldcg(reinterpret_cast<const float4*>(p.params + p.b_off[1] + S2 * j + c));
        b2r[c] = v.x; b2r[c + 1] = v.y; b2r[c + 2] = v.z; b2r[c + 3] = v.w;
      }
    }
#pragma unroll
    for (int c = 0; c < 16; c += 4) {
      const float4 v = __ldcg(reinterpret_cast<const float4*>(p.params + p.b_off[LAST] + c));
      blr[c] = v.x; blr[c + 1] = v.y; blr[c + 2] = v.z; blr[c + 3] = v.w;
    }
  }
  float h1[S1];
  float h2[kMid ? S2 : 1];
  if (worker) {
    mbar_wait(&bar_mma, mma_phase);
    tc_fence_after_sync();
    // reduce-scatter: columns [S1*i, S1*i+S1) of my partial go to CTA i
#pragma unroll 1
    for (int i = 0; i < kFC; ++i) {
      uint32_t r[16];
      tmem_ld_x16(tm + tq + cF1 + S1 * i, r);
      tmem_ld_wait();
      if (rok) {
        const uint32_t dst = mapa(sb + oRS1 + (j * 64 + row) * (S1 * 4), i);
        const uint32_t bar = mapa(smem_u32(&bar_ex[0]), i);
#pragma unroll
        for (int c = 0; c < S1; c += 4)
          st_async_f4(dst + c * 4, bar, __uint_as_float(r[c]), __uint_as_float(r[c + 1]), __uint_as_float(r[c + 2]),
                      __uint_as_float(r[c + 3]));
      }
    }
  }
  mma_phase ^= 1;
  FPROF(2);

  // ------------------------------------------------------------------ P1: H1_j, layer-2 partial
  if (worker) {
    mbar_wait(&bar_ex[0], sp);
    FPROF(3);
#pragma unroll
    for (int c = 0; c < S1; ++c) h1[c] = 0.f;
    if (rok) {
      #pragma unroll 1
      for (int s = 0; s < kFC; ++s) {  // fixed order: deterministic sums
        const float4* src = reinterpret_cast<const float4*>(smem + oRS1 + (s * 64 + row) * (S1 * 4));
#pragma unroll
        for (int c = 0; c < S1; c += 4) {
          const float4 v = src[c >> 2];
          h1[c] += v.x; h1[c + 1] += v.y; h1[c + 2] += v.z; h1[c + 3] += v.w;
        }
      }
#pragma unroll
      for (int c = 0; c < S1; ++c) h1[c] = fmaxf(h1[c] + b1r[c], 0.f);
#pragma unroll
      for (int c = 0; c < S1; c += 4) {
        *reinterpret_cast<float4*>(smem + oH1K + koff(row, c)) = make_float4(h1[c], h1[c + 1], h1[c + 2], h1[c + 3]);
        *reinterpret_cast<float4*>(smem + oH1M + moff(row, c)) = make_float4(h1[c], h1[c + 1], h1[c + 2], h1[c + 3]);
      }
      // "ones" column: row S1 of H1_j^T . dZ is then the column sum of dZ = a bias gradient, for free
      *reinterpret_cast<float*>(smem + oH1M + moff(row, S1)) = 1.f;
      if (!GX && p.dbg)
        for (int c = 0; c < S1; ++c) p.dbg[row * H1 + S1 * j + c] = h1[c];
    }
    // X^T tile: column 127 lies outside every CTA's K slice (<= 104 wide) -> make it the ones column
    mbar_wait(&bar_tma1, sp);
    if (rok) *reinterpret_cast<float*>(smem + oXT + moff(row, 127)) = 1.f;
    fence_proxy_async_smem();
  }
  __syncthreads();
  if (kMid) {
    if (warp == 0) {
      mbar_wait(&bar_tma1, sp);
      fence_proxy_async_smem();
      tc_fence_after_sync();
      if (elect_one()) {
        const uint32_t idesc = make_instr_desc(4, false, true, H2);
#pragma unroll
        for (int t = 0; t < S1 / 8; ++t)
          tc_mma<4>(tm + cF2, kdesc(sb + oH1K, t), mdesc(sb + oW2B, t, S1 * 128), idesc, t != 0);
        tc_commit(&bar_mma);
      }
      __syncwarp();
    }
    if (worker) {
      mbar_wait(&bar_mma, mma_phase);
      tc_fence_after_sync();
#pragma unroll 1
      for (int i = 0; i < kFC; ++i) {
        uint32_t r[8];
        tmem_ld_x8(tm + tq + cF2 + S2 * i, r);
        tmem_ld_wait();
        if (rok) {
          const uint32_t dst = mapa(sb + oRS2 + (j * 64 + row) * (S2 * 4), i);
          const uint32_t bar = mapa(smem_u32(&bar_ex[1]), i);
          st_async_f4(dst, bar, __uint_as_float(r[0]), __uint_as_float(r[1]), __uint_as_float(r[2]), __uint_as_float(r[3]));
          st_async_f4(dst + 16, bar, __uint_as_float(r[4]), __uint_as_float(r[5]), __uint_as_float(r[6]), __uint_as_float(r[7]));
        }
      }
    }
    mma_phase ^= 1;
    FPROF(4);
    // ---------------------------------------------------------------- P2: H2_j
    if (worker) {
      mbar_wait(&bar_ex[1], sp);
      FPROF(5);
#pragma unroll
      for (int c = 0; c < S2; ++c) h2[c] = 0.f;
      if (rok) {
        #pragma unroll 1
        #pragma unroll 1
      for (int s = 0; s < kFC; ++s) {
          const float4* src = reinterpret_cast<const float4*>(smem + oRS2 + (s * 64 + row) * (S2 * 4));
#pragma unroll
          for (int c = 0; c < S2; c += 4) {
            const float4 v = src[c >> 2];
            h2[c] += v.x; h2[c + 1] += v.y; h2[c + 2] += v.z; h2[c + 3] += v.w;
          }
        }
#pragma unroll
        for (int c = 0; c < S2; ++c) h2[c] = fmaxf(h2[c] + b2r[kMid ? c : 0], 0.f);
#pragma unroll
        for (int c = 0; c < S2; c += 4) {
          *reinterpret_cast<float4*>(smem + oH2K + koff(row, c)) = make_float4(h2[c], h2[c + 1], h2[c + 2], h2[c + 3]);
          *reinterpret_cast<float4*>(smem + oH2M + moff(row, c)) = make_float4(h2[c], h2[c + 1], h2[c + 2], h2[c + 3]);
        }
        *reinterpret_cast<float*>(smem + oH2M + moff(row, S2)) = 1.f;
        if (!GX && p.dbg)
          for (int c = 0; c < S2; ++c) p.dbg[64 * H1 + row * H2 + S2 * j + c] = h2[c];
      }
      fence_proxy_async_smem();
    }
    __syncthreads();
  }
  // ------------------------------------------------------------------ last layer partial -> all-reduce
  const uint32_t oHLK = kMid ? oH2K : oH1K;
  const uint32_t oDZ1 = kMid ? oDZ1M : oDZ1M2;
  const uint32_t oHLM = kMid ? oH2M : oH1M;
  if (warp == 0) {
    if (!kMid) mbar_wait(&bar_tma1, sp);
    fence_proxy_async_smem();
    tc_fence_after_sync();
    if (elect_one()) {
      const uint32_t idesc = make_instr_desc(4, false, true, 16);
#pragma unroll
      for (int t = 0; t < SL / 8; ++t)
        tc_mma<4>(tm + cF3, kdesc(sb + oHLK, t), mdesc(sb + oWLB, t, SL * 128), idesc, t != 0);
      tc_commit(&bar_mma);
    }
    __syncwarp();
  }
  if (worker) {
    mbar_wait(&bar_mma, mma_phase);
    tc_fence_after_sync();
    uint32_t r[16];
    tmem_ld_x16(tm + tq + cF3, r);
    tmem_ld_wait();
    if (rok) {
#pragma unroll 1
      for (int i = 0; i < kFC; ++i) {
        const uint32_t dst = mapa(sb + oAR3 + (j * 64 + row) * 64, i);
        const uint32_t bar = mapa(smem_u32(&bar_ex[2]), i);
#pragma unroll
        for (int c = 0; c < 16; c += 4)
          st_async_f4(dst + c * 4, bar, __uint_as_float(r[c]), __uint_as_float(r[c + 1]), __uint_as_float(r[c + 2]),
                      __uint_as_float(r[c + 3]));
      }
    }
  }
  mma_phase ^= 1;
  FPROF(6);

  // ------------------------------------------------------------------ P3: softmax-CE, last-layer grads
  float dz3[16];
  if (worker) {
    mbar_wait(&bar_ex[2], sp);
    FPROF(7);
    float z[16];
#pragma unroll
    for (int c = 0; c < 16; ++c) z[c] = 0.f;
    float loss = 0.f;
    int hit = 0;
    if (rok) {
      #pragma unroll 1
      for (int s = 0; s < kFC; ++s) {
        const float4* src = reinterpret_cast<const float4*>(smem + oAR3 + (s * 64 + row) * 64);
#pragma unroll
        for (int c = 0; c < 16; c += 4) {
          const float4 v = src[c >> 2];
          z[c] += v.x; z[c + 1] += v.y; z[c + 2] += v.z; z[c + 3] += v.w;
        }
      }
      float mx = -INFINITY;
      int arg = 0;
#pragma unroll
      for (int c = 0; c < 16; ++c) {
        z[c] += blr[c];
        if (c < p.C && z[c] > mx) { mx = z[c]; arg = c; }
      }
      float sum = 0.f;
#pragma unroll
      for (int c = 0; c < 16; ++c) {
        z[c] = (c < p.C) ? __expf(z[c] - mx) : 0.f;
        sum += z[c];
      }
      const float inv = 1.f / sum;
      const int y = ylab;
      const float scale = 1.f / static_cast<float>(B);
#pragma unroll
      for (int c = 0; c < 16; ++c) {
        const float pr = z[c] * inv;
        if (c == y) loss = -__logf(pr + 1e-10f);
        dz3[c] = (c < p.C) ? (pr - (c == y ? 1.f : 0.f)) * scale : 0.f;
      }
      hit = (arg == y) ? 1 : 0;
#pragma unroll
      for (int c = 0; c < 16; c += 4) {
        *reinterpret_cast<float4*>(smem + oDZ3K + koff(row, c)) = make_float4(dz3[c], dz3[c + 1], dz3[c + 2], dz3[c + 3]);
        *reinterpret_cast<float4*>(smem + oDZ3M + moff(row, c)) = make_float4(dz3[c], dz3[c + 1], dz3[c + 2], dz3[c + 3]);
      }
      if (!GX && p.dbg && j == 0)
        for (int c = 0; c < 16; ++c) p.dbg[64 * H1 + 64 * (kMid ? H2 : 0) + row * 16 + c] = dz3[c];
    } else {
#pragma unroll
      for (int c = 0; c < 16; ++c) dz3[c] = 0.f;
    }
    fence_proxy_async_smem();
    if (j == 0) {  // loss and hits are produced once (CTA 0)
#pragma unroll
      for (int o = 16; o > 0; o >>= 1) {
        loss += __shfl_xor_sync(0xffffffffu, loss, o);
        hit += __shfl_xor_sync(0xffffffffu, hit, o);
      }
      if (lane == 0) { red[wwarp] = loss; reinterpret_cast<int*>(red)[8 + wwarp] = hit; }
    }
  }
  __syncthreads();
  const int nb8 = B / 8;  // K steps of the weight-gradient GEMMs (K = batch)
  if (warp == 0) {
    fence_proxy_async_smem();
    tc_fence_after_sync();
    if (elect_one()) {
      // critical: dHL_j[B x SL] = dZ3 . W_last_j^T  (A and B K-major, K = 16)
      const uint32_t idg = make_instr_desc(4, false, false, 16);
#pragma unroll
      for (int t = 0; t < 2; ++t) tc_mma<4>(tm + cDGL, kdesc(sb + oDZ3K, t), kdesc(sb + oWLK, t), idg, t != 0);
      tc_commit(&bar_mma);
      // beside the chain: dW_last[SL x 16] = HL_j^T . dZ3   (A and B MN-major, K = batch)
      const uint32_t idw = make_instr_desc(4, true, true, 16);
#pragma unroll
      for (int t = 0; t < 8; ++t)
        if (t < nb8) tc_mma<4>(tm + cWGL, mdesc(sb + oHLM, t, kColM), mdesc(sb + oDZ3M, t, kColM), idw, t != 0);
      tc_commit(&bar_wgl);
    }
    __syncwarp();
  }
  if (worker && j == 0 && wwarp == 0 && lane == 0) {  // (red[] was filled before the __syncthreads above)
    p.metrics[2 * st] = red[0] + red[1] + red[2] + red[3];
    reinterpret_cast<int*>(p.metrics)[2 * st + 1] = reinterpret_cast<int*>(red)[8] + reinterpret_cast<int*>(red)[9] +
                                           reinterpret_cast<int*>(red)[10] + reinterpret_cast<int*>(red)[11];
  }
  float dzl[SL];
  if (worker) {
    mbar_wait(&bar_mma, mma_phase);
    tc_fence_after_sync();
    uint32_t r[16];
    tmem_ld_x16(tm + tq + cDGL, r);
    tmem_ld_wait();
#pragma unroll
    for (int c = 0; c < SL; ++c) {
      const float hl = kMid ? h2[kMid ? c : 0] : h1[kMid ? 0 : c];
      dzl[c] = (rok && hl > 0.f) ? __uint_as_float(r[c]) : 0.f;
    }
  }
  mma_phase ^= 1;

  float dz1[S1];
  if (kMid) {
    // all-gather dZ2_j (= dzl) into every CTA: K-major (A of dgrad2) and MN-major (B of wgrad2)
    if (rok) {
#pragma unroll 1
      for (int i = 0; i < kFC; ++i) {
        const uint32_t bar = mapa(smem_u32(&bar_ex[3]), i);
        const uint32_t dm = mapa(sb + oDZ2M + moff(row, S2 * j), i);
        st_async_f4(mapa(sb + oDZ2K + koff(row, S2 * j), i), bar, dzl[0], dzl[1], dzl[2], dzl[3]);
        st_async_f4(mapa(sb + oDZ2K + koff(row, S2 * j + 4), i), bar, dzl[4], dzl[5], dzl[6], dzl[7]);
        st_async_f4(dm, bar, dzl[0], dzl[1], dzl[2], dzl[3]);
        st_async_f4(dm + 16, bar, dzl[4], dzl[5], dzl[6], dzl[7]);
      }
      if (!GX && p.dbg)
        for (int c = 0; c < S2; ++c) p.dbg[64 * H1 + 64 * H2 + 64 * 16 + row * H2 + S2 * j + c] = dzl[c];
    }
    FPROF(8);
    // ---------------------------------------------------------------- P4: layer-2 grads
    if (warp == 0) {
      mbar_wait(&bar_ex[3], sp);
      fence_proxy_async_smem();
      tc_fence_after_sync();
      if (elect_one()) {
        const uint32_t idg = make_instr_desc(4, false, false, 16);
#pragma unroll
        for (int t = 0; t < H2 / 8; ++t)
          tc_mma<4>(tm + cDG2, kdesc(sb + oDZ2K, t), make_smem_desc(sb + oW2K + (t >> 2) * 4096 + (t & 3) * 32, 16, 1024, 2), idg, t != 0);
        tc_commit(&bar_mma);
        const uint32_t idw = make_instr_desc(4, true, true, H2);
#pragma unroll
        for (int t = 0; t < 8; ++t)
          if (t < nb8) tc_mma<4>(tm + cWG2, mdesc(sb + oH1M, t, kColM), mdesc(sb + oDZ2M, t, kColM), idw, t != 0);
        tc_commit(&bar_wg2);
      }
      __syncwarp();
    }
    if (worker) {
      mbar_wait(&bar_mma, mma_phase);
      FPROF(9);
      tc_fence_after_sync();
      uint32_t r[16];
      tmem_ld_x16(tm + tq + cDG2, r);
      tmem_ld_wait();
#pragma unroll
      for (int c = 0; c < S1; ++c) dz1[c] = (rok && h1[c] > 0.f) ? __uint_as_float(r[c]) : 0.f;
    }
    mma_phase ^= 1;
  } else {
#pragma unroll
    for (int c = 0; c < S1; ++c) dz1[c] = dzl[kMid ? 0 : c];
  }
  // global-batch mode: the slice goes to every peer first (longest way to travel), as block j of the peers' dZ1 receive
  // image: [8 source CTAs][64 rows][16 columns], 4 KiB contiguous per CTA. Rows of a warp are transposed through shared
  // memory so that each store instruction covers 512 contiguous bytes (see oSTG).
  FPROFT(32, 16);
  FPROFT(128, 32);
  if (gmode && worker && (warp & 3) < 2) {  // the two warps that hold batch rows 0..63
    const uint32_t stg = sb + oSTG + (warp & 3) * 2048;
    if (elect_one()) bulk_wait_read();  // (the previous step's copies out of this warp's staging half)
    __syncwarp();
    if (rok) {
      sts128(stg + lane * 64, dz1[0], dz1[1], dz1[2], dz1[3]);
      sts128(stg + lane * 64 + 16, dz1[4], dz1[5], dz1[6], dz1[7]);
      sts128(stg + lane * 64 + 32, dz1[8], dz1[9], dz1[10], dz1[11]);
      sts128(stg + lane * 64 + 48, dz1[12], dz1[13], dz1[14], dz1[15]);
    }
    fence_proxy_async_smem();  // staged with ordinary stores, read by the bulk copy engine
    __syncwarp();
    if (gpush >= 0 && elect_one()) {
      // one bulk copy per peer: the copy engine moves it, this SM's load/store path stays free (SM-issued stores to a
      // peer stall the issuing warp for about a microsecond)
      const int rows = B - (warp & 3) * 32 < 32 ? B - (warp & 3) * 32 : 32;
      if (rows > 0) {
        const size_t o = static_cast<size_t>(gslot) + p.w_off[0] + j * 1024 + (warp & 3) * 512;
        #pragma unroll 1
        for (int q = 0; q < W; ++q)
          if (q != p.rank) bulk_s2g(p.xrecv[q] + o, stg, rows * 64);
        bulk_commit();
      }
    }
  }
  FPROFT(32, 17);
  FPROFT(128, 33);
  // all-gather dZ1_j into every CTA (MN-major: B operand of the layer-1 weight gradient)
  if (rok) {
#pragma unroll 1
    for (int i = 0; i < kFC; ++i) {
      const uint32_t bar = mapa(smem_u32(&bar_ex[4]), i);
      const uint32_t dm = mapa(sb + oDZ1 + moff(row, S1 * j), i);
      st_async_f4(dm, bar, dz1[0], dz1[1], dz1[2], dz1[3]);
      st_async_f4(dm + 16, bar, dz1[4], dz1[5], dz1[6], dz1[7]);
      const uint32_t dm2 = mapa(sb + oDZ1 + moff(row, S1 * j + 8), i);
      st_async_f4(dm2, bar, dz1[8], dz1[9], dz1[10], dz1[11]);
      st_async_f4(dm2 + 16, bar, dz1[12], dz1[13], dz1[14], dz1[15]);
    }
    if (!GX && p.dbg)
      for (int c = 0; c < S1; ++c) p.dbg[64 * H1 + 128 * (kMid ? H2 : 0) + 64 * 16 + row * H1 + S1 * j + c] = dz1[c];
  }
  FPROF(10);

  // ------------------------------------------------------------------ P5: dW1_j = X_j^T . dZ1
  if (warp == 0 && gmode) {
    // Global batch: dW1_j = sum over ranks r (in rank order, own rank in its place) of X_{r,j}^T . dZ1_r, accumulated in
    // ONE TMEM accumulator. Peers' operands pass through two shared-memory stages: this warp loads the X^T tile (TMA from
    // the gathered input buffer in this GPU's memory), the workers copy the dZ1 image out of the receive area.
    // The stage regions are free once the weight-gradient GEMMs issued after the last dgrad have completed.
    mbar_wait(&bar_wgl, sp);
    if (kMid) mbar_wait(&bar_wg2, sp);
    if (st == 0 && !gdead) {
      // the peers' inputs for every step of this launch have been delivered (one ready word per copy CTA and source rank)
      for (int i = lane; i < W * kXPushCtas; i += 32) {
        const int src = i / kXPushCtas;
        if (src != p.rank && !pf_wait(p.xpad[p.rank] + kPadXReady + i, p.xneed, p, src, 11u, gdeadline)) s_xfail = 1;
      }
      __threadfence_system();
      __syncwarp();
      fence_proxy_async_all();  // written by the peers with ordinary stores, read here by TMA (async proxy)
    }
    const uint32_t gx0 = sb + (kMid ? oGX0 : oGX0b);
    auto fill = [&](int q) {
      const int s = q & 1;
      const uint32_t u = s ? guse1 : guse0;
      if (u > 0) mbar_wait_long(&bar_ge[s], (u - 1) & 1);  // the stage's previous MMAs have read it
      if (elect_one()) {
        const int r = q < p.rank ? q : q + 1;
        const int xrow = (r * p.xslots + p.slot0 + st) * B;
        mbar_arrive_expect_tx(&bar_gx[s], 4 * kColM);
        uint8_t* dst = smem + (s ? oGX1 : (kMid ? oGX0 : oGX0b));
        for (int a = 0; a < 4; ++a) tma_load_2d(dst + a * kColM, &tmXT, &bar_gx[s], k0 + 32 * a, xrow);
      }
      __syncwarp();
      if (s) ++guse1; else ++guse0;
    };
    for (int q = 0; q < npeer && q < 2; ++q) fill(q);
    const uint32_t idw = make_instr_desc(4, true, true, H1);
    int pending = -1;
    for (int r = 0; r < W; ++r) {
      uint32_t xt, dzt;
      int q = -1, s = 0;
      if (r == p.rank) {
        mbar_wait(&bar_ex[4], sp);
        xt = sb + oXT;
        dzt = sb + oDZ1;
      } else {
        q = r < p.rank ? r : r - 1;
        s = q & 1;
        mbar_wait_long(&bar_gf[s], (s ? gmma1 : gmma0) & 1u);
        xt = s ? sb + oGX1 : gx0;
        dzt = sb + (s ? oGD1 : oGD0);
      }
      fence_proxy_async_smem();
      tc_fence_after_sync();
      if (elect_one()) {
#pragma unroll
        for (int t = 0; t < 8; ++t)
          if (t < nb8) tc_mma<4>(tm + cWG1, mdesc(xt, t, kColM), mdesc(dzt, t, kColM), idw, (r | t) != 0);
        if (q >= 0) tc_commit(&bar_ge[s]);
        if (r == W - 1) tc_commit(&bar_mma);
      }
      __syncwarp();
      if (pending >= 0) { fill(pending); pending = -1; }  // (after this rank's MMAs are queued: the wait inside overlaps them)
      if (q >= 0) {
        if (s) ++gmma1; else ++gmma0;
        if (q + 2 < npeer) pending = q + 2;
      }
    }
  } else if (warp == 0) {
    mbar_wait(&bar_ex[4], sp);
    fence_proxy_async_smem();
    tc_fence_after_sync();
    if (elect_one()) {
      const uint32_t idw = make_instr_desc(4, true, true, H1);
#pragma unroll
      for (int t = 0; t < 8; ++t)
        if (t < nb8) tc_mma<4>(tm + cWG1, mdesc(sb + oXT, t, kColM), mdesc(sb + oDZ1, t, kColM), idw, t != 0);
      tc_commit(&bar_mma);
    }
    __syncwarp();
  }
  float dbl[16];  // last-layer bias gradient (valid in the thread that owns accumulator row SL)
  // Global-batch mode: the worker warps split up. The warp that owns TMEM lanes 0..31 (the only one whose rows of the side
  // weight gradients are real) drains and pushes them; the other three stage the peers' dZ1 at once.
  if (worker && (!gmode || (warp & 3) == 0)) {
    // while dW1 runs on the tensor core: drain the weight-gradient GEMMs that ran beside the chain
    FPROFT(32, 18);
    FPROFT(128, 34);
    mbar_wait(&bar_wgl, sp);
    FPROFT(128, 35);
    tc_fence_after_sync();
    {
      uint32_t r[16];
      tmem_ld_x16(tm + tq + cWGL, r);
      tmem_ld_wait();
      if (gmode) {
        // pushed as one contiguous SL x 64 B block (rows < SL sit in the lanes of the warp that owns TMEM lanes 0..31)
        if ((warp & 3) == 0) {
          if (elect_one()) bulk_wait_read();  // this warp's dZ1 copies have left the staging half
          __syncwarp();
          if (row <= SL) {  // rows < SL: the dW slice; row SL (the ones row): the last bias gradient
#pragma unroll
            for (int c = 0; c < 16; c += 4)
              sts128(sb + oSTG + row * 64 + c * 4, __uint_as_float(r[c]), __uint_as_float(r[c + 1]), __uint_as_float(r[c + 2]), __uint_as_float(r[c + 3]));
          }
          fence_proxy_async_smem();
          __syncwarp();
          if (gpush >= 0 && elect_one()) {
            const size_t o = static_cast<size_t>(gpush) + p.w_off[LAST] + static_cast<size_t>(SL * j) * 16;
            #pragma unroll 1
            for (int q = 0; q < W; ++q) {
              bulk_s2g(p.xrecv[q] + o, sb + oSTG, SL * 64);
              if (j == 0) bulk_s2g(p.xrecv[q] + gpush + p.b_off[LAST], sb + oSTG + SL * 64, 64);
            }
            bulk_commit();
          }
        }
      } else if (row < SL) {
        const size_t off = p.w_off[LAST] + static_cast<size_t>(SL * j + row) * 16;
#pragma unroll
        for (int c = 0; c < 16; c += 4)
          emit_grad4(p, step, off + c, xs, kXsWL + row * 64 + c * 4, __uint_as_float(r[c]), __uint_as_float(r[c + 1]), __uint_as_float(r[c + 2]), __uint_as_float(r[c + 3]), gdst, gpush);
      }
      if (row == SL) {  // the ones row: column sums of dZ_last = gradient of the last bias
#pragma unroll
        for (int c = 0; c < 16; ++c) dbl[c] = __uint_as_float(r[c]);
        // (global-batch mode: pushed with the dW slice above - a push changes no weight, so it need not wait for wgrad1)
      }
    }
    FPROFT(128, 36);
    if (kMid) {
      mbar_wait(&bar_wg2, sp);
      FPROFT(128, 37);
      tc_fence_after_sync();
#pragma unroll 1
      for (int cc = 0; cc < H2; cc += 16) {
        uint32_t r[16];
        tmem_ld_x16(tm + tq + cWG2 + cc, r);
        tmem_ld_wait();
        if (gmode) {
          if (row < S1) {  // staged: the slice is pushed as one contiguous block after the loop
#pragma unroll
            for (int c = 0; c < 16; c += 4)
              sts128(sb + oSTG2 + row * (H2 * 4) + (cc + c) * 4, __uint_as_float(r[c]), __uint_as_float(r[c + 1]), __uint_as_float(r[c + 2]), __uint_as_float(r[c + 3]));
          }
        } else if (row < S1) {
          const size_t off = p.w_off[1] + static_cast<size_t>(S1 * j + row) * H2 + cc;
#pragma unroll
          for (int c = 0; c < 16; c += 4)
            emit_grad4(p, step, off + c, xs, kXsW2 + row * (H2 * 4) + (cc + c) * 4, __uint_as_float(r[c]), __uint_as_float(r[c + 1]), __uint_as_float(r[c + 2]), __uint_as_float(r[c + 3]), gdst, gpush);
        }
        if (row == S1 && (S2 * j) / 16 == cc / 16) {  // ones row: db2; this CTA owns b2[S2*j, S2*j+S2)
          const int c0 = (S2 * j) % 16;
          const size_t off = p.b_off[1] + S2 * j;
#pragma unroll
          for (int c = 0; c < 16; c += 4)
            if (c >= c0 && c < c0 + S2) {
              if (gmode)  // staged behind the last-layer block of this warp's staging half, pushed with the dW2 slice
                sts128(sb + oSTG + 1280 + (c - c0) * 4, __uint_as_float(r[c]), __uint_as_float(r[c + 1]), __uint_as_float(r[c + 2]), __uint_as_float(r[c + 3]));
              else
                emit_grad4(p, step, off + (c - c0), xs, kXsB2 + (c - c0) * 4, __uint_as_float(r[c]), __uint_as_float(r[c + 1]), __uint_as_float(r[c + 2]), __uint_as_float(r[c + 3]), gdst, gpush);
            }
        }
      }
    }
    if (gmode && kMid && (warp & 3) == 0) {
      fence_proxy_async_smem();
      __syncwarp();
      if (gpush >= 0 && elect_one()) {
        #pragma unroll 1
        for (int q = 0; q < W; ++q) {
          bulk_s2g(p.xrecv[q] + gpush + p.w_off[1] + static_cast<size_t>(S1 * j) * H2, sb + oSTG2, S1 * H2 * 4);
          bulk_s2g(p.xrecv[q] + gpush + p.b_off[1] + S2 * j, sb + oSTG + 1280, S2 * 4);
        }
        bulk_commit();
      }
    }
    FPROFT(128, 38);
  }
  if (worker) {
    if (gmode && (warp & 3) != 0) {
      FPROF(15);
      const int wt = static_cast<int>(threadIdx.x) - 32;  // 0..95: warps 1..3
      // ---- peers' dZ1 images -> stages (the control warp loads the matching X^T tiles and issues the MMAs)
      for (int q = 0; q < npeer; ++q) {
        const int s = q & 1, r = q < p.rank ? q : q + 1;
        const uint32_t u = s ? guse1 : guse0;
        if (u > 0) mbar_wait_long(&bar_ge[s], (u - 1) & 1);
        if (q == 0) { FPROFT(32, 19); FPROFT(128, 39); }
        const float* img = p.xrecv[p.rank] + (static_cast<size_t>(xpar) * W + r) * p.total + p.w_off[0];
        const uint32_t dzt = sb + (s ? oGD1 : oGD0);
        // 21-22 chunks of 16 bytes per thread, copied global -> shared asynchronously (no registers, all in flight at once,
        // rolled loops: the code executed per step is refetched from L2 every time, so its size counts) straight into
        // their place in the operand image, then validated there; chunks that were still empty are copied again.
        {
          uint32_t pend = 0;
          for (int k = 0; k < 22; ++k) {
            const int c = wt + 96 * k;  // 2048 chunks over 96 threads
            if (c < 2048 && ((c & 255) >> 2) < B) pend |= 1u << k;  // rows >= B are never sent
          }
          uint32_t spins = 0;
          while (pend != 0u) {
            for (int k = 0; k < 22; ++k) {
              // chunk c of the receive image = source CTA c / 256, batch row (c % 256) / 4, columns 4 * (c % 4) ..
              const int c = wt + 96 * k;
              if ((pend >> k) & 1u) cp_async16(dzt + moff((c & 255) >> 2, S1 * (c >> 8) + 4 * (c & 3)), img + 4 * c);
            }
            cp_async_wait_all();
            for (int k = 0; k < 22; ++k) {
              const int c = wt + 96 * k;
              if ((pend >> k) & 1u) {
                const float4 v = lds128(dzt + moff((c & 255) >> 2, S1 * (c >> 8) + 4 * (c & 3)));
                if (gx_valid(v)) pend &= ~(1u << k);
              }
            }
            if (*xfail != 0) break;
            if (pend != 0u && (++spins & 0xF) == 0 && gx_timeout(p, r, 9u, gdeadline)) *xfail = 1;
          }
          if (q == 0) { FPROFT(32, 20); FPROFT(128, 40); }
        }
        mbar_wait_long(&bar_gx[s], u & 1);  // the X^T tile has landed: its (unused) column 127 becomes the ones column
        if (q == 0) { FPROFT(32, 22); FPROFT(128, 42); }
        if (wt < 64) {
          const uint32_t xt = sb + (s ? oGX1 : (kMid ? oGX0 : oGX0b));
          asm volatile("st.shared.f32 [%0], %1;" ::"r"(xt + moff(wt, 127)), "f"(1.f) : "memory");
        }
        fence_proxy_async_smem();
        mbar_arrive(&bar_gf[s]);
        if (s) ++guse1; else ++guse0;
      }
      FPROF(0);
    }
  }
  if (gmode) {
    // ---- small tensors (W2 / last-layer W slices, b2, last bias): summed while the last wgrad1 MMAs run, by ALL 160
    // threads (the control warp has issued everything, the staging warps are done, the pushing warp too): at most two
    // 16-byte items per thread, every rank's contribution (own included) of both in flight at once, ranks in groups of
    // four added one group after the other (rank order). Applied with a fire-and-forget reduction, w += -(lr/world)*sum:
    // every element is touched by one thread, so the result equals w - (lr/world)*sum bit for bit on every rank.
    constexpr int nW2 = kMid ? S1 * H2 / 4 : 0, nWL = SL * 16 / 4, nB2 = kMid ? S2 / 4 : 0;
    static_assert(nW2 + nWL + nB2 + 4 <= 2 * kFThreads, "more than two small-tensor items per thread");
    const int n_items = nW2 + nWL + nB2 + (j == 0 ? 4 : 0);
    const int tid = static_cast<int>(threadIdx.x);
    float* rb = p.xrecv[p.rank] + static_cast<size_t>(xpar) * W * p.total;
    const float4 e4 = make_float4(__uint_as_float(kGxEmpty), __uint_as_float(kGxEmpty), __uint_as_float(kGxEmpty), __uint_as_float(kGxEmpty));
    size_t offs[2];
    float4 acc[2];
#pragma unroll
    for (int i = 0; i < 2; ++i) {
      int k = tid + kFThreads * i;
      size_t off;
      if (k < nW2) off = p.w_off[1] + static_cast<size_t>(S1 * j) * H2;
      else if ((k -= nW2) < nWL) off = p.w_off[LAST] + static_cast<size_t>(SL * j) * 16;
      else if ((k -= nWL) < nB2) off = p.b_off[1] + S2 * j;
      else { k -= nB2; off = p.b_off[LAST]; }
      offs[i] = off + static_cast<size_t>(k) * 4;
      acc[i] = make_float4(0.f, 0.f, 0.f, 0.f);
    }
    for (int r0 = 0; r0 < W; r0 += 4) {
      uint32_t pend = 0;
#pragma unroll
      for (int i = 0; i < 2; ++i)
#pragma unroll
        for (int rr = 0; rr < 4; ++rr)
          if (tid + kFThreads * i < n_items && r0 + rr < W) pend |= 1u << (i * 4 + rr);
      float4 v[2][4];
      uint32_t spins = 0;
      while (pend != 0u) {
#pragma unroll
        for (int i = 0; i < 2; ++i)
#pragma unroll
          for (int rr = 0; rr < 4; ++rr)
            if ((pend >> (i * 4 + rr)) & 1u) v[i][rr] = ld_poll_f4(rb + static_cast<size_t>(r0 + rr) * p.total + offs[i]);
#pragma unroll
        for (int i = 0; i < 2; ++i)
#pragma unroll
          for (int rr = 0; rr < 4; ++rr)
            if (((pend >> (i * 4 + rr)) & 1u) && (gx_valid(v[i][rr]) || *xfail != 0)) pend &= ~(1u << (i * 4 + rr));
        // (on a timeout the record names the rank whose contribution is still missing: bit i*4+rr = rank r0+rr)
        if (pend != 0u && (++spins & 0xF) == 0 && gx_timeout(p, r0 + ((__ffs(pend) - 1) & 3), 10u, gdeadline)) *xfail = 1;
      }
#pragma unroll
      for (int rr = 0; rr < 4; ++rr)
#pragma unroll
        for (int i = 0; i < 2; ++i)
          if (tid + kFThreads * i < n_items && r0 + rr < W) {
            acc[i].x += v[i][rr].x; acc[i].y += v[i][rr].y; acc[i].z += v[i][rr].z; acc[i].w += v[i][rr].w;
            *reinterpret_cast<float4*>(rb + static_cast<size_t>(r0 + rr) * p.total + offs[i]) = e4;
          }
    }
#pragma unroll
    for (int i = 0; i < 2; ++i)
      if (tid + kFThreads * i < n_items)
        asm volatile("red.global.add.v4.f32 [%0], {%1, %2, %3, %4};" ::"l"(p.wparams + offs[i]), "f"(-step * acc[i].x),
                     "f"(-step * acc[i].y), "f"(-step * acc[i].z), "f"(-step * acc[i].w) : "memory");
    FPROF(1);
  }
  if (worker) {
    mbar_wait(&bar_mma, mma_phase);
    FPROF(11);
    tc_fence_after_sync();
#pragma unroll 1
    for (int cc = 0; cc < H1; cc += 16) {
      uint32_t r[16];
      tmem_ld_x16(tm + tq + cWG1 + cc, r);
      tmem_ld_wait();
      if (row < nk1 * 8) {
        const size_t off = p.w_off[0] + static_cast<size_t>(k0 + row) * H1 + cc;
#pragma unroll
        for (int c = 0; c < 16; c += 4)
          emit_grad4(p, step, off + c, xs, kXsW1 + row * (H1 * 4) + (cc + c) * 4, __uint_as_float(r[c]), __uint_as_float(r[c + 1]), __uint_as_float(r[c + 2]), __uint_as_float(r[c + 3]), gdst);
      }
      if (row == 127 && cc == S1 * j) {  // ones row: db1; this CTA owns b1[S1*j, S1*j+S1)
        const size_t off = p.b_off[0] + S1 * j;
#pragma unroll
        for (int c = 0; c < 16; c += 4)
          emit_grad4(p, step, off + c, xs, kXsB1 + c * 4, __uint_as_float(r[c]), __uint_as_float(r[c + 1]), __uint_as_float(r[c + 2]), __uint_as_float(r[c + 3]), gdst);
      }
    }
  }
  mma_phase ^= 1;
  // Last-layer bias: every CTA consumed its copy before delivering dZ1, and wgrad1 (just waited for) was issued
  // after all dZ1 arrived -> CTA 0 may update it now; the end-of-step barrier publishes it.
  if (!gmode && j == 0 && worker && row == SL) {
    const size_t off = p.b_off[LAST];
#pragma unroll
    for (int c = 0; c < 16; c += 4) emit_grad4(p, step, off + c, xs, kXsBL + c * 4, dbl[c], dbl[c + 1], dbl[c + 2], dbl[c + 3], gdst);
  }
  FPROF(12);
  if (pmode) {
    // gradients performed at GPU scope (they sit in this GPU's L2, which is where peer reads are served from) before the
    // flag: one fence by one thread after the CTA barrier (cumulative), not a MEMBAR.SYS per thread (~4 us)
    __syncthreads();
    const bool pdead = p.fault_step >= 0 && static_cast<long long>(p.g0) + st >= p.fault_step;  // fault injection
    if (threadIdx.x < p.world && !pdead) {  // one thread per rank (own included): that rank's reducers poll [my rank][my CTA] locally
      if (p.strict == 1) __threadfence_system(); else
      __threadfence();
      st_relaxed_sys(p.xpad[threadIdx.x] + kPadFusedFlags + kPfRedFlags + p.rank * kFC + j, epoch);
    }
  }
  if (bmode) {
    // Single GPU: the staged -lr*g images are added to the weights by the bulk copy engine
    // (cp.reduce.async.bulk ... add.f32): one instruction per range instead of ~3,300 red.global.add.v4 from the
    // epilogue threads.
    fence_proxy_async_smem();
    __syncthreads();
    if (warp == 0 && elect_one()) {
      bulk_red_add_f32(p.wparams + p.w_off[0] + static_cast<size_t>(k0) * H1, sb + kXsW1, nk1 * 8 * H1 * 4);
      if (kMid) {
        bulk_red_add_f32(p.wparams + p.w_off[1] + static_cast<size_t>(S1 * j) * H2, sb + kXsW2, S1 * H2 * 4);
        bulk_red_add_f32(p.wparams + p.b_off[1] + S2 * j, sb + kXsB2, S2 * 4);
      }
      bulk_red_add_f32(p.wparams + p.w_off[LAST] + static_cast<size_t>(SL * j) * 16, sb + kXsWL, SL * 16 * 4);
      bulk_red_add_f32(p.wparams + p.b_off[0] + S1 * j, sb + kXsB1, S1 * 4);
      if (j == 0) bulk_red_add_f32(p.wparams + p.b_off[LAST], sb + kXsBL, 64);
      bulk_commit();
      bulk_wait_all();  // performed before the end-of-step barrier lets the next step load the weights
    }
  }
  if (xmode) {
    // ---------------------------------------------------------------- gradient exchange (data parallel)
    // This CTA's gradient slices sit in a staging image in shared memory. The bulk copy engine pushes the image
    // into slot [step parity][my rank] of every PEER's receive area (posted writes over NVLink: no round trips,
    // one instruction per range). One arrival flag per (source rank, CTA): CTA j only needs the peers' CTA j
    // (the last bias belongs to CTA 0 on every rank). Then the peers' slices and my weights are bulk-loaded
    // back into shared memory chunk by chunk, summed in RANK ORDER — the same order on every rank, so the
    // replicas stay bit-identical — w -= lr/world * sum is applied and the weights are bulk-stored.
    // Slot reuse two steps later is safe: a peer's CTA j cannot send step s+1 before it has consumed step s.
    fence_proxy_async_smem();  // staged gradients (generic stores) -> bulk copies (async proxy)
    __syncthreads();
    const int W = p.world;
    const size_t slot_me = (static_cast<size_t>(xpar) * W + p.rank) * p.total;
    const float* rbase = p.xrecv[p.rank] + static_cast<size_t>(xpar) * W * p.total;  // my receive area, this parity
    const size_t w1_off = p.w_off[0] + static_cast<size_t>(k0) * H1;
    // The layer-1 slice goes through the receive stages in 4 (world <= 4) or 8 chunks; a stage holds my weights
    // chunk + one chunk per peer, and as many stages as fit into the 8 x 13 KiB of dead shared memory are in flight
    // (world 2: the whole slice at once).
    const int nchunk = W <= 4 ? 4 : 8;
    const uint32_t chunk_rows = nk1 * 8 / nchunk, chunk_bytes = chunk_rows * H1 * 4;
    const uint32_t buf_stride = W <= 4 ? kXChunk : kXChunk / 2;
    int nstage = static_cast<int>((8 * kXChunk) / (W * buf_stride));
    if (nstage > 4) nstage = 4;
    // receive stage s: buffer 0 = my weights chunk, buffers 1.. = the peers' gradient chunks in rank order
    auto stage_buf = [&](int sidx, int b) -> uint32_t { return sb + kXr + (sidx * W + b) * buf_stride; };
    auto issue_chunk = [&](int c) {  // elected lane of warp 0
      const int sidx = c % nstage;
      mbar_arrive_expect_tx(&bar_xr[sidx], W * chunk_bytes);
      const size_t o = w1_off + static_cast<size_t>(c) * chunk_rows * H1;
      bulk_g2s(stage_buf(sidx, 0), p.wparams + o, chunk_bytes, &bar_xr[sidx]);
      int b = 1;
      for (int r = 0; r < W; ++r)
        if (r != p.rank) bulk_g2s(stage_buf(sidx, b++), rbase + static_cast<size_t>(r) * p.total + o, chunk_bytes, &bar_xr[sidx]);
    };
    if (warp == 0) {
      if (elect_one()) {
        for (int r = 0; r < W; ++r) {
          if (r == p.rank) continue;
          float* d = p.xrecv[r] + slot_me;
          bulk_s2g(d + w1_off, sb + kXsW1, nk1 * 8 * H1 * 4);
          if (kMid) {
            bulk_s2g(d + p.w_off[1] + static_cast<size_t>(S1 * j) * H2, sb + kXsW2, S1 * H2 * 4);
            bulk_s2g(d + p.b_off[1] + S2 * j, sb + kXsB2, S2 * 4);
          }
          bulk_s2g(d + p.w_off[LAST] + static_cast<size_t>(SL * j) * 16, sb + kXsWL, SL * 16 * 4);
          bulk_s2g(d + p.b_off[0] + S1 * j, sb + kXsB1, S1 * 4);
          if (j == 0) bulk_s2g(d + p.b_off[LAST], sb + kXsBL, 64);
        }
        bulk_commit();
        bulk_wait_all();         // the pushes are complete ...
        fence_proxy_async_all();
        __threadfence_system();  // ... and ordered before the flags
        for (int r = 0; r < W; ++r)
          if (r != p.rank) st_relaxed_sys(p.xpad[r] + kPadFusedFlags + p.rank * 8 + j, epoch);
      }
      __syncwarp();
      if (lane < W && lane != p.rank) {
        const int peer = lane;
        const uint32_t* f = p.xpad[p.rank] + kPadFusedFlags + peer * 8 + j;
        const unsigned long long deadline = globaltimer_ns() + p.xtimeout_ns;
        uint32_t spins = 0;
        while (static_cast<int32_t>(ld_relaxed_sys(f) - epoch) < 0) {
          if ((++spins & 0x3F) == 0) {
            const uint32_t ab = ld_relaxed_sys(p.xpad[p.rank] + kPadAbortWord);
            if (ab != 0u) { record_abort(p.xstatus, p.rank, ab, 5u); s_xfail = 1; break; }
            if (globaltimer_ns() > deadline) {
              if (p.xstatus != nullptr) {
                p.xstatus[1] = static_cast<uint32_t>(p.rank);
                p.xstatus[2] = static_cast<uint32_t>(peer);
                p.xstatus[3] = 5u;  // phase: fused-step gradient exchange
                __threadfence_system();
                p.xstatus[0] = DMLP_ERR_TIMEOUT;
              }
              for (int q = 0; q < W; ++q) st_relaxed_sys(p.xpad[q] + kPadAbortWord, abort_word_for(p.rank, peer));
              s_xfail = 1;
              break;
            }
          }
        }
      }
      if (p.strict == 1) __threadfence_system();
      __syncwarp();
    }
    __syncthreads();
    FPROF(15);
    if (s_xfail == 0) {
      const float sc = step / static_cast<float>(W);
      if (warp == 0 && elect_one()) {
        for (int c = 0; c < nstage; ++c) issue_chunk(c);
      }
      FPROF(0);  // (slots 0/1 are re-used inside the step loop: first loads issued / layer-1 slice reduced)
      // the layer-1 slice: 4 chunks through the receive stages
      const int items = chunk_rows * H1 / 4;
      for (int c = 0; c < nchunk; ++c) {
        const int sidx = c % nstage;
        mbar_wait(&bar_xr[sidx], (xr_phase >> sidx) & 1u);
        xr_phase ^= 1u << sidx;
        for (int i = threadIdx.x; i < items; i += kFThreads) {
          float4 acc = make_float4(0.f, 0.f, 0.f, 0.f);
          int b = 1;
#pragma unroll
          for (int r = 0; r < 8; ++r)
            if (r < W) {
              const float4 g = (r == p.rank) ? lds128(sb + kXsW1 + c * chunk_bytes + i * 16) : lds128(stage_buf(sidx, b++) + i * 16);
              acc.x += g.x; acc.y += g.y; acc.z += g.z; acc.w += g.w;
            }
          float4 w = lds128(stage_buf(sidx, 0) + i * 16);
          w.x -= sc * acc.x; w.y -= sc * acc.y; w.z -= sc * acc.z; w.w -= sc * acc.w;
          sts128(stage_buf(sidx, 0) + i * 16, w.x, w.y, w.z, w.w);
        }
        fence_proxy_async_smem();
        __syncthreads();
        if (warp == 0 && elect_one()) {
          bulk_s2g(p.wparams + w1_off + static_cast<size_t>(c) * chunk_rows * H1, stage_buf(sidx, 0), chunk_bytes);
          bulk_commit();
          if (c + nstage < nchunk) {
            bulk_wait_read();  // the engine has read this stage's weights buffer: the stage may be refilled
            issue_chunk(c + nstage);
          }
        }
      }
      FPROF(1);
      // Small ranges (layer-2/3 slices, bias slices) through registers while the first chunks travel: one flat item
      // list so that every thread issues all its loads at once (mine from the staging image, the peers' from my
      // receive area with L1-bypassing loads: a peer wrote them).
      {
        constexpr int nW2 = kMid ? S1 * H2 / 4 : 0, nWL = SL * 16 / 4, nB1 = S1 / 4, nB2 = kMid ? S2 / 4 : 0;
        const int n_items = nW2 + nWL + nB1 + nB2 + (j == 0 ? 4 : 0);
        for (int it = threadIdx.x; it < n_items; it += kFThreads) {
          size_t off;
          uint32_t soff;
          int k = it;
          if (k < nW2) { off = p.w_off[1] + static_cast<size_t>(S1 * j) * H2; soff = kXsW2; }
          else if ((k -= nW2) < nWL) { off = p.w_off[LAST] + static_cast<size_t>(SL * j) * 16; soff = kXsWL; }
          else if ((k -= nWL) < nB1) { off = p.b_off[0] + S1 * j; soff = kXsB1; }
          else if ((k -= nB1) < nB2) { off = p.b_off[1] + S2 * j; soff = kXsB2; }
          else { k -= nB2; off = p.b_off[LAST]; soff = kXsBL; }
          off += static_cast<size_t>(k) * 4;
          float4 w = __ldcg(reinterpret_cast<const float4*>(p.wparams + off));
          float4 acc = make_float4(0.f, 0.f, 0.f, 0.f);
#pragma unroll
          for (int r = 0; r < 8; ++r)
            if (r < W) {
              const float4 g = (r == p.rank) ? lds128(sb + soff + k * 16)
                                             : __ldcg(reinterpret_cast<const float4*>(rbase + static_cast<size_t>(r) * p.total + off));
              acc.x += g.x; acc.y += g.y; acc.z += g.z; acc.w += g.w;
            }
          w.x -= sc * acc.x; w.y -= sc * acc.y; w.z -= sc * acc.z; w.w -= sc * acc.w;
          *reinterpret_cast<float4*>(p.wparams + off) = w;
        }
      }
      if (warp == 0 && elect_one()) bulk_wait_all();  // updated weights are in memory before the step ends
    }
  }
  // ---- end of step: my weight updates are performed (fence), my reads of exchanged data are done, the
  // exchange barriers are re-armed BEFORE any peer can pass the barrier and start sending the next step's data.
  if (st + 1 < nsteps) {
    __threadfence();
    fence_proxy_async_all();  // the next step's TMA loads (async proxy) re-read weights updated by red.global
    if (threadIdx.x == 0) arm_exchanges();
  }
  tc_fence_before_sync();
  __syncthreads();
  cluster_sync_all();  // last step: nobody leaves while a peer could still be writing into its shared memory
  tc_fence_after_sync();
  FPROF(14);
  }  // step loop
  if (gmode && worker && elect_one()) bulk_wait_read();  // pushes still reading this CTA's staging area
  if ((xmode || pmode) && j == 0 && threadIdx.x == 0) p.xpad[p.rank][kPadFusedEpoch] = epoch0 + static_cast<uint32_t>(nsteps);
  // global-batch mode: this rank no longer reads the input slots of these steps (every CTA is past the last barrier):
  // the peers' copy kernels may overwrite them
  if (gmode && j == 0 && threadIdx.x < W && static_cast<int>(threadIdx.x) != p.rank &&
      !(p.fault_step >= 0 && static_cast<long long>(p.g0) + nsteps > p.fault_step))
    st_relaxed_sys(p.xpad[threadIdx.x] + kPadXDone + p.rank, static_cast<uint32_t>(p.g0) + static_cast<uint32_t>(nsteps));
  if (warp == 0) {
    tc_fence_after_sync();
    tmem_dealloc(tm, 512);
  }
}

struct FusedPlan {
  CUtensorMap tmXA, tmXT, tmW1, tmW2B, tmW2K, tmWLB, tmWLK;
  FusedParams p;
  int layers;
  int slots;
};

}  // namespace dmlp

using namespace dmlp;

extern "C" int dmlp_mlp_fused_supported(int num_layers, const int* dims, int batch) {
  if (num_layers != 2 && num_layers != 3) return 0;
  if (batch < 8 || batch > 64 || batch % 8) return 0;
  if (dims[0] % 8 || dims[0] > 832 || dims[0] < 64) return 0;
  if (dims[1] != 128) return 0;
  if (num_layers == 3 && dims[2] != 64) return 0;
  if (dims[num_layers] > 16) return 0;
  return 1;
}

extern "C" int dmlp_mlp_fused_plan_create(const dmlp_mlp_fused_desc_t* d, dmlp_mlp_fused_plan_t* out) {
  DMLP_REQUIRE(d && out, "dmlp_mlp_fused_plan_create: null argument");
  DMLP_REQUIRE(dmlp_mlp_fused_supported(d->num_layers, d->dims, d->batch),
               "fused step supports batch<=64 (x8), in<=832 (x8), hidden 128[,64], classes<=16");
  std::unique_ptr<FusedPlan> owner(new (std::nothrow) FusedPlan());  // freed on every early return below
  FusedPlan* pl = owner.get();
  if (!pl) return set_error(DMLP_ERR_CUDA, "out of host memory");
  const int L = d->num_layers, in0 = d->dims[0], H1 = d->dims[1], H2 = L == 3 ? d->dims[2] : 0;
  const int SL = L == 3 ? H2 / 8 : H1 / 8;
  const float* P = d->params_dev;
  int rc = 0;
#define MK(expr) do { rc = (expr); if (rc) return rc; } while (0)
  // X as K-major A (forward) and as MN-major A (layer-1 weight gradient)
  const int slots = d->x_slots > 1 ? d->x_slots : 1;
  DMLP_REQUIRE(slots == 1 || d->sgd_params_dev != nullptr, "multi-step launches need in-place SGD (sgd_params_dev)");
  // global-batch mode: x_dev is the gathered input buffer [world][slots][batch][in0] (own rank's ring at its rank index)
  const int xranks = (d->gx_enable && d->xchg_world > 1) ? d->xchg_world : 1;
  MK(make_tensor_map_2d(&pl->tmXA, d->x_dev, 4, false, in0, static_cast<uint64_t>(d->batch) * slots * xranks, in0, 32, kFBK, 1));
  MK(make_tensor_map_2d(&pl->tmXT, d->x_dev, 4, false, in0, static_cast<uint64_t>(d->batch) * slots * xranks, in0, 32, kFBK, 2));
  MK(make_tensor_map_2d(&pl->tmW1, P + d->w_off[0], 4, false, H1, in0, H1, 32, kKGroupsMax * 8, 2));
  const int last = L - 1;
  if (L == 3) {
    MK(make_tensor_map_2d(&pl->tmW2B, P + d->w_off[1], 4, false, H2, H1, H2, 32, H1 / 8, 2));
    MK(make_tensor_map_2d(&pl->tmW2K, P + d->w_off[1], 4, false, H2, H1, H2, 32, H1 / 8, 1));
  } else {
    pl->tmW2B = pl->tmW1;
    pl->tmW2K = pl->tmW1;
  }
  const int HL = L == 3 ? H2 : H1;
  MK(make_tensor_map_2d(&pl->tmWLB, P + d->w_off[last], 4, false, 16, HL, 16, 32, SL, 2));
  MK(make_tensor_map_2d(&pl->tmWLK, P + d->w_off[last], 4, false, 16, HL, 16, 32, SL, 1));
#undef MK
  FusedParams& p = pl->p;
  memset(&p, 0, sizeof(p));
  p.B = d->batch; p.in0 = in0; p.C = d->dims[L];
  p.params = P; p.grads = d->grads_dev;
  for (int l = 0; l < L; ++l) { p.w_off[l] = d->w_off[l]; p.b_off[l] = d->b_off[l]; }
  p.labels = d->labels_dev;
  p.metrics = d->metrics_dev;
  p.dbg = d->debug_dev;
  p.wparams = d->sgd_params_dev;
  p.lr = d->lr_dev;
  p.prof = reinterpret_cast<unsigned long long*>(d->prof_dev);
  pl->layers = L;
  pl->slots = slots;
  if (d->xchg_world > 1) {
    DMLP_REQUIRE(d->xchg_world <= 8 && d->xchg_rank >= 0 && d->xchg_rank < d->xchg_world, "fused step: bad exchange rank/world");
    DMLP_REQUIRE(d->sgd_params_dev != nullptr && d->lr_dev != nullptr && d->param_count > 0,
                 "fused step: the in-kernel gradient exchange applies SGD itself (sgd_params_dev, lr_dev, param_count)");
    p.world = d->xchg_world;
    p.rank = d->xchg_rank;
    for (int r = 0; r < d->xchg_world; ++r) {
      DMLP_REQUIRE((d->xchg_recv_dev[r] || (d->xchg_reducer_ctas > 0 && !d->gx_enable)) && d->xchg_pad_dev[r], "fused step: rank %d's receive area / signal pad is not mapped", r);
      p.xrecv[r] = d->xchg_recv_dev[r];
      p.xpad[r] = d->xchg_pad_dev[r];
    }
    p.xstatus = d->xchg_status;
    p.xtimeout_ns = d->xchg_timeout_ns ? d->xchg_timeout_ns : 10ull * 1000 * 1000 * 1000;
    p.total = d->param_count;
    if (d->gx_enable) {
      DMLP_REQUIRE(d->xchg_reducer_ctas == 0, "fused step: the global-batch mode has no reducer CTAs");
      p.gx = 1;
      p.xslots = slots;
    }
    if (d->xchg_reducer_ctas > 0) {
      DMLP_REQUIRE(d->xchg_reducer_ctas % kFC == 0 && d->xchg_reducer_ctas <= kMaxRedCtas && d->param_count % 4 == 0,
                   "fused step: reducer CTAs must come in clusters of %d (<= %d) and the parameter count be a multiple of 4", kFC,
                   kMaxRedCtas);
      p.red_ctas = d->xchg_reducer_ctas;
      for (int par = 0; par < 2; ++par)
        for (int r = 0; r < d->xchg_world; ++r) {
          DMLP_REQUIRE(d->xchg_grads_dev[par][r] != nullptr, "fused step: rank %d's gradient buffer is not mapped", r);
          p.xgrads[par][r] = d->xchg_grads_dev[par][r];
        }
    }
  }
  p.nsteps = 1;
  p.slot0 = 0;
  p.fault_step = -1;
  {
    const char* e = getenv("DMLP_COLL_STRICT");
    p.strict = e ? atoi(e) : 2;  // default: acquire loads on the flag polls (+0.05 us per rendezvous, measured at 2 GPUs)
  }
  {
    const char* e = getenv("DMLP_FUSED_BULK");
    p.bulk_update = e ? atoi(e) : 0;  // measured: 22.3 us/step with bulk reduce-adds vs 21.3 with red.global (default)
  }
  *out = reinterpret_cast<dmlp_mlp_fused_plan_t>(owner.release());
  return DMLP_OK;
}

extern "C" int dmlp_mlp_fused_plan_destroy(dmlp_mlp_fused_plan_t h) {
  delete reinterpret_cast<FusedPlan*>(h);
  return DMLP_OK;
}

template <int S1, int S2, bool GX>
static int launch_fused(const FusedPlan* pl, const FusedParams& prm, cudaStream_t st) {
  auto kern = mlp_fused_step_kernel<S1, S2, GX>;
  static unsigned long long attr_mask = 0;
  int dev = 0;
  DMLP_CUDA_CHECK(cudaGetDevice(&dev));
  if (!(attr_mask & (1ull << dev))) {
    DMLP_CUDA_CHECK(cudaFuncSetAttribute(kern, cudaFuncAttributeMaxDynamicSharedMemorySize, kFusedSmem));
    attr_mask |= 1ull << dev;
  }
  cudaLaunchConfig_t cfg = {};
  cfg.gridDim = dim3(kFC + (prm.red_ctas > 0 ? prm.red_ctas : 0));
  cfg.blockDim = dim3(kFThreads);
  cfg.dynamicSmemBytes = kFusedSmem;
  cfg.stream = st;
  cudaLaunchAttribute attr[1];
  attr[0].id = cudaLaunchAttributeProgrammaticStreamSerialization;
  attr[0].val.programmaticStreamSerializationAllowed = 1;
  cfg.attrs = attr;
  cfg.numAttrs = 1;
  if (prm.red_ctas > 0) {
    // The compute cluster and the reducer CTAs wait on each other: every cluster of the launch must be resident at once.
    static int max_clusters[64] = {0};
    if (dev < 64 && max_clusters[dev] == 0) {
      int n = 0;
      DMLP_CUDA_CHECK(cudaOccupancyMaxActiveClusters(&n, kern, &cfg));
      max_clusters[dev] = n > 0 ? n : -1;
    }
    const int need = 1 + prm.red_ctas / kFC;
    DMLP_REQUIRE(dev >= 64 || max_clusters[dev] >= need, "fused step: %d clusters of %d CTAs cannot be co-resident on this device (max %d)",
                 need, kFC, dev < 64 ? max_clusters[dev] : 0);
  }
  DMLP_CUDA_CHECK(cudaLaunchKernelEx(&cfg, kern, pl->tmXA, pl->tmXT, pl->tmW1, pl->tmW2B, pl->tmW2K, pl->tmWLB, pl->tmWLK, prm));
  DMLP_CUDA_CHECK(cudaGetLastError());
  return DMLP_OK;
}

extern "C" int dmlp_mlp_fused_step(dmlp_mlp_fused_plan_t h, void* stream) {
  const FusedPlan* pl = reinterpret_cast<const FusedPlan*>(h);
  DMLP_REQUIRE(pl, "null fused plan");
  if (pl->layers == 3) return launch_fused<16, 8, false>(pl, pl->p, as_stream(stream));
  return launch_fused<16, 0, false>(pl, pl->p, as_stream(stream));
}

// Global-batch data-parallel mode (plan created with gx_enable): `global_step` is the index of the first of these steps
// since the trainer started (identical on every rank; it selects the parity of the receive areas) and `inputs_ready` the
// step count the peers' input-ready words must have reached (dmlp_input_exchange) before their inputs are read.
extern "C" int dmlp_mlp_fused_steps_gx(dmlp_mlp_fused_plan_t h, int slot0, int nsteps, unsigned long long global_step,
                                       unsigned int inputs_ready, void* stream) {
  const FusedPlan* pl = reinterpret_cast<const FusedPlan*>(h);
  DMLP_REQUIRE(pl && (pl->p.gx || pl->p.red_ctas > 0), "dmlp_mlp_fused_steps_gx: the plan has no data-parallel mode with a global step");
  DMLP_REQUIRE(nsteps >= 1 && slot0 >= 0 && slot0 + nsteps <= pl->slots, "dmlp_mlp_fused_steps_gx: slots [%d, %d) outside the plan's %d",
               slot0, slot0 + nsteps, pl->slots);
  FusedParams prm = pl->p;
  prm.nsteps = nsteps;
  prm.slot0 = slot0;
  prm.g0 = global_step;
  prm.xneed = inputs_ready;
  prm.labels = pl->p.labels + static_cast<size_t>(slot0) * pl->p.B;
  prm.metrics = pl->p.metrics + 2 * static_cast<size_t>(slot0);
  if (!pl->p.gx) {  // persistent compute + reducer launch: the global step only feeds the fault injection
    if (pl->layers == 3) return launch_fused<16, 8, false>(pl, prm, as_stream(stream));
    return launch_fused<16, 0, false>(pl, prm, as_stream(stream));
  }
  if (pl->layers == 3) return launch_fused<16, 8, true>(pl, prm, as_stream(stream));
  return launch_fused<16, 0, true>(pl, prm, as_stream(stream));
}

// Fault injection for the global-batch mode: from global step `step` on this rank behaves as if it had died (sends
// nothing, waits for nothing); its peers' watchdogs then fire. -1 switches it off. Also refreshes the watchdog budget.
extern "C" int dmlp_mlp_fused_set_fault(dmlp_mlp_fused_plan_t h, long long step, unsigned long long timeout_ns) {
  FusedPlan* pl = reinterpret_cast<FusedPlan*>(h);
  DMLP_REQUIRE(pl, "null fused plan");
  pl->p.fault_step = step;
  if (timeout_ns) pl->p.xtimeout_ns = timeout_ns;
  return DMLP_OK;
}

extern "C" int dmlp_mlp_fused_steps(dmlp_mlp_fused_plan_t h, int slot0, int nsteps, void* stream) {
  const FusedPlan* pl = reinterpret_cast<const FusedPlan*>(h);
  DMLP_REQUIRE(pl, "null fused plan");
  DMLP_REQUIRE(nsteps >= 1 && slot0 >= 0 && slot0 + nsteps <= pl->slots, "dmlp_mlp_fused_steps: slots [%d, %d) outside the plan's %d",
               slot0, slot0 + nsteps, pl->slots);
  DMLP_REQUIRE(nsteps == 1 || pl->p.wparams != nullptr, "dmlp_mlp_fused_steps: several steps per launch need in-place SGD");
  FusedParams prm = pl->p;
  prm.nsteps = nsteps;
  prm.slot0 = slot0;
  prm.labels = pl->p.labels + static_cast<size_t>(slot0) * pl->p.B;
  prm.metrics = pl->p.metrics + 2 * static_cast<size_t>(slot0);
  if (pl->layers == 3) return launch_fused<16, 8, false>(pl, prm, as_stream(stream));
  return launch_fused<16, 0, false>(pl, prm, as_stream(stream));
}
