// Native data-parallel MLP trainer: owns the device buffers, pre-encodes every TMA descriptor,
// captures the whole training step into CUDA graphs and runs the input pipeline and the step loop
// without returning to Python.
//
// Replaces the reference's training loop (DSML/client/client.go:516-659): per batch it ran
// forwardPass / backwardPass on the CPU, pushed the gradient to every "device" over gRPC, called
// AllReduceRing, read the gradient back, applied SGD on the CPU and re-uploaded the weights
// (8 client RPCs + ~60 coordinator RPCs per batch). Here one step is:
//   H2D(x, y) on a copy stream  ->  graph { fwd GEMMs (+bias+ReLU), fused softmax-CE, dgrad/wgrad
//   GEMMs (+ReLU mask, +bias-gradient), all-reduce fused with average+SGD over peer memory }
//   -> D2H(loss, correct).
// Gradients are double-buffered by step parity, so the one-shot all-reduce needs a single
// rendezvous per step; with `overlap` on, each layer's bucket is reduced on a second stream while
// backward continues (layer-by-layer overlap).
#include "common.h"
#include "gemm_plan.h"
#include "wgrad_rs.h"

#include <string.h>

#include <new>
#include <vector>

// Implemented in collectives.cu: hands out the communicator's pads / state / watchdog words.
extern "C" int dmlp_comm_internal_view(dmlp_comm_t comm, int local_idx, uint32_t* pad[DMLP_MAX_RANKS], uint32_t** state,
                                       uint32_t** host_flags, unsigned long long* timeout_ns, uint32_t* dead_mask,
                                       int* rank, int* world, int* dev);

namespace dmlp {

constexpr int kMaxLayers = DMLP_MAX_LAYERS;
constexpr int kRingChunk = 8, kRingChunks = 2, kRingSlots = kRingChunk * kRingChunks;

static size_t align_up(size_t v, size_t a) { return (v + a - 1) / a * a; }

struct ParamLayout {
  int L = 0;
  int in[kMaxLayers], out[kMaxLayers], out_pad[kMaxLayers];
  size_t w_off[kMaxLayers], b_off[kMaxLayers];  // in floats
  size_t total = 0;                              // floats, multiple of 64
};

static void make_layout(const dmlp_trainer_config_t& c, ParamLayout* lay) {
  lay->L = c.num_layers;
  size_t cur = 0;
  for (int l = 0; l < c.num_layers; ++l) {
    lay->in[l] = c.dims[l];
    lay->out[l] = c.dims[l + 1];
    lay->out_pad[l] = static_cast<int>(align_up(c.dims[l + 1], 16));
    lay->w_off[l] = cur;
    cur = align_up(cur + static_cast<size_t>(lay->in[l]) * lay->out_pad[l], 64);
    lay->b_off[l] = cur;
    cur = align_up(cur + lay->out_pad[l], 64);
  }
  lay->total = cur;
}

struct Trainer {
  dmlp_trainer_config_t cfg;
  ParamLayout lay;
  int dev = 0;
  dmlp_comm_t comm = nullptr;
  int local_idx = 0, rank = 0, world = 1;
  bool bf16 = false;
  // peer-visible (arena) or private buffers
  float* params = nullptr;
  __nv_bfloat16* shadow = nullptr;
  float* grads[2] = {nullptr, nullptr};
  float* peer_params[DMLP_MAX_RANKS];
  void* peer_shadow[DMLP_MAX_RANKS];
  float* peer_grads[2][DMLP_MAX_RANKS];
  float* peer_recv[DMLP_MAX_RANKS];  // fused step with in-kernel gradient exchange: [2][world][total] receive areas
  unsigned long long* fused_prof = nullptr;  // DMLP_FUSED_PROF=1: [8 CTAs][64] phase stamps of the ring plan's last step
  bool gx = false;                    // global-batch mode (default on real GPUs): dZ1 / input exchange, see mlp_fused.cu
  float* peer_xall[DMLP_MAX_RANKS];   // rank r's gathered input buffer [world][kRingSlots][batch][in0] as mapped here
  unsigned long long gsteps = 0;      // steps run in global-batch mode (identical on every rank)
  unsigned long long ring_end[2] = {0, 0};  // gsteps at the end of the chunk that last used ring chunk c
  uint32_t* gx_pads[DMLP_MAX_RANKS];
  uint32_t* gx_status = nullptr;
  long long fault_step = -1;          // fault injection: this rank goes silent from this global step on
  bool failed = false;                // a collective watchdog fired: the trainer refuses further steps
  bool pdp = false;                   // persistent data-parallel launches (fused == 4): compute cluster + reducer CTAs
  bool xchg = false;                  // ... is in use (multi-GPU, fused step, fused != 2, no NVLS arena)
  bool owns_params = false;
  bool nvls = false;  // params/grads live in multicast-bound memory; all-reduce by multimem
  float* mc_params = nullptr;
  float* mc_grads[2] = {nullptr, nullptr};
  void* mc_shadow = nullptr;
  // private buffers
  float* x[2] = {nullptr, nullptr};
  __nv_bfloat16* x16 = nullptr;
  int* labels[2] = {nullptr, nullptr};
  float* act[kMaxLayers];            // fp32 activations (tf32 mode) / logits (always, last layer)
  __nv_bfloat16* act16[kMaxLayers];  // bf16 activations (bf16 mode)
  float* dz[kMaxLayers];
  __nv_bfloat16* dz16[kMaxLayers];
  float* metrics_dev[2] = {nullptr, nullptr};  // {loss_sum, correct(int bits)}
  float* metrics_host = nullptr;               // pinned [2][2]
  float* lr_dev = nullptr;
  float* lr_host = nullptr;                    // pinned staging for lr updates
  cudaStream_t s_main = nullptr, s_comm = nullptr, s_copy = nullptr;
  cudaEvent_t ev_h2d[2], ev_done[2], ev_fork, ev_join, ev_bucket[kMaxLayers];
  bool launched[2] = {false, false};
  cudaGraph_t graph[2] = {nullptr, nullptr};
  cudaGraphExec_t gexec[2] = {nullptr, nullptr};
  GemmPlan fwd[2][kMaxLayers], wgrad[2][kMaxLayers], dgrad[kMaxLayers], fwd_eval_last;
  bool fused_step = false;  // whole fwd+loss+bwd in one cluster launch (mlp_fused.cu)
  dmlp_mlp_fused_plan_t fplan[2] = {nullptr, nullptr};
  // fused path input ring: kRingChunks chunks of kRingChunk steps; one H2D + one event per chunk so the
  // step kernels run back to back with no stream operation between them
  float* x_ring = nullptr;
  int* y_ring = nullptr;
  float* metrics_ring = nullptr;  // [kRingSlots][2] on the device, copied to the host once per chunk
  dmlp_mlp_fused_plan_t rplan[kRingSlots];
  // single-GPU mode: ONE plan over the whole ring; a chunk's steps run inside one launch (weights updated in place
  // between steps inside the kernel), see dmlp_mlp_fused_steps
  dmlp_mlp_fused_plan_t ring_plan = nullptr;
  cudaEvent_t ev_ring_h2d[2], ev_ring_done[2];
  bool ring_launched[2] = {false, false};
  bool fused_ce = false;   // softmax-CE runs in the last forward GEMM's epilogue
  bool direct = false;     // batch fits one M tile: reductions are stored, no zeroing needed
  cudaStream_t s_aux = nullptr;  // wgrad branch (runs beside the dgrad chain)
  cudaEvent_t ev_dz[kMaxLayers], ev_aux_done;
  // wgrad -> reduce-scatter -> SGD -> all-gather pipeline of the large layers (wgrad_rs.cuh): the layer's wgrad GEMM pushes
  // its tiles to their owners; consumer warps of the same kernel reduce, update and broadcast the tiles this rank owns
  bool emulated = false;
  bool rs_on[kMaxLayers] = {};
  int rs_layers = 0;
  cudaEvent_t ev_rs = nullptr, ev_rs_comm = nullptr;
  int kernels_per_step = 0;
  long long launches_last_run = 0;  // kernel launches of the most recent run (fused path: a launch may cover several steps)
  unsigned long long steps_done = 0;
  std::vector<void*> owned;
};

static int dev_alloc(Trainer* t, size_t bytes, void** out) {
  int rc = dmlp_malloc(t->dev, bytes, out);
  if (rc == DMLP_OK) t->owned.push_back(*out);
  return rc;
}

#define TRY(expr)        \
  do {                   \
    int _rc = (expr);    \
    if (_rc) return _rc; \
  } while (0)

static_assert(dmlp::kPadXDone + DMLP_MAX_RANKS <= dmlp::kPadRsBase && (dmlp::kPadRsBase + dmlp::kPadRsWords) * 4 <= 64 * 1024,
              "the reduce-scatter counters must sit behind the other users of the signal pad and inside its 64 KiB");

// Picks the layers whose weight gradient goes through the fused wgrad -> reduce-scatter -> SGD -> all-gather pipeline
// (layer-by-layer overlap on a real multi-GPU communicator, big layers whose tiles divide among the ranks) and rewires
// their wgrad plans. DMLP_WGRAD_RS=0 turns it off, DMLP_WGRAD_RS_MIN_KB moves the size threshold (default 8 MiB).
static int setup_wgrad_rs(Trainer* t) {
  const ParamLayout& L = t->lay;
  t->rs_layers = 0;
  for (int l = 0; l < kMaxLayers; ++l) t->rs_on[l] = false;
  if (t->comm == nullptr || t->nvls || t->cfg.overlap == 0 || t->fused_step || t->emulated) return DMLP_OK;
  const char* e = getenv("DMLP_WGRAD_RS");
  if (e && atoi(e) == 0) return DMLP_OK;
  const char* mk = getenv("DMLP_WGRAD_RS_MIN_KB");
  const size_t min_bytes = (mk ? static_cast<size_t>(atol(mk)) : 8192u) * 1024u;
  for (int r = 0; r < t->world; ++r)
    if (!t->peer_grads[0][r] || !t->peer_grads[1][r] || !t->peer_params[r] || (t->bf16 && !t->peer_shadow[r])) return DMLP_OK;
  uint32_t* pads[DMLP_MAX_RANKS];
  uint32_t *state = nullptr, *host_flags = nullptr, dead = 0;
  unsigned long long timeout_ns = 0;
  int crank = 0, cworld = 0, cdev = 0;
  TRY(dmlp_comm_internal_view(t->comm, t->local_idx, pads, &state, &host_flags, &timeout_ns, &dead, &crank, &cworld, &cdev));
  int used = 0;
  for (int l = 0; l < L.L; ++l) {
    const size_t wbytes = static_cast<size_t>(L.in[l]) * L.out_pad[l] * 4;
    if (wbytes < min_bytes || !gemm_plan_rs_eligible(&t->wgrad[0][l], t->world)) continue;
    const int words = kRsCounts + (t->wgrad[0][l].total_tiles + 15) / 16 * 16;
    if (used + words > kPadRsWords) continue;
    for (int par = 0; par < 2; ++par) {
      float* slots[DMLP_MAX_RANKS];
      uint32_t* counts[DMLP_MAX_RANKS];
      for (int r = 0; r < t->world; ++r) {
        slots[r] = t->peer_grads[par][r] + L.w_off[l];
        counts[r] = pads[r] + kPadRsBase + used + kRsCounts;
      }
      GemmRsGeometry g;
      TRY(gemm_plan_enable_rs(&t->wgrad[par][l], t->world, t->rank, slots, counts, &g));
      RsReduceParams a;
      memset(&a, 0, sizeof(a));
      a.world = t->world; a.rank = t->rank;
      a.ld = L.out_pad[l];
      a.tm = g.tm; a.bn = g.bn; a.subs = g.subs;
      a.tiles_m = g.tiles_m; a.tiles_n = g.tiles_n; a.raster_g = g.raster_g;
      a.tiles_per_owner = g.total_tiles / t->world;
      a.slots = slots[t->rank];
      a.counts = counts[t->rank];
      a.per_tile = 8u * static_cast<uint32_t>(g.subs);
      for (int r = 0; r < t->world; ++r) {
        a.w[r] = t->peer_params[r] + L.w_off[l];
        a.w16[r] = t->bf16 ? static_cast<__nv_bfloat16*>(t->peer_shadow[r]) + L.w_off[l] : nullptr;
        a.done[r] = pads[r] + kPadRsBase + used + kRsDone;
        a.pad[r] = pads[r];
      }
      a.lr = t->lr_dev;
      a.scale = 1.f / static_cast<float>(t->world);
      a.epoch = pads[t->rank] + kPadRsBase + used + kRsEpoch;
      a.status = host_flags + 4;
      a.timeout_ns = timeout_ns ? timeout_ns : 10ull * 1000 * 1000 * 1000;
      { const char* pe = getenv("DMLP_RS_PROF"); a.prof = (pe && atoi(pe) > 0) ? 1 : 0; }
      t->wgrad[par][l].p.rsq = a;
    }
    t->rs_on[l] = true;
    ++t->rs_layers;
    used += words;
  }
  // counters start from zero for this trainer (the ranks meet in a barrier before the first step)
  if (t->rs_layers > 0) DMLP_CUDA_CHECK(cudaMemset(pads[t->rank] + kPadRsBase, 0, sizeof(uint32_t) * kPadRsWords));
  return DMLP_OK;
}

// Builds every GEMM plan (TMA descriptors included) for both input parities.
static int build_plans(Trainer* t) {
  const ParamLayout& L = t->lay;
  const int B = t->cfg.batch;
  const int dt = t->bf16 ? DMLP_BF16 : DMLP_F32;
  for (int par = 0; par < 2; ++par) {
    for (int l = 0; l < L.L; ++l) {
      const bool last = l == L.L - 1;
      dmlp_gemm_desc_t d;
      // ---- forward: act[l] = f(in_l . W_l + b_l)
      memset(&d, 0, sizeof(d));
      d.dtype = dt; d.a_major = 0; d.b_major = 1;
      d.M = B; d.N = L.out_pad[l]; d.K = L.in[l];
      if (l == 0) { d.A_dev = t->bf16 ? static_cast<void*>(t->x16) : static_cast<void*>(t->x[par]); d.lda = L.in[0]; }
      else { d.A_dev = t->bf16 ? static_cast<void*>(t->act16[l - 1]) : static_cast<void*>(t->act[l - 1]); d.lda = L.out_pad[l - 1]; }
      d.B_dev = t->bf16 ? static_cast<void*>(t->shadow + L.w_off[l]) : static_cast<void*>(t->params + L.w_off[l]);
      d.ldb = L.out_pad[l];
      d.bias_dev = t->params + L.b_off[l];
      d.relu = last ? 0 : 1;
      if (last || !t->bf16) { d.out_f32_dev = t->act[l]; d.ld_out = L.out_pad[l]; }
      if (!last && t->bf16) { d.out_bf16_dev = t->act16[l]; d.ld_out_bf16 = L.out_pad[l]; }
      if (last && t->fused_ce) {
        // logits stay in TMEM/registers: the epilogue writes dlogits, db, loss and hit count
        d.out_f32_dev = t->bf16 ? nullptr : t->dz[l];
        d.ld_out = L.out_pad[l];
        d.out_bf16_dev = t->bf16 ? static_cast<void*>(t->dz16[l]) : nullptr;
        d.ld_out_bf16 = L.out_pad[l];
        d.ce_labels_dev = t->labels[par];
        d.ce_classes = L.out[l];
        d.ce_scale = 1.f / static_cast<float>(B);
        d.colsum_dev = t->grads[par] + L.b_off[l];
        d.ce_loss_sum_dev = t->metrics_dev[par];
        d.ce_correct_dev = reinterpret_cast<int32_t*>(t->metrics_dev[par] + 1);
        if (par == 0) {  // evaluation: same kernel, no gradient outputs
          dmlp_gemm_desc_t e = d;
          e.out_f32_dev = nullptr; e.out_bf16_dev = nullptr; e.colsum_dev = nullptr;
          TRY(gemm_plan_create(&e, &t->fwd_eval_last));
        }
      }
      TRY(gemm_plan_create(&d, &t->fwd[par][l]));
      // ---- wgrad: dW_l = in_l^T . dZ_l  -> grads[par]
      memset(&d, 0, sizeof(d));
      d.dtype = dt; d.a_major = 1; d.b_major = 1;
      d.M = L.in[l]; d.N = L.out_pad[l]; d.K = B;
      if (l == 0) { d.A_dev = t->bf16 ? static_cast<void*>(t->x16) : static_cast<void*>(t->x[par]); d.lda = L.in[0]; }
      else { d.A_dev = t->bf16 ? static_cast<void*>(t->act16[l - 1]) : static_cast<void*>(t->act[l - 1]); d.lda = L.out_pad[l - 1]; }
      d.B_dev = t->bf16 ? static_cast<void*>(t->dz16[l]) : static_cast<void*>(t->dz[l]);
      d.ldb = L.out_pad[l];
      d.out_f32_dev = t->grads[par] + L.w_off[l];
      d.ld_out = L.out_pad[l];
      TRY(gemm_plan_create(&d, &t->wgrad[par][l]));
    }
  }
  TRY(setup_wgrad_rs(t));
  for (int l = 1; l < L.L; ++l) {
    // ---- dgrad: dZ_{l-1} = (dZ_l . W_l^T) * (act_{l-1} > 0); column sums -> db_{l-1} (set per parity at launch)
    dmlp_gemm_desc_t d;
    memset(&d, 0, sizeof(d));
    d.dtype = dt; d.a_major = 0; d.b_major = 0;
    d.M = B; d.N = L.in[l]; d.K = L.out_pad[l];
    d.A_dev = t->bf16 ? static_cast<void*>(t->dz16[l]) : static_cast<void*>(t->dz[l]);
    d.lda = L.out_pad[l];
    d.B_dev = t->bf16 ? static_cast<void*>(t->shadow + L.w_off[l]) : static_cast<void*>(t->params + L.w_off[l]);
    d.ldb = L.out_pad[l];
    if (t->bf16) { d.out_bf16_dev = t->dz16[l - 1]; d.ld_out_bf16 = L.out_pad[l - 1]; d.mask_dev = t->act16[l - 1]; d.mask_dtype = DMLP_BF16; }
    else { d.out_f32_dev = t->dz[l - 1]; d.ld_out = L.out_pad[l - 1]; d.mask_dev = t->act[l - 1]; d.mask_dtype = DMLP_F32; }
    d.ld_mask = L.out_pad[l - 1];
    d.colsum_dev = t->grads[0] + L.b_off[l - 1];  // patched per parity in enqueue_step
    TRY(gemm_plan_create(&d, &t->dgrad[l]));
  }
  if (t->fused_step) {
    for (int par = 0; par < 2; ++par) {
      dmlp_mlp_fused_desc_t f;
      memset(&f, 0, sizeof(f));
      f.num_layers = L.L;
      for (int l = 0; l <= L.L; ++l) f.dims[l] = t->cfg.dims[l];
      f.batch = B;
      f.x_dev = t->x[par];
      f.labels_dev = t->labels[par];
      f.params_dev = t->params;
      f.grads_dev = t->grads[par];
      for (int l = 0; l < L.L; ++l) { f.w_off[l] = L.w_off[l]; f.b_off[l] = L.b_off[l]; }
      // metrics go straight to pinned host memory (mapped): no D2H copy node in the step graph
      f.metrics_dev = t->metrics_host + 2 * par;
      if (t->comm == nullptr && t->cfg.fused != 2) { f.sgd_params_dev = t->params; f.lr_dev = t->lr_dev; }
      TRY(dmlp_mlp_fused_plan_create(&f, &t->fplan[par]));
    }
    for (int slot = 0; slot < kRingSlots; ++slot) {
      dmlp_mlp_fused_desc_t f;
      memset(&f, 0, sizeof(f));
      f.num_layers = L.L;
      for (int l = 0; l <= L.L; ++l) f.dims[l] = t->cfg.dims[l];
      f.batch = B;
      f.x_dev = t->x_ring + static_cast<size_t>(slot) * B * L.in[0];
      f.labels_dev = t->y_ring + static_cast<size_t>(slot) * B;
      f.params_dev = t->params;
      f.grads_dev = t->grads[slot & 1];
      for (int l = 0; l < L.L; ++l) { f.w_off[l] = L.w_off[l]; f.b_off[l] = L.b_off[l]; }
      f.metrics_dev = t->metrics_ring + 2 * slot;  // device memory: a write to mapped host memory stalls kernel completion
      if (t->comm == nullptr && t->cfg.fused != 2) { f.sgd_params_dev = t->params; f.lr_dev = t->lr_dev; }
      TRY(dmlp_mlp_fused_plan_create(&f, &t->rplan[slot]));
    }
    if ((t->comm == nullptr && t->cfg.fused != 2) || t->xchg || t->pdp || t->gx) {
      dmlp_mlp_fused_desc_t f;
      memset(&f, 0, sizeof(f));
      f.num_layers = L.L;
      for (int l = 0; l <= L.L; ++l) f.dims[l] = t->cfg.dims[l];
      f.batch = B;
      f.x_dev = t->gx ? t->peer_xall[t->rank] : t->x_ring;
      f.labels_dev = t->y_ring;
      f.params_dev = t->params;
      f.grads_dev = t->grads[0];
      for (int l = 0; l < L.L; ++l) { f.w_off[l] = L.w_off[l]; f.b_off[l] = L.b_off[l]; }
      f.metrics_dev = t->metrics_ring;
      f.sgd_params_dev = t->params;
      f.lr_dev = t->lr_dev;
      f.x_slots = kRingSlots;
      {
        const char* e = getenv("DMLP_FUSED_PROF");
        if (e && atoi(e) > 0) {
          void* pp = nullptr;
          TRY(dev_alloc(t, 8 * 64 * 8, &pp));
          t->fused_prof = static_cast<unsigned long long*>(pp);
          f.prof_dev = pp;
        }
      }
      if (t->xchg || t->pdp || t->gx) {
        uint32_t* pads[DMLP_MAX_RANKS];
        uint32_t *state = nullptr, *host_flags = nullptr, dead = 0;
        unsigned long long timeout_ns = 0;
        int crank = 0, cworld = 0, cdev = 0;
        TRY(dmlp_comm_internal_view(t->comm, t->local_idx, pads, &state, &host_flags, &timeout_ns, &dead, &crank, &cworld, &cdev));
        f.xchg_world = t->world;
        f.xchg_rank = t->rank;
        for (int r = 0; r < t->world; ++r) { f.xchg_recv_dev[r] = t->peer_recv[r]; f.xchg_pad_dev[r] = pads[r]; }
        f.xchg_status = host_flags + 4;
        f.xchg_timeout_ns = timeout_ns;
        f.param_count = L.total;
        if (t->gx) {
          f.gx_enable = 1;
          for (int r = 0; r < t->world; ++r) t->gx_pads[r] = pads[r];
          t->gx_status = host_flags + 4;
        }
        if (t->pdp) {
          static int red_env = -1;
          if (red_env < 0) { const char* e = getenv("DMLP_FUSED_REDUCERS"); red_env = (e && atoi(e) > 0) ? atoi(e) : 56; }
          f.xchg_reducer_ctas = red_env / 8 * 8;
          for (int par = 0; par < 2; ++par)
            for (int r = 0; r < t->world; ++r) f.xchg_grads_dev[par][r] = t->peer_grads[par][r];
        }
      }
      TRY(dmlp_mlp_fused_plan_create(&f, &t->ring_plan));
    }
  }
  return DMLP_OK;
}

static int enqueue_allreduce(Trainer* t, int par, size_t off, size_t count, cudaStream_t st) {
  if (t->comm == nullptr) {
    return dmlp_sgd(t->params + off, t->grads[par] + off, count, t->lr_dev, 1.f,
                    t->bf16 ? static_cast<void*>(t->shadow + off) : nullptr, st);
  }
  if (t->nvls) {
    dmlp_nvls_allreduce_desc_t n;
    memset(&n, 0, sizeof(n));
    n.count = count;
    n.mc_in = t->mc_grads[par] + off;
    n.mc_out = t->mc_params + off;
    n.mc_out_bf16 = t->bf16 ? static_cast<void*>(static_cast<__nv_bfloat16*>(t->mc_shadow) + off) : nullptr;
    n.uc_out = t->params + off;
    n.lr = t->lr_dev;
    n.scale = 1.f / static_cast<float>(t->world);
    n.fuse_sgd = 1;
    // small (latency-bound) gradients: one-shot pull through the switch; gradients are double-buffered by parity
    n.oneshot = (!t->bf16 && count * 4 <= (1u << 20)) ? 1 : 0;
    return dmlp_allreduce_nvls(t->comm, t->local_idx, &n, st);
  }
  dmlp_allreduce_desc_t d;
  memset(&d, 0, sizeof(d));
  d.algo = t->cfg.allreduce_algo;
  d.dtype = DMLP_F32;
  d.op = DMLP_SUM;
  d.fuse_sgd = 1;
  d.skip_end_barrier = 1;  // gradients are double-buffered by step parity
  d.scale = 1.f / static_cast<float>(t->world);
  d.count = count;
  for (int r = 0; r < t->world; ++r) {
    d.in[r] = t->peer_grads[par][r] ? t->peer_grads[par][r] + off : nullptr;
    d.out[r] = t->peer_params[r] ? t->peer_params[r] + off : nullptr;
    d.out_bf16[r] = (t->bf16 && t->peer_shadow[r]) ? static_cast<void*>(static_cast<__nv_bfloat16*>(t->peer_shadow[r]) + off) : nullptr;
  }
  d.lr[t->rank] = t->lr_dev;
  return dmlp_allreduce(t->comm, t->local_idx, &d, st);
}

// Enqueues one full training step for input parity `par` on the trainer's streams.
// Stream layout (captured as parallel graph branches):
//   s_main : fwd_1 .. fwd_L(+softmax-CE) -> dgrad_L -> ... -> dgrad_2 ........ -> all-reduce+SGD
//   s_aux  :                      wgrad_L (after dZ_L) ... wgrad_1 (after dZ_1) --^
//   s_comm : (overlap mode) bucket_l all-reduce+SGD once dgrad_l and wgrad_l are done
static int enqueue_step(Trainer* t, int par, bool count_kernels) {
  const ParamLayout& L = t->lay;
  const int B = t->cfg.batch;
  cudaStream_t S = t->s_main, A = t->s_aux;
  int nk = 0;
  const int last = L.L - 1;
  if (t->fused_step) {
    // one cluster launch produces every gradient, the loss and the hit count; then one kernel
    // reduces + averages + applies SGD (or plain SGD on a single GPU)
    TRY(dmlp_mlp_fused_step(t->fplan[par], S));
    ++nk;
    if (t->comm != nullptr || t->cfg.fused == 2) {  // single GPU: the fused kernel already applied SGD in place
      TRY(enqueue_allreduce(t, par, 0, L.total, S));
      ++nk;
    }
    if (count_kernels) t->kernels_per_step = nk;
    return DMLP_OK;
  }
  if (!(t->fused_ce && t->direct)) DMLP_CUDA_CHECK(cudaMemsetAsync(t->metrics_dev[par], 0, 8, S));
  if (!t->direct)
    for (int l = 0; l < L.L; ++l)
      DMLP_CUDA_CHECK(cudaMemsetAsync(t->grads[par] + L.b_off[l], 0, sizeof(float) * L.out_pad[l], S));
  if (t->bf16) { TRY(dmlp_cast_f32_to_bf16(t->x[par], t->x16, static_cast<size_t>(B) * L.in[0], S)); ++nk; }
  for (int l = 0; l < L.L; ++l) { TRY(gemm_plan_run(&t->fwd[par][l], S)); ++nk; }
  if (!t->fused_ce) {
    if (t->direct) DMLP_CUDA_CHECK(cudaMemsetAsync(t->grads[par] + L.b_off[last], 0, sizeof(float) * L.out_pad[last], S));
    TRY(dmlp_softmax_ce(t->act[last], L.out_pad[last], t->labels[par], B, L.out[last], 1.f / B,
                        t->bf16 ? nullptr : t->dz[last], t->bf16 ? static_cast<void*>(t->dz16[last]) : nullptr,
                        t->metrics_dev[par], reinterpret_cast<int32_t*>(t->metrics_dev[par] + 1),
                        t->grads[par] + L.b_off[last], S));
    ++nk;
  }
  const bool overlap = t->cfg.overlap != 0 && t->comm != nullptr;
  for (int l = last; l >= 0; --l) {
    // dZ_l is ready on S: the wgrad branch may start
    DMLP_CUDA_CHECK(cudaEventRecord(t->ev_dz[l], S));
    DMLP_CUDA_CHECK(cudaStreamWaitEvent(A, t->ev_dz[l], 0));
    if (!t->rs_on[l]) {  // (the pipelined layers run their wgrad last: see below)
      TRY(gemm_plan_run(&t->wgrad[par][l], A));
      ++nk;
    }
    if (l > 0) {
      GemmPlan pl = t->dgrad[l];
      pl.p.colsum = t->grads[par] + L.b_off[l - 1];
      TRY(gemm_plan_run(&pl, S));
      ++nk;
    }
    if (overlap && t->rs_on[l]) {
      // only the bias goes through the bucket all-reduce (its gradient came from the producer of dZ_l)
      const size_t hi = (l + 1 < L.L) ? L.w_off[l + 1] : L.total;
      DMLP_CUDA_CHECK(cudaStreamWaitEvent(t->s_comm, t->ev_dz[l], 0));
      TRY(enqueue_allreduce(t, par, L.b_off[l], hi - L.b_off[l], t->s_comm));
      ++nk;
    } else if (overlap) {
      // bucket l = [W_l | b_l]: needs wgrad_l (A) and dgrad_l (S, it reads W_l which the fused SGD rewrites);
      // b_l came from the producer of dZ_l, which precedes both.
      const size_t lo = L.w_off[l];
      const size_t hi = (l + 1 < L.L) ? L.w_off[l + 1] : L.total;
      DMLP_CUDA_CHECK(cudaEventRecord(t->ev_bucket[l], A));
      DMLP_CUDA_CHECK(cudaStreamWaitEvent(t->s_comm, t->ev_bucket[l], 0));
      if (l > 0) {
        DMLP_CUDA_CHECK(cudaEventRecord(t->ev_fork, S));
        DMLP_CUDA_CHECK(cudaStreamWaitEvent(t->s_comm, t->ev_fork, 0));
      }
      TRY(enqueue_allreduce(t, par, lo, hi - lo, t->s_comm));
      ++nk;
    }
  }
  if (overlap && t->rs_layers > 0) {
    // Pipelined layers: wgrad GEMM with the reduce-scatter epilogue and its consumer warps, after every dgrad of this step
    // (a pushed tile tells the peers that this rank no longer reads the layer's weights). The GEMMs of one step run one
    // after the other anyway (each fills the GPU); putting the big weight gradient last lets its whole exchange hide
    // behind its own tensor-core time.
    uint32_t* pads[DMLP_MAX_RANKS];
    uint32_t *state = nullptr, *host_flags = nullptr, dead = 0;
    unsigned long long timeout_ns = 0;
    int crank = 0, cworld = 0, cdev = 0;
    TRY(dmlp_comm_internal_view(t->comm, t->local_idx, pads, &state, &host_flags, &timeout_ns, &dead, &crank, &cworld, &cdev));
    DMLP_CUDA_CHECK(cudaEventRecord(t->ev_rs, S));
    DMLP_CUDA_CHECK(cudaStreamWaitEvent(A, t->ev_rs, 0));
    // ... and after every bucket all-reduce of this step has COMPLETED. Both kinds of kernel wait for their peers: if one
    // rank's GPU ran the all-reduce first (its CTAs spinning at the entry barrier, keeping some of the GEMM's persistent CTAs
    // off their SMs) while the other ran the GEMM first (all SMs taken, its all-reduce not yet scheduled), each side would
    // wait for a kernel the other cannot start. Kernels that wait for peers never overlap each other here.
    DMLP_CUDA_CHECK(cudaEventRecord(t->ev_rs_comm, t->s_comm));
    DMLP_CUDA_CHECK(cudaStreamWaitEvent(A, t->ev_rs_comm, 0));
    if (!((dead >> t->rank) & 1u)) {  // (a rank marked dead goes silent: the survivors' watchdogs report it)
      for (int l = last; l >= 0; --l) {
        if (!t->rs_on[l]) continue;
        TRY(gemm_plan_run(&t->wgrad[par][l], A));  // (GEMM + reduce-scatter + SGD + all-gather: one kernel)
        ++nk;
      }
    }
  }
  if (overlap) {
    DMLP_CUDA_CHECK(cudaEventRecord(t->ev_join, t->s_comm));
    DMLP_CUDA_CHECK(cudaStreamWaitEvent(S, t->ev_join, 0));
    DMLP_CUDA_CHECK(cudaEventRecord(t->ev_aux_done, A));
    DMLP_CUDA_CHECK(cudaStreamWaitEvent(S, t->ev_aux_done, 0));
  } else {
    DMLP_CUDA_CHECK(cudaEventRecord(t->ev_aux_done, A));
    DMLP_CUDA_CHECK(cudaStreamWaitEvent(S, t->ev_aux_done, 0));
    TRY(enqueue_allreduce(t, par, 0, L.total, S));
    ++nk;
  }
  DMLP_CUDA_CHECK(cudaMemcpyAsync(t->metrics_host + 2 * par, t->metrics_dev[par], 8, cudaMemcpyDeviceToHost, S));
  if (count_kernels) t->kernels_per_step = nk;
  return DMLP_OK;
}

static int capture_graphs(Trainer* t) {
  NvtxRange nvtx_cap("dmlp.capture_graphs");
  for (int par = 0; par < 2; ++par) {
    DMLP_CUDA_CHECK(cudaStreamBeginCapture(t->s_main, cudaStreamCaptureModeRelaxed));
    int rc = enqueue_step(t, par, false);
    cudaGraph_t g = nullptr;
    cudaError_t e = cudaStreamEndCapture(t->s_main, &g);
    if (rc) { if (g) cudaGraphDestroy(g); return rc; }
    if (e != cudaSuccess) return set_error(DMLP_ERR_CUDA, "cudaStreamEndCapture: %s", cudaGetErrorString(e));
    t->graph[par] = g;
    DMLP_CUDA_CHECK(cudaGraphInstantiate(&t->gexec[par], g, 0));
  }
  return DMLP_OK;
}

// Fused-path step loop. Inputs are staged chunk-wise into a device ring (one H2D per chunk when the
// batches are contiguous in the source), so the step kernels of a chunk are launched back to back
// with no event between them; loss / hit count arrive in mapped host memory, one slot per step.
static int run_fused_ring(Trainer* t, const float* x_host, const int32_t* y_host, size_t num_batches,
                          size_t first_batch, size_t batch_stride, int steps, const float* lr_schedule,
                          float* metrics_out, void* start_event, void* stop_event) {
  const int B = t->cfg.batch;
  const size_t xin = t->lay.in[0];
  const ParamLayout& L = t->lay;
  int chunk_first_step[kRingChunks] = {-1, -1}, chunk_n[kRingChunks] = {0, 0};
  int chunk_off_rec[kRingChunks] = {0, 0};
  auto collect = [&](int c) {
    if (chunk_first_step[c] < 0 || !metrics_out) return;
    for (int i = 0; i < chunk_n[c]; ++i) {
      const int slot = c * kRingChunk + chunk_off_rec[c] + i, s = chunk_first_step[c] + i;
      metrics_out[2 * s] = t->metrics_host[4 + 2 * slot];
      metrics_out[2 * s + 1] = static_cast<float>(reinterpret_cast<int*>(t->metrics_host)[4 + 2 * slot + 1]);
    }
    chunk_first_step[c] = -1;
  };
  unsigned long long gx_timeout_ns = 0;
  if (t->gx || t->pdp) {
    // the watchdog budget and the fault-injection step are read at every run (dmlp_comm_set_timeout_ms may have changed)
    uint32_t* pads[DMLP_MAX_RANKS];
    uint32_t *state = nullptr, *host_flags = nullptr, dead = 0;
    int crank = 0, cworld = 0, cdev = 0;
    TRY(dmlp_comm_internal_view(t->comm, t->local_idx, pads, &state, &host_flags, &gx_timeout_ns, &dead, &crank, &cworld, &cdev));
    long long fs = t->fault_step;
    if (fs < 0 && ((dead >> t->rank) & 1u)) fs = 0;  // dmlp_comm_inject_dead_ranks: dead from now on
    TRY(dmlp_mlp_fused_set_fault(t->ring_plan, fs, gx_timeout_ns));
  }
  if (start_event) DMLP_CUDA_CHECK(cudaEventRecord(reinterpret_cast<cudaEvent_t>(start_event), t->s_main));
  // Gradient buffers alternate with the GLOBAL step parity (slot parity == step parity), also across
  // calls: a call that starts on an odd step begins at slot 1 of its first chunk.
  int s0 = 0, ci = 0, last_c = -1;
  long long launches = 0;
  while (s0 < steps) {
    const int c = ci % kRingChunks;
    const int off = (ci == 0) ? static_cast<int>(t->steps_done & 1) : 0;
    const int n = steps - s0 < kRingChunk - off ? steps - s0 : kRingChunk - off;
    if (t->ring_launched[c]) {
      DMLP_CUDA_CHECK(cudaEventSynchronize(t->ev_ring_done[c]));
      collect(c);
    }
    const size_t b0 = (first_batch + static_cast<size_t>(s0) * batch_stride) % num_batches;
    float* xdst = t->x_ring + static_cast<size_t>(c * kRingChunk + off) * B * xin;
    int* ydst = t->y_ring + static_cast<size_t>(c * kRingChunk + off) * B;
    if (batch_stride == 1 && b0 + n <= num_batches) {
      DMLP_CUDA_CHECK(cudaMemcpyAsync(xdst, x_host + b0 * B * xin, sizeof(float) * n * B * xin, cudaMemcpyDefault, t->s_copy));
      DMLP_CUDA_CHECK(cudaMemcpyAsync(ydst, y_host + b0 * B, sizeof(int) * n * B, cudaMemcpyDefault, t->s_copy));
    } else {
      for (int i = 0; i < n; ++i) {
        const size_t bi = (first_batch + static_cast<size_t>(s0 + i) * batch_stride) % num_batches;
        DMLP_CUDA_CHECK(cudaMemcpyAsync(xdst + static_cast<size_t>(i) * B * xin, x_host + bi * B * xin, sizeof(float) * B * xin,
                                        cudaMemcpyDefault, t->s_copy));
        DMLP_CUDA_CHECK(cudaMemcpyAsync(ydst + static_cast<size_t>(i) * B, y_host + bi * B, sizeof(int) * B, cudaMemcpyDefault, t->s_copy));
      }
    }
    DMLP_CUDA_CHECK(cudaEventRecord(t->ev_ring_h2d[c], t->s_copy));
    DMLP_CUDA_CHECK(cudaStreamWaitEvent(t->s_main, t->ev_ring_h2d[c], 0));
    const unsigned long long g_chunk0 = t->gsteps, g_chunk1 = t->gsteps + static_cast<unsigned long long>(n);
    if (t->gx) {
      // all-gather the chunk's inputs: into the same ring slots of every peer, once the peers are done with the steps
      // that last used them; the step kernels wait for the peers' ready words (= g_chunk1) before reading peer inputs
      void* dsts[DMLP_MAX_RANKS];
      const size_t ring_off = static_cast<size_t>(t->rank) * kRingSlots * B * xin + static_cast<size_t>(c * kRingChunk + off) * B * xin;
      for (int r = 0; r < t->world; ++r) dsts[r] = t->peer_xall[r] + ring_off;
      TRY(dmlp_input_exchange(xdst, dsts, sizeof(float) * n * B * xin, t->gx_pads, t->rank, t->world,
                              static_cast<uint32_t>(t->ring_end[c]), static_cast<uint32_t>(g_chunk1), gx_timeout_ns, t->gx_status,
                              t->s_copy));
      t->ring_end[c] = g_chunk1;
    }
    static int per_launch_env = -1;  // experiment knob: cap on steps per launch (default: the whole chunk)
    if (per_launch_env < 0) { const char* e = getenv("DMLP_FUSED_STEPS_PER_LAUNCH"); per_launch_env = (e && atoi(e) > 0) ? atoi(e) : kRingChunk; }
    for (int i = 0; i < n;) {
      const int s = s0 + i;
      if (lr_schedule && lr_schedule[s] != t->lr_host[0]) {
        DMLP_CUDA_CHECK(cudaStreamSynchronize(t->s_main));
        t->lr_host[0] = lr_schedule[s];
        DMLP_CUDA_CHECK(cudaMemcpyAsync(t->lr_dev, t->lr_host, 4, cudaMemcpyHostToDevice, t->s_main));
      }
      const int slot = c * kRingChunk + off + i;
      if (t->ring_plan != nullptr) {
        // as many steps as share this learning rate, in one launch
        int m = 1;
        while (i + m < n && m < per_launch_env && (!lr_schedule || lr_schedule[s + m] == t->lr_host[0])) ++m;
        if (t->gx || t->pdp) {
          TRY(dmlp_mlp_fused_steps_gx(t->ring_plan, slot, m, g_chunk0 + static_cast<unsigned long long>(i),
                                      static_cast<unsigned int>(g_chunk1), t->s_main));
        } else {
          TRY(dmlp_mlp_fused_steps(t->ring_plan, slot, m, t->s_main));
        }
        ++launches;
        i += m;
      } else {
        TRY(dmlp_mlp_fused_step(t->rplan[slot], t->s_main));
        TRY(enqueue_allreduce(t, slot & 1, 0, L.total, t->s_main));
        launches += 2;
        i += 1;
      }
    }
    DMLP_CUDA_CHECK(cudaMemcpyAsync(t->metrics_host + 4 + 2 * (c * kRingChunk + off), t->metrics_ring + 2 * (c * kRingChunk + off),
                                    8 * n, cudaMemcpyDeviceToHost, t->s_main));
    DMLP_CUDA_CHECK(cudaEventRecord(t->ev_ring_done[c], t->s_main));
    t->ring_launched[c] = true;
    chunk_first_step[c] = s0;
    chunk_n[c] = n;
    chunk_off_rec[c] = off;
    last_c = c;
    s0 += n;
    ++ci;
    if (t->gx || t->pdp) t->gsteps = g_chunk1;
  }
  if (stop_event) DMLP_CUDA_CHECK(cudaEventRecord(reinterpret_cast<cudaEvent_t>(stop_event), t->s_main));
  DMLP_CUDA_CHECK(cudaStreamSynchronize(t->s_main));
  // drain in launch order
  if (last_c >= 0) { collect((last_c + 1) % kRingChunks); collect(last_c); }
  t->kernels_per_step = t->ring_plan != nullptr ? 1 : 2;
  t->launches_last_run = launches;
  t->steps_done += steps;
  if (t->comm) {
    uint32_t st[5];
    TRY(dmlp_comm_status(t->comm, st));
    if (st[0] != 0) {
      t->failed = true;  // sticky, like a FAILED communicator of the reference (gpu_coordinator_server.go:281-285)
      return set_error(DMLP_ERR_TIMEOUT, "collective watchdog: rank %u waited for rank %u (phase %u)", st[1], st[2], st[3]);
    }
  }
  return DMLP_OK;
}

}  // namespace dmlp

using namespace dmlp;

extern "C" int dmlp_trainer_param_count(const dmlp_trainer_config_t* cfg, size_t* floats) {
  DMLP_REQUIRE(cfg && floats && cfg->num_layers >= 1 && cfg->num_layers <= kMaxLayers, "bad trainer config");
  ParamLayout lay;
  make_layout(*cfg, &lay);
  *floats = lay.total;
  return DMLP_OK;
}

// Bytes of communicator arena the trainer needs: params + 2 gradient buffers (+ bf16 shadow).
extern "C" int dmlp_trainer_arena_bytes(const dmlp_trainer_config_t* cfg, size_t* bytes) {
  size_t n = 0;
  TRY(dmlp_trainer_param_count(cfg, &n));
  size_t b = 3 * align_up(n * 4, 256);
  if (cfg->dtype == DMLP_BF16) b += align_up(n * 2, 256);
  // fused step with in-kernel gradient exchange: receive slots [2 parities][up to 8 ranks][params]
  const bool fused_ok = cfg->dtype == DMLP_F32 && dmlp_mlp_fused_supported(cfg->num_layers, cfg->dims, cfg->batch);
  if (fused_ok && (cfg->fused == 3 || cfg->fused == 5)) b += 2 * DMLP_MAX_RANKS * align_up(n * 4, 256);
  // global-batch mode: the gathered input buffer [ranks][ring slots][batch][in0]
  if (fused_ok && cfg->fused == 5)
    b += align_up(static_cast<size_t>(DMLP_MAX_RANKS) * kRingSlots * cfg->batch * cfg->dims[0] * 4, 256);
  *bytes = b;
  return DMLP_OK;
}

extern "C" int dmlp_trainer_layout(const dmlp_trainer_config_t* cfg, int layer, size_t* w_off, size_t* b_off,
                                   int* ld) {
  DMLP_REQUIRE(cfg && layer >= 0 && layer < cfg->num_layers, "bad layer index");
  ParamLayout lay;
  make_layout(*cfg, &lay);
  if (w_off) *w_off = lay.w_off[layer];
  if (b_off) *b_off = lay.b_off[layer];
  if (ld) *ld = lay.out_pad[layer];
  return DMLP_OK;
}

extern "C" int dmlp_trainer_destroy(dmlp_trainer_t h) {
  Trainer* t = reinterpret_cast<Trainer*>(h);
  if (!t) return DMLP_OK;
  DeviceGuard g(t->dev);
  cudaDeviceSynchronize();
  for (int p = 0; p < 2; ++p) {
    if (t->gexec[p]) cudaGraphExecDestroy(t->gexec[p]);
    if (t->graph[p]) cudaGraphDestroy(t->graph[p]);
    if (t->ev_h2d[p]) cudaEventDestroy(t->ev_h2d[p]);
    if (t->ev_done[p]) cudaEventDestroy(t->ev_done[p]);
  }
  if (t->ev_join) cudaEventDestroy(t->ev_join);
  if (t->ev_fork) cudaEventDestroy(t->ev_fork);
  if (t->ev_aux_done) cudaEventDestroy(t->ev_aux_done);
  if (t->ev_rs) cudaEventDestroy(t->ev_rs);
  if (t->ev_rs_comm) cudaEventDestroy(t->ev_rs_comm);
  for (int l = 0; l < kMaxLayers; ++l)
    if (t->ev_dz[l]) cudaEventDestroy(t->ev_dz[l]);
  for (int l = 0; l < kMaxLayers; ++l)
    if (t->ev_bucket[l]) cudaEventDestroy(t->ev_bucket[l]);
  if (t->s_main) cudaStreamDestroy(t->s_main);
  if (t->s_comm) cudaStreamDestroy(t->s_comm);
  if (t->s_copy) cudaStreamDestroy(t->s_copy);
  if (t->s_aux) cudaStreamDestroy(t->s_aux);
  for (int p = 0; p < 2; ++p)
    if (t->fplan[p]) dmlp_mlp_fused_plan_destroy(t->fplan[p]);
  for (int sl = 0; sl < kRingSlots; ++sl)
    if (t->rplan[sl]) dmlp_mlp_fused_plan_destroy(t->rplan[sl]);
  if (t->ring_plan) dmlp_mlp_fused_plan_destroy(t->ring_plan);
  for (int c = 0; c < kRingChunks; ++c) {
    if (t->ev_ring_h2d[c]) cudaEventDestroy(t->ev_ring_h2d[c]);
    if (t->ev_ring_done[c]) cudaEventDestroy(t->ev_ring_done[c]);
  }
  for (void* p : t->owned) cudaFree(p);
  if (t->metrics_host) cudaFreeHost(t->metrics_host);
  if (t->lr_host) cudaFreeHost(t->lr_host);
  cudaGetLastError();
  delete t;
  return DMLP_OK;
}

extern "C" int dmlp_trainer_create(const dmlp_trainer_config_t* cfg, int device, dmlp_comm_t comm, int local_idx,
                                   int rank, int world, size_t arena_offset, dmlp_trainer_t* out) {
  DMLP_REQUIRE(cfg && out, "dmlp_trainer_create: null argument");
  DMLP_REQUIRE(cfg->num_layers >= 1 && cfg->num_layers <= kMaxLayers, "num_layers %d not in [1,%d]", cfg->num_layers, kMaxLayers);
  DMLP_REQUIRE(cfg->batch >= 1, "batch must be positive");
  DMLP_REQUIRE(cfg->dtype == DMLP_F32 || cfg->dtype == DMLP_BF16, "dtype must be F32 (tf32 math) or BF16");
  DMLP_REQUIRE(cfg->dims[0] % 8 == 0, "input width %d must be a multiple of 8 (TMA row pitch)", cfg->dims[0]);
  for (int l = 1; l < cfg->num_layers; ++l)
    DMLP_REQUIRE(cfg->dims[l] % 16 == 0, "hidden width %d must be a multiple of 16", cfg->dims[l]);
  DMLP_REQUIRE(world >= 1 && world <= DMLP_MAX_RANKS && rank >= 0 && rank < world, "bad rank/world");
  DMLP_REQUIRE((world == 1) || comm != nullptr, "world > 1 needs a communicator");
  Trainer* t = new (std::nothrow) Trainer();
  if (!t) return set_error(DMLP_ERR_CUDA, "out of host memory");
  memset(t->ev_h2d, 0, sizeof(t->ev_h2d));
  memset(t->rplan, 0, sizeof(t->rplan));
  memset(t->ev_ring_h2d, 0, sizeof(t->ev_ring_h2d));
  memset(t->ev_ring_done, 0, sizeof(t->ev_ring_done));
  memset(t->ev_done, 0, sizeof(t->ev_done));
  memset(t->ev_bucket, 0, sizeof(t->ev_bucket));
  t->ev_fork = t->ev_join = nullptr;
  memset(t->act, 0, sizeof(t->act));
  memset(t->act16, 0, sizeof(t->act16));
  memset(t->dz, 0, sizeof(t->dz));
  memset(t->dz16, 0, sizeof(t->dz16));
  memset(t->peer_params, 0, sizeof(t->peer_params));
  memset(t->peer_shadow, 0, sizeof(t->peer_shadow));
  memset(t->peer_grads, 0, sizeof(t->peer_grads));
  memset(t->peer_recv, 0, sizeof(t->peer_recv));
  memset(t->peer_xall, 0, sizeof(t->peer_xall));
  memset(t->gx_pads, 0, sizeof(t->gx_pads));
  t->cfg = *cfg;
  t->dev = device;
  t->comm = (world > 1) ? comm : nullptr;
  t->local_idx = local_idx;
  t->rank = rank;
  t->world = world;
  t->bf16 = cfg->dtype == DMLP_BF16;
  make_layout(*cfg, &t->lay);
  const ParamLayout& L = t->lay;
  const int B = cfg->batch;
  t->fused_ce = L.out_pad[L.L - 1] <= 32;
  t->fused_step = !t->bf16 && cfg->fused != 0 && dmlp_mlp_fused_supported(cfg->num_layers, cfg->dims, cfg->batch) != 0;
  if ((cfg->fused == 1 || cfg->fused == 3 || cfg->fused == 4 || cfg->fused == 5) && !t->fused_step) {
    set_error(DMLP_ERR_UNSUPPORTED, "fused step kernel does not support this shape / precision");
    delete t;
    return DMLP_ERR_UNSUPPORTED;
  }
  t->direct = B <= 128;
  memset(t->ev_dz, 0, sizeof(t->ev_dz));
  t->ev_aux_done = nullptr;
  DeviceGuard guard(device);
  int rc = DMLP_OK;
  auto fail = [&](int code) { dmlp_trainer_destroy(reinterpret_cast<dmlp_trainer_t>(t)); return code; };
#define TRYF(expr) do { rc = (expr); if (rc) return fail(rc); } while (0)
#define CUF(expr) do { cudaError_t _e = (expr); if (_e != cudaSuccess) { set_error(DMLP_ERR_CUDA, "%s: %s", #expr, cudaGetErrorString(_e)); return fail(DMLP_ERR_CUDA); } } while (0)

  const size_t pbytes = align_up(L.total * 4, 256);
  const size_t sbytes = align_up(L.total * 2, 256);
  t->nvls = t->comm != nullptr && cfg->allreduce_algo == DMLP_ALGO_NVLS;
  if (t->nvls) {
    const size_t need = 3 * pbytes + (t->bf16 ? sbytes : 0);
    if (!cfg->nvls_uc || !cfg->nvls_mc || cfg->nvls_bytes < need) {
      set_error(DMLP_ERR_INVALID, "NVLS all-reduce needs a multicast-bound arena of at least %zu bytes", need);
      return fail(DMLP_ERR_INVALID);
    }
    uint8_t* u = static_cast<uint8_t*>(cfg->nvls_uc);
    uint8_t* m = static_cast<uint8_t*>(cfg->nvls_mc);
    t->params = reinterpret_cast<float*>(u);
    t->grads[0] = reinterpret_cast<float*>(u + pbytes);
    t->grads[1] = reinterpret_cast<float*>(u + 2 * pbytes);
    t->mc_params = reinterpret_cast<float*>(m);
    t->mc_grads[0] = reinterpret_cast<float*>(m + pbytes);
    t->mc_grads[1] = reinterpret_cast<float*>(m + 2 * pbytes);
    if (t->bf16) {
      t->shadow = reinterpret_cast<__nv_bfloat16*>(u + 3 * pbytes);
      t->mc_shadow = m + 3 * pbytes;
    }
  } else if (t->comm) {
    for (int r = 0; r < world; ++r) {
      void* base = nullptr;
      size_t cap = 0;
      int e = dmlp_comm_arena_ptr(t->comm, local_idx, r, &base, &cap);
      if (e != DMLP_OK) { if (r == rank) return fail(e); continue; }  // (a peer that is not mapped: dmlp_allreduce will refuse later)
      const size_t xall_bytes = align_up(static_cast<size_t>(world) * kRingSlots * B * L.in[0] * 4, 256);
      const size_t base_need = arena_offset + 3 * pbytes + (t->bf16 ? sbytes : 0);
      const size_t recv_bytes = 2 * static_cast<size_t>(world) * pbytes;
      const bool want_g = t->fused_step && cfg->fused == 5;
      const bool want_x = t->fused_step && (cfg->fused == 3 || want_g);
      size_t need = base_need + (want_x ? recv_bytes : 0) + (want_g ? xall_bytes : 0);
      if (cap < need) { set_error(DMLP_ERR_INVALID, "communicator arena too small: %zu < %zu", cap, need); return fail(DMLP_ERR_INVALID); }
      uint8_t* b = static_cast<uint8_t*>(base) + arena_offset;
      t->peer_params[r] = reinterpret_cast<float*>(b);
      t->peer_grads[0][r] = reinterpret_cast<float*>(b + pbytes);
      t->peer_grads[1][r] = reinterpret_cast<float*>(b + 2 * pbytes);
      if (t->bf16) t->peer_shadow[r] = b + 3 * pbytes;
      if (want_x) t->peer_recv[r] = reinterpret_cast<float*>(b + 3 * pbytes);
      if (want_g) t->peer_xall[r] = reinterpret_cast<float*>(b + 3 * pbytes + recv_bytes);
    }
    // Persistent data-parallel launches (compute cluster + reducer CTAs in one kernel, several steps per launch) are the
    // default for the cluster-fused shapes on real multi-GPU communicators; fused == 1 keeps the two-kernel chain.
    int emulated = 0;
    TRYF(dmlp_comm_is_emulated(t->comm, &emulated));
    t->emulated = emulated != 0;
    t->pdp = t->fused_step && !emulated && (cfg->fused == 4 || cfg->fused == -1);
    for (int r = 0; r < world; ++r) t->pdp = t->pdp && t->peer_grads[0][r] != nullptr;
    t->xchg = t->fused_step && cfg->fused == 3;
    for (int r = 0; r < world; ++r) t->xchg = t->xchg && t->peer_recv[r] != nullptr;
    // Global-batch mode (fused == 5, opt-in): measured against the persistent launch in profiles/r2_dp_modes_scaling.md
    // (level at 2 GPUs, slower from 4 GPUs on: its per-peer staging is serial inside each CTA).
    t->gx = t->fused_step && !emulated && cfg->fused == 5;
    for (int r = 0; r < world; ++r) t->gx = t->gx && t->peer_recv[r] != nullptr && t->peer_xall[r] != nullptr;
    if (cfg->fused == 5 && !t->gx) {
      set_error(DMLP_ERR_UNSUPPORTED, "global-batch mode needs one GPU per rank and every rank's arena mapped");
      return fail(DMLP_ERR_UNSUPPORTED);
    }
    if (t->gx) t->pdp = false;
    t->params = t->peer_params[rank];
    t->grads[0] = t->peer_grads[0][rank];
    t->grads[1] = t->peer_grads[1][rank];
    t->shadow = reinterpret_cast<__nv_bfloat16*>(t->peer_shadow[rank]);
  } else {
    void* p = nullptr;
    TRYF(dev_alloc(t, pbytes, &p)); t->params = static_cast<float*>(p);
    TRYF(dev_alloc(t, pbytes, &p)); t->grads[0] = static_cast<float*>(p);
    TRYF(dev_alloc(t, pbytes, &p)); t->grads[1] = static_cast<float*>(p);
    if (t->bf16) { TRYF(dev_alloc(t, sbytes, &p)); t->shadow = static_cast<__nv_bfloat16*>(p); }
  }
  void* p = nullptr;
  for (int par = 0; par < 2; ++par) {
    TRYF(dev_alloc(t, static_cast<size_t>(B) * L.in[0] * 4, &p)); t->x[par] = static_cast<float*>(p);
    TRYF(dev_alloc(t, static_cast<size_t>(B) * 4, &p)); t->labels[par] = static_cast<int*>(p);
    TRYF(dev_alloc(t, 16, &p)); t->metrics_dev[par] = static_cast<float*>(p);
  }
  if (t->bf16) { TRYF(dev_alloc(t, static_cast<size_t>(B) * L.in[0] * 2, &p)); t->x16 = static_cast<__nv_bfloat16*>(p); }
  for (int l = 0; l < L.L; ++l) {
    const size_t n = static_cast<size_t>(B) * L.out_pad[l];
    const bool last = l == L.L - 1;
    if (!t->bf16 || last) { TRYF(dev_alloc(t, n * 4, &p)); t->act[l] = static_cast<float*>(p); }
    if (t->bf16 && !last) { TRYF(dev_alloc(t, n * 2, &p)); t->act16[l] = static_cast<__nv_bfloat16*>(p); }
    if (t->bf16) { TRYF(dev_alloc(t, n * 2, &p)); t->dz16[l] = static_cast<__nv_bfloat16*>(p); }
    else { TRYF(dev_alloc(t, n * 4, &p)); t->dz[l] = static_cast<float*>(p); }
  }
  TRYF(dev_alloc(t, 16, &p)); t->lr_dev = static_cast<float*>(p);
  CUF(cudaHostAlloc(reinterpret_cast<void**>(&t->metrics_host), 1024, cudaHostAllocPortable | cudaHostAllocMapped));
  CUF(cudaHostAlloc(reinterpret_cast<void**>(&t->lr_host), 64, cudaHostAllocPortable));
  memset(t->metrics_host, 0, 512);
  t->lr_host[0] = cfg->lr;
  CUF(cudaMemcpy(t->lr_dev, t->lr_host, 4, cudaMemcpyHostToDevice));
  CUF(cudaStreamCreateWithFlags(&t->s_main, cudaStreamNonBlocking));
  CUF(cudaStreamCreateWithFlags(&t->s_comm, cudaStreamNonBlocking));
  CUF(cudaStreamCreateWithFlags(&t->s_copy, cudaStreamNonBlocking));
  CUF(cudaStreamCreateWithFlags(&t->s_aux, cudaStreamNonBlocking));
  for (int par = 0; par < 2; ++par) {
    CUF(cudaEventCreateWithFlags(&t->ev_h2d[par], cudaEventDisableTiming));
    CUF(cudaEventCreateWithFlags(&t->ev_done[par], cudaEventDisableTiming));
  }
  CUF(cudaEventCreateWithFlags(&t->ev_join, cudaEventDisableTiming));
  CUF(cudaEventCreateWithFlags(&t->ev_fork, cudaEventDisableTiming));
  CUF(cudaEventCreateWithFlags(&t->ev_aux_done, cudaEventDisableTiming));
  CUF(cudaEventCreateWithFlags(&t->ev_rs, cudaEventDisableTiming));
  CUF(cudaEventCreateWithFlags(&t->ev_rs_comm, cudaEventDisableTiming));
  for (int l = 0; l < L.L; ++l) CUF(cudaEventCreateWithFlags(&t->ev_dz[l], cudaEventDisableTiming));
  for (int l = 0; l < L.L; ++l) CUF(cudaEventCreateWithFlags(&t->ev_bucket[l], cudaEventDisableTiming));
  if (t->fused_step) {
    if (t->gx) {
      // own ring = this rank's part of its gathered input buffer; receive areas start out "empty" (0xFF bytes), the
      // exchange words of the signal pad at zero. Python barriers across the ranks before the first step.
      t->x_ring = t->peer_xall[rank] + static_cast<size_t>(rank) * kRingSlots * B * L.in[0];
      CUF(cudaMemset(t->peer_recv[rank], 0xFF, 2 * static_cast<size_t>(world) * pbytes));
      uint32_t* pads[DMLP_MAX_RANKS];
      uint32_t *state = nullptr, *host_flags = nullptr, dead = 0;
      unsigned long long timeout_ns = 0;
      int crank = 0, cworld = 0, cdev = 0;
      TRYF(dmlp_comm_internal_view(t->comm, t->local_idx, pads, &state, &host_flags, &timeout_ns, &dead, &crank, &cworld, &cdev));
      CUF(cudaMemset(pads[rank] + kPadXReady, 0, 4 * DMLP_MAX_RANKS * kXPushCtas));
      CUF(cudaMemset(pads[rank] + kPadXDone, 0, 4 * DMLP_MAX_RANKS));
      CUF(cudaDeviceSynchronize());
    } else {
      TRYF(dev_alloc(t, static_cast<size_t>(kRingSlots) * B * L.in[0] * 4, &p)); t->x_ring = static_cast<float*>(p);
    }
    TRYF(dev_alloc(t, static_cast<size_t>(kRingSlots) * B * 4, &p)); t->y_ring = static_cast<int*>(p);
    TRYF(dev_alloc(t, static_cast<size_t>(kRingSlots) * 8, &p)); t->metrics_ring = static_cast<float*>(p);
    for (int c = 0; c < kRingChunks; ++c) {
      CUF(cudaEventCreateWithFlags(&t->ev_ring_h2d[c], cudaEventDisableTiming));
      CUF(cudaEventCreateWithFlags(&t->ev_ring_done[c], cudaEventDisableTiming));
    }
  }
  TRYF(build_plans(t));
  *out = reinterpret_cast<dmlp_trainer_t>(t);
  return DMLP_OK;
#undef TRYF
#undef CUF
}

extern "C" int dmlp_trainer_params(dmlp_trainer_t h, float** params_dev, size_t* count) {
  Trainer* t = reinterpret_cast<Trainer*>(h);
  DMLP_REQUIRE(t, "null trainer");
  if (params_dev) *params_dev = t->params;
  if (count) *count = t->lay.total;
  return DMLP_OK;
}

// Call after the parameters were (re)written from the host: refreshes the bf16 shadow.
extern "C" int dmlp_trainer_params_updated(dmlp_trainer_t h) {
  Trainer* t = reinterpret_cast<Trainer*>(h);
  DMLP_REQUIRE(t, "null trainer");
  DeviceGuard g(t->dev);
  if (t->bf16) TRY(dmlp_cast_f32_to_bf16(t->params, t->shadow, t->lay.total, t->s_main));
  DMLP_CUDA_CHECK(cudaStreamSynchronize(t->s_main));
  return DMLP_OK;
}

extern "C" int dmlp_trainer_set_lr(dmlp_trainer_t h, float lr) {
  Trainer* t = reinterpret_cast<Trainer*>(h);
  DMLP_REQUIRE(t, "null trainer");
  DeviceGuard g(t->dev);
  DMLP_CUDA_CHECK(cudaStreamSynchronize(t->s_main));
  t->lr_host[0] = lr;
  DMLP_CUDA_CHECK(cudaMemcpyAsync(t->lr_dev, t->lr_host, 4, cudaMemcpyHostToDevice, t->s_main));
  DMLP_CUDA_CHECK(cudaStreamSynchronize(t->s_main));
  return DMLP_OK;
}

// Fault injection (BASELINE config 5): this rank stops taking part in the exchange from global step `step` on
// (counted from the trainer's creation; -1 = never), as if its GPU had died mid-run. Its peers' in-kernel watchdogs
// fire after the communicator's timeout, dmlp_trainer_run returns DMLP_ERR_TIMEOUT there and the trainer turns FAILED.
// Mirrors the reference's device-failure test (gpu_coordinator_server_test.go:370-429).
extern "C" int dmlp_trainer_set_fault(dmlp_trainer_t h, long long step) {
  Trainer* t = reinterpret_cast<Trainer*>(h);
  DMLP_REQUIRE(t, "null trainer");
  t->fault_step = step;
  return DMLP_OK;
}

// 0 = usable, 1 = FAILED (sticky). mode: 0 single GPU / per-layer, 1 two-kernel chain, 3 in-cluster exchange,
// 4 persistent compute + reducer launch, 5 global-batch mode.
extern "C" int dmlp_trainer_state(dmlp_trainer_t h, int* failed, int* mode) {
  Trainer* t = reinterpret_cast<Trainer*>(h);
  DMLP_REQUIRE(t, "null trainer");
  if (failed) *failed = t->failed ? 1 : 0;
  if (mode) *mode = t->gx ? 5 : t->pdp ? 4 : t->xchg ? 3 : (t->fused_step && t->comm) ? 1 : 0;
  return DMLP_OK;
}

// Layers whose weight gradient runs through the wgrad -> reduce-scatter -> SGD -> all-gather pipeline (wgrad_rs.cuh).
extern "C" int dmlp_trainer_pipelined_layers(dmlp_trainer_t h, int* n) {
  Trainer* t = reinterpret_cast<Trainer*>(h);
  DMLP_REQUIRE(t && n, "bad arguments");
  *n = t->rs_layers;
  return DMLP_OK;
}

extern "C" int dmlp_trainer_launches_last_run(dmlp_trainer_t h, long long* n) {
  Trainer* t = reinterpret_cast<Trainer*>(h);
  DMLP_REQUIRE(t && n, "bad arguments");
  *n = t->launches_last_run;
  return DMLP_OK;
}

extern "C" int dmlp_trainer_kernels_per_step(dmlp_trainer_t h, int* n) {
  Trainer* t = reinterpret_cast<Trainer*>(h);
  DMLP_REQUIRE(t && n, "null argument");
  *n = t->kernels_per_step;
  return DMLP_OK;
}

// Runs `steps` training steps. Batch for step s is rows [(first_batch + s*batch_stride) mod
// num_batches] * batch of x_host/y_host (pinned host memory gives true async copies; the pointers
// may also name a device-resident dataset — the copy direction is inferred — which is how the
// kernel-only timing streams inputs from a dataset larger than L2). start/stop events (nullable)
// are recorded on the compute stream around the loop. Per-step
// {loss_sum, correct} pairs are written to metrics_out (2*steps floats; correct as float) if non-null.
// lr_schedule (nullable, `steps` entries) is applied before each step.
extern "C" int dmlp_trainer_run(dmlp_trainer_t h, const float* x_host, const int32_t* y_host, size_t num_batches,
                                size_t first_batch, size_t batch_stride, int steps, const float* lr_schedule,
                                float* metrics_out, int use_graph, void* start_event, void* stop_event) {
  Trainer* t = reinterpret_cast<Trainer*>(h);
  DMLP_REQUIRE(t && x_host && y_host && num_batches > 0 && steps >= 0, "dmlp_trainer_run: bad arguments");
  NvtxRange nvtx_run("dmlp.trainer_run");
  if (t->failed)
    return set_error(DMLP_ERR_ABORTED, "the trainer's communicator is FAILED (a collective watchdog fired earlier); create a new one");
  DeviceGuard g(t->dev);
  const int B = t->cfg.batch;
  const size_t xin = t->lay.in[0];
  if (t->fused_step)
    return run_fused_ring(t, x_host, y_host, num_batches, first_batch, batch_stride, steps, lr_schedule, metrics_out,
                          start_event, stop_event);
  if (use_graph && t->gexec[0] == nullptr) {
    // One eager warm-up pass per parity sets function attributes / loads modules outside capture.
    // It must not change the model: run it with lr = 0.
    // (Collectives stay matched across ranks: every rank performs the same two extra steps.)
    float saved = t->lr_host[0];
    t->lr_host[0] = 0.f;
    DMLP_CUDA_CHECK(cudaMemcpy(t->lr_dev, t->lr_host, 4, cudaMemcpyHostToDevice));
    // Parities continue the alternation of the real steps (gradient buffers are double-buffered by step parity and the
    // one-shot all-reduce skips its trailing rendezvous: reusing the parity of the step just finished would overwrite a
    // buffer a slower peer may still be reading).
    for (int i = 0; i < 2; ++i) {
      const int par = static_cast<int>((t->steps_done + i) & 1);
      DMLP_CUDA_CHECK(cudaMemcpyAsync(t->x[par], x_host, sizeof(float) * B * xin, cudaMemcpyDefault, t->s_main));
      DMLP_CUDA_CHECK(cudaMemcpyAsync(t->labels[par], y_host, sizeof(int) * B, cudaMemcpyDefault, t->s_main));
      TRY(enqueue_step(t, par, true));
    }
    DMLP_CUDA_CHECK(cudaStreamSynchronize(t->s_main));
    t->lr_host[0] = saved;
    DMLP_CUDA_CHECK(cudaMemcpy(t->lr_dev, t->lr_host, 4, cudaMemcpyHostToDevice));
    TRY(capture_graphs(t));
  }
  int pending[2] = {-1, -1};  // step index whose metrics sit in metrics_host[par]
  if (start_event) DMLP_CUDA_CHECK(cudaEventRecord(reinterpret_cast<cudaEvent_t>(start_event), t->s_main));
  for (int s = 0; s < steps; ++s) {
    const int par = static_cast<int>((t->steps_done + s) & 1);
    if (t->launched[par]) {
      DMLP_CUDA_CHECK(cudaEventSynchronize(t->ev_done[par]));
      if (pending[par] >= 0 && metrics_out) {
        metrics_out[2 * pending[par]] = t->metrics_host[2 * par];
        metrics_out[2 * pending[par] + 1] = static_cast<float>(reinterpret_cast<int*>(t->metrics_host)[2 * par + 1]);
      }
    }
    const size_t bidx = (first_batch + static_cast<size_t>(s) * batch_stride) % num_batches;
    DMLP_CUDA_CHECK(cudaMemcpyAsync(t->x[par], x_host + bidx * B * xin, sizeof(float) * B * xin,
                                    cudaMemcpyDefault, t->s_copy));
    DMLP_CUDA_CHECK(cudaMemcpyAsync(t->labels[par], y_host + bidx * B, sizeof(int) * B, cudaMemcpyDefault, t->s_copy));
    DMLP_CUDA_CHECK(cudaEventRecord(t->ev_h2d[par], t->s_copy));
    DMLP_CUDA_CHECK(cudaStreamWaitEvent(t->s_main, t->ev_h2d[par], 0));
    if (lr_schedule && lr_schedule[s] != t->lr_host[0]) {
      // lr_host is read by the async copy: wait for the previous update to land first
      DMLP_CUDA_CHECK(cudaStreamSynchronize(t->s_main));
      t->lr_host[0] = lr_schedule[s];
      DMLP_CUDA_CHECK(cudaMemcpyAsync(t->lr_dev, t->lr_host, 4, cudaMemcpyHostToDevice, t->s_main));
    }
    if (use_graph) {
      DMLP_CUDA_CHECK(cudaGraphLaunch(t->gexec[par], t->s_main));
    } else {
      TRY(enqueue_step(t, par, true));
    }
    DMLP_CUDA_CHECK(cudaEventRecord(t->ev_done[par], t->s_main));
    t->launched[par] = true;
    pending[par] = s;
  }
  if (stop_event) DMLP_CUDA_CHECK(cudaEventRecord(reinterpret_cast<cudaEvent_t>(stop_event), t->s_main));
  DMLP_CUDA_CHECK(cudaStreamSynchronize(t->s_main));
  for (int par = 0; par < 2; ++par) {
    if (pending[par] >= 0 && metrics_out) {
      metrics_out[2 * pending[par]] = t->metrics_host[2 * par];
      metrics_out[2 * pending[par] + 1] = static_cast<float>(reinterpret_cast<int*>(t->metrics_host)[2 * par + 1]);
    }
  }
  t->steps_done += steps;
  if (t->comm) {
    uint32_t st[5];
    TRY(dmlp_comm_status(t->comm, st));
    if (st[0] != 0) {
      t->failed = true;  // sticky: further runs are refused with DMLP_ERR_ABORTED
      return set_error(DMLP_ERR_TIMEOUT, "collective watchdog: rank %u waited for rank %u (phase %u)", st[1], st[2], st[3]);
    }
  }
  return DMLP_OK;
}

// Device-resident variant used for kernel-only timing: inputs already sit in the trainer's two
// input buffers (filled by dmlp_trainer_load_inputs); no H2D/D2H happens inside the loop.
extern "C" int dmlp_trainer_load_inputs(dmlp_trainer_t h, int parity, const float* x_host, const int32_t* y_host) {
  Trainer* t = reinterpret_cast<Trainer*>(h);
  DMLP_REQUIRE(t && x_host && y_host && (parity == 0 || parity == 1), "bad arguments");
  DeviceGuard g(t->dev);
  DMLP_CUDA_CHECK(cudaMemcpy(t->x[parity], x_host, sizeof(float) * t->cfg.batch * t->lay.in[0], cudaMemcpyHostToDevice));
  DMLP_CUDA_CHECK(cudaMemcpy(t->labels[parity], y_host, sizeof(int) * t->cfg.batch, cudaMemcpyHostToDevice));
  return DMLP_OK;
}

extern "C" int dmlp_trainer_run_resident(dmlp_trainer_t h, int steps, int use_graph, void* start_event, void* stop_event) {
  Trainer* t = reinterpret_cast<Trainer*>(h);
  DMLP_REQUIRE(t && steps >= 0, "bad arguments");
  DeviceGuard g(t->dev);
  if (use_graph && t->gexec[0] == nullptr) {
    float saved = t->lr_host[0];
    t->lr_host[0] = 0.f;
    DMLP_CUDA_CHECK(cudaMemcpy(t->lr_dev, t->lr_host, 4, cudaMemcpyHostToDevice));
    for (int i = 0; i < 2; ++i) TRY(enqueue_step(t, static_cast<int>((t->steps_done + i) & 1), true));  // (see dmlp_trainer_run)
    DMLP_CUDA_CHECK(cudaStreamSynchronize(t->s_main));
    t->lr_host[0] = saved;
    DMLP_CUDA_CHECK(cudaMemcpy(t->lr_dev, t->lr_host, 4, cudaMemcpyHostToDevice));
    TRY(capture_graphs(t));
  }
  if (start_event) DMLP_CUDA_CHECK(cudaEventRecord(reinterpret_cast<cudaEvent_t>(start_event), t->s_main));
  for (int s = 0; s < steps; ++s) {
    const int par = static_cast<int>((t->steps_done + s) & 1);
    if (use_graph) DMLP_CUDA_CHECK(cudaGraphLaunch(t->gexec[par], t->s_main));
    else TRY(enqueue_step(t, par, true));
  }
  if (stop_event) DMLP_CUDA_CHECK(cudaEventRecord(reinterpret_cast<cudaEvent_t>(stop_event), t->s_main));
  DMLP_CUDA_CHECK(cudaStreamSynchronize(t->s_main));
  t->steps_done += steps;
  return DMLP_OK;
}

// Forward-only evaluation over n_batches full batches (reference testModel, client.go:462-502).
extern "C" int dmlp_trainer_eval(dmlp_trainer_t h, const float* x_host, const int32_t* y_host, size_t n_batches,
                                 double* loss_sum, long long* correct) {
  Trainer* t = reinterpret_cast<Trainer*>(h);
  DMLP_REQUIRE(t && x_host && y_host, "bad arguments");
  NvtxRange nvtx_eval("dmlp.trainer_eval");
  DeviceGuard g(t->dev);
  const ParamLayout& L = t->lay;
  const int B = t->cfg.batch;
  const int last = L.L - 1;
  double ls = 0;
  long long cs = 0;
  DMLP_CUDA_CHECK(cudaStreamSynchronize(t->s_main));
  for (size_t b = 0; b < n_batches; ++b) {
    cudaStream_t S = t->s_main;
    DMLP_CUDA_CHECK(cudaMemcpyAsync(t->x[0], x_host + b * B * L.in[0], sizeof(float) * B * L.in[0], cudaMemcpyHostToDevice, S));
    DMLP_CUDA_CHECK(cudaMemcpyAsync(t->labels[0], y_host + b * B, sizeof(int) * B, cudaMemcpyHostToDevice, S));
    DMLP_CUDA_CHECK(cudaMemsetAsync(t->metrics_dev[0], 0, 8, S));
    if (t->bf16) TRY(dmlp_cast_f32_to_bf16(t->x[0], t->x16, static_cast<size_t>(B) * L.in[0], S));
    for (int l = 0; l < L.L; ++l) {
      if (l == last && t->fused_ce) TRY(gemm_plan_run(&t->fwd_eval_last, S));
      else TRY(gemm_plan_run(&t->fwd[0][l], S));
    }
    if (!t->fused_ce)
      TRY(dmlp_softmax_ce(t->act[last], L.out_pad[last], t->labels[0], B, L.out[last], 1.f / B, nullptr, nullptr,
                          t->metrics_dev[0], reinterpret_cast<int32_t*>(t->metrics_dev[0] + 1), nullptr, S));
    DMLP_CUDA_CHECK(cudaMemcpyAsync(t->metrics_host, t->metrics_dev[0], 8, cudaMemcpyDeviceToHost, S));
    DMLP_CUDA_CHECK(cudaStreamSynchronize(S));
    ls += t->metrics_host[0];
    cs += reinterpret_cast<int*>(t->metrics_host)[1];
  }
  if (loss_sum) *loss_sum = ls;
  if (correct) *correct = cs;
  return DMLP_OK;
}

// Debug/test access to intermediate buffers: kind 0 = activation l (fp32), 1 = dZ l (fp32),
// 2 = gradient buffer `l` (parity), 3 = logits.
extern "C" int dmlp_trainer_buffer(dmlp_trainer_t h, int kind, int l, void** ptr) {
  Trainer* t = reinterpret_cast<Trainer*>(h);
  DMLP_REQUIRE(t && ptr, "bad arguments");
  switch (kind) {
    case 0: *ptr = t->act[l]; break;
    case 1: *ptr = t->dz[l]; break;
    case 2: *ptr = t->grads[l & 1]; break;
    case 3: *ptr = t->act[t->lay.L - 1]; break;
    case 9: *ptr = t->fused_prof; break;  // null unless DMLP_FUSED_PROF=1
    default: return set_error(DMLP_ERR_INVALID, "unknown buffer kind %d", kind);
  }
  return DMLP_OK;
}
