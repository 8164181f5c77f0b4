/*
 * dmlp_b200 — C ABI of the B200-native (sm_100a) collective runtime and MLP kernels.
 *
 * Every entry point returns 0 on success and a non-zero code on failure;
 * dmlp_last_error() returns a thread-local human-readable description of the last failure.
 * Pointers named *_dev are CUDA device pointers; `stream` is a cudaStream_t passed as void*.
 *
 * The reference implementation this library replaces is a Go gRPC simulator
 * (Helenbzbz/Distributed-Machine-Learning-Pipeline). Each section cites the reference
 * file:line whose behaviour the entry points provide on real GPUs.
 */
#ifndef DMLP_B200_H_
#define DMLP_B200_H_

#include <stddef.h>
#include <stdint.h>

#ifdef __cplusplus
extern "C" {
#endif

#define DMLP_OK 0
#define DMLP_ERR_INVALID 1   /* bad argument (maps to gRPC InvalidArgument) */
#define DMLP_ERR_CUDA 2      /* CUDA runtime / driver failure (gRPC Internal) */
#define DMLP_ERR_NOT_FOUND 3 /* unknown handle / slot (gRPC NotFound) */
#define DMLP_ERR_TIMEOUT 4   /* collective watchdog fired: a peer never arrived */
#define DMLP_ERR_ABORTED 5   /* communicator was aborted / is FAILED (FailedPrecondition) */
#define DMLP_ERR_UNSUPPORTED 6

#define DMLP_MAX_RANKS 8

const char* dmlp_last_error(void);
/* Library build description: arch, version, compile flags. */
const char* dmlp_build_info(void);
/* Debug aid: per-tile timeline (8 globaltimer stamps per tile, first 64 tiles) of CTA 0 of the last GEMM whose
 * plan was created with DMLP_GEMM_PROF=1 in the environment. See csrc/gemm_tcgen05.cu for the slot meanings. */
int dmlp_gemm_prof_read(unsigned long long* host, int n_words);

/* ------------------------------------------------------------------ data types / ops */
typedef enum { DMLP_F32 = 0, DMLP_BF16 = 1, DMLP_U8 = 2, DMLP_I32 = 3, DMLP_F16 = 4 } dmlp_dtype_t;
/* Same numbering as gpu_sim.ReduceOp (reference DSML/proto/gpu_sim.proto:162-168); AVG is the
 * extension the reference left commented out. */
typedef enum { DMLP_SUM = 0, DMLP_PROD = 1, DMLP_MIN = 2, DMLP_MAX = 3, DMLP_AVG = 4 } dmlp_op_t;

/* ------------------------------------------------------------------ tcgen05 GEMM
 * D[M,N] = epilogue(A[M,K] * B[K,N]) on the 5th-gen tensor cores (tcgen05.mma, accumulators in
 * TMEM, operands staged by TMA with 128-byte swizzle). Covers the three GEMM forms of the MLP:
 *   forward  relu(X*W+b)      reference DSML/client/client.go:112-141
 *   dgrad    dZ*W^T (.) mask  reference DSML/client/client.go:178-189
 *   wgrad    A^T*dZ           reference DSML/client/client.go:167-176,191-199
 * Operand "major": 0 = K-major (the K index is contiguous in memory), 1 = MN-major (the M index
 * of A, resp. the N index of B, is contiguous). Weights keep the reference's [in][out] row-major
 * layout, which is MN-major for forward, K-major for dgrad; no transposes are materialised.
 */
typedef struct {
  int dtype;          /* DMLP_F32 (tf32 tensor-core math on fp32 storage) or DMLP_BF16 */
  int a_major;        /* 0: A stored [M][lda] ; 1: A stored [K][lda] (M contiguous) */
  int b_major;        /* 0: B stored [N][ldb] ; 1: B stored [K][ldb] (N contiguous) */
  int M, N, K;
  const void* A_dev;
  int64_t lda;        /* leading dimensions in elements */
  const void* B_dev;
  int64_t ldb;
  float* out_f32_dev; /* nullable */
  int64_t ld_out;
  void* out_bf16_dev; /* nullable: bf16 copy of the result (activation / weight shadow) */
  int64_t ld_out_bf16;
  const float* bias_dev; /* nullable, length N, added before activation */
  int relu;              /* apply max(x,0) */
  const void* mask_dev;  /* nullable: result is zeroed where mask[m][n] <= 0 (ReLU backward) */
  int mask_dtype;        /* DMLP_F32 or DMLP_BF16 */
  int64_t ld_mask;
  float* colsum_dev;     /* nullable, length N: column sums of the result (bias gradient): stored when
                            M <= 128 (one M tile sees all rows), else atomically accumulated */
  int accumulate;        /* out_f32 += result instead of overwrite */
  int block_n;           /* N tile (multiple of 16, <= 256); 0 = choose automatically */
  /* Fused softmax cross-entropy epilogue (reference client.go:76-93,144-165,352-382), enabled by a
   * non-null ce_labels_dev; needs ce_classes <= N <= 32. The GEMM result + bias are the logits; the
   * kernel writes dlogits = (softmax - onehot) * ce_scale to out_f32/out_bf16 (zero in padded
   * columns), the bias gradient to colsum_dev, sum of -log(p_true + 1e-10) to ce_loss_sum_dev and the
   * number of argmax hits to ce_correct_dev. With M <= 128 these reductions are stored (no prior
   * zeroing needed), otherwise accumulated atomically. ce_logits_dev (nullable) receives the logits. */
  const int32_t* ce_labels_dev;
  int ce_classes;
  float ce_scale;
  float* ce_logits_dev;
  float* ce_loss_sum_dev;
  int32_t* ce_correct_dev;
} dmlp_gemm_desc_t;

int dmlp_gemm(const dmlp_gemm_desc_t* desc, void* stream);

/* ------------------------------------------------------------------ fused softmax cross-entropy
 * reference DSML/client/client.go:76-93,144-165 (softmax, loss, dLogits) and :352-382 (accuracy).
 * logits [B][ld] fp32 (ld >= C); labels int32 [B]. Writes dlogits[b][c] = (p - onehot) * scale for
 * c < C and 0 for C <= c < ld; adds sum_b -log(p_true + 1e-10) to loss_sum[0] and the number of
 * argmax hits to correct[0]; optionally adds column sums of dlogits to colsum (bias gradient).
 */
int dmlp_softmax_ce(const float* logits_dev, int64_t ld, const int32_t* labels_dev, int B, int C,
                    float scale, float* dlogits_dev, void* dlogits_bf16_dev, float* loss_sum_dev,
                    int32_t* correct_dev, float* colsum_dev, void* stream);

/* ------------------------------------------------------------------ SGD (single-GPU path / oracle)
 * reference DSML/client/client.go:254-267. w -= lr * scale * g, lr read from device memory so a
 * host-side schedule works under CUDA graphs. Optionally refreshes a bf16 shadow of w.
 */
int dmlp_sgd(float* w_dev, const float* g_dev, size_t n, const float* lr_dev, float scale,
             void* w_bf16_dev, void* stream);

/* Fill / convert helpers used by the framework layer. */
int dmlp_fill_u32(void* dst_dev, uint32_t value, size_t n_words, void* stream);
int dmlp_cast_f32_to_bf16(const float* src_dev, void* dst_bf16_dev, size_t n, void* stream);

/* ------------------------------------------------------------------ devices, memory, streams
 * The real-GPU counterpart of the reference's simulated device memory
 * (DSML/gpu_device_service/gpu_device_server.go:26-49) and of CommInit's per-device connection
 * set-up (DSML/gpu_coordinator_service/gpu_coordinator_server.go:131-164). */
typedef struct {
  char name[128];
  char uuid[48];
  int sm_count;
  int cc_major, cc_minor;
  int clock_khz;
  int multicast_supported; /* NVLS: cuMulticast* usable on this device */
  int smem_per_block_optin;
  size_t total_mem;
  size_t l2_bytes;
} dmlp_device_info_t;

int dmlp_device_count(int* n);
int dmlp_device_info(int dev, dmlp_device_info_t* out);
/* n x n matrices (row-major): peer access possible, native atomics, performance rank. */
int dmlp_peer_matrix(int n, const int* devs, int* access, int* atomics, int* perf_rank);
int dmlp_enable_peer_access(int n, const int* devs);

int dmlp_malloc(int dev, size_t bytes, void** ptr_dev); /* zero-initialised */
int dmlp_free(int dev, void* ptr_dev);
int dmlp_host_alloc(size_t bytes, void** ptr_host);    /* pinned, portable, mapped */
int dmlp_host_free(void* ptr_host);
/* stream == NULL: synchronous copy. Reference Memcpy H2D/D2H: gpu_device_server.go:195-230. */
int dmlp_memcpy_h2d(int dev, void* dst_dev, const void* src_host, size_t bytes, void* stream);
int dmlp_memcpy_d2h(int dev, void* dst_host, const void* src_dev, size_t bytes, void* stream);
int dmlp_memcpy_d2d(int dev, void* dst_dev, const void* src_dev, size_t bytes, void* stream);
int dmlp_memset(int dev, void* dst_dev, int value, size_t bytes, void* stream);

int dmlp_stream_create(int dev, void** stream);
int dmlp_stream_destroy(int dev, void* stream);
int dmlp_stream_sync(int dev, void* stream);
int dmlp_stream_query(int dev, void* stream, int* busy);
int dmlp_stream_wait_event(int dev, void* stream, void* event);
int dmlp_device_sync(int dev);
int dmlp_event_create(int dev, int timing, void** event);
int dmlp_event_destroy(void* event);
int dmlp_event_record(int dev, void* event, void* stream);
int dmlp_event_query(void* event, int* busy); /* GetStreamStatus: gpu_device_server.go:183-193 */
int dmlp_event_sync(void* event);
int dmlp_event_elapsed_ms(void* start, void* stop, float* ms);

/* CUDA IPC: lets one-process-per-GPU jobs map each other's buffers. */
int dmlp_ipc_get_handle(int dev, void* ptr_dev, unsigned char out[64]);
int dmlp_ipc_open_handle(int dev, const unsigned char in[64], void** ptr_dev);
int dmlp_ipc_close_handle(int dev, void* ptr_dev);

/* ------------------------------------------------------------------ communicator + collectives
 * Reference: communicator registry and CommInit (gpu_coordinator_server.go:25-52,121-192), ring
 * all-reduce (:272-566), naive all-reduce (:611-717), device failure handling (:91-119).
 *
 * A communicator spans `world` ranks; `nlocal` of them are driven by the calling process
 * (all of them in the single-process coordinator; exactly one under torchrun). Each rank owns a
 * peer-visible segment: [signal pad][ring staging][arena]. Other processes' segments are mapped
 * with dmlp_comm_export / dmlp_comm_import. If every rank is local and on the SAME device the
 * communicator is "emulated": collectives run as one cooperative launch with gridDim.y = world.
 */
typedef struct dmlp_comm_opaque* dmlp_comm_t;

int dmlp_comm_create(int world, int nlocal, const int* local_ranks, const int* local_devs,
                     size_t staging_bytes, size_t arena_bytes, dmlp_comm_t* out);
int dmlp_comm_destroy(dmlp_comm_t comm);
int dmlp_comm_export(dmlp_comm_t comm, int local_idx, unsigned char out[64]);
int dmlp_comm_import(dmlp_comm_t comm, int rank, const unsigned char in[64]);
/* Base of rank `rank`'s arena as addressable from local rank `local_idx`. */
int dmlp_comm_arena_ptr(dmlp_comm_t comm, int local_idx, int rank, void** ptr_dev, size_t* bytes);
int dmlp_comm_set_timeout_ms(dmlp_comm_t comm, uint64_t ms); /* in-kernel watchdog, default 10 s */
int dmlp_comm_abort(dmlp_comm_t comm);                       /* releases every spinning kernel */
/* out = {failure code, reporting rank, awaited peer, phase, abort flag}; code 0 = healthy. */
int dmlp_comm_is_emulated(dmlp_comm_t comm, int* emulated);  /* all ranks on one device (single cooperative launches) */
int dmlp_comm_status(dmlp_comm_t comm, uint32_t out[5]);
int dmlp_comm_inject_dead_ranks(dmlp_comm_t comm, uint32_t rank_mask); /* fault injection */

typedef enum {
  DMLP_ALGO_AUTO = 0,
  DMLP_ALGO_ONESHOT = 1,
  DMLP_ALGO_TWOSHOT = 2,
  DMLP_ALGO_RING = 3,
  DMLP_ALGO_NVLS = 4
} dmlp_algo_t;

typedef struct {
  int algo;  /* dmlp_algo_t */
  int dtype; /* dmlp_dtype_t */
  int op;    /* dmlp_op_t */
  int fuse_sgd;         /* out (weights) -= lr * scale * sum(in) instead of out = reduce(in); f32 SUM */
  int skip_end_barrier; /* one-shot only: caller double-buffers `in`, drop the trailing rendezvous */
  float scale;          /* 0 is read as 1; AVG multiplies by 1/world on top */
  size_t count;         /* elements */
  /* Per-rank buffers as addressable from the launching GPU, 16-byte aligned. one-shot needs every
   * in[]; two-shot every in[] and out[]; ring only the caller's own in/out (it moves data through
   * the communicator's staging). */
  const void* in[DMLP_MAX_RANKS];
  void* out[DMLP_MAX_RANKS];
  void* out_bf16[DMLP_MAX_RANKS];  /* optional bf16 shadow of the updated weights (fuse_sgd) */
  const float* lr[DMLP_MAX_RANKS]; /* device scalar per rank (fuse_sgd) */
  void* streams[DMLP_MAX_RANKS];   /* local_idx < 0: per-local-rank streams (NULL = `stream` arg) */
  void* prof_dev;                  /* optional: [blocks][4] u64 globaltimer stamps {entry, after barrier-in,
                                      after data, exit} written by thread 0 of each CTA (one-shot kernel) */
} dmlp_allreduce_desc_t;

/* Launches the all-reduce for local rank `local_idx` (or for every local rank when < 0) and
 * returns immediately. Failures of peers surface through dmlp_comm_status. */
int dmlp_allreduce(dmlp_comm_t comm, int local_idx, const dmlp_allreduce_desc_t* desc, void* stream);
int dmlp_allreduce_pick_algo(int world, size_t bytes);
/* Same with NVLS in the table: nvls_available = both buffers live in multicast-bound memory (dmlp_nvls_*), then
 * DMLP_ALGO_NVLS is returned from the size on at which the in-switch reduction wins (run it with dmlp_allreduce_nvls). */
int dmlp_allreduce_pick_algo_ex(int world, size_t bytes, int nvls_available);
/* Overrides rows of that table for this process: "world:oneshot_max:nvls_min[,world:...]", byte counts with an optional
 * K / M / G suffix, "never" = NVLS is not picked; NULL or "" restores the built-in table (+ the DMLP_ALLREDUCE_TABLE
 * environment variable, same syntax, read at first use). */
int dmlp_allreduce_set_table(const char* spec);

/* P2P push copy kernel (StreamSend on real GPUs, gpu_device_server.go:112-181): launched on
 * src_dev, dst may be peer memory. */
int dmlp_p2p_copy(int src_dev, void* dst_dev, const void* src_dev_ptr, size_t bytes, void* stream);
/* dst = reduce(srcs[0..n)) on one device (root step of NaiveAllReduce, :681-686). */
int dmlp_reduce_n(int dev, void* dst_dev, const void* const* srcs_dev, int n, size_t count, int dtype,
                  int op, void* stream);

/* ------------------------------------------------------------------ NVLS (NVSwitch multicast) all-reduce
 * The multicast variant of AllReduceRing (gpu_coordinator_server.go:272-566): rank r reduces slice r
 * with multimem.ld_reduce (the switch sums all replicas) and broadcasts the result / the updated
 * weights with multimem.st. Buffers must live in memory bound to one multicast object:
 *   creator:  dmlp_nvls_create (returns a POSIX fd for other processes)   others: dmlp_nvls_import
 *   everyone: dmlp_nvls_add_device for its device(s)  -- barrier --  dmlp_nvls_bind  -- barrier --
 * Barriers and the watchdog reuse the communicator's signal pads. f32 SUM (+ fused average/SGD). */
typedef struct dmlp_nvls_opaque* dmlp_nvls_t;
int dmlp_nvls_supported(int dev, int* supported);
int dmlp_nvls_create(int world, size_t bytes, int* fd_out, dmlp_nvls_t* out);
int dmlp_nvls_import(int world, size_t bytes, int fd, dmlp_nvls_t* out);
int dmlp_nvls_size(dmlp_nvls_t nvls, size_t* bytes);
int dmlp_nvls_add_device(dmlp_nvls_t nvls, int dev);
int dmlp_nvls_bind(dmlp_nvls_t nvls, int dev, void** uc_ptr_dev, void** mc_ptr_dev);
int dmlp_nvls_destroy(dmlp_nvls_t nvls);
typedef struct {
  size_t count;          /* floats */
  const void* mc_in;     /* multicast address of the inputs (gradients) */
  void* mc_out;          /* multicast address of the outputs (weights when fuse_sgd) */
  void* mc_out_bf16;     /* optional multicast address of the bf16 weight shadow */
  const void* uc_out;    /* caller's own replica of the outputs (read by the fused SGD) */
  const float* lr;       /* device scalar (fuse_sgd) */
  float scale;
  int fuse_sgd;
  int oneshot;           /* 1: latency-bound flavour - every rank pulls the switch-reduced sum of all elements into
                          * its own replica (uc_out required, no bf16 shadow); no broadcast, one rendezvous; the
                          * caller must double-buffer the inputs across consecutive calls */
} dmlp_nvls_allreduce_desc_t;
int dmlp_allreduce_nvls(dmlp_comm_t comm, int local_idx, const dmlp_nvls_allreduce_desc_t* desc, void* stream);

/* ------------------------------------------------------------------ all-reduce without allocation
 * (No counterpart in the reference: its coordinator allocates freely on the host.)
 *
 * out[r][i] = sum over ranks of in[r'][i] for `world` logical ranks whose buffers all live on the
 * CURRENT device, using the ring kernel — but with every byte of working memory (signal pads,
 * sequence state, watchdog words, staging slots) taken from a fixed array compiled into the
 * library, so the call performs no cudaMalloc / cudaFree / cudaHostAlloc and creates no
 * communicator. It exists for callers that must not allocate at call time: a process that has
 * already handed the whole GPU to its own memory pool (a late cudaMalloc may fail), or a loop
 * that cannot tolerate the device-wide synchronisation cudaMalloc and cudaFree imply.
 *
 *  - in_dev / out_dev are HOST arrays of `world` device pointers, 16-byte aligned; out may alias in.
 *  - Limit: count <= world * 65536 floats (one 256 KiB staging slot per rank per step); larger
 *    counts are rejected with DMLP_ERR_INVALID — use dmlp_allreduce for those. world <= 8.
 *  - Cost: about 4.1 MiB of device memory on every GPU that loads the library, held until exit.
 *  - The first call on each device looks up the array's device address; make that first call
 *    outside of any stream capture. Later calls only enqueue a memset and one kernel on `stream`.
 *  - All calls on a device share the one array (the control words are cleared at the start of
 *    each call). Two calls whose GPU execution overlaps corrupt each other: keep them on one
 *    stream or order them with events.
 */
int dmlp_allreduce_sum_f32_noalloc(int world, const float* const* in_dev, float* const* out_dev,
                                   size_t count, void* stream);
/* Watchdog record {code, rank, peer, phase} of the most recent dmlp_allreduce_sum_f32_noalloc on the current device
 * (it lives in the static array, out of the host's sight). Call after synchronising that call's stream: out[0] == 0
 * means it completed; DMLP_ERR_TIMEOUT means a rank gave up after 10 s and `out_dev` holds garbage. One small
 * synchronous device-to-host copy, no allocation. Make the first call on a device (either entry point) at start-up,
 * before the GPU's memory is handed to a pool and outside stream capture: with lazy module loading that first call is
 * what materialises the 4.1 MiB array. No counterpart in the reference. */
int dmlp_allreduce_noalloc_status(uint32_t out[4]);

/* ------------------------------------------------------------------ cluster-fused MLP step
 * Forward + softmax-CE + backward of one batch of the reference-sized MLP in ONE launch of an 8-CTA
 * thread-block cluster (K split across CTAs, partial sums exchanged through distributed shared
 * memory, every product a tcgen05.mma). Reference: DSML/client/client.go:112-202.
 * Shapes: batch <= 64 (x8), dims[0] % 8 == 0 and <= 832, dims[1] == 128, optional dims[2] == 64,
 * classes <= 16. Writes all gradients (flat parameter layout), loss sum and hit count. */
typedef struct {
  int num_layers;            /* 2 or 3 */
  int dims[4];
  int batch;
  const float* x_dev;        /* [batch][dims[0]] */
  const int32_t* labels_dev;
  const float* params_dev;   /* flat parameters (see trainer layout) */
  float* grads_dev;          /* flat gradients, same layout */
  size_t w_off[3], b_off[3]; /* in floats */
  float* metrics_dev;        /* [0] loss sum (float), [1] hit count (int32) */
  float* debug_dev;          /* optional dump of intermediates */
  void* prof_dev;            /* optional [8][16] u64 globaltimer stamps at the phase boundaries */
  float* sgd_params_dev;     /* non-null (single GPU): apply w -= lr*g in place (== params_dev) and write no gradients */
  const float* lr_dev;       /* device scalar, with sgd_params_dev */
  int x_slots;               /* 0/1: x_dev is one batch. n > 1: x_dev / labels_dev / metrics_dev are rings of n slots
                              * ([n][batch][dims[0]], [n][batch], [n][2]) addressed by dmlp_mlp_fused_steps */
  /* Data-parallel mode (xchg_world > 1, needs sgd_params_dev): the kernel itself exchanges the gradients. Every rank
   * pushes its gradient slices over NVLink into slot [step parity][its rank] of every rank's receive area
   * ([2][world][param_count] floats, peer-mapped), raises one flag per CTA in the peers' signal pads, waits for the
   * peers' flags, sums the slots in rank order and applies w -= lr/world * sum. No separate all-reduce kernel. */
  int xchg_world, xchg_rank;
  float* xchg_recv_dev[DMLP_MAX_RANKS];     /* rank r's receive area as addressable from this device (own rank: local) */
  uint32_t* xchg_pad_dev[DMLP_MAX_RANKS];   /* rank r's communicator signal pad */
  uint32_t* xchg_status;                    /* host-mapped {code, rank, peer, phase} failure record (nullable) */
  unsigned long long xchg_timeout_ns;       /* watchdog for the flag wait */
  size_t param_count;                       /* floats in the flat parameter vector (slot size) */
  /* Persistent data-parallel mode (xchg_reducer_ctas > 0, multiple of 8, <= 136; all 8 + n CTAs must be co-resident): the launch is one compute cluster plus
   * that many reducer CTAs. The compute cluster writes gradients to xchg_grads_dev[parity][rank]; the reducers run the
   * one-shot all-reduce + average + SGD over peer memory inside the same launch (local flags both ways), so several
   * steps per launch work on several GPUs too. xchg_recv_dev is not used in this mode. */
  int xchg_reducer_ctas;
  float* xchg_grads_dev[2][DMLP_MAX_RANKS];
  /* Global-batch mode (gx_enable != 0 with xchg_world > 1 and xchg_reducer_ctas == 0; the default on real GPUs): the
   * ranks exchange the FACTORS of the layer-1 weight gradient instead of the gradient: inputs are all-gathered ahead of
   * time (dmlp_input_exchange; x_dev is then [world][x_slots][batch][dims[0]], own ring at index xchg_rank), every CTA
   * pushes its dZ1 slice into all peers' receive areas (xchg_recv_dev, pre-filled with 0xFF bytes) and each rank
   * accumulates dW1 = sum_r X_r^T dZ1_r itself in rank order; W2/W3/biases are all-reduced in the kernel by push +
   * rank-ordered sum. Arrival is detected on the data, no flags. Launch with dmlp_mlp_fused_steps_gx. */
  int gx_enable;
} dmlp_mlp_fused_desc_t;
typedef struct dmlp_mlp_fused_opaque* dmlp_mlp_fused_plan_t;
int dmlp_mlp_fused_supported(int num_layers, const int* dims, int batch);
int dmlp_mlp_fused_plan_create(const dmlp_mlp_fused_desc_t* desc, dmlp_mlp_fused_plan_t* out);
/* Global-batch mode launches: global_step = index of the first step (same on every rank), inputs_ready = value the
 * peers' input-ready words must have reached. dmlp_mlp_fused_set_fault: fault injection (rank goes silent from that
 * global step on; -1 = off) and watchdog budget (0 keeps it). Reference for the failure semantics:
 * DSML/gpu_coordinator_service/gpu_coordinator_server.go:339-346. */
int dmlp_mlp_fused_steps_gx(dmlp_mlp_fused_plan_t plan, int slot0, int nsteps, unsigned long long global_step,
                            unsigned int inputs_ready, void* stream);
int dmlp_mlp_fused_set_fault(dmlp_mlp_fused_plan_t plan, long long step, unsigned long long timeout_ns);
/* Input all-gather for the global-batch mode (csrc/input_exchange.cu): writes `bytes` (x16) from src_local into the
 * same place of every peer's gathered input buffer (dst_peers[r], r != rank, peer-mapped) once all peers' "done"
 * words in this rank's signal pad reached need_done, then publishes ready_value in the peers' "ready" words.
 * No counterpart in the reference (inputs never leave the client there, DSML/client/client.go:596-608). */
int dmlp_input_exchange(const void* src_local, void* const* dst_peers, size_t bytes, uint32_t* const* pads, int rank,
                        int world, uint32_t need_done, uint32_t ready_value, unsigned long long timeout_ns,
                        uint32_t* status, void* stream);
/* Runs `nsteps` consecutive optimisation steps on ring slots slot0 .. slot0+nsteps-1 in ONE launch (plan created with
 * x_slots > 1 and sgd_params_dev: the weights are updated in place between the steps inside the kernel). */
int dmlp_mlp_fused_steps(dmlp_mlp_fused_plan_t plan, int slot0, int nsteps, void* stream);
int dmlp_mlp_fused_plan_destroy(dmlp_mlp_fused_plan_t plan);
int dmlp_mlp_fused_step(dmlp_mlp_fused_plan_t plan, void* stream);

/* ------------------------------------------------------------------ native data-parallel MLP trainer
 * Reference: the client's model, step and epoch loop (DSML/client/client.go:22-58 model/init,
 * :112-202 forward/backward, :254-267 SGD, :462-502 eval, :516-659 training loop).
 *
 * Parameters live in one flat fp32 buffer, per layer W[in][out_pad] then b[out_pad] (out_pad = out
 * rounded up to 16, padding kept at zero; every tensor starts on a 256-byte boundary). This is the
 * reference's W1|b1|W2|b2(|W3|b3) order; the unpadded little-endian byte layout the reference
 * uploads to MemAddr 0x2000 (client.go:60-66,204-218) is produced/consumed by the Python
 * checkpoint helpers. With a communicator the buffer sits in the symmetric arena so peers can
 * read gradients / write updated weights directly.
 */
#define DMLP_MAX_LAYERS 8
typedef struct {
  int num_layers;
  int dims[DMLP_MAX_LAYERS + 1]; /* e.g. 784,128,64,10 */
  int batch;                     /* per-GPU batch */
  int dtype;                     /* DMLP_F32: tf32 tensor-core math, fp32 storage; DMLP_BF16 */
  float lr;
  int allreduce_algo;            /* dmlp_algo_t */
  int overlap;                   /* 1: per-layer buckets reduced on a second stream during backward */
  void* nvls_uc;                 /* allreduce_algo == DMLP_ALGO_NVLS: this rank's replica of the multicast-bound */
  void* nvls_mc;                 /* arena and its multicast address; params/grads/shadow are placed there */
  size_t nvls_bytes;
  int fused;                     /* cluster-fused step kernel: -1 auto (when the shape allows), 0 off, 1 on,
                                    2 on but always write gradients + run the separate update kernel,
                                    3 (world > 1) on with the gradient exchange INSIDE the step kernel: bulk-copy push into
                                      the peers' receive slots, rank-ordered reduce + SGD; one kernel per step, several
                                      steps per launch. Measured slower than 1 + all-reduce kernel at this model size (only
                                      8 SMs drive NVLink, profiles/r1_fused_exchange.md): opt-in */
} dmlp_trainer_config_t;
typedef struct dmlp_trainer_opaque* dmlp_trainer_t;

int dmlp_trainer_param_count(const dmlp_trainer_config_t* cfg, size_t* floats);
int dmlp_trainer_arena_bytes(const dmlp_trainer_config_t* cfg, size_t* bytes);
int dmlp_trainer_layout(const dmlp_trainer_config_t* cfg, int layer, size_t* w_off, size_t* b_off, int* ld);
int dmlp_trainer_create(const dmlp_trainer_config_t* cfg, int device, dmlp_comm_t comm, int local_idx,
                        int rank, int world, size_t arena_offset, dmlp_trainer_t* out);
int dmlp_trainer_destroy(dmlp_trainer_t t);
int dmlp_trainer_params(dmlp_trainer_t t, float** params_dev, size_t* count);
int dmlp_trainer_params_updated(dmlp_trainer_t t);
int dmlp_trainer_set_lr(dmlp_trainer_t t, float lr);
int dmlp_trainer_kernels_per_step(dmlp_trainer_t t, int* n);
/* Kernel launches issued by the most recent dmlp_trainer_run on the fused path (a launch may cover several steps);
 * 0 when the per-layer path ran (use kernels_per_step * steps there). */
int dmlp_trainer_launches_last_run(dmlp_trainer_t t, long long* n);
/* Fault injection on the training path (BASELINE config 5; reference test: gpu_coordinator_server_test.go:370-429):
 * this rank goes silent from global step `step` on (-1: never). dmlp_trainer_state: failed = 1 once a collective
 * watchdog fired (sticky: dmlp_trainer_run then returns DMLP_ERR_ABORTED, like AllReduceRing on a FAILED communicator,
 * gpu_coordinator_server.go:281-285); mode = data-parallel variant in use (5 = global-batch mode). */
int dmlp_trainer_set_fault(dmlp_trainer_t trainer, long long step);
int dmlp_trainer_state(dmlp_trainer_t trainer, int* failed, int* mode);
/* Layers whose weight gradient is exchanged by the GEMM's reduce-scatter epilogue and its consumer kernel (multi-GPU,
 * overlap on, big layers; DMLP_WGRAD_RS=0 disables, DMLP_WGRAD_RS_MIN_KB sets the size threshold, default 8192). */
int dmlp_trainer_pipelined_layers(dmlp_trainer_t trainer, int* n);
int dmlp_trainer_run(dmlp_trainer_t t, const float* x_host, const int32_t* y_host, size_t num_batches,
                     size_t first_batch, size_t batch_stride, int steps, const float* lr_schedule,
                     float* metrics_out, int use_graph, void* start_event, void* stop_event);
int dmlp_trainer_load_inputs(dmlp_trainer_t t, int parity, const float* x_host, const int32_t* y_host);
int dmlp_trainer_run_resident(dmlp_trainer_t t, int steps, int use_graph, void* start_event, void* stop_event);
int dmlp_trainer_eval(dmlp_trainer_t t, const float* x_host, const int32_t* y_host, size_t n_batches,
                      double* loss_sum, long long* correct);
int dmlp_trainer_buffer(dmlp_trainer_t t, int kind, int index, void** ptr_dev);

#ifdef __cplusplus
}
#endif
#endif /* DMLP_B200_H_ */
