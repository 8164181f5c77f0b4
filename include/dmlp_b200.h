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
  float* colsum_dev;     /* nullable, length N: += column sums of the result (bias gradient) */
  int accumulate;        /* out_f32 += result instead of overwrite */
  int block_n;           /* N tile (multiple of 16, <= 256); 0 = choose automatically */
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

#ifdef __cplusplus
}
#endif
#endif /* DMLP_B200_H_ */
