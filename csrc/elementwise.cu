// Non-GEMM device kernels of the MLP step: fused softmax cross-entropy forward+backward,
// standalone SGD (single-GPU path; the multi-GPU path fuses SGD into the all-reduce),
// fills and casts. Reference: DSML/client/client.go:76-93,144-165,254-267,352-382.
#include "common.h"

#include <cuda_bf16.h>
#include <string.h>

namespace dmlp {

static thread_local char g_err[1024] = "";

int set_error(int code, const char* fmt, ...) {
  va_list ap;
  va_start(ap, fmt);
  vsnprintf(g_err, sizeof(g_err), fmt, ap);
  va_end(ap);
  return code;
}

// One warp per row. C (number of classes) <= ld <= 32*kMaxPerLane.
constexpr int kCeMaxPerLane = 8;  // up to 256 classes

__global__ void softmax_ce_kernel(const float* __restrict__ logits, long long ld,
                                  const int* __restrict__ labels, int B, int C, float scale,
                                  float* __restrict__ dlogits, __nv_bfloat16* __restrict__ dlogits_bf16,
                                  float* __restrict__ loss_sum, int* __restrict__ correct,
                                  float* __restrict__ colsum) {
  const int warps_per_block = blockDim.x >> 5;
  const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
  float loss_acc = 0.f;
  int hit_acc = 0;
  float cs[kCeMaxPerLane];
#pragma unroll
  for (int i = 0; i < kCeMaxPerLane; ++i) cs[i] = 0.f;

  for (int b = blockIdx.x * warps_per_block + warp; b < B; b += gridDim.x * warps_per_block) {
    const float* row = logits + static_cast<long long>(b) * ld;
    float x[kCeMaxPerLane];
    float mx = -INFINITY;
    int arg = 0x7fffffff;
#pragma unroll
    for (int i = 0; i < kCeMaxPerLane; ++i) {
      const int c = lane + 32 * i;
      x[i] = (c < C) ? row[c] : -INFINITY;
      if (x[i] > mx) { mx = x[i]; arg = c; }
    }
    // warp arg-max (first maximal index wins, like the reference's strict '>' scan)
#pragma unroll
    for (int o = 16; o > 0; o >>= 1) {
      const float omx = __shfl_xor_sync(0xffffffffu, mx, o);
      const int oarg = __shfl_xor_sync(0xffffffffu, arg, o);
      if (omx > mx || (omx == mx && oarg < arg)) { mx = omx; arg = oarg; }
    }
    float e[kCeMaxPerLane];
    float sum = 0.f;
#pragma unroll
    for (int i = 0; i < kCeMaxPerLane; ++i) {
      const int c = lane + 32 * i;
      e[i] = (c < C) ? __expf(x[i] - mx) : 0.f;
      sum += e[i];
    }
#pragma unroll
    for (int o = 16; o > 0; o >>= 1) sum += __shfl_xor_sync(0xffffffffu, sum, o);
    const float inv = 1.f / sum;
    const int y = labels[b];
    float* drow = dlogits ? dlogits + static_cast<long long>(b) * ld : nullptr;
    __nv_bfloat16* drow16 = dlogits_bf16 ? dlogits_bf16 + static_cast<long long>(b) * ld : nullptr;
#pragma unroll
    for (int i = 0; i < kCeMaxPerLane; ++i) {
      const int c = lane + 32 * i;
      if (c < ld) {
        const float pr = e[i] * inv;
        float g = 0.f;
        if (c < C) {
          g = (pr - (c == y ? 1.f : 0.f)) * scale;
          if (c == y) loss_acc += -__logf(pr + 1e-10f);
        }
        if (drow) drow[c] = g;
        if (drow16) drow16[c] = __float2bfloat16_rn(g);
        cs[i] += g;
      }
    }
    if (lane == 0 && arg == y) hit_acc += 1;
  }

  // Block-level reduction of loss / hits, then one atomic per block.
  __shared__ float s_loss[32];
  __shared__ int s_hit[32];
#pragma unroll
  for (int o = 16; o > 0; o >>= 1) loss_acc += __shfl_xor_sync(0xffffffffu, loss_acc, o);
  if (lane == 0) { s_loss[warp] = loss_acc; s_hit[warp] = hit_acc; }
  __syncthreads();
  if (warp == 0) {
    float l = lane < warps_per_block ? s_loss[lane] : 0.f;
    int h = lane < warps_per_block ? s_hit[lane] : 0;
#pragma unroll
    for (int o = 16; o > 0; o >>= 1) {
      l += __shfl_xor_sync(0xffffffffu, l, o);
      h += __shfl_xor_sync(0xffffffffu, h, o);
    }
    if (lane == 0) {
      if (loss_sum) atomicAdd(loss_sum, l);
      if (correct && h) atomicAdd(correct, h);
    }
  }
  if (colsum != nullptr) {
    // per-lane column partials -> smem across warps -> atomics (ld <= 256 columns)
    __shared__ float s_cs[32][33];
#pragma unroll
    for (int i = 0; i < kCeMaxPerLane; ++i) {
      __syncthreads();
      s_cs[warp][lane] = cs[i];
      __syncthreads();
      if (warp == 0) {
        float t = 0.f;
        for (int w = 0; w < warps_per_block; ++w) t += s_cs[w][lane];
        const int c = lane + 32 * i;
        if (c < C && t != 0.f) atomicAdd(colsum + c, t);
      }
    }
  }
}

__global__ void sgd_kernel(float* __restrict__ w, const float* __restrict__ g, size_t n,
                           const float* __restrict__ lr_ptr, float scale,
                           __nv_bfloat16* __restrict__ w16) {
  const float step = lr_ptr[0] * scale;
  const size_t tid = blockIdx.x * static_cast<size_t>(blockDim.x) + threadIdx.x;
  const size_t stride = static_cast<size_t>(gridDim.x) * blockDim.x;
  const size_t n4 = n / 4;
  const bool aligned = ((reinterpret_cast<uintptr_t>(w) | reinterpret_cast<uintptr_t>(g)) & 15) == 0 &&
                       (w16 == nullptr || (reinterpret_cast<uintptr_t>(w16) & 7) == 0);
  if (aligned) {
    float4* w4 = reinterpret_cast<float4*>(w);
    const float4* g4 = reinterpret_cast<const float4*>(g);
    for (size_t i = tid; i < n4; i += stride) {
      float4 a = w4[i];
      const float4 b = __ldg(g4 + i);
      a.x -= step * b.x; a.y -= step * b.y; a.z -= step * b.z; a.w -= step * b.w;
      w4[i] = a;
      if (w16) {
        __nv_bfloat162 lo = __floats2bfloat162_rn(a.x, a.y), hi = __floats2bfloat162_rn(a.z, a.w);
        uint2 pk;
        pk.x = *reinterpret_cast<uint32_t*>(&lo);
        pk.y = *reinterpret_cast<uint32_t*>(&hi);
        reinterpret_cast<uint2*>(w16)[i] = pk;
      }
    }
    for (size_t i = n4 * 4 + tid; i < n; i += stride) {
      const float a = w[i] - step * g[i];
      w[i] = a;
      if (w16) w16[i] = __float2bfloat16_rn(a);
    }
  } else {
    for (size_t i = tid; i < n; i += stride) {
      const float a = w[i] - step * g[i];
      w[i] = a;
      if (w16) w16[i] = __float2bfloat16_rn(a);
    }
  }
}

__global__ void fill_u32_kernel(uint32_t* __restrict__ dst, uint32_t v, size_t n) {
  const size_t stride = static_cast<size_t>(gridDim.x) * blockDim.x;
  for (size_t i = blockIdx.x * static_cast<size_t>(blockDim.x) + threadIdx.x; i < n; i += stride)
    dst[i] = v;
}

__global__ void cast_f32_bf16_kernel(const float* __restrict__ src, __nv_bfloat16* __restrict__ dst,
                                     size_t n) {
  const size_t stride = static_cast<size_t>(gridDim.x) * blockDim.x;
  for (size_t i = blockIdx.x * static_cast<size_t>(blockDim.x) + threadIdx.x; i < n; i += stride)
    dst[i] = __float2bfloat16_rn(src[i]);
}

static int grid_for(size_t n, int threads, int per_thread) {
  size_t blocks = (n + static_cast<size_t>(threads) * per_thread - 1) / (static_cast<size_t>(threads) * per_thread);
  if (blocks < 1) blocks = 1;
  if (blocks > 148 * 8) blocks = 148 * 8;
  return static_cast<int>(blocks);
}

}  // namespace dmlp

using namespace dmlp;

extern "C" const char* dmlp_last_error(void) { return g_err; }

extern "C" const char* dmlp_build_info(void) {
  return "dmlp_b200 0.1 | sm_100a | tcgen05+TMEM+TMA GEMM, peer-memory collectives | CUDA "
#define DMLP_STR2(x) #x
#define DMLP_STR(x) DMLP_STR2(x)
      DMLP_STR(CUDART_VERSION);
}

extern "C" int dmlp_softmax_ce(const float* logits, int64_t ld, const int32_t* labels, int B, int C,
                               float scale, float* dlogits, void* dlogits_bf16, float* loss_sum,
                               int32_t* correct, float* colsum, void* stream) {
  DMLP_REQUIRE(logits && labels, "dmlp_softmax_ce: null input");
  DMLP_REQUIRE(B > 0 && C > 0 && C <= ld && ld <= 32 * kCeMaxPerLane,
               "dmlp_softmax_ce: need 0 < C <= ld <= %d (got C=%d ld=%lld)", 32 * kCeMaxPerLane, C,
               (long long)ld);
  const int threads = 256;  // 8 rows per block
  int blocks = (B + 7) / 8;
  if (blocks > 148 * 4) blocks = 148 * 4;
  softmax_ce_kernel<<<blocks, threads, 0, as_stream(stream)>>>(
      logits, ld, labels, B, C, scale, dlogits, reinterpret_cast<__nv_bfloat16*>(dlogits_bf16),
      loss_sum, correct, colsum);
  DMLP_CUDA_CHECK(cudaGetLastError());
  return DMLP_OK;
}

extern "C" int dmlp_sgd(float* w, const float* g, size_t n, const float* lr, float scale,
                        void* w_bf16, void* stream) {
  DMLP_REQUIRE(w && g && lr, "dmlp_sgd: null pointer");
  if (n == 0) return DMLP_OK;
  sgd_kernel<<<grid_for(n, 256, 8), 256, 0, as_stream(stream)>>>(
      w, g, n, lr, scale, reinterpret_cast<__nv_bfloat16*>(w_bf16));
  DMLP_CUDA_CHECK(cudaGetLastError());
  return DMLP_OK;
}

extern "C" int dmlp_fill_u32(void* dst, uint32_t value, size_t n_words, void* stream) {
  DMLP_REQUIRE(dst != nullptr || n_words == 0, "dmlp_fill_u32: null pointer");
  if (n_words == 0) return DMLP_OK;
  fill_u32_kernel<<<grid_for(n_words, 256, 8), 256, 0, as_stream(stream)>>>(
      reinterpret_cast<uint32_t*>(dst), value, n_words);
  DMLP_CUDA_CHECK(cudaGetLastError());
  return DMLP_OK;
}

extern "C" int dmlp_cast_f32_to_bf16(const float* src, void* dst, size_t n, void* stream) {
  DMLP_REQUIRE((src && dst) || n == 0, "dmlp_cast_f32_to_bf16: null pointer");
  if (n == 0) return DMLP_OK;
  cast_f32_bf16_kernel<<<grid_for(n, 256, 8), 256, 0, as_stream(stream)>>>(
      src, reinterpret_cast<__nv_bfloat16*>(dst), n);
  DMLP_CUDA_CHECK(cudaGetLastError());
  return DMLP_OK;
}
