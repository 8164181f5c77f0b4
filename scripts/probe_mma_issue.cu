// Microbenchmark: what bounds a short chain of tcgen05.mma instructions (operands in shared memory)?
//   nvcc -gencode arch=compute_100a,code=sm_100a -O3 -std=c++17 -I csrc -o build/probe_mma_issue scripts/probe_mma_issue.cu
// One CTA, one thread issues R MMAs and one commit, then waits for the mbarrier; clock64 around it.
// Variants: N (16..256), element size (tf32 / bf16), M (128 / 64), number of accumulators the chain
// rotates over (1 = every MMA depends on the previous one's accumulator), A/B K-major tiles of zeros.
#include <cstdio>
#include <cstdint>
#include <cuda.h>
#include <cuda_runtime.h>
#include "ptx_sm100.cuh"

using namespace dmlp;

template <int EB, bool UNI>
__global__ void __launch_bounds__(128, 1) probe(int R, int N, int M, int nacc, int same_k, long long* out) {
  extern __shared__ uint8_t raw[];
  uint8_t* smem = reinterpret_cast<uint8_t*>((reinterpret_cast<uintptr_t>(raw) + 1023) & ~uintptr_t(1023));
  __shared__ __align__(8) uint64_t bar;
  __shared__ uint32_t slot;
  for (int i = threadIdx.x; i < (16384 + 32768) / 4; i += blockDim.x) reinterpret_cast<uint32_t*>(smem)[i] = 0;
  if (threadIdx.x == 0) {
    mbar_init(&bar, 1);
    fence_mbar_init();
  }
  if (threadIdx.x < 32) {
    tmem_alloc(&slot, 512);
    tmem_relinquish();
  }
  fence_proxy_async_smem();
  tc_fence_before_sync();
  __syncthreads();
  tc_fence_after_sync();
  const uint32_t tm = slot;
  const uint32_t sb = smem_u32(smem);
  if (UNI ? (threadIdx.x < 32) : (threadIdx.x == 0)) {
    uint32_t idesc = make_instr_desc(EB, false, false, N);
    idesc = (idesc & ~(0x1Fu << 24)) | (static_cast<uint32_t>(M >> 4) << 24);
    uint32_t phase = 0;
    for (int rep = 0; rep < 5; ++rep) {
      const long long t0 = clock64();
      for (int t = 0; t < R; ++t) {
        const int ks = same_k ? 0 : (t & 3);
        const uint64_t ad = make_smem_desc(sb + ks * 32, 16, 1024, 2);
        const uint64_t bd = make_smem_desc(sb + 16384 + ks * 32, 16, 1024, 2);
        if (UNI) {
          if (elect_one()) tc_mma<EB>(tm + (t % nacc) * (512 / nacc), ad, bd, idesc, t >= nacc);
        } else {
          tc_mma<EB>(tm + (t % nacc) * (512 / nacc), ad, bd, idesc, t >= nacc);
        }
      }
      if (UNI) {
        if (elect_one()) tc_commit(&bar);
        __syncwarp();
      } else {
        tc_commit(&bar);
      }
      const long long t1 = clock64();
      mbar_wait(&bar, phase);
      const long long t2 = clock64();
      phase ^= 1;
      if (threadIdx.x == 0) {
        out[rep * 2] = t1 - t0;
        out[rep * 2 + 1] = t2 - t0;
      }
    }
  }
  tc_fence_before_sync();
  __syncthreads();
  if (threadIdx.x < 32) {
    tc_fence_after_sync();
    tmem_dealloc(tm, 512);
  }
}

template <int EB, bool UNI = false>
static void run(int R, int N, int M, int nacc, int same_k, long long* dout) {
  cudaFuncSetAttribute(probe<EB, UNI>, cudaFuncAttributeMaxDynamicSharedMemorySize, 64 * 1024);
  probe<EB, UNI><<<1, 128, 64 * 1024>>>(R, N, M, nacc, same_k, dout);
  long long h[10];
  cudaError_t e = cudaMemcpy(h, dout, sizeof(h), cudaMemcpyDeviceToHost);
  if (e != cudaSuccess) {
    printf("error %s\n", cudaGetErrorString(e));
    return;
  }
  printf("uni=%d eb=%d M=%3d N=%3d R=%2d nacc=%d same_k=%d : issue %5lld cyc, total %6lld cyc -> %6.1f cyc/mma\n", int(UNI), EB, M, N, R,
         nacc, same_k, h[8], h[9], double(h[9]) / R);
}


// Fully unrolled issue: descriptors precomputed, compile-time trip count, no integer division.
template <int EB, int R, int NACC, bool UNI>
__global__ void __launch_bounds__(128, 1) probe_unrolled(int N, int M, long long* out) {
  extern __shared__ uint8_t raw[];
  uint8_t* smem = reinterpret_cast<uint8_t*>((reinterpret_cast<uintptr_t>(raw) + 1023) & ~uintptr_t(1023));
  __shared__ __align__(8) uint64_t bar;
  __shared__ uint32_t slot;
  for (int i = threadIdx.x; i < (16384 + 32768) / 4; i += blockDim.x) reinterpret_cast<uint32_t*>(smem)[i] = 0;
  if (threadIdx.x == 0) {
    mbar_init(&bar, 1);
    fence_mbar_init();
  }
  if (threadIdx.x < 32) {
    tmem_alloc(&slot, 512);
    tmem_relinquish();
  }
  fence_proxy_async_smem();
  tc_fence_before_sync();
  __syncthreads();
  tc_fence_after_sync();
  const uint32_t tm = UNI ? __shfl_sync(0xffffffffu, slot, 0) : slot;
  const uint32_t sb = smem_u32(smem);
  const int warp_u = __shfl_sync(0xffffffffu, threadIdx.x >> 5, 0);
  if (UNI ? (warp_u == 0) : (threadIdx.x == 0)) {
    uint32_t idesc = make_instr_desc(EB, false, false, N);
    idesc = (idesc & ~(0x1Fu << 24)) | (static_cast<uint32_t>(M >> 4) << 24);
    const uint64_t ad0 = make_smem_desc(sb, 16, 1024, 2);
    const uint64_t bd0 = make_smem_desc(sb + 16384, 16, 1024, 2);
    uint32_t phase = 0;
    for (int rep = 0; rep < 5; ++rep) {
      const long long t0 = clock64();
      if (!UNI || elect_one()) {
#pragma unroll
        for (int t = 0; t < R; ++t)
          tc_mma<EB>(tm + (t % NACC) * (512 / NACC), ad0 + 2 * (t & 3), bd0 + 2 * (t & 3), idesc, t >= NACC);
        tc_commit(&bar);
      }
      if (UNI) __syncwarp();
      const long long t1 = clock64();
      mbar_wait(&bar, phase);
      const long long t2 = clock64();
      phase ^= 1;
      if (threadIdx.x == 0) {
        out[rep * 2] = t1 - t0;
        out[rep * 2 + 1] = t2 - t0;
      }
    }
  }
  tc_fence_before_sync();
  __syncthreads();
  if (threadIdx.x < 32) {
    tc_fence_after_sync();
    tmem_dealloc(tm, 512);
  }
}

template <int EB, int R, int NACC, bool UNI>
static void run_unrolled(int N, int M, long long* dout) {
  cudaFuncSetAttribute(probe_unrolled<EB, R, NACC, UNI>, cudaFuncAttributeMaxDynamicSharedMemorySize, 64 * 1024);
  probe_unrolled<EB, R, NACC, UNI><<<1, 128, 64 * 1024>>>(N, M, dout);
  long long h[10];
  cudaError_t e = cudaMemcpy(h, dout, sizeof(h), cudaMemcpyDeviceToHost);
  if (e != cudaSuccess) {
    printf("error %s\n", cudaGetErrorString(e));
    return;
  }
  printf("UNROLLED uni=%d eb=%d M=%3d N=%3d R=%2d nacc=%d : issue %5lld cyc, total %6lld cyc -> %6.1f cyc/mma\n", int(UNI),
         EB, M, N, R, NACC, h[8], h[9], double(h[9]) / R);
}

int main() {
  long long* dout;
  cudaMalloc(&dout, 80);
  const int Ns[] = {16, 32, 64, 128, 256};
  for (int M : {128, 64})
    for (int N : Ns) {
      // nacc > 1 places accumulators 256 / 128 columns apart: N <= 128 for nacc 4, N <= 256 for nacc 2
      run<4>(16, N, M, 1, 0, dout);
      run<4>(16, N, M, 2, 0, dout);
      if (N <= 128) run<4>(16, N, M, 4, 0, dout);
    }
  run<4>(1, 128, 128, 1, 0, dout);
  run<4>(2, 128, 128, 1, 0, dout);
  run<4>(4, 128, 128, 1, 0, dout);
  run<4>(8, 128, 128, 1, 0, dout);
  run<4>(32, 128, 128, 1, 0, dout);
  run<4>(16, 128, 128, 1, 1, dout);
  run<2>(16, 128, 128, 1, 0, dout);
  run<2>(16, 256, 128, 1, 0, dout);
  run<2>(16, 32, 128, 1, 0, dout);
  run<2>(16, 32, 128, 2, 0, dout);
  run<2>(16, 32, 64, 1, 0, dout);
  for (int N : Ns) {
    run<4, true>(16, N, 128, 1, 0, dout);
    run<4, true>(16, N, 128, 2, 0, dout);
    run<4, true>(16, N, 64, 1, 0, dout);
  }
  for (int R : {1, 2, 4, 8, 32, 64}) run<4, true>(R, 128, 128, 1, 0, dout);
  for (int R : {1, 2, 8, 32}) run<4, true>(R, 16, 128, 1, 0, dout);
  run<2, true>(32, 256, 128, 1, 0, dout);
  run<2, true>(32, 128, 128, 1, 0, dout);
  run<2, true>(32, 32, 128, 1, 0, dout);
  for (int N : Ns) {
    run_unrolled<4, 16, 1, false>(N, 128, dout);
    run_unrolled<4, 16, 2, false>(N, 128, dout);
    run_unrolled<4, 16, 1, true>(N, 128, dout);
    run_unrolled<2, 16, 1, false>(N, 128, dout);
    run_unrolled<4, 16, 1, false>(N, 64, dout);
  }
  run_unrolled<4, 1, 1, false>(128, 128, dout);
  run_unrolled<4, 2, 1, false>(128, 128, dout);
  run_unrolled<4, 4, 1, false>(128, 128, dout);
  run_unrolled<4, 8, 1, false>(128, 128, dout);
  run_unrolled<4, 32, 1, false>(128, 128, dout);
  run_unrolled<4, 32, 1, false>(16, 128, dout);
  run_unrolled<2, 32, 1, false>(256, 128, dout);
  run_unrolled<2, 32, 2, false>(256, 128, dout);
  return 0;
}
