// desc_shift.cu -- hardware probe: can a tcgen05 A-operand descriptor start ONE ROW (128 B) above or below a
// 1024-byte aligned SWIZZLE_128B K-major tile, so that MMA row r reads tile row r +- 1?  (Word-order channels of the
// GGNN step: acts[t] += h[t-1] @ W_prev + h[t+1] @ W_next on the resident h tile.)  Tries the descriptor's
// base-offset field (bits 49-51) = 0, the row phase of the start address, and its complement.
//   nvcc -gencode arch=compute_100a,code=sm_100a -O3 -std=c++17 -I ggnn_b200/csrc scripts/micro/desc_shift.cu -o scripts/micro/desc_shift
#include <cstdio>
#include <cstdlib>
#include <cmath>
#include <vector>
#include <cuda_fp16.h>
#include "tc.cuh"

constexpr int M = 128, N = 128, K = 64;
constexpr int CHUNK = M * K * 2;   // 16 KB

__device__ __forceinline__ uint32_t sw_off(int r, int k) {   // byte offset of element (r, k) in a swizzled [128 x 64] f16 chunk
  return (r >> 3) * 1024 + (r & 7) * 128 + ((((k >> 3) ^ (r & 7))) << 4) + (k & 7) * 2;
}

__global__ void __launch_bounds__(128, 1)
probe_kernel(const __half* __restrict__ A, const __half* __restrict__ W, float* __restrict__ out, int delta_rows, int bo) {
  extern __shared__ __align__(1024) uint8_t smem[];
  const int tid = threadIdx.x, warp = tid >> 5, lane = tid & 31;
  const uint32_t base = tc::smem_u32(smem);
  const uint32_t sA = base + 1024, sW = sA + CHUNK + 1024, bar = sW + CHUNK, slot = bar + 8;
  for (int i = tid; i < (CHUNK + 2048 + CHUNK) / 4; i += 128) reinterpret_cast<uint32_t*>(smem)[i] = 0u;   // pads = 0
  __syncthreads();
  for (int i = tid; i < M * K; i += 128) {
    const int r = i / K, k = i % K;
    *reinterpret_cast<__half*>(smem + 1024 + sw_off(r, k)) = A[i];
    *reinterpret_cast<__half*>(smem + 1024 + CHUNK + 1024 + sw_off(r, k)) = W[i];   // W[n][k], n = r
  }
  if (tid == 0) { tc::mbar_init(bar, 1); tc::fence_barrier_init(); }
  if (warp == 0) tc::tmem_alloc(slot, 128);
  tc::fence_proxy_async();
  tc::fence_before_sync();
  __syncthreads();
  tc::fence_after_sync();
  const uint32_t tmem = *reinterpret_cast<volatile uint32_t*>(smem + (slot - base));
  if (tid == 0) {
    const uint32_t idesc = tc::umma_idesc(0, M, N);
    const uint32_t a0 = sA + delta_rows * 128;
#pragma unroll
    for (int k = 0; k < 4; ++k) {
      uint64_t da = tc::umma_desc_sw128(a0 + k * 32) | ((uint64_t)(bo & 7) << 49);
      tc::umma_f16(tmem, da, tc::umma_desc_sw128(sW + k * 32), idesc, k != 0);
    }
    tc::umma_commit(bar);
  }
  tc::mbar_wait(bar, 0);
  tc::fence_after_sync();
  const int r = warp * 32 + lane;
  for (int cc = 0; cc < 4; ++cc) {
    uint32_t acc[32];
    tc::tmem_ld32(tmem + ((uint32_t)(warp * 32) << 16) + cc * 32, acc);
    tc::tmem_ld_wait();
    for (int j = 0; j < 32; ++j) out[r * N + cc * 32 + j] = __uint_as_float(acc[j]);
  }
  tc::fence_before_sync();
  __syncthreads();
  if (warp == 0) tc::tmem_dealloc(tmem, 128);
}

int main() {
  std::vector<__half> A(M * K), W(N * K);
  std::vector<float> Af(M * K), Wf(N * K);
  srand(1);
  for (int i = 0; i < M * K; ++i) { float v = (rand() % 2001 - 1000) / 1000.f; A[i] = __float2half(v); Af[i] = __half2float(A[i]); }
  for (int i = 0; i < N * K; ++i) { float v = (rand() % 2001 - 1000) / 1000.f; W[i] = __float2half(v); Wf[i] = __half2float(W[i]); }
  __half *dA, *dW; float* dO;
  cudaMalloc(&dA, M * K * 2); cudaMalloc(&dW, N * K * 2); cudaMalloc(&dO, M * N * 4);
  cudaMemcpy(dA, A.data(), M * K * 2, cudaMemcpyHostToDevice);
  cudaMemcpy(dW, W.data(), N * K * 2, cudaMemcpyHostToDevice);
  const int smem_bytes = 1024 + CHUNK + 1024 + CHUNK + 1024;
  cudaFuncSetAttribute(probe_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, smem_bytes);
  std::vector<float> out(M * N);
  for (int delta = -1; delta <= 1; ++delta) {
    const int phase = delta & 7;                     // (start >> 7) & 7 for a 1024-aligned tile
    const int bos[3] = {0, phase, (8 - phase) & 7};
    for (int v = 0; v < (delta == 0 ? 1 : 3); ++v) {
      cudaMemset(dO, 0, M * N * 4);
      probe_kernel<<<1, 128, smem_bytes>>>(dA, dW, dO, delta, bos[v]);
      cudaError_t e = cudaDeviceSynchronize();
      if (e != cudaSuccess) { printf("delta=%+d base_offset=%d: CUDA error %s\n", delta, bos[v], cudaGetErrorString(e)); return 1; }
      cudaMemcpy(out.data(), dO, M * N * 4, cudaMemcpyDeviceToHost);
      double worst = 0; int bad_rows = 0;
      for (int r = 0; r < M; ++r) {
        double row_worst = 0;
        for (int n = 0; n < N; ++n) {
          double ref = 0;
          const int sr = r + delta;
          if (sr >= 0 && sr < M) for (int k = 0; k < K; ++k) ref += (double)Af[sr * K + k] * Wf[n * K + k];
          row_worst = fmax(row_worst, fabs(ref - out[r * N + n]));
        }
        if (row_worst > 1e-3) ++bad_rows;
        worst = fmax(worst, row_worst);
      }
      printf("delta=%+d base_offset=%d: max|err|=%.3e bad_rows=%d -> %s\n", delta, bos[v], worst, bad_rows,
             bad_rows == 0 ? "OK" : "MISMATCH");
    }
  }
  return 0;
}
