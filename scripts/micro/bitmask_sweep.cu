// Standalone sweep of the dense-adjacency -> bitmask pass (development aid): thread count, loads in flight,
// grid size, 128- vs 256-bit loads, L2 eviction hint.  nvcc -O3 -gencode arch=compute_100a,code=sm_100a
#include <cstdio>
#include <cstdint>
#include <cuda_runtime.h>

__device__ __forceinline__ float4 ld128(const float4* p) {
  float4 r;
  asm volatile("ld.global.nc.L1::no_allocate.v4.f32 {%0,%1,%2,%3}, [%4];"
               : "=f"(r.x), "=f"(r.y), "=f"(r.z), "=f"(r.w) : "l"(p));
  return r;
}
__device__ __forceinline__ float4 ld128_ef(const float4* p) {
  float4 r;
  asm volatile("ld.global.cs.v4.f32 {%0,%1,%2,%3}, [%4];"
               : "=f"(r.x), "=f"(r.y), "=f"(r.z), "=f"(r.w) : "l"(p));
  return r;
}
struct f8 { float v[8]; };
__device__ __forceinline__ f8 ld256(const float* p) {
  f8 r;
  uint32_t u[8];
  asm volatile("ld.global.nc.L1::no_allocate.L2::evict_first.v8.b32 {%0,%1,%2,%3,%4,%5,%6,%7}, [%8];"
               : "=r"(u[0]), "=r"(u[1]), "=r"(u[2]), "=r"(u[3]), "=r"(u[4]), "=r"(u[5]), "=r"(u[6]), "=r"(u[7])
               : "l"(p));
#pragma unroll
  for (int q = 0; q < 8; ++q) r.v[q] = __uint_as_float(u[q]);
  return r;
}

template <int THREADS, int UNROLL, int HINT>
__global__ void __launch_bounds__(THREADS) k128(const float4* __restrict__ adj, int64_t n4, uint32_t* __restrict__ bits) {
  const int lane = threadIdx.x & 31;
  const int64_t stride = (int64_t)gridDim.x * THREADS;
  const int64_t iters = (n4 + stride * UNROLL - 1) / (stride * UNROLL);
  int64_t i0 = (int64_t)blockIdx.x * THREADS + threadIdx.x;
  for (int64_t it = 0; it < iters; ++it, i0 += stride * UNROLL) {
    float4 x[UNROLL];
#pragma unroll
    for (int u = 0; u < UNROLL; ++u) {
      int64_t i = i0 + u * stride;
      x[u] = i < n4 ? (HINT ? ld128_ef(adj + i) : ld128(adj + i)) : make_float4(0.f, 0.f, 0.f, 0.f);
    }
#pragma unroll
    for (int u = 0; u < UNROLL; ++u) {
      int64_t i = i0 + u * stride;
      uint32_t nib = (x[u].x != 0.f ? 1u : 0u) | (x[u].y != 0.f ? 2u : 0u) | (x[u].z != 0.f ? 4u : 0u) | (x[u].w != 0.f ? 8u : 0u);
      uint32_t w = nib << ((lane & 7) * 4);
      w |= __shfl_xor_sync(0xffffffffu, w, 1);
      w |= __shfl_xor_sync(0xffffffffu, w, 2);
      w |= __shfl_xor_sync(0xffffffffu, w, 4);
      if ((lane & 7) == 0 && i < n4) bits[i >> 3] = w;
    }
  }
}

// 256-bit loads: 8 floats per thread -> one byte; 4 adjacent lanes -> one word
template <int THREADS, int UNROLL>
__global__ void __launch_bounds__(THREADS) k256(const float* __restrict__ adj, int64_t n8, uint32_t* __restrict__ bits) {
  const int lane = threadIdx.x & 31;
  const int64_t stride = (int64_t)gridDim.x * THREADS;
  const int64_t iters = (n8 + stride * UNROLL - 1) / (stride * UNROLL);
  int64_t i0 = (int64_t)blockIdx.x * THREADS + threadIdx.x;
  for (int64_t it = 0; it < iters; ++it, i0 += stride * UNROLL) {
    f8 x[UNROLL];
#pragma unroll
    for (int u = 0; u < UNROLL; ++u) {
      int64_t i = i0 + u * stride;
      if (i < n8) x[u] = ld256(adj + i * 8);
      else { for (int q = 0; q < 8; ++q) x[u].v[q] = 0.f; }
    }
#pragma unroll
    for (int u = 0; u < UNROLL; ++u) {
      int64_t i = i0 + u * stride;
      uint32_t by = 0;
#pragma unroll
      for (int q = 0; q < 8; ++q) by |= (x[u].v[q] != 0.f ? 1u : 0u) << q;
      uint32_t w = by << ((lane & 3) * 8);
      w |= __shfl_xor_sync(0xffffffffu, w, 1);
      w |= __shfl_xor_sync(0xffffffffu, w, 2);
      if ((lane & 3) == 0 && i < n8) bits[i >> 2] = w;
    }
  }
}

template <class F> float timeit(F f, int reps) {
  cudaEvent_t a, b; cudaEventCreate(&a); cudaEventCreate(&b);
  f(); cudaDeviceSynchronize();
  cudaEventRecord(a);
  for (int r = 0; r < reps; ++r) f();
  cudaEventRecord(b); cudaEventSynchronize(b);
  float ms; cudaEventElapsedTime(&ms, a, b);
  return ms / reps;
}

int main() {
  const int64_t graphs = 4096, per = 92LL * 128 * 128;
  const int64_t total = graphs * per;                         // floats
  float* adj; uint32_t* bits; uint32_t* bits2;
  cudaMalloc(&adj, total * 4); cudaMalloc(&bits, total / 8); cudaMalloc(&bits2, total / 8);
  cudaMemset(adj, 0, total * 4);
  // sprinkle ones so that the result is checkable
  {
    float one = 1.f;
    for (int64_t i = 0; i < total; i += 2957) cudaMemcpy(adj + i, &one, 4, cudaMemcpyHostToDevice);
  }
  const double gb = total * 4.0 / 1e9 * (1.0 + 1.0 / 32);
  const int sm = 148;
  const int64_t n4 = total / 4, n8 = total / 8;
  k128<256, 4, 0><<<sm * 16, 256>>>(reinterpret_cast<const float4*>(adj), n4, bits);   // reference result
  cudaDeviceSynchronize();
#define RUN128(T, U, H, G)                                                                                   \
  {                                                                                                          \
    float ms = timeit([&] { k128<T, U, H><<<sm * G, T>>>(reinterpret_cast<const float4*>(adj), n4, bits2); }, 5); \
    printf("128b threads=%d unroll=%d hint=%d grid=148x%d : %.3f ms  %.1f GB/s\n", T, U, H, G, ms, gb / ms * 1e3); \
  }
#define RUN256(T, U, G)                                                                                      \
  {                                                                                                          \
    float ms = timeit([&] { k256<T, U><<<sm * G, T>>>(adj, n8, bits2); }, 5);                                  \
    printf("256b threads=%d unroll=%d grid=148x%d : %.3f ms  %.1f GB/s\n", T, U, G, ms, gb / ms * 1e3);      \
  }
  RUN128(256, 4, 0, 16) RUN128(256, 4, 1, 16) RUN128(256, 8, 0, 8) RUN128(256, 8, 1, 8) RUN128(256, 8, 1, 16)
  RUN128(512, 4, 1, 4) RUN128(512, 4, 1, 8) RUN128(256, 4, 1, 32) RUN128(128, 8, 1, 16) RUN128(256, 2, 1, 32)
  RUN128(1024, 2, 1, 2) RUN128(1024, 4, 1, 2)
  RUN256(256, 2, 16) RUN256(256, 4, 8) RUN256(256, 4, 16) RUN256(512, 2, 8) RUN256(128, 4, 16) RUN256(256, 2, 32)
  RUN256(256, 4, 12) RUN256(256, 4, 24) RUN256(256, 4, 32) RUN256(256, 3, 16) RUN256(256, 6, 16) RUN256(512, 4, 8) RUN256(512, 4, 4)
  RUN256(384, 4, 10) RUN256(256, 4, 64) RUN256(256, 4, 16) RUN256(1024, 2, 2) RUN256(1024, 4, 2)
  // check the 256-bit variant against the reference bitmask
  k256<256, 4><<<sm * 8, 256>>>(adj, n8, bits2);
  cudaDeviceSynchronize();
  uint32_t* h1 = (uint32_t*)malloc(total / 8); uint32_t* h2 = (uint32_t*)malloc(total / 8);
  cudaMemcpy(h1, bits, total / 8, cudaMemcpyDeviceToHost); cudaMemcpy(h2, bits2, total / 8, cudaMemcpyDeviceToHost);
  int64_t bad = 0; for (int64_t i = 0; i < total / 32; ++i) bad += h1[i] != h2[i];
  printf("256b result mismatches: %lld  (%s)\n", (long long)bad, cudaGetErrorString(cudaGetLastError()));
  return 0;
}
