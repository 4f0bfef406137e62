// bg_stream.cu -- can a small-footprint streaming CTA (12 warps x 40 registers, no shared memory: what fits next to the
// persistent GRU kernel) be lifted from ~4.5 TB/s towards the HBM rate by pulling its data into L2 ahead of its loads?
// Same inner loop as compact.cu's bitmask_queue_kernel (256-bit streaming loads, 4 per thread in flight, one bitmask
// byte per load), static slice assignment.  Modes: 0 none; 1 thread 0 bulk-prefetches the CTA's next slice piece by
// piece; 2 every 4th lane prefetch.global.L2 one slice ahead; 3 = 1 with two slices of distance.
//   nvcc -gencode arch=compute_100a,code=sm_100a -O3 -std=c++17 scripts/micro/bg_stream.cu -o scripts/micro/bg_stream
#include <cstdio>
#include <cstdint>
#include <cuda_runtime.h>

constexpr int64_t Q_SLICE = 12288;     // 32-byte items per slice (384 KB)
constexpr int UNROLL = 4;
struct f32x8 { uint32_t u[8]; };
__device__ __forceinline__ f32x8 ldg8(const float* p) {
  f32x8 r;
  asm volatile("ld.global.nc.L1::no_allocate.L2::evict_first.v8.b32 {%0,%1,%2,%3,%4,%5,%6,%7}, [%8];"
               : "=r"(r.u[0]), "=r"(r.u[1]), "=r"(r.u[2]), "=r"(r.u[3]), "=r"(r.u[4]), "=r"(r.u[5]), "=r"(r.u[6]), "=r"(r.u[7]) : "l"(p));
  return r;
}

// dist = how many pieces (nthr x 4 loads x 32 B) ahead of its own loads thread 0 bulk-prefetches, following this CTA's
// stream across slice boundaries; 0 = no prefetch
__global__ void __maxnreg__(40) stream_kernel(const float* __restrict__ adj, int64_t n_items, uint32_t* __restrict__ bits, int dist) {
  const int lane = threadIdx.x & 31, nthr = blockDim.x;
  const int64_t n_slices = (n_items + Q_SLICE - 1) / Q_SLICE;
  const int piece = nthr * UNROLL, ppslice = (int)Q_SLICE / piece;
  for (int64_t cur = blockIdx.x; cur < n_slices; cur += gridDim.x) {
    const int64_t g0 = cur * Q_SLICE + threadIdx.x;
#pragma unroll 1
    for (int k = 0; k < ppslice; ++k) {
      if (dist && threadIdx.x == 0) {
        const int kk = k + dist;
        const int64_t sl = cur + (int64_t)(kk / ppslice) * gridDim.x;
        if (sl < n_slices) {
          const char* p = reinterpret_cast<const char*>(adj) + (sl * Q_SLICE + (int64_t)(kk % ppslice) * piece) * 32;
          asm volatile("cp.async.bulk.prefetch.L2.global [%0], %1;" ::"l"(p), "r"(piece * 32) : "memory");
        }
      }
      const int64_t i0 = g0 + (int64_t)k * piece;
      f32x8 x[UNROLL];
#pragma unroll
      for (int u = 0; u < UNROLL; ++u) x[u] = ldg8(adj + (i0 + u * nthr) * 8);
#pragma unroll
      for (int u = 0; u < UNROLL; ++u) {
        const int64_t i = i0 + u * nthr;
        uint32_t by = 0;
#pragma unroll
        for (int q = 0; q < 8; ++q) by |= (__uint_as_float(x[u].u[q]) != 0.f ? 1u : 0u) << q;
        uint32_t w = by << ((lane & 3) * 8);
        w |= __shfl_xor_sync(0xffffffffu, w, 1);
        w |= __shfl_xor_sync(0xffffffffu, w, 2);
        if ((lane & 3) == 0) bits[i >> 2] = w;
      }
    }
  }
}

static void run(const float* adj, int64_t n_items, uint32_t* bits, int sms, int threads, int dist) {
  cudaEvent_t a, b;
  cudaEventCreate(&a); cudaEventCreate(&b);
  stream_kernel<<<sms, threads>>>(adj, n_items, bits, dist);
  float best = 1e9f, worst = 0.f;
  for (int rep = 0; rep < 4; ++rep) {
    cudaEventRecord(a);
    stream_kernel<<<sms, threads>>>(adj, n_items, bits, dist);
    cudaEventRecord(b);
    cudaDeviceSynchronize();
    float ms = 0;
    cudaEventElapsedTime(&ms, a, b);
    best = ms < best ? ms : best; worst = ms > worst ? ms : worst;
  }
  cudaError_t e = cudaGetLastError();
  printf("threads %d, prefetch %2d pieces (%4d KB) ahead: %6.3f .. %6.3f ms per pass = %5.2f .. %5.2f TB/s (%s)\n", threads, dist,
         dist * threads * UNROLL * 32 / 1024, best, worst, n_items * 32.0 / worst / 1e9, n_items * 32.0 / best / 1e9,
         e == cudaSuccess ? "ok" : cudaGetErrorString(e));
}

int main() {
  int sms = 0;
  cudaDeviceGetAttribute(&sms, cudaDevAttrMultiProcessorCount, 0);
  const int64_t n_slices = 148 * 128;                  // 7.4 GB: far beyond L2
  const int64_t n_items = n_slices * Q_SLICE;
  float* adj; uint32_t* bits;
  cudaMalloc(&adj, n_items * 32); cudaMalloc(&bits, n_items);
  cudaMemset(adj, 0, n_items * 32);
  for (int threads : {384, 256, 128})
    for (int dist : {0, 1, 2, 4, 8, 16}) run(adj, n_items, bits, sms, threads, dist);
  return 0;
}
