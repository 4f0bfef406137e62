// mc_stream.cu -- does TMA multicast lift the L2 -> SM bound of the wide-h contractions?
// The FP32-operand GEMMs of the C5 backward and the K = 896 FP16 gate contraction at h = 400 need 64-125 bytes per clock and
// SM from L2, and one SM takes ~42 (scripts/micro/s9_stream.cu: 80 GB/s).  If two CTAs of a cluster hold different
// column slices of the weights and stream the SAME activation chunks, each chunk can be fetched once and multicast to
// both.  This probe measures what an SM RECEIVES per microsecond in that arrangement against every CTA fetching
// everything itself: all 148 SMs, 16 KB bulk copies out of an L2-resident 3 MB set, 8-slot ring, nothing consumed.
//   mode 0  every CTA issues every chunk for itself                                   (cp.async.bulk)
//   mode 1  clusters of 2: CTA r issues the chunks with (i & 1) == r to BOTH CTAs     (cp.async.bulk ... .multicast::cluster)
// Slot reuse in mode 1 is guarded the way a real kernel must: a consumer releases a slot on the empty barrier of the CTA
// that will refill it (remote mbarrier arrive), and the issuer waits for both releases.
//   nvcc -gencode arch=compute_100a,code=sm_100a -O3 -std=c++17 scripts/micro/mc_stream.cu -o scripts/micro/mc_stream
#include <cstdio>
#include <cstdlib>
#include <cstdint>
#include <cuda_runtime.h>

#ifndef MC_CHUNK
#define MC_CHUNK 16384
#endif
#ifndef MC_STAGES
#define MC_STAGES 8
#endif
constexpr int CHUNK = MC_CHUNK;       // (-DMC_CHUNK=8192 -DMC_STAGES=16: same shared memory, twice the copies in flight)
constexpr int STAGES = MC_STAGES;

__device__ __forceinline__ uint32_t smem_u32(const void* p) { return (uint32_t)__cvta_generic_to_shared(p); }
__device__ __forceinline__ void wait_parity(uint32_t bar, unsigned parity) {
  asm volatile(
      "{\n\t.reg .pred p;\n\tW: mbarrier.try_wait.parity.shared::cta.b64 p, [%0], %1, 2000;\n\t@p bra D;\n\tbra W;\n\tD:\n\t}"
      ::"r"(bar), "r"(parity) : "memory");
}
__device__ __forceinline__ uint32_t cta_rank() {
  uint32_t r;
  asm volatile("mov.u32 %0, %%cluster_ctarank;" : "=r"(r));
  return r;
}
__device__ __forceinline__ void cluster_sync() {
  asm volatile("barrier.cluster.arrive.release.aligned;\n\tbarrier.cluster.wait.acquire.aligned;" ::: "memory");
}
// arrive on the barrier at the same shared-memory offset in CTA `target` of the cluster
__device__ __forceinline__ void remote_arrive(uint32_t bar, uint32_t target) {
  uint32_t remote;
  asm volatile("mapa.shared::cluster.u32 %0, %1, %2;" : "=r"(remote) : "r"(bar), "r"(target));
  asm volatile("mbarrier.arrive.release.cluster.shared::cluster.b64 _, [%0];" ::"r"(remote) : "memory");
}

template <int MODE>
__global__ void __launch_bounds__(128, 1)
stream_kernel(const uint8_t* __restrict__ src_base_, int n_chunks_set, long long total, unsigned long long* out_clk,
              size_t private_bytes) {
  extern __shared__ __align__(1024) uint8_t smem[];
  __shared__ __align__(8) uint64_t full[STAGES];
  __shared__ __align__(8) uint64_t empty[STAGES];
  const int tid = threadIdx.x;
  const uint32_t rank = MODE ? cta_rank() : 0u;
  // private_bytes != 0: every CTA (pair) streams its OWN region, so no two SMs ever ask for the same line
  const uint8_t* src_base = src_base_ + (size_t)(blockIdx.x >> (MODE ? 1 : 0)) * private_bytes;
  if (tid == 0) {
    for (int s = 0; s < STAGES; ++s) {
      asm volatile("mbarrier.init.shared::cta.b64 [%0], 1;" ::"r"(smem_u32(&full[s])));
      asm volatile("mbarrier.init.shared::cta.b64 [%0], 2;" ::"r"(smem_u32(&empty[s])));
    }
    asm volatile("fence.mbarrier_init.release.cluster;" ::: "memory");
  }
  __syncthreads();
  if (MODE) cluster_sync();
  const unsigned long long t0 = clock64();
  if (tid == 0) {
    long long issued = 0, done = 0;             // chunk indices of the PAIR's common stream
    unsigned rng = 12345u + (blockIdx.x >> (MODE ? 1 : 0)) * 7919u;
    while (done < total) {
      while (issued < total && issued - done < STAGES) {
        const int s = (int)(issued % STAGES);
        rng = rng * 1664525u + 1013904223u;     // both CTAs of a pair walk the same sequence
        const uint8_t* src = src_base + (size_t)((rng >> 8) % (unsigned)n_chunks_set) * CHUNK;
        // every CTA expects the chunk on its own barrier, whoever issues it
        asm volatile("mbarrier.arrive.expect_tx.shared::cta.b64 _, [%0], %1;" ::"r"(smem_u32(&full[s])), "r"(CHUNK) : "memory");
        if (MODE == 0) {
          asm volatile("cp.async.bulk.shared::cluster.global.mbarrier::complete_tx::bytes [%0], [%1], %2, [%3];"
                       ::"r"(smem_u32(smem + s * CHUNK)), "l"(src), "r"(CHUNK), "r"(smem_u32(&full[s])) : "memory");
        } else if ((uint32_t)(issued & 1) == rank) {
          // the slot must be free in BOTH CTAs: two releases per use (from the second use of a slot on)
          if (issued >= STAGES) wait_parity(smem_u32(&empty[s]), (unsigned)(((issued / STAGES) - 1) & 1));
          asm volatile("cp.async.bulk.shared::cluster.global.mbarrier::complete_tx::bytes.multicast::cluster [%0], [%1], %2, [%3], %4;"
                       ::"r"(smem_u32(smem + s * CHUNK)), "l"(src), "r"(CHUNK), "r"(smem_u32(&full[s])), "h"((unsigned short)3) : "memory");
        }
        ++issued;
      }
      const int s = (int)(done % STAGES);
      wait_parity(smem_u32(&full[s]), (unsigned)((done / STAGES) & 1));
      if (MODE) {                               // "consumed": release the slot to the CTA that refills it
        const uint32_t refiller = (uint32_t)((done + STAGES) & 1);   // STAGES is even: same CTA as this use
        remote_arrive(smem_u32(&empty[s]), refiller);
      }
      ++done;
    }
  }
  __syncthreads();
  if (MODE) cluster_sync();
  if (tid == 0) out_clk[blockIdx.x] = clock64() - t0;
}

// mode 2: unicast like mode 0, but TWO threads of the CTA (different warps) each drive half of the ring: is the ~0.2 us per
// copy a cost of the issuing thread or of the SM's copy engine?
__global__ void __launch_bounds__(128, 1)
stream2_kernel(const uint8_t* __restrict__ src_base_, int n_chunks_set, long long total, size_t private_bytes) {
  extern __shared__ __align__(1024) uint8_t smem[];
  __shared__ __align__(8) uint64_t full[STAGES];
  const int tid = threadIdx.x;
  const uint8_t* src_base = src_base_ + (size_t)blockIdx.x * private_bytes;
  if (tid == 0) {
    for (int s = 0; s < STAGES; ++s) asm volatile("mbarrier.init.shared::cta.b64 [%0], 1;" ::"r"(smem_u32(&full[s])));
    asm volatile("fence.mbarrier_init.release.cluster;" ::: "memory");
  }
  __syncthreads();
  if (tid == 0 || tid == 32) {
    const int half = tid >> 5, HS = STAGES / 2;
    long long issued = 0, done = 0;
    const long long my_total = total / 2;
    unsigned rng = 12345u + blockIdx.x * 7919u + half * 104729u;
    while (done < my_total) {
      while (issued < my_total && issued - done < HS) {
        const int s = half * HS + (int)(issued % HS);
        rng = rng * 1664525u + 1013904223u;
        const uint8_t* src = src_base + (size_t)((rng >> 8) % (unsigned)n_chunks_set) * CHUNK;
        asm volatile("mbarrier.arrive.expect_tx.shared::cta.b64 _, [%0], %1;" ::"r"(smem_u32(&full[s])), "r"(CHUNK) : "memory");
        asm volatile("cp.async.bulk.shared::cluster.global.mbarrier::complete_tx::bytes [%0], [%1], %2, [%3];"
                     ::"r"(smem_u32(smem + s * CHUNK)), "l"(src), "r"(CHUNK), "r"(smem_u32(&full[s])) : "memory");
        ++issued;
      }
      const int s = half * HS + (int)(done % HS);
      wait_parity(smem_u32(&full[s]), (unsigned)((done / HS) & 1));
      ++done;
    }
  }
  __syncthreads();
}

int main() {
  int dev = 0, sms = 0;
  cudaDeviceGetAttribute(&sms, cudaDevAttrMultiProcessorCount, dev);
  const int n_chunks_set = 3 * 1024 * 1024 / CHUNK;
  const size_t private_bytes = 512 * 1024;        // 148 x 512 KB = 74 MB: L2-resident after the warm-up launch
  uint8_t* w;
  cudaMalloc(&w, (size_t)sms * private_bytes + (size_t)n_chunks_set * CHUNK);
  cudaMemset(w, 1, (size_t)sms * private_bytes + (size_t)n_chunks_set * CHUNK);
  unsigned long long* clk;
  cudaMalloc(&clk, sms * sizeof(unsigned long long));
  const int smem_bytes = STAGES * CHUNK;
  cudaFuncSetAttribute(stream_kernel<0>, cudaFuncAttributeMaxDynamicSharedMemorySize, smem_bytes);
  cudaFuncSetAttribute(stream_kernel<1>, cudaFuncAttributeMaxDynamicSharedMemorySize, smem_bytes);
  const long long total = (64ll << 20) / CHUNK;  // chunks received per CTA: 64 MB
  printf("SMs %d; %d-slot ring of %d KB bulk copies out of an L2-resident 3 MB set; %lld chunks received per CTA\n", sms, STAGES,
         CHUNK / 1024, total);
  cudaFuncSetAttribute(stream2_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, smem_bytes);
  for (int rep = 0; rep < 3; ++rep) {
    cudaEvent_t a, b;
    cudaEventCreate(&a); cudaEventCreate(&b);
    cudaEventRecord(a);
    stream2_kernel<<<sms, 128, smem_bytes>>>(w, (int)(private_bytes / CHUNK), total, private_bytes);
    cudaEventRecord(b);
    if (cudaDeviceSynchronize() != cudaSuccess) { printf("CUDA error\n"); return 1; }
    float ms = 0;
    cudaEventElapsedTime(&ms, a, b);
    if (rep == 2)
      printf("private regions, mode 2 (two issuing threads, half a ring each): %.3f ms -> each SM RECEIVES %.1f GB/s\n", ms,
             (double)total * CHUNK / ms / 1e6);
  }
  for (int priv = 0; priv < 2; ++priv)
  for (int mode = 0; mode < 2; ++mode) {
    const size_t pb = priv ? private_bytes : 0;
    const int nset = priv ? (int)(private_bytes / CHUNK) : n_chunks_set;
    for (int rep = 0; rep < 3; ++rep) {
      cudaEvent_t a, b;
      cudaEventCreate(&a); cudaEventCreate(&b);
      cudaEventRecord(a);
      if (mode == 0) {
        stream_kernel<0><<<sms, 128, smem_bytes>>>(w, nset, total, clk, pb);
      } else {
        cudaLaunchConfig_t cfg = {};
        cfg.gridDim = dim3(sms & ~1); cfg.blockDim = dim3(128); cfg.dynamicSmemBytes = smem_bytes;
        cudaLaunchAttribute at[1];
        at[0].id = cudaLaunchAttributeClusterDimension;
        at[0].val.clusterDim.x = 2; at[0].val.clusterDim.y = 1; at[0].val.clusterDim.z = 1;
        cfg.attrs = at; cfg.numAttrs = 1;
        cudaLaunchKernelEx(&cfg, stream_kernel<1>, (const uint8_t*)w, nset, total, clk, pb);
      }
      cudaEventRecord(b);
      cudaError_t e = cudaDeviceSynchronize();
      if (e != cudaSuccess) { printf("CUDA error %s\n", cudaGetErrorString(e)); return 1; }
      float ms = 0;
      cudaEventElapsedTime(&ms, a, b);
      if (rep < 2) continue;                    // the first launches warm L2
      const double recv = (double)total * CHUNK;
      printf("%s, mode %d (%s): %.3f ms -> each SM RECEIVES %.1f GB/s = %.1f B/clk at 1.9 GHz; L2 reads per SM %.1f GB/s\n",
             priv ? "a private 512 KB region per CTA / pair" : "one 3 MB set shared by all SMs", mode,
             mode ? "pairs, multicast" : "every CTA for itself", ms, recv / ms / 1e6, recv / ms / 1e6 / 1.9,
             recv / ms / 1e6 / (mode ? 2 : 1));
    }
  }
  return 0;
}
