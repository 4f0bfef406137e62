// s9_stream.cu -- evidence for the state-resident design of north_star item 4 / SURVEY 7.1 S9.
// A persistent CTA that keeps ONE graph tile's node states in shared memory across all T steps has to bring, every
// step, the weights its graph uses to its SM: a C3 tree (127 words, Zipf labels over 45 types, forward + reverse +
// two word-order channels) uses ~40-60 distinct W_e of 32 KB (FP16 128x128) plus the 192 KB of GRU weights.
// This probe measures the one thing that bounds that design from below: how fast ALL 148 SMs together can stream
// such a weight set out of L2 into shared memory (cp.async.bulk ring, nothing else running, no MMA, no epilogue),
// as microseconds per graph-step, for several set sizes.  The shipped per-timestep kernels spend
//   (msg 0.71 ms + gru 2.72 ms) / 8 steps / (4096 graphs / 148 SMs) = 15.5 us per graph-step per SM in total.
//   nvcc -gencode arch=compute_100a,code=sm_100a -O3 -std=c++17 scripts/micro/s9_stream.cu -o scripts/micro/s9_stream
#include <cstdio>
#include <cstdlib>
#include <cstdint>
#include <cuda_runtime.h>

constexpr int CHUNK = 16384;      // bytes per bulk copy
constexpr int STAGES = 8;

__device__ __forceinline__ uint32_t smem_u32(const void* p) { return (uint32_t)__cvta_generic_to_shared(p); }

__global__ void __launch_bounds__(128, 1)
stream_kernel(const uint8_t* __restrict__ weights, int n_types_total, int types_per_step, int steps, int graphs_per_cta,
              unsigned long long* out_clk) {
  extern __shared__ __align__(1024) uint8_t smem[];
  __shared__ uint64_t full[STAGES];
  const int tid = threadIdx.x;
  if (tid == 0) {
    for (int s = 0; s < STAGES; ++s) asm volatile("mbarrier.init.shared::cta.b64 [%0], 1;" ::"r"(smem_u32(&full[s])));
    asm volatile("fence.mbarrier_init.release.cluster;" ::: "memory");
  }
  __syncthreads();
  const int chunks_per_type = 32768 / CHUNK;
  const long long total = (long long)graphs_per_cta * steps * (types_per_step + 6) * chunks_per_type;   // + 192 KB GRU
  unsigned long long t0 = clock64();
  if (tid == 0) {
    // issue up to STAGES ahead, wait in order; "consumption" is just the wait (an MMA would read the slot here)
    long long issued = 0, done = 0;
    unsigned rng = 12345u + blockIdx.x * 7919u;
    while (done < total) {
      while (issued < total && issued - done < STAGES) {
        const int s = (int)(issued % STAGES);
        rng = rng * 1664525u + 1013904223u;
        const int type = (int)((rng >> 8) % (unsigned)n_types_total);       // which W_e: spread over the whole 3 MB set
        const uint8_t* src = weights + (size_t)type * 32768 + (size_t)(issued % chunks_per_type) * CHUNK;
        asm volatile("mbarrier.arrive.expect_tx.shared::cta.b64 _, [%0], %1;" ::"r"(smem_u32(&full[s])), "r"(CHUNK) : "memory");
        asm volatile("cp.async.bulk.shared::cluster.global.mbarrier::complete_tx::bytes [%0], [%1], %2, [%3];"
                     ::"r"(smem_u32(smem + s * CHUNK)), "l"(src), "r"(CHUNK), "r"(smem_u32(&full[s])) : "memory");
        ++issued;
      }
      const int s = (int)(done % STAGES);
      const unsigned parity = (unsigned)((done / STAGES) & 1);
      asm volatile(
          "{\n\t.reg .pred p;\n\tW: mbarrier.try_wait.parity.shared::cta.b64 p, [%0], %1, 2000;\n\t@p bra D;\n\tbra W;\n\tD:\n\t}"
          ::"r"(smem_u32(&full[s])), "r"(parity) : "memory");
      ++done;
    }
  }
  __syncthreads();
  if (tid == 0) out_clk[blockIdx.x] = clock64() - t0;
}

int main() {
  int dev = 0, sms = 0;
  cudaDeviceGetAttribute(&sms, cudaDevAttrMultiProcessorCount, dev);
  const int n_types = 92;
  uint8_t* w;
  cudaMalloc(&w, (size_t)(n_types + 8) * 32768);
  cudaMemset(w, 1, (size_t)(n_types + 8) * 32768);
  unsigned long long* clk;
  cudaMalloc(&clk, sms * sizeof(unsigned long long));
  const int smem_bytes = STAGES * CHUNK;
  cudaFuncSetAttribute(stream_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, smem_bytes);
  const int steps = 8, graphs_per_cta = 8;
  printf("SMs %d; %d-slot ring of %d KB bulk copies; weights 92 x 32 KB FP16 resident in L2\n", sms, STAGES, CHUNK / 1024);
  for (int types : {20, 40, 60, 92}) {
    cudaEvent_t a, b;
    cudaEventCreate(&a); cudaEventCreate(&b);
    stream_kernel<<<sms, 128, smem_bytes>>>(w, n_types, types, steps, graphs_per_cta, clk);     // warm-up (fills L2)
    cudaEventRecord(a);
    stream_kernel<<<sms, 128, smem_bytes>>>(w, n_types, types, steps, graphs_per_cta, clk);
    cudaEventRecord(b);
    cudaError_t e = cudaDeviceSynchronize();
    if (e != cudaSuccess) { printf("CUDA error %s\n", cudaGetErrorString(e)); return 1; }
    float ms = 0;
    cudaEventElapsedTime(&ms, a, b);
    const double bytes_step = (types + 6) * 32768.0;
    const double us_step = ms * 1e3 / (steps * graphs_per_cta);
    printf("distinct W_e per graph-step %2d (+192 KB GRU): %7.1f KB per graph-step -> %6.2f us per graph-step per SM, "
           "%5.1f GB/s per SM, %5.2f TB/s over the chip\n",
           types, bytes_step / 1024, us_step, bytes_step / us_step / 1e3, bytes_step * sms / us_step / 1e6);
  }
  return 0;
}
