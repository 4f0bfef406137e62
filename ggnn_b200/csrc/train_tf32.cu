// train_tf32.cu -- backward of the propagation path on the tensor cores (tcgen05 kind::tf32, FP32
// accumulators in TMEM).  Same recurrence and the same pointwise kernels as train_fp32.cu; the
// contractions change:
//   NT form  C = A @ B^T   ([dx|drh] = dzc @ Wc^T, [dx2|dh2] = [dzr,dzu] @ Wg^T, dh[src] += dx[tgt] @ W_e^T)
//            B^T of a row-major weight IS the K-major "transposed weight" operand of the forward
//            kernel, so these run on tf32.cu's gemm_tf32_kernel<MODE_NT> over re-padded copies of
//            the weights in their own layout (no transpose).
//   TN form  C += X^T @ D  (dWc, dWg, dW_e): the reduction runs over the ROWS of two row-major
//            matrices, so both operands are MN-major in shared memory: tn_tc_kernel below.
// Operands are TF32-rounded where they are produced (tape: forward kernels; dz*: pointwise kernels).
#include "plan.cuh"
#include "train.cuh"
#include "tf32.cuh"
#include "tc.cuh"

namespace tn {

constexpr int BMo = 256;                  // output rows per CTA (columns of X): two UMMA M = 128 halves
constexpr int BNo = 256;                  // max output columns per CTA (columns of D) = UMMA N
constexpr int RK = 32;                    // reduction rows per stage (4 MMAs of K = 8 per half)
constexpr int STAGES = 3;
constexpr int GROUP_BYTES = RK * 128;     // one 32-float M/N group of a stage: 32 rows x 128 B
constexpr int A_BYTES = (BMo / 32) * GROUP_BYTES;   // 32 KB
constexpr int B_BYTES = (BNo / 32) * GROUP_BYTES;   // 32 KB
constexpr int NPROD = 128;
constexpr int NTHREADS = NPROD + 32;
constexpr int TMEM_COLS = 512;            // two 128 x 256 FP32 accumulators: the whole tensor memory of the SM
constexpr int SMEM_BYTES = STAGES * (A_BYTES + B_BYTES) + 1024 + 256;

struct params {
  const float* X; int ldx; const int32_t* x_rowmap; int K1;   // X [R, K1], rows optionally gathered (-1 = zero)
  const float* D; int ldd; const int32_t* d_rowmap; int N;    // D [R, N]
  int bn;                                                     // output columns per CTA (multiple of 32, <= 256)
  int64_t R; int rows_per_cta;                                // multiple of 128
  float* C; int ldc;                                          // [K1, N] (+ type * c_type_stride when grouped)
  const int32_t* tile_type; const int32_t* n_tiles; int64_t c_type_stride;
};

// end of the run of message tiles that share the type of the tile at `seg_begin`
__device__ __forceinline__ int64_t segment_end(const params& p, int64_t seg_begin, int64_t r_end) {
  if (!p.tile_type) return r_end;
  int64_t t = seg_begin / GGNN_TILE_M;
  const int type = p.tile_type[t];
  ++t;
  while (t * GGNN_TILE_M < r_end && p.tile_type[t] == type) ++t;
  const int64_t e = t * GGNN_TILE_M;
  return e < r_end ? e : r_end;
}

__global__ void __launch_bounds__(NTHREADS, 1)
tn_tc_kernel(const params p) {
  extern __shared__ uint8_t smem_raw[];
  const int tid = threadIdx.x, warp = tid >> 5, lane = tid & 31;
  const int n0 = blockIdx.x * p.bn, k0 = blockIdx.y * BMo;
  int64_t r_begin = (int64_t)blockIdx.z * p.rows_per_cta, r_end = r_begin + p.rows_per_cta;
  int64_t R = p.R;
  if (p.tile_type) R = (int64_t)(*p.n_tiles) * GGNN_TILE_M;
  if (r_end > R) r_end = R;
  if (r_begin >= r_end) return;                         // uniform
  const int n_valid = min(p.bn, ((p.N - n0) + 31) & ~31);          // UMMA N of this CTA (multiple of 32)
  const int halves = (p.K1 - k0 > 128) ? 2 : 1;                     // active 128-row halves of the output tile
  const int a_groups = halves * 4, b_groups = n_valid >> 5;

  const uint32_t base = (tc::smem_u32(smem_raw) + 1023u) & ~1023u;
  const uint32_t sA = base, sB = base + STAGES * A_BYTES;
  const uint32_t bars = sB + STAGES * B_BYTES;
  auto full_bar = [&](int s) { return bars + 8u * s; };
  auto empty_bar = [&](int s) { return bars + 8u * (STAGES + s); };
  const uint32_t accum_bar = bars + 8u * (2 * STAGES);
  const uint32_t tmem_slot = bars + 8u * (2 * STAGES + 1);
  uint8_t* smem_al = smem_raw + (base - tc::smem_u32(smem_raw));

  if (warp == 4 && lane == 0) {
    for (int s = 0; s < STAGES; ++s) {
      tc::mbar_init(full_bar(s), NPROD);
      tc::mbar_init(empty_bar(s), 1);
    }
    tc::mbar_init(accum_bar, 1);
    tc::fence_barrier_init();
  }
  if (warp == 0) tc::tmem_alloc(tmem_slot, TMEM_COLS);
  tc::fence_before_sync();
  __syncthreads();
  tc::fence_after_sync();
  const uint32_t tmem_base = *reinterpret_cast<volatile uint32_t*>(smem_al + (tmem_slot - base));

  if (warp < 4) {
    // ===== producers: warp w fills M/N groups w and w + 4 (32 floats each) of both operands;
    //       lane = (row % 4, 16-byte chunk of the 128-byte row)
    const int c16 = lane & 7, rsub = lane >> 3;
    uint32_t off[8];
#pragma unroll
    for (int j = 0; j < 8; ++j) {
      const int rr = rsub + 4 * j;
      off[j] = rr * 128 + ((((c16 >> 1) ^ (rr & 3)) << 5) | ((c16 & 1) << 4));   // 32-byte-base swizzle
    }
    int it = 0, seg = 0;
    for (int64_t seg_begin = r_begin; seg_begin < r_end; ++seg) {
      const int64_t seg_end = segment_end(p, seg_begin, r_end);
      for (int64_t r0 = seg_begin; r0 < seg_end; r0 += RK, ++it) {
        const int s = it % STAGES;
        if (it >= STAGES) tc::mbar_wait(empty_bar(s), ((it / STAGES) - 1) & 1);
        const uint32_t a_st = sA + s * A_BYTES, b_st = sB + s * B_BYTES;
        int64_t xs[8], ds[8];
#pragma unroll
        for (int j = 0; j < 8; ++j) {
          const int64_t r = r0 + rsub + 4 * j;
          xs[j] = -1; ds[j] = -1;
          if (r < seg_end) {
            xs[j] = p.x_rowmap ? (int64_t)p.x_rowmap[r] : r;
            ds[j] = p.d_rowmap ? (int64_t)p.d_rowmap[r] : r;
          }
        }
#pragma unroll
        for (int gg = 0; gg < 2; ++gg) {
          const int g = warp + 4 * gg;
          if (g < a_groups) {
            const int xcol = k0 + g * 32 + c16 * 4;
            const bool in = xcol < p.K1;                           // K1 multiple of 4
#pragma unroll
            for (int j = 0; j < 8; ++j) {
              const bool ok = in && xs[j] >= 0;
              tc::cp_async16(a_st + g * GROUP_BYTES + off[j],
                             ok ? (const void*)(p.X + xs[j] * p.ldx + xcol) : (const void*)p.X, ok ? 16u : 0u);
            }
          }
          if (g < b_groups) {
            const int dcol = n0 + g * 32 + c16 * 4;
            const bool in = dcol < p.N;
#pragma unroll
            for (int j = 0; j < 8; ++j) {
              const bool ok = in && ds[j] >= 0;
              tc::cp_async16(b_st + g * GROUP_BYTES + off[j],
                             ok ? (const void*)(p.D + ds[j] * p.ldd + dcol) : (const void*)p.D, ok ? 16u : 0u);
            }
          }
        }
        tc::cp_async_arrive(full_bar(s));
      }
      // ===== epilogue of this segment: C[type] += accumulators (TMEM lane = output row)
      tc::mbar_wait(accum_bar, seg & 1);
      tc::fence_after_sync();
      const int type = p.tile_type ? p.tile_type[seg_begin / GGNN_TILE_M] : 0;
      for (int hf = 0; hf < halves; ++hf) {
        const int krow = k0 + hf * 128 + warp * 32 + lane;
        float* crow = p.C + (size_t)type * p.c_type_stride + (size_t)krow * p.ldc;
#pragma unroll 1
        for (int cc = 0; cc < b_groups; ++cc) {
          uint32_t acc[32];
          tc::tmem_ld32(tmem_base + ((uint32_t)(warp * 32) << 16) + (uint32_t)(hf * BNo + cc * 32), acc);
          tc::tmem_ld_wait();
          if (krow < p.K1) {
#pragma unroll
            for (int q = 0; q < 8; ++q) {
              const int n = n0 + cc * 32 + q * 4;
              if (n >= p.N) break;
              atomicAdd(reinterpret_cast<float4*>(crow + n),
                        make_float4(__uint_as_float(acc[q * 4 + 0]), __uint_as_float(acc[q * 4 + 1]),
                                    __uint_as_float(acc[q * 4 + 2]), __uint_as_float(acc[q * 4 + 3])));
            }
          }
        }
      }
      tc::fence_before_sync();       // TMEM reads done before the next segment's MMAs (ordered by full_bar)
      seg_begin = seg_end;
    }
  } else {
    // ===== MMA issuer
    if (lane == 0) {
      const uint32_t idesc = tc::umma_idesc_mn(/*TF32*/ 2, 128, n_valid);
      int it = 0;
      for (int64_t seg_begin = r_begin; seg_begin < r_end;) {
        const int64_t seg_end = segment_end(p, seg_begin, r_end);
        bool first = true;
        for (int64_t r0 = seg_begin; r0 < seg_end; r0 += RK, ++it) {
          const int s = it % STAGES;
          tc::mbar_wait(full_bar(s), (it / STAGES) & 1);
          tc::fence_proxy_async();                      // cp.async (generic proxy) writes -> async-proxy reads
          tc::fence_after_sync();
          const uint32_t a0 = sA + s * A_BYTES, b0 = sB + s * B_BYTES;
          for (int hf = 0; hf < halves; ++hf) {
#pragma unroll
            for (int k = 0; k < RK / 8; ++k) {
              tc::umma_tf32(tmem_base + hf * BNo,
                            tc::umma_desc_mn_sw128_32b(a0 + hf * 4 * GROUP_BYTES + k * 1024, GROUP_BYTES, 512),
                            tc::umma_desc_mn_sw128_32b(b0 + k * 1024, GROUP_BYTES, 512), idesc, !(first && k == 0));
            }
          }
          first = false;
          tc::umma_commit(empty_bar(s));
        }
        tc::umma_commit(accum_bar);
        seg_begin = seg_end;
      }
    }
    __syncwarp();
  }

  tc::fence_before_sync();
  __syncthreads();
  if (warp == 0) tc::tmem_dealloc(tmem_base, TMEM_COLS);
}

}  // namespace tn

int ggnn_tf32_tn(ggnn_ctx* ctx, const float* X, int ldx, const int32_t* x_rowmap, int K1, const float* D, int ldd,
                 const int32_t* d_rowmap, int N, int64_t R, int rows_per_cta, float* C, int ldc,
                 const int32_t* tile_type, const int32_t* n_tiles, int64_t c_type_stride, cudaStream_t st) {
  if (R <= 0 || K1 <= 0 || N <= 0) return GGNN_OK;
  GGNN_ENSURE_SMEM(ctx, GGNN_ATTR_TN, tn::tn_tc_kernel, tn::SMEM_BYTES);
  tn::params p{};
  p.X = X; p.ldx = ldx; p.x_rowmap = x_rowmap; p.K1 = K1;
  p.D = D; p.ldd = ldd; p.d_rowmap = d_rowmap; p.N = N;
  // balanced column tiles: as few as 256 allows, equal width rounded up to 32
  const int ntiles = (N + tn::BNo - 1) / tn::BNo;
  p.bn = (int)align_up((size_t)((N + ntiles - 1) / ntiles), 32);
  const int ktiles = (K1 + tn::BMo - 1) / tn::BMo;
  const int gx = (N + p.bn - 1) / p.bn;
  if (rows_per_cta <= 0) {          // ungrouped: about two waves of equal CTAs
    const int64_t chunks = (2 * (int64_t)ctx->sm_count) / ((int64_t)gx * ktiles) > 0 ? (2 * (int64_t)ctx->sm_count) / ((int64_t)gx * ktiles) : 1;
    rows_per_cta = (int)align_up((size_t)ceil_div64(R, chunks), 128);
  }
  p.R = R; p.rows_per_cta = rows_per_cta; p.C = C; p.ldc = ldc;
  p.tile_type = tile_type; p.n_tiles = n_tiles; p.c_type_stride = c_type_stride;
  dim3 grid(gx, ktiles, (unsigned)ceil_div64(R, rows_per_cta));
  tn::tn_tc_kernel<<<grid, tn::NTHREADS, tn::SMEM_BYTES, st>>>(p);
  GGNN_LAUNCHED(ctx);
  return GGNN_OK;
}
