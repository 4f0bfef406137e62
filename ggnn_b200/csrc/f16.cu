// f16.cu -- GGNN_PREC_F16: the TF32-mode pipeline with every tensor-core operand held in a 16-bit
// container.  An FP16 value has the same 11-bit significand as a TF32 operand, so as long as
// nothing leaves the FP16 range (checked by the caller, see include/ggnn_b200.h) the products
// entering the FP32 accumulators are the ones the TF32 mode forms -- but the node states, the
// gathered message operands and the streamed weights move half the bytes and the tensor pipe runs
// kind::f16 at twice the kind::tf32 rate.
//
//   msg_f16_kernel   M[r,:] = f16(h[msg_src[r],:]) @ f16(W[type]) + B[type]  -> FP16 rows
//                    (persistent grouped GEMM, see msg_tc.cu; 4 x 32 KB gather stages)
//   gru_f16_kernel   aggregation + GRU cell per 128-node tile (see gru_fused.cu), reset gate first:
//                    only [x,h] @ Wg[:, :h] precedes the first epilogue pass; the update-gate and
//                    candidate contractions of x run on the tensor pipe underneath it; the x tile
//                    is double-buffered so the aggregation of the next tile never waits for them.
#include "f16.cuh"
#include "tf32.cuh"
#include "tc.cuh"
#include <cuda_fp16.h>
#include <stdlib.h>

namespace {

constexpr int HP = 128;
constexpr int BM = 128;
constexpr int BKH = 64;                           // halves per K chunk = one 128-byte swizzle row
constexpr int CHUNK_BYTES = BM * BKH * 2;         // 16 KB: one [128 x 64] FP16 K-major chunk
constexpr int TILE16_BYTES = 2 * CHUNK_BYTES;     // 32 KB: [128 x 128] FP16

// ---- packing ----------------------------------------------------------------------------
// dst[e][n][k] (n, k < 128) = f16(src[e][k][n]) for k, n < h, else 0
__global__ void pack_edge16_kernel(const float* __restrict__ W, __half* __restrict__ Wt, int two_e, int h) {
  const int64_t total = (int64_t)two_e * HP * HP;
  for (int64_t i = (int64_t)blockIdx.x * blockDim.x + threadIdx.x; i < total; i += (int64_t)gridDim.x * blockDim.x) {
    const int k = (int)(i % HP), n = (int)((i / HP) % HP);
    const int64_t e = i / ((int64_t)HP * HP);
    Wt[i] = __float2half_rn((k < h && n < h) ? W[(e * h + k) * h + n] : 0.f);
  }
}
// GRU kernels: src [2h, ncols*h] -> dst [ncols*128, 256]: dst[nb*128+n][kb*128+k] = f16(src[kb*h+k][nb*h+n])
__global__ void pack_gru16_kernel(const float* __restrict__ src, __half* __restrict__ dst, int ncols, int h) {
  const int Kp = 2 * HP, Np = ncols * HP;
  const int64_t total = (int64_t)Np * Kp;
  for (int64_t i = (int64_t)blockIdx.x * blockDim.x + threadIdx.x; i < total; i += (int64_t)gridDim.x * blockDim.x) {
    const int kk = (int)(i % Kp), nn = (int)(i / Kp);
    const int kb = kk / HP, k = kk % HP, nb = nn / HP, n = nn % HP;
    dst[i] = __float2half_rn((k < h && n < h) ? src[(size_t)(kb * h + k) * (ncols * h) + nb * h + n] : 0.f);
  }
}
// state16[n][c] = f16(h[n][c]) (c < h) or 0; one thread per 4 columns.  Also max|h| -> *absmax (float bits).
__global__ void state_in_kernel(const float* __restrict__ hs, __half* __restrict__ s16, int64_t N, int h,
                                unsigned* __restrict__ absmax) {
  const int64_t total = N * (HP / 4);
  float m = 0.f;
  for (int64_t i = (int64_t)blockIdx.x * blockDim.x + threadIdx.x; i < total; i += (int64_t)gridDim.x * blockDim.x) {
    const int64_t n = i / (HP / 4);
    const int c = (int)(i % (HP / 4)) * 4;
    float4 v = make_float4(0.f, 0.f, 0.f, 0.f);
    if (c < h) v = *reinterpret_cast<const float4*>(hs + n * h + c);          // h % 4 == 0
    m = fmaxf(fmaxf(m, fmaxf(fabsf(v.x), fabsf(v.y))), fmaxf(fabsf(v.z), fabsf(v.w)));
    __half2 lo = __floats2half2_rn(v.x, v.y), hi = __floats2half2_rn(v.z, v.w);
    uint2 pk;
    pk.x = *reinterpret_cast<uint32_t*>(&lo); pk.y = *reinterpret_cast<uint32_t*>(&hi);
    *reinterpret_cast<uint2*>(s16 + n * HP + c) = pk;
  }
#pragma unroll
  for (int o = 16; o > 0; o >>= 1) m = fmaxf(m, __shfl_xor_sync(0xffffffffu, m, o));
  if (absmax && (threadIdx.x & 31) == 0 && m > 0.f) atomicMax(absmax, __float_as_uint(m));
}
// ---- word-order channels folded through the GRU's input weights (gru16::gru_f16_kernel) ------------------------
// F[d][n][k] = f16( (W[e_d] @ [Wg[:h, :2h] | Wc[:h, :h]])[k][n] ),  e_0 = E-1 (t-1 -> t), e_1 = 2E-1 (t+1 -> t);
// n = 128-column blocks r | u | cand, zero padded like the other packs.  One CTA per (d, n), thread = k.
__global__ void fold_kernel(const float* __restrict__ W, const float* __restrict__ Wg, const float* __restrict__ Wc,
                            __half* __restrict__ F, int two_e, int h) {
  const int d = blockIdx.x / (3 * HP), n = blockIdx.x % (3 * HP), k = threadIdx.x;
  const int e = d == 0 ? two_e / 2 - 1 : two_e - 1;
  const int nb = n / HP, nn = n % HP;
  float acc = 0.f;
  if (k < h && nn < h) {
    const float* wrow = W + ((size_t)e * h + k) * h;
    for (int j = 0; j < h; ++j) {
      const float wx = nb < 2 ? Wg[(size_t)j * 2 * h + nb * h + nn] : Wc[(size_t)j * h + nn];
      acc = fmaf(wrow[j], wx, acc);
    }
  }
  F[((size_t)d * 3 * HP + n) * HP + k] = __float2half_rn(acc);
}
// scaled bias vectors of a word-order tile [3][384]: inner rows (both word-order messages), first row (t+1 -> t
// only), last row (t-1 -> t only); scale = -log2e for the gates, 2 log2e for the candidate (the epilogue's forms)
__global__ void fold_bias_kernel(const float* __restrict__ B, const float* __restrict__ Wg, const float* __restrict__ Wc,
                                 const float* __restrict__ bg, const float* __restrict__ bc, float* __restrict__ out,
                                 int two_e, int h) {
  const int n = blockIdx.x * blockDim.x + threadIdx.x;
  if (n >= 3 * HP) return;
  const int nb = n / HP, nn = n % HP;
  float plain = 0.f, dp = 0.f, dn = 0.f;
  if (nn < h) {
    plain = nb < 2 ? bg[nb * h + nn] : bc[nn];
    if (B) {
      const float* bp = B + (size_t)(two_e / 2 - 1) * h;
      const float* bn = B + (size_t)(two_e - 1) * h;
      for (int j = 0; j < h; ++j) {
        const float wx = nb < 2 ? Wg[(size_t)j * 2 * h + nb * h + nn] : Wc[(size_t)j * h + nn];
        dp = fmaf(bp[j], wx, dp);
        dn = fmaf(bn[j], wx, dn);
      }
    }
  }
  const float sc = nb < 2 ? -1.4426950408889634f : 2.0f * 1.4426950408889634f;
  out[n] = sc * (plain + dp + dn);
  out[3 * HP + n] = sc * (plain + dn);
  out[6 * HP + n] = sc * (plain + dp);
}

__global__ void absmax_kernel(const float* __restrict__ x, int64_t n, unsigned* __restrict__ out) {
  float m = 0.f;
  for (int64_t i = (int64_t)blockIdx.x * blockDim.x + threadIdx.x; i < n; i += (int64_t)gridDim.x * blockDim.x)
    m = fmaxf(m, fabsf(x[i]));                       // NaN is dropped by fmaxf; Inf propagates
#pragma unroll
  for (int o = 16; o > 0; o >>= 1) m = fmaxf(m, __shfl_xor_sync(0xffffffffu, m, o));
  if ((threadIdx.x & 31) == 0 && m > 0.f) atomicMax(out, __float_as_uint(m));
}

}  // namespace

// =========================================================================================
// message transform (persistent grouped GEMM, FP16 operands) -- structure of msg_tc.cu
// =========================================================================================
namespace msg16 {

constexpr int ASTAGES = 4;
constexpr int A_BYTES = TILE16_BYTES;           // 32 KB per stage
constexpr int W_BYTES = TILE16_BYTES;           // 32 KB
constexpr int O_BYTES = 2 * TILE16_BYTES;       // 2 x 32 KB FP16 staging (two 64-column boxes each): the epilogue of tile i+1
                                                // fills one buffer while the TMA store of tile i still reads the other
constexpr int NPROD = 128;
constexpr int WARP_MMA = 4;
constexpr int WARP_EPI0 = 5;
constexpr int NTHREADS = 9 * 32;
constexpr int SMEM_BYTES = ASTAGES * A_BYTES + W_BYTES + O_BYTES + 1024 /*align*/ + 256 /*barriers*/ + 512 /*bias*/;

enum {
  BAR_A_FULL = 0,                       // [ASTAGES] 128 cp.async arrivals
  BAR_A_EMPTY = BAR_A_FULL + ASTAGES,   // [ASTAGES] tcgen05.commit
  BAR_W_FULL = BAR_A_EMPTY + ASTAGES,   // TMA bytes
  BAR_ACC_FULL,                         // [2]
  BAR_ACC_EMPTY = BAR_ACC_FULL + 2,     // [2] 4 epilogue warps
  BAR_COUNT = BAR_ACC_EMPTY + 2
};

struct params {
  const __half* state16;      // [N, 128]
  const int32_t* msg_src;     // [tiles*128] source row per message row, -1 = padding
  const int32_t* tile_type;   // [tiles]
  const int32_t* n_tiles;
  const float* bias;          // [two_e, h] or null
  int h;
};

__global__ void __launch_bounds__(NTHREADS, 1)
msg_f16_kernel(const params p, const __grid_constant__ CUtensorMap tmW,
               const __grid_constant__ CUtensorMap tmOut) {
  extern __shared__ uint8_t smem_raw[];
  const int tid = threadIdx.x, warp = tid >> 5, lane = tid & 31;
  const uint32_t base = (tc::smem_u32(smem_raw) + 1023u) & ~1023u;
  uint8_t* smem_al = smem_raw + (base - tc::smem_u32(smem_raw));
  const uint32_t sA = base;
  const uint32_t sW = base + ASTAGES * A_BYTES;
  const uint32_t sO = sW + W_BYTES;
  const uint32_t bars = sO + O_BYTES;
  auto bar = [&](int i) { return bars + 8u * i; };
  const uint32_t tmem_slot = bars + 8u * BAR_COUNT;
  float* s_bias = reinterpret_cast<float*>(smem_al + (bars + 256 - base));   // [128]

  const int n_tiles = *p.n_tiles;
  const int per = (n_tiles + (int)gridDim.x - 1) / (int)gridDim.x;
  const int t_begin = min(n_tiles, (int)blockIdx.x * per);
  const int t_end = min(n_tiles, t_begin + per);
  const int my_tiles = t_end - t_begin;

  if (warp == WARP_MMA && lane == 0) {
    for (int s = 0; s < ASTAGES; ++s) {
      tc::mbar_init(bar(BAR_A_FULL + s), NPROD);
      tc::mbar_init(bar(BAR_A_EMPTY + s), 1);
    }
    tc::mbar_init(bar(BAR_W_FULL), 1);
    for (int s = 0; s < 2; ++s) {
      tc::mbar_init(bar(BAR_ACC_FULL + s), 1);
      tc::mbar_init(bar(BAR_ACC_EMPTY + s), 4);
    }
    tc::fence_barrier_init();
  }
  if (warp == 0) {
    if (lane == 0) { tc::tma_prefetch_desc(&tmW); tc::tma_prefetch_desc(&tmOut); }
    __syncwarp();
    tc::tmem_alloc(tmem_slot, 256);
  }
  tc::fence_before_sync();
  __syncthreads();
  tc::fence_after_sync();
  const uint32_t tmem_base = *reinterpret_cast<volatile uint32_t*>(smem_al + (tmem_slot - base));

  if (warp < 4) {
    // ---- gather producers: a 256-byte state row = 16 slots of 16 bytes; 16 lanes per row
    const int c16 = tid & 15;                 // slot within the row: K chunk c16 >> 3, slot c16 & 7
    const int rbase = tid >> 4;               // rows rbase + 8*j, j = 0..15
    const uint32_t kc_off = (c16 >> 3) * CHUNK_BYTES;
    int src[16], nsrc[16];
    auto load_idx = [&](int tile, int (&dst)[16]) {
#pragma unroll
      for (int j = 0; j < 16; ++j) dst[j] = __ldg(p.msg_src + (size_t)tile * BM + rbase + 8 * j);
    };
    if (my_tiles > 0) load_idx(t_begin, src);
    int cur_type = -1;
    for (int i = 0; i < my_tiles; ++i) {
      const int tile = t_begin + i;
      const int s = i % ASTAGES;
      if (i + 1 < my_tiles) load_idx(tile + 1, nsrc);
      if (i >= ASTAGES) tc::mbar_wait(bar(BAR_A_EMPTY + s), ((i / ASTAGES) - 1) & 1);
      if (tid == 0) {
        const int type = __ldg(p.tile_type + tile);
        if (type != cur_type) {
          // every MMA reading the old W must have retired: tiles <= i-ASTAGES are known complete
          // (stage wait above); wait for the ones still in flight on their own stage barriers
          for (int d = 1; d < ASTAGES && d <= i; ++d)
            tc::mbar_wait(bar(BAR_A_EMPTY + (i - d) % ASTAGES), ((i - d) / ASTAGES) & 1);
          tc::mbar_arrive_expect_tx(bar(BAR_W_FULL), W_BYTES);
          tc::tma_load_2d(sW, &tmW, 0, type * HP, bar(BAR_W_FULL));
          tc::tma_load_2d(sW + CHUNK_BYTES, &tmW, BKH, type * HP, bar(BAR_W_FULL));
          cur_type = type;
        }
      }
      const uint32_t a_st = sA + s * A_BYTES + kc_off;
#pragma unroll
      for (int j = 0; j < 16; ++j) {
        const int r = rbase + 8 * j;
        const uint32_t off = (r >> 3) * 1024 + (r & 7) * 128 + (((c16 & 7) ^ (r & 7)) << 4);
        const bool ok = src[j] >= 0;
        tc::cp_async16(a_st + off, ok ? (const void*)(p.state16 + (size_t)src[j] * HP + c16 * 8) : (const void*)p.state16,
                       ok ? 16u : 0u);
      }
      tc::cp_async_arrive(bar(BAR_A_FULL + s));
#pragma unroll
      for (int j = 0; j < 16; ++j) src[j] = nsrc[j];
    }
  } else if (warp == WARP_MMA) {
    if (lane == 0) {
      const uint32_t idesc = tc::umma_idesc(0 /*F16*/, BM, HP);
      int cur_type = -1, w_gen = 0;
      for (int i = 0; i < my_tiles; ++i) {
        const int s = i % ASTAGES, buf = i & 1;
        const int type = __ldg(p.tile_type + t_begin + i);
        if (type != cur_type) {
          tc::mbar_wait(bar(BAR_W_FULL), w_gen & 1);
          cur_type = type;
          ++w_gen;
        }
        if (i >= 2) tc::mbar_wait(bar(BAR_ACC_EMPTY + buf), ((i >> 1) - 1) & 1);
        tc::mbar_wait(bar(BAR_A_FULL + s), (i / ASTAGES) & 1);
        tc::fence_after_sync();
        const uint32_t a0 = sA + s * A_BYTES, d = tmem_base + buf * HP;
#pragma unroll
        for (int kc = 0; kc < 2; ++kc)
#pragma unroll
          for (int k = 0; k < 4; ++k)                 // UMMA K = 16 halves = 32 bytes
            tc::umma_f16(d, tc::umma_desc_sw128(a0 + kc * CHUNK_BYTES + k * 32),
                         tc::umma_desc_sw128(sW + kc * CHUNK_BYTES + k * 32), idesc, (kc | k) != 0);
        tc::umma_commit(bar(BAR_A_EMPTY + s));
        tc::umma_commit(bar(BAR_ACC_FULL + buf));
      }
    }
    __syncwarp();
  } else {
    // ---- epilogue (warps 5-8): identical to msg_tc.cu
    const int q = warp & 3;
    const int r = q * 32 + lane;
    const int etid = tid - WARP_EPI0 * 32;
    const uint32_t t_lane = tmem_base + ((uint32_t)(q * 32) << 16);
    int cur_type = -1;
    for (int i = 0; i < my_tiles; ++i) {
      const int tile = t_begin + i, buf = i & 1;
      const int type = __ldg(p.tile_type + tile);
      if (type != cur_type) {
        tc::named_bar_sync(2, 128);
        s_bias[etid] = (p.bias && etid < p.h) ? __ldg(p.bias + (size_t)type * p.h + etid) : 0.f;
        cur_type = type;
      }
      tc::mbar_wait(bar(BAR_ACC_FULL + buf), (i >> 1) & 1);
      tc::fence_after_sync();
      if (etid == 0) asm volatile("cp.async.bulk.wait_group.read 1;" ::: "memory");   // the store of tile i-2 has read this buffer
      tc::named_bar_sync(1, 128);
      const uint32_t so_buf = sO + (uint32_t)(i & 1) * TILE16_BYTES;
      uint8_t* orow = smem_al + (so_buf - base) + r * 128;
#pragma unroll
      for (int cc = 0; cc < 4; ++cc) {
        uint32_t acc[32];
        tc::tmem_ld32(t_lane + buf * HP + cc * 32, acc);
        tc::tmem_ld_wait();
        uint8_t* obox = orow + (cc >> 1) * CHUNK_BYTES;
#pragma unroll
        for (int g = 0; g < 4; ++g) {
          const int col = cc * 32 + g * 8;
          const float4 b0 = *reinterpret_cast<const float4*>(s_bias + col);
          const float4 b1 = *reinterpret_cast<const float4*>(s_bias + col + 4);
          __half2 h0 = __floats2half2_rn(__uint_as_float(acc[g * 8 + 0]) + b0.x, __uint_as_float(acc[g * 8 + 1]) + b0.y);
          __half2 h1 = __floats2half2_rn(__uint_as_float(acc[g * 8 + 2]) + b0.z, __uint_as_float(acc[g * 8 + 3]) + b0.w);
          __half2 h2 = __floats2half2_rn(__uint_as_float(acc[g * 8 + 4]) + b1.x, __uint_as_float(acc[g * 8 + 5]) + b1.y);
          __half2 h3 = __floats2half2_rn(__uint_as_float(acc[g * 8 + 6]) + b1.z, __uint_as_float(acc[g * 8 + 7]) + b1.w);
          uint4 pk;
          pk.x = *reinterpret_cast<uint32_t*>(&h0); pk.y = *reinterpret_cast<uint32_t*>(&h1);
          pk.z = *reinterpret_cast<uint32_t*>(&h2); pk.w = *reinterpret_cast<uint32_t*>(&h3);
          const int chunk = (cc & 1) * 4 + g;
          *reinterpret_cast<uint4*>(obox + ((chunk ^ (r & 7)) << 4)) = pk;
        }
      }
      tc::fence_before_sync();
      __syncwarp();
      if (lane == 0) tc::mbar_arrive(bar(BAR_ACC_EMPTY + buf));
      tc::fence_proxy_async();
      tc::named_bar_sync(1, 128);
      if (etid == 0) {
        tc::tma_store_2d(&tmOut, 0, tile * BM, so_buf);
        tc::tma_store_2d(&tmOut, 64, tile * BM, so_buf + CHUNK_BYTES);
        tc::tma_store_commit();
      }
    }
    if (etid == 0) asm volatile("cp.async.bulk.wait_group 0;" ::: "memory");
  }

  tc::fence_before_sync();
  __syncthreads();
  if (warp == 0) tc::tmem_dealloc(tmem_base, 256);
}

}  // namespace msg16

// =========================================================================================
// aggregation + GRU cell, one persistent kernel per timestep (FP16 operands) -- roles of gru_fused.cu
// =========================================================================================
namespace gru16 {

constexpr float kLog2e = 1.4426950408889634f;
constexpr int WSTAGES = 7;                               // weight ring: 7 x 16 KB [128 n x 64 k] slots
constexpr int N_AGG_WARPS = 6;                           // rows of a tile: 22, 22, 21, 21, 21, 21
constexpr int N_EPI_WARPS = 8;
constexpr int WARP_EPI0 = 0;                             // 0..7 (warp % 4 = TMEM lane quarter)
constexpr int WARP_MMA = 8;
constexpr int WARP_TMA = 9;
constexpr int WARP_AGG0 = 10;                            // 10..15
// 16 warps = 4 per scheduler partition (the register file is split four ways: 16 384 registers per partition):
// 4 x 32 x 96 = 12 288 registers used per partition, 4 096 free, and > 1 KB of shared memory left over -- one
// background CTA of the adjacency bitmask queue (compact.cu: 12 warps x 32 x 40 registers, 3 warps per partition)
// fits next to this kernel, so the next batch's adjacency streams from HBM while this kernel waits on its
// latency chains.  (With 18 or 20 warps one partition holds 5 warps and has no room for a single queue warp.)
constexpr int NTHREADS = (WARP_AGG0 + N_AGG_WARPS) * 32; // 512
// The h tile sits between 1 KB gaps that stay ZERO: an A-operand descriptor that starts one 128-byte row below /
// above a 128B-swizzled chunk makes MMA row r read tile row r-1 / r+1 (the swizzle is a function of the absolute
// shared-memory address: scripts/micro/desc_shift.cu), and the row that falls off the tile reads the gap.
constexpr int HGAP = 1024;
constexpr int HCHUNK_STRIDE = CHUNK_BYTES + HGAP;
constexpr int H_BYTES = 2 * HCHUNK_STRIDE + HGAP;        // gap | chunk 0 | gap | chunk 1 | gap
constexpr int NB = 3 * HP;                               // bias vector: r | u | cand
constexpr int SMEM_BYTES = 2 * TILE16_BYTES + H_BYTES + WSTAGES * CHUNK_BYTES + 256 /*barriers*/ +
                           4 * NB * 4 /*biases: plain, word-order tile: inner rows, first row, last row*/;   // 222 464

enum {
  BAR_W_FULL = 0,                    // [WSTAGES]
  BAR_W_EMPTY = BAR_W_FULL + WSTAGES,
  BAR_X_FULL = BAR_W_EMPTY + WSTAGES,   // [2]
  BAR_X_EMPTY = BAR_X_FULL + 2,         // [2] every MMA reading this x buffer (x, later r*h) has retired
  BAR_H_FULL = BAR_X_EMPTY + 2,
  BAR_H_EMPTY,
  BAR_R_DONE,                        // r accumulator complete
  BAR_X_READ,                        // the MMAs reading x of this tile have retired: the buffer may take r*h
  BAR_UC_DONE,                       // u accumulator complete
  BAR_RH_FULL,
  BAR_CAND_DONE,
  BAR_TMEM_EMPTY,
  BAR_COUNT
};
static_assert(8 * BAR_COUNT + 8 <= 256, "barrier area");

struct params {
  const __half* msg;        // [*, 128] FP16 message rows
  const int32_t* tptr;      // [N+1]
  const int32_t* tmsg;      // [nnz]
  const int32_t* regular;   // [n_tiles] 1 = word-order tile (see plan_bits.cu) or null
  const float* fold_bias;   // [3][384] scaled biases of a word-order tile: inner rows, first row, last row (or null)
  float* h_out;             // [N, h] FP32 result (written on the last step only) or null
  const float* bg_pad;      // [256]
  const float* bc;          // [h]
  int64_t N;
  int h;
  int n_tiles;
};

// byte offset of 16-byte slot `slot` (0..7) of row r inside one 128B-swizzled [128 x 64] FP16 chunk
__device__ __forceinline__ uint32_t sw_off(int r, int slot) {
  return (r >> 3) * 1024 + (r & 7) * 128 + ((slot ^ (r & 7)) << 4);
}

// one aggregated row (this lane's 4 columns) -> FP16 -> its 8 bytes of the swizzled x tile
__device__ __noinline__ void emit_row(uint8_t* xtile, int r, int lane, float4 acc) {
  const int kc = lane >> 4, l16 = lane & 15;
  __half2 lo = __floats2half2_rn(acc.x, acc.y), hi = __floats2half2_rn(acc.z, acc.w);
  uint2 pk;
  pk.x = *reinterpret_cast<uint32_t*>(&lo); pk.y = *reinterpret_cast<uint32_t*>(&hi);
  *reinterpret_cast<uint2*>(xtile + kc * CHUNK_BYTES + sw_off(r, l16 >> 1) + (l16 & 1) * 8) = pk;
}

__device__ __forceinline__ void unpack8(const uint4 v, float (&f)[8]) {
  const __half2* hp = reinterpret_cast<const __half2*>(&v);
#pragma unroll
  for (int i = 0; i < 4; ++i) { const float2 t = __half22float2(hp[i]); f[2 * i] = t.x; f[2 * i + 1] = t.y; }
}
__device__ __forceinline__ uint4 pack8(const float (&f)[8]) {
  uint4 v;
  __half2* hp = reinterpret_cast<__half2*>(&v);
#pragma unroll
  for (int i = 0; i < 4; ++i) hp[i] = __floats2half2_rn(f[2 * i], f[2 * i + 1]);
  return v;
}

// Per 128-node tile:
//   x      = sum of the incoming FP16 message rows (aggregation warps, fixed order)            -> x buffer
//   r      = sigmoid([x, h] @ Wg[:, :h] + bg_r);  r*h (FP16) overwrites the x buffer once its readers have retired
//   u      = sigmoid([x, h] @ Wg[:, h:] + bg_u);  cand = tanh(x @ Wc[:h] + (r*h) @ Wc[h:] + bc)
//   h'     = cand + u * (h - cand)                                                              -> h tile -> TMA store
// Word-order tile (params::regular): the messages of channels E-1 (t-1 -> t) and 2E-1 (t+1 -> t) never exist as
// rows; their contribution to the three pre-activations is formed from the RESIDENT h tile with row-shifted
// A descriptors and the folded weights F = W_channel @ [Wg[:h] | Wc[:h]] (packed by fold_kernel).
__global__ void __maxnreg__(96)      // 4 warps x 32 x 96 registers per scheduler partition (see NTHREADS)
gru_f16_kernel(const params p, const __grid_constant__ CUtensorMap tmH, const __grid_constant__ CUtensorMap tmWg,
               const __grid_constant__ CUtensorMap tmWc, const __grid_constant__ CUtensorMap tmF) {
  extern __shared__ __align__(1024) uint8_t smem_raw[];
  const int tid = threadIdx.x, warp = tid >> 5, lane = tid & 31;

  const uint32_t base = tc::smem_u32(smem_raw);          // 1024-byte aligned (__align__ on the dynamic array)
  uint8_t* smem_al = smem_raw;
  const uint32_t sX = base;                              // two x tiles
  const uint32_t sHg = base + 2 * TILE16_BYTES;          // gap | h chunk 0 | gap | h chunk 1 | gap
  const uint32_t sW = sHg + H_BYTES;                     // weight ring
  const uint32_t bars = sW + WSTAGES * CHUNK_BYTES;
  auto bar = [&](int i) { return bars + 8u * i; };
  auto sHc = [&](int kc) { return sHg + HGAP + (uint32_t)kc * HCHUNK_STRIDE; };   // K chunk kc of the h tile
  const uint32_t tmem_slot = bars + 8u * BAR_COUNT;
  float* s_bias = reinterpret_cast<float*>(smem_al + (bars + 256 - base));   // [4][384]: -log2e*bg (256) | 2log2e*bc (128)

  if (warp == WARP_MMA && lane == 0) {
    for (int s = 0; s < WSTAGES; ++s) {
      tc::mbar_init(bar(BAR_W_FULL + s), 1);
      tc::mbar_init(bar(BAR_W_EMPTY + s), 1);
    }
    for (int s = 0; s < 2; ++s) {
      tc::mbar_init(bar(BAR_X_FULL + s), N_AGG_WARPS);
      tc::mbar_init(bar(BAR_X_EMPTY + s), 1);
    }
    tc::mbar_init(bar(BAR_H_FULL), 1);
    tc::mbar_init(bar(BAR_H_EMPTY), 1);
    tc::mbar_init(bar(BAR_R_DONE), 1);
    tc::mbar_init(bar(BAR_X_READ), 1);
    tc::mbar_init(bar(BAR_UC_DONE), 1);
    tc::mbar_init(bar(BAR_RH_FULL), N_EPI_WARPS);
    tc::mbar_init(bar(BAR_CAND_DONE), 1);
    tc::mbar_init(bar(BAR_TMEM_EMPTY), N_EPI_WARPS);
    tc::fence_barrier_init();
  }
  if (warp == WARP_TMA) {
    if (lane == 0) {
      tc::tma_prefetch_desc(&tmH);
      tc::tma_prefetch_desc(&tmWg);
      tc::tma_prefetch_desc(&tmWc);
      tc::tma_prefetch_desc(&tmF);
    }
    __syncwarp();
    tc::tmem_alloc(tmem_slot, 512);
  }
  for (int i = tid; i < 3 * HGAP / 4; i += NTHREADS)      // the three gaps around the h chunks: zero, for good
    reinterpret_cast<uint32_t*>(smem_al + (sHg - base) + (i / (HGAP / 4)) * HCHUNK_STRIDE)[i % (HGAP / 4)] = 0u;
  for (int i = tid; i < NB; i += NTHREADS) {
    const float pl = i < 2 * HP ? -kLog2e * p.bg_pad[i] : ((i - 2 * HP) < p.h ? 2.0f * kLog2e * p.bc[i - 2 * HP] : 0.f);
    s_bias[i] = pl;
#pragma unroll
    for (int q = 0; q < 3; ++q) s_bias[(q + 1) * NB + i] = p.fold_bias ? p.fold_bias[q * NB + i] : pl;
  }
  tc::fence_proxy_async();                                // the zeroed gaps are read by the tensor core
  tc::fence_before_sync();
  __syncthreads();
  tc::fence_after_sync();
  const uint32_t tmem_base = *reinterpret_cast<volatile uint32_t*>(smem_al + (tmem_slot - base));

  if (warp >= WARP_AGG0) {
    // =============================== aggregation ===============================
    const int aw = warp - WARP_AGG0;
    const int R = BM / N_AGG_WARPS + (aw < BM % N_AGG_WARPS ? 1 : 0);          // rows of this warp (<= 31)
    const int row0 = aw * (BM / N_AGG_WARPS) + min(aw, BM % N_AGG_WARPS);      // first of them
    auto load_ptr = [&](int tile) -> int {                    // tptr[r0 .. r0+R] -> lanes 0..R (clamped at N)
      int64_t rr = (int64_t)tile * BM + row0 + lane;
      if (rr > p.N) rr = p.N;
      return lane <= R ? __ldg(p.tptr + rr) : 0x7fffffff;
    };
    int it = 0;
    int my_ptr = blockIdx.x < p.n_tiles ? load_ptr(blockIdx.x) : 0;
    for (int tile = blockIdx.x; tile < p.n_tiles; tile += gridDim.x, ++it) {
      const int xb = it & 1;
      uint8_t* xtile = smem_al + (sX - base) + xb * TILE16_BYTES;
      const int P0 = __shfl_sync(0xffffffffu, my_ptr, 0);
      const int P1 = __shfl_sync(0xffffffffu, my_ptr, R);
      int my_idx = (P0 + lane < P1) ? __ldg(p.tmsg + P0 + lane) : 0;
      int nidx = (P0 + 32 + lane < P1) ? __ldg(p.tmsg + P0 + 32 + lane) : 0;
      const int next_tile = tile + gridDim.x;
      const int nxt_ptr = next_tile < p.n_tiles ? load_ptr(next_tile) : 0;
      auto load32 = [&](uint2 (&v)[32], int pb) {             // messages pb .. pb+31 (re-reads past P1)
#pragma unroll
        for (int j = 0; j < 32; ++j) {
          const int idx = __shfl_sync(0xffffffffu, my_idx, (pb + j < P1) ? j : 0);
          v[j] = __ldg(reinterpret_cast<const uint2*>(p.msg + (size_t)idx * HP) + lane);
        }
      };
      uint2 cur[32];
      if (P0 < P1) load32(cur, P0);
      if (it >= 2) tc::mbar_wait_relaxed(bar(BAR_X_EMPTY + xb), ((it >> 1) - 1) & 1);
      {   // rows without any message are written as zeros up front
        const int nxt_p = __shfl_down_sync(0xffffffffu, my_ptr, 1);
        unsigned empty = __ballot_sync(0xffffffffu, lane < R && nxt_p == my_ptr);
        while (empty) {
          const int r = __ffs(empty) - 1;
          empty &= empty - 1;
          emit_row(xtile, row0 + r, lane, make_float4(0.f, 0.f, 0.f, 0.f));
        }
      }
      float4 acc = make_float4(0.f, 0.f, 0.f, 0.f);
#pragma unroll 1
      for (int pb = P0; pb < P1; pb += 32) {
        const int pos_l = pb + lane;
        int rowid = 0;
#pragma unroll
        for (int k = 1; k <= BM / N_AGG_WARPS + 1; ++k)       // lanes above R hold INT_MAX: never counted
          rowid += (__shfl_sync(0xffffffffu, my_ptr, k) <= pos_l) ? 1 : 0;
        const int rend = __shfl_sync(0xffffffffu, my_ptr, min(rowid + 1, R));
        const unsigned lastmask = __ballot_sync(0xffffffffu, pos_l < P1 && pos_l + 1 == rend);
        const unsigned valid = (P1 - pb >= 32) ? 0xffffffffu : ((1u << (P1 - pb)) - 1u);
#pragma unroll
        for (int j = 0; j < 32; ++j) {
          if (valid & (1u << j)) {
            const float2 lo = __half22float2(*reinterpret_cast<const __half2*>(&cur[j].x));
            const float2 hi = __half22float2(*reinterpret_cast<const __half2*>(&cur[j].y));
            acc.x += lo.x; acc.y += lo.y; acc.z += hi.x; acc.w += hi.y;
            if (lastmask & (1u << j)) {
              emit_row(xtile, row0 + __shfl_sync(0xffffffffu, rowid, j), lane, acc);
              acc = make_float4(0.f, 0.f, 0.f, 0.f);
            }
          }
        }
        if (pb + 32 < P1) {
          my_idx = nidx;
          nidx = (pb + 64 + lane < P1) ? __ldg(p.tmsg + pb + 64 + lane) : 0;
          load32(cur, pb + 32);
        }
      }
      tc::fence_proxy_async();
      __syncwarp();
      if (lane == 0) tc::mbar_arrive(bar(BAR_X_FULL + xb));
      my_ptr = nxt_ptr;
    }
  } else if (warp == WARP_TMA) {
    // =============================== TMA producer ===============================
    if (lane == 0) {
      int fill = 0;
      auto w_fill = [&](const CUtensorMap* tm, int kcol, int nrow) {       // one [128 n x 64 k] FP16 chunk
        const int s = fill % WSTAGES;
        if (fill >= WSTAGES) tc::mbar_wait_relaxed(bar(BAR_W_EMPTY + s), ((fill / WSTAGES) - 1) & 1);
        tc::mbar_arrive_expect_tx(bar(BAR_W_FULL + s), CHUNK_BYTES);
        tc::tma_load_2d(sW + s * CHUNK_BYTES, tm, kcol, nrow, bar(BAR_W_FULL + s));
        ++fill;
      };
      auto w_pair = [&](const CUtensorMap* tm, int kcol, int nrow) {       // both K chunks of a [128 x 128] block
        w_fill(tm, kcol, nrow);
        w_fill(tm, kcol + BKH, nrow);
      };
      int it = 0;
      for (int tile = blockIdx.x; tile < p.n_tiles; tile += gridDim.x, ++it) {   // order = the MMA issuer's
        const bool reg = p.regular && __ldg(p.regular + tile) != 0;
        w_pair(&tmWg, 0, 0);                                                // x      -> r
        if (it > 0) tc::mbar_wait_relaxed(bar(BAR_H_EMPTY), (it - 1) & 1); // previous h' stored
        tc::mbar_arrive_expect_tx(bar(BAR_H_FULL), TILE16_BYTES);
        tc::tma_load_2d(sHc(0), &tmH, 0, tile * BM, bar(BAR_H_FULL));
        tc::tma_load_2d(sHc(1), &tmH, BKH, tile * BM, bar(BAR_H_FULL));
        w_pair(&tmWg, HP, 0);                                               // h      -> r
        if (reg) { w_pair(&tmF, 0, 0); w_pair(&tmF, 0, NB); }               // h[t-1], h[t+1] -> r
        w_pair(&tmWg, 0, HP);                                               // x      -> u
        w_pair(&tmWc, 0, 0);                                                // x      -> cand
        w_pair(&tmWg, HP, HP);                                              // h      -> u
        if (reg) { w_pair(&tmF, 0, HP); w_pair(&tmF, 0, NB + HP); }         // h[t-1], h[t+1] -> u
        if (reg) { w_pair(&tmF, 0, 2 * HP); w_pair(&tmF, 0, NB + 2 * HP); } // h[t-1], h[t+1] -> cand
        w_pair(&tmWc, HP, 0);                                               // r*h    -> cand
      }
    }
    __syncwarp();
  } else if (warp == WARP_MMA) {
    // =============================== MMA issuer ===============================
    if (lane == 0) {
      const uint32_t idesc = tc::umma_idesc(0 /*F16*/, BM, 128);
      const uint32_t d_r = tmem_base, d_u = tmem_base + 128, d_cand = tmem_base + 256;
      int use = 0;
      auto chunk_mma = [&](uint32_t d, uint32_t a, bool first) {   // D[128x128] (+)= A chunk [128x64] @ W slot
        const int s = use % WSTAGES;
        tc::mbar_wait(bar(BAR_W_FULL + s), (use / WSTAGES) & 1);
        tc::fence_after_sync();
        const uint32_t w = sW + s * CHUNK_BYTES;
#pragma unroll
        for (int k = 0; k < 4; ++k)
          tc::umma_f16(d, tc::umma_desc_sw128(a + k * 32), tc::umma_desc_sw128(w + k * 32), idesc,
                       (first && k == 0) ? 0u : 1u);
        tc::umma_commit(bar(BAR_W_EMPTY + s));
        ++use;
      };
      // both K chunks of an A tile whose chunks are `stride` bytes apart; shift = -128 / +128: rows t-1 / t+1
      auto tile_mma = [&](uint32_t d, uint32_t a0, uint32_t stride, bool first) {
        chunk_mma(d, a0, first);
        chunk_mma(d, a0 + stride, false);
      };
      int it = 0;
      for (int tile = blockIdx.x; tile < p.n_tiles; tile += gridDim.x, ++it) {
        const bool reg = p.regular && __ldg(p.regular + tile) != 0;
        const int xb = it & 1;
        const uint32_t x0 = sX + xb * TILE16_BYTES, h0 = sHc(0);
        if (it > 0) tc::mbar_wait(bar(BAR_TMEM_EMPTY), (it - 1) & 1);
        tc::mbar_wait(bar(BAR_X_FULL + xb), (it >> 1) & 1);
        tc::fence_after_sync();
        tile_mma(d_r, x0, CHUNK_BYTES, true);                               // x -> r
        tc::mbar_wait(bar(BAR_H_FULL), it & 1);
        tc::fence_after_sync();
        tile_mma(d_r, h0, HCHUNK_STRIDE, false);                            // h -> r
        if (reg) {
          tile_mma(d_r, h0 - 128, HCHUNK_STRIDE, false);                    // h[t-1] -> r
          tile_mma(d_r, h0 + 128, HCHUNK_STRIDE, false);                    // h[t+1] -> r
        }
        tc::umma_commit(bar(BAR_R_DONE));
        tile_mma(d_u, x0, CHUNK_BYTES, true);                               // x -> u
        tile_mma(d_cand, x0, CHUNK_BYTES, true);                            // x -> cand
        tc::umma_commit(bar(BAR_X_READ));                                   // the x buffer may take r*h
        tile_mma(d_u, h0, HCHUNK_STRIDE, false);                            // h -> u
        if (reg) {
          tile_mma(d_u, h0 - 128, HCHUNK_STRIDE, false);
          tile_mma(d_u, h0 + 128, HCHUNK_STRIDE, false);
        }
        tc::umma_commit(bar(BAR_UC_DONE));                                  // u complete: the update-gate pass may start
        if (reg) {
          tile_mma(d_cand, h0 - 128, HCHUNK_STRIDE, false);
          tile_mma(d_cand, h0 + 128, HCHUNK_STRIDE, false);
        }
        tc::mbar_wait(bar(BAR_RH_FULL), it & 1);                            // f16(r*h) is in the x buffer
        tc::fence_after_sync();
        tile_mma(d_cand, x0, CHUNK_BYTES, false);                           // r*h -> cand
        tc::umma_commit(bar(BAR_X_EMPTY + xb));                             // this x buffer may be refilled
        tc::umma_commit(bar(BAR_CAND_DONE));
      }
    }
    __syncwarp();
  }
  else if (warp < WARP_EPI0 + N_EPI_WARPS) {
    // =============================== epilogue (warps 0-7) ===============================
    const int ew = warp & 3;                                  // TMEM lane quarter this warp may access
    const int chalf = (warp - WARP_EPI0) >> 2;                // which 64 columns = which K chunk of the tiles
    const int r = ew * 32 + lane;                             // row within the tile = TMEM lane
    const uint32_t t_lane = tmem_base + ((uint32_t)(ew * 32) << 16);
    uint8_t* hrow = smem_al + (sHc(chalf) - base);            // + sw_off(r, slot): this thread's h row, its 64 columns
    const bool leader = tid == WARP_EPI0 * 32;
    int it = 0;
    for (int tile = blockIdx.x; tile < p.n_tiles; tile += gridDim.x, ++it) {
      const bool reg = p.regular && __ldg(p.regular + tile) != 0;
      // biases of this row: plain, or (word-order tile) with the folded edge biases of the word-order messages it gets
      const float* s_bg = s_bias + (!reg ? 0 : (r == 0 ? 2 * NB : (r == BM - 1 ? 3 * NB : NB)));
      const float* s_bc = s_bg + 2 * HP;
      uint8_t* xrow = smem_al + (sX - base) + (it & 1) * TILE16_BYTES + chalf * CHUNK_BYTES;
      // ---- EPI1a: reset gate; f16(r*h) goes into the x buffer (its readers have retired); h stays where it is
      tc::mbar_wait(bar(BAR_R_DONE), it & 1);
      tc::fence_after_sync();
#pragma unroll 1
      for (int c16 = 0; c16 < 4; ++c16) {
        const int col0 = chalf * 64 + c16 * 16;
        uint32_t ar[16];
        tc::tmem_ld16(t_lane + col0, ar);
        float hh[16];
        {
          float t0[8], t1[8];
          unpack8(*reinterpret_cast<const uint4*>(hrow + sw_off(r, 2 * c16)), t0);
          unpack8(*reinterpret_cast<const uint4*>(hrow + sw_off(r, 2 * c16 + 1)), t1);
#pragma unroll
          for (int i = 0; i < 8; ++i) { hh[i] = t0[i]; hh[8 + i] = t1[i]; }
        }
        tc::tmem_ld_wait();
        float o0[8], o1[8];
#pragma unroll
        for (int i = 0; i < 16; ++i) {
          const float rg = tc::rcp_approx(1.0f + tc::ex2_approx(fmaf(__uint_as_float(ar[i]), -kLog2e, s_bg[col0 + i])));
          const float v = rg * hh[i];
          if (i < 8) o0[i] = v; else o1[i - 8] = v;
        }
        if (c16 == 0) tc::mbar_wait(bar(BAR_X_READ), it & 1);
        *reinterpret_cast<uint4*>(xrow + sw_off(r, 2 * c16)) = pack8(o0);
        *reinterpret_cast<uint4*>(xrow + sw_off(r, 2 * c16 + 1)) = pack8(o1);
      }
      tc::fence_proxy_async();
      __syncwarp();
      if (lane == 0) tc::mbar_arrive(bar(BAR_RH_FULL));
      // ---- EPI1b (overlaps the r*h MMAs): update gate, in place in TMEM
      tc::mbar_wait(bar(BAR_UC_DONE), it & 1);
      tc::fence_after_sync();
#pragma unroll 1
      for (int c16 = 0; c16 < 4; ++c16) {
        const int col0 = chalf * 64 + c16 * 16;
        uint32_t au[16];
        tc::tmem_ld16(t_lane + 128 + col0, au);
        tc::tmem_ld_wait();
#pragma unroll
        for (int i = 0; i < 16; ++i)
          au[i] = __float_as_uint(tc::rcp_approx(1.0f + tc::ex2_approx(fmaf(__uint_as_float(au[i]), -kLog2e, s_bg[HP + col0 + i]))));
        tc::tmem_st16(t_lane + 128 + col0, au);
      }
      tc::tmem_st_wait();
      // ---- EPI2: candidate and new state h' = c + u*(h - c): FP16 over h in the h tile (every MMA reading it has
      //      retired) for the TMA store of the state; the unrounded FP32 straight to h_out on the last step
      tc::mbar_wait(bar(BAR_CAND_DONE), it & 1);
      tc::fence_after_sync();
      const int64_t grow = (int64_t)tile * BM + r;
#pragma unroll 1
      for (int c16 = 0; c16 < 4; ++c16) {
        const int col0 = chalf * 64 + c16 * 16;
        uint32_t ac[16], uu[16];
        tc::tmem_ld16(t_lane + 256 + col0, ac);
        tc::tmem_ld16(t_lane + 128 + col0, uu);
        uint4* s0 = reinterpret_cast<uint4*>(hrow + sw_off(r, 2 * c16));
        uint4* s1 = reinterpret_cast<uint4*>(hrow + sw_off(r, 2 * c16 + 1));
        float hh[16];
        {
          float t0[8], t1[8];
          unpack8(*s0, t0); unpack8(*s1, t1);
#pragma unroll
          for (int i = 0; i < 8; ++i) { hh[i] = t0[i]; hh[8 + i] = t1[i]; }
        }
        tc::tmem_ld_wait();
        float o0[8], o1[8];
#pragma unroll
        for (int i = 0; i < 16; ++i) {
          const float c = fmaf(-2.0f, tc::rcp_approx(1.0f + tc::ex2_approx(
                                          fmaf(__uint_as_float(ac[i]), 2.0f * kLog2e, s_bc[col0 + i]))), 1.0f);
          const float v = fmaf(__uint_as_float(uu[i]), hh[i] - c, c);
          if (i < 8) o0[i] = v; else o1[i - 8] = v;
        }
        *s0 = pack8(o0);
        *s1 = pack8(o1);
        if (p.h_out && grow < p.N) {
          float* dst = p.h_out + grow * p.h + col0;
#pragma unroll
          for (int q = 0; q < 4; ++q) {
            if (col0 + q * 4 < p.h) {
              const float* src = q < 2 ? o0 + q * 4 : o1 + (q - 2) * 4;
              *reinterpret_cast<float4*>(dst + q * 4) = make_float4(src[0], src[1], src[2], src[3]);
            }
          }
        }
      }
      tc::fence_before_sync();                                // TMEM reads done
      tc::fence_proxy_async();                                // smem writes -> visible to the TMA unit
      __syncwarp();
      if (lane == 0) tc::mbar_arrive(bar(BAR_TMEM_EMPTY));
      tc::named_bar_sync(1, N_EPI_WARPS * 32);
      if (leader) {
        tc::tma_store_2d(&tmH, 0, tile * BM, sHc(0));         // rows >= N clipped by the map
        tc::tma_store_2d(&tmH, BKH, tile * BM, sHc(1));
        tc::tma_store_commit();
        tc::tma_store_wait_read();
        tc::mbar_arrive(bar(BAR_H_EMPTY));                    // h tile may be refilled
      }
    }
    if (leader) asm volatile("cp.async.bulk.wait_group 0;" ::: "memory");
  }

  tc::fence_before_sync();
  __syncthreads();
  if (warp == WARP_TMA) tc::tmem_dealloc(tmem_base, 512);
}

}  // namespace gru16

// =========================================================================================
bool ggnn_f16_supported(const ggnn_ctx* ctx) { return ctx->hp == HP && ctx->h % 4 == 0; }

static int realloc_h(ggnn_ctx* ctx, void*& p, size_t n_halves) {
  if (p && ctx->repack_in_place) return GGNN_OK;
  if (p) { cudaFree(p); p = nullptr; }
  GGNN_CUDA(ctx, cudaMalloc(&p, n_halves * 2));
  return GGNN_OK;
}

// folded word-order weights of one slot (needs the slot's edge weights and the GRU cell)
static int fold_slot(ggnn_ctx* ctx, int slot, cudaStream_t st) {
  ggnn_edge_weights& ew = ctx->edge[slot];
  ew.fold_ok = false;
  if (ctx->two_e < 2 || ctx->two_e % 2 || !ctx->Wg || !ctx->Wc || !ew.W) return GGNN_OK;
  GGNN_TRY(realloc_h(ctx, ew.fold_f16, (size_t)2 * 3 * HP * HP));
  if (!ew.fold_bias) GGNN_CUDA(ctx, cudaMalloc(&ew.fold_bias, 9 * HP * sizeof(float)));
  fold_kernel<<<2 * 3 * HP, HP, 0, st>>>(ew.W, ctx->Wg, ctx->Wc, reinterpret_cast<__half*>(ew.fold_f16), ctx->two_e, ctx->h);
  GGNN_LAUNCHED(ctx);
  fold_bias_kernel<<<3, HP, 0, st>>>(ew.B, ctx->Wg, ctx->Wc, ctx->bg, ctx->bc, ew.fold_bias, ctx->two_e, ctx->h);
  GGNN_LAUNCHED(ctx);
  GGNN_TRY(ggnn_make_tmap_2d_f16(ctx, &ew.tm_fold, ew.fold_f16, 2 * 3 * HP, HP, BKH, HP));
  ew.fold_ok = true;
  return GGNN_OK;
}

bool ggnn_f16_word_fold_ready(const ggnn_ctx* ctx, int slot) {
  static const bool off = getenv("GGNN_NO_WORD_FOLD") != nullptr;      // A-B switch
  return !off && ggnn_f16_supported(ctx) && ctx->edge[slot].fold_ok;
}

int ggnn_f16_pack_weights(ggnn_ctx* ctx, int slot, bool with_gru, cudaStream_t st) {
  if (!ggnn_f16_supported(ctx)) { ctx->edge[slot].fold_ok = false; return GGNN_OK; }
  const int h = ctx->h, two_e = ctx->two_e;
  ggnn_edge_weights& ew = ctx->edge[slot];
  ew.absmax_w = -1.f;
  GGNN_TRY(realloc_h(ctx, ew.Wt_f16, (size_t)two_e * HP * HP));
  pack_edge16_kernel<<<ctx->sm_count * 4, 256, 0, st>>>(ew.W, reinterpret_cast<__half*>(ew.Wt_f16), two_e, h);
  GGNN_LAUNCHED(ctx);
  GGNN_TRY(ggnn_make_tmap_2d_f16(ctx, &ew.tm_Wt16, ew.Wt_f16, (int64_t)two_e * HP, HP, BKH, HP));
  GGNN_TRY(ggnn_w16_pack_bias(ctx, slot, HP, st));
  if (with_gru) {
    GGNN_TRY(realloc_h(ctx, ctx->Wgt_f16, (size_t)4 * HP * HP));
    GGNN_TRY(realloc_h(ctx, ctx->Wct_f16, (size_t)2 * HP * HP));
    pack_gru16_kernel<<<ctx->sm_count * 2, 256, 0, st>>>(ctx->Wg, reinterpret_cast<__half*>(ctx->Wgt_f16), 2, h);
    GGNN_LAUNCHED(ctx);
    pack_gru16_kernel<<<ctx->sm_count * 2, 256, 0, st>>>(ctx->Wc, reinterpret_cast<__half*>(ctx->Wct_f16), 1, h);
    GGNN_LAUNCHED(ctx);
    GGNN_TRY(ggnn_make_tmap_2d_f16(ctx, &ctx->tm_Wg16, ctx->Wgt_f16, 2 * HP, 2 * HP, BKH, HP));
    GGNN_TRY(ggnn_make_tmap_2d_f16(ctx, &ctx->tm_Wc16, ctx->Wct_f16, HP, 2 * HP, BKH, HP));
    for (int s2 = 0; s2 < GGNN_MAX_SLOTS; ++s2)          // a new GRU cell re-folds every slot that is already set
      if (s2 != slot && ctx->edge[s2].set) GGNN_TRY(fold_slot(ctx, s2, st));
  }
  GGNN_TRY(fold_slot(ctx, slot, st));
  GGNN_TRY(ggnn_gru_pair_pack(ctx, slot, with_gru, st));
  return GGNN_OK;
}

int ggnn_f16_state_in(ggnn_ctx* ctx, const float* h, void* state16, int64_t N, unsigned* absmax_bits,
                      cudaStream_t st) {
  if (N == 0) return GGNN_OK;
  prof_scope ps(ctx, GGNN_STAGE_OTHER, st);
  state_in_kernel<<<ctx->sm_count * 8, 256, 0, st>>>(h, reinterpret_cast<__half*>(state16), N, ctx->h, absmax_bits);
  GGNN_LAUNCHED(ctx);
  return GGNN_OK;
}

// max over (e, n) of sum_k |W[e][k][n]|: the largest L1 norm of a column of any edge weight matrix.  |h @ W_e| <=
// colnorm * max|h| element-wise, which is what the FP16 range guard needs -- h * max|W| is the same bound with every
// |W| replaced by the largest one.  One thread per (e, n), coalesced over n; fixed order, no float atomics on the sum.
__global__ void colnorm_kernel(const float* __restrict__ W, int two_e, int h, unsigned* __restrict__ out) {
  const int64_t i = (int64_t)blockIdx.x * blockDim.x + threadIdx.x;
  float s = 0.f;
  if (i < (int64_t)two_e * h) {
    const int64_t e = i / h;
    const int n = (int)(i % h);
    const float* p = W + e * h * h + n;
    for (int k = 0; k < h; ++k) s += fabsf(p[(size_t)k * h]);
  }
#pragma unroll
  for (int o = 16; o > 0; o >>= 1) s = fmaxf(s, __shfl_xor_sync(0xffffffffu, s, o));
  if ((threadIdx.x & 31) == 0 && s > 0.f) atomicMax(out, __float_as_uint(s));     // NaN fails the `> 0`; Inf propagates
}

int ggnn_colnorm_max(ggnn_ctx* ctx, const float* W, int two_e, int h, unsigned* out_bits, cudaStream_t st) {
  const int64_t n = (int64_t)two_e * h;
  if (n <= 0) return GGNN_OK;
  colnorm_kernel<<<(unsigned)ceil_div64(n, 128), 128, 0, st>>>(W, two_e, h, out_bits);
  GGNN_LAUNCHED(ctx);
  return GGNN_OK;
}

int ggnn_absmax(ggnn_ctx* ctx, const float* x, int64_t n, unsigned* out_bits, cudaStream_t st) {
  if (n <= 0) return GGNN_OK;
  absmax_kernel<<<ctx->sm_count * 4, 256, 0, st>>>(x, n, out_bits);
  GGNN_LAUNCHED(ctx);
  return GGNN_OK;
}

int ggnn_step_f16(ggnn_ctx* ctx, int slot, const ggnn_plan& plan, void* msg16, void* state16,
                  float* h_out, bool last_step, cudaStream_t st) {
  if (plan.N == 0) return GGNN_OK;
  if (!(ctx->attr_mask & (1u << GGNN_ATTR_F16))) {
    GGNN_CUDA(ctx, cudaFuncSetAttribute(msg16::msg_f16_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, msg16::SMEM_BYTES));
    GGNN_CUDA(ctx, cudaFuncSetAttribute(gru16::gru_f16_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, gru16::SMEM_BYTES));
    ctx->attr_mask |= 1u << GGNN_ATTR_F16;
  }
  const ggnn_edge_weights& ew = ctx->edge[slot];
  static const bool msg_w16 = getenv("GGNN_MSG_W16") != nullptr;      // A-B switch
  if (msg_w16) {
    GGNN_TRY(ggnn_w16_msg(ctx, slot, plan, msg16, state16, HP, st));
  } else if (plan.max_tiles > 0 && plan.nnz > 0) {
    CUtensorMap tmOut;
    GGNN_TRY(ggnn_make_tmap_2d_f16(ctx, &tmOut, msg16, plan.max_tiles * (int64_t)BM, HP, BKH, BM));
    msg16::params p{};
    p.state16 = reinterpret_cast<const __half*>(state16); p.msg_src = plan.msg_src; p.tile_type = plan.tile_type;
    p.n_tiles = plan.n_tiles; p.bias = ew.B; p.h = ctx->h;
    const int grid = (int)(plan.max_tiles < ctx->sm_count ? plan.max_tiles : ctx->sm_count);
    prof_scope ps(ctx, GGNN_STAGE_MSG, st);
    msg16::msg_f16_kernel<<<grid, msg16::NTHREADS, msg16::SMEM_BYTES, st>>>(p, ew.tm_Wt16, tmOut);
    GGNN_LAUNCHED(ctx);
  }
  if (ggnn_gru_pair_enabled(ctx, slot)) {                // CTA pairs, resident weights (gru_pair.cu)
    prof_scope ps(ctx, GGNN_STAGE_GRU_GATES, st);
    return ggnn_gru_pair_launch(ctx, slot, plan, msg16, state16, last_step ? h_out : nullptr, st);
  }
  CUtensorMap tmH;
  GGNN_TRY(ggnn_make_tmap_2d_f16(ctx, &tmH, state16, plan.N, HP, BKH, BM));
  gru16::params g{};
  g.msg = reinterpret_cast<const __half*>(msg16); g.tptr = plan.tptr; g.tmsg = plan.tmsg;
  g.h_out = last_step ? h_out : nullptr; g.bg_pad = ctx->bg_pad; g.bc = ctx->bc;
  g.regular = plan.tile_regular;                       // non-null only when the plan left the word-order rows out
  g.fold_bias = plan.tile_regular ? ew.fold_bias : nullptr;
  if (plan.tile_regular && !ew.fold_ok) GGNN_FAIL(ctx, GGNN_E_STATE, "internal: word-order plan without folded weights");
  g.N = plan.N; g.h = ctx->h; g.n_tiles = (int)ceil_div64(plan.N, BM);
  const int grid = g.n_tiles < ctx->sm_count ? g.n_tiles : ctx->sm_count;
  prof_scope ps(ctx, GGNN_STAGE_GRU_GATES, st);
  gru16::gru_f16_kernel<<<grid, gru16::NTHREADS, gru16::SMEM_BYTES, st>>>(g, tmH, ctx->tm_Wg16, ctx->tm_Wc16,
                                                                          ew.fold_ok ? ew.tm_fold : ctx->tm_Wc16);
  GGNN_LAUNCHED(ctx);
  return GGNN_OK;
}
