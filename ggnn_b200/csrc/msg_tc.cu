// msg_tc.cu -- TF32 mode, hidden size padded to 128: the message transform
//     M[r, :] = tf32(h[msg_src[r], :]) @ tf32(W[type(tile)]) + B[type]   (chem_tensorflow_dense.py:397-412)
// as ONE persistent, warp-specialised kernel per timestep (an MoE-style grouped GEMM: the 2E
// edge types are the experts, message rows arrive sorted by type in 128-row tiles).
//
// Each CTA owns a CONTIGUOUS range of tiles, so W_e^T (64 KB) stays resident in shared memory
// until the edge type changes (once or twice per CTA at C3) instead of being re-streamed per
// tile.  Roles (288 threads, 1 CTA / SM, 225 KB smem, 256 TMEM columns):
//   warps 0-3  gather producers: 128 state rows x 512 B per tile by 16-byte cp.async straight
//              into the 128B-swizzled K-major operand tile; 2 x 64 KB stages
//   warp  4    one lane issues 16 tcgen05.mma (M128 N128 K8, kind::tf32) per tile into one of two
//              TMEM accumulators and commits to the stage-empty / accumulator-full barriers
//   warps 5-8  epilogue: tcgen05.ld, + bias, -> FP16, swizzled smem staging, TMA store of the
//              [128 x 128] FP16 tile (coalesced 128-byte lines instead of per-row 8-byte stores)
// The epilogue of tile i overlaps the MMAs of tile i+1 and the gather of tiles i+2, i+3.
#include "tf32.cuh"
#include "tc.cuh"
#include <cuda_fp16.h>

namespace {

constexpr int HP = 128;
constexpr int BM = 128;
constexpr int BKF = 32;
constexpr int CHUNK_BYTES = BM * BKF * 4;       // 16 KB
constexpr int A_BYTES = 4 * CHUNK_BYTES;        // 64 KB per stage
constexpr int W_BYTES = 4 * CHUNK_BYTES;        // 64 KB
constexpr int O_BYTES = BM * HP * 2;            // 32 KB FP16 staging (two 64-column boxes)
constexpr int ASTAGES = 2;
constexpr int NPROD = 128;
constexpr int WARP_MMA = 4;
constexpr int WARP_EPI0 = 5;
constexpr int NTHREADS = 9 * 32;
constexpr int SMEM_BYTES = ASTAGES * A_BYTES + W_BYTES + O_BYTES + 1024 /*align*/ + 512 /*barriers + bias*/ + 512;

enum {
  BAR_A_FULL = 0,                       // [ASTAGES] 128 cp.async arrivals
  BAR_A_EMPTY = BAR_A_FULL + ASTAGES,   // [ASTAGES] tcgen05.commit
  BAR_W_FULL = BAR_A_EMPTY + ASTAGES,   // TMA bytes
  BAR_MMA_DONE,                         // tcgen05.commit per tile (W may be replaced after it)
  BAR_ACC_FULL,                         // [2]
  BAR_ACC_EMPTY = BAR_ACC_FULL + 2,     // [2] 4 epilogue warps
  BAR_COUNT = BAR_ACC_EMPTY + 2
};

struct mparams {
  const float* h_state;       // [N, h] TF32-rounded node states
  const int32_t* msg_src;     // [tiles*128] source row per message row, -1 = padding
  const int32_t* tile_type;   // [tiles]
  const int32_t* n_tiles;     // device-side tile count
  const float* bias;          // [two_e, h] or null
  int h;
};

__global__ void __launch_bounds__(NTHREADS, 1)
msg_tc_kernel(const mparams p, const __grid_constant__ CUtensorMap tmW,
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
  float* s_bias = reinterpret_cast<float*>(smem_al + (bars + 128 - base));   // [128]

  // this CTA's contiguous tile range
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
    tc::mbar_init(bar(BAR_MMA_DONE), 1);
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
    // =============================== gather producers ===============================
    const int c16 = tid & 7;                 // 16-byte slot within a 128-byte row chunk
    const int rbase = tid >> 3;              // rows rbase + 16*j
    uint32_t a_off[8];
#pragma unroll
    for (int j = 0; j < 8; ++j) {
      const int r = rbase + 16 * j;
      a_off[j] = (r >> 3) * 1024 + (r & 7) * 128 + ((c16 ^ (r & 7)) << 4);
    }
    int src[8], nsrc[8];
    auto load_idx = [&](int tile, int (&dst)[8]) {
#pragma unroll
      for (int j = 0; j < 8; ++j) dst[j] = __ldg(p.msg_src + (size_t)tile * BM + rbase + 16 * j);
    };
    if (my_tiles > 0) load_idx(t_begin, src);
    int cur_type = -1;
    for (int i = 0; i < my_tiles; ++i) {
      const int tile = t_begin + i;
      const int s = i % ASTAGES;
      if (i + 1 < my_tiles) load_idx(tile + 1, nsrc);
      if (i >= ASTAGES) tc::mbar_wait(bar(BAR_A_EMPTY + s), ((i / ASTAGES) - 1) & 1);
      if (tid == 0) {                         // weights: reload when the edge type changes
        const int type = __ldg(p.tile_type + tile);
        if (type != cur_type) {
          // MMAs reading the old W must be done.  Tile i-2 is known complete (stage wait above), so
          // this thread is at most one phase behind MMA_DONE and the parity wait is unambiguous.
          if (i > 0) tc::mbar_wait(bar(BAR_MMA_DONE), (i - 1) & 1);
          tc::mbar_arrive_expect_tx(bar(BAR_W_FULL), W_BYTES);
#pragma unroll
          for (int kc = 0; kc < 4; ++kc)
            tc::tma_load_2d(sW + kc * CHUNK_BYTES, &tmW, kc * BKF, type * HP, bar(BAR_W_FULL));
          cur_type = type;
        }
      }
      const uint32_t a_st = sA + s * A_BYTES;
#pragma unroll
      for (int kc = 0; kc < 4; ++kc) {
        const int kloc = kc * BKF + c16 * 4;
        const bool kin = kloc < p.h;
#pragma unroll
        for (int j = 0; j < 8; ++j) {
          const bool ok = kin && src[j] >= 0;
          tc::cp_async16(a_st + kc * CHUNK_BYTES + a_off[j],
                         ok ? (const void*)(p.h_state + (size_t)src[j] * p.h + kloc) : (const void*)p.h_state,
                         ok ? 16u : 0u);
        }
      }
      tc::cp_async_arrive(bar(BAR_A_FULL + s));
#pragma unroll
      for (int j = 0; j < 8; ++j) src[j] = nsrc[j];
    }
  } else if (warp == WARP_MMA) {
    // =============================== MMA issuer ===============================
    if (lane == 0) {
      const uint32_t idesc = tc::umma_idesc(2, BM, HP);
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
        for (int kc = 0; kc < 4; ++kc)
#pragma unroll
          for (int k = 0; k < 4; ++k)
            tc::umma_tf32(d, tc::umma_desc_sw128(a0 + kc * CHUNK_BYTES + k * 32),
                          tc::umma_desc_sw128(sW + kc * CHUNK_BYTES + k * 32), idesc, (kc | k) != 0);
        tc::umma_commit(bar(BAR_A_EMPTY + s));
        tc::umma_commit(bar(BAR_ACC_FULL + buf));
        tc::umma_commit(bar(BAR_MMA_DONE));
      }
    }
    __syncwarp();
  } else {
    // =============================== epilogue (warps 5-8) ===============================
    const int q = warp & 3;                       // TMEM lane quarter this warp may access
    const int r = q * 32 + lane;                  // row within the tile
    const int etid = tid - WARP_EPI0 * 32;        // 0..127
    const uint32_t t_lane = tmem_base + ((uint32_t)(q * 32) << 16);
    uint8_t* orow = smem_al + (sO - base) + r * 128;   // + box * 16 KB + (chunk ^ (r & 7)) * 16
    int cur_type = -1;
    for (int i = 0; i < my_tiles; ++i) {
      const int tile = t_begin + i, buf = i & 1;
      const int type = __ldg(p.tile_type + tile);
      if (type != cur_type) {                     // uniform across the 4 warps
        tc::named_bar_sync(2, 128);               // everyone is done reading the old bias
        s_bias[etid] = (p.bias && etid < p.h) ? __ldg(p.bias + (size_t)type * p.h + etid) : 0.f;
        cur_type = type;
      }
      tc::mbar_wait(bar(BAR_ACC_FULL + buf), (i >> 1) & 1);
      tc::fence_after_sync();
      if (etid == 0) tc::tma_store_wait_read();   // previous tile's store has left the staging buffer
      tc::named_bar_sync(1, 128);                 // (also publishes s_bias)
#pragma unroll
      for (int cc = 0; cc < 4; ++cc) {            // 32 columns at a time
        uint32_t acc[32];
        tc::tmem_ld32(t_lane + buf * HP + cc * 32, acc);
        tc::tmem_ld_wait();
        uint8_t* obox = orow + (cc >> 1) * (BM * 128);
#pragma unroll
        for (int g = 0; g < 4; ++g) {             // 8 columns -> one 16-byte chunk
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
          const int chunk = (cc & 1) * 4 + g;     // 16-byte chunk within the 128-byte box row
          *reinterpret_cast<uint4*>(obox + ((chunk ^ (r & 7)) << 4)) = pk;
        }
      }
      tc::fence_before_sync();                    // TMEM reads done -> accumulator may be overwritten
      __syncwarp();
      if (lane == 0) tc::mbar_arrive(bar(BAR_ACC_EMPTY + buf));
      tc::fence_proxy_async();                    // staging writes -> visible to the TMA unit
      tc::named_bar_sync(1, 128);
      if (etid == 0) {
        tc::tma_store_2d(&tmOut, 0, tile * BM, sO);
        tc::tma_store_2d(&tmOut, 64, tile * BM, sO + BM * 128);
        tc::tma_store_commit();
      }
    }
    if (etid == 0) asm volatile("cp.async.bulk.wait_group 0;" ::: "memory");
  }

  tc::fence_before_sync();
  __syncthreads();
  if (warp == 0) tc::tmem_dealloc(tmem_base, 256);
}

}  // namespace

bool ggnn_msg_tc_supported(const ggnn_ctx* ctx) { return ctx->hp == HP && ctx->h % 4 == 0; }

// msg_half: [max_tiles*128, 128] FP16 (row pitch 128 halves whatever h is)
int ggnn_msg_tc_step(ggnn_ctx* ctx, int slot, const ggnn_plan& plan, const float* h_state,
                     void* msg_half, cudaStream_t st) {
  if (plan.max_tiles == 0 || plan.nnz == 0) return GGNN_OK;
  GGNN_ENSURE_SMEM(ctx, GGNN_ATTR_MSG_TC, msg_tc_kernel, SMEM_BYTES);
  const ggnn_edge_weights& ew = ctx->edge[slot];
  CUtensorMap tmOut;
  GGNN_TRY(ggnn_make_tmap_2d_f16(ctx, &tmOut, msg_half, plan.max_tiles * (int64_t)BM, HP, 64, BM));
  mparams p{};
  p.h_state = h_state; p.msg_src = plan.msg_src; p.tile_type = plan.tile_type; p.n_tiles = plan.n_tiles;
  p.bias = ew.B; p.h = ctx->h;
  const int grid = (int)(plan.max_tiles < ctx->sm_count ? plan.max_tiles : ctx->sm_count);
  prof_scope ps(ctx, GGNN_STAGE_MSG, st);
  msg_tc_kernel<<<grid, NTHREADS, SMEM_BYTES, st>>>(p, ew.tm_Wt, tmOut);
  GGNN_LAUNCHED(ctx);
  return GGNN_OK;
}
