// gru_fused.cu -- TF32 mode, hidden size <= 128: aggregation + the whole GRU cell in ONE
// persistent kernel per timestep.  Per 128-node tile (one graph at v = 128):
//
//   AGG    x[t,:]  = sum of the incoming message rows (segmented gather-sum, target-CSR order),
//                    rounded to TF32 and written straight into the MMA operand tile in smem
//   MMA1   gates   = [x, h] @ Wg        (TMEM cols   0..255: r | u)
//          cand_x  =  x     @ Wc[:h]    (TMEM cols 256..383)
//   EPI1   r = sigmoid(gates_r + b); rh = tf32(r * h) written over the h tile in smem, h parked
//          in TMEM; then, while MMA2 runs, u = sigmoid(gates_u + b) in place in TMEM
//   MMA2   cand   += rh @ Wc[h:]
//   EPI2   c = tanh(cand + b); h' = c + u*(h - c) -> smem -> TMA store
//
// acts, r*h and u never touch HBM (the three-kernel path moves 5 x 64 KB more per tile).
// Warp roles (640 threads, 1 CTA / SM, ~224 KB smem, 512 TMEM columns):
//   warps 0-7    epilogue (TMEM lanes 32*(warp%4)..), thread = node row, 64 columns each
//   warp  8      tcgen05.mma issuer
//   warp  9      TMA: h tile (4 boxes) and the weight ring (2 x 48 KB stages, 12 fills per tile)
//   warps 12-19  aggregation producers (generic-proxy smem stores + fence.proxy.async)
// AGG of tile i+1 overlaps MMA/EPI of tile i (x tile is released by a tcgen05.commit as soon
// as MMA1 has consumed it).
#include "tf32.cuh"
#include "tc.cuh"
#include <cuda_fp16.h>

namespace {

constexpr float kLog2e = 1.4426950408889634f;
constexpr int HP = 128;                      // padded hidden size this kernel is built for
constexpr int BM = 128;
constexpr int BKF = 32;
constexpr int CHUNK_BYTES = BM * BKF * 4;    // 16 KB: one [128 x 32] K-major operand chunk
constexpr int TILE_BYTES = 4 * CHUNK_BYTES;  // 64 KB: x tile / h tile
constexpr int WSTAGES = 2;
constexpr int WSTAGE_BYTES = 48 * 1024;      // [256 gate rows | 128 cand rows] x 32 k
// Warp ids double as scheduler priority (higher id wins the issue slot), so the roles that do
// the most instruction work get the highest ids and the mostly-waiting ones the lowest.
constexpr int N_AGG_WARPS = 8;
constexpr int N_EPI_WARPS = 8;   // two per TMEM lane quarter, each takes half of the columns
constexpr int WARP_EPI0 = 0;                             // 0..7 (4-aligned: warp % 4 = TMEM lane quarter)
constexpr int WARP_MMA = 8;
constexpr int WARP_TMA = 9;                              // warps 10, 11 only take part in the block syncs
constexpr int WARP_AGG0 = 12;                            // 12..19
constexpr int NTHREADS = (WARP_AGG0 + N_AGG_WARPS) * 32; // 640
constexpr int SMEM_BYTES = 2 * TILE_BYTES + WSTAGES * WSTAGE_BYTES + 1024 /*align*/ + 128 /*barriers*/ +
                           3 * HP * 4 /*biases*/;   // 232 064 <= 232 448 (227 KB opt-in limit)

struct gparams {
  const __half* msg;        // [*, 128] message rows (pitch HP halves whatever h is; FP16: same 11-bit significand as a TF32 operand)
  const int32_t* tptr;      // [N+1]
  const int32_t* tmsg;      // [nnz]
  float* h_state;           // [N, h] in/out (TF32-rounded between steps)
  const float* bg_pad;      // [2*HP]
  const float* bc;          // [h]
  int64_t N;
  int h;
  int n_tiles;
  int round_out;
  unsigned long long* dbg;  // optional [64 events x 16 iterations] globaltimer trace of CTA 0
};

// (compiled out unless built with GGNN_NVCC_EXTRA=-DGGNN_GRU_FUSED_TRACE: hooks that are never taken still cost the
//  CTA-pair kernel of gru_pair.cu 4-9 % -- measured there)
__device__ __forceinline__ void trace(const gparams& p, int it, int ev) {
#ifdef GGNN_GRU_FUSED_TRACE
  if (p.dbg && blockIdx.x == 0 && it < 16) {
    unsigned long long t;
    asm volatile("mov.u64 %0, %%globaltimer;" : "=l"(t));
    p.dbg[it * 64 + ev] = t;
  }
#else
  (void)p; (void)it; (void)ev;
#endif
}

// barrier slots (8 bytes each) after the operand buffers
enum {
  BAR_W_FULL = 0,                    // [WSTAGES]
  BAR_W_EMPTY = BAR_W_FULL + WSTAGES,
  BAR_X_FULL = BAR_W_EMPTY + WSTAGES,
  BAR_X_EMPTY,
  BAR_H_FULL,
  BAR_H_EMPTY,
  BAR_GATES_DONE,
  BAR_RH_FULL,
  BAR_CAND_DONE,
  BAR_TMEM_EMPTY,
  BAR_COUNT
};

// one aggregated row (this lane's 4 columns) -> TF32 -> its slot in the swizzled x tile.
// Deliberately not inlined: it sits inside a 16x unrolled loop and the kernel must stay small
// enough for the instruction cache (18 warps run five different code paths).
__device__ __noinline__ void emit_row(uint8_t* xtile, int r, int lane, float4 acc, bool col_ok) {
  const int kc = lane >> 3, c16 = lane & 7;
  const uint32_t off = kc * CHUNK_BYTES + (r >> 3) * 1024 + (r & 7) * 128 + ((c16 ^ (r & 7)) << 4);
  float4 o = col_ok ? acc : make_float4(0.f, 0.f, 0.f, 0.f);
  o.x = tc::to_tf32_fast(o.x); o.y = tc::to_tf32_fast(o.y); o.z = tc::to_tf32_fast(o.z); o.w = tc::to_tf32_fast(o.w);
  *reinterpret_cast<float4*>(xtile + off) = o;
}

__global__ void __launch_bounds__(NTHREADS, 1)
gru_fused_kernel(const gparams p, const __grid_constant__ CUtensorMap tmH,
                 const __grid_constant__ CUtensorMap tmWg, const __grid_constant__ CUtensorMap tmWc) {
  extern __shared__ uint8_t smem_raw[];
  const int tid = threadIdx.x, warp = tid >> 5, lane = tid & 31;

  const uint32_t base = (tc::smem_u32(smem_raw) + 1023u) & ~1023u;
  uint8_t* smem_al = smem_raw + (base - tc::smem_u32(smem_raw));
  const uint32_t sX = base;                         // x tile: 4 chunks
  const uint32_t sH = base + TILE_BYTES;            // h tile, later r*h
  const uint32_t sW = base + 2 * TILE_BYTES;        // weight ring
  const uint32_t bars = sW + WSTAGES * WSTAGE_BYTES;
  auto bar = [&](int i) { return bars + 8u * i; };
  const uint32_t tmem_slot = bars + 8u * BAR_COUNT;

  if (warp == WARP_MMA && lane == 0) {
    for (int s = 0; s < WSTAGES; ++s) {
      tc::mbar_init(bar(BAR_W_FULL + s), 1);
      tc::mbar_init(bar(BAR_W_EMPTY + s), 1);
    }
    tc::mbar_init(bar(BAR_X_FULL), N_AGG_WARPS);
    tc::mbar_init(bar(BAR_X_EMPTY), 1);
    tc::mbar_init(bar(BAR_H_FULL), 1);
    tc::mbar_init(bar(BAR_H_EMPTY), 1);
    tc::mbar_init(bar(BAR_GATES_DONE), 1);
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
    }
    __syncwarp();
    tc::tmem_alloc(tmem_slot, 512);
  }
  {   // biases -> smem (read by every epilogue thread for every tile)
    float* s_bias = reinterpret_cast<float*>(smem_al + (tmem_slot + 16 - base));
    // pre-scaled so that sigmoid / tanh need one FFMA before ex2:
    //   sigmoid(a + b) = 1 / (1 + 2^(-log2e*a - log2e*b)),  tanh(a + b) = 1 - 2 / (1 + 2^(2log2e*a + 2log2e*b))
    if (tid < 2 * HP) s_bias[tid] = -1.4426950408889634f * p.bg_pad[tid];
    else if (tid < 3 * HP) s_bias[tid] = (tid - 2 * HP) < p.h ? 2.8853900817779268f * p.bc[tid - 2 * HP] : 0.f;
  }
  tc::fence_before_sync();
  __syncthreads();
  tc::fence_after_sync();
  const uint32_t tmem_base = *reinterpret_cast<volatile uint32_t*>(smem_al + (tmem_slot - base));
  const int h = p.h;

  if (warp >= WARP_AGG0) {
    const int aw = warp - WARP_AGG0;                          // 0..7: rows 16*aw .. 16*aw+15
    // =============================== aggregation ===============================
    // warp w owns rows 16w .. 16w+15 of the tile; their message lists are one contiguous
    // range of tmsg.  lane -> (K chunk = lane / 8, 16-byte slot = lane % 8) of the row.
    const bool col_ok = lane * 4 < h;
    const int lane_c = col_ok ? lane : 0;
    auto load_ptr = [&](int tile) -> int {           // tptr[r0 .. r0+16] -> lanes 0..16 (clamped at N)
      int64_t rr = (int64_t)tile * BM + aw * 16 + lane;
      if (rr > p.N) rr = p.N;
      return lane <= 16 ? __ldg(p.tptr + rr) : 0;
    };
    uint8_t* xtile = smem_al + (sX - base);
    int it = 0;
    int my_ptr = blockIdx.x < p.n_tiles ? load_ptr(blockIdx.x) : 0;
    for (int tile = blockIdx.x; tile < p.n_tiles; tile += gridDim.x, ++it) {
      const int P0 = __shfl_sync(0xffffffffu, my_ptr, 0);
      const int P1 = __shfl_sync(0xffffffffu, my_ptr, 16);
      // index window: lane l holds tmsg[win + l]; one window feeds two 16-message batches
      int my_idx = (P0 + lane < P1) ? __ldg(p.tmsg + P0 + lane) : 0;
      int nidx = (P0 + 32 + lane < P1) ? __ldg(p.tmsg + P0 + 32 + lane) : 0;
      const int next_tile = tile + gridDim.x;
      const int nxt_ptr = next_tile < p.n_tiles ? load_ptr(next_tile) : 0;

      // Messages land in registers, 32 rows (8 KB per warp) per batch; the first batch is issued
      // before waiting for the x tile, so its latency overlaps MMA1 of the previous tile.
      auto load32 = [&](uint2 (&v)[32], int pb) {             // messages pb .. pb+31 (re-reads past P1)
#pragma unroll
        for (int j = 0; j < 32; ++j) {
          const int idx = __shfl_sync(0xffffffffu, my_idx, (pb + j < P1) ? j : 0);
          v[j] = __ldg(reinterpret_cast<const uint2*>(p.msg + (size_t)idx * HP) + lane_c);
        }
      };
      uint2 cur[32];
      if (P0 < P1) load32(cur, P0);
      if (it > 0) tc::mbar_wait_relaxed(bar(BAR_X_EMPTY), (it - 1) & 1);
      if (lane == 0) trace(p, it, aw * 2);
      {   // rows without any message are written as zeros up front (rare: padding nodes)
        const int nxt_p = __shfl_down_sync(0xffffffffu, my_ptr, 1);
        unsigned empty = __ballot_sync(0xffffffffu, lane < 16 && nxt_p == my_ptr);
        while (empty) {
          const int r = __ffs(empty) - 1;
          empty &= empty - 1;
          emit_row(xtile, aw * 16 + r, lane, make_float4(0.f, 0.f, 0.f, 0.f), col_ok);
        }
      }
      float4 acc = make_float4(0.f, 0.f, 0.f, 0.f);
#pragma unroll 1
      for (int pb = P0; pb < P1; pb += 32) {
        // once per batch: row of message pb+lane (count of row ends <= position) and whether
        // it is the last message of that row; per message only a mask bit is tested
        const int pos_l = pb + lane;
        int rowid = 0;
#pragma unroll
        for (int k = 1; k <= 16; ++k) rowid += (__shfl_sync(0xffffffffu, my_ptr, k) <= pos_l) ? 1 : 0;
        const int rend = __shfl_sync(0xffffffffu, my_ptr, min(rowid + 1, 16));
        const unsigned lastmask = __ballot_sync(0xffffffffu, pos_l < P1 && pos_l + 1 == rend);
        const unsigned valid = (P1 - pb >= 32) ? 0xffffffffu : ((1u << (P1 - pb)) - 1u);
#pragma unroll
        for (int j = 0; j < 32; ++j) {
          if (valid & (1u << j)) {
            const float2 lo = __half22float2(*reinterpret_cast<const __half2*>(&cur[j].x));
            const float2 hi = __half22float2(*reinterpret_cast<const __half2*>(&cur[j].y));
            acc.x += lo.x; acc.y += lo.y; acc.z += hi.x; acc.w += hi.y;
            if (lastmask & (1u << j)) {
              emit_row(xtile, aw * 16 + __shfl_sync(0xffffffffu, rowid, j), lane, acc, col_ok);
              acc = make_float4(0.f, 0.f, 0.f, 0.f);
            }
          }
        }
        if (aw == 0 && lane == 0) trace(p, it, 40 + ((pb - P0) >> 5) * 2);
        if (pb + 32 < P1) {                                   // next window (prefetched indices)
          my_idx = nidx;
          nidx = (pb + 64 + lane < P1) ? __ldg(p.tmsg + pb + 64 + lane) : 0;
          load32(cur, pb + 32);
          if (aw == 0 && lane == 0) trace(p, it, 41 + ((pb - P0) >> 5) * 2);
        }
      }
      tc::fence_proxy_async();
      __syncwarp();
      if (lane == 0) { tc::mbar_arrive(bar(BAR_X_FULL)); trace(p, it, aw * 2 + 1); }
      my_ptr = nxt_ptr;
    }
  }
  else if (warp == WARP_TMA) {
    // =============================== TMA producer ===============================
    if (lane == 0) {
      int fill = 0;                                           // running weight-stage fill count
      auto w_fill = [&](int gate_col, int cand_col) {
        const int s = fill % WSTAGES;
        if (fill >= WSTAGES) tc::mbar_wait_relaxed(bar(BAR_W_EMPTY + s), ((fill / WSTAGES) - 1) & 1);
        const uint32_t dst = sW + s * WSTAGE_BYTES;
        const uint32_t bytes = (gate_col >= 0 ? 2 * CHUNK_BYTES : 0) + (cand_col >= 0 ? CHUNK_BYTES : 0);
        tc::mbar_arrive_expect_tx(bar(BAR_W_FULL + s), bytes);
        if (gate_col >= 0) tc::tma_load_2d(dst, &tmWg, gate_col, 0, bar(BAR_W_FULL + s));                 // 256 rows
        if (cand_col >= 0) tc::tma_load_2d(dst + 2 * CHUNK_BYTES, &tmWc, cand_col, 0, bar(BAR_W_FULL + s)); // 128 rows
        ++fill;
      };
      int it = 0;
      for (int tile = blockIdx.x; tile < p.n_tiles; tile += gridDim.x, ++it) {
        // h tile (free once MMA2 of the previous tile has read r*h out of it)
        if (it > 0) tc::mbar_wait_relaxed(bar(BAR_H_EMPTY), (it - 1) & 1);
        tc::mbar_arrive_expect_tx(bar(BAR_H_FULL), TILE_BYTES);
#pragma unroll
        for (int kc = 0; kc < 4; ++kc)
          tc::tma_load_2d(sH + kc * CHUNK_BYTES, &tmH, kc * BKF, tile * BM, bar(BAR_H_FULL));
        // weights, in the order the MMA warp consumes them
        for (int kc = 0; kc < 4; ++kc) w_fill(kc * BKF, kc * BKF);            // x chunks: gates + cand_x
        for (int kc = 0; kc < 4; ++kc) w_fill(HP + kc * BKF, -1);             // h chunks: gates
        for (int kc = 0; kc < 4; ++kc) w_fill(-1, HP + kc * BKF);             // r*h chunks: cand
      }
    }
    __syncwarp();
  } else if (warp == WARP_MMA) {
    // =============================== MMA issuer ===============================
    if (lane == 0) {
      const uint32_t idesc256 = tc::umma_idesc(2, BM, 256);
      const uint32_t idesc128 = tc::umma_idesc(2, BM, 128);
      const uint32_t d_gates = tmem_base, d_cand = tmem_base + 256;
      int use = 0;                                            // running weight-stage use count
      auto w_wait = [&]() {
        const int s = use % WSTAGES;
        tc::mbar_wait(bar(BAR_W_FULL + s), (use / WSTAGES) & 1);
        tc::fence_after_sync();
        return sW + s * WSTAGE_BYTES;
      };
      auto w_release = [&]() {
        tc::umma_commit(bar(BAR_W_EMPTY + (use % WSTAGES)));
        ++use;
      };
      int it = 0;
      for (int tile = blockIdx.x; tile < p.n_tiles; tile += gridDim.x, ++it) {
        if (it > 0) tc::mbar_wait(bar(BAR_TMEM_EMPTY), (it - 1) & 1);
        trace(p, it, 16);
        tc::mbar_wait(bar(BAR_X_FULL), it & 1);
        tc::fence_after_sync();
        trace(p, it, 17);
        for (int kc = 0; kc < 4; ++kc) {                      // x chunks -> gates and cand
          const uint32_t w = w_wait();
          const uint32_t a = sX + kc * CHUNK_BYTES;
#pragma unroll
          for (int k = 0; k < 4; ++k)
            tc::umma_tf32(d_gates, tc::umma_desc_sw128(a + k * 32), tc::umma_desc_sw128(w + k * 32),
                          idesc256, (kc | k) != 0);
#pragma unroll
          for (int k = 0; k < 4; ++k)
            tc::umma_tf32(d_cand, tc::umma_desc_sw128(a + k * 32),
                          tc::umma_desc_sw128(w + 2 * CHUNK_BYTES + k * 32), idesc128, (kc | k) != 0);
          w_release();
        }
        tc::umma_commit(bar(BAR_X_EMPTY));                    // x tile may be refilled (next tile's AGG)
        trace(p, it, 18);
        tc::mbar_wait(bar(BAR_H_FULL), it & 1);
        tc::fence_after_sync();
        trace(p, it, 19);
        for (int kc = 0; kc < 4; ++kc) {                      // h chunks -> gates
          const uint32_t w = w_wait();
          const uint32_t a = sH + kc * CHUNK_BYTES;
#pragma unroll
          for (int k = 0; k < 4; ++k)
            tc::umma_tf32(d_gates, tc::umma_desc_sw128(a + k * 32), tc::umma_desc_sw128(w + k * 32),
                          idesc256, 1);
          w_release();
        }
        tc::umma_commit(bar(BAR_GATES_DONE));
        trace(p, it, 20);
        tc::mbar_wait(bar(BAR_RH_FULL), it & 1);              // epilogue replaced h by tf32(r*h)
        tc::fence_after_sync();
        trace(p, it, 21);
        for (int kc = 0; kc < 4; ++kc) {                      // r*h chunks -> cand
          const uint32_t w = w_wait();
          const uint32_t a = sH + kc * CHUNK_BYTES;
#pragma unroll
          for (int k = 0; k < 4; ++k)
            tc::umma_tf32(d_cand, tc::umma_desc_sw128(a + k * 32),
                          tc::umma_desc_sw128(w + 2 * CHUNK_BYTES + k * 32), idesc128, 1);
          w_release();
        }
        tc::umma_commit(bar(BAR_CAND_DONE));
        trace(p, it, 22);
      }
    }
    __syncwarp();
  }
  else if (warp < WARP_EPI0 + N_EPI_WARPS) {
    // =============================== epilogue (warps 0-7) ===============================
    const int ew = warp & 3;                                  // TMEM lane quarter this warp may access
    const int chalf = (warp - WARP_EPI0) >> 2;                // which 64 of the 128 columns
    const int r = ew * 32 + lane;                             // row within the tile = TMEM lane
    const uint32_t t_lane = tmem_base + ((uint32_t)(ew * 32) << 16);
    const float* s_bg = reinterpret_cast<const float*>(smem_al + (tmem_slot + 16 - base));
    const float* s_bc = s_bg + 2 * HP;
    uint8_t* hrow = smem_al + (sH - base) + (r >> 3) * 1024 + (r & 7) * 128;
    const bool leader = tid == WARP_EPI0 * 32;
    int it = 0;
    for (int tile = blockIdx.x; tile < p.n_tiles; tile += gridDim.x, ++it) {
      // ---- EPI1: both gates.  r*h (TF32) replaces h in smem as the next MMA operand;
      //      u*h and (1-u) are parked in TMEM (cols 384.. and 128..) so EPI2 never re-reads h.
      tc::mbar_wait(bar(BAR_GATES_DONE), it & 1);
      tc::fence_after_sync();
      if (leader) trace(p, it, 32);
#pragma unroll 1
      for (int c16 = 0; c16 < 4; ++c16) {                     // pass A: reset gate only (MMA2 waits on it)
        const int col0 = chalf * 64 + c16 * 16;               // first of 16 columns
        const int cc = col0 >> 5, q0 = (col0 & 31) >> 2;      // K chunk, first 16-byte slot
        uint32_t ar[16];
        tc::tmem_ld16(t_lane + col0, ar);
        float4 hv[4];
#pragma unroll
        for (int q = 0; q < 4; ++q)
          hv[q] = *reinterpret_cast<const float4*>(hrow + cc * CHUNK_BYTES + (((q0 + q) ^ (r & 7)) << 4));
        tc::tmem_ld_wait();
#pragma unroll
        for (int q = 0; q < 4; ++q) {
          const float4 sb = *reinterpret_cast<const float4*>(s_bg + col0 + q * 4);
          const float hh[4] = {hv[q].x, hv[q].y, hv[q].z, hv[q].w};
          const float bb[4] = {sb.x, sb.y, sb.z, sb.w};
          float o[4];
#pragma unroll
          for (int i = 0; i < 4; ++i) {
            const float rg = tc::rcp_approx(1.0f + tc::ex2_approx(fmaf(__uint_as_float(ar[q * 4 + i]), -kLog2e, bb[i])));
            o[i] = tc::to_tf32_fast(rg * hh[i]);
            ar[q * 4 + i] = __float_as_uint(hh[i]);           // park h itself for EPI2
          }
          *reinterpret_cast<float4*>(hrow + cc * CHUNK_BYTES + (((q0 + q) ^ (r & 7)) << 4)) =
              make_float4(o[0], o[1], o[2], o[3]);
        }
        tc::tmem_st16(t_lane + 384 + col0, ar);
      }
      tc::fence_proxy_async();
      __syncwarp();
      if (lane == 0) tc::mbar_arrive(bar(BAR_RH_FULL));
      if (leader) trace(p, it, 33);
#pragma unroll 1
      for (int c16 = 0; c16 < 4; ++c16) {                     // pass B (overlaps MMA2): update gate, in place
        const int col0 = chalf * 64 + c16 * 16;
        uint32_t au[16];
        tc::tmem_ld16(t_lane + 128 + col0, au);
        tc::tmem_ld_wait();
#pragma unroll
        for (int q = 0; q < 4; ++q) {
          const float4 sb = *reinterpret_cast<const float4*>(s_bg + HP + col0 + q * 4);
          const float bb[4] = {sb.x, sb.y, sb.z, sb.w};
#pragma unroll
          for (int i = 0; i < 4; ++i)
            au[q * 4 + i] = __float_as_uint(
                tc::rcp_approx(1.0f + tc::ex2_approx(fmaf(__uint_as_float(au[q * 4 + i]), -kLog2e, bb[i]))));
        }
        tc::tmem_st16(t_lane + 128 + col0, au);
      }
      tc::tmem_st_wait();

      // ---- EPI2: candidate and new state h' = c + u*(h - c), staged in the (now free) h tile
      //      and written with TMA
      tc::mbar_wait(bar(BAR_CAND_DONE), it & 1);
      tc::fence_after_sync();
      if (leader) trace(p, it, 34);
#pragma unroll 1
      for (int c16 = 0; c16 < 4; ++c16) {
        const int col0 = chalf * 64 + c16 * 16;
        const int cc = col0 >> 5, q0 = (col0 & 31) >> 2;
        uint32_t ac[16], uu[16], hh[16];
        tc::tmem_ld16(t_lane + 256 + col0, ac);
        tc::tmem_ld16(t_lane + 128 + col0, uu);
        tc::tmem_ld16(t_lane + 384 + col0, hh);
        tc::tmem_ld_wait();
#pragma unroll
        for (int q = 0; q < 4; ++q) {
          const float4 sb = *reinterpret_cast<const float4*>(s_bc + col0 + q * 4);
          const float bb[4] = {sb.x, sb.y, sb.z, sb.w};
          float o[4];
#pragma unroll
          for (int i = 0; i < 4; ++i) {
            const float c = fmaf(-2.0f, tc::rcp_approx(1.0f + tc::ex2_approx(
                                            fmaf(__uint_as_float(ac[q * 4 + i]), 2.0f * kLog2e, bb[i]))), 1.0f);
            o[i] = fmaf(__uint_as_float(uu[q * 4 + i]), __uint_as_float(hh[q * 4 + i]) - c, c);
            if (p.round_out) o[i] = tc::to_tf32_fast(o[i]);
          }
          *reinterpret_cast<float4*>(hrow + cc * CHUNK_BYTES + (((q0 + q) ^ (r & 7)) << 4)) =
              make_float4(o[0], o[1], o[2], o[3]);
        }
      }
      tc::fence_before_sync();                                // TMEM reads done
      tc::fence_proxy_async();                                // smem writes -> visible to the TMA unit
      __syncwarp();
      if (lane == 0) tc::mbar_arrive(bar(BAR_TMEM_EMPTY));
      tc::named_bar_sync(1, N_EPI_WARPS * 32);
      if (leader) {
#pragma unroll
        for (int cc = 0; cc < 4; ++cc)
          tc::tma_store_2d(&tmH, cc * BKF, tile * BM, sH + cc * CHUNK_BYTES);   // rows >= N, cols >= h clipped
        tc::tma_store_commit();
        tc::tma_store_wait_read();
        tc::mbar_arrive(bar(BAR_H_EMPTY));                    // h tile may be refilled
        trace(p, it, 35);
      }
    }
    if (leader) asm volatile("cp.async.bulk.wait_group 0;" ::: "memory");
  }

  tc::fence_before_sync();
  __syncthreads();
  if (warp == WARP_TMA) tc::tmem_dealloc(tmem_base, 512);
}

}  // namespace

bool ggnn_gru_fused_supported(const ggnn_ctx* ctx) { return ctx->hp == HP && ctx->h % 4 == 0; }

int ggnn_gru_fused_step(ggnn_ctx* ctx, const ggnn_plan& plan, const float* msg, float* h_state,
                        bool last_step, cudaStream_t st) {
  if (plan.N == 0) return GGNN_OK;
  GGNN_ENSURE_SMEM(ctx, GGNN_ATTR_GRU_FUSED, gru_fused_kernel, SMEM_BYTES);
  CUtensorMap tmH;
  GGNN_TRY(ggnn_make_tmap_2d(ctx, &tmH, h_state, plan.N, ctx->h, BM));
  gparams p{};
  p.msg = reinterpret_cast<const __half*>(msg); p.tptr = plan.tptr; p.tmsg = plan.tmsg; p.h_state = h_state;
  p.bg_pad = ctx->bg_pad; p.bc = ctx->bc;
  p.N = plan.N; p.h = ctx->h;
  p.n_tiles = (int)ceil_div64(plan.N, BM);
  p.round_out = last_step ? 0 : 1;
  p.dbg = ctx->dbg_trace;
  const int grid = p.n_tiles < ctx->sm_count ? p.n_tiles : ctx->sm_count;
  prof_scope ps(ctx, GGNN_STAGE_GRU_GATES, st);
  gru_fused_kernel<<<grid, NTHREADS, SMEM_BYTES, st>>>(p, tmH, ctx->tm_Wg256, ctx->tm_Wc);
  GGNN_LAUNCHED(ctx);
  return GGNN_OK;
}
