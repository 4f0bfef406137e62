// gru_pair.cu -- GGNN_PREC_F16, h <= 128: aggregation + GRU cell of one timestep on CTA PAIRS (tcgen05 cta_group::2)
// with every weight RESIDENT in shared memory for the whole launch.
//
// Why: gru16::gru_f16_kernel (f16.cu) re-streams 256-384 KB of weights from L2 for every 128-node tile and is bound by
// the latency of that stream (scripts/experiments/README.md).  A CTA pair working on two adjacent tiles as one 256-row
// MMA tile needs only HALF of every weight matrix's output columns per CTA:
//     [x,h] -> r, u  (N = 128 each: 64 columns of r and 64 of u per CTA)                      64 KB per CTA
//     [x,r*h] -> cand (N = 128: 64 columns each)                                             32 KB
//     W[E-1], W[2E-1] (the two word-order channels, N = 128 each: 64 columns each)           32 KB
// = 128 KB, loaded once.  What moves per tile is the state tile (32 KB in, 32 KB out) and the message rows.
//
// Per pair of tiles `it` (both CTAs run the same roles on their own tile; the even CTA's thread issues all MMAs):
//   TMA warp     lane 0: h tile -> H[it & 1] (two tiles deep); after the epilogue the TMA store of h' from the same buffer,
//                then its refill.  Lanes 1-31: L2 prefetch of the message rows three tiles ahead.
//   agg warps    x = sum of the tile's FP16 message rows (24 rows per warp in flight through registers, branch-free
//                summation) -> x tile.  The x tile is free again as soon as the MMAs reading it have retired (r*h does NOT
//                go there), so the gather of tile it+1 runs under the whole epilogue of tile it.
//   MMA          word-order messages of a regular tile: Zp = h[t-1] @ W[E-1], Zn = h[t+1] @ W[2E-1] (row-shifted A
//                descriptors on the resident h tile, scripts/micro/desc_shift.cu) into the r accumulator columns and the
//                spare 128 TMEM columns -- issued one tile AHEAD, right after the previous tile's candidate
//   epilogue     (after its last pass on tile it-1) z = Zp + b[E-1] (rows >= 1) + Zn + b[2E-1] (rows <= 126), split into
//                FP16 hi + lo and written back to TMEM in the tensor core's A-operand form (complete_z)
//   MMA          r = x @ Wr_x + h @ Wr_h + z @ Wr_x   (z: A operand FROM TENSOR MEMORY, same weights as x: the word-order
//                messages are part of x in the reference, but never pass through x's FP16 rounding here), h @ Wu_h;
//                then u's x and z parts, cand = (x + z) @ Wc_x
//   epilogue     1a: r -> f16(r*h) over h in the h tile (h stays in registers);  1b: u (16-bit fixed point) | h -> u's TMEM
//                word;  MMA cand += (r*h) @ Wc_h;  2: h' = c + u (h - c) over r*h in the h tile -> TMA store
// The row that falls off a shifted tile reads whatever lies next to the h tile in shared memory; that row of Zp / Zn is
// masked out in complete_z (row 0 has no t-1, row 127 no t+1), so no zeroed gaps are needed.
//
// Shared memory: x 32 KB | h 2 x 32 KB | weights 128 KB | barriers, the two edge biases: 230 656 B -- one background CTA
// of compact.cu's bitmask queue still fits next to it.  The GRU biases live in registers (6 per epilogue lane, broadcast by
// shuffle): no shared memory is left for them and both L1 and the constant bank cost a cache round trip per 16 columns.
// Phase trace: scripts/gpu_pair_trace.py (GP_TRACE).
#include "f16.cuh"
#include "tf32.cuh"
#include "tc.cuh"
#include <cuda_fp16.h>
#include <stdlib.h>
#include <string.h>

namespace gp {

constexpr float kLog2e = 1.4426950408889634f;
constexpr int HP = 128, BM = 128, BKH = 64;
constexpr int CHUNK = BM * BKH * 2;               // 16 KB: [128 x 64] FP16, 128B-swizzled
constexpr int TILE = 2 * CHUNK;                   // 32 KB
constexpr int N_AGG = 6, N_EPI = 8;
constexpr int AB = 24;                            // message rows in flight per aggregation warp (2 registers per lane each)
constexpr int WARP_MMA = 8, WARP_TMA = 9, WARP_AGG0 = 10;
constexpr int NTHREADS = (WARP_AGG0 + N_AGG) * 32;   // 512: 4 warps per scheduler partition at 96 registers (see f16.cu)

// shared memory map (bytes from the 1024-aligned base)
constexpr int OFF_X = 0;                          // x tile, later f16(r*h)
constexpr int OFF_H = TILE;                       // two h tiles
constexpr int OFF_W = 3 * TILE;
constexpr int W_RUX = 0;                          // two K chunks of 16 KB, each [64 n of r | 64 n of u] x 64 k: x -> r, u
constexpr int W_RUH = W_RUX + TILE;               // h -> r, u
constexpr int W_CX = W_RUH + TILE;                // [64 n x 128 k] as two K chunks of 8 KB: x -> cand
constexpr int W_CRH = W_CX + TILE / 2;            // r*h -> cand
constexpr int W_ZP = W_CRH + TILE / 2;            // W[E-1]
constexpr int W_ZN = W_ZP + TILE / 2;             // W[2E-1]
constexpr int W_BYTES = W_ZN + TILE / 2;          // 128 KB
constexpr int OFF_BAR = OFF_W + W_BYTES;
constexpr int OFF_EB = OFF_BAR + 256;             // b[E-1] | b[2E-1], FP32
constexpr int SMEM_BYTES = OFF_EB + 2 * HP * 4;   // 230 656: leaves room for one background CTA of compact.cu's queue
static_assert(W_BYTES == 128 * 1024, "weight area");

enum {
  BAR_W_READY = 0,      // even CTA: the resident weights of both CTAs have landed
  BAR_H_FULL,           // [2] even CTA: both h tiles of the pair
  BAR_H_STORE = BAR_H_FULL + 2,   // [2] local: the 8 epilogue warps have written h' into the h tile
  BAR_Z_DONE = BAR_H_STORE + 2,   // multicast commit: Zp / Zn of the tile are in TMEM
  BAR_X_FULL,           // even CTA: 2 x 6 aggregation warps: both x tiles are whole
  BAR_R_DONE,           // multicast commit: reset gate complete (it goes first: r*h is on the critical path)
  BAR_U_DONE,           // multicast commit: update gate complete
  BAR_X_READ,           // multicast commit: every MMA reading x has retired: the buffer may take the next tile's sums
  BAR_RH_FULL,          // even CTA: 16 epilogue warps
  BAR_CAND_DONE,        // multicast commit: candidate complete
  BAR_TMEM_FREE,        // even CTA: 16 epilogue warps: u / cand of the pair consumed AND z of the next pair in A-operand form
  BAR_COUNT
};
static_assert(8 * BAR_COUNT + 8 <= 256, "barrier area");

struct params {
  const __half* msg;        // [*, 128] FP16 message rows
  const int32_t* tptr;      // [N+1]
  const int32_t* tmsg;      // [nnz]
  const int32_t* regular;   // [n_tiles] 1 = word-order tile (plan_bits.cu) or null
  const float* btab;        // [640] bias table of the slot (bias_tab below), device memory
  float* h_out;             // [N, h] FP32 result (last step) or null
  int64_t N;
  int h;
  int n_tiles;
  int zp_row, zn_row;       // first row of W[E-1] / W[2E-1] in the packed edge weights
  unsigned long long* dbg;  // phase trace of cluster 0 (ggnn_debug_trace) or null
  int dbg_wide;             // the buffer also has the whole-grid region (GP_WIDE)
  int dbg_cluster;          // the cluster whose phases are traced (env GGNN_PAIR_TRACE_CLUSTER, default 0)
};

// Every bias the kernel adds.  There is no shared memory left for the GRU biases and reading them through L1 (which the
// message rows stream through) or the constant bank (indexed LDC) cost a cache round trip per 16 columns -- measured 1.5-2.4 us
// per group -- so every epilogue lane keeps 6 of its warp's 192 values in registers and the warp broadcasts them by shuffle.
// The two edge biases of the completion pass (1 KB) do live in shared memory.
struct bias_tab {
  float g[2 * HP];          // -log2e * bg (r | u)
  float c[HP];              // 2 log2e * bc
  float bp[HP];             // bias of channel E-1 (zero padded)
  float bn[HP];             // bias of channel 2E-1
};

// development aid: event `ev` of iteration `it`, CTA `blockIdx.x` (0 or 1), nanoseconds.
// GP_TRACE exists in builds with -DGGNN_PAIR_TRACE (GGNN_NVCC_EXTRA; scripts/gpu_pair_trace.py).
// Built with -DGGNN_PAIR_WIDE_TRACE (GGNN_NVCC_EXTRA, scripts/gpu_pair_wide.py) the traced cluster can be chosen
// (GGNN_PAIR_TRACE_CLUSTER) and GP_WIDE records every CTA's entry / set-up / weights-in / pair starts / exit; in the shipped
// build those hooks are compiled out -- even never taken they cost the kernel 5 % when it runs alone (3.00 -> 3.15 ms per
// 4 096 graphs x 8 steps, measured)
#ifdef GGNN_PAIR_WIDE_TRACE
#define GP_TRACED_CLUSTER ((int)(blockIdx.x >> 1) == p.dbg_cluster)
#else
#define GP_TRACED_CLUSTER (blockIdx.x < 2)
#endif
#if defined(GGNN_PAIR_TRACE) || defined(GGNN_PAIR_WIDE_TRACE)
#define GP_TRACE(ev)                                                                            \
  do {                                                                                          \
    if (p.dbg && GP_TRACED_CLUSTER && it < 12) {                                                \
      unsigned long long t_;                                                                    \
      asm volatile("mov.u64 %0, %%globaltimer;" : "=l"(t_));                                    \
      p.dbg[(((blockIdx.x & 1) * 12 + it) * 32 + (ev))] = t_;                                   \
    }                                                                                           \
  } while (0)
#else
#define GP_TRACE(ev) do { } while (0)       // (the ~25 hooks are compiled out of the shipped build: another 4 %, 3.06 -> 2.93 ms)
#endif

// whole-grid trace behind the cluster-0 one: per CTA 48 slots (0 entry, 1 set-up done, 2 weights in, 3 exit, 8 + it: MMAs of pair it start)
#ifndef GGNN_PAIR_WIDE_TRACE
#define GP_WIDE(slot) do { } while (0)
#else
#define GP_WIDE(slot)                                                                           \
  do {                                                                                          \
    if (p.dbg && p.dbg_wide) {                                                                  \
      unsigned long long t_;                                                                    \
      asm volatile("mov.u64 %0, %%globaltimer;" : "=l"(t_));                                    \
      p.dbg[2 * 12 * 32 + blockIdx.x * 48 + (slot)] = t_;                                       \
    }                                                                                           \
  } while (0)
#endif

__device__ __forceinline__ uint32_t sw_off(int r, int slot) {     // 16-byte slot `slot` of row r in a swizzled chunk
  return (r >> 3) * 1024 + (r & 7) * 128 + ((slot ^ (r & 7)) << 4);
}
__device__ __noinline__ void emit_row(uint8_t* xtile, int r, int lane, float4 acc) {
  const int kc = lane >> 4, l16 = lane & 15;
  __half2 lo = __floats2half2_rn(acc.x, acc.y), hi = __floats2half2_rn(acc.z, acc.w);
  uint2 pk;
  pk.x = *reinterpret_cast<uint32_t*>(&lo); pk.y = *reinterpret_cast<uint32_t*>(&hi);
  *reinterpret_cast<uint2*>(xtile + kc * CHUNK + sw_off(r, l16 >> 1) + (l16 & 1) * 8) = pk;
}
// st.shared.v2.b32 under a predicate (guaranteed predication: no branch)
__device__ __forceinline__ void st_shared_v2_if(uint32_t addr, uint32_t a, uint32_t b, uint32_t pred) {
  asm volatile(
      "{\n\t"
      ".reg .pred p;\n\t"
      "setp.ne.u32 p, %3, 0;\n\t"
      "@p st.shared.v2.b32 [%0], {%1, %2};\n\t"
      "}" ::"r"(addr), "r"(a), "r"(b), "r"(pred) : "memory");
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

// The word-order messages of a regular tile, z = Zp + b[E-1] (rows >= 1) + Zn + b[2E-1] (rows <= 126), become a SECOND A
// operand of the gate MMAs instead of being added to the FP16 x tile (that sum, rounded to FP16, doubled the mode's error on
// word-order graphs): z is split into FP16 hi + lo (22 significant bits) and written back into the Zn columns in the tensor
// core's A-operand form -- K-group c (16 columns of z) = 8 columns of hi at 16c, 8 columns of lo at 16c + 8, exactly the
// columns this iteration has just read.  Done by the EPILOGUE warps (two per TMEM lane quarter, 64 columns each, thread =
// row) for the NEXT tile, right after the current tile's last pass: its accumulators are long complete by then and the
// aggregation warps are left alone.  reg = false (the other tile of the pair is the regular one): zeros.
__device__ __forceinline__ void complete_z(uint32_t tmem_base, int q4, int lane, int c_begin, const float* s_eb, bool reg,
                                           uint32_t zbar, uint32_t parity) {
  const int r = q4 * 32 + lane;
  const uint32_t t_lane = tmem_base + ((uint32_t)(q4 * 32) << 16);
  // row 0 has no t-1, row 127 no t+1: masks instead of branches (what sits in those accumulator rows may be NaN)
  const uint32_t mp = (reg && r >= 1) ? 0xffffffffu : 0u, mn = (reg && r <= BM - 2) ? 0xffffffffu : 0u;
  const float fp = (reg && r >= 1) ? 1.f : 0.f, fn = (reg && r <= BM - 2) ? 1.f : 0.f;
  tc::mbar_wait(zbar, parity);
  tc::fence_after_sync();
  uint32_t zpv[2][16], znv[2][16];
  tc::tmem_ld16(t_lane + c_begin * 16, zpv[0]);
  tc::tmem_ld16(t_lane + 384 + c_begin * 16, znv[0]);
#pragma unroll
  for (int cc = 0; cc < 4; ++cc) {                        // this warp's 64 columns of z; the next group's loads go first
    const int col0 = (c_begin + cc) * 16;
    uint32_t (&zp)[16] = zpv[cc & 1];
    uint32_t (&zn)[16] = znv[cc & 1];
    tc::tmem_ld_wait();
    if (cc < 3) {
      tc::tmem_ld16(t_lane + col0 + 16, zpv[(cc + 1) & 1]);
      tc::tmem_ld16(t_lane + 384 + col0 + 16, znv[(cc + 1) & 1]);
    }
    uint32_t wh[8], wl[8];
#pragma unroll
    for (int i = 0; i < 16; i += 4) {
      const float4 bp = *reinterpret_cast<const float4*>(s_eb + col0 + i);
      const float4 bn = *reinterpret_cast<const float4*>(s_eb + HP + col0 + i);
      const float bb[4] = {fmaf(fp, bp.x, fn * bn.x), fmaf(fp, bp.y, fn * bn.y), fmaf(fp, bp.z, fn * bn.z), fmaf(fp, bp.w, fn * bn.w)};
#pragma unroll
      for (int j = 0; j < 4; j += 2) {
      const float za = (__uint_as_float(zp[i + j] & mp) + __uint_as_float(zn[i + j] & mn)) + bb[j];
      const float zb = (__uint_as_float(zp[i + j + 1] & mp) + __uint_as_float(zn[i + j + 1] & mn)) + bb[j + 1];
      const __half2 hi = __floats2half2_rn(za, zb);
      const float2 hf = __half22float2(hi);
      const __half2 lo = __floats2half2_rn(za - hf.x, zb - hf.y);
      wh[(i + j) >> 1] = *reinterpret_cast<const uint32_t*>(&hi);
      wl[(i + j) >> 1] = *reinterpret_cast<const uint32_t*>(&lo);
      }
    }
    tc::tmem_st8(t_lane + 384 + col0, wh);
    tc::tmem_st8(t_lane + 384 + col0 + 8, wl);
  }
  tc::tmem_st_wait();
  tc::fence_before_sync();                            // TMEM traffic done before the MMAs read / overwrite these columns
}

__global__ void __cluster_dims__(2, 1, 1) __maxnreg__(96)
gru_pair_kernel(const params p, const __grid_constant__ CUtensorMap tmH, const __grid_constant__ CUtensorMap tmWg,
                const __grid_constant__ CUtensorMap tmWcq, const __grid_constant__ CUtensorMap tmWeq) {
  extern __shared__ __align__(1024) uint8_t smem_raw[];
  const int tid = threadIdx.x, warp = tid >> 5, lane = tid & 31;
  const uint32_t base = tc::smem_u32(smem_raw);          // 1024-byte aligned (no static shared memory in this kernel)
  uint8_t* smem_al = smem_raw;
  const uint32_t sX = base + OFF_X, sW = base + OFF_W, bars = base + OFF_BAR;
  auto sH = [&](int hb) { return base + OFF_H + (uint32_t)hb * TILE; };
  auto bar = [&](int i) { return bars + 8u * i; };
  const uint32_t tmem_slot = bars + 8u * BAR_COUNT;
  const uint32_t rank = tc::cluster_ctarank();
  const bool lead = rank == 0;
  // pair j of this cluster = tiles (tile0 - rank) + j * gridDim.x and the one after it; this CTA's is tile0 + j * gridDim.x.
  // A pair exists while its EVEN tile does (the odd one may lie past the end: loaded as zeros, stored clipped).
  const int tile0 = (int)blockIdx.x;                     // = 2 * (blockIdx.x >> 1) + rank
  if (tid == 0) GP_WIDE(0);
#ifdef GGNN_PAIR_WIDE_TRACE
  if (tid == 0 && p.dbg && p.dbg_wide) {
    unsigned smid_;
    asm volatile("mov.u32 %0, %%smid;" : "=r"(smid_));
    p.dbg[2 * 12 * 32 + blockIdx.x * 48 + 4] = smid_;
  }
#endif
  auto live = [&](int tile) { return tile - (int)rank < p.n_tiles; };
  auto is_reg = [&](int tile) { return p.regular && tile < p.n_tiles && __ldg(p.regular + tile) != 0; };

  if (warp == WARP_MMA && lane == 0) {
    tc::mbar_init(bar(BAR_W_READY), 1);
    for (int s = 0; s < 2; ++s) {
      tc::mbar_init(bar(BAR_H_FULL + s), 1);
      tc::mbar_init(bar(BAR_H_STORE + s), N_EPI);
    }
    tc::mbar_init(bar(BAR_Z_DONE), 1);
    tc::mbar_init(bar(BAR_X_FULL), 2 * N_AGG);
    tc::mbar_init(bar(BAR_R_DONE), 1);
    tc::mbar_init(bar(BAR_U_DONE), 1);
    tc::mbar_init(bar(BAR_X_READ), 1);
    tc::mbar_init(bar(BAR_RH_FULL), 2 * N_EPI);
    tc::mbar_init(bar(BAR_CAND_DONE), 1);
    tc::mbar_init(bar(BAR_TMEM_FREE), 2 * N_EPI);
    tc::fence_barrier_init();
  }
  if (warp == WARP_TMA) {
    if (lane == 0) {
      tc::tma_prefetch_desc(&tmH); tc::tma_prefetch_desc(&tmWg); tc::tma_prefetch_desc(&tmWcq); tc::tma_prefetch_desc(&tmWeq);
    }
    __syncwarp();
    tc::tmem_alloc_pair(tmem_slot, 512);
  }
  float* s_eb = reinterpret_cast<float*>(smem_al + OFF_EB);
  for (int i = tid; i < 2 * HP; i += NTHREADS) s_eb[i] = p.btab[3 * HP + i];
  __syncthreads();
  tc::fence_before_sync();
  tc::cluster_sync_all();                                 // barriers of both CTAs initialised, TMEM allocated
  tc::fence_after_sync();
  const uint32_t tmem_base = *reinterpret_cast<volatile uint32_t*>(smem_al + (tmem_slot - base));
  if (tid == 0) GP_WIDE(1);

  if (warp >= WARP_AGG0) {
    // =============================== aggregation (6 warps), completion (4 of them) ===============================
    const int aw = warp - WARP_AGG0;
    const int R = BM / N_AGG + (aw < BM % N_AGG ? 1 : 0);                      // rows of this warp: 22, 22, 21 x 4
    const int row0 = aw * (BM / N_AGG) + min(aw, BM % N_AGG);
    auto load_ptr = [&](int tile) -> int {                    // tptr[r0 .. r0+R] -> lanes 0..R (clamped at N)
      int64_t rr = (int64_t)tile * BM + row0 + lane;
      if (rr > p.N) rr = p.N;
      return lane <= R ? __ldg(p.tptr + rr) : 0x7fffffff;
    };
    uint8_t* xtile = smem_al + OFF_X;
    const int slot_l = (lane & 15) >> 1;                      // this lane's 4 columns: chunk lane >> 4, 16-byte slot, half of it
    const uint32_t x_lane = sX + (uint32_t)((lane >> 4) * CHUNK + (lane & 1) * 8);
    int my_ptr = live(tile0) ? load_ptr(tile0) : 0;
    int it = 0;
    for (int tile = tile0; live(tile); tile += gridDim.x, ++it) {
      const int P0 = __shfl_sync(0xffffffffu, my_ptr, 0);
      const int P1 = __shfl_sync(0xffffffffu, my_ptr, R);
      int my_idx = (P0 + lane < P1) ? __ldg(p.tmsg + P0 + lane) : 0;
      int nidx = (P0 + AB + lane < P1) ? __ldg(p.tmsg + P0 + AB + lane) : 0;
      const int nxt_ptr = live(tile + gridDim.x) ? load_ptr(tile + gridDim.x) : 0;
      auto load32 = [&](uint2 (&v)[AB], int pb) {             // messages pb .. pb+AB-1 (re-reads past P1)
#pragma unroll
        for (int j = 0; j < AB; ++j) {
          const int idx = __shfl_sync(0xffffffffu, my_idx, (pb + j < P1) ? j : 0);
          v[j] = __ldg(reinterpret_cast<const uint2*>(p.msg + (size_t)idx * HP) + lane);
        }
      };
      uint2 cur[AB];
      if (P0 < P1) load32(cur, P0);                           // in flight while the x buffer is still busy
      if (warp == WARP_AGG0 && lane == 0) GP_TRACE(16);
      if (it > 0) tc::mbar_wait_relaxed(bar(BAR_X_READ), (it - 1) & 1);        // the previous tile's x has been consumed
      if (warp == WARP_AGG0 && lane == 0) GP_TRACE(17);
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
      for (int pb = P0; pb < P1; pb += AB) {
        if (warp == WARP_AGG0 && lane == 0) { if (pb == P0) GP_TRACE(27); else if (pb == P0 + AB) GP_TRACE(28); else GP_TRACE(29); }
        const int pos_l = pb + lane;
        int rowid = 0;
#pragma unroll
        for (int k = 1; k <= BM / N_AGG + 1; ++k)             // lanes above R hold INT_MAX: never counted
          rowid += (__shfl_sync(0xffffffffu, my_ptr, k) <= pos_l) ? 1 : 0;
        const int rend = __shfl_sync(0xffffffffu, my_ptr, min(rowid + 1, R));
        const unsigned lastmask = __ballot_sync(0xffffffffu, pos_l < P1 && pos_l + 1 == rend);
        const unsigned valid = (P1 - pb >= AB) ? ((1u << AB) - 1u) : ((1u << (P1 - pb)) - 1u);
        // Straight-line code, no branch per message (one warp per scheduler cannot hide branch latency: the branchy
        // form of this loop cost ~250 cycles per message): EVERY valid message stores the running sum of its row --
        // the row's last message stores last and wins -- and the sum restarts after a row's last message.  Messages
        // past P1 (re-read duplicates) only pollute an accumulator nobody stores.
#pragma unroll
        for (int j = 0; j < AB; ++j) {
          const float2 lo = __half22float2(*reinterpret_cast<const __half2*>(&cur[j].x));
          const float2 hi = __half22float2(*reinterpret_cast<const __half2*>(&cur[j].y));
          acc.x += lo.x; acc.y += lo.y; acc.z += hi.x; acc.w += hi.y;
          const int rj = row0 + __shfl_sync(0xffffffffu, rowid, j);
          const uint32_t dst = x_lane + (uint32_t)((rj >> 3) * 1024 + (rj & 7) * 128) + ((uint32_t)(slot_l ^ (rj & 7)) << 4);
          const __half2 plo = __floats2half2_rn(acc.x, acc.y), phi = __floats2half2_rn(acc.z, acc.w);
          st_shared_v2_if(dst, *reinterpret_cast<const uint32_t*>(&plo), *reinterpret_cast<const uint32_t*>(&phi), (valid >> j) & 1u);
          const bool last = (lastmask >> j) & 1u;
          acc.x = last ? 0.f : acc.x; acc.y = last ? 0.f : acc.y; acc.z = last ? 0.f : acc.z; acc.w = last ? 0.f : acc.w;
        }
        if (warp == WARP_AGG0 && lane == 0) { if (pb == P0) GP_TRACE(30); else if (pb == P0 + AB) GP_TRACE(31); }
        if (pb + AB < P1) {
          my_idx = nidx;
          nidx = (pb + 2 * AB + lane < P1) ? __ldg(p.tmsg + pb + 2 * AB + lane) : 0;
          load32(cur, pb + AB);
        }
      }
      tc::fence_proxy_async();                                // every writer: its rows -> visible to the tensor core
      if (warp == WARP_AGG0 && lane == 0) GP_TRACE(18);
      __syncwarp();
      if (lane == 0) tc::mbar_arrive_cta0(bar(BAR_X_FULL));
      if (warp == WARP_AGG0 && lane == 0) GP_TRACE(20);
      my_ptr = nxt_ptr;
    }
  }
  else if (warp == WARP_TMA) {
    // =============================== TMA: resident weights, h tiles in, h' tiles out ===============================
    if (lane == 0) {
      if (lead) tc::mbar_arrive_expect_tx(bar(BAR_W_READY), 2u * W_BYTES);
      const int r64 = (int)rank * 64;
      for (int kc = 0; kc < 2; ++kc) {
        tc::tma_load_2d_pair(sW + W_RUX + kc * CHUNK, &tmWg, kc * BKH, r64, bar(BAR_W_READY));                          // r
        tc::tma_load_2d_pair(sW + W_RUX + kc * CHUNK + CHUNK / 2, &tmWg, kc * BKH, HP + r64, bar(BAR_W_READY));         // u
        tc::tma_load_2d_pair(sW + W_RUH + kc * CHUNK, &tmWg, HP + kc * BKH, r64, bar(BAR_W_READY));
        tc::tma_load_2d_pair(sW + W_RUH + kc * CHUNK + CHUNK / 2, &tmWg, HP + kc * BKH, HP + r64, bar(BAR_W_READY));
        tc::tma_load_2d_pair(sW + W_CX + kc * (CHUNK / 2), &tmWcq, kc * BKH, r64, bar(BAR_W_READY));
        tc::tma_load_2d_pair(sW + W_CRH + kc * (CHUNK / 2), &tmWcq, HP + kc * BKH, r64, bar(BAR_W_READY));
        tc::tma_load_2d_pair(sW + W_ZP + kc * (CHUNK / 2), &tmWeq, kc * BKH, p.zp_row + r64, bar(BAR_W_READY));
        tc::tma_load_2d_pair(sW + W_ZN + kc * (CHUNK / 2), &tmWeq, kc * BKH, p.zn_row + r64, bar(BAR_W_READY));
      }
      auto load_h = [&](int j, int tile) {
        const int hb = j & 1;
        if (lead) tc::mbar_arrive_expect_tx(bar(BAR_H_FULL + hb), 2u * TILE);
        tc::tma_load_2d_pair(sH(hb), &tmH, 0, tile * BM, bar(BAR_H_FULL + hb));            // rows >= N: zero filled
        tc::tma_load_2d_pair(sH(hb) + CHUNK, &tmH, BKH, tile * BM, bar(BAR_H_FULL + hb));
      };
      if (live(tile0)) load_h(0, tile0);
      if (live(tile0 + gridDim.x)) load_h(1, tile0 + gridDim.x);
      int it = 0;
      for (int tile = tile0; live(tile); tile += gridDim.x, ++it) {
        const int hb = it & 1;
        tc::mbar_wait_relaxed(bar(BAR_H_STORE + hb), (it >> 1) & 1);     // h' written (and fenced) by the 8 epilogue warps
        GP_TRACE(24);
        tc::tma_store_2d(&tmH, 0, tile * BM, sH(hb));                    // rows >= N clipped by the map
        tc::tma_store_2d(&tmH, BKH, tile * BM, sH(hb) + CHUNK);
        tc::tma_store_commit();
        if (live(tile + 2 * gridDim.x)) {
          tc::tma_store_wait_read();                                     // the buffer has been read: refill it
          GP_TRACE(25);
          load_h(it + 2, tile + 2 * gridDim.x);
        }
      }
      asm volatile("cp.async.bulk.wait_group 0;" ::: "memory");
    } else {
      // lanes 1-31: pull the message rows of the tile THREE iterations ahead into L2 (they are 254 random 256-byte rows
      // of a buffer larger than L2; the aggregation warps can keep only 24 rows each in flight, so what a batch costs
      // is the latency of its slowest row)
      auto pf = [&](int tile) {
        if (tile >= p.n_tiles) return;
        const int64_t r0 = (int64_t)tile * BM, r1 = min(p.N, r0 + BM);
        const int P0 = __ldg(p.tptr + r0), P1 = __ldg(p.tptr + r1);
        for (int q = P0 + lane - 1; q < P1; q += 31) {
          const char* row = reinterpret_cast<const char*>(p.msg + (size_t)__ldg(p.tmsg + q) * HP);
          asm volatile("prefetch.global.L2 [%0];" ::"l"(row));
          asm volatile("prefetch.global.L2 [%0];" ::"l"(row + 128));
        }
      };
      pf(tile0); pf(tile0 + gridDim.x); pf(tile0 + 2 * gridDim.x);
      int it = 0;
      for (int tile = tile0; live(tile); tile += gridDim.x, ++it) {
        tc::mbar_wait_relaxed(bar(BAR_H_STORE + (it & 1)), (it >> 1) & 1);
        pf(tile + 3 * gridDim.x);
      }
    }
    __syncwarp();
  } else if (warp == WARP_MMA) {
    // =============================== MMA issuer (even CTA) ===============================
    if (lead && lane == 0) {
      const uint32_t idesc128 = tc::umma_idesc(0 /*F16*/, 2 * BM, 128);
      const uint32_t d_r = tmem_base, d_u = tmem_base + 128, d_cand = tmem_base + 256, d_zp = tmem_base, d_zn = tmem_base + 384;
      // D (+)= A tile [256 x 128] @ W^T, both K chunks; wstride = bytes between the two K chunks of this CTA's W half
      auto tile_mma = [&](uint32_t d, uint32_t a0, uint32_t w0, uint32_t wstride, uint32_t idesc, bool first) {
#pragma unroll
        for (int kc = 0; kc < 2; ++kc)
#pragma unroll
          for (int k = 0; k < 4; ++k)
            tc::umma_f16_pair(d, tc::umma_desc_sw128(a0 + kc * CHUNK + k * 32), tc::umma_desc_sw128(w0 + kc * wstride + k * 32),
                              idesc, (first && kc == 0 && k == 0) ? 0u : 1u);
      };
      // D += z @ W^T with z as the A operand in tensor memory (complete_z's layout): the same weights as x
      auto z_mma = [&](uint32_t d, uint32_t w0, uint32_t wstride, uint32_t idesc) {
#pragma unroll
        for (int c = 0; c < 8; ++c) {
          const uint64_t bd = tc::umma_desc_sw128(w0 + (c >> 2) * wstride + (c & 3) * 32);
          tc::umma_f16_pair_ts(d, tmem_base + 384 + 16 * c, bd, idesc, 1u);
          tc::umma_f16_pair_ts(d, tmem_base + 384 + 16 * c + 8, bd, idesc, 1u);
        }
      };
      auto z_phase = [&](int it, int t0) {                    // word-order messages of pair `it` (its h tiles must be in)
        const int hb = it & 1;
        tc::mbar_wait_cluster(bar(BAR_H_FULL + hb), (it >> 1) & 1);
        tc::fence_after_sync();
        if (is_reg(t0) || is_reg(t0 + 1)) {
          tile_mma(d_zp, sH(hb) - 128, sW + W_ZP, CHUNK / 2, idesc128, true);       // row t reads h[t-1]
          tile_mma(d_zn, sH(hb) + 128, sW + W_ZN, CHUNK / 2, idesc128, true);       // row t reads h[t+1]
        }
        tc::umma_commit_pair(bar(BAR_Z_DONE));
      };
      tc::mbar_wait_cluster(bar(BAR_W_READY), 0);
      tc::fence_after_sync();
      GP_WIDE(2);
      if (live(tile0)) z_phase(0, tile0);
      int it = 0;
      for (int tile = tile0; live(tile); tile += gridDim.x, ++it) {
        const uint32_t h0 = sH(it & 1);
        tc::mbar_wait_cluster(bar(BAR_TMEM_FREE), it & 1);        // u, cand of the previous pair consumed; z of this one in place
        GP_TRACE(0);
        tc::mbar_wait_cluster(bar(BAR_X_FULL), it & 1);                            // x final in both CTAs, Zp / Zn consumed
        tc::fence_after_sync();
        GP_TRACE(1);
        if (it < 40) GP_WIDE(8 + it);
        const bool preg = is_reg(tile) || is_reg(tile + 1);
        tile_mma(d_r, sX, sW + W_RUX, CHUNK, idesc128, true);                       // x -> r
        tile_mma(d_r, h0, sW + W_RUH, CHUNK, idesc128, false);                      // h -> r
        if (preg) z_mma(d_r, sW + W_RUX, CHUNK, idesc128);                          // z (hi, lo) -> r
        tile_mma(d_u, h0, sW + W_RUH + CHUNK / 2, CHUNK, idesc128, true);           // h -> u (h's last reader: r*h will replace h)
        tc::umma_commit_pair(bar(BAR_R_DONE));
        tile_mma(d_u, sX, sW + W_RUX + CHUNK / 2, CHUNK, idesc128, false);          // x -> u
        if (preg) z_mma(d_u, sW + W_RUX + CHUNK / 2, CHUNK, idesc128);              // z (hi, lo) -> u
        tc::umma_commit_pair(bar(BAR_U_DONE));
        tile_mma(d_cand, sX, sW + W_CX, CHUNK / 2, idesc128, true);                 // x -> cand
        if (preg) z_mma(d_cand, sW + W_CX, CHUNK / 2, idesc128);                    // z (hi, lo) -> cand
        tc::umma_commit_pair(bar(BAR_X_READ));                                      // the x buffers may take the next tile's sums
        GP_TRACE(2);
        tc::mbar_wait_cluster(bar(BAR_RH_FULL), it & 1);                            // r drained, f16(r*h) in both x buffers
        tc::fence_after_sync();
        GP_TRACE(3);
        tile_mma(d_cand, h0, sW + W_CRH, CHUNK / 2, idesc128, false);               // r*h (over h in the h tile) -> cand
        tc::umma_commit_pair(bar(BAR_CAND_DONE));
        GP_TRACE(4);
        if (live(tile + gridDim.x)) z_phase(it + 1, tile + gridDim.x);
        GP_TRACE(5);                                         // under this pair's last epilogue pass
      }
    }
    __syncwarp();
  }
  else {
    // =============================== epilogue (warps 0-7 of both CTAs) ===============================
    const int ew = warp & 3;                                  // TMEM lane quarter this warp may access
    const int chalf = warp >> 2;                              // which 64 columns = which K chunk of the tiles
    const int r = ew * 32 + lane;                             // row within the tile = TMEM lane
    const uint32_t t_lane = tmem_base + ((uint32_t)(ew * 32) << 16);
    // this warp's 3 x 64 biases, two per lane and gate; value of column c (0..63): lane c & 31 of register c >> 5
    const float* bt = p.btab + chalf * 64 + lane;
    const float br0 = __ldg(bt), br1 = __ldg(bt + 32), bu0 = __ldg(bt + HP), bu1 = __ldg(bt + HP + 32);
    const float bc0 = __ldg(bt + 2 * HP), bc1 = __ldg(bt + 2 * HP + 32);
    auto z_ready = [&](int it, int tile) {                    // z of pair `it` -> A operand form; then "TMEM free" phase `it`
      if (is_reg(tile) || is_reg(tile ^ 1))
        complete_z(tmem_base, ew, lane, chalf * 4, s_eb, is_reg(tile), bar(BAR_Z_DONE), it & 1);
      __syncwarp();
      if (lane == 0) tc::mbar_arrive_cta0(bar(BAR_TMEM_FREE));
    };
    if (live(tile0)) z_ready(0, tile0);
    int it = 0;
    for (int tile = tile0; live(tile); tile += gridDim.x, ++it) {
      const int hb = it & 1;
      uint8_t* hrow = smem_al + OFF_H + hb * TILE + chalf * CHUNK;      // + sw_off(r, slot): this thread's h row, its 64 columns
      // ---- pass 1a: reset gate.  f16(r*h) replaces h in the h tile (every MMA reading h has retired with r; the x
      //      buffer is already being refilled for the next tile); the h it displaces stays in registers for pass 1b
      tc::mbar_wait(bar(BAR_R_DONE), it & 1);
      tc::fence_after_sync();
      if (tid == 0) GP_TRACE(8);
      // (all passes: the TMEM loads of group c16 + 1 are issued before the math of group c16 -- a load takes ~0.3 us)
      uint32_t hkeep[4][8];
      uint32_t av[2][16];
      tc::tmem_ld16(t_lane + chalf * 64, av[0]);
#pragma unroll
      for (int c16 = 0; c16 < 4; ++c16) {
        uint32_t (&ar)[16] = av[c16 & 1];
        uint4* s0 = reinterpret_cast<uint4*>(hrow + sw_off(r, 2 * c16));
        uint4* s1 = reinterpret_cast<uint4*>(hrow + sw_off(r, 2 * c16 + 1));
        const uint4 h0 = *s0, h1 = *s1;
        hkeep[c16][0] = h0.x; hkeep[c16][1] = h0.y; hkeep[c16][2] = h0.z; hkeep[c16][3] = h0.w;
        hkeep[c16][4] = h1.x; hkeep[c16][5] = h1.y; hkeep[c16][6] = h1.z; hkeep[c16][7] = h1.w;
        const float brs = c16 < 2 ? br0 : br1;
        const int l0 = (c16 & 1) * 16;
        tc::tmem_ld_wait();
        if (c16 < 3) tc::tmem_ld16(t_lane + chalf * 64 + (c16 + 1) * 16, av[(c16 + 1) & 1]);
        uint32_t rh[8];
#pragma unroll
        for (int i = 0; i < 16; i += 2) {
          const float2 hv = __half22float2(*reinterpret_cast<const __half2*>(&hkeep[c16][i >> 1]));
          const float ra = tc::rcp_approx(1.0f + tc::ex2_approx(fmaf(__uint_as_float(ar[i]), -kLog2e, __shfl_sync(0xffffffffu, brs, l0 + i))));
          const float rb = tc::rcp_approx(1.0f + tc::ex2_approx(fmaf(__uint_as_float(ar[i + 1]), -kLog2e, __shfl_sync(0xffffffffu, brs, l0 + i + 1))));
          const __half2 pr = __floats2half2_rn(ra * hv.x, rb * hv.y);
          rh[i >> 1] = *reinterpret_cast<const uint32_t*>(&pr);
        }
        *s0 = make_uint4(rh[0], rh[1], rh[2], rh[3]);
        *s1 = make_uint4(rh[4], rh[5], rh[6], rh[7]);
      }
      tc::fence_before_sync();                                // r has been read: the next pair's Zp may land there
      tc::fence_proxy_async();
      __syncwarp();
      if (lane == 0) tc::mbar_arrive_cta0(bar(BAR_RH_FULL));
      if (tid == 0) GP_TRACE(9);
      // ---- pass 1b (under the r*h MMAs): update gate; u (16-bit fixed point) and h go into u's TMEM word
      tc::mbar_wait(bar(BAR_U_DONE), it & 1);
      tc::fence_after_sync();
      tc::tmem_ld16(t_lane + 128 + chalf * 64, av[0]);
#pragma unroll
      for (int c16 = 0; c16 < 4; ++c16) {
        const int col0 = chalf * 64 + c16 * 16;
        uint32_t (&au)[16] = av[c16 & 1];
        const float bus = c16 < 2 ? bu0 : bu1;
        const int l0 = (c16 & 1) * 16;
        tc::tmem_ld_wait();
        if (c16 < 3) tc::tmem_ld16(t_lane + 128 + col0 + 16, av[(c16 + 1) & 1]);
#pragma unroll
        for (int i = 0; i < 16; i += 2) {
          const float ua = tc::rcp_approx(1.0f + tc::ex2_approx(fmaf(__uint_as_float(au[i]), -kLog2e, __shfl_sync(0xffffffffu, bus, l0 + i))));
          const float ub = tc::rcp_approx(1.0f + tc::ex2_approx(fmaf(__uint_as_float(au[i + 1]), -kLog2e, __shfl_sync(0xffffffffu, bus, l0 + i + 1))));
          // u in (0, 1) as 16-bit fixed point (|error| <= 2^-17; FP16 would lose up to 2^-12): 2^23 + 65535 u, low 16 bits
          au[i] = (__float_as_uint(fmaf(ua, 65535.0f, 8388608.0f)) & 0xffffu) | (hkeep[c16][i >> 1] << 16);
          au[i + 1] = (__float_as_uint(fmaf(ub, 65535.0f, 8388608.0f)) & 0xffffu) | (hkeep[c16][i >> 1] & 0xffff0000u);
        }
        tc::tmem_st16(t_lane + 128 + col0, au);
      }
      tc::tmem_st_wait();
      if (tid == 0) GP_TRACE(10);
      // ---- pass 2: candidate and new state h' = c + u (h - c): FP16 over r*h in the h tile (the MMAs reading it have
      //      retired) for the TMA store; the unrounded FP32 straight to h_out on the last step
      tc::mbar_wait(bar(BAR_CAND_DONE), it & 1);
      tc::fence_after_sync();
      if (tid == 0) GP_TRACE(11);
      const int64_t grow = (int64_t)tile * BM + r;
      uint32_t uv[2][16];
      tc::tmem_ld16(t_lane + 256 + chalf * 64, av[0]);
      tc::tmem_ld16(t_lane + 128 + chalf * 64, uv[0]);
#pragma unroll
      for (int c16 = 0; c16 < 4; ++c16) {
        const int col0 = chalf * 64 + c16 * 16;
        uint32_t (&ac)[16] = av[c16 & 1];
        uint32_t (&uu)[16] = uv[c16 & 1];
        const float bcs = c16 < 2 ? bc0 : bc1;
        const int l0 = (c16 & 1) * 16;
        tc::tmem_ld_wait();
        if (c16 < 3) {
          tc::tmem_ld16(t_lane + 256 + col0 + 16, av[(c16 + 1) & 1]);
          tc::tmem_ld16(t_lane + 128 + col0 + 16, uv[(c16 + 1) & 1]);
        }
        float o0[8], o1[8];
#pragma unroll
        for (int i = 0; i < 16; ++i) {
          const float c = fmaf(-2.0f, tc::rcp_approx(1.0f + tc::ex2_approx(
                                          fmaf(__uint_as_float(ac[i]), 2.0f * kLog2e, __shfl_sync(0xffffffffu, bcs, l0 + i)))), 1.0f);
          const float u = (__uint_as_float(0x4B000000u | (uu[i] & 0xffffu)) - 8388608.0f) * (1.0f / 65535.0f);
          const float hv = __half2float(__ushort_as_half((unsigned short)(uu[i] >> 16)));
          const float v = fmaf(u, hv - c, c);
          if (i < 8) o0[i] = v; else o1[i - 8] = v;
        }
        *reinterpret_cast<uint4*>(hrow + sw_off(r, 2 * c16)) = pack8(o0);
        *reinterpret_cast<uint4*>(hrow + sw_off(r, 2 * c16 + 1)) = pack8(o1);
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
      tc::fence_proxy_async();                                // h' -> visible to the TMA unit
      __syncwarp();
      if (lane == 0) tc::mbar_arrive(bar(BAR_H_STORE + hb));
      if (tid == 0) GP_TRACE(12);
      if (live(tile + gridDim.x)) z_ready(it + 1, tile + gridDim.x);          // (nobody waits for a phase after the last pair)
      if (tid == 0) GP_TRACE(19);
    }
  }

  tc::fence_before_sync();
  tc::cluster_sync_all();                                     // both CTAs are done with the pair's TMEM, barriers and shared memory
  if (warp == WARP_TMA) tc::tmem_dealloc_pair(tmem_base, 512);
  if (tid == 0) GP_WIDE(3);
}

// the bias_tab of one slot, built on the device (the weights may never have been on the host): the epilogue's exponent forms
__global__ void bias_tab_kernel(const float* __restrict__ bg_pad, const float* __restrict__ bc, const float* __restrict__ b_prev,
                                const float* __restrict__ b_next, int h, float* __restrict__ out) {
  const int i = blockIdx.x * blockDim.x + threadIdx.x;
  if (i >= 5 * HP) return;
  float v;
  if (i < 2 * HP) v = -kLog2e * bg_pad[i];
  else if (i < 3 * HP) v = (i - 2 * HP) < h ? 2.0f * kLog2e * bc[i - 2 * HP] : 0.f;
  else if (i < 4 * HP) v = b_prev ? b_prev[i - 3 * HP] : 0.f;
  else v = b_next ? b_next[i - 4 * HP] : 0.f;
  out[i] = v;
}
static_assert(sizeof(bias_tab) == 5 * HP * sizeof(float), "bias_tab layout");

}  // namespace gp

// =========================================================================================
// can a 2-CTA cluster of this kernel be resident at all (asked once per ctx; no: gru16::gru_f16_kernel serves the step)
static bool pair_fits(ggnn_ctx* ctx) {
  if (ctx->pair_fit < 0) {
    ctx->pair_fit = 0;
    if (cudaFuncSetAttribute(gp::gru_pair_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, gp::SMEM_BYTES) == cudaSuccess) {
      ctx->attr_mask |= 1u << GGNN_ATTR_GRU_PAIR;
      cudaLaunchConfig_t cfg = {};
      cfg.gridDim = dim3(2, 1, 1); cfg.blockDim = dim3(gp::NTHREADS, 1, 1); cfg.dynamicSmemBytes = gp::SMEM_BYTES;
      cudaLaunchAttribute at[1];
      at[0].id = cudaLaunchAttributeClusterDimension;
      at[0].val.clusterDim.x = 2; at[0].val.clusterDim.y = 1; at[0].val.clusterDim.z = 1;
      cfg.attrs = at; cfg.numAttrs = 1;
      int n = 0;
      if (cudaOccupancyMaxActiveClusters(&n, gp::gru_pair_kernel, &cfg) == cudaSuccess && n > 0) ctx->pair_fit = 1;
    }
    cudaGetLastError();                                      // a refusal is an answer, not an error of the call
  }
  return ctx->pair_fit == 1;
}

bool ggnn_gru_pair_enabled(ggnn_ctx* ctx, int slot) {
  const char* e = getenv("GGNN_GRU_PAIR");                   // A-B switch, read per call (tests flip it): 0 = off
  if (e && e[0] == '0') return false;
  return ggnn_f16_supported(ctx) && ctx->pair_ok && ctx->edge[slot].Wt_f16 && ctx->edge[slot].B_pad16 &&
         ctx->two_e >= 2 && ctx->two_e % 2 == 0 && ctx->sm_count >= 2 && pair_fits(ctx);
}

int ggnn_gru_pair_pack(ggnn_ctx* ctx, int slot, bool with_gru, cudaStream_t st) {
  using namespace gp;
  ggnn_edge_weights& ew = ctx->edge[slot];
  GGNN_TRY(ggnn_make_tmap_2d_f16(ctx, &ew.tm_Wt16q, ew.Wt_f16, (int64_t)ctx->two_e * HP, HP, BKH, HP / 2));
  ctx->btab_ok[slot] = false;                               // rebuilt by the next launch
  if (with_gru) {
    ctx->pair_ok = false;
    for (int s2 = 0; s2 < GGNN_MAX_SLOTS; ++s2) ctx->btab_ok[s2] = false;
    GGNN_TRY(ggnn_make_tmap_2d_f16(ctx, &ctx->tm_Wc16q, ctx->Wct_f16, HP, 2 * HP, BKH, HP / 2));
    GGNN_TRY(ggnn_make_tmap_2d_f16(ctx, &ctx->tm_Wg16q, ctx->Wgt_f16, 2 * HP, 2 * HP, BKH, HP / 2));
    ctx->pair_ok = true;
  }
  return GGNN_OK;
}

int ggnn_gru_pair_launch(ggnn_ctx* ctx, int slot, const ggnn_plan& plan, const void* msg16, void* state16, float* h_out,
                         cudaStream_t st) {
  using namespace gp;
  GGNN_ENSURE_SMEM(ctx, GGNN_ATTR_GRU_PAIR, gru_pair_kernel, SMEM_BYTES);
  const ggnn_edge_weights& ew = ctx->edge[slot];
  CUtensorMap tmH;
  GGNN_TRY(ggnn_make_tmap_2d_f16(ctx, &tmH, state16, plan.N, HP, BKH, BM));
  const int e_prev = ctx->two_e / 2 - 1, e_next = ctx->two_e - 1;
  constexpr int NT = (int)(sizeof(bias_tab) / sizeof(float));
  if (!ctx->btab16) GGNN_CUDA(ctx, cudaMalloc(&ctx->btab16, (size_t)GGNN_MAX_SLOTS * NT * sizeof(float)));
  float* d = ctx->btab16 + (size_t)slot * NT;
  if (!ctx->btab_ok[slot]) {      // once per set_weights
    bias_tab_kernel<<<5, HP, 0, st>>>(ctx->bg_pad, ctx->bc, ew.B ? ew.B_pad16 + (size_t)e_prev * HP : nullptr,
                                      ew.B ? ew.B_pad16 + (size_t)e_next * HP : nullptr, ctx->h, d);
    GGNN_LAUNCHED(ctx);
    ctx->btab_ok[slot] = true;
  }
  params g{};
  g.msg = reinterpret_cast<const __half*>(msg16); g.tptr = plan.tptr; g.tmsg = plan.tmsg;
  g.regular = plan.tile_regular;
  g.btab = d;
  g.h_out = h_out;
  g.N = plan.N; g.h = ctx->h; g.n_tiles = (int)((plan.N + BM - 1) / BM);
  g.zp_row = e_prev * HP; g.zn_row = e_next * HP;
  g.dbg_wide = ctx->dbg_trace_wide;
  { const char* e = getenv("GGNN_PAIR_TRACE_CLUSTER"); g.dbg_cluster = (e && ctx->dbg_trace) ? atoi(e) : 0; }
  g.dbg = h_out ? nullptr : ctx->dbg_trace;               // trace a step that is not the last one (no FP32 output)
  const int n_pairs = (g.n_tiles + 1) / 2;
  const int clusters = n_pairs < ctx->sm_count / 2 ? n_pairs : ctx->sm_count / 2;
  gru_pair_kernel<<<2 * clusters, NTHREADS, SMEM_BYTES, st>>>(g, tmH, ctx->tm_Wg16q, ctx->tm_Wc16q, ew.tm_Wt16q);
  GGNN_LAUNCHED(ctx);
  return GGNN_OK;
}
