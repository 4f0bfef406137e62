// wide16.cu -- GGNN_PREC_F16 for hidden sizes above 128 (C4: h = 256; the WSJ shapes: h = 400).
//
// The three contractions of a timestep
//   MSG    M[r, :]  = f16(h[src(r), :]) @ f16(W[type]) + B[type]                      -> FP16 rows
//   GATES  g = sigmoid([x, h] @ Wg + bg);  rh = f16(g[:, :h] * h);  u = f16(g[:, h:])
//   CAND   c = tanh([x, rh] @ Wc + bc);    h' = c + u * (h - c)                        -> FP16 state (+ FP32 result)
// run on ONE persistent, warp-specialised tcgen05 kernel (wgemm16_kernel; kind::f16, FP32 accumulators in TMEM).
// The hidden size is padded to hw = a multiple of 64 (one 128-byte swizzle row of FP16).  For hw <= 256 the GRU half
// (GATES + CAND) runs as one fused kernel instead (gru_w16_kernel below; gru_w16x2_kernel is its CTA-pair variant).
// With a training tape the unfused kernels also store x, r, u, c of the step as FP32 rows (forward of BASELINE config 5).
//
// Tiling of wgemm16_kernel: the weight matrix is cut into column slices of NW <= 256 output columns whose K-major FP16
// panel [NW x K] fits 128 KB of shared memory.  A CTA owns a contiguous range of (slice, 128-row tile)
// work items, slice-major, so the panel stays RESIDENT while the CTA streams row tiles under it: only
// the activations move through the 4 x 16 KB chunk ring (one [128 x 64] FP16 chunk per stage, by TMA
// when the rows are contiguous, by 16-byte cp.async when they are gathered through msg_src).  At
// h = 256 a both-operands-streamed 128 x 256 tile needs 96 B/clk/SM from L2 at the kind::f16 rate --
// more than twice what the L2 delivers per SM -- while the resident panel needs 32 (MSG) / 64 (GRU).
//
// Warp roles (13 warps): 0-3 activation producers; 4 MMA issuer (also (re)loads the weight panel when the
// (edge type, slice) key changes); 5-12 two epilogue teams of four warps (TMEM lane quarter = warp % 4),
// each team owning alternate 64-column groups and its own 16 KB swizzled staging box that leaves by TMA
// store.  Two TMEM accumulators (2 x 256 columns): the epilogue of item i runs under the MMAs of i+1.
#include "f16.cuh"
#include "tf32.cuh"
#include "tc.cuh"
#include "pack_t.cuh"
#include <cuda_fp16.h>
#include <stdlib.h>

namespace w16 {

constexpr int BM = 128;
constexpr int BKH = 64;                         // halves per K chunk = one 128-byte swizzle row
constexpr int CHUNK = BM * BKH * 2;             // 16 KB
constexpr int STAGES = 4;
constexpr int W_BYTES = 128 * 1024;             // resident weight panel
constexpr int NPROD = 128;
constexpr int WARP_MMA = 4;
constexpr int WARP_EPI0 = 5;
constexpr int N_EPI_WARPS = 8;
constexpr int NTHREADS = (WARP_EPI0 + N_EPI_WARPS) * 32;     // 416
constexpr int MAX_NW = 256;
constexpr int SMEM_BYTES = W_BYTES + STAGES * CHUNK + 2 * CHUNK + 1024 /*align*/ + 256 /*barriers*/ + MAX_NW * 4;
static_assert(SMEM_BYTES <= 232448, "shared memory budget");
constexpr float kLog2e = 1.4426950408889634f;

enum { MODE_MSG = 0, MODE_GATES = 1, MODE_CAND = 2 };

enum {
  BAR_A_FULL = 0,                       // [STAGES]
  BAR_A_EMPTY = BAR_A_FULL + STAGES,    // [STAGES] tcgen05.commit
  BAR_W_FULL = BAR_A_EMPTY + STAGES,    // TMA bytes of the panel
  BAR_W_FREE,                           // every MMA reading the old panel has retired
  BAR_ACC_FULL,                         // [2]
  BAR_ACC_EMPTY = BAR_ACC_FULL + 2,     // [2] 8 epilogue warps
  BAR_COUNT = BAR_ACC_EMPTY + 2
};
static_assert(8 * BAR_COUNT + 8 <= 256, "barrier area");

struct params {
  const __half* A_gather;     // MSG, gathered rows: state16 [N, hw]
  const int32_t* gather;      // msg_src [tiles*128] (-1 = padding) or null when the rows come by TMA
  int tiles_per_type;         // MSG by TMA (dense-row plan): tile t reads node rows (t % tiles_per_type) * 128
  const int32_t* tile_type;   // MSG: [tiles]
  const int32_t* n_tiles_dev; // MSG: device-side tile count, else null
  int n_tiles;                // row tiles when n_tiles_dev is null
  int KC, KC1;                // K chunks in total / in the first K segment
  int Ntot, NW, S;            // output columns, slice width, number of slices
  int NB;                     // row blocks (>= 1): the activation sweeps run block by block, see item_iter
  const float* bias;          // [types][Ntot] zero-padded FP32
  int hw, h;
  int64_t N;                  // GATES / CAND: valid node rows
  const __half* state16;      // GATES, CAND: h  [N, hw]
  const __half* u16;          // CAND: u [N, hw]
  float* h_out;               // CAND, last step: FP32 result [N, h] or null
  float* tape_g;              // training tape, [N, h] FP32 or null: GATES r, CAND c
  float* tape_u;              // GATES: u
};

__device__ __forceinline__ uint32_t sw_off(int r, int slot) {
  return (r >> 3) * 1024 + (r & 7) * 128 + ((slot ^ (r & 7)) << 4);
}
__device__ __forceinline__ void unpack8(const uint4 v, float* f) {
  const __half2* hp = reinterpret_cast<const __half2*>(&v);
#pragma unroll
  for (int i = 0; i < 4; ++i) { const float2 t = __half22float2(hp[i]); f[2 * i] = t.x; f[2 * i + 1] = t.y; }
}
__device__ __forceinline__ uint4 pack8(const float* f) {
  uint4 v;
  __half2* hp = reinterpret_cast<__half2*>(&v);
#pragma unroll
  for (int i = 0; i < 4; ++i) hp[i] = __floats2half2_rn(f[2 * i], f[2 * i + 1]);
  return v;
}

// This CTA's work items.  An item = (slice of the weights, 128-row tile); the activations are swept once per slice.
// With ONE sweep order over all rows (slice-major), S > 1 sweeps of activations that do not stay in L2 come from DRAM
// every time -- at h = 400 (98 MB of [x | h] against an L2 that holds ~60 MB of data every SM reads) ncu counted
// 700 MB of DRAM reads per CAND launch (7 sweeps) and 656 MB per GATES launch.  So the rows are cut into NB blocks whose
// activations do fit, blocks outermost; inside a block every CTA takes a contiguous range of the block's (slice, tile)
// items (slice-major: the resident panel still changes only once or twice per block), and because every CTA owns a share
// of EVERY block the whole chip works on one block at a time.  NB == 1 is the plain slice-major order.
struct item_iter {
  int nb, tb, n_tiles, S, blk, tiles_b;
  unsigned cta, grid;
  int64_t j, je;
  __device__ __forceinline__ item_iter(int nb_, int n_tiles_, int S_, unsigned cta_, unsigned grid_)
      : nb(nb_), tb((n_tiles_ + nb_ - 1) / nb_), n_tiles(n_tiles_), S(S_), blk(-1), tiles_b(1), cta(cta_), grid(grid_), j(0), je(0) {
    next_block();
  }
  __device__ __forceinline__ void next_block() {
    for (++blk; blk < nb; ++blk) {
      tiles_b = min(tb, n_tiles - blk * tb);
      if (tiles_b <= 0) { blk = nb; break; }
      const int64_t total = (int64_t)tiles_b * S;
      j = total * cta / grid;
      je = total * (cta + 1) / grid;
      if (j < je) return;
    }
  }
  __device__ __forceinline__ bool valid() const { return blk < nb; }
  __device__ __forceinline__ void advance() { if (++j >= je) next_block(); }
  __device__ __forceinline__ int slice() const { return (int)(j / tiles_b); }
  __device__ __forceinline__ int tile() const { return blk * tb + (int)(j % tiles_b); }
};

template <int MODE>
__global__ void __launch_bounds__(NTHREADS, 1)
wgemm16_kernel(const params p, const __grid_constant__ CUtensorMap tmA1, const __grid_constant__ CUtensorMap tmA2,
               const __grid_constant__ CUtensorMap tmW, const __grid_constant__ CUtensorMap tmO0,
               const __grid_constant__ CUtensorMap tmO1) {
  extern __shared__ uint8_t smem_raw[];
  const int tid = threadIdx.x, warp = tid >> 5, lane = tid & 31;
  const uint32_t base = (tc::smem_u32(smem_raw) + 1023u) & ~1023u;
  uint8_t* smem_al = smem_raw + (base - tc::smem_u32(smem_raw));
  const uint32_t sW = base;
  const uint32_t sA = base + W_BYTES;
  const uint32_t sO = sA + STAGES * CHUNK;
  const uint32_t bars = sO + 2 * CHUNK;
  auto bar = [&](int i) { return bars + 8u * i; };
  const uint32_t tmem_slot = bars + 8u * BAR_COUNT;
  float* s_bias = reinterpret_cast<float*>(smem_al + (bars + 256 - base));   // [NW]

  const bool a_tma = p.gather == nullptr;
  const int n_tiles = p.n_tiles_dev ? *p.n_tiles_dev : p.n_tiles;
  const int nb = p.NB > 0 ? p.NB : 1;

  if (warp == WARP_MMA && lane == 0) {
    for (int s = 0; s < STAGES; ++s) {
      tc::mbar_init(bar(BAR_A_FULL + s), a_tma ? 1 : NPROD);
      tc::mbar_init(bar(BAR_A_EMPTY + s), 1);
    }
    tc::mbar_init(bar(BAR_W_FULL), 1);
    tc::mbar_init(bar(BAR_W_FREE), 1);
    for (int s = 0; s < 2; ++s) {
      tc::mbar_init(bar(BAR_ACC_FULL + s), 1);
      tc::mbar_init(bar(BAR_ACC_EMPTY + s), N_EPI_WARPS);
    }
    tc::fence_barrier_init();
  }
  if (warp == 0) {
    if (lane == 0) {
      tc::tma_prefetch_desc(&tmA1); tc::tma_prefetch_desc(&tmA2); tc::tma_prefetch_desc(&tmW);
      tc::tma_prefetch_desc(&tmO0); tc::tma_prefetch_desc(&tmO1);
    }
    __syncwarp();
    tc::tmem_alloc(tmem_slot, 512);
  }
  tc::fence_before_sync();
  __syncthreads();
  tc::fence_after_sync();
  const uint32_t tmem_base = *reinterpret_cast<volatile uint32_t*>(smem_al + (tmem_slot - base));

  if (warp < 4) {
    // =============================== activation producers ===============================
    uint32_t c = 0;                                       // running chunk counter of this CTA
    if (a_tma) {
      if (tid == 0) {
        for (item_iter q(nb, n_tiles, p.S, blockIdx.x, gridDim.x); q.valid(); q.advance()) {
          const int t = q.tile();
          const int row0 = (MODE == MODE_MSG ? (t % p.tiles_per_type) : t) * BM;
          for (int kc = 0; kc < p.KC; ++kc, ++c) {
            const int s = c % STAGES;
            if (c >= STAGES) tc::mbar_wait(bar(BAR_A_EMPTY + s), ((c / STAGES) - 1) & 1);
            tc::mbar_arrive_expect_tx(bar(BAR_A_FULL + s), CHUNK);
            const bool seg2 = kc >= p.KC1;
            tc::tma_load_2d(sA + s * CHUNK, seg2 ? &tmA2 : &tmA1, (seg2 ? kc - p.KC1 : kc) * BKH, row0,
                            bar(BAR_A_FULL + s));
          }
        }
      }
    } else {
      const int slot = tid & 7;                           // 16-byte slot within the 128-byte chunk row
      const int rbase = tid >> 3;                         // rows rbase + 16*j, j = 0..7
      uint32_t a_off[8];
#pragma unroll
      for (int j = 0; j < 8; ++j) a_off[j] = sw_off(rbase + 16 * j, slot);
      for (item_iter q(nb, n_tiles, p.S, blockIdx.x, gridDim.x); q.valid(); q.advance()) {
        const int t = q.tile();
        int src[8];
#pragma unroll
        for (int j = 0; j < 8; ++j) src[j] = __ldg(p.gather + (size_t)t * BM + rbase + 16 * j);
        for (int kc = 0; kc < p.KC; ++kc, ++c) {
          const int s = c % STAGES;
          if (c >= STAGES) tc::mbar_wait(bar(BAR_A_EMPTY + s), ((c / STAGES) - 1) & 1);
          const uint32_t a_st = sA + s * CHUNK;
#pragma unroll
          for (int j = 0; j < 8; ++j) {
            const bool ok = src[j] >= 0;
            tc::cp_async16(a_st + a_off[j],
                           ok ? (const void*)(p.A_gather + (size_t)src[j] * p.hw + kc * BKH + slot * 8)
                              : (const void*)p.A_gather,
                           ok ? 16u : 0u);
          }
          tc::cp_async_arrive(bar(BAR_A_FULL + s));
        }
      }
    }
  } else if (warp == WARP_MMA) {
    // =============================== MMA issuer + weight panel ===============================
    if (lane == 0) {
      uint32_t c = 0;
      int it = 0, cur_key = -1, w_gen = 0;
      const uint32_t panel_chunk = (uint32_t)p.NW * 128u;  // bytes of one K chunk of the panel
      for (item_iter q(nb, n_tiles, p.S, blockIdx.x, gridDim.x); q.valid(); q.advance(), ++it) {
        const int sl = q.slice(), t = q.tile();
        const int type = MODE == MODE_MSG ? __ldg(p.tile_type + t) : 0;
        const int key = type * p.S + sl;
        const int n0 = sl * p.NW;
        const int nw = min(p.NW, p.Ntot - n0);
        if (key != cur_key) {
          if (w_gen > 0) {                                 // MMAs still reading the old panel
            tc::umma_commit(bar(BAR_W_FREE));
            tc::mbar_wait(bar(BAR_W_FREE), (w_gen - 1) & 1);
          }
          tc::mbar_arrive_expect_tx(bar(BAR_W_FULL), (uint32_t)p.KC * panel_chunk);
          for (int kc = 0; kc < p.KC; ++kc)
            tc::tma_load_2d(sW + kc * panel_chunk, &tmW, kc * BKH, type * p.Ntot + n0, bar(BAR_W_FULL));
          tc::mbar_wait(bar(BAR_W_FULL), w_gen & 1);
          ++w_gen;
          cur_key = key;
        }
        const int buf = it & 1;
        if (it >= 2) tc::mbar_wait(bar(BAR_ACC_EMPTY + buf), ((it >> 1) - 1) & 1);
        tc::fence_after_sync();
        const uint32_t d = tmem_base + buf * MAX_NW;
        const uint32_t idesc = tc::umma_idesc(0 /*F16*/, BM, nw);
        for (int kc = 0; kc < p.KC; ++kc, ++c) {
          const int s = c % STAGES;
          tc::mbar_wait(bar(BAR_A_FULL + s), (c / STAGES) & 1);
          tc::fence_after_sync();
          const uint32_t a0 = sA + s * CHUNK, w0 = sW + kc * panel_chunk;
#pragma unroll
          for (int k = 0; k < 4; ++k)                      // UMMA K = 16 halves = 32 bytes
            tc::umma_f16(d, tc::umma_desc_sw128(a0 + k * 32), tc::umma_desc_sw128(w0 + k * 32), idesc,
                         (kc | k) != 0);
          tc::umma_commit(bar(BAR_A_EMPTY + s));
        }
        tc::umma_commit(bar(BAR_ACC_FULL + buf));
      }
    }
    __syncwarp();
  } else {
    // =============================== epilogue teams ===============================
    const int q = warp & 3;                                // TMEM lane quarter this warp may access
    const int team = (warp - WARP_EPI0) >> 2;
    const int r = q * 32 + lane;                           // row within the tile = TMEM lane
    const int e256 = tid - WARP_EPI0 * 32;                 // 0..255 over both teams
    const bool leader = (e256 & 127) == 0;
    const uint32_t t_lane = tmem_base + ((uint32_t)(q * 32) << 16);
    const uint32_t box_addr = sO + team * CHUNK;
    uint8_t* box = smem_al + (box_addr - base);
    int it = 0, cur_key = -1;
    for (item_iter qi(nb, n_tiles, p.S, blockIdx.x, gridDim.x); qi.valid(); qi.advance(), ++it) {
      const int sl = qi.slice(), t = qi.tile();
      const int type = MODE == MODE_MSG ? __ldg(p.tile_type + t) : 0;
      const int key = type * p.S + sl;
      const int n0 = sl * p.NW;
      const int nw = min(p.NW, p.Ntot - n0);
      if (key != cur_key) {
        tc::named_bar_sync(3, N_EPI_WARPS * 32);
        for (int j = e256; j < nw; j += N_EPI_WARPS * 32) s_bias[j] = __ldg(p.bias + (size_t)type * p.Ntot + n0 + j);
        tc::named_bar_sync(3, N_EPI_WARPS * 32);
        cur_key = key;
      }
      const int buf = it & 1;
      const int row0 = t * BM;
      const int64_t grow = (int64_t)row0 + r;
      const bool row_ok = MODE == MODE_MSG ? true : grow < p.N;
      const int ngroups = nw >> 6;
      // A slice of ONE 64-column group (K = 2 hw > 512, i.e. h = 400: the resident panel is only 64 columns wide) would
      // leave team 1 without work and make the kernel epilogue-bound (measured: 355 TFLOP/s in GATES at h = 400): the
      // teams then alternate ITEMS instead of groups -- item `it` uses accumulator it & 1 and belongs to team it & 1
      const bool alt = ngroups == 1;
      const bool mine = !alt || (it & 1) == team;
      const int g_first = alt ? 0 : team, g_step = alt ? 1 : 2;
      // GATES / CAND: this row's h (and u) values of a 64-column group; the first group's are requested BEFORE the
      // accumulator wait so their latency is not part of the per-item epilogue chain
      uint4 hv[8], uv[8];
      auto fetch = [&](int g) {
        const int colbase = n0 + g * 64;
        const bool is_r = colbase < p.hw;
        const int hcol = is_r ? colbase : colbase - p.hw;
        if (MODE == MODE_GATES && is_r) {
#pragma unroll
          for (int j = 0; j < 8; ++j)
            hv[j] = row_ok ? __ldg(reinterpret_cast<const uint4*>(p.state16 + grow * p.hw + hcol) + j)
                           : make_uint4(0u, 0u, 0u, 0u);
        } else if (MODE == MODE_CAND) {
#pragma unroll
          for (int j = 0; j < 8; ++j) {
            hv[j] = row_ok ? __ldg(reinterpret_cast<const uint4*>(p.state16 + grow * p.hw + hcol) + j)
                           : make_uint4(0u, 0u, 0u, 0u);
            uv[j] = row_ok ? __ldg(reinterpret_cast<const uint4*>(p.u16 + grow * p.hw + hcol) + j)
                           : make_uint4(0u, 0u, 0u, 0u);
          }
        }
      };
      if (mine && g_first < ngroups) fetch(g_first);
      tc::mbar_wait(bar(BAR_ACC_FULL + buf), (it >> 1) & 1);      // (also the team that skips the item: its arrival
      tc::fence_after_sync();                                     //  below must not run ahead of the accumulator)
      for (int g = mine ? g_first : ngroups; g < ngroups; g += g_step) {
        const int colbase = n0 + g * 64;                   // column in [0, Ntot)
        const bool is_r = colbase < p.hw;                  // GATES: reset-gate half
        const int hcol = is_r ? colbase : colbase - p.hw;  // column in [0, hw)
        if (g != g_first) fetch(g);
        if (leader) tc::tma_store_wait_read();             // the team's box is free again
        tc::named_bar_sync(1 + team, 128);
#pragma unroll
        for (int c16 = 0; c16 < 4; ++c16) {
          uint32_t acc[16];
          tc::tmem_ld16(t_lane + buf * MAX_NW + g * 64 + c16 * 16, acc);
          tc::tmem_ld_wait();
          const float* sb = s_bias + g * 64 + c16 * 16;
          float o[16];
          if (MODE == MODE_MSG) {
#pragma unroll
            for (int k = 0; k < 16; ++k) o[k] = __uint_as_float(acc[k]) + sb[k];
          } else if (MODE == MODE_GATES) {
            float hh[16];
            if (is_r) { unpack8(hv[2 * c16], hh); unpack8(hv[2 * c16 + 1], hh + 8); }
#pragma unroll
            for (int k = 0; k < 16; ++k) {
              const float gate = tc::rcp_approx(1.0f + tc::ex2_approx(-kLog2e * (__uint_as_float(acc[k]) + sb[k])));
              o[k] = is_r ? gate * hh[k] : gate;
              if (p.tape_g) acc[k] = __float_as_uint(gate);     // the tape keeps the gate itself
            }
            if (p.tape_g && row_ok) {
              float* dst = (is_r ? p.tape_g : p.tape_u) + grow * p.h + hcol + c16 * 16;
#pragma unroll
              for (int k4 = 0; k4 < 4; ++k4)
                if (hcol + c16 * 16 + k4 * 4 < p.h)
                  __stcs(reinterpret_cast<float4*>(dst + k4 * 4),          // streaming: the tape must not push the operands out of L2
                         make_float4(__uint_as_float(acc[k4 * 4]), __uint_as_float(acc[k4 * 4 + 1]),
                                     __uint_as_float(acc[k4 * 4 + 2]), __uint_as_float(acc[k4 * 4 + 3])));
            }
          } else {
            float hh[16], uu[16];
            unpack8(hv[2 * c16], hh); unpack8(hv[2 * c16 + 1], hh + 8);
            unpack8(uv[2 * c16], uu); unpack8(uv[2 * c16 + 1], uu + 8);
#pragma unroll
            for (int k = 0; k < 16; ++k) {
              const float cand = fmaf(-2.0f, tc::rcp_approx(1.0f + tc::ex2_approx(
                                                 2.0f * kLog2e * (__uint_as_float(acc[k]) + sb[k]))), 1.0f);
              o[k] = fmaf(uu[k], hh[k] - cand, cand);
              if (p.tape_g) acc[k] = __float_as_uint(cand);
            }
            if (p.tape_g && row_ok) {
              float* dst = p.tape_g + grow * p.h + hcol + c16 * 16;
#pragma unroll
              for (int k4 = 0; k4 < 4; ++k4)
                if (hcol + c16 * 16 + k4 * 4 < p.h)
                  __stcs(reinterpret_cast<float4*>(dst + k4 * 4),          // streaming: the tape must not push the operands out of L2
                         make_float4(__uint_as_float(acc[k4 * 4]), __uint_as_float(acc[k4 * 4 + 1]),
                                     __uint_as_float(acc[k4 * 4 + 2]), __uint_as_float(acc[k4 * 4 + 3])));
            }
            if (p.h_out && row_ok) {
              float* dst = p.h_out + grow * p.h + hcol + c16 * 16;
#pragma unroll
              for (int k4 = 0; k4 < 4; ++k4)
                if (hcol + c16 * 16 + k4 * 4 < p.h)
                  __stcs(reinterpret_cast<float4*>(dst + k4 * 4), make_float4(o[k4 * 4], o[k4 * 4 + 1], o[k4 * 4 + 2], o[k4 * 4 + 3]));
            }
          }
          *reinterpret_cast<uint4*>(box + sw_off(r, 2 * c16)) = pack8(o);
          *reinterpret_cast<uint4*>(box + sw_off(r, 2 * c16 + 1)) = pack8(o + 8);
        }
        tc::fence_proxy_async();
        tc::named_bar_sync(1 + team, 128);
        if (leader) {
          if (MODE == MODE_GATES && !is_r) tc::tma_store_2d(&tmO1, hcol, row0, box_addr);
          else tc::tma_store_2d(&tmO0, hcol, row0, box_addr);
          tc::tma_store_commit();
        }
      }
      tc::fence_before_sync();
      __syncwarp();
      if (lane == 0) tc::mbar_arrive(bar(BAR_ACC_EMPTY + buf));
    }
    if (leader) asm volatile("cp.async.bulk.wait_group 0;" ::: "memory");
  }

  tc::fence_before_sync();
  __syncthreads();
  if (warp == 0) tc::tmem_dealloc(tmem_base, 512);
}

// ---- packing ------------------------------------------------------------------------------
// edge weights: dst[e][n][k] (n, k < hw) = f16(src[e][k][n]) for k, n < h, else 0 -- packt::edge_t_kernel (pack_t.cuh)
// src [2h, ncols*h] -> dst [ncols*hw, 2hw]: dst[nb*hw+n][kb*hw+k] = f16(src[kb*h+k][nb*h+n])
__global__ void pack_gru_kernel(const float* __restrict__ src, __half* __restrict__ dst, int ncols, int h, int hw) {
  const int Kp = 2 * hw, Np = ncols * hw;
  const int64_t total = (int64_t)Np * Kp;
  for (int64_t i = (int64_t)blockIdx.x * blockDim.x + threadIdx.x; i < total; i += (int64_t)gridDim.x * blockDim.x) {
    const int kk = (int)(i % Kp), nn = (int)(i / Kp);
    const int kb = kk / hw, k = kk % hw, nb = nn / hw, n = nn % hw;
    dst[i] = __float2half_rn((k < h && n < h) ? src[(size_t)(kb * h + k) * (ncols * h) + nb * h + n] : 0.f);
  }
}
// dst[b][n] (n < hw) = src[b][n] (n < h) else 0; src null -> zeros
__global__ void pad_bias_kernel(const float* __restrict__ src, float* __restrict__ dst, int nblocks, int h, int hw) {
  const int i = blockIdx.x * blockDim.x + threadIdx.x;
  if (i >= nblocks * hw) return;
  const int nb = i / hw, n = i % hw;
  dst[i] = (src && n < h) ? src[(size_t)nb * h + n] : 0.f;
}
// state16[n][c] = f16(h[n][c]) (c < h) or 0; one thread per 4 columns; max|h| -> *absmax (float bits)
__global__ void state_in_kernel(const float* __restrict__ hs, __half* __restrict__ s16, int64_t N, int h, int hw,
                                unsigned* __restrict__ absmax) {
  const int q = hw / 4;
  const int64_t total = N * q;
  float m = 0.f;
  for (int64_t i = (int64_t)blockIdx.x * blockDim.x + threadIdx.x; i < total; i += (int64_t)gridDim.x * blockDim.x) {
    const int64_t n = i / q;
    const int c = (int)(i % q) * 4;
    float4 v = make_float4(0.f, 0.f, 0.f, 0.f);
    if (c < h) v = *reinterpret_cast<const float4*>(hs + n * h + c);
    m = fmaxf(fmaxf(m, fmaxf(fabsf(v.x), fabsf(v.y))), fmaxf(fabsf(v.z), fabsf(v.w)));
    __half2 lo = __floats2half2_rn(v.x, v.y), hi = __floats2half2_rn(v.z, v.w);
    uint2 pk;
    pk.x = *reinterpret_cast<uint32_t*>(&lo); pk.y = *reinterpret_cast<uint32_t*>(&hi);
    *reinterpret_cast<uint2*>(s16 + n * hw + c) = pk;
  }
#pragma unroll
  for (int o = 16; o > 0; o >>= 1) m = fmaxf(m, __shfl_xor_sync(0xffffffffu, m, o));
  if (absmax && (threadIdx.x & 31) == 0 && m > 0.f) atomicMax(absmax, __float_as_uint(m));
}

// ---- aggregation: x[n, :] = sum of the FP16 message rows of node n (fixed order), FP32 sum -> FP16 ----
__global__ void __launch_bounds__(256)
aggregate_kernel(const int32_t* __restrict__ tptr, const int32_t* __restrict__ tmsg, const __half* __restrict__ msg,
                 __half* __restrict__ acts, int64_t N, int hw, float* __restrict__ acts32, int h) {
  const int lane = threadIdx.x & 31;
  const int64_t node = ((int64_t)blockIdx.x * blockDim.x + threadIdx.x) >> 5;
  if (node >= N) return;
  const int j0 = tptr[node], j1 = tptr[node + 1];
  const int h8 = hw >> 3;
  for (int c = lane; c < h8; c += 32) {
    float acc[8];
#pragma unroll
    for (int k = 0; k < 8; ++k) acc[k] = 0.f;
    int j = j0;
    for (; j + 4 <= j1; j += 4) {
      const int r0 = __ldg(tmsg + j), r1 = __ldg(tmsg + j + 1), r2 = __ldg(tmsg + j + 2), r3 = __ldg(tmsg + j + 3);
      const uint4 m0 = __ldg(reinterpret_cast<const uint4*>(msg + (size_t)r0 * hw) + c);
      const uint4 m1 = __ldg(reinterpret_cast<const uint4*>(msg + (size_t)r1 * hw) + c);
      const uint4 m2 = __ldg(reinterpret_cast<const uint4*>(msg + (size_t)r2 * hw) + c);
      const uint4 m3 = __ldg(reinterpret_cast<const uint4*>(msg + (size_t)r3 * hw) + c);
      float f[8];
      unpack8(m0, f);
#pragma unroll
      for (int k = 0; k < 8; ++k) acc[k] += f[k];
      unpack8(m1, f);
#pragma unroll
      for (int k = 0; k < 8; ++k) acc[k] += f[k];
      unpack8(m2, f);
#pragma unroll
      for (int k = 0; k < 8; ++k) acc[k] += f[k];
      unpack8(m3, f);
#pragma unroll
      for (int k = 0; k < 8; ++k) acc[k] += f[k];
    }
    for (; j < j1; ++j) {
      const int r0 = __ldg(tmsg + j);
      float f[8];
      unpack8(__ldg(reinterpret_cast<const uint4*>(msg + (size_t)r0 * hw) + c), f);
#pragma unroll
      for (int k = 0; k < 8; ++k) acc[k] += f[k];
    }
    const uint4 pk = pack8(acc);
    reinterpret_cast<uint4*>(acts + (size_t)node * hw)[c] = pk;
    if (acts32 && c * 8 < h) {                 // training tape: the FP16-rounded operand, widened (h % 4 == 0)
      float f[8];
      unpack8(pk, f);
      float* dst = acts32 + (size_t)node * h + c * 8;
      *reinterpret_cast<float4*>(dst) = make_float4(f[0], f[1], f[2], f[3]);
      if (c * 8 + 4 < h) *reinterpret_cast<float4*>(dst + 4) = make_float4(f[4], f[5], f[6], f[7]);
    }
  }
}

// dst [N, h] FP32 = state16 [N, hw] (training tape: the state a step started from)
__global__ void state_out_kernel(const __half* __restrict__ s16, float* __restrict__ dst, int64_t N, int h, int hw) {
  const int q = h / 4;
  const int64_t total = N * q;
  for (int64_t i = (int64_t)blockIdx.x * blockDim.x + threadIdx.x; i < total; i += (int64_t)gridDim.x * blockDim.x) {
    const int64_t n = i / q;
    const int c = (int)(i % q) * 4;
    const uint2 v = *reinterpret_cast<const uint2*>(s16 + n * hw + c);
    const float2 lo = __half22float2(*reinterpret_cast<const __half2*>(&v.x));
    const float2 hi = __half22float2(*reinterpret_cast<const __half2*>(&v.y));
    *reinterpret_cast<float4*>(dst + n * h + c) = make_float4(lo.x, lo.y, hi.x, hi.y);
  }
}
// state16 = f16(state16 * mask * scale): DropoutWrapper(state_keep_prob) on the new state
__global__ void state_mask_kernel(__half* __restrict__ s16, const float* __restrict__ mask, float scale, int64_t N, int h,
                                  int hw) {
  const int q = h / 4;
  const int64_t total = N * q;
  for (int64_t i = (int64_t)blockIdx.x * blockDim.x + threadIdx.x; i < total; i += (int64_t)gridDim.x * blockDim.x) {
    const int64_t n = i / q;
    const int c = (int)(i % q) * 4;
    uint2 v = *reinterpret_cast<uint2*>(s16 + n * hw + c);
    const float4 m = *reinterpret_cast<const float4*>(mask + n * h + c);
    const float2 lo = __half22float2(*reinterpret_cast<const __half2*>(&v.x));
    const float2 hi = __half22float2(*reinterpret_cast<const __half2*>(&v.y));
    __half2 a = __floats2half2_rn(lo.x * m.x * scale, lo.y * m.y * scale);
    __half2 b = __floats2half2_rn(hi.x * m.z * scale, hi.y * m.w * scale);
    v.x = *reinterpret_cast<uint32_t*>(&a); v.y = *reinterpret_cast<uint32_t*>(&b);
    *reinterpret_cast<uint2*>(s16 + n * hw + c) = v;
  }
}

// slice width: the widest multiple of 64 (<= 256) whose [NW x K] FP16 panel fits the resident 128 KB
int slice_width(int K, int Ntot) {
  int nw = (W_BYTES / (K * 2)) / 64 * 64;
  if (nw > MAX_NW) nw = MAX_NW;
  if (nw > Ntot) nw = Ntot;
  return nw;
}

// row blocks of item_iter: blocks whose activations ([rows x K] FP16) stay below ~32 MB when they are swept more than once
int row_blocks(int64_t n_row_tiles, int K, int S) {
  static const int64_t block_bytes = getenv("GGNN_W16_BLOCK_MB") ? (int64_t)atoi(getenv("GGNN_W16_BLOCK_MB")) << 20 : (int64_t)32 << 20;
  if (S <= 1 || block_bytes <= 0) return 1;
  const int64_t bytes = n_row_tiles * BM * (int64_t)K * 2;
  int64_t nb = (bytes + block_bytes - 1) / block_bytes;
  if (nb < 1) nb = 1;
  if (nb > n_row_tiles) nb = n_row_tiles;
  return (int)nb;
}

template <int MODE>
int launch(ggnn_ctx* ctx, const params& p, const CUtensorMap& a1, const CUtensorMap& a2, const CUtensorMap& w,
           const CUtensorMap& o0, const CUtensorMap& o1, int64_t max_items, cudaStream_t st) {
  GGNN_ENSURE_SMEM(ctx, GGNN_ATTR_W16_0 + MODE, wgemm16_kernel<MODE>, SMEM_BYTES);
  const int grid = (int)(max_items < ctx->sm_count ? max_items : ctx->sm_count);
  if (grid <= 0) return GGNN_OK;
  wgemm16_kernel<MODE><<<grid, NTHREADS, SMEM_BYTES, st>>>(p, a1, a2, w, o0, o1);
  GGNN_LAUNCHED(ctx);
  return GGNN_OK;
}

}  // namespace w16

// =========================================================================================
// GRU cell of a 128-node tile in ONE persistent kernel, hw <= 256 (the wide twin of f16.cu's gru_f16_kernel, without
// the aggregation): x and h tiles resident in shared memory, the 6 hw^2 FP16 weights streamed from L2 through a
// 3-slot ring of [hw n x 64 k] chunks, gates in TMEM.  r, u, r*h and the candidate never touch HBM:
// per timestep the unfused GATES + CAND kernels move 9 node-state-sized arrays through HBM, this one 3
// (x and h in, h' out).
//   MMA order   u_x, u_h, r_h | r_x | cand_x, cand_rh          (u: TMEM columns 256.., r and cand: columns 0..)
//   EPI-u       after u_h / r_h retire: u = sigmoid(.), packed with the row's h as (f16 u, f16 h) into u's own TMEM
//               words -- h survives there while the h tile becomes r*h
//   EPI-r       after r_x: r = sigmoid(.), f16(r*h) overwrites the h tile (the K operand of cand_rh)
//   EPI-c       c = tanh(.), h' = c + u (h - c) -> FP16 into the (dead) x tile -> TMA store; FP32 result on the last step
namespace gw16 {
using namespace w16;

constexpr int WSLOT = 32 * 1024;                 // one [256 n x 64 k] FP16 chunk
constexpr int WSLOTS = 3;
constexpr int TILE_BYTES = 4 * CHUNK;            // 64 KB: [128 x 256] FP16 as four swizzled 64-column chunks
constexpr int WARP_TMA = 8, WARP_MMA = 9;        // warps 0-7: epilogue (lane quarter = warp % 4, column half = warp / 4)
constexpr int NTHREADS = 10 * 32;
constexpr int SMEM_BYTES = 2 * TILE_BYTES + WSLOTS * WSLOT + 1024 /*align*/ + 256 /*barriers*/;
static_assert(SMEM_BYTES <= 232448, "shared memory budget");

enum {
  BAR_W_FULL = 0,                     // [WSLOTS]
  BAR_W_EMPTY = BAR_W_FULL + WSLOTS,  // [WSLOTS]
  BAR_X_FULL = BAR_W_EMPTY + WSLOTS,
  BAR_H_FULL,
  BAR_X_EMPTY,                        // h' of the previous tile has left the x tile (leader, after the store has read it)
  BAR_U_DONE,                         // u complete, every MMA reading the h tile retired
  BAR_R_DONE,
  BAR_RH_FULL,                        // 8 epilogue warps: r drained from TMEM, r*h in the h tile
  BAR_C_DONE,                         // candidate complete (h tile free)
  BAR_TMEM_FREE,                      // 8 epilogue warps: c and (u, h) read
  BAR_COUNT
};
static_assert(8 * BAR_COUNT + 8 <= 256, "barrier area");

struct params {
  const float* bg;          // [2 hw] zero-padded
  const float* bc;          // [hw]
  float* h_out;             // [N, h] FP32 result (last step) or null
  int64_t N;
  int h, hw, n_tiles;
};

__global__ void __launch_bounds__(NTHREADS, 1)
gru_w16_kernel(const params p, const __grid_constant__ CUtensorMap tmX, const __grid_constant__ CUtensorMap tmH,
               const __grid_constant__ CUtensorMap tmWg, const __grid_constant__ CUtensorMap tmWc) {
  extern __shared__ uint8_t smem_raw[];
  const int tid = threadIdx.x, warp = tid >> 5, lane = tid & 31;
  const uint32_t base = (tc::smem_u32(smem_raw) + 1023u) & ~1023u;
  uint8_t* smem_al = smem_raw + (base - tc::smem_u32(smem_raw));
  const uint32_t sX = base, sH = base + TILE_BYTES, sW = sH + TILE_BYTES;
  const uint32_t bars = sW + WSLOTS * WSLOT;
  auto bar = [&](int i) { return bars + 8u * i; };
  const uint32_t tmem_slot = bars + 8u * BAR_COUNT;
  const int hw = p.hw, KC1 = hw / BKH;
  const uint32_t wbytes = (uint32_t)hw * 128u;             // bytes one weight chunk delivers

  if (warp == WARP_MMA && lane == 0) {
    for (int s = 0; s < WSLOTS; ++s) { tc::mbar_init(bar(BAR_W_FULL + s), 1); tc::mbar_init(bar(BAR_W_EMPTY + s), 1); }
    tc::mbar_init(bar(BAR_X_FULL), 1);
    tc::mbar_init(bar(BAR_H_FULL), 1);
    tc::mbar_init(bar(BAR_X_EMPTY), 1);
    tc::mbar_init(bar(BAR_U_DONE), 1);
    tc::mbar_init(bar(BAR_R_DONE), 1);
    tc::mbar_init(bar(BAR_RH_FULL), 8);
    tc::mbar_init(bar(BAR_C_DONE), 1);
    tc::mbar_init(bar(BAR_TMEM_FREE), 8);
    tc::fence_barrier_init();
  }
  if (warp == WARP_TMA) {
    if (lane == 0) {
      tc::tma_prefetch_desc(&tmX); tc::tma_prefetch_desc(&tmH); tc::tma_prefetch_desc(&tmWg); tc::tma_prefetch_desc(&tmWc);
    }
    __syncwarp();
    tc::tmem_alloc(tmem_slot, 512);
  }
  tc::fence_before_sync();
  __syncthreads();
  tc::fence_after_sync();
  const uint32_t tmem_base = *reinterpret_cast<volatile uint32_t*>(smem_al + (tmem_slot - base));

  if (warp == WARP_TMA) {
    // =============================== TMA producer ===============================
    if (lane == 0) {
      uint32_t fill = 0;
      auto w_fill = [&](const CUtensorMap* tm, int kcol, int nrow) {
        const int s = fill % WSLOTS;
        if (fill >= WSLOTS) tc::mbar_wait(bar(BAR_W_EMPTY + s), ((fill / WSLOTS) - 1) & 1);
        tc::mbar_arrive_expect_tx(bar(BAR_W_FULL + s), wbytes);
        tc::tma_load_2d(sW + s * WSLOT, tm, kcol, nrow, bar(BAR_W_FULL + s));
        ++fill;
      };
      int it = 0;
      for (int tile = blockIdx.x; tile < p.n_tiles; tile += gridDim.x, ++it) {
        // the first ring-full of weights does not depend on the tiles being free; the h tile is free as soon as the
        // previous candidate MMAs retire, the x tile (h' staging) only after the previous store has read it
        int issued = 0;
        auto tiles_in = [&]() {
          if (it > 0) tc::mbar_wait(bar(BAR_C_DONE), (it - 1) & 1);
          tc::mbar_arrive_expect_tx(bar(BAR_H_FULL), (uint32_t)KC1 * CHUNK);
          for (int kc = 0; kc < KC1; ++kc) tc::tma_load_2d(sH + kc * CHUNK, &tmH, kc * BKH, tile * BM, bar(BAR_H_FULL));
          if (it > 0) tc::mbar_wait(bar(BAR_X_EMPTY), (it - 1) & 1);
          tc::mbar_arrive_expect_tx(bar(BAR_X_FULL), (uint32_t)KC1 * CHUNK);
          for (int kc = 0; kc < KC1; ++kc) tc::tma_load_2d(sX + kc * CHUNK, &tmX, kc * BKH, tile * BM, bar(BAR_X_FULL));
        };
        auto step = [&](const CUtensorMap* tm, int kcol, int nrow) {
          w_fill(tm, kcol, nrow);
          if (++issued == WSLOTS) tiles_in();
        };
        for (int kc = 0; kc < KC1; ++kc) step(&tmWg, hw + kc * BKH, hw);          // h   -> u
        for (int kc = 0; kc < KC1; ++kc) step(&tmWg, hw + kc * BKH, 0);           // h   -> r
        for (int kc = 0; kc < KC1; ++kc) step(&tmWg, kc * BKH, hw);               // x   -> u
        for (int kc = 0; kc < KC1; ++kc) step(&tmWg, kc * BKH, 0);                // x   -> r
        for (int kc = 0; kc < KC1; ++kc) step(&tmWc, kc * BKH, 0);                // x   -> cand
        for (int kc = 0; kc < KC1; ++kc) step(&tmWc, hw + kc * BKH, 0);           // r*h -> cand
      }
    }
    __syncwarp();
  } else if (warp == WARP_MMA) {
    // =============================== MMA issuer ===============================
    if (lane == 0) {
      const uint32_t idesc = tc::umma_idesc(0 /*F16*/, BM, hw);
      const uint32_t d_rc = tmem_base, d_u = tmem_base + 256;
      uint32_t use = 0;
      auto chunk_mma = [&](uint32_t d, uint32_t a, bool first) {
        const int s = use % WSLOTS;
        tc::mbar_wait(bar(BAR_W_FULL + s), (use / WSLOTS) & 1);
        tc::fence_after_sync();
        const uint32_t w = sW + s * WSLOT;
#pragma unroll
        for (int k = 0; k < 4; ++k)
          tc::umma_f16(d, tc::umma_desc_sw128(a + k * 32), tc::umma_desc_sw128(w + k * 32), idesc, (first && k == 0) ? 0u : 1u);
        tc::umma_commit(bar(BAR_W_EMPTY + s));
        ++use;
      };
      int it = 0;
      for (int tile = blockIdx.x; tile < p.n_tiles; tile += gridDim.x, ++it) {
        if (it > 0) tc::mbar_wait(bar(BAR_TMEM_FREE), (it - 1) & 1);
        tc::mbar_wait(bar(BAR_H_FULL), it & 1);
        tc::fence_after_sync();
        for (int kc = 0; kc < KC1; ++kc) chunk_mma(d_u, sH + kc * CHUNK, kc == 0);      // h -> u
        for (int kc = 0; kc < KC1; ++kc) chunk_mma(d_rc, sH + kc * CHUNK, kc == 0);     // h -> r
        tc::mbar_wait(bar(BAR_X_FULL), it & 1);                                         // x arrives under the h MMAs
        tc::fence_after_sync();
        for (int kc = 0; kc < KC1; ++kc) chunk_mma(d_u, sX + kc * CHUNK, false);        // x -> u
        tc::umma_commit(bar(BAR_U_DONE));
        for (int kc = 0; kc < KC1; ++kc) chunk_mma(d_rc, sX + kc * CHUNK, false);       // x -> r
        tc::umma_commit(bar(BAR_R_DONE));
        tc::mbar_wait(bar(BAR_RH_FULL), it & 1);                                        // r drained, r*h in the h tile
        tc::fence_after_sync();
        for (int kc = 0; kc < KC1; ++kc) chunk_mma(d_rc, sX + kc * CHUNK, kc == 0);     // x -> cand
        for (int kc = 0; kc < KC1; ++kc) chunk_mma(d_rc, sH + kc * CHUNK, false);       // r*h -> cand
        tc::umma_commit(bar(BAR_C_DONE));
      }
    }
    __syncwarp();
  } else {
    // =============================== epilogue (warps 0-7) ===============================
    const int q = warp & 3, half = warp >> 2;
    const int r = q * 32 + lane;
    const uint32_t t_lane = tmem_base + ((uint32_t)(q * 32) << 16);
    const int ng = hw / 16;                                   // 16-column groups
    const int g_begin = half * (ng / 2), g_end = half ? ng : ng / 2;
    uint8_t* xt = smem_al + (sX - base);
    uint8_t* ht = smem_al + (sH - base);
    const bool leader = tid == 0;
    int it = 0;
    for (int tile = blockIdx.x; tile < p.n_tiles; tile += gridDim.x, ++it) {
      const int64_t grow = (int64_t)tile * BM + r;
      // ---- EPI-u
      tc::mbar_wait(bar(BAR_U_DONE), it & 1);
      tc::fence_after_sync();
#pragma unroll 1
      for (int g = g_begin; g < g_end; ++g) {
        const int col0 = g * 16;
        uint32_t au[16];
        tc::tmem_ld16(t_lane + 256 + col0, au);
        const uint8_t* hp = ht + (col0 >> 6) * CHUNK;
        const int s0 = (col0 & 63) >> 3;
        const uint4 h0 = *reinterpret_cast<const uint4*>(hp + sw_off(r, s0));
        const uint4 h1 = *reinterpret_cast<const uint4*>(hp + sw_off(r, s0 + 1));
        const uint32_t hb[8] = {h0.x, h0.y, h0.z, h0.w, h1.x, h1.y, h1.z, h1.w};
        float sb[16];
#pragma unroll
        for (int k4 = 0; k4 < 4; ++k4) {
          const float4 b4 = __ldg(reinterpret_cast<const float4*>(p.bg + hw + col0) + k4);
          sb[4 * k4] = b4.x; sb[4 * k4 + 1] = b4.y; sb[4 * k4 + 2] = b4.z; sb[4 * k4 + 3] = b4.w;
        }
        tc::tmem_ld_wait();
#pragma unroll
        for (int k = 0; k < 16; ++k) {
          const float u = tc::rcp_approx(1.0f + tc::ex2_approx(-kLog2e * (__uint_as_float(au[k]) + sb[k])));
          const uint32_t hbits = (k & 1) ? (hb[k >> 1] >> 16) : (hb[k >> 1] & 0xffffu);
          au[k] = (uint32_t)__half_as_ushort(__float2half_rn(u)) | (hbits << 16);     // (f16 u, f16 h)
        }
        tc::tmem_st16(t_lane + 256 + col0, au);
      }
      tc::tmem_st_wait();
      // ---- EPI-r: f16(r*h) replaces h in shared memory
      tc::mbar_wait(bar(BAR_R_DONE), it & 1);
      tc::fence_after_sync();
#pragma unroll 1
      for (int g = g_begin; g < g_end; ++g) {
        const int col0 = g * 16;
        uint32_t ar[16];
        tc::tmem_ld16(t_lane + col0, ar);
        uint8_t* hp = ht + (col0 >> 6) * CHUNK;
        const int s0 = (col0 & 63) >> 3;
        uint4* p0 = reinterpret_cast<uint4*>(hp + sw_off(r, s0));
        uint4* p1 = reinterpret_cast<uint4*>(hp + sw_off(r, s0 + 1));
        float hh[16];
        unpack8(*p0, hh); unpack8(*p1, hh + 8);
        float sb[16];
#pragma unroll
        for (int k4 = 0; k4 < 4; ++k4) {
          const float4 b4 = __ldg(reinterpret_cast<const float4*>(p.bg + col0) + k4);
          sb[4 * k4] = b4.x; sb[4 * k4 + 1] = b4.y; sb[4 * k4 + 2] = b4.z; sb[4 * k4 + 3] = b4.w;
        }
        tc::tmem_ld_wait();
        float o[16];
#pragma unroll
        for (int k = 0; k < 16; ++k)
          o[k] = tc::rcp_approx(1.0f + tc::ex2_approx(-kLog2e * (__uint_as_float(ar[k]) + sb[k]))) * hh[k];
        *p0 = pack8(o);
        *p1 = pack8(o + 8);
      }
      tc::fence_before_sync();
      tc::fence_proxy_async();
      __syncwarp();
      if (lane == 0) tc::mbar_arrive(bar(BAR_RH_FULL));
      // ---- EPI-c: candidate and the new state
      tc::mbar_wait(bar(BAR_C_DONE), it & 1);
      tc::fence_after_sync();
#pragma unroll 1
      for (int g = g_begin; g < g_end; ++g) {
        const int col0 = g * 16;
        uint32_t ac[16], uh[16];
        tc::tmem_ld16(t_lane + col0, ac);
        tc::tmem_ld16(t_lane + 256 + col0, uh);
        float sb[16];
#pragma unroll
        for (int k4 = 0; k4 < 4; ++k4) {
          const float4 b4 = __ldg(reinterpret_cast<const float4*>(p.bc + col0) + k4);
          sb[4 * k4] = b4.x; sb[4 * k4 + 1] = b4.y; sb[4 * k4 + 2] = b4.z; sb[4 * k4 + 3] = b4.w;
        }
        tc::tmem_ld_wait();
        float o[16];
#pragma unroll
        for (int k = 0; k < 16; ++k) {
          const float c = fmaf(-2.0f, tc::rcp_approx(1.0f + tc::ex2_approx(2.0f * kLog2e * (__uint_as_float(ac[k]) + sb[k]))), 1.0f);
          const float u = __half2float(__ushort_as_half((unsigned short)(uh[k] & 0xffffu)));
          const float hv = __half2float(__ushort_as_half((unsigned short)(uh[k] >> 16)));
          o[k] = fmaf(u, hv - c, c);
        }
        uint8_t* xp = xt + (col0 >> 6) * CHUNK;
        const int s0 = (col0 & 63) >> 3;
        *reinterpret_cast<uint4*>(xp + sw_off(r, s0)) = pack8(o);
        *reinterpret_cast<uint4*>(xp + sw_off(r, s0 + 1)) = pack8(o + 8);
        if (p.h_out && grow < p.N) {
          float* dst = p.h_out + grow * p.h + col0;
#pragma unroll
          for (int k4 = 0; k4 < 4; ++k4)
            if (col0 + k4 * 4 < p.h)
              *reinterpret_cast<float4*>(dst + k4 * 4) = make_float4(o[k4 * 4], o[k4 * 4 + 1], o[k4 * 4 + 2], o[k4 * 4 + 3]);
        }
      }
      tc::fence_before_sync();
      tc::fence_proxy_async();
      __syncwarp();
      if (lane == 0) tc::mbar_arrive(bar(BAR_TMEM_FREE));
      tc::named_bar_sync(1, 256);
      if (leader) {
        for (int kc = 0; kc < KC1; ++kc) tc::tma_store_2d(&tmH, kc * BKH, tile * BM, sX + kc * CHUNK);   // rows >= N clipped
        tc::tma_store_commit();
        tc::tma_store_wait_read();
        tc::mbar_arrive(bar(BAR_X_EMPTY));
      }
    }
    if (leader) asm volatile("cp.async.bulk.wait_group 0;" ::: "memory");
  }

  tc::fence_before_sync();
  __syncthreads();
  if (warp == WARP_TMA) tc::tmem_dealloc(tmem_base, 512);
}

}  // namespace gw16

// =========================================================================================
// The same GRU cell on CTA PAIRS (tcgen05 cta_group::2): the two CTAs of a cluster take two adjacent 128-node tiles as one
// 256-row MMA tile.  Each CTA keeps its own x / h tiles and epilogues, but only HALF of every weight chunk ([hw/2 n x 64 k]):
// the weight bytes an SM pulls from L2 per tile halve (768 -> 384 KB at hw = 256, the measured limiter of gru_w16_kernel)
// and the ring holds six chunks instead of three.  The even CTA issues every MMA; TMA completions of both CTAs are counted on
// its barriers, MMA completions are multicast to both.
namespace gw16x2 {
using namespace w16;

constexpr int WSLOT = 16 * 1024;                 // one [128 n x 64 k] FP16 half chunk
constexpr int WSLOTS = 6;
constexpr int TILE_BYTES = 4 * CHUNK;
constexpr int WARP_TMA = 8, WARP_MMA = 9;
constexpr int NTHREADS = 10 * 32;
constexpr int SMEM_BYTES = 2 * TILE_BYTES + WSLOTS * WSLOT + 1024 + 256;

enum {
  BAR_W_FULL = 0,                     // [WSLOTS] even CTA: bytes of both halves
  BAR_W_EMPTY = BAR_W_FULL + WSLOTS,  // [WSLOTS] multicast commit
  BAR_X_FULL = BAR_W_EMPTY + WSLOTS,  // even CTA: both x tiles
  BAR_H_FULL,                         // even CTA: both h tiles
  BAR_X_EMPTY,                        // local
  BAR_U_DONE, BAR_R_DONE, BAR_C_DONE, // multicast commits
  BAR_RH_FULL,                        // even CTA: 16 epilogue warps
  BAR_TMEM_FREE,                      // even CTA: 16 epilogue warps
  BAR_COUNT
};
static_assert(8 * BAR_COUNT + 8 <= 256, "barrier area");

__global__ void __cluster_dims__(2, 1, 1) __launch_bounds__(NTHREADS, 1)
gru_w16x2_kernel(const gw16::params p, const __grid_constant__ CUtensorMap tmX, const __grid_constant__ CUtensorMap tmH,
                 const __grid_constant__ CUtensorMap tmWg, const __grid_constant__ CUtensorMap tmWc) {
  extern __shared__ uint8_t smem_raw[];
  const int tid = threadIdx.x, warp = tid >> 5, lane = tid & 31;
  const uint32_t base = (tc::smem_u32(smem_raw) + 1023u) & ~1023u;
  uint8_t* smem_al = smem_raw + (base - tc::smem_u32(smem_raw));
  const uint32_t sX = base, sH = base + TILE_BYTES, sW = sH + TILE_BYTES;
  const uint32_t bars = sW + WSLOTS * WSLOT;
  auto bar = [&](int i) { return bars + 8u * i; };
  const uint32_t tmem_slot = bars + 8u * BAR_COUNT;
  const int hw = p.hw, KC1 = hw / BKH, hh = hw / 2;
  const uint32_t rank = tc::cluster_ctarank();
  const bool lead = rank == 0;
  const int n_pairs = (p.n_tiles + 1) / 2;
  const int pair0 = blockIdx.x >> 1, pair_stride = gridDim.x >> 1;
  const uint32_t wbytes = (uint32_t)hh * 128u;              // bytes one CTA's half chunk delivers

  if (warp == WARP_MMA && lane == 0) {
    for (int s = 0; s < WSLOTS; ++s) { tc::mbar_init(bar(BAR_W_FULL + s), 1); tc::mbar_init(bar(BAR_W_EMPTY + s), 1); }
    tc::mbar_init(bar(BAR_X_FULL), 1);
    tc::mbar_init(bar(BAR_H_FULL), 1);
    tc::mbar_init(bar(BAR_X_EMPTY), 1);
    tc::mbar_init(bar(BAR_U_DONE), 1);
    tc::mbar_init(bar(BAR_R_DONE), 1);
    tc::mbar_init(bar(BAR_C_DONE), 1);
    tc::mbar_init(bar(BAR_RH_FULL), 16);
    tc::mbar_init(bar(BAR_TMEM_FREE), 16);
    tc::fence_barrier_init();
  }
  if (warp == WARP_TMA) {
    if (lane == 0) {
      tc::tma_prefetch_desc(&tmX); tc::tma_prefetch_desc(&tmH); tc::tma_prefetch_desc(&tmWg); tc::tma_prefetch_desc(&tmWc);
    }
    __syncwarp();
    tc::tmem_alloc_pair(tmem_slot, 512);
  }
  tc::fence_before_sync();
  tc::cluster_sync_all();                                   // barriers of both CTAs initialised, TMEM allocated
  tc::fence_after_sync();
  const uint32_t tmem_base = *reinterpret_cast<volatile uint32_t*>(smem_al + (tmem_slot - base));

  if (warp == WARP_TMA) {
    // =============================== TMA producer (both CTAs) ===============================
    if (lane == 0) {
      uint32_t fill = 0;
      auto w_fill = [&](const CUtensorMap* tm, int kcol, int nrow) {
        const int s = fill % WSLOTS;
        if (fill >= WSLOTS) tc::mbar_wait(bar(BAR_W_EMPTY + s), ((fill / WSLOTS) - 1) & 1);
        if (lead) tc::mbar_arrive_expect_tx(bar(BAR_W_FULL + s), 2u * wbytes);
        tc::tma_load_2d_pair(sW + s * WSLOT, tm, kcol, nrow + (int)rank * hh, bar(BAR_W_FULL + s));
        ++fill;
      };
      int it = 0;
      for (int pr = pair0; pr < n_pairs; pr += pair_stride, ++it) {
        const int tile = 2 * pr + (int)rank;
        int issued = 0;
        auto tiles_in = [&]() {
          if (it > 0) tc::mbar_wait(bar(BAR_C_DONE), (it - 1) & 1);
          if (lead) tc::mbar_arrive_expect_tx(bar(BAR_H_FULL), 2u * (uint32_t)KC1 * CHUNK);
          for (int kc = 0; kc < KC1; ++kc) tc::tma_load_2d_pair(sH + kc * CHUNK, &tmH, kc * BKH, tile * BM, bar(BAR_H_FULL));
          if (it > 0) tc::mbar_wait(bar(BAR_X_EMPTY), (it - 1) & 1);
          if (lead) tc::mbar_arrive_expect_tx(bar(BAR_X_FULL), 2u * (uint32_t)KC1 * CHUNK);
          for (int kc = 0; kc < KC1; ++kc) tc::tma_load_2d_pair(sX + kc * CHUNK, &tmX, kc * BKH, tile * BM, bar(BAR_X_FULL));
        };
        auto step = [&](const CUtensorMap* tm, int kcol, int nrow) {
          w_fill(tm, kcol, nrow);
          if (++issued == WSLOTS) tiles_in();
        };
        for (int kc = 0; kc < KC1; ++kc) step(&tmWg, hw + kc * BKH, hw);          // h   -> u
        for (int kc = 0; kc < KC1; ++kc) step(&tmWg, hw + kc * BKH, 0);           // h   -> r
        for (int kc = 0; kc < KC1; ++kc) step(&tmWg, kc * BKH, hw);               // x   -> u
        for (int kc = 0; kc < KC1; ++kc) step(&tmWg, kc * BKH, 0);                // x   -> r
        for (int kc = 0; kc < KC1; ++kc) step(&tmWc, kc * BKH, 0);                // x   -> cand
        for (int kc = 0; kc < KC1; ++kc) step(&tmWc, hw + kc * BKH, 0);           // r*h -> cand
      }
    }
    __syncwarp();
  } else if (warp == WARP_MMA) {
    // =============================== MMA issuer (even CTA only) ===============================
    if (lead && lane == 0) {
      const uint32_t idesc = tc::umma_idesc(0 /*F16*/, 2 * BM, hw);
      const uint32_t d_rc = tmem_base, d_u = tmem_base + 256;
      uint32_t use = 0;
      auto chunk_mma = [&](uint32_t d, uint32_t a, bool first) {
        const int s = use % WSLOTS;
        tc::mbar_wait_cluster(bar(BAR_W_FULL + s), (use / WSLOTS) & 1);
        tc::fence_after_sync();
        const uint32_t w = sW + s * WSLOT;
#pragma unroll
        for (int k = 0; k < 4; ++k)
          tc::umma_f16_pair(d, tc::umma_desc_sw128(a + k * 32), tc::umma_desc_sw128(w + k * 32), idesc, (first && k == 0) ? 0u : 1u);
        tc::umma_commit_pair(bar(BAR_W_EMPTY + s));
        ++use;
      };
      int it = 0;
      for (int pr = pair0; pr < n_pairs; pr += pair_stride, ++it) {
        if (it > 0) tc::mbar_wait_cluster(bar(BAR_TMEM_FREE), (it - 1) & 1);
        tc::mbar_wait_cluster(bar(BAR_H_FULL), it & 1);
        tc::fence_after_sync();
        for (int kc = 0; kc < KC1; ++kc) chunk_mma(d_u, sH + kc * CHUNK, kc == 0);      // h -> u
        for (int kc = 0; kc < KC1; ++kc) chunk_mma(d_rc, sH + kc * CHUNK, kc == 0);     // h -> r
        tc::mbar_wait_cluster(bar(BAR_X_FULL), it & 1);
        tc::fence_after_sync();
        for (int kc = 0; kc < KC1; ++kc) chunk_mma(d_u, sX + kc * CHUNK, false);        // x -> u
        tc::umma_commit_pair(bar(BAR_U_DONE));
        for (int kc = 0; kc < KC1; ++kc) chunk_mma(d_rc, sX + kc * CHUNK, false);       // x -> r
        tc::umma_commit_pair(bar(BAR_R_DONE));
        tc::mbar_wait_cluster(bar(BAR_RH_FULL), it & 1);                                // both CTAs: r drained, r*h in place
        tc::fence_after_sync();
        for (int kc = 0; kc < KC1; ++kc) chunk_mma(d_rc, sX + kc * CHUNK, kc == 0);     // x -> cand
        for (int kc = 0; kc < KC1; ++kc) chunk_mma(d_rc, sH + kc * CHUNK, false);       // r*h -> cand
        tc::umma_commit_pair(bar(BAR_C_DONE));
      }
    }
    __syncwarp();
  } else {
    // =============================== epilogue (warps 0-7, both CTAs; gru_w16_kernel's) ===============================
    const int q = warp & 3, half = warp >> 2;
    const int r = q * 32 + lane;
    const uint32_t t_lane = tmem_base + ((uint32_t)(q * 32) << 16);
    const int ng = hw / 16;
    const int g_begin = half * (ng / 2), g_end = half ? ng : ng / 2;
    uint8_t* xt = smem_al + (sX - base);
    uint8_t* ht = smem_al + (sH - base);
    const bool leader = tid == 0;
    int it = 0;
    for (int pr = pair0; pr < n_pairs; pr += pair_stride, ++it) {
      const int tile = 2 * pr + (int)rank;
      const int64_t grow = (int64_t)tile * BM + r;
      // ---- EPI-u
      tc::mbar_wait(bar(BAR_U_DONE), it & 1);
      tc::fence_after_sync();
#pragma unroll 1
      for (int g = g_begin; g < g_end; ++g) {
        const int col0 = g * 16;
        uint32_t au[16];
        tc::tmem_ld16(t_lane + 256 + col0, au);
        const uint8_t* hp = ht + (col0 >> 6) * CHUNK;
        const int s0 = (col0 & 63) >> 3;
        const uint4 h0 = *reinterpret_cast<const uint4*>(hp + sw_off(r, s0));
        const uint4 h1 = *reinterpret_cast<const uint4*>(hp + sw_off(r, s0 + 1));
        const uint32_t hb[8] = {h0.x, h0.y, h0.z, h0.w, h1.x, h1.y, h1.z, h1.w};
        float sb[16];
#pragma unroll
        for (int k4 = 0; k4 < 4; ++k4) {
          const float4 b4 = __ldg(reinterpret_cast<const float4*>(p.bg + hw + col0) + k4);
          sb[4 * k4] = b4.x; sb[4 * k4 + 1] = b4.y; sb[4 * k4 + 2] = b4.z; sb[4 * k4 + 3] = b4.w;
        }
        tc::tmem_ld_wait();
#pragma unroll
        for (int k = 0; k < 16; ++k) {
          const float u = tc::rcp_approx(1.0f + tc::ex2_approx(-kLog2e * (__uint_as_float(au[k]) + sb[k])));
          const uint32_t hbits = (k & 1) ? (hb[k >> 1] >> 16) : (hb[k >> 1] & 0xffffu);
          au[k] = (uint32_t)__half_as_ushort(__float2half_rn(u)) | (hbits << 16);
        }
        tc::tmem_st16(t_lane + 256 + col0, au);
      }
      tc::tmem_st_wait();
      // ---- EPI-r
      tc::mbar_wait(bar(BAR_R_DONE), it & 1);
      tc::fence_after_sync();
#pragma unroll 1
      for (int g = g_begin; g < g_end; ++g) {
        const int col0 = g * 16;
        uint32_t ar[16];
        tc::tmem_ld16(t_lane + col0, ar);
        uint8_t* hp = ht + (col0 >> 6) * CHUNK;
        const int s0 = (col0 & 63) >> 3;
        uint4* p0 = reinterpret_cast<uint4*>(hp + sw_off(r, s0));
        uint4* p1 = reinterpret_cast<uint4*>(hp + sw_off(r, s0 + 1));
        float hv[16];
        unpack8(*p0, hv); unpack8(*p1, hv + 8);
        float sb[16];
#pragma unroll
        for (int k4 = 0; k4 < 4; ++k4) {
          const float4 b4 = __ldg(reinterpret_cast<const float4*>(p.bg + col0) + k4);
          sb[4 * k4] = b4.x; sb[4 * k4 + 1] = b4.y; sb[4 * k4 + 2] = b4.z; sb[4 * k4 + 3] = b4.w;
        }
        tc::tmem_ld_wait();
        float o[16];
#pragma unroll
        for (int k = 0; k < 16; ++k)
          o[k] = tc::rcp_approx(1.0f + tc::ex2_approx(-kLog2e * (__uint_as_float(ar[k]) + sb[k]))) * hv[k];
        *p0 = pack8(o);
        *p1 = pack8(o + 8);
      }
      tc::fence_before_sync();
      tc::fence_proxy_async();
      __syncwarp();
      if (lane == 0) tc::mbar_arrive_cta0(bar(BAR_RH_FULL));
      // ---- EPI-c
      tc::mbar_wait(bar(BAR_C_DONE), it & 1);
      tc::fence_after_sync();
#pragma unroll 1
      for (int g = g_begin; g < g_end; ++g) {
        const int col0 = g * 16;
        uint32_t ac[16], uh[16];
        tc::tmem_ld16(t_lane + col0, ac);
        tc::tmem_ld16(t_lane + 256 + col0, uh);
        float sb[16];
#pragma unroll
        for (int k4 = 0; k4 < 4; ++k4) {
          const float4 b4 = __ldg(reinterpret_cast<const float4*>(p.bc + col0) + k4);
          sb[4 * k4] = b4.x; sb[4 * k4 + 1] = b4.y; sb[4 * k4 + 2] = b4.z; sb[4 * k4 + 3] = b4.w;
        }
        tc::tmem_ld_wait();
        float o[16];
#pragma unroll
        for (int k = 0; k < 16; ++k) {
          const float c = fmaf(-2.0f, tc::rcp_approx(1.0f + tc::ex2_approx(2.0f * kLog2e * (__uint_as_float(ac[k]) + sb[k]))), 1.0f);
          const float u = __half2float(__ushort_as_half((unsigned short)(uh[k] & 0xffffu)));
          const float hv = __half2float(__ushort_as_half((unsigned short)(uh[k] >> 16)));
          o[k] = fmaf(u, hv - c, c);
        }
        uint8_t* xp = xt + (col0 >> 6) * CHUNK;
        const int s0 = (col0 & 63) >> 3;
        *reinterpret_cast<uint4*>(xp + sw_off(r, s0)) = pack8(o);
        *reinterpret_cast<uint4*>(xp + sw_off(r, s0 + 1)) = pack8(o + 8);
        if (p.h_out && grow < p.N) {
          float* dst = p.h_out + grow * p.h + col0;
#pragma unroll
          for (int k4 = 0; k4 < 4; ++k4)
            if (col0 + k4 * 4 < p.h)
              *reinterpret_cast<float4*>(dst + k4 * 4) = make_float4(o[k4 * 4], o[k4 * 4 + 1], o[k4 * 4 + 2], o[k4 * 4 + 3]);
        }
      }
      tc::fence_before_sync();
      tc::fence_proxy_async();
      __syncwarp();
      if (lane == 0) tc::mbar_arrive_cta0(bar(BAR_TMEM_FREE));
      tc::named_bar_sync(1, 256);
      if (leader) {
        if (tile < p.n_tiles) {
          for (int kc = 0; kc < KC1; ++kc) tc::tma_store_2d(&tmH, kc * BKH, tile * BM, sX + kc * CHUNK);
          tc::tma_store_commit();
          tc::tma_store_wait_read();
        }
        tc::mbar_arrive(bar(BAR_X_EMPTY));
      }
    }
    if (leader) asm volatile("cp.async.bulk.wait_group 0;" ::: "memory");
  }

  tc::fence_before_sync();
  tc::cluster_sync_all();                                   // both CTAs are done with the pair's TMEM and barriers
  if (warp == WARP_TMA) tc::tmem_dealloc_pair(tmem_base, 512);
}

}  // namespace gw16x2

// =========================================================================================
bool ggnn_w16_supported(const ggnn_ctx* ctx) {
  return ctx->h > 128 && ctx->h % 4 == 0 && ctx->hw16 >= 192 && ctx->hw16 <= 512;
}

static int realloc_bytes(ggnn_ctx* ctx, void*& p, size_t bytes) {
  if (p && ctx->repack_in_place) return GGNN_OK;
  if (p) { cudaFree(p); p = nullptr; }
  GGNN_CUDA(ctx, cudaMalloc(&p, bytes));
  return GGNN_OK;
}

int ggnn_w16_pack_weights(ggnn_ctx* ctx, int slot, bool with_gru, cudaStream_t st) {
  if (!ggnn_w16_supported(ctx)) return GGNN_OK;
  const int h = ctx->h, hw = ctx->hw16, two_e = ctx->two_e;
  ggnn_edge_weights& ew = ctx->edge[slot];
  ew.absmax_w = -1.f;
  GGNN_TRY(realloc_bytes(ctx, ew.Wt_f16, (size_t)two_e * hw * hw * 2));
  packt::launch_edge_t<packt::OUT_F16>(ew.W, nullptr, 1.f, ew.Wt_f16, two_e, h, hw, st);
  GGNN_LAUNCHED(ctx);
  void* bp = ew.B_pad16;
  GGNN_TRY(realloc_bytes(ctx, bp, (size_t)two_e * hw * 4));
  ew.B_pad16 = reinterpret_cast<float*>(bp);
  w16::pad_bias_kernel<<<(two_e * hw + 255) / 256, 256, 0, st>>>(ew.B, ew.B_pad16, two_e, h, hw);
  GGNN_LAUNCHED(ctx);
  GGNN_TRY(ggnn_make_tmap_2d_f16(ctx, &ew.tm_Wt16, ew.Wt_f16, (int64_t)two_e * hw, hw, w16::BKH,
                                 w16::slice_width(hw, hw)));
  if (with_gru) {
    GGNN_TRY(realloc_bytes(ctx, ctx->Wgt_f16, (size_t)4 * hw * hw * 2));
    GGNN_TRY(realloc_bytes(ctx, ctx->Wct_f16, (size_t)2 * hw * hw * 2));
    w16::pack_gru_kernel<<<ctx->sm_count * 2, 256, 0, st>>>(ctx->Wg, reinterpret_cast<__half*>(ctx->Wgt_f16), 2, h, hw);
    GGNN_LAUNCHED(ctx);
    w16::pack_gru_kernel<<<ctx->sm_count * 2, 256, 0, st>>>(ctx->Wc, reinterpret_cast<__half*>(ctx->Wct_f16), 1, h, hw);
    GGNN_LAUNCHED(ctx);
    void* g = ctx->bg_w16; void* c = ctx->bc_w16;
    GGNN_TRY(realloc_bytes(ctx, g, (size_t)2 * hw * 4));
    GGNN_TRY(realloc_bytes(ctx, c, (size_t)hw * 4));
    ctx->bg_w16 = reinterpret_cast<float*>(g); ctx->bc_w16 = reinterpret_cast<float*>(c);
    w16::pad_bias_kernel<<<(2 * hw + 255) / 256, 256, 0, st>>>(ctx->bg, ctx->bg_w16, 2, h, hw);
    GGNN_LAUNCHED(ctx);
    w16::pad_bias_kernel<<<(hw + 255) / 256, 256, 0, st>>>(ctx->bc, ctx->bc_w16, 1, h, hw);
    GGNN_LAUNCHED(ctx);
    GGNN_TRY(ggnn_make_tmap_2d_f16(ctx, &ctx->tm_Wg16, ctx->Wgt_f16, 2 * hw, 2 * hw, w16::BKH,
                                   w16::slice_width(2 * hw, 2 * hw)));
    GGNN_TRY(ggnn_make_tmap_2d_f16(ctx, &ctx->tm_Wc16, ctx->Wct_f16, hw, 2 * hw, w16::BKH,
                                   w16::slice_width(2 * hw, hw)));
    if (hw <= 256) {     // fused GRU kernel: [hw n x 64 k] chunks; CTA pairs: [hw/2 n x 64 k] half chunks
      GGNN_TRY(ggnn_make_tmap_2d_f16(ctx, &ctx->tm_Wg16f, ctx->Wgt_f16, 2 * hw, 2 * hw, w16::BKH, hw));
      GGNN_TRY(ggnn_make_tmap_2d_f16(ctx, &ctx->tm_Wc16f, ctx->Wct_f16, hw, 2 * hw, w16::BKH, hw));
      GGNN_TRY(ggnn_make_tmap_2d_f16(ctx, &ctx->tm_Wg16h, ctx->Wgt_f16, 2 * hw, 2 * hw, w16::BKH, hw / 2));
      GGNN_TRY(ggnn_make_tmap_2d_f16(ctx, &ctx->tm_Wc16h, ctx->Wct_f16, hw, 2 * hw, w16::BKH, hw / 2));
    }
  }
  return GGNN_OK;
}

int ggnn_w16_state_in(ggnn_ctx* ctx, const float* h, void* state16, int64_t N, unsigned* absmax_bits,
                      cudaStream_t st) {
  if (N == 0) return GGNN_OK;
  prof_scope ps(ctx, GGNN_STAGE_OTHER, st);
  w16::state_in_kernel<<<ctx->sm_count * 8, 256, 0, st>>>(h, reinterpret_cast<__half*>(state16), N, ctx->h,
                                                           ctx->hw16, absmax_bits);
  GGNN_LAUNCHED(ctx);
  return GGNN_OK;
}

// message transform alone, any hw in {128, 192, .., 512}: used by f16.cu (hw = 128) as an A-B alternative to msg_f16_kernel
int ggnn_w16_msg(ggnn_ctx* ctx, int slot, const ggnn_plan& plan, void* msg16, void* state16, int hw, cudaStream_t st) {
  using namespace w16;
  const ggnn_edge_weights& ew = ctx->edge[slot];
  if (plan.max_tiles <= 0 || plan.nnz <= 0) return GGNN_OK;
  const int n_row_tiles = (int)ceil_div64(plan.N, BM);
  CUtensorMap tmState, tmMsg;
  GGNN_TRY(ggnn_make_tmap_2d_f16(ctx, &tmState, state16, plan.N, hw, BKH, BM));
  GGNN_TRY(ggnn_make_tmap_2d_f16(ctx, &tmMsg, msg16, plan.max_tiles * (int64_t)BM, hw, BKH, BM));
  params p{};
  p.A_gather = reinterpret_cast<const __half*>(state16);
  if (plan.dense_rows) { p.gather = nullptr; p.tiles_per_type = n_row_tiles; }
  else { p.gather = plan.msg_src; p.tiles_per_type = 1; }
  p.tile_type = plan.tile_type; p.n_tiles_dev = plan.n_tiles;
  p.KC = hw / BKH; p.KC1 = p.KC;
  p.Ntot = hw; p.NW = slice_width(hw, hw); p.S = (hw + p.NW - 1) / p.NW;
  p.bias = ew.B_pad16; p.hw = hw; p.h = ctx->h; p.N = plan.N;
  prof_scope ps(ctx, GGNN_STAGE_MSG, st);
  return launch<MODE_MSG>(ctx, p, tmState, tmState, ew.tm_Wt16, tmMsg, tmMsg, plan.max_tiles * p.S, st);
}

// padded FP32 biases [two_e, hw] of `slot` (also needed by ggnn_w16_msg at hw = 128)
int ggnn_w16_pack_bias(ggnn_ctx* ctx, int slot, int hw, cudaStream_t st) {
  ggnn_edge_weights& ew = ctx->edge[slot];
  void* bp = ew.B_pad16;
  GGNN_TRY(realloc_bytes(ctx, bp, (size_t)ctx->two_e * hw * 4));
  ew.B_pad16 = reinterpret_cast<float*>(bp);
  w16::pad_bias_kernel<<<(ctx->two_e * hw + 255) / 256, 256, 0, st>>>(ew.B, ew.B_pad16, ctx->two_e, ctx->h, hw);
  GGNN_LAUNCHED(ctx);
  return GGNN_OK;
}

// One timestep.  Scratch (all FP16, pitch hw): msg16 [max_tiles*128], acts16 / rh16 / u16 [N]; state16 [N] in/out.
int ggnn_step_w16(ggnn_ctx* ctx, int slot, const ggnn_plan& plan, void* msg16, void* acts16, void* rh16, void* u16,
                  void* state16, float* h_out, bool last_step, cudaStream_t st, const ggnn_w16_tape* tape) {
  using namespace w16;
  if (plan.N == 0) return GGNN_OK;
  const int h = ctx->h, hw = ctx->hw16;
  const ggnn_edge_weights& ew = ctx->edge[slot];
  const int n_row_tiles = (int)ceil_div64(plan.N, BM);
  CUtensorMap tmState, tmActs, tmRh, tmU, tmMsg;
  GGNN_TRY(ggnn_make_tmap_2d_f16(ctx, &tmState, state16, plan.N, hw, BKH, BM));
  GGNN_TRY(ggnn_make_tmap_2d_f16(ctx, &tmActs, acts16, plan.N, hw, BKH, BM));
  GGNN_TRY(ggnn_make_tmap_2d_f16(ctx, &tmRh, rh16, plan.N, hw, BKH, BM));
  GGNN_TRY(ggnn_make_tmap_2d_f16(ctx, &tmU, u16, plan.N, hw, BKH, BM));
  if (plan.max_tiles > 0 && plan.nnz > 0) {
    GGNN_TRY(ggnn_make_tmap_2d_f16(ctx, &tmMsg, msg16, plan.max_tiles * (int64_t)BM, hw, BKH, BM));
    params p{};
    p.A_gather = reinterpret_cast<const __half*>(state16);
    if (plan.dense_rows) { p.gather = nullptr; p.tiles_per_type = n_row_tiles; }
    else { p.gather = plan.msg_src; p.tiles_per_type = 1; }
    p.tile_type = plan.tile_type; p.n_tiles_dev = plan.n_tiles;
    p.KC = hw / BKH; p.KC1 = p.KC;
    p.Ntot = hw; p.NW = slice_width(hw, hw); p.S = (hw + p.NW - 1) / p.NW;
    p.bias = ew.B_pad16; p.hw = hw; p.h = h; p.N = plan.N;
    prof_scope ps(ctx, GGNN_STAGE_MSG, st);
    GGNN_TRY(launch<MODE_MSG>(ctx, p, tmState, tmState, ew.tm_Wt16, tmMsg, tmMsg, plan.max_tiles * p.S, st));
  }
  {
    prof_scope ps(ctx, GGNN_STAGE_AGG, st);
    const int64_t threads = plan.N * 32;
    aggregate_kernel<<<(unsigned)ceil_div64(threads, 256), 256, 0, st>>>(
        plan.tptr, plan.tmsg, reinterpret_cast<const __half*>(msg16), reinterpret_cast<__half*>(acts16), plan.N, hw,
        tape ? tape->x : nullptr, h);
    GGNN_LAUNCHED(ctx);
  }
  if (hw <= 256 && !ctx->no_fused && !tape) {       // GATES + CAND in one kernel, gates and r*h on chip
    GGNN_ENSURE_SMEM(ctx, GGNN_ATTR_W16_0 + 3, gw16::gru_w16_kernel, gw16::SMEM_BYTES);
    gw16::params g{};
    g.bg = ctx->bg_w16; g.bc = ctx->bc_w16; g.h_out = last_step ? h_out : nullptr;
    g.N = plan.N; g.h = h; g.hw = hw; g.n_tiles = n_row_tiles;
    static const bool pair = getenv("GGNN_GRU_2CTA") != nullptr;     // A-B switch
    if (pair && hw == 256) {
      GGNN_ENSURE_SMEM(ctx, GGNN_ATTR_W16_0 + 4, gw16x2::gru_w16x2_kernel, gw16x2::SMEM_BYTES);
      const int n_pairs = (n_row_tiles + 1) / 2;
      const int clusters = n_pairs < ctx->sm_count / 2 ? n_pairs : ctx->sm_count / 2;
      prof_scope ps(ctx, GGNN_STAGE_GRU_GATES, st);
      gw16x2::gru_w16x2_kernel<<<2 * clusters, gw16x2::NTHREADS, gw16x2::SMEM_BYTES, st>>>(g, tmActs, tmState, ctx->tm_Wg16h,
                                                                                           ctx->tm_Wc16h);
      GGNN_LAUNCHED(ctx);
      return GGNN_OK;
    }
    const int grid = n_row_tiles < ctx->sm_count ? n_row_tiles : ctx->sm_count;
    prof_scope ps(ctx, GGNN_STAGE_GRU_GATES, st);
    gw16::gru_w16_kernel<<<grid, gw16::NTHREADS, gw16::SMEM_BYTES, st>>>(g, tmActs, tmState, ctx->tm_Wg16f, ctx->tm_Wc16f);
    GGNN_LAUNCHED(ctx);
    return GGNN_OK;
  }
  {
    params p{};
    p.n_tiles = n_row_tiles;
    p.KC = 2 * hw / BKH; p.KC1 = hw / BKH;
    p.Ntot = 2 * hw; p.NW = slice_width(2 * hw, 2 * hw); p.S = (p.Ntot + p.NW - 1) / p.NW;
    p.bias = ctx->bg_w16; p.hw = hw; p.h = h; p.N = plan.N;
    p.state16 = reinterpret_cast<const __half*>(state16);
    if (tape) { p.tape_g = tape->r; p.tape_u = tape->u; }
    prof_scope ps(ctx, GGNN_STAGE_GRU_GATES, st);
    p.NB = row_blocks(n_row_tiles, 2 * hw, p.S);
    GGNN_TRY(launch<MODE_GATES>(ctx, p, tmActs, tmState, ctx->tm_Wg16, tmRh, tmU, (int64_t)n_row_tiles * p.S, st));
  }
  {
    params p{};
    p.n_tiles = n_row_tiles;
    p.KC = 2 * hw / BKH; p.KC1 = hw / BKH;
    p.Ntot = hw; p.NW = slice_width(2 * hw, hw); p.S = (p.Ntot + p.NW - 1) / p.NW;
    p.bias = ctx->bc_w16; p.hw = hw; p.h = h; p.N = plan.N;
    p.state16 = reinterpret_cast<const __half*>(state16);
    p.u16 = reinterpret_cast<const __half*>(u16);
    p.h_out = last_step ? h_out : nullptr;
    if (tape) p.tape_g = tape->c;
    prof_scope ps(ctx, GGNN_STAGE_GRU_CAND, st);
    p.NB = row_blocks(n_row_tiles, 2 * hw, p.S);
    GGNN_TRY(launch<MODE_CAND>(ctx, p, tmActs, tmRh, ctx->tm_Wc16, tmState, tmState, (int64_t)n_row_tiles * p.S, st));
  }
  return GGNN_OK;
}

// ---- training tape helpers (run_steps, SAVE_TAPE with GGNN_PREC_F16) ----------------------------------------
int ggnn_w16_state_out(ggnn_ctx* ctx, const void* state16, float* dst, int64_t N, cudaStream_t st) {
  if (N == 0) return GGNN_OK;
  w16::state_out_kernel<<<ctx->sm_count * 8, 256, 0, st>>>(reinterpret_cast<const __half*>(state16), dst, N, ctx->h, ctx->hw16);
  GGNN_LAUNCHED(ctx);
  return GGNN_OK;
}
int ggnn_w16_state_mask(ggnn_ctx* ctx, void* state16, const float* mask, float scale, int64_t N, cudaStream_t st) {
  if (N == 0) return GGNN_OK;
  w16::state_mask_kernel<<<ctx->sm_count * 8, 256, 0, st>>>(reinterpret_cast<__half*>(state16), mask, scale, N, ctx->h, ctx->hw16);
  GGNN_LAUNCHED(ctx);
  return GGNN_OK;
}
// edge weights of `slot` re-packed from W_src (null = the slot's own); mask != null: this step's effective weights
// W_src * mask * scale (edge-weight dropout), formed inside the pack
int ggnn_w16_repack_edge(ggnn_ctx* ctx, int slot, const float* W_src, const float* mask, float scale, cudaStream_t st) {
  ggnn_edge_weights& ew = ctx->edge[slot];
  packt::launch_edge_t<packt::OUT_F16>(W_src ? W_src : ew.W, mask, scale, ew.Wt_f16, ctx->two_e, ctx->h, ctx->hw16, st);
  GGNN_LAUNCHED(ctx);
  return GGNN_OK;
}
