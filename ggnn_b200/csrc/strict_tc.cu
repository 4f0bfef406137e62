// strict_tc.cu -- GGNN_PREC_FP32 on the tensor cores for hidden sizes up to 128 (hp <= 128).
//
// tcgen05 has no FP32-operand kind, so every operand a is carried as a_hi = tf32(a) and a_lo = tf32(a - a_hi) and a
// product is formed as a_hi*b_lo + a_lo*b_hi + a_hi*b_hi: three kind::tf32 MMAs into one FP32 accumulator in TMEM (the
// dropped a_lo*b_lo term is ~2^-22 relative).  Measured 6e-7 .. 2e-6 norm-wise against the FP64 oracle at h <= 128,
// like the FFMA kernels of step_fp32.cu.  The scheme stops at hp = 128 on purpose: the tensor core's FP32 accumulation
// is not exact enough over longer reductions -- at h = 400 (K = 800 in the GRU contractions) the same split measured
// 1.43e-5, outside the 1e-5 the strict mode promises -- so larger hidden sizes keep the FFMA kernels.
//
// ONE persistent, warp-specialised kernel serves the three contractions of a timestep (same math and operand layouts as
// tf32.cu):   MSG    M[r,:] = h[msg_src[r],:] @ W[type] + B[type]                    -> FP32 message rows
//             GATES  g = sigmoid([acts, h] @ Wg + bg);  rh = g[:, :h]*h;  u = g[:, h:]
//             CAND   c = tanh([acts, rh] @ Wc + bc);    h' = u*h + (1-u)*c             (state updated in place)
//   warps 0-3  producers: the unrounded FP32 activations arrive by 16-byte cp.async in a 6-deep ring of raw 16 KB chunks
//              (up to 80 KB per SM in flight, no registers held); each thread then re-reads its own pieces, splits them
//              and writes the two 128B-swizzled K-major tiles; one lane streams the hi / lo weight boxes by TMA.
//              Stage = A_hi | A_lo | B_hi | B_lo = 64 KB, 2 stages.
//   warp  4    one lane issues the MMAs (3 per K step): a_hi*b_hi into the main accumulator, the two cross terms into a
//              second one (two such pairs = all 512 TMEM columns, so N <= 128 per work item: GATES runs as two halves).
//              Why two: the tensor core ACCUMULATES with truncation, a bias that grows with the number of accumulator
//              updates -- 96 updates per GATES row in one accumulator measured 9e-6 end to end at T = 8; with the small
//              cross terms kept apart the main accumulator sees a third of the updates
//   warps 5-8  epilogue of the other accumulator: tcgen05.ld, bias, gate math (ex2.approx / rcp.approx: ~2e-7), stores
// Each CTA owns a contiguous range of 128-row tiles; the chunk ring runs across tile boundaries, so the loads of tile
// i+1 are in flight while tile i's MMAs and epilogue run.
#include "tf32.cuh"
#include "tc.cuh"
#include <stdlib.h>

namespace stc {

constexpr int BM = 128;
constexpr int BKF = 32;                       // floats per K chunk = one 128-byte swizzle row
constexpr int A_BYTES = BM * BKF * 4;         // 16 KB
constexpr int NST = 2;                        // operand stages: A_hi | A_lo | B_hi | B_lo = 64 KB each (N <= 128)
constexpr int RAW = 4;                        // staging ring of raw FP32 activation chunks (16 KB each) filled by cp.async
constexpr int STAGE_BYTES = 4 * A_BYTES;      // b_alloc = 16 KB: N <= 128
constexpr int EPI_BYTES = 2 * A_BYTES;        // epilogue: two [128 rows x 32 columns] FP32 transposition tiles
constexpr int OPER_BYTES = NST * STAGE_BYTES + RAW * A_BYTES + EPI_BYTES;      // 128 + 64 + 32 = 224 KB
constexpr int NPW = 4;                        // producer warps (8 measured no faster: the floor is the weight boxes re-streamed per tile)
constexpr int NPROD = NPW * 32;
constexpr int RPT = 1024 / NPROD;             // rows per producer thread and chunk (128 rows x 8 pieces / threads)
constexpr int RSTEP = NPROD / 8;              // ... which are RSTEP rows apart
constexpr int WARP_MMA = NPW;
constexpr int WARP_EPI0 = NPW + 1;            // 4 warps; warp % 4 = TMEM lane quarter (any 4 consecutive warps cover all)
constexpr int NTHREADS = (NPW + 5) * 32;
constexpr int SMEM_BYTES = OPER_BYTES + 256;

enum { MODE_MSG = 0, MODE_GATES = 1, MODE_CAND = 2 };
enum { BAR_FULL = 0 /*[3]*/, BAR_EMPTY = 3 /*[3]*/, BAR_ACC_FULL = 6 /*[2]*/, BAR_ACC_EMPTY = 8 /*[2]*/ };

struct params {
  const float* A1;          // [*, h] K segment 1 source rows
  const float* A2;          // [*, h] K segment 2 source rows (GATES: h, CAND: rh) or null
  const int32_t* gather;    // MSG: source row per message row (-1 = padding)
  const float* bias;        // MSG: [two_e, h]; GATES: padded [2hp]; CAND: [h]; or null
  int h, hp, Kp, Np;
  int64_t M;                // rows
  const int32_t* tile_type; // MSG
  const int32_t* n_tiles;   // MSG: device-side tile count
  int n_tiles_host;         // GATES / CAND
  float* out0;              // MSG: msg;  GATES: rh;  CAND: h state (in/out)
  float* out1;              // GATES: u
  const float* hstate;      // GATES: h
  const float* ubuf;        // CAND: u
  int b_bytes;              // bytes one TMA box of the weight maps delivers per K chunk
  int combined;             // tmHi maps the [hi | lo] side-by-side pack (hp == 128): ONE 32 KB box per K chunk instead of two
                            // 16 KB ones -- a bulk copy costs ~0.2 us whatever its size (scripts/micro/mc_stream.cu)
  int dbg;                  // timing experiments (GGNN_STC_DBG): 1 = epilogue touches no global memory, 2 = no split math,
                            // 4 = one MMA per K step instead of three
  unsigned long long* trace;// development aid (ggnn_debug_trace): %globaltimer of CTA 0's roles, MODE_GATES launches only
};
// the timing experiments of GGNN_STC_DBG exist only in builds with -DGGNN_STC_TRACE as well (a constant 0 otherwise)
#ifdef GGNN_STC_TRACE
#define STC_DBG(p) ((p).dbg)
#else
#define STC_DBG(p) 0
#endif
#ifdef GGNN_STC_TRACE                 // development build (GGNN_NVCC_EXTRA=-DGGNN_STC_TRACE, scripts/gpu_stc_trace.py)
#define STC_TRACE(slot, idx)                                                                     \
  do {                                                                                           \
    if (MODE == MODE_GATES && p.trace && blockIdx.x == 0 && (idx) < 64) {                        \
      unsigned long long t_;                                                                     \
      asm volatile("mov.u64 %0, %%globaltimer;" : "=l"(t_));                                     \
      p.trace[(slot) * 64 + (idx)] = t_;                                                         \
    }                                                                                            \
  } while (0)
#else
#define STC_TRACE(slot, idx) do { } while (0)
#endif

// gate math: ex2.approx / rcp.approx forms (~2e-7 relative).  Measured: accurate expf / tanhf change the end-to-end
// error by < 2 % (it is set by the tensor core's accumulation, see the header) and cost 3 ms per 4 096-graph chunk.
__device__ __forceinline__ float stc_sigmoid(float x) { return tc::fast_sigmoid(x); }
__device__ __forceinline__ float stc_tanh(float x) { return tc::fast_tanh(x); }

template <int MODE>
__global__ void __maxnreg__(128)   // 3 warps x 32 x 128 registers in the fullest scheduler partition: a background bitmask-queue CTA (3 x 32 x 40) fits beside it
split_gemm_kernel(const params p, const __grid_constant__ CUtensorMap tmHi, const __grid_constant__ CUtensorMap tmLo) {
  extern __shared__ __align__(1024) uint8_t smem[];
  const int tid = threadIdx.x, warp = tid >> 5, lane = tid & 31;
  const uint32_t base = tc::smem_u32(smem);
  const uint32_t b_alloc = A_BYTES;                   // N <= 128 rows of 128 bytes
  const uint32_t st_bytes = STAGE_BYTES;
  const uint32_t sRaw = base + NST * STAGE_BYTES;
  const uint32_t sEpi = sRaw + RAW * A_BYTES;
  const uint32_t bars = base + OPER_BYTES;
  auto bar = [&](int i) { return bars + 8u * i; };
  const uint32_t tmem_slot = bars + 8u * 10;

  // GATES runs as two N = hp halves per row tile (r columns, u columns): work item = 2 * row tile + half
  const int n_tiles = MODE == MODE_MSG ? *p.n_tiles : p.n_tiles_host;
  const int per = (n_tiles + (int)gridDim.x - 1) / (int)gridDim.x;
  const int t_begin = min(n_tiles, (int)blockIdx.x * per);
  const int t_end = min(n_tiles, t_begin + per);
  const int KC = p.Kp / BKF, KC1 = p.hp / BKF;

  if (warp == WARP_MMA && lane == 0) {
    for (int s = 0; s < 3; ++s) {
      tc::mbar_init(bar(BAR_FULL + s), NPROD + 1);
      tc::mbar_init(bar(BAR_EMPTY + s), 1);
    }
    for (int s = 0; s < 2; ++s) {
      tc::mbar_init(bar(BAR_ACC_FULL + s), 1);
      tc::mbar_init(bar(BAR_ACC_EMPTY + s), 4);
    }
    tc::fence_barrier_init();
  }
  if (warp == 0) {
    if (lane == 0) { tc::tma_prefetch_desc(&tmHi); tc::tma_prefetch_desc(&tmLo); }
    __syncwarp();
    tc::tmem_alloc(tmem_slot, 512);
  }
  tc::fence_before_sync();
  __syncthreads();
  tc::fence_after_sync();
  const uint32_t tmem_base = *reinterpret_cast<volatile uint32_t*>(smem + (tmem_slot - base));

  if (warp < NPW) {
    // =============================== producers ===============================
    const int c16 = tid & 7;                          // 16-byte slot within the 128-byte row
    const int rbase = tid >> 3;                       // rows rbase + RSTEP*j, j = 0..RPT-1
    uint32_t a_off[RPT];
#pragma unroll
    for (int j = 0; j < RPT; ++j) {
      const int r = rbase + RSTEP * j;
      a_off[j] = (r >> 3) * 1024 + (r & 7) * 128 + ((c16 ^ (r & 7)) << 4);
    }
    // Raw chunks are fetched RAW-1 ahead of their use with 16-byte cp.async (zero-filled for padding rows / columns); a
    // thread later reads back exactly the pieces it copied itself, so cp.async.wait_group is all the ordering needed.
    const uint32_t raw_off = (uint32_t)rbase * 128 + (uint32_t)c16 * 16;      // + j * RSTEP * 128
    const long long total = (long long)(t_end - t_begin) * KC;
    auto item_rows = [&](int item, int64_t (&src)[RPT]) {
      const int tile = MODE == MODE_GATES ? item >> 1 : item;
#pragma unroll
      for (int j = 0; j < RPT; ++j) {
        const int64_t grow = (int64_t)tile * BM + rbase + RSTEP * j;
        src[j] = -1;
        if (grow < p.M) src[j] = MODE == MODE_MSG ? (int64_t)__ldg(p.gather + grow) : grow;
      }
    };
    int64_t isrc[RPT], nsrc[RPT];                         // rows of the item being ISSUED, and of the one after it
    int i_item = t_begin, i_kc = 0;
    if (t_begin < t_end) item_rows(t_begin, isrc);
    if (t_begin + 1 < t_end) item_rows(t_begin + 1, nsrc);
    auto issue = [&](long long g) {                   // raw chunk g -> ring slot g % RAW
      if (g < total) {
        const bool seg2 = i_kc >= KC1;
        const float* Asrc = seg2 ? p.A2 : p.A1;
        const int kloc = (seg2 ? i_kc - KC1 : i_kc) * BKF + c16 * 4;
        const bool kin = kloc < p.h;                  // h % 4 == 0
        const uint32_t dst = sRaw + (uint32_t)(g % RAW) * A_BYTES + raw_off;
#pragma unroll
        for (int j = 0; j < RPT; ++j) {
          const bool ok = kin && isrc[j] >= 0;
          tc::cp_async16(dst + j * (RSTEP * 128), ok ? (const void*)(Asrc + isrc[j] * p.h + kloc) : (const void*)p.A1, ok ? 16u : 0u);
        }
        if (++i_kc == KC) {
          i_kc = 0;
          ++i_item;
#pragma unroll
          for (int j = 0; j < RPT; ++j) isrc[j] = nsrc[j];
          if (i_item + 1 < t_end) item_rows(i_item + 1, nsrc);
        }
      }
      asm volatile("cp.async.commit_group;" ::: "memory");
    };
    for (long long g = 0; g < RAW - 1; ++g) issue(g);
    int item = t_begin, kc = 0;
    for (long long c = 0; c < total; ++c) {
      if (tid == 0) STC_TRACE(0, c);
      issue(c + RAW - 1);
      asm volatile("cp.async.wait_group %0;" ::"n"(RAW - 1) : "memory");      // raw chunk c has landed (this thread's pieces)
      if (tid == 0) STC_TRACE(1, c);
      const uint8_t* raw = smem + (sRaw - base) + (size_t)(c % RAW) * A_BYTES + raw_off;
      float4 x[RPT];
#pragma unroll
      for (int j = 0; j < RPT; ++j) x[j] = *reinterpret_cast<const float4*>(raw + j * (RSTEP * 128));
      const int s = (int)(c % NST);
      if (c >= NST) tc::mbar_wait(bar(BAR_EMPTY + s), (uint32_t)((c / NST) - 1) & 1);
      if (tid == 0) STC_TRACE(2, c);
      const uint32_t st0 = base + (uint32_t)s * st_bytes;
      // (the weight boxes of this stage are requested by the MMA thread, one chunk ahead: a producer thread that also
      //  issued the two TMA loads made its whole warp -- and so every stage -- ~0.3 us late)
      uint8_t* a_hi = smem + (st0 - base);
#pragma unroll
      for (int j = 0; j < RPT; ++j) {
        float4 hi, lo;
        if (STC_DBG(p) & 2) { hi = x[j]; lo = x[j]; } else {
        // round-half-up in two integer ops (tc::to_tf32_fast) instead of cvt.rna (a conversion-unit instruction, 16 per
        // clock per SM): x - hi is exact either way and lo only has to be SOME TF32 number within 2^-11 of it
        hi.x = tc::to_tf32_fast(x[j].x); hi.y = tc::to_tf32_fast(x[j].y); hi.z = tc::to_tf32_fast(x[j].z); hi.w = tc::to_tf32_fast(x[j].w);
        lo.x = tc::to_tf32_fast(x[j].x - hi.x); lo.y = tc::to_tf32_fast(x[j].y - hi.y);
        lo.z = tc::to_tf32_fast(x[j].z - hi.z); lo.w = tc::to_tf32_fast(x[j].w - hi.w); }
        *reinterpret_cast<float4*>(a_hi + a_off[j]) = hi;
        *reinterpret_cast<float4*>(a_hi + A_BYTES + a_off[j]) = lo;
      }
      tc::fence_proxy_async();
      tc::mbar_arrive(bar(BAR_FULL + s));
      if (++kc == KC) { kc = 0; ++item; }
    }
    (void)item;
  } else if (warp == WARP_MMA) {
    // =============================== MMA issuer ===============================
    if (lane == 0) {
      const uint32_t idesc = tc::umma_idesc(/*TF32*/ 2, BM, p.Np);
      long long c = 0;
      const long long total = (long long)(t_end - t_begin) * KC;
      // weights of chunk cc = K chunk kcw of item itw: hi and lo boxes [Np rows x 32 k], swizzled by the TMA unit
      int itw = t_begin, kcw = 0;
      auto issue_w = [&](long long cc) {
        const int s = (int)(cc % NST);
        if (cc >= NST) tc::mbar_wait(bar(BAR_EMPTY + s), (uint32_t)((cc / NST) - 1) & 1);      // its last reader has retired
        const int b_row = MODE == MODE_MSG ? __ldg(p.tile_type + itw) * p.Np : (MODE == MODE_GATES ? (itw & 1) * p.Np : 0);
        const uint32_t st0 = base + (uint32_t)s * st_bytes;
        tc::mbar_arrive_expect_tx(bar(BAR_FULL + s), 2u * (uint32_t)p.b_bytes);
        if (p.combined) {       // rows [2 b_row, 2 b_row + Np) = hi -> B_hi, the next Np rows = lo -> B_lo (b_alloc == Np * 128 here)
          tc::tma_load_2d(st0 + 2 * A_BYTES, &tmHi, kcw * BKF, 2 * b_row, bar(BAR_FULL + s));
        } else {
          tc::tma_load_2d(st0 + 2 * A_BYTES, &tmHi, kcw * BKF, b_row, bar(BAR_FULL + s));
          tc::tma_load_2d(st0 + 2 * A_BYTES + b_alloc, &tmLo, kcw * BKF, b_row, bar(BAR_FULL + s));
        }
        if (++kcw == KC) { kcw = 0; ++itw; }
      };
      if (total > 0) issue_w(0);
      for (int tile = t_begin, i = 0; tile < t_end; ++tile, ++i) {
        const int buf = i & 1;
        if (i >= 2) tc::mbar_wait(bar(BAR_ACC_EMPTY + buf), (uint32_t)((i >> 1) - 1) & 1);
        tc::fence_after_sync();
        const uint32_t d = tmem_base + buf * 256, dx = d + 128;      // main | cross-term accumulator
        for (int kc = 0; kc < KC; ++kc, ++c) {
          const int s = (int)(c % NST);
          if (c + 1 < total) issue_w(c + 1);
          tc::mbar_wait(bar(BAR_FULL + s), (uint32_t)(c / NST) & 1);
          tc::fence_after_sync();
          STC_TRACE(3, c);
          const uint32_t a0 = base + (uint32_t)s * st_bytes, b0 = a0 + 2 * A_BYTES;
#pragma unroll
          for (int k = 0; k < BKF / 8; ++k) {         // UMMA K = 8 for TF32 (32 bytes)
            if (!(STC_DBG(p) & 4)) {
            tc::umma_tf32(dx, tc::umma_desc_sw128(a0 + k * 32), tc::umma_desc_sw128(b0 + b_alloc + k * 32), idesc, (kc | k) != 0);
            tc::umma_tf32(dx, tc::umma_desc_sw128(a0 + A_BYTES + k * 32), tc::umma_desc_sw128(b0 + k * 32), idesc, 1u); }
            tc::umma_tf32(d, tc::umma_desc_sw128(a0 + k * 32), tc::umma_desc_sw128(b0 + k * 32), idesc, (kc | k) != 0);
          }
          tc::umma_commit(bar(BAR_EMPTY + s));
        }
        tc::umma_commit(bar(BAR_ACC_FULL + buf));
      }
    }
    __syncwarp();
  } else {
    // =============================== epilogue (warps 5-8) ===============================
    // TMEM hands a thread ONE ROW (its lane) x 32 columns; global memory wants consecutive lanes on consecutive
    // addresses (measured: per-row accesses at a 512-byte stride made the epilogue, not the MMAs, the bottleneck: 18.2
    // -> 8.9 ms per 8 steps without them).  So every 32-column group goes through a swizzled [128 x 32] FP32 tile in
    // shared memory (two of them, alternating: one named barrier per group) and the bias / gate math / loads of h and
    // u / stores run in a second pass where 8 lanes cover one 128-byte row segment.
    const int q4 = warp & 3;                          // TMEM lane quarter this warp may access
    const int r = q4 * 32 + lane;
    const int etid = tid - WARP_EPI0 * 32;            // 0..127
    const int prow = etid >> 3, pc4 = etid & 7;       // second pass: rows prow + 16*k, 16-byte piece pc4
    const uint32_t t_lane = tmem_base + ((uint32_t)(q4 * 32) << 16);
    uint8_t* stg = smem + (sEpi - base);
    long long grp = 0;                                // 32-column groups staged so far (all items)
    for (int item = t_begin, i = 0; item < t_end; ++item, ++i) {
      const int tile = MODE == MODE_GATES ? item >> 1 : item;
      const int buf = i & 1;
      const int half = MODE == MODE_GATES ? (item & 1) : 0;
      bool ok[8];                                     // second-pass rows of this thread that exist (and are no padding)
#pragma unroll
      for (int k = 0; k < 8; ++k) {
        const int64_t row = (int64_t)tile * BM + prow + 16 * k;
        ok[k] = row < p.M && !(STC_DBG(p) & 1);
        if (MODE == MODE_MSG && ok[k]) ok[k] = __ldg(p.gather + row) >= 0;
      }
      const float* bias = p.bias;
      if (MODE == MODE_MSG && bias) bias += (size_t)__ldg(p.tile_type + tile) * p.h;
      if (MODE == MODE_GATES) bias += half * p.hp;
      if (etid == 0) STC_TRACE(4, i);
      tc::mbar_wait(bar(BAR_ACC_FULL + buf), (uint32_t)(i >> 1) & 1);
      tc::fence_after_sync();
      if (etid == 0) STC_TRACE(5, i);
      const int ngroups = p.Np / 32;
      for (int cc = 0; cc < ngroups; ++cc, ++grp) {
        uint32_t acc[32], accx[32];
        tc::tmem_ld32(t_lane + buf * 256 + cc * 32, acc);
        tc::tmem_ld32(t_lane + buf * 256 + 128 + cc * 32, accx);
        tc::tmem_ld_wait();
        if (cc == ngroups - 1) {                      // the accumulators are in registers: hand the TMEM buffer back
          tc::fence_before_sync();
          __syncwarp();
          if (lane == 0) tc::mbar_arrive(bar(BAR_ACC_EMPTY + buf));
        }
        uint8_t* sb = stg + (grp & 1) * A_BYTES;
#pragma unroll
        for (int q = 0; q < 8; ++q) {
          float4 v;
          v.x = __uint_as_float(acc[q * 4 + 0]) + __uint_as_float(accx[q * 4 + 0]);
          v.y = __uint_as_float(acc[q * 4 + 1]) + __uint_as_float(accx[q * 4 + 1]);
          v.z = __uint_as_float(acc[q * 4 + 2]) + __uint_as_float(accx[q * 4 + 2]);
          v.w = __uint_as_float(acc[q * 4 + 3]) + __uint_as_float(accx[q * 4 + 3]);
          *reinterpret_cast<float4*>(sb + r * 128 + ((q ^ (r & 7)) << 4)) = v;
        }
        tc::named_bar_sync(1, 128);                   // tile complete (and, two groups on, free again: see above)
        const int col = cc * 32 + pc4 * 4;            // column within this item's N = hp block
        if (col < p.h) {
          float4 bb = make_float4(0.f, 0.f, 0.f, 0.f);
          if (bias) bb = __ldg(reinterpret_cast<const float4*>(bias + col));
          // all loads of the 8 pieces first (the stores below may alias them as far as the compiler knows: issued one
          // by one behind the stores they cost a full memory latency per piece), then the math and the stores
          float4 v[8], hv[8], uv[8];
#pragma unroll
          for (int k = 0; k < 8; ++k) {
            const int lr = prow + 16 * k;
            const int64_t row = (int64_t)tile * BM + lr;
            v[k] = *reinterpret_cast<const float4*>(sb + lr * 128 + ((pc4 ^ (lr & 7)) << 4));
            hv[k] = uv[k] = make_float4(0.f, 0.f, 0.f, 0.f);
            if (ok[k]) {
              if (MODE == MODE_GATES && half == 0) hv[k] = *reinterpret_cast<const float4*>(p.hstate + row * p.h + col);
              if (MODE == MODE_CAND) {
                hv[k] = *reinterpret_cast<const float4*>(p.out0 + row * p.h + col);
                uv[k] = *reinterpret_cast<const float4*>(p.ubuf + row * p.h + col);
              }
            }
          }
#pragma unroll
          for (int k = 0; k < 8; ++k) {
            if (!ok[k]) continue;
            const int64_t row = (int64_t)tile * BM + prow + 16 * k;
            const float4 z = make_float4(v[k].x + bb.x, v[k].y + bb.y, v[k].z + bb.z, v[k].w + bb.w);
            if (MODE == MODE_MSG) {
              *reinterpret_cast<float4*>(p.out0 + row * p.h + col) = z;
            } else if (MODE == MODE_GATES) {
              const float4 g = make_float4(stc_sigmoid(z.x), stc_sigmoid(z.y), stc_sigmoid(z.z), stc_sigmoid(z.w));
              if (half == 0)
                *reinterpret_cast<float4*>(p.out0 + row * p.h + col) =
                    make_float4(g.x * hv[k].x, g.y * hv[k].y, g.z * hv[k].z, g.w * hv[k].w);
              else
                *reinterpret_cast<float4*>(p.out1 + row * p.h + col) = g;
            } else {
              const float c0 = stc_tanh(z.x), c1 = stc_tanh(z.y), c2 = stc_tanh(z.z), c3 = stc_tanh(z.w);
              float4 o;
              o.x = uv[k].x * hv[k].x + (1.0f - uv[k].x) * c0;
              o.y = uv[k].y * hv[k].y + (1.0f - uv[k].y) * c1;
              o.z = uv[k].z * hv[k].z + (1.0f - uv[k].z) * c2;
              o.w = uv[k].w * hv[k].w + (1.0f - uv[k].w) * c3;
              *reinterpret_cast<float4*>(p.out0 + row * p.h + col) = o;
            }
          }
        }
      }
    }
  }

  tc::fence_before_sync();
  __syncthreads();
  if (warp == 0) tc::tmem_dealloc(tmem_base, 512);
}

template <int MODE>
int launch(ggnn_ctx* ctx, const params& p, const CUtensorMap& hi, const CUtensorMap& lo, int64_t tiles_bound, cudaStream_t st) {
  if (tiles_bound <= 0) return GGNN_OK;
  if (!(ctx->attr_mask & (1u << GGNN_ATTR_STRICT_TC))) {
    GGNN_CUDA(ctx, cudaFuncSetAttribute(split_gemm_kernel<MODE_MSG>, cudaFuncAttributeMaxDynamicSharedMemorySize, SMEM_BYTES));
    GGNN_CUDA(ctx, cudaFuncSetAttribute(split_gemm_kernel<MODE_GATES>, cudaFuncAttributeMaxDynamicSharedMemorySize, SMEM_BYTES));
    GGNN_CUDA(ctx, cudaFuncSetAttribute(split_gemm_kernel<MODE_CAND>, cudaFuncAttributeMaxDynamicSharedMemorySize, SMEM_BYTES));
    ctx->attr_mask |= 1u << GGNN_ATTR_STRICT_TC;
  }
  const int grid = (int)(tiles_bound < ctx->sm_count ? tiles_bound : ctx->sm_count);
  static const int dbg = getenv("GGNN_STC_DBG") ? atoi(getenv("GGNN_STC_DBG")) : 0;
  params pd = p;
  pd.dbg = dbg;
  pd.trace = ctx->dbg_trace;
  split_gemm_kernel<MODE><<<grid, NTHREADS, SMEM_BYTES, st>>>(pd, hi, lo);
  GGNN_LAUNCHED(ctx);
  return GGNN_OK;
}

}  // namespace stc

bool ggnn_fp32_tc_enabled(const ggnn_ctx* ctx) {
  const bool ffma = getenv("GGNN_FP32_FFMA") != nullptr;               // second opinion / A-B switch (read per call)
  return !ffma && ggnn_tf32_supported(ctx) && ctx->hp <= 128 && ctx->Wgt_tf32_lo != nullptr;
}

int ggnn_step_fp32_tc(ggnn_ctx* ctx, int slot, const ggnn_plan& plan, const ggnn_step_bufs& bufs, float* h_state,
                      cudaStream_t st) {
  using namespace stc;
  const int h = ctx->h, hp = ctx->hp;
  const ggnn_edge_weights& ew = ctx->edge[slot];
  if (plan.N == 0) return GGNN_OK;
  if (!ew.Wt_tf32_lo) GGNN_FAIL(ctx, GGNN_E_STATE, "internal: strict tensor-core mode without residual weight packs");
  static const bool split_boxes = getenv("GGNN_STC_SPLIT_BOXES") != nullptr;      // A-B switch: two 16 KB weight boxes per chunk
  const bool hl = !split_boxes && hp == 128 && ctx->hl_ok && ew.Wt_tf32_hl != nullptr;
  if (plan.max_tiles > 0 && plan.nnz > 0) {
    prof_scope ps(ctx, GGNN_STAGE_MSG, st);
    params p{};
    p.A1 = h_state; p.gather = plan.msg_src; p.bias = ew.B;
    p.h = h; p.hp = hp; p.Kp = hp; p.Np = hp;
    p.M = plan.max_tiles * GGNN_TILE_M;
    p.tile_type = plan.tile_type; p.n_tiles = plan.n_tiles;
    p.out0 = bufs.msg;
    p.b_bytes = ew.tm_rows * BKF * 4;
    p.combined = hl ? 1 : 0;
    GGNN_TRY(launch<MODE_MSG>(ctx, p, hl ? ew.tm_Wt_hl : ew.tm_Wt, ew.tm_Wt_lo, plan.max_tiles, st));
  }
  GGNN_TRY(ggnn_aggregate(ctx, plan, bufs.msg, bufs.acts, h, /*round_tf32=*/false, st));
  const int64_t mt = ceil_div64(plan.N, BM);
  {
    prof_scope ps(ctx, GGNN_STAGE_GRU_GATES, st);
    params p{};
    p.A1 = bufs.acts; p.A2 = h_state; p.bias = ctx->bg_pad;
    p.h = h; p.hp = hp; p.Kp = 2 * hp; p.Np = hp; p.M = plan.N; p.n_tiles_host = (int)(2 * mt);
    p.out0 = bufs.rh; p.out1 = bufs.u; p.hstate = h_state;
    p.b_bytes = hp * BKF * 4;
    p.combined = hl ? 1 : 0;
    GGNN_TRY(launch<MODE_GATES>(ctx, p, hl ? ctx->tm_Wg_hl : ctx->tm_Wg_half, ctx->tm_Wg_lo, 2 * mt, st));
  }
  {
    prof_scope ps(ctx, GGNN_STAGE_GRU_CAND, st);
    params p{};
    p.A1 = bufs.acts; p.A2 = bufs.rh; p.bias = ctx->bc;
    p.h = h; p.hp = hp; p.Kp = 2 * hp; p.Np = hp; p.M = plan.N; p.n_tiles_host = (int)mt;
    p.out0 = h_state; p.ubuf = bufs.u;
    p.b_bytes = ctx->tm_rows_c * BKF * 4;
    p.combined = hl ? 1 : 0;
    GGNN_TRY(launch<MODE_CAND>(ctx, p, hl ? ctx->tm_Wc_hl : ctx->tm_Wc, ctx->tm_Wc_lo, mt, st));
  }
  return GGNN_OK;
}
