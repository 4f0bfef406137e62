// plan_bits.cu -- dense path, fast route: build the propagation plan STRAIGHT from the bitmask of
// the adjacency, without materialising the canonical CSR.
//
// The canonical CSR has one row per (graph, channel, target): 48 M rows for 2 M edges at C3
// (4 096 graphs), and every kernel that walks it moves ~200 MB.  Here one CTA per graph (one thread
// per target node) walks its graph's 2E x v rows of the bitmask once, so the only per-row array is
// a 16-bit count:
//   count  (before the planning sync)  per-row popcounts -> rowcnt16; per-(channel, graph) segment
//          lengths -> segposT; per-node in-degree -> tptr;  two scans; nnz = last(tptr)
//   place  (after it)                  per channel, rank of every row inside its segment (warp
//          scans over rowcnt16 in shared memory); thread t then emits, channel by channel, the
//          message rows of its node: msg_src[row] = source, tmsg[cursor++] = row
// The message-row numbering and the per-node (channel, source) order of tmsg are exactly those of
// plan.cu (same tile table, same segment offsets), so results are bit-identical to the CSR route.
#include "plan.cuh"
#include "plan_bits.cuh"
#include <stdlib.h>

namespace {

constexpr int MAX_THREADS = 1024;

// bits [lo, hi) of the flat bitmask -> popcount
__device__ __forceinline__ int row_popc(const uint32_t* __restrict__ bits, int64_t lo, int v) {
  if ((v & 127) == 0 && (lo & 127) == 0) {          // aligned fast path: 128-bit loads
    int c = 0;
    const uint4* p = reinterpret_cast<const uint4*>(bits + (lo >> 5));
    for (int q = 0; q < (v >> 7); ++q) {
      const uint4 m = __ldg(p + q);
      c += __popc(m.x) + __popc(m.y) + __popc(m.z) + __popc(m.w);
    }
    return c;
  }
  const int64_t hi = lo + v;
  int c = 0;
  for (int64_t w = lo >> 5; w <= (hi - 1) >> 5; ++w) {
    uint32_t m = __ldg(bits + w);
    const int64_t wb = w << 5;
    if (lo > wb) m &= 0xffffffffu << (int)(lo - wb);
    if (hi < wb + 32) m &= 0xffffffffu >> (int)(wb + 32 - hi);
    c += __popc(m);
  }
  return c;
}

// one CTA per graph, thread t = target node.  smem: uint16 part[nwarps][two_e] (a warp's 32 rows hold at most 32 * 255
// edges).  16-bit on purpose: at v = 128 the CTA then needs 736 bytes of shared memory and fits beside a CTA of the
// 226 KB GRU kernels when the counts of the NEXT call run on the prefetch stream (ggnn_dense_prefetch)
__global__ void __launch_bounds__(MAX_THREADS)
graph_count_kernel(const uint32_t* __restrict__ bits, int64_t b, int two_e, int v,
                   uint16_t* __restrict__ rowcnt, int32_t* __restrict__ seglenT,
                   int32_t* __restrict__ tcnt, int word_fold, int32_t* __restrict__ regular) {
  extern __shared__ uint16_t part[];
  const int64_t g = blockIdx.x;
  const int t = threadIdx.x, lane = t & 31, wid = t >> 5, nw = blockDim.x >> 5;
  // word-order graph (v == 128 = blockDim.x): row t of channel E-1 is exactly {t-1}, of channel 2E-1 exactly {t+1}
  const int cp = two_e / 2 - 1, cn = two_e - 1;
  int reg = 0;
  if (word_fold) {
    const uint4 mp = __ldg(reinterpret_cast<const uint4*>(bits) + ((g * two_e + cp) * v + t));   // 128 bits = one row
    const uint4 mn = __ldg(reinterpret_cast<const uint4*>(bits) + ((g * two_e + cn) * v + t));
    const uint32_t wp[4] = {mp.x, mp.y, mp.z, mp.w}, wn[4] = {mn.x, mn.y, mn.z, mn.w};
    bool ok = true;
#pragma unroll
    for (int q = 0; q < 4; ++q) {
      const uint32_t ep = (t >= 1 && ((t - 1) >> 5) == q) ? 1u << ((t - 1) & 31) : 0u;
      const uint32_t en = (t <= 126 && ((t + 1) >> 5) == q) ? 1u << ((t + 1) & 31) : 0u;
      ok = ok && wp[q] == ep && wn[q] == en;
    }
    reg = __syncthreads_and(ok ? 1 : 0);
    if (t == 0) regular[g] = reg;
  }
  int deg = 0;
#pragma unroll 8
  for (int e = 0; e < two_e; ++e) {
    const int64_t r = (g * two_e + e) * v + t;
    int c = 0;
    if (t < v) {
      c = (reg && (e == cp || e == cn)) ? 0 : row_popc(bits, r * v, v);
      rowcnt[r] = (uint16_t)c;
    }
    deg += c;
    const int s = (int)__reduce_add_sync(0xffffffffu, (unsigned)c);      // redux.sync: one instruction per channel
    if (lane == 0) part[wid * two_e + e] = (uint16_t)s;
  }
  if (t < v) tcnt[g * v + t] = deg;
  if (g == b - 1 && t == 0) { tcnt[b * v] = 0; seglenT[b * two_e] = 0; }
  __syncthreads();
  for (int e = t; e < two_e; e += blockDim.x) {
    int s = 0;
    for (int w = 0; w < nw; ++w) s += part[w * two_e + e];
    seglenT[(int64_t)e * b + g] = s;
  }
}

// one CTA per graph, thread t = target node.  smem: uint16 rank[two_e][vt] (+ uint32 nzmap[vt][nzw] when wide)
// wide (v % 8 == 0 and the extra 4 * vt * nzw bytes fit): the row counts come in as 16-byte loads -- 8 targets of one
// channel per load, two_e * v / 8 loads spread over the CTA, all independent -- instead of two_e 2-byte loads per thread
// issued eight at a time; which channels of a target are non-empty is then collected in shared memory (atomicOr per
// non-empty row, ~4 per node) instead of in registers.
__global__ void __launch_bounds__(MAX_THREADS)
graph_place_kernel(const uint32_t* __restrict__ bits, const uint16_t* __restrict__ rowcnt, int64_t b,
                   int two_e, int v, const int32_t* __restrict__ segposT,
                   const int32_t* __restrict__ tile_base, const int32_t* __restrict__ tptr,
                   int32_t* __restrict__ msg_src, int32_t* __restrict__ tmsg, int wide) {
  extern __shared__ __align__(16) uint16_t rank[];
  const int64_t g = blockIdx.x;
  const int t = threadIdx.x, lane = t & 31, wid = t >> 5, nw = blockDim.x >> 5, vt = blockDim.x;
  // load this target's row counts (independent loads, several in flight) and remember which channels are
  // non-empty: the emit loop below then visits ~4 rows per node instead of all 2E
  unsigned long long nz[4] = {0ull, 0ull, 0ull, 0ull};          // two_e <= 256
  if (wide) {
    const int nzw = (two_e + 31) >> 5;
    uint32_t* nzmap = reinterpret_cast<uint32_t*>(rank + (size_t)two_e * vt);      // [vt][nzw]
    for (int i = t; i < vt * nzw; i += vt) nzmap[i] = 0u;
    if (vt > v)                                                  // padding targets: count 0 in every channel
      for (int i = t; i < two_e * (vt - v); i += vt) rank[(i / (vt - v)) * vt + v + i % (vt - v)] = 0;
    __syncthreads();
    const int v8 = v >> 3, items = two_e * v8;
#pragma unroll 4
    for (int i = t; i < items; i += vt) {
      const int e = i / v8, j = i - e * v8;
      const uint4 c8 = __ldg(reinterpret_cast<const uint4*>(rowcnt + (g * two_e + e) * v) + j);
      *reinterpret_cast<uint4*>(rank + e * vt + 8 * j) = c8;
      const uint32_t w[4] = {c8.x, c8.y, c8.z, c8.w};
#pragma unroll
      for (int k = 0; k < 8; ++k)
        if ((w[k >> 1] >> (16 * (k & 1))) & 0xffffu) atomicOr(nzmap + (8 * j + k) * nzw + (e >> 5), 1u << (e & 31));
    }
    __syncthreads();
    for (int q = 0; q < nzw; ++q) {                              // this target's non-empty channels -> registers
      const unsigned long long m = nzmap[t * nzw + q];
      const int q64 = q >> 1;
      const unsigned long long bitsq = (q & 1) ? m << 32 : m;
      nz[0] |= q64 == 0 ? bitsq : 0ull; nz[1] |= q64 == 1 ? bitsq : 0ull;
      nz[2] |= q64 == 2 ? bitsq : 0ull; nz[3] |= q64 == 3 ? bitsq : 0ull;
    }
  } else {
#pragma unroll 8
  for (int e = 0; e < two_e; ++e) {
    const uint16_t c = t < v ? rowcnt[(g * two_e + e) * v + t] : (uint16_t)0;
    rank[e * vt + t] = c;
    const unsigned long long bit = c ? 1ull << (e & 63) : 0ull;
    const int q = e >> 6;                                       // selects keep nz[] in registers
    nz[0] |= q == 0 ? bit : 0ull; nz[1] |= q == 1 ? bit : 0ull;
    nz[2] |= q == 2 ? bit : 0ull; nz[3] |= q == 3 ? bit : 0ull;
  }
  }
  __syncthreads();
  // exclusive scan of every channel's row counts over the target index (warp per channel)
  for (int e = wid; e < two_e; e += nw) {
    int carry = 0;
    for (int t0 = 0; t0 < vt; t0 += 32) {
      const int c = rank[e * vt + t0 + lane];
      int incl = c;
#pragma unroll
      for (int d = 1; d < 32; d <<= 1) {
        const int y = __shfl_up_sync(0xffffffffu, incl, d);
        if (lane >= d) incl += y;
      }
      rank[e * vt + t0 + lane] = (uint16_t)(carry + incl - c);
      carry += __shfl_sync(0xffffffffu, incl, 31);
    }
  }
  __syncthreads();
  if (t >= v) return;
  int cursor = tptr[g * v + t];
  const int end = tptr[g * v + t + 1];
#pragma unroll
  for (int q = 0; q < 4; ++q) {
    unsigned long long m64 = nz[q];
    while (m64) {
      const int e = q * 64 + __ffsll((long long)m64) - 1;
      m64 &= m64 - 1;
      const int64_t r = (g * two_e + e) * v + t;
      int mrow = tile_base[e] * GGNN_TILE_M + (segposT[(int64_t)e * b + g] - segposT[(int64_t)e * b]) +
                 rank[e * vt + t];
      const int64_t lo = r * v, hi = lo + v;
      if ((v & 127) == 0) {                       // rows of whole 128-bit groups: one load per group instead of four
        const uint4* p4 = reinterpret_cast<const uint4*>(bits + (lo >> 5));
        for (int q4 = 0; q4 < (v >> 7); ++q4) {
          const uint4 m4 = __ldg(p4 + q4);
          const uint32_t mw[4] = {m4.x, m4.y, m4.z, m4.w};
#pragma unroll
          for (int k = 0; k < 4; ++k) {
            uint32_t m = mw[k];
            while (m) {
              const int bpos = __ffs(m) - 1;
              m &= m - 1;
              msg_src[mrow] = (int32_t)(g * v + (q4 * 128 + k * 32 + bpos));
              tmsg[cursor++] = mrow++;
            }
          }
        }
        continue;
      }
      for (int64_t w = lo >> 5; w <= (hi - 1) >> 5; ++w) {
        uint32_t m = __ldg(bits + w);
        const int64_t wb = w << 5;
        if (lo > wb) m &= 0xffffffffu << (int)(lo - wb);
        if (hi < wb + 32) m &= 0xffffffffu >> (int)(wb + 32 - hi);
        while (m) {
          const int bpos = __ffs(m) - 1;
          m &= m - 1;
          msg_src[mrow] = (int32_t)(g * v + (wb + bpos - lo));
          tmsg[cursor++] = mrow++;
        }
      }
    }
  }
}

// same tile table as plan.cu's tiles_kernel (kept separate: that one is file-local)
__global__ void bits_tiles_kernel(const int32_t* __restrict__ segposT, int64_t b, int two_e,
                                  int32_t* __restrict__ tile_base, int32_t* __restrict__ tile_type,
                                  int32_t* __restrict__ n_tiles, int64_t max_tiles) {
  extern __shared__ int32_t sh[];   // [two_e + 1]
  if (threadIdx.x == 0) {
    int32_t run = 0;
    for (int e = 0; e < two_e; ++e) {
      const int32_t cnt = segposT[(int64_t)(e + 1) * b] - segposT[(int64_t)e * b];
      sh[e] = run;
      run += (cnt + GGNN_TILE_M - 1) / GGNN_TILE_M;
    }
    sh[two_e] = run;
    *n_tiles = run;
  }
  __syncthreads();
  for (int e = threadIdx.x; e <= two_e; e += blockDim.x) tile_base[e] = sh[e];
  for (int e = 0; e < two_e; ++e)
    for (int t = sh[e] + threadIdx.x; t < sh[e + 1]; t += blockDim.x)
      if (t < max_tiles) tile_type[t] = e;
}

inline int threads_for(int v) { return (int)align_up((size_t)v, 32); }

}  // namespace

bool ggnn_plan_bits_supported(int two_e, int v) {
  const int vt = threads_for(v);
  return v >= 1 && v <= 255 /* ranks inside a segment fit 16 bits */ && vt <= MAX_THREADS && two_e <= 256 &&
         (size_t)two_e * vt * 2 <= 48 * 1024 &&
         (size_t)(vt / 32) * two_e * 4 <= 48 * 1024;
}

void ggnn_plan_bits_take_pre(ws_alloc& a, int64_t b, int two_e, int v, ggnn_plan_bits_pre* pre) {
  const int64_t rows = b * two_e * v, N = b * v, nseg = b * two_e;
  pre->bits = a.take<uint32_t>((size_t)ceil_div64(rows * v, 32) + 4);
  pre->rowcnt = a.take<uint16_t>((size_t)rows + 2);
  pre->segposT = a.take<int32_t>((size_t)nseg + 1);
  pre->tptr = a.take<int32_t>((size_t)N + 1);
  pre->regular = a.take<int32_t>((size_t)b + 1);
  pre->scan_elems = ggnn_scan_tmp_elems((nseg > N ? nseg : N) + 1);
  pre->scan_tmp = a.take<int32_t>(pre->scan_elems);
}

int ggnn_plan_bits_count(ggnn_ctx* ctx, const ggnn_plan_bits_pre& pre, int64_t b, int two_e, int v, bool word_fold,
                         int64_t* nnz_out_dev, cudaStream_t st) {
  prof_scope ps(ctx, GGNN_STAGE_COMPACT_REST, st);
  const int vt = threads_for(v);
  if (word_fold && (v != GGNN_TILE_M || two_e < 2 || two_e % 2)) GGNN_FAIL(ctx, GGNN_E_ARG, "internal: word_fold needs v == 128 and an even channel count");
  graph_count_kernel<<<(unsigned)b, vt, (size_t)(vt / 32) * two_e * 2, st>>>(pre.bits, b, two_e, v, pre.rowcnt,
                                                                            pre.segposT, pre.tptr, word_fold ? 1 : 0,
                                                                            pre.regular);
  GGNN_LAUNCHED(ctx);
  GGNN_TRY(ggnn_exclusive_scan_i32(ctx, pre.segposT, pre.segposT, b * two_e + 1, nullptr, pre.scan_tmp, pre.scan_elems, st));
  GGNN_TRY(ggnn_exclusive_scan_i32(ctx, pre.tptr, pre.tptr, b * v + 1, nnz_out_dev, pre.scan_tmp, pre.scan_elems, st));
  return GGNN_OK;
}

int ggnn_plan_bits_place(ggnn_ctx* ctx, const ggnn_plan_bits_pre& pre, int64_t b, int two_e, int v,
                         int64_t nnz, ggnn_plan* plan, ws_alloc& a, bool dry, cudaStream_t st) {
  plan->N = b * v;
  plan->nnz = nnz;
  plan->dense_rows = false;
  plan->max_tiles = nnz / GGNN_TILE_M + two_e + 1;
  plan->n_tiles = a.take<int32_t>(4);
  plan->msg_src = a.take<int32_t>((size_t)plan->max_tiles * GGNN_TILE_M);
  plan->tile_type = a.take<int32_t>((size_t)plan->max_tiles);
  plan->tmsg = a.take<int32_t>((size_t)nnz + 1);
  int32_t* tile_base = a.take<int32_t>((size_t)two_e + 1);
  plan->tptr = pre.tptr;
  if (dry) return GGNN_OK;
  prof_scope ps(ctx, GGNN_STAGE_PLAN, st);
  bits_tiles_kernel<<<1, 256, (two_e + 1) * sizeof(int32_t), st>>>(pre.segposT, b, two_e, tile_base, plan->tile_type,
                                                                  plan->n_tiles, plan->max_tiles);
  GGNN_LAUNCHED(ctx);
  GGNN_CUDA(ctx, cudaMemsetAsync(plan->msg_src, 0xFF, (size_t)plan->max_tiles * GGNN_TILE_M * sizeof(int32_t), st));
  const int vt = threads_for(v);
  const size_t rank_bytes = (size_t)two_e * vt * 2, nz_bytes = (size_t)vt * ((two_e + 31) / 32) * 4;
  static const bool no_wide = getenv("GGNN_PLACE_NARROW") != nullptr;       // A-B switch
  const bool wide = !no_wide && v % 8 == 0 && rank_bytes + nz_bytes <= 48 * 1024;
  graph_place_kernel<<<(unsigned)b, vt, rank_bytes + (wide ? nz_bytes : 0), st>>>(pre.bits, pre.rowcnt, b, two_e, v,
                                                                                  pre.segposT, tile_base, pre.tptr,
                                                                                  plan->msg_src, plan->tmsg, wide ? 1 : 0);
  GGNN_LAUNCHED(ctx);
  return GGNN_OK;
}
