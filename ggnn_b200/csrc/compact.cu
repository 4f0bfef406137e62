// compact.cu -- A1: dense 0/1 adjacency [b, 2E, v, v] -> canonical CSR, bit-exact with
// np.argwhere(adj != 0) (reference edge set: chem_tensorflow_dense.py:65-83).
//
// Row-major order of [b][e][t][s] already IS the canonical (graph, channel, target,
// source) order, so this is a stable stream compaction, no sort:
//   P1  stream the dense tensor once (128-bit loads, several in flight per thread) and
//       emit one bit per element -> bitmask (1/32 .. 1/8 of the input bytes).  HBM-bound;
//       this is the only pass that touches the dense data.
//   P2  per-row popcounts of the bitmask -> rowptr (as counts)
//   P3  exclusive scan of rowptr in place (scan.cu) -> nnz
//   P4  per-row bit iteration -> col
#include "common.cuh"
#include <stdlib.h>
#include <mutex>

namespace {

constexpr int P1_THREADS = 256;
constexpr int P1_UNROLL = 4;

__device__ __forceinline__ float4 ldg_stream_f4(const float4* p) {
  float4 r;
  asm volatile("ld.global.nc.L1::no_allocate.v4.f32 {%0,%1,%2,%3}, [%4];"
               : "=f"(r.x), "=f"(r.y), "=f"(r.z), "=f"(r.w) : "l"(p));
  return r;
}
__device__ __forceinline__ uint4 ldg_stream_u4(const uint4* p) {
  uint4 r;
  asm volatile("ld.global.nc.L1::no_allocate.v4.u32 {%0,%1,%2,%3}, [%4];"
               : "=r"(r.x), "=r"(r.y), "=r"(r.z), "=r"(r.w) : "l"(p));
  return r;
}

// 256-bit streaming load (sm_100): 8 floats, no L1 allocation, L2 evict-first -- the dense tensor is read once
struct f32x8 { uint32_t u[8]; };
__device__ __forceinline__ f32x8 ldg_stream_f8(const float* p) {
  f32x8 r;
  asm volatile("ld.global.nc.L1::no_allocate.L2::evict_first.v8.b32 {%0,%1,%2,%3,%4,%5,%6,%7}, [%8];"
               : "=r"(r.u[0]), "=r"(r.u[1]), "=r"(r.u[2]), "=r"(r.u[3]), "=r"(r.u[4]), "=r"(r.u[5]), "=r"(r.u[6]), "=r"(r.u[7])
               : "l"(p));
  return r;
}

// 8 floats per thread -> one byte; 4 adjacent lanes -> one 32-bit word of the bitmask.  n8 = number of
// 8-float groups (a multiple of 4: whole words only).  32-byte aligned input.  Measured 6.7-6.8 TB/s against
// 6.2 TB/s for the 128-bit kernel below (scripts/micro/bitmask_sweep.cu).
__global__ void __launch_bounds__(P1_THREADS)
bitmask_f32_v8_kernel(const float* __restrict__ adj, int64_t n8, uint32_t* __restrict__ bits) {
  const int lane = threadIdx.x & 31;
  const int64_t stride = (int64_t)gridDim.x * P1_THREADS;
  const int64_t iters = (n8 + stride * P1_UNROLL - 1) / (stride * P1_UNROLL);
  int64_t i0 = (int64_t)blockIdx.x * P1_THREADS + threadIdx.x;
  for (int64_t it = 0; it < iters; ++it, i0 += stride * P1_UNROLL) {
    f32x8 x[P1_UNROLL];
#pragma unroll
    for (int u = 0; u < P1_UNROLL; ++u) {
      const int64_t i = i0 + u * stride;
      if (i < n8) x[u] = ldg_stream_f8(adj + i * 8);
      else {
#pragma unroll
        for (int q = 0; q < 8; ++q) x[u].u[q] = 0u;
      }
    }
#pragma unroll
    for (int u = 0; u < P1_UNROLL; ++u) {
      const int64_t i = i0 + u * stride;
      uint32_t by = 0;
#pragma unroll
      for (int q = 0; q < 8; ++q) by |= (__uint_as_float(x[u].u[q]) != 0.f ? 1u : 0u) << q;   // -0.0 is zero, NaN is not
      uint32_t w = by << ((lane & 3) * 8);
      w |= __shfl_xor_sync(0xffffffffu, w, 1);
      w |= __shfl_xor_sync(0xffffffffu, w, 2);
      if ((lane & 3) == 0 && i < n8) bits[i >> 2] = w;
    }
  }
}

// 4 floats per thread -> nibble; 8 adjacent lanes -> one 32-bit word of the bitmask.
// n4 = number of float4 groups (total elements / 4); words = n4 / 8 rounded up.
__global__ void __launch_bounds__(P1_THREADS)
bitmask_f32_kernel(const float4* __restrict__ adj, int64_t n4, uint32_t* __restrict__ bits) {
  const int lane = threadIdx.x & 31;
  const int64_t stride = (int64_t)gridDim.x * P1_THREADS;
  // every warp iterates the same number of times so the shuffles stay converged
  const int64_t iters = (n4 + stride * P1_UNROLL - 1) / (stride * P1_UNROLL);
  int64_t i0 = (int64_t)blockIdx.x * P1_THREADS + threadIdx.x;
  for (int64_t it = 0; it < iters; ++it, i0 += stride * P1_UNROLL) {
    float4 x[P1_UNROLL];
#pragma unroll
    for (int u = 0; u < P1_UNROLL; ++u) {
      int64_t i = i0 + u * stride;
      x[u] = i < n4 ? ldg_stream_f4(adj + i) : make_float4(0.f, 0.f, 0.f, 0.f);
    }
#pragma unroll
    for (int u = 0; u < P1_UNROLL; ++u) {
      int64_t i = i0 + u * stride;
      uint32_t nib = (x[u].x != 0.f ? 1u : 0u) | (x[u].y != 0.f ? 2u : 0u) |
                     (x[u].z != 0.f ? 4u : 0u) | (x[u].w != 0.f ? 8u : 0u);
      uint32_t w = nib << ((lane & 7) * 4);
      w |= __shfl_xor_sync(0xffffffffu, w, 1);
      w |= __shfl_xor_sync(0xffffffffu, w, 2);
      w |= __shfl_xor_sync(0xffffffffu, w, 4);
      if ((lane & 7) == 0 && i < n4) bits[i >> 3] = w;
    }
  }
}

// uint8 adjacency, 256-bit streaming loads: 32 bytes per thread -> one whole word of the bitmask, no shuffles.
// n32 = number of words.  32-byte aligned input.
__global__ void __launch_bounds__(P1_THREADS)
bitmask_u8_v8_kernel(const uint8_t* __restrict__ adj, int64_t n32, uint32_t* __restrict__ bits) {
  const int64_t stride = (int64_t)gridDim.x * P1_THREADS;
  const int64_t iters = (n32 + stride * P1_UNROLL - 1) / (stride * P1_UNROLL);
  int64_t i0 = (int64_t)blockIdx.x * P1_THREADS + threadIdx.x;
  for (int64_t it = 0; it < iters; ++it, i0 += stride * P1_UNROLL) {
    f32x8 x[P1_UNROLL];
#pragma unroll
    for (int u = 0; u < P1_UNROLL; ++u) {
      const int64_t i = i0 + u * stride;
      if (i < n32) x[u] = ldg_stream_f8(reinterpret_cast<const float*>(adj + i * 32));
      else {
#pragma unroll
        for (int q = 0; q < 8; ++q) x[u].u[q] = 0u;
      }
    }
#pragma unroll
    for (int u = 0; u < P1_UNROLL; ++u) {
      const int64_t i = i0 + u * stride;
      uint32_t m = 0;
#pragma unroll
      for (int q = 0; q < 8; ++q) {
#pragma unroll
        for (int by = 0; by < 4; ++by)
          m |= ((x[u].u[q] >> (8 * by)) & 0xffu) ? (1u << (q * 4 + by)) : 0u;
      }
      if (i < n32) bits[i] = m;
    }
  }
}

// 16 bytes per thread -> 16 bits; 2 adjacent lanes -> one word.  n16 = total / 16.
__global__ void __launch_bounds__(P1_THREADS)
bitmask_u8_kernel(const uint4* __restrict__ adj, int64_t n16, uint32_t* __restrict__ bits) {
  const int lane = threadIdx.x & 31;
  const int64_t stride = (int64_t)gridDim.x * P1_THREADS;
  const int64_t iters = (n16 + stride * P1_UNROLL - 1) / (stride * P1_UNROLL);
  int64_t i0 = (int64_t)blockIdx.x * P1_THREADS + threadIdx.x;
  for (int64_t it = 0; it < iters; ++it, i0 += stride * P1_UNROLL) {
    uint4 x[P1_UNROLL];
#pragma unroll
    for (int u = 0; u < P1_UNROLL; ++u) {
      int64_t i = i0 + u * stride;
      x[u] = i < n16 ? ldg_stream_u4(adj + i) : make_uint4(0, 0, 0, 0);
    }
#pragma unroll
    for (int u = 0; u < P1_UNROLL; ++u) {
      int64_t i = i0 + u * stride;
      uint32_t wd[4] = {x[u].x, x[u].y, x[u].z, x[u].w};
      uint32_t m = 0;
#pragma unroll
      for (int q = 0; q < 4; ++q) {
#pragma unroll
        for (int by = 0; by < 4; ++by)
          m |= ((wd[q] >> (8 * by)) & 0xffu) ? (1u << (q * 4 + by)) : 0u;
      }
      uint32_t w = m << ((lane & 1) * 16);
      w |= __shfl_xor_sync(0xffffffffu, w, 1);
      if ((lane & 1) == 0 && i < n16) bits[i >> 1] = w;
    }
  }
}

// ---- the same pass as a WORK QUEUE (ggnn_dense_prefetch) ---------------------------------------------------
// The bitmask pass is pure HBM streaming; the propagation kernels of the PREVIOUS batch are latency-bound and leave
// most of the HBM bandwidth idle.  So the pass is also available as slices of 256 KB handed out by an atomic
// counter to two kinds of CTAs:
//   background  launched on the ctx's side stream while the previous batch's msg / GRU kernels run: at most ONE
//               live CTA per SM (a per-SM claim; surplus CTAs exit at once) with a footprint (256 threads x 40
//               registers, no dynamic shared memory) that fits NEXT TO one CTA of those persistent kernels;
//   drain       launched on the caller's stream by the propagate call that consumes the prefetch, at full occupancy:
//               takes whatever slices are left.
// Nobody waits on anybody (the queue is a counter), so the two launches need not run concurrently.
constexpr int64_t Q_SLICE = 12288;   // work items (32-byte groups) per slice = 384 KB: whole iterations for 256- and 384-thread CTAs

// ctl[0] = next slice, ctl[8 + smid] = background CTAs that have claimed SM smid
template <bool U8>
__global__ void __maxnreg__(40)      // (not combinable with __launch_bounds__; always launched with P1_THREADS)
bitmask_queue_kernel(const void* __restrict__ adj_v, int64_t n_items, uint32_t* __restrict__ bits,
                     unsigned* __restrict__ ctl, int one_per_sm, int l2pf) {
  __shared__ unsigned s_next;
  const int lane = threadIdx.x & 31;
  if (threadIdx.x == 0) {
    unsigned claim = 0;
    if (one_per_sm) {
      unsigned smid;
      asm volatile("mov.u32 %0, %%smid;" : "=r"(smid));
      claim = atomicAdd(ctl + 8 + (smid & 255u), 1u);
    }
    s_next = claim ? 0xffffffffu : atomicAdd(ctl, 1u);
  }
  __syncthreads();
  unsigned cur = s_next;
  __syncthreads();
  const int64_t n_slices = (n_items + Q_SLICE - 1) / Q_SLICE;
  const float* adj = reinterpret_cast<const float*>(adj_v);
  while ((int64_t)cur < n_slices) {
    unsigned nxt = 0;
    if (threadIdx.x == 0) nxt = atomicAdd(ctl, 1u);          // in flight while this slice streams
    const int64_t g0 = (int64_t)cur * Q_SLICE + threadIdx.x;
    const int nthr = blockDim.x;                               // 256 or 384: Q_SLICE is a multiple of 4 * nthr
#pragma unroll 1
    for (int off = 0; off < (int)Q_SLICE; off += nthr * P1_UNROLL) {
      // One piece (nthr x 4 loads) AHEAD of the CTA's own loads, thread 0 pulls the next piece -- of this slice, then the
      // first of the next one -- into L2 (no registers, no shared memory).  A CTA of this footprint holds only 32-48 KB in
      // flight; with its loads hitting L2 that is enough for the full HBM rate (scripts/micro/bg_stream.cu: 384 threads
      // 6.1 -> 6.9 TB/s, 256 threads 4.9 -> 6.1; farther ahead than one piece is erratic).
      if (l2pf && threadIdx.x == 0) {
        const int noff = off + nthr * P1_UNROLL;
        const int64_t it0 = noff < (int)Q_SLICE ? (int64_t)cur * Q_SLICE + noff : (int64_t)nxt * Q_SLICE;
        if (it0 < n_items) {
          const int64_t left = n_items - it0;
          const uint32_t bytes = (uint32_t)(left < nthr * P1_UNROLL ? left : nthr * P1_UNROLL) * 32u;
          asm volatile("cp.async.bulk.prefetch.L2.global [%0], %1;" ::"l"(reinterpret_cast<const char*>(adj_v) + it0 * 32), "r"(bytes) : "memory");
        }
      }
      const int64_t i0 = g0 + off;
      f32x8 x[P1_UNROLL];
#pragma unroll
      for (int u = 0; u < P1_UNROLL; ++u) {
        const int64_t i = i0 + u * nthr;
        if (i < n_items) x[u] = ldg_stream_f8(adj + i * 8);
        else {
#pragma unroll
          for (int q = 0; q < 8; ++q) x[u].u[q] = 0u;
        }
      }
#pragma unroll
      for (int u = 0; u < P1_UNROLL; ++u) {
        const int64_t i = i0 + u * nthr;
        if (U8) {                       // 32 bytes -> one whole word
          uint32_t m = 0;
#pragma unroll
          for (int q = 0; q < 8; ++q) {
#pragma unroll
            for (int by = 0; by < 4; ++by) m |= ((x[u].u[q] >> (8 * by)) & 0xffu) ? (1u << (q * 4 + by)) : 0u;
          }
          if (i < n_items) bits[i] = m;
        } else {                        // 8 floats -> one byte; 4 adjacent lanes -> one word
          uint32_t by = 0;
#pragma unroll
          for (int q = 0; q < 8; ++q) by |= (__uint_as_float(x[u].u[q]) != 0.f ? 1u : 0u) << q;
          uint32_t w = by << ((lane & 3) * 8);
          w |= __shfl_xor_sync(0xffffffffu, w, 1);
          w |= __shfl_xor_sync(0xffffffffu, w, 2);
          if ((lane & 3) == 0 && i < n_items) bits[i >> 2] = w;
        }
      }
    }
    if (threadIdx.x == 0) s_next = nxt;
    __syncthreads();
    cur = s_next;
    __syncthreads();
  }
}

// scalar tail / unaligned fallback: one thread per bitmask word over [first_word, words)
template <typename T>
__global__ void bitmask_scalar_kernel(const T* __restrict__ adj, int64_t total,
                                      int64_t first_word, int64_t words,
                                      uint32_t* __restrict__ bits) {
  int64_t w = first_word + (int64_t)blockIdx.x * blockDim.x + threadIdx.x;
  if (w >= words) return;
  uint32_t m = 0;
  for (int j = 0; j < 32; ++j) {
    int64_t i = w * 32 + j;
    if (i < total && adj[i] != (T)0) m |= 1u << j;
  }
  bits[w] = m;
}

// popcount of bits [lo, hi) of the flat bitmask
__device__ __forceinline__ int popc_range(const uint32_t* __restrict__ bits, int64_t lo, int64_t hi) {
  int c = 0;
  for (int64_t w = lo >> 5; w <= (hi - 1) >> 5; ++w) {
    uint32_t m = __ldg(bits + w);
    int64_t wb = w << 5;
    if (lo > wb) m &= 0xffffffffu << (int)(lo - wb);
    if (hi < wb + 32) m &= 0xffffffffu >> (int)(wb + 32 - hi);
    c += __popc(m);
  }
  return c;
}

__global__ void rowcount_kernel(const uint32_t* __restrict__ bits, int64_t rows, int v,
                                int32_t* __restrict__ rowptr) {
  int64_t r = (int64_t)blockIdx.x * blockDim.x + threadIdx.x;
  if (r > rows) return;
  if (r == rows) { rowptr[r] = 0; return; }
  int c;
  if (v == 128) {           // aligned fast path: 4 words per row, one 128-bit load
    uint4 m = __ldg(reinterpret_cast<const uint4*>(bits) + r);
    c = __popc(m.x) + __popc(m.y) + __popc(m.z) + __popc(m.w);
  } else {
    c = popc_range(bits, r * v, (r + 1) * v);
  }
  rowptr[r] = c;
}

__global__ void fill_kernel(const uint32_t* __restrict__ bits, int64_t rows, int v,
                            const int32_t* __restrict__ rowptr, int32_t* __restrict__ col,
                            int64_t cap) {
  int64_t r = (int64_t)blockIdx.x * blockDim.x + threadIdx.x;
  if (r >= rows) return;
  int start = rowptr[r];
  int cnt = rowptr[r + 1] - start;
  if (cnt == 0) return;
  const int64_t lo = r * v, hi = lo + v;
  int64_t pos = start;
  for (int64_t w = lo >> 5; w <= (hi - 1) >> 5; ++w) {
    uint32_t m = __ldg(bits + w);
    int64_t wb = w << 5;
    if (lo > wb) m &= 0xffffffffu << (int)(lo - wb);
    if (hi < wb + 32) m &= 0xffffffffu >> (int)(wb + 32 - hi);
    while (m) {
      int bpos = __ffs(m) - 1;
      m &= m - 1;
      if (pos < cap) col[pos] = (int32_t)(wb + bpos - lo);
      ++pos;
    }
  }
}

}  // namespace

// tmp layout: [bitmask words][scan tmp]
size_t ggnn_compact_tmp_bytes(int64_t total_elems, int64_t rows) {
  size_t words = (size_t)ceil_div64(total_elems, 32);
  return align_up(words * 4 + 16, 256) + ggnn_scan_tmp_elems(rows + 1) * 4 + 256;
}

// P1: dense adjacency -> bitmask (the HBM-bound pass)
int ggnn_compact_bitmask(ggnn_ctx* ctx, const void* adj, int adj_dtype, int64_t total, uint32_t* bits,
                         cudaStream_t st) {
  const int64_t words = ceil_div64(total, 32);
  if (total > 0) {
    prof_scope ps(ctx, GGNN_STAGE_COMPACT_MASK, st);
    const bool aligned = (reinterpret_cast<uintptr_t>(adj) & 15u) == 0;
    int64_t vec_words = 0;   // words produced by the vector kernel
    if (adj_dtype == GGNN_ADJ_F32) {
      if (aligned) {
        vec_words = total / 32;                       // whole words only
        int64_t n4 = vec_words * 8;
        if (n4 > 0 && (reinterpret_cast<uintptr_t>(adj) & 31u) == 0) {      // 256-bit loads
          const int64_t n8 = vec_words * 4;
          int64_t want = ceil_div64(n8, (int64_t)P1_THREADS * P1_UNROLL);
          int grid = (int)(want < (int64_t)ctx->sm_count * 64 ? want : (int64_t)ctx->sm_count * 64);
          bitmask_f32_v8_kernel<<<grid, P1_THREADS, 0, st>>>(reinterpret_cast<const float*>(adj), n8, bits);
          GGNN_LAUNCHED(ctx);
        } else if (n4 > 0) {
          int64_t want = ceil_div64(n4, (int64_t)P1_THREADS * P1_UNROLL);
          int grid = (int)(want < (int64_t)ctx->sm_count * 16 ? want : (int64_t)ctx->sm_count * 16);
          bitmask_f32_kernel<<<grid, P1_THREADS, 0, st>>>(
              reinterpret_cast<const float4*>(adj), n4, bits);
          GGNN_LAUNCHED(ctx);
        }
      }
      if (vec_words < words) {
        int64_t n = words - vec_words;
        bitmask_scalar_kernel<float><<<(unsigned)ceil_div64(n, 256), 256, 0, st>>>(
            reinterpret_cast<const float*>(adj), total, vec_words, words, bits);
        GGNN_LAUNCHED(ctx);
      }
    } else {
      if (aligned) {
        vec_words = total / 32;
        int64_t n16 = vec_words * 2;
        if (n16 > 0 && (reinterpret_cast<uintptr_t>(adj) & 31u) == 0) {      // 256-bit loads
          int64_t want = ceil_div64(vec_words, (int64_t)P1_THREADS * P1_UNROLL);
          int grid = (int)(want < (int64_t)ctx->sm_count * 64 ? want : (int64_t)ctx->sm_count * 64);
          bitmask_u8_v8_kernel<<<grid, P1_THREADS, 0, st>>>(reinterpret_cast<const uint8_t*>(adj), vec_words, bits);
          GGNN_LAUNCHED(ctx);
        } else if (n16 > 0) {
          int64_t want = ceil_div64(n16, (int64_t)P1_THREADS * P1_UNROLL);
          int grid = (int)(want < (int64_t)ctx->sm_count * 16 ? want : (int64_t)ctx->sm_count * 16);
          bitmask_u8_kernel<<<grid, P1_THREADS, 0, st>>>(
              reinterpret_cast<const uint4*>(adj), n16, bits);
          GGNN_LAUNCHED(ctx);
        }
      }
      if (vec_words < words) {
        int64_t n = words - vec_words;
        bitmask_scalar_kernel<uint8_t><<<(unsigned)ceil_div64(n, 256), 256, 0, st>>>(
            reinterpret_cast<const uint8_t*>(adj), total, vec_words, words, bits);
        GGNN_LAUNCHED(ctx);
      }
    }
  }
  return GGNN_OK;
}

// work-queue form: only for 32-byte-aligned inputs made of whole words (anything else takes the plain pass)
bool ggnn_bitmask_queue_supported(const void* adj, int64_t total) {
  return total > 0 && total % 32 == 0 && (reinterpret_cast<uintptr_t>(adj) & 31u) == 0;
}
size_t ggnn_bitmask_queue_ctl_bytes() { return (8 + 256) * sizeof(unsigned); }

// background == true: persistent one-CTA-per-SM launch (side stream); false: the drain at full occupancy
int ggnn_bitmask_queue_launch(ggnn_ctx* ctx, const void* adj, int adj_dtype, int64_t total, uint32_t* bits,
                              unsigned* ctl, bool background, cudaStream_t st) {
  const int64_t n_items = adj_dtype == GGNN_ADJ_F32 ? total / 8 : total / 32;
  const int64_t n_slices = ceil_div64(n_items, Q_SLICE);
  if (!(ctx->attr_mask & (1u << GGNN_ATTR_BMQ))) {
    // the same shared-memory configuration as the persistent kernels it has to sit next to
    GGNN_CUDA(ctx, cudaFuncSetAttribute(bitmask_queue_kernel<false>, cudaFuncAttributePreferredSharedMemoryCarveout, 100));
    GGNN_CUDA(ctx, cudaFuncSetAttribute(bitmask_queue_kernel<true>, cudaFuncAttributePreferredSharedMemoryCarveout, 100));
    ctx->attr_mask |= 1u << GGNN_ATTR_BMQ;
  }
  int64_t grid = background ? (int64_t)ctx->sm_count * 4 : (int64_t)ctx->sm_count * 8;
  if (!background && grid > n_slices) grid = n_slices;
  if (grid < 1) grid = 1;
  // background CTA: 12 warps x 32 x 40 registers = 3 warps per scheduler partition, what the 16-warp GRU kernel leaves free
  static const int bg_threads = getenv("GGNN_BMQ_THREADS") ? atoi(getenv("GGNN_BMQ_THREADS")) : 384;
  const int threads = background && (bg_threads == 256 || bg_threads == 384) ? bg_threads : P1_THREADS;
  static const int l2pf_env = getenv("GGNN_BMQ_L2PF") ? atoi(getenv("GGNN_BMQ_L2PF")) : 1;    // A-B switch: 0 off, 1 background, 2 drain too
  const int l2pf = background ? l2pf_env >= 1 : l2pf_env >= 2;
  if (adj_dtype == GGNN_ADJ_F32)
    bitmask_queue_kernel<false><<<(unsigned)grid, threads, 0, st>>>(adj, n_items, bits, ctl, background ? 1 : 0, l2pf);
  else
    bitmask_queue_kernel<true><<<(unsigned)grid, threads, 0, st>>>(adj, n_items, bits, ctl, background ? 1 : 0, l2pf);
  GGNN_LAUNCHED(ctx);
  return GGNN_OK;
}

// P1..P3: bitmask, row counts, scan.  rowptr is final afterwards, *nnz_out_dev holds nnz.
int ggnn_compact_count(ggnn_ctx* ctx, const void* adj, int adj_dtype, int64_t rows, int v,
                       int32_t* rowptr, int64_t* nnz_out_dev, char* tmp, cudaStream_t st) {
  const int64_t total = rows * (int64_t)v;
  const int64_t words = ceil_div64(total, 32);
  uint32_t* bits = reinterpret_cast<uint32_t*>(tmp);
  int32_t* scan_tmp = reinterpret_cast<int32_t*>(tmp + align_up((size_t)words * 4 + 16, 256));
  const size_t scan_elems = ggnn_scan_tmp_elems(rows + 1);
  GGNN_TRY(ggnn_compact_bitmask(ctx, adj, adj_dtype, total, bits, st));
  prof_scope ps(ctx, GGNN_STAGE_COMPACT_REST, st);
  rowcount_kernel<<<(unsigned)ceil_div64(rows + 1, 256), 256, 0, st>>>(bits, rows, v, rowptr);
  GGNN_LAUNCHED(ctx);
  GGNN_TRY(ggnn_exclusive_scan_i32(ctx, rowptr, rowptr, rows + 1, nnz_out_dev, scan_tmp,
                                   scan_elems, st));
  return GGNN_OK;
}

// P4: needs the bitmask left in tmp by ggnn_compact_count.
int ggnn_compact_fill(ggnn_ctx* ctx, int64_t rows, int v, const int32_t* rowptr, int32_t* col,
                      int64_t col_capacity, const char* tmp, cudaStream_t st) {
  const uint32_t* bits = reinterpret_cast<const uint32_t*>(tmp);
  prof_scope ps(ctx, GGNN_STAGE_COMPACT_REST, st);
  if (rows > 0) {
    fill_kernel<<<(unsigned)ceil_div64(rows, 256), 256, 0, st>>>(bits, rows, v, rowptr, col,
                                                                col_capacity);
    GGNN_LAUNCHED(ctx);
  }
  return GGNN_OK;
}

// ---------------------------------------------------------------------------------------------------------------
// ggnn_compact_adjacency_static: the same compaction with its scratch (bitmask + scan partials) in a file-scope
// device array instead of the ctx's cudaMalloc-ed arena, for callers that must not allocate (CUDA graph capture,
// tight serving loops).  No counterpart in the reference.
// Limits: b*two_e*v*v <= GGNN_STATIC_MAX_ELEMS adjacency elements and b*two_e*v <= GGNN_STATIC_MAX_ROWS rows.
// Every GPU that loads the library holds the array (2 MB + 4 KB) until the process exits.
// ---------------------------------------------------------------------------------------------------------------
namespace {

constexpr int64_t STATIC_MAX_ELEMS = GGNN_STATIC_MAX_ELEMS;       // 16 777 216 -> 2 MB of bitmask
constexpr int64_t STATIC_MAX_ROWS = GGNN_STATIC_MAX_ROWS;         //  1 048 576 -> 1.5 KB of scan partials
constexpr size_t STATIC_SCRATCH_BYTES = (size_t)(STATIC_MAX_ELEMS / 32) * 4 + 4096;

static __device__ __align__(256) unsigned char g_static_scratch[STATIC_SCRATCH_BYTES];

// In host code `g_static_scratch` is only a stand-in for the symbol, NOT a device address: the address each device
// gives the array is asked for once per device, with that device current, and handed to the kernels as a pointer.
std::mutex g_scratch_mu;
char* g_scratch_ptr[64] = {nullptr};

}  // namespace

extern "C" int ggnn_compact_adjacency_static(ggnn_ctx* ctx, const void* adj, int adj_dtype, int b, int two_e, int v,
                                             int32_t* rowptr, int32_t* col, int64_t col_capacity,
                                             int64_t* nnz_out_dev, void* stream) {
  if (!ctx) return GGNN_E_ARG;
  if (b < 0 || two_e <= 0 || v <= 0) GGNN_FAIL(ctx, GGNN_E_ARG, "compact_static: bad sizes b=%d two_e=%d v=%d", b, two_e, v);
  if ((b > 0 && !adj) || !rowptr || !nnz_out_dev || (col_capacity > 0 && !col) || col_capacity < 0)
    GGNN_FAIL(ctx, GGNN_E_ARG, "compact_static: null pointer");
  if (adj_dtype != GGNN_ADJ_F32 && adj_dtype != GGNN_ADJ_U8) GGNN_FAIL(ctx, GGNN_E_ARG, "compact_static: bad adj_dtype %d", adj_dtype);
  const int64_t rows = (int64_t)b * two_e * v;
  const int64_t total = rows * (int64_t)v;
  if (total > STATIC_MAX_ELEMS || rows > STATIC_MAX_ROWS || ggnn_compact_tmp_bytes(total, rows) > STATIC_SCRATCH_BYTES)
    GGNN_FAIL(ctx, GGNN_E_UNSUPPORTED,
              "compact_static: %lld adjacency elements / %lld rows exceed the static scratch (at most %lld elements and "
              "%lld rows per call); split the batch or use ggnn_compact_adjacency",
              (long long)total, (long long)rows, (long long)STATIC_MAX_ELEMS, (long long)STATIC_MAX_ROWS);
  if (ctx->device < 0 || ctx->device >= 64) GGNN_FAIL(ctx, GGNN_E_UNSUPPORTED, "compact_static: device index %d", ctx->device);
  int prev = -1;
  GGNN_CUDA(ctx, cudaGetDevice(&prev));
  if (prev != ctx->device) GGNN_CUDA(ctx, cudaSetDevice(ctx->device));
  int rc = GGNN_OK;
  char* tmp = nullptr;
  {
    std::lock_guard<std::mutex> lk(g_scratch_mu);
    if (!g_scratch_ptr[ctx->device]) {            // set-up, once per device, that device current
      void* ptr = nullptr;
      const cudaError_t e = cudaGetSymbolAddress(&ptr, g_static_scratch);
      if (e != cudaSuccess || !ptr) {
        snprintf(ctx->err, sizeof(ctx->err), "compact_static: cudaGetSymbolAddress failed: %s", cudaGetErrorString(e));
        rc = GGNN_E_CUDA;
      } else {
        g_scratch_ptr[ctx->device] = static_cast<char*>(ptr);
      }
    }
    tmp = g_scratch_ptr[ctx->device];
  }
  if (rc == GGNN_OK) {
    // Nothing here allocates: kernel launches and one cudaMemsetAsync.  (The per-stage profiler creates CUDA events on
    // first use, so it is held off for the duration of the call.)  Every scratch word this call reads it has written
    // itself first -- the bitmask pass covers every word, the scan writes its partials before reading them -- so the
    // previous call's contents need no clearing.
    const bool prof = ctx->prof_on;
    ctx->prof_on = false;
    cudaStream_t st = (cudaStream_t)stream;
    rc = ggnn_compact_count(ctx, adj, adj_dtype, rows, v, rowptr, nnz_out_dev, tmp, st);
    if (rc == GGNN_OK) rc = ggnn_compact_fill(ctx, rows, v, rowptr, col, col_capacity, tmp, st);
    ctx->prof_on = prof;
  }
  if (prev != ctx->device) cudaSetDevice(prev);
  return rc;
}
