// plan_edges.cu -- build the propagation plan from an edge list (A9, the "sparse" form).
//
// edges int32 [nnz, 3] = (src, channel, dst), any order, duplicates allowed (each counts).
// Two layouts for the message rows:
//   edge mode   (nnz <  N*2E): one message row per edge, grouped by channel;
//   dense mode  (nnz >= N*2E): one message row per (channel, node) pair -- every node is
//                transformed once per channel and edges only index the result
//                (min(nnz, N*2E) row-products).
// Slot assignment inside a channel uses integer atomics; that only permutes rows of the
// message buffer, never the values.  The per-target lists are then sorted by (channel, src),
// so the FP32 summation order -- and therefore the result -- is deterministic.
#include "plan.cuh"

namespace {

__global__ void edge_hist_kernel(const int32_t* __restrict__ edges, int64_t nnz, int64_t N,
                                 int two_e, int32_t* __restrict__ chcount,
                                 int32_t* __restrict__ tcnt, int64_t* __restrict__ bad) {
  int64_t k = (int64_t)blockIdx.x * blockDim.x + threadIdx.x;
  if (k >= nnz) return;
  int s = edges[3 * k], c = edges[3 * k + 1], d = edges[3 * k + 2];
  if (s < 0 || s >= N || d < 0 || d >= N || c < 0 || c >= two_e) {
    atomicAdd(reinterpret_cast<unsigned long long*>(bad), 1ull);
    return;
  }
  if (chcount) atomicAdd(chcount + c, 1);
  atomicAdd(tcnt + d, 1);
}

__global__ void tiles_from_counts_kernel(const int32_t* __restrict__ chcount, int two_e,
                                         int32_t* __restrict__ cursor,
                                         int32_t* __restrict__ tile_type,
                                         int32_t* __restrict__ n_tiles, int64_t max_tiles) {
  extern __shared__ int32_t sh[];
  if (threadIdx.x == 0) {
    int32_t run = 0;
    for (int e = 0; e < two_e; ++e) {
      sh[e] = run;
      run += (chcount[e] + GGNN_TILE_M - 1) / GGNN_TILE_M;
    }
    sh[two_e] = run;
    *n_tiles = run;
  }
  __syncthreads();
  for (int e = threadIdx.x; e < two_e; e += blockDim.x) cursor[e] = sh[e] * GGNN_TILE_M;
  for (int e = 0; e < two_e; ++e)
    for (int t = sh[e] + threadIdx.x; t < sh[e + 1]; t += blockDim.x)
      if (t < max_tiles) tile_type[t] = e;
}

// dense mode: message row (c, i) = c*Np + i
__global__ void dense_rows_kernel(int64_t N, int64_t Np, int two_e, int32_t* __restrict__ msg_src,
                                  int32_t* __restrict__ tile_type, int32_t* __restrict__ n_tiles) {
  int64_t r = (int64_t)blockIdx.x * blockDim.x + threadIdx.x;
  const int64_t total = Np * two_e;
  if (r == 0) *n_tiles = (int32_t)(total / GGNN_TILE_M);
  if (r >= total) return;
  int64_t i = r % Np;
  msg_src[r] = i < N ? (int32_t)i : -1;
  if ((r % GGNN_TILE_M) == 0) tile_type[r / GGNN_TILE_M] = (int32_t)(r / Np);
}

__global__ void edge_assign_kernel(const int32_t* __restrict__ edges, int64_t nnz, int64_t N,
                                   int two_e, int64_t Np_dense /* 0 = edge mode */,
                                   int32_t* __restrict__ cursor, int32_t* __restrict__ msg_src,
                                   const int32_t* __restrict__ tptr, int32_t* __restrict__ tcur,
                                   unsigned long long* __restrict__ tkey,
                                   int32_t* __restrict__ tval) {
  int64_t k = (int64_t)blockIdx.x * blockDim.x + threadIdx.x;
  if (k >= nnz) return;
  int s = edges[3 * k], c = edges[3 * k + 1], d = edges[3 * k + 2];
  if (s < 0 || s >= N || d < 0 || d >= N || c < 0 || c >= two_e) return;
  int32_t mrow;
  if (Np_dense) {
    mrow = (int32_t)((int64_t)c * Np_dense + s);
  } else {
    mrow = atomicAdd(cursor + c, 1);
    msg_src[mrow] = s;
  }
  int pos = tptr[d] + atomicAdd(tcur + d, 1);
  tkey[pos] = ((unsigned long long)(unsigned)c << 32) | (unsigned)s;
  tval[pos] = mrow;
}

// Sorting a node's (key, val) segment by (channel, source).  Slots inside a segment were handed out with atomicAdd, so
// edges with EQUAL keys (duplicate (src, channel, dst) triples) end up in arbitrary relative order -- harmless: their
// message values are identical, so the sum does not depend on it.
// Two kernels: nodes with at most HUB_DEG incoming edges (every node of the BASELINE configurations) are sorted by one
// thread each (insertion sort, O(d^2)); a hub -- e.g. a ROOT node shared by a packed batch, 1e5 in-edges -- would keep a
// single thread busy for seconds that way, so hubs are left to target_sort_hub_kernel: a whole CTA per hub, bitonic
// network in shared memory (d <= HUB_SMEM) or in place in global memory (L2) above that, O(d log^2 d / threads).
constexpr int HUB_DEG = 64;
constexpr int HUB_SMEM = 2048;
constexpr int HUB_THREADS = 512;

__global__ void target_sort_kernel(const int32_t* __restrict__ tptr, int64_t N,
                                   unsigned long long* __restrict__ tkey,
                                   int32_t* __restrict__ tval) {
  int64_t n = (int64_t)blockIdx.x * blockDim.x + threadIdx.x;
  if (n >= N) return;
  const int j0 = tptr[n], j1 = tptr[n + 1];
  if (j1 - j0 > HUB_DEG) return;                      // target_sort_hub_kernel
  for (int i = j0 + 1; i < j1; ++i) {
    unsigned long long k = tkey[i];
    int32_t v = tval[i];
    int j = i - 1;
    while (j >= j0 && tkey[j] > k) {
      tkey[j + 1] = tkey[j];
      tval[j + 1] = tval[j];
      --j;
    }
    tkey[j + 1] = k;
    tval[j + 1] = v;
  }
}

// Bitonic network in its "flip + disperse" form: every compare-exchange moves the smaller key to the LOWER index, so a
// length that is not a power of two needs no padding -- the missing elements behave as +inf at the top and never move.
template <typename K, typename V>
__device__ __forceinline__ void bitonic_block(K* key, V* val, int n) {
  for (int k = 2; (k >> 1) < n; k <<= 1) {
    for (int i = threadIdx.x; i < n; i += blockDim.x) {          // flip: partner inside the block of k elements
      const int j = i ^ (k - 1);
      if (j > i && j < n) {
        const K a = key[i], b = key[j];
        if (a > b) { key[i] = b; key[j] = a; const V t = val[i]; val[i] = val[j]; val[j] = t; }
      }
    }
    __syncthreads();
    for (int d = k >> 2; d > 0; d >>= 1) {                        // disperse
      for (int i = threadIdx.x; i < n; i += blockDim.x) {
        const int j = i | d;
        if ((i & d) == 0 && j < n) {
          const K a = key[i], b = key[j];
          if (a > b) { key[i] = b; key[j] = a; const V t = val[i]; val[i] = val[j]; val[j] = t; }
        }
      }
      __syncthreads();
    }
  }
}

__global__ void __launch_bounds__(HUB_THREADS)
target_sort_hub_kernel(const int32_t* __restrict__ tptr, int64_t N, unsigned long long* tkey, int32_t* tval) {
  __shared__ unsigned long long s_key[HUB_SMEM];
  __shared__ int32_t s_val[HUB_SMEM];
  __shared__ int s_hub[HUB_THREADS];
  __shared__ int s_nhub;
  for (int64_t span = (int64_t)blockIdx.x * HUB_THREADS; span < N; span += (int64_t)gridDim.x * HUB_THREADS) {
    if (threadIdx.x == 0) s_nhub = 0;
    __syncthreads();
    const int64_t n = span + threadIdx.x;
    if (n < N && tptr[n + 1] - tptr[n] > HUB_DEG) s_hub[atomicAdd(&s_nhub, 1)] = (int)threadIdx.x;
    __syncthreads();
    const int nhub = s_nhub;
    for (int q = 0; q < nhub; ++q) {                              // (list order is arbitrary; every hub is independent)
      const int64_t node = span + s_hub[q];
      const int j0 = tptr[node], d = tptr[node + 1] - j0;
      if (d <= HUB_SMEM) {
        for (int i = threadIdx.x; i < d; i += blockDim.x) { s_key[i] = tkey[j0 + i]; s_val[i] = tval[j0 + i]; }
        __syncthreads();
        bitonic_block(s_key, s_val, d);
        for (int i = threadIdx.x; i < d; i += blockDim.x) { tkey[j0 + i] = s_key[i]; tval[j0 + i] = s_val[i]; }
        __syncthreads();
      } else {
        bitonic_block(tkey + j0, tval + j0, d);                   // in place; __syncthreads orders the CTA's global accesses
      }
    }
    __syncthreads();                                              // s_hub / s_nhub are rewritten by the next span
  }
}

}  // namespace

int ggnn_plan_from_edges(ggnn_ctx* ctx, const int32_t* edges, int64_t nnz, int64_t N, int two_e,
                         ggnn_plan* plan, ws_alloc& a, bool dry, cudaStream_t st) {
  const bool dense_mode = nnz >= N * (int64_t)two_e;
  const int64_t Np = (int64_t)align_up((size_t)N, GGNN_TILE_M);
  plan->N = N;
  plan->nnz = nnz;
  plan->dense_rows = dense_mode;
  plan->max_tiles = dense_mode ? Np / GGNN_TILE_M * two_e : nnz / GGNN_TILE_M + two_e + 1;
  if (plan->max_tiles * GGNN_TILE_M >= 0x7fffffffLL || nnz >= 0x7fffffffLL)
    GGNN_FAIL(ctx, GGNN_E_UNSUPPORTED, "edge plan too large for int32 row ids; split the batch");
  plan->n_tiles = a.take<int32_t>(4);
  plan->msg_src = a.take<int32_t>((size_t)plan->max_tiles * GGNN_TILE_M);
  plan->tile_type = a.take<int32_t>((size_t)plan->max_tiles);
  plan->tptr = a.take<int32_t>((size_t)N + 1);
  plan->tmsg = a.take<int32_t>((size_t)nnz + 1);
  int32_t* chcount = a.take<int32_t>((size_t)two_e + 1);
  int32_t* cursor = a.take<int32_t>((size_t)two_e + 1);
  int32_t* tcur = a.take<int32_t>((size_t)N + 1);
  unsigned long long* tkey = a.take<unsigned long long>((size_t)nnz + 1);
  const size_t scan_elems = ggnn_scan_tmp_elems(N + 1);
  int32_t* scan_tmp = a.take<int32_t>(scan_elems);
  if (dry) return GGNN_OK;
  prof_scope ps(ctx, GGNN_STAGE_PLAN, st);

  int64_t* bad = ctx->d_mail + 2;
  GGNN_CUDA(ctx, cudaMemsetAsync(bad, 0, sizeof(int64_t), st));
  GGNN_CUDA(ctx, cudaMemsetAsync(chcount, 0, ((size_t)two_e + 1) * 4, st));
  GGNN_CUDA(ctx, cudaMemsetAsync(plan->tptr, 0, ((size_t)N + 1) * 4, st));
  GGNN_CUDA(ctx, cudaMemsetAsync(tcur, 0, ((size_t)N + 1) * 4, st));
  const unsigned eg = (unsigned)ceil_div64(nnz > 0 ? nnz : 1, 256);
  if (nnz > 0) {
    edge_hist_kernel<<<eg, 256, 0, st>>>(edges, nnz, N, two_e, dense_mode ? nullptr : chcount,
                                         plan->tptr, bad);
    GGNN_LAUNCHED(ctx);
  }
  GGNN_TRY(ggnn_exclusive_scan_i32(ctx, plan->tptr, plan->tptr, N + 1, nullptr, scan_tmp, scan_elems, st));
  if (dense_mode) {
    dense_rows_kernel<<<(unsigned)ceil_div64(Np * two_e, 256), 256, 0, st>>>(
        N, Np, two_e, plan->msg_src, plan->tile_type, plan->n_tiles);
    GGNN_LAUNCHED(ctx);
  } else {
    GGNN_CUDA(ctx, cudaMemsetAsync(plan->msg_src, 0xFF, (size_t)plan->max_tiles * GGNN_TILE_M * 4, st));
    tiles_from_counts_kernel<<<1, 256, (two_e + 1) * sizeof(int32_t), st>>>(
        chcount, two_e, cursor, plan->tile_type, plan->n_tiles, plan->max_tiles);
    GGNN_LAUNCHED(ctx);
  }
  if (nnz > 0) {
    edge_assign_kernel<<<eg, 256, 0, st>>>(edges, nnz, N, two_e, dense_mode ? Np : 0, cursor,
                                           plan->msg_src, plan->tptr, tcur, tkey, plan->tmsg);
    GGNN_LAUNCHED(ctx);
    target_sort_kernel<<<(unsigned)ceil_div64(N, 128), 128, 0, st>>>(plan->tptr, N, tkey, plan->tmsg);
    GGNN_LAUNCHED(ctx);
    if (nnz > HUB_DEG) {                // a node with more than HUB_DEG incoming edges can exist
      const int64_t spans = ceil_div64(N, HUB_THREADS);
      const unsigned hg = (unsigned)(spans < 2 * (int64_t)ctx->sm_count ? spans : 2 * (int64_t)ctx->sm_count);
      target_sort_hub_kernel<<<hg, HUB_THREADS, 0, st>>>(plan->tptr, N, tkey, plan->tmsg);
      GGNN_LAUNCHED(ctx);
    }
  }
  return GGNN_OK;
}
