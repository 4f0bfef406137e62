// plan.cu -- build the propagation plan from the canonical CSR (dense path).
//
// Message rows are the edges re-ordered channel-major across the whole batch (so one GEMM
// tile of 128 rows multiplies by a single W_e), each channel padded to a multiple of 128.
// The target CSR lists, per node, the message rows it sums, in ascending (channel, source)
// order -- the same order the reference's [b, v, 2E*v] x [b, 2E*v, h] matmul sums them
// (chem_tensorflow_dense.py:414-428), so FP32 summation order is reproducible.
#include "plan.cuh"

namespace {

// seglenT[e*b + g] = number of edges of channel e in graph g; seglenT[2E*b] = 0
__global__ void seglen_kernel(const int32_t* __restrict__ rowptr, int64_t b, int two_e, int v,
                              int32_t* __restrict__ seglenT) {
  int64_t i = (int64_t)blockIdx.x * blockDim.x + threadIdx.x;
  const int64_t n = b * two_e;
  if (i > n) return;
  if (i == n) { seglenT[i] = 0; return; }
  int64_t e = i / b, g = i % b;
  int64_t r0 = (g * two_e + e) * v;
  seglenT[i] = rowptr[r0 + v] - rowptr[r0];
}

// one block: per-channel tile counts -> tile_base[e] (in tiles), tile_type[], n_tiles
__global__ void tiles_kernel(const int32_t* __restrict__ segposT, int64_t b, int two_e,
                             int32_t* __restrict__ tile_base, int32_t* __restrict__ tile_type,
                             int32_t* __restrict__ n_tiles, int64_t max_tiles) {
  extern __shared__ int32_t sh[];   // [two_e + 1]
  if (threadIdx.x == 0) {
    int32_t run = 0;
    for (int e = 0; e < two_e; ++e) {
      int32_t cnt = segposT[(int64_t)(e + 1) * b] - segposT[(int64_t)e * b];
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

// one thread per canonical row (g, e, t): place its edges into message rows
__global__ void place_kernel(const int32_t* __restrict__ rowptr, const int32_t* __restrict__ col,
                             int64_t b, int two_e, int v, const int32_t* __restrict__ segposT,
                             const int32_t* __restrict__ tile_base,
                             int32_t* __restrict__ msg_src, int32_t* __restrict__ edge_mrow) {
  int64_t r = (int64_t)blockIdx.x * blockDim.x + threadIdx.x;
  if (r >= b * two_e * v) return;
  int k0 = rowptr[r], k1 = rowptr[r + 1];
  if (k0 == k1) return;
  int64_t g = r / ((int64_t)two_e * v);
  int e = (int)((r / v) % two_e);
  int seg0 = rowptr[(g * two_e + e) * v];
  int32_t base = tile_base[e] * GGNN_TILE_M + (segposT[(int64_t)e * b + g] - segposT[(int64_t)e * b]);
  for (int k = k0; k < k1; ++k) {
    int32_t mrow = base + (k - seg0);
    msg_src[mrow] = (int32_t)(g * v + col[k]);
    edge_mrow[k] = mrow;
  }
}

// one thread per node (g, t): in-degree over all channels; tcnt[N] = 0
__global__ void tcount_kernel(const int32_t* __restrict__ rowptr, int64_t b, int two_e, int v,
                              int32_t* __restrict__ tcnt) {
  int64_t n = (int64_t)blockIdx.x * blockDim.x + threadIdx.x;
  const int64_t N = b * v;
  if (n > N) return;
  if (n == N) { tcnt[n] = 0; return; }
  int64_t g = n / v, t = n % v;
  int c = 0;
  for (int e = 0; e < two_e; ++e) {
    int64_t r = (g * two_e + e) * v + t;
    c += rowptr[r + 1] - rowptr[r];
  }
  tcnt[n] = c;
}

__global__ void tfill_kernel(const int32_t* __restrict__ rowptr, int64_t b, int two_e, int v,
                             const int32_t* __restrict__ tptr,
                             const int32_t* __restrict__ edge_mrow, int32_t* __restrict__ tmsg) {
  int64_t n = (int64_t)blockIdx.x * blockDim.x + threadIdx.x;
  if (n >= b * v) return;
  int pos = tptr[n];
  if (tptr[n + 1] == pos) return;
  int64_t g = n / v, t = n % v;
  for (int e = 0; e < two_e; ++e) {
    int64_t r = (g * two_e + e) * v + t;
    for (int k = rowptr[r]; k < rowptr[r + 1]; ++k) tmsg[pos++] = edge_mrow[k];
  }
}

}  // namespace

int ggnn_plan_from_csr(ggnn_ctx* ctx, const int32_t* rowptr, const int32_t* col, int64_t b,
                       int two_e, int v, int64_t nnz, ggnn_plan* plan, ws_alloc& a, bool dry,
                       cudaStream_t st) {
  const int64_t N = b * v;
  plan->N = N;
  plan->nnz = nnz;
  plan->max_tiles = nnz / GGNN_TILE_M + two_e + 1;
  plan->n_tiles = a.take<int32_t>(4);
  plan->msg_src = a.take<int32_t>((size_t)plan->max_tiles * GGNN_TILE_M);
  plan->tile_type = a.take<int32_t>((size_t)plan->max_tiles);
  plan->tptr = a.take<int32_t>((size_t)N + 1);
  plan->tmsg = a.take<int32_t>((size_t)nnz + 1);
  // builder temporaries
  const int64_t nseg = b * two_e;
  int32_t* segposT = a.take<int32_t>((size_t)nseg + 1);
  int32_t* tile_base = a.take<int32_t>((size_t)two_e + 1);
  int32_t* edge_mrow = a.take<int32_t>((size_t)nnz + 1);
  const size_t scan_elems = ggnn_scan_tmp_elems((nseg > N ? nseg : N) + 1);
  int32_t* scan_tmp = a.take<int32_t>(scan_elems);
  if (dry) return GGNN_OK;
  prof_scope ps(ctx, GGNN_STAGE_PLAN, st);

  seglen_kernel<<<(unsigned)ceil_div64(nseg + 1, 256), 256, 0, st>>>(rowptr, b, two_e, v, segposT);
  GGNN_LAUNCHED(ctx);
  GGNN_TRY(ggnn_exclusive_scan_i32(ctx, segposT, segposT, nseg + 1, nullptr, scan_tmp, scan_elems, st));
  tiles_kernel<<<1, 256, (two_e + 1) * sizeof(int32_t), st>>>(segposT, b, two_e, tile_base,
                                                             plan->tile_type, plan->n_tiles,
                                                             plan->max_tiles);
  GGNN_LAUNCHED(ctx);
  GGNN_CUDA(ctx, cudaMemsetAsync(plan->msg_src, 0xFF,
                                 (size_t)plan->max_tiles * GGNN_TILE_M * sizeof(int32_t), st));
  const int64_t rows = b * two_e * v;
  place_kernel<<<(unsigned)ceil_div64(rows, 256), 256, 0, st>>>(rowptr, col, b, two_e, v, segposT,
                                                               tile_base, plan->msg_src, edge_mrow);
  GGNN_LAUNCHED(ctx);
  tcount_kernel<<<(unsigned)ceil_div64(N + 1, 256), 256, 0, st>>>(rowptr, b, two_e, v, plan->tptr);
  GGNN_LAUNCHED(ctx);
  GGNN_TRY(ggnn_exclusive_scan_i32(ctx, plan->tptr, plan->tptr, N + 1, nullptr, scan_tmp, scan_elems, st));
  tfill_kernel<<<(unsigned)ceil_div64(N, 256), 256, 0, st>>>(rowptr, b, two_e, v, plan->tptr,
                                                            edge_mrow, plan->tmsg);
  GGNN_LAUNCHED(ctx);
  return GGNN_OK;
}
