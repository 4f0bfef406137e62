// scan.cu -- device-wide exclusive prefix sum over int32 (reduce / scan-sums / downsweep).
// Used for rowptr (compaction) and for the plan offsets.  Deterministic, no spinning.
#include "common.cuh"

namespace {

constexpr int SCAN_THREADS = 512;
constexpr int SCAN_ITEMS = 8;
constexpr int SCAN_TILE = SCAN_THREADS * SCAN_ITEMS;  // 4096

__device__ __forceinline__ int warp_incl_scan(int x) {
  const int lane = threadIdx.x & 31;
#pragma unroll
  for (int d = 1; d < 32; d <<= 1) {
    int y = __shfl_up_sync(0xffffffffu, x, d);
    if (lane >= d) x += y;
  }
  return x;
}

// block-wide exclusive scan of one value per thread; returns exclusive prefix, total in *total
__device__ __forceinline__ int block_excl_scan(int x, int* total) {
  __shared__ int warp_sums[SCAN_THREADS / 32];
  const int lane = threadIdx.x & 31, wid = threadIdx.x >> 5;
  int incl = warp_incl_scan(x);
  if (lane == 31) warp_sums[wid] = incl;
  __syncthreads();
  if (wid == 0) {
    int s = lane < SCAN_THREADS / 32 ? warp_sums[lane] : 0;
    s = warp_incl_scan(s);
    if (lane < SCAN_THREADS / 32) warp_sums[lane] = s;
  }
  __syncthreads();
  int base = wid ? warp_sums[wid - 1] : 0;
  *total = warp_sums[SCAN_THREADS / 32 - 1];
  __syncthreads();
  return base + incl - x;
}

__global__ void __launch_bounds__(SCAN_THREADS)
tile_reduce_kernel(const int32_t* __restrict__ in, int64_t n, int32_t* __restrict__ sums,
                   unsigned long long* __restrict__ total64) {
  const int64_t base = (int64_t)blockIdx.x * SCAN_TILE;
  int acc = 0;
#pragma unroll
  for (int i = 0; i < SCAN_ITEMS; ++i) {
    int64_t idx = base + (int64_t)i * SCAN_THREADS + threadIdx.x;
    if (idx < n) acc += in[idx];
  }
  int total;
  block_excl_scan(acc, &total);
  if (threadIdx.x == 0) {
    sums[blockIdx.x] = total;
    // the grand total is carried in 64 bits (integer atomics: order-independent), so a sum >= 2^31 is
    // REPORTED instead of wrapping silently; the int32 prefix array is invalid in that case and callers
    // check the total before using it
    if (total64) atomicAdd(total64, (unsigned long long)(unsigned)total);
  }
}

// offsets == nullptr: single tile (n <= SCAN_TILE).  total_out (int64) optional.
__global__ void __launch_bounds__(SCAN_THREADS)
tile_scan_kernel(const int32_t* in, int32_t* out, int64_t n,   // in may alias out
                 const int32_t* __restrict__ offsets, int64_t* __restrict__ total_out) {
  const int64_t base = (int64_t)blockIdx.x * SCAN_TILE + (int64_t)threadIdx.x * SCAN_ITEMS;
  int v[SCAN_ITEMS];
  int acc = 0;
#pragma unroll
  for (int i = 0; i < SCAN_ITEMS; ++i) {
    int64_t idx = base + i;
    v[i] = idx < n ? in[idx] : 0;
    acc += v[i];
  }
  int total;
  int excl = block_excl_scan(acc, &total);
  int off = offsets ? offsets[blockIdx.x] : 0;
  int run = off + excl;
#pragma unroll
  for (int i = 0; i < SCAN_ITEMS; ++i) {
    int64_t idx = base + i;
    if (idx < n) out[idx] = run;
    run += v[i];
  }
  if (total_out && blockIdx.x == gridDim.x - 1 && threadIdx.x == 0)
    *total_out = (int64_t)off + (int64_t)total;
}

}  // namespace

size_t ggnn_scan_tmp_elems(int64_t n) {
  size_t tot = 0;
  while (n > SCAN_TILE) {
    n = ceil_div64(n, SCAN_TILE);
    tot += align_up((size_t)n, 64);
  }
  return tot + 64;
}

int ggnn_exclusive_scan_i32(ggnn_ctx* ctx, const int32_t* in, int32_t* out, int64_t n,
                            int64_t* total_out_dev, int32_t* tmp, size_t tmp_elems,
                            cudaStream_t st) {
  if (n <= 0) {
    if (total_out_dev) GGNN_CUDA(ctx, cudaMemsetAsync(total_out_dev, 0, sizeof(int64_t), st));
    return GGNN_OK;
  }
  if (n <= SCAN_TILE) {
    tile_scan_kernel<<<1, SCAN_THREADS, 0, st>>>(in, out, n, nullptr, total_out_dev);
    GGNN_LAUNCHED(ctx);
    return GGNN_OK;
  }
  const int64_t tiles = ceil_div64(n, SCAN_TILE);
  if (tiles > 0x7fffffffLL) GGNN_FAIL(ctx, GGNN_E_UNSUPPORTED, "scan too large");
  if (tmp_elems < align_up((size_t)tiles, 64)) GGNN_FAIL(ctx, GGNN_E_CAPACITY, "scan tmp too small");
  int32_t* sums = tmp;
  if (total_out_dev) GGNN_CUDA(ctx, cudaMemsetAsync(total_out_dev, 0, sizeof(int64_t), st));
  tile_reduce_kernel<<<(unsigned)tiles, SCAN_THREADS, 0, st>>>(in, n, sums, reinterpret_cast<unsigned long long*>(total_out_dev));
  GGNN_LAUNCHED(ctx);
  GGNN_TRY(ggnn_exclusive_scan_i32(ctx, sums, sums, tiles, nullptr,
                                   tmp + align_up((size_t)tiles, 64),
                                   tmp_elems - align_up((size_t)tiles, 64), st));
  tile_scan_kernel<<<(unsigned)tiles, SCAN_THREADS, 0, st>>>(in, out, n, sums, nullptr);
  GGNN_LAUNCHED(ctx);
  return GGNN_OK;
}
