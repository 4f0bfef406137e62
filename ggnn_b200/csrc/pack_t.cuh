// pack_t.cuh -- the weight packs of the tensor-core paths as tiled copies (shared by tf32.cu and wide16.cu).
//
// The MMA operands are K-major: dst[e][n][k] = cvt(W[e][k][n]), zero padded from h to hp.  The first form of these packs
// was one thread per OUTPUT element, i.e. a 4 h-byte stride between the reads of neighbouring threads: 0.9 TB/s.  With
// edge-weight dropout (BASELINE config 5) the pack runs once per unrolled timestep in the forward and in the backward,
// on W * mask_t / keep -- the product is formed here ((w * m) * scale, the order ggnn_mask_scale uses, so the packed
// values are bit-identical to packing a materialised W_eff) instead of in a separate pass over a 59 MB scratch tensor.
#pragma once
#include <cuda_fp16.h>
#include <stdint.h>
#include "tc.cuh"

namespace packt {

enum { OUT_F16 = 0, OUT_TF32 = 1, OUT_TF32_LO = 2 };     // TF32_LO: the residual tf32(w - tf32(w)) of the 3xTF32 split

__device__ __forceinline__ float cvt_tf32(float w, int lo) {
  const float hi = tc::to_tf32(w);
  return lo ? tc::to_tf32(w - hi) : hi;
}

// One CTA = a [64 k x 32 n] tile of matrix blockIdx.z: 128-byte row reads, 128-byte (F16) / 2 x 128-byte (TF32) row writes.
// grid (hp / 32, ceil(hp / 64), matrices); 256 threads.  hp % 32 == 0; OUT_F16 needs hp % 2 == 0.
template <int OUT>
__global__ void __launch_bounds__(256)
edge_t_kernel(const float* __restrict__ W, const float* __restrict__ mask, float scale, void* __restrict__ dst_, int h, int hp) {
  __shared__ float tile[64][33];
  const int e = blockIdx.z, k0 = blockIdx.y * 64, n0 = blockIdx.x * 32;
  const int tx = threadIdx.x & 31, ty = threadIdx.x >> 5;
  const size_t mbase = (size_t)e * h * h;
#pragma unroll
  for (int r = ty; r < 64; r += 8) {
    const int k = k0 + r, n = n0 + tx;
    float v = 0.f;
    if (k < h && n < h) {
      const size_t i = mbase + (size_t)k * h + n;
      v = W[i];
      if (mask) v = v * mask[i] * scale;
    }
    tile[r][tx] = v;
  }
  __syncthreads();
  const size_t obase = (size_t)e * hp * hp;
#pragma unroll
  for (int r = ty; r < 32; r += 8) {
    const int n = n0 + r;                                  // < hp: hp % 32 == 0
    if (OUT == OUT_F16) {
      const int k = k0 + 2 * tx;
      if (k < hp) {
        const __half2 pr = __floats2half2_rn(tile[2 * tx][r], tile[2 * tx + 1][r]);
        *reinterpret_cast<__half2*>(reinterpret_cast<__half*>(dst_) + obase + (size_t)n * hp + k) = pr;
      }
    } else {
      float* dst = reinterpret_cast<float*>(dst_) + obase + (size_t)n * hp;
      if (k0 + tx < hp) dst[k0 + tx] = cvt_tf32(tile[tx][r], OUT == OUT_TF32_LO);
      if (k0 + 32 + tx < hp) dst[k0 + 32 + tx] = cvt_tf32(tile[tx + 32][r], OUT == OUT_TF32_LO);
    }
  }
}

template <int OUT>
inline void launch_edge_t(const float* W, const float* mask, float scale, void* dst, int two_e, int h, int hp, cudaStream_t st) {
  dim3 grid((unsigned)(hp / 32), (unsigned)((hp + 63) / 64), (unsigned)two_e);
  edge_t_kernel<OUT><<<grid, 256, 0, st>>>(W, mask, scale, dst, h, hp);
}

// Block copy without a transpose (the NT operands of the backward): four columns per thread.
// dst[b][rb*hp + r][cb*hp + c] = tf32(src[b][(rb*h + r) * src_ld + cb*h + c] (* mask[same] * scale)), zero padding.
// h % 4 == 0, hp % 4 == 0, src_ld % 4 == 0, 16-byte aligned bases.
static __global__ void __launch_bounds__(256)
blocks4_kernel(const float* __restrict__ src, const float* __restrict__ mask, float scale, int src_ld, float* __restrict__ dst,
               int nbatch, int rblocks, int cblocks, int h, int hp) {
  const int cols4 = cblocks * hp / 4, rows = rblocks * hp;
  const int64_t total = (int64_t)nbatch * rows * cols4;
  for (int64_t i = (int64_t)blockIdx.x * blockDim.x + threadIdx.x; i < total; i += (int64_t)gridDim.x * blockDim.x) {
    const int c4 = (int)(i % cols4);
    const int64_t t = i / cols4;
    const int rr = (int)(t % rows);
    const int64_t b = t / rows;
    const int cc = c4 * 4, rb = rr / hp, r = rr - rb * hp, cb = cc / hp, c = cc - cb * hp;
    float4 v = make_float4(0.f, 0.f, 0.f, 0.f);
    if (r < h && c < h) {
      const size_t idx = ((size_t)b * rblocks * h + (size_t)rb * h + r) * src_ld + (size_t)cb * h + c;
      v = *reinterpret_cast<const float4*>(src + idx);
      if (mask) {
        const float4 m = *reinterpret_cast<const float4*>(mask + idx);
        v.x = v.x * m.x * scale; v.y = v.y * m.y * scale; v.z = v.z * m.z * scale; v.w = v.w * m.w * scale;
      }
      v.x = tc::to_tf32(v.x); v.y = tc::to_tf32(v.y); v.z = tc::to_tf32(v.z); v.w = tc::to_tf32(v.w);
    }
    reinterpret_cast<float4*>(dst)[i] = v;
  }
}

}  // namespace packt
