// step_fp32.cu -- strict-FP32 timestep: FFMA SGEMMs with gather / concat A operands and
// fused epilogues, plus the segmented gather-sum.  This is GGNN_PREC_FP32: every product
// and sum is IEEE FP32 (no TF32), so it tracks the FP64 oracle to ~5e-7 norm-wise.
//
//   MSG    M[r,:]  = h[msg_src[r],:] @ W[type(tile)] + B[type]          (chem_tensorflow_dense.py:397-412)
//   AGG    acts[n] = sum_{j in tptr[n]..tptr[n+1]} M[tmsg[j],:]         (:414-428, as a sparse sum)
//   GATES  g = sigmoid([acts, h] @ Wg + bg); rh = g[:, :h] * h; u = g[:, h:]     (TF GRUCell)
//   CAND   c = tanh([acts, rh] @ Wc + bc);  h' = u*h + (1-u)*c
#include "plan.cuh"
#include <cuda_fp16.h>
#include "tc.cuh"
#include "heads.cuh"

namespace {

constexpr int BM = 128, BN = 128, BK = 8, NT = 256;

enum { MODE_MSG = 0, MODE_GATES = 1, MODE_CAND = 2, MODE_PLAIN = 3 };

struct sgemm_params {
  const float* A1;        // [*, lda] first K1 columns of A
  const float* A2;        // [*, lda] next K2 columns (row-aligned with the output rows) or null
  const int32_t* gather;  // row index into A1 per output row (MSG) or null
  int K1, K2, lda;
  const float* Bmat;      // [K, N] row-major (per type for MSG)
  const float* bias;      // [N] (per type for MSG) or null
  int N;
  int64_t M;              // rows
  const int32_t* tile_type;  // MSG
  const int32_t* n_tiles;    // MSG: device-side tile count
  // outputs / epilogue operands
  float* out0;            // MSG: msg [M, N];  GATES: rh [M, h];  CAND: h state [M, h] (in/out)
  float* out1;            // GATES: u [M, h]
  const float* hstate;    // GATES: h [M, h]
  const float* ubuf;      // CAND: u [M, h]
  float* out_r;           // GATES, training tape: r [M, h] (or null)
  float* out_c;           // CAND, training tape: c [M, h] (or null)
  int h;
};

__device__ __forceinline__ float sigmoid_acc(float x) { return 1.0f / (1.0f + expf(-x)); }

template <int MODE>
__global__ void __launch_bounds__(NT)
sgemm_kernel(const sgemm_params p) {
  __shared__ __align__(16) float As[2][BK][BM + 4];
  __shared__ __align__(16) float Bs[2][BK][BN];

  const int tid = threadIdx.x;
  // row tiles ride on gridDim.x (limit 2^31-1), column tiles on gridDim.y: a 65 536 x 128-node batch has 65 536 row tiles
  const int64_t row0 = (int64_t)blockIdx.x * BM;
  const int col0 = blockIdx.y * BN;
  const float* Bmat = p.Bmat;
  const float* bias = p.bias;
  if (MODE == MODE_MSG) {
    if ((int)blockIdx.x >= *p.n_tiles) return;
    const int type = p.tile_type[blockIdx.x];
    Bmat += (size_t)type * p.K1 * p.N;
    if (bias) bias += (size_t)type * p.N;
  }
  const int K = p.K1 + p.K2;

  // ---- A loader: thread -> (row = tid/2, 4 consecutive k at (tid%2)*4)
  const int a_row = tid >> 1;
  const int a_k4 = (tid & 1) * 4;
  int64_t a_grow = row0 + a_row;          // output row
  int64_t a_src1 = -1;                    // row in A1
  bool a_valid = a_grow < p.M;
  if (a_valid) {
    if (MODE == MODE_MSG) {
      int s = p.gather[a_grow];
      a_src1 = s;                         // -1 => padding row
    } else {
      a_src1 = a_grow;
    }
  }
  // ---- B loader: thread -> (k = tid/32, 4 consecutive n at (tid%32)*4)
  const int b_k = tid >> 5;
  const int b_n4 = (tid & 31) * 4;

  auto load_a = [&](int k0) -> float4 {
    int k = k0 + a_k4;
    float4 v = make_float4(0.f, 0.f, 0.f, 0.f);
    if (k < K) {
      if (k < p.K1) {
        if (a_src1 >= 0) v = *reinterpret_cast<const float4*>(p.A1 + a_src1 * p.lda + k);
      } else if (a_valid) {
        v = *reinterpret_cast<const float4*>(p.A2 + a_grow * p.lda + (k - p.K1));
      }
    }
    return v;
  };
  auto load_b = [&](int k0) -> float4 {
    int k = k0 + b_k, n = col0 + b_n4;
    float4 v = make_float4(0.f, 0.f, 0.f, 0.f);
    if (k < K && n < p.N) v = __ldg(reinterpret_cast<const float4*>(Bmat + (size_t)k * p.N + n));
    return v;
  };
  auto store_a = [&](int buf, float4 v) {
    As[buf][a_k4 + 0][a_row] = v.x;
    As[buf][a_k4 + 1][a_row] = v.y;
    As[buf][a_k4 + 2][a_row] = v.z;
    As[buf][a_k4 + 3][a_row] = v.w;
  };
  auto store_b = [&](int buf, float4 v) {
    *reinterpret_cast<float4*>(&Bs[buf][b_k][b_n4]) = v;
  };

  const int tx = tid & 15, ty = tid >> 4;   // 16 x 16 threads, each 8x8 = (4+4) x (4+4)
  float acc[8][8];
#pragma unroll
  for (int i = 0; i < 8; ++i)
#pragma unroll
    for (int j = 0; j < 8; ++j) acc[i][j] = 0.f;

  float4 ra = load_a(0), rb = load_b(0);
  store_a(0, ra);
  store_b(0, rb);
  __syncthreads();
  const int ktiles = (K + BK - 1) / BK;
  for (int kt = 0; kt < ktiles; ++kt) {
    const int buf = kt & 1;
    if (kt + 1 < ktiles) {
      ra = load_a((kt + 1) * BK);
      rb = load_b((kt + 1) * BK);
    }
#pragma unroll
    for (int kk = 0; kk < BK; ++kk) {
      float4 a0 = *reinterpret_cast<const float4*>(&As[buf][kk][ty * 4]);
      float4 a1 = *reinterpret_cast<const float4*>(&As[buf][kk][64 + ty * 4]);
      float4 b0 = *reinterpret_cast<const float4*>(&Bs[buf][kk][tx * 4]);
      float4 b1 = *reinterpret_cast<const float4*>(&Bs[buf][kk][64 + tx * 4]);
      float a[8] = {a0.x, a0.y, a0.z, a0.w, a1.x, a1.y, a1.z, a1.w};
      float b[8] = {b0.x, b0.y, b0.z, b0.w, b1.x, b1.y, b1.z, b1.w};
#pragma unroll
      for (int i = 0; i < 8; ++i)
#pragma unroll
        for (int j = 0; j < 8; ++j) acc[i][j] = fmaf(a[i], b[j], acc[i][j]);
    }
    if (kt + 1 < ktiles) {
      store_a(buf ^ 1, ra);
      store_b(buf ^ 1, rb);
    }
    __syncthreads();
  }

  // ---- epilogue
#pragma unroll
  for (int i = 0; i < 8; ++i) {
    const int64_t row = row0 + (i < 4 ? ty * 4 + i : 64 + ty * 4 + (i - 4));
    if (row >= p.M) continue;
    if (MODE == MODE_MSG && p.gather[row] < 0) continue;
#pragma unroll
    for (int jh = 0; jh < 2; ++jh) {
      const int n = col0 + (jh ? 64 + tx * 4 : tx * 4);
      if (n >= p.N) continue;
      float v[4] = {acc[i][jh * 4 + 0], acc[i][jh * 4 + 1], acc[i][jh * 4 + 2], acc[i][jh * 4 + 3]};
      if (bias) {
        float4 bb = __ldg(reinterpret_cast<const float4*>(bias + n));
        v[0] += bb.x; v[1] += bb.y; v[2] += bb.z; v[3] += bb.w;
      }
      if (MODE == MODE_MSG || MODE == MODE_PLAIN) {
        *reinterpret_cast<float4*>(p.out0 + row * p.N + n) = make_float4(v[0], v[1], v[2], v[3]);
      } else if (MODE == MODE_GATES) {
        float g[4];
#pragma unroll
        for (int q = 0; q < 4; ++q) g[q] = sigmoid_acc(v[q]);
        if (n < p.h) {   // reset gate: rh = r * h     (h % 4 == 0, so a float4 never straddles)
          float4 hv = *reinterpret_cast<const float4*>(p.hstate + row * p.h + n);
          *reinterpret_cast<float4*>(p.out0 + row * p.h + n) =
              make_float4(g[0] * hv.x, g[1] * hv.y, g[2] * hv.z, g[3] * hv.w);
          if (p.out_r) *reinterpret_cast<float4*>(p.out_r + row * p.h + n) = make_float4(g[0], g[1], g[2], g[3]);
        } else {         // update gate
          *reinterpret_cast<float4*>(p.out1 + row * p.h + (n - p.h)) = make_float4(g[0], g[1], g[2], g[3]);
        }
      } else {           // MODE_CAND
        float4 hv = *reinterpret_cast<const float4*>(p.out0 + row * p.h + n);
        float4 uv = *reinterpret_cast<const float4*>(p.ubuf + row * p.h + n);
        float hh[4] = {hv.x, hv.y, hv.z, hv.w}, uu[4] = {uv.x, uv.y, uv.z, uv.w}, o[4], cv[4];
#pragma unroll
        for (int q = 0; q < 4; ++q) { cv[q] = tanhf(v[q]); o[q] = uu[q] * hh[q] + (1.0f - uu[q]) * cv[q]; }
        if (p.out_c) *reinterpret_cast<float4*>(p.out_c + row * p.h + n) = make_float4(cv[0], cv[1], cv[2], cv[3]);
        *reinterpret_cast<float4*>(p.out0 + row * p.h + n) = make_float4(o[0], o[1], o[2], o[3]);
      }
    }
  }
}

// Segmented gather-sum: one warp per target node, lanes across the feature dimension with
// 128-bit loads; messages are added in tmsg order (ascending channel, source) -- no atomics,
// so the result is bit-reproducible.
// one message row chunk (4 columns) as FP32, from FP32 or FP16 storage
__device__ __forceinline__ float4 ld_msg4(const float* msg, size_t row, int h, int c) {
  return __ldg(reinterpret_cast<const float4*>(msg + row * h) + c);
}
__device__ __forceinline__ float4 ld_msg4(const __half* msg, size_t row, int h, int c) {
  const uint2 v = __ldg(reinterpret_cast<const uint2*>(msg + row * h) + c);
  const float2 lo = __half22float2(*reinterpret_cast<const __half2*>(&v.x));
  const float2 hi = __half22float2(*reinterpret_cast<const __half2*>(&v.y));
  return make_float4(lo.x, lo.y, hi.x, hi.y);
}

template <typename MT>
__global__ void __launch_bounds__(256)
aggregate_kernel(const int32_t* __restrict__ tptr, const int32_t* __restrict__ tmsg,
                 const MT* __restrict__ msg, float* __restrict__ acts, int64_t N, int h,
                 int round_tf32) {
  const int lane = threadIdx.x & 31;
  const int64_t node = ((int64_t)blockIdx.x * blockDim.x + threadIdx.x) >> 5;
  if (node >= N) return;
  const int j0 = tptr[node], j1 = tptr[node + 1];
  const int h4 = h >> 2;
  for (int c = lane; c < h4; c += 32) {
    float4 acc = make_float4(0.f, 0.f, 0.f, 0.f);
    int j = j0;
    for (; j + 4 <= j1; j += 4) {
      int r0 = __ldg(tmsg + j), r1 = __ldg(tmsg + j + 1), r2 = __ldg(tmsg + j + 2), r3 = __ldg(tmsg + j + 3);
      float4 m0 = ld_msg4(msg, (size_t)r0, h, c);
      float4 m1 = ld_msg4(msg, (size_t)r1, h, c);
      float4 m2 = ld_msg4(msg, (size_t)r2, h, c);
      float4 m3 = ld_msg4(msg, (size_t)r3, h, c);
      acc.x += m0.x; acc.y += m0.y; acc.z += m0.z; acc.w += m0.w;
      acc.x += m1.x; acc.y += m1.y; acc.z += m1.z; acc.w += m1.w;
      acc.x += m2.x; acc.y += m2.y; acc.z += m2.z; acc.w += m2.w;
      acc.x += m3.x; acc.y += m3.y; acc.z += m3.z; acc.w += m3.w;
    }
    for (; j < j1; ++j) {
      int r0 = __ldg(tmsg + j);
      float4 m0 = ld_msg4(msg, (size_t)r0, h, c);
      acc.x += m0.x; acc.y += m0.y; acc.z += m0.z; acc.w += m0.w;
    }
    if (round_tf32) {   // acts only feeds tensor-core operands in TF32 mode: round once, here
      acc.x = tc::to_tf32(acc.x); acc.y = tc::to_tf32(acc.y);
      acc.z = tc::to_tf32(acc.z); acc.w = tc::to_tf32(acc.w);
    }
    reinterpret_cast<float4*>(acts + (size_t)node * h)[c] = acc;
  }
}

}  // namespace

int ggnn_aggregate(ggnn_ctx* ctx, const ggnn_plan& plan, const float* msg, float* acts, int h,
                   bool round_tf32, cudaStream_t st) {
  if (plan.N == 0) return GGNN_OK;
  prof_scope ps(ctx, GGNN_STAGE_AGG, st);
  const int64_t threads = plan.N * 32;
  aggregate_kernel<float><<<(unsigned)ceil_div64(threads, 256), 256, 0, st>>>(plan.tptr, plan.tmsg, msg,
                                                                             acts, plan.N, h, round_tf32 ? 1 : 0);
  GGNN_LAUNCHED(ctx);
  return GGNN_OK;
}

// same, message rows stored as FP16 with pitch h (TF32 mode, hidden sizes the fused kernels do not cover)
int ggnn_aggregate_f16(ggnn_ctx* ctx, const ggnn_plan& plan, const void* msg_half, float* acts, int h,
                       bool round_tf32, cudaStream_t st) {
  if (plan.N == 0) return GGNN_OK;
  prof_scope ps(ctx, GGNN_STAGE_AGG, st);
  const int64_t threads = plan.N * 32;
  aggregate_kernel<__half><<<(unsigned)ceil_div64(threads, 256), 256, 0, st>>>(
      plan.tptr, plan.tmsg, reinterpret_cast<const __half*>(msg_half), acts, plan.N, h, round_tf32 ? 1 : 0);
  GGNN_LAUNCHED(ctx);
  return GGNN_OK;
}

int ggnn_sgemm_concat(ggnn_ctx* ctx, const float* A1, const float* A2, int K1, int K2, int lda,
                      const float* B, const float* bias, int N, int64_t M, float* out, cudaStream_t st) {
  if (M == 0 || N == 0) return GGNN_OK;
  sgemm_params p{};
  p.A1 = A1; p.A2 = A2; p.gather = nullptr; p.K1 = K1; p.K2 = K2; p.lda = lda;
  p.Bmat = B; p.bias = bias; p.N = N; p.M = M; p.out0 = out; p.h = N;
  dim3 grid((unsigned)ceil_div64(M, BM), (N + BN - 1) / BN);
  sgemm_kernel<MODE_PLAIN><<<grid, NT, 0, st>>>(p);
  GGNN_LAUNCHED(ctx);
  return GGNN_OK;
}

int ggnn_step_fp32(ggnn_ctx* ctx, int slot, const ggnn_plan& plan, const ggnn_step_bufs& bufs,
                   float* h_state, cudaStream_t st) {
  const int h = ctx->h;
  const ggnn_edge_weights& ew = ctx->edge[slot];
  if (plan.N == 0) return GGNN_OK;
  // MSG
  if (plan.max_tiles > 0 && plan.nnz > 0) {
    prof_scope ps(ctx, GGNN_STAGE_MSG, st);
    sgemm_params p{};
    p.A1 = h_state; p.A2 = nullptr; p.gather = plan.msg_src;
    p.K1 = h; p.K2 = 0; p.lda = h;
    p.Bmat = bufs.W_override ? bufs.W_override : ew.W; p.bias = ew.B; p.N = h;
    p.M = plan.max_tiles * GGNN_TILE_M;
    p.tile_type = plan.tile_type; p.n_tiles = plan.n_tiles;
    p.out0 = bufs.msg; p.h = h;
    dim3 grid((unsigned)plan.max_tiles, (h + BN - 1) / BN);
    sgemm_kernel<MODE_MSG><<<grid, NT, 0, st>>>(p);
    GGNN_LAUNCHED(ctx);
  }
  GGNN_TRY(ggnn_aggregate(ctx, plan, bufs.msg, bufs.acts, h, false, st));
  const unsigned mt = (unsigned)ceil_div64(plan.N, BM);
  {  // GATES
    prof_scope ps(ctx, GGNN_STAGE_GRU_GATES, st);
    sgemm_params p{};
    p.A1 = bufs.acts; p.A2 = h_state; p.gather = nullptr;
    p.K1 = h; p.K2 = h; p.lda = h;
    p.Bmat = ctx->Wg; p.bias = ctx->bg; p.N = 2 * h; p.M = plan.N;
    p.out0 = bufs.rh; p.out1 = bufs.u; p.hstate = h_state; p.h = h; p.out_r = bufs.tape_r;
    dim3 grid(mt, (2 * h + BN - 1) / BN);
    sgemm_kernel<MODE_GATES><<<grid, NT, 0, st>>>(p);
    GGNN_LAUNCHED(ctx);
  }
  {  // CAND + state update (in place)
    prof_scope ps(ctx, GGNN_STAGE_GRU_CAND, st);
    sgemm_params p{};
    p.A1 = bufs.acts; p.A2 = bufs.rh; p.gather = nullptr;
    p.K1 = h; p.K2 = h; p.lda = h;
    p.Bmat = ctx->Wc; p.bias = ctx->bc; p.N = h; p.M = plan.N;
    p.out0 = h_state; p.ubuf = bufs.u; p.h = h; p.out_c = bufs.tape_c;
    dim3 grid(mt, (h + BN - 1) / BN);
    sgemm_kernel<MODE_CAND><<<grid, NT, 0, st>>>(p);
    GGNN_LAUNCHED(ctx);
  }
  return GGNN_OK;
}
