// train_fp32.cu -- backward of the propagation path (BASELINE config 5), strict FP32.
//
// Forward (per step, tape = h, x, r, u, c):     reference chem_tensorflow_dense.py:391-437 + TF GRUCell
//   m_j = h[src_j] @ W_e(j) + b_e(j);  x[n] = sum_{j -> n} m_j
//   [r,u] = sigmoid([x,h] @ Wg + bg);  c = tanh([x, r*h] @ Wc + bc);  h' = u*h + (1-u)*c
// Backward (per step, in reverse), given dh' (gradient w.r.t. the new state):
//   dzc = dh' (1-u)(1-c^2)          dzu = dh' (h-c) u (1-u)            dh  = dh' u
//   [dx | drh] = dzc @ Wc^T         dzr = drh h r (1-r)                dh += drh r
//   [dx2 | dh2] = [dzr,dzu] @ Wg^T  dx += dx2                          dh += dh2
//   dWc += [x, r*h]^T dzc           dWg += [x, h]^T [dzr,dzu]          dbc, dbg += column sums
//   dm_j = dx[tgt_j]                dW_e += h[src_j]^T dm_j            db_e += dm_j
//   dh[src_j] += dm_j @ W_e^T
// Two FFMA GEMM kernels carry all contractions:
//   nt_kernel   C[M,N]   = A[M,K] @ B[N,K]^T   (A rows optionally gathered, two K segments;
//                                               store, or atomic scatter-add by a row map)
//   tn_kernel   C[K1,N] += X[R,K1]^T @ D[R,N]  (rows of X and D optionally gathered, grouped by
//                                               message tile type; split over row blocks, atomics)
#include "plan.cuh"
#include "train.cuh"
#include "tf32.cuh"
#include "tc.cuh"

namespace {

constexpr int BM = 128, BN = 128, BK = 8, NT = 256;

// ---------------------------------------------------------------------------------------
struct nt_params {
  const float* A1; const float* A2;     // K segments (A2 may be null); row-major, lda
  int K1, K2, lda;
  const int32_t* a_rowmap;              // row of A1 per output row (-1 = zero row) or null
  const float* B;                       // [N, K] row-major (per type when tile_type != null)
  int ldb; int64_t b_type_stride;
  int N; int64_t M;
  const int32_t* tile_type; const int32_t* n_tiles;   // grouped mode (one type per 128-row tile)
  float* C; int ldc;                    // store mode
  const int32_t* c_rowmap;              // atomic scatter mode: C row per output row (-1 = skip)
};

__global__ void __launch_bounds__(NT)
nt_kernel(const nt_params p) {
  __shared__ __align__(16) float As[2][BK][BM + 4];
  __shared__ __align__(16) float Bs[2][BK][BN + 4];
  const int tid = threadIdx.x;
  const int64_t row0 = (int64_t)blockIdx.x * BM;          // row tiles on gridDim.x (limit 2^31-1)
  const int col0 = blockIdx.y * BN;
  const float* Bm = p.B;
  if (p.tile_type) {
    if ((int)blockIdx.x >= *p.n_tiles) return;
    Bm += (size_t)p.tile_type[blockIdx.x] * p.b_type_stride;
  }
  const int K = p.K1 + p.K2;
  const int lrow = tid >> 1, lk4 = (tid & 1) * 4;        // both loaders: row = tid/2, 4 consecutive k
  const int64_t a_grow = row0 + lrow;
  int64_t a_src = -1;
  if (a_grow < p.M) a_src = p.a_rowmap ? (int64_t)p.a_rowmap[a_grow] : a_grow;
  const int b_n = col0 + lrow;

  auto load_a = [&](int k0) -> float4 {
    const int k = k0 + lk4;
    float4 v = make_float4(0.f, 0.f, 0.f, 0.f);
    if (k < K && a_src >= 0) {
      if (k < p.K1) v = *reinterpret_cast<const float4*>(p.A1 + a_src * p.lda + k);
      else v = *reinterpret_cast<const float4*>(p.A2 + a_src * p.lda + (k - p.K1));
    }
    return v;
  };
  auto load_b = [&](int k0) -> float4 {
    const int k = k0 + lk4;
    float4 v = make_float4(0.f, 0.f, 0.f, 0.f);
    if (k < K && b_n < p.N) v = __ldg(reinterpret_cast<const float4*>(Bm + (size_t)b_n * p.ldb + k));
    return v;
  };
  auto store_t = [&](float (*S)[BM + 4], float4 v) {
    S[lk4 + 0][lrow] = v.x; S[lk4 + 1][lrow] = v.y; S[lk4 + 2][lrow] = v.z; S[lk4 + 3][lrow] = v.w;
  };

  const int tx = tid & 15, ty = tid >> 4;
  float acc[8][8];
#pragma unroll
  for (int i = 0; i < 8; ++i)
#pragma unroll
    for (int j = 0; j < 8; ++j) acc[i][j] = 0.f;

  float4 ra = load_a(0), rb = load_b(0);
  store_t(As[0], ra); store_t(Bs[0], rb);
  __syncthreads();
  const int ktiles = (K + BK - 1) / BK;
  for (int kt = 0; kt < ktiles; ++kt) {
    const int buf = kt & 1;
    if (kt + 1 < ktiles) { ra = load_a((kt + 1) * BK); rb = load_b((kt + 1) * BK); }
#pragma unroll
    for (int kk = 0; kk < BK; ++kk) {
      const float4 a0 = *reinterpret_cast<const float4*>(&As[buf][kk][ty * 4]);
      const float4 a1 = *reinterpret_cast<const float4*>(&As[buf][kk][64 + ty * 4]);
      const float4 b0 = *reinterpret_cast<const float4*>(&Bs[buf][kk][tx * 4]);
      const float4 b1 = *reinterpret_cast<const float4*>(&Bs[buf][kk][64 + tx * 4]);
      const float a[8] = {a0.x, a0.y, a0.z, a0.w, a1.x, a1.y, a1.z, a1.w};
      const float b[8] = {b0.x, b0.y, b0.z, b0.w, b1.x, b1.y, b1.z, b1.w};
#pragma unroll
      for (int i = 0; i < 8; ++i)
#pragma unroll
        for (int j = 0; j < 8; ++j) acc[i][j] = fmaf(a[i], b[j], acc[i][j]);
    }
    if (kt + 1 < ktiles) { store_t(As[buf ^ 1], ra); store_t(Bs[buf ^ 1], rb); }
    __syncthreads();
  }

#pragma unroll
  for (int i = 0; i < 8; ++i) {
    const int64_t row = row0 + (i < 4 ? ty * 4 + i : 64 + ty * 4 + (i - 4));
    if (row >= p.M) continue;
    int64_t crow = row;
    if (p.c_rowmap) { crow = p.c_rowmap[row]; if (crow < 0) continue; }
#pragma unroll
    for (int jh = 0; jh < 2; ++jh) {
      const int n = col0 + (jh ? 64 + tx * 4 : tx * 4);
      if (n >= p.N) continue;                       // N % 4 == 0
      float* dst = p.C + crow * p.ldc + n;
      if (p.c_rowmap) {
#pragma unroll
        for (int q = 0; q < 4; ++q) atomicAdd(dst + q, acc[i][jh * 4 + q]);
      } else {
        *reinterpret_cast<float4*>(dst) = make_float4(acc[i][jh * 4 + 0], acc[i][jh * 4 + 1],
                                                      acc[i][jh * 4 + 2], acc[i][jh * 4 + 3]);
      }
    }
  }
}

// ---------------------------------------------------------------------------------------
struct tn_params {
  const float* X; int ldx; const int32_t* x_rowmap;   // [R, K1] (rows optionally gathered, -1 = zero)
  const float* D; int ldd; const int32_t* d_rowmap;   // [R, N]
  int K1, N; int64_t R;
  int rows_per_cta;                                   // multiple of 128
  float* C; int ldc;                                  // [K1, N] (+ type * c_type_stride in grouped mode)
  const int32_t* tile_type; const int32_t* n_tiles; int64_t c_type_stride;
};

__global__ void __launch_bounds__(NT)
tn_kernel(const tn_params p) {
  __shared__ __align__(16) float Xs[2][BK][BM];
  __shared__ __align__(16) float Ds[2][BK][BN];
  const int tid = threadIdx.x;
  const int m0 = blockIdx.y * BM, n0 = blockIdx.x * BN;
  int64_t r_begin = (int64_t)blockIdx.z * p.rows_per_cta;
  int64_t r_end = r_begin + p.rows_per_cta;
  int64_t R = p.R;
  if (p.tile_type) R = (int64_t)(*p.n_tiles) * GGNN_TILE_M;
  if (r_end > R) r_end = R;
  if (r_begin >= r_end) return;

  const int lk = tid >> 5, lc4 = (tid & 31) * 4;       // loaders: row k = tid/32, 4 consecutive columns
  auto load_x = [&](int64_t r0) -> float4 {
    const int64_t r = r0 + lk;
    float4 v = make_float4(0.f, 0.f, 0.f, 0.f);
    if (r < r_end && m0 + lc4 < p.K1) {
      const int64_t src = p.x_rowmap ? (int64_t)p.x_rowmap[r] : r;
      if (src >= 0) v = *reinterpret_cast<const float4*>(p.X + src * p.ldx + m0 + lc4);
    }
    return v;
  };
  auto load_d = [&](int64_t r0) -> float4 {
    const int64_t r = r0 + lk;
    float4 v = make_float4(0.f, 0.f, 0.f, 0.f);
    if (r < r_end && n0 + lc4 < p.N) {
      const int64_t src = p.d_rowmap ? (int64_t)p.d_rowmap[r] : r;
      if (src >= 0) v = *reinterpret_cast<const float4*>(p.D + src * p.ldd + n0 + lc4);
    }
    return v;
  };

  const int tx = tid & 15, ty = tid >> 4;
  float acc[8][8];
  auto zero = [&]() {
#pragma unroll
    for (int i = 0; i < 8; ++i)
#pragma unroll
      for (int j = 0; j < 8; ++j) acc[i][j] = 0.f;
  };
  auto flush = [&](int type) {
    float* Cb = p.C + (size_t)type * p.c_type_stride;
#pragma unroll
    for (int i = 0; i < 8; ++i) {
      const int m = m0 + (i < 4 ? ty * 4 + i : 64 + ty * 4 + (i - 4));
      if (m >= p.K1) continue;
#pragma unroll
      for (int j = 0; j < 8; ++j) {
        const int n = n0 + (j < 4 ? tx * 4 + j : 64 + tx * 4 + (j - 4));
        if (n < p.N && acc[i][j] != 0.f) atomicAdd(Cb + (size_t)m * p.ldc + n, acc[i][j]);
      }
    }
  };
  zero();
  int cur_type = p.tile_type ? p.tile_type[r_begin / GGNN_TILE_M] : 0;

  float4 rx = load_x(r_begin), rd = load_d(r_begin);
  *reinterpret_cast<float4*>(&Xs[0][lk][lc4]) = rx;
  *reinterpret_cast<float4*>(&Ds[0][lk][lc4]) = rd;
  __syncthreads();
  int buf = 0;
  for (int64_t r0 = r_begin; r0 < r_end; r0 += BK) {
    const bool more = r0 + BK < r_end;
    if (more) { rx = load_x(r0 + BK); rd = load_d(r0 + BK); }
#pragma unroll
    for (int kk = 0; kk < BK; ++kk) {
      const float4 a0 = *reinterpret_cast<const float4*>(&Xs[buf][kk][ty * 4]);
      const float4 a1 = *reinterpret_cast<const float4*>(&Xs[buf][kk][64 + ty * 4]);
      const float4 b0 = *reinterpret_cast<const float4*>(&Ds[buf][kk][tx * 4]);
      const float4 b1 = *reinterpret_cast<const float4*>(&Ds[buf][kk][64 + tx * 4]);
      const float a[8] = {a0.x, a0.y, a0.z, a0.w, a1.x, a1.y, a1.z, a1.w};
      const float b[8] = {b0.x, b0.y, b0.z, b0.w, b1.x, b1.y, b1.z, b1.w};
#pragma unroll
      for (int i = 0; i < 8; ++i)
#pragma unroll
        for (int j = 0; j < 8; ++j) acc[i][j] = fmaf(a[i], b[j], acc[i][j]);
    }
    if (more) {
      *reinterpret_cast<float4*>(&Xs[buf ^ 1][lk][lc4]) = rx;
      *reinterpret_cast<float4*>(&Ds[buf ^ 1][lk][lc4]) = rd;
    }
    __syncthreads();
    buf ^= 1;
    // grouped mode: a new 128-row tile may belong to another edge type -> flush and restart
    if (p.tile_type && more && ((r0 + BK) % GGNN_TILE_M) == 0) {
      const int t = p.tile_type[(r0 + BK) / GGNN_TILE_M];
      if (t != cur_type) { flush(cur_type); zero(); cur_type = t; }
    }
  }
  flush(cur_type);
}

// column sums of D rows (optionally gathered), grouped by tile type: out[type][n] += sum_r D[map(r)][n].
// One CTA = 128 rows (one message tile in grouped mode, so one type) x 128 columns; thread (x, y) sums the
// float4 column group x over rows y, y + 8, ..., then the 8 partial sums are combined in shared memory.
__global__ void __launch_bounds__(256)
colsum_kernel(const float* __restrict__ D, int ldd, const int32_t* __restrict__ d_rowmap, int N,
              int64_t R, float* __restrict__ out, const int32_t* __restrict__ tile_type,
              const int32_t* __restrict__ n_tiles, int out_type_stride) {
  __shared__ float4 part[8][32];
  const int tx = threadIdx.x & 31, ty = threadIdx.x >> 5;
  const int64_t r_begin = (int64_t)blockIdx.x * GGNN_TILE_M;   // row tiles on gridDim.x
  if (tile_type) R = (int64_t)(*n_tiles) * GGNN_TILE_M;
  int64_t r_end = r_begin + GGNN_TILE_M;
  if (r_end > R) r_end = R;
  if (r_begin >= r_end) return;
  const int n = blockIdx.y * 128 + tx * 4;                     // N % 4 == 0
  float4 s = make_float4(0.f, 0.f, 0.f, 0.f);
  if (n < N) {
#pragma unroll 4
    for (int64_t r = r_begin + ty; r < r_end; r += 8) {
      const int64_t src = d_rowmap ? (int64_t)d_rowmap[r] : r;
      if (src >= 0) {
        const float4 v = *reinterpret_cast<const float4*>(D + src * ldd + n);
        s.x += v.x; s.y += v.y; s.z += v.z; s.w += v.w;
      }
    }
  }
  part[ty][tx] = s;
  __syncthreads();
  if (ty == 0 && n < N) {
#pragma unroll
    for (int y = 1; y < 8; ++y) { const float4 v = part[y][tx]; s.x += v.x; s.y += v.y; s.z += v.z; s.w += v.w; }
    const int type = tile_type ? tile_type[r_begin / GGNN_TILE_M] : 0;
    if (s.x != 0.f || s.y != 0.f || s.z != 0.f || s.w != 0.f)
      atomicAdd(reinterpret_cast<float4*>(out + (size_t)type * out_type_stride + n), s);
  }
}

// ---- pointwise pieces of the GRU backward -------------------------------------------------
// dzc = dh'(1-u)(1-c^2);  dzu -> dzg[:, h:];  dh = dh' u;  rh = r*h (recomputed operand of dWc)
// smask != null: dh' arrives w.r.t. the DROPPED state h' * s_t / keep, so it is (dhn * smask) * sscale here (the product
// ggnn_mask_scale would have materialised, same order of operations)
__global__ void gru_bwd_pw1_kernel(const float* __restrict__ dhn, const float* __restrict__ smask, float sscale,
                                   const float* __restrict__ hs,
                                   const float* __restrict__ r, const float* __restrict__ u,
                                   const float* __restrict__ c, float* __restrict__ dzc,
                                   float* __restrict__ dzg, float* __restrict__ dh, float* __restrict__ rh,
                                   int64_t N, int h, int rnd) {   // rnd: GEMM operands rounded to TF32
  const int64_t total = N * h;
  for (int64_t i = (int64_t)blockIdx.x * blockDim.x + threadIdx.x; i < total; i += (int64_t)gridDim.x * blockDim.x) {
    const int64_t n = i / h; const int k = (int)(i % h);
    const float g = smask ? dhn[i] * smask[i] * sscale : dhn[i];
    const float uu = u[i], cc = c[i], hh = hs[i];
    float a = g * (1.f - uu) * (1.f - cc * cc), b = g * (hh - cc) * uu * (1.f - uu), q = r[i] * hh;
    if (rnd) { a = tc::to_tf32(a); b = tc::to_tf32(b); q = tc::to_tf32(q); }
    dzc[i] = a;
    dzg[n * 2 * h + h + k] = b;
    dh[i] = g * uu;
    rh[i] = q;
  }
}
// dxr = [dx | drh]:  dzr -> dzg[:, :h];  dh += drh r;  dx = dxr[:, :h]
__global__ void gru_bwd_pw2_kernel(const float* __restrict__ dxr, const float* __restrict__ hs,
                                   const float* __restrict__ r, float* __restrict__ dzg,
                                   float* __restrict__ dh, float* __restrict__ dx, int64_t N, int h, int rnd) {
  const int64_t total = N * h;
  for (int64_t i = (int64_t)blockIdx.x * blockDim.x + threadIdx.x; i < total; i += (int64_t)gridDim.x * blockDim.x) {
    const int64_t n = i / h; const int k = (int)(i % h);
    const float drh = dxr[n * 2 * h + h + k], rr = r[i];
    const float dzr = drh * hs[i] * rr * (1.f - rr);
    dzg[n * 2 * h + k] = rnd ? tc::to_tf32(dzr) : dzr;
    dh[i] += drh * rr;
    dx[i] = dxr[n * 2 * h + k];
  }
}
// dxh = [dx2 | dh2]:  dx += dx2;  dh += dh2
__global__ void gru_bwd_pw3_kernel(const float* __restrict__ dxh, float* __restrict__ dx,
                                   float* __restrict__ dh, int64_t N, int h, int rnd) {
  const int64_t total = N * h;
  for (int64_t i = (int64_t)blockIdx.x * blockDim.x + threadIdx.x; i < total; i += (int64_t)gridDim.x * blockDim.x) {
    const int64_t n = i / h; const int k = (int)(i % h);
    const float v = dx[i] + dxh[n * 2 * h + k];
    dx[i] = rnd ? tc::to_tf32(v) : v;
    dh[i] += dxh[n * 2 * h + h + k];
  }
}

// out = (accumulate ? out : 0) + x * mask * scale
__global__ void mask_scale_kernel(const float* x, const float* __restrict__ mask, float scale,
                                  float* out, int64_t n, int accumulate) {   // x may alias out
  for (int64_t i = (int64_t)blockIdx.x * blockDim.x + threadIdx.x; i < n; i += (int64_t)gridDim.x * blockDim.x) {
    const float v = x[i] * mask[i] * scale;
    out[i] = accumulate ? out[i] + v : v;
  }
}

// Philox4x32-10 (Salmon et al., SC'11): counter = (i / 4, stream), key = seed; element i of a mask is
// word i % 4 of the block; mask = 1 where u32 * 2^-32 < keep.  Counter-based, so the backward regenerates
// nothing -- the masks are materialised once per forward and read by both passes.
__device__ __forceinline__ void philox_round(uint32_t (&c)[4], uint32_t k0, uint32_t k1) {
  const uint32_t hi0 = __umulhi(0xD2511F53u, c[0]), lo0 = 0xD2511F53u * c[0];
  const uint32_t hi1 = __umulhi(0xCD9E8D57u, c[2]), lo1 = 0xCD9E8D57u * c[2];
  const uint32_t n0 = hi1 ^ c[1] ^ k0, n2 = hi0 ^ c[3] ^ k1;
  c[0] = n0; c[1] = lo1; c[2] = n2; c[3] = lo0;
}
__global__ void __launch_bounds__(256)
dropout_mask_kernel(float* __restrict__ out, int64_t n, unsigned long long seed, unsigned long long stream, float keep) {
  const int64_t nblk = (n + 3) / 4;
  for (int64_t b = (int64_t)blockIdx.x * blockDim.x + threadIdx.x; b < nblk; b += (int64_t)gridDim.x * blockDim.x) {
    uint32_t c[4] = {(uint32_t)b, (uint32_t)((unsigned long long)b >> 32), (uint32_t)stream, (uint32_t)(stream >> 32)};
    uint32_t k0 = (uint32_t)seed, k1 = (uint32_t)(seed >> 32);
#pragma unroll
    for (int r = 0; r < 10; ++r) {
      philox_round(c, k0, k1);
      k0 += 0x9E3779B9u; k1 += 0xBB67AE85u;
    }
#pragma unroll
    for (int q = 0; q < 4; ++q) {
      const int64_t i = b * 4 + q;
      if (i < n) out[i] = ((float)c[q] * 2.3283064365386963e-10f < keep) ? 1.f : 0.f;
    }
  }
}

// msg_tgt[message row] = target node (from the target CSR); padding rows keep -1
__global__ void msg_tgt_kernel(const int32_t* __restrict__ tptr, const int32_t* __restrict__ tmsg,
                               int64_t N, int32_t* __restrict__ msg_tgt) {
  const int64_t n = (int64_t)blockIdx.x * blockDim.x + threadIdx.x;
  if (n >= N) return;
  for (int j = tptr[n]; j < tptr[n + 1]; ++j) msg_tgt[tmsg[j]] = (int32_t)n;
}

}  // namespace

// ---------------------------------------------------------------------------------------
size_t ggnn_grad_count(int two_e, int h) {
  return (size_t)two_e * h * h + (size_t)two_e * h + (size_t)4 * h * h + 2 * h + (size_t)2 * h * h + h;
}

int ggnn_mask_scale(ggnn_ctx* ctx, const float* x, const float* mask, float scale, float* out, int64_t n,
                    bool accumulate, cudaStream_t st) {
  if (n <= 0) return GGNN_OK;
  mask_scale_kernel<<<ctx->sm_count * 8, 256, 0, st>>>(x, mask, scale, out, n, accumulate ? 1 : 0);
  GGNN_LAUNCHED(ctx);
  return GGNN_OK;
}

int ggnn_dropout_mask_fill(ggnn_ctx* ctx, float* out, int64_t n, unsigned long long seed, unsigned long long stream,
                           float keep, cudaStream_t st) {
  if (n <= 0) return GGNN_OK;
  dropout_mask_kernel<<<ctx->sm_count * 8, 256, 0, st>>>(out, n, seed, stream, keep);
  GGNN_LAUNCHED(ctx);
  return GGNN_OK;
}

void ggnn_bwd_take_scratch(ws_alloc& a, const ggnn_plan& plan, int h, int two_e, bool with_w_mask, ggnn_bwd_scratch* s) {
  const size_t nh = (size_t)plan.N * h;
  s->dh_a = a.take<float>(nh);
  s->dh_b = a.take<float>(nh);
  s->dx = a.take<float>(nh);
  s->dzc = a.take<float>(nh);
  s->rh = a.take<float>(nh);
  s->dzg = a.take<float>(2 * nh);
  s->dxr = a.take<float>(2 * nh);
  s->msg_tgt = a.take<int32_t>((size_t)plan.max_tiles * GGNN_TILE_M);
  s->gw_tmp = with_w_mask ? a.take<float>((size_t)two_e * h * h) : nullptr;
}
void ggnn_bwd_take_nt_weights(ws_alloc& a, const ggnn_ctx* ctx, ggnn_bwd_scratch* s) { ggnn_tf32_nt_take(a, ctx, &s->ntw); }

namespace {

int launch_nt(ggnn_ctx* ctx, const nt_params& p, unsigned mtiles, cudaStream_t st) {
  if (mtiles == 0 || p.N == 0) return GGNN_OK;
  dim3 grid(mtiles, (p.N + BN - 1) / BN);
  nt_kernel<<<grid, NT, 0, st>>>(p);
  GGNN_LAUNCHED(ctx);
  return GGNN_OK;
}

int launch_tn(ggnn_ctx* ctx, const tn_params& p, int64_t rows_bound, cudaStream_t st) {
  if (rows_bound <= 0) return GGNN_OK;
  dim3 grid((p.N + BN - 1) / BN, (p.K1 + BM - 1) / BM, (unsigned)ceil_div64(rows_bound, p.rows_per_cta));
  tn_kernel<<<grid, NT, 0, st>>>(p);
  GGNN_LAUNCHED(ctx);
  return GGNN_OK;
}

int launch_colsum(ggnn_ctx* ctx, const float* D, int ldd, const int32_t* map, int N, int64_t R,
                  float* out, const int32_t* tile_type, const int32_t* n_tiles, int stride, cudaStream_t st) {
  if (R <= 0) return GGNN_OK;
  dim3 grid((unsigned)ceil_div64(R, GGNN_TILE_M), (N + 127) / 128);
  colsum_kernel<<<grid, 256, 0, st>>>(D, ldd, map, N, R, out, tile_type, n_tiles, stride);
  GGNN_LAUNCHED(ctx);
  return GGNN_OK;
}

}  // namespace

int ggnn_backward_fp32(ggnn_ctx* ctx, const ggnn_tape& tape, const ggnn_bwd_scratch& s,
                       const float* d_hT, float* d_h0, float* grads, cudaStream_t st) {
  const ggnn_plan& plan = tape.plan;
  const int h = ctx->h, two_e = ctx->two_e;
  const int64_t N = plan.N;
  const ggnn_edge_weights& ew = ctx->edge[tape.slot];
  float* gW = grads;
  float* gB = gW + (size_t)two_e * h * h;
  float* gWg = gB + (size_t)two_e * h;
  float* gbg = gWg + (size_t)4 * h * h;
  float* gWc = gbg + 2 * h;
  float* gbc = gWc + (size_t)2 * h * h;
  const size_t nh = (size_t)N * h;
  if (N == 0) return GGNN_OK;
  if (tape.T == 0) {
    GGNN_CUDA(ctx, cudaMemcpyAsync(d_h0, d_hT, nh * 4, cudaMemcpyDeviceToDevice, st));
    return GGNN_OK;
  }
  prof_scope ps(ctx, GGNN_STAGE_OTHER, st);
  const int64_t mrows = plan.max_tiles * GGNN_TILE_M;
  const bool have_msgs = plan.max_tiles > 0 && plan.nnz > 0;
  if (have_msgs) {
    GGNN_CUDA(ctx, cudaMemsetAsync(s.msg_tgt, 0xFF, (size_t)mrows * 4, st));
    msg_tgt_kernel<<<(unsigned)ceil_div64(N, 256), 256, 0, st>>>(plan.tptr, plan.tmsg, N, s.msg_tgt);
    GGNN_LAUNCHED(ctx);
  }
  const unsigned mt = (unsigned)ceil_div64(N, BM);
  const int pw_grid = ctx->sm_count * 8;
  const bool tf = tape.tf32;                 // contractions on the tensor cores (train_tf32.cu)
  const int rnd = tf ? 1 : 0, hp = ctx->hp;
  ggnn_nt_weights ntw = s.ntw;
  if (tf) {
    GGNN_TRY(ggnn_tf32_pack_nt_gru(ctx, &ntw, st));
    if (have_msgs && !tape.w_mask) GGNN_TRY(ggnn_tf32_pack_nt_edge(ctx, ew.W, nullptr, 1.f, &ntw, st));
  }
  const int tn_rows = 2048;
  const float* dhn = d_hT;
  float* dh = s.dh_a;
  for (int t = tape.T - 1; t >= 0; --t) {
    const float *ht = tape.at(t, TAPE_H), *xt = tape.at(t, TAPE_X), *rt = tape.at(t, TAPE_R),
                *ut = tape.at(t, TAPE_U), *ct = tape.at(t, TAPE_C);
    const float inv_keep = 1.0f / tape.keep;
    // (dhn is w.r.t. the dropped state h' * s_t / keep: the first pointwise kernel applies the mask on the way in)
    const float* smask = tape.state_mask ? tape.state_mask + (size_t)t * nh : nullptr;
    const float* Wt = ew.W;
    if (tape.w_mask && !tf) {   // this step's effective weights W * m_t / keep (tensor-core path: formed inside the operand pack)
      GGNN_TRY(ggnn_mask_scale(ctx, ew.W, tape.w_mask + (size_t)t * two_e * h * h, inv_keep, tape.w_eff,
                               (int64_t)two_e * h * h, false, st));
      Wt = tape.w_eff;
    }
    gru_bwd_pw1_kernel<<<pw_grid, 256, 0, st>>>(dhn, smask, inv_keep, ht, rt, ut, ct, s.dzc, s.dzg, dh, s.rh, N, h, rnd);
    GGNN_LAUNCHED(ctx);
    if (tf) {
      if (tape.w_mask && have_msgs)
        GGNN_TRY(ggnn_tf32_pack_nt_edge(ctx, ew.W, tape.w_mask + (size_t)t * two_e * h * h, inv_keep, &ntw, st));
      GGNN_TRY(ggnn_tf32_nt(ctx, s.dzc, nullptr, h, nullptr, N, mt, ntw.Wc, ntw.tmC, ntw.rowsC, hp, 2 * hp, nullptr,
                            nullptr, s.dxr, 2 * h, nullptr, st));
      for (int part = 0; part < 2; ++part)
        GGNN_TRY(ggnn_tf32_tn(ctx, part ? s.rh : xt, h, nullptr, h, s.dzc, h, nullptr, h, N, 0,
                              gWc + (size_t)part * h * h, h, nullptr, nullptr, 0, st));
    } else {
    {  // [dx | drh] = dzc @ Wc^T
      nt_params p{};
      p.A1 = s.dzc; p.K1 = h; p.lda = h; p.B = ctx->Wc; p.ldb = h; p.N = 2 * h; p.M = N;
      p.C = s.dxr; p.ldc = 2 * h;
      GGNN_TRY(launch_nt(ctx, p, mt, st));
    }
    for (int part = 0; part < 2; ++part) {   // dWc += [x, rh]^T dzc
      tn_params p{};
      p.X = part ? s.rh : xt; p.ldx = h; p.D = s.dzc; p.ldd = h; p.K1 = h; p.N = h; p.R = N;
      p.rows_per_cta = tn_rows; p.C = gWc + (size_t)part * h * h; p.ldc = h;
      GGNN_TRY(launch_tn(ctx, p, N, st));
    }
    }
    GGNN_TRY(launch_colsum(ctx, s.dzc, h, nullptr, h, N, gbc, nullptr, nullptr, 0, st));
    gru_bwd_pw2_kernel<<<pw_grid, 256, 0, st>>>(s.dxr, ht, rt, s.dzg, dh, s.dx, N, h, rnd);
    GGNN_LAUNCHED(ctx);
    if (tf) {
      GGNN_TRY(ggnn_tf32_nt(ctx, s.dzg, s.dzg + h, 2 * h, nullptr, N, mt, ntw.Wg, ntw.tmG, ntw.rowsG, 2 * hp, 2 * hp,
                            nullptr, nullptr, s.dxr, 2 * h, nullptr, st));
      for (int part = 0; part < 2; ++part)
        GGNN_TRY(ggnn_tf32_tn(ctx, part ? ht : xt, h, nullptr, h, s.dzg, 2 * h, nullptr, 2 * h, N, 0,
                              gWg + (size_t)part * h * 2 * h, 2 * h, nullptr, nullptr, 0, st));
    } else {
    {  // [dx2 | dh2] = [dzr, dzu] @ Wg^T
      nt_params p{};
      p.A1 = s.dzg; p.K1 = 2 * h; p.lda = 2 * h; p.B = ctx->Wg; p.ldb = 2 * h; p.N = 2 * h; p.M = N;
      p.C = s.dxr; p.ldc = 2 * h;
      GGNN_TRY(launch_nt(ctx, p, mt, st));
    }
    for (int part = 0; part < 2; ++part) {   // dWg += [x, h]^T [dzr, dzu]
      tn_params p{};
      p.X = part ? ht : xt; p.ldx = h; p.D = s.dzg; p.ldd = 2 * h; p.K1 = h; p.N = 2 * h; p.R = N;
      p.rows_per_cta = tn_rows; p.C = gWg + (size_t)part * h * 2 * h; p.ldc = 2 * h;
      GGNN_TRY(launch_tn(ctx, p, N, st));
    }
    }
    GGNN_TRY(launch_colsum(ctx, s.dzg, 2 * h, nullptr, 2 * h, N, gbg, nullptr, nullptr, 0, st));
    gru_bwd_pw3_kernel<<<pw_grid, 256, 0, st>>>(s.dxr, s.dx, dh, N, h, rnd);
    GGNN_LAUNCHED(ctx);
    if (have_msgs && tf) {
      float* gw = tape.w_mask ? s.gw_tmp : gW;
      GGNN_TRY(ggnn_tf32_nt(ctx, s.dx, nullptr, h, s.msg_tgt, mrows, (unsigned)plan.max_tiles, ntw.We, ntw.tmE, ntw.rowsE,
                            hp, hp, plan.tile_type, plan.n_tiles, dh, h, plan.msg_src, st));
      if (tape.w_mask) GGNN_CUDA(ctx, cudaMemsetAsync(s.gw_tmp, 0, (size_t)two_e * h * h * 4, st));
      GGNN_TRY(ggnn_tf32_tn(ctx, ht, h, plan.msg_src, h, s.dx, h, s.msg_tgt, h, mrows, 8 * GGNN_TILE_M, gw, h,
                            plan.tile_type, plan.n_tiles, (int64_t)h * h, st));
      if (tape.w_mask)
        GGNN_TRY(ggnn_mask_scale(ctx, s.gw_tmp, tape.w_mask + (size_t)t * two_e * h * h, inv_keep, gW,
                                 (int64_t)two_e * h * h, true, st));
      GGNN_TRY(launch_colsum(ctx, s.dx, h, s.msg_tgt, h, mrows, gB, plan.tile_type, plan.n_tiles, h, st));
    } else if (have_msgs) {
      {  // dh[src_j] += dx[tgt_j] @ W_e^T
        nt_params p{};
        p.A1 = s.dx; p.K1 = h; p.lda = h; p.a_rowmap = s.msg_tgt;
        p.B = Wt; p.ldb = h; p.b_type_stride = (int64_t)h * h; p.N = h; p.M = mrows;
        p.tile_type = plan.tile_type; p.n_tiles = plan.n_tiles;
        p.C = dh; p.ldc = h; p.c_rowmap = plan.msg_src;
        GGNN_TRY(launch_nt(ctx, p, (unsigned)plan.max_tiles, st));
      }
      {  // dW_e += h[src_j]^T dx[tgt_j]
        tn_params p{};
        p.X = ht; p.ldx = h; p.x_rowmap = plan.msg_src; p.D = s.dx; p.ldd = h; p.d_rowmap = s.msg_tgt;
        p.K1 = h; p.N = h; p.R = mrows; p.rows_per_cta = 8 * GGNN_TILE_M;
        p.C = tape.w_mask ? s.gw_tmp : gW; p.ldc = h; p.tile_type = plan.tile_type; p.n_tiles = plan.n_tiles;
        p.c_type_stride = (int64_t)h * h;
        if (tape.w_mask) GGNN_CUDA(ctx, cudaMemsetAsync(s.gw_tmp, 0, (size_t)two_e * h * h * 4, st));
        GGNN_TRY(launch_tn(ctx, p, mrows, st));
        if (tape.w_mask)     // dW += dW_t * m_t / keep
          GGNN_TRY(ggnn_mask_scale(ctx, s.gw_tmp, tape.w_mask + (size_t)t * two_e * h * h, inv_keep, gW,
                                   (int64_t)two_e * h * h, true, st));
      }
      GGNN_TRY(launch_colsum(ctx, s.dx, h, s.msg_tgt, h, mrows, gB, plan.tile_type, plan.n_tiles, h, st));
    }
    dhn = dh;
    dh = (dh == s.dh_a) ? s.dh_b : s.dh_a;
  }
  GGNN_CUDA(ctx, cudaMemcpyAsync(d_h0, dhn, nh * 4, cudaMemcpyDeviceToDevice, st));
  return GGNN_OK;
}

// ---------------------------------------------------------------------------------------
// N4: per-tensor clip_by_norm + Adam over the flat gradient buffer (make_train_step,
// chem_tensorflow.py:483-506: AdamOptimizer(lr, beta1=0.9, beta2=0.999), tf.clip_by_norm per variable)
namespace {

// Fixed-order two-stage norm: every block writes ONE partial (no float atomics), and every block of the Adam kernel
// re-reduces the same partials in the same order, so the clip scale is bit-identical from run to run and from rank
// to rank (the all-reduced gradient is) and data-parallel replicas cannot drift apart.
constexpr int SUMSQ_MAX_BLOCKS = 2048;

__device__ __forceinline__ float block_sum_fixed(float s, float* ws /* [8] */) {
#pragma unroll
  for (int o = 16; o > 0; o >>= 1) s += __shfl_xor_sync(0xffffffffu, s, o);
  if ((threadIdx.x & 31) == 0) ws[threadIdx.x >> 5] = s;
  __syncthreads();
  float t = 0.f;
#pragma unroll
  for (int w = 0; w < 8; ++w) t += ws[w];
  return t;
}

__global__ void __launch_bounds__(256)
sumsq_kernel(const float* __restrict__ g, int64_t n, float* __restrict__ partial) {
  __shared__ float ws[8];
  float s = 0.f;
  for (int64_t i = (int64_t)blockIdx.x * blockDim.x + threadIdx.x; i < n; i += (int64_t)gridDim.x * blockDim.x)
    s = fmaf(g[i], g[i], s);
  const float t = block_sum_fixed(s, ws);
  if (threadIdx.x == 0) partial[blockIdx.x] = t;
}

// g_c = g * clip / max(||g||, clip);  m, v, p updated in place (TF1 AdamOptimizer: eps outside the sqrt)
__global__ void __launch_bounds__(256)
adam_kernel(float* __restrict__ p, const float* __restrict__ g, float* __restrict__ m, float* __restrict__ v,
            int64_t n, const float* __restrict__ partial, int n_partial, float clip, float lr_t, float b1, float b2,
            float eps) {
  __shared__ float ws[8];
  float s = 0.f;
  for (int i = threadIdx.x; i < n_partial; i += 256) s += partial[i];
  const float norm = sqrtf(block_sum_fixed(s, ws));
  const float scale = clip > 0.f ? clip / fmaxf(norm, clip) : 1.f;
  for (int64_t i = (int64_t)blockIdx.x * blockDim.x + threadIdx.x; i < n; i += (int64_t)gridDim.x * blockDim.x) {
    const float gc = g[i] * scale;
    const float mi = b1 * m[i] + (1.f - b1) * gc;
    const float vi = b2 * v[i] + (1.f - b2) * gc * gc;
    m[i] = mi; v[i] = vi;
    p[i] -= lr_t * mi / (sqrtf(vi) + eps);
  }
}

}  // namespace

int ggnn_adam_segment(ggnn_ctx* ctx, float* param, const float* grad, float* m, float* v, int64_t n,
                      float* partials /* [GGNN_SUMSQ_SLOT] */, float clip, float lr_t, float b1, float b2, float eps,
                      cudaStream_t st) {
  if (n <= 0 || !param) return GGNN_OK;
  int64_t want = ceil_div64(n, 256);
  if (want > (int64_t)ctx->sm_count * 8) want = (int64_t)ctx->sm_count * 8;
  const int grid = (int)(want < SUMSQ_MAX_BLOCKS ? want : SUMSQ_MAX_BLOCKS);
  static_assert(SUMSQ_MAX_BLOCKS == GGNN_SUMSQ_SLOT, "partial-sum slot size");
  sumsq_kernel<<<grid, 256, 0, st>>>(grad, n, partials);
  GGNN_LAUNCHED(ctx);
  adam_kernel<<<grid, 256, 0, st>>>(param, grad, m, v, n, partials, grid, clip, lr_t, b1, b2, eps);
  GGNN_LAUNCHED(ctx);
  return GGNN_OK;
}
