// tf32.cu -- GGNN_PREC_TF32 timestep: the three contractions of a timestep on the 5th-gen
// tensor cores (tcgen05.mma kind::tf32, FP32 accumulators in TMEM).
//
//   MSG    M[r,:]  = tf32(h[msg_src[r],:]) @ tf32(W[type]) + B[type]
//   GATES  g = sigmoid([tf32(acts), tf32(h)] @ tf32(Wg) + bg);  rh = g[:, :h]*h;  u = g[:, h:]
//   CAND   c = tanh([tf32(acts), tf32(rh)] @ tf32(Wc) + bc);    h' = u*h + (1-u)*c
//
// One CTA = one 128-row x (<=256)-column output tile (two CTAs per SM: 2 x 48 KB stages, 256 TMEM columns each).  Warp roles:
//   warps 0-3  operand producers: activations by 16-byte cp.async straight into the
//              128B-swizzled tile (no register staging), weights by one TMA box per K chunk;
//              then the epilogue (tcgen05.ld, fused bias / sigmoid / tanh / state update)
//   warp  4    one elected lane issues tcgen05.mma and commits to the mbarriers
// Operands are K-major; weights are pre-transposed, TF32-rounded and zero-padded to hp
// (a multiple of 32) once in ggnn_set_weights, so the mainloop needs no K/N edge cases.
#include "tf32.cuh"
#include "tc.cuh"
#include "pack_t.cuh"
#include <cuda_fp16.h>

namespace {

constexpr int BM = 128;           // rows per tile  (UMMA M, one TMEM lane per row)
constexpr int BN = 256;           // max columns per tile (UMMA N, TMEM columns)
constexpr int BKF = 32;           // floats per K chunk = one 128-byte swizzle row
constexpr int STAGES = 2;           // 2 x 48 KB: two CTAs per SM, so one CTA's epilogue overlaps the other's mainloop
constexpr int A_BYTES = BM * BKF * 4;     // 16 KB
constexpr int B_BYTES = BN * BKF * 4;     // 32 KB
constexpr int NPROD = 128;                // producer / epilogue threads (warps 0-3)
constexpr int NTHREADS = NPROD + 32;      // + MMA warp
constexpr int SMEM_BYTES = STAGES * (A_BYTES + B_BYTES) + 1024 /*align*/ + 256 /*barriers*/;

enum { MODE_MSG = 0, MODE_GATES = 1, MODE_CAND = 2, MODE_NT = 3 };

struct tparams {
  const float* A1;          // [*, h]  K segment 1 source rows
  const float* A2;          // [*, h]  K segment 2 source rows (GATES: h, CAND: rh) or null
  const int32_t* gather;    // MSG: source row per message row (-1 = padding)
  const float* Bt;          // [Np, Kp] transposed, TF32-rounded, padded weights (per type for MSG)
  const float* bias;        // [Np] padded bias (per type: [h] unpadded for MSG) or null
  int h, hp;                // real / padded hidden size
  int Kp;                   // padded K = hp * (segments)
  int Np;                   // padded N
  int64_t M;
  const int32_t* tile_type;
  const int32_t* n_tiles;
  float* out0;              // MSG: msg;  GATES: rh;  CAND: h state (in/out)
  float* out1;              // GATES: u
  const float* hstate;      // GATES: h
  const float* ubuf;        // CAND: u
  int b_bytes;              // bytes one TMA box of the weight map delivers per K chunk
  __half* out_half;         // MSG: when set, messages are stored as FP16 [M, h] instead of out0
  int round_out;            // GATES: always rounds rh; CAND: round h' to TF32 (not on the last step)
  int lda1, lda2;           // row strides of A1 / A2 in floats (0 = h)
  float* tape_r;            // GATES: also store r (training tape) or null
  float* tape_c;            // CAND: also store the candidate c (training tape) or null
  // NT (backward, C = A @ B^T with B in its own row-major layout): padded output column n maps to
  // out0[crow * ldc + (n / hp) * h + n % hp]; crow = row, or c_rowmap[row] with atomic accumulation
  int ldc;
  const int32_t* c_rowmap;
};

__device__ __forceinline__ float sigmoid_f(float x) { return 1.0f / (1.0f + __expf(-x)); }
__device__ __forceinline__ float tanh_f(float x) {
  // tanh(x) = 1 - 2/(exp(2x)+1); __expf relative error ~2^-21, far below the TF32 noise floor
  float e = __expf(2.0f * x);
  return 1.0f - 2.0f / (e + 1.0f);
}

template <int MODE>
__global__ void __launch_bounds__(NTHREADS)
gemm_tf32_kernel(const tparams p, const __grid_constant__ CUtensorMap tmapB) {
  extern __shared__ uint8_t smem_raw[];
  const int tid = threadIdx.x, warp = tid >> 5, lane = tid & 31;
  // row tiles on gridDim.x (limit 2^31-1), column tiles on gridDim.y
  const int64_t row0 = (int64_t)blockIdx.x * BM;
  const int n0 = blockIdx.y * BN;
  const int n_tile = min(BN, p.Np - n0);              // multiple of 32

  const float* Bt = p.Bt;
  const float* bias = p.bias;
  const bool grouped = MODE == MODE_MSG || (MODE == MODE_NT && p.tile_type != nullptr);
  if (grouped) {
    if ((int)blockIdx.x >= *p.n_tiles) return;        // uniform across the CTA
    const int type = p.tile_type[blockIdx.x];
    Bt += (size_t)type * p.Np * p.Kp;
    if (bias) bias += (size_t)type * p.h;
  }

  // ---- shared memory carve-up (1024-byte aligned operand tiles)
  const uint32_t base = (tc::smem_u32(smem_raw) + 1023u) & ~1023u;
  const uint32_t sA = base, sB = base + STAGES * A_BYTES;
  const uint32_t bars = sB + STAGES * B_BYTES;
  auto full_bar = [&](int s) { return bars + 8u * s; };
  auto empty_bar = [&](int s) { return bars + 8u * (STAGES + s); };
  const uint32_t accum_bar = bars + 8u * (2 * STAGES);
  const uint32_t tmem_slot = bars + 8u * (2 * STAGES + 1);
  uint8_t* smem_al = smem_raw + (base - tc::smem_u32(smem_raw));

  if (warp == 4 && lane == 0) {
    for (int s = 0; s < STAGES; ++s) {
      tc::mbar_init(full_bar(s), NPROD + 1);
      tc::mbar_init(empty_bar(s), 1);
    }
    tc::mbar_init(accum_bar, 1);
    tc::fence_barrier_init();
  }
  if (warp == 0) tc::tmem_alloc(tmem_slot, BN);
  if (warp == 1 && lane == 0) tc::tma_prefetch_desc(&tmapB);
  tc::fence_before_sync();
  __syncthreads();
  tc::fence_after_sync();
  const uint32_t tmem_base = *reinterpret_cast<volatile uint32_t*>(smem_al + (tmem_slot - base));

  const int KC = p.Kp / BKF;                          // K chunks
  const int KC1 = p.hp / BKF;                         // chunks of segment 1

  if (warp < 4) {
    // ================= producers =================
    const int c16 = tid & 7;                          // 16-byte chunk within the 128-byte row
    const int rbase = tid >> 3;                       // rows rbase + 16*j, j = 0..7
    int64_t src1[8], src2[8];
#pragma unroll
    for (int j = 0; j < 8; ++j) {
      const int64_t grow = row0 + rbase + 16 * j;
      src1[j] = -1; src2[j] = -1;
      if (grow < p.M) {
        if (MODE == MODE_MSG || (MODE == MODE_NT && p.gather)) { src1[j] = p.gather[grow]; src2[j] = src1[j]; }
        else { src1[j] = grow; src2[j] = grow; }
      }
    }
    int b_row = n0;
    if (grouped) b_row += p.tile_type[blockIdx.x] * p.Np;
    uint32_t a_off[8];
#pragma unroll
    for (int j = 0; j < 8; ++j) {
      const int r = rbase + 16 * j;
      a_off[j] = (r >> 3) * 1024 + (r & 7) * 128 + ((c16 ^ (r & 7)) << 4);
    }
    for (int kc = 0; kc < KC; ++kc) {
      const int s = kc % STAGES;
      if (kc >= STAGES) tc::mbar_wait(empty_bar(s), ((kc / STAGES) - 1) & 1);
      if (tid == 0) {        // weights: one TMA box [b rows x 32 k] per chunk, 128B-swizzled by the TMA unit
        tc::mbar_arrive_expect_tx(full_bar(s), (uint32_t)p.b_bytes);
        tc::tma_load_2d(sB + s * B_BYTES, &tmapB, kc * BKF, b_row, full_bar(s));
      }
      // activations: 16-byte cp.async straight into the swizzled tile (operands are already
      // TF32-rounded where they were produced, so no register staging is needed)
      const bool seg2 = kc >= KC1;
      const float* Asrc = seg2 ? p.A2 : p.A1;
      const int lda = seg2 ? p.lda2 : p.lda1;
      const int kloc = (seg2 ? kc - KC1 : kc) * BKF + c16 * 4;     // column within the segment
      const bool kin = kloc < p.h;                                 // h % 4 == 0
      const uint32_t a_st = sA + s * A_BYTES;
#pragma unroll
      for (int j = 0; j < 8; ++j) {
        const int64_t r = seg2 ? src2[j] : src1[j];
        const bool ok = kin && r >= 0;
        tc::cp_async16(a_st + a_off[j], ok ? (const void*)(Asrc + r * lda + kloc) : (const void*)Asrc,
                       ok ? 16u : 0u);
      }
      tc::cp_async_arrive(full_bar(s));
    }

    // ================= epilogue =================
    tc::mbar_wait(accum_bar, 0);
    tc::fence_after_sync();
    const int64_t row = row0 + warp * 32 + lane;
    bool row_ok = row < p.M;
    if (MODE == MODE_MSG && row_ok) row_ok = p.gather[row] >= 0;
    int64_t crow = row;
    if (MODE == MODE_NT && row_ok && p.c_rowmap) { crow = p.c_rowmap[row]; row_ok = crow >= 0; }
    for (int cc = 0; cc < n_tile / 32; ++cc) {
      uint32_t acc[32];
      tc::tmem_ld32(tmem_base + ((uint32_t)(warp * 32) << 16) + (uint32_t)(cc * 32), acc);
      tc::tmem_ld_wait();
      const int n = n0 + cc * 32;                      // padded column index of acc[0]
      if (!row_ok) {
        // nothing to store for this row (keep the warp converged for the next tcgen05.ld)
      } else if (MODE == MODE_NT) {
        const int nb = n / p.hp, cbase = n - nb * p.hp;
        float* dst = p.out0 + crow * p.ldc + nb * p.h + cbase;
#pragma unroll
        for (int q = 0; q < 8; ++q) {
          if (cbase + q * 4 >= p.h) break;
          const float4 v = make_float4(__uint_as_float(acc[q * 4 + 0]), __uint_as_float(acc[q * 4 + 1]),
                                       __uint_as_float(acc[q * 4 + 2]), __uint_as_float(acc[q * 4 + 3]));
          if (p.c_rowmap) atomicAdd(reinterpret_cast<float4*>(dst + q * 4), v);
          else *reinterpret_cast<float4*>(dst + q * 4) = v;
        }
      } else if (MODE == MODE_MSG || MODE == MODE_CAND) {
#pragma unroll
        for (int q = 0; q < 8; ++q) {
          const int col = n + q * 4;
          if (col >= p.h) break;
          float v[4];
#pragma unroll
          for (int i = 0; i < 4; ++i) v[i] = __uint_as_float(acc[q * 4 + i]);
          if (bias) {
            const float4 bb = __ldg(reinterpret_cast<const float4*>(bias + col));
            v[0] += bb.x; v[1] += bb.y; v[2] += bb.z; v[3] += bb.w;
          }
          if (MODE == MODE_MSG) {
            if (p.out_half) {
              __half2 lo = __floats2half2_rn(v[0], v[1]), hi = __floats2half2_rn(v[2], v[3]);
              uint2 pk;
              pk.x = *reinterpret_cast<uint32_t*>(&lo);
              pk.y = *reinterpret_cast<uint32_t*>(&hi);
              *reinterpret_cast<uint2*>(p.out_half + row * p.h + col) = pk;
            } else {
              *reinterpret_cast<float4*>(p.out0 + row * p.h + col) = make_float4(v[0], v[1], v[2], v[3]);
            }
          } else {
            const float4 hv = *reinterpret_cast<const float4*>(p.out0 + row * p.h + col);
            const float4 uv = *reinterpret_cast<const float4*>(p.ubuf + row * p.h + col);
            const float4 cv = make_float4(tanh_f(v[0]), tanh_f(v[1]), tanh_f(v[2]), tanh_f(v[3]));
            if (p.tape_c) *reinterpret_cast<float4*>(p.tape_c + row * p.h + col) = cv;
            float4 o;
            o.x = uv.x * hv.x + (1.0f - uv.x) * cv.x;
            o.y = uv.y * hv.y + (1.0f - uv.y) * cv.y;
            o.z = uv.z * hv.z + (1.0f - uv.z) * cv.z;
            o.w = uv.w * hv.w + (1.0f - uv.w) * cv.w;
            if (p.round_out) {
              o.x = tc::to_tf32(o.x); o.y = tc::to_tf32(o.y); o.z = tc::to_tf32(o.z); o.w = tc::to_tf32(o.w);
            }
            *reinterpret_cast<float4*>(p.out0 + row * p.h + col) = o;
          }
        }
      } else {  // GATES: padded columns [0,hp) = r, [hp,2hp) = u
        const bool is_r = n < p.hp;
        const int cbase = is_r ? n : n - p.hp;
#pragma unroll
        for (int q = 0; q < 8; ++q) {
          const int col = cbase + q * 4;
          if (col >= p.h) break;
          const float4 bb = __ldg(reinterpret_cast<const float4*>(bias + n + q * 4));
          float g[4];
          g[0] = sigmoid_f(__uint_as_float(acc[q * 4 + 0]) + bb.x);
          g[1] = sigmoid_f(__uint_as_float(acc[q * 4 + 1]) + bb.y);
          g[2] = sigmoid_f(__uint_as_float(acc[q * 4 + 2]) + bb.z);
          g[3] = sigmoid_f(__uint_as_float(acc[q * 4 + 3]) + bb.w);
          if (is_r) {
            const float4 hv = *reinterpret_cast<const float4*>(p.hstate + row * p.h + col);
            *reinterpret_cast<float4*>(p.out0 + row * p.h + col) =
                make_float4(tc::to_tf32(g[0] * hv.x), tc::to_tf32(g[1] * hv.y),
                            tc::to_tf32(g[2] * hv.z), tc::to_tf32(g[3] * hv.w));
            if (p.tape_r) *reinterpret_cast<float4*>(p.tape_r + row * p.h + col) = make_float4(g[0], g[1], g[2], g[3]);
          } else {
            *reinterpret_cast<float4*>(p.out1 + row * p.h + col) = make_float4(g[0], g[1], g[2], g[3]);
          }
        }
      }
    }
  } else {
    // ================= MMA issuer (warp 4, one lane) =================
    if (lane == 0) {
      const uint32_t idesc = tc::umma_idesc(/*TF32*/ 2, BM, n_tile);
      for (int kc = 0; kc < KC; ++kc) {
        const int s = kc % STAGES;
        tc::mbar_wait(full_bar(s), (kc / STAGES) & 1);
        tc::fence_after_sync();
        const uint32_t a0 = sA + s * A_BYTES, b0 = sB + s * B_BYTES;
#pragma unroll
        for (int k = 0; k < BKF / 8; ++k) {           // UMMA K = 8 for TF32 (32 bytes)
          tc::umma_tf32(tmem_base, tc::umma_desc_sw128(a0 + k * 32), tc::umma_desc_sw128(b0 + k * 32),
                        idesc, (kc | k) != 0);
        }
        tc::umma_commit(empty_bar(s));                // frees the smem stage when the MMAs retire
      }
      tc::umma_commit(accum_bar);                     // accumulator complete -> epilogue
    }
    __syncwarp();
  }

  tc::fence_before_sync();
  __syncthreads();
  if (warp == 0) tc::tmem_dealloc(tmem_base, BN);
}

// ---- weight packing -------------------------------------------------------------------
// edge weights: dst[e][n][k] (n, k < hp) = tf32(src[e][k][n]) for k, n < h, else 0 -- packt::edge_t_kernel (pack_t.cuh)
// lo != 0: the residual pack of the 3xTF32 split (strict_tc.cu), tf32(w - tf32(w))
__device__ __forceinline__ float pack_val(float w, int lo) {
  const float hi = tc::to_tf32(w);
  return lo ? tc::to_tf32(w - hi) : hi;
}
// GRU kernels: src [2h, ncols*h] (rows: x part then h part; column blocks of width h)
// dst [ncols*hp, 2hp]: dst[nb*hp + n][kb*hp + k] = tf32(src[kb*h + k][nb*h + n])
__global__ void pack_gru_kernel(const float* __restrict__ src, float* __restrict__ dst, int ncols,
                                int h, int hp, int lo) {
  const int Kp = 2 * hp, Np = ncols * hp;
  const int64_t total = (int64_t)Np * Kp;
  for (int64_t i = (int64_t)blockIdx.x * blockDim.x + threadIdx.x; i < total;
       i += (int64_t)gridDim.x * blockDim.x) {
    const int kk = (int)(i % Kp), nn = (int)(i / Kp);
    const int kb = kk / hp, k = kk % hp, nb = nn / hp, n = nn % hp;
    dst[i] = (k < h && n < h) ? pack_val(src[(size_t)(kb * h + k) * (ncols * h) + nb * h + n], lo) : 0.f;
  }
}

__global__ void pack_bias_kernel(const float* __restrict__ src, float* __restrict__ dst, int nblocks,
                                 int h, int hp) {
  const int i = blockIdx.x * blockDim.x + threadIdx.x;
  if (i >= nblocks * hp) return;
  const int nb = i / hp, n = i % hp;
  dst[i] = n < h ? src[nb * h + n] : 0.f;
}

int realloc_f(ggnn_ctx* ctx, float*& p, size_t n) {
  if (p && ctx->repack_in_place) return GGNN_OK;
  if (p) { cudaFree(p); p = nullptr; }
  GGNN_CUDA(ctx, cudaMalloc(&p, n * sizeof(float)));
  return GGNN_OK;
}

// ---- TMA tensor maps over the packed weights ---------------------------------------------
typedef CUresult (*encode_tiled_fn)(CUtensorMap*, CUtensorMapDataType, cuuint32_t, void*, const cuuint64_t*,
                                    const cuuint64_t*, const cuuint32_t*, const cuuint32_t*,
                                    CUtensorMapInterleave, CUtensorMapSwizzle, CUtensorMapL2promotion,
                                    CUtensorMapFloatOOBfill);

encode_tiled_fn get_encode() {
  static encode_tiled_fn fn = nullptr;
  if (!fn) {
    void* p = nullptr;
    cudaDriverEntryPointQueryResult q;
    if (cudaGetDriverEntryPoint("cuTensorMapEncodeTiled", &p, cudaEnableDefault, &q) == cudaSuccess &&
        q == cudaDriverEntryPointSuccess)
      fn = reinterpret_cast<encode_tiled_fn>(p);
  }
  return fn;
}

__global__ void round_copy_kernel(const float4* __restrict__ src, float4* __restrict__ dst, int64_t n4) {
  for (int64_t i = (int64_t)blockIdx.x * blockDim.x + threadIdx.x; i < n4; i += (int64_t)gridDim.x * blockDim.x) {
    float4 v = src[i];
    v.x = tc::to_tf32(v.x); v.y = tc::to_tf32(v.y); v.z = tc::to_tf32(v.z); v.w = tc::to_tf32(v.w);
    dst[i] = v;
  }
}

template <int MODE>
int launch(ggnn_ctx* ctx, const tparams& p_in, const CUtensorMap& tm, unsigned mtiles, cudaStream_t st) {
  tparams p = p_in;
  if (p.lda1 == 0) p.lda1 = p.h;
  if (p.lda2 == 0) p.lda2 = p.h;
  GGNN_ENSURE_SMEM(ctx, GGNN_ATTR_TF32_GEMM0 + MODE, gemm_tf32_kernel<MODE>, SMEM_BYTES);
  dim3 grid(mtiles, (p.Np + BN - 1) / BN);
  gemm_tf32_kernel<MODE><<<grid, NTHREADS, SMEM_BYTES, st>>>(p, tm);
  GGNN_LAUNCHED(ctx);
  return GGNN_OK;
}

}  // namespace

int ggnn_make_tmap_2d(ggnn_ctx* ctx, CUtensorMap* tm, const float* base, int64_t rows, int64_t cols,
                      int box_rows) {
  encode_tiled_fn enc = get_encode();
  if (!enc) GGNN_FAIL(ctx, GGNN_E_CUDA, "cuTensorMapEncodeTiled entry point not found");
  cuuint64_t gdim[2] = {(cuuint64_t)cols, (cuuint64_t)rows};
  cuuint64_t gstride[1] = {(cuuint64_t)cols * 4};
  cuuint32_t box[2] = {(cuuint32_t)BKF, (cuuint32_t)box_rows};
  cuuint32_t estr[2] = {1, 1};
  CUresult r = enc(tm, CU_TENSOR_MAP_DATA_TYPE_FLOAT32, 2, const_cast<float*>(base), gdim, gstride, box,
                   estr, CU_TENSOR_MAP_INTERLEAVE_NONE, CU_TENSOR_MAP_SWIZZLE_128B,
                   CU_TENSOR_MAP_L2_PROMOTION_L2_256B, CU_TENSOR_MAP_FLOAT_OOB_FILL_NONE);
  if (r != CUDA_SUCCESS) GGNN_FAIL(ctx, GGNN_E_CUDA, "cuTensorMapEncodeTiled failed (%d)", (int)r);
  return GGNN_OK;
}

int ggnn_make_tmap_2d_f16(ggnn_ctx* ctx, CUtensorMap* tm, const void* base, int64_t rows, int64_t cols,
                          int box_cols, int box_rows) {
  encode_tiled_fn enc = get_encode();
  if (!enc) GGNN_FAIL(ctx, GGNN_E_CUDA, "cuTensorMapEncodeTiled entry point not found");
  cuuint64_t gdim[2] = {(cuuint64_t)cols, (cuuint64_t)rows};
  cuuint64_t gstride[1] = {(cuuint64_t)cols * 2};
  cuuint32_t box[2] = {(cuuint32_t)box_cols, (cuuint32_t)box_rows};
  cuuint32_t estr[2] = {1, 1};
  CUresult r = enc(tm, CU_TENSOR_MAP_DATA_TYPE_FLOAT16, 2, const_cast<void*>(base), gdim, gstride, box,
                   estr, CU_TENSOR_MAP_INTERLEAVE_NONE, CU_TENSOR_MAP_SWIZZLE_128B,
                   CU_TENSOR_MAP_L2_PROMOTION_L2_256B, CU_TENSOR_MAP_FLOAT_OOB_FILL_NONE);
  if (r != CUDA_SUCCESS) GGNN_FAIL(ctx, GGNN_E_CUDA, "cuTensorMapEncodeTiled (f16) failed (%d)", (int)r);
  return GGNN_OK;
}

// np = output columns one GEMM sees (per edge type for the message weights): the box never needs more rows
static int make_weight_map(ggnn_ctx* ctx, CUtensorMap* tm, float* base, int64_t rows, int64_t cols,
                           int* box_rows, int np) {
  *box_rows = (int)(rows < BN ? rows : BN);
  if (*box_rows > np) *box_rows = np;
  return ggnn_make_tmap_2d(ctx, tm, base, rows, cols, *box_rows);
}

// ---- backward (train_tf32.cu): NT contractions on the forward kernel ----------------------
size_t ggnn_tf32_nt_weight_floats(const ggnn_ctx* ctx) {
  const size_t hp = ctx->hp;
  return (size_t)ctx->two_e * hp * hp + 4 * hp * hp + 2 * hp * hp + 64;
}

void ggnn_tf32_nt_take(ws_alloc& a, const ggnn_ctx* ctx, ggnn_nt_weights* w) {
  const size_t hp = ctx->hp;
  w->We = a.take<float>((size_t)ctx->two_e * hp * hp);
  w->Wg = a.take<float>(4 * hp * hp);
  w->Wc = a.take<float>(2 * hp * hp);
}

// mask != null: this step's effective weights W_src * mask * scale (edge-weight dropout), formed inside the pack
int ggnn_tf32_pack_nt_edge(ggnn_ctx* ctx, const float* W_src, const float* mask, float scale, ggnn_nt_weights* w,
                           cudaStream_t st) {
  const int h = ctx->h, hp = ctx->hp, two_e = ctx->two_e;
  packt::blocks4_kernel<<<ctx->sm_count * 8, 256, 0, st>>>(W_src, mask, scale, h, w->We, two_e, 1, 1, h, hp);
  GGNN_LAUNCHED(ctx);
  return make_weight_map(ctx, &w->tmE, w->We, (int64_t)two_e * hp, hp, &w->rowsE, hp);
}

int ggnn_tf32_pack_nt_gru(ggnn_ctx* ctx, ggnn_nt_weights* w, cudaStream_t st) {
  const int h = ctx->h, hp = ctx->hp;
  packt::blocks4_kernel<<<ctx->sm_count * 2, 256, 0, st>>>(ctx->Wg, nullptr, 1.f, 2 * h, w->Wg, 1, 2, 2, h, hp);
  GGNN_LAUNCHED(ctx);
  packt::blocks4_kernel<<<ctx->sm_count * 2, 256, 0, st>>>(ctx->Wc, nullptr, 1.f, h, w->Wc, 1, 2, 1, h, hp);
  GGNN_LAUNCHED(ctx);
  GGNN_TRY(make_weight_map(ctx, &w->tmG, w->Wg, 2 * hp, 2 * hp, &w->rowsG, 2 * hp));
  return make_weight_map(ctx, &w->tmC, w->Wc, 2 * hp, hp, &w->rowsC, 2 * hp);
}

int ggnn_tf32_nt(ggnn_ctx* ctx, const float* A1, const float* A2, int lda, const int32_t* gather, int64_t M,
                 unsigned mtiles, const float* Bt, const CUtensorMap& tm, int tm_rows, int Kp, int Np,
                 const int32_t* tile_type, const int32_t* n_tiles, float* out, int ldc,
                 const int32_t* c_rowmap, cudaStream_t st) {
  if (mtiles == 0) return GGNN_OK;
  tparams p{};
  p.A1 = A1; p.A2 = A2; p.lda1 = lda; p.lda2 = lda; p.gather = gather;
  p.Bt = Bt; p.h = ctx->h; p.hp = ctx->hp; p.Kp = Kp; p.Np = Np; p.M = M;
  p.tile_type = tile_type; p.n_tiles = n_tiles;
  p.out0 = out; p.ldc = ldc; p.c_rowmap = c_rowmap;
  p.b_bytes = tm_rows * BKF * 4;
  return launch<MODE_NT>(ctx, p, tm, mtiles, st);
}

// edge weights of `slot` re-packed from W_src (dropout: this step's effective weights; null = the slot's own)
int ggnn_tf32_repack_edge(ggnn_ctx* ctx, int slot, const float* W_src, cudaStream_t st) {
  ggnn_edge_weights& ew = ctx->edge[slot];
  packt::launch_edge_t<packt::OUT_TF32>(W_src ? W_src : ew.W, nullptr, 1.f, ew.Wt_tf32, ctx->two_e, ctx->h, ctx->hp, st);
  GGNN_LAUNCHED(ctx);
  return GGNN_OK;
}

bool ggnn_tf32_supported(const ggnn_ctx* ctx) { return ctx->h % 4 == 0 && ctx->hp >= 32; }

// h (any FP32) -> TF32-rounded copy; n = number of floats (multiple of 4). src may equal dst.
int ggnn_tf32_round_copy(ggnn_ctx* ctx, const float* src, float* dst, int64_t n, cudaStream_t st) {
  if (n <= 0) return GGNN_OK;
  prof_scope ps(ctx, GGNN_STAGE_OTHER, st);
  round_copy_kernel<<<ctx->sm_count * 8, 256, 0, st>>>(reinterpret_cast<const float4*>(src),
                                                      reinterpret_cast<float4*>(dst), n / 4);
  GGNN_LAUNCHED(ctx);
  return GGNN_OK;
}

int ggnn_tf32_pack_weights(ggnn_ctx* ctx, int slot, bool with_gru, cudaStream_t st) {
  const int h = ctx->h, hp = ctx->hp, two_e = ctx->two_e;
  ggnn_edge_weights& ew = ctx->edge[slot];
  GGNN_TRY(realloc_f(ctx, ew.Wt_tf32, (size_t)two_e * hp * hp));
  packt::launch_edge_t<packt::OUT_TF32>(ew.W, nullptr, 1.f, ew.Wt_tf32, two_e, h, hp, st);
  GGNN_LAUNCHED(ctx);
  GGNN_TRY(make_weight_map(ctx, &ew.tm_Wt, ew.Wt_tf32, (int64_t)two_e * hp, hp, &ew.tm_rows, hp));
  const bool strict_tc = hp <= 128;      // residual packs of the strict (3xTF32) mode, strict_tc.cu
  if (strict_tc) {
    GGNN_TRY(realloc_f(ctx, ew.Wt_tf32_lo, (size_t)two_e * hp * hp));
    packt::launch_edge_t<packt::OUT_TF32_LO>(ew.W, nullptr, 1.f, ew.Wt_tf32_lo, two_e, h, hp, st);
    GGNN_LAUNCHED(ctx);
    int rows_lo;
    GGNN_TRY(make_weight_map(ctx, &ew.tm_Wt_lo, ew.Wt_tf32_lo, (int64_t)two_e * hp, hp, &rows_lo, hp));
    if (hp == 128) {     // both packs of a type side by side: strict_tc.cu fetches them as ONE 32 KB box per K chunk
      GGNN_TRY(realloc_f(ctx, ew.Wt_tf32_hl, (size_t)two_e * 2 * hp * hp));
      const size_t blk = (size_t)hp * hp * 4;
      GGNN_CUDA(ctx, cudaMemcpy2DAsync(ew.Wt_tf32_hl, 2 * blk, ew.Wt_tf32, blk, blk, two_e, cudaMemcpyDeviceToDevice, st));
      GGNN_CUDA(ctx, cudaMemcpy2DAsync(ew.Wt_tf32_hl + (size_t)hp * hp, 2 * blk, ew.Wt_tf32_lo, blk, blk, two_e, cudaMemcpyDeviceToDevice, st));
      GGNN_TRY(ggnn_make_tmap_2d(ctx, &ew.tm_Wt_hl, ew.Wt_tf32_hl, (int64_t)two_e * 2 * hp, hp, 2 * hp));
    }
  }
  if (with_gru) {
    GGNN_TRY(realloc_f(ctx, ctx->Wgt_tf32, (size_t)4 * hp * hp));
    GGNN_TRY(realloc_f(ctx, ctx->Wct_tf32, (size_t)2 * hp * hp));
    GGNN_TRY(realloc_f(ctx, ctx->bg_pad, (size_t)2 * hp));
    GGNN_TRY(realloc_f(ctx, ctx->bc_pad, (size_t)hp));
    pack_gru_kernel<<<ctx->sm_count * 2, 256, 0, st>>>(ctx->Wg, ctx->Wgt_tf32, 2, h, hp, 0);
    GGNN_LAUNCHED(ctx);
    pack_gru_kernel<<<ctx->sm_count * 2, 256, 0, st>>>(ctx->Wc, ctx->Wct_tf32, 1, h, hp, 0);
    GGNN_LAUNCHED(ctx);
    if (strict_tc) {
      GGNN_TRY(realloc_f(ctx, ctx->Wgt_tf32_lo, (size_t)4 * hp * hp));
      GGNN_TRY(realloc_f(ctx, ctx->Wct_tf32_lo, (size_t)2 * hp * hp));
      pack_gru_kernel<<<ctx->sm_count * 2, 256, 0, st>>>(ctx->Wg, ctx->Wgt_tf32_lo, 2, h, hp, 1);
      GGNN_LAUNCHED(ctx);
      pack_gru_kernel<<<ctx->sm_count * 2, 256, 0, st>>>(ctx->Wc, ctx->Wct_tf32_lo, 1, h, hp, 1);
      GGNN_LAUNCHED(ctx);
      int r_;
      // strict_tc.cu runs the gate contraction as two N = hp halves: boxes of hp rows
      GGNN_TRY(ggnn_make_tmap_2d(ctx, &ctx->tm_Wg_lo, ctx->Wgt_tf32_lo, 2 * hp, 2 * hp, hp));
      GGNN_TRY(ggnn_make_tmap_2d(ctx, &ctx->tm_Wg_half, ctx->Wgt_tf32, 2 * hp, 2 * hp, hp));
      GGNN_TRY(make_weight_map(ctx, &ctx->tm_Wc_lo, ctx->Wct_tf32_lo, hp, 2 * hp, &r_, hp));
      ctx->hl_ok = false;
      if (hp == 128) {
        GGNN_TRY(realloc_f(ctx, ctx->Wgt_tf32_hl, (size_t)8 * hp * hp));
        GGNN_TRY(realloc_f(ctx, ctx->Wct_tf32_hl, (size_t)4 * hp * hp));
        const size_t half = (size_t)hp * 2 * hp * 4;         // one N = hp half of the gate weights, or all of Wc
        GGNN_CUDA(ctx, cudaMemcpy2DAsync(ctx->Wgt_tf32_hl, 2 * half, ctx->Wgt_tf32, half, half, 2, cudaMemcpyDeviceToDevice, st));
        GGNN_CUDA(ctx, cudaMemcpy2DAsync(ctx->Wgt_tf32_hl + (size_t)hp * 2 * hp, 2 * half, ctx->Wgt_tf32_lo, half, half, 2, cudaMemcpyDeviceToDevice, st));
        GGNN_CUDA(ctx, cudaMemcpyAsync(ctx->Wct_tf32_hl, ctx->Wct_tf32, half, cudaMemcpyDeviceToDevice, st));
        GGNN_CUDA(ctx, cudaMemcpyAsync(ctx->Wct_tf32_hl + (size_t)hp * 2 * hp, ctx->Wct_tf32_lo, half, cudaMemcpyDeviceToDevice, st));
        GGNN_TRY(ggnn_make_tmap_2d(ctx, &ctx->tm_Wg_hl, ctx->Wgt_tf32_hl, 4 * hp, 2 * hp, 2 * hp));
        GGNN_TRY(ggnn_make_tmap_2d(ctx, &ctx->tm_Wc_hl, ctx->Wct_tf32_hl, 2 * hp, 2 * hp, 2 * hp));
        ctx->hl_ok = true;
      }
    }
    pack_bias_kernel<<<(2 * hp + 255) / 256, 256, 0, st>>>(ctx->bg, ctx->bg_pad, 2, h, hp);
    GGNN_LAUNCHED(ctx);
    pack_bias_kernel<<<(hp + 255) / 256, 256, 0, st>>>(ctx->bc, ctx->bc_pad, 1, h, hp);
    GGNN_LAUNCHED(ctx);
    GGNN_TRY(make_weight_map(ctx, &ctx->tm_Wg, ctx->Wgt_tf32, 2 * hp, 2 * hp, &ctx->tm_rows_g, 2 * hp));
    GGNN_TRY(make_weight_map(ctx, &ctx->tm_Wc, ctx->Wct_tf32, hp, 2 * hp, &ctx->tm_rows_c, hp));
    if (2 * hp <= 256) GGNN_TRY(ggnn_make_tmap_2d(ctx, &ctx->tm_Wg256, ctx->Wgt_tf32, 2 * hp, 2 * hp, 2 * hp));
  }
  return GGNN_OK;
}

// One timestep.  The state buffer holds TF32-rounded values between steps (so every MMA operand
// is exact in TF32 and can be copied asynchronously); the last step stores the unrounded FP32.
int ggnn_step_tf32(ggnn_ctx* ctx, int slot, const ggnn_plan& plan, const ggnn_step_bufs& bufs,
                   float* h_state, bool last_step, cudaStream_t st) {
  const int h = ctx->h, hp = ctx->hp;
  const ggnn_edge_weights& ew = ctx->edge[slot];
  if (plan.N == 0) return GGNN_OK;
  const bool fused = ggnn_gru_fused_supported(ctx) && !ctx->no_fused && !bufs.tape_r;   // the tape needs x, r, u, c
  if (fused) {
    GGNN_TRY(ggnn_msg_tc_step(ctx, slot, plan, h_state, bufs.msg, st));
  } else if (plan.max_tiles > 0 && plan.nnz > 0) {
    prof_scope ps(ctx, GGNN_STAGE_MSG, st);
    tparams p{};
    p.A1 = h_state; p.gather = plan.msg_src;
    p.Bt = ew.Wt_tf32; p.bias = ew.B;
    p.h = h; p.hp = hp; p.Kp = hp; p.Np = hp;
    p.M = plan.max_tiles * GGNN_TILE_M;
    p.tile_type = plan.tile_type; p.n_tiles = plan.n_tiles;
    p.out0 = bufs.msg;
    // messages as FP16 (the 11-bit significand of a TF32 operand, half the bytes for the aggregation), as in the
    // fused kernels; the training tape keeps FP32 messages
    if (!bufs.tape_r) p.out_half = reinterpret_cast<__half*>(bufs.msg);
    p.b_bytes = ew.tm_rows * BKF * 4;
    GGNN_TRY(launch<MODE_MSG>(ctx, p, ew.tm_Wt, (unsigned)plan.max_tiles, st));
  }
  if (fused) return ggnn_gru_fused_step(ctx, plan, bufs.msg, h_state, last_step, st);
  if (!bufs.tape_r) GGNN_TRY(ggnn_aggregate_f16(ctx, plan, bufs.msg, bufs.acts, h, /*round_tf32=*/true, st));
  else GGNN_TRY(ggnn_aggregate(ctx, plan, bufs.msg, bufs.acts, h, /*round_tf32=*/true, st));
  const unsigned mt = (unsigned)ceil_div64(plan.N, BM);
  {
    prof_scope ps(ctx, GGNN_STAGE_GRU_GATES, st);
    tparams p{};
    p.A1 = bufs.acts; p.A2 = h_state;
    p.Bt = ctx->Wgt_tf32; p.bias = ctx->bg_pad;
    p.h = h; p.hp = hp; p.Kp = 2 * hp; p.Np = 2 * hp; p.M = plan.N;
    p.out0 = bufs.rh; p.out1 = bufs.u; p.hstate = h_state; p.tape_r = bufs.tape_r;
    p.b_bytes = ctx->tm_rows_g * BKF * 4;
    GGNN_TRY(launch<MODE_GATES>(ctx, p, ctx->tm_Wg, mt, st));
  }
  {
    prof_scope ps(ctx, GGNN_STAGE_GRU_CAND, st);
    tparams p{};
    p.A1 = bufs.acts; p.A2 = bufs.rh;
    p.Bt = ctx->Wct_tf32; p.bias = ctx->bc;
    p.h = h; p.hp = hp; p.Kp = 2 * hp; p.Np = hp; p.M = plan.N;
    p.out0 = h_state; p.ubuf = bufs.u; p.tape_c = bufs.tape_c;
    p.b_bytes = ctx->tm_rows_c * BKF * 4;
    p.round_out = last_step ? 0 : 1;
    GGNN_TRY(launch<MODE_CAND>(ctx, p, ctx->tm_Wc, mt, st));
  }
  return GGNN_OK;
}
