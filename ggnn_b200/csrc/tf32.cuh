// tf32.cuh -- tcgen05 (kind::tf32) timestep: declarations.
#pragma once
#include "plan.cuh"

// (re)pack the weights of `slot` (and the GRU cell when with_gru) into the K-major,
// TF32-rounded, padded layouts the tcgen05 kernels read.
int ggnn_tf32_pack_weights(ggnn_ctx* ctx, int slot, bool with_gru, cudaStream_t st);

// one timestep in TF32 mode; same buffers as the FP32 step
int ggnn_step_tf32(ggnn_ctx* ctx, int slot, const ggnn_plan& plan, const ggnn_step_bufs& bufs,
                   float* h_state, bool last_step, cudaStream_t st);

// dst = TF32-rounded src (n floats, multiple of 4); src may equal dst
int ggnn_tf32_round_copy(ggnn_ctx* ctx, const float* src, float* dst, int64_t n, cudaStream_t st);

bool ggnn_tf32_supported(const ggnn_ctx* ctx);

// 2-D FP32 row-major tensor [rows, cols] -> TMA map with a 32-column (128-byte) x box_rows box,
// 128B swizzle (the K-major SWIZZLE_128B operand layout of tc.cuh)
int ggnn_make_tmap_2d(ggnn_ctx* ctx, CUtensorMap* tm, const float* base, int64_t rows, int64_t cols,
                      int box_rows);

// fused aggregation + GRU timestep (gru_fused.cu), hidden size padded to 128
bool ggnn_gru_fused_supported(const ggnn_ctx* ctx);
int ggnn_gru_fused_step(ggnn_ctx* ctx, const ggnn_plan& plan, const float* msg, float* h_state,
                        bool last_step, cudaStream_t st);

// FP16 row-major [rows, cols] -> TMA map, box_cols x box_rows box (box_cols * 2 = 128 bytes), 128B swizzle
int ggnn_make_tmap_2d_f16(ggnn_ctx* ctx, CUtensorMap* tm, const void* base, int64_t rows, int64_t cols,
                          int box_cols, int box_rows);

// persistent TF32 message transform with FP16 output rows of pitch 128 (msg_tc.cu), hp == 128
bool ggnn_msg_tc_supported(const ggnn_ctx* ctx);
int ggnn_msg_tc_step(ggnn_ctx* ctx, int slot, const ggnn_plan& plan, const float* h_state,
                     void* msg_half, cudaStream_t st);

// ---- backward on the tensor cores (train_tf32.cu) -----------------------------------------
// the weights in their OWN row-major layout, TF32-rounded and padded to hp: B^T of C = A @ B^T is then the
// K-major operand the forward kernel already reads
struct ggnn_nt_weights {
  float *We, *Wg, *Wc;          // [2E*hp, hp], [2hp, 2hp], [2hp, hp]
  CUtensorMap tmE, tmG, tmC;
  int rowsE, rowsG, rowsC;
};
size_t ggnn_tf32_nt_weight_floats(const ggnn_ctx* ctx);
void ggnn_tf32_nt_take(ws_alloc& a, const ggnn_ctx* ctx, ggnn_nt_weights* w);
int ggnn_tf32_pack_nt_edge(ggnn_ctx* ctx, const float* W_src, const float* mask, float scale, ggnn_nt_weights* w,
                           cudaStream_t st);
int ggnn_tf32_pack_nt_gru(ggnn_ctx* ctx, ggnn_nt_weights* w, cudaStream_t st);
// out[crow, (n / hp) * h + n % hp] (=|+=) sum_k A[row, k] * Bt[n, k];  A = [A1 | A2] (K segments of hp), rows
// optionally gathered; crow = row (store) or c_rowmap[row] (atomic add); grouped by tile type when tile_type != null
int ggnn_tf32_nt(ggnn_ctx* ctx, const float* A1, const float* A2, int lda, const int32_t* gather, int64_t M,
                 unsigned mtiles, const float* Bt, const CUtensorMap& tm, int tm_rows, int Kp, int Np,
                 const int32_t* tile_type, const int32_t* n_tiles, float* out, int ldc,
                 const int32_t* c_rowmap, cudaStream_t st);
int ggnn_tf32_repack_edge(ggnn_ctx* ctx, int slot, const float* W_src, cudaStream_t st);
// C[type][K1, N] += X[R, K1]^T @ D[R, N] (rows optionally gathered; grouped by message tile type)
int ggnn_tf32_tn(ggnn_ctx* ctx, const float* X, int ldx, const int32_t* x_rowmap, int K1, const float* D, int ldd,
                 const int32_t* d_rowmap, int N, int64_t R, int rows_per_cta, float* C, int ldc,
                 const int32_t* tile_type, const int32_t* n_tiles, int64_t c_type_stride, cudaStream_t st);

// GGNN_PREC_FP32 on the tensor cores (3xTF32 split, strict_tc.cu), hidden sizes up to 128; GGNN_FP32_FFMA=1 keeps the FFMA
// kernels of step_fp32.cu
bool ggnn_fp32_tc_enabled(const ggnn_ctx* ctx);
int ggnn_step_fp32_tc(ggnn_ctx* ctx, int slot, const ggnn_plan& plan, const ggnn_step_bufs& bufs, float* h_state,
                      cudaStream_t st);
