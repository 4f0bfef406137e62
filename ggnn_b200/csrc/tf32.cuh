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
