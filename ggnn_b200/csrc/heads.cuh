// heads.cuh -- front end (N1) and output heads + decode (N2): declarations.
#pragma once
#include "common.cuh"

int ggnn_embed_launch(ggnn_ctx* ctx, const int32_t* idx, int64_t N, const float* loc_tab, int loc_rows,
                      int loc_dim, const float* pos_tab, int pos_rows, int pos_dim, const float* word_tab,
                      int word_rows, int word_dim, int h, float* h0, int* bad, cudaStream_t st);
int ggnn_decode_launch(ggnn_ctx* ctx, const float* logits, int ld, int64_t N, int n_head, int n_lab,
                       int lab_off, const int32_t* n_active, int32_t* pred, float* p_head, float* p_lab,
                       cudaStream_t st);
// out[M, N] = [A1 | A2] @ B + bias   (strict FP32 FFMA; A1, A2 [M, K1] / [M, K2] with row pitch lda;
// B [K1+K2, N] row-major, N % 4 == 0)   -- step_fp32.cu
int ggnn_sgemm_concat(ggnn_ctx* ctx, const float* A1, const float* A2, int K1, int K2, int lda,
                      const float* B, const float* bias, int N, int64_t M, float* out, cudaStream_t st);
