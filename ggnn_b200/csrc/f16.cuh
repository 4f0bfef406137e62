// f16.cuh -- GGNN_PREC_F16 timestep (tcgen05 kind::f16, FP16 operands in 16-bit containers,
// FP32 accumulate), hidden size padded to 128: declarations.
#pragma once
#include "plan.cuh"

bool ggnn_f16_supported(const ggnn_ctx* ctx);
// the plan may leave the word-order rows of regular graphs to the GRU kernel (folded weights are packed, not switched off)
bool ggnn_f16_word_fold_ready(const ggnn_ctx* ctx, int slot);
// pack the weights of `slot` (and the GRU cell) into K-major FP16 tiles + TMA maps
int ggnn_f16_pack_weights(ggnn_ctx* ctx, int slot, bool with_gru, cudaStream_t st);
// state16 [N, 128] FP16 = h [N, h] FP32 rounded to nearest, columns >= h zero; max|h| is
// atomicMax-ed (as float bits) into *absmax_bits when given
int ggnn_f16_state_in(ggnn_ctx* ctx, const float* h, void* state16, int64_t N, unsigned* absmax_bits,
                      cudaStream_t st);
// max|x| over n floats, accumulated into *out_bits (float bits, atomicMax; caller zeroes)
int ggnn_absmax(ggnn_ctx* ctx, const float* x, int64_t n, unsigned* out_bits, cudaStream_t st);
// max over edge types and output columns of the column's L1 norm sum_k |W[e][k][n]| (float bits, atomicMax; caller zeroes)
int ggnn_colnorm_max(ggnn_ctx* ctx, const float* W, int two_e, int h, unsigned* out_bits, cudaStream_t st);
// one timestep: msg16 [max_tiles*128, 128] FP16 scratch; state16 in/out; on the last step the
// unrounded FP32 state is also written to h_out [N, h]
int ggnn_step_f16(ggnn_ctx* ctx, int slot, const ggnn_plan& plan, void* msg16, void* state16,
                  float* h_out, bool last_step, cudaStream_t st);

// ---- gru_pair.cu: the GRU kernel of the h <= 128 timestep on CTA pairs with resident weights (the default; GGNN_GRU_PAIR=0: gru_f16_kernel) ----
bool ggnn_gru_pair_enabled(ggnn_ctx* ctx, int slot);
int ggnn_gru_pair_pack(ggnn_ctx* ctx, int slot, bool with_gru, cudaStream_t st);   // after ggnn_f16_pack_weights' packs
int ggnn_gru_pair_launch(ggnn_ctx* ctx, int slot, const ggnn_plan& plan, const void* msg16, void* state16, float* h_out,
                         cudaStream_t st);

// ---- hidden sizes above 128 (wide16.cu): FP16 buffers of pitch ctx->hw16 = ceil64(h) <= 512 ----
bool ggnn_w16_supported(const ggnn_ctx* ctx);
int ggnn_w16_pack_weights(ggnn_ctx* ctx, int slot, bool with_gru, cudaStream_t st);
int ggnn_w16_state_in(ggnn_ctx* ctx, const float* h, void* state16, int64_t N, unsigned* absmax_bits,
                      cudaStream_t st);
// one timestep; msg16 [max_tiles*128, hw16], acts16 / rh16 / u16 [N, hw16] FP16 scratch; state16 in/out; on the
// last step the FP32 state is also written to h_out [N, h]
// tape != null (training forward): the unfused kernels also store x (FP16-rounded), r, u, c of this step as FP32 [N, h]
struct ggnn_w16_tape { float *x, *r, *u, *c; };
int ggnn_step_w16(ggnn_ctx* ctx, int slot, const ggnn_plan& plan, void* msg16, void* acts16, void* rh16, void* u16,
                  void* state16, float* h_out, bool last_step, cudaStream_t st, const ggnn_w16_tape* tape = nullptr);
int ggnn_w16_state_out(ggnn_ctx* ctx, const void* state16, float* dst, int64_t N, cudaStream_t st);
int ggnn_w16_state_mask(ggnn_ctx* ctx, void* state16, const float* mask, float scale, int64_t N, cudaStream_t st);
int ggnn_w16_repack_edge(ggnn_ctx* ctx, int slot, const float* W_src, const float* mask, float scale, cudaStream_t st);
// message transform alone through wgemm16_kernel<MSG> (hw = 128 too) and the padded biases it reads
int ggnn_w16_msg(ggnn_ctx* ctx, int slot, const ggnn_plan& plan, void* msg16, void* state16, int hw, cudaStream_t st);
int ggnn_w16_pack_bias(ggnn_ctx* ctx, int slot, int hw, cudaStream_t st);
