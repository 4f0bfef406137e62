// train.cuh -- training tape + backward of the propagation path (strict FP32), declarations.
#pragma once
#include "plan.cuh"
#include "tf32.cuh"

// per-step saved activations, each [N, h]: state before the step, aggregated messages, gates, candidate
enum { TAPE_H = 0, TAPE_X = 1, TAPE_R = 2, TAPE_U = 3, TAPE_C = 4, TAPE_ARRAYS = 5 };

struct ggnn_tape {
  bool valid = false;
  int T = 0, slot = 0;
  bool tf32 = false;           // forward ran in GGNN_PREC_TF32: tape holds TF32-rounded operands, backward uses the tensor cores
  ggnn_plan plan;              // lives in the step arena of the forward call
  float* acts = nullptr;       // [T][TAPE_ARRAYS][N*h]
  float* at(int t, int which) const { return acts + ((size_t)t * TAPE_ARRAYS + which) * (size_t)plan.N * h; }
  int h = 0;
  // A8: training-time dropout with caller-supplied 0/1 masks (null = keep everything)
  const float* w_mask = nullptr;      // [T, 2E, h, h]  tf.nn.dropout on edge_weights, fresh per step (:397-403)
  const float* state_mask = nullptr;  // [T, N, h]      DropoutWrapper(state_keep_prob) on the new state (:239-240)
  float keep = 1.f;
  float* w_eff = nullptr;             // [2E, h, h] scratch: W * mask_t / keep
};

// scratch the backward needs, in floats / int32s (for the bump allocator)
struct ggnn_bwd_scratch {
  float *dh_a, *dh_b, *dx, *dzc, *rh, *dzg, *dxr;
  int32_t* msg_tgt;
  float* gw_tmp;      // [2E, h, h] per-step dW before the mask is applied (only with w_mask)
  ggnn_nt_weights ntw; // TF32 backward: weights in their own layout, rounded + padded (only when tape.tf32)
};
void ggnn_bwd_take_nt_weights(ws_alloc& a, const ggnn_ctx* ctx, ggnn_bwd_scratch* s);
void ggnn_bwd_take_scratch(ws_alloc& a, const ggnn_plan& plan, int h, int two_e, bool with_w_mask, ggnn_bwd_scratch* s);

// elementwise helpers of the dropout path
int ggnn_mask_scale(ggnn_ctx* ctx, const float* x, const float* mask, float scale, float* out, int64_t n, bool accumulate, cudaStream_t st);

// Philox4x32-10 0/1 mask: out[i] = 1 where uniform(seed, stream, i) < keep
int ggnn_dropout_mask_fill(ggnn_ctx* ctx, float* out, int64_t n, unsigned long long seed, unsigned long long stream,
                           float keep, cudaStream_t st);

// flat gradient layout: [dW 2E*h*h | dB 2E*h | dWg 2h*2h | dbg 2h | dWc 2h*h | dbc h]
size_t ggnn_grad_count(int two_e, int h);

// accumulates into `grads` (caller zeroes it); writes d_h0 [N, h]
int ggnn_backward_fp32(ggnn_ctx* ctx, const ggnn_tape& tape, const ggnn_bwd_scratch& s,
                       const float* d_hT, float* d_h0, float* grads, cudaStream_t st);

// N4: one variable of the flat buffer: clip_by_norm(grad, clip) then Adam; `partials` (GGNN_SUMSQ_SLOT floats)
// receives the per-block partial sums of ||grad||^2, reduced in a fixed order (bit-repeatable clip scale)
#define GGNN_SUMSQ_SLOT 2048
int ggnn_adam_segment(ggnn_ctx* ctx, float* param, const float* grad, float* m, float* v, int64_t n,
                      float* partials, float clip, float lr_t, float b1, float b2, float eps, cudaStream_t st);
