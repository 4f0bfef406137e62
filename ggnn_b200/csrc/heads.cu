// heads.cu -- the rows either side of the propagation path, evaluation mode.
//
// N1  get_initial_node_representation (chem_tensorflow_dense.py:264-310): four embedding lookups
//     concat(loc[80] | pos[50] | word[100] | loc[head][80]) zero-padded to h.  Dropouts are
//     identities at evaluation (:800-801).
// N2  gated_regression, btb branch (:439-505): logits = [h_T | h_0] @ W + b for the 150 head
//     positions and for the output labels (utils.MLP with no hidden layer returns the
//     pre-activation, utils.py:77-84), softmax over each group, the reference's masks
//     (get_mask :633-661: row t and head position s must both be < n_active), and the decode of
//     adj_mat_to_target (:106-131): FIRST maximum of the masked probabilities.
#include "plan.cuh"
#include "heads.cuh"

namespace {

// one warp per node; lanes stride over the h output columns
__global__ void __launch_bounds__(256)
embed_kernel(const int32_t* __restrict__ idx, int64_t N, const float* __restrict__ loc_tab, int loc_rows,
             int loc_dim, const float* __restrict__ pos_tab, int pos_rows, int pos_dim,
             const float* __restrict__ word_tab, int word_rows, int word_dim, int h,
             float* __restrict__ h0, int* __restrict__ bad) {
  const int lane = threadIdx.x & 31;
  const int64_t n = ((int64_t)blockIdx.x * blockDim.x + threadIdx.x) >> 5;
  if (n >= N) return;
  const int il = idx[n * 4 + 0], ip = idx[n * 4 + 1], iw = idx[n * 4 + 2], ih = idx[n * 4 + 3];
  if (il < 0 || il >= loc_rows || ip < 0 || ip >= pos_rows || iw < 0 || iw >= word_rows || ih < 0 || ih >= loc_rows) {
    if (lane == 0) atomicAdd(bad, 1);
    return;
  }
  const int o1 = loc_dim, o2 = o1 + pos_dim, o3 = o2 + word_dim, o4 = o3 + loc_dim;
  for (int c = lane; c < h; c += 32) {
    float v = 0.f;
    if (c < o1) v = loc_tab[(size_t)il * loc_dim + c];
    else if (c < o2) v = pos_tab[(size_t)ip * pos_dim + (c - o1)];
    else if (c < o3) v = word_tab[(size_t)iw * word_dim + (c - o2)];
    else if (c < o4) v = loc_tab[(size_t)ih * loc_dim + (c - o3)];
    h0[n * h + c] = v;
  }
}

// warp-wide helpers
__device__ __forceinline__ float warp_max(float v) {
#pragma unroll
  for (int o = 16; o > 0; o >>= 1) v = fmaxf(v, __shfl_xor_sync(0xffffffffu, v, o));
  return v;
}
__device__ __forceinline__ float warp_sum(float v) {
#pragma unroll
  for (int o = 16; o > 0; o >>= 1) v += __shfl_xor_sync(0xffffffffu, v, o);
  return v;
}

// softmax over `n` logits of one row (exp(x - max) / sum, as tf.nn.softmax), then the FIRST index
// whose probability equals the maximum among the first `n_valid` entries.  Optionally stores the
// probabilities.  Returns -1 when n_valid == 0.
__device__ int row_softmax_argmax(const float* __restrict__ x, int n, int n_valid, float* __restrict__ probs,
                                  int lane) {
  float mx = -INFINITY;
  for (int c = lane; c < n; c += 32) mx = fmaxf(mx, x[c]);
  mx = warp_max(mx);
  float s = 0.f;
  for (int c = lane; c < n; c += 32) s += expf(x[c] - mx);
  s = warp_sum(s);
  float best = -1.f;
  int best_i = 0x7fffffff;
  for (int c = lane; c < n; c += 32) {
    const float p = expf(x[c] - mx) / s;
    if (probs) probs[c] = c < n_valid ? p : 0.f;
    if (c < n_valid && p > best) { best = p; best_i = c; }      // ascending c: first maximum per lane
  }
#pragma unroll
  for (int o = 16; o > 0; o >>= 1) {
    const float ob = __shfl_xor_sync(0xffffffffu, best, o);
    const int oi = __shfl_xor_sync(0xffffffffu, best_i, o);
    if (ob > best || (ob == best && oi < best_i)) { best = ob; best_i = oi; }
  }
  return n_valid > 0 ? best_i : -1;
}

// one warp per node row: logits [N, ld] = [heads 0..n_head) | pad | labels lab_off..lab_off+n_lab)
__global__ void __launch_bounds__(256)
decode_kernel(const float* __restrict__ logits, int ld, int64_t N, int n_head, int n_lab, int lab_off,
              const int32_t* __restrict__ n_active, int32_t* __restrict__ pred,
              float* __restrict__ p_head, float* __restrict__ p_lab) {
  const int lane = threadIdx.x & 31;
  const int64_t n = ((int64_t)blockIdx.x * blockDim.x + threadIdx.x) >> 5;
  if (n >= N) return;
  const int na = n_active[n];                // 0 for a padded row: everything masked
  const float* row = logits + n * ld;
  const int nh = na < n_head ? na : n_head;
  const int hd = row_softmax_argmax(row, n_head, nh, p_head ? p_head + n * n_head : nullptr, lane);
  const int lb = row_softmax_argmax(row + lab_off, n_lab, na > 0 ? n_lab : 0,
                                    p_lab ? p_lab + n * n_lab : nullptr, lane);
  if (lane == 0) { pred[n * 2 + 0] = hd; pred[n * 2 + 1] = lb; }
}

}  // namespace

int ggnn_embed_launch(ggnn_ctx* ctx, const int32_t* idx, int64_t N, const float* loc_tab, int loc_rows,
                      int loc_dim, const float* pos_tab, int pos_rows, int pos_dim, const float* word_tab,
                      int word_rows, int word_dim, int h, float* h0, int* bad, cudaStream_t st) {
  if (N == 0) return GGNN_OK;
  embed_kernel<<<(unsigned)ceil_div64(N * 32, 256), 256, 0, st>>>(idx, N, loc_tab, loc_rows, loc_dim, pos_tab,
                                                                 pos_rows, pos_dim, word_tab, word_rows,
                                                                 word_dim, h, h0, bad);
  GGNN_LAUNCHED(ctx);
  return GGNN_OK;
}

int ggnn_decode_launch(ggnn_ctx* ctx, const float* logits, int ld, int64_t N, int n_head, int n_lab,
                       int lab_off, const int32_t* n_active, int32_t* pred, float* p_head, float* p_lab,
                       cudaStream_t st) {
  if (N == 0) return GGNN_OK;
  decode_kernel<<<(unsigned)ceil_div64(N * 32, 256), 256, 0, st>>>(logits, ld, N, n_head, n_lab, lab_off,
                                                                  n_active, pred, p_head, p_lab);
  GGNN_LAUNCHED(ctx);
  return GGNN_OK;
}
