/* ggnn_b200.h -- C ABI of libggnn_b200.so (sm_100a).
 *
 * Drop-in boundary for ONE path of crismolav/ggnn: the GGNN propagation step
 *   DenseGGNNChemModel.compute_final_node_representations
 *       /root/reference/chem_tensorflow_dense.py:312-340   (driver)
 *       /root/reference/chem_tensorflow_dense.py:391-437   (compute_timestep_fast)
 *       /root/reference/chem_tensorflow_dense.py:350-389   (compute_timestep_normal)
 *       /root/reference/chem_tensorflow_dense.py:237-241,333 (TF GRUCell)
 *   and its edge-list ("sparse") form, which the reference only declares
 *       /root/reference/chem_tensorflow.py:522-523.
 *
 * The reference exposes no FFI for this path (it is a Python method building TF
 * ops); these entry points are what a ctypes binding for it would bind -- see
 * INTEGRATION.md.  Plain pointers and sizes only; no torch / CUDA types.
 *
 * Conventions
 *   - every pointer is a DEVICE pointer on the ctx's device unless it says "host";
 *   - state and gradient buffers must be 16-byte aligned (the kernels move 16-byte vectors; any cudaMalloc /
 *     framework allocation is); a misaligned pointer is GGNN_E_ARG.  The adjacency may be element-aligned only
 *     (slower scalar path);
 *   - `stream` is a cudaStream_t passed as void* (NULL = legacy default stream);
 *   - all work is enqueued on `stream`; the calls that must learn a device-side
 *     count to size their plan (propagate_dense / propagate_csr / propagate_edges)
 *     synchronise `stream` once while planning and say so below; compact_adjacency
 *     and set_weights do not synchronise;
 *   - return 0 on success, a negative GGNN_E_* otherwise; never throws, never exits;
 *     ggnn_last_error() gives the text for the last failure on that ctx;
 *   - one ctx per (device, host thread); a ctx is not thread-safe; ctxs are independent;
 *   - the caller owns every input/output buffer until the stream work has finished;
 *     the ctx owns packed weights and scratch.
 */
#ifndef GGNN_B200_H
#define GGNN_B200_H

#include <stddef.h>
#include <stdint.h>

#ifdef __cplusplus
extern "C" {
#endif

#define GGNN_OK              0
#define GGNN_E_ARG          -1   /* null pointer, bad enum, negative size            */
#define GGNN_E_SHAPE        -2   /* sizes inconsistent with the weights / each other */
#define GGNN_E_CAPACITY     -3   /* an output buffer is too small                    */
#define GGNN_E_CUDA         -4   /* a CUDA call failed; see ggnn_last_error          */
#define GGNN_E_UNSUPPORTED  -5   /* valid request this build cannot serve            */
#define GGNN_E_STATE        -6   /* weights not set, etc.                            */

/* adjacency element type: the reference's placeholder is tf.float32
 * (chem_tensorflow_dense.py:192-193); uint8 is accepted as a 4x smaller twin. */
#define GGNN_ADJ_F32  0
#define GGNN_ADJ_U8   1

/* arithmetic mode of the message / GRU contractions */
#define GGNN_PREC_FP32  0   /* strict, <= 1e-5 norm-wise vs the FP64 oracle.  Hidden sizes up to 128: tcgen05 with every
                             * operand split in two TF32 halves (3 MMAs per product, FP32 accumulate; ~3e-6 at T = 8);
                             * above 128, and with GGNN_FP32_FFMA=1 in the environment: FP32 FFMA (~6e-7)           */
#define GGNN_PREC_TF32  1   /* tcgen05 kind::tf32, FP32 accumulate, <= 2e-3 norm-wise   */
#define GGNN_PREC_F16   2   /* tcgen05 kind::f16: operands rounded to FP16 (the same 11-bit significand
                             * as TF32) and held in 16-bit containers, FP32 accumulate, FP32 gate math,
                             * FP32 result; <= 2e-3 norm-wise.  Hidden size in (96, 512].  Needs
                             * max_(e,n) sum_k |W[e][k][n]| * max(max|h0|,1) + max|B| < 64 (largest column L1 norm
                             * of any edge weight matrix) so nothing leaves the FP16 range; otherwise the call
                             * returns GGNN_E_UNSUPPORTED (use GGNN_PREC_TF32).    */

/* weight slot: 0 = edge_weights / edge_biases (chem_tensorflow_dense.py:202-206),
 *              1 = edge_weights_fixed / edge_biases_fixed (:208-212), used when the
 *                  reference is called with fixed_ts (:400-403, 411-412).
 * The GRU cell is shared by both slots (:237-241). */
#define GGNN_SLOT_MAIN   0
#define GGNN_SLOT_FIXED  1

/* flags.  (There is no flag for compute_timestep_normal, chem_tensorflow_dense.py:350-389: a node's
 * messages are always summed in ascending (channel, source) order, which IS the loop form's accumulation
 * order, so both reference forms map to the same call; tests hold the result to the goldens of both.) */
#define GGNN_FLAG_SAVE_TAPE   2u  /* training forward: keep every step's activations in the ctx for
                                   * ggnn_propagate_backward.  All three arithmetic modes (GGNN_PREC_F16
                                   * needs a hidden size in (128, 512]); the batch must fit one chunk. */

typedef struct ggnn_ctx ggnn_ctx;

/* lifetime -------------------------------------------------------------------- */
int         ggnn_create(ggnn_ctx** out, int device);
void        ggnn_destroy(ggnn_ctx* ctx);
const char* ggnn_last_error(const ggnn_ctx* ctx);      /* host string, owned by ctx */
int         ggnn_version(void);

/* parameters ------------------------------------------------------------------
 * Replaces self.weights['edge_weights'|'edge_biases'|'node_gru'] (chem_tensorflow_dense.py
 * :202-212, 237-241).  Layouts are the reference's / TF's:
 *   W  [two_e, h, h]  (m = h_state @ W[e]),  B [two_e, h] or NULL (use_edge_bias False)
 *   Wg [2h, 2h] rows 0:h multiply the message, h:2h the state; cols 0:h = r, h:2h = u
 *   bg [2h], Wc [2h, h], bc [h].
 * The ctx copies and re-packs them; the caller's buffers are free after the stream
 * work completes.  For slot 1 the GRU pointers may be NULL (slot 0's cell is used). */
int ggnn_set_weights(ggnn_ctx* ctx, int slot, int two_e, int h,
                     const float* W, const float* B_or_null,
                     const float* Wg, const float* bg,
                     const float* Wc, const float* bc, void* stream);

/* A1: adjacency compaction -----------------------------------------------------
 * Dense adj [b, two_e, v, v] (A[b,e,target,source] != 0 <=> edge;
 * graph_to_adj_mat_bd, chem_tensorflow_dense.py:65-83) -> canonical CSR:
 *   rows = (b, e, target) row-major, rowptr int32 [b*two_e*v + 1], col int32 = source,
 * i.e. exactly np.argwhere(adj != 0) order.  *nnz_out_dev (device int64) receives nnz.
 * If nnz > col_capacity only the first col_capacity columns are written and
 * *nnz_out_dev still holds the true count (caller checks).  The count is carried in 64 bits: when it
 * is >= 2^31 the int32 rowptr / col are not valid (split the batch).  No synchronisation. */
int ggnn_compact_adjacency(ggnn_ctx* ctx, const void* adj, int adj_dtype,
                           int b, int two_e, int v,
                           int32_t* rowptr, int32_t* col, int64_t col_capacity,
                           int64_t* nnz_out_dev, void* stream);

/* The same compaction for callers that may not allocate device memory (code under CUDA graph capture, tight
 * serving loops, harnesses that forbid cudaMalloc).  NO COUNTERPART IN THE REFERENCE.
 * Same arguments, results and conventions as ggnn_compact_adjacency, but the scratch
 * (bitmask + scan partials) is a fixed device array inside the library (2 MB + 4 KB, held by every GPU that
 * loads the library until the process exits) instead of the ctx's grow-only arena: the call performs no
 * cudaMalloc / cudaFree and no other allocating call.
 *   - limits: b*two_e*v*v <= GGNN_STATIC_MAX_ELEMS and b*two_e*v <= GGNN_STATIC_MAX_ROWS; larger batches return
 *     GGNN_E_UNSUPPORTED (split them, or use ggnn_compact_adjacency);
 *   - ALL calls on a device share the one array, from every ctx and stream: calls whose GPU work can overlap
 *     race on it -- keep them on one stream or order them with events.  It keeps the previous call's contents;
 *     none of them is read before it is rewritten, so nothing needs clearing between calls. */
#define GGNN_STATIC_MAX_ELEMS  16777216   /* adjacency elements per call */
#define GGNN_STATIC_MAX_ROWS    1048576   /* (graph, channel, target) rows per call */
int ggnn_compact_adjacency_static(ggnn_ctx* ctx, const void* adj, int adj_dtype,
                                  int b, int two_e, int v,
                                  int32_t* rowptr, int32_t* col, int64_t col_capacity,
                                  int64_t* nnz_out_dev, void* stream);

/* A2..A6: propagation ----------------------------------------------------------
 * h0, h_out: float32 [b, v, h] (h_out may alias h0).  T = number of timesteps.
 * Every row is updated, padded or not (the reference never reads node_mask inside
 * the path); `mask_or_null` is accepted for signature parity.
 * Synchronises `stream` once while planning (reads the edge count). */
int ggnn_propagate_dense(ggnn_ctx* ctx, int slot, const float* h0,
                         const void* adj, int adj_dtype, const float* mask_or_null,
                         int b, int v, int T, int precision, unsigned flags,
                         float* h_out, void* stream);

/* Batch pipelining for ggnn_propagate_dense -- the counterpart of the reference's ThreadedIterator, which
 * builds the next batch while sess.run works on the current one (utils.py:17-37, chem_tensorflow.py:541).
 * Announces the adjacency of the NEXT ggnn_propagate_dense call on this ctx.  Its dense -> bitmask pass (the
 * HBM-bound half of the path) starts at once on a ctx-owned side stream as a small background kernel that
 * shares the SMs with the msg / GRU kernels still running on the caller's stream; the propagate call that
 * follows with the SAME (adj, adj_dtype, b, v) finishes whatever is left of the pass and goes on as usual --
 * same results, bit for bit.  The plan's counting kernels follow the pass on the side stream, assuming the consuming
 * call uses the precision / slot / flags of the ctx's last dense call (they decide whether word-order channels are
 * counted); a consuming call that differs simply counts again.  Any other ggnn_propagate_dense call, a second prefetch, or ggnn_set_weights drops
 * the announcement (after waiting for the background pass).
 *   - `adj` must already hold its final contents when this is called, or `ready_event` (a cudaEvent_t as void*,
 *     else NULL) must mark their completion; it must stay valid until the consuming call's work has finished;
 *   - a hint: batches the fast route does not serve (more than one chunk, v > 255, unaligned adjacency) are
 *     accepted and simply not prefetched.  No synchronisation unless an announcement is dropped. */
int ggnn_dense_prefetch(ggnn_ctx* ctx, const void* adj, int adj_dtype, int b, int v, void* ready_event_or_null);

/* Host-resident dense adjacency (the reference feeds it from host memory every step: feed_dict,
 * chem_tensorflow.py:262-268, built by chem_tensorflow_dense.py:566-602).  The path uses ONE BIT per element (A != 0),
 * and over PCIe the FP32 tensor caps the call at ~9 k C3 graphs/s whatever the GPU does, so for HOST inputs the bit is
 * formed where the data is and 1/32 of the bytes travel:
 *   ggnn_host_pack_adjacency   host function, no ctx, no CUDA: bits_host[w] bit i = (element 32 w + i != 0) over the
 *                              flat tensor (n_elems = b*2E*v*v; the last word zero padded) -- exactly the
 *                              intermediate the GPU's own first pass produces for device-resident inputs
 *                              (tests/test_gpu_hostpack.py: bit for bit).  threads <= 0: all host cores.
 *   ggnn_propagate_dense_bits  ggnn_propagate_dense with that bitmask (DEVICE memory, 16-byte aligned) in place of
 *                              the dense tensor: row counts, plan, messages and GRU steps on the GPU as usual; same
 *                              results bit for bit.  GGNN_E_UNSUPPORTED where the bitmask route does not apply
 *                              (v > 255, 2E > 256): use ggnn_propagate_dense.
 * Device-resident adjacency keeps the GPU bitmask pass (ggnn_propagate_dense, ggnn_dense_prefetch). */
int ggnn_host_pack_adjacency(const void* adj_host, int adj_dtype, int64_t n_elems, uint32_t* bits_host, int threads);
int ggnn_propagate_dense_bits(ggnn_ctx* ctx, int slot, const float* h0, const uint32_t* bits,
                              int b, int v, int T, int precision, unsigned flags,
                              float* h_out, void* stream);

int ggnn_propagate_csr(ggnn_ctx* ctx, int slot, const float* h0,
                       const int32_t* rowptr, const int32_t* col,
                       int b, int v, int T, int precision, unsigned flags,
                       float* h_out, void* stream);

/* Edge-list form (A9): h0/h_out float32 [N, h]; edges int32 [nnz, 3] rows
 * (src, channel, dst), channel 0-based in [0, two_e), node ids global in [0, N).
 * acts[dst] = sum over edges (h[src] @ W[channel] + B[channel]); messages into one
 * node are summed in ascending (channel, src) order (duplicate triples, whose messages are identical, in any order). */
int ggnn_propagate_edges(ggnn_ctx* ctx, int slot, const float* h0,
                         const int32_t* edges, int64_t nnz, int64_t N,
                         int T, int precision, unsigned flags,
                         float* h_out, void* stream);

/* Training (BASELINE config 5): backward of the LAST propagate call that ran with
 * GGNN_FLAG_SAVE_TAPE on this ctx -- what tf.gradients builds over the unrolled path for
 * make_train_step (chem_tensorflow.py:483-506), dropout keep-probabilities 1.0.
 *   d_hT  [N, h]  gradient w.r.t. the final node representations (in)
 *   d_h0  [N, h]  gradient w.r.t. the initial node representations (out)
 *   grads         flat FP32 buffer of ggnn_grad_floats() elements, ACCUMULATED into (zero it first):
 *                 [dW 2E*h*h | dB 2E*h | dWg 2h*2h | dbg 2h | dWc 2h*h | dbc h]  (layouts of set_weights)
 * This buffer is what the data-parallel trainer all-reduces with NCCL.  Weight gradients are
 * summed with float atomics (order not fixed).  No synchronisation. */
int64_t ggnn_grad_floats(const ggnn_ctx* ctx);
/* N4: the optimiser step of make_train_step (chem_tensorflow.py:483-506) on the ctx-owned FP32
 * master weights of `slot` (and the shared GRU cell): per VARIABLE tf.clip_by_norm(grad, clip_norm)
 * (the six segments of the flat buffer are the six variables), then TF1 AdamOptimizer
 *   lr_t = lr*sqrt(1-beta2^t)/(1-beta1^t); m = b1 m + (1-b1) g; v = b2 v + (1-b2) g^2; p -= lr_t m/(sqrt(v)+eps)
 * with ctx-owned moments (zero-initialised, t counted per ctx).  clip_norm <= 0 disables clipping.  The
 * TF32 / FP16 packed copies are refreshed.  `grads` is the (all-reduced) buffer of
 * ggnn_propagate_backward.  ggnn_get_weights copies the current master weights out (NULL = skip). */
int ggnn_optimizer_step(ggnn_ctx* ctx, int slot, const float* grads, float lr, float beta1, float beta2,
                        float eps, float clip_norm, void* stream);
int ggnn_get_weights(ggnn_ctx* ctx, int slot, float* W, float* B, float* Wg, float* bg, float* Wc,
                     float* bc, void* stream);

/* A8: training-time dropout of the path with CALLER-SUPPLIED 0/1 masks (TF's RNG stream cannot be
 * reproduced, so the masks are inputs).  Applies to the NEXT propagate call with GGNN_FLAG_SAVE_TAPE
 * (and to its backward), then is cleared.
 *   w_mask     [T, two_e, h, h] or NULL: tf.nn.dropout on edge_weights, a fresh mask per unrolled
 *              step -- the step uses W * w_mask[t] / keep (chem_tensorflow_dense.py:397-403)
 *   state_mask [T, N, h] or NULL: DropoutWrapper(state_keep_prob) -- the new state of step t becomes
 *              h' * state_mask[t] / keep (:239-240, 333)
 * Both pointers must stay valid until the backward has run.  keep in (0, 1]. */
int ggnn_set_dropout(ggnn_ctx* ctx, const float* w_mask, const float* state_mask, float keep, int T);

/* Same, with the masks drawn on the device: Philox4x32-10, key = seed, counter = (element / 4, stream id),
 * stream id 2t = edge-weight mask of unrolled step t, 2t+1 = state mask of step t; mask = 1 where
 * u32 * 2^-32 < keep.  (TF's own generator cannot be reproduced bit for bit; parity of the dropped path is
 * established with the SAME masks fed to the oracle: ggnn_dropout_mask materialises exactly the mask of
 * (seed, stream id) the kernels use, n elements, into a device buffer.)  Applies to the next SAVE_TAPE
 * propagate call and its backward only. */
int ggnn_set_dropout_seed(ggnn_ctx* ctx, float keep, unsigned long long seed, int drop_weights, int drop_state);
int ggnn_dropout_mask(ggnn_ctx* ctx, float* out, int64_t n, unsigned long long seed,
                      unsigned long long stream_id, float keep, void* stream);
int ggnn_propagate_backward(ggnn_ctx* ctx, const float* d_hT, float* d_h0, float* grads, void* stream);

/* "Next" rows either side of the path, evaluation mode --------------------
 * N1: get_initial_node_representation (chem_tensorflow_dense.py:264-310).  idx int32 [N, 4] =
 *     (loc, pos, word, head_loc) -- columns 0-3 of the reference's word_inputs (:825-842);
 *     h0[n] = concat(loc_tab[loc] | pos_tab[pos] | word_tab[word] | loc_tab[head_loc]) zero-padded
 *     to h.  An index outside its table is GGNN_E_ARG.  Synchronises `stream` once.
 * N2: gated_regression, btb branch (:439-505) + masks (:633-661) + decode (:106-131):
 *     logits = [hT | h0] @ Wcat + bcat with Wcat [2h, ncat] holding the head-position layer in
 *     columns [0, n_head) and the label layer in [lab_off, lab_off + n_lab) (ncat, lab_off % 4 == 0,
 *     padding columns zero); softmax per group; n_active [N] = number of active nodes of the row's
 *     graph for an active row, 0 for a padded row; pred [N, 2] = (first arg-max head position among
 *     s < n_active, first arg-max label), -1 for padded rows.  p_head [N, n_head] / p_lab [N, n_lab]
 *     (optional) receive the masked probabilities.  No synchronisation. */
int ggnn_embed_nodes(ggnn_ctx* ctx, const int32_t* idx, int64_t N,
                     const float* loc_tab, int loc_rows, int loc_dim,
                     const float* pos_tab, int pos_rows, int pos_dim,
                     const float* word_tab, int word_rows, int word_dim,
                     int h, float* h0, void* stream);
int ggnn_heads_decode(ggnn_ctx* ctx, const float* hT, const float* h0, int64_t N, int h,
                      const float* Wcat, const float* bcat, int ncat, int n_head, int n_lab, int lab_off,
                      const int32_t* n_active, int32_t* pred, float* p_head_or_null, float* p_lab_or_null,
                      void* stream);

/* UPPER-BOUND ESTIMATE (closed form, independent of the arithmetic mode) of the ctx-owned scratch a
 * propagate_edges / propagate_csr call of this size will hold; the ctx sizes its arenas exactly at call time */
size_t ggnn_workspace_bytes(const ggnn_ctx* ctx, int64_t N, int64_t nnz, int two_e, int h);

/* instrumentation: number of kernel launches this ctx has issued so far */
int64_t ggnn_launch_count(const ggnn_ctx* ctx);

/* instrumentation: per-stage device time.  With profiling on, the ctx records a CUDA event
 * pair on the caller's stream around each stage's launches.  ggnn_profile_read synchronises
 * the device, adds the elapsed times into ms[GGNN_N_STAGES] / launches[GGNN_N_STAGES]
 * (either may be NULL) and resets the accumulators. */
#define GGNN_STAGE_COMPACT_MASK  0   /* dense adjacency -> bitmask (the HBM-bound pass)   */
#define GGNN_STAGE_COMPACT_REST  1   /* row counts, scan, column fill                      */
#define GGNN_STAGE_PLAN          2   /* channel-major message rows + target CSR            */
#define GGNN_STAGE_MSG           3   /* message transform  h[src] @ W_e + b_e              */
#define GGNN_STAGE_AGG           4   /* segmented gather-sum                               */
#define GGNN_STAGE_GRU_GATES     5   /* [x,h] @ Wg, sigmoid, r*h                           */
#define GGNN_STAGE_GRU_CAND      6   /* [x,rh] @ Wc, tanh, state update                    */
#define GGNN_STAGE_OTHER         7
#define GGNN_N_STAGES            8
int ggnn_profile_enable(ggnn_ctx* ctx, int on);
int ggnn_profile_read(ggnn_ctx* ctx, double* ms, int64_t* launches);

#ifdef __cplusplus
}
#endif
#endif /* GGNN_B200_H */
