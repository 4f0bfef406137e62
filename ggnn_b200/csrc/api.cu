// api.cu -- the C ABI of libggnn_b200.so (see include/ggnn_b200.h for the contract).
#include "plan.cuh"
#include "tf32.cuh"
#include "train.cuh"
#include "heads.cuh"
#include "f16.cuh"
#include "plan_bits.cuh"

#include <new>
#include <stdlib.h>

namespace {

struct arena {
  char* p = nullptr;
  size_t bytes = 0;
};

// Two grow-only arenas per ctx: [0] compaction (bitmask, rowptr, scan), [1] col + plan + step
// buffers.  They are separate because [1] is sized only after the edge count is known and
// must not move [0].
// [2] training tape (per-step activations), [3] backward scratch, [4] FP16 node states,
// [5] effective edge weights of the dropout path, [6] Adam moments + norm slots,
// [7], [8] bitmask + queue control of a prefetched adjacency (ggnn_dense_prefetch), used alternately.
struct prefetch_rec {
  bool active = false;
  const void* adj = nullptr;
  int dtype = 0, b = 0, v = 0, two_e = 0, buf = 0;
  bool counted = false;      // the row counts / scans of the plan were run on the side stream too, with ...
  bool word_fold = false;    // ... this word-order setting (the consuming call re-counts from the bitmask if its own differs)
};

struct ctx_ext {
  arena ar[9];
  // ggnn_dense_prefetch: side stream the background bitmask pass runs on, its completion events, the record
  cudaStream_t side = nullptr;
  cudaEvent_t pf_done[2] = {nullptr, nullptr};
  prefetch_rec pf;
  int pf_last_buf = 1;       // buffer of the last CONSUMED prefetch: the next one goes to the other
  // the counting half of the plan (plan_bits.cu count + scans) depends on whether the word-order channels are folded
  // into the GRU kernel, i.e. on the precision / slot / flags of the CONSUMING call, which a prefetch does not know: it
  // assumes the next call looks like the last one on the bitmask route
  bool last_known = false, last_word_fold = false;
  bool pf_count = true;      // GGNN_PF_COUNT=0: prefetch the bitmask only (A-B switch)
  int64_t adam_n = 0;       // floats the moment buffers were sized (and zeroed) for
  int64_t adam_t = 0;       // steps taken
  ggnn_tape tape;
  // A8: masks for the next SAVE_TAPE forward (ggnn_set_dropout)
  const float* drop_w = nullptr;
  const float* drop_s = nullptr;
  float drop_keep = 1.f;
  int drop_T = -1;
  // ... or generated on the device from a seed (ggnn_set_dropout_seed); stream ids: 2t = weights, 2t+1 = state of step t
  bool drop_gen = false, drop_gen_w = false, drop_gen_s = false;
  unsigned long long drop_seed = 0;
};

ctx_ext* ext_of(ggnn_ctx* c) { return reinterpret_cast<ctx_ext*>(c->ws); }

// the kernels move 16-byte vectors (float4, cp.async, TMA): every state / gradient pointer of the ABI must be
// 16-byte aligned (any cudaMalloc or framework allocation is); the adjacency may be element-aligned only (the
// bitmask pass falls back to scalar loads)
inline bool misaligned16(const void* p) { return (reinterpret_cast<uintptr_t>(p) & 15u) != 0; }

int arena_reserve(ggnn_ctx* ctx, int which, size_t bytes, cudaStream_t st) {
  arena& a = ext_of(ctx)->ar[which];
  if (a.bytes >= bytes) return GGNN_OK;
  if (a.p) {
    GGNN_CUDA(ctx, cudaStreamSynchronize(st));
    GGNN_CUDA(ctx, cudaFree(a.p));
    a.p = nullptr;
    a.bytes = 0;
  }
  size_t want = align_up(bytes + bytes / 8, (size_t)1 << 20);
  cudaError_t e = cudaMalloc(&a.p, want);
  if (e != cudaSuccess) {
    want = align_up(bytes, (size_t)1 << 20);
    GGNN_CUDA(ctx, cudaMalloc(&a.p, want));
  }
  a.bytes = want;
  return GGNN_OK;
}

int free_dev(ggnn_ctx* ctx, float*& p) {
  if (p) { cudaFree(p); p = nullptr; }
  return GGNN_OK;
}

int dup_dev(ggnn_ctx* ctx, float*& dst, const float* src, size_t n, cudaStream_t st) {
  free_dev(ctx, dst);
  GGNN_CUDA(ctx, cudaMalloc(&dst, n * sizeof(float)));
  GGNN_CUDA(ctx, cudaMemcpyAsync(dst, src, n * sizeof(float), cudaMemcpyDeviceToDevice, st));
  return GGNN_OK;
}

struct device_guard {
  int prev = -1;
  explicit device_guard(int dev) { cudaGetDevice(&prev); if (prev != dev) cudaSetDevice(dev); else prev = -1; }
  ~device_guard() { if (prev >= 0) cudaSetDevice(prev); }
};

// graphs per chunk of the dense path: canonical rows stay well inside int32 and scratch stays bounded
int64_t dense_chunk(int two_e, int v) {
  int64_t chunk = ((int64_t)1 << 26) / ((int64_t)two_e * v);
  if (const char* e = getenv("GGNN_CHUNK")) { const int64_t c = atoll(e); if (c > 0 && c < chunk) chunk = c; }   // A-B switch
  if (chunk < 1) chunk = 1;
  if (chunk >= 16) chunk &= ~(int64_t)15;     // chunk starts stay 16-byte aligned for any v, 2E and adjacency dtype
  return chunk;
}

// drop an unconsumed prefetch: wait for its background pass so that its buffer can be reused
void prefetch_cancel(ggnn_ctx* ctx) {
  ctx_ext* ex = ext_of(ctx);
  if (ex->pf.active) {
    cudaStreamSynchronize(ex->side);
    ex->pf.active = false;
  }
}

}  // namespace

extern "C" int ggnn_version(void) { return 100; }

extern "C" int ggnn_create(ggnn_ctx** out, int device) {
  if (!out) return GGNN_E_ARG;
  *out = nullptr;
  int ndev = 0;
  if (cudaGetDeviceCount(&ndev) != cudaSuccess || device < 0 || device >= ndev) return GGNN_E_CUDA;
  ggnn_ctx* c = new (std::nothrow) ggnn_ctx();
  if (!c) return GGNN_E_CUDA;
  c->device = device;
  device_guard g(device);
  cudaDeviceProp prop;
  if (cudaGetDeviceProperties(&prop, device) != cudaSuccess) { delete c; return GGNN_E_CUDA; }
  if (prop.major != 10) {   // sm_100a only: no fallback path exists
    delete c;
    return GGNN_E_UNSUPPORTED;
  }
  c->sm_count = prop.multiProcessorCount;
  c->no_fused = getenv("GGNN_NO_FUSED") != nullptr;
  c->no_bits_plan = getenv("GGNN_NO_BITS_PLAN") != nullptr;
  c->ws = reinterpret_cast<char*>(new (std::nothrow) ctx_ext());
  if (cudaMallocHost(&c->h_mail, 8 * sizeof(int64_t)) != cudaSuccess ||
      cudaMalloc(&c->d_mail, 8 * sizeof(int64_t)) != cudaSuccess) {
    ggnn_destroy(c);
    return GGNN_E_CUDA;
  }
  *out = c;
  return GGNN_OK;
}

extern "C" void ggnn_destroy(ggnn_ctx* c) {
  if (!c) return;
  device_guard g(c->device);
  cudaDeviceSynchronize();
  for (int s = 0; s < GGNN_MAX_SLOTS; ++s) {
    free_dev(c, c->edge[s].W);
    free_dev(c, c->edge[s].B);
    free_dev(c, c->edge[s].Wt_tf32);
    free_dev(c, c->edge[s].Wt_tf32_lo);
    free_dev(c, c->edge[s].Wt_tf32_hl);
    if (c->edge[s].Wt_f16) { cudaFree(c->edge[s].Wt_f16); c->edge[s].Wt_f16 = nullptr; }
    free_dev(c, c->edge[s].B_pad16);
    if (c->edge[s].fold_f16) { cudaFree(c->edge[s].fold_f16); c->edge[s].fold_f16 = nullptr; }
    free_dev(c, c->edge[s].fold_bias);
  }
  free_dev(c, c->Wg); free_dev(c, c->bg); free_dev(c, c->Wc); free_dev(c, c->bc);
  if (c->Wgt_f16) cudaFree(c->Wgt_f16);
  if (c->Wct_f16) cudaFree(c->Wct_f16);
  free_dev(c, c->Wgt_tf32); free_dev(c, c->Wct_tf32); free_dev(c, c->Wgt_tf32_lo); free_dev(c, c->Wct_tf32_lo); free_dev(c, c->Wgt_tf32_hl); free_dev(c, c->Wct_tf32_hl); free_dev(c, c->bg_pad); free_dev(c, c->bc_pad);
  if (c->btab16) cudaFree(c->btab16);
  free_dev(c, c->bg_w16); free_dev(c, c->bc_w16);
  if (c->ws) {
    ctx_ext* e = ext_of(c);
    for (auto& a : e->ar) if (a.p) cudaFree(a.p);
    if (e->side) cudaStreamDestroy(e->side);
    for (auto& ev : e->pf_done) if (ev) cudaEventDestroy(ev);
    delete e;
  }
  for (auto& r : c->prof_recs) { cudaEventDestroy(r.a); cudaEventDestroy(r.b); }
  for (auto& e : c->prof_pool) cudaEventDestroy(e);
  if (c->h_mail) cudaFreeHost(c->h_mail);
  if (c->d_mail) cudaFree(c->d_mail);
  delete c;
}

extern "C" const char* ggnn_last_error(const ggnn_ctx* c) { return c ? c->err : "null ctx"; }
extern "C" int64_t ggnn_launch_count(const ggnn_ctx* c) { return c ? c->launches : 0; }

// development aid (not in the public header): device buffer the persistent kernels write
// globaltimer phase traces of CTA 0 into; pass NULL to switch tracing off
extern "C" int ggnn_debug_trace(ggnn_ctx* c, unsigned long long* dev_buf) {
  if (!c) return GGNN_E_ARG;
  c->dbg_trace = dev_buf;
  { const char* e = getenv("GGNN_PAIR_TRACE_WIDE"); c->dbg_trace_wide = (e && e[0] == '1') ? 1 : 0; }
  return GGNN_OK;
}

extern "C" int ggnn_profile_enable(ggnn_ctx* c, int on) {
  if (!c) return GGNN_E_ARG;
  c->prof_on = on != 0;
  return GGNN_OK;
}

extern "C" int ggnn_profile_read(ggnn_ctx* c, double* ms, int64_t* launches) {
  if (!c) return GGNN_E_ARG;
  device_guard g(c->device);
  GGNN_CUDA(c, cudaDeviceSynchronize());
  for (auto& r : c->prof_recs) {
    float t = 0.f;
    if (cudaEventElapsedTime(&t, r.a, r.b) == cudaSuccess) c->prof_ms[r.stage] += t;
    c->prof_pool.push_back(r.a);
    c->prof_pool.push_back(r.b);
  }
  c->prof_recs.clear();
  for (int i = 0; i < GGNN_N_STAGES; ++i) {
    if (ms) ms[i] = c->prof_ms[i];
    if (launches) launches[i] = c->prof_launches[i];
    c->prof_ms[i] = 0;
    c->prof_launches[i] = 0;
  }
  return GGNN_OK;
}

extern "C" int ggnn_set_weights(ggnn_ctx* ctx, int slot, int two_e, int h, const float* W,
                                const float* B, const float* Wg, const float* bg,
                                const float* Wc, const float* bc, void* stream) {
  if (!ctx) return GGNN_E_ARG;
  if (slot < 0 || slot >= GGNN_MAX_SLOTS || !W) GGNN_FAIL(ctx, GGNN_E_ARG, "set_weights: bad slot or null W");
  if (two_e <= 0 || h <= 0) GGNN_FAIL(ctx, GGNN_E_ARG, "set_weights: two_e and h must be positive");
  if (h % 4 != 0) GGNN_FAIL(ctx, GGNN_E_UNSUPPORTED, "hidden size %d is not a multiple of 4", h);
  const bool has_gru = Wg && bg && Wc && bc;
  if (slot == GGNN_SLOT_MAIN && !has_gru) GGNN_FAIL(ctx, GGNN_E_ARG, "set_weights: slot 0 needs the GRU cell");
  if (slot != GGNN_SLOT_MAIN && (!ctx->gru_set || two_e != ctx->two_e || h != ctx->h))
    GGNN_FAIL(ctx, GGNN_E_SHAPE, "set_weights: slot %d must match slot 0 (two_e=%d, h=%d)", slot, ctx->two_e, ctx->h);
  device_guard g(ctx->device);
  cudaStream_t st = (cudaStream_t)stream;
  prefetch_cancel(ctx);
  ext_of(ctx)->tape.valid = false;     // a tape belongs to the weights it was recorded with
  if (slot == GGNN_SLOT_MAIN) {
    if (ctx->two_e != two_e || ctx->h != h) ctx->edge[GGNN_SLOT_FIXED].set = false;
    ctx->two_e = two_e;
    ctx->h = h;
    ctx->hp = (int)align_up((size_t)h, 32);
    const int hw = h > 128 ? (int)align_up((size_t)h, 64) : 0;
    if (hw != ctx->hw16 || (h <= 128) != (ctx->hp_f16_narrow)) {   // the FP16 packs change layout: drop them
      for (int s2 = 0; s2 < GGNN_MAX_SLOTS; ++s2) {
        if (ctx->edge[s2].Wt_f16) { cudaFree(ctx->edge[s2].Wt_f16); ctx->edge[s2].Wt_f16 = nullptr; }
        free_dev(ctx, ctx->edge[s2].B_pad16);
      }
      if (ctx->Wgt_f16) { cudaFree(ctx->Wgt_f16); ctx->Wgt_f16 = nullptr; }
      ctx->pair_ok = false;
      if (ctx->Wct_f16) { cudaFree(ctx->Wct_f16); ctx->Wct_f16 = nullptr; }
      free_dev(ctx, ctx->bg_w16); free_dev(ctx, ctx->bc_w16);
    }
    ctx->hw16 = hw;
    ctx->hp_f16_narrow = h <= 128;
  }
  ggnn_edge_weights& ew = ctx->edge[slot];
  GGNN_TRY(dup_dev(ctx, ew.W, W, (size_t)two_e * h * h, st));
  if (B) GGNN_TRY(dup_dev(ctx, ew.B, B, (size_t)two_e * h, st)); else free_dev(ctx, ew.B);
  if (has_gru) {
    GGNN_TRY(dup_dev(ctx, ctx->Wg, Wg, (size_t)4 * h * h, st));
    GGNN_TRY(dup_dev(ctx, ctx->bg, bg, (size_t)2 * h, st));
    GGNN_TRY(dup_dev(ctx, ctx->Wc, Wc, (size_t)2 * h * h, st));
    GGNN_TRY(dup_dev(ctx, ctx->bc, bc, (size_t)h, st));
    ctx->gru_set = true;
  }
  GGNN_TRY(ggnn_tf32_pack_weights(ctx, slot, has_gru, st));
  GGNN_TRY(ggnn_f16_pack_weights(ctx, slot, has_gru, st));
  GGNN_TRY(ggnn_w16_pack_weights(ctx, slot, has_gru, st));
  ew.set = true;
  return GGNN_OK;
}

// ---------------------------------------------------------------------------------------
// A1
extern "C" int ggnn_compact_adjacency(ggnn_ctx* ctx, const void* adj, int adj_dtype, int b,
                                      int two_e, int v, int32_t* rowptr, int32_t* col,
                                      int64_t col_capacity, int64_t* nnz_out_dev, void* stream) {
  if (!ctx) return GGNN_E_ARG;
  if (b < 0 || two_e <= 0 || v <= 0) GGNN_FAIL(ctx, GGNN_E_ARG, "compact: bad sizes b=%d two_e=%d v=%d", b, two_e, v);
  if ((b > 0 && !adj) || !rowptr || !nnz_out_dev || (col_capacity > 0 && !col) || col_capacity < 0)
    GGNN_FAIL(ctx, GGNN_E_ARG, "compact: null pointer");
  if (adj_dtype != GGNN_ADJ_F32 && adj_dtype != GGNN_ADJ_U8) GGNN_FAIL(ctx, GGNN_E_ARG, "compact: bad adj_dtype %d", adj_dtype);
  const int64_t rows = (int64_t)b * two_e * v;
  if (rows >= 0x7fffffffLL) GGNN_FAIL(ctx, GGNN_E_UNSUPPORTED, "compact: b*two_e*v = %lld rows needs int64 rowptr; split the batch", (long long)rows);
  device_guard g(ctx->device);
  cudaStream_t st = (cudaStream_t)stream;
  ext_of(ctx)->tape.valid = false;     // a taped dense forward keeps its target CSR in the compaction arena
  GGNN_TRY(arena_reserve(ctx, 0, ggnn_compact_tmp_bytes(rows * v, rows), st));
  char* tmp = ext_of(ctx)->ar[0].p;
  GGNN_TRY(ggnn_compact_count(ctx, adj, adj_dtype, rows, v, rowptr, nnz_out_dev, tmp, st));
  GGNN_TRY(ggnn_compact_fill(ctx, rows, v, rowptr, col, col_capacity, tmp, st));
  return GGNN_OK;
}

// ---------------------------------------------------------------------------------------
namespace {

int check_ready(ggnn_ctx* ctx, int slot, int precision) {
  if (slot < 0 || slot >= GGNN_MAX_SLOTS) GGNN_FAIL(ctx, GGNN_E_ARG, "bad weight slot %d", slot);
  if (!ctx->gru_set || !ctx->edge[slot].set) GGNN_FAIL(ctx, GGNN_E_STATE, "weights for slot %d not set", slot);
  if (precision != GGNN_PREC_FP32 && precision != GGNN_PREC_TF32 && precision != GGNN_PREC_F16)
    GGNN_FAIL(ctx, GGNN_E_ARG, "bad precision %d", precision);
  if (precision == GGNN_PREC_F16 && !ggnn_f16_supported(ctx) && !ggnn_w16_supported(ctx))
    GGNN_FAIL(ctx, GGNN_E_UNSUPPORTED, "F16 path needs a hidden size in (96, 512], got h=%d", ctx->h);
  if (precision == GGNN_PREC_TF32 && !ggnn_tf32_supported(ctx))
    GGNN_FAIL(ctx, GGNN_E_UNSUPPORTED, "TF32 path does not support h=%d", ctx->h);
  return GGNN_OK;
}

// read one slot of the device mailbox; the whole mailbox (8 slots) comes back with it so the
// FP16 range statistics (slots 4-6) ride on the same synchronisation
int read_i64(ggnn_ctx* ctx, const int64_t* dev, int64_t* out, cudaStream_t st) {
  GGNN_CUDA(ctx, cudaMemcpyAsync(ctx->h_mail, ctx->d_mail, 8 * sizeof(int64_t), cudaMemcpyDeviceToHost, st));
  GGNN_CUDA(ctx, cudaStreamSynchronize(st));
  *out = ctx->h_mail[dev - ctx->d_mail];
  return GGNN_OK;
}

// GGNN_PREC_F16: before the planning synchronisation, convert the states to FP16 and collect
// max|h0| (slot 6) and, once per weight set, max|W| / max|B| (slots 4, 5)
int f16_prologue(ggnn_ctx* ctx, int slot, const float* h0, int64_t N, cudaStream_t st) {
  const bool wide = ggnn_w16_supported(ctx);
  GGNN_TRY(arena_reserve(ctx, 4, (size_t)N * (wide ? ctx->hw16 * 2 : 256) + 256, st));
  GGNN_CUDA(ctx, cudaMemsetAsync(ctx->d_mail + 4, 0, 3 * sizeof(int64_t), st));
  ggnn_edge_weights& ew = ctx->edge[slot];
  if (ew.absmax_w < 0.f) {
    GGNN_TRY(ggnn_colnorm_max(ctx, ew.W, ctx->two_e, ctx->h, reinterpret_cast<unsigned*>(ctx->d_mail + 4), st));
    if (ew.B) GGNN_TRY(ggnn_absmax(ctx, ew.B, (int64_t)ctx->two_e * ctx->h, reinterpret_cast<unsigned*>(ctx->d_mail + 5), st));
  }
  if (wide) return ggnn_w16_state_in(ctx, h0, ext_of(ctx)->ar[4].p, N, reinterpret_cast<unsigned*>(ctx->d_mail + 6), st);
  return ggnn_f16_state_in(ctx, h0, ext_of(ctx)->ar[4].p, N, reinterpret_cast<unsigned*>(ctx->d_mail + 6), st);
}

// after the synchronisation: refuse inputs that could leave the FP16 range (include/ggnn_b200.h)
int f16_check(ggnn_ctx* ctx, int slot) {
  auto f = [&](int i) { unsigned u = (unsigned)ctx->h_mail[i]; float v; memcpy(&v, &u, 4); return v; };
  ggnn_edge_weights& ew = ctx->edge[slot];
  if (ew.absmax_w < 0.f) { ew.absmax_w = f(4); ew.absmax_b = f(5); }
  const float hmax = f(6);
  // |h @ W_e + b_e| <= (largest column L1 norm of any W_e) * max|h| + max|B|; states stay within max(max|h0|, 1)
  const float bound = ew.absmax_w * (hmax > 1.f ? hmax : 1.f) + ew.absmax_b;
  if (!(bound < 64.f))
    GGNN_FAIL(ctx, GGNN_E_UNSUPPORTED, "F16 mode: max column L1 norm of W (%g) * max(max|h0|,1) (max|h0|=%g) + max|B| = %g >= 64: values may leave the FP16 range; use GGNN_PREC_TF32",
              ew.absmax_w, hmax, bound);
  return GGNN_OK;
}

// step buffers after the plan, from arena 1
void take_step_bufs(ws_alloc& a, const ggnn_plan& plan, int h, ggnn_step_bufs* sb) {
  sb->msg = a.take<float>((size_t)plan.max_tiles * GGNN_TILE_M * (h < 64 ? 64 : h));   // >= 256 B per row: FP16 rows of pitch 128
  sb->acts = a.take<float>((size_t)plan.N * h);
  sb->rh = a.take<float>((size_t)plan.N * h);
  sb->u = a.take<float>((size_t)plan.N * h);
}

int run_steps(ggnn_ctx* ctx, int slot, const ggnn_plan& plan, const ggnn_step_bufs& sb,
              float* h_state, int T, int precision, unsigned flags, cudaStream_t st) {
  ggnn_tape& tape = ext_of(ctx)->tape;
  tape.valid = false;
  if (flags & GGNN_FLAG_SAVE_TAPE) {      // training forward: keep h, x, r, u, c of every step
    // GGNN_PREC_F16 (h > 128): forward on the FP16 kernels of wide16.cu, tape in FP32, backward on the TF32 tensor-core path
    const bool f16t = precision == GGNN_PREC_F16;
    if (f16t && !ggnn_w16_supported(ctx))
      GGNN_FAIL(ctx, GGNN_E_UNSUPPORTED, "SAVE_TAPE with GGNN_PREC_F16 needs a hidden size in (128, 512], got h=%d; use GGNN_PREC_TF32", ctx->h);
    const bool tf = precision == GGNN_PREC_TF32;
    void* state16 = ext_of(ctx)->ar[4].p;
    if (plan.dense_rows) GGNN_FAIL(ctx, GGNN_E_UNSUPPORTED, "SAVE_TAPE: edge list is in dense-row mode (nnz >= N*2E)");
    const size_t nh = (size_t)plan.N * ctx->h;
    GGNN_TRY(arena_reserve(ctx, 2, (size_t)T * TAPE_ARRAYS * nh * 4 + 256, st));
    tape.plan = plan; tape.T = T; tape.slot = slot; tape.h = ctx->h; tape.tf32 = tf || f16t;
    tape.acts = reinterpret_cast<float*>(ext_of(ctx)->ar[2].p);
    ctx_ext* ex = ext_of(ctx);
    tape.w_mask = tape.state_mask = nullptr; tape.keep = 1.f; tape.w_eff = nullptr;
    if (ex->drop_T >= 0) {                 // masks registered by ggnn_set_dropout apply to this call only
      if (ex->drop_T != T) GGNN_FAIL(ctx, GGNN_E_SHAPE, "dropout masks were set for T=%d, propagate has T=%d", ex->drop_T, T);
      tape.w_mask = ex->drop_w; tape.state_mask = ex->drop_s; tape.keep = ex->drop_keep;
      ex->drop_T = -1;
    }
    const size_t wn = (size_t)ctx->two_e * ctx->h * ctx->h;
    if (ex->drop_gen) {                    // device-generated masks: materialise [T, 2E, h, h] and [T, N, h] once
      ex->drop_gen = false;
      const size_t wm = ex->drop_gen_w ? (size_t)T * wn : 0, sm = ex->drop_gen_s ? (size_t)T * nh : 0;
      GGNN_TRY(arena_reserve(ctx, 5, (wn + wm + sm) * 4 + 1024, st));
      float* base = reinterpret_cast<float*>(ex->ar[5].p);
      tape.w_eff = base;
      tape.keep = ex->drop_keep;
      if (wm) {
        float* m = base + wn;
        for (int t = 0; t < T; ++t)
          GGNN_TRY(ggnn_dropout_mask_fill(ctx, m + (size_t)t * wn, (int64_t)wn, ex->drop_seed, 2ull * t, tape.keep, st));
        tape.w_mask = m;
      }
      if (sm) {
        float* m = base + wn + wm;
        for (int t = 0; t < T; ++t)
          GGNN_TRY(ggnn_dropout_mask_fill(ctx, m + (size_t)t * nh, (int64_t)nh, ex->drop_seed, 2ull * t + 1, tape.keep, st));
        tape.state_mask = m;
      }
    } else if (tape.w_mask) {
      GGNN_TRY(arena_reserve(ctx, 5, wn * 4 + 256, st));
      tape.w_eff = reinterpret_cast<float*>(ex->ar[5].p);
    }
    if (tf && T > 0) GGNN_TRY(ggnn_tf32_round_copy(ctx, h_state, h_state, (int64_t)nh, st));
    for (int t = 0; t < T; ++t) {
      if (f16t) GGNN_TRY(ggnn_w16_state_out(ctx, state16, tape.at(t, TAPE_H), plan.N, st));
      else GGNN_CUDA(ctx, cudaMemcpyAsync(tape.at(t, TAPE_H), h_state, nh * 4, cudaMemcpyDeviceToDevice, st));
      ggnn_step_bufs tb = sb;
      tb.acts = tape.at(t, TAPE_X); tb.u = tape.at(t, TAPE_U);
      tb.tape_r = tape.at(t, TAPE_R); tb.tape_c = tape.at(t, TAPE_C);
      if (tape.w_mask && f16t) {           // tf.nn.dropout(W): W * m_t / keep, a fresh mask per unrolled step --
        // formed inside the pack of this step's FP16 operand (the FP32 product is never written)
        GGNN_TRY(ggnn_w16_repack_edge(ctx, slot, nullptr, tape.w_mask + (size_t)t * wn, 1.0f / tape.keep, st));
      } else if (tape.w_mask) {
        GGNN_TRY(ggnn_mask_scale(ctx, ctx->edge[slot].W, tape.w_mask + (size_t)t * wn, 1.0f / tape.keep, tape.w_eff,
                                 (int64_t)wn, false, st));
        tb.W_override = tape.w_eff;
        if (tf) GGNN_TRY(ggnn_tf32_repack_edge(ctx, slot, tape.w_eff, st));   // this step's packed operand
      }
      if (f16t) {
        const ggnn_w16_tape wt{tape.at(t, TAPE_X), tape.at(t, TAPE_R), tape.at(t, TAPE_U), tape.at(t, TAPE_C)};
        GGNN_TRY(ggnn_step_w16(ctx, slot, plan, sb.msg, sb.acts, sb.rh, sb.u, state16, h_state, t == T - 1, st, &wt));
        if (tape.state_mask) {
          if (t < T - 1) GGNN_TRY(ggnn_w16_state_mask(ctx, state16, tape.state_mask + (size_t)t * nh, 1.0f / tape.keep, plan.N, st));
          else GGNN_TRY(ggnn_mask_scale(ctx, h_state, tape.state_mask + (size_t)t * nh, 1.0f / tape.keep, h_state, (int64_t)nh, false, st));
        }
        continue;
      }
      if (tf) GGNN_TRY(ggnn_step_tf32(ctx, slot, plan, tb, h_state, t == T - 1, st));
      else GGNN_TRY(ggnn_step_fp32(ctx, slot, plan, tb, h_state, st));
      if (tape.state_mask) {               // DropoutWrapper(state_keep_prob): the new state is dropped
        GGNN_TRY(ggnn_mask_scale(ctx, h_state, tape.state_mask + (size_t)t * nh, 1.0f / tape.keep, h_state,
                                 (int64_t)nh, false, st));
        if (tf && t < T - 1) GGNN_TRY(ggnn_tf32_round_copy(ctx, h_state, h_state, (int64_t)nh, st));
      }
    }
    if (tf && tape.w_mask) GGNN_TRY(ggnn_tf32_repack_edge(ctx, slot, nullptr, st));   // back to the undropped weights
    if (f16t && tape.w_mask) GGNN_TRY(ggnn_w16_repack_edge(ctx, slot, nullptr, nullptr, 1.f, st));
    tape.valid = true;
    return GGNN_OK;
  }
  if (precision == GGNN_PREC_F16) {           // states already converted by f16_prologue (arena 4)
    const bool wide = ggnn_w16_supported(ctx);
    for (int t = 0; t < T; ++t) {
      if (wide) GGNN_TRY(ggnn_step_w16(ctx, slot, plan, sb.msg, sb.acts, sb.rh, sb.u, ext_of(ctx)->ar[4].p, h_state, t == T - 1, st));
      else GGNN_TRY(ggnn_step_f16(ctx, slot, plan, sb.msg, ext_of(ctx)->ar[4].p, h_state, t == T - 1, st));
    }
    return GGNN_OK;
  }
  if (precision == GGNN_PREC_TF32 && T > 0)   // state is kept TF32-rounded between steps
    GGNN_TRY(ggnn_tf32_round_copy(ctx, h_state, h_state, plan.N * ctx->h, st));
  const bool strict_tc = precision == GGNN_PREC_FP32 && ggnn_fp32_tc_enabled(ctx);
  for (int t = 0; t < T; ++t) {
    if (precision == GGNN_PREC_TF32) GGNN_TRY(ggnn_step_tf32(ctx, slot, plan, sb, h_state, t == T - 1, st));
    else if (strict_tc) GGNN_TRY(ggnn_step_fp32_tc(ctx, slot, plan, sb, h_state, st));
    else GGNN_TRY(ggnn_step_fp32(ctx, slot, plan, sb, h_state, st));
  }
  return GGNN_OK;
}

// propagate `bc` graphs whose canonical CSR is (rowptr, col) with nnz edges; state in place.
// `extra` = bytes at the front of arena 1 already in use (col of the dense path).
int propagate_csr_chunk(ggnn_ctx* ctx, int slot, float* h_state, const int32_t* rowptr,
                        const int32_t* col, int64_t bc, int v, int64_t nnz, int T, int precision,
                        unsigned flags, size_t extra, cudaStream_t st) {
  ggnn_plan plan;
  ggnn_step_bufs sb;
  {
    ws_alloc dry(nullptr);
    dry.off = extra;
    GGNN_TRY(ggnn_plan_from_csr(ctx, rowptr, col, bc, ctx->two_e, v, nnz, &plan, dry, true, st));
    take_step_bufs(dry, plan, ctx->h, &sb);
    if (ext_of(ctx)->ar[1].bytes < dry.off + 256) {
      if (extra) GGNN_FAIL(ctx, GGNN_E_CAPACITY, "internal: arena 1 under-reserved");
      GGNN_TRY(arena_reserve(ctx, 1, dry.off + 256, st));
    }
  }
  ws_alloc a(ext_of(ctx)->ar[1].p);
  a.off = extra;
  GGNN_TRY(ggnn_plan_from_csr(ctx, rowptr, col, bc, ctx->two_e, v, nnz, &plan, a, false, st));
  take_step_bufs(a, plan, ctx->h, &sb);
  return run_steps(ctx, slot, plan, sb, h_state, T, precision, flags, st);
}

size_t csr_chunk_bytes(ggnn_ctx* ctx, int64_t bc, int v, int64_t nnz, size_t extra) {
  ggnn_plan plan;
  ggnn_step_bufs sb;
  ws_alloc dry(nullptr);
  dry.off = extra;
  ggnn_plan_from_csr(ctx, nullptr, nullptr, bc, ctx->two_e, v, nnz, &plan, dry, true, 0);
  take_step_bufs(dry, plan, ctx->h, &sb);
  return dry.off + 256;
}

}  // namespace

// bits_in != null: the adjacency arrives as its bitmask (device memory; word w, bit i <=> element 32 w + i of the dense
// [b, 2E, v, v] tensor is non-zero) and `adj` is not read -- ggnn_propagate_dense_bits
static int propagate_dense_impl(ggnn_ctx* ctx, int slot, const float* h0, const void* adj, int adj_dtype,
                                const uint32_t* bits_in, int b, int v, int T, int precision, unsigned flags,
                                float* h_out, void* stream) {
  if (!ctx) return GGNN_E_ARG;
  if (b < 0 || v <= 0 || T < 0) GGNN_FAIL(ctx, GGNN_E_ARG, "propagate_dense: bad sizes b=%d v=%d T=%d", b, v, T);
  if (b > 0 && (!h0 || (!adj && !bits_in) || !h_out)) GGNN_FAIL(ctx, GGNN_E_ARG, "propagate_dense: null pointer");
  if (misaligned16(h0) || misaligned16(h_out)) GGNN_FAIL(ctx, GGNN_E_ARG, "propagate_dense: state pointers must be 16-byte aligned");
  if (adj_dtype != GGNN_ADJ_F32 && adj_dtype != GGNN_ADJ_U8) GGNN_FAIL(ctx, GGNN_E_ARG, "bad adj_dtype %d", adj_dtype);
  GGNN_TRY(check_ready(ctx, slot, precision));
  if (b == 0) return GGNN_OK;
  device_guard g(ctx->device);
  cudaStream_t st = (cudaStream_t)stream;
  const int two_e = ctx->two_e, h = ctx->h;
  const size_t esz = adj_dtype == GGNN_ADJ_F32 ? 4 : 1;
  // GGNN_PREC_F16 keeps the state in its own FP16 buffer (converted straight from h0) and its last step writes every
  // element of h_out: no FP32 working copy of h0 is needed (268 MB of copy per 4 096 C3 graphs)
  const bool f16_direct = precision == GGNN_PREC_F16 && T > 0 && !(flags & GGNN_FLAG_SAVE_TAPE);
  if (h_out != h0 && !f16_direct)
    GGNN_CUDA(ctx, cudaMemcpyAsync(h_out, h0, (size_t)b * v * h * sizeof(float), cudaMemcpyDeviceToDevice, st));
  // chunk so that canonical rows stay well inside int32 and scratch stays bounded
  const int64_t rows_per_graph = (int64_t)two_e * v;
  int64_t chunk = dense_chunk(two_e, v);
  if (chunk > b) chunk = b;
  if (bits_in) {
    if (!ggnn_plan_bits_supported(two_e, v) || ctx->no_bits_plan)
      GGNN_FAIL(ctx, GGNN_E_UNSUPPORTED, "propagate_dense_bits: the bitmask route does not serve 2E=%d, v=%d", two_e, v);
    if (chunk < b && (chunk * rows_per_graph * v) % 32 != 0)
      GGNN_FAIL(ctx, GGNN_E_UNSUPPORTED, "propagate_dense_bits: %d graphs need chunks that do not start on a bitmask word; split the batch", b);
  }
  // a prefetched adjacency (ggnn_dense_prefetch) is consumed by the call it was announced for, and only by that one
  ctx_ext* ex = ext_of(ctx);
  const prefetch_rec pf = ex->pf;
  const bool use_pf = !bits_in && pf.active && pf.adj == adj && pf.dtype == adj_dtype && pf.b == b && pf.v == v &&
                      pf.two_e == two_e && chunk == b && ggnn_plan_bits_supported(two_e, v) && !ctx->no_bits_plan;
  if (pf.active && !use_pf) prefetch_cancel(ctx);
  ex->pf.active = false;
  if ((flags & GGNN_FLAG_SAVE_TAPE) && chunk < b)
    GGNN_FAIL(ctx, GGNN_E_UNSUPPORTED, "SAVE_TAPE: batch of %d graphs needs more than one chunk (%lld graphs)", b, (long long)chunk);
  for (int64_t g0 = 0; g0 < b; g0 += chunk) {
    const int64_t bc = (b - g0 < chunk) ? b - g0 : chunk;
    const int64_t rows = bc * rows_per_graph;
    const char* adj_c = reinterpret_cast<const char*>(adj) + (size_t)g0 * rows_per_graph * v * esz;
    float* h_chunk = h_out + (size_t)g0 * v * h;
    if (ggnn_plan_bits_supported(two_e, v) && !ctx->no_bits_plan) {
      // fast route: bitmask -> plan, the 2E*v-rows-per-graph CSR is never materialised (plan_bits.cu)
      ggnn_plan_bits_pre pre;
      {
        ws_alloc dry(nullptr);
        ggnn_plan_bits_take_pre(dry, bc, two_e, v, &pre);
        GGNN_TRY(arena_reserve(ctx, 0, dry.off + 256, st));
      }
      ws_alloc a0(ext_of(ctx)->ar[0].p);
      ggnn_plan_bits_take_pre(a0, bc, two_e, v, &pre);
      // GGNN_PREC_F16, h <= 128, v == 128: graphs whose word-order channels are exactly t-1 -> t / t+1 -> t keep those
      // edges out of the plan; the GRU kernel forms them from the resident h tile (f16.cu)
      const bool word_fold = precision == GGNN_PREC_F16 && v == GGNN_TILE_M && !(flags & GGNN_FLAG_SAVE_TAPE) &&
                             ggnn_f16_word_fold_ready(ctx, slot);
      ex->last_known = true; ex->last_word_fold = word_fold;
      bool counted = false;
      const int64_t* pf_nnz = nullptr;
      if (use_pf) {
        // the background pass has been streaming this adjacency since the prefetch call: take the slices that are
        // left at full occupancy -- FIRST, before the state conversion below: co-resident with the previous call's
        // kernels the background pass ends ~0.1 ms after them, and the side stream's counting kernels wait for it
        // (the plan's per-row arrays live behind the bitmask in the prefetch buffer: the side stream fills them)
        char* pb = ex->ar[7 + pf.buf].p;
        ws_alloc ap(pb + 4096);
        ggnn_plan_bits_take_pre(ap, bc, two_e, v, &pre);
        pf_nnz = ap.take<int64_t>(1);
        prof_scope ps(ctx, GGNN_STAGE_COMPACT_MASK, st);
        GGNN_TRY(ggnn_bitmask_queue_launch(ctx, adj_c, adj_dtype, rows * v, pre.bits, reinterpret_cast<unsigned*>(pb), false, st));
      }
      if (precision == GGNN_PREC_F16)
        GGNN_TRY(f16_prologue(ctx, slot, f16_direct ? h0 + (size_t)g0 * v * h : h_chunk, bc * v, st));
      if (bits_in) {
        pre.bits = const_cast<uint32_t*>(bits_in) + (size_t)g0 * rows_per_graph * v / 32;     // read only from here on
      } else if (use_pf) {
        // ... then wait for the background CTAs' last slices (and the counts that follow them on the side stream)
        GGNN_CUDA(ctx, cudaStreamWaitEvent(st, ex->pf_done[pf.buf], 0));
        ex->pf_last_buf = pf.buf;
        if (pf.counted && pf.word_fold == word_fold) {      // counts and scans are there already; the edge count joins the mailbox
          GGNN_CUDA(ctx, cudaMemcpyAsync(ctx->d_mail, pf_nnz, sizeof(int64_t), cudaMemcpyDeviceToDevice, st));
          counted = true;
        }
      } else {
        GGNN_TRY(ggnn_compact_bitmask(ctx, adj_c, adj_dtype, rows * v, pre.bits, st));
      }
      if (!counted) GGNN_TRY(ggnn_plan_bits_count(ctx, pre, bc, two_e, v, word_fold, ctx->d_mail, st));
      int64_t nnz = 0;
      GGNN_TRY(read_i64(ctx, ctx->d_mail, &nnz, st));     // the one planning sync
      if (precision == GGNN_PREC_F16) GGNN_TRY(f16_check(ctx, slot));
      if (nnz >= 0x7fffffffLL) GGNN_FAIL(ctx, GGNN_E_UNSUPPORTED, "chunk has %lld edges", (long long)nnz);
      ggnn_plan plan;
      ggnn_step_bufs sb;
      {
        ws_alloc dry(nullptr);
        GGNN_TRY(ggnn_plan_bits_place(ctx, pre, bc, two_e, v, nnz, &plan, dry, true, st));
        take_step_bufs(dry, plan, h, &sb);
        GGNN_TRY(arena_reserve(ctx, 1, dry.off + 256, st));
      }
      ws_alloc a1(ext_of(ctx)->ar[1].p);
      GGNN_TRY(ggnn_plan_bits_place(ctx, pre, bc, two_e, v, nnz, &plan, a1, false, st));
      if (word_fold) plan.tile_regular = pre.regular;
      take_step_bufs(a1, plan, h, &sb);
      GGNN_TRY(run_steps(ctx, slot, plan, sb, h_chunk, T, precision, flags, st));
      continue;
    }
    // arena 0: [rowptr][compaction tmp]
    const size_t rp_bytes = align_up((size_t)(rows + 1) * 4, 256);
    GGNN_TRY(arena_reserve(ctx, 0, rp_bytes + ggnn_compact_tmp_bytes(rows * v, rows), st));
    int32_t* rowptr = reinterpret_cast<int32_t*>(ext_of(ctx)->ar[0].p);
    char* tmp = ext_of(ctx)->ar[0].p + rp_bytes;
    if (precision == GGNN_PREC_F16)
      GGNN_TRY(f16_prologue(ctx, slot, (f16_direct ? h0 : h_out) + (size_t)g0 * v * h, bc * v, st));
    GGNN_TRY(ggnn_compact_count(ctx, adj_c, adj_dtype, rows, v, rowptr, ctx->d_mail, tmp, st));
    int64_t nnz = 0;
    GGNN_TRY(read_i64(ctx, ctx->d_mail, &nnz, st));     // the one planning sync
    if (precision == GGNN_PREC_F16) GGNN_TRY(f16_check(ctx, slot));
    if (nnz >= 0x7fffffffLL) GGNN_FAIL(ctx, GGNN_E_UNSUPPORTED, "chunk has %lld edges", (long long)nnz);
    // arena 1: [col][plan][step buffers]
    const size_t col_bytes = align_up((size_t)(nnz + 1) * 4, 256);
    GGNN_TRY(arena_reserve(ctx, 1, csr_chunk_bytes(ctx, bc, v, nnz, col_bytes), st));
    int32_t* col = reinterpret_cast<int32_t*>(ext_of(ctx)->ar[1].p);
    GGNN_TRY(ggnn_compact_fill(ctx, rows, v, rowptr, col, nnz, tmp, st));
    GGNN_TRY(propagate_csr_chunk(ctx, slot, h_out + (size_t)g0 * v * h, rowptr, col, bc, v, nnz,
                                 T, precision, flags, col_bytes, st));
  }
  return GGNN_OK;
}

extern "C" int ggnn_propagate_dense(ggnn_ctx* ctx, int slot, const float* h0, const void* adj,
                                    int adj_dtype, const float* mask_or_null, int b, int v, int T,
                                    int precision, unsigned flags, float* h_out, void* stream) {
  (void)mask_or_null;
  if (ctx && b > 0 && !adj) GGNN_FAIL(ctx, GGNN_E_ARG, "propagate_dense: null pointer");
  return propagate_dense_impl(ctx, slot, h0, adj, adj_dtype, nullptr, b, v, T, precision, flags, h_out, stream);
}

extern "C" int ggnn_propagate_dense_bits(ggnn_ctx* ctx, int slot, const float* h0, const uint32_t* bits, int b, int v,
                                         int T, int precision, unsigned flags, float* h_out, void* stream) {
  if (ctx && b > 0 && !bits) GGNN_FAIL(ctx, GGNN_E_ARG, "propagate_dense_bits: null pointer");
  if (ctx && (reinterpret_cast<uintptr_t>(bits) & 15u)) GGNN_FAIL(ctx, GGNN_E_ARG, "propagate_dense_bits: the bitmask must be 16-byte aligned");
  return propagate_dense_impl(ctx, slot, h0, nullptr, GGNN_ADJ_F32, bits, b, v, T, precision, flags, h_out, stream);
}

// Announce the adjacency of the NEXT ggnn_propagate_dense call: its bitmask pass starts now, on the ctx's own side
// stream, as a background kernel that shares the SMs with whatever the caller's stream is running (header).
extern "C" int ggnn_dense_prefetch(ggnn_ctx* ctx, const void* adj, int adj_dtype, int b, int v, void* ready_event) {
  if (!ctx) return GGNN_E_ARG;
  if (b < 0 || v <= 0) GGNN_FAIL(ctx, GGNN_E_ARG, "dense_prefetch: bad sizes b=%d v=%d", b, v);
  if (b > 0 && !adj) GGNN_FAIL(ctx, GGNN_E_ARG, "dense_prefetch: null pointer");
  if (adj_dtype != GGNN_ADJ_F32 && adj_dtype != GGNN_ADJ_U8) GGNN_FAIL(ctx, GGNN_E_ARG, "bad adj_dtype %d", adj_dtype);
  if (!ctx->gru_set) GGNN_FAIL(ctx, GGNN_E_STATE, "dense_prefetch: weights not set");
  device_guard g(ctx->device);
  ctx_ext* ex = ext_of(ctx);
  const int two_e = ctx->two_e;
  const int64_t total = (int64_t)b * two_e * v * v;
  // a hint: shapes the fast route does not serve are simply not prefetched
  if (b == 0 || !ggnn_plan_bits_supported(two_e, v) || ctx->no_bits_plan || dense_chunk(two_e, v) < b ||
      !ggnn_bitmask_queue_supported(adj, total)) {
    prefetch_cancel(ctx);
    return GGNN_OK;
  }
  if (!ex->side) {
    // lowest priority: the side stream's short count CTAs must not get in front of the caller's persistent kernels
    int prio_lo = 0, prio_hi = 0;
    GGNN_CUDA(ctx, cudaDeviceGetStreamPriorityRange(&prio_lo, &prio_hi));
    GGNN_CUDA(ctx, cudaStreamCreateWithPriority(&ex->side, cudaStreamNonBlocking, prio_lo));
    { const char* e = getenv("GGNN_PF_COUNT"); ex->pf_count = !(e && e[0] == '0'); }
    for (auto& e : ex->pf_done) GGNN_CUDA(ctx, cudaEventCreateWithFlags(&e, cudaEventDisableTiming));
  }
  int buf = ex->pf_last_buf ^ 1;
  if (ex->pf.active) { buf = ex->pf.buf; prefetch_cancel(ctx); }   // replaced before it was used: same buffer
  ggnn_plan_bits_pre pre;
  {
    ws_alloc dry(nullptr);
    ggnn_plan_bits_take_pre(dry, b, two_e, v, &pre);
    dry.take<int64_t>(1);
    GGNN_TRY(arena_reserve(ctx, 7 + buf, 4096 + dry.off + 256, ex->side));
  }
  char* pb = ex->ar[7 + buf].p;
  ws_alloc ap(pb + 4096);
  ggnn_plan_bits_take_pre(ap, b, two_e, v, &pre);          // pre.bits = pb + 4096
  int64_t* pf_nnz = ap.take<int64_t>(1);
  GGNN_CUDA(ctx, cudaMemsetAsync(pb, 0, ggnn_bitmask_queue_ctl_bytes(), ex->side));
  if (ready_event) GGNN_CUDA(ctx, cudaStreamWaitEvent(ex->side, (cudaEvent_t)ready_event, 0));
  GGNN_TRY(ggnn_bitmask_queue_launch(ctx, adj, adj_dtype, total, pre.bits, reinterpret_cast<unsigned*>(pb), true, ex->side));
  // The counting half of the plan follows the bitmask on the side stream (the background CTAs finish the queue on their
  // own; the consuming call's drain launch only helps): 0.22 ms per 4 096 C3 graphs that the consuming call no longer
  // runs in front of its first timestep.  The buffer is safe to reuse: it alternates, and the call in between has
  // synchronised the caller's stream (planning sync) after the last reader of this one.
  const bool fold = ex->last_word_fold && v == GGNN_TILE_M && two_e >= 2 && two_e % 2 == 0;
  const bool count = ex->pf_count && ex->last_known;
  if (count) GGNN_TRY(ggnn_plan_bits_count(ctx, pre, b, two_e, v, fold, pf_nnz, ex->side));
  GGNN_CUDA(ctx, cudaEventRecord(ex->pf_done[buf], ex->side));
  ex->pf.counted = count; ex->pf.word_fold = fold;
  ex->pf.active = true; ex->pf.adj = adj; ex->pf.dtype = adj_dtype; ex->pf.b = b; ex->pf.v = v;
  ex->pf.two_e = two_e; ex->pf.buf = buf;
  return GGNN_OK;
}

extern "C" int ggnn_propagate_csr(ggnn_ctx* ctx, int slot, const float* h0, const int32_t* rowptr,
                                  const int32_t* col, int b, int v, int T, int precision,
                                  unsigned flags, float* h_out, void* stream) {
  if (!ctx) return GGNN_E_ARG;
  if (b < 0 || v <= 0 || T < 0) GGNN_FAIL(ctx, GGNN_E_ARG, "propagate_csr: bad sizes");
  if (b > 0 && (!h0 || !rowptr || !h_out)) GGNN_FAIL(ctx, GGNN_E_ARG, "propagate_csr: null pointer");
  if (misaligned16(h0) || misaligned16(h_out)) GGNN_FAIL(ctx, GGNN_E_ARG, "propagate_csr: state pointers must be 16-byte aligned");
  GGNN_TRY(check_ready(ctx, slot, precision));
  if (b == 0) return GGNN_OK;
  const int64_t rows = (int64_t)b * ctx->two_e * v;
  if (rows >= 0x7fffffffLL) GGNN_FAIL(ctx, GGNN_E_UNSUPPORTED, "propagate_csr: too many rows; split the batch");
  device_guard g(ctx->device);
  cudaStream_t st = (cudaStream_t)stream;
  if (h_out != h0)
    GGNN_CUDA(ctx, cudaMemcpyAsync(h_out, h0, (size_t)b * v * ctx->h * sizeof(float), cudaMemcpyDeviceToDevice, st));
  if (precision == GGNN_PREC_F16) {
    int64_t dummy = 0;
    GGNN_TRY(f16_prologue(ctx, slot, h_out, (int64_t)b * v, st));
    GGNN_TRY(read_i64(ctx, ctx->d_mail, &dummy, st));
    GGNN_TRY(f16_check(ctx, slot));
  }
  // nnz = rowptr[rows]
  GGNN_CUDA(ctx, cudaMemcpyAsync(ctx->h_mail + 1, rowptr + rows, sizeof(int32_t), cudaMemcpyDeviceToHost, st));
  GGNN_CUDA(ctx, cudaStreamSynchronize(st));
  const int64_t nnz = *reinterpret_cast<int32_t*>(ctx->h_mail + 1);
  if (nnz < 0) GGNN_FAIL(ctx, GGNN_E_ARG, "propagate_csr: rowptr[last] is negative");
  if (nnz > 0 && !col) GGNN_FAIL(ctx, GGNN_E_ARG, "propagate_csr: null col");
  return propagate_csr_chunk(ctx, slot, h_out, rowptr, col, b, v, nnz, T, precision, flags, 0, st);
}

extern "C" int ggnn_propagate_edges(ggnn_ctx* ctx, int slot, const float* h0, const int32_t* edges,
                                    int64_t nnz, int64_t N, int T, int precision, unsigned flags,
                                    float* h_out, void* stream) {
  if (!ctx) return GGNN_E_ARG;
  if (nnz < 0 || N < 0 || T < 0) GGNN_FAIL(ctx, GGNN_E_ARG, "propagate_edges: bad sizes");
  if ((N > 0 && (!h0 || !h_out)) || (nnz > 0 && !edges)) GGNN_FAIL(ctx, GGNN_E_ARG, "propagate_edges: null pointer");
  if (misaligned16(h0) || misaligned16(h_out)) GGNN_FAIL(ctx, GGNN_E_ARG, "propagate_edges: state pointers must be 16-byte aligned");
  GGNN_TRY(check_ready(ctx, slot, precision));
  if (N == 0) return GGNN_OK;
  if (N >= 0x7fffffffLL) GGNN_FAIL(ctx, GGNN_E_UNSUPPORTED, "propagate_edges: N too large for int32 node ids");
  device_guard g(ctx->device);
  cudaStream_t st = (cudaStream_t)stream;
  const bool f16_direct = precision == GGNN_PREC_F16 && T > 0 && !(flags & GGNN_FLAG_SAVE_TAPE);   // see ggnn_propagate_dense
  if (h_out != h0 && !f16_direct)
    GGNN_CUDA(ctx, cudaMemcpyAsync(h_out, h0, (size_t)N * ctx->h * sizeof(float), cudaMemcpyDeviceToDevice, st));
  if (precision == GGNN_PREC_F16) GGNN_TRY(f16_prologue(ctx, slot, f16_direct ? h0 : h_out, N, st));
  ggnn_plan plan;
  ggnn_step_bufs sb;
  {
    ws_alloc dry(nullptr);
    GGNN_TRY(ggnn_plan_from_edges(ctx, edges, nnz, N, ctx->two_e, &plan, dry, true, st));
    take_step_bufs(dry, plan, ctx->h, &sb);
    GGNN_TRY(arena_reserve(ctx, 1, dry.off + 256, st));
  }
  ws_alloc a(ext_of(ctx)->ar[1].p);
  GGNN_TRY(ggnn_plan_from_edges(ctx, edges, nnz, N, ctx->two_e, &plan, a, false, st));
  take_step_bufs(a, plan, ctx->h, &sb);
  int64_t bad = 0;
  GGNN_TRY(read_i64(ctx, ctx->d_mail + 2, &bad, st));   // planning sync: edge validation result
  if (bad) GGNN_FAIL(ctx, GGNN_E_ARG, "propagate_edges: %lld edges have node or channel ids out of range", (long long)bad);
  if (precision == GGNN_PREC_F16) GGNN_TRY(f16_check(ctx, slot));
  return run_steps(ctx, slot, plan, sb, h_out, T, precision, flags, st);
}

// ---------------------------------------------------------------------------------------
// config 5: backward of the last SAVE_TAPE forward
extern "C" int64_t ggnn_grad_floats(const ggnn_ctx* ctx) {
  if (!ctx || !ctx->gru_set) return 0;
  return (int64_t)ggnn_grad_count(ctx->two_e, ctx->h);
}

// N4 ------------------------------------------------------------------------------------
extern "C" int ggnn_optimizer_step(ggnn_ctx* ctx, int slot, const float* grads, float lr, float beta1,
                                   float beta2, float eps, float clip_norm, void* stream) {
  if (!ctx) return GGNN_E_ARG;
  if (slot < 0 || slot >= GGNN_MAX_SLOTS || !grads) GGNN_FAIL(ctx, GGNN_E_ARG, "optimizer_step: bad slot or null grads");
  if (misaligned16(grads)) GGNN_FAIL(ctx, GGNN_E_ARG, "optimizer_step: grads must be 16-byte aligned");
  if (!ctx->gru_set || !ctx->edge[slot].set) GGNN_FAIL(ctx, GGNN_E_STATE, "optimizer_step: weights for slot %d not set", slot);
  device_guard g(ctx->device);
  cudaStream_t st = (cudaStream_t)stream;
  ctx_ext* ex = ext_of(ctx);
  const int h = ctx->h, two_e = ctx->two_e;
  const int64_t n = (int64_t)ggnn_grad_count(two_e, h);
  if (ex->adam_n != n) {                      // (re)create zeroed moments
    GGNN_TRY(arena_reserve(ctx, 6, (size_t)(2 * n + 6 * GGNN_SUMSQ_SLOT) * 4, st));
    GGNN_CUDA(ctx, cudaMemsetAsync(ex->ar[6].p, 0, (size_t)(2 * n + 6 * GGNN_SUMSQ_SLOT) * 4, st));
    ex->adam_n = n;
    ex->adam_t = 0;
  }
  float* m = reinterpret_cast<float*>(ex->ar[6].p);
  float* v = m + n;
  float* norms = v + n;                       // 6 slots of per-block partial sums
  const int64_t t = ++ex->adam_t;
  const float lr_t = lr * sqrtf(1.f - powf(beta2, (float)t)) / (1.f - powf(beta1, (float)t));
  ggnn_edge_weights& ew = ctx->edge[slot];
  float* params[6] = {ew.W, ew.B, ctx->Wg, ctx->bg, ctx->Wc, ctx->bc};
  const int64_t sizes[6] = {(int64_t)two_e * h * h, (int64_t)two_e * h, (int64_t)4 * h * h, 2 * h, (int64_t)2 * h * h, h};
  int64_t off = 0;
  for (int i = 0; i < 6; ++i) {
    GGNN_TRY(ggnn_adam_segment(ctx, params[i], grads + off, m + off, v + off, sizes[i], norms + (size_t)i * GGNN_SUMSQ_SLOT, clip_norm, lr_t,
                               beta1, beta2, eps, st));
    off += sizes[i];
  }
  // the tensor-core copies follow the FP32 master weights
  ctx->repack_in_place = true;
  int rc = ggnn_tf32_pack_weights(ctx, slot, true, st);
  if (rc == GGNN_OK) rc = ggnn_f16_pack_weights(ctx, slot, true, st);
  if (rc == GGNN_OK) rc = ggnn_w16_pack_weights(ctx, slot, true, st);
  ctx->repack_in_place = false;
  if (rc != GGNN_OK) return rc;
  ext_of(ctx)->tape.valid = false;
  return GGNN_OK;
}

extern "C" int ggnn_get_weights(ggnn_ctx* ctx, int slot, float* W, float* B, float* Wg, float* bg, float* Wc,
                                float* bc, void* stream) {
  if (!ctx) return GGNN_E_ARG;
  if (slot < 0 || slot >= GGNN_MAX_SLOTS) GGNN_FAIL(ctx, GGNN_E_ARG, "get_weights: bad slot");
  if (!ctx->gru_set || !ctx->edge[slot].set) GGNN_FAIL(ctx, GGNN_E_STATE, "get_weights: weights for slot %d not set", slot);
  device_guard g(ctx->device);
  cudaStream_t st = (cudaStream_t)stream;
  const int h = ctx->h, two_e = ctx->two_e;
  const ggnn_edge_weights& ew = ctx->edge[slot];
  auto cp = [&](float* dst, const float* src, size_t n) -> int {
    if (dst && src) GGNN_CUDA(ctx, cudaMemcpyAsync(dst, src, n * 4, cudaMemcpyDeviceToDevice, st));
    return GGNN_OK;
  };
  GGNN_TRY(cp(W, ew.W, (size_t)two_e * h * h));
  GGNN_TRY(cp(B, ew.B, (size_t)two_e * h));
  GGNN_TRY(cp(Wg, ctx->Wg, (size_t)4 * h * h));
  GGNN_TRY(cp(bg, ctx->bg, (size_t)2 * h));
  GGNN_TRY(cp(Wc, ctx->Wc, (size_t)2 * h * h));
  GGNN_TRY(cp(bc, ctx->bc, (size_t)h));
  return GGNN_OK;
}

extern "C" int ggnn_set_dropout_seed(ggnn_ctx* ctx, float keep, unsigned long long seed, int drop_weights,
                                     int drop_state) {
  if (!ctx) return GGNN_E_ARG;
  if (!(keep > 0.f && keep <= 1.f)) GGNN_FAIL(ctx, GGNN_E_ARG, "set_dropout_seed: keep must be in (0, 1]");
  ctx_ext* ex = ext_of(ctx);
  ex->drop_T = -1; ex->drop_w = ex->drop_s = nullptr;
  ex->drop_gen = (drop_weights || drop_state) && keep < 1.f;
  ex->drop_gen_w = drop_weights != 0; ex->drop_gen_s = drop_state != 0;
  ex->drop_keep = keep; ex->drop_seed = seed;
  return GGNN_OK;
}

extern "C" int ggnn_dropout_mask(ggnn_ctx* ctx, float* out, int64_t n, unsigned long long seed,
                                 unsigned long long stream_id, float keep, void* stream) {
  if (!ctx) return GGNN_E_ARG;
  if (!out && n > 0) GGNN_FAIL(ctx, GGNN_E_ARG, "dropout_mask: null output");
  device_guard g(ctx->device);
  return ggnn_dropout_mask_fill(ctx, out, n, seed, stream_id, keep, (cudaStream_t)stream);
}

extern "C" int ggnn_set_dropout(ggnn_ctx* ctx, const float* w_mask, const float* state_mask, float keep, int T) {
  if (!ctx) return GGNN_E_ARG;
  if (!(keep > 0.f && keep <= 1.f) || T < 0) GGNN_FAIL(ctx, GGNN_E_ARG, "set_dropout: keep must be in (0, 1], T >= 0");
  ctx_ext* ex = ext_of(ctx);
  ex->drop_gen = false;
  ex->drop_w = w_mask; ex->drop_s = state_mask; ex->drop_keep = keep;
  ex->drop_T = (w_mask || state_mask) ? T : -1;
  return GGNN_OK;
}

extern "C" int ggnn_propagate_backward(ggnn_ctx* ctx, const float* d_hT, float* d_h0, float* grads,
                                       void* stream) {
  if (!ctx) return GGNN_E_ARG;
  ggnn_tape& tape = ext_of(ctx)->tape;
  if (!tape.valid) GGNN_FAIL(ctx, GGNN_E_STATE, "propagate_backward: no tape (run a propagate call with GGNN_FLAG_SAVE_TAPE first)");
  if (tape.plan.N > 0 && (!d_hT || !d_h0 || !grads)) GGNN_FAIL(ctx, GGNN_E_ARG, "propagate_backward: null pointer");
  if (misaligned16(d_hT) || misaligned16(d_h0) || misaligned16(grads))
    GGNN_FAIL(ctx, GGNN_E_ARG, "propagate_backward: pointers must be 16-byte aligned");
  device_guard g(ctx->device);
  cudaStream_t st = (cudaStream_t)stream;
  ggnn_bwd_scratch sc;
  {
    ws_alloc dry(nullptr);
    ggnn_bwd_take_scratch(dry, tape.plan, ctx->h, ctx->two_e, tape.w_mask != nullptr, &sc);
    if (tape.tf32) ggnn_bwd_take_nt_weights(dry, ctx, &sc);
    GGNN_TRY(arena_reserve(ctx, 3, dry.off + 256, st));
  }
  ws_alloc a(ext_of(ctx)->ar[3].p);
  ggnn_bwd_take_scratch(a, tape.plan, ctx->h, ctx->two_e, tape.w_mask != nullptr, &sc);
  if (tape.tf32) ggnn_bwd_take_nt_weights(a, ctx, &sc);
  return ggnn_backward_fp32(ctx, tape, sc, d_hT, d_h0, grads, st);
}

// ---------------------------------------------------------------------------------------
// N1 / N2: the rows either side of the path (evaluation mode)
extern "C" int ggnn_embed_nodes(ggnn_ctx* ctx, const int32_t* idx, int64_t N, const float* loc_tab,
                                int loc_rows, int loc_dim, const float* pos_tab, int pos_rows, int pos_dim,
                                const float* word_tab, int word_rows, int word_dim, int h, float* h0,
                                void* stream) {
  if (!ctx) return GGNN_E_ARG;
  if (N < 0 || h <= 0 || loc_dim < 0 || pos_dim < 0 || word_dim < 0 || loc_rows <= 0 || pos_rows <= 0 || word_rows <= 0)
    GGNN_FAIL(ctx, GGNN_E_ARG, "embed_nodes: bad sizes");
  if (2 * loc_dim + pos_dim + word_dim > h)
    GGNN_FAIL(ctx, GGNN_E_SHAPE, "embed_nodes: embeddings (%d) wider than the hidden size %d (the reference pads with h - width zeros, chem_tensorflow_dense.py:304-306)",
              2 * loc_dim + pos_dim + word_dim, h);
  if (N > 0 && (!idx || !loc_tab || !pos_tab || !word_tab || !h0)) GGNN_FAIL(ctx, GGNN_E_ARG, "embed_nodes: null pointer");
  if (N == 0) return GGNN_OK;
  device_guard g(ctx->device);
  cudaStream_t st = (cudaStream_t)stream;
  int* bad = reinterpret_cast<int*>(ctx->d_mail + 3);
  GGNN_CUDA(ctx, cudaMemsetAsync(bad, 0, sizeof(int64_t), st));
  GGNN_TRY(ggnn_embed_launch(ctx, idx, N, loc_tab, loc_rows, loc_dim, pos_tab, pos_rows, pos_dim, word_tab,
                             word_rows, word_dim, h, h0, bad, st));
  int64_t nbad = 0;
  GGNN_TRY(read_i64(ctx, ctx->d_mail + 3, &nbad, st));
  if (nbad) GGNN_FAIL(ctx, GGNN_E_ARG, "embed_nodes: %lld nodes have an index outside its embedding table", (long long)nbad);
  return GGNN_OK;
}

extern "C" int ggnn_heads_decode(ggnn_ctx* ctx, const float* hT, const float* h0, int64_t N, int h,
                                 const float* Wcat, const float* bcat, int ncat, int n_head, int n_lab,
                                 int lab_off, const int32_t* n_active, int32_t* pred, float* p_head,
                                 float* p_lab, void* stream) {
  if (!ctx) return GGNN_E_ARG;
  if (N < 0 || h <= 0 || h % 4 || ncat <= 0 || ncat % 4 || n_head <= 0 || n_lab <= 0 || lab_off < n_head ||
      lab_off % 4 || lab_off + n_lab > ncat)
    GGNN_FAIL(ctx, GGNN_E_ARG, "heads_decode: bad sizes (h=%d ncat=%d n_head=%d n_lab=%d lab_off=%d)", h, ncat, n_head, n_lab, lab_off);
  if (N > 0 && (!hT || !h0 || !Wcat || !bcat || !n_active || !pred)) GGNN_FAIL(ctx, GGNN_E_ARG, "heads_decode: null pointer");
  if (N == 0) return GGNN_OK;
  device_guard g(ctx->device);
  cudaStream_t st = (cudaStream_t)stream;
  GGNN_TRY(arena_reserve(ctx, 3, (size_t)N * ncat * 4 + 256, st));
  float* logits = reinterpret_cast<float*>(ext_of(ctx)->ar[3].p);
  prof_scope ps(ctx, GGNN_STAGE_OTHER, st);
  GGNN_TRY(ggnn_sgemm_concat(ctx, hT, h0, h, h, h, Wcat, bcat, ncat, N, logits, st));
  return ggnn_decode_launch(ctx, logits, ncat, N, n_head, n_lab, lab_off, n_active, pred, p_head, p_lab, st);
}

extern "C" size_t ggnn_workspace_bytes(const ggnn_ctx* ctx, int64_t N, int64_t nnz, int two_e, int h) {
  (void)ctx;
  if (N < 0 || nnz < 0 || two_e <= 0 || h <= 0) return 0;
  const int64_t tiles_edge = nnz / GGNN_TILE_M + two_e + 1;
  const int64_t tiles_dense = (int64_t)align_up((size_t)N, GGNN_TILE_M) / GGNN_TILE_M * two_e;
  const int64_t tiles = nnz >= N * (int64_t)two_e ? tiles_dense : tiles_edge;
  size_t bytes = (size_t)tiles * GGNN_TILE_M * (h + 1) * 4;       // messages + msg_src
  bytes += (size_t)N * h * 4 * 3;                                  // acts, rh, u
  bytes += (size_t)nnz * (4 + 4 + 8) + (size_t)N * 12;             // col/tmsg/keys, tptr...
  return bytes + ((size_t)4 << 20);
}
