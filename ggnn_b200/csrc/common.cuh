// common.cuh -- ctx, error plumbing and small device helpers shared by the kernels.
#pragma once

#include <cuda.h>
#include <cuda_runtime.h>
#include <stdint.h>
#include <stdio.h>
#include <string.h>
#include <vector>

#include "../../include/ggnn_b200.h"

#define GGNN_MAX_SLOTS 2

// One set of edge parameters (slot) as the kernels want them.
struct ggnn_edge_weights {
  float* W = nullptr;      // [two_e, h, h]  FP32, reference layout (K rows, N cols)
  float* B = nullptr;      // [two_e, h] or nullptr
  float* Wt_tf32 = nullptr;// [two_e, hp, hp] TF32-rounded, transposed ([n][k]), zero padded to hp
  float* Wt_tf32_lo = nullptr; // same layout, tf32(W - tf32(W)): the residual pack of the strict (3xTF32) mode
  CUtensorMap tm_Wt_lo;
  float* Wt_tf32_hl = nullptr; // hp == 128: [two_e][hi | lo][hp][hp], one 32 KB TMA box brings both packs of a K chunk
  CUtensorMap tm_Wt_hl;
  void*  Wt_f16 = nullptr; // [two_e, P, P] FP16, transposed ([n][k]), zero padded; P = 128 (hp == 128) or hw16 (h > 128)
  CUtensorMap tm_Wt16;     // TMA view of Wt_f16 as [two_e*128 rows, 128 cols], box 64 x 128
  CUtensorMap tm_Wt16q;    // same, box 64 x 64 rows (gru_pair.cu: half of a matrix's output columns per CTA)
  float* B_pad16 = nullptr;// [two_e, hw16] FP32 zero-padded biases (wide16.cu, h > 128)
  // word-order channels E-1 / 2E-1 folded through the GRU input weights (f16.cu, hp == 128, two_e even)
  void*  fold_f16 = nullptr;   // [2][384][128] FP16, K-major
  float* fold_bias = nullptr;  // [3][384] scaled biases of a word-order tile
  CUtensorMap tm_fold;
  bool   fold_ok = false;
  float  absmax_w = -1.f;  // largest column L1 norm of any W_e, max|B| of this slot (host copy, < 0 = not computed yet)
  float  absmax_b = 0.f;
  CUtensorMap tm_Wt;       // TMA view of Wt_tf32 as [two_e*hp rows, hp cols], box 32 x tm_rows
  int    tm_rows = 0;
  bool   set = false;
};

struct ggnn_ctx {
  int device = 0;
  int sm_count = 148;
  unsigned attr_mask = 0;   // kernels whose dynamic-smem attribute has been set on THIS ctx's device (GGNN_ATTR_*)
  int two_e = 0;
  int h = 0;
  int hp = 0;              // h padded to the tensor-core tile granularity
  bool hp_f16_narrow = true; // the FP16 packs are the 128-pitch ones of f16.cu
  int hw16 = 0;            // h > 128: h padded to 64 (pitch of the FP16 buffers of wide16.cu), else 0
  ggnn_edge_weights edge[GGNN_MAX_SLOTS];
  // GRU cell (shared by both slots)
  float* Wg = nullptr;     // [2h, 2h]
  float* bg = nullptr;     // [2h]
  float* Wc = nullptr;     // [2h, h]
  float* bc = nullptr;     // [h]
  float* Wgt_tf32 = nullptr; // [2hp, 2hp] transposed + padded + TF32-rounded
  float* Wct_tf32 = nullptr; // [hp, 2hp]
  float* Wgt_tf32_lo = nullptr, *Wct_tf32_lo = nullptr;   // residual packs (strict mode)
  CUtensorMap tm_Wg_lo, tm_Wc_lo, tm_Wg_half;   // (tm_Wg_lo / tm_Wg_half: boxes of hp rows)
  float* Wgt_tf32_hl = nullptr, *Wct_tf32_hl = nullptr;   // hp == 128: [half][hi | lo][hp][2hp] and [hi | lo][hp][2hp]
  CUtensorMap tm_Wg_hl, tm_Wc_hl;
  bool hl_ok = false;
  void* Wgt_f16 = nullptr;   // [256, 256] FP16 packed like Wgt_tf32 (hp == 128 only)
  void* Wct_f16 = nullptr;   // [128, 256]
  CUtensorMap tm_Wg16, tm_Wc16;
  CUtensorMap tm_Wc16q, tm_Wg16q;   // boxes of 64 x 64 rows (gru_pair.cu)
  float* btab16 = nullptr;   // [slots][640] gru_pair.cu's bias table (gp::bias_tab)
  bool   btab_ok[GGNN_MAX_SLOTS] = {false, false};
  bool   pair_ok = false;    // tm_Wc16q / tm_Wg16q are current
  int    pair_fit = -1;      // a CTA pair of gru_pair_kernel can be resident on this device: 1 / 0, -1 = not asked yet
  CUtensorMap tm_Wg16h, tm_Wc16h;   // same, boxes of hw16 / 2 rows (CTA pairs)
  CUtensorMap tm_Wg16f, tm_Wc16f;   // wide16.cu fused GRU kernel (hw16 <= 256): boxes of hw16 rows
  float* bg_w16 = nullptr;   // [2*hw16], [hw16] zero-padded GRU biases (wide16.cu)
  float* bc_w16 = nullptr;
  CUtensorMap tm_Wg, tm_Wc, tm_Wg256;  // TMA views of the packed GRU kernels (256-row box: fused kernel)
  unsigned long long* dbg_trace = nullptr;   // device buffer for kernel phase traces (ggnn_debug_trace)
  int dbg_trace_wide = 0;                    // env GGNN_PAIR_TRACE_WIDE=1: the buffer also holds gru_pair.cu's whole-grid region
  bool   no_fused = false;   // debugging / A-B switch (GGNN_NO_FUSED=1)
  bool   repack_in_place = false; // optimizer step: shapes unchanged, packed weight buffers are reused
  bool   no_bits_plan = false; // A-B switch (GGNN_NO_BITS_PLAN=1): dense path goes through the canonical CSR
  int    tm_rows_g = 0, tm_rows_c = 0;
  float* bg_pad = nullptr;   // [2hp]
  float* bc_pad = nullptr;   // [hp]
  bool   gru_set = false;
  // grow-only scratch arena
  char*  ws = nullptr;
  size_t ws_bytes = 0;
  // small pinned host mailbox for device->host counts
  int64_t* h_mail = nullptr;
  int64_t* d_mail = nullptr;
  int64_t launches = 0;
  char err[512] = {0};
  // optional per-stage CUDA-event instrumentation (ggnn_profile_*)
  bool prof_on = false;
  struct prof_rec { int stage; cudaEvent_t a, b; };
  std::vector<prof_rec> prof_recs;
  std::vector<cudaEvent_t> prof_pool;
  double prof_ms[GGNN_N_STAGES] = {0};
  int64_t prof_launches[GGNN_N_STAGES] = {0};
};

// cudaFuncSetAttribute is per device: remember per ctx (a process may own one ctx per GPU), not in a static
enum { GGNN_ATTR_TF32_GEMM0 = 0 /* + mode 0..3 */, GGNN_ATTR_GRU_FUSED = 4, GGNN_ATTR_MSG_TC = 5, GGNN_ATTR_F16 = 6,
       GGNN_ATTR_TN = 7, GGNN_ATTR_W16_0 = 8 /* + mode 0..2, 3 = fused GRU, 4 = fused GRU on CTA pairs */, GGNN_ATTR_BMQ = 13, GGNN_ATTR_STRICT_TC = 14,
       GGNN_ATTR_GRU_PAIR = 15 };
#define GGNN_ENSURE_SMEM(ctx, bit, kernel, bytes)                                                          \
  do {                                                                                                     \
    if (!((ctx)->attr_mask & (1u << (bit)))) {                                                             \
      GGNN_CUDA(ctx, cudaFuncSetAttribute(kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, (bytes))); \
      (ctx)->attr_mask |= 1u << (bit);                                                                     \
    }                                                                                                      \
  } while (0)

// RAII stage timer: records an event pair around the launches of one stage when profiling is on
struct prof_scope {
  ggnn_ctx* c; cudaStream_t st; cudaEvent_t b = nullptr; int stage; int64_t l0;
  prof_scope(ggnn_ctx* ctx, int stage_, cudaStream_t s) : c(ctx), st(s), stage(stage_), l0(ctx->launches) {
    if (!c->prof_on) return;
    cudaEvent_t a;
    auto get = [&]() { cudaEvent_t e; if (c->prof_pool.empty()) cudaEventCreate(&e); else { e = c->prof_pool.back(); c->prof_pool.pop_back(); } return e; };
    a = get(); b = get();
    cudaEventRecord(a, st);
    c->prof_recs.push_back({stage, a, b});
  }
  ~prof_scope() {
    if (!c->prof_on || !b) return;
    cudaEventRecord(b, st);
    c->prof_launches[stage] += c->launches - l0;
  }
};

#define GGNN_FAIL(ctx, code, ...)                                   \
  do {                                                              \
    if (ctx) snprintf((ctx)->err, sizeof((ctx)->err), __VA_ARGS__); \
    return (code);                                                  \
  } while (0)

#define GGNN_CUDA(ctx, expr)                                                         \
  do {                                                                               \
    cudaError_t _e = (expr);                                                         \
    if (_e != cudaSuccess) {                                                         \
      GGNN_FAIL(ctx, GGNN_E_CUDA, "%s failed: %s (%s:%d)", #expr,                    \
                cudaGetErrorString(_e), __FILE__, __LINE__);                         \
    }                                                                                \
  } while (0)

#define GGNN_TRY(expr)            \
  do {                            \
    int _r = (expr);              \
    if (_r != GGNN_OK) return _r; \
  } while (0)

// post-launch check + launch counter
#define GGNN_LAUNCHED(ctx)                 \
  do {                                     \
    (ctx)->launches++;                     \
    GGNN_CUDA(ctx, cudaGetLastError());    \
  } while (0)

static inline size_t align_up(size_t x, size_t a) { return (x + a - 1) / a * a; }
static inline int64_t ceil_div64(int64_t a, int64_t b) { return (a + b - 1) / b; }

// Bump allocator over the ctx scratch arena.  Sizes are computed first (dry run),
// the arena is grown once, then the same sequence hands out pointers.
struct ws_alloc {
  char* base;
  size_t off = 0;
  explicit ws_alloc(char* b) : base(b) {}
  template <typename T>
  T* take(size_t n) {
    off = align_up(off, 256);
    T* p = base ? reinterpret_cast<T*>(base + off) : nullptr;
    off += n * sizeof(T);
    return p;
  }
};


// ---- scan.cu
int ggnn_exclusive_scan_i32(ggnn_ctx* ctx, const int32_t* in, int32_t* out, int64_t n,
                            int64_t* total_out_dev, int32_t* tmp, size_t tmp_elems,
                            cudaStream_t st);
size_t ggnn_scan_tmp_elems(int64_t n);

// ---- compact.cu
size_t ggnn_compact_tmp_bytes(int64_t total_elems, int64_t rows);
int ggnn_compact_bitmask(ggnn_ctx* ctx, const void* adj, int adj_dtype, int64_t total, uint32_t* bits,
                         cudaStream_t st);
bool ggnn_bitmask_queue_supported(const void* adj, int64_t total);
size_t ggnn_bitmask_queue_ctl_bytes();
int ggnn_bitmask_queue_launch(ggnn_ctx* ctx, const void* adj, int adj_dtype, int64_t total, uint32_t* bits,
                              unsigned* ctl, bool background, cudaStream_t st);
int ggnn_compact_count(ggnn_ctx* ctx, const void* adj, int adj_dtype, int64_t rows, int v,
                       int32_t* rowptr, int64_t* nnz_out_dev, char* tmp, cudaStream_t st);
int ggnn_compact_fill(ggnn_ctx* ctx, int64_t rows, int v, const int32_t* rowptr, int32_t* col,
                      int64_t col_capacity, const char* tmp, cudaStream_t st);
