// plan.cuh -- the per-batch "plan": everything the per-timestep kernels need to know
// about the graph structure, derived once from the canonical CSR (dense path) or from
// an edge list (sparse path).
#pragma once
#include "common.cuh"

constexpr int GGNN_TILE_M = 128;   // message rows per GEMM tile (one edge type per tile)

struct ggnn_plan {
  int64_t N = 0;         // nodes (b*v for the dense path)
  int64_t nnz = 0;       // edges
  int64_t max_tiles = 0; // host-side bound on the number of message tiles
  bool dense_rows = false; // message rows are ALL (node, channel) pairs and may feed several targets
  int32_t* n_tiles = nullptr;   // [1]  device: actual number of message tiles
  int32_t* msg_src = nullptr;   // [max_tiles*128] global source node of a message row, -1 = padding
  int32_t* tile_type = nullptr; // [max_tiles]     edge channel of a tile
  int32_t* tptr = nullptr;      // [N+1]           CSR by target node ...
  int32_t* tmsg = nullptr;      // [nnz]           ... of message-row ids, ascending (channel, source)
  // GGNN_PREC_F16, h <= 128, dense route, v == 128: [N/128] 1 = the graph's word-order channels E-1 / 2E-1 are
  // exactly t-1 -> t and t+1 -> t; their edges are NOT in msg_src / tmsg / nnz, the GRU kernel forms them from the
  // resident h tile.  nullptr = every edge is a message row.
  int32_t* tile_regular = nullptr;
};

// Builders take their arrays from the bump allocator in a fixed order.  With dry=true
// they only account for the bytes (a.base == nullptr) and launch nothing.
int ggnn_plan_from_csr(ggnn_ctx* ctx, const int32_t* rowptr, const int32_t* col, int64_t b,
                       int two_e, int v, int64_t nnz, ggnn_plan* plan, ws_alloc& a, bool dry,
                       cudaStream_t st);

int ggnn_plan_from_edges(ggnn_ctx* ctx, const int32_t* edges, int64_t nnz, int64_t N, int two_e,
                         ggnn_plan* plan, ws_alloc& a, bool dry, cudaStream_t st);

// ---- step kernels (fp32 path), step_fp32.cu
struct ggnn_step_bufs {
  float* msg = nullptr;   // [max_tiles*128, h]
  float* acts = nullptr;  // [N, h]
  float* rh = nullptr;    // [N, h]
  float* u = nullptr;     // [N, h]
  float* tape_r = nullptr;  // training tape (FP32 path only): reset gate r of this step, or null
  float* tape_c = nullptr;  // training tape: candidate c of this step, or null
  const float* W_override = nullptr;  // training with edge-weight dropout: W * mask / keep of this step
};
int ggnn_step_fp32(ggnn_ctx* ctx, int slot, const ggnn_plan& plan, const ggnn_step_bufs& bufs,
                   float* h_state, cudaStream_t st);
int ggnn_aggregate(ggnn_ctx* ctx, const ggnn_plan& plan, const float* msg, float* acts, int h,
                   bool round_tf32, cudaStream_t st);
int ggnn_aggregate_f16(ggnn_ctx* ctx, const ggnn_plan& plan, const void* msg_half, float* acts, int h,
                       bool round_tf32, cudaStream_t st);
