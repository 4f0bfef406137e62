// plan_bits.cuh -- plan straight from the adjacency bitmask (dense path fast route): declarations.
#pragma once
#include "plan.cuh"

// arrays that do not depend on the edge count (allocated before the planning synchronisation)
struct ggnn_plan_bits_pre {
  uint32_t* bits = nullptr;      // 1 bit per adjacency element
  uint16_t* rowcnt = nullptr;    // [b*2E*v] popcount of every canonical row
  int32_t* segposT = nullptr;    // [2E*b + 1] exclusive scan of per-(channel, graph) edge counts
  int32_t* tptr = nullptr;       // [N + 1]    exclusive scan of node in-degrees (becomes plan.tptr)
  int32_t* regular = nullptr;    // [b]        1 = word-order graph whose channels E-1 / 2E-1 were left out (word_fold)
  int32_t* scan_tmp = nullptr;
  size_t scan_elems = 0;
};

bool ggnn_plan_bits_supported(int two_e, int v);
void ggnn_plan_bits_take_pre(ws_alloc& a, int64_t b, int two_e, int v, ggnn_plan_bits_pre* pre);
// needs pre.bits filled (ggnn_compact_bitmask); *nnz_out_dev receives the edge count
// word_fold (v == 128 only): a graph whose channel E-1 holds exactly the edges t-1 -> t (t = 1..127) and whose channel
// 2E-1 holds exactly t+1 -> t (t = 0..126) -- the word-order channels of graph_to_adj_mat_bd for a 127-word sentence,
// chem_tensorflow_dense.py:76-80 -- is marked in pre.regular and those two channels count as EMPTY in everything
// the plan is built from
int ggnn_plan_bits_count(ggnn_ctx* ctx, const ggnn_plan_bits_pre& pre, int64_t b, int two_e, int v, bool word_fold,
                         int64_t* nnz_out_dev, cudaStream_t st);
int ggnn_plan_bits_place(ggnn_ctx* ctx, const ggnn_plan_bits_pre& pre, int64_t b, int two_e, int v,
                         int64_t nnz, ggnn_plan* plan, ws_alloc& a, bool dry, cudaStream_t st);
