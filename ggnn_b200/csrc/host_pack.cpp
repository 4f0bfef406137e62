// host_pack.cpp -- the adjacency bitmask of a HOST-resident dense adjacency, computed on the host cores before the
// tensor crosses the bus (ggnn_host_pack_adjacency, include/ggnn_b200.h).
//
// Why: the reference feeds its dense FP32 adjacency from host memory every step (feed_dict, chem_tensorflow.py:262-268 /
// chem_tensorflow_dense.py:566-602): 6.03 MB per C3 graph, of which the path uses one bit per element (A != 0).  Over
// PCIe that is 9.1 k graphs/s at the measured 55.6 GB/s whatever the GPU does.  The bit per element is the SAME
// intermediate the GPU's own first pass produces (compact.cu bitmask kernels: word w, bit i <=> element 32 w + i != 0;
// tests/test_gpu_hostpack.py compares the two bit for bit), so for host inputs it is formed where the data is and 1/32
// of the bytes travel.  Everything after the bitmask -- row counts, plan, messages, GRU -- runs on the GPU as for
// device-resident inputs, which keep the GPU bitmask pass.
//
// Plain C++ (std::thread), AVX2 / AVX-512 picked at run time; no CUDA in this file.
#include <stdint.h>
#include <string.h>
#include <atomic>
#include <thread>
#include <vector>
#include <immintrin.h>
#include "../../include/ggnn_b200.h"

namespace {

// ---- FP32: 32 floats -> one word --------------------------------------------------------------------------------
void pack_f32_scalar(const float* a, int64_t words, uint32_t* out) {
  for (int64_t w = 0; w < words; ++w) {
    uint32_t m = 0;
    for (int i = 0; i < 32; ++i) m |= (a[w * 32 + i] != 0.0f ? 1u : 0u) << i;
    out[w] = m;
  }
}
__attribute__((target("avx2"))) void pack_f32_avx2(const float* a, int64_t words, uint32_t* out) {
  const __m256 z = _mm256_setzero_ps();
  for (int64_t w = 0; w < words; ++w) {
    const float* p = a + w * 32;
    const uint32_t m0 = (uint32_t)_mm256_movemask_ps(_mm256_cmp_ps(_mm256_loadu_ps(p), z, _CMP_NEQ_UQ));
    const uint32_t m1 = (uint32_t)_mm256_movemask_ps(_mm256_cmp_ps(_mm256_loadu_ps(p + 8), z, _CMP_NEQ_UQ));
    const uint32_t m2 = (uint32_t)_mm256_movemask_ps(_mm256_cmp_ps(_mm256_loadu_ps(p + 16), z, _CMP_NEQ_UQ));
    const uint32_t m3 = (uint32_t)_mm256_movemask_ps(_mm256_cmp_ps(_mm256_loadu_ps(p + 24), z, _CMP_NEQ_UQ));
    out[w] = m0 | (m1 << 8) | (m2 << 16) | (m3 << 24);
  }
}
__attribute__((target("avx512f"))) void pack_f32_avx512(const float* a, int64_t words, uint32_t* out) {
  const __m512 z = _mm512_setzero_ps();
  for (int64_t w = 0; w < words; ++w) {
    const float* p = a + w * 32;
    const uint32_t m0 = (uint32_t)_mm512_cmp_ps_mask(_mm512_loadu_ps(p), z, _CMP_NEQ_UQ);
    const uint32_t m1 = (uint32_t)_mm512_cmp_ps_mask(_mm512_loadu_ps(p + 16), z, _CMP_NEQ_UQ);
    out[w] = m0 | (m1 << 16);
  }
}
// ---- uint8: 32 bytes -> one word --------------------------------------------------------------------------------
void pack_u8_scalar(const uint8_t* a, int64_t words, uint32_t* out) {
  for (int64_t w = 0; w < words; ++w) {
    uint32_t m = 0;
    for (int i = 0; i < 32; ++i) m |= (a[w * 32 + i] != 0 ? 1u : 0u) << i;
    out[w] = m;
  }
}
__attribute__((target("avx2"))) void pack_u8_avx2(const uint8_t* a, int64_t words, uint32_t* out) {
  const __m256i z = _mm256_setzero_si256();
  for (int64_t w = 0; w < words; ++w) {
    const __m256i x = _mm256_loadu_si256(reinterpret_cast<const __m256i*>(a + w * 32));
    out[w] = ~(uint32_t)_mm256_movemask_epi8(_mm256_cmpeq_epi8(x, z));
  }
}

template <typename T>
void pack_range(const T* a, int64_t words, uint32_t* out) {
  static const bool has512 = __builtin_cpu_supports("avx512f"), has2 = __builtin_cpu_supports("avx2");
  if (sizeof(T) == 4) {
    const float* f = reinterpret_cast<const float*>(a);
    if (has512) pack_f32_avx512(f, words, out);
    else if (has2) pack_f32_avx2(f, words, out);
    else pack_f32_scalar(f, words, out);
  } else {
    const uint8_t* u = reinterpret_cast<const uint8_t*>(a);
    if (has2) pack_u8_avx2(u, words, out);
    else pack_u8_scalar(u, words, out);
  }
}

}  // namespace

extern "C" int ggnn_host_pack_adjacency(const void* adj_host, int adj_dtype, int64_t n_elems, uint32_t* bits_host, int threads) {
  if (n_elems < 0 || (n_elems > 0 && (!adj_host || !bits_host))) return GGNN_E_ARG;
  if (adj_dtype != GGNN_ADJ_F32 && adj_dtype != GGNN_ADJ_U8) return GGNN_E_ARG;
  const int64_t words = n_elems / 32, tail = n_elems % 32;
  if (threads <= 0) {
    threads = (int)std::thread::hardware_concurrency();
    if (threads <= 0) threads = 1;
  }
  if (threads > 64) threads = 64;
  const int64_t min_words = 1 << 14;                       // 2 MB of FP32 per thread at least
  if ((int64_t)threads > (words + min_words - 1) / min_words) threads = (int)((words + min_words - 1) / min_words);
  if (threads < 1) threads = 1;
  // blocks of 64 K words (8 MB of FP32) handed out by a counter: a core that is busy with something else (the caller's
  // own thread, another rank) delays one block, not a fixed 1/threads share of the tensor
  const int64_t blk = 1 << 16;
  std::atomic<int64_t> next(0);
  auto work = [&](int) {
    for (;;) {
      const int64_t w0 = next.fetch_add(blk, std::memory_order_relaxed);
      if (w0 >= words) break;
      const int64_t w1 = w0 + blk < words ? w0 + blk : words;
      if (adj_dtype == GGNN_ADJ_F32) pack_range(reinterpret_cast<const float*>(adj_host) + w0 * 32, w1 - w0, bits_host + w0);
      else pack_range(reinterpret_cast<const uint8_t*>(adj_host) + w0 * 32, w1 - w0, bits_host + w0);
    }
  };
  if (threads == 1) {
    work(0);
  } else {
    std::vector<std::thread> pool;
    pool.reserve(threads - 1);
    for (int t = 1; t < threads; ++t) pool.emplace_back(work, t);
    work(0);
    for (auto& th : pool) th.join();
  }
  if (tail) {                                              // last partial word (zero padded)
    uint32_t m = 0;
    for (int64_t i = 0; i < tail; ++i) {
      const bool nz = adj_dtype == GGNN_ADJ_F32 ? reinterpret_cast<const float*>(adj_host)[words * 32 + i] != 0.0f
                                                : reinterpret_cast<const uint8_t*>(adj_host)[words * 32 + i] != 0;
      m |= (nz ? 1u : 0u) << i;
    }
    bits_host[words] = m;
  }
  return GGNN_OK;
}
