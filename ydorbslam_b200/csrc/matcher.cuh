// Matcher-side handles (opaque to C callers).
#pragma once
#include "common.cuh"

struct YdMatcher {
  int device;
  int sm_count;
  int top_ctas_per_sm;   // grid size of the chunked top-2 kernel, in CTAs per SM
  cudaStream_t stream;
  cudaStream_t own_stream;
  void* d_scratch;       // kernel scratch (per-chunk partial results, candidate lists)
  size_t scratch_bytes;
  void* d_io;            // staging for host entry points
  size_t io_bytes;
  void* d_pairs;         // fused top-2 scan: [2][pairs_cap] u64 (best | second, all-ones between calls) + ticket counters
  int pairs_cap;
};

namespace ydorb {
int matcher_reserve(YdMatcher* m, size_t bytes);
int matcher_reserve_io(YdMatcher* m, size_t bytes);
}  // namespace ydorb
