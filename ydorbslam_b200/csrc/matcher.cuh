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
  cudaEvent_t ev_dep;     // orders the matcher's stream after the extractor's (no host wait in between)
  cudaEvent_t ev_block;   // blocking-sync event: batch calls sleep on it instead of spinning in cudaStreamSynchronize
};

namespace ydorb {
int matcher_reserve(YdMatcher* m, size_t bytes);
int matcher_reserve_io(YdMatcher* m, size_t bytes);
// caller-supplied octaves / levels index fixed per-level tables on the device: reject anything outside [0, n_levels)
inline bool octaves_in_range(const YdKeyPoint* k, int n, int n_levels) {
  for (int i = 0; i < n; i++)
    if (k[i].octave < 0 || k[i].octave >= n_levels) return false;
  return true;
}
}  // namespace ydorb
