// M2 OrbMatcher::searchByProjectionInLastAndCurrentFrame      (src/orbMatcher.cpp:65-155)
// M3 OrbMatcher::searchByProjectionInKeyFrameAndCurrentFrame  (:156-239)
// M4 searchByProjectionInSim (:240-302), searchBySim3 (:566-681, one call per direction), and the candidate search
//    of fuseByProjection (:682-745) / fuseBySim3 (:746-807)
// M7 computeThreeMaxima (:827-854)
//
// One kernel serves all of them: a thread owns one projected map point (YdWindowQuery), walks the grid cells of
// Frame::getKeyPointsInArea in the reference's order (ix outer, iy inner, insertion order inside a cell), applies the
// gates that differ between the functions (YdWindowParams) and keeps the best-1 candidate under `d < best`.
//
// Greedy exclusion.  In the reference a keypoint claimed by an earlier map point is skipped by later
// ones.  As in frame.cu, every query is evaluated in parallel against a blocker table B[idx] = lowest blocking
// claimant of keypoint idx, and the pass is repeated until B stops changing; by induction on the query index the
// fixed point is unique and equals the sequential loop.  Functions without exclusion take exactly one pass.
//
// Orientation check: one CTA bins every accepted query (bin = round(rotation / 30), :121-129), thread 0 runs
// computeThreeMaxima, and claims in the other bins are cleared (-2), which is what :144-152 does after the loop.
#include <algorithm>

#include "frame.cuh"

namespace ydorb {

constexpr int kHistLen = 30;  // OrbMatcher::m_int_histLen src/orbMatcher.cpp:9

// bin of the rotation histogram.  add_f32: searchForTriangulation adds 360.0f (:521), every other caller 360.0 (:123)
__device__ __forceinline__ int rotation_bin(float a, float b, bool add_f32) {
  const float factor = (float)(1.0 / kHistLen);  // :78
  float rot = __fsub_rn(a, b);
  if (rot < 0.0f) rot = add_f32 ? __fadd_rn(rot, 360.0f) : (float)((double)rot + 360.0);
  int bin = (int)round((double)__fmul_rn(rot, factor));
  if (bin == kHistLen) bin = 0;
  return bin;
}

__host__ __device__ inline void three_maxima(const int* sizes, int len, int& i1, int& i2, int& i3) {
  int max1 = 0, max2 = 0, max3 = 0;
  i1 = i2 = i3 = -1;
  for (int i = 0; i < len; i++) {
    const int n = sizes[i];
    if (n > max1) { max3 = max2; max2 = max1; max1 = n; i3 = i2; i2 = i1; i1 = i; }
    else if (n > max2) { max3 = max2; max2 = n; i3 = i2; i2 = i; }
    else if (n > max3) { max3 = n; i3 = i; }
  }
  if (max2 < max1 / 10) { i3 = -1; i2 = -1; }
  else if (max3 < max1 / 10) { i3 = -1; }
}

struct WindowArgs {
  GridParams gp;
  const YdKeyPoint* kps;
  const uint4* desc;
  const float* right_x;
  const int* cell_start;
  const int* items;
  const YdWindowQuery* q;
  const uint4* q_desc;
  int nq;
  int max_dist, strict, exclusive, chi2;
  float inv_sf2[kMaxLevels];
  const int* blocker;
  int* blocker_next;
  int* last_writer;
  int* best_of;
  int* flags;  // see k_window_search
};

// flags: [0]/[1] "round changed the blocker table" (even / odd rounds), [2] accepted count, [3] fixed point reached.
// Rounds are launched ahead of the host in batches; once [3] is set the remaining kernels of a batch return at once.
__global__ void __launch_bounds__(128) k_window_search(WindowArgs A) {
  const int j = blockIdx.x * blockDim.x + threadIdx.x;
  if (j >= A.nq || A.flags[3]) return;
  const YdWindowQuery q = A.q[j];
  int best_idx = -1;
  if (q.flags & YDORB_Q_ACTIVE) {
    const uint4 q0 = A.q_desc[2 * j], q1 = A.q_desc[2 * j + 1];
    int bestDist = 256;
    CellRange cr;
    if (cell_range(A.gp, q.x, q.y, q.r, cr)) {
      for (int ix = cr.x0; ix <= cr.x1; ix++)
        for (int iy = cr.y0; iy <= cr.y1; iy++) {
          const int c = ix * kGridRows + iy;
          const int e = A.cell_start[c + 1];
          for (int t = A.cell_start[c]; t < e; t++) {
            const int idx = A.items[t];
            const YdKeyPoint kp = A.kps[idx];
            if (!in_window(kp, q.x, q.y, q.r, q.min_level, q.max_level)) continue;
            if (A.exclusive && A.blocker[idx] < j) continue;                             // :106-107, :194, :284
            const float rx = A.right_x[idx];
            if (q.right_r >= 0.f && !(rx <= 0.f || fabsf(__fsub_rn(q.right_x, rx)) <= q.right_r)) continue;  // :108-109
            if (kp.octave < q.oct_lo || kp.octave > q.oct_hi) continue;                  // :285-286, :616, :716, :782
            if (A.chi2) {                                                                // :713-719
              const double dx = (double)__fsub_rn(kp.x, q.x), dy = (double)__fsub_rn(kp.y, q.y);
              const float mono = (float)(dx * dx + dy * dy);
              const double dr = (double)__fsub_rn(rx, q.right_x);
              const float stereo = (float)((double)mono + dr * dr);
              const float w = A.inv_sf2[kp.octave];
              const bool ok = (rx >= 0.f && (double)__fmul_rn(stereo, w) <= 7.81) ||
                              (rx < 0.f && (double)__fmul_rn(mono, w) <= 5.99);
              if (!ok) continue;
            }
            const int d = hamming256(q0, q1, A.desc[2 * idx], A.desc[2 * idx + 1]);
            if (d < bestDist) { bestDist = d; best_idx = idx; }                          // :111-114
          }
        }
    }
    const bool accept = best_idx >= 0 && (A.strict ? bestDist < A.max_dist : bestDist <= A.max_dist);
    if (!accept) best_idx = -1;
  }
  A.best_of[j] = best_idx;
  if (best_idx >= 0) {
    atomicAdd(&A.flags[2], 1);
    atomicMax(&A.last_writer[best_idx], j);
    if (A.exclusive && (q.flags & YDORB_Q_BLOCKS)) atomicMin(&A.blocker_next[best_idx], j);
  }
}

__global__ void k_win_blocker_init(const uint8_t* __restrict__ occupied, int n, int* __restrict__ b) {
  const int i = blockIdx.x * blockDim.x + threadIdx.x;
  if (i < n) b[i] = (occupied && occupied[i]) ? -1 : kNoBlocker;
}

__global__ void k_win_round_init(const int* __restrict__ b0, int n, int* __restrict__ bnext, int* __restrict__ last,
                                 int* __restrict__ flags, int round) {
  const int i = blockIdx.x * blockDim.x + threadIdx.x;
  // every thread derives the same verdict from values no thread of this kernel changes
  const bool done = flags[3] || (round > 0 && flags[(round - 1) & 1] == 0);
  if (i == 0) {
    if (done) flags[3] = 1;
    else { flags[round & 1] = 0; flags[2] = 0; }
  }
  if (done) return;
  if (i < n) {
    bnext[i] = (b0[i] == -1) ? -1 : kNoBlocker;
    last[i] = -1;
  }
}

__global__ void k_win_blocker_diff(const int* __restrict__ a, const int* __restrict__ b, int n, int* __restrict__ flags,
                                   int round) {
  const int i = blockIdx.x * blockDim.x + threadIdx.x;
  if (flags[3]) return;
  if (i < n && a[i] != b[i]) flags[round & 1] = 1;
}

// rotation histogram + three maxima + clearing, one CTA (:138-153).  src_angle[j] is the query's source angle;
// key_of[j] = best_of[j] is the slot that was pushed into the histogram.
__global__ void __launch_bounds__(1024)
k_orientation_cull(const YdWindowQuery* __restrict__ q, const YdKeyPoint* __restrict__ kps, const int* __restrict__ best_of,
                   int nq, int* __restrict__ last_writer, int* __restrict__ flags) {
  __shared__ int s_hist[kHistLen];
  __shared__ int s_keep[3];
  __shared__ int s_culled;
  if (threadIdx.x < kHistLen) s_hist[threadIdx.x] = 0;
  if (threadIdx.x == 0) s_culled = 0;
  __syncthreads();
  for (int j = threadIdx.x; j < nq; j += blockDim.x) {
    const int b = best_of[j];
    if (b >= 0) atomicAdd(&s_hist[rotation_bin(q[j].angle, kps[b].angle, false)], 1);
  }
  __syncthreads();
  if (threadIdx.x == 0) three_maxima(s_hist, kHistLen, s_keep[0], s_keep[1], s_keep[2]);
  __syncthreads();
  int culled = 0;
  for (int j = threadIdx.x; j < nq; j += blockDim.x) {
    const int b = best_of[j];
    if (b < 0) continue;
    const int bin = rotation_bin(q[j].angle, kps[b].angle, false);
    if (bin != s_keep[0] && bin != s_keep[1] && bin != s_keep[2]) {
      last_writer[b] = -2;  // :147 (cleared, whoever the last claimant was)
      culled++;
    }
  }
  if (culled) atomicAdd(&s_culled, culled);
  __syncthreads();
  if (threadIdx.x == 0) flags[2] -= s_culled;  // :148
}

}  // namespace ydorb

using namespace ydorb;

extern "C" {

void ydorb_three_maxima(const int32_t* sizes, int len, int32_t* idx1, int32_t* idx2, int32_t* idx3) {
  int a, b, c;
  three_maxima(sizes, len, a, b, c);
  *idx1 = a; *idx2 = b; *idx3 = c;
}

int ydorb_search_windows(YdFrame* f, const YdWindowQuery* queries, const uint8_t* q_descriptors, int nq,
                         const YdWindowParams* prm, const uint8_t* kp_occupied, int32_t* best_of, int32_t* match_out,
                         int32_t* n_matches) {
  YD_REQUIRE(f && prm && n_matches && nq >= 0, "bad argument");
  YD_REQUIRE(nq == 0 || (queries && q_descriptors), "null query arrays");
  YdMatcher* m = f->m;
  YD_CUDA(cudaSetDevice(m->device));
  cudaStream_t s = m->stream;
  const int n = f->n;
  if (match_out) for (int i = 0; i < n; i++) match_out[i] = -1;
  if (best_of) for (int j = 0; j < nq; j++) best_of[j] = -1;
  *n_matches = 0;
  if (n == 0 || nq == 0) return YDORB_OK;
  auto up256 = [](size_t v) { return (v + 255) / 256 * 256; };
  const size_t o_q = 0, o_desc = up256(sizeof(YdWindowQuery) * (size_t)nq), o_occ = o_desc + up256((size_t)32 * nq),
               o_best = o_occ + up256((size_t)n), total = o_best + sizeof(int) * (size_t)nq;
  int rc = matcher_reserve_io(m, total);
  if (rc != YDORB_OK) return rc;
  uint8_t* io = reinterpret_cast<uint8_t*>(m->d_io);
  YD_CUDA(cudaMemcpyAsync(io + o_q, queries, sizeof(YdWindowQuery) * (size_t)nq, cudaMemcpyHostToDevice, s));
  YD_CUDA(cudaMemcpyAsync(io + o_desc, q_descriptors, (size_t)32 * nq, cudaMemcpyHostToDevice, s));
  if (kp_occupied) YD_CUDA(cudaMemcpyAsync(io + o_occ, kp_occupied, (size_t)n, cudaMemcpyHostToDevice, s));
  const int gb = (n + 255) / 256;
  k_win_blocker_init<<<gb, 256, 0, s>>>(kp_occupied ? io + o_occ : nullptr, n, f->d_blocker);
  count_launches(1);
  WindowArgs A;
  A.gp = f->gp;
  A.kps = f->d_kps;
  A.desc = f->d_desc;
  A.right_x = f->d_right_x;
  A.cell_start = f->d_cell_start;
  A.items = f->d_cell_items;
  A.q = reinterpret_cast<const YdWindowQuery*>(io + o_q);
  A.q_desc = reinterpret_cast<const uint4*>(io + o_desc);
  A.nq = nq;
  A.max_dist = prm->max_dist;
  A.strict = prm->strict;
  A.exclusive = prm->exclusive;
  A.chi2 = prm->chi2_gate;
  for (int i = 0; i < kMaxLevels; i++) A.inv_sf2[i] = prm->inv_sf2[i];
  A.last_writer = f->d_last;
  A.best_of = reinterpret_cast<int*>(io + o_best);
  A.flags = f->d_flags;
  int* cur = f->d_blocker;
  int* nxt = f->d_blocker2;
  int flags[4] = {0, 0, 0, 0};
  YD_CUDA(cudaMemsetAsync(f->d_flags, 0, sizeof flags, s));
  const int batch = prm->exclusive ? 6 : 1;  // no greedy dependence: one pass is the answer
  for (int round = 0; !flags[3];) {
    for (int b = 0; b < batch; b++, round++) {
      k_win_round_init<<<gb, 256, 0, s>>>(f->d_blocker, n, nxt, f->d_last, f->d_flags, round);
      A.blocker = cur;
      A.blocker_next = nxt;
      k_window_search<<<(nq + 127) / 128, 128, 0, s>>>(A);
      count_launches(2);
      if (!prm->exclusive) break;
      k_win_blocker_diff<<<gb, 256, 0, s>>>(cur, nxt, n, f->d_flags, round);
      count_launches(1);
      std::swap(cur, nxt);
    }
    if (!prm->exclusive) break;
    // a seventh init only folds the last round's verdict into flags[3]; it resets nothing unless another round is due
    k_win_round_init<<<gb, 256, 0, s>>>(f->d_blocker, n, nxt, f->d_last, f->d_flags, round);
    count_launches(1);
    YD_CUDA(cudaGetLastError());
    YD_CUDA(cudaMemcpyAsync(flags, f->d_flags, sizeof flags, cudaMemcpyDeviceToHost, s));
    YD_CUDA(cudaStreamSynchronize(s));
    if (round > nq + 8) { set_error("window search did not converge"); return YDORB_E_STATE; }
  }
  if (prm->check_orientation) {
    k_orientation_cull<<<1, 1024, 0, s>>>(A.q, f->d_kps, A.best_of, nq, f->d_last, f->d_flags);
    count_launches(1);
  }
  YD_CUDA(cudaGetLastError());
  YD_CUDA(cudaMemcpyAsync(flags, f->d_flags, sizeof flags, cudaMemcpyDeviceToHost, s));
  if (best_of) YD_CUDA(cudaMemcpyAsync(best_of, A.best_of, sizeof(int) * (size_t)nq, cudaMemcpyDeviceToHost, s));
  if (match_out) YD_CUDA(cudaMemcpyAsync(match_out, f->d_last, sizeof(int) * (size_t)n, cudaMemcpyDeviceToHost, s));
  YD_CUDA(cudaStreamSynchronize(s));
  *n_matches = flags[2];
  return YDORB_OK;
}

}  // extern "C"
