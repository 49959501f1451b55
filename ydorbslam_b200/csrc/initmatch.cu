// EXTRA  searchForInitialization as `ydorb_match_windowed_top2`.
//
// `north_star` names searchForInitialization; the reference has no such function (absent from
// src/orbMatcher.hpp:24-66), so there is nothing to be bit-exact against.  What is built here is the published algorithm
// the reference descends from -- ORB-SLAM2, ORBmatcher::SearchForInitialization (ORB_SLAM2 src/ORBmatcher.cc) -- on this
// reference's own pieces: Frame::getKeyPointsInArea (src/frame.cpp:337-361; area_mode 0 keeps its predicate, SURVEY
// F9, area_mode 1 uses ORB-SLAM2's |dx| < r && |dy| < r and minLevel <= octave <= maxLevel), the descriptor distance
// (src/orbMatcher.cpp:11-23), the rotation histogram convention of src/orbMatcher.cpp:120-131 and computeThreeMaxima
// (:827-854).  Parity is against oracle/ydorb_oracle.cpp orc_search_for_initialization, labelled "no reference parity".
//
// The sequential loop is greedy: query i1 skips a candidate i2 whose recorded match distance is already <= its own
// (vMatchedDistance[i2] <= dist), and that record is written by earlier queries.  As in search.cu every query is
// evaluated in parallel against the claims of the PREVIOUS round (per keypoint of frame 2 a linked list of its
// claimants: a candidate is blocked iff a claimant with a lower index holds it at a distance <= dist); rounds repeat
// until no decision changes.  By induction on the query index the fixed point is unique and equals the loop.
// The final owner of a keypoint is its highest-index claimant (a later claim un-matches the earlier owner); the
// orientation histogram counts every claim, owners that were displaced included, exactly as the loop pushes them.
#include <algorithm>

#include "frame.cuh"

namespace ydorb {

constexpr int kInitHistLen = 30;  // OrbMatcher::m_int_histLen src/orbMatcher.cpp:9

__device__ __forceinline__ int init_rotation_bin(float a, float b) {  // src/orbMatcher.cpp:121-129
  const float factor = (float)(1.0 / kInitHistLen);
  float rot = __fsub_rn(a, b);
  if (rot < 0.0f) rot = (float)((double)rot + 360.0);
  int bin = (int)round((double)__fmul_rn(rot, factor));
  if (bin == kInitHistLen) bin = 0;
  return bin;
}

struct InitArgs {
  GridParams gp;
  const YdKeyPoint* kps2;
  const uint4* desc2;
  const int* cell_start;
  const int* items;
  const YdKeyPoint* kps1;
  const uint4* desc1;
  const float2* prev;
  int n1, n2;
  float r, ratio;
  int th_low, area_mode;
  const int* match_prev;  // decisions of the previous round
  const int* dist_prev;
  const int* head;        // per keypoint of frame 2: first claimant of the previous round, -1 = none
  const int* next;        // per query: next claimant of the same keypoint
  int* match_cur;
  int* dist_cur;
  int* flags;             // [0]/[1] round changed (even / odd), [2] match count, [3] fixed point reached
  int round;
};

__global__ void __launch_bounds__(128) k_init_round(InitArgs A) {
  const int i1 = blockIdx.x * blockDim.x + threadIdx.x;
  if (i1 >= A.n1 || A.flags[3]) return;
  const YdKeyPoint k1 = A.kps1[i1];
  int best_idx = -1, best = 0x7FFFFFFF, best2 = 0x7FFFFFFF;
  if (k1.octave <= 0) {  // level1 > 0: continue
    const float2 pm = A.prev[i1];
    const uint4 q0 = A.desc1[2 * i1], q1 = A.desc1[2 * i1 + 1];
    CellRange cr;
    if (cell_range(A.gp, pm.x, pm.y, A.r, cr)) {
      for (int ix = cr.x0; ix <= cr.x1; ix++)
        for (int iy = cr.y0; iy <= cr.y1; iy++) {
          const int c = ix * kGridRows + iy;
          const int e = A.cell_start[c + 1];
          for (int t = A.cell_start[c]; t < e; t++) {
            const int i2 = A.items[t];
            const YdKeyPoint kp = A.kps2[i2];
            if (A.area_mode == 0) {
              if (!in_window(kp, pm.x, pm.y, A.r, k1.octave, k1.octave)) continue;
            } else {
              if (kp.octave != k1.octave) continue;
              if (!(fabsf(__fsub_rn(kp.x, pm.x)) < A.r && fabsf(__fsub_rn(kp.y, pm.y)) < A.r)) continue;
            }
            const int d = hamming256(q0, q1, A.desc2[2 * i2], A.desc2[2 * i2 + 1]);
            bool blocked = false;  // vMatchedDistance[i2] <= dist, as earlier queries left it
            for (int j = A.head[i2]; j >= 0; j = A.next[j])
              if (j < i1 && A.dist_prev[j] <= d) { blocked = true; break; }
            if (blocked) continue;
            if (d < best) { best2 = best; best = d; best_idx = i2; }
            else if (d < best2) best2 = d;
          }
        }
    }
    if (!(best_idx >= 0 && best <= A.th_low && (float)best < __fmul_rn((float)best2, A.ratio))) best_idx = -1;
  }
  if (best_idx < 0) best = 0x7FFFFFFF;
  A.match_cur[i1] = best_idx;
  A.dist_cur[i1] = best;
  if (A.round == 0 || A.match_prev[i1] != best_idx || A.dist_prev[i1] != best) A.flags[A.round & 1] = 1;
}

// between rounds: fold the verdict of round-1 into flags[3], reset this round's flag and the claim lists
__global__ void k_init_reset(int* __restrict__ head, int n2, int* __restrict__ flags, int round) {
  const int i = blockIdx.x * blockDim.x + threadIdx.x;
  const bool done = flags[3] || (round > 0 && flags[(round - 1) & 1] == 0);
  if (i == 0) {
    if (done) flags[3] = 1;
    else flags[round & 1] = 0;
  }
  if (done) return;
  if (i < n2) head[i] = -1;
}

// claim lists of the decisions in `match`
__global__ void k_init_link(const int* __restrict__ match, int n1, int* __restrict__ head, int* __restrict__ next,
                            const int* __restrict__ flags) {
  const int i1 = blockIdx.x * blockDim.x + threadIdx.x;
  if (i1 >= n1 || flags[3]) return;
  const int m = match[i1];
  if (m >= 0) next[i1] = atomicExch(&head[m], i1);
}

// owner of a keypoint = its highest claimant; histogram over every claim; cull; write-back.  One CTA.
__global__ void __launch_bounds__(1024)
k_init_final(const int* __restrict__ match, const YdKeyPoint* __restrict__ kps1, const YdKeyPoint* __restrict__ kps2, int n1,
             int n2, int* __restrict__ owner, int check_orientation, int* __restrict__ matches12, float2* __restrict__ prev,
             int* __restrict__ flags) {
  __shared__ int s_hist[kInitHistLen];
  __shared__ int s_keep[3];
  __shared__ int s_count;
  if (threadIdx.x < kInitHistLen) s_hist[threadIdx.x] = 0;
  if (threadIdx.x == 0) s_count = 0;
  for (int i = threadIdx.x; i < n2; i += blockDim.x) owner[i] = -1;
  __syncthreads();
  for (int i1 = threadIdx.x; i1 < n1; i1 += blockDim.x) {
    const int m = match[i1];
    if (m >= 0) {
      atomicMax(&owner[m], i1);
      if (check_orientation) atomicAdd(&s_hist[init_rotation_bin(kps1[i1].angle, kps2[m].angle)], 1);
    }
  }
  __syncthreads();
  if (threadIdx.x == 0) {
    // computeThreeMaxima, src/orbMatcher.cpp:827-854
    int max1 = 0, max2 = 0, max3 = 0, a = -1, b = -1, c = -1;
    for (int i = 0; i < kInitHistLen; i++) {
      const int n = s_hist[i];
      if (n > max1) { max3 = max2; max2 = max1; max1 = n; c = b; b = a; a = i; }
      else if (n > max2) { max3 = max2; max2 = n; c = b; b = i; }
      else if (n > max3) { max3 = n; c = i; }
    }
    if (max2 < max1 / 10) { c = -1; b = -1; }
    else if (max3 < max1 / 10) { c = -1; }
    s_keep[0] = a; s_keep[1] = b; s_keep[2] = c;
  }
  __syncthreads();
  int count = 0;
  for (int i1 = threadIdx.x; i1 < n1; i1 += blockDim.x) {
    int m = match[i1];
    if (m >= 0 && owner[m] != i1) m = -1;  // un-matched by a later claim of the same keypoint
    if (m >= 0 && check_orientation) {
      const int bin = init_rotation_bin(kps1[i1].angle, kps2[m].angle);
      if (bin != s_keep[0] && bin != s_keep[1] && bin != s_keep[2]) m = -1;
    }
    matches12[i1] = m;
    if (m >= 0) {
      count++;
      prev[i1] = make_float2(kps2[m].x, kps2[m].y);
    }
  }
  if (count) atomicAdd(&s_count, count);
  __syncthreads();
  if (threadIdx.x == 0) flags[2] = s_count;
}

}  // namespace ydorb

using namespace ydorb;

extern "C" int ydorb_match_windowed_top2(YdFrame* f2, const YdKeyPoint* kps1, const uint8_t* desc1, int n1,
                                         float* prev_matched, int window_size, float nn_ratio, int th_low,
                                         int check_orientation, int area_mode, int32_t* matches12, int32_t* n_matches) {
  YD_REQUIRE(f2 && n_matches && n1 >= 0, "bad argument");
  YD_REQUIRE(n1 == 0 || (kps1 && desc1 && prev_matched && matches12), "null query arrays");
  YD_REQUIRE(area_mode == 0 || area_mode == 1, "area_mode must be 0 (reference predicate) or 1 (ORB-SLAM2 predicate)");
  YdMatcher* m = f2->m;
  YD_CUDA(cudaSetDevice(m->device));
  cudaStream_t s = m->stream;
  const int n2 = f2->n;
  *n_matches = 0;
  for (int i = 0; i < n1; i++) matches12[i] = -1;
  if (n1 == 0 || n2 == 0) return YDORB_OK;
  auto up256 = [](size_t v) { return (v + 255) / 256 * 256; };
  const size_t o_kp = 0, o_desc = up256(sizeof(YdKeyPoint) * (size_t)n1), o_prev = o_desc + up256((size_t)32 * n1),
               o_m0 = o_prev + up256(sizeof(float2) * (size_t)n1), o_m1 = o_m0 + up256(sizeof(int) * (size_t)n1),
               o_d0 = o_m1 + up256(sizeof(int) * (size_t)n1), o_d1 = o_d0 + up256(sizeof(int) * (size_t)n1),
               o_next = o_d1 + up256(sizeof(int) * (size_t)n1), o_out = o_next + up256(sizeof(int) * (size_t)n1),
               total = o_out + up256(sizeof(int) * (size_t)n1);
  int rc = matcher_reserve_io(m, total);
  if (rc != YDORB_OK) return rc;
  uint8_t* io = reinterpret_cast<uint8_t*>(m->d_io);
  YD_CUDA(cudaMemcpyAsync(io + o_kp, kps1, sizeof(YdKeyPoint) * (size_t)n1, cudaMemcpyHostToDevice, s));
  YD_CUDA(cudaMemcpyAsync(io + o_desc, desc1, (size_t)32 * n1, cudaMemcpyHostToDevice, s));
  YD_CUDA(cudaMemcpyAsync(io + o_prev, prev_matched, sizeof(float2) * (size_t)n1, cudaMemcpyHostToDevice, s));
  YD_CUDA(cudaMemsetAsync(f2->d_flags, 0, 4 * sizeof(int), s));
  InitArgs A;
  A.gp = f2->gp;
  A.kps2 = f2->d_kps;
  A.desc2 = f2->d_desc;
  A.cell_start = f2->d_cell_start;
  A.items = f2->d_cell_items;
  A.kps1 = reinterpret_cast<const YdKeyPoint*>(io + o_kp);
  A.desc1 = reinterpret_cast<const uint4*>(io + o_desc);
  A.prev = reinterpret_cast<const float2*>(io + o_prev);
  A.n1 = n1;
  A.n2 = n2;
  A.r = (float)window_size;
  A.ratio = nn_ratio;
  A.th_low = th_low;
  A.area_mode = area_mode;
  A.head = f2->d_blocker;  // [n2] scratch of the frame handle
  A.next = reinterpret_cast<int*>(io + o_next);
  A.flags = f2->d_flags;
  int* mbuf[2] = {reinterpret_cast<int*>(io + o_m0), reinterpret_cast<int*>(io + o_m1)};
  int* dbuf[2] = {reinterpret_cast<int*>(io + o_d0), reinterpret_cast<int*>(io + o_d1)};
  const int g1 = (n1 + 127) / 128, g2 = (n2 + 255) / 256;
  int flags[4] = {0, 0, 0, 0};
  int round = 0;
  while (!flags[3]) {
    for (int b = 0; b < 4; b++, round++) {  // rounds are launched ahead of the host; finished ones return at once
      k_init_reset<<<g2, 256, 0, s>>>(f2->d_blocker, n2, f2->d_flags, round);
      if (round > 0) k_init_link<<<g1, 128, 0, s>>>(mbuf[(round - 1) & 1], n1, f2->d_blocker, reinterpret_cast<int*>(io + o_next), f2->d_flags);
      A.match_prev = mbuf[(round - 1) & 1];
      A.dist_prev = dbuf[(round - 1) & 1];
      A.match_cur = mbuf[round & 1];
      A.dist_cur = dbuf[round & 1];
      A.round = round;
      k_init_round<<<g1, 128, 0, s>>>(A);
      count_launches(round > 0 ? 3 : 2);
    }
    k_init_reset<<<g2, 256, 0, s>>>(f2->d_blocker, n2, f2->d_flags, round);  // folds the last verdict into flags[3]
    count_launches(1);
    YD_CUDA(cudaGetLastError());
    YD_CUDA(cudaMemcpyAsync(flags, f2->d_flags, sizeof flags, cudaMemcpyDeviceToHost, s));
    YD_CUDA(cudaStreamSynchronize(s));
    if (round > n1 + 8) { set_error("windowed top-2 did not converge"); return YDORB_E_STATE; }
  }
  // Which buffer holds the fixed point?  Once a round changes nothing, both buffers agree; rounds after flags[3] was
  // set did not run.  The last buffer written is the one of the last round that ran, and it equals the other.
  k_init_final<<<1, 1024, 0, s>>>(mbuf[0], A.kps1, f2->d_kps, n1, n2, f2->d_blocker, check_orientation,
                                  reinterpret_cast<int*>(io + o_out), reinterpret_cast<float2*>(io + o_prev), f2->d_flags);
  count_launches(1);
  YD_CUDA(cudaGetLastError());
  YD_CUDA(cudaMemcpyAsync(flags, f2->d_flags, sizeof flags, cudaMemcpyDeviceToHost, s));
  YD_CUDA(cudaMemcpyAsync(matches12, io + o_out, sizeof(int) * (size_t)n1, cudaMemcpyDeviceToHost, s));
  YD_CUDA(cudaMemcpyAsync(prev_matched, io + o_prev, sizeof(float2) * (size_t)n1, cudaMemcpyDeviceToHost, s));
  YD_CUDA(cudaStreamSynchronize(s));
  *n_matches = flags[2];
  return YDORB_OK;
}
