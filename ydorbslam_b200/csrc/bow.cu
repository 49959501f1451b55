// M5 OrbMatcher::searchByBowInKeyFrameAndFrame (src/orbMatcher.cpp:303-379), searchByBowInTwoKeyFrames (:380-462)
// M6 OrbMatcher::searchForTriangulation (:463-565) with isEpipolarLineDistCorrect (:808-819)
//
// The reference merge-walks two DBoW3 FeatureVectors (std::map<NodeId, vector<idx>>) and, inside every shared node,
// runs a brute-force top-2 (or gated best-1) of each first-side feature against the second-side features that are not
// matched yet.  The "not matched yet" exclusion only couples features of the same node (a feature lives in exactly
// one node), so nodes are independent: one warp owns one shared node, walks the first-side features in list order
// (the sequential part) and spreads the second-side features over its lanes.  Per-lane running top-2 are packed keys
// (distance << 20 | position) merged by shuffles; the position in the low bits reproduces "first candidate wins
// ties" of `d < best` (:327) -- and, inverted, "last candidate wins ties" of `d <= best` in searchForTriangulation
// (:502).  The rotation-histogram check is a second, single-CTA kernel.
#include <algorithm>

#include "matcher.cuh"

namespace ydorb {

constexpr int kBowHistLen = 30;
constexpr uint32_t kBowNoKey = 0xFFFFFFFFu;
constexpr int kBowPosBits = 20;
constexpr uint32_t kBowPosMask = (1u << kBowPosBits) - 1;

__device__ __forceinline__ int bow_hamming(const uint4& a0, const uint4& a1, const uint4& b0, const uint4& b1) {
  return __popc(a0.x ^ b0.x) + __popc(a0.y ^ b0.y) + __popc(a0.z ^ b0.z) + __popc(a0.w ^ b0.w) +
         __popc(a1.x ^ b1.x) + __popc(a1.y ^ b1.y) + __popc(a1.z ^ b1.z) + __popc(a1.w ^ b1.w);
}

__device__ __forceinline__ int bow_rotation_bin(float a, float b, bool add_f32) {
  const float factor = (float)(1.0 / kBowHistLen);  // :311
  float rot = __fsub_rn(a, b);
  if (rot < 0.0f) rot = add_f32 ? __fadd_rn(rot, 360.0f) : (float)((double)rot + 360.0);  // :341 / :521
  int bin = (int)round((double)__fmul_rn(rot, factor));
  if (bin == kBowHistLen) bin = 0;
  return bin;
}

struct BowSide {
  const YdKeyPoint* kps;
  const uint4* desc;
  const uint8_t* valid;    // M5: map point present and not bad; M6: has_mp (getMapPoint != null)
  const float* right_x;    // M6 only
  const int* ids;
  const int* start;
  const int* items;
  int n_nodes;
};

struct BowArgs {
  BowSide a, b;
  int mode;        // 0 KeyFrame/Frame, 1 two key frames, 2 triangulation
  float ratio;
  volatile uint8_t* matched2;
  int* match_out;
  // M6
  float F[9];
  float ex, ey;
  float sf[kMaxLevels], sf2[kMaxLevels];
  int stereo_only;
};

// isEpipolarLineDistCorrect :808-819 (float arithmetic, one rounding per operation; squared denominator as written)
__device__ __forceinline__ bool epipolar_ok(const BowArgs& A, const YdKeyPoint& k1, const YdKeyPoint& k2) {
  const float a = __fadd_rn(__fadd_rn(__fmul_rn(A.F[0], k1.x), __fmul_rn(A.F[3], k1.y)), A.F[6]);
  const float b = __fadd_rn(__fadd_rn(__fmul_rn(A.F[1], k1.x), __fmul_rn(A.F[4], k1.y)), A.F[7]);
  const float c = __fadd_rn(__fadd_rn(__fmul_rn(A.F[2], k1.x), __fmul_rn(A.F[5], k1.y)), A.F[8]);
  const float den = __fadd_rn(__fmul_rn(a, a), __fmul_rn(b, b));
  if (!(den > 0.f)) return false;
  const float num = __fadd_rn(__fadd_rn(__fmul_rn(a, k2.x), __fmul_rn(b, k2.y)), c);
  const float lhs = __fdiv_rn(__fmul_rn(num, num), __fmul_rn(den, den));
  return (double)lhs < 3.841 * (double)A.sf2[k2.octave];
}

__global__ void __launch_bounds__(128) k_bow_nodes(BowArgs A) {
  const int warp = (blockIdx.x * blockDim.x + threadIdx.x) >> 5, lane = threadIdx.x & 31;
  if (warp >= A.a.n_nodes) return;
  // std::map lookup of the same node id on the other side (the merge walk of :317-361 visits exactly the shared ids)
  const int id = A.a.ids[warp];
  int lo = 0, hi = A.b.n_nodes;
  while (lo < hi) {
    const int mid = (lo + hi) >> 1;
    if (A.b.ids[mid] < id) lo = mid + 1; else hi = mid;
  }
  if (lo >= A.b.n_nodes || A.b.ids[lo] != id) return;
  const int s1 = A.a.start[warp], e1 = A.a.start[warp + 1];
  const int s2 = A.b.start[lo], e2 = A.b.start[lo + 1];
  for (int p = s1; p < e1; p++) {
    const int i1 = A.a.items[p];
    YdKeyPoint k1;
    bool stereo1 = false;
    if (A.mode == 2) {
      stereo1 = A.a.right_x[i1] >= 0.f;                               // :489
      if (A.a.valid[i1] || (A.stereo_only && !stereo1)) continue;     // :491
      k1 = A.a.kps[i1];
    } else if (!A.a.valid[i1]) {
      continue;                                                       // :320, :400
    }
    const uint4 q0 = A.a.desc[2 * i1], q1 = A.a.desc[2 * i1 + 1];
    uint32_t best = kBowNoKey, second = kBowNoKey;
    for (int t = s2 + lane; t < e2; t += 32) {
      const int i2 = A.b.items[t];
      if (A.matched2[i2]) continue;                                   // :325, :405, :497
      uint32_t key;
      if (A.mode == 2) {
        const bool stereo2 = A.b.right_x[i2] >= 0.f;                  // :496
        if (A.b.valid[i2] || (A.stereo_only && !stereo2)) continue;   // :498-499
        const int d = bow_hamming(q0, q1, A.b.desc[2 * i2], A.b.desc[2 * i2 + 1]);
        if (d > 50) continue;                                         // :502
        const YdKeyPoint k2 = A.b.kps[i2];
        if (!(stereo1 || stereo2)) {                                  // :502-505 epipole distance
          const double dx = (double)__fsub_rn(A.ex, k2.x), dy = (double)__fsub_rn(A.ey, k2.y);
          if (!(dx * dx + dy * dy >= (double)__fmul_rn(100.f, A.sf[k2.octave]))) continue;
        }
        if (!epipolar_ok(A, k1, k2)) continue;                        // :506
        key = ((uint32_t)d << kBowPosBits) | (kBowPosMask - (uint32_t)(t - s2));  // `d <= best`: last wins ties
      } else {
        if (A.mode == 1 && !A.b.valid[i2]) continue;                  // :406-407
        const int d = bow_hamming(q0, q1, A.b.desc[2 * i2], A.b.desc[2 * i2 + 1]);
        key = ((uint32_t)d << kBowPosBits) | (uint32_t)(t - s2);      // `d < best`: first wins ties
      }
      second = min(second, max(best, key));
      best = min(best, key);
    }
#pragma unroll
    for (int o = 16; o > 0; o >>= 1) {
      const uint32_t ob = __shfl_xor_sync(0xffffffffu, best, o), os = __shfl_xor_sync(0xffffffffu, second, o);
      second = min(min(second, os), max(best, ob));
      best = min(best, ob);
    }
    if (best != kBowNoKey) {
      const int bestDist = (int)(best >> kBowPosBits);
      const uint32_t pos = best & kBowPosMask;
      const int i2 = A.b.items[s2 + (int)(A.mode == 2 ? kBowPosMask - pos : pos)];
      bool accept;
      if (A.mode == 2) {
        accept = true;  // every key already passed d <= 50 and both gates (:513)
      } else {
        const int secondDist = second == kBowNoKey ? 256 : (int)(second >> kBowPosBits);
        accept = bestDist <= 50 && (float)bestDist < __fmul_rn(A.ratio, (float)secondDist);  // :336, :418
      }
      if (accept && lane == 0) {
        A.matched2[i2] = 1;
        if (A.mode == 0) A.match_out[i2] = i1; else A.match_out[i1] = i2;  // :337, :419, :516
      }
    }
    __syncwarp();
  }
}

// rotation histogram over the output slots, computeThreeMaxima, clearing, final count (:362-377, :445-460, :539-554)
__global__ void __launch_bounds__(1024)
k_bow_finish(const YdKeyPoint* __restrict__ kps1, const YdKeyPoint* __restrict__ kps2, int mode, int check_orientation,
             int* __restrict__ match_out, int n_out, int* __restrict__ count) {
  __shared__ int s_hist[kBowHistLen];
  __shared__ int s_keep[3];
  __shared__ int s_count;
  if (threadIdx.x < kBowHistLen) s_hist[threadIdx.x] = 0;
  if (threadIdx.x == 0) s_count = 0;
  __syncthreads();
  auto bin_of = [&](int slot, int partner) {
    const int i1 = mode == 0 ? partner : slot, i2 = mode == 0 ? slot : partner;
    return bow_rotation_bin(kps1[i1].angle, kps2[i2].angle, mode == 2);
  };
  if (check_orientation) {
    for (int i = threadIdx.x; i < n_out; i += blockDim.x) {
      const int p = match_out[i];
      if (p >= 0) atomicAdd(&s_hist[bin_of(i, p)], 1);
    }
    __syncthreads();
    if (threadIdx.x == 0) {
      int max1 = 0, max2 = 0, max3 = 0, i1 = -1, i2 = -1, i3 = -1;  // computeThreeMaxima :827-854
      for (int i = 0; i < kBowHistLen; i++) {
        const int n = s_hist[i];
        if (n > max1) { max3 = max2; max2 = max1; max1 = n; i3 = i2; i2 = i1; i1 = i; }
        else if (n > max2) { max3 = max2; max2 = n; i3 = i2; i2 = i; }
        else if (n > max3) { max3 = n; i3 = i; }
      }
      if (max2 < max1 / 10) { i3 = -1; i2 = -1; }
      else if (max3 < max1 / 10) { i3 = -1; }
      s_keep[0] = i1; s_keep[1] = i2; s_keep[2] = i3;
    }
    __syncthreads();
  }
  int kept = 0;
  for (int i = threadIdx.x; i < n_out; i += blockDim.x) {
    const int p = match_out[i];
    if (p < 0) continue;
    if (check_orientation) {
      const int bin = bin_of(i, p);
      if (bin != s_keep[0] && bin != s_keep[1] && bin != s_keep[2]) { match_out[i] = -1; continue; }
    }
    kept++;
  }
  if (kept) atomicAdd(&s_count, kept);
  __syncthreads();
  if (threadIdx.x == 0) *count = s_count;
}

struct BowHostSide {
  const YdKeyPoint* kps; const uint8_t* desc; const uint8_t* valid; const float* right_x; int n; const YdFeatureVector* fv;
};

// uploads one side into the staging buffer at *ofs (256 B aligned pieces) and fills the device view
static int bow_upload(YdMatcher* m, uint8_t* io, size_t& ofs, const BowHostSide& h, BowSide& d, bool dry) {
  auto piece = [&](const void* src, size_t bytes) -> uint8_t* {
    uint8_t* dst = io + ofs;
    ofs += (bytes + 255) / 256 * 256;
    if (!dry && src && bytes) cudaMemcpyAsync(dst, src, bytes, cudaMemcpyHostToDevice, m->stream);
    return dst;
  };
  const int items = h.fv->n_nodes > 0 ? h.fv->node_start[h.fv->n_nodes] : 0;
  d.kps = reinterpret_cast<const YdKeyPoint*>(piece(h.kps, sizeof(YdKeyPoint) * (size_t)h.n));
  d.desc = reinterpret_cast<const uint4*>(piece(h.desc, (size_t)32 * h.n));
  d.valid = piece(h.valid, (size_t)h.n);
  d.right_x = reinterpret_cast<const float*>(piece(h.right_x, h.right_x ? sizeof(float) * (size_t)h.n : 0));
  d.ids = reinterpret_cast<const int*>(piece(h.fv->node_ids, sizeof(int) * (size_t)h.fv->n_nodes));
  d.start = reinterpret_cast<const int*>(piece(h.fv->node_start, sizeof(int) * ((size_t)h.fv->n_nodes + 1)));
  d.items = reinterpret_cast<const int*>(piece(h.fv->items, sizeof(int) * (size_t)items));
  d.n_nodes = h.fv->n_nodes;
  return YDORB_OK;
}

static int bow_check_fv(const YdFeatureVector* fv, int n) {
  YD_REQUIRE(fv && fv->n_nodes >= 0, "null feature vector");
  if (fv->n_nodes == 0) return YDORB_OK;
  YD_REQUIRE(fv->node_ids && fv->node_start && fv->node_start[0] == 0, "malformed feature vector");
  for (int i = 0; i < fv->n_nodes; i++) {
    YD_REQUIRE(fv->node_start[i + 1] >= fv->node_start[i], "feature vector offsets not monotone");
    YD_REQUIRE(i == 0 || fv->node_ids[i] > fv->node_ids[i - 1], "feature vector node ids not ascending");
    YD_REQUIRE(fv->node_start[i + 1] - fv->node_start[i] <= (int)kBowPosMask, "feature vector node too large");
  }
  const int items = fv->node_start[fv->n_nodes];
  YD_REQUIRE(items == 0 || fv->items, "null feature vector items");
  for (int i = 0; i < items; i++) YD_REQUIRE(fv->items[i] >= 0 && fv->items[i] < n, "feature index out of range");
  return YDORB_OK;
}

static int bow_run(YdMatcher* m, const BowHostSide& h1, const BowHostSide& h2, BowArgs& A, int check_orientation,
                   int32_t* match_out, int32_t* n_matches) {
  int rc;
  if ((rc = bow_check_fv(h1.fv, h1.n)) != YDORB_OK || (rc = bow_check_fv(h2.fv, h2.n)) != YDORB_OK) return rc;
  const int n_out = A.mode == 0 ? h2.n : h1.n;
  for (int i = 0; i < n_out; i++) match_out[i] = -1;
  *n_matches = 0;
  if (h1.n == 0 || h2.n == 0 || h1.fv->n_nodes == 0 || h2.fv->n_nodes == 0) return YDORB_OK;
  YD_CUDA(cudaSetDevice(m->device));
  cudaStream_t s = m->stream;
  size_t ofs = 0;
  BowSide tmp;
  bow_upload(m, nullptr, ofs, h1, tmp, true);
  bow_upload(m, nullptr, ofs, h2, tmp, true);
  const size_t o_matched = ofs, o_out = o_matched + ((size_t)h2.n + 255) / 256 * 256,
               o_count = o_out + (sizeof(int) * (size_t)n_out + 255) / 256 * 256, total = o_count + 256;
  if ((rc = matcher_reserve_io(m, total)) != YDORB_OK) return rc;
  uint8_t* io = reinterpret_cast<uint8_t*>(m->d_io);
  ofs = 0;
  bow_upload(m, io, ofs, h1, A.a, false);
  bow_upload(m, io, ofs, h2, A.b, false);
  A.matched2 = io + o_matched;
  A.match_out = reinterpret_cast<int*>(io + o_out);
  YD_CUDA(cudaMemsetAsync(io + o_matched, 0, (size_t)h2.n, s));
  YD_CUDA(cudaMemsetAsync(io + o_out, 0xFF, sizeof(int) * (size_t)n_out, s));
  const int warps_per_cta = 4;
  k_bow_nodes<<<(A.a.n_nodes + warps_per_cta - 1) / warps_per_cta, 32 * warps_per_cta, 0, s>>>(A);
  k_bow_finish<<<1, 1024, 0, s>>>(A.a.kps, A.b.kps, A.mode, check_orientation, A.match_out, n_out,
                                  reinterpret_cast<int*>(io + o_count));
  count_launches(2);
  YD_CUDA(cudaGetLastError());
  YD_CUDA(cudaMemcpyAsync(match_out, io + o_out, sizeof(int) * (size_t)n_out, cudaMemcpyDeviceToHost, s));
  YD_CUDA(cudaMemcpyAsync(n_matches, io + o_count, sizeof(int), cudaMemcpyDeviceToHost, s));
  YD_CUDA(cudaStreamSynchronize(s));
  return YDORB_OK;
}

}  // namespace ydorb

using namespace ydorb;

extern "C" {

int ydorb_search_by_bow(YdMatcher* m, const YdKeyPoint* kps1, const uint8_t* desc1, const uint8_t* valid1, int n1,
                        const YdFeatureVector* fv1, const YdKeyPoint* kps2, const uint8_t* desc2, const uint8_t* valid2,
                        int n2, const YdFeatureVector* fv2, int mode, float ratio, int check_orientation,
                        int32_t* match_out, int32_t* n_matches) {
  YD_REQUIRE(m && match_out && n_matches && n1 >= 0 && n2 >= 0, "bad argument");
  YD_REQUIRE(mode == 0 || mode == 1, "mode must be 0 (key frame / frame) or 1 (two key frames)");
  YD_REQUIRE(n1 == 0 || (kps1 && desc1 && valid1), "null first-side arrays");
  YD_REQUIRE(n2 == 0 || (kps2 && desc2 && (mode == 0 || valid2)), "null second-side arrays");
  BowArgs A;
  memset(&A, 0, sizeof A);
  A.mode = mode;
  A.ratio = ratio;
  const BowHostSide h1{kps1, desc1, valid1, nullptr, n1, fv1}, h2{kps2, desc2, mode == 1 ? valid2 : nullptr, nullptr, n2, fv2};
  return bow_run(m, h1, h2, A, check_orientation, match_out, n_matches);
}

int ydorb_search_for_triangulation(YdMatcher* m, const YdKeyPoint* kps1, const uint8_t* desc1, const float* right_x1,
                                   const uint8_t* has_mp1, int n1, const YdFeatureVector* fv1, const YdKeyPoint* kps2,
                                   const uint8_t* desc2, const float* right_x2, const uint8_t* has_mp2, int n2,
                                   const YdFeatureVector* fv2, const float* F, float epipole_x, float epipole_y,
                                   const float* sf, const float* sf2, int n_levels, int stereo_only,
                                   int check_orientation, int32_t* match_out, int32_t* n_matches) {
  YD_REQUIRE(m && match_out && n_matches && n1 >= 0 && n2 >= 0 && F && sf && sf2, "bad argument");
  YD_REQUIRE(n_levels >= 1 && n_levels <= kMaxLevels, "n_levels out of range");
  YD_REQUIRE(n1 == 0 || (kps1 && desc1 && right_x1 && has_mp1), "null first-side arrays");
  YD_REQUIRE(n2 == 0 || (kps2 && desc2 && right_x2 && has_mp2), "null second-side arrays");
  YD_REQUIRE(octaves_in_range(kps1, n1, n_levels) && octaves_in_range(kps2, n2, n_levels), "keypoint octave outside [0, n_levels)");
  BowArgs A;
  memset(&A, 0, sizeof A);
  A.mode = 2;
  for (int i = 0; i < 9; i++) A.F[i] = F[i];
  A.ex = epipole_x;
  A.ey = epipole_y;
  for (int i = 0; i < kMaxLevels; i++) { A.sf[i] = sf[std::min(i, n_levels - 1)]; A.sf2[i] = sf2[std::min(i, n_levels - 1)]; }
  A.stereo_only = stereo_only;
  const BowHostSide h1{kps1, desc1, has_mp1, right_x1, n1, fv1}, h2{kps2, desc2, has_mp2, right_x2, n2, fv2};
  return bow_run(m, h1, h2, A, check_orientation, match_out, n_matches);
}

}  // extern "C"
