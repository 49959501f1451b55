// S1  Frame::computeStereoMatches  (src/frame.cpp:362-472)
//
// Reference behaviour kept in mode 0: the descriptor row and the output slot of a left keypoint are
// `leftIdx`, a counter that is NOT advanced when one of the four `continue`s fires, so it lags the keypoint index;
// the SAD minimum is tracked in an `int bestDist = 256` that shadows the Hamming one, so only SADs < 256 move it;
// the final outlier loop only ever fires when the median SAD is 0 (then it invalidates every match).
//
// Everything that does not depend on the lag is computed in parallel first:
//   A. per left keypoint k: the ordered list of right keypoints passing the row-band / octave / x-range filters
//      (:373-379, :393-394) and whether the row table entry is non-empty at all (:388);
//   B. per (k, candidate) pair: the whole SAD refinement (:405-459) -> {state, rightX, depth, SAD};
//   C. the lag chain: one warp walks k = 0..n-1, evaluates the Hamming arg-min over the candidate list with the
//      descriptor row the lag dictates (lanes over candidates, shuffle arg-min, first minimum wins) and looks the
//      pair result up.  In mode 1 (corrected: row = slot = k) C is parallel over k as well.
#include <algorithm>
#include <cmath>
#include <cstdlib>

#include "matcher.cuh"

namespace ydorb {

struct StereoGeom {
  int n_levels, rows;
  int lw[kMaxLevels], lh[kMaxLevels], stride[kMaxLevels];
  int64_t origin[kMaxLevels];
  float sf[kMaxLevels], inv_sf[kMaxLevels];
};

struct PairResult {
  float right_x, depth;
  int sad;
  int state;  // 0: `continue` (leftIdx not advanced), 1: advanced, nothing stored, 2: advanced, values stored
};

__device__ __forceinline__ bool stereo_candidate(const YdKeyPoint& kl, const YdKeyPoint& kr, const StereoGeom& g,
                                                 float max_d, bool* covers_row) {
  // row band of the right keypoint (:376): iy from max(floor(y-2sf),0) while iy <= min(ceil(y+2sf), rows-1)
  const float rad = __fmul_rn(2.0f, g.sf[kr.octave]);
  const float lo = fmaxf((float)floor((double)__fsub_rn(kr.y, rad)), 0.0f);
  const float hi = fminf((float)ceil((double)__fadd_rn(kr.y, rad)), __fsub_rn((float)g.rows, 1.0f));
  const int row = (int)kl.y;  // vRowIndices[keyPoint.pt.y] (:388)
  const bool cov = (int)lo <= row && (float)row <= hi;
  *covers_row = cov;
  if (!cov) return false;
  return kr.octave >= kl.octave - 1 && kr.octave <= kl.octave + 1 && kr.x >= __fsub_rn(kl.x, max_d) &&
         kr.x <= __fsub_rn(kl.x, 0.0f);  // :393-394
}

// A1/A2: warp per left keypoint; pass 0 counts, pass 1 writes the ordered candidate list
__global__ void __launch_bounds__(128) k_stereo_candidates(StereoGeom g, const YdKeyPoint* __restrict__ kl, int nl,
                                                           const YdKeyPoint* __restrict__ kr, int nr, float max_d,
                                                           int* __restrict__ cand_count, uint8_t* __restrict__ row_nonempty,
                                                           const int* __restrict__ cand_start, int* __restrict__ cand,
                                                           int* __restrict__ pair_left, int pass) {
  const int k = (blockIdx.x * blockDim.x + threadIdx.x) >> 5, lane = threadIdx.x & 31;
  if (k >= nl) return;
  const YdKeyPoint kp = kl[k];
  int pos = pass ? cand_start[k] : 0;
  bool any_row = false;
  for (int base = 0; base < nr; base += 32) {
    const int r = base + lane;
    bool ok = false, cov = false;
    if (r < nr) ok = stereo_candidate(kp, kr[r], g, max_d, &cov);
    const uint32_t bal = __ballot_sync(0xffffffffu, ok);
    any_row |= __ballot_sync(0xffffffffu, cov) != 0;
    if (pass && ok) {
      const int slot = pos + __popc(bal & ((1u << lane) - 1u));
      cand[slot] = r;
      pair_left[slot] = k;
    }
    pos += __popc(bal);
  }
  if (!pass && lane == 0) {
    cand_count[k] = pos;
    row_nonempty[k] = (any_row && kp.x >= 0.0f) ? 1 : 0;  // :388
  }
}

// exclusive scan, single CTA
__global__ void __launch_bounds__(1024) k_scan_single(const int* __restrict__ in, int n, int* __restrict__ out) {
  __shared__ int s[1024];
  __shared__ int carry;
  const int tid = threadIdx.x;
  if (tid == 0) carry = 0;
  __syncthreads();
  for (int base = 0; base < n; base += 1024) {
    const int i = base + tid;
    const int v = i < n ? in[i] : 0;
    s[tid] = v;
    __syncthreads();
    for (int o = 1; o < 1024; o <<= 1) {
      const int t = tid >= o ? s[tid - o] : 0;
      __syncthreads();
      s[tid] += t;
      __syncthreads();
    }
    const int c = carry;
    if (i < n) out[i] = c + s[tid] - v;
    __syncthreads();
    if (tid == 1023) carry = c + s[1023];
    __syncthreads();
  }
  if (tid == 0) out[n] = carry;
}

// B: SAD refinement of one (left keypoint, right candidate) pair, :405-459
__global__ void __launch_bounds__(128) k_stereo_sad(StereoGeom g, const uint8_t* __restrict__ pyr_l, const uint8_t* __restrict__ pyr_r,
                                                    const YdKeyPoint* __restrict__ kl, const YdKeyPoint* __restrict__ kr,
                                                    const int* __restrict__ cand, const int* __restrict__ pair_left, int npairs,
                                                    float bf, float max_d, PairResult* __restrict__ res) {
  const int p = blockIdx.x * blockDim.x + threadIdx.x;
  if (p >= npairs) return;
  const YdKeyPoint kp = kl[pair_left[p]];
  const YdKeyPoint rp = kr[cand[p]];
  PairResult out;
  out.right_x = -1.f; out.depth = -1.f; out.sad = 0; out.state = 0;
  const int o = kp.octave;
  const float lsx = (float)round((double)__fmul_rn(kp.x, g.inv_sf[o]));  // :407
  const float lsy = (float)round((double)__fmul_rn(kp.y, g.inv_sf[o]));  // :408
  const float rsx = (float)round((double)__fmul_rn(rp.x, g.inv_sf[o]));  // :409
  const int W = 5, S = 5;
  const int lw = g.lw[o], lh = g.lh[o], st = g.stride[o];
  if ((lsy - W) < 0 || (lsy + W + 1) >= lh || (lsx - W) < 0 || (lsx + W + 1) >= lw) { res[p] = out; return; }  // :412-414
  if ((rsx + S - W) < 0 || (rsx + S + W + 1) >= lw) { res[p] = out; return; }                                  // :423-425
  const uint8_t* L = pyr_l + g.origin[o];
  const uint8_t* R = pyr_r + g.origin[o];
  const int ly0 = (int)(lsy - W), lx0 = (int)(lsx - W);
  const int lc = L[(ptrdiff_t)(ly0 + W) * st + lx0 + W];
  int bestSad = 256, bestCol = 0;  // :418-419 (the shadowing int)
  float dists[11];
#pragma unroll 1
  for (int i = -S; i <= S; i++) {
    const int rx0 = (int)(rsx + (float)i - (float)W);  // >= -10: inside the 19 px pad
    const int rc = R[(ptrdiff_t)(ly0 + W) * st + rx0 + W];
    int acc = 0;
    for (int yy = 0; yy < 11; yy++) {
      const uint8_t* lrow = L + (ptrdiff_t)(ly0 + yy) * st + lx0;
      const uint8_t* rrow = R + (ptrdiff_t)(ly0 + yy) * st + rx0;
#pragma unroll
      for (int xx = 0; xx < 11; xx++) acc += abs(((int)lrow[xx] - lc) - ((int)rrow[xx] - rc));
    }
    const float dist = (float)acc;  // cv::norm(L1) of integer-valued float patches: exact
    if (dist < (float)bestSad) { bestSad = (int)dist; bestCol = i; }  // :431-434
    dists[i + S] = dist;
  }
  if (bestCol == -S || bestCol == S) { res[p] = out; return; }  // :437-439
  float d1 = 0.f, d2 = 0.f, d3 = 0.f;
#pragma unroll
  for (int t = 1; t < 10; t++)
    if (t == S + bestCol) { d1 = dists[t - 1]; d2 = dists[t]; d3 = dists[t + 1]; }
  // (d1-d3)/(2.0*(d1+d3-2.0*d2)) : float sub/add, then double (:441-443)
  const double den = __dmul_rn(2.0, __dsub_rn((double)__fadd_rn(d1, d3), __dmul_rn(2.0, (double)d2)));
  const float delta = (float)__ddiv_rn((double)__fsub_rn(d1, d3), den);
  if (delta < -1 || delta > 1) { res[p] = out; return; }  // :444-446 (NaN passes, as in the reference)
  float bestRightX = __fmul_rn(g.sf[o], __fadd_rn(__fadd_rn(rsx, delta), (float)bestCol));  // :448
  float disparity = __fsub_rn(kp.x, bestRightX);
  out.state = 1;
  if (disparity >= 0.0f && disparity < max_d) {  // :450
    if (disparity <= 0) {
      disparity = 0.01f;
      bestRightX = (float)__dsub_rn((double)kp.x, 0.01);
    }
    out.depth = __fdiv_rn(bf, disparity);  // :455
    out.right_x = bestRightX;
    out.sad = bestSad;
    out.state = 2;
  }
  res[p] = out;
}

struct ChainArgs {
  const uint4* dl;
  const uint4* dr;
  const int* cand_start;
  const int* cand;
  const uint8_t* row_nonempty;
  const PairResult* res;
  int nl;
  float* right_x;
  float* depth;
  int* rec_sad;     // [nl] SAD recorded at a slot, -1 = none
  int* counters;    // [0] recorded, [1] recorded with SAD 0, [2] lag window overflowed, [3] final lag
};

// Hamming arg-min of left descriptor row `row` over the candidate list of keypoint k (warp-cooperative)
__device__ __forceinline__ void stereo_best(const ChainArgs& A, int k, int row, int lane, int* best_dist, int* best_pos) {
  const uint4 a0 = A.dl[2 * row], a1 = A.dl[2 * row + 1];
  const int b = A.cand_start[k], e = A.cand_start[k + 1];
  uint32_t key = 0xFFFFFFFFu;  // dist << 22 | position : minimum = smallest distance, first position
  for (int t = b + lane; t < e; t += 32) {
    const int r = A.cand[t];
    const uint4 b0 = A.dr[2 * r], b1 = A.dr[2 * r + 1];
    const int d = __popc(a0.x ^ b0.x) + __popc(a0.y ^ b0.y) + __popc(a0.z ^ b0.z) + __popc(a0.w ^ b0.w) +
                  __popc(a1.x ^ b1.x) + __popc(a1.y ^ b1.y) + __popc(a1.z ^ b1.z) + __popc(a1.w ^ b1.w);
    key = min(key, ((uint32_t)d << 22) | (uint32_t)(t - b));
  }
#pragma unroll
  for (int o = 16; o > 0; o >>= 1) key = min(key, __shfl_xor_sync(0xffffffffu, key, o));
  if (key == 0xFFFFFFFFu) { *best_dist = 256; *best_pos = -1; }
  else { *best_dist = (int)(key >> 22); *best_pos = (int)(key & 0x3FFFFFu); }
}

__device__ __forceinline__ int stereo_step(const ChainArgs& A, int k, int slot, int lane) {
  // returns 1 if leftIdx advances
  if (!A.row_nonempty[k]) return 0;  // :388
  int bd, bp;
  stereo_best(A, k, slot, lane, &bd, &bp);
  if (bd < 75) {  // orbDistThd (:365, :405)
    const PairResult pr = A.res[A.cand_start[k] + bp];
    if (pr.state == 0) return 0;
    if (pr.state == 2 && lane == 0) {
      A.depth[slot] = pr.depth;  // :455-457
      A.right_x[slot] = pr.right_x;
      A.rec_sad[slot] = pr.sad;
      atomicAdd(&A.counters[0], 1);
      if (pr.sad == 0) atomicAdd(&A.counters[1], 1);
    }
  }
  return 1;
}

// C (reference mode), general form: one warp walks the lag chain.  Only runs when the windowed form below overflowed.
__global__ void __launch_bounds__(32) k_stereo_chain(ChainArgs A) {
  if (!A.counters[2]) return;
  const int lane = threadIdx.x;
  int left_idx = 0;
  for (int k = 0; k < A.nl; k++) left_idx += stereo_step(A, k, left_idx, lane);
}
// C (corrected mode, lag == nullptr): slot = k.  Reference mode: slot = k - lag[k] from the lag walk.
__global__ void __launch_bounds__(128) k_stereo_direct(ChainArgs A, const int* __restrict__ lag) {
  const int k = (blockIdx.x * blockDim.x + threadIdx.x) >> 5, lane = threadIdx.x & 31;
  if (k >= A.nl) return;
  if (lag && A.counters[2]) return;  // window overflowed: k_stereo_chain does the work
  stereo_step(A, k, lag ? k - lag[k] : k, lane);
}

// C (reference mode), windowed form.  lag(k) = k - leftIdx(k) only grows when a step does not advance, and once it
// is > 0 the descriptor row belongs to another keypoint, so it stays small in practice.  Whether step k advances is
// a function of (k, lag) alone: tabulate it for lag < 64 with the whole GPU (bit `lag` of adv[k]), then resolve the
// chain as a walk over those bits.
constexpr int kLagWindow = 64;
constexpr int kLagBlock = 32;      // keypoints per walk block
constexpr int kLagMaxBlocks = 1024;  // 64 KB of shared memory: n_left <= 32768, else the general form runs

__global__ void __launch_bounds__(128) k_stereo_lag_table(ChainArgs A, uint32_t* __restrict__ adv) {
  // warp per (keypoint, half of the window), lane per lag
  const int wid = (blockIdx.x * blockDim.x + threadIdx.x) >> 5, lane = threadIdx.x & 31;
  const int k = wid >> 1, h = wid & 1;
  if (k >= A.nl) return;
  if (!A.row_nonempty[k]) {  // :388, never advances
    if (lane == 0) adv[wid] = 0u;
    return;
  }
  const int b = A.cand_start[k], e = A.cand_start[k + 1];
  const int row = k - (lane + 32 * h);
  bool bit = true;  // lags beyond k are never reached
  if (row >= 0) {
    const uint4 a0 = A.dl[2 * row], a1 = A.dl[2 * row + 1];
    uint32_t key = 0xFFFFFFFFu;
#pragma unroll 4
    for (int t = b; t < e; t++) {  // every lane walks the same list: broadcast loads
      const int r = A.cand[t];
      const uint4 b0 = A.dr[2 * r], b1 = A.dr[2 * r + 1];
      const int d = __popc(a0.x ^ b0.x) + __popc(a0.y ^ b0.y) + __popc(a0.z ^ b0.z) + __popc(a0.w ^ b0.w) +
                    __popc(a1.x ^ b1.x) + __popc(a1.y ^ b1.y) + __popc(a1.z ^ b1.z) + __popc(a1.w ^ b1.w);
      key = min(key, ((uint32_t)d << 22) | (uint32_t)(t - b));
    }
    if (key != 0xFFFFFFFFu && (key >> 22) < 75u) bit = A.res[b + (int)(key & 0x3FFFFFu)].state != 0;
  }
  const uint32_t word = __ballot_sync(0xffffffffu, bit);
  if (lane == 0) adv[wid] = word;  // little-endian halves of the 64-bit row
}

__device__ __forceinline__ int lag_walk(const unsigned long long* __restrict__ adv, int k0, int k1, int lag, int* out) {
  // lag == kLagWindow is the sticky overflow state
#pragma unroll 4
  for (int k = k0; k < k1; k++) {
    const unsigned long long bits = adv[k];
    if (out) out[k] = lag;
    if (lag < kLagWindow) lag += 1 - (int)((bits >> lag) & 1ull);
  }
  return lag;
}

// single CTA: (1) per block of 32 keypoints and per entry lag, the exit lag; (2) one thread chains the blocks;
// (3) per block, the walk from its now known entry lag writes lag[k].  Serial depth 32 + nblk + 32 instead of n.
__global__ void __launch_bounds__(1024) k_stereo_lag_walk(const unsigned long long* __restrict__ adv, int nl,
                                                          int* __restrict__ lag, int* __restrict__ counters) {
  extern __shared__ uint8_t s_exit[];  // [nblk][64]
  __shared__ int s_entry[kLagMaxBlocks];
  const int nblk = (nl + kLagBlock - 1) / kLagBlock;
  for (int w = threadIdx.x; w < nblk * kLagWindow; w += blockDim.x) {
    const int blk = w >> 6, l0 = w & 63, k0 = blk * kLagBlock;
    s_exit[w] = (uint8_t)lag_walk(adv, k0, min(k0 + kLagBlock, nl), l0, nullptr);
  }
  __syncthreads();
  if (threadIdx.x == 0) {
    int l = 0;
    for (int blk = 0; blk < nblk; blk++) {
      s_entry[blk] = l;
      if (l < kLagWindow) l = s_exit[blk * kLagWindow + l];
    }
    counters[2] = l >= kLagWindow;
    counters[3] = l;
  }
  __syncthreads();
  for (int blk = threadIdx.x; blk < nblk; blk += blockDim.x) {
    const int k0 = blk * kLagBlock;
    lag_walk(adv, k0, min(k0 + kLagBlock, nl), s_entry[blk], lag);
  }
}

__global__ void k_stereo_init(int nl, float* __restrict__ right_x, float* __restrict__ depth, int* __restrict__ rec_sad,
                              int* __restrict__ counters) {
  const int i = blockIdx.x * blockDim.x + threadIdx.x;
  if (i < nl) { right_x[i] = -1.0f; depth[i] = -1.0f; rec_sad[i] = -1; }  // :363-364
  if (i == 0) { counters[0] = 0; counters[1] = 0; counters[2] = 0; counters[3] = 0; }
}

__global__ void k_stereo_force_chain(int* counters) { counters[2] = 1; }

// :463-471 -- the ascending loop breaks at once unless the median SAD is 0, in which case every match goes to -2
__global__ void k_stereo_outliers(int nl, const int* __restrict__ rec_sad, const int* __restrict__ counters,
                                  float* __restrict__ right_x, float* __restrict__ depth) {
  const int i = blockIdx.x * blockDim.x + threadIdx.x;
  if (i >= nl) return;
  const int size = counters[0], zeros = counters[1];
  if (size > 0 && zeros >= size / 2 + 1 && rec_sad[i] >= 0) { right_x[i] = -2.0f; depth[i] = -2.0f; }
}

}  // namespace ydorb

using namespace ydorb;

extern "C" int ydorb_stereo_match(YdMatcher* m, const YdExtractor* left, int left_frame, const YdExtractor* right,
                                  int right_frame, const YdKeyPoint* kp_left, const uint8_t* desc_left, int n_left,
                                  const YdKeyPoint* kp_right, const uint8_t* desc_right, int n_right, float bf,
                                  float max_d, int mode, float* right_x, float* depth) {
  YD_REQUIRE(m && left && right && right_x && depth, "null argument");
  YD_REQUIRE(n_left >= 0 && n_right >= 0, "negative size");
  YD_REQUIRE(mode == 0 || mode == 1, "mode must be 0 (reference) or 1 (corrected)");
  if (left->cur_w <= 0 || right->cur_w <= 0) { set_error("extract must run on both handles first"); return YDORB_E_STATE; }
  YD_REQUIRE(left->cur_w == right->cur_w && left->cur_h == right->cur_h && left->geom.n_levels == right->geom.n_levels,
             "left/right pyramids differ in geometry");
  YD_REQUIRE(left_frame >= 0 && left_frame < left->cur_batch && right_frame >= 0 && right_frame < right->cur_batch,
             "frame index out of range");
  YD_REQUIRE(left->device == m->device && right->device == m->device, "handles live on different devices");
  for (int i = 0; i < n_left; i++) { right_x[i] = -1.0f; depth[i] = -1.0f; }
  if (n_left == 0) return YDORB_OK;
  YD_REQUIRE(kp_left && desc_left && (n_right == 0 || (kp_right && desc_right)), "null keypoints/descriptors");
  YD_REQUIRE(octaves_in_range(kp_left, n_left, left->geom.n_levels) && octaves_in_range(kp_right, n_right, left->geom.n_levels),
             "keypoint octave outside [0, n_levels)");
  YD_CUDA(cudaSetDevice(m->device));
  // the pyramids were produced on the extractors' streams
  YD_CUDA(cudaStreamSynchronize(left->stream));
  YD_CUDA(cudaStreamSynchronize(right->stream));
  cudaStream_t s = m->stream;
  StereoGeom g;
  memset(&g, 0, sizeof g);
  g.n_levels = left->geom.n_levels;
  g.rows = left->geom.lv[0].h;  // :366
  for (int l = 0; l < g.n_levels; l++) {
    const LevelGeom& lv = left->geom.lv[l];
    g.lw[l] = lv.w; g.lh[l] = lv.h; g.stride[l] = lv.stride; g.origin[l] = level_origin(lv);
    g.sf[l] = left->sf[l]; g.inv_sf[l] = left->inv_sf[l];
  }
  const uint8_t* pyr_l = left->d_pyr + (size_t)left_frame * left->geom.pyr_bytes;
  const uint8_t* pyr_r = right->d_pyr + (size_t)right_frame * right->geom.pyr_bytes;
  const int nl = n_left, nr = n_right;
  auto al = [](size_t v) { return (v + 255) / 256 * 256; };
  // io layout
  const size_t o_kl = 0, o_kr = o_kl + al(sizeof(YdKeyPoint) * (size_t)nl), o_dl = o_kr + al(sizeof(YdKeyPoint) * (size_t)std::max(nr, 1)),
               o_dr = o_dl + al((size_t)32 * nl), o_rx = o_dr + al((size_t)32 * std::max(nr, 1)), o_dp = o_rx + al(sizeof(float) * (size_t)nl),
               o_cnt = o_dp + al(sizeof(float) * (size_t)nl), o_start = o_cnt + al(sizeof(int) * (size_t)nl),
               o_row = o_start + al(sizeof(int) * ((size_t)nl + 1)), o_rec = o_row + al((size_t)nl),
               o_ctr = o_rec + al(sizeof(int) * (size_t)nl), o_adv = o_ctr + 256,
               o_lag = o_adv + al(sizeof(unsigned long long) * (size_t)nl), io_total = o_lag + al(sizeof(int) * (size_t)nl);
  int rc = matcher_reserve_io(m, io_total);
  if (rc != YDORB_OK) return rc;
  uint8_t* io = reinterpret_cast<uint8_t*>(m->d_io);
  YdKeyPoint* d_kl = reinterpret_cast<YdKeyPoint*>(io + o_kl);
  YdKeyPoint* d_kr = reinterpret_cast<YdKeyPoint*>(io + o_kr);
  float* d_rx = reinterpret_cast<float*>(io + o_rx);
  float* d_dp = reinterpret_cast<float*>(io + o_dp);
  int* d_cnt = reinterpret_cast<int*>(io + o_cnt);
  int* d_start = reinterpret_cast<int*>(io + o_start);
  uint8_t* d_row = io + o_row;
  int* d_rec = reinterpret_cast<int*>(io + o_rec);
  int* d_ctr = reinterpret_cast<int*>(io + o_ctr);
  unsigned long long* d_adv = reinterpret_cast<unsigned long long*>(io + o_adv);
  int* d_lag = reinterpret_cast<int*>(io + o_lag);
  YD_CUDA(cudaMemcpyAsync(d_kl, kp_left, sizeof(YdKeyPoint) * (size_t)nl, cudaMemcpyHostToDevice, s));
  YD_CUDA(cudaMemcpyAsync(io + o_dl, desc_left, (size_t)32 * nl, cudaMemcpyHostToDevice, s));
  if (nr > 0) {
    YD_CUDA(cudaMemcpyAsync(d_kr, kp_right, sizeof(YdKeyPoint) * (size_t)nr, cudaMemcpyHostToDevice, s));
    YD_CUDA(cudaMemcpyAsync(io + o_dr, desc_right, (size_t)32 * nr, cudaMemcpyHostToDevice, s));
  }
  const int wblocks = (nl * 32 + 127) / 128;
  k_stereo_init<<<(nl + 255) / 256, 256, 0, s>>>(nl, d_rx, d_dp, d_rec, d_ctr);
  k_stereo_candidates<<<wblocks, 128, 0, s>>>(g, d_kl, nl, d_kr, nr, max_d, d_cnt, d_row, nullptr, nullptr, nullptr, 0);
  k_scan_single<<<1, 1024, 0, s>>>(d_cnt, nl, d_start);
  int npairs = 0;
  YD_CUDA(cudaMemcpyAsync(&npairs, d_start + nl, sizeof(int), cudaMemcpyDeviceToHost, s));
  YD_CUDA(cudaStreamSynchronize(s));
  const size_t s_cand = 0, s_pl = al(sizeof(int) * (size_t)std::max(npairs, 1)), s_res = s_pl + al(sizeof(int) * (size_t)std::max(npairs, 1)),
               s_total = s_res + sizeof(PairResult) * (size_t)std::max(npairs, 1);
  rc = matcher_reserve(m, s_total);
  if (rc != YDORB_OK) return rc;
  uint8_t* sc = reinterpret_cast<uint8_t*>(m->d_scratch);
  int* d_cand = reinterpret_cast<int*>(sc + s_cand);
  int* d_pl = reinterpret_cast<int*>(sc + s_pl);
  PairResult* d_res = reinterpret_cast<PairResult*>(sc + s_res);
  if (npairs > 0) {
    k_stereo_candidates<<<wblocks, 128, 0, s>>>(g, d_kl, nl, d_kr, nr, max_d, d_cnt, d_row, d_start, d_cand, d_pl, 1);
    k_stereo_sad<<<(npairs + 127) / 128, 128, 0, s>>>(g, pyr_l, pyr_r, d_kl, d_kr, d_cand, d_pl, npairs, bf, max_d, d_res);
  }
  ChainArgs A;
  A.dl = reinterpret_cast<const uint4*>(io + o_dl);
  A.dr = reinterpret_cast<const uint4*>(io + o_dr);
  A.cand_start = d_start;
  A.cand = d_cand;
  A.row_nonempty = d_row;
  A.res = d_res;
  A.nl = nl;
  A.right_x = d_rx;
  A.depth = d_dp;
  A.rec_sad = d_rec;
  A.counters = d_ctr;
  if (mode == 0) {
    const int nblk = (nl + kLagBlock - 1) / kLagBlock;
    static const bool force_chain = getenv("YDORB_STEREO_CHAIN") != nullptr;  // probe: the general form only
    if (nblk <= kLagMaxBlocks && !force_chain) {
      if ((size_t)nblk * kLagWindow > 40 * 1024)
        YD_CUDA(cudaFuncSetAttribute(k_stereo_lag_walk, cudaFuncAttributeMaxDynamicSharedMemorySize,
                                     kLagMaxBlocks * kLagWindow));
      k_stereo_lag_table<<<(nl * 64 + 127) / 128, 128, 0, s>>>(A, reinterpret_cast<uint32_t*>(d_adv));
      k_stereo_lag_walk<<<1, 1024, (size_t)nblk * kLagWindow, s>>>(d_adv, nl, d_lag, d_ctr);
      k_stereo_direct<<<wblocks, 128, 0, s>>>(A, d_lag);
    } else {
      k_stereo_force_chain<<<1, 1, 0, s>>>(d_ctr);
    }
    k_stereo_chain<<<1, 32, 0, s>>>(A);
  } else {
    k_stereo_direct<<<wblocks, 128, 0, s>>>(A, nullptr);
  }
  k_stereo_outliers<<<(nl + 255) / 256, 256, 0, s>>>(nl, d_rec, d_ctr, d_rx, d_dp);
  YD_CUDA(cudaGetLastError());
  YD_CUDA(cudaMemcpyAsync(right_x, d_rx, sizeof(float) * (size_t)nl, cudaMemcpyDeviceToHost, s));
  YD_CUDA(cudaMemcpyAsync(depth, d_dp, sizeof(float) * (size_t)nl, cudaMemcpyDeviceToHost, s));
  YD_CUDA(cudaStreamSynchronize(s));
  return YDORB_OK;
}
