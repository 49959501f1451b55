// Frame feature state on the device (keypoints, descriptors, rightX, the 64 x 48 grid as CSR) and the device helpers
// every window search shares.  G1 Frame::assignKeyPointsToGrid (src/frame.cpp:249-264), G2 getKeyPointsInArea (:337-361).
#pragma once
#include "matcher.cuh"

namespace ydorb {

constexpr int kGridCols = 64, kGridRows = 48;  // src/frame.hpp:137-138
constexpr int kGridCells = kGridCols * kGridRows;
constexpr int kNoBlocker = 0x7FFFFFFF;

struct GridParams {
  float min_x, min_y, inv_w, inv_h;
  float sf[kMaxLevels];
};

}  // namespace ydorb

struct YdFrame {
  YdMatcher* m;
  int n;
  int n_levels;
  ydorb::GridParams gp;
  YdKeyPoint* d_kps;
  uint4* d_desc;
  float* d_right_x;
  int* d_cell_start;  // [kGridCells + 1]
  int* d_cell_items;  // [n]
  int* d_cell_of;     // [n] scratch
  int* d_blocker;     // [n] two buffers + changed flag
  int* d_blocker2;
  int* d_last;        // [n] last writer
  int* d_flags;       // [4]
};

namespace ydorb {

__device__ __forceinline__ int round_half_away(float v) { return (int)round((double)v); }

__device__ __forceinline__ int grid_cell(const GridParams& gp, float x, float y) {
  const int lx = round_half_away(__fmul_rn(__fsub_rn(x, gp.min_x), gp.inv_w));  // :328
  const int ly = round_half_away(__fmul_rn(__fsub_rn(y, gp.min_x), gp.inv_h));  // :329 (minX, as the reference)
  if (lx < 0 || lx >= kGridCols || ly < 0 || ly >= kGridRows) return -1;        // :331
  return lx * kGridRows + ly;
}

struct CellRange { int x0, x1, y0, y1; };

__device__ __forceinline__ bool cell_range(const GridParams& gp, float x, float y, float r, CellRange& cr) {
  cr.x0 = max(0, (int)floor((double)__fmul_rn(__fsub_rn(__fsub_rn(x, gp.min_x), r), gp.inv_w)));               // :340
  cr.x1 = min(kGridCols - 1, (int)ceil((double)__fmul_rn(__fadd_rn(__fsub_rn(x, gp.min_x), r), gp.inv_w)));    // :341
  cr.y0 = max(0, (int)floor((double)__fmul_rn(__fsub_rn(__fsub_rn(y, gp.min_y), r), gp.inv_h)));               // :342
  cr.y1 = min(kGridRows - 1, (int)ceil((double)__fmul_rn(__fadd_rn(__fsub_rn(y, gp.min_y), r), gp.inv_h)));    // :343
  return cr.x0 < kGridCols && cr.x1 >= 0 && cr.y0 < kGridRows && cr.y1 >= 0;                                   // :344
}

// level filter + window predicate exactly as written in the reference
__device__ __forceinline__ bool in_window(const YdKeyPoint& kp, float x, float y, float r, int minL, int maxL) {
  if (minL > 0 || maxL >= 0) {                                                    // :348
    if (kp.octave < minL || (maxL >= 0 && kp.octave < maxL)) return false;        // :349
  }
  return fabsf(__fsub_rn(kp.x, x)) > r && fabsf(__fsub_rn(kp.y, y)) < r;          // :353
}

__device__ __forceinline__ int hamming256(const uint4& a0, const uint4& a1, const uint4& b0, const uint4& b1) {
  return __popc(a0.x ^ b0.x) + __popc(a0.y ^ b0.y) + __popc(a0.z ^ b0.z) + __popc(a0.w ^ b0.w) +
         __popc(a1.x ^ b1.x) + __popc(a1.y ^ b1.y) + __popc(a1.z ^ b1.z) + __popc(a1.w ^ b1.w);
}

}  // namespace ydorb
