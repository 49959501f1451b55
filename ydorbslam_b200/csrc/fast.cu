// E3  per-cell cv::FAST(TYPE_9_16, nonmax=true) with the iniThFAST -> minThFAST fallback
//     (OrbExtractor::computeKeyPointsPyramid, src/orbExtractor.cpp:562-593; OpenCV model: SURVEY Appendix A3).
//
// One warp owns one reference cell: its ROI (<= 66x66 px incl. the 3 px apron) is staged in shared memory, the
// corner test and score run from there, non-max suppression sees only the cell's own scores (a cv::Mat ROI behaves
// like a standalone image, so NMS never looks across a cell seam), and warp ballots give the row-major order the
// reference's candidate list has inside a cell.  Cells reserve a contiguous slice of the frame's candidate buffer
// with one atomicAdd; the global (cell-major) order is restored later from the per-cell counts.
#include "kernels.cuh"

namespace ydorb {

constexpr int kTileMax = 66;     // cell_w <= 59 (width/cols with cols=floor(width/30)), +6 apron, +1 spare
constexpr int kTileStride = 68;  // bytes per smem tile row
constexpr int kWarpsPerCta = 4;

__device__ __forceinline__ bool has_arc9(uint32_t m16) {
  // 9 contiguous set bits in a circular 16-bit mask
  uint32_t m = m16 | (m16 << 16);
  uint32_t t = m & (m >> 1);
  t &= t >> 2;
  t &= t >> 4;       // bit i set <=> bits i..i+7 set
  t &= m >> 8;       // ... and bit i+8
  return (t & 0xFFFFu) != 0;
}

// max over the 16 arcs of 9 consecutive circle pixels of min(d) -- log-tree of running minima
__device__ __forceinline__ int arc9_max_min(const int (&d)[16]) {
  int w2[16], w4[16], w8[16];
#pragma unroll
  for (int k = 0; k < 16; k++) w2[k] = min(d[k], d[(k + 1) & 15]);
#pragma unroll
  for (int k = 0; k < 16; k++) w4[k] = min(w2[k], w2[(k + 2) & 15]);
#pragma unroll
  for (int k = 0; k < 16; k++) w8[k] = min(w4[k], w4[(k + 4) & 15]);
  int best = -256;
#pragma unroll
  for (int k = 0; k < 16; k++) best = max(best, min(w8[k], d[(k + 8) & 15]));
  return best;
}

// FAST-9/16 score m of the pixel at tile position p (0 if not a corner at threshold th): corner iff m > th,
// OpenCV's response is m-1.
__device__ __forceinline__ int fast_score(const uint8_t* p, int th) {
  // circle offsets in OpenCV order (dx,dy): (0,3)(1,3)(2,2)(3,1)(3,0)(3,-1)(2,-2)(1,-3)(0,-3)(-1,-3)(-2,-2)(-3,-1)(-3,0)(-3,1)(-2,2)(-1,3)
  constexpr int S = kTileStride;
  const int c = p[0];
  int d[16];
  d[0] = c - p[3 * S];       d[1] = c - p[3 * S + 1];   d[2] = c - p[2 * S + 2];   d[3] = c - p[S + 3];
  d[4] = c - p[3];           d[5] = c - p[-S + 3];      d[6] = c - p[-2 * S + 2];  d[7] = c - p[-3 * S + 1];
  d[8] = c - p[-3 * S];      d[9] = c - p[-3 * S - 1];  d[10] = c - p[-2 * S - 2]; d[11] = c - p[-S - 3];
  d[12] = c - p[-3];         d[13] = c - p[S - 3];      d[14] = c - p[2 * S - 2];  d[15] = c - p[3 * S - 1];
  uint32_t dark = 0, bright = 0;  // d>th : circle pixel darker than centre-th ; d<-th : brighter
#pragma unroll
  for (int k = 0; k < 16; k++) {
    dark |= (uint32_t)(d[k] > th) << k;
    bright |= (uint32_t)(d[k] < -th) << k;
  }
  if (!has_arc9(dark) && !has_arc9(bright)) return 0;
  int m = arc9_max_min(d);
#pragma unroll
  for (int k = 0; k < 16; k++) d[k] = -d[k];
  m = max(m, arc9_max_min(d));
  return m;  // > th by construction
}

__global__ void __launch_bounds__(kWarpsPerCta * 32)
k_fast_cells(const FrameGeomDev* __restrict__ g, const int4* __restrict__ cells, const uint8_t* __restrict__ pyr,
             int ini_th, int min_th, uint32_t* __restrict__ raw, int* __restrict__ raw_counter,
             int* __restrict__ cell_count, int* __restrict__ cell_start, int* __restrict__ status) {
  __shared__ __align__(16) uint8_t s_tile[kWarpsPerCta][kTileMax * kTileStride];
  __shared__ __align__(16) uint8_t s_score[kWarpsPerCta][kTileMax * kTileStride];
  const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
  const int cell = blockIdx.x * kWarpsPerCta + warp;
  const int frame = blockIdx.y;
  if (cell >= g->n_cells) return;
  const int4 roi = cells[cell];
  const int tw = roi.z - roi.x, th_ = roi.w - roi.y;
  int* out_count = cell_count + (size_t)frame * g->n_cells + cell;
  int* out_start = cell_start + (size_t)frame * g->n_cells + cell;
  if (tw < 7 || th_ < 7) {  // skipped cell or nothing testable
    if (lane == 0) { *out_count = 0; *out_start = 0; }
    return;
  }
  int level = 0;
  for (int l = 1; l < g->n_levels; l++)
    if (cell >= g->lv[l].cell_base) level = l;
  const LevelGeom lv = g->lv[level];
  const uint8_t* src = pyr + (size_t)frame * g->pyr_bytes + level_origin(lv) + (size_t)roi.y * lv.stride + roi.x;
  uint8_t* tile = s_tile[warp];
  uint8_t* score = s_score[warp];
  for (int r = 0; r < th_; r++) {
    const uint8_t* srow = src + (size_t)r * lv.stride;
    for (int c = lane; c < tw; c += 32) tile[r * kTileStride + c] = srow[c];
    for (int c = lane; c < tw; c += 32) score[r * kTileStride + c] = 0;
  }
  __syncwarp();
  const int iw = tw - 6, ih = th_ - 6, total = iw * ih;
  constexpr int S = kTileStride;
  // NMS predicate (strictly greater than all 8 neighbours; untested pixels count 0), row-major pixel i
  auto nms_keep = [&](int i, int& r, int& c, int& m) -> bool {
    if (i >= total) return false;
    r = i / iw + 3;
    c = i - (r - 3) * iw + 3;
    const uint8_t* s = score + r * S + c;
    m = s[0];
    return m > 0 && m > s[-1] && m > s[1] && m > s[-S - 1] && m > s[-S] && m > s[-S + 1] && m > s[S - 1] &&
           m > s[S] && m > s[S + 1];
  };
  int count = 0;
  for (int attempt = 0; attempt < 2 && count == 0; attempt++) {
    if (attempt == 1 && min_th == ini_th) break;  // same threshold, same (empty) answer (reference: SURVEY F3)
    const int th = min(max(attempt == 0 ? ini_th : min_th, 0), 255);
    for (int i = lane; i < total; i += 32) {
      const int r = i / iw + 3, c = i - (r - 3) * iw + 3;
      score[r * S + c] = (uint8_t)fast_score(tile + r * S + c, th);
    }
    __syncwarp();
    // keyPointsInCell.empty() is judged after NMS (src/orbExtractor.cpp:582)
    for (int base = 0; base < total; base += 32) {
      int r, c, m;
      count += __popc(__ballot_sync(0xffffffffu, nms_keep(base + lane, r, c, m)));
    }
    __syncwarp();
  }
  if (count == 0) {
    if (lane == 0) { *out_count = 0; *out_start = 0; }
    return;
  }
  int start = 0;
  if (lane == 0) start = atomicAdd(raw_counter + frame, count);
  start = __shfl_sync(0xffffffffu, start, 0);
  if (start + count > g->raw_cap) {  // candidate buffer exhausted: report, drop the cell
    if (lane == 0) { atomicOr(status, kStatusRawOverflow); *out_count = 0; *out_start = 0; }
    return;
  }
  int pos = start;
  for (int base = 0; base < total; base += 32) {
    int r = 0, c = 0, m = 0;
    const bool keep = nms_keep(base + lane, r, c, m);
    const uint32_t bal = __ballot_sync(0xffffffffu, keep);
    if (keep) {
      const int slot = pos + __popc(bal & ((1u << lane) - 1u));
      // coordinates relative to the 16 px border, as distributeQuadTree sees them (:586-590); 12 bits each
      raw[(size_t)frame * g->raw_cap + slot] =
          ((uint32_t)(r + roi.y - kBorder) << 20) | ((uint32_t)(c + roi.x - kBorder) << 8) | (uint32_t)(m - 1);
    }
    pos += __popc(bal);
  }
  if (lane == 0) { *out_count = count; *out_start = start; }
}

int launch_fast(const YdExtractor* ex, int batch, cudaStream_t s) {
  YD_CUDA(cudaMemsetAsync(ex->d_raw_counter, 0, sizeof(int) * batch, s));
  const int ini = ex->params.ini_th_fast;
  const int mn = ex->params.reference_min_th_quirk ? ini : ex->params.min_th_fast;
  dim3 grid((ex->geom.n_cells + kWarpsPerCta - 1) / kWarpsPerCta, batch);
  k_fast_cells<<<grid, kWarpsPerCta * 32, 0, s>>>(ex->d_geom, ex->d_cells, ex->d_pyr, ini, mn, ex->d_raw, ex->d_raw_counter,
                                                  ex->d_cell_count, ex->d_cell_start, ex->d_status);
  count_launches(1);
  YD_CUDA(cudaGetLastError());
  return YDORB_OK;
}

}  // namespace ydorb
