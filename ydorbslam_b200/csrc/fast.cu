// E3  per-cell cv::FAST(TYPE_9_16, nonmax=true) with the iniThFAST -> minThFAST fallback
//     (OrbExtractor::computeKeyPointsPyramid, src/orbExtractor.cpp:562-593; OpenCV model: SURVEY Appendix A3).
//
// One warp owns one reference cell: its ROI (incl. the 3 px apron) is staged in shared memory, non-max suppression
// sees only the cell's own scores (a cv::Mat ROI behaves like a standalone image, so NMS never looks across a cell
// seam), and warp ballots give the row-major order the reference's candidate list has inside a cell.  Cells reserve
// a contiguous slice of the frame's candidate buffer with one atomicAdd; the global (cell-major) order is restored
// later from the per-cell counts.
//
// Scoring is branch-free and exact for EVERY tested pixel: with d_k = I(p) - I(circle_k),
//     m = max( max_arcs min_arc d ,  -min_arcs max_arc d ),   corner <=> m > t,   response = m - 1,
// which does not depend on the threshold, so one scoring pass serves both thresholds of the fallback.  A thread
// scores two horizontally adjacent pixels at once on packed s16x2 values with the DPX 3-input min/max
// (VIMNMX3.S16x2): the tile is stored as 16-bit pixel pairs, d is kept biased by +256 so packed subtraction never
// borrows across halves, the 9-arc minimum is min3(min3) over a circular array (16+16 ops), the max over arcs 8 ops.
// This is integer-pipe work (about 75 instructions per pixel), not HBM-bound: the level is read from HBM once.
#include "kernels.cuh"

namespace ydorb {

constexpr int kWarpsPerCta = 4;

// circular 9-arc extremum: r[k] = op3 over D[k..k+8]
#define YD_ARC_TREE(OP3, D, OUT)                                                   \
  {                                                                                \
    uint32_t w3[16];                                                               \
    _Pragma("unroll") for (int k = 0; k < 16; k++) w3[k] = OP3(D[k], D[(k + 1) & 15], D[(k + 2) & 15]); \
    _Pragma("unroll") for (int k = 0; k < 16; k++) OUT[k] = OP3(w3[k], w3[(k + 3) & 15], w3[(k + 6) & 15]); \
  }

__device__ __forceinline__ uint32_t max16_s16x2(const uint32_t (&v)[16]) {
  uint32_t a = __vimax3_s16x2(v[0], v[1], v[2]), b = __vimax3_s16x2(v[3], v[4], v[5]);
  uint32_t c = __vimax3_s16x2(v[6], v[7], v[8]), d = __vimax3_s16x2(v[9], v[10], v[11]);
  uint32_t e = __vimax3_s16x2(v[12], v[13], v[14]);
  a = __vimax3_s16x2(a, b, c);
  d = __vimax3_s16x2(d, e, v[15]);
  return __vmaxs2(a, d);
}
__device__ __forceinline__ uint32_t min16_s16x2(const uint32_t (&v)[16]) {
  uint32_t a = __vimin3_s16x2(v[0], v[1], v[2]), b = __vimin3_s16x2(v[3], v[4], v[5]);
  uint32_t c = __vimin3_s16x2(v[6], v[7], v[8]), d = __vimin3_s16x2(v[9], v[10], v[11]);
  uint32_t e = __vimin3_s16x2(v[12], v[13], v[14]);
  a = __vimin3_s16x2(a, b, c);
  d = __vimin3_s16x2(d, e, v[15]);
  return __vmins2(a, d);
}

// FAST-9/16 score m of the two pixels held in the 16-bit pair word at `row` (tile columns 2j-1+? see k_fast_cells).
// Returns (m_lo, m_hi) as two ints in [-255, 255].
__device__ __forceinline__ void fast_score_pair(const uint32_t* __restrict__ row, int wpr, int& m_lo, int& m_hi) {
  // pair word j covers tile columns (2j-1, 2j); this pair starts at an odd tile column c, word j = (c+1)/2.
  // even dx: word j + dx/2; odd dx: funnel shift of words j+(dx-1)/2 and j+(dx+1)/2.
  auto fsr = [](uint32_t lo, uint32_t hi) { return __funnelshift_r(lo, hi, 16); };
  uint32_t P[16];
  {
    const uint32_t* r = row + 3 * wpr;  // dy = +3: k0 (0,3) k1 (1,3) k15 (-1,3)
    const uint32_t a = r[-1], b = r[0], c = r[1];
    P[0] = b; P[1] = fsr(b, c); P[15] = fsr(a, b);
  }
  {
    const uint32_t* r = row - 3 * wpr;  // dy = -3: k8 (0,-3) k7 (1,-3) k9 (-1,-3)
    const uint32_t a = r[-1], b = r[0], c = r[1];
    P[8] = b; P[7] = fsr(b, c); P[9] = fsr(a, b);
  }
  P[2] = row[2 * wpr + 1];    // (2,2)
  P[14] = row[2 * wpr - 1];   // (-2,2)
  P[6] = row[-2 * wpr + 1];   // (2,-2)
  P[10] = row[-2 * wpr - 1];  // (-2,-2)
  P[3] = fsr(row[wpr + 1], row[wpr + 2]);      // (3,1)
  P[13] = fsr(row[wpr - 2], row[wpr - 1]);     // (-3,1)
  P[5] = fsr(row[-wpr + 1], row[-wpr + 2]);    // (3,-1)
  P[11] = fsr(row[-wpr - 2], row[-wpr - 1]);   // (-3,-1)
  P[4] = fsr(row[1], row[2]);                  // (3,0)
  P[12] = fsr(row[-2], row[-1]);               // (-3,0)
  const uint32_t cb = row[0] + 0x01000100u;    // centre + 256 in both halves
  uint32_t D[16];
#pragma unroll
  for (int k = 0; k < 16; k++) D[k] = cb - P[k];  // halves in [1, 511]: no borrow between halves
  uint32_t lo9[16], hi9[16];
  YD_ARC_TREE(__vimin3_s16x2, D, lo9)
  YD_ARC_TREE(__vimax3_s16x2, D, hi9)
  const uint32_t dark = max16_s16x2(lo9);    // max over arcs of min d (+256)
  const uint32_t bright = min16_s16x2(hi9);  // min over arcs of max d (+256)
  m_lo = max((int)(dark & 0xFFFFu) - 256, 256 - (int)(bright & 0xFFFFu));
  m_hi = max((int)(dark >> 16) - 256, 256 - (int)(bright >> 16));
}

struct FastSmemLayout {
  int wpr;          // pair words per tile row
  int sstride;      // bytes per score row
  int tile_words;   // per warp
  int score_bytes;  // per warp
  int warp_bytes;
};

__host__ __device__ inline FastSmemLayout fast_layout(int max_tw, int max_th) {
  FastSmemLayout L;
  L.wpr = (max_tw + 1) / 2 + 2;              // word j covers cols (2j-1, 2j); +1 word of slack each side
  L.sstride = (max_tw + 2 + 3) & ~3;
  L.tile_words = L.wpr * max_th;
  L.score_bytes = L.sstride * (max_th + 2);
  L.warp_bytes = (L.tile_words * 4 + L.score_bytes + 128 * 4 + 15) & ~15;  // + ballot words
  return L;
}

__global__ void __launch_bounds__(kWarpsPerCta * 32)
k_fast_cells(const FrameGeomDev* __restrict__ g, const int4* __restrict__ cells, const uint8_t* __restrict__ pyr,
             int ini_th, int min_th, int max_tw, int max_th, uint32_t* __restrict__ raw, int* __restrict__ raw_counter,
             int* __restrict__ cell_count, int* __restrict__ cell_start, int* __restrict__ status) {
  extern __shared__ __align__(16) uint8_t smem_raw[];
  const FastSmemLayout L = fast_layout(max_tw, max_th);
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
  uint8_t* wbase = smem_raw + (size_t)warp * L.warp_bytes;
  uint32_t* tile = reinterpret_cast<uint32_t*>(wbase);                 // [th][wpr] pair words, word 1 = cols (-1,0)
  uint8_t* score = wbase + L.tile_words * 4 + L.sstride;               // [th][sstride], one zero row above
  uint32_t* ballots = reinterpret_cast<uint32_t*>(wbase + L.tile_words * 4 + L.score_bytes);
  const int wpr = L.wpr, S = L.sstride;
  // stage the tile as 16-bit pixel pairs: word (r, 1 + j) = (col 2j-1, col 2j); columns outside the tile read 0
  const int nwords = (tw + 2) / 2;  // j = 0 .. nwords-1 covers cols -1 .. tw (at most)
  for (int r = 0; r < th_; r++) {
    const uint8_t* srow = src + (size_t)r * lv.stride;
    for (int j = lane; j < nwords; j += 32) {
      const int c0 = 2 * j - 1, c1 = 2 * j;
      const uint32_t v0 = (c0 >= 0 && c0 < tw) ? srow[c0] : 0u;
      const uint32_t v1 = (c1 < tw) ? srow[c1] : 0u;
      tile[r * wpr + 1 + j] = v0 | (v1 << 16);
    }
    for (int c = lane; c < S; c += 32) score[r * S + c] = 0;
  }
  for (int c = lane; c < S; c += 32) { score[-S + c] = 0; score[th_ * S + c] = 0; }
  __syncwarp();
  // score every tested pixel; a lane handles the pair at tile columns (c, c+1), c odd, rows 3..th-4
  const int iw = tw - 6, ih = th_ - 6;
  const int ipw = (iw + 1) / 2;  // pairs per row
  const int npairs = ipw * ih;
  const int low_th = min(max(min(ini_th, min_th), 0), 255);
  {
    int r = 0, jj = lane;  // pair (r, jj) without a division per step
    while (jj >= ipw) { jj -= ipw; r++; }
    for (int p = lane; p < npairs; p += 32) {
      const int c = 3 + 2 * jj, rr = r + 3;
      int m0, m1;
      fast_score_pair(tile + rr * wpr + 1 + (c + 1) / 2, wpr, m0, m1);
      // stored score: m if corner at the lower threshold, else 0 (m <= 255)
      score[rr * S + c] = (uint8_t)(m0 > low_th ? m0 : 0);
      if (c + 1 < tw - 3) score[rr * S + c + 1] = (uint8_t)(m1 > low_th ? m1 : 0);
      jj += 32;
      while (jj >= ipw) { jj -= ipw; r++; }
    }
  }
  __syncwarp();
  // NMS at threshold th: corners are pixels with score > th; strictly greater than all 8 neighbours, where
  // non-corners (and untested pixels) count 0.  Row-major pixel index i over the tested area.
  const int total = iw * ih;
  auto nms_keep = [&](int i, int th, int& r, int& c, int& m) -> bool {
    if (i >= total) return false;
    r = i / iw + 3;
    c = i - (r - 3) * iw + 3;
    const uint8_t* s = score + r * S + c;
    m = s[0];
    if (m <= th) return false;
    auto e = [&](int v) { return v > th ? v : 0; };
    return m > e(s[-1]) && m > e(s[1]) && m > e(s[-S - 1]) && m > e(s[-S]) && m > e(s[-S + 1]) && m > e(s[S - 1]) &&
           m > e(s[S]) && m > e(s[S + 1]);
  };
  int count = 0, th_used = 0;
  for (int attempt = 0; attempt < 2 && count == 0; attempt++) {
    if (attempt == 1 && min_th == ini_th) break;  // same threshold, same (empty) answer (reference: SURVEY F3)
    th_used = min(max(attempt == 0 ? ini_th : min_th, 0), 255);
    // keyPointsInCell.empty() is judged after NMS (src/orbExtractor.cpp:582)
    for (int base = 0, w = 0; base < total; base += 32, w++) {
      int r, c, m;
      const uint32_t bal = __ballot_sync(0xffffffffu, nms_keep(base + lane, th_used, r, c, m));
      if (lane == 0) ballots[w] = bal;
      count += __popc(bal);
    }
  }
  __syncwarp();
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
  for (int base = 0, w = 0; base < total; base += 32, w++) {
    const uint32_t bal = ballots[w];
    if (bal & (1u << lane)) {
      const int i = base + lane;
      const int r = i / iw + 3, c = i - (r - 3) * iw + 3;
      const int m = score[r * S + c];
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
  const FastSmemLayout L = fast_layout(ex->max_cell_w, ex->max_cell_h);
  const size_t smem = (size_t)L.warp_bytes * kWarpsPerCta;
  if (smem > 48 * 1024) YD_CUDA(cudaFuncSetAttribute(k_fast_cells, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem));
  dim3 grid((ex->geom.n_cells + kWarpsPerCta - 1) / kWarpsPerCta, batch);
  k_fast_cells<<<grid, kWarpsPerCta * 32, smem, s>>>(ex->d_geom, ex->d_cells, ex->d_pyr, ini, mn, ex->max_cell_w,
                                                     ex->max_cell_h, ex->d_raw, ex->d_raw_counter, ex->d_cell_count,
                                                     ex->d_cell_start, ex->d_status);
  count_launches(1);
  YD_CUDA(cudaGetLastError());
  return YDORB_OK;
}

}  // namespace ydorb
