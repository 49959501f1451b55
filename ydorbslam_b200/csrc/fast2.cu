// E3  per-cell cv::FAST(TYPE_9_16, nonmax=true) with the iniThFAST -> minThFAST fallback, CTA-per-block-of-cells form
//     (OrbExtractor::computeKeyPointsPyramid, src/orbExtractor.cpp:562-593; OpenCV model: SURVEY Appendix A3).
//
// The reference's cells overlap by 6 px only in what they READ; the pixels they TEST tile the level exactly
// (cell j tests [iniX_j + 3, iniX_j + wCell + 3)).  The FAST score of a pixel does not depend on the cell it is tested
// in -- only non-max suppression (a cv::Mat ROI is a standalone image: neighbours outside the cell count as 0), the
// empty-cell threshold fallback and the order of the candidate list are per cell.  So a CTA owns a block of up to
// 4 x 1 cells (YD_F2_CELL_COLS x YD_F2_CELL_ROWS) and works in dense phases over the whole block:
//   stage    one TMA box load (cp.async.bulk.tensor.3d) of the block's ROI (the box starts on a 16-byte boundary left
//            of the ROI: TMA needs that); bytes -> s16x2 pixel-pair plane, shifted so that word j of a row holds ROI
//            columns (2j - 1, 2j): the first tested column (3) is then the low half of word 2 and a row of up to
//            128 tested pixels is exactly 64 fully tested pairs;
//   reject   a warp per row, a thread per two adjacent pairs (one LDS.64 feeds both): antipodal necessary condition
//            (fast_common.cuh); survivors go to the warp's own queue (running count in a register, no atomics);
//   score    exact branch-free score of the queued pairs on full warps -> score plane, max(m - low_th, 0) per half;
//   nms      queued pairs with a non-zero score: 8-neighbour maximum with the neighbours across a cell seam masked
//            to 0 (per-column / per-row seam masks); survivors set their bit in a per-row keep bitmap;
//   cells    warp per cell, lane per row: popcount of the row's keep bits inside the cell, warp scan -> row offsets;
//            an empty cell re-runs NMS at minThFAST (:582-585); the cell reserves a slice of the frame's candidate
//            buffer and every lane walks its row's bits, which is the reference's row-major order.
// Results are bit-identical to fast.cu's warp-per-cell kernel (tests run both); that kernel remains the choice for
// launches too small to fill the GPU with these CTAs (single frames).
#include <algorithm>
#include <cstdlib>

#include "fast_common.cuh"
#include "tma.cuh"

namespace ydorb {

// Block shape (compile-time, -DYD_F2_*): threads per CTA, rows of cells per block, resident CTAs per SM the register
// budget is set for.  One row of cells and 4 warps: the phases of a CTA are separated by barriers, so many small CTAs
// per SM interleave better than few large ones (measured per 256 frames: 8 warps x 4 x 2 cells 0.614 ms, 4 warps x 4 x 1
// cells 0.584 ms, 4 warps x 2 x 1 cells 0.635 ms -- the 6 px halo of every block is scored twice).
#ifndef YD_F2_THREADS
#define YD_F2_THREADS 128
#endif
#ifndef YD_F2_CELL_ROWS
#define YD_F2_CELL_ROWS 1
#endif
#ifndef YD_F2_CELL_COLS
#define YD_F2_CELL_COLS 4
#endif
#ifndef YD_F2_CTAS
#define YD_F2_CTAS 8
#endif
constexpr int kF2Threads = YD_F2_THREADS;
constexpr int kF2Warps = kF2Threads / 32;
constexpr int kF2CellRows = YD_F2_CELL_ROWS;
constexpr int kF2CellCols = YD_F2_CELL_COLS;
constexpr int kF2MaxW = 32 * kF2CellCols + 6, kF2MaxH = 70;  // largest block ROI: 4 (2) cells of 32 + 6 wide, 2 cells of 32 + 6 (1 of 64 + 6) high
constexpr int kF2BoxW = (15 + kF2MaxW + 4 + 15) & ~15;  // TMA box: up to 15 + 1 bytes left of the ROI, one word of look-ahead (160 / 96)
constexpr int kF2Wpr = kF2MaxW / 2 + 1;     // pair words per plane row: word j = ROI columns (2j - 1, 2j), j <= 67 (35)
constexpr int kF2KeepWords = (kF2BoxW + 31) / 32;  // keep bitmap words per row (tile columns 0..159)
constexpr int kF2SeamWords = 2 * kF2Wpr + 2 * kF2MaxH;  // cmL[68] cmR[68] rmU[70] rmD[70] of one level

// shared memory map (bytes).  The small tables have room for kF2MaxH rows; the queue and the two planes are sized for
// the tallest block ROI of the CURRENT geometry (`rows`, a launch parameter; also the height of the TMA box): with one
// row of cells per block that is 36-46 rows for the usual image sizes, ~30 KB per CTA.
constexpr int kF2OffTab = 128;                                         // cmL[68] cmR[68] rmU[70] rmD[70]
constexpr int kF2OffKeep = kF2OffTab + 4 * (2 * kF2Wpr + 2 * kF2MaxH); // keep bitmap [70][5]
constexpr int kF2OffQueue = (kF2OffKeep + 4 * kF2MaxH * kF2KeepWords + 15) & ~15;  // u16 [warps][f2_queue_cap(rows)] survivors of the reject
// entries of one warp's survivor queue: a warp does at most ceil(tested rows / warps) reject steps of <= 64 pairs each
__host__ __device__ constexpr int f2_queue_cap(int rows) { return 64 * ((rows - 6 + kF2Warps - 1) / kF2Warps); }
__host__ __device__ constexpr int f2_plane_bytes(int rows) { return 4 * kF2Wpr * rows; }
__host__ __device__ constexpr int f2_off_pair(int rows) { return (kF2OffQueue + 2 * f2_queue_cap(rows) * kF2Warps + 127) & ~127; }
__host__ __device__ constexpr int f2_off_score(int rows) { return (f2_off_pair(rows) + f2_plane_bytes(rows) + 127) & ~127; }
__host__ __device__ constexpr int f2_smem_bytes(int rows) { return f2_off_score(rows) + f2_plane_bytes(rows); }
static_assert(kF2BoxW <= 4 * kF2Wpr, "the byte tile aliases the score plane");

struct Fast2Maps {
  CUtensorMap lv[kMaxLevels];
};

// reject test of the two adjacent pairs at words j (even) and j + 1 of row `row` (= &pair[r][j])
__device__ __forceinline__ void fast2_quick2(const uint32_t* __restrict__ row, uint32_t k1, bool& passA, bool& passB) {
  constexpr int S = kF2Wpr;
  const uint2 u3 = *reinterpret_cast<const uint2*>(row + 3 * S), d3 = *reinterpret_cast<const uint2*>(row - 3 * S);
  const uint2 u2 = *reinterpret_cast<const uint2*>(row + 2 * S), d2 = *reinterpret_cast<const uint2*>(row - 2 * S);
  const uint32_t u2l = row[2 * S - 1], u2r = row[2 * S + 2], d2l = row[-2 * S - 1], d2r = row[-2 * S + 2];
  const uint2 c0 = *reinterpret_cast<const uint2*>(row - 2), c1 = *reinterpret_cast<const uint2*>(row),
              c2 = *reinterpret_cast<const uint2*>(row + 2);
  {
    // pair A at word j: (0,3) (0,-3) | (3,0) (-3,0) | (2,2) (-2,-2) | (2,-2) (-2,2)
    const uint32_t cb = c1.x + 0x01000100u;
    const uint32_t p0 = u3.x, p8 = d3.x, p4 = fsr16(c1.y, c2.x), p12 = fsr16(c0.x, c0.y);
    const uint32_t p2 = u2.y, p10 = d2l, p6 = d2.y, p14 = u2l;
    const uint32_t hi = __vimax3_s16x2(__vmins2(p0, p8), __vmins2(p4, p12), __vmaxs2(__vmins2(p2, p10), __vmins2(p6, p14)));
    const uint32_t lo = __vimin3_s16x2(__vmaxs2(p0, p8), __vmaxs2(p4, p12), __vmins2(__vmaxs2(p2, p10), __vmaxs2(p6, p14)));
    passA = ((((cb - hi) + k1) | ((lo + 0x02000200u - cb) + k1)) & 0x80008000u) != 0u;
  }
  {
    // pair B at word j + 1
    const uint32_t cb = c1.y + 0x01000100u;
    const uint32_t p0 = u3.y, p8 = d3.y, p4 = fsr16(c2.x, c2.y), p12 = fsr16(c0.y, c1.x);
    const uint32_t p2 = u2r, p10 = d2.x, p6 = d2r, p14 = u2.x;
    const uint32_t hi = __vimax3_s16x2(__vmins2(p0, p8), __vmins2(p4, p12), __vmaxs2(__vmins2(p2, p10), __vmins2(p6, p14)));
    const uint32_t lo = __vimin3_s16x2(__vmaxs2(p0, p8), __vmaxs2(p4, p12), __vmins2(__vmaxs2(p2, p10), __vmaxs2(p6, p14)));
    passB = ((((cb - hi) + k1) | ((lo + 0x02000200u - cb) + k1)) & 0x80008000u) != 0u;
  }
}

// Weaker, cheaper variant: only the vertical and the horizontal antipodal pair (4 circle pixels).  Lets more pairs
// through to the exact score (which decides anyway) for 5 loads and 24 ALU operations per two pairs instead of 13 and 40.
__device__ __forceinline__ void fast2_quick2_vh(const uint32_t* __restrict__ row, uint32_t k1, bool& passA, bool& passB) {
  constexpr int S = kF2Wpr;
  const uint2 u3 = *reinterpret_cast<const uint2*>(row + 3 * S), d3 = *reinterpret_cast<const uint2*>(row - 3 * S);
  const uint2 c0 = *reinterpret_cast<const uint2*>(row - 2), c1 = *reinterpret_cast<const uint2*>(row),
              c2 = *reinterpret_cast<const uint2*>(row + 2);
  {
    const uint32_t cb = c1.x + 0x01000100u;
    const uint32_t p0 = u3.x, p8 = d3.x, p4 = fsr16(c1.y, c2.x), p12 = fsr16(c0.x, c0.y);
    const uint32_t hi = __vmaxs2(__vmins2(p0, p8), __vmins2(p4, p12));
    const uint32_t lo = __vmins2(__vmaxs2(p0, p8), __vmaxs2(p4, p12));
    passA = ((((cb - hi) + k1) | ((lo + 0x02000200u - cb) + k1)) & 0x80008000u) != 0u;
  }
  {
    const uint32_t cb = c1.y + 0x01000100u;
    const uint32_t p0 = u3.y, p8 = d3.y, p4 = fsr16(c2.x, c2.y), p12 = fsr16(c0.y, c1.x);
    const uint32_t hi = __vmaxs2(__vmins2(p0, p8), __vmins2(p4, p12));
    const uint32_t lo = __vmins2(__vmaxs2(p0, p8), __vmaxs2(p4, p12));
    passB = ((((cb - hi) + k1) | ((lo + 0x02000200u - cb) + k1)) & 0x80008000u) != 0u;
  }
}

// 8-neighbour maximum of the pair at p inside its cell, thresholded, and the keep flags (bit 15: low half survives,
// bit 31: high half).  c = the pair's own thresholded score (non-zero).
__device__ __forceinline__ uint32_t fast2_nms_flags(const uint32_t* __restrict__ p, uint32_t c, uint32_t kd, uint32_t mL,
                                                    uint32_t mR, uint32_t mU, uint32_t mD) {
  constexpr int S = kF2Wpr;
  const uint32_t cl = p[-1], c0 = p[0], cr = p[1];
  const uint32_t ul = p[-S - 1], um = p[-S], ur = p[-S + 1];
  const uint32_t dl = p[S - 1], dm = p[S], dr = p[S + 1];
  // left-shifted words hold columns (x-1, x) of the pair (x, x+1), right-shifted (x+1, x+2): the left / right
  // neighbours of both halves
  const uint32_t L = __vimax3_s16x2(fsr16(ul, um) & mU, fsr16(cl, c0), fsr16(dl, dm) & mD) & mL;
  const uint32_t R = __vimax3_s16x2(fsr16(um, ur) & mU, fsr16(c0, cr), fsr16(dm, dr) & mD) & mR;
  uint32_t n = __vimax3_s16x2(L, R, __vmaxs2(um & mU, dm & mD));
  n = __vmaxs2(n, kd) - kd;
  uint32_t flags = (c + 0x7FFF7FFFu - n) & 0x80008000u;  // half of c > half of n (all values in [0, 255])
  if ((c & 0xFFFFu) == 0u) flags &= 0x80000000u;         // a zero half is not a corner at this threshold
  if ((c >> 16) == 0u) flags &= 0x00008000u;
  return flags;
}

// bits [xa, xb) of a keep-bitmap row
__device__ __forceinline__ int fast2_row_count(const uint32_t* __restrict__ krow, int xa, int xb) {
  int n = 0;
  for (int k = xa >> 5; k <= (xb - 1) >> 5; k++) {
    uint32_t m = krow[k];
    if (k == (xa >> 5)) m &= 0xFFFFFFFFu << (xa & 31);
    if (k == ((xb - 1) >> 5) && (xb & 31)) m &= (1u << (xb & 31)) - 1u;
    n += __popc(m);
  }
  return n;
}

template <int REJECT>  // 4: four antipodal pairs in the reject, 2: vertical + horizontal only
__global__ void __launch_bounds__(kF2Threads, YD_F2_CTAS)
k_fast_blocks(const __grid_constant__ Fast2Maps maps, int n_cells, int raw_cap, const int4* __restrict__ cells,
              const FastBlock* __restrict__ blocks, const uint32_t* __restrict__ seams, int rows_cap, int frame0, int ini_th, int min_th, uint32_t* __restrict__ raw,
              int* __restrict__ raw_counter, int* __restrict__ cell_count, int* __restrict__ cell_start,
              int* __restrict__ status) {
  // Up to pdl_wait() only launch parameters and geometry tables are read (never the pyramid the predecessor wrote):
  // that part overlaps the predecessor's tail, and thread 0 issues the box load the moment the pyramid is complete.
  extern __shared__ __align__(128) uint8_t smem[];
  uint64_t* bar = reinterpret_cast<uint64_t*>(smem);
  int* s_qn = reinterpret_cast<int*>(smem + 8);  // [kF2Warps] entries of the warps' survivor queues
  uint32_t* cmL = reinterpret_cast<uint32_t*>(smem + kF2OffTab);  // [68] seam masks of the left / right neighbours
  uint32_t* cmR = cmL + kF2Wpr;
  uint32_t* rmU = cmR + kF2Wpr;                                   // [70] row above / below in the same cell
  uint32_t* rmD = rmU + kF2MaxH;
  uint32_t* keep = reinterpret_cast<uint32_t*>(smem + kF2OffKeep);       // [70][5] bit x of row r: pixel survives NMS
  uint32_t* pair = reinterpret_cast<uint32_t*>(smem + f2_off_pair(rows_cap));        // [rows_cap][68]
  uint32_t* score = reinterpret_cast<uint32_t*>(smem + f2_off_score(rows_cap));      // [rows_cap][68]
  const uint32_t* tile8 = score;  // the TMA byte tile lands in the score plane first
  const int tid = threadIdx.x, warp = tid >> 5, lane = tid & 31;
  uint16_t* queue = reinterpret_cast<uint16_t*>(smem + kF2OffQueue) + warp * f2_queue_cap(rows_cap);  // this warp's survivors
  const uint32_t lt_mask = (1u << lane) - 1u;
  const int frame = blockIdx.y;
  const FastBlock b = blocks[blockIdx.x];
  constexpr int S = kF2Wpr;
  const int low_th = min(max(min(ini_th, min_th), 0), 255);
  const int W = b.w, H = b.h;

  if (W > 0) {
    // the box starts on the 16-byte boundary at or left of ROI column -1 (plane word j = columns 2j - 1, 2j)
    const int bx = kPad + b.x0 - 1, box_x = bx & ~15, a = bx & 15;
    if (tid == 0) {
      mbar_init(bar, 1);
      pdl_wait();
      mbar_expect_tx(bar, (uint32_t)(kF2BoxW * rows_cap));
      tma_load_3d(score, &maps.lv[b.level], bar, box_x, kPad + b.y0, frame0 + frame);
    }
    // ---- while the box is in flight: keep bitmap := 0, this level's seam masks (host-built, fast2_build_blocks)
    {
      for (int i = tid; i < kF2MaxH * kF2KeepWords; i += kF2Threads) keep[i] = 0u;
      const uint32_t* src = seams + b.level * kF2SeamWords;
      for (int i = tid; i < kF2SeamWords; i += kF2Threads) cmL[i] = src[i];
    }
    if (tid != 0) pdl_wait();
    __syncthreads();  // thread 0's barrier init is visible
    mbar_wait(bar, 0);
    // ---- bytes -> pixel pairs: tile bytes a + 4w .. a + 4w + 3 of row r = columns 4w - 1 .. 4w + 2 -> words 2w, 2w + 1
    // words needed: 0 .. j1 + 2 (the reject / score of the last tested pair reads two words to its right)
    const int c_hi = W - 4;                // last tested column
    const int j1 = (c_hi + 1) >> 1;        // its word
    // a thread converts 16 bytes: five tile words -> four shifted words -> eight plane words (row = i / conv_n by a
    // host-made reciprocal)
    {
      const int k0 = a >> 2, ash = 8 * (a & 3);
      const int ng = b.conv_n, total = ng * H;
      for (int i = tid; i < total; i += kF2Threads) {
        const int r = (int)(((uint32_t)i * (uint32_t)b.conv_magic) >> 16), gq = i - r * ng;
        const uint32_t* t = tile8 + r * (kF2BoxW / 4) + k0 + 4 * gq;
        const uint32_t t0 = t[0], t1 = t[1], t2 = t[2], t3 = t[3], t4 = t[4];
        const uint32_t v0 = __funnelshift_r(t0, t1, ash), v1 = __funnelshift_r(t1, t2, ash),
                       v2 = __funnelshift_r(t2, t3, ash), v3 = __funnelshift_r(t3, t4, ash);
        uint4* o = reinterpret_cast<uint4*>(pair + r * S + 8 * gq);
        o[0] = make_uint4(__byte_perm(v0, 0, 0x4140), __byte_perm(v0, 0, 0x4342), __byte_perm(v1, 0, 0x4140), __byte_perm(v1, 0, 0x4342));
        if (8 * gq + 4 < S)
          o[1] = make_uint4(__byte_perm(v2, 0, 0x4140), __byte_perm(v2, 0, 0x4342), __byte_perm(v3, 0, 0x4140), __byte_perm(v3, 0, 0x4342));
      }
    }
    __syncthreads();  // pair plane complete; the byte tile is dead, its memory becomes the score plane
    {
      uint4* z = reinterpret_cast<uint4*>(score);
      for (int i = tid; i < f2_plane_bytes(rows_cap) / 16; i += kF2Threads) z[i] = make_uint4(0u, 0u, 0u, 0u);
    }
    // tested pixels: columns [3, W - 4] = words [2, j1], rows [3, H - 4]
    const int ipw = j1 - 1, ih = H - 6;
    // ---- reject: a thread per two adjacent pairs, tpr threads per row; survivors into the warp's own queue
    int qn = 0;  // entries of this warp's queue (warp-uniform)
    {
      const uint32_t k1 = (uint32_t)(0x7FFF - (low_th + 256)) * 0x10001u;
      const int ipw2 = (ipw + 1) >> 1;
      const int sh = ipw2 <= 8 ? 3 : ipw2 <= 16 ? 4 : 5;  // log2(threads per row)
      const int q = tid & ((1 << sh) - 1), rstep = kF2Threads >> sh;
      const int j = 2 + 2 * q;
      const bool colA = j <= j1, colB = j + 1 <= j1;
      for (int rb = 3; rb < 3 + ih; rb += rstep) {  // same trip count for every thread: the ballots need full warps
        const int r = rb + (tid >> sh);
        bool pa = false, pb = false;
        if (colA && r < 3 + ih) {
          if (REJECT == 2) fast2_quick2_vh(pair + r * S + j, k1, pa, pb);
          else fast2_quick2(pair + r * S + j, k1, pa, pb);
          pb = pb && colB;
        }
        const uint32_t ba = __ballot_sync(0xffffffffu, pa), bb = __ballot_sync(0xffffffffu, pb);
        const int na = __popc(ba);
        const uint32_t e = (uint32_t)((r << 7) | j);
        if (pa) queue[qn + __popc(ba & lt_mask)] = (uint16_t)e;
        if (pb) queue[qn + na + __popc(bb & lt_mask)] = (uint16_t)(e + 1u);
        qn += na + __popc(bb);
      }
      if (lane == 0) s_qn[warp] = qn;
    }
    __syncthreads();  // the score plane is zero everywhere (all warps cleared their share) before any score lands in it
    // the warps' queues as one list (balanced over the CTA): entry i lives in queue w at i - first[w]
    int first[kF2Warps + 1];
    first[0] = 0;
#pragma unroll
    for (int w = 0; w < kF2Warps; w++) first[w + 1] = first[w] + s_qn[w];
    const int total = first[kF2Warps];
    const uint16_t* qbase = reinterpret_cast<const uint16_t*>(smem + kF2OffQueue);
    const int qcap = f2_queue_cap(rows_cap);
    auto entry = [&](int i) -> uint32_t {
      int w = 0;
#pragma unroll
      for (int k = 1; k < kF2Warps; k++) w += (i >= first[k]) ? 1 : 0;
      int f = first[0];
#pragma unroll
      for (int k = 1; k < kF2Warps; k++) f = (i >= first[k]) ? first[k] : f;
      return qbase[w * qcap + i - f];
    };
    // ---- exact score of the survivors
    {
      const uint32_t kc = (uint32_t)(256 + low_th) * 0x10001u;
      for (int i = tid; i < total; i += kF2Threads) {
        const uint32_t e = entry(i);
        const int r = (int)(e >> 7), j = (int)(e & 127u);
        const uint32_t mm = fast_score_pair(pair + r * S + j, S);  // m + 256
        uint32_t s = __vmaxs2(mm, kc) - kc;                        // max(m - low_th, 0), no borrow
        if (2 * j > c_hi) s &= 0x0000FFFFu;                        // column W - 3 is not tested
        if (s != 0u) score[r * S + j] = s;
      }
    }
    __syncthreads();
    // ---- non-max suppression at iniThFAST over the survivors -> keep bitmap
    {
      const uint32_t kd = (uint32_t)(min(max(ini_th, 0), 255) - low_th) * 0x10001u;
      for (int i = tid; i < total; i += kF2Threads) {
        const uint32_t e = entry(i);
        const int r = (int)(e >> 7), j = (int)(e & 127u);
        const uint32_t* p = score + r * S + j;
        const uint32_t c = __vmaxs2(p[0], kd) - kd;
        if (c != 0u) {
          const uint32_t fl = fast2_nms_flags(p, c, kd, cmL[j], cmR[j], rmU[r], rmD[r]);
          if (fl & 0x8000u) atomicOr(&keep[r * kF2KeepWords + ((2 * j - 1) >> 5)], 1u << ((2 * j - 1) & 31));
          if (fl & 0x80000000u) atomicOr(&keep[r * kF2KeepWords + ((2 * j) >> 5)], 1u << ((2 * j) & 31));
        }
      }
    }
    __syncthreads();
  } else {
    pdl_wait();  // nothing to test in this block: its cells still report zero candidates below
  }

  // ---- cells: warp per cell; lane = rpl consecutive rows (rpl = 1 up to 32 tested rows, 2 up to 64), so that one warp
  // scan gives every lane the offset of its first row and the lane that counted a row is the one that emits it
  for (int ci = warp; ci < b.ncx * b.ncy; ci += kF2Warps) {
    const int cyi = ci >= b.ncx ? 1 : 0, cxi = ci - cyi * b.ncx;  // at most two rows of cells
    const int cell = b.cell0 + cyi * b.cols + cxi;
    int* out_count = cell_count + (size_t)frame * n_cells + cell;
    int* out_start = cell_start + (size_t)frame * n_cells + cell;
    const int4 roi = cells[cell];
    if (roi.z - roi.x < 7 || roi.w - roi.y < 7) {  // skipped cell (:565, :574) or nothing testable
      if (lane == 0) { *out_count = 0; *out_start = 0; }
      continue;
    }
    // tested columns [xa, xb), rows [ya, yb) in tile coordinates
    const int xa = roi.x - b.x0 + 3, xb = roi.z - b.x0 - 3, ya = roi.y - b.y0 + 3, yb = roi.w - b.y0 - 3;
    const int nrows = yb - ya;
    const int rpl = (nrows + 31) >> 5, row_lo = min(lane * rpl, nrows), row_hi = min(row_lo + rpl, nrows);
    int lane_first = 0;  // survivors of the cell in the rows above this lane's
    auto count_rows = [&]() -> int {
      int n = 0;
      for (int rr = row_lo; rr < row_hi; rr++) n += fast2_row_count(keep + (ya + rr) * kF2KeepWords, xa, xb);
      int inc = n;
#pragma unroll
      for (int d = 1; d < 32; d <<= 1) {
        const int t = __shfl_up_sync(0xffffffffu, inc, d);
        if (lane >= d) inc += t;
      }
      lane_first = inc - n;
      return __shfl_sync(0xffffffffu, inc, 31);
    };
    int count = count_rows();
    if (count == 0 && min_th != ini_th) {
      // keyPointsInCell.empty() is judged after NMS (src/orbExtractor.cpp:582): second FAST call at minThFAST.  Rare:
      // dense pass over the cell's words; the other half of a word on a seam belongs to the neighbouring cell's warp
      const uint32_t kd = (uint32_t)(min(max(min_th, 0), 255) - low_th) * 0x10001u;
      const int jA = (xa + 1) >> 1, jB = xb >> 1;
      for (int r = ya; r < yb; r++)
        for (int j = jA + lane; j <= jB; j += 32) {
          const uint32_t* p = score + r * S + j;
          uint32_t c0 = p[0];
          if (2 * j - 1 < xa) c0 &= 0xFFFF0000u;
          if (2 * j >= xb) c0 &= 0x0000FFFFu;
          const uint32_t c = __vmaxs2(c0, kd) - kd;
          if (c != 0u) {
            const uint32_t fl = fast2_nms_flags(p, c, kd, cmL[j], cmR[j], rmU[r], rmD[r]);
            if (fl & 0x8000u) atomicOr(&keep[r * kF2KeepWords + ((2 * j - 1) >> 5)], 1u << ((2 * j - 1) & 31));
            if (fl & 0x80000000u) atomicOr(&keep[r * kF2KeepWords + ((2 * j) >> 5)], 1u << ((2 * j) & 31));
          }
        }
      __syncwarp();
      count = count_rows();
    }
    if (count == 0) {
      if (lane == 0) { *out_count = 0; *out_start = 0; }
      continue;
    }
    int start = 0;
    if (lane == 0) start = atomicAdd(raw_counter + frame, count);
    start = __shfl_sync(0xffffffffu, start, 0);
    if (start + count > raw_cap) {  // candidate buffer exhausted: report, drop the cell
      if (lane == 0) { atomicOr(status, kStatusRawOverflow); *out_count = 0; *out_start = 0; }
      continue;
    }
    // ---- emit: lane = row, walking the row's keep bits gives the reference's row-major order.
    // y:12 | x:12 | response:8, coordinates relative to the 16 px border as distributeQuadTree sees them (:586-590);
    // response = m - 1 = stored + low_th - 1
    {
      const uint32_t pos_base = ((uint32_t)(b.y0 - kBorder) << 20) + ((uint32_t)(b.x0 - kBorder) << 8) + (uint32_t)(low_th - 1);
      uint32_t* dst = raw + (size_t)frame * raw_cap + start;
      uint32_t* o = dst + lane_first;
      for (int rr = row_lo; rr < row_hi; rr++) {
        const int r = ya + rr;
        const uint32_t* krow = keep + r * kF2KeepWords;
        for (int k = xa >> 5; k <= (xb - 1) >> 5; k++) {
          uint32_t m = krow[k];
          if (k == (xa >> 5)) m &= 0xFFFFFFFFu << (xa & 31);
          if (k == ((xb - 1) >> 5) && (xb & 31)) m &= (1u << (xb & 31)) - 1u;
          while (m) {
            const int x = 32 * k + __ffs(m) - 1;
            m &= m - 1u;
            const uint32_t v = score[r * S + ((x + 1) >> 1)];  // column x: low half of word (x + 1) / 2 if x is odd
            *o++ = pos_base + ((uint32_t)r << 20) + ((uint32_t)x << 8) + ((x & 1) ? (v & 0xFFu) : ((v >> 16) & 0xFFu));
          }
        }
      }
    }
    if (lane == 0) { *out_count = count; *out_start = start; }
  }
}

// Host: partition every level's cell grid into blocks whose ROI fits the shared-memory tile.
int fast2_build_blocks(const FrameGeomDev& g, const std::vector<int4>& cells, std::vector<FastBlock>* out,
                       std::vector<uint32_t>* seams, int* max_w, int* max_h) {
  constexpr int kMaxW = kF2MaxW, kMaxH = kF2MaxH;
  *max_w = 16;
  *max_h = 8;
  seams->assign((size_t)kMaxLevels * kF2SeamWords, 0xFFFFFFFFu);
  for (int l = 0; l < g.n_levels; l++) {
    const LevelGeom& lv = g.lv[l];
    if (lv.cell_cols <= 0 || lv.cell_rows <= 0) continue;
    {
      // seam masks.  first(x): tile column (row) x is the first tested column of a cell -- cells test
      // [k * pitch + 3, (k + 1) * pitch + 3) and every block starts at a cell's origin.  Plane word j holds columns
      // (2j - 1, 2j): cmL masks the left neighbours of its two halves, cmR the right ones; rmU / rmD the rows.
      auto first = [](int x, int pitch) { return x >= 3 && (x - 3) % pitch == 0; };
      uint32_t* t = seams->data() + (size_t)l * kF2SeamWords;
      for (int j = 0; j < kF2Wpr; j++) {
        const bool fl = first(2 * j - 1, lv.cell_w), fh = first(2 * j, lv.cell_w), fn = first(2 * j + 1, lv.cell_w);
        t[j] = (fl ? 0u : 0x0000FFFFu) | (fh ? 0u : 0xFFFF0000u);
        t[kF2Wpr + j] = (fh ? 0u : 0x0000FFFFu) | (fn ? 0u : 0xFFFF0000u);
      }
      for (int r = 0; r < kF2MaxH; r++) {
        t[2 * kF2Wpr + r] = first(r, lv.cell_h) ? 0u : 0xFFFFFFFFu;
        t[2 * kF2Wpr + kF2MaxH + r] = first(r + 1, lv.cell_h) ? 0u : 0xFFFFFFFFu;
      }
    }
    const int bx = std::max(1, std::min(kF2CellCols, (kMaxW - 6) / lv.cell_w)), by = std::max(1, std::min(kF2CellRows, (kMaxH - 6) / lv.cell_h));
    for (int cy = 0; cy < lv.cell_rows; cy += by)
      for (int cx = 0; cx < lv.cell_cols; cx += bx) {
        FastBlock b = {};
        b.level = l;
        b.cell0 = lv.cell_base + cy * lv.cell_cols + cx;
        b.ncx = std::min(bx, lv.cell_cols - cx);
        b.ncy = std::min(by, lv.cell_rows - cy);
        b.cols = lv.cell_cols;
        int x0 = 1 << 30, y0 = 1 << 30, x1 = -1, y1 = -1;
        for (int i = 0; i < b.ncy; i++)
          for (int j = 0; j < b.ncx; j++) {
            const int4 c = cells[b.cell0 + i * lv.cell_cols + j];
            if (c.z - c.x < 7 || c.w - c.y < 7) continue;
            x0 = std::min(x0, c.x); y0 = std::min(y0, c.y);
            x1 = std::max(x1, c.z); y1 = std::max(y1, c.w);
          }
        if (x1 > x0) {
          b.x0 = x0; b.y0 = y0; b.w = x1 - x0; b.h = y1 - y0;
          // the block's first cell must be the one at its ROI origin (seam masks count cells from column / row 3)
          const int4 c0 = cells[b.cell0];
          if (c0.x != x0 || c0.y != y0 || c0.z - c0.x < 7 || c0.w - c0.y < 7) {
            set_error("internal: FAST block does not start at a valid cell");
            return YDORB_E_INVALID;
          }
          if (b.w > kF2MaxW || b.h > kF2MaxH || b.w < 7 || b.h < 7) {
            set_error("internal: FAST block larger than the shared-memory tile");
            return YDORB_E_INVALID;
          }
          {
            // conversion plan: plane words 0 .. j1 + 2 are needed, 8 per thread
            const int j1 = (b.w - 3) >> 1, nw4 = ((j1 + 2) >> 1) + 1;
            b.conv_n = (nw4 + 3) / 4;
            b.conv_magic = 65536 / b.conv_n + 1;
            for (int i = 0; i < b.conv_n * kF2MaxH; i++)
              if ((int)(((uint32_t)i * (uint32_t)b.conv_magic) >> 16) != i / b.conv_n) {
                set_error("internal: FAST block reciprocal");
                return YDORB_E_INVALID;
              }
          }
          *max_w = std::max(*max_w, b.w);
          *max_h = std::max(*max_h, b.h);
        }
        out->push_back(b);
      }
  }
  return YDORB_OK;
}

size_t fast2_smem_bytes(int, int max_h) { return (size_t)f2_smem_bytes(std::max(max_h, 8)); }

void fast2_box(int, int max_h, int* box_w, int* box_h) {
  *box_w = kF2BoxW;
  *box_h = std::max(max_h, 8);
}

int launch_fast_blocks(const YdExtractor* ex, int batch, cudaStream_t s) {
  const int ini = ex->params.ini_th_fast;
  const int mn = ex->params.reference_min_th_quirk ? ini : ex->params.min_th_fast;
  const size_t smem = fast2_smem_bytes(ex->fb_max_w, ex->fb_max_h);
  // reject test: 2 = vertical + horizontal antipodal pairs only (weaker, so more pairs reach the exact score, but half
  // the arithmetic on every pair: 0.569 -> 0.556 ms per 256 frames), 4 = all four pairs.  Same results either way.
  static const int reject = getenv("YDORB_FAST_REJECT") ? atoi(getenv("YDORB_FAST_REJECT")) : 2;
  auto kernel = reject == 2 ? k_fast_blocks<2> : k_fast_blocks<4>;
  if (smem > 48 * 1024) YD_CUDA(cudaFuncSetAttribute(kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem));
  Fast2Maps maps;
  for (int l = 0; l < kMaxLevels; l++) maps.lv[l] = ex->tmap_level[l];
  const int frame0 = (int)((ex->d_pyr - ex->pyr_base) / ex->geom.pyr_bytes);
  dim3 grid(ex->n_fast_blocks, batch);
  YD_CUDA(launch_chain(kernel, grid, dim3(kF2Threads), smem, s, maps, ex->geom.n_cells, ex->geom.raw_cap, ex->d_cells, ex->d_fast_blocks, ex->d_fast_seams, std::max(ex->fb_max_h, 8), frame0,
                       ini, mn, ex->d_raw, ex->d_raw_counter, ex->d_cell_count, ex->d_cell_start,
                       ex->d_status));
  count_launches(1);
  YD_CUDA(cudaGetLastError());
  return YDORB_OK;
}

}  // namespace ydorb
