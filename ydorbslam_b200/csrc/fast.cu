// E3  per-cell cv::FAST(TYPE_9_16, nonmax=true) with the iniThFAST -> minThFAST fallback
//     (OrbExtractor::computeKeyPointsPyramid, src/orbExtractor.cpp:562-593; OpenCV model: SURVEY Appendix A3).
//
// One warp owns one reference cell: its ROI (incl. the 3 px apron) is staged in shared memory, non-max suppression
// sees only the cell's own scores (a cv::Mat ROI behaves like a standalone image, so NMS never looks across a cell
// seam), and warp ballots give the row-major order the reference's candidate list has inside a cell.  Cells reserve
// a contiguous slice of the frame's candidate buffer with one atomicAdd; the global (cell-major) order is restored
// later from the per-cell counts.
//
// Everything is done on packed s16x2 pixel pairs (two horizontally adjacent pixels per 32-bit word, DPX VIMNMX3):
//   stage   aligned 32-bit global loads -> two pair words per load (PRMT), tile column 0 is 4-byte aligned in HBM;
//   reject  necessary condition for a 9-arc: every antipodal pair (k, k+8) of the circle has a member on the arc, so
//           min_k max(d_k, d_k+8) > t  or  max_k min(d_k, d_k+8) < -t  must hold; tested on k = 0,2,4,6 (8 loads);
//           pairs that pass are compacted through a small per-warp queue so the exact score runs on full warps;
//   score   exact and branch-free: with d_k = I(p) - I(circle_k),
//               m = max( max_arcs min_arc d ,  -min_arcs max_arc d ),   corner <=> m > t,   response = m - 1;
//           d is kept biased by +256 so packed subtraction never borrows across halves, the 9-arc minimum is
//           min3(min3) over a circular array (16+16 ops), the max over arcs 8 ops.  Stored value: max(m - t_low, 0);
//   nms     packed: the 8-neighbour maximum of a pair is 4 funnel shifts + 4 VIMNMX; keep <=> centre > neighbours.
// This is integer-pipe work, not HBM-bound: the level is read from HBM once.
#include <cstdlib>
#include "fast_common.cuh"

namespace ydorb {

constexpr int kWarpsPerCta = 4;
constexpr int kFastQueue = 64;  // compaction queue entries per warp

struct FastSmemLayout {
  int wpr;          // pair words per tile row (even)
  int plane_words;  // words of one [max_th][wpr] plane incl. 2 slack words on either side
  int nz_cap;       // tested pairs per cell (entries of the non-zero-score list, 2 bytes each)
  int warp_bytes;
};

__host__ __device__ inline FastSmemLayout fast_layout(int max_tw, int max_th) {
  FastSmemLayout L;
  const int nw4 = (4 + max_tw + 3) / 4;  // aligned 32-bit loads per row, worst-case alignment
  L.wpr = 2 * nw4;
  L.plane_words = L.wpr * max_th + 4;
  L.nz_cap = ((max_tw - 6) / 2 + 2) * (max_th - 6) + 32;
  L.warp_bytes = (2 * L.plane_words * 4 + L.nz_cap * 2 + kFastQueue * 2 + 15) & ~15;
  return L;
}

constexpr int kFastPf = 14;       // 32-bit words per lane prefetched for the next cell (covers 36 x 36 px cells)

struct FastCell {
  int valid;
  int cell;
  int4 roi;
  int tw, th, a, nw4;      // ROI extents, tile column of ROI column 0, aligned 32-bit loads per row
  const uint8_t* src_al;   // 4-byte aligned address of tile (0, 0)
  int stride;
};

__device__ __forceinline__ FastCell fast_cell_setup(const FrameGeomDev* __restrict__ g, const int4* __restrict__ cells,
                                                    const uint8_t* __restrict__ pyr, int cell, int frame, int lane,
                                                    int* __restrict__ cell_count, int* __restrict__ cell_start) {
  FastCell c;
  c.valid = 0;
  c.cell = cell;
  if (cell >= g->n_cells) return c;
  c.roi = cells[cell];
  c.tw = c.roi.z - c.roi.x;
  c.th = c.roi.w - c.roi.y;
  if (c.tw < 7 || c.th < 7) {  // skipped cell or nothing testable
    if (lane == 0) {
      cell_count[(size_t)frame * g->n_cells + cell] = 0;
      cell_start[(size_t)frame * g->n_cells + cell] = 0;
    }
    return c;
  }
  int level = 0;
  for (int l = 1; l < g->n_levels; l++)
    if (cell >= g->lv[l].cell_base) level = l;
  c.stride = g->lv[level].stride;
  const uint8_t* src = pyr + (size_t)frame * g->pyr_bytes + g->lv[level].offset + (size_t)(kPad + c.roi.y) * c.stride + kPad + c.roi.x;
  // ROI column 0 is tile column a, a in [1, 4]: tile column 0 is 4-byte aligned in HBM and every word a tested pixel's
  // LOW half depends on is a staged pixel (<= 255), so packed subtraction never borrows into a tested half.
  c.a = (int)((reinterpret_cast<uintptr_t>(src) - 1u) & 3u) + 1;
  c.nw4 = (c.a + c.tw + 3) >> 2;
  c.src_al = src - c.a;
  c.valid = 1;
  return c;
}

// issue the global loads of a cell's tile into registers (lane i takes words i, i+32, ... in row-major order)
__device__ __forceinline__ void fast_prefetch(const FastCell& c, int lane, uint32_t (&v)[kFastPf]) {
  const int sr = 32 / c.nw4, sw = 32 - sr * c.nw4;
  int r = lane / c.nw4, w = lane - r * c.nw4;
#pragma unroll
  for (int i = 0; i < kFastPf; i++) {
    v[i] = 0;
    if (r < c.th) v[i] = __ldg(reinterpret_cast<const uint32_t*>(c.src_al + (size_t)r * c.stride) + w);
    r += sr;
    w += sw;
    if (w >= c.nw4) { w -= c.nw4; r++; }
  }
}

// registers -> shared tile: word (r, 2w) = cols (4w, 4w+1), word (r, 2w+1) = cols (4w+2, 4w+3) of the aligned row;
// words beyond the prefetch depth (cells larger than usual) are loaded here
__device__ __forceinline__ void fast_stage(const FastCell& c, int lane, const uint32_t (&v)[kFastPf], uint32_t* __restrict__ tile,
                                           int wpr) {
  const int sr = 32 / c.nw4, sw = 32 - sr * c.nw4;
  int r = lane / c.nw4, w = lane - r * c.nw4;
#pragma unroll
  for (int i = 0; i < kFastPf; i++) {
    if (r < c.th)
      *reinterpret_cast<uint2*>(tile + r * wpr + 2 * w) = make_uint2(__byte_perm(v[i], 0, 0x4140), __byte_perm(v[i], 0, 0x4342));
    r += sr;
    w += sw;
    if (w >= c.nw4) { w -= c.nw4; r++; }
  }
  while (r < c.th) {
    const uint32_t x = __ldg(reinterpret_cast<const uint32_t*>(c.src_al + (size_t)r * c.stride) + w);
    *reinterpret_cast<uint2*>(tile + r * wpr + 2 * w) = make_uint2(__byte_perm(x, 0, 0x4140), __byte_perm(x, 0, 0x4342));
    r += sr;
    w += sw;
    if (w >= c.nw4) { w -= c.nw4; r++; }
  }
}

struct FastOut {
  const FrameGeomDev* g;
  uint32_t* raw;
  int* raw_counter;
  int* cell_count;
  int* cell_start;
  int* status;
};

// reject, score, NMS and emit of one staged cell (all 32 lanes)
__device__ void fast_cell_compute(const FastCell& fc, int frame, int lane, int ini_th, int min_th, const FastSmemLayout& L,
                                  uint32_t* __restrict__ tile, uint32_t* __restrict__ score, uint16_t* __restrict__ nz,
                                  uint16_t* __restrict__ queue, const FastOut& O) {
  const FrameGeomDev* g = O.g;
  uint32_t* raw = O.raw;
  int* raw_counter = O.raw_counter;
  int* status = O.status;
  const uint32_t lt_mask = (1u << lane) - 1u;
  const int4 roi = fc.roi;
  const int tw = fc.tw, th_ = fc.th, a = fc.a, wpr = L.wpr;
  int* out_count = O.cell_count + (size_t)frame * g->n_cells + fc.cell;
  int* out_start = O.cell_start + (size_t)frame * g->n_cells + fc.cell;
  // tested pixels: tile columns [a+3, a+tw-4], rows [3, th-4]
  const int c_lo = a + 3, c_hi = a + tw - 4;
  const int j0 = c_lo >> 1, j1 = c_hi >> 1;
  const int ipw = j1 - j0 + 1, ih = th_ - 6;
  const int npairs = ipw * ih;
  // lane's walk over the tested pairs in row-major order, 32 pairs per step
  const int step_r = 32 / ipw, step_j = 32 - step_r * ipw;
  const int r_first = 3 + lane / ipw, j_first = lane - (lane / ipw) * ipw;
  // score plane := 0 (only pairs that hold a corner are written; NMS reads one pair / one row around them)
  {
    uint2* z = reinterpret_cast<uint2*>(score - 2);
    for (int i = lane; i < L.plane_words / 2; i += 32) z[i] = make_uint2(0u, 0u);
  }
  __syncwarp();

  const int low_th = min(max(min(ini_th, min_th), 0), 255);
  int nzn = 0;  // pairs with a non-zero stored score, in row-major order: nz[i] = r << 6 | j
  {
    // ---- reject + exact score
    const uint32_t k1 = (uint32_t)(0x7FFF - (low_th + 256)) * 0x10001u;
    const uint32_t kc = (uint32_t)(256 + low_th) * 0x10001u;
    auto score_batch = [&](int nvalid) {
      uint32_t s = 0;
      int r = 0, j = 0;
      if (lane < nvalid) {
        const uint32_t e = queue[lane];
        r = (int)(e >> 8);
        j = (int)(e & 255u);
        const uint32_t mm = fast_score_pair(tile + r * wpr + j, wpr);  // m + 256
        s = __vmaxs2(mm, kc) - kc;                                     // max(m - low_th, 0), no borrow
        if (2 * j < c_lo) s &= 0xFFFF0000u;                            // untested half of an edge pair
        if (2 * j + 1 > c_hi) s &= 0x0000FFFFu;
        if (s != 0u) score[r * wpr + j] = s;
      }
      const uint32_t bal = __ballot_sync(0xffffffffu, s != 0u);
      if (s != 0u) nz[nzn + __popc(bal & lt_mask)] = (uint16_t)((r << 6) | j);
      nzn += __popc(bal);
    };
    int qn = 0;
    int r = r_first, jj = j_first;
    for (int base = 0; base < npairs; base += 32) {
      bool pass = false;
      if (base + lane < npairs) pass = fast_quick_pair(tile + r * wpr + j0 + jj, wpr, k1);
      const uint32_t bal = __ballot_sync(0xffffffffu, pass);
      if (pass) queue[qn + __popc(bal & lt_mask)] = (uint16_t)((r << 8) | (j0 + jj));
      qn += __popc(bal);
      __syncwarp();
      if (qn >= 32) {
        score_batch(32);
        const int rem = qn - 32;
        uint16_t t = 0;
        if (lane < rem) t = queue[32 + lane];
        __syncwarp();
        if (lane < rem) queue[lane] = t;
        qn = rem;
        __syncwarp();
      }
      r += step_r;
      jj += step_j;
      if (jj >= ipw) { jj -= ipw; r++; }
    }
    score_batch(qn);
    __syncwarp();
  }

  // ---- NMS over the non-zero pairs only: strictly greater than all 8 neighbours, non-corners (and untested pixels)
  // count 0.  The plane holds max(m - low_th, 0); for a threshold th >= low_th the values max(m - th, 0) are
  // max(v, kd) - kd with kd = th - low_th, which commutes with the neighbourhood maximum.  Two neighbouring pixels
  // cannot both survive, so a pair keeps at most one: flag bits 13 (low half) / 14 (high half) of the nz entry.
  auto nms_pass = [&](int th) -> int {
    const uint32_t kd = (uint32_t)(th - low_th) * 0x10001u;
    int count = 0;
    for (int base = 0; base < nzn; base += 32) {
      uint32_t flags = 0;
      uint32_t e = 0;
      if (base + lane < nzn) {
        e = nz[base + lane] & 0x1FFFu;
        const uint32_t* p = score + (int)(e >> 6) * wpr + (int)(e & 63u);
        const uint32_t c = __vmaxs2(p[0], kd) - kd;
        if (c != 0u) {
          const uint32_t cl = p[-1], cr = p[1];
          const uint32_t ul = p[-wpr - 1], um = p[-wpr], ur = p[-wpr + 1];
          const uint32_t dl = p[wpr - 1], dm = p[wpr], dr = p[wpr + 1];
          uint32_t n = __vmaxs2(fsr16(cl, p[0]), fsr16(p[0], cr));
          n = __vimax3_s16x2(n, __vimax3_s16x2(fsr16(ul, um), um, fsr16(um, ur)),
                             __vimax3_s16x2(fsr16(dl, dm), dm, fsr16(dm, dr)));
          n = __vmaxs2(n, kd) - kd;
          flags = (c + 0x7FFF7FFFu - n) & 0x80008000u;  // half of c > half of n (all values in [0, 255])
          if ((c & 0xFFFFu) == 0u) flags &= 0x80000000u;  // a zero half is not a corner at this threshold
          if ((c >> 16) == 0u) flags &= 0x00008000u;
        }
        nz[base + lane] = (uint16_t)(e | ((flags & 0x8000u) >> 2) | ((flags >> 31) << 14));
      }
      count += __popc(__ballot_sync(0xffffffffu, flags != 0u));
    }
    return count;
  };
  int count = 0, th_used = low_th;
  for (int attempt = 0; attempt < 2 && count == 0; attempt++) {
    if (attempt == 1 && min_th == ini_th) break;  // same threshold, same (empty) answer (reference: SURVEY F3)
    th_used = min(max(attempt == 0 ? ini_th : min_th, 0), 255);
    // keyPointsInCell.empty() is judged after NMS (src/orbExtractor.cpp:582)
    count = nms_pass(th_used);
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
  // ---- emit in row-major order.  y:12 | x:12 | response:8, coordinates relative to the 16 px border as
  // distributeQuadTree sees them (:586-590); response = m - 1 = stored + low_th - 1.
  // (the x term may be negative before the column is added: sum modulo 2^32, never OR)
  {
    const uint32_t pos_base = ((uint32_t)(roi.y - kBorder) << 20) + ((uint32_t)(roi.x - a - kBorder) << 8) + (uint32_t)(low_th - 1);
    uint32_t* dst = raw + (size_t)frame * g->raw_cap + start;
    int done = 0;
    for (int base = 0; base < nzn; base += 32) {
      uint32_t e = 0;
      if (base + lane < nzn) e = nz[base + lane];
      const bool keep = (e & 0x6000u) != 0u;
      const uint32_t bal = __ballot_sync(0xffffffffu, keep);
      if (keep) {
        const int r = (int)((e >> 6) & 127u), j = (int)(e & 63u);
        const uint32_t v = score[r * wpr + j];
        const bool hi = (e & 0x4000u) != 0u;
        dst[done + __popc(bal & lt_mask)] =
            pos_base + ((uint32_t)r << 20) + ((uint32_t)(2 * j + (hi ? 1 : 0)) << 8) + (hi ? (v >> 16) : (v & 0xFFFFu));
      }
      done += __popc(bal);
    }
  }
  if (lane == 0) { *out_count = count; *out_start = start; }
}

// CPW = consecutive cells a warp works through (the next cell's pixels prefetched into registers while this one
// computes); 1 spreads a single frame's ~1000 cells over all SMs and is what launch_fast uses
template <int CPW>
__global__ void __launch_bounds__(kWarpsPerCta * 32)
k_fast_cells(const FrameGeomDev* __restrict__ g, const int4* __restrict__ cells, const uint8_t* __restrict__ pyr,
             int ini_th, int min_th, int max_tw, int max_th, uint32_t* __restrict__ raw, int* __restrict__ raw_counter,
             int* __restrict__ cell_count, int* __restrict__ cell_start, int* __restrict__ status) {
  pdl_wait();
  extern __shared__ __align__(16) uint8_t smem_raw[];
  const FastSmemLayout L = fast_layout(max_tw, max_th);
  const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
  const int frame = blockIdx.y;
  uint8_t* wbase = smem_raw + (size_t)warp * L.warp_bytes;
  uint32_t* tile = reinterpret_cast<uint32_t*>(wbase) + 2;                    // [th][wpr] pixel pairs
  uint32_t* score = tile + L.plane_words;                                      // [th][wpr] stored scores, pairs
  uint16_t* nz = reinterpret_cast<uint16_t*>(reinterpret_cast<uint32_t*>(wbase) + 2 * L.plane_words);
  uint16_t* queue = nz + L.nz_cap;
  const FastOut O{g, raw, raw_counter, cell_count, cell_start, status};
  // software pipeline over the warp's cells: the next cell's pixels travel HBM -> registers while this one computes
  const int cell0 = (blockIdx.x * kWarpsPerCta + warp) * CPW;
  uint32_t v[kFastPf];
  FastCell cur = fast_cell_setup(g, cells, pyr, cell0, frame, lane, cell_count, cell_start);
  if (cur.valid) fast_prefetch(cur, lane, v);
  for (int k = 0; k < CPW; k++) {
    if (cur.valid) fast_stage(cur, lane, v, tile, L.wpr);
    FastCell nxt;
    nxt.valid = 0;
    if (k + 1 < CPW) {
      nxt = fast_cell_setup(g, cells, pyr, cell0 + k + 1, frame, lane, cell_count, cell_start);
      if (nxt.valid) fast_prefetch(nxt, lane, v);
    }
    if (cur.valid) fast_cell_compute(cur, frame, lane, ini_th, min_th, L, tile, score, nz, queue, O);
    __syncwarp();
    cur = nxt;
  }
}

int launch_fast(const YdExtractor* ex, int batch, cudaStream_t s) {
  YD_CUDA(cudaMemsetAsync(ex->d_raw_counter, 0, sizeof(int) * batch, s));
  // launches of a few frames or more take the CTA-per-cell-block kernel (fast2.cu); a single frame's ~230 4-warp blocks
  // leave SMs idle, the warp-per-cell kernel below spreads its ~800 cells over all of them (640x480, measured: 1 frame
  // 29 us against 45 us, 2-4 frames equal, 16 frames 76 us against 68 us).
  // YDORB_FAST_KERNEL=cells|blocks forces one of them (tests run both and compare).
  {
    const char* force = getenv("YDORB_FAST_KERNEL");  // read per call: tests flip it in-process
    static const int min_ctas = getenv("YDORB_FAST_BLOCKS_MIN_CTAS") ? atoi(getenv("YDORB_FAST_BLOCKS_MIN_CTAS")) : 592;
    bool blocks = (long long)ex->n_fast_blocks * batch >= min_ctas;
    if (force && force[0] == 'c') blocks = false;
    if (force && force[0] == 'b') blocks = true;
    if (blocks && ex->n_fast_blocks > 0) return launch_fast_blocks(ex, batch, s);
  }
  const int ini = ex->params.ini_th_fast;
  const int mn = ex->params.reference_min_th_quirk ? ini : ex->params.min_th_fast;
  const FastSmemLayout L = fast_layout(ex->max_cell_w, ex->max_cell_h);
  const size_t smem = (size_t)L.warp_bytes * kWarpsPerCta;
  // measured on a B200 (bench.py --batch 16/64/256, tools/latency_probe.py): one cell per warp is fastest at every
  // batch size (0.747 vs 0.764 ms per 256 frames, 22 vs 50 us for a single frame); 2 and 4 stay as a tuning knob
  static const int force_cpw = getenv("YDORB_FAST_CPW") ? atoi(getenv("YDORB_FAST_CPW")) : 0;
  const int cpw = force_cpw ? force_cpw : 1;
  auto launch = [&](auto kernel, int c) -> int {
    if (smem > 48 * 1024) YD_CUDA(cudaFuncSetAttribute(kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem));
    const int per_cta = kWarpsPerCta * c;
    dim3 grid((ex->geom.n_cells + per_cta - 1) / per_cta, batch);
    YD_CUDA(launch_chain(kernel, grid, dim3(kWarpsPerCta * 32), smem, s, ex->d_geom, ex->d_cells, ex->d_pyr, ini, mn,
                         ex->max_cell_w, ex->max_cell_h, ex->d_raw, ex->d_raw_counter, ex->d_cell_count, ex->d_cell_start,
                         ex->d_status));
    return YDORB_OK;
  };
  const int rc = cpw == 4 ? launch(k_fast_cells<4>, 4) : cpw == 2 ? launch(k_fast_cells<2>, 2) : launch(k_fast_cells<1>, 1);
  if (rc != YDORB_OK) return rc;
  count_launches(1);
  YD_CUDA(cudaGetLastError());
  return YDORB_OK;
}

}  // namespace ydorb
