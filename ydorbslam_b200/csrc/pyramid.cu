// E2  OrbExtractor::computePyramid  (src/orbExtractor.cpp:605-621)
//
// Level l is cv::resize(level l-1, INTER_LINEAR) written into the centre of a padded buffer, then
// cv::copyMakeBorder(REFLECT_101, 19 px).  Arithmetic is OpenCV's 11-bit fixed point bilinear
// (SURVEY Appendix A1): per-axis (offset, a0, a1) tables are built on the host exactly as OpenCV builds them,
// the kernel does   dst = (((b0*(R0>>4))>>16) + ((b1*(R1>>4))>>16) + 2) >> 2   in int32.
//
// Layout: each thread produces 4 horizontally adjacent bytes of the PADDED destination (one aligned 32-bit
// store); border pixels are produced by evaluating the reflected interior coordinate, so the pad never needs a
// second pass over memory.  HBM traffic per level = read of level l-1 (L2/L1 absorbs the 4-tap reuse) + one write.
#include <algorithm>

#include "kernels.cuh"

namespace ydorb {

__device__ __forceinline__ int reflect101(int i, int n) {
  // |pad| < n is guaranteed by the host (levels narrower than the pad are rejected)
  if (i < 0) i = -i;
  if (i >= n) i = 2 * (n - 1) - i;
  return i;
}

// Inclusive range [lo, hi] of reflect101(i, n) over i in [a, b] (a <= b, |pad| < n).
__device__ __forceinline__ void reflect_range(int a, int b, int n, int& lo, int& hi) {
  const int ra = reflect101(a, n), rb = reflect101(b, n);
  lo = (a < 0 && b >= 0) ? 0 : min(ra, rb);
  hi = (a <= n - 1 && b >= n - 1) ? n - 1 : max(ra, rb);
}

// level 0: copy the input image into the centre of the padded buffer and synthesise the border.
// A thread owns 4 padded bytes (one aligned store).  When all four are interior pixels and the source rows are
// 4-byte aligned (`aligned4`), they come from two aligned 32-bit loads and a funnel shift; otherwise byte by byte.
__global__ void __launch_bounds__(256) k_pyr_level0(const FrameGeomDev* __restrict__ g, const uint8_t* __restrict__ images,
                                                    int in_stride, size_t frame_stride, int aligned4,
                                                    uint8_t* __restrict__ pyr) {
  const LevelGeom lv = g->lv[0];
  const int frame = blockIdx.z;
  const int py = blockIdx.y;                                    // padded row
  const int px4 = (blockIdx.x * blockDim.x + threadIdx.x) * 4;  // padded column of the first of 4 bytes
  const int pw = lv.w + 2 * kPad;
  if (px4 >= lv.stride) return;
  const uint8_t* src = images + (size_t)frame * frame_stride;
  const int sy = reflect101(py - kPad, lv.h);
  const uint8_t* srow = src + (size_t)sy * in_stride;
  const int x = px4 - kPad;
  uint32_t out = 0;
  if (aligned4 && x >= 0 && ((x >> 2) + 2) * 4 <= lv.w) {
    const uint32_t* w = reinterpret_cast<const uint32_t*>(srow) + (x >> 2);
    out = __funnelshift_r(__ldg(w), __ldg(w + 1), 8 * (x & 3));
  } else {
#pragma unroll
    for (int k = 0; k < 4; k++) {
      const int px = px4 + k;
      uint32_t v = 0;
      if (px < pw) v = srow[reflect101(px - kPad, lv.w)];
      out |= v << (8 * k);
    }
  }
  uint8_t* drow = pyr + (size_t)frame * g->pyr_bytes + lv.offset + (size_t)py * lv.stride;
  *reinterpret_cast<uint32_t*>(drow + px4) = out;
}

// level l >= 1 from level l-1, separable like OpenCV's own resize (hresize into an int row buffer, then vresize):
// a CTA owns a kPyTW x th tile of the PADDED destination.
//   phase A  every source row the tile needs is interpolated horizontally ONCE for the tile's destination columns
//            into shared memory as (S[sx0]*a0 + S[sx1]*a1) >> 4 (the value the vertical pass consumes); a thread
//            keeps its column's (sx0, sx1, a0, a1) in registers and walks down the source rows.
//   phase B  a thread produces 4 adjacent padded bytes per row: two LDS.128 (rows r0, r1 of the buffer), 4 x
//            (((b0*H0)>>16) + ((b1*H1)>>16) + 2) >> 2, one aligned 32-bit store.  Pad pixels read the reflected
//            column/row of the same buffer (REFLECT_101 of the resized level == resize evaluated at the reflected
//            coordinate), so the border costs no second pass over memory.
// The buffer's column origin is chosen = 1 (mod 4) so that interior 4-byte groups are 16-byte aligned in it.
constexpr int kPyTW = 128;             // padded destination columns per CTA
constexpr int kPyHS = kPyTW + 4;       // row stride of the horizontal-pass buffer (ints)
constexpr int kPyMaxSrcRows = 48;      // source rows held per tile; the host picks th so that they fit
constexpr int kPyMaxTH = 32;

__global__ void __launch_bounds__(256) k_pyr_resize(const FrameGeomDev* __restrict__ g, int level, int th,
                                                    const int* __restrict__ tab_ofs, const short2* __restrict__ tab_a,
                                                    uint8_t* __restrict__ pyr) {
  __shared__ __align__(16) int s_h[kPyMaxSrcRows * kPyHS];
  __shared__ int4 s_row[kPyMaxTH];  // per tile row: {r0*kPyHS, r1*kPyHS, b0, b1}
  const LevelGeom lv = g->lv[level];
  const LevelGeom sv = g->lv[level - 1];
  const int tid = threadIdx.x;
  const int frame = blockIdx.z;
  const int px0 = blockIdx.x * kPyTW, py0 = blockIdx.y * th;
  const int pw = lv.w + 2 * kPad, ph = lv.h + 2 * kPad;
  uint8_t* fbase = pyr + (size_t)frame * g->pyr_bytes;
  const uint8_t* s = fbase + level_origin(sv);
  const int xb = lv.tab_base, yb = lv.tab_base + lv.w;

  int lo, hi, rlo, rhi;
  reflect_range(px0 - kPad, max(min(px0 + kPyTW, pw) - 1, px0) - kPad, lv.w, lo, hi);
  reflect_range(py0 - kPad, min(py0 + th, ph) - 1 - kPad, lv.h, rlo, rhi);
  const int org = lo - ((lo - 1) & 3);  // buffer column 0 <-> destination column org, org = 1 (mod 4)
  const int sy_lo = tab_ofs[yb + rlo];
  const int nsr = min(tab_ofs[yb + rhi] + 1, sv.h - 1) - sy_lo + 1;

  if (tid < th) {
    const int py = py0 + tid;
    int4 r = make_int4(0, 0, 0, 0);
    if (py < ph) {
      const int dy = reflect101(py - kPad, lv.h);
      const int sy0 = tab_ofs[yb + dy];
      const short2 b = tab_a[yb + dy];
      r = make_int4((sy0 - sy_lo) * kPyHS, (min(sy0 + 1, sv.h - 1) - sy_lo) * kPyHS, b.x, b.y);
    }
    s_row[tid] = r;
  }
  // ---- phase A
  {
    const int c = tid & (kPyTW - 1);
    const int dx = org + c;
    if (dx >= lo && dx <= hi) {
      const int sx0 = tab_ofs[xb + dx];
      const short2 a = tab_a[xb + dx];
      const int sx1 = min(sx0 + 1, sv.w - 1);
      const uint8_t* p = s + (size_t)sy_lo * sv.stride;
#pragma unroll 4
      for (int sr = tid >> 7; sr < nsr; sr += 2) {
        const uint8_t* q = p + (size_t)sr * sv.stride;
        s_h[sr * kPyHS + c] = (q[sx0] * a.x + q[sx1] * a.y) >> 4;
      }
    }
    const int c2 = kPyTW + (tid & 3), dx2 = org + c2;  // the up to 3 columns the alignment shift adds
    if (dx2 <= hi) {
      const int sx0 = tab_ofs[xb + dx2];
      const short2 a = tab_a[xb + dx2];
      const int sx1 = min(sx0 + 1, sv.w - 1);
      for (int sr = tid >> 2; sr < nsr; sr += 64) {
        const uint8_t* q = s + (size_t)(sy_lo + sr) * sv.stride;
        s_h[sr * kPyHS + c2] = (q[sx0] * a.x + q[sx1] * a.y) >> 4;
      }
    }
  }
  __syncthreads();
  // ---- phase B
  const int px4 = px0 + 4 * (tid & 31);
  if (px4 >= lv.stride) return;
  const bool straight = px4 - kPad >= 0 && px4 + 3 - kPad <= lv.w - 1;
  int c[4];
#pragma unroll
  for (int k = 0; k < 4; k++) c[k] = (px4 + k < pw) ? reflect101(px4 + k - kPad, lv.w) - org : -1;
  uint8_t* dcol = fbase + lv.offset + px4;
  for (int i = tid >> 5; i < th; i += 8) {
    const int py = py0 + i;
    if (py >= ph) break;
    const int4 r = s_row[i];
    int h0[4], h1[4];
    if (straight) {
      const int4 u = *reinterpret_cast<const int4*>(&s_h[r.x + c[0]]);
      const int4 v = *reinterpret_cast<const int4*>(&s_h[r.y + c[0]]);
      h0[0] = u.x; h0[1] = u.y; h0[2] = u.z; h0[3] = u.w;
      h1[0] = v.x; h1[1] = v.y; h1[2] = v.z; h1[3] = v.w;
    } else {
#pragma unroll
      for (int k = 0; k < 4; k++) {
        h0[k] = c[k] >= 0 ? s_h[r.x + c[k]] : 0;
        h1[k] = c[k] >= 0 ? s_h[r.y + c[k]] : 0;
      }
    }
    uint32_t out = 0;
#pragma unroll
    for (int k = 0; k < 4; k++) {
      uint32_t v = (uint32_t)((((r.z * h0[k]) >> 16) + ((r.w * h1[k]) >> 16) + 2) >> 2);
      if (c[k] < 0) v = 0;
      out |= v << (8 * k);
    }
    *reinterpret_cast<uint32_t*>(dcol + (size_t)py * lv.stride) = out;
  }
}

// rows of the destination tile such that the source rows it needs fit the shared buffer (any scale factor)
static int resize_tile_rows(const LevelGeom& sv, const LevelGeom& lv) {
  const double scale = (double)sv.h / (double)lv.h;
  int th = (int)((kPyMaxSrcRows - 4) / scale);
  return std::max(1, std::min(th, kPyMaxTH));
}

int launch_pyramid(const YdExtractor* ex, const uint8_t* d_images, int stride, size_t frame_stride, int batch,
                   cudaStream_t s) {
  const FrameGeomDev& g = ex->geom;
  const int aligned4 = ((uintptr_t)d_images % 4 == 0 && stride % 4 == 0 && frame_stride % 4 == 0) ? 1 : 0;
  for (int l = 0; l < g.n_levels; l++) {
    const LevelGeom& lv = g.lv[l];
    if (l == 0) {
      dim3 grid((lv.stride / 4 + 255) / 256, lv.h + 2 * kPad, batch);
      k_pyr_level0<<<grid, 256, 0, s>>>(ex->d_geom, d_images, stride, frame_stride, aligned4, ex->d_pyr);
    } else {
      const int th = resize_tile_rows(g.lv[l - 1], lv);
      dim3 grid((lv.stride + kPyTW - 1) / kPyTW, (lv.h + 2 * kPad + th - 1) / th, batch);
      k_pyr_resize<<<grid, 256, 0, s>>>(ex->d_geom, l, th, ex->d_tab_ofs, reinterpret_cast<const short2*>(ex->d_tab_a),
                                        ex->d_pyr);
    }
  }
  count_launches(g.n_levels);
  YD_CUDA(cudaGetLastError());
  return YDORB_OK;
}

}  // namespace ydorb
