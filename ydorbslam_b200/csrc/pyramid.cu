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
#include "tma.cuh"

namespace ydorb {

__host__ __device__ __forceinline__ int reflect101(int i, int n) {
  // |pad| < n is guaranteed by the host (levels narrower than the pad are rejected)
  if (i < 0) i = -i;
  if (i >= n) i = 2 * (n - 1) - i;
  return i;
}

// Inclusive range [lo, hi] of reflect101(i, n) over i in [a, b] (a <= b, |pad| < n).
static void reflect_range(int a, int b, int n, int& lo, int& hi) {
  const int ra = reflect101(a, n), rb = reflect101(b, n);
  lo = (a < 0 && b >= 0) ? 0 : std::min(ra, rb);
  hi = (a <= n - 1 && b >= n - 1) ? n - 1 : std::max(ra, rb);
}

// level 0: copy the input image into the centre of the padded buffer and synthesise the border.
// A thread owns 16 padded bytes of kPy0Rows consecutive rows (one aligned 128-bit store per row).
//   interior role (blockIdx.x < gridDim.x-1): 16-byte groups made only of image pixels -- five aligned 32-bit loads
//     and four funnel shifts (any row stride; needs a 4-byte aligned frame base, `aligned4`);
//   border role (last blockIdx.x): the few groups per row that touch the pad go byte by byte through reflect101.
// Keeping the two roles in different CTAs keeps every warp on one path.
constexpr int kPy0Rows = 8;
__global__ void __launch_bounds__(256) k_pyr_level0(const uint8_t* __restrict__ images, int in_stride, size_t frame_stride,
                                                    int aligned4, int w, int h, int dstride, long long doff,
                                                    long long pyr_bytes, uint8_t* __restrict__ pyr) {
  pdl_wait();
  const int frame = blockIdx.z;
  const int pw = w + 2 * kPad, ph = h + 2 * kPad;
  const uint8_t* src = images + (size_t)frame * frame_stride;
  uint8_t* dst = pyr + (size_t)frame * pyr_bytes + doff;
  // groups g_lo..g_hi are interior: x = 16 g - 19 >= 0 and x + 15 <= w - 1
  const int ngroups = dstride / 16, g_lo = 2, g_hi = aligned4 ? (w + 3) / 16 : 1;
  if (blockIdx.x + 1 < gridDim.x) {
    const int grp = g_lo + blockIdx.x * 64 + (threadIdx.x & 63);
    const int py0 = (blockIdx.y * 4 + (threadIdx.x >> 6)) * kPy0Rows;
    if (grp > g_hi) return;
    const int x = grp * 16 - kPad;
    const long long frame_bytes = (long long)(h - 1) * in_stride + w;
#pragma unroll 2
    for (int i = 0; i < kPy0Rows; i++) {
      const int py = py0 + i;
      if (py >= ph) break;
      const int sy = reflect101(py - kPad, h);
      const long long fo = (long long)sy * in_stride + x;  // byte offset of the first source pixel in the frame
      uint4 out;
      if ((fo & ~3ll) + 20 <= frame_bytes) {
        const uint32_t* wp = reinterpret_cast<const uint32_t*>(src + (fo & ~3ll));
        const uint32_t w0 = __ldg(wp), w1 = __ldg(wp + 1), w2 = __ldg(wp + 2), w3 = __ldg(wp + 3), w4 = __ldg(wp + 4);
        const int sh = 8 * (int)(fo & 3);
        out = make_uint4(__funnelshift_r(w0, w1, sh), __funnelshift_r(w1, w2, sh), __funnelshift_r(w2, w3, sh),
                         __funnelshift_r(w3, w4, sh));
      } else {  // the last few bytes of a frame: stay inside the caller's buffer
        const uint8_t* sp = src + fo;
        uint32_t o[4] = {0, 0, 0, 0};
#pragma unroll
        for (int k = 0; k < 16; k++) o[k >> 2] |= (uint32_t)sp[k] << (8 * (k & 3));
        out = make_uint4(o[0], o[1], o[2], o[3]);
      }
      *reinterpret_cast<uint4*>(dst + (size_t)py * dstride + grp * 16) = out;
    }
  } else {
    // border groups: 0, 1 and g_hi+1 .. ngroups-1; a thread owns one 4-byte word of one group of one row
    const int nb = 2 + (ngroups - 1 - g_hi);
    const int nwords = nb * 4;
    const int rows_per_cta = 4 * kPy0Rows;
    for (int idx = threadIdx.x; idx < nwords * rows_per_cta; idx += 256) {
      const int r = idx / nwords, wd = idx - r * nwords;
      const int py = blockIdx.y * rows_per_cta + r;
      if (py >= ph) break;
      const int b = wd >> 2;
      const int grp = b < 2 ? b : g_hi + 1 + (b - 2);
      const int px4 = grp * 16 + (wd & 3) * 4;
      const uint8_t* srow = src + (size_t)reflect101(py - kPad, h) * in_stride;
      uint32_t o = 0;
#pragma unroll
      for (int k = 0; k < 4; k++) {
        const int px = px4 + k;
        if (px < pw) o |= (uint32_t)srow[reflect101(px - kPad, w)] << (8 * k);
      }
      *reinterpret_cast<uint32_t*>(dst + (size_t)py * dstride + px4) = o;
    }
  }
}

// level l >= 1 from level l-1, separable like OpenCV's own resize (hresize into an int row buffer, then vresize):
// a CTA owns a kPyTW x th tile of the PADDED destination.
//   stage    the source rows/columns the tile needs are copied to shared memory with aligned 128-bit loads, twice:
//            copy A as is and copy B shifted by one byte, so that the pixel pair (S[sx], S[sx+1]) of any sx is ONE
//            aligned 16-bit shared load (from A when sx is even, from B when it is odd).
//   phase A  every staged source row is interpolated horizontally ONCE for the tile's destination columns into the
//            int buffer H as (S[sx0]*a0 + S[sx0+1]*a1) >> 4 -- one LDS.U16 + one DP2A per value; a thread keeps its
//            column's offset and packed weights in registers and walks down the rows.  (Where OpenCV clamps
//            sx1 = sx0 the weight a1 is 0, so reading the pad byte at sx0+1 is harmless.)
//   phase B  a thread produces 4 adjacent padded bytes per row: two LDS.128 (rows r0, r1 of H), 4 x
//            (((b0*H0)>>16) + ((b1*H1)>>16) + 2) >> 2, one aligned 32-bit store.  Pad pixels read the reflected
//            column/row of H (REFLECT_101 of the resized level == resize evaluated at the reflected coordinate),
//            so the border costs no second pass over memory.
// H's column origin is chosen = 1 (mod 4) so that interior 4-byte groups are 16-byte aligned in it; H column
// kPyHS-1 is kept zero and stands in for the stride padding beyond the padded width.
constexpr int kPyTW = 128;             // padded destination columns per CTA
constexpr int kPyHS = kPyTW + 4;       // row stride of H (ints)
// CTA shape (measured per 256 frames of 640x480, all 7 resize launches + level 0): 256 threads x 32 rows 0.324 ms,
// 128 x 24 0.330, 128 x 16 0.316, 128 x 12 0.319, 256 x 16 0.371.  A version that walked several tiles per CTA with the
// next tile's TMA box in flight (double-buffered staging, 4 CTAs/SM instead of 5) was slower: 0.345 ms -- the hardware's
// interleaving of many small CTAs hides the box latency better than a software pipeline inside fewer, larger ones.
#ifndef YD_PYR_THREADS
#define YD_PYR_THREADS 128
#endif
#ifndef YD_PYR_TH
#define YD_PYR_TH 16
#endif
constexpr int kPyThreads = YD_PYR_THREADS;  // 128 or 256
constexpr int kPyStreams = kPyThreads / kPyTW;  // source rows a phase-A pass covers
constexpr int kPyMaxTH = YD_PYR_TH;    // destination rows per CTA (the host shrinks it for large scale factors)

struct PyrArgs {
  int dw, dh, dstride;   // destination level
  int sw, sh, sstride;   // source level
  long long doff;        // byte offset of the destination padded buffer in a frame's pyramid block
  long long sorg;        // byte offset of source pixel (0,0)
  long long pyr_bytes;
  int xb, yb;            // table bases (x then y)
  int th;                // destination rows per tile
  int nsr_cap;           // source rows the shared buffers hold
  int tile_base;         // index (in the int table array) of this level's tile descriptors, see pyramid_tile_tables
};

#ifndef YD_PYR_STAGE_ILP
#define YD_PYR_STAGE_ILP 3
#endif
struct PyrMap {
  CUtensorMap m;  // source level as u8 [frame][row][byte], box = 192 bytes x nsr_cap rows
};
template <int SS>  // SS: row stride (bytes) of the staged source copies; 192 (the reference's scale factors): staged by TMA
__global__ void __launch_bounds__(kPyThreads) k_pyr_resize(const __grid_constant__ PyrMap map, const __grid_constant__ PyrStart st,
                                                    int frame0, const PyrArgs a, const int* __restrict__ tab_ofs, const short2* __restrict__ tab_a,
                                                    uint8_t* __restrict__ pyr) {
  extern __shared__ __align__(128) uint8_t smem[];
  uint8_t* s_a = smem;                                                       // [nsr_cap][SS]  (TMA destination)
  uint8_t* s_b = s_a + a.nsr_cap * SS;                                       // [nsr_cap][SS], s_b[j] = s_a[j+1]
  int* s_h = reinterpret_cast<int*>(s_b + a.nsr_cap * SS);                   // [nsr_cap][kPyHS]
  int4* s_row = reinterpret_cast<int4*>(s_h + a.nsr_cap * kPyHS);            // [kPyMaxTH] {r0*kPyHS, r1*kPyHS, b0, b1}
  uint64_t* bar = reinterpret_cast<uint64_t*>(s_row + kPyMaxTH);
  const int tid = threadIdx.x;
  const int px0 = blockIdx.x * kPyTW, py0 = blockIdx.y * a.th;
  const int pw = a.dw + 2 * kPad, ph = a.dh + 2 * kPad;
  uint8_t* fbase = pyr + (size_t)blockIdx.z * a.pyr_bytes;
  const uint8_t* src0 = fbase + a.sorg;

  // tile geometry precomputed on the host (pyramid_tile_tables): nothing below depends on the thread
  const int4* td = reinterpret_cast<const int4*>(tab_ofs + a.tile_base);
  // Everything up to pdl_wait() reads launch parameters and tables written when the geometry was set, never the
  // predecessor's output: it overlaps the tail of the previous level's launch.
  if (SS == 192 && tid == 0) {
    // one TMA box load of the source rows (192 bytes x nsr_cap rows from the 16-byte aligned column the host chose),
    // issued before anything else; copy B, shifted by one byte, is derived from shared memory once it has landed
    int bx_src, by_src;
    if (st.in_params) {
      bx_src = st.sxa[blockIdx.x];
      by_src = st.sy_lo[blockIdx.y];
    } else {
      bx_src = __ldg(td + 2 * blockIdx.x).w;
      by_src = __ldg(td + 2 * gridDim.x + blockIdx.y).x;
    }
    mbar_init(bar, 1);
    pdl_wait();
    mbar_expect_tx(bar, (uint32_t)(SS * a.nsr_cap));
    tma_load_3d(s_a, &map.m, bar, kPad + bx_src, kPad + by_src, frame0 + (int)blockIdx.z);
  }
  const int4 X0 = __ldg(td + 2 * blockIdx.x), X1 = __ldg(td + 2 * blockIdx.x + 1), Y = __ldg(td + 2 * gridDim.x + blockIdx.y);
  const int sxa = X0.w, nchunk = X1.x;
  const int sy_lo = Y.x, nsr = Y.y;

  if (tid < a.th) {
    const int py = py0 + tid;
    int4 r = make_int4(0, 0, 0, 0);
    if (py < ph) {
      const int dy = reflect101(py - kPad, a.dh);
      const int sy0 = tab_ofs[a.yb + dy];
      const short2 b = tab_a[a.yb + dy];
      r = make_int4((sy0 - sy_lo) * kPyHS, (min(sy0 + 1, a.sh - 1) - sy_lo) * kPyHS, b.x, b.y);
    }
    s_row[tid] = r;
  }
  if (SS != 192 || tid != 0) pdl_wait();
  // ---- stage
  if (SS == 192) {
    __syncthreads();  // the barrier thread 0 initialised is visible
    mbar_wait(bar, 0);
    for (int c = tid & 15; c < nchunk; c += 16)
      for (int r = tid >> 4; r < nsr; r += kPyThreads / 16) {
        const uint4 v = *reinterpret_cast<const uint4*>(s_a + r * SS + 16 * c);
        const uint32_t nx = (c + 1 < nchunk) ? *reinterpret_cast<const uint32_t*>(s_a + r * SS + 16 * c + 16) : 0u;
        *reinterpret_cast<uint4*>(s_b + r * SS + 16 * c) =
            make_uint4(__funnelshift_r(v.x, v.y, 8), __funnelshift_r(v.y, v.z, 8), __funnelshift_r(v.z, v.w, 8),
                       __funnelshift_r(v.w, nx, 8));
      }
  } else {
    constexpr int kStageIlp = YD_PYR_STAGE_ILP;  // rows whose loads are in flight together per thread
    for (int c = tid & 15; c < nchunk; c += 16) {
      for (int r0 = tid >> 4; r0 < nsr; r0 += (kPyThreads / 16) * kStageIlp) {
        uint4 v[kStageIlp];
        uint32_t nx[kStageIlp];
#pragma unroll
        for (int k = 0; k < kStageIlp; k++) {
          const int r = r0 + (kPyThreads / 16) * k;
          if (r < nsr) {
            const uint8_t* gp = src0 + (long long)(sy_lo + r) * a.sstride + sxa + 16 * c;
            v[k] = *reinterpret_cast<const uint4*>(gp);
            nx[k] = (c + 1 < nchunk) ? *reinterpret_cast<const uint32_t*>(gp + 16) : 0u;
          }
        }
#pragma unroll
        for (int k = 0; k < kStageIlp; k++) {
          const int r = r0 + (kPyThreads / 16) * k;
          if (r < nsr) {
            *reinterpret_cast<uint4*>(s_a + r * SS + 16 * c) = v[k];
            *reinterpret_cast<uint4*>(s_b + r * SS + 16 * c) =
                make_uint4(__funnelshift_r(v[k].x, v[k].y, 8), __funnelshift_r(v[k].y, v[k].z, 8),
                           __funnelshift_r(v[k].z, v[k].w, 8), __funnelshift_r(v[k].w, nx[k], 8));
          }
        }
      }
    }
  }
  __syncthreads();
  // ---- phase A: H column c = PADDED destination column px0 + c, evaluated at its reflected interior coordinate
  // (REFLECT_101 of the resized level == the resize evaluated at the reflected coordinate), 0 beyond the padded width
  {
    const int c = tid & (kPyTW - 1);
    const int px = px0 + c;
    int* hp = s_h + (tid >> 7) * kPyHS + c;
    const int n = (nsr - (tid >> 7) + kPyStreams - 1) / kPyStreams;  // rows tid>>7, +kPyStreams, ...
    if (px < pw) {
      const int dx = reflect101(px - kPad, a.dw);
      const int o0 = tab_ofs[a.xb + dx] - sxa;
      const short2 wa = tab_a[a.xb + dx];
      const uint32_t wpk = ((uint32_t)(uint16_t)wa.y << 16) | (uint16_t)wa.x;
      const uint8_t* sp = ((o0 & 1) ? s_b : s_a) + (o0 & ~1) + (tid >> 7) * SS;
#pragma unroll 4
      for (int i = 0; i < n; i++)
        hp[i * kPyStreams * kPyHS] = (int)(__dp2a_lo(wpk, (uint32_t) * reinterpret_cast<const uint16_t*>(sp + i * kPyStreams * SS), 0u) >> 4);
    } else {
      for (int i = 0; i < n; i++) hp[i * kPyStreams * kPyHS] = 0;
    }
  }
  __syncthreads();
  // ---- phase B: 4 adjacent padded bytes per thread and row, two aligned LDS.128 of H each
  const int c0 = 4 * (tid & 31);
  const int px4 = px0 + c0;
  if (px4 >= a.dstride) return;
  uint8_t* dcol = fbase + a.doff + px4;
  for (int i = tid >> 5; i < a.th; i += kPyThreads / 32) {
    const int py = py0 + i;
    if (py >= ph) break;
    const int4 r = s_row[i];
    const int4 u = *reinterpret_cast<const int4*>(&s_h[r.x + c0]);
    const int4 w = *reinterpret_cast<const int4*>(&s_h[r.y + c0]);
    const int h0[4] = {u.x, u.y, u.z, u.w}, h1[4] = {w.x, w.y, w.z, w.w};
    uint32_t v[4];
#pragma unroll
    for (int k = 0; k < 4; k++)  // the "+2" rides on the first product as 2 << 16
      v[k] = (uint32_t)((((r.z * h0[k] + 0x20000) >> 16) + ((r.w * h1[k]) >> 16)) >> 2);
    const uint32_t out = __byte_perm(__byte_perm(v[0], v[1], 0x0040), __byte_perm(v[2], v[3], 0x0040), 0x5410);
    *reinterpret_cast<uint32_t*>(dcol + (size_t)py * a.dstride) = out;
  }
}

// tile rows and shared-memory sizing of one resize launch (any scale factor up to kPyMaxScale)
struct ResizePlan {
  int th, nsr_cap, ss;
  size_t smem;
};
static bool plan_resize(const LevelGeom& sv, const LevelGeom& lv, ResizePlan* p) {
  const double sy = (double)sv.h / (double)lv.h, sx = (double)sv.w / (double)lv.w;
  int th = kPyMaxTH;
  while (th > 1 && (int)((th - 1) * sy) + 4 > 48) th--;
  p->th = th;
  p->nsr_cap = (int)((th - 1) * sy) + 4;
  // bytes staged per row: <= 15 (alignment) + floor(127*sx) + 3 source columns, in whole 16-byte chunks
  const int need = ((int)((kPyTW - 1) * sx) + 19 + 15) / 16 * 16;
  p->ss = need <= 192 ? 192 : 512;
  p->smem = (size_t)p->nsr_cap * kPyHS * 4 + sizeof(int4) * kPyMaxTH + 2 * (size_t)p->nsr_cap * p->ss + 16;
  return need <= 512;
}

// Appends the per-tile descriptors of every resize launch to the int table array and records their base per level.
//   per tile column bx: {lo, hi, org, sxa} {nchunk, 0, 0, 0}   then per tile row by: {sy_lo, nsr, 0, 0}
// lo..hi = destination columns (after REFLECT_101) the tile needs, org = H column origin (= 1 mod 4),
// sxa = first staged source column (16-byte aligned address), nchunk = 16-byte chunks staged per source row,
// sy_lo/nsr = staged source rows.
int pyramid_tile_tables(FrameGeomDev* g, std::vector<int>* tab) {
  for (int l = 1; l < g->n_levels; l++) {
    LevelGeom& lv = g->lv[l];
    const LevelGeom& sv = g->lv[l - 1];
    ResizePlan p;
    YD_REQUIRE(plan_resize(sv, lv, &p), "scale factor too large for the resize kernel (max ~3.5)");
    const int pw = lv.w + 2 * kPad, ph = lv.h + 2 * kPad;
    const int nbx = (lv.stride + kPyTW - 1) / kPyTW, nby = (ph + p.th - 1) / p.th;
    const int mis = (int)(level_origin(sv) & 15);
    tab->resize((tab->size() + 3) / 4 * 4);
    lv.tile_base = (int)tab->size();
    const int* xo = tab->data() + lv.tab_base;
    std::vector<int> d;
    for (int bx = 0; bx < nbx; bx++) {
      const int px0 = bx * kPyTW;
      int lo, hi;
      reflect_range(px0 - kPad, std::max(std::min(px0 + kPyTW, pw) - 1, px0) - kPad, lv.w, lo, hi);
      const int org = lo - ((lo - 1) & 3);
      const int sx_lo = xo[lo], sx_hi = xo[hi] + 1;
      const int sxa = sx_lo - ((sx_lo + mis) & 15);
      const int nchunk = (sx_hi - sxa) / 16 + 1;
      YD_REQUIRE(nchunk * 16 <= p.ss, "internal: resize tile exceeds its shared buffers");
      const int v[8] = {lo, hi, org, sxa, nchunk, 0, 0, 0};
      d.insert(d.end(), v, v + 8);
    }
    for (int by = 0; by < nby; by++) {
      const int py0 = by * p.th;
      int rlo, rhi;
      reflect_range(py0 - kPad, std::min(py0 + p.th, ph) - 1 - kPad, lv.h, rlo, rhi);
      const int sy_lo = xo[lv.w + rlo];
      const int nsr = std::min(xo[lv.w + rhi] + 1, sv.h - 1) - sy_lo + 1;
      YD_REQUIRE(nsr <= p.nsr_cap, "internal: resize tile exceeds its shared buffers");
      const int v[4] = {sy_lo, nsr, 0, 0};
      d.insert(d.end(), v, v + 4);
    }
    tab->insert(tab->end(), d.begin(), d.end());  // (xo is dead from here: the vector may have moved)
  }
  return YDORB_OK;
}

// per destination level: the tile start coordinates of pyramid_tile_tables, in the form the kernel takes as a parameter
void pyramid_tile_starts(const FrameGeomDev& g, const std::vector<int>& tab, PyrStart* out) {
  for (int l = 1; l < g.n_levels; l++) {
    const LevelGeom& lv = g.lv[l];
    ResizePlan p;
    PyrStart& st = out[l];
    st.in_params = 0;
    if (!plan_resize(g.lv[l - 1], lv, &p)) continue;
    const int nbx = (lv.stride + kPyTW - 1) / kPyTW, nby = (lv.h + 2 * kPad + p.th - 1) / p.th;
    if (nbx > kPyStartX || nby > kPyStartY) continue;
    const int* td = tab.data() + lv.tile_base;
    for (int bx = 0; bx < nbx; bx++) st.sxa[bx] = (short)td[8 * bx + 3];
    for (int by = 0; by < nby; by++) st.sy_lo[by] = (short)td[8 * nbx + 4 * by];
    st.in_params = 1;
  }
}

// box of the TMA map that stages the SOURCE level of destination level l (0 x 0: that launch does not use TMA)
void pyramid_box(const FrameGeomDev& g, int l, int* box_w, int* box_h) {
  ResizePlan p;
  *box_w = *box_h = 0;
  if (l < 1 || !plan_resize(g.lv[l - 1], g.lv[l], &p) || p.ss != 192) return;
  *box_w = 192;
  *box_h = p.nsr_cap;
}

int launch_pyramid(const YdExtractor* ex, const uint8_t* d_images, int stride, size_t frame_stride, int batch,
                   cudaStream_t s) {
  const FrameGeomDev& g = ex->geom;
  const int aligned4 = ((uintptr_t)d_images % 4 == 0 && frame_stride % 4 == 0) ? 1 : 0;
  YD_REQUIRE(g.pyr_bytes % 16 == 0, "pyramid block must be a multiple of 16 bytes");
  for (int l = 0; l < g.n_levels; l++) {
    const LevelGeom& lv = g.lv[l];
    if (l == 0) {
      const int n_int = std::max(aligned4 ? (lv.w + 3) / 16 - 1 : 0, 0);  // interior 16-byte groups per row
      dim3 grid((n_int + 63) / 64 + 1, (lv.h + 2 * kPad + 4 * kPy0Rows - 1) / (4 * kPy0Rows), batch);
      YD_CUDA(launch_chain(k_pyr_level0, grid, dim3(256), 0, s, d_images, stride, frame_stride, aligned4, lv.w, lv.h, lv.stride,
                           lv.offset, g.pyr_bytes, ex->d_pyr));
    } else {
      const LevelGeom& sv = g.lv[l - 1];
      ResizePlan p;
      YD_REQUIRE(plan_resize(sv, lv, &p), "scale factor too large for the resize kernel (max ~3.5)");
      YD_REQUIRE(sv.stride % 16 == 0 && lv.offset % 16 == 0 && sv.offset % 16 == 0, "pyramid levels must be 16-byte aligned");
      PyrArgs a;
      a.dw = lv.w; a.dh = lv.h; a.dstride = lv.stride;
      a.sw = sv.w; a.sh = sv.h; a.sstride = sv.stride;
      a.doff = lv.offset; a.sorg = level_origin(sv); a.pyr_bytes = g.pyr_bytes;
      a.xb = lv.tab_base; a.yb = lv.tab_base + lv.w;
      a.th = p.th; a.nsr_cap = p.nsr_cap; a.tile_base = lv.tile_base;
      dim3 grid((lv.stride + kPyTW - 1) / kPyTW, (lv.h + 2 * kPad + p.th - 1) / p.th, batch);
      const short2* ta = reinterpret_cast<const short2*>(ex->d_tab_a);
      if (p.smem > 48 * 1024)  // only reached with scale factors well above the reference's 1.2
        YD_CUDA(cudaFuncSetAttribute(k_pyr_resize<512>, cudaFuncAttributeMaxDynamicSharedMemorySize, 100 * 1024));
      PyrMap map;
      map.m = ex->tmap_pyr[l];
      const PyrStart& st = ex->pyr_start[l];
      const int frame0 = (int)((ex->d_pyr - ex->pyr_base) / g.pyr_bytes);
      if (p.ss == 192)
        YD_CUDA(launch_chain(k_pyr_resize<192>, grid, dim3(kPyThreads), p.smem, s, map, st, frame0, a, ex->d_tab_ofs, ta, ex->d_pyr));
      else
        YD_CUDA(launch_chain(k_pyr_resize<512>, grid, dim3(kPyThreads), p.smem, s, map, st, frame0, a, ex->d_tab_ofs, ta, ex->d_pyr));
    }
  }
  count_launches(g.n_levels);
  YD_CUDA(cudaGetLastError());
  return YDORB_OK;
}

}  // namespace ydorb
