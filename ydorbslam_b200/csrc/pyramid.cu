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
#include "kernels.cuh"

namespace ydorb {

__device__ __forceinline__ int reflect101(int i, int n) {
  // |pad| < n is guaranteed by the host (levels narrower than the pad are rejected)
  if (i < 0) i = -i;
  if (i >= n) i = 2 * (n - 1) - i;
  return i;
}

// level 0: copy the input image into the centre of the padded buffer and synthesise the border
__global__ void __launch_bounds__(256) k_pyr_level0(const FrameGeomDev* __restrict__ g, const uint8_t* __restrict__ images,
                                                    int in_stride, size_t frame_stride, uint8_t* __restrict__ pyr) {
  const LevelGeom lv = g->lv[0];
  const int frame = blockIdx.z;
  const int py = blockIdx.y;                                    // padded row
  const int px4 = (blockIdx.x * blockDim.x + threadIdx.x) * 4;  // padded column of the first of 4 bytes
  const int pw = lv.w + 2 * kPad;
  if (px4 >= lv.stride) return;
  const uint8_t* src = images + (size_t)frame * frame_stride;
  const int sy = reflect101(py - kPad, lv.h);
  const uint8_t* srow = src + (size_t)sy * in_stride;
  uint32_t out = 0;
#pragma unroll
  for (int k = 0; k < 4; k++) {
    const int px = px4 + k;
    uint32_t v = 0;
    if (px < pw) v = srow[reflect101(px - kPad, lv.w)];
    out |= v << (8 * k);
  }
  uint8_t* drow = pyr + (size_t)frame * g->pyr_bytes + lv.offset + (size_t)py * lv.stride;
  *reinterpret_cast<uint32_t*>(drow + px4) = out;
}

// level l >= 1 from level l-1
__global__ void __launch_bounds__(256) k_pyr_resize(const FrameGeomDev* __restrict__ g, int level,
                                                    const int* __restrict__ tab_ofs, const short2* __restrict__ tab_a,
                                                    uint8_t* __restrict__ pyr) {
  const LevelGeom lv = g->lv[level];
  const LevelGeom sv = g->lv[level - 1];
  const int frame = blockIdx.z;
  const int py = blockIdx.y;
  const int px4 = (blockIdx.x * blockDim.x + threadIdx.x) * 4;
  const int pw = lv.w + 2 * kPad;
  if (px4 >= lv.stride) return;
  uint8_t* fbase = pyr + (size_t)frame * g->pyr_bytes;
  const uint8_t* s = fbase + level_origin(sv);
  const int dy = reflect101(py - kPad, lv.h);
  const int ytab = lv.tab_base + lv.w + dy;
  const int sy0 = tab_ofs[ytab];
  const short2 yb = tab_a[ytab];
  const int sy1 = min(sy0 + 1, sv.h - 1);
  const uint8_t* r0 = s + (size_t)sy0 * sv.stride;
  const uint8_t* r1 = s + (size_t)sy1 * sv.stride;
  uint32_t out = 0;
#pragma unroll
  for (int k = 0; k < 4; k++) {
    const int px = px4 + k;
    uint32_t v = 0;
    if (px < pw) {
      const int dx = reflect101(px - kPad, lv.w);
      const int sx0 = tab_ofs[lv.tab_base + dx];
      const short2 xa = tab_a[lv.tab_base + dx];
      const int sx1 = min(sx0 + 1, sv.w - 1);
      const int h0 = r0[sx0] * xa.x + r0[sx1] * xa.y;
      const int h1 = r1[sx0] * xa.x + r1[sx1] * xa.y;
      v = (uint32_t)((((yb.x * (h0 >> 4)) >> 16) + ((yb.y * (h1 >> 4)) >> 16) + 2) >> 2);
    }
    out |= v << (8 * k);
  }
  uint8_t* drow = fbase + lv.offset + (size_t)py * lv.stride;
  *reinterpret_cast<uint32_t*>(drow + px4) = out;
}

int launch_pyramid(const YdExtractor* ex, const uint8_t* d_images, int stride, size_t frame_stride, int batch,
                   cudaStream_t s) {
  const FrameGeomDev& g = ex->geom;
  for (int l = 0; l < g.n_levels; l++) {
    const LevelGeom& lv = g.lv[l];
    dim3 block(256);
    dim3 grid((lv.stride / 4 + 255) / 256, lv.h + 2 * kPad, batch);
    if (l == 0)
      k_pyr_level0<<<grid, block, 0, s>>>(ex->d_geom, d_images, stride, frame_stride, ex->d_pyr);
    else
      k_pyr_resize<<<grid, block, 0, s>>>(ex->d_geom, l, ex->d_tab_ofs, reinterpret_cast<const short2*>(ex->d_tab_a),
                                          ex->d_pyr);
  }
  count_launches(g.n_levels);
  YD_CUDA(cudaGetLastError());
  return YDORB_OK;
}

}  // namespace ydorb
