// the pre-/post-processing steps around extractAndCompute, kept on the device so that a frame does
// not bounce through host memory between them:
//   cv::cvtColor(.., COLOR_{RGB,BGR,RGBA,BGRA}2GRAY)  src/tracking.cpp:73-89, :97-107        -> k_cvt_gray
//   Frame::undistortKeyPoints / computeImageBounds     src/frame.cpp:193-213, :224-248        -> k_undistort
//   depthImage.convertTo(CV_32F, f) + Frame::computeStereoFromRGBD  src/tracking.cpp:109-111, src/frame.cpp:214-223
//                                                                                              -> k_stereo_from_rgbd
// All three restate OpenCV arithmetic that was pinned against cv2 4.13.0 (tests/test_oracle_preprocess.py).
#include <algorithm>

#include "matcher.cuh"

namespace ydorb {

// OpenCV's 8-bit RGB2Gray (color_rgb.simd.hpp, RGB2Gray<uchar>): 15-bit coefficients, round to nearest.
__device__ __forceinline__ uint32_t gray15(uint32_t r, uint32_t g, uint32_t b) {
  return (r * 9798u + g * 19235u + b * 3735u + (1u << 14)) >> 15;
}

// thread = 4 consecutive destination pixels of one row (one 32-bit store); the source bytes of those pixels are
// 12 (3 channels) or 16 (4 channels) consecutive bytes, fetched as aligned 32-bit words when the row allows it.
template <int CH>
__global__ void __launch_bounds__(256) k_cvt_gray(const uint8_t* __restrict__ src, int width, int height, int src_stride,
                                                  size_t src_frame_stride, int swap_rb, uint8_t* __restrict__ dst, int dst_stride,
                                                  size_t dst_frame_stride) {
  const int x4 = (blockIdx.x * blockDim.x + threadIdx.x) * 4, y = blockIdx.y, b = blockIdx.z;
  if (x4 >= width) return;
  const uint8_t* s = src + (size_t)b * src_frame_stride + (size_t)y * src_stride + (size_t)x4 * CH;
  uint8_t* d = dst + (size_t)b * dst_frame_stride + (size_t)y * dst_stride + x4;
  const int n = min(4, width - x4);
  uint32_t px[4][3];
  if (n == 4 && (((uintptr_t)s) & 3) == 0) {
    uint32_t w[CH];
#pragma unroll
    for (int i = 0; i < CH; i++) w[i] = __ldg(reinterpret_cast<const uint32_t*>(s) + i);
#pragma unroll
    for (int p = 0; p < 4; p++)
#pragma unroll
      for (int c = 0; c < 3; c++) {
        const int byte = p * CH + c;
        px[p][c] = (w[byte >> 2] >> (8 * (byte & 3))) & 0xffu;
      }
  } else {
#pragma unroll
    for (int p = 0; p < 4; p++)
#pragma unroll
      for (int c = 0; c < 3; c++) px[p][c] = p < n ? s[p * CH + c] : 0u;
  }
  uint32_t out = 0;
#pragma unroll
  for (int p = 0; p < 4; p++) {
    const uint32_t r = swap_rb ? px[p][2] : px[p][0], bl = swap_rb ? px[p][0] : px[p][2];
    out |= gray15(r, px[p][1], bl) << (8 * p);
  }
  if (n == 4 && (((uintptr_t)d) & 3) == 0) {
    *reinterpret_cast<uint32_t*>(d) = out;
  } else {
    for (int p = 0; p < n; p++) d[p] = (uint8_t)(out >> (8 * p));
  }
}

struct UndistortArgs {
  double fx, fy, cx, cy, ifx, ify;
  double k[12];
  double rr[9];  // P(3x3) * R with R = I
};

// cvUndistortPointsInternal with TermCriteria(MAX_ITER, 5, 0.01) (calib3d undistort.dispatch.cpp): every double
// operation rounded separately, in OpenCV's expression order (the CPU build has no FMA contraction).
__device__ __forceinline__ void undistort_one(const UndistortArgs& a, float fx_in, float fy_in, float* ox, float* oy) {
  const double u = (double)fx_in, v = (double)fy_in;
  double x = __dmul_rn(__dsub_rn(u, a.cx), a.ifx), y = __dmul_rn(__dsub_rn(v, a.cy), a.ify);
  const double x0 = x, y0 = y;
  const double* k = a.k;
  for (int j = 0; j < 5; j++) {
    const double r2 = __dadd_rn(__dmul_rn(x, x), __dmul_rn(y, y));
    const double num = __dadd_rn(1.0, __dmul_rn(__dadd_rn(__dmul_rn(__dadd_rn(__dmul_rn(k[7], r2), k[6]), r2), k[5]), r2));
    const double den = __dadd_rn(1.0, __dmul_rn(__dadd_rn(__dmul_rn(__dadd_rn(__dmul_rn(k[4], r2), k[1]), r2), k[0]), r2));
    const double icdist = __ddiv_rn(num, den);
    if (icdist < 0) {  // test: undistortPoints.regression_14583
      x = __dmul_rn(__dsub_rn(u, a.cx), a.ifx);
      y = __dmul_rn(__dsub_rn(v, a.cy), a.ify);
      break;
    }
    // deltaX = 2*k[2]*x*y + k[3]*(r2 + 2*x*x) + k[8]*r2 + k[9]*r2*r2   (left-to-right)
    const double dx = __dadd_rn(__dadd_rn(__dadd_rn(__dmul_rn(__dmul_rn(__dmul_rn(2.0, k[2]), x), y),
                                                    __dmul_rn(k[3], __dadd_rn(r2, __dmul_rn(__dmul_rn(2.0, x), x)))),
                                          __dmul_rn(k[8], r2)),
                                __dmul_rn(__dmul_rn(k[9], r2), r2));
    // deltaY = k[2]*(r2 + 2*y*y) + 2*k[3]*x*y + k[10]*r2 + k[11]*r2*r2
    const double dy = __dadd_rn(__dadd_rn(__dadd_rn(__dmul_rn(k[2], __dadd_rn(r2, __dmul_rn(__dmul_rn(2.0, y), y))),
                                                    __dmul_rn(__dmul_rn(__dmul_rn(2.0, k[3]), x), y)),
                                          __dmul_rn(k[10], r2)),
                                __dmul_rn(__dmul_rn(k[11], r2), r2));
    x = __dmul_rn(__dsub_rn(x0, dx), icdist);
    y = __dmul_rn(__dsub_rn(y0, dy), icdist);
  }
  const double xx = __dadd_rn(__dadd_rn(__dmul_rn(a.rr[0], x), __dmul_rn(a.rr[1], y)), a.rr[2]);
  const double yy = __dadd_rn(__dadd_rn(__dmul_rn(a.rr[3], x), __dmul_rn(a.rr[4], y)), a.rr[5]);
  const double ww = __ddiv_rn(1.0, __dadd_rn(__dadd_rn(__dmul_rn(a.rr[6], x), __dmul_rn(a.rr[7], y)), a.rr[8]));
  *ox = (float)__dmul_rn(xx, ww);
  *oy = (float)__dmul_rn(yy, ww);
}

// (x, y) pairs at `stride_floats` apart: 2 for a packed point list, 7 for YdKeyPoint arrays (in place on pt.x, pt.y:
// in and out may be the same array, hence no __restrict__)
__global__ void __launch_bounds__(256) k_undistort(const UndistortArgs a, const float* in, float* out,
                                                   int stride_floats, const int32_t* __restrict__ n_per_frame, int capacity,
                                                   int total) {
  const int t = blockIdx.x * blockDim.x + threadIdx.x;
  if (t >= total) return;
  if (n_per_frame) {
    const int b = t / capacity;
    if (t - b * capacity >= n_per_frame[b]) return;
  }
  const size_t o = (size_t)t * stride_floats;
  float x, y;
  undistort_one(a, in[o], in[o + 1], &x, &y);
  out[o] = x;
  out[o + 1] = y;
}

__global__ void __launch_bounds__(256) k_stereo_from_rgbd(const YdKeyPoint* __restrict__ kps, const int32_t* __restrict__ n_per_frame,
                                                          int capacity, int total, const uint8_t* __restrict__ depth_maps,
                                                          int depth_type, int width, int height, int depth_stride,
                                                          size_t depth_frame_stride, float factor, int apply_factor, float bf,
                                                          float* __restrict__ right_x, float* __restrict__ depth) {
  const int t = blockIdx.x * blockDim.x + threadIdx.x;
  if (t >= total) return;
  const int b = t / capacity;
  if (n_per_frame && t - b * capacity >= n_per_frame[b]) return;
  const float px = kps[t].x, py = kps[t].y;
  float rx = -1.f, dz = -1.f;  // src/frame.cpp:215-216
  const int ix = (int)px, iy = (int)py;  // at<float>(float, float): implicit float -> int truncation (:218)
  if (px > -1.f && py > -1.f && ix >= 0 && iy >= 0 && ix < width && iy < height) {
    const uint8_t* row = depth_maps + (size_t)b * depth_frame_stride + (size_t)iy * depth_stride;
    float d;
    if (depth_type == YDORB_DEPTH_U16) {
      d = __fmul_rn((float)__ldg(reinterpret_cast<const uint16_t*>(row) + ix), factor);  // convertTo(CV_32F, factor)
    } else {
      d = __ldg(reinterpret_cast<const float*>(row) + ix);
      if (apply_factor) d = __fmul_rn(d, factor);
    }
    if (d > 0.f) {  // :219
      dz = d;
      rx = __fsub_rn(px, __fdiv_rn(bf, d));  // :221
    }
  }
  right_x[t] = rx;
  depth[t] = dz;
}

}  // namespace ydorb

using namespace ydorb;

namespace {

int make_undistort_args(const float* K, const float* dist, int n_dist, const float* P, UndistortArgs* a) {
  YD_REQUIRE(K && dist && P, "null camera argument");
  YD_REQUIRE(n_dist == 4 || n_dist == 5 || n_dist == 8 || n_dist == 12, "n_dist must be 4, 5, 8 or 12");
  a->fx = (double)K[0];
  a->fy = (double)K[4];
  a->cx = (double)K[2];
  a->cy = (double)K[5];
  a->ifx = 1. / a->fx;
  a->ify = 1. / a->fy;
  for (int i = 0; i < 12; i++) a->k[i] = i < n_dist ? (double)dist[i] : 0.0;
  for (int i = 0; i < 9; i++) a->rr[i] = (double)P[i];  // RR = P * R, R = identity: exact
  return YDORB_OK;
}

}  // namespace

extern "C" {

int ydorb_cvt_gray_device(YdMatcher* m, const uint8_t* d_src, int width, int height, int src_stride, size_t src_frame_stride,
                          int code, uint8_t* d_dst, int dst_stride, size_t dst_frame_stride, int batch) {
  YD_REQUIRE(m && batch >= 0 && width >= 0 && height >= 0, "bad argument");
  YD_REQUIRE(code >= YDORB_RGB2GRAY && code <= YDORB_BGRA2GRAY, "code must be one of YDORB_{RGB,BGR,RGBA,BGRA}2GRAY");
  if (batch == 0 || width == 0 || height == 0) return YDORB_OK;
  YD_REQUIRE(d_src && d_dst, "null image pointer");
  const int ch = code >= YDORB_RGBA2GRAY ? 4 : 3;
  YD_REQUIRE(src_stride >= width * ch && dst_stride >= width, "stride smaller than a row");
  YD_REQUIRE(height <= 65535 && batch <= 65535, "image height / batch too large for one launch");
  YD_CUDA(cudaSetDevice(m->device));
  const int swap = code == YDORB_BGR2GRAY || code == YDORB_BGRA2GRAY;
  dim3 grid(((width + 3) / 4 + 255) / 256, height, batch);
  if (ch == 3)
    k_cvt_gray<3><<<grid, 256, 0, m->stream>>>(d_src, width, height, src_stride, src_frame_stride, swap, d_dst, dst_stride,
                                               dst_frame_stride);
  else
    k_cvt_gray<4><<<grid, 256, 0, m->stream>>>(d_src, width, height, src_stride, src_frame_stride, swap, d_dst, dst_stride,
                                               dst_frame_stride);
  count_launches(1);
  YD_CUDA(cudaGetLastError());
  return YDORB_OK;
}

int ydorb_cvt_gray(YdMatcher* m, const uint8_t* src, int width, int height, int src_stride, int code, uint8_t* dst,
                   int dst_stride) {
  YD_REQUIRE(m && width >= 0 && height >= 0, "bad argument");
  YD_REQUIRE(code >= YDORB_RGB2GRAY && code <= YDORB_BGRA2GRAY, "code must be one of YDORB_{RGB,BGR,RGBA,BGRA}2GRAY");
  if (width == 0 || height == 0) return YDORB_OK;
  YD_REQUIRE(src && dst, "null image pointer");
  const int ch = code >= YDORB_RGBA2GRAY ? 4 : 3;
  YD_REQUIRE(src_stride >= width * ch && dst_stride >= width, "stride smaller than a row");
  YD_CUDA(cudaSetDevice(m->device));
  const size_t srow = ((size_t)width * ch + 15) / 16 * 16, drow = ((size_t)width + 15) / 16 * 16;
  const size_t sb = (srow * height + 255) / 256 * 256;
  int rc = matcher_reserve_io(m, sb + drow * height);
  if (rc != YDORB_OK) return rc;
  uint8_t* io = reinterpret_cast<uint8_t*>(m->d_io);
  YD_CUDA(cudaMemcpy2DAsync(io, srow, src, src_stride, (size_t)width * ch, height, cudaMemcpyHostToDevice, m->stream));
  rc = ydorb_cvt_gray_device(m, io, width, height, (int)srow, 0, code, io + sb, (int)drow, 0, 1);
  if (rc != YDORB_OK) return rc;
  YD_CUDA(cudaMemcpy2DAsync(dst, dst_stride, io + sb, drow, width, height, cudaMemcpyDeviceToHost, m->stream));
  YD_CUDA(cudaStreamSynchronize(m->stream));
  return YDORB_OK;
}

int ydorb_undistort_points(YdMatcher* m, const float* xy, int n, const float* K, const float* dist, int n_dist,
                           const float* P, float* xy_out) {
  YD_REQUIRE(m && n >= 0, "bad argument");
  UndistortArgs a;
  int rc = make_undistort_args(K, dist, n_dist, P, &a);
  if (rc != YDORB_OK) return rc;
  if (n == 0) return YDORB_OK;
  YD_REQUIRE(xy && xy_out, "null point array");
  YD_CUDA(cudaSetDevice(m->device));
  const size_t bytes = sizeof(float) * 2 * (size_t)n;
  rc = matcher_reserve_io(m, bytes);
  if (rc != YDORB_OK) return rc;
  float* io = reinterpret_cast<float*>(m->d_io);
  YD_CUDA(cudaMemcpyAsync(io, xy, bytes, cudaMemcpyHostToDevice, m->stream));
  k_undistort<<<(n + 255) / 256, 256, 0, m->stream>>>(a, io, io, 2, nullptr, 0, n);
  count_launches(1);
  YD_CUDA(cudaGetLastError());
  YD_CUDA(cudaMemcpyAsync(xy_out, io, bytes, cudaMemcpyDeviceToHost, m->stream));
  YD_CUDA(cudaStreamSynchronize(m->stream));
  return YDORB_OK;
}

int ydorb_undistort_keypoints_device(YdMatcher* m, YdKeyPoint* d_keypoints, const int32_t* d_n, int capacity, int batch,
                                     const float* K, const float* dist, int n_dist, const float* P) {
  YD_REQUIRE(m && batch >= 0 && capacity >= 0, "bad argument");
  UndistortArgs a;
  int rc = make_undistort_args(K, dist, n_dist, P, &a);
  if (rc != YDORB_OK) return rc;
  if (dist[0] == 0.0f) return YDORB_OK;  // if (_cvMat_imageDistCoef.at<float>(0) != 0.0)  (src/frame.cpp:194)
  if (batch == 0 || capacity == 0) return YDORB_OK;
  YD_REQUIRE(d_keypoints && d_n, "null device pointer");
  YD_REQUIRE((long long)capacity * batch < (1ll << 31), "batch x capacity overflows");
  YD_CUDA(cudaSetDevice(m->device));
  const int total = capacity * batch;
  static_assert(sizeof(YdKeyPoint) == 7 * sizeof(float), "YdKeyPoint layout");
  float* p = reinterpret_cast<float*>(d_keypoints);
  k_undistort<<<(total + 255) / 256, 256, 0, m->stream>>>(a, p, p, 7, d_n, capacity, total);
  count_launches(1);
  YD_CUDA(cudaGetLastError());
  return YDORB_OK;
}

int ydorb_stereo_from_rgbd_device(YdMatcher* m, const YdKeyPoint* d_keypoints, const int32_t* d_n, int capacity, int batch,
                                  const void* d_depth_maps, int depth_type, int width, int height, int depth_stride,
                                  size_t depth_frame_stride, float depth_map_factor, float bf, float* d_right_x,
                                  float* d_depth) {
  YD_REQUIRE(m && batch >= 0 && capacity >= 0 && width >= 0 && height >= 0, "bad argument");
  YD_REQUIRE(depth_type == YDORB_DEPTH_U16 || depth_type == YDORB_DEPTH_F32, "depth_type must be YDORB_DEPTH_U16 or _F32");
  if (batch == 0 || capacity == 0) return YDORB_OK;
  YD_REQUIRE(d_keypoints && d_right_x && d_depth && (d_depth_maps || width == 0 || height == 0), "null device pointer");
  const int elem = depth_type == YDORB_DEPTH_U16 ? 2 : 4;
  YD_REQUIRE(depth_stride >= width * elem && depth_stride % elem == 0, "bad depth stride");
  YD_REQUIRE(((uintptr_t)d_depth_maps % elem) == 0 && depth_frame_stride % elem == 0, "depth map misaligned");
  YD_REQUIRE((long long)capacity * batch < (1ll << 31), "batch x capacity overflows");
  YD_CUDA(cudaSetDevice(m->device));
  const int total = capacity * batch;
  // src/tracking.cpp:109: convertTo runs when |factor - 1| > 1e-5 or the map is not CV_32F
  const int apply = depth_type == YDORB_DEPTH_U16 || std::fabs(depth_map_factor - 1.0f) > 1e-5;
  k_stereo_from_rgbd<<<(total + 255) / 256, 256, 0, m->stream>>>(d_keypoints, d_n, capacity, total,
                                                                 reinterpret_cast<const uint8_t*>(d_depth_maps), depth_type, width,
                                                                 height, depth_stride, depth_frame_stride, depth_map_factor, apply,
                                                                 bf, d_right_x, d_depth);
  count_launches(1);
  YD_CUDA(cudaGetLastError());
  return YDORB_OK;
}

int ydorb_stereo_from_rgbd(YdMatcher* m, const YdKeyPoint* keypoints, int n, const void* depth_map, int depth_type,
                           int width, int height, int depth_stride, float depth_map_factor, float bf, float* right_x,
                           float* depth) {
  YD_REQUIRE(m && n >= 0 && width >= 0 && height >= 0, "bad argument");
  YD_REQUIRE(depth_type == YDORB_DEPTH_U16 || depth_type == YDORB_DEPTH_F32, "depth_type must be YDORB_DEPTH_U16 or _F32");
  if (n == 0) return YDORB_OK;
  YD_REQUIRE(keypoints && right_x && depth && (depth_map || width == 0 || height == 0), "null pointer");
  const int elem = depth_type == YDORB_DEPTH_U16 ? 2 : 4;
  YD_REQUIRE(depth_stride >= width * elem, "bad depth stride");
  YD_CUDA(cudaSetDevice(m->device));
  auto al = [](size_t b) { return (b + 255) / 256 * 256; };
  const size_t row = al((size_t)width * elem), o_kp = al(row * std::max(height, 1)), o_rx = o_kp + al(sizeof(YdKeyPoint) * (size_t)n),
               o_d = o_rx + al(sizeof(float) * (size_t)n);
  int rc = matcher_reserve_io(m, o_d + al(sizeof(float) * (size_t)n));
  if (rc != YDORB_OK) return rc;
  uint8_t* io = reinterpret_cast<uint8_t*>(m->d_io);
  if (width > 0 && height > 0)
    YD_CUDA(cudaMemcpy2DAsync(io, row, depth_map, depth_stride, (size_t)width * elem, height, cudaMemcpyHostToDevice, m->stream));
  YD_CUDA(cudaMemcpyAsync(io + o_kp, keypoints, sizeof(YdKeyPoint) * (size_t)n, cudaMemcpyHostToDevice, m->stream));
  rc = ydorb_stereo_from_rgbd_device(m, reinterpret_cast<const YdKeyPoint*>(io + o_kp), nullptr, n, 1, io, depth_type, width, height,
                                     (int)row, 0, depth_map_factor, bf, reinterpret_cast<float*>(io + o_rx),
                                     reinterpret_cast<float*>(io + o_d));
  if (rc != YDORB_OK) return rc;
  YD_CUDA(cudaMemcpyAsync(right_x, io + o_rx, sizeof(float) * (size_t)n, cudaMemcpyDeviceToHost, m->stream));
  YD_CUDA(cudaMemcpyAsync(depth, io + o_d, sizeof(float) * (size_t)n, cudaMemcpyDeviceToHost, m->stream));
  YD_CUDA(cudaStreamSynchronize(m->stream));
  return YDORB_OK;
}

}  // extern "C"
